"""CPU self-checks of the BA oracle: symmetric systems, MT vs ST branch agreement (the MT
branch is not run-to-run deterministic, so the ST branch is the parity oracle), state decisions,
J vs numeric differentiation."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def ba(pkg, oracle, window_small):
    W = window_small
    pyr = [oracle.make_images(img, 1)[0][0] for img in W.images]
    B = oracle.BA(W, pyr)
    lin = B.linearize()
    act, jp = B.apply_res()
    return W, B, lin, act, jp


def test_states(ba):
    W, B, lin, act, jp = ba
    st = lin["newState"]
    assert set(np.unique(st)).issubset({0, 1, 2})
    assert (st == 0).sum() > 0.3 * W.nR
    assert np.array_equal(act.astype(bool), st == 0)
    # OOB residuals return before writing energies
    assert (lin["newEnergyWithOutlier"][st == 1] == -1).all()
    # outliers are clamped to the frame threshold
    assert (lin["newEnergy"][st == 2] == 512).all()
    assert (lin["newEnergy"][st == 0] <= 512).all()


def test_systems_symmetric_and_mt_equals_st(ba):
    W, B, lin, act, jp = ba
    A_st = B.accumulate_A(False)
    S_st = B.accumulate_SC(False)
    A_mt = B.accumulate_A(True, 6)
    S_mt = B.accumulate_SC(True, 6)
    N = 8 * W.nF + 4
    assert A_st["H"].shape == (N, N)
    assert np.allclose(A_st["H"], A_st["H"].T, rtol=1e-12, atol=1e-9 * np.abs(A_st["H"]).max())
    assert A_st["resInA"] == int(act.sum()) == A_mt["resInA"]
    sc = np.abs(A_st["H"]).max()
    assert np.abs(A_st["H"] - A_mt["H"]).max() <= 1e-5 * sc
    assert np.abs(A_st["b"] - A_mt["b"]).max() <= 1e-5 * np.abs(A_st["b"]).max()
    sc = np.abs(S_st["H"]).max()
    assert np.abs(S_st["H"] - S_mt["H"]).max() <= 1e-5 * sc
    # Schur complement of a PSD system: H_A - H_sc stays PSD on the pose block (up to float noise)
    M = (A_st["H"] - S_st["H"])[4:, 4:]
    M = 0.5 * (M + M.T)
    assert np.linalg.eigvalsh(M).min() > -1e-3 * np.abs(M).max()


def test_point_sums(ba):
    W, B, lin, act, jp = ba
    A = B.accumulate_A(False)
    S = B.accumulate_SC(False)
    has = np.zeros(W.nP, dtype=bool)
    np.logical_or.at(has, W.res_point, act.astype(bool))
    assert (S["HdiF"][~has] == 0).all()
    H = A["Hdd"] + W.priorF
    ok = has & (H >= 1e-10)
    assert np.allclose(S["HdiF"][ok], 1.0 / H[ok], rtol=1e-6)
    assert np.allclose(S["idepth_hessian"][ok], H[ok], rtol=1e-6)


def test_residual_jacobian_numeric(pkg, oracle, window_small):
    """d(resF)/d(idepth) from J (JIdx * Jpdd) vs a finite difference through the oracle itself"""
    import copy
    W = copy.copy(window_small)
    pyr = [oracle.make_images(img, 1)[0][0] for img in W.images]
    W.idepth_zero_scaled = W.idepth_scaled.copy()       # linearise at the current depth: FEJ point == state
    B0 = oracle.BA(W, pyr)
    l0 = B0.linearize()
    eps = 1e-3
    Wp = copy.copy(W)
    Wp.idepth_scaled = (W.idepth_scaled * (1 + eps)).astype(np.float32)
    Wp.idepth_zero_scaled = Wp.idepth_scaled.copy()
    lp = oracle.BA(Wp, pyr).linearize()
    both = (l0["newState"] == 0) & (lp["newState"] == 0)
    J = l0["J"][both]
    dr = (lp["J"][both][:, :8] - J[:, :8])
    dd = (Wp.idepth_scaled - W.idepth_scaled)[W.res_point][both]
    pred = (J[:, 30:38] * J[:, 28:29] + J[:, 38:46] * J[:, 29:30]) * dd[:, None]   # JIdx0*Jpdd0 + JIdx1*Jpdd1
    # the interpolated central-difference gradient is not the exact derivative of the bilinear intensity,
    # so this is a sanity check (sign / magnitude / correlation), not an equality
    good = np.abs(J[:, :8]).max(axis=1) < 4
    p, d = pred[good].reshape(-1), dr[good].reshape(-1)
    assert np.corrcoef(p, d)[0, 1] > 0.9
    assert 0.7 < np.sum(p * d) / np.sum(p * p) < 1.3
