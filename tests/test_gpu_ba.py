"""K3-K5 parity: PointFrameResidual::linearize / applyRes / accumulateAF_MT / accumulateSCF_MT on the GPU vs the
oracle's single-thread branch (the reference's deterministic path).

Bars: states, energies and the RawResidualJacobian bit-exact (the 8-term sums are taken in the reference's order);
H and b <= 1e-4 relative per entry (relative to max(|H_ij|, sqrt(H_ii H_jj)); b_i relative to max|b|).
"""
import numpy as np
import pytest
from helpers import rel_err_matrix

pytestmark = pytest.mark.gpu


def _setup(pkg, oracle, W):
    pyr = [oracle.make_images(img, 1)[0][0] for img in W.images]
    B = oracle.BA(W, pyr)
    ctx = pkg.Context(0, W.w, W.h, W.nF + 1)
    for i, img in enumerate(W.images):
        ctx.make_images(i, img, 1, want_host=False)
    ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
    g = B.glue()
    ctx.ba_set_glue(g["precalc"], g["adHost"], g["adTarget"], g["adHTdeltaF"], g["cDeltaF"])
    ctx.ba_set_points(W)
    ctx.ba_set_residuals(W)
    return B, ctx


def _check_system(Hg, bg, Ho, bo, tol=1e-4):
    assert rel_err_matrix(Hg, Ho).max() <= tol
    assert np.abs(bg - bo).max() <= tol * np.abs(bo).max()


@pytest.mark.parametrize("which", ["small", "vga", "large"])
def test_ba_pipeline_parity(pkg, oracle, request, which):
    """small: 320x240 / 4 key frames / 500 points; vga: BASELINE configs[1]; large: one configs[3] size (1280x1024, ~18 k
    points, ~54 k residuals)"""
    W = request.getfixturevalue("window_" + which)
    B, ctx = _setup(pkg, oracle, W)
    lo = B.linearize()
    lg = ctx.ba_linearize()
    assert np.array_equal(lo["newState"], lg["newState"])
    ok = lo["newState"] != 1
    assert np.array_equal(lo["newEnergy"].view(np.uint32), lg["newEnergy"].view(np.uint32))
    assert np.array_equal(lo["newEnergyWithOutlier"].view(np.uint32), lg["newEnergyWithOutlier"].view(np.uint32))
    assert np.array_equal(lo["J"][ok].view(np.uint32), lg["J"][ok].view(np.uint32))
    assert np.array_equal(lo["projectedTo"][ok].view(np.uint32), lg["projectedTo"][ok].view(np.uint32))
    assert np.array_equal(lo["centerProjectedTo"].view(np.uint32), lg["centerProjectedTo"].view(np.uint32))
    assert abs(lo["total"] - lg["total"]) <= 1e-10 * abs(lo["total"])
    assert (lo["newState"] == 0).sum() > 0.3 * W.nR and (lo["newState"] == 1).sum() > 0     # IN and OOB both exercised

    ao, jo = B.apply_res()
    ag, jg = ctx.ba_apply_res()
    assert np.array_equal(ao, ag)
    assert np.array_equal(jo[ao == 1].view(np.uint32), jg[ag == 1].view(np.uint32))

    Ao = B.accumulate_A(False)
    Ag = ctx.ba_accumulate_A()
    assert Ao["resInA"] == Ag["resInA"] == int(ao.sum())
    assert np.array_equal(Ao["Hdd"].view(np.uint32), Ag["Hdd"].view(np.uint32))
    assert np.array_equal(Ao["bd"].view(np.uint32), Ag["bd"].view(np.uint32))
    assert np.array_equal(Ao["Hcd"].view(np.uint32), Ag["Hcd"].view(np.uint32))
    _check_system(Ag["H"], Ag["b"], Ao["H"], Ao["b"])
    assert np.allclose(Ag["H"], Ag["H"].T, rtol=0, atol=1e-9 * np.abs(Ag["H"]).max())

    So = B.accumulate_SC(False)
    Sg = ctx.ba_accumulate_SC()
    assert np.array_equal(So["HdiF"].view(np.uint32), Sg["HdiF"].view(np.uint32))
    assert np.array_equal(So["bdSumF"].view(np.uint32), Sg["bdSumF"].view(np.uint32))
    assert np.array_equal(So["idepth_hessian"].view(np.uint32), Sg["idepth_hessian"].view(np.uint32))
    _check_system(Sg["H"], Sg["b"], So["H"], So["b"])

    # second GN-style pass: states now carry over (OOB stays OOB, J buffers flip back)
    lo2, lg2 = B.linearize(), ctx.ba_linearize()
    assert np.array_equal(lo2["newState"], lg2["newState"])
    ok2 = lo2["newState"] != 1
    assert np.array_equal(lo2["J"][ok2].view(np.uint32), lg2["J"][ok2].view(np.uint32))
    B.apply_res(); ctx.ba_apply_res()
    Ao2, Ag2 = B.accumulate_A(False), ctx.ba_accumulate_A()
    _check_system(Ag2["H"], Ag2["b"], Ao2["H"], Ao2["b"])
    ctx.close()


def test_ba_fused_step_equals_separate_calls(pkg, oracle, window_small):
    W = window_small
    B, ctx = _setup(pkg, oracle, W)
    st = ctx.ba_step()
    B.linearize(); B.apply_res()
    Ao, So = B.accumulate_A(False), B.accumulate_SC(False)
    _check_system(st["H_A"], st["b_A"], Ao["H"], Ao["b"])
    _check_system(st["H_sc"], st["b_sc"], So["H"], So["b"])
    assert st["resInA"] == Ao["resInA"]
    # the per-point terms the fused step leaves on the device (per-point sums written with the Jacobians, summed and prepared
    # inside the Schur accumulation): bit for bit the oracle's
    pt = ctx.ba_get_point_terms()
    for key, ref in (("Hdd", Ao["Hdd"]), ("bd", Ao["bd"]), ("Hcd", Ao["Hcd"]), ("HdiF", So["HdiF"]), ("bdSumF", So["bdSumF"]),
                     ("idepth_hessian", So["idepth_hessian"])):
        assert np.array_equal(pt[key].reshape(-1).view(np.uint32), np.asarray(ref, dtype=np.float32).reshape(-1).view(np.uint32)), key
    ctx.close()


def test_ba_input_validation(pkg, oracle, window_small):
    import copy
    W = copy.copy(window_small)
    B, ctx = _setup(pkg, oracle, W)
    W2 = copy.copy(W)
    W2.res_point = W.res_point[::-1].copy()           # not grouped by point
    with pytest.raises(pkg.DsoError):
        ctx.ba_set_residuals(W2)
    W3 = copy.copy(W)
    W3.res_isLinearized = np.ones(W.nR, dtype=np.uint8)
    W3.res_isActive = np.ones(W.nR, dtype=np.uint8)   # active linearised residuals: not implemented (N2), must fail loudly
    with pytest.raises(pkg.DsoError):
        ctx.ba_set_residuals(W3)
    ctx.close()


@pytest.mark.parametrize("which", ["small", "vga"])
def test_ba_set_frames_matches_oracle_glue(pkg, oracle, window_small, window_vga, which):
    """FrameHessian::setState + FrameFramePrecalc::set (HessianBlocks.h:183-199, HessianBlocks.cpp:213-250) +
    EnergyFunctional::setAdjointsF / setDeltaF (EnergyFunctional.cpp:47-114, 178-205) inside the library vs the oracle
    (itself bit-identical to the reference build, test_oracle_vs_reference.py::test_ba_linearize_accumulate): every
    float / double of the glue bit for bit, and therefore the same linearisation and system as the set_glue path.
    This is the path bench.py and host/dso_b200_host.h (setPrecalcValues) use."""
    W = window_small if which == "small" else window_vga
    B, ctx = _setup(pkg, oracle, W)
    g = B.glue()
    out = ctx.ba_set_frames(W, want_out=True)
    assert np.array_equal(out["precalc"].view(np.uint32), np.ascontiguousarray(g["precalc"]).reshape(-1, 28).view(np.uint32))
    assert np.array_equal(out["adHost"].reshape(-1).view(np.uint64), np.ascontiguousarray(g["adHost"]).reshape(-1).view(np.uint64))
    assert np.array_equal(out["adTarget"].reshape(-1).view(np.uint64), np.ascontiguousarray(g["adTarget"]).reshape(-1).view(np.uint64))
    # and the pipeline driven by it: states / energies / Jacobians bit-exact, the system within the H, b bar
    lo, lg = B.linearize(), ctx.ba_linearize()
    assert np.array_equal(lo["newState"], lg["newState"])
    ok = lo["newState"] != 1
    assert np.array_equal(lo["J"][ok].view(np.uint32), lg["J"][ok].view(np.uint32))
    assert np.array_equal(lo["newEnergy"].view(np.uint32), lg["newEnergy"].view(np.uint32))
    st = ctx.ba_step()
    B.linearize(); B.apply_res()
    Ao, So = B.accumulate_A(False), B.accumulate_SC(False)
    _check_system(st["H_A"], st["b_A"], Ao["H"], Ao["b"])
    _check_system(st["H_sc"], st["b_sc"], So["H"], So["b"])
    ctx.close()


@pytest.mark.parametrize("modes", [(-1.0, -1.0), (1e12, -1.0), (-1.0, 1e8)])
def test_ba_fixed_affine_modes(pkg, oracle, window_small, modes):
    """Residuals.cpp:257-258: the stored JabF rows are zero for a fixed parameter; everything downstream follows"""
    W = window_small
    B, ctx = _setup(pkg, oracle, W)
    try:
        oracle.set_affine_modes(*modes)
        ctx.ba_set_affine_modes(*modes)
        lo, lg = B.linearize(), ctx.ba_linearize()
        ok = lo["newState"] != 1
        assert np.array_equal(lo["newState"], lg["newState"])
        assert np.array_equal(lo["J"][ok].view(np.uint32), lg["J"][ok].view(np.uint32))
        B.apply_res(); ctx.ba_apply_res()
        Ao, Ag = B.accumulate_A(False), ctx.ba_accumulate_A()
        _check_system(Ag["H"], Ag["b"], Ao["H"], Ao["b"])
        So, Sg = B.accumulate_SC(False), ctx.ba_accumulate_SC()
        _check_system(Sg["H"], Sg["b"], So["H"], So["b"])
    finally:
        oracle.set_affine_modes()
        ctx.close()


@pytest.mark.parametrize("which", ["small", "vga"])
def test_per_object_linearize_forwarder(pkg, oracle, request, which):
    """dso_b200_ba_linearize_residuals: the per-object PointFrameResidual::linearize (Residuals.h:95) / resetOOB / applyRes
    sequence of FullSystem::flagPointsForRemoval (FullSystem.cpp:776-787) as one batch of k residuals, against the oracle
    (pinned to the reference build in test_oracle_vs_reference.py::test_per_object_linearize_apply)"""
    W = request.getfixturevalue("window_" + which)
    B, ctx = _setup(pkg, oracle, W)
    B.linearize(); B.apply_res()
    ctx.ba_linearize(); ctx.ba_apply_res()
    idx = np.nonzero(W.res_point % 7 == 3)[0].astype(np.int32)
    po = B.linearize_residuals(idx, reset_oob=True, apply=True)
    pg = ctx.ba_linearize_residuals(idx, reset_oob=True, apply=True)
    assert np.array_equal(po["newState"], pg["newState"]) and np.array_equal(po["isActive"], pg["isActive"])
    assert np.array_equal(po["newEnergy"].view(np.uint32), pg["newEnergy"].view(np.uint32))
    assert (po["newState"] == 0).sum() > 0
    # the rest of the window is untouched and the swapped-in Jacobians feed the next accumulation
    Ao, Ag = B.accumulate_A(False), ctx.ba_accumulate_A()
    assert Ao["resInA"] == Ag["resInA"]
    for key in ("Hdd", "bd", "Hcd"):
        assert np.array_equal(Ao[key].view(np.uint32), Ag[key].view(np.uint32)), key
    _check_system(Ag["H"], Ag["b"], Ao["H"], Ao["b"])
    # a batch of one
    p1o = B.linearize_residuals(idx[:1], reset_oob=True, apply=False)
    p1g = ctx.ba_linearize_residuals(idx[:1], reset_oob=True, apply=False)
    assert p1o["newState"][0] == p1g["newState"][0] and p1o["newEnergy"].view(np.uint32)[0] == p1g["newEnergy"].view(np.uint32)[0]
    ctx.close()
