"""Next rows (SURVEY.md 8f, N1 + the accumulate half of N2) on the GPU vs the oracle (which is pinned against the
reference build in test_oracle_vs_reference.py): fixLinearizationF, accumulateLF_MT, the solveSystemF tail,
resubstituteF_MT and the point half of doStepFromBackup.

Bars: res_toZeroF and the per-point *_accLF sums bit-exact; H_L, b_L and the assembled HFinal/bFinal <= 1e-4 relative
(same bar as H_A / H_sc); the solve itself is checked against numpy on the GPU's own system (<= 1e-9), x and the
per-point steps against the oracle within the conditioning of the system.
"""
import copy
import numpy as np
import pytest
from helpers import rel_err_matrix
from test_gpu_ba import _setup, _check_system
from test_oracle_vs_reference import _solve_inputs

pytestmark = pytest.mark.gpu


def _np_solve(HF, bF):
    sv = 1.0 / np.sqrt(np.diag(HF) + 10)
    return sv * np.linalg.solve(sv[:, None] * HF * sv[None, :], sv * bF)


@pytest.mark.parametrize("win_name,iteration,with_lin", [("small", 0, False), ("small", 3, False), ("small", 3, True),
                                                      ("vga", 0, False), ("vga", 3, True)])
def test_solve_parity(pkg, oracle, request, win_name, iteration, with_lin):
    W = request.getfixturevalue("window_" + win_name)
    B, ctx = _setup(pkg, oracle, W)
    inp = _solve_inputs(W)
    B.linearize(); ctx.ba_linearize()
    act, _ = B.apply_res(); ctx.ba_apply_res()
    if with_lin:
        which = (np.arange(W.nR) % 3 == 0).astype(np.uint8)
        to, tg = B.fix_linearization(which), ctx.ba_fix_linearization(which)
        sel = (which == 1) & (act == 1)
        assert sel.sum() > 100 and np.array_equal(to[sel].view(np.uint32), tg[sel].view(np.uint32))
    B.set_priors(*inp[:6]); B.set_nullspaces(inp[6])
    ctx.ba_set_priors(*inp[:6]); ctx.ba_set_nullspaces(inp[6])
    if with_lin:
        Lo, Lg = B.accumulate_L(False), ctx.ba_accumulate_L()
        assert Lo["resInL"] == Lg["resInL"] == int(sel.sum())
        for key in ("Hdd", "bd", "Hcd"):
            assert np.array_equal(Lo[key].view(np.uint32), Lg[key].view(np.uint32)), key
        _check_system(Lg["H"], Lg["b"], Lo["H"], Lo["b"])
    so = B.solve(iteration, 0.1)
    sg = ctx.ba_solve(iteration, 0.1, debug=True)
    _check_system(sg["HFinal"], sg["bFinal"], so["HFinal"], so["bFinal"])
    # the device LDL^T against the oracle's (Eigen's algorithm: same pivots, lower triangle only) on the device's own system
    if iteration < 2:
        xl = oracle.scaled_ldlt_solve(sg["HFinal"], sg["bFinal"])
        assert np.abs(sg["x"] - xl).max() <= 1e-11 * np.abs(xl).max()
        xn = _np_solve(sg["HFinal"], sg["bFinal"])           # general LU on the full (1e-11-asymmetric) matrix
        assert np.abs(sg["x"] - xn).max() <= 1e-3 * np.abs(xn).max()
    # against the oracle: the two systems differ by the FP32 accumulation order of H_A / H_sc (<= 1e-4 per entry, checked
    # above), which the conditioning of the window amplifies.  Without the null-space projection the difference of the
    # solutions must be the first-order image of the difference of the systems, dx = H^-1 (db - dH x), to 1e-4 of |x|
    # (anything else would be a solver fault); with it (iteration >= 2) the bound is the amplification seen on these windows.
    if iteration < 2:
        dH, db = sg["HFinal"] - so["HFinal"], sg["bFinal"] - so["bFinal"]
        pred = _np_solve(so["HFinal"], db - dH @ so["x"])
        assert np.abs(sg["x"] - so["x"] - pred).max() <= 3e-4 * np.abs(so["x"]).max()      # second-order remainder
    assert np.abs(sg["x"] - so["x"]).max() <= (2e-3 if win_name == "small" else 1e-2) * np.abs(so["x"]).max()
    good = np.abs(so["step"]) > 1e-3 * np.abs(so["step"]).max()
    # the point steps inherit the difference of x through Hcd / Hdd (resubstituteFPt, EnergyFunctional.cpp:292-330)
    assert np.abs(sg["step"] - so["step"])[good].max() <= (2e-3 if win_name == "small" else 3e-2) * np.abs(so["step"]).max()
    assert np.array_equal(sg["step"] == 0, so["step"] == 0)
    ctx.close()


def test_resubstitute_matches_oracle_on_same_x(pkg, oracle, window_small):
    """lambda = 10 without SOLVER_FIX_LAMBDA: a well-conditioned damped system, so x and the per-point steps of
    resubstituteFPt can be compared tightly"""
    W = window_small
    B, ctx = _setup(pkg, oracle, W)
    B.linearize(); ctx.ba_linearize()
    B.apply_res(); ctx.ba_apply_res()
    inp = _solve_inputs(W)
    B.set_priors(*inp[:6]); ctx.ba_set_priors(*inp[:6])
    so = B.solve(0, 10.0, solver_mode=0)
    sg = ctx.ba_solve(0, 10.0, solver_mode=0)
    assert np.abs(sg["x"] - so["x"]).max() <= 1e-4 * np.abs(so["x"]).max()
    assert np.abs(sg["step"] - so["step"]).max() <= 1e-4 * np.abs(so["step"]).max()
    ctx.close()


def test_points_backup_step_restore(pkg, oracle, window_small):
    W = window_small
    B, ctx = _setup(pkg, oracle, W)
    ctx.ba_linearize(); ctx.ba_apply_res()
    ctx.ba_set_priors(np.full(4, 5e9), np.tile([1e10] * 3 + [1e11] * 3 + [1e14, 1e14], W.nF))
    s = ctx.ba_solve(0, 1e-5)
    ctx.ba_points_backup()
    sums, idp = ctx.ba_points_step(0.5)
    assert np.array_equal(idp, (W.idepth_scaled + np.float32(0.5) * s["step"]).astype(np.float32))
    assert abs(sums[0] - float((s["step"].astype(np.float64) ** 2).sum())) <= 1e-6 * sums[0]
    assert abs(sums[1] - float(np.abs(W.idepth_scaled.astype(np.float64)).sum())) <= 1e-6 * sums[1]
    ctx.ba_points_restore()
    _, idp2 = ctx.ba_points_step(0.0)
    assert np.array_equal(idp2, W.idepth_scaled)
    ctx.close()


def test_gauss_newton_iterations_reduce_energy(pkg, oracle, window_small):
    """device-side optimize() loop: linearize + applyRes + solveSystemF + point step per call, frame states updated on
    the host from -x (FullSystem::doStepFromBackup) and re-uploaded through dso_b200_ba_set_frames"""
    W = copy.deepcopy(window_small)
    B, ctx = _setup(pkg, oracle, W)
    cP = np.full(4, 5e9)
    fP = np.zeros((W.nF, 8))
    fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]       # first frame held by its prior (HessianBlocks.h:285-313)
    ctx.ba_set_priors(cP, fP.reshape(-1))
    ctx.ba_set_calib_delta(np.zeros(4))
    energies = []
    for it in range(5):
        ctx.ba_set_frames(W)
        r = ctx.ba_iteration(it, 1e-5)
        energies.append(r["energy"])
        assert np.isfinite(r["x"]).all()
        for f in range(W.nF):
            W.state[f, :8] -= r["x"][4 + 8 * f: 12 + 8 * f]
    ctx.ba_set_frames(W)
    energies.append(ctx.ba_linearize(want_J=False)["total"])
    assert energies[-1] < 0.8 * energies[0], energies
    assert all(b <= a * 1.02 for a, b in zip(energies, energies[1:])), energies
    ctx.close()


def test_marginalize_points_and_energies(pkg, oracle, window_small):
    """rest of N2 on the device: calcLEnergyF_MT, calcMEnergyF, marginalizePointsF vs the oracle (pinned against the reference
    build in test_oracle_vs_reference.py::test_marginalize_points_and_energies)"""
    W = window_small
    B, ctx = _setup(pkg, oracle, W)
    inp = _solve_inputs(W)
    pflags = (np.arange(W.nP) % 3 == 1).astype(np.uint8)
    rflags = pflags[W.res_point]
    B.linearize(); ctx.ba_linearize()
    B.apply_res(); ctx.ba_apply_res()
    B.fix_linearization(rflags); ctx.ba_fix_linearization(rflags)
    B.set_priors(*inp[:6]); ctx.ba_set_priors(*inp[:6])
    Lo, Mo = B.calc_L_energy(1), B.calc_M_energy()
    Lg, Mg = ctx.ba_calc_energies()
    assert abs(Mg - Mo) <= 1e-12 * abs(Mo) and Mo != 0
    assert abs(Lg - Lo) <= 1e-5 * abs(Lo) and Lo != 0           # FP64 tree vs the reference's float tiers
    mo, mg = B.marginalize_points(pflags), ctx.ba_marginalize_points(pflags)
    assert mo["resInM"] == mg["resInM"] > 100
    sel = pflags == 1
    assert np.array_equal(mo["idepth_hessian"][sel].view(np.uint32), mg["idepth_hessian"][sel].view(np.uint32))
    _check_system(mg["H"], mg["b"], mo["H"], mo["b"])
    _check_system(mg["HM"], mg["bM"], mo["HM"], mo["bM"])
    assert np.allclose(mg["HM"], inp[4] + 0.25 * mg["H"], rtol=0, atol=1e-9 * np.abs(mg["HM"]).max())
    ctx.close()


from test_oracle_vs_reference import (_no_first_frame_inputs, SOLVER_DEFAULT, SOLVER_SVD, SOLVER_SVD_CUT7, SOLVER_ORTH_SYSTEM,
                                      SOLVER_ORTH_POINTMARG, SOLVER_ORTH_FULL)


@pytest.mark.parametrize("win_name,mode,first", [("small", SOLVER_DEFAULT | SOLVER_ORTH_SYSTEM, False), ("small", SOLVER_DEFAULT | SOLVER_ORTH_SYSTEM, True),
                                                 ("small", SOLVER_DEFAULT | SOLVER_SVD, True), ("small", SOLVER_DEFAULT | SOLVER_SVD | SOLVER_SVD_CUT7, True),
                                                 ("small", SOLVER_DEFAULT | SOLVER_SVD | SOLVER_ORTH_SYSTEM, False),
                                                 ("vga", SOLVER_DEFAULT | SOLVER_SVD | SOLVER_ORTH_SYSTEM, False), ("vga", SOLVER_DEFAULT | SOLVER_SVD, True)])
def test_solve_non_default_branches(pkg, oracle, request, win_name, mode, first):
    """SOLVER_ORTHOGONALIZE_SYSTEM / SOLVER_SVD / SOLVER_SVD_CUT7 of solveSystemF (EnergyFunctional.cpp:832-900) on the device
    vs the oracle (pinned against the reference build in test_oracle_vs_reference.py::test_solve_system_non_default_branches).
    HFinal / bFinal to the H_A / H_sc bar; x to the first-order image of the system difference, as in test_solve_parity."""
    W = request.getfixturevalue("window_" + win_name)
    pyr = [oracle.make_images(img, 1)[0][0] for img in W.images]
    inp = _solve_inputs(W) if first else _no_first_frame_inputs(W, pyr)
    B, ctx = _setup(pkg, oracle, W)
    B.linearize(); ctx.ba_linearize()
    B.apply_res(); ctx.ba_apply_res()
    B.set_priors(*inp[:6]); B.set_nullspaces(inp[6]); B.set_solver_state(have_first_frame=first)
    ctx.ba_set_priors(*inp[:6]); ctx.ba_set_nullspaces(inp[6]); ctx.ba_set_solver_state(have_first_frame=first)
    so = B.solve(2, 0.1, solver_mode=mode)
    sg = ctx.ba_solve(2, 0.1, solver_mode=mode, debug=True)
    _check_system(sg["HFinal"], sg["bFinal"], so["HFinal"], so["bFinal"])
    assert np.isfinite(sg["x"]).all()
    scale = np.abs(so["x"]).max()
    assert np.abs(sg["x"] - so["x"]).max() <= (2e-3 if win_name == "small" else 1e-2) * scale
    if mode & SOLVER_SVD:
        # the device SVD on the device's own system against numpy's: same truncated pseudo-inverse
        H, b = sg["HFinal"], sg["bFinal"]
        sv = 1.0 / np.sqrt(np.diag(H))
        U, S, Vt = np.linalg.svd(sv[:, None] * H * sv[None, :])
        Ub = U.T @ (sv * b)
        N = len(b)
        for i in range(N):
            if S[i] < 1e-5 * S.max():
                Ub[i] = 0
            if (mode & SOLVER_SVD_CUT7) and i >= N - 7:
                Ub[i] = 0
            else:
                Ub[i] /= S[i]
        xn = sv * (Vt.T @ Ub)
        if first:          # without orthogonalize(x): iteration 2 with X_LATER projects x afterwards, so compare before it
            sg0 = ctx.ba_solve(0, 0.1, solver_mode=mode, debug=True)
            assert np.abs(sg0["x"] - xn).max() <= 1e-9 * np.abs(xn).max()
    good = np.abs(so["step"]) > 1e-3 * np.abs(so["step"]).max()
    assert np.abs(sg["step"] - so["step"])[good].max() <= (1e-2 if win_name == "small" else 3e-2) * np.abs(so["step"]).max()
    ctx.close()


@pytest.mark.parametrize("mode", [SOLVER_DEFAULT | SOLVER_ORTH_POINTMARG, SOLVER_DEFAULT | SOLVER_ORTH_FULL,
                                  SOLVER_DEFAULT | SOLVER_ORTH_POINTMARG | SOLVER_ORTH_FULL])
def test_marginalize_points_orthogonalize_branches(pkg, oracle, window_small, mode):
    """SOLVER_ORTHOGONALIZE_POINTMARG / _FULL of marginalizePointsF (EnergyFunctional.cpp:684-699) on the device vs the oracle"""
    W = window_small
    inp = _solve_inputs(W)
    B, ctx = _setup(pkg, oracle, W)
    pflags = (np.arange(W.nP) % 3 == 1).astype(np.uint8)
    rflags = pflags[W.res_point]
    B.linearize(); ctx.ba_linearize()
    B.apply_res(); ctx.ba_apply_res()
    B.fix_linearization(rflags); ctx.ba_fix_linearization(rflags)
    B.set_priors(*inp[:6]); B.set_nullspaces(inp[6]); B.set_solver_state(have_first_frame=False, marg_solver_mode=mode)
    ctx.ba_set_priors(*inp[:6]); ctx.ba_set_nullspaces(inp[6]); ctx.ba_set_solver_state(have_first_frame=False, marg_solver_mode=mode)
    mo, mg = B.marginalize_points(pflags), ctx.ba_marginalize_points(pflags)
    assert mo["resInM"] == mg["resInM"] > 100
    _check_system(mg["H"], mg["b"], mo["H"], mo["b"])
    _check_system(mg["HM"], mg["bM"], mo["HM"], mo["bM"])
    ctx.close()
