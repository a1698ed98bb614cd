"""Pins the oracle restatement against the REAL reference code: oracle/_ref/libdso_ref.so is the reference's own
translation units (CoarseTracker.cpp, HessianBlocks.cpp, Residuals.cpp, EnergyFunctional*.cpp, Accumulated*.cpp,
MatrixAccumulators.h, ...) compiled unmodified from /root/reference against stand-in Eigen/Sophus/Boost headers
(oracle/refbuild).  Everything the reference computes in its own source is compared bit for bit; only results that
pass through the stand-in third-party arithmetic (LDLT / SE3::exp inside trackNewestCoarse) get a tolerance."""
import numpy as np
import pytest
from helpers import flat

import __graft_entry__ as g

O = g.load_oracle()
pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref/libdso_ref.so not built (needs /root/reference)")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32 if a.dtype == np.float32 else np.uint64)


@pytest.mark.parametrize("w,h,levels", [(640, 480, 4), (320, 240, 3), (424, 320, 3)])
def test_make_images_bit_exact(w, h, levels):
    rng = np.random.default_rng(w)
    img = rng.uniform(0, 255, (h, w)).astype(np.float32)
    B = (np.arange(256, dtype=np.float32) ** 1.05).astype(np.float32)
    for lut in (None, B):
        d0, a0 = O.make_images(img, levels, lut)
        d1, a1 = O.make_images(img, levels, lut, kind="ref")
        for l in range(levels):
            assert np.array_equal(bits(d0[l]), bits(d1[l])) and np.array_equal(bits(a0[l]), bits(a1[l]))


@pytest.fixture(scope="module", params=["small", "vga"])
def win(request):
    """every pin runs on the small window (320x240, 4 key frames, 500 points) AND on the BASELINE window
    (640x480, 7 key frames, 2002 points, 12012 residuals) the GPU parity tests and bench.py use"""
    return request.getfixturevalue("window_" + request.param)


@pytest.fixture(scope="module")
def trackers(pkg, win):
    W = win
    syn = pkg.synthetic
    levels = syn.pyr_levels_for(W.w, W.h)
    ref_pyr = O.make_images(W.images[-1], levels)
    cu, cv, cid, hdi = syn.tracker_reference_inputs(W)
    a0, b0 = W.aff_true[-1]
    frames = syn.tracking_sequence(W, 2)
    out = []
    for kind in ("oracle", "ref"):
        t = O.Tracker(W.w, W.h, levels, *W.K, kind=kind)
        t.set_ref(cu, cv, cid, hdi, flat(ref_pyr[0]), a0, b0, 1.0)
        out.append(t)
    return W, syn, levels, frames, out[0], out[1], (a0, b0)


def test_make_coarse_depth_and_K_bit_exact(trackers):
    W, syn, levels, frames, to, tr, _ = trackers
    for l in range(levels):
        co, cr = to.get_cloud(l), tr.get_cloud(l)
        assert len(co[0]) == len(cr[0]) > 0
        for a, b in zip(co, cr):
            assert np.array_equal(bits(a), bits(b))
        (ko, Kio), (kr, Kir) = to.get_K(l), tr.get_K(l)
        assert ko == kr and np.array_equal(bits(Kio), bits(Kir))


def test_calc_res_and_gs_bit_exact(trackers):
    W, syn, levels, frames, to, tr, (a0, b0) = trackers
    rng = np.random.default_rng(3)
    for k, (img, rel, aff) in enumerate(frames):
        pyr = O.make_images(img, levels)
        to.set_new_frame(flat(pyr[0]), 1.0)
        tr.set_new_frame(flat(pyr[0]), 1.0)
        for pose in ((np.eye(3), np.zeros(3)), rel, syn.se3_mul(syn.se3_exp(rng.normal(0, 3e-3, 6)), rel)):
            for lvl in range(levels):
                for cutoff in (20.0, 80.0):
                    p = syn.pose12(pose)
                    ro = to.calc_res(lvl, p, aff[0], aff[1], cutoff)
                    rr = tr.calc_res(lvl, p, aff[0], aff[1], cutoff)
                    assert np.array_equal(bits(ro), bits(rr)), (k, lvl, ro, rr)
                    no, bo = to.get_warped()
                    nr, br = tr.get_warped()
                    assert no == nr
                    for x, y in zip(bo, br):
                        assert np.array_equal(bits(x), bits(y))
                    Ho, go = to.calc_gs(lvl, aff[0], aff[1])
                    Hr, gr = tr.calc_gs(lvl, aff[0], aff[1])
                    assert np.array_equal(bits(Ho), bits(Hr)) and np.array_equal(bits(go), bits(gr))   # Accumulator9 tiers and lanes included


def test_track_newest_coarse_agrees(trackers):
    """the LM loop goes through the stand-in LDLT / SE3::exp in both builds (same algorithms) -> tight tolerance"""
    W, syn, levels, frames, to, tr, (a0, b0) = trackers
    ident = syn.pose12((np.eye(3), np.zeros(3)))
    for k, (img, rel, aff) in enumerate(frames):
        pyr = O.make_images(img, levels)
        to.set_new_frame(flat(pyr[0]), 1.0)
        tr.set_new_frame(flat(pyr[0]), 1.0)
        ro = to.track(ident, a0, b0, levels - 1)
        rr = tr.track(ident, a0, b0, levels - 1)
        assert ro.ok == rr.ok == 1
        assert np.abs(np.array(ro.pose) - np.array(rr.pose)).max() < 1e-9
        assert abs(ro.aff_a - rr.aff_a) < 1e-9 and abs(ro.aff_b - rr.aff_b) < 1e-6
        assert np.allclose(list(ro.lastResiduals)[:levels], list(rr.lastResiduals)[:levels], rtol=1e-6)
        assert np.allclose(list(ro.lastFlowIndicators), list(rr.lastFlowIndicators), rtol=1e-6)
        thr = np.array(ro.lastResiduals) * 0.3
        assert to.track(ident, a0, b0, levels - 1, thr).ok == tr.track(ident, a0, b0, levels - 1, thr).ok == 0


def test_ba_linearize_accumulate(pkg, win):
    W = win
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    Bo, Br = O.BA(W, pyr), O.BA(W, pyr, kind="ref")
    go, gr = Bo.glue(), Br.glue()
    for key in ("precalc", "adHost", "adTarget", "adHTdeltaF", "cDeltaF"):
        assert np.array_equal(bits(go[key]), bits(gr[key])), key          # FrameFramePrecalc::set, setAdjointsF
    lo, lr = Bo.linearize(), Br.linearize()
    assert np.array_equal(lo["newState"], lr["newState"])
    ok = lo["newState"] != 1
    for key in ("newEnergy", "newEnergyWithOutlier"):
        assert np.array_equal(bits(lo[key]), bits(lr[key])), key
    assert np.array_equal(bits(lo["J"][ok]), bits(lr["J"][ok]))
    assert np.array_equal(bits(lo["projectedTo"][ok]), bits(lr["projectedTo"][ok]))
    assert lo["total"] == lr["total"]
    ao, jo = Bo.apply_res()
    ar, jr = Br.apply_res()
    assert np.array_equal(ao, ar) and np.array_equal(bits(jo[ao == 1]), bits(jr[ar == 1]))
    Ao, Ar = Bo.accumulate_A(False), Br.accumulate_A(False)
    assert Ao["resInA"] == Ar["resInA"]
    for key in ("Hdd", "bd", "Hcd"):
        assert np.array_equal(bits(Ao[key]), bits(Ar[key])), key
    sc = np.abs(Ar["H"]).max()
    assert np.abs(Ao["H"] - Ar["H"]).max() <= 1e-12 * sc and np.abs(Ao["b"] - Ar["b"]).max() <= 1e-12 * np.abs(Ar["b"]).max()
    So, Sr = Bo.accumulate_SC(False), Br.accumulate_SC(False)
    for key in ("HdiF", "bdSumF", "idepth_hessian"):
        assert np.array_equal(bits(So[key]), bits(Sr[key])), key
    sc = np.abs(Sr["H"]).max()
    assert np.abs(So["H"] - Sr["H"]).max() <= 1e-12 * sc and np.abs(So["b"] - Sr["b"]).max() <= 1e-12 * np.abs(Sr["b"]).max()
    # the reference's thread-pool branch (6 workers, chunks of 50) agrees with its single-thread branch to float noise
    Am = Br.accumulate_A(True)
    assert np.abs(Am["H"] - Ar["H"]).max() <= 1e-5 * np.abs(Ar["H"]).max()


@pytest.mark.parametrize("modes", [(-1.0, -1.0), (1e12, -1.0), (-1.0, 1e8)])
def test_fixed_affine_modes_agree(trackers, pkg, win, modes):
    """setting_affineOptModeA/B < 0 (CoarseTracker.cpp:621-643, :743-744; Residuals.cpp:257-258)"""
    W, syn, levels, frames, to, tr, (a0, b0) = trackers
    ident = syn.pose12((np.eye(3), np.zeros(3)))
    img = frames[0][0]
    pyr = O.make_images(img, levels)
    to.set_new_frame(flat(pyr[0]), 1.0)
    tr.set_new_frame(flat(pyr[0]), 1.0)
    try:
        O.set_affine_modes(*modes)
        O.set_affine_modes(*modes, kind="ref")
        ro = to.track(ident, a0, b0, levels - 1)
        rr = tr.track(ident, a0, b0, levels - 1)
        assert ro.ok == rr.ok
        assert np.abs(np.array(ro.pose) - np.array(rr.pose)).max() < 1e-9
        assert abs(ro.aff_a - rr.aff_a) < 1e-9 and abs(ro.aff_b - rr.aff_b) < 1e-6
        if modes[0] < 0:
            assert ro.aff_a == rr.aff_a == 0
        if modes[1] < 0:
            assert ro.aff_b == rr.aff_b == 0
        p0 = [O.make_images(im, 1)[0][0] for im in win.images]
        lo, lr = O.BA(win, p0).linearize(), O.BA(win, p0, kind="ref").linearize()
        ok = lo["newState"] != 1
        assert np.array_equal(bits(lo["J"][ok]), bits(lr["J"][ok]))
        J = lo["J"][lo["newState"] == 0]
        assert len(J) > 0
        if modes[0] < 0:
            assert not J[:, 46:54].any()
        else:
            assert J[:, 46:54].any()
        if modes[1] < 0:
            assert not J[:, 54:62].any()
    finally:
        O.set_affine_modes()
        O.set_affine_modes(kind="ref")


def _solve_inputs(W, seed=0):
    """priors as FrameHessian::getPrior gives them for frameID 0 (HessianBlocks.h:285-313), a synthetic marginalisation
    prior HM/bM, frame deltas and seven null-space vectors"""
    rng = np.random.default_rng(seed)
    nF, N = W.nF, 8 * W.nF + 4
    cP = np.full(4, 5e9)
    fP = np.zeros((nF, 8))
    fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]
    dP = rng.normal(0, 1e-4, (nF, 8))
    dl = rng.normal(0, 1e-3, (nF, 8))
    A = rng.normal(0, 30.0, (N, 12))
    HM = A @ A.T
    HM[:4, :] *= 10
    HM[:, :4] *= 10
    bM = rng.normal(0, 50.0, N)
    ns = rng.normal(0, 1.0, (7, N))
    return cP, fP.reshape(-1), dP.reshape(-1), dl.reshape(-1), HM, bM, ns


@pytest.mark.parametrize("iteration,with_lin", [(0, False), (3, False), (3, True)])
def test_solve_system_and_resubstitute(pkg, win, iteration, with_lin):
    """EnergyFunctional::solveSystemF + resubstituteF_MT (EnergyFunctional.cpp:802-924, 276-330), default solver mode;
    with_lin also runs fixLinearizationF on a third of the residuals so that accumulateLF_MT (addPoint<1>) has work"""
    W = win
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    Bo, Br = O.BA(W, pyr), O.BA(W, pyr, kind="ref")
    inp = _solve_inputs(W)
    outs = []
    for B in (Bo, Br):
        B.linearize()
        act, _ = B.apply_res()
        if with_lin:
            which = (np.arange(W.nR) % 3 == 0).astype(np.uint8)
            outs.append(B.fix_linearization(which)[(which == 1) & (act == 1)])      # the rest is uninitialised in the reference
        B.set_priors(*inp[:6])
        B.set_nullspaces(inp[6])
    if with_lin:
        assert np.array_equal(bits(outs[0]), bits(outs[1])) and np.abs(outs[0]).max() > 0
        Lo, Lr = Bo.accumulate_L(False), Br.accumulate_L(False)
        assert Lo["resInL"] == Lr["resInL"] > 0
        for key in ("Hdd", "bd", "Hcd"):
            assert np.array_equal(bits(Lo[key]), bits(Lr[key])), key
        assert np.abs(Lo["H"] - Lr["H"]).max() <= 1e-12 * np.abs(Lr["H"]).max()
        assert np.abs(Lo["b"] - Lr["b"]).max() <= 1e-12 * np.abs(Lr["b"]).max()
    so, sr = Bo.solve(iteration, 1e-1), Br.solve(iteration, 1e-1)
    assert np.abs(so["x"]).max() > 0
    assert np.abs(so["x"] - sr["x"]).max() <= 1e-9 * np.abs(sr["x"]).max()
    assert np.abs(so["step"] - sr["step"]).max() <= 1e-5 * np.abs(sr["step"]).max()
    assert np.isfinite(so["step"]).all() and np.abs(so["step"]).max() > 0


SOLVER_SVD, SOLVER_ORTH_SYSTEM, SOLVER_ORTH_POINTMARG, SOLVER_ORTH_FULL, SOLVER_SVD_CUT7 = 1, 2, 4, 8, 16
SOLVER_DEFAULT = 128 | 2048


def _no_first_frame_inputs(W, pyr):
    """a window whose frame 0 is not the sequence's first frame: no gauge prior on any frame (FrameHessian::getPrior gives
    zeros for frameID != 0), the seven near-null directions of the active system as null spaces (what
    FullSystem::getNullspaces hands over is that space), and a marginalisation prior of full rank"""
    inp = list(_solve_inputs(W))
    N = 8 * W.nF + 4
    inp[1] = np.zeros_like(inp[1])
    B = O.BA(W, pyr)
    B.linearize(); B.apply_res()
    B.set_priors(inp[0], inp[1], inp[2], inp[3], np.zeros((N, N)), np.zeros(N))
    H0 = B.solve(0, 0.0, solver_mode=64)["HFinal"]
    w, V = np.linalg.eigh(0.5 * (H0 + H0.T))
    assert w[6] < 1e-6 * w[-1]
    inp[6] = np.ascontiguousarray(V[:, :7].T)
    inp[4] = inp[4] + np.eye(N) * 1e3
    return inp


@pytest.mark.parametrize("mode,first", [(SOLVER_DEFAULT | SOLVER_ORTH_SYSTEM, False), (SOLVER_DEFAULT | SOLVER_ORTH_SYSTEM, True),
                                        (SOLVER_DEFAULT | SOLVER_SVD, True), (SOLVER_DEFAULT | SOLVER_SVD | SOLVER_SVD_CUT7, True),
                                        (SOLVER_DEFAULT | SOLVER_SVD | SOLVER_ORTH_SYSTEM, False)])
def test_solve_system_non_default_branches(pkg, win, mode, first):
    """the SVD / SVD_CUT7 / ORTHOGONALIZE_SYSTEM branches of solveSystemF (EnergyFunctional.cpp:832-900) with and without the
    gauge-fixing first frame in the window.  (The reference build's JacobiSVD is the stand-in Hestenes SVD, so these branches
    are pinned up to the SVD algorithm: x agrees to 1e-7, not to rounding.)"""
    W = win
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    inp = _solve_inputs(W) if first else _no_first_frame_inputs(W, pyr)
    Bo, Br = O.BA(W, pyr), O.BA(W, pyr, kind="ref")
    for B in (Bo, Br):
        B.linearize()
        B.apply_res()
        B.set_priors(*inp[:6])
        B.set_nullspaces(inp[6])
        B.set_solver_state(have_first_frame=first)
    so, sr = Bo.solve(2, 1e-1, solver_mode=mode), Br.solve(2, 1e-1, solver_mode=mode)
    assert np.isfinite(so["x"]).all() and np.abs(so["x"]).max() > 0
    assert np.abs(so["x"] - sr["x"]).max() <= 1e-7 * np.abs(sr["x"]).max()
    assert np.abs(so["step"] - sr["step"]).max() <= 1e-4 * np.abs(sr["step"]).max()
    # the branch really is a different computation
    sd = Bo.solve(2, 1e-1)
    assert np.abs(so["x"] - sd["x"]).max() > 1e-9 * np.abs(sd["x"]).max()


@pytest.mark.parametrize("mode", [SOLVER_DEFAULT | SOLVER_ORTH_POINTMARG, SOLVER_DEFAULT | SOLVER_ORTH_FULL])
def test_marginalize_points_orthogonalize_branches(pkg, win, mode):
    """SOLVER_ORTHOGONALIZE_POINTMARG / _FULL of marginalizePointsF (EnergyFunctional.cpp:684-699), first frame gone"""
    W = win
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    inp = _solve_inputs(W)            # random "null-space" vectors: with the true ones H P ~ 0 and the projection is a no-op
    pflags = (np.arange(W.nP) % 3 == 1).astype(np.uint8)
    rflags = pflags[W.res_point]

    def run(kind, m):
        B = O.BA(W, pyr, kind=kind)
        B.linearize()
        B.apply_res()
        B.fix_linearization(rflags)
        B.set_priors(*inp[:6])
        B.set_nullspaces(inp[6])
        B.set_solver_state(have_first_frame=False, marg_solver_mode=m)
        return B.marginalize_points(pflags)
    mo, mr, md = run("oracle", mode), run("ref", mode), run("oracle", SOLVER_DEFAULT)
    assert np.abs(mo["HM"] - mr["HM"]).max() <= 1e-9 * np.abs(mr["HM"]).max()
    assert np.abs(mo["bM"] - mr["bM"]).max() <= 1e-9 * np.abs(mr["bM"]).max()
    assert np.abs(mo["HM"] - md["HM"]).max() > 1e-9 * np.abs(md["HM"]).max()       # the projection changed the prior


def test_marginalize_points_and_energies(pkg, win):
    """rest of N2: calcLEnergyF_MT, calcMEnergyF and marginalizePointsF (EnergyFunctional.cpp:334-428, 642-703).  A third of
    the points are flagged; their residuals are linearised first (FullSystem::flagPointsForRemoval does the same)."""
    W = win
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    Bo, Br = O.BA(W, pyr), O.BA(W, pyr, kind="ref")
    inp = _solve_inputs(W)
    pflags = (np.arange(W.nP) % 3 == 1).astype(np.uint8)
    rflags = pflags[W.res_point]
    res = []
    for B in (Bo, Br):
        B.linearize()
        B.apply_res()
        B.fix_linearization(rflags)
        B.set_priors(*inp[:6])
        res.append((B.calc_L_energy(6), B.calc_M_energy()))
    (Lo, Mo), (Lr, Mr) = res
    assert abs(Mo - Mr) <= 1e-12 * abs(Mr) and Mr != 0
    assert abs(Lo - Lr) <= 1e-5 * abs(Lr) and Lr != 0          # float tiers; the reference's chunk -> worker assignment is dynamic
    mo, mr = Bo.marginalize_points(pflags), Br.marginalize_points(pflags)
    assert mo["resInM"] == mr["resInM"] > 100
    sel = pflags == 1
    assert np.array_equal(bits(mo["idepth_hessian"][sel]), bits(mr["idepth_hessian"][sel]))
    sc = np.abs(mr["H"]).max()
    assert np.abs(mo["H"] - mr["H"]).max() <= 1e-9 * sc and sc > 0      # recovered from HM differences in the reference build
    assert np.abs(mo["b"] - mr["b"]).max() <= 1e-9 * np.abs(mr["b"]).max()
    assert np.abs(mo["HM"] - mr["HM"]).max() <= 1e-12 * np.abs(mr["HM"]).max()
    assert np.abs(mo["bM"] - mr["bM"]).max() <= 1e-12 * np.abs(mr["bM"]).max()


def test_immature_point_init_and_trace_on_bit_exact(pkg, win):
    """next row N3: ImmaturePoint's constructor and traceOn (FullSystem/ImmaturePoint.cpp:25-57, 77-438) over three
    successive frames (interval (0, NaN) first, then the refined intervals), every output bit for bit"""
    W = win
    sc = pkg.synthetic.immature_scenario(W, host_idx=1, n_points=600, n_frames=3)
    hostI = O.make_images(sc["host_image"], 1)[0][0]
    fo, fr = O.immature_init(hostI, sc["u"], sc["v"]), O.immature_init(hostI, sc["u"], sc["v"], kind="ref")
    for k in fo:
        assert np.array_equal(bits(fo[k]), bits(fr[k])), k
    so, sr = O.trace_state(len(sc["u"])), O.trace_state(len(sc["u"]))
    seen = set()
    for fr_ in sc["frames"]:
        dI = O.make_images(fr_["image"], 1)[0][0]
        O.trace_on(dI, sc["u"], sc["v"], fo, so, fr_["KRKi"], fr_["Kt"], fr_["aff"])
        O.trace_on(dI, sc["u"], sc["v"], fo, sr, fr_["KRKi"], fr_["Kt"], fr_["aff"], kind="ref")
        assert np.array_equal(so["status"], sr["status"])
        for k in ("idepth_min", "idepth_max", "quality", "uv", "interval"):
            assert np.array_equal(bits(so[k]), bits(sr[k])), k
        seen |= set(so["status"].tolist())
    assert {0, 1}.issubset(seen) and len(seen) >= 3, seen          # GOOD, OOB and at least one of SKIPPED / BADCONDITION / OUTLIER
    good = so["status"] == 0
    assert good.sum() > (100 if W.w == 320 else 10)      # the VGA scenario's longer baseline leaves mostly OOB / OUTLIER / SKIPPED exits
    mid = 0.5 * (so["idepth_min"][good] + so["idepth_max"][good])
    if W.w == 320:          # scenario sanity (not parity): on the small window the search finds the true depth
        assert np.median(np.abs(mid - sc["idepth_true"][good]) / sc["idepth_true"][good]) < 0.05


def test_optimize_immature_point_bit_exact(pkg, win):
    """next row N3: FullSystem::optimizeImmaturePoint + ImmaturePoint::linearizeResidual (FullSystemOptPoint.cpp:51-209,
    ImmaturePoint.cpp:499-570): result code, activated inverse depth and residual states, bit for bit"""
    W = win
    sc = pkg.synthetic.activation_scenario(W, host_idx=1, n_points=600)
    fields = O.immature_init(O.make_images(sc["host_image"], 1)[0][0], sc["u"], sc["v"])
    dIs = [O.make_images(im, 1)[0][0] for im in sc["images"]]
    args = (W.K, dIs, sc["RTll"], sc["tTll"], sc["aff"], sc["u"], sc["v"], fields, sc["idepth_min"], sc["idepth_max"])
    ro, io, so = O.optimize_immature(*args, min_obs=1)
    rr, ir, sr = O.optimize_immature(*args, min_obs=1, kind="ref")
    assert np.array_equal(ro, rr)
    ok = ro == 1
    assert ok.sum() > 200 and (ro == 0).sum() + (ro == -1).sum() > 10, np.bincount(ro + 1)
    assert np.array_equal(bits(io[ok]), bits(ir[ok]))
    assert np.array_equal(so[ro != 0], sr[rr != 0])
    good = ok & ~sc["wrong"]
    assert np.median(np.abs(io[good] - sc["idepth_true"][good]) / sc["idepth_true"][good]) < 0.01      # converges to the true depth


def test_non_default_settings_bit_exact(pkg, window_small):
    """the run-time settings of util/settings.cpp that the hot path reads (huberTH :127, outlierTHSumComponent :91, maxPixSearch
    :156, the setting_trace_* family :160-165, minTraceTestRadius :158, GNItsOnPointActivation :159, minIdepthH_act :67) changed
    in both builds: the restatement must still match the reference's own code bit for bit"""
    W = window_small
    syn = pkg.synthetic
    s = O.Settings(6.0, 30.0 * 30.0, 0.02, 3, 2, 0.8, 2, 0.2, 1.5, 2.5, 3.0, 50.0)
    try:
        O.set_settings(s)
        O.set_settings(s, kind="ref")
        # tracker: calcRes / calcGSSSE with another Huber threshold
        levels = syn.pyr_levels_for(W.w, W.h)
        ref_pyr = O.make_images(W.images[-1], levels)
        cu, cv, cid, hdi = syn.tracker_reference_inputs(W)
        a0, b0 = W.aff_true[-1]
        img, rel, aff = syn.tracking_sequence(W, 1)[0]
        pyr = O.make_images(img, levels)
        res = []
        for kind in ("oracle", "ref"):
            t = O.Tracker(W.w, W.h, levels, *W.K, kind=kind)
            t.set_ref(cu, cv, cid, hdi, flat(ref_pyr[0]), a0, b0, 1.0)
            t.set_new_frame(flat(pyr[0]), 1.0)
            r = t.calc_res(1, syn.pose12(rel), aff[0], aff[1], 20.0)
            H, b = t.calc_gs(1, aff[0], aff[1])
            res.append((r, H, b))
        for x, y in zip(res[0], res[1]):
            assert np.array_equal(bits(np.asarray(x)), bits(np.asarray(y)))
        # BA: linearize with other huber / gradient-weight thresholds
        p0 = [O.make_images(im, 1)[0][0] for im in W.images]
        lo, lr = O.BA(W, p0).linearize(), O.BA(W, p0, kind="ref").linearize()
        ok = lo["newState"] != 1
        assert np.array_equal(lo["newState"], lr["newState"]) and np.array_equal(bits(lo["J"][ok]), bits(lr["J"][ok]))
        # immature points: traceOn and optimizeImmaturePoint
        sc = syn.immature_scenario(W, host_idx=1, n_points=400, n_frames=2)
        hostI = O.make_images(sc["host_image"], 1)[0][0]
        fo = O.immature_init(hostI, sc["u"], sc["v"])
        so, sr = O.trace_state(len(sc["u"])), O.trace_state(len(sc["u"]))
        for fr_ in sc["frames"]:
            dI = O.make_images(fr_["image"], 1)[0][0]
            O.trace_on(dI, sc["u"], sc["v"], fo, so, fr_["KRKi"], fr_["Kt"], fr_["aff"])
            O.trace_on(dI, sc["u"], sc["v"], fo, sr, fr_["KRKi"], fr_["Kt"], fr_["aff"], kind="ref")
            assert np.array_equal(so["status"], sr["status"])
            for k in ("idepth_min", "idepth_max", "quality", "uv", "interval"):
                assert np.array_equal(bits(so[k]), bits(sr[k])), k
        ac = syn.activation_scenario(W, host_idx=1, n_points=400)
        fields = O.immature_init(O.make_images(ac["host_image"], 1)[0][0], ac["u"], ac["v"])
        dIs = [O.make_images(im, 1)[0][0] for im in ac["images"]]
        args = (W.K, dIs, ac["RTll"], ac["tTll"], ac["aff"], ac["u"], ac["v"], fields, ac["idepth_min"], ac["idepth_max"])
        ro, io, _ = O.optimize_immature(*args, min_obs=1)
        rr, ir, _ = O.optimize_immature(*args, min_obs=1, kind="ref")
        assert np.array_equal(ro, rr) and np.array_equal(bits(io[ro == 1]), bits(ir[rr == 1]))
    finally:
        O.set_settings()
        O.set_settings(kind="ref")
    # and the change was not a no-op: the default settings give other Jacobians
    ld = O.BA(W, p0).linearize()
    assert not np.array_equal(bits(ld["J"]), bits(lo["J"]))


def test_per_object_linearize_apply(pkg, win):
    """the serial per-object calls of FullSystem::flagPointsForRemoval (FullSystem.cpp:776-787): r->resetOOB();
    r->linearize(&Hcalib); r->applyRes(true) on the residuals of a few points, after a regular linearizeAll + applyRes"""
    W = win
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    Bo, Br = O.BA(W, pyr), O.BA(W, pyr, kind="ref")
    idx = np.nonzero(W.res_point % 7 == 3)[0].astype(np.int32)
    assert len(idx) > 20
    outs = []
    for B in (Bo, Br):
        B.linearize(); B.apply_res()
        outs.append(B.linearize_residuals(idx, reset_oob=True, apply=True))
        outs.append(B.accumulate_A(False))
    (po, Ao, pr, Ar) = outs[0], outs[1], outs[2], outs[3]
    assert np.array_equal(po["newState"], pr["newState"]) and np.array_equal(po["isActive"], pr["isActive"])
    assert np.array_equal(bits(po["newEnergy"]), bits(pr["newEnergy"]))
    assert (po["newState"] == 0).sum() > 0
    for key in ("Hdd", "bd", "Hcd"):
        assert np.array_equal(bits(Ao[key]), bits(Ar[key])), key      # the swapped-in Jacobians feed the next accumulation


@pytest.mark.parametrize("lvl,unit_depth", [(0, False), (1, False), (2, True), (3, False)])
def test_initializer_calc_res_and_gs(pkg, window_vga, lvl, unit_depth):
    """next row N4: CoarseInitializer::calcResAndGS (FullSystem/CoarseInitializer.cpp:370-732) and makeK (:1126-1155) -- the
    restatement against the reference's own translation unit: per-point outputs, H, b, H_sc, b_sc and the returned Vec3f bit for
    bit (Accumulator9 / Accumulator11 lanes and tiers included)"""
    W = window_vga
    sc = pkg.synthetic.initializer_scenario(W, lvl=lvl, n_points=2500, unit_depth=unit_depth)
    ko, kr = O.init_make_K(W.K, lvl), O.init_make_K(W.K, lvl, kind="ref")
    assert np.array_equal(bits(ko[0]), bits(kr[0])) and np.array_equal(bits(ko[1]), bits(kr[1]))
    first = O.make_images(sc["first"], lvl + 1)[0][lvl]
    new = O.make_images(sc["new"], lvl + 1)[0][lvl]
    ro = O.init_calc_res_and_gs(W.K, first, new, sc)
    rr = O.init_calc_res_and_gs(W.K, first, new, sc, kind="ref")
    assert np.array_equal(ro["isGood_new"], rr["isGood_new"])
    good = ro["isGood_new"] == 1
    n_in = int((sc["isGood"] == 1).sum())
    assert good.sum() > 0.5 * n_in and (~good & (sc["isGood"] == 1)).sum() > 5, (good.sum(), n_in)     # kept, and rejected by bounds / outlierTH
    assert np.array_equal(bits(ro["energy_new"]), bits(rr["energy_new"]))
    for k in ("maxstep", "lastHessian_new"):
        assert np.array_equal(bits(ro[k][good]), bits(rr[k][good])), k
    assert np.array_equal(bits(ro["Jb"][good]), bits(rr["Jb"][good]))
    for k in ("H", "b", "Hsc", "bsc", "res"):
        assert np.array_equal(bits(ro[k]), bits(rr[k])), (k, ro[k], rr[k])
    assert ro["res"][2] == 2 * len(sc["u"])          # E.num counts every point twice (:617-633 feed E, not EAlpha)
