"""K2 parity: calcRes / calcGSSSE / trackNewestCoarse on the GPU vs the CPU oracle.

Bars (BASELINE.json north_star): residual counts and inlier / out-of-bounds / cut-off masks bit-exact,
warped values bit-exact, H and b <= 1e-4 relative per entry (relative to the entry's scale
max(|H_ij|, sqrt(H_ii H_jj)); b_i relative to sqrt(H_ii * sum w r^2)), poses <= 1e-5 for matching
accept/reject traces.
"""
import numpy as np
import pytest
from helpers import TrackScenario, compact_warped, rel_err_matrix

pytestmark = pytest.mark.gpu

TOL_H = 1e-4
TOL_POSE = 1e-5


@pytest.fixture(scope="module")
def scen_vga(pkg, oracle, window_vga):
    return TrackScenario(pkg, oracle, window_vga, n_frames=3)


@pytest.fixture(scope="module")
def ctx_vga(pkg, scen_vga):
    ctx = pkg.Context(0, 640, 480, 6)
    scen_vga.gpu_setup(ctx)
    yield ctx
    ctx.close()


def _check_eval(ctx, scen, k, lvl, pose, a, b, cutoff):
    scen.oracle_set_frame(k)
    ro = scen.trk.calc_res(lvl, pose, a, b, cutoff)
    Ho, bo = scen.trk.calc_gs(lvl, a, b)
    n_o, bufs_o = scen.trk.get_warped()
    n_pts = len(scen.clouds[lvl][0])
    st_o = scen.trk.get_point_status(n_pts)

    rg = ctx.calc_res(0, 1 + k, lvl, pose, a, b, cutoff)
    Hg, bg = ctx.calc_gs(0)
    st_g, wp_g = ctx.calc_res_debug(0, 1 + k, lvl, n_pts, pose, a, b, cutoff)

    # masks and counts: bit-exact
    assert np.array_equal(st_o, st_g), "status mask differs at %s" % np.nonzero(st_o != st_g)[0][:5]
    assert rg[1] == ro[1] and rg[5] == ro[5]
    n_g, bufs_g = compact_warped(st_g, wp_g)
    assert n_g == n_o
    for kk in range(8):
        assert np.array_equal(bufs_o[kk].view(np.uint32), bufs_g[kk].view(np.uint32)), "warped buffer %d" % kk
    # energy: float sum in a different order
    assert abs(rg[0] - ro[0]) <= 2e-5 * max(abs(ro[0]), 1.0)
    assert abs(rg[2] - ro[2]) <= 1e-5 * max(abs(ro[2]), 1e-3) and abs(rg[4] - ro[4]) <= 1e-5 * max(abs(ro[4]), 1e-3)
    if n_o > 0:
        assert rel_err_matrix(Hg, Ho).max() <= TOL_H
        # b_i = sum w J_i r: scale sqrt(H_ii * sum w r^2)
        sc = np.array([10 * 10 if i == 6 else 1 for i in range(8)])
        wr2 = float(np.sum(bufs_o[6].astype(np.float64) * bufs_o[5].astype(np.float64) ** 2)) / n_o
        scale = np.sqrt(np.abs(np.diag(Ho)) * wr2) + 1e-30
        assert (np.abs(bg - bo) / scale).max() <= TOL_H
    return ro, rg


@pytest.mark.parametrize("lvl", [0, 1, 2, 3])
def test_calc_res_gs_parity(ctx_vga, scen_vga, lvl):
    for k in range(2):
        a, b = scen_vga.frames[k][2]
        for pose in scen_vga.poses(k):
            _check_eval(ctx_vga, scen_vga, k, lvl, pose, a, b, 20.0)


def test_calc_res_cutoffs_and_affine(ctx_vga, scen_vga):
    pose = scen_vga.poses(0)[0]
    for cutoff in (5.0, 40.0, 160.0):
        _check_eval(ctx_vga, scen_vga, 0, 1, pose, 0.3, -25.0, cutoff)


def test_calc_res_all_out_of_bounds(ctx_vga, scen_vga):
    """a pose that throws every point out of the image: zero terms, NaN ratio like the reference"""
    syn = scen_vga.syn
    far = syn.pose12(syn.se3_exp(np.array([50.0, 0, 0, 0, 0, 0])))
    scen_vga.oracle_set_frame(0)
    ro = scen_vga.trk.calc_res(0, far, 0.0, 0.0, 20.0)
    rg = ctx_vga.calc_res(0, 1, 0, far, 0.0, 0.0, 20.0)
    assert ro[1] == 0 and rg[1] == 0 and np.isnan(ro[5]) and np.isnan(rg[5])


def _pose_err(syn, p1, p2):
    return np.abs(syn.se3_log(syn.se3_mul(syn.from12(np.array(p1)), syn.se3_inv(syn.from12(np.array(p2)))))).max()


TIE = 1e-6      # |E'n - En'| / (En') below which an accept test (CoarseTracker.cpp:669) counts as a float-rounding tie


def _compare_track(syn, ro, rg, levels):
    """GPU result vs oracle result of one trackNewestCoarse.  Same accept/reject trace: pose <= 1e-5, affine, residuals, pass
    structure must match.  Different trace: legitimate ONLY if the first differing decision was a tie on BOTH sides (the
    energies are float sums taken in different orders); anything else is a bug and fails.  Returns True for a full match."""
    assert rg.ok == ro.ok
    tr_o, tr_g = list(ro.trace[:ro.trace_len]), list(rg.trace[:rg.trace_len])
    if tr_o == tr_g:
        assert _pose_err(syn, rg.pose, ro.pose) <= TOL_POSE
        assert abs(rg.aff_a - ro.aff_a) <= 1e-5 and abs(rg.aff_b - ro.aff_b) <= 1e-3
        assert rg.n_evals == ro.n_evals and rg.n_passes == ro.n_passes
        for l in range(levels):
            assert abs(rg.lastResiduals[l] - ro.lastResiduals[l]) <= 1e-4 * ro.lastResiduals[l]
        return True
    d = next((i for i, (x, y) in enumerate(zip(tr_o, tr_g)) if x != y), None)
    assert d is not None, "traces differ in length only (%d vs %d): not an accept/reject tie" % (len(tr_o), len(tr_g))
    assert d < 64, "first differing decision %d lies beyond the recorded margins" % d
    m_o = abs(ro.trace_dE[d] / ro.trace_E[d])
    m_g = abs(rg.trace_dE[d] / rg.trace_E[d])
    assert m_o < TIE and m_g < TIE, "decision %d flipped without a tie: |dE|/E oracle %.3e, gpu %.3e" % (d, m_o, m_g)
    # up to the flip both sides evaluated the same states: the recorded energies agree to float-sum accuracy
    assert abs(ro.trace_E[d] - rg.trace_E[d]) <= 1e-5 * abs(ro.trace_E[d])
    assert _pose_err(syn, rg.pose, ro.pose) < 2e-3
    return False


def _track_sequence(ctx, scen, mode, n_frames, frame_slot0=1):
    syn = scen.syn
    prev = syn.pose12((np.eye(3), np.zeros(3)))
    a, b = scen.a0, scen.b0
    n_match = 0
    for k in range(n_frames):
        scen.oracle_set_frame(k)
        ro = scen.trk.track(prev, a, b, scen.levels - 1)
        if mode == "hostloop":
            rg = ctx.track_hostloop(0, frame_slot0 + k, prev, a, b, scen.levels - 1)
        else:
            rg = ctx.track_batched(0, frame_slot0 + k, prev, a, b, scen.levels - 1)[0]
        assert rg.ok == ro.ok == 1
        # both must recover the generating motion
        assert _pose_err(syn, rg.pose, syn.pose12(scen.frames[k][1])) < 5e-3
        n_match += _compare_track(syn, ro, rg, scen.levels)
        prev = np.array(ro.pose)
        a, b = ro.aff_a, ro.aff_b
    return n_match


@pytest.mark.parametrize("mode", ["hostloop", "persistent"])
def test_track_newest_coarse_parity(ctx_vga, scen_vga, mode):
    n_match = _track_sequence(ctx_vga, scen_vga, mode, 3)
    assert n_match >= 2, "accept/reject traces diverged on most frames"


@pytest.mark.parametrize("mode", ["hostloop", "persistent"])
def test_track_forced_five_levels(pkg, oracle, window_vga, mode):
    """BASELINE configs[0] as worded ("5-level pyramid"): the reference picks 4 levels at 640x480; forcing a fifth
    (40x30) runs the coarsest-level path with ~900 points and maxIterations[4] = 50"""
    scen = TrackScenario(pkg, oracle, window_vga, levels=5, n_frames=2)
    assert len(scen.clouds) == 5 and 300 < len(scen.clouds[4][0]) < 1500
    ctx = pkg.Context(0, 640, 480, 4)
    scen.gpu_setup(ctx)
    try:
        for lvl in (4, 3):
            a, b = scen.frames[0][2]
            _check_eval(ctx, scen, 0, lvl, scen.poses(0)[1], a, b, 20.0)
        assert _track_sequence(ctx, scen, mode, 2) >= 1
    finally:
        ctx.close()


@pytest.mark.parametrize("nc", [2, 4])
def test_track_multi_cluster_forced(ctx_vga, scen_vga, nc):
    """the multi-cluster path of track_persistent_kernel (one hypothesis spread over nc clusters, cluster totals through
    global memory and an arrival counter), forced on the VGA cloud"""
    ctx_vga.track_set_tuning(nclusters=nc)
    try:
        n_match = _track_sequence(ctx_vga, scen_vga, "persistent", 3)
        assert ctx_vga.track_last_shape()[2] == nc
        assert n_match >= 2
    finally:
        ctx_vga.track_set_tuning()
    ident = scen_vga.syn.pose12((np.eye(3), np.zeros(3)))
    ctx_vga.track_batched(0, 1, ident, scen_vga.a0, scen_vga.b0, scen_vga.levels - 1)
    assert ctx_vga.track_last_shape()[2] == 1


def test_track_large_cloud_multi_cluster(pkg, oracle, window_large):
    """BASELINE configs[3], 1280x1024 with ~18 k points: the level-0 cloud (> 32768 points) selects the multi-cluster path by
    itself.  calcRes masks / warped values bit-exact on the two finest levels, trackNewestCoarse vs the oracle."""
    scen = TrackScenario(pkg, oracle, window_large, levels=5, n_frames=2)
    assert len(scen.clouds[0][0]) >= 32768
    ctx = pkg.Context(0, 1280, 1024, 4)
    scen.gpu_setup(ctx)
    try:
        a, b = scen.frames[0][2]
        for lvl in (0, 1):
            _check_eval(ctx, scen, 0, lvl, scen.poses(0)[1], a, b, 20.0)
        n_match = _track_sequence(ctx, scen, "persistent", 2)
        assert ctx.track_last_shape()[2] > 1, "multi-cluster path not taken: %s" % (ctx.track_last_shape(),)
        assert n_match >= 1
    finally:
        ctx.close()


def test_track_batched_31_hypotheses(ctx_vga, scen_vga):
    """FullSystem.cpp:311-354: all 31 pose priors in one launch; each must equal its single run."""
    syn = scen_vga.syn
    rel0 = scen_vga.frames[0][1]
    # slast = frame 0, sprelast = the key frame itself: lastF_2_slast = rel0, fh_2_slast := slast_2_sprelast = rel0^-1
    hyps = syn.motion_hypotheses(syn.se3_inv(rel0), rel0)
    assert len(hyps) == 31
    P = np.stack([syn.pose12(h) for h in hyps])
    res = ctx_vga.track_batched(0, 2, P, scen_vga.a0, scen_vga.b0, scen_vga.levels - 1)
    single = ctx_vga.track_batched(0, 2, P[7], scen_vga.a0, scen_vga.b0, scen_vga.levels - 1)[0]
    # a different cluster size changes the summation tree, nothing else
    if list(res[7].trace[:res[7].trace_len]) == list(single.trace[:single.trace_len]):
        assert np.abs(np.array(res[7].pose) - np.array(single.pose)).max() < 1e-6 and res[7].n_evals == single.n_evals   # FP32 sums, two trees
    scen_vga.oracle_set_frame(1)
    good = 0
    for i in (0, 3, 12):
        ro = scen_vga.trk.track(P[i], scen_vga.a0, scen_vga.b0, scen_vga.levels - 1)
        if list(ro.trace[:ro.trace_len]) == list(res[i].trace[:res[i].trace_len]):
            assert _pose_err(syn, res[i].pose, ro.pose) <= TOL_POSE
            good += 1
    assert good >= 2
    # sequential semantics of FullSystem::trackNewCoarse replayed from the per-pass RMSEs
    achieved = np.full(5, np.nan)
    ok0, lr0 = ctx_vga.replay_abort(res[0], achieved)
    assert ok0 == 1
    tight = np.array(lr0) * 0.5
    ok1, lr1 = ctx_vga.replay_abort(res[4], np.where(np.isnan(tight), np.nan, tight))
    assert ok1 in (0, 1)


def test_track_replay_abort_matches_oracle(ctx_vga, scen_vga):
    """abort thresholds (CoarseTracker.cpp:713): host loop with minResForAbort == oracle with the same"""
    syn = scen_vga.syn
    ident = syn.pose12((np.eye(3), np.zeros(3)))
    scen_vga.oracle_set_frame(2)
    free = scen_vga.trk.track(ident, scen_vga.a0, scen_vga.b0, scen_vga.levels - 1)
    thr = np.array(free.lastResiduals) * 0.2     # impossible to meet -> abort at the coarsest level
    ro = scen_vga.trk.track(ident, scen_vga.a0, scen_vga.b0, scen_vga.levels - 1, thr)
    rg = ctx_vga.track_hostloop(0, 3, ident, scen_vga.a0, scen_vga.b0, scen_vga.levels - 1, thr)
    assert ro.ok == 0 and rg.ok == 0 and ro.aborted_level == rg.reserved == scen_vga.levels - 1
    full = ctx_vga.track_batched(0, 3, ident, scen_vga.a0, scen_vga.b0, scen_vga.levels - 1)[0]
    ok, lr = ctx_vga.replay_abort(full, thr)
    assert ok == 0 and np.isnan(lr[0])


@pytest.mark.parametrize("modes", [(-1.0, -1.0), (1e12, -1.0), (-1.0, 1e8)])
def test_track_fixed_affine_modes(ctx_vga, scen_vga, oracle, modes):
    """setting_affineOptModeA/B < 0: 6x6 / 7x7 solves (CoarseTracker.cpp:621-643) and the zeroed outputs (:743-744)"""
    syn = scen_vga.syn
    ident = syn.pose12((np.eye(3), np.zeros(3)))
    a, b = scen_vga.a0, scen_vga.b0
    opts = ctx_vga.default_opts()
    opts.affineOptModeA, opts.affineOptModeB = modes
    scen_vga.oracle_set_frame(0)
    try:
        oracle.set_affine_modes(*modes)
        ro = scen_vga.trk.track(ident, a, b, scen_vga.levels - 1)
    finally:
        oracle.set_affine_modes()
    for rg in (ctx_vga.track_batched(0, 1, ident, a, b, scen_vga.levels - 1, opts=opts)[0],
               ctx_vga.track_hostloop(0, 1, ident, a, b, scen_vga.levels - 1, opts=opts)):
        assert rg.ok == ro.ok
        if modes[0] < 0:
            assert rg.aff_a == 0 == ro.aff_a
        if modes[1] < 0:
            assert rg.aff_b == 0 == ro.aff_b
        _compare_track(syn, ro, rg, scen_vga.levels)


@pytest.mark.gpu
def test_track_and_prefetch_equals_the_three_calls(ctx_vga, scen_vga):
    """dso_b200_track_and_prefetch = track_batched_async + make_images_async(next image) + track_fetch: same tracking
    result, and the prefetched slot tracks exactly like the slot built by dso_b200_make_images from the same image."""
    syn = scen_vga.syn
    pose = np.ascontiguousarray(syn.pose12(scen_vga.frames[0][1]), dtype=np.float64).reshape(-1)
    lv = scen_vga.levels
    ref = ctx_vga.track_batched(0, 2, pose, scen_vga.a0, scen_vga.b0, lv - 1)[0]
    nxt = np.ascontiguousarray(scen_vga.frames[2][0], dtype=np.float32)
    got = ctx_vga.track_and_prefetch(0, 2, pose, scen_vga.a0, scen_vga.b0, lv - 1, 5, nxt.ctypes.data, 640, 480, lv)
    assert got.ok == ref.ok and got.n_evals == ref.n_evals and list(got.pose) == list(ref.pose)
    assert list(got.trace[:got.trace_len]) == list(ref.trace[:ref.trace_len])
    a = ctx_vga.track_batched(0, 3, pose, scen_vga.a0, scen_vga.b0, lv - 1)[0]       # slot 3: frame 2 through make_images
    b = ctx_vga.track_batched(0, 5, pose, scen_vga.a0, scen_vga.b0, lv - 1)[0]       # slot 5: the same image, prefetched
    assert list(a.pose) == list(b.pose) and a.n_evals == b.n_evals
    # no next image: plain tracking
    got2 = ctx_vga.track_and_prefetch(0, 2, pose, scen_vga.a0, scen_vga.b0, lv - 1, -1, None, 640, 480, lv)
    assert list(got2.pose) == list(ref.pose)
