"""CPU self-checks of the tracker oracle (no external truth exists for this path):
convergence to the generating motion, H symmetric PSD, b vs a finite difference of the energy,
LDLT / SE3 helpers vs numpy."""
import numpy as np
import pytest
from helpers import TrackScenario


@pytest.fixture(scope="module")
def scen(pkg, oracle, window_small):
    return TrackScenario(pkg, oracle, window_small, n_frames=2)


def test_cloud_shapes(scen):
    n = [len(c[0]) for c in scen.clouds]
    assert n[0] > n[-1] > 50
    u, v, idp, col = scen.clouds[0]
    assert (u >= 2).all() and (u < scen.W.w - 2).all() and (idp > 0).all() and np.isfinite(col).all()


def test_tracks_generating_motion(scen):
    syn = scen.syn
    scen.oracle_set_frame(0)
    r = scen.trk.track(syn.pose12((np.eye(3), np.zeros(3))), scen.a0, scen.b0, scen.levels - 1)
    assert r.ok == 1
    est = syn.from12(np.array(r.pose))
    err = syn.se3_log(syn.se3_mul(est, syn.se3_inv(scen.frames[0][1])))
    assert np.abs(err).max() < 5e-3
    assert r.lastResiduals[0] < 6.0
    assert r.n_passes in (scen.levels, scen.levels + 1) and r.trace_len == r.n_iterations   # +1: one-shot level repeat (:717-722)


def test_H_symmetric_psd_and_b_is_gradient(scen):
    syn = scen.syn
    scen.oracle_set_frame(0)
    lvl = 1
    pose = syn.pose12(scen.frames[0][1])
    a, b = scen.frames[0][2]
    r0 = scen.trk.calc_res(lvl, pose, a, b, 1e6)
    H, g = scen.trk.calc_gs(lvl, a, b)
    assert np.allclose(H, H.T)
    assert np.linalg.eigvalsh(H).min() > -1e-6 * np.abs(H).max()
    # finite difference of E/n wrt the (unscaled) affine offset b: dE/db = 2 * sum(w r dr/db)/n, dr/db = -1 (J7)
    eps = 0.05
    rp = scen.trk.calc_res(lvl, pose, a, b + eps, 1e6)
    rm = scen.trk.calc_res(lvl, pose, a, b - eps, 1e6)
    n, _ = scen.trk.get_warped()
    fd = (rp[0] - rm[0]) / (2 * eps) / n
    assert rp[1] == rm[1] == r0[1]
    assert abs(fd - 2 * g[7] / 1000.0) < 0.05 * abs(fd) + 1e-3     # SCALE_B = 1000 folded into g[7]


def test_abort_threshold(scen):
    syn = scen.syn
    scen.oracle_set_frame(1)
    ident = syn.pose12((np.eye(3), np.zeros(3)))
    free = scen.trk.track(ident, scen.a0, scen.b0, scen.levels - 1)
    thr = np.array(free.lastResiduals) * 0.1
    r = scen.trk.track(ident, scen.a0, scen.b0, scen.levels - 1, thr)
    assert r.ok == 0 and r.aborted_level == scen.levels - 1 and list(r.pose) == list(ident)
    assert np.isnan(r.lastResiduals[0])


def test_padding_to_multiple_of_four(scen):
    scen.oracle_set_frame(0)
    scen.trk.calc_res(0, scen.syn.pose12(scen.frames[0][1]), 0.0, 0.0, 20.0)
    n, bufs = scen.trk.get_warped()
    assert n % 4 == 0
    st = scen.trk.get_point_status(len(scen.clouds[0][0]))
    kept = int((st == 2).sum())
    assert n == (kept + 3) // 4 * 4
    assert not bufs[6][kept:].any()
