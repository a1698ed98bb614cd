"""Next row N3 (first half): CoarseTracker::makeCoarseDepthL0 on the device vs the oracle's (which is bit-identical to the
reference build, test_oracle_vs_reference.py::test_make_coarse_depth_and_K_bit_exact): the per-level clouds
pc_u / pc_v / pc_idepth / pc_color and pc_n must match bit for bit, in the reference's raster order."""
import numpy as np
import pytest
from helpers import TrackScenario, flat

pytestmark = pytest.mark.gpu


def _collisions(W, cu, cv):
    key = (cv + 0.5).astype(np.int64) * W.w + (cu + 0.5).astype(np.int64)
    _, cnt = np.unique(key, return_counts=True)
    return cnt.max()


@pytest.mark.parametrize("which", ["small", "vga"])
def test_coarse_depth_cloud_bit_exact(pkg, oracle, window_small, window_vga, which):
    W = window_small if which == "small" else window_vga
    scen = TrackScenario(pkg, oracle, W, n_frames=1)
    cu, cv, cid, hdi = pkg.synthetic.tracker_reference_inputs(W)
    assert _collisions(W, cu, cv) >= (3 if which == "small" else 1)   # pixels hit by 3+ points take the ordered path
    ctx = pkg.Context(0, W.w, W.h, 4)
    ctx.make_images(0, W.images[-1], scen.levels, want_host=False)
    pc_n = ctx.tracker_set_ref_points(0, 0, scen.levels, W.K, cu, cv, cid, hdi, scen.a0, scen.b0, 1.0)
    for l in range(scen.levels):
        u, v, idp, col = scen.clouds[l]
        assert pc_n[l] == len(u) > 0
        g = ctx.tracker_get_cloud(0, l)
        ref = np.stack([u, v, idp, col], axis=1).astype(np.float32)
        assert np.array_equal(g.view(np.uint32), ref.view(np.uint32)), l
    # and the tracker runs on it exactly as on the uploaded cloud
    ctx.make_images(1, scen.frames[0][0], scen.levels, want_host=False)
    ident = pkg.synthetic.pose12((np.eye(3), np.zeros(3)))
    r1 = ctx.track_batched(0, 1, ident, scen.a0, scen.b0, scen.levels - 1)[0]
    ctx.tracker_set_ref(1, W.w, W.h, scen.levels, W.K, scen.clouds, scen.a0, scen.b0, 1.0)
    r2 = ctx.track_batched(1, 1, ident, scen.a0, scen.b0, scen.levels - 1)[0]
    assert r1.ok == r2.ok == 1 and list(r1.trace[:r1.trace_len]) == list(r2.trace[:r2.trace_len])
    assert np.array_equal(np.array(r1.pose), np.array(r2.pose))
    ctx.close()


def test_coarse_depth_edge_cases(pkg, oracle, window_small):
    W = window_small
    levels = pkg.synthetic.pyr_levels_for(W.w, W.h)
    ctx = pkg.Context(0, W.w, W.h, 2)
    ctx.make_images(0, W.images[-1], levels, want_host=False)
    # no points at all: empty clouds on every level
    z = np.zeros(0, dtype=np.float32)
    pc_n = ctx.tracker_set_ref_points(0, 0, levels, W.K, z, z, z, z, 0.0, 0.0)
    assert not pc_n.any()
    # points outside the image and on the border are ignored / never compacted (x, y in [2, w-2))
    cu = np.array([-5.0, W.w + 3.0, 0.0, W.w - 1.0, 10.0], dtype=np.float32)
    cv = np.array([3.0, 3.0, 0.0, W.h - 1.0, 12.0], dtype=np.float32)
    cid = np.full(5, 0.5, dtype=np.float32)
    hdi = np.full(5, 1e-3, dtype=np.float32)
    pc_n = ctx.tracker_set_ref_points(0, 0, levels, W.K, cu, cv, cid, hdi, 0.0, 0.0)
    g = ctx.tracker_get_cloud(0, 0)
    # level 0: the point at (10, 12) and its four diagonal neighbours from the dilation pass
    assert pc_n[0] == 5 and sorted(map(tuple, g[:, :2].astype(int))) == [(9, 11), (9, 13), (10, 12), (11, 11), (11, 13)]
    assert np.allclose(g[:, 2], 0.5)
    ctx.close()
