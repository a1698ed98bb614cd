"""Minimal driver for one `ncu --set full` capture: 2 frames of makeImages + trackNewestCoarse, makeCoarseDepthL0, one BA
step and one full GN iteration (every kernel of the library appears once or twice; run on the GPU box)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import __graft_entry__ as entry

pkg = entry.load_package()
wl = bench.build_workload(0, 2, 2000, with_oracle=False)
W, levels = wl["W"], wl["levels"]
ctx = pkg.Context(0, W.w, W.h, W.nF + 4)
ctx.make_images(W.nF, W.images[-1], levels, want_host=False)
pc_n = ctx.tracker_set_ref_points(0, W.nF, levels, W.K, *wl["ref_inputs"], wl["a0"], wl["b0"], 1.0)
for k in range(2):
    a, b = wl["affs"][k]
    r = ctx.track_frame(0, W.nF + 1 + k, np.ascontiguousarray(wl["frames"][k][0]), W.w, W.h, levels, wl["inits"][k], a, b, levels - 1)[0]
    assert r.ok == 1
for i, img in enumerate(W.images):
    ctx.make_images(i, img, 1, want_host=False)
ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
ctx.ba_set_frames(W)
ctx.ba_set_points(W)
ctx.ba_set_residuals(W)
os.environ.setdefault("DSO_B200_NO_GRAPH", "1")
st = ctx.ba_step()
fP = np.zeros((W.nF, 8))
fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]
ctx.ba_set_priors(np.full(4, 5e9), fP.reshape(-1))
it = ctx.ba_iteration(0, 1e-5, stepfacD=0.0)
print("profile_small ok: pc_n", list(pc_n), "launches", ctx.launch_count(), "resInA", st["resInA"], "energy", it["energy"])
ctx.close()
