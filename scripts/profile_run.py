"""Short driver for ncu: a few frames of makeImages + trackNewestCoarse, a few BA steps and a few full GN iterations (run on the GPU box)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import __graft_entry__ as entry

pkg = entry.load_package()
wl = bench.build_workload(0, 4, 2000)
W, levels = wl["W"], wl["levels"]
ctx = pkg.Context(0, W.w, W.h, W.nF + 6)
ctx.tracker_set_ref(0, W.w, W.h, levels, W.K, wl["clouds"], wl["a0"], wl["b0"], 1.0)
for rep in range(3):
    for k in range(4):
        a, b = wl["affs"][k]
        r = ctx.track_frame(0, W.nF + 1 + k, np.ascontiguousarray(wl["frames"][k][0]), W.w, W.h, levels, wl["inits"][k], a, b, levels - 1)[0]
        assert r.ok == 1
for i, img in enumerate(W.images):
    ctx.make_images(i, img, 1, want_host=False)
ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
ctx.ba_set_frames(W)
ctx.ba_set_points(W)
ctx.ba_set_residuals(W)
for rep in range(4):
    st = ctx.ba_step()
fP = np.zeros((W.nF, 8))
fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]
ctx.ba_set_priors(np.full(4, 5e9), fP.reshape(-1))
for rep in range(4):
    it = ctx.ba_iteration(rep, 1e-5, stepfacD=0.0)
print("profile_run ok: launches", ctx.launch_count(), "resInA", st["resInA"], "energy", it["energy"])
ctx.close()
