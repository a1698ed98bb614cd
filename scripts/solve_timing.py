import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import bench, __graft_entry__ as entry
pkg = entry.load_package()
wl = bench.build_workload(0, 4, 2000)
W = wl["W"]
ctx = pkg.Context(0, W.w, W.h, W.nF + 2)
for i, img in enumerate(W.images):
    ctx.make_images(i, img, 1, want_host=False)
ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
ctx.ba_set_frames(W); ctx.ba_set_points(W); ctx.ba_set_residuals(W)
fP = np.zeros((W.nF, 8)); fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]
ctx.ba_set_priors(np.full(4, 5e9), fP.reshape(-1))
ctx.ba_set_nullspaces(np.random.default_rng(0).normal(0, 1, (7, 8 * W.nF + 4)))
ctx.ba_linearize(want_J=False); ctx.ba_apply_res()
for it in (0, 0, 3, 3):
    ctx.ba_solve(it, 1e-5)
ctx.close()
