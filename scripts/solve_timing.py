"""Phase cycles of ba_solve_kernel, printed by the library on stderr (DSO_B200_SOLVE_TIMING=1 must be set before the
context's solver state is created; this script sets it itself).  Run on the GPU box."""
import os
import sys
os.environ["DSO_B200_SOLVE_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as entry

pkg = entry.load_package()
W = pkg.synthetic.make_window(seed=0, n_points=2000)
ctx = pkg.Context(0, W.w, W.h, W.nF + 1)
for i, img in enumerate(W.images):
    ctx.make_images(i, img, 1, want_host=False)
ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
ctx.ba_set_frames(W)
ctx.ba_set_points(W)
ctx.ba_set_residuals(W)
fP = np.zeros((W.nF, 8))
fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]
ctx.ba_set_priors(np.full(4, 5e9), fP.reshape(-1))
ctx.ba_linearize()
ctx.ba_apply_res()
for it in range(4):
    ctx.ba_solve(0, 1e-5)          # dso_b200_ba_solve prints the stamps (the graphed dso_b200_ba_iteration does not)
ctx.close()
