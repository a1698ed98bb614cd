"""Driver for a per-kernel launch list of the BA step on a larger window (1920x1080, ~40 k points, 4 key frames); run on the GPU box."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as entry

os.environ.setdefault("DSO_B200_NO_GRAPH", "1")
pkg = entry.load_package()
syn = pkg.synthetic
W = syn.make_window(seed=0, w=1920, h=1080, nF=4, n_points=40000, kf_step_t=0.04, kf_step_deg=1.0)
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
for _ in range(3):
    st = ctx.ba_step()
it = ctx.ba_iteration(0, 1e-5, stepfacD=0.0)
print("profile_ba_large ok: residuals", W.nR, "resInA", st["resInA"])
ctx.close()
