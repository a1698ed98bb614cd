"""Driver for the ncu capture of the bandwidth-regime launches: makeImages at 3840x2160 / 5 levels and one trackNewestCoarse on the
1280x1024 / ~18 k-point window, whose level-0 cloud (> 32768 points) takes the multi-cluster path (run on the GPU box)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
import __graft_entry__ as entry

pkg = entry.load_package()
syn = pkg.synthetic
w, h, levels = 3840, 2160, 5
ctx = pkg.Context(0, w, h, 1)
img = torch.rand(h, w, device="cuda") * 255
for _ in range(3):
    ctx.make_images_dev(0, img.data_ptr(), w, h, levels)
ctx.sync()
ctx.close()
del img

W = syn.make_window(seed=0, w=1280, h=1024, nF=4, n_points=18000, kf_step_t=0.04, kf_step_deg=1.0)
levels = 5
cu, cv, cid, hdi = syn.tracker_reference_inputs(W)
a0, b0 = W.aff_true[-1]
frames = syn.tracking_sequence(W, 2)
ctx = pkg.Context(0, W.w, W.h, 4)
ctx.make_images(0, W.images[-1], levels, want_host=False)
pc_n = ctx.tracker_set_ref_points(0, 0, levels, W.K, cu, cv, cid, hdi, a0, b0, 1.0)
ident = syn.pose12((np.eye(3), np.zeros(3)))
for k, f in enumerate(frames):
    ctx.make_images(1 + k, f[0], levels, want_host=False)
    r = ctx.track_batched(0, 1 + k, ident, a0, b0, levels - 1)[0]
    assert r.ok == 1
print("profile_large ok: pc_n", [int(x) for x in pc_n], "shape", ctx.track_last_shape())
ctx.close()
