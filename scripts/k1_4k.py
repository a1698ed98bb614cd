"""makeImages at 3840x2160, 5 levels, input resident (for ncu launch lists of the two pyramid kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as entry
pkg = entry.load_package()
w, h, levels = 3840, 2160, 5
ctx = pkg.Context(0, w, h, 2)
img = torch.rand(h, w, device="cuda") * 255
stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
for _ in range(3):
    ctx.make_images_dev(0, img.data_ptr(), w, h, levels)
ctx.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for _ in range(20):
    ctx.make_images_dev(0, img.data_ptr(), w, h, levels)
e1.record(stream); e1.synchronize()
px = sum((w >> l) * (h >> l) for l in range(levels))
ms = e0.elapsed_time(e1) / 20
print("K1 4K: %.4f ms, %.0f GB/s" % (ms, (4 * w * h + 16 * px) / ms / 1e6))
ctx.close()
# the same at 640x480, 4 levels (launch-bound regime)
w, h, levels = 640, 480, 4
ctx = pkg.Context(0, w, h, 2)
img = torch.rand(h, w, device="cuda") * 255
stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
for _ in range(3):
    ctx.make_images_dev(0, img.data_ptr(), w, h, levels)
ctx.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for _ in range(50):
    ctx.make_images_dev(0, img.data_ptr(), w, h, levels)
e1.record(stream); e1.synchronize()
print("K1 VGA: %.4f ms per frame (back to back)" % (e0.elapsed_time(e1) / 50))
ctx.close()
