"""Per-phase cycle breakdown of the persistent tracker (debug aid; run on the GPU box)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import __graft_entry__ as entry

pkg = entry.load_package()
wl = bench.build_workload(0, 6, 2000)
W, levels = wl["W"], wl["levels"]
ctx = pkg.Context(0, W.w, W.h, 10)
ctx.tracker_set_ref(0, W.w, W.h, levels, W.K, wl["clouds"], wl["a0"], wl["b0"], 1.0)
for k in range(6):
    ctx.make_images(1 + k, wl["frames"][k][0], levels, want_host=False)
ctx.phase_cycles(True)
import torch
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda:0")
for k in range(6):
    a, b = wl["affs"][k]
    if "--no-flush" not in sys.argv:
        flush.zero_(); torch.cuda.synchronize()
    r = ctx.track_batched(0, 1 + k, wl["inits"][k], a, b, levels - 1)[0]
    c = ctx.phase_cycles(True, read=True)
    n = max(c[4], 1)
    print("frame", k, "evals", r.n_evals, "cycles/eval: eval %.0f reduce %.0f xchg %.0f lm %.0f total %.0f" %
          (c[0] / n, c[1] / n, c[2] / n, c[3] / n, c[:4].sum() / n), "| kernel cycles", c[5], "ns", c[6], "| lm split: take/store+level %.0f solve %.0f exp+pose %.0f params %.0f | sums->H %.0f control %.0f" % tuple(c[8:14] / n),
          "| outside the evaluations: prologue %d staging %d epilogue %d other %d" % (c[7], c[14], c[15], c[5] - c[:4].sum() - c[7] - c[14] - c[15]),
          "| prologue: LM set-up %d, +staging/syncthreads %d, +cluster.sync %d | epilogue: finalize %d, +result stores %d, +system fence %d, end %d" %
          (c[16], c[17], c[7], c[18], c[19], c[20], c[15]))
    if "--per-eval" in sys.argv:
        for e in range(min(int(n), 24)):
            print("    eval %2d: eval %5d reduce %5d xchg %5d lm %5d" % ((e,) + tuple(c[24 + 4 * e:28 + 4 * e])))
hy = np.stack([wl["inits"][2]] * 31)
ctx.track_batched(0, 3, hy, *wl["affs"][2], levels - 1)
c = ctx.phase_cycles(True, read=True)
n = max(c[4], 1)
print("31 hyps: cycles/eval: eval %.0f reduce %.0f xchg %.0f lm %.0f total %.0f" % (c[0] / n, c[1] / n, c[2] / n, c[3] / n, c[:4].sum() / n),
      "| kernel ns", c[6])
