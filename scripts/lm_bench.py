"""Cycles per LM step of the persistent tracker in isolation (one warp, warm instruction cache); run on the GPU box."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry

pkg = entry.load_package()
ctx = pkg.Context(0, 640, 480, 2)
for n in (100, 2000):
    c = ctx.debug_lm_bench(n)
    print("iters %d: %.0f cycles/step | level %.0f solve %.0f exp+pose %.0f params %.0f sums->H %.0f control %.0f" % ((n,) + tuple(c)))
ctx.close()
