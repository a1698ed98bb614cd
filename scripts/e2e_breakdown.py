"""Wall-clock split of the per-frame path (run on the GPU box)."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench, __graft_entry__ as entry
pkg = entry.load_package()
wl = bench.build_workload(0, 6, 2000)
W, levels = wl["W"], wl["levels"]
ctx = pkg.Context(0, W.w, W.h, 10)
ctx.tracker_set_ref(0, W.w, W.h, levels, W.K, wl["clouds"], wl["a0"], wl["b0"], 1.0)
L = pkg.capi.load()
npx = W.w * W.h
pin = []
for k in range(6):
    p = L.dso_b200_host_alloc(npx * 4)
    np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_float)), shape=(npx,))[:] = wl["frames"][k][0].reshape(-1)
    pin.append(p)
    ctx.make_images_ptr(1 + k, p, W.w, W.h, levels)


def t(fn, n=300):
    for _ in range(20):
        fn(0)
    t0 = time.perf_counter()
    for i in range(n):
        fn(i % 6)
    return (time.perf_counter() - t0) / n * 1e6


a, b = wl["affs"][2]
print("sync only                 %.1f us" % t(lambda k: ctx.sync()))
print("H2D + K1 + sync           %.1f us" % t(lambda k: ctx.make_images_ptr(7, pin[k], W.w, W.h, levels)))
print("track_batched (resident)  %.1f us" % t(lambda k: ctx.track_batched(0, 1 + k, wl["inits"][k], *wl["affs"][k], levels - 1)))
print("track_frame (e2e)         %.1f us" % t(lambda k: ctx.track_frame(0, 7, pin[k], W.w, W.h, levels, wl["inits"][k], *wl["affs"][k], levels - 1)))
ctx.phase_cycles(True)
ctx.track_batched(0, 1, wl["inits"][0], *wl["affs"][0], levels - 1)
c = ctx.phase_cycles(True, read=True)
print("kernel ns (in-kernel globaltimer)", c[6])
