"""The reference tracks and maps on two threads at once: FullSystem::addActiveFrame runs under trackMutex
(FullSystem.cpp:838), FullSystem::makeKeyFrame / optimize under mapMutex (:1098), and the two CoarseTracker objects are
swapped between them (:898-902, :1180-1182).  One context must therefore serve a tracking thread and a mapping thread
concurrently: tracker slot 0 on its own stream, the windowed BA on the mapping stream.  Results must be those of the serial
run, bit for bit, and a tracking call must not queue behind the BA iterations of the other thread."""
import threading
import time
import numpy as np
import pytest
from helpers import TrackScenario

pytestmark = pytest.mark.gpu

N_CALLS = 50


def test_tracking_and_mapping_threads(pkg, oracle, window_vga):
    W = window_vga
    syn = pkg.synthetic
    scen = TrackScenario(pkg, oracle, W, n_frames=3)
    ctx = pkg.Context(0, W.w, W.h, W.nF + 5)
    # mapping side: the window's key frames in slots 0..nF-1; tracking side: new frames in slots nF+1...
    for i, img in enumerate(W.images):
        ctx.make_images(i, img, 1, want_host=False)
    ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
    ctx.ba_set_frames(W)
    ctx.ba_set_points(W)
    ctx.ba_set_residuals(W)
    fP = np.zeros((W.nF, 8))
    fP[0] = [1e10] * 3 + [1e11] * 3 + [1e14, 1e14]
    ctx.ba_set_priors(np.full(4, 5e9), fP.reshape(-1))
    ctx.tracker_set_ref(0, W.w, W.h, scen.levels, W.K, scen.clouds, scen.a0, scen.b0, 1.0)
    for k, f in enumerate(scen.frames):
        ctx.make_images(W.nF + 1 + k, f[0], scen.levels, want_host=False)
    ident = syn.pose12((np.eye(3), np.zeros(3)))

    def track(k):
        return ctx.track_batched(0, W.nF + 1 + k % 3, ident, scen.a0, scen.b0, scen.levels - 1)[0]

    def ba():
        return ctx.ba_iteration(0, 1e-5, stepfacD=0.0)        # stepfacD = 0: the window state is the same for every call

    try:
        for k in range(3):
            track(k); ba()
        # ---- serial reference ----
        t0 = time.perf_counter()
        ser_track = [track(k) for k in range(N_CALLS)]
        t_track_serial = (time.perf_counter() - t0) / N_CALLS
        ser_ba = [ba() for _ in range(N_CALLS)]
        # ---- two threads ----
        out = {}
        errs = []
        go = threading.Barrier(2)

        def tracking_thread():
            try:
                go.wait()
                t = time.perf_counter()
                out["track"] = [track(k) for k in range(N_CALLS)]
                out["t_track"] = (time.perf_counter() - t) / N_CALLS
            except Exception as e:      # noqa: BLE001 - reported below
                errs.append(e)

        def mapping_thread():
            try:
                go.wait()
                res = []
                t_end = None
                # keep the mapping stream busy for as long as the tracking thread runs, at least N_CALLS iterations
                while len(res) < N_CALLS or "track" not in out:
                    res.append(ba())
                    if len(res) > 100 * N_CALLS:
                        break
                out["ba"] = res
            except Exception as e:      # noqa: BLE001
                errs.append(e)

        th = [threading.Thread(target=tracking_thread), threading.Thread(target=mapping_thread)]
        for t in th:
            t.start()
        for t in th:
            t.join(120)
        assert not errs, errs
        assert "track" in out and "ba" in out
        # results identical to the serial run (every kernel is deterministic and the two groups share no buffers)
        for a, b in zip(ser_track, out["track"]):
            assert a.ok == b.ok == 1
            assert list(a.pose) == list(b.pose) and a.aff_a == b.aff_a and a.aff_b == b.aff_b
            assert list(a.trace[:a.trace_len]) == list(b.trace[:b.trace_len])
        for a, b in zip(ser_ba, out["ba"]):
            assert np.array_equal(a["x"], b["x"]) and a["energy"] == b["energy"]
        # a tracking call does not wait for the other thread's BA iterations (~0.2 ms each, back to back): with one stream
        # per context every call would sit behind at least one of them
        print("track ms/call: serial %.3f, with concurrent BA %.3f (%d BA iterations meanwhile)"
              % (t_track_serial * 1e3, out["t_track"] * 1e3, len(out["ba"])))
        assert out["t_track"] < t_track_serial + 0.15e-3
    finally:
        ctx.close()
