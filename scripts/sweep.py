"""Resolution / density sweep (BASELINE.json configs[3]) and the 31-hypothesis batch (configs[2]); run on the GPU box.

For each size: makeImages (K1) GB/s against the algorithmic bytes, one trackNewestCoarse (K2b) with its gather GB/s,
one BA step (K3-K5), each next to the CPU oracle on the same inputs and checked against it.  Prints one JSON line per
configuration (collected into profiles/ and BASELINE.md).
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import __graft_entry__ as entry

pkg = entry.load_package()
O = entry.load_oracle()
syn = pkg.synthetic
PEAK, _ = bench.peaks()


def ev_time(stream, fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn()
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


def run(w, h, npts, nF, levels):
    t0 = time.time()
    W = syn.make_window(seed=0, w=w, h=h, nF=nF, n_points=npts, kf_step_t=0.04, kf_step_deg=1.0)
    frames = syn.tracking_sequence(W, 2)
    ctx = pkg.Context(0, w, h, nF + 3)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    out = {"config": "%dx%d, %d active points, %d key frames, %d levels" % (w, h, W.nP, nF, levels), "gen_s": round(time.time() - t0, 1)}

    # ---- K1 ----
    img = torch.from_numpy(np.ascontiguousarray(frames[0][0])).to(dev)
    ms = ev_time(stream, lambda: ctx.make_images_dev(nF + 1, img.data_ptr(), w, h, levels), 20)
    px = sum((w >> l) * (h >> l) for l in range(levels))
    k1_bytes = 4 * w * h + 16 * px
    gd, ga = ctx.make_images(nF + 1, frames[0][0], levels)
    od, oa = O.make_images(frames[0][0], levels)
    k1_ok = all(np.array_equal(a, b) for a, b in zip(gd, od)) and all(np.array_equal(a, b) for a, b in zip(ga, oa))
    t1 = time.perf_counter()
    O.make_images(frames[0][0], levels, out=(od, oa))
    cpu_k1 = (time.perf_counter() - t1) * 1e3
    out["K1"] = {"ms": ms, "algorithmic_MB": k1_bytes / 1e6, "GBps": k1_bytes / ms / 1e6, "frac_of_measured_hbm": k1_bytes / ms / 1e6 / PEAK,
                 "bit_exact": bool(k1_ok), "cpu_ms": cpu_k1}

    # ---- K2 ----
    ref_pyr = O.make_images(W.images[-1], levels)
    trk = O.Tracker(w, h, levels, *W.K)
    cu, cv, cid, hdi = syn.tracker_reference_inputs(W)
    a0, b0 = W.aff_true[-1]
    trk.set_ref(cu, cv, cid, hdi, [x.reshape(-1) for x in ref_pyr[0]], a0, b0, 1.0)
    clouds = [trk.get_cloud(l) for l in range(levels)]
    ctx.tracker_set_ref(0, w, h, levels, W.K, clouds, a0, b0, 1.0)
    ident = syn.pose12((np.eye(3), np.zeros(3)))
    trk.set_new_frame([x.reshape(-1) for x in od], 1.0)
    t1 = time.perf_counter()
    ro = trk.track(ident, a0, b0, levels - 1)
    cpu_k2 = (time.perf_counter() - t1) * 1e3
    rg = ctx.track_batched(0, nF + 1, ident, a0, b0, levels - 1)[0]
    tstream = torch.cuda.ExternalStream(ctx.tracker_stream(0), device=dev)
    ms = ev_time(tstream, lambda: ctx.track_kernel_only(0, nF + 1, 1, levels - 1), 10)     # poses of the call above
    same = list(ro.trace[:ro.trace_len]) == list(rg.trace[:rg.trace_len])
    perr = float(np.abs(np.array(ro.pose) - np.array(rg.pose)).max())
    gb = rg.point_evals * bench.POINT_BYTES
    out["K2b"] = {"ms": ms, "pc_n": [len(c[0]) for c in clouds], "evals": rg.n_evals, "point_evals": rg.point_evals,
                  "gather_GBps": gb / ms / 1e6, "frac_of_measured_hbm": gb / ms / 1e6 / PEAK, "cpu_ms": cpu_k2,
                  "trace_match": bool(same), "pose_diff": perr, "ok": int(rg.ok)}

    # ---- K3-K5 ----
    pyr0 = [O.make_images(im, 1)[0][0] for im in W.images]
    B = O.BA(W, pyr0)
    for i, im in enumerate(W.images):
        ctx.make_images(i, im, 1, want_host=False)
    ctx.ba_set_window(np.arange(nF), w, h, W.K, W.frameEnergyTH)
    ctx.ba_set_frames(W)
    ctx.ba_set_points(W)
    ctx.ba_set_residuals(W)
    st = ctx.ba_step()
    t1 = time.perf_counter()
    B.linearize_mt(6); B.L.dso_oracle_ba_apply_res(B.p, None, None)
    Ao = B.accumulate_A(True, 6); So = B.accumulate_SC(True, 6)
    cpu_ba = (time.perf_counter() - t1) * 1e3
    errA = float(np.abs(st["H_A"] - Ao["H"]).max() / np.abs(Ao["H"]).max())
    errS = float(np.abs(st["H_sc"] - So["H"]).max() / np.abs(So["H"]).max())
    bstream = torch.cuda.ExternalStream(ctx.ba_stream, device=dev)
    ms = ev_time(bstream, lambda: ctx.ba_step(), 5)
    out["BA"] = {"ms": ms, "residuals": int(W.nR), "active": int(st["resInA"]), "algorithmic_MB": W.nR * 850 / 1e6,
                 "GBps": W.nR * 850 / ms / 1e6, "cpu_mt6_ms": cpu_ba, "relerr_HA": errA, "relerr_Hsc": errS}
    ctx.close()
    print(json.dumps(out), flush=True)


def batch31():
    wl = bench.build_workload(0, 4, 2000)
    W, levels = wl["W"], wl["levels"]
    ctx = pkg.Context(0, W.w, W.h, 8)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    ctx.tracker_set_ref(0, W.w, W.h, levels, W.K, wl["clouds"], wl["a0"], wl["b0"], 1.0)
    for k in range(4):
        ctx.make_images(1 + k, wl["frames"][k][0], levels, want_host=False)
    rel0, rel1 = wl["frames"][1][1], wl["frames"][2][1]
    hyps = syn.motion_hypotheses(syn.se3_mul(rel0, syn.se3_inv(rel1)), rel1)     # slast = frame 2, sprelast = frame 1
    P = np.stack([syn.pose12(x) for x in hyps])
    a, b = wl["affs"][3]
    res = ctx.track_batched(0, 4, P, a, b, levels - 1)
    tstream = torch.cuda.ExternalStream(ctx.tracker_stream(0), device=dev)
    ctx.track_batched(0, 4, P, a, b, levels - 1)
    ms31 = ev_time(tstream, lambda: ctx.track_kernel_only(0, 4, 31, levels - 1), 10)
    ctx.track_batched(0, 4, P[0], a, b, levels - 1)
    ms1 = ev_time(tstream, lambda: ctx.track_kernel_only(0, 4, 1, levels - 1), 10)
    trk = wl["oracle_tracker"]
    pyr = O.make_images(wl["frames"][3][0], levels)
    trk.set_new_frame([x.reshape(-1) for x in pyr[0]], 1.0)
    t1 = time.perf_counter()
    for i in range(31):
        trk.track(P[i], a, b, levels - 1)
    cpu = (time.perf_counter() - t1) * 1e3
    print(json.dumps({"config": "configs[2]: 31 pose priors of FullSystem::trackNewCoarse in one launch, 640x480",
                      "gpu_ms_31_hypotheses": ms31, "gpu_ms_1_hypothesis": ms1, "cpu_ms_31_sequential_no_abort": cpu,
                      "n_ok": int(sum(r.ok for r in res)), "evals_total": int(sum(r.n_evals for r in res))}), flush=True)
    ctx.close()


if __name__ == "__main__":
    batch31()
    for (w, h, n, lv) in [(640, 480, 2000, 4), (1280, 1024, 20000, 5), (1920, 1080, 50000, 4), (3840, 2160, 100000, 5)]:
        if len(sys.argv) > 1 and str(w) not in sys.argv[1:]:
            continue
        run(w, h, n, 4 if w > 640 else 7, lv)
