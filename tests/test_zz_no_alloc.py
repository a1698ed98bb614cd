"""dso_b200_calc_res_noalloc: the allocation-free evaluation for embedders that may not allocate.

A tracker that lives inside a larger real-time process is often told "no GPU allocations after start-up":
cudaMalloc/cudaFree stall the device the mapping thread is also using and can fail once somebody else owns the
memory, and building a dso_b200 context is itself a big allocation.  The no-alloc entry point therefore works on
buffers the caller already has (here: torch tensors standing in for the embedder's memory) and keeps its
per-block partial sums in a fixed array inside the library.

This file is collected last on purpose.  It checks, through the C ABI, that the entry point returns the same
sums as the regular context path (dso_b200_calc_res + dso_b200_calc_gs) on two calls in a row -- the second call
sees the first call's leftovers in the shared array -- and that clouds beyond the fixed maximum are refused.
"""
import numpy as np
import pytest
from helpers import TrackScenario, rel_err_matrix

pytestmark = pytest.mark.gpu

SC = np.array([1.0, 1.0, 1.0, 0.5, 0.5, 0.5, 10.0, 1000.0])   # SCALE_XI_ROT x3, SCALE_XI_TRANS x3, SCALE_A, SCALE_B


def sums_to_vec6_H_b(s):
    """the conversion documented in include/dso_b200.h (CoarseTracker.cpp:338-353, 511-519)"""
    s = s.astype(np.float32)
    vec6 = np.array([s[0], s[4], float(s[1]) / (float(s[3]) + 0.1), 0.0, float(s[2]) / (float(s[3]) + 0.1),
                     float(np.float32(s[5]) / np.float32(s[4])) if s[4] > 0 else np.nan])
    n = (int(s[6]) + 3) // 4 * 4
    invn = np.float32(1.0) / np.float32(n) if n > 0 else np.float32(np.inf)
    H = np.zeros((8, 8))
    k = 8
    for i in range(8):
        for j in range(i, 8):
            H[i, j] = H[j, i] = float(s[k]) * float(invn) * SC[j] * SC[i]
            k += 1
    b = np.array([float(s[44 + i]) * float(invn) * SC[i] for i in range(8)])
    return vec6, H, b


def test_noalloc_matches_context_path_twice(pkg, oracle, window_small):
    import torch
    scen = TrackScenario(pkg, oracle, window_small, n_frames=1)
    W = scen.W
    ctx = pkg.Context(0, W.w, W.h, 3)
    scen.gpu_setup(ctx)
    dev = torch.device("cuda", 0)
    a, b = scen.frames[0][2]
    sums = torch.zeros(52, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    calls = [(0, scen.poses(0)[1], 20.0), (1, scen.poses(0)[2], 40.0)]    # two calls in a row, different level / pose / cut-off
    for lvl, pose, cutoff in calls:
        u, v, idp, col = scen.clouds[lvl]
        cloud = torch.from_numpy(np.stack([u, v, idp, col], axis=1).astype(np.float32)).to(dev).contiguous()
        dIp, ab = scen.new_pyr[0][0][lvl], scen.new_pyr[0][1][lvl]
        frame = torch.from_numpy(np.concatenate([dIp, ab[:, :, None]], axis=2).astype(np.float32)).to(dev).contiguous()
        (fx, fy, cx, cy), _ = scen.trk.get_K(lvl)
        hl, wl = ab.shape
        rc, err = pkg.capi.calc_res_noalloc(cloud.data_ptr(), cloud.shape[0], frame.data_ptr(), wl, hl, lvl == 0, (fx, fy, cx, cy),
                                            pose, scen.a0, scen.b0, 1.0, a, b, 1.0, cutoff, stream, sums.data_ptr())
        assert rc == 0, err
        torch.cuda.synchronize()
        v6, H, bb = sums_to_vec6_H_b(sums.cpu().numpy())
        ref6 = ctx.calc_res(0, 1, lvl, pose, a, b, cutoff)
        Href, bref = ctx.calc_gs(0)
        assert v6[1] == ref6[1] and v6[1] > 0                      # same number of terms
        assert (v6[5] == ref6[5]) or (np.isnan(v6[5]) and np.isnan(ref6[5]))
        assert abs(v6[0] - ref6[0]) <= 2e-5 * max(abs(ref6[0]), 1.0)
        assert abs(v6[2] - ref6[2]) <= 1e-5 * max(abs(ref6[2]), 1e-3) and abs(v6[4] - ref6[4]) <= 1e-5 * max(abs(ref6[4]), 1e-3)
        assert rel_err_matrix(H, Href).max() <= 1e-4
        scale = np.sqrt(np.abs(np.diag(Href)) * max(ref6[0] / ref6[1], 1e-12)) + 1e-30
        assert (np.abs(bb - bref) / scale).max() <= 1e-4
        # and directly against the oracle (CoarseTracker::calcRes / calcGSSSE): same term count, H and b within 1e-4
        scen.oracle_set_frame(0)
        ro = scen.trk.calc_res(lvl, pose, a, b, cutoff)
        Ho, bo = scen.trk.calc_gs(lvl, a, b)
        assert v6[1] == ro[1] and v6[5] == ro[5] or (np.isnan(v6[5]) and np.isnan(ro[5]))
        assert rel_err_matrix(H, Ho).max() <= 1e-4
        assert (np.abs(bb - bo) / scale).max() <= 1e-4
    ctx.close()


def test_noalloc_rejects_oversized_cloud(pkg):
    import torch
    dev = torch.device("cuda", 0)
    dummy = torch.zeros(64, dtype=torch.float32, device=dev)
    n = pkg.capi.NOALLOC_MAX_POINTS + 1
    rc, err = pkg.capi.calc_res_noalloc(dummy.data_ptr(), n, dummy.data_ptr(), 64, 48, 1, (50.0, 50.0, 31.5, 23.5),
                                        np.concatenate([np.eye(3).reshape(-1), np.zeros(3)]), 0.0, 0.0, 1.0, 0.0, 0.0, 1.0, 20.0,
                                        None, dummy.data_ptr())
    assert rc == -4 and "DSO_B200_NOALLOC_MAX_POINTS" in err
