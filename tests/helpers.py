"""Shared scenario builders for the parity tests (oracle = checker, never the thing under test)."""
import numpy as np


def flat(pyr):
    return [np.ascontiguousarray(x).reshape(-1) for x in pyr]


class TrackScenario:
    """Reference cloud (oracle makeCoarseDepthL0) + new frames for one window."""

    def __init__(self, pkg, O, W, levels=None, n_frames=2):
        syn = pkg.synthetic
        self.W, self.syn, self.O = W, syn, O
        self.levels = levels or syn.pyr_levels_for(W.w, W.h)
        self.ref_pyr = O.make_images(W.images[-1], self.levels)
        self.trk = O.Tracker(W.w, W.h, self.levels, *W.K)
        cu, cv, cid, hdi = syn.tracker_reference_inputs(W)
        self.a0, self.b0 = W.aff_true[-1]
        self.trk.set_ref(cu, cv, cid, hdi, flat(self.ref_pyr[0]), self.a0, self.b0, 1.0)
        self.clouds = [self.trk.get_cloud(l) for l in range(self.levels)]
        self.frames = syn.tracking_sequence(W, n_frames)
        self.new_pyr = [O.make_images(f[0], self.levels) for f in self.frames]

    def oracle_set_frame(self, k):
        self.trk.set_new_frame(flat(self.new_pyr[k][0]), 1.0)

    def gpu_setup(self, ctx, trk=0):
        ctx.tracker_set_ref(trk, self.W.w, self.W.h, self.levels, self.W.K, self.clouds, self.a0, self.b0, 1.0)
        for k, f in enumerate(self.frames):
            ctx.make_images(1 + k, f[0], self.levels, want_host=False)

    def poses(self, k):
        """a few poses around the true relative pose of frame k"""
        syn = self.syn
        rel = self.frames[k][1]
        ident = (np.eye(3), np.zeros(3))
        rng = np.random.default_rng(100 + k)
        out = [ident, rel]
        for _ in range(2):
            out.append(syn.se3_mul(syn.se3_exp(rng.normal(0, 3e-3, 6)), rel))
        return [syn.pose12(p) for p in out]


def compact_warped(status, warped8):
    """device per-point debug output -> the reference's compacted, zero-padded buf_warped_* arrays"""
    keep = status == 2
    w = warped8[keep]
    n = w.shape[0]
    npad = (n + 3) // 4 * 4
    out = np.zeros((npad, 8), dtype=np.float32)
    out[:n] = w
    return npad, [out[:, k].copy() for k in range(8)]


def rel_err_matrix(A, B):
    """|A-B| relative to the scale of the entry: max(|B_ij|, sqrt(B_ii B_jj)) for square symmetric B."""
    A = np.asarray(A, dtype=np.float64)
    B = np.asarray(B, dtype=np.float64)
    d = np.sqrt(np.abs(np.diag(B)))
    scale = np.maximum(np.abs(B), np.outer(d, d))
    scale = np.where(scale > 0, scale, 1.0)
    return np.abs(A - B) / scale
