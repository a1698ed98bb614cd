"""ctypes binding of libdso_b200.so (the C ABI in include/dso_b200.h).

This is plumbing for the tests and bench.py; the product is the shared library.  Loading fails loudly
when the library has not been built, and every compute call raises DsoError when there is no CUDA device:
there is no CPU fallback anywhere in this package.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdso_b200.so")
MAX_LEVELS, MAX_FRAMES, MAX_HYP, MAX_PASSES = 6, 8, 32, 8

f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
vp = C.c_void_p


class DsoError(RuntimeError):
    pass


class TrackOpts(C.Structure):
    _fields_ = [("huberTH", C.c_float), ("coarseCutoffTH", C.c_float), ("affineOptModeA", C.c_float),
                ("affineOptModeB", C.c_float), ("maxIterations", C.c_int * 5)]


class TrackResult(C.Structure):
    _fields_ = [("ok", C.c_int), ("reserved", C.c_int), ("pose", C.c_double * 12), ("aff_a", C.c_double),
                ("aff_b", C.c_double), ("lastResiduals", C.c_double * 5), ("lastFlowIndicators", C.c_double * 3),
                ("n_passes", C.c_int), ("pass_level", C.c_int * MAX_PASSES), ("pass_rmse", C.c_double * MAX_PASSES),
                ("n_iterations", C.c_int), ("n_evals", C.c_int), ("trace_len", C.c_int), ("trace", C.c_ubyte * 256),
                ("point_evals", C.c_double), ("trace_dE", C.c_float * 64), ("trace_E", C.c_float * 64)]


class Settings(C.Structure):
    _fields_ = [("huberTH", C.c_float), ("outlierTHSumComponent", C.c_float), ("maxPixSearch", C.c_float),
                ("minTraceTestRadius", C.c_int), ("GNItsOnPointActivation", C.c_int), ("trace_stepsize", C.c_float),
                ("trace_GNIterations", C.c_int), ("trace_GNThreshold", C.c_float), ("trace_extraSlackOnTH", C.c_float),
                ("trace_slackInterval", C.c_float), ("trace_minImprovementFactor", C.c_float), ("minIdepthH_act", C.c_float)]


class Precalc(C.Structure):
    _fields_ = [("PRE_RTll_0", C.c_float * 9), ("PRE_tTll_0", C.c_float * 3), ("PRE_KRKiTll", C.c_float * 9),
                ("PRE_KtTll", C.c_float * 3), ("PRE_aff_mode", C.c_float * 2), ("PRE_b0_mode", C.c_float), ("pad", C.c_float)]


# every symbol include/dso_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "dso_b200_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "dso_b200_destroy": (None, [vp]),
    "dso_b200_last_error": (C.c_char_p, []),
    "dso_b200_settings_default": (None, [C.POINTER(Settings)]),
    "dso_b200_set_settings": (C.c_int, [vp, C.POINTER(Settings)]),
    "dso_b200_stream": (vp, [vp]),
    "dso_b200_stream_for": (vp, [vp, C.c_int]),
    "dso_b200_sync": (C.c_int, [vp]),
    "dso_b200_host_alloc": (vp, [C.c_size_t]),
    "dso_b200_host_free": (None, [vp]),
    "dso_b200_launch_count": (C.c_longlong, [vp]),
    "dso_b200_make_images": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, vp, C.POINTER(vp), C.POINTER(vp)]),
    "dso_b200_tracker_set_ref_points": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int,
                                                  f32p, f32p, f32p, f32p, C.c_double, C.c_double, C.c_float, vp]),
    "dso_b200_tracker_get_cloud": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, C.POINTER(C.c_int)]),
    "dso_b200_trace_set_points": (C.c_int, [vp, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, f32p, f32p, f32p, i32p, vp, vp]),
    "dso_b200_trace_new_coarse": (C.c_int, [vp, C.c_int, C.c_int, f32p, f32p, f32p, vp, vp, vp, vp, vp, vp]),
    "dso_b200_optimize_immature": (C.c_int, [vp, C.c_int, i32p, C.c_float, C.c_float, C.c_float, C.c_float, f32p, f32p, f32p, C.c_int,
                                             f32p, f32p, f32p, f32p, f32p, f32p, f32p, C.c_int, i32p, f32p, i32p]),
    "dso_b200_make_images_async": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int]),
    "dso_b200_make_images_dev": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, vp]),
    "dso_b200_tracker_set_ref": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                           i32p, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                           C.c_double, C.c_double, C.c_float]),
    "dso_b200_calc_res": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, f64p, C.c_double, C.c_double, C.c_float, C.c_float, f64p]),
    "dso_b200_calc_gs": (C.c_int, [vp, C.c_int, f64p, f64p]),
    "dso_b200_calc_res_noalloc": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                            f64p, C.c_double, C.c_double, C.c_float, C.c_double, C.c_double, C.c_float, C.c_float,
                                            vp, vp]),
    "dso_b200_calc_res_debug": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, f64p, C.c_double, C.c_double, C.c_float, C.c_float, u8p, f32p]),
    "dso_b200_track_opts_default": (None, [C.POINTER(TrackOpts)]),
    "dso_b200_track_batched": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, f64p, C.c_double, C.c_double, C.c_float, C.c_int,
                                         C.POINTER(TrackOpts), C.POINTER(TrackResult)]),
    "dso_b200_track_batched_async": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, f64p, C.c_double, C.c_double, C.c_float, C.c_int,
                                               C.POINTER(TrackOpts)]),
    "dso_b200_track_fetch": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(TrackResult)]),
    "dso_b200_track_and_prefetch": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, C.c_double, C.c_double, C.c_float, C.c_int, vp,
                                              C.c_int, vp, C.c_int, C.c_int, C.c_int, vp]),
    "dso_b200_track_frame": (C.c_int, [vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, f64p, C.c_double, C.c_double,
                                       C.c_float, C.c_int, C.POINTER(TrackOpts), C.POINTER(TrackResult)]),
    "dso_b200_track_phase_cycles": (C.c_int, [vp, C.c_int, vp]),
    "dso_b200_track_set_tuning": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
    "dso_b200_debug_lm_bench": (C.c_int, [vp, C.c_int, vp]),
    "dso_b200_track_last_shape": (C.c_int, [vp, i32p]),
    "dso_b200_track_replay_abort": (C.c_int, [C.POINTER(TrackResult), f64p, f64p]),
    "dso_b200_track_hostloop": (C.c_int, [vp, C.c_int, C.c_int, f64p, C.c_double, C.c_double, C.c_float, C.c_int, vp,
                                          C.POINTER(TrackOpts), C.POINTER(TrackResult)]),
    "dso_b200_ba_set_affine_modes": (C.c_int, [vp, C.c_float, C.c_float]),
    "dso_b200_init_calc_res_and_gs": (C.c_int, [vp, C.c_int, C.c_int, C.c_int] + [C.c_float] * 4 + [f64p, C.c_double, C.c_double, C.c_int] +
                                      [vp] * 7 + [C.c_float] * 3 + [vp] * 10),
    "dso_b200_ba_linearize_residuals": (C.c_int, [vp, C.c_int, i32p, C.c_int, C.c_int, vp, vp, vp]),
    "dso_b200_ba_set_calib_delta": (C.c_int, [vp, f64p]),
    "dso_b200_ba_set_priors": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "dso_b200_ba_set_nullspaces": (C.c_int, [vp, vp, C.c_int, C.c_double]),
    "dso_b200_ba_set_solver_state": (C.c_int, [vp, C.c_int, C.c_int, C.c_double]),
    "dso_b200_ba_get_point_terms": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "dso_b200_ba_fix_linearization": (C.c_int, [vp, vp, vp]),
    "dso_b200_ba_accumulate_L": (C.c_int, [vp, f64p, f64p, C.POINTER(C.c_int), vp, vp, vp]),
    "dso_b200_ba_solve": (C.c_int, [vp, C.c_int, C.c_double, C.c_int, vp, vp, vp, vp]),
    "dso_b200_ba_points_backup": (C.c_int, [vp]),
    "dso_b200_ba_points_step": (C.c_int, [vp, C.c_float, vp, vp]),
    "dso_b200_ba_points_restore": (C.c_int, [vp]),
    "dso_b200_ba_marginalize_points": (C.c_int, [vp, vp, C.c_float, C.c_double, vp, vp, vp, vp]),
    "dso_b200_ba_get_marg_prior": (C.c_int, [vp, f64p, f64p]),
    "dso_b200_ba_calc_energies": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "dso_b200_ba_iteration": (C.c_int, [vp, C.c_int, C.c_double, C.c_int, C.c_float, vp, vp, vp]),
    "dso_b200_ba_set_window": (C.c_int, [vp, C.c_int, i32p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, f32p]),
    "dso_b200_ba_set_glue": (C.c_int, [vp, vp, f64p, f64p, f32p, f32p]),
    "dso_b200_ba_set_frames": (C.c_int, [vp, f64p, f64p, f64p, f32p, vp, vp, vp]),
    "dso_b200_ba_set_points": (C.c_int, [vp, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, f32p, f32p]),
    "dso_b200_ba_set_residuals": (C.c_int, [vp, C.c_int, i32p, i32p, i32p, f64p, u8p, u8p]),
    "dso_b200_ba_linearize": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp]),
    "dso_b200_ba_apply_res": (C.c_int, [vp, vp, vp]),
    "dso_b200_ba_accumulate_A": (C.c_int, [vp, f64p, f64p, C.POINTER(C.c_int), vp, vp, vp]),
    "dso_b200_ba_accumulate_SC": (C.c_int, [vp, f64p, f64p, vp, vp, vp]),
    "dso_b200_ba_step": (C.c_int, [vp, f64p, f64p, f64p, f64p, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
}

_LIB = None


def load():
    """Loads the shared library (no CUDA call is made)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise DsoError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


NOALLOC_MAX_POINTS = 65536


def calc_res_noalloc(cloud_ptr, n_points, frame_ptr, w, h, is_level0, K_level, pose12, ref_a, ref_b, ref_exposure,
                     a, b, new_exposure, cutoff, stream, sums_ptr):
    """Raw call of dso_b200_calc_res_noalloc (device addresses as ints); returns (rc, error text)."""
    L = load()
    fx, fy, cx, cy = K_level
    rc = L.dso_b200_calc_res_noalloc(cloud_ptr, n_points, frame_ptr, w, h, int(is_level0), fx, fy, cx, cy,
                                     np.ascontiguousarray(pose12, dtype=np.float64), ref_a, ref_b, ref_exposure,
                                     a, b, new_exposure, cutoff, stream, sums_ptr)
    return rc, L.dso_b200_last_error().decode()


def _ptrs(arrs):
    a = (vp * len(arrs))()
    for i, x in enumerate(arrs):
        a[i] = None if x is None else x.ctypes.data
    return a


def _opt(a):
    return None if a is None else a.ctypes.data


class Context:
    def __init__(self, device=0, max_w=640, max_h=480, n_frame_slots=10):
        self.L = load()
        h = vp()
        self._check(self.L.dso_b200_create(device, max_w, max_h, n_frame_slots, C.byref(h)))
        self.h = h
        self.max_w, self.max_h = max_w, max_h

    def _check(self, rc):
        if rc != 0:
            raise DsoError("dso_b200 error %d: %s" % (rc, self.L.dso_b200_last_error().decode()))

    def close(self):
        if getattr(self, "h", None):
            self.L.dso_b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.L.dso_b200_stream(self.h)

    def default_settings(self):
        s = Settings()
        self.L.dso_b200_settings_default(C.byref(s))
        return s

    def set_settings(self, s):
        self._check(self.L.dso_b200_set_settings(self.h, C.byref(s)))

    def stream_for(self, which):
        """0 main (makeImages), 1 / 2 tracker slot 0 / 1, 3 mapping (BA, immature points), 4 upload"""
        return self.L.dso_b200_stream_for(self.h, which)

    def tracker_stream(self, trk=0):
        return self.stream_for(1 + trk)

    @property
    def ba_stream(self):
        return self.stream_for(3)

    def sync(self):
        self._check(self.L.dso_b200_sync(self.h))

    def launch_count(self):
        return self.L.dso_b200_launch_count(self.h)

    # ---- K1 ----
    def make_images(self, slot, color, levels, B=None, want_host=True):
        """color: [h,w] float32 (numpy) -> (dIp list [h_l,w_l,3], absgrad list [h_l,w_l]) or None."""
        color = np.ascontiguousarray(color, dtype=np.float32)
        h, w = color.shape
        Bp = None
        if B is not None:
            B = np.ascontiguousarray(B, dtype=np.float32)
            Bp = B.ctypes.data
        if want_host:
            dIp = [np.zeros((h >> l, w >> l, 3), dtype=np.float32) for l in range(levels)]
            ab = [np.zeros((h >> l, w >> l), dtype=np.float32) for l in range(levels)]
            self._check(self.L.dso_b200_make_images(self.h, slot, color.ctypes.data, w, h, levels, Bp, _ptrs(dIp), _ptrs(ab)))
            return dIp, ab
        self._check(self.L.dso_b200_make_images(self.h, slot, color.ctypes.data, w, h, levels, Bp, None, None))
        return None

    def make_images_ptr(self, slot, host_ptr, w, h, levels):
        """host pointer variant (e.g. pinned memory), no copy back."""
        self._check(self.L.dso_b200_make_images(self.h, slot, host_ptr, w, h, levels, None, None, None))

    def make_images_dev(self, slot, dev_ptr, w, h, levels, B=None):
        Bp = None
        if B is not None:
            B = np.ascontiguousarray(B, dtype=np.float32)
            Bp = B.ctypes.data
        self._check(self.L.dso_b200_make_images_dev(self.h, slot, dev_ptr, w, h, levels, Bp))

    # ---- K2 ----
    def tracker_set_ref(self, trk, w, h, levels, K, clouds, ref_a, ref_b, ref_exposure):
        """clouds: per level (u, v, idepth, color) float32 arrays."""
        fx, fy, cx, cy = K
        n = np.array([len(c[0]) for c in clouds], dtype=np.int32)
        cl = [[np.ascontiguousarray(a, dtype=np.float32) for a in c] for c in clouds]
        self._check(self.L.dso_b200_tracker_set_ref(self.h, trk, w, h, levels, fx, fy, cx, cy, n,
                                                    _ptrs([c[0] for c in cl]), _ptrs([c[1] for c in cl]),
                                                    _ptrs([c[2] for c in cl]), _ptrs([c[3] for c in cl]),
                                                    ref_a, ref_b, ref_exposure))

    def calc_res(self, trk, slot, lvl, pose12, a, b, cutoff, new_exposure=1.0):
        out = np.zeros(6)
        self._check(self.L.dso_b200_calc_res(self.h, trk, slot, lvl, np.ascontiguousarray(pose12, dtype=np.float64),
                                             a, b, new_exposure, cutoff, out))
        return out

    def calc_gs(self, trk):
        H = np.zeros(64)
        b = np.zeros(8)
        self._check(self.L.dso_b200_calc_gs(self.h, trk, H, b))
        return H.reshape(8, 8), b

    def calc_res_debug(self, trk, slot, lvl, n, pose12, a, b, cutoff, new_exposure=1.0):
        st = np.zeros(n + 16, dtype=np.uint8)
        wp = np.zeros((n + 1) * 8, dtype=np.float32)
        self._check(self.L.dso_b200_calc_res_debug(self.h, trk, slot, lvl, np.ascontiguousarray(pose12, dtype=np.float64),
                                                   a, b, new_exposure, cutoff, st, wp))
        return st[:n], wp[:n * 8].reshape(n, 8)

    def default_opts(self):
        o = TrackOpts()
        self.L.dso_b200_track_opts_default(C.byref(o))
        return o

    def track_hostloop(self, trk, slot, pose12, a, b, coarsest, min_res=None, new_exposure=1.0, opts=None):
        res = TrackResult()
        mr = None
        if min_res is not None:
            mr = np.ascontiguousarray(min_res, dtype=np.float64)
        self._check(self.L.dso_b200_track_hostloop(self.h, trk, slot, np.ascontiguousarray(pose12, dtype=np.float64), a, b,
                                                   new_exposure, coarsest, _opt(mr), None if opts is None else C.byref(opts),
                                                   C.byref(res)))
        return res

    def tracker_set_ref_points(self, trk, ref_slot, levels, K, cu, cv, cid, hdi, ref_a, ref_b, ref_exposure=1.0):
        """makeCoarseDepthL0 on the device; returns pc_n per level"""
        fx, fy, cx, cy = K
        arrs = [np.ascontiguousarray(x, dtype=np.float32) for x in (cu, cv, cid, hdi)]
        pc_n = np.zeros(8, dtype=np.int32)
        self._check(self.L.dso_b200_tracker_set_ref_points(self.h, trk, ref_slot, levels, fx, fy, cx, cy, len(arrs[0]), *arrs,
                                                           ref_a, ref_b, ref_exposure, pc_n.ctypes.data))
        return pc_n[:levels].copy()

    def tracker_get_cloud(self, trk, lvl):
        n = C.c_int(0)
        self._check(self.L.dso_b200_tracker_get_cloud(self.h, trk, lvl, 0, None, C.byref(n)))
        out = np.zeros((n.value, 4), dtype=np.float32)
        if n.value:
            self._check(self.L.dso_b200_tracker_get_cloud(self.h, trk, lvl, n.value, out.ctypes.data, C.byref(n)))
        return out

    def trace_set_points(self, u, v, fields, host_idx, state):
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        self._trace_n = len(u)
        self._check(self.L.dso_b200_trace_set_points(self.h, len(u), f(u), f(v), f(fields["color"]).reshape(-1), f(fields["weights"]).reshape(-1),
                                                     f(fields["gradH"]).reshape(-1), f(fields["energyTH"]),
                                                     np.ascontiguousarray(host_idx, dtype=np.int32), f(state["idepth_min"]), f(state["idepth_max"]),
                                                     f(state["quality"]), np.ascontiguousarray(state["status"], dtype=np.int32),
                                                     f(state["uv"]).ctypes.data, f(state["interval"]).ctypes.data))

    def trace_new_coarse(self, slot, KRKi, Kt, aff, fetch=True):
        KRKi = np.ascontiguousarray(KRKi, dtype=np.float32).reshape(-1, 9)
        n = self._trace_n
        out = dict(idepth_min=np.zeros(n, dtype=np.float32), idepth_max=np.zeros(n, dtype=np.float32), quality=np.zeros(n, dtype=np.float32),
                   status=np.zeros(n, dtype=np.int32), uv=np.zeros((n, 2), dtype=np.float32), interval=np.zeros(n, dtype=np.float32))
        ptrs = [out[k].ctypes.data if fetch else None for k in ("idepth_min", "idepth_max", "quality", "status", "uv", "interval")]
        self._check(self.L.dso_b200_trace_new_coarse(self.h, slot, KRKi.shape[0], KRKi.reshape(-1), np.ascontiguousarray(Kt, dtype=np.float32).reshape(-1),
                                                     np.ascontiguousarray(aff, dtype=np.float32).reshape(-1), *ptrs))
        return out

    def optimize_immature(self, target_slots, K, RTll, tTll, aff, u, v, fields, idepth_min, idepth_max, min_obs=1):
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
        fx, fy, cx, cy = K
        n, nT = len(u), len(target_slots)
        result = np.zeros(n, dtype=np.int32); idp = np.zeros(n, dtype=np.float32); rs = np.zeros(n * nT, dtype=np.int32)
        self._check(self.L.dso_b200_optimize_immature(self.h, nT, np.ascontiguousarray(target_slots, dtype=np.int32), fx, fy, cx, cy,
                                                      f(RTll), f(tTll), f(aff), n, f(u), f(v), f(fields["color"]), f(fields["weights"]),
                                                      f(fields["energyTH"]), f(idepth_min), f(idepth_max), min_obs, result, idp, rs))
        return result, idp, rs.reshape(n, nT)

    def make_images_async(self, slot, color_ptr, w, h, levels):
        """color_ptr: host address (pinned) or numpy array that the caller keeps alive until the slot is consumed"""
        if isinstance(color_ptr, np.ndarray):
            color_ptr = color_ptr.ctypes.data
        self._check(self.L.dso_b200_make_images_async(self.h, slot, color_ptr, w, h, levels))

    def track_batched(self, trk, slot, poses12, a, b, coarsest, new_exposure=1.0, opts=None):
        poses12 = np.ascontiguousarray(poses12, dtype=np.float64).reshape(-1, 12)
        n = poses12.shape[0]
        res = (TrackResult * n)()
        self._check(self.L.dso_b200_track_batched(self.h, trk, slot, n, poses12.reshape(-1), a, b, new_exposure, coarsest,
                                                  None if opts is None else C.byref(opts), res))
        return list(res)

    def track_kernel_only(self, trk, slot, n_hyp, coarsest, new_exposure=1.0):
        f = self.L.dso_b200_track_batched_async
        old = f.argtypes
        f.argtypes = old[:4] + [vp] + old[5:]
        try:
            self._check(f(self.h, trk, slot, n_hyp, None, 0.0, 0.0, new_exposure, coarsest, None))
        finally:
            f.argtypes = old

    def track_batched_async(self, trk, slot, poses12, a, b, coarsest, new_exposure=1.0, opts=None):
        poses12 = np.ascontiguousarray(poses12, dtype=np.float64).reshape(-1, 12)
        self._check(self.L.dso_b200_track_batched_async(self.h, trk, slot, poses12.shape[0], poses12.reshape(-1), a, b,
                                                        new_exposure, coarsest, None if opts is None else C.byref(opts)))

    def track_and_prefetch(self, trk, slot, pose12, a, b, coarsest, next_slot, next_color_ptr, w, h, levels, out=None,
                           new_exposure=1.0):
        """one hypothesis; pose12: contiguous float64 array of 12 kept alive by the caller for the call (or its address as an
        int); out: a TrackResult to fill (re-used by a caller that tracks in a loop)"""
        if out is None:
            out = TrackResult()
        rc = self.L.dso_b200_track_and_prefetch(self.h, trk, slot, 1, pose12 if isinstance(pose12, int) else pose12.ctypes.data, a, b,
                                                new_exposure, coarsest, None, next_slot, next_color_ptr, w, h, levels, C.addressof(out))
        if rc:
            self._check(rc)
        return out

    def track_fetch(self, trk, n=1):
        res = (TrackResult * n)()
        self._check(self.L.dso_b200_track_fetch(self.h, trk, n, res))
        return list(res)

    def track_frame(self, trk, slot, color_ptr, w, h, levels, poses12, a, b, coarsest, new_exposure=1.0, opts=None):
        """color_ptr: host address of w*h floats (numpy array or pinned pointer)"""
        if isinstance(color_ptr, np.ndarray):
            color_ptr = color_ptr.ctypes.data
        poses12 = np.ascontiguousarray(poses12, dtype=np.float64).reshape(-1, 12)
        n = poses12.shape[0]
        res = (TrackResult * n)()
        self._check(self.L.dso_b200_track_frame(self.h, trk, slot, color_ptr, w, h, levels, n, poses12.reshape(-1), a, b,
                                                new_exposure, coarsest, None if opts is None else C.byref(opts), res))
        return list(res)

    def phase_cycles(self, enable=True, read=False):
        out = np.zeros(120, dtype=np.int64)
        self._check(self.L.dso_b200_track_phase_cycles(self.h, int(enable), out.ctypes.data if read else None))
        return out

    def debug_lm_bench(self, iters=1000):
        out = np.zeros(7, dtype=np.int64)
        self._check(self.L.dso_b200_debug_lm_bench(self.h, iters, out.ctypes.data))
        return out / iters

    def track_set_tuning(self, threads=0, cluster=0, nclusters=0):
        self._check(self.L.dso_b200_track_set_tuning(self.h, threads, cluster, nclusters))

    def track_last_shape(self):
        out = np.zeros(3, dtype=np.int32)
        self._check(self.L.dso_b200_track_last_shape(self.h, out))
        return tuple(int(x) for x in out)

    def replay_abort(self, res, min_res):
        lr = np.zeros(5)
        ok = self.L.dso_b200_track_replay_abort(C.byref(res), np.ascontiguousarray(min_res, dtype=np.float64), lr)
        return ok, lr

    # ---- K3-K5 ----
    def ba_set_affine_modes(self, mode_a=1e12, mode_b=1e8):
        self._check(self.L.dso_b200_ba_set_affine_modes(self.h, mode_a, mode_b))

    def ba_set_window(self, slots, w, h, K, frameEnergyTH):
        fx, fy, cx, cy = K
        slots = np.ascontiguousarray(slots, dtype=np.int32)
        self.nF = len(slots)
        self._check(self.L.dso_b200_ba_set_window(self.h, len(slots), slots, w, h, fx, fy, cx, cy,
                                                  np.ascontiguousarray(frameEnergyTH, dtype=np.float32)))

    def ba_set_glue(self, precalc28, adHost, adTarget, adHTdeltaF, cDeltaF):
        pre = np.ascontiguousarray(precalc28, dtype=np.float32)
        self._check(self.L.dso_b200_ba_set_glue(self.h, pre.ctypes.data, np.ascontiguousarray(adHost, dtype=np.float64).reshape(-1),
                                                np.ascontiguousarray(adTarget, dtype=np.float64).reshape(-1),
                                                np.ascontiguousarray(adHTdeltaF, dtype=np.float32).reshape(-1),
                                                np.ascontiguousarray(cDeltaF, dtype=np.float32)))

    def ba_set_frames(self, W, want_out=False):
        nF = self.nF
        ev = np.ascontiguousarray(W.evalPT, dtype=np.float64).reshape(-1)
        st = np.ascontiguousarray(W.state, dtype=np.float64).reshape(-1)
        sz = np.ascontiguousarray(W.state_zero, dtype=np.float64).reshape(-1)
        ex = np.ascontiguousarray(W.ab_exposure, dtype=np.float32)
        if not want_out:
            self._check(self.L.dso_b200_ba_set_frames(self.h, ev, st, sz, ex, None, None, None))
            return None
        pre = np.zeros(nF * nF * 28, dtype=np.float32)
        adH = np.zeros(nF * nF * 64)
        adT = np.zeros(nF * nF * 64)
        self._check(self.L.dso_b200_ba_set_frames(self.h, ev, st, sz, ex, pre.ctypes.data, adH.ctypes.data, adT.ctypes.data))
        return dict(precalc=pre.reshape(nF * nF, 28), adHost=adH.reshape(nF * nF, 8, 8), adTarget=adT.reshape(nF * nF, 8, 8))

    def ba_set_points(self, W):
        self.nP = W.nP
        self._check(self.L.dso_b200_ba_set_points(self.h, W.nP, W.u, W.v, W.idepth_scaled, W.idepth_zero_scaled,
                                                  W.color.reshape(-1), W.weights.reshape(-1), W.host, W.priorF, W.deltaF))

    def ba_set_residuals(self, W):
        self.nR = W.nR
        self._check(self.L.dso_b200_ba_set_residuals(self.h, W.nR, W.res_point, W.res_target, W.res_state, W.res_energy,
                                                     W.res_isLinearized, W.res_isActive))

    def ba_linearize(self, want_J=True):
        nR = self.nR
        out = dict(newState=np.zeros(nR, dtype=np.int32), newEnergy=np.zeros(nR, dtype=np.float32),
                   newEnergyWithOutlier=np.zeros(nR, dtype=np.float32), centerProjectedTo=np.zeros(nR * 3, dtype=np.float32),
                   projectedTo=np.zeros(nR * 16, dtype=np.float32), J=np.zeros(nR * 74 if want_J else 1, dtype=np.float32))
        tot = np.zeros(1)
        self._check(self.L.dso_b200_ba_linearize(self.h, out["newState"].ctypes.data, out["newEnergy"].ctypes.data,
                                                 out["newEnergyWithOutlier"].ctypes.data, out["centerProjectedTo"].ctypes.data,
                                                 out["projectedTo"].ctypes.data, out["J"].ctypes.data if want_J else None,
                                                 tot.ctypes.data))
        out["total"] = float(tot[0])
        if want_J:
            out["J"] = out["J"].reshape(nR, 74)
        out["centerProjectedTo"] = out["centerProjectedTo"].reshape(nR, 3)
        out["projectedTo"] = out["projectedTo"].reshape(nR, 8, 2)
        return out

    def ba_apply_res(self):
        act = np.zeros(self.nR, dtype=np.uint8)
        jp = np.zeros(self.nR * 8, dtype=np.float32)
        self._check(self.L.dso_b200_ba_apply_res(self.h, act.ctypes.data, jp.ctypes.data))
        return act, jp.reshape(self.nR, 8)

    def ba_accumulate_A(self):
        N = 8 * self.nF + 4
        H = np.zeros(N * N)
        b = np.zeros(N)
        n = C.c_int(0)
        Hdd = np.zeros(self.nP, dtype=np.float32)
        bd = np.zeros(self.nP, dtype=np.float32)
        Hcd = np.zeros(self.nP * 4, dtype=np.float32)
        self._check(self.L.dso_b200_ba_accumulate_A(self.h, H, b, C.byref(n), Hdd.ctypes.data, bd.ctypes.data, Hcd.ctypes.data))
        return dict(H=H.reshape(N, N), b=b, resInA=n.value, Hdd=Hdd, bd=bd, Hcd=Hcd.reshape(self.nP, 4))

    def ba_accumulate_SC(self):
        N = 8 * self.nF + 4
        H = np.zeros(N * N)
        b = np.zeros(N)
        HdiF = np.zeros(self.nP, dtype=np.float32)
        bdSumF = np.zeros(self.nP, dtype=np.float32)
        ih = np.zeros(self.nP, dtype=np.float32)
        self._check(self.L.dso_b200_ba_accumulate_SC(self.h, H, b, HdiF.ctypes.data, bdSumF.ctypes.data, ih.ctypes.data))
        return dict(H=H.reshape(N, N), b=b, HdiF=HdiF, bdSumF=bdSumF, idepth_hessian=ih)

    # ---- next rows: solveSystemF tail, resubstituteF_MT, linearised residuals ----
    def ba_set_calib_delta(self, d4):
        self._check(self.L.dso_b200_ba_set_calib_delta(self.h, np.ascontiguousarray(d4, dtype=np.float64)))

    def ba_set_priors(self, cPrior=None, framePrior=None, frameDeltaPrior=None, frameDelta=None, HM=None, bM=None):
        arrs = [None if x is None else np.ascontiguousarray(x, dtype=np.float64) for x in
                (cPrior, framePrior, frameDeltaPrior, frameDelta, HM, bM)]
        self._check(self.L.dso_b200_ba_set_priors(self.h, *[None if x is None else x.ctypes.data for x in arrs]))

    def ba_set_nullspaces(self, ns, delta=1e-5):
        N = 8 * self.nF + 4
        ns = np.ascontiguousarray(ns, dtype=np.float64).reshape(-1, N)
        self._check(self.L.dso_b200_ba_set_nullspaces(self.h, ns.ctypes.data, ns.shape[0], delta))

    def ba_set_solver_state(self, have_first_frame=True, marg_solver_mode=128 | 2048, delta=1e-5):
        self._check(self.L.dso_b200_ba_set_solver_state(self.h, int(have_first_frame), int(marg_solver_mode), float(delta)))

    def ba_fix_linearization(self, which=None):
        out = np.zeros(self.nR * 8, dtype=np.float32)
        w = None if which is None else np.ascontiguousarray(which, dtype=np.uint8)
        self._check(self.L.dso_b200_ba_fix_linearization(self.h, None if w is None else w.ctypes.data, out.ctypes.data))
        return out.reshape(self.nR, 8)

    def ba_accumulate_L(self):
        N = 8 * self.nF + 4
        H, b = np.zeros(N * N), np.zeros(N)
        n = C.c_int(0)
        Hdd = np.zeros(self.nP, dtype=np.float32)
        bd = np.zeros(self.nP, dtype=np.float32)
        Hcd = np.zeros(self.nP * 4, dtype=np.float32)
        self._check(self.L.dso_b200_ba_accumulate_L(self.h, H, b, C.byref(n), Hdd.ctypes.data, bd.ctypes.data, Hcd.ctypes.data))
        return dict(H=H.reshape(N, N), b=b, resInL=n.value, Hdd=Hdd, bd=bd, Hcd=Hcd.reshape(self.nP, 4))

    def ba_solve(self, iteration, lam, solver_mode=128 | 2048, debug=False):
        N = 8 * self.nF + 4
        x = np.zeros(N)
        step = np.zeros(self.nP, dtype=np.float32)
        HF = np.zeros(N * N) if debug else None
        bF = np.zeros(N) if debug else None
        self._check(self.L.dso_b200_ba_solve(self.h, iteration, lam, solver_mode, x.ctypes.data, step.ctypes.data,
                                             None if HF is None else HF.ctypes.data, None if bF is None else bF.ctypes.data))
        out = dict(x=x, step=step)
        if debug:
            out.update(HFinal=HF.reshape(N, N), bFinal=bF)
        return out

    def ba_points_backup(self):
        self._check(self.L.dso_b200_ba_points_backup(self.h))

    def ba_points_step(self, stepfacD=1.0, want_idepth=True):
        sums = np.zeros(2, dtype=np.float32)
        idp = np.zeros(self.nP, dtype=np.float32) if want_idepth else None
        self._check(self.L.dso_b200_ba_points_step(self.h, stepfacD, sums.ctypes.data, None if idp is None else idp.ctypes.data))
        return sums, idp

    def ba_points_restore(self):
        self._check(self.L.dso_b200_ba_points_restore(self.h))

    def ba_iteration(self, iteration, lam, solver_mode=128 | 2048, stepfacD=1.0):
        N = 8 * self.nF + 4
        x = np.zeros(N)
        e = C.c_double(0)
        sums = np.zeros(2, dtype=np.float32)
        self._check(self.L.dso_b200_ba_iteration(self.h, iteration, lam, solver_mode, stepfacD, x.ctypes.data, C.byref(e), sums.ctypes.data))
        return dict(x=x, energy=e.value, step_sums=sums)

    def init_calc_res_and_gs(self, first_slot, new_slot, K, sc):
        """CoarseInitializer::calcResAndGS; sc: the scenario dict of synthetic.initializer_scenario"""
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
        n = len(sc["u"])
        a = dict(u=f(sc["u"]), v=f(sc["v"]), id=f(sc["idepth_new"]), g=np.ascontiguousarray(sc["isGood"], dtype=np.uint8), e=f(sc["energy"]),
                 o=f(sc["outlierTH"]), r=f(sc["iR"]))
        out = dict(H=np.zeros(64, dtype=np.float32), b=np.zeros(8, dtype=np.float32), Hsc=np.zeros(64, dtype=np.float32),
                   bsc=np.zeros(8, dtype=np.float32), res=np.zeros(3, dtype=np.float32), isGood_new=np.zeros(n, dtype=np.uint8),
                   energy_new=np.zeros(2 * n, dtype=np.float32), maxstep=np.zeros(n, dtype=np.float32),
                   lastHessian_new=np.zeros(n, dtype=np.float32), Jb=np.zeros(10 * n, dtype=np.float32))
        p = lambda x: x.ctypes.data
        self._check(self.L.dso_b200_init_calc_res_and_gs(
            self.h, first_slot, new_slot, sc["lvl"], K[0], K[1], K[2], K[3], np.ascontiguousarray(sc["pose12"], dtype=np.float64),
            sc["aff"][0], sc["aff"][1], n, p(a["u"]), p(a["v"]), p(a["id"]), p(a["g"]), p(a["e"]), p(a["o"]), p(a["r"]),
            sc["alphaW"], sc["alphaK"], sc["couplingWeight"], p(out["H"]), p(out["b"]), p(out["Hsc"]), p(out["bsc"]), p(out["res"]),
            p(out["isGood_new"]), p(out["energy_new"]), p(out["maxstep"]), p(out["lastHessian_new"]), p(out["Jb"])))
        out["H"] = out["H"].reshape(8, 8); out["Hsc"] = out["Hsc"].reshape(8, 8)
        out["energy_new"] = out["energy_new"].reshape(n, 2); out["Jb"] = out["Jb"].reshape(n, 10)
        return out

    def ba_linearize_residuals(self, idx, reset_oob=True, apply=True):
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        k = len(idx)
        ns, ne, ia = np.zeros(k, dtype=np.int32), np.zeros(k, dtype=np.float32), np.zeros(k, dtype=np.uint8)
        self._check(self.L.dso_b200_ba_linearize_residuals(self.h, k, idx, int(reset_oob), int(apply), ns.ctypes.data, ne.ctypes.data, ia.ctypes.data))
        return dict(newState=ns, newEnergy=ne, isActive=ia)

    def ba_marginalize_points(self, flags, prior_marg_fac=600.0 * 600.0, marg_weight_fac=0.25):
        N = 8 * self.nF + 4
        flags = np.ascontiguousarray(flags, dtype=np.uint8)
        H, b = np.zeros(N * N), np.zeros(N)
        n = C.c_int(0)
        ih = np.zeros(self.nP, dtype=np.float32)
        self._check(self.L.dso_b200_ba_marginalize_points(self.h, flags.ctypes.data, prior_marg_fac, marg_weight_fac, H.ctypes.data,
                                                          b.ctypes.data, C.addressof(n), ih.ctypes.data))
        HM, bM = np.zeros(N * N), np.zeros(N)
        self._check(self.L.dso_b200_ba_get_marg_prior(self.h, HM, bM))
        return dict(H=H.reshape(N, N), b=b, resInM=n.value, idepth_hessian=ih, HM=HM.reshape(N, N), bM=bM)

    def ba_calc_energies(self):
        l, m = C.c_double(0), C.c_double(0)
        self._check(self.L.dso_b200_ba_calc_energies(self.h, C.byref(l), C.byref(m)))
        return l.value, m.value

    def ba_get_point_terms(self):
        f = lambda n: np.zeros(n, dtype=np.float32)
        out = dict(Hdd=f(self.nP), bd=f(self.nP), Hcd=f(4 * self.nP), HdiF=f(self.nP), bdSumF=f(self.nP), idepth_hessian=f(self.nP))
        self._check(self.L.dso_b200_ba_get_point_terms(self.h, *[out[k].ctypes.data for k in ("Hdd", "bd", "Hcd", "HdiF", "bdSumF", "idepth_hessian")]))
        out["Hcd"] = out["Hcd"].reshape(self.nP, 4)
        return out

    def ba_step(self):
        N = 8 * self.nF + 4
        HA, bA, HS, bS = np.zeros(N * N), np.zeros(N), np.zeros(N * N), np.zeros(N)
        e = C.c_double(0)
        n = C.c_int(0)
        self._check(self.L.dso_b200_ba_step(self.h, HA, bA, HS, bS, C.byref(e), C.byref(n)))
        return dict(H_A=HA.reshape(N, N), b_A=bA, H_sc=HS.reshape(N, N), b_sc=bS, energy=e.value, resInA=n.value)
