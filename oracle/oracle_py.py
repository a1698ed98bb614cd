"""TEST INFRASTRUCTURE ONLY -- ctypes binding of the CPU oracle (oracle/libdso_oracle.so).

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product package.  PARITY UNPINNED: see oracle/dso_oracle.h.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


class TrackResult(C.Structure):
    _fields_ = [("ok", C.c_int), ("aborted_level", C.c_int), ("pose", C.c_double * 12),
                ("aff_a", C.c_double), ("aff_b", C.c_double), ("lastResiduals", C.c_double * 5),
                ("lastFlowIndicators", C.c_double * 3), ("n_passes", C.c_int), ("pass_level", C.c_int * 8),
                ("pass_rmse", C.c_double * 8), ("n_iterations", C.c_int), ("n_evals", C.c_int),
                ("trace_len", C.c_int), ("trace", C.c_ubyte * 256), ("point_evals", C.c_double),
                ("trace_dE", C.c_float * 64), ("trace_E", C.c_float * 64)]


def build(force=False):
    so = os.path.join(_HERE, "libdso_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle_tracker.cpp", "oracle_ba.cpp", "oracle_math.h", "dso_oracle.h", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "CXX=g++"], stdout=subprocess.DEVNULL)
    return so


class _Proxy:
    """Maps the dso_oracle_* names onto another library that exports the same C signatures under a different
    prefix (oracle/_ref/libdso_ref.so: the unmodified reference sources built by oracle/refbuild)."""

    def __init__(self, cdll, prefix):
        self._cdll, self._prefix = cdll, prefix

    def __getattr__(self, name):
        return getattr(self._cdll, name.replace("dso_oracle_", self._prefix))


REF_SO = os.path.join(_HERE, "_ref", "libdso_ref.so")
_LIBS = {}


def ref_available():
    return os.path.exists(REF_SO)


def lib(kind="oracle"):
    if kind in _LIBS:
        return _LIBS[kind]
    if kind == "oracle":
        so = os.path.join(_HERE, "libdso_oracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
    else:
        if not ref_available():
            raise RuntimeError("oracle/_ref/libdso_ref.so not built (make -C oracle/refbuild; needs /root/reference)")
        L = _Proxy(C.CDLL(REF_SO), "dso_ref_")
    vp = C.c_void_p

    def sig(name, argtypes, restype=None):
        try:
            f = getattr(L, name)
        except AttributeError:
            return          # the reference build has no counterpart for a few debug helpers
        f.argtypes = argtypes
        if restype is not None:
            f.restype = restype

    sig("dso_oracle_set_affine_modes", [C.c_float, C.c_float])
    sig("dso_oracle_set_settings", [vp])
    sig("dso_oracle_init_make_K", [C.c_float] * 4 + [C.c_int, f64p, f64p])
    sig("dso_oracle_init_calc_res_and_gs", [C.c_int, C.c_int, C.c_int] + [C.c_float] * 4 + [f32p, f32p, f64p, C.c_double, C.c_double,
                                            C.c_int, f32p, f32p, f32p, u8p, f32p, f32p, f32p, C.c_float, C.c_float, C.c_float,
                                            f32p, f32p, f32p, f32p, f32p, u8p, f32p, f32p, f32p, f32p])
    sig("dso_oracle_make_images", [f32p, C.c_int, C.c_int, C.c_int, vp, C.POINTER(vp), C.POINTER(vp)])
    sig("dso_oracle_tracker_create", [C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float], vp)
    sig("dso_oracle_tracker_destroy", [vp])
    sig("dso_oracle_tracker_get_K", [vp, C.c_int] + [C.POINTER(C.c_float)] * 4 + [vp])
    sig("dso_oracle_tracker_set_ref", [vp, C.c_int, f32p, f32p, f32p, f32p, C.POINTER(vp), C.c_double, C.c_double, C.c_float])
    sig("dso_oracle_tracker_set_cloud", [vp, C.c_int, C.c_int, f32p, f32p, f32p, f32p])
    sig("dso_oracle_tracker_set_ref_aff", [vp, C.c_double, C.c_double, C.c_float])
    sig("dso_oracle_tracker_get_cloud", [vp, C.c_int, vp, vp, vp, vp], C.c_int)
    sig("dso_oracle_tracker_set_new_frame", [vp, C.POINTER(vp), C.c_float])
    sig("dso_oracle_calc_res", [vp, C.c_int, f64p, C.c_double, C.c_double, C.c_float, f64p])
    sig("dso_oracle_calc_gs", [vp, C.c_int, C.c_double, C.c_double, f64p, f64p])
    sig("dso_oracle_get_warped", [vp, C.POINTER(vp)], C.c_int)
    sig("dso_oracle_get_point_status", [vp, u8p])
    sig("dso_oracle_track_newest_coarse", [vp, f64p, C.c_double, C.c_double, C.c_int, f64p, C.POINTER(TrackResult)])
    sig("dso_oracle_ba_create", [C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float], vp)
    sig("dso_oracle_ba_destroy", [vp])
    sig("dso_oracle_ba_set_frame", [vp, C.c_int, f64p, f64p, f64p, C.c_float, C.c_float, f32p])
    sig("dso_oracle_ba_set_calib_delta", [vp, f64p])
    sig("dso_oracle_ba_prepare", [vp])
    sig("dso_oracle_ba_get_glue", [vp, f32p, f64p, f64p, f32p, f32p])
    sig("dso_oracle_ba_set_points", [vp, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, f32p, f32p])
    sig("dso_oracle_ba_set_residuals", [vp, C.c_int, i32p, i32p, i32p, f64p, u8p, u8p])
    sig("dso_oracle_ba_linearize", [vp, vp, vp, vp, vp, vp, vp, vp], C.c_double)
    sig("dso_oracle_ba_linearize_mt", [vp, C.c_int], C.c_double)
    sig("dso_oracle_ba_apply_res", [vp, vp, vp])
    sig("dso_oracle_ba_linearize_residuals", [vp, C.c_int, vp, C.c_int, C.c_int, vp, vp, vp])
    sig("dso_oracle_ba_accumulate_A", [vp, C.c_int, C.c_int, f64p, f64p, C.POINTER(C.c_int), f32p, f32p, f32p])
    sig("dso_oracle_ba_accumulate_SC", [vp, C.c_int, C.c_int, f64p, f64p, f32p, f32p, f32p])
    sig("dso_oracle_ba_accumulate_L", [vp, C.c_int, C.c_int, f64p, f64p, C.POINTER(C.c_int), f32p, f32p, f32p])
    sig("dso_oracle_ba_set_priors", [vp, vp, vp, vp, vp, vp, vp])
    sig("dso_oracle_ba_set_nullspaces", [vp, f64p, C.c_int])
    sig("dso_oracle_ba_fix_linearization", [vp, vp, vp])
    sig("dso_oracle_ba_calc_L_energy", [vp, C.c_int], C.c_double)
    sig("dso_oracle_ba_calc_M_energy", [vp], C.c_double)
    sig("dso_oracle_ba_marginalize_points", [vp, vp, C.c_float, C.c_double, vp, vp, vp, vp])
    sig("dso_oracle_ba_get_marg_prior", [vp, f64p, f64p])
    sig("dso_oracle_immature_init", [C.c_int, C.c_int, f32p, C.c_int, i32p, i32p, f32p, f32p, f32p, f32p])
    sig("dso_oracle_trace_on", [C.c_int, C.c_int, f32p, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f32p, i32p,
                                f32p, f32p, f32p, f32p, f32p])
    sig("dso_oracle_optimize_immature", [C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.POINTER(vp), f32p, f32p, f32p,
                                         C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, f32p, C.c_int, i32p, f32p, i32p])
    sig("dso_oracle_scaled_ldlt_solve", [f64p, f64p, C.c_int, f64p])
    sig("dso_oracle_ba_solve", [vp, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int, C.c_int, vp, vp, vp, vp], C.c_int)
    _LIBS[kind] = L
    return L


class Settings(C.Structure):
    """util/settings.cpp values read by the restated functions; same layout as dso_b200_settings"""
    _fields_ = [("huberTH", C.c_float), ("outlierTHSumComponent", C.c_float), ("maxPixSearch", C.c_float),
                ("minTraceTestRadius", C.c_int), ("GNItsOnPointActivation", C.c_int), ("trace_stepsize", C.c_float),
                ("trace_GNIterations", C.c_int), ("trace_GNThreshold", C.c_float), ("trace_extraSlackOnTH", C.c_float),
                ("trace_slackInterval", C.c_float), ("trace_minImprovementFactor", C.c_float), ("minIdepthH_act", C.c_float)]

    @staticmethod
    def defaults():
        return Settings(9, 50 * 50, 0.027, 2, 3, 1.0, 3, 0.1, 1.2, 1.5, 2, 100)


def set_settings(s=None, kind="oracle"):
    """process-wide like the reference's globals: call with no argument to restore the defaults after use.  `s` may be any
    ctypes structure with the layout of Settings (e.g. the product binding's)."""
    lib(kind).dso_oracle_set_settings(C.byref(s) if s is not None else None)


def set_affine_modes(mode_a=1e12, mode_b=1e8, kind="oracle"):
    """setting_affineOptModeA/B (settings.cpp:119-120); process-wide: restore the defaults after use."""
    lib(kind).dso_oracle_set_affine_modes(mode_a, mode_b)


def scaled_ldlt_solve(HF, bF):
    """EnergyFunctional.cpp:905-908 on a given system (oracle build only)"""
    HF = np.ascontiguousarray(HF, dtype=np.float64)
    bF = np.ascontiguousarray(bF, dtype=np.float64)
    x = np.zeros(len(bF))
    lib("oracle").dso_oracle_scaled_ldlt_solve(HF.reshape(-1), bF, len(bF), x)
    return x


def immature_init(dI, u, v, kind="oracle"):
    """ImmaturePoint constructor fields for integer pixels (u, v) of the AoS level-0 image dI [h, w, 3]"""
    h, w = dI.shape[:2]
    n = len(u)
    out = dict(color=np.zeros((n, 8), dtype=np.float32), weights=np.zeros((n, 8), dtype=np.float32),
               gradH=np.zeros((n, 4), dtype=np.float32), energyTH=np.zeros(n, dtype=np.float32))
    lib(kind).dso_oracle_immature_init(w, h, np.ascontiguousarray(dI, dtype=np.float32).reshape(-1), n,
                                       np.ascontiguousarray(u, dtype=np.int32), np.ascontiguousarray(v, dtype=np.int32),
                                       out["color"].reshape(-1), out["weights"].reshape(-1), out["gradH"].reshape(-1), out["energyTH"])
    return out


def trace_on(dI, u, v, fields, state, KRKi, Kt, aff, kind="oracle"):
    """ImmaturePoint::traceOn on all points; state: dict(idepth_min, idepth_max, quality, status, uv, interval), updated in place"""
    h, w = dI.shape[:2]
    lib(kind).dso_oracle_trace_on(w, h, np.ascontiguousarray(dI, dtype=np.float32).reshape(-1), len(u),
                                  np.ascontiguousarray(u, dtype=np.float32), np.ascontiguousarray(v, dtype=np.float32),
                                  fields["color"].reshape(-1), fields["weights"].reshape(-1), fields["gradH"].reshape(-1), fields["energyTH"],
                                  state["idepth_min"], state["idepth_max"], state["quality"], state["status"],
                                  np.ascontiguousarray(KRKi, dtype=np.float32).reshape(-1), np.ascontiguousarray(Kt, dtype=np.float32),
                                  np.ascontiguousarray(aff, dtype=np.float32), state["uv"].reshape(-1), state["interval"])
    return state


def trace_state(n):
    """a fresh immature point: idepth_min = 0, idepth_max = NaN, quality = 10000, IPS_UNINITIALIZED (ImmaturePoint.cpp:25-57, .h)"""
    return dict(idepth_min=np.zeros(n, dtype=np.float32), idepth_max=np.full(n, np.nan, dtype=np.float32),
                quality=np.full(n, 10000.0, dtype=np.float32), status=np.full(n, 5, dtype=np.int32),
                uv=np.full((n, 2), -1.0, dtype=np.float32), interval=np.zeros(n, dtype=np.float32))


def optimize_immature(K, target_dIs, RTll, tTll, aff, u, v, fields, idepth_min, idepth_max, min_obs=1, kind="oracle"):
    """FullSystem::optimizeImmaturePoint on every point; returns (result, idepth, res_state [n, nT])"""
    h, w = target_dIs[0].shape[:2]
    fx, fy, cx, cy = K
    dI = [np.ascontiguousarray(x, dtype=np.float32) for x in target_dIs]
    n, nT = len(u), len(dI)
    result = np.zeros(n, dtype=np.int32); idp = np.zeros(n, dtype=np.float32); rs = np.zeros(n * nT, dtype=np.int32)
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
    lib(kind).dso_oracle_optimize_immature(w, h, fx, fy, cx, cy, nT, _ptr_array(dI), f(RTll), f(tTll), f(aff), n, f(u), f(v),
                                           f(fields["color"]), f(fields["weights"]), f(fields["energyTH"]), f(idepth_min), f(idepth_max),
                                           min_obs, result, idp, rs)
    return result, idp, rs.reshape(n, nT)


def init_make_K(K, lvl, kind="oracle"):
    """CoarseInitializer::makeK for one level: (fx, fy, cx, cy) and K^-1 in double"""
    k4, Ki = np.zeros(4), np.zeros(9)
    lib(kind).dso_oracle_init_make_K(K[0], K[1], K[2], K[3], lvl, k4, Ki)
    return k4, Ki.reshape(3, 3)


def init_calc_res_and_gs(K, first_dIp, new_dIp, sc, kind="oracle"):
    """CoarseInitializer::calcResAndGS on level sc["lvl"]; first_dIp / new_dIp: dIp[lvl] ([h_l, w_l, 3]) of the two frames;
    sc: the scenario dict of synthetic.initializer_scenario"""
    lvl = sc["lvl"]
    hl, wl = first_dIp.shape[:2]
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
    n = len(sc["u"])
    out = dict(H=np.zeros(64, dtype=np.float32), b=np.zeros(8, dtype=np.float32), Hsc=np.zeros(64, dtype=np.float32),
               bsc=np.zeros(8, dtype=np.float32), res=np.zeros(3, dtype=np.float32), isGood_new=np.zeros(n, dtype=np.uint8),
               energy_new=np.zeros(2 * n, dtype=np.float32), maxstep=np.zeros(n, dtype=np.float32),
               lastHessian_new=np.zeros(n, dtype=np.float32), Jb=np.zeros(10 * n, dtype=np.float32))
    lib(kind).dso_oracle_init_calc_res_and_gs(wl << lvl, hl << lvl, lvl, K[0], K[1], K[2], K[3], f(first_dIp), f(new_dIp),
                                              np.ascontiguousarray(sc["pose12"], dtype=np.float64), sc["aff"][0], sc["aff"][1],
                                              n, f(sc["u"]), f(sc["v"]), f(sc["idepth_new"]), np.ascontiguousarray(sc["isGood"], dtype=np.uint8),
                                              f(sc["energy"]), f(sc["outlierTH"]), f(sc["iR"]), sc["alphaW"], sc["alphaK"], sc["couplingWeight"],
                                              out["H"], out["b"], out["Hsc"], out["bsc"], out["res"], out["isGood_new"], out["energy_new"],
                                              out["maxstep"], out["lastHessian_new"], out["Jb"])
    out["H"] = out["H"].reshape(8, 8); out["Hsc"] = out["Hsc"].reshape(8, 8)
    out["energy_new"] = out["energy_new"].reshape(n, 2); out["Jb"] = out["Jb"].reshape(n, 10)
    return out


def _ptr_array(arrs):
    a = (C.c_void_p * len(arrs))()
    for i, x in enumerate(arrs):
        a[i] = x.ctypes.data
    return a


def make_images(color, levels, B=None, out=None, kind="oracle"):
    """Returns (dIp list of [h_l, w_l, 3] float32, absgrad list of [h_l, w_l] float32); untouched elements are 0.
    out: a previous return value to overwrite (skips the allocation, as the reference's new[] does not zero)."""
    h, w = color.shape
    color = np.ascontiguousarray(color, dtype=np.float32)
    if out is not None:
        dIp, ab = out
    else:
        dIp = [np.zeros((h >> l, w >> l, 3), dtype=np.float32) for l in range(levels)]
        ab = [np.zeros((h >> l, w >> l), dtype=np.float32) for l in range(levels)]
    Bp = None
    if B is not None:
        B = np.ascontiguousarray(B, dtype=np.float32)
        Bp = B.ctypes.data
    lib(kind).dso_oracle_make_images(color, w, h, levels, Bp, _ptr_array(dIp), _ptr_array(ab))
    return dIp, ab


class Tracker:
    def __init__(self, w, h, levels, fx, fy, cx, cy, kind="oracle"):
        self.L = lib(kind)
        self.w, self.h, self.levels = w, h, levels
        self.p = self.L.dso_oracle_tracker_create(w, h, levels, fx, fy, cx, cy)
        self._keep = []

    def __del__(self):
        if getattr(self, "p", None):
            self.L.dso_oracle_tracker_destroy(self.p)
            self.p = None

    def get_K(self, lvl):
        v = [C.c_float() for _ in range(4)]
        Ki = np.zeros(9, dtype=np.float32)
        self.L.dso_oracle_tracker_get_K(self.p, lvl, *[C.byref(x) for x in v], Ki.ctypes.data)
        return [x.value for x in v], Ki

    def set_ref(self, cu, cv, cid, hdi, ref_dIp, ref_a, ref_b, ref_exposure):
        self._ref = [np.ascontiguousarray(x, dtype=np.float32) for x in ref_dIp]
        self.L.dso_oracle_tracker_set_ref(self.p, len(cu), cu, cv, cid, hdi, _ptr_array(self._ref), ref_a, ref_b, ref_exposure)

    def set_cloud(self, lvl, u, v, idepth, color):
        self.L.dso_oracle_tracker_set_cloud(self.p, lvl, len(u), u, v, idepth, color)

    def set_ref_aff(self, a, b, exposure):
        self.L.dso_oracle_tracker_set_ref_aff(self.p, a, b, exposure)

    def get_cloud(self, lvl):
        n = self.L.dso_oracle_tracker_get_cloud(self.p, lvl, None, None, None, None)
        arrs = [np.zeros(n, dtype=np.float32) for _ in range(4)]
        self.L.dso_oracle_tracker_get_cloud(self.p, lvl, *[a.ctypes.data for a in arrs])
        return arrs

    def set_new_frame(self, dIp, exposure=1.0):
        self._new = [np.ascontiguousarray(x, dtype=np.float32) for x in dIp]
        self.L.dso_oracle_tracker_set_new_frame(self.p, _ptr_array(self._new), exposure)

    def calc_res(self, lvl, pose12, a, b, cutoff):
        out = np.zeros(6)
        self.L.dso_oracle_calc_res(self.p, lvl, np.ascontiguousarray(pose12, dtype=np.float64), a, b, cutoff, out)
        return out

    def calc_gs(self, lvl, a, b):
        H = np.zeros(64)
        bb = np.zeros(8)
        self.L.dso_oracle_calc_gs(self.p, lvl, a, b, H, bb)
        return H.reshape(8, 8), bb

    def get_warped(self):
        bufs = [np.zeros(self.w * self.h + 8, dtype=np.float32) for _ in range(8)]
        n = self.L.dso_oracle_get_warped(self.p, _ptr_array(bufs))
        return n, [b[:n] for b in bufs]

    def get_point_status(self, n):
        s = np.zeros(self.w * self.h, dtype=np.uint8)
        self.L.dso_oracle_get_point_status(self.p, s)
        return s[:n]

    def track(self, pose12, a, b, coarsest, min_res_for_abort=None):
        if min_res_for_abort is None:
            min_res_for_abort = np.full(5, np.nan)
        res = TrackResult()
        self.L.dso_oracle_track_newest_coarse(self.p, np.ascontiguousarray(pose12, dtype=np.float64), a, b, coarsest,
                                              np.ascontiguousarray(min_res_for_abort, dtype=np.float64), C.byref(res))
        return res


class BA:
    def __init__(self, W, dI_list, kind="oracle"):
        """W: synthetic.Window; dI_list: per frame level-0 AoS image [h,w,3] float32."""
        self.L = lib(kind)
        fx, fy, cx, cy = W.K
        self.nF, self.nP, self.nR = W.nF, W.nP, W.nR
        self.p = self.L.dso_oracle_ba_create(W.w, W.h, W.nF, fx, fy, cx, cy)
        self._dI = [np.ascontiguousarray(x, dtype=np.float32) for x in dI_list]
        for i in range(W.nF):
            self.L.dso_oracle_ba_set_frame(self.p, i, np.ascontiguousarray(W.evalPT[i]), np.ascontiguousarray(W.state[i]),
                                           np.ascontiguousarray(W.state_zero[i]), float(W.ab_exposure[i]),
                                           float(W.frameEnergyTH[i]), self._dI[i].reshape(-1))
        self.L.dso_oracle_ba_set_calib_delta(self.p, np.ascontiguousarray(W.calib_delta, dtype=np.float64))
        self.L.dso_oracle_ba_prepare(self.p)
        self.L.dso_oracle_ba_set_points(self.p, W.nP, W.u, W.v, W.idepth_scaled, W.idepth_zero_scaled,
                                        W.color.reshape(-1), W.weights.reshape(-1), W.host, W.priorF, W.deltaF)
        self.set_residuals(W)

    def set_residuals(self, W):
        self.L.dso_oracle_ba_set_residuals(self.p, W.nR, W.res_point, W.res_target, W.res_state, W.res_energy,
                                           W.res_isLinearized, W.res_isActive)

    def __del__(self):
        if getattr(self, "p", None):
            self.L.dso_oracle_ba_destroy(self.p)
            self.p = None

    def glue(self):
        nF = self.nF
        pre = np.zeros(nF * nF * 28, dtype=np.float32)
        adH = np.zeros(nF * nF * 64)
        adT = np.zeros(nF * nF * 64)
        adHTd = np.zeros(nF * nF * 8, dtype=np.float32)
        cD = np.zeros(4, dtype=np.float32)
        self.L.dso_oracle_ba_get_glue(self.p, pre, adH, adT, adHTd, cD)
        return dict(precalc=pre.reshape(nF * nF, 28), adHost=adH.reshape(nF * nF, 8, 8), adTarget=adT.reshape(nF * nF, 8, 8),
                    adHTdeltaF=adHTd.reshape(nF * nF, 8), cDeltaF=cD)

    def linearize(self):
        nR = self.nR
        out = dict(newState=np.zeros(nR, dtype=np.int32), newEnergy=np.zeros(nR, dtype=np.float32),
                   newEnergyWithOutlier=np.zeros(nR, dtype=np.float32), centerProjectedTo=np.zeros(nR * 3, dtype=np.float32),
                   projectedTo=np.zeros(nR * 16, dtype=np.float32), J=np.zeros(nR * 74, dtype=np.float32),
                   energy=np.zeros(nR, dtype=np.float64))
        out["total"] = self.L.dso_oracle_ba_linearize(self.p, *[out[k].ctypes.data for k in
                                                      ("newState", "newEnergy", "newEnergyWithOutlier", "centerProjectedTo", "projectedTo", "J", "energy")])
        out["J"] = out["J"].reshape(nR, 74)
        out["centerProjectedTo"] = out["centerProjectedTo"].reshape(nR, 3)
        out["projectedTo"] = out["projectedTo"].reshape(nR, 8, 2)
        return out

    def linearize_mt(self, nthreads):
        return self.L.dso_oracle_ba_linearize_mt(self.p, nthreads)

    def linearize_residuals(self, idx, reset_oob=True, apply=True):
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        k = len(idx)
        ns, ne, ia = np.zeros(k, dtype=np.int32), np.zeros(k, dtype=np.float32), np.zeros(k, dtype=np.uint8)
        self.L.dso_oracle_ba_linearize_residuals(self.p, k, idx.ctypes.data, int(reset_oob), int(apply), ns.ctypes.data, ne.ctypes.data, ia.ctypes.data)
        return dict(newState=ns, newEnergy=ne, isActive=ia)

    def apply_res(self):
        act = np.zeros(self.nR, dtype=np.uint8)
        jp = np.zeros(self.nR * 8, dtype=np.float32)
        self.L.dso_oracle_ba_apply_res(self.p, act.ctypes.data, jp.ctypes.data)
        return act, jp.reshape(self.nR, 8)

    def accumulate_A(self, MT=False, nthreads=6):
        N = 8 * self.nF + 4
        H = np.zeros(N * N)
        b = np.zeros(N)
        n = C.c_int(0)
        Hdd = np.zeros(self.nP, dtype=np.float32)
        bd = np.zeros(self.nP, dtype=np.float32)
        Hcd = np.zeros(self.nP * 4, dtype=np.float32)
        self.L.dso_oracle_ba_accumulate_A(self.p, int(MT), nthreads, H, b, C.byref(n), Hdd, bd, Hcd)
        return dict(H=H.reshape(N, N), b=b, resInA=n.value, Hdd=Hdd, bd=bd, Hcd=Hcd.reshape(self.nP, 4))

    def accumulate_SC(self, MT=False, nthreads=6):
        N = 8 * self.nF + 4
        H = np.zeros(N * N)
        b = np.zeros(N)
        HdiF = np.zeros(self.nP, dtype=np.float32)
        bdSumF = np.zeros(self.nP, dtype=np.float32)
        ih = np.zeros(self.nP, dtype=np.float32)
        self.L.dso_oracle_ba_accumulate_SC(self.p, int(MT), nthreads, H, b, HdiF, bdSumF, ih)
        return dict(H=H.reshape(N, N), b=b, HdiF=HdiF, bdSumF=bdSumF, idepth_hessian=ih)

    def accumulate_L(self, MT=False, nthreads=6):
        N = 8 * self.nF + 4
        H = np.zeros(N * N)
        b = np.zeros(N)
        n = C.c_int(0)
        Hdd = np.zeros(self.nP, dtype=np.float32)
        bd = np.zeros(self.nP, dtype=np.float32)
        Hcd = np.zeros(self.nP * 4, dtype=np.float32)
        self.L.dso_oracle_ba_accumulate_L(self.p, int(MT), nthreads, H, b, C.byref(n), Hdd, bd, Hcd)
        return dict(H=H.reshape(N, N), b=b, resInL=n.value, Hdd=Hdd, bd=bd, Hcd=Hcd.reshape(self.nP, 4))

    def set_priors(self, cPrior, framePrior, frameDeltaPrior, frameDelta, HM, bM):
        self._pri = [None if x is None else np.ascontiguousarray(x, dtype=np.float64) for x in
                     (cPrior, framePrior, frameDeltaPrior, frameDelta, HM, bM)]
        self.L.dso_oracle_ba_set_priors(self.p, *[None if x is None else x.ctypes.data for x in self._pri])

    def set_nullspaces(self, ns):
        ns = np.ascontiguousarray(ns, dtype=np.float64).reshape(-1, 8 * self.nF + 4)
        self.L.dso_oracle_ba_set_nullspaces(self.p, ns.reshape(-1), ns.shape[0])

    def set_solver_state(self, have_first_frame=True, marg_solver_mode=128 | 2048, marg_delta=1e-5):
        self.L.dso_oracle_ba_set_solver_state.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        self.L.dso_oracle_ba_set_solver_state(self.p, int(have_first_frame), int(marg_solver_mode), float(marg_delta))

    def fix_linearization(self, which=None):
        out = np.zeros(self.nR * 8, dtype=np.float32)
        w = None if which is None else np.ascontiguousarray(which, dtype=np.uint8)
        self.L.dso_oracle_ba_fix_linearization(self.p, None if w is None else w.ctypes.data, out.ctypes.data)
        return out.reshape(self.nR, 8)

    def solve(self, iteration, lam, solver_mode=128 | 2048, delta=1e-5, MT=False, nthreads=6):
        N = 8 * self.nF + 4
        x = np.zeros(N)
        step = np.zeros(self.nP, dtype=np.float32)
        HF = np.zeros(N * N)
        bF = np.zeros(N)
        rc = self.L.dso_oracle_ba_solve(self.p, iteration, lam, solver_mode, delta, int(MT), nthreads, x.ctypes.data,
                                        step.ctypes.data, HF.ctypes.data, bF.ctypes.data)
        if rc != 0:
            raise RuntimeError("solver mode not restated")
        return dict(x=x, step=step, HFinal=HF.reshape(N, N), bFinal=bF)

    def calc_L_energy(self, nthreads=1):
        return self.L.dso_oracle_ba_calc_L_energy(self.p, nthreads)

    def calc_M_energy(self):
        return self.L.dso_oracle_ba_calc_M_energy(self.p)

    def marginalize_points(self, flags, prior_marg_fac=600.0 * 600.0, marg_weight_fac=0.25):
        N = 8 * self.nF + 4
        flags = np.ascontiguousarray(flags, dtype=np.uint8)
        H, b = np.zeros(N * N), np.zeros(N)
        n = C.c_int(0)
        ih = np.zeros(self.nP, dtype=np.float32)
        self.L.dso_oracle_ba_marginalize_points(self.p, flags.ctypes.data, prior_marg_fac, marg_weight_fac, H.ctypes.data, b.ctypes.data,
                                                C.addressof(n), ih.ctypes.data)
        HM, bM = np.zeros(N * N), np.zeros(N)
        self.L.dso_oracle_ba_get_marg_prior(self.p, HM, bM)
        return dict(H=H.reshape(N, N), b=b, resInM=n.value, idepth_hessian=ih, HM=HM.reshape(N, N), bM=bM)
