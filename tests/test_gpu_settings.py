"""dso_b200_set_settings: the reference reads setting_huberTH, setting_outlierTHSumComponent, setting_maxPixSearch, the
setting_trace_* family, setting_minTraceTestRadius, setting_GNItsOnPointActivation and setting_minIdepthH_act
(util/settings.cpp:67, 91, 127, 156-165) as global variables at call time.  With values other than the defaults the GPU
path must still equal the oracle (itself pinned against the reference build with the same values,
test_oracle_vs_reference.py::test_non_default_settings_bit_exact) bit for bit."""
import numpy as np
import pytest
from helpers import TrackScenario, compact_warped
from test_gpu_ba import _setup

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_non_default_settings(pkg, oracle, window_small):
    W = window_small
    syn = pkg.synthetic
    vals = (6.0, 30.0 * 30.0, 0.02, 3, 2, 0.8, 2, 0.2, 1.5, 2.5, 3.0, 50.0)
    so = oracle.Settings(*vals)
    try:
        oracle.set_settings(so)
        # ---- tracker: calcRes masks / warped values with another Huber threshold ----
        scen = TrackScenario(pkg, oracle, W, n_frames=1)
        ctx = pkg.Context(0, W.w, W.h, W.nF + 2)
        sg = ctx.default_settings()
        for (name, _), v in zip(sg._fields_, vals):
            setattr(sg, name, v)
        ctx.set_settings(sg)
        scen.gpu_setup(ctx)
        scen.oracle_set_frame(0)
        a, b = scen.frames[0][2]
        pose = scen.poses(0)[1]
        ro = scen.trk.calc_res(1, pose, a, b, 20.0)
        n_o, bufs_o = scen.trk.get_warped()
        rg = ctx.calc_res(0, 1, 1, pose, a, b, 20.0)
        st_g, wp_g = ctx.calc_res_debug(0, 1, 1, len(scen.clouds[1][0]), pose, a, b, 20.0)
        n_g, bufs_g = compact_warped(st_g, wp_g)
        assert n_g == n_o and rg[1] == ro[1]
        for k in range(8):
            assert np.array_equal(_bits(bufs_o[k]), _bits(bufs_g[k])), k
        # the persistent tracker takes the threshold through its options (default: the context's setting)
        ident = syn.pose12((np.eye(3), np.zeros(3)))
        to = scen.trk.track(ident, scen.a0, scen.b0, scen.levels - 1)
        tg = ctx.track_batched(0, 1, ident, scen.a0, scen.b0, scen.levels - 1)[0]
        assert to.ok == tg.ok
        if list(to.trace[:to.trace_len]) == list(tg.trace[:tg.trace_len]):
            assert np.abs(np.array(to.pose) - np.array(tg.pose)).max() <= 1e-5
        ctx.close()
        # ---- BA: linearize with other Huber / gradient-weight thresholds ----
        B, ctx = _setup(pkg, oracle, W)
        ctx.set_settings(sg)
        lo, lg = B.linearize(), ctx.ba_linearize()
        ok = lo["newState"] != 1
        assert np.array_equal(lo["newState"], lg["newState"])
        assert np.array_equal(_bits(lo["J"][ok]), _bits(lg["J"][ok]))
        assert np.array_equal(_bits(lo["newEnergy"]), _bits(lg["newEnergy"]))
        ctx.close()
        # ---- immature points: traceOn over two frames, optimizeImmaturePoint ----
        sc = syn.immature_scenario(W, host_idx=1, n_points=500, n_frames=2)
        fields = oracle.immature_init(oracle.make_images(sc["host_image"], 1)[0][0], sc["u"], sc["v"])
        st_o = oracle.trace_state(len(sc["u"]))
        ctx = pkg.Context(0, W.w, W.h, W.nF + 2)
        ctx.set_settings(sg)
        ctx.trace_set_points(sc["u"], sc["v"], fields, np.zeros(len(sc["u"]), dtype=np.int32), oracle.trace_state(len(sc["u"])))
        for fr in sc["frames"]:
            dI = oracle.make_images(fr["image"], 1)[0][0]
            ctx.make_images(0, fr["image"], 1, want_host=False)
            oracle.trace_on(dI, sc["u"], sc["v"], fields, st_o, fr["KRKi"], fr["Kt"], fr["aff"])
            g = ctx.trace_new_coarse(0, fr["KRKi"].reshape(1, 9), fr["Kt"].reshape(1, 3), fr["aff"].reshape(1, 2))
            assert np.array_equal(g["status"], st_o["status"])
            for key in ("idepth_min", "idepth_max", "quality", "uv", "interval"):
                assert np.array_equal(_bits(g[key]), _bits(st_o[key])), key
        ac = syn.activation_scenario(W, host_idx=1, n_points=500)
        fields = oracle.immature_init(oracle.make_images(ac["host_image"], 1)[0][0], ac["u"], ac["v"])
        dIs = [oracle.make_images(im, 1)[0][0] for im in ac["images"]]
        ro, io, so_ = oracle.optimize_immature(W.K, dIs, ac["RTll"], ac["tTll"], ac["aff"], ac["u"], ac["v"], fields,
                                               ac["idepth_min"], ac["idepth_max"], min_obs=1)
        for i, im in enumerate(ac["images"]):
            ctx.make_images(i, im, 1, want_host=False)
        rg, ig, sg_ = ctx.optimize_immature(np.arange(len(dIs)), W.K, ac["RTll"], ac["tTll"], ac["aff"], ac["u"], ac["v"], fields,
                                            ac["idepth_min"], ac["idepth_max"], min_obs=1)
        assert np.array_equal(rg, ro) and np.array_equal(_bits(ig[ro == 1]), _bits(io[ro == 1]))
        ctx.close()
    finally:
        oracle.set_settings()
