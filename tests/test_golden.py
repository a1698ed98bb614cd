"""Golden vectors (tests/golden/golden_small.npz, made by tests/golden/make_golden.py from the oracle):
CPU: the oracle still reproduces them bit for bit; GPU: the CUDA path meets the parity bars against them."""
import importlib.util
import os
import zlib
import numpy as np
import pytest
from helpers import rel_err_matrix

HERE = os.path.dirname(os.path.abspath(__file__))


def _mk():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(os.path.join(HERE, "golden", "golden_small.npz")))


def test_oracle_reproduces_golden(pkg, oracle, golden):
    out = _mk().compute(pkg, oracle)
    assert np.array_equal(out["image_crc"], golden["image_crc"]), "synthetic generator drifted: regenerate the fixture"
    for k, v in golden.items():
        if v.dtype.kind == "f":
            assert np.array_equal(np.asarray(out[k]), v, equal_nan=True), k
        else:
            assert np.array_equal(out[k], v), k


@pytest.mark.gpu
def test_gpu_against_golden(pkg, oracle, golden):
    m = _mk()
    W, scen = m.build(pkg, oracle)
    crc = np.array([zlib.crc32(np.ascontiguousarray(im).tobytes()) for im in W.images], dtype=np.uint32)
    assert np.array_equal(crc, golden["image_crc"]), "synthetic generator drifted: regenerate the fixture"
    ctx = pkg.Context(0, W.w, W.h, W.nF + 4)
    d, a = ctx.make_images(W.nF + 3, W.images[0], scen.levels)
    for l in range(scen.levels):
        assert zlib.crc32(d[l].tobytes()) == golden["pyr_crc_%d" % l][0] and zlib.crc32(a[l].tobytes()) == golden["pyr_crc_%d" % l][1]
    scen.gpu_setup(ctx)
    aa, bb = golden["eval_aff"]
    for lvl in range(scen.levels):
        r = ctx.calc_res(0, 1, lvl, golden["eval_pose"], aa, bb, 20.0)
        H, b = ctx.calc_gs(0)
        st, _ = ctx.calc_res_debug(0, 1, lvl, int(golden["pc_n"][lvl]), golden["eval_pose"], aa, bb, 20.0)
        assert zlib.crc32(st.tobytes()) == golden["status_crc_%d" % lvl][0]
        g = golden["res_%d" % lvl]
        assert r[1] == g[1] and abs(r[0] - g[0]) <= 2e-5 * abs(g[0])
        assert rel_err_matrix(H, golden["H_%d" % lvl]).max() <= 1e-4
    ident = pkg.synthetic.pose12((np.eye(3), np.zeros(3)))
    tr = ctx.track_batched(0, 1, ident, scen.a0, scen.b0, scen.levels - 1)[0]
    if list(tr.trace[:tr.trace_len]) == list(golden["track_trace"]):
        assert np.abs(np.array(tr.pose) - golden["track_pose"]).max() <= 1e-5
    else:
        assert np.abs(np.array(tr.pose) - golden["track_pose"]).max() <= 2e-3    # documented accept/reject tie
    for i, img in enumerate(W.images):
        ctx.make_images(i, img, 1, want_host=False)
    ctx.ba_set_window(np.arange(W.nF), W.w, W.h, W.K, W.frameEnergyTH)
    gl = oracle.BA(W, [oracle.make_images(img, 1)[0][0] for img in W.images]).glue()
    ctx.ba_set_glue(gl["precalc"], gl["adHost"], gl["adTarget"], gl["adHTdeltaF"], gl["cDeltaF"])
    ctx.ba_set_points(W)
    ctx.ba_set_residuals(W)
    lin = ctx.ba_linearize()
    assert np.array_equal(lin["newState"].astype(np.int8), golden["ba_newState"])
    assert np.array_equal(lin["newEnergy"], golden["ba_energy"])
    assert zlib.crc32(lin["J"][lin["newState"] != 1].tobytes()) == golden["ba_J_crc"][0]
    ctx.ba_apply_res()
    A = ctx.ba_accumulate_A()
    S = ctx.ba_accumulate_SC()
    assert rel_err_matrix(A["H"], golden["ba_HA"]).max() <= 1e-4 and np.abs(A["b"] - golden["ba_bA"]).max() <= 1e-4 * np.abs(golden["ba_bA"]).max()
    assert rel_err_matrix(S["H"], golden["ba_HS"]).max() <= 1e-4 and np.abs(S["b"] - golden["ba_bS"]).max() <= 1e-4 * np.abs(golden["ba_bS"]).max()
    assert np.array_equal(S["HdiF"], golden["ba_HdiF"])
    ctx.close()
