"""Next row N4: CoarseInitializer::calcResAndGS (FullSystem/CoarseInitializer.cpp:370-732) on the device vs the oracle, which is
bit-identical to the reference's own translation unit (test_oracle_vs_reference.py::test_initializer_calc_res_and_gs).

Bars: isGood_new, energy_new, maxstep, lastHessian_new and JbBuffer_new bit-exact (8-term sums in pattern order); H, b, H_sc,
b_sc <= 1e-4 relative per entry (FP32 sums in a different order than the reference's Accumulator9 tiers); the returned Vec3f:
energy sum <= 1e-5 relative, alphaEnergy and the count exact."""
import numpy as np
import pytest
from helpers import rel_err_matrix

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("lvl,unit_depth", [(0, False), (1, False), (2, True), (3, False)])
def test_init_calc_res_and_gs(pkg, oracle, window_vga, lvl, unit_depth):
    W = window_vga
    sc = pkg.synthetic.initializer_scenario(W, lvl=lvl, n_points=2500, unit_depth=unit_depth)
    first = oracle.make_images(sc["first"], lvl + 1)[0][lvl]
    new = oracle.make_images(sc["new"], lvl + 1)[0][lvl]
    ro = oracle.init_calc_res_and_gs(W.K, first, new, sc)
    ctx = pkg.Context(0, W.w, W.h, 2)
    ctx.make_images(0, sc["first"], lvl + 1, want_host=False)
    ctx.make_images(1, sc["new"], lvl + 1, want_host=False)
    rg = ctx.init_calc_res_and_gs(0, 1, W.K, sc)
    ctx.close()
    assert np.array_equal(ro["isGood_new"], rg["isGood_new"]), np.nonzero(ro["isGood_new"] != rg["isGood_new"])[0][:10]
    good = ro["isGood_new"] == 1
    assert good.sum() > 100
    assert np.array_equal(_bits(ro["energy_new"]), _bits(rg["energy_new"]))
    for k in ("maxstep", "lastHessian_new"):
        assert np.array_equal(_bits(ro[k][good]), _bits(rg[k][good])), k
    assert np.array_equal(_bits(ro["Jb"][good]), _bits(rg["Jb"][good]))
    for Hk, bk in (("H", "b"), ("Hsc", "bsc")):
        Ho, Hg = ro[Hk].astype(np.float64), rg[Hk].astype(np.float64)
        assert rel_err_matrix(Hg, Ho).max() <= 1e-4, (Hk, rel_err_matrix(Hg, Ho).max())
        scale = np.sqrt(np.abs(np.diag(Ho)) * max(np.abs(ro[bk]).max() ** 2 / np.abs(np.diag(Ho)).max(), 1e-30)) + 1e-30
        assert (np.abs(rg[bk].astype(np.float64) - ro[bk]) <= 1e-4 * np.maximum(np.abs(ro[bk]).max(), scale)).all(), bk
    assert abs(float(rg["res"][0]) - float(ro["res"][0])) <= 1e-5 * abs(float(ro["res"][0]))
    assert rg["res"][1] == ro["res"][1] and rg["res"][2] == ro["res"][2]
