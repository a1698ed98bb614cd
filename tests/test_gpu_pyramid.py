"""K1 parity: dso_b200_make_images vs the oracle's FrameHessian::makeImages (bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZES = [(640, 480, 4), (640, 480, 5), (320, 240, 3), (1280, 1024, 5), (1920, 1080, 4), (424, 320, 3), (96, 64, 2)]


def _image(w, h, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = 127 + 60 * np.sin(xx * 0.05) * np.cos(yy * 0.07) + rng.normal(0, 8, (h, w))
    return np.clip(img, 0, 255).astype(np.float32)


@pytest.mark.parametrize("w,h,levels", SIZES)
def test_make_images_bit_exact(pkg, oracle, w, h, levels):
    ctx = pkg.Context(0, w, h, 2)
    img = _image(w, h, w + h)
    od, oa = oracle.make_images(img, levels)
    gd, ga = ctx.make_images(0, img, levels)
    for l in range(levels):
        assert np.array_equal(od[l].view(np.uint32), gd[l].view(np.uint32)), "dIp level %d" % l
        assert np.array_equal(oa[l].view(np.uint32), ga[l].view(np.uint32)), "absSquaredGrad level %d" % l
    ctx.close()


def test_make_images_gamma_lut_and_nonfinite(pkg, oracle):
    w, h, levels = 640, 480, 4
    ctx = pkg.Context(0, w, h, 2)
    img = _image(w, h, 7)
    img[100, 200] = np.inf          # HessianBlocks.cpp:194-195: non-finite gradients become 0
    img[300, 10] = np.nan
    B = (np.arange(256, dtype=np.float32) ** 1.1).astype(np.float32)
    od, oa = oracle.make_images(img, levels, B)
    gd, ga = ctx.make_images(0, img, levels, B)
    for l in range(levels):
        for o, g in ((od[l], gd[l]), (oa[l], ga[l])):
            fin = np.isfinite(o)
            assert np.array_equal(fin, np.isfinite(g)) and np.array_equal(np.isnan(o), np.isnan(g))
            assert np.array_equal(o[fin].view(np.uint32), g[fin].view(np.uint32))   # NaN payloads may differ
            assert np.array_equal(o[np.isinf(o)], g[np.isinf(o)])
    ctx.close()


def test_make_images_row_wrap_and_borders(pkg):
    """x=0 / x=w-1 use the wrapped neighbours of the adjacent rows; rows 0 and h-1 stay 0."""
    w, h = 64, 32
    ctx = pkg.Context(0, w, h, 1)
    img = np.arange(w * h, dtype=np.float32).reshape(h, w) % 251
    gd, ga = ctx.make_images(0, img, 1)
    f = img.reshape(-1)
    idx = 5 * w            # x = 0 on row 5
    assert gd[0].reshape(-1, 3)[idx, 1] == np.float32(0.5) * (f[idx + 1] - f[idx - 1])
    idx = 5 * w + w - 1    # x = w-1
    assert gd[0].reshape(-1, 3)[idx, 1] == np.float32(0.5) * (f[idx + 1] - f[idx - 1])
    assert not gd[0][0, :, 1:].any() and not gd[0][-1, :, 1:].any()
    assert not ga[0][0].any() and not ga[0][-1].any()
    ctx.close()


def test_make_images_full_size_properties(pkg):
    """3840x2160 (BASELINE configs[3]): level-l intensity equals an independent numpy box pyramid."""
    w, h, levels = 3840, 2160, 5
    ctx = pkg.Context(0, w, h, 1)
    img = _image(w, h, 11)
    gd, ga = ctx.make_images(0, img, levels)
    ref = pkg.synthetic.numpy_pyramid(img, levels)
    for l in range(levels):
        assert np.array_equal(ref[l], gd[l][:, :, 0])
        dx = gd[l][1:-1, 1:-1, 1]
        assert np.array_equal(dx, np.float32(0.5) * (ref[l][1:-1, 2:] - ref[l][1:-1, :-2]))
        assert np.array_equal(ga[l][1:-1], gd[l][1:-1, :, 1] ** 2 + gd[l][1:-1, :, 2] ** 2)
    ctx.close()
