"""CPU checks of the oracle's makeImages restatement (HessianBlocks.cpp:137-211)."""
import numpy as np


def _img(w, h, seed=0):
    rng = np.random.default_rng(seed)
    return rng.uniform(0, 255, (h, w)).astype(np.float32)


def test_box_pyramid_matches_numpy(pkg, oracle):
    img = _img(640, 480)
    d, a = oracle.make_images(img, 5)
    ref = pkg.synthetic.numpy_pyramid(img, 5)
    for l in range(5):
        assert np.array_equal(d[l][:, :, 0], ref[l])


def test_gradients_and_wrap(oracle):
    img = _img(64, 48, 1)
    d, a = oracle.make_images(img, 3)
    for l in range(3):
        I = d[l][:, :, 0]
        h, w = I.shape
        f = I.reshape(-1)
        dx = d[l][:, :, 1].reshape(-1)
        dy = d[l][:, :, 2].reshape(-1)
        idx = np.arange(w, w * (h - 1))
        assert np.array_equal(dx[idx], np.float32(0.5) * (f[idx + 1] - f[idx - 1]))   # includes the row wrap at x=0, w-1
        assert np.array_equal(dy[idx], np.float32(0.5) * (f[idx + w] - f[idx - w]))
        assert np.array_equal(a[l].reshape(-1)[idx], dx[idx] * dx[idx] + dy[idx] * dy[idx])
        assert not dx[:w].any() and not dx[w * (h - 1):].any()     # never written (zero-initialised by the binding)


def test_gamma_weight(oracle):
    img = _img(64, 48, 2)
    B = np.linspace(0, 510, 256).astype(np.float32)      # slope 2 -> gw = 2 -> abs *= 4
    d0, a0 = oracle.make_images(img, 1)
    d1, a1 = oracle.make_images(img, 1, B)
    assert np.array_equal(d0[0], d1[0])
    assert np.allclose(a1[0], a0[0] * 4, rtol=1e-6)


def test_pyr_levels_rule(pkg):
    f = pkg.synthetic.pyr_levels_for
    assert f(640, 480) == 4 and f(1280, 1024) == 6 and f(1920, 1080) == 4 and f(3840, 2160) == 5
