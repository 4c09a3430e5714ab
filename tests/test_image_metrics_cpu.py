"""tests/image_metrics.py (MSE / LDR-FLIP) against properties the paper states."""
import numpy as np

from . import image_metrics as im


def _img(seed, h=48, w=64):
    return np.random.default_rng(seed).random((h, w, 3))


def test_identical_images_have_zero_distance():
    a = _img(1)
    assert im.mse(a, a) == 0.0
    assert im.flip(a, a) == 0.0


def test_black_against_white_is_in_the_top_band():
    z, o = np.zeros((32, 32, 3)), np.ones((32, 32, 3))
    # |dL| = 100 -> 100**0.7 = 25.1, above pc*cmax = 16.5: mapped into [0.95, 1]
    assert 0.95 < im.flip(z, o) < 1.0
    assert im.flip_map(z, o).max() <= 1.0


def test_flip_is_bounded_symmetric_in_range_and_monotone_in_error():
    a = _img(2)
    n = np.random.default_rng(3).normal(size=a.shape)
    f1, f2, f3 = (im.flip(a, np.clip(a + s * n, 0, 1)) for s in (0.01, 0.05, 0.2))
    assert 0.0 < f1 < f2 < f3 < 1.0
    m = im.flip_map(a, np.clip(a + 0.05 * n, 0, 1))
    assert m.shape == a.shape[:2] and m.min() >= 0.0 and m.max() <= 1.0


def test_colour_extremes_map_to_one():
    # the paper's re-mapping: green against blue is the largest Hunt-adjusted HyAB distance
    g = np.zeros((24, 24, 3)); g[..., 1] = 1.0
    b = np.zeros((24, 24, 3)); b[..., 2] = 1.0
    assert abs(im.flip(g, b) - 1.0) < 1e-6


def test_display_range_and_oetf():
    x = np.linspace(0, 1, 257)
    assert np.allclose(im.srgb_decode(im.srgb_encode(x)), x, atol=1e-12)
    assert abs(im.srgb_encode(0.5) - 0.735357) < 1e-5
    # values above 1 are display-clamped: no FLIP difference between 2 and 5
    assert im.flip(np.full((16, 16, 3), 2.0), np.full((16, 16, 3), 5.0)) == 0.0


def test_filters_are_normalised():
    for k in im._csf_kernels(im.PPD):
        assert abs(k.sum() - 1.0) < 1e-12
    e, p = im._feature_kernels(im.PPD)
    assert abs(e[e > 0].sum() - 1.0) < 1e-12 and abs(e[e < 0].sum() + 1.0) < 1e-12
    assert abs(p[p > 0].sum() - 1.0) < 1e-12 and abs(p[p < 0].sum() + 1.0) < 1e-12
