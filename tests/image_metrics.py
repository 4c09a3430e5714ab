"""MSE and LDR-FLIP between two linear-RGB images — the two distances the
reference's README quotes for its renders (README.md:40-48, Images/Diffs.txt;
measured there with pbrt's `imgtool diff`).

Test infrastructure (numpy + scipy on the host), never on the product path.

* mse(): mean squared difference on linear RGB, all channels.
* flip(): LDR-FLIP (Andersson et al. 2020, "FLIP: A Difference Evaluator for
  Alternating Images"), restated from the paper: CSF pre-filter in YCxCz,
  Hunt-adjusted HyAB colour difference re-mapped to [0,1], edge/point feature
  difference, combined as dE_c ** (1 - dE_f); 67 pixels per degree (the paper's
  default: 0.7 m from a 0.7 m wide 3840-pixel monitor).  Inputs are linear RGB;
  they are clamped to [0,1] and encoded with the sRGB OETF the reference's
  display pass uses (source/Renderer/OpenGL/Shader/Final.frag:12-17).
"""
import numpy as np
from scipy import ndimage

PPD = 0.7 * 3840 / 0.7 * np.pi / 180.0   # 67.02 pixels per degree

_RGB2XYZ = np.array([
    [10135552 / 24577794, 8788810 / 24577794, 4435075 / 24577794],
    [2613072 / 12288897, 8788810 / 12288897, 887015 / 12288897],
    [1425312 / 73733382, 8788810 / 73733382, 70074185 / 73733382]], dtype=np.float64)
_XYZ2RGB = np.linalg.inv(_RGB2XYZ)
_WHITE = _RGB2XYZ @ np.ones(3)


def mse(a, b) -> float:
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(((a - b) ** 2).mean())


def srgb_encode(lin):
    lin = np.clip(np.asarray(lin, dtype=np.float64), 0.0, 1.0)
    return np.where(lin <= 0.0031308, 12.92 * lin, 1.055 * np.power(lin, 1.0 / 2.4) - 0.055)


def srgb_decode(s):
    s = np.clip(np.asarray(s, dtype=np.float64), 0.0, 1.0)
    return np.where(s <= 0.04045, s / 12.92, np.power((s + 0.055) / 1.055, 2.4))


def _lin_to_ycxcz(lin):
    xyz = lin @ _RGB2XYZ.T / _WHITE
    y = 116.0 * xyz[..., 1] - 16.0
    cx = 500.0 * (xyz[..., 0] - xyz[..., 1])
    cz = 200.0 * (xyz[..., 1] - xyz[..., 2])
    return np.stack([y, cx, cz], axis=-1)


def _ycxcz_to_lin(ycc):
    y = (ycc[..., 0] + 16.0) / 116.0
    x = y + ycc[..., 1] / 500.0
    z = y - ycc[..., 2] / 200.0
    xyz = np.stack([x, y, z], axis=-1) * _WHITE
    return xyz @ _XYZ2RGB.T


def _lin_to_lab(lin):
    t = lin @ _RGB2XYZ.T / _WHITE
    d = 6.0 / 29.0
    f = np.where(t > d ** 3, np.cbrt(np.maximum(t, 0.0)), t / (3.0 * d * d) + 4.0 / 29.0)
    return np.stack([116.0 * f[..., 1] - 16.0, 500.0 * (f[..., 0] - f[..., 1]), 200.0 * (f[..., 1] - f[..., 2])],
                    axis=-1)


def _hunt(lab):
    out = lab.copy()
    out[..., 1] *= 0.01 * lab[..., 0]
    out[..., 2] *= 0.01 * lab[..., 0]
    return out


def _hyab(a, b):
    d = a - b
    return np.abs(d[..., 0]) + np.sqrt(d[..., 1] ** 2 + d[..., 2] ** 2)


def _csf_kernels(ppd):
    par = {"A": (1.0, 0.0047, 0.0, 1e-5), "RG": (1.0, 0.0053, 0.0, 1e-5), "BY": (34.1, 0.04, 13.5, 0.025)}
    r = int(np.ceil(3.0 * np.sqrt(0.04 / (2.0 * np.pi ** 2)) * ppd))
    ax = np.arange(-r, r + 1) / ppd
    z = ax[None, :] ** 2 + ax[:, None] ** 2
    ks = []
    for ch in ("A", "RG", "BY"):
        a1, b1, a2, b2 = par[ch]
        g = a1 * np.sqrt(np.pi / b1) * np.exp(-np.pi ** 2 * z / b1) + a2 * np.sqrt(np.pi / b2) * np.exp(-np.pi ** 2 * z / b2)
        ks.append(g / g.sum())
    return ks


def _conv(img, k):
    return ndimage.correlate(img, k, mode="nearest")


def _spatial_filter(ycc, ppd):
    ks = _csf_kernels(ppd)
    f = np.stack([_conv(ycc[..., c], ks[c]) for c in range(3)], axis=-1)
    return np.clip(_ycxcz_to_lin(f), 0.0, 1.0)


def _feature_kernels(ppd):
    sd = 0.5 * 0.082 * ppd
    r = int(np.ceil(3.0 * sd))
    ax = np.arange(-r, r + 1, dtype=np.float64)
    x, y = np.meshgrid(ax, ax)
    g = np.exp(-(x * x + y * y) / (2.0 * sd * sd))

    def norm(k):   # positive and negative lobes each sum to one in magnitude
        pos, neg = k[k > 0].sum(), -k[k < 0].sum()
        return np.where(k > 0, k / pos, k / neg)

    edge = norm(-x * g)
    point = norm((x * x / (sd * sd) - 1.0) * g)
    return edge, point


def _features(ycc, ppd):
    ynorm = (ycc[..., 0] + 16.0) / 116.0
    edge, point = _feature_kernels(ppd)
    e = np.hypot(_conv(ynorm, edge), _conv(ynorm, edge.T))
    p = np.hypot(_conv(ynorm, point), _conv(ynorm, point.T))
    return e, p


def flip_map(reference_linear, test_linear, ppd: float = PPD):
    """Per-pixel LDR-FLIP error in [0,1] between two linear-RGB images (H,W,3)."""
    qc, qf, pc, pt = 0.7, 0.5, 0.4, 0.95
    ref = srgb_decode(srgb_encode(reference_linear))   # display-referred: clamp to [0,1]
    tst = srgb_decode(srgb_encode(test_linear))
    yr, yt = _lin_to_ycxcz(ref), _lin_to_ycxcz(tst)
    # colour pipeline
    hr = _hunt(_lin_to_lab(_spatial_filter(yr, ppd)))
    ht = _hunt(_lin_to_lab(_spatial_filter(yt, ppd)))
    de = _hyab(hr, ht) ** qc
    green = _hunt(_lin_to_lab(np.array([[0.0, 1.0, 0.0]])))
    blue = _hunt(_lin_to_lab(np.array([[0.0, 0.0, 1.0]])))
    cmax = float(_hyab(green, blue)[0]) ** qc
    pcc = pc * cmax
    dec = np.where(de < pcc, (pt / pcc) * de, pt + (de - pcc) / (cmax - pcc) * (1.0 - pt))
    # feature pipeline
    er, pr = _features(yr, ppd)
    et, pt_ = _features(yt, ppd)
    def_ = np.maximum(np.abs(er - et), np.abs(pr - pt_))
    def_ = np.clip(def_ / np.sqrt(2.0), 0.0, 1.0) ** qf
    return np.clip(dec, 0.0, 1.0) ** (1.0 - def_)


def flip(reference_linear, test_linear, ppd: float = PPD) -> float:
    """Mean LDR-FLIP error — the single number `imgtool diff --metric FLIP` prints."""
    return float(flip_map(reference_linear, test_linear, ppd).mean())


# README.md:40-48 / Images/Diffs.txt: the reference's distance to pbrt-v4 per scene (MSE, FLIP)
README_ENVELOPE = {
    "conductor": (2.571821e-7, 0.0017468039),
    "diffuse": (4.586374e-6, 0.006595257),
    "dielectric_dark": (1.2437624e-7, 0.00084382464),
    "dielectric_bright": (4.9750368e-5, 0.007553334),
    "layered": (8.8446933e-7, 0.0032398894),
}
