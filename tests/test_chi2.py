"""Chi-square goodness of fit of the BSDF samplers against their own pdfs.

The reference sketches this test and leaves it commented out (UnitTests/SpherGeom_Test.cpp:258-298: bin sampled
directions on the sphere, compare with the integrated pdf); this finishes it, pbrt-style:

  * the sphere of directions is cut into NZ x NP (cos theta, phi) cells of equal solid angle;
  * N directions are drawn with `*_sample` (Dielectric::Sample_f, LambertDiffuse::Sample_f as restated in
    csrc/bsdf.cuh) for a fixed wo; failed samples are dropped, as the integrator drops them;
  * the expected count of a cell is N x the integral of `*_pdf` (Dielectric::PDF, LambertDiffuse::PDF) over it
    (midpoint rule on a SUB x SUB sub-grid; the integral over the sphere must equal the success rate, which is
    checked first — a pdf that is not the sampler's density fails there already);
  * cells expecting fewer than 5 samples are pooled, Pearson's statistic is referred to chi2(dof) and the p-values
    of all cases of a run are judged together with a Sidak-corrected threshold.

It is run for BOTH generators: the reference's LCG (bit-exact tier's instantiation) and Philox, the production
instantiation (`lcg_mode = PHILOX` of the hooks draws from PhiloxRng keyed (seed, row) exactly as K7 does) — a
sampler is only as good as the uniforms it is fed.  CPU tier: the per-thread code on the host (tests/hostsim);
GPU tier: the same rows through b2pt_bsdf_sample / b2pt_bsdf_pdf.  A deliberately wrong pdf must be rejected
(the test has power).
"""
import ctypes as C

import numpy as np
import pytest
from scipy import stats

from tests.util import HostSim

f32 = np.float32
DIFFUSE, DIELECTRIC = 2, 3
PHILOX, LCG_HOST = 0, 1
NZ, NP, SUB = 24, 48, 32
N_SAMPLES = 400_000
ALPHA = 1e-3          # family-wise false-alarm rate of one test function


def unit(v):
    v = np.asarray(v, dtype=np.float64)
    return (v / np.linalg.norm(v)).astype(f32)


# (bsdf, roughness, wo): Lambert from above (from below: see test_lambert_sampler_ignores_the_side_of_wo); the rough
# dielectric from near-normal, grazing and inside, from a wide lobe (roughness 1 -> alpha 1) to a narrow one
# (roughness 0.3 -> alpha 0.09)
CASES = [
    (DIFFUSE, 0.5, unit((0.3, 0.1, 0.9))),
    (DIFFUSE, 0.5, unit((-0.6, 0.75, 0.28))),
    (DIELECTRIC, 1.0, unit((0.3, 0.1, 0.9))),
    (DIELECTRIC, 0.6, unit((0.3, 0.1, 0.9))),
    (DIELECTRIC, 0.6, unit((0.8, -0.2, 0.3))),
    (DIELECTRIC, 0.6, unit((-0.3, -0.2, -0.8))),
    (DIELECTRIC, 0.3, unit((0.3, 0.1, 0.9))),
    (DIELECTRIC, 0.3, unit((-0.6, 0.3, -0.5))),
]


def _cell_directions():
    z = (np.arange(NZ * SUB) + 0.5) / (NZ * SUB) * 2.0 - 1.0
    ph = (np.arange(NP * SUB) + 0.5) / (NP * SUB) * 2.0 * np.pi
    Z, P = np.meshgrid(z, ph, indexing="ij")
    s = np.sqrt(np.maximum(0.0, 1.0 - Z * Z))
    return np.stack([s * np.cos(P), s * np.sin(P), Z], -1).reshape(-1, 3).astype(f32)


_DIRS = None


def expected_mass(pdf_fn, bsdf, rough, wo):
    """integral of the pdf over each cell: (NZ, NP), summing to the sampler's success rate"""
    global _DIRS
    if _DIRS is None:
        _DIRS = _cell_directions()
    p = pdf_fn(bsdf, rough, np.tile(wo, (len(_DIRS), 1)), _DIRS).astype(np.float64)
    dw = (2.0 / (NZ * SUB)) * (2.0 * np.pi / (NP * SUB))
    return p.reshape(NZ, SUB, NP, SUB).sum((1, 3)) * dw


def chi2_p(sample_fn, pdf_fn, bsdf, rough, wo, seed, n=N_SAMPLES):
    seeds = ((np.arange(n, dtype=np.uint64) * 2654435761 + seed) % (1 << 32)).astype(np.uint32)
    ok, wi = sample_fn(bsdf, rough, np.tile(wo, (n, 1)), seeds)
    wi = wi[ok > 0].astype(np.float64)
    assert (np.abs(np.linalg.norm(wi, axis=1) - 1.0) < 1e-4).all()
    mass = expected_mass(pdf_fn, bsdf, rough, wo)
    rate = len(wi) / n
    # the pdf is the density of the SUCCESSFUL samples: its integral is the success rate
    assert abs(mass.sum() - rate) < 4.5 * np.sqrt(max(rate * (1 - rate), 1e-6) / n) + 2e-3, \
        f"pdf integrates to {mass.sum():.5f}, success rate {rate:.5f} (bsdf {bsdf} roughness {rough} wo {wo})"
    iz = np.minimum(((wi[:, 2] + 1.0) * 0.5 * NZ).astype(np.int64), NZ - 1)
    ip = np.minimum((np.mod(np.arctan2(wi[:, 1], wi[:, 0]), 2.0 * np.pi) / (2.0 * np.pi) * NP).astype(np.int64), NP - 1)
    obs = np.bincount(iz * NP + ip, minlength=NZ * NP).astype(np.float64)
    exp = (mass / mass.sum() * len(wi)).ravel()
    order = np.argsort(exp)                      # pool the sparse cells (expected < 5) into one
    exp, obs = exp[order], obs[order]
    k = int(np.searchsorted(np.cumsum(exp), 5.0)) + 1
    exp = np.concatenate([[exp[:k].sum()], exp[k:]])
    obs = np.concatenate([[obs[:k].sum()], obs[k:]])
    keep = exp >= 5.0
    assert keep.sum() > 20
    stat = ((obs[keep] - exp[keep]) ** 2 / exp[keep]).sum()
    return float(stats.chi2.sf(stat, keep.sum() - 1))


def judge(pvals, what):
    thr = 1.0 - (1.0 - ALPHA) ** (1.0 / len(pvals))     # Sidak
    bad = [(c, p) for c, p in pvals if p < thr]
    assert not bad, f"{what}: sampler and pdf disagree (p < {thr:.2e}): {bad}"


# ---- host tier --------------------------------------------------------------------------------------------
def _host_sample(mode):
    hs = HostSim()

    def fn(bsdf, rough, wo, seeds):
        ok, _, wi, _, _, _ = hs.bsdf_sample(bsdf, (0.8, 0.8, 0.8), rough, wo, seeds, lcg_mode=mode)
        return ok, wi
    return fn


def _host_pdf(bsdf, rough, wo, wi):
    return HostSim().bsdf_pdf(bsdf, rough, wo, wi)


@pytest.mark.parametrize("mode", [LCG_HOST, PHILOX], ids=["lcg_host", "philox"])
def test_samplers_follow_their_pdfs_host(mode):
    fn = _host_sample(mode)
    judge([((b, r, tuple(w)), chi2_p(fn, _host_pdf, b, r, w, seed=17 + i)) for i, (b, r, w) in enumerate(CASES)],
          "host, " + ("Philox" if mode == PHILOX else "LCG"))


def test_lambert_sampler_ignores_the_side_of_wo_as_the_reference_does():
    """A trait of the reference, kept because the port is bit-exact: LambertDiffuse::Sample_f (PBRT/LambertDiffuse.h:
    111-133) always returns a direction with z > 0 — its flip tests `direction.z < 0`, which a cosine-hemisphere
    sample never is; pbrt flips when WO is below — while LambertDiffuse::PDF (:135-141) is non-zero only on wo's side.
    For wo below the surface the sampler and the pdf therefore live on opposite hemispheres.  This is what the
    chi-square test finds when given such a wo; it is asserted here so that a silent "fix" on either side shows."""
    hs = HostSim()
    n = 20000
    wo = unit((-0.5, 0.4, -0.3))
    for mode in (LCG_HOST, PHILOX):
        ok, _, wi, pdf, _, _ = hs.bsdf_sample(DIFFUSE, (0.8, 0.8, 0.8), 0.5, np.tile(wo, (n, 1)), np.arange(n, dtype=np.uint32), lcg_mode=mode)
        assert ok.all() and (wi[:, 2] > 0).all() and (pdf > 0).all()
        assert not hs.bsdf_pdf(DIFFUSE, 0.5, np.tile(wo, (n, 1)), wi).any()
    mass = expected_mass(_host_pdf, DIFFUSE, 0.5, wo)
    assert abs(mass[: NZ // 2].sum() - 1.0) < 1e-3 and mass[NZ // 2:].sum() == 0.0


def test_chi2_rejects_a_wrong_pdf():
    """power: the Lambert sampler against the pdf of a slightly rougher dielectric lobe, and the dielectric sampler
    against the pdf of a neighbouring roughness, must fail"""
    fn = _host_sample(PHILOX)
    b, r, w = CASES[3]

    def shifted(bsdf, rough, wo, wi):
        return HostSim().bsdf_pdf(bsdf, 0.66, wo, wi)      # sampler at 0.6
    try:
        p = chi2_p(fn, shifted, b, r, w, seed=5)
    except AssertionError:
        return                                             # already caught by the normalisation check
    assert p < 1e-6


def test_philox_draws_are_uniform_in_every_slot():
    """the four 24-bit uniforms of a Philox block, as the samplers consume them (u = x >> 8): uniform and pairwise
    uncorrelated.  Seen through the Lambert sampler, whose wi is an invertible map of (u1, u2): wi.z^2 and the
    azimuth are independent U(0,1) variables when the draws are."""
    hs = HostSim()
    n = 1_000_000
    wo = np.tile(unit((0, 0, 1)), (n, 1))
    ok, _, wi, _, _, _ = hs.bsdf_sample(DIFFUSE, (0.8, 0.8, 0.8), 0.5, wo, np.arange(n, dtype=np.uint32), lcg_mode=PHILOX)
    wi = wi[ok > 0].astype(np.float64)
    a = wi[:, 2] ** 2                                             # cosine-weighted: z^2 ~ U(0,1)
    b = np.mod(np.arctan2(wi[:, 1], wi[:, 0]), 2 * np.pi) / (2 * np.pi)
    for v in (a, b):
        h = np.bincount(np.minimum((v * 256).astype(np.int64), 255), minlength=256)
        assert stats.chisquare(h).pvalue > 1e-4
    h2 = np.bincount(np.minimum((a * 32).astype(np.int64), 31) * 32 + np.minimum((b * 32).astype(np.int64), 31), minlength=1024)
    assert stats.chisquare(h2).pvalue > 1e-4
    assert abs(np.corrcoef(a, b)[0, 1]) < 4.5 / np.sqrt(len(a))


# ---- GPU tier ---------------------------------------------------------------------------------------------
def _p(a, t=C.c_float):
    return a.ctypes.data_as(C.POINTER(t))


def _gpu_sample(mode):
    from optixpathtracer_b200 import _lib
    lib = _lib.load()

    def fn(bsdf, rough, wo, seeds):
        n = len(wo)
        a = np.ascontiguousarray(np.broadcast_to(f32(0.8), (n, 3)), f32)
        r = np.ascontiguousarray(np.broadcast_to(f32(rough), (n,)), f32)
        wo = np.ascontiguousarray(wo, f32)
        ok = np.empty(n, np.uint8); f = np.empty((n, 3), f32); wi = np.empty((n, 3), f32)
        pdf = np.empty(n, f32); fl = np.empty(n, np.uint8)
        rc = lib.b2pt_bsdf_sample(0, bsdf, mode, n, _p(a), _p(r), _p(wo), _p(seeds, C.c_uint32), _p(ok, C.c_uint8), _p(f),
                                  _p(wi), _p(pdf), _p(fl, C.c_uint8), None)
        assert rc == 0
        return ok, wi
    return fn


def _gpu_pdf(bsdf, rough, wo, wi):
    from optixpathtracer_b200 import _lib
    lib = _lib.load()
    n = len(wo)
    r = np.ascontiguousarray(np.broadcast_to(f32(rough), (n,)), f32)
    wo = np.ascontiguousarray(wo, f32); wi = np.ascontiguousarray(wi, f32)
    out = np.empty(n, f32)
    assert lib.b2pt_bsdf_pdf(0, bsdf, n, _p(r), _p(wo), _p(wi), 1, 1, _p(out)) == 0
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [LCG_HOST, PHILOX], ids=["lcg_host", "philox"])
def test_samplers_follow_their_pdfs_gpu(mode):
    fn = _gpu_sample(mode)
    judge([((b, r, tuple(w)), chi2_p(fn, _gpu_pdf, b, r, w, seed=91 + i)) for i, (b, r, w) in enumerate(CASES)],
          "B200, " + ("Philox" if mode == PHILOX else "LCG"))
