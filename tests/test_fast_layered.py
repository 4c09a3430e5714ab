"""The layered walk engine of the production (Philox) generator against the reference's own GlossyDiffuse.

csrc/fast_layered.cuh re-organises the reference's stochastic layered BSDF (PBRT/GlossyDiffuse.h:141-367 `f`,
:372-514 `Sample_f`) into independent walk tasks with analytically simplified sample weights and approximate
device arithmetic, so it cannot be compared bit for bit (that is what the LCG modes and bsdf.cuh are for).  It is
the SAME ESTIMATOR, which is checked here the only way a stochastic function can be — by its moments:

  * E[f(wo, wi)] per colour channel over many random streams equals the reference's, within 4.5 combined
    standard errors, over a lattice of roughness (smooth coating .. fully rough), wo (normal .. grazing, both
    sides) and wi (incl. near the mirror direction, where the deterministic entrance reflection dominates:
    there the bar is 5e-5 relative, the fp32 evaluation difference of a peaked GGX lobe);
  * the sample walk: E[w g(wi)] for w = f |cos| / pdf and test functions g in {1, x, y, z, z^2, xy} equals the
    reference's — i.e. the directional distribution of the weighted samples, the success rate included.

CPU tier: the per-lane logic runs on the host (tests/hostsim).  GPU tier (same file, marked): the same rows
through the C-ABI hook b2pt_bsdf_f / b2pt_bsdf_sample with rng PHILOX, i.e. compiled by nvcc with
--use_fast_math exactly as the engine kernel is, and whole renders of the engine against the serial kernel.
"""
import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from tests.util import HostSim, RefOracle

f32 = np.float32
ROOT = Path(__file__).resolve().parent.parent
LAYERED, PHILOX = 4, 0
ALBEDO = np.array((0.8, 0.3, 0.05), f32)
ROUGHNESS = (0.0, 0.02, 0.1, 0.3, 0.6, 1.0)


def unit(v):
    v = np.asarray(v, dtype=np.float64)
    return (v / np.linalg.norm(v)).astype(f32)


WOS = [unit((0.3, 0.1, 0.95)), unit((0.8, -0.2, 0.3)), unit((0.1, 0.9, 0.08)), unit((-0.3, -0.2, -0.8))]
WIS = [unit((-0.3, -0.05, 0.9)), unit((0.5, 0.5, 0.4)), unit((-0.85, 0.2, 0.25))]


@pytest.fixture(scope="module")
def ref():
    if not RefOracle.available():
        pytest.skip("oracle/_ref not built")
    return RefOracle()


def _hostsim_f(albedo, rough, wo, wi, seeds):
    return HostSim().bsdf_f(LAYERED, albedo, rough, wo, wi, seeds, lcg_mode=PHILOX)[0]


def _hostsim_sample(albedo, rough, wo, seeds):
    ok, f, wi, pdf, _, _ = HostSim().bsdf_sample(LAYERED, albedo, rough, wo, seeds, lcg_mode=PHILOX)
    return ok, f, wi, pdf


def _p(a, t=C.c_float):
    return a.ctypes.data_as(C.POINTER(t))


def _gpu_f(albedo, rough, wo, wi, seeds):
    from optixpathtracer_b200 import _lib
    lib = _lib.load()
    n = len(wo)
    a = np.ascontiguousarray(np.broadcast_to(albedo, (n, 3)), f32)
    r = np.ascontiguousarray(np.broadcast_to(f32(rough), (n,)), f32)
    out = np.empty((n, 3), f32)
    rc = lib.b2pt_bsdf_f(0, LAYERED, PHILOX, n, _p(a), _p(r), _p(wo), _p(wi), _p(seeds, C.c_uint32), _p(out), None)
    assert rc == 0
    return out


def _gpu_sample(albedo, rough, wo, seeds):
    from optixpathtracer_b200 import _lib
    lib = _lib.load()
    n = len(wo)
    a = np.ascontiguousarray(np.broadcast_to(albedo, (n, 3)), f32)
    r = np.ascontiguousarray(np.broadcast_to(f32(rough), (n,)), f32)
    ok = np.empty(n, np.uint8); f = np.empty((n, 3), f32); wi = np.empty((n, 3), f32)
    pdf = np.empty(n, f32); fl = np.empty(n, np.uint8)
    rc = lib.b2pt_bsdf_sample(0, LAYERED, PHILOX, n, _p(a), _p(r), _p(wo), _p(seeds, C.c_uint32), _p(ok, C.c_uint8), _p(f),
                              _p(wi), _p(pdf), _p(fl, C.c_uint8), None)
    assert rc == 0
    return ok, f, wi, pdf


def check_f_moments(ref, fast_f, n, roughness=ROUGHNESS, zmax=4.5):
    rng = np.random.default_rng(7)
    for rough in roughness:
        for wo in WOS:
            for wi in WIS:
                wi = -wi if wo[2] < 0 else wi
                WO, WI = np.tile(wo, (n, 1)), np.tile(wi, (n, 1))
                seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
                fr, _ = ref.bsdf_f(LAYERED, ALBEDO, rough, WO, WI, seeds)
                ff = fast_f(ALBEDO, rough, WO, WI, seeds)
                assert np.isfinite(ff).all()
                mr, mf = fr.mean(0, dtype=np.float64), ff.mean(0, dtype=np.float64)
                se = np.sqrt(fr.var(0, dtype=np.float64) / n + ff.var(0, dtype=np.float64) / n)
                tol = zmax * se + 5e-5 * np.abs(mr) + 1e-9
                assert (np.abs(mr - mf) <= tol).all(), f"roughness {rough} wo {wo} wi {wi}: E[f] {mf} vs reference {mr} (se {se})"
    # light on the other side of the surface: exactly zero, as the reference's walks all end at the opaque base
    wo, wi = WOS[0], -WIS[0]
    z = fast_f(ALBEDO, 0.3, np.tile(wo, (64, 1)), np.tile(wi, (64, 1)), np.arange(64, dtype=np.uint32))
    zr, _ = ref.bsdf_f(LAYERED, ALBEDO, 0.3, np.tile(wo, (64, 1)), np.tile(wi, (64, 1)), np.arange(64, dtype=np.uint32))
    assert not z.any() and not zr.any()


def check_sample_moments(ref, fast_sample, n, roughness=ROUGHNESS, zmax=4.5):
    rng = np.random.default_rng(11)
    tests = (lambda w: np.ones(len(w)), lambda w: w[:, 2], lambda w: w[:, 0], lambda w: w[:, 1],
             lambda w: w[:, 2] ** 2, lambda w: w[:, 0] * w[:, 1])
    for rough in roughness:
        for wo in WOS:
            WO = np.tile(wo, (n, 1))
            seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
            ok, f, wi, pdf, _, _ = ref.bsdf_sample(LAYERED, ALBEDO, rough, WO, seeds)
            wr = np.where(ok[:, None] > 0, f.astype(np.float64) * np.abs(wi[:, 2:3]) / np.maximum(pdf[:, None], 1e-30), 0.0)
            ok2, f2, wi2, pdf2 = fast_sample(ALBEDO, rough, WO, seeds)
            wf = np.where(ok2[:, None] > 0, f2.astype(np.float64) * np.abs(wi2[:, 2:3]) / np.maximum(pdf2[:, None], 1e-30), 0.0)
            assert np.isfinite(wf).all()
            assert (np.abs(np.linalg.norm(wi2[ok2 > 0], axis=1) - 1.0) < 1e-4).all()
            assert abs(ok.mean() - ok2.mean()) <= zmax * np.sqrt(0.5 / n)
            for g in tests:
                a, b = wr * g(wi)[:, None], wf * g(wi2)[:, None]
                se = np.sqrt(a.var(0) / n + b.var(0) / n) + 1e-9
                assert (np.abs(a.mean(0) - b.mean(0)) <= zmax * se).all(), f"roughness {rough} wo {wo}: weighted moment differs"


def test_fast_layered_f_has_the_reference_mean_host(ref):
    check_f_moments(ref, _hostsim_f, 12000)


def test_fast_layered_sample_has_the_reference_moments_host(ref):
    check_sample_moments(ref, _hostsim_sample, 60000)


def test_draws_are_uniform_and_distinct():
    """five 24-bit uniforms per Philox block: the bit slicing must not correlate them (checked on the host build)"""
    hs = HostSim()
    n = 200000
    # the sample walk of a smooth coating at normal incidence reflects with probability R(1) = 0.04 exactly when uc < R
    wo = np.tile(unit((0, 0, 1)), (n, 1))
    ok, f, wi, pdf, fl, _ = hs.bsdf_sample(LAYERED, ALBEDO, 0.0, wo, np.arange(n, dtype=np.uint32), lcg_mode=PHILOX)
    mirror = (ok > 0) & (np.abs(wi[:, 2] - 1.0) < 1e-6)
    assert abs(mirror.mean() - 0.04) < 4.5 * np.sqrt(0.04 * 0.96 / n)


# ---------------------------------------------------------------------------------------------------
# GPU tier
# ---------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_fast_layered_f_has_the_reference_mean_gpu(ref):
    check_f_moments(ref, _gpu_f, 40000)


@pytest.mark.gpu
def test_fast_layered_sample_has_the_reference_moments_gpu(ref):
    check_sample_moments(ref, _gpu_sample, 200000)


PROBE = r'''
import sys, numpy as np
sys.path.insert(0, %r)
from optixpathtracer_b200 import OptixRenderer, scenes
sc = scenes.spheres_scene(segments=32, rings=16, metallic=0.0)
with OptixRenderer(None, sc, bsdf_mode="layered", rng_mode="philox", seed=int(sys.argv[2]), max_bounces=8) as r:
    r.Resize((160, 160)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
    a = r.RenderAccumulate(256)
    st = r.Stats()
np.save(sys.argv[1], a)
print("RAYS", st["radiance_rays"], st["shadow_rays"], st["shadow_rays_traced"])
'''


@pytest.mark.gpu
def test_engine_image_equals_the_serial_kernel_image(tmp_path):
    """The same scene rendered by the walk engine and by the serial k_shade<LAYERED> (B2PT_LAYERED_ENGINE=0), both
    with Philox: independent estimates of one integral.  Their difference must be noise — no larger than the
    difference between two SERIAL renders with different seeds (the noise floor at this sample count), by the
    same measures: global mean, per-tile means (median and 99th percentile of the relative difference), MSE of
    the display-referred image.  And the engine's image must not depend on how tasks were handed to lanes."""
    from tests import image_metrics as im
    imgs = {}
    for tag, env, seed in (("engine", {}, 1), ("engine_refill1", {"B2PT_ENGINE_REFILL": "1"}, 1),
                           ("serial_a", {"B2PT_LAYERED_ENGINE": "0"}, 2), ("serial_b", {"B2PT_LAYERED_ENGINE": "0"}, 3)):
        e = {k: v for k, v in os.environ.items() if not k.startswith("B2PT_")}
        e.update(env)
        out = tmp_path / f"{tag}.npy"
        p = subprocess.run([sys.executable, "-c", PROBE % str(ROOT), str(out), str(seed)], env=e, capture_output=True, text=True, timeout=600)
        assert p.returncode == 0, p.stderr[-2000:]
        imgs[tag] = np.load(out).astype(np.float64)
    assert np.array_equal(imgs["engine"], imgs["engine_refill1"]), "the engine's image depends on how tasks were handed to lanes"
    assert np.isfinite(imgs["engine"]).all()

    def measures(a, b):
        t = lambda x: x.reshape(20, 8, 20, 8, 3).mean((1, 3))
        rel = np.abs(t(a) - t(b)) / np.maximum(0.5 * (t(a) + t(b)), 1e-3)
        return (abs(a.mean() - b.mean()) / b.mean(), np.median(rel), np.percentile(rel, 99),
                im.mse(np.clip(a, 0, 1), np.clip(b, 0, 1)))

    floor = measures(imgs["serial_a"], imgs["serial_b"])
    for other in ("serial_a", "serial_b"):
        got = measures(imgs["engine"], imgs[other])
        assert got[0] <= max(3.0 * floor[0], 0.004), ("global mean", got, floor)
        assert got[1] <= 1.5 * floor[1] + 1e-3, ("median tile difference", got, floor)
        assert got[2] <= 2.0 * floor[2] + 1e-2, ("99th percentile tile difference", got, floor)
        assert got[3] <= 1.5 * floor[3] + 1e-7, ("display MSE", got, floor)
