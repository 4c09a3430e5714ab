"""CPU tier: BSDF rows and whole frames.

1. the golden fixtures (tests/golden/*.npz, written by tools/make_golden.py from
   the reference's own code) pin the oracle: where oracle/_ref/liboracle_ref.so
   is present it must reproduce them bit for bit;
2. the product's BSDF and integrator logic (bsdf.cuh / integrator.cuh, run on the
   host by tests/hostsim with the LCG_HOST policy) must equal the reference BIT
   FOR BIT: f, Sample_f (ok, f, wi, pdf, flags), PDF, RNG state after the call,
   and whole frames incl. ray counts — in all five BSDF modes.
"""
from pathlib import Path

import numpy as np
import pytest

from tests import cases
from tests.util import HostSim, RefOracle, bits

GOLD = Path(__file__).resolve().parent / "golden"
f32 = np.float32


@pytest.fixture(scope="module")
def gold_rows():
    return dict(np.load(GOLD / "bsdf_rows.npz"))


@pytest.fixture(scope="module")
def gold_frames():
    return dict(np.load(GOLD / "frames.npz"))


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype == f32:
        return np.array_equal(bits(a), bits(b))
    return np.array_equal(a, b)


@pytest.mark.skipif(not RefOracle.available(), reason="oracle/_ref not built")
def test_ref_oracle_reproduces_goldens(gold_rows, gold_frames):
    ro = RefOracle()
    g = gold_rows
    for b in (1, 2, 3, 4):
        f, so = ro.bsdf_f(b, g["albedo"], g["roughness"], g["wo"], g["wi"], g["seed"])
        assert _same(f, g[f"f{b}"]) and _same(so, g[f"f{b}_seed"])
        out = ro.bsdf_sample(b, g["albedo"], g["roughness"], g["wo"], g["seed"])
        for got, key in zip(out, ("ok", "f", "wi", "pdf", "flags", "seed")):
            assert _same(got, g[f"s{b}_{key}"]), (b, key)
    sc = cases.frame_scenes()["mixed"]
    _, fr, cnt = ro.render(sc, cases.FRAME_W, cases.FRAME_H, bsdf_mode=0, max_bounces=cases.FRAME_BOUNCES,
                           n_frames=cases.FRAME_N, want_frames=True)
    assert _same(fr, gold_frames["mixed_m0"])
    # AUTO through the run-time-mode restatement == the reference's unmodified closest-hit program
    _, fr2, _ = ro.render(sc, cases.FRAME_W, cases.FRAME_H, bsdf_mode=0, max_bounces=cases.FRAME_BOUNCES,
                          n_frames=cases.FRAME_N, want_frames=True, restated=True)
    assert _same(fr, fr2)


@pytest.mark.parametrize("bsdf", [1, 2, 3, 4])
def test_bsdf_rows_bit_exact(hostsim, gold_rows, bsdf):
    g = gold_rows
    f, so = hostsim.bsdf_f(bsdf, g["albedo"], g["roughness"], g["wo"], g["wi"], g["seed"])
    assert _same(f, g[f"f{bsdf}"]) and _same(so, g[f"f{bsdf}_seed"])
    assert (g[f"f{bsdf}"] != 0).any(axis=1).mean() > 0.3          # the rows are not trivially zero
    out = hostsim.bsdf_sample(bsdf, g["albedo"], g["roughness"], g["wo"], g["seed"])
    for got, key in zip(out, ("ok", "f", "wi", "pdf", "flags", "seed")):
        assert _same(got, g[f"s{bsdf}_{key}"]), key
    if bsdf in (2, 3):
        for re, tr in ((1, 1), (1, 0), (0, 1)):
            p = hostsim.bsdf_pdf(bsdf, g["roughness"], g["wo"], g["wi"], re, tr)
            assert _same(p, g[f"p{bsdf}_{re}{tr}"])


@pytest.mark.parametrize("name", ["plain", "mixed", "textured"])
def test_frames_bit_exact(hostsim, gold_frames, name):
    sc = cases.frame_scenes()[name]
    for mode in range(5):
        fr, cnt = hostsim.render(sc, cases.FRAME_W, cases.FRAME_H, bsdf_mode=mode, rng_mode=2,
                                 max_bounces=cases.FRAME_BOUNCES, n_frames=cases.FRAME_N)
        assert _same(fr, gold_frames[f"{name}_m{mode}"]), (name, mode)
        rays = gold_frames[f"{name}_m{mode}_rays"]
        assert (cnt["radiance_rays"], cnt["shadow_rays"]) == (int(rays[0]), int(rays[1]))
        assert fr.mean() > 1e-4


def test_lcg_device_flavour_differs_only_where_documented(hostsim, gold_rows):
    """LCG_DEVICE = nvcc's evaluation order for vec2(rnd, rnd): the conductor's
    smooth branch draws nothing, so it is identical; rough sampling differs."""
    g = gold_rows
    smooth = g["roughness"] < 0.03
    h = hostsim.bsdf_sample(1, g["albedo"], g["roughness"], g["wo"], g["seed"], lcg_mode=1)
    assert _same(h[2][smooth], g["s1_wi"][smooth])
    assert not _same(h[2][~smooth], g["s1_wi"][~smooth])
    assert _same(h[5], g["s1_seed"])                               # same number of draws


def test_energy_conservation_furnace(hostsim):
    """The reference's own unit tests (UnitTests/SpherGeom_Test.cpp:28-252): mean of
    f*|cos|/pdf over 16384 samples must stay below 1.01 for albedo 1."""
    rng = np.random.default_rng(5)
    n = 16384
    for bsdf in (1, 4):
        for rough in (0.0, 0.5, 1.0):
            for _ in range(3):
                wo = rng.normal(size=3)
                wo[2] = abs(wo[2]) + 0.05
                wo = (wo / np.linalg.norm(wo)).astype(f32)
                seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
                ok, f, wi, pdf, _, _ = hostsim.bsdf_sample(bsdf, (1, 1, 1), rough, np.tile(wo, (n, 1)), seeds)
                w = np.where(ok[:, None] > 0, f * np.abs(wi[:, 2:3]) / np.maximum(pdf[:, None], 1e-30), 0)
                assert w.mean(axis=0).max() < 1.01, (bsdf, rough)


def test_philox_is_counter_based(hostsim):
    """Production RNG: a frame depends only on (seed, pixel, sample index) — rendering
    frames [1,2] together or separately gives the same bits; other seeds differ."""
    sc = cases.frame_scenes()["mixed"]
    a, _ = hostsim.render(sc, 32, 18, bsdf_mode=0, rng_mode=0, max_bounces=4, frame_first=1, n_frames=2)
    b, _ = hostsim.render(sc, 32, 18, bsdf_mode=0, rng_mode=0, max_bounces=4, frame_first=2, n_frames=1)
    c, _ = hostsim.render(sc, 32, 18, bsdf_mode=0, rng_mode=0, seed=99, max_bounces=4, frame_first=2, n_frames=1)
    assert _same(a[1], b[0]) and not _same(b[0], c[0])


def test_hostmath_is_the_host_library():
    hs = HostSim()
    """hostmath.cuh (the sin/cos/exp the shading kernels use) against this host's libm, bit for bit."""
    import ctypes as C
    rng = np.random.default_rng(11)
    sweeps = {
        0: np.concatenate([rng.uniform(-0.2, 6.5, 400000), rng.uniform(-119, 119, 100000), [0.0, 1e-5, 0.785398, 3.14159274]]),
        1: np.concatenate([rng.uniform(-0.2, 6.5, 400000), rng.uniform(-119, 119, 100000), [0.0, 1e-5, 0.785398, 3.14159274]]),
        2: np.concatenate([-rng.uniform(0, 40, 400000), rng.uniform(-110, 90, 100000), [0.0, -np.inf, np.inf, -1e-30, -103.9, -104.1]]),
    }
    for kind, x in sweeps.items():
        x = np.ascontiguousarray(x, dtype=np.float32)
        a, b = np.empty_like(x), np.empty_like(x)
        hs.lib.hostsim_hostmath(C.c_int(kind), C.c_int64(len(x)), x.ctypes.data_as(C.c_void_p),
                                a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p))
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"kind {kind}: {(a.view(np.uint32) != b.view(np.uint32)).sum()} differ"
