"""GPU tier: the BASELINE.json configs at their OWN size, depth and light colour.

The other parity tests run on small stand-ins (64x36 frames, 160^2 images); these
run what bench.py runs (bench.CONFIGS), through the C-ABI, against the same two
oracles:

  * C1-C4 (README scenes, 142 854 triangles): frames 1-2 of `Render()` at the
    config's width x height, depth and BSDF mode, with the reference's random
    sequence (lcg_host), against oracle/_ref (the reference's own device code
    on the host).  Untextured scenes: every pixel bit-equal, ray counts equal.
  * C5 (2.2 M-triangle terrain + spheres, AUTO mode): closest hits / occlusion
    of camera, random and bounce-like rays against the fp64 brute force,
    bit-exact, and one 480x270 AUTO-mode frame against oracle/_ref (the deepest
    trees, the 63-bit Morton keys at scale and the mixed material queues).
"""
import numpy as np
import pytest

import bench
from optixpathtracer_b200 import camera_rays, scenes
from tests.util import Oracle, RefOracle, bits

pytestmark = pytest.mark.gpu
f32 = np.float32
MODE_ID = {"auto": 0, "conductor": 1, "diffuse": 2, "dielectric": 3, "layered": 4}


@pytest.fixture(scope="module")
def ref():
    if not RefOracle.available():
        pytest.skip("oracle/_ref not built")
    return RefOracle()


def _frames_gpu(scene, w, h, mode, depth, n_frames):
    from optixpathtracer_b200 import OptixRenderer
    with OptixRenderer(None, scene, bsdf_mode=mode, rng_mode="lcg_host", max_bounces=depth) as r:
        r.Resize((w, h))
        r.SetCamera(scene["camera"])
        r.SetLights(scene["lights"])
        out = np.stack([r.Render() for _ in range(n_frames)])
        return out, r.Stats()


@pytest.mark.parametrize("cid", ["c1", "c2", "c3", "c4"])
def test_readme_config_frames_are_the_reference_frames(ref, cid):
    label, spec, w, h, _, depth, mode = bench.CONFIGS[cid]
    scene = bench.make_scene(spec)
    n = 2
    got, st = _frames_gpu(scene, w, h, mode, depth, n)
    _, want, cnt = ref.render(scene, w, h, bsdf_mode=MODE_ID[mode], max_bounces=depth, frame_first=1, n_frames=n,
                              restated=(mode != "auto"), want_frames=True)
    same = bits(got) == bits(want)
    both_nan = np.isnan(got) & np.isnan(want)
    bad = ~(same | both_nan)
    assert not bad.any(), f"{label}: {bad.any(axis=-1).sum()} of {n * w * h} pixels differ from the reference frames"
    assert st["radiance_rays"] == cnt["radiance_rays"], f"{label}: radiance rays {st['radiance_rays']} vs {cnt['radiance_rays']}"
    assert st["shadow_rays"] == cnt["shadow_rays"], f"{label}: shadow rays {st['shadow_rays']} vs {cnt['shadow_rays']}"


@pytest.fixture(scope="module")
def terrain():
    sc = scenes.terrain_scene()
    soup, _ = scenes.scene_triangle_soup(sc)
    return sc, soup


def _terrain_rays(sc, soup, rng, n_cam=1536, n_rand=1536, n_bounce=1536):
    """Camera rays of C5's own camera (a strided subset of the 3840x2160 frame), random rays through the
    scene's bounds, and bounce-like rays: origins on triangles (offset along the normal as the integrator
    does, devicePrograms.cu:503-506) with cosine-ish directions."""
    w, h = 3840, 2160
    pos, iv, ip = sc["camera"].launch_matrices(w, h)
    o, d = camera_rays(pos, iv, ip, w, h)
    pick = rng.choice(len(o), n_cam, replace=False)
    o1, d1 = o[pick], d[pick]
    lo, hi = soup.reshape(-1, 3).min(0), soup.reshape(-1, 3).max(0)
    o2 = rng.uniform(lo - 0.1, hi + 0.1, (n_rand, 3)).astype(f32)
    d2 = rng.normal(size=(n_rand, 3))
    d2 = (d2 / np.linalg.norm(d2, axis=1, keepdims=True)).astype(f32)
    tri = soup.reshape(-1, 3, 3)[rng.choice(len(soup), n_bounce, replace=False)].astype(np.float64)
    bary = rng.dirichlet((1, 1, 1), n_bounce)
    p = (tri * bary[:, :, None]).sum(1)
    nrm = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-30)
    d3 = rng.normal(size=(n_bounce, 3))
    d3 /= np.linalg.norm(d3, axis=1, keepdims=True)
    d3 = np.where((d3 * nrm).sum(1, keepdims=True) < 0, -d3, d3)
    o3 = (p + 1e-3 * nrm).astype(f32)
    return np.concatenate([o1, o2, o3]), np.concatenate([d1, d2, d3.astype(f32)])


def test_c5_terrain_closest_and_any_hits_match_the_brute_force(terrain):
    from optixpathtracer_b200 import OptixRenderer
    sc, soup = terrain
    rng = np.random.default_rng(55)
    o, d = _terrain_rays(sc, soup, rng)
    assert len(o) >= 4096
    oracle = Oracle()
    with OptixRenderer(None, sc, bsdf_mode="auto") as r:
        st = r.Stats()
        assert st["n_triangles"] == len(soup) > 2_000_000
        gp, gm, gt, gu, gv = r.TraceClosest(o, d)
        op, ot, ou, ov, _ = oracle.closest(soup, o, d)
        assert np.array_equal(gp, op), f"{(gp != op).sum()} primitive ids differ"
        assert np.array_equal(bits(gt), bits(ot)) and np.array_equal(bits(gu), bits(ou)) and np.array_equal(bits(gv), bits(ov))
        assert (op >= 0).mean() > 0.3          # the rays do hit the scene
        # occlusion towards C5's lights from the hit points (LightVisibility, devicePrograms.cu:216-241)
        hit = op >= 0
        p = (o[hit].astype(np.float64) + d[hit].astype(np.float64) * ot[hit, None]).astype(f32)
        lights = np.array([l[0] for l in sc["lights"]], dtype=f32)
        lp = lights[rng.integers(0, len(lights), len(p))]
        ld = lp - p
        tmax = np.linalg.norm(ld, axis=1).astype(f32)
        sd = (ld / np.maximum(tmax[:, None], 1e-20)).astype(f32)
        go = r.TraceAny(p, sd, tmax)
        oo = oracle.any(soup, p, sd, tmax)
        assert np.array_equal(go, oo), f"{(go != oo).sum()} occlusion answers differ"


def test_c5_auto_mode_frame_is_the_reference_frame(ref, terrain):
    sc, _ = terrain
    w, h, depth = 480, 270, 8
    got, st = _frames_gpu(sc, w, h, "auto", depth, 1)
    _, want, cnt = ref.render(sc, w, h, bsdf_mode=0, max_bounces=depth, frame_first=1, n_frames=1, want_frames=True)
    bad = ~((bits(got) == bits(want)) | (np.isnan(got) & np.isnan(want)))
    assert not bad.any(), f"C5 480x270: {bad.any(axis=-1).sum()} of {w * h} pixels differ from the reference frame"
    assert st["radiance_rays"] == cnt["radiance_rays"] and st["shadow_rays"] == cnt["shadow_rays"]
