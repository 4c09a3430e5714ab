"""GPU tier: the relaxed-arithmetic build of the Philox kernels (csrc/wavefront_relaxed.cu, the default of rng_mode
"philox") against the exact build (B2PT_RELAXED=0: the arithmetic that reproduces the reference bit for bit in the
LCG modes).

Both builds draw the same Philox numbers, so their images differ only where a rounding difference sends a path
somewhere else.  Per README scene: the distance between the two images must be well below the noise floor at this
sample count (the distance between two exact renders of disjoint sample sets), the means must agree, and ray
counts must agree to a fraction of a percent.  The switch is read once per process: the exact images come from a
subprocess."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
W = H = 160
SPP = 32
SCENES = {
    # name: (scene kwargs, bsdf mode, depth)
    "conductor": (dict(metallic=1.0), 1, 16),
    "diffuse": (dict(metallic=0.0), 2, 8),
    "dielectric_bright": (dict(metallic=0.0, light_color=20.0), 3, 32),
    "auto_half_metal": (dict(metallic=0.5), 0, 8),
}

PROBE = r"""
import json, sys
import numpy as np
sys.path.insert(0, {root!r})
from optixpathtracer_b200 import OptixRenderer, scenes
out = {{}}
for name, (kw, mode, depth) in {scenes!r}.items():
    sc = scenes.spheres_scene(segments=32, rings=16, **kw)
    with OptixRenderer(None, sc, bsdf_mode=mode, rng_mode="philox", max_bounces=depth) as r:
        r.Resize(({w}, {h})); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
        a = r.RenderAccumulate({spp})
        st = r.Stats()
        r.ClearAccumulation()
        b = r.RenderAccumulate({spp})          # the next {spp} samples: an independent estimate
    np.save({tmp!r} + "/" + name + "_a.npy", a); np.save({tmp!r} + "/" + name + "_b.npy", b)
    out[name] = [int(st["radiance_rays"]), int(st["shadow_rays"])]
print("PROBE " + json.dumps(out))
"""


def _run(tmp, relaxed):
    env = {k: v for k, v in os.environ.items() if not k.startswith("B2PT_")}
    env["B2PT_RELAXED"] = "1" if relaxed else "0"
    code = PROBE.format(root=str(ROOT), scenes=SCENES, w=W, h=H, spp=SPP, tmp=str(tmp))
    p = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=900, cwd=str(ROOT))
    assert p.returncode == 0, p.stderr[-2000:]
    rays = json.loads([l for l in p.stdout.splitlines() if l.startswith("PROBE ")][-1][6:])
    imgs = {n: (np.load(tmp / f"{n}_a.npy"), np.load(tmp / f"{n}_b.npy")) for n in SCENES}
    return imgs, rays


@pytest.fixture(scope="module")
def both(tmp_path_factory):
    te, tr = tmp_path_factory.mktemp("exact"), tmp_path_factory.mktemp("relaxed")
    return _run(te, False), _run(tr, True)


@pytest.mark.parametrize("name", sorted(SCENES))
def test_relaxed_kernels_match_the_exact_ones_within_noise(both, name):
    (exact, rays_e), (relaxed, rays_r) = both
    ea, eb = exact[name]
    ra, _ = relaxed[name]
    assert np.isfinite(ra).all()
    disp = lambda a: np.clip(a, 0.0, 1.0)   # display-referred, as the image tier does (fireflies of the bright scene)
    floor = float(np.mean((disp(ea) - disp(eb)) ** 2))          # two exact renders, disjoint samples
    dist = float(np.mean((disp(ea) - disp(ra)) ** 2))           # exact vs relaxed, SAME samples
    assert dist <= 0.25 * floor + 1e-10, f"{name}: relaxed-vs-exact MSE {dist:.3g}, noise floor {floor:.3g}"
    me, mr = float(disp(ea).mean()), float(disp(ra).mean())
    assert abs(mr - me) <= 0.005 * me, f"{name}: mean {mr:.6g} vs {me:.6g}"
    for k in range(2):
        assert abs(rays_r[name][k] - rays_e[name][k]) <= 0.005 * rays_e[name][k], f"{name}: ray counts {rays_r[name]} vs {rays_e[name]}"
