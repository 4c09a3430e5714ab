"""Radiance / shadow rays per bounce of a bench config (differences of the stats of renders with max_bounces = 1..depth)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import bench
from optixpathtracer_b200 import OptixRenderer

cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
label, spec, w, h, spp, depth, mode = bench.CONFIGS[cfg]
scene = bench.make_scene(spec)
prev = (0, 0)
for k in range(1, depth + 1):
    with OptixRenderer(None, scene, bsdf_mode=mode, rng_mode="philox", max_bounces=k) as r:
        r.Resize((w, h)); r.SetCamera(scene["camera"]); r.SetLights(scene["lights"])
        r.RenderAccumulate(1, download=False)
        st = r.Stats()
    cur = (st["radiance_rays"], st["shadow_rays_traced"])
    print(f"bounce {k:2d}: radiance {cur[0]-prev[0]:9d} ({(cur[0]-prev[0])/(w*h):.4f}/path)  shadow {cur[1]-prev[1]:9d}")
    prev = cur
