#!/usr/bin/env python
"""How much does K8 lose to the four-way light divergence of a warp's shadow rays?  Same scene, same frame, once with all
lights and once with only the first one (every shadow ray of a warp then aims at the same point): K8 time per traced shadow ray."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench
from optixpathtracer_b200 import OptixRenderer

for cid in ("c2", "c5", "c4"):
    label, spec, w, h, _, depth, mode = bench.CONFIGS[cid]
    sc = bench.make_scene(spec)
    for nl in (len(sc["lights"]), 1):
        lights = sc["lights"][:nl]
        for bounces in (1, depth):
            with OptixRenderer(None, sc, bsdf_mode=mode, rng_mode="philox", max_bounces=bounces) as r:
                r.Resize((w, h)); r.SetCamera(sc["camera"]); r.SetLights(lights)
                r.RenderAccumulate(16, download=False)
                r.ResetStats(); r.SetKernelTiming(True)
                r.RenderAccumulate(16, download=False)
                st = r.Stats()
            ms = st["device_ms_kernel"]
            print(f"{cid} lights {nl} bounces {bounces}: K8 {ms[1]:.2f} ms for {st['shadow_rays_traced'] / 1e6:.1f} M shadow rays = "
                  f"{ms[1] * 1e6 / max(st["shadow_rays_traced"], 1):.1f} ns/ray | K5 {ms[0]:.2f} ms for {st['radiance_rays'] / 1e6:.1f} M rays = "
                  f"{ms[0] * 1e6 / max(st['radiance_rays'], 1):.1f} ns/ray")
