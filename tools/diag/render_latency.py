#!/usr/bin/env python
"""Latency of the reference's own call pattern: Render() = one sample per pixel per call, frame downloaded to the host
(OptixRenderer::Render + downloadPixels, OptixRenderer.cpp:617-647), on the README scenes.  Prints ms per frame and frames/s
for a few resolutions; the wave kernels are built for throughput (64 spp per call), this is what the interactive
OptixView loop sees."""
import sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from optixpathtracer_b200 import OptixRenderer, scenes

for name, kw, mode, depth in (("conductor", dict(metallic=1.0), 1, 16), ("layered", dict(metallic=0.0), 4, 8)):
    sc = scenes.spheres_scene(**kw)
    for w, h in ((512, 512), (1024, 1024), (1920, 1080)):
        with OptixRenderer(None, sc, bsdf_mode=mode, rng_mode="philox", max_bounces=depth) as r:
            r.Resize((w, h)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
            for _ in range(5):
                r.Render()
            t0 = time.perf_counter()
            n = 40
            for _ in range(n):
                img = r.Render()
            dt = (time.perf_counter() - t0) / n
            st = r.Stats()
        print(f"{name} {w}x{h} depth {depth}: {dt * 1e3:.2f} ms per Render() ({1 / dt:.0f} frames/s, {w * h / dt / 1e6:.0f} Msamples/s), "
              f"{st['kernel_launches'] / (n + 5):.0f} launches per frame")
