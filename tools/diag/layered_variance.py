#!/usr/bin/env python
"""Per-pixel variance of the layered scene's estimators: K independent renders of SPP samples each at SIZE^2 per
variant (reference sequence = rng lcg_host; Philox through the serial kernel; Philox through the walk engine; the
latter with the exact / relaxed K6), printed as mean / percentiles / max of the per-sample pixel variance and the
image mean.  Diagnostic for profiles/r02_image_parity.json (why the engine's 64-spp image sits above the noise floor).

    python tools/diag/layered_variance.py            # runs every variant in a subprocess (env switches are per process)
"""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
SIZE, SPP, K = 256, 16, 24

VARIANTS = {
    "lcg_host (reference sequence, serial kernel)": ({}, "lcg_host"),
    "philox serial kernel, exact arithmetic": ({"B2PT_LAYERED_ENGINE": "0", "B2PT_RELAXED": "0"}, "philox"),
    "philox walk engine, exact K6": ({"B2PT_RELAXED": "0"}, "philox"),
    "philox walk engine, relaxed K6 (default)": ({}, "philox"),
    "philox walk engine, no NEE culling": ({"B2PT_CULL_NEE": "0"}, "philox"),
}


def probe(rng_mode):
    from optixpathtracer_b200 import OptixRenderer, scenes
    sc = scenes.spheres_scene(segments=32, rings=16, metallic=0.0)
    imgs = []
    with OptixRenderer(None, sc, bsdf_mode=4, rng_mode=rng_mode, max_bounces=8) as r:
        r.Resize((SIZE, SIZE)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
        for k in range(K):
            r.ClearAccumulation()
            r.SetFrameId(1000 * k)
            imgs.append(r.RenderAccumulate(SPP).astype(np.float64))
    a = np.stack(imgs)                       # K, H, W, 3
    var = a.var(axis=0, ddof=1) * SPP        # per-sample variance per pixel and channel
    lum = var.mean(axis=2)
    out = {"mean_var": float(lum.mean()), "p50": float(np.percentile(lum, 50)), "p99": float(np.percentile(lum, 99)),
           "p999": float(np.percentile(lum, 99.9)), "max": float(lum.max()), "image_mean": float(a.mean()),
           "max_sample_mean": float(a.max())}
    print("PROBE " + json.dumps(out))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        probe(sys.argv[1])
        sys.exit(0)
    for name, (env_extra, rng_mode) in VARIANTS.items():
        env = {k: v for k, v in os.environ.items() if not k.startswith("B2PT_") or k == "B2PT_NO_REBUILD"}
        env.update(env_extra)
        p = subprocess.run([sys.executable, __file__, rng_mode], env=env, capture_output=True, text=True, cwd=str(ROOT))
        line = [l for l in p.stdout.splitlines() if l.startswith("PROBE ")]
        print(name, "|", line[-1][6:] if line else p.stderr[-500:])
