#!/usr/bin/env python
"""Image-level parity table for BASELINE.md §4 / DESIGN.md: for each README scene
  ref     = the reference's device code on the host (oracle/_ref), frames 1..N
  ref2    = the same with frames N+1..2N (noise floor of two independent estimates)
  same    = B200, rng lcg_host (the reference's sequence), frames 1..N
  philox  = B200, rng philox, N samples
and MSE / FLIP of each against ref, beside the README row.  A second, GPU-only
part renders lcg_host and philox at the config's full resolution and a larger N
and a larger N and checks that their distance shrinks as 1/N (no bias term).

    python tools/image_parity.py --size 256 --spp 64 --out profiles/r01_image_parity.json
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from optixpathtracer_b200 import OptixRenderer, scenes   # noqa: E402
from tests import image_metrics as im                    # noqa: E402
from tests.util import RefOracle                         # noqa: E402

SCENES = {
    "conductor": (dict(metallic=1.0), 1, 16),
    "diffuse": (dict(metallic=0.0), 2, 8),
    "dielectric_dark": (dict(metallic=0.0), 3, 32),
    "dielectric_bright": (dict(metallic=0.0, light_color=20.0), 3, 32),
    "layered": (dict(metallic=0.0), 4, 8),
}


def gpu(scene, w, h, mode, depth, rng, spp):
    with OptixRenderer(None, scene, bsdf_mode=mode, rng_mode=rng, max_bounces=depth) as r:
        r.Resize((w, h))
        r.SetCamera(scene["camera"])
        r.SetLights(scene["lights"])
        return r.RenderAccumulate(spp)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--full-size", type=int, default=1024)
    ap.add_argument("--full-spp", type=int, default=1024)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    ref = RefOracle()
    rows = {}
    for name, (kw, mode, depth) in SCENES.items():
        sc = scenes.spheres_scene(**kw)
        t0 = time.time()
        ra, _, cnt = ref.render(sc, a.size, a.size, bsdf_mode=mode, max_bounces=depth, frame_first=1, n_frames=a.spp)
        rb, _, _ = ref.render(sc, a.size, a.size, bsdf_mode=mode, max_bounces=depth, frame_first=a.spp + 1,
                              n_frames=a.spp)
        t_ref = time.time() - t0
        same = gpu(sc, a.size, a.size, mode, depth, "lcg_host", a.spp)
        phil = gpu(sc, a.size, a.size, mode, depth, "philox", a.spp)
        env = im.README_ENVELOPE[name]
        row = {
            "size": a.size, "spp": a.spp, "depth": depth, "readme_mse": env[0], "readme_flip": env[1],
            "same_sequence": {"mse": im.mse(ra, same), "flip": im.flip(ra, same),
                              "pixels_bit_equal": float((ra.view(np.uint32) == same.view(np.uint32)).all(axis=2).mean()),
                              "max_abs": float(np.abs(ra - same).max())},
            "noise_floor_ref_vs_ref": {"mse": im.mse(ra, rb), "flip": im.flip(ra, rb)},
            "philox": {"mse": im.mse(ra, phil), "flip": im.flip(ra, phil),
                       "mean_ratio": float(phil.mean() / ra.mean())},
            "reference_host_seconds": round(t_ref, 2), "reference_threads": int(cnt["threads"]),
        }
        # GPU only: two independent unbiased estimators at the config resolution
        fs, fn = a.full_size, a.full_spp
        big_a = gpu(sc, fs, fs, mode, depth, "lcg_host", fn)
        big_b = gpu(sc, fs, fs, mode, depth, "philox", fn)
        half = gpu(sc, fs, fs, mode, depth, "philox", fn // 4)
        row["converged_gpu"] = {
            "size": fs, "spp": fn, "mse_lcg_vs_philox": im.mse(big_a, big_b), "flip_lcg_vs_philox": im.flip(big_a, big_b),
            # MSE(N vs N) = 2 var/N, MSE(N vs N/4) = 5 var/N: the two variance estimates agree when
            # neither estimator is biased (a bias would add a term that does not shrink with N)
            "pixel_variance_from_full": im.mse(big_a, big_b) * fn / 2.0,
            "pixel_variance_from_quarter": im.mse(big_a, half) * fn / 5.0,
            "mean_ratio": float(big_b.mean() / big_a.mean()),
        }
        rows[name] = row
        print(name, json.dumps(row), flush=True)
    if a.out:
        Path(a.out).write_text(json.dumps(rows, indent=1) + "\n")


if __name__ == "__main__":
    main()
