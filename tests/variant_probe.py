"""Runs one build variant of the traversal kernels (selected by environment variables, which the library reads
once per process) on fixed inputs and prints one JSON line: whether the closest hits equal the fp64 brute-force
oracle bit for bit, the occlusion answers, and digests of a 1-spp frame in the reference's random sequence, of
the accumulated image of a small wave-capacity sweep and of the same image with the default capacity.
Used by tests/test_variants_gpu.py (one subprocess per variant)."""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    from optixpathtracer_b200 import OptixRenderer, camera_rays, scenes
    from tests import cases
    from tests.util import Oracle, bits, random_rays

    out = {}
    sc = cases.frame_scenes()["mixed"]
    soup, _ = scenes.scene_triangle_soup(sc)
    w, h = cases.FRAME_W, cases.FRAME_H
    pos, iv, ip = sc["camera"].launch_matrices(w, h)
    o1, d1 = camera_rays(pos, iv, ip, w, h)
    lo, hi = soup.reshape(-1, 3).min(0), soup.reshape(-1, 3).max(0)
    o2, d2 = random_rays(np.random.default_rng(7), 4096, lo - 0.5, hi + 0.5)
    o = np.concatenate([o1, o2]).astype(np.float32)
    d = np.concatenate([d1, d2]).astype(np.float32)
    with OptixRenderer(None, sc, bsdf_mode="auto", rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES) as r:
        r.Resize((w, h)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
        prim, _, t, u, v = r.TraceClosest(o, d)
        op, ot, ou, ov, _ = Oracle().closest(soup, o, d)
        out["closest_equals_oracle"] = bool(np.array_equal(prim, op) and np.array_equal(bits(t), bits(ot)) and
                                            np.array_equal(bits(u), bits(ou)) and np.array_equal(bits(v), bits(ov)))
        occ = r.TraceAny(o, d, np.full(len(o), 3.0, dtype=np.float32))
        out["any"] = hashlib.sha256(np.asarray(occ).tobytes()).hexdigest()
        out["frame"] = hashlib.sha256(r.Render().tobytes()).hexdigest()
        r.SetRng("philox")
        r.SetFrameId(0)
        out["accum_auto"] = hashlib.sha256(r.RenderAccumulate(6).tobytes()).hexdigest()
        st = r.Stats()
        out["node_bytes"] = st["node_bytes"]
    # the alpha cut-out scene: the any-hit predicate runs inside whichever traversal kernel the variant selects
    # (the thread-per-ray and binary-node variants are the first implementation, pinned against the reference's
    # frames in tests/test_render_gpu.py; the others must give their bits)
    tx = cases.frame_scenes()["textured"]
    with OptixRenderer(None, tx, bsdf_mode="auto", rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES) as r:
        r.Resize((w, h)); r.SetCamera(tx["camera"]); r.SetLights(tx["lights"])
        prim, _, t, u, v = r.TraceClosest(o, d)
        out["closest_textured"] = hashlib.sha256(prim.tobytes() + t.tobytes() + u.tobytes() + v.tobytes()).hexdigest()
        occ = r.TraceAny(o, d, np.full(len(o), 3.0, dtype=np.float32))
        out["any_textured"] = hashlib.sha256(np.asarray(occ).tobytes()).hexdigest()
        out["frame_textured"] = hashlib.sha256(r.Render().tobytes()).hexdigest()
    with OptixRenderer(None, sc, bsdf_mode="auto", rng_mode="philox", max_bounces=cases.FRAME_BOUNCES,
                       wave_paths_log2=12) as r:   # many small waves
        r.Resize((w, h)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
        out["accum_small_waves"] = hashlib.sha256(r.RenderAccumulate(6).tobytes()).hexdigest()
    print("VARIANT " + json.dumps(out))


if __name__ == "__main__":
    main()
