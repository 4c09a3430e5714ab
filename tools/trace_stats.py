#!/usr/bin/env python
"""Per-ray traversal counters of one config (stats build of the kernels): tools/trace_stats.py c2"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench
from optixpathtracer_b200 import OptixRenderer
cid = sys.argv[1] if len(sys.argv) > 1 else "c2"
label, spec, w, h, spp, depth, mode = bench.CONFIGS[cid]
sc = bench.make_scene(spec)
with OptixRenderer(None, sc, bsdf_mode=mode, rng_mode="philox", max_bounces=depth, collect_stats=True) as r:
    r.Resize((w, h)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
    r.RenderAccumulate(2, download=False)
    st = r.Stats()
rr, sr = st["radiance_rays"], max(st["shadow_rays_traced"], 1)
print(cid, "closest: nodes/ray %.2f tris/ray %.2f exact/ray %.3f simd node %.3f tri %.3f | any: nodes/ray %.2f tris/ray %.2f exact/ray %.3f | culled %.3f" % (
    st["nodes_visited"] / rr, st["triangles_tested"] / rr, st["exact_tests"] / rr,
    st["nodes_visited"] / max(st["warp_slots_node"], 1), st["triangles_tested"] / max(st["warp_slots_tri"], 1),
    st["nodes_visited_shadow"] / sr, st["triangles_tested_shadow"] / sr, st["exact_tests_shadow"] / sr,
    1 - st["shadow_rays_traced"] / max(st["shadow_rays"], 1)))
