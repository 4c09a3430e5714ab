#!/usr/bin/env python
"""Write the synthetic scenes of BASELINE.json as glTF 2.0: the README "Spheres" scene
(C1-C4) or the 2 M-triangle terrain (C5).

    python tools/make_scene_gltf.py spheres --metallic 1 out/spheres_conductor.gltf
    python tools/make_scene_gltf.py terrain out/terrain.gltf
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from optixpathtracer_b200 import gltf, scenes   # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("kind", choices=["spheres", "terrain"])
ap.add_argument("out")
ap.add_argument("--metallic", type=float, default=0.0)
ap.add_argument("--light-color", type=float, default=1.0)
a = ap.parse_args()
sc = scenes.spheres_scene(metallic=a.metallic, light_color=a.light_color) if a.kind == "spheres" else scenes.terrain_scene()
Path(a.out).parent.mkdir(parents=True, exist_ok=True)
p = gltf.save_gltf(sc, a.out)
print(p, sum(len(m["indices"]) for m in sc["meshes"]), "triangles in", len(sc["meshes"]), "meshes")
