"""Traversal micro-benchmark through the C-ABI hooks (host buffers; the device
time of the kernels is read from CUDA-event stats is not available here, so this
reports wall time of the whole hook call incl. copies — a first sanity number)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from optixpathtracer_b200 import OptixRenderer, camera_rays, scenes

sc = scenes.spheres_scene()
w = h = 1024
pos, iv, ip = sc["camera"].launch_matrices(w, h)
o, d = camera_rays(pos, iv, ip, w, h)
rng = np.random.default_rng(0)
o2 = rng.uniform([-0.5, -0.5, -2], [3, 2.5, 2], (len(o), 3)).astype(np.float32)
d2 = rng.normal(size=(len(o), 3)); d2 = (d2 / np.linalg.norm(d2, axis=1, keepdims=True)).astype(np.float32)
for stats in (False, True):
    with OptixRenderer(None, sc, collect_stats=stats) as r:
        print("build ms", r.Stats()["device_ms_build"], "tris", r.Stats()["n_triangles"])
        for name, (oo, dd) in {"primary": (o, d), "random": (o2, d2)}.items():
            r.TraceClosest(oo[:1000], dd[:1000])
            r.ResetStats()
            t0 = time.time(); prim, *_ = r.TraceClosest(oo, dd); dt = time.time() - t0
            st = r.Stats()
            print(name, "stats" if stats else "fast", f"wall {dt*1e3:.1f} ms  hit {np.mean(prim>=0):.3f}",
                  {k: st[k] / len(oo) for k in ("nodes_visited", "triangles_tested", "exact_tests")} if stats else "")
