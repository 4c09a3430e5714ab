"""Per-scene / per-mode pixel error of GPU frames against the golden frames (diagnostic)."""
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests import cases  # noqa: E402
from optixpathtracer_b200 import OptixRenderer  # noqa: E402

gold = dict(np.load(ROOT / "tests" / "golden" / "frames.npz"))
for name, sc in cases.frame_scenes().items():
    for mode in range(5):
        with OptixRenderer(None, sc, bsdf_mode=mode, rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES) as r:
            r.Resize((cases.FRAME_W, cases.FRAME_H)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
            for k in range(cases.FRAME_N):
                img = r.Render()
                ref = gold[f"{name}_m{mode}"][k]
                e = np.abs(img - ref) / (1e-6 + 1e-4 * np.abs(ref))
                px = e.max(axis=2)
                exact = (img.view(np.uint32) == ref.view(np.uint32)).all(axis=2).mean()
                print(f"{name} m{mode} f{k}: bit-exact px {exact:.4f}  px>tol {(px > 1).mean():.4f}  median err/tol {np.median(px):.3g}")
            st = r.Stats()
            print("   rays", st["radiance_rays"], st["shadow_rays"], "gold", gold[f"{name}_m{mode}_rays"])
