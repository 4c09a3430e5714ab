"""Generate tests/golden/*.npz from the reference's own code.

Runs oracle/_ref/liboracle_ref.so (the reference's device code compiled for the
host by oracle/Makefile from /root/reference — see oracle/ref_harness.cpp) on
seeded inputs and stores inputs + outputs.  The fixtures let the parity tests
run where neither /root/reference nor the prebuilt _ref library exists.

    python tools/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests import cases  # noqa: E402
from tests.util import RefOracle  # noqa: E402


def main():
    ro = RefOracle()
    out = ROOT / "tests" / "golden"
    out.mkdir(exist_ok=True)
    rows = cases.bsdf_rows(seed=7, seeds_per_case=1)
    data = {k: v for k, v in rows.items()}
    for bsdf in (1, 2, 3, 4):
        f, so = ro.bsdf_f(bsdf, rows["albedo"], rows["roughness"], rows["wo"], rows["wi"], rows["seed"])
        data[f"f{bsdf}"], data[f"f{bsdf}_seed"] = f, so
        ok, sf, wi, pdf, flags, so = ro.bsdf_sample(bsdf, rows["albedo"], rows["roughness"], rows["wo"], rows["seed"])
        data[f"s{bsdf}_ok"], data[f"s{bsdf}_f"], data[f"s{bsdf}_wi"] = ok, sf, wi
        data[f"s{bsdf}_pdf"], data[f"s{bsdf}_flags"], data[f"s{bsdf}_seed"] = pdf, flags, so
    for bsdf in (2, 3):
        for re, tr in ((1, 1), (1, 0), (0, 1)):
            data[f"p{bsdf}_{re}{tr}"] = ro.bsdf_pdf(bsdf, rows["roughness"], rows["wo"], rows["wi"], re, tr)
    np.savez_compressed(out / "bsdf_rows.npz", **data)

    frames = {}
    for name, sc in cases.frame_scenes().items():
        for mode in range(5):
            _, fr, cnt = ro.render(sc, cases.FRAME_W, cases.FRAME_H, bsdf_mode=mode, max_bounces=cases.FRAME_BOUNCES,
                                   n_frames=cases.FRAME_N, want_frames=True)
            frames[f"{name}_m{mode}"] = fr
            frames[f"{name}_m{mode}_rays"] = np.array([cnt["radiance_rays"], cnt["shadow_rays"]], dtype=np.uint64)
    np.savez_compressed(out / "frames.npz", **frames)
    print("wrote", [p.name for p in out.glob("*.npz")])


if __name__ == "__main__":
    main()
