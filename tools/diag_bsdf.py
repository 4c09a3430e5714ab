"""Error distribution of the GPU BSDF hooks against the golden rows (diagnostic)."""
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests.test_render_gpu import gpu_bsdf_f, gpu_bsdf_sample  # noqa: E402

g = dict(np.load(ROOT / "tests" / "golden" / "bsdf_rows.npz"))


def rel(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-6 / 1e-5 * 1e-5)


for b in (1, 2, 3, 4):
    f, so = gpu_bsdf_f(b, g)
    e = (np.abs(f - g[f"f{b}"]) / (1e-6 + 1e-5 * np.abs(g[f"f{b}"]))).max(1)
    print(f"f{b}: rows>tol {(e > 1).sum()}  worst x tol {e.max():.2f}  seeds differ {(so != g[f'f{b}_seed']).sum()}")
    ok, sf, wi, pdf, fl, so = gpu_bsdf_sample(b, g)
    m = (ok == g[f"s{b}_ok"]) & (fl == g[f"s{b}_flags"]) & (so == g[f"s{b}_seed"])
    print(f"s{b}: branch flips {(~m).sum()}")
    with np.errstate(all="ignore"):
        wg = sf * np.abs(wi[:, 2:3]) / pdf[:, None]
        wr = g[f"s{b}_f"] * np.abs(g[f"s{b}_wi"][:, 2:3]) / g[f"s{b}_pdf"][:, None]
    wg, wr = np.nan_to_num(wg), np.nan_to_num(wr)
    for name, got, ref in (("f", sf, g[f"s{b}_f"]), ("wi", wi, g[f"s{b}_wi"]), ("pdf", pdf[:, None], g[f"s{b}_pdf"][:, None]), ("weight", wg, wr)):
        e = (np.abs(got - ref) / (1e-6 + 1e-5 * np.abs(ref))).max(1)
        e = np.where(m, e, 0)
        bad = np.nonzero(e > 1)[0]
        print(f"   {name}: rows>tol {len(bad)} worst x tol {e.max():.2f}")
        for i in bad[:6]:
            print(f"      row {i} rough {g['roughness'][i]:.2f} wo {g['wo'][i]} got {got[i]} ref {ref[i]}")
