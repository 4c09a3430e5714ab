"""Per-kernel totals and shares of an ncu launch list (--metrics gpu__time_duration.sum --csv)."""
import csv, re, sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]
kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[kn])
    name = re.sub(r"^void ", "", name).replace("b2pt::", "").replace("(anonymous namespace)::", "")[:48]
    v = float(r[mv].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[mu], 1e-6)
    tot[name] += v
    cnt[name] += 1
s = sum(tot.values())
for k in sorted(tot, key=tot.get, reverse=True):
    print(f"{k:48s} launches {cnt[k]:4d}  total {tot[k]:9.3f} ms  share {tot[k] / s:6.3f}")
