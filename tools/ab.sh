#!/bin/bash
# A/B helper (run under gpurun): tools/ab.sh "ENV=.. ENV2=.." ...   one bench run per argument (AB_ARGS = bench flags)
for v in "$@"; do
  env $v python bench.py ${AB_ARGS:---steps 4 --warmup 3} --no-cpu-baseline > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err
  python - "$v" <<'PY'
import json,sys
d=json.load(open("gpurun_out/ab.json")); r=d["roofline"]
print(sys.argv[1], "|", d["config"]["config_id"], round(d["value"],1),"Msamples/s", round(d["mrays_per_s"]),"Mrays/s", {k:round(x,2) for k,x in r.get("ms_per_step_by_class", r["share_of_step"]).items()}, "ms/step by class |", "nodes/ray", round(r.get("traversal", r)["nodes_per_ray"],2), "tris/ray", round(r.get("traversal", r)["tris_per_ray"],2), "culled", round(d["shadow_rays_culled_frac"],3), "mean", round(d["image_mean"],5))
PY
done
