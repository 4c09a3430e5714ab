#!/bin/bash
# round 2, final-build measurement (calls 16 and 20): the round's final build - whole GPU tier, the driver's bench line (c4) + reference arm, every other config,
# task census of the engine launch the ncu capture shows (bounce 0, 16 spp)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c16_pytest_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -2 gpurun_out/c16_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/c16_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c16_pytest.log
timeout 900 python bench.py > gpurun_out/c16_bench_c4.json 2> gpurun_out/c16_bench_c4.err; echo "bench c4 exit $?"; tail -3 gpurun_out/c16_bench_c4.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c16_bench_ref_c4.json 2> gpurun_out/c16_bench_ref.err; echo "ref exit $?"; cut -c1-300 gpurun_out/c16_bench_ref_c4.json
for c in c1 c2 c3 c5 tex auto; do
  timeout 600 python bench.py --config $c --steps 4 --warmup 3 > gpurun_out/c16_bench_$c.json 2> gpurun_out/c16_bench_$c.err; echo "bench $c exit $?"
done
python - <<'PY'
import json
for c in ("c4", "c1", "c2", "c3", "c5", "tex", "auto"):
    try:
        d = json.load(open(f"gpurun_out/c16_bench_{c}.json")); r = d["roofline"]; k = r.get("traversal", r)
        print(c, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), {a: round(b, 2) for a, b in r["ms_per_step_by_class"].items()},
              "nodes/ray", round(k["nodes_per_ray"], 2), "build ms", round(d["device"]["bvh_build_ms"], 2),
              "wide nodes", d["device"]["bvh_wide_nodes"], "cpu", d.get("cpu_baseline", {}).get("value"), "mrays", round(d["mrays_per_s"]))
    except Exception as e:
        print(c, "unreadable:", e)
PY
python - <<'PY' 2>&1 | tee gpurun_out/c16_engine_tasks.txt
import bench
from optixpathtracer_b200 import OptixRenderer
label, spec, w, h, _, depth, mode = bench.CONFIGS["c4"]
sc = bench.make_scene(spec)
with OptixRenderer(None, sc, bsdf_mode=mode, rng_mode="philox", max_bounces=1) as r:
    r.Resize((w, h)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
    r.RenderAccumulate(16, download=False)
    st = r.Stats()
    print("ENGINE_TASKS bounce0 16spp", {k: st[k] for k in ("walk_f_hits", "walk_s_hits", "radiance_rays", "shadow_rays", "shadow_rays_traced")})
PY
# evidence captures (each after its own command exited 0 without ncu): launch list, one full capture per kernel (bounce 0 of a
# 16-spp wave), the bounce-2 launches of the traversal kernels (incoherent rays); summaries: tools/ncu_summary.py, tools/launch_summary.py
if [ "${B2PT_CAPTURE:-0}" = 1 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/final_launches_c4.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/final_ncu_launches.log 2>&1; echo "ncu launches exit $?"
  for k in k_layered_engine k_closest_w k_any_w k_surface k_layered_prep k_layered_finish k_raygen k_accumulate; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/final_$k python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/final_ncu_$k.log 2>&1; echo "ncu $k exit $?"
  done
  for k in k_closest_w k_any_w; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -o gpurun_out/final_${k}_b2 python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/final_ncu_${k}_b2.log 2>&1; echo "ncu $k b2 exit $?"
  done
  timeout 1500 python tools/image_parity.py --size 256 --spp 64 --out gpurun_out/final_image_parity.json > gpurun_out/final_image_parity.log 2>&1; echo "image parity exit $?"
fi
