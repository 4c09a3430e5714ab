#!/bin/bash
# round 2, GPU call 4: compacted BVH + stack guard tests, new bench.py line (c4) and reference arm, all configs, evidence captures
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/c4_pytest.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/c4_pytest.log
timeout 900 python bench.py > gpurun_out/c4_bench_c4.json 2> gpurun_out/c4_bench_c4.err; echo "bench c4 exit $?"; tail -3 gpurun_out/c4_bench_c4.err; python -c "
import json; d=json.load(open('gpurun_out/c4_bench_c4.json')); print('c4', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['share_of_step'], 'frac', d['roofline']['frac'], d['device'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c4_bench_ref_c4.json 2> gpurun_out/c4_bench_ref.err; echo "ref exit $?"; cut -c1-300 gpurun_out/c4_bench_ref_c4.json
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_REFILL=16 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=16 B2PT_ENGINE_OCC=6" 2>&1 | tee gpurun_out/c4_ab_engine.txt
for c in c1 c2 c3 c5; do
  timeout 600 python bench.py --config $c --steps 4 --warmup 3 > gpurun_out/c4_bench_$c.json 2> gpurun_out/c4_bench_$c.err; echo "bench $c exit $?"; python -c "
import json; d=json.load(open('gpurun_out/c4_bench_$c.json')); r=d['roofline']; print('$c', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), r['share_of_step'], 'K5 ms/launch', r.get('ms_per_launch'), 'nodes/ray', r.get('nodes_per_ray'), d['device'], 'cpu', d['cpu_baseline']['value'])"
done
python - <<'PY' 2>&1 | tee gpurun_out/c4_engine_tasks.txt
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
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/c4_launches_c4.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c4_ncu_launches.log 2>&1; echo "ncu launches exit $?"
for k in k_accumulate k_raygen k_any_w k_closest_w k_layered_finish k_layered_prep k_surface; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/c4_$k python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c4_ncu_$k.log 2>&1; echo "ncu $k exit $?"
done
