#!/bin/bash
# round 2, GPU call 14: ncu captures of the current build (engine with source view, K5, K8, K6, prep, finish), launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/c14_launches_c4.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c14_ncu_launches.log 2>&1; echo "ncu launches exit $?"
for k in k_layered_engine k_closest_w k_any_w k_surface k_layered_prep k_layered_finish; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/c14_$k python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c14_ncu_$k.log 2>&1; echo "ncu $k exit $?"
done
# bounce-2 launches (incoherent rays) of K5 and K8
for k in k_closest_w k_any_w; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -o gpurun_out/c14_${k}_b2 python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c14_ncu_${k}_b2.log 2>&1; echo "ncu $k b2 exit $?"
done
