#!/bin/bash
# round 2, GPU call 6: light-ordered shadow rays (K6 -> K8 work list) and the load-first prep / finish kernels:
# full GPU tier, A/B of the work list on c4 / c2 / c5, bench lines, ncu of the changed kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c6_pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/c6_pytest.log
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_SORT_SHADOW=0" "B2PT_SORT_SHADOW=1" 2>&1 | tee gpurun_out/c6_ab_sort_c4.txt
export AB_ARGS="--config c2 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_SORT_SHADOW=0" "B2PT_SORT_SHADOW=1" 2>&1 | tee gpurun_out/c6_ab_sort_c2.txt
export AB_ARGS="--config c5 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_SORT_SHADOW=0" "B2PT_SORT_SHADOW=1" 2>&1 | tee gpurun_out/c6_ab_sort_c5.txt
timeout 900 python bench.py > gpurun_out/c6_bench_c4.json 2> gpurun_out/c6_bench_c4.err; echo "bench c4 exit $?"; tail -3 gpurun_out/c6_bench_c4.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/c6_launches_c4.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c6_ncu_launches.log 2>&1; echo "ncu launches exit $?"
for k in k_any_w k_layered_finish k_layered_prep; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/c6_$k python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c6_ncu_$k.log 2>&1; echo "ncu $k exit $?"
done
