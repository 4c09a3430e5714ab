#!/bin/bash
# round 2, GPU call 10: finish kernel with every load in one round (partial sums by slot, has-f flag in the task record), queue entries fetched a trip ahead
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c10_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c10_pytest.log
for c in c4 c5 c2; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_X=1" 2>&1 | tee gpurun_out/c10_ab_$c.txt
done
for k in k_layered_finish k_layered_prep; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/c10_$k python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c10_ncu_$k.log 2>&1; echo "ncu $k exit $?"
done
