#!/bin/bash
# round 2, GPU call 7: block-aggregated light buckets (K6 -> K8), multi-triangle leaves (B2PT_LEAF / bvh_leaf_size), wave 2^27
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c7_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c7_pytest.log
for L in 2 4 8; do
  B2PT_LEAF=$L timeout 1200 python -m pytest tests/test_trace_gpu.py tests/test_render_gpu.py tests/test_configs_gpu.py tests/test_variants_gpu.py -m gpu -q -x > gpurun_out/c7_pytest_leaf$L.log 2>&1; echo "pytest leaf $L exit $?"; tail -3 gpurun_out/c7_pytest_leaf$L.log
done
for c in c4 c2 c5; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_SORT_SHADOW=0" "B2PT_SORT_SHADOW=1" 2>&1 | tee gpurun_out/c7_ab_sort_$c.txt
done
for c in c2 c4 c5; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_LEAF=1" "B2PT_LEAF=2" "B2PT_LEAF=3" "B2PT_LEAF=4" 2>&1 | tee gpurun_out/c7_ab_leaf_$c.txt
done
export AB_ARGS="--config c4 --steps 3 --warmup 2 --wave-log2 27"
bash tools/ab.sh "B2PT_X=1" 2>&1 | tee gpurun_out/c7_ab_wave27_c4.txt
