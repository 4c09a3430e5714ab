#!/bin/bash
# round 2, GPU call 17: triangle records prefetched into L1 when a leaf is parked (B2PT_TRI_PREFETCH)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_trace_gpu.py tests/test_variants_gpu.py tests/test_render_gpu.py -m gpu -q -x > gpurun_out/c17_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c17_pytest.log
for c in c2 c4 c5; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_TRI_PREFETCH=0" "B2PT_TRI_PREFETCH=1" 2>&1 | tee gpurun_out/c17_ab_$c.txt
done
