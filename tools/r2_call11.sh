#!/bin/bash
# round 2, GPU call 11: block-aggregated queue appends (one atomic per block and trip instead of one per warp) in K6, K7, K7d-p, K7d-f
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c11_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c11_pytest.log
for c in c4 c2 c5 c1; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_X=1" 2>&1 | tee gpurun_out/c11_ab_$c.txt
done
