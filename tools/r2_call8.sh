#!/bin/bash
# round 2, GPU call 8: default leaf size 2, bvh_leaf_size tests, new variants; A/B of the relaxed-arithmetic Philox kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c8_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c8_pytest.log
for c in c2 c1 c3 c5 c4 tex; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_RELAXED=0" "B2PT_RELAXED=1" 2>&1 | tee gpurun_out/c8_ab_relaxed_$c.txt
done
