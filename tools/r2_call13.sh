#!/bin/bash
# round 2, GPU call 13: K6 / K7 / K7d-p / K7d-f as 256-thread blocks (one queue atomic per 8 warps)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c13_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c13_pytest.log
for c in c4 c2 c5 c1; do
  export AB_ARGS="--config $c --steps 3 --warmup 2"
  bash tools/ab.sh "B2PT_X=1" 2>&1 | tee gpurun_out/c13_ab_$c.txt
done
