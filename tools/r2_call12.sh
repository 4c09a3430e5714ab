#!/bin/bash
# round 2, GPU call 12: the walk engine reserves 128 tasks per atomic
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c12_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c12_pytest.log
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_REFILL=16" "B2PT_ENGINE_REFILL=8" "B2PT_ENGINE_REFILL=4" "B2PT_ENGINE_REFILL=8 B2PT_ENGINE_OCC=8" 2>&1 | tee gpurun_out/c12_ab_c4.txt
export AB_ARGS="--config c5 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_REFILL=16" "B2PT_ENGINE_REFILL=8" 2>&1 | tee gpurun_out/c12_ab_c5.txt
