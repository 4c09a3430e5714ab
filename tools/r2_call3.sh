#!/bin/bash
# round 2, GPU call 3: structured walk_step (A/B of the refill threshold again), multi-GPU entry points on one GPU
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/c3_pytest.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/c3_pytest.log
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_REFILL=32 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=16 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=8 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=1 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=32 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=8 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=1 B2PT_ENGINE_OCC=8" 2>&1 | tee gpurun_out/c3_ab_engine.txt
B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=8 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_layered_engine -c 1 -o gpurun_out/c3_k_layered_engine_r4 python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c3_ncu_engine_r4.log 2>&1; echo "ncu engine r4 exit $?"
B2PT_ENGINE_REFILL=32 B2PT_ENGINE_OCC=8 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_layered_engine -c 1 -o gpurun_out/c3_k_layered_engine_r32 python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c3_ncu_engine_r32.log 2>&1; echo "ncu engine r32 exit $?"
