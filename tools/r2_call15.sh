#!/bin/bash
# round 2, GPU call 15: walk engine with per-lane task prefetch (cp.async into shared memory) against the warp-wide fetch
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_variants_gpu.py tests/test_fast_layered.py -m gpu -q -x > gpurun_out/c15_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/c15_pytest.log
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_PREFETCH=0" "B2PT_ENGINE_PREFETCH=1" "B2PT_ENGINE_PREFETCH=1 B2PT_ENGINE_OCC=5" "B2PT_ENGINE_PREFETCH=1 B2PT_ENGINE_OCC=4" 2>&1 | tee gpurun_out/c15_ab_c4.txt
export AB_ARGS="--config c5 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_PREFETCH=0" "B2PT_ENGINE_PREFETCH=1" 2>&1 | tee gpurun_out/c15_ab_c5.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_layered_engine -c 1 -o gpurun_out/c15_k_layered_engine_pf python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c15_ncu.log 2>&1; echo "ncu exit $?"
