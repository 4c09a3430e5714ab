#!/bin/bash
# round 2, GPU call 2: full GPU test tier, engine A/B with per-hit task records, ncu of the engine
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/c2_pytest.log
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_ENGINE_REFILL=32 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=16 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=8 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=1 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=32 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=8 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=1 B2PT_ENGINE_OCC=8" 2>&1 | tee gpurun_out/c2_ab_engine.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_layered_engine -c 1 -o gpurun_out/c2_k_layered_engine python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c2_ncu_engine.log 2>&1; echo "ncu engine exit $?"
B2PT_ENGINE_REFILL=4 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_layered_engine -c 1 -o gpurun_out/c2_k_layered_engine_r4 python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c2_ncu_engine_r4.log 2>&1; echo "ncu engine r4 exit $?"
for c in c1 c2 c3 c5; do
  timeout 600 python bench.py --config $c --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/c2_bench_$c.json 2> gpurun_out/c2_bench_$c.err; echo "bench $c exit $?"; python -c "
import json; d=json.load(open('gpurun_out/c2_bench_$c.json')); print('$c', round(d['value'],1), d['roofline']['share_of_step'])"
done
