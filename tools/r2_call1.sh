#!/bin/bash
# round 2, GPU call 1: the new parity tiers, the layered walk engine A/B, its ncu capture, sanitizer, C5 traffic
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader; nproc
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/c1_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
export AB_ARGS="--config c4 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_LAYERED_ENGINE=0" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=1 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=8 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=16 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=32 B2PT_ENGINE_OCC=6" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=8" "B2PT_ENGINE_REFILL=4 B2PT_ENGINE_OCC=4" 2>&1 | tee gpurun_out/c1_ab_engine.txt
export AB_ARGS="--config c5 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_LAYERED_ENGINE=0" "B2PT_LAYERED_ENGINE=1" 2>&1 | tee gpurun_out/c1_ab_engine_c5.txt
timeout 900 python bench.py > gpurun_out/c1_bench_c4.json 2> gpurun_out/c1_bench_c4.err; echo "bench c4 exit $?"; cut -c1-600 gpurun_out/c1_bench_c4.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c1_bench_ref_c4.json 2> gpurun_out/c1_bench_ref.err; echo "ref exit $?"; cut -c1-400 gpurun_out/c1_bench_ref_c4.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/c1_launches_c4.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c1_ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_layered_engine -c 1 -o gpurun_out/c1_k_layered_engine python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c1_ncu_engine.log 2>&1; echo "ncu engine exit $?"
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"k_closest|k_any" -c 16 --csv --log-file gpurun_out/c1_traffic_c5.csv python bench.py --config c5 --steps 1 --warmup 0 --spp-per-step 8 --no-cpu-baseline > gpurun_out/c1_ncu_c5.log 2>&1; echo "ncu c5 exit $?"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c1_memcheck.log 2>&1; echo "memcheck exit $?"; tail -4 gpurun_out/c1_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c1_racecheck.log 2>&1; echo "racecheck exit $?"; tail -4 gpurun_out/c1_racecheck.log
