timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v2_pytest.log 2>&1; tail -3 gpurun_out/v2_pytest.log
timeout 600 python bench.py > gpurun_out/v2_bench_c2.json 2> gpurun_out/v2_bench_c2.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/v2_bench_c2.json')); print(d['value'], d['e2e'], d['ms_per_step'], d['roofline']['frac'], d['gpu_launches'], d['cpu_baseline']['value'])"
export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_TRIQ=0" "B2PT_TRIQ=1" "B2PT_SMALL_CHUNKS=0" "B2PT_REFILL=4" "B2PT_REFILL=12" "B2PT_BREAK=12" "B2PT_BREAK=20" "B2PT_OCC=8" "B2PT_OCC=4" 2>&1 | tee gpurun_out/v2_ab.txt
