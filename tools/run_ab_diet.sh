timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/diet_pytest.log 2>&1; tail -3 gpurun_out/diet_pytest.log
export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_DIET=0" "B2PT_DIET=1" "B2PT_DIET=0" "B2PT_DIET=1" "B2PT_DIET=1 B2PT_BREAK=20" "B2PT_DIET=1 B2PT_REFILL=4" "B2PT_DIET=1 B2PT_OCC=8" 2>&1 | tee gpurun_out/diet_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c3" bash tools/ab.sh "B2PT_DIET=0" "B2PT_DIET=1" 2>&1 | tee -a gpurun_out/diet_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c5" bash tools/ab.sh "B2PT_DIET=0" "B2PT_DIET=1" 2>&1 | tee -a gpurun_out/diet_ab.txt
