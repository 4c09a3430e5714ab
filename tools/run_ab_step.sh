B2PT_STEP=4 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/step_pytest.log 2>&1; tail -3 gpurun_out/step_pytest.log
export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_STEP=1" "B2PT_STEP=4" "B2PT_STEP=4 B2PT_OCC=8" "B2PT_STEP=4 B2PT_BREAK=20" "B2PT_STEP=1" "B2PT_STEP=4" 2>&1 | tee gpurun_out/step4_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c3" bash tools/ab.sh "B2PT_STEP=1" "B2PT_STEP=4" 2>&1 | tee -a gpurun_out/step4_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c5" bash tools/ab.sh "B2PT_STEP=1" "B2PT_STEP=4" 2>&1 | tee -a gpurun_out/step4_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c1" bash tools/ab.sh "B2PT_STEP=1" "B2PT_STEP=4" 2>&1 | tee -a gpurun_out/step4_ab.txt
