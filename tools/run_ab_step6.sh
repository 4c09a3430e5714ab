timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/step6_pytest.log 2>&1; tail -3 gpurun_out/step6_pytest.log
export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_STEP=6" "B2PT_STEP=6" 2>&1 | tee gpurun_out/step6c_ab.txt
for c in c1 c3 c5; do AB_ARGS="--steps 2 --warmup 3 --config $c" bash tools/ab.sh "B2PT_STEP=6" 2>&1 | tee -a gpurun_out/step6c_ab.txt; done
