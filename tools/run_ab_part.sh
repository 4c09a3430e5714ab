timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/part_pytest.log 2>&1; tail -3 gpurun_out/part_pytest.log
for c in c4 c5 c2; do AB_ARGS="--steps 2 --warmup 2 --config $c" bash tools/ab.sh "B2PT_PARTITION=0" "B2PT_PARTITION=1" 2>&1 | tee -a gpurun_out/part_ab.txt; done
