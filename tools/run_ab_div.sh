timeout 900 python -m pytest tests/test_trace_gpu.py tests/test_render_gpu.py tests/test_variants_gpu.py -m gpu -x -q 2>&1 | tail -3
export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_X=1" "B2PT_X=2" 2>&1 | tee gpurun_out/div_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c3" bash tools/ab.sh "B2PT_X=1" 2>&1 | tee -a gpurun_out/div_ab.txt
