export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_STEP=4" "B2PT_STEP=5" "B2PT_STEP=4" "B2PT_STEP=5" 2>&1 | tee gpurun_out/step5_ab.txt
timeout 300 python -m pytest tests/test_gltf_gpu.py tests/test_render_gpu.py -m gpu -x -q 2>&1 | tail -2
