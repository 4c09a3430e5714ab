export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_NEAREST_ONLY=0" "B2PT_NEAREST_ONLY=1" 2>&1 | tee gpurun_out/nearest_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c3" bash tools/ab.sh "B2PT_NEAREST_ONLY=0" "B2PT_NEAREST_ONLY=1" 2>&1 | tee -a gpurun_out/nearest_ab.txt
