export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_SHADE_OCC3=6" "B2PT_SHADE_OCC3=8" "B2PT_SHADE_OCC3=6" "B2PT_SHADE_OCC3=8" 2>&1 | tee gpurun_out/shade_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c3" bash tools/ab.sh "B2PT_SHADE_OCC3=6" "B2PT_SHADE_OCC3=8" 2>&1 | tee -a gpurun_out/shade_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c1" bash tools/ab.sh "B2PT_SHADE_OCC3=6" "B2PT_SHADE_OCC3=8" 2>&1 | tee -a gpurun_out/shade_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c5" bash tools/ab.sh "B2PT_SHADE_OCC3=6" "B2PT_SHADE_OCC3=8" 2>&1 | tee -a gpurun_out/shade_ab.txt
