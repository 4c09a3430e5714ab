export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_OCC=6" "B2PT_OCC=7" "B2PT_OCC=8" "B2PT_OCC=7 B2PT_BREAK=16" 2>&1 | tee gpurun_out/occ7_ab.txt
AB_ARGS="--steps 2 --warmup 3 --config c3" bash tools/ab.sh "B2PT_OCC=6" "B2PT_OCC=7" 2>&1 | tee -a gpurun_out/occ7_ab.txt
