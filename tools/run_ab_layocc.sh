for v in "B2PT_SHADE_OCC=8" "B2PT_SHADE_OCC=10" "B2PT_SHADE_OCC=12"; do AB_ARGS="--steps 1 --warmup 1 --config c4 --spp-per-step 16" bash tools/ab.sh "$v"; done 2>&1 | tee gpurun_out/layocc_ab2.txt
