export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_BREAK=16 B2PT_ANY_UNSORTED=1" 2>&1 | tee gpurun_out/policy2_ab.txt
for c in c1 c3 c5; do AB_ARGS="--steps 2 --warmup 3 --config $c" bash tools/ab.sh "B2PT_BREAK=20" "B2PT_BREAK=16 B2PT_ANY_UNSORTED=1" 2>&1 | tee -a gpurun_out/policy2_ab.txt; done
