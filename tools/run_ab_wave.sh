export B2PT_TRIQ=0
for a in "--wave-log2 22" "--wave-log2 24" "--wave-log2 24 --spp-per-step 64" "--wave-log2 26 --spp-per-step 64"; do
  echo "== $a"
  AB_ARGS="--steps 4 --warmup 3 $a" bash tools/ab.sh "B2PT_SMALL_CHUNKS=1"
done 2>&1 | tee gpurun_out/wave_ab.txt
for c in c1 c3 c4 c5; do
 for a in "--wave-log2 22" "--wave-log2 25 --spp-per-step 32"; do
  echo "== $c $a"
  AB_ARGS="--steps 2 --warmup 3 --config $c $a" bash tools/ab.sh "B2PT_SMALL_CHUNKS=1"
 done
done 2>&1 | tee -a gpurun_out/wave_ab.txt
