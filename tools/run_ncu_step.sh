for st in 1 2; do
B2PT_STEP=$st timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_closest -s 1 -c 1 -o gpurun_out/s18_k_closest_step$st python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/s18_ncu_step$st.log 2>&1
tail -2 gpurun_out/s18_ncu_step$st.log
done
