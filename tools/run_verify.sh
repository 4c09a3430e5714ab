set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s18_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s18_pytest.log
timeout 300 python bench.py --steps 8 --warmup 3 > gpurun_out/s18_bench_c2.json 2> gpurun_out/s18_bench_c2.err; echo "bench exit $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_closest -s 4 -c 2 -o gpurun_out/s18_k_closest_wide python bench.py --steps 1 --warmup 1 --spp-per-step 4 --no-cpu-baseline > gpurun_out/s18_ncu1.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_any -s 4 -c 2 -o gpurun_out/s18_k_any_wide python bench.py --steps 1 --warmup 1 --spp-per-step 4 --no-cpu-baseline > gpurun_out/s18_ncu2.log 2>&1
tail -3 gpurun_out/s18_pytest.log
