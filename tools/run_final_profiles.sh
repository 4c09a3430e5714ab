set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; tail -2 gpurun_out/f_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for c in c2 c1 c3 c4 c5; do
  timeout 900 python bench.py --config $c > gpurun_out/f_bench_$c.json 2> gpurun_out/f_bench_$c.err; echo "bench $c exit $?"
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/f_bench_ref_c2.json 2> gpurun_out/f_bench_ref.err; echo "ref exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/f_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_closest -s 1 -c 1 -o gpurun_out/f_k_closest_final python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/f_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 1 -c 1 -o gpurun_out/f_k_shade_layered_final python bench.py --config c4 --steps 1 --warmup 0 --spp-per-step 4 --no-cpu-baseline > gpurun_out/f_ncu3.log 2>&1
