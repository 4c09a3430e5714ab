for c in c1 c3 c4 c5; do
  timeout 900 python bench.py --config $c > gpurun_out/r_bench_$c.json 2> gpurun_out/r_bench_$c.err; echo "bench $c exit $?"
done
