timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v4_pytest.log 2>&1; tail -2 gpurun_out/v4_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/v4_bench_c2.json 2> gpurun_out/v4_bench_c2.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/v4_bench_c2.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['traffic'], d['clocks'], d['step_ms_host'])"
