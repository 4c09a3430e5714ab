timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v3_pytest.log 2>&1; tail -2 gpurun_out/v3_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/v3_bench_c2.json 2> gpurun_out/v3_bench_c2.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/v3_bench_c2.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['traffic'], d['clocks'], d['wall_ms'], d['device_ms_render_max'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/v3_bench_c2_n2.json 2> gpurun_out/v3_bench_n2.err; echo "n2 exit $?"; tail -c 600 gpurun_out/v3_bench_c2_n2.json | head -c 300
