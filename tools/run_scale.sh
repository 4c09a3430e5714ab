for n in 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 4 --warmup 3 > gpurun_out/scale_c2_n$n.json 2> gpurun_out/scale_n$n.err; echo "n$n exit $?"
python -c "
import json; d=json.load(open('gpurun_out/scale_c2_n$n.json')); print(d['n_gpus'], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['ms_per_step'], d['step_ms_host'])"
done
