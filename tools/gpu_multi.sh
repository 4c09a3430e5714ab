#!/bin/bash
# round 2, multi-GPU call (gpurun --gpus N): the multi-device tests on real peers, then strong scaling of C5 (the north star's
# configs[4]: 3840x2160 split over the GPUs with an NCCL reduce) and of C1 (the smallest frame: where wave tails would show)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l); echo "GPUs: $NG"
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q > gpurun_out/m_pytest_multi.log 2>&1; echo "pytest multi exit $?"; tail -5 gpurun_out/m_pytest_multi.log
run() {  # config scaling n
  local c=$1 sc=$2 n=$3 out=gpurun_out/m_bench_${1}_${2}_n${3}.json
  if [ "$n" = 1 ]; then
    CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --config $c --scaling $sc --steps 4 --warmup 3 --no-cpu-baseline > $out 2> ${out%.json}.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --config $c --scaling $sc --steps 4 --warmup 3 > $out 2> ${out%.json}.err
  fi
  echo "bench $c $sc n=$n exit $?"
  python -c "
import json; d=json.load(open('$out')); print('$c $sc n=$n', round(d['value'],1), 'Msamples/s e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],2), d['config']['spp_per_step_per_gpu'], 'spp/gpu/step')" 2>/dev/null || tail -3 ${out%.json}.err
}
for n in ${NS:-2 4 8}; do
  [ $n -le $NG ] || continue
  run c5 strong $n
  run c1 strong $n
done
for n in ${NS:-2 4 8}; do
  [ $n -le $NG ] || continue
  run c4 weak $n
done
