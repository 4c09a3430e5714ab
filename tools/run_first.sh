for i in 1 2 3; do
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/first_$i.json 2>gpurun_out/first_$i.err
python - <<PY
import json; d=json.load(open('gpurun_out/first_$i.json')); print($i, round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['step_ms_host'], d['clocks'], round(d['wall_ms'],1), round(d['device_ms_render_max'],1))
PY
done
