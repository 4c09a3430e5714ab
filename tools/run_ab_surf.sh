export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_SURF_OCC=8" "B2PT_SURF_OCC=10" "B2PT_SURF_OCC=12" "B2PT_SURF_OCC=8" 2>&1 | tee gpurun_out/surf_ab.txt
python - <<'PY'
import json; d=json.load(open('gpurun_out/ab.json')); print('step_ms_host', d['step_ms_host'], d['clocks'])
PY
