timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/chunks_pytest.log 2>&1; tail -3 gpurun_out/chunks_pytest.log
export AB_ARGS="--steps 4 --warmup 3"
bash tools/ab.sh "B2PT_TRIQ=0 B2PT_SMALL_CHUNKS=0" "B2PT_TRIQ=0 B2PT_SMALL_CHUNKS=1" "B2PT_TRIQ=1 B2PT_SMALL_CHUNKS=1" 2>&1 | tee gpurun_out/chunks_ab.txt
AB_ARGS="--steps 3 --warmup 3 --config c3" bash tools/ab.sh "B2PT_TRIQ=0 B2PT_SMALL_CHUNKS=0" "B2PT_TRIQ=0 B2PT_SMALL_CHUNKS=1" 2>&1 | tee -a gpurun_out/chunks_ab.txt
AB_ARGS="--steps 3 --warmup 3 --config c1" bash tools/ab.sh "B2PT_TRIQ=0 B2PT_SMALL_CHUNKS=0" "B2PT_TRIQ=0 B2PT_SMALL_CHUNKS=1" 2>&1 | tee -a gpurun_out/chunks_ab.txt
