mkdir -p gpurun_out; L=gpurun_out/diag.log
: > $L
date >> $L
timeout 30 nvidia-smi >> $L 2>&1; echo "smi rc=$?" >> $L
date >> $L
timeout 90 python -X faulthandler -c "
import faulthandler,sys,time; faulthandler.dump_traceback_later(60, exit=True)
t=time.time()
import ctypes
lib=ctypes.CDLL('libcudart.so.12') if False else None
import torch; print('torch import', time.time()-t, flush=True)
x=torch.zeros(4,device='cuda'); torch.cuda.synchronize(); print('cuda ok', time.time()-t, flush=True)
" >> $L 2>&1; echo "torch rc=$?" >> $L
date >> $L
timeout 120 python -X faulthandler -c "
import faulthandler,sys,time; faulthandler.dump_traceback_later(90, exit=True)
t=time.time()
from tests import cases
from optixpathtracer_b200 import OptixRenderer
print('import', time.time()-t, flush=True)
sc=cases.small_spheres(); print('scene', time.time()-t, flush=True)
r=OptixRenderer(None, sc); print('create', time.time()-t, flush=True)
print(r.Stats()); print('stats', time.time()-t, flush=True)
" >> $L 2>&1; echo "b2pt rc=$?" >> $L
date >> $L
cat $L
