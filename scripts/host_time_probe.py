import sys, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import torch, helpers, nntoolkitcore_b200 as nn
from nntoolkitcore_b200 import layers as L
lib=nn.lib(); lib.NNTKInit(0); lib.NNTKSetStream(torch.cuda.current_stream().cuda_stream); lib.NNTKSetMathMode(1)
B,T,H=4096,500,512
x=torch.randn(B,T,H,device='cuda'); y=torch.empty(B,T,H,device='cuda')
l=L.Recurrent(lib,'lstm',H,H,T); helpers.fill_recurrent(l,1); L.freeze(lib,l)
l.apply_batch(x,y); torch.cuda.synchronize()
for _ in range(2):
    t0=time.perf_counter(); l.apply_batch(x,y); t1=time.perf_counter(); torch.cuda.synchronize(); t2=time.perf_counter()
    print(f"host enqueue {1e3*(t1-t0):.1f} ms, total {1e3*(t2-t0):.1f} ms")
