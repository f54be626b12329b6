"""Controller tail forward at small row counts (graph replays): the small-R kernel (R <= 255) against the tiled FP32 product it replaced
there (49 us for any R <= 128) and the tensor-core path (from 256 rows)."""
import sys; sys.path.insert(0,'/root/repo')
import torch
from ddsp_pytorch_b200 import ops
dev='cuda'
for R in (1,4,16,17,64,128,129,255,256):
    U,K,NB=512,180,195
    x=torch.randn(R,U,device=dev)
    ps=[torch.randn(K,U,device=dev)*0.05,torch.randn(K,device=dev),torch.randn(1,U,device=dev)*0.05,torch.randn(1,device=dev),torch.randn(NB,U,device=dev)*0.05,torch.randn(NB,device=dev)]
    with torch.no_grad():
        for _ in range(3): ops.controller_heads(x,*ps)
        torch.cuda.synchronize()
        g=torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            ops.controller_heads(x,*ps)
        for _ in range(3): g.replay()
        torch.cuda.synchronize()
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(100): g.replay()
        e1.record(); torch.cuda.synchronize()
        print(R, round(e0.elapsed_time(e1)/100*1e3,1),'us')
