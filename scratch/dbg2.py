import sys, torch
sys.path.insert(0, '.')
from ddsp_pytorch_b200 import ops
from oracle import ddsp_oracle as O
def l2(a,b): return float((a.double().cpu()-b).norm()/b.norm())
B,N=4,64000
gen=torch.Generator().manual_seed(N)
xt=(torch.randn(B,N,generator=gen)*0.1).clamp(-1,1); xp=0.5*xt+torch.randn(B,N,generator=gen)*0.05
for n_ffts in ([2048],[512],[64]):
  for alpha,eps in ((1.0,1e-7),(0.0,1e-7),(1.0,1e-2),(1.0,1.0)):
    a=xp.double().requires_grad_(True); O.mss_loss(a,xt.double(),n_ffts,alpha=alpha,eps=eps).backward()
    r=xp.clone().requires_grad_(True); O.mss_loss(r,xt,n_ffts,alpha=alpha,eps=eps).backward()
    b=xp.cuda().requires_grad_(True); ops.mss_loss(b,xt.cuda(),n_ffts,alpha=alpha,eps=eps).backward()
    e=(b.grad.double().cpu()-a.grad)
    # where is the error? per-sample energy near edges vs middle
    print(n_ffts,'alpha',alpha,'eps',eps,'cuda',l2(b.grad,a.grad),'ref32',l2(r.grad,a.grad),'err edge frac',float(e[:,:4096].norm()**2/e.norm()**2), 'maxerr idx', int(e.abs().argmax())%N)
