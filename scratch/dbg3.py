import sys, torch
sys.path.insert(0, '.')
from ddsp_pytorch_b200 import ops
from oracle import ddsp_oracle as O
def l2(a,b): return float((a.double().cpu()-b).norm()/b.norm())
for (B,N) in ((4,64000),(1,8192)):
  gen=torch.Generator().manual_seed(N)
  xt=(torch.randn(B,N,generator=gen)*0.1).clamp(-1,1); xp=0.5*xt+torch.randn(B,N,generator=gen)*0.05
  for n in (256,512,1024,2048,4096):
    a=xp.double().requires_grad_(True); O.mss_loss(a,xt.double(),[n]).backward()
    r=xp.clone().requires_grad_(True); O.mss_loss(r,xt,[n]).backward()
    b=xp.cuda().requires_grad_(True); ops.mss_loss(b,xt.cuda(),[n],windows=[torch.hann_window(n)]).backward()
    b2=xp.cuda().requires_grad_(True); ops.mss_loss(b2,xt.cuda(),[n]).backward()
    print(B,N,n,'cuda',l2(b.grad,a.grad),'ref32',l2(r.grad,a.grad),'run-to-run',float((b.grad-b2.grad).norm()/b.grad.norm()))
