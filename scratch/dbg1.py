import sys, numpy as np, torch
sys.path.insert(0, '.')
import ddsp_pytorch_b200 as d
from ddsp_pytorch_b200 import ops
from oracle import ddsp_oracle as O
for (B,N,n_ffts) in [(3,5000,[64,128,256]),(1,2050,[4096,2048]),(1,2050,[2048]),(4,64000,[2048,1024,512,256,128,64]),(4,64000,[64]),(4,64000,[2048])]:
    gen=torch.Generator().manual_seed(N)
    xt=(torch.randn(B,N,generator=gen)*0.1).clamp(-1,1); xp=0.5*xt+torch.randn(B,N,generator=gen)*0.05
    a=xp.double().requires_grad_(True); want=O.mss_loss(a,xt.double(),n_ffts); want.backward()
    b=xp.cuda().requires_grad_(True); got=ops.mss_loss(b,xt.cuda(),n_ffts); got.backward()
    e=b.grad.double().cpu()-a.grad
    print(B,N,n_ffts,'loss rel',abs(float(got)-float(want))/float(want),'grad maxabs rel',float(e.abs().max()/a.grad.abs().max()),'L2 rel',float(e.norm()/a.grad.norm()))
g=dict(np.load('tests/golden/ref_path.npz'))
cu=lambda x,gr=False: torch.as_tensor(np.asarray(x),dtype=torch.float32).cuda().requires_grad_(gr)
K,NB,sr,hop=[int(g[k]) for k in ('K','NB','sr','hop')]
conf=d.Config(n_harmonics=K,n_noise_filters=NB,sample_rate=sr,hop_length=hop)
bank,fn,rv=d.OscillatorBank(conf).cuda(),d.FilteredNoise(conf),d.Reverb(conf,float(g['rev_wet']),float(g['rev_decay'])).cuda()
rv.noise.data=cu(g['rev_noise'])
c,a,H=cu(g['c'],True),cu(g['a'],True),cu(g['H'],True)
ctrl={'f0':cu(g['f0']),'c':c,'a':a,'H':H}
dry=bank(ctrl)+fn(ctrl,noise=cu(g['noise'])); dry.retain_grad()
y=rv(dry); y.retain_grad()
loss=d.MSSLoss([int(n) for n in g['n_ffts']])(y,cu(g['target']))
loss.backward()
for nm,t,ref in (('y.grad',y.grad,None),('dry.grad',dry.grad,None),('c',c.grad,g['gc']),('a',a.grad,g['ga']),('H',H.grad,g['gH']),('rn',rv.noise.grad,g['g_rev_noise'])):
    print(nm, float(t.abs().max()), None if ref is None else (float(np.abs(ref).max()), float(np.abs(t.double().cpu().numpy()-ref).max())))
print('gc ref nonzero idx', np.argwhere(np.abs(g['gc'])>0)[:3], 'got nonzero', int((c.grad!=0).sum()), int((np.abs(g['gc'])>0).sum()))
