import sys, torch
sys.path.insert(0, '.')
import ddsp_pytorch_b200 as d
from ddsp_pytorch_b200.train import Trainer, synthetic_batch
torch.manual_seed(0)
tr = Trainer(d.SHIPPED, "cuda:0", amp=True, noise_seed=7)
batch = synthetic_batch(d.SHIPPED, 4, "cuda:0", seed=2)
for i in range(6):
    with torch.autocast("cuda", dtype=torch.float16):
        ctrl = tr.zak.model.controller(batch)
    print(i, {k: (str(v.dtype), bool(torch.isfinite(v).all()), float(v.float().min()), float(v.float().max())) for k, v in ctrl.items() if k in ("c", "a", "H")})
    l = tr.training_step(batch)
    print("  loss", float(l), "scale", tr.scaler.get_scale(), "params finite", all(bool(torch.isfinite(p).all()) for p in tr.zak.parameters()))
