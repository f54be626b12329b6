import torch, sys
sys.path.insert(0, ".")
import ddsp_pytorch_b200 as d
from ddsp_pytorch_b200.train import Trainer, synthetic_batch
conf = d.SHIPPED
torch.manual_seed(0)
tr = Trainer(conf, "cuda:0", amp=True, noise_seed=1)
batches = [synthetic_batch(conf, 8, "cuda:0", seed=s) for s in range(4)]
for i in range(80):
    loss = tr.training_step(batches[i % 4])
    if i % 5 == 0 or i > 70:
        print(i, float(loss), tr.scaler.get_scale())
bad = [n for n, p in tr.zak.named_parameters() if not torch.isfinite(p).all()]
print("non-finite params:", bad[:5])
