import torch, time
from ddsp_pytorch_b200.config import Config
from ddsp_pytorch_b200.controller import Controller
conf = Config(); dev = "cuda"
c = Controller(conf).to(dev)
B, T = 32, 172
batch = {"normalized_cents": torch.rand(B, T, 1, device=dev), "loudness": torch.rand(B, T, 1, device=dev), "f0": torch.rand(B, T, 1, device=dev) * 400 + 100}
def step():
    out = c(batch)
    (out["c"].sum() + out["a"].sum() + out["H"].sum()).backward()
for _ in range(5): step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as p:
    for _ in range(5): step()
    torch.cuda.synchronize()
print(p.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
# GRU alone
g = c.gru; x = torch.rand(B, T, 1024, device=dev, requires_grad=True)
for nm, fn in (("gru fwd", lambda: g(x)), ("gru fwd+bwd", lambda: g(x)[0].sum().backward())):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(10): fn()
    torch.cuda.synchronize(); print(nm, (time.perf_counter() - t) * 100, "ms")
