"""Per-kernel times of ctc_loss(logits) fwd+bwd at cfg5 (B=64 T=1000 V=5000 fp32)."""
import sys; sys.path.insert(0, '.')
import torch
import wenet_celoss_b200 as pkg
from torch.profiler import profile, ProfilerActivity
dev = torch.device('cuda:0')
g = torch.Generator().manual_seed(4321)
B5, T5, L5, V = 64, 1000, 100, 5000
yl = torch.randint(50, L5 + 1, (B5,), generator=g)
ys = torch.randint(1, V, (B5, L5), generator=g)
ys = torch.where(torch.arange(L5)[None] < yl[:, None], ys, torch.full_like(ys, -1)).to(dev)
hl = torch.full((B5,), T5, dtype=torch.int32, device=dev)
yl = yl.to(dev)
logits = torch.randn(B5, T5, V, generator=g).to(dev).requires_grad_(True)
def step():
    logits.grad = None
    l = pkg.ctc_loss(logits, ys, hl, yl, reduction="sum") / B5
    l.backward()
    return l
for _ in range(3): step()
ts = []
for _ in range(10):
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record(); step(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
print("step ms:", " ".join(f"{t:.3f}" for t in ts))
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(5): step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))

if len(sys.argv) > 1 and sys.argv[1] == "module":
    # the fused module (ctc_lo + log_softmax + loss): per-kernel device times of CTC.forward + backward
    E = 256
    m = pkg.CTC(V, E).to(dev)
    hs = torch.randn(B5, T5, E, generator=g).to(dev).requires_grad_(True)
    def mstep():
        hs.grad = None
        m.zero_grad(set_to_none=True)
        loss = m(hs, hl, ys, yl)
        loss.backward()
    for _ in range(3): mstep()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(5): mstep()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=24, max_name_column_width=70))
