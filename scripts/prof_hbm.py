"""One fwd+bwd of the HBM-bound paths for ncu captures: rnnt_loss over materialised bf16 logits at cfg3
(argv[1] = B, default 32) and ctc_loss over fp32 logits at cfg5."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import wenet_celoss_b200 as pkg
torch.cuda.set_device(0)
cfg = dict(bench.CFG3)
cfg["B"] = int(sys.argv[1]) if len(sys.argv) > 1 else 32
wl = bench.RnntLogitsWorkload(cfg, 0)
for _ in range(2):
    l = wl.step()
torch.cuda.synchronize()
print("rnnt_logits loss", float(l))
del wl
torch.cuda.empty_cache()
dev = "cuda:0"
g = torch.Generator().manual_seed(4321)
B5, T5, L5, V = 64, 1000, 100, 5000
logits = torch.randn(B5, T5, V, generator=g).to(dev).requires_grad_(True)
yl = torch.randint(50, L5 + 1, (B5,), generator=g)
ys = torch.randint(1, V, (B5, L5), generator=g)
ys = torch.where(torch.arange(L5)[None] < yl[:, None], ys, torch.full_like(ys, -1)).to(dev)
hl = torch.full((B5,), T5, dtype=torch.int32, device=dev)
for _ in range(2):
    logits.grad = None
    l = pkg.ctc_loss(logits, ys, hl, yl.to(dev), reduction="sum") / B5
    l.backward()
torch.cuda.synchronize()
print("ctc loss", float(l))
