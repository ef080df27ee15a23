import sys; sys.path.insert(0, '.')
import numpy as np, torch, torchaudio
import wenet_celoss_b200 as pkg
from oracle.c_oracle import rnnt_loss_c
dev='cuda:0'
g = torch.Generator().manual_seed(0)
B, T, U, V = 4, 50, 10, 500
logits = torch.randn(B, T, U + 1, V, generator=g) * 2
targets = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32)
tl = torch.tensor([50, 31, 44, 27], dtype=torch.int32)
ul = torch.tensor([7, 10, 5, 9], dtype=torch.int32)
x = logits.to(dev).requires_grad_(True)
loss = pkg.rnnt_loss(x, targets.to(dev), tl.to(dev), ul.to(dev), blank=0, reduction="mean")
loss.backward()
xr = logits.clone().requires_grad_(True)
ref = torchaudio.functional.rnnt_loss(xr, targets, tl, ul, blank=0, reduction="mean"); ref.backward()
c64, g64, a64, b64 = rnnt_loss_c(logits, targets, tl, ul, blank=0, dy=np.full(B, 1.0 / B))
mine = x.grad.cpu().numpy(); ta = xr.grad.numpy()
print("loss", loss.item(), ref.item(), c64.mean())
for name, r in (("torchaudio", ta), ("c64", g64)):
    err = np.abs(mine - r); tol = 1e-6 + 1e-4*np.abs(r)
    bad = err > tol
    print(name, "max abs", err.max(), "nbad", bad.sum(), "of", bad.size)
    idx = np.argwhere(bad)[:10]
    for i in idx:
        i = tuple(i); print("  ", i, mine[i], r[i], "target" if (i[2] < U and targets[i[0], i[2]] == i[3]) else ("blank" if i[3]==0 else "other"))
err = np.abs(ta - g64); print("torchaudio vs c64 max abs", err.max(), (err > 1e-6 + 1e-4*np.abs(g64)).sum())
