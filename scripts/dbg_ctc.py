import sys; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, torch, torch.nn.functional as F
from test_ctc_gpu import _case, _mine
from oracle.c_oracle import ctc_loss_c
for shape in [(4, 60, 10, 500), (3, 50, 7, 4233), (2, 9, 2, 5), (5, 120, 40, 1030)]:
    B, T, L, V = shape
    logits, ys, hl, yl = _case(B, T, L, V, seed=T * 7 + V)
    scale = torch.linspace(0.5, 1.5, B)
    nll, grad = _mine(logits, ys, hl, yl, scale)
    x = logits.clone().requires_grad_(True)
    ref = F.ctc_loss(x.transpose(0, 1).log_softmax(2), ys, hl, yl, blank=0, reduction="none")
    (ref * scale).sum().backward(); ta = x.grad.numpy()
    xd = logits.double().requires_grad_(True)
    refd = F.ctc_loss(xd.transpose(0, 1).log_softmax(2), ys, hl, yl, blank=0, reduction="none")
    (refd * scale.double()).sum().backward(); td = xd.grad.numpy()
    g64 = np.zeros_like(ta); n64 = np.zeros(B)
    for b in range(B):
        n1, g1 = ctc_loss_c(logits[b:b+1], ys[b:b+1], hl[b:b+1], yl[b:b+1], blank=0, scale=float(scale[b]))
        n64[b], g64[b] = n1[0], g1[0]
    def st(a, b):
        e = np.abs(a - b); i = np.unravel_index(e.argmax(), e.shape)
        return f"max {e.max():.3e} at {i} vals {a[i]:.6e} {b[i]:.6e}"
    print(shape, "nll", nll, n64)
    print("  mine-c64 ", st(grad, g64)); print("  mine-aten", st(grad, ta)); print("  aten-c64 ", st(ta, g64))
    print("  aten64-c64", st(td.astype(np.float32), g64)); print("  mine-aten64", st(grad, td.astype(np.float32)))
    print("  rowsum mine", np.abs(grad.sum(-1)).max(), "aten", np.abs(ta.sum(-1)).max(), "c64", np.abs(g64.sum(-1)).max())
