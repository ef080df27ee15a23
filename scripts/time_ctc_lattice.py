"""ctc_lattice_kernel alone at cfg5 (B=64 T=1000 L in [50,100]): CUDA-event time per launch."""
import sys; sys.path.insert(0, '.')
import statistics, torch
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
dev = torch.device('cuda:0')
g = torch.Generator().manual_seed(4321)
B, T, L, V = 64, 1000, 100, 5000
yl = torch.randint(50, L + 1, (B,), generator=g)
ys = torch.randint(1, V, (B, L), generator=g)
ys = torch.where(torch.arange(L)[None] < yl[:, None], ys, torch.full_like(ys, -1)).to(dev)
hl = torch.full((B,), T, dtype=torch.int32, device=dev); yl = yl.to(dev).int()
S = 2 * L + 1
lp = (-8.5 + torch.randn(B, T, S, generator=g)).to(dev)
al = torch.empty(B, T, S, device=dev); be = torch.empty(B, T, S, device=dev)
nll = torch.empty(B, device=dev); leader = torch.empty(B, S, dtype=torch.int32, device=dev)
off = torch.zeros(B, 2, T + 1, dtype=torch.float64, device=dev)
st = _lib.stream_handle(dev)
f = lambda: _lib.call("b200_ctc_lattice", ptr(lp), ptr(ys), ptr(hl), ptr(yl), i32(B), i32(T), i32(L), i32(0),
                      ptr(al), ptr(be), ptr(nll), ptr(leader), ptr(off), st)
f(); torch.cuda.synchronize(); ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); f(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
print(f"ctc_lattice: {statistics.median(ts) * 1e3:.1f} us  nll0 {float(nll[0]):.3f}")
