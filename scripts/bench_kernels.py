"""Per-kernel timings (CUDA events) of the small kernels; run on the GPU box."""
import sys; sys.path.insert(0, '.')
import statistics, torch
import wenet_celoss_b200 as pkg
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
dev = 'cuda:0'
def timed(fn, reps=10):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)
B, T, U1 = 32, 500, 101
g = torch.Generator(device=dev).manual_seed(0)
lpb = -torch.rand(B, T, U1, generator=g, device=dev) * 3 - 0.1
lpe = -torch.rand(B, T, U1, generator=g, device=dev) * 9 - 5
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U1 - 1, dtype=torch.int32, device=dev)
alpha = torch.empty_like(lpb); beta = torch.empty_like(lpb); costs = torch.empty(B, device=dev)
doff = torch.empty(B, 2, T + U1, dtype=torch.float64, device=dev)
st = _lib.stream_handle(torch.device(dev))
f = lambda: _lib.call("b200_rnnt_lattice", ptr(lpb), ptr(lpe), ptr(tl), ptr(ul), i32(B), i32(T), i32(U1), ptr(alpha), ptr(beta), ptr(costs), ptr(doff), st)
print("rnnt_lattice B32 T500 U100: %.3f ms" % timed(f), "cost0", float(costs[0]))
