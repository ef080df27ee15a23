"""CUDA-event time of joint_kernel<FWD> (+H store) and joint_kernel<GRAD> at the cfg3 launch size
(B=32 unless argv[1]); B200_LIB=... selects a variant build (scripts/build_variant.sh)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
from wenet_celoss_b200.joint import _i64, _pad_bias, joint_logprobs
from wenet_celoss_b200.rnnt import rnnt_grad_coef, rnnt_lattice
dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T, U, V, J = 500, 100, 5000, 512
U1 = U + 1
g = torch.Generator().manual_seed(0)
e = (torch.randn(B, T, J, generator=g) * 0.6).to(dev); p = (torch.randn(B, U1, J, generator=g) * 0.6).to(dev)
w = ((torch.rand(V, J, generator=g) * 2 - 1) / J ** 0.5).to(dev).to(torch.bfloat16).contiguous()
bias = _pad_bias(((torch.rand(V, generator=g) * 2 - 1) / J ** 0.5).to(dev), V)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
n = B * T * U1
H = torch.empty(n, J, dtype=torch.bfloat16, device=dev)
st = _lib.stream_handle(dev)
def timed(fn, it=4):
    ts = []
    for _ in range(it + 1):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    return sum(ts[1:]) / it
box = {}
def fwd():
    box["lat"] = joint_logprobs(e, p, w, bias, tgt, tl, ul, V, 0, H=H)
t_fwd = timed(fwd)
lse, lpb, lpe = box["lat"]
alpha, beta, costs, doff = rnnt_lattice(lpb, lpe, tl, ul)
dy = torch.full((B,), 1.0 / B, device=dev)
coef = rnnt_grad_coef(lse, alpha, beta, doff, dy, tgt, tl, ul, 0)
ldg = (V + 15) // 16 * 16
G = torch.empty(n, ldg, dtype=torch.bfloat16, device=dev)
dbp = torch.zeros(16, (V + 127) // 128 * 128, device=dev)
def grad():
    _lib.call("b200_joint_grad_chunk_h", ptr(e), ptr(p), ptr(w), ptr(bias), ptr(tgt), ptr(coef), i32(B), i32(T),
              i32(U1), i32(V), i32(J), i32(0), _i64(0), _i64(n), ptr(G), i32(ldg), ptr(H), i32(J), ptr(dbp), i32(0), st)
t_grad = timed(grad)
from wenet_celoss_b200.joint import joint_dw
dW = torch.zeros(V, J, device=dev); de = torch.zeros(B, T, J, device=dev); dp = torch.zeros(B, U1, J, device=dev)
t_dw = timed(lambda: joint_dw(G, H, n, V, J, dW))
def dh():
    _lib.call("b200_joint_dh", ptr(G), i32(ldg), ptr(w), ptr(H), i32(J), i32(B), i32(T), i32(U1), i32(V), i32(J),
              ptr(tl), ptr(ul), ptr(de), ptr(dp), i32(0), st)
t_dh = timed(dh)
flops = 2.0 * n * J * V
print(f"   dW {t_dw:.3f} ms ({flops / t_dw / 1e9:.0f} TF/s)  dH {t_dh:.3f} ms ({flops / t_dh / 1e9:.0f} TF/s)")
print(f"{os.environ.get('B200_LIB', 'default'):40s} B={B} FWD {t_fwd:.3f} ms ({flops / t_fwd / 1e9:.0f} TF/s)  "
      f"GRAD {t_grad:.3f} ms ({flops / t_grad / 1e9:.0f} TF/s)  costs[0] {costs[0].item():.3f}", flush=True)
