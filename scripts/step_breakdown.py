"""Per-phase CUDA-event timing of one fused joint+loss step at cfg3 (phases of the backward loop)."""
import sys; sys.path.insert(0, '.')
import collections, torch
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
from wenet_celoss_b200 import joint as jm
from wenet_celoss_b200.rnnt import rnnt_grad_coef, rnnt_lattice
dev = torch.device('cuda:0')
B, T, U, V, J = 32, 500, 100, 5000, 512
U1 = U + 1
g = torch.Generator().manual_seed(0)
e = torch.randn(B, T, J, generator=g).to(dev); p = torch.randn(B, U1, J, generator=g).to(dev)
w = (torch.randn(V, J, generator=g) / 16).to(dev).to(torch.bfloat16).contiguous()
bias = jm._pad_bias(torch.zeros(V, device=dev), V)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 1
st = _lib.stream_handle(dev)
acc = collections.OrderedDict()
class Ph:
    def __init__(s, name): s.name = name
    def __enter__(s):
        s.a = torch.cuda.Event(True); s.b = torch.cuda.Event(True); s.a.record()
    def __exit__(s, *x):
        s.b.record(); acc.setdefault(s.name, []).append((s.a, s.b))
def step():
    with Ph('fwd joint'):
        lse, lpb, lpe = jm.joint_logprobs(e, p, w, bias, tgt, tl, ul, V, 0)
    with Ph('lattice'):
        alpha, beta, costs, diag_off = rnnt_lattice(lpb, lpe, tl, ul)
    dy = torch.full((B,), 1.0 / B, device=dev)
    with Ph('coef'):
        coef = rnnt_grad_coef(lse, alpha, beta, diag_off, dy, tgt, tl, ul, 0)
    per_utt = T * U1; ldg = (V + 15) // 16 * 16; ldh = J
    with Ph('alloc/fill'):
        G = torch.empty(nb * per_utt, ldg, dtype=torch.bfloat16, device=dev)
        H = torch.empty(nb * per_utt, ldh, dtype=torch.bfloat16, device=dev)
        dbp = torch.zeros(16, (V + 127) // 128 * 128, device=dev)
        dW = torch.zeros(V, J, dtype=torch.float32, device=dev)
        de = torch.empty(B, T, J, device=dev); dp = torch.zeros(B, U1, J, device=dev)
    for b0 in range(0, B, nb):
        b1 = min(B, b0 + nb); n = (b1 - b0) * per_utt
        with Ph('grad joint (+H)'):
            _lib.call("b200_joint_grad_chunk", ptr(e), ptr(p), ptr(w), ptr(bias), ptr(tgt), ptr(coef), i32(B), i32(T),
                      i32(U1), i32(V), i32(J), i32(0), jm._i64(b0 * per_utt), jm._i64(n), ptr(G), i32(ldg), ptr(H),
                      i32(ldh), ptr(dbp), i32(0), st)
        Gv = G[:n, :V]
        with Ph('dH gemm'):
            dH = torch.mm(Gv, w, out_dtype=torch.float32)
        with Ph('dW gemm'):
            jm.joint_dw(G, H, n, V, J, dW)
        with Ph('dz reduce'):
            _lib.call("b200_joint_dz_reduce", ptr(dH), ptr(H), i32(ldh), i32(b1 - b0), i32(T), i32(U1), i32(J),
                      ptr(de[b0:b1]), ptr(dp[b0:b1]), st)
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for it in range(iters):
    acc.clear()
    t0 = torch.cuda.Event(True); t1 = torch.cuda.Event(True)
    t0.record(); step(); t1.record(); torch.cuda.synchronize()
print(f"nb={nb} step {t0.elapsed_time(t1):.2f} ms")
for k, v in acc.items():
    print(f"  {k:18s} {sum(a.elapsed_time(b) for a, b in v):8.3f} ms  ({len(v)} launches)")
