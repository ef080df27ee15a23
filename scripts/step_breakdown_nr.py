"""Per-phase CUDA-event timing of one fused joint+loss step at cfg3, no-recompute backward
(forward keeps exp(logit); csrc/joint_nr.cu): python scripts/step_breakdown_nr.py [B] [iters]"""
import sys; sys.path.insert(0, '.')
import collections, torch
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
from wenet_celoss_b200 import joint as jm
from wenet_celoss_b200.rnnt import rnnt_grad_coef, rnnt_lattice
dev = torch.device('cuda:0')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T, U, V, J = 500, 100, 5000, 512
U1 = U + 1
g = torch.Generator().manual_seed(0)
e = torch.randn(B, T, J, generator=g).to(dev); p = torch.randn(B, U1, J, generator=g).to(dev)
w = (torch.randn(V, J, generator=g) / 16).to(dev).to(torch.bfloat16).contiguous()
bias = jm._pad_bias(torch.zeros(V, device=dev), V)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
st = _lib.stream_handle(dev)
acc = collections.OrderedDict()
class Ph:
    def __init__(s, name): s.name = name
    def __enter__(s):
        s.a = torch.cuda.Event(True); s.b = torch.cuda.Event(True); s.a.record()
    def __exit__(s, *x):
        s.b.record(); acc.setdefault(s.name, []).append((s.a, s.b))
cells = B * T * U1; ldg = (V + 15) // 16 * 16
bound = jm.joint_logit_bound(w, bias, V)
print("bound*log2e", float(bound) * 1.4427)
def step():
    with Ph('alloc'):
        H = torch.empty(cells, J, dtype=torch.bfloat16, device=dev)
        Pt = torch.empty(cells, ldg, dtype=torch.bfloat16, device=dev)
        lat = torch.empty(3, B, T, U1, dtype=torch.float32, device=dev)
        lse, lpb, lpe = lat.unbind(0)
    with Ph('fwd_p (H kernel + joint)'):
        _lib.call("b200_joint_fwd_p", ptr(e), ptr(p), ptr(w), ptr(bias), ptr(tgt), ptr(tl), ptr(ul), i32(B), i32(T),
                  i32(U1), i32(V), i32(J), i32(0), ptr(lse), ptr(lpb), ptr(lpe), ptr(bound), ptr(H), i32(J), ptr(Pt),
                  i32(ldg), i32(0), st)
    with Ph('lattice'):
        alpha, beta, costs, diag_off = rnnt_lattice(lpb, lpe, tl, ul)
    dy = torch.full((B,), 1.0 / B, device=dev)
    with Ph('coef'):
        coef = rnnt_grad_coef(lse, alpha, beta, diag_off, dy, tgt, tl, ul, 0)
    with Ph('alloc/fill bwd'):
        rowscale = torch.empty(cells, dtype=torch.float32, device=dev)
        Hs = torch.empty_like(H)
        dW = torch.zeros(V, J, dtype=torch.float32, device=dev); db = torch.zeros(V, device=dev)
        de = torch.zeros(B, T, J, device=dev); dp = torch.zeros(B, U1, J, device=dev)
    with Ph('rowscale'):
        _lib.call("b200_joint_rowscale", ptr(coef), ptr(lse), ptr(lpb), ptr(lpe), ptr(tgt), i32(B), i32(T), i32(U1),
                  i32(V), i32(J), i32(0), ptr(Pt), i32(ldg), ptr(H), i32(J), ptr(Hs), ptr(rowscale), st)
    with Ph('dW + db'):
        _lib.call("b200_joint_dw_db", ptr(Pt), i32(ldg), ptr(Hs), i32(J), jm._i64(cells), i32(V), i32(J), ptr(dW),
                  ptr(rowscale), ptr(db), i32(0), st)
    with Ph('dH scaled'):
        _lib.call("b200_joint_dh_scaled", ptr(Pt), i32(ldg), ptr(w), ptr(H), i32(J), ptr(rowscale), i32(B), i32(T),
                  i32(U1), i32(V), i32(J), ptr(tl), ptr(ul), ptr(de), ptr(dp), i32(0), st)
    return costs
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 6
for it in range(iters):
    acc.clear()
    t0 = torch.cuda.Event(True); t1 = torch.cuda.Event(True)
    t0.record(); c = step(); t1.record(); torch.cuda.synchronize()
    print(f"B={B} step {t0.elapsed_time(t1):.2f} ms  cost0 {float(c[0]):.3f}")
for k, v in acc.items():
    print(f"  {k:26s} {sum(a.elapsed_time(b) for a, b in v):8.3f} ms  ({len(v)} launches)")
