"""One GRAD fused-joint chunk launch (one cfg3 utterance), for ncu captures."""
import sys; sys.path.insert(0, '.')
import ctypes, os, pathlib, torch
from wenet_celoss_b200 import _lib
if os.environ.get("B200_DBG_LIB"):   # knock-out builds of the GRAD epilogue (scripts/_dbg, -DB200_GRAD_DBG=n)
    _lib.LIB_PATH = pathlib.Path(os.environ["B200_DBG_LIB"])
from wenet_celoss_b200._lib import i32, ptr
from wenet_celoss_b200.joint import _pad_bias
dev = 'cuda:0'
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 1
B, T, U, V, J = max(2, nb), 500, 100, 5000, 512
U1 = U + 1
g = torch.Generator().manual_seed(0)
e = torch.randn(B, T, J, generator=g).to(dev); p = torch.randn(B, U1, J, generator=g).to(dev)
w = (torch.randn(V, J, generator=g) / 16).to(dev).to(torch.bfloat16).contiguous()
bias = _pad_bias(torch.zeros(V, device=dev), V)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
coef = torch.zeros(B, T, U1, 4, device=dev); coef[..., 0] = -9.0; coef[..., 1] = -3.0; coef[..., 2] = -3.0; coef[..., 3] = 1.0
per_utt = T * U1; ldg = (V + 15) // 16 * 16; ldh = J + 8
dbg = int(sys.argv[2]) if len(sys.argv) > 2 else 0
G = torch.empty(nb * per_utt, ldg, dtype=torch.bfloat16, device=dev)
H = torch.zeros(nb * per_utt, ldh, dtype=torch.bfloat16, device=dev)
dbp = torch.zeros(16, (V + 127) // 128 * 128, device=dev) if dbg else None
st = _lib.stream_handle(torch.device(dev))
ts = []
for _ in range(4):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    _lib.call("b200_joint_grad_chunk", ptr(e), ptr(p), ptr(w), ptr(bias), ptr(tgt), ptr(coef), i32(B), i32(T),
              i32(U1), i32(V), i32(J), i32(0), ctypes.c_longlong(0), ctypes.c_longlong(nb * per_utt), ptr(G),
              i32(ldg), ptr(H), i32(ldh), ptr(dbp), i32(0), st)
    e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
tiles = -(-nb * per_utt // 128)
print(f"GRAD dbg={dbg} chunk nb={nb}: {min(ts):.3f} ms, tiles {tiles}, waves {-(-tiles // 148)}, per wave {min(ts) * 1e3 / -(-tiles // 148):.1f} us")
