"""joint_dw_kernel with / without the bias-gradient column sums at the cfg3 size: python scripts/time_dw.py [B]"""
import sys; sys.path.insert(0, '.')
import torch
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
from wenet_celoss_b200 import joint as jm
dev = torch.device('cuda:0')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T, U1, V, J = 500, 101, 5000, 512
n = B * T * U1; ldg = 5008
G = (torch.randn(n, ldg, device=dev, dtype=torch.bfloat16))
H = torch.randn(n, J, device=dev, dtype=torch.bfloat16)
rs = torch.rand(n, device=dev)
dW = torch.zeros(V, J, device=dev); db = torch.zeros(V, device=dev)
st = _lib.stream_handle(dev)
def run(with_db):
    if with_db:
        _lib.call("b200_joint_dw_db", ptr(G), i32(ldg), ptr(H), i32(J), jm._i64(n), i32(V), i32(J), ptr(dW), ptr(rs),
                  ptr(db), i32(0), st)
    else:
        _lib.call("b200_joint_dw", ptr(G), i32(ldg), ptr(H), i32(J), jm._i64(n), i32(V), i32(J), ptr(dW), i32(0), st)
for with_db in (False, True) * 4:
    ts = []
    for i in range(5):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(); run(with_db); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    print("db" if with_db else "plain", " ".join(f"{t:.2f}" for t in ts), "ms")
m = min(n, 64 * 4096)
ref = (G[:m, :V].float() * rs[:m, None]).sum(0)
db.zero_()
_lib.call("b200_joint_dw_db", ptr(G), i32(ldg), ptr(H), i32(J), jm._i64(m), i32(V), i32(J), ptr(dW), ptr(rs), ptr(db),
          i32(0), st)
torch.cuda.synchronize()
print("db check rel err", float((db - ref).norm() / ref.norm()))
