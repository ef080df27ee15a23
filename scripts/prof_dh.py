"""joint_dh_kernel alone on a cfg3 slice of nb utterances (argv[1], default 8): CUDA-event timing, TF/s,
and a check against torch on a few frames.  Also the target of the ncu captures under profiles/."""
import sys; sys.path.insert(0, '.')
import statistics, torch
from wenet_celoss_b200 import _lib
from wenet_celoss_b200._lib import i32, ptr
dev = torch.device('cuda:0')
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
T, U1, V, J, ldg = 500, 101, 5000, 512, 5008
n = nb * T * U1
G = torch.empty(n, ldg, device=dev, dtype=torch.bfloat16)
for i in range(0, n, 1 << 20):
    G[i:i + (1 << 20)] = (torch.randn(min(1 << 20, n - i), ldg, device=dev) * 0.01).to(torch.bfloat16)
H = torch.tanh(torch.randn(n, J, device=dev)).to(torch.bfloat16)
W = (torch.randn(V, J, device=dev) * 0.05).to(torch.bfloat16)
de = torch.zeros(nb, T, J, device=dev); dp = torch.zeros(nb, U1, J, device=dev)
st = _lib.stream_handle(dev)
def run():
    _lib.call("b200_joint_dh", ptr(G), i32(ldg), ptr(W), ptr(H), i32(J), i32(nb), i32(T), i32(U1), i32(V), i32(J),
              None, None, ptr(de), ptr(dp), i32(0), st)
run(); torch.cuda.synchronize()
ref = ((G[:U1 * 3, :V].float() @ W.float()) * (1 - H[:U1 * 3].float() ** 2)).view(3, U1, J).sum(1)
err = ((de[0, :3] - ref).abs().max() / ref.abs().max()).item()
ts = []
for _ in range(reps):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); run(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
ms = statistics.median(ts)
print(f"joint_dh nb={nb}: {ms:.3f} ms  {2.0 * n * J * V / ms / 1e9:.0f} TF/s  de rel err (3 frames) {err:.2e}")
