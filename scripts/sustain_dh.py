"""Does a tensor kernel keep its speed when it runs back to back for seconds (power cap)?  argv[1] = dh | dw | fwd"""
import sys; sys.path.insert(0, '.')
import torch
from wenet_celoss_b200 import _lib, joint as jm
from wenet_celoss_b200._lib import i32, ptr
dev = torch.device('cuda:0')
which = sys.argv[1] if len(sys.argv) > 1 else "dh"
nb, T, U1, V, J, ldg = 32, 500, 101, 5000, 512, 5008
n = nb * T * U1
G = torch.empty(n, ldg, device=dev, dtype=torch.bfloat16)
for i in range(0, n, 1 << 20):
    G[i:i + (1 << 20)] = (torch.randn(min(1 << 20, n - i), ldg, device=dev) * 0.01).to(torch.bfloat16)
H = torch.tanh(torch.randn(n, J, device=dev)).to(torch.bfloat16)
W = (torch.randn(V, J, device=dev) * 0.05).to(torch.bfloat16)
de = torch.zeros(nb, T, J, device=dev); dp = torch.zeros(nb, U1, J, device=dev)
dW = torch.zeros(V, J, device=dev)
st = _lib.stream_handle(dev)
def dh():
    _lib.call("b200_joint_dh", ptr(G), i32(ldg), ptr(W), ptr(H), i32(J), i32(nb), i32(T), i32(U1), i32(V), i32(J),
              None, None, ptr(de), ptr(dp), i32(0), st)
def dw():
    jm.joint_dw(G, H, n, V, J, dW)
fn = {"dh": dh, "dw": dw}[which]
fn(); torch.cuda.synchronize()
import pynvml
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
for blk in range(12):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(25): fn()
    e1.record()
    clk = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM); pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1e3
    e1.synchronize()
    print(f"{which} block {blk}: {e0.elapsed_time(e1) / 25:.3f} ms/launch  sm {clk} MHz  {pw:.0f} W")
