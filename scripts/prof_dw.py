"""Single launches of joint_dw_kernel (cfg3 slice of nb utterances) for an ncu --set full capture."""
import sys; sys.path.insert(0, '.')
import torch
from wenet_celoss_b200 import joint as jm
dev = torch.device('cuda:0')
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n, V, J, ldg = nb * 50500, 5000, 512, 5008
G = (torch.randn(n, ldg, device=dev) * 0.01).to(torch.bfloat16)
H = torch.tanh(torch.randn(n, J, device=dev)).to(torch.bfloat16)
dW = torch.zeros(V, J, device=dev)
for _ in range(3):
    jm.joint_dw(G, H, n, V, J, dW)
torch.cuda.synchronize()
print("ok", float(dW.abs().max()))
