"""One FWD (and optionally GRAD) fused-joint launch at a cfg3-shaped slice, for ncu captures."""
import sys; sys.path.insert(0, '.')
import torch
from wenet_celoss_b200.joint import joint_logprobs, _pad_bias
dev = 'cuda:0'
B, T, U, V, J = (int(sys.argv[1]) if len(sys.argv) > 1 else 6), 500, 100, 5000, 512   # B=6: 2367 tiles = 16 per SM
g = torch.Generator().manual_seed(0)
e = torch.randn(B, T, J, generator=g).to(dev); p = torch.randn(B, U + 1, J, generator=g).to(dev)
w = (torch.randn(V, J, generator=g) / 16).to(dev).to(torch.bfloat16).contiguous()
bias = _pad_bias(torch.zeros(V, device=dev), V)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
for _ in range(2):
    joint_logprobs(e, p, w, bias, tgt, tl, ul, V, 0, 0)
torch.cuda.synchronize()
