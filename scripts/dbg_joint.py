"""Debug/diagnostic run of the fused joint kernels against a torch reference (GPU)."""
import sys; sys.path.insert(0, '.')
import torch, numpy as np
import wenet_celoss_b200 as pkg
from wenet_celoss_b200 import _lib
from wenet_celoss_b200.joint import joint_logprobs, _pad_bias
dev = 'cuda:0'
def run(B, T, U, V, J, max_ctas=0, seed=0, use_bound=True):
    g = torch.Generator().manual_seed(seed)
    U1 = U + 1
    e = torch.randn(B, T, J, generator=g).to(dev)
    p = torch.randn(B, U1, J, generator=g).to(dev)
    w = (torch.randn(V, J, generator=g) / J ** 0.5 * 2).to(dev)
    bias = torch.randn(V, generator=g).to(dev)
    tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
    tl = torch.full((B,), T, dtype=torch.int32, device=dev)
    ul = torch.full((B,), U, dtype=torch.int32, device=dev)
    wb = w.to(torch.bfloat16).contiguous()
    lse, lpb, lpe = joint_logprobs(e, p, wb, _pad_bias(bias, V), tgt, tl, ul, V, 0, max_ctas, use_bound)
    torch.cuda.synchronize()
    h = torch.tanh(e[:, :, None] + p[:, None]).to(torch.bfloat16).float()
    logits = h @ wb.float().t() + bias
    rl = torch.logsumexp(logits, -1)
    rb = logits[..., 0] - rl
    idx = torch.cat([tgt.long(), torch.zeros(B, 1, dtype=torch.long, device=dev)], 1)[:, None, :, None].expand(B, T, U1, 1)
    re = logits.gather(-1, idx)[..., 0] - rl
    print(f"bound={int(use_bound)} B{B} T{T} U{U} V{V} J{J} ctas{max_ctas}: lse err {float((lse-rl).abs().max()):.3e} "
          f"lpb err {float((lpb-rb).abs().max()):.3e} lpe err {float((lpe-re)[:, :, :U].abs().max()):.3e} "
          f"(lse range {float(rl.min()):.2f}..{float(rl.max()):.2f})", flush=True)
    return float((lse - rl).abs().max())
for args in [(1, 8, 7, 256, 64), (1, 8, 7, 256, 128), (2, 20, 7, 300, 128), (2, 20, 7, 300, 512),
             (2, 40, 15, 1000, 512), (3, 50, 20, 5000, 512)]:
    run(*args)
    run(*args, use_bound=False)
run(4, 100, 30, 5000, 512, max_ctas=7)
run(4, 100, 30, 4233, 512, max_ctas=3)
