"""FWD joint kernel time vs number of CTAs: is the W_out stream from L2 the limiter?"""
import sys; sys.path.insert(0, '.')
import statistics, torch
from wenet_celoss_b200.joint import joint_logprobs, _pad_bias
dev = 'cuda:0'
B, T, U, V, J = 8, 500, 100, 5000, 512
g = torch.Generator().manual_seed(0)
e = torch.randn(B, T, J, generator=g).to(dev); p = torch.randn(B, U + 1, J, generator=g).to(dev)
w = (torch.randn(V, J, generator=g) / 16).to(dev).to(torch.bfloat16).contiguous()
bias = _pad_bias(torch.zeros(V, device=dev), V)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
tiles = B * T * (U + 1) // 128
for ctas in (148, 111, 74, 37, 16):
    ts = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); joint_logprobs(e, p, w, bias, tgt, tl, ul, V, 0, ctas); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = statistics.median(ts[1:])
    waves = -(-tiles // ctas)
    print(f"ctas {ctas:4d}: {ms:7.3f} ms  per-tile-wave {ms / waves * 1e3:7.1f} us  TF/s {2.0*B*T*(U+1)*J*V/ms/1e9:7.1f}")
