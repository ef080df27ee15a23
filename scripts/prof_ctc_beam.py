"""CTC prefix beam search (beam 10) at the cfg5 shape, short T: timing and a target for ncu."""
import sys; sys.path.insert(0, '.')
import torch
from wenet_celoss_b200.ctc_search import ctc_prefix_beam_search_batch
dev = torch.device('cuda:0')
T = int(sys.argv[1]) if len(sys.argv) > 1 else 200
B, V = 64, 5000
g = torch.Generator().manual_seed(11)
lp = torch.log_softmax(torch.randn(B, T, V, generator=g) * 3.0, dim=2).to(dev)
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
for _ in range(2):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    hyps = ctc_prefix_beam_search_batch(lp, lens, 10)
    e1.record(); e1.synchronize()
print(f"T={T}: {e0.elapsed_time(e1):.2f} ms, {e0.elapsed_time(e1) * 1e3 / T:.1f} us/frame (incl. top-k kernel, D2H)")
