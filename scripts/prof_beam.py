"""Prefix beam search (beam 10) at the bench's cfg4 shape, short T: for ncu launch lists."""
import sys; sys.path.insert(0, '.')
import torch
import bench
import wenet_celoss_b200 as pkg
dev = torch.device('cuda:0')
cfg = dict(bench.CFG3)
torch.manual_seed(4321)
g = torch.Generator().manual_seed(4321)
T = int(sys.argv[1]) if len(sys.argv) > 1 else 20
model = bench._SearchModel(cfg).to(dev).eval()
enc = (torch.randn(64, T, cfg["E"], generator=g) * 2.0).to(dev)
lens = torch.full((64,), T, dtype=torch.int32, device=dev)
for _ in range(2):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    beams = pkg.prefix_beam_search_batch(model, enc, lens, beam_size=10)
    e1.record(); e1.synchronize()
print(f"T={T}: {e0.elapsed_time(e1):.2f} ms, {e0.elapsed_time(e1) * 1e3 / T:.1f} us/frame")
