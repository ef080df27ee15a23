"""Gated hot-word greedy (context_filter_state='on') at the bench's cfg4 shape, short T: for ncu launch lists."""
import sys; sys.path.insert(0, '.')
import torch
import bench
import wenet_celoss_b200 as pkg
dev = torch.device('cuda:0')
cfg = dict(bench.CFG3)
torch.manual_seed(4321)
g = torch.Generator().manual_seed(4321)
T = int(sys.argv[1]) if len(sys.argv) > 1 else 40
model = bench._SearchModel(cfg).to(dev).eval()
model.context_bias = bench._ContextBias(cfg["P"], 100).to(dev).eval()
with torch.no_grad():
    model.joint.ffn_out.bias[0] -= 2.0
enc = (torch.randn(64, T, cfg["E"], generator=g) * 2.0).to(dev)
lens = torch.full((64,), T, dtype=torch.int32, device=dev)
ctx_list = torch.zeros(100, 4, dtype=torch.int32, device=dev)
ctx_len = torch.full((100,), 4, dtype=torch.int32, device=dev)
for _ in range(2):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    hyps, dec, rounds = pkg.greedy_search_batch(model, enc, lens, n_steps=64, context_list=ctx_list,
                                                context_lengths=ctx_len, context_filter_state="on",
                                                return_decisions=True, return_rounds=True)
    e1.record(); e1.synchronize()
print(f"T={T}: {e0.elapsed_time(e1):.2f} ms, rounds {rounds}, {e0.elapsed_time(e1) * 1e3 / rounds:.1f} us/round, "
      f"tokens {sum(len(h) for h in hyps)}, decisions {sum(len(d) for d in dec)}")
