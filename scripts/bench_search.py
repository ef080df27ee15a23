"""cfg4 greedy (and optionally beam) timing, persistent kernel vs round kernels."""
import os, sys; sys.path.insert(0, '.')
import statistics, torch
import bench
import wenet_celoss_b200 as pkg
dev = torch.device('cuda:0')
cfg = dict(bench.CFG3) if hasattr(bench, 'CFG3') else dict(B=32, T=500, U=100, V=5000, E=256, P=256, J=512)
torch.manual_seed(4321); g = torch.Generator().manual_seed(4321)
B4, T4 = 64, 500
model = bench._SearchModel(cfg).to(dev).eval()
enc = (torch.randn(B4, T4, cfg["E"], generator=g) * 2.0).to(dev)
lens = torch.full((B4,), T4, dtype=torch.int32, device=dev)
def timed(fn, reps=3):
    fn(); ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); r = fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    return statistics.median(ts), r
res = {}
for mode in ("0", "1"):
    os.environ["B200_GREEDY_ROUND_KERNELS"] = mode
    ms, (hyps, rounds) = timed(lambda: pkg.greedy_search_batch(model, enc, lens, n_steps=64, return_rounds=True))
    res[mode] = hyps
    print(f"greedy round_kernels={mode}: {ms:.2f} ms, {B4 / ms * 1e3:.0f} utt/s, rounds {rounds}, us/round {ms * 1e3 / rounds:.1f}, tokens {sum(len(h) for h in hyps)}")
print("identical hypotheses:", res["0"] == res["1"])
if len(sys.argv) > 1:
    ms, beams = timed(lambda: pkg.prefix_beam_search_batch(model, enc, lens, beam_size=10), reps=2)
    print(f"beam10: {ms:.1f} ms, {B4 / ms * 1e3:.1f} utt/s")
