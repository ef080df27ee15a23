"""Sustained run of the bench step (fused_joint_rnnt_loss fwd+bwd at cfg3) with per-step times, to
separate clock droop under the power cap from per-kernel costs."""
import sys; sys.path.insert(0, '.')
import torch, bench
wl = bench.RnntFusedWorkload(dict(bench.CFG3), 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
evs = []
for i in range(n):
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record(); wl.step(); b.record(); evs.append((a, b))
torch.cuda.synchronize()
print("per-step ms:", " ".join(f"{a.elapsed_time(b):.2f}" for a, b in evs))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(3):
        wl.step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
