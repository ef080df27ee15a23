"""Per-kernel device time of bench.py's fused step (torch profiler, 3 steps after warm-up):
python scripts/prof_step.py [workload]"""
import sys; sys.path.insert(0, '.')
import torch
from torch.profiler import ProfilerActivity, profile
import bench
torch.cuda.set_device(0)
name = sys.argv[1] if len(sys.argv) > 1 else "rnnt_fused"
wl = bench.WORKLOADS[name](dict(bench.CFG3), 0)
for _ in range(4):
    wl.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(10):
    wl.step()
e1.record(); e1.synchronize()
print(f"step (10 back-to-back, no profiler): {e0.elapsed_time(e1) / 10:.3f} ms")
n = 3
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(n):
        wl.step()
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / n, e.count // n) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows if not r[0].startswith(("aten::", "autograd::", "_Fused", "torch::")))
print(f"sum of kernel time per step: {tot / 1e3:.3f} ms")
for k, t, c in rows[:40]:
    if k.startswith(("aten::", "autograd::", "_Fused", "torch::")):
        continue
    print(f"{t / 1e3:9.3f} ms  x{c:<3d} {k[:110]}")
