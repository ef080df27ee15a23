"""Real (warm-cache, in-stream) per-kernel durations of the RNN-T prefix beam search: torch profiler."""
import sys; sys.path.insert(0, '.')
import runpy, torch
from torch.profiler import ProfilerActivity, profile
sys.argv = ["prof_beam.py", sys.argv[1] if len(sys.argv) > 1 else "100"]
runpy.run_path("scripts/prof_beam.py")
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    runpy.run_path("scripts/prof_beam.py")
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / max(e.count, 1), e.count, e.device_time_total) for e in prof.key_averages()
        if e.device_time_total > 0]
rows.sort(key=lambda r: -r[3])
for k, t, c, tot in rows[:10]:
    print(f"{t:8.1f} us x{c:<5d} total {tot / 1e3:8.2f} ms  {k[:100]}")
