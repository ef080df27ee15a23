"""Fused step with the no-recompute backward vs the recompute backward, same process, interleaved:
python scripts/ab_nr.py [cfg3|cfg2] [reps] [bf16|fp32]"""
import sys; sys.path.insert(0, '.')
import torch
import bench
from wenet_celoss_b200 import joint as jm
torch.cuda.set_device(0)
cfg = dict(bench.CFG2 if (len(sys.argv) > 1 and sys.argv[1] == "cfg2") else bench.CFG3)
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
if len(sys.argv) > 3:
    cfg["precision"] = sys.argv[3]  # bf16 | fp32
wl = bench.RnntFusedWorkload(cfg, 0)
def run(nr, n=10):
    jm.NO_RECOMPUTE = nr
    for _ in range(3):
        wl.step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(n):
        wl.step()
    b.record(); b.synchronize()
    return a.elapsed_time(b) / n, jm.LAST_BACKWARD_PATH
for r in range(reps):
    for nr in (True, False):
        ms, path = run(nr)
        print(f"{path:13s} {ms:7.3f} ms/step  {cfg['B'] / ms * 1e3:8.1f} utt/s")
