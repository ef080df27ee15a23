"""Print the numbers profiles/README.md quotes from an `ncu -i X.ncu-rep --page raw --csv` export:
duration, DRAM bytes, achieved GB/s against the measured HBM peak, tensor-pipe activity, clocks."""
import csv, json, os, sys

WANT = [("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("sm__cycles_elapsed.avg.per_second", "clk"), ("lts__t_sector_hit_rate.pct", "l2hit"),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_el"),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor_act"),
        ("sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "uni"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue"),
        ("launch__registers_per_thread", "regs")]
SCALE = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9,
         "Ghz": 1e9, "Mhz": 1e6}


def main(path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        peak = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6556.8
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {}
    for name, key in WANT:
        for i, h in enumerate(hdr):
            if h == name or h.endswith("." + name):
                col[key] = i
                break
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        v = {}
        for key, i in col.items():
            try:
                v[key] = float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)
            except ValueError:
                pass
        line = f"{r[ki][:48]:48s} {v.get('dur', 0) * 1e6:9.1f} us"
        if "rd" in v:
            gb = (v["rd"] + v.get("wr", 0)) / 1e9
            line += f"  dram rd {v['rd'] / 1e9:7.3f} wr {v.get('wr', 0) / 1e9:7.3f} GB = {gb / v['dur']:6.0f} GB/s ({gb / v['dur'] / peak:.2f} of {peak:.0f})"
        for key in ("tensor_el", "tensor_act", "issue", "l2hit", "regs"):
            if key in v:
                line += f"  {key} {v[key]:.1f}"
        if "clk" in v:
            line += f"  {v['clk'] / 1e9:.2f} GHz"
        print(line)


if __name__ == "__main__":
    main(sys.argv[1])
