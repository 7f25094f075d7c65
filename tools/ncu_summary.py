"""Summarise ncu outputs brought back in gpurun_out/ (run here, no GPU needed).
  python tools/ncu_summary.py launches gpurun_out/launches.csv
  python tools/ncu_summary.py rep gpurun_out/prof_gemm.ncu-rep
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
        "smsp__cycles_active.avg", "gpc__cycles_elapsed.max", "sm__pipe_tensor_op", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_xu", "lts__t_sector_hit_rate.pct", "dram__cycles_active", "gpu__dram_throughput"]


def launches(path):
    rows = [l for l in open(path) if l.startswith('"')]
    rd = csv.DictReader(io.StringIO("".join(rows)))
    agg = defaultdict(list)
    for r in rd:
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        agg[r["Kernel Name"][:90]].append(us)
    total = sum(sum(v) for v in agg.values())
    print(f"{'kernel':92s} {'n':>4s} {'mean_us':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:92s} {len(v):4d} {sum(v) / len(v):10.1f} {sum(v) / total:7.1%}")


def rep(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units = rd[0], rd[1]
    for row in rd[2:]:
        print("==", row[hdr.index("Kernel Name")][:80], "grid", row[hdr.index("Grid Size")])
        for i, h in enumerate(hdr):
            if any(h.startswith(k) for k in KEYS):
                print(f"   {h:75s} {row[i]:>18s} {units[i]}")


if __name__ == "__main__":
    {"launches": launches, "rep": rep}[sys.argv[1]](sys.argv[2])
