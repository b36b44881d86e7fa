"""Summarise ncu output for profiles/:  launches CSV -> per-kernel time shares;  .ncu-rep -> key metrics per launch.
    python tools/ncu_summary.py launches gpurun_out/x_launches.csv
    python tools/ncu_summary.py rep gpurun_out/x.ncu-rep
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "launch__grid_size",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum",
        # what the attention kernels are argued to be bound by: the MUFU (xu) pipe and shared-memory wavefronts
        "sm__inst_executed_pipe_xu.sum", "sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg, tot = collections.defaultdict(lambda: [0, 0.0]), 0.0
    for row in csv.DictReader(lines):
        try:
            t = float(row["Metric Value"].replace(",", ""))
        except Exception:
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])[:100]
        agg[name][0] += 1
        agg[name][1] += t
        tot += t
    print(f"total {tot / 1e6:.3f} ms over {sum(v[0] for v in agg.values())} launches (ncu: serialised, cold cache -> compare SHARES)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1] / 1e6:9.3f} ms {100 * v[1] / tot:5.1f}%  n={v[0]:5d}  avg={v[1] / v[0] / 1e3:8.1f} us  {k}")


def rep(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("----", r[col["Kernel Name"]][:110], "grid", r[col.get("Grid Size", 0)] if "Grid Size" in col else "")
        for k in KEYS:
            if k in col:
                print(f"   {k:72s} {r[col[k]]:>16s} {units[col[k]]}")


if __name__ == "__main__":
    {"launches": launches, "rep": rep}[sys.argv[1]](sys.argv[2])
