#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into text files under profiles/.

usage: python profiles/summarize.py <launches.csv> <prof.ncu-rep> <out_prefix>
  launch list : ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file launches.csv <cmd>
  full capture: ncu --set full --clock-control none --import-source on -k regex:... -o prof <cmd>
"""
import collections
import csv
import subprocess
import sys


def launches(path, out):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    H, data = rows[hdr], rows[hdr + 1:]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in data:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        u = r[ui]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        agg.setdefault(r[ki], []).append(v)
    ours = {k: v for k, v in agg.items() if "rf::" in k or k.startswith("k_")}
    tot_all = sum(sum(v) for v in agg.values())
    tot = sum(sum(v) for v in ours.values())
    with open(out, "w") as f:
        f.write(f"# per-launch device time (us), cold-cache & serialised under ncu: compare SHARES, not absolutes\n")
        f.write(f"# all kernels {tot_all:.1f} us; rfgpu kernels {tot:.1f} us (the rest is torch set-up: sort / de-dup of the index build)\n")
        f.write(f"{'kernel':70s} {'n':>4s} {'sum_us':>10s} {'share_of_rfgpu':>14s} {'max_us':>9s}\n")
        for k, v in sorted(ours.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"{k[:70]:70s} {len(v):4d} {sum(v):10.1f} {sum(v) / tot:14.3f} {max(v):9.1f}\n")


WANT = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def full(path, out):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    H = rows[0]
    idx = [H.index(w) for w in WANT if w in H]
    with open(out, "w") as f:
        f.write("# ncu --set full --clock-control none, one row per captured launch\n")
        f.write(" | ".join(H[i] for i in idx) + "\n")
        f.write(" | ".join(rows[1][i] for i in idx) + "\n")
        for r in rows[2:]:
            f.write(" | ".join(r[i][:60] for i in idx) + "\n")


if __name__ == "__main__":
    l, p, prefix = sys.argv[1:4]
    launches(l, prefix + "_launches.txt")
    full(p, prefix + "_ncu_full.txt")
