#!/usr/bin/env python3
"""Summarise ncu output brought back in gpurun_out/ into the tables kept under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches_<workload>.csv ...
    python tools/ncu_summary.py full gpurun_out/full_<kernel>.raw.csv ...
"""
import collections
import csv
import sys

FULL = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__bytes_read.sum.per_second", "dram read rate"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64 pipe active %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm throughput %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2 throughput %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1/tex throughput %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__cluster_size", "cluster"),
    ("launch__registers_per_thread", "regs/thread"),
    ("sm__cycles_elapsed.max.per_second", "sm clock"),
]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        nm = r[ki].split("(")[0].replace("void ", "")
        v = float(r[vi].replace(",", ""))
        us = v / 1e3 if r[ui].startswith("n") else (v if r[ui].startswith("u") else v * 1e3)
        a = agg.setdefault(nm, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print(f"\n### {path}  (sum of kernel time {tot / 1e3:.2f} ms)\n")
    print("| kernel | launches | mean us | share |\n|---|---|---|---|")
    for nm, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{nm}` | {c} | {t / c:.1f} | {100 * t / tot:.1f} % |")


def full(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print(f"\n### {path}: `{r[hdr.index('Kernel Name')][:70]}`\n")
        print("| metric | value |\n|---|---|")
        for key, label in FULL:
            if key in hdr:
                i = hdr.index(key)
                print(f"| {label} (`{key}`) | {r[i]} {units[i]} |")


if __name__ == "__main__":
    fn = {"launches": launches, "full": full}[sys.argv[1]]
    for p in sys.argv[2:]:
        fn(p)
