#!/usr/bin/env python3
"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed here.

    python profiles/summarize.py launches gpurun_out/launches.csv  > profiles/r01_launches.txt
    python profiles/summarize.py raw gpurun_out/prof.ncu-rep       > profiles/r01_kernel_raw.txt
"""
import collections
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
    "launch__shared_mem_per_block_dynamic", "sm__maximum_warps_per_active_cycle_pct",
]


def launches(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr, data = rows[hi], rows[hi + 1:]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    big = []
    for r in data:
        if len(r) <= mv:
            continue
        v = float(r[mv].replace(",", ""))
        v = v / 1e3 if r[mu] == "ns" else v * 1e3 if r[mu] == "ms" else v * 1e6 if r[mu] == "s" else v
        a = agg.setdefault(r[kn], [0, 0.0, 0.0])
        a[0] += 1; a[1] += v; a[2] = max(a[2], v)
        if v > 200:
            big.append((r[0], r[kn][:60], v))
    tot = sum(a[1] for a in agg.values())
    print(f"# ncu launch list summary of {path}: {len(data)} launches, {tot / 1e3:.2f} ms of kernel time (cold-cache, serialised)")
    for k, (n, t, mx) in agg.items():
        print(f"{n:6d} launches  total {t:12.1f} us  avg {t / n:9.2f} us  max {mx:9.2f} us  share {100 * t / tot:5.1f}%  {k}")
    print("# launches longer than 200 us (id, kernel, us):")
    for b in big:
        print(f"  {b[0]:>5} {b[1]:<60} {b[2]:10.1f}")


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print(f"# ncu --set full capture {path}: {len(data)} launches")
    for i, r in enumerate(data):
        print(f"## launch {i}: {r[hdr.index('Kernel Name')]}")
        for m in METRICS:
            if m in hdr:
                j = hdr.index(m)
                print(f"{m:75s} {r[j]:>18s} {units[j]}")


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
