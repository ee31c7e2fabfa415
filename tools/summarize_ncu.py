#!/usr/bin/env python
"""Turns the ncu outputs brought back in gpurun_out/ into the small CSV summaries kept under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches_a.csv  > profiles/rNN_launches_summary.csv
  python tools/summarize_ncu.py full     gpurun_out/gemm_full_a.ncu-rep > profiles/rNN_gemm_full_summary.csv
"""
import collections
import csv
import subprocess
import sys

FULL_METRICS = [
    "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct",
    "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "lts__t_sectors_srcunit_tex_op_red.sum", "lts__t_sectors_srcunit_tex_op_red.avg.pct_of_peak_sustained_elapsed",
]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for x in csv.DictReader(lines):
        if x["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(x["Metric Value"].replace(",", ""))
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}[x["Metric Unit"]]
        k = x["Kernel Name"].split("(")[0].replace(", ", ";")       # template arguments: keep the CSV columns aligned
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print("kernel,launches,total_us,avg_us,share")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%s,%d,%.1f,%.2f,%.3f" % (k, v[0], v[1], v[1] / v[0], v[1] / tot))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    hdr, units, rows = r[0], r[1], r[2:]
    idx = [hdr.index(m) for m in FULL_METRICS if m in hdr]
    print(",".join(hdr[i] for i in idx))
    print(",".join(units[i] for i in idx))
    for x in rows:
        print(",".join(x[i].split("(")[0].replace(", ", ";") if hdr[i] == "Kernel Name" else x[i].replace(",", "") for i in idx))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
