#!/usr/bin/env python
"""Turns gpurun_out/*.ncu-rep + launch-list CSVs into the short text summaries kept in profiles/.
usage: python profiles/summarize.py <launches.csv> <prof.ncu-rep> <out.txt>"""
import csv
import subprocess
import sys
from collections import defaultdict

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "lts__t_sector_hit_rate.pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size", "sm__cycles_elapsed.max", "smsp__cycles_active.avg"]


def launches(path, out):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    c = {h: i for i, h in enumerate(rows[hi])}
    agg = defaultdict(list)
    for r in rows[hi + 2:]:
        if len(r) >= len(rows[hi]):
            agg[r[c["Kernel Name"]].split("(")[0]].append(float(r[c["Metric Value"]]))
    tot = sum(sum(v) for v in agg.values())
    out.write("== launch list (%s): gpu__time_duration per kernel, cold-cache serialised; compare SHARES\n" % path)
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.write("%-64s n=%4d mean=%8.2f us share=%5.1f%%\n" % (k[:64], len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / tot))


def full(path, out):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out.write("\n== ncu --set full (%s)\n" % path)
    for d in rows[2:]:
        out.write("-- %s\n" % d[idx["Kernel Name"]][:100])
        for m in METRICS:
            if m in idx:
                out.write("   %-66s %s %s\n" % (m, d[idx[m]], units[idx[m]]))


if __name__ == "__main__":
    with open(sys.argv[3], "w") as f:
        launches(sys.argv[1], f)
        full(sys.argv[2], f)
    print(open(sys.argv[3]).read())
