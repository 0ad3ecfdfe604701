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
           "launch__block_size", "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "sm__inst_executed.avg.per_cycle_elapsed",
           "smsp__inst_executed.sum", "l1tex__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
           "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.per_cycle_active"]


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
        stalls = []
        for h, i in idx.items():
            if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
                try:
                    stalls.append((float(d[i]), h[len("smsp__pcsamp_warps_issue_stalled_"):]))
                except ValueError:
                    pass
        tot = sum(v for v, _ in stalls) or 1.0
        out.write("   warp-state samples: " + ", ".join("%s %.0f%%" % (n, 100 * v / tot) for v, n in sorted(stalls, reverse=True)[:7]) + "\n")


if __name__ == "__main__":
    with open(sys.argv[3], "w") as f:
        if sys.argv[1] != "-":
            launches(sys.argv[1], f)
        for rep in sys.argv[2].split(","):
            full(rep, f)
    print(open(sys.argv[3]).read())
