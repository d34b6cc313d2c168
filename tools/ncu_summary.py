#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): headline counters, opcode mix,
stall reasons and the hottest SASS lines.  usage: tools/ncu_summary.py rep [kernel-index]"""
import csv
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__inst_executed.sum",
        "smsp__inst_executed.sum", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "lts__t_sector_hit_rate.pct", "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]
for vals in rows[2:]:
    print("=" * 60)
    for h, u, v in zip(hdr, units, vals):
        if h in want:
            print("%-64s %-10s %s" % (h, u, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
# several kernels: blocks start with a "Kernel Name" row
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
for b in blocks:
    hdr, data = b["rows"][0], [r for r in b["rows"][1:] if len(r) == len(b["rows"][0])]
    ix = {h: i for i, h in enumerate(hdr)}
    ti = sum(int(r[ix["Instructions Executed"]]) for r in data)
    ts = sum(int(r[ix["# Samples"]]) for r in data) or 1
    print("=" * 60)
    print(b["name"][:100], "warp-instructions", ti, "samples", ts)
    c, cs = Counter(), Counter()
    for r in data:
        parts = r[ix["Source"]].split()
        op = parts[1] if parts[0].startswith("@") else parts[0]
        c[op.split(".")[0]] += int(r[ix["Instructions Executed"]])
        cs[op.split(".")[0]] += int(r[ix["# Samples"]])
    print("opcode mix:", ", ".join("%s %.1f%%(s%.0f%%)" % (k, 100 * v / ti, 100 * cs[k] / ts) for k, v in c.most_common(16)))
    st = []
    for h in hdr:
        if h.startswith("stall_") and "Not" not in h:
            v = sum(int(r[ix[h]]) for r in data)
            if v > 0.02 * ts:
                st.append("%s %.1f%%" % (h[6:], 100 * v / ts))
    print("stalls:", ", ".join(st))
    hot = sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:14]
    for r in hot:
        reasons = [(int(r[ix[h]]), h[6:]) for h in hdr if h.startswith("stall_") and "Not" not in h]
        reasons.sort(reverse=True)
        print("  %5.1f%%  %-60s %s" % (100 * int(r[ix["# Samples"]]) / ts, r[ix["Source"]].strip()[:60],
                                    ",".join("%s" % n for v, n in reasons[:2] if v)))
    ex = sum(int(r[ix["L1 Wavefronts Shared Excessive"]]) for r in data)
    tot = sum(int(r[ix["L1 Wavefronts Shared"]]) for r in data)
    print("shared wavefronts", tot, "excessive", ex)
