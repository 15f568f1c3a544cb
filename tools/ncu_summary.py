#!/usr/bin/env python
"""Compact text summary of every kernel in an ncu report: the raw metrics the read-kernel work was steered by, the stall
samples by instruction pattern (unrolled copies merged) and the single instructions that wait longest.

  python tools/ncu_summary.py <report.ncu-rep> [top]
"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "sm__icc_request_hit_rate.pct"]


def main():
    report = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    raw = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)))
    hdr, units = raw[0], raw[1]
    src = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", report, "--page", "source", "--csv"], capture_output=True, text=True).stdout)))
    groups, cur = [], None
    for r in src:
        if r and r[0] == "Address":
            cur = [r]; groups.append(cur)
        elif cur is not None and len(r) > 5:
            cur.append(r)
    for n, row in enumerate(raw[2:]):
        print(f"== {row[hdr.index('Kernel Name')]}  grid {row[hdr.index('Grid Size')]} block {row[hdr.index('Block Size')]}")
        for k in KEYS:
            if k in hdr:
                print(f"   {k} = {row[hdr.index(k)]} {units[hdr.index(k)]}")
        for k in hdr:
            if "issue_stalled" in k and k.endswith("_per_issue_active.ratio"):
                try:
                    v = float(row[hdr.index(k)])
                except ValueError:
                    continue
                if v >= 0.2:
                    print(f"   stall {k.split('issue_stalled_')[1].split('_per_issue')[0]} = {v:.2f} warps per issue")
        if n >= len(groups):
            continue
        g = groups[n]
        ci, cs = g[0].index("Instructions Executed"), g[0].index("# Samples")
        data = g[1:]
        tot = sum(int(r[cs] or 0) for r in data) or 1
        toti = sum(int(r[ci] or 0) for r in data) or 1
        agg = defaultdict(lambda: [0, 0, 0])
        for r in data:
            t = re.sub(r"0x[0-9a-f]+", "IMM", r[1].strip())
            t = re.sub(r"\bU?R\d+", "R", t); t = re.sub(r"\bU?P\d", "P", t); t = re.sub(r"\bB\d", "B", t)
            a = agg[t]; a[0] += int(r[cs] or 0); a[1] += int(r[ci] or 0); a[2] += 1
        print(f"   {len(data)} SASS instructions, {toti} warp instructions executed, {tot} stall samples; by instruction pattern:")
        for t, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
            print(f"     {100 * a[0] / tot:6.2f}% samples {100 * a[1] / toti:6.2f}% instr  x{a[2]:<4d} {t[:80]}")
        print("   single instructions that wait longest:")
        for i, r in sorted(enumerate(data), key=lambda ir: -int(ir[1][cs] or 0))[:top]:
            print(f"     #{i:<6d} {100 * int(r[cs] or 0) / tot:6.2f}% samples {int(r[ci] or 0):>12d} executions  {r[1].strip()[:80]}")


if __name__ == "__main__":
    main()
