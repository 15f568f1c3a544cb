#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an ncu report.

`ncu --page source --csv` prints the SASS view (one row per instruction with "Instructions Executed" and the stall
samples) but no source line; `nvdisasm -g` on the cubin prints the same instructions with `//## File ..., line N`
markers. The two listings are joined by instruction order.

  python tools/ncu_lines.py <report.ncu-rep> <lib.so> <kernel-substring> [top]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def sass_rows(report):
    out = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci = hdr.index("Instructions Executed")
    cs = hdr.index("# Samples")
    return [(r[1].strip(), int(r[ci] or 0), int(r[cs] or 0)) for r in rows[hdr_i + 1:] if len(r) > ci]


def line_table(so, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    for f in sorted(os.listdir(tmp)):
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        if kernel not in txt:
            continue
        lines, cur, inside = [], None, False
        for ln in txt.splitlines():
            if ln.startswith(".text.") or re.match(r"\s*\.section\s+\.text\.", ln):
                inside = kernel in ln
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                lines.append(cur)
        if lines:
            return lines
    raise SystemExit("kernel not found in any cubin")


def main():
    report, so, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    sass = sass_rows(report)
    lt = line_table(so, kernel)
    if len(sass) != len(lt):
        print(f"warning: {len(sass)} SASS rows in the report, {len(lt)} in the cubin (rebuilt since the capture?)")
    agg, smp = defaultdict(int), defaultdict(int)
    for (txt, n, s), loc in zip(sass, lt):
        agg[loc] += n
        smp[loc] += s
    tot, tots = sum(agg.values()), sum(smp.values())
    print(f"total warp instructions {tot}, samples {tots}")
    for loc, n in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{100 * n / tot:6.2f}% inst {100 * smp[loc] / max(tots, 1):6.2f}% samples  {loc}")


if __name__ == "__main__":
    main()
