#!/usr/bin/env python
"""DRAM bytes of bench.py's own materialise_kernel launches -> profiles/r02_traffic.json (what bench.py's `roofline.traffic` quotes).

On the GPU box (one call; the plain run first, as the profiling recipe asks):
  python bench.py --steps 6 --warmup 3 --no-cpu-baseline --read-haplotypes 0 --length-bins 0 --e2e-haplotypes 0 --file-emission 0 \
      > gpurun_out/traffic_plain.log 2>&1 && \
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:materialise_kernel \
      --csv --log-file gpurun_out/r02_traffic_launches.csv python bench.py --steps 6 --warmup 3 --no-cpu-baseline \
      --read-haplotypes 0 --length-bins 0 --e2e-haplotypes 0 --file-emission 0 > gpurun_out/traffic_ncu.log 2>&1
Here:  python tools/capture_traffic.py gpurun_out/r02_traffic_launches.csv gpurun_out/traffic_plain.log
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    launches_csv, plain_log = sys.argv[1], sys.argv[2]
    rows = [r for r in csv.reader(open(launches_csv)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    iid, iname, imetric, iunit, ival = (hdr.index(k) for k in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value"))
    per = {}
    for r in rows[1:]:
        if "materialise_kernel" not in r[iname]:
            continue
        v = float(r[ival].replace(",", ""))
        u = r[iunit].lower()
        scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0,
                 "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}.get(u, 1)
        per.setdefault(r[iid], {})[r[imetric]] = v * scale
    line = json.loads([l for l in open(plain_log).read().splitlines() if l.startswith("{")][-1])
    # the launches of the timed steps (the batch launches; the read/e2e legs are switched off in this capture)
    full = [p for p in per.values() if p.get("dram__bytes_write.sum", 0) > 1e9]
    rd = sum(p["dram__bytes_read.sum"] for p in full) / len(full)
    wr = sum(p["dram__bytes_write.sum"] for p in full) / len(full)
    ms = sum(p["gpu__time_duration.sum"] for p in full) / len(full) * 1e3
    out = {"capture": os.path.basename(launches_csv) + " (ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum of bench.py's own steps)",
           "kernel": "materialise_kernel", "launches_averaged": len(full),
           "batch": line["arm"]["haplotypes_per_step_per_gpu"], "reference_length": 248956422,
           "bases_per_launch": line["roofline"]["bases_per_launch"],
           "dram_bytes_read_per_launch": rd, "dram_bytes_write_per_launch": wr, "dram_bytes_per_launch": rd + wr,
           "ncu_ms_per_launch_cold_serialised": ms,
           "dram_bytes_per_output_base": (rd + wr) / line["roofline"]["bases_per_launch"]}
    with open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
