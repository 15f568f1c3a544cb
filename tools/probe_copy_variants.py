"""A/B of the copy-kernel designs on the shape of bench.py's step (125 chr1-sized slots from one shared source, region-major,
source shifted by 9 bytes as a FASTA header does): stores only, the register path the product uses, and the bulk-copy engine
(cp.async.bulk + mbarrier). One JSON line; numbers are bytes WRITTEN per second (the DRAM side of the launch is writes)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "population-sv-simulator_b200"))
from synthpop import lib as sp  # noqa: E402

L = int(os.environ.get("PROBE_L", 248_956_422))
N = int(os.environ.get("PROBE_SLOTS", 125))
out = {"slot_bytes": L, "n_slots": N}
with sp.Context(0, 42) as ctx:
    for kind, name in ((0, "store_only"), (1, "register_copy"), (2, "bulk_copy_mbarrier"), (3, "bulk_copy_warp_pipelines"),
                       (4, "source_stationary_all_slots"), (5, "source_stationary_25_slots"), (6, "source_stationary_5_slots"),
                       (7, "store_only_hashed_data")):
        if os.environ.get("PROBE_KINDS") and str(kind) not in os.environ["PROBE_KINDS"].split(","):
            continue
        for mis in ((0,) if kind in (0, 4, 5, 6, 7) else (0, 9)):
            ms = ctx.probe_bandwidth(kind, L, N, mis, int(os.environ.get("PROBE_REPS", 5)))
            out[f"{name}_misalign{mis}"] = {"ms": round(ms, 4), "written_GBps": round(L * N / ms / 1e6, 1)}
            print(name, mis, out[f"{name}_misalign{mis}"], file=sys.stderr, flush=True)
print(json.dumps(out))
