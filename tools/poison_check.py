"""Debug aid: does any result of the read path depend on memory the library never wrote? Poisons most of the free device
memory with a byte value (allocate, fill, free), then runs the scenario of
tests/test_reads_gpu.py::test_reads_from_a_materialised_haplotype_never_leave_the_device against the oracle, for several
poison values in one process."""
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "population-sv-simulator_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import oracle_helpers as oh  # noqa: E402
from synthpop import lib as sp  # noqa: E402
import test_reads_gpu as T  # noqa: E402

sp.load()
oracle_lib = oh.load_oracle()
tmp = tempfile.mkdtemp()


class P:   # pathlib-like tmp_path for T._model
    def __truediv__(self, o):
        import pathlib
        return pathlib.Path(tmp) / o


def poison(value, gib=40):
    bufs = []
    for _ in range(gib):
        bufs.append(torch.full((1 << 30,), value, dtype=torch.uint8, device="cuda"))
    torch.cuda.synchronize()
    del bufs
    torch.cuda.empty_cache()


rng = np.random.default_rng(77)
case = oh.random_edit_lists(rng, 400_000, 200, 2, max_len=3000)
model = T._model(P(), "hmm")
p = dict(T.CCS, depth=3.0)
wants = []
for h in range(2):
    hap = oh.build_haplotype(oracle_lib, case["ref"], case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"], None, case["lists"][h])
    wants.append(oh.pbsim_simulate(oracle_lib, hap, model, seed=21, per_read_streams=True, ref_num=h + 1, **p))
bad_any = False
for value in (0x00, 0xA5, 0xFF, 0x4E, 0x41, 0x01):
    poison(value)
    with sp.Context(0, 42) as ctx:
        ctx.set_reference(case["ref"])
        ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
        ctx.set_haplotype_lists(case["list_off"], case["ids"])
        ctx.plan_haplotypes(0, 1, False)
        ctx.materialise()
        T._configure(sp, ctx, p, 21, model, False, sp.READS_PER_READ_STREAMS)
        for h in range(2):
            want, log = wants[h]
            res = ctx.reads_simulate_haplotype(h, ref_num=h + 1)
            got = ctx.reads_copy(res["file_bytes"])
            ok = res["n_reads"] == log["n_reads"] and got.size == want.size and bool((got == want).all())
            msg = ""
            if not ok:
                bad_any = True
                if got.size == want.size:
                    bad = np.nonzero(got != want)[0]
                    msg = f"first diff at {int(bad[0])} of {got.size}, {bad.size} bytes differ: got {got[max(0, bad[0]-40):bad[0]+40].tobytes()!r} want {want[max(0, bad[0]-40):bad[0]+40].tobytes()!r}"
                else:
                    msg = f"sizes {got.size} vs {want.size}, reads {res['n_reads']} vs {log['n_reads']}, res {res}"
            print(f"poison 0x{value:02X} haplotype {h}: {'ok' if ok else 'MISMATCH ' + msg}", flush=True)
print("RESULT", "mismatch" if bad_any else "all ok")
