"""Small driver for ncu captures of the tiled copy kernel in its record-gather role (read mix): two 40 Mbp haplotypes,
10x CCS reads each, one mixed file of ~0.8 GB. Usage under ncu:
  ncu --set full --clock-control none --import-source on -k regex:materialise_kernel --launch-skip 1 --launch-count 1 \
      -o gpurun_out/mix python tools/profile_mix.py"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "population-sv-simulator_b200"))
from synthpop import lib as sp, synth  # noqa: E402

L = int(os.environ.get("PROFILE_L", 40_000_000))
depth = int(os.environ.get("PROFILE_DEPTH", 10))
work = tempfile.mkdtemp()
model = os.path.join(work, "m.model")
synth.write_hmm_model(model, quality="phred")
rng = np.random.default_rng(1)
ref = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=L, dtype=np.uint8)].copy()
with sp.Context(0, 42) as ctx:
    ctx.set_reference(ref)
    z = np.zeros(0, np.int32)
    ctx.set_variants(np.zeros(0, np.int8), z, z, np.zeros(1, np.int64), np.zeros(0, np.uint8))
    ctx.set_haplotype_lists(np.zeros(3, np.int64), z)
    opts = ["--length-min", "5000", "--length-max", "30000", "--accuracy-min", "0.99", "--accuracy-max", "1.0",
            "--difference-ratio", "6:50:54", "--length-mean", "15000", "--length-sd", "2000", "--accuracy-mean", "0.995",
            "--depth", str(depth), "--hmm_model", model, "--seed", "42"]
    args, _ = sp.ReadArgs.from_options(opts, stream_mode=sp.READS_PER_READ_STREAMS)
    ctx.reads_configure(args)
    n, b = ctx.haplotype2reads(0, 42)
    t = ctx.reads_mix_timing()
    print(f"mixed {n} reads, {b} bytes, order {t['order_ms']:.3f} ms, copy {t['copy_ms']:.3f} ms, "
          f"{2 * b / t['copy_ms'] / 1e6:.0f} GB/s algorithmic")
