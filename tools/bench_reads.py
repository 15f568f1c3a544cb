"""Read-kernel timing on one synthetic haplotype (tuning aid; not the bench). PROFILE_L bases, PROFILE_DEPTH x."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "population-sv-simulator_b200"))
from synthpop import lib as sp, synth  # noqa: E402

L = int(os.environ.get("PROFILE_L", 248_956_422))
depth = int(os.environ.get("PROFILE_DEPTH", 30))
work = tempfile.mkdtemp()
model = os.path.join(work, "m.model")
synth.write_hmm_model(model, quality=os.environ.get("PROFILE_QUALITY", "phred"))
bases, _, _ = synth.random_reference(L, 7)
with sp.Context(0, 42) as ctx:
    opts = ["--length-min", "5000", "--length-max", "30000", "--accuracy-min", "0.99", "--accuracy-max", "1.0",
            "--difference-ratio", os.environ.get("PROFILE_RATIO", "6:50:54"), "--length-mean", "15000", "--length-sd", "2000", "--accuracy-mean", "0.995",
            "--depth", str(depth), "--hmm_model", model, "--seed", "42"]
    args, _ = sp.ReadArgs.from_options(opts, stream_mode=sp.READS_PER_READ_STREAMS)
    ctx.reads_configure(args)
    # PROFILE_VARIANTS: comma list of <design>:<min CTAs per SM>[:<prefetch bits>] (csrc/reads.cu pick_read_kernel), A/B in one process
    for variant in os.environ.get("PROFILE_VARIANTS", "2:10").split(","):
        design, min_ctas, pf = (variant.split(":") + ["0"])[:3]
        os.environ["SYNTHPOP_READS_KERNEL"] = design
        os.environ["SYNTHPOP_READS_MIN_CTAS"] = min_ctas
        os.environ["SYNTHPOP_READS_PF"] = pf
        for it in range(int(os.environ.get("PROFILE_ITERS", 3))):
            r = ctx.reads_simulate_sequence(bases)
            t = ctx.reads_timing()
            print(f"design {design} min_ctas {min_ctas} pf {pf}: reads {r['n_reads']} bases {r['bases']} kernel {t['kernel_ms']:.2f} ms "
                  f"plan {t['plan_ms']:.2f} ms -> {r['bases'] / t['kernel_ms'] / 1e6:.1f} Gbp/s", flush=True)
