"""The C++ command-line host above the C ABI (population-sv-simulator_b200/host/synthpop_main.cpp): the reference's class
names and positional argv (src/PrintHaplotypes.java:7-12, docker/haplotype2reads.sh:16-24,53-64), checked against the
oracle's files."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import read_mix as rm  # noqa: E402
import oracle_helpers as oh  # noqa: E402

pytestmark = pytest.mark.gpu
CLI = os.path.join(ROOT, "population-sv-simulator_b200", "synthpop_cli")


def test_print_haplotypes_and_haplotype2reads_commands(sp, oracle_bin, oracle_lib, tmp_path):
    from synthpop import synth
    assert os.path.exists(CLI), "synthpop_cli is missing: run __graft_entry__.build()"
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 1_200_000, n_insertions=1500, seed=8)
    want = str(tmp_path / "want"); hw = str(tmp_path / "hw"); hg = str(tmp_path / "hg")
    for d in (want, hw, hg):
        os.makedirs(d)
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(want, 3, n_frequent=200), check=True)
    files = [inp.reference_file, want + "/haplotype2variants.txt", want + "/variants.txt"]
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "2", "3"] + files + [hw], check=True)
    subprocess.run([CLI, "PrintHaplotypes", "2", "3"] + files + [hg], check=True)
    for h in (2, 3):
        assert open(f"{hw}/haplotype_{h}.fa", "rb").read() == open(f"{hg}/haplotype_{h}.fa", "rb").read()
    # a range outside the population fails like the Java (uncaught exception, exit status 1)
    r = subprocess.run([CLI, "PrintHaplotypes", "5", "9"] + files + [hg], capture_output=True)
    assert r.returncode == 1 and b"Exception in thread" in r.stderr
    # haplotype2reads: individual (2,3) at 4x, seed 11 for pbsim and the shuffle
    model = str(tmp_path / "hmm.model"); synth.write_hmm_model(model)
    out = str(tmp_path / "reads_i2_i3.fa")
    subprocess.run([CLI, "--seed", "11", "haplotype2reads", "2", "5000", "30000", "15000", "2000", "4"] + files + [model, out], check=True)
    p = dict(depth=4.0, len_mean=15000.0, len_sd=2000.0, len_min=5000, len_max=30000, accuracy_mean_cli=0.995, ratio=(6, 50, 54))
    recs = []
    for h in (2, 3):
        hap = np.frombuffer(open(f"{hw}/haplotype_{h}.fa", "rb").read().split(b"\n")[1], np.uint8)
        w, _ = oh.pbsim_simulate(oracle_lib, hap, model, seed=sp.haplotype_read_seed(11, h), per_read_streams=True, ref_num=1, **p)
        recs.append(w.tobytes())
    assert open(out, "rb").read() == rm.mix_reads(recs, 11)


def test_whole_genome_driver_small(tmp_path):
    """tools/whole_genome.py (BASELINE config 3) at 0.4 % scale: three chromosomes, rank 1 of 2 owns individuals 4..7."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "whole_genome.py"), "--scale", "0.004", "--chromosomes", "3",
                        "--individuals", "8", "--as-world", "2", "--as-rank", "1"], capture_output=True, check=True, timeout=300)
    line = json.loads(r.stdout.decode().strip().splitlines()[-1])
    rows = line["chromosomes"]
    assert [c["chromosome"] for c in rows] == ["chr1", "chr2", "chr3"]
    assert all(c["haplotypes"] == 8 and c["variants"] > 0 for c in rows)
    assert line["bases"] == sum(c["bases"] for c in rows) and line["gpu_launches"] > 0
    assert all(c["bases"] > 0 and c["copy_ms"] > 0 for c in rows) and line["wall_s"] > 0 and line["prologue_threads"] >= 1
