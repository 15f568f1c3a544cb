"""SURVEY.md §8e on real GPUs: SimulateHaplotypes sharded by individual over 2 (or more) ranks with the NCCL all-gather
of carrier counts must write exactly the files of the single-process CPU restatement. Skipped on a 1-GPU box."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_two_rank_population_matches_oracle(sp, oracle_bin, tmp_path):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 4 if n >= 4 else 2
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 3_000_000, n_insertions=3000)
    want, got, hw, hg = (str(tmp_path / d) for d in ("want", "got", "hw", "hg"))
    for d in (want, got, hw, hg):
        os.makedirs(d)
    n_ind = 7  # uneven shards
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(want, n_ind, n_frequent=400), check=True)
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", str(2 * n_ind - 1), inp.reference_file,
                    want + "/haplotype2variants.txt", want + "/variants.txt", hw], check=True)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(HERE, "dist_population_worker.py"), hg] + inp.simulate_argv(got, n_ind, n_frequent=400)
    subprocess.run(cmd, check=True, timeout=600)
    names = ["groundTruth_joint.vcf", "haplotype2variants.txt", "variants.txt", "histogram_nVariantsPerHaplotype.txt",
             "histogram_startPositions.txt"] + [f"groundTruth_individual_{2 * k}.vcf" for k in range(n_ind)]
    for name in names:
        assert open(os.path.join(got, name), "rb").read() == open(os.path.join(want, name), "rb").read(), name
    for h in range(2 * n_ind):
        assert open(f"{hg}/haplotype_{h}.fa", "rb").read() == open(f"{hw}/haplotype_{h}.fa", "rb").read(), h


def test_owner_broadcast_population_matches_oracle(sp, oracle_bin, tmp_path):
    """One owner per run (tools/whole_genome.py --mode owner): the last rank runs SimulateHaplotypes alone and broadcasts the
    reference (device to device) and the tables; every rank's haplotypes must be the oracle's PrintHaplotypes."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 4 if n >= 4 else 2
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 2_500_000, n_insertions=3000, seed=31)
    want, got, hw, hg = (str(tmp_path / d) for d in ("want", "got", "hw", "hg"))
    for d in (want, got, hw, hg):
        os.makedirs(d)
    n_ind = 9
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(want, n_ind, n_frequent=400), check=True)
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", str(2 * n_ind - 1), inp.reference_file,
                    want + "/haplotype2variants.txt", want + "/variants.txt", hw], check=True)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29733", os.path.join(HERE, "dist_population_worker.py"), hg] + inp.simulate_argv(got, n_ind, n_frequent=400)
    subprocess.run(cmd, check=True, timeout=600, env=dict(os.environ, SYNTHPOP_TEST_OWNER_MODE="1"))
    for name in ("groundTruth_joint.vcf", "haplotype2variants.txt", "variants.txt"):
        assert open(os.path.join(got, name), "rb").read() == open(os.path.join(want, name), "rb").read(), name
    for h in range(2 * n_ind):
        assert open(f"{hg}/haplotype_{h}.fa", "rb").read() == open(f"{hw}/haplotype_{h}.fa", "rb").read(), h
