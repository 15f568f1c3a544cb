"""Pins the oracle's pbsim2 restatement (oracle/pbsim_port.cpp, incl. its restated glibc rand()) against the
reference itself: oracle/_ref/pbsim2 and pbsim2noError are the UNMODIFIED reference sources compiled in place.
Byte-identical read files for the same arguments and seed. CPU only."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
PORT = os.path.join(ROOT, "oracle", "_build", "pbsim_port")

pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "pbsim2")),
                                reason="oracle/_ref/pbsim2 not built (needs /root/reference or the prebuilt binary)")

BASE = ["--length-min", "5000", "--length-max", "30000", "--accuracy-min", "0.99", "--accuracy-max", "1.0",
        "--difference-ratio", "6:50:54", "--length-mean", "15000", "--length-sd", "2000", "--accuracy-mean", "0.995"]


def _haplotype(tmp_path, L=600_000, seed=3):
    from synthpop import synth
    bases, _, _ = synth.random_reference(L, seed)
    fa = str(tmp_path / "haplotype_0.fa")
    with open(fa, "wb") as f:  # PrintHaplotypes layout: one header, one long sequence line
        f.write(b">chr1_0\n" + bases.tobytes() + b"\n")
    return fa


@pytest.mark.parametrize("model,noerr,depth,seed", [("empty", False, 6, 42), ("empty", True, 4, 7), ("hmm", False, 5, 42),
                                                     ("hmm", True, 3, 1), ("hmm_partial", False, 5, 99)])
def test_port_matches_compiled_reference(oracle_lib, tmp_path, model, noerr, depth, seed):
    from synthpop import synth
    fa = _haplotype(tmp_path)
    mf = str(tmp_path / "m.model")
    if model == "empty":
        open(mf, "w").close()
    elif model == "hmm":
        synth.write_hmm_model(mf)
    else:
        synth.write_hmm_model(mf, accuracies=range(90, 99))  # other accuracy classes take the uniform branch
    args = BASE + ["--depth", str(depth), "--seed", str(seed), "--hmm_model", mf]
    exe = os.path.join(REF, "pbsim2noError" if noerr else "pbsim2")
    subprocess.run([exe] + args + ["--prefix", str(tmp_path / "ref"), fa], check=True, stderr=subprocess.DEVNULL)
    subprocess.run([PORT] + args + (["--no-error"] if noerr else []) + ["--prefix", str(tmp_path / "port"),
                                                                      "--log", str(tmp_path / "log.tsv"), fa], check=True)
    a = open(tmp_path / "ref_0001.fastq", "rb").read()
    b = open(tmp_path / "port_0001.fastq", "rb").read()
    assert len(a) > 1_000_000 and a == b
    log = np.loadtxt(tmp_path / "log.tsv", dtype=np.int64, ndmin=2)
    assert (np.diff(log[:, 2]) > 0).all() and log[0, 2] == 0
    assert log[:, 3].sum() >= int(depth * 600_000)


def test_port_other_parameters(oracle_lib, tmp_path):
    """The reference's own runs use sd 4000 (wdl/inputs/SimulateHaplotypes_i300_noPav.json:10-13); plus pbsim defaults."""
    fa = _haplotype(tmp_path, 400_000, seed=11)
    mf = str(tmp_path / "m.model"); open(mf, "w").close()
    for k, args in enumerate([["--length-min", "5000", "--length-max", "30000", "--length-mean", "12500", "--length-sd", "4000",
                               "--accuracy-mean", "0.995", "--difference-ratio", "6:50:54", "--depth", "5", "--seed", "5"],
                              ["--depth", "3", "--seed", "8"],
                              ["--length-sd", "0", "--length-mean", "7000", "--accuracy-mean", "0.9", "--depth", "4", "--seed", "3"]]):
        args = args + ["--hmm_model", mf]
        subprocess.run([os.path.join(REF, "pbsim2")] + args + ["--prefix", str(tmp_path / f"r{k}"), fa], check=True, stderr=subprocess.DEVNULL)
        subprocess.run([PORT] + args + ["--prefix", str(tmp_path / f"p{k}"), fa], check=True)
        assert open(tmp_path / f"r{k}_0001.fastq", "rb").read() == open(tmp_path / f"p{k}_0001.fastq", "rb").read(), k
