"""Golden vectors of row P (tests/golden/*, written by tests/golden/make_golden.py from the reference's own pbsim2 /
pbsim2noError binaries): the CPU restatement must reproduce them on any box (no /root/reference needed), and the GPU
read kernel in replay mode must too."""
import gzip
import json
import os

import numpy as np
import pytest

import oracle_helpers as oh

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(d for d in os.listdir(GOLD) if os.path.isdir(os.path.join(GOLD, d)))


def _load(name, tmp_path):
    d = os.path.join(GOLD, name)
    case = json.load(open(os.path.join(d, "case.json")))
    fa = gzip.open(os.path.join(d, "hap.fa.gz")).read()
    lines = fa.split(b"\n")
    hap = np.frombuffer(b"".join(lines[1:]), dtype=np.uint8)
    model = str(tmp_path / "m.model")
    with open(model, "wb") as f:
        f.write(gzip.open(os.path.join(d, "model.txt.gz")).read())
    want = gzip.open(os.path.join(d, "out_0001.fastq.gz")).read()
    o = dict(zip(case["opts"][::2], case["opts"][1::2]))
    params = dict(depth=float(o["--depth"]), len_mean=float(o["--length-mean"]), len_sd=float(o["--length-sd"]),
                  len_min=int(o["--length-min"]), len_max=int(o["--length-max"]), accuracy_mean_cli=float(o["--accuracy-mean"]),
                  ratio=tuple(int(x) for x in o["--difference-ratio"].split(":")))
    return case, hap, model, want, params


def test_golden_cases_exist():
    assert len(CASES) >= 3


@pytest.mark.parametrize("name", CASES)
def test_restatement_reproduces_reference_output(oracle_lib, tmp_path, name):
    case, hap, model, want, params = _load(name, tmp_path)
    got, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=case["seed"], no_error=case["no_error"], **params)
    assert got.tobytes() == want


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_replay_reproduces_reference_output(sp, oracle_lib, tmp_path, name):
    case, hap, model, want, params = _load(name, tmp_path)
    _, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=case["seed"], no_error=case["no_error"], **params)
    draws = oh.glibc_rand(oracle_lib, case["seed"], log["total_draws"])
    args, _ = sp.ReadArgs.from_options(case["opts"] + ["--hmm_model", model, "--seed", str(case["seed"])],
                                       no_error=case["no_error"], stream_mode=sp.READS_REPLAY)
    with sp.Context(0, 42) as ctx:
        ctx.reads_configure(args)
        ctx.reads_set_replay(draws, log["draw_offset"])
        res = ctx.reads_simulate_sequence(hap)
        got = ctx.reads_copy(res["file_bytes"])
    assert got.tobytes() == want
