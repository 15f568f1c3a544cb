"""The read kernel's per-read code (population-sv-simulator_b200/csrc/reads_core.cuh) compiled for the HOST and run one
lane at a time (tests/emu/reads_emu.cu), against the pbsim2 restatement (oracle/pbsim_port.cpp, itself pinned to the
reference binary by tests/test_pbsim_port.py). A logic check that needs no GPU: ring rotation, one draw per step, error
events, the record's aligned output words, the dropped-read rule. The emulator is test infrastructure — the library never
runs this code on the host; what only a GPU can show (warp votes, shared memory, the launch) is tests/test_reads_gpu.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_helpers as oh
from test_reads_gpu import CCS, _cli, _haplotype, _model

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_DIR = os.path.join(ROOT, "tests", "emu")


@pytest.fixture(scope="module")
def emu():
    subprocess.run(["make", "-C", EMU_DIR], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(EMU_DIR, "libreads_emu.so"))
    lib.emu_reads.restype = C.c_int
    return lib


def _run(emu, params, seed, model, hap, no_error, mode, ref_num=1, draws=None, offsets=None):
    from synthpop import lib as sp
    args, _ = sp.ReadArgs.from_options(_cli(params, seed, model), no_error=no_error, stream_mode=mode)
    cap = int(hap.size * params["depth"] * 1.3) + (1 << 20)
    log_cap = int(hap.size * params["depth"] / max(1, params["len_min"])) + 1024
    out = np.empty(cap, np.uint8)
    ln = np.zeros(log_cap, np.int32); dr = np.zeros(log_cap, np.uint8); us = np.zeros(log_cap, np.uint32)
    nb = C.c_int64(); nr = C.c_int64()
    err = C.create_string_buffer(512)
    d = np.ascontiguousarray(draws, np.int32) if draws is not None else np.zeros(1, np.int32)
    o = np.ascontiguousarray(offsets, np.int64) if offsets is not None else np.zeros(1, np.int64)
    hap = np.ascontiguousarray(hap, np.uint8)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = emu.emu_reads(C.byref(args), vp(hap), C.c_int64(hap.size), C.c_int64(ref_num), vp(d),
                       C.c_int64(d.size if draws is not None else 0), vp(o), C.c_int64(o.size if offsets is not None else 0),
                       vp(out), C.c_int64(cap), C.byref(nb), C.byref(nr), vp(ln), vp(dr), vp(us), C.c_int64(log_cap), err, 512)
    if rc != 0:
        raise RuntimeError(err.value.decode())
    n = nr.value
    return out[:nb.value].copy(), ln[:n].copy(), dr[:n].copy(), us[:n].copy()


CLR = dict(depth=3.0, len_mean=9000.0, len_sd=7000.0, len_min=100, len_max=1000000, accuracy_mean_cli=0.85, ratio=(6, 54, 40))


@pytest.mark.parametrize("kind,no_error,seed,params", [
    ("empty", False, 42, CCS), ("hmm", False, 42, CCS), ("hmm", True, 1, CCS), ("hmm_sparse", False, 5, CCS),
    ("empty", False, 8, CLR), ("hmm", False, 3, dict(CCS, len_mean=7000.0, len_sd=0.0, accuracy_mean_cli=0.9, depth=4.0)),
])
def test_replay_through_the_kernel_code_on_the_host(emu, oracle_lib, tmp_path, kind, no_error, seed, params):
    from synthpop import lib as sp
    hap = _haplotype(300_000)
    model = _model(tmp_path, kind)
    want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=seed, no_error=no_error, **params)
    draws = oh.glibc_rand(oracle_lib, seed, log["total_draws"])
    got, ln, dr, used = _run(emu, params, seed, model, hap, no_error, sp.READS_REPLAY, draws=draws, offsets=log["draw_offset"])
    assert np.array_equal(ln, log["length"]) and np.array_equal(dr, log["dropped"])
    assert np.array_equal(used.astype(np.int64), log["draws_used"])
    assert got.size == want.size and np.array_equal(got, want)


@pytest.mark.parametrize("kind,no_error,seed,params", [
    ("empty", False, 11, CCS), ("hmm", False, 12, CCS), ("hmm", True, 13, CCS), ("hmm_sparse", False, 14, CCS),
    ("hmm", False, 15, CLR), ("empty", False, 16, dict(CCS, len_min=2, len_mean=40.0, len_sd=30.0, len_max=200, depth=0.5)),
    ("hmm", False, 77, dict(CCS, len_min=8, len_max=30000, len_mean=2500.0, len_sd=3000.0, depth=3.0)),
    ("hmm", True, 78, dict(CCS, len_min=8, len_max=30000, len_mean=2500.0, len_sd=3000.0, depth=3.0)),
])
def test_per_read_streams_through_the_kernel_code_on_the_host(emu, oracle_lib, tmp_path, kind, no_error, seed, params):
    """The generator in its shared-memory-column form: compile-time ring positions, two rotations per read."""
    from synthpop import lib as sp
    hap = _haplotype(300_000, seed=9)
    model = _model(tmp_path, kind)
    want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=seed, no_error=no_error, per_read_streams=True, ref_num=3, **params)
    got, ln, dr, used = _run(emu, params, seed, model, hap, no_error, sp.READS_PER_READ_STREAMS, ref_num=3)
    assert np.array_equal(ln, log["length"]) and np.array_equal(dr, log["dropped"])
    assert np.array_equal(used.astype(np.int64), log["draws_used"])
    assert got.size == want.size and np.array_equal(got, want)
