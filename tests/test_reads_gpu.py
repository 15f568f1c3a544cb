"""Row P on the GPU: the read kernel against the pbsim2 restatement (oracle/pbsim_port.cpp, itself pinned byte-for-byte
against the reference's own docker/pbsim2.cpp compiled unmodified, tests/test_pbsim_port.py) and, where the prebuilt
binary travelled with the snapshot, against that reference binary directly.

Replay mode consumes the reference's draw stream (the values rand() returned after srand(seed)) and the index of
each read's first draw: the output must be the reference's file, byte for byte. Per-read-stream mode is checked against
the restatement run with the same per-read seeds."""
import os
import subprocess

import numpy as np
import pytest

import oracle_helpers as oh

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref")

CCS = dict(depth=5.0, len_mean=15000.0, len_sd=2000.0, len_min=5000, len_max=30000, accuracy_mean_cli=0.995, ratio=(6, 50, 54))


def _cli(p, seed, model, prefix=None):
    a = ["--depth", str(p["depth"]), "--length-mean", str(p["len_mean"]), "--length-sd", str(p["len_sd"]),
         "--length-min", str(p["len_min"]), "--length-max", str(p["len_max"]), "--accuracy-mean", str(p["accuracy_mean_cli"]),
         "--accuracy-min", "0.99", "--accuracy-max", "1.0", "--difference-ratio", ":".join(str(x) for x in p["ratio"]),
         "--hmm_model", model, "--seed", str(seed)]
    if prefix:
        a += ["--prefix", prefix]
    return a


def _model(tmp_path, kind):
    from synthpop import synth
    mf = str(tmp_path / f"{kind}.model")
    if kind == "empty":
        open(mf, "w").close()
    elif kind == "hmm":
        synth.write_hmm_model(mf)
    elif kind == "hmm_partial":
        synth.write_hmm_model(mf, accuracies=range(90, 99))
    elif kind == "hmm_sparse":  # state indices 3, 7, 20, 41: exercises the renumbering of visited states
        _states_model(mf, (3, 7, 20, 41))
    elif kind == "hmm_wide":    # 20 states: beyond the nibble maps, serial tile path
        _states_model(mf, tuple(range(1, 21)))
    return mf


def _states_model(path, states):
    """HMM over the given state indices; emission windows stay below Q30 so that the deletion rejection loop
    (acceptance 10^(-q/10), docker/pbsim2.cpp:2284-2292) keeps the draw log of a test small."""
    rng = np.random.default_rng(5)
    with open(path, "w") as f:
        for acc in range(74, 101):
            ip = rng.dirichlet([2.0] * len(states))
            for s, p in zip(states, ip):
                f.write(f"{acc} IP {s} {p:.6f}\n")
            for s in states:
                w = np.zeros(94); c = int(rng.integers(5, 22)); w[c:c + 8] = rng.random(8) + 0.05; w /= w.sum()
                f.write(f"{acc} EP {s} " + " ".join(f"{x:.6f}" for x in w) + "\n")
            for s in states:
                row = np.zeros(max(states))
                row[[t - 1 for t in states]] = rng.dirichlet([1.5] * len(states))
                f.write(f"{acc} TP {s} " + " ".join(f"{x:.6f}" for x in row) + "\n")


def _haplotype(L, seed=3):
    from synthpop import synth
    bases, _, _ = synth.random_reference(L, seed)  # N runs at both ends and in a centromere block, soft-masked repeats
    return bases


def _configure(sp, ctx, p, seed, model, no_error, mode):
    args, _ = sp.ReadArgs.from_options(_cli(p, seed, model), no_error=no_error, stream_mode=mode)
    ctx.reads_configure(args)


@pytest.mark.parametrize("kind,no_error,seed,params", [
    ("empty", False, 42, CCS), ("hmm", False, 42, CCS), ("hmm", True, 1, CCS), ("hmm_partial", False, 99, CCS),
    ("empty", True, 7, CCS), ("hmm_sparse", False, 5, CCS), ("hmm_wide", False, 6, dict(CCS, depth=2.0)),
    ("empty", False, 8, dict(depth=3.0, len_mean=9000.0, len_sd=7000.0, len_min=100, len_max=1000000, accuracy_mean_cli=0.85, ratio=(6, 54, 40))),
    ("hmm", False, 3, dict(CCS, len_mean=7000.0, len_sd=0.0, accuracy_mean_cli=0.9, depth=4.0)),
])
def test_replay_is_byte_identical_to_the_reference(sp, oracle_lib, tmp_path, kind, no_error, seed, params):
    L = 600_000
    hap = _haplotype(L)
    model = _model(tmp_path, kind)
    want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=seed, no_error=no_error, **params)
    assert log["n_reads"] > 20 and len(want) > 500_000
    exe = os.path.join(REF_BIN, "pbsim2noError" if no_error else "pbsim2")
    if os.path.exists(exe):  # the reference itself, compiled from its own sources (oracle/Makefile)
        fa = str(tmp_path / "hap.fa")
        with open(fa, "wb") as f:
            f.write(b">chr1_0\n" + hap.tobytes() + b"\n")
        subprocess.run([exe] + _cli(params, seed, model, str(tmp_path / "ref")) + [fa], check=True, stderr=subprocess.DEVNULL)
        assert open(tmp_path / "ref_0001.fastq", "rb").read() == want.tobytes()
    draws = oh.glibc_rand(oracle_lib, seed, log["total_draws"])
    with sp.Context(0, 42) as ctx:
        _configure(sp, ctx, params, seed, model, no_error, sp.READS_REPLAY)
        ctx.reads_set_replay(draws, log["draw_offset"])
        res = ctx.reads_simulate_sequence(hap)
        assert res["n_reads"] == log["n_reads"]
        assert res["n_dropped"] == int(log["dropped"].sum())
        got = ctx.reads_copy(res["file_bytes"])
        ln, dr, used = ctx.reads_log(res["n_reads"])
    assert np.array_equal(ln, log["length"]) and np.array_equal(dr, log["dropped"])
    assert np.array_equal(used.astype(np.int64), log["draws_used"])
    assert got.size == want.size and np.array_equal(got, want)


@pytest.mark.parametrize("kind,no_error,seed", [("empty", False, 11), ("hmm", False, 12), ("hmm", True, 13), ("hmm_sparse", False, 14)])
def test_per_read_streams_match_the_restatement(sp, oracle_lib, tmp_path, kind, no_error, seed):
    hap = _haplotype(500_000, seed=9)
    model = _model(tmp_path, kind)
    want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=seed, no_error=no_error, per_read_streams=True, ref_num=3, **CCS)
    with sp.Context(0, 42) as ctx:
        _configure(sp, ctx, CCS, seed, model, no_error, sp.READS_PER_READ_STREAMS)
        res = ctx.reads_simulate_sequence(hap, ref_num=3)
        got = ctx.reads_copy(res["file_bytes"])
        ln, dr, used = ctx.reads_log(res["n_reads"])
    assert res["n_reads"] == log["n_reads"] and res["bases"] == int(log["length"].sum())
    assert np.array_equal(used.astype(np.int64), log["draws_used"])
    assert np.array_equal(dr, log["dropped"]) and dr.sum() > 0  # reads over the N runs are dropped (:1711)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("design,min_ctas,pf", [("1", "10", "0"), ("1", "10", "10"), ("1", "10", "26"), ("1", "10", "58"), ("2", "10", "0"), ("2", "9", "0"), ("2", "8", "0")])
def test_every_kernel_variant_writes_the_same_bytes(sp, oracle_lib, tmp_path, monkeypatch, design, min_ctas, pf):
    """Every kernel csrc/reads.cu pick_read_kernel can return — the first design of the per-read code with its variant bits
    (58 is the default every other test runs) and the second design at 10, 9 and 8 CTAs per SM — against the restatement,
    per-read streams and replay."""
    monkeypatch.setenv("SYNTHPOP_READS_KERNEL", design)
    monkeypatch.setenv("SYNTHPOP_READS_MIN_CTAS", min_ctas)
    monkeypatch.setenv("SYNTHPOP_READS_PF", pf)
    hap = _haplotype(400_000, seed=4)
    model = _model(tmp_path, "hmm")
    want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=31, per_read_streams=True, ref_num=2, **CCS)
    want_r, log_r = oh.pbsim_simulate(oracle_lib, hap, model, seed=32, **CCS)
    draws = oh.glibc_rand(oracle_lib, 32, log_r["total_draws"])
    with sp.Context(0, 42) as ctx:
        _configure(sp, ctx, CCS, 31, model, False, sp.READS_PER_READ_STREAMS)
        res = ctx.reads_simulate_sequence(hap, ref_num=2)
        got = ctx.reads_copy(res["file_bytes"])
        _, dr, used = ctx.reads_log(res["n_reads"])
        assert np.array_equal(used.astype(np.int64), log["draws_used"]) and np.array_equal(dr, log["dropped"])
        assert np.array_equal(got, want)
        _configure(sp, ctx, CCS, 32, model, False, sp.READS_REPLAY)
        ctx.reads_set_replay(draws, log_r["draw_offset"])
        res = ctx.reads_simulate_sequence(hap)
        got = ctx.reads_copy(res["file_bytes"])
        assert np.array_equal(got, want_r)


@pytest.mark.parametrize("design", ["1", "2"])
@pytest.mark.parametrize("no_error", [False, True])
def test_short_and_ragged_reads(sp, oracle_lib, tmp_path, monkeypatch, design, no_error):
    """Lengths from 8 to 30 000 in one run (reads that end inside their first output word, reads shorter than a template
    chunk or a scratch row, warps whose lanes finish thousands of bases apart), on a sequence short enough that windows
    touch both of its ends: the requests made ahead of need (next template chunk, scratch rows) stay inside what they may read."""
    monkeypatch.setenv("SYNTHPOP_READS_KERNEL", design)
    p = dict(CCS, len_min=8, len_max=30000, len_mean=2500.0, len_sd=3000.0, depth=6.0)
    hap = _haplotype(120_000, seed=2)
    model = _model(tmp_path, "hmm")
    want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=77, no_error=no_error, per_read_streams=True, ref_num=1, **p)
    assert log["length"].min() < 64 and log["length"].max() > 8000
    with sp.Context(0, 42) as ctx:
        _configure(sp, ctx, p, 77, model, no_error, sp.READS_PER_READ_STREAMS)
        res = ctx.reads_simulate_sequence(hap)
        got = ctx.reads_copy(res["file_bytes"])
        ln, dr, used = ctx.reads_log(res["n_reads"])
    assert np.array_equal(ln, log["length"]) and np.array_equal(dr, log["dropped"])
    assert np.array_equal(used.astype(np.int64), log["draws_used"])
    assert np.array_equal(got, want)


def test_reads_from_a_materialised_haplotype_never_leave_the_device(sp, oracle_lib, tmp_path):
    """PrintHaplotypes -> pbsim without the .fa file: plan + materialise + read kernel on the resident haplotype."""
    rng = np.random.default_rng(77)
    case = oh.random_edit_lists(rng, 400_000, 200, 2, max_len=3000)
    model = _model(tmp_path, "hmm")
    p = dict(CCS, depth=3.0)
    with sp.Context(0, 42) as ctx:
        ctx.set_reference(case["ref"])
        ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
        ctx.set_haplotype_lists(case["list_off"], case["ids"])
        lens, stride = ctx.plan_haplotypes(0, 1, False)
        ctx.materialise()
        _configure(sp, ctx, p, 21, model, False, sp.READS_PER_READ_STREAMS)
        for h in range(2):
            hap = oh.build_haplotype(oracle_lib, case["ref"], case["sv_type"], case["position"], case["length"],
                                     case["payload_off"], case["payload"], None, case["lists"][h])
            assert ctx.copy_haplotype(h, stride).tobytes() == hap.tobytes(), ("haplotype", h, int(lens[h]), hap.size)
            want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=21, per_read_streams=True, ref_num=h + 1, **p)
            res = ctx.reads_simulate_haplotype(h, ref_num=h + 1)
            got = ctx.reads_copy(res["file_bytes"])
            if res["n_reads"] != log["n_reads"] or got.size != want.size:
                # Seen in about one of ten full-suite runs, never alone: the FIRST in-process oracle call of this test returned
                # a different read count than the same call repeated (and than the oracle run standalone, which agrees with the
                # device). The oracle is recomputed twice; the test goes on only if both repeats agree with each other, and says so.
                import warnings
                first = (log["n_reads"], want.size)
                want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=21, per_read_streams=True, ref_num=h + 1, **p)
                want3, log3 = oh.pbsim_simulate(oracle_lib, hap, model, seed=21, per_read_streams=True, ref_num=h + 1, **p)
                details = (h, "device", res, "oracle first call", first, "oracle repeated", (log["n_reads"], want.size), (log3["n_reads"], want3.size),
                           "haplotype bytes", hap.size, "model bytes", os.path.getsize(model), p)
                assert want.tobytes() == want3.tobytes() and (log["n_reads"], want.size) != first, details
                warnings.warn(f"oracle pbsim call was not repeatable in this process: {details}")
            assert res["n_reads"] == log["n_reads"], (h, res, log["n_reads"])
            assert got.size == want.size, (h, got.size, want.size, res)
            bad = np.nonzero(got != want)[0]
            assert bad.size == 0, (h, int(bad[0]), int(bad.size), got[max(0, bad[0] - 30):bad[0] + 30].tobytes(), want[max(0, bad[0] - 30):bad[0] + 30].tobytes())


def test_pbsim_command_writes_the_reference_file_layout(sp, oracle_lib, tmp_path):
    """sp_simulate_reads = the `pbsim` command of docker/haplotype2reads.sh:53 on a two-sequence FASTA."""
    a, b = _haplotype(300_000, seed=1), _haplotype(200_000, seed=2)
    fa = str(tmp_path / "two.fa")
    with open(fa, "wb") as f:
        f.write(b">first\n" + a.tobytes() + b"\n>second\n")
        for i in range(0, b.size, 70):  # line-wrapped record
            f.write(b[i:i + 70].tobytes() + b"\n")
    model = _model(tmp_path, "hmm")
    p = dict(CCS, depth=4.0)
    with sp.Context(0, 42) as ctx:
        args, _ = sp.ReadArgs.from_options(_cli(p, 5, model, str(tmp_path / "reads_7")))
        ctx.simulate_reads(args, fa)
    for k, seq in enumerate((a, b)):
        want, _ = oh.pbsim_simulate(oracle_lib, seq, model, seed=5, per_read_streams=True, ref_num=k + 1, **p)
        assert open(tmp_path / f"reads_7_{k + 1:04d}.fastq", "rb").read() == want.tobytes()


def test_pbsim_argument_errors(sp, tmp_path):
    with sp.Context(0, 42) as ctx:
        args, _ = sp.ReadArgs.from_options(["--depth", "5"])
        with pytest.raises(sp.SynthpopError, match="hmm_model"):
            ctx.reads_configure(args)
        mf = str(tmp_path / "m"); open(mf, "w").close()
        args, _ = sp.ReadArgs.from_options(["--length-min", "9000", "--length-max", "100", "--hmm_model", mf])
        with pytest.raises(sp.SynthpopError, match="greater than max"):
            ctx.reads_configure(args)
        args, _ = sp.ReadArgs.from_options(["--hmm_model", str(tmp_path / "missing")])
        with pytest.raises(sp.SynthpopError, match="Cannot open file"):
            ctx.reads_configure(args)
        args, _ = sp.ReadArgs.from_options(["--hmm_model", mf, "--depth", "2"])
        ctx.reads_configure(args)
        with pytest.raises(sp.SynthpopError, match="too short"):
            ctx.reads_simulate_sequence(np.frombuffer(b"ACGT" * 10, np.uint8))


@pytest.mark.parametrize("design", ["1", "2"])
def test_reads_the_reference_cannot_simulate_are_refused(sp, tmp_path, monkeypatch, design):
    """A read of length 1 makes the reference compute rand() % 0 (docker/pbsim2.cpp:2284); a read at least as long as its
    sequence makes it emit fewer bases than it accounted for. Both kernels stop with an error that names the read; the context
    stays usable."""
    monkeypatch.setenv("SYNTHPOP_READS_KERNEL", design)
    mf = str(tmp_path / "m"); open(mf, "w").close()
    hap = _haplotype(20_000, seed=1)
    with sp.Context(0, 42) as ctx:
        args, _ = sp.ReadArgs.from_options(["--hmm_model", mf, "--depth", "1", "--length-min", "1", "--length-max", "40", "--length-mean", "2",
                                            "--length-sd", "2", "--seed", "3"])
        ctx.reads_configure(args)
        with pytest.raises(sp.SynthpopError, match="length 1"):
            ctx.reads_simulate_sequence(hap)
        args, _ = sp.ReadArgs.from_options(["--hmm_model", mf, "--depth", "3", "--length-min", "15000", "--length-max", "90000", "--length-mean",
                                            "30000", "--length-sd", "9000", "--seed", "3"])
        ctx.reads_configure(args)
        with pytest.raises(sp.SynthpopError, match="as long as the sequence"):
            ctx.reads_simulate_sequence(hap)
        args, _ = sp.ReadArgs.from_options(["--hmm_model", mf, "--depth", "2", "--length-min", "100", "--length-max", "4000", "--length-mean",
                                            "1500", "--length-sd", "500", "--seed", "3"])
        ctx.reads_configure(args)
        res = ctx.reads_simulate_sequence(hap)
        assert res["n_reads"] > 10 and res["bases"] >= 2 * hap.size


def test_full_size_30x_chr1_properties(sp, tmp_path):
    """BASELINE configs[3] at full size (30x CCS over one chr1-sized haplotype, ~7.5 Gbp of reads): too large for the
    CPU restatement, so the size-independent properties of simulate_by_model are checked instead — the quota rule
    (docker/pbsim2.cpp:1641-1650), the record layout (:1710-1711), the NNNNN filter (:1709), determinism, and independence of the
    per-read streams from what else ran on the context."""
    from synthpop import synth
    L = 248_956_422
    hap, _, _ = synth.random_reference(L, 11)
    model = str(tmp_path / "hmm.model"); synth.write_hmm_model(model, quality="phred")
    p = dict(CCS, depth=30.0)
    with sp.Context(0, 42) as ctx:
        _configure(sp, ctx, p, 42, model, False, sp.READS_PER_READ_STREAMS)
        res = ctx.reads_simulate_sequence(hap)
        quota = int(30.0 * L)
        assert quota <= res["bases"] < quota + p["len_min"]            # the last read is clipped to the quota or to len_min
        ln, dr, used = ctx.reads_log(res["n_reads"])
        assert int(ln.sum()) == res["bases"] and int(dr.sum()) == res["n_dropped"] > 0
        assert ln.min() >= 2 and ln[:-1].min() >= p["len_min"] and ln.max() <= p["len_max"]
        assert int(used.astype(np.int64).sum()) >= 3 * res["bases"]    # 2 quality draws + 1 error draw per base at least
        buf = ctx.reads_copy(res["file_bytes"])
        kept = res["n_reads"] - res["n_dropped"]
        assert int(np.count_nonzero(buf == ord("\n"))) == kept and int(np.count_nonzero(buf == ord("Z"))) == kept
        assert int(np.count_nonzero(buf == ord(">"))) == kept and buf[0] == ord(">") and buf[-1] == ord("\n")
        hist = np.bincount(buf, minlength=256)
        allowed = np.zeros(256, bool); allowed[list(b"ACGTN>SZ_\n0123456789")] = True
        assert hist[~allowed].sum() == 0
        # sequence bytes = file bytes minus the names: the bases of the kept reads
        seq_bytes = int(hist[list(b"ACGTN")].sum())
        assert seq_bytes == int(ln[dr == 0].sum())
        # no kept record contains NNNNN (bytes.find scans at memory speed)
        assert buf.tobytes().find(b"NNNNN") == -1
        checksum = int(np.add.reduce(buf[::7].astype(np.uint64) * np.arange(buf[::7].size, dtype=np.uint64) % np.uint64(1000003)))
        # determinism, and independence from the context's history: another sequence in between, then the same again
        ctx.reads_simulate_sequence(hap[: 3_000_000])
        res2 = ctx.reads_simulate_sequence(hap)
        assert res2 == res
        buf2 = ctx.reads_copy(res2["file_bytes"])
        assert int(np.add.reduce(buf2[::7].astype(np.uint64) * np.arange(buf2[::7].size, dtype=np.uint64) % np.uint64(1000003))) == checksum
        assert np.array_equal(buf2[:50_000_000], buf[:50_000_000]) and np.array_equal(buf2[-50_000_000:], buf[-50_000_000:])
