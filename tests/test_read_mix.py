"""`cat | shuf | tr 'Z' '\\n'` (docker/haplotype2reads.sh:59-63) and the 1x chunks of
docker/reads2svs.sh:65-74. CPU part: the restatement (oracle/read_mix.py) against the coreutils binaries the reference
script runs. GPU part: sp_reads_mix_* against the restatement, byte for byte."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import read_mix as rm  # noqa: E402
import oracle_helpers as oh  # noqa: E402


def _fake_pbsim_file(rng, n, tag=b"S1_"):
    recs = []
    for i in range(n):
        seq = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=int(rng.integers(5, 90))))
        recs.append(b">" + tag + str(i + 1).encode() + b"Z" + seq + b"\n")
    return b"".join(recs)


def test_java_lcg_keys_known_answers():
    # java.util.Random(42).nextInt() = -1170105035, 234785527 are the top 32 bits of states 1 and 2
    k = rm.shuffle_keys(42, 3)
    assert [int(np.int32(np.uint32(int(x) >> 16))) for x in k[:2]] == [-1170105035, 234785527]
    assert len(set(int(x) for x in rm.shuffle_keys(7, 5000))) == 5000


@pytest.mark.parametrize("n1,n2,max_cov", [(37, 41, 5), (200, 180, 30), (8, 8, 4), (3, 0, 2)])
def test_restatement_against_coreutils(tmp_path, n1, n2, max_cov):
    rng = np.random.default_rng(n1 * 1000 + n2)
    f1, f2 = _fake_pbsim_file(rng, n1), _fake_pbsim_file(rng, n2)
    mixed = rm.mix_reads([f1, f2], seed=99)
    # cat + tr: same multiset of two-line records
    (tmp_path / "a.fa").write_bytes(f1); (tmp_path / "b.fa").write_bytes(f2)
    cat_tr = subprocess.run(f"cat {tmp_path}/a.fa {tmp_path}/b.fa | tr 'Z' '\\n'", shell=True, capture_output=True, check=True).stdout
    recs = lambda b: sorted(b.split(b">")[1:])
    assert recs(mixed) == recs(cat_tr) and len(mixed) == len(cat_tr)
    if n1 + n2 > 4:
        assert mixed != cat_tr  # it is not the identity permutation
    # split -d -l N_ROWS_1X: chunk-00.. and the coverage prefixes
    (tmp_path / "reads.fa").write_bytes(mixed)
    n_rows = int(subprocess.run(f"wc -l < {tmp_path}/reads.fa", shell=True, capture_output=True, check=True).stdout)
    r1 = n_rows // max_cov
    if r1 % 4 != 0:
        r1 = (r1 // 4 + 1) * 4   # docker/reads2svs.sh:68-71
    assert r1 == rm.rows_1x(n_rows, max_cov)
    subprocess.run(["split", "-d", "-l", str(r1), str(tmp_path / "reads.fa"), str(tmp_path / "chunk-")], check=True)
    acc = b""
    for c in range(1, max_cov + 1):
        p = tmp_path / f"chunk-{c - 1:02d}"
        if p.exists():
            acc += p.read_bytes()
        assert rm.coverage_prefix(mixed, max_cov, c) == acc, c


CCS = dict(depth=6.0, len_mean=15000.0, len_sd=2000.0, len_min=5000, len_max=30000, accuracy_mean_cli=0.995, ratio=(6, 50, 54))


@pytest.mark.gpu
@pytest.mark.parametrize("seed,shuffle_seed,max_cov", [(21, 5, 6), (4, 123456789, 3)])
def test_gpu_mix_matches_the_restatement(sp, oracle_lib, tmp_path, seed, shuffle_seed, max_cov):
    from synthpop import synth
    import test_reads_gpu as trg
    model = str(tmp_path / "hmm.model"); synth.write_hmm_model(model)
    haps = [trg._haplotype(420_000, seed=1), trg._haplotype(390_000, seed=2)]
    files = []
    with sp.Context(0, 42) as ctx:
        trg._configure(sp, ctx, CCS, seed, model, False, sp.READS_PER_READ_STREAMS)
        ctx.reads_mix_begin()
        dropped = 0
        for hap in haps:
            want, log = oh.pbsim_simulate(oracle_lib, hap, model, seed=seed, per_read_streams=True, ref_num=1, **CCS)
            files.append(want.tobytes())
            res = ctx.reads_simulate_sequence(hap, ref_num=1)
            dropped += res["n_dropped"]
            ctx.reads_mix_add()
        assert dropped > 0   # reads over the N runs never reach the file (docker/pbsim2.cpp:1711)
        n, nbytes = ctx.reads_mix_finish(shuffle_seed)
        want = rm.mix_reads(files, shuffle_seed)
        assert nbytes == len(want) and n == want.count(b">")
        got = ctx.reads_mix_copy(0, nbytes).tobytes()
        assert got == want
        off = ctx.reads_mix_offsets(n)
        assert off[0] == 0 and off[-1] == nbytes and all(got[o:o + 1] == b">" for o in off[:-1])
        for c in range(0, max_cov + 2):
            pre = ctx.reads_mix_coverage_prefix(max_cov, c)
            w = rm.coverage_prefix(want, max_cov, c)
            assert pre["bytes"] == len(w) and pre["rows"] == w.count(b"\n"), c
            assert pre["rows_1x"] == rm.rows_1x(2 * n, max_cov)
        # a second individual on the same context reuses the buffers
        ctx.reads_mix_begin()
        ctx.reads_simulate_sequence(haps[1], ref_num=1)
        ctx.reads_mix_add()
        n2, b2 = ctx.reads_mix_finish(shuffle_seed)
        assert ctx.reads_mix_copy(0, b2).tobytes() == rm.mix_reads(files[1:], shuffle_seed)


@pytest.mark.gpu
def test_haplotype2reads_whole_script(sp, oracle_lib, tmp_path):
    """sp_haplotype2reads = docker/haplotype2reads.sh:53-64 for one individual, haplotypes resident on the device."""
    from synthpop import synth
    import test_reads_gpu as trg
    rng = np.random.default_rng(5)
    case = oh.random_edit_lists(rng, 400_000, 200, 2, max_len=3000)
    model = str(tmp_path / "hmm.model"); synth.write_hmm_model(model)
    p = dict(CCS, depth=3.0)
    out = str(tmp_path / "reads_i0_i1.fa")
    with sp.Context(0, 42) as ctx:
        ctx.set_reference(case["ref"])
        ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
        ctx.set_haplotype_lists(case["list_off"], case["ids"])
        trg._configure(sp, ctx, p, 21, model, False, sp.READS_PER_READ_STREAMS)
        n, nbytes = ctx.haplotype2reads(0, 77, out)
    files = []
    for h in range(2):
        hap = oh.build_haplotype(oracle_lib, case["ref"], case["sv_type"], case["position"], case["length"],
                                 case["payload_off"], case["payload"], None, case["lists"][h])
        # one pbsim process per haplotype in the reference, each with its own seed (docker/haplotype2reads.sh:53-58)
        want, _ = oh.pbsim_simulate(oracle_lib, hap, model, seed=sp.haplotype_read_seed(21, h), per_read_streams=True, ref_num=1, **p)
        files.append(want.tobytes())
    assert open(out, "rb").read() == rm.mix_reads(files, 77)
    assert files[0][:4000] != files[1][:4000] and sp.haplotype_read_seed(21, 0) != sp.haplotype_read_seed(21, 1)


@pytest.mark.gpu
def test_pair_in_flight_equals_one_by_one(sp, tmp_path):
    """sp_reads_mix_add_pair (both read kernels of an individual in flight, second state and stream) against the one-by-one
    calls, which the tests above pin to the oracle: same statistics per haplotype, same mixed file, twice in a row."""
    from synthpop import synth
    import test_reads_gpu as trg
    rng = np.random.default_rng(9)
    case = oh.random_edit_lists(rng, 500_000, 250, 4, max_len=3000)
    model = str(tmp_path / "hmm.model"); synth.write_hmm_model(model)
    p = dict(CCS, depth=4.0)
    with sp.Context(0, 42) as ctx:
        ctx.set_reference(case["ref"])
        ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
        ctx.set_haplotype_lists(case["list_off"], case["ids"])
        trg._configure(sp, ctx, p, 33, model, False, sp.READS_PER_READ_STREAMS)
        ctx.plan_haplotypes(0, 3, False)
        ctx.materialise()
        for h0 in (2, 0, 2):
            ctx.reads_mix_begin()
            one = []
            for h in (h0, h0 + 1):
                one.append(ctx.reads_simulate_haplotype(h, ref_num=1))
                ctx.reads_mix_add()
            n, nbytes = ctx.reads_mix_finish(5 + h0)
            want = ctx.reads_mix_copy(0, nbytes).tobytes()
            ctx.reads_mix_begin()
            l0 = ctx.kernel_launches
            res, span = ctx.reads_mix_add_pair(h0, h0 + 1, 1, 1)
            assert span > 0 and ctx.kernel_launches > l0
            for a, b in zip(one, res):
                assert {k: a[k] for k in b} == b
            n2, b2 = ctx.reads_mix_finish(5 + h0)
            assert (n2, b2) == (n, nbytes)
            assert ctx.reads_mix_copy(0, b2).tobytes() == want
        # per-haplotype seeds change the reads (and sp_haplotype2reads, which is pinned to the oracle above, runs this path)
        ctx.reads_mix_begin()
        res2, _ = ctx.reads_mix_add_pair(2, 3, 1, 1, seed_per_haplotype=True)
        n3, b3 = ctx.reads_mix_finish(7)
        assert ctx.reads_mix_copy(0, b3).tobytes() != want
        # a reconfiguration reaches the second state too
        trg._configure(sp, ctx, dict(p, depth=2.0), 34, model, False, sp.READS_PER_READ_STREAMS)
        ctx.reads_mix_begin()
        one = []
        for h in (0, 1):
            one.append(ctx.reads_simulate_haplotype(h, ref_num=1))
            ctx.reads_mix_add()
        n, nbytes = ctx.reads_mix_finish(1)
        want = ctx.reads_mix_copy(0, nbytes).tobytes()
        ctx.reads_mix_begin()
        res, _ = ctx.reads_mix_add_pair(0, 1, 1, 1)
        assert [r["n_reads"] for r in res] == [r["n_reads"] for r in one]
        n2, b2 = ctx.reads_mix_finish(1)
        assert ctx.reads_mix_copy(0, b2).tobytes() == want
