"""Row M parity: the CUDA materialisation path (through the C ABI) against the CPU oracle's restatement of
SimulateHaplotypes.buildHaplotypes (reference src/SimulateHaplotypes.java:917-960). Bit-exact."""
import os

import numpy as np
import pytest

import oracle_helpers as oh

pytestmark = pytest.mark.gpu


def _run_case(sp, oracle_lib, case, framing, batch=None):
    ctx = sp.Context(0, 42)
    ctx.set_reference(case["ref"])
    ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
    ctx.set_haplotype_lists(case["list_off"], case["ids"])
    H = len(case["lists"])
    batch = batch or H
    for a in range(0, H, batch):
        b = min(H, a + batch) - 1
        lens, stride = ctx.plan_haplotypes(a, b, framing)
        ctx.materialise()
        ctx.synchronize()
        for h in range(a, b + 1):
            want = oh.build_haplotype(oracle_lib, case["ref"], case["sv_type"], case["position"], case["length"],
                                      case["payload_off"], case["payload"], None, case["lists"][h])
            if framing:
                want = np.concatenate([np.frombuffer(f">chr1_{h}\n".encode(), np.uint8), want, np.frombuffer(b"\n", np.uint8)])
            got = ctx.copy_haplotype(h, int(stride))
            assert got.size == want.size == lens[h - a]
            if not np.array_equal(got, want):
                bad = int(np.nonzero(got != want)[0][0])
                raise AssertionError(f"haplotype {h}: first mismatch at byte {bad}: got {got[bad]} want {want[bad]}")
            ptr, n = ctx.haplotype_device(h)
            assert n == want.size - (len(f">chr1_{h}\n") + 1 if framing else 0)
    ctx.close()


@pytest.mark.parametrize("framing", [False, True])
@pytest.mark.parametrize("L,nv,nh,maxlen", [(5000, 40, 3, 200), (200_000, 600, 8, 3000), (3_000_000, 400, 5, 60000),
                                            (70_000, 6000, 4, 12)])
def test_random_edit_lists(sp, oracle_lib, L, nv, nh, maxlen, framing):
    rng = np.random.default_rng(L + nv)
    case = oh.random_edit_lists(rng, L, nv, nh, max_len=maxlen)
    _run_case(sp, oracle_lib, case, framing)
    _run_case(sp, oracle_lib, case, framing, batch=2)


def test_empty_and_ragged(sp, oracle_lib):
    rng = np.random.default_rng(5)
    case = oh.random_edit_lists(rng, 40_000, 50, 4)
    case["lists"][0] = np.zeros(0, np.int32)                       # haplotype with no variant: plain copy
    case["lists"][2] = np.arange(case["sv_type"].size, dtype=np.int32)  # carries everything
    case["list_off"] = np.concatenate(([0], np.cumsum([l.size for l in case["lists"]]))).astype(np.int64)
    case["ids"] = np.concatenate(case["lists"]).astype(np.int32)
    _run_case(sp, oracle_lib, case, False)
    _run_case(sp, oracle_lib, case, True)


def test_each_type_alone_and_edges(sp, oracle_lib):
    L = 50_000
    rng = np.random.default_rng(9)
    ref = np.frombuffer(b"ACGTacgtNnRy", dtype=np.uint8)[rng.integers(0, 12, size=L)].copy()
    payload = np.frombuffer(b"ACGTTTGACCA" * 30, dtype=np.uint8).copy()
    #           DEL at 0   DUP      INS at 0-len gap  INV      INS at L (after the end)  DEL to the end
    sv_type = np.array([0, 1, 2, 3, 2, 0, 3], dtype=np.int8)
    position = np.array([0, 100, 1000, 20_000, L, L - 70, 17], dtype=np.int32)
    length = np.array([17, 333, 200, 16_400, 130, 70, 1], dtype=np.int32)
    payload_off = np.array([0, 0, 0, 200, 200, 330, 330, 330], dtype=np.int64)
    lists = [np.array([i], np.int32) for i in range(7)] + [np.array([0, 6, 1, 2, 3], np.int32),
                                                            np.array([2, 5, 4], np.int32), np.array([3, 5], np.int32)]
    case = dict(ref=ref, sv_type=sv_type, position=position, length=length, payload_off=payload_off, payload=payload,
                lists=lists, list_off=np.concatenate(([0], np.cumsum([l.size for l in lists]))).astype(np.int64),
                ids=np.concatenate(lists).astype(np.int32))
    _run_case(sp, oracle_lib, case, False)
    _run_case(sp, oracle_lib, case, True)


def test_overlapping_edits_raise_like_java(sp, oracle_lib):
    L = 10_000
    ref = np.full(L, ord("A"), dtype=np.uint8)
    sv_type = np.array([0, 0, 1], dtype=np.int8)
    position = np.array([100, 150, 9_990], dtype=np.int32)
    length = np.array([100, 10, 50], dtype=np.int32)
    ctx = sp.Context(0, 42)
    ctx.set_reference(ref)
    ctx.set_variants(sv_type, position, length)
    for ids, msg in ((np.array([0, 1], np.int32), "start > end"), (np.array([2], np.int32), "reference length"),
                     (np.array([7], np.int32), "variant id")):
        ctx.set_haplotype_lists(np.array([0, ids.size], np.int64), ids)
        with pytest.raises(sp.SynthpopError) as ei:
            ctx.plan_haplotypes(0, 0)
        assert ei.value.code == sp.SP_ERR_JAVA_EXCEPTION and msg in ei.value.message
        with pytest.raises(RuntimeError):  # the oracle (Java semantics) throws on the same input
            oh.build_haplotype(oracle_lib, ref, sv_type, position, length, np.zeros(4, np.int64), np.zeros(1, np.uint8), None, ids)
    ctx.close()


def test_print_haplotypes_files_match_oracle(sp, oracle_bin, tmp_path):
    """PrintHaplotypes end to end on generated SimulateHaplotypes outputs: byte-identical haplotype_<i>.fa."""
    import subprocess
    from synthpop import synth
    root = str(tmp_path / "in")
    inp = synth.generate_population_inputs(root, 1_500_000, n_insertions=1500)
    out = str(tmp_path / "pop"); os.makedirs(out)
    subprocess.run([oracle_bin, "--quiet", "--seed", "7", "SimulateHaplotypes"] + inp.simulate_argv(out, 4, n_frequent=250), check=True)
    want_dir = str(tmp_path / "want"); got_dir = str(tmp_path / "got")
    os.makedirs(want_dir); os.makedirs(got_dir)
    args = [inp.reference_file, out + "/haplotype2variants.txt", out + "/variants.txt"]
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", "7"] + args + [want_dir], check=True)
    ctx = sp.Context(0, 42)
    ctx.print_haplotypes(0, 7, *args, got_dir)
    for i in range(8):
        a = open(f"{want_dir}/haplotype_{i}.fa", "rb").read()
        b = open(f"{got_dir}/haplotype_{i}.fa", "rb").read()
        assert a == b, f"haplotype_{i}.fa differs"
    with pytest.raises(sp.SynthpopError):
        ctx.print_haplotypes(6, 9, *args, got_dir)  # haplotype 8 does not exist: ArrayIndexOutOfBounds like the Java
    ctx.close()


def test_full_size_chr1_against_oracle(sp, oracle_lib):
    """BASELINE configs[1] sizes (248 956 422 bp): bit-exact on sampled haplotypes, length identity on all of the batch."""
    import bench
    L = bench.CHR1
    ref = bench.random_reference(L)
    pop = bench.synth_population(L, 16)
    ctx = sp.Context(0, 42)
    ctx.set_reference(ref)
    ctx.set_variants(pop["sv_type"], pop["position"], pop["length"], pop["payload_off"], pop["payload"])
    ctx.set_haplotype_lists(pop["list_off"], pop["ids"])
    lens, stride = ctx.plan_haplotypes(0, 15, False)
    ctx.materialise()
    ctx.synchronize()
    t, ln = pop["sv_type"].astype(np.int64), pop["length"].astype(np.int64)
    delta = np.where(t == 0, -ln, np.where(t == 3, 0, ln))  # DEL -len, DUP +len, INS +len, INV 0
    for h in range(16):
        assert lens[h] == L + int(delta[pop["lists"][h]].sum())
    for h in (0, 7, 15):
        want = oh.build_haplotype(oracle_lib, ref, pop["sv_type"], pop["position"], pop["length"], pop["payload_off"],
                                  pop["payload"], None, pop["lists"][h])
        got = ctx.copy_haplotype(h, int(stride))
        assert np.array_equal(got, want), f"haplotype {h} differs at chr1 size"
    ctx.close()


def test_copy_design_probes_copy_the_right_bytes(sp):
    """csrc/probe.cu (the register path, the bulk-copy engine under an mbarrier, warp-private bulk pipelines): each probe
    checks the bytes it wrote against the shifted source before it reports a time."""
    with sp.Context(0, 42) as ctx:
        for kind in (0, 1, 2, 3):
            for mis in ((0,) if kind == 0 else (0, 9, 15)):
                assert ctx.probe_bandwidth(kind, 3_000_001, 7, mis, 2) > 0.0
