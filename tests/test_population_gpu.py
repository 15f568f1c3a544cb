"""Rows S1/G/C/N/I/R1/W/Z parity: sp_simulate_population (CUDA genotype kernel + collision fixer + host prologue and
writers, through the C ABI) against the CPU oracle's restatement of SimulateHaplotypes.main on the same seed and
inputs. Every output file must be byte-identical."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FILES = ["groundTruth_joint.vcf", "haplotype2variants.txt", "variants.txt", "histogram_nVariantsPerHaplotype.txt",
         "histogram_startPositions.txt"]


def _cmp_dirs(want, got, n_ind):
    names = FILES + [f"groundTruth_individual_{2 * k}.vcf" for k in range(n_ind)]
    for n in names:
        a = open(os.path.join(want, n), "rb").read()
        b = open(os.path.join(got, n), "rb").read()
        if a != b:
            k = next(i for i in range(min(len(a), len(b))) if a[i] != b[i]) if a[:min(len(a), len(b))] != b[:min(len(a), len(b))] else min(len(a), len(b))
            raise AssertionError(f"{n} differs at byte {k}: want {a[max(0,k-60):k+60]!r} got {b[max(0,k-60):k+60]!r}")


def _parse_lists(path):
    rows = open(path).read().split("\n")
    H = int(rows[0])
    out = []
    for r in rows[1:1 + H]:
        t = r.split(",")
        out.append(np.array([int(x) for x in t[1:1 + int(t[0]) + 1]], dtype=np.int32))
    return out


@pytest.mark.parametrize("L,n_ind,n_freq,seed", [(1_500_000, 5, 250, 42), (6_000_000, 16, 900, 7), (800_000, 3, 120, 123456789)])
def test_simulate_haplotypes_outputs_byte_identical(sp, oracle_bin, tmp_path, L, n_ind, n_freq, seed):
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), L, n_insertions=2500, seed=seed)
    want = str(tmp_path / "want"); got = str(tmp_path / "got")
    os.makedirs(want); os.makedirs(got)
    pre = str(tmp_path / "pre.txt"); log = str(tmp_path / "draws.txt")
    subprocess.run([oracle_bin, "--quiet", "--seed", str(seed), "--dump-pre-collision", pre, "--draw-log", log,
                    "SimulateHaplotypes"] + inp.simulate_argv(want, n_ind, n_frequent=n_freq), check=True)
    os.environ["SYNTHPOP_KEEP_PRE_COLLISION"] = "1"
    ctx = sp.Context(0, seed)
    ctx.simulate_haplotypes(inp.simulate_argv(got, n_ind, n_frequent=n_freq))
    # row S1: same number of candidates and LCG steps as the oracle's draw log
    dl = dict(l.split("=") for l in open(log).read().split())
    mine = ctx.population_draw_log()
    assert mine["steps_after_sample_variants"] == int(dl["steps_after_sampleVariants"])
    assert mine["candidates"] == int(dl["candidates"])
    # row G: the raw genotype matrix (lists before collision fixing)
    off, ids = ctx.population_pre_collision_lists(2 * n_ind)
    for h, w in enumerate(_parse_lists(pre)):
        assert np.array_equal(ids[off[h]:off[h + 1]], w), f"genotype row of haplotype {h} differs"
    # rows C/N/I/R1/W/Z: every output file
    _cmp_dirs(want, got, n_ind)
    # the context is ready for materialisation: haplotypes equal the oracle's PrintHaplotypes
    hw = str(tmp_path / "hw"); os.makedirs(hw)
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", "3", inp.reference_file, want + "/haplotype2variants.txt",
                    want + "/variants.txt", hw], check=True)
    lens, stride = ctx.plan_haplotypes(0, 3, True)
    ctx.materialise()
    for h in range(4):
        assert ctx.copy_haplotype(h, stride).tobytes() == open(f"{hw}/haplotype_{h}.fa", "rb").read()
    ctx.close()


def test_sharded_population_matches_single(sp, oracle_bin, tmp_path):
    """Two 'ranks' on one GPU, run one after the other: shard [0..5] and [6..11], exchange counts on the host."""
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 2_000_000, n_insertions=2000)
    want = str(tmp_path / "want"); got = str(tmp_path / "got")
    os.makedirs(want); os.makedirs(got)
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(want, 6, n_frequent=300), check=True)
    argv = inp.simulate_argv(got, 6, n_frequent=300)
    a = sp.Context(0, 42); b = sp.Context(0, 42)
    a.population_begin(argv, 0, 5); b.population_begin(argv, 6, 11)
    counts = a.population_local_counts() + b.population_local_counts()
    oa, ia = a.population_local_lists(6); ob, ib = b.population_local_lists(6)
    all_off = np.concatenate([oa, ob[1:] + oa[-1]]); all_ids = np.concatenate([ia, ib])
    a.population_finish(counts, all_off, all_ids, True)
    b.population_finish(counts, None, None, False)
    _cmp_dirs(want, got, 6)
    a.close(); b.close()


def test_baseline_config1_chr21_ten_individuals(sp, oracle_bin, tmp_path):
    """BASELINE.json configs[0]: chr21-sized reference (46 709 983 bp), 10 diploid individuals, 5 000 frequent SVs,
    30 000 insertion strings — every SimulateHaplotypes output and every haplotype_<i>.fa byte-identical."""
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 46_709_983, n_insertions=30000)
    want = str(tmp_path / "want"); got = str(tmp_path / "got")
    os.makedirs(want); os.makedirs(got)
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(want, 10, n_frequent=5000), check=True)
    ctx = sp.Context(0, 42)
    ctx.simulate_haplotypes(inp.simulate_argv(got, 10, n_frequent=5000))
    _cmp_dirs(want, got, 10)
    hw = str(tmp_path / "hw"); hg = str(tmp_path / "hg"); os.makedirs(hw); os.makedirs(hg)
    files = [inp.reference_file, want + "/haplotype2variants.txt", want + "/variants.txt"]
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", "19"] + files + [hw], check=True)
    ctx.print_haplotypes(0, 19, *files, hg)
    for h in range(20):
        assert open(f"{hw}/haplotype_{h}.fa", "rb").read() == open(f"{hg}/haplotype_{h}.fa", "rb").read(), h
    ctx.close()


def test_global_memory_collision_fixer_fallback(sp, oracle_bin, tmp_path, monkeypatch):
    """The fallback used when a haplotype's list does not fit in shared memory gives the same files."""
    from synthpop import synth
    monkeypatch.setenv("SYNTHPOP_FORCE_GLOBAL_COLLISION_FIX", "1")
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 3_000_000, n_insertions=2000, seed=99)
    want = str(tmp_path / "want"); got = str(tmp_path / "got")
    os.makedirs(want); os.makedirs(got)
    subprocess.run([oracle_bin, "--quiet", "--seed", "5", "SimulateHaplotypes"] + inp.simulate_argv(want, 8, n_frequent=600), check=True)
    ctx = sp.Context(0, 5)
    ctx.simulate_haplotypes(inp.simulate_argv(got, 8, n_frequent=600))
    _cmp_dirs(want, got, 8)
    ctx.close()


def test_pipelined_batches_on_population_stream(sp, oracle_bin, tmp_path):
    """sp_set_population_stream: rows G/C/N of batch s+1 run on a second stream while batch s is being materialised;
    every haplotype of every batch still equals the oracle's PrintHaplotypes output."""
    import torch
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 3_000_000, n_insertions=2000, seed=11)
    want = str(tmp_path / "want"); hw = str(tmp_path / "hw")
    os.makedirs(want); os.makedirs(hw)
    n_ind = 12
    argv = inp.simulate_argv(want, n_ind, n_frequent=500)
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + argv, check=True)
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", str(2 * n_ind - 1), inp.reference_file,
                    want + "/haplotype2variants.txt", want + "/variants.txt", hw], check=True)
    ctx = sp.Context(0, 42)
    main, side = torch.cuda.Stream(), torch.cuda.Stream(priority=-1)
    ctx.set_stream(main.cuda_stream)
    ctx.population_begin(inp.simulate_argv(str(tmp_path / "got"), n_ind, n_frequent=500), 0, -1)
    ctx.population_finish(None, None, None, 2)
    ctx.set_population_stream(side.cuda_stream)
    batch = 6
    pending = None
    for a in range(0, 2 * n_ind, batch):
        ctx.population_genotypes(a, a + batch - 1)     # overlaps the previous batch's materialisation kernel
        if pending is not None:                        # the previous batch is read back only now
            pa, pstride = pending
            for h in range(pa, pa + batch):
                assert ctx.copy_haplotype(h, pstride).tobytes() == open(f"{hw}/haplotype_{h}.fa", "rb").read(), h
        lens, stride = ctx.plan_haplotypes(a, a + batch - 1, True)
        ctx.materialise()
        pending = (a, stride)
    pa, pstride = pending
    for h in range(pa, pa + batch):
        assert ctx.copy_haplotype(h, pstride).tobytes() == open(f"{hw}/haplotype_{h}.fa", "rb").read(), h
    ctx.set_population_stream(None)
    ctx.close()
