"""Row T of SURVEY.md §8a (BASELINE config 5): a Mendelian child and its coverage sweep. The reference has no code that
derives a child (its trios are real families), so the ROW IS PARITY-UNPINNED BY CONSTRUCTION; what is checked here is
the definition in include/synthpop_cuda.h against pieces the oracle does pin: the child's haplotypes are byte-identical
to the transmitted parental haplotypes the oracle's PrintHaplotypes writes, and the child's truth VCF is the reference's
buildVCF merge (src/SimulateHaplotypes.java:1027-1066) of records the oracle wrote for the parents."""
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


def _transmitted(vcf_path, allele):
    """records of an individual's truth VCF carried by haplotype `allele` (0/1): {variant index: line without GT}"""
    out, header = {}, []
    for line in open(vcf_path, "rb").read().split(b"\n"):
        if not line:
            continue
        if line.startswith(b"#"):
            header.append(line)
            continue
        cols = line.split(b"\t")
        gt = cols[-1].split(b"|")
        if gt[allele] == b"1":
            out[int(cols[2].split(b"_")[1])] = b"\t".join(cols[:-1])
    return out, header


def test_child_haplotypes_vcf_and_coverage_sweep(sp, oracle_bin, oracle_lib, tmp_path):
    from synthpop import synth
    import test_reads_gpu as trg
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 2_500_000, n_insertions=2000, seed=3)
    want = str(tmp_path / "want"); got = str(tmp_path / "got"); hw = str(tmp_path / "hw")
    for d in (want, got, hw):
        os.makedirs(d)
    n_ind = 6
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(want, n_ind, n_frequent=400), check=True)
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", str(2 * n_ind - 1), inp.reference_file,
                    want + "/haplotype2variants.txt", want + "/variants.txt", hw], check=True)
    with sp.Context(0, 42) as ctx:
        ctx.simulate_haplotypes(inp.simulate_argv(got, n_ind, n_frequent=400))
        father, mother = 1, 4
        child = ctx.trio_add_child(father, mother, 42)
        # java.util.Random(42): nextBoolean() = true, false (top bits of -1170105035, 234785527)
        assert child == dict(child_first_haplotype=12, from_father_haplotype=3, from_mother_haplotype=8)
        lens, stride = ctx.plan_haplotypes(12, 13, True)
        ctx.materialise()
        for ch, parent in ((12, 3), (13, 8)):
            seq = open(f"{hw}/haplotype_{parent}.fa", "rb").read().split(b"\n", 1)[1]
            assert ctx.copy_haplotype(ch, stride).tobytes() == f">chr1_{ch}\n".encode() + seq, ch
        # truth VCF: the buildVCF merge of the transmitted records
        fa, header = _transmitted(f"{want}/groundTruth_individual_{2 * father}.vcf", 1)
        mo, _ = _transmitted(f"{want}/groundTruth_individual_{2 * mother}.vcf", 0)
        header[-1] = header[-1].rsplit(b"\t", 1)[0] + b"\tindividual_12"
        lines = list(header)
        for idx in sorted(set(fa) | set(mo)):
            gt = b"1|1" if idx in fa and idx in mo else (b"1|0" if idx in fa else b"0|1")
            lines.append((fa.get(idx) or mo[idx]) + b"\t" + gt)
        assert open(f"{got}/groundTruth_individual_12.vcf", "rb").read() == b"\n".join(lines) + b"\n"
        assert len(fa) > 50 and len(mo) > 50
        # a grandchild: the child (individual 6) as a parent
        g = ctx.trio_add_child(6, 0, 7, write_vcf=False)
        assert g["child_first_haplotype"] == 14 and g["from_father_haplotype"] in (12, 13)
        # BASELINE config 5: reads of the child at the maximum coverage, then the 1x/2x/3x prefixes of the shuffled file
        model = str(tmp_path / "hmm.model"); synth.write_hmm_model(model)
        p = dict(depth=3.0, len_mean=15000.0, len_sd=2000.0, len_min=5000, len_max=30000, accuracy_mean_cli=0.995, ratio=(6, 50, 54))
        trg._configure(sp, ctx, p, 21, model, False, sp.READS_PER_READ_STREAMS)
        n, nbytes = ctx.haplotype2reads(12, 9)
        mixed = ctx.reads_mix_copy(0, nbytes).tobytes()
        files = []
        for ch, parent in ((12, 3), (13, 8)):
            hap = np.frombuffer(open(f"{hw}/haplotype_{parent}.fa", "rb").read().split(b"\n")[1], np.uint8)
            # the child's haplotype is sequenced in its own pbsim run: its reads are NOT the parent's reads
            w, _ = oh.pbsim_simulate(oracle_lib, hap, model, seed=sp.haplotype_read_seed(21, ch), per_read_streams=True, ref_num=1, **p)
            files.append(w.tobytes())
            wp, _ = oh.pbsim_simulate(oracle_lib, hap, model, seed=sp.haplotype_read_seed(21, parent), per_read_streams=True, ref_num=1, **p)
            assert wp.tobytes() != files[-1]
        assert mixed == rm.mix_reads(files, 9)
        sizes = [ctx.reads_mix_coverage_prefix(3, c)["bytes"] for c in (1, 2, 3)]
        assert sizes == [len(rm.coverage_prefix(mixed, 3, c)) for c in (1, 2, 3)] and sizes[0] < sizes[1] < sizes[2] <= nbytes


def test_trio_argument_errors(sp, tmp_path):
    with sp.Context(0, 42) as ctx:
        with pytest.raises(sp.SynthpopError, match="sp_population_finish"):
            ctx.trio_add_child(0, 1, 1)
