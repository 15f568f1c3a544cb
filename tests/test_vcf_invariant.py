"""The invariant DESIGN.md §5 names as a pin that shares no code with either implementation: applying an individual's truth
VCF (row W, src/SimulateHaplotypes.java:1027-1066, :1243-1279) to the reference must reproduce both of its haplotypes
(row M, :917-960). The applier below is written from the VCF semantics alone (POS/REF/ALT/SVTYPE/END/GT), not from
buildHaplotypes: DEL removes REF[1:] after the anchor at POS; INS (POS = first base after the splice point, REF = that base,
ALT = REF + payload) puts ALT[1:] in front of POS, which is where the reference splices it (:937-941) although a VCF reader
would put it behind; DUP repeats the 1-based interval [POS..END]; INV replaces [POS..END] by its reverse complement."""
import os
import subprocess

import numpy as np
import pytest

import oracle_helpers as oh  # noqa: F401

COMP = np.full(256, ord("N"), np.uint8)
for a, b in zip(b"acgtACGT", b"tgcaTGCA"):
    COMP[a] = b


def apply_vcf(ref: np.ndarray, vcf_path: str, allele: int) -> np.ndarray:
    """Haplotype `allele` (0/1) of the single sample of an individual truth VCF, applied to `ref` (uint8)."""
    pieces = []
    cursor = 0  # next reference base (0-based) not yet emitted
    for line in open(vcf_path, "rb"):
        if line.startswith(b"#"):
            continue
        c = line.rstrip(b"\n").split(b"\t")
        if c[-1].split(b"|")[allele] != b"1":
            continue
        pos = int(c[1])
        info = dict(kv.split(b"=", 1) for kv in c[7].split(b";") if b"=" in kv)
        svtype = info[b"SVTYPE"]
        end = int(info[b"END"])
        if svtype == b"DEL":
            # REF = anchor + deleted bases; 1-based POS of the anchor = pos -> deleted bases are 0-based [pos, pos+len)
            assert ref[pos - 1:pos - 1 + len(c[3])].tobytes() == c[3], "REF of a DEL must be the reference substring"
            assert c[4] == c[3][:1] and end == pos + len(c[3]) - 1 and int(info[b"SVLEN"]) == -(len(c[3]) - 1)
            pieces.append(ref[cursor:pos]); cursor = pos + len(c[3]) - 1
        elif svtype == b"INS":
            # POS = position+1 (1-based) of the base AFTER which... the reference writes REF=ref[position], ALT=REF+payload and
            # splices the payload BEFORE ref[position] (:937-941): the anchor convention is checked, the splice point is POS-1
            assert ref[pos - 1:pos].tobytes() == c[3] and c[4][:1] == c[3] and int(info[b"SVLEN"]) == len(c[4]) - 1
            pieces.append(ref[cursor:pos - 1]); pieces.append(np.frombuffer(c[4][1:], np.uint8)); cursor = pos - 1
        elif svtype == b"DUP":
            assert c[4] == b"<DUP>"
            assert ref[pos - 1:pos].tobytes() == c[3]
            pieces.append(ref[cursor:end]); pieces.append(ref[pos - 1:end]); cursor = end
        elif svtype == b"INV":
            assert c[4] == b"<INV>"
            assert ref[pos - 1:pos].tobytes() == c[3]
            pieces.append(ref[cursor:pos - 1]); pieces.append(COMP[ref[pos - 1:end]][::-1]); cursor = end
        else:
            raise AssertionError(svtype)
    pieces.append(ref[cursor:])
    return np.concatenate(pieces)


def _reference_bases(inp):
    return np.frombuffer(open(inp.reference_file, "rb").read().split(b"\n", 1)[1].replace(b"\n", b""), np.uint8)


def _check(out_dir, hap_dir, ref, n_ind):
    n_records = 0
    for k in range(n_ind):
        vcf = os.path.join(out_dir, f"groundTruth_individual_{2 * k}.vcf")
        n_records += sum(1 for l in open(vcf, "rb") if not l.startswith(b"#"))
        for allele in (0, 1):
            h = 2 * k + allele
            want = open(os.path.join(hap_dir, f"haplotype_{h}.fa"), "rb").read()
            got = f">chr1_{h}\n".encode() + apply_vcf(ref, vcf, allele).tobytes() + b"\n"
            assert got == want, f"haplotype {h}: the truth VCF applied to the reference is not the haplotype"
    return n_records


def test_oracle_vcf_applied_to_reference_is_the_haplotype(oracle_bin, tmp_path):
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 1_800_000, n_insertions=2000, seed=12)
    out, haps = str(tmp_path / "out"), str(tmp_path / "haps")
    os.makedirs(out); os.makedirs(haps)
    n_ind = 5
    subprocess.run([oracle_bin, "--quiet", "--seed", "7", "SimulateHaplotypes"] + inp.simulate_argv(out, n_ind, n_frequent=400), check=True)
    subprocess.run([oracle_bin, "--quiet", "PrintHaplotypes", "0", str(2 * n_ind - 1), inp.reference_file,
                    out + "/haplotype2variants.txt", out + "/variants.txt", haps], check=True)
    assert _check(out, haps, _reference_bases(inp), n_ind) > 300


@pytest.mark.gpu
def test_gpu_vcf_applied_to_reference_is_the_haplotype(sp, tmp_path):
    """The same on the product's own files: SimulateHaplotypes + PrintHaplotypes through the C ABI, no oracle involved."""
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), 3_000_000, n_insertions=3000, seed=21)
    out, haps = str(tmp_path / "out"), str(tmp_path / "haps")
    os.makedirs(out); os.makedirs(haps)
    n_ind = 6
    with sp.Context(0, 99) as ctx:
        ctx.simulate_haplotypes(inp.simulate_argv(out, n_ind, n_frequent=500))
        ctx.print_haplotypes(0, 2 * n_ind - 1, inp.reference_file, out + "/haplotype2variants.txt", out + "/variants.txt", haps)
    assert _check(out, haps, _reference_bases(inp), n_ind) > 500
