"""SURVEY.md §8(f) rank 4 — BuildModel (reference src/BuildModel.java:21-47), the builder of the model directory.

CPU half: the oracle's twelve files on a hand-worked input whose tracks, interval classes and length bins are derived by hand
below (including the reference's forward-only cursor running past another contig's intervals, :981-984), evaluated here with
an independent formulation (windowed sums instead of the carried sum, per-bin overlaps instead of first/middle/last bins).
GPU half: the CUDA path's files against the oracle's, byte for byte, on seeded gnomAD-SV-shaped inputs."""
import os
import subprocess

import numpy as np
import pytest

import oracle_helpers as oh  # noqa: F401
from test_oracle_kat import java_layout

TYPES = ("DEL", "DUP", "INS", "INV")
POPS = ("ALL", "AFR", "AMR", "EAS", "EUR", "OTH")
AF_KEYS = ("AF", "AFR_AF", "AMR_AF", "EAS_AF", "EUR_AF", "OTH_AF")
N_LEN = 20000
Q = 100000
CHROMS = [f"chr{i}" for i in range(1, 23)] + ["chrX", "chrY"]
LENGTH = {c: (3_000_000 if c == "chr1" else 2_000_000 + 10_000 * i) for i, c in enumerate(CHROMS)}
P_ARM = {c: (1_200_000 if c == "chr1" else 800_000 + 7_000 * i) for i, c in enumerate(CHROMS)}


def _hand_case(tmp_path):
    """Inputs small enough to classify by hand. 0-based closed coordinates after the loaders' -1 / -2 (:121-122, :526).

    RepeatMasker: [1000,1099] Alu and [1049,1299] simple repeat overlap by >= 25 -> one interval of class 2; two L1
    fragments 11 bp apart whose consensus ranges abut (orientation C) are compacted (:159-190) -> [5000,5399] class 0;
    chr2 [2000,2499] satellite. TRF: [8000,8099] and [8049,8199] overlap by >= 10 -> [8000,8199] class 1.
    Segdups: chr1 [20000,20999] + [20499,21998] -> [20000,21998]; chr2 [30000,30998]."""
    d = tmp_path
    rm = ["900,1.0,0.0,0.0,chr1,1001,1100,(0),+,AluSx,SINE/Alu,1,100,(200),1",
          "300,1.0,0.0,0.0,chr1,1050,1300,(0),+,(CA)n,Simple_repeat,1,251,(0),2",
          "5000,1.0,0.0,0.0,chr1,5001,5200,(0),C,L1PA3,LINE/L1,(5100),900,701,3",
          "5000,1.0,0.0,0.0,chr1,5211,5400,(0),C,L1PA3,LINE/L1,(5300),700,511,4",
          "700,1.0,0.0,0.0,chr2,2001,2500,(0),+,ALR/Alpha,Satellite/centr,1,500,(0),5",
          "700,1.0,0.0,0.0,chr2_KI270715v1_random,1,50,(0),+,ALR/Alpha,Satellite/centr,1,50,(0),6",
          "700,1.0,0.0,0.0,chr3,1,50,(0),+,ALR/Alpha,Satellite/centr,1,50,(0),7"]   # after the break: never loaded
    (d / "rm.csv").write_text("\n".join(rm) + "\n")
    (d / "trf.csv").write_text("chr1,8001,8100,trf,2,50\nchr1,8050,8200,trf,3,60\n")
    (d / "seg.tsv").write_text("#bin\tchrom\tchromStart\tchromEnd\n1\tchr2\t30001\t31000\n1\tchr1\t20500\t22000\n1\tchr1\t20001\t21001\n")
    (d / "ref.fai").write_text("".join(f"{c}\t{LENGTH[c]}\t5\t60\t61\n" for c in CHROMS) + "chrM\t16569\t5\t60\t61\n")
    (d / "cyto.txt").write_text("".join(f"{c}\t0\t{P_ARM[c] // 2}\tp12\tgneg\n{c}\t{P_ARM[c] // 2}\t{P_ARM[c]}\tp11\tacen\n{c}\t{P_ARM[c]}\t{LENGTH[c]}\tq11\tacen\n"
                                        for c in CHROMS) + "chrM\t0\t16569\tx\tgneg\n")

    def info(t, svlen, ev, afs, with_len=True):
        s = f"END=1;SVTYPE={t};" + (f"SVLEN={svlen};" if with_len else "") + f"EVIDENCE={ev};AN=100"
        return s + "".join(f";{k}={v}" for k, v in zip(AF_KEYS, afs)) + ";POPMAX_AF=0.5"

    def afs(x):
        return [x, x / 2, x / 4, "1e-05", x / 8, 0.0005]

    # (contig, POS, FILTER, type, SVLEN, EVIDENCE, AF) -> hand-derived (class, length bin) or None when the record is skipped
    rec = [("1", 1010, "PASS", "DEL", -120, "SR", 0.02, (2, 1)),          # 25 bp reach: inside interval 0
           ("1", 1330, "PASS", "INS", 300, "PE", 0.3, (2, 5)),            # 250 bp reach: 1299+250 >= 1330
           ("1", 4700, "PASS", "DUP", 1000, "RD", 0.04, (0, 19)),         # 1000 bp reach: cursor leaves interval 0, 5000 <= 5700
           ("1", 4800, "PASS", "BND", 100, "SR", 0.5, None),              # unsupported type
           ("1", 7000, "UNRESOLVED", "DEL", 60, "SR", 0.5, None),         # onlyPass
           ("1", 7100, "pass", "INV", 5000, "PE,SR", 0.004, (4, 99)),     # SR wins: 25 bp; 8000 > 7125; segdup 20000 > 7125
           ("1", 20010, "PASS", "DEL", 49, "SR", 0.5, None),              # shorter than 50
           ("1", 20015, "PASS", "DEL", None, "SR", 0.5, None),            # no SVLEN
           ("1", 20020, "MULTIALLELIC", "DEL", 2000, "RD", 0.7, (3, 39)),  # cursor runs past chr2's [2000,2499] (2499+1000 < 20020): class 4, then segdup
           ("2", 2100, "PASS", "INS", 75, "SR", 0.011, (4, 0)),           # inside chr2's satellite, but the cursor is spent
           ("2", 30500, "PASS", "DUP", 123456, "PE", 0.25, (3, 2468)),
           ("2", 40000, "PASS", "INV", 2000000, "SR", 0.9, (4, 19999)),   # last length bin takes everything longer
           ("X", 5000, "PASS", "DEL", 500, "SR", 1.3, (4, 9)),            # counted, but no allele frequencies from X
           ("GL000192.1", 100, "PASS", "DEL", 500, "SR", 0.5, None),      # ends the scan
           ("3", 100, "PASS", "DEL", 500, "SR", 0.5, None)]
    with open(d / "sv.vcf", "w") as f:
        f.write("##fileformat=VCFv4.2\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
        for c, pos, flt, t, svlen, ev, x, _ in rec:
            f.write(f"{c}\t{pos}\tid\tN\t<{t}>\t9\t{flt}\t{info(t, svlen, ev, afs(x), svlen is not None)}\n")
    out = d / "out"
    out.mkdir()
    argv = [d / "rm.csv", 7, d / "trf.csv", 2, d / "seg.tsv", 3, d / "ref.fai", d / "cyto.txt", d / "sv.vcf", 1, out, 0]
    counted = [(r[0], r[1], r[3], r[6], r[7]) for r in rec if r[7] is not None]
    return [str(x) for x in argv], str(out), counted, afs


def _smooth(a, radius):
    """Centred moving average of counts, ends kept (:1024-1056), as windowed sums."""
    out = a.copy()
    w = 2 * radius + 1
    for k in range(radius, len(a) - radius):
        out[k] = float(a[k - radius:k + radius + 1].sum()) / w
    return out


def _seq_sum(v):
    s = 0.0
    for x in v:
        s += float(x)
    return s


def _fmt(x):
    x = float(x)
    return "NaN" if x != x else java_layout(x)


def _row(vals):
    return "".join(_fmt(x) + "," for x in vals) + "\n"


def _expected_files(counted, afs):
    length = np.zeros((4, 5, N_LEN))
    bins = {c: (LENGTH[c] + Q) // Q for c in CHROMS}
    pos = {(t, c): np.zeros(bins[c]) for t in range(4) for c in CHROMS}
    af = {(p, t): [] for p in range(6) for t in range(4)}
    for contig, position, t, x, (cls, lb) in counted:
        ti = TYPES.index(t)
        length[ti, cls, lb] += 1
        pos[(ti, "chr" + contig)][position // Q] += 1
        if contig not in ("X", "Y"):
            for p, v in enumerate(afs(x)):
                af[(p, ti)].append(float(v))
    files = {}
    cls_sum = np.zeros((4, 5))
    for t in range(4):
        for c in range(5):
            length[t, c] = _smooth(length[t, c], 4)
            cls_sum[t, c] = _seq_sum(length[t, c])
            with np.errstate(invalid="ignore"):
                length[t, c] = length[t, c] / cls_sum[t, c]   # 0/0 = NaN where a class is empty, as in the Java
    type_sum = np.array([_seq_sum(cls_sum[t]) for t in range(4)])
    cls_p = cls_sum / type_sum[:, None]
    type_p = type_sum / _seq_sum(type_sum)
    files["statistics_svType.txt"] = "4\n" + _row(np.cumsum(type_p))
    files["statistics_intervalType.txt"] = "4,5\n" + "".join(_row(np.cumsum(cls_p[t])) for t in range(4))
    files["statistics_length.txt"] = f"4,5,{N_LEN}\n" + "".join(_row(np.cumsum(length[t, c])) for t in range(4) for c in range(5))
    # allele frequencies (:1076-1094): bin j is written over element j of the sorted array while it is still being read
    for p in range(6):
        text = "4,10000\n"
        for t in range(4):
            arr = sorted(af[(p, t)])
            n = len(arr)
            arr = arr + [0.0] * (10000 - n)
            lo, hi = arr[0], arr[n - 1]
            quantum = (hi - lo) / 9999
            k = 0
            for j in range(10000):
                boundary = lo + j * quantum
                s = 0.0
                while k < n and arr[k] < boundary:
                    s += 1
                    k += 1
                arr[j] = s / n
            text += _fmt(lo) + "," + _fmt(hi) + "," + _row(np.cumsum(np.array(arr)))
        files[f"statistics_alleleFrequency_{POPS[p]}.txt"] = text
    # positions (:740-802): every autosome's arms stretched onto chr1's; a source bin's mass spread evenly over its image
    p1, l1 = P_ARM["chr1"], LENGTH["chr1"]
    text = f"4,{bins['chr1']}\n"
    for t in range(4):
        sm = {c: _smooth(pos[(t, c)], 5) for c in CHROMS}
        out = sm["chr1"].copy()

        def spread(frm, to, per_bp):
            for h in range(frm // Q, to // Q + 1):
                out[h] += per_bp * (min(to, (h + 1) * Q - 1) - max(frm, h * Q) + 1)

        for c in CHROMS[1:22]:
            pj, lj = P_ARM[c], LENGTH[c]
            rp, rq = p1 / pj, (l1 - p1) / (lj - pj)
            for k in range(bins[c]):
                first, last = k * Q, min((k + 1) * Q, lj) - 1
                cnt = float(sm[c][k])
                parts = []   # (is_p, first, last, mass)
                if last < pj:
                    parts.append((True, first, last, cnt))
                elif first < pj:
                    parts.append((True, first, pj - 1, ((pj - first) * cnt) / Q))
                    parts.append((False, pj, last, ((last - pj + 1) * cnt) / Q))
                else:
                    parts.append((False, first, last, cnt))
                for is_p, a, b, mass in parts:
                    if is_p:
                        frm, to = int(a * rp), int(b * rp)
                        if frm < p1:
                            spread(frm, min(to, p1 - 1), mass / (to - frm + 1))
                    else:
                        frm, to = p1 + int((a - pj) * rq), p1 + int((b - pj) * rq)
                        spread(frm, min(to, l1 - 1), mass / (to - frm + 1))
        text += _row(np.cumsum(out / _seq_sum(out)))
    files["statistics_position.txt"] = text
    files["repeatIntervals.txt"] = "3\n2,1000,1299\n0,5000,5399\n1,8000,8199\n"
    files["segdups.txt"] = "1\n20000,21998\n"
    return files


def test_oracle_build_model_on_a_hand_worked_case(oracle_bin, tmp_path):
    argv, out, counted, afs = _hand_case(tmp_path)
    p = subprocess.run([oracle_bin, "BuildModel"] + argv, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert "Total SVs in chr1: 5. With allele frequency >=0.01: 4" in p.stderr   # A, B, C, E, G; E's 0.004 is below the line
    want = _expected_files(counted, afs)
    assert sorted(os.listdir(out)) == sorted(want)
    for name, text in want.items():
        got = open(os.path.join(out, name)).read()
        assert got == text, name


def test_oracle_build_model_rejects_what_the_java_rejects(oracle_bin, tmp_path):
    argv, out, _, _ = _hand_case(tmp_path)
    # fewer segdup rows than declared: Arrays.sort over null entries (:528)
    bad = list(argv)
    bad[5] = "4"
    p = subprocess.run([oracle_bin, "BuildModel"] + bad, capture_output=True, text=True)
    assert p.returncode == 1 and "NullPointerException" in p.stderr
    # an INFO value that is not closed by ';' (:950-951)
    vcf = open(argv[8]).read().replace(";POPMAX_AF=0.5", "")
    open(argv[8], "w").write(vcf)
    p = subprocess.run([oracle_bin, "BuildModel"] + argv, capture_output=True, text=True)
    assert p.returncode == 1 and "StringIndexOutOfBoundsException" in p.stderr


def test_synthetic_inputs_have_the_shape_the_loaders_expect(tmp_path):
    from synthpop import synth
    r = synth.write_build_model_inputs(str(tmp_path), n_records=2000, n_repeats=4000, n_trf=800, n_segdups=100)
    rows = [ln.split(",") for ln in open(r["files"]["rm"]).read().splitlines()]
    assert len(rows) == r["n_repeatmasker"] and all(len(x) == 15 for x in rows)
    contigs = [x[4] for x in rows]
    assert contigs[:-2] == sorted(contigs[:-2]) and contigs[-2:] == ["chr1_KI270706v1_random", "chrM"]
    starts = {}
    for x in rows[:-2]:
        assert int(x[5]) >= starts.get(x[4], 0)   # sorted by start inside a contig
        starts[x[4]] = int(x[5])
    body = [ln.split("\t") for ln in open(r["files"]["vcf"]).read().splitlines() if not ln.startswith("#")]
    assert len(body) == r["n_vcf_records"] + 1 and body[-1][0] == "GL000192.1"
    assert all(";AF=" in x[7] and ";OTH_AF=" in x[7] and x[7].count("EVIDENCE=") == 1 for x in body[:-1])


def _compare_dirs(want, got):
    from synthpop import synth
    assert sorted(os.listdir(got)) == sorted(synth.BUILD_MODEL_OUTPUTS)
    for name in synth.BUILD_MODEL_OUTPUTS:
        a, b = open(os.path.join(want, name), "rb").read(), open(os.path.join(got, name), "rb").read()
        assert a == b, name


@pytest.mark.gpu
def test_gpu_build_model_hand_case_and_errors(sp, oracle_bin, tmp_path):
    argv, out, counted, afs = _hand_case(tmp_path)
    with sp.Context(0, 1) as ctx:
        ctx.build_model(argv)
        tm = ctx.build_model_timing()
        assert tm["records"] == len(counted) == 9 and tm["allele_frequencies"] == 8 * 6
        want = _expected_files(counted, afs)
        for name, text in want.items():
            assert open(os.path.join(out, name)).read() == text, name
        bad = list(argv)
        bad[5] = "4"
        with pytest.raises(sp.SynthpopError, match="NullPointerException"):
            ctx.build_model(bad)
        vcf = open(argv[8]).read().replace(";POPMAX_AF=0.5", "")
        open(argv[8], "w").write(vcf)
        with pytest.raises(sp.SynthpopError, match="StringIndexOutOfBoundsException"):
            ctx.build_model(argv)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,only_pass", [(1, 1), (2, 0)])
def test_gpu_build_model_is_byte_identical_to_the_oracle(sp, oracle_bin, tmp_path, seed, only_pass):
    from synthpop import synth
    r = synth.write_build_model_inputs(str(tmp_path), seed=seed)
    argv = [str(x) for x in r["argv"]]
    argv[9] = str(only_pass)
    want, got = str(tmp_path / "want"), str(tmp_path / "got")
    os.makedirs(want), os.makedirs(got)
    subprocess.run([oracle_bin, "BuildModel"] + argv[:10] + [want, "0"], check=True, capture_output=True)
    with sp.Context(0, 1) as ctx:
        l0 = ctx.kernel_launches
        ctx.build_model(argv[:10] + [got, "0"])
        assert ctx.kernel_launches - l0 == 4   # count, keys, segmented sort, keys back
        tm = ctx.build_model_timing()
        assert 20000 < tm["records"] <= r["n_vcf_records"] and tm["allele_frequencies"] % 6 == 0
    _compare_dirs(want, got)
    # the files are what SimulateHaplotypes' loaders read: cumulative rows ending at 1
    last = [float(x) for x in open(os.path.join(got, "statistics_svType.txt")).read().split("\n")[1].split(",")[:-1]]
    assert abs(last[-1] - 1.0) < 1e-12 and last == sorted(last)


@pytest.mark.gpu
def test_gpu_build_model_dies_like_the_java_when_a_type_has_no_autosomal_record(sp, oracle_bin, tmp_path):
    from synthpop import synth
    r = synth.write_build_model_inputs(str(tmp_path), seed=3, n_records=3000, n_repeats=6000, n_trf=1000, n_segdups=100, drop_type="INV")
    argv = [str(x) for x in r["argv"]]
    p = subprocess.run([oracle_bin, "BuildModel"] + argv, capture_output=True, text=True)
    assert p.returncode == 1 and "ArrayIndexOutOfBoundsException" in p.stderr
    with sp.Context(0, 1) as ctx:
        with pytest.raises(sp.SynthpopError, match="ArrayIndexOutOfBoundsException"):
            ctx.build_model(argv)


@pytest.mark.gpu
def test_gpu_built_model_feeds_simulate_haplotypes(sp, oracle_bin, tmp_path):
    """BuildModel -> SimulateHaplotypes through the files: the model directory, repeatIntervals.txt and segdups.txt written
    by the CUDA path are loaded by sp_simulate_population and by the oracle's SimulateHaplotypes; both must agree."""
    from synthpop import synth
    from test_population_gpu import _cmp_dirs
    r = synth.write_build_model_inputs(str(tmp_path / "bm"), seed=5, scale=25, max_sv_length=10000)
    with sp.Context(0, 11) as ctx:
        ctx.build_model(r["argv"])
    model, L = r["out_dir"], r["chr1_length"]
    bases, _, _ = synth.random_reference(L, 5)
    ref, ins = str(tmp_path / "chr1.fa"), str(tmp_path / "insertions.txt")
    synth.write_fasta(ref, "chr1", bases)
    synth.write_insertions(ins, 2500, 5, max_len=10500)   # the smoothing leaks INS mass four bins past 10000
    n_ind = 4
    want, got = str(tmp_path / "want"), str(tmp_path / "got")
    os.makedirs(want), os.makedirs(got)
    argv = [model, "EUR", ref, model + "/repeatIntervals.txt", model + "/segdups.txt", ins, "2500", "0.01", "400", str(n_ind)]
    subprocess.run([oracle_bin, "--quiet", "--seed", "77", "SimulateHaplotypes"] + argv + [want], check=True)
    with sp.Context(0, 77) as ctx:
        ctx.simulate_haplotypes(argv + [got])
    _cmp_dirs(want, got, n_ind)
    assert os.path.getsize(os.path.join(got, "groundTruth_joint.vcf")) > 20000
