"""Seeded generator of synthetic inputs in the reference's on-disk formats.

Formats follow the reference's serialisers / parsers:
  * model directory ........ src/BuildModel.java:1135-1197 (serializeStatistics) / :1200-1282 (parser)
  * repeatIntervals.txt .... src/BuildModel.java:464-480   ("count", then "type,start,end")
  * segdups.txt ............ src/BuildModel.java:581-598   ("count", then "start,end")
  * insertion strings ...... src/SimulateHaplotypes.java:313-337 (one string per line, sorted by length)
  * reference FASTA ........ src/SimulateHaplotypes.java:119-134 (one header line, then sequence lines)
  * length-bin files ....... src/BuildReadLengthBins.java:42-105 ("<prefix><i>.bin", 4-line FASTQ records)
  * pbsim2 HMM model ....... docker/pbsim2.cpp:2583-2647 ("<acc> IP|EP|TP <state> <probs...>")

Everything is a pure function of (shape parameters, seed); numpy only.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np

GENERATOR_SEED = 20261104
SV_TYPES = ("DEL", "DUP", "INS", "INV")
AF_NAMES = ("ALL", "AFR", "AMR", "EAS", "EUR", "OTH")
N_LENGTH_BINS = 1 + (1000000 - 50 + 1) // 50  # src/BuildModel.java:919
AF_NBINS = 10000

GRCH38_LENGTHS = {
    "chr1": 248956422, "chr2": 242193529, "chr3": 198295559, "chr4": 190214555, "chr5": 181538259,
    "chr6": 170805979, "chr7": 159345973, "chr8": 145138636, "chr9": 138394717, "chr10": 133797422,
    "chr11": 135086622, "chr12": 133275309, "chr13": 114364328, "chr14": 107043718, "chr15": 101991189,
    "chr16": 90338345, "chr17": 83257441, "chr18": 80373285, "chr19": 58617616, "chr20": 64444167,
    "chr21": 46709983, "chr22": 50818468, "chrX": 156040895, "chrY": 57227415,
}


@dataclass
class SynthInputs:
    root: str
    reference_file: str
    model_dir: str
    repeat_intervals_file: str
    segdups_file: str
    insertions_file: str
    n_insertions: int
    reference_length: int
    extra: dict = field(default_factory=dict)

    def simulate_argv(self, out_dir: str, n_individuals: int, n_frequent: int = 5000,
                      af_threshold: str = "0.01", population: str = "ALL") -> list[str]:
        """Positional argv of SimulateHaplotypes (src/SimulateHaplotypes.java:76-86)."""
        return [self.model_dir, population, self.reference_file, self.repeat_intervals_file, self.segdups_file,
                self.insertions_file, str(self.n_insertions), af_threshold, str(n_frequent), str(n_individuals),
                out_dir]


def random_reference(length: int, seed: int = GENERATOR_SEED):
    """uint8[length] of i.i.d. ACGT with N runs at both ends and a centromere block, plus interval tables.

    Returns (bases, repeat_intervals[int32 n x 3: type,start,end], segdups[int32 m x 2]).
    """
    rng = np.random.default_rng(seed)
    bases = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=length, dtype=np.uint8)].copy()
    end_n = min(10000, max(1, length // 100))
    bases[:end_n] = ord("N")
    bases[length - end_n:] = ord("N")
    c0 = int(0.49 * length)
    c1 = min(length - end_n, c0 + int(0.07 * length))
    bases[c0:c1] = ord("N")

    # repeat intervals: alternating gap / interval lengths ~ Geometric(mean 300)
    n_est = int(length / 600 * 1.3) + 16
    gaps = rng.geometric(1.0 / 300.0, size=n_est).astype(np.int64) + 30
    lens = rng.geometric(1.0 / 300.0, size=n_est).astype(np.int64)
    starts = np.cumsum(gaps + np.concatenate(([0], lens[:-1]))) + end_n
    ends = starts + lens - 1
    keep = ends < length - end_n - 1
    starts, ends = starts[keep], ends[keep]
    types = rng.choice(3, size=starts.size, p=[0.85, 0.10, 0.05]).astype(np.int64)
    repeat_intervals = np.stack([types, starts, ends], axis=1).astype(np.int32)

    # soft-mask (lower-case) exactly on repeat intervals
    delta = np.zeros(length + 1, dtype=np.int32)
    np.add.at(delta, starts, 1)
    np.add.at(delta, ends + 1, -1)
    mask = np.cumsum(delta[:-1]) > 0
    mask &= bases != ord("N")
    bases[mask] |= 0x20

    # segdups: ~5 % of the sequence, lengths log-uniform in [1 kbp, 100 kbp] (scaled down for tiny references)
    hi = min(100000.0, max(2000.0, length / 50))
    target = int(0.05 * length)
    seg = []
    total = 0
    while total < target:
        ln = int(np.exp(rng.uniform(np.log(1000.0), np.log(hi))))
        st = int(rng.integers(end_n, max(end_n + 1, length - end_n - ln - 1)))
        seg.append((st, st + ln - 1))
        total += ln
    seg.sort()
    merged = []
    for s, e in seg:
        if merged and s - merged[-1][1] < 25:
            merged[-1] = (merged[-1][0], max(merged[-1][1], e))
        else:
            merged.append((s, e))
    segdups = np.array(merged, dtype=np.int32).reshape(-1, 2)
    return bases, repeat_intervals, segdups


def write_fasta(path: str, header: str, bases: np.ndarray, width: int = 60) -> None:
    n = bases.size
    full = n // width
    with open(path, "wb") as f:
        f.write(b">" + header.encode() + b"\n")
        if full:
            block = np.empty((full, width + 1), dtype=np.uint8)
            block[:, :width] = bases[: full * width].reshape(full, width)
            block[:, width] = ord("\n")
            f.write(block.tobytes())
        if n % width:
            f.write(bases[full * width:].tobytes() + b"\n")


def _cum_row(weights: np.ndarray) -> np.ndarray:
    c = np.cumsum(weights.astype(np.float64))
    c /= c[-1]
    c[-1] = 1.0
    nz = np.nonzero(weights)[0]
    if nz.size:
        c[nz[-1]:] = 1.0  # every cumulative row reaches exactly 1.0 at its last non-empty bin
    return c


def _fmt_row(vals) -> str:
    return ",".join(repr(float(v)) for v in vals) + ","


def write_model(model_dir: str, seed: int = GENERATOR_SEED, max_sv_length: int = 1000000,
                max_ins_length: int = 10000, frac_frequent: float | None = None) -> None:
    """BuildModel-format model directory. `max_sv_length` caps the mass of the length CDFs (tests on small
    references); INS rows carry zero mass above `max_ins_length`."""
    os.makedirs(model_dir, exist_ok=True)
    rng = np.random.default_rng(seed + 1)
    with open(os.path.join(model_dir, "statistics_svType.txt"), "w") as f:
        f.write("4\n" + _fmt_row([0.52, 0.67, 0.998, 1.0]) + "\n")
    with open(os.path.join(model_dir, "statistics_intervalType.txt"), "w") as f:
        f.write("4,5\n")
        for _ in range(4):
            w = rng.dirichlet([2.0] * 5)
            f.write(_fmt_row(_cum_row(w)) + "\n")
    centers = 50.0 + 50.0 * np.arange(N_LENGTH_BINS) + 25.0
    base = 1.0 / centers  # log-uniform density integrated over equal-width bins
    with open(os.path.join(model_dir, "statistics_length.txt"), "w") as f:
        f.write(f"4,5,{N_LENGTH_BINS}\n")
        for t in range(4):
            for _ in range(5):
                w = base * rng.uniform(0.8, 1.2)
                w = w + 40.0 * base.sum() / N_LENGTH_BINS * np.exp(-0.5 * ((centers - 300.0) / 60.0) ** 2)
                w = w + 8.0 * base.sum() / N_LENGTH_BINS * np.exp(-0.5 * ((centers - 6000.0) / 300.0) ** 2)
                cap = max_ins_length if t == 2 else max_sv_length
                # INS: length/50 must stay below nBins = 1 + max_ins_length/50  =>  p <= max_ins_length/50 - 1
                last_bin = (cap // 50 - 1) if t == 2 else max(0, (cap - 50 - 49) // 50)
                w[last_bin + 1:] = 0.0
                f.write(_fmt_row(_cum_row(w)) + "\n")
    af_min, af_max = 4.6e-5, 1.0
    q = (af_max - af_min) / (AF_NBINS - 1)
    af_centers = af_min + q * (np.arange(AF_NBINS) + 0.5)
    for name in AF_NAMES:
        with open(os.path.join(model_dir, f"statistics_alleleFrequency_{name}.txt"), "w") as f:
            f.write(f"4,{AF_NBINS}\n")
            for _ in range(4):
                w = 1.0 / af_centers  # mass ∝ 1/AF
                if frac_frequent is not None:
                    below = af_centers < 0.01
                    w[below] *= (1.0 - frac_frequent) / w[below].sum()
                    w[~below] *= frac_frequent / w[~below].sum()
                f.write(repr(af_min) + "," + repr(af_max) + "," + _fmt_row(_cum_row(w)) + "\n")
    with open(os.path.join(model_dir, "statistics_position.txt"), "w") as f:
        f.write("4,10\n")
        for _ in range(4):
            f.write(_fmt_row(_cum_row(np.ones(10))) + "\n")


def write_insertions(path: str, n: int, seed: int = GENERATOR_SEED, max_len: int = 10000) -> int:
    """`n` distinct ACGT strings sorted by length (50..max_len), at least one in the top 50-bp bin."""
    rng = np.random.default_rng(seed + 2)
    lens = np.exp(rng.uniform(np.log(50.0), np.log(max_len), size=n)).astype(np.int64)
    lens[: max(1, n // 200)] = max_len  # the top bin is never empty
    lens = np.sort(np.clip(lens, 50, max_len))
    alphabet = np.frombuffer(b"ACGT", dtype=np.uint8)
    blob = alphabet[rng.integers(0, 4, size=int(lens.sum()), dtype=np.uint8)]
    out = np.empty(blob.size + n, dtype=np.uint8)
    ends = np.cumsum(lens) + np.arange(n)
    out[:] = 0
    out[ends] = ord("\n")
    out[out == 0] = blob
    with open(path, "wb") as f:
        f.write(out.tobytes())
    return n


def generate_population_inputs(root: str, length: int, seed: int = GENERATOR_SEED, n_insertions: int = 30000,
                               max_sv_length: int | None = None, frac_frequent: float | None = None,
                               reuse: bool = False) -> SynthInputs:
    """reuse=True: the files under `root` were already written by another process with the same arguments (the ranks of one
    box share them); only the in-memory arrays are rebuilt."""
    os.makedirs(root, exist_ok=True)
    bases, rep, seg = random_reference(length, seed)
    ref = os.path.join(root, "reference.fa")
    repf = os.path.join(root, "repeatIntervals.txt")
    segf = os.path.join(root, "segdups.txt")
    model_dir = os.path.join(root, "model")
    insf = os.path.join(root, "insertions.txt")
    if not reuse:
        write_fasta(ref, "chr1", bases)
        with open(repf, "w") as f:
            f.write(f"{rep.shape[0]}\n")
            np.savetxt(f, rep, fmt="%d", delimiter=",")
        with open(segf, "w") as f:
            f.write(f"{seg.shape[0]}\n")
            np.savetxt(f, seg, fmt="%d", delimiter=",")
        if max_sv_length is None:
            max_sv_length = min(1000000, max(100, length // 40))
        max_ins = min(10000, max(100, (length // 100) // 50 * 50))
        write_model(model_dir, seed, max_sv_length=max_sv_length, max_ins_length=max_ins, frac_frequent=frac_frequent)
        write_insertions(insf, n_insertions, seed, max_len=max_ins)
    return SynthInputs(root=root, reference_file=ref, model_dir=model_dir, repeat_intervals_file=repf,
                       segdups_file=segf, insertions_file=insf, n_insertions=n_insertions,
                       reference_length=length, extra={"bases": bases, "repeat_intervals": rep, "segdups": seg})


def write_hmm_model(path: str, seed: int = GENERATOR_SEED, accuracies=range(74, 101), n_states: int = 5,
                    n_columns: int = 94, quality: str = "high") -> None:
    """Synthetic quality-code HMM in the text format set_hmm_model parses (docker/pbsim2.cpp:2583-2647):
    "<acc> IP <state> <p>", "<acc> EP <state> <94 emission probabilities>", "<acc> TP <state> <p_1..p_n>".
    (The real P6C4.model is not part of the reference repository.)

    quality="phred": the emission windows of accuracy class a are centred on the Phred value of its error rate,
    -10*log10((100.5-a)/100) (Q6 at 74 %, Q13 at 95 %, Q23 at 100 %), shifted by 2 per state — what a model trained on
    real reads looks like, since PBSIM2 groups training reads by the accuracy their qualities imply.
    quality="high": centres at (a-70)*0.9+3*state (Q16..Q42): a stress case for the deletion rejection loop, whose
    acceptance probability is 10^(-q/10) (docker/pbsim2.cpp:2284-2292)."""
    import math
    rng = np.random.default_rng(seed + 3)
    with open(path, "w") as f:
        for acc in accuracies:
            ip = rng.dirichlet([2.0] * n_states)
            for s in range(1, n_states + 1):
                f.write(f"{acc} IP {s} {ip[s - 1]:.6f}\n")
            for s in range(1, n_states + 1):
                if quality == "phred":
                    centre = int(round(-10.0 * math.log10((100.5 - acc) / 100.0))) + 2 * (s - (n_states + 1) // 2)
                else:
                    centre = int(round((acc - 70) * 0.9 + 3 * s))
                centre = min(n_columns - 2, max(2, centre))
                w = np.zeros(n_columns)
                lo, hi = max(0, centre - 6), min(n_columns, centre + 7)
                w[lo:hi] = rng.random(hi - lo) + 0.05
                w /= w.sum()
                f.write(f"{acc} EP {s} " + " ".join(f"{x:.6f}" for x in w) + "\n")
            for s in range(1, n_states + 1):
                tp = rng.dirichlet([1.5] * n_states)
                f.write(f"{acc} TP {s} " + " ".join(f"{x:.6f}" for x in tp) + "\n")


def cumulative_bimodal(n_bins: int, bin_length: int, mean_left: float, std_left: float, mean_right: float,
                       std_right: float, weight_left: float) -> np.ndarray:
    """buildCumulativeBimodal (src/SampleReadsFromLengthBins.java:87-102) with libm erf in place of commons-math3's
    NormalDistribution.probability — the table is an INPUT of the native side (the Java host computes it with
    commons-math; last-ulp differences between the two erf implementations are outside the drop-in boundary)."""
    import math

    def prob(mean, sd, x0, x1):
        d = sd * math.sqrt(2.0)
        return 0.5 * (math.erf((x1 - mean) / d) - math.erf((x0 - mean) / d))

    c = np.array([prob(mean_left, std_left, i * bin_length, (i + 1) * bin_length) * weight_left +
                  prob(mean_right, std_right, i * bin_length, (i + 1) * bin_length) * (1.0 - weight_left)
                  for i in range(n_bins)], dtype=np.float64)
    s = 0.0
    for v in c:  # same summation order as the Java loop
        s += float(v)
    c = c / s
    for i in range(1, n_bins):
        c[i] = c[i] + c[i - 1]
    return c


def write_length_bins(prefix: str, n_bins: int, bin_length: int, records_per_bin, seed: int = GENERATOR_SEED,
                      with_names: bool = True) -> int:
    """Length-bin files in the format BuildReadLengthBins writes (src/BuildReadLengthBins.java:42-105): file
    <prefix><i>.bin holds 4-line FASTQ records whose sequence length lies in [i*bin_length, (i+1)*bin_length).
    records_per_bin: int or sequence of n_bins ints. Returns the total number of sequence bases written."""
    rng = np.random.default_rng(seed + 11)
    per = [int(records_per_bin)] * n_bins if np.isscalar(records_per_bin) else [int(x) for x in records_per_bin]
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    total = 0
    for b in range(n_bins):
        parts = []
        for j in range(per[b]):
            lo = max(1, b * bin_length)
            ln = int(rng.integers(lo, (b + 1) * bin_length))
            seq = acgt[rng.integers(0, 4, size=ln)].tobytes()
            qual = (rng.integers(33 + 20, 33 + 60, size=ln).astype(np.uint8)).tobytes()
            name = (f"@m64_{b}_{j}/ccs" if with_names else "@r").encode()
            parts += [name, b"\n", seq, b"\n+\n", qual, b"\n"]
            total += ln
        with open(f"{prefix}{b}.bin", "wb") as f:
            f.write(b"".join(parts))
    return total


GRCH38_P_ARM_END = {   # approximate last p band end of each chromosome (cytoband file of the reference build)
    "chr1": 123400000, "chr2": 93900000, "chr3": 90900000, "chr4": 50000000, "chr5": 48800000, "chr6": 59800000,
    "chr7": 60100000, "chr8": 45200000, "chr9": 43000000, "chr10": 39800000, "chr11": 53400000, "chr12": 35500000,
    "chr13": 17700000, "chr14": 17200000, "chr15": 19000000, "chr16": 36800000, "chr17": 25100000, "chr18": 18500000,
    "chr19": 26200000, "chr20": 28100000, "chr21": 12000000, "chr22": 15000000, "chrX": 60600000, "chrY": 10400000,
}


def write_build_model_inputs(root: str, seed: int = GENERATOR_SEED, n_records: int = 40000, n_repeats: int = 120000,
                             n_trf: int = 30000, n_segdups: int = 3000, drop_type: str | None = None, scale: int = 1,
                             max_sv_length: int = 3000000) -> dict:
    """The nine inputs of BuildModel.main (src/BuildModel.java:22-33) in the formats its loaders parse: a cleaned RepeatMasker
    table (:101-157: comma-separated, chromosome at column 4, strand at 8, name 9, family 10, consensus range 11-13), a
    cleaned TRF table (:233-247), a UCSC segdups table with a header (:519-526), reference .fai and cytobands (:666-708)
    and a gnomAD-SV-style VCF (:933-1020: CHROM without "chr", INFO with SVLEN, SVTYPE, EVIDENCE and the six *_AF keys).
    Contigs come in the order the Java assumes (chr1, chr10..chr19, chr2, chr20..22, chr3..chr9, chrX, chrY, other).
    drop_type leaves one SV type out of the autosomes (the case the Java dies on). scale divides every chromosome length
    (a karyotype small enough to run SimulateHaplotypes on the result); max_sv_length caps SVLEN. Returns the argv and the
    row counts."""
    os.makedirs(root, exist_ok=True)
    rng = np.random.default_rng(seed + 23)
    GRCH38_LENGTHS = {c: v // scale for c, v in globals()["GRCH38_LENGTHS"].items()}
    GRCH38_P_ARM_END = {c: v // scale for c, v in globals()["GRCH38_P_ARM_END"].items()}
    order = sorted(GRCH38_LENGTHS)   # the lexicographic contig order
    total = float(sum(GRCH38_LENGTHS.values()))
    families = [("L1PA3", "LINE/L1"), ("AluSx", "SINE/Alu"), ("ALR/Alpha", "Satellite/centr"), ("(CA)n", "Simple_repeat"),
                ("AT_rich", "Low_complexity"), ("MLT1B", "LTR/ERVL-MaLR"), ("Charlie1", "DNA/hAT-Charlie"), ("HSATII", "Satellite")]
    rm_lines, trf_lines, repeat_spots = [], [], {}
    for chrom in order:
        L = GRCH38_LENGTHS[chrom]
        n = max(20, int(n_repeats * L / total))
        starts = np.sort(rng.integers(1000, L - 20000, size=n))
        spots = []
        k = 0
        while k < n:
            name, fam = families[int(rng.integers(0, len(families)))]
            strand = "+" if rng.random() < 0.5 else "C"
            pos = int(starts[k])
            cons = int(rng.integers(1, 3000))
            pieces = 1 + int(rng.random() < 0.3) + int(rng.random() < 0.1)   # fragments the compaction joins
            for _ in range(pieces):
                ln = int(rng.integers(30, 3000))
                a, b = pos + 1, pos + ln
                ca, cb = cons, cons + ln - 1
                left = f"({int(rng.integers(0, 500))})"
                cols = [str(int(rng.integers(200, 30000))), "12.1", "0.5", "1.2", chrom, str(a), str(b), f"({L - b})", strand, name, fam]
                cols += [str(ca), str(cb), left] if strand == "+" else [left, str(cb), str(ca)]
                cols.append(str(len(rm_lines) + 1))
                rm_lines.append(",".join(cols))
                spots.append(a)
                if rng.random() < 0.3 and ln > 80:   # another element lying inside or across this one (what buildRepeatIntervals joins)
                    n2, f2 = families[int(rng.integers(0, len(families)))]
                    a2 = a + int(rng.integers(1, ln - 40))
                    b2 = a2 + int(rng.integers(30, 1500))
                    rm_lines.append(",".join([str(int(rng.integers(200, 30000))), "9.0", "0.0", "0.3", chrom, str(a2), str(b2), f"({L - b2})", "+",
                                              n2, f2, "1", str(b2 - a2 + 1), "(0)", str(len(rm_lines) + 1)]))
                    b = max(b, b2)
                gap = int(rng.integers(0, 60))   # <= 25 joins, above does not
                pos = b + gap
                cons = cb + 1 + int(rng.integers(0, 40)) if strand == "+" else max(1, ca - ln - int(rng.integers(0, 40)))
            k += 1
            while k < n and int(starts[k]) <= pos:   # keep the table sorted by start
                k += 1
        repeat_spots[chrom] = np.array(spots, dtype=np.int64)
        m = max(5, int(n_trf * L / total))
        tstarts = np.sort(rng.integers(1000, L - 5000, size=m))
        for j in range(m):
            a = int(tstarts[j]) if rng.random() < 0.7 else int(repeat_spots[chrom][int(rng.integers(0, len(spots)))]) + int(rng.integers(-40, 40))
            a = max(1, a)
            trf_lines.append((chrom, a, a + int(rng.integers(25, 600))))
    # TRF rows sorted by start inside a contig, as the UCSC table is
    trf_lines.sort(key=lambda t: (order.index(t[0]), t[1]))
    for extra in ("chr1_KI270706v1_random", "chrM"):
        rm_lines.append(",".join(["300", "1.0", "0.0", "0.0", extra, "10", "200", "(0)", "+", "L1PA3", "LINE/L1", "1", "191", "(0)", "1"]))
    seg_rows = []
    for _ in range(n_segdups):
        chrom = order[int(rng.integers(0, len(order)))]
        a = int(rng.integers(10000, GRCH38_LENGTHS[chrom] - 400000))
        seg_rows.append((chrom, a, a + int(rng.integers(1000, 300000))))
    seg_rows.append(("chrUn_KI270742v1", 100, 5000))
    files = {k: os.path.join(root, v) for k, v in dict(rm="repeatmasker.csv", trf="trf.csv", seg="segdups.tsv", fai="reference.fa.fai",
                                                        cyto="cytobands.txt", vcf="gnomad_sv.vcf").items()}
    with open(files["rm"], "w") as f:
        f.write("\n".join(rm_lines) + "\n")
    with open(files["trf"], "w") as f:
        f.write("".join(f"{c},{a},{b},trf,{int(rng.integers(1, 60))},{b - a}\n" for c, a, b in trf_lines))
    with open(files["seg"], "w") as f:
        f.write("#bin\tchrom\tchromStart\tchromEnd\tname\n")
        f.write("".join(f"{585 + i % 7}\t{c}\t{a}\t{b}\t{c}:{a}\n" for i, (c, a, b) in enumerate(seg_rows)))
    with open(files["fai"], "w") as f:
        for chrom in list(GRCH38_LENGTHS) + ["chrM", "chr1_KI270706v1_random"]:
            f.write(f"{chrom}\t{GRCH38_LENGTHS.get(chrom, 16569)}\t6\t60\t61\n")
    with open(files["cyto"], "w") as f:
        for chrom in GRCH38_LENGTHS:
            p, L = GRCH38_P_ARM_END[chrom], GRCH38_LENGTHS[chrom]
            f.write(f"{chrom}\t0\t{p // 2}\tp21.1\tgneg\n{chrom}\t{p // 2}\t{p}\tp11.1\tacen\n{chrom}\t{p}\t{(p + L) // 2}\tq11\tacen\n{chrom}\t{(p + L) // 2}\t{L}\tq21\tgpos50\n")
        f.write("chrM\t0\t16569\t\t\n")
    type_pool = ["DEL"] * 45 + ["DUP"] * 20 + ["INS"] * 25 + ["INV"] * 3 + ["MCNV"] * 2 + ["BND"] * 3 + ["CPX"] * 2
    evidence_pool = ["SR", "PE", "RD", "PE,SR", "BAF,RD", "PE,RD,SR", "RD,SR"]
    filter_pool = ["PASS"] * 8 + ["MULTIALLELIC", "UNRESOLVED", "PCRPLUS_ENRICHED", "PESR_GT_OVERDISPERSION;UNSTABLE_AF_PCRMINUS"]
    n_vcf = 0
    with open(files["vcf"], "w") as f:
        f.write("##fileformat=VCFv4.2\n##source=synthetic gnomAD-SV style records\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
        for chrom in order:
            L = GRCH38_LENGTHS[chrom]
            n = max(50, int(n_records * L / total))
            near = repeat_spots[chrom][rng.integers(0, len(repeat_spots[chrom]), size=n)] + rng.integers(-1500, 1500, size=n)
            free = rng.integers(1, L - 1, size=n)
            pos = np.sort(np.clip(np.where(rng.random(n) < 0.6, near, free), 1, L - 1))
            autosome = chrom not in ("chrX", "chrY")
            for j in range(n):
                t = type_pool[int(rng.integers(0, len(type_pool)))]
                if t == drop_type and autosome:
                    t = "DEL"
                u = rng.random()
                ln = int(rng.integers(50, 400)) if u < 0.5 else (int(rng.integers(400, 20000)) if u < 0.9 else int(rng.integers(20000, 3000000)))
                ln = min(ln, max_sv_length)
                if t == "DEL" and 5000 <= ln < 60000:
                    ln = (ln // 1000) * 1000   # the read-depth caller's 1 kb steps the smoothing is there for
                if rng.random() < 0.03:
                    ln = int(rng.integers(-1, 50))
                svlen = -ln if (t == "DEL" and rng.random() < 0.2) else ln
                afs = np.minimum(1.0, rng.beta(0.25, 6.0, size=6) + 1.0 / 21694.0)
                if not autosome and rng.random() < 0.2:
                    afs[0] = 1.3   # hemizygous calls: the reason the Java leaves X and Y out
                fmt = [repr(float(x)) if rng.random() < 0.5 else f"{x:.6g}" for x in afs]
                info = (f"END={int(pos[j]) + abs(ln)};SVTYPE={t};" + (f"SVLEN={svlen};" if rng.random() > 0.01 else "") +
                        f"CHR2={chrom[3:]};ALGORITHMS=manta;EVIDENCE={evidence_pool[int(rng.integers(0, len(evidence_pool)))]};PROTEIN_CODING__INTERGENIC;"
                        f"AN=21694;AC={int(afs[0] * 21694)};AF={fmt[0]};N_BI_GENOS=10847;AFR_AN=9534;AFR_AF={fmt[1]};AMR_AN=1952;AMR_AF={fmt[2]};"
                        f"EAS_AN=2416;EAS_AF={fmt[3]};EUR_AN=7624;EUR_AF={fmt[4]};OTH_AN=168;OTH_AF={fmt[5]};POPMAX_AF={max(afs):.6g}")
                f.write(f"{chrom[3:]}\t{int(pos[j])}\tgnomAD-SV_v2.1_{t}_{chrom[3:]}_{j}\tN\t<{t}>\t{int(rng.integers(1, 999))}\t{filter_pool[int(rng.integers(0, len(filter_pool)))]}\t{info}\n")
                n_vcf += 1
        f.write("GL000192.1\t100\tx\tN\t<DEL>\t1\tPASS\tEND=300;SVTYPE=DEL;SVLEN=200;EVIDENCE=SR;AF=0.1;\n")
    out_dir = os.path.join(root, "model_out")
    os.makedirs(out_dir, exist_ok=True)
    argv = [files["rm"], len(rm_lines), files["trf"], len(trf_lines), files["seg"], len(seg_rows), files["fai"], files["cyto"], files["vcf"], 1, out_dir, 0]
    return dict(argv=argv, files=files, out_dir=out_dir, n_vcf_records=n_vcf, chr1_length=GRCH38_LENGTHS["chr1"], n_repeatmasker=len(rm_lines), n_trf=len(trf_lines), n_segdups=len(seg_rows))


BUILD_MODEL_OUTPUTS = ("statistics_svType.txt", "statistics_intervalType.txt", "statistics_length.txt", "statistics_position.txt",
                       "repeatIntervals.txt", "segdups.txt") + tuple(f"statistics_alleleFrequency_{p}.txt" for p in AF_NAMES)
