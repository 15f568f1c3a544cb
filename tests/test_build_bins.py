"""SURVEY.md §8(f) rank 3 — BuildReadLengthBins + Fastq2LengthHistogram (reference src/BuildReadLengthBins.java,
src/Fastq2LengthHistogram.java). CPU half: the parallel form of the Fisher-Yates the CUDA path uses (swap forest) against the
sequential loop, and the oracle's files against a pure-Python evaluation with a from-the-spec java.util.Random. GPU half:
byte-identical .bin / .histogram / .stats files against the oracle, then row D (SampleReadsFromLengthBins) fed from them."""
import bisect
import os
import subprocess

import numpy as np
import pytest

import oracle_helpers as oh  # noqa: F401


class JavaRandom:   # java.util.Random from its published algorithm
    def __init__(self, seed):
        self.s = (seed ^ 0x5DEECE66D) & ((1 << 48) - 1)
        self.steps = 0

    def next(self, bits):
        self.s = (self.s * 0x5DEECE66D + 0xB) & ((1 << 48) - 1)
        self.steps += 1
        v = self.s >> (48 - bits)
        return v - (1 << bits) if bits == 32 and v >= (1 << 31) else v

    def next_int(self, bound):
        r = self.next(31)
        m = bound - 1
        if bound & m == 0:
            return (bound * r) >> 31
        u = r
        while True:
            r = u % bound
            if u - r + m < (1 << 31):   # no int overflow
                return r
            u = self.next(31)


def _fastq_files(tmp_path, rng, crlf=False):
    files = []
    def record(k, ln):
        seq = "".join("ACGT"[int(x)] for x in rng.integers(0, 4, size=ln))
        qual = "".join(chr(33 + int(x)) for x in rng.integers(0, 40, size=ln))
        return [f"@read_{k}/ccs", seq, "+", qual]
    k = 0
    # file 0: many tiny reads -> one bin of ~180 000 records (n^2 / 2^32 ~ 7.5 expected redraws of nextInt)
    lines = []
    for _ in range(180_000):
        lines += record(k, int(rng.integers(1, 12))); k += 1
    files.append(("tiny.fastq", "\n".join(lines) + "\n"))
    # file 1: a bimodal spread, last line without terminator, plus reads beyond MAX_READ_LENGTH
    lines = []
    for _ in range(9000):
        ln = int(abs(rng.normal(1750, 350))) if rng.random() < 0.45 else int(abs(rng.normal(3750, 350)))
        lines += record(k, max(1, ln)); k += 1
    for _ in range(5):
        lines += record(k, 9000); k += 1   # beyond MAX_READ_LENGTH: the last bin takes them
    files.append(("spread.fastq", "\n".join(lines)))
    # file 2: CRLF line ends (readLine drops the '\r'), two records in a bin of their own
    lines = []
    for ln in (5100, 5150, 5300):
        lines += record(k, ln); k += 1
    files.append(("crlf.fastq", ("\r\n" if crlf else "\n").join(lines) + ("\r\n" if crlf else "\n")))
    paths = []
    for name, text in files:
        p = tmp_path / name
        p.write_bytes(text.encode())
        paths.append(str(p))
    lst = tmp_path / "list.txt"
    lst.write_text("".join(p + "\n" for p in paths))
    return str(lst), paths


def test_swap_forest_equals_sequential_fisher_yates():
    """final[i] = C(next(j_i, i)) with C resolved by pointer jumping (csrc/length_bin_builder.cu) against the loop of
    BuildReadLengthBins.shuffle (:108-116)."""
    rng = np.random.default_rng(2)
    for trial in range(400):
        n = int(rng.integers(1, 80)) if trial < 390 else int(rng.integers(2000, 9000))
        J = [0] + [int(rng.integers(0, i + 1)) for i in range(1, n)]
        a = list(range(n))
        for i in range(n - 1, 0, -1):
            a[i], a[J[i]] = a[J[i]], a[i]
        pairs = sorted((J[y], y) for y in range(1, n))

        def nxt_writer(p, t):
            k = bisect.bisect_left(pairs, (p, t + 1))
            return pairs[k][1] if k < len(pairs) and pairs[k][0] == p else None
        nxt = [x if nxt_writer(x, x) is None else nxt_writer(x, x) for x in range(n)]
        while True:
            new = [nxt[nxt[x]] for x in range(n)]
            if new == nxt:
                break
            nxt = new
        final = []
        for i in range(n):
            p = J[i] if i > 0 else 0
            q = nxt_writer(p, i)
            final.append(p if q is None else nxt[q])
        assert final == a


def test_oracle_against_python_evaluation(oracle_bin, tmp_path):
    rng = np.random.default_rng(4)
    lst, paths = _fastq_files(tmp_path, rng, crlf=True)
    prefix = str(tmp_path / "bins_")
    seed, bl, mx = 17, 500, 6000
    subprocess.run([oracle_bin, "--seed", str(seed), "BuildReadLengthBins", lst, str(bl), str(mx), "3100000", prefix], check=True)
    n_bins = (mx + bl - 1) // bl
    recs = []
    for p in paths:
        rows = open(p, "rb").read().decode().replace("\r\n", "\n").split("\n")
        if rows[-1] == "":
            rows.pop()
        recs += [rows[i:i + 4] for i in range(0, len(rows), 4)]
    bins = [[] for _ in range(n_bins)]
    for r in recs:
        bins[min(len(r[1]) // bl, n_bins - 1)].append(r)
    jr = JavaRandom(seed)
    for b in range(n_bins):
        a = bins[b]
        if len(a) >= 3:
            for i in range(len(a) - 1, 0, -1):
                j = jr.next_int(i + 1)
                a[i], a[j] = a[j], a[i]
        want = "".join("\n".join(r) + "\n" for r in a)
        assert open(f"{prefix}{b}.bin").read() == want, b
    assert jr.steps > sum(len(a) - 1 for a in bins if len(a) >= 3)   # at least one redraw happened: the fixpoint path is exercised
    assert open(prefix + ".histogram").read() == "".join(f"{b * bl + bl // 2},{len(bins[b])}\n" for b in range(n_bins))
    stats = open(prefix + ".stats").read().split("\n")
    assert stats[0] == "2" and len(stats) == 12 and stats[1] == "1750" and stats[3] == "3750"
    assert all("." in stats[k] for k in (2, 4, 7, 8, 9, 10))   # the Double.toString lines


@pytest.mark.gpu
@pytest.mark.parametrize("crlf", [False, True])
def test_gpu_bins_are_byte_identical_to_the_oracle(sp, oracle_bin, tmp_path, crlf):
    from synthpop import synth
    rng = np.random.default_rng(9)
    lst, paths = _fastq_files(tmp_path, rng, crlf=crlf)
    want, got = str(tmp_path / "want_"), str(tmp_path / "got_")
    seed, bl, mx = 23, 500, 6000
    n_bins = (mx + bl - 1) // bl
    subprocess.run([oracle_bin, "--seed", str(seed), "BuildReadLengthBins", lst, str(bl), str(mx), "3100000", want], check=True)
    with sp.Context(0, seed) as ctx:
        ctx.build_read_length_bins([lst, bl, mx, 3100000, got])
        tm = ctx.build_bins_timing()
        assert tm["records"] == 189_008 and tm["bytes"] == sum(os.path.getsize(f"{want}{b}.bin") for b in range(n_bins))
        for name in [f"{b}.bin" for b in range(n_bins)] + [".histogram", ".stats"]:
            assert open(got + name, "rb").read() == open(want + name, "rb").read(), name
        # Fastq2LengthHistogram on one of the inputs
        for d, tool in ((want, None), (got, ctx)):
            argv = [paths[1], 100, 5000, d + "f2l.hist", d + "f2l.max"]
            if tool is None:
                subprocess.run([oracle_bin, "Fastq2LengthHistogram"] + [str(x) for x in argv], check=True)
            else:
                tool.fastq2length_histogram(argv)
        for name in ("f2l.hist", "f2l.max"):
            assert open(got + name, "rb").read() == open(want + name, "rb").read(), name
        # row D fed by these bins (src/SampleReadsFromLengthBins.java:19-81): product against oracle, same seed
        cum = synth.cumulative_bimodal(n_bins, bl, 1750, 350, 3750, 350, 0.45)
        argv = ["1750", "350", "3750", "350", "0.45", str(bl), str(mx), got, "200000", "3.0", str(tmp_path / "sample_got.fastq")]
        ctx.sample_reads_from_length_bins(argv, cum)
    cdf = tmp_path / "cdf.txt"
    cdf.write_text("".join(repr(float(v)) + "\n" for v in cum))
    argv_w = argv[:7] + [want] + argv[8:10] + [str(tmp_path / "sample_want.fastq")]
    subprocess.run([oracle_bin, "--seed", str(seed), "--cdf", str(cdf), "SampleReadsFromLengthBins"] + argv_w, check=True)
    assert open(tmp_path / "sample_got.fastq", "rb").read() == open(tmp_path / "sample_want.fastq", "rb").read()
    # a file that ends inside a record: the Java dereferences a null line
    bad = tmp_path / "bad.fastq"; bad.write_text("@r\nACGT\n+\n")
    lst2 = tmp_path / "list2.txt"; lst2.write_text(str(bad) + "\n")
    assert subprocess.run([oracle_bin, "BuildReadLengthBins", str(lst2), "500", "5000", "1000", want + "bad_"], capture_output=True).returncode == 1
    with sp.Context(0, seed) as ctx:
        with pytest.raises(sp.SynthpopError, match="NullPointerException"):
            ctx.build_read_length_bins([str(lst2), 500, 5000, 1000, got + "bad_"])
