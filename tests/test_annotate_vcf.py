"""SURVEY.md §8(f) rank 2 — AnnotateVCF (reference src/AnnotateVCF.java:96-510).

CPU half: the oracle restatement (oracle/annotate_vcf.cpp) against a brute-force evaluation of the definition (per-position
class arrays built with numpy straight from the loadPositions rules, src/SimulateHaplotypes.java:143-218) and against
hand-computed known answers of the Java string handling. GPU half: libsynthpop_cuda's sp_annotate_vcf must write the same
bytes as the oracle on the truth VCFs of row W and on caller-like records."""
import os
import subprocess

import numpy as np
import pytest

import oracle_helpers as oh  # noqa: F401  (path set-up)

HEADER = "##fileformat=VCFv4.2\n##contig=<ID=chr1,length=1>\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n"


def _inputs(tmp_path, L=200_000, seed=5):
    from synthpop import synth
    return synth.generate_population_inputs(str(tmp_path / "in"), L, n_insertions=300, seed=seed)


def _caller_like_records(rng, L, n=400):
    """Records the way callers write them: every branch of annotateVCF (:150-176) and getConfidenceInterval (:299-433)."""
    out = []
    for k in range(n):
        pos = int(rng.integers(1, L - 5000))
        ln = int(rng.integers(50, 4000))
        kind = k % 16
        info = []
        alt = "N"
        if kind == 0:
            info = [f"SVTYPE=DEL", f"END={pos + ln}", f"SVLEN=-{ln}"]
        elif kind == 1:
            info = ["SVTYPE=DEL", f"SVLEN=-{ln}"]                              # no END: end = pos+len-1
        elif kind == 2:
            info = ["SVTYPE=INS", f"SVLEN={ln}"]                              # insertion without END: end = pos
        elif kind == 3:
            info = ["SVTYPE=INS", f"END={pos + 1}", f"SVLEN={ln}"]
        elif kind == 4:
            info = ["SVTYPE=DUP:TANDEM", f"END={pos + ln}", f"CIPOS=-{rng.integers(0, 200)},{rng.integers(0, 200)}",
                    f"CIEND=-{rng.integers(0, 200)},{rng.integers(0, 200)}"]
        elif kind == 5:
            info = [f"CIEND=-10,10", "SVTYPE=INV", f"END={pos + ln}"]        # END after CIEND: the guard of getField (:213-216)
        elif kind == 6:
            info = ["SVTYPE=BND", f"END={pos + ln}"]                          # not handled: copied through
        elif kind == 7:
            alt = "<DEL>"; info = [f"END={pos + ln}", "IMPRECISE"]            # type from ALT (:151)
        elif kind == 8:
            alt = "<dup:tandem>"; info = [f"SVLEN={ln}", f"STD_quant_start={rng.uniform(0, 30):.3f}", f"STD_quant_stop={rng.uniform(0, 30):.3f}"]
        elif kind == 9:
            info = ["SVTYPE=del", f"END={pos + ln}", f"STD_POS1={rng.uniform(0, 9):.2f}", f"STD_POS2={rng.uniform(0, 9):.2f}"]
        elif kind == 10:
            info = ["SVTYPE=DEL", f"END={pos + ln}", f"CILEN={rng.integers(0, 50)},{rng.integers(0, 50)}"]
        elif kind == 11:
            info = ["SVTYPE=DEL"]                                             # neither END nor SVLEN (:173-175)
        elif kind == 12:
            info = ["SVTYPE=INV", f"END={pos + ln}", "CIPOS=.,5"]             # non-digit first value -> 0 (:317)
        elif kind == 13:
            info = ["PRECISE", "SVTYPE=DEL", f"END={max(1, pos - 20)}"]       # END before POS: negative length
        elif kind == 14:
            info = ["SVTYPE=DEL/INV", f"END={pos + ln}"]                      # known but unhandled type: no ALT fallback
        else:
            info = ["SVTYPE=DUP", f"END={pos + ln}", f"SVLEN={ln + 500}"]     # SVLEN != END-POS: warning only
        tail = "\tGT\t0/1" if k % 3 == 0 else ("\t\t" if k % 7 == 0 else "")  # trailing empty columns vanish (String.split)
        out.append(f"chr1\t{pos}\tcall_{k}\tN\t{alt}\t60\tPASS\t{';'.join(info)}{tail}")
    return out


def _brute_force_classes(inp, L):
    """positions[0..4] as boolean arrays, from the rules of loadPositions (src/SimulateHaplotypes.java:143-218)."""
    ref = np.frombuffer(open(inp.reference_file, "rb").read().split(b"\n", 1)[1].replace(b"\n", b""), np.uint8)
    assert ref.size == L
    raw = np.zeros((4, L), bool)
    seen = [False] * 4
    rows = open(inp.repeat_intervals_file).read().split("\n")
    for line in rows[1:1 + int(rows[0])]:
        t, s, e = (int(x) for x in line.split(","))
        raw[t, max(0, s - 25):min(L - 1, e + 25) + 1] = True; seen[t] = True
    rows = open(inp.segdups_file).read().split("\n")
    for line in rows[1:1 + int(rows[0])]:
        s, e = (int(x) for x in line.split(","))
        raw[3, max(0, s - 25):min(L - 1, e + 25) + 1] = True; seen[3] = True
    for t in range(4):
        if not seen[t]:
            raw[t, 0] = True   # the de-duplication loop leaves {0} for an empty class (:170-174)
    both = raw[2] | (raw[0] & raw[1])
    rep, sat = raw[0] & ~both, raw[1] & ~both
    seg = raw[3] & ~rep & ~sat & ~both
    none = ~(rep | sat | both | seg)
    ok = np.isin(ref, np.frombuffer(b"ACGTacgt", np.uint8))
    return [c & ok for c in (rep, sat, both, seg, none)]


def _expected_fields(classes, pos, end, length, ci_pos, ci_end, dmin, dmax):
    L = classes[0].size

    def any_in(c, a, b):
        a, b = max(a, 0), min(b, L - 1)
        return a <= b and bool(c[a:b + 1].any())

    def rtype(a, b):
        f = [any_in(c, a, b) for c in classes]
        return 2 if f[2] or (f[0] and f[1]) else 0 if f[0] else 1 if f[1] else 3 if f[3] else 4

    def clamp(a, b):
        a = max(a, pos - dmax); a = min(a, pos - dmin)
        b = min(b, pos + dmax); b = max(b, pos + dmin)
        return a, b
    start = rtype(*clamp(pos + ci_pos[0], pos + ci_pos[1]))
    if end is None:
        return start, 4, 0.0
    e = rtype(*clamp(end + ci_end[0], end + ci_end[1]))
    if length is None:
        return start, e, 0.0
    a, b = max(pos, 0), min(end, L - 1)
    surface = sum(int(c[a:b + 1].sum()) for c in classes[:4]) if a <= b else 0
    return start, e, max(surface, 1) / length


def _run_oracle(oracle_bin, files, inp, dmin, dmax, tmp_path):
    lst = tmp_path / "list.txt"
    lst.write_text("".join(f + "\n" for f in files))
    r = subprocess.run([oracle_bin, "AnnotateVCF", str(lst), str(dmin), str(dmax), inp.reference_file, inp.repeat_intervals_file,
                        inp.segdups_file], capture_output=True)
    return r


def test_oracle_matches_brute_force_definition(oracle_bin, tmp_path):
    L = 120_000
    inp = _inputs(tmp_path, L)
    classes = _brute_force_classes(inp, L)
    rng = np.random.default_rng(11)
    cases = []
    for k in range(300):
        pos = int(rng.integers(30, L - 3000)); ln = int(rng.integers(1, 2500))
        cp = (-int(rng.integers(0, 150)), int(rng.integers(0, 150))); ce = (-int(rng.integers(0, 150)), int(rng.integers(0, 150)))
        cases.append((pos, pos + ln - 1, ln, cp, ce))
    vcf = tmp_path / "calls.vcf"
    with open(vcf, "w") as f:
        f.write(HEADER)
        for i, (pos, end, ln, cp, ce) in enumerate(cases):
            f.write(f"chr1\t{pos}\tc{i}\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END={end + 1};CIPOS={cp[0]},{cp[1]};CIEND={ce[0]},{ce[1]}\n")
    assert _run_oracle(oracle_bin, [str(vcf)], inp, 10, 100, tmp_path).returncode == 0
    got = [l for l in open(str(vcf).replace(".vcf", "_annotated.vcf")).read().split("\n") if l and l[0] != "#"]
    assert len(got) == len(cases)
    seen_types = set()
    for line, (pos, end, ln, cp, ce) in zip(got, cases):
        s, e, frac = _expected_fields(classes, pos, end, ln, cp, ce, 10, 100)
        seen_types.update((s, e))
        want = f";REPEATS_START={s};REPEATS_END={e};REPEATS_FRACTION={frac:,.4f}"
        assert line.endswith(want), (line, want)
    assert len(seen_types) >= 3


def test_oracle_known_answers_of_the_string_handling(oracle_bin, tmp_path):
    inp = _inputs(tmp_path, 60_000)
    vcf = tmp_path / "k.vcf"
    recs = ["chr1\t1000\ta\tN\t<INS>\t.\tPASS\tSVLEN=300\t\t",            # type from ALT, trailing empty columns dropped
            "chr1\t1000\tb\tN\tN\t.\tPASS\tSVTYPE=BND;END=2000",           # not handled: unchanged
            "chr1\t1000\tc\tN\tN\t.\tPASS\tCIEND=-5,5;SVTYPE=DEL",          # END= only inside CIEND= -> no END, no SVLEN
            "chr1\t1000\td\tN\tN\t.\tPASS\tSVTYPE=TRA"]
    vcf.write_text(HEADER + "\n".join(recs) + "\n")
    assert _run_oracle(oracle_bin, [str(vcf)], inp, 10, 100, tmp_path).returncode == 0
    out = open(str(vcf).replace(".vcf", "_annotated.vcf")).read().split("\n")
    body = [l for l in out if l and l[0] != "#"]
    assert out[:3] == HEADER.split("\n")[:3]
    assert body[0].startswith("chr1\t1000\ta\tN\t<INS>\t.\tPASS\tSVLEN=300;REPEATS_START=") and body[0].endswith(";REPEATS_FRACTION=0.0000")
    assert body[1] == recs[1] and body[3] == recs[3]
    assert body[2].startswith(recs[2] + ";REPEATS_START=") and body[2].endswith(";REPEATS_END=4;REPEATS_FRACTION=0.0000")
    # uncaught Java exceptions: an empty line (charAt(0)), a record with fewer than 8 columns, a non-numeric POS
    for bad in ("\n", "chr1\t5\tx\n", "chr1\tabc\tx\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END=9\n"):
        vcf.write_text(HEADER + bad)
        r = _run_oracle(oracle_bin, [str(vcf)], inp, 10, 100, tmp_path)
        assert r.returncode == 1 and b"Exception in thread" in r.stderr
    # a listed path without ".vcf": substring(0,-1) throws (:119)
    other = tmp_path / "calls.txt"; other.write_text(HEADER)
    assert _run_oracle(oracle_bin, [str(other)], inp, 10, 100, tmp_path).returncode == 1


@pytest.mark.gpu
def test_gpu_annotation_is_byte_identical_to_the_oracle(sp, oracle_bin, tmp_path):
    L = 1_500_000
    inp = _inputs(tmp_path, L, seed=9)
    truth = tmp_path / "truth"; truth.mkdir()
    subprocess.run([oracle_bin, "--quiet", "--seed", "42", "SimulateHaplotypes"] + inp.simulate_argv(str(truth), 4, n_frequent=300), check=True)
    rng = np.random.default_rng(3)
    calls = tmp_path / "calls.vcf"
    calls.write_text(HEADER + "\n".join(_caller_like_records(rng, L, 800)) + "\n")
    files = [str(truth / "groundTruth_joint.vcf"), str(truth / "groundTruth_individual_0.vcf"), str(calls)]
    want_dir = tmp_path / "want"; got_dir = tmp_path / "got"
    outs = {}
    for d in (want_dir, got_dir):
        d.mkdir()
        outs[d] = []
        for f in files:
            dst = d / os.path.basename(f)
            dst.write_bytes(open(f, "rb").read())
            outs[d].append(str(dst))
    for dmin, dmax in ((10, 100), (0, 0), (50, 20)):
        assert _run_oracle(oracle_bin, outs[want_dir], inp, dmin, dmax, tmp_path).returncode == 0
        lst = tmp_path / "got_list.txt"
        lst.write_text("".join(f + "\n" for f in outs[got_dir]))
        with sp.Context(0, 42) as ctx:
            ctx.annotate_vcf([str(lst), dmin, dmax, inp.reference_file, inp.repeat_intervals_file, inp.segdups_file])
            tm = ctx.annotate_timing()
            assert tm["records"] > 500 and ctx.kernel_launches >= 3
        for w, g in zip(outs[want_dir], outs[got_dir]):
            a = open(w.replace(".vcf", "_annotated.vcf"), "rb").read()
            b = open(g.replace(".vcf", "_annotated.vcf"), "rb").read()
            assert a == b, (w, dmin, dmax)
            assert a.count(b"REPEATS_FRACTION=") > 100
    # error behaviour: the Java's uncaught exceptions come back as SP_ERR_JAVA_EXCEPTION
    bad = tmp_path / "bad.vcf"; bad.write_text(HEADER + "chr1\t5\tx\n")
    with sp.Context(0, 42) as ctx:
        ctx.annotate_load(inp.reference_file, inp.repeat_intervals_file, inp.segdups_file)
        with pytest.raises(sp.SynthpopError, match="ArrayIndexOutOfBounds"):
            ctx.annotate_file(str(bad), str(tmp_path / "bad_out.vcf"), 10, 100)
