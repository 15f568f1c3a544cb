"""Row R1 (annotateVariantsWithRepeats, src/SimulateHaplotypes.java:412-571) against its DEFINITION, evaluated here with
numpy from the input tables alone: the repetitive context of a position is decided by the repeat intervals within
MAX_DISTANCE_TO_REPEAT_INTERVAL (:88) of it (one type, or 2 when they differ, :83; else 3 when a segmental duplication is
that close, else 4), and REPEATS_FRACTION is the length of the union of the repeat and segdup intervals inside the variant.
The restatement (oracle/) and the product (csrc/population_sample.cu) transliterate the same Java and so could share a
misreading; this check shares nothing with either. Where the Java itself departs from the definition the outcome is
history-dependent and is left to the byte parity of the two implementations: a variant that meets no interval reads the
stale tmpPairs[0] of an earlier call (:513), one that meets more than 100 goes through the grow path that clobbers its own
loop variable (:544-566) — test_dense_repeats_take_the_grow_path builds that case."""
import os
import subprocess

import numpy as np
import pytest

MAX_DISTANCE = 100   # src/SimulateHaplotypes.java:88


def _variants(path):
    rows = open(path, "rb").read().split(b"\n")
    n = int(rows[0]) + 1
    out = []
    for i in range(n):
        r = rows[1 + 10 * i: 11 + 10 * i]
        out.append(dict(svType=int(r[0]), length=int(r[2]), position=int(r[4]), it_start=int(r[7]), it_end=int(r[8]), surface=int(r[9])))
    return out


def _context(pos, rep, seg):
    near = rep[(rep[:, 2] + MAX_DISTANCE >= pos) & (rep[:, 1] <= pos + MAX_DISTANCE)]
    if near.shape[0]:
        kinds = set(int(t) for t in near[:, 0])
        return kinds.pop() if len(kinds) == 1 else 2
    return 3 if np.any((seg[:, 1] + MAX_DISTANCE >= pos) & (seg[:, 0] <= pos + MAX_DISTANCE)) else 4


def _surface(first, last, rep, seg):
    """(number of intervals met, bases of [first, last] they cover)"""
    a = rep[(rep[:, 2] >= first) & (rep[:, 1] <= last)][:, 1:3]
    b = seg[(seg[:, 1] >= first) & (seg[:, 0] <= last)]
    iv = np.concatenate([a, b]) if a.size or b.size else np.zeros((0, 2), np.int64)
    if iv.shape[0] == 0:
        return 0, 0
    covered = np.zeros(last - first + 2, np.int32)
    np.add.at(covered, np.maximum(iv[:, 0], first) - first, 1)
    np.add.at(covered, np.minimum(iv[:, 1], last) - first + 1, -1)
    return iv.shape[0], int(np.count_nonzero(np.cumsum(covered[:-1]) > 0))


def _check_against_definition(variants_file, rep, seg, min_checked):
    rep = rep.astype(np.int64); seg = seg.astype(np.int64)
    checked = skipped = 0
    for v in _variants(variants_file):
        assert v["it_start"] == _context(v["position"], rep, seg), v
        if v["svType"] == 2:   # INS: one position (:422-425)
            assert v["it_end"] == v["it_start"] and v["surface"] == 0, v
            continue
        last = v["position"] + v["length"] - 1
        assert v["it_end"] == _context(last, rep, seg), v
        n_met, covered = _surface(v["position"], last, rep, seg)
        if n_met == 0 or n_met > 100:
            skipped += 1   # stale tmpPairs[0] / grow path: not a function of this variant alone
            continue
        assert v["surface"] == covered, (v, n_met)
        checked += 1
    assert checked >= min_checked, (checked, skipped)


@pytest.mark.parametrize("L,seed", [(3_000_000, 42), (1_200_000, 7)])
def test_restatement_annotates_as_defined(oracle_bin, tmp_path, L, seed):
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), L, n_insertions=2000, seed=seed)
    out = str(tmp_path / "out"); os.makedirs(out)
    subprocess.run([oracle_bin, "--quiet", "--seed", str(seed), "SimulateHaplotypes"] + inp.simulate_argv(out, 4, n_frequent=400), check=True)
    _check_against_definition(os.path.join(out, "variants.txt"), inp.extra["repeat_intervals"], inp.extra["segdups"], 150)


@pytest.mark.gpu
@pytest.mark.parametrize("L,seed", [(3_000_000, 42), (1_200_000, 7)])
def test_product_annotates_as_defined(sp, tmp_path, L, seed):
    from synthpop import synth
    inp = synth.generate_population_inputs(str(tmp_path / "in"), L, n_insertions=2000, seed=seed)
    out = str(tmp_path / "out"); os.makedirs(out)
    with sp.Context(0, seed) as ctx:
        ctx.simulate_haplotypes(inp.simulate_argv(out, 4, n_frequent=400))
    _check_against_definition(os.path.join(out, "variants.txt"), inp.extra["repeat_intervals"], inp.extra["segdups"], 150)


def _dense_inputs(root, L, seed):
    """Repeat intervals of 20 bases every 40 over 40 % of the chromosome: a variant of more than 4 kbp there meets more than
    100 of them."""
    from synthpop import synth
    inp = synth.generate_population_inputs(root, L, n_insertions=2000, seed=seed)
    lo, hi = L // 10, L // 2   # dense in this stretch only: every interval class keeps positions to sample from
    starts = np.arange(lo, hi, 40, dtype=np.int64)
    dense = np.stack([(starts // 40) % 3, starts, starts + 19], axis=1)
    old = inp.extra["repeat_intervals"].astype(np.int64)
    old = old[(old[:, 2] < lo - 200) | (old[:, 1] > hi + 200)]
    rep = np.concatenate([old, dense])
    rep = rep[np.argsort(rep[:, 1], kind="stable")]
    with open(inp.repeat_intervals_file, "w") as f:
        f.write(f"{rep.shape[0]}\n")
        np.savetxt(f, rep, fmt="%d", delimiter=",")
    return inp, rep


def test_dense_repeats_restatement_departs_from_the_definition_only_beyond_100(oracle_bin, tmp_path):
    """With dense intervals the definition still holds for every variant that meets at most 100 of them."""
    inp, rep = _dense_inputs(str(tmp_path / "in"), 2_000_000, 11)
    out = str(tmp_path / "out"); os.makedirs(out)
    r = subprocess.run([oracle_bin, "--quiet", "--seed", "11", "SimulateHaplotypes"] + inp.simulate_argv(out, 3, n_frequent=300),
                       capture_output=True, text=True)
    if r.returncode != 0:   # the grow path may walk out of the interval table: a Java exception, same in both implementations
        assert "Exception" in (r.stderr + r.stdout)
        return
    _check_against_definition(os.path.join(out, "variants.txt"), rep, inp.extra["segdups"], 20)


@pytest.mark.gpu
def test_dense_repeats_take_the_grow_path(sp, oracle_bin, tmp_path):
    """More than 100 intersecting intervals: tmpPairs doubles and the Java assigns the new capacity to its loop variable.
    Whatever that does — rows skipped, rows repeated, an ArrayIndexOutOfBoundsException — both implementations do it."""
    from test_population_gpu import _cmp_dirs
    inp, rep = _dense_inputs(str(tmp_path / "in"), 2_000_000, 11)
    want = str(tmp_path / "want"); got = str(tmp_path / "got")
    os.makedirs(want); os.makedirs(got)
    r = subprocess.run([oracle_bin, "--quiet", "--seed", "11", "SimulateHaplotypes"] + inp.simulate_argv(want, 3, n_frequent=300),
                       capture_output=True, text=True)
    seg = inp.extra["segdups"].astype(np.int64)
    n_beyond = sum(1 for v in (_variants(os.path.join(want, "variants.txt")) if r.returncode == 0 else [])
                   if v["svType"] != 2 and _surface(v["position"], v["position"] + v["length"] - 1, rep, seg)[0] > 100)
    with sp.Context(0, 11) as ctx:
        if r.returncode != 0:
            with pytest.raises(RuntimeError, match="Exception"):
                ctx.simulate_haplotypes(inp.simulate_argv(got, 3, n_frequent=300))
            return
        ctx.simulate_haplotypes(inp.simulate_argv(got, 3, n_frequent=300))
    assert n_beyond >= 5, "the inputs no longer reach the grow path"
    _cmp_dirs(want, got, 3)
