"""TEST INFRASTRUCTURE ONLY (CPU oracle; never imported by the product).

Restatement of the tail of the reference's docker/haplotype2reads.sh:59-63 and of the 1x chunking of
docker/reads2svs.sh:65-74,98-128:

    cat reads_ID2.fa >> reads_ID1.fa ; shuf ; tr 'Z' '\\n'          -> one two-line FASTA record per read
    N_ROWS_1X = N_ROWS / MAX_COVERAGE (rounded up to a multiple of 4) ; split -d -l ; coverage c = first c chunks

Parity: `cat`, `tr` and `split` are pinned against the coreutils binaries themselves (tests/test_read_mix.py runs
them). `shuf` is unseeded in the reference (docker/haplotype2reads.sh:61) and therefore PARITY UNPINNED by
construction: the permutation is this repository's definition — line k of the concatenated file gets the key "state of
java.util.Random(seed) after k+1 steps" (s' = s*0x5DEECE66D + 0xB mod 2^48, s0 = (seed ^ 0x5DEECE66D) mod 2^48) and the
lines are emitted in ascending key order — checked as a permutation (same multiset of lines as `cat | tr`).
"""
import numpy as np

A = 0x5DEECE66D
C = 0xB
MASK = (1 << 48) - 1


def shuffle_keys(seed: int, n: int) -> np.ndarray:
    s = (seed ^ A) & MASK
    out = np.empty(n, dtype=np.uint64)
    for k in range(n):
        s = (s * A + C) & MASK
        out[k] = s
    return out


def mix_reads(files, seed: int) -> bytes:
    """files: the pbsim outputs (bytes, lines '>S1_<n>Z<seq>') of ID1 and ID2, in that order."""
    lines = []
    for f in files:
        lines += [l for l in bytes(f).split(b"\n") if l]
    order = np.argsort(shuffle_keys(seed, len(lines)), kind="stable")
    return b"".join(lines[i].replace(b"Z", b"\n") + b"\n" for i in order)


def rows_1x(n_rows: int, max_coverage: int) -> int:
    r = n_rows // max_coverage
    if r % 4 != 0:
        r = (r // 4 + 1) * 4
    return r


def coverage_prefix(mixed: bytes, max_coverage: int, coverage: int) -> bytes:
    rows = mixed.split(b"\n")[:-1]
    k = min(len(rows), coverage * rows_1x(len(rows), max_coverage))
    return b"".join(r + b"\n" for r in rows[:k])
