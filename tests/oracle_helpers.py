"""Access to the CPU oracle (oracle/) from tests. Never imported by product code."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_lib = None


def ensure_oracle_built():
    so = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
    exe = os.path.join(ORACLE_DIR, "_build", "synthpop_oracle")
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".h"))]
    newest = max(os.path.getmtime(s) for s in srcs)
    if not (os.path.exists(so) and os.path.exists(exe)) or min(os.path.getmtime(so), os.path.getmtime(exe)) < newest:
        subprocess.run(["make", "-C", ORACLE_DIR, "-s"], check=True, stdout=subprocess.DEVNULL)
    return so


def load_oracle():
    global _lib
    if _lib is None:
        _lib = C.CDLL(ensure_oracle_built())
        _lib.oracle_build_haplotype.restype = C.c_int64
        _lib.oracle_java_random_state_after.restype = C.c_uint64
        _lib.oracle_java_random_state_after.argtypes = [C.c_int64, C.c_uint64]
    return _lib


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def build_haplotype(lib, ref, sv_type, position, length, payload_off, payload_blob, has_payload, ids):
    """oracle restatement of the buildHaplotypes loop body for one edit list; returns bytes or raises."""
    ref = np.ascontiguousarray(ref, dtype=np.uint8)
    sv_type = np.ascontiguousarray(sv_type, dtype=np.int8)
    position = np.ascontiguousarray(position, dtype=np.int32)
    length = np.ascontiguousarray(length, dtype=np.int32)
    payload_off = np.ascontiguousarray(payload_off, dtype=np.int64)
    payload_blob = np.ascontiguousarray(payload_blob, dtype=np.uint8)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    hp = None if has_payload is None else np.ascontiguousarray(has_payload, dtype=np.int8)
    cap = int(ref.size + 2 * np.abs(length.astype(np.int64)).sum() + payload_blob.size + 4 * sv_type.size + 64)
    out = np.empty(cap, dtype=np.uint8)
    err = C.create_string_buffer(512)
    n = lib.oracle_build_haplotype(_vp(ref), C.c_int(ref.size), C.c_int(sv_type.size), _vp(sv_type), _vp(position),
                                   _vp(length), _vp(payload_off), _vp(payload_blob), None if hp is None else _vp(hp),
                                   _vp(ids), C.c_int(ids.size), _vp(out), C.c_int64(cap), err, C.c_int(512))
    if n == -1:
        raise RuntimeError(err.value.decode())
    if n < 0:
        raise RuntimeError("oracle output buffer too small")
    return out[:n].copy()


def random_edit_lists(rng, L, n_variants, n_hap, p_carry=0.5, max_len=2000, ins_frac=0.45):
    """Globally non-overlapping variants (so any subset is a valid per-haplotype edit list) + payloads."""
    ref = np.frombuffer(b"ACGTacgtN", dtype=np.uint8)[rng.integers(0, 9, size=L)].copy()
    gaps = rng.integers(0, max(2, (L // max(n_variants, 1))), size=n_variants)
    types, pos, length = [], [], []
    p = int(rng.integers(0, 20))
    for g in gaps:
        p += int(g)
        t = int(rng.choice([0, 1, 2, 3], p=[(1 - ins_frac) / 2, (1 - ins_frac) / 4, ins_frac, (1 - ins_frac) / 4]))
        ln = int(rng.integers(1, max_len))
        if t != 2 and p + ln > L:
            break
        if p > L:
            break
        types.append(t); pos.append(p); length.append(ln)
        p += 0 if t == 2 else ln
    n = len(types)
    types = np.array(types, dtype=np.int8); pos = np.array(pos, dtype=np.int32); length = np.array(length, dtype=np.int32)
    plen = np.where(types == 2, length, 0).astype(np.int64)
    payload_off = np.concatenate(([0], np.cumsum(plen))).astype(np.int64)
    payload = np.frombuffer(b"ACGTacgt", dtype=np.uint8)[rng.integers(0, 8, size=max(int(plen.sum()), 1))].copy()
    lists = [np.nonzero(rng.random(n) < p_carry)[0].astype(np.int32) for _ in range(n_hap)]
    list_off = np.concatenate(([0], np.cumsum([l.size for l in lists]))).astype(np.int64)
    ids = np.concatenate(lists).astype(np.int32) if lists else np.zeros(0, np.int32)
    return dict(ref=ref, sv_type=types, position=pos, length=length, payload_off=payload_off, payload=payload,
                lists=lists, list_off=list_off, ids=ids)


# ---- row P: the pbsim2 restatement ------------------------------------------------------------------------------
def glibc_rand(lib, seed: int, n: int) -> np.ndarray:
    """First n values of rand() after srand(seed) (restated glibc TYPE_3 generator, pinned by test_oracle_kat)."""
    assert 0 <= n < 2**31 - 400, "draw log too long for a test"
    out = np.empty(max(n, 1), dtype=np.int32)
    lib.oracle_glibc_rand(C.c_uint(seed), C.c_int(n), _vp(out))
    return out[:n]


def pbsim_simulate(lib, seq: np.ndarray, hmm_model: str, depth, len_mean, len_sd, len_min, len_max, accuracy_mean_cli,
                   ratio=(6, 54, 40), seed=0, id_prefix="S", no_error=False, per_read_streams=False, ref_num=1):
    """oracle/pbsim_port.cpp on one in-memory sequence. accuracy_mean_cli is the command-line value (int(x*100)*0.01 is
    applied here as set_sim_param does). Returns (file bytes, dict of per-read log arrays)."""
    seq = np.ascontiguousarray(seq, dtype=np.uint8)
    acc = int(accuracy_mean_cli * 100) * 0.01
    cap = int(seq.size * depth * 1.3) + (1 << 20)
    log_cap = int(seq.size * depth / max(1, len_min)) + 1024
    out = np.empty(cap, dtype=np.uint8)
    off = np.zeros(log_cap, np.uint64); ln = np.zeros(log_cap, np.int64); dr = np.zeros(log_cap, np.uint8)
    us = np.zeros(log_cap, np.uint64)
    n_reads = C.c_int64(); total = C.c_uint64()
    err = C.create_string_buffer(512)
    lib.oracle_pbsim_simulate.restype = C.c_int64
    lib.oracle_pbsim_simulate.argtypes = [C.c_char_p, C.c_double, C.c_double, C.c_double, C.c_long, C.c_long, C.c_double,
                                          C.c_long, C.c_long, C.c_long, C.c_uint, C.c_char_p, C.c_int, C.c_int, C.c_void_p,
                                          C.c_int64, C.c_long, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_uint64), C.c_char_p, C.c_int]
    n = lib.oracle_pbsim_simulate(hmm_model.encode(), depth, len_mean, len_sd, len_min, len_max, acc, ratio[0], ratio[1],
                                  ratio[2], seed, id_prefix.encode(), int(no_error), int(per_read_streams), _vp(seq), seq.size,
                                  ref_num, _vp(out), cap, C.byref(n_reads), _vp(off), _vp(ln), _vp(dr), _vp(us), log_cap,
                                  C.byref(total), err, 512)
    if n == -1:
        raise RuntimeError(err.value.decode())
    if n < 0:
        raise RuntimeError("oracle pbsim buffers too small")
    R = n_reads.value
    return out[:n].copy(), dict(n_reads=R, draw_offset=off[:R].astype(np.int64), length=ln[:R].copy(), dropped=dr[:R].copy(),
                                draws_used=us[:R].astype(np.int64), total_draws=int(total.value))
