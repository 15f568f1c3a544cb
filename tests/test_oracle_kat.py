"""Pins the oracle's JDK / glibc primitives to known-answer vectors. CPU only."""
import ctypes as C

import numpy as np


def draws(lib, seed, kind, n, bound=0):
    out = np.zeros(n, dtype=np.int64)
    st = C.c_uint64()
    lib.oracle_java_random_draws(C.c_int64(seed), kind, bound, n, out.ctypes.data_as(C.c_void_p), C.byref(st))
    return out, st.value


def test_java_random_nextint(oracle_lib):
    assert list(draws(oracle_lib, 0, 0, 3)[0]) == [-1155484576, -723955400, 1033096058]
    assert list(draws(oracle_lib, 42, 0, 3)[0]) == [-1170105035, 234785527, -1360544799]


def test_java_random_nextdouble(oracle_lib):
    d0 = draws(oracle_lib, 0, 1, 3)[0] * 2.0 ** -53
    assert list(d0) == [0.730967787376657, 0.24053641567148587, 0.6374174253501083]
    d42 = draws(oracle_lib, 42, 1, 3)[0] * 2.0 ** -53
    assert list(d42) == [0.7275636800328681, 0.6832234717598454, 0.30871945533265976]


def test_java_random_bounded(oracle_lib):
    assert list(draws(oracle_lib, 0, 2, 8, 50)[0]) == [10, 48, 29, 47, 15, 3, 41, 11]
    assert list(draws(oracle_lib, 0, 2, 5, 100)[0]) == [60, 48, 29, 47, 15]
    assert list(draws(oracle_lib, 42, 2, 8, 50)[0]) == [30, 13, 48, 34, 20, 25, 5, 18]
    assert list(draws(oracle_lib, 42, 2, 5, 100)[0]) == [30, 63, 48, 84, 70]
    # power-of-two bound takes the multiply-shift shortcut
    assert list(draws(oracle_lib, 42, 2, 4, 16)[0]) == [int((16 * ((v & 0xFFFFFFFF) >> 1)) >> 31) for v in draws(oracle_lib, 42, 0, 4)[0]]


def test_java_random_boolean(oracle_lib):
    assert list(draws(oracle_lib, 0, 3, 8)[0]) == [1, 1, 0, 1, 1, 0, 1, 0]


def test_java_random_jump_ahead_identity(oracle_lib):
    # s_n = A^n s + C (A^n - 1)/(A - 1) mod 2^48, checked at n = 1 000 003 from seed 42
    A, Cc, M = 0x5DEECE66D, 0xB, (1 << 48) - 1
    s = (42 ^ A) & M
    a, c, n = 1, 0, 1000003
    ba, bc = A, Cc
    while n:
        if n & 1:
            a, c = (a * ba) & M, (c * ba + bc) & M
        ba, bc = (ba * ba) & M, (bc * ba + bc) & M
        n >>= 1
    assert ((a * s + c) & M) == oracle_lib.oracle_java_random_state_after(42, 1000003) == 0x87E51934F5B0


def test_glibc_rand_matches_libc(oracle_lib):
    libc = C.CDLL("libc.so.6")
    for seed, first in ((1, [1804289383, 846930886, 1681692777, 1714636915, 1957747793]),
                        (42, [71876166, 708592740, 1483128881, 907283241, 442951012])):
        out = np.zeros(2000, dtype=np.int32)
        oracle_lib.oracle_glibc_rand(C.c_uint(seed), 2000, out.ctypes.data_as(C.c_void_p))
        assert list(out[:5]) == first
        libc.srand(seed)
        assert [libc.rand() for _ in range(2000)] == list(out)


def fmt(lib, fn, x):
    buf = C.create_string_buffer(600)
    getattr(lib, fn)(C.c_double(x), buf, 600)
    return buf.value.decode()


def test_double_to_string(oracle_lib):
    cases = {0.05: "0.05", 1.0: "1.0", 5.0e-4: "5.0E-4", 1 / 6000: "1.6666666666666666E-4", 0.001: "0.001",
             9999999.0: "9999999.0", 1.0e7: "1.0E7", 0.3: "0.3", 2.0 / 3: "0.6666666666666666", 100.0: "100.0",
             1.0e-5: "1.0E-5", 123456.789: "123456.789", 4.9e-324: "4.9E-324", 0.0: "0.0", 12345678.9: "1.23456789E7",
             0.1 + 0.2: "0.30000000000000004", 1.0e21: "1.0E21", 0.55: "0.55"}
    for x, s in cases.items():
        assert fmt(oracle_lib, "oracle_double_to_string", x) == s
    # every value round-trips and is the shortest representation python's repr also finds
    rng = np.random.default_rng(1)
    for x in np.concatenate([rng.random(2000), rng.random(500) * 1e-4, (np.arange(1, 400) / 400.0)]):
        s = fmt(oracle_lib, "oracle_double_to_string", float(x))
        assert float(s.replace("E", "e")) == float(x)
        mant = s.split("E")[0].replace(".", "").replace("-", "").strip("0")
        assert mant == repr(float(x)).split("e")[0].replace(".", "").replace("-", "").strip("0")


def test_number_format4(oracle_lib):
    cases = {0.0: "0.0000", 1.0: "1.0000", 0.03125: "0.0312", 0.09375: "0.0938", 0.12345: "0.1235", 0.5: "0.5000",
             1234.5: "1,234.5000", 1234567.891: "1,234,567.8910", -0.25: "-0.2500", -1234.56789: "-1,234.5679",
             0.00005: "0.0001", 2.5e-5: "0.0000", 999.99995: "1,000.0000"}
    for x, s in cases.items():
        assert fmt(oracle_lib, "oracle_number_format4", x) == s, x


def test_binary_search(oracle_lib):
    a = np.array([0.1, 0.2, 0.2, 0.7, 1.0, 1.0])
    f = oracle_lib.oracle_binary_search
    p = a.ctypes.data_as(C.c_void_p)
    assert f(p, 0, 6, C.c_double(0.05)) == -1
    assert f(p, 0, 6, C.c_double(0.15)) == -2
    assert f(p, 0, 6, C.c_double(0.7)) == 3
    assert f(p, 0, 6, C.c_double(0.99)) == -5
    assert f(p, 0, 6, C.c_double(1.5)) == -7
    assert f(p, 0, 6, C.c_double(0.2)) == 2  # (0+5)>>>1 = 2 is hit first


def test_priority_queue_matches_heapq_model(oracle_lib):
    """The JDK heap restatement against an independent model: poll order by degree, arbitrary remove."""
    rng = np.random.default_rng(7)
    for trial in range(50):
        n = int(rng.integers(1, 40))
        deg = rng.integers(0, 6, size=n).astype(np.int32)
        ops = [(0 << 24) | i for i in range(n)]
        alive = set(range(n))
        for _ in range(int(rng.integers(0, 30))):
            r = rng.random()
            if r < 0.4:
                ops.append(1 << 24)
            elif r < 0.8:
                ops.append((2 << 24) | int(rng.integers(0, n)))
            else:
                ops.append((0 << 24) | int(rng.integers(0, n)))
        ops = np.array(ops, dtype=np.int32)
        cap = n + 40
        trace = np.zeros((ops.size, cap), dtype=np.int32)
        res = np.zeros(ops.size, dtype=np.int32)
        oracle_lib.oracle_pq_script(n, deg.ctypes.data_as(C.c_void_p), ops.size, ops.ctypes.data_as(C.c_void_p), cap,
                                    trace.ctypes.data_as(C.c_void_p), res.ctypes.data_as(C.c_void_p))
        content = []
        for k, op in enumerate(ops):
            code, idv = int(op) >> 24, int(op) & 0xFFFFFF
            if code == 0:
                content.append(idv)
            elif code == 1:
                if content:
                    m = min(deg[c] for c in content)
                    assert deg[res[k]] == m and res[k] in content
                    content.remove(res[k])
                else:
                    assert res[k] == -1
            elif code == 2:
                assert res[k] == (1 if idv in content else 0)
                if idv in content:
                    content.remove(idv)
            heap = [x for x in trace[k] if x >= 0]
            assert sorted(heap) == sorted(content)
            for i in range(1, len(heap)):  # heap invariant
                assert deg[heap[(i - 1) >> 1]] <= deg[heap[i]]
