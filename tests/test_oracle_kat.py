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


def _formatters():
    """(name, double->str for Double.toString, double->str for NumberFormat4): the oracle's copy and the product's copy
    (csrc/api.cu) — the same vectors must hold for both, they are separate pieces of code."""
    import oracle_helpers as oh
    from synthpop import lib as splib
    o, p = oh.load_oracle(), splib.load()
    return [("oracle", lambda x: fmt(o, "oracle_double_to_string", x), lambda x: fmt(o, "oracle_number_format4", x)),
            ("product", lambda x: fmt(p, "sp_test_double_to_string", x), lambda x: fmt(p, "sp_test_number_format4", x))]


def java_layout(x: float) -> str:
    """Double.toString's layout rules (Javadoc) applied to the shortest round-trip digits CPython's repr finds (David Gay's
    dtoa, an implementation independent of both restatements): plain decimal for 1e-3 <= |x| < 1e7, else d.dddE[-]n; at
    least one fractional digit."""
    if x == 0:
        return "-0.0" if str(x).startswith("-") else "0.0"
    r = repr(abs(x))
    if "e" in r:
        mant, e = r.split("e")
        exp = int(e)
    else:
        mant, exp = r, 0
    ip, _, fp = mant.partition(".")
    digits = (ip + fp).lstrip("0")
    exp10 = exp + len(ip) - 1 if ip.strip("0") else exp - (len(fp) - len(fp.lstrip("0"))) - 1
    digits = digits.rstrip("0") or "0"
    if len(digits) == 1:   # at least two digits are shown: the two-digit decimal closest to x (4.9E-324, not 5.0E-324)
        m2, e2 = f"{abs(x):.1e}".split("e")
        digits, exp10 = m2.replace(".", "").rstrip("0") or "0", int(e2)
    sign = "-" if x < 0 else ""
    if 1e-3 <= abs(x) < 1e7:
        if exp10 >= 0:
            whole = (digits + "0" * (exp10 + 1))[: exp10 + 1]
            frac = digits[exp10 + 1:] or "0"
            return f"{sign}{whole}.{frac}"
        return f"{sign}0.{'0' * (-exp10 - 1)}{digits}"
    return f"{sign}{digits[0]}.{digits[1:] or '0'}E{exp10}"


def test_double_to_string():
    # known answers: the Javadoc's own examples and the published regression values of the JDK-19 rewrite (Raffaello Giulietti,
    # JDK-4511638): boundaries of the plain/scientific switch, the largest, smallest normal and smallest subnormal doubles, a
    # power of two whose shortest form older JDKs missed (2^-44), 1e23 (older JDKs printed 1.0E23 only by luck of their digit loop)
    cases = {0.05: "0.05", 1.0: "1.0", 5.0e-4: "5.0E-4", 1 / 6000: "1.6666666666666666E-4", 0.001: "0.001",
             9999999.0: "9999999.0", 1.0e7: "1.0E7", 0.3: "0.3", 2.0 / 3: "0.6666666666666666", 100.0: "100.0",
             1.0e-5: "1.0E-5", 123456.789: "123456.789", 4.9e-324: "4.9E-324", 0.0: "0.0", 12345678.9: "1.23456789E7",
             0.1 + 0.2: "0.30000000000000004", 1.0e21: "1.0E21", 0.55: "0.55",
             2.0e-3: "0.002", float(np.nextafter(0.001, 0)): "9.999999999999998E-4", 9999999.999999998: "9999999.999999998",
             1.0e23: "1.0E23", 1.7976931348623157e308: "1.7976931348623157E308",
             2.2250738585072014e-308: "2.2250738585072014E-308", 2.0 ** -44: "5.684341886080802E-14", 2.0e-322: "2.0E-322",
             8.41e21: "8.41E21", 2.82879384806159e17: "2.82879384806159E17", 1.387364135037754e18: "1.387364135037754E18",
             1.45800632428665e17: "1.45800632428665E17", 123456789.0: "1.23456789E8", -0.25: "-0.25", -1.0e-10: "-1.0E-10"}
    rng = np.random.default_rng(1)
    randoms = np.concatenate([rng.random(2000), rng.random(500) * 1e-4, (np.arange(1, 400) / 400.0),
                              np.exp(rng.uniform(-700, 700, size=1500)), np.ldexp(rng.random(300), rng.integers(-1070, -1020, size=300))])
    for name, d2s, _ in _formatters():
        for x, s in cases.items():
            assert d2s(x) == s, (name, x)
            assert java_layout(x) == s, x   # the layout helper below agrees with the published strings
        for x in randoms:
            x = float(x)
            s = d2s(x)
            assert float(s.replace("E", "e")) == x, (name, x)
            assert s == java_layout(x), (name, x)


def test_number_format4():
    cases = {0.0: "0.0000", 1.0: "1.0000", 0.03125: "0.0312", 0.09375: "0.0938", 0.12345: "0.1235", 0.5: "0.5000",
             1234.5: "1,234.5000", 1234567.891: "1,234,567.8910", -0.25: "-0.2500", -1234.56789: "-1,234.5679",
             0.00005: "0.0001", 2.5e-5: "0.0000", 999.99995: "1,000.0000", 0.00015: "0.0001", 0.00025: "0.0003",
             1e9 + 0.5: "1,000,000,000.5000"}
    import decimal
    rng = np.random.default_rng(3)
    randoms = np.concatenate([rng.random(500), rng.random(300) * 5000 - 2500, rng.integers(0, 20000, size=300) / 20000.0])
    for name, _, nf4 in _formatters():
        for x, s in cases.items():
            assert nf4(x) == s, (name, x)
        for x in randoms:   # HALF_EVEN on the exact binary value, grouping by thousands: python's decimal does the same from the exact value
            x = float(x)
            q = decimal.Decimal(x).quantize(decimal.Decimal("0.0001"), rounding=decimal.ROUND_HALF_EVEN)
            want = f"{q:,.4f}"
            if want.startswith("-") and float(q) == 0:
                want = "-" + want[1:]   # NumberFormat prints "-0.0000" for a negative value that rounds to zero
            assert nf4(x) == want, (name, x)


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
