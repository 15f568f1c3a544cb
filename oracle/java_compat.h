// ORACLE — TEST INFRASTRUCTURE ONLY. Not linked into libsynthpop_cuda.so; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build or run it.
//
// JDK behaviours that the reference (fabio-cunial/population-sv-simulator, src/*.java) relies on,
// restated from their public specifications because no JVM exists in this image.
// Parity status: the JDK primitives are pinned by known-answer vectors (tests/test_oracle_kat.py);
// the end-to-end Java path is "parity unpinned" (the reference ships no test vectors and cannot run here).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include <stdexcept>
#include <algorithm>
#include <array>
#include <strings.h>

namespace jc {

// java.util.Random (48-bit LCG). Used at src/SimulateHaplotypes.java:71,386-398,624,802,809,839-873
// and src/SampleReadsFromLengthBins.java:41,57.
struct JavaRandom {
    static constexpr uint64_t MULT = 0x5DEECE66DULL, ADD = 0xBULL, MASK = (1ULL << 48) - 1;
    uint64_t seed;
    uint64_t steps = 0;  // LCG steps taken so far (used by tests of the index-addressable phases)
    explicit JavaRandom(int64_t s) { seed = ((uint64_t)s ^ MULT) & MASK; }
    int32_t next(int bits) {
        seed = (seed * MULT + ADD) & MASK;
        steps++;
        return (int32_t)(int64_t)(seed >> (48 - bits));
    }
    int32_t nextInt() { return next(32); }
    int32_t nextInt(int32_t bound) {
        if (bound <= 0) throw std::runtime_error("IllegalArgumentException: bound must be positive");
        int32_t r = next(31);
        int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        // for (int u = r; u - (r = u % bound) + m < 0; u = next(31));   with int32 wrap-around
        int32_t u = r;
        for (;;) {
            r = u % bound;
            int32_t t = (int32_t)((uint32_t)u - (uint32_t)r + (uint32_t)m);
            if (t >= 0) break;
            u = next(31);
        }
        return r;
    }
    double nextDouble() {
        int64_t hi = (int64_t)next(26);
        int64_t lo = (int64_t)next(27);
        return (double)((hi << 27) + lo) * 0x1.0p-53;
    }
    bool nextBoolean() { return next(1) != 0; }
};

// Arrays.binarySearch(double[] a, int from, int to, double key) for non-NaN finite keys
// (src/SimulateHaplotypes.java:386-395, src/SampleReadsFromLengthBins.java:57).
// The JDK compares with < and >, then falls back to bit comparison; -0.0 vs 0.0 never arises here
// because the key comes from nextDouble() (>= +0.0) and the tables are parsed non-negative decimals.
inline int binarySearch(const double* a, int from, int to, double key) {
    int lo = from, hi = to - 1;
    while (lo <= hi) {
        int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
        double v = a[mid];
        if (v < key) lo = mid + 1;
        else if (v > key) hi = mid - 1;
        else {
            int64_t mb, kb;
            memcpy(&mb, &v, 8); memcpy(&kb, &key, 8);
            if (mb == kb) return mid;
            else if (mb < kb) lo = mid + 1;
            else hi = mid - 1;
        }
    }
    return -(lo + 1);
}

// Arrays.binarySearch(Object[] a, Object key) on Strings (src/SimulateHaplotypes.java:101).
inline int binarySearchStr(const char* const* a, int n, const std::string& key) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
        int c = strcmp(a[mid], key.c_str());
        if (c < 0) lo = mid + 1; else if (c > 0) hi = mid - 1; else return mid;
    }
    return -(lo + 1);
}

// Double.toString(double) (JDK >= 19: shortest decimal that round-trips, closest to the true value;
// format rules of the Javadoc). Used for "AF=" (src/SimulateHaplotypes.java:1275) and variants.txt (:1285).
inline std::string doubleToString(double d) {
    if (std::isnan(d)) return "NaN";
    if (std::isinf(d)) return d > 0 ? "Infinity" : "-Infinity";
    if (d == 0.0) return std::signbit(d) ? "-0.0" : "0.0";
    char buf[64];
    // The shortest decimal that reads back as d, at least two significant digits (4.9E-324). It is not always the
    // correctly rounded one: next to a power of two the gap below is half the gap above, and the nearest n-digit decimal
    // can fall outside the narrow side while its neighbour inside the wide side still reads back (2^-44 =
    // 5.6843418860808015e-14 prints as 5.684341886080802E-14). So at every length the rounded decimal and its two
    // neighbours in the last place are tried, and of those that read back the one closest to d wins.
    std::string best;
    for (int prec = 1; prec <= 16 && best.empty(); prec++) {
        snprintf(buf, sizeof buf, "%.*e", prec, d);
        const std::string r(buf);
        const size_t ep = r.find('e');
        long double bestDist = 0;
        for (int k = 0; k < 3; k++) {
            const int delta = k == 0 ? 0 : (k == 1 ? -1 : 1);   // the correctly rounded one first: it wins a tie (half-even, as the JDK)
            std::string c = r;
            if (delta != 0) {   // one unit in the last place of the mantissa, as decimal string arithmetic
                std::string m;
                for (size_t i = 0; i < ep; i++) if (c[i] != '.' && c[i] != '-') m.push_back(c[i]);
                long long v = atoll(m.c_str()) + delta;
                int e10 = atoi(c.c_str() + ep + 1);
                std::string mv = std::to_string(v);
                if (mv.size() > m.size()) { e10++; mv.pop_back(); }            // 9.99 + 0.01 -> 10.0
                else if (mv.size() < m.size()) { e10--; mv.push_back('9'); }    // 1.00 - 0.01 -> 0.99 -> 9.99e-1 (one more 9 keeps the length)
                snprintf(buf, sizeof buf, "%s%c.%se%d", d < 0 ? "-" : "", mv[0], mv.c_str() + 1, e10);
                c = buf;
            }
            if (strtod(c.c_str(), nullptr) != d) continue;
            const long double dist = fabsl(strtold(c.c_str(), nullptr) - (long double)d);
            if (best.empty() || dist < bestDist) { best = c; bestDist = dist; }
        }
    }
    if (best.empty()) { snprintf(buf, sizeof buf, "%.16e", d); best = buf; }
    // parse "[-]D.DDDDe[+-]XX"
    std::string s(best);
    bool neg = s[0] == '-';
    if (neg) s = s.substr(1);
    size_t epos = s.find('e');
    int exp10 = atoi(s.c_str() + epos + 1);
    std::string digits;
    for (size_t i = 0; i < epos; i++) if (s[i] != '.') digits.push_back(s[i]);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out = neg ? "-" : "";
    double ad = std::fabs(d);
    if (ad >= 1e-3 && ad < 1e7) {
        if (exp10 >= 0) {
            std::string ip, fp;
            for (int i = 0; i <= exp10; i++) ip.push_back(i < (int)digits.size() ? digits[i] : '0');
            if ((int)digits.size() > exp10 + 1) fp = digits.substr(exp10 + 1);
            if (fp.empty()) fp = "0";
            out += ip + "." + fp;
        } else {
            out += "0.";
            for (int i = 0; i < -exp10 - 1; i++) out.push_back('0');
            out += digits;
        }
    } else {
        out.push_back(digits[0]);
        out.push_back('.');
        if (digits.size() > 1) out += digits.substr(1); else out.push_back('0');
        out += "E" + std::to_string(exp10);
    }
    return out;
}

// NumberFormat.getInstance() with min=max=4 fraction digits (src/SimulateHaplotypes.java:1010-1011,1278):
// DecimalFormat, RoundingMode.HALF_EVEN on the exact binary value, grouping size 3 with ','.
inline std::string numberFormat4(double x) {
    if (std::isnan(x)) return "NaN";
    if (std::isinf(x)) return x > 0 ? "\xE2\x88\x9E" : "-\xE2\x88\x9E";
    char buf[512];
    snprintf(buf, sizeof buf, "%.4f", x);  // glibc: exact value, round-half-even in the default rounding mode
    std::string s(buf);
    bool neg = s[0] == '-';
    if (neg) s = s.substr(1);
    size_t dot = s.find('.');
    std::string ip = s.substr(0, dot), fp = s.substr(dot);
    std::string g;
    for (size_t i = 0; i < ip.size(); i++) {
        if (i > 0 && (ip.size() - i) % 3 == 0) g.push_back(',');
        g.push_back(ip[i]);
    }
    return (neg ? "-" : "") + g + fp;
}

// BufferedReader.readLine(): lines end at \n, \r or \r\n; String.trim(): strip chars <= ' ' on both ends.
struct LineReader {
    FILE* f;
    bool pendingCR = false;
    explicit LineReader(const std::string& path) {
        f = fopen(path.c_str(), "rb");
        if (!f) throw std::runtime_error("FileNotFoundException: " + path);
        setvbuf(f, nullptr, _IOFBF, 1 << 20);
    }
    ~LineReader() { if (f) fclose(f); }
    bool readLine(std::string& out) {
        out.clear();
        int c;
        bool any = false;
        while ((c = getc_unlocked(f)) != EOF) {
            if (pendingCR) { pendingCR = false; if (c == '\n') continue; }
            any = true;
            if (c == '\n') return true;
            if (c == '\r') { pendingCR = true; return true; }
            out.push_back((char)c);
        }
        return any;
    }
};
inline std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && (unsigned char)s[a] <= ' ') a++;
    while (b > a && (unsigned char)s[b - 1] <= ' ') b--;
    return s.substr(a, b - a);
}
// String.split(","): trailing empty strings are removed.
inline std::vector<std::string> splitComma(const std::string& s) {
    std::vector<std::string> t;
    size_t p = 0;
    for (;;) {
        size_t q = s.find(',', p);
        if (q == std::string::npos) { t.push_back(s.substr(p)); break; }
        t.push_back(s.substr(p, q - p));
        p = q + 1;
    }
    while (!t.empty() && t.back().empty()) t.pop_back();
    return t;
}
inline int parseInt(const std::string& s) {
    // Integer.parseInt: optional sign, decimal digits only, no whitespace.
    if (s.empty()) throw std::runtime_error("NumberFormatException: empty");
    size_t i = 0; bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; }
    if (i == s.size()) throw std::runtime_error("NumberFormatException: " + s);
    int64_t v = 0;
    for (; i < s.size(); i++) {
        if (s[i] < '0' || s[i] > '9') throw std::runtime_error("NumberFormatException: " + s);
        v = v * 10 + (s[i] - '0');
        if (v > (int64_t)INT32_MAX + 1) throw std::runtime_error("NumberFormatException: " + s);
    }
    v = neg ? -v : v;
    if (v > INT32_MAX || v < INT32_MIN) throw std::runtime_error("NumberFormatException: " + s);
    return (int)v;
}
inline double parseDouble(const std::string& s) {
    // Double.parseDouble trims whitespace and is correctly rounded, as is glibc strtod.
    std::string t = trim(s);
    char* end = nullptr;
    double d = strtod(t.c_str(), &end);
    if (t.empty() || end != t.c_str() + t.size()) throw std::runtime_error("NumberFormatException: " + s);
    return d;
}

// java.util.PriorityQueue<E> on (degree) keys with equals() by id, restated from the JDK sources
// (SURVEY.md Appendix A.2). Elements are ids into external degree[] (the QueueElement objects are shared).
struct JavaPriorityQueue {
    std::vector<int> q;  // Object[] queue
    int size = 0;
    const int* degree;   // QueueElement.degree, indexed by variantID
    explicit JavaPriorityQueue(int cap, const int* deg) : q(cap > 0 ? cap : 1), degree(deg) {}
    int cmp(int a, int b) const { return degree[a] < degree[b] ? -1 : (degree[a] > degree[b] ? 1 : 0); }
    void clear() { size = 0; }
    bool isEmpty() const { return size == 0; }
    void siftUp(int k, int x) {
        while (k > 0) {
            int parent = (int)(((unsigned)(k - 1)) >> 1);
            int e = q[parent];
            if (cmp(x, e) >= 0) break;
            q[k] = e; k = parent;
        }
        q[k] = x;
    }
    void siftDown(int k, int x, int n) {
        int half = (int)(((unsigned)n) >> 1);
        while (k < half) {
            int child = (k << 1) + 1;
            int c = q[child];
            int right = child + 1;
            if (right < n && cmp(c, q[right]) > 0) c = q[child = right];
            if (cmp(x, c) <= 0) break;
            q[k] = c; k = child;
        }
        q[k] = x;
    }
    void add(int x) {
        int i = size;
        if (i >= (int)q.size()) q.resize(q.size() * 2);
        siftUp(i, x);
        size = i + 1;
    }
    int poll() {
        int result = q[0];
        int n = --size;
        int x = q[n];
        if (n > 0) siftDown(0, x, n);
        return result;
    }
    bool remove(int o) {
        int i = -1;
        for (int k = 0; k < size; k++) if (q[k] == o) { i = k; break; }
        if (i < 0) return false;
        int s = --size;
        if (s != i) {
            int moved = q[s];
            siftDown(i, moved, s);
            if (q[i] == moved) siftUp(i, moved);
        }
        return true;
    }
};

}  // namespace jc
