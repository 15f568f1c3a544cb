// Host-side text readers/writers of the reference's file contracts (product code, C++; no CUDA here).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "common.cuh"

namespace sp {

// Whole file in memory.
inline std::vector<char> slurp(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw Error(SP_ERR_JAVA_EXCEPTION, "FileNotFoundException: " + path);
    fseeko(f, 0, SEEK_END);
    const off_t n = ftello(f);
    fseeko(f, 0, SEEK_SET);
    std::vector<char> buf((size_t)n);
    if (n > 0 && fread(buf.data(), 1, (size_t)n, f) != (size_t)n) { fclose(f); throw Error(SP_ERR, "short read: " + path); }
    fclose(f);
    return buf;
}

// Iterates lines the way java.io.BufferedReader.readLine does (\n, \r or \r\n end a line).
struct Lines {
    const char* p;
    const char* end;
    Lines(const char* b, const char* e) : p(b), end(e) {}
    bool next(const char*& b, const char*& e) {
        if (p >= end) return false;
        b = p;
        const char* q = p;
        while (q < end && *q != '\n' && *q != '\r') q++;
        e = q;
        if (q < end) {
            if (*q == '\r' && q + 1 < end && q[1] == '\n') q += 2; else q += 1;
        }
        p = q;
        return true;
    }
};

inline void trim_span(const char*& b, const char*& e) {  // String.trim()
    while (b < e && (unsigned char)*b <= ' ') b++;
    while (e > b && (unsigned char)e[-1] <= ' ') e--;
}

inline int32_t parse_int(const char* b, const char* e, const char* what) {  // Integer.parseInt
    if (b >= e) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: empty ") + what);
    bool neg = false;
    const char* q = b;
    if (*q == '-' || *q == '+') { neg = *q == '-'; q++; }
    if (q >= e) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: ") + what);
    int64_t v = 0;
    for (; q < e; q++) {
        if (*q < '0' || *q > '9') throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: ") + std::string(b, e) + " (" + what + ")");
        v = v * 10 + (*q - '0');
        if (v > (int64_t)INT32_MAX + 1) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: ") + what);
    }
    v = neg ? -v : v;
    if (v > INT32_MAX || v < INT32_MIN) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: ") + what);
    return (int32_t)v;
}

inline double parse_double(const char* b, const char* e, const char* what) {  // Double.parseDouble (correctly rounded)
    trim_span(b, e);
    std::string s(b, e);
    char* endp = nullptr;
    const double d = strtod(s.c_str(), &endp);
    if (s.empty() || endp != s.c_str() + s.size()) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: ") + s + " (" + what + ")");
    return d;
}

// String.split(","): fields between commas, trailing empty fields dropped.
inline void split_comma(const char* b, const char* e, std::vector<std::pair<const char*, const char*>>& out) {
    out.clear();
    const char* q = b;
    for (;;) {
        const char* c = (const char*)memchr(q, ',', (size_t)(e - q));
        if (!c) { out.emplace_back(q, e); break; }
        out.emplace_back(q, c);
        q = c + 1;
    }
    while (!out.empty() && out.back().first == out.back().second) out.pop_back();
}

// Double.toString (JDK >= 19 shortest round-trip digits) and NumberFormat(4 digits, HALF_EVEN, grouping).
std::string java_double_to_string(double d);
std::string java_number_format4(double x);

}  // namespace sp
