// ORACLE — TEST INFRASTRUCTURE ONLY (see java_compat.h). CPU restatement of the reference's src/AnnotateVCF.java
// (SURVEY.md §8f rank 2): same control flow, same string handling, on the materialised position arrays of
// SimulateHaplotypes.loadPositions. Parity unpinned by the reference (it ships no tests and no JVM exists here); the
// JDK string primitives it leans on (String.split, indexOf, Integer.parseInt, Math.round, NumberFormat) are restated
// from their specifications and pinned by tests/test_annotate_vcf.py's known answers.
// Restriction (stated, not hidden): Character.isDigit is taken over ASCII; Double.parseDouble over what strtod accepts.
#include "simulate_haplotypes.h"
#include <climits>

namespace oracle {

namespace {

const int TYPE_INSERTION = 1, TYPE_DELETION = 2, TYPE_DEL_INV = 3, TYPE_INVERSION = 4, TYPE_INV_DUP = 5,
          TYPE_DUPLICATION = 6, TYPE_CNV = 7, TYPE_BREAKEND = 8, TYPE_TRANSLOCATION = 9;  // src/AnnotateVCF.java:30-38

bool equalsIgnoreCase(const std::string& a, const char* b) { return a.size() == strlen(b) && strcasecmp(a.c_str(), b) == 0; }

// String.indexOf(String, fromIndex)
int indexOf(const std::string& s, const std::string& what, int from = 0) {
    if (from < 0) from = 0;
    if ((size_t)from > s.size()) return -1;
    size_t p = s.find(what, (size_t)from);
    return p == std::string::npos ? -1 : (int)p;
}

// String.substring(begin,end): StringIndexOutOfBoundsException outside 0<=begin<=end<=length
std::string substring(const std::string& s, int b, int e) {
    if (b < 0 || e > (int)s.size() || b > e) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: begin " + std::to_string(b) + ", end " + std::to_string(e) + ", length " + std::to_string(s.size()));
    return s.substr((size_t)b, (size_t)(e - b));
}
std::string substring(const std::string& s, int b) { return substring(s, b, (int)s.size()); }
char charAt(const std::string& s, int i) {
    if (i < 0 || i >= (int)s.size()) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: index " + std::to_string(i) + ", length " + std::to_string(s.size()));
    return s[(size_t)i];
}

// Integer.parseInt(String): optional sign, decimal digits, int range
int parseInt(const std::string& s) {
    auto bad = [&]() { return std::runtime_error("java.lang.NumberFormatException: For input string: \"" + s + "\""); };
    if (s.empty()) throw bad();
    size_t i = 0;
    bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; if (s.size() == 1) throw bad(); }
    int64_t v = 0;
    for (; i < s.size(); i++) {
        if (s[i] < '0' || s[i] > '9') throw bad();
        v = v * 10 + (s[i] - '0');
        if (v > (int64_t)INT_MAX + 1) throw bad();
    }
    if (neg) v = -v;
    if (v > INT_MAX || v < INT_MIN) throw bad();
    return (int)v;
}

// (int)Math.round(Double.parseDouble(s) * 3): floor(x + 1/2) as a long, then the narrowing cast
int roundTimes3(const std::string& s) {
    char* end = nullptr;
    std::string t = s;
    // Double.parseDouble trims whitespace
    size_t a = 0, b = t.size();
    while (a < b && (unsigned char)t[a] <= ' ') a++;
    while (b > a && (unsigned char)t[b - 1] <= ' ') b--;
    t = t.substr(a, b - a);
    if (t.empty()) throw std::runtime_error("java.lang.NumberFormatException: empty String");
    double d = strtod(t.c_str(), &end);
    if (*end != 0) throw std::runtime_error("java.lang.NumberFormatException: For input string: \"" + s + "\"");
    double x = d * 3;
    if (x != x) return 0;
    double f = std::floor(x + 0.5);
    long long r;
    if (f >= 9.2233720368547758e18) r = LLONG_MAX; else if (f <= -9.2233720368547758e18) r = LLONG_MIN; else r = (long long)f;
    return (int)(int32_t)(uint32_t)(uint64_t)r;
}

// String.split("\t") (limit 0): trailing empty strings are removed; an empty input gives one empty token
std::vector<std::string> splitTab(const std::string& s) {
    std::vector<std::string> out;
    size_t a = 0;
    for (;;) {
        size_t p = s.find('\t', a);
        if (p == std::string::npos) { out.push_back(s.substr(a)); break; }
        out.push_back(s.substr(a, p - a));
        a = p + 1;
    }
    while (out.size() > 1 && out.back().empty()) out.pop_back();
    if (out.size() == 1 && out[0].empty() && !s.empty()) out.clear();  // "\t\t" -> no tokens at all
    return out;
}

// BufferedReader.readLine: lines end at \n, \r or \r\n; a final line without terminator counts; EOF -> false
struct LineReader {
    std::string buf;
    size_t pos = 0;
    explicit LineReader(const std::string& path) {
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) throw std::runtime_error("java.io.FileNotFoundException: " + path + " (No such file or directory)");
        char tmp[1 << 16];
        size_t n;
        while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.append(tmp, n);
        fclose(f);
    }
    bool next(std::string& line) {
        if (pos >= buf.size()) return false;
        size_t e = pos;
        while (e < buf.size() && buf[e] != '\n' && buf[e] != '\r') e++;
        line.assign(buf, pos, e - pos);
        if (e < buf.size()) { if (buf[e] == '\r' && e + 1 < buf.size() && buf[e + 1] == '\n') e++; e++; }
        pos = e;
        return true;
    }
};

}  // namespace

struct AnnotateVCF {
    SimulateHaplotypes& sh;
    explicit AnnotateVCF(SimulateHaplotypes& s) : sh(s) {}

    // :209-220; the boolean says whether the Java's `null` was returned
    static bool getField(const std::string& str, const std::string& field, std::string& out) {
        const int FIELD_LENGTH = (int)field.size() + 1;
        int p = indexOf(str, field + "=");
        if (p < 0) return false;
        if (equalsIgnoreCase(field, "END")) {
            while (p >= 2 && equalsIgnoreCase(substring(str, p - 2, p - 2 + 5), "CIEND")) p = indexOf(str, field + "=", p + 1);
            if (p < 0) return false;
        }
        const int q = indexOf(str, ";", p + FIELD_LENGTH);
        out = substring(str, p + FIELD_LENGTH, q < 0 ? (int)str.size() : q);
        return true;
    }

    static int getType_infoField(bool isNull, const std::string& type) {  // :226-266
        if (isNull || type.empty()) return -1;
        if (equalsIgnoreCase(type, "DEL") || equalsIgnoreCase(type, "DEL:ME")) return TYPE_DELETION;
        else if (equalsIgnoreCase(type, "DEL/INV")) return TYPE_DEL_INV;
        else if (equalsIgnoreCase(type, "INS") || equalsIgnoreCase(type, "INS:ME") || equalsIgnoreCase(type, "INS:NOVEL")) return TYPE_INSERTION;
        else if (equalsIgnoreCase(type, "DUP") || equalsIgnoreCase(type, "DUP:TANDEM") || equalsIgnoreCase(type, "DUP:INT")) return TYPE_DUPLICATION;
        else if (equalsIgnoreCase(type, "INV")) return TYPE_INVERSION;
        else if (equalsIgnoreCase(type, "INVDUP")) return TYPE_INV_DUP;
        else if (equalsIgnoreCase(type, "CNV")) return TYPE_CNV;
        else if (equalsIgnoreCase(type, "BND")) return TYPE_BREAKEND;
        else if (equalsIgnoreCase(type, "TRA")) return TYPE_TRANSLOCATION;
        else return -1;
    }

    static int getType_altField(const std::string& str) {  // :269-290
        if (str.empty()) return -1;
        if (equalsIgnoreCase(str, "<DEL>") || equalsIgnoreCase(str, "<DEL:ME>")) return TYPE_DELETION;
        else if (equalsIgnoreCase(str, "<DEL/INV>")) return TYPE_DEL_INV;
        else if (equalsIgnoreCase(str, "<INS>") || equalsIgnoreCase(str, "<INS:ME>") || equalsIgnoreCase(str, "<INS:NOVEL>")) return TYPE_INSERTION;
        else if (equalsIgnoreCase(str, "<DUP>") || equalsIgnoreCase(str, "<DUP:TANDEM>") || equalsIgnoreCase(str, "<DUP:INT>")) return TYPE_DUPLICATION;
        else if (equalsIgnoreCase(str, "<INV>")) return TYPE_INVERSION;
        else if (equalsIgnoreCase(str, "<INVDUP>")) return TYPE_INV_DUP;
        else if (equalsIgnoreCase(str, "<CNV>")) return TYPE_CNV;
        else if (equalsIgnoreCase(str, "<BND>")) return TYPE_BREAKEND;
        else if (equalsIgnoreCase(str, "<TRA>")) return TYPE_TRANSLOCATION;
        else return -1;
    }

    // the CIPOS/CIEND/CILEN pattern (:306-327, :363-403): two comma-separated integers after the key
    static void pairAfter(const std::string& str, int p, int keyLength, bool allowMinus, int* out) {
        p += keyLength + 1;
        int q = indexOf(str, ",", p);
        char c = charAt(str, p);
        if ((c >= '0' && c <= '9') || (allowMinus && c == '-')) out[0] = q >= 0 ? parseInt(substring(str, p, q)) : parseInt(substring(str, p));
        else out[0] = 0;
        p = q + 1; q = indexOf(str, ";", p);
        c = charAt(str, p);
        if ((c >= '0' && c <= '9') || (allowMinus && c == '-')) out[1] = q >= 0 ? parseInt(substring(str, p, q)) : parseInt(substring(str, p));
        else out[1] = 0;
    }
    // the STD_* pattern (:328-359, :404-431): one standard deviation, times SIGMA_MULTIPLE = 3
    static void sigmaAfter(const std::string& str, int p, int keyLength, int* out) {
        p += keyLength + 1;
        const int q = indexOf(str, ";", p);
        const char c = charAt(str, p);
        if (c >= '0' && c <= '9') {
            const int quantum = q >= 0 ? roundTimes3(substring(str, p, q)) : roundTimes3(substring(str, p));
            out[0] = -quantum; out[1] = quantum;
        } else { out[0] = 0; out[1] = 0; }
    }

    static void getConfidenceInterval(const std::string& str, bool posID, int* out) {  // :299-433
        int p;
        if (!posID) {
            if ((p = indexOf(str, "CIPOS")) >= 0) { pairAfter(str, p, 5, true, out); return; }
            if ((p = indexOf(str, "STD_quant_start")) >= 0) { sigmaAfter(str, p, 15, out); return; }
            if ((p = indexOf(str, "STD_POS1")) >= 0) { sigmaAfter(str, p, 8, out); return; }
        } else {
            if ((p = indexOf(str, "CIEND")) >= 0) { pairAfter(str, p, 5, true, out); return; }
            if ((p = indexOf(str, "CILEN")) >= 0) { pairAfter(str, p, 5, false, out); return; }
            if ((p = indexOf(str, "STD_quant_stop")) >= 0) { sigmaAfter(str, p, 14, out); return; }
            if ((p = indexOf(str, "STD_POS2")) >= 0) { sigmaAfter(str, p, 8, out); return; }
        }
        out[0] = 0; out[1] = 0;
    }

    static bool nonemptyIntersection(const std::vector<int>& x1, int from1, int last1, int from2, int last2) {  // :463-472
        int i1 = from1, i2 = from2;
        while (i1 <= last1 && i2 <= last2) {
            if (x1[i1] < i2) i1++;
            else if (x1[i1] > i2) i2++;
            else return true;
        }
        return false;
    }

    static int intersection(const std::vector<int>& x1, int from1, int last1, int from2, int last2, std::vector<int>& out, int lastOut) {  // :498-510
        int i1 = from1, i2 = from2, j = lastOut;
        while (i1 <= last1 && i2 <= last2) {
            if (x1[i1] < i2) i1++;
            else if (x1[i1] > i2) i2++;
            else {
                if (j + 1 >= (int)out.size()) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index " + std::to_string(j + 1) + " out of bounds for length " + std::to_string(out.size()));
                out[++j] = i2;
                i1++; i2++;
            }
        }
        return j;
    }

    static int binarySearchInt(const std::vector<int>& a, int from, int to, int key) {  // Arrays.binarySearch(int[],from,to,key)
        int lo = from, hi = to - 1;
        while (lo <= hi) {
            const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
            if (a[mid] < key) lo = mid + 1;
            else if (a[mid] > key) hi = mid - 1;
            else return mid;
        }
        return -(lo + 1);
    }

    int getRepeatType(int first, int last, bool* tmpArray) {  // :437-457
        for (int i = 0; i <= 4; i++) {
            int j = binarySearchInt(sh.positions[i], 0, sh.positions_last[i] + 1, first);
            if (j < 0) j = -1 - j;
            tmpArray[i] = nonemptyIntersection(sh.positions[i], j, sh.positions_last[i], first, last);
        }
        if (tmpArray[2] || (tmpArray[0] && tmpArray[1])) return 2;
        else if (tmpArray[0]) return 0;
        else if (tmpArray[1]) return 1;
        else if (tmpArray[3]) return 3;
        else if (tmpArray[4]) return 4;
        else return 4;  // after a warning on stderr
    }

    int getRepeatSurface(int first, int last, std::vector<int>& tmpArray) {  // :480-495
        int j = -1, k;
        for (int i = 0; i <= 3; i++) {
            k = binarySearchInt(sh.positions[i], 0, sh.positions_last[i] + 1, first);
            if (k < 0) k = -1 - k;
            j = intersection(sh.positions[i], k, sh.positions_last[i], first, last, tmpArray, j);
        }
        if (j == 0) return 1;
        std::sort(tmpArray.begin(), tmpArray.begin() + (j + 1));
        k = 0;
        for (int i = 1; i <= j; i++) {
            if (tmpArray[i] != tmpArray[k]) tmpArray[++k] = tmpArray[i];
        }
        return k + 1;
    }

    void annotateVCF(const std::string& inputFile, const std::string& outputFile, int minDistance, int maxDistance) {  // :127-205
        const int MAX_LENGTH_DIFFERENCE = 100;
        int position, end, length, lengthPrime, type;
        int intervalFirst, intervalLast, repeatsStart, repeatsEnd;
        double repeatsFraction;
        std::string str, strPrime;
        bool tmpBoolean[5];
        std::vector<int> tmpInt(20000);
        int out[2];
        LineReader br(inputFile);
        FILE* bw = fopen(outputFile.c_str(), "wb");
        if (!bw) throw std::runtime_error("java.io.FileNotFoundException: " + outputFile);
        struct Closer { FILE* f; ~Closer() { fclose(f); } } closer{bw};
        while (br.next(str)) {
            if (charAt(str, 0) == '#') {
                fwrite(str.data(), 1, str.size(), bw); fputc('\n', bw);
                continue;
            }
            std::vector<std::string> tokens = splitTab(str);
            auto tok = [&](size_t i) -> std::string& {
                if (i >= tokens.size()) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index " + std::to_string(i) + " out of bounds for length " + std::to_string(tokens.size()));
                return tokens[i];
            };
            bool has = getField(tok(7), "SVTYPE", strPrime);
            type = getType_infoField(!has, strPrime);
            if (type == -1) type = getType_altField(tok(4));
            if (type != TYPE_DELETION && type != TYPE_DUPLICATION && type != TYPE_INSERTION && type != TYPE_INVERSION) {
                fwrite(str.data(), 1, str.size(), bw); fputc('\n', bw);
                continue;
            }
            position = parseInt(tok(1));
            if (getField(tok(7), "END", strPrime)) {
                end = (int)((uint32_t)parseInt(strPrime) - 1u);
                if (type != TYPE_INSERTION) {
                    length = (int)((uint32_t)end - (uint32_t)position + 1u);
                    if (getField(tok(7), "SVLEN", strPrime)) {
                        lengthPrime = parseInt(strPrime);
                        if (lengthPrime < 0) lengthPrime = -lengthPrime;
                        (void)MAX_LENGTH_DIFFERENCE;  // only a warning on stderr depends on it
                    }
                } else length = INT_MAX;
            } else if (type != TYPE_INSERTION) {
                if (getField(tok(7), "SVLEN", strPrime)) {
                    length = parseInt(strPrime);
                    if (length < 0) length = (int)(0u - (uint32_t)length);
                    end = (int)((uint32_t)position + (uint32_t)length - 1u);
                } else { end = INT_MAX; length = INT_MAX; }
            } else { end = position; length = INT_MAX; }
            auto add = [](int a, int b) { return (int)((uint32_t)a + (uint32_t)b); };  // Java int arithmetic wraps
            getConfidenceInterval(tok(7), false, out);
            intervalFirst = add(position, out[0]); intervalLast = add(position, out[1]);
            if (intervalFirst < add(position, -maxDistance)) intervalFirst = add(position, -maxDistance);
            if (intervalFirst > add(position, -minDistance)) intervalFirst = add(position, -minDistance);
            if (intervalLast > add(position, maxDistance)) intervalLast = add(position, maxDistance);
            if (intervalLast < add(position, minDistance)) intervalLast = add(position, minDistance);
            repeatsStart = getRepeatType(intervalFirst, intervalLast, tmpBoolean);
            if (end != INT_MAX) {
                getConfidenceInterval(tok(7), true, out);
                intervalFirst = add(end, out[0]); intervalLast = add(end, out[1]);
                // (the clamps below are relative to `position`, not to `end`: :184-187, as written)
                if (intervalFirst < add(position, -maxDistance)) intervalFirst = add(position, -maxDistance);
                if (intervalFirst > add(position, -minDistance)) intervalFirst = add(position, -minDistance);
                if (intervalLast > add(position, maxDistance)) intervalLast = add(position, maxDistance);
                if (intervalLast < add(position, minDistance)) intervalLast = add(position, minDistance);
                repeatsEnd = getRepeatType(intervalFirst, intervalLast, tmpBoolean);
                if (length != INT_MAX) {
                    const int need = (int)(((uint32_t)end - (uint32_t)position + 1u) * 4u);
                    if (need > (int)tmpInt.size()) tmpInt.assign((size_t)need, 0);
                    repeatsFraction = getRepeatSurface(position, end, tmpInt) / ((double)length);
                } else repeatsFraction = 0.0;
            } else { repeatsFraction = 0.0; repeatsEnd = 4; }
            tok(7) += ";REPEATS_START=" + std::to_string(repeatsStart) + ";REPEATS_END=" + std::to_string(repeatsEnd) +
                      ";REPEATS_FRACTION=" + jc::numberFormat4(repeatsFraction);
            fwrite(tokens[0].data(), 1, tokens[0].size(), bw);
            for (size_t i = 1; i < tokens.size(); i++) { fputc('\t', bw); fwrite(tokens[i].data(), 1, tokens[i].size(), bw); }
            fputc('\n', bw);
        }
    }
};

// AnnotateVCF.main (src/AnnotateVCF.java:96-124): args = INPUT_FILES_LIST MIN_DISTANCE MAX_DISTANCE REFERENCE_FILE
// REPEAT_INTERVALS_FILE SEGDUPS_FILE
int annotateVcfMain(const std::vector<std::string>& args) {
    if (args.size() < 6) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index " + std::to_string(args.size()) + " out of bounds for length " + std::to_string(args.size()));
    const int MIN_DISTANCE = parseInt(args[1]), MAX_DISTANCE = parseInt(args[2]);
    SimulateHaplotypes sh(0);
    sh.verbose = false;
    sh.loadReference(args[3]);
    sh.model.deserializeRepeatIntervals(args[4]);
    sh.model.deserializeSegdups(args[5]);
    sh.loadPositions(sh.referenceLength, BuildModel::MAX_DISTANCE_SR);
    AnnotateVCF av(sh);
    LineReader br(args[0]);
    std::string str;
    while (br.next(str)) {
        const int p = indexOf(str, ".vcf");
        av.annotateVCF(str, substring(str, 0, p) + "_annotated.vcf", MIN_DISTANCE, MAX_DISTANCE);
    }
    return 0;
}

}  // namespace oracle
