// Host-side model loaders and interval algebra (see model.h). Pure host code, compiled with nvcc for one .so.
#include "model.h"
#include "textio.h"
#include <algorithm>

namespace sp {

const char* const SV_TYPE_NAMES[4] = {"DEL", "DUP", "INS", "INV"};
const char* const AF_NAMES[6] = {"ALL", "AFR", "AMR", "EAS", "EUR", "OTH"};
const int ADD_TYPE[3][3] = {{0, 2, 2}, {2, 1, 2}, {2, 2, 2}};

namespace {
using Tok = std::vector<std::pair<const char*, const char*>>;

struct FileLines {
    std::vector<char> buf;
    Lines lines;
    std::string path;
    explicit FileLines(const std::string& p) : buf(slurp(p)), lines(buf.data(), buf.data() + buf.size()), path(p) {}
    void must(const char*& b, const char*& e) {
        if (!lines.next(b, e)) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: unexpected end of " + path);
    }
    void row(Tok& t) {
        const char *b, *e;
        must(b, e);
        split_comma(b, e, t);
    }
};
const std::pair<const char*, const char*>& at(const Tok& t, size_t i) {
    if (i >= t.size()) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: " + std::to_string(i));
    return t[i];
}
int32_t tok_int(const Tok& t, size_t i) { return parse_int(at(t, i).first, at(t, i).second, "model header"); }
double tok_double(const Tok& t, size_t i) { return parse_double(at(t, i).first, at(t, i).second, "model value"); }
}  // namespace

void Model::load_statistics(const std::string& dir) {
    Tok t;
    const char *b, *e;
    {
        FileLines f(dir + "/statistics_svType.txt");
        f.must(b, e);
        const int n = parse_int(b, e, "statistics_svType length");
        f.row(t);
        svType.assign(n, 0.0);
        for (int i = 0; i < n; i++) svType[i] = tok_double(t, i);
    }
    {
        FileLines f(dir + "/statistics_intervalType.txt");
        f.row(t);
        const int ni = tok_int(t, 0), nj = tok_int(t, 1);
        intervalType.assign(ni, std::vector<double>(nj, 0.0));
        for (int i = 0; i < ni; i++) {
            f.row(t);
            for (int j = 0; j < nj; j++) intervalType[i][j] = tok_double(t, j);
        }
    }
    {
        FileLines f(dir + "/statistics_length.txt");
        f.row(t);
        const int ni = tok_int(t, 0), nj = tok_int(t, 1), nk = tok_int(t, 2);
        length.assign(ni, std::vector<std::vector<double>>(nj));
        for (int i = 0; i < ni; i++)
            for (int j = 0; j < nj; j++) {
                f.row(t);
                std::vector<double>& r = length[i][j];
                r.assign(nk, 0.0);
                for (int k = 0; k < nk; k++) r[k] = tok_double(t, k);
            }
    }
    af.assign(6, {});
    af_minmax.assign(6, {});
    for (int p = 0; p < 6; p++) {
        FileLines f(dir + "/statistics_alleleFrequency_" + AF_NAMES[p] + ".txt");
        f.row(t);
        const int ni = tok_int(t, 0), nj = tok_int(t, 1);
        af[p].assign(ni, std::vector<double>(nj, 0.0));
        af_minmax[p].assign(ni, {0.0, 0.0});
        for (int i = 0; i < ni; i++) {
            f.row(t);
            af_minmax[p][i][0] = tok_double(t, 0);
            af_minmax[p][i][1] = tok_double(t, 1);
            for (int j = 0; j < nj; j++) af[p][i][j] = tok_double(t, 2 + j);
        }
    }
    {   // statistics_position.txt is parsed by the reference (and must exist and be well formed) but never used
        FileLines f(dir + "/statistics_position.txt");
        f.row(t);
        const int ni = tok_int(t, 0), nj = tok_int(t, 1);
        for (int i = 0; i < ni; i++) {
            f.row(t);
            for (int j = 0; j < nj; j++) tok_double(t, j);
        }
    }
}

void Model::load_repeat_intervals(const std::string& path) {
    FileLines f(path);
    const char *b, *e;
    f.must(b, e);
    const int n = parse_int(b, e, "repeat interval count");
    if (n < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: repeatIntervals");
    repeats.assign(n, {0, 0, 0});
    Tok t;
    for (int i = 0; i < n; i++) {
        f.row(t);
        for (int j = 0; j < 3; j++) repeats[i][j] = tok_int(t, j);
    }
}

void Model::load_segdups(const std::string& path) {
    FileLines f(path);
    const char *b, *e;
    f.must(b, e);
    const int n = parse_int(b, e, "segdup count");
    if (n < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: segdups");
    segdups.assign(n, {0, 0});
    Tok t;
    for (int i = 0; i < n; i++) {
        f.row(t);
        segdups[i][0] = tok_int(t, 0);
        segdups[i][1] = tok_int(t, 1);
    }
}

// ---------------------------------------------------------------------------------------------- intervals
void IntervalSet::finalise() {
    before.assign(iv.size() + 1, 0);
    for (size_t i = 0; i < iv.size(); i++) before[i + 1] = before[i] + ((int64_t)iv[i].second - iv[i].first + 1);
}

void IntervalSet::normalise() {
    std::sort(iv.begin(), iv.end());
    size_t w = 0;
    for (size_t i = 0; i < iv.size(); i++) {
        if (iv[i].second < iv[i].first) continue;
        if (w > 0 && (int64_t)iv[i].first <= (int64_t)iv[w - 1].second + 1) iv[w - 1].second = std::max(iv[w - 1].second, iv[i].second);
        else iv[w++] = iv[i];
    }
    iv.resize(w);
    finalise();
}

int32_t IntervalSet::kth(int64_t k) const {
    // last i with before[i] <= k
    size_t lo = 0, hi = iv.size();
    while (hi - lo > 1) {
        const size_t mid = (lo + hi) >> 1;
        if (before[mid] <= k) lo = mid; else hi = mid;
    }
    return (int32_t)(iv[lo].first + (k - before[lo]));
}

IntervalSet IntervalSet::unite(const IntervalSet& a, const IntervalSet& b) {
    IntervalSet r;
    r.iv.reserve(a.iv.size() + b.iv.size());
    r.iv.insert(r.iv.end(), a.iv.begin(), a.iv.end());
    r.iv.insert(r.iv.end(), b.iv.begin(), b.iv.end());
    r.normalise();
    return r;
}

IntervalSet IntervalSet::intersect(const IntervalSet& a, const IntervalSet& b) {
    IntervalSet r;
    size_t i = 0, j = 0;
    while (i < a.iv.size() && j < b.iv.size()) {
        const int32_t lo = std::max(a.iv[i].first, b.iv[j].first), hi = std::min(a.iv[i].second, b.iv[j].second);
        if (lo <= hi) r.iv.emplace_back(lo, hi);
        if (a.iv[i].second < b.iv[j].second) i++; else j++;
    }
    r.normalise();
    return r;
}

IntervalSet IntervalSet::minus(const IntervalSet& a, const IntervalSet& b) {
    IntervalSet r;
    size_t j = 0;
    for (size_t i = 0; i < a.iv.size(); i++) {
        int64_t cur = a.iv[i].first;
        const int64_t end = a.iv[i].second;
        while (j < b.iv.size() && b.iv[j].second < cur) j++;
        size_t k = j;
        while (cur <= end) {
            if (k >= b.iv.size() || b.iv[k].first > end) { r.iv.emplace_back((int32_t)cur, (int32_t)end); break; }
            if (b.iv[k].first > cur) r.iv.emplace_back((int32_t)cur, b.iv[k].first - 1);
            cur = (int64_t)b.iv[k].second + 1;
            k++;
        }
    }
    r.normalise();
    return r;
}

// src/SimulateHaplotypes.java:143-218 on interval sets. Quirk kept: a repeat class with no interval at all ends
// up as the one-element set {0}, because the Java de-duplication loop leaves positions_last=0 on an empty,
// zero-initialised array (:170-174).
void PositionSets::build(const Model& m, const uint8_t* reference, int32_t L, int max_distance) {
    IntervalSet raw[4];
    bool any[4] = {false, false, false, false};
    auto add = [&](int type, int64_t s, int64_t e) {
        int64_t from = s - max_distance, to = e + max_distance;
        if (from < 0) from = 0;
        if (to >= L) to = (int64_t)L - 1;
        if (from <= to) { raw[type].iv.emplace_back((int32_t)from, (int32_t)to); any[type] = true; }
    };
    int64_t appended[4] = {0, 0, 0, 0};
    for (const auto& r : m.repeats) {
        if (r[0] < 0 || r[0] >= N_INTERVAL_TYPES) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: repeat interval type " + std::to_string(r[0]));
        const int type = r[0];
        if (type > 3) {  // positions[4] is legal in the Java but later overwritten; only the capacity check applies
            continue;
        }
        add(type, r[1], r[2]);
        int64_t from = std::max<int64_t>(0, (int64_t)r[1] - max_distance), to = std::min<int64_t>((int64_t)L - 1, (int64_t)r[2] + max_distance);
        if (to >= from) appended[type] += to - from + 1;
    }
    for (const auto& s : m.segdups) {
        add(3, s[0], s[1]);
        int64_t from = std::max<int64_t>(0, (int64_t)s[0] - max_distance), to = std::min<int64_t>((int64_t)L - 1, (int64_t)s[1] + max_distance);
        if (to >= from) appended[3] += to - from + 1;
    }
    for (int t = 0; t < 4; t++) {
        // the Java appends every dilated position (duplicates included) into an int[L] before de-duplicating
        if (appended[t] > (int64_t)L) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: positions[" + std::to_string(t) + "]");
        if (!any[t] && L > 0) raw[t].iv.emplace_back(0, 0);
        raw[t].normalise();
    }
    const IntervalSet both = IntervalSet::unite(IntervalSet::intersect(raw[0], raw[1]), raw[2]);
    const IntervalSet rep = IntervalSet::minus(raw[0], both);
    const IntervalSet sat = IntervalSet::minus(raw[1], both);
    const IntervalSet seg = IntervalSet::minus(IntervalSet::minus(IntervalSet::minus(raw[3], rep), sat), both);
    IntervalSet all;
    all.iv.emplace_back(0, L - 1);
    all.normalise();
    const IntervalSet none = IntervalSet::minus(all, IntervalSet::unite(IntervalSet::unite(IntervalSet::unite(rep, sat), both), seg));
    // positions whose character is not in ACGTacgt are dropped (:207-217)
    IntervalSet bad;
    {
        bool lut[256] = {};
        for (const char* p = "ACGTacgt"; *p; p++) lut[(unsigned char)*p] = true;
        int32_t run = -1;
        for (int32_t i = 0; i < L; i++) {
            if (!lut[reference[i]]) { if (run < 0) run = i; }
            else if (run >= 0) { bad.iv.emplace_back(run, i - 1); run = -1; }
        }
        if (run >= 0) bad.iv.emplace_back(run, L - 1);
        bad.finalise();
    }
    set[0] = IntervalSet::minus(rep, bad);
    set[1] = IntervalSet::minus(sat, bad);
    set[2] = IntervalSet::minus(both, bad);
    set[3] = IntervalSet::minus(seg, bad);
    set[4] = IntervalSet::minus(none, bad);
}

}  // namespace sp
