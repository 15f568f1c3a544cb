// BuildModel (src/BuildModel.java:21-47) — the builder of the statistics files, repeatIntervals.txt
// and segdups.txt that SimulateHaplotypes.loadPositions/deserializeStatistics read.
//
// What runs where. The annotation tracks (a few million rows, sorted merges with carried state) and the gnomAD-SV record
// scan (a cursor over the repeat intervals that only moves forward, :981-1006, so a record's class depends on the records
// before it) are host code. The per-record counting and the 24 allele-frequency sorts are the data-parallel part:
//   count_kernel      one thread per record: (SV type, interval class, length bin) and (SV type, chromosome, 100 kb bin)
//                     histograms by integer atomics (exact: the Java's `++` on doubles counts integers below 2^53)
//   cub segmented radix sort of the 6 populations x 4 SV types allele-frequency arrays (Arrays.sort(double[]), :1079) on
//                     order-preserving 64-bit keys
// Everything after the counts follows the Java's floating-point operation order on the host (sums in index order; the
// allele-frequency histogram written over its own sorted input, :1086-1091), because the outputs are printed with
// Double.toString and compared byte for byte.
#include "common.cuh"
#include "textio.h"
#include <cub/device/device_segmented_radix_sort.cuh>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <strings.h>

namespace sp {

namespace {

constexpr int MAX_DISTANCE_REPEATMASKER = 25, MIN_OVERLAP_TRF = 10, MIN_OVERLAP_INTERVALS = 25;                  // :12-14
constexpr int MAX_DISTANCE_SR = 25, MAX_DISTANCE_PE = 250, MAX_DISTANCE_OTHER = 1000, MAX_DISTANCE_SEGDUPS = 25; // :15-18
constexpr int N_CHR = 24, N_AUTOSOMES = 22, POSITION_QUANTUM = 100000, POSITION_RADIUS = 5;                      // :628-637
constexpr int N_TYPES = 4, N_CLASSES = 5, N_POPULATIONS = 6;                                                     // :837,:842,:854
constexpr int MIN_SV_LENGTH = 50, MAX_SV_LENGTH = 1000000, LENGTH_QUANTUM = 50, AF_BINS = 10000, LENGTH_RADIUS = 4;  // :838-841,:871
constexpr int N_LENGTH_BINS = 1 + (MAX_SV_LENGTH - MIN_SV_LENGTH + 1) / LENGTH_QUANTUM;                          // :921
constexpr double MIN_FREQUENCY = 0.01;                                                                           // :873
const char* const TYPE_NAMES[N_TYPES] = {"DEL", "DUP", "INS", "INV"};
const char* const AF_KEYS[N_POPULATIONS] = {";AF=", ";AFR_AF=", ";AMR_AF=", ";EAS_AF=", ";EUR_AF=", ";OTH_AF="};   // :844
const char* const AF_FILES[N_POPULATIONS] = {"statistics_alleleFrequency_ALL", "statistics_alleleFrequency_AFR", "statistics_alleleFrequency_AMR",
                                             "statistics_alleleFrequency_EAS", "statistics_alleleFrequency_EUR", "statistics_alleleFrequency_OTH"};
const char SHORT_PERIOD[] = "ARTIFICIAL_SHORT_PERIOD";                                                           // :332
__host__ __device__ inline int merge_class(int a, int b) { return a == b && a != 2 ? a : 2; }                    // ADD_TYPE, :83

struct Text {   // a view into a slurped file
    const char* b;
    const char* e;
    int size() const { return (int)(e - b); }
    std::string str() const { return std::string(b, e); }
    bool is(const char* s) const { const size_t n = strlen(s); return (size_t)(e - b) == n && memcmp(b, s, n) == 0; }
    bool is_nocase(const char* s) const { const size_t n = strlen(s); return (size_t)(e - b) == n && strncasecmp(b, s, n) == 0; }
};

[[noreturn]] void java(const std::string& what) { throw Error(SP_ERR_JAVA_EXCEPTION, what); }

// String.split(sep) with limit 0
void fields(Text line, char sep, std::vector<Text>& out) {
    out.clear();
    const char* q = line.b;
    for (;;) {
        const char* c = (const char*)memchr(q, sep, (size_t)(line.e - q));
        if (!c) { out.push_back({q, line.e}); break; }
        out.push_back({q, c});
        q = c + 1;
    }
    while (out.size() > 1 && out.back().b == out.back().e) out.pop_back();
}
Text at(const std::vector<Text>& f, size_t i) {
    if (i >= f.size()) java("ArrayIndexOutOfBoundsException: Index " + std::to_string(i) + " out of bounds for length " + std::to_string(f.size()));
    return f[i];
}
// String.indexOf(pattern, from)
int find(Text s, const char* pat, int from = 0) {
    if (from < 0) from = 0;
    const int n = s.size(), m = (int)strlen(pat);
    for (int i = from; i + m <= n; i++) if (memcmp(s.b + i, pat, (size_t)m) == 0) return i;
    return -1;
}
// info.substring(p + keylen, info.indexOf(";", p + keylen))
Text value_after(Text info, int p, int keylen) {
    const int a = p + keylen, q = find(info, ";", a);
    if (a < 0 || q > info.size() || a > q) java("StringIndexOutOfBoundsException: begin " + std::to_string(a) + ", end " + std::to_string(q) + ", length " + std::to_string(info.size()));
    return {info.b + a, info.b + q};
}
std::string lower(Text t) { std::string s = t.str(); for (auto& ch : s) if (ch >= 'A' && ch <= 'Z') ch = (char)(ch + 32); return s; }
bool same_nocase(const std::string& a, const std::string& b) { return a.size() == b.size() && strcasecmp(a.c_str(), b.c_str()) == 0; }

struct Annotation {   // one row of repeatTrack (:58-70) / one Tuple (:331-396)
    std::string chr, repeat, family;
    int32_t first, last, rep_first, rep_last, forward;
    bool satellite;
};
struct Interval { std::string chr; int32_t cls, first, last; };   // repeatIntervals (:77-81); segdups use first/last only

// :715-729
int chromosome_index(Text s) {
    char c;
    if (s.size() > 3) {
        if (strncasecmp(s.b, "chr", 3) != 0 || s.size() > 5) return -1;
        c = s.b[3];
    } else {
        if (s.size() == 0) java("StringIndexOutOfBoundsException: empty chromosome name");
        c = s.b[0];
    }
    if (c == 'M' || c == 'm') return -1;
    if (c == 'X' || c == 'x') return N_CHR - 2;
    if (c == 'Y' || c == 'y') return N_CHR - 1;
    const int id = (s.size() > 3 ? parse_int(s.b + 3, s.e, "chromosome") : parse_int(s.b, s.e, "chromosome")) - 1;
    if (id < 0 || id >= N_CHR) java("ArrayIndexOutOfBoundsException: chromosome " + s.str());
    return id;
}

// :101-192
std::vector<Annotation> load_repeatmasker(const char* path, int declared_rows) {
    const std::vector<char> buf = slurp(path);
    Lines lines(buf.data(), buf.data() + buf.size());
    std::vector<Annotation> rows;
    std::vector<Text> f;
    const char *b, *e;
    while (lines.next(b, e)) {
        fields({b, e}, ',', f);
        const Text chr = at(f, 4);
        if (chr.size() > 5 || chr.is_nocase("chrM")) break;
        if ((int)rows.size() >= declared_rows) java("ArrayIndexOutOfBoundsException: more RepeatMasker rows than declared");
        Annotation a;
        a.chr = chr.str();
        a.first = parse_int(at(f, 5).b, at(f, 5).e, "genoStart") - 1;
        a.last = parse_int(at(f, 6).b, at(f, 6).e, "genoEnd") - 1;
        if (at(f, 8).size() == 0) java("StringIndexOutOfBoundsException: strand");
        a.forward = f[8].b[0] == '+' ? 1 : 0;
        if (a.forward) { a.rep_first = parse_int(at(f, 11).b, at(f, 11).e, "repStart") - 1; a.rep_last = parse_int(at(f, 12).b, at(f, 12).e, "repEnd") - 1; }
        else { a.rep_last = parse_int(at(f, 12).b, at(f, 12).e, "repEnd") - 1; a.rep_first = parse_int(at(f, 13).b, at(f, 13).e, "repLeft") - 1; }
        a.repeat = lower(at(f, 9));
        a.family = lower(at(f, 10));
        a.satellite = a.family.find("sat") != std::string::npos || a.family.find("simple") != std::string::npos || a.family.find("low_complexity") != std::string::npos;
        rows.push_back(std::move(a));
    }
    // fragments of one repeat copy a few bases apart (:159-190)
    const int d = MAX_DISTANCE_REPEATMASKER;
    size_t keep = 0;
    for (size_t i = 1; i < rows.size(); i++) {
        Annotation& head = rows[keep];
        const Annotation& cur = rows[i];
        const bool consensus_close = cur.forward ? std::abs(cur.rep_first - head.rep_last) <= d : std::abs(cur.rep_last - head.rep_first) <= d;
        if (cur.repeat == head.repeat && cur.family == head.family && cur.forward == head.forward && same_nocase(cur.chr, head.chr) &&
            std::abs(cur.first - head.last) <= d && (cur.satellite || consensus_close)) {
            head.last = cur.last;
            head.rep_first = std::min(head.rep_first, cur.rep_first);
            head.rep_last = std::max(head.rep_last, cur.rep_last);
        } else {
            keep++;
            if (keep != i) rows[keep] = rows[i];
        }
    }
    if (!rows.empty()) rows.resize(keep + 1);
    return rows;
}

// :220-323: TRF rows, overlapping ones merged, then the union with the RepeatMasker rows
void add_trf(const char* path, int declared_rows, std::vector<Annotation>& track) {
    const std::vector<char> buf = slurp(path);
    Lines lines(buf.data(), buf.data() + buf.size());
    std::vector<Annotation> trf;
    std::vector<Text> f;
    const char *b, *e;
    while (lines.next(b, e)) {
        fields({b, e}, ',', f);
        if ((int)trf.size() >= declared_rows) java("ArrayIndexOutOfBoundsException: more TRF rows than declared");
        Annotation a;
        a.chr = at(f, 0).str();
        a.first = parse_int(at(f, 1).b, at(f, 1).e, "chromStart") - 1;
        a.last = parse_int(at(f, 2).b, at(f, 2).e, "chromEnd") - 1;
        a.rep_first = a.rep_last = -1; a.forward = 1; a.satellite = true; a.repeat = SHORT_PERIOD;
        trf.push_back(std::move(a));
    }
    if (trf.empty()) java("ArrayIndexOutOfBoundsException: Index 0 out of bounds (empty TRF track)");
    size_t keep = 0;
    for (size_t i = 1; i < trf.size(); i++) {
        if (same_nocase(trf[i].chr, trf[keep].chr) && trf[i].first <= trf[keep].last - MIN_OVERLAP_TRF) trf[keep].last = std::max(trf[keep].last, trf[i].last);
        else { keep++; if (keep != i) trf[keep] = trf[i]; }
    }
    trf.resize(keep + 1);
    track.insert(track.end(), trf.begin(), trf.end());
    std::stable_sort(track.begin(), track.end(), [](const Annotation& x, const Annotation& y) {   // Tuple.compareTo, :380-390
        const int c = x.chr.compare(y.chr);
        if (c != 0) return c < 0;
        if (x.first != y.first) return x.first < y.first;
        return x.last < y.last;
    });
    keep = 0;
    for (size_t i = 1; i < track.size(); i++) {
        Annotation& head = track[keep];
        const Annotation& cur = track[i];
        const bool head_short = head.repeat == SHORT_PERIOD, cur_short = cur.repeat == SHORT_PERIOD;
        if ((head_short || cur_short) && cur.chr == head.chr && cur.first <= head.last - MIN_OVERLAP_TRF) {   // Tuple.merge, :352-374
            bool grown = false;
            if (cur.first < head.first) { head.first = cur.first; grown = true; }
            if (cur.last > head.last) { head.last = cur.last; grown = true; }
            head.satellite = head.satellite || cur.satellite;
            if (grown) {
                head.rep_first = head.rep_last = -1;
                if (!head_short && cur_short) { head.repeat = SHORT_PERIOD; head.family.clear(); head.forward = 1; }
            }
        } else {
            keep++;
            if (keep != i) std::swap(track[keep], track[i]);
        }
    }
    track.resize(keep + 1);
}

// :401-425
std::vector<Interval> repeat_intervals(const std::vector<Annotation>& track) {
    if (track.size() < 2) java("ArrayIndexOutOfBoundsException: Index 0 out of bounds for length 0 (repeat track with fewer than two rows)");
    const size_t capacity = track.size() - 1;   // new int[lastRepeatTrack][3]
    std::vector<Interval> out;
    out.push_back({track[0].chr, track[0].satellite ? 1 : 0, track[0].first, track[0].last});
    for (size_t i = 1; i < track.size(); i++) {
        Interval& cur = out.back();
        // the Java compares against chromosomeIDs[lastRepeatInterval], the TRACK row with the interval's index (:413)
        if (track[i].chr == track[out.size() - 1].chr && track[i].first <= cur.last - MIN_OVERLAP_INTERVALS) {
            cur.last = std::max(cur.last, track[i].last);
            cur.cls = merge_class(cur.cls, track[i].satellite ? 1 : 0);
        } else {
            if (out.size() >= capacity) java("ArrayIndexOutOfBoundsException: repeatIntervals");
            out.push_back({track[i].chr, track[i].satellite ? 1 : 0, track[i].first, track[i].last});
        }
    }
    return out;
}

// :506-541
std::vector<Interval> load_segdups(const char* path, int declared_rows) {
    const std::vector<char> buf = slurp(path);
    Lines lines(buf.data(), buf.data() + buf.size());
    std::vector<Interval> rows;
    std::vector<Text> f;
    const char *b, *e;
    lines.next(b, e);   // header
    while (lines.next(b, e)) {
        fields({b, e}, '\t', f);
        if ((int)rows.size() >= declared_rows) java("ArrayIndexOutOfBoundsException: more segdup rows than declared");
        rows.push_back({at(f, 1).str(), 0, parse_int(at(f, 2).b, at(f, 2).e, "chromStart") - 1, parse_int(at(f, 3).b, at(f, 3).e, "chromEnd") - 2});
    }
    if ((int)rows.size() != declared_rows) java("NullPointerException: fewer segdup rows than declared (Arrays.sort over null entries)");
    std::stable_sort(rows.begin(), rows.end(), [](const Interval& x, const Interval& y) {
        const int c = x.chr.compare(y.chr);
        if (c != 0) return c < 0;
        if (x.first != y.first) return x.first < y.first;
        return x.last < y.last;
    });
    size_t keep = 0;   // note: no chromosome test in the Java's merge (:530-533)
    for (size_t i = 1; i < rows.size(); i++) {
        if (rows[i].first > rows[keep].last + MAX_DISTANCE_SEGDUPS) { keep++; if (keep != i) rows[keep] = rows[i]; }
        else rows[keep].last = std::max(rows[keep].last, rows[i].last);
    }
    if (!rows.empty()) rows.resize(keep + 1);
    return rows;
}

struct Karyotype { int32_t length[N_CHR], p_arm[N_CHR]; };

// :666-708
Karyotype load_karyotype(const char* fai, const char* cytobands) {
    Karyotype k;
    memset(&k, 0, sizeof k);
    std::vector<Text> f;
    const char *b, *e;
    {
        const std::vector<char> buf = slurp(fai);
        Lines lines(buf.data(), buf.data() + buf.size());
        while (lines.next(b, e)) {
            fields({b, e}, '\t', f);
            const int i = chromosome_index(at(f, 0));
            if (i >= 0) k.length[i] = parse_int(at(f, 1).b, at(f, 1).e, "fai length");
        }
    }
    {
        const std::vector<char> buf = slurp(cytobands);
        Lines lines(buf.data(), buf.data() + buf.size());
        while (lines.next(b, e)) {
            fields({b, e}, '\t', f);
            const int i = chromosome_index(at(f, 0));
            if (i < 0) continue;
            const int end = parse_int(at(f, 2).b, at(f, 2).e, "band end");
            if (at(f, 3).size() == 0) java("StringIndexOutOfBoundsException: band name");
            if (f[3].b[0] == 'p' && end > k.p_arm[i]) k.p_arm[i] = end;
        }
    }
    return k;
}

// What the device needs of one accepted gnomAD-SV record.
struct Counted { int32_t length_slot; int32_t position_slot; };

__global__ void count_kernel(const Counted* __restrict__ rec, int64_t n, uint32_t* __restrict__ length_counts, uint32_t* __restrict__ position_counts) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const Counted r = rec[i];
        atomicAdd(&length_counts[r.length_slot], 1u);
        atomicAdd(&position_counts[r.position_slot], 1u);
    }
}

// IEEE doubles -> unsigned keys in the order of Double.compare (what Arrays.sort(double[]) uses), and back
__global__ void af_keys_kernel(uint64_t* __restrict__ v, int64_t n, int back) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t x = v[i];
        v[i] = back ? ((x >> 63) ? x ^ 0x8000000000000000ull : ~x) : ((x >> 63) ? ~x : x ^ 0x8000000000000000ull);
    }
}

// the carried-sum moving average of :1028-1039 and :1042-1055 on counts
void smooth(std::vector<double>& a, int radius, std::vector<double>& scratch) {
    const int n = (int)a.size(), window = 2 * radius + 1;
    if (n < window + 1) java("ArrayIndexOutOfBoundsException: array shorter than the smoothing window");
    double sum = 0.0;
    for (int k = 0; k < window; k++) sum += a[k];
    for (int k = radius; k < n - radius; k++) {
        scratch[k] = sum / window;
        if (k + radius + 1 < n) sum += a[k + radius + 1] - a[k - radius];
    }
    for (int k = radius; k < n - radius; k++) a[k] = scratch[k];
}

// the mass of one source bin spread over chr1's bins [from..to] at `per_bp` (:771-780, :787-794)
void spread(std::vector<double>& out, int from, int to, double per_bp) {
    const int Q = POSITION_QUANTUM, b0 = from / Q, b1 = to / Q;
    if (b0 < 0 || b1 >= (int)out.size()) java("ArrayIndexOutOfBoundsException: projected bin");
    if (b0 == b1) { out[b0] += per_bp * (to - from + 1); return; }
    out[b0] += per_bp * ((b0 + 1) * Q - from);
    for (int h = b0 + 1; h < b1; h++) out[h] += per_bp * Q;
    out[b1] += per_bp * (to + 1 - b1 * Q);
}

// cumulative rows are long runs of one value: format a value once per run
void put(std::string& s, double d) {
    static thread_local uint64_t last_bits = 0;
    static thread_local std::string last_text;
    uint64_t bits;
    memcpy(&bits, &d, 8);
    if (last_text.empty() || bits != last_bits) { last_text = java_double_to_string(d); last_bits = bits; }
    s += last_text; s += ',';
}
void write_file(const std::string& path, const std::string& text) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) java("FileNotFoundException: " + path);
    const bool ok = fwrite(text.data(), 1, text.size(), f) == text.size();
    if (fclose(f) != 0 || !ok) throw Error(SP_ERR, "short write: " + path);
}

}  // namespace

struct BuildModelTiming { double parse_s = 0, device_ms = 0, finish_s = 0; int64_t records = 0, allele_frequencies = 0; };
static BuildModelTiming g_last_timing;   // per process; reported by sp_build_model_timing

void build_model(sp_ctx* c, const sp_build_model_args* a) {
    if (!a || !a->repeatmasker_file || !a->trf_file || !a->segdups_file || !a->reference_fai || !a->reference_cytobands || !a->gnomad_vcf || !a->output_dir)
        throw Error(SP_ERR_ARG, "BuildModel: missing argument");
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<Annotation> track = load_repeatmasker(a->repeatmasker_file, a->repeatmasker_rows);
    add_trf(a->trf_file, a->trf_rows, track);
    const std::vector<Interval> repeats = repeat_intervals(track);
    track.clear(); track.shrink_to_fit();
    const std::vector<Interval> segdups = load_segdups(a->segdups_file, a->segdups_rows);
    const Karyotype kar = load_karyotype(a->reference_fai, a->reference_cytobands);

    // flattened (type, chromosome) -> first position bin
    int64_t position_off[N_TYPES][N_CHR];
    int32_t position_bins[N_CHR];
    int64_t n_position = 0;
    for (int j = 0; j < N_CHR; j++) position_bins[j] = (kar.length[j] + POSITION_QUANTUM) / POSITION_QUANTUM;
    for (int i = 0; i < N_TYPES; i++) for (int j = 0; j < N_CHR; j++) { position_off[i][j] = n_position; n_position += position_bins[j]; }

    // the record scan, :933-1020
    std::vector<Counted> counted;
    std::vector<double> af[N_POPULATIONS][N_TYPES];
    int32_t n_chr1 = 0, n_frequent_chr1 = 0;
    {
        const std::vector<char> buf = slurp(a->gnomad_vcf);
        Lines lines(buf.data(), buf.data() + buf.size());
        std::vector<Text> f;
        const char *b, *e;
        size_t rk = 0, sk = 0;   // the two forward-only cursors k and x
        std::string chr;
        const int64_t R = (int64_t)repeats.size(), S = (int64_t)segdups.size();
        while (lines.next(b, e)) {
            if (b == e) java("StringIndexOutOfBoundsException: index 0, length 0 (empty line)");
            if (*b == '#') continue;
            fields({b, e}, '\t', f);
            const Text contig = f[0];
            if (contig.size() > 5 || contig.is_nocase("chrM")) break;
            if (a->only_pass && !at(f, 6).is_nocase("PASS") && !f[6].is_nocase("MULTIALLELIC")) continue;
            const Text info = at(f, 7);
            int p = find(info, "SVLEN=");
            if (p < 0) continue;
            const Text svlen = value_after(info, p, 6);
            int32_t length = parse_int(svlen.b, svlen.e, "SVLEN");
            if (length < 0) length = (int32_t)(0u - (uint32_t)length);
            if (length < MIN_SV_LENGTH) continue;
            p = find(info, "SVTYPE=");
            if (p < 0) continue;
            const Text type_name = value_after(info, p, 7);
            int type = -1;
            for (int t = 0; t < N_TYPES; t++) if (type_name.is(TYPE_NAMES[t])) type = t;
            if (type < 0) continue;
            chr.assign("chr"); chr.append(contig.b, contig.e);
            const bool is_chr1 = contig.is_nocase("1");
            if (is_chr1) n_chr1++;
            const int32_t position = parse_int(at(f, 1).b, at(f, 1).e, "POS");
            p = find(info, "EVIDENCE");
            const Text evidence = value_after(info, p, 8);
            const int reach = find(evidence, "SR") >= 0 ? MAX_DISTANCE_SR : (find(evidence, "PE") >= 0 ? MAX_DISTANCE_PE : MAX_DISTANCE_OTHER);
            while ((int64_t)rk < R && (repeats[rk].chr.compare(chr) < 0 || repeats[rk].last + reach < position)) rk++;
            int cls = 4;
            if ((int64_t)rk < R && repeats[rk].chr.compare(chr) <= 0 && repeats[rk].first <= position + reach) {
                cls = repeats[rk].cls;
                for (int64_t h = (int64_t)rk - 1; h >= 0 && repeats[h].last + reach >= position; h--) cls = merge_class(cls, repeats[h].cls);   // no chromosome test, as :996-999
                for (int64_t h = (int64_t)rk + 1; h < R && repeats[h].first <= position + reach; h++) cls = merge_class(cls, repeats[h].cls);
            }
            if (cls == 4) {
                while ((int64_t)sk < S && (segdups[sk].chr.compare(chr) < 0 || segdups[sk].last + MAX_DISTANCE_SEGDUPS < position)) sk++;
                if ((int64_t)sk < S && segdups[sk].chr.compare(chr) <= 0 && segdups[sk].first <= position + MAX_DISTANCE_SEGDUPS) cls = 3;
            }
            const int length_bin = std::min((length - MIN_SV_LENGTH) / LENGTH_QUANTUM, N_LENGTH_BINS - 1);
            if (!contig.is_nocase("X") && !contig.is_nocase("Y")) {
                bool frequent = false;
                for (int k = 0; k < N_POPULATIONS; k++) {
                    const int keylen = (int)strlen(AF_KEYS[k]);
                    const Text v = value_after(info, find(info, AF_KEYS[k]), keylen);
                    const double x = parse_double(v.b, v.e, AF_KEYS[k]);
                    if (is_chr1 && x >= MIN_FREQUENCY) frequent = true;
                    af[k][type].push_back(x);
                }
                if (frequent) n_frequent_chr1++;
            }
            const int chr_id = chromosome_index(contig);
            if (chr_id < 0) java("ArrayIndexOutOfBoundsException: Index -1 out of bounds for length 24");
            if (position < 0 || position / POSITION_QUANTUM >= position_bins[chr_id]) java("ArrayIndexOutOfBoundsException: position bin of POS " + std::to_string(position));
            counted.push_back({(type * N_CLASSES + cls) * N_LENGTH_BINS + length_bin, (int32_t)(position_off[type][chr_id] + position / POSITION_QUANTUM)});
        }
    }
    if (a->verbose >= 0) fprintf(stderr, "Total SVs in chr1: %d. With allele frequency >=%s: %d\n", n_chr1, java_double_to_string(MIN_FREQUENCY).c_str(), n_frequent_chr1);
    for (int k = 0; k < N_POPULATIONS; k++)
        for (int i = 0; i < N_TYPES; i++)
            if (af[k][i].empty()) java(std::string("ArrayIndexOutOfBoundsException: Index -1 out of bounds (no autosomal record of type ") + TYPE_NAMES[i] + ")");
    const auto t1 = std::chrono::steady_clock::now();

    // device: the two histograms and the 24 sorts
    const int64_t n = (int64_t)counted.size();
    const int64_t n_length = (int64_t)N_TYPES * N_CLASSES * N_LENGTH_BINS;
    int64_t seg_off[N_POPULATIONS * N_TYPES + 1];
    seg_off[0] = 0;
    for (int k = 0; k < N_POPULATIONS; k++) for (int i = 0; i < N_TYPES; i++) seg_off[k * N_TYPES + i + 1] = seg_off[k * N_TYPES + i] + (int64_t)af[k][i].size();
    const int64_t n_af = seg_off[N_POPULATIONS * N_TYPES];
    if (n_af > INT32_MAX) throw Error(SP_ERR_ARG, "BuildModel: more than 2^31-1 allele frequencies");
    std::vector<double> af_flat((size_t)n_af);
    for (int k = 0; k < N_POPULATIONS; k++) for (int i = 0; i < N_TYPES; i++) std::copy(af[k][i].begin(), af[k][i].end(), af_flat.begin() + seg_off[k * N_TYPES + i]);
    DevBuf<Counted> d_rec;
    DevBuf<uint32_t> d_counts;
    DevBuf<uint64_t> d_af, d_af2;
    DevBuf<int64_t> d_seg;
    DevBuf<uint8_t> d_tmp;
    d_rec.reserve((size_t)std::max<int64_t>(n, 1), 0.0);
    d_counts.reserve((size_t)(n_length + n_position), 0.0);
    d_af.reserve((size_t)n_af, 0.0); d_af2.reserve((size_t)n_af, 0.0);
    d_seg.reserve(N_POPULATIONS * N_TYPES + 1, 0.0);
    cudaStream_t st = c->stream;
    size_t tmp_bytes = 0;
    cub::DoubleBuffer<uint64_t> keys(d_af.p, d_af2.p);
    SP_CUDA(cub::DeviceSegmentedRadixSort::SortKeys(nullptr, tmp_bytes, keys, (int)n_af, N_POPULATIONS * N_TYPES, d_seg.p, d_seg.p + 1, 0, 64, st));
    d_tmp.reserve(tmp_bytes + 16, 0.0);
    cudaEvent_t ev0, ev1;
    SP_CUDA(cudaEventCreate(&ev0)); SP_CUDA(cudaEventCreate(&ev1));
    std::vector<uint32_t> counts((size_t)(n_length + n_position));
    float ms = 0.f;
    try {
        SP_CUDA(cudaMemcpyAsync(d_rec.p, counted.data(), (size_t)n * sizeof(Counted), cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(d_af.p, af_flat.data(), (size_t)n_af * 8, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(d_seg.p, seg_off, sizeof seg_off, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaEventRecord(ev0, st));
        SP_CUDA(cudaMemsetAsync(d_counts.p, 0, (size_t)(n_length + n_position) * 4, st));
        const int grid = c->sm_count * 8;
        if (n > 0) { count_kernel<<<grid, 256, 0, st>>>(d_rec.p, n, d_counts.p, d_counts.p + n_length); c->launches++; }
        af_keys_kernel<<<grid, 256, 0, st>>>(d_af.p, n_af, 0); c->launches++;
        SP_CUDA(cub::DeviceSegmentedRadixSort::SortKeys(d_tmp.p, tmp_bytes, keys, (int)n_af, N_POPULATIONS * N_TYPES, d_seg.p, d_seg.p + 1, 0, 64, st));
        c->launches++;
        af_keys_kernel<<<grid, 256, 0, st>>>(keys.Current(), n_af, 1); c->launches++;
        SP_CUDA(cudaGetLastError());
        SP_CUDA(cudaEventRecord(ev1, st));
        SP_CUDA(cudaMemcpyAsync(counts.data(), d_counts.p, counts.size() * 4, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaMemcpyAsync(af_flat.data(), keys.Current(), (size_t)n_af * 8, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        SP_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    } catch (...) { cudaEventDestroy(ev0); cudaEventDestroy(ev1); throw; }
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    const auto t2 = std::chrono::steady_clock::now();

    // counts -> smoothed -> probabilities, :1024-1075
    std::vector<double> scratch((size_t)std::max<int>(N_LENGTH_BINS, position_bins[0]));
    std::vector<double> length_p[N_TYPES][N_CLASSES];
    double class_p[N_TYPES][N_CLASSES], type_p[N_TYPES];
    for (int i = 0; i < N_TYPES; i++)
        for (int j = 0; j < N_CLASSES; j++) {
            std::vector<double>& v = length_p[i][j];
            const uint32_t* src = counts.data() + (size_t)(i * N_CLASSES + j) * N_LENGTH_BINS;
            v.assign(src, src + N_LENGTH_BINS);
            smooth(v, LENGTH_RADIUS, scratch);
        }
    std::vector<double> per_position[N_TYPES][N_CHR];
    for (int i = 0; i < N_TYPES; i++)
        for (int j = 0; j < N_CHR; j++) {
            if (position_bins[j] > position_bins[0]) java("ArrayIndexOutOfBoundsException: a chromosome with more position bins than chr1");
            const uint32_t* src = counts.data() + n_length + position_off[i][j];
            per_position[i][j].assign(src, src + position_bins[j]);
            smooth(per_position[i][j], POSITION_RADIUS, scratch);
        }
    for (int i = 0; i < N_TYPES; i++)
        for (int j = 0; j < N_CLASSES; j++) {
            double sum = 0.0;
            for (double x : length_p[i][j]) sum += x;
            for (double& x : length_p[i][j]) x /= sum;
            class_p[i][j] = sum;
        }
    for (int i = 0; i < N_TYPES; i++) {
        double sum = 0.0;
        for (int j = 0; j < N_CLASSES; j++) sum += class_p[i][j];
        for (int j = 0; j < N_CLASSES; j++) class_p[i][j] /= sum;
        type_p[i] = sum;
    }
    {
        double sum = 0.0;
        for (int i = 0; i < N_TYPES; i++) sum += type_p[i];
        for (int i = 0; i < N_TYPES; i++) type_p[i] /= sum;
    }

    // allele-frequency histograms, :1076-1094. The Java writes bin j over element j of the sorted array it is still
    // reading, so a read at an index below j sees a fraction, not a frequency; kept.
    double af_minmax[N_POPULATIONS][N_TYPES][2];
    std::vector<double> af_p[N_POPULATIONS][N_TYPES];
    for (int k = 0; k < N_POPULATIONS; k++)
        for (int i = 0; i < N_TYPES; i++) {
            const int64_t cnt = seg_off[k * N_TYPES + i + 1] - seg_off[k * N_TYPES + i];
            std::vector<double>& v = af_p[k][i];
            size_t cap = AF_BINS;
            while ((int64_t)cap < cnt) cap <<= 1;
            v.assign(cap, 0.0);
            std::copy(af_flat.begin() + seg_off[k * N_TYPES + i], af_flat.begin() + seg_off[k * N_TYPES + i + 1], v.begin());
            const double lo = v[0], hi = v[(size_t)cnt - 1], quantum = (hi - lo) / (AF_BINS - 1);
            af_minmax[k][i][0] = lo; af_minmax[k][i][1] = hi;
            int64_t next = 0;
            for (int j = 0; j < AF_BINS; j++) {
                const double boundary = lo + j * quantum;
                double below = 0.0;
                while (next < cnt && v[(size_t)next] < boundary) { below++; next++; }
                v[(size_t)j] = below / (double)cnt;
            }
        }

    // positions projected on chr1, :740-802
    std::vector<double> position_p[N_TYPES];
    for (int j = 1; j < N_AUTOSOMES; j++)
        if (kar.p_arm[j] <= 0 || kar.p_arm[j] >= kar.length[j] || kar.p_arm[0] <= 0)
            throw Error(SP_ERR_ARG, "BuildModel: autosome " + std::to_string(j + 1) + " has no p-arm band or no length (the Java would divide by zero)");
    for (int i = 0; i < N_TYPES; i++) {
        std::vector<double>& out = position_p[i];
        out = per_position[i][0];
        const int p1 = kar.p_arm[0], l1 = kar.length[0];
        for (int j = 1; j < N_AUTOSOMES; j++) {
            const int pj = kar.p_arm[j], lj = kar.length[j];
            const double ratio_p = ((double)p1) / pj, ratio_q = ((double)(l1 - p1)) / (lj - pj);
            for (int k = 0; k < position_bins[j]; k++) {
                const int first = k * POSITION_QUANTUM, last = std::min((k + 1) * POSITION_QUANTUM, lj) - 1;
                const double count = per_position[i][j][k];
                int fp = -1, lp = -1, fq = -1, lq = -1;
                double mass_p = 0.0, mass_q = 0.0;
                if (last < pj) { fp = first; lp = last; mass_p = count; }
                else if (first < pj) {
                    fp = first; lp = pj - 1; fq = pj; lq = last;
                    mass_p = ((double)((lp - fp + 1) * count)) / POSITION_QUANTUM;
                    mass_q = ((double)((lq - fq + 1) * count)) / POSITION_QUANTUM;
                } else { fq = first; lq = last; mass_q = count; }
                if (fp != -1) {
                    const int from = (int)(fp * ratio_p), to = (int)(lp * ratio_p);
                    if (from < p1) spread(out, from, std::min(to, p1 - 1), mass_p / (to - from + 1));
                }
                if (fq != -1) {
                    const int from = p1 + (int)((fq - pj) * ratio_q), to = p1 + (int)((lq - pj) * ratio_q);
                    spread(out, from, std::min(to, l1 - 1), mass_q / (to - from + 1));
                }
            }
        }
        double sum = 0.0;
        for (double x : out) sum += x;
        for (double& x : out) x /= sum;
    }

    // cumulative, :1100-1132 (the allele-frequency arrays over their whole capacity, as the Java)
    for (int i = 1; i < N_TYPES; i++) type_p[i] += type_p[i - 1];
    for (int i = 0; i < N_TYPES; i++) for (int j = 1; j < N_CLASSES; j++) class_p[i][j] += class_p[i][j - 1];
    for (int i = 0; i < N_TYPES; i++) for (int j = 0; j < N_CLASSES; j++) for (size_t k = 1; k < length_p[i][j].size(); k++) length_p[i][j][k] += length_p[i][j][k - 1];
    for (int k = 0; k < N_POPULATIONS; k++) for (int i = 0; i < N_TYPES; i++) for (int j = 1; j < AF_BINS; j++) af_p[k][i][(size_t)j] += af_p[k][i][(size_t)j - 1];
    for (int i = 0; i < N_TYPES; i++) for (size_t j = 1; j < position_p[i].size(); j++) position_p[i][j] += position_p[i][j - 1];

    // files, :1135-1197, :446-458, :563-577
    const std::string dir = a->output_dir;
    std::string s;
    s = std::to_string(N_TYPES) + "\n";
    for (int i = 0; i < N_TYPES; i++) put(s, type_p[i]);
    s += '\n';
    write_file(dir + "/statistics_svType.txt", s);
    s = std::to_string(N_TYPES) + "," + std::to_string(N_CLASSES) + "\n";
    for (int i = 0; i < N_TYPES; i++) { for (int j = 0; j < N_CLASSES; j++) put(s, class_p[i][j]); s += '\n'; }
    write_file(dir + "/statistics_intervalType.txt", s);
    s = std::to_string(N_TYPES) + "," + std::to_string(N_CLASSES) + "," + std::to_string(N_LENGTH_BINS) + "\n";
    for (int i = 0; i < N_TYPES; i++) for (int j = 0; j < N_CLASSES; j++) { for (double x : length_p[i][j]) put(s, x); s += '\n'; }
    write_file(dir + "/statistics_length.txt", s);
    for (int k = 0; k < N_POPULATIONS; k++) {
        s = std::to_string(N_TYPES) + "," + std::to_string(AF_BINS) + "\n";
        for (int i = 0; i < N_TYPES; i++) {
            put(s, af_minmax[k][i][0]); put(s, af_minmax[k][i][1]);
            for (int j = 0; j < AF_BINS; j++) put(s, af_p[k][i][(size_t)j]);
            s += '\n';
        }
        write_file(dir + "/" + AF_FILES[k] + ".txt", s);
    }
    s = std::to_string(N_TYPES) + "," + std::to_string(position_p[0].size()) + "\n";
    for (int i = 0; i < N_TYPES; i++) { for (double x : position_p[i]) put(s, x); s += '\n'; }
    write_file(dir + "/statistics_position.txt", s);
    {
        size_t n1 = 0;
        while (n1 < repeats.size() && same_nocase(repeats[n1].chr, "chr1")) n1++;
        s = std::to_string(n1) + "\n";
        for (size_t i = 0; i < n1; i++) s += std::to_string(repeats[i].cls) + "," + std::to_string(repeats[i].first) + "," + std::to_string(repeats[i].last) + "\n";
        write_file(dir + "/repeatIntervals.txt", s);
        n1 = 0;
        while (n1 < segdups.size() && same_nocase(segdups[n1].chr, "chr1")) n1++;
        s = std::to_string(n1) + "\n";
        for (size_t i = 0; i < n1; i++) s += std::to_string(segdups[i].first) + "," + std::to_string(segdups[i].last) + "\n";
        write_file(dir + "/segdups.txt", s);
    }
    const auto t3 = std::chrono::steady_clock::now();
    g_last_timing.parse_s = std::chrono::duration<double>(t1 - t0).count();
    g_last_timing.device_ms = ms;
    g_last_timing.finish_s = std::chrono::duration<double>(t3 - t2).count();
    g_last_timing.records = n;
    g_last_timing.allele_frequencies = n_af;
}

}  // namespace sp

extern "C" {

int sp_build_model(sp_ctx* ctx, const sp_build_model_args* args) {
    SP_API_BEGIN
    sp::build_model(ctx, args);
    SP_API_END
}

int sp_build_model_timing(sp_ctx* ctx, double* parse_seconds, double* device_ms, double* finish_seconds, int64_t* n_records, int64_t* n_allele_frequencies) {
    SP_API_BEGIN
    if (parse_seconds) *parse_seconds = sp::g_last_timing.parse_s;
    if (device_ms) *device_ms = sp::g_last_timing.device_ms;
    if (finish_seconds) *finish_seconds = sp::g_last_timing.finish_s;
    if (n_records) *n_records = sp::g_last_timing.records;
    if (n_allele_frequencies) *n_allele_frequencies = sp::g_last_timing.allele_frequencies;
    SP_API_END
}

}  // extern "C"
