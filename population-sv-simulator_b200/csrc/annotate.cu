// SURVEY.md §8(f) rank 2 — AnnotateVCF (reference src/AnnotateVCF.java:96-205, helpers :209-510): adds
// REPEATS_START / REPEATS_END / REPEATS_FRACTION to every DEL/DUP/INS/INV record of callers' VCFs.
//
// The reference materialises loadPositions' five int[L] arrays (<= 22 GB, :16) and answers every record with a binary
// search plus a merge walk over positions. Here the five sets are row A3's sorted disjoint intervals with prefix counts
// (csrc/model.cu), resident in HBM, and a record is three rank queries:
//   "does set t hold a position in [a,b]"  = rank_t(b+1) - rank_t(a) > 0                      (getRepeatType, :437-457)
//   "how many positions of sets 0..3 in [a,b]" = sum of the same differences (the sets are disjoint by construction,
//   so the reference's sort + de-duplication never removes anything)                           (getRepeatSurface, :480-495)
// One thread per record (annotate_kernel); the host parses the text into seven integers per record and formats the
// three results. The VCF text handling follows the Java string calls literally (indexOf without '=', the CIEND guard
// of getField, String.split's dropped trailing fields, Integer.parseInt, Math.round).
#include "common.cuh"
#include "model.h"
#include "textio.h"
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>

namespace sp {

struct AnnotRecord {   // what the kernel needs of one VCF record
    int32_t position, end;       // end == INT_MAX: no END and no SVLEN (:174), repeatsEnd = 4 and fraction 0
    int32_t ci_pos[2], ci_end[2];
    int32_t has_length;          // 0: length == Integer.MAX_VALUE (insertions, or END/SVLEN missing)
    int32_t pad;
};
struct AnnotResult { int32_t repeats_start, repeats_end, surface, pad; };

struct AnnotateState {
    Model model;
    PositionSets sets;
    int32_t set_off[N_INTERVAL_TYPES + 1] = {0, 0, 0, 0, 0, 0};
    DevBuf<int32_t> d_first, d_last;
    DevBuf<int64_t> d_before;          // set t: entries [set_off[t] + t, set_off[t+1] + t]
    DevBuf<AnnotRecord> d_rec;
    DevBuf<AnnotResult> d_res;
    bool loaded = false;
    double kernel_ms = 0.0, load_s = 0.0;
    int64_t records_total = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    ~AnnotateState() { if (ev0) cudaEventDestroy(ev0); if (ev1) cudaEventDestroy(ev1); }
};

void load_reference_fasta(sp_ctx* c, const char* path);

void free_annotate(sp_ctx* c) {
    delete c->annot;
    c->annot = nullptr;
}

struct SetsDev {
    const int32_t* first;
    const int32_t* last;
    const int64_t* before;
    int32_t off[N_INTERVAL_TYPES + 1];
};

// number of positions of set t that are < x
__device__ __forceinline__ int64_t rank_below(const SetsDev& S, int t, int64_t x) {
    const int32_t a = S.off[t], n = S.off[t + 1] - a;
    int32_t lo = 0, hi = n;   // first interval whose last position is >= x
    while (lo < hi) {
        const int32_t mid = (lo + hi) >> 1;
        if ((int64_t)S.last[a + mid] < x) lo = mid + 1; else hi = mid;
    }
    int64_t r = S.before[a + t + lo];
    if (lo < n) { const int64_t f = S.first[a + lo]; if (x > f) r += x - f; }
    return r;
}
__device__ __forceinline__ int64_t count_in(const SetsDev& S, int t, int64_t a, int64_t b) {
    return a > b ? 0 : rank_below(S, t, b + 1) - rank_below(S, t, a);
}
// getRepeatType (:437-457)
__device__ __forceinline__ int32_t repeat_type(const SetsDev& S, int64_t first, int64_t last) {
    bool in[N_INTERVAL_TYPES];
#pragma unroll
    for (int t = 0; t < N_INTERVAL_TYPES; t++) in[t] = count_in(S, t, first, last) > 0;
    if (in[2] || (in[0] && in[1])) return 2;
    if (in[0]) return 0;
    if (in[1]) return 1;
    if (in[3]) return 3;
    return 4;
}
// Java int arithmetic wraps
__device__ __forceinline__ int32_t jadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }

__global__ void __launch_bounds__(256)
annotate_kernel(const AnnotRecord* __restrict__ rec, int64_t n, int32_t min_distance, int32_t max_distance, SetsDev S,
                AnnotResult* __restrict__ res) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const AnnotRecord r = rec[i];
    const int32_t lo_far = jadd(r.position, -max_distance), lo_near = jadd(r.position, -min_distance);
    const int32_t hi_far = jadd(r.position, max_distance), hi_near = jadd(r.position, min_distance);
    auto clamp = [&](int32_t& f, int32_t& l) {   // :177-180 and, with the same bounds around POS, :184-187
        if (f < lo_far) f = lo_far;
        if (f > lo_near) f = lo_near;
        if (l > hi_far) l = hi_far;
        if (l < hi_near) l = hi_near;
    };
    AnnotResult o;
    int32_t f = jadd(r.position, r.ci_pos[0]), l = jadd(r.position, r.ci_pos[1]);
    clamp(f, l);
    o.repeats_start = repeat_type(S, f, l);
    o.repeats_end = 4; o.surface = 0; o.pad = 0;
    if (r.end != INT_MAX) {
        f = jadd(r.end, r.ci_end[0]); l = jadd(r.end, r.ci_end[1]);
        clamp(f, l);
        o.repeats_end = repeat_type(S, f, l);
        if (r.has_length) {
            int64_t s = 0;
#pragma unroll
            for (int t = 0; t < 4; t++) s += count_in(S, t, r.position, r.end);
            o.surface = (int32_t)(s <= 1 ? 1 : s);   // :488 and the k+1 of :494 (one when nothing intersects, too)
        }
    }
    res[i] = o;
}

// ------------------------------------------------------------------------------------------------- host: text
namespace {

struct Span {
    const char* b;
    const char* e;
    int size() const { return (int)(e - b); }
};

int index_of(Span s, const char* what, int from = 0) {   // String.indexOf(String, fromIndex)
    const int n = s.size(), m = (int)strlen(what);
    if (from < 0) from = 0;
    if (from > n) return -1;
    const void* p = memmem(s.b + from, (size_t)(n - from), what, (size_t)m);
    return p ? (int)((const char*)p - s.b) : -1;
}
bool ieq(const char* b, int n, const char* lit) { return (int)strlen(lit) == n && strncasecmp(b, lit, (size_t)n) == 0; }

[[noreturn]] void throw_sioobe(int index, int length) {
    throw Error(SP_ERR_JAVA_EXCEPTION, "StringIndexOutOfBoundsException: index " + std::to_string(index) + ", length " + std::to_string(length));
}
char char_at(Span s, int i) {
    if (i < 0 || i >= s.size()) throw_sioobe(i, s.size());
    return s.b[i];
}
Span sub(Span s, int b, int e) {
    if (b < 0 || e > s.size() || b > e) throw_sioobe(b, s.size());
    return Span{s.b + b, s.b + e};
}

// getField (:209-220): false = the Java's null
bool get_field(Span info, const char* key_eq /* "KEY=" */, bool is_end, Span& out) {
    const int klen = (int)strlen(key_eq);
    int p = index_of(info, key_eq);
    if (p < 0) return false;
    if (is_end) {
        while (p >= 2 && ieq(info.b + p - 2, 5, "CIEND")) p = index_of(info, key_eq, p + 1);
        if (p < 0) return false;
    }
    const int q = index_of(info, ";", p + klen);
    out = sub(info, p + klen, q < 0 ? info.size() : q);
    return true;
}

enum { T_NONE = -1, T_INS = 1, T_DEL = 2, T_INV = 4, T_DUP = 6, T_OTHER = 99 };   // :30-38 (the four handled ones by value)

int type_of_label(const char* b, int n) {   // getType_infoField / getType_altField without the brackets (:226-290)
    static const struct { const char* s; int t; } table[] = {
        {"DEL", T_DEL}, {"DEL:ME", T_DEL}, {"DEL/INV", T_OTHER}, {"INS", T_INS}, {"INS:ME", T_INS}, {"INS:NOVEL", T_INS},
        {"DUP", T_DUP}, {"DUP:TANDEM", T_DUP}, {"DUP:INT", T_DUP}, {"INV", T_INV}, {"INVDUP", T_OTHER}, {"CNV", T_OTHER},
        {"BND", T_OTHER}, {"TRA", T_OTHER}};
    for (const auto& r : table)
        if (ieq(b, n, r.s)) return r.t;
    return T_NONE;
}

int round3(Span v) {   // (int)Math.round(Double.parseDouble(v) * 3)
    const double d = parse_double(v.b, v.e, "standard deviation");
    const double x = d * 3;
    if (x != x) return 0;
    const double f = std::floor(x + 0.5);
    long long r = f >= 9.2233720368547758e18 ? LLONG_MAX : (f <= -9.2233720368547758e18 ? LLONG_MIN : (long long)f);
    return (int)(int32_t)(uint32_t)(uint64_t)r;
}

// getConfidenceInterval (:299-433): the keys are searched WITHOUT '=' and the value starts key.length()+1 further on
void confidence_interval(Span info, bool at_end, int32_t out[2]) {
    struct Key { const char* name; int kind; };   // kind 0: "a,b" signed, 1: "a,b" digits only, 2: one sigma
    static const Key pos_keys[] = {{"CIPOS", 0}, {"STD_quant_start", 2}, {"STD_POS1", 2}};
    static const Key end_keys[] = {{"CIEND", 0}, {"CILEN", 1}, {"STD_quant_stop", 2}, {"STD_POS2", 2}};
    const Key* keys = at_end ? end_keys : pos_keys;
    const int nk = at_end ? 4 : 3;
    out[0] = 0; out[1] = 0;
    for (int k = 0; k < nk; k++) {
        int p = index_of(info, keys[k].name);
        if (p < 0) continue;
        p += (int)strlen(keys[k].name) + 1;
        if (keys[k].kind == 2) {
            const int q = index_of(info, ";", p);
            const char c = char_at(info, p);
            if (c >= '0' && c <= '9') { const int v = round3(sub(info, p, q >= 0 ? q : info.size())); out[0] = (int32_t)(0u - (uint32_t)v); out[1] = v; }
            return;
        }
        const bool minus_ok = keys[k].kind == 0;
        int q = index_of(info, ",", p);
        char c = char_at(info, p);
        if ((c >= '0' && c <= '9') || (minus_ok && c == '-')) { const Span v = sub(info, p, q >= 0 ? q : info.size()); out[0] = parse_int(v.b, v.e, keys[k].name); }
        p = q + 1; q = index_of(info, ";", p);
        c = char_at(info, p);
        if ((c >= '0' && c <= '9') || (minus_ok && c == '-')) { const Span v = sub(info, p, q >= 0 ? q : info.size()); out[1] = parse_int(v.b, v.e, keys[k].name); }
        return;
    }
}

struct ParsedLine {
    Span line;
    int64_t record;       // index into the record array, -1: copied through unchanged
    const char* tail_end; // end of the last token that survives String.split's removal of trailing empty fields
    Span info;            // tokens[7]
};

// tokens of str.split("\t"), trailing empty strings removed
void split_tabs(Span s, std::vector<Span>& tok) {
    tok.clear();
    const char* q = s.b;
    for (;;) {
        const char* t = (const char*)memchr(q, '\t', (size_t)(s.e - q));
        if (!t) { tok.push_back(Span{q, s.e}); break; }
        tok.push_back(Span{q, t});
        q = t + 1;
    }
    while (!tok.empty() && tok.back().b == tok.back().e && (tok.size() > 1 || s.size() > 0)) tok.pop_back();
}

Span token(const std::vector<Span>& tok, size_t i) {
    if (i >= tok.size()) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: Index " + std::to_string(i) + " out of bounds for length " + std::to_string(tok.size()));
    return tok[i];
}

}  // namespace

static AnnotateState& annot_of(sp_ctx* c) {
    if (!c->annot || !c->annot->loaded) throw Error(SP_ERR_ARG, "sp_annotate_load has not been called");
    return *c->annot;
}

void annotate_load(sp_ctx* c, const char* reference_file, const char* repeats_file, const char* segdups_file) {
    if (!reference_file || !repeats_file || !segdups_file) throw Error(SP_ERR_ARG, "AnnotateVCF: missing file argument");
    free_annotate(c);
    c->annot = new AnnotateState();
    AnnotateState& A = *c->annot;
    const auto t0 = std::chrono::steady_clock::now();
    load_reference_fasta(c, reference_file);                                   // :112
    A.model.load_repeat_intervals(repeats_file);                               // :113
    A.model.load_segdups(segdups_file);                                        // :114
    A.sets.build(A.model, c->h_ref.data(), (int32_t)c->ref_len, MAX_DISTANCE_SR);  // :115
    std::vector<int32_t> first, last;
    std::vector<int64_t> before;
    for (int t = 0; t < N_INTERVAL_TYPES; t++) {
        const IntervalSet& s = A.sets.set[t];
        A.set_off[t] = (int32_t)first.size();
        for (const auto& iv : s.iv) { first.push_back(iv.first); last.push_back(iv.second); }
        int64_t acc = 0;
        for (const auto& iv : s.iv) { before.push_back(acc); acc += (int64_t)iv.second - iv.first + 1; }
        before.push_back(acc);
    }
    A.set_off[N_INTERVAL_TYPES] = (int32_t)first.size();
    A.d_first.reserve(std::max<size_t>(first.size(), 1)); A.d_last.reserve(std::max<size_t>(last.size(), 1));
    A.d_before.reserve(before.size());
    cudaStream_t st = c->stream;
    if (!first.empty()) {
        SP_CUDA(cudaMemcpyAsync(A.d_first.p, first.data(), first.size() * 4, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(A.d_last.p, last.data(), last.size() * 4, cudaMemcpyHostToDevice, st));
    }
    SP_CUDA(cudaMemcpyAsync(A.d_before.p, before.data(), before.size() * 8, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaStreamSynchronize(st));
    SP_CUDA(cudaEventCreate(&A.ev0)); SP_CUDA(cudaEventCreate(&A.ev1));
    A.load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    A.loaded = true;
}

// annotateVCF (:127-205) for one file
int64_t annotate_file(sp_ctx* c, const char* input_file, const char* output_file, int32_t min_distance, int32_t max_distance) {
    AnnotateState& A = annot_of(c);
    if (!input_file || !output_file) throw Error(SP_ERR_ARG, "AnnotateVCF: missing file argument");
    if (max_distance < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: " + std::to_string((1 + (int64_t)max_distance * 2) * 5));  // new boolean[(1+maxDistance*2)*5], :136
    const std::vector<char> buf = slurp(input_file);
    Lines lines(buf.data(), buf.data() + buf.size());
    std::vector<ParsedLine> parsed;
    std::vector<AnnotRecord> recs;
    std::vector<Span> tok;
    const char *b, *e;
    while (lines.next(b, e)) {
        ParsedLine pl{Span{b, e}, -1, b, Span{b, b}};
        if (b == e) throw_sioobe(0, 0);                       // str.charAt(0) of an empty line (:148)
        if (*b != '#') {
            split_tabs(pl.line, tok);
            const Span info = token(tok, 7);
            Span v;
            int type = T_NONE;
            if (get_field(info, "SVTYPE=", false, v) && v.size() > 0) type = type_of_label(v.b, v.size());
            if (type == T_NONE) {
                const Span alt = token(tok, 4);
                if (alt.size() >= 2 && alt.b[0] == '<' && alt.e[-1] == '>') type = type_of_label(alt.b + 1, alt.size() - 2);
            }
            if (type == T_DEL || type == T_DUP || type == T_INS || type == T_INV) {
                AnnotRecord r{};
                const Span pos = token(tok, 1);
                r.position = parse_int(pos.b, pos.e, "POS");
                r.has_length = 0;
                if (get_field(info, "END=", true, v)) {
                    r.end = (int32_t)((uint32_t)parse_int(v.b, v.e, "END") - 1u);
                    if (type != T_INS) {
                        r.has_length = 1;                                   // length = end - position + 1 (:162)
                        if (get_field(info, "SVLEN=", false, v)) parse_int(v.b, v.e, "SVLEN");  // parsed (may throw), only a warning uses it
                    }
                } else if (type != T_INS) {
                    if (get_field(info, "SVLEN=", false, v)) {
                        int32_t len = parse_int(v.b, v.e, "SVLEN");
                        if (len < 0) len = (int32_t)(0u - (uint32_t)len);
                        r.end = (int32_t)((uint32_t)r.position + (uint32_t)len - 1u);
                        r.has_length = 2;                                   // length = |SVLEN| (:170-172)
                        r.pad = len;
                    } else r.end = INT_MAX;
                } else r.end = r.position;
                confidence_interval(info, false, r.ci_pos);
                if (r.end != INT_MAX) confidence_interval(info, true, r.ci_end);
                pl.record = (int64_t)recs.size();
                pl.tail_end = tok.back().e;
                pl.info = info;
                recs.push_back(r);
            }
        }
        parsed.push_back(pl);
    }
    const int64_t n = (int64_t)recs.size();
    std::vector<AnnotResult> res((size_t)n);
    A.kernel_ms = 0.0;
    if (n > 0) {
        A.d_rec.reserve((size_t)n); A.d_res.reserve((size_t)n);
        cudaStream_t st = c->stream;
        SP_CUDA(cudaMemcpyAsync(A.d_rec.p, recs.data(), (size_t)n * sizeof(AnnotRecord), cudaMemcpyHostToDevice, st));
        SetsDev S;
        S.first = A.d_first.p; S.last = A.d_last.p; S.before = A.d_before.p;
        for (int t = 0; t <= N_INTERVAL_TYPES; t++) S.off[t] = A.set_off[t];
        SP_CUDA(cudaEventRecord(A.ev0, st));
        annotate_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(A.d_rec.p, n, min_distance, max_distance, S, A.d_res.p);
        c->launches++;
        SP_CUDA(cudaGetLastError());
        SP_CUDA(cudaEventRecord(A.ev1, st));
        SP_CUDA(cudaMemcpyAsync(res.data(), A.d_res.p, (size_t)n * sizeof(AnnotResult), cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        float ms = 0.f;
        SP_CUDA(cudaEventElapsedTime(&ms, A.ev0, A.ev1));
        A.kernel_ms = ms;
    }
    // output: comment and unhandled lines unchanged, handled records with the three fields appended to INFO and any
    // trailing empty columns gone (String.split)
    std::string out;
    out.reserve(buf.size() + (size_t)n * 64);
    for (const ParsedLine& pl : parsed) {
        if (pl.record < 0) { out.append(pl.line.b, pl.line.e); out.push_back('\n'); continue; }
        const AnnotRecord& r = recs[(size_t)pl.record];
        const AnnotResult& o = res[(size_t)pl.record];
        double fraction = 0.0;
        if (r.end != INT_MAX && r.has_length) {
            const int32_t length = r.has_length == 2 ? r.pad : (int32_t)((uint32_t)r.end - (uint32_t)r.position + 1u);
            fraction = (double)o.surface / (double)length;
        }
        out.append(pl.line.b, pl.info.e);
        out += ";REPEATS_START=" + std::to_string(o.repeats_start) + ";REPEATS_END=" + std::to_string(o.repeats_end) +
               ";REPEATS_FRACTION=" + java_number_format4(fraction);
        if (pl.tail_end > pl.info.e) out.append(pl.info.e, pl.tail_end);   // the columns after INFO that survive the split
        out.push_back('\n');
    }
    FILE* f = fopen(output_file, "wb");
    if (!f) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("FileNotFoundException: ") + output_file);
    const size_t w = fwrite(out.data(), 1, out.size(), f);
    fclose(f);
    if (w != out.size()) throw Error(SP_ERR, std::string("IOException: short write ") + output_file);
    A.records_total += n;
    return n;
}

}  // namespace sp

using sp::Error;

extern "C" {

int sp_annotate_load(sp_ctx* ctx, const char* reference_file, const char* repeat_intervals_file, const char* segdups_file) {
    SP_API_BEGIN
    sp::annotate_load(ctx, reference_file, repeat_intervals_file, segdups_file);
    SP_API_END
}

int sp_annotate_file(sp_ctx* ctx, const char* input_vcf, const char* output_vcf, int32_t min_distance, int32_t max_distance,
                     int64_t* n_annotated) {
    SP_API_BEGIN
    const int64_t n = sp::annotate_file(ctx, input_vcf, output_vcf, min_distance, max_distance);
    if (n_annotated) *n_annotated = n;
    SP_API_END
}

int sp_annotate_vcf(sp_ctx* ctx, const sp_annotate_args* args) {
    SP_API_BEGIN
    if (!args || !args->input_files_list) throw Error(SP_ERR_ARG, "null sp_annotate_args");
    sp::annotate_load(ctx, args->reference_file, args->repeat_intervals_file, args->segdups_file);
    const std::vector<char> list = sp::slurp(args->input_files_list);
    sp::Lines lines(list.data(), list.data() + list.size());
    const char *b, *e;
    while (lines.next(b, e)) {
        const std::string path(b, e);
        const size_t p = path.find(".vcf");
        if (p == std::string::npos) throw Error(SP_ERR_JAVA_EXCEPTION, "StringIndexOutOfBoundsException: begin 0, end -1, length " + std::to_string(path.size()));
        sp::annotate_file(ctx, path.c_str(), (path.substr(0, p) + "_annotated.vcf").c_str(), args->min_distance, args->max_distance);
    }
    SP_API_END
}

int sp_annotate_timing(sp_ctx* ctx, double* load_seconds, double* last_kernel_ms, int64_t* records_total) {
    SP_API_BEGIN
    sp::AnnotateState& A = sp::annot_of(ctx);
    if (load_seconds) *load_seconds = A.load_s;
    if (last_kernel_ms) *last_kernel_ms = A.kernel_ms;
    if (records_total) *records_total = A.records_total;
    SP_API_END
}

}  // extern "C"
