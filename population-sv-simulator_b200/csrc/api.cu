// C ABI of libsynthpop_cuda.so: context, reference, variant table, haplotype lists, materialisation.
#include "common.cuh"
#include "textio.h"
#include <algorithm>
#include "file_sink.h"

namespace sp {
void plan_haplotypes(sp_ctx* c, int32_t first, int32_t last, int32_t framing, int64_t* out_lengths, int64_t* slot_stride);
void materialise(sp_ctx* c);
void free_population(sp_ctx* c);
void free_reads(sp_ctx* c);
void free_length_bins(sp_ctx* c);
void free_read_mix(sp_ctx* c);
void free_annotate(sp_ctx* c);

// Double.toString (JDK >= 19): the shortest decimal (at least two digits) that reads back as d and, among those of that
// length, the one closest to d. Worked from the EXACT expansion of d (glibc's printf is exact at any precision): at length n
// the only candidates are the expansion cut after n digits and that value plus one unit in the last place; which of the two
// is closer is read off the digits that were cut.
std::string java_double_to_string(double d) {
    if (d != d) return "NaN";
    if (d == 1.0 / 0.0) return "Infinity";
    if (d == -1.0 / 0.0) return "-Infinity";
    if (d == 0.0) return std::signbit(d) ? "-0.0" : "0.0";
    const bool neg = d < 0;
    const double ad = neg ? -d : d;
    char ex[900];
    snprintf(ex, sizeof ex, "%.800e", ad);   // 801 significant digits: more than any double has
    std::string all;
    const char* e = strchr(ex, 'e');
    for (const char* q = ex; q < e; q++) if (*q != '.') all.push_back(*q);
    int exp10 = atoi(e + 1);
    std::string digits;
    int dexp = exp10;
    for (size_t n = 2; n <= 17 && digits.empty(); n++) {
        const std::string lo = all.substr(0, n);
        std::string hi = lo;   // lo + 1 unit in the last place
        int hexp = exp10;
        {
            int k = (int)n - 1;
            while (k >= 0 && hi[(size_t)k] == '9') hi[(size_t)k--] = '0';
            if (k >= 0) hi[(size_t)k]++; else { hi = "1" + hi.substr(0, n - 1); hexp++; }
        }
        auto reads_back = [&](const std::string& m, int x) {
            const std::string t = m.substr(0, 1) + "." + m.substr(1) + "e" + std::to_string(x);
            return strtod(t.c_str(), nullptr) == ad;
        };
        const bool lo_ok = reads_back(lo, exp10), hi_ok = reads_back(hi, hexp);
        if (!lo_ok && !hi_ok) continue;
        bool take_hi = hi_ok && !lo_ok;
        if (lo_ok && hi_ok) {   // the cut-off tail against one half: above -> hi is closer
            const std::string tail = all.substr(n);
            if (tail[0] > '5') take_hi = true;
            else if (tail[0] == '5') take_hi = tail.find_first_not_of('0', 1) != std::string::npos || ((lo[n - 1] - '0') & 1);
        }
        digits = take_hi ? hi : lo;
        dexp = take_hi ? hexp : exp10;
    }
    if (digits.empty()) { digits = all.substr(0, 17); dexp = exp10; }
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out = neg ? "-" : "";
    if (ad >= 1e-3 && ad < 1e7) {
        if (dexp >= 0) {
            for (int i = 0; i <= dexp; i++) out.push_back(i < (int)digits.size() ? digits[(size_t)i] : '0');
            out.push_back('.');
            if ((int)digits.size() > dexp + 1) out += digits.substr((size_t)dexp + 1); else out.push_back('0');
        } else {
            out += "0.";
            out.append((size_t)(-dexp - 1), '0');
            out += digits;
        }
    } else {
        out.push_back(digits[0]);
        out.push_back('.');
        if (digits.size() > 1) out += digits.substr(1); else out.push_back('0');
        out += "E" + std::to_string(dexp);
    }
    return out;
}

std::string java_number_format4(double x) {
    if (x != x) return "NaN";
    char buf[512];
    snprintf(buf, sizeof buf, "%.4f", x);
    std::string s(buf);
    const bool neg = s[0] == '-';
    if (neg) s.erase(0, 1);
    const size_t dot = s.find('.');
    std::string out = neg ? "-" : "";
    for (size_t i = 0; i < dot; i++) {
        if (i > 0 && (dot - i) % 3 == 0) out.push_back(',');
        out.push_back(s[i]);
    }
    return out + s.substr(dot);
}

static void upload_variants(sp_ctx* c) {
    const int32_t n = c->n_variants;
    c->d_type.reserve(std::max(n, 1)); c->d_pos.reserve(std::max(n, 1)); c->d_len.reserve(std::max(n, 1));
    c->d_payload_off.reserve(n + 1);
    c->d_payload.reserve(std::max<size_t>(c->h_payload.size(), 1));
    cudaStream_t st = c->stream;
    if (n > 0) {
        SP_CUDA(cudaMemcpyAsync(c->d_type.p, c->h_type.data(), n, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(c->d_pos.p, c->h_pos.data(), n * 4, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(c->d_len.p, c->h_len.data(), n * 4, cudaMemcpyHostToDevice, st));
    }
    SP_CUDA(cudaMemcpyAsync(c->d_payload_off.p, c->h_payload_off.data(), (n + 1) * 8, cudaMemcpyHostToDevice, st));
    if (!c->h_payload.empty())
        SP_CUDA(cudaMemcpyAsync(c->d_payload.p, c->h_payload.data(), c->h_payload.size(), cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaStreamSynchronize(st));
}

static void upload_lists(sp_ctx* c) {
    const int32_t H = c->n_haplotypes;
    c->d_list_off.reserve(H + 1);
    c->d_list_ids.reserve(std::max<size_t>(c->h_list_ids.size(), 1));
    SP_CUDA(cudaMemcpyAsync(c->d_list_off.p, c->h_list_off.data(), (H + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    if (!c->h_list_ids.empty())
        SP_CUDA(cudaMemcpyAsync(c->d_list_ids.p, c->h_list_ids.data(), c->h_list_ids.size() * 4, cudaMemcpyHostToDevice, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
    c->view_own_lists();
    c->plan_last = -1; c->plan_first = 0;
}

static void set_reference(sp_ctx* c) {
    c->ref_len = (int64_t)c->h_ref.size();
    c->d_ref.reserve(std::max<size_t>(c->h_ref.size(), 1));
    if (c->ref_len > 0)
        SP_CUDA(cudaMemcpyAsync(c->d_ref.p, c->h_ref.data(), c->h_ref.size(), cudaMemcpyHostToDevice, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
    c->plan_last = -1; c->plan_first = 0;
}

// SimulateHaplotypes.loadReference, src/SimulateHaplotypes.java:119-134: skip one header line, String.trim() every
// following line and concatenate. Lines are found with memchr and copied with memcpy (a chr1-sized FASTA is four million
// 60-byte lines; this loop is most of row A1's time).
void load_reference_fasta(sp_ctx* c, const char* path) {
    std::vector<char> buf = slurp(path);
    const char* p = buf.data();
    const char* const end = p + buf.size();
    auto next_line = [&](const char*& b, const char*& e) -> bool {   // BufferedReader.readLine: \n, \r or \r\n end a line
        if (p >= end) return false;
        b = p;
        const char* q = (const char*)memchr(p, '\n', (size_t)(end - p));
        const char* stop = q ? q : end;
        const char* r = (const char*)memchr(p, '\r', (size_t)(stop - p));
        if (r) {   // a carriage return before the line feed: the line ends there ("\r\n" counts as one terminator)
            e = r;
            p = (r + 1 < end && r[1] == '\n') ? r + 2 : r + 1;
        } else {
            e = stop;
            p = q ? q + 1 : end;
        }
        return true;
    };
    const char *b, *e;
    next_line(b, e);  // Skipping header
    c->h_ref.clear();
    c->h_ref.resize(buf.size());
    uint8_t* out = c->h_ref.data();
    while (next_line(b, e)) {
        if (b < e && ((unsigned char)*b <= ' ' || (unsigned char)e[-1] <= ' ')) trim_span(b, e);
        memcpy(out, b, (size_t)(e - b));
        out += e - b;
    }
    c->h_ref.resize((size_t)(out - c->h_ref.data()));
    if (c->h_ref.size() > (size_t)INT32_MAX) throw Error(SP_ERR_ARG, "reference longer than 2^31-1 (the reference uses int coordinates)");
    set_reference(c);
}

// SimulateHaplotypes.deserialize, src/SimulateHaplotypes.java:1147-1178 + Variant.deserialize :1294-1306
void load_population(sp_ctx* c, const char* h2v_file, const char* variants_file) {
    std::vector<std::pair<const char*, const char*>> tok;
    const char *b, *e;
    {
        std::vector<char> buf = slurp(h2v_file);
        Lines lines(buf.data(), buf.data() + buf.size());
        if (!lines.next(b, e)) throw Error(SP_ERR_JAVA_EXCEPTION, "NumberFormatException: null (haplotype2variants header)");
        const int32_t H = parse_int(b, e, "N_HAPLOTYPES");
        if (H < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: N_HAPLOTYPES");
        c->h_list_off.assign((size_t)H + 1, 0);
        c->h_list_ids.clear();
        for (int32_t i = 0; i < H; i++) {
            if (!lines.next(b, e)) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: truncated haplotype2variants file");
            split_comma(b, e, tok);
            if (tok.empty()) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: 0");
            const int32_t last = parse_int(tok[0].first, tok[0].second, "haplotype2variants_last");
            if (last + 1 < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException");
            if ((int64_t)tok.size() < (int64_t)last + 2) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: haplotype2variants row " + std::to_string(i));
            for (int32_t j = 0; j <= last; j++) c->h_list_ids.push_back(parse_int(tok[1 + j].first, tok[1 + j].second, "variant id"));
            c->h_list_off[i + 1] = (int64_t)c->h_list_ids.size();
        }
        c->n_haplotypes = H;
    }
    {
        std::vector<char> buf = slurp(variants_file);
        Lines lines(buf.data(), buf.data() + buf.size());
        auto must = [&](const char* what) {
            if (!lines.next(b, e)) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("NumberFormatException: null (") + what + ")");
        };
        must("lastVariant");
        const int32_t n = parse_int(b, e, "lastVariant") + 1;
        if (n < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: lastVariant");
        c->n_variants = n;
        c->h_type.assign(n, 0); c->h_pos.assign(n, 0); c->h_len.assign(n, 0);
        c->h_payload_off.assign((size_t)n + 1, 0);
        c->h_payload.clear();
        for (int32_t i = 0; i < n; i++) {
            must("svType"); const int32_t t = parse_int(b, e, "svType");
            must("intervalType"); parse_int(b, e, "intervalType");
            must("length"); c->h_len[i] = parse_int(b, e, "length");
            must("alleleFrequency"); parse_double(b, e, "alleleFrequency");
            must("position"); c->h_pos[i] = parse_int(b, e, "position");
            must("isActive");
            must("sequence");
            const bool is_null = (e - b) == 4 && strncasecmp(b, "null", 4) == 0;
            if (t < -128 || t > 127) throw Error(SP_ERR_ARG, "invalid svType");
            c->h_type[i] = (int8_t)t;
            // an INS whose sequence is null splices the four characters "null" (StringBuilder.append((String)null))
            if (!is_null || t == SP_INS) c->h_payload.insert(c->h_payload.end(), (const uint8_t*)b, (const uint8_t*)e);
            c->h_payload_off[i + 1] = (int64_t)c->h_payload.size();
            must("intervalType_start"); parse_int(b, e, "intervalType_start");
            must("intervalType_end"); parse_int(b, e, "intervalType_end");
            must("intervalSurface"); parse_int(b, e, "intervalSurface");
        }
    }
    upload_variants(c);
    upload_lists(c);
}

// Streams device bytes into files through a ring of pinned chunks: the caller's thread issues the D2H copies (stream
// order), a writer thread waits for each chunk's event and appends it to the current file. One file at a time, in order.
FileSink::FileSink(int device, int n_slots, int64_t chunk_bytes) : device_(device), chunk_(chunk_bytes), slots_((size_t)n_slots) {
    for (Slot& s : slots_) {
        SP_CUDA(cudaHostAlloc((void**)&s.buf, (size_t)chunk_, cudaHostAllocDefault));
        SP_CUDA(cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming));
    }
    writer_ = std::thread([this] { this->run(); });
}

FileSink::~FileSink() {
    { std::lock_guard<std::mutex> g(m_); stop_ = true; }
    cv_.notify_all();
    if (writer_.joinable()) writer_.join();
    for (Slot& s : slots_) { if (s.buf) cudaFreeHost(s.buf); if (s.ev) cudaEventDestroy(s.ev); }
    if (file_) fclose(file_);
}

void FileSink::run() {
    cudaSetDevice(device_);
    for (;;) {
        Task t;
        {
            std::unique_lock<std::mutex> g(m_);
            cv_.wait(g, [&] { return stop_ || !queue_.empty(); });
            if (queue_.empty()) return;
            t = queue_.front(); queue_.pop_front();
        }
        Slot& s = slots_[(size_t)t.slot];
        bool ok = cudaEventSynchronize(s.ev) == cudaSuccess;
        if (ok && !failed_) {
            if (t.open) {
                file_ = fopen(t.path.c_str(), "wb");
                if (!file_) { std::lock_guard<std::mutex> g(m_); failed_ = true; error_ = "FileNotFoundException: " + t.path; }
            }
            if (file_ && fwrite(s.buf, 1, (size_t)t.bytes, file_) != (size_t)t.bytes) {
                std::lock_guard<std::mutex> g(m_); failed_ = true; error_ = "IOException: short write " + t.path;
            }
            if (t.close && file_) { fclose(file_); file_ = nullptr; }
        } else if (!ok) { std::lock_guard<std::mutex> g(m_); failed_ = true; error_ = "device-to-host copy failed"; }
        { std::lock_guard<std::mutex> g(m_); s.busy = false; pending_--; }
        cv_.notify_all();
    }
}

void FileSink::write_file(const std::string& path, const uint8_t* d_src, int64_t n, cudaStream_t st) {
    int64_t off = 0;
    bool first = true;
    do {
        const int64_t nb = std::min(chunk_, n - off);
        int slot;
        {
            std::unique_lock<std::mutex> g(m_);
            cv_.wait(g, [&] { return !slots_[(size_t)next_].busy; });
            if (failed_) {   // drain() reports it: nothing more is queued for this file
                g.unlock();
                drain();
            }
            slot = next_; next_ = (next_ + 1) % (int)slots_.size();
            slots_[(size_t)slot].busy = true; pending_++;
        }
        if (nb > 0) SP_CUDA(cudaMemcpyAsync(slots_[(size_t)slot].buf, d_src + off, (size_t)nb, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaEventRecord(slots_[(size_t)slot].ev, st));
        off += nb;
        {
            std::lock_guard<std::mutex> g(m_);
            queue_.push_back(Task{slot, nb, first, off >= n, path});
        }
        cv_.notify_all();
        first = false;
    } while (off < n);
}

void FileSink::drain() {
    std::unique_lock<std::mutex> g(m_);
    cv_.wait(g, [&] { return pending_ == 0; });
    if (failed_) {   // report once; the sink stays usable for the next file
        failed_ = false;
        const std::string what = error_;
        throw Error(SP_ERR_JAVA_EXCEPTION, what);
    }
}

void FileSink::finish() { drain(); }

FileSink& file_sink(sp_ctx* c) {
    if (!c->sink) c->sink = new FileSink(c->device, 4, 8ll << 20);
    return *c->sink;
}
void free_file_sink(sp_ctx* c) {
    delete c->sink;
    c->sink = nullptr;
}

static void check_planned(sp_ctx* c, int32_t h) {
    if (c->plan_last < c->plan_first || h < c->plan_first || h > c->plan_last)
        throw Error(SP_ERR_ARG, "haplotype " + std::to_string(h) + " is not in the planned batch");
}

}  // namespace sp

using sp::Error;


extern "C" {

const char* sp_version(void) { return "synthpop-cuda 0.2 (sm_100a)"; }

// Test hooks: the product's own Double.toString / NumberFormat(4 digits) restatements (host text formatting of rows W, Z
// and AnnotateVCF), so that the known-answer vectors run against this library and not only against the oracle's copy.
int sp_test_double_to_string(double x, char* out, int32_t capacity) {
    const std::string s = sp::java_double_to_string(x);
    snprintf(out, (size_t)capacity, "%s", s.c_str());
    return (int)s.size();
}
int sp_test_number_format4(double x, char* out, int32_t capacity) {
    const std::string s = sp::java_number_format4(x);
    snprintf(out, (size_t)capacity, "%s", s.c_str());
    return (int)s.size();
}

int sp_create(sp_ctx** out, int device, int64_t seed) {
    if (!out) return SP_ERR_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return SP_ERR_CUDA;  // no CPU fallback
    sp_ctx* c = new sp_ctx();
    c->device = device; c->seed = seed;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c;
        return SP_ERR_CUDA;
    }
    c->stream = c->own_stream;
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
    *out = c;
    return SP_OK;
}

void sp_destroy(sp_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStream_t s = ctx->own_stream;
    sp::free_population(ctx);
    sp::free_reads(ctx);
    sp::free_length_bins(ctx);
    sp::free_read_mix(ctx);
    sp::free_annotate(ctx);
    sp::free_file_sink(ctx);
    if (ctx->ev_planned) cudaEventDestroy(ctx->ev_planned);
    for (auto& b : ctx->plan_set) if (b.ev_used) cudaEventDestroy(b.ev_used);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    delete ctx;
    if (s) cudaStreamDestroy(s);
}

const char* sp_last_error(const sp_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int sp_set_stream(sp_ctx* ctx, void* cuda_stream) {
    SP_API_BEGIN
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    SP_API_END
}

int sp_set_population_stream(sp_ctx* ctx, void* cuda_stream) {
    SP_API_BEGIN
    ctx->pop_stream = (cudaStream_t)cuda_stream;
    SP_API_END
}

int64_t sp_kernel_launches(const sp_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sp_load_reference_fasta(sp_ctx* ctx, const char* path) {
    SP_API_BEGIN
    sp::load_reference_fasta(ctx, path);
    SP_API_END
}

int sp_set_reference(sp_ctx* ctx, const uint8_t* bases, int64_t length) {
    SP_API_BEGIN
    if (length < 0 || (length > 0 && !bases) || length > INT32_MAX) throw Error(SP_ERR_ARG, "bad reference buffer");
    // straight H2D from the caller's (ideally pinned) buffer; no host copy is kept. A population begun on the previous
    // reference is dropped: its writers would read the host copy that is no longer there.
    sp::free_population(ctx);
    ctx->h_ref.clear();
    ctx->ref_len = length;
    ctx->d_ref.reserve(std::max<size_t>((size_t)length, 1));
    if (length > 0) SP_CUDA(cudaMemcpyAsync(ctx->d_ref.p, bases, (size_t)length, cudaMemcpyHostToDevice, ctx->stream));
    SP_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->plan_last = -1; ctx->plan_first = 0;
    SP_API_END
}

int64_t sp_reference_length(const sp_ctx* ctx) { return ctx ? ctx->ref_len : -1; }

int sp_reference_device(sp_ctx* ctx, void** device_ptr, int64_t* length) {
    SP_API_BEGIN
    if (!device_ptr) throw Error(SP_ERR_ARG, "null output pointer");
    if (ctx->ref_len <= 0 || !ctx->d_ref.p) throw Error(SP_ERR_ARG, "no reference is resident");
    SP_CUDA(cudaStreamSynchronize(ctx->stream));   // the caller reads it on a stream of its own
    *device_ptr = ctx->d_ref.p;
    if (length) *length = ctx->ref_len;
    SP_API_END
}

int sp_reference_device_buffer(sp_ctx* ctx, int64_t length, void** device_ptr) {
    SP_API_BEGIN
    if (length <= 0 || length > INT32_MAX || !device_ptr) throw Error(SP_ERR_ARG, "bad reference buffer");
    sp::free_population(ctx);
    ctx->h_ref.clear();
    SP_CUDA(cudaStreamSynchronize(ctx->stream));   // nothing of this context may still read the old reference
    ctx->ref_len = length;
    ctx->d_ref.reserve((size_t)length);
    ctx->plan_last = -1; ctx->plan_first = 0;
    *device_ptr = ctx->d_ref.p;
    SP_API_END
}

int sp_set_variants(sp_ctx* ctx, int32_t n, const int8_t* sv_type, const int32_t* position, const int32_t* length,
                    const int64_t* payload_off, const uint8_t* payload_blob, const int8_t* has_payload) {
    SP_API_BEGIN
    if (n < 0 || (n > 0 && (!sv_type || !position || !length))) throw Error(SP_ERR_ARG, "bad variant table");
    sp::free_population(ctx);   // its tables are being replaced
    ctx->n_variants = n;
    ctx->h_type.assign(sv_type, sv_type + n);
    ctx->h_pos.assign(position, position + n);
    ctx->h_len.assign(length, length + n);
    ctx->h_payload_off.assign((size_t)n + 1, 0);
    ctx->h_payload.clear();
    for (int32_t i = 0; i < n; i++) {
        const int64_t a = payload_off ? payload_off[i] : 0, b = payload_off ? payload_off[i + 1] : 0;
        if (b < a) throw Error(SP_ERR_ARG, "payload offsets must be non-decreasing");
        const bool has = has_payload ? has_payload[i] != 0 : b > a;
        if (has) ctx->h_payload.insert(ctx->h_payload.end(), payload_blob + a, payload_blob + b);
        else if (sv_type[i] == SP_INS) { const char* s = "null"; ctx->h_payload.insert(ctx->h_payload.end(), s, s + 4); }
        ctx->h_payload_off[i + 1] = (int64_t)ctx->h_payload.size();
    }
    sp::upload_variants(ctx);
    ctx->plan_last = -1; ctx->plan_first = 0;
    SP_API_END
}

int sp_set_haplotype_lists(sp_ctx* ctx, int32_t n_haplotypes, const int64_t* list_off, const int32_t* ids) {
    SP_API_BEGIN
    if (n_haplotypes < 0 || !list_off || list_off[0] != 0) throw Error(SP_ERR_ARG, "bad haplotype lists");
    for (int32_t i = 0; i < n_haplotypes; i++)
        if (list_off[i + 1] < list_off[i]) throw Error(SP_ERR_ARG, "list offsets must be non-decreasing");
    sp::free_population(ctx);   // its lists are being replaced
    ctx->n_haplotypes = n_haplotypes;
    ctx->h_list_off.assign(list_off, list_off + n_haplotypes + 1);
    ctx->h_list_ids.assign(ids, ids + list_off[n_haplotypes]);
    sp::upload_lists(ctx);
    SP_API_END
}

int sp_get_variants(sp_ctx* ctx, int32_t* n, int64_t* payload_bytes, int8_t* sv_type, int32_t* position, int32_t* length,
                    int64_t* payload_off, uint8_t* payload_blob) {
    SP_API_BEGIN
    const int32_t nv = ctx->n_variants;
    if (n) *n = nv;
    if (payload_bytes) *payload_bytes = (int64_t)ctx->h_payload.size();
    if (sv_type && nv) memcpy(sv_type, ctx->h_type.data(), nv);
    if (position && nv) memcpy(position, ctx->h_pos.data(), (size_t)nv * 4);
    if (length && nv) memcpy(length, ctx->h_len.data(), (size_t)nv * 4);
    if (payload_off && !ctx->h_payload_off.empty()) memcpy(payload_off, ctx->h_payload_off.data(), ((size_t)nv + 1) * 8);
    if (payload_blob && !ctx->h_payload.empty()) memcpy(payload_blob, ctx->h_payload.data(), ctx->h_payload.size());
    SP_API_END
}

int sp_get_haplotype_lists(sp_ctx* ctx, int32_t* n_haplotypes, int64_t* n_ids, int64_t* list_off, int32_t* ids) {
    SP_API_BEGIN
    if (n_haplotypes) *n_haplotypes = ctx->n_haplotypes;
    if (n_ids) *n_ids = (int64_t)ctx->h_list_ids.size();
    if (list_off && !ctx->h_list_off.empty()) memcpy(list_off, ctx->h_list_off.data(), ctx->h_list_off.size() * 8);
    if (ids && !ctx->h_list_ids.empty()) memcpy(ids, ctx->h_list_ids.data(), ctx->h_list_ids.size() * 4);
    SP_API_END
}

int sp_load_population(sp_ctx* ctx, const char* haplotype2variants_file, const char* variants_file) {
    SP_API_BEGIN
    sp::load_population(ctx, haplotype2variants_file, variants_file);
    SP_API_END
}

int sp_plan_haplotypes(sp_ctx* ctx, int32_t first, int32_t last, int32_t fasta_framing, int64_t* out_lengths, int64_t* slot_stride) {
    SP_API_BEGIN
    sp::plan_haplotypes(ctx, first, last, fasta_framing, out_lengths, slot_stride);
    SP_API_END
}

int sp_materialise(sp_ctx* ctx) {
    SP_API_BEGIN
    sp::materialise(ctx);
    SP_API_END
}

int sp_synchronize(sp_ctx* ctx) {
    SP_API_BEGIN
    SP_CUDA(cudaStreamSynchronize(ctx->stream));
    SP_API_END
}

int sp_haplotype_device(sp_ctx* ctx, int32_t h, void** device_ptr, int64_t* sequence_length) {
    SP_API_BEGIN
    sp::check_planned(ctx, h);
    const int i = h - ctx->plan_first;
    const int64_t hdr = ctx->h_hdr_len[i];
    if (device_ptr) *device_ptr = ctx->d_out.p + (int64_t)i * ctx->slot_stride + hdr;
    if (sequence_length) *sequence_length = ctx->h_hap_len[i] - hdr - (hdr > 0 ? 1 : 0);
    SP_API_END
}

int sp_copy_haplotype(sp_ctx* ctx, int32_t h, uint8_t* host_out, int64_t capacity, int64_t* written) {
    SP_API_BEGIN
    sp::check_planned(ctx, h);
    if (!ctx->materialised) throw Error(SP_ERR_ARG, "sp_materialise has not been called");
    const int i = h - ctx->plan_first;
    const int64_t n = ctx->h_hap_len[i];
    if (n > capacity) throw Error(SP_ERR_ARG, "host buffer too small");
    SP_CUDA(cudaMemcpyAsync(host_out, ctx->d_out.p + (int64_t)i * ctx->slot_stride, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    SP_CUDA(cudaStreamSynchronize(ctx->stream));
    if (written) *written = n;
    SP_API_END
}

int sp_build_haplotypes_host(sp_ctx* ctx, int32_t first, int32_t last, int32_t fasta_framing, uint8_t* host_out,
                             int64_t capacity, int64_t* offsets) {
    SP_API_BEGIN
    sp::plan_haplotypes(ctx, first, last, fasta_framing, nullptr, nullptr);
    sp::materialise(ctx);
    const int nh = last - first + 1;
    int64_t off = 0;
    for (int i = 0; i < nh; i++) {
        const int64_t n = ctx->h_hap_len[i];
        if (off + n > capacity) throw Error(SP_ERR_ARG, "host buffer too small");
        SP_CUDA(cudaMemcpyAsync(host_out + off, ctx->d_out.p + (int64_t)i * ctx->slot_stride, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
        if (offsets) offsets[i] = off;
        off += n;
    }
    if (offsets) offsets[nh] = off;
    SP_CUDA(cudaStreamSynchronize(ctx->stream));
    SP_API_END
}

int sp_print_haplotypes(sp_ctx* ctx, int32_t first, int32_t last, const char* reference_file,
                        const char* haplotype2variants_file, const char* variants_file, const char* output_dir) {
    SP_API_BEGIN
    sp::load_reference_fasta(ctx, reference_file);
    sp::load_population(ctx, haplotype2variants_file, variants_file);
    if (first > last) return SP_OK;  // the Java for-loop simply does not execute
    if (first < 0 || first >= ctx->n_haplotypes)
        throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: haplotype " + std::to_string(first));
    const int32_t requested_last = last;
    last = std::min(last, ctx->n_haplotypes - 1);  // the Java writes the valid ones, then throws at the first invalid index
    // batches sized to the free HBM; each slot is the exact byte content of haplotype_<i>.fa
    size_t free_b = 0, total_b = 0;
    SP_CUDA(cudaMemGetInfo(&free_b, &total_b));
    const int64_t per_hap = ctx->ref_len * 2 + (64 << 20);
    int32_t batch = (int32_t)std::max<int64_t>(1, std::min<int64_t>(64, (int64_t)(free_b / 2) / std::max<int64_t>(per_hap, 1)));
    // Device -> pinned ring -> files: the copy engine fills chunk k+1 (and the copy kernel materialises the next batch)
    // while a writer thread pushes chunk k into the page cache.
    sp::FileSink& sink = sp::file_sink(ctx);
    for (int32_t a = first; a <= last; a += batch) {
        const int32_t b = std::min(last, a + batch - 1);
        sink.drain();   // the slots are about to be overwritten: every copy out of them must have been issued AND done
        sp::plan_haplotypes(ctx, a, b, 1, nullptr, nullptr);
        sp::materialise(ctx);
        for (int32_t h = a; h <= b; h++)
            sink.write_file(std::string(output_dir) + "/haplotype_" + std::to_string(h) + ".fa",
                            ctx->d_out.p + (int64_t)(h - a) * ctx->slot_stride, ctx->h_hap_len[h - a], ctx->stream);
    }
    sink.finish();
    if (requested_last > last)
        throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: haplotype " + std::to_string(last + 1));
    SP_API_END
}

}  // extern "C"
