// SURVEY.md §8(f) rank 3 — BuildReadLengthBins (reference src/BuildReadLengthBins.java:21-116, printStats :139-239,
// src/Fastq.java) and Fastq2LengthHistogram (src/Fastq2LengthHistogram.java:11-51): the producer of row D's bin files.
//
// The reference streams every FASTQ record into its length bin, re-reads every bin, permutes it with a sequential
// Fisher-Yates (`j = random.nextInt(i+1)` for i = last..1, ONE java.util.Random shared by all bins, :108-116) and writes it
// back. Here the FASTQ bytes go to HBM once and everything is index work on the device until one gather writes all bins:
//   line index ...... terminator positions by count + scan + scatter (line_count_kernel, line_scatter_kernel)
//   records ......... record r = lines 4r..4r+3; bin = min(len(sequence) / BIN_LENGTH, N_BINS-1) (:57-58); per-bin counts
//                     and sequence bases by atomics; stable radix sort by bin = the bins in input order
//   draws ........... call g of nextInt() starts g + (redraws before it) steps into the LCG stream. The redraws are rare
//                     (probability (2^31 mod n)/2^31 per call), so the offsets are found as a fixpoint: every call is
//                     evaluated at its assumed step by jump-ahead, the redraw counts are scanned into new offsets, until
//                     nothing moves (each pass fixes at least everything up to the next real redraw).
//   swaps ........... position i is final after step i and receives what position j_i holds then. With
//                     next(p, t) = min{y > t : j_y = p} (binary search in the sorted pairs (j_y, y)), the content of
//                     position x just before its own step is C(x) = C(next(x, x)), or record x if no later step writes x:
//                     a forest, resolved by pointer jumping; final[i] = C(next(j_i, i)) or record j_i. No thread ever
//                     performs a dependent swap. (Checked against the sequential loop in tests/test_build_bins.py.)
//   gather .......... all bins back to back through the tiled copy kernel of row M (one segment per record, 2 B per byte),
//                     then device -> pinned ring -> <prefix><i>.bin.
#include "common.cuh"
#include "java_random.cuh"
#include "textio.h"
#include "file_sink.h"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>
#include <cmath>

namespace sp {

void launch_tiled_copy(sp_ctx* c, const uint8_t* src, const uint8_t* payload, const uint8_t* aux, const int64_t* seg_dst,
                       const int64_t* seg_src, const int64_t* seg_base, int64_t n_segments, uint8_t* out, int64_t stride, int n_slots);
void reserve_tiled_copy(sp_ctx* c, int64_t stride, int n_slots, int64_t n_segments);

namespace {

constexpr int LI_BYTES = 4096;   // bytes per CTA of the line index kernels (256 threads x 16)

__device__ __forceinline__ bool is_term(const uint8_t* b, int64_t i, int64_t n) { return i < n && b[i] == '\n'; }

// terminators per CTA ('\n' only: the host turned "\r\n" and lone '\r' into '\n' while reading, as readLine does)
__global__ void __launch_bounds__(256)
line_count_kernel(const uint8_t* __restrict__ blob, int64_t n, int64_t* __restrict__ block_count) {
    __shared__ int32_t ws[8];
    const int64_t i0 = (int64_t)blockIdx.x * LI_BYTES + threadIdx.x * 16;
    int32_t c = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) c += is_term(blob, i0 + k, n) ? 1 : 0;
    for (int d = 16; d > 0; d >>= 1) c += __shfl_down_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) { int32_t s = 0; for (int k = 0; k < 8; k++) s += ws[k]; block_count[blockIdx.x] = s; }
}

__global__ void __launch_bounds__(256)
line_scatter_kernel(const uint8_t* __restrict__ blob, int64_t n, const int64_t* __restrict__ block_base, int64_t* __restrict__ term_pos) {
    __shared__ int32_t ws[8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i0 = (int64_t)blockIdx.x * LI_BYTES + threadIdx.x * 16;
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) m |= is_term(blob, i0 + k, n) ? 1u << k : 0u;
    const int32_t v = __popc(m);
    int32_t incl = v;
    for (int d = 1; d < 32; d <<= 1) { const int32_t o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
    if (lane == 31) ws[w] = incl;
    __syncthreads();
    int32_t wp = 0;
    for (int q = 0; q < w; q++) wp += ws[q];
    int64_t p = block_base[blockIdx.x] + wp + incl - v;
    while (m) { const int k = __ffs(m) - 1; m &= m - 1; term_pos[p++] = i0 + k; }
}

// record r = lines 4r .. 4r+3 (header, sequence, separator, quality)
__global__ void __launch_bounds__(256)
record_kernel(const int64_t* __restrict__ term_pos, int64_t n_records, int32_t bin_length, int32_t n_bins, int64_t* __restrict__ rec_start,
              int32_t* __restrict__ rec_bytes, int32_t* __restrict__ bin, int32_t* __restrict__ idx, unsigned long long* __restrict__ hist,
              unsigned long long* __restrict__ bases, int32_t* __restrict__ err) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_records) return;
    const int64_t s0 = r == 0 ? 0 : term_pos[4 * r - 1] + 1;
    const int64_t t0 = term_pos[4 * r], t1 = term_pos[4 * r + 1], t3 = term_pos[4 * r + 3];
    const int64_t seq_len = t1 - t0 - 1;
    const int64_t nb = t3 + 1 - s0;
    if (seq_len > INT32_MAX || nb > INT32_MAX) { atomicExch(err, 1); return; }
    int32_t b = (int32_t)seq_len / bin_length;     // :57
    if (b >= n_bins) b = n_bins - 1;               // :58
    if (b < 0) { atomicExch(err, 2); return; }     // buffers_last[-1]: ArrayIndexOutOfBoundsException
    rec_start[r] = s0; rec_bytes[r] = (int32_t)nb;
    bin[r] = b; idx[r] = (int32_t)r;
    atomicAdd(&hist[b], 1ull);
    atomicAdd(&bases[b], (unsigned long long)seq_len);
}

struct CallMap {   // the nextInt() calls of all shuffled bins, in the order the reference makes them
    const int64_t* call_base;   // n_bins+1: first call of each bin (bins with fewer than 3 records make none)
    const int64_t* bin_off;     // n_bins+1: first output position of each bin
    int32_t n_bins;
};
__device__ __forceinline__ int32_t bin_of(const int64_t* base, int32_t n_bins, int64_t g) {   // last b with base[b] <= g, base[b+1] > g
    int32_t lo = 0, hi = n_bins;
    while (hi - lo > 1) { const int32_t mid = (lo + hi) >> 1; if (base[mid] <= g) lo = mid; else hi = mid; }
    return lo;
}

// call g at its assumed position in the stream: j and the number of redraws (Random.nextInt(bound), bound = i+1)
__global__ void __launch_bounds__(256)
draw_kernel(CallMap M, int64_t n_calls, uint64_t state0, const int64_t* __restrict__ offset, int32_t* __restrict__ j_out,
            int64_t* __restrict__ extra) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_calls) return;
    const int32_t b = bin_of(M.call_base, M.n_bins, g);
    const int64_t nb = M.bin_off[b + 1] - M.bin_off[b];
    const int32_t bound = (int32_t)(nb - (g - M.call_base[b]));   // i + 1 with i = last - k
    uint64_t s = affine_apply(lcg_jump((uint64_t)(g + offset[g])), state0);
    auto next31 = [&]() { s = (s * LCG_A + LCG_C) & LCG_MASK; return (int32_t)(uint32_t)(s >> 17); };
    int32_t r = next31();
    int64_t redraws = 0;
    const int32_t m = bound - 1;
    if ((bound & m) == 0) r = (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
    else {
        for (int32_t u = r;; u = next31(), redraws++) {
            r = u % bound;
            if ((int32_t)((uint32_t)u - (uint32_t)r + (uint32_t)m) >= 0) break;
        }
    }
    j_out[g] = r;
    extra[g] = redraws;
}

__global__ void __launch_bounds__(256)
differ_kernel(const int64_t* __restrict__ a, const int64_t* __restrict__ b, int64_t n, int32_t* __restrict__ flag) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n && a[g] != b[g]) *flag = 1;
}

// key of step y of a bin: (output position its j points at) << 32 | y
__global__ void __launch_bounds__(256)
pair_kernel(CallMap M, int64_t n_calls, const int32_t* __restrict__ j, uint64_t* __restrict__ key) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_calls) return;
    const int32_t b = bin_of(M.call_base, M.n_bins, g);
    const int64_t nb = M.bin_off[b + 1] - M.bin_off[b];
    const uint64_t y = (uint64_t)(nb - 1 - (g - M.call_base[b]));
    key[g] = ((uint64_t)(M.bin_off[b] + j[g]) << 32) | y;
}

// min{y > t : j_y = position X} of X's bin, as a global position; -1 if no later step writes X
__device__ __forceinline__ int64_t next_writer(const uint64_t* __restrict__ key, int64_t n_calls, int64_t X, int64_t t, int64_t bin_first) {
    const uint64_t want = ((uint64_t)X << 32) | (uint64_t)(t + 1);
    int64_t lo = 0, hi = n_calls;
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (key[mid] < want) lo = mid + 1; else hi = mid; }
    if (lo < n_calls && (int64_t)(key[lo] >> 32) == X) return bin_first + (int64_t)(key[lo] & 0xffffffffull);
    return -1;
}

__global__ void __launch_bounds__(256)
parent_kernel(CallMap M, int64_t n_pos, const uint64_t* __restrict__ key, int64_t n_calls, int64_t* __restrict__ nxt) {
    const int64_t X = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (X >= n_pos) return;
    const int32_t b = bin_of(M.bin_off, M.n_bins, X);
    const int64_t q = next_writer(key, n_calls, X, X - M.bin_off[b], M.bin_off[b]);
    nxt[X] = q < 0 ? X : q;
}

__global__ void __launch_bounds__(256)
jump_kernel(const int64_t* __restrict__ in, int64_t* __restrict__ out, int64_t n, int32_t* __restrict__ flag) {
    const int64_t X = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (X >= n) return;
    const int64_t a = in[X], b = in[a];
    out[X] = b;
    if (b != a) *flag = 1;
}

// output position X receives record order[src]
__global__ void __launch_bounds__(256)
final_kernel(CallMap M, int64_t n_pos, const uint64_t* __restrict__ key, int64_t n_calls, const int32_t* __restrict__ j,
             const int64_t* __restrict__ root, const int32_t* __restrict__ order, const int32_t* __restrict__ rec_bytes,
             int32_t* __restrict__ out_rec, int64_t* __restrict__ out_len) {
    const int64_t X = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (X > n_pos) return;
    if (X == n_pos) { out_len[X] = 0; return; }
    const int32_t b = bin_of(M.bin_off, M.n_bins, X);
    const int64_t first = M.bin_off[b], nb = M.bin_off[b + 1] - first, i = X - first;
    int64_t src = X;
    if (M.call_base[b + 1] > M.call_base[b]) {   // a shuffled bin
        const int64_t p = i > 0 ? (int64_t)j[M.call_base[b] + (nb - 1 - i)] : 0;
        const int64_t q = next_writer(key, n_calls, first + p, i, first);
        src = q < 0 ? first + p : root[q];
    }
    const int32_t rec = order[src];
    out_rec[X] = rec;
    out_len[X] = rec_bytes[rec];
}

__global__ void __launch_bounds__(256)
segment_kernel(int64_t n_pos, const int32_t* __restrict__ out_rec, const int64_t* __restrict__ out_off, const int64_t* __restrict__ rec_start,
               int64_t* __restrict__ seg_dst, int64_t* __restrict__ seg_src) {
    const int64_t X = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (X > n_pos + 1) return;
    if (X == n_pos) { seg_dst[X] = out_off[n_pos]; seg_src[X] = int64_t(SEG_PAD) << SEG_KIND_SHIFT; return; }
    if (X == n_pos + 1) { seg_dst[X] = INT64_MAX; seg_src[X] = int64_t(SEG_PAD) << SEG_KIND_SHIFT; return; }
    seg_dst[X] = out_off[X];
    seg_src[X] = (int64_t(SEG_REF_FWD) << SEG_KIND_SHIFT) | rec_start[out_rec[X]];
}

// A file the way BufferedReader.readLine sees it: "\r\n" and a lone '\r' end a line like '\n', and a last line without
// terminator still counts. Appended to `blob` with every line ending in exactly one '\n'.
void append_normalised(const std::string& path, std::vector<uint8_t>& blob) {
    const std::vector<char> buf = slurp(path);
    const size_t n = buf.size();
    if (n == 0) return;
    const size_t at = blob.size();
    if (!memchr(buf.data(), '\r', n)) {
        blob.insert(blob.end(), (const uint8_t*)buf.data(), (const uint8_t*)buf.data() + n);
    } else {
        blob.resize(at + n);
        size_t o = at;
        for (size_t i = 0; i < n; i++) {
            if (buf[i] == '\r') { blob[o++] = '\n'; if (i + 1 < n && buf[i + 1] == '\n') i++; }
            else blob[o++] = (uint8_t)buf[i];
        }
        blob.resize(o);
    }
    if (blob.back() != '\n') blob.push_back('\n');
}

struct LineIndex {
    DevBuf<uint8_t> d_blob;
    DevBuf<int64_t> d_block, d_term;
    DevBuf<uint8_t> d_tmp;
    int64_t n_bytes = 0, n_lines = 0;
};

// blob -> HBM, terminator positions
void build_line_index(sp_ctx* c, const std::vector<uint8_t>& blob, LineIndex& I) {
    cudaStream_t st = c->stream;
    I.n_bytes = (int64_t)blob.size();
    I.d_blob.reserve(std::max<size_t>(blob.size(), 1));
    if (I.n_bytes > 0) SP_CUDA(cudaMemcpyAsync(I.d_blob.p, blob.data(), blob.size(), cudaMemcpyHostToDevice, st));
    const int64_t nblk = (I.n_bytes + LI_BYTES - 1) / LI_BYTES;
    I.d_block.reserve((size_t)nblk + 1);
    I.n_lines = 0;
    if (nblk == 0) return;
    line_count_kernel<<<(unsigned)nblk, 256, 0, st>>>(I.d_blob.p, I.n_bytes, I.d_block.p);
    SP_CUDA(cudaMemsetAsync(I.d_block.p + nblk, 0, 8, st));
    size_t tmp = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp, I.d_block.p, I.d_block.p, (int)(nblk + 1), st);
    I.d_tmp.reserve(tmp + 256);
    tmp = I.d_tmp.cap;
    SP_CUDA(cub::DeviceScan::ExclusiveSum(I.d_tmp.p, tmp, I.d_block.p, I.d_block.p, (int)(nblk + 1), st));
    SP_CUDA(cudaMemcpyAsync(&I.n_lines, I.d_block.p + nblk, 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    I.d_term.reserve((size_t)std::max<int64_t>(I.n_lines, 1));
    line_scatter_kernel<<<(unsigned)nblk, 256, 0, st>>>(I.d_blob.p, I.n_bytes, I.d_block.p, I.d_term.p);
    c->launches += 3;
    SP_CUDA(cudaGetLastError());
}

void write_text(const std::string& path, const std::string& s) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw Error(SP_ERR_JAVA_EXCEPTION, "FileNotFoundException: " + path);
    const size_t w = fwrite(s.data(), 1, s.size(), f);
    fclose(f);
    if (w != s.size()) throw Error(SP_ERR, "IOException: short write " + path);
}

constexpr int MIN_READS_IN_LOCAL_MAXIMUM = 500, MIN_LOCAL_MAXIMUM = 1000, MAX_LOCAL_MAXIMUM = 25000;  // :14-16

// printStats :139-208 and std :219-239, on counts (= buffers_last + 1) and sequence bases per bin
std::string stats_text(const std::vector<int64_t>& count, const std::vector<int64_t>& bases, int32_t bin_length, int64_t genome_haploid) {
    const int n = (int)count.size(), half = bin_length >> 1;
    auto last_of = [&](int i) { return (int32_t)(count[(size_t)i] - 1); };   // buffers_last[i]
    std::vector<int32_t> smoothed((size_t)n);
    for (int i = 0; i < n; i++) {
        if (i < 1 || i >= n - 1) smoothed[(size_t)i] = last_of(i);
        else smoothed[(size_t)i] = (last_of(i - 1) + last_of(i) + last_of(i + 1)) / 3;
    }
    std::string out;
    int n_maxima = 0, last_max = -1, prev_max = -1;
    for (int i = 1; i < n - 1; i++) {
        if ((i + 1) * bin_length - 1 >= MIN_LOCAL_MAXIMUM && i * bin_length <= MAX_LOCAL_MAXIMUM && count[(size_t)i] >= MIN_READS_IN_LOCAL_MAXIMUM &&
            smoothed[(size_t)i] > smoothed[(size_t)i - 1] && smoothed[(size_t)i] > smoothed[(size_t)i + 1]) {
            n_maxima++; prev_max = last_max; last_max = i;
        }
    }
    out += std::to_string(n_maxima) + "\n";
    if (n_maxima < 2) return out;
    auto first_local_min = [&](int from, int to, int none) {
        for (int i = from; i < to; i++)
            if (smoothed[(size_t)i] < smoothed[(size_t)i - 1] && smoothed[(size_t)i] < smoothed[(size_t)i + 1]) return i;
        return none;
    };
    const int left_right_min = first_local_min(prev_max + 1, last_max, -1);
    const int left_min = first_local_min(1, prev_max, 0);
    auto one_sided_std = [&](int from, int to, int bin) {
        const double mean = bin * bin_length + half;
        double sd = 0.0, m = 0.0;
        for (int i = from; i <= to; i++) {
            m += (int32_t)count[(size_t)i];
            const double k = i * bin_length + half - mean;
            sd += (int32_t)count[(size_t)i] * k * k;
        }
        m = m + (int32_t)count[(size_t)bin] + m;
        sd = sd + 0 + sd;
        return std::sqrt(sd / m);
    };
    out += std::to_string(prev_max * bin_length + half) + "\n";
    out += java_double_to_string(one_sided_std(left_min, prev_max - 1, prev_max)) + "\n";
    out += std::to_string(last_max * bin_length + half) + "\n";
    out += java_double_to_string(one_sided_std(last_max + 1, n - 1, last_max)) + "\n";
    out += std::to_string(left_min) + "\n" + std::to_string(left_right_min) + "\n";
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += (int32_t)count[(size_t)i];
    double mass = 0.0;
    for (int i = left_min; i <= left_right_min; i++) mass += (int32_t)count[(size_t)i];
    out += java_double_to_string(mass / sum) + "\n";
    mass = 0.0;
    for (int i = left_right_min + 1; i < n; i++) mass += (int32_t)count[(size_t)i];
    out += java_double_to_string(mass / sum) + "\n";
    const double genome2 = (double)(genome_haploid << 1);
    sum = 0.0;
    for (int i = left_min; i <= left_right_min; i++) sum += (double)bases[(size_t)i];
    out += java_double_to_string(sum / genome2) + "\n";
    sum = 0.0;
    for (int i = left_right_min + 1; i < n; i++) sum += (double)bases[(size_t)i];
    out += java_double_to_string(sum / genome2) + "\n";
    return out;
}

double ms_between(cudaEvent_t a, cudaEvent_t b) { float ms = 0.f; cudaEventElapsedTime(&ms, a, b); return ms; }

}  // namespace

// BuildReadLengthBins.main (:21-105)
void build_read_length_bins(sp_ctx* c, const sp_build_bins_args* a) {
    if (!a || !a->input_files_list || !a->output_prefix) throw Error(SP_ERR_ARG, "null sp_build_bins_args");
    if (a->bin_length == 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArithmeticException: / by zero");
    const int32_t n_bins = (a->max_read_length + a->bin_length - 1) / a->bin_length;   // :28
    if (n_bins < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: " + std::to_string(n_bins));
    const int32_t half = a->bin_length >> 1;
    const std::string prefix = a->output_prefix;
    cudaStream_t st = c->stream;
    cudaEvent_t ev[5];
    for (auto& e : ev) SP_CUDA(cudaEventCreate(&e));
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 5; i++) cudaEventDestroy(e[i]); } } guard{ev};

    // the listed FASTQ files, back to back (the bins are first created empty, :44)
    for (int32_t i = 0; i < n_bins; i++) write_text(prefix + std::to_string(i) + ".bin", "");
    std::vector<uint8_t> blob;
    std::vector<int64_t> file_end;
    {
        const std::vector<char> list = slurp(a->input_files_list);
        Lines lines(list.data(), list.data() + list.size());
        const char *b, *e;
        while (lines.next(b, e)) { append_normalised(std::string(b, e), blob); file_end.push_back((int64_t)blob.size()); }
    }
    if (blob.size() >= ((size_t)1 << 40)) throw Error(SP_ERR_ARG, "BuildReadLengthBins: more than 1 TiB of FASTQ in one call");
    SP_CUDA(cudaEventRecord(ev[0], st));
    LineIndex I;
    build_line_index(c, blob, I);
    // a file whose line count is not a multiple of four ends inside a record: the Java dereferences a null line
    {
        std::vector<int64_t> h_term;   // only the terminators at the file ends are needed: count lines by binary search
        int64_t lines_before = 0;
        for (int64_t fe : file_end) {
            if (fe == 0) continue;
            // the file's last byte is a terminator: its index in term_pos + 1 = lines up to here. Found by bisection on the device array.
            int64_t lo = lines_before, hi = I.n_lines;
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                int64_t v;
                SP_CUDA(cudaMemcpyAsync(&v, I.d_term.p + mid, 8, cudaMemcpyDeviceToHost, st));
                SP_CUDA(cudaStreamSynchronize(st));
                if (v < fe - 1) lo = mid + 1; else hi = mid;
            }
            const int64_t upto = lo + 1;
            if ((upto - lines_before) % 4 != 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: a FASTQ file ends inside a record");
            lines_before = upto;
        }
    }
    const int64_t R = I.n_lines / 4;
    if (R > INT32_MAX) throw Error(SP_ERR_ARG, "BuildReadLengthBins: more than 2^31-1 records");
    if (R > 0 && n_bins == 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: Index -1 out of bounds for length 0");
    const size_t R1 = (size_t)std::max<int64_t>(R, 1), NB1 = (size_t)std::max(n_bins, 1);
    DevBuf<int64_t> d_rec_start, d_out_len, d_out_off, d_seg_dst, d_seg_src, d_seg_base, d_call_base, d_bin_off, d_off_a, d_off_b, d_extra, d_nxt_a, d_nxt_b;
    DevBuf<int32_t> d_rec_bytes, d_bin, d_bin2, d_idx, d_order, d_j, d_out_rec, d_flag;
    DevBuf<unsigned long long> d_hist;
    DevBuf<uint64_t> d_key, d_key2;
    DevBuf<uint8_t> d_tmp, d_out;
    d_rec_start.reserve(R1); d_rec_bytes.reserve(R1); d_bin.reserve(R1); d_bin2.reserve(R1); d_idx.reserve(R1); d_order.reserve(R1);
    d_hist.reserve(2 * NB1); d_flag.reserve(2);
    SP_CUDA(cudaMemsetAsync(d_hist.p, 0, 2 * NB1 * 8, st));
    SP_CUDA(cudaMemsetAsync(d_flag.p, 0, 8, st));
    std::vector<int64_t> count((size_t)n_bins, 0), bases((size_t)n_bins, 0);
    if (R > 0) {
        record_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(I.d_term.p, R, a->bin_length, n_bins, d_rec_start.p, d_rec_bytes.p, d_bin.p,
                                                                    d_idx.p, d_hist.p, d_hist.p + NB1, d_flag.p);
        int bits = 1;
        while ((1 << bits) < n_bins) bits++;
        size_t tmp = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tmp, d_bin.p, d_bin2.p, d_idx.p, d_order.p, (int)R, 0, bits, st);
        d_tmp.reserve(tmp + 256);
        tmp = d_tmp.cap;
        SP_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp, d_bin.p, d_bin2.p, d_idx.p, d_order.p, (int)R, 0, bits, st));   // stable: input order inside a bin
        c->launches += 2;
        std::vector<unsigned long long> h((size_t)2 * NB1);
        int32_t err = 0;
        SP_CUDA(cudaMemcpyAsync(h.data(), d_hist.p, h.size() * 8, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaMemcpyAsync(&err, d_flag.p, 4, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        if (err == 1) throw Error(SP_ERR_ARG, "BuildReadLengthBins: a record longer than 2^31-1 bytes");
        if (err == 2) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: negative bin");
        for (int32_t b = 0; b < n_bins; b++) { count[(size_t)b] = (int64_t)h[(size_t)b]; bases[(size_t)b] = (int64_t)h[NB1 + (size_t)b]; }
    }
    SP_CUDA(cudaEventRecord(ev[1], st));
    // <prefix>.histogram (:77-84) and <prefix>.stats
    {
        std::string hs;
        for (int32_t i = 0; i < n_bins; i++) hs += std::to_string(i * a->bin_length + half) + "," + std::to_string(count[(size_t)i]) + "\n";
        write_text(prefix + ".histogram", hs);
        if (n_bins > 0) write_text(prefix + ".stats", stats_text(count, bases, a->bin_length, a->genome_length_haploid));
        else throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException or ArrayIndexOutOfBoundsException in printStats with no bins");
    }
    if (R == 0) { c->bb_ms[0] = ms_between(ev[0], ev[1]); c->bb_ms[1] = c->bb_ms[2] = 0; c->bb_records = 0; return; }

    // the calls of the shuffles: bins with at least three records (`if (last<=1) continue`, :95), in bin order
    std::vector<int64_t> call_base((size_t)n_bins + 1, 0), bin_off((size_t)n_bins + 1, 0);
    for (int32_t b = 0; b < n_bins; b++) {
        bin_off[(size_t)b + 1] = bin_off[(size_t)b] + count[(size_t)b];
        call_base[(size_t)b + 1] = call_base[(size_t)b] + (count[(size_t)b] >= 3 ? count[(size_t)b] - 1 : 0);
    }
    const int64_t G = call_base[(size_t)n_bins];
    const size_t G1 = (size_t)std::max<int64_t>(G, 1);
    d_call_base.reserve((size_t)n_bins + 1); d_bin_off.reserve((size_t)n_bins + 1);
    SP_CUDA(cudaMemcpyAsync(d_call_base.p, call_base.data(), call_base.size() * 8, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemcpyAsync(d_bin_off.p, bin_off.data(), bin_off.size() * 8, cudaMemcpyHostToDevice, st));
    const CallMap M{d_call_base.p, d_bin_off.p, n_bins};
    d_j.reserve(G1); d_off_a.reserve(G1 + 1); d_off_b.reserve(G1 + 1); d_extra.reserve(G1 + 1); d_key.reserve(G1); d_key2.reserve(G1);
    d_nxt_a.reserve(R1); d_nxt_b.reserve(R1); d_out_rec.reserve(R1); d_out_len.reserve(R1 + 1); d_out_off.reserve(R1 + 1);
    const unsigned gg = (unsigned)((G + 255) / 256), gr = (unsigned)((R + 255) / 256);
    const uint64_t state0 = lcg_scramble(c->seed);
    const uint64_t* keys = d_key.p;
    if (G > 0) {
        size_t tmp_scan = 0, tmp_sort = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, d_extra.p, d_off_b.p, (int)(G + 1), st);
        cub::DeviceRadixSort::SortKeys(nullptr, tmp_sort, d_key.p, d_key2.p, (int)G, 0, 64, st);
        d_tmp.reserve(std::max(tmp_scan, tmp_sort) + 256);
        SP_CUDA(cudaMemsetAsync(d_off_a.p, 0, (G1 + 1) * 8, st));
        int64_t* cur = d_off_a.p;
        int64_t* nxt = d_off_b.p;
        for (int pass = 0;; pass++) {
            if (pass > 100000) throw Error(SP_ERR, "BuildReadLengthBins: the draw offsets did not settle");
            draw_kernel<<<gg, 256, 0, st>>>(M, G, state0, cur, d_j.p, d_extra.p);
            SP_CUDA(cudaMemsetAsync(d_extra.p + G, 0, 8, st));
            size_t tmp = d_tmp.cap;
            SP_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp, d_extra.p, nxt, (int)(G + 1), st));
            SP_CUDA(cudaMemsetAsync(d_flag.p, 0, 4, st));
            differ_kernel<<<gg, 256, 0, st>>>(cur, nxt, G, d_flag.p);
            c->launches += 3;
            int32_t moved = 0;
            SP_CUDA(cudaMemcpyAsync(&moved, d_flag.p, 4, cudaMemcpyDeviceToHost, st));
            SP_CUDA(cudaStreamSynchronize(st));
            std::swap(cur, nxt);
            if (!moved) break;
        }
        pair_kernel<<<gg, 256, 0, st>>>(M, G, d_j.p, d_key.p);
        size_t tmp = d_tmp.cap;
        cub::DoubleBuffer<uint64_t> kb(d_key.p, d_key2.p);
        SP_CUDA(cub::DeviceRadixSort::SortKeys(d_tmp.p, tmp, kb, (int)G, 0, 64, st));
        keys = kb.Current();
        c->launches += 2;
    }
    // the swap forest: parents, pointer jumping until every position points at the record that ends up written through it
    parent_kernel<<<gr, 256, 0, st>>>(M, R, keys, G, d_nxt_a.p);
    int64_t* na = d_nxt_a.p;
    int64_t* nb2 = d_nxt_b.p;
    for (int round = 0; round < 64; round++) {
        SP_CUDA(cudaMemsetAsync(d_flag.p, 0, 4, st));
        jump_kernel<<<gr, 256, 0, st>>>(na, nb2, R, d_flag.p);
        c->launches++;
        int32_t moved = 0;
        SP_CUDA(cudaMemcpyAsync(&moved, d_flag.p, 4, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        std::swap(na, nb2);
        if (!moved) break;
    }
    final_kernel<<<(unsigned)((R + 256) / 256), 256, 0, st>>>(M, R, keys, G, d_j.p, na, d_order.p, d_rec_bytes.p, d_out_rec.p, d_out_len.p);
    {
        size_t tmp = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp, d_out_len.p, d_out_off.p, (int)(R + 1), st);
        d_tmp.reserve(tmp + 256);
        tmp = d_tmp.cap;
        SP_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp, d_out_len.p, d_out_off.p, (int)(R + 1), st));
    }
    d_seg_dst.reserve((size_t)R + 2); d_seg_src.reserve((size_t)R + 2); d_seg_base.reserve(2);
    segment_kernel<<<(unsigned)((R + 2 + 255) / 256), 256, 0, st>>>(R, d_out_rec.p, d_out_off.p, d_rec_start.p, d_seg_dst.p, d_seg_src.p);
    c->launches += 4;
    SP_CUDA(cudaGetLastError());
    const int64_t seg_base[2] = {0, R + 2};
    SP_CUDA(cudaMemcpyAsync(d_seg_base.p, seg_base, sizeof seg_base, cudaMemcpyHostToDevice, st));
    // byte offset of every bin in the gathered output = offset of its first position
    std::vector<int64_t> h_bin_byte((size_t)n_bins + 1, 0);
    for (int32_t b = 0; b <= n_bins; b++)
        SP_CUDA(cudaMemcpyAsync(&h_bin_byte[(size_t)b], d_out_off.p + bin_off[(size_t)b], 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    const int64_t total = h_bin_byte[(size_t)n_bins];
    const int64_t stride = std::max<int64_t>(TILE_BYTES, (total + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES);
    d_out.reserve((size_t)stride, 0.0);
    reserve_tiled_copy(c, stride, 1, R + 2);   // allocations are not part of the timed gather
    SP_CUDA(cudaEventRecord(ev[2], st));
    launch_tiled_copy(c, I.d_blob.p, I.d_blob.p, I.d_blob.p, d_seg_dst.p, d_seg_src.p, d_seg_base.p, R + 2, d_out.p, stride, 1);
    SP_CUDA(cudaEventRecord(ev[3], st));
    {
        FileSink& sink = file_sink(c);
        for (int32_t b = 0; b < n_bins; b++)
            sink.write_file(prefix + std::to_string(b) + ".bin", d_out.p + h_bin_byte[(size_t)b], h_bin_byte[(size_t)b + 1] - h_bin_byte[(size_t)b], st);
        sink.finish();
    }
    SP_CUDA(cudaEventSynchronize(ev[3]));
    c->bb_ms[0] = ms_between(ev[0], ev[1]);
    c->bb_ms[1] = ms_between(ev[1], ev[2]);
    c->bb_ms[2] = ms_between(ev[2], ev[3]);
    c->bb_records = R;
    c->bb_bytes = total;
}

// Fastq2LengthHistogram.main (src/Fastq2LengthHistogram.java:11-51)
void fastq2length_histogram(sp_ctx* c, const char* fastq_file, int32_t bin_length, int32_t max_read_length, const char* out_hist, const char* out_max) {
    if (!fastq_file || !out_hist || !out_max) throw Error(SP_ERR_ARG, "Fastq2LengthHistogram: missing file argument");
    if (bin_length == 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArithmeticException: / by zero");
    const int32_t n_bins = (max_read_length + bin_length - 1) / bin_length;
    if (n_bins < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: " + std::to_string(n_bins));
    std::vector<uint8_t> blob;
    append_normalised(fastq_file, blob);
    // the loop reads a header and a sequence, then two lines that may be missing (:31-38): pad the last record
    LineIndex I;
    {
        int64_t lines = 0;
        for (uint8_t ch : blob) lines += ch == '\n';
        if (lines % 4 == 1) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: a header without a sequence line");
        while (lines % 4 != 0) { blob.push_back('\n'); lines++; }
    }
    build_line_index(c, blob, I);
    const int64_t R = I.n_lines / 4;
    if (R > INT32_MAX) throw Error(SP_ERR_ARG, "Fastq2LengthHistogram: more than 2^31-1 records");
    if (R > 0 && n_bins == 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: Index -1 out of bounds for length 0");
    cudaStream_t st = c->stream;
    const size_t R1 = (size_t)std::max<int64_t>(R, 1), NB1 = (size_t)std::max(n_bins, 1);
    DevBuf<int64_t> d_rec_start;
    DevBuf<int32_t> d_rec_bytes, d_bin, d_idx, d_flag;
    DevBuf<unsigned long long> d_hist;
    d_rec_start.reserve(R1); d_rec_bytes.reserve(R1); d_bin.reserve(R1); d_idx.reserve(R1); d_hist.reserve(2 * NB1); d_flag.reserve(2);
    SP_CUDA(cudaMemsetAsync(d_hist.p, 0, 2 * NB1 * 8, st));
    SP_CUDA(cudaMemsetAsync(d_flag.p, 0, 8, st));
    std::vector<unsigned long long> h((size_t)2 * NB1, 0);
    if (R > 0) {
        record_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(I.d_term.p, R, bin_length, n_bins, d_rec_start.p, d_rec_bytes.p, d_bin.p, d_idx.p,
                                                                    d_hist.p, d_hist.p + NB1, d_flag.p);
        c->launches++;
        int32_t err = 0;
        SP_CUDA(cudaMemcpyAsync(h.data(), d_hist.p, h.size() * 8, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaMemcpyAsync(&err, d_flag.p, 4, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        if (err) throw Error(err == 1 ? SP_ERR_ARG : SP_ERR_JAVA_EXCEPTION, err == 1 ? "a record longer than 2^31-1 bytes" : "ArrayIndexOutOfBoundsException: negative bin");
    }
    std::string mx, hs;
    for (int i = 1; i < n_bins - 1; i++) {
        const long long a = (long long)h[(size_t)i], l = (long long)h[(size_t)i - 1], r = (long long)h[(size_t)i + 1];
        if ((i + 1) * bin_length - 1 >= MIN_LOCAL_MAXIMUM && i * bin_length <= MAX_LOCAL_MAXIMUM && a + 1 >= MIN_READS_IN_LOCAL_MAXIMUM && a > l && a > r)
            mx += std::to_string(i * bin_length) + "," + std::to_string((i + 1) * bin_length - 1) + "\n";
    }
    for (int i = 0; i < n_bins; i++) hs += std::to_string(i * bin_length) + "," + std::to_string((long long)h[(size_t)i]) + "\n";
    write_text(out_max, mx);
    write_text(out_hist, hs);
}

}  // namespace sp

using sp::Error;

extern "C" {

int sp_build_read_length_bins(sp_ctx* ctx, const sp_build_bins_args* args) {
    SP_API_BEGIN
    sp::build_read_length_bins(ctx, args);
    SP_API_END
}

int sp_fastq2length_histogram(sp_ctx* ctx, const char* fastq_file, int32_t bin_length, int32_t max_read_length,
                              const char* output_file_histogram, const char* output_file_max) {
    SP_API_BEGIN
    sp::fastq2length_histogram(ctx, fastq_file, bin_length, max_read_length, output_file_histogram, output_file_max);
    SP_API_END
}

int sp_build_bins_timing(sp_ctx* ctx, double* index_ms, double* shuffle_ms, double* copy_ms, int64_t* n_records, int64_t* n_bytes) {
    SP_API_BEGIN
    if (index_ms) *index_ms = ctx->bb_ms[0];
    if (shuffle_ms) *shuffle_ms = ctx->bb_ms[1];
    if (copy_ms) *copy_ms = ctx->bb_ms[2];
    if (n_records) *n_records = ctx->bb_records;
    if (n_bytes) *n_bytes = ctx->bb_bytes;
    SP_API_END
}

}  // extern "C"
