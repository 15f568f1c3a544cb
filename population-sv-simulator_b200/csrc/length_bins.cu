// Row D of SURVEY.md §8a — SampleReadsFromLengthBins.main (reference src/SampleReadsFromLengthBins.java:19-81),
// re-designed for B200. The reference draws one nextDouble() per attempt, maps it to a length bin through the cumulative
// bimodal table, takes the next unread record of that bin (a bin that ran dry makes the attempt fail) and stops when
// the sequence bases taken reach the coverage target. Every attempt costs exactly two LCG steps, so attempt t is
// index-addressable (jump-ahead), and "next unread record of its bin" is the rank of t among the attempts that drew the
// same bin:
//   bin_draw_kernel ....... attempt t -> draw (jump-ahead) -> Arrays.binarySearch(cumulativeBimodal) -> bin id;
//                           per-CTA bin histogram in shared memory
//   bin_base_kernel ....... exclusive scan of the CTA histograms per bin
//   bin_rank_kernel ....... rank of every attempt inside its bin (warp match + per-warp counters), acceptance
//                           (rank < records in the bin), record id; per-CTA sums of {accepted, bases, bytes}
//   block_scan_kernel ..... scan of the per-CTA sums
//   cut_kernel ............ prefixes per attempt; the cut (first accepted attempt whose inclusive base count reaches the
//                           target, :55/:73) and the first run of N_BINS*10 failed attempts (:56-67, exit status 3)
//   emit_kernel ........... kept attempts -> "@read<k>\n" header text + two segments per read (header, record body),
//                           histogram[bin]++ (:68)
//   tiled copy ............ the materialisation kernel of row M on those segments (HBM-bound, 2 B per output byte)
#include "common.cuh"
#include "java_random.cuh"
#include "textio.h"
#include <algorithm>

namespace sp {

void launch_tiled_copy(sp_ctx* c, const uint8_t* src, const uint8_t* payload, const uint8_t* aux, const int64_t* seg_dst,
                       const int64_t* seg_src, const int64_t* seg_base, int64_t n_segments, uint8_t* out, int64_t stride, int n_slots);

constexpr int LB_CTA = 256;  // attempts per CTA

struct LengthBinState {
    int32_t n_bins = 0;
    int64_t n_records = 0, blob_bytes = 0, total_seq = 0;
    std::vector<int64_t> h_bin_first;  // n_bins+1: first record of each bin
    std::vector<int32_t> h_bin_starts; // records a bin can start (complete ones + a truncated tail, if any)
    DevBuf<uint8_t> d_blob;            // the bin files back to back
    DevBuf<int64_t> d_bin_first;
    DevBuf<int32_t> d_bin_complete, d_bin_starts;
    DevBuf<int64_t> d_rec_off;         // byte offset of the record's 2nd line (the sequence) in the blob
    DevBuf<int32_t> d_rec_seq, d_rec_bytes;  // sequence length; bytes of "seq\nline3\nline4\n"
    // one sampling run
    DevBuf<double> d_cdf;
    DevBuf<int32_t> d_bin, d_rec, d_hist_blocks, d_hist;
    DevBuf<int64_t> d_block_sums, d_acc_incl, d_cov_incl, d_bytes_excl, d_result;
    DevBuf<int64_t> d_seg_dst, d_seg_src, d_seg_base;
    DevBuf<uint8_t> d_hdr, d_out;
    int64_t n_reads = 0, out_bytes = 0, attempts = 0, coverage = 0;
    std::vector<int32_t> h_hist;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    float sample_ms = 0.f, copy_ms = 0.f;
};

void free_length_bins(sp_ctx* c) {
    if (c->bins) {
        if (c->bins->ev0) { cudaEventDestroy(c->bins->ev0); cudaEventDestroy(c->bins->ev1); cudaEventDestroy(c->bins->ev2); }
        delete c->bins;
    }
    c->bins = nullptr;
}

// Arrays.binarySearch(double[] a, double key) on [0, n) for finite keys, mapped as :58-59.
__device__ __forceinline__ int bin_of_draw(const double* cdf, int n, double key) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        const double v = cdf[mid];
        if (v < key) lo = mid + 1;
        else if (v > key) hi = mid - 1;
        else return mid;  // found: `else if (bin==N_BINS)` can never fire here, mid < n
    }
    return lo;  // -1-(-(lo+1)); may equal n: the reference then indexes bins[N_BINS] and throws
}

__global__ void __launch_bounds__(LB_CTA)
bin_draw_kernel(const double* __restrict__ cdf_g, int n_bins, uint64_t state0, int64_t n_attempts, int32_t* __restrict__ bin,
                int32_t* __restrict__ hist_blocks, int64_t* __restrict__ result) {
    extern __shared__ __align__(8) uint8_t lb_smem[];
    double* cdf = (double*)lb_smem;
    int32_t* hist = (int32_t*)(cdf + n_bins);
    for (int i = threadIdx.x; i < n_bins; i += LB_CTA) { cdf[i] = cdf_g[i]; hist[i] = 0; }
    __syncthreads();
    const int64_t t = (int64_t)blockIdx.x * LB_CTA + threadIdx.x;
    if (t < n_attempts) {
        const uint64_t s = affine_apply(lcg_jump(2ull * (uint64_t)t), state0);
        const double key = (double)(int64_t)draw53_from_state(s) * 0x1.0p-53;  // nextDouble()
        const int b = bin_of_draw(cdf, n_bins, key);
        bin[t] = b;
        if (b >= n_bins) atomicMin((unsigned long long*)&result[2], (unsigned long long)t);  // first out-of-range attempt
        else atomicAdd(&hist[b], 1);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_bins; i += LB_CTA) hist_blocks[(int64_t)blockIdx.x * n_bins + i] = hist[i];
}

// hist_blocks[blk][b] -> number of attempts with bin b in earlier CTAs (in place)
__global__ void __launch_bounds__(128)
bin_base_kernel(int32_t* __restrict__ hist_blocks, int n_bins, int n_blocks) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_bins) return;
    int32_t run = 0;
    for (int k = 0; k < n_blocks; k++) {
        const int32_t v = hist_blocks[(int64_t)k * n_bins + b];
        hist_blocks[(int64_t)k * n_bins + b] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(LB_CTA)
bin_rank_kernel(const int32_t* __restrict__ bin, const int32_t* __restrict__ hist_blocks, int n_bins, int64_t n_attempts,
                const int64_t* __restrict__ bin_first, const int32_t* __restrict__ bin_starts,
                const int32_t* __restrict__ rec_seq, const int32_t* __restrict__ rec_bytes, int32_t* __restrict__ rec,
                int64_t* __restrict__ block_sums) {
    extern __shared__ __align__(8) uint8_t lb_smem[];
    int32_t* wh = (int32_t*)lb_smem;  // [LB_CTA/32][n_bins]
    __shared__ int64_t red[3][LB_CTA / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (LB_CTA / 32) * n_bins; i += LB_CTA) wh[i] = 0;
    __syncthreads();
    const int64_t t = (int64_t)blockIdx.x * LB_CTA + threadIdx.x;
    const int b = t < n_attempts ? bin[t] : -1;
    const bool live = b >= 0 && b < n_bins;
    const unsigned peers = __match_any_sync(0xffffffffu, live ? b : -1 - lane);
    const int in_warp = __popc(peers & ((1u << lane) - 1));
    if (live && in_warp == 0) wh[w * n_bins + b] = __popc(peers);
    __syncthreads();
    int64_t acc = 0, cov = 0, bytes = 0;
    int32_t r = -1;
    if (live) {
        int32_t rank = hist_blocks[(int64_t)blockIdx.x * n_bins + b] + in_warp;
        for (int k = 0; k < w; k++) rank += wh[k * n_bins + b];
        if (rank < bin_starts[b]) {  // bins[bin].read() != -1 (:60-61)
            r = (int32_t)(bin_first[b] + rank);
            acc = 1; cov = rec_seq[r]; bytes = rec_bytes[r];
        }
    }
    if (t < n_attempts) rec[t] = r;
    for (int d = 16; d > 0; d >>= 1) {
        acc += __shfl_down_sync(0xffffffffu, acc, d);
        cov += __shfl_down_sync(0xffffffffu, cov, d);
        bytes += __shfl_down_sync(0xffffffffu, bytes, d);
    }
    if (lane == 0) { red[0][w] = acc; red[1][w] = cov; red[2][w] = bytes; }
    __syncthreads();
    if (threadIdx.x < 3) {
        int64_t s = 0;
        for (int k = 0; k < LB_CTA / 32; k++) s += red[threadIdx.x][k];
        block_sums[(int64_t)blockIdx.x * 3 + threadIdx.x] = s;
    }
}

// exclusive scan of block_sums[n][3] in place, single CTA
__global__ void __launch_bounds__(1024)
block_scan_kernel(int64_t* __restrict__ sums, int n) {
    __shared__ int64_t ws[3][32];
    __shared__ int64_t carry[3];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 3) carry[threadIdx.x] = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int k = base + threadIdx.x;
        int64_t v[3], incl[3];
        for (int q = 0; q < 3; q++) { v[q] = k < n ? sums[(int64_t)k * 3 + q] : 0; incl[q] = v[q]; }
        for (int d = 1; d < 32; d <<= 1)
            for (int q = 0; q < 3; q++) {
                const int64_t o = __shfl_up_sync(0xffffffffu, incl[q], d);
                if (lane >= d) incl[q] += o;
            }
        if (lane == 31) for (int q = 0; q < 3; q++) ws[q][warp] = incl[q];
        __syncthreads();
        for (int q = 0; q < 3; q++) {
            int64_t wp = 0;
            for (int u = 0; u < warp; u++) wp += ws[q][u];
            if (k < n) sums[(int64_t)k * 3 + q] = carry[q] + wp + incl[q] - v[q];
        }
        __syncthreads();
        if (threadIdx.x < 3) {
            int64_t tot = 0;
            for (int u = 0; u < 32; u++) tot += ws[threadIdx.x][u];
            carry[threadIdx.x] += tot;
        }
        __syncthreads();
    }
}

// bytes of the headers "@read<i>\n" of reads 0..k-1 (:69)
__host__ __device__ inline int64_t header_bytes_before(int64_t k) {
    int64_t total = 6 * k, lo = 0, p = 10;
    for (int d = 1; d <= 18 && lo < k; d++) {
        const int64_t hi = k < p ? k : p;
        total += (int64_t)d * (hi - lo);
        lo = p; p *= 10;
    }
    return total;
}

// result: [0] cut attempt, [1] first attempt that ends a run of max_fail failures, [2] first out-of-range bin,
//         [3] first truncated record taken, [4] total output bytes, [5] reads, [6] bases
__global__ void __launch_bounds__(LB_CTA)
cut_kernel(const int32_t* __restrict__ rec, const int32_t* __restrict__ rec_seq, const int32_t* __restrict__ rec_bytes,
           const int64_t* __restrict__ block_sums, int64_t n_attempts, int64_t target_bp, int64_t* __restrict__ acc_incl,
           int64_t* __restrict__ cov_incl, int64_t* __restrict__ bytes_excl, int64_t* __restrict__ result) {
    __shared__ int64_t ws[3][LB_CTA / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t t = (int64_t)blockIdx.x * LB_CTA + threadIdx.x;
    const int32_t r = t < n_attempts ? rec[t] : -1;
    int64_t v[3] = {0, 0, 0}, incl[3];
    if (r >= 0) { v[0] = 1; v[1] = rec_seq[r]; v[2] = rec_bytes[r]; }
    for (int q = 0; q < 3; q++) incl[q] = v[q];
    for (int d = 1; d < 32; d <<= 1)
        for (int q = 0; q < 3; q++) {
            const int64_t o = __shfl_up_sync(0xffffffffu, incl[q], d);
            if (lane >= d) incl[q] += o;
        }
    if (lane == 31) for (int q = 0; q < 3; q++) ws[q][w] = incl[q];
    __syncthreads();
    for (int q = 0; q < 3; q++) {
        int64_t wp = block_sums[(int64_t)blockIdx.x * 3 + q];
        for (int u = 0; u < w; u++) wp += ws[q][u];
        incl[q] += wp;
    }
    if (t >= n_attempts) return;
    acc_incl[t] = incl[0];
    cov_incl[t] = incl[1];
    bytes_excl[t] = incl[2] - v[2];
    if (r >= 0) {
        if (incl[1] >= target_bp && incl[1] - v[1] < target_bp) {  // exactly one attempt satisfies this
            result[0] = t; result[5] = incl[0]; result[6] = incl[1];
            result[4] = header_bytes_before(incl[0]) + incl[2];
        }
        if (v[2] < 0) atomicMin((unsigned long long*)&result[3], (unsigned long long)t);
    }
}

__global__ void __launch_bounds__(256)
fail_run_kernel(const int64_t* __restrict__ acc_incl, int64_t n_attempts, int64_t max_fail, int64_t* __restrict__ result) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_attempts || t + 1 < max_fail) return;
    const int64_t before = t - max_fail >= 0 ? acc_incl[t - max_fail] : 0;
    if (acc_incl[t] == before) atomicMin((unsigned long long*)&result[1], (unsigned long long)t);
}

__global__ void __launch_bounds__(256)
emit_kernel(const int32_t* __restrict__ bin, const int32_t* __restrict__ rec, const int64_t* __restrict__ rec_off,
            const int32_t* __restrict__ rec_seq, const int64_t* __restrict__ acc_incl, const int64_t* __restrict__ cov_incl,
            const int64_t* __restrict__ bytes_excl, const int32_t* __restrict__ rec_bytes, int64_t t_cut,
            int64_t* __restrict__ seg_dst, int64_t* __restrict__ seg_src, uint8_t* __restrict__ hdr,
            int32_t* __restrict__ hist) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t > t_cut) return;
    const int32_t r = rec[t];
    if (r < 0) return;
    const int64_t k = acc_incl[t] - 1;
    const int64_t h0 = header_bytes_before(k);
    // "@read" + k + "\n"
    uint8_t* p = hdr + h0;
    p[0] = '@'; p[1] = 'r'; p[2] = 'e'; p[3] = 'a'; p[4] = 'd';
    char dig[20];
    int nd = 0;
    int64_t x = k;
    do { dig[nd++] = (char)('0' + x % 10); x /= 10; } while (x > 0);
    for (int i = 0; i < nd; i++) p[5 + i] = (uint8_t)dig[nd - 1 - i];
    p[5 + nd] = '\n';
    const int64_t hl = 6 + nd;
    const int64_t dst = h0 + bytes_excl[t];
    seg_dst[2 * k] = dst;
    seg_src[2 * k] = (int64_t(SEG_AUX) << SEG_KIND_SHIFT) | h0;
    seg_dst[2 * k + 1] = dst + hl;
    seg_src[2 * k + 1] = (int64_t(SEG_REF_FWD) << SEG_KIND_SHIFT) | rec_off[r];
    atomicAdd(&hist[bin[t]], 1);
    if (t == t_cut) {
        const int64_t total = dst + hl + rec_bytes[r];
        seg_dst[2 * k + 2] = total;     seg_src[2 * k + 2] = int64_t(SEG_PAD) << SEG_KIND_SHIFT;
        seg_dst[2 * k + 3] = INT64_MAX; seg_src[2 * k + 3] = int64_t(SEG_PAD) << SEG_KIND_SHIFT;
    }
}

// ------------------------------------------------------------------------------------------------- host
static LengthBinState& bins_of(sp_ctx* c) {
    if (!c->bins || c->bins->n_bins <= 0) throw Error(SP_ERR_ARG, "no length bins loaded (sp_length_bins_load / sp_length_bins_set)");
    return *c->bins;
}

// Index the records of the bin files (4 lines each: name, sequence, separator, qualities — the format
// BuildReadLengthBins writes, src/BuildReadLengthBins.java:42-105) and upload blob + tables.
static void set_bins(sp_ctx* c, int32_t n_bins, const uint8_t* blob, const int64_t* bin_off) {
    if (n_bins <= 0 || n_bins > 4096) throw Error(SP_ERR_ARG, "N_BINS must be in [1, 4096]");
    if (!c->bins) c->bins = new LengthBinState();
    LengthBinState& B = *c->bins;
    B.n_bins = n_bins;
    std::vector<int64_t> rec_off;
    std::vector<int32_t> rec_seq, rec_bytes, complete(n_bins), starts(n_bins);
    B.h_bin_first.assign(n_bins + 1, 0);
    // device copy of the blob: every bin is followed by one '\n' so that a last line without a line terminator still
    // yields the bytes BufferedWriter.newLine() would add
    std::vector<int64_t> dev_off(n_bins + 1, 0);
    for (int b = 0; b < n_bins; b++) dev_off[b + 1] = dev_off[b] + (bin_off[b + 1] - bin_off[b]) + 1;
    B.blob_bytes = dev_off[n_bins];
    B.d_blob.reserve((size_t)B.blob_bytes + 64, 0.0);
    cudaStream_t st = c->stream;
    static const uint8_t nl = '\n';
    for (int b = 0; b < n_bins; b++) {
        const uint8_t* p = blob + bin_off[b];
        const int64_t n = bin_off[b + 1] - bin_off[b];
        if (n > 0) SP_CUDA(cudaMemcpyAsync(B.d_blob.p + dev_off[b], p, (size_t)n, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(B.d_blob.p + dev_off[b] + n, &nl, 1, cudaMemcpyHostToDevice, st));
        B.h_bin_first[b] = (int64_t)rec_off.size();
        int64_t q = 0;
        int32_t nrec = 0;
        bool truncated = false;
        while (q < n) {
            int64_t ls[5];  // starts of the 4 lines and the end
            ls[0] = q;
            int k = 0;
            for (; k < 4 && q < n; k++) {
                const uint8_t* e = (const uint8_t*)memchr(p + q, '\n', (size_t)(n - q));
                if (e && e > p + q && e[-1] == '\r') throw Error(SP_ERR_ARG, "length bins with \\r\\n line ends are not supported");
                q = e ? (e - p) + 1 : n + 1;  // a last line without '\n' ends at the '\n' appended on the device
                ls[k + 1] = q;
            }
            if (k < 4) {  // the reference would take this record and die with a NullPointerException (:70-73)
                truncated = true;
                rec_off.push_back(dev_off[b] + (k >= 2 ? ls[1] : ls[0]));
                rec_seq.push_back(k >= 2 ? (int32_t)(ls[2] - ls[1] - 1) : 0);
                rec_bytes.push_back(-1);
                break;
            }
            if (ls[2] - ls[1] - 1 > INT32_MAX / 4) throw Error(SP_ERR_ARG, "read longer than 2^29 in a length bin");
            rec_off.push_back(dev_off[b] + ls[1]);
            rec_seq.push_back((int32_t)(ls[2] - ls[1] - 1));
            rec_bytes.push_back((int32_t)(ls[4] - ls[1]));
            nrec++;
        }
        complete[b] = nrec;
        starts[b] = nrec + (truncated ? 1 : 0);
    }
    B.h_bin_first[n_bins] = (int64_t)rec_off.size();
    B.total_seq = 0;
    for (int32_t v : rec_seq) B.total_seq += v;
    B.h_bin_starts = starts;
    B.n_records = (int64_t)rec_off.size();
    if (B.n_records >= INT32_MAX) throw Error(SP_ERR_ARG, "too many records in the length bins");
    const size_t nr = (size_t)std::max<int64_t>(B.n_records, 1);
    B.d_bin_first.reserve(n_bins + 1); B.d_bin_complete.reserve(n_bins); B.d_bin_starts.reserve(n_bins);
    B.d_rec_off.reserve(nr); B.d_rec_seq.reserve(nr); B.d_rec_bytes.reserve(nr);
    SP_CUDA(cudaMemcpyAsync(B.d_bin_first.p, B.h_bin_first.data(), (n_bins + 1) * 8, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemcpyAsync(B.d_bin_complete.p, complete.data(), n_bins * 4, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemcpyAsync(B.d_bin_starts.p, starts.data(), n_bins * 4, cudaMemcpyHostToDevice, st));
    if (B.n_records > 0) {
        SP_CUDA(cudaMemcpyAsync(B.d_rec_off.p, rec_off.data(), B.n_records * 8, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(B.d_rec_seq.p, rec_seq.data(), B.n_records * 4, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(B.d_rec_bytes.p, rec_bytes.data(), B.n_records * 4, cudaMemcpyHostToDevice, st));
    }
    SP_CUDA(cudaStreamSynchronize(st));
}

// The sampling loop of main (:53-76) for a fresh `new Random(seed)`. Returns SP_OK or SP_ERR_COVERAGE (exit status 3).
static int sample_bins(sp_ctx* c, const double* cdf, int32_t n_bins, int64_t target_bp) {
    LengthBinState& B = bins_of(c);
    if (n_bins != B.n_bins) throw Error(SP_ERR_ARG, "cumulative table has " + std::to_string(n_bins) + " entries, " + std::to_string(B.n_bins) + " bins are loaded");
    if (!cdf) throw Error(SP_ERR_ARG, "null cumulative table");
    cudaStream_t st = c->stream;
    if (!B.ev0) { SP_CUDA(cudaEventCreate(&B.ev0)); SP_CUDA(cudaEventCreate(&B.ev1)); SP_CUDA(cudaEventCreate(&B.ev2)); }
    B.n_reads = 0; B.out_bytes = 0; B.attempts = 0; B.coverage = 0;
    B.h_hist.assign(n_bins, 0);
    B.sample_ms = B.copy_ms = 0.f;
    if (target_bp <= 0) return SP_OK;  // while (coverage<TARGET_COVERAGE_BP) never entered: empty FASTQ, zero histogram
    B.d_cdf.reserve(n_bins); B.d_hist.reserve(n_bins); B.d_result.reserve(8);
    SP_CUDA(cudaMemcpyAsync(B.d_cdf.p, cdf, n_bins * 8, cudaMemcpyHostToDevice, st));
    const int64_t max_fail = (int64_t)n_bins * 10;  // MAX_SAMPLING_ATTEMPTS (:38)
    const uint64_t state0 = lcg_scramble(c->seed);
    // attempts to look at: every record could be taken once, plus one failing run; start from an estimate and grow
    const int64_t T_max = B.n_records + (B.n_records + 1) * max_fail + 1;  // beyond this the loop must have ended
    const double mean_seq = B.n_records > 0 ? std::max(1.0, (double)B.total_seq / (double)B.n_records) : 1.0;
    int64_t T = std::min<int64_t>(T_max, std::max<int64_t>(1 << 14, (int64_t)((double)target_bp / mean_seq * 1.25) + max_fail));
    const size_t rank_smem = (size_t)n_bins * 4 * (LB_CTA / 32);
    if (rank_smem > 48 * 1024) SP_CUDA(cudaFuncSetAttribute(bin_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rank_smem));
    SP_CUDA(cudaEventRecord(B.ev0, st));
    int64_t res[8];
    const int64_t NONE = INT64_MAX;
    for (;;) {
        const int nblk = (int)((T + LB_CTA - 1) / LB_CTA);
        B.d_bin.reserve(T); B.d_rec.reserve(T); B.d_hist_blocks.reserve((size_t)nblk * n_bins);
        B.d_block_sums.reserve((size_t)nblk * 3); B.d_acc_incl.reserve(T); B.d_cov_incl.reserve(T); B.d_bytes_excl.reserve(T);
        for (int i = 0; i < 8; i++) res[i] = i < 4 ? NONE : 0;
        SP_CUDA(cudaMemcpyAsync(B.d_result.p, res, sizeof res, cudaMemcpyHostToDevice, st));
        bin_draw_kernel<<<nblk, LB_CTA, (size_t)n_bins * 12, st>>>(B.d_cdf.p, n_bins, state0, T, B.d_bin.p, B.d_hist_blocks.p, B.d_result.p);
        bin_base_kernel<<<(n_bins + 127) / 128, 128, 0, st>>>(B.d_hist_blocks.p, n_bins, nblk);
        bin_rank_kernel<<<nblk, LB_CTA, rank_smem, st>>>(B.d_bin.p, B.d_hist_blocks.p, n_bins, T, B.d_bin_first.p,
                                                                            B.d_bin_starts.p, B.d_rec_seq.p, B.d_rec_bytes.p, B.d_rec.p, B.d_block_sums.p);
        block_scan_kernel<<<1, 1024, 0, st>>>(B.d_block_sums.p, nblk);
        cut_kernel<<<nblk, LB_CTA, 0, st>>>(B.d_rec.p, B.d_rec_seq.p, B.d_rec_bytes.p, B.d_block_sums.p, T, target_bp, B.d_acc_incl.p,
                                            B.d_cov_incl.p, B.d_bytes_excl.p, B.d_result.p);
        fail_run_kernel<<<(unsigned)((T + 255) / 256), 256, 0, st>>>(B.d_acc_incl.p, T, max_fail, B.d_result.p);
        c->launches += 6;
        SP_CUDA(cudaGetLastError());
        SP_CUDA(cudaMemcpyAsync(res, B.d_result.p, sizeof res, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        if (res[0] != NONE || res[1] != NONE || res[2] != NONE) break;
        if (T >= T_max) throw Error(SP_ERR, "length-bin sampling did not terminate");
        T = std::min<int64_t>(T * 4, T_max);
    }
    // what happens first in attempt order?
    const int64_t t_cut = res[0], t_fail = res[1], t_oob = res[2], t_trunc = res[3];
    const int64_t t_end = std::min(t_cut, t_fail);
    if (t_oob <= t_end)
        throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: bins[" + std::to_string(n_bins) + "] (a draw above the last cumulative value, attempt " + std::to_string(t_oob) + ")");
    if (t_trunc <= t_end) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: truncated record in a length bin (attempt " + std::to_string(t_trunc) + ")");
    if (t_fail < t_cut) { B.attempts = t_fail + 1; return SP_ERR_COVERAGE; }
    B.attempts = t_cut + 1; B.out_bytes = res[4]; B.n_reads = res[5]; B.coverage = res[6];
    const int64_t nseg = 2 * B.n_reads + 2;
    const int64_t stride = (B.out_bytes + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES;
    B.d_seg_dst.reserve(nseg); B.d_seg_src.reserve(nseg); B.d_seg_base.reserve(2);
    B.d_hdr.reserve((size_t)header_bytes_before(B.n_reads) + 64);
    B.d_out.reserve((size_t)stride, 0.02);
    const int64_t seg_base[2] = {0, nseg};
    SP_CUDA(cudaMemcpyAsync(B.d_seg_base.p, seg_base, sizeof seg_base, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemsetAsync(B.d_hist.p, 0, n_bins * 4, st));
    emit_kernel<<<(unsigned)((t_cut + 256) / 256), 256, 0, st>>>(B.d_bin.p, B.d_rec.p, B.d_rec_off.p, B.d_rec_seq.p, B.d_acc_incl.p,
                                                                B.d_cov_incl.p, B.d_bytes_excl.p, B.d_rec_bytes.p, t_cut, B.d_seg_dst.p,
                                                                B.d_seg_src.p, B.d_hdr.p, B.d_hist.p);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaEventRecord(B.ev1, st));
    launch_tiled_copy(c, B.d_blob.p, B.d_blob.p, B.d_hdr.p, B.d_seg_dst.p, B.d_seg_src.p, B.d_seg_base.p, nseg, B.d_out.p, stride, 1);
    SP_CUDA(cudaEventRecord(B.ev2, st));
    SP_CUDA(cudaMemcpyAsync(B.h_hist.data(), B.d_hist.p, n_bins * 4, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    SP_CUDA(cudaEventElapsedTime(&B.sample_ms, B.ev0, B.ev1));
    SP_CUDA(cudaEventElapsedTime(&B.copy_ms, B.ev1, B.ev2));
    return SP_OK;
}

}  // namespace sp

using sp::Error;


extern "C" {

int sp_length_bins_set(sp_ctx* ctx, int32_t n_bins, const uint8_t* blob, const int64_t* bin_off) {
    SP_API_BEGIN
    if (!blob || !bin_off) throw Error(SP_ERR_ARG, "null bin buffers");
    sp::set_bins(ctx, n_bins, blob, bin_off);
    SP_API_END
}

int sp_length_bins_load(sp_ctx* ctx, const char* bins_prefix, int32_t n_bins) {
    SP_API_BEGIN
    if (!bins_prefix) throw Error(SP_ERR_ARG, "null bins prefix");
    if (n_bins <= 0 || n_bins > 4096) throw Error(SP_ERR_ARG, "N_BINS must be in [1, 4096]");
    std::vector<uint8_t> blob;
    std::vector<int64_t> off(n_bins + 1, 0);
    for (int i = 0; i < n_bins; i++) {  // new FileReader(BINS_PREFIX+i+".bin") (:49)
        const std::vector<char> f = sp::slurp(std::string(bins_prefix) + std::to_string(i) + ".bin");
        blob.insert(blob.end(), f.begin(), f.end());
        off[i + 1] = (int64_t)blob.size();
    }
    if (blob.empty()) blob.push_back(0);
    sp::set_bins(ctx, n_bins, blob.data(), off.data());
    SP_API_END
}

int sp_length_bins_sample(sp_ctx* ctx, const double* cumulative, int32_t n_bins, int64_t target_coverage_bp, int64_t* n_reads,
                          int64_t* fastq_bytes, int64_t* attempts, int64_t* bases, int32_t* histogram) {
    if (!ctx) return SP_ERR_ARG;
    try {
        cudaSetDevice(ctx->device);
        const int rc = sp::sample_bins(ctx, cumulative, n_bins, target_coverage_bp);
        sp::LengthBinState& B = *ctx->bins;
        if (n_reads) *n_reads = B.n_reads;
        if (fastq_bytes) *fastq_bytes = B.out_bytes;
        if (attempts) *attempts = B.attempts;
        if (bases) *bases = B.coverage;
        if (histogram && rc == SP_OK) memcpy(histogram, B.h_hist.data(), (size_t)n_bins * 4);
        if (rc == SP_ERR_COVERAGE) ctx->err = "Not enough reads to sample the target coverage from the given bins and distribution.";
        return rc;
    }
    catch (const sp::Error& e) { ctx->err = e.what(); return e.code; }
    catch (const std::exception& e) { ctx->err = e.what(); return SP_ERR; }
}

int sp_length_bins_copy(sp_ctx* ctx, uint8_t* host_out, int64_t capacity, int64_t* written) {
    SP_API_BEGIN
    sp::LengthBinState& B = sp::bins_of(ctx);
    if (B.out_bytes > capacity) throw Error(SP_ERR_ARG, "host buffer too small");
    if (B.out_bytes > 0) {
        SP_CUDA(cudaMemcpyAsync(host_out, B.d_out.p, (size_t)B.out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
        SP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    if (written) *written = B.out_bytes;
    SP_API_END
}

int sp_length_bins_timing(sp_ctx* ctx, double* sample_ms, double* copy_ms) {
    SP_API_BEGIN
    sp::LengthBinState& B = sp::bins_of(ctx);
    if (sample_ms) *sample_ms = B.sample_ms;
    if (copy_ms) *copy_ms = B.copy_ms;
    SP_API_END
}

int sp_sample_reads_from_bins(sp_ctx* ctx, const sp_length_bin_args* a) {
    if (!ctx) return SP_ERR_ARG;
    try {
        cudaSetDevice(ctx->device);
        if (!a || !a->bins_prefix || !a->output_fastq_file || !a->cumulative) throw Error(SP_ERR_ARG, "null argument");
        if (a->bin_length <= 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArithmeticException: / by zero");
        const int32_t n_bins = (a->max_read_length + a->bin_length - 1) / a->bin_length;  // :36
        const int64_t target = (int64_t)((double)a->genome_length_each_haplotype * (a->target_coverage_each_haplotype * 2));  // :37
        int rc = sp_length_bins_load(ctx, a->bins_prefix, n_bins);
        if (rc != SP_OK) return rc;
        // new FileWriter(OUTPUT_FASTQ_FILE) happens before the loop (:53): the file exists even when the run fails
        FILE* f = fopen(a->output_fastq_file, "wb");
        if (!f) throw Error(SP_ERR_JAVA_EXCEPTION, std::string("FileNotFoundException: ") + a->output_fastq_file);
        rc = sp::sample_bins(ctx, a->cumulative, n_bins, target);
        if (rc == SP_ERR_COVERAGE) {
            fclose(f);
            ctx->err = "Not enough reads to sample " + sp::java_double_to_string(a->target_coverage_each_haplotype) +
                       "x coverage for each haplotype from the given bins and distribution.";
            return rc;
        }
        sp::LengthBinState& B = *ctx->bins;
        std::vector<uint8_t> host((size_t)std::max<int64_t>(B.out_bytes, 1));
        int64_t w = 0;
        rc = sp_length_bins_copy(ctx, host.data(), (int64_t)host.size(), &w);
        if (rc != SP_OK) { fclose(f); return rc; }
        const size_t wr = fwrite(host.data(), 1, (size_t)w, f);
        fclose(f);
        if (wr != (size_t)w) throw Error(SP_ERR, "IOException: short write");
        f = fopen((std::string(a->output_fastq_file) + ".histogram").c_str(), "wb");  // :78-80
        if (!f) throw Error(SP_ERR_JAVA_EXCEPTION, "FileNotFoundException: histogram");
        for (int i = 0; i < n_bins; i++) fprintf(f, "%d\n", B.h_hist[i]);
        fclose(f);
        return SP_OK;
    }
    catch (const sp::Error& e) { ctx->err = e.what(); return e.code; }
    catch (const std::exception& e) { ctx->err = e.what(); return SP_ERR; }
}

}  // extern "C"
