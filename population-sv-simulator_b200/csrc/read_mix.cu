// SURVEY.md §8(f) rank 1 — the tail of docker/haplotype2reads.sh and the coverage chunks of docker/reads2svs.sh,
// kept on the device so the reads of an individual never make a host round trip between pbsim and the aligner:
//
//   cat reads_${ID2}.fa >> reads_${ID1}.fa        (haplotype2reads.sh:59)   records of ID1, then the records of ID2
//   shuf reads_${ID1}.fa > reads_${ID1}_${ID2}.1  (:61)                     a permutation of the LINES (one per read)
//   tr 'Z' '\n' < ... > ${OUTPUT_FILE}            (:63)                     ">S1_7Z<seq>\n" -> ">S1_7\n<seq>\n"
//   N_ROWS_1X = N_ROWS / MAX_COVERAGE, rounded up to a multiple of 4; split -d -l ${N_ROWS_1X}; coverage c = the
//   first c chunks                                (reads2svs.sh:65-74,98-128)
//
// `shuf` is unseeded in the reference, so its permutation cannot be reproduced; here it is DEFINED from a seed, in a
// way every thread can evaluate on its own: line k of the concatenated file (k counts kept records, ID1's first) gets
// the key  s_(k+1) = state of java.util.Random(shuffle_seed) after k+1 LCG steps  (jump-ahead, java_random.cuh), and
// the lines are written in ascending key order. An LCG of full period never repeats a state, so there are no ties.
//
//   mix_keys_kernel ...... key per record (dropped records get the sentinel 2^48 and sort to the end)
//   cub radix sort ....... 49-bit keys, record index as value (plumbing; ~2 M records per individual)
//   mix_len_kernel + cub exclusive scan ... output offset of every line
//   mix_segments_kernel .. three segments per record: name, the '\n' that replaces 'Z', sequence + '\n'
//   tiled copy ........... the materialisation kernel of row M on those segments (HBM-bound, 2 B per file byte)
#include "common.cuh"
#include "java_random.cuh"
#include "reads.h"
#include "file_sink.h"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>

namespace sp {

void launch_tiled_copy(sp_ctx* c, const uint8_t* src, const uint8_t* payload, const uint8_t* aux, const int64_t* seg_dst,
                       const int64_t* seg_src, const int64_t* seg_base, int64_t n_segments, uint8_t* out, int64_t stride, int n_slots);

struct MixSource {
    DevBuf<uint8_t> records, dropped;
    DevBuf<int64_t> rec_off;
    int64_t n_reads = 0, n_kept = 0, bytes = 0;
    int32_t id_len = 0;  // strlen(">" id_prefix ref_num "_")
};

struct ReadMixState {
    MixSource src[2];
    int n_sources = 0;
    DevBuf<uint64_t> d_key, d_key2;
    DevBuf<int32_t> d_idx, d_idx2;
    DevBuf<int64_t> d_len, d_out_off, d_seg_dst, d_seg_src, d_seg_base;
    DevBuf<uint8_t> d_aux, d_tmp, d_out;
    int64_t n_reads = 0, out_bytes = 0;
    std::vector<int64_t> h_out_off;
    bool mixed = false;
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    float order_ms = 0.f, copy_ms = 0.f;
};

void free_read_mix(sp_ctx* c) {
    if (!c->mix) return;
    for (auto& e : c->mix->ev) if (e) cudaEventDestroy(e);
    delete c->mix;
    c->mix = nullptr;
}

__device__ __forceinline__ int digits_of(int64_t v) {
    int n = 1;
    while (v >= 10) { v /= 10; n++; }
    return n;
}

// k = index in `cat` order over all simulated records; kept_rank[k] = its line number in the concatenated file (reads
// that contained NNNNN were never written, docker/pbsim2.cpp:1709-1711, so they do not count as lines).
__global__ void __launch_bounds__(256)
mix_keys_kernel(int64_t n0, int64_t n1, const uint8_t* __restrict__ dropped0, const uint8_t* __restrict__ dropped1,
                const int64_t* __restrict__ kept_rank, uint64_t state0, uint64_t* __restrict__ key, int32_t* __restrict__ idx) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n0 + n1) return;
    const bool d = k < n0 ? dropped0[k] != 0 : dropped1[k - n0] != 0;
    idx[k] = (int32_t)k;
    key[k] = d ? (1ull << 48) : affine_apply(lcg_jump((uint64_t)kept_rank[k] + 1), state0);
}

__global__ void __launch_bounds__(256)
mix_flag_kernel(int64_t n0, int64_t n1, const uint8_t* __restrict__ dropped0, const uint8_t* __restrict__ dropped1,
                int64_t* __restrict__ flag) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n0 + n1) return;
    flag[k] = k == n0 + n1 ? 0 : ((k < n0 ? dropped0[k] : dropped1[k - n0]) ? 0 : 1);
}

__global__ void __launch_bounds__(256)
mix_len_kernel(const int32_t* __restrict__ perm, int64_t n_kept, int64_t n0, const int64_t* __restrict__ off0,
               const int64_t* __restrict__ off1, int64_t* __restrict__ len) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_kept) return;
    if (i == n_kept) { len[i] = 0; return; }
    const int64_t k = perm[i];
    len[i] = k < n0 ? off0[k + 1] - off0[k] : off1[k - n0 + 1] - off1[k - n0];
}

__global__ void __launch_bounds__(256)
mix_segments_kernel(const int32_t* __restrict__ perm, int64_t n_kept, int64_t n0, const int64_t* __restrict__ off0,
                    const int64_t* __restrict__ off1, int32_t id_len0, int32_t id_len1, const int64_t* __restrict__ out_off,
                    int64_t* __restrict__ seg_dst, int64_t* __restrict__ seg_src) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_kept) return;
    if (i == n_kept) {
        seg_dst[3 * i] = out_off[n_kept]; seg_src[3 * i] = int64_t(SEG_PAD) << SEG_KIND_SHIFT;
        seg_dst[3 * i + 1] = INT64_MAX;   seg_src[3 * i + 1] = int64_t(SEG_PAD) << SEG_KIND_SHIFT;
        return;
    }
    const int64_t k = perm[i];
    const bool second = k >= n0;
    const int64_t r = second ? k - n0 : k;
    const int64_t a = second ? off1[r] : off0[r];
    const int64_t name = (second ? id_len1 : id_len0) + digits_of(r + 1);  // ">S1_" + read number (docker/pbsim2.cpp:1710)
    const int64_t kind = int64_t(second ? SEG_PAYLOAD : SEG_REF_FWD) << SEG_KIND_SHIFT;
    const int64_t dst = out_off[i];
    seg_dst[3 * i] = dst;                seg_src[3 * i] = kind | a;
    seg_dst[3 * i + 1] = dst + name;     seg_src[3 * i + 1] = (int64_t(SEG_AUX) << SEG_KIND_SHIFT) | 0;  // 'Z' -> '\n'
    seg_dst[3 * i + 2] = dst + name + 1; seg_src[3 * i + 2] = kind | (a + name + 1);
}

// ------------------------------------------------------------------------------------------------- host
static ReadMixState& mix_of(sp_ctx* c) {
    if (!c->mix) c->mix = new ReadMixState();
    return *c->mix;
}

static void mix_begin(sp_ctx* c) {
    ReadMixState& M = mix_of(c);
    M.n_sources = 0; M.mixed = false; M.n_reads = 0; M.out_bytes = 0; M.h_out_off.clear();
}

// Takes the device records of the last sp_reads_simulate_* call (no copy: the buffers change owner).
static void mix_add(sp_ctx* c) {
    if (!c->reads || !c->reads->configured) throw Error(SP_ERR_ARG, "sp_reads_configure has not been called");
    mix_add_from(c, *c->reads);
}

void mix_add_from(sp_ctx* c, ReadState& R) {
    ReadMixState& M = mix_of(c);
    if (M.n_sources >= 2) throw Error(SP_ERR_ARG, "read mix: an individual has two haplotypes");
    MixSource& S = M.src[M.n_sources++];
    S.records.swap(R.d_out); S.dropped.swap(R.d_dropped); S.rec_off.swap(R.d_rec_off);
    S.n_reads = R.n_reads; S.n_kept = R.n_reads - R.n_dropped; S.bytes = R.out_bytes;
    S.id_len = (int32_t)(1 + R.par.id_prefix.size() + std::to_string(R.ref_num).size() + 1);
    R.n_reads = 0; R.n_dropped = 0; R.out_bytes = 0; R.emitted_bases = 0; R.results_on_host = false;
    R.h_len.clear(); R.h_rec_off.assign(1, 0); R.plan_on_device = false;
    M.mixed = false;
}

static void mix_finish(sp_ctx* c, int64_t shuffle_seed) {
    ReadMixState& M = mix_of(c);
    if (M.n_sources == 0) throw Error(SP_ERR_ARG, "read mix: no reads added");
    cudaStream_t st = c->stream;
    const MixSource& A = M.src[0];
    const MixSource& B = M.src[M.n_sources > 1 ? 1 : 0];
    const int64_t n0 = A.n_reads, n1 = M.n_sources > 1 ? B.n_reads : 0;
    const int64_t n = n0 + n1, n_kept = A.n_kept + (M.n_sources > 1 ? B.n_kept : 0);
    if (n >= (int64_t)INT32_MAX) throw Error(SP_ERR_ARG, "read mix: too many reads");
    M.n_reads = n_kept;
    M.out_bytes = A.bytes + (M.n_sources > 1 ? B.bytes : 0);
    for (auto& e : M.ev) if (!e) SP_CUDA(cudaEventCreate(&e));
    const int64_t stride = std::max<int64_t>(TILE_BYTES, (M.out_bytes + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES);
    M.d_out.reserve((size_t)stride, 0.02);
    SP_CUDA(cudaEventRecord(M.ev[0], st));
    const size_t n1s = (size_t)std::max<int64_t>(n, 1);
    M.d_key.reserve(n1s); M.d_key2.reserve(n1s); M.d_idx.reserve(n1s); M.d_idx2.reserve(n1s);
    M.d_len.reserve(n1s + 1); M.d_out_off.reserve(n1s + 1);
    const unsigned gb = (unsigned)((n + 256) / 256);
    size_t tmp_sort = 0, tmp_scan = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, M.d_key.p, M.d_key2.p, M.d_idx.p, M.d_idx2.p, (int)n, 0, 49, st);
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, M.d_len.p, M.d_out_off.p, (int)(n + 1), st);
    M.d_tmp.reserve(std::max(tmp_sort, tmp_scan) + 256);
    size_t tmp_bytes = M.d_tmp.cap;
    // line numbers of the kept records in `cat` order (exclusive scan of the kept flags; d_len/d_out_off are free here)
    mix_flag_kernel<<<gb, 256, 0, st>>>(n0, n1, A.dropped.p, B.dropped.p, M.d_len.p);
    SP_CUDA(cub::DeviceScan::ExclusiveSum(M.d_tmp.p, tmp_bytes, M.d_len.p, M.d_out_off.p, (int)(n + 1), st));
    mix_keys_kernel<<<gb, 256, 0, st>>>(n0, n1, A.dropped.p, B.dropped.p, M.d_out_off.p, lcg_scramble(shuffle_seed), M.d_key.p, M.d_idx.p);
    tmp_bytes = M.d_tmp.cap;
    SP_CUDA(cub::DeviceRadixSort::SortPairs(M.d_tmp.p, tmp_bytes, M.d_key.p, M.d_key2.p, M.d_idx.p, M.d_idx2.p, (int)n, 0, 49, st));
    const unsigned gk = (unsigned)((n_kept + 256) / 256);
    mix_len_kernel<<<gk, 256, 0, st>>>(M.d_idx2.p, n_kept, n0, A.rec_off.p, B.rec_off.p, M.d_len.p);
    tmp_bytes = M.d_tmp.cap;
    SP_CUDA(cub::DeviceScan::ExclusiveSum(M.d_tmp.p, tmp_bytes, M.d_len.p, M.d_out_off.p, (int)(n_kept + 1), st));
    const int64_t nseg = 3 * n_kept + 2;
    M.d_seg_dst.reserve((size_t)nseg); M.d_seg_src.reserve((size_t)nseg); M.d_seg_base.reserve(2);
    M.d_aux.reserve(16);
    const int64_t seg_base[2] = {0, nseg};
    const uint8_t nl = '\n';
    SP_CUDA(cudaMemcpyAsync(M.d_seg_base.p, seg_base, sizeof seg_base, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemcpyAsync(M.d_aux.p, &nl, 1, cudaMemcpyHostToDevice, st));
    mix_segments_kernel<<<gk, 256, 0, st>>>(M.d_idx2.p, n_kept, n0, A.rec_off.p, B.rec_off.p, A.id_len, B.id_len, M.d_out_off.p,
                                            M.d_seg_dst.p, M.d_seg_src.p);
    c->launches += 4 + 3;  // four kernels of this file + cub's scans and sort (counted as one each)
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaEventRecord(M.ev[1], st));
    launch_tiled_copy(c, A.records.p, B.records.p, M.d_aux.p, M.d_seg_dst.p, M.d_seg_src.p, M.d_seg_base.p, nseg, M.d_out.p, stride, 1);
    SP_CUDA(cudaEventRecord(M.ev[2], st));
    int64_t total = 0;
    SP_CUDA(cudaMemcpyAsync(&total, M.d_out_off.p + n_kept, 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    SP_CUDA(cudaEventElapsedTime(&M.order_ms, M.ev[0], M.ev[1]));
    SP_CUDA(cudaEventElapsedTime(&M.copy_ms, M.ev[1], M.ev[2]));
    if (total != M.out_bytes) throw Error(SP_ERR, "read mix: " + std::to_string(total) + " bytes laid out, " + std::to_string(M.out_bytes) + " expected");
    M.h_out_off.clear();
    M.mixed = true;
}

static ReadMixState& mixed_of(sp_ctx* c) {
    if (!c->mix || !c->mix->mixed) throw Error(SP_ERR_ARG, "sp_reads_mix_finish has not been called");
    return *c->mix;
}

static void fetch_offsets(sp_ctx* c, ReadMixState& M) {
    if (!M.h_out_off.empty()) return;
    M.h_out_off.assign((size_t)M.n_reads + 1, 0);
    SP_CUDA(cudaMemcpyAsync(M.h_out_off.data(), M.d_out_off.p, ((size_t)M.n_reads + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
}

}  // namespace sp

using sp::Error;


extern "C" {

int sp_reads_mix_begin(sp_ctx* ctx) {
    SP_API_BEGIN
    sp::mix_begin(ctx);
    SP_API_END
}

int sp_reads_mix_add(sp_ctx* ctx) {
    SP_API_BEGIN
    sp::mix_add(ctx);
    SP_API_END
}

int sp_reads_mix_finish(sp_ctx* ctx, int64_t shuffle_seed, int64_t* n_reads, int64_t* file_bytes) {
    SP_API_BEGIN
    sp::mix_finish(ctx, shuffle_seed);
    if (n_reads) *n_reads = ctx->mix->n_reads;
    if (file_bytes) *file_bytes = ctx->mix->out_bytes;
    SP_API_END
}

// reads2svs.sh:65-74: N_ROWS = wc -l (two rows per read after tr), N_ROWS_1X = N_ROWS / MAX_COVERAGE, made a multiple
// of 4 by rounding up; :98-128: coverage c = chunk-00 .. chunk-(c-1) = the first c*N_ROWS_1X rows of the file.
int sp_reads_mix_coverage_prefix(sp_ctx* ctx, int32_t max_coverage, int32_t coverage, int64_t* n_rows_1x, int64_t* n_rows,
                                 int64_t* n_bytes) {
    SP_API_BEGIN
    sp::ReadMixState& M = sp::mixed_of(ctx);
    if (max_coverage <= 0 || coverage < 0) throw Error(SP_ERR_ARG, "coverage prefix: MAX_COVERAGE must be positive, COVERAGE non-negative");
    const int64_t rows_total = 2 * M.n_reads;
    int64_t rows_1x = rows_total / max_coverage;
    if (rows_1x % 4 != 0) rows_1x = (rows_1x / 4 + 1) * 4;
    const int64_t rows = std::min<int64_t>(rows_total, (int64_t)coverage * rows_1x);
    sp::fetch_offsets(ctx, M);
    if (n_rows_1x) *n_rows_1x = rows_1x;
    if (n_rows) *n_rows = rows;
    if (n_bytes) *n_bytes = M.h_out_off[(size_t)(rows / 2)];
    SP_API_END
}

int sp_reads_mix_copy(sp_ctx* ctx, int64_t offset, int64_t n_bytes, uint8_t* host_out) {
    SP_API_BEGIN
    sp::ReadMixState& M = sp::mixed_of(ctx);
    if (offset < 0 || n_bytes < 0 || offset + n_bytes > M.out_bytes) throw Error(SP_ERR_ARG, "read mix: byte range outside the file");
    if (n_bytes > 0) SP_CUDA(cudaMemcpyAsync(host_out, M.d_out.p + offset, (size_t)n_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SP_CUDA(cudaStreamSynchronize(ctx->stream));
    SP_API_END
}

int sp_reads_mix_device(sp_ctx* ctx, void** file_bytes, void** read_offsets, int64_t* n_reads, int64_t* n_bytes) {
    SP_API_BEGIN
    sp::ReadMixState& M = sp::mixed_of(ctx);
    if (file_bytes) *file_bytes = M.d_out.p;
    if (read_offsets) *read_offsets = M.d_out_off.p;
    if (n_reads) *n_reads = M.n_reads;
    if (n_bytes) *n_bytes = M.out_bytes;
    SP_API_END
}

int sp_reads_mix_offsets(sp_ctx* ctx, int64_t* read_offsets) {
    SP_API_BEGIN
    sp::ReadMixState& M = sp::mixed_of(ctx);
    sp::fetch_offsets(ctx, M);
    memcpy(read_offsets, M.h_out_off.data(), M.h_out_off.size() * 8);
    SP_API_END
}

int sp_reads_mix_timing(sp_ctx* ctx, double* order_ms, double* copy_ms) {
    SP_API_BEGIN
    sp::ReadMixState& M = sp::mixed_of(ctx);
    if (order_ms) *order_ms = M.order_ms;
    if (copy_ms) *copy_ms = M.copy_ms;
    SP_API_END
}

uint32_t sp_haplotype_read_seed(uint32_t seed, int32_t haplotype) {
    uint64_t z = (uint64_t)seed * 0x9E3779B97F4A7C15ULL + ((uint64_t)(uint32_t)haplotype + 1) * 0xD6E8FEB86659FD93ULL;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27; z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return (uint32_t)(1 + z % 2147483646ULL);
}

// docker/haplotype2reads.sh:53-64 for the individual whose haplotypes are id1 and id1+1: pbsim on both haplotypes
// (each its own process in the reference, hence ref_num 1 and an independent seed for each), cat, shuf, tr, one file.
int sp_haplotype2reads(sp_ctx* ctx, int32_t id1, int64_t shuffle_seed, const char* output_file, int64_t* n_reads,
                       int64_t* file_bytes) {
    SP_API_BEGIN
    int rc = sp_plan_haplotypes(ctx, id1, id1 + 1, 0, nullptr, nullptr);
    if (rc == SP_OK) rc = sp_materialise(ctx);
    if (rc == SP_OK) rc = sp_reads_mix_begin(ctx);
    if (rc == SP_OK && (!ctx->reads || !ctx->reads->configured)) throw Error(SP_ERR_ARG, "sp_reads_configure has not been called");
    // one pbsim run per haplotype, each with its own seed (sp_haplotype_read_seed), both read kernels in flight
    if (rc == SP_OK) rc = sp_reads_mix_add_pair(ctx, id1, id1 + 1, 1, 1, 1, nullptr, nullptr, nullptr, nullptr, nullptr);
    if (rc == SP_OK) rc = sp_reads_mix_finish(ctx, shuffle_seed, n_reads, file_bytes);
    if (rc != SP_OK) return rc;
    if (output_file) {   // device -> pinned ring -> file, the writer thread overlapping the copies (csrc/file_sink.h)
        sp::ReadMixState& M = *ctx->mix;
        sp::FileSink& sink = sp::file_sink(ctx);
        sink.write_file(output_file, M.d_out.p, M.out_bytes, ctx->stream);
        sink.finish();
    }
    SP_API_END
}

}  // extern "C"
