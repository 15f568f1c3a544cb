// Bandwidth probes behind the design decisions of the tiled copy kernel (row M). Not on any product path: test hooks for
// tools/probe_copy_variants.py, which runs them in one call on the shape of bench.py's step (n_slots haplotype-sized slots
// filled from one shared source, region-major) so that "registers against the bulk-copy engine" and "how far is the kernel
// from what the memory system takes for a write-only stream" are measurements, not arguments:
//   kind 0  store only ........ 16-byte st.global.cs, no loads: the ceiling of a write-dominated stream
//   kind 1  register copy ..... the uniform path of materialise_kernel: two aligned 16-byte __ldg per chunk, funnel shifts,
//                               16-byte st.global.cs; one 16 KiB tile per CTA
//   kind 2  bulk copy ......... persistent CTAs, the 16-byte-aligned superset window of each tile brought to shared memory
//                               by cp.async.bulk (the TMA engine, SASS UBLKCP) under an mbarrier, three tiles in flight,
//                               realigned from shared memory, 16-byte st.global.cs
//   kind 3  bulk copy, warp-private pipelines: every warp owns a ring of 2 KiB windows and its own barriers (no CTA-wide
//                               barrier anywhere), lane 0 issues the warp's bulk copies
#include "common.cuh"

namespace sp {

constexpr int PB_TILE = 16384;

__device__ __forceinline__ uint4 pb_shift(uint4 A, uint4 B, unsigned m) {
    const unsigned s = (m & 3u) * 8u;
    uint32_t x0, x1, x2, x3, x4;
    switch (m >> 2) {
        case 0: x0 = A.x; x1 = A.y; x2 = A.z; x3 = A.w; x4 = B.x; break;
        case 1: x0 = A.y; x1 = A.z; x2 = A.w; x3 = B.x; x4 = B.y; break;
        case 2: x0 = A.z; x1 = A.w; x2 = B.x; x3 = B.y; x4 = B.z; break;
        default: x0 = A.w; x1 = B.x; x2 = B.y; x3 = B.z; x4 = B.w; break;
    }
    return make_uint4(__funnelshift_r(x0, x1, s), __funnelshift_r(x1, x2, s), __funnelshift_r(x2, x3, s), __funnelshift_r(x3, x4, s));
}

__global__ void __launch_bounds__(256, 5)
probe_store_kernel(uint8_t* __restrict__ out, int64_t stride, int64_t tiles_per_slot, int n_slots) {
    const int64_t b = blockIdx.x;
    const int hb = (int)(b % n_slots);
    const int64_t t = b / n_slots;
    uint4* o = (uint4*)(out + (int64_t)hb * stride + t * PB_TILE);
    const uint4 v = make_uint4((uint32_t)b, 0x41434754u, (uint32_t)t, threadIdx.x);
#pragma unroll
    for (int u = 0; u < 4; u++) __stcs(o + u * 256 + threadIdx.x, v);
}

__global__ void __launch_bounds__(256, 5)
probe_register_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ out, int64_t stride, int64_t tiles_per_slot, int n_slots, int misalign) {
    const int64_t b = blockIdx.x;
    const int hb = (int)(b % n_slots);
    const int64_t t = b / n_slots;
    const uint8_t* p = src + t * PB_TILE + misalign;
    const unsigned m = (unsigned)((uintptr_t)p & 15u);
    const uint4* q = (const uint4*)(p - m);
    uint4* o = (uint4*)(out + (int64_t)hb * stride + t * PB_TILE);
    const int tid = threadIdx.x;
    uint4 A[4], B[4];
#pragma unroll
    for (int u = 0; u < 4; u++) { A[u] = __ldg(q + u * 256 + tid); B[u] = __ldg(q + u * 256 + tid + 1); }
#pragma unroll
    for (int u = 0; u < 4; u++) __stcs(o + u * 256 + tid, m ? pb_shift(A[u], B[u], m) : A[u]);
}

// kind 7: store only, but every 16 bytes differ (a hash of the address) instead of three constants and the thread index
__global__ void __launch_bounds__(256, 5)
probe_store_hashed_kernel(uint8_t* __restrict__ out, int64_t stride, int64_t tiles_per_slot, int n_slots) {
    const int64_t b = blockIdx.x;
    const int hb = (int)(b % n_slots);
    const int64_t t = b / n_slots;
    uint4* o = (uint4*)(out + (int64_t)hb * stride + t * PB_TILE);
    uint32_t x = (uint32_t)b * 2654435761u ^ (threadIdx.x * 40503u + 0x9E3779B9u);
#pragma unroll
    for (int u = 0; u < 4; u++) {
        uint4 v;
        x ^= x << 13; x ^= x >> 17; x ^= x << 5; v.x = x;
        x ^= x << 13; x ^= x >> 17; x ^= x << 5; v.y = x;
        x ^= x << 13; x ^= x >> 17; x ^= x << 5; v.z = x;
        x ^= x << 13; x ^= x >> 17; x ^= x << 5; v.w = x;
        __stcs(o + u * 256 + threadIdx.x, v);
    }
}

// kinds 4-6: source-stationary. A CTA loads its 16 KiB source tile once and stores it into `group` slots (all of them, 25
// or 5): what the stream would cost if the 124 repeat reads of a reference tile never went to L2 at all.
__global__ void __launch_bounds__(256, 5)
probe_stationary_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ out, int64_t stride, int64_t tiles_per_slot, int n_slots, int group) {
    const int n_groups = (n_slots + group - 1) / group;
    const int64_t b = blockIdx.x;
    const int g = (int)(b % n_groups);
    const int64_t t = b / n_groups;
    const uint4* q = (const uint4*)(src + t * PB_TILE);
    const int tid = threadIdx.x;
    uint4 A[4];
#pragma unroll
    for (int u = 0; u < 4; u++) A[u] = __ldg(q + u * 256 + tid);
    const int h1 = min(n_slots, (g + 1) * group);
    for (int hb = g * group; hb < h1; hb++) {
        uint4* o = (uint4*)(out + (int64_t)hb * stride + t * PB_TILE);
#pragma unroll
        for (int u = 0; u < 4; u++) __stcs(o + u * 256 + tid, A[u]);
    }
}

__global__ void probe_fill_kernel(uint8_t* p, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = (uint8_t)(i % 251);
}

constexpr int PB_STAGES = 3;
constexpr int PB_WINDOW = PB_TILE + 16;   // the 16-byte-aligned superset of a misaligned tile

__device__ __forceinline__ uint32_t pb_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// returns false if the barrier did not flip within ~2^26 polls (a wrong phase must not hang the GPU)
__device__ __forceinline__ bool pb_wait(uint32_t bar, uint32_t phase) {
    for (uint32_t spin = 0; spin < (1u << 26); spin++) {
        uint32_t ok;
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(bar), "r"(phase) : "memory");
        if (ok) return true;
    }
    return false;
}

__global__ void __launch_bounds__(256, 4)
probe_bulk_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ out, int64_t stride, int64_t tiles_per_slot, int n_slots, int misalign,
                  int32_t* __restrict__ err) {
    extern __shared__ __align__(128) uint8_t pb_buf[];   // PB_STAGES windows, then PB_STAGES barriers
    uint64_t* bars = (uint64_t*)(pb_buf + PB_STAGES * PB_WINDOW);
    const int tid = threadIdx.x;
    const int64_t n_tiles = tiles_per_slot * n_slots;
    if (tid == 0) {
        for (int s = 0; s < PB_STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(pb_smem(bars + s)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int64_t tile_idx, int s) {   // one thread: arm the barrier with the byte count, start the bulk copy
        const int64_t t = tile_idx / n_slots;
        const uint8_t* p = src + t * PB_TILE + misalign;
        const uint8_t* g = p - ((uintptr_t)p & 15u);
        const uint32_t bar = pb_smem(bars + s), dst = pb_smem(pb_buf + s * PB_WINDOW);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)PB_WINDOW) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(g), "r"((uint32_t)PB_WINDOW), "r"(bar) : "memory");
    };
    // region-major walk with a grid stride: consecutive CTAs take the same tile index of different slots
    int64_t mine = blockIdx.x;
    if (tid == 0) {
        int64_t k = mine;
        for (int s = 0; s < PB_STAGES && k < n_tiles; s++, k += gridDim.x) issue(k, s);
    }
    const unsigned m = (unsigned)(((uintptr_t)src + (unsigned)misalign) & 15u);   // the same for every tile: tiles are 16 KiB apart
    int s = 0;
    uint32_t phase = 0;
    for (; mine < n_tiles; mine += gridDim.x) {
        if (!pb_wait(pb_smem(bars + s), phase)) { if (tid == 0) atomicExch(err, 1); return; }
        const int hb = (int)(mine % n_slots);
        const int64_t t = mine / n_slots;
        const uint4* w = (const uint4*)(pb_buf + s * PB_WINDOW);
        uint4* o = (uint4*)(out + (int64_t)hb * stride + t * PB_TILE);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint4 A = w[u * 256 + tid], B = w[u * 256 + tid + 1];
            __stcs(o + u * 256 + tid, m ? pb_shift(A, B, m) : A);
        }
        __syncthreads();   // every thread has read stage s: it may be refilled
        if (tid == 0) {
            const int64_t nxt = mine + (int64_t)PB_STAGES * gridDim.x;
            if (nxt < n_tiles) issue(nxt, s);
        }
        if (++s == PB_STAGES) { s = 0; phase ^= 1u; }
    }
}

constexpr int PW_SUB = 2048;             // output bytes per warp work unit
constexpr int PW_WINDOW = PW_SUB + 16;
constexpr int PW_STAGES = 3;

__global__ void __launch_bounds__(256, 4)
probe_bulk_warp_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ out, int64_t stride, int64_t tiles_per_slot, int n_slots, int misalign,
                       int32_t* __restrict__ err) {
    extern __shared__ __align__(128) uint8_t pb_buf[];   // per warp: PW_STAGES windows; then all barriers
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint8_t* ring = pb_buf + w * (PW_STAGES * PW_WINDOW);
    uint64_t* bars = (uint64_t*)(pb_buf + 8 * PW_STAGES * PW_WINDOW) + w * PW_STAGES;
    const int64_t n_tiles = tiles_per_slot * n_slots;
    if (lane == 0) {
        for (int s = 0; s < PW_STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(pb_smem(bars + s)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto issue = [&](int64_t tile_idx, int s) {
        const int64_t t = tile_idx / n_slots;
        const uint8_t* p = src + t * PB_TILE + w * PW_SUB + misalign;
        const uint8_t* g = p - ((uintptr_t)p & 15u);
        const uint32_t bar = pb_smem(bars + s), dst = pb_smem(ring + s * PW_WINDOW);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)PW_WINDOW) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(g), "r"((uint32_t)PW_WINDOW), "r"(bar) : "memory");
    };
    int64_t mine = blockIdx.x;
    if (lane == 0) {
        int64_t k = mine;
        for (int s = 0; s < PW_STAGES && k < n_tiles; s++, k += gridDim.x) issue(k, s);
    }
    const unsigned m = (unsigned)(((uintptr_t)src + (unsigned)misalign) & 15u);
    int s = 0;
    uint32_t phase = 0;
    for (; mine < n_tiles; mine += gridDim.x) {
        if (!pb_wait(pb_smem(bars + s), phase)) { if (lane == 0) atomicExch(err, 1); return; }
        const int hb = (int)(mine % n_slots);
        const int64_t t = mine / n_slots;
        const uint4* win = (const uint4*)(ring + s * PW_WINDOW);
        uint4* o = (uint4*)(out + (int64_t)hb * stride + t * PB_TILE + w * PW_SUB);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint4 A = win[u * 32 + lane], B = win[u * 32 + lane + 1];
            __stcs(o + u * 32 + lane, m ? pb_shift(A, B, m) : A);
        }
        __syncwarp();   // the warp has read stage s: it may be refilled
        if (lane == 0) {
            const int64_t nxt = mine + (int64_t)PW_STAGES * gridDim.x;
            if (nxt < n_tiles) issue(nxt, s);
        }
        if (++s == PW_STAGES) { s = 0; phase ^= 1u; }
    }
}

void probe_bandwidth(sp_ctx* c, int kind, int64_t slot_bytes, int n_slots, int misalign, int reps, double* ms_out) {
    if (slot_bytes <= 0 || n_slots <= 0 || reps <= 0 || misalign < 0 || misalign > 4096 || kind < 0 || kind > 7 || (kind >= 4 && misalign != 0))
        throw Error(SP_ERR_ARG, "probe: bad arguments");
    const int64_t tiles_per_slot = (slot_bytes + PB_TILE - 1) / PB_TILE, stride = tiles_per_slot * PB_TILE;
    const int64_t n_tiles = tiles_per_slot * n_slots;
    if (n_tiles >= INT32_MAX) throw Error(SP_ERR_ARG, "probe: too many tiles");
    DevBuf<uint8_t> src, out;
    DevBuf<int32_t> err;
    src.reserve((size_t)stride + 8192);
    out.reserve((size_t)stride * n_slots, 0.0);
    err.reserve(1);
    cudaStream_t st = c->stream;
    probe_fill_kernel<<<c->sm_count * 8, 256, 0, st>>>(src.p, stride + 8192);
    SP_CUDA(cudaMemsetAsync(err.p, 0, 4, st));
    cudaEvent_t e0, e1;
    SP_CUDA(cudaEventCreate(&e0)); SP_CUDA(cudaEventCreate(&e1));
    const size_t smem = PB_STAGES * PB_WINDOW + PB_STAGES * 8 + 64;
    if (kind == 2) SP_CUDA(cudaFuncSetAttribute(probe_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const size_t smem_w = 8 * PW_STAGES * PW_WINDOW + 8 * PW_STAGES * 8 + 64;
    if (kind == 3) SP_CUDA(cudaFuncSetAttribute(probe_bulk_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w));
    auto launch = [&]() {
        if (kind == 0) probe_store_kernel<<<(unsigned)n_tiles, 256, 0, st>>>(out.p, stride, tiles_per_slot, n_slots);
        else if (kind == 1) probe_register_kernel<<<(unsigned)n_tiles, 256, 0, st>>>(src.p, out.p, stride, tiles_per_slot, n_slots, misalign);
        else if (kind == 2) probe_bulk_kernel<<<c->sm_count * 4, 256, smem, st>>>(src.p, out.p, stride, tiles_per_slot, n_slots, misalign, err.p);
        else if (kind == 3) probe_bulk_warp_kernel<<<c->sm_count * 4, 256, smem_w, st>>>(src.p, out.p, stride, tiles_per_slot, n_slots, misalign, err.p);
        else if (kind == 7) probe_store_hashed_kernel<<<(unsigned)n_tiles, 256, 0, st>>>(out.p, stride, tiles_per_slot, n_slots);
        else {
            const int group = kind == 4 ? n_slots : (kind == 5 ? 25 : 5);
            const int64_t n_cta = tiles_per_slot * ((n_slots + group - 1) / group);
            probe_stationary_kernel<<<(unsigned)n_cta, 256, 0, st>>>(src.p, out.p, stride, tiles_per_slot, n_slots, group);
        }
        c->launches++;
    };
    launch();   // warm-up
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaEventRecord(e0, st));
    for (int r = 0; r < reps; r++) launch();
    SP_CUDA(cudaEventRecord(e1, st));
    SP_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    int32_t h_err = 0;
    SP_CUDA(cudaMemcpy(&h_err, err.p, 4, cudaMemcpyDeviceToHost));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (h_err) throw Error(SP_ERR, "probe: the bulk-copy barrier did not complete");
    if (kind >= 1 && kind != 7) {   // the copy must be right before its time means anything: out[slot][x] = src[x + misalign]
        uint8_t probe[4096];
        const int64_t at[3] = {0, (tiles_per_slot / 2) * PB_TILE + 5000, stride - 4096};
        for (int slot : {0, n_slots - 1})
            for (int64_t x0 : at) {
                SP_CUDA(cudaMemcpy(probe, out.p + (size_t)stride * slot + x0, sizeof probe, cudaMemcpyDeviceToHost));
                for (int i = 0; i < 4096; i++)
                    if (probe[i] != (uint8_t)((x0 + i + misalign) % 251)) throw Error(SP_ERR, "probe: wrong bytes copied at slot " + std::to_string(slot) + " offset " + std::to_string(x0 + i));
            }
    }
    if (ms_out) *ms_out = (double)ms / reps;
}

}  // namespace sp

extern "C" int sp_probe_bandwidth(sp_ctx* ctx, int32_t kind, int64_t slot_bytes, int32_t n_slots, int32_t misalign, int32_t reps, double* ms_per_launch) {
    SP_API_BEGIN
    sp::probe_bandwidth(ctx, kind, slot_bytes, n_slots, misalign, reps, ms_per_launch);
    SP_API_END
}
