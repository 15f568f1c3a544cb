// Row M of SURVEY.md §8a — haplotype materialisation (SimulateHaplotypes.buildHaplotypes,
// reference src/SimulateHaplotypes.java:917-960), re-designed for B200:
//
//   plan_kernel ......... one CTA per haplotype; every edit of the sorted edit list becomes three segment
//                         slots (reference gap, body, second DUP copy) and a block-wide exclusive scan over the
//                         per-edit output deltas yields each segment's destination offset (the "segmented scan").
//   tile_count/fill ..... one thread per 16 KiB output tile: the tile's RUNS (pieces of one segment inside the tile),
//                         the first one in the 16-byte tile descriptor, the others in a run array (count, scan, fill).
//   materialise_kernel .. one CTA per output tile, tiles ordered region-major across the haplotypes of the
//                         batch so that the reference window they share stays in the 126 MB L2. A run is a plain
//                         copy: 16-byte aligned streaming stores, the arbitrary byte shift between source and
//                         destination resolved in registers (two aligned 16-byte loads + funnel shifts), a few byte
//                         stores at the ends of a run that does not start or end on a 16-byte boundary. No segment
//                         table, no search, no barrier in the copy kernel.
//
// HBM-bound: 1 B read + 1 B written per output base (2 B/bp algorithmic).
#include "common.cuh"
#include <cub/device/device_scan.cuh>

namespace sp {

__device__ __forceinline__ uint8_t complement_base(uint8_t c) {  // reverseComplement, :886-896
    switch (c) {
        case 'a': return 't'; case 'c': return 'g'; case 'g': return 'c'; case 't': return 'a';
        case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A';
        default: return 'N';
    }
}

// Exclusive scan of one int64 per thread over a 256-thread CTA; returns the exclusive prefix, total in `total`.
__device__ __forceinline__ int64_t block_exclusive_scan_256(int64_t v, int64_t* warp_sums /*[8]*/, int64_t& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    int64_t warp_prefix = 0, t = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        int64_t s = warp_sums[w];
        if (w < warp) warp_prefix += s;
        t += s;
    }
    total = t;
    __syncthreads();
    return warp_prefix + incl - v;
}

// err[0] = code (1 bad variant id, 2 negative gap / start>end, 3 end beyond reference, 4 bad type), err[1] = haplotype,
// err[2] = index in list.
__global__ void __launch_bounds__(256)
plan_kernel(const int8_t* __restrict__ type, const int32_t* __restrict__ pos, const int32_t* __restrict__ len,
            const int64_t* __restrict__ payload_off, int32_t n_variants, const int64_t* __restrict__ list_off,
            const int32_t* __restrict__ ids, int64_t ref_len, int first_hap, const int64_t* __restrict__ aux_off,
            const int64_t* __restrict__ seg_base, int64_t* __restrict__ seg_dst, int64_t* __restrict__ seg_src,
            int64_t* __restrict__ hap_len, int32_t* __restrict__ err) {
    __shared__ int64_t warp_sums[8];
    __shared__ int64_t carry_s;
    const int hb = blockIdx.x, h = first_hap + hb, tid = threadIdx.x;  // list_off points at the batch's first row
    const int64_t lo = list_off[hb], n = list_off[hb + 1] - lo;
    const int64_t sb = seg_base[hb];
    const int64_t hdr_off = aux_off[hb], hdr_len = aux_off[hb + 1] - hdr_off;
    if (tid == 0) {
        seg_dst[sb] = 0;
        seg_src[sb] = (int64_t(SEG_AUX) << SEG_KIND_SHIFT) | hdr_off;
        carry_s = hdr_len;
    }
    __syncthreads();
    auto fail = [&](int code, int64_t k) {
        if (atomicCAS(&err[0], 0, code) == 0) { err[1] = h; err[2] = (int32_t)k; }
    };
    for (int64_t base = 0; base < n; base += 256) {
        const int64_t k = base + tid;
        int64_t gap = 0, b1 = 0, b2 = 0, src_gap = 0, src_b1 = 0, src_b2 = 0;
        if (k < n) {
            const int32_t id = ids[lo + k];
            if (id < 0 || id >= n_variants) fail(1, k);
            else {
                const int t = type[id];
                const int64_t p = pos[id], l = len[id];
                int64_t prev_after = 0;
                if (k > 0) {
                    const int32_t pid = ids[lo + k - 1];
                    if (pid >= 0 && pid < n_variants)
                        prev_after = (type[pid] == SP_INS) ? (int64_t)pos[pid] : (int64_t)pos[pid] + len[pid];
                }
                gap = p - prev_after;
                // out.append(reference,firstInReference,variant.position) throws unless 0<=start<=end<=length
                if (gap < 0 || prev_after < 0) { fail(2, k); gap = 0; }
                else if (p > ref_len) { fail(3, k); gap = 0; }
                src_gap = (int64_t(SEG_REF_FWD) << SEG_KIND_SHIFT) | (prev_after & SEG_OFF_MASK);
                if (t == SP_DEL) {
                } else if (t == SP_DUP || t == SP_INV) {
                    if (l < 0) fail(2, k);
                    else if (p + l > ref_len) fail(3, k);
                    else if (t == SP_DUP) {
                        b1 = l; b2 = l;
                        src_b1 = src_b2 = (int64_t(SEG_REF_FWD) << SEG_KIND_SHIFT) | p;
                    } else {
                        b1 = l;
                        src_b1 = (int64_t(SEG_REF_RC) << SEG_KIND_SHIFT) | ((p + l - 1) & SEG_OFF_MASK);
                    }
                } else if (t == SP_INS) {
                    b1 = payload_off[id + 1] - payload_off[id];
                    src_b1 = (int64_t(SEG_PAYLOAD) << SEG_KIND_SHIFT) | payload_off[id];
                } else fail(4, k);
            }
        }
        int64_t total;
        const int64_t excl = block_exclusive_scan_256(gap + b1 + b2, warp_sums, total) + carry_s;
        if (k < n) {
            const int64_t s = sb + 1 + 3 * k;
            seg_dst[s] = excl;            seg_src[s] = src_gap;
            seg_dst[s + 1] = excl + gap;  seg_src[s + 1] = src_b1;
            seg_dst[s + 2] = excl + gap + b1;  seg_src[s + 2] = src_b2;
        }
        __syncthreads();
        if (tid == 0) carry_s += total;
        __syncthreads();
    }
    if (tid == 0) {
        int64_t last_after = 0;
        if (n > 0) {
            const int32_t pid = ids[lo + n - 1];
            if (pid >= 0 && pid < n_variants)
                last_after = (type[pid] == SP_INS) ? (int64_t)pos[pid] : (int64_t)pos[pid] + len[pid];
        }
        int64_t c = carry_s;
        const int64_t s = sb + 1 + 3 * n;
        const int64_t tail = last_after < ref_len ? ref_len - last_after : 0;  // :953
        seg_dst[s] = c;  seg_src[s] = (int64_t(SEG_REF_FWD) << SEG_KIND_SHIFT) | (last_after & SEG_OFF_MASK);
        c += tail;
        seg_dst[s + 1] = c;  seg_src[s + 1] = (int64_t(SEG_AUX) << SEG_KIND_SHIFT) | 0;  // "\n" (bw.newLine, :956)
        c += hdr_len > 0 ? 1 : 0;
        seg_dst[s + 2] = c;  seg_src[s + 2] = int64_t(SEG_PAD) << SEG_KIND_SHIFT;
        seg_dst[s + 3] = INT64_MAX;  seg_src[s + 3] = int64_t(SEG_PAD) << SEG_KIND_SHIFT;
        hap_len[hb] = c;
    }
}

// One thread per 16 KiB output tile: the runs of the tile = the non-empty, non-PAD intersections of the slot's segments
// with the tile, in output order. Pass 1 counts them (and remembers the first segment), a scan assigns each tile its
// place in the run array, pass 2 writes the descriptor (first run) and the run array (the others).
__device__ __forceinline__ int32_t first_segment_of_tile(const int64_t* __restrict__ seg_dst, int64_t n, int64_t x) {
    int64_t lo = 0, hi = n;   // first local index with dst > x
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (seg_dst[mid] <= x) lo = mid + 1; else hi = mid;
    }
    return (int32_t)(lo - 1);
}

template <bool FILL>
__global__ void __launch_bounds__(256)
tile_runs_kernel(const int64_t* __restrict__ seg_dst, const int64_t* __restrict__ seg_src, const int64_t* __restrict__ seg_base,
                 int n_hap, int64_t tiles_per_slot, int32_t* __restrict__ s0_io, int32_t* __restrict__ more, const int32_t* __restrict__ first_more,
                 TileDesc* __restrict__ desc, RunItem* __restrict__ items) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= tiles_per_slot * n_hap) return;
    const int hb = (int)(idx / tiles_per_slot);
    const int64_t x = (idx % tiles_per_slot) * TILE_BYTES;
    const int64_t sb = seg_base[hb];
    const int64_t* gd = seg_dst + sb;
    const int64_t* gs = seg_src + sb;
    const int32_t s0 = FILL ? s0_io[idx] : first_segment_of_tile(gd, seg_base[hb + 1] - sb, x);  // the sentinel (INT64_MAX) is never <= x
    int32_t n_runs = 0;
    TileDesc d;
    d.srcw0 = int64_t(SEG_PAD) << SEG_KIND_SHIFT; d.len0_more = 0; d.item_first = 0;
    int32_t at = FILL ? first_more[idx] : 0;
    int64_t cur = gd[s0];
    for (int32_t s = s0; cur < x + TILE_BYTES; s++) {
        const int64_t nxt = gd[s + 1];
        const int64_t a = cur > x ? cur : x, b = nxt < x + TILE_BYTES ? nxt : x + TILE_BYTES;
        const int64_t sw = gs[s];
        const int kind = (int)(sw >> SEG_KIND_SHIFT);
        if (b > a && kind != SEG_PAD) {
            if (FILL) {
                const int64_t r = a - cur, o = sw & SEG_OFF_MASK;
                const int64_t srcw = (int64_t(kind) << SEG_KIND_SHIFT) | ((kind == SEG_REF_RC ? o - r : o + r) & SEG_OFF_MASK);
                if (n_runs == 0) {
                    if (a == x) { d.srcw0 = srcw; d.len0_more = (uint32_t)(b - a); }
                    else { RunItem it; it.srcw = srcw; it.off = (int32_t)(a - x); it.len = (int32_t)(b - a); items[at++] = it; }
                } else { RunItem it; it.srcw = srcw; it.off = (int32_t)(a - x); it.len = (int32_t)(b - a); items[at++] = it; }
            }
            // a first run that does not start at the tile's first byte goes to the run array too (the descriptor's run is empty)
            n_runs += (n_runs == 0 && a != x) ? 2 : 1;
        }
        cur = nxt;
    }
    const int32_t n_more = n_runs > 0 ? n_runs - 1 : 0;
    if (FILL) {
        d.len0_more |= (uint32_t)n_more << 15;
        d.item_first = first_more[idx];
        desc[idx] = d;
    } else { s0_io[idx] = s0; more[idx] = n_more; }
}

__device__ __forceinline__ uint4 shift_combine(uint4 A, uint4 B, unsigned m) {
    const unsigned s = (m & 3u) * 8u;
    uint32_t x0, x1, x2, x3, x4;
    switch (m >> 2) {
        case 0: x0 = A.x; x1 = A.y; x2 = A.z; x3 = A.w; x4 = B.x; break;
        case 1: x0 = A.y; x1 = A.z; x2 = A.w; x3 = B.x; x4 = B.y; break;
        case 2: x0 = A.z; x1 = A.w; x2 = B.x; x3 = B.y; x4 = B.z; break;
        default: x0 = A.w; x1 = B.x; x2 = B.y; x3 = B.z; x4 = B.w; break;
    }
    return make_uint4(__funnelshift_r(x0, x1, s), __funnelshift_r(x1, x2, s), __funnelshift_r(x2, x3, s),
                      __funnelshift_r(x3, x4, s));
}

// One run: output bytes [a, a+n) of the tile from one source. Every thread of the CTA takes part.
__device__ __noinline__ void copy_run(const uint8_t* __restrict__ ref, const uint8_t* __restrict__ payload, const uint8_t* __restrict__ aux,
                                         int64_t srcw, int32_t a, int32_t n, uint8_t* __restrict__ dst_tile) {
    const int kind = (int)(srcw >> SEG_KIND_SHIFT);
    if (kind == SEG_PAD || n <= 0) return;
    const int tid = threadIdx.x;
    uint8_t* d = dst_tile + a;
    if (kind == SEG_REF_RC) {   // an inversion (2 per thousand variants): byte by byte, walking the reference downwards
        const uint8_t* sp = ref + (srcw & SEG_OFF_MASK);
        for (int32_t i = tid; i < n; i += 256) d[i] = complement_base(__ldg(sp - i));
        return;
    }
    const uint8_t* p = (kind == SEG_REF_FWD ? ref : (kind == SEG_PAYLOAD ? payload : aux)) + (srcw & SEG_OFF_MASK);
    const int32_t head = min(n, (16 - (a & 15)) & 15);   // bytes up to the first 16-byte boundary of the output (tiles are 16 KiB aligned)
    const int32_t nbody = (n - head) >> 4, tail = (n - head) & 15;
    if (tid < head) d[tid] = __ldg(p + tid);
    if (tid >= 32 && tid - 32 < tail) d[head + nbody * 16 + (tid - 32)] = __ldg(p + head + nbody * 16 + (tid - 32));
    const uint8_t* pb = p + head;
    const unsigned m = (unsigned)((uintptr_t)pb & 15u);
    const uint4* q = (const uint4*)(pb - m);
    uint4* o = (uint4*)(d + head);
    for (int32_t c0 = 0; c0 < nbody; c0 += 1024) {   // all loads of a batch before its stores: 8 in flight per thread
        uint4 A[4], B[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int32_t c = c0 + u * 256 + tid;
            if (c < nbody) { A[u] = __ldg(q + c); if (m) B[u] = __ldg(q + c + 1); }   // buffers carry >= 32 bytes of slack past their end
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int32_t c = c0 + u * 256 + tid;
            if (c < nbody) __stcs(o + c, m ? shift_combine(A[u], B[u], m) : A[u]);
        }
    }
}

__global__ void __launch_bounds__(256, 5)
materialise_kernel(const uint8_t* __restrict__ ref, const uint8_t* __restrict__ payload, const uint8_t* __restrict__ aux,
                   const TileDesc* __restrict__ desc, const RunItem* __restrict__ items, uint8_t* __restrict__ out, int64_t stride,
                   int64_t tiles_per_slot, int n_hap) {
    // region-major order: consecutive CTAs work on the same tile index of different haplotypes
    const int64_t b = blockIdx.x;
    const int hb = (int)(b % n_hap);
    const int64_t t = b / n_hap;
    const int64_t tile = (int64_t)hb * tiles_per_slot + t;
    const int4 dw = __ldg((const int4*)(desc + tile));
    uint8_t* __restrict__ dst_tile = out + (int64_t)hb * stride + t * TILE_BYTES;
    const int tid = threadIdx.x;
    constexpr int CH = TILE_BYTES / 16 / 256;
    const int64_t srcw0 = ((int64_t)(uint32_t)dw.y << 32) | (uint32_t)dw.x;
    const uint32_t len0 = (uint32_t)dw.z & 0x7fffu, n_more = (uint32_t)dw.z >> 15;
    const int kind0 = (int)(srcw0 >> SEG_KIND_SHIFT);

    if (len0 == TILE_BYTES && kind0 != SEG_REF_RC) {
        // ---- the whole tile lies inside one forward segment (n_more is 0 then)
        const uint8_t* base = kind0 == SEG_REF_FWD ? ref : (kind0 == SEG_PAYLOAD ? payload : aux);
        const uint8_t* p = base + (srcw0 & SEG_OFF_MASK);
        const unsigned m = (unsigned)((uintptr_t)p & 15u);
        const uint4* q = (const uint4*)(p - m);
        uint4* o = (uint4*)dst_tile;
        if (m == 0) {
            uint4 v[CH];
#pragma unroll
            for (int u = 0; u < CH; u++) v[u] = __ldg(q + u * 256 + tid);
#pragma unroll
            for (int u = 0; u < CH; u++) __stcs(o + u * 256 + tid, v[u]);
        } else {
            uint4 A[CH], B[CH];
#pragma unroll
            for (int u = 0; u < CH; u++) { A[u] = __ldg(q + u * 256 + tid); B[u] = __ldg(q + u * 256 + tid + 1); }
#pragma unroll
            for (int u = 0; u < CH; u++) __stcs(o + u * 256 + tid, shift_combine(A[u], B[u], m));
        }
        return;
    }
    // ---- several runs (or a short / reverse-complemented one): the run array is read while the first run is copied
    const int4* it = (const int4*)(items + dw.w);
    int4 nxt = make_int4(0, 0, 0, 0);
    if (n_more) nxt = __ldg(it);
    copy_run(ref, payload, aux, srcw0, 0, (int32_t)len0, dst_tile);
#pragma unroll 1
    for (uint32_t j = 0; j < n_more; j++) {
        const int4 cur = nxt;
        if (j + 1 < n_more) nxt = __ldg(it + j + 1);
        copy_run(ref, payload, aux, ((int64_t)(uint32_t)cur.y << 32) | (uint32_t)cur.x, cur.z, cur.w, dst_tile);
    }
}

// Buffers of build_run_tables, so that a caller that times its launches can have the allocations behind it first.
static void reserve_run_tables(sp_ctx* c, int64_t ntiles, int64_t n_segments, DevBuf<TileDesc>& desc, DevBuf<RunItem>& items) {
    desc.reserve((size_t)ntiles);
    items.reserve((size_t)(n_segments + ntiles + 64));   // every run but the first of a tile: fewer than segments + tiles
    c->d_tile_more.reserve((size_t)ntiles + 1); c->d_tile_s0.reserve((size_t)ntiles); c->d_tile_scan.reserve((size_t)ntiles + 1);
    size_t tmp = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp, c->d_tile_more.p, c->d_tile_scan.p, (int)ntiles, c->stream);
    c->d_tile_tmp.reserve(tmp + 256);
}
void reserve_tiled_copy(sp_ctx* c, int64_t stride, int n_slots, int64_t n_segments) {
    reserve_run_tables(c, stride / TILE_BYTES * n_slots, n_segments, c->d_copy_desc, c->d_copy_items);
}

// Run tables for n_hap slots of tiles_per_slot tiles (count, scan, fill) on stream st.
static void build_run_tables(sp_ctx* c, cudaStream_t st, const int64_t* seg_dst, const int64_t* seg_src, const int64_t* seg_base, int n_hap,
                             int64_t tiles_per_slot, int64_t n_segments, DevBuf<TileDesc>& desc, DevBuf<RunItem>& items) {
    const int64_t ntiles = tiles_per_slot * n_hap;
    reserve_run_tables(c, ntiles, n_segments, desc, items);
    size_t tmp = 0;
    const unsigned grid = (unsigned)((ntiles + 255) / 256);
    tile_runs_kernel<false><<<grid, 256, 0, st>>>(seg_dst, seg_src, seg_base, n_hap, tiles_per_slot, c->d_tile_s0.p, c->d_tile_more.p, nullptr, nullptr, nullptr);
    tmp = c->d_tile_tmp.cap;
    SP_CUDA(cub::DeviceScan::ExclusiveSum(c->d_tile_tmp.p, tmp, c->d_tile_more.p, c->d_tile_scan.p, (int)ntiles, st));
    tile_runs_kernel<true><<<grid, 256, 0, st>>>(seg_dst, seg_src, seg_base, n_hap, tiles_per_slot, c->d_tile_s0.p, nullptr, c->d_tile_scan.p, desc.p, items.p);
    c->launches += 3;
    SP_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------- host
static int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

static void launch_materialise(cudaStream_t st, int64_t ntiles, const uint8_t* ref, const uint8_t* payload, const uint8_t* aux,
                               const TileDesc* desc, const RunItem* items, uint8_t* out, int64_t stride, int64_t tiles_per_slot, int n_slots) {
    materialise_kernel<<<(unsigned)ntiles, 256, 0, st>>>(ref, payload, aux, desc, items, out, stride, tiles_per_slot, n_slots);
}

void plan_haplotypes(sp_ctx* c, int32_t first, int32_t last, int32_t framing, int64_t* out_lengths, int64_t* slot_stride) {
    if (c->ref_len <= 0) throw Error(SP_ERR_ARG, "no reference loaded");
    if (c->v_n <= 0 || !c->v_off_dev) throw Error(SP_ERR_ARG, "no haplotype lists loaded");
    if (first < c->v_h0 || last >= c->v_h0 + c->v_n || first > last)
        throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: haplotype range [" + std::to_string(first) + ".." + std::to_string(last) + "]");
    const int nh = last - first + 1;
    c->materialised = false;
    c->plan_last = -1; c->plan_first = 0;   // no valid plan while this one is being built
    // aux blob: "\n" at offset 0, then the FASTA headers ">chr1_<i>\n" (src/SimulateHaplotypes.java:955)
    std::vector<uint8_t> aux(1, (uint8_t)'\n');
    std::vector<int64_t> aux_off(nh + 1);
    c->h_hdr_len.assign(nh, 0);
    for (int i = 0; i < nh; i++) {
        aux_off[i] = (int64_t)aux.size();
        if (framing) {
            std::string hdr = ">chr1_" + std::to_string(first + i) + "\n";
            aux.insert(aux.end(), hdr.begin(), hdr.end());
            c->h_hdr_len[i] = (int32_t)hdr.size();
        }
    }
    aux_off[nh] = (int64_t)aux.size();
    std::vector<int64_t> seg_base(nh + 1);
    const int64_t* lo_host = c->v_off_host + (first - c->v_h0);
    const int64_t l0 = lo_host[0];
    for (int i = 0; i <= nh; i++) seg_base[i] = 3 * (lo_host[i] - l0) + 5 * (int64_t)i;
    const int64_t nseg = seg_base[nh];
    // The plan goes into the set the copy kernel in flight is not reading, on the planning stream when there is one: the
    // round trip below then waits for the plan kernel alone, never for the copy kernel of the batch before.
    cudaStream_t st = c->pop_stream ? c->pop_stream : c->stream;
    const int ps = c->plan_cur ^ 1;
    sp_ctx::PlanSet& B = c->plan_set[ps];
    if (!B.ev_used) SP_CUDA(cudaEventCreateWithFlags(&B.ev_used, cudaEventDisableTiming));
    else if (st != c->stream) SP_CUDA(cudaStreamWaitEvent(st, B.ev_used, 0));   // the copy kernel that last read this set
    if (!c->ev_planned) SP_CUDA(cudaEventCreateWithFlags(&c->ev_planned, cudaEventDisableTiming));
    B.aux.reserve(aux.size());
    B.aux_off.reserve(nh + 1);
    B.seg_base.reserve(nh + 1);
    B.seg_dst.reserve(nseg);
    B.seg_src.reserve(nseg);
    c->d_hap_len.reserve(nh);
    c->d_err.reserve(4);
    SP_CUDA(cudaMemcpyAsync(B.aux.p, aux.data(), aux.size(), cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemcpyAsync(B.aux_off.p, aux_off.data(), (nh + 1) * 8, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemcpyAsync(B.seg_base.p, seg_base.data(), (nh + 1) * 8, cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaMemsetAsync(c->d_err.p, 0, 16, st));
    plan_kernel<<<nh, 256, 0, st>>>(c->d_type.p, c->d_pos.p, c->d_len.p, c->d_payload_off.p, c->n_variants,
                                    c->v_off_dev + (first - c->v_h0), c->v_ids_dev, c->ref_len, first, B.aux_off.p,
                                    B.seg_base.p, B.seg_dst.p, B.seg_src.p, c->d_hap_len.p, c->d_err.p);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    int64_t* hp = c->pinned((size_t)nh * 8 + 16);
    int32_t* err = (int32_t*)(hp + nh);
    SP_CUDA(cudaMemcpyAsync(hp, c->d_hap_len.p, (size_t)nh * 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaMemcpyAsync(err, c->d_err.p, 16, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    if (err[0] != 0) {
        static const char* what[] = {"", "ArrayIndexOutOfBoundsException: variant id", "IndexOutOfBoundsException: start > end",
                                     "IndexOutOfBoundsException: end > reference length", "invalid svType"};
        throw Error(err[0] == 4 ? SP_ERR_ARG : SP_ERR_JAVA_EXCEPTION,
                    std::string(what[err[0]]) + " (haplotype " + std::to_string(err[1]) + ", list index " + std::to_string(err[2]) + ")");
    }
    c->h_hap_len.assign(hp, hp + nh);
    int64_t mx = 0;
    for (int i = 0; i < nh; i++) mx = std::max(mx, c->h_hap_len[i]);
    c->slot_stride = round_up(std::max<int64_t>(mx, 1), TILE_BYTES);
    c->tiles_per_slot = c->slot_stride / TILE_BYTES;
    const int64_t ntiles = c->tiles_per_slot * nh;
    if (ntiles >= (int64_t)INT32_MAX) throw Error(SP_ERR_ARG, "batch too large: reduce the haplotype range");
    build_run_tables(c, st, B.seg_dst.p, B.seg_src.p, B.seg_base.p, nh, c->tiles_per_slot, nseg, B.tile_desc, B.items);
    SP_CUDA(cudaEventRecord(c->ev_planned, st));
    c->d_out.reserve((size_t)c->slot_stride * nh, 0.01);
    c->plan_cur = ps;
    c->plan_first = first; c->plan_last = last; c->plan_framing = framing;
    if (out_lengths) for (int i = 0; i < nh; i++) out_lengths[i] = c->h_hap_len[i];
    if (slot_stride) *slot_stride = c->slot_stride;
}

void materialise(sp_ctx* c) {
    if (c->plan_last < c->plan_first) throw Error(SP_ERR_ARG, "sp_plan_haplotypes has not been called");
    const int nh = c->plan_last - c->plan_first + 1;
    const int64_t ntiles = c->tiles_per_slot * nh;
    sp_ctx::PlanSet& B = c->plan_set[c->plan_cur];
    SP_CUDA(cudaStreamWaitEvent(c->stream, c->ev_planned, 0));   // no-op when the plan ran on this stream
    launch_materialise(c->stream, ntiles, c->d_ref.p, c->d_payload.p, B.aux.p, B.tile_desc.p, B.items.p, c->d_out.p, c->slot_stride,
                       c->tiles_per_slot, nh);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaEventRecord(B.ev_used, c->stream));
    c->materialised = true;
}

// The tiled copy on caller-built segment tables (rows D, W, F1, F3 reuse it: records -> file bytes).
// n_slots slots of `stride` bytes; slot s owns segments [seg_base[s], seg_base[s+1]) laid out as plan_kernel does:
// ascending seg_dst, then a PAD segment at the slot's total length and a sentinel with dst = INT64_MAX; n_segments = their
// total number.
void launch_tiled_copy(sp_ctx* c, const uint8_t* src, const uint8_t* payload, const uint8_t* aux, const int64_t* seg_dst,
                       const int64_t* seg_src, const int64_t* seg_base, int64_t n_segments, uint8_t* out, int64_t stride, int n_slots) {
    const int64_t tiles_per_slot = stride / TILE_BYTES;
    const int64_t ntiles = tiles_per_slot * n_slots;
    if (ntiles <= 0) return;
    if (ntiles >= (int64_t)INT32_MAX) throw Error(SP_ERR_ARG, "tiled copy: too many tiles");
    build_run_tables(c, c->stream, seg_dst, seg_src, seg_base, n_slots, tiles_per_slot, n_segments, c->d_copy_desc, c->d_copy_items);
    launch_materialise(c->stream, ntiles, src, payload, aux, c->d_copy_desc.p, c->d_copy_items.p, out, stride, tiles_per_slot, n_slots);
    c->launches++;
    SP_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------- no-allocation variant
// sp_materialise_haplotype_static: one haplotype through the same plan_kernel / tile_runs_kernel / materialise_kernel, with
// every plan table (segment offsets, per-tile counts, tile descriptors, run array, the scan's temporary storage) in ONE
// file-scope device array instead of the context's cudaMalloc'ed buffers. For callers that may not allocate (CUDA graph
// capture, serving loops). Every GPU that loads the library holds the array until exit; every call shares it (calls that
// overlap on the GPU race) and it keeps the last call's contents, so the words that must start at zero are cleared per call.
constexpr int32_t STATIC_MAX_EDITS = SP_STATIC_MAX_EDITS;
constexpr int64_t STATIC_MAX_TILES = SP_STATIC_MAX_OUTPUT_BYTES / TILE_BYTES;
constexpr int64_t STATIC_NSEG = 3 * (int64_t)STATIC_MAX_EDITS + 5;
constexpr size_t static_up(size_t x) { return (x + 255) / 256 * 256; }
// header (256 B): int64 list_off[2] @0, seg_base[2] @16, aux_off[2] @32, hap_len @48; int32 err[4] @64; aux bytes @128 ('\n' + slack)
constexpr size_t ST_HDR = 0;
constexpr size_t ST_SEG_DST = 256;
constexpr size_t ST_SEG_SRC = ST_SEG_DST + static_up(STATIC_NSEG * 8);
constexpr size_t ST_TILE_S0 = ST_SEG_SRC + static_up(STATIC_NSEG * 8);
constexpr size_t ST_TILE_MORE = ST_TILE_S0 + static_up(STATIC_MAX_TILES * 4);
constexpr size_t ST_TILE_SCAN = ST_TILE_MORE + static_up((STATIC_MAX_TILES + 1) * 4);
constexpr size_t ST_DESC = ST_TILE_SCAN + static_up((STATIC_MAX_TILES + 1) * 4);
constexpr size_t ST_ITEMS = ST_DESC + static_up(STATIC_MAX_TILES * sizeof(TileDesc));
constexpr size_t ST_CUB = ST_ITEMS + static_up((STATIC_NSEG + STATIC_MAX_TILES + 64) * sizeof(RunItem));
constexpr size_t ST_CUB_BYTES = 64 << 10;
constexpr size_t ST_BYTES = ST_CUB + ST_CUB_BYTES;   // about 3.1 MB
static_assert(ST_BYTES <= (4u << 20), "the static scratch is meant to stay modest");

static __device__ __align__(256) uint8_t g_static_scratch[ST_BYTES];

// In host code `g_static_scratch` is a stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static uint8_t* static_scratch_address(int device) {
    static uint8_t* addr[64] = {};
    static bool known[64] = {};
    if (device < 0 || device >= 64) throw Error(SP_ERR_ARG, "static scratch: device index out of range");
    if (!known[device]) {
        void* p = nullptr;
        SP_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) throw Error(SP_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        addr[device] = (uint8_t*)p;
        known[device] = true;
    }
    return addr[device];
}

__global__ void static_plan_init_kernel(uint8_t* __restrict__ s, int32_t n_ids) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int64_t* h = (int64_t*)s;
    h[0] = 0; h[1] = n_ids;                       // list_off of the one haplotype
    h[2] = 0; h[3] = 3 * (int64_t)n_ids + 5;      // seg_base
    h[4] = 1; h[5] = 1;                           // aux_off: no FASTA header (the '\n' of bw.newLine sits at aux[0])
    h[6] = 0;                                     // hap_len
    int32_t* e = (int32_t*)(s + 64);
    e[0] = e[1] = e[2] = e[3] = 0;                // the plan's error words must start at zero
    s[128] = (uint8_t)'\n';
}

void materialise_haplotype_static(sp_ctx* c, const int32_t* device_ids, int32_t n_ids, void* device_out, int64_t capacity, int64_t* length) {
    if (c->ref_len <= 0 || !c->d_ref.p) throw Error(SP_ERR_ARG, "no reference loaded");
    if (n_ids < 0 || (n_ids > 0 && !device_ids)) throw Error(SP_ERR_ARG, "bad edit list");
    if (n_ids > 0 && c->n_variants <= 0) throw Error(SP_ERR_ARG, "no variant table loaded");
    if (n_ids > STATIC_MAX_EDITS)
        throw Error(SP_ERR_ARG, "sp_materialise_haplotype_static: " + std::to_string(n_ids) + " edits exceed the static scratch (SP_STATIC_MAX_EDITS = " +
                                    std::to_string(STATIC_MAX_EDITS) + "); use sp_plan_haplotypes + sp_materialise");
    if (!device_out || capacity <= 0 || ((uintptr_t)device_out & 15u)) throw Error(SP_ERR_ARG, "the output must be a 16-byte aligned device buffer");
    uint8_t* s = static_scratch_address(c->device);
    cudaStream_t st = c->stream;
    int64_t* hdr = (int64_t*)(s + ST_HDR);
    int64_t* seg_dst = (int64_t*)(s + ST_SEG_DST);
    int64_t* seg_src = (int64_t*)(s + ST_SEG_SRC);
    static_plan_init_kernel<<<1, 32, 0, st>>>(s, n_ids);
    plan_kernel<<<1, 256, 0, st>>>(c->d_type.p, c->d_pos.p, c->d_len.p, c->d_payload_off.p, c->n_variants, hdr + 0, device_ids, c->ref_len, 0,
                                   hdr + 4, hdr + 2, seg_dst, seg_src, hdr + 6, (int32_t*)(s + 64));
    c->launches += 2;
    SP_CUDA(cudaGetLastError());
    int64_t h_len = 0;
    int32_t h_err[4] = {0, 0, 0, 0};
    SP_CUDA(cudaMemcpyAsync(&h_len, hdr + 6, 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaMemcpyAsync(h_err, s + 64, 16, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    if (h_err[0] != 0) {
        static const char* what[] = {"", "ArrayIndexOutOfBoundsException: variant id", "IndexOutOfBoundsException: start > end",
                                     "IndexOutOfBoundsException: end > reference length", "invalid svType"};
        throw Error(h_err[0] == 4 ? SP_ERR_ARG : SP_ERR_JAVA_EXCEPTION, std::string(what[h_err[0]]) + " (list index " + std::to_string(h_err[2]) + ")");
    }
    if (length) *length = h_len;
    if (h_len > (int64_t)SP_STATIC_MAX_OUTPUT_BYTES)
        throw Error(SP_ERR_ARG, "sp_materialise_haplotype_static: a haplotype of " + std::to_string(h_len) + " bytes exceeds the static scratch (SP_STATIC_MAX_OUTPUT_BYTES = " +
                                    std::to_string((int64_t)SP_STATIC_MAX_OUTPUT_BYTES) + ")");
    if (h_len > capacity) throw Error(SP_ERR_ARG, "device buffer too small: the haplotype has " + std::to_string(h_len) + " bytes");
    const int64_t tiles = round_up(std::max<int64_t>(h_len, 1), TILE_BYTES) / TILE_BYTES;
    int32_t* s0 = (int32_t*)(s + ST_TILE_S0);
    int32_t* more = (int32_t*)(s + ST_TILE_MORE);
    int32_t* scan = (int32_t*)(s + ST_TILE_SCAN);
    TileDesc* desc = (TileDesc*)(s + ST_DESC);
    RunItem* items = (RunItem*)(s + ST_ITEMS);
    size_t tmp = 0;
    SP_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, more, scan, (int)tiles, st));   // size query: allocates nothing
    if (tmp > ST_CUB_BYTES) throw Error(SP_ERR, "static scratch: the scan needs " + std::to_string(tmp) + " bytes of temporary storage");
    const unsigned grid = (unsigned)((tiles + 255) / 256);
    tile_runs_kernel<false><<<grid, 256, 0, st>>>(seg_dst, seg_src, hdr + 2, 1, tiles, s0, more, nullptr, nullptr, nullptr);
    tmp = ST_CUB_BYTES;
    SP_CUDA(cub::DeviceScan::ExclusiveSum(s + ST_CUB, tmp, more, scan, (int)tiles, st));
    tile_runs_kernel<true><<<grid, 256, 0, st>>>(seg_dst, seg_src, hdr + 2, 1, tiles, s0, nullptr, scan, desc, items);
    launch_materialise(st, tiles, c->d_ref.p, c->d_payload.p, s + 128, desc, items, (uint8_t*)device_out, tiles * TILE_BYTES, tiles, 1);
    c->launches += 4;
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaStreamSynchronize(st));   // the output is complete and the scratch is free again when the call returns
}

}  // namespace sp

extern "C" int sp_materialise_haplotype_static(sp_ctx* ctx, const int32_t* device_ids, int32_t n_ids, void* device_out, int64_t capacity,
                                               int64_t* length) {
    SP_API_BEGIN
    sp::materialise_haplotype_static(ctx, device_ids, n_ids, device_out, capacity, length);
    SP_API_END
}
