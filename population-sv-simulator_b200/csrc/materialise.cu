// Row M of SURVEY.md §8a — haplotype materialisation (SimulateHaplotypes.buildHaplotypes,
// reference src/SimulateHaplotypes.java:917-960), re-designed for B200:
//
//   plan_kernel ......... one CTA per haplotype; every edit of the sorted edit list becomes three segment
//                         slots (reference gap, body, second DUP copy) and a block-wide exclusive scan over the
//                         per-edit output deltas yields each segment's destination offset (the "segmented scan").
//   tile_table_kernel ... one thread per 16 KiB output tile: index of the first segment that intersects it.
//   materialise_kernel .. one CTA per output tile, tiles ordered region-major across the haplotypes of the
//                         batch so that the reference window they share stays in the 126 MB L2; 16-byte
//                         aligned streaming stores; the arbitrary byte shift between source and destination
//                         is resolved in registers (two aligned 16-byte loads + funnel shifts).
//
// HBM-bound: 1 B read + 1 B written per output base (2 B/bp algorithmic).
#include "common.cuh"

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

// One thread per 16 KiB output tile: which segments intersect it, and — for the common tile that lies inside one
// forward segment — where its source bytes start, so that the copy kernel needs ONE 16-byte descriptor load before its
// data loads instead of a chain through tile_first -> seg_dst/seg_src.
__global__ void __launch_bounds__(256)
tile_table_kernel(const int64_t* __restrict__ seg_dst, const int64_t* __restrict__ seg_src, const int64_t* __restrict__ seg_base,
                  int n_hap, int64_t tiles_per_slot, TileDesc* __restrict__ desc, int inline_entries) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= tiles_per_slot * n_hap) return;
    const int hb = (int)(idx / tiles_per_slot);
    const int64_t x = (idx % tiles_per_slot) * TILE_BYTES;
    const int64_t sb = seg_base[hb];
    const int64_t n = seg_base[hb + 1] - sb;  // includes the PAD segment and the INT64_MAX sentinel
    int64_t lo = 0, hi = n;                   // first local index with dst > x
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (seg_dst[sb + mid] <= x) lo = mid + 1; else hi = mid;
    }
    const int64_t s0 = lo - 1;
    int64_t lo2 = lo, hi2 = n;                // first local index with dst >= x + TILE_BYTES (the sentinel at the latest)
    while (lo2 < hi2) {
        const int64_t mid = (lo2 + hi2) >> 1;
        if (seg_dst[sb + mid] < x + TILE_BYTES) lo2 = mid + 1; else hi2 = mid;
    }
    const int64_t s1 = lo2 - 1;
    const int64_t src = seg_src[sb + s0];
    const int kind = (int)(src >> SEG_KIND_SHIFT);
    TileDesc d;
    d.s0 = (int32_t)s0;
    if (s1 == s0 && kind != SEG_REF_RC) {
        d.nseg = 0;  // uniform tile
        d.srcw = (int64_t(kind) << SEG_KIND_SHIFT) | (((src & SEG_OFF_MASK) + (x - seg_dst[sb + s0])) & SEG_OFF_MASK);
    } else {
        d.nseg = (int32_t)(s1 - s0 + 1);
        d.srcw = 0;
        if (inline_entries && d.nseg <= INLINE_SEGS) {
            auto clamp = [](int64_t r) { return (int32_t)(r < -(1 << 30) ? -(1 << 30) : (r > (1 << 30) ? (1 << 30) : r)); };
            SegEntry* e = (SegEntry*)(desc + tiles_per_slot * n_hap) + idx * INLINE_SEGS;
            for (int i = 0; i < d.nseg; i++) {
                const int64_t r = seg_dst[sb + s0 + i] - x;
                const int64_t sw = seg_src[sb + s0 + i];
                const int k = (int)(sw >> SEG_KIND_SHIFT);
                SegEntry v;
                v.rel = clamp(r); v.kind = k;
                v.adj = k == SEG_REF_RC ? (sw & SEG_OFF_MASK) + r : (sw & SEG_OFF_MASK) - r;
                e[i] = v;
            }
            d.srcw = TILE_INLINE_FLAG | (int64_t)(uint32_t)clamp(seg_dst[sb + s0 + d.nseg] - x);
        }
    }
    desc[idx] = d;
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

constexpr int MAX_TILE_SEGS = 96;
constexpr int COUNT_SEARCH_SEGS = 12;  // up to here the segment of a chunk is found by counting (independent loads)

// Tile with more than MAX_TILE_SEGS segments (thousands of tiny records): rare, kept simple. One 16-byte chunk at a
// time, segment found by binary search in global memory, bytes gathered one by one.
__device__ __noinline__ void copy_tile_many_segments(const uint8_t* __restrict__ ref, const uint8_t* __restrict__ payload,
                                                     const uint8_t* __restrict__ aux, const int64_t* __restrict__ gd,
                                                     const int64_t* __restrict__ gs, int nseg, int64_t x0,
                                                     uint8_t* __restrict__ dst_tile) {
    for (int c = threadIdx.x; c < TILE_BYTES / 16; c += 256) {
        const int64_t x = x0 + (int64_t)c * 16;
        int i = 0, hi = nseg;  // last i in [0,nseg) with gd[i] <= x
        while (hi - i > 1) {
            const int mid = (i + hi) >> 1;
            if (gd[mid] <= x) i = mid; else hi = mid;
        }
        int64_t dcur = gd[i], dnext = gd[i + 1], src = gs[i];
        uint32_t w[4] = {0, 0, 0, 0};
        bool any = false;
#pragma unroll 1
        for (int j = 0; j < 16; j++) {
            const int64_t xb = x + j;
            while (xb >= dnext) { i++; dcur = dnext; dnext = gd[i + 1]; src = gs[i]; }
            const int kind = (int)(src >> SEG_KIND_SHIFT);
            const int64_t o = src & SEG_OFF_MASK, r = xb - dcur;
            uint8_t byte = 0;
            if (kind == SEG_REF_FWD) byte = ref[o + r];
            else if (kind == SEG_REF_RC) byte = complement_base(ref[o - r]);
            else if (kind == SEG_PAYLOAD) byte = payload[o + r];
            else if (kind == SEG_AUX) byte = aux[o + r];
            if (kind != SEG_PAD) any = true;
            w[j >> 2] |= (uint32_t)byte << ((j & 3) * 8);
        }
        if (any) __stcs((uint4*)(dst_tile + (int64_t)c * 16), make_uint4(w[0], w[1], w[2], w[3]));
    }
}

__global__ void __launch_bounds__(256, 5)
materialise_kernel(const uint8_t* __restrict__ ref, const uint8_t* __restrict__ payload,
                   const uint8_t* __restrict__ aux, const int64_t* __restrict__ seg_dst,
                   const int64_t* __restrict__ seg_src, const int64_t* __restrict__ seg_base,
                   const TileDesc* __restrict__ desc, uint8_t* __restrict__ out, int64_t stride,
                   int64_t tiles_per_slot, int n_hap, int inline_entries) {
    // region-major order: consecutive CTAs work on the same tile index of different haplotypes
    const int64_t b = blockIdx.x;
    const int hb = (int)(b % n_hap);
    const int64_t t = b / n_hap;
    const int64_t x0 = t * TILE_BYTES;
    const int64_t tile = (int64_t)hb * tiles_per_slot + t;
    const int4 dw = __ldg((const int4*)(desc + tile));
    // record-gather launches: nearly every tile is crossed by record boundaries, so the tile's inline segment entries are
    // requested together with the descriptor (one memory round trip before the data loads instead of two)
    int4 ev = make_int4(0, 0, 0, 0);
    if (inline_entries && threadIdx.x < INLINE_SEGS)
        ev = __ldg((const int4*)((const SegEntry*)(desc + tiles_per_slot * n_hap) + tile * INLINE_SEGS + threadIdx.x));
    uint8_t* __restrict__ dst_tile = out + (int64_t)hb * stride + x0;
    const int tid = threadIdx.x;
    constexpr int CH = TILE_BYTES / 16 / 256;
    const int nseg = dw.w;

    if (nseg == 0) {
        // ---- uniform path: the whole tile lies inside one forward segment
        const int64_t srcw = ((int64_t)(uint32_t)dw.y << 32) | (uint32_t)dw.x;
        const int kind = (int)(srcw >> SEG_KIND_SHIFT);
        if (kind == SEG_PAD) return;
        const uint8_t* base = kind == SEG_REF_FWD ? ref : (kind == SEG_PAYLOAD ? payload : aux);
        const uint8_t* p = base + (srcw & SEG_OFF_MASK);
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

    // ---- general path: several segments (or a reverse-complemented one) intersect the tile
    const int64_t sb = seg_base[hb] + dw.z;
    if (nseg > MAX_TILE_SEGS) {
        copy_tile_many_segments(ref, payload, aux, seg_dst + sb, seg_src + sb, nseg, x0, dst_tile);
        return;
    }
    // Segment table of the tile in shared memory, relative to the tile: rel[i] = first output byte of segment i minus x0
    // (clamped to 32 bits), adj[i] = source offset that corresponds to x0 (forward: src + (x - x0), reverse
    // complement: src - (x - x0)), kind[i].
    __shared__ int32_t s_rel[MAX_TILE_SEGS + 1];
    __shared__ int64_t s_adj[MAX_TILE_SEGS];
    __shared__ int32_t s_kind[MAX_TILE_SEGS];
    const int64_t srcw = ((int64_t)(uint32_t)dw.y << 32) | (uint32_t)dw.x;
    if (inline_entries && (srcw & TILE_INLINE_FLAG)) {
        if (tid < nseg) { s_rel[tid] = ev.x; s_kind[tid] = ev.y; s_adj[tid] = ((int64_t)(uint32_t)ev.w << 32) | (uint32_t)ev.z; }
        if (tid == 0) s_rel[nseg] = (int32_t)(uint32_t)dw.x;
    } else
    for (int i = tid; i <= nseg; i += 256) {
        const int64_t r = seg_dst[sb + i] - x0;
        s_rel[i] = (int32_t)(r < -(1 << 30) ? -(1 << 30) : (r > (1 << 30) ? (1 << 30) : r));
        if (i < nseg) {
            const int64_t src = seg_src[sb + i];
            const int kind = (int)(src >> SEG_KIND_SHIFT);
            s_kind[i] = kind;
            s_adj[i] = kind == SEG_REF_RC ? (src & SEG_OFF_MASK) + r : (src & SEG_OFF_MASK) - r;
        }
    }
    __syncthreads();
    auto find_segment = [&](int xr) -> int {  // last i in [0,nseg) with rel[i] <= xr (rel[0] <= 0 always)
        int lo = 0;
        if (nseg <= COUNT_SEARCH_SEGS) {
            for (int i = 1; i < nseg; i++) lo += s_rel[i] <= xr ? 1 : 0;
        } else {
            int hi = nseg;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (s_rel[mid] <= xr) lo = mid; else hi = mid;
            }
        }
        return lo;
    };
    // Phase 1: every thread resolves its four 16-byte chunks and issues the loads of all the chunks that lie inside one
    // forward segment (8 aligned 16-byte loads in flight per thread, as in the uniform path). Phase 2: realign + store.
    // Phase 3: the few chunks that straddle a segment boundary or belong to a reverse-complemented segment: sixteen
    // independent byte loads each.
    uint4 A[CH], B[CH];
    unsigned mis[CH];
    unsigned vec_mask = 0, slow_mask = 0;
#pragma unroll
    for (int u = 0; u < CH; u++) {
        const int xr = (tid + u * 256) * 16;
        const int lo = find_segment(xr);
        const int kind = s_kind[lo];
        mis[u] = 0;
        if (xr + 16 > s_rel[lo + 1] || kind == SEG_REF_RC) slow_mask |= 1u << u;
        else if (kind != SEG_PAD) {
            const uint8_t* base = kind == SEG_REF_FWD ? ref : (kind == SEG_PAYLOAD ? payload : aux);
            const uint8_t* p = base + (s_adj[lo] + xr);
            const unsigned m = (unsigned)((uintptr_t)p & 15u);
            const uint4* q = (const uint4*)(p - m);
            A[u] = __ldg(q);
            B[u] = __ldg(q + 1);   // buffers carry >= 32 bytes of slack past their end
            mis[u] = m; vec_mask |= 1u << u;
        }
    }
#pragma unroll
    for (int u = 0; u < CH; u++)
        if ((vec_mask >> u) & 1u) __stcs((uint4*)(dst_tile + (tid + u * 256) * 16), shift_combine(A[u], B[u], mis[u]));
#pragma unroll 1
    for (int u = 0; u < CH; u++) {
        if (!((slow_mask >> u) & 1u)) continue;
        const int xr = (tid + u * 256) * 16;
        int i = find_segment(xr);
        int rnext = s_rel[i + 1], kind = s_kind[i];
        int64_t adj = s_adj[i];
        uint8_t by[16];
        unsigned rc_mask = 0, any_mask = 0;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const int xb = xr + j;
            while (xb >= rnext) {  // zero-length segments are skipped here as well
                i++;
                rnext = s_rel[i + 1]; kind = s_kind[i]; adj = s_adj[i];
            }
            const uint8_t* p = kind == SEG_REF_RC ? ref + (adj - xb)
                                                  : (kind == SEG_REF_FWD ? ref : (kind == SEG_PAYLOAD ? payload : aux)) + (adj + xb);
            by[j] = 0;
            if (kind != SEG_PAD) { by[j] = __ldg(p); any_mask |= 1u << j; }
            if (kind == SEG_REF_RC) rc_mask |= 1u << j;
        }
        if (!any_mask) continue;
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const uint8_t byte = (rc_mask >> j) & 1u ? complement_base(by[j]) : by[j];
            w[j >> 2] |= (uint32_t)byte << ((j & 3) * 8);
        }
        __stcs((uint4*)(dst_tile + (tid + u * 256) * 16), make_uint4(w[0], w[1], w[2], w[3]));
    }
}

// ------------------------------------------------------------------------------------------------- host
static int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

static void launch_materialise(cudaStream_t st, int64_t ntiles, const uint8_t* ref, const uint8_t* payload, const uint8_t* aux,
                               const int64_t* seg_dst, const int64_t* seg_src, const int64_t* seg_base, const TileDesc* desc,
                               uint8_t* out, int64_t stride, int64_t tiles_per_slot, int n_slots, int inline_entries) {
    materialise_kernel<<<(unsigned)ntiles, 256, 0, st>>>(ref, payload, aux, seg_dst, seg_src, seg_base, desc, out, stride, tiles_per_slot, n_slots, inline_entries);
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
    B.tile_desc.reserve(ntiles);
    tile_table_kernel<<<(unsigned)((ntiles + 255) / 256), 256, 0, st>>>(B.seg_dst.p, B.seg_src.p, B.seg_base.p, nh,
                                                                        c->tiles_per_slot, B.tile_desc.p, 0);
    c->launches++;
    SP_CUDA(cudaGetLastError());
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
    launch_materialise(c->stream, ntiles, c->d_ref.p, c->d_payload.p, B.aux.p, B.seg_dst.p, B.seg_src.p, B.seg_base.p,
                       B.tile_desc.p, c->d_out.p, c->slot_stride, c->tiles_per_slot, nh, 0);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaEventRecord(B.ev_used, c->stream));
    c->materialised = true;
}

// The tiled copy on caller-built segment tables (row D reuses it: records of the length-bin blob -> FASTQ bytes).
// n_slots slots of `stride` bytes; slot s owns segments [seg_base[s], seg_base[s+1]) laid out as plan_kernel does:
// ascending seg_dst, then a PAD segment at the slot's total length and a sentinel with dst = INT64_MAX.
void launch_tiled_copy(sp_ctx* c, const uint8_t* src, const uint8_t* payload, const uint8_t* aux, const int64_t* seg_dst,
                       const int64_t* seg_src, const int64_t* seg_base, TileDesc* tile_desc, uint8_t* out, int64_t stride,
                       int n_slots) {
    const int64_t tiles_per_slot = stride / TILE_BYTES;
    const int64_t ntiles = tiles_per_slot * n_slots;
    if (ntiles <= 0) return;
    if (ntiles >= (int64_t)INT32_MAX) throw Error(SP_ERR_ARG, "tiled copy: too many tiles");
    // tile_desc holds ntiles descriptors followed by ntiles * INLINE_SEGS inline segment entries (both 16 bytes each)
    tile_table_kernel<<<(unsigned)((ntiles + 255) / 256), 256, 0, c->stream>>>(seg_dst, seg_src, seg_base, n_slots, tiles_per_slot, tile_desc, 1);
    launch_materialise(c->stream, ntiles, src, payload, aux, seg_dst, seg_src, seg_base, tile_desc, out, stride, tiles_per_slot, n_slots, 1);
    c->launches += 2;
    SP_CUDA(cudaGetLastError());
}

}  // namespace sp
