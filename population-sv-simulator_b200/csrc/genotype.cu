// Rows G, C and N of SURVEY.md §8a — genotype matrix, per-haplotype collision fixer, carrier counts
// (SimulateHaplotypes.buildPopulation, reference src/SimulateHaplotypes.java:603-718, and the counting half of
// buildVariant2haplotypes :967-987).
//
//   genotype_bits_kernel .... cell (variant j, haplotype i) consumes LCG steps S1+2(j*H+i)+1,+2 of the single
//                             java.util.Random stream (:624). Each thread jumps to its first cell in O(log n) and
//                             then strides by 2H steps per variant; carrier iff draw53 <= floor(AF*2^53)
//                             (integer form of nextDouble()<=AF). One thread = 32 variants of one haplotype =
//                             one word of the haplotype-major bit matrix.
//   count/expand kernels .... popcount + scan -> ascending per-haplotype variant lists (CSR).
//   collision_count_kernel .. forward edges per list entry inside the reference's scan window (:647-658).
//   collision_fix_kernel .... one warp-sized CTA per haplotype, lane 0 replays java.util.PriorityQueue
//                             (siftUp / siftDown / removeAt) exactly, with an index map instead of indexOf's scan.
//   compact_count_kernel .... survivors -> contiguous CSR + per-variant carrier counts (atomics).
#include "common.cuh"
#include "java_random.cuh"
#include "population.h"

namespace sp {

__global__ void __launch_bounds__(256)
genotype_bits_kernel(const int64_t* __restrict__ thr, int32_t V, int32_t W, int32_t H_total, int32_t h0, int32_t H_loc,
                     uint64_t state0, Affine stride, uint32_t* __restrict__ bits) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)W * H_loc) return;
    const int32_t w = (int32_t)(t % W), il = (int32_t)(t / W);
    const int64_t i = h0 + il;
    const int32_t j0 = w * 32;
    uint64_t s = affine_apply(lcg_jump(2ull * ((uint64_t)j0 * (uint64_t)H_total + (uint64_t)i)), state0);
    uint32_t mask = 0;
    const int32_t nb = min(32, V - j0);
    for (int32_t b = 0; b < nb; b++) {
        const int64_t d = (int64_t)draw53_from_state(s);
        if (d <= thr[j0 + b]) mask |= 1u << b;
        s = affine_apply(stride, s);
    }
    bits[(int64_t)il * W + w] = mask;
}

__global__ void __launch_bounds__(256)
count_bits_kernel(const uint32_t* __restrict__ bits, int32_t W, int32_t* __restrict__ cnt) {
    __shared__ int32_t ws[8];
    const int il = blockIdx.x;
    int32_t c = 0;
    for (int32_t w = threadIdx.x; w < W; w += 256) c += __popc(bits[(int64_t)il * W + w]);
    for (int d = 16; d > 0; d >>= 1) c += __shfl_down_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int32_t s = 0;
        for (int k = 0; k < 8; k++) s += ws[k];
        cnt[il] = s;
    }
}

// Exclusive scan of int32 counts into int64 offsets (n+1 entries), single CTA.
__global__ void __launch_bounds__(1024)
scan_counts_kernel(const int32_t* __restrict__ cnt, int32_t n, int64_t* __restrict__ off, int32_t* __restrict__ max_out = nullptr) {
    __shared__ int64_t ws[32];
    __shared__ int64_t carry;
    __shared__ int32_t mx_s;
    if (threadIdx.x == 0) { carry = 0; mx_s = 0; }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int32_t base = 0; base < n; base += 1024) {
        const int32_t k = base + threadIdx.x;
        const int64_t v = k < n ? cnt[k] : 0;
        if (max_out && v > 0) atomicMax(&mx_s, (int32_t)v);
        int64_t incl = v;
        for (int d = 1; d < 32; d <<= 1) {
            const int64_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) ws[warp] = incl;
        __syncthreads();
        int64_t wp = 0, tot = 0;
        for (int q = 0; q < 32; q++) { const int64_t s = ws[q]; if (q < warp) wp += s; tot += s; }
        if (k < n) off[k] = carry + wp + incl - v;
        __syncthreads();
        if (threadIdx.x == 0) carry += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) { off[n] = carry; if (max_out) *max_out = mx_s; }
}

__global__ void __launch_bounds__(256)
expand_bits_kernel(const uint32_t* __restrict__ bits, int32_t W, const int64_t* __restrict__ off, int32_t* __restrict__ ids) {
    __shared__ int32_t ws[8];
    __shared__ int32_t carry;
    const int il = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    int32_t* out = ids + off[il];
    for (int32_t base = 0; base < W; base += 256) {
        const int32_t w = base + threadIdx.x;
        uint32_t m = w < W ? bits[(int64_t)il * W + w] : 0u;
        const int32_t v = __popc(m);
        int32_t incl = v;
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) ws[warp] = incl;
        __syncthreads();
        int32_t wp = 0, tot = 0;
        for (int q = 0; q < 8; q++) { const int32_t s = ws[q]; if (q < warp) wp += s; tot += s; }
        int32_t p = carry + wp + incl - v;
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            out[p++] = w * 32 + b;
        }
        __syncthreads();
        if (threadIdx.x == 0) carry += tot;
        __syncthreads();
    }
}

// Variant.collidesWith(this=K, other=J, 5), src/SimulateHaplotypes.java:1214-1234 (non-INS pairs always collide).
__device__ __forceinline__ bool collides_with(int tK, int pK, int lK, int tJ, int pJ, int lJ, int d) {
    if (tK == SP_INS) {
        if (tJ == SP_INS) return abs(pK - pJ) <= d;
        return pK >= pJ - d && pK <= pJ + lJ - 1 + d;
    }
    if (tJ == SP_INS) return pJ >= pK - d && pJ <= pK + lK - 1 + d;
    const int first1 = min(pK, pJ), last1 = max(pK + lK - 1, pJ + lJ - 1);
    if (first1 <= last1) return true;
    const int f1 = pK, l1 = pK + lK - 1, f2 = pJ, l2 = pJ + lJ - 1;
    return (f1 > l2 && f1 <= l2 + d) || (l1 < f2 && l1 >= f2 - d);
}

constexpr int COLLISION_DISTANCE = 5;  // :605

// fdeg[off+j] = number of forward edges of list entry j; nedges[il] = their sum.
__global__ void __launch_bounds__(128)
collision_count_kernel(const int8_t* __restrict__ type, const int32_t* __restrict__ pos, const int32_t* __restrict__ len,
                       const int64_t* __restrict__ off, const int32_t* __restrict__ ids, int32_t* __restrict__ fdeg,
                       int32_t* __restrict__ nedges) {
    __shared__ int32_t ws[4];
    const int il = blockIdx.x;
    const int64_t lo = off[il];
    const int32_t n = (int32_t)(off[il + 1] - lo);
    int32_t total = 0;
    for (int32_t j = threadIdx.x; j < n; j += 128) {
        const int32_t vj = ids[lo + j];
        const int tJ = type[vj], pJ = pos[vj], lJ = len[vj];
        const int lastPosition = pJ + (tJ == SP_INS ? 0 : lJ - 1);
        int32_t c = 0;
        for (int32_t k = j + 1; k < n; k++) {
            const int32_t vk = ids[lo + k];
            const int pK = pos[vk];
            if (pK > lastPosition + COLLISION_DISTANCE) break;
            if (collides_with(type[vk], pK, len[vk], tJ, pJ, lJ, COLLISION_DISTANCE)) c++;
        }
        fdeg[lo + j] = c;
        total += c;
    }
    for (int d = 16; d > 0; d >>= 1) total += __shfl_down_sync(0xffffffffu, total, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = total;
    __syncthreads();
    if (threadIdx.x == 0) nedges[il] = ws[0] + ws[1] + ws[2] + ws[3];
}

// java.util.PriorityQueue on QueueElement (ordered by degree, identity by node), SURVEY.md Appendix A.2.
// `where` replaces indexOf()'s linear scan: every node is in the queue at most once, so "first index whose
// element equals o" is simply the node's current slot.
struct JavaHeap {
    int32_t* q;
    int32_t* where;
    const int32_t* degree;
    int32_t size;
    __device__ __forceinline__ void put(int32_t k, int32_t x) { q[k] = x; where[x] = k; }
    __device__ void siftUp(int32_t k, int32_t x) {
        while (k > 0) {
            const int32_t parent = (k - 1) >> 1;
            const int32_t e = q[parent];
            if (degree[x] >= degree[e]) break;
            put(k, e);
            k = parent;
        }
        put(k, x);
    }
    __device__ void siftDown(int32_t k, int32_t x, int32_t n) {
        const int32_t half = n >> 1;
        while (k < half) {
            int32_t child = (k << 1) + 1;
            int32_t c = q[child];
            const int32_t right = child + 1;
            if (right < n && degree[c] > degree[q[right]]) c = q[child = right];
            if (degree[x] <= degree[c]) break;
            put(k, c);
            k = child;
        }
        put(k, x);
    }
    __device__ void add(int32_t x) { const int32_t i = size; siftUp(i, x); size = i + 1; }
    __device__ int32_t poll() {
        const int32_t result = q[0];
        const int32_t n = --size;
        const int32_t x = q[n];
        where[result] = -1;
        if (n > 0) siftDown(0, x, n);
        return result;
    }
    __device__ bool remove(int32_t o) {
        const int32_t i = where[o];
        if (i < 0) return false;
        const int32_t s = --size;
        where[o] = -1;
        if (s != i) {
            const int32_t moved = q[s];
            siftDown(i, moved, s);
            if (q[i] == moved) siftUp(i, moved);
        }
        return true;
    }
};

// One CTA (one warp) per haplotype. Scratch per list entry (global index lo+j): fdeg (in), adj_off/where/q/degree
// (int32), removed (int8); adjacency adj[2*E] at eoff[il]. Survivors are compacted in place; cnt2[il] = new count.
__global__ void __launch_bounds__(32)
collision_fix_kernel(const int8_t* __restrict__ type, const int32_t* __restrict__ pos, const int32_t* __restrict__ len,
                     const int64_t* __restrict__ off, int32_t* __restrict__ ids, const int32_t* __restrict__ fdeg,
                     const int64_t* __restrict__ eoff, int32_t* __restrict__ adj, int32_t* __restrict__ adj_off_g,
                     int32_t* __restrict__ q_g, int32_t* __restrict__ where_g, int32_t* __restrict__ degree_g,
                     int8_t* __restrict__ removed_g, int32_t* __restrict__ cnt2, int32_t* __restrict__ n_fixed) {
    const int il = blockIdx.x;
    const int64_t lo = off[il];
    const int32_t n = (int32_t)(off[il + 1] - lo);
    const int64_t e0 = eoff[il];
    const int32_t E = (int32_t)(eoff[il + 1] - e0);
    if (E == 0) {  // if (nEdges==0) continue;  (:659)
        if (threadIdx.x == 0) cnt2[il] = n;
        return;
    }
    int32_t* adj_off = adj_off_g + lo + il;  // n+1 entries per haplotype
    int32_t* q = q_g + lo;
    int32_t* where = where_g + lo;
    int32_t* degree = degree_g + lo;
    int8_t* removed = removed_g + lo;
    int32_t* A = adj + 2 * e0;
    if (threadIdx.x == 0) {
        atomicAdd(n_fixed, 1);
        // degrees = forward + backward edges; adjacency in the order addEdge() builds it (:654, :721-736):
        // for node x first the earlier partners (ascending), then the later ones (ascending).
        for (int32_t j = 0; j < n; j++) degree[j] = 0;
        for (int32_t j = 0; j < n; j++) {
            const int32_t vj = ids[lo + j];
            const int tJ = type[vj], pJ = pos[vj], lJ = len[vj];
            const int lastPosition = pJ + (tJ == SP_INS ? 0 : lJ - 1);
            for (int32_t k = j + 1; k < n; k++) {
                const int32_t vk = ids[lo + k];
                const int pK = pos[vk];
                if (pK > lastPosition + COLLISION_DISTANCE) break;
                if (collides_with(type[vk], pK, len[vk], tJ, pJ, lJ, COLLISION_DISTANCE)) degree[k]++;  // backward count
            }
        }
        int32_t acc = 0;
        for (int32_t j = 0; j < n; j++) {
            adj_off[j] = acc;
            const int32_t b = degree[j];
            acc += b + fdeg[lo + j];
            where[j] = b;       // temporarily: number of backward partners (forward part starts after them)
            q[j] = 0;           // temporarily: backward fill cursor
        }
        adj_off[n] = acc;
        for (int32_t j = 0; j < n; j++) {
            const int32_t vj = ids[lo + j];
            const int tJ = type[vj], pJ = pos[vj], lJ = len[vj];
            const int lastPosition = pJ + (tJ == SP_INS ? 0 : lJ - 1);
            int32_t r = 0;
            for (int32_t k = j + 1; k < n; k++) {
                const int32_t vk = ids[lo + k];
                const int pK = pos[vk];
                if (pK > lastPosition + COLLISION_DISTANCE) break;
                if (collides_with(type[vk], pK, len[vk], tJ, pJ, lJ, COLLISION_DISTANCE)) {
                    A[adj_off[j] + where[j] + r++] = k;
                    A[adj_off[k] + q[k]++] = j;
                }
            }
        }
        // queue.clear(); add every variant of the haplotype in list order (:661-667)
        JavaHeap heap{q, where, degree, 0};
        for (int32_t j = 0; j < n; j++) { degree[j] = adj_off[j + 1] - adj_off[j]; removed[j] = 0; where[j] = -1; }
        for (int32_t j = 0; j < n; j++) heap.add(j);
        while (heap.size > 0) {  // :668-683
            const int32_t x = heap.poll();
            if (removed[x]) continue;
            for (int32_t k = adj_off[x]; k < adj_off[x + 1]; k++) {
                const int32_t nb = A[k];
                if (removed[nb]) continue;
                removed[nb] = 1;
                for (int32_t h = adj_off[nb]; h < adj_off[nb + 1]; h++) {
                    const int32_t e = A[h];
                    if (!heap.remove(e)) continue;
                    degree[e]--;
                    heap.add(e);
                }
            }
        }
        int32_t k = -1;
        for (int32_t j = 0; j < n; j++)
            if (!removed[j]) ids[lo + (++k)] = ids[lo + j];
        cnt2[il] = k + 1;
    }
}

// Shared-memory variant of collision_fix_kernel for lists of up to `cap` entries (cap*38+64 bytes of dynamic
// shared memory): same algorithm, but the heap (ids and a copy of each element's degree, which is valid because the
// Java only changes an element's degree while it is outside the queue, :678-680), the index map, the adjacency
// offsets and the per-entry variant fields live on chip, so the serial replay of java.util.PriorityQueue runs at
// shared-memory latency instead of L2 latency.
__global__ void __launch_bounds__(32)
collision_fix_smem_kernel(const int8_t* __restrict__ type, const int32_t* __restrict__ pos, const int32_t* __restrict__ len,
                          const int64_t* __restrict__ off, int32_t* __restrict__ ids, const int32_t* __restrict__ fdeg,
                          const int64_t* __restrict__ eoff, int32_t* __restrict__ adj, int32_t cap,
                          int32_t* __restrict__ cnt2, int32_t* __restrict__ n_fixed) {
    extern __shared__ int32_t sm[];
    const int il = blockIdx.x;
    const int64_t lo = off[il];
    const int32_t n = (int32_t)(off[il + 1] - lo);
    const int64_t e0 = eoff[il];
    const int32_t E = (int32_t)(eoff[il + 1] - e0);
    if (E == 0) {
        if (threadIdx.x == 0) cnt2[il] = n;
        return;
    }
    int32_t* s_pos = sm;                  // [cap]
    int32_t* s_len = s_pos + cap;         // [cap]
    int32_t* s_adjoff = s_len + cap;      // [cap+1]
    int32_t* s_qid = s_adjoff + cap + 1;  // [cap] heap: element
    int32_t* s_qdeg = s_qid + cap;        // [cap] heap: its degree
    int32_t* s_where = s_qdeg + cap;      // [cap]
    int32_t* s_deg = s_where + cap;       // [cap]
    int32_t* s_id = s_deg + cap;          // [cap] variant ids of the list
    int32_t* s_fdeg = s_id + cap;         // [cap] forward edge counts
    int8_t* s_type = (int8_t*)(s_fdeg + cap);  // [cap]
    int8_t* s_removed = s_type + cap;          // [cap]
    for (int32_t j = threadIdx.x; j < n; j += 32) {
        const int32_t v = ids[lo + j];
        s_id[j] = v; s_fdeg[j] = fdeg[lo + j];
        s_pos[j] = pos[v]; s_len[j] = len[v]; s_type[j] = type[v];
        s_deg[j] = 0; s_removed[j] = 0; s_where[j] = -1;
    }
    __syncwarp();
    int32_t* A = adj + 2 * e0;
    if (threadIdx.x == 0) {
    atomicAdd(n_fixed, 1);
    auto window_end = [&](int32_t j) { return s_pos[j] + (s_type[j] == SP_INS ? 0 : s_len[j] - 1) + COLLISION_DISTANCE; };
    for (int32_t j = 0; j < n; j++) {
        if (s_fdeg[j] == 0) continue;
        const int lim = window_end(j);
        for (int32_t k = j + 1; k < n && s_pos[k] <= lim; k++)
            if (collides_with(s_type[k], s_pos[k], s_len[k], s_type[j], s_pos[j], s_len[j], COLLISION_DISTANCE)) s_deg[k]++;
    }
    int32_t acc = 0;
    for (int32_t j = 0; j < n; j++) {
        s_adjoff[j] = acc;
        const int32_t b = s_deg[j];
        acc += b + s_fdeg[j];
        s_where[j] = b;   // temporarily: number of earlier partners
        s_qid[j] = 0;     // temporarily: fill cursor of the earlier partners
    }
    s_adjoff[n] = acc;
    for (int32_t j = 0; j < n; j++) {
        if (s_fdeg[j] == 0) continue;
        const int lim = window_end(j);
        int32_t r = 0;
        for (int32_t k = j + 1; k < n && s_pos[k] <= lim; k++)
            if (collides_with(s_type[k], s_pos[k], s_len[k], s_type[j], s_pos[j], s_len[j], COLLISION_DISTANCE)) {
                A[s_adjoff[j] + s_where[j] + r++] = k;
                A[s_adjoff[k] + s_qid[k]++] = j;
            }
    }
    for (int32_t j = 0; j < n; j++) { s_deg[j] = s_adjoff[j + 1] - s_adjoff[j]; s_where[j] = -1; }
    int32_t size = 0;
    auto put = [&](int32_t k, int32_t x, int32_t d) { s_qid[k] = x; s_qdeg[k] = d; s_where[x] = k; };
    auto siftUp = [&](int32_t k, int32_t x, int32_t d) {
        while (k > 0) {
            const int32_t parent = (k - 1) >> 1;
            const int32_t ed = s_qdeg[parent];
            if (d >= ed) break;
            put(k, s_qid[parent], ed);
            k = parent;
        }
        put(k, x, d);
    };
    auto siftDown = [&](int32_t k, int32_t x, int32_t d, int32_t m) {
        const int32_t half = m >> 1;
        while (k < half) {
            int32_t child = (k << 1) + 1;
            const int32_t right = child + 1;
            int32_t cd = s_qdeg[child];
            if (right < m) { const int32_t rd = s_qdeg[right]; if (cd > rd) { child = right; cd = rd; } }
            if (d <= cd) break;
            put(k, s_qid[child], cd);
            k = child;
        }
        put(k, x, d);
    };
    for (int32_t j = 0; j < n; j++) { siftUp(size, j, s_deg[j]); size++; }
    while (size > 0) {
        const int32_t x = s_qid[0];
        const int32_t m = --size;
        s_where[x] = -1;
        if (m > 0) siftDown(0, s_qid[m], s_qdeg[m], m);
        if (s_removed[x]) continue;
        for (int32_t k = s_adjoff[x]; k < s_adjoff[x + 1]; k++) {
            const int32_t nb = A[k];
            if (s_removed[nb]) continue;
            s_removed[nb] = 1;
            for (int32_t h = s_adjoff[nb]; h < s_adjoff[nb + 1]; h++) {
                const int32_t e = A[h];
                const int32_t i = s_where[e];
                if (i < 0) continue;            // if (!queue.remove(element)) continue;
                const int32_t s2 = --size;      // removeAt(i)
                s_where[e] = -1;
                if (s2 != i) {
                    const int32_t moved = s_qid[s2], md = s_qdeg[s2];
                    siftDown(i, moved, md, s2);
                    if (s_qid[i] == moved) siftUp(i, moved, md);
                }
                const int32_t nd = --s_deg[e];  // element.degree--; queue.add(element);
                siftUp(size, e, nd);
                size++;
            }
        }
    }
    }  // lane 0
    __syncwarp();
    // survivors keep their order (:684-689): warp-wide ballot compaction back into the list
    int32_t kept = 0;
    for (int32_t base = 0; base < n; base += 32) {
        const int32_t j = base + threadIdx.x;
        const bool keep = j < n && !s_removed[j];
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) ids[lo + kept + __popc(m & ((1u << threadIdx.x) - 1u))] = s_id[j];
        kept += __popc(m);
    }
    if (threadIdx.x == 0) cnt2[il] = kept;
}

// Survivors -> contiguous CSR (off2) and per-variant carrier counts.
__global__ void __launch_bounds__(256)
compact_count_kernel(const int64_t* __restrict__ off, const int32_t* __restrict__ ids, const int32_t* __restrict__ cnt2,
                     const int64_t* __restrict__ off2, int32_t* __restrict__ ids2, int32_t* __restrict__ counts) {
    const int il = blockIdx.x;
    const int64_t a = off[il], b = off2[il];
    const int32_t n = cnt2[il];
    for (int32_t j = threadIdx.x; j < n; j += 256) {
        const int32_t v = ids[a + j];
        ids2[b + j] = v;
        atomicAdd(&counts[v], 1);
    }
}

// ------------------------------------------------------------------------------------------------ host
// Host copies of the local lists and carrier counts: only the writers and the list getters need them, a batch that goes
// straight from rows G/C/N to the plan kernel never moves them off the device.
void fetch_host_lists(sp_ctx* c, PopulationState& P) {
    if (P.host_lists_valid) return;
    cudaStream_t st = c->pop_stream ? c->pop_stream : c->stream;
    const int32_t V = P.n_variants, Hl = P.h1 - P.h0;
    P.h_ids.assign((size_t)P.h_off[Hl], 0);
    if (P.h_off[Hl] > 0) SP_CUDA(cudaMemcpyAsync(P.h_ids.data(), P.d_ids2.p, (size_t)P.h_off[Hl] * 4, cudaMemcpyDeviceToHost, st));
    P.h_counts_local.assign(V, 0);
    if (V > 0) SP_CUDA(cudaMemcpyAsync(P.h_counts_local.data(), P.d_counts.p, (size_t)V * 4, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    P.host_lists_valid = true;
}

void run_genotype_and_collisions(sp_ctx* c, PopulationState& P) {
    // On the planning stream when there is one (sp_set_population_stream): the plan kernel that read the previous batch's
    // lists ran on the same stream, so they are free to be overwritten in stream order.
    cudaStream_t st = c->pop_stream ? c->pop_stream : c->stream;
    const int32_t V = P.n_variants, H = P.n_haplotypes_total, h0 = P.h0, Hl = P.h1 - P.h0;
    const int32_t W = (V + 31) / 32;
    P.host_lists_valid = false;
    P.d_thr.reserve(std::max(V, 1));
    P.d_bits.reserve(std::max<size_t>((size_t)W * Hl, 1));
    P.d_cnt.reserve(std::max(Hl, 1));
    P.d_off.reserve(Hl + 1);
    P.d_scalars.reserve(4);
    if (!P.thr_uploaded) {
        SP_CUDA(cudaMemcpyAsync(P.d_thr.p, P.h_thr.data(), (size_t)V * 8, cudaMemcpyHostToDevice, st));
        P.thr_uploaded = true;
    }
    const Affine stride = lcg_jump(2ull * (uint64_t)H);
    const int64_t nthreads = (int64_t)W * Hl;
    if (nthreads > 0) {
        genotype_bits_kernel<<<(unsigned)((nthreads + 255) / 256), 256, 0, st>>>(P.d_thr.p, V, W, H, h0, Hl, P.state_after_sampling, stride, P.d_bits.p);
        count_bits_kernel<<<Hl, 256, 0, st>>>(P.d_bits.p, W, P.d_cnt.p);
        c->launches += 2;
    }
    scan_counts_kernel<<<1, 1024, 0, st>>>(P.d_cnt.p, Hl, P.d_off.p, P.d_scalars.p);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    // The list buffers are sized by the bound V*Hl when that is small (a batch), so that the number of carriers does not
    // have to come back to the host before the lists are expanded; a whole population takes the exact total first.
    int64_t* hp = c->pinned((size_t)(Hl + 1) * 8 + 64);
    int64_t total_cap = (int64_t)V * Hl;
    const bool bounded = total_cap <= ((int64_t)1 << 24);
    if (!bounded) {
        SP_CUDA(cudaMemcpyAsync(hp, P.d_off.p + Hl, 8, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        total_cap = hp[0];
    }
    P.d_ids.reserve(std::max<int64_t>(total_cap, 1));
    P.d_fdeg.reserve(std::max<int64_t>(total_cap, 1));
    P.d_nedges.reserve(std::max(Hl, 1));
    P.d_eoff.reserve(Hl + 1);
    if (Hl > 0) {
        expand_bits_kernel<<<Hl, 256, 0, st>>>(P.d_bits.p, W, P.d_off.p, P.d_ids.p);
        collision_count_kernel<<<Hl, 128, 0, st>>>(c->d_type.p, c->d_pos.p, c->d_len.p, P.d_off.p, P.d_ids.p, P.d_fdeg.p, P.d_nedges.p);
        c->launches += 2;
    }
    scan_counts_kernel<<<1, 1024, 0, st>>>(P.d_nedges.p, Hl, P.d_eoff.p);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    // one round trip: total carriers, longest list, total edges (+ the offsets when the raw matrix is kept for tests)
    int32_t* hmax = (int32_t*)(hp + 2);
    SP_CUDA(cudaMemcpyAsync(hp, P.d_off.p + Hl, 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaMemcpyAsync(hp + 1, P.d_eoff.p + Hl, 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaMemcpyAsync(hmax, P.d_scalars.p, 4, cudaMemcpyDeviceToHost, st));
    if (P.keep_pre_collision) {
        P.h_pre_off.assign(Hl + 1, 0);
        SP_CUDA(cudaMemcpyAsync(P.h_pre_off.data(), P.d_off.p, (size_t)(Hl + 1) * 8, cudaMemcpyDeviceToHost, st));
    }
    SP_CUDA(cudaStreamSynchronize(st));
    const int64_t total = hp[0], etotal = hp[1];
    const int32_t max_n = hmax[0];
    P.pre_total = total;
    if (P.keep_pre_collision) {
        P.h_pre_ids.assign((size_t)total, 0);
        if (total > 0) SP_CUDA(cudaMemcpyAsync(P.h_pre_ids.data(), P.d_ids.p, (size_t)total * 4, cudaMemcpyDeviceToHost, st));
    }
    P.d_adj.reserve(std::max<int64_t>(2 * etotal, 1));
    P.d_cnt2.reserve(std::max(Hl, 1));
    P.d_off2.reserve(Hl + 1);
    P.d_counts.reserve(std::max(V, 1));
    SP_CUDA(cudaMemsetAsync(P.d_scalars.p + 1, 0, 4, st));
    int32_t* d_nfixed = P.d_scalars.p + 1;
    SP_CUDA(cudaMemsetAsync(P.d_counts.p, 0, (size_t)std::max(V, 1) * 4, st));
    const size_t smem_bytes = (size_t)max_n * 38 + 64;
    const bool force_global = getenv("SYNTHPOP_FORCE_GLOBAL_COLLISION_FIX") != nullptr;  // test hook
    if (Hl > 0 && smem_bytes <= 200 * 1024 && !force_global) {
        // the attribute is per device and per context: set on every launch (cheap), not once per process
        SP_CUDA(cudaFuncSetAttribute(collision_fix_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        collision_fix_smem_kernel<<<Hl, 32, smem_bytes, st>>>(c->d_type.p, c->d_pos.p, c->d_len.p, P.d_off.p, P.d_ids.p, P.d_fdeg.p,
                                                              P.d_eoff.p, P.d_adj.p, max_n, P.d_cnt2.p, d_nfixed);
        c->launches++;
    } else if (Hl > 0) {
        P.d_adj_off.reserve(std::max<int64_t>(total + Hl + 1, 1));
        P.d_q.reserve(std::max<int64_t>(total, 1));
        P.d_where.reserve(std::max<int64_t>(total, 1));
        P.d_degree.reserve(std::max<int64_t>(total, 1));
        P.d_removed.reserve(std::max<int64_t>(total, 1));
        collision_fix_kernel<<<Hl, 32, 0, st>>>(c->d_type.p, c->d_pos.p, c->d_len.p, P.d_off.p, P.d_ids.p, P.d_fdeg.p, P.d_eoff.p,
                                                P.d_adj.p, P.d_adj_off.p, P.d_q.p, P.d_where.p, P.d_degree.p, P.d_removed.p,
                                                P.d_cnt2.p, d_nfixed);
        c->launches++;
    }
    scan_counts_kernel<<<1, 1024, 0, st>>>(P.d_cnt2.p, Hl, P.d_off2.p);
    c->launches++;
    P.d_ids2.reserve(std::max<int64_t>(bounded ? total_cap : total, 1));
    if (Hl > 0) {
        compact_count_kernel<<<Hl, 256, 0, st>>>(P.d_off.p, P.d_ids.p, P.d_cnt2.p, P.d_off2.p, P.d_ids2.p, P.d_counts.p);
        c->launches++;
    }
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaMemcpyAsync(hp, P.d_off2.p, (size_t)(Hl + 1) * 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaMemcpyAsync(hp + Hl + 1, d_nfixed, 4, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    P.h_off.assign(hp, hp + Hl + 1);
    P.n_haplotypes_with_collision = *(int32_t*)(hp + Hl + 1);
}

}  // namespace sp
