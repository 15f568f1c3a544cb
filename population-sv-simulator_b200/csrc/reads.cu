// Row P of SURVEY.md §8a — the read kernel: the reference's patched PBSIM2 in model mode
// (simulate_by_model docker/pbsim2.cpp:1641-1711, mutate :2260-2413), re-designed for B200.
//
// Every random decision of the reference is one glibc rand() call and a read consumes its draws in a data-dependent
// order (Appendix A.6): quality chain, deletion rejection loop, start offset, then one draw per base plus 1-2 per
// error event. The work per base is a few dozen dependent integer operations and no reuse, so the design is
// ONE THREAD PER READ (first version: one warp per read around a shared-memory draw ring — 16 warp instructions per
// base, most of them cooperation overhead; see profiles/r01_reads_kernel_v1_lines.txt .. r01c_reads_kernel_lines.txt):
//   * the thread runs the reference's loops as they are, with the glibc TYPE_3 generator r[i] = r[i-31] + r[i-3] as
//     a 32-slot ring in shared memory (csrc/reads_core.cuh: value i in slot i & 31 reads slots i+1 and i-3; positions
//     wrap by masking) laid out [slot][thread] (bank = thread: conflict-free whatever each thread's phase), one add per
//     draw;
//   * a warp takes 32 reads that the host has grouped by (accuracy class, length), so its threads run the same
//     phases for nearly the same trip counts;
//   * per-read scratch (quality codes and deletion counters, one byte per position) lives in HBM interleaved
//     [word][lane], so the sequential passes are 128-byte coalesced and only the rejection loop's lookups scatter;
//   * the haplotype window is read in aligned 8-byte chunks and the read is written in aligned 8-byte chunks.
//   * the quality loops are unrolled by one lap of the ring (16 bases at two draws each, or 32): every read enters them
//     after the same number of draws, so the slot of each quality draw is a compile-time constant there (on a 31-slot
//     ring the lap was 31 bases and 22 KB of code; the instruction cache served 88 % of the requests);
//   * the pass over the bases asks for its scratch rows two loads ahead and for the next template chunk in one common
//     step per 8 bases, the deletion loop keeps two attempts in flight;
//   * the deletion rejection loop skips the scattered lookup of q[pos] whenever the draw is already above the largest
//     threshold any position of the read can have (same decisions, same draws);
//   * in per-read-stream mode the loop header (length, accuracy, quota cut, record offsets, work order) is computed on
//     the device as map + scan + cut (plan_*_kernel), the host only learns four scalars.
// Replay mode feeds the same code from the reference's logged draws instead of the generator.
// Everything is integer arithmetic; the only floating-point expressions of the reference's loop (:1682, :2282) are
// evaluated on the host with the reference's own types when the per-read plan is made.
#include "reads_core.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>
#include <chrono>

namespace sp {

// The oriented, upper-cased template (:2316-2341): seq[t] for t in [0, len'), fetched through aligned 8-byte chunks.
// The strand is data (a step of +-1 and an xor mask), not a branch: the threads of a warp alternate strands.
struct Template {
    const unsigned long long* base8;  // 8-byte aligned address at or below seq[0]
    int32_t o0;                       // seq[0] - base8, 0..7
    int32_t dir;                      // +1 / -1
    const uint8_t* lut;               // shared memory: byte -> upper-cased base (+) or its complement (-), :2316-2341
    int32_t cur;                      // index (relative to base8) of the cached 8-byte chunk
    uint32_t lo, hi;
    int32_t nw;                       // PF bit 3: index of the chunk held in (nlo, nhi), requested at the last base index
    uint32_t nlo, nhi;                // that is a multiple of 8 — by every lane in the same step
    __device__ __forceinline__ void begin(const uint8_t* w, int32_t lp, bool forward, const uint8_t* s_lut) {
        const uint8_t* origin = forward ? w : w + (lp - 1);  // first base of the window (+) or its last base (-)
        o0 = (int32_t)((uintptr_t)origin & 7);
        base8 = (const unsigned long long*)(origin - o0);
        dir = forward ? 1 : -1; lut = forward ? s_lut : s_lut + LUT_REV_OFFSET; cur = INT32_MIN; lo = hi = 0;
        nw = INT32_MIN; nlo = nhi = 0;
    }
    // PF bit 3: the lanes of a warp cross their chunk boundaries in different steps, so with a load at the crossing some lane
    // makes the warp wait for memory in EVERY step (a quarter of all stall samples, profiles/r02j). Here every lane requests the
    // chunk after its current one in the SAME step, once per 8 bases (k is the warp's common base index), into a second
    // register pair; a crossing is then a register move, and only the crossings right behind the common request still wait.
    // A lane that jumps over a whole chunk (deletions) or starts a read falls back to the load at the crossing.
    template <int PF>
    __device__ __forceinline__ uint8_t at_k(int32_t t, int32_t k, int32_t lp) {
        const int32_t rel = o0 + dir * t;
        const int32_t w = rel >> 3;
        if (w != cur) {
            if (w == nw) { lo = nlo; hi = nhi; }
            else {
                const unsigned long long chunk = __ldg(base8 + w);
                lo = (uint32_t)chunk; hi = (uint32_t)(chunk >> 32);
            }
            cur = w;
        }
        if ((k & 7) == 0) {
            nw = w + dir;
            if (dir * (nw - ((o0 + dir * (lp - 1)) >> 3)) <= 0) {   // still a chunk of this read's window (which lies inside the sequence)
                const unsigned long long chunk = __ldg(base8 + nw);
                nlo = (uint32_t)chunk; nhi = (uint32_t)(chunk >> 32);
            } else nw = INT32_MIN;
        }
        return lut[__byte_perm(lo, hi, (uint32_t)(rel & 7)) & 0xffu];
    }
    // PF bit 0: two bases before the lane leaves its chunk, the next chunk is requested into L1 (a prefetch has no
    // destination register, so no lane's request makes the warp wait); bit 2: on entering a 128-byte line, the next line is
    // requested into L2.
    template <int PF>
    __device__ __forceinline__ uint8_t at(int32_t t) {
        const int32_t rel = o0 + dir * t;   // 32-bit: a read is far shorter than 2^31 bases
        const int32_t w = rel >> 3;         // floor, also for the negative offsets of the reverse strand
        if (w != cur) {
            const unsigned long long chunk = __ldg(base8 + w);
            lo = (uint32_t)chunk; hi = (uint32_t)(chunk >> 32); cur = w;
            if (PF & 4) {
                if ((w & 15) == (dir > 0 ? 0 : 15)) asm volatile("prefetch.global.L2 [%0];" ::"l"(base8 + (w + 16 * dir)));
            }
        }
        if (PF & 1) {
            if ((rel & 7) == ((7 + 5 * dir) >> 1)) asm volatile("prefetch.global.L1 [%0];" ::"l"(base8 + (w + dir)));
        }
        return lut[__byte_perm(lo, hi, (uint32_t)(rel & 7)) & 0xffu];
    }
};

// The read's bases go out in aligned 8-byte words at a pace that depends on the base index only (every 8th base), so
// the threads of a warp stay in step whatever the alignment of their records: word m of the record's 8-byte grid is
// assembled from the previous and the current group of 8 bases by a funnel shift.
struct SeqOut {
    uint8_t* out;
    int64_t seq0;    // global offset of base 0
    uint32_t a8;     // 8 * (seq0 & 7)
    uint64_t prev, cur;
    __device__ __forceinline__ void begin(uint8_t* o, int64_t at) { out = o; seq0 = at; a8 = 8u * (uint32_t)(at & 7); prev = 0; cur = 0; }
    __device__ __forceinline__ void put(int32_t k, uint8_t c) {
        cur |= (uint64_t)c << (8 * (k & 7));
        if ((k & 7) == 7) {
            const int64_t g = (seq0 & ~7ll) + (k - 7);  // aligned address of the word that can be completed now
            if (k == 7) {                                // the first word shares its low bytes with the header
                for (uint32_t b = a8; b < 64; b += 8) out[g + (b >> 3)] = (uint8_t)(cur >> (b - a8));
            } else {
                const uint64_t w = a8 ? (prev >> (64 - a8)) | (cur << a8) : cur;
                *(uint64_t*)(out + g) = w;
            }
            prev = cur; cur = 0;
        }
    }
    // the same with the accumulator as a shift register (two funnel shifts instead of a variable 64-bit shift and two ORs)
    __device__ __forceinline__ void put_shift(int32_t k, uint8_t c) {
        cur = (cur >> 8) | ((uint64_t)c << 56);
        if ((k & 7) == 7) {
            const int64_t g = (seq0 & ~7ll) + (k - 7);
            if (k == 7) {
                for (uint32_t b = a8; b < 64; b += 8) out[g + (b >> 3)] = (uint8_t)(cur >> (b - a8));
            } else {
                const uint64_t w = a8 ? (prev >> (64 - a8)) | (cur << a8) : cur;
                *(uint64_t*)(out + g) = w;
            }
            prev = cur; cur = 0;
        }
    }
    __device__ __forceinline__ void finish_shift(int32_t n) {
        if (n & 7) cur >>= 8 * (8 - (n & 7));   // the pending bases sit in the top bytes of the shift register
        finish(n);
    }
    // bases [0, n) were put; writes what is still pending, then the record's closing '\n'
    __device__ __forceinline__ void finish(int32_t n) {
        const int32_t done = n >= 8 ? (n & ~7) - (int32_t)(a8 >> 3) : 0;  // stream bytes already in memory
        for (int32_t i = done; i < n; i++) {
            const uint64_t src = (i >> 3) == (n >> 3) ? cur : prev;
            out[seq0 + i] = (uint8_t)(src >> (8 * (i & 7)));
        }
        out[seq0 + n] = '\n';
    }
};

// One read, start to end, by one thread (simulate_by_model :1657-1711 + mutate :2260-2413).
template <bool REPLAY, int PF>
__device__ __forceinline__ void simulate_read(const ReadJob& J, const ReadTablesDev& T, const int32_t* s_thre, Draws<REPLAY>& S,
                                              uint32_t* qw, uint32_t* dw, int64_t r, unsigned mask) {
    // `mask` = the lanes of this warp that hold a read. Trip counts differ a little between the threads of a warp; the
    // __syncwarp(mask) after every phase makes them start the next one together (independent thread scheduling would
    // otherwise let them drift apart for the rest of the read, one thread per issue slot).
    int32_t len = J.len[r], ndel = J.ndel[r];
    const int acc = J.acc[r];
    const int32_t lenp = len + ndel;
    const int64_t rec0 = J.rec_off[r];
    const bool bad = len < 2 || lenp >= J.L;  // rand() % (len-1) with len 1 traps in the reference; len' >= ref.len changes the output length
    if (bad) {
        job_error(J, len < 2 ? 1 : 2, r);
        len = 0; ndel = 0;  // every loop below runs zero times; the thread still joins the warp's phase barriers
    }
    if (REPLAY) { S.replay = J.replay + J.replay_off[r]; S.used = 0; }
    else S.seed(read_seed(J.seed, J.ref_num, r));
    S.next(); S.next();  // the length and accuracy draws (:1642, :1654), evaluated on the host when the plan was made
    __syncwarp(mask);

    // ---- quality codes (:1657-1680), four per scratch word; the deletion counters are cleared on the way
    uint32_t qmin4 = 0xffffffffu;  // bytewise minimum of the quality words: the read's lowest quality code bounds every
                                   // acceptance threshold of the deletion loop below
    // Every read reaches this point after the same number of draws, so the slot of each quality draw is a function of
    // the base index alone: one lap of the ring is unrolled with compile-time slots.
    if (T.exist_hmm[acc]) {
        const int a51 = acc * 51;
        uint32_t st = __ldg(T.init2state + acc * 101 + fastmod(S.next(), T.mod_init + acc) + 1);  // third draw, slot 1
        uint32_t word = 0;
        if (len > 0) word = hmm_emis(T, a51, st, S.next_at(2));
        // base p >= 1 draws at slots 1+2p (transition) and 2+2p (emission) of the 32-slot ring: 16 bases per lap
#pragma unroll 1
        for (int32_t p0 = 1; p0 < len; p0 += 16) {
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const int32_t p = p0 + j;
                if (p < len) {
                    st = hmm_tran(T, a51, st, S.next_at((3 + 2 * j) & 31));
                    word |= hmm_emis(T, a51, st, S.next_at((4 + 2 * j) & 31)) << (8 * (p & 3));
                    if ((p & 3) == 3) { qw[(p >> 2) * 32] = word; dw[(p >> 2) * 32] = 0; qmin4 = __vminu4(qmin4, word); word = 0; }
                }
            }
        }
        if (len & 3) { qw[(len >> 2) * 32] = word; dw[(len >> 2) * 32] = 0; qmin4 = __vminu4(qmin4, word | (0xffffffffu << (8 * (len & 3)))); }
        if (len > 0) S.skip(2u * (uint32_t)len - 1u);
    } else {
        const uint4 mf = __ldg((const uint4*)(T.mod_freq + acc));
        const uint8_t* f2q = T.freq2qc + acc * 1001;
        uint32_t word = 0;
        // two draws so far: base p draws at slot 1+p mod 32
#pragma unroll 1
        for (int32_t p0 = 0; p0 < len; p0 += 32) {
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const int32_t p = p0 + j;
                if (p < len) {
                    word |= (uint32_t)__ldg(f2q + fastmod(S.next_at((1 + j) & 31), mf.x, mf.y, mf.z) + 1) << (8 * (p & 3));
                    if ((p & 3) == 3) { qw[(p >> 2) * 32] = word; dw[(p >> 2) * 32] = 0; qmin4 = __vminu4(qmin4, word); word = 0; }
                }
            }
        }
        if (len & 3) { qw[(len >> 2) * 32] = word; dw[(len >> 2) * 32] = 0; qmin4 = __vminu4(qmin4, word | (0xffffffffu << (8 * (len & 3)))); }
        S.skip((uint32_t)len);
    }
    __syncwarp(mask);

    // ---- deletions by rejection (:2282-2292). The loop is driven by a warp vote so that the threads stay in step.
    {
        const FastMod ml = make_fastmod_dev((uint32_t)max(len - 1, 1));
        const uint32_t qm = min(min(qmin4 & 255u, (qmin4 >> 8) & 255u), min((qmin4 >> 16) & 255u, qmin4 >> 24));
        const int32_t tmax = s_thre[3 * PB_NQ + min(qm, (uint32_t)(PB_NQ - 1))];
        int32_t num = 0;
        if (PF & 16) {
            // Two attempts in flight: the scattered read of q[pos] is a trip to HBM for one byte, and the attempt after it
            // cannot start before it is answered (15 % of all stall samples, profiles/r02k). The draws of the attempts do
            // not depend on the answers, only their NUMBER does: attempts j+1 and j+2 are drawn (and their reads requested)
            // before attempt j is decided, and when attempt j turns out to be the last one the four draws made beyond it
            // are taken back (undraw). Same rand() calls in the same order for everything that counts. Each attempt has its
            // own registers (the loop is unrolled by three), so no attempt's read is waited for by another attempt's code.
            uint32_t p0 = 0, p1 = 0, p2 = 0, q0 = 0, q1 = 0, q2 = 0;
            int32_t v0 = 0, v1 = 0, v2 = 0;
            bool live = num < ndel;
#define SP_START(P, V, Q)                                                           \
    {                                                                               \
        P = fastmod(S.next(), ml.M, ml.sh, ml.d);                                   \
        V = (int32_t)(S.next() % 1000000u);                                         \
        if (V < tmax) Q = qw[(P >> 2) * 32];                                        \
    }
#define SP_RESOLVE(P, V, Q)                                                         \
    {                                                                               \
        if (V < tmax) {                                                             \
            const uint32_t sh = 8 * (P & 3);                                        \
            if (V < s_thre[(Q >> sh) & 255u]) {                                     \
                asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(dw + (P >> 2) * 32), "r"(1u << sh) : "memory"); \
                num++;                                                              \
            }                                                                       \
        }                                                                           \
        if (num >= ndel) { S.undraw(4); live = false; }                             \
    }
            if (live) { SP_START(p0, v0, q0) SP_START(p1, v1, q1) }
#pragma unroll 1
            for (;;) {
                if (!__any_sync(mask, live)) break;
                if (live) { SP_START(p2, v2, q2) SP_RESOLVE(p0, v0, q0) }
                if (!__any_sync(mask, live)) break;
                if (live) { SP_START(p0, v0, q0) SP_RESOLVE(p1, v1, q1) }
                if (!__any_sync(mask, live)) break;
                if (live) { SP_START(p1, v1, q1) SP_RESOLVE(p2, v2, q2) }
            }
#undef SP_START
#undef SP_RESOLVE
        } else
#pragma unroll 1
        for (;;) {
            const bool more = num < ndel;
            if (!__any_sync(mask, more)) break;
            if (more) {
                const uint32_t pos = fastmod(S.next(), ml.M, ml.sh, ml.d);
                const int32_t v = (int32_t)(S.next() % 1000000u);
                // v >= the largest threshold any position of this read can have: rejected whatever q[pos] is, and the
                // scattered read of q[pos] (one HBM access for one byte) is not made. Same decision, same draws.
                if (v < tmax) {   // (no `continue` here: every lane must reach the one vote instruction of this loop)
                    const uint32_t sh = 8 * (pos & 3);
                    const int qv = (qw[(pos >> 2) * 32] >> sh) & 255u;
                    if (v < s_thre[qv]) {
                        // one more deletion after position pos: a fire-and-forget reduction (no load to wait for). A
                        // counter byte that wraps (more than 255 deletions at one position) shows up as a wrong total in
                        // the pass over the bases below.
                        asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(dw + (pos >> 2) * 32), "r"(1u << sh) : "memory");
                        num++;
                    }
                }
            }
        }
    }

    // ---- start offset (:2306-2311), header, bases (:2348-2401)
    const uint32_t offset = bad ? 0u : S.next() % (uint32_t)(J.L - lenp + 1);
    const bool fwd = (r & 1) == 0;  // res_num % 2 == 0 before the increment (:2316)
    Template Tm;
    Tm.begin(J.seq + offset, lenp, fwd, (const uint8_t*)(s_thre + 4 * PB_NQ));
    int64_t o = rec0;
    if (!bad) {  // ">S<ref>_<n>Z", a dozen bytes
        for (int i = 0; i < J.id_len; i++) J.out[o++] = (uint8_t)J.id[i];
        char dig[20];
        int nd = 0;
        for (long long v = r + 1; v > 0; v /= 10) dig[nd++] = (char)('0' + v % 10);
        while (nd > 0) J.out[o++] = (uint8_t)dig[--nd];
        J.out[o++] = 'Z';
    }
    SeqOut O;
    O.begin(J.out, o);
    const bool emit = !J.no_error;
    int nrun = 0;
    bool drop = false;
    int32_t t = 0;  // template index: advances with every base except after an insertion, and over deleted bases
    uint32_t wq = 0, wd = 0;
    int32_t dsum = 0;   // deletions met along the read: must come to ndel
    __syncwarp(mask);
#pragma unroll 1
    for (int32_t k = 0;; k++) {
        const bool more = k < len;
        if (!__any_sync(mask, more)) break;
        if (more) {
            if ((k & 3) == 0) {
                wq = qw[(k >> 2) * 32]; wd = ndel > 0 ? dw[(k >> 2) * 32] : 0u;
                if (PF & 2) {   // the rows two loads ahead, into L1 (the scratch has two rows of slack at its end)
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(qw + ((k >> 2) + 2) * 32));
                    if (ndel > 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(dw + ((k >> 2) + 2) * 32));
                }
            }
            const int qv = wq & 255u; wq >>= 8;
            const int32_t dc = wd & 255u; wd >>= 8;
            dsum += dc;
            const uint8_t nt = (PF & 8) ? Tm.at_k<PF>(t, k, lenp) : Tm.at<PF>(t);
            const int32_t d = (int32_t)(S.next() % 1000000u);
            uint8_t outc = nt;
            int32_t adv = 1;
            if (d < s_thre[2 * PB_NQ + qv]) {  // an error event: ins_thre >= sub_thre (:1455-1459), so one test screens both
                if ((PF & 32) && d < s_thre[PB_NQ + qv]) {   // the same tables as shifts of a constant: no load on this path
                    const uint32_t sh = 8 * (S.next() % 3u);
                    if (nt == 'A') outc = (uint8_t)(0x434754u >> sh);        // "TGC"
                    else if (nt == 'T') outc = (uint8_t)(0x434741u >> sh);   // "AGC"
                    else if (nt == 'G') outc = (uint8_t)(0x435441u >> sh);   // "ATC"
                    else if (nt == 'C') outc = (uint8_t)(0x475441u >> sh);   // "ATG"
                    else outc = (uint8_t)(0x43475441u >> (8 * (S.next() % 4u)));   // "ATGC"
                } else if (PF & 32) {
                    const uint32_t idx = S.next() % 8u;
                    if (idx < 4) outc = (uint8_t)(0x43475441u >> (8 * idx));
                    adv = 0;
                } else
                if (d < s_thre[PB_NQ + qv]) {  // substitution (:2362-2378), per-base tables of set_mut :2211-2214
                    uint32_t idx = S.next() % 3u;
                    if (nt == 'A') outc = "TGC"[idx];
                    else if (nt == 'T') outc = "AGC"[idx];
                    else if (nt == 'G') outc = "ATC"[idx];
                    else if (nt == 'C') outc = "ATG"[idx];
                    else { idx = S.next() % 4u; outc = "ATGC"[idx]; }
                } else {                       // insertion (:2379-2387): the template base is not consumed
                    const uint32_t idx = S.next() % 8u;
                    if (idx < 4) outc = "ATGC"[idx];
                    adv = 0;
                }
            }
            t += adv + dc;
            if (emit) {
                if (PF & 32) {
                    if (outc != 'N') nrun = k + 1;        // here nrun is the index after the last base that is not 'N'
                    else if (k + 1 - nrun >= 5) drop = true;
                    O.put_shift(k, outc);
                } else {
                    nrun = outc == 'N' ? nrun + 1 : 0;
                    drop |= nrun >= 5;
                    O.put(k, outc);
                }
            }
        }
    }
    __syncwarp(mask);
    if (bad) { J.dropped[r] = 1; J.used[r] = 0; return; }
    if (dsum != ndel) job_error(J, 3, r);
    if (emit) { if (PF & 32) O.finish_shift(len); else O.finish(len); }
    else {  // pbsim2noError.cpp:1711 prints mut.seq, the oriented template of length len'
#pragma unroll 1
        for (int32_t i = 0; i < lenp; i++) {
            const uint8_t c = Tm.at<0>(i);
            nrun = c == 'N' ? nrun + 1 : 0;
            drop |= nrun >= 5;
            O.put(i, c);
        }
        O.finish(lenp);
    }
    J.dropped[r] = drop ? 1 : 0;
    J.used[r] = S.used;
    if (drop) {
        atomicAdd(&J.ctr[3], 1ull);
        atomicAdd(&J.ctr[4], (unsigned long long)(J.rec_off[r + 1] - rec0));
    }
}

// 10 CTAs per SM (48 registers, a few spilled words): measured best against 8, 9 and 12 with the default carve-out
template <bool REPLAY, int PF>
__global__ void __launch_bounds__(RD_THREADS, 10)
reads_kernel(const ReadJob J, const ReadTablesDev T) {
    __shared__ uint32_t s_state[REPLAY ? 32 : RING * RD_THREADS];
    __shared__ int32_t s_thre[4 * PB_NQ + (LUT_REV_OFFSET + 256) / 4];  // err_thre, sub_thre, ins_thre, max(err_thre[q..93]), strand tables
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 3 * PB_NQ; i += blockDim.x) s_thre[i] = T.thre[i];
    __syncthreads();
    {   // toupper and toupper+complement (non-ACGT bytes keep their upper-cased value on both strands)
        uint8_t* lut = (uint8_t*)(s_thre + 4 * PB_NQ);
        for (int i = threadIdx.x; i < 256; i += blockDim.x) {
            uint32_t ch = (uint32_t)i;
            ch -= (ch - 'a' <= 'z' - 'a') ? 32u : 0u;
            lut[i] = (uint8_t)ch;
            lut[LUT_REV_OFFSET + i] = (uint8_t)(ch == 'A' ? 'T' : ch == 'T' ? 'A' : ch == 'C' ? 'G' : ch == 'G' ? 'C' : ch);
        }
    }
    if (threadIdx.x == 0) {
        int32_t m = 0;
        for (int q = PB_NQ - 1; q >= 0; q--) { m = max(m, s_thre[q]); s_thre[3 * PB_NQ + q] = m; }
    }
    __syncthreads();
    const int64_t gw = (int64_t)blockIdx.x * (RD_THREADS / 32) + w;
    uint32_t* qw = J.scratch + gw * 2 * J.scratch_words * 32 + lane;
    uint32_t* dw = qw + J.scratch_words * 32;
    Draws<REPLAY> S;
    S.st = s_state + (REPLAY ? 0 : threadIdx.x);
    S.replay = nullptr; S.used = 0; S.pos = 31;
    for (;;) {
        unsigned long long b0 = 0;
        if (lane == 0) b0 = atomicAdd(&J.ctr[0], 32ull);
        const int64_t base = (int64_t)__shfl_sync(FULL, b0, 0);
        if (base >= J.n_reads) break;
        const int64_t my = base + lane;
        const unsigned mask = __ballot_sync(FULL, my < J.n_reads);
        if (my < J.n_reads) simulate_read<REPLAY, PF>(J, T, s_thre, S, qw, dw, (int64_t)J.order[my], mask);
        __syncwarp();
    }
}

// The second design of the per-read code (csrc/reads_core.cuh: one draw per step at a ring position the compiler knows,
// quality codes and deletion counters of four positions in one 8-byte scratch row). Same launch shape and scratch size.
template <bool REPLAY, bool NOERR, int MIN_CTAS>
__global__ void __launch_bounds__(RD_THREADS, MIN_CTAS)
reads_sync_kernel(const ReadJob J, const ReadTablesDev T) {
    __shared__ uint32_t s_state[REPLAY ? 32 : RING * RD_THREADS];
    __shared__ int32_t s_thre[RD_TABLE_WORDS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    fill_cta_tables(s_thre, T.thre, threadIdx.x, blockDim.x, 0);
    __syncthreads();
    fill_cta_tables(s_thre, T.thre, threadIdx.x, blockDim.x, 1);
    __syncthreads();
    CtaTables C;
    C.bind(s_thre);
    const int64_t gw = (int64_t)blockIdx.x * (RD_THREADS / 32) + w;
    uint2* qd = (uint2*)(J.scratch + gw * 2 * J.scratch_words * 32) + lane;
    Draws<REPLAY> S;
    S.st = s_state + (REPLAY ? 0 : threadIdx.x);
    S.replay = nullptr; S.used = 0; S.pos = 31;
    for (;;) {
        unsigned long long b0 = 0;
        if (lane == 0) b0 = atomicAdd(&J.ctr[0], 32ull);
        const int64_t base = (int64_t)__shfl_sync(FULL, b0, 0);
        if (base >= J.n_reads) break;
        const int64_t my = base + lane;
        const unsigned mask = __ballot_sync(FULL, my < J.n_reads);
        if (my < J.n_reads) simulate_read_sync<REPLAY, NOERR>(J, T, C, S, qd, J.order[my], mask);
        __syncwarp();
    }
}

// Which per-read code runs: 1 = the first design (reads_kernel, default, with the variant bits of SYNTHPOP_READS_PF: 2 scratch
// rows requested two loads ahead, 8 template chunk requested by all lanes in the same step, 16 two deletion attempts in
// flight, 32 shift-register output and table-free error path), 2 = reads_sync_kernel (csrc/reads_core.cuh).
// SYNTHPOP_READS_KERNEL, SYNTHPOP_READS_PF and SYNTHPOP_READS_MIN_CTAS are A/B hooks for bench and profiling runs; every
// choice produces the same bytes (tests/test_reads_gpu.py::test_every_kernel_variant_writes_the_same_bytes).
using ReadKernel = void (*)(const ReadJob, const ReadTablesDev);
static ReadKernel pick_read_kernel(bool replay, bool no_error) {
    int version = 1, min_ctas = 10;
    if (const char* e = getenv("SYNTHPOP_READS_KERNEL")) version = atoi(e) == 2 ? 2 : 1;
    if (const char* e = getenv("SYNTHPOP_READS_MIN_CTAS")) min_ctas = atoi(e);
    if (version == 1) {
        int pf = 58;
        if (const char* e = getenv("SYNTHPOP_READS_PF")) pf = atoi(e);   // variants of the first design (bits: see simulate_read)
        if (replay) return pf == 0 ? reads_kernel<true, 0> : reads_kernel<true, 58>;   // replay runs the default variant's code: the golden vectors of the reference binary pin it
        return pf == 0 ? reads_kernel<false, 0> : pf == 10 ? reads_kernel<false, 10> : pf == 26 ? reads_kernel<false, 26> : reads_kernel<false, 58>;
    }
    if (replay) return no_error ? reads_sync_kernel<true, true, 10> : reads_sync_kernel<true, false, 10>;
    if (no_error) return reads_sync_kernel<false, true, 10>;
    return min_ctas == 8 ? reads_sync_kernel<false, false, 8> : min_ctas == 9 ? reads_sync_kernel<false, false, 9> : reads_sync_kernel<false, false, 10>;
}

// First two draws of reads [r0, r0+n) in per-read-stream mode (length and accuracy, :1642, :1654).
__global__ void __launch_bounds__(128)
first_draws_kernel(uint32_t seed, int64_t ref_num, int64_t r0, int64_t n, int32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t r[31];
    glibc_seed(read_seed(seed, ref_num, r0 + i), r);
    r[3] += r[0];
    out[2 * i] = (int32_t)(r[3] >> 1);
    out[2 * i + 1] = (int32_t)((r[4] + r[1]) >> 1);
}

// Kept records -> contiguous file bytes (only launched when some read was dropped).
__global__ void __launch_bounds__(256)
compact_records_kernel(const uint8_t* __restrict__ in, const int64_t* __restrict__ rec_off, const int64_t* __restrict__ koff,
                       int64_t n, uint8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = wid; r < n; r += nw) {
        const int64_t a = rec_off[r], len = koff[r + 1] - koff[r], b = koff[r];
        for (int64_t i = lane; i < len; i += 32) out[b + i] = in[a + i];
    }
}

// ------------------------------------------------------------------------------------------------- host
static ReadState& reads_of(sp_ctx* c) {
    if (!c->reads || !c->reads->configured) throw Error(SP_ERR_ARG, "sp_reads_configure has not been called");
    return *c->reads;
}

static int count_digits(long long v) {
    int n = 1;
    while (v >= 10) { v /= 10; n++; }
    return n;
}

// ---- the per-read plan on the device (per-read-stream mode) ---------------------------------------------------------
// simulate_by_model's loop header (:1641-1655) decides, from the first two draws of every read, its length and accuracy
// class; the loop ends with the read that reaches the quota. With one stream per read the draws of read i do not depend
// on the reads before it, so the header is a map + an inclusive scan + a cut, not a sequential loop.
__global__ void __launch_bounds__(128)
plan_lengths_kernel(uint32_t seed, int64_t ref_num, int64_t n, const int32_t* __restrict__ prob2len,
                    const uint8_t* __restrict__ prob2acc, uint32_t len_rv, uint32_t acc_rv, int64_t* __restrict__ len64,
                    uint8_t* __restrict__ acc) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t r[31];
    glibc_seed(read_seed(seed, ref_num, i), r);
    r[3] += r[0];
    const uint32_t d0 = r[3] >> 1, d1 = (r[4] + r[1]) >> 1;
    len64[i] = prob2len[d0 % len_rv + 1];       // :1642
    acc[i] = prob2acc[d1 % acc_rv + 1];         // :1654-1655
}

// res[0] = number of reads (the first read whose inclusive length sum reaches the quota is the last, :1641), res[1] =
// bases emitted; the last read is clipped to the quota, but not below len_min (:1643-1650).
__global__ void __launch_bounds__(256)
plan_cut_kernel(int64_t n, const int64_t* __restrict__ scan, int64_t* __restrict__ len64, int64_t quota, int64_t len_min,
                int64_t* __restrict__ res) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t prev = i ? scan[i - 1] : 0;
    if (prev < quota && scan[i] >= quota) {
        int64_t len = len64[i];
        if (prev + len > quota) {
            len = quota - prev;
            if (len < len_min) len = len_min;
        }
        len64[i] = len;
        res[0] = i + 1;
        res[1] = prev + len;
    }
}

// Per read: del_num_total = int((float)len * del_thre / 1000000 + 0.5) (:2282; the reference's own types: a float product
// and quotient, then a double sum), the record length of ">" id n "Z" seq "\n", the longest read.
__global__ void __launch_bounds__(256)
plan_finish_kernel(int64_t n_reads, const int64_t* __restrict__ len64, const uint8_t* __restrict__ acc,
                   const int32_t* __restrict__ delthre, int32_t no_error, int32_t id_len, int32_t* __restrict__ len,
                   int32_t* __restrict__ ndel, int64_t* __restrict__ reclen, int64_t* __restrict__ res) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_reads) return;
    if (i == n_reads) { reclen[i] = 0; return; }
    const int32_t l = (int32_t)len64[i];
    const float q = __fdiv_rn(__fmul_rn((float)(long long)l, (float)(long long)delthre[acc[i]]), 1000000.0f);
    const int32_t nd = (int32_t)((double)q + 0.5);
    len[i] = l; ndel[i] = nd;
    int digits = 1;
    for (int64_t v = i + 1; v >= 10; v /= 10) digits++;
    reclen[i] = (int64_t)id_len + digits + 1 + (no_error ? (int64_t)l + nd : (int64_t)l) + 1;
    atomicMax((unsigned long long*)&res[2], (unsigned long long)l);
}

// Grouping key of the read kernel's work order: longest first, then by accuracy class (stable in the read index).
__global__ void __launch_bounds__(256)
plan_keys_kernel(int64_t n_reads, const int32_t* __restrict__ len, const uint8_t* __restrict__ acc,
                 const int64_t* __restrict__ res, int32_t* __restrict__ key, int32_t* __restrict__ idx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    const int32_t max_len = max((int32_t)res[2], 2);
    const int32_t width = max(1, (max_len + 63) / 64);
    key[i] = (63 - min(63, len[i] / width)) * 101 + (int32_t)acc[i];
    idx[i] = (int32_t)i;
}

// Fills d_len, d_ndel, d_acc, d_rec_off, d_order for one sequence; returns the number of reads.
static int64_t plan_reads_on_device(sp_ctx* c, ReadState& R, int64_t ref_num, long long quota, int32_t id_len,
                                    int32_t* max_len_out, int64_t* bases_out, int64_t* bytes_out) {
    const ReadParams& p = R.par;
    const ReadTablesHost& t = R.tab;
    cudaStream_t st = c->stream;
    if (R.plan_tables_stale) {
        R.d_prob2len.reserve(t.prob2len.size()); R.d_prob2acc.reserve(t.prob2accuracy.size()); R.d_delthre.reserve(PB_ACCURACY_MAX + 1);
        int32_t delthre[PB_ACCURACY_MAX + 1];
        for (int a = 0; a <= PB_ACCURACY_MAX; a++) delthre[a] = (int32_t)(long)int(((100 - a) * p.del_rate) * 10000 + 0.5);  // :1682
        SP_CUDA(cudaMemcpyAsync(R.d_prob2len.p, t.prob2len.data(), t.prob2len.size() * 4, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(R.d_prob2acc.p, t.prob2accuracy.data(), t.prob2accuracy.size(), cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(R.d_delthre.p, delthre, sizeof delthre, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaStreamSynchronize(st));
        R.plan_tables_stale = false;
    }
    R.d_planres.reserve(4);
    int64_t res[4] = {-1, 0, 0, 0};
    if (quota <= 0) { *max_len_out = 2; *bases_out = 0; *bytes_out = 0; return 0; }
    const double mean_len = std::max(1.0, std::min(p.len_mean, (double)p.len_max));
    int64_t n = (int64_t)((double)quota / mean_len * 1.05) + 1024;
    for (;;) {
        n = std::min<int64_t>(n, (int64_t)1 << 27);
        R.d_len64.reserve((size_t)n); R.d_scan.reserve((size_t)n); R.d_acc.reserve((size_t)n);
        size_t tmp = 0;
        cub::DeviceScan::InclusiveSum(nullptr, tmp, R.d_len64.p, R.d_scan.p, (int)n, st);
        R.d_cubtmp.reserve(tmp + 256);
        SP_CUDA(cudaMemcpyAsync(R.d_planres.p, res, sizeof res, cudaMemcpyHostToDevice, st));
        plan_lengths_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(p.seed, ref_num, n, R.d_prob2len.p, R.d_prob2acc.p,
                                                                         (uint32_t)t.len_rand_value, (uint32_t)t.accuracy_rand_value,
                                                                         R.d_len64.p, R.d_acc.p);
        tmp = R.d_cubtmp.cap;
        SP_CUDA(cub::DeviceScan::InclusiveSum(R.d_cubtmp.p, tmp, R.d_len64.p, R.d_scan.p, (int)n, st));
        plan_cut_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, R.d_scan.p, R.d_len64.p, quota, p.len_min, R.d_planres.p);
        c->launches += 3;
        SP_CUDA(cudaGetLastError());
        SP_CUDA(cudaMemcpyAsync(res, R.d_planres.p, sizeof res, cudaMemcpyDeviceToHost, st));
        SP_CUDA(cudaStreamSynchronize(st));
        if (res[0] >= 0) break;
        if (n >= ((int64_t)1 << 27)) throw Error(SP_ERR_ARG, "read plan: more than 2^27 reads in one sequence");
        res[0] = -1; res[1] = 0; res[2] = 0; res[3] = 0;
        n *= 2;
    }
    const int64_t n_reads = res[0];
    const size_t n1 = (size_t)std::max<int64_t>(n_reads, 1);
    R.d_len.reserve(n1); R.d_ndel.reserve(n1); R.d_reclen.reserve(n1 + 1); R.d_rec_off.reserve(n1 + 1);
    R.d_key.reserve(n1); R.d_key2.reserve(n1); R.d_idx.reserve(n1); R.d_order.reserve(n1);
    size_t tmp_scan = 0, tmp_sort = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, R.d_reclen.p, R.d_rec_off.p, (int)(n_reads + 1), st);
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, R.d_key.p, R.d_key2.p, R.d_idx.p, R.d_order.p, (int)n_reads, 0, 13, st);
    R.d_cubtmp.reserve(std::max(tmp_scan, tmp_sort) + 256);
    const unsigned g = (unsigned)((n_reads + 256) / 256);
    plan_finish_kernel<<<g, 256, 0, st>>>(n_reads, R.d_len64.p, R.d_acc.p, R.d_delthre.p, p.no_error ? 1 : 0, id_len, R.d_len.p,
                                          R.d_ndel.p, R.d_reclen.p, R.d_planres.p);
    size_t tmp = R.d_cubtmp.cap;
    SP_CUDA(cub::DeviceScan::ExclusiveSum(R.d_cubtmp.p, tmp, R.d_reclen.p, R.d_rec_off.p, (int)(n_reads + 1), st));
    plan_keys_kernel<<<g, 256, 0, st>>>(n_reads, R.d_len.p, R.d_acc.p, R.d_planres.p, R.d_key.p, R.d_idx.p);
    tmp = R.d_cubtmp.cap;
    SP_CUDA(cub::DeviceRadixSort::SortPairs(R.d_cubtmp.p, tmp, R.d_key.p, R.d_key2.p, R.d_idx.p, R.d_order.p, (int)n_reads, 0, 13, st));
    c->launches += 4;
    SP_CUDA(cudaGetLastError());
    int64_t total = 0;
    SP_CUDA(cudaMemcpyAsync(res, R.d_planres.p, sizeof res, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaMemcpyAsync(&total, R.d_rec_off.p + n_reads, 8, cudaMemcpyDeviceToHost, st));
    SP_CUDA(cudaStreamSynchronize(st));
    *max_len_out = std::max<int32_t>(2, (int32_t)res[2]);
    *bases_out = res[1];
    *bytes_out = total;
    return n_reads;
}

// h_len / h_rec_off of the current batch, fetched only when a caller asks for the per-read log or the file bytes
static void ensure_host_plan(sp_ctx* c, ReadState& R) {
    if (!R.plan_on_device) return;
    R.h_len.assign((size_t)R.n_reads, 0);
    R.h_rec_off.assign((size_t)R.n_reads + 1, 0);
    if (R.n_reads > 0) SP_CUDA(cudaMemcpyAsync(R.h_len.data(), R.d_len.p, (size_t)R.n_reads * 4, cudaMemcpyDeviceToHost, c->stream));
    SP_CUDA(cudaMemcpyAsync(R.h_rec_off.data(), R.d_rec_off.p, ((size_t)R.n_reads + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
    R.plan_on_device = false;
}

// simulate_by_model's loop header and per-read plan (:1641-1655, :1682, :2282) for every read of one sequence, then the
// read kernel. d_seq must stay valid until the stream has run the kernel.
// One batch (= all reads of one sequence) in three phases, so that two batches can be in flight (simulate_pair below):
// plan_batch (the per-read plan; synchronises the context's stream), launch_batch (the read kernel on stream ks, nothing
// waits), collect_batch (waits for ks, reads the counters). simulate_on_device is the three in a row on the context's stream.
static bool plan_batch(sp_ctx* c, ReadState& R, const uint8_t* d_seq, int64_t L, int64_t ref_num) {
    const ReadParams& p = R.par;
    const ReadTablesHost& t = R.tab;
    cudaStream_t st = c->stream;
    const auto t_plan0 = std::chrono::steady_clock::now();
    if (L < 100) throw Error(SP_ERR_ARG, "ERROR: Reference is too short. Acceptable length >= 100.");
    if (L > 1000000000) throw Error(SP_ERR_ARG, "ERROR: Reference is too long. Acceptable length <= 1000000000.");
    const long long quota = (long long)(p.depth * (long)L);  // :566
    const bool replay = p.stream_mode == SP_READS_REPLAY;
    if (replay && !R.have_replay) throw Error(SP_ERR_ARG, "replay mode: sp_reads_set_replay has not been called");
    R.h_len.clear(); R.h_acc.clear(); R.h_ndel.clear();
    R.results_on_host = false;
    const std::string id = ">" + p.id_prefix + std::to_string(ref_num) + "_";  // record: id n "Z" seq "\n"
    int64_t n_reads = 0;
    int32_t max_len = 2;
    if (!replay) {
        // per-read streams: the plan is a map + scan + cut on the device, nothing but four scalars comes back
        int64_t bases = 0, bytes = 0;
        n_reads = plan_reads_on_device(c, R, ref_num, quota, (int32_t)id.size(), &max_len, &bases, &bytes);
        R.n_reads = n_reads; R.ref_num = (long)ref_num; R.emitted_bases = bases; R.out_bytes = bytes;
        R.plan_on_device = true;
        const size_t n1 = (size_t)std::max<int64_t>(n_reads, 1);
        R.d_dropped.reserve(n1); R.d_used.reserve(n1);
        R.d_ctr.reserve(16);
        R.d_out.reserve((size_t)R.out_bytes + 64, 0.05);
        SP_CUDA(cudaMemsetAsync(R.d_ctr.p, 0, 64, st));
        R.n_dropped = 0;
        if (n_reads == 0) return false;
    } else {
        R.plan_on_device = false;
        long long len_total = 0;
        std::vector<int32_t> first;  // two draws per read of the current batch
        int64_t batch0 = 0, batch_n = 0;
        auto fetch_batch = [&](int64_t r0) {
            if (replay) {
                batch0 = 0; batch_n = (int64_t)R.replay_off.size();
                first.resize((size_t)batch_n * 2);
                for (int64_t r = 0; r < batch_n; r++) {
                    const int64_t o = R.replay_off[r];
                    if (o < 0 || o + 1 >= (int64_t)R.replay_draws.size()) throw Error(SP_ERR_ARG, "replay: read offset outside the draw log");
                    first[2 * r] = R.replay_draws[o]; first[2 * r + 1] = R.replay_draws[o + 1];
                }
                return;
            }
            const double mean_len = std::max(1.0, std::min(p.len_mean, (double)p.len_max));
            int64_t n = (int64_t)((double)(quota - len_total) / mean_len * 1.05) + 1024;
            n = std::min<int64_t>(n, 1 << 26);
            R.d_first.reserve((size_t)n * 2);
            first_draws_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(p.seed, ref_num, r0, n, R.d_first.p);
            c->launches++;
            SP_CUDA(cudaGetLastError());
            first.resize((size_t)n * 2);
            SP_CUDA(cudaMemcpyAsync(first.data(), R.d_first.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
            SP_CUDA(cudaStreamSynchronize(st));
            batch0 = r0; batch_n = n;
        };
        int64_t r = 0;
        while (len_total < quota) {  // :1641
            if (r >= batch0 + batch_n) {
                if (replay && batch_n > 0) throw Error(SP_ERR_ARG, "replay: the log has fewer reads than the quota needs");
                fetch_batch(r);
                if (batch_n == 0) throw Error(SP_ERR_ARG, "replay: empty log");
            }
            const long d0 = first[2 * (r - batch0)], d1 = first[2 * (r - batch0) + 1];
            long len = t.prob2len[d0 % t.len_rand_value + 1];
            if (len_total + len > quota) {
                len = quota - len_total;
                if (len < p.len_min) len = p.len_min;
            }
            const int accuracy = t.prob2accuracy[d1 % t.accuracy_rand_value + 1];
            const long del_thre = int(((100 - accuracy) * p.del_rate) * 10000 + 0.5);      // :1682
            const long del_num_total = int((float)len * del_thre / 1000000 + 0.5);         // :2282 (float32 product)
            R.h_len.push_back((int32_t)len); R.h_acc.push_back((uint8_t)accuracy); R.h_ndel.push_back((int32_t)del_num_total);
            len_total += len;  // strlen(mut.new_seq) == len whenever len' < ref.len (the kernel rejects the other case)
            r++;
        }
        if (replay && r != (int64_t)R.replay_off.size())
            throw Error(SP_ERR_ARG, "replay: the log has " + std::to_string(R.replay_off.size()) + " reads, the quota ends after " + std::to_string(r));
        n_reads = r;
        R.n_reads = n_reads; R.ref_num = (long)ref_num; R.emitted_bases = len_total;
        // record offsets: ">" id_prefix ref_num "_" n "Z" seq "\n"
        R.h_rec_off.assign((size_t)n_reads + 1, 0);
        max_len = 2;
        for (int64_t i = 0; i < n_reads; i++) {
            const int64_t seqlen = p.no_error ? (int64_t)R.h_len[i] + R.h_ndel[i] : R.h_len[i];
            R.h_rec_off[i + 1] = R.h_rec_off[i] + (int64_t)id.size() + count_digits(i + 1) + 1 + seqlen + 1;
            max_len = std::max(max_len, R.h_len[i]);
        }
        R.out_bytes = R.h_rec_off[n_reads];
        const size_t n1 = (size_t)std::max<int64_t>(n_reads, 1);
        R.d_len.reserve(n1); R.d_ndel.reserve(n1); R.d_acc.reserve(n1); R.d_rec_off.reserve(n1 + 1);
        R.d_dropped.reserve(n1); R.d_used.reserve(n1);
        R.d_ctr.reserve(16);
        R.d_out.reserve((size_t)R.out_bytes + 64, 0.05);
        if (n_reads > 0) {
            SP_CUDA(cudaMemcpyAsync(R.d_len.p, R.h_len.data(), n_reads * 4, cudaMemcpyHostToDevice, st));
            SP_CUDA(cudaMemcpyAsync(R.d_ndel.p, R.h_ndel.data(), n_reads * 4, cudaMemcpyHostToDevice, st));
            SP_CUDA(cudaMemcpyAsync(R.d_acc.p, R.h_acc.data(), n_reads, cudaMemcpyHostToDevice, st));
        }
        SP_CUDA(cudaMemcpyAsync(R.d_rec_off.p, R.h_rec_off.data(), (n_reads + 1) * 8, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemsetAsync(R.d_ctr.p, 0, 64, st));
        R.n_dropped = 0;
        if (n_reads == 0) return false;

        // group the reads so that the 32 threads of a warp run nearly the same trip counts: longest first (tail balance),
        // then by accuracy class (same tables, similar deletion loops). Counting sort, O(n).
        {
            const int32_t width = std::max(1, (max_len + 63) / 64);
            std::vector<int32_t> count(64 * 101 + 1, 0);
            auto key = [&](int64_t i) { return (63 - std::min(63, R.h_len[i] / width)) * 101 + (int)R.h_acc[i]; };
            for (int64_t i = 0; i < n_reads; i++) count[key(i) + 1]++;
            for (size_t k = 1; k < count.size(); k++) count[k] += count[k - 1];
            R.h_order.resize((size_t)n_reads);
            for (int64_t i = 0; i < n_reads; i++) R.h_order[(size_t)count[key(i)]++] = (int32_t)i;
            R.d_order.reserve(n1);
            SP_CUDA(cudaMemcpyAsync(R.d_order.p, R.h_order.data(), n_reads * 4, cudaMemcpyHostToDevice, st));
        }
    }
    R.cur_seq = d_seq; R.cur_L = L; R.cur_max_len = max_len; R.cur_replay = replay;
    R.plan_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_plan0).count();
    return true;
}

static void launch_batch(sp_ctx* c, ReadState& R, cudaStream_t ks) {
    const ReadParams& p = R.par;
    const bool replay = R.cur_replay;
    const int64_t n_reads = R.n_reads, ref_num = R.ref_num;
    const std::string id = ">" + p.id_prefix + std::to_string(ref_num) + "_";
    int ctas_per_sm = 1;
    {   // Shared memory against L1: the generator state columns are 17 KB per CTA, but the table walks, the template
        // chunks and the scratch words live in L1. Measured at 30x chr1: the driver's default split (5 CTAs per SM, large
        // L1) 88 ms, the largest carve-out (8 CTAs per SM) 119 ms. SYNTHPOP_READS_CARVEOUT (percent) is a tuning hook.
        if (const char* e = getenv("SYNTHPOP_READS_CARVEOUT"))   // the attribute is per device: set on every launch (cheap)
            SP_CUDA(cudaFuncSetAttribute(pick_read_kernel(replay, p.no_error), cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e)));
    }
    const ReadKernel kernel = pick_read_kernel(replay, p.no_error);
    SP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kernel, RD_THREADS, 0));
    ctas_per_sm = std::max(1, ctas_per_sm);
    if (const char* e = getenv("SYNTHPOP_READS_CTAS_PER_SM")) ctas_per_sm = std::max(1, std::min(ctas_per_sm, atoi(e)));  // tuning hook
    const int64_t want_ctas = (n_reads + RD_THREADS - 1) / RD_THREADS;
    const int64_t grid_full = std::min<int64_t>(want_ctas, (int64_t)c->sm_count * ctas_per_sm);
    const int64_t words = ((int64_t)R.cur_max_len + 3) / 4 + 1;
    const size_t cta_scratch = (size_t)(RD_THREADS / 32) * 2 * words * 32 * sizeof(uint32_t);
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(grid_full, (int64_t)((16ull << 30) / cta_scratch)));  // <= 16 GiB of scratch
    R.d_scratch.reserve(cta_scratch * grid + 4096, 0.0);   // slack: row prefetches run two rows past a lane's last row
    if (!R.h_ctr) SP_CUDA(cudaHostAlloc((void**)&R.h_ctr, 8 * sizeof(unsigned long long), cudaHostAllocDefault));

    ReadJob J{};
    J.seq = R.cur_seq; J.L = R.cur_L; J.n_reads = n_reads; J.order = R.d_order.p;
    J.len = R.d_len.p; J.acc = R.d_acc.p; J.ndel = R.d_ndel.p; J.rec_off = R.d_rec_off.p;
    J.seed = p.seed; J.ref_num = ref_num; J.no_error = p.no_error ? 1 : 0;
    J.replay = replay ? R.d_replay.p : nullptr;
    J.replay_off = replay ? R.d_replay_off.p : nullptr;
    J.out = R.d_out.p; J.dropped = R.d_dropped.p; J.used = R.d_used.p;
    J.scratch = (uint32_t*)R.d_scratch.p; J.scratch_words = words;
    J.ctr = (unsigned long long*)R.d_ctr.p;
    if (id.size() >= sizeof J.id) throw Error(SP_ERR_ARG, "read id prefix too long");
    memcpy(J.id, id.data(), id.size());
    J.id_len = (int32_t)id.size();
    if (!R.ev0) { SP_CUDA(cudaEventCreate(&R.ev0)); SP_CUDA(cudaEventCreate(&R.ev1)); }
    SP_CUDA(cudaEventRecord(R.ev0, ks));
    kernel<<<grid, RD_THREADS, 0, ks>>>(J, R.dev);
    c->launches++;
    SP_CUDA(cudaGetLastError());
    SP_CUDA(cudaEventRecord(R.ev1, ks));
    SP_CUDA(cudaMemcpyAsync(R.h_ctr, R.d_ctr.p, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ks));   // pinned: does not wait
}

static void collect_batch(sp_ctx* c, ReadState& R, cudaStream_t ks) {
    SP_CUDA(cudaStreamSynchronize(ks));
    SP_CUDA(cudaEventElapsedTime(&R.kernel_ms, R.ev0, R.ev1));
    const unsigned long long* ctr = R.h_ctr;
    if (R.cur_replay) R.have_replay = false;
    if (ctr[1] != 0) {
        static const char* what[] = {"", "a read of length 1 (the reference computes rand() % 0)",
                                     "a read at least as long as the sequence it is drawn from (unsupported: the reference truncates it)",
                                     "more than 255 deletions at one position (unsupported)"};
        throw Error(SP_ERR_ARG, std::string("read kernel: ") + what[std::min<unsigned long long>(ctr[1], 3)] + ", read " + std::to_string(ctr[2] + 1));
    }
    R.n_dropped = (int64_t)ctr[3];
    R.out_bytes -= (int64_t)ctr[4];
}

static void simulate_on_device(sp_ctx* c, const uint8_t* d_seq, int64_t L, int64_t ref_num) {
    ReadState& R = reads_of(c);
    if (!plan_batch(c, R, d_seq, L, ref_num)) return;
    launch_batch(c, R, c->stream);
    collect_batch(c, R, c->stream);
}

// The same configuration on a second set of batch buffers. The device tables are shared (read-only); the host tables are
// copied because the plan reads them.
static ReadState& second_state(sp_ctx* c) {
    ReadState& A = reads_of(c);
    if (!c->reads_b) c->reads_b = new ReadState();
    ReadState& B = *c->reads_b;
    if (B.config_epoch != A.config_epoch) {
        B.par = A.par; B.tab = A.tab; B.dev = A.dev;
        B.have_replay = false; B.n_reads = 0; B.plan_tables_stale = true; B.configured = true;
        B.config_epoch = A.config_epoch;
    }
    if (!B.kernel_stream) SP_CUDA(cudaStreamCreateWithFlags(&B.kernel_stream, cudaStreamNonBlocking));
    if (!B.ev_ready) SP_CUDA(cudaEventCreateWithFlags(&B.ev_ready, cudaEventDisableTiming));
    return B;
}

// docker/haplotype2reads.sh:53-58 sequences haplotype ID1, then ID2, then concatenates. Here both read kernels are in flight:
// a launch is one persistent wave of CTAs that fetch reads until none is left, so its last third runs with fewer and fewer
// warps (35 of 40 resident on average, profiles/r02a_reads_kernel_base_raw.csv); the second haplotype's CTAs take the slots
// the first one's give up. Reads are independent of each other (per-read streams), so every record is what the one-by-one
// calls produce. Both batches become sources of the mix, in order.
static void simulate_pair_into_mix(sp_ctx* c, const uint8_t* seq0, int64_t L0, int64_t ref_num0, const uint8_t* seq1, int64_t L1,
                                   int64_t ref_num1, const uint32_t* seeds /*[2] or null: the configured seed for both*/, int64_t* n_reads,
                                   int64_t* file_bytes, int64_t* n_dropped, int64_t* bases, double* span_ms) {
    ReadState& A = reads_of(c);
    if (A.par.stream_mode == SP_READS_REPLAY) throw Error(SP_ERR_ARG, "sp_reads_mix_add_pair: per-read streams only (a replay log belongs to one sequence)");
    ReadState& B = second_state(c);
    struct SeedGuard {   // the configured seed comes back whatever happens below
        ReadState &A, &B; uint32_t seed;
        ~SeedGuard() { A.par.seed = seed; B.par.seed = seed; }
    } guard{A, B, A.par.seed};
    if (seeds) { A.par.seed = seeds[0]; B.par.seed = seeds[1]; }
    const bool a = plan_batch(c, A, seq0, L0, ref_num0);
    const bool b = plan_batch(c, B, seq1, L1, ref_num1);
    SP_CUDA(cudaEventRecord(B.ev_ready, c->stream));              // both plans (and whatever made the sequences) are done
    SP_CUDA(cudaStreamWaitEvent(B.kernel_stream, B.ev_ready, 0));
    if (a) launch_batch(c, A, c->stream);
    if (b) launch_batch(c, B, B.kernel_stream);
    try {
        if (a) collect_batch(c, A, c->stream);
    } catch (...) {
        cudaStreamSynchronize(B.kernel_stream);   // nothing of this call is still running when the error leaves
        throw;
    }
    if (b) collect_batch(c, B, B.kernel_stream);
    A.pair_span_ms = 0.f;
    if (a && b) SP_CUDA(cudaEventElapsedTime(&A.pair_span_ms, A.ev0, B.ev1));
    else if (a) A.pair_span_ms = A.kernel_ms;
    else if (b) A.pair_span_ms = B.kernel_ms;
    ReadState* r[2] = {&A, &B};
    for (int k = 0; k < 2; k++) {
        if (n_reads) n_reads[k] = r[k]->n_reads;
        if (file_bytes) file_bytes[k] = r[k]->out_bytes;
        if (n_dropped) n_dropped[k] = r[k]->n_dropped;
        if (bases) bases[k] = r[k]->emitted_bases;
    }
    if (span_ms) *span_ms = A.pair_span_ms;
    mix_add_from(c, A);
    mix_add_from(c, B);
}

static void fetch_read_log(sp_ctx* c, ReadState& R) {
    if (R.results_on_host) return;
    R.h_dropped.assign((size_t)R.n_reads, 0);
    R.h_used.assign((size_t)R.n_reads, 0);
    if (R.n_reads > 0) {
        SP_CUDA(cudaMemcpyAsync(R.h_dropped.data(), R.d_dropped.p, (size_t)R.n_reads, cudaMemcpyDeviceToHost, c->stream));
        SP_CUDA(cudaMemcpyAsync(R.h_used.data(), R.d_used.p, (size_t)R.n_reads * 4, cudaMemcpyDeviceToHost, c->stream));
        SP_CUDA(cudaStreamSynchronize(c->stream));
    }
    R.results_on_host = true;
}

// file bytes of the last batch -> host (dropped records squeezed out on the device first)
static int64_t copy_reads(sp_ctx* c, uint8_t* host_out, int64_t capacity) {
    ReadState& R = reads_of(c);
    if (R.out_bytes > capacity) throw Error(SP_ERR_ARG, "host buffer too small");
    if (R.n_reads == 0 || R.out_bytes == 0) return 0;
    const uint8_t* src = R.d_out.p;
    if (R.n_dropped > 0) {
        fetch_read_log(c, R);
        ensure_host_plan(c, R);
        std::vector<int64_t> koff((size_t)R.n_reads + 1, 0);
        for (int64_t r = 0; r < R.n_reads; r++)
            koff[r + 1] = koff[r] + (R.h_dropped[r] ? 0 : R.h_rec_off[r + 1] - R.h_rec_off[r]);
        R.d_koff.reserve((size_t)R.n_reads + 1);
        R.d_compact.reserve((size_t)R.out_bytes + 64, 0.05);
        SP_CUDA(cudaMemcpyAsync(R.d_koff.p, koff.data(), koff.size() * 8, cudaMemcpyHostToDevice, c->stream));
        compact_records_kernel<<<c->sm_count * 8, 256, 0, c->stream>>>(R.d_out.p, R.d_rec_off.p, R.d_koff.p, R.n_reads, R.d_compact.p);
        c->launches++;
        SP_CUDA(cudaGetLastError());
        src = R.d_compact.p;
    }
    SP_CUDA(cudaMemcpyAsync(host_out, src, (size_t)R.out_bytes, cudaMemcpyDeviceToHost, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
    return R.out_bytes;
}

// get_ref_inf / get_ref_seq (:618-750): a line that starts with '>' opens a sequence, every other line is appended.
static void read_fasta_sequences(const char* path, std::vector<std::vector<uint8_t>>& seqs) {
    FILE* fp = fopen(path, "rb");
    if (!fp) throw Error(SP_ERR, std::string("ERROR: Cannot open file: ") + path);
    std::vector<char> buf(1 << 22);
    bool at_line_start = true, in_header = false, have = false;
    seqs.clear();
    size_t n;
    while ((n = fread(buf.data(), 1, buf.size(), fp)) > 0) {
        for (size_t i = 0; i < n; i++) {
            const char ch = buf[i];
            if (at_line_start && ch == '>') {
                seqs.emplace_back(); have = true; in_header = true; at_line_start = false;
                continue;
            }
            if (ch == '\n') { in_header = false; at_line_start = true; continue; }
            at_line_start = false;
            if (in_header) continue;
            if (!have) { seqs.emplace_back(); have = true; }
            seqs.back().push_back((uint8_t)ch);
        }
    }
    fclose(fp);
}

}  // namespace sp

using sp::Error;


static void report(sp::ReadState& R, int64_t* n_reads, int64_t* file_bytes, int64_t* n_dropped, int64_t* bases) {
    if (n_reads) *n_reads = R.n_reads;
    if (file_bytes) *file_bytes = R.out_bytes;
    if (n_dropped) *n_dropped = R.n_dropped;
    if (bases) *bases = R.emitted_bases;
}

extern "C" {

int sp_reads_configure(sp_ctx* ctx, const sp_read_args* args) {
    SP_API_BEGIN
    sp::reads_configure(ctx, args);
    SP_API_END
}

int sp_reads_set_replay(sp_ctx* ctx, const int32_t* draws, int64_t n_draws, const int64_t* read_offsets, int64_t n_reads) {
    SP_API_BEGIN
    sp::ReadState& R = sp::reads_of(ctx);
    if (!draws || !read_offsets || n_draws < 2 || n_reads < 1) throw Error(SP_ERR_ARG, "bad replay log");
    R.replay_draws.assign(draws, draws + n_draws);
    R.replay_off.assign(read_offsets, read_offsets + n_reads);
    const size_t pad = 64;
    R.d_replay.reserve((size_t)n_draws + pad);
    R.d_replay_off.reserve((size_t)n_reads);
    SP_CUDA(cudaMemsetAsync(R.d_replay.p + n_draws, 0, pad * 4, ctx->stream));
    SP_CUDA(cudaMemcpyAsync(R.d_replay.p, draws, (size_t)n_draws * 4, cudaMemcpyHostToDevice, ctx->stream));
    SP_CUDA(cudaMemcpyAsync(R.d_replay_off.p, read_offsets, (size_t)n_reads * 8, cudaMemcpyHostToDevice, ctx->stream));
    SP_CUDA(cudaStreamSynchronize(ctx->stream));
    R.have_replay = true;
    SP_API_END
}

int sp_reads_simulate_haplotype(sp_ctx* ctx, int32_t h, int64_t ref_num, int64_t* n_reads, int64_t* file_bytes,
                                int64_t* n_dropped, int64_t* bases) {
    SP_API_BEGIN
    void* p = nullptr;
    int64_t L = 0;
    if (!ctx->materialised) throw Error(SP_ERR_ARG, "sp_materialise has not been called");
    const int rc = sp_haplotype_device(ctx, h, &p, &L);
    if (rc != SP_OK) return rc;
    sp::simulate_on_device(ctx, (const uint8_t*)p, L, ref_num);
    report(sp::reads_of(ctx), n_reads, file_bytes, n_dropped, bases);
    SP_API_END
}

int sp_reads_mix_add_pair(sp_ctx* ctx, int32_t h0, int32_t h1, int64_t ref_num0, int64_t ref_num1, int32_t seed_per_haplotype,
                          int64_t* n_reads, int64_t* file_bytes, int64_t* n_dropped, int64_t* bases, double* kernels_span_ms) {
    SP_API_BEGIN
    void *p0 = nullptr, *p1 = nullptr;
    int64_t L0 = 0, L1 = 0;
    if (!ctx->materialised) throw Error(SP_ERR_ARG, "sp_materialise has not been called");
    int rc = sp_haplotype_device(ctx, h0, &p0, &L0);
    if (rc != SP_OK) return rc;
    rc = sp_haplotype_device(ctx, h1, &p1, &L1);
    if (rc != SP_OK) return rc;
    uint32_t seeds[2] = {0, 0};
    if (seed_per_haplotype) {
        const uint32_t seed0 = sp::reads_of(ctx).par.seed;
        seeds[0] = sp_haplotype_read_seed(seed0, h0); seeds[1] = sp_haplotype_read_seed(seed0, h1);
    }
    sp::simulate_pair_into_mix(ctx, (const uint8_t*)p0, L0, ref_num0, (const uint8_t*)p1, L1, ref_num1, seed_per_haplotype ? seeds : nullptr,
                               n_reads, file_bytes, n_dropped, bases, kernels_span_ms);
    SP_API_END
}

int sp_reads_simulate_device(sp_ctx* ctx, const void* device_seq, int64_t seq_len, int64_t ref_num, int64_t* n_reads,
                             int64_t* file_bytes, int64_t* n_dropped, int64_t* bases) {
    SP_API_BEGIN
    if (!device_seq) throw Error(SP_ERR_ARG, "null device sequence");
    sp::simulate_on_device(ctx, (const uint8_t*)device_seq, seq_len, ref_num);
    report(sp::reads_of(ctx), n_reads, file_bytes, n_dropped, bases);
    SP_API_END
}

int sp_reads_simulate_sequence(sp_ctx* ctx, const uint8_t* seq, int64_t seq_len, int64_t ref_num, int64_t* n_reads,
                               int64_t* file_bytes, int64_t* n_dropped, int64_t* bases) {
    SP_API_BEGIN
    sp::ReadState& R = sp::reads_of(ctx);
    if (!seq || seq_len <= 0) throw Error(SP_ERR_ARG, "bad sequence buffer");
    R.d_seq.reserve((size_t)seq_len);
    SP_CUDA(cudaMemcpyAsync(R.d_seq.p, seq, (size_t)seq_len, cudaMemcpyHostToDevice, ctx->stream));
    sp::simulate_on_device(ctx, R.d_seq.p, seq_len, ref_num);
    report(R, n_reads, file_bytes, n_dropped, bases);
    SP_API_END
}

int sp_reads_copy(sp_ctx* ctx, uint8_t* host_out, int64_t capacity, int64_t* written) {
    SP_API_BEGIN
    const int64_t n = sp::copy_reads(ctx, host_out, capacity);
    if (written) *written = n;
    SP_API_END
}

int sp_reads_log(sp_ctx* ctx, int32_t* length, uint8_t* dropped, uint32_t* draws_used) {
    SP_API_BEGIN
    sp::ReadState& R = sp::reads_of(ctx);
    sp::fetch_read_log(ctx, R);
    sp::ensure_host_plan(ctx, R);
    const size_t n = (size_t)R.n_reads;
    if (length && n) memcpy(length, R.h_len.data(), n * 4);
    if (dropped && n) memcpy(dropped, R.h_dropped.data(), n);
    if (draws_used && n) memcpy(draws_used, R.h_used.data(), n * 4);
    SP_API_END
}

int sp_reads_timing(sp_ctx* ctx, double* plan_ms, double* kernel_ms) {
    SP_API_BEGIN
    sp::ReadState& R = sp::reads_of(ctx);
    if (plan_ms) *plan_ms = R.plan_ms;
    if (kernel_ms) *kernel_ms = R.kernel_ms;
    SP_API_END
}

int sp_reads_device(sp_ctx* ctx, void** records, void** record_offsets, void** dropped_flags) {
    SP_API_BEGIN
    sp::ReadState& R = sp::reads_of(ctx);
    if (records) *records = R.d_out.p;
    if (record_offsets) *record_offsets = R.d_rec_off.p;
    if (dropped_flags) *dropped_flags = R.d_dropped.p;
    SP_API_END
}

int sp_simulate_reads(sp_ctx* ctx, const sp_read_args* args, const char* fasta_path) {
    SP_API_BEGIN
    if (!fasta_path) throw Error(SP_ERR_ARG, "null FASTA path");
    sp::reads_configure(ctx, args);
    sp::ReadState& R = sp::reads_of(ctx);
    if (R.par.stream_mode == SP_READS_REPLAY) throw Error(SP_ERR_ARG, "sp_simulate_reads runs in per-read-stream mode; replay goes through sp_reads_set_replay");
    std::vector<std::vector<uint8_t>> seqs;
    sp::read_fasta_sequences(fasta_path, seqs);
    if (seqs.size() > 9999) throw Error(SP_ERR_ARG, "ERROR: References are too many. Max number of reference is 9999.");
    for (auto& s : seqs)
        if (s.size() < 100) throw Error(SP_ERR_ARG, "ERROR: Reference is too short. Acceptable length >= 100.");
    std::vector<uint8_t> host;
    for (size_t k = 0; k < seqs.size(); k++) {
        R.d_seq.reserve(seqs[k].size());
        SP_CUDA(cudaMemcpyAsync(R.d_seq.p, seqs[k].data(), seqs[k].size(), cudaMemcpyHostToDevice, ctx->stream));
        sp::simulate_on_device(ctx, R.d_seq.p, (int64_t)seqs[k].size(), (int64_t)k + 1);
        host.resize((size_t)R.out_bytes);
        sp::copy_reads(ctx, host.data(), (int64_t)host.size());
        char name[4096];
        snprintf(name, sizeof name, "%s_%04ld.fastq", R.par.prefix.c_str(), (long)k + 1);  // main :552
        FILE* f = fopen(name, "wb");
        if (!f) throw Error(SP_ERR, std::string("ERROR: Cannot open output file: ") + name);
        const size_t w = fwrite(host.data(), 1, host.size(), f);
        fclose(f);
        if (w != host.size()) throw Error(SP_ERR, std::string("ERROR: short write: ") + name);
    }
    SP_API_END
}

}  // extern "C"
