// Row P of SURVEY.md §8a — the per-read code of the read kernel (docker/pbsim2.cpp simulate_by_model :1641-1711, mutate
// :2260-2413), written so that it compiles for the device (csrc/reads.cu) AND for the host (tests/emu/reads_emu.cu runs the
// very same functions one lane at a time on the CPU and compares them with the restatement — a logic check that needs no
// GPU; it is test infrastructure, the library never runs this code on the host).
//
// Second design of the per-read code ("one draw per step"): the glibc generator r[i] = r[i-31] + r[i-3] lives in a
// shared-memory column per thread, and what made the first design expensive was the RING POSITION of a draw being a
// run-time value in two of the three phases (12 instructions per draw, 5 with a position the compiler knows). Here the
// ring is physically rotated to position 0 before the deletion loop and again before the pass over the bases (31 loads +
// 31 stores, twice per read), and both loops are unrolled over 31 ring positions:
//   * deletion loop: an attempt is two draws, 31 attempts are two laps of the ring;
//   * pass over the bases: every lane of a warp makes exactly ONE draw per step, at the same ring position. What the draw
//     means is the lane's own business: the error decision of its next base (the common case), or — after an error event —
//     the substitution / insertion index the reference draws next (:2362-2387). A lane therefore falls behind the step
//     counter by one base per extra draw, keeps its own base index, and loads its quality word / stores its output word
//     under its own predicate. Nothing is ever drawn out of order: lane by lane the sequence of rand() calls is the
//     reference's.
// Output words are aligned to the RECORD's position in the file (the accumulator is a shift register, a full aligned
// 8-byte word is stored as it is), so there is no funnel shift per word and no second accumulator.
#pragma once
#include "reads.h"

#if defined(__CUDACC__)
#define SP_HD __host__ __device__ __forceinline__
#define SP_HD_NOINLINE __host__ __device__ __noinline__
#else
#define SP_HD inline
#define SP_HD_NOINLINE inline
#endif

namespace sp {

constexpr int RD_THREADS = 128;  // reads in flight per CTA
constexpr unsigned FULL = 0xffffffffu;
constexpr int LUT_REV_OFFSET = 288;  // bytes between the two strand tables (a shift of 8 banks: 'A' on one strand and
                                     // 'A' on the other fall into different banks)

// ---- the few intrinsics the per-read code needs, with a host meaning (one lane = one warp of its own) -----------------
template <class T>
SP_HD T ld_ro(const T* p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
SP_HD uint4 ld_ro4(const void* p) {
#ifdef __CUDA_ARCH__
    return __ldg((const uint4*)p);
#else
    return *(const uint4*)p;
#endif
}
SP_HD uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) {
#ifdef __CUDA_ARCH__
    return __byte_perm(a, b, s);
#else
    const uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((s >> (4 * i)) & 7u))) & 0xffu) << (8 * i);
    return r;
#endif
}
SP_HD uint32_t vminu4(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    return __vminu4(a, b);
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) {
        const uint32_t x = (a >> (8 * i)) & 255u, y = (b >> (8 * i)) & 255u;
        r |= (x < y ? x : y) << (8 * i);
    }
    return r;
#endif
}
SP_HD bool warp_any(unsigned mask, bool p) {
#ifdef __CUDA_ARCH__
    return __any_sync(mask, p);
#else
    (void)mask;
    return p;
#endif
}
SP_HD void warp_sync(unsigned mask) {
#ifdef __CUDA_ARCH__
    __syncwarp(mask);
#else
    (void)mask;
#endif
}
SP_HD void red_add_u32(uint32_t* p, uint32_t v) {  // fire-and-forget: nothing waits for the old value
#ifdef __CUDA_ARCH__
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#else
    *p += v;
#endif
}
SP_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {  // low word of (hi:lo) >> sh, 0 <= sh < 32
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, sh);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> (sh & 31u));
#endif
}
// The value as the compiler cannot re-derive it: a cheap-to-recompute value (a shared-window base, a strand sign) that it
// would otherwise rebuild from scratch at every use to save a register, four instructions per step.
SP_HD uint32_t keep_in_register(uint32_t v) {
#ifdef __CUDA_ARCH__
    asm volatile("mov.b32 %0, %0;" : "+r"(v));
#endif
    return v;
}
SP_HD int clz32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}

// Seed of read `read_index` (0-based) of sequence `ref_num` in per-read-stream mode: splitmix64 finaliser -> [1, 2^31-2].
SP_HD uint32_t read_seed(uint32_t seed, int64_t ref_num, int64_t read_index) {
    uint64_t z = (uint64_t)seed * 0x9E3779B97F4A7C15ULL + (uint64_t)ref_num * 0xBF58476D1CE4E5B9ULL +
                 (uint64_t)read_index * 0x94D049BB133111EBULL;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27; z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return (uint32_t)(1 + z % 2147483646ULL);
}

SP_HD uint32_t fastmod(uint32_t a, uint32_t M, uint32_t sh, uint32_t d) {
    const uint32_t q = (uint32_t)(((uint64_t)a * M) >> sh);
    return a - q * d;
}
SP_HD uint32_t fastmod(uint32_t a, const FastMod* p) {
    const uint4 v = ld_ro4(p);
    return fastmod(a, v.x, v.y, v.z);
}
SP_HD FastMod make_fastmod_dev(uint32_t d) {
    FastMod f;
    const uint32_t l = d <= 1 ? 0 : 32 - clz32(d - 1);
    f.sh = 31 + l;
    f.M = (uint32_t)(((1ull << f.sh) + d - 1) / d);
    f.d = d; f.pad = 0;
    return f;
}

// glibc srand(s), 1 <= s <= 2^31-2 (random_r.c TYPE_3: degree 31, separation 3): r[i] = 16807 * r[i-1] mod (2^31-1), then
// 310 outputs of r[f] += r[b] (f = 3, b = 0 at the start) are discarded; afterwards f is 3 again.
SP_HD void glibc_seed(uint32_t s, uint32_t (&r)[31]) {
    uint64_t w = s;
    r[0] = s;
#pragma unroll
    for (int i = 1; i < 31; i++) { w = (w * 16807ull) % 2147483647ull; r[i] = (uint32_t)w; }
#pragma unroll 1
    for (int round = 0; round < 10; round++) {
#pragma unroll
        for (int j = 0; j < 31; j++) r[(j + 3) % 31] += r[j];
    }
}

struct ReadJob {
    const uint8_t* seq;  // the sequence reads are drawn from (a haplotype resident in HBM)
    int64_t L;
    int64_t n_reads;
    const int32_t* order; // read indices grouped by (accuracy class, length): 32 consecutive entries share a warp
    const int32_t* len;   // emitted length per read (:1642-1650, quota clipping applied)
    const uint8_t* acc;   // accuracy class (:1654-1655)
    const int32_t* ndel;  // del_num_total (:2282)
    const int64_t* rec_off;  // n_reads+1 record offsets in `out`
    uint32_t seed;
    int64_t ref_num;
    int32_t no_error;
    const int32_t* replay;      // nullptr in per-read-stream mode
    const int64_t* replay_off;
    uint8_t* out;
    uint8_t* dropped;
    uint32_t* used;
    uint32_t* scratch;          // per warp: quality codes and deletion counters, one byte per position each (layout: see the kernels)
    int64_t scratch_words;      // words per lane per array (4 positions each)
    unsigned long long* ctr;    // [0] next read, [1] error code, [2] error read, [3] dropped reads, [4] dropped bytes
    char id[64];                // ">" + id_prefix + ref_num + "_"
    int32_t id_len;
};

SP_HD void job_error(const ReadJob& J, int code, int64_t r) {
#ifdef __CUDA_ARCH__
    if (atomicCAS(&J.ctr[1], 0ull, (unsigned long long)code) == 0ull) J.ctr[2] = (unsigned long long)r;
#else
    if (J.ctr[1] == 0ull) { J.ctr[1] = (unsigned long long)code; J.ctr[2] = (unsigned long long)r; }
#endif
}
SP_HD void job_count_dropped(const ReadJob& J, unsigned long long bytes) {
#ifdef __CUDA_ARCH__
    atomicAdd(&J.ctr[3], 1ull);
    atomicAdd(&J.ctr[4], bytes);
#else
    J.ctr[3] += 1ull; J.ctr[4] += bytes;
#endif
}

// Per-CTA tables in shared memory: err_thre, sub_thre, ins_thre (3 x PB_NQ, :1455-1459), max(err_thre[q..93]) (PB_NQ), then
// the two strand tables (byte -> upper-cased base, byte -> upper-cased complement; non-ACGT bytes keep their upper-cased
// value on both strands, :2316-2341). Thread `tid` of `nthr` fills its share; part 1 needs part 0 complete.
constexpr int RD_TABLE_WORDS = 4 * PB_NQ + (LUT_REV_OFFSET + 256) / 4;
SP_HD void fill_cta_tables(int32_t* s_thre, const int32_t* thre, int tid, int nthr, int part) {
    if (part == 0) {
        for (int i = tid; i < 3 * PB_NQ; i += nthr) s_thre[i] = thre[i];
        uint8_t* lut = (uint8_t*)(s_thre + 4 * PB_NQ);
        for (int i = tid; i < 256; i += nthr) {
            uint32_t ch = (uint32_t)i;
            ch -= (ch - 'a' <= (uint32_t)('z' - 'a')) ? 32u : 0u;
            lut[i] = (uint8_t)ch;
            lut[LUT_REV_OFFSET + i] = (uint8_t)(ch == 'A' ? 'T' : ch == 'T' ? 'A' : ch == 'C' ? 'G' : ch == 'G' ? 'C' : ch);
        }
    } else if (tid == 0) {
        int32_t m = 0;
        for (int q = PB_NQ - 1; q >= 0; q--) { m = m > s_thre[q] ? m : s_thre[q]; s_thre[3 * PB_NQ + q] = m; }
    }
}

// The read's rand(): the thread's own glibc generator (state column: word j of this thread is st[j * RD_THREADS]) or the
// replayed log.
template <bool REPLAY>
struct Draws {
    uint32_t* st;
    const int32_t* replay;  // first logged draw of this read
    uint32_t used;          // rand() calls so far
    int f;                  // glibc's front index (rear = f - 3 mod 31)
    SP_HD void seed(uint32_t s) {
        uint32_t r[31];
        glibc_seed(s, r);
#pragma unroll
        for (int j = 0; j < 31; j++) st[j * RD_THREADS] = r[j];
        f = 3; used = 0;
    }
    // rand() at a ring position the caller knows at compile time: no index arithmetic; f and used are brought up to date by
    // the caller once the unrolled block is over
    SP_HD uint32_t next_at(int fi) {
        if (REPLAY) return (uint32_t)ld_ro(replay + used++);
        const int b = fi >= 3 ? fi - 3 : fi + 28;
        const uint32_t v = st[fi * RD_THREADS] + st[b * RD_THREADS];
        st[fi * RD_THREADS] = v;
        return v >> 1;
    }
    SP_HD void skip(uint32_t n) {  // account for n draws made through next_at
        if (REPLAY) return;
        used += n;
        f = (int)((uint32_t)(f + n) % 31u);
    }
    SP_HD uint32_t next() {
        if (REPLAY) return (uint32_t)ld_ro(replay + used++);
        const int b = f >= 3 ? f - 3 : f + 28;
        const uint32_t v = st[f * RD_THREADS] + st[b * RD_THREADS];
        st[f * RD_THREADS] = v;
        f = f == 30 ? 0 : f + 1;
        used++;
        return v >> 1;
    }
    // Physical rotation of the ring: word (j + f) mod 31 moves to word j, afterwards the front index is 0. 31 is prime, so
    // for f != 0 the moves form one cycle 0 <- f <- 2f <- ... ; one temporary.
    SP_HD void rotate() {
        if (REPLAY) return;
        const int s = f;
        if (s != 0) {
            const uint32_t t0 = st[0];
            int j = 0;
#pragma unroll 1
            for (int i = 0; i < 30; i++) {
                int nj = j + s;
                nj = nj >= 31 ? nj - 31 : nj;
                st[j * RD_THREADS] = st[nj * RD_THREADS];
                j = nj;
            }
            st[j * RD_THREADS] = t0;
        }
        f = 0;
    }
};

// The CTA's tables as the per-read code sees them: a shared-window address on the device (32 bits, loads with an
// immediate offset), a plain pointer on the host.
struct CtaTables {
#ifdef __CUDA_ARCH__
    uint32_t a;
    __device__ __forceinline__ void bind(const int32_t* s) { a = keep_in_register((uint32_t)__cvta_generic_to_shared(s)); }
    __device__ __forceinline__ int32_t w(uint32_t byte_off) const {
        int32_t v;
        asm("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a + byte_off));
        return v;
    }
    __device__ __forceinline__ uint32_t b(uint32_t byte_off) const {
        uint32_t v;
        asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a + byte_off));
        return v;
    }
#else
    const uint8_t* a;
    void bind(const int32_t* s) { a = (const uint8_t*)s; }
    int32_t w(uint32_t byte_off) const { return *(const int32_t*)(a + byte_off); }
    uint32_t b(uint32_t byte_off) const { return a[byte_off]; }
#endif
    SP_HD int32_t err_thre(uint32_t q) const { return w(4 * q); }
    SP_HD int32_t sub_thre(uint32_t q) const { return w(4 * q + 4 * PB_NQ); }
    SP_HD int32_t ins_thre(uint32_t q) const { return w(4 * q + 8 * PB_NQ); }
    SP_HD int32_t err_thre_max_from(uint32_t q) const { return w(4 * q + 12 * PB_NQ); }
    SP_HD uint32_t base_of(uint32_t strand_off, uint32_t byte) const { return b(16 * PB_NQ + strand_off + byte); }
};

// The oriented, upper-cased template (:2316-2341): seq[t] for t in [0, len'), fetched through aligned 8-byte chunks.
// The strand is data (a step of +-1 and a table offset), not a branch: the threads of a warp alternate strands.
struct TemplateWindow {
    const unsigned long long* base8;  // 8-byte aligned address at or below seq[0]
    int32_t o0;                       // seq[0] - base8, 0..7
    int32_t dir;                      // +1 / -1
    uint32_t strand_off;              // 0 / LUT_REV_OFFSET
    int32_t cur;                      // index (relative to base8) of the cached 8-byte chunk
    uint32_t lo, hi;
    SP_HD void begin(const uint8_t* w, int32_t lp, bool forward) {
        const uint8_t* origin = forward ? w : w + (lp - 1);  // first base of the window (+) or its last base (-)
        o0 = (int32_t)((uintptr_t)origin & 7);
        base8 = (const unsigned long long*)(origin - o0);
        dir = (int32_t)keep_in_register(forward ? 1u : 0xffffffffu);
        strand_off = forward ? 0u : (uint32_t)LUT_REV_OFFSET; cur = INT32_MIN; lo = hi = 0;
    }
    SP_HD uint32_t at(int32_t t, const CtaTables& C) {
        const int32_t rel = o0 + dir * t;   // 32-bit: a read is far shorter than 2^31 bases
        const int32_t w = rel >> 3;         // floor, also for the negative offsets of the reverse strand
        if (w != cur) {
            const unsigned long long chunk = ld_ro(base8 + w);
            lo = (uint32_t)chunk; hi = (uint32_t)(chunk >> 32); cur = w;
        }
        return C.base_of(strand_off, byte_perm(lo, hi, (uint32_t)(rel & 7)) & 0xffu);
    }
};

// bytes [b0, b1) of the 8-byte value (hi:lo) to q[b0..b1): the ends of a record, once or twice per read, out of line
SP_HD_NOINLINE void store_bytes(uint8_t* q, uint32_t lo, uint32_t hi, uint32_t b0, uint32_t b1) {
    for (uint32_t b = b0; b < b1; b++) q[b] = (uint8_t)((b < 4 ? lo >> (8 * b) : hi >> (8 * (b - 4))) & 0xffu);
}

// The record's bases, written as aligned 8-byte words of the FILE: the accumulator is a shift register that starts
// (offset of base 0) & 7 bytes into a word, so that a word is complete whenever kk = (that offset) + (bases put) is a
// multiple of 8 and goes out as it is. The first word of a record shares its low bytes with the header (or with the record
// before): byte stores. The "NNNNN" rule of :1711 is tracked here: `last` is kk after the most recent base that is not 'N'.
struct RecordWords {
    uint8_t* p;      // 8-byte aligned address of the word that holds base 0
    uint32_t a, kk;  // a = (offset of base 0) & 7, kk = a + bases put
    uint32_t lo, hi;
    uint32_t last;
    bool drop;
    SP_HD void begin(uint8_t* out, int64_t at) {
        p = out + (at & ~7ll); a = (uint32_t)(at & 7); kk = a; lo = hi = 0; last = kk; drop = false;
    }
    SP_HD void put(uint32_t c) {
        lo = funnel_r(lo, hi, 8);
        hi = (hi >> 8) | (c << 24);
        kk++;
        if (c != 'N') last = kk;
        else if (kk - last >= 5) drop = true;
        if ((kk & 7u) == 0) {
            if (kk == 8) {   // the record owns bytes a..7 of its first word
                store_bytes(p, lo, hi, a, 8);
            } else {
                *(uint2*)(p + (kk - 8)) = make_uint2(lo, hi);
            }
        }
    }
    // writes what is still in the accumulator, then the record's closing '\n'
    SP_HD void finish() {
        const uint32_t nb = kk < 8 ? kk - a : (kk & 7u);   // pending bases: the top nb bytes of (hi:lo)
        store_bytes(p + ((int64_t)kk - 8), lo, hi, 8 - nb, 8);   // kk < 8: a record that ends inside its first word
        p[kk] = '\n';
    }
};

SP_HD uint32_t hmm_tran(const ReadTablesDev& T, int a51, uint32_t st, uint32_t d) {
    return ld_ro(T.tran2state + (a51 + st) * 101 + fastmod(d, T.mod_tran + a51 + st) + 1);
}
SP_HD uint32_t hmm_emis(const ReadTablesDev& T, int a51, uint32_t st, uint32_t d) {
    return ld_ro(T.emis2qc + (a51 + st) * 101 + fastmod(d, T.mod_emis + a51 + st) + 1);
}

// The draw AFTER an error decision (:2362-2387). `pend` = mode | nt << 8 | dc << 16. mode 1: substitution index for the
// template base nt (per-base tables of set_mut :2211-2214); a base that is none of ACGT draws again (mode 2: one of
// "ATGC"). mode 3: insertion, the template base is not consumed. Returns outc | adv << 8 | finished << 9. A pure function
// of two scalars, kept out of line: it runs for 0.3 % of the draws and would otherwise be copied into each of the 31
// unrolled steps.
SP_HD_NOINLINE uint32_t error_draw(uint32_t pend, uint32_t d) {
    const uint32_t mode = pend & 255u, nt = (pend >> 8) & 255u;
    uint32_t outc = nt, adv = 1, fin = 1;
    if (mode == 1) {
        const uint32_t sh = 8 * (d % 3u);
        if (nt == 'A') outc = (0x434754u >> sh) & 255u;        // "TGC"
        else if (nt == 'T') outc = (0x434741u >> sh) & 255u;   // "AGC"
        else if (nt == 'G') outc = (0x435441u >> sh) & 255u;   // "ATC"
        else if (nt == 'C') outc = (0x475441u >> sh) & 255u;   // "ATG"
        else fin = 0;
    } else if (mode == 2) {
        outc = (0x43475441u >> (8 * (d % 4u))) & 255u;         // "ATGC"
    } else {
        const uint32_t idx = d % 8u;
        if (idx < 4) outc = (0x43475441u >> (8 * idx)) & 255u;
        adv = 0;
    }
    return outc | adv << 8 | fin << 9;
}

// One read, start to end, by one thread. qd = this lane's scratch: row i holds {4 quality codes, 4 deletion counters} of
// positions 4i..4i+3 as one uint2, rows interleaved by lane (qd[i * 32]), so the sequential passes are coalesced 8-byte
// accesses and one load brings both.
template <bool REPLAY, bool NOERR>
SP_HD void simulate_read_sync(const ReadJob& J, const ReadTablesDev& T, const CtaTables& C, Draws<REPLAY>& S, uint2* qd,
                              int32_t r, unsigned mask) {
    // `mask` = the lanes of this warp that hold a read; the warp_sync after every phase makes them start the next together.
    int32_t len = J.len[r], ndel = J.ndel[r];
    const int acc = J.acc[r];
    const bool bad = len < 2 || len + ndel >= J.L;  // rand() % (len-1) with len 1 traps in the reference; len' >= ref.len changes the output length
    if (bad) {
        job_error(J, len < 2 ? 1 : 2, r);
        len = 0; ndel = 0;  // every loop below runs zero times; the thread still joins the warp's phase barriers
    }
    if (REPLAY) { S.replay = J.replay + J.replay_off[r]; S.used = 0; }
    else S.seed(read_seed(J.seed, J.ref_num, r));
    S.next(); S.next();  // the length and accuracy draws (:1642, :1654), evaluated when the plan was made
    warp_sync(mask);

    // ---- quality codes (:1657-1680), four per scratch row; the deletion counters are cleared on the way. Every read
    // reaches this point after the same number of draws: blocks of 31 bases are unrolled with compile-time ring positions.
    uint32_t qmin4 = 0xffffffffu;  // bytewise minimum of the quality words: bounds every acceptance threshold below
    if (T.exist_hmm[acc]) {
        const int a51 = acc * 51;
        uint32_t st = ld_ro(T.init2state + acc * 101 + fastmod(S.next(), T.mod_init + acc) + 1);  // third draw: f is 6 now
        uint32_t word = 0;
        if (len > 0) word = hmm_emis(T, a51, st, S.next_at(6));
        // base p >= 1 draws at ring positions 5+2p (transition) and 6+2p (emission), mod 31
#pragma unroll 1
        for (int32_t p0 = 1; p0 < len; p0 += 31) {
#pragma unroll
            for (int j = 0; j < 31; j++) {
                const int32_t p = p0 + j;
                if (p < len) {
                    st = hmm_tran(T, a51, st, S.next_at((7 + 2 * j) % 31));
                    word |= hmm_emis(T, a51, st, S.next_at((8 + 2 * j) % 31)) << (8 * (p & 3));
                    if ((p & 3) == 3) { qd[(p >> 2) * 32] = make_uint2(word, 0u); qmin4 = vminu4(qmin4, word); word = 0; }
                }
            }
        }
        if (len & 3) { qd[(len >> 2) * 32] = make_uint2(word, 0u); qmin4 = vminu4(qmin4, word | (0xffffffffu << (8 * (len & 3)))); }
        if (len > 0) S.skip(2u * (uint32_t)len - 1u);
    } else {
        const uint4 mf = ld_ro4(T.mod_freq + acc);
        const uint8_t* f2q = T.freq2qc + acc * 1001;
        uint32_t word = 0;
        // two draws so far (f is 5): base p draws at ring position 5+p mod 31
#pragma unroll 1
        for (int32_t p0 = 0; p0 < len; p0 += 31) {
#pragma unroll
            for (int j = 0; j < 31; j++) {
                const int32_t p = p0 + j;
                if (p < len) {
                    word |= (uint32_t)ld_ro(f2q + fastmod(S.next_at((5 + j) % 31), mf.x, mf.y, mf.z) + 1) << (8 * (p & 3));
                    if ((p & 3) == 3) { qd[(p >> 2) * 32] = make_uint2(word, 0u); qmin4 = vminu4(qmin4, word); word = 0; }
                }
            }
        }
        if (len & 3) { qd[(len >> 2) * 32] = make_uint2(word, 0u); qmin4 = vminu4(qmin4, word | (0xffffffffu << (8 * (len & 3)))); }
        S.skip((uint32_t)len);
    }
    S.rotate();
    warp_sync(mask);

    // ---- deletions by rejection (:2282-2292): an attempt is two draws, 31 attempts are two laps of the ring
    {
        const FastMod ml = make_fastmod_dev((uint32_t)(len - 1 > 1 ? len - 1 : 1));
        const uint32_t q0 = qmin4 & 255u, q1 = (qmin4 >> 8) & 255u, q2 = (qmin4 >> 16) & 255u, q3 = qmin4 >> 24;
        uint32_t qm = q0 < q1 ? q0 : q1;
        qm = qm < q2 ? qm : q2;
        qm = qm < q3 ? qm : q3;
        const int32_t tmax = C.err_thre_max_from(qm < (uint32_t)(PB_NQ - 1) ? qm : (uint32_t)(PB_NQ - 1));
        int32_t num = 0;
        uint32_t att = 0;
#pragma unroll 1
        for (;;) {
            if (!warp_any(mask, num < ndel)) break;
#pragma unroll
            for (int j = 0; j < 31; j++) {
                if (num < ndel) {
                    const uint32_t pos = fastmod(S.next_at((2 * j) % 31), ml.M, ml.sh, ml.d);
                    const int32_t v = (int32_t)(S.next_at((2 * j + 1) % 31) % 1000000u);
                    att++;
                    // v >= the largest threshold any position of this read can have: rejected whatever q[pos] is, and the
                    // scattered read of q[pos] (one HBM access for one byte) is not made. Same decision, same draws.
                    if (v < tmax) {
                        const uint32_t sh = 8 * (pos & 3);
                        uint32_t* row = (uint32_t*)(qd + (pos >> 2) * 32);
                        const uint32_t qv = (row[0] >> sh) & 255u;
                        if (v < C.err_thre(qv)) {
                            // one more deletion after position pos: a reduction, no load to wait for. A counter byte that
                            // wraps (more than 255 deletions at one position) shows up as a wrong total below.
                            red_add_u32(row + 1, 1u << sh);
                            num++;
                        }
                    }
                }
            }
        }
        if (!REPLAY) { S.used += 2u * att; S.f = (int)((2u * att) % 31u); }
    }

    // ---- start offset (:2306-2311), header
    const uint32_t offset = bad ? 0u : S.next() % (uint32_t)(J.L - (len + ndel) + 1);
    S.rotate();
    TemplateWindow Tm;
    Tm.begin(J.seq + offset, len + ndel, (r & 1) == 0);  // res_num % 2 == 0 before the increment (:2316)
    RecordWords O;
    {
        int64_t o = J.rec_off[r];
        if (!bad) {  // ">S<ref>_<n>Z", a dozen bytes
            for (int i = 0; i < J.id_len; i++) J.out[o++] = (uint8_t)J.id[i];
            char dig[12];
            int nd = 0;
            for (int32_t v = r + 1; v > 0; v /= 10) dig[nd++] = (char)('0' + v % 10);
            while (nd > 0) J.out[o++] = (uint8_t)dig[--nd];
            J.out[o++] = 'Z';
        }
        O.begin(J.out, o);
    }

    // ---- the bases (:2348-2401), one draw per step for every lane
    int32_t k = 0;      // this lane's next base
    int32_t t = 0;      // template index: advances with every base except after an insertion, and over deleted bases
    uint32_t pend = 0;  // 0: the step's draw decides base k; else an error event of base k waits for its draw (error_draw)
    uint32_t wq = 0, wd = 0;
    uint32_t nrare = 0, nins = 0;  // draws that went to error_draw, insertions
    warp_sync(mask);
#pragma unroll 1
    for (;;) {
        if (!warp_any(mask, k < len)) break;
#pragma unroll
        for (int s = 0; s < 31; s++) {
            if (k < len) {
                const uint32_t d = S.next_at(s);
                if (pend == 0) {
                    if ((k & 3) == 0) { const uint2 w = qd[(k >> 2) * 32]; wq = w.x; wd = w.y; }
                    const uint32_t qv = wq & 255u, dc = wd & 255u;
                    wq >>= 8; wd >>= 8;
                    const uint32_t nt = Tm.at(t, C);
                    const int32_t dm = (int32_t)(d % 1000000u);
                    if (dm < C.ins_thre(qv)) {  // an error event: ins_thre >= sub_thre (:1455-1459), one test screens both
                        pend = (dm < C.sub_thre(qv) ? 1u : 3u) | nt << 8 | dc << 16;
                    } else {
                        t += (int32_t)(1u + dc);
                        if (!NOERR) O.put(nt);
                        k++;
                    }
                } else {
                    const uint32_t e = error_draw(pend, d);
                    nrare++;
                    if (e & 0x200u) {
                        const uint32_t adv = (e >> 8) & 1u;
                        nins += adv ^ 1u;
                        t += (int32_t)(adv + (pend >> 16));
                        if (!NOERR) O.put(e & 255u);
                        k++;
                        pend = 0;
                    } else {
                        pend = (pend & ~255u) | 2u;   // a template base that is none of ACGT: one more draw
                    }
                }
            }
        }
    }
    warp_sync(mask);
    if (bad) { J.dropped[r] = 1; J.used[r] = 0; return; }
    len = J.len[r]; ndel = J.ndel[r];
    if (!REPLAY) S.used += (uint32_t)len + nrare;
    // every deletion was met along the read: t = (bases that consumed a template base) + (deletions); a counter byte that
    // wrapped loses 256 here
    if (t != len - (int32_t)nins + ndel) job_error(J, 3, r);
    if (!NOERR) O.finish();
    else {  // pbsim2noError.cpp:1711 prints mut.seq, the oriented template of length len'
#pragma unroll 1
        for (int32_t i = 0; i < len + ndel; i++) O.put(Tm.at(i, C));
        O.finish();
    }
    J.dropped[r] = O.drop ? 1 : 0;
    J.used[r] = S.used;
    if (O.drop) job_count_dropped(J, (unsigned long long)(J.rec_off[r + 1] - J.rec_off[r]));
}

}  // namespace sp
