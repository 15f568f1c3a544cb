// Row P of SURVEY.md §8a — the per-read code of the read kernel (docker/pbsim2.cpp simulate_by_model :1641-1711, mutate
// :2260-2413), written so that it compiles for the device (csrc/reads.cu) AND for the host (tests/emu/reads_emu.cu runs the
// very same functions one lane at a time on the CPU and compares them with the restatement — a logic check that needs no
// GPU; it is test infrastructure, the library never runs this code on the host).
//
// Second design of the per-read code ("one draw per step"). The glibc generator r[i] = r[i-31] + r[i-3] lives in a
// shared-memory column per thread; in the first design (csrc/reads.cu) the ring position of a draw is a run-time value of
// each thread in two of the three phases (12 instructions per draw), and the pass over the bases costs 129 warp
// instructions per base because 55 % of the steps have some lane in the error path (two more draws with run-time
// positions). Here:
//   * the ring has 32 slots (value i lives in slot i & 31; it reads slots i+1 = i-31 and i-3 and writes slot i), so positions
//     wrap by masking and a block of four consecutive draws that starts at a multiple of four needs no wrap inside;
//   * the ring is physically rotated to slot 0 before the deletion loop and again before the pass over the bases (32 moves,
//     twice per read), after which every lane of a warp is at the SAME slot: the slot is loop state of the warp, not of the
//     thread, and inside a block of four draws every address is a base register plus a constant;
//   * in the pass over the bases every lane makes exactly ONE draw per step. What the draw means is the lane's own business:
//     the error decision of its next base (the common case), or — after an error event — the substitution / insertion
//     index the reference draws next (:2362-2387). A lane falls behind the step counter by one base per extra draw, keeps
//     its own base index, and loads its quality row / stores its output word under its own predicate. Lane by lane the
//     sequence of rand() calls is the reference's.
// Output words are aligned to the RECORD's position in the file (the accumulator is a shift register, a full aligned
// 8-byte word is stored as it is), so there is no funnel shift per word and no second accumulator.
// (A first version unrolled both loops over all 31 positions of a 31-slot ring: 41 % fewer instructions in the deletion
// loop, but 94 ms against 75 ms — 60 KB of straight-line code, sm__icc_request_hit_rate fell and stall_no_instruction rose
// from 0.24 to 3.6 warps per issue, profiles/r02j.)
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
SP_HD void prefetch_l1(const void* p) {  // a request with no destination register: no lane's request makes the warp wait
#ifdef __CUDA_ARCH__
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
    (void)p;
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

// The read's rand(): the thread's own glibc generator (state column: slot j of this thread is st[j * RD_THREADS], 32
// slots) or the replayed log. `pos` = the slot the next value goes to.
constexpr int RING = 32;
template <bool REPLAY>
struct Draws {
    uint32_t* st;
    const int32_t* replay;  // first logged draw of this read
    uint32_t used;          // rand() calls so far
    int pos;
    SP_HD void seed(uint32_t s) {
        uint32_t r[31];
        glibc_seed(s, r);   // glibc's front index is 3 afterwards: its next output is r[3] += r[0], so r[(3 + j) % 31] is value j of 31
#pragma unroll
        for (int j = 0; j < 31; j++) st[j * RD_THREADS] = r[(3 + j) % 31];
        pos = 31; used = 0;
    }
    // rand() at a slot the caller knows at compile time; pos and used are brought up to date by the caller
    SP_HD uint32_t next_at(int slot) {
        if (REPLAY) return (uint32_t)ld_ro(replay + used++);
        const uint32_t v = st[((slot + 1) & 31) * RD_THREADS] + st[((slot + 29) & 31) * RD_THREADS];
        st[slot * RD_THREADS] = v;
        return v >> 1;
    }
    SP_HD void skip(uint32_t n) {  // account for n draws made through next_at
        if (REPLAY) return;
        used += n;
        pos = (int)((uint32_t)(pos + n) & 31u);
    }
    SP_HD uint32_t next() {
        if (REPLAY) return (uint32_t)ld_ro(replay + used++);
        const uint32_t v = st[((pos + 1) & 31) * RD_THREADS] + st[((pos + 29) & 31) * RD_THREADS];
        st[pos * RD_THREADS] = v;
        pos = (pos + 1) & 31;
        used++;
        return v >> 1;
    }
    // Takes the last n (< 28) rand() calls back. Values i .. i+n-1 went to slots s .. s+n-1 and overwrote values i-32 ..
    // i-33+n, of which all but the first are read again (value m reads value m-31): value i-32+k = value i+k-1 - value
    // i+k-4, restored from the highest k down so that every operand is still in place.
    SP_HD void undraw(int n) {
        used -= (uint32_t)n;
        if (!REPLAY) {
            const int s = (pos - n) & 31;
            for (int k = n - 1; k >= 1; k--)
                st[((s + k) & 31) * RD_THREADS] = st[((s + k - 1) & 31) * RD_THREADS] - st[((s + k - 4) & 31) * RD_THREADS];
            pos = s;
        }
    }
    // Physical rotation of the ring: slot (j + pos) & 31 moves to slot j, afterwards pos is 0. The moves form gcd(32, pos)
    // cycles; one flat loop of 32 moves walks them one after the other (c = the cycle's first slot, t0 = its old value).
    SP_HD void rotate() {
        if (REPLAY) return;
        const int s = pos;
        if (s != 0) {
            int c = 0, j = 0;
            uint32_t t0 = st[0];
#pragma unroll 1
            for (int i = 0; i < 32; i++) {
                const int nj = (j + s) & 31;
                if (nj == c) {
                    st[j * RD_THREADS] = t0;
                    c++; j = c;
                    t0 = st[(c & 31) * RD_THREADS];
                } else {
                    st[j * RD_THREADS] = st[nj * RD_THREADS];
                    j = nj;
                }
            }
        }
        pos = 0;
    }
};

// Four consecutive draws at slots p..p+3, p a multiple of four known to the whole warp: every address is one of three base
// pointers (the blocks before, at and after p) plus a constant.
template <bool REPLAY>
struct DrawBlock {
    uint32_t *prev, *cur, *next;
    SP_HD void begin(Draws<REPLAY>& S) {   // S.pos == 0
        prev = S.st + 28 * RD_THREADS; cur = S.st; next = S.st + 4 * RD_THREADS;
    }
    SP_HD void advance(Draws<REPLAY>& S, int p_after_next) {   // p_after_next = (p + 8) & 31 of the block that ends
        prev = cur; cur = next; next = S.st + p_after_next * RD_THREADS;
    }
    template <int J>
    SP_HD uint32_t draw(Draws<REPLAY>& S) {   // draw J (0..3) of the block
        if (REPLAY) return (uint32_t)ld_ro(S.replay + S.used++);
        const uint32_t ahead = J < 3 ? cur[(J + 1) * RD_THREADS] : next[0];
        const uint32_t back = J >= 3 ? cur[(J - 3) * RD_THREADS] : prev[(J + 1) * RD_THREADS];
        const uint32_t v = ahead + back;
        cur[J * RD_THREADS] = v;
        return v >> 1;
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
    int32_t nw;                       // index of the chunk in (nlo, nhi): request_next
    uint32_t nlo, nhi;
    SP_HD void begin(const uint8_t* w, int32_t lp, bool forward) {
        const uint8_t* origin = forward ? w : w + (lp - 1);  // first base of the window (+) or its last base (-)
        o0 = (int32_t)((uintptr_t)origin & 7);
        base8 = (const unsigned long long*)(origin - o0);
        dir = (int32_t)keep_in_register(forward ? 1u : 0xffffffffu);
        strand_off = forward ? 0u : (uint32_t)LUT_REV_OFFSET; cur = INT32_MIN; lo = hi = 0;
        nw = INT32_MIN; nlo = nhi = 0;
    }
    SP_HD uint32_t at(int32_t t, const CtaTables& C) {
        const int32_t rel = o0 + dir * t;   // 32-bit: a read is far shorter than 2^31 bases
        const int32_t w = rel >> 3;         // floor, also for the negative offsets of the reverse strand
        if (w != cur) {
            if (w == nw) { lo = nlo; hi = nhi; }   // the chunk every lane requested in the same step (request_next)
            else {
                const unsigned long long chunk = ld_ro(base8 + w);
                lo = (uint32_t)chunk; hi = (uint32_t)(chunk >> 32);
            }
            cur = w;
        }
        return C.base_of(strand_off, byte_perm(lo, hi, (uint32_t)(rel & 7)) & 0xffu);
    }
    // The lanes of a warp cross their chunk boundaries in different steps: with a load at the crossing, some lane makes the
    // warp wait for memory in every step. Called by every lane in the same step, once per 8 steps: the chunk after the
    // current one goes into a second register pair, and a crossing is a register move. lp = the window's length.
    SP_HD void request_next(int32_t lp) {
        const int32_t w = cur + dir;
        if (cur != INT32_MIN && dir * (w - ((o0 + dir * (lp - 1)) >> 3)) <= 0) {   // still a chunk of this read's window
            const unsigned long long chunk = ld_ro(base8 + w);
            nlo = (uint32_t)chunk; nhi = (uint32_t)(chunk >> 32); nw = w;
        }
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

// The draw AFTER an error decision (:2362-2387), as arithmetic on constants (the reference's per-base tables of set_mut
// :2211-2214 are three characters each). `pend` = mode | nt << 8 | dc << 16. mode 1: substitution index for the template base
// nt; a base that is none of ACGT draws again (mode 2: one of "ATGC"). mode 3: insertion, the template base is not
// consumed. Returns outc | adv << 8 | finished << 9.
SP_HD uint32_t error_draw(uint32_t pend, uint32_t d) {
    const uint32_t mode = pend & 255u, nt = (pend >> 8) & 255u;
    if (mode == 1) {
        const uint32_t tab = nt == 'A' ? 0x434754u      // "TGC"
                           : nt == 'T' ? 0x434741u      // "AGC"
                           : nt == 'G' ? 0x435441u      // "ATC"
                           : nt == 'C' ? 0x475441u      // "ATG"
                           : 0u;
        return tab ? (((tab >> (8 * (d % 3u))) & 255u) | 0x300u) : (nt | 0x100u);
    }
    if (mode == 2) return ((0x43475441u >> (8 * (d & 3u))) & 255u) | 0x300u;   // "ATGC"
    return ((d & 4u) ? nt : ((0x43475441u >> (8 * (d & 3u))) & 255u)) | 0x200u;   // insertion: idx = d % 8, idx < 4 draws a base
}

// The pass over the bases (:2348-2401) as the state of one lane; step<J>() is one draw of a block of four.
template <bool REPLAY, bool NOERR>
struct BaseLoop {
    int32_t k, len;     // this lane's next base, the read's length
    int32_t t;          // template index: advances with every base except after an insertion, and over deleted bases
    uint32_t pend;      // 0: the step's draw decides base k; else an error event of base k waits for its draw (error_draw)
    uint32_t wq, wd;    // what is left of the current scratch row: quality codes, deletion counters
    uint32_t nq, nd;    // the next row, requested by every lane that had used up its last one at the start of a block of four
    bool have;          // steps (a lane enters at most one new row per block), so that entering a row is a register move
    uint32_t nrare, nins;   // draws that went to error_draw, insertions
    const uint2* row;   // the lane's next scratch row
    TemplateWindow Tm;
    RecordWords O;

    SP_HD void request_row() {
        if (!have) {
            const uint2 w = *row;   // at most one row past the read's last: inside the lane's scratch
            nq = w.x; nd = w.y;
            row += 32;
            have = true;
        }
    }
    template <int J>
    SP_HD void step(Draws<REPLAY>& S, DrawBlock<REPLAY>& B, const CtaTables& C) {
        if (k < len) {
            const uint32_t d = B.template draw<J>(S);
            uint32_t outc, adv, dc;
            bool fin;
            if (pend == 0) {
                if ((k & 3) == 0) { wq = nq; wd = nd; have = false; }
                const uint32_t qv = wq & 255u;
                dc = wd & 255u;
                wq >>= 8; wd >>= 8;
                outc = Tm.at(t, C);
                adv = 1;
                const int32_t dm = (int32_t)(d % 1000000u);
                fin = dm >= C.ins_thre(qv);   // ins_thre >= sub_thre (:1455-1459): one test screens both kinds of event
                if (!fin) pend = (dm < C.sub_thre(qv) ? 1u : 3u) | outc << 8 | dc << 16;
            } else {
                const uint32_t e = error_draw(pend, d);
                nrare++;
                outc = e & 255u; adv = (e >> 8) & 1u; fin = (e & 0x200u) != 0; dc = pend >> 16;
                nins += fin ? (adv ^ 1u) : 0u;
                pend = fin ? 0u : ((pend & ~255u) | 2u);   // a template base that is none of ACGT: one more draw
            }
            if (fin) {
                t += (int32_t)(adv + dc);
                if (!NOERR) O.put(outc);
                k++;
            }
        }
    }
};

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
    S.next(); S.next();  // the length and accuracy draws (:1642, :1654), evaluated when the plan was made: the next slot is 1
    warp_sync(mask);

    // ---- quality codes (:1657-1680), four per scratch row; the deletion counters are cleared on the way. Every read
    // reaches this point after the same number of draws, so the slot of each quality draw is a function of the base index
    // alone: blocks of 16 (two draws per base) or 32 bases are unrolled with compile-time slots.
    uint32_t qmin4 = 0xffffffffu;  // bytewise minimum of the quality words: bounds every acceptance threshold below
    if (T.exist_hmm[acc]) {
        const int a51 = acc * 51;
        uint32_t st = ld_ro(T.init2state + acc * 101 + fastmod(S.next(), T.mod_init + acc) + 1);  // third draw, slot 1
        uint32_t word = 0;
        if (len > 0) word = hmm_emis(T, a51, st, S.next_at(2));
        // base p >= 1 draws at slots 1+2p (transition) and 2+2p (emission), mod 32
#pragma unroll 1
        for (int32_t p0 = 1; p0 < len; p0 += 16) {
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const int32_t p = p0 + j;
                if (p < len) {
                    st = hmm_tran(T, a51, st, S.next_at((3 + 2 * j) & 31));
                    word |= hmm_emis(T, a51, st, S.next_at((4 + 2 * j) & 31)) << (8 * (p & 3));
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
        // two draws so far: base p draws at slot 1+p mod 32
#pragma unroll 1
        for (int32_t p0 = 0; p0 < len; p0 += 32) {
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const int32_t p = p0 + j;
                if (p < len) {
                    word |= (uint32_t)ld_ro(f2q + fastmod(S.next_at((1 + j) & 31), mf.x, mf.y, mf.z) + 1) << (8 * (p & 3));
                    if ((p & 3) == 3) { qd[(p >> 2) * 32] = make_uint2(word, 0u); qmin4 = vminu4(qmin4, word); word = 0; }
                }
            }
        }
        if (len & 3) { qd[(len >> 2) * 32] = make_uint2(word, 0u); qmin4 = vminu4(qmin4, word | (0xffffffffu << (8 * (len & 3)))); }
        S.skip((uint32_t)len);
    }
    S.rotate();
    warp_sync(mask);

    // ---- deletions by rejection (:2282-2292): an attempt is two draws, a block of four draws is two attempts
    {
        const FastMod ml = make_fastmod_dev((uint32_t)(len - 1 > 1 ? len - 1 : 1));
        const uint32_t q0 = qmin4 & 255u, q1 = (qmin4 >> 8) & 255u, q2 = (qmin4 >> 16) & 255u, q3 = qmin4 >> 24;
        uint32_t qm = q0 < q1 ? q0 : q1;
        qm = qm < q2 ? qm : q2;
        qm = qm < q3 ? qm : q3;
        const int32_t tmax = C.err_thre_max_from(qm < (uint32_t)(PB_NQ - 1) ? qm : (uint32_t)(PB_NQ - 1));
        int32_t num = 0;
        uint32_t att = 0;
        int p = 0;   // the block's first slot: the same for every lane of the warp
        DrawBlock<REPLAY> B;
        B.begin(S);
#define SP_ATTEMPT(J0, J1)                                                                                              \
    if (num < ndel) {                                                                                                   \
        const uint32_t pos = fastmod(B.template draw<J0>(S), ml.M, ml.sh, ml.d);                                        \
        const int32_t v = (int32_t)(B.template draw<J1>(S) % 1000000u);                                                 \
        att++;                                                                                                          \
        /* v >= the largest threshold any position of this read can have: rejected whatever q[pos] is, and the */      \
        /* scattered read of q[pos] (one HBM access for one byte) is not made. Same decision, same draws. */            \
        if (v < tmax) {                                                                                                 \
            const uint32_t sh = 8 * (pos & 3);                                                                          \
            uint32_t* rw = (uint32_t*)(qd + (pos >> 2) * 32);                                                           \
            const uint32_t qv = (rw[0] >> sh) & 255u;                                                                   \
            if (v < C.err_thre(qv)) {                                                                                   \
                /* one more deletion after position pos: a reduction, no load to wait for. A counter byte that wraps */ \
                /* (more than 255 deletions at one position) shows up as a wrong total below. */                        \
                red_add_u32(rw + 1, 1u << sh);                                                                          \
                num++;                                                                                                  \
            }                                                                                                           \
        }                                                                                                               \
    }
#pragma unroll 1
        for (;;) {
            if (!warp_any(mask, num < ndel)) break;
            SP_ATTEMPT(0, 1)
            SP_ATTEMPT(2, 3)
            B.advance(S, (p + 8) & 31);
            p = (p + 4) & 31;
        }
#undef SP_ATTEMPT
        if (!REPLAY) { S.used += 2u * att; S.pos = (int)((2u * att) & 31u); }
    }

    // ---- start offset (:2306-2311), header
    const uint32_t offset = bad ? 0u : S.next() % (uint32_t)(J.L - (len + ndel) + 1);
    S.rotate();
    BaseLoop<REPLAY, NOERR> W;
    W.Tm.begin(J.seq + offset, len + ndel, (r & 1) == 0);  // res_num % 2 == 0 before the increment (:2316)
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
        W.O.begin(J.out, o);
    }

    // ---- the bases (:2348-2401), one draw per step for every lane, four steps per block
    W.k = 0; W.len = len; W.t = 0; W.pend = 0; W.wq = 0; W.wd = 0; W.nrare = 0; W.nins = 0; W.row = qd;
    W.nq = 0; W.nd = 0; W.have = false;
    {
        int p = 0;
        DrawBlock<REPLAY> B;
        B.begin(S);
        warp_sync(mask);
#pragma unroll 1
        for (;;) {
            if (!warp_any(mask, W.k < len)) break;
            W.request_row();
            if ((p & 4) == 0) W.Tm.request_next(len + ndel);
            W.template step<0>(S, B, C);
            W.template step<1>(S, B, C);
            W.template step<2>(S, B, C);
            W.template step<3>(S, B, C);
            B.advance(S, (p + 8) & 31);
            p = (p + 4) & 31;
        }
    }
    warp_sync(mask);
    if (bad) { J.dropped[r] = 1; J.used[r] = 0; return; }
    len = J.len[r]; ndel = J.ndel[r];
    if (!REPLAY) S.used += (uint32_t)len + W.nrare;
    // every deletion was met along the read: t = (bases that consumed a template base) + (deletions); a counter byte that
    // wrapped loses 256 here
    if (W.t != len - (int32_t)W.nins + ndel) job_error(J, 3, r);
    if (!NOERR) W.O.finish();
    else {  // pbsim2noError.cpp:1711 prints mut.seq, the oriented template of length len'
#pragma unroll 1
        for (int32_t i = 0; i < len + ndel; i++) W.O.put(W.Tm.at(i, C));
        W.O.finish();
    }
    J.dropped[r] = W.O.drop ? 1 : 0;
    J.used[r] = S.used;
    if (W.O.drop) job_count_dropped(J, (unsigned long long)(J.rec_off[r + 1] - J.rec_off[r]));
}

}  // namespace sp
