// The reference's RNG: java.util.Random, a 48-bit LCG (s' = s*0x5DEECE66D + 0xB mod 2^48), usable on host and
// device, with O(log n) jump-ahead so that any thread can reproduce its slice of the reference draw stream
// (reference: `private static Random random = new Random()`, src/SimulateHaplotypes.java:71).
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define SP_HD __host__ __device__ __forceinline__
#else
#define SP_HD inline
#endif

namespace sp {

constexpr uint64_t LCG_A = 0x5DEECE66DULL;
constexpr uint64_t LCG_C = 0xBULL;
constexpr uint64_t LCG_MASK = (1ULL << 48) - 1;

// Affine map x -> a*x + c (mod 2^48); composition is again affine, which is what makes the LCG jumpable.
struct Affine {
    uint64_t a, c;
};
SP_HD Affine affine_identity() { return Affine{1, 0}; }
// (g o f)(x) = g.a*(f.a*x + f.c) + g.c
SP_HD Affine affine_compose(const Affine& g, const Affine& f) {
    return Affine{(g.a * f.a) & LCG_MASK, (g.a * f.c + g.c) & LCG_MASK};
}
SP_HD uint64_t affine_apply(const Affine& f, uint64_t x) { return (f.a * x + f.c) & LCG_MASK; }

// n LCG steps as one affine map, by binary exponentiation (<= 64 iterations).
SP_HD Affine lcg_jump(uint64_t n) {
    Affine r = affine_identity();
    Affine b{LCG_A, LCG_C};
    while (n) {
        if (n & 1) r = affine_compose(b, r);
        b = affine_compose(b, b);
        n >>= 1;
    }
    return r;
}

SP_HD uint64_t lcg_scramble(int64_t seed) { return ((uint64_t)seed ^ LCG_A) & LCG_MASK; }

// java.util.Random restated from its specification.
struct JavaRandom {
    uint64_t state;
    uint64_t steps;  // number of LCG steps since construction (the index of the next draw in the stream)
    SP_HD explicit JavaRandom(int64_t seed) : state(lcg_scramble(seed)), steps(0) {}
    SP_HD int32_t next(int bits) {
        state = (state * LCG_A + LCG_C) & LCG_MASK;
        steps++;
        return (int32_t)(uint32_t)(state >> (48 - bits));
    }
    SP_HD void skip(uint64_t n) {
        state = affine_apply(lcg_jump(n), state);
        steps += n;
    }
    SP_HD double nextDouble() {
        const int64_t hi = next(26);
        const int64_t lo = next(27);
        return (double)((hi << 27) + lo) * 0x1.0p-53;
    }
    SP_HD bool nextBoolean() { return next(1) != 0; }
    // nextInt(bound), bound > 0: multiply-shift for powers of two, else the rejection loop on int32 overflow.
    SP_HD int32_t nextInt(int32_t bound) {
        int32_t r = next(31);
        const int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        for (int32_t u = r;; u = next(31)) {
            r = u % bound;
            if ((int32_t)((uint32_t)u - (uint32_t)r + (uint32_t)m) >= 0) break;
        }
        return r;
    }
};

// 53-bit integer of the nextDouble() that starts `n` steps into the stream of `state0` (draw = value * 2^-53).
SP_HD uint64_t draw53_from_state(uint64_t s /* state BEFORE the two steps */) {
    const uint64_t s1 = (s * LCG_A + LCG_C) & LCG_MASK;
    const uint64_t s2 = (s1 * LCG_A + LCG_C) & LCG_MASK;
    return ((s1 >> 22) << 27) + (s2 >> 21);
}

}  // namespace sp
