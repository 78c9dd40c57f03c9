/* TEST INFRASTRUCTURE ONLY (oracle/): Philox4x32-10 (Salmon et al., SC'11, Random123) in plain C.
 *
 * The reference draws from two racy process-wide std::default_random_engine instances
 * (material.h:16-20, main.cpp:91-92,127-128), which is not reproducible under OpenMP.  The GPU path
 * replaces them by a counter-based stream; this header is the CPU statement of that stream contract
 * so that the oracle, the real reference (through oracle/shim/rng_inject.h) and the CUDA kernels
 * all see the same numbers:
 *
 *   key     = (seed, 0)
 *   counter = (pixel = y*W + x, sample, slot, 0)      slot 0 = camera jitter, slot 1+depth = bounce
 *   the k-th Random() call inside one slot returns word k (k = 0..3) of that block
 *   u       = (word >> 8) * 2^-24  in [0,1)
 *   jitter  = (double)u - 0.5      (the reference's uniform_real_distribution<double>(-0.5,0.5));
 *             slot 0: word 0 -> y jitter, word 1 -> x jitter (g++ evaluates the constructor
 *             arguments of main.cpp:126-128 right to left)
 *   word 3 of a bounce slot is reserved for Russian roulette (an extension; off in parity runs)
 */
#ifndef PT_ORACLE_PHILOX_H
#define PT_ORACLE_PHILOX_H
#include <stdint.h>

static inline void pt_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* u in (0,1): the top 24 bits times 2^-24, with the all-zero pattern mapped to 2^-25.  Zero is excluded because the
 * reference's Material::Process indexes functions_[-1] on a zero draw (material.h:43-48: `while (sample > 0)` never
 * runs) — undefined behaviour that the reference's own engine reaches once in 2^31 draws and a 24-bit stream once
 * in 2^24, i.e. in every full-size render. */
static inline float pt_u01(uint32_t w) {
    const float u = (float)(w >> 8) * (1.0f / 16777216.0f);
    return u > 0.0f ? u : (1.0f / 33554432.0f);
}

#endif
