/* TEST INFRASTRUCTURE (oracle side).  Counter-based random numbers shared by oracle/frt_oracle.c
 * and the -DFRT_REF_CTR_RNG build of the reference driver.  The CUDA product has its own
 * statement of the same generator (first_raytracer_b200/csrc/frt_rng.cuh); tests check that the
 * two agree.  The reference uses std::mt19937 (sampler.h:9-35), which cannot be reproduced on
 * a GPU per-path; SURVEY §8 a20 replaces it by Philox4x32-10 keyed by (seed, pixel, sample, block).
 *
 * Block layout of one path (pixel, sample):
 *   block 0          : [pixel jitter x, pixel jitter y, lens u, lens v]        path.cpp:131-133
 *   block 1 + 2*k    : [scatter sample 0,1,2, light pick]  at path vertex k    path.cpp:36,39
 *   block 2 + 2*k    : [light sample u,v, BSDF sample u,v] at path vertex k    path.cpp:45,99
 */
#ifndef FRT_ORACLE_RNG_H
#define FRT_ORACLE_RNG_H
#include <stdint.h>

static inline void frt_philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int round = 0; round < 10; ++round)
    {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

static inline void frt_rng_block(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t block, uint32_t out[4])
{
    out[0] = pixel; out[1] = sample; out[2] = block; out[3] = 0x46525431u; /* "FRT1" */
    frt_philox4x32_10(out, seed, 0x9E3779B9u);
}

/* 24-bit uniform in [0,1): exactly representable in float32 and float64. */
static inline double frt_rng_unit(uint32_t x)
{
    return (double)(x >> 8) * (1.0 / 16777216.0);
}

#endif
