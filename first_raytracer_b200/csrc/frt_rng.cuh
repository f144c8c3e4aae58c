// Per-path counter-based random numbers: Philox4x32-10 keyed by (seed), counter =
// (pixel, sample, block, tag).  Replaces the per-thread std::mt19937 of sampler.h:9-35, which
// has no per-path form.  One call yields one block of four uniforms:
//   block 0       : pixel jitter x, y, lens u, v            (path.cpp:131-133)
//   block 1 + 2k  : scatter sample 0..2, light pick          (path.cpp:36,39)   at vertex k
//   block 2 + 2k  : light sample u,v, BSDF sample u,v        (path.cpp:45,99)   at vertex k
#pragma once
#include <stdint.h>

__device__ __forceinline__ uint4 frt_philox4x32_10(uint4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r)
    {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}

// Four uniforms in [0,1) with 24 bits each (exact in float32).
__device__ __forceinline__ float4 frt_rng_block(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t block)
{
    const uint4 r = frt_philox4x32_10(make_uint4(pixel, sample, block, 0x46525431u), seed, 0x9E3779B9u);
    const float s = 1.0f / 16777216.0f;
    return make_float4((r.x >> 8) * s, (r.y >> 8) * s, (r.z >> 8) * s, (r.w >> 8) * s);
}
