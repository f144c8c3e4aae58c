// Device-side Morton ordering of a ray batch (an option of frt_trace_rays, timed inside the call): rays that start in
// the same part of the scene and point into the same octant are traced by the same warp, so the nodes a warp fetches
// are shared between its lanes and stay in L1 / L2.  Key = 27-bit Morton code of the origin (9 bits per axis inside
// the scene box) << 3 | direction octant; a hand-written stable LSD radix sort (8-bit digits, 4 passes: histogram per
// block -> one scan -> stable scatter) orders the ray INDICES, the rays themselves stay where the caller put them.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FRT_SORT_THREADS 256
#define FRT_SORT_MAX_BLOCKS 1024

__device__ __forceinline__ uint32_t frt_spread3(uint32_t v) // 9 bits -> every third bit
{
    v &= 0x1FFu;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

// Key layout: `ob` bits per axis of the origin's cell (Morton-interleaved, high part) and `db` bits per axis of the unit
// direction (Morton-interleaved, low part); 3 * (ob + db) <= 32.  db = 1 is the direction octant.
__global__ void k_morton_keys(const float2 *rays, unsigned int n, float3 lo, float3 inv_ext, uint32_t *keys, uint32_t *idx, int ob, int db)
{
    const float oscale = (float)(1u << ob), dscale = (float)(1u << db);
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const float2 r0 = __ldg(rays + 3 * (size_t)i), r1 = __ldg(rays + 3 * (size_t)i + 1), r2 = __ldg(rays + 3 * (size_t)i + 2);
        const float fx = fminf(fmaxf((r0.x - lo.x) * inv_ext.x, 0.0f), 0.999999f), fy = fminf(fmaxf((r0.y - lo.y) * inv_ext.y, 0.0f), 0.999999f),
                    fz = fminf(fmaxf((r1.x - lo.z) * inv_ext.z, 0.0f), 0.999999f);
        const uint32_t m = frt_spread3((uint32_t)(fx * oscale)) | (frt_spread3((uint32_t)(fy * oscale)) << 1) | (frt_spread3((uint32_t)(fz * oscale)) << 2);
        // direction components scaled by the largest one: [-1, 1] -> [0, 1); db = 1 gives the sign bits (octant)
        const float dm = fmaxf(fmaxf(fabsf(r1.y), fabsf(r2.x)), fmaxf(fabsf(r2.y), 1e-30f));
        const float gx = fminf(fmaxf(0.5f + 0.5f * r1.y / dm, 0.0f), 0.999999f), gy = fminf(fmaxf(0.5f + 0.5f * r2.x / dm, 0.0f), 0.999999f),
                    gz = fminf(fmaxf(0.5f + 0.5f * r2.y / dm, 0.0f), 0.999999f);
        const uint32_t dcode = frt_spread3((uint32_t)(gx * dscale)) | (frt_spread3((uint32_t)(gy * dscale)) << 1) | (frt_spread3((uint32_t)(gz * dscale)) << 2);
        keys[i] = (m << (3 * db)) | dcode;
        idx[i] = i;
    }
}

// hist[digit * n_blocks + block] = how many keys of the block's range carry `digit` in this pass
__global__ void __launch_bounds__(FRT_SORT_THREADS) k_radix_hist(const uint32_t *keys, unsigned int n, unsigned int per_block, int shift,
                                                                 unsigned int *hist)
{
    __shared__ unsigned int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned int b0 = blockIdx.x * per_block, b1 = min(n, b0 + per_block);
    for (unsigned int i = b0 + threadIdx.x; i < b1; i += FRT_SORT_THREADS) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
    __syncthreads();
    hist[threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of the 256 * n_blocks counts in digit-major order, one block
__global__ void __launch_bounds__(1024) k_radix_scan(unsigned int *hist, unsigned int total)
{
    __shared__ unsigned int part[1024];
    const unsigned int per = (total + 1023u) / 1024u, b0 = threadIdx.x * per, b1 = min(total, b0 + per);
    unsigned int s = 0;
    for (unsigned int i = b0; i < b1; ++i) s += hist[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (unsigned int off = 1; off < 1024u; off <<= 1)
    {
        const unsigned int v = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned int run = part[threadIdx.x] - s;
    for (unsigned int i = b0; i < b1; ++i) { const unsigned int c = hist[i]; hist[i] = run; run += c; }
}

// stable scatter: the block walks its range in chunks of 256 keys; a key's place = its digit's running offset +
// the number of equal digits before it in the chunk (match.any inside the warp, per-warp counts across warps)
__global__ void __launch_bounds__(FRT_SORT_THREADS) k_radix_scatter(const uint32_t *keys_in, const uint32_t *idx_in, uint32_t *keys_out,
                                                                    uint32_t *idx_out, unsigned int n, unsigned int per_block, int shift,
                                                                    const unsigned int *hist)
{
    __shared__ unsigned int base[256];
    __shared__ unsigned int cnt[FRT_SORT_THREADS / 32][256];
    const unsigned int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    base[tid] = hist[tid * gridDim.x + blockIdx.x];
    const unsigned int b0 = blockIdx.x * per_block, b1 = min(n, b0 + per_block);
    for (unsigned int c0 = b0; c0 < b1; c0 += FRT_SORT_THREADS)
    {
        for (int w = 0; w < FRT_SORT_THREADS / 32; ++w) cnt[w][tid] = 0;
        __syncthreads();
        const unsigned int i = c0 + tid;
        const bool valid = i < b1;
        uint32_t key = 0, id = 0, digit = 256u + lane; // invalid lanes: a digit nobody shares
        if (valid) { key = keys_in[i]; id = idx_in[i]; digit = (key >> shift) & 255u; }
        const unsigned int peers = __match_any_sync(0xFFFFFFFFu, digit);
        const unsigned int rank = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank == 0) cnt[warp][digit] = __popc(peers);
        __syncthreads();
        if (valid)
        {
            unsigned int before = 0;
            for (unsigned int w = 0; w < warp; ++w) before += cnt[w][digit];
            const unsigned int dst = base[digit] + before + rank;
            keys_out[dst] = key;
            idx_out[dst] = id;
        }
        __syncthreads();
        unsigned int add = 0;
        for (int w = 0; w < FRT_SORT_THREADS / 32; ++w) add += cnt[w][tid];
        base[tid] += add;
        __syncthreads();
    }
}

// Scratch layout in `scratch` (uint32 words): keys A | idx A | keys B | idx B | hist (256 * FRT_SORT_MAX_BLOCKS)
inline size_t frt_sort_scratch_words(size_t n) { return 4 * n + 256 * (size_t)FRT_SORT_MAX_BLOCKS; }

// Leaves the sorted ray indices in *order_out (inside scratch).  4 passes; all launches on `stream`.
inline cudaError_t frt_sort_rays(const float2 *rays, unsigned int n, const float lo[3], const float hi[3], uint32_t *scratch,
                                 const uint32_t **order_out, cudaStream_t stream, unsigned long long *launches, int origin_bits = 9,
                                 int dir_bits = 1)
{
    uint32_t *keys[2] = {scratch, scratch + 2 * (size_t)n}, *idx[2] = {scratch + n, scratch + 3 * (size_t)n};
    unsigned int *hist = scratch + 4 * (size_t)n;
    const float3 l = make_float3(lo[0], lo[1], lo[2]);
    const float3 inv = make_float3(hi[0] > lo[0] ? 1.0f / (hi[0] - lo[0]) : 0.0f, hi[1] > lo[1] ? 1.0f / (hi[1] - lo[1]) : 0.0f,
                                   hi[2] > lo[2] ? 1.0f / (hi[2] - lo[2]) : 0.0f);
    const unsigned int kb = (unsigned int)((n + 255u) / 256u);
    k_morton_keys<<<kb < 4096u ? kb : 4096u, 256, 0, stream>>>(rays, n, l, inv, keys[0], idx[0], origin_bits, dir_bits);
    unsigned int n_blocks = (n + 4095u) / 4096u;
    if (n_blocks > FRT_SORT_MAX_BLOCKS) n_blocks = FRT_SORT_MAX_BLOCKS;
    unsigned int per_block = (n + n_blocks - 1u) / n_blocks;
    per_block = (per_block + FRT_SORT_THREADS - 1u) / FRT_SORT_THREADS * FRT_SORT_THREADS;
    n_blocks = (n + per_block - 1u) / per_block;
    int cur = 0;
    for (int pass = 0; pass < 4; ++pass)
    {
        const int shift = 8 * pass;
        k_radix_hist<<<n_blocks, FRT_SORT_THREADS, 0, stream>>>(keys[cur], n, per_block, shift, hist);
        k_radix_scan<<<1, 1024, 0, stream>>>(hist, 256u * n_blocks);
        k_radix_scatter<<<n_blocks, FRT_SORT_THREADS, 0, stream>>>(keys[cur], idx[cur], keys[cur ^ 1], idx[cur ^ 1], n, per_block, shift, hist);
        cur ^= 1;
    }
    if (launches) *launches += 13;
    *order_out = idx[cur];
    return cudaGetLastError();
}
