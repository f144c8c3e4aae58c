// Wavefront path-tracing kernels: generate → extend → shade → connect, one launch each per
// iteration.  Together they restate path::Render's sample loop and the recursion of path::Li
// (path.cpp:4-148) as `rad += thr * Le; thr *= f` over a pool of path slots kept in HBM/L2 as
// SoA float4 arrays.  A slot whose path ended is pushed on a free queue (warp ballot + one
// atomic per warp) and refilled with the next camera sample by `generate`, so the pool stays
// full until the tile's samples run out.
#pragma once
#include "frt_device.cuh"
#include "frt_rng.cuh"
#include "frt_traverse.cuh"
#include "frt_bsdf.cuh"

#define FRT_FLAG_ALIVE 1u
#define FRT_FLAG_PREV_UNWEIGHTED 2u
#define FRT_FLAG_FINISH 4u          // ao: the occlusion ray is in flight, splat at the next shade (no more vertices) // previous vertex was modified_phong / metal / dielectric (path.cpp:18-22)
#define FRT_DEPTH_SHIFT 8

struct PathPool
{
    float4 *ray_o;  // origin.xyz, w = pdf of the BSDF sample that produced this ray (prev_bsdf_pdf)
    float4 *ray_d;  // direction.xyz (not normalised), w = bits: flags | depth << 8
    float4 *hit;    // t, b1, b2, w = bits of RayHit::prim
    float4 *thr;    // throughput.rgb, w = bits of pixel index (y * nx + x)
    float4 *rad;    // radiance gathered so far .rgb, w = bits of sample index
    float4 *prev_p; // previous vertex position (prev_hrec.p, path.cpp:25)
    float4 *sh_a;   // shadow ray: origin.xyz, w = to_light.x
    float4 *sh_b;   // to_light.y, to_light.z, contribution.r, contribution.g
    float *sh_c;    // contribution.b
    unsigned int *free_q[2];
    unsigned int *shadow_q;
    unsigned int n_slots;
};

struct TileParams
{
    unsigned int nx, ny, x0, y0, w, h;
    unsigned int spp_begin; // first sample index of this device's share
    unsigned int seed, max_depth;
    unsigned int integrator; // frt_integrator
};

struct KernelArgs
{
    DevScene sc;
    PathPool pool;
    TileParams tile;
    Counters *ctr;
    float *film; // nx*ny*3 radiance sums, row 0 = bottom
};

__device__ __forceinline__ unsigned int lane_id() { return threadIdx.x & 31u; }

// Warp-aggregated append: every lane with `pred` gets a distinct index from *counter.
__device__ __forceinline__ unsigned int warp_append(bool pred, unsigned int *counter)
{
    const unsigned int mask = __ballot_sync(__activemask(), pred);
    if (!pred) return 0;
    const unsigned int leader = __ffs(mask) - 1;
    unsigned int base = 0;
    if (lane_id() == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(mask, base, leader);
    return base + __popc(mask & ((1u << lane_id()) - 1u));
}

// ---------------------------------------------------------------------------------------------
// generate: path.cpp:129-133 + camera::get_ray camera.h:30-35 + random_in_unit_disk util.h:21-41
// ---------------------------------------------------------------------------------------------
// Persistent-warp ray casting over a queue of `n_items` work items behind an atomic cursor (shared or
// global memory).  One ray per lane; between traversal rounds the warp ballots its lanes and, once
// `refill_below` or fewer are still traversing, hands every idle lane the next item (dynamic fetch, one
// atomic per warp), so a warp is not held by its slowest ray.  (Letting lanes run several rounds between two
// looks at the warp was measured: no gain.)  Io::load(item, o, d) fetches a ray,
// Io::store(sc, walker) consumes the finished traversal; both run for the lanes of one refill together.
template <bool ANY, bool COUNT, class Io>
__device__ __forceinline__ void frt_persistent_trace(const DevScene &sc, Io &io, const unsigned int n_items, unsigned int *cursor,
                                                     const int refill_below, unsigned int &n_rays, unsigned int &n_nodes,
                                                     unsigned int &n_tris)
{
    if (n_items == 0) return;
    const unsigned int lane = lane_id();
    Walker<ANY, COUNT> w;
    uint2 stk[FRT_STACK_SIZE];
    w.active = false;
    bool pending = false, more = true;
    while (true)
    {
        if (more)
        {
            const unsigned int idle = __ballot_sync(0xFFFFFFFFu, !w.active);
            unsigned int base = 0;
            if (lane == 0) base = atomicAdd(cursor, (unsigned int)__popc(idle));
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            const unsigned int my = base + __popc(idle & ((1u << lane) - 1u));
            if (!w.active && my < n_items)
            {
                float3 o, d;
                io.load(my, o, d);
                w.begin(sc, o, d);
                pending = true;
                ++n_rays;
            }
            more = base + __popc(idle) < n_items;
        }
        unsigned int act = __ballot_sync(0xFFFFFFFFu, w.active);
        while (act)
        {
            if (w.active) w.round(sc, stk, n_nodes, n_tris);
            act = __ballot_sync(0xFFFFFFFFu, w.active);
            if (more && __popc(act) <= refill_below) break;
        }
        if (pending && !w.active)
        {
            io.store(sc, w);
            pending = false;
        }
        if (!more && act == 0) break;
    }
}

// Camera path number `idx` of the tile (sample idx / npix of pixel idx % npix) into slot `slot`.
__device__ __forceinline__ void frt_generate_path(const KernelArgs &a, const PathPool &pool, const unsigned int slot,
                                                  const unsigned long long idx)
{
    const unsigned int npix = a.tile.w * a.tile.h;
    const unsigned int s = a.tile.spp_begin + (unsigned int)(idx / npix);
    const unsigned int pix = (unsigned int)(idx % npix);
    const unsigned int px = a.tile.x0 + pix % a.tile.w, py = a.tile.y0 + pix / a.tile.w;
    const unsigned int pixel = py * a.tile.nx + px;

    const float4 r = frt_rng_block(a.tile.seed, pixel, s, 0);
    const float u = ((float)px + r.x) / (float)a.tile.nx;
    const float v = ((float)py + r.y) / (float)a.tile.ny;
    const frt_camera &c = a.sc.cam;
    float rdx = 0.0f, rdy = 0.0f;
    if (c.lens_radius != 0.0f)
    {
        const float ax = r.z * 2.0f - 1.0f, ay = r.w * 2.0f - 1.0f;
        if (!(ax == 0.0f && ay == 0.0f))
        {
            float phi, rr;
            if (ax * ax > ay * ay) { rr = ax; phi = (FRT_PI / 4.0f) * (ay / ax); }
            else { rr = ay; phi = (FRT_PI / 2.0f) - (FRT_PI / 4.0f) * (ax / ay); }
            rdx = c.lens_radius * rr * cosf(phi);
            rdy = c.lens_radius * rr * sinf(phi);
        }
    }
    const float3 offset = arr3(c.u) * rdx + arr3(c.v) * rdy;
    const float3 o = arr3(c.origin) + offset;
    const float3 d = arr3(c.lower_left_corner) + u * arr3(c.horizontal) + v * arr3(c.vertical) - arr3(c.origin) - offset;

    pool.ray_o[slot] = make_float4(o.x, o.y, o.z, 0.0f);
    pool.ray_d[slot] = make_float4(d.x, d.y, d.z, __uint_as_float(FRT_FLAG_ALIVE));
    pool.thr[slot] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(pixel));
    pool.rad[slot] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(s));
    pool.prev_p[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

__global__ void __launch_bounds__(256) k_generate(const KernelArgs a, const int parity)
{
    Counters *ctr = a.ctr;
    const unsigned int n_free = ctr->n_free[parity];
    const unsigned int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid == 0)
    {
        ctr->n_free[parity ^ 1] = 0; // filled by this iteration's shade
        ctr->n_shadow = 0;
        ctr->extend_next = 0;
        ctr->connect_next = 0;
    }
    const unsigned long long total = ctr->total_samples;
    for (unsigned int i = tid; i < ((n_free + 31u) & ~31u); i += gridDim.x * blockDim.x)
    {
        const bool want = i < n_free;
        // one atomic per warp hands out consecutive sample indices
        const unsigned int mask = __ballot_sync(0xFFFFFFFFu, want);
        unsigned long long base = 0;
        if (lane_id() == 0 && mask) base = atomicAdd(&ctr->next_sample, (unsigned long long)__popc(mask));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        const unsigned long long idx = base + __popc(mask & ((1u << lane_id()) - 1u));
        const bool start = want && idx < total;
        const unsigned int started = __popc(__ballot_sync(0xFFFFFFFFu, start));
        if (lane_id() == 0 && started)
        {
            atomicAdd(&ctr->n_alive, (int)started);
            atomicAdd(&ctr->samples_started, (unsigned long long)started);
        }
        if (!start) continue;
        frt_generate_path(a, a.pool, a.pool.free_q[parity][i], idx);
    }
}

// ---------------------------------------------------------------------------------------------
// extend: world->hit(r, EPSILON, FLT_MAX, hrec)  path.cpp:10.  Persistent warps pull batches of
// 32 slots from a global cursor so that long traversals do not hold a block's other warps.
// ---------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(128) k_extend(const KernelArgs a)
{
    unsigned int n_rays = 0, n_nodes = 0, n_tris = 0;
    const unsigned int n = a.pool.n_slots;
    while (true)
    {
        unsigned int base = 0;
        if (lane_id() == 0) base = atomicAdd(&a.ctr->extend_next, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n) break;
        const unsigned int slot = base + lane_id();
        if (slot >= n) continue;
        const float4 dw = a.pool.ray_d[slot];
        if (!(__float_as_uint(dw.w) & FRT_FLAG_ALIVE)) continue;
        const float4 ow = a.pool.ray_o[slot];
        const RayHit h = frt_traverse<false, COUNT>(a.sc, f3(ow), f3(dw), FRT_FLT_MAX, n_nodes, n_tris);
        a.pool.hit[slot] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
        ++n_rays;
    }
    // one set of atomics per warp
    for (int off = 16; off > 0; off >>= 1)
    {
        n_rays += __shfl_down_sync(0xFFFFFFFFu, n_rays, off);
        if (COUNT)
        {
            n_nodes += __shfl_down_sync(0xFFFFFFFFu, n_nodes, off);
            n_tris += __shfl_down_sync(0xFFFFFFFFu, n_tris, off);
        }
    }
    if (lane_id() == 0)
    {
        if (n_rays) atomicAdd(&a.ctr->closest_rays, (unsigned long long)n_rays);
        if (COUNT)
        {
            atomicAdd(&a.ctr->node_visits, (unsigned long long)n_nodes);
            atomicAdd(&a.ctr->tri_tests, (unsigned long long)n_tris);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// connect: !world->hit(shadow_ray, EPSILON, 1 - SHADOW_EPSILON, lrec)  path.cpp:50, any-hit
// with early exit; the contribution computed by shade is committed when unoccluded (path.cpp:73)
// ---------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(128) k_connect(const KernelArgs a)
{
    unsigned int n_rays = 0, n_nodes = 0, n_tris = 0;
    const unsigned int n = a.ctr->n_shadow;
    while (true)
    {
        unsigned int base = 0;
        if (lane_id() == 0) base = atomicAdd(&a.ctr->connect_next, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n) break;
        const unsigned int i = base + lane_id();
        if (i >= n) continue;
        const unsigned int slot = a.pool.shadow_q[i];
        const float4 sa = a.pool.sh_a[slot], sb = a.pool.sh_b[slot];
        const RayHit h = frt_traverse<true, COUNT>(a.sc, f3(sa), f3(sa.w, sb.x, sb.y), 1.0f - FRT_SHADOW_EPSILON,
                                                   n_nodes, n_tris);
        ++n_rays;
        if (h.prim == 0xFFFFFFFFu)
        {
            float4 r = a.pool.rad[slot];
            r.x += sb.z; r.y += sb.w; r.z += a.pool.sh_c[slot];
            a.pool.rad[slot] = r;
        }
    }
    for (int off = 16; off > 0; off >>= 1)
    {
        n_rays += __shfl_down_sync(0xFFFFFFFFu, n_rays, off);
        if (COUNT)
        {
            n_nodes += __shfl_down_sync(0xFFFFFFFFu, n_nodes, off);
            n_tris += __shfl_down_sync(0xFFFFFFFFu, n_tris, off);
        }
    }
    if (lane_id() == 0)
    {
        if (n_rays) atomicAdd(&a.ctr->shadow_rays, (unsigned long long)n_rays);
        if (COUNT)
        {
            atomicAdd(&a.ctr->node_visits, (unsigned long long)n_nodes);
            atomicAdd(&a.ctr->tri_tests, (unsigned long long)n_tris);
        }
    }
}
