// Building blocks of the wavefront path tracer (frt_wavefront.cuh) and of the ray-cast entry (k_trace): the path
// pool layout, the persistent-warp trace loop with dynamic fetch, and camera-sample generation.  Together with
// frt_shade.cuh they restate path::Render's sample loop and the recursion of path::Li (path.cpp:4-148) as
// `rad += thr * Le; thr *= f` over a pool of path slots kept as SoA float4 arrays in SHARED memory.
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
    TileParams tile;
    Counters *ctr;
    float *film;                // nx*ny*3 radiance sums, row 0 = bottom
    unsigned int epoch;         // this tile's launch number: frt_cancel stops the tiles with epoch <= Counters::cancel_epoch
                                // (viewer::to_exit, viewer.cpp:101-104 / path.cpp:127)
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
template <bool ANY, bool COUNT, class Io, class Stk, int WIDE = 4, bool RT = false>
__device__ __forceinline__ void frt_persistent_trace(const DevScene &sc, Io &io, Stk &stk, const unsigned int n_items, unsigned int *cursor,
                                                     const int refill_below, unsigned int &n_rays, unsigned int &n_nodes,
                                                     unsigned int &n_tris)
{
    if (n_items == 0) return;
    const unsigned int lane = lane_id();
    Walker<ANY, COUNT, WIDE, RT> w;
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
