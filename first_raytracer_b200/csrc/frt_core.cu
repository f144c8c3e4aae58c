// libfrt.so — context, scene upload, tile launch, film readback and the ray-cast entry points
// behind include/frt.h.  sm_100a only; no CPU path exists in this library.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>
#include "frt_shade.cuh"
#include "frt_wavefront.cuh"
#include "frt_wide8.cuh"
#include "frt_sort.cuh"

// Scratch of frt_trace_rays_device_noalloc (include/frt.h): the sort's keys, ray indices and histograms for up to
// FRT_TRACE_NOALLOC_MAX_RAYS rays, part of the library's device image so that the call allocates nothing.  In host code
// this NAME is only a stand-in for the symbol: frt_create asks every device for its address (cudaGetSymbolAddress,
// with that device current) and the launches get that pointer.
#define FRT_NOALLOC_SCRATCH_WORDS (4u * FRT_TRACE_NOALLOC_MAX_RAYS + 256u * FRT_SORT_MAX_BLOCKS)
static __device__ __align__(256) uint32_t g_noalloc_sort_scratch[FRT_NOALLOC_SCRATCH_WORDS];

namespace {

thread_local std::string g_create_error;

struct Device
{
    int ordinal = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t side_stream = nullptr; // frt_cancel's 4-byte copy: runs beside the persistent kernel
    cudaStream_t in_stream = nullptr, out_stream = nullptr; // frt_trace_rays: chunked copies in / out beside the casts
    std::vector<cudaEvent_t> chunk_ev;                      // three per chunk (grown on demand)
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_sum = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
    // scene
    void *d_nodes = nullptr, *d_tris = nullptr, *d_normals = nullptr, *d_spheres = nullptr, *d_mats = nullptr,
         *d_lights = nullptr, *d_env = nullptr, *d_textures = nullptr, *d_texels = nullptr, *d_tri_uv = nullptr,
         *d_nodes8 = nullptr;
    Counters *d_ctr = nullptr;
    Counters *h_ctr = nullptr; // pinned
    float *d_film = nullptr;
    float *d_film_stage = nullptr; // device 0 only: staging for a peer's film when peer access is unavailable
    size_t film_floats = 0;
    // ray casting: device buffers kept between calls (grow-only), sort scratch
    void *d_rays = nullptr, *d_hits = nullptr, *d_sort = nullptr;
    uint32_t *d_noalloc_scratch = nullptr; // this device's address of g_noalloc_sort_scratch (never freed: not an allocation)
    size_t trace_cap = 0, sort_cap = 0;
    int sms = 0;
    int blocks_trace[2][2][4] = {}; // resident CTAs of trace_kernels[wide8][any][.][occupancy]
    int blocks_wavefront[2] = {0, 0}, blocks_wavefront_other = 0; // [0] small scene shape, [1] large scene shape
    unsigned int wf_ctas_launched = 0;
    bool rendered = false;
    unsigned long long launches = 0;
    std::map<void *, size_t> cap; // bytes allocated behind each scene-array pointer above (keyed by the member's address)
};

} // namespace

struct frt_ctx
{
    std::vector<Device> dev;
    std::string error;
    uint32_t flags = 0;
    bool have_scene = false;
    bool peer_ok = true;
    DevScene host_view{}; // counts + camera; pointers filled per device
    uint32_t nx = 0, ny = 0;
    unsigned int wf_slots = 1280; // general kernel (ao / normals / direct / textured): slots per CTA (multiple of 32): 185 KB of the SM's 228 KB; the rest stays L1 for the scene and the
                                  // traversal stacks (1024 -> 1280: +2 % Cornell, +8 % veach_mi, -1 % teapot; 1568 leaves no L1: -25 %)
    // The path kernel's launch shape follows the size of the scene (profiles/r02_wf_sweep_pool_size.log): a scene that fits
    // the SM's L1 beside the pool (Cornell, veach_mi: a few KB) runs fastest as ONE 512-thread CTA with the largest pool;
    // a mesh scene (teapot: 0.2 MB of nodes and triangles) wants the L1 for the tree, and several small CTAs hide its
    // node fetches and each other's stage tails better: two 256-thread CTAs of 512 slots +8 %, four 128-thread CTAs of
    // 256 slots another +5 % (profiles/r02_wf_sweep_128x4.log; the same shape costs Cornell 23 % and veach 43 %: their
    // shade stage needs many paths per CTA to fill warps per material).  Index 0 = small, 1 = large.
    int wf_config[2] = {1, 4};                // index into wf_configs[]
    unsigned int wf_slots_of[2] = {1280, 256}; // slots per CTA of the path kernel, per shape
    size_t wf_big_scene_bytes = 128u << 10;   // nodes + triangles + spheres above this: the large-scene shape
    size_t scene_bytes = 0;
    int wf_refill = 20;           // dynamic fetch threshold (lanes still traversing)
    int wf_sort = -1;             // bit 0: bin queued rays by where they start; bit 1: shade grouped by material;
                                  // -1 = bit 0 always, bit 1 when the scene has more than one BSDF type
    int trace_occ[2] = {2, 2};    // ray casting: index into {4, 6, 8, 12} CTAs per SM for closest / occlusion casts (a large
                                  // incoherent scene is latency-bound: 8 CTAs of 64 registers measured +25 % over 4 of 128)
    int trace_refill[2] = {20, 20}; // dynamic fetch threshold for closest / occlusion casts (-1: plain batches of 32); re-measured once no
                                    // pathological ray hid the differences (profiles/r02_soup_sweep_clean.log): occlusion casts +35 % over -1
    int sort_bits[2] = {9, 1};     // FRT_TRACE_SORT key: bits per axis of the origin cell / of the direction (FRT_SORT_BITS=o,d)
    size_t trace_chunk = 1u << 21; // frt_trace_rays with pinned buffers: rays per pipelined chunk (FRT_TRACE_CHUNK)
    int wide8 = 0;                // ray casting over the 8-wide compressed nodes (FRT_WIDE8=1): off by default, measured slower
    bool have_wide8 = false;
    uint32_t n_nodes8 = 0;
    bool multi_bsdf = false;
    float ao_t_max = 0.f;
    float bounds_lo[3] = {0, 0, 0}, bounds_hi[3] = {0, 0, 0}; // scene box (root node): Morton keys of frt_trace_rays' sort
    unsigned int *h_cancel = nullptr; // pinned: the epoch frt_cancel copies into every device's Counters::cancel_epoch
    unsigned int epoch = 0;           // launch number of the last tile
    // host staging of the film (pinned, ctx-owned): float sums and the 8-bit tonemapped image
    float *h_film = nullptr;
    unsigned char *h_ldr = nullptr;
    size_t h_film_floats = 0;
    unsigned char *d_ldr = nullptr; // device 0
    frt_stats stats{};
};

namespace {

int fail(frt_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->error = buf; else g_create_error = buf;
    return code;
}

#define CU(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess)                                                                             \
            return fail(ctx, FRT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// Scene arrays keep their device allocation between uploads (renderer<gpu_path_integrator> uploads the scene every
// frame: ten cudaFree + cudaMalloc pairs per device cost more than a 1080p x 128 spp share of an 8-GPU frame); the
// copy is enqueued on the device's stream, frt_upload_scene synchronises every device once at the end.
template <class T>
int upload_array(frt_ctx *ctx, Device &d, void **dst, const T *src, size_t n)
{
    const size_t bytes = n * sizeof(T);
    size_t &cap = d.cap[(void *)dst];
    if (n == 0)
    {
        if (*dst) { cudaFree(*dst); *dst = nullptr; }
        cap = 0;
        return FRT_OK;
    }
    if (!*dst || cap < bytes)
    {
        if (*dst) { cudaFree(*dst); *dst = nullptr; }
        cap = 0;
        CU(cudaMalloc(dst, bytes));
        cap = bytes;
    }
    CU(cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, d.stream));
    return FRT_OK;
}

// Work split over host threads for the film copies (a 1080p film is 25 MB of sums in, 50 MB of doubles out).
template <class F>
void parallel_chunks(size_t n, size_t min_chunk, F fn)
{
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t want = std::max<size_t>(1, n / std::max<size_t>(1, min_chunk));
    const unsigned nt = (unsigned)std::min<size_t>(std::min<unsigned>(hw, 16u), want);
    if (nt <= 1) { fn((size_t)0, n); return; }
    std::vector<std::thread> workers;
    for (unsigned w = 1; w < nt; ++w) workers.emplace_back(fn, n * w / nt, n * (w + 1) / nt);
    fn((size_t)0, n / nt);
    for (std::thread &t : workers) t.join();
}

__global__ void k_reset_tile(Counters *ctr, unsigned long long total)
{
    ctr->next_sample = 0;
    ctr->total_samples = total;
}

// In-process multi-GPU film sum as a reduce-scatter over peer memory (NVLink loads / stores): the device that runs
// this sums float4 range [begin, end) of EVERY device's film and stores the result into device 0's film.  Range g is
// read and written by device g only, so the sum is in place; device 0 then holds the whole film.
struct PeerFilms { const float *p[16]; int n; };
__global__ void k_sum_slice(float *dst, PeerFilms films, size_t begin4, size_t end4, size_t tail_begin, size_t tail_end)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = begin4 + t0; i < end4; i += stride)
    {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = 0; k < films.n; ++k)
        {
            const float4 v = reinterpret_cast<const float4 *>(films.p[k])[i];
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        reinterpret_cast<float4 *>(dst)[i] = s;
    }
    for (size_t i = tail_begin + t0; i < tail_end; i += stride) // the < 4 floats a film size leaves over
    {
        float s = 0.f;
        for (int k = 0; k < films.n; ++k) s += films.p[k][i];
        dst[i] = s;
    }
}

// viewer::add_sample's tone map for the whole film (viewer.cpp:111-126): c = sum * (1 / ns) in double
// (geometry.h:533-538), 8-bit = (GLubyte) int(pow(1 - exp(-c), 1 / 2.2) * 255 + .5).
__global__ void k_tonemap(const float *sum, unsigned char *ldr, size_t n, double inv_ns)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    {
        const double c = (double)sum[i] * inv_ns;
        const double v = pow(1.0 - exp(-c), 1.0 / 2.2) * 255.0 + 0.5;
        // int(NaN) (a negative mean, e.g. the normals integrator) is INT_MIN on the reference's x86 build -> byte 0
        ldr[i] = (v >= -2147483648.0 && v < 2147483648.0) ? (unsigned char)(int)v : (unsigned char)0;
    }
}

// Ray source / hit sink of the ray-cast entry points: frt_ray array in, frt_hit array out (global memory).  With
// `order` (frt_trace_rays' Morton sort) work item k is ray order[k]; hits go back to the ray's own index.
struct TraceIo
{
    const float2 *rays;
    frt_hit *hits;
    const unsigned int *order;
    unsigned int i;
    __device__ __forceinline__ void load(const unsigned int item, float3 &o, float3 &d)
    {
        i = order ? __ldg(order + item) : item;
        const float2 r0 = __ldg(rays + 3 * (size_t)i), r1 = __ldg(rays + 3 * (size_t)i + 1), r2 = __ldg(rays + 3 * (size_t)i + 2);
        o = f3(r0.x, r0.y, r1.x); d = f3(r1.y, r2.x, r2.y);
    }
    __device__ __forceinline__ void put(const DevScene &sc, const RayHit h)
    {
        int prim_id = -1;
        if (h.prim != 0xFFFFFFFFu)
        {
            if (h.prim & FRT_CHILD_SPHERE)
                prim_id = (int)__float_as_uint(__ldg(sc.spheres + 2 * (h.prim & FRT_LEAF_FIRST_MASK) + 1).y);
            else
                prim_id = (int)__float_as_uint(__ldg(sc.tris + 3 * (size_t)h.prim).w);
        }
        reinterpret_cast<float4 *>(hits)[i] = make_float4(h.t, h.u, h.v, __int_as_float(prim_id));
    }
    template <class W>
    __device__ __forceinline__ void store(const DevScene &sc, W &w) { put(sc, w.finish(sc)); }
};

// Ray counts of a ray-cast launch: into the device counters frt_get_counters reads (closest_rays for closest-hit
// casts, shadow_rays for occlusion casts), one set of atomics per warp.
template <bool ANY, bool COUNT>
__device__ __forceinline__ void frt_trace_count(Counters *ctr, unsigned int n_rays, unsigned int n_nodes, unsigned int n_tris)
{
    unsigned long long r = n_rays, nn = n_nodes, nt = n_tris;
    for (int off = 16; off > 0; off >>= 1)
    {
        r += __shfl_down_sync(0xFFFFFFFFu, r, off);
        if (COUNT)
        {
            nn += __shfl_down_sync(0xFFFFFFFFu, nn, off);
            nt += __shfl_down_sync(0xFFFFFFFFu, nt, off);
        }
    }
    if (lane_id() == 0)
    {
        if (r) atomicAdd(ANY ? &ctr->shadow_rays : &ctr->closest_rays, r);
        if (COUNT)
        {
            atomicAdd(&ctr->node_visits, nn);
            atomicAdd(&ctr->tri_tests, nt);
        }
    }
}

// frt_trace_rays over the 4-wide nodes (scenes below ctx->wide8_min_tris).  Closest hit: persistent warps over a
// global cursor with dynamic fetch (frt_persistent_trace).  Occlusion queries end early and carry little state; for
// them plain batches of 32 rays measured faster, kept.  Register budget: closest hit 4 CTAs / SM (no spills);
// occlusion queries 8 (64 registers).
// OCC = CTAs of 128 threads per SM the register budget is cut for (4: 128 registers, 6: 85, 8: 64, 12: 42).
// The first FRT_TRACE_SMEM_STACK stack entries of every thread live in shared memory (HybridStack).
#ifndef FRT_TRACE_SMEM_STACK
#define FRT_TRACE_SMEM_STACK 12
#endif
template <bool ANY, bool COUNT, int OCC, int WIDE>
__device__ __forceinline__ void frt_trace_body(const DevScene &sc, Counters *ctr, const float2 *rays, frt_hit *hits, const unsigned int *order,
                                               const unsigned int n, const int refill_below)
{
    __shared__ uint2 shared_stack[FRT_TRACE_SMEM_STACK * 128];
    uint2 overflow[(WIDE == 8 ? FRT_STACK8_SIZE : FRT_STACK_SIZE) - FRT_TRACE_SMEM_STACK];
    HybridStack<FRT_TRACE_SMEM_STACK, 128> stk{shared_stack + threadIdx.x, overflow};
    unsigned int n_rays = 0, n_nodes = 0, n_tris = 0;
    TraceIo io{rays, hits, order, 0u};
    if (ANY && refill_below < 0)
    {
        // plain batches of 32 rays: occlusion queries end early and carry little state
        while (true)
        {
            unsigned int base = 0;
            if (lane_id() == 0) base = atomicAdd(&ctr->trace_next, 32u);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base >= n) break;
            if (base + lane_id() >= n) continue;
            float3 o, d;
            io.load(base + lane_id(), o, d);
            Walker<ANY, COUNT, WIDE, true> w;
            w.begin(sc, o, d);
            while (w.active) w.round(sc, stk, n_nodes, n_tris);
            io.store(sc, w);
            ++n_rays;
        }
    }
    else
        frt_persistent_trace<ANY, COUNT, TraceIo, HybridStack<FRT_TRACE_SMEM_STACK, 128>, WIDE, true>(
            sc, io, stk, n, &ctr->trace_next, refill_below < 0 ? 0 : refill_below, n_rays, n_nodes, n_tris);
    frt_trace_count<ANY, COUNT>(ctr, n_rays, n_nodes, n_tris);
}

template <bool ANY, bool COUNT, int OCC>
__global__ void __launch_bounds__(128, OCC) k_trace(DevScene sc, Counters *ctr, const float2 *rays, frt_hit *hits,
                                                          const unsigned int *order, unsigned int n, int refill_below)
{
    frt_trace_body<ANY, COUNT, OCC, 4>(sc, ctr, rays, hits, order, n, refill_below);
}

// The same entry over the 8-wide compressed nodes (frt_wide8.cuh).
template <bool ANY, bool COUNT, int OCC>
__global__ void __launch_bounds__(128, OCC) k_trace8(DevScene sc, Counters *ctr, const float2 *rays, frt_hit *hits,
                                                           const unsigned int *order, unsigned int n, int refill_below)
{
    frt_trace_body<ANY, COUNT, OCC, 8>(sc, ctr, rays, hits, order, n, refill_below);
}

// The launch shapes of the ray-cast kernels compiled into the library: [wide8][any][count][occupancy 4 / 6 / 8]
typedef void (*TraceKernel)(DevScene, Counters *, const float2 *, frt_hit *, const unsigned int *, unsigned int, int);
#define TRACE_OCC(K, A, C) {K<A, C, 4>, K<A, C, 6>, K<A, C, 8>, K<A, C, 12>}
const TraceKernel trace_kernels[2][2][2][4] = {{{TRACE_OCC(k_trace, false, false), TRACE_OCC(k_trace, false, true)},
                                                {TRACE_OCC(k_trace, true, false), TRACE_OCC(k_trace, true, true)}},
                                               {{TRACE_OCC(k_trace8, false, false), TRACE_OCC(k_trace8, false, true)},
                                                {TRACE_OCC(k_trace8, true, false), TRACE_OCC(k_trace8, true, true)}}};

// Launch shapes of k_wavefront compiled into the library (threads per CTA, CTAs per SM the register budget is cut
// for); FRT_WF_CONFIG picks one, the default is the measured best (profiles/r01_sweep_launch_shapes.log, r02_wf_sweep*.log).
// Measured and rejected in round 2 (profiles/r02_wf_sweep_warp_autonomous.log): a warp-autonomous pipeline — every warp
// running generate / trace / shade on its own 64-slot slice, no CTA barrier — lost 40 % on Cornell and 65 % on
// veach_mi (warps in different stages evict each other's code from the instruction caches, and 64 slots leave too few
// paths per material to fill a warp), gained 5 % on the teapot.
typedef void (*WfKernel)(const KernelArgs, const unsigned int, const int, const int);
struct WfConfig
{
    int threads, min_blocks;
    WfKernel kernel[2]; // [counting]
};
#define WF_CFG(T, B, O) {T, B, {k_wavefront<false, T, B, O>, k_wavefront<true, T, B, O>}}
const WfConfig wf_configs[] = {WF_CFG(256, 2, false), WF_CFG(512, 1, false), WF_CFG(384, 1, false), WF_CFG(320, 2, false),
                                WF_CFG(128, 4, false)};
const int n_wf_configs = (int)(sizeof wf_configs / sizeof wf_configs[0]);
// The general kernel: the integrators next to `path` (ao, normals_renderer, ass_1_renderer) and `path` over textured
// surfaces.  Its own instantiation, so that the path kernel of the untextured scenes carries none of that code
// (measured: -11 % when it did), in one shape.
const WfConfig wf_other = WF_CFG(512, 1, true);

// `any_tend` = upper end of occlusion queries: path.cpp:50 passes 1 - SHADOW_EPSILON (a float literal, util.h:11)
DevScene device_view(const frt_ctx *ctx, const Device &d, double any_tend = 1.0 - (double)1e-3f)
{
    DevScene sc = ctx->host_view;
    sc.any_tend = any_tend;
    const float f = (float)any_tend; // round to nearest, then step outwards: float32 brackets of the double
    sc.any_tend_lo = (double)f <= any_tend ? f : std::nextafter(f, -3.0e38f);
    sc.any_tend_hi = (double)f >= any_tend ? f : std::nextafter(f, 3.0e38f);
    sc.nodes = (const float4 *)d.d_nodes;
    sc.tris = (const float4 *)d.d_tris;
    sc.tri_normals = (const float4 *)d.d_normals;
    sc.spheres = (const float4 *)d.d_spheres;
    sc.materials = (const frt_material *)d.d_mats;
    sc.lights = (const frt_light *)d.d_lights;
    sc.env_img = (const float *)d.d_env;
    sc.textures = (const frt_texture *)d.d_textures;
    sc.tex_texels = (const float *)d.d_texels;
    sc.tri_uv = (const float *)d.d_tri_uv;
    sc.nodes8 = (const uint4 *)d.d_nodes8;
    return sc;
}

} // namespace

#include "frt_api.inl"
#include "frt_render.inl"
