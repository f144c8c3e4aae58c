// libfrt.so — context, scene upload, wavefront driver loop, film readback and the ray-cast
// entry points behind include/frt.h.  sm_100a only; no CPU path exists in this library.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "frt_shade.cuh"
#include "frt_wavefront.cuh"

namespace {

thread_local std::string g_create_error;

struct KernelTimer
{
    std::vector<cudaEvent_t> ev; // pairs
    std::vector<int> kind;
    double ms[4] = {0, 0, 0, 0};
};

struct Device
{
    int ordinal = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    // scene
    void *d_nodes = nullptr, *d_tris = nullptr, *d_normals = nullptr, *d_spheres = nullptr, *d_mats = nullptr,
         *d_lights = nullptr, *d_env = nullptr, *d_textures = nullptr, *d_texels = nullptr, *d_tri_uv = nullptr;
    // pool
    PathPool pool{};
    std::vector<void *> pool_allocs;
    Counters *d_ctr = nullptr;
    Counters *h_ctr = nullptr; // pinned
    float *d_film = nullptr;
    float *d_film_peer = nullptr; // staging when peer access is unavailable
    size_t film_floats = 0;
    int sms = 0;
    int blocks_extend = 0, blocks_connect = 0, blocks_trace = 0, blocks_trace_any = 0, blocks_wavefront = 0, blocks_wavefront_other = 0;
    unsigned int wf_ctas_launched = 0;
    bool rendered = false;
    // per-render bookkeeping
    bool done = true;
    unsigned long long iterations = 0, launches = 0, tile_iter = 0;
    KernelTimer timers;
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
    unsigned int n_slots = 1u << 20; // streamed (global-memory) pool
    unsigned int wf_slots = 1024;    // CTA-resident pool: slots per CTA (multiple of 32)
    int wf_config = 4;               // index into wf_configs[]: 512 threads, one CTA per SM (measured best)
    int wf_refill = 24;              // dynamic fetch threshold (lanes still traversing)
    int wf_sort = -1;                // bit 0: bin queued rays by where they start; bit 1: shade grouped by material;
                                     // -1 = bit 0 always, bit 1 when the scene has more than one BSDF type
    bool multi_bsdf = false;
    float ao_t_max = 0.f;
    int poll_every = 16;
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

template <class T>
int upload_array(frt_ctx *ctx, void **dst, const T *src, size_t n)
{
    if (*dst) { cudaFree(*dst); *dst = nullptr; }
    if (n == 0) return FRT_OK;
    CU(cudaMalloc(dst, n * sizeof(T)));
    CU(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
    return FRT_OK;
}

int alloc_pool(frt_ctx *ctx, Device &d, unsigned int n)
{
    for (void *p : d.pool_allocs) cudaFree(p);
    d.pool_allocs.clear();
    auto grab = [&](size_t bytes, void **out) -> cudaError_t {
        cudaError_t e = cudaMalloc(out, bytes);
        if (e == cudaSuccess) d.pool_allocs.push_back(*out);
        return e;
    };
    PathPool &p = d.pool;
    p.n_slots = n;
    CU(grab(sizeof(float4) * n, (void **)&p.ray_o));
    CU(grab(sizeof(float4) * n, (void **)&p.ray_d));
    CU(grab(sizeof(float4) * n, (void **)&p.hit));
    CU(grab(sizeof(float4) * n, (void **)&p.thr));
    CU(grab(sizeof(float4) * n, (void **)&p.rad));
    CU(grab(sizeof(float4) * n, (void **)&p.prev_p));
    CU(grab(sizeof(float4) * n, (void **)&p.sh_a));
    CU(grab(sizeof(float4) * n, (void **)&p.sh_b));
    CU(grab(sizeof(float) * n, (void **)&p.sh_c));
    CU(grab(sizeof(unsigned int) * n, (void **)&p.free_q[0]));
    CU(grab(sizeof(unsigned int) * n, (void **)&p.free_q[1]));
    CU(grab(sizeof(unsigned int) * n, (void **)&p.shadow_q));
    return FRT_OK;
}

__global__ void k_reset_pool(PathPool pool, Counters *ctr, unsigned long long total)
{
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < pool.n_slots)
    {
        pool.free_q[0][i] = i;
        pool.ray_d[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (i == 0)
    {
        ctr->next_sample = 0;
        ctr->total_samples = total;
        ctr->n_alive = 0;
        ctr->n_free[0] = pool.n_slots;
        ctr->n_free[1] = 0;
        ctr->n_shadow = 0;
        ctr->extend_next = ctr->connect_next = 0;
    }
}

__global__ void k_reset_tile(Counters *ctr, unsigned long long total)
{
    ctr->next_sample = 0;
    ctr->total_samples = total;
}

// Sum the other devices' films into device 0's over NVLink peer loads (one pass, no staging).
struct PeerFilms { const float *p[16]; int n; };
__global__ void k_sum_films(float *dst, PeerFilms peers, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    {
        float s = dst[i];
        for (int k = 0; k < peers.n; ++k) s += peers.p[k][i];
        dst[i] = s;
    }
}

// Ray source / hit sink of the ray-cast entry points: frt_ray array in, frt_hit array out (global memory).
struct TraceIo
{
    const float2 *rays;
    frt_hit *hits;
    unsigned int i;
    __device__ __forceinline__ void load(const unsigned int item, float3 &o, float3 &d)
    {
        i = item;
        const float2 r0 = __ldg(rays + 3 * (size_t)i), r1 = __ldg(rays + 3 * (size_t)i + 1), r2 = __ldg(rays + 3 * (size_t)i + 2);
        o = f3(r0.x, r0.y, r1.x); d = f3(r1.y, r2.x, r2.y);
    }
    template <class W>
    __device__ __forceinline__ void store(const DevScene &sc, W &w)
    {
        const RayHit h = w.finish(sc);
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
};

// frt_trace_rays: persistent warps over a global cursor with dynamic fetch (frt_persistent_trace).
// frt_trace_rays.  Closest hit: persistent warps over a global cursor with dynamic fetch
// (frt_persistent_trace): on the 10M-triangle soup 308 -> 467 Mrays/s.  Occlusion queries end early and
// carry little state; for them plain batches of 32 rays measured faster (357 vs 333 Mrays/s), kept.
// Register budget: closest hit 4 CTAs / SM (no spills); occlusion queries 8 (64 registers; measured on the soup:
// 351 -> 386 Mrays/s; for closest hit 5 / 6 / 8 CTAs measured within 3 % of 4).
template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(128, ANY ? 8 : 4) k_trace(DevScene sc, Counters *ctr, const float2 *rays, frt_hit *hits,
                                                                    unsigned int n, int refill_below)
{
    unsigned int n_rays = 0, n_nodes = 0, n_tris = 0;
    TraceIo io{rays, hits, 0u};
    if (ANY)
    {
        while (true)
        {
            unsigned int base = 0;
            if (lane_id() == 0) base = atomicAdd(&ctr->trace_next, 32u);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base >= n) break;
            if (base + lane_id() >= n) continue;
            float3 o, d;
            io.load(base + lane_id(), o, d);
            Walker<ANY, COUNT> w;
            uint2 stk[FRT_STACK_SIZE];
            w.begin(sc, o, d);
            while (w.active) w.round(sc, stk, n_nodes, n_tris);
            io.store(sc, w);
        }
    }
    else
        frt_persistent_trace<ANY, COUNT>(sc, io, n, &ctr->trace_next, refill_below, n_rays, n_nodes, n_tris);
    if (COUNT)
    {
        for (int off = 16; off > 0; off >>= 1)
        {
            n_nodes += __shfl_down_sync(0xFFFFFFFFu, n_nodes, off);
            n_tris += __shfl_down_sync(0xFFFFFFFFu, n_tris, off);
        }
        if (lane_id() == 0)
        {
            atomicAdd(&ctr->node_visits, (unsigned long long)n_nodes);
            atomicAdd(&ctr->tri_tests, (unsigned long long)n_tris);
        }
    }
}

// Launch shapes of k_wavefront compiled into the library (threads per CTA, CTAs per SM the register
// budget is cut for); FRT_WF_CONFIG picks one, the default is the measured best.
struct WfConfig
{
    int threads, min_blocks;
    void (*plain)(const KernelArgs, const unsigned int, const int, const int);
    void (*counting)(const KernelArgs, const unsigned int, const int, const int);
};
#define WF_CFG(T, B) {T, B, k_wavefront<false, T, B, false>, k_wavefront<true, T, B, false>}
const WfConfig wf_configs[] = {WF_CFG(256, 2), WF_CFG(128, 4), WF_CFG(128, 5), WF_CFG(256, 3), WF_CFG(512, 1), WF_CFG(128, 3),
                               WF_CFG(192, 3), WF_CFG(160, 4), WF_CFG(320, 2), WF_CFG(224, 3), WF_CFG(288, 2),
                               WF_CFG(384, 1), WF_CFG(448, 1), WF_CFG(192, 2)};
const int n_wf_configs = (int)(sizeof wf_configs / sizeof wf_configs[0]);
// The same kernel with the shade stage of the integrators next to `path` (ao, normals_renderer, ass_1_renderer): its
// own instantiation, so that the path kernel carries none of their code (measured: -11 % when it did), in one shape.
const WfConfig wf_other = {512, 1, k_wavefront<false, 512, 1, true>, k_wavefront<true, 512, 1, true>};

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
    return sc;
}

} // namespace

#include "frt_api.inl"
#include "frt_render.inl"
