// CTA-resident wavefront: the whole generate → extend → shade → connect pipeline of one tile in ONE
// persistent kernel.  Every CTA owns a private pool of P path slots that lives in SHARED MEMORY for
// the whole render (144 B per slot: 8 float4 SoA arrays + shadow contribution + three queues), so
// path state never touches HBM; the only global traffic is scene fetches (L1/L2 resident for the
// reference's scenes), one sample-counter atomic per CTA iteration and the film splats.
//
//   generate : one global atomic hands the CTA a run of consecutive camera samples for the slots
//              the last shade stage freed (path.cpp:129-133, camera.h:30-35)
//   trace    : persistent warps pull rays from the CTA's extension queue, then from its shadow
//              queue (smem cursors), one ray per lane; a lane whose ray finishes takes the next
//              queue entry as soon as enough lanes of its warp are idle (dynamic fetch), so a warp
//              is not held by its slowest ray (path.cpp:10 and :50)
//   shade    : everything path::Li does at a vertex (path.cpp:11-115); continuing / ended /
//              shadow-casting slots are compacted into the three queues by warp ballot + one
//              shared-memory atomic per warp
// CTAs run unsynchronised: while one is in its (latency-bound) shade stage its SM neighbour is
// usually in its (issue-bound) trace stage.
#pragma once
#include "frt_shade.cuh"

#define FRT_WF_SLOT_BYTES 148u // + origin-bin key (1 B), material key (1 B), shade order (2 B) per slot

#define FRT_WF_BINS 64
struct WfShared
{
    unsigned int hist_ext[FRT_WF_BINS], hist_sh[FRT_WF_BINS]; // ray-origin bins: counts, then write offsets
    unsigned int hist_mat[8];                                 // shade order: miss + the six material types
    unsigned long long base; // first sample index of this iteration's run
    unsigned int n_ext, n_free, n_shadow, n_new, cur_ext, cur_shadow, exhausted, cancelled;
    unsigned int cancel_seen;  // frt_cancel's flag as the CTA's last warp read it during the previous shade stage
    unsigned int shadow_first; // this iteration traces the shadow queue before the extension queue
};

__device__ __forceinline__ unsigned long long frt_globaltimer()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Ray source / hit sink of the CTA-resident pool: queue entry -> slot -> ray; result -> slot.
template <bool ANY>
struct WfQueueIo
{
    const PathPool &lp;
    const unsigned int *q;
    unsigned int slot;
    __device__ __forceinline__ void load(const unsigned int item, float3 &o, float3 &d)
    {
        slot = q[item];
        if (ANY)
        {
            const float4 sa = lp.sh_a[slot], sb = lp.sh_b[slot];
            o = f3(sa); d = f3(sa.w, sb.x, sb.y);
        }
        else
        {
            o = f3(lp.ray_o[slot]); d = f3(lp.ray_d[slot]);
        }
    }
    template <class W>
    __device__ __forceinline__ void store(const DevScene &, W &w)
    {
        if (ANY)
        {
            // path.cpp:73 — the contribution shade computed is committed when nothing is in the way
            if (w.best.prim == 0xFFFFFFFFu)
            {
                const float4 sb = lp.sh_b[slot];
                float4 r = lp.rad[slot];
                r.x += sb.z; r.y += sb.w; r.z += lp.sh_c[slot];
                lp.rad[slot] = r;
            }
        }
        else
        {
            const RayHit h = w.finish_lazy(); // float64 t, u, v are resolved in shade, with the warp converged
            lp.hit[slot] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
        }
    }
};

template <bool ANY, bool COUNT>
__device__ __forceinline__ void wf_trace_queue(const DevScene &sc, const PathPool &lp, const unsigned int *q,
                                               const unsigned int n_items, unsigned int *cursor, const int refill_below,
                                               unsigned int &n_rays, unsigned int &n_nodes, unsigned int &n_tris)
{
    WfQueueIo<ANY> io{lp, q, 0u};
    uint2 stack_mem[FRT_STACK_SIZE]; // local memory; L1-resident next to a scene that fits the caches
    LocalStack stk{stack_mem};
    frt_persistent_trace<ANY, COUNT>(sc, io, stk, n_items, cursor, refill_below, n_rays, n_nodes, n_tris);
}

// Statistics: one set of 64-bit atomics per warp; the per-thread counters restart from zero.
template <bool COUNT>
__device__ __forceinline__ void wf_flush_counters(Counters *ctr, unsigned int &n_closest, unsigned int &n_shadow_rays, unsigned int &n_started,
                                                  unsigned int &nonfinite, unsigned int &n_nodes, unsigned int &n_tris)
{
    unsigned long long c = n_closest, s = n_shadow_rays, g = n_started, f = nonfinite, nn = n_nodes, nt = n_tris;
    for (int off = 16; off > 0; off >>= 1)
    {
        c += __shfl_down_sync(0xFFFFFFFFu, c, off);
        s += __shfl_down_sync(0xFFFFFFFFu, s, off);
        g += __shfl_down_sync(0xFFFFFFFFu, g, off);
        f += __shfl_down_sync(0xFFFFFFFFu, f, off);
        if (COUNT)
        {
            nn += __shfl_down_sync(0xFFFFFFFFu, nn, off);
            nt += __shfl_down_sync(0xFFFFFFFFu, nt, off);
        }
    }
    if (lane_id() == 0)
    {
        if (c) atomicAdd(&ctr->closest_rays, c);
        if (s) atomicAdd(&ctr->shadow_rays, s);
        if (g) atomicAdd(&ctr->samples_started, g);
        if (f) atomicAdd(&ctr->nonfinite, f);
        if (COUNT)
        {
            atomicAdd(&ctr->node_visits, nn);
            atomicAdd(&ctr->tri_tests, nt);
        }
    }
    n_closest = n_shadow_rays = n_started = nonfinite = n_nodes = n_tris = 0;
}

// P = slots per CTA (a multiple of 32); refill_below = a warp fetches new rays once this many lanes or
// fewer are still traversing.
template <bool COUNT, int THREADS, int MIN_BLOCKS, bool OTHER>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) k_wavefront(const KernelArgs a, const unsigned int P, const int refill_below,
                                                                   const int sort_rays)
{
    extern __shared__ float4 wf_smem[];
    __shared__ WfShared sh;
    PathPool lp;
    lp.n_slots = P;
    lp.ray_o = wf_smem; lp.ray_d = lp.ray_o + P; lp.hit = lp.ray_d + P; lp.thr = lp.hit + P; lp.rad = lp.thr + P;
    lp.prev_p = lp.rad + P; lp.sh_a = lp.prev_p + P; lp.sh_b = lp.sh_a + P;
    lp.sh_c = reinterpret_cast<float *>(lp.sh_b + P);
    unsigned int *ext_q = reinterpret_cast<unsigned int *>(lp.sh_c + P), *free_q = ext_q + P, *shadow_q = free_q + P;

    const unsigned int tid = threadIdx.x;
    for (unsigned int i = tid; i < P; i += THREADS)
    {
        lp.ray_d[i] = make_float4(0.f, 0.f, 0.f, 0.f); // dead
        free_q[i] = i;
    }
    if (tid == 0)
    {
        sh.n_ext = 0; sh.n_free = P; sh.n_shadow = 0; sh.n_new = 0; sh.exhausted = 0; sh.cancel_seen = 0; sh.shadow_first = 0;
    }
    __syncthreads();

    const unsigned long long total = a.ctr->total_samples;
    unsigned int n_closest = 0, n_shadow_rays = 0, n_nodes = 0, n_tris = 0, nonfinite = 0, n_started = 0, n_iter = 0;
    unsigned long long t_gen = 0, t_ext = 0, t_con = 0, t_shade = 0, t0 = 0;
    if (tid == 0) t0 = frt_globaltimer();

    while (true)
    {
        // ---- generate -------------------------------------------------------------------------
        if (tid == 0)
        {
            const unsigned int want = sh.n_free;
            unsigned long long base = total;
            // frt_cancel (viewer::to_exit, viewer.cpp:101-104; path.cpp:127 leaves its row loop): polled once per
            // CTA iteration (by another warp, during the shade stage, so that nobody waits for the load here); a
            // cancelled tile takes no more samples and drops the paths in flight
            const bool cancelled = sh.cancel_seen != 0;
            sh.cancelled = cancelled;
            // the two trace stages run back to back without a barrier, so only the second one's tail (its last rays,
            // most lanes of the CTA idle) is exposed: put the stage that has been taking longer first
            sh.shadow_first = t_con > t_ext ? 1u : 0u;
            if (want && !sh.exhausted && !cancelled)
            {
                base = atomicAdd(&a.ctr->next_sample, (unsigned long long)want);
                if (base + want >= total) sh.exhausted = 1;
            }
            sh.base = base;
            sh.n_new = base < total ? (unsigned int)min((unsigned long long)want, total - base) : 0u;
            sh.cur_ext = 0; sh.cur_shadow = 0;
        }
        __syncthreads();
        if (sh.cancelled) break;
        const unsigned int n_new = sh.n_new, n_bounce = sh.n_ext, n_sh = sh.n_shadow;
        const unsigned long long base = sh.base;
        for (unsigned int i = tid; i < n_new; i += THREADS)
        {
            const unsigned int slot = free_q[i];
            frt_generate_path(a, lp, slot, base + i);
            ext_q[n_bounce + i] = slot; // bounce rays first (the long ones), camera rays after them
            ++n_started;
        }
        const unsigned int n_ext = n_bounce + n_new;
        if (n_ext == 0) break; // nothing in flight and no samples left: this CTA is done
        __syncthreads();
        if (tid == 0) { const unsigned long long t = frt_globaltimer(); t_gen += t - t0; t0 = t; }

        // ---- extend + connect -----------------------------------------------------------------
        const bool shadow_first = sh.shadow_first != 0;
#pragma unroll 1
        for (int pass = 0; pass < 2; ++pass) // a loop, so that each trace loop is compiled into the kernel once
        {
            if ((pass == 0) == shadow_first)
                wf_trace_queue<true, COUNT>(a.sc, lp, shadow_q, n_sh, &sh.cur_shadow, refill_below, n_shadow_rays, n_nodes, n_tris);
            else
                wf_trace_queue<false, COUNT>(a.sc, lp, ext_q, n_ext, &sh.cur_ext, refill_below, n_closest, n_nodes, n_tris);
            if (pass == 0 && tid == 0)
            {
                const unsigned long long t = frt_globaltimer();
                if (shadow_first) t_con += t - t0; else t_ext += t - t0;
                t0 = t;
            }
        }
        __syncthreads();
        if (tid == 0)
        {
            sh.n_ext = 0; sh.n_free = 0; sh.n_shadow = 0;
            const unsigned long long t = frt_globaltimer();
            if (shadow_first) t_ext += t - t0; else t_con += t - t0;
            t0 = t;
        }
        if (tid == THREADS - 32) sh.cancel_seen = *(const volatile unsigned int *)&a.ctr->cancel_epoch >= a.epoch; // device memory: an L2 hit
        __syncthreads();

        // ---- shade + queue compaction -----------------------------------------------------------
        if (sort_rays)
        {
            // Rays are queued binned by where they start (the leaf-order index of the primitive the vertex
            // lies on is a spatial key), so a warp of the next trace stage pulls rays that leave the same
            // part of the scene: counting sort in shared memory, hist -> exclusive scan -> scatter.
            unsigned char *key = reinterpret_cast<unsigned char *>(shadow_q + P); // P bytes: bin | 0x80 shadow, 0xFF = not queued
            unsigned char *mkey = key + P;                                        // P bytes: what the path hit (material type)
            unsigned short *order = reinterpret_cast<unsigned short *>(mkey + P); // P entries: slots in shading order
            const bool by_material = sort_rays & 2;
            if (tid < FRT_WF_BINS) { sh.hist_ext[tid] = 0; sh.hist_sh[tid] = 0; }
            if (by_material)
            {
                // Scenes with several BSDFs: shade the slots grouped by the material they hit, so a warp runs one
                // BSDF's code (veach_mi: conductor plates + lambertian floor + lights ran at 8/32 lanes before).
                // Counting sort over 8 keys, one shared-memory atomic per distinct key per warp (match.any).
                if (tid < 8) sh.hist_mat[tid] = 0;
                for (unsigned int slot = tid; slot < P; slot += THREADS) key[slot] = 0xFFu;
                __syncthreads();
                const unsigned int n32 = (n_ext + 31u) & ~31u;
                for (int pass = 0; pass < 2; ++pass)
                {
                    for (unsigned int i = tid; i < n32; i += THREADS)
                    {
                        unsigned int k = 0xFFu, slot = 0;
                        if (i < n_ext)
                        {
                            slot = ext_q[i];
                            if (pass == 0)
                            {
                                const unsigned int prim = __float_as_uint(lp.hit[slot].w);
                                k = 0;
                                if (prim != 0xFFFFFFFFu)
                                {
                                    const unsigned int idx = prim & FRT_LEAF_FIRST_MASK;
                                    const unsigned int mat = (prim & FRT_CHILD_SPHERE)
                                                                 ? __float_as_uint(__ldg(a.sc.spheres + 2 * idx + 1).x)
                                                                 : __float_as_uint(__ldg(a.sc.tris + 3 * (size_t)idx + 1).w);
                                    k = 1u + min(a.sc.materials[mat].type, 6u);
                                }
                                mkey[slot] = (unsigned char)k;
                            }
                            else
                                k = mkey[slot];
                        }
                        const unsigned int peers = __match_any_sync(0xFFFFFFFFu, k);
                        const unsigned int leader = __ffs(peers) - 1, rank = __popc(peers & ((1u << lane_id()) - 1u));
                        unsigned int base_k = 0;
                        if (k != 0xFFu && lane_id() == leader) base_k = atomicAdd(&sh.hist_mat[k], (unsigned int)__popc(peers));
                        base_k = __shfl_sync(0xFFFFFFFFu, base_k, leader);
                        if (pass == 1 && k != 0xFFu) order[base_k + rank] = (unsigned short)slot;
                    }
                    __syncthreads();
                    if (pass == 0)
                    {
                        if (tid == 0)
                        {
                            unsigned int run = 0;
                            for (int b = 0; b < 8; ++b) { const unsigned int c = sh.hist_mat[b]; sh.hist_mat[b] = run; run += c; }
                        }
                        __syncthreads();
                    }
                }
            }
            else
                __syncthreads();
            const unsigned int bin_div = (max(a.sc.n_tris, a.sc.n_spheres) + FRT_WF_BINS - 1) / FRT_WF_BINS;
            const unsigned int n_shade = by_material ? ((n_ext + 31u) & ~31u) : P;
            for (unsigned int s0 = 0; s0 < n_shade; s0 += THREADS)
            {
                const unsigned int idx = s0 + tid; // bounds are multiples of 32: whole warps fall off the end together
                if (idx >= n_shade) break;
                const bool valid = !by_material || idx < n_ext;
                const unsigned int slot = by_material ? (valid ? order[idx] : 0u) : idx;
                const unsigned int r = valid ? frt_shade_vertex<OTHER>(a, lp, slot, nonfinite) : 0u;
                const bool ended = r & FRT_SHADE_TERMINATED, goes_on = (r & FRT_SHADE_WAS_ALIVE) && !ended;
                const unsigned int fi = warp_append(ended, &sh.n_free);
                if (ended) free_q[fi] = slot;
                unsigned int k = 0xFFu;
                if (goes_on)
                {
                    k = min((r >> FRT_SHADE_PRIM_SHIFT) / max(bin_div, 1u), FRT_WF_BINS - 1u);
                    atomicAdd(&sh.hist_ext[k], 1u);
                    if (r & FRT_SHADE_SHADOW) { atomicAdd(&sh.hist_sh[k], 1u); k |= 0x80u; }
                }
                if (valid) key[slot] = (unsigned char)k;
            }
            __syncthreads();
            if (tid < 32)
            {
                // exclusive scan of both histograms by one warp (two bins per lane)
                unsigned int e0 = sh.hist_ext[2 * tid], e1 = sh.hist_ext[2 * tid + 1];
                unsigned int h0 = sh.hist_sh[2 * tid], h1 = sh.hist_sh[2 * tid + 1];
                unsigned int es = e0 + e1, hs = h0 + h1;
                for (int off = 1; off < 32; off <<= 1)
                {
                    const unsigned int ev = __shfl_up_sync(0xFFFFFFFFu, es, off), hv = __shfl_up_sync(0xFFFFFFFFu, hs, off);
                    if ((int)tid >= off) { es += ev; hs += hv; }
                }
                sh.hist_ext[2 * tid] = es - e0 - e1; sh.hist_ext[2 * tid + 1] = es - e1;
                sh.hist_sh[2 * tid] = hs - h0 - h1; sh.hist_sh[2 * tid + 1] = hs - h1;
                if (tid == 31) { sh.n_ext = es; sh.n_shadow = hs; }
            }
            __syncthreads();
            for (unsigned int slot = tid; slot < P; slot += THREADS)
            {
                const unsigned int k = key[slot];
                if (k == 0xFFu) continue;
                ext_q[atomicAdd(&sh.hist_ext[k & 0x7Fu], 1u)] = slot;
                if (k & 0x80u) shadow_q[atomicAdd(&sh.hist_sh[k & 0x7Fu], 1u)] = slot;
            }
        }
        else
        for (unsigned int s0 = 0; s0 < P; s0 += THREADS)
        {
            const unsigned int slot = s0 + tid; // P is a multiple of 32: whole warps fall off the end together
            if (slot >= P) break;
            const unsigned int r = frt_shade_vertex<OTHER>(a, lp, slot, nonfinite);
            const bool ended = r & FRT_SHADE_TERMINATED, goes_on = (r & FRT_SHADE_WAS_ALIVE) && !ended;
            const unsigned int ei = warp_append(goes_on, &sh.n_ext);
            if (goes_on) ext_q[ei] = slot;
            const unsigned int fi = warp_append(ended, &sh.n_free);
            if (ended) free_q[fi] = slot;
            const unsigned int si = warp_append(r & FRT_SHADE_SHADOW, &sh.n_shadow);
            if (r & FRT_SHADE_SHADOW) shadow_q[si] = slot;
        }
        __syncthreads();
        if (tid == 0) { const unsigned long long t = frt_globaltimer(); t_shade += t - t0; t0 = t; }
        ++n_iter;
        // the per-thread counters are 32-bit: fold them into the 64-bit device counters long before they can wrap
        if ((n_iter & 0xFFFFu) == 0u) wf_flush_counters<COUNT>(a.ctr, n_closest, n_shadow_rays, n_started, nonfinite, n_nodes, n_tris);
    }

    wf_flush_counters<COUNT>(a.ctr, n_closest, n_shadow_rays, n_started, nonfinite, n_nodes, n_tris);
    if (tid == 0)
    {
        atomicAdd(&a.ctr->stage_ns[0], t_gen);
        atomicAdd(&a.ctr->stage_ns[1], t_ext);
        atomicAdd(&a.ctr->stage_ns[2], t_shade);
        atomicAdd(&a.ctr->stage_ns[3], t_con);
        atomicMax(&a.ctr->wf_iterations, n_iter);
    }
}
