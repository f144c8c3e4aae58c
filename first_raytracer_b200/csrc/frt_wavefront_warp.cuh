// Warp-autonomous wavefront: the same generate -> extend / connect -> shade pipeline as frt_wavefront.cuh, but every
// WARP owns its slice of the CTA's shared-memory pool (P / warps slots) and runs the pipeline on it by itself — no
// CTA-wide barrier anywhere.  Why: with one pool per CTA every stage ends at a __syncthreads, so the 16 warps of an SM
// wait for the slowest ray of the stage (ncu, teapot scene: barrier 3.46 stall cycles per issued instruction, 9 of 32
// lanes active in the traversal; Cornell: 1.09).  Here a long ray holds back only the 31 other lanes of its own warp
// while the other warps of the SM keep issuing, and because the warps drift apart the latency-bound shade stage of one
// overlaps the issue-bound trace stage of another on the same scheduler.
// Samples are handed to a warp in chunks of consecutive indices (one global atomic per chunk), queues are compacted
// with warp ballots only (no shared-memory atomics), paths are shaded grouped by the material they hit when the scene
// has several BSDFs.
#pragma once
#include "frt_wavefront.cuh"

template <bool COUNT, int THREADS, int MIN_BLOCKS, bool OTHER>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) k_wavefront_warp(const KernelArgs a, const unsigned int P, const int refill_below,
                                                                        const int sort_rays)
{
    extern __shared__ float4 wf_smem[];
    constexpr unsigned int N_WARPS = THREADS / 32;
    __shared__ unsigned int cursors[N_WARPS][2];
    PathPool lp;
    lp.n_slots = P;
    lp.ray_o = wf_smem; lp.ray_d = lp.ray_o + P; lp.hit = lp.ray_d + P; lp.thr = lp.hit + P; lp.rad = lp.thr + P;
    lp.prev_p = lp.rad + P; lp.sh_a = lp.prev_p + P; lp.sh_b = lp.sh_a + P;
    lp.sh_c = reinterpret_cast<float *>(lp.sh_b + P);
    const unsigned int lane = lane_id(), warp = threadIdx.x >> 5;
    const unsigned int S = (P / N_WARPS) & ~31u, slot0 = warp * S; // this warp's slots: [slot0, slot0 + S)
    unsigned int *q_base = reinterpret_cast<unsigned int *>(lp.sh_c + P);
    unsigned int *ext_q = q_base + slot0, *free_q = ext_q + P, *shadow_q = free_q + P;
    // shading order: the area the CTA-wide kernel keeps its sort keys and order in (2 + 2 bytes per slot behind the queues)
    unsigned short *order = reinterpret_cast<unsigned short *>(q_base + 3 * P + P / 2) + slot0;
    const unsigned int lt = (1u << lane) - 1u;

    for (unsigned int i = lane; i < S; i += 32)
    {
        lp.ray_d[slot0 + i] = make_float4(0.f, 0.f, 0.f, 0.f); // dead
        free_q[i] = slot0 + i;
    }
    __syncwarp();

    const unsigned long long total = a.ctr->total_samples;
    // consecutive sample indices in chunks: large enough to make the global atomic rare, small enough to balance the tail
    const unsigned long long all_warps = (unsigned long long)gridDim.x * N_WARPS;
    const unsigned int chunk = (unsigned int)min(2048ull, max(32ull, total / (all_warps * 8ull)));
    unsigned long long my_next = 0, my_end = 0;
    bool exhausted = false;
    unsigned int n_ext = 0, n_free = S, n_sh = 0; // warp-uniform queue lengths
    unsigned int n_closest = 0, n_shadow_rays = 0, n_nodes = 0, n_tris = 0, nonfinite = 0, n_started = 0, n_iter = 0;
    unsigned long long t_gen = 0, t_ext = 0, t_con = 0, t_shade = 0, t0 = 0;
    if (lane == 0) t0 = frt_globaltimer();
    const bool by_material = sort_rays & 2;

    while (S != 0)
    {
        // ---- generate: the freed slots take the next samples of this warp's chunk ----------------------------------
        // frt_cancel (viewer::to_exit, viewer.cpp:101-104; path.cpp:127): looked at every 8th iteration of every warp
        if ((n_iter & 7u) == 0u)
        {
            unsigned int cancelled = 0;
            if (lane == 0) cancelled = *(const volatile unsigned int *)&a.ctr->cancel_epoch >= a.epoch;
            if (__shfl_sync(0xFFFFFFFFu, cancelled, 0)) break;
        }
        const unsigned int want = n_free;
        unsigned int n_new = 0;
        unsigned long long base = 0;
        if (want && !exhausted)
        {
            if (my_next == my_end)
            {
                unsigned long long b = 0;
                if (lane == 0) b = atomicAdd(&a.ctr->next_sample, (unsigned long long)chunk);
                b = __shfl_sync(0xFFFFFFFFu, b, 0);
                if (b >= total) exhausted = true;
                else { my_next = b; my_end = min(b + chunk, total); }
            }
            n_new = (unsigned int)min((unsigned long long)want, my_end - my_next);
            base = my_next;
            my_next += n_new;
        }
        // free_q is a stack: the new paths take its top n_new entries, the others stay where they are
        for (unsigned int i = lane; i < n_new; i += 32)
        {
            const unsigned int slot = free_q[want - n_new + i];
            frt_generate_path(a, lp, slot, base + i);
            ext_q[n_ext + i] = slot; // bounce rays first (the long ones), camera rays after them
            ++n_started;
        }
        n_free = want - n_new;
        n_ext += n_new;
        if (n_ext == 0) break; // nothing in flight and no samples left: this warp is done
        if (lane == 0) { cursors[warp][0] = 0; cursors[warp][1] = 0; }
        __syncwarp();
        if (lane == 0) { const unsigned long long t = frt_globaltimer(); t_gen += t - t0; t0 = t; }

        // ---- extend + connect: this warp's two queues, dynamic fetch inside the warp ---------------------------------
        wf_trace_queue<false, COUNT>(a.sc, lp, ext_q, n_ext, &cursors[warp][0], refill_below, n_closest, n_nodes, n_tris);
        if (lane == 0) { const unsigned long long t = frt_globaltimer(); t_ext += t - t0; t0 = t; }
        wf_trace_queue<true, COUNT>(a.sc, lp, shadow_q, n_sh, &cursors[warp][1], refill_below, n_shadow_rays, n_nodes, n_tris);
        __syncwarp();
        if (lane == 0) { const unsigned long long t = frt_globaltimer(); t_con += t - t0; t0 = t; }

        // ---- shade + queue compaction (ballots only) ------------------------------------------------------------------
        const unsigned int n_live = n_ext, rounds = (n_live + 31u) >> 5;
        if (by_material)
        {
            // shading order grouped by the material the path hit, so that the warp runs one BSDF's code at a time:
            // a stable counting sort over 8 keys, one ballot per key and round
            unsigned int run = 0;
            for (unsigned int key = 0; key < 8u; ++key)
                for (unsigned int r = 0; r < rounds; ++r)
                {
                    const unsigned int i = (r << 5) + lane;
                    unsigned int k = 0xFFu, slot = 0;
                    if (i < n_live)
                    {
                        slot = ext_q[i];
                        const unsigned int prim = __float_as_uint(lp.hit[slot].w);
                        k = 0;
                        if (prim != 0xFFFFFFFFu)
                        {
                            const unsigned int idx = prim & FRT_LEAF_FIRST_MASK;
                            const unsigned int mat = (prim & FRT_CHILD_SPHERE) ? __float_as_uint(__ldg(a.sc.spheres + 2 * idx + 1).x)
                                                                               : __float_as_uint(__ldg(a.sc.tris + 3 * (size_t)idx + 1).w);
                            k = 1u + min(a.sc.materials[mat].type, 6u);
                        }
                    }
                    const unsigned int m = __ballot_sync(0xFFFFFFFFu, k == key);
                    if (k == key) order[run + __popc(m & lt)] = (unsigned short)slot;
                    run += __popc(m);
                }
            __syncwarp();
        }
        unsigned int new_ext = 0, new_sh = 0;
        for (unsigned int r = 0; r < rounds; ++r)
        {
            const unsigned int i = (r << 5) + lane;
            const bool valid = i < n_live;
            const unsigned int slot = valid ? (by_material ? (unsigned int)order[i] : ext_q[i]) : 0u;
            __syncwarp(); // every lane holds its slot before the queues are rewritten below
            const unsigned int res = valid ? frt_shade_vertex<OTHER>(a, lp, slot, nonfinite) : 0u;
            const bool ended = res & FRT_SHADE_TERMINATED, goes_on = (res & FRT_SHADE_WAS_ALIVE) && !ended, shadow = res & FRT_SHADE_SHADOW;
            const unsigned int m_on = __ballot_sync(0xFFFFFFFFu, goes_on), m_end = __ballot_sync(0xFFFFFFFFu, ended),
                               m_sh = __ballot_sync(0xFFFFFFFFu, shadow);
            if (goes_on) ext_q[new_ext + __popc(m_on & lt)] = slot;
            if (ended) free_q[n_free + __popc(m_end & lt)] = slot;
            if (shadow) shadow_q[new_sh + __popc(m_sh & lt)] = slot;
            new_ext += __popc(m_on); n_free += __popc(m_end); new_sh += __popc(m_sh);
        }
        n_ext = new_ext; n_sh = new_sh;
        __syncwarp();
        if (lane == 0) { const unsigned long long t = frt_globaltimer(); t_shade += t - t0; t0 = t; }
        ++n_iter;
        if ((n_iter & 0xFFFFu) == 0u) wf_flush_counters<COUNT>(a.ctr, n_closest, n_shadow_rays, n_started, nonfinite, n_nodes, n_tris);
    }

    wf_flush_counters<COUNT>(a.ctr, n_closest, n_shadow_rays, n_started, nonfinite, n_nodes, n_tris);
    if (lane == 0)
    {
        // stage times are reported as the mean over CTAs: every warp adds its share
        atomicAdd(&a.ctr->stage_ns[0], t_gen / N_WARPS);
        atomicAdd(&a.ctr->stage_ns[1], t_ext / N_WARPS);
        atomicAdd(&a.ctr->stage_ns[2], t_shade / N_WARPS);
        atomicAdd(&a.ctr->stage_ns[3], t_con / N_WARPS);
        atomicMax(&a.ctr->wf_iterations, n_iter);
    }
}
