// Part of frt_core.cu: wavefront driver loop, film readback, ray-cast entry points.

namespace {

void timer_mark(Device &d, int kind, bool begin)
{
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, d.stream);
    d.timers.ev.push_back(e);
    if (begin) d.timers.kind.push_back(kind);
}

void timer_collect(Device &d)
{
    for (size_t i = 0; i + 1 < d.timers.ev.size(); i += 2)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, d.timers.ev[i], d.timers.ev[i + 1]);
        d.timers.ms[d.timers.kind[i / 2]] += ms;
        cudaEventDestroy(d.timers.ev[i]);
        cudaEventDestroy(d.timers.ev[i + 1]);
    }
    d.timers.ev.clear();
    d.timers.kind.clear();
}

int ensure_film(frt_ctx *ctx, uint32_t nx, uint32_t ny)
{
    const size_t n = (size_t)nx * ny * 3;
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        if (d.film_floats != n)
        {
            if (d.d_film) cudaFree(d.d_film);
            if (d.d_film_peer) { cudaFree(d.d_film_peer); d.d_film_peer = nullptr; }
            d.d_film = nullptr;
            CU(cudaMalloc((void **)&d.d_film, n * sizeof(float)));
            d.film_floats = n;
            CU(cudaMemsetAsync(d.d_film, 0, n * sizeof(float), d.stream));
        }
    }
    ctx->nx = nx; ctx->ny = ny;
    return FRT_OK;
}

// One wavefront iteration on one device: generate → extend → shade → connect.
template <bool COUNT>
void launch_iteration(frt_ctx *ctx, Device &d, const KernelArgs &a)
{
    const int parity = (int)(d.tile_iter & 1ull);
    const unsigned int n = d.pool.n_slots;
    const bool timed = ctx->flags & FRT_FLAG_KERNEL_TIMERS;
    const int gs_blocks = std::min<unsigned int>((n + 255u) / 256u, (unsigned int)d.sms * 8u);
    const int sh_blocks = std::min<unsigned int>((n + 127u) / 128u, (unsigned int)d.sms * 16u);
    if (timed) timer_mark(d, 0, true);
    k_generate<<<gs_blocks, 256, 0, d.stream>>>(a, parity);
    if (timed) { timer_mark(d, 0, false); timer_mark(d, 1, true); }
    k_extend<COUNT><<<d.blocks_extend, 128, 0, d.stream>>>(a);
    if (timed) { timer_mark(d, 1, false); timer_mark(d, 2, true); }
    k_shade<<<sh_blocks, 128, 0, d.stream>>>(a, parity ^ 1);
    if (timed) { timer_mark(d, 2, false); timer_mark(d, 3, true); }
    k_connect<COUNT><<<d.blocks_connect, 128, 0, d.stream>>>(a);
    if (timed) timer_mark(d, 3, false);
    d.iterations += 1;
    d.tile_iter += 1;
    d.launches += 4;
}

} // namespace

extern "C" {

int frt_clear_film(frt_ctx *ctx)
{
    if (!ctx) return FRT_ERR_INVALID;
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        if (d.d_film) CU(cudaMemsetAsync(d.d_film, 0, d.film_floats * sizeof(float), d.stream));
        CU(cudaMemsetAsync(d.d_ctr, 0, sizeof(Counters), d.stream));
        d.iterations = d.launches = 0;
        d.wf_ctas_launched = 0;
        d.rendered = false;
        for (double &m : d.timers.ms) m = 0;
    }
    memset(&ctx->stats, 0, sizeof ctx->stats);
    return FRT_OK;
}

int frt_render_tile(frt_ctx *ctx, const frt_tile *t)
{
    if (!ctx || !t) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_render_tile: no scene uploaded");
    if (t->nx == 0 || t->ny == 0 || t->w == 0 || t->h == 0 || t->x0 + t->w > t->nx || t->y0 + t->h > t->ny)
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: tile outside the film");
    if (t->spp_end < t->spp_begin) return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: empty sample range");
    if (t->max_depth > 200) return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: max_depth too large");
    if (t->integrator > FRT_INTEGRATOR_DIRECT || t->reserved != 0)
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: unknown integrator %u", t->integrator);
    if (t->integrator == FRT_INTEGRATOR_AO && !(ctx->ao_t_max > 0.f))
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: ao needs frt_scene_desc::ao_t_max > 0");
    if (t->integrator != FRT_INTEGRATOR_PATH && (ctx->flags & FRT_FLAG_STREAMED_POOL))
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: the streamed pool runs the path integrator only");
    int rc = ensure_film(ctx, t->nx, t->ny);
    if (rc != FRT_OK) return rc;

    const int nd = (int)ctx->dev.size();
    const uint32_t spp = t->spp_end - t->spp_begin;
    const bool streamed = ctx->flags & FRT_FLAG_STREAMED_POOL;
    const bool count = ctx->flags & FRT_FLAG_COUNTERS;
    std::vector<KernelArgs> args(nd);
    for (int i = 0; i < nd; ++i)
    {
        Device &d = ctx->dev[i];
        CU(cudaSetDevice(d.ordinal));
        // spp range split evenly over the devices
        const uint32_t s0 = t->spp_begin + (uint32_t)(((uint64_t)spp * i) / nd);
        const uint32_t s1 = t->spp_begin + (uint32_t)(((uint64_t)spp * (i + 1)) / nd);
        KernelArgs &a = args[i];
        a.sc = t->integrator == FRT_INTEGRATOR_AO ? device_view(ctx, d, (double)ctx->ao_t_max) : device_view(ctx, d);
        a.pool = d.pool;
        a.tile.nx = t->nx; a.tile.ny = t->ny; a.tile.x0 = t->x0; a.tile.y0 = t->y0; a.tile.w = t->w; a.tile.h = t->h;
        a.tile.spp_begin = s0; a.tile.seed = t->seed; a.tile.max_depth = t->max_depth; a.tile.integrator = t->integrator;
        a.ctr = d.d_ctr;
        a.film = d.d_film;
        const unsigned long long total = (unsigned long long)t->w * t->h * (s1 - s0);
        d.done = total == 0;
        if (d.done) continue;
        if (!d.rendered) CU(cudaEventRecord(d.ev_begin, d.stream));
        d.rendered = true;
        if (!streamed)
        {
            // CTA-resident pool: the whole tile is one persistent launch; nothing to poll
            const size_t smem = (size_t)ctx->wf_slots * FRT_WF_SLOT_BYTES;
            // no more CTAs than can be kept busy: one CTA-iteration consumes up to wf_slots samples
            const unsigned long long want = (total + ctx->wf_slots - 1) / ctx->wf_slots;
            // the general kernel: the integrators next to `path`, and `path` itself over textured surfaces
            const bool other = t->integrator != FRT_INTEGRATOR_PATH || ctx->host_view.n_textures != 0;
            const WfConfig &wc = other ? wf_other : wf_configs[ctx->wf_config];
            const int resident = other ? d.blocks_wavefront_other : d.blocks_wavefront;
            if (resident < 1) return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: FRT_WF_SLOTS=%u leaves no room for this integrator's kernel", ctx->wf_slots);
            const unsigned int grid = (unsigned int)std::min<unsigned long long>((unsigned long long)resident, want);
            k_reset_tile<<<1, 1, 0, d.stream>>>(d.d_ctr, total);
            (count ? wc.counting : wc.plain)<<<grid, wc.threads, smem, d.stream>>>(a, ctx->wf_slots, ctx->wf_refill,
                                                                                     ctx->wf_sort >= 0 ? ctx->wf_sort : (ctx->multi_bsdf ? 3 : 1));
            CU(cudaGetLastError());
            CU(cudaEventRecord(d.ev_end, d.stream));
            d.launches += 2;
            d.wf_ctas_launched += grid;
            d.done = true;
            continue;
        }
        if (d.pool.n_slots != ctx->n_slots)
        {
            int prc = alloc_pool(ctx, d, ctx->n_slots);
            if (prc != FRT_OK) return prc;
            a.pool = d.pool;
        }
        d.tile_iter = 0; // queue parity restarts with the pool reset
        k_reset_pool<<<(d.pool.n_slots + 255) / 256, 256, 0, d.stream>>>(d.pool, d.d_ctr, total);
        d.launches += 1;
    }
    // streamed pool: lock-step issue across devices; a device is polled every poll_every iterations
    bool all_done = true;
    for (int i = 0; i < nd; ++i) all_done = all_done && ctx->dev[i].done;
    while (!all_done)
    {
        for (int it = 0; it < ctx->poll_every; ++it)
            for (int i = 0; i < nd; ++i)
            {
                Device &d = ctx->dev[i];
                if (d.done) continue;
                cudaSetDevice(d.ordinal);
                if (count) launch_iteration<true>(ctx, d, args[i]); else launch_iteration<false>(ctx, d, args[i]);
            }
        for (int i = 0; i < nd; ++i)
        {
            Device &d = ctx->dev[i];
            if (d.done) continue;
            CU(cudaSetDevice(d.ordinal));
            CU(cudaMemcpyAsync(d.h_ctr, d.d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
        }
        all_done = true;
        for (int i = 0; i < nd; ++i)
        {
            Device &d = ctx->dev[i];
            if (d.done) continue;
            CU(cudaSetDevice(d.ordinal));
            CU(cudaStreamSynchronize(d.stream));
            if (ctx->flags & FRT_FLAG_KERNEL_TIMERS) timer_collect(d);
            if (d.h_ctr->next_sample >= d.h_ctr->total_samples && d.h_ctr->n_alive == 0)
            {
                d.done = true;
                CU(cudaEventRecord(d.ev_end, d.stream));
            }
            else
                all_done = false;
        }
    }
    return FRT_OK;
}

int frt_restore_film(frt_ctx *ctx, uint32_t nx, uint32_t ny, const float *rgb_sum, size_t n_floats)
{
    if (!ctx || !rgb_sum || nx == 0 || ny == 0) return FRT_ERR_INVALID;
    if (n_floats != (size_t)nx * ny * 3)
        return fail(ctx, FRT_ERR_INVALID, "frt_restore_film: expected %zu floats, got %zu", (size_t)nx * ny * 3, n_floats);
    int rc = ensure_film(ctx, nx, ny);
    if (rc != FRT_OK) return rc;
    rc = frt_clear_film(ctx);
    if (rc != FRT_OK) return rc;
    Device &d0 = ctx->dev[0];
    CU(cudaSetDevice(d0.ordinal));
    CU(cudaMemcpyAsync(d0.d_film, rgb_sum, n_floats * sizeof(float), cudaMemcpyHostToDevice, d0.stream));
    CU(cudaStreamSynchronize(d0.stream)); // the caller's buffer may go away after the call
    return FRT_OK;
}

int frt_get_counters(frt_ctx *ctx, frt_stats *stats)
{
    if (!ctx || !stats) return FRT_ERR_INVALID;
    memset(stats, 0, sizeof *stats);
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaMemcpyAsync(d.h_ctr, d.d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        stats->samples += d.h_ctr->samples_started;
        stats->closest_rays += d.h_ctr->closest_rays;
        stats->shadow_rays += d.h_ctr->shadow_rays;
        stats->node_visits += d.h_ctr->node_visits;
        stats->tri_tests += d.h_ctr->tri_tests;
        stats->kernel_launches += d.launches;
    }
    return FRT_OK;
}

int frt_film_device_ptr(frt_ctx *ctx, int dev, void **ptr, size_t *n_floats)
{
    if (!ctx || dev < 0 || dev >= (int)ctx->dev.size() || !ptr) return FRT_ERR_INVALID;
    Device &d = ctx->dev[dev];
    if (!d.d_film) return fail(ctx, FRT_ERR_STATE, "frt_film_device_ptr: no film yet");
    CU(cudaSetDevice(d.ordinal));
    CU(cudaStreamSynchronize(d.stream));
    *ptr = d.d_film;
    if (n_floats) *n_floats = d.film_floats;
    return FRT_OK;
}

int frt_stream_handle(frt_ctx *ctx, int dev, void **stream)
{
    if (!ctx || dev < 0 || dev >= (int)ctx->dev.size() || !stream) return FRT_ERR_INVALID;
    *stream = (void *)ctx->dev[dev].stream;
    return FRT_OK;
}

int frt_readback(frt_ctx *ctx, float *rgb_sum, size_t n_floats, frt_stats *stats)
{
    if (!ctx) return FRT_ERR_INVALID;
    Device &d0 = ctx->dev[0];
    if (!d0.d_film) return fail(ctx, FRT_ERR_STATE, "frt_readback: nothing rendered");
    if (rgb_sum && n_floats != d0.film_floats)
        return fail(ctx, FRT_ERR_INVALID, "frt_readback: expected %zu floats, got %zu", d0.film_floats, n_floats);
    frt_stats st;
    memset(&st, 0, sizeof st);
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaMemcpyAsync(d.h_ctr, d.d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        st.samples += d.h_ctr->samples_started;
        st.closest_rays += d.h_ctr->closest_rays;
        st.shadow_rays += d.h_ctr->shadow_rays;
        st.node_visits += d.h_ctr->node_visits;
        st.tri_tests += d.h_ctr->tri_tests;
        st.iterations = std::max<uint64_t>(st.iterations, d.iterations + d.h_ctr->wf_iterations);
        st.kernel_launches += d.launches;
        if (d.rendered)
        {
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
            st.render_ms = std::max<double>(st.render_ms, ms);
        }
        const double per_cta = d.wf_ctas_launched ? 1e-6 / d.wf_ctas_launched : 0.0; // ns summed over CTAs -> mean ms
        st.generate_ms = std::max(st.generate_ms, d.timers.ms[0] + d.h_ctr->stage_ns[0] * per_cta);
        st.extend_ms = std::max(st.extend_ms, d.timers.ms[1] + d.h_ctr->stage_ns[1] * per_cta);
        st.shade_ms = std::max(st.shade_ms, d.timers.ms[2] + d.h_ctr->stage_ns[2] * per_cta);
        st.connect_ms = std::max(st.connect_ms, d.timers.ms[3] + d.h_ctr->stage_ns[3] * per_cta);
    }
    // in-process multi-GPU: device 0 sums its peers' films (P2P loads over NVLink when enabled)
    if (ctx->dev.size() > 1)
    {
        CU(cudaSetDevice(d0.ordinal));
        PeerFilms peers;
        peers.n = 0;
        for (size_t i = 1; i < ctx->dev.size(); ++i)
        {
            Device &d = ctx->dev[i];
            if (!d.rendered) continue;
            if (ctx->peer_ok)
                peers.p[peers.n++] = d.d_film;
            else
            {
                if (!d0.d_film_peer) CU(cudaMalloc((void **)&d0.d_film_peer, d0.film_floats * sizeof(float)));
                CU(cudaMemcpyPeerAsync(d0.d_film_peer, d0.ordinal, d.d_film, d.ordinal, d0.film_floats * sizeof(float), d0.stream));
                PeerFilms one;
                one.n = 1; one.p[0] = d0.d_film_peer;
                k_sum_films<<<d0.sms * 4, 256, 0, d0.stream>>>(d0.d_film, one, d0.film_floats);
            }
        }
        if (peers.n) k_sum_films<<<d0.sms * 4, 256, 0, d0.stream>>>(d0.d_film, peers, d0.film_floats);
        CU(cudaStreamSynchronize(d0.stream));
        // the peers' partial sums are now folded into device 0: clear them so a second readback
        // (or further tiles) does not count them twice
        for (size_t i = 1; i < ctx->dev.size(); ++i)
        {
            Device &d = ctx->dev[i];
            CU(cudaSetDevice(d.ordinal));
            CU(cudaMemsetAsync(d.d_film, 0, d.film_floats * sizeof(float), d.stream));
        }
        CU(cudaSetDevice(d0.ordinal));
    }
    if (rgb_sum) CU(cudaMemcpy(rgb_sum, d0.d_film, n_floats * sizeof(float), cudaMemcpyDeviceToHost));
    ctx->stats = st;
    if (stats) *stats = st;
    return FRT_OK;
}

int frt_trace_rays_device(frt_ctx *ctx, int dev, const void *d_rays, size_t n, void *d_hits, int any_hit, double *device_ms)
{
    if (!ctx || dev < 0 || dev >= (int)ctx->dev.size()) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_trace_rays: no scene uploaded");
    if (n > 0xFFFFFF00ull) return fail(ctx, FRT_ERR_INVALID, "frt_trace_rays: too many rays in one call");
    if (n == 0) { if (device_ms) *device_ms = 0; return FRT_OK; }
    Device &d = ctx->dev[dev];
    CU(cudaSetDevice(d.ordinal));
    const DevScene sc = device_view(ctx, d);
    const bool count = ctx->flags & FRT_FLAG_COUNTERS;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    CU(cudaMemsetAsync(&d.d_ctr->trace_next, 0, sizeof(unsigned int), d.stream));
    CU(cudaEventRecord(e0, d.stream));
    const float2 *r = (const float2 *)d_rays;
    frt_hit *h = (frt_hit *)d_hits;
    if (any_hit)
    {
        if (count) k_trace<true, true><<<d.blocks_trace_any, 128, 0, d.stream>>>(sc, d.d_ctr, r, h, (unsigned int)n, ctx->wf_refill);
        else k_trace<true, false><<<d.blocks_trace_any, 128, 0, d.stream>>>(sc, d.d_ctr, r, h, (unsigned int)n, ctx->wf_refill);
    }
    else
    {
        if (count) k_trace<false, true><<<d.blocks_trace, 128, 0, d.stream>>>(sc, d.d_ctr, r, h, (unsigned int)n, ctx->wf_refill);
        else k_trace<false, false><<<d.blocks_trace, 128, 0, d.stream>>>(sc, d.d_ctr, r, h, (unsigned int)n, ctx->wf_refill);
    }
    CU(cudaEventRecord(e1, d.stream));
    CU(cudaEventSynchronize(e1));
    CU(cudaGetLastError());
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    d.launches += 1;
    if (device_ms) *device_ms = ms;
    return FRT_OK;
}

int frt_trace_rays(frt_ctx *ctx, const frt_ray *rays, size_t n, frt_hit *hits, int any_hit, double *device_ms)
{
    if (!ctx || (n && (!rays || !hits))) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_trace_rays: no scene uploaded");
    const int nd = (int)ctx->dev.size();
    double worst = 0;
    // block partition over devices (no exchange)
    for (int i = 0; i < nd; ++i)
    {
        const size_t b = n * i / nd, e = n * (i + 1) / nd;
        if (e == b) continue;
        Device &d = ctx->dev[i];
        CU(cudaSetDevice(d.ordinal));
        void *dr = nullptr, *dh = nullptr;
        CU(cudaMalloc(&dr, (e - b) * sizeof(frt_ray)));
        CU(cudaMalloc(&dh, (e - b) * sizeof(frt_hit)));
        CU(cudaMemcpyAsync(dr, rays + b, (e - b) * sizeof(frt_ray), cudaMemcpyHostToDevice, d.stream));
        double ms = 0;
        int rc = frt_trace_rays_device(ctx, i, dr, e - b, dh, any_hit, &ms);
        if (rc == FRT_OK)
        {
            cudaError_t ce = cudaMemcpyAsync(hits + b, dh, (e - b) * sizeof(frt_hit), cudaMemcpyDeviceToHost, d.stream);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(d.stream);
            if (ce != cudaSuccess) rc = fail(ctx, FRT_ERR_CUDA, "frt_trace_rays: copy back failed: %s", cudaGetErrorString(ce));
        }
        cudaFree(dr);
        cudaFree(dh);
        if (rc != FRT_OK) return rc;
        worst = std::max(worst, ms);
    }
    if (device_ms) *device_ms = worst;
    return FRT_OK;
}

} // extern "C"
