// Part of frt_core.cu: tile launch, film sum / readback / tone map, cancellation, ray-cast entry points.

namespace {

int ensure_film(frt_ctx *ctx, uint32_t nx, uint32_t ny)
{
    const size_t n = (size_t)nx * ny * 3;
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        if (d.film_floats != n)
        {
            if (d.d_film) cudaFree(d.d_film);
            if (d.d_film_stage) { cudaFree(d.d_film_stage); d.d_film_stage = nullptr; }
            d.d_film = nullptr;
            CU(cudaMalloc((void **)&d.d_film, n * sizeof(float)));
            d.film_floats = n;
            CU(cudaMemsetAsync(d.d_film, 0, n * sizeof(float), d.stream));
        }
    }
    ctx->nx = nx; ctx->ny = ny;
    return FRT_OK;
}

// ctx-owned pinned host staging of the film: sums (float) and the 8-bit image
int ensure_host_film(frt_ctx *ctx, size_t n)
{
    if (ctx->h_film_floats == n) return FRT_OK;
    if (ctx->h_film) { cudaFreeHost(ctx->h_film); ctx->h_film = nullptr; }
    if (ctx->h_ldr) { cudaFreeHost(ctx->h_ldr); ctx->h_ldr = nullptr; }
    ctx->h_film_floats = 0;
    CU(cudaSetDevice(ctx->dev[0].ordinal));
    if (ctx->d_ldr) { cudaFree(ctx->d_ldr); ctx->d_ldr = nullptr; }
    CU(cudaHostAlloc((void **)&ctx->h_film, n * sizeof(float), cudaHostAllocPortable));
    CU(cudaHostAlloc((void **)&ctx->h_ldr, n, cudaHostAllocPortable));
    CU(cudaMalloc((void **)&ctx->d_ldr, n));
    ctx->h_film_floats = n;
    return FRT_OK;
}

// Waits for every device's tiles, gathers the counters and leaves the SUM of the per-device films in device 0's film
// (stream-ordered on device 0's stream: the caller enqueues its copies behind it).
int finish_render(frt_ctx *ctx, frt_stats *out)
{
    frt_stats st;
    memset(&st, 0, sizeof st);
    Device &d0 = ctx->dev[0];
    const int nd = (int)ctx->dev.size();
    // the film sum first, so that it overlaps the counter copies: every device sums its range of all the films over
    // peer loads as soon as ALL renders are done, and stores it into device 0's film
    int n_rendered = 0;
    for (Device &d : ctx->dev) n_rendered += d.rendered ? 1 : 0;
    const bool need_sum = nd > 1 && d0.d_film && (n_rendered > 1 || (n_rendered == 1 && !d0.rendered));
    if (need_sum && ctx->peer_ok)
    {
        PeerFilms films;
        films.n = 0;
        for (Device &d : ctx->dev) films.p[films.n++] = d.d_film; // a device that rendered nothing holds zeros
        const size_t n4 = d0.film_floats / 4;
        for (int g = 0; g < nd; ++g)
        {
            Device &d = ctx->dev[g];
            CU(cudaSetDevice(d.ordinal));
            for (int k = 0; k < nd; ++k)
                if (k != g && ctx->dev[k].rendered) CU(cudaStreamWaitEvent(d.stream, ctx->dev[k].ev_end, 0));
            const size_t b4 = n4 * g / nd, e4 = n4 * (g + 1) / nd;
            const size_t tb = g == nd - 1 ? 4 * n4 : 0, te = g == nd - 1 ? d0.film_floats : 0;
            k_sum_slice<<<d.sms * 2, 256, 0, d.stream>>>(d0.d_film, films, b4, e4, tb, te);
            CU(cudaGetLastError());
            CU(cudaEventRecord(d.ev_sum, d.stream));
            d.launches += 1;
        }
        CU(cudaSetDevice(d0.ordinal));
        for (int g = 1; g < nd; ++g) CU(cudaStreamWaitEvent(d0.stream, ctx->dev[g].ev_sum, 0));
    }
    else if (need_sum)
    {
        // no peer access: the films are copied to device 0 one after the other and added there
        CU(cudaSetDevice(d0.ordinal));
        if (!d0.d_film_stage) CU(cudaMalloc((void **)&d0.d_film_stage, d0.film_floats * sizeof(float)));
        for (int g = 1; g < nd; ++g)
        {
            Device &d = ctx->dev[g];
            if (!d.rendered) continue;
            CU(cudaStreamWaitEvent(d0.stream, d.ev_end, 0));
            CU(cudaMemcpyPeerAsync(d0.d_film_stage, d0.ordinal, d.d_film, d.ordinal, d0.film_floats * sizeof(float), d0.stream));
            PeerFilms two;
            two.n = 2; two.p[0] = d0.d_film; two.p[1] = d0.d_film_stage;
            k_sum_slice<<<d0.sms * 2, 256, 0, d0.stream>>>(d0.d_film, two, 0, d0.film_floats / 4, d0.film_floats / 4 * 4, d0.film_floats);
            CU(cudaGetLastError());
            d0.launches += 1;
        }
    }
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaMemcpyAsync(d.h_ctr, d.d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
    }
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
        st.samples += d.h_ctr->samples_started;
        st.closest_rays += d.h_ctr->closest_rays;
        st.shadow_rays += d.h_ctr->shadow_rays;
        st.node_visits += d.h_ctr->node_visits;
        st.tri_tests += d.h_ctr->tri_tests;
        st.iterations = std::max<uint64_t>(st.iterations, d.h_ctr->wf_iterations);
        st.kernel_launches += d.launches;
        if (d.rendered)
        {
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
            st.render_ms = std::max<double>(st.render_ms, ms);
        }
        const double per_cta = d.wf_ctas_launched ? 1e-6 / d.wf_ctas_launched : 0.0; // ns summed over CTAs -> mean ms
        st.generate_ms = std::max(st.generate_ms, d.h_ctr->stage_ns[0] * per_cta);
        st.extend_ms = std::max(st.extend_ms, d.h_ctr->stage_ns[1] * per_cta);
        st.shade_ms = std::max(st.shade_ms, d.h_ctr->stage_ns[2] * per_cta);
        st.connect_ms = std::max(st.connect_ms, d.h_ctr->stage_ns[3] * per_cta);
    }
    if (need_sum)
    {
        // the partial sums are folded into device 0: clear the others so that a second readback (or further tiles)
        // does not count them twice
        for (int g = 1; g < nd; ++g)
        {
            Device &d = ctx->dev[g];
            CU(cudaSetDevice(d.ordinal));
            CU(cudaMemsetAsync(d.d_film, 0, d.film_floats * sizeof(float), d.stream));
            d.rendered = false;
        }
        d0.rendered = true;
        CU(cudaSetDevice(d0.ordinal));
    }
    ctx->stats = st;
    if (out) *out = st;
    return FRT_OK;
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
        d.launches = 0;
        d.wf_ctas_launched = 0;
        d.rendered = false;
    }
    memset(&ctx->stats, 0, sizeof ctx->stats);
    return FRT_OK;
}

int frt_render_tile(frt_ctx *ctx, const frt_tile *t)
{
    if (!ctx || !t) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_render_tile: no scene uploaded");
    // bounds without 32-bit wrap-around: x0 = 0xFFFFFFF0, w = 0x20 must not pass
    if (t->nx == 0 || t->ny == 0 || t->w == 0 || t->h == 0 || t->w > t->nx || t->x0 > t->nx - t->w || t->h > t->ny || t->y0 > t->ny - t->h)
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: tile outside the film");
    if ((uint64_t)t->nx * t->ny >= (1ull << 32) / 3)
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: film of %u x %u pixels is too large (pixel indices are 32-bit)", t->nx, t->ny);
    if (t->spp_end < t->spp_begin) return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: empty sample range");
    if (t->max_depth > 200) return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: max_depth too large");
    if (t->integrator > FRT_INTEGRATOR_DIRECT || t->reserved != 0)
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: unknown integrator %u", t->integrator);
    if (t->integrator == FRT_INTEGRATOR_AO && !(ctx->ao_t_max > 0.f))
        return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: ao needs frt_scene_desc::ao_t_max > 0");
    int rc = ensure_film(ctx, t->nx, t->ny);
    if (rc != FRT_OK) return rc;
    ctx->epoch += 1; // frt_cancel stops the tiles launched up to its call: later tiles carry a larger epoch

    const int nd = (int)ctx->dev.size();
    const uint32_t spp = t->spp_end - t->spp_begin;
    const bool count = ctx->flags & FRT_FLAG_COUNTERS;
    for (int i = 0; i < nd; ++i)
    {
        Device &d = ctx->dev[i];
        CU(cudaSetDevice(d.ordinal));
        // spp range split evenly over the devices
        const uint32_t s0 = t->spp_begin + (uint32_t)(((uint64_t)spp * i) / nd);
        const uint32_t s1 = t->spp_begin + (uint32_t)(((uint64_t)spp * (i + 1)) / nd);
        KernelArgs a;
        a.sc = t->integrator == FRT_INTEGRATOR_AO ? device_view(ctx, d, (double)ctx->ao_t_max) : device_view(ctx, d);
        a.tile.nx = t->nx; a.tile.ny = t->ny; a.tile.x0 = t->x0; a.tile.y0 = t->y0; a.tile.w = t->w; a.tile.h = t->h;
        a.tile.spp_begin = s0; a.tile.seed = t->seed; a.tile.max_depth = t->max_depth; a.tile.integrator = t->integrator;
        a.ctr = d.d_ctr;
        a.film = d.d_film;
        a.epoch = ctx->epoch;
        const unsigned long long total = (unsigned long long)t->w * t->h * (s1 - s0);
        if (total == 0) continue;
        if (!d.rendered) CU(cudaEventRecord(d.ev_begin, d.stream));
        d.rendered = true;
        // the whole tile is one persistent launch: path state lives in the CTAs' shared memory
        // the general kernel: the integrators next to `path`, and `path` itself over textured surfaces
        const bool other = t->integrator != FRT_INTEGRATOR_PATH || ctx->host_view.n_textures != 0;
        // the path kernel's launch shape follows the scene's size (frt_ctx::wf_config)
        const int shape = ctx->scene_bytes > ctx->wf_big_scene_bytes ? 1 : 0;
        const unsigned int slots = other ? ctx->wf_slots : ctx->wf_slots_of[shape];
        const size_t smem = (size_t)slots * FRT_WF_SLOT_BYTES;
        // no more CTAs than can be kept busy: one CTA-iteration consumes up to `slots` samples
        const unsigned long long want = (total + slots - 1) / slots;
        const WfConfig &wc = other ? wf_other : wf_configs[ctx->wf_config[shape]];
        const int resident = other ? d.blocks_wavefront_other : d.blocks_wavefront[shape];
        if (resident < 1) return fail(ctx, FRT_ERR_INVALID, "frt_render_tile: FRT_WF_SLOTS=%u leaves no room for this integrator's kernel", slots);
        const unsigned int grid = (unsigned int)std::min<unsigned long long>((unsigned long long)resident, want);
        k_reset_tile<<<1, 1, 0, d.stream>>>(d.d_ctr, total);
        wc.kernel[count ? 1 : 0]<<<grid, wc.threads, smem, d.stream>>>(a, slots, ctx->wf_refill,
                                                                               ctx->wf_sort >= 0 ? ctx->wf_sort : (ctx->multi_bsdf ? 3 : 1));
        CU(cudaGetLastError());
        CU(cudaEventRecord(d.ev_end, d.stream));
        d.launches += 2;
        d.wf_ctas_launched += grid;
    }
    return FRT_OK;
}

// viewer::to_exit (viewer.cpp:101-104) / path.cpp:127: the tiles launched so far stop taking samples and drop their
// paths in flight; frt_readback then returns whatever the film holds.  Callable from any thread.
int frt_cancel(frt_ctx *ctx)
{
    if (!ctx || !ctx->h_cancel) return FRT_ERR_INVALID;
    // The flag the kernels poll lives in DEVICE memory (a host-mapped flag costs every CTA a PCIe round trip per
    // iteration: measured 4x on the whole render); the request travels as a 4-byte asynchronous copy on a side
    // stream, which the copy engine runs while the persistent kernel occupies the SMs.
    *ctx->h_cancel = ctx->epoch;
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaMemcpyAsync(&d.d_ctr->cancel_epoch, ctx->h_cancel, sizeof(unsigned int), cudaMemcpyHostToDevice, d.side_stream));
    }
    return FRT_OK;
}

// Non-blocking: *done = 1 when every tile launched so far has finished on every device.
int frt_poll(frt_ctx *ctx, int *done)
{
    if (!ctx || !done) return FRT_ERR_INVALID;
    *done = 1;
    for (Device &d : ctx->dev)
    {
        if (!d.rendered) continue;
        const cudaError_t e = cudaEventQuery(d.ev_end);
        if (e == cudaErrorNotReady) { *done = 0; cudaGetLastError(); }
        else if (e != cudaSuccess) return fail(ctx, FRT_ERR_CUDA, "frt_poll: %s", cudaGetErrorString(e));
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
    int rc = finish_render(ctx, stats);
    if (rc != FRT_OK) return rc;
    if (!rgb_sum) { CU(cudaStreamSynchronize(d0.stream)); return FRT_OK; }
    // device -> pinned staging at link speed, then out to the caller's (pageable) buffer on several host threads
    if ((rc = ensure_host_film(ctx, d0.film_floats)) != FRT_OK) return rc;
    CU(cudaSetDevice(d0.ordinal));
    CU(cudaMemcpyAsync(ctx->h_film, d0.d_film, n_floats * sizeof(float), cudaMemcpyDeviceToHost, d0.stream));
    CU(cudaStreamSynchronize(d0.stream));
    const float *src = ctx->h_film;
    parallel_chunks(n_floats, 1u << 19, [=](size_t b, size_t e) { memcpy(rgb_sum + b, src + b, (e - b) * sizeof(float)); });
    return FRT_OK;
}

// Replaces the per-pixel viewer::add_sample calls of path.cpp:143 for the whole film (viewer.cpp:109-132): the mean
// c = sum * (1 / ns) goes to fout_image (double) and / or mean_rgb (float32), the tone-mapped
// int(pow(1 - exp(-c), 1 / 2.2) * 255 + .5) — evaluated in double on device 0 — to out_image.  Each of the three
// outputs may be NULL.  Layout nx * ny * 3, row 0 = bottom, like the viewer's buffers.
int frt_readback_film(frt_ctx *ctx, uint64_t ns, double *fout_image, unsigned char *out_image, float *mean_rgb, frt_stats *stats)
{
    if (!ctx || ns == 0) return FRT_ERR_INVALID;
    Device &d0 = ctx->dev[0];
    if (!d0.d_film) return fail(ctx, FRT_ERR_STATE, "frt_readback_film: nothing rendered");
    int rc = finish_render(ctx, stats);
    if (rc != FRT_OK) return rc;
    const size_t n = d0.film_floats;
    if ((rc = ensure_host_film(ctx, n)) != FRT_OK) return rc;
    CU(cudaSetDevice(d0.ordinal));
    const double inv_ns = 1.0 / (double)ns; // geometry.h:533-538: operator/= multiplies by the reciprocal
    if (out_image)
    {
        k_tonemap<<<d0.sms * 4, 256, 0, d0.stream>>>(d0.d_film, ctx->d_ldr, n, inv_ns);
        CU(cudaGetLastError());
        d0.launches += 1;
        ctx->stats.kernel_launches += 1;
        if (stats) stats->kernel_launches += 1;
        CU(cudaMemcpyAsync(ctx->h_ldr, ctx->d_ldr, n, cudaMemcpyDeviceToHost, d0.stream));
    }
    if (fout_image || mean_rgb) CU(cudaMemcpyAsync(ctx->h_film, d0.d_film, n * sizeof(float), cudaMemcpyDeviceToHost, d0.stream));
    CU(cudaStreamSynchronize(d0.stream));
    const float *sums = ctx->h_film;
    const unsigned char *ldr = ctx->h_ldr;
    parallel_chunks(n, 1u << 18, [=](size_t b, size_t e) {
        if (fout_image) for (size_t i = b; i < e; ++i) fout_image[i] = (double)sums[i] * inv_ns;
        if (mean_rgb) for (size_t i = b; i < e; ++i) mean_rgb[i] = (float)((double)sums[i] * inv_ns);
        if (out_image) memcpy(out_image + b, ldr + b, e - b);
    });
    return FRT_OK;
}

int frt_trace_rays_device(frt_ctx *ctx, int dev, const void *d_rays, size_t n, void *d_hits, int mode, double *device_ms)
{
    if (!ctx || dev < 0 || dev >= (int)ctx->dev.size()) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_trace_rays: no scene uploaded");
    if (n > 0xFFFFFF00ull) return fail(ctx, FRT_ERR_INVALID, "frt_trace_rays: too many rays in one call");
    if (mode & ~(FRT_TRACE_ANY_HIT | FRT_TRACE_SORT)) return fail(ctx, FRT_ERR_INVALID, "frt_trace_rays: unknown mode bits %d", mode);
    if (n == 0) { if (device_ms) *device_ms = 0; return FRT_OK; }
    Device &d = ctx->dev[dev];
    CU(cudaSetDevice(d.ordinal));
    const DevScene sc = device_view(ctx, d);
    const bool count = ctx->flags & FRT_FLAG_COUNTERS, any_hit = mode & FRT_TRACE_ANY_HIT;
    const float2 *r = (const float2 *)d_rays;
    frt_hit *h = (frt_hit *)d_hits;
    const uint32_t *order = nullptr;
    CU(cudaMemsetAsync(&d.d_ctr->trace_next, 0, sizeof(unsigned int), d.stream));
    if (mode & FRT_TRACE_SORT)
    {
        const size_t words = frt_sort_scratch_words(n);
        if (d.sort_cap < words)
        {
            if (d.d_sort) { cudaFree(d.d_sort); d.d_sort = nullptr; d.sort_cap = 0; }
            CU(cudaMalloc(&d.d_sort, words * sizeof(uint32_t)));
            d.sort_cap = words;
        }
    }
    CU(cudaEventRecord(d.ev_t0, d.stream));
    if (mode & FRT_TRACE_SORT) // the ordering is part of the call and of its device time
        CU(frt_sort_rays(r, (unsigned int)n, ctx->bounds_lo, ctx->bounds_hi, (uint32_t *)d.d_sort, &order, d.stream, &d.launches, ctx->sort_bits[0], ctx->sort_bits[1]));
    {
        const int a = any_hit ? 1 : 0, w8 = ctx->have_wide8 ? 1 : 0, occ = ctx->trace_occ[a];
        trace_kernels[w8][a][count ? 1 : 0][occ]<<<d.blocks_trace[w8][a][occ], 128, 0, d.stream>>>(sc, d.d_ctr, r, h, order, (unsigned int)n,
                                                                                             ctx->trace_refill[a]);
    }
    CU(cudaGetLastError());
    CU(cudaEventRecord(d.ev_t1, d.stream));
    d.launches += 1;
    if (device_ms) // NULL = asynchronous: the caller synchronises the stream (frt_trace_rays does, once, for all devices)
    {
        CU(cudaEventSynchronize(d.ev_t1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, d.ev_t0, d.ev_t1));
        *device_ms = ms;
    }
    return FRT_OK;
}

// The Morton-ordered cast with the sort's scratch in g_noalloc_sort_scratch: nothing here (or below it) allocates —
// stream-ordered memset, event records and kernel launches only.  A separate body on purpose: frt_trace_rays_device
// stays as it is.
int frt_trace_rays_device_noalloc(frt_ctx *ctx, int dev, const void *d_rays, size_t n, void *d_hits, int mode, double *device_ms)
{
    if (!ctx || dev < 0 || dev >= (int)ctx->dev.size()) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_trace_rays_device_noalloc: no scene uploaded");
    if (mode & ~(FRT_TRACE_ANY_HIT | FRT_TRACE_SORT)) return fail(ctx, FRT_ERR_INVALID, "frt_trace_rays_device_noalloc: unknown mode bits %d", mode);
    if (n > FRT_TRACE_NOALLOC_MAX_RAYS)
        return fail(ctx, FRT_ERR_INVALID, "frt_trace_rays_device_noalloc: %zu rays exceed the static scratch (FRT_TRACE_NOALLOC_MAX_RAYS = %u); "
                                          "split the batch or use frt_trace_rays_device", n, (unsigned int)FRT_TRACE_NOALLOC_MAX_RAYS);
    if (n && (!d_rays || !d_hits)) return fail(ctx, FRT_ERR_INVALID, "frt_trace_rays_device_noalloc: NULL buffer");
    if (n == 0) { if (device_ms) *device_ms = 0; return FRT_OK; }
    Device &d = ctx->dev[dev];
    if (!d.d_noalloc_scratch) return fail(ctx, FRT_ERR_STATE, "frt_trace_rays_device_noalloc: the scratch symbol was not resolved on device %d", d.ordinal);
    CU(cudaSetDevice(d.ordinal));
    const DevScene sc = device_view(ctx, d);
    const bool count = ctx->flags & FRT_FLAG_COUNTERS, any_hit = mode & FRT_TRACE_ANY_HIT;
    const uint32_t *order = nullptr;
    // the one word that must start at zero: the persistent kernel's work cursor.  Keys, indices and histograms are
    // written in full by each pass before they are read, so the previous call's contents do not matter.
    CU(cudaMemsetAsync(&d.d_ctr->trace_next, 0, sizeof(unsigned int), d.stream));
    CU(cudaEventRecord(d.ev_t0, d.stream));
    CU(frt_sort_rays((const float2 *)d_rays, (unsigned int)n, ctx->bounds_lo, ctx->bounds_hi, d.d_noalloc_scratch, &order, d.stream, &d.launches, ctx->sort_bits[0], ctx->sort_bits[1]));
    const int a = any_hit ? 1 : 0, w8 = ctx->have_wide8 ? 1 : 0, occ = ctx->trace_occ[a];
    trace_kernels[w8][a][count ? 1 : 0][occ]<<<d.blocks_trace[w8][a][occ], 128, 0, d.stream>>>(sc, d.d_ctr, (const float2 *)d_rays, (frt_hit *)d_hits, order,
                                                                                         (unsigned int)n, ctx->trace_refill[a]);
    CU(cudaGetLastError());
    CU(cudaEventRecord(d.ev_t1, d.stream));
    d.launches += 1;
    if (device_ms)
    {
        CU(cudaEventSynchronize(d.ev_t1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, d.ev_t0, d.ev_t1));
        *device_ms = ms;
    }
    return FRT_OK;
}

int frt_trace_rays(frt_ctx *ctx, const frt_ray *rays, size_t n, frt_hit *hits, int mode, double *device_ms)
{
    if (!ctx || (n && (!rays || !hits))) return FRT_ERR_INVALID;
    if (!ctx->have_scene) return fail(ctx, FRT_ERR_STATE, "frt_trace_rays: no scene uploaded");
    const int nd = (int)ctx->dev.size();
    if (device_ms) *device_ms = 0;
    if (n == 0) return FRT_OK;
    // Block partition over the devices (no exchange).  Every device's copy in, kernel and copy out are
    // enqueued on its own stream before anything is waited for, so the devices work at the same time.  Pinned caller
    // buffers (frt_host_alloc, cudaHostAlloc, cudaHostRegister) are copied asynchronously straight from / to the
    // caller; pageable ones are staged by the runtime inside cudaMemcpyAsync, which blocks the issuing thread, so
    // each device then gets its own host thread.
    struct Part { size_t b, e; int rc; size_t chunks; };
    std::vector<Part> parts(nd);
    for (int i = 0; i < nd; ++i) { parts[i].b = n * i / nd; parts[i].e = n * (i + 1) / nd; parts[i].rc = FRT_OK; parts[i].chunks = 0; }
    for (int i = 0; i < nd; ++i)
    {
        Device &d = ctx->dev[i];
        const size_t m = parts[i].e - parts[i].b;
        if (m == 0 || d.trace_cap >= m) continue;
        CU(cudaSetDevice(d.ordinal));
        if (d.d_rays) { cudaFree(d.d_rays); d.d_rays = nullptr; }
        if (d.d_hits) { cudaFree(d.d_hits); d.d_hits = nullptr; }
        d.trace_cap = 0;
        CU(cudaMalloc(&d.d_rays, m * sizeof(frt_ray)));
        CU(cudaMalloc(&d.d_hits, m * sizeof(frt_hit)));
        d.trace_cap = m;
    }
    cudaPointerAttributes pa_r{}, pa_h{};
    const bool pinned = cudaPointerGetAttributes(&pa_r, rays) == cudaSuccess && pa_r.type == cudaMemoryTypeHost &&
                        cudaPointerGetAttributes(&pa_h, hits) == cudaSuccess && pa_h.type == cudaMemoryTypeHost;
    cudaGetLastError();
    auto enqueue = [&](int i) {
        Device &d = ctx->dev[i];
        const size_t b = parts[i].b, m = parts[i].e - b;
        if (m == 0) return;
        if (cudaSetDevice(d.ordinal) != cudaSuccess || cudaMemcpyAsync(d.d_rays, rays + b, m * sizeof(frt_ray), cudaMemcpyHostToDevice, d.stream) != cudaSuccess)
        { parts[i].rc = FRT_ERR_CUDA; return; }
        parts[i].rc = frt_trace_rays_device(ctx, i, d.d_rays, m, d.d_hits, mode, nullptr);
        if (parts[i].rc != FRT_OK) return;
        if (cudaMemcpyAsync(hits + b, d.d_hits, m * sizeof(frt_hit), cudaMemcpyDeviceToHost, d.stream) != cudaSuccess) parts[i].rc = FRT_ERR_CUDA;
        if (!pinned && cudaStreamSynchronize(d.stream) != cudaSuccess) parts[i].rc = FRT_ERR_CUDA;
    };
    // Pinned buffers: a device's share moves in chunks on three streams, so that the copy in of chunk k+1 and the copy
    // out of chunk k-1 run under the cast of chunk k (PCIe both ways beside the kernel); only the first copy in and the
    // last copy out stay exposed.  Three events per chunk: rays arrived, cast begun, cast ended (the last two give the
    // chunk's device time).  FRT_TRACE_SORT orders every chunk by itself.
    const size_t chunk = ctx->trace_chunk;
    auto pipeline = [&](int i) {
        Device &d = ctx->dev[i];
        const size_t b = parts[i].b, m = parts[i].e - b;
        if (m == 0) return;
        const size_t nc = (m + chunk - 1) / chunk;
        bool ok = cudaSetDevice(d.ordinal) == cudaSuccess;
        while (ok && d.chunk_ev.size() < 3 * nc)
        {
            cudaEvent_t ev = nullptr;
            ok = cudaEventCreate(&ev) == cudaSuccess;
            if (ok) d.chunk_ev.push_back(ev);
        }
        // the copy streams must not start before earlier work on the main stream (a previous call's casts) is done
        ok = ok && cudaEventRecord(d.ev_t0, d.stream) == cudaSuccess && cudaStreamWaitEvent(d.in_stream, d.ev_t0, 0) == cudaSuccess;
        for (size_t c = 0; ok && c < nc; ++c)
        {
            const size_t o = c * chunk, k = std::min(chunk, m - o);
            cudaEvent_t arrived = d.chunk_ev[3 * c], begun = d.chunk_ev[3 * c + 1], ended = d.chunk_ev[3 * c + 2];
            ok = cudaMemcpyAsync((frt_ray *)d.d_rays + o, rays + b + o, k * sizeof(frt_ray), cudaMemcpyHostToDevice, d.in_stream) == cudaSuccess &&
                 cudaEventRecord(arrived, d.in_stream) == cudaSuccess && cudaStreamWaitEvent(d.stream, arrived, 0) == cudaSuccess &&
                 cudaEventRecord(begun, d.stream) == cudaSuccess;
            if (!ok) break;
            parts[i].rc = frt_trace_rays_device(ctx, i, (frt_ray *)d.d_rays + o, k, (frt_hit *)d.d_hits + o, mode, nullptr);
            if (parts[i].rc != FRT_OK) return;
            ok = cudaSetDevice(d.ordinal) == cudaSuccess && cudaEventRecord(ended, d.stream) == cudaSuccess &&
                 cudaStreamWaitEvent(d.out_stream, ended, 0) == cudaSuccess &&
                 cudaMemcpyAsync(hits + b + o, (frt_hit *)d.d_hits + o, k * sizeof(frt_hit), cudaMemcpyDeviceToHost, d.out_stream) == cudaSuccess;
        }
        if (!ok) parts[i].rc = FRT_ERR_CUDA;
        parts[i].chunks = nc;
    };
    if (pinned)
        for (int i = 0; i < nd; ++i) pipeline(i);
    else if (nd == 1)
        enqueue(0);
    else
    {
        std::vector<std::thread> workers;
        for (int i = 1; i < nd; ++i) workers.emplace_back(enqueue, i);
        enqueue(0);
        for (std::thread &t : workers) t.join();
    }
    double worst = 0;
    for (int i = 0; i < nd; ++i)
    {
        Device &d = ctx->dev[i];
        if (parts[i].e == parts[i].b) continue;
        if (parts[i].rc != FRT_OK) return parts[i].rc == FRT_ERR_CUDA ? fail(ctx, FRT_ERR_CUDA, "frt_trace_rays: device %d: %s", i, cudaGetErrorString(cudaGetLastError())) : parts[i].rc;
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
        double total = 0;
        if (parts[i].chunks)
        {
            CU(cudaStreamSynchronize(d.out_stream));
            for (size_t c = 0; c < parts[i].chunks; ++c)
            {
                float ms = 0.f;
                CU(cudaEventElapsedTime(&ms, d.chunk_ev[3 * c + 1], d.chunk_ev[3 * c + 2]));
                total += ms;
            }
        }
        else
        {
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, d.ev_t0, d.ev_t1));
            total = ms;
        }
        worst = std::max(worst, total);
    }
    if (device_ms) *device_ms = worst;
    return FRT_OK;
}

} // extern "C"
