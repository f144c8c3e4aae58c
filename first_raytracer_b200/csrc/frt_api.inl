// Part of frt_core.cu: context lifetime and scene upload (extern "C" entry points).

extern "C" {

const char *frt_last_error(const frt_ctx *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int frt_device_count(const frt_ctx *ctx) { return ctx ? (int)ctx->dev.size() : 0; }

int frt_set_flags(frt_ctx *ctx, uint32_t flags)
{
    if (!ctx) return FRT_ERR_INVALID;
    if (flags & ~FRT_FLAG_COUNTERS) return fail(ctx, FRT_ERR_INVALID, "frt_set_flags: unknown flag bits 0x%x", flags & ~FRT_FLAG_COUNTERS);
    ctx->flags = flags;
    return FRT_OK;
}

int frt_create(frt_ctx **out, const int *devices, int n)
{
    frt_ctx *ctx = nullptr; // for CU(): errors go to g_create_error
    if (!out || n < 1 || n > 16) return fail(nullptr, FRT_ERR_INVALID, "frt_create: bad arguments");
    *out = nullptr;
    int have = 0;
    cudaError_t e = cudaGetDeviceCount(&have);
    if (e != cudaSuccess || have == 0)
        return fail(nullptr, FRT_ERR_CUDA, "frt_create: no CUDA device (%s); this library has no CPU path",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    for (int i = 0; i < n; ++i)
    {
        const int o = devices ? devices[i] : i;
        if (o < 0 || o >= have) return fail(nullptr, FRT_ERR_INVALID, "frt_create: device %d not present (%d visible)", o, have);
    }
    // owned until the end: every error path below gives the streams, events and buffers back
    std::unique_ptr<frt_ctx, void (*)(frt_ctx *)> guard(new frt_ctx, frt_destroy);
    frt_ctx *c = guard.get();
    c->dev.resize(n);
    // FRT_WF_SLOTS / FRT_WF_CONFIG force ONE shape for every scene (the sweeps of scripts/wf_sweep.py)
    if (const char *s = getenv("FRT_WF_SLOTS"))
        c->wf_slots = c->wf_slots_of[0] = c->wf_slots_of[1] = (unsigned int)std::min(1568, std::max(32, atoi(s) / 32 * 32));
    if (const char *s = getenv("FRT_WF_CONFIG")) c->wf_config[0] = c->wf_config[1] = std::min(n_wf_configs - 1, std::max(0, atoi(s)));
    if (const char *s = getenv("FRT_WF_BIG_SCENE")) c->wf_big_scene_bytes = (size_t)std::max(0ll, atoll(s));
    if (const char *s = getenv("FRT_WF_SORT")) c->wf_sort = atoi(s) & 3;
    if (const char *s = getenv("FRT_WF_REFILL")) c->wf_refill = std::min(31, std::max(0, atoi(s)));
    if (const char *s = getenv("FRT_WIDE8")) c->wide8 = atoi(s) ? 1 : 0;
    // ray-cast tuning knobs (measured defaults; the sweeps of scripts/soup_sweep.py set them)
    if (const char *s = getenv("FRT_TRACE_OCC")) c->trace_occ[0] = std::min(3, std::max(0, atoi(s)));
    if (const char *s = getenv("FRT_TRACE_OCC_ANY")) c->trace_occ[1] = std::min(3, std::max(0, atoi(s)));
    if (const char *s = getenv("FRT_TRACE_REFILL")) c->trace_refill[0] = std::min(31, std::max(0, atoi(s)));
    if (const char *s = getenv("FRT_SORT_BITS"))
    {
        int o = 9, dd = 1;
        if ((sscanf(s, "%d,%d", &o, &dd) == 2 || sscanf(s, "%d:%d", &o, &dd) == 2) && o >= 1 && o <= 9 && dd >= 1 && dd <= 9 && 3 * (o + dd) <= 32) { c->sort_bits[0] = o; c->sort_bits[1] = dd; }
    }
    if (const char *s = getenv("FRT_TRACE_CHUNK")) c->trace_chunk = (size_t)std::max(1024ll, atoll(s));
    if (const char *s = getenv("FRT_TRACE_REFILL_ANY")) c->trace_refill[1] = std::min(31, std::max(-1, atoi(s)));

    CU(cudaHostAlloc((void **)&c->h_cancel, sizeof(unsigned int), cudaHostAllocPortable));
    *c->h_cancel = 0;
    for (int i = 0; i < n; ++i)
    {
        Device &d = c->dev[i];
        d.ordinal = devices ? devices[i] : i;
        CU(cudaSetDevice(d.ordinal));
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, d.ordinal));
        if (prop.major < 10)
            return fail(nullptr, FRT_ERR_CUDA, "frt_create: device %d is sm_%d%d; this library is built for sm_100a only",
                        d.ordinal, prop.major, prop.minor);
        d.sms = prop.multiProcessorCount;
        CU(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&d.side_stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&d.in_stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&d.out_stream, cudaStreamNonBlocking));
        for (cudaEvent_t *ev : {&d.ev_begin, &d.ev_end, &d.ev_sum, &d.ev_t0, &d.ev_t1}) CU(cudaEventCreate(ev));
        CU(cudaMalloc((void **)&d.d_ctr, sizeof(Counters)));
        CU(cudaMemset(d.d_ctr, 0, sizeof(Counters)));
        CU(cudaMallocHost((void **)&d.h_ctr, sizeof(Counters)));
        memset(d.h_ctr, 0, sizeof(Counters));
        // frt_trace_rays_device_noalloc: where the static scratch lives on THIS device
        CU(cudaGetSymbolAddress((void **)&d.d_noalloc_scratch, g_noalloc_sort_scratch));
        int per_sm = 0;
        for (int w8 = 0; w8 < 2; ++w8)
            for (int any = 0; any < 2; ++any)
                for (int occ = 0; occ < 4; ++occ)
                {
                    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_kernels[w8][any][0][occ], 128, 0));
                    d.blocks_trace[w8][any][occ] = std::max(1, per_sm) * d.sms;
                }
        for (int shape = 0; shape < 2; ++shape)
        {
            const size_t smem = (size_t)c->wf_slots_of[shape] * FRT_WF_SLOT_BYTES;
            const WfConfig &wc = wf_configs[c->wf_config[shape]];
            // the attribute is per kernel: when both shapes use one kernel it must cover the larger pool
            const size_t smem_max = c->wf_config[0] == c->wf_config[1]
                                        ? (size_t)std::max(c->wf_slots_of[0], c->wf_slots_of[1]) * FRT_WF_SLOT_BYTES : smem;
            for (int cnt = 0; cnt < 2; ++cnt)
                CU(cudaFuncSetAttribute(wc.kernel[cnt], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wc.kernel[0], wc.threads, smem));
            if (per_sm < 1)
                return fail(nullptr, FRT_ERR_CUDA, "frt_create: k_wavefront does not fit an SM with %u slots per CTA", c->wf_slots_of[shape]);
            d.blocks_wavefront[shape] = per_sm * d.sms;
        }
        const size_t wf_smem = (size_t)c->wf_slots * FRT_WF_SLOT_BYTES;
        for (int cnt = 0; cnt < 2; ++cnt)
            CU(cudaFuncSetAttribute(wf_other.kernel[cnt], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wf_smem));
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wf_other.kernel[0], wf_other.threads, wf_smem));
        d.blocks_wavefront_other = per_sm * d.sms; // 0: only the path integrator fits this FRT_WF_SLOTS
    }
    // peer access between every pair of the context's devices: the film reduce-scatter of frt_readback reads
    // every device's film from every device (the same ordinal listed twice needs none)
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j)
        {
            if (c->dev[i].ordinal == c->dev[j].ordinal) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, c->dev[i].ordinal, c->dev[j].ordinal);
            if (!can) { c->peer_ok = false; continue; }
            cudaSetDevice(c->dev[i].ordinal);
            cudaError_t pe = cudaDeviceEnablePeerAccess(c->dev[j].ordinal, 0);
            if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) c->peer_ok = false;
            cudaGetLastError();
        }
    if (getenv("FRT_NO_PEER")) c->peer_ok = false; // exercise the staged fallback
    *out = guard.release();
    return FRT_OK;
}

void frt_destroy(frt_ctx *ctx)
{
    if (!ctx) return;
    for (Device &d : ctx->dev)
    {
        cudaSetDevice(d.ordinal);
        if (d.stream) cudaStreamSynchronize(d.stream);
        void *ptrs[] = {d.d_nodes, d.d_tris, d.d_normals, d.d_spheres, d.d_mats, d.d_lights, d.d_env, d.d_ctr, d.d_film, d.d_film_stage,
                        d.d_textures, d.d_texels, d.d_tri_uv, d.d_nodes8, d.d_rays, d.d_hits, d.d_sort};
        for (void *p : ptrs) if (p) cudaFree(p);
        if (d.h_ctr) cudaFreeHost(d.h_ctr);
        for (cudaEvent_t ev : {d.ev_begin, d.ev_end, d.ev_sum, d.ev_t0, d.ev_t1}) if (ev) cudaEventDestroy(ev);
        if (d.stream) cudaStreamDestroy(d.stream);
        if (d.side_stream) cudaStreamDestroy(d.side_stream);
        if (d.in_stream) cudaStreamDestroy(d.in_stream);
        if (d.out_stream) cudaStreamDestroy(d.out_stream);
        for (cudaEvent_t ev : d.chunk_ev) cudaEventDestroy(ev);
    }
    if (!ctx->dev.empty()) cudaSetDevice(ctx->dev[0].ordinal);
    if (ctx->d_ldr) cudaFree(ctx->d_ldr);
    if (ctx->h_film) cudaFreeHost(ctx->h_film);
    if (ctx->h_ldr) cudaFreeHost(ctx->h_ldr);
    if (ctx->h_cancel) cudaFreeHost(ctx->h_cancel);
    delete ctx;
}

// Pinned host memory for callers that want their ray / hit / film buffers to move at link speed and asynchronously
// (frt_trace_rays recognises such pointers and skips its staging).
void *frt_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void frt_host_free(void *p) { if (p) cudaFreeHost(p); }

// The 8-wide compressed nodes frt_upload_scene derives for large scenes, as a host utility (no device needed): lets
// the CPU tests check the conversion.  Writes up to `cap_words` uint32 words (24 per node) and the node count.
int frt_build_wide8(const frt_scene_desc *s, uint32_t *words, size_t cap_words, uint32_t *n_nodes8, int *depth)
{
    if (!s || !n_nodes8 || (s->n_nodes && !s->nodes)) return FRT_ERR_INVALID;
    std::vector<uint32_t> w;
    float m = 0.f;
    for (uint32_t i = 0; i < s->n_nodes; ++i)
        for (int k = 0; k < 4; ++k)
            if (s->nodes[i].child[k] != FRT_CHILD_EMPTY)
                for (const float v : {s->nodes[i].lo_x[k], s->nodes[i].hi_x[k], s->nodes[i].lo_y[k], s->nodes[i].hi_y[k], s->nodes[i].lo_z[k], s->nodes[i].hi_z[k]})
                    if (std::isfinite(v)) m = std::fmax(m, std::fabs(v));
    for (int a = 0; a < 3; ++a) m = std::fmax(m, std::fabs(s->camera.origin[a]));
    frt_wide8::build(s->nodes, s->n_nodes, 8e-7f * m, w, depth);
    *n_nodes8 = (uint32_t)(w.size() / 24);
    if (words)
    {
        if (cap_words < w.size()) return FRT_ERR_INVALID;
        memcpy(words, w.data(), w.size() * sizeof(uint32_t));
    }
    return FRT_OK;
}

static int validate_scene(frt_ctx *ctx, const frt_scene_desc *s)
{
    if (!s || s->struct_size != sizeof(frt_scene_desc)) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: struct_size mismatch");
    if (s->n_nodes && !s->nodes) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: nodes is NULL");
    if (s->n_tris && !s->tris) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: tris is NULL");
    if (s->n_spheres && !s->spheres) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: spheres is NULL");
    if (s->n_materials == 0 || !s->materials) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: no materials");
    if (s->n_lights && !s->lights) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: lights is NULL");
    if (s->env_image && (s->env_nx == 0 || s->env_ny == 0 || (uint64_t)s->env_nx * s->env_ny > (1ull << 28)))
        return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: environment image of %u x %u texels", s->env_nx, s->env_ny);
    if (s->n_tris > FRT_LEAF_FIRST_MASK || s->n_spheres > FRT_LEAF_FIRST_MASK)
        return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: too many primitives");
    if ((s->n_tris || s->n_spheres) && s->n_nodes == 0) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: primitives without nodes");
    std::vector<uint8_t> level(s->n_nodes, 1); // inner levels above and including node i (children follow parents)
    for (uint32_t i = 0; i < s->n_nodes; ++i)
        for (int k = 0; k < 4; ++k)
        {
            const uint32_t c = s->nodes[i].child[k];
            if (c == FRT_CHILD_EMPTY) continue;
            if (!(c & FRT_CHILD_LEAF) && c < s->n_nodes && c > i)
            {
                // the traversal pushes at most three children per inner level on a FRT_STACK_SIZE-entry stack
                if (3u * (level[i] + 1u) > FRT_STACK_SIZE)
                    return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: tree deeper than %d inner levels", FRT_STACK_SIZE / 3);
                level[c] = std::max<uint8_t>(level[c], (uint8_t)(level[i] + 1));
            }
            if (c & FRT_CHILD_LEAF)
            {
                const uint32_t first = c & FRT_LEAF_FIRST_MASK, count = ((c >> FRT_LEAF_COUNT_SHIFT) & 15u) + 1u;
                const uint32_t limit = (c & FRT_CHILD_SPHERE) ? s->n_spheres : s->n_tris;
                if (first + count > limit) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: node %u child %d leaf range out of bounds", i, k);
            }
            else if (c >= s->n_nodes || c <= i)
                return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: node %u child %d index %u invalid (children must follow parents)", i, k, c);
        }
    bool need_normals = false;
    for (uint32_t i = 0; i < s->n_tris; ++i)
    {
        uint32_t mat, flags;
        memcpy(&mat, &s->tris[3 * (size_t)i + 1].w, 4);
        memcpy(&flags, &s->tris[3 * (size_t)i + 2].w, 4);
        if (mat >= s->n_materials) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: triangle %u material %u out of range", i, mat);
        if (!(flags & FRT_TRI_GEOMETRY_NORMALS)) need_normals = true;
    }
    if (need_normals && !s->tri_normals) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: tri_normals is NULL but smooth triangles exist");
    for (uint32_t i = 0; i < s->n_spheres; ++i)
        if (s->spheres[i].material >= s->n_materials) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: sphere %u material out of range", i);
    for (uint32_t i = 0; i < s->n_materials; ++i)
        if (s->materials[i].type > FRT_MAT_DIFFUSE_LIGHT) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: material %u has unknown type", i);
    for (uint32_t i = 0; i < s->n_lights; ++i)
        if (s->lights[i].material >= s->n_materials || s->lights[i].kind > FRT_LIGHT_SPHERE)
            return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: light %u invalid", i);
    // surface textures: table entries, checker children and image extents must stay inside their arrays
    if (s->n_textures && !s->textures) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: textures is NULL");
    if (s->n_texels && !s->tex_texels) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: tex_texels is NULL");
    for (uint32_t i = 0; i < s->n_textures; ++i)
    {
        const frt_texture &t = s->textures[i];
        if (t.kind > FRT_TEX_IMAGE) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: texture %u has unknown kind", i);
        if (t.kind == FRT_TEX_CHECKER && (t.a >= s->n_textures || t.b >= s->n_textures || t.a == i || t.b == i))
            return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: checker texture %u refers to texture %u / %u", i, t.a, t.b);
        if (t.kind == FRT_TEX_IMAGE && (t.b == 0 || t.c == 0 || t.b > 65536 || t.c > 65536 || (uint64_t)t.a + (uint64_t)t.b * t.c > s->n_texels))
            return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: image texture %u (%u x %u at texel %u) outside tex_texels", i, t.b, t.c, t.a);
    }
    for (uint32_t i = 0; i < s->n_materials; ++i)
        for (int slot : {FRT_MAT_TEX_A, FRT_MAT_TEX_B})
        {
            uint32_t ref;
            memcpy(&ref, &s->materials[i].p[slot], 4);
            if (ref > s->n_textures) return fail(ctx, FRT_ERR_INVALID, "frt_upload_scene: material %u refers to texture %u of %u", i, ref - 1, s->n_textures);
        }
    return FRT_OK;
}

int frt_upload_scene(frt_ctx *ctx, const frt_scene_desc *s)
{
    if (!ctx) return FRT_ERR_INVALID;
    int rc = validate_scene(ctx, s);
    if (rc != FRT_OK) return rc;
    ctx->have_scene = false;
    DevScene &v = ctx->host_view;
    memset(&v, 0, sizeof v);
    v.n_nodes = s->n_nodes; v.n_tris = s->n_tris; v.n_spheres = s->n_spheres;
    v.n_materials = s->n_materials; v.n_lights = s->n_lights;
    v.cam = s->camera;
    memcpy(v.env, s->env_color, sizeof v.env);
    ctx->ao_t_max = s->ao_t_max;
    v.env_nx = s->env_image ? s->env_nx : 0; v.env_ny = s->env_image ? s->env_ny : 0;
    v.n_textures = s->n_textures;
    for (Device &d : ctx->dev)
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
        if ((rc = upload_array(ctx, d, &d.d_nodes, s->nodes, s->n_nodes)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_tris, s->tris, 3 * (size_t)s->n_tris)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_normals, s->tri_normals, s->tri_normals ? 3 * (size_t)s->n_tris : 0)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_spheres, s->spheres, s->n_spheres)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_mats, s->materials, s->n_materials)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_lights, s->lights, s->n_lights)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_env, s->env_image, s->env_image ? 3 * (size_t)s->env_nx * s->env_ny : 0)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_textures, s->textures, s->n_textures)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_texels, s->tex_texels, 3 * (size_t)s->n_texels)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, d, &d.d_tri_uv, s->tri_uv, (s->tri_uv && s->n_textures) ? 6 * (size_t)s->n_tris : 0)) != FRT_OK) return rc;
    }
    for (Device &d : ctx->dev) // the caller's arrays may go away after the call
    {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
    }
    // scene box (children of the root): the Morton keys of frt_trace_rays' sort are taken inside it
    for (int a = 0; a < 3; ++a) { ctx->bounds_lo[a] = 0.f; ctx->bounds_hi[a] = 0.f; }
    if (s->n_nodes)
    {
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        const frt_bvh4_node &r = s->nodes[0];
        for (int k = 0; k < 4; ++k)
        {
            if (r.child[k] == FRT_CHILD_EMPTY) continue;
            lo[0] = std::fmin(lo[0], r.lo_x[k]); lo[1] = std::fmin(lo[1], r.lo_y[k]); lo[2] = std::fmin(lo[2], r.lo_z[k]);
            hi[0] = std::fmax(hi[0], r.hi_x[k]); hi[1] = std::fmax(hi[1], r.hi_y[k]); hi[2] = std::fmax(hi[2], r.hi_z[k]);
        }
        for (int a = 0; a < 3; ++a) if (std::isfinite(lo[a]) && std::isfinite(hi[a])) { ctx->bounds_lo[a] = lo[a]; ctx->bounds_hi[a] = hi[a]; }
    }
    // large scenes: ray casting walks 8-wide compressed nodes derived from the 4-wide ones (frt_wide8.cuh)
    ctx->have_wide8 = false;
    ctx->n_nodes8 = 0;
    // Opt-in (FRT_WIDE8=1): on the 10^7-triangle soup the 8-wide walk fetches 25 % fewer nodes at 3/4 of the bytes each but
    // measured 15 % slower than the 4-wide one (492 vs 589 Mrays/s closest hit; profiles/r02_soup_sweep.log) — the byte
    // decode and the 8-key sort cost more issue slots than the shorter dependent chain saves.
    const bool want8 = ctx->wide8 > 0;
    if (want8 && s->n_nodes)
    {
        std::vector<uint32_t> words;
        uint32_t n8 = 0;
        int depth8 = 0;
        float m = 0.f;
        for (int a = 0; a < 3; ++a) m = std::fmax(m, std::fmax(std::fabs(ctx->bounds_lo[a]), std::fabs(ctx->bounds_hi[a])));
        for (int a = 0; a < 3; ++a) m = std::fmax(m, std::fabs(s->camera.origin[a]));
        frt_wide8::build(s->nodes, s->n_nodes, 8e-7f * m, words, &depth8);
        n8 = (uint32_t)(words.size() / 24);
        // FRT_WIDE8_STRIDE=8: every node on its own 128 B line (a packed 96 B node straddles two lines half of the time)
        const uint32_t stride = (getenv("FRT_WIDE8_STRIDE") && atoi(getenv("FRT_WIDE8_STRIDE")) == 8) ? 8u : 6u;
        if (stride == 8)
        {
            std::vector<uint32_t> padded((size_t)n8 * 32, 0u);
            for (uint32_t i = 0; i < n8; ++i) memcpy(&padded[(size_t)i * 32], &words[(size_t)i * 24], 96);
            words.swap(padded);
        }
        v.node8_stride = stride;
        if (7 * depth8 + 1 <= FRT_STACK8_SIZE)
        {
            for (Device &d : ctx->dev)
            {
                CU(cudaSetDevice(d.ordinal));
                if ((rc = upload_array(ctx, d, &d.d_nodes8, words.data(), words.size())) != FRT_OK) return rc;
            }
            for (Device &d : ctx->dev) // `words` is a local
            {
                CU(cudaSetDevice(d.ordinal));
                CU(cudaStreamSynchronize(d.stream));
            }
            ctx->have_wide8 = true;
            ctx->n_nodes8 = n8;
        }
    }
    if (!ctx->have_wide8)
        for (Device &d : ctx->dev)
            if (d.d_nodes8) { cudaSetDevice(d.ordinal); cudaFree(d.d_nodes8); d.d_nodes8 = nullptr; }
    uint32_t bsdf_types = 0;
    for (uint32_t i = 0; i < s->n_materials; ++i)
        if (s->materials[i].type != FRT_MAT_DIFFUSE_LIGHT) bsdf_types |= 1u << s->materials[i].type;
    ctx->multi_bsdf = (bsdf_types & (bsdf_types - 1)) != 0;
    ctx->scene_bytes = (size_t)s->n_nodes * sizeof(frt_bvh4_node) + (size_t)s->n_tris * 48 + (size_t)s->n_spheres * sizeof(frt_sphere);
    ctx->have_scene = true;
    return FRT_OK;
}

} // extern "C"
