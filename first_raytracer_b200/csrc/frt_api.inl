// Part of frt_core.cu: context lifetime and scene upload (extern "C" entry points).

extern "C" {

const char *frt_last_error(const frt_ctx *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int frt_device_count(const frt_ctx *ctx) { return ctx ? (int)ctx->dev.size() : 0; }

int frt_set_flags(frt_ctx *ctx, uint32_t flags)
{
    if (!ctx) return FRT_ERR_INVALID;
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
    frt_ctx *c = new frt_ctx;
    c->dev.resize(n);
    if (const char *s = getenv("FRT_POOL_SLOTS")) c->n_slots = (unsigned int)std::max(1024ll, atoll(s));
    if (const char *s = getenv("FRT_POLL_EVERY")) c->poll_every = std::max(1, atoi(s));
    if (const char *s = getenv("FRT_WF_SLOTS")) c->wf_slots = (unsigned int)std::min(1568, std::max(32, atoi(s) / 32 * 32));
    if (const char *s = getenv("FRT_WF_CONFIG")) c->wf_config = std::min(n_wf_configs - 1, std::max(0, atoi(s)));
    if (const char *s = getenv("FRT_WF_SORT")) c->wf_sort = atoi(s) & 3;
    if (const char *s = getenv("FRT_WF_REFILL")) c->wf_refill = std::min(31, std::max(0, atoi(s)));
    for (int i = 0; i < n; ++i)
    {
        Device &d = c->dev[i];
        d.ordinal = devices ? devices[i] : i;
        CU(cudaSetDevice(d.ordinal));
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, d.ordinal));
        if (prop.major < 10)
        {
            delete c;
            return fail(nullptr, FRT_ERR_CUDA, "frt_create: device %d is sm_%d%d; this library is built for sm_100a only",
                        d.ordinal, prop.major, prop.minor);
        }
        d.sms = prop.multiProcessorCount;
        CU(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
        CU(cudaEventCreate(&d.ev_begin));
        CU(cudaEventCreate(&d.ev_end));
        CU(cudaMalloc((void **)&d.d_ctr, sizeof(Counters)));
        CU(cudaMemset(d.d_ctr, 0, sizeof(Counters)));
        CU(cudaMallocHost((void **)&d.h_ctr, sizeof(Counters)));
        memset(d.h_ctr, 0, sizeof(Counters));
        int per_sm = 0;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_extend<false>, 128, 0));
        d.blocks_extend = std::max(1, per_sm) * d.sms;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_connect<false>, 128, 0));
        d.blocks_connect = std::max(1, per_sm) * d.sms;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace<false, false>, 128, 0));
        d.blocks_trace = std::max(1, per_sm) * d.sms;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace<true, false>, 128, 0));
        d.blocks_trace_any = std::max(1, per_sm) * d.sms;
        const size_t wf_smem = (size_t)c->wf_slots * FRT_WF_SLOT_BYTES;
        const WfConfig &wc = wf_configs[c->wf_config];
        CU(cudaFuncSetAttribute(wc.plain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wf_smem));
        CU(cudaFuncSetAttribute(wc.counting, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wf_smem));
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wc.plain, wc.threads, wf_smem));
        if (per_sm < 1)
        {
            delete c;
            return fail(nullptr, FRT_ERR_CUDA, "frt_create: k_wavefront does not fit an SM with %u slots per CTA", c->wf_slots);
        }
        d.blocks_wavefront = per_sm * d.sms;
        CU(cudaFuncSetAttribute(wf_other.plain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wf_smem));
        CU(cudaFuncSetAttribute(wf_other.counting, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wf_smem));
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wf_other.plain, wf_other.threads, wf_smem));
        d.blocks_wavefront_other = per_sm * d.sms; // 0: only the path integrator fits this FRT_WF_SLOTS
    }
    // peer access for the in-process film reduction
    for (int i = 1; i < n; ++i)
    {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, c->dev[0].ordinal, c->dev[i].ordinal);
        if (!can) { c->peer_ok = false; continue; }
        cudaSetDevice(c->dev[0].ordinal);
        cudaError_t pe = cudaDeviceEnablePeerAccess(c->dev[i].ordinal, 0);
        if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) c->peer_ok = false;
        cudaGetLastError();
    }
    *out = c;
    return FRT_OK;
}

void frt_destroy(frt_ctx *ctx)
{
    if (!ctx) return;
    for (Device &d : ctx->dev)
    {
        cudaSetDevice(d.ordinal);
        cudaStreamSynchronize(d.stream);
        for (void *p : d.pool_allocs) cudaFree(p);
        void *ptrs[] = {d.d_nodes, d.d_tris, d.d_normals, d.d_spheres, d.d_mats, d.d_lights, d.d_env, d.d_ctr, d.d_film, d.d_film_peer,
                        d.d_textures, d.d_texels, d.d_tri_uv};
        for (void *p : ptrs) if (p) cudaFree(p);
        if (d.h_ctr) cudaFreeHost(d.h_ctr);
        for (cudaEvent_t e : d.timers.ev) cudaEventDestroy(e);
        if (d.ev_begin) cudaEventDestroy(d.ev_begin);
        if (d.ev_end) cudaEventDestroy(d.ev_end);
        if (d.stream) cudaStreamDestroy(d.stream);
    }
    delete ctx;
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
        if ((rc = upload_array(ctx, &d.d_nodes, s->nodes, s->n_nodes)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_tris, s->tris, 3 * (size_t)s->n_tris)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_normals, s->tri_normals, s->tri_normals ? 3 * (size_t)s->n_tris : 0)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_spheres, s->spheres, s->n_spheres)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_mats, s->materials, s->n_materials)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_lights, s->lights, s->n_lights)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_env, s->env_image, s->env_image ? 3 * (size_t)s->env_nx * s->env_ny : 0)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_textures, s->textures, s->n_textures)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_texels, s->tex_texels, 3 * (size_t)s->n_texels)) != FRT_OK) return rc;
        if ((rc = upload_array(ctx, &d.d_tri_uv, s->tri_uv, (s->tri_uv && s->n_textures) ? 6 * (size_t)s->n_tris : 0)) != FRT_OK) return rc;
    }
    uint32_t bsdf_types = 0;
    for (uint32_t i = 0; i < s->n_materials; ++i)
        if (s->materials[i].type != FRT_MAT_DIFFUSE_LIGHT) bsdf_types |= 1u << s->materials[i].type;
    ctx->multi_bsdf = (bsdf_types & (bsdf_types - 1)) != 0;
    ctx->have_scene = true;
    return FRT_OK;
}

} // extern "C"
