/* frt.h — C ABI of the B200 ray-tracing core (libfrt.so, first_raytracer_b200/csrc).
 *
 * The reference (jammm/first_raytracer) has no FFI: its path integrator calls C++ virtuals in
 * one address space.  This header is the boundary a `gpu_path_integrator` (the duck-typed
 * integrator concept of integrator.h:11-47) binds instead of running path::Render / path::Li
 * (path.cpp:4-148) on cpp-taskflow workers.  Each entry point cites what it replaces.
 * Names frt_upload_scene / frt_render_tile / frt_readback are fixed by BASELINE.json.
 *
 * Conventions: plain C, POD structs, caller-owned host memory (copied during the call),
 * ctx-owned device memory and streams.  Every function returns FRT_OK (0) or a negative
 * frt_status; frt_last_error() gives the text.  No exception crosses the boundary.  A ctx is
 * not thread-safe (it is driven from the single task the integrator emplaces on the taskflow).
 * There is no CPU fallback: without a CUDA device frt_create fails with FRT_ERR_CUDA.
 */
#ifndef FRT_H
#define FRT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define FRT_API __attribute__((visibility("default")))
#else
#define FRT_API
#endif

typedef enum frt_status {
    FRT_OK = 0,
    FRT_ERR_INVALID = -1, /* bad argument / malformed scene description */
    FRT_ERR_CUDA = -2,    /* CUDA runtime error or no device */
    FRT_ERR_STATE = -3,   /* call out of order (e.g. render before upload) */
    FRT_ERR_COMM = -4     /* multi-GPU reduce failed */
} frt_status;

typedef struct frt_ctx frt_ctx;

typedef struct frt_float4 { float x, y, z, w; } frt_float4;

/* ---- acceleration structure -------------------------------------------------------------
 * Wide node built by flattening what parallel_bvh_node (parallel_bvh.h:67-160) built on the
 * host.  128 B, 16 B aligned, child boxes stored SoA so one node is eight 128-bit loads.
 * Replaces the pointer tree {left,right,box} of parallel_bvh.h:19-21. */
#define FRT_CHILD_EMPTY 0xFFFFFFFFu
#define FRT_CHILD_LEAF 0x80000000u   /* bit 31: leaf                                   */
#define FRT_CHILD_SPHERE 0x40000000u /* bit 30: leaf indexes spheres[] instead of tris */
#define FRT_LEAF_COUNT_SHIFT 26      /* bits 26..29: primitive count - 1 (1..16)       */
#define FRT_LEAF_FIRST_MASK 0x03FFFFFFu
typedef struct frt_bvh4_node {
    float lo_x[4], hi_x[4], lo_y[4], hi_y[4], lo_z[4], hi_z[4];
    uint32_t child[4]; /* inner: node index; leaf: flags | (count-1)<<26 | first; or EMPTY */
    uint32_t reserved[4];
} frt_bvh4_node;

/* Triangles, in leaf order: three float4 per triangle, what triangle::hit (triangle.h:69-118)
 * reads through mesh->vertices / edge1 / edge2:
 *   [3i+0] = v0.xyz, w = bits of primitive id (creation order, what hit_record::obj identifies)
 *   [3i+1] = v1.xyz, w = bits of material index
 *   [3i+2] = v2.xyz, w = bits of (mesh->nTriangles << 1) | FRT_TRI_GEOMETRY_NORMALS?
 *            (nTriangles gives inv_area = 1/(area*mesh->nTriangles), triangle.h:65)
 * The three VERTICES are stored (exact in float32: mesh files hold float32 coordinates), not the
 * edges: edge1/edge2 (triangle.h:60-61) are formed on the device, in float32 for the candidate
 * filter and in float64 — exactly the reference's `vertices[V[1]] - vertices[V[0]]` — for the
 * deciding test.  Triangles are listed in the depth-first, left-to-right leaf order of the host
 * BVH, the order in which parallel_bvh_node::hit (parallel_bvh.h:53-61) meets them.
 * tri_normals (may be NULL when every triangle has FRT_TRI_GEOMETRY_NORMALS):
 *   [3i+k] = vertex normal k .xyz */
#define FRT_TRI_GEOMETRY_NORMALS 1u

typedef struct frt_sphere { /* sphere.h:20-23 */
    float cx, cy, cz, radius;
    uint32_t material, prim_id, pad0, pad1;
} frt_sphere;

/* material.h: one POD per material object; parameters as the reference's public members */
typedef enum frt_material_type {
    FRT_MAT_LAMBERTIAN = 0,      /* p[0..2] albedo                         material.h:50-73   */
    FRT_MAT_PHONG = 1,           /* p[0..2] Kd, p[3..5] Ks, p[6] exponent  material.h:75-108  */
    FRT_MAT_DIELECTRIC = 2,      /* p[0..2] Ks, p[6] ior                   material.h:133-177 */
    FRT_MAT_ROUGH_CONDUCTOR = 3, /* p[0..2] reflectance, p[3..5] eta, p[6] alpha,
                                    p[7..9] k, p[10] 0=ggx 1=beckmann      material.h:246-315 */
    FRT_MAT_METAL = 4,           /* p[0..2] albedo                         material.h:111-130 */
    FRT_MAT_DIFFUSE_LIGHT = 5    /* p[0..2] emitted radiance               material.h:179-192 */
} frt_material_type;
typedef struct frt_material {
    uint32_t type;
    float p[15];
} frt_material;

/* texture.h:17-112 — a material colour that is not a constant_texture.  The material keeps the constant colour in
 * p[0..2] / p[3..5]; a slot that is textured instead carries 1 + the texture's index (as uint32 bits) in
 * p[FRT_MAT_TEX_A] (colour p[0..2]: lambertian albedo, phong Kd, dielectric Ks, conductor reflectance) or
 * p[FRT_MAT_TEX_B] (colour p[3..5]: phong Ks); 0 = constant.  Evaluated at the hit's texture coordinates
 * hit_record::u,v: interpolated mesh uv for triangles (triangle.h:104-113), get_sphere_uv of the UN-normalised
 * p - center for spheres (hitable.h:15-21, sphere.h:51 — as written in the reference). */
typedef enum frt_texture_kind {
    FRT_TEX_CONSTANT = 0, /* p[0..2] colour                                           texture.h:17-28  */
    FRT_TEX_CHECKER = 1,  /* a, b = texture indices tex0, tex1; p[0], p[1] = u, v scale  texture.h:30-49  */
    FRT_TEX_IMAGE = 2     /* a = first texel in tex_texels, b = nx, c = ny; nearest texel, wrap, no filtering
                             (texture.h:60-86); texels are linear RGB floats (LDR files: FromSrgb(byte / 255)) */
} frt_texture_kind;
typedef struct frt_texture {
    uint32_t kind, a, b, c;
    float p[4];
} frt_texture;
#define FRT_MAT_TEX_A 11
#define FRT_MAT_TEX_B 12

/* Scene::lights (Scene.h:13): self-contained copies, because the reference's light list may
 * hold objects that are not the ones in `world` (veach_mis, main.cpp:292-302). */
typedef enum frt_light_kind { FRT_LIGHT_TRIANGLE = 0, FRT_LIGHT_SPHERE = 1 } frt_light_kind;
typedef struct frt_light {
    uint32_t kind, material, flags, pad;
    float v0[4]; /* triangle: v0.xyz, w = inv_area;  sphere: center.xyz, w = radius */
    float v1[4], v2[4];
    float n0[4], n1[4], n2[4]; /* vertex normals when !(flags & FRT_TRI_GEOMETRY_NORMALS) */
    float reserved[4];
} frt_light;

typedef struct frt_camera { /* camera.h:42-48 */
    float origin[3], lower_left_corner[3], horizontal[3], vertical[3], u[3], v[3], w[3];
    float lens_radius;
} frt_camera;

typedef struct frt_scene_desc {
    uint32_t struct_size; /* = sizeof(frt_scene_desc) */
    uint32_t n_nodes;
    const frt_bvh4_node *nodes;
    uint32_t n_tris;
    uint32_t n_spheres;
    const frt_float4 *tris;
    const frt_float4 *tri_normals;
    const frt_sphere *spheres;
    uint32_t n_materials;
    uint32_t n_lights;
    const frt_material *materials;
    const frt_light *lights;
    frt_camera camera;
    float env_color[3]; /* environment_map over a constant_texture, material.h:219-232 */
    float ao_t_max;     /* 0.5 * world bounding box size.y: the occlusion range of ao::Li (ao.cpp:19-22) */
    /* environment_map over an HDR image_texture (material.h:209-232, texture.h:51-86): env_nx * env_ny * 3 floats,
     * row 0 = top of the lat-long image (theta = 0, direction +y), nearest texel, no filtering.  NULL = the
     * constant env_color above. */
    const float *env_image;
    uint32_t env_nx, env_ny;
    /* surface textures: all may be 0 / NULL for a scene of constant colours */
    uint32_t n_textures;
    uint32_t reserved0;
    const frt_texture *textures;
    const float *tex_texels; /* 3 floats per texel, every image back to back; row 0 = image::data row 0 */
    uint64_t n_texels;
    const float *tri_uv;     /* 6 floats per triangle (mesh->uv of V[0], V[1], V[2]), leaf order; NULL = all zero */
} frt_scene_desc;

/* One unit of rendering work: the pixel rectangle and sample range path::Render's row loop
 * (path.cpp:120-147) would cover.  Sample s of pixel (x,y) is identified by (seed, y*nx+x, s)
 * whatever tile / device / rank renders it, so splitting [spp_begin,spp_end) across GPUs
 * gives the same multiset of samples as one GPU. */
/* Which Li the tile runs (the integrators next to `path`, same trace kernels) */
typedef enum frt_integrator {
    FRT_INTEGRATOR_PATH = 0,   /* path::Li, path.cpp:4-116 (NEE + MIS) */
    FRT_INTEGRATOR_AO = 1,     /* ao::Li, ao.cpp:4-28: one BSDF-sampled occlusion ray of length ao_t_max per sample */
    FRT_INTEGRATOR_NORMALS = 2, /* normals_renderer::Li, debug_renderer.h:8-17: the hit normal as colour */
    FRT_INTEGRATOR_DIRECT = 3   /* ass_1_renderer::Li, assignment_1.cpp:4-38: every light sampled once, no visibility test */
} frt_integrator;

typedef struct frt_tile {
    uint32_t nx, ny;         /* film size (viewer::nx, ny) */
    uint32_t x0, y0, w, h;   /* pixel rectangle, row 0 = bottom (viewer.cpp:109-132) */
    uint32_t spp_begin, spp_end;
    uint32_t seed;
    uint32_t max_depth;      /* deepest vertex that still scatters + 1; reference: 34 (path.cpp:36) */
    uint32_t integrator;     /* frt_integrator; 0 = path */
    uint32_t reserved;       /* must be 0 */
} frt_tile;

typedef struct frt_stats {
    uint64_t samples;        /* camera paths started (all counters are 64-bit: a 1080p x 1024 spp frame traces 1.4e10 rays) */
    uint64_t closest_rays;   /* world->hit(r, EPSILON, FLT_MAX) equivalents, path.cpp:10 */
    uint64_t shadow_rays;    /* world->hit(shadow_ray, EPSILON, 1-SHADOW_EPSILON), path.cpp:50 */
    uint64_t node_visits;    /* wide-node fetches (0 unless counters enabled) */
    uint64_t tri_tests;      /* triangle fetches (0 unless counters enabled) */
    uint64_t iterations;     /* wavefront iterations */
    uint64_t kernel_launches;
    double render_ms;        /* device time between first and last kernel, max over devices */
    double extend_ms, shade_ms, connect_ms, generate_ms; /* per stage of the persistent kernel: mean over its CTAs */
} frt_stats;

typedef struct frt_ray { float ox, oy, oz, dx, dy, dz; } frt_ray; /* ray.h:6-17, d not normalised */
typedef struct frt_hit {
    float t, b1, b2;  /* hit_record::t and barycentric uv (hitable.h:23-36) */
    int32_t prim_id;  /* creation-order primitive id, -1 = miss */
} frt_hit;

#define FRT_FLAG_COUNTERS 1u /* count node visits / triangle tests (slower); bits 1 and 2 belonged to the round-1
                                global-memory path pool, which is gone: frt_set_flags rejects them */

/* Create a context on `n` CUDA devices (ordinals in `devices`, NULL = device 0..n-1). */
FRT_API int frt_create(frt_ctx **out, const int *devices, int n);
FRT_API void frt_destroy(frt_ctx *ctx);
FRT_API const char *frt_last_error(const frt_ctx *ctx); /* ctx may be NULL: last creation error */
FRT_API int frt_set_flags(frt_ctx *ctx, uint32_t flags);
FRT_API int frt_device_count(const frt_ctx *ctx);

/* Replaces handing `Scene*` to the integrator (integrator.h:14, Scene.h:10-26): copies the
 * flattened scene to every device of the context.  Validates indices. */
FRT_API int frt_upload_scene(frt_ctx *ctx, const frt_scene_desc *scene);

/* Replaces path::Render + path::Li (path.cpp:4-148) for the tile: asynchronous; the sample
 * range is split evenly over the context's devices.  Radiance is accumulated (summed, not
 * averaged) into the context's film, which is (re)allocated and zeroed when nx/ny change. */
FRT_API int frt_render_tile(frt_ctx *ctx, const frt_tile *tile);
FRT_API int frt_clear_film(frt_ctx *ctx);

/* Replaces viewer::add_sample's store (viewer.cpp:109-132): waits for rendering, sums the
 * per-device films into device 0's, copies nx*ny*3 float32 radiance SUMS (row 0 = bottom) to
 * `rgb_sum` (may be NULL to only synchronise) and fills `stats` (may be NULL).
 * The sum over the context's devices is a reduce-scatter over peer memory (device g adds range g of every film with
 * NVLink loads and stores it into device 0's film), NOT an NCCL call: libfrt.so links no NCCL.  BASELINE.json's "one
 * NCCL reduce" is what a one-process-per-GPU launcher does between frt_film_device_ptr and frt_readback (bench.py
 * under torchrun does); both give the same film. */
FRT_API int frt_readback(frt_ctx *ctx, float *rgb_sum, size_t n_floats, frt_stats *stats);

/* The same wait + sum, then the whole of viewer::add_sample for every pixel (viewer.cpp:109-132) instead of one call
 * per pixel from path.cpp:143: mean c = sum * (1 / ns) to fout_image (double, the viewer's own buffer) and / or
 * mean_rgb (float32); out_image = (GLubyte) int(pow(1 - exp(-c), 1 / 2.2) * 255 + .5), evaluated in double on
 * device 0.  Each of the three outputs may be NULL; nx*ny*3 elements, row 0 = bottom. */
FRT_API int frt_readback_film(frt_ctx *ctx, uint64_t ns, double *fout_image, unsigned char *out_image, float *mean_rgb,
                              frt_stats *stats);

/* viewer::to_exit (viewer.cpp:101-104), which path.cpp:127 polls per row: the tiles launched so far stop taking
 * samples and drop their paths in flight (each CTA looks at the flag once per wavefront iteration); frt_readback
 * then returns within milliseconds with what the film holds.  The next frt_render_tile clears the request.
 * Callable from any thread while another one is inside frt_readback / frt_poll. */
FRT_API int frt_cancel(frt_ctx *ctx);
/* Non-blocking: *done = 1 when every tile launched so far has finished on every device (what a caller polls while
 * it watches viewer::to_exit). */
FRT_API int frt_poll(frt_ctx *ctx, int *done);

/* Resume: load radiance SUMS saved from an earlier frt_readback (nx*ny*3 float32, row 0 =
 * bottom) as the film's starting state, so that rendering the remaining sample range continues an interrupted
 * render; (re)allocates the film.  The sums go to device 0, the other devices' films start at zero. */
FRT_API int frt_restore_film(frt_ctx *ctx, uint32_t nx, uint32_t ny, const float *rgb_sum, size_t n_floats);

/* Counters only (rays, node visits, triangle tests since the last frt_clear_film), summed over the
 * devices; synchronises.  Works without a film, i.e. after frt_trace_rays alone: the integrator's
 * statistics block (integrator.h:33-43) for a pure ray-cast run. */
FRT_API int frt_get_counters(frt_ctx *ctx, frt_stats *stats);

/* Device address of device `dev`'s film (nx*ny*3 float32) so that a multi-process launcher can
 * reduce it in place with its own communicator before frt_readback.  Synchronises `dev`. */
FRT_API int frt_film_device_ptr(frt_ctx *ctx, int dev, void **ptr, size_t *n_floats);

/* The CUDA stream (cudaStream_t) device `dev` of the context launches on, so that a harness can
 * bracket calls with its own events or enqueue a collective behind the render. */
FRT_API int frt_stream_handle(frt_ctx *ctx, int dev, void **stream);

/* Test / ray-cast entry: what scene->world->hit(r, EPSILON, FLT_MAX, rec) (mode 0, path.cpp:10) or
 * world->hit(r, EPSILON, 1 - SHADOW_EPSILON, rec) (FRT_TRACE_ANY_HIT, path.cpp:50; only prim_id >= 0 / -1 is
 * meaningful) return for each ray.  Host buffers; the rays are block-partitioned over the context's devices, which
 * copy, trace and copy back AT THE SAME TIME (per-device streams and ctx-owned device buffers that persist between
 * calls).  Buffers from frt_host_alloc (or any pinned memory) move in chunks of 2^21 rays on separate copy streams, so
 * that the copy in of chunk k+1 and the copy out of chunk k-1 run under the cast of chunk k; pageable ones go through
 * one host thread per device.  FRT_TRACE_SORT: the rays are first put in Morton order of (origin cell,
 * direction octant) by a device radix sort — only the order they are traced in changes, every hit lands at its ray's
 * index.  `device_ms` (may be NULL) = sort + kernel time, max over devices.  With FRT_WIDE8=1 in the environment the
 * casts walk 8-wide compressed nodes derived at upload (same hits; measured slower than the 4-wide walk, off by default). */
#define FRT_TRACE_ANY_HIT 1
#define FRT_TRACE_SORT 2
FRT_API int frt_trace_rays(frt_ctx *ctx, const frt_ray *rays, size_t n, frt_hit *hits, int mode, double *device_ms);

/* Same on device-resident buffers of device `dev` (bench: inputs already in HBM).  device_ms == NULL: asynchronous
 * on the device's stream (frt_stream_handle). */
FRT_API int frt_trace_rays_device(frt_ctx *ctx, int dev, const void *d_rays, size_t n, void *d_hits, int mode,
                          double *device_ms);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * frt_trace_rays_device with FRT_TRACE_SORT (implied here: the Morton-ordered cast is the kernel sequence that needs
 * scratch — keys, ray indices and radix histograms), but the scratch is a fixed array in the library's device image
 * instead of a grow-only cudaMalloc: the call allocates NOTHING (no cudaMalloc / cudaFree / allocating library
 * routine); the caller passes device-resident rays and hits of device `dev` as before.  Same hits as
 * frt_trace_rays_device, bit for bit.
 *   - n > FRT_TRACE_NOALLOC_MAX_RAYS is rejected with FRT_ERR_INVALID and a message in frt_last_error.
 *   - The array (17 MiB) is held by every GPU that loads libfrt.so until the process exits.
 *   - Every call on a GPU shares the one array, which keeps the previous call's contents (each pass overwrites what
 *     it reads; the work cursor is cleared per call).  Calls that OVERLAP on the same GPU — two contexts, two host
 *     threads, or one ordinal listed twice in a context — race on it: serialise them. */
#define FRT_TRACE_NOALLOC_MAX_RAYS (1u << 20)
FRT_API int frt_trace_rays_device_noalloc(frt_ctx *ctx, int dev, const void *d_rays, size_t n, void *d_hits, int mode,
                                          double *device_ms);

/* Pinned host memory for ray / hit / film buffers (cudaHostAlloc, portable); NULL on failure. */
FRT_API void *frt_host_alloc(size_t bytes);
FRT_API void frt_host_free(void *p);

/* Host utility (no device needed): the 8-wide compressed nodes frt_upload_scene derives from `scene->nodes` for large
 * scenes (24 uint32 words per node, layout in csrc/frt_wide8.cuh).  words may be NULL to only count. */
FRT_API int frt_build_wide8(const frt_scene_desc *scene, uint32_t *words, size_t cap_words, uint32_t *n_nodes8, int *depth);

#ifdef __cplusplus
}
#endif
#endif /* FRT_H */
