// Device-side scene view and small float3 helpers shared by the traversal and shading code.
// All device arithmetic is float32 (the reference is double; tolerances in tests/).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/frt.h"

#define FRT_EPSILON 1e-4f        // util.h:10
#define FRT_SHADOW_EPSILON 1e-3f // util.h:11
#define FRT_DELTA_EPSILON 1e-3f  // util.h:12
#define FRT_PI 3.14159265358979323846f
#define FRT_FLT_MAX 3.402823466e+38f

struct DevScene
{
    const float4 *nodes;       // 8 float4 per frt_bvh4_node
    const uint4 *nodes8;       // 6 uint4 per 8-wide compressed node (frt_wide8.cuh) or nullptr
    uint32_t node8_stride;     // uint4 per 8-wide node: 6 (96 B, packed) or 8 (128 B, one cache line each)
    const float4 *tris;        // 3 float4 per triangle
    const float4 *tri_normals; // 3 float4 per triangle or nullptr
    const float4 *spheres;     // 2 float4 per frt_sphere
    const frt_material *materials;
    const frt_light *lights;
    uint32_t n_nodes, n_tris, n_spheres, n_materials, n_lights;
    frt_camera cam;
    float env[3];
    const float *env_img; // lat-long HDR environment image (3 floats per texel, row 0 = top) or nullptr
    uint32_t env_nx, env_ny;
    // surface textures (texture.h:30-112): table, texels, per-triangle mesh uv (6 floats); all null without textures
    const frt_texture *textures;
    const float *tex_texels;
    const float *tri_uv;
    uint32_t n_textures;
    // upper end of an occlusion query (uniform per launch): 1 - SHADOW_EPSILON for path.cpp:50 and the ray-cast
    // entry, ao_t_max for ao.cpp:22; with its float32 brackets
    double any_tend;
    float any_tend_lo, any_tend_hi;
};

struct Counters // one per device, device memory; mirrored to pinned host memory by readback
{
    unsigned long long next_sample;   // next unissued sample index of the current tile
    unsigned long long total_samples; // samples in the current tile (this device's share)
    unsigned long long closest_rays, shadow_rays, node_visits, tri_tests, samples_started, nonfinite;
    unsigned int trace_next;          // k_trace: persistent-warp work cursor
    unsigned int wf_iterations;       // wavefront: most iterations any CTA ran
    unsigned int cancel_epoch;        // frt_cancel: tiles launched with an epoch <= this stop (written by an async copy)
    unsigned int pad_;
    unsigned long long stage_ns[4];   // wavefront: generate, extend, shade, connect time summed over CTAs
};

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 f3(const float4 &v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
// one IEEE reciprocal and three multiplies instead of three IEEE divisions (differs by <= 1 ulp per component)
__device__ __forceinline__ float3 operator/(float3 a, float s) { const float inv = 1.0f / s; return f3(a.x * inv, a.y * inv, a.z * inv); }
__device__ __forceinline__ float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross(float3 a, float3 b)
{
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float length(float3 a) { return sqrtf(dot(a, a)); }
// geometry.h:445 (v / v.length()) as one reciprocal square root and three multiplies: within 2 ulp per component
__device__ __forceinline__ float3 unit_vector(float3 a) { return a * rsqrtf(dot(a, a)); }
__device__ __forceinline__ bool nonzero(float3 a) { return a.x != 0.f || a.y != 0.f || a.z != 0.f; }
__device__ __forceinline__ float3 arr3(const float *p) { return f3(p[0], p[1], p[2]); }

// environment_map::eval material.h:219-232 — the colour a ray that leaves the scene gathers.  Over a constant_texture
// it is sc.env; over an HDR image_texture (texture.h:60-86) the lat-long texel the direction falls in: phi =
// atan2(x, -z) wrapped to [0, 2 pi), theta = acos(y), u = phi / 2 pi, v = theta / pi, texel (int(u nx), int(v ny)),
// the last row / column absorbing u = 1 / v = 1, no filtering.  Out of line: one call per path at most, and the
// path kernel's register budget stays what the traversal needs.
__device__ __noinline__ float3 frt_env_image(const float *img, uint32_t nx, uint32_t ny, float dx, float dy, float dz)
{
    const float inv = rsqrtf(dx * dx + dy * dy + dz * dz);
    dx *= inv; dy *= inv; dz *= inv;
    float phi = atan2f(dx, -dz);
    const float theta = acosf(fminf(1.0f, fmaxf(-1.0f, dy)));
    if (phi < 0.0f) phi += 6.28318530717958647692f;
    const float u = phi * 0.15915494309189533577f, v = theta * 0.31830988618379067154f;
    const int i = min(max((int)(u * (float)nx), 0), (int)nx - 1), j = min(max((int)(v * (float)ny), 0), (int)ny - 1);
    const float *t = img + 3 * ((size_t)j * nx + (size_t)i);
    return f3(__ldg(t), __ldg(t + 1), __ldg(t + 2));
}
__device__ __forceinline__ float3 frt_env(const DevScene &sc, float3 d)
{
    return sc.env_img ? frt_env_image(sc.env_img, sc.env_nx, sc.env_ny, d.x, d.y, d.z) : arr3(sc.env);
}

// ---- surface textures -----------------------------------------------------------------------
// (int)x as the reference's x86 build converts it (cvttsd2si): NaN and out-of-range values give INT_MIN.  The
// 1000-radius ground sphere of random_scene feeds asin(1000) = NaN through this (sphere.h:51 passes the
// UN-normalised p - center to get_sphere_uv).
__device__ __forceinline__ int frt_f2i(float x) { return (x >= -2147483648.0f && x < 2147483648.0f) ? (int)x : (int)0x80000000; }
__device__ __forceinline__ int frt_modulo(int a, int b) { const int r = a % b; return r < 0 ? r + b : r; } // util.h:125-128
// get_sphere_uv hitable.h:15-21 on p = hit point - center (not divided by the radius, sphere.h:51)
__device__ __forceinline__ void frt_sphere_uv(float3 p, float &u, float &v)
{
    const float phi = atan2f(p.z, p.x), theta = asinf(p.y);
    u = 1.0f - (phi + FRT_PI) / (2.0f * FRT_PI);
    v = (theta + FRT_PI / 2.0f) / FRT_PI;
}
// texture::value(hrec) for texture reference `ref` (1 + index): constant_texture texture.h:17-28, checker_texture
// texture.h:30-49 (picks tex1 when the two cell parities agree, else tex0, and evaluates that one), image_texture
// texture.h:60-86 (nearest texel, indices outside [0, n] wrapped with modulo, i == n clamped to n - 1).
__device__ __noinline__ float3 frt_tex_eval(const DevScene &sc, uint32_t ref, float u, float v)
{
    uint32_t idx = ref - 1u;
    for (int level = 0; level < 8; ++level)
    {
        const frt_texture t = sc.textures[idx];
        if (t.kind == FRT_TEX_CHECKER)
        {
            const int x = 2 * frt_modulo(frt_f2i(u * t.p[0] * 2.0f), 2) - 1, y = 2 * frt_modulo(frt_f2i(v * t.p[1] * 2.0f), 2) - 1;
            idx = (x * y == 1) ? t.b : t.a;
            continue;
        }
        if (t.kind == FRT_TEX_IMAGE)
        {
            const int nx = (int)t.b, ny = (int)t.c;
            int i = frt_f2i(u * (float)nx), j = frt_f2i(v * (float)ny);
            if (i < 0 || i > nx) i = frt_modulo(i, nx);
            if (j < 0 || j > ny) j = frt_modulo(j, ny);
            if (i == nx) i = nx - 1;
            if (j == ny) j = ny - 1;
            const float *q = sc.tex_texels + 3 * ((size_t)t.a + (size_t)j * nx + i);
            return f3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
        }
        return f3(t.p[0], t.p[1], t.p[2]);
    }
    return f3(0.f, 0.f, 0.f);
}
