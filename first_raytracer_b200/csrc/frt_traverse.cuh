// Closest-hit / any-hit traversal of the 4-wide flattened BVH.
//
// Semantics follow the reference's live (non-SSE) code, evaluated in float32:
//   aabb::hit            aabb.h:14-31       slab test, reject iff tmax < tmin, NaNs ignored
//   triangle::hit        triangle.h:69-118  Möller–Trumbore, a == 0 rejects, strict t bounds
//   sphere::hit          sphere.h:26-56     near root, far root if near < t_min, non-strict bounds
//   parallel_bvh_node::hit parallel_bvh.h:39-64  origin-scaled t_min; closest hit wins,
//                        traversal order is free (only exact ties depend on it).
// One thread owns one ray; the short stack lives in local memory (L1-resident).
#pragma once
#include "frt_device.cuh"

#define FRT_STACK_SIZE 96

struct RayHit
{
    float t, u, v;
    uint32_t prim; // index into tris[] (leaf order), | FRT_CHILD_SPHERE for spheres[]; 0xFFFFFFFF = miss
};

// parallel_bvh.h:46-51: t_min = max(EPSILON, EPSILON * max(|o.x|,|o.y|,|o.z|,EPSILON))
__device__ __forceinline__ float frt_ray_tmin(float3 o)
{
    const float m = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
    return FRT_EPSILON * fmaxf(m, 1.0f);
}

__device__ __forceinline__ uint32_t frt_pick(const uint4 &c, uint32_t i)
{
    return i == 0 ? c.x : (i == 1 ? c.y : (i == 2 ? c.z : c.w));
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ RayHit frt_traverse(const DevScene &sc, const float3 o, const float3 d, float tmax,
                                               unsigned int &n_nodes, unsigned int &n_tris)
{
    RayHit best;
    best.t = 0.f; best.u = 0.f; best.v = 0.f; best.prim = 0xFFFFFFFFu;
    if (sc.n_nodes == 0) return best;

    const float tmin = frt_ray_tmin(o);
    const float idx = 1.0f / d.x, idy = 1.0f / d.y, idz = 1.0f / d.z;
    const float oox = o.x * idx, ooy = o.y * idy, ooz = o.z * idz;
    const int sx = idx < 0.0f ? 1 : 0, sy = idy < 0.0f ? 1 : 0, sz = idz < 0.0f ? 1 : 0;

    uint32_t stack[FRT_STACK_SIZE];
    int sp = 0;
    uint32_t cur = 0;

    while (true)
    {
        if (cur == FRT_CHILD_EMPTY)
        {
            if (sp == 0) break;
            cur = stack[--sp];
            continue;
        }
        if (cur & FRT_CHILD_LEAF)
        {
            const uint32_t first = cur & FRT_LEAF_FIRST_MASK;
            const uint32_t count = ((cur >> FRT_LEAF_COUNT_SHIFT) & 15u) + 1u;
            if (cur & FRT_CHILD_SPHERE)
            {
                for (uint32_t k = 0; k < count; ++k)
                {
                    const float4 s = __ldg(sc.spheres + 2 * (first + k));
                    const float3 oc = o - f3(s);
                    const float a = dot(d, d), b = dot(oc, d), c = dot(oc, oc) - s.w * s.w;
                    float disc = b * b - a * c;
                    if (disc >= 0.0f)
                    {
                        disc = sqrtf(disc);
                        float t = (-b - disc) / a;
                        if (t < tmin) t = (-b + disc) / a;
                        if (!(t < tmin || t > tmax))
                        {
                            best.t = t; best.u = 0.f; best.v = 0.f; best.prim = (first + k) | FRT_CHILD_SPHERE;
                            if (ANY) return best;
                            tmax = t;
                        }
                    }
                }
            }
            else
            {
                for (uint32_t k = 0; k < count; ++k)
                {
                    const float4 *tp = sc.tris + 3 * (size_t)(first + k);
                    const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
                    if (COUNT) ++n_tris;
                    const float3 v0 = f3(q0), e1 = f3(q1), e2 = f3(q2);
                    const float3 h = cross(d, e2);
                    const float a = dot(e1, h);
                    if (a == 0.0f) continue;
                    const float f = 1.0f / a;
                    const float3 s = o - v0;
                    const float u = f * dot(s, h);
                    if (u < 0.0f || u > 1.0f) continue;
                    const float3 q = cross(s, e1);
                    const float v = f * dot(d, q);
                    if (v >= 0.0f && u + v <= 1.0f)
                    {
                        const float t = f * dot(e2, q);
                        if (t > tmin && t < tmax)
                        {
                            best.t = t; best.u = u; best.v = v; best.prim = first + k;
                            if (ANY) return best;
                            tmax = t;
                        }
                    }
                }
            }
            cur = FRT_CHILD_EMPTY;
            continue;
        }

        // inner node: eight 128-bit loads, near/far planes picked by ray-direction sign
        const float4 *n = sc.nodes + 8 * (size_t)cur;
        if (COUNT) ++n_nodes;
        const float4 nx = __ldg(n + (0 ^ sx)), fx = __ldg(n + (1 ^ sx));
        const float4 ny = __ldg(n + (2 ^ sy)), fy = __ldg(n + (3 ^ sy));
        const float4 nz = __ldg(n + (4 ^ sz)), fz = __ldg(n + (5 ^ sz));
        const uint4 ch = __ldg(reinterpret_cast<const uint4 *>(n + 6));

        // fmaxf / fminf drop a NaN operand, like aabb.h:24-25's `t0 > tmin ? t0 : tmin`
        const float tn0 = fmaxf(fmaxf(nx.x * idx - oox, ny.x * idy - ooy), fmaxf(nz.x * idz - ooz, tmin));
        const float tn1 = fmaxf(fmaxf(nx.y * idx - oox, ny.y * idy - ooy), fmaxf(nz.y * idz - ooz, tmin));
        const float tn2 = fmaxf(fmaxf(nx.z * idx - oox, ny.z * idy - ooy), fmaxf(nz.z * idz - ooz, tmin));
        const float tn3 = fmaxf(fmaxf(nx.w * idx - oox, ny.w * idy - ooy), fmaxf(nz.w * idz - ooz, tmin));
        const float tf0 = fminf(fminf(fx.x * idx - oox, fy.x * idy - ooy), fminf(fz.x * idz - ooz, tmax));
        const float tf1 = fminf(fminf(fx.y * idx - oox, fy.y * idy - ooy), fminf(fz.y * idz - ooz, tmax));
        const float tf2 = fminf(fminf(fx.z * idx - oox, fy.z * idy - ooy), fminf(fz.z * idz - ooz, tmax));
        const float tf3 = fminf(fminf(fx.w * idx - oox, fy.w * idy - ooy), fminf(fz.w * idz - ooz, tmax));

        // sort the (up to four) intersected children near-to-far: key = quantised tnear | slot
        uint32_t k0 = (tf0 >= tn0) ? ((__float_as_uint(tn0) & ~3u) | 0u) : 0xFFFFFFFFu;
        uint32_t k1 = (tf1 >= tn1) ? ((__float_as_uint(tn1) & ~3u) | 1u) : 0xFFFFFFFFu;
        uint32_t k2 = (tf2 >= tn2) ? ((__float_as_uint(tn2) & ~3u) | 2u) : 0xFFFFFFFFu;
        uint32_t k3 = (tf3 >= tn3) ? ((__float_as_uint(tn3) & ~3u) | 3u) : 0xFFFFFFFFu;
        uint32_t a, b;
        a = min(k0, k1); b = max(k0, k1); k0 = a; k1 = b;
        a = min(k2, k3); b = max(k2, k3); k2 = a; k3 = b;
        a = min(k0, k2); b = max(k0, k2); k0 = a; k2 = b;
        a = min(k1, k3); b = max(k1, k3); k1 = a; k3 = b;
        a = min(k1, k2); b = max(k1, k2); k1 = a; k2 = b;

        if (k3 != 0xFFFFFFFFu && sp < FRT_STACK_SIZE) stack[sp++] = frt_pick(ch, k3 & 3u);
        if (k2 != 0xFFFFFFFFu && sp < FRT_STACK_SIZE) stack[sp++] = frt_pick(ch, k2 & 3u);
        if (k1 != 0xFFFFFFFFu && sp < FRT_STACK_SIZE) stack[sp++] = frt_pick(ch, k1 & 3u);
        cur = (k0 != 0xFFFFFFFFu) ? frt_pick(ch, k0 & 3u) : FRT_CHILD_EMPTY;
    }
    return best;
}
