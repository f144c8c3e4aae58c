// shade: everything path::Li does at one vertex between two world->hit calls (path.cpp:11-115),
// plus the film splat of viewer::add_sample (viewer.cpp:109-132) when the path ends.
#pragma once
#include "frt_kernels.cuh"

struct LightSample // what sample_direct leaves in lrec + its return value
{
    float3 to_light; // un-normalised
    float3 n;        // lrec.normal
};

// triangle::sample_direct triangle.h:145-175 / sphere::sample_direct sphere.h:80-107
__device__ __forceinline__ LightSample light_sample_direct(const frt_light &L, float3 o, float u0, float u1)
{
    LightSample ls;
    if (L.kind == FRT_LIGHT_TRIANGLE)
    {
        const float3 v0 = arr3(L.v0), v1 = arr3(L.v1), v2 = arr3(L.v2);
        const float su0 = sqrtf(u0);
        const float b0 = 1.0f - su0, b1 = u1 * su0;
        const float3 pt = (1.0f - b0 - b1) * v0 + b0 * v1 + b1 * v2;
        if (L.flags & FRT_TRI_GEOMETRY_NORMALS)
            ls.n = unit_vector(cross(v1 - v0, v2 - v0));
        else
            ls.n = unit_vector((1.0f - b0 - b1) * arr3(L.n0) + b0 * arr3(L.n1) + b1 * arr3(L.n2));
        ls.to_light = pt - o;
        return ls;
    }
    const float3 c = arr3(L.v0);
    const float radius = L.v0[3];
    const float3 direction = c - o;
    const float d2 = dot(direction, direction);
    if (d2 <= radius * radius)
    {
        // uniform_sample_sphere pdf.h:38-44
        const float z = 1.0f - 2.0f * u0;
        const float r = sqrtf(fmaxf(0.0f, 1.0f - z * z));
        float sp, cp;
        sincospif(2.0f * u1, &sp, &cp);
        const float3 p = c + radius * f3(r * cp, r * sp, z);
        ls.n = unit_vector(c - p);
        ls.to_light = p - o;
        return ls;
    }
    // random_to_sphere pdf.h:46-56 pushed through the un-normalised basis (sphere.h:84-85,100)
    const float z = 1.0f + u1 * (sqrtf(1.0f - radius * radius / d2) - 1.0f);
    float sp, cp;
    sincospif(2.0f * u0, &sp, &cp);
    const float s = sqrtf(1.0f - z * z);
    const float3 p = Onb(direction).from_local(f3(cp * s, sp * s, z));
    ls.n = unit_vector(p);
    ls.to_light = p;
    return ls;
}

// sphere::pdf_direct_sampling sphere.h:64-78 with the record/direction the caller passes
__device__ __forceinline__ float sphere_pdf_direct(float3 c, float radius, float3 rec_p, float rec_t, float3 rec_n,
                                                   float3 to_light)
{
    const float3 o = rec_p + to_light * (-rec_t);
    const float3 direction = c - o;
    const float d2 = dot(direction, direction);
    const float r2 = radius * radius;
    if (d2 <= r2) return 1.0f / (4.0f * FRT_PI * radius * radius);
    const float cos_theta_max = sqrtf(1.0f - r2 / d2);
    const float solid_angle = 2.0f * FRT_PI * (1.0f - cos_theta_max);
    return (1.0f / solid_angle) * fabsf(dot(to_light, rec_n)) / d2;
}

// ass_1_renderer::Li assignment_1.cpp:13-34 at a surface that scatters: every light sampled once (one get2d each, after
// the get3d that scatter() consumes), NO visibility test, to_light left un-normalised, the light's area-measure pdf
// used as is.  Out of line: the other integrators never call it and must not pay registers for it.
__device__ __noinline__ float3 frt_direct_lighting(const DevScene &sc, const frt_material &m, const Surface h, unsigned int seed,
                                                   unsigned int pixel, unsigned int sample)
{
    float3 Le = f3(0.f, 0.f, 0.f);
    const float3 so = h.p + FRT_EPSILON * h.n;
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned int i = 0; i < sc.n_lights; ++i)
    {
        if ((i & 1u) == 0) r = frt_rng_block(seed, pixel, sample, 2 + (i >> 1)); // sampler layout: oracle/frt_rng.h
        const float u0 = (i & 1u) ? r.z : r.x, u1 = (i & 1u) ? r.w : r.y;
        const frt_light &L = sc.lights[i];
        const LightSample ls = light_sample_direct(L, so, u0, u1);
        const float3 bsdf = mat_eval(m, h, ls.to_light);
        const float cos_wi = fmaxf(0.0f, dot(h.n, unit_vector(ls.to_light)));
        float light_pdf;
        if (L.kind == FRT_LIGHT_TRIANGLE) light_pdf = L.v0[3];
        else light_pdf = sphere_pdf_direct(arr3(L.v0), L.v0[3], h.p, h.t, h.n, ls.to_light);
        const frt_material &lm = sc.materials[L.material];
        if (lm.type == FRT_MAT_DIFFUSE_LIGHT && dot(ls.n, ls.to_light) < 0.0f) // diffuse_light::emitted material.h:184-190
            Le = Le + bsdf * arr3(lm.p) * cos_wi / light_pdf;
    }
    return Le;
}

__device__ __forceinline__ void end_path(const KernelArgs &a, const PathPool &pool, unsigned int slot, float3 rad,
                                         unsigned int pixel, unsigned int &nonfinite)
{
    if (isfinite(rad.x) && isfinite(rad.y) && isfinite(rad.z))
    {
        // adding zero is the identity: paths that gathered nothing (misses on a black environment,
        // 44 % of the 16:9 Cornell frame) cost no film traffic
        float *px = a.film + 3 * (size_t)pixel;
        if (rad.x != 0.0f) atomicAdd(px + 0, rad.x);
        if (rad.y != 0.0f) atomicAdd(px + 1, rad.y);
        if (rad.z != 0.0f) atomicAdd(px + 2, rad.z);
    }
    else
        ++nonfinite;
    pool.ray_d[slot].w = __uint_as_float(0u);
}

// One path vertex of slot `slot` of `pool` (the CTA's shared-memory pool): bit 0 of the result = the path ended here (splatted, slot dead), bit 1 =
// a shadow ray was written to sh_a / sh_b / sh_c.  A dead slot returns 0.
#define FRT_SHADE_TERMINATED 1u
#define FRT_SHADE_SHADOW 2u
#define FRT_SHADE_WAS_ALIVE 4u
// Bits 8.. of the result = the primitive the vertex lies on (leaf-order index), a spatial key for
// binning the rays that leave it.
#define FRT_SHADE_PRIM_SHIFT 3
// OTHER = false: path::Li only (the tile's integrator field is not read: the hot kernel carries none of the code
// below it does not run); OTHER = true: ao / normals_renderer / ass_1_renderer, selected by a.tile.integrator.
template <bool OTHER>
__device__ __forceinline__ unsigned int frt_shade_vertex(const KernelArgs &a, const PathPool &pool, const unsigned int slot,
                                                         unsigned int &nonfinite)
{
        unsigned int origin_prim = 0;
        bool terminated = false, want_shadow = false;
        const float4 dw = pool.ray_d[slot];
        const unsigned int flags = __float_as_uint(dw.w);
        if (flags & FRT_FLAG_ALIVE)
        {
            const float4 ow = pool.ray_o[slot];
            float4 hw = pool.hit[slot];
            float4 tw = pool.thr[slot];
            float4 rw = pool.rad[slot];
            const float3 o = f3(ow), d = f3(dw);
            const float prev_pdf = ow.w;
            const unsigned int depth = flags >> FRT_DEPTH_SHIFT;
            unsigned int prim = __float_as_uint(hw.w);
            // a winner taken on its float32 interval gets the reference's float64 t, u, v here, where the
            // warp is converged (the CTA-resident wavefront defers it out of the divergent traversal)
            if (prim != 0xFFFFFFFFu && (prim & FRT_HIT_INEXACT)) frt_resolve_hit(a.sc, o, d, hw.x, hw.y, hw.z, prim);
            origin_prim = prim & FRT_LEAF_FIRST_MASK;
            const unsigned int pixel = __float_as_uint(tw.w), sample = __float_as_uint(rw.w);
            float3 thr = f3(tw), rad = f3(rw);

            if (OTHER && (flags & FRT_FLAG_FINISH))
            {
                // ao: the occlusion ray of the only vertex has been resolved by the trace stage
                end_path(a, pool, slot, rad, pixel, nonfinite);
                terminated = true;
            }
            else if (prim == 0xFFFFFFFFu)
            {
                // path.cpp:115 / ao.cpp:27 / debug_renderer.h:17 — environment_map::eval material.h:219-232 (constant colour or HDR image)
                rad = rad + thr * frt_env(a.sc, d);
                end_path(a, pool, slot, rad, pixel, nonfinite);
                terminated = true;
            }
            else
            {
                // ---- hit_record as the intersector would have filled it ----
                Surface h;
                h.t = hw.x;
                h.p = o + d * h.t;
                h.wi = -unit_vector(d);
                unsigned int mat_id;
                float3 e1 = f3(0.f, 0.f, 0.f), e2 = e1, sph_c = e1;
                float sph_r = 0.0f;
                unsigned int tri_flags = 0;
                const bool is_sphere = prim & FRT_CHILD_SPHERE;
                if (is_sphere)
                {
                    const unsigned int si = prim & FRT_LEAF_FIRST_MASK;
                    const float4 s0 = __ldg(a.sc.spheres + 2 * si), s1 = __ldg(a.sc.spheres + 2 * si + 1);
                    sph_c = f3(s0); sph_r = s0.w;
                    mat_id = __float_as_uint(s1.x);
                    h.n = (h.p - sph_c) / sph_r; // sphere.h:45-48
                    const float3 oc = o - sph_c;
                    if (dot(oc, oc) < sph_r * sph_r) h.n = -h.n;
                }
                else
                {
                    const float4 *tp = a.sc.tris + 3 * (size_t)prim;
                    const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
                    e1 = f3(q1) - f3(q0); e2 = f3(q2) - f3(q0); // triangle.h:60-61
                    mat_id = __float_as_uint(q1.w);
                    tri_flags = __float_as_uint(q2.w);
                    if (tri_flags & FRT_TRI_GEOMETRY_NORMALS) // triangle.h:100-103
                        h.n = unit_vector(cross(e1, e2));
                    else
                    {
                        const float4 *np = a.sc.tri_normals + 3 * (size_t)prim;
                        h.n = unit_vector((1.0f - hw.y - hw.z) * f3(__ldg(np)) + hw.y * f3(__ldg(np + 1)) + hw.z * f3(__ldg(np + 2)));
                    }
                }
                const frt_material *mp = &a.sc.materials[mat_id];
                frt_material mtex;
                if (OTHER && a.sc.n_textures)
                {
                    // textured colours (texture.h:30-112) are looked up at hit_record::u, v and put where the constant
                    // colour would be; only the general kernel carries this (the path kernel of untextured scenes does not)
                    const unsigned int ta = __float_as_uint(mp->p[FRT_MAT_TEX_A]), tb = __float_as_uint(mp->p[FRT_MAT_TEX_B]);
                    if (ta | tb)
                    {
                        float tu = 0.0f, tv = 0.0f;
                        if (is_sphere) frt_sphere_uv(h.p - sph_c, tu, tv); // sphere.h:51
                        else if (a.sc.tri_uv)
                        {
                            // triangle.h:104-106: (1 - u - v) uv[V0] + u uv[V1] + v uv[V2]
                            const float *q = a.sc.tri_uv + 6 * (size_t)prim;
                            const float w0 = 1.0f - hw.y - hw.z;
                            tu = w0 * __ldg(q) + hw.y * __ldg(q + 2) + hw.z * __ldg(q + 4);
                            tv = w0 * __ldg(q + 1) + hw.y * __ldg(q + 3) + hw.z * __ldg(q + 5);
                        }
                        mtex = *mp;
                        if (ta) { const float3 c = frt_tex_eval(a.sc, ta, tu, tv); mtex.p[0] = c.x; mtex.p[1] = c.y; mtex.p[2] = c.z; }
                        if (tb) { const float3 c = frt_tex_eval(a.sc, tb, tu, tv); mtex.p[3] = c.x; mtex.p[4] = c.y; mtex.p[5] = c.z; }
                        mp = &mtex;
                    }
                }
                const frt_material &m = *mp;
                const unsigned int mtype = m.type;

                if (OTHER && a.tile.integrator == FRT_INTEGRATOR_NORMALS)
                {
                    // normals_renderer::Li debug_renderer.h:11-15: hrec.normal as the colour (can be negative)
                    end_path(a, pool, slot, h.n, pixel, nonfinite);
                    terminated = true;
                }
                else if (OTHER && a.tile.integrator == FRT_INTEGRATOR_DIRECT)
                {
                    // a surface that does not scatter (diffuse_light) returns black: assignment_1.cpp:15,35
                    float3 Le = f3(0.f, 0.f, 0.f);
                    if (mtype != FRT_MAT_DIFFUSE_LIGHT) Le = frt_direct_lighting(a.sc, m, h, a.tile.seed, pixel, sample);
                    end_path(a, pool, slot, Le, pixel, nonfinite);
                    terminated = true;
                }
                else if (OTHER && a.tile.integrator == FRT_INTEGRATOR_AO)
                {
                    // ao::Li ao.cpp:4-28: one direction drawn from the BSDF's pdf (get3d for scatter, then get2d for
                    // generate), occluded within 0.5 * world size.y -> black, else the environment colour.  A
                    // material that does not scatter (diffuse_light) also returns the environment.
                    const float3 env = frt_env(a.sc, d); // evaluated for the camera ray (ao.cpp:27)
                    if (mtype == FRT_MAT_DIFFUSE_LIGHT || !nonzero(env))
                    {
                        end_path(a, pool, slot, mtype == FRT_MAT_DIFFUSE_LIGHT ? env : f3(0.f, 0.f, 0.f), pixel, nonfinite);
                        terminated = true;
                    }
                    else
                    {
                        const float4 r1 = frt_rng_block(a.tile.seed, pixel, sample, 1);
                        const float4 r2 = frt_rng_block(a.tile.seed, pixel, sample, 2);
                        ScatterRec srec;
                        mat_scatter(m, h, d, r1.x, r1.y, srec);
                        float3 dir;
                        if (srec.is_specular) // the specular pdfs' generate is what their scatter calls (material.h:88-89, ...)
                        {
                            ScatterRec again;
                            mat_scatter(m, h, d, r2.x, r2.y, again);
                            dir = again.dir;
                        }
                        else // cosine_pdf::generate pdf.h:91-94
                        {
                            const float r = sqrtf(r2.x);
                            float sp, cp;
                            sincospif(2.0f * r2.y, &sp, &cp);
                            dir = Onb(h.n).from_local(f3(r * cp, r * sp, sqrtf(1.0f - r2.x)));
                        }
                        want_shadow = true; // ray(hrec.p, dir): no offset along the normal (ao.cpp:17)
                        pool.sh_a[slot] = make_float4(h.p.x, h.p.y, h.p.z, dir.x);
                        pool.sh_b[slot] = make_float4(dir.y, dir.z, env.x, env.y);
                        pool.sh_c[slot] = env.z;
                        // the slot lives one more iteration to splat what the occlusion ray leaves: its
                        // "extension ray" starts far outside every scene box and misses at the root
                        pool.ray_o[slot] = make_float4(3.0e30f, 3.0e30f, 3.0e30f, 0.0f);
                        pool.ray_d[slot] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(FRT_FLAG_ALIVE | FRT_FLAG_FINISH));
                    }
                }
                else
                {
                // ---- emitted: diffuse_light::emitted material.h:184-190 ----
                float3 Le = f3(0.f, 0.f, 0.f);
                if (mtype == FRT_MAT_DIFFUSE_LIGHT && dot(h.n, d) < 0.0f) Le = arr3(m.p);

                if (nonzero(Le))
                {
                    // path.cpp:16-33
                    if (depth == 0 || (flags & FRT_FLAG_PREV_UNWEIGHTED))
                        rad = rad + thr * Le;
                    else
                    {
                        const float cos_wo = dot(h.n, -unit_vector(d));
                        const float3 dp = h.p - f3(pool.prev_p[slot]);
                        float dist2 = dot(dp, dp);
                        if (dist2 <= FRT_EPSILON) dist2 = FRT_EPSILON;
                        float pdf_area;
                        if (is_sphere)
                            pdf_area = sphere_pdf_direct(sph_c, sph_r, h.p, h.t, h.n, d);
                        else // triangle.h:65,139-144
                            pdf_area = 1.0f / (0.5f * length(cross(e1, e2)) * (float)(tri_flags >> 1));
                        const float light_pdf = pdf_area * dist2 / fabsf(cos_wo);
                        rad = rad + thr * Le * mi_weight(prev_pdf, light_pdf);
                    }
                    end_path(a, pool, slot, rad, pixel, nonfinite);
                    terminated = true;
                }
                else if (depth < a.tile.max_depth && mtype != FRT_MAT_DIFFUSE_LIGHT)
                {
                    // path.cpp:36 — scatter
                    const float4 r1 = frt_rng_block(a.tile.seed, pixel, sample, 1 + 2 * depth);
                    const float4 r2 = frt_rng_block(a.tile.seed, pixel, sample, 2 + 2 * depth);
                    ScatterRec srec;
                    mat_scatter(m, h, d, r1.x, r1.y, srec);

                    // ---- continuation first: a zero pdf voids this vertex's NEE too (path.cpp:87-90,103-106)
                    float3 next_o, next_d, f;
                    float pdf;
                    if (srec.is_specular)
                    {
                        pdf = mat_pdf_value(m, h, srec.dir);
                        if (srec.sampled_pdf > 0.0f) pdf = srec.sampled_pdf;
                        f = mat_eval(m, h, srec.dir);
                        const bool outside = dot(h.n, srec.dir) > 0.0f;
                        next_o = outside ? (h.p + FRT_EPSILON * h.n) : (h.p - FRT_EPSILON * h.n);
                        next_d = srec.dir;
                    }
                    else
                    {
                        // cosine_pdf::generate pdf.h:91-94 + hemisphere_to_cosine_direction pdf.h:13-23
                        const float r = sqrtf(r2.z);
                        float sp, cp;
                        sincospif(2.0f * r2.w, &sp, &cp);
                        next_d = Onb(h.n).from_local(f3(r * cp, r * sp, sqrtf(1.0f - r2.z)));
                        next_o = h.p + FRT_EPSILON * h.n;
                        pdf = mat_pdf_value(m, h, next_d);
                        f = mat_eval(m, h, next_d) * fabsf(dot(h.n, unit_vector(next_d)));
                    }

                    if (pdf == 0.0f)
                    {
                        end_path(a, pool, slot, rad, pixel, nonfinite);
                        terminated = true;
                    }
                    else
                    {
                        // ---- next-event estimation path.cpp:39-75 ----
                        const unsigned int nl = a.sc.n_lights;
                        if (nl > 0 && mtype != FRT_MAT_DIELECTRIC)
                        {
                            int index = (int)(r1.w * (float)nl); // hitable_list::pick_sample
                            if (index == (int)nl) index -= 1;
                            const frt_light &L = a.sc.lights[index];
                            const float3 so = h.p + FRT_EPSILON * h.n;
                            const LightSample ls = light_sample_direct(L, so, r2.x, r2.y);
                            const float dist = length(ls.to_light);
                            const float3 tl = ls.to_light / dist; // make_unit_vector
                            float3 bsdf = mat_eval(m, h, tl);
                            const float cos_wi = dot(h.n, unit_vector(tl));
                            const float cos_wo = dot(ls.n, -unit_vector(tl));
                            float3 contrib = f3(0.f, 0.f, 0.f);
                            if (cos_wo != 0.0f)
                            {
                                const float dist2 = dist * dist;
                                if (!srec.is_specular) bsdf = bsdf * cos_wi;
                                float pdf_area;
                                if (L.kind == FRT_LIGHT_TRIANGLE) pdf_area = L.v0[3];
                                else pdf_area = sphere_pdf_direct(arr3(L.v0), L.v0[3], h.p, h.t, h.n, tl);
                                const float light_pdf = pdf_area * dist2 / fabsf(cos_wo);
                                const float bsdf_pdf = mat_pdf_value(m, h, tl);
                                const float weight = mi_weight(light_pdf, bsdf_pdf);
                                const frt_material &lm = a.sc.materials[L.material];
                                float3 emit = f3(0.f, 0.f, 0.f);
                                if (lm.type == FRT_MAT_DIFFUSE_LIGHT && dot(ls.n, tl) < 0.0f) emit = arr3(lm.p);
                                contrib = thr * (emit * bsdf * weight / light_pdf);
                            }
                            // a zero contribution needs no visibility test: adding 0 is the identity
                            if (nonzero(contrib) || !(isfinite(contrib.x) && isfinite(contrib.y) && isfinite(contrib.z)))
                            {
                                want_shadow = true;
                                pool.sh_a[slot] = make_float4(so.x, so.y, so.z, ls.to_light.x);
                                pool.sh_b[slot] = make_float4(ls.to_light.y, ls.to_light.z, contrib.x, contrib.y);
                                pool.sh_c[slot] = contrib.z;
                            }
                        }
                        // ---- continue the path: L = Le + f * Li(next) [* cos] / pdf ----
                        thr = thr * f / pdf;
                        unsigned int nf = FRT_FLAG_ALIVE | ((depth + 1) << FRT_DEPTH_SHIFT);
                        if (mtype == FRT_MAT_PHONG || mtype == FRT_MAT_METAL || mtype == FRT_MAT_DIELECTRIC)
                            nf |= FRT_FLAG_PREV_UNWEIGHTED;
                        pool.ray_o[slot] = make_float4(next_o.x, next_o.y, next_o.z, pdf);
                        pool.ray_d[slot] = make_float4(next_d.x, next_d.y, next_d.z, __uint_as_float(nf));
                        pool.thr[slot] = make_float4(thr.x, thr.y, thr.z, tw.w);
                        pool.prev_p[slot] = make_float4(h.p.x, h.p.y, h.p.z, 0.0f);
                    }
                }
                else
                {
                    // path.cpp:112 — depth cap reached, or the back of a light: returns Le == 0
                    end_path(a, pool, slot, rad, pixel, nonfinite);
                    terminated = true;
                }
                } // integrator == path
            }
        }
        return (terminated ? FRT_SHADE_TERMINATED : 0u) | (want_shadow ? FRT_SHADE_SHADOW : 0u) |
               ((flags & FRT_FLAG_ALIVE) ? FRT_SHADE_WAS_ALIVE : 0u) | (origin_prim << FRT_SHADE_PRIM_SHIFT);
}
