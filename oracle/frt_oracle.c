/* TEST INFRASTRUCTURE — see frt_oracle.h.  Plain C restatement of the reference's hot path; every
 * function cites the reference lines (relative to /root/reference/first_ray/) it follows.
 * Arithmetic type: double (the reference's), or float with -DFRT_ORACLE_F32. */
#include "frt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "frt_rng.h"

#ifdef FRT_ORACLE_F32
typedef float real;
#define R(x) x##f
#define rsqrt_(x) sqrtf(x)
#define rabs(x) fabsf(x)
#define rpow(a, b) powf(a, b)
#define rexp(x) expf(x)
#define rlog(x) logf(x)
#define rsin(x) sinf(x)
#define rcos(x) cosf(x)
#define rtan(x) tanf(x)
#define racos(x) acosf(x)
#define rasin(x) asinf(x)
#define ratan2(a, b) atan2f(a, b)
#define rcopysign(a, b) copysignf(a, b)
#else
typedef double real;
#define R(x) x
#define rsqrt_(x) sqrt(x)
#define rabs(x) fabs(x)
#define rpow(a, b) pow(a, b)
#define rexp(x) exp(x)
#define rlog(x) log(x)
#define rsin(x) sin(x)
#define rcos(x) cos(x)
#define rtan(x) tan(x)
#define racos(x) acos(x)
#define rasin(x) asin(x)
#define ratan2(a, b) atan2(a, b)
#define rcopysign(a, b) copysign(a, b)
#endif

#define EPSILON R(1e-4)                 /* util.h:10 */
#define SHADOW_EPSILON ((real)1e-3f)    /* util.h:11 (a float literal in the reference) */
#define DELTA_EPSILON ((real)1e-3f)     /* util.h:12 */
#define PI ((real)3.14159265358979323846)
#define REAL_FLT_MAX ((real)3.402823466e+38)

int frt_oracle_is_f32(void) { return sizeof(real) == 4; }

typedef struct { real x, y, z; } v3;
static inline v3 V(real x, real y, real z) { v3 r = {x, y, z}; return r; }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 divv(v3 a, v3 b) { return V(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline v3 scl(v3 a, real s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 dvs(v3 a, real s) { return V(a.x / s, a.y / s, a.z / s); }
static inline v3 neg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline real dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 cross(v3 a, v3 b) { return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline real len(v3 a) { return rsqrt_(dot(a, a)); }
static inline v3 unit(v3 a) { return dvs(a, len(a)); } /* geometry.h unit_vector */
static inline int nonzero(v3 a) { return a.x != 0 || a.y != 0 || a.z != 0; }
static inline v3 f3(const float *p) { return V((real)p[0], (real)p[1], (real)p[2]); }

/* environment_map::eval material.h:219-232: unit direction -> phi = atan2(x, -z) (+2 pi if negative), theta = acos(y)
 * -> u = phi / 2 pi, v = theta / pi -> env_map_tex->value.  constant_texture texture.h:17-28 returns its colour;
 * image_texture::value texture.h:60-86 over an HDR image: i = int(u nx), j = int(v ny), wrapped with modulo when
 * outside [0, n], i == nx -> nx - 1 (same for j), the three floats at (j nx + i) 3, no filtering. */
static int modulo_(int a, int b) { const int r = a % b; return r < 0 ? r + b : r; } /* util.h modulo */
static v3 env_eval(const frt_scene_desc *sc, v3 d)
{
    if (!sc->env_image) return f3(sc->env_color);
    const v3 dir = unit(d);
    real phi = ratan2(dir.x, -dir.z), theta = racos(dir.y);
    phi = (phi < 0) ? (phi + PI * 2) : phi;
    theta = (theta < 0) ? (theta + PI) : theta;
    const real u = phi / (R(2.0) * PI), v = theta / PI;
    const int nx = (int)sc->env_nx, ny = (int)sc->env_ny;
    int i = (int)(u * nx), j = (int)(v * ny);
    if (i < 0 || i > nx) i = modulo_(i, nx);
    if (j < 0 || j > ny) j = modulo_(j, ny);
    if (i == nx) i = nx - 1;
    if (j == ny) j = ny - 1;
    return f3(sc->env_image + ((size_t)j * nx + i) * 3);
}
static inline v3 q3(frt_float4 q) { return V((real)q.x, (real)q.y, (real)q.z); }
static inline uint32_t wbits(frt_float4 q) { uint32_t u; memcpy(&u, &q.w, 4); return u; }
static inline real comp(v3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

/* ---- what the intersector leaves in hit_record (hitable.h:23-36) ------------------------- */
#define PRIM_NONE 0xFFFFFFFFu
typedef struct {
    real t, b1, b2;
    uint32_t prim; /* triangle index (leaf order) or sphere index | FRT_CHILD_SPHERE */
    v3 p, wi, n;
    uint32_t mat;
    real u, v; /* hit_record::u, v — texture coordinates */
} hitrec;

typedef struct {
    const frt_scene_desc *sc;
    int use_bvh;
    frt_oracle_stats st;
} octx;

/* parallel_bvh.h:46-51 — t_min = EPSILON scaled by max(|o.x|,|o.y|,|o.z|,EPSILON) when larger */
static real ray_tmin(v3 o)
{
    real m = rabs(o.x);
    if (rabs(o.y) > m) m = rabs(o.y);
    if (rabs(o.z) > m) m = rabs(o.z);
    if (EPSILON > m) m = EPSILON;
    const real scaled = EPSILON * m;
    return scaled > EPSILON ? scaled : EPSILON;
}

/* triangle::hit triangle.h:69-118 (geometry part; the record is completed by fill_hit) */
static int tri_hit(const frt_scene_desc *sc, uint32_t i, v3 o, v3 d, real tmin, real tmax, hitrec *h)
{
    const v3 v0 = q3(sc->tris[3 * (size_t)i]);
    const v3 e1 = sub(q3(sc->tris[3 * (size_t)i + 1]), v0), e2 = sub(q3(sc->tris[3 * (size_t)i + 2]), v0); /* triangle.h:60-61 */
    const v3 hh = cross(d, e2);
    const real a = dot(e1, hh);
    if (a == 0) return 0;
    const real f = R(1.0) / a;
    const v3 s = sub(o, v0);
    const real u = f * dot(s, hh);
    if (u < 0 || u > 1) return 0;
    const v3 q = cross(s, e1);
    const real v = f * dot(d, q);
    if (v >= 0 && u + v <= 1)
    {
        const real t = f * dot(e2, q);
        if (t > tmin && t < tmax) { h->t = t; h->b1 = u; h->b2 = v; h->prim = i; return 1; }
    }
    return 0;
}

/* sphere::hit sphere.h:26-56 */
static int sphere_hit(const frt_scene_desc *sc, uint32_t i, v3 o, v3 d, real tmin, real tmax, hitrec *h)
{
    const frt_sphere *s = &sc->spheres[i];
    const v3 oc = sub(o, V((real)s->cx, (real)s->cy, (real)s->cz));
    const real r = (real)s->radius;
    const real a = dot(d, d), b = dot(oc, d), c = dot(oc, oc) - r * r;
    real disc = b * b - a * c;
    if (disc >= 0)
    {
        disc = rsqrt_(disc);
        real t = (-b - disc) / a;
        if (t < tmin) t = (-b + disc) / a;
        if (t < tmin || t > tmax) return 0;
        h->t = t; h->b1 = 0; h->b2 = 0; h->prim = i | FRT_CHILD_SPHERE;
        return 1;
    }
    return 0;
}

/* aabb::hit aabb.h:14-31 on one child box of a wide node */
static int box_hit(const frt_bvh4_node *nd, int k, v3 o, v3 d, real tmin, real tmax)
{
    const real lo[3] = {(real)nd->lo_x[k], (real)nd->lo_y[k], (real)nd->lo_z[k]};
    const real hi[3] = {(real)nd->hi_x[k], (real)nd->hi_y[k], (real)nd->hi_z[k]};
    for (int i = 0; i < 3; ++i)
    {
        const real invD = R(1.0) / comp(d, i);
        real t0 = (lo[i] - comp(o, i)) * invD, t1 = (hi[i] - comp(o, i)) * invD;
        if (invD < 0) { const real tmp = t0; t0 = t1; t1 = tmp; }
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax < tmin) return 0;
    }
    return 1;
}

/* world->hit(r, EPSILON, tmax): closest hit (any = 0) or first hit found (any = 1).
 * use_bvh = 0 is hitable_list::hit (hitable_list.cpp:4-21) over every primitive; use_bvh = 1 is
 * parallel_bvh_node::hit's recursion (parallel_bvh.h:39-64) over the flattened wide nodes, children
 * in stored order, interval shrinking as hits are found. */
static int world_hit(octx *cx, v3 o, v3 d, real tmax, int any, hitrec *best)
{
    const frt_scene_desc *sc = cx->sc;
    const real tmin = ray_tmin(o);
    int found = 0;
    best->prim = PRIM_NONE;
    if (!cx->use_bvh)
    {
        for (uint32_t i = 0; i < sc->n_tris; ++i)
        {
            cx->st.tri_tests++;
            if (tri_hit(sc, i, o, d, tmin, tmax, best)) { found = 1; tmax = best->t; if (any) return 1; }
        }
        for (uint32_t i = 0; i < sc->n_spheres; ++i)
            if (sphere_hit(sc, i, o, d, tmin, tmax, best)) { found = 1; tmax = best->t; if (any) return 1; }
        return found;
    }
    if (sc->n_nodes == 0) return 0;
    uint32_t stack[256];
    int sp = 0;
    stack[sp++] = 0;
    while (sp)
    {
        const uint32_t cur = stack[--sp];
        if (cur & FRT_CHILD_LEAF)
        {
            const uint32_t first = cur & FRT_LEAF_FIRST_MASK, count = ((cur >> FRT_LEAF_COUNT_SHIFT) & 15u) + 1u;
            for (uint32_t k = 0; k < count; ++k)
            {
                int h;
                if (cur & FRT_CHILD_SPHERE) h = sphere_hit(sc, first + k, o, d, tmin, tmax, best);
                else { cx->st.tri_tests++; h = tri_hit(sc, first + k, o, d, tmin, tmax, best); }
                if (h) { found = 1; tmax = best->t; if (any) return 1; }
            }
            continue;
        }
        const frt_bvh4_node *nd = &sc->nodes[cur];
        cx->st.node_visits++;
        for (int k = 3; k >= 0; --k) /* pushed in reverse so that child 0 is visited first */
            if (nd->child[k] != FRT_CHILD_EMPTY && box_hit(nd, k, o, d, tmin, tmax) && sp < 255)
                stack[sp++] = nd->child[k];
    }
    return found;
}

static int32_t prim_id_of(const frt_scene_desc *sc, uint32_t prim)
{
    if (prim == PRIM_NONE) return -1;
    if (prim & FRT_CHILD_SPHERE) return (int32_t)sc->spheres[prim & FRT_LEAF_FIRST_MASK].prim_id;
    return (int32_t)wbits(sc->tris[3 * (size_t)prim]);
}

/* the rest of the record: triangle.h:92-112 / sphere.h:43-52 */
static void fill_hit(const frt_scene_desc *sc, v3 o, v3 d, hitrec *h)
{
    h->p = add(o, scl(d, h->t));
    h->wi = neg(unit(d));
    if (h->prim & FRT_CHILD_SPHERE)
    {
        const frt_sphere *s = &sc->spheres[h->prim & FRT_LEAF_FIRST_MASK];
        const v3 c = V((real)s->cx, (real)s->cy, (real)s->cz);
        const real r = (real)s->radius;
        h->n = dvs(sub(h->p, c), r);
        const v3 oc = sub(o, c);
        if (dot(oc, oc) < r * r) h->n = neg(h->n);
        h->mat = s->material;
        /* get_sphere_uv hitable.h:15-21, called with the UN-normalised p - center (sphere.h:51) */
        {
            const v3 q = sub(h->p, c);
            const real phi = ratan2(q.z, q.x), theta = rasin(q.y);
            h->u = 1 - (phi + PI) / (2 * PI);
            h->v = (theta + PI / 2) / PI;
        }
        return;
    }
    const size_t i = h->prim;
    const v3 v0 = q3(sc->tris[3 * i]), e1 = sub(q3(sc->tris[3 * i + 1]), v0), e2 = sub(q3(sc->tris[3 * i + 2]), v0);
    h->mat = wbits(sc->tris[3 * i + 1]);
    if (wbits(sc->tris[3 * i + 2]) & FRT_TRI_GEOMETRY_NORMALS)
        h->n = unit(cross(e1, e2));
    else
    {
        const v3 n0 = q3(sc->tri_normals[3 * i]), n1 = q3(sc->tri_normals[3 * i + 1]), n2 = q3(sc->tri_normals[3 * i + 2]);
        h->n = unit(add(add(scl(n0, 1 - h->b1 - h->b2), scl(n1, h->b1)), scl(n2, h->b2)));
    }
    /* triangle.h:104-106: uvhit = (1 - u - v) uv[V0] + u uv[V1] + v uv[V2] */
    h->u = h->v = 0;
    if (sc->tri_uv)
    {
        const float *q = sc->tri_uv + 6 * i;
        const real w0 = 1 - h->b1 - h->b2;
        h->u = w0 * (real)q[0] + h->b1 * (real)q[2] + h->b2 * (real)q[4];
        h->v = w0 * (real)q[1] + h->b1 * (real)q[3] + h->b2 * (real)q[5];
    }
}

/* (int)x as the reference's x86-64 build converts a double (cvttsd2si): NaN and out-of-range give INT_MIN */
static int f2i(real x) { return (x >= (real)-2147483648.0 && x < (real)2147483648.0) ? (int)x : (int)0x80000000; }

/* texture::value(hrec): constant_texture texture.h:17-28, checker_texture texture.h:30-49, image_texture texture.h:60-86 */
static v3 tex_eval(const frt_scene_desc *sc, uint32_t ref, real u, real v)
{
    uint32_t idx = ref - 1u;
    for (int level = 0; level < 8; ++level)
    {
        const frt_texture *t = &sc->textures[idx];
        if (t->kind == FRT_TEX_CHECKER)
        {
            const int x = 2 * modulo_(f2i(u * (real)t->p[0] * 2), 2) - 1, y = 2 * modulo_(f2i(v * (real)t->p[1] * 2), 2) - 1;
            idx = (x * y == 1) ? t->b : t->a;
            continue;
        }
        if (t->kind == FRT_TEX_IMAGE)
        {
            const int nx = (int)t->b, ny = (int)t->c;
            int i = f2i(u * nx), j = f2i(v * ny);
            if (i < 0 || i > nx) i = modulo_(i, nx);
            if (j < 0 || j > ny) j = modulo_(j, ny);
            if (i == nx) i = nx - 1;
            if (j == ny) j = ny - 1;
            return f3(sc->tex_texels + 3 * ((size_t)t->a + (size_t)j * nx + i));
        }
        return f3(t->p);
    }
    return V(0, 0, 0);
}

/* the material of a hit with its textured colours (albedo->value(hrec) etc., material.h:64,98,158,297) put where the
 * constant colours are */
static frt_material material_at(const frt_scene_desc *sc, const hitrec *h)
{
    frt_material m = sc->materials[h->mat];
    uint32_t ta, tb;
    memcpy(&ta, &m.p[FRT_MAT_TEX_A], 4);
    memcpy(&tb, &m.p[FRT_MAT_TEX_B], 4);
    if (sc->n_textures && ta) { const v3 c = tex_eval(sc, ta, h->u, h->v); m.p[0] = (float)c.x; m.p[1] = (float)c.y; m.p[2] = (float)c.z; }
    if (sc->n_textures && tb) { const v3 c = tex_eval(sc, tb, h->u, h->v); m.p[3] = (float)c.x; m.p[4] = (float)c.y; m.p[5] = (float)c.z; }
    return m;
}

int frt_oracle_trace(const frt_scene_desc *scene, const frt_ray *rays, size_t n, int any_hit, int use_bvh,
                     int32_t *prim, double *t, double *u, double *v)
{
    if (!scene || (n && (!rays || !prim))) return FRT_ERR_INVALID;
    octx cx;
    memset(&cx, 0, sizeof cx);
    cx.sc = scene; cx.use_bvh = use_bvh;
    for (size_t i = 0; i < n; ++i)
    {
        hitrec h;
        const v3 o = V((real)rays[i].ox, (real)rays[i].oy, (real)rays[i].oz), d = V((real)rays[i].dx, (real)rays[i].dy, (real)rays[i].dz);
        const int hit = world_hit(&cx, o, d, any_hit ? R(1.0) - SHADOW_EPSILON : REAL_FLT_MAX, any_hit, &h);
        prim[i] = hit ? prim_id_of(scene, h.prim) : -1;
        if (t) t[i] = hit ? (double)h.t : 0.0;
        if (u) u[i] = hit ? (double)h.b1 : 0.0;
        if (v) v[i] = hit ? (double)h.b2 : 0.0;
    }
    return FRT_OK;
}

/* ---- shading helpers ------------------------------------------------------------------------ */
typedef struct { v3 u, v, w; } onb;
/* onb::build_from_w onb.h:18-30 — w is NOT normalised */
static onb onb_from_w(v3 n)
{
    onb b;
    b.w = n;
    if (rabs(n.x) > rabs(n.y)) { const real inv = R(1.0) / rsqrt_(n.x * n.x + n.z * n.z); b.v = V(n.z * inv, 0, -n.x * inv); }
    else { const real inv = R(1.0) / rsqrt_(n.y * n.y + n.z * n.z); b.v = V(0, n.z * inv, -n.y * inv); }
    b.u = cross(b.v, b.w);
    return b;
}
static v3 from_local(const onb *b, v3 a) { return add(add(scl(b->u, a.x), scl(b->v, a.y)), scl(b->w, a.z)); }
static v3 to_local(const onb *b, v3 a) { return V(dot(a, b->u), dot(a, b->v), dot(a, b->w)); }

static real safe_sqrt(real x) { const real s = rsqrt_(x); return s > 0 ? s : 0; } /* util.h:43-46: max(0, sqrt(x)); NaN -> 0 */
static real mi_weight(real a, real b) { a *= a; b *= b; return a / (a + b); }      /* util.h:55-60 */
static v3 reflect_(v3 v, v3 n) { return unit(sub(v, scl(n, 2 * dot(v, n)))); }     /* util.h:73-76 */
/* util.h:79-84 */
static v3 refract_(v3 wi, v3 n, real eta, real cosThetaT)
{
    if (cosThetaT < 0) eta = 1 / eta;
    return unit(sub(scl(n, dot(wi, n) * eta + cosThetaT), scl(wi, eta)));
}
/* util.h:134-144, non-GNU branch (what the author's Windows build ran and what oracle/_ref is
 * compiled with): sin and cos evaluated in float32 */
static void sincos_ref(real theta, real *s, real *c) { *s = (real)sinf((float)theta); *c = (real)cosf((float)theta); }

/* util.h:86-118 */
static real fresnel_dielectric_ext(real cosThetaI_, real *cosThetaT_, real eta)
{
    if (eta == 1) { *cosThetaT_ = -cosThetaI_; return 0; }
    const real scale = (cosThetaI_ > 0) ? 1 / eta : eta;
    const real cosThetaTSqr = 1 - (1 - cosThetaI_ * cosThetaI_) * (scale * scale);
    if (cosThetaTSqr <= 0) { *cosThetaT_ = 0; return 1; }
    const real cosThetaI = rabs(cosThetaI_), cosThetaT = rsqrt_(cosThetaTSqr);
    const real Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
    const real Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
    *cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
    return R(0.5) * (Rs * Rs + Rp * Rp);
}

/* microfacet.h:8-31 */
static v3 fresnel_conductor_exact(real cosThetaI, v3 eta, v3 k)
{
    const real c2 = cosThetaI * cosThetaI, s2 = 1 - c2, s4 = s2 * s2;
    real out[3];
    for (int i = 0; i < 3; ++i)
    {
        const real e = comp(eta, i), kk = comp(k, i);
        const real temp1 = e * e - kk * kk - s2;
        const real a2pb2 = safe_sqrt(temp1 * temp1 + kk * kk * e * e * 4);
        const real a = safe_sqrt((a2pb2 + temp1) * R(0.5));
        const real term1 = a2pb2 + c2, term2 = a * (2 * cosThetaI);
        const real Rs2 = (term1 - term2) / (term1 + term2);
        const real term3 = a2pb2 * c2 + s4, term4 = term2 * s2;
        const real Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
        out[i] = R(0.5) * (Rp2 + Rs2);
    }
    return V(out[0], out[1], out[2]);
}

/* util.h:241-254 */
static real hypot2(real a, real b)
{
    real r;
    if (rabs(a) > rabs(b)) { r = b / a; r = rabs(a) * rsqrt_(1 + r * r); }
    else if (b != 0) { r = a / b; r = rabs(b) * rsqrt_(1 + r * r); }
    else r = 0;
    return r;
}

/* microfacet::smithG1 microfacet.h:47-88; alphaU == alphaV (material.h:251) so projectRoughness
 * (microfacet.h:33-45) returns alpha */
static real smith_g1(v3 v, v3 m, v3 n, real alpha, int beckmann)
{
    const real cosTheta = dot(n, v);
    if (dot(v, m) * cosTheta <= 0) return 0;
    const real temp = 1 - cosTheta * cosTheta;
    if (temp <= 0) return 1;
    const real tanTheta = rsqrt_(temp) / cosTheta;
    if (beckmann)
    {
        const real a = 1 / (alpha * tanTheta);
        if (a >= (real)1.6f) return 1;
        const real aSqr = a * a;
        return ((real)3.535f * a + (real)2.181f * aSqr) / (1 + (real)2.276f * a + (real)2.577f * aSqr);
    }
    return 2 / (1 + hypot2(1, alpha * tanTheta));
}

/* microfacet::eval microfacet.h:90-128 */
static real mf_eval(v3 m, v3 n, real alpha, int beckmann)
{
    const onb b = onb_from_w(n);
    const v3 ml = to_local(&b, m);
    const real cosTheta = ml.z;
    if (cosTheta <= 0) return 0;
    const real cos2 = cosTheta * cosTheta;
    const real ex = ((ml.x * ml.x) / (alpha * alpha) + (ml.y * ml.y) / (alpha * alpha)) / cos2;
    real result;
    if (beckmann) result = rexp(-ex) / (PI * alpha * alpha * cos2 * cos2);
    else { const real root = (1 + ex) * cos2; result = 1 / (PI * alpha * alpha * root * root); }
    if (result * cosTheta < (real)1e-20f) result = 0;
    return result;
}

/* util.h:185-221 */
static real erfinv_(real x)
{
    real w = -rlog((1 - x) * (1 + x)), p;
    if (w < 5)
    {
        w = w - R(2.5);
        p = R(2.81022636e-08);
        p = R(3.43273939e-07) + p * w; p = R(-3.5233877e-06) + p * w; p = R(-4.39150654e-06) + p * w;
        p = R(0.00021858087) + p * w; p = R(-0.00125372503) + p * w; p = R(-0.00417768164) + p * w;
        p = R(0.246640727) + p * w; p = R(1.50140941) + p * w;
    }
    else
    {
        w = rsqrt_(w) - 3;
        p = R(-0.000200214257);
        p = R(0.000100950558) + p * w; p = R(0.00134934322) + p * w; p = R(-0.00367342844) + p * w;
        p = R(0.00573950773) + p * w; p = R(-0.0076224613) + p * w; p = R(0.00943887047) + p * w;
        p = R(1.00167406) + p * w; p = R(2.83297682) + p * w;
    }
    return p * x;
}

/* util.h:223-239 */
static real erf_(real x)
{
    const real sign = (real)copysignf(1.0f, (float)x);
    x = rabs(x);
    const real t = 1 / (1 + R(0.3275911) * x);
    const real y = 1 - (((((R(1.061405429) * t + R(-1.453152027)) * t) + R(1.421413741)) * t + R(-0.284496736)) * t + R(0.254829592)) * t * rexp(-x * x);
    return sign * y;
}

/* roughconductor_pdf::sampleVisible11 pdf.h:280-403 */
static void sample_visible_11(real thetaI, real sx, real sy, int beckmann, real *slope_x, real *slope_y)
{
    const real SQRT_PI_INV = 1 / rsqrt_(PI);
    if (beckmann)
    {
        if (thetaI < (real)1e-4f)
        {
            real sp, cp;
            const real r = rsqrt_(-rlog(1 - sx));
            sincos_ref(2 * PI * sy, &sp, &cp);
            *slope_x = r * cp; *slope_y = r * sp;
            return;
        }
        const real tanThetaI = rtan(thetaI), cotThetaI = 1 / tanThetaI;
        real a = -1, c = erf_(cotThetaI);
        const real sample_x = sx > (real)1e-6f ? sx : (real)1e-6f;
        const real fit = 1 + thetaI * ((real)-0.876f + thetaI * ((real)0.4265f - (real)0.0594f * thetaI));
        real b = c - (1 + c) * rpow(1 - sample_x, fit);
        const real normalization = 1 / (1 + c + SQRT_PI_INV * tanThetaI * rexp(-cotThetaI * cotThetaI));
        int it = 0;
        while (++it < 10)
        {
            if (!(b >= a && b <= c)) b = R(0.5) * (a + c);
            const real invErf = erfinv_(b);
            const real value = normalization * (1 + b + SQRT_PI_INV * tanThetaI * rexp(-invErf * invErf)) - sample_x;
            const real derivative = normalization * (1 - invErf * tanThetaI);
            if (rabs(value) < (real)1e-5f) break;
            if (value > 0) c = b; else a = b;
            b -= value / derivative;
        }
        *slope_x = erfinv_(b);
        *slope_y = erfinv_(2 * (sy > (real)1e-6f ? sy : (real)1e-6f) - 1);
        return;
    }
    if (thetaI < (real)1e-4f)
    {
        real sp, cp;
        const real r = safe_sqrt(sx / (1 - sx));
        sincos_ref(2 * PI * sy, &sp, &cp);
        *slope_x = r * cp; *slope_y = r * sp;
        return;
    }
    const real tanThetaI = rtan(thetaI);
    const real a = 1 / tanThetaI;
    const real G1 = 2 / (1 + safe_sqrt(1 + 1 / (a * a)));
    real A = 2 * sx / G1 - 1;
    if (rabs(A) == 1) A -= (real)copysignf(1.0f, (float)A) * EPSILON;
    const real tmp = 1 / (A * A - 1);
    const real B = tanThetaI;
    const real D = safe_sqrt(B * B * tmp * tmp - (A * A - B * B) * tmp);
    const real slope_x_1 = B * tmp - D, slope_x_2 = B * tmp + D;
    *slope_x = (A < 0 || slope_x_2 > 1 / tanThetaI) ? slope_x_1 : slope_x_2;
    real S;
    if (sy > R(0.5)) { S = 1; sy = 2 * (sy - R(0.5)); }
    else { S = -1; sy = 2 * (R(0.5) - sy); }
    const real z = (sy * (sy * (sy * (-R(0.365728915865723)) + R(0.790235037209296)) - R(0.424965825137544)) + R(0.000152998850436920)) /
                   (sy * (sy * (sy * (sy * R(0.169507819808272) - R(0.397203533833404)) - R(0.232500544458471)) + 1) - R(0.539825872510702));
    *slope_y = S * z * rsqrt_(1 + *slope_x * *slope_x);
}

/* roughconductor_pdf::sampleVisible pdf.h:405-453 */
static v3 sample_visible(v3 wi_local, real alpha, real sx, real sy, int beckmann)
{
    const v3 wi = unit(V(alpha * wi_local.x, alpha * wi_local.y, wi_local.z));
    real theta = 0, phi = 0;
    if (wi.z < R(0.99999)) { theta = racos(wi.z); phi = ratan2(wi.y, wi.x); }
    real sinPhi, cosPhi;
    sincos_ref(phi, &sinPhi, &cosPhi);
    real slx, sly;
    sample_visible_11(theta, sx, sy, beckmann, &slx, &sly);
    if (!isfinite(slx)) slx = 0;
    const real rx = cosPhi * slx - sinPhi * sly, ry = sinPhi * slx + cosPhi * sly;
    slx = rx * alpha; sly = ry * alpha;
    const real normalization = 1 / rsqrt_(slx * slx + sly * sly + 1);
    return V(-slx * normalization, -sly * normalization, normalization);
}

/* ---- materials: scatter / pdf value / eval_bsdf --------------------------------------------- */
typedef struct { int is_specular; v3 dir; real sampled_pdf; } scatrec; /* scatter_record pdf.h:60-70 */
static v3 mp3(const frt_material *m, int at) { return V((real)m->p[at], (real)m->p[at + 1], (real)m->p[at + 2]); }

/* material::scatter; returns 0 for diffuse_light (material.h:183).  s0,s1 = sample[0], sample[1] */
static int mat_scatter(const frt_material *m, const hitrec *h, v3 d_in, real s0, real s1, scatrec *sr)
{
    sr->sampled_pdf = -1; /* pdf.h:69 */
    sr->dir = V(0, 0, 0);
    switch (m->type)
    {
    case FRT_MAT_LAMBERTIAN: /* material.h:55-60 */
        sr->is_specular = 0;
        return 1;
    case FRT_MAT_PHONG: /* material.h:82-93, cosine_power_pdf::generate pdf.h:116-131 */
    {
        sr->is_specular = 1;
        const v3 Rr = reflect_(neg(h->wi), h->n);
        const real e = (real)m->p[6];
        const real sinAlpha = rsqrt_(1 - rpow(s1, 2 / (e + 1))), cosAlpha = rpow(s1, 1 / (e + 1));
        const real phi = (2 * PI) * s0;
        const onb b = onb_from_w(Rr);
        sr->dir = from_local(&b, V(sinAlpha * rcos(phi), sinAlpha * rsin(phi), cosAlpha));
        return 1;
    }
    case FRT_MAT_DIELECTRIC: /* material.h:139-146, dielectric_pdf::generate pdf.h:163-180 */
    {
        sr->is_specular = 1;
        real cosThetaT;
        const real F = fresnel_dielectric_ext(dot(h->wi, h->n), &cosThetaT, (real)m->p[6]);
        sr->dir = (s0 <= F) ? reflect_(neg(h->wi), h->n) : refract_(h->wi, h->n, (real)m->p[6], cosThetaT);
        return 1;
    }
    case FRT_MAT_ROUGH_CONDUCTOR: /* material.h:262-270, roughconductor_pdf::generate pdf.h:464-480, pdfVisible pdf.cpp:5-12 */
    {
        sr->is_specular = 1;
        const real alpha = (real)m->p[6];
        const int beck = m->p[10] != 0.0f;
        const onb uvw = onb_from_w(h->n);
        const v3 wi_l = to_local(&uvw, h->wi);
        const v3 ml = sample_visible(wi_l, alpha, s0, s1, beck);
        const v3 mw = from_local(&uvw, ml);
        real pdf = 0;
        if (wi_l.z != 0)
            pdf = smith_g1(from_local(&uvw, wi_l), mw, h->n, alpha, beck) * rabs(dot(wi_l, ml)) * mf_eval(mw, h->n, alpha, beck) / rabs(wi_l.z);
        const v3 wo = reflect_(neg(h->wi), mw);
        sr->sampled_pdf = pdf / (4 * dot(wo, mw));
        sr->dir = unit(wo);
        return 1;
    }
    case FRT_MAT_METAL: /* material.h:115-122 */
        sr->is_specular = 1;
        sr->dir = reflect_(unit(d_in), h->n);
        return 1;
    default: /* diffuse_light */
        return 0;
    }
}

/* srec.pdf_ptr->value(hrec, wo) */
static real mat_pdf_value(const frt_material *m, const hitrec *h, v3 wo)
{
    switch (m->type)
    {
    case FRT_MAT_LAMBERTIAN: /* cosine_pdf::value pdf.h:84-89 */
    {
        const real c = dot(h->n, unit(wo));
        return (c > 0 ? c : 0) / PI;
    }
    case FRT_MAT_PHONG: /* cosine_power_pdf::value pdf.h:104-114 */
    {
        if (dot(h->n, wo) <= 0 || dot(h->n, h->wi) <= 0) return 0;
        real alpha = dot(reflect_(neg(h->wi), h->n), wo);
        if (!(alpha > 0)) alpha = 0;
        return rpow(alpha, (real)m->p[6]) * ((real)m->p[6] + 1) / (2 * PI);
    }
    case FRT_MAT_DIELECTRIC: /* dielectric_pdf::value pdf.h:142-161 */
    {
        real cosThetaT;
        const real F = fresnel_dielectric_ext(dot(h->wi, h->n), &cosThetaT, (real)m->p[6]);
        if (dot(h->wi, h->n) * dot(wo, h->n) >= 0)
        {
            if (rabs(dot(reflect_(neg(h->wi), h->n), wo) - 1) > DELTA_EPSILON) return 0;
            return F;
        }
        if (rabs(dot(refract_(h->wi, h->n, (real)m->p[6], cosThetaT), wo) - 1) > DELTA_EPSILON) return 0;
        return 1 - F;
    }
    case FRT_MAT_ROUGH_CONDUCTOR: /* roughconductor_pdf::value pdf.h:237-250 */
    {
        if (dot(h->n, wo) <= 0 || dot(h->n, h->wi) <= 0) return 0;
        const v3 H = unit(add(wo, h->wi));
        const int beck = m->p[10] != 0.0f;
        return mf_eval(H, h->n, (real)m->p[6], beck) * smith_g1(h->wi, H, h->n, (real)m->p[6], beck) / (4 * dot(h->wi, h->n));
    }
    default: /* constant_pdf(1.0) pdf.h:186-201 */
        return 1;
    }
}

/* material::eval_bsdf(r_in, hrec, wo) */
static v3 mat_eval(const frt_material *m, const hitrec *h, v3 wo)
{
    switch (m->type)
    {
    case FRT_MAT_LAMBERTIAN: /* material.h:62-65 */
        return dvs(mp3(m, 0), PI);
    case FRT_MAT_PHONG: /* material.h:95-103 */
    {
        const v3 reflected = unit(reflect_(neg(h->wi), h->n));
        real alpha = dot(reflected, wo);
        if (!(alpha > 0)) alpha = 0;
        const real e = (real)m->p[6];
        const v3 result = add(dvs(mp3(m, 0), PI), scl(mp3(m, 3), ((e + 2) / (2 * PI)) * rpow(alpha, e)));
        return scl(result, dot(h->n, wo));
    }
    case FRT_MAT_DIELECTRIC: /* material.h:147-172 */
    {
        real cosThetaT;
        const real ior = (real)m->p[6];
        const real F = fresnel_dielectric_ext(dot(h->wi, h->n), &cosThetaT, ior);
        if (dot(h->wi, h->n) * dot(wo, h->n) >= 0)
        {
            if (rabs(dot(reflect_(neg(h->wi), h->n), wo) - 1) > DELTA_EPSILON) return V(0, 0, 0);
            return scl(mp3(m, 0), F);
        }
        if (rabs(dot(refract_(h->wi, h->n, ior, cosThetaT), wo) - 1) > DELTA_EPSILON) return V(0, 0, 0);
        const real factor = cosThetaT < 0 ? (1 / ior) : ior;
        return scl(mp3(m, 0), factor * factor * (1 - F));
    }
    case FRT_MAT_ROUGH_CONDUCTOR: /* material.h:276-306 */
    {
        const real cosWi = dot(h->wi, h->n);
        if (cosWi <= 0 || dot(wo, h->n) <= 0) return V(0, 0, 0);
        const v3 H = unit(add(wo, h->wi));
        const real alpha = (real)m->p[6];
        const int beck = m->p[10] != 0.0f;
        const real D = mf_eval(H, h->n, alpha, beck);
        if (D == 0) return V(0, 0, 0);
        const v3 F = mul(fresnel_conductor_exact(dot(h->wi, H), mp3(m, 3), mp3(m, 7)), mp3(m, 0));
        const real G = smith_g1(h->wi, H, h->n, alpha, beck) * smith_g1(wo, H, h->n, alpha, beck);
        return scl(F, D * G / (4 * cosWi));
    }
    case FRT_MAT_METAL: /* material.h:123-126 */
        return mp3(m, 0);
    default:
        return V(0, 0, 0);
    }
}

/* ---- lights --------------------------------------------------------------------------------- */
typedef struct { v3 to_light, n; } lightsample;

/* triangle::sample_direct triangle.h:145-175 / sphere::sample_direct sphere.h:80-107 */
static lightsample light_sample_direct(const frt_light *L, v3 o, real u0, real u1)
{
    lightsample ls;
    if (L->kind == FRT_LIGHT_TRIANGLE)
    {
        const v3 v0 = f3(L->v0), v1 = f3(L->v1), v2 = f3(L->v2);
        const real su0 = rsqrt_(u0), b0 = 1 - su0, b1 = u1 * su0;
        const v3 pt = add(add(scl(v0, 1 - b0 - b1), scl(v1, b0)), scl(v2, b1));
        if (L->flags & FRT_TRI_GEOMETRY_NORMALS) ls.n = unit(cross(sub(v1, v0), sub(v2, v0)));
        else ls.n = unit(add(add(scl(f3(L->n0), 1 - b0 - b1), scl(f3(L->n1), b0)), scl(f3(L->n2), b1)));
        ls.to_light = sub(pt, o);
        return ls;
    }
    const v3 c = f3(L->v0);
    const real radius = (real)L->v0[3];
    const v3 direction = sub(c, o);
    const real d2 = dot(direction, direction);
    if (d2 <= radius * radius)
    {
        /* uniform_sample_sphere pdf.h:38-44 */
        const real z = 1 - 2 * u0;
        real r = 1 - z * z;
        r = rsqrt_(r > 0 ? r : 0);
        const real phi = 2 * PI * u1;
        const v3 p = add(c, scl(V(r * rcos(phi), r * rsin(phi), z), radius));
        ls.n = unit(sub(c, p));
        ls.to_light = sub(p, o);
        return ls;
    }
    /* random_to_sphere pdf.h:46-56 through onb(center - o), whose u and w are not unit (sphere.h:84-85,100) */
    const real z = 1 + u1 * (rsqrt_(1 - radius * radius / d2) - 1);
    const real phi = 2 * PI * u0;
    const real s = rsqrt_(1 - z * z);
    const onb b = onb_from_w(direction);
    const v3 p = from_local(&b, V(rcos(phi) * s, rsin(phi) * s, z));
    ls.n = unit(p);
    ls.to_light = p;
    return ls;
}

/* sphere::pdf_direct_sampling sphere.h:64-78 (called with the SURFACE record, path.cpp:30,65) */
static real sphere_pdf_direct(v3 c, real radius, v3 rec_p, real rec_t, v3 rec_n, v3 to_light)
{
    const v3 o = add(rec_p, scl(to_light, -rec_t));
    const v3 direction = sub(c, o);
    const real d2 = dot(direction, direction), r2 = radius * radius;
    if (d2 <= r2) return 1 / (4 * PI * radius * radius);
    const real cos_theta_max = rsqrt_(1 - r2 / d2);
    const real solid_angle = 2 * PI * (1 - cos_theta_max);
    return (1 / solid_angle) * rabs(dot(to_light, rec_n)) / d2;
}

/* obj->pdf_direct_sampling for the primitive a ray hit (path.cpp:30) */
static real hit_pdf_direct(const frt_scene_desc *sc, const hitrec *h, v3 d)
{
    if (h->prim & FRT_CHILD_SPHERE)
    {
        const frt_sphere *s = &sc->spheres[h->prim & FRT_LEAF_FIRST_MASK];
        return sphere_pdf_direct(V((real)s->cx, (real)s->cy, (real)s->cz), (real)s->radius, h->p, h->t, h->n, d);
    }
    const size_t i = h->prim;
    const v3 v0 = q3(sc->tris[3 * i]), e1 = sub(q3(sc->tris[3 * i + 1]), v0), e2 = sub(q3(sc->tris[3 * i + 2]), v0);
    const real n_tris = (real)(wbits(sc->tris[3 * i + 2]) >> 1);
    return 1 / (R(0.5) * len(cross(e1, e2)) * n_tris); /* triangle.h:65,139-144 */
}

/* ---- sampler: the counter-based stand-in of sampler.h:9-35 (layout in frt_rng.h) ------------- */
typedef struct { uint32_t seed, pixel, sample; int vertex, cam_1d, n2d, has_lights; } sampler;
static real sdim(const sampler *s, uint32_t block, int lane)
{
    uint32_t r[4];
    frt_rng_block(s->seed, s->pixel, s->sample, block, r);
    return (real)frt_rng_unit(r[lane]);
}
static real get1d(sampler *s) { return s->vertex < 0 ? sdim(s, 0, s->cam_1d++) : sdim(s, 1 + 2 * s->vertex, 3); }
static void get2d(sampler *s, real *a, real *b)
{
    if (s->vertex < 0) { *a = sdim(s, 0, 2); *b = sdim(s, 0, 3); return; }
    int k = s->n2d++;
    if (!s->has_lights) k += 1;
    const uint32_t block = (uint32_t)(2 + 2 * s->vertex + (k >> 1)); /* k >= 2 only in Li_direct (one draw per light) */
    *a = sdim(s, block, 2 * (k & 1)); *b = sdim(s, block, 2 * (k & 1) + 1);
}
static void get3d(sampler *s, real *a, real *b, real *c)
{
    ++s->vertex; s->n2d = 0;
    *a = sdim(s, 1 + 2 * s->vertex, 0); *b = sdim(s, 1 + 2 * s->vertex, 1); *c = sdim(s, 1 + 2 * s->vertex, 2);
}

/* ---- path::Li path.cpp:4-116 ------------------------------------------------------------------ */
static v3 Li(octx *cx, v3 o, v3 d, int depth, const hitrec *prev, uint32_t prev_mat_type, real prev_bsdf_pdf,
             int max_depth, sampler *rs)
{
    const frt_scene_desc *sc = cx->sc;
    hitrec h;
    cx->st.closest_rays++;
    if (!world_hit(cx, o, d, REAL_FLT_MAX, 0, &h)) /* path.cpp:10 */
        return env_eval(sc, d);                    /* path.cpp:115, environment_map::eval material.h:219-232 */
    fill_hit(sc, o, d, &h);
    const frt_material mm = material_at(sc, &h);
    const frt_material *m = &mm;
    /* diffuse_light::emitted material.h:184-190 */
    v3 Le = V(0, 0, 0);
    if (m->type == FRT_MAT_DIFFUSE_LIGHT && dot(h.n, d) < 0) Le = mp3(m, 0);
    if (nonzero(Le)) /* path.cpp:16-33 */
    {
        if (depth == 0 || prev_mat_type == FRT_MAT_PHONG || prev_mat_type == FRT_MAT_METAL || prev_mat_type == FRT_MAT_DIELECTRIC)
            return Le;
        const real cos_wo = dot(h.n, neg(unit(d)));
        const v3 dp = sub(h.p, prev->p);
        real distance_squared = dot(dp, dp);
        if (distance_squared <= EPSILON) distance_squared = EPSILON;
        const real light_pdf = hit_pdf_direct(sc, &h, d) * distance_squared / rabs(cos_wo);
        return scl(Le, mi_weight(prev_bsdf_pdf, light_pdf));
    }
    scatrec sr;
    real s0, s1, s2;
    if (depth < max_depth && m->type != FRT_MAT_DIFFUSE_LIGHT) /* path.cpp:36: depth <= 33 && scatter() */
    {
        get3d(rs, &s0, &s1, &s2);
        (void)s2;
        mat_scatter(m, &h, d, s0, s1, &sr);
        /* hitable_list::pick_sample hitable_list.cpp:64-70 */
        int index = (int)(get1d(rs) * (real)sc->n_lights);
        if (index == (int)sc->n_lights) index -= 1;
        if (index >= 0 && m->type != FRT_MAT_DIELECTRIC) /* path.cpp:40 */
        {
            const frt_light *L = &sc->lights[index];
            const v3 offset_origin = add(h.p, scl(h.n, EPSILON));
            real u0, u1;
            get2d(rs, &u0, &u1);
            const lightsample ls = light_sample_direct(L, offset_origin, u0, u1);
            v3 to_light = ls.to_light;
            const real dist_to_light = len(to_light);
            hitrec occ;
            cx->st.shadow_rays++;
            if (!world_hit(cx, offset_origin, to_light, R(1.0) - SHADOW_EPSILON, 1, &occ)) /* path.cpp:50 */
            {
                to_light = dvs(to_light, dist_to_light); /* make_unit_vector */
                v3 surface_bsdf = mat_eval(m, &h, to_light);
                const real cos_wi = dot(h.n, unit(to_light));
                const real cos_wo = dot(ls.n, neg(unit(to_light)));
                if (cos_wo != 0)
                {
                    const real distance_squared = dist_to_light * dist_to_light;
                    if (!sr.is_specular) surface_bsdf = scl(surface_bsdf, cos_wi);
                    real pdf_area;
                    if (L->kind == FRT_LIGHT_TRIANGLE) pdf_area = (real)L->v0[3];
                    else pdf_area = sphere_pdf_direct(f3(L->v0), (real)L->v0[3], h.p, h.t, h.n, to_light);
                    const real light_pdf = pdf_area * distance_squared / rabs(cos_wo);
                    const real surface_bsdf_pdf = mat_pdf_value(m, &h, to_light);
                    const real weight = mi_weight(light_pdf, surface_bsdf_pdf);
                    const frt_material *lm = &sc->materials[L->material];
                    v3 emit = V(0, 0, 0);
                    if (lm->type == FRT_MAT_DIFFUSE_LIGHT && dot(ls.n, to_light) < 0) emit = mp3(lm, 0);
                    Le = add(Le, dvs(scl(mul(emit, surface_bsdf), weight), light_pdf)); /* path.cpp:73 */
                }
            }
        }
        if (sr.is_specular) /* path.cpp:76-95 */
        {
            real surface_bsdf_pdf = mat_pdf_value(m, &h, sr.dir);
            if (sr.sampled_pdf > 0) surface_bsdf_pdf = sr.sampled_pdf;
            const v3 surface_bsdf = mat_eval(m, &h, sr.dir);
            if (surface_bsdf_pdf == 0) return V(0, 0, 0);
            const int outside = dot(h.n, sr.dir) > 0;
            const v3 no = outside ? add(h.p, scl(h.n, EPSILON)) : sub(h.p, scl(h.n, EPSILON));
            const v3 Lnext = Li(cx, no, sr.dir, depth + 1, &h, m->type, surface_bsdf_pdf, max_depth, rs);
            return add(Le, dvs(mul(surface_bsdf, Lnext), surface_bsdf_pdf));
        }
        /* path.cpp:96-110: cosine_pdf::generate pdf.h:91-94, hemisphere_to_cosine_direction pdf.h:13-23 */
        real r0, r1;
        get2d(rs, &r0, &r1);
        const real r = rsqrt_(r0), phi = 2 * PI * r1;
        const onb b = onb_from_w(h.n);
        const v3 wo = from_local(&b, V(r * rcos(phi), r * rsin(phi), rsqrt_(1 - r0)));
        const v3 no = add(h.p, scl(h.n, EPSILON));
        const real surface_bsdf_pdf = mat_pdf_value(m, &h, wo);
        const v3 surface_bsdf = mat_eval(m, &h, wo);
        if (surface_bsdf_pdf == 0) return V(0, 0, 0);
        const real cos_wo = rabs(dot(h.n, unit(wo)));
        const v3 Lnext = Li(cx, no, wo, depth + 1, &h, m->type, surface_bsdf_pdf, max_depth, rs);
        return add(Le, dvs(scl(mul(surface_bsdf, Lnext), cos_wo), surface_bsdf_pdf));
    }
    return Le; /* path.cpp:112 */
}

/* ---- ao::Li ao.cpp:4-28 ------------------------------------------------------------------------- */
static v3 Li_ao(octx *cx, v3 o, v3 d, sampler *rs)
{
    const frt_scene_desc *sc = cx->sc;
    hitrec h;
    cx->st.closest_rays++;
    if (world_hit(cx, o, d, REAL_FLT_MAX, 0, &h)) /* ao.cpp:9 */
    {
        fill_hit(sc, o, d, &h);
        const frt_material mm = material_at(sc, &h);
        const frt_material *m = &mm;
        real s0, s1, s2;
        scatrec sr;
        get3d(rs, &s0, &s1, &s2);
        (void)s2;
        if (mat_scatter(m, &h, d, s0, s1, &sr)) /* ao.cpp:13 */
        {
            /* srec.pdf_ptr->generate(get2d(), srec) ao.cpp:17: cosine_pdf::generate pdf.h:91-94 for lambertian;
             * the specular pdfs' generate is what their scatter calls with (sample[0], sample[1]) */
            real r0, r1;
            v3 dir;
            get2d(rs, &r0, &r1);
            if (sr.is_specular)
            {
                scatrec again;
                mat_scatter(m, &h, d, r0, r1, &again);
                dir = again.dir;
            }
            else
            {
                const real r = rsqrt_(r0), phi = 2 * PI * r1;
                const onb b = onb_from_w(h.n);
                dir = from_local(&b, V(r * rcos(phi), r * rsin(phi), rsqrt_(1 - r0)));
            }
            /* t_max = bbox.size.y() * 0.50f ao.cpp:19-22; ray from hrec.p, not offset */
            hitrec occ;
            cx->st.shadow_rays++;
            if (world_hit(cx, h.p, dir, (real)sc->ao_t_max, 1, &occ)) return V(0, 0, 0);
        }
    }
    return env_eval(sc, d); /* ao.cpp:27: evaluated for the camera ray */
}

/* ---- normals_renderer::Li debug_renderer.h:8-17 ------------------------------------------------- */
/* ---- ass_1_renderer::Li assignment_1.cpp:4-38: every light sampled once, NO visibility test, area-measure pdf -- */
static v3 Li_direct(octx *cx, v3 o, v3 d, sampler *rs)
{
    const frt_scene_desc *sc = cx->sc;
    hitrec h;
    cx->st.closest_rays++;
    if (!world_hit(cx, o, d, REAL_FLT_MAX, 0, &h)) return env_eval(sc, d);
    fill_hit(sc, o, d, &h);
    const frt_material mm = material_at(sc, &h);
    const frt_material *m = &mm;
    v3 Le = V(0, 0, 0);
    if (m->type == FRT_MAT_DIFFUSE_LIGHT) return Le; /* scatter() false: assignment_1.cpp:15 */
    real s0, s1, s2;
    get3d(rs, &s0, &s1, &s2); /* consumed by scatter(); its record is not used */
    (void)s0; (void)s1; (void)s2;
    const v3 offset_origin = add(h.p, scl(h.n, EPSILON));
    for (uint32_t i = 0; i < sc->n_lights; ++i)
    {
        const frt_light *L = &sc->lights[i];
        real u0, u1;
        get2d(rs, &u0, &u1);
        const lightsample ls = light_sample_direct(L, offset_origin, u0, u1);
        const v3 to_light = ls.to_light; /* un-normalised throughout: assignment_1.cpp:22-31 */
        const v3 surface_bsdf = mat_eval(m, &h, to_light);
        real cos_wi = dot(h.n, unit(to_light));
        if (!(cos_wi > 0)) cos_wi = 0; /* std::max(0.0, x) */
        real light_pdf;
        if (L->kind == FRT_LIGHT_TRIANGLE) light_pdf = (real)L->v0[3];
        else light_pdf = sphere_pdf_direct(f3(L->v0), (real)L->v0[3], h.p, h.t, h.n, to_light);
        const frt_material *lm = &sc->materials[L->material];
        v3 emit = V(0, 0, 0);
        if (lm->type == FRT_MAT_DIFFUSE_LIGHT && dot(ls.n, to_light) < 0) emit = mp3(lm, 0);
        Le = add(Le, dvs(scl(mul(surface_bsdf, emit), cos_wi), light_pdf));
    }
    return Le;
}

static v3 Li_normals(octx *cx, v3 o, v3 d)
{
    hitrec h;
    cx->st.closest_rays++;
    if (world_hit(cx, o, d, REAL_FLT_MAX, 0, &h))
    {
        fill_hit(cx->sc, o, d, &h);
        return h.n;
    }
    return env_eval(cx->sc, d);
}

/* ---- path::Render's sample loop path.cpp:129-143 over a tile; camera::get_ray camera.h:30-35 -- */
int frt_oracle_render(const frt_scene_desc *scene, const frt_tile *t, int use_bvh, double *rgb_sum, frt_oracle_stats *stats)
{
    if (!scene || !t || !rgb_sum) return FRT_ERR_INVALID;
    if (t->x0 + t->w > t->nx || t->y0 + t->h > t->ny) return FRT_ERR_INVALID;
    octx cx;
    memset(&cx, 0, sizeof cx);
    cx.sc = scene; cx.use_bvh = use_bvh;
    const frt_camera *c = &scene->camera;
    const v3 origin = f3(c->origin), llc = f3(c->lower_left_corner), hor = f3(c->horizontal), ver = f3(c->vertical);
    const v3 cu = f3(c->u), cv = f3(c->v);
    const real lens_radius = (real)c->lens_radius;
    hitrec none;
    memset(&none, 0, sizeof none);
    for (uint32_t y = t->y0; y < t->y0 + t->h; ++y)
        for (uint32_t x = t->x0; x < t->x0 + t->w; ++x)
        {
            double col[3] = {0, 0, 0};
            for (uint32_t s = t->spp_begin; s < t->spp_end; ++s)
            {
                sampler rs;
                rs.seed = t->seed; rs.pixel = y * t->nx + x; rs.sample = s;
                rs.vertex = -1; rs.cam_1d = 0; rs.n2d = 0; rs.has_lights = scene->n_lights > 0;
                const real u = ((real)x + get1d(&rs)) / (real)t->nx;
                const real v = ((real)y + get1d(&rs)) / (real)t->ny;
                real l0, l1, rdx = 0, rdy = 0;
                get2d(&rs, &l0, &l1);
                if (lens_radius != 0)
                {
                    /* random_in_unit_disk util.h:21-41 */
                    const real ax = l0 * 2 - 1, ay = l1 * 2 - 1;
                    if (!(ax == 0 && ay == 0))
                    {
                        real phi, rr;
                        if (ax * ax > ay * ay) { rr = ax; phi = (PI / 4) * (ay / ax); }
                        else { rr = ay; phi = (PI / 2) - (PI / 4) * (ax / ay); }
                        rdx = lens_radius * rr * rcos(phi);
                        rdy = lens_radius * rr * rsin(phi);
                    }
                }
                const v3 offset = add(scl(cu, rdx), scl(cv, rdy));
                const v3 ro = add(origin, offset);
                const v3 rd = sub(sub(add(add(llc, scl(hor, u)), scl(ver, v)), origin), offset);
                cx.st.samples++;
                const v3 L = t->integrator == FRT_INTEGRATOR_AO ? Li_ao(&cx, ro, rd, &rs)
                             : t->integrator == FRT_INTEGRATOR_NORMALS ? Li_normals(&cx, ro, rd)
                             : t->integrator == FRT_INTEGRATOR_DIRECT ? Li_direct(&cx, ro, rd, &rs)
                             : Li(&cx, ro, rd, 0, &none, 0xFFFFFFFFu, 0, (int)t->max_depth, &rs);
                if (!(isfinite(L.x) && isfinite(L.y) && isfinite(L.z))) { cx.st.nonfinite++; continue; }
                col[0] += (double)L.x; col[1] += (double)L.y; col[2] += (double)L.z;
            }
            double *px = rgb_sum + 3 * ((size_t)y * t->nx + x);
            px[0] += col[0]; px[1] += col[1]; px[2] += col[2];
        }
    if (stats) *stats = cx.st;
    return FRT_OK;
}

int frt_oracle_bsdf(const frt_material *m, const double n_[3], const double wi_[3], const double s[3],
                    const double wo_probe[3], double out[12])
{
    if (!m || !n_ || !wi_ || !s || !wo_probe || !out) return FRT_ERR_INVALID;
    hitrec h;
    memset(&h, 0, sizeof h);
    h.n = V((real)n_[0], (real)n_[1], (real)n_[2]);
    h.wi = V((real)wi_[0], (real)wi_[1], (real)wi_[2]);
    h.t = 1;
    scatrec sr;
    if (!mat_scatter(m, &h, neg(h.wi), (real)s[0], (real)s[1], &sr)) { memset(out, 0, 12 * sizeof(double)); return FRT_OK; }
    v3 dir = sr.dir;
    if (!sr.is_specular)
    {
        const real r = rsqrt_((real)s[0]), phi = 2 * PI * (real)s[1];
        const onb b = onb_from_w(h.n);
        dir = from_local(&b, V(r * rcos(phi), r * rsin(phi), rsqrt_(1 - (real)s[0])));
    }
    const v3 probe = V((real)wo_probe[0], (real)wo_probe[1], (real)wo_probe[2]);
    const v3 f = mat_eval(m, &h, dir), fp = mat_eval(m, &h, probe);
    out[0] = dir.x; out[1] = dir.y; out[2] = dir.z;
    out[3] = sr.sampled_pdf;
    out[4] = mat_pdf_value(m, &h, dir);
    out[5] = f.x; out[6] = f.y; out[7] = f.z;
    out[8] = mat_pdf_value(m, &h, probe);
    out[9] = fp.x; out[10] = fp.y; out[11] = fp.z;
    return FRT_OK;
}

/* For each ray: hit_record::t of ONE given primitive (creation-order id) tested alone with the
 * closest-hit interval, or NaN.  Lets the parity tests recognise exact geometric ties (coincident
 * surfaces such as the Cornell boxes standing on the floor), where the reference's winner is decided
 * by its traversal order and last-bit rounding rather than by the scene. */
int frt_oracle_prim_t(const frt_scene_desc *scene, const frt_ray *rays, const int32_t *prim_id, size_t n, double *t)
{
    if (!scene || (n && (!rays || !prim_id || !t))) return FRT_ERR_INVALID;
    /* id -> index maps */
    uint32_t max_id = 0;
    for (uint32_t i = 0; i < scene->n_tris; ++i) { const uint32_t id = wbits(scene->tris[3 * (size_t)i]); if (id > max_id) max_id = id; }
    for (uint32_t i = 0; i < scene->n_spheres; ++i) if (scene->spheres[i].prim_id > max_id) max_id = scene->spheres[i].prim_id;
    uint32_t *map = (uint32_t *)malloc(((size_t)max_id + 1) * sizeof(uint32_t));
    if (!map) return FRT_ERR_INVALID;
    for (uint32_t i = 0; i <= max_id; ++i) map[i] = PRIM_NONE;
    for (uint32_t i = 0; i < scene->n_tris; ++i) map[wbits(scene->tris[3 * (size_t)i])] = i;
    for (uint32_t i = 0; i < scene->n_spheres; ++i) map[scene->spheres[i].prim_id] = i | FRT_CHILD_SPHERE;
    for (size_t i = 0; i < n; ++i)
    {
        t[i] = NAN;
        if (prim_id[i] < 0 || (uint32_t)prim_id[i] > max_id || map[prim_id[i]] == PRIM_NONE) continue;
        const uint32_t p = map[prim_id[i]];
        const v3 o = V((real)rays[i].ox, (real)rays[i].oy, (real)rays[i].oz), d = V((real)rays[i].dx, (real)rays[i].dy, (real)rays[i].dz);
        hitrec h;
        const int hit = (p & FRT_CHILD_SPHERE) ? sphere_hit(scene, p & FRT_LEAF_FIRST_MASK, o, d, ray_tmin(o), REAL_FLT_MAX, &h)
                                               : tri_hit(scene, p, o, d, ray_tmin(o), REAL_FLT_MAX, &h);
        if (hit) t[i] = (double)h.t;
    }
    free(map);
    return FRT_OK;
}
