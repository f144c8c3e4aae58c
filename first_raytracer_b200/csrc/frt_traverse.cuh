// Closest-hit / any-hit traversal of the 4-wide flattened BVH.
//
// Semantics are the reference's live (non-SSE) code:
//   aabb::hit            aabb.h:14-31       slab test, reject iff tmax < tmin, NaNs ignored
//   triangle::hit        triangle.h:69-118  Möller–Trumbore, a == 0 rejects, strict t bounds
//   sphere::hit          sphere.h:26-56     near root, far root if near < t_min, non-strict bounds
//   parallel_bvh_node::hit parallel_bvh.h:39-64  origin-scaled t_min; closest hit wins; of two hits
//                        at exactly the same t the one met first (left subtree first) wins.
// Precision split: boxes and a conservative triangle FILTER run in float32; a triangle that passes
// the filter is DECIDED by the reference's own formula in float64 on the exact float32 vertices
// (edge = v1 - v0 in float64, as triangle.h:60-61), so hit / miss, t, u, v and the winner among
// near-coincident surfaces are the reference's, not a float32 approximation of them.  Triangles
// are stored in the reference's depth-first leaf order, so "met first" is "lower index": the tie
// rule holds whatever order this traversal visits the nodes in.
// One thread owns one ray; the short stack lives in local memory (L1-resident).
#pragma once
#include "frt_device.cuh"

#define FRT_STACK_SIZE 64 // 4-wide: 3 pushes per level, trees up to 21 levels deep
#define FRT_STACK8_SIZE 96 // 8-wide: 7 pushes per level, trees up to 13 levels deep
// Closest hit whose t, u, v are still the float32 estimate carries this bit in RayHit::prim; the
// consumer (shade, with the whole warp converged) calls frt_resolve_hit before using them.
#define FRT_HIT_INEXACT 0x20000000u

struct RayHit
{
    float t, u, v;
    uint32_t prim; // index into tris[] (leaf order), | FRT_CHILD_SPHERE for spheres[]; 0xFFFFFFFF = miss
};

__device__ __forceinline__ uint32_t frt_pick(const uint4 &c, uint32_t i)
{
    return i == 0 ? c.x : (i == 1 ? c.y : (i == 2 ? c.z : c.w));
}

// triangle::hit (triangle.h:69-118) in float64 on the exact float32 vertices of triangle `prim`;
// t < 0 means "no hit" (hits have t > t_min > 0); the caller applies the upper bound.  Kept out of
// line: it runs off the fast path (ambiguous candidates) and once per ray for the winner.
struct ExactHit { double t; float u, v; };
// parallel_bvh.h:46-51 (the caller passed EPSILON): t_min = max(1e-4, 1e-4 * max(|o.x|, |o.y|, |o.z|))
__device__ __forceinline__ double frt_tmin_exact(const float3 o)
{
    const double m = fmax(fmax(fabs((double)o.x), fabs((double)o.y)), fmax(fabs((double)o.z), 1e-4));
    return fmax(1e-4, 1e-4 * m);
}
__device__ __noinline__ ExactHit frt_tri_exact(const float4 *tris, uint32_t prim, float ox, float oy, float oz, float dxf,
                                               float dyf, float dzf, double tmin)
{
    const float4 *tp = tris + 3 * (size_t)prim;
    const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
    ExactHit r;
    r.t = -1.0; r.u = 0.f; r.v = 0.f;
    const double v0x = q0.x, v0y = q0.y, v0z = q0.z;
    const double e1x = (double)q1.x - v0x, e1y = (double)q1.y - v0y, e1z = (double)q1.z - v0z;
    const double e2x = (double)q2.x - v0x, e2y = (double)q2.y - v0y, e2z = (double)q2.z - v0z;
    const double dx = dxf, dy = dyf, dz = dzf;
    const double hx = dy * e2z - dz * e2y, hy = dz * e2x - dx * e2z, hz = dx * e2y - dy * e2x;
    const double a = e1x * hx + e1y * hy + e1z * hz;
    if (a == 0.0) return r;
    const double f = 1.0 / a;
    const double sx = (double)ox - v0x, sy = (double)oy - v0y, sz = (double)oz - v0z;
    const double u = f * (sx * hx + sy * hy + sz * hz);
    if (u < 0.0 || u > 1.0) return r;
    const double qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
    const double v = f * (dx * qx + dy * qy + dz * qz);
    if (!(v >= 0.0 && u + v <= 1.0)) return r;
    const double t = f * (e2x * qx + e2y * qy + e2z * qz);
    if (t > tmin) { r.t = t; r.u = (float)u; r.v = (float)v; }
    return r;
}

#define FRT_FILTER_K 1e-5f // ~170 float32 ulps on each sum of absolute products
// sum of |a_i * b_i| as one multiply and two fused multiply-adds (the absolute values are operand modifiers)
__device__ __forceinline__ float frt_absdot(const float3 a, const float3 b)
{
    return fmaf(fabsf(a.z), fabsf(b.z), fmaf(fabsf(a.y), fabsf(b.y), fabsf(a.x) * fabsf(b.x)));
}

// Closest hit so far.  A hit that the float32 test places clearly inside the triangle and clearly
// in front of everything found so far is taken on its float32 interval [lo, hi] ("lazy"); float64
// is evaluated only when two candidates' intervals overlap or a candidate sits on an edge / on the
// t bounds, and once at the end for the winner (converged across the warp), which gives t, u, v.
struct BestHit
{
    uint32_t prim;
    float lo, hi; // bounds on t (the rounded float64 value once exact)
    float u, v;
    double t;     // valid iff exact
    bool exact;
};

// One ray's traversal state.  begin() sets the ray up, round() runs ONE while-while round (descend
// to a leaf, test its primitives, pop the next postponed child), finish() gives the hit.  Splitting
// the loop this way lets a persistent warp look at its lanes between rounds and hand finished
// lanes a new ray (dynamic fetch) instead of idling until the slowest lane is done.
// The traversal's short stack of postponed children: x = child reference, y = bits of its entry distance (to drop it
// at pop time once something closer was found).  LocalStack is a per-thread array in local memory: L1-resident for
// the scenes that fit the caches (the wavefront kernels).  HybridStack keeps the first K entries of every thread in
// SHARED memory (entry e of thread t at [e * threads + t]: conflict-free whatever the lanes' depths) and only deeper
// ones in local memory: on a scene that streams gigabytes through L1 the local-memory stack lines are evicted
// between push and pop (ncu, 10^7-triangle soup: 0.77 KB of DRAM WRITES per ray, and pops that wait for HBM).
struct LocalStack
{
    uint2 *p;
    __device__ __forceinline__ void push(int &sp, const uint2 v) { p[sp++] = v; }
    __device__ __forceinline__ uint2 pop(int &sp) { return p[--sp]; }
};
template <int K, int THREADS>
struct HybridStack
{
    uint2 *sm; // this thread's column of the CTA's shared stack area
    uint2 *lm; // overflow (FRT_STACK_SIZE - K entries of local memory)
    __device__ __forceinline__ void push(int &sp, const uint2 v)
    {
        if (sp < K) sm[sp * THREADS] = v; else lm[sp - K] = v;
        ++sp;
    }
    __device__ __forceinline__ uint2 pop(int &sp)
    {
        --sp;
        return sp < K ? sm[sp * THREADS] : lm[sp - K];
    }
};

// RT = instantiated by the ray-cast entry (k_trace), where occlusion queries walk the children near-to-far exactly like
// closest-hit queries (on the 10^7-triangle soup: 379 against 359 Mrays/s; the nearest entered leaf is the likeliest
// to hold an occluder); the wavefront kernels (RT = false) keep the cheaper unordered occlusion walk, measured faster
// on the scenes that fit the caches.
template <bool ANY, bool COUNT, int WIDE = 4, bool RT = false>
struct Walker
{
    float3 o, d;
    float idx, idy, idz, oox, ooy, ooz;
    int sx, sy, sz;
    float tmin, tmin_hi, tmax;
    BestHit best;
    uint32_t cur;
    int sp;
    bool active;
    // The short stack is NOT a member: a dynamically indexed array inside the struct would put every
    // member in local memory.  The caller declares `uint2 stk[FRT_STACK_SIZE]` (x = child reference,
    // y = bits of its entry distance, to drop it at pop time once something closer was found) and
    // hands it to round().  frt_upload_scene rejects trees deeper than the stack can serve.

    // upper end of the t interval: FLT_MAX for closest hit (path.cpp:10), DevScene::any_tend (uniform) for occlusion
    static __device__ __forceinline__ double tend_d(const DevScene &sc) { return ANY ? sc.any_tend : (double)FRT_FLT_MAX; }
    __device__ __forceinline__ double tmin_exact() const { return frt_tmin_exact(o); }

    __device__ __forceinline__ void begin(const DevScene &sc, const float3 o_, const float3 d_)
    {
        o = o_; d = d_;
        // parallel_bvh.h:46-51: t_min = max(1e-4, 1e-4 * max|o_i|) in double.  The traversal keeps float32
        // brackets of it (1e-4f < 1e-4 < next float; products rounded outwards); the double itself is
        // formed only where a decision is made in float64 (tmin_exact()).
        const float m = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
        tmin = fmaxf(9.99999974737875e-05f, __fmul_rd(9.99999974737875e-05f, m));
        tmin_hi = fmaxf(1.00000004749745e-04f, __fmul_ru(1.00000004749745e-04f, m));
        best.prim = 0xFFFFFFFFu; best.u = best.v = 0.f; best.t = tend_d(sc); best.exact = true;
        best.lo = ANY ? sc.any_tend_lo : FRT_FLT_MAX; best.hi = ANY ? sc.any_tend_hi : FRT_FLT_MAX;
        tmax = best.hi; // culling bound
        // correctly rounded reciprocals (the library is built with -prec-div=false): the padding of the float32
        // boxes (host/flatten.h) is sized for a slab test whose only error is rounding
        // A zero (or vanishing) direction component would give an infinite reciprocal and NaN slab products, which
        // fmaxf / fminf drop: that axis would stop culling and the ray would visit every node its other two axes touch
        // (measured on the 10^7-triangle soup: ONE such ray among 2^24 — torch.randn returns exact zeros — walked 9387
        // nodes and 29286 triangles and held the whole launch for 15.8 ms).  +-2^-60 keeps the products finite and, as
        // a power of two, exact, so the sign of (plane - origin) decides, as (min - origin) / 0 = +-inf does in
        // aabb.h:18-19.  Only the slab test sees the substitute; hits are decided with the true direction.
        const float tiny = 8.67361737988403547e-19f; // 2^-60
        idx = __frcp_rn(fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x);
        idy = __frcp_rn(fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y);
        idz = __frcp_rn(fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
        oox = o.x * idx; ooy = o.y * idy; ooz = o.z * idz;
        sx = idx < 0.0f ? 1 : 0; sy = idy < 0.0f ? 1 : 0; sz = idz < 0.0f ? 1 : 0;
        sp = 0;
        cur = 0;
        // A NaN in the ray misses everything in the reference too (every comparison of aabb.h:24-28 and
        // triangle.h:80-89 is false: no box is rejected, no triangle accepted) — after visiting the whole tree.
        // The same answer without the walk.
        const float s = o.x + o.y + o.z + d.x + d.y + d.z;
        active = sc.n_nodes != 0 && s == s;
    }

    // ---- while-while, part 1 over the 4-wide nodes: descend through inner nodes (EMPTY carries the leaf bit and
    // ends the loop)
    template <class Stk>
    __device__ __forceinline__ void descend4(const DevScene &sc, Stk &stk, unsigned int &n_nodes)
    {
        while (!(cur & FRT_CHILD_LEAF))
        {
            // eight 128-bit loads, near/far planes picked by ray-direction sign
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
            if (ANY && !RT)
            {
                // an occlusion query looks at every child the segment enters unless an occluder turns up
                // first, so the order is free: no sort, no entry distances on the stack
                uint32_t next = FRT_CHILD_EMPTY;
                if (tf3 >= tn3) next = ch.w;
                if (tf2 >= tn2) { if (next != FRT_CHILD_EMPTY) stk.push(sp, make_uint2(next, 0u)); next = ch.z; }
                if (tf1 >= tn1) { if (next != FRT_CHILD_EMPTY) stk.push(sp, make_uint2(next, 0u)); next = ch.y; }
                if (tf0 >= tn0) { if (next != FRT_CHILD_EMPTY) stk.push(sp, make_uint2(next, 0u)); next = ch.x; }
                cur = next;
                continue;
            }
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
            // invalid keys sort last: push the far ones (farthest first), continue with the nearest
            if (k3 != 0xFFFFFFFFu) stk.push(sp, make_uint2(frt_pick(ch, k3 & 3u), k3 & ~3u));
            if (k2 != 0xFFFFFFFFu) stk.push(sp, make_uint2(frt_pick(ch, k2 & 3u), k2 & ~3u));
            if (k1 != 0xFFFFFFFFu) stk.push(sp, make_uint2(frt_pick(ch, k1 & 3u), k1 & ~3u));
            cur = (k0 != 0xFFFFFFFFu) ? frt_pick(ch, k0 & 3u) : FRT_CHILD_EMPTY;
        }
    }

    // ---- the same over the 8-wide compressed nodes of frt_wide8.cuh (96 B: box origin + per-axis power-of-two
    // scale, 8 x 6 byte-quantised planes, 8 child references).  Plane k of an axis is origin + q * scale, so its
    // slab distance is q * (scale / d) + (origin / d - o / d): one FFMA per plane after three per node.  The
    // quantised boxes contain the float32 boxes of the 4-wide tree (floor / ceil), hits are decided by the same
    // triangle code, so results do not depend on which tree is walked.
    template <class Stk>
    __device__ __forceinline__ void descend8(const DevScene &sc, Stk &stk, unsigned int &n_nodes)
    {
        while (!(cur & FRT_CHILD_LEAF))
        {
            const uint4 *n = sc.nodes8 + sc.node8_stride * (size_t)cur;
            if (COUNT) ++n_nodes;
            const uint4 hd = __ldg(n), qa = __ldg(n + 1), qb = __ldg(n + 2), qc = __ldg(n + 3);
            const uint4 c0 = __ldg(n + 4), c1 = __ldg(n + 5);
            const float sx_ = __uint_as_float((hd.w & 0xFFu) << 23) * idx, sy_ = __uint_as_float(((hd.w >> 8) & 0xFFu) << 23) * idy,
                        sz_ = __uint_as_float(((hd.w >> 16) & 0xFFu) << 23) * idz;
            const float ax = fmaf(__uint_as_float(hd.x), idx, -oox), ay = fmaf(__uint_as_float(hd.y), idy, -ooy),
                        az = fmaf(__uint_as_float(hd.z), idz, -ooz);
            // low / high plane bytes: x lo = qa.xy, y lo = qa.zw, z lo = qb.xy, x hi = qb.zw, y hi = qc.xy, z hi = qc.zw
            const uint2 nxw = sx ? make_uint2(qb.z, qb.w) : make_uint2(qa.x, qa.y), fxw = sx ? make_uint2(qa.x, qa.y) : make_uint2(qb.z, qb.w);
            const uint2 nyw = sy ? make_uint2(qc.x, qc.y) : make_uint2(qa.z, qa.w), fyw = sy ? make_uint2(qa.z, qa.w) : make_uint2(qc.x, qc.y);
            const uint2 nzw = sz ? make_uint2(qc.z, qc.w) : make_uint2(qb.x, qb.y), fzw = sz ? make_uint2(qb.x, qb.y) : make_uint2(qc.z, qc.w);
            uint32_t key[8];
#pragma unroll
            for (int k = 0; k < 8; ++k)
            {
                const int sh = 8 * (k & 3);
                const float qnx = (float)(((k < 4 ? nxw.x : nxw.y) >> sh) & 0xFFu), qfx = (float)(((k < 4 ? fxw.x : fxw.y) >> sh) & 0xFFu);
                const float qny = (float)(((k < 4 ? nyw.x : nyw.y) >> sh) & 0xFFu), qfy = (float)(((k < 4 ? fyw.x : fyw.y) >> sh) & 0xFFu);
                const float qnz = (float)(((k < 4 ? nzw.x : nzw.y) >> sh) & 0xFFu), qfz = (float)(((k < 4 ? fzw.x : fzw.y) >> sh) & 0xFFu);
                const float tn = fmaxf(fmaxf(fmaf(qnx, sx_, ax), fmaf(qny, sy_, ay)), fmaxf(fmaf(qnz, sz_, az), tmin));
                const float tf = fminf(fminf(fmaf(qfx, sx_, ax), fmaf(qfy, sy_, ay)), fminf(fmaf(qfz, sz_, az), tmax));
                const uint32_t ref = k < 4 ? frt_pick(c0, k) : frt_pick(c1, k - 4);
                // an empty slot stores inverted planes (255 / 0) AND the EMPTY reference: never entered
                key[k] = (tf >= tn && ref != FRT_CHILD_EMPTY) ? ((__float_as_uint(tn) & ~7u) | (uint32_t)k) : 0xFFFFFFFFu;
            }
            if (ANY && !RT)
            {
                uint32_t next = FRT_CHILD_EMPTY;
#pragma unroll
                for (int k = 7; k >= 0; --k)
                    if (key[k] != 0xFFFFFFFFu)
                    {
                        if (next != FRT_CHILD_EMPTY) stk.push(sp, make_uint2(next, 0u));
                        next = k < 4 ? frt_pick(c0, k) : frt_pick(c1, k - 4);
                    }
                cur = next;
                continue;
            }
            // sort the entered children near-to-far: 19-comparator network on tnear | slot keys
#define FRT_CSWAP(i, j) { const uint32_t a_ = min(key[i], key[j]), b_ = max(key[i], key[j]); key[i] = a_; key[j] = b_; }
            FRT_CSWAP(0, 1) FRT_CSWAP(2, 3) FRT_CSWAP(4, 5) FRT_CSWAP(6, 7)
            FRT_CSWAP(0, 2) FRT_CSWAP(1, 3) FRT_CSWAP(4, 6) FRT_CSWAP(5, 7)
            FRT_CSWAP(1, 2) FRT_CSWAP(5, 6) FRT_CSWAP(0, 4) FRT_CSWAP(3, 7)
            FRT_CSWAP(1, 5) FRT_CSWAP(2, 6)
            FRT_CSWAP(1, 4) FRT_CSWAP(3, 6)
            FRT_CSWAP(2, 4) FRT_CSWAP(3, 5)
            FRT_CSWAP(3, 4)
#undef FRT_CSWAP
            // invalid keys sort last: push the far ones (farthest first), continue with the nearest
#pragma unroll
            for (int j = 7; j >= 1; --j)
                if (key[j] != 0xFFFFFFFFu)
                {
                    const uint32_t slot = key[j] & 7u;
                    const uint32_t ref = slot < 4u ? frt_pick(c0, slot) : frt_pick(c1, slot - 4u);
                    stk.push(sp, make_uint2(ref, key[j] & ~7u));
                }
            if (key[0] != 0xFFFFFFFFu)
            {
                const uint32_t slot = key[0] & 7u;
                cur = slot < 4u ? frt_pick(c0, slot) : frt_pick(c1, slot - 4u);
            }
            else
                cur = FRT_CHILD_EMPTY;
        }
    }

    template <class Stk>
    __device__ __forceinline__ void round(const DevScene &sc, Stk &stk, unsigned int &n_nodes, unsigned int &n_tris)
    {
        if (WIDE == 8) descend8(sc, stk, n_nodes); else descend4(sc, stk, n_nodes);

        // ---- part 2: one leaf
        if (cur != FRT_CHILD_EMPTY)
        {
            const uint32_t first = cur & FRT_LEAF_FIRST_MASK;
            const uint32_t count = ((cur >> FRT_LEAF_COUNT_SHIFT) & 15u) + 1u;
            const bool is_sphere = cur & FRT_CHILD_SPHERE;
            for (uint32_t k = 0; k < count; ++k)
            {
                const uint32_t prim = is_sphere ? ((first + k) | FRT_CHILD_SPHERE) : first + k;
                ExactHit e;
                if (is_sphere)
                {
                    // sphere.h:26-56 entirely in float64 (few spheres; the float32 discriminant cancels badly)
                    const float4 s = __ldg(sc.spheres + 2 * (first + k));
                    const double ocx = (double)o.x - s.x, ocy = (double)o.y - s.y, ocz = (double)o.z - s.z;
                    const double dx = d.x, dy = d.y, dz = d.z, r = s.w;
                    const double a = dx * dx + dy * dy + dz * dz, b = ocx * dx + ocy * dy + ocz * dz;
                    const double c = ocx * ocx + ocy * ocy + ocz * ocz - r * r;
                    double disc = b * b - a * c;
                    if (!(disc >= 0.0)) continue;
                    disc = sqrt(disc);
                    double t = (-b - disc) / a;
                    const double tmin_d = tmin_exact();
                    if (t < tmin_d) t = (-b + disc) / a;
                    if (t < tmin_d || t > tend_d(sc)) continue;
                    e.t = t; e.u = 0.f; e.v = 0.f;
                }
                else
                {
                    const float4 *tp = sc.tris + 3 * (size_t)prim;
                    const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
                    if (COUNT) ++n_tris;
                    // float32 pass: the same formula with error margins from the sums of absolute products,
                    // evaluated without branches (a warp almost always has a lane that needs every term) and
                    // classified once: certainly missed / certainly the closest hit so far / decide in float64
                    const float3 v0 = f3(q0), e1 = f3(q1) - v0, e2 = f3(q2) - v0;
                    const float3 h = cross(d, e2);
                    const float a = dot(e1, h);
                    const float abs_a = frt_absdot(e1, h);
                    const float f = __fdividef(1.0f, a), af = fabsf(f); // 2 ulp: far inside the margins
                    const float kf = FRT_FILTER_K * af;
                    const float rel = kf * abs_a + 1e-6f;
                    const float3 s = o - v0;
                    const float uf = f * dot(s, h);
                    const float du = fmaf(kf, frt_absdot(s, h), fmaf(fabsf(uf), rel, 1e-6f));
                    const float3 q = cross(s, e1);
                    const float vf = f * dot(d, q);
                    const float dv = fmaf(kf, frt_absdot(d, q), fmaf(fabsf(vf), rel, 1e-6f));
                    const float tf = f * dot(e2, q);
                    const float dt = fmaf(kf, frt_absdot(e2, q), fabsf(tf) * rel);
                    const float wf = uf + vf, dw = du + dv;
                    // too ill-conditioned for float32 (or a == 0): neither verdict, decide exactly
                    const bool conditioned = fabsf(a) > 16.0f * FRT_FILTER_K * abs_a;
                    const bool missed = uf < -du || uf > 1.0f + du || vf < -dv || wf > 1.0f + dw // outside the triangle
                                        || tf < tmin - dt || tf > tmax + dt;                  // outside (t_min, closest so far)
                    if (conditioned && missed) continue;
                    if (conditioned && uf > du && vf > dv && wf < 1.0f - dw && tf - dt > tmin_hi && tf + dt < best.lo)
                    {
                        // certainly a hit and certainly the closest so far (an occlusion query keeps only the
                        // occluder's index: its Walker carries no hit interval, t, u, v between rounds)
                        best.prim = prim;
                        if (ANY) { active = false; return; }
                        best.lo = tf - dt; best.hi = tf + dt; best.u = uf; best.v = vf; best.exact = false;
                        tmax = best.hi;
                        continue;
                    }
                    e = frt_tri_exact(sc.tris, prim, o.x, o.y, o.z, d.x, d.y, d.z, tmin_exact());
                    if (e.t < 0.0) continue;
                }
                // float64 decision: strict t < t_max for triangles (triangle.h:88-89), t <= t_max for spheres
                // (sphere.h:41); at an exact tie the triangle the reference meets first (lower index) stays
                if (ANY)
                {
                    if (is_sphere || e.t < tend_d(sc)) { best.prim = prim; active = false; return; }
                    continue;
                }
                if (!best.exact) // the closest so far was taken on its float32 interval: pin it down now
                {
                    const ExactHit b = frt_tri_exact(sc.tris, best.prim, o.x, o.y, o.z, d.x, d.y, d.z, tmin_exact());
                    if (b.t >= 0.0) { best.t = b.t; best.u = b.u; best.v = b.v; }
                    else best.t = 0.5 * ((double)best.lo + (double)best.hi); // not reachable while the margins hold
                    best.exact = true;
                }
                // Exact ties as the reference's visiting order decides them: a sphere accepts t <= t_max (sphere.h:41), a
                // triangle only t < t_max (triangle.h:88-89), so at equal t a sphere beats a triangle whichever is met
                // first, the later-met (higher leaf index) of two spheres stays, and the earlier-met of two triangles.
                const bool best_none = best.prim == 0xFFFFFFFFu, best_sph = !best_none && (best.prim & FRT_CHILD_SPHERE);
                const bool closer = e.t < best.t ||
                                    (e.t == best.t && (is_sphere ? (best_none || !best_sph || prim > best.prim)
                                                                 : (!best_none && !best_sph && prim < best.prim)));
                if (closer)
                {
                    best.prim = prim; best.t = e.t; best.exact = true; best.u = e.u; best.v = e.v;
                    best.lo = __double2float_rd(e.t); best.hi = __double2float_ru(e.t);
                    tmax = best.hi;
                }
            }
        }

        // ---- pop the nearest postponed child that can still hold something closer
        active = false;
        while (sp > 0)
        {
            const uint2 e = stk.pop(sp);
            if (__uint_as_float(e.y) <= tmax) { cur = e.x; active = true; break; }
        }
    }

    // t, u, v of the winner: one float64 evaluation per ray, outside the loop (closest hit only;
    // an any-hit query reports just the occluder's index)
    __device__ __forceinline__ RayHit finish(const DevScene &sc)
    {
        RayHit out;
        out.t = 0.f; out.u = 0.f; out.v = 0.f; out.prim = best.prim;
        if (best.prim == 0xFFFFFFFFu) return out;
        if (ANY) return out; // only "occluded by prim" is meaningful (path.cpp:50)
        if (!best.exact)
        {
            const ExactHit b = frt_tri_exact(sc.tris, best.prim, o.x, o.y, o.z, d.x, d.y, d.z, tmin_exact());
            if (b.t >= 0.0) { best.t = b.t; best.u = b.u; best.v = b.v; }
            else best.t = 0.5 * ((double)best.lo + (double)best.hi);
        }
        out.t = (float)best.t; out.u = best.u; out.v = best.v;
        return out;
    }

    // The same without the float64 evaluation: an inexact winner is flagged for frt_resolve_hit.
    __device__ __forceinline__ RayHit finish_lazy() const
    {
        RayHit out;
        out.prim = best.prim;
        if (best.prim == 0xFFFFFFFFu || best.exact) { out.t = (float)best.t; out.u = best.u; out.v = best.v; return out; }
        out.t = 0.5f * (best.lo + best.hi); out.u = best.u; out.v = best.v;
        // The interval is the conservative margin (1e-5-scale), at least ten times the real float32 error: when
        // it is below 1e-4 t the float32 t, u, v are within ~1e-6 of the float64 ones, which is all shading
        // needs (it continues in float32 anyway).  Only wider intervals are resolved in float64.
        if (best.hi - best.lo > 1e-4f * best.lo) out.prim = best.prim | FRT_HIT_INEXACT;
        return out;
    }
};

__device__ __forceinline__ void frt_resolve_hit(const DevScene &sc, const float3 o, const float3 d, float &t, float &u, float &v,
                                                uint32_t &prim)
{
    prim &= ~FRT_HIT_INEXACT;
    const ExactHit b = frt_tri_exact(sc.tris, prim, o.x, o.y, o.z, d.x, d.y, d.z, frt_tmin_exact(o));
    if (b.t >= 0.0) { t = (float)b.t; u = b.u; v = b.v; }
}
