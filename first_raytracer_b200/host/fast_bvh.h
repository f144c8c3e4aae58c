// Fast host builder for large triangle soups: binned-SAH binary build over
// an index array, parallel over subtrees, collapsed straight into the 4-wide node layout of
// frt.h.  The reference's builder (parallel_bvh.h:67-160) sorts every node's range with qsort +
// virtual calls and keeps a primitive list per node; at 10^7 triangles that is minutes and GBs.
// Closest-hit results do not depend on the tree shape (only exact ties do), so parity with the
// reference is unaffected; tests compare both builders on the same rays.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <future>
#include <limits>
#include <vector>
#include "flatten.h"

namespace frt_fast {

struct box3
{
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<float>::infinity(); hi[a] = -lo[a]; } }
    void grow(const box3 &o) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], o.lo[a]); hi[a] = std::max(hi[a], o.hi[a]); } }
    void grow(const float *p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    float half_area() const
    {
        const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return x * y + y * z + z * x;
    }
};

struct bin_node
{
    box3 box;
    int left = -1, right = -1; // children (node indices) or -1
    uint32_t first = 0, count = 0; // range in the index array
};

struct builder
{
    const float *tri9;
    size_t n;
    int max_leaf;
    std::vector<box3> pbox;
    std::vector<float> cent; // 3 per prim
    std::vector<uint32_t> idx;

    // Builds the subtree over idx[first, first+count) into `out` (local indices), returns root.
    int build(std::vector<bin_node> &out, uint32_t first, uint32_t count, int depth)
    {
        const int self = (int)out.size();
        out.emplace_back();
        box3 b, cb;
        b.reset(); cb.reset();
        for (uint32_t i = first; i < first + count; ++i) { b.grow(pbox[idx[i]]); cb.grow(&cent[3 * (size_t)idx[i]]); }
        out[self].box = b; out[self].first = first; out[self].count = count;
        if ((int)count <= max_leaf) return self;

        constexpr int NB = 16;
        int best_axis = -1, best_split = -1;
        float best_cost = std::numeric_limits<float>::infinity();
        for (int axis = 0; axis < 3; ++axis)
        {
            const float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0.f)) continue;
            box3 bb[NB]; uint32_t bc[NB];
            for (int k = 0; k < NB; ++k) { bb[k].reset(); bc[k] = 0; }
            const float scale = NB * (1.f - 1e-6f) / ext;
            for (uint32_t i = first; i < first + count; ++i)
            {
                int k = (int)((cent[3 * (size_t)idx[i] + axis] - cb.lo[axis]) * scale);
                k = k < 0 ? 0 : (k >= NB ? NB - 1 : k);
                bb[k].grow(pbox[idx[i]]); ++bc[k];
            }
            float ra[NB]; box3 acc; acc.reset();
            for (int k = NB - 1; k > 0; --k) { acc.grow(bb[k]); ra[k] = acc.half_area(); }
            acc.reset();
            uint32_t lc = 0;
            for (int k = 0; k < NB - 1; ++k)
            {
                acc.grow(bb[k]); lc += bc[k];
                if (lc == 0 || lc == count) continue;
                const float cost = acc.half_area() * lc + ra[k + 1] * (count - lc);
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = k; }
            }
        }
        uint32_t mid;
        if (best_axis < 0)
            mid = first + count / 2; // all centroids coincide: split the range in half
        else
        {
            const float ext = cb.hi[best_axis] - cb.lo[best_axis];
            const float scale = NB * (1.f - 1e-6f) / ext;
            const float lo = cb.lo[best_axis];
            const int ax = best_axis, sp = best_split;
            auto it = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t p) {
                int k = (int)((cent[3 * (size_t)p + ax] - lo) * scale);
                k = k < 0 ? 0 : (k >= NB ? NB - 1 : k);
                return k <= sp; });
            mid = (uint32_t)(it - idx.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        const uint32_t nl = mid - first, nr = count - nl;
        if (depth < 5 && nl > 50000 && nr > 50000)
        {
            // left subtree on another thread into its own vector, spliced in afterwards
            auto fut = std::async(std::launch::async, [this, first, nl, depth] {
                std::vector<bin_node> sub;
                build(sub, first, nl, depth + 1);
                return sub; });
            const int r = build(out, mid, nr, depth + 1);
            std::vector<bin_node> sub = fut.get();
            const int base = (int)out.size();
            for (bin_node &s : sub) { if (s.left >= 0) s.left += base; if (s.right >= 0) s.right += base; out.push_back(s); }
            out[self].left = base; out[self].right = r;
        }
        else
        {
            const int l = build(out, first, nl, depth + 1);
            const int r = build(out, mid, nr, depth + 1);
            out[self].left = l; out[self].right = r;
        }
        return self;
    }
};

struct collapser
{
    const std::vector<bin_node> &bn;
    const builder &b;
    frt_flat_scene &out;
    float pad;
    uint32_t mesh_tris;

    float down(float x) const { return std::nextafter(x - pad, -std::numeric_limits<float>::infinity()); }
    float up(float x) const { return std::nextafter(x + pad, std::numeric_limits<float>::infinity()); }
    bool leaf(int i) const { return bn[i].left < 0; }

    uint32_t emit_leaf(int i)
    {
        const uint32_t first = (uint32_t)(out.tris.size() / 3);
        for (uint32_t k = bn[i].first; k < bn[i].first + bn[i].count; ++k)
        {
            const uint32_t p = b.idx[k];
            const float *t = b.tri9 + 9 * (size_t)p;
            frt_float4 v0{t[0], t[1], t[2], 0}, e1{t[3], t[4], t[5], 0}, e2{t[6], t[7], t[8], 0}; // v0, v1, v2
            const uint32_t mat = 0, flags = (mesh_tris << 1) | FRT_TRI_GEOMETRY_NORMALS;
            std::memcpy(&v0.w, &p, 4); std::memcpy(&e1.w, &mat, 4); std::memcpy(&e2.w, &flags, 4);
            out.tris.push_back(v0); out.tris.push_back(e1); out.tris.push_back(e2);
        }
        return FRT_CHILD_LEAF | ((bn[i].count - 1) << FRT_LEAF_COUNT_SHIFT) | first;
    }

    uint32_t emit_wide(int i, int depth)
    {
        if (depth > out.max_depth_wide) out.max_depth_wide = depth;
        const uint32_t self = (uint32_t)out.nodes.size();
        out.nodes.emplace_back();
        int kids[4], nk = 0;
        if (leaf(i)) kids[nk++] = i; else { kids[nk++] = bn[i].left; kids[nk++] = bn[i].right; }
        while (nk < 4)
        {
            int pick = -1; float best = -1.f;
            for (int k = 0; k < nk; ++k)
                if (!leaf(kids[k]) && bn[kids[k]].box.half_area() > best) { best = bn[kids[k]].box.half_area(); pick = k; }
            if (pick < 0) break;
            const int open = kids[pick];
            for (int k = nk; k > pick + 1; --k) kids[k] = kids[k - 1];
            kids[pick] = bn[open].left; kids[pick + 1] = bn[open].right;
            ++nk;
        }
        frt_bvh4_node node;
        const float inf = std::numeric_limits<float>::infinity();
        for (int k = 0; k < 4; ++k)
        {
            node.lo_x[k] = node.lo_y[k] = node.lo_z[k] = inf;
            node.hi_x[k] = node.hi_y[k] = node.hi_z[k] = -inf;
            node.child[k] = FRT_CHILD_EMPTY; node.reserved[k] = 0;
        }
        for (int k = 0; k < nk; ++k)
        {
            const box3 &bx = bn[kids[k]].box;
            node.lo_x[k] = down(bx.lo[0]); node.hi_x[k] = up(bx.hi[0]);
            node.lo_y[k] = down(bx.lo[1]); node.hi_y[k] = up(bx.hi[1]);
            node.lo_z[k] = down(bx.lo[2]); node.hi_z[k] = up(bx.hi[2]);
            node.child[k] = leaf(kids[k]) ? emit_leaf(kids[k]) : emit_wide(kids[k], depth + 1);
        }
        out.nodes[self] = node;
        return self;
    }
};

// tri9: n*9 floats; primitive id = index in tri9.  One lambertian material, no lights.
inline frt_flat_scene build_soup(const float *tri9, size_t n, int max_leaf)
{
    builder b;
    b.tri9 = tri9; b.n = n; b.max_leaf = max_leaf < 1 ? 1 : (max_leaf > 16 ? 16 : max_leaf);
    b.pbox.resize(n); b.cent.resize(3 * n); b.idx.resize(n);
    float m = 0.f;
    for (size_t i = 0; i < n; ++i)
    {
        box3 bx; bx.reset();
        for (int v = 0; v < 3; ++v) bx.grow(tri9 + 9 * i + 3 * v);
        b.pbox[i] = bx;
        for (int a = 0; a < 3; ++a)
        {
            b.cent[3 * i + a] = 0.5f * (bx.lo[a] + bx.hi[a]);
            m = std::max(m, std::max(std::fabs(bx.lo[a]), std::fabs(bx.hi[a])));
        }
        b.idx[i] = (uint32_t)i;
    }
    std::vector<bin_node> bn;
    bn.reserve(2 * n / (size_t)b.max_leaf + 16);
    frt_flat_scene out;
    if (n)
    {
        b.build(bn, 0, (uint32_t)n, 0);
        out.nodes.reserve(bn.size() / 2 + 16);
        out.tris.reserve(3 * n);
        collapser c{bn, b, out, 4e-7f * std::max(m, 4.0f), (uint32_t)std::min<size_t>(n, 0x7FFFFFFFu)};
        c.emit_wide(0, 1);
    }
    frt_material mat;
    std::memset(&mat, 0, sizeof mat);
    mat.type = FRT_MAT_LAMBERTIAN; mat.p[0] = mat.p[1] = mat.p[2] = 0.5f;
    out.materials.push_back(mat);
    out.env[0] = out.env[1] = out.env[2] = 1.f;
    return out;
}

} // namespace frt_fast
