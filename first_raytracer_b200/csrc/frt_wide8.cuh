// 8-wide compressed nodes for scenes that do not fit the caches (the 10^7-triangle soup: 0.9 GB).
//
// frt_upload_scene derives them from the 4-wide nodes of the scene description (include/frt.h keeps one node format
// at the boundary): a node takes its children and keeps opening the inner child of largest area while the result
// fits 8 slots, which halves the number of DEPENDENT node fetches per ray (what bounds an incoherent walk of a tree
// in HBM: each fetch waits for the one before), and the child boxes are stored as bytes relative to the node's own
// box, so a node is 96 B for 8 children instead of 128 B for 4:
//
//   uint4 0 : origin.x, origin.y, origin.z (float bits), w = ex | ey << 8 | ez << 16 (biased float exponents: the
//             plane spacing of an axis is 2^(e - 127), exact to decode)
//   uint4 1 : lo_x[0..7] bytes (x, y), lo_y[0..7] (z, w)
//   uint4 2 : lo_z[0..7] (x, y), hi_x[0..7] (z, w)
//   uint4 3 : hi_y[0..7] (x, y), hi_z[0..7] (z, w)
//   uint4 4, 5 : child[0..7] — inner: index of an 8-wide node; leaf / empty: the 4-wide tree's reference as is
//
// Child box = origin + byte * spacing, low planes rounded down and high planes rounded up after a margin for the
// float32 decode on the device, so it contains the (already conservative) float32 box of the 4-wide tree.  Hits are
// decided by the triangle / sphere code both walks share: a looser box only costs visits.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../include/frt.h"

namespace frt_wide8 {

struct Kid
{
    uint32_t ref;
    float lo[3], hi[3];
};

inline int kids_of(const frt_bvh4_node &n, Kid *out)
{
    int k = 0;
    for (int i = 0; i < 4; ++i)
    {
        if (n.child[i] == FRT_CHILD_EMPTY) continue;
        Kid &c = out[k++];
        c.ref = n.child[i];
        c.lo[0] = n.lo_x[i]; c.lo[1] = n.lo_y[i]; c.lo[2] = n.lo_z[i];
        c.hi[0] = n.hi_x[i]; c.hi[1] = n.hi_y[i]; c.hi[2] = n.hi_z[i];
    }
    return k;
}

inline double area(const Kid &c)
{
    const double x = (double)c.hi[0] - c.lo[0], y = (double)c.hi[1] - c.lo[1], z = (double)c.hi[2] - c.lo[2];
    return x * y + y * z + z * x;
}

struct Builder
{
    const frt_bvh4_node *n4;
    uint32_t n_nodes4;
    float margin; // added around every child box before quantising
    std::vector<uint32_t> out; // 24 words per node
    int max_depth = 0;

    uint32_t emit(uint32_t node4, int depth)
    {
        if (depth > max_depth) max_depth = depth;
        const uint32_t self = (uint32_t)(out.size() / 24);
        out.resize(out.size() + 24);
        Kid kids[8];
        int nk = kids_of(n4[node4], kids);
        for (;;)
        {
            int pick = -1;
            double best = -1.0;
            Kid sub[4];
            for (int i = 0; i < nk; ++i)
            {
                if (kids[i].ref & FRT_CHILD_LEAF) continue;
                const int m = kids_of(n4[kids[i].ref], sub);
                if (m == 0 || nk - 1 + m > 8) continue;
                const double a = area(kids[i]);
                if (a > best) { best = a; pick = i; }
            }
            if (pick < 0) break;
            const int m = kids_of(n4[kids[pick].ref], sub);
            // open in place: the children keep the depth-first order of the 4-wide tree
            for (int i = nk - 1; i > pick; --i) kids[i + m - 1] = kids[i];
            for (int i = 0; i < m; ++i) kids[pick + i] = sub[i];
            nk += m - 1;
        }
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int i = 0; i < nk; ++i)
            for (int a = 0; a < 3; ++a)
            {
                lo[a] = std::fmin(lo[a], kids[i].lo[a] - margin);
                hi[a] = std::fmax(hi[a], kids[i].hi[a] + margin);
            }
        uint32_t w[24];
        std::memset(w, 0, sizeof w);
        uint32_t exps = 0;
        float spacing[3];
        for (int a = 0; a < 3; ++a)
        {
            // smallest power of two with origin + 255 * spacing >= hi (one more step when rounding bites)
            const double ext = nk ? std::fmax((double)hi[a] - (double)lo[a], 1e-30) : 1.0;
            int e = (int)std::ceil(std::log2(ext / 255.0));
            while (std::ldexp(255.0, e) < ext) ++e;
            e = std::min(127, std::max(-126, e));
            spacing[a] = std::ldexp(1.0f, e);
            exps |= (uint32_t)(e + 127) << (8 * a);
            if (!nk) lo[a] = 0.f;
            std::memcpy(&w[a], &lo[a], 4);
        }
        w[3] = exps;
        unsigned char *q = reinterpret_cast<unsigned char *>(&w[4]); // 48 plane bytes: lo_x, lo_y, lo_z, hi_x, hi_y, hi_z
        for (int k = 0; k < 8; ++k)
        {
            for (int a = 0; a < 3; ++a) { q[8 * a + k] = 255; q[8 * (3 + a) + k] = 0; } // empty: inverted planes
            w[16 + k] = FRT_CHILD_EMPTY;
        }
        for (int k = 0; k < nk; ++k)
            for (int a = 0; a < 3; ++a)
            {
                const double l = std::floor(((double)kids[k].lo[a] - margin - (double)lo[a]) / spacing[a]);
                const double h = std::ceil(((double)kids[k].hi[a] + margin - (double)lo[a]) / spacing[a]);
                q[8 * a + k] = (unsigned char)std::min(255.0, std::max(0.0, l));
                q[8 * (3 + a) + k] = (unsigned char)std::min(255.0, std::max(0.0, h));
            }
        for (int k = 0; k < nk; ++k)
            w[16 + k] = (kids[k].ref & FRT_CHILD_LEAF) ? kids[k].ref : emit(kids[k].ref, depth + 1);
        std::memcpy(&out[(size_t)self * 24], w, sizeof w);
        return self;
    }
};

// `margin` covers the float32 decode of a plane distance on the device (three roundings on values of the size of the
// scene's coordinates): a few ulps of the largest coordinate.
inline void build(const frt_bvh4_node *nodes4, uint32_t n_nodes4, float margin, std::vector<uint32_t> &words, int *depth)
{
    Builder b{nodes4, n_nodes4, margin, {}, 0};
    if (n_nodes4) b.emit(0, 1);
    words.swap(b.out);
    if (depth) *depth = b.max_depth;
}

} // namespace frt_wide8
