// Host SAH BVH build with the reference's class name, members and split rule
// (parallel_bvh.h:8-27,67-175; comparators bvh.h:6-55), so that the tree the GPU flattener
// receives has the shape the reference would traverse:
//   * bounds = union of primitive boxes; split axis = aabb::longest_axis() (aabb.h:33-43)
//   * primitives ordered by box.min[axis]; full sweep; cost_i = i*A_L[i] + (n-i-1)*A_R[i+1]
//     for i in [0, n-2], first minimum wins (parallel_bvh.h:117-127)
//   * left = l[0..idx], right = l[idx+1..n), one primitive per leaf
// Differences that do not change the tree: boxes are computed once per primitive instead of inside
// the comparator (the sort itself is the same qsort call), subtrees are built on std::async tasks instead of taskflow subflows, and a
// node does not keep a copy of its primitive list (parallel_bvh.h:68).
// Traversal is NOT here: it happens on the GPU (frt_traverse.cuh).
#pragma once
#include <cstdlib>
#include <future>
#include <thread>
#include "scene_api.h"

class parallel_bvh_node : public hitable_list
{
public:
    parallel_bvh_node() {}
    bool bounding_box(double, double, aabb &b) const override { b = box; return true; }
    static parallel_bvh_node *create_bvh(hitable **l, int n, double time0, double time1);

    std::unique_ptr<hitable> left, right;
    aabb box;

private:
    struct item { hitable *h; aabb b; };
    static parallel_bvh_node *build(item *it, hitable **l, int n, int depth);
    template <int AXIS>
    static int cmp_axis(const void *a, const void *b) // box_{x,y,z}_compare, bvh.h:6-55
    {
        const item *ia = *(const item *const *)a, *ib = *(const item *const *)b;
        return ia->b.min[AXIS] - ib->b.min[AXIS] < 0.0 ? -1 : 1;
    }
};

inline parallel_bvh_node *parallel_bvh_node::build(item *it, hitable **l, int n, int depth)
{
    parallel_bvh_node *node = new parallel_bvh_node;
    ++Scene::num_bvh_nodes;
    aabb main_box = it[0].b;
    for (int i = 1; i < n; ++i) main_box = surrounding_box(it[i].b, main_box);
    node->box = main_box;
    if (n == 1)
    {
        // a one-primitive scene: the reference would read l[1] (parallel_bvh.h:143-149); here the
        // single primitive becomes the left leaf and the right child stays empty
        ++Scene::num_bvh_nodes; ++Scene::num_bvh_leaf_nodes;
        node->left.reset(it[0].h);
        return node;
    }
    const int axis = main_box.longest_axis();
    // Same call as parallel_bvh.h:93-98: the C library's qsort over 8-byte elements with a comparator
    // that never returns 0 (bvh.h:6-55).  Primitives with equal box.min (the Cornell boxes stand on
    // the floor) end up in whatever order that qsort leaves them in, and that order decides which of
    // two coincident triangles parallel_bvh_node::hit meets first — so the same routine is used here.
    {
        std::vector<const item *> order(n);
        for (int i = 0; i < n; ++i) order[i] = it + i;
        std::qsort(order.data(), (size_t)n, sizeof(const item *), axis == 0 ? cmp_axis<0> : (axis == 1 ? cmp_axis<1> : cmp_axis<2>));
        std::vector<item> sorted(n);
        for (int i = 0; i < n; ++i) sorted[i] = *order[i];
        std::copy(sorted.begin(), sorted.end(), it);
    }
    for (int i = 0; i < n; ++i) l[i] = it[i].h;

    std::vector<double> left_area(n), right_area(n);
    aabb acc = it[0].b;
    left_area[0] = acc.area();
    for (int i = 1; i < n - 1; ++i) { acc = surrounding_box(acc, it[i].b); left_area[i] = acc.area(); }
    acc = it[n - 1].b;
    right_area[n - 1] = acc.area();
    for (int i = n - 2; i > 0; --i) { acc = surrounding_box(acc, it[i].b); right_area[i] = acc.area(); }
    double best = FLT_MAX;
    int idx = 0;
    for (int i = 0; i < n - 1; ++i)
    {
        const double sah = i * left_area[i] + (n - i - 1) * right_area[i + 1];
        if (sah < best) { idx = i; best = sah; }
    }

    const int nl = idx + 1, nr = n - idx - 1;
    std::future<parallel_bvh_node *> pending;
    const bool spawn = depth < 6 && nl > 4096 && nr > 4096;
    if (idx == 0) { ++Scene::num_bvh_nodes; ++Scene::num_bvh_leaf_nodes; node->left.reset(it[0].h); }
    else if (spawn) pending = std::async(std::launch::async, build, it, l, nl, depth + 1);
    else node->left.reset(build(it, l, nl, depth + 1));
    if (idx == n - 2) { ++Scene::num_bvh_nodes; ++Scene::num_bvh_leaf_nodes; node->right.reset(it[idx + 1].h); }
    else node->right.reset(build(it + nl, l + nl, nr, depth + 1));
    if (pending.valid()) node->left.reset(pending.get());
    return node;
}

inline parallel_bvh_node *parallel_bvh_node::create_bvh(hitable **l, int n, double time0, double time1)
{
    if (n <= 0) return nullptr;
    std::vector<item> items(n);
    for (int i = 0; i < n; ++i) { items[i].h = l[i]; l[i]->bounding_box(time0, time1, items[i].b); }
    return build(items.data(), l, n, 0);
}
