// renderer<I>: the framework side of the reference's duck-typed integrator concept
// (integrator.h:11-47).  An integrator provides
//     void Render(Scene*, viewer*, tf::Taskflow&)   — enqueue work on the taskflow
//     static constexpr bool using_custom_viewer
// and renderer<I>::Render opens the (here: headless) viewer, dispatches the taskflow, waits,
// prints the timing block and returns the elapsed seconds.
#pragma once
#include <chrono>
#include <iostream>
#include "scene_api.h"
#include "taskflow_shim.h"
#include "viewer.h"

template <typename integrator>
struct renderer : public integrator
{
private:
    // integrators with a `stats` member (frt_stats, 64-bit) vs. the reference's own (32-bit statics of Scene)
    template <class I> static auto ray_count(I *i) -> decltype((unsigned long long)i->stats.closest_rays) { return i->stats.closest_rays + i->stats.shadow_rays; }
    static unsigned long long ray_count(...) { return Scene::num_rays_traversed; }
    template <class I> static auto box_count(I *i) -> decltype((unsigned long long)i->stats.node_visits) { return i->stats.node_visits; }
    static unsigned long long box_count(...) { return Scene::num_ray_box_intersections; }
    template <class I> static auto tri_count(I *i) -> decltype((unsigned long long)i->stats.tri_tests) { return i->stats.tri_tests; }
    static unsigned long long tri_count(...) { return Scene::num_ray_tri_intersections; }

public:
    double Render(Scene *scene, viewer &film_viewer)
    {
        GLFWwindow *window = film_viewer.init();
        tf::Taskflow tf;
        integrator::Render(scene, &film_viewer, tf);
        const auto t1 = std::chrono::high_resolution_clock::now();
        auto future = tf.dispatch();
        if constexpr (integrator::using_custom_viewer)
            integrator::background_thread(future, window, &film_viewer);
        else
            film_viewer.background_thread(future, window);
        future.get(); // a failed task (e.g. no CUDA device) surfaces here instead of being dropped
        const auto t2 = std::chrono::high_resolution_clock::now();
        const double secs = std::chrono::duration<double>(t2 - t1).count();
        std::cout << "\nIt took me " << secs << " seconds to render." << std::endl;
        // the reference prints its 32-bit atomics here; a GPU frame overflows them (1.4e10 rays at 1080p x 1024 spp),
        // so integrators that carry 64-bit device counters (frt_stats) are printed from those
        const unsigned long long rays = ray_count(this), boxes = box_count(this), tris = tri_count(this);
        std::cout << "\n Number of ray traversals: " << rays;
        std::cout << "\n Number of ray-box intersections: " << boxes;
        std::cout << "\n Number of ray-triangle intersections: " << tris;
        std::cout << "\n Number of BVH nodes: " << Scene::num_bvh_nodes;
        std::cout << "\n Number of BVH leaves: " << Scene::num_bvh_leaf_nodes;
        std::cout << "\n Rays/second: " << ((double)rays / secs) / 1e+6 << "M/sec" << std::endl;
        return secs;
    }
};
