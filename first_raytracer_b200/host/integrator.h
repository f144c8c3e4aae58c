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
        std::cout << "\n Number of ray traversals: " << Scene::num_rays_traversed;
        std::cout << "\n Number of ray-box intersections: " << Scene::num_ray_box_intersections;
        std::cout << "\n Number of ray-triangle intersections: " << Scene::num_ray_tri_intersections;
        std::cout << "\n Number of BVH nodes: " << Scene::num_bvh_nodes;
        std::cout << "\n Number of BVH leaves: " << Scene::num_bvh_leaf_nodes;
        std::cout << "\n Rays/second: " << (Scene::num_rays_traversed / secs) / 1e+6 << "M/sec" << std::endl;
        return secs;
    }
};
