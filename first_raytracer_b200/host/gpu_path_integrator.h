// gpu_path_integrator — drop-in for the reference's `path` integrator (path.h:8-18) behind
// renderer<> (integrator.h:11-47): select it with renderer<gpu_path_integrator> where main.cpp:517
// says renderer<path>.  Render() emplaces ONE task that flattens the Scene, uploads it through the
// C ABI (include/frt.h), renders the whole film on the context's GPUs and feeds the film exactly
// as path.cpp:143 does: one add_sample(pixel, sum over ns) per pixel.
// If libfrt cannot create a CUDA context the task throws: there is no CPU fallback.
#pragma once
#include <stdexcept>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "../../include/frt.h"
#include "flatten.h"
#ifndef FRT_WITH_REFERENCE_HEADERS
#include "taskflow_shim.h"
#include "viewer.h"
#endif

struct gpu_path_integrator
{
    // knobs the reference hard-codes: depth cap path.cpp:36, seed path.cpp:122
    int n_gpus = 1;
    uint32_t seed = 1;
    uint32_t max_depth = 34;
    int max_leaf = 0; // triangles per wide-BVH leaf; 0 = chosen by scene size (flatten.h: 2, or 4 for mesh scenes)
    uint32_t flags = 0;
    uint32_t integrator = FRT_INTEGRATOR_PATH; // which Li the tile runs; see gpu_ao_integrator / gpu_normals_integrator
    // resume: film sums of samples [0, spp_begin) from an earlier run; the render continues at
    // spp_begin and add_sample receives the sum over all film->ns samples.  film_sums_out receives the final sums.
    uint32_t spp_begin = 0;
    const std::vector<float> *resume_sums = nullptr;
    std::vector<float> *film_sums_out = nullptr;
    float *mean_rgb_out = nullptr; // optional float32 copy of the mean image (nx * ny * 3, row 0 = bottom)
    bool cancelled = false;        // the last render was stopped by viewer::to_exit
    frt_stats stats{};
    const std::unordered_map<const hitable *, int> *prim_ids = nullptr;

    void Render(Scene *scene, viewer *film_viewer, tf::Taskflow &tf)
    {
        tf.emplace([this, scene, film_viewer]() { run(scene, film_viewer); });
    }

    constexpr static bool using_custom_viewer = false;

    // The device context (streams, counters, peer access between the GPUs) is kept between renders of one
    // process: creating it costs more than a 1080p x 128 spp share of an 8-GPU render.  Scene and film are
    // per render.  Call this before exit (frt_main does) to give the devices back explicitly.
    static void release_cached_context()
    {
        std::lock_guard<std::mutex> lock(cache().mutex);
        if (cache().ctx) frt_destroy(cache().ctx);
        cache().ctx = nullptr;
    }

private:
    struct ctx_cache { std::mutex mutex; frt_ctx *ctx = nullptr; int n_gpus = 0; };
    static ctx_cache &cache() { static ctx_cache c; return c; }
    static frt_ctx *acquire(int n)
    {
        std::lock_guard<std::mutex> lock(cache().mutex);
        frt_ctx *ctx = cache().ctx;
        cache().ctx = nullptr;
        if (ctx && cache().n_gpus == n) return ctx;
        if (ctx) frt_destroy(ctx);
        ctx = nullptr;
        check(nullptr, frt_create(&ctx, nullptr, n), "frt_create");
        return ctx;
    }
    static void give_back(frt_ctx *ctx, int n)
    {
        std::lock_guard<std::mutex> lock(cache().mutex);
        if (cache().ctx) { frt_destroy(ctx); return; } // another render finished first: keep one
        cache().ctx = ctx;
        cache().n_gpus = n;
    }

    static void check(frt_ctx *ctx, int rc, const char *what)
    {
        if (rc == FRT_OK) return;
        const std::string msg = std::string(what) + ": " + frt_last_error(ctx);
        if (ctx) frt_destroy(ctx);
        throw std::runtime_error(msg);
    }

    void run(Scene *scene, viewer *film)
    {
        using clk = std::chrono::steady_clock;
        const bool timing = std::getenv("FRT_HOST_TIMING") != nullptr;
        auto t_prev = clk::now();
        auto lap = [&](const char *what) {
            if (!timing) return;
            const auto now = clk::now();
            std::fprintf(stderr, "[frt host] %-10s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
            t_prev = now;
        };
        cancelled = false;
        frt_flattener flattener(prim_ids, max_leaf);
        const frt_flat_scene flat = flattener.flatten(scene);
        const frt_scene_desc desc = flat.desc();
        lap("flatten");
        frt_ctx *ctx = acquire(n_gpus);
        lap("context");
        check(ctx, frt_set_flags(ctx, flags), "frt_set_flags");
        if (resume_sums)
            check(ctx, frt_restore_film(ctx, (uint32_t)film->nx, (uint32_t)film->ny, resume_sums->data(), resume_sums->size()), "frt_restore_film");
        else
            check(ctx, frt_clear_film(ctx), "frt_clear_film");
        check(ctx, frt_upload_scene(ctx, &desc), "frt_upload_scene");
        frt_tile tile{};
        tile.integrator = integrator;
        tile.nx = (uint32_t)film->nx; tile.ny = (uint32_t)film->ny;
        tile.x0 = tile.y0 = 0; tile.w = tile.nx; tile.h = tile.ny;
        tile.spp_begin = spp_begin; tile.spp_end = (uint32_t)film->ns;
        tile.seed = seed; tile.max_depth = max_depth;
        lap("upload");
        check(ctx, frt_render_tile(ctx, &tile), "frt_render_tile");
        lap("launch");
        // viewer::to_exit (set by the window's close button, viewer.cpp:101-104) is what path.cpp:127 polls per row:
        // here the host watches it while the tile runs and forwards it to the device (frt_cancel), so a render
        // of any length stops within milliseconds
        for (;;)
        {
            int done = 0;
            check(ctx, frt_poll(ctx, &done), "frt_poll");
            if (done) break;
            if (film->to_exit) { check(ctx, frt_cancel(ctx), "frt_cancel"); cancelled = true; break; }
            std::this_thread::sleep_for(std::chrono::microseconds(200));
        }
        // path.cpp:143 / viewer.cpp:109-132 for every pixel at once: the film's mean goes to viewer::fout_image, the
        // tone-mapped bytes (computed on the device, in double like the reference) to viewer::out_image; with
        // film_sums_out the raw sums are fetched as well (checkpoints)
        if (film_sums_out)
        {
            film_sums_out->resize((size_t)film->nx * film->ny * 3);
            check(ctx, frt_readback(ctx, film_sums_out->data(), film_sums_out->size(), nullptr), "frt_readback");
        }
        check(ctx, frt_readback_film(ctx, (uint64_t)film->ns, film->fout_image.get(), film->out_image.get(), mean_rgb_out, &stats), "frt_readback_film");
        lap("readback");
        give_back(ctx, n_gpus);
        // the reference's statistics block reads these 32-bit statics (integrator.h:33-43); a 1080p x 1024 spp frame
        // traces 1.4e10 rays, so they saturate instead of wrapping and `stats` (64-bit) is what the renderer prints
        auto sat = [](uint64_t v) { return (unsigned int)std::min<uint64_t>(v, 0xFFFFFFFFull); };
        Scene::num_rays_traversed = sat(stats.closest_rays + stats.shadow_rays);
        Scene::num_ray_box_intersections = sat(stats.node_visits);
        Scene::num_ray_tri_intersections = sat(stats.tri_tests);
    }
};

// The reference's other integrators that are one shade stage over the same trace kernels:
// drop-ins for `ao` (ao.h:8-47, ao.cpp:4-28) and `normals_renderer` (debug_renderer.h:6-47) behind renderer<>.
struct gpu_ao_integrator : gpu_path_integrator { gpu_ao_integrator() { integrator = FRT_INTEGRATOR_AO; } };
struct gpu_normals_integrator : gpu_path_integrator { gpu_normals_integrator() { integrator = FRT_INTEGRATOR_NORMALS; } };
// `ass_1_renderer` (assignment_1.hpp:8-19, assignment_1.cpp:4-38): unshadowed direct lighting, every light sampled once.
struct gpu_ass_1_integrator : gpu_path_integrator { gpu_ass_1_integrator() { integrator = FRT_INTEGRATOR_DIRECT; } };
