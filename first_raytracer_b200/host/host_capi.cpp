// C ABI of the host layer (include/frt_host.h).
#include <thread>
#include <vector>
#include <algorithm>
#include "../../include/frt_host.h"
#include <chrono>
#include <cstring>
#include <memory>
#include <string>
#include "fast_bvh.h"
#include "gpu_path_integrator.h"
#include "integrator.h"
#include "scenes.h"

struct frth_scene
{
    std::unique_ptr<Scene> scene;
    frt_scene_ids ids;
    frt_flat_scene flat;
    frt_scene_desc desc;
    frth_info info;
    int max_leaf;
};

namespace {
thread_local std::string g_error;
double ms_since(std::chrono::steady_clock::time_point t0)
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

int finish_scene(frth_scene *s, double build_ms, frth_scene **out)
{
    const auto t1 = std::chrono::steady_clock::now();
    frt_flattener fl(&s->ids.id, s->max_leaf);
    s->flat = fl.flatten(s->scene.get());
    s->desc = s->flat.desc();
    std::memset(&s->info, 0, sizeof s->info);
    s->info.flatten_ms = ms_since(t1);
    s->info.build_ms = build_ms;
    s->info.n_tris = s->desc.n_tris; s->info.n_spheres = s->desc.n_spheres;
    s->info.n_prims = s->desc.n_tris + s->desc.n_spheres;
    s->info.n_lights = s->desc.n_lights; s->info.n_materials = s->desc.n_materials;
    s->info.n_wide_nodes = s->desc.n_nodes; s->info.wide_depth = (uint32_t)s->flat.max_depth_wide;
    s->info.bvh_nodes = Scene::num_bvh_nodes; s->info.bvh_leaves = Scene::num_bvh_leaf_nodes;
    aabb box;
    if (s->scene->world && s->scene->world->bounding_box(0, 0, box))
        for (int a = 0; a < 3; ++a) { s->info.bounds_lo[a] = (float)box.min[a]; s->info.bounds_hi[a] = (float)box.max[a]; }
    *out = s;
    return FRT_OK;
}
} // namespace

extern "C" {

const char *frth_last_error(void) { return g_error.c_str(); }

int frth_scene_create(const char *name, double aspect, int max_leaf, frth_scene **out)
{
    if (!name || !out) return FRT_ERR_INVALID;
    *out = nullptr;
    try
    {
        std::unique_ptr<frth_scene> s(new frth_scene);
        s->max_leaf = max_leaf;
        Scene::num_bvh_nodes = 0; Scene::num_bvh_leaf_nodes = 0;
        const auto t0 = std::chrono::steady_clock::now();
        s->scene.reset(make_scene(name, aspect, &s->ids));
        return finish_scene(s.release(), ms_since(t0), out);
    }
    catch (const std::exception &e) { g_error = e.what(); return FRT_ERR_INVALID; }
}

int frth_scene_create_soup(const float *tri9, size_t n, double aspect, int max_leaf, frth_scene **out)
{
    if (!tri9 || !out || n == 0) return FRT_ERR_INVALID;
    *out = nullptr;
    try
    {
        std::unique_ptr<frth_scene> s(new frth_scene);
        s->max_leaf = max_leaf;
        Scene::num_bvh_nodes = 0; Scene::num_bvh_leaf_nodes = 0;
        const auto t0 = std::chrono::steady_clock::now();
        s->scene.reset(soup_scene(aspect, tri9, n, &s->ids));
        return finish_scene(s.release(), ms_since(t0), out);
    }
    catch (const std::exception &e) { g_error = e.what(); return FRT_ERR_INVALID; }
}

int frth_bvh4_build(const float *tri9, size_t n, int max_leaf, double aspect, frth_scene **out)
{
    if (!tri9 || !out || n == 0 || n > FRT_LEAF_FIRST_MASK) return FRT_ERR_INVALID;
    *out = nullptr;
    try
    {
        std::unique_ptr<frth_scene> s(new frth_scene);
        s->max_leaf = max_leaf;
        const auto t0 = std::chrono::steady_clock::now();
        s->flat = frt_fast::build_soup(tri9, n, max_leaf);
        // same camera as soup_scene() (scenes.cpp)
        const camera c(Vector3f(0.5, 0.5, 3.0), Vector3f(0.5, 0.5, 0.5), Vector3f(0, 1, 0), 30.0, aspect, 0.0, 10.0);
        auto put = [](float *d, const Vector3f &v) { d[0] = (float)v[0]; d[1] = (float)v[1]; d[2] = (float)v[2]; };
        frt_camera &fc = s->flat.camera;
        put(fc.origin, c.origin); put(fc.lower_left_corner, c.lower_left_corner); put(fc.horizontal, c.horizontal);
        put(fc.vertical, c.vertical); put(fc.u, c.u); put(fc.v, c.v); put(fc.w, c.w);
        fc.lens_radius = 0.f;
        s->desc = s->flat.desc();
        std::memset(&s->info, 0, sizeof s->info);
        s->info.build_ms = ms_since(t0);
        s->info.n_tris = s->info.n_prims = s->desc.n_tris;
        s->info.n_materials = s->desc.n_materials;
        s->info.n_wide_nodes = s->desc.n_nodes; s->info.wide_depth = (uint32_t)s->flat.max_depth_wide;
        *out = s.release();
        return FRT_OK;
    }
    catch (const std::exception &e) { g_error = e.what(); return FRT_ERR_INVALID; }
}

void frth_scene_destroy(frth_scene *s) { delete s; }

const frt_scene_desc *frth_scene_desc(const frth_scene *s) { return s ? &s->desc : nullptr; }

int frth_scene_info(const frth_scene *s, frth_info *info)
{
    if (!s || !info) return FRT_ERR_INVALID;
    *info = s->info;
    return FRT_OK;
}

int frth_soup_triangles(size_t n, uint64_t seed, float edge, float *out)
{
    if (!out) return FRT_ERR_INVALID;
    std::vector<float> t;
    soup_triangles(n, seed, edge, t);
    std::memcpy(out, t.data(), t.size() * sizeof(float));
    return FRT_OK;
}

int frth_camera_rays(const frth_scene *s, int nx, int ny, frt_ray *out)
{
    if (!s || !out || nx <= 0 || ny <= 0 || !s->scene) return FRT_ERR_INVALID;
    const camera &c = s->scene->cam;
    for (int y = 0; y < ny; ++y)
        for (int x = 0; x < nx; ++x)
        {
            const double u = (x + 0.5) / nx, v = (y + 0.5) / ny;
            const Vector3f d = c.lower_left_corner + u * c.horizontal + v * c.vertical - c.origin;
            frt_ray &r = out[(size_t)y * nx + x];
            r.ox = (float)c.origin[0]; r.oy = (float)c.origin[1]; r.oz = (float)c.origin[2];
            r.dx = (float)d[0]; r.dy = (float)d[1]; r.dz = (float)d[2];
        }
    return FRT_OK;
}

int frth_render(frth_scene *s, int n_gpus, int nx, int ny, uint32_t ns, uint32_t seed, uint32_t flags, float *rgb_mean,
                frt_stats *stats, double *wall_seconds)
{
    if (!s || nx <= 0 || ny <= 0 || ns == 0) return FRT_ERR_INVALID;
    if (!s->scene) { g_error = "frth_render: scene was built by frth_bvh4_build (ray casting only)"; return FRT_ERR_STATE; }
    try
    {
        viewer film(nx, ny, ns, 3);
        renderer<gpu_path_integrator> render;
        render.n_gpus = n_gpus;
        render.seed = seed;
        render.flags = flags & 0xFFu;
        render.integrator = (flags >> 8) & 0xFFu; // FRTH_RENDER_* in frt_host.h
        render.max_leaf = s->max_leaf;
        render.prim_ids = &s->ids.id;
        const double secs = render.Render(s->scene.get(), film);
        if (rgb_mean)
        {
            // viewer::fout_image (double) -> caller's float32 buffer, dealt to the host cores
            const size_t n = (size_t)nx * ny * 3;
            const unsigned n_workers = std::max(1u, std::min(std::thread::hardware_concurrency(), (unsigned)(n >> 18) + 1u));
            auto convert = [&](unsigned w) {
                for (size_t i = n * w / n_workers, e = n * (w + 1) / n_workers; i < e; ++i) rgb_mean[i] = (float)film.fout_image[i];
            };
            std::vector<std::thread> workers;
            for (unsigned w = 1; w < n_workers; ++w) workers.emplace_back(convert, w);
            convert(0);
            for (std::thread &t : workers) t.join();
        }
        if (stats) *stats = render.stats;
        if (wall_seconds) *wall_seconds = secs;
        return FRT_OK;
    }
    catch (const std::exception &e) { g_error = e.what(); return FRT_ERR_CUDA; }
}

} // extern "C"
