// C ABI of the host layer (include/frt_host.h).
#include <thread>
#include <vector>
#include <algorithm>
#include "../../include/frt_host.h"
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include "fast_bvh.h"
#include "gpu_path_integrator.h"
#include "integrator.h"
#include "png_writer.h"
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

// The film of the render in flight (frth_cancel sets its to_exit, what the window's close button does in the
// reference, viewer.cpp:101-104) and a film kept between renders of one size: constructing a viewer zero-fills
// 56 MB at 1080p (viewer.cpp:10-16), more than the device needs to copy a finished frame back.
namespace {
std::mutex g_film_mutex;
viewer *g_rendering = nullptr;
std::unique_ptr<viewer> g_film_cache;
}

int frth_cancel(void)
{
    std::lock_guard<std::mutex> lock(g_film_mutex);
    if (!g_rendering) return FRT_ERR_STATE;
    g_rendering->to_exit = true;
    return FRT_OK;
}

int frth_render(frth_scene *s, int n_gpus, int nx, int ny, uint32_t ns, uint32_t seed, uint32_t flags, float *rgb_mean,
                frt_stats *stats, double *wall_seconds)
{
    if (!s || nx <= 0 || ny <= 0 || ns == 0) return FRT_ERR_INVALID;
    if (!s->scene) { g_error = "frth_render: scene was built by frth_bvh4_build (ray casting only)"; return FRT_ERR_STATE; }
    try
    {
        std::unique_ptr<viewer> film;
        {
            std::lock_guard<std::mutex> lock(g_film_mutex);
            if (g_film_cache && g_film_cache->nx == nx && g_film_cache->ny == ny && g_film_cache->ns == ns) film = std::move(g_film_cache);
        }
        if (!film) film.reset(new viewer(nx, ny, ns, 3));
        film->to_exit = false;
        renderer<gpu_path_integrator> render;
        render.n_gpus = n_gpus;
        render.seed = seed;
        render.flags = flags & 0xFFu;
        render.integrator = (flags >> 8) & 0xFFu; // FRTH_RENDER_* in frt_host.h
        render.max_leaf = s->max_leaf;
        render.prim_ids = &s->ids.id;
        render.mean_rgb_out = rgb_mean; // viewer::fout_image as float32, written by frt_readback_film beside the film
        { std::lock_guard<std::mutex> lock(g_film_mutex); g_rendering = film.get(); }
        double secs = 0;
        try { secs = render.Render(s->scene.get(), *film); }
        catch (...) { std::lock_guard<std::mutex> lock(g_film_mutex); g_rendering = nullptr; throw; }
        { std::lock_guard<std::mutex> lock(g_film_mutex); g_rendering = nullptr; g_film_cache = std::move(film); }
        if (stats) *stats = render.stats;
        if (wall_seconds) *wall_seconds = secs;
        return render.cancelled ? FRTH_CANCELLED : FRT_OK;
    }
    catch (const std::exception &e) { g_error = e.what(); return FRT_ERR_CUDA; }
}

// viewer::out_image of the last frth_render (the 8-bit tone-mapped film, viewer.cpp:117-125): nx * ny * 3 bytes
int frth_last_ldr(unsigned char *out, size_t n_bytes)
{
    std::lock_guard<std::mutex> lock(g_film_mutex);
    if (!g_film_cache || !out || n_bytes != (size_t)g_film_cache->nx * g_film_cache->ny * 3) return FRT_ERR_STATE;
    std::memcpy(out, g_film_cache->out_image.get(), n_bytes);
    return FRT_OK;
}

// Host utility: an 8-bit RGB film (row 0 = bottom, what frth_last_ldr returns) as a PNG file
int frth_write_png(const char *path, const unsigned char *rgb, int nx, int ny)
{
    if (!path || !rgb || nx <= 0 || ny <= 0) return FRT_ERR_INVALID;
    if (!frt_png::write_rgb8(path, rgb, nx, ny)) { g_error = std::string("cannot write ") + path; return FRT_ERR_INVALID; }
    return FRT_OK;
}

} // extern "C"
