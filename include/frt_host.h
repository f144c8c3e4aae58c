/* frt_host.h — C ABI of the C++ host layer (libfrt_host.so, first_raytracer_b200/host), for
 * non-C++ callers (the Python tests and bench.py).  It exposes what a C++ user gets from
 * scenes.h + flatten.h + renderer<gpu_path_integrator>: build one of the north-star scenes with
 * the reference-shaped classes and host SAH BVH (parallel_bvh.h:67-175), flatten it to the POD
 * description of frt.h, and render through the integrator concept (integrator.h:11-47). */
#ifndef FRT_HOST_H
#define FRT_HOST_H
#include "frt.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct frth_scene frth_scene;

typedef struct frth_info {
    uint32_t n_prims, n_tris, n_spheres, n_lights, n_materials;
    uint32_t n_wide_nodes, wide_depth;
    uint32_t bvh_nodes, bvh_leaves; /* Scene::num_bvh_nodes / num_bvh_leaf_nodes after the build */
    double build_ms, flatten_ms;
    float bounds_lo[3], bounds_hi[3];
} frth_info;

FRT_API const char *frth_last_error(void);
/* name: cornell | cornell_ao | cornell_glossy | cornell_sphere | cornell_metal | cornell_lens | veach | veach_beckmann | teapot |
 * furnace | textured | random_scene | soup:<N>, optionally followed by +sky or +hdr (environment); assets are read
 * from $FRT_ASSETS (default ./assets).  max_leaf = triangles per wide-BVH leaf (1..16; 0 = chosen by scene size). */
FRT_API int frth_scene_create(const char *name, double aspect, int max_leaf, frth_scene **out);
/* Triangle soup given as n*9 floats (v0,v1,v2). */
FRT_API int frth_scene_create_soup(const float *tri9, size_t n, double aspect, int max_leaf, frth_scene **out);
/* Fast path for large soups: binned-SAH build straight into wide nodes,
 * without the reference-shaped classes.  Ray casting only (frth_render returns FRT_ERR_STATE). */
FRT_API int frth_bvh4_build(const float *tri9, size_t n, int max_leaf, double aspect, frth_scene **out);
FRT_API void frth_scene_destroy(frth_scene *s);
FRT_API const frt_scene_desc *frth_scene_desc(const frth_scene *s);
FRT_API int frth_scene_info(const frth_scene *s, frth_info *info);
/* SURVEY §8d config 5 generator (splitmix64 stream): n*9 floats */
FRT_API int frth_soup_triangles(size_t n, uint64_t seed, float edge, float *out);
/* Pinhole camera rays through the pixel centres of an nx*ny grid (camera::get_ray_as_pinhole,
 * camera.h:37-40), computed in double and rounded to float: n = nx*ny rays. */
FRT_API int frth_camera_rays(const frth_scene *s, int nx, int ny, frt_ray *out);
/* renderer<gpu_path_integrator>::Render on a headless viewer(nx, ny, ns): fills rgb_mean
 * (nx*ny*3, viewer::fout_image as float, row 0 = bottom). */
#define FRTH_RENDER_AO (FRT_INTEGRATOR_AO << 8)           /* or into `flags`: renderer<gpu_ao_integrator> */
#define FRTH_RENDER_NORMALS (FRT_INTEGRATOR_NORMALS << 8) /* or into `flags`: renderer<gpu_normals_integrator> */
#define FRTH_RENDER_DIRECT (FRT_INTEGRATOR_DIRECT << 8)   /* or into `flags`: renderer<gpu_ass_1_integrator> */
FRT_API int frth_render(frth_scene *s, int n_gpus, int nx, int ny, uint32_t ns, uint32_t seed, uint32_t flags,
                        float *rgb_mean, frt_stats *stats, double *wall_seconds);
/* What closing the preview window does in the reference (viewer.cpp:101-104 sets viewer::to_exit, path.cpp:127 polls
 * it): stops the frth_render in flight on another thread, which then returns FRTH_CANCELLED with a partial film.
 * FRT_ERR_STATE when nothing is rendering. */
#define FRTH_CANCELLED 1
FRT_API int frth_cancel(void);
/* viewer::out_image of the last frth_render: the 8-bit tone-mapped film (viewer.cpp:117-125), nx*ny*3 bytes. */
FRT_API int frth_last_ldr(unsigned char *out, size_t n_bytes);

/* The window of the reference (viewer.cpp:18-107 shows viewer::out_image through GL) does not exist on a headless
 * box: the 8-bit tone-mapped film goes to a PNG file instead (8-bit RGB, rows flipped so that row 0 = bottom ends up
 * at the bottom of the picture; image.cpp:30 flips on write too).  `frt_main --png FILE` calls it after the render. */
FRT_API int frth_write_png(const char *path, const unsigned char *rgb, int nx, int ny);

#ifdef __cplusplus
}
#endif
#endif
