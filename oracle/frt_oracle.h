/* TEST INFRASTRUCTURE — CPU restatement (plain C, float64) of the reference's hot path:
 * path::Li (path.cpp:4-116) over triangle::hit / sphere::hit / aabb::hit and the material / pdf
 * classes, fed the same POD scene description the CUDA core receives (include/frt.h) and the
 * same counter-based random numbers (oracle/frt_rng.h).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library, and
 * only as the checker.  The product (first_raytracer_b200/) never links, loads or calls it.
 *
 * Parity pin: tests/test_oracle.py checks this restatement against the reference's own object
 * code (oracle/_ref/frt_ref*, compiled from /root/reference by oracle/Makefile) and against the
 * golden vectors under tests/golden/ that were generated from it (scripts/make_golden.py).
 */
#ifndef FRT_ORACLE_H
#define FRT_ORACLE_H
#include "../include/frt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct frt_oracle_stats {
    uint64_t samples, closest_rays, shadow_rays, nonfinite;
    uint64_t node_visits, tri_tests;
} frt_oracle_stats;

/* world->hit(r, EPSILON, FLT_MAX) (any_hit = 0, path.cpp:10) or world->hit(r, EPSILON,
 * 1 - SHADOW_EPSILON) (any_hit = 1, path.cpp:50) for each ray.  use_bvh = 0: linear closest-hit
 * loop over every primitive (hitable_list::hit, hitable_list.cpp:4-21), independent of the
 * flattened nodes; use_bvh = 1: walks scene->nodes in float64.  prim = creation-order id or -1. */
FRT_API int frt_oracle_trace(const frt_scene_desc *scene, const frt_ray *rays, size_t n, int any_hit, int use_bvh,
                             int32_t *prim, double *t, double *u, double *v);

/* path.cpp:120-147 over the tile (full rows) calling the restated Li; rgb_sum
 * (nx*ny*3 doubles, row 0 = bottom) receives the SUM over [spp_begin, spp_end) of each pixel. */
FRT_API int frt_oracle_render(const frt_scene_desc *scene, const frt_tile *tile, int use_bvh, double *rgb_sum,
                              frt_oracle_stats *stats);

/* Unit entry for BSDF parity: material m at a surface with unit normal n and
 * unit wi (pointing away from the surface), incoming ray direction -wi, uniforms s[0..2]:
 *   out[0..2] scattered direction (specular materials: scatter(); lambertian: generate(s[0],s[1]))
 *   out[3] sampled_pdf (or -1)   out[4] pdf value(dir)   out[5..7] eval_bsdf(dir)
 *   out[8] pdf value(wo_probe)   out[9..11] eval_bsdf(wo_probe) */
FRT_API int frt_oracle_bsdf(const frt_material *m, const double n[3], const double wi[3], const double s[3],
                            const double wo_probe[3], double out[12]);

/* t of one given primitive (creation-order id) tested alone per ray, NaN when it is not hit: used by
 * the parity tests to recognise exact geometric ties between coincident surfaces. */
FRT_API int frt_oracle_prim_t(const frt_scene_desc *scene, const frt_ray *rays, const int32_t *prim_id, size_t n, double *t);

/* 1 when built with -DFRT_ORACLE_F32 (same code in float32: isolates rounding from logic) */
FRT_API int frt_oracle_is_f32(void);

#ifdef __cplusplus
}
#endif
#endif
