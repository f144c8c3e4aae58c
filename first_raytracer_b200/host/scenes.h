// Scene builders for the north-star configurations (see scenes.cpp).
#pragma once
#include <string>
#include <unordered_map>
#include <vector>
#include "scene_api.h"

// creation-order primitive ids (the order the scene function filled its list in), the id space
// the parity tests and frt_hit::prim_id use
struct frt_scene_ids { std::unordered_map<const hitable *, int> id; };

Scene *cornell_box_obj(const double &aspect, const char *variant = "Original", frt_scene_ids *ids = nullptr);
Scene *veach_mis(const double &aspect, frt_scene_ids *ids = nullptr, const char *distribution = "ggx");
Scene *textured_scene(const double &aspect, frt_scene_ids *ids = nullptr);
Scene *teapot_scene(const double &aspect, frt_scene_ids *ids = nullptr);
Scene *soup_scene(const double &aspect, const float *tri9, size_t n, frt_scene_ids *ids = nullptr);
void soup_triangles(size_t n, uint64_t seed, float edge, std::vector<float> &out);
Scene *make_scene(const std::string &name, double aspect, frt_scene_ids *ids = nullptr);
