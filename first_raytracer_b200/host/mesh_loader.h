// OBJ import + triangle factories; interface of mesh_loader.h:18-22 and triangle.h:189-190.
#pragma once
#include <memory>
#include <string>
#include <vector>
#include "scene_api.h"

class mesh_loader
{
public:
    static std::vector<std::shared_ptr<triangle_mesh>> load_obj(std::string file, bool use_geometry_normals);
};

std::vector<std::shared_ptr<hitable>> create_triangle_mesh(const std::string &file, std::vector<hitable *> &lights,
                                                           bool use_geometry_normals = false);
std::vector<std::shared_ptr<hitable>> create_triangle_mesh(const std::string &file, const Matrix4x4 &toWorld, material *bsdf,
                                                           std::vector<hitable *> &lights, bool use_geometry_normals = false);
