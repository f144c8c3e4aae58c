// Host-side scene description classes with the reference's names and public members
// (hitable.h, hitable_list.h, aabb.h, triangle.h, sphere.h, texture.h, material.h, camera.h,
// Scene.h), so that the reference's scene-building functions (main.cpp:222-314) can be written
// against them verbatim and the flattener (flatten.h) walks the same fields it would walk in
// the reference tree.
//
// These classes only DESCRIBE a scene.  They have no hit(), scatter() or eval_bsdf(): in this
// product every ray is traced and every BSDF evaluated by the CUDA core (libfrt.so) — there is
// no CPU rendering path to fall back to.
#pragma once
#include <algorithm>
#include <atomic>
#include <cctype>
#include <cfloat>
#include <cmath>
#include <memory>
#include <string>
#include <vector>
#include "geometry.h"

constexpr double EPSILON = 1e-4; // util.h:10

// util.h:62-71
inline double FromSrgb(double v) { return v <= 0.04045 ? v * (1.0 / 12.92) : std::pow((v + 0.055) * (1.0 / 1.055), 2.4); }
inline Vector3f FromSrgb(const Vector3f &v) { return {FromSrgb(v[0]), FromSrgb(v[1]), FromSrgb(v[2])}; }

enum class microfacet_distributions { ggx = 0, beckmann = 1 }; // util.h:48-52

// ---- aabb.h -----------------------------------------------------------------------------------
class aabb
{
public:
    aabb() {}
    aabb(const Vector3f &lo, const Vector3f &hi) : min(lo), max(hi), size(hi - lo) {}
    // aabb.h:33-43 — kept as is: z is only considered when y does not beat x
    int longest_axis() const { return size.y() > size.x() ? 1 : (size.z() > size.x() ? 2 : 0); }
    double area() const { return 2 * (size.x() * size.y() + size.y() * size.z() + size.x() * size.z()); }
    Vector3f min, max, size;
};
inline aabb surrounding_box(const aabb &a, const aabb &b)
{
    return aabb(Vector3f(std::fmin(a.min.x(), b.min.x()), std::fmin(a.min.y(), b.min.y()), std::fmin(a.min.z(), b.min.z())),
                Vector3f(std::fmax(a.max.x(), b.max.x()), std::fmax(a.max.y(), b.max.y()), std::fmax(a.max.z(), b.max.z())));
}

// ---- hitable.h ----------------------------------------------------------------------------------
class material;
class hitable;
struct hit_record // only what constant textures and the flattener look at
{
    double t = 0, u = 0, v = 0;
    Vector2f uv;
    Vector3f p, wi, normal;
    material *mat_ptr = nullptr;
    hitable *obj = nullptr;
};

class hitable
{
public:
    virtual bool bounding_box(double t0, double t1, aabb &box) const = 0;
    virtual ~hitable() {}
};

// ---- hitable_list.h -------------------------------------------------------------------------------
class hitable_list : public hitable
{
public:
    hitable_list() : list_size(0) {}
    hitable_list(const std::vector<hitable *> &l, int n) : list_size(n), list(l) {}
    bool bounding_box(double t0, double t1, aabb &box) const override
    {
        if (list_size < 1) return false;
        aabb tmp;
        for (int i = 0; i < list_size; ++i)
        {
            if (!list[i]->bounding_box(t0, t1, tmp)) return false;
            box = i == 0 ? tmp : surrounding_box(box, tmp);
        }
        return true;
    }
    hitable *operator[](int i) const { return list[i]; }
    int list_size;
    std::vector<hitable *> list;
};

// ---- texture.h ----------------------------------------------------------------------------------
class texture
{
public:
    virtual Vector3f value(const hit_record &rec) const = 0;
    virtual ~texture() {}
};
class constant_texture : public texture
{
public:
    constant_texture() {}
    explicit constant_texture(const Vector3f &c) : color(c) {}
    Vector3f value(const hit_record &) const override { return color; }
    Vector3f color;
};

// ---- material.h ---------------------------------------------------------------------------------
class material
{
public:
    virtual ~material() {}
};
class lambertian : public material
{
public:
    explicit lambertian(texture *a) : albedo(a) {}
    texture *albedo;
};
class modified_phong : public material
{
public:
    modified_phong(texture *kd, texture *ks, double exponent)
        : diffuse_reflectance(kd), specular_reflectance(ks), specular_exponent(exponent) {}
    texture *diffuse_reflectance, *specular_reflectance;
    const double specular_exponent;
};
class metal : public material
{
public:
    explicit metal(const Vector3f &a) : albedo(a) {}
    Vector3f albedo;
};
class dielectric : public material
{
public:
    dielectric(double ior, texture *ks, texture *kt) : ref_idx(ior), specular_reflectance(ks), specular_transmittance(kt) {}
    double ref_idx;
    texture *specular_reflectance, *specular_transmittance;
};
class diffuse_light : public material
{
public:
    explicit diffuse_light(texture *a) : emit(a) {}
    texture *emit;
};
class rough_conductor : public material
{
public:
    rough_conductor(double alpha_, double extEta_, const Vector3f &eta_, const Vector3f &k_, texture *refl,
                    const std::string &distribution_)
        : alpha(alpha_), extEta(extEta_), alphaU(alpha_), alphaV(alpha_), eta(eta_), k(k_), specular_reflectance(refl),
          distribution(distribution_)
    {
        for (char &c : distribution) c = (char)std::tolower((unsigned char)c);
        distribution_type = distribution == "ggx" ? microfacet_distributions::ggx : microfacet_distributions::beckmann;
    }
    double alpha, extEta, alphaU, alphaV;
    Vector3f eta, k;
    texture *specular_reflectance;
    std::string distribution;
    microfacet_distributions distribution_type;
};
class environment_map : public material
{
public:
    explicit environment_map(std::unique_ptr<texture> e) : env_map_tex(std::move(e)) {}
    std::string env_map_filename;
    std::unique_ptr<texture> env_map_tex;
};

// ---- triangle.h ---------------------------------------------------------------------------------
class triangle_mesh
{
public:
    triangle_mesh() : nTriangles(0), nVertices(0), use_geometry_normals(false) {}
    // Takes ownership of v / n / uv_ when shallow_copy (what mesh_loader does), else copies.
    triangle_mesh(int nTriangles_, int nVertices_, Vector3f *v, const int *idx, Vector3f *n, Vector2f *uv_,
                  std::unique_ptr<material> m, bool use_geometry_normals_, std::string name_, bool shallow_copy = false)
        : nTriangles(nTriangles_), nVertices(nVertices_), indices(idx, idx + 3 * nTriangles_), mat(std::move(m)),
          use_geometry_normals(use_geometry_normals_), name(std::move(name_))
    {
        if (shallow_copy) { vertices.reset(v); normals.reset(n); uv.reset(uv_); }
        else
        {
            vertices.reset(new Vector3f[nVertices]); normals.reset(new Vector3f[nVertices]); uv.reset(new Vector2f[nVertices]);
            std::copy(v, v + nVertices, vertices.get());
            std::copy(n, n + nVertices, normals.get());
            std::copy(uv_, uv_ + nVertices, uv.get());
        }
    }
    int nTriangles, nVertices;
    std::vector<int> indices;
    std::unique_ptr<Vector3f[]> vertices, normals;
    std::unique_ptr<Vector2f[]> uv;
    std::unique_ptr<material> mat;
    bool use_geometry_normals;
    std::string name;
};

class triangle : public hitable
{
public:
    triangle(const std::shared_ptr<triangle_mesh> &mesh_, int tri_num)
        : V(&mesh_->indices[3 * tri_num]),
          edge1(mesh_->vertices[V[1]] - mesh_->vertices[V[0]]), edge2(mesh_->vertices[V[2]] - mesh_->vertices[V[0]]),
          mesh(mesh_), mat_ptr(mesh_->mat.get())
    {
        inv_area = 1 / (0.5 * cross(edge1, edge2).length() * mesh->nTriangles); // triangle.h:65
    }
    bool bounding_box(double, double, aabb &b) const override
    {
        const Vector3f &a = mesh->vertices[V[0]], &c = mesh->vertices[V[1]], &d = mesh->vertices[V[2]];
        b = aabb(Vector3f(std::fmin(std::fmin(a.x(), c.x()), d.x()), std::fmin(std::fmin(a.y(), c.y()), d.y()),
                          std::fmin(std::fmin(a.z(), c.z()), d.z())),
                 Vector3f(std::fmax(std::fmax(a.x(), c.x()), d.x()), std::fmax(std::fmax(a.y(), c.y()), d.y()),
                          std::fmax(std::fmax(a.z(), c.z()), d.z())));
        return true;
    }
    const int *V;
    double inv_area;
    const Vector3f edge1, edge2;
    std::shared_ptr<triangle_mesh> mesh;
    material *mat_ptr;
};

// ---- sphere.h -----------------------------------------------------------------------------------
class sphere : public hitable
{
public:
    sphere(const Vector3f &cen, double r, material *m) : center(cen), radius(r), mat_ptr(m) {}
    bool bounding_box(double, double, aabb &b) const override
    {
        b = aabb(center - Vector3f(radius, radius, radius), center + Vector3f(radius, radius, radius));
        return true;
    }
    Vector3f center;
    double radius;
    material *mat_ptr;
};

// ---- camera.h -----------------------------------------------------------------------------------
struct camera
{
    camera() : lens_radius(0), dist(0) {}
    camera(Vector3f lookfrom, Vector3f lookat, Vector3f vup, double vfov, double aspect, double aperture, double focus_dist)
    {
        lens_radius = aperture / 2;
        const double half_height = std::tan(vfov * M_PI / 180.0 / 2);
        const double half_width = aspect * half_height;
        origin = lookfrom;
        w = unit_vector(lookfrom - lookat);
        u = unit_vector(cross(vup, w));
        v = cross(w, u);
        lower_left_corner = origin - half_width * focus_dist * u - half_height * focus_dist * v - focus_dist * w;
        horizontal = 2 * half_width * focus_dist * u;
        vertical = 2 * half_height * focus_dist * v;
        dist = 512.0 / (2 * half_height);
    }
    Vector3f origin, lower_left_corner, horizontal, vertical, u, v, w;
    double lens_radius, dist;
};

// ---- Scene.h ------------------------------------------------------------------------------------
struct Scene
{
    std::unique_ptr<hitable> world;
    std::unique_ptr<environment_map> env_map;
    hitable_list lights;
    camera cam;
    // The reference bumps these on every node visit (parallel_bvh.h:41,45); here only the host
    // BVH build touches the node counters, the ray counters come back from the device.
    static std::atomic<unsigned int> num_rays_traversed, num_ray_tri_intersections, num_ray_box_intersections,
        num_bvh_nodes, num_bvh_leaf_nodes;
    Scene(hitable *world_, environment_map *env, const camera &cam_, const std::vector<hitable *> &l)
        : world(world_), env_map(env), lights(l, (int)l.size()), cam(cam_) {}
};
