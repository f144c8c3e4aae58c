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
#include <cstdint>
#include <cstdio>
#include <stdexcept>
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
// texture.h:30-49: the flattener reads tex0 / tex1 / u_scale / v_scale; the lookup itself runs on the device
class checker_texture : public texture
{
public:
    checker_texture() {}
    checker_texture(texture *t0, texture *t1, const double u_scale_, const double v_scale_)
        : tex0(t0), tex1(t1), u_scale(u_scale_), v_scale(v_scale_) {}
    Vector3f value(const hit_record &) const override { throw std::runtime_error("checker_texture is evaluated on the device"); }
    texture *tex0 = nullptr, *tex1 = nullptr;
    double u_scale = 1, v_scale = 1;
};

// ---- image.h / image.cpp / texture.h:51-112 -------------------------------------------------------
// The reference loads HDR environment images through stb_image (image.cpp:11-22, stbi_loadf).  Here: a reader
// for the Radiance RGBE container itself (run-length coded or flat scanlines, "-Y h +X w" orientation), decoded the
// way stbi_loadf does: channel = mantissa * 2^(exponent - 136), exponent 0 = black; row 0 = top of the picture.
enum class formats { STBI_JPG, STBI_PNG, STBI_BMP, STBI_HDR };
class image
{
public:
    image() {}
    image(const std::string &filename_, const formats &format = formats::STBI_HDR) : filename(filename_), type(format)
    {
        if (format == formats::STBI_HDR) { load_rgbe(); dataf = pixels.data(); }
        else if (format == formats::STBI_BMP) { load_bmp(); data = bytes.data(); }
        else throw std::runtime_error("image: only Radiance .hdr and 24-bit .bmp files are read (" + filename + ")");
    }
    image(const float *rgb, int nx_, int ny_) : filename("memory"), type(formats::STBI_HDR), nx(nx_), ny(ny_), nn(3)
    {
        pixels.assign(rgb, rgb + (size_t)nx * ny * 3);
        dataf = pixels.data();
    }
    std::string filename;
    unsigned char *data = nullptr;
    formats type = formats::STBI_HDR;
    float *dataf = nullptr;
    int nx = 0, ny = 0, nn = 3;

private:
    std::vector<float> pixels;
    std::vector<unsigned char> bytes;
    // image.cpp:19-21: stbi_load with stbi_set_flip_vertically_on_load(true) -> row 0 of `data` is the BOTTOM row of
    // the picture, RGB order.  An uncompressed 24-bit BMP with positive height stores exactly that order (as BGR).
    void load_bmp()
    {
        FILE *f = std::fopen(filename.c_str(), "rb");
        if (!f) throw std::runtime_error("image: cannot open " + filename);
        unsigned char h[54];
        if (std::fread(h, 1, 54, f) != 54 || h[0] != 'B' || h[1] != 'M') { std::fclose(f); throw std::runtime_error("image: not a BMP file: " + filename); }
        auto u32 = [&](int at) { return (uint32_t)h[at] | ((uint32_t)h[at + 1] << 8) | ((uint32_t)h[at + 2] << 16) | ((uint32_t)h[at + 3] << 24); };
        const uint32_t offset = u32(10);
        const int w = (int)u32(18), hh = (int)u32(22);
        const int bpp = h[28] | (h[29] << 8);
        if (u32(14) < 40 || bpp != 24 || u32(30) != 0 || w <= 0 || hh == 0)
        { std::fclose(f); throw std::runtime_error("image: only uncompressed 24-bit BMP is read: " + filename); }
        nx = w; ny = hh < 0 ? -hh : hh; nn = 3;
        const size_t stride = ((size_t)nx * 3 + 3) & ~(size_t)3;
        std::vector<unsigned char> row(stride);
        bytes.resize((size_t)nx * ny * 3);
        std::fseek(f, (long)offset, SEEK_SET);
        for (int j = 0; j < ny; ++j)
        {
            if (std::fread(row.data(), 1, stride, f) != stride) { std::fclose(f); throw std::runtime_error("image: truncated " + filename); }
            const int dst = hh > 0 ? j : ny - 1 - j; // top-down files (negative height) are turned bottom-up
            for (int i = 0; i < nx; ++i)
            {
                unsigned char *o = &bytes[((size_t)dst * nx + i) * 3];
                o[0] = row[3 * i + 2]; o[1] = row[3 * i + 1]; o[2] = row[3 * i];
            }
        }
        std::fclose(f);
    }
    void load_rgbe()
    {
        FILE *f = std::fopen(filename.c_str(), "rb");
        if (!f) throw std::runtime_error("image: cannot open " + filename);
        std::vector<unsigned char> buf;
        unsigned char chunk[65536];
        size_t got;
        while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) buf.insert(buf.end(), chunk, chunk + got);
        std::fclose(f);
        size_t pos = 0;
        auto line = [&]() {
            std::string l;
            while (pos < buf.size() && buf[pos] != '\n') l.push_back((char)buf[pos++]);
            ++pos;
            return l;
        };
        const std::string magic = line();
        if (magic != "#?RADIANCE" && magic != "#?RGBE") throw std::runtime_error("image: not a Radiance file: " + filename);
        bool rgbe = false;
        for (;;)
        {
            const std::string l = line();
            if (l.empty()) break;
            if (l == "FORMAT=32-bit_rle_rgbe") rgbe = true;
            if (pos >= buf.size()) break;
        }
        if (!rgbe) throw std::runtime_error("image: unsupported Radiance format in " + filename);
        const std::string res = line();
        if (std::sscanf(res.c_str(), "-Y %d +X %d", &ny, &nx) != 2 || nx <= 0 || ny <= 0)
            throw std::runtime_error("image: unsupported orientation '" + res + "' in " + filename);
        nn = 3;
        pixels.resize((size_t)nx * ny * 3);
        std::vector<unsigned char> scan((size_t)nx * 4);
        auto need = [&](size_t n) { if (pos + n > buf.size()) throw std::runtime_error("image: truncated " + filename); };
        auto convert = [&](const unsigned char *rgbe4, float *out) {
            if (rgbe4[3] != 0)
            {
                const float f1 = std::ldexp(1.0f, (int)rgbe4[3] - (128 + 8));
                out[0] = rgbe4[0] * f1; out[1] = rgbe4[1] * f1; out[2] = rgbe4[2] * f1;
            }
            else out[0] = out[1] = out[2] = 0.f;
        };
        if (nx < 8 || nx >= 32768 || !(pos + 4 <= buf.size() && buf[pos] == 2 && buf[pos + 1] == 2 && !(buf[pos + 2] & 0x80)))
        {
            need((size_t)nx * ny * 4); // flat
            for (size_t i = 0; i < (size_t)nx * ny; ++i) convert(&buf[pos + 4 * i], &pixels[3 * i]);
            return;
        }
        for (int j = 0; j < ny; ++j)
        {
            need(4);
            if (buf[pos] != 2 || buf[pos + 1] != 2 || ((buf[pos + 2] << 8) | buf[pos + 3]) != nx)
                throw std::runtime_error("image: bad scanline header in " + filename);
            pos += 4;
            for (int k = 0; k < 4; ++k)
            {
                int i = 0;
                while (i < nx)
                {
                    need(1);
                    int count = buf[pos++];
                    if (count > 128)
                    {
                        count -= 128;
                        need(1);
                        const unsigned char value = buf[pos++];
                        if (count == 0 || i + count > nx) throw std::runtime_error("image: corrupt run in " + filename);
                        for (int z = 0; z < count; ++z) scan[(size_t)(i++) * 4 + k] = value;
                    }
                    else
                    {
                        if (count == 0 || i + count > nx) throw std::runtime_error("image: corrupt run in " + filename);
                        need((size_t)count);
                        for (int z = 0; z < count; ++z) scan[(size_t)(i++) * 4 + k] = buf[pos++];
                    }
                }
            }
            for (int i = 0; i < nx; ++i) convert(&scan[(size_t)i * 4], &pixels[((size_t)j * nx + i) * 3]);
        }
    }
};
class image_texture : public texture
{
public:
    image_texture() {}
    explicit image_texture(std::unique_ptr<image> img_) : img(std::move(img_)) {}
    Vector3f value(const hit_record &rec) const override // texture.h:60-86 (HDR branch; LDR images are looked up on the device)
    {
        if (img->type != formats::STBI_HDR) throw std::runtime_error("image_texture: LDR lookups run on the device");
        const int nx = img->nx, ny = img->ny;
        int i = (int)(rec.u * nx), j = (int)(rec.v * ny);
        if (i < 0 || i > nx) i = ((i % nx) + nx) % nx;
        if (j < 0 || j > ny) j = ((j % ny) + ny) % ny;
        if (i == nx) i = nx - 1;
        if (j == ny) j = ny - 1;
        const float *t = img->dataf + ((size_t)j * nx + i) * 3;
        return Vector3f(t[0], t[1], t[2]);
    }
    std::unique_ptr<image> img;
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
    explicit environment_map(std::string env_map_filename_) : env_map_filename(std::move(env_map_filename_)) // material.h:209-213
    {
        env_map_tex = std::make_unique<image_texture>(std::make_unique<image>(env_map_filename, formats::STBI_HDR));
    }
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
