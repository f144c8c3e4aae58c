// OBJ / MTL import for the host layer.  Same interface as the reference's mesh_loader
// (mesh_loader.h:18-22) and create_triangle_mesh factories (triangle.h:189-190, triangle.cpp:9-60)
// but self-contained: the reference delegates parsing to Assimp 5.0.1, which is not part of this
// product.  What is kept from the reference is what decides pixel values:
//   * one triangle_mesh per (group, material) in first-appearance order, polygons fan-triangulated
//   * MTL → material mapping of mesh_loader.cpp:59-112 (Ke → diffuse_light; Ks & d<1 →
//     dielectric; Ks → modified_phong; else lambertian; colours through FromSrgb except Ke)
//   * angle-weighted vertex normals when the file has none (mesh_loader.cpp:114-155)
//   * emissive triangles are appended to `lights` in creation order (triangle.cpp:15-20)
#include "mesh_loader.h"
#include <array>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>

namespace {

struct mtl_record
{
    float Kd[3] = {0.6f, 0.6f, 0.6f}, Ks[3] = {0.f, 0.f, 0.f}, Ke[3] = {0.f, 0.f, 0.f};
    float Ns = 0.f, Ni = 1.f, d = 1.f;
};

void parse_mtl(const std::string &path, std::map<std::string, mtl_record> &out)
{
    std::ifstream in(path);
    std::string line, key, current;
    while (std::getline(in, line))
    {
        const size_t c = line.find('#');
        if (c != std::string::npos) line.erase(c);
        std::istringstream ls(line);
        if (!(ls >> key)) continue;
        if (key == "newmtl") { ls >> current; out[current] = mtl_record(); continue; }
        if (current.empty()) continue;
        mtl_record &r = out[current];
        float *dst3 = key == "Kd" ? r.Kd : key == "Ks" ? r.Ks : key == "Ke" ? r.Ke : nullptr;
        if (dst3) ls >> dst3[0] >> dst3[1] >> dst3[2];
        else if (key == "Ns") ls >> r.Ns;
        else if (key == "Ni") ls >> r.Ni;
        else if (key == "d") ls >> r.d;
    }
}

bool any_nonzero(const float *c) { return c[0] != 0.f || c[1] != 0.f || c[2] != 0.f; }
Vector3f srgb3(const float *c) { return Vector3f(FromSrgb(c[0]), FromSrgb(c[1]), FromSrgb(c[2])); }

std::unique_ptr<material> make_material(const mtl_record *r)
{
    if (!r) return std::make_unique<lambertian>(new constant_texture(Vector3f(0.5, 0.5, 0.5)));
    if (any_nonzero(r->Ke)) return std::make_unique<diffuse_light>(new constant_texture(Vector3f(r->Ke[0], r->Ke[1], r->Ke[2])));
    if (any_nonzero(r->Ks))
    {
        if ((double)r->d < 1.0)
            return std::make_unique<dielectric>((double)r->Ni, new constant_texture(srgb3(r->Ks)),
                                                new constant_texture(Vector3f(1.0, 1.0, 1.0)));
        return std::make_unique<modified_phong>(new constant_texture(srgb3(r->Kd)), new constant_texture(srgb3(r->Ks)), (double)r->Ns);
    }
    return std::make_unique<lambertian>(new constant_texture(srgb3(r->Kd)));
}

struct corner_key
{
    int v, t, n;
    bool operator<(const corner_key &o) const { return v != o.v ? v < o.v : (t != o.t ? t < o.t : n < o.n); }
};

struct mesh_builder
{
    std::string group, mtl;
    std::map<corner_key, int> seen;
    std::vector<Vector3f> pos, nrm;
    std::vector<Vector2f> tex;
    std::vector<int> idx;
    bool all_normals = true;
};

int resolve(int i, size_t count) { return i < 0 ? (int)count + i + 1 : i; }

// util.h:14-19
double unit_angle(const Vector3f &u, const Vector3f &v)
{
    if (dot(u, v) < 0) return M_PI - 2 * std::asin(0.5 * (v + u).length());
    return 2 * std::asin(0.5 * (v - u).length());
}

void facet_normals(mesh_builder &m)
{
    const int nt = (int)m.idx.size() / 3;
    std::fill(m.nrm.begin(), m.nrm.end(), Vector3f(0, 0, 0));
    for (int f = 0; f < nt; ++f)
    {
        const int *V = &m.idx[3 * f];
        Vector3f fn(0, 0, 0);
        for (int c = 0; c < 3; ++c)
        {
            const Vector3f &a = m.pos[V[c]], &b = m.pos[V[(c + 1) % 3]], &d = m.pos[V[(c + 2) % 3]];
            const Vector3f ab = b - a, ad = d - a;
            if (c == 0)
            {
                fn = cross(ab, ad);
                const double len = fn.length();
                if (len == 0) break;
                fn /= len;
            }
            m.nrm[V[c]] += fn * unit_angle(unit_vector(ab), unit_vector(ad));
        }
    }
    for (Vector3f &n : m.nrm)
    {
        const double len = n.length();
        if (len != 0) n /= len; else n = Vector3f(1, 0, 0);
    }
}

} // namespace

std::vector<std::shared_ptr<triangle_mesh>> mesh_loader::load_obj(std::string file, bool use_geometry_normals)
{
    std::ifstream in(file);
    if (!in.is_open()) throw std::runtime_error("mesh_loader: cannot open " + file);
    const size_t slash = file.find_last_of('/');
    const std::string dir = slash == std::string::npos ? std::string() : file.substr(0, slash + 1);

    std::vector<std::array<float, 3>> P, N;
    std::vector<std::array<float, 2>> T;
    std::map<std::string, mtl_record> library;
    std::vector<mesh_builder> builders;
    std::map<std::pair<std::string, std::string>, size_t> index_of;
    std::string group = "default", mtl, line, key;

    while (std::getline(in, line))
    {
        const size_t c = line.find('#');
        if (c != std::string::npos) line.erase(c);
        std::istringstream ls(line);
        if (!(ls >> key)) continue;
        if (key == "v") { std::array<float, 3> p{}; ls >> p[0] >> p[1] >> p[2]; P.push_back(p); }
        else if (key == "vn") { std::array<float, 3> p{}; ls >> p[0] >> p[1] >> p[2]; N.push_back(p); }
        else if (key == "vt") { std::array<float, 2> p{}; ls >> p[0] >> p[1]; T.push_back(p); }
        else if (key == "g" || key == "o") { group = "default"; ls >> group; }
        else if (key == "usemtl") { mtl.clear(); ls >> mtl; }
        else if (key == "mtllib") { std::string f; ls >> f; parse_mtl(dir + f, library); }
        else if (key == "f")
        {
            const auto mk = std::make_pair(group, mtl);
            auto found = index_of.find(mk);
            if (found == index_of.end())
            {
                found = index_of.emplace(mk, builders.size()).first;
                builders.emplace_back();
                builders.back().group = group;
                builders.back().mtl = mtl;
            }
            mesh_builder &m = builders[found->second];
            std::vector<int> poly;
            std::string tok;
            while (ls >> tok)
            {
                corner_key k{0, 0, 0};
                const char *s = tok.c_str();
                char *end = nullptr;
                k.v = (int)std::strtol(s, &end, 10);
                if (*end == '/')
                {
                    s = end + 1;
                    if (*s != '/') { k.t = (int)std::strtol(s, &end, 10); } else end = const_cast<char *>(s);
                    if (*end == '/') k.n = (int)std::strtol(end + 1, &end, 10);
                }
                k.v = resolve(k.v, P.size()); k.t = resolve(k.t, T.size()); k.n = resolve(k.n, N.size());
                if (k.v < 1 || k.v > (int)P.size()) throw std::runtime_error("mesh_loader: bad vertex index in " + file);
                auto it = m.seen.find(k);
                if (it == m.seen.end())
                {
                    it = m.seen.emplace(k, (int)m.pos.size()).first;
                    const auto &p = P[k.v - 1];
                    m.pos.emplace_back(p[0], p[1], p[2]);
                    if (k.t >= 1 && k.t <= (int)T.size()) m.tex.emplace_back(T[k.t - 1][0], T[k.t - 1][1]);
                    else m.tex.emplace_back(0.0, 0.0);
                    if (k.n >= 1 && k.n <= (int)N.size()) m.nrm.emplace_back(N[k.n - 1][0], N[k.n - 1][1], N[k.n - 1][2]);
                    else { m.nrm.emplace_back(0.0, 0.0, 0.0); m.all_normals = false; }
                }
                poly.push_back(it->second);
            }
            for (size_t i = 1; i + 1 < poly.size(); ++i)
            {
                m.idx.push_back(poly[0]);
                m.idx.push_back(poly[i]);
                m.idx.push_back(poly[i + 1]);
            }
        }
    }

    std::vector<std::shared_ptr<triangle_mesh>> meshes;
    for (mesh_builder &m : builders)
    {
        if (!m.all_normals) facet_normals(m);
        const mtl_record *rec = nullptr;
        if (!m.mtl.empty())
        {
            auto it = library.find(m.mtl);
            if (it != library.end())
            {
                rec = &it->second;
                std::printf("Loading %s\n", m.mtl.c_str()); // mesh_loader.cpp:66
            }
        }
        const int nv = (int)m.pos.size(), nt = (int)m.idx.size() / 3;
        Vector3f *v = new Vector3f[nv];
        Vector3f *n = new Vector3f[nv];
        Vector2f *t = new Vector2f[nv];
        std::copy(m.pos.begin(), m.pos.end(), v);
        std::copy(m.nrm.begin(), m.nrm.end(), n);
        std::copy(m.tex.begin(), m.tex.end(), t);
        // `name` carries the OBJ group so scene builders can address sub-meshes (veach_mis plates)
        meshes.push_back(std::make_shared<triangle_mesh>(nt, nv, v, m.idx.data(), n, t, make_material(rec),
                                                         use_geometry_normals, m.group, true));
    }
    return meshes;
}

namespace {
void append_triangles(const std::vector<std::shared_ptr<triangle_mesh>> &meshes,
                      std::vector<std::shared_ptr<hitable>> &out, std::vector<hitable *> &lights)
{
    for (const auto &mesh : meshes)
        for (int i = 0; i < mesh->nTriangles; ++i)
        {
            // The BVH leaf owns the triangle through its unique_ptr (parallel_bvh.h:134,148);
            // the shared_ptr handed back to the scene builder therefore must not delete it.
            out.push_back(std::shared_ptr<hitable>(new triangle(mesh, i), [](hitable *) {}));
            if (dynamic_cast<diffuse_light *>(mesh->mat.get())) lights.push_back(out.back().get());
        }
}
} // namespace

std::vector<std::shared_ptr<hitable>> create_triangle_mesh(const std::string &file, std::vector<hitable *> &lights,
                                                           bool use_geometry_normals)
{
    std::vector<std::shared_ptr<hitable>> tris;
    append_triangles(mesh_loader::load_obj(file, use_geometry_normals), tris, lights);
    return tris;
}

std::vector<std::shared_ptr<hitable>> create_triangle_mesh(const std::string &file, const Matrix4x4 &toWorld, material *bsdf,
                                                           std::vector<hitable *> &lights, bool use_geometry_normals)
{
    auto meshes = mesh_loader::load_obj(file, use_geometry_normals);
    Matrix4x4 inv;
    if (!toWorld.invert(inv)) throw std::runtime_error("Cannot compute inverse. Matrix isn't singular."); // triangle.cpp:34
    for (auto &mesh : meshes)
    {
        if (bsdf) mesh->mat.reset(bsdf);
        for (int j = 0; j < mesh->nVertices; ++j)
        {
            mesh->vertices[j] = toWorld * mesh->vertices[j];
            mesh->normals[j] = inv.normal_transform(mesh->normals[j]);
        }
    }
    std::vector<std::shared_ptr<hitable>> tris;
    append_triangles(meshes, tris, lights);
    return tris;
}
