// OBJ / MTL import for the host layer.  Same interface as the reference's mesh_loader
// (mesh_loader.h:18-22) and create_triangle_mesh factories (triangle.h:189-190, triangle.cpp:9-60)
// but self-contained: the reference delegates parsing to its vendored Assimp 5.0.1
// (mesh_loader.cpp:9: TargetRealtime_Fast without GenNormals, plus GenSmoothNormals), which is not
// part of this product.  What decides pixel values is restated here and checked against the
// reference's own loader (oracle/_ref/frt_ref_assimp, tests/test_host.py):
//   * numbers are read the way Assimp's fast_atoreal_move<float> reads them (fast_atof.h:232-338):
//     float(integer part) + float(double(up to 15 fraction digits) * 10^-k) — not the correctly
//     rounded float, so vertices can differ from strtof by one ulp
//   * one triangle_mesh per (group, material) in first-appearance order; quads are split from
//     their concave corner if they have one, else from corner 0 (TriangulateProcess.cpp:227-277);
//     larger polygons are fanned (Assimp ear-clips them: none in the shipped scenes)
//   * a file without normals gets Assimp's GenSmoothNormals (GenVertexNormalsProcess.cpp:115-236
//     with the default 175 degree limit): unweighted mean of the unit facet normals of every
//     corner within 1e-4 of the mesh's box diagonal, in float.  The angle-weighted normals of
//     mesh_loader.cpp:114-155 never run in the reference: Assimp has already made normals
//   * MTL -> material mapping of mesh_loader.cpp:59-112 (Ke -> diffuse_light; Ks & d<1 ->
//     dielectric; Ks -> modified_phong; else lambertian; colours through FromSrgb except Ke).
//     The reference reads opacity / ior / shininess into `double`s, which aiMaterial::Get refuses
//     (4-byte properties), so for Ks != 0 it branches on uninitialised stack; the mapping here is
//     the one the code spells out
//   * emissive triangles are appended to `lights` in creation order (triangle.cpp:15-20)
#include "mesh_loader.h"
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
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

// fast_atof.h:232-338 for Real = float (the sign, digits, fraction and exponent forms the shipped files use)
float assimp_atof(const char *c, const char **end = nullptr)
{
    static const double neg_pow10[16] = {0.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001, 0.00000001, 0.000000001,
                                         0.0000000001, 0.00000000001, 0.000000000001, 0.0000000000001, 0.00000000000001,
                                         0.000000000000001};
    float f = 0.f;
    const bool neg = *c == '-';
    if (neg || *c == '+') ++c;
    auto digits = [&c](unsigned int limit, unsigned int *count) {
        uint64_t v = 0;
        unsigned int n = 0;
        while (*c >= '0' && *c <= '9')
        {
            v = v * 10u + (uint64_t)(*c - '0');
            ++c;
            if (++n == limit) { while (*c >= '0' && *c <= '9') ++c; break; }
        }
        if (count) *count = n;
        return v;
    };
    if (*c != '.') f = (float)digits(0, nullptr);
    if (*c == '.' && c[1] >= '0' && c[1] <= '9')
    {
        ++c;
        unsigned int n = 0;
        double frac = (double)digits(15, &n);
        frac *= neg_pow10[n];
        f += (float)frac;
    }
    else if (*c == '.') ++c;
    if (*c == 'e' || *c == 'E')
    {
        ++c;
        const bool eneg = *c == '-';
        if (eneg || *c == '+') ++c;
        float e = (float)digits(0, nullptr);
        if (eneg) e = -e;
        f *= std::pow(10.0f, e);
    }
    if (end) *end = c;
    return neg ? -f : f;
}

// the next `n` blank-separated numbers of a line (missing ones stay as they are)
void read_floats(std::istringstream &ls, float *dst, int n)
{
    std::string tok;
    for (int i = 0; i < n && (ls >> tok); ++i)
    {
        const char *c = tok.c_str();
        const char *d = *c == '-' || *c == '+' ? c + 1 : c;
        if ((*d >= '0' && *d <= '9') || (*d == '.' && d[1] >= '0' && d[1] <= '9')) dst[i] = assimp_atof(c);
    }
}

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
        if (dst3) read_floats(ls, dst3, 3);
        else if (key == "Ns") read_floats(ls, &r.Ns, 1);
        else if (key == "Ni") read_floats(ls, &r.Ni, 1);
        else if (key == "d") read_floats(ls, &r.d, 1);
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
    // Assimp's view of the mesh for smooth_normals: one vertex per polygon corner in file order (`corner_vertex` = its
    // index into pos), and for every triangle corner the polygon corner it came from
    std::vector<int> corner_vertex, tri_corner;
    bool all_normals = true;
};

int resolve(int i, size_t count) { return i < 0 ? (int)count + i + 1 : i; }

struct f3
{
    float x, y, z;
};
inline f3 sub(const f3 &a, const f3 &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot3(const f3 &a, const f3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline f3 cross3(const f3 &a, const f3 &b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length3(const f3 &a) { return std::sqrt(dot3(a, a)); }
inline f3 scaled(const f3 &a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline f3 as_f3(const Vector3f &v) { return {(float)v[0], (float)v[1], (float)v[2]}; }

// TriangulateProcess.cpp:232-254: the corner a quad is fanned from — its concave corner (the two angles between the
// edges and the diagonal add up to more than pi, in float), else corner 0
int quad_fan_start(const f3 q[4])
{
    for (int i = 0; i < 4; ++i)
    {
        const f3 &v = q[i];
        f3 left = sub(q[(i + 3) % 4], v), diag = sub(q[(i + 2) % 4], v), right = sub(q[(i + 1) % 4], v);
        left = scaled(left, 1.0f / length3(left));
        diag = scaled(diag, 1.0f / length3(diag));
        right = scaled(right, 1.0f / length3(right));
        const float angle = std::acos(dot3(left, diag)) + std::acos(dot3(right, diag));
        if (angle > 3.1415926538f) return i;
    }
    return 0;
}

// GenVertexNormalsProcess.cpp:115-236 (no angle limit).  Assimp's mesh has one vertex per POLYGON corner; triangulation
// (which runs first) leaves them shared between a quad's two triangles.  Every triangle writes its unit facet normal to
// its three vertices, so a vertex ends up with the normal of the LAST triangle that uses it; then vertices closer than
// 1e-4 x the mesh's box diagonal (ProcessHelper.cpp:147-155) to the first vertex not yet visited form a group and
// share the normalised sum of their normals.  Float throughout, like aiVector3D; the sum itself is taken in double
// (Assimp adds in the order of its unstable spatial sort, so the last bit is not defined by the source).
void smooth_normals(mesh_builder &m)
{
    const int nt = (int)m.idx.size() / 3, nc = (int)m.corner_vertex.size();
    std::vector<f3> pos(nc);
    for (int c = 0; c < nc; ++c) pos[c] = as_f3(m.pos[m.corner_vertex[c]]);
    std::vector<f3> facet(nc, f3{0.f, 0.f, 0.f});
    for (int f = 0; f < nt; ++f)
    {
        const f3 &a = pos[m.tri_corner[3 * f]], &b = pos[m.tri_corner[3 * f + 1]], &c = pos[m.tri_corner[3 * f + 2]];
        f3 n = cross3(sub(b, a), sub(c, a));
        const float len = length3(n);
        if (len > 0.f) n = scaled(n, 1.0f / len);
        for (int k = 0; k < 3; ++k) facet[m.tri_corner[3 * f + k]] = n;
    }
    f3 lo{1e10f, 1e10f, 1e10f}, hi{-1e10f, -1e10f, -1e10f};
    for (const f3 &p : pos)
    {
        lo = {std::min(lo.x, p.x), std::min(lo.y, p.y), std::min(lo.z, p.z)};
        hi = {std::max(hi.x, p.x), std::max(hi.y, p.y), std::max(hi.z, p.z)};
    }
    const float eps = length3(sub(hi, lo)) * 1e-4f, eps2 = eps * eps;
    // vertices ordered along x: the candidates of a vertex are a window of this order
    std::vector<int> order(nc), rank(nc);
    for (int c = 0; c < nc; ++c) order[c] = c;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return pos[a].x < pos[b].x; });
    for (int r = 0; r < nc; ++r) rank[order[r]] = r;
    std::vector<char> had(nc, 0);
    std::vector<int> found;
    for (int c = 0; c < nc; ++c)
    {
        if (had[c]) continue;
        const f3 &p = pos[c];
        found.clear();
        int r0 = rank[c];
        while (r0 > 0 && p.x - pos[order[r0 - 1]].x <= eps) --r0;
        for (int r = r0; r < nc && pos[order[r]].x - p.x <= eps; ++r)
        {
            const f3 d = sub(pos[order[r]], p);
            if (dot3(d, d) < eps2) found.push_back(order[r]);
        }
        double sx = 0, sy = 0, sz = 0;
        for (int k : found) { sx += facet[k].x; sy += facet[k].y; sz += facet[k].z; }
        f3 n{(float)sx, (float)sy, (float)sz};
        const float len = length3(n);
        if (len > 0.f) n = scaled(n, 1.0f / len);
        for (int k : found) { m.nrm[m.corner_vertex[k]] = Vector3f(n.x, n.y, n.z); had[k] = 1; }
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
        if (key == "v") { std::array<float, 3> p{}; read_floats(ls, p.data(), 3); P.push_back(p); }
        else if (key == "vn") { std::array<float, 3> p{}; read_floats(ls, p.data(), 3); N.push_back(p); }
        else if (key == "vt") { std::array<float, 2> p{}; read_floats(ls, p.data(), 2); T.push_back(p); }
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
            const int corner0 = (int)m.corner_vertex.size();
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
                m.corner_vertex.push_back(it->second);
            }
            size_t first = 0;
            if (poly.size() == 4)
            {
                const f3 q[4] = {as_f3(m.pos[poly[0]]), as_f3(m.pos[poly[1]]), as_f3(m.pos[poly[2]]), as_f3(m.pos[poly[3]])};
                first = (size_t)quad_fan_start(q);
            }
            for (size_t i = 1; i + 1 < poly.size(); ++i)
            {
                m.idx.push_back(poly[first]);
                m.idx.push_back(poly[(first + i) % poly.size()]);
                m.idx.push_back(poly[(first + i + 1) % poly.size()]);
                m.tri_corner.push_back(corner0 + (int)first);
                m.tri_corner.push_back(corner0 + (int)((first + i) % poly.size()));
                m.tri_corner.push_back(corner0 + (int)((first + i + 1) % poly.size()));
            }
        }
    }

    std::vector<std::shared_ptr<triangle_mesh>> meshes;
    for (mesh_builder &m : builders)
    {
        if (!m.all_normals) smooth_normals(m);
        // no `usemtl` at all: Assimp's "DefaultMaterial" -> the placeholder of mesh_loader.cpp:105-109.  A name the
        // library does not hold: Assimp makes a default material (Kd 0.6) under that name (ObjFileParser.cpp:570-576)
        const mtl_record *rec = nullptr;
        static const mtl_record unlisted;
        if (!m.mtl.empty())
        {
            auto it = library.find(m.mtl);
            rec = it != library.end() ? &it->second : &unlisted;
            std::printf("Loading %s\n", m.mtl.c_str()); // mesh_loader.cpp:66
        }
        const int nv = (int)m.pos.size(), nt = (int)m.idx.size() / 3;
        Vector3f *v = new Vector3f[nv];
        Vector3f *n = new Vector3f[nv];
        Vector2f *t = new Vector2f[nv];
        std::copy(m.pos.begin(), m.pos.end(), v);
        std::copy(m.nrm.begin(), m.nrm.end(), n);
        std::copy(m.tex.begin(), m.tex.end(), t);
        // `name` is the MATERIAL's name, as in the reference (mesh_loader.cpp:62,158); "unknown" without one (:108).
        // In CornellBox-Original.obj the boxes' faces come before their `g` lines, so the group would name them wrongly
        meshes.push_back(std::make_shared<triangle_mesh>(nt, nv, v, m.idx.data(), n, t, make_material(rec),
                                                         use_geometry_normals, rec ? m.mtl : std::string("unknown"), true));
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
