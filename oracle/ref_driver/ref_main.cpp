// TEST INFRASTRUCTURE — driver around the UNMODIFIED reference hot path.
//
// oracle/Makefile compiles this file together with /root/reference/first_ray/{path,hitable,
// hitable_list,scene,pdf,material,texture,image}.cpp (from where they lie) into
// oracle/_ref/frt_ref (std::mt19937 sampler, as shipped) and oracle/_ref/frt_ref_ctr
// (-DFRT_REF_CTR_RNG, counter-based sampler from ref_prelude.h).  Everything that decides a
// pixel value or a hit — path::Li, parallel_bvh_node, triangle, sphere, material, pdf,
// microfacet, camera — is the reference's own object code.
//
// What is NOT the reference here (documented deviations):
//  * OBJ files go through the reference's own mesh_loader.cpp + its vendored Assimp 5.0.1; the one
//    override (materials with Ks != 0, where the reference reads uninitialised doubles) is
//    explained above read_obj.  triangle.cpp's create_triangle_mesh is restated in push_triangles
//    (it is six lines around the loader call).
//  * path::Render only covers rows 101..209 (path.cpp:120); the row loop below covers the
//    whole frame like ao.h:17 does, on the reference's own tf::Taskflow.
//  * veach_mis(): plates become rough_conductor (alphas from veach_mi/mi.xml), root becomes
//    parallel_bvh_node (main.cpp:313 alternative), as BASELINE config 3 asks.
//
// Commands (all print one JSON line on stdout):
//   frt_ref trace  <scene> <aspect> closest|shadow <rays.f32> <out.bin>
//   frt_ref render <scene> <w> <h> <spp> <out.pfm> [--threads N] [--seed S] [--rows y0 y1] [--ystep K]
//                  [--integrator path|ao|normals|direct]
//   frt_ref dump   <scene> <aspect> <out.bin>
//   frt_ref bsdf   <in.f64> <out.f64>
#include "Scene.h"
#include "sphere.h"
#include "triangle.h"
#include "material.h"
#include "camera.h"
#include "texture.h"
#include "util.h"
#include "bvh.h"
#include "parallel_bvh.h"
#include "pdf.h"
#include "sampler.h"
#include "integrator.h"
#include "path.h"
#include "ao.h"
#include "assignment_1.hpp"
#include "debug_renderer.h"

#include "mesh_loader.h"
#include <assimp/fast_atof.h>

#ifdef FRT_DROPIN
// The drop-in build: the product's integrator and flattener compiled against the REFERENCE's own headers
// (Scene.h, triangle.h, material.h, texture.h, integrator.h ...), the way a maintainer of the reference would add
// them: renderer<gpu_path_integrator> is the reference's renderer<> (integrator.h:11-47).
#define FRT_WITH_REFERENCE_HEADERS
#include "../../first_raytracer_b200/host/gpu_path_integrator.h"
#endif

static std::string g_assets = "assets";

// ------------------------------------------------------------------------------------------
// OBJ import: the reference's own mesh_loader::load_obj (mesh_loader.cpp:5-162) over its vendored Assimp 5.0.1
// (oracle/Makefile target `assimp`: plain g++ over the sources where they lie).
//
// One override.  For a material with Ks != 0, mesh_loader.cpp:81-96 reads AI_MATKEY_OPACITY / REFRACTI / SHININESS
// into `double`s; aiMaterial::Get<double> refuses a 4-byte property and leaves them uninitialised, so the reference
// picks dielectric or modified_phong (and its ior / exponent) from whatever the stack holds — undefined behaviour, not
// an algorithm.  Those materials are rebuilt here the way the code spells the mapping out (d < 1 -> dielectric(Ni),
// else modified_phong(Kd, Ks, Ns)), from the MTL numbers read with Assimp's own fast_atof.
// ------------------------------------------------------------------------------------------
struct mtl_entry
{
    float Kd[3] = {0.6f, 0.6f, 0.6f}, Ks[3] = {0, 0, 0}, Ke[3] = {0, 0, 0};
    float Ns = 0.f, Ni = 1.f, d = 1.f;
};

static void mtl_numbers(std::istringstream &ss, float *dst, int n)
{
    std::string tok;
    for (int i = 0; i < n && (ss >> tok); ++i) dst[i] = Assimp::fast_atof(tok.c_str());
}

static std::map<std::string, mtl_entry> read_mtl(const std::string &path)
{
    std::map<std::string, mtl_entry> out;
    std::ifstream is(path);
    std::string line, cur;
    while (std::getline(is, line))
    {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.resize(hash);
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key)) continue;
        if (key == "newmtl") { ss >> cur; out[cur] = mtl_entry(); continue; }
        if (cur.empty()) continue;
        mtl_entry &m = out[cur];
        if (key == "Kd") mtl_numbers(ss, m.Kd, 3);
        else if (key == "Ks") mtl_numbers(ss, m.Ks, 3);
        else if (key == "Ke") mtl_numbers(ss, m.Ke, 3);
        else if (key == "Ns") mtl_numbers(ss, &m.Ns, 1);
        else if (key == "Ni") mtl_numbers(ss, &m.Ni, 1);
        else if (key == "d") mtl_numbers(ss, &m.d, 1);
    }
    return out;
}

static std::vector<std::shared_ptr<triangle_mesh>> read_obj(const std::string &file, bool use_geometry_normals)
{
    std::vector<std::shared_ptr<triangle_mesh>> meshes = mesh_loader::load_obj(file, use_geometry_normals);
    std::ifstream is(file);
    const std::string dir = file.substr(0, file.find_last_of('/') + 1);
    std::map<std::string, mtl_entry> mtls;
    std::string line;
    while (std::getline(is, line))
    {
        std::istringstream ss(line);
        std::string key, f;
        if ((ss >> key) && key == "mtllib" && (ss >> f))
        {
            auto m = read_mtl(dir + f);
            mtls.insert(m.begin(), m.end());
        }
    }
    for (auto &mesh : meshes) // mesh->name is the material's name (mesh_loader.cpp:62,158)
    {
        auto f = mtls.find(mesh->name);
        if (f == mtls.end()) continue;
        const mtl_entry &m = f->second;
        const bool emissive = m.Ke[0] != 0 || m.Ke[1] != 0 || m.Ke[2] != 0;
        const bool specular = m.Ks[0] != 0 || m.Ks[1] != 0 || m.Ks[2] != 0;
        if (emissive || !specular) continue;
        const Vector3f ks(FromSrgb(m.Ks[0]), FromSrgb(m.Ks[1]), FromSrgb(m.Ks[2]));
        if (double(m.d) < 1.0)
            mesh->mat = std::make_unique<dielectric>(double(m.Ni), new constant_texture(ks), new constant_texture(Vector3f(1.0, 1.0, 1.0)));
        else
        {
            const Vector3f kd(FromSrgb(m.Kd[0]), FromSrgb(m.Kd[1]), FromSrgb(m.Kd[2]));
            mesh->mat = std::make_unique<modified_phong>(new constant_texture(kd), new constant_texture(ks), double(m.Ns));
        }
    }
    return meshes;
}

// triangle.cpp:9-26 without the Assimp call
static void push_triangles(const std::vector<std::shared_ptr<triangle_mesh>> &meshes,
                           std::vector<std::shared_ptr<hitable>> &tris, std::vector<hitable *> &lights)
{
    for (auto &mesh : meshes)
        for (int i = 0; i < mesh->nTriangles; ++i)
        {
            tris.push_back(std::make_shared<triangle>(mesh, i));
            if (dynamic_cast<diffuse_light *>(mesh->mat.get()))
                lights.push_back(tris.back().get());
        }
}

static std::shared_ptr<triangle_mesh> quad_mesh(const Vector3f &a, const Vector3f &b, const Vector3f &c,
                                                const Vector3f &d, std::unique_ptr<material> mat, const char *name)
{
    Vector3f *v = new Vector3f[4]{a, b, c, d};
    const Vector3f nn = unit_vector(cross(b - a, c - a));
    Vector3f *n = new Vector3f[4]{nn, nn, nn, nn};
    Vector2f *uv = new Vector2f[4]{Vector2f(0, 0), Vector2f(1, 0), Vector2f(1, 1), Vector2f(0, 1)};
    const int idx[6] = {0, 1, 2, 0, 2, 3};
    return std::make_shared<triangle_mesh>(2, 4, v, idx, n, uv, std::move(mat), true, name, true);
}

// ------------------------------------------------------------------------------------------
// Scenes
// ------------------------------------------------------------------------------------------
struct built_scene
{
    std::unique_ptr<Scene> scene;
    std::vector<std::shared_ptr<hitable>> keep;          // owners of triangles
    std::vector<hitable *> creation_order;                // primitive id = index here
    std::vector<std::shared_ptr<triangle_mesh>> meshes;
};

static built_scene &build_scene(const std::string &full_name, double aspect)
{
    // "<scene>+sky": same scene under a constant sky-blue environment (gives ao::Li something to return)
    const bool sky = full_name.size() > 4 && full_name.compare(full_name.size() - 4, 4, "+sky") == 0;
    // "<scene>+hdr": same scene under the HDR environment image of main.cpp:170 (the reference's own stb loader)
    const bool hdr = full_name.size() > 4 && full_name.compare(full_name.size() - 4, 4, "+hdr") == 0;
    const std::string name = (sky || hdr) ? full_name.substr(0, full_name.size() - 4) : full_name;
    built_scene &b = *new built_scene;
    // The HDR environment is loaded FIRST: image.cpp:19-21 switches stb_image's process-global vertical flip on when
    // any LDR image is loaded, and a later stbi_loadf would then flip the environment too (as shipped, main.cpp:170
    // loads the environment of a scene without LDR textures, i.e. unflipped).
    environment_map *hdr_env = hdr ? new environment_map(g_assets + "/data/test.hdr") : nullptr;
    std::vector<hitable *> lights;
    std::vector<hitable *> list;
    camera cam;
    Vector3f env(0, 0, 0);

    if (name == "cornell" || name == "cornell_glossy" || name == "cornell_sphere" || name == "cornell_metal" || name == "cornell_lens" ||
        name == "cornell_ao")
    {
        // main.cpp:222-252 (cornell_box_obj)
        // cornell_metal: the tall box becomes a `metal` mirror (material.h:111-130, only random_scene uses it as shipped)
        // cornell_lens : thin-lens camera, aperture 0.2 focused on the boxes (camera.h:10-35, util.h:21-41)
        const char *file = name == "cornell_glossy" ? "/CornellBox/CornellBox-Glossy.obj"
                         : name == "cornell_sphere" ? "/CornellBox/CornellBox-Sphere.obj"
                                                    : "/CornellBox/CornellBox-Original.obj";
        // cornell_ao  : cornell_box_ao, main.cpp:254-280 — the same file WITHOUT use_geometry_normals, so shading uses the
        //                vertex normals Assimp's GenSmoothNormals makes (box corners get rounded), white environment
        b.meshes = read_obj(g_assets + file, name != "cornell_ao");
        if (name == "cornell_ao") env = Vector3f(1.0, 1.0, 1.0);
        // CornellBox-Glossy.mtl has no `light` entry (the shipped scene is black): same emitter as Original
        for (auto &m : b.meshes)
            if (m->name == "light" && !dynamic_cast<diffuse_light *>(m->mat.get()))
                m->mat.reset(new diffuse_light(new constant_texture(Vector3f(17, 12, 4))));
        if (name == "cornell_metal")
            for (auto &m : b.meshes)
                if (m->name == "tallBox") m->mat.reset(new metal(Vector3f(0.8, 0.7, 0.6)));
        push_triangles(b.meshes, b.keep, lights);
        for (auto &t : b.keep) list.push_back(t.get());
        if (name == "cornell_lens")
            cam = camera(Vector3f(0, 1, 3.9f), Vector3f(0, 1, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.2, 4.0);
        else
            cam = camera(Vector3f(0, 1, 3.9f), Vector3f(0, 1, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.0, 10.0);
    }
    else if (name == "veach" || name == "veach_beckmann")
    {
        // veach_beckmann: the same plates with the Beckmann distribution (pdf.h:280-403, microfacet.h:47-128)
        const char *distribution = name == "veach" ? "ggx" : "beckmann";
        // main.cpp:281-314 (veach_mis) + SURVEY §8c(iii)
        // veach_mi.obj restarts its face indices per object, which (OBJ indices being global) puts
        // every plate on top of the floor triangles; the per-object files hold the intended geometry.
        const double alpha[4] = {0.005, 0.02, 0.05, 0.1}; // veach_mi/mi.xml:84,95,106,117
        for (int p = -1; p < 4; ++p)
        {
            const std::string file = p < 0 ? "/veach_mi/floor.obj" : "/veach_mi/plate" + std::to_string(p + 1) + ".obj";
            for (auto &m : read_obj(g_assets + file, false))
            {
                if (p < 0)
                    m->mat.reset(new lambertian(new constant_texture(FromSrgb(Vector3f(0.4, 0.4, 0.4)))));
                else
                    m->mat.reset(new rough_conductor(alpha[p], 1.0, Vector3f(1.65746, 0.880369, 0.521229),
                                                     Vector3f(9.22387, 6.26952, 4.837),
                                                     new constant_texture(Vector3f(1, 1, 1)), distribution));
                b.meshes.push_back(m);
            }
        }
        push_triangles(b.meshes, b.keep, lights);
        for (auto &t : b.keep) list.push_back(t.get());
        const Vector3f c[5] = {Vector3f(10, 10, 4), Vector3f(-1.25f, 0, 0), Vector3f(-3.75f, 0, 0),
                               Vector3f(1.25f, 0, 0), Vector3f(3.75f, 0, 0)};
        const double r[5] = {0.5, 0.1f, 0.03333f, 0.3f, 0.9f};
        const double e[5] = {800, 100, 901.803f, 11.1111f, 1.23457};
        for (int i = 0; i < 5; ++i)
        {
            list.push_back(new sphere(c[i], r[i], new diffuse_light(new constant_texture(Vector3f(e[i], e[i], e[i])))));
            lights.push_back(new sphere(c[i], r[i], new diffuse_light(new constant_texture(Vector3f(e[i], e[i], e[i])))));
        }
        cam = camera(Vector3f(0, 2, 15), Vector3f(0, -2, 2.5), Vector3f(0, 1, 0), 28.0, aspect, 0.0, 50.0);
    }
    else if (name == "teapot")
    {
        // stand-in for the unshipped Mitsuba matpreview mesh; camera from
        // furnace_test_scene main.cpp:207-212
        b.meshes = read_obj(g_assets + "/cube/teapot.obj", false);
        b.meshes.push_back(quad_mesh(Vector3f(-150, 0, -150), Vector3f(-150, 0, 150), Vector3f(150, 0, 150),
                                     Vector3f(150, 0, -150),
                                     std::make_unique<lambertian>(new constant_texture(Vector3f(0.4, 0.4, 0.4))), "floor"));
        b.meshes.push_back(quad_mesh(Vector3f(-40, 160, -40), Vector3f(40, 160, -40), Vector3f(40, 160, 40),
                                     Vector3f(-40, 160, 40),
                                     std::make_unique<diffuse_light>(new constant_texture(Vector3f(30, 30, 30))), "light"));
        push_triangles(b.meshes, b.keep, lights);
        for (auto &t : b.keep) list.push_back(t.get());
        cam = camera(Vector3f(0, 125, 175.0), Vector3f(0, 50, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.0, 300.0);
        env = Vector3f(0.05, 0.06, 0.08);
    }
    else if (name == "furnace")
    {
        // furnace_test_scene, main.cpp:177-221: cube/teapot.obj through create_triangle_mesh(file, toWorld = I, bsdf,
        // lights) (triangle.cpp:26-60, restated: every mesh gets the dielectric and its vertices / normals go through
        // the matrix), NO lights (pick_sample returns -1: path.cpp:40 skips NEE), camera main.cpp:207-212.  The
        // environment of the shipped scene is data/ennis.hdr, which the repository does not hold: "furnace+hdr" puts
        // the shipped data/test.hdr there (main.cpp:170 uses that one).
        b.meshes = read_obj(g_assets + "/cube/teapot.obj", false);
        Matrix4x4 I, I_inv;
        I.set_identity();
        if (!I.invert(I_inv)) throw std::runtime_error("Cannot compute inverse. Matrix isn't singular.");
        for (auto &m : b.meshes)
        {
            m->mat.reset(new dielectric(1.5, new constant_texture(Vector3f(1.0, 1.0, 1.0)), new constant_texture(Vector3f(1.0, 1.0, 1.0))));
            for (int j = 0; j < m->nVertices; ++j)
            {
                m->vertices[j] = I * m->vertices[j];
                m->normals[j] = I_inv.normal_transform(m->normals[j]);
            }
        }
        push_triangles(b.meshes, b.keep, lights);
        for (auto &t : b.keep) list.push_back(t.get());
        cam = camera(Vector3f(0, 125, 175.0), Vector3f(0, 50, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.0, 300.0);
    }
    else if (name == "textured")
    {
        // A reduced random_scene (main.cpp:55-113): the 1000-radius ground sphere under a checker_texture, the sphere
        // light, small spheres of every material random_scene draws (lambertian / metal / dielectric) on a fixed
        // grid, plus what exercises the texture classes on every primitive type: checker_texture (texture.h:30-49)
        // and image_texture (texture.h:51-112, LDR file through the reference's own stb loader) on spheres
        // (get_sphere_uv hitable.h:15-21) and on triangles (uv interpolation triangle.h:104-113).
        texture *checker = new checker_texture(new constant_texture(Vector3f(0.2f, 0.3f, 0.1f)),
                                               new constant_texture(Vector3f(0.99f, 0.99f, 0.99f)), 1, 1);
        auto tiles = [&]() { return new image_texture(std::make_unique<image>(g_assets + "/data/tiles.bmp", formats::STBI_BMP)); };
        list.push_back(new sphere(Vector3f(0, -1000, 0), 1000, new lambertian(checker)));
        list.push_back(new sphere(Vector3f(0, 1.5, 0), 1.5, new diffuse_light(new constant_texture(Vector3f(10, 10, 10)))));
        lights.push_back(new sphere(Vector3f(0, 1.5, 0), 1.5, new diffuse_light(new constant_texture(Vector3f(10, 10, 10)))));
        int k = 0;
        for (int a = -2; a <= 2; ++a)
            for (int c = -2; c <= 2; ++c, ++k)
            {
                if (a == 0 && c == 0) continue;
                // float32-representable centres: the C ABI carries spheres as float32 (frt_sphere)
                const Vector3f center((float)(1.6 * a + 0.3 * ((k * 7) % 5) / 5.0), 0.4f, (float)(1.6 * c + 0.3 * ((k * 3) % 5) / 5.0));
                material *m;
                switch (k % 6)
                {
                case 0: m = new lambertian(new constant_texture(Vector3f(0.1 + 0.15 * (k % 5), 0.6 - 0.1 * (k % 4), 0.3))); break;
                case 1: m = new metal(Vector3f(0.5 * (1 + 0.2 * (k % 5)), 0.5 * (1 + 0.25 * (k % 3)), 0.5 * (1 + 0.1 * (k % 7)))); break;
                case 2: m = new dielectric(1.5, new constant_texture(Vector3f(1.0, 1.0, 1.0)), new constant_texture(Vector3f(1.0, 1.0, 1.0))); break;
                case 3: m = new lambertian(new checker_texture(new constant_texture(Vector3f(0.8, 0.1, 0.1)),
                                                               new constant_texture(Vector3f(0.1, 0.1, 0.8)), 4, 16)); break;
                case 4: m = new lambertian(tiles()); break;
                default: m = new modified_phong(new checker_texture(tiles(), new constant_texture(Vector3f(0.05, 0.05, 0.05)), 2, 8),
                                                new constant_texture(Vector3f(0.3, 0.3, 0.3)), 40.0); break;
                }
                list.push_back(new sphere(center, 0.4f, m));
            }
        // textured triangles: a picture wall (image_texture) and a checker wall standing behind the spheres
        b.meshes.push_back(quad_mesh(Vector3f(5, 0, -4), Vector3f(5, 0, 4), Vector3f(5, 4, 4), Vector3f(5, 4, -4),
                                     std::make_unique<lambertian>(tiles()), "picture"));
        b.meshes.push_back(quad_mesh(Vector3f(-4, 0, -5), Vector3f(5, 0, -5), Vector3f(5, 4, -5), Vector3f(-4, 4, -5),
                                     std::make_unique<lambertian>(new checker_texture(new constant_texture(Vector3f(0.9, 0.8, 0.2)),
                                                                                     tiles(), 3, 2)), "checkerwall"));
        push_triangles(b.meshes, b.keep, lights);
        for (auto &t : b.keep) list.push_back(t.get());
        cam = camera(Vector3f(-9, 3, 5), Vector3f(0.5, 0.6, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.1f, 10.0);
        env = Vector3f(0.15, 0.18, 0.24);
    }
    else if (name == "random_scene")
    {
        // random_scene, main.cpp:55-113, as written: 22 x 22 cells, one 0.2-radius sphere per cell unless it would touch the
        // spot at (4, 0.2, 0); lambertian / metal / dielectric by a uniform draw; the checkered 1000-radius ground sphere,
        // the sphere light, the thin-lens camera (aperture 0.1).  The reference seeds its sampler with time(0)
        // (sampler.h:14), so every run of it places other spheres: this is the instance of std::mt19937(1), with the
        // draws taken one per statement (the reference leaves their order to the compiler: constructor arguments).
        // Centres are rounded to float32 (the C ABI carries spheres as float32).  Its environment, data/grace.hdr, is not
        // in the repository: "random_scene+hdr" uses the shipped data/test.hdr.

        std::mt19937 engine(1);
        auto get1d = [&engine]() { return std::uniform_real_distribution<>(0.0, 1.0)(engine); };
        texture *checker = new checker_texture(new constant_texture(Vector3f(0.2f, 0.3f, 0.1f)),
                                               new constant_texture(Vector3f(0.99f, 0.99f, 0.99f)), 1, 1);
        list.push_back(new sphere(Vector3f(0, -1000, 0), 1000, new lambertian(checker)));
        list.push_back(new sphere(Vector3f(0, 1.5, 0), 1.5, new diffuse_light(new constant_texture(Vector3f(10, 10, 10)))));
        lights.push_back(new sphere(Vector3f(0, 1.5, 0), 1.5, new diffuse_light(new constant_texture(Vector3f(10, 10, 10)))));
        for (int a = -11; a < 11; a++)
            for (int c = -11; c < 11; c++)
            {
                const double choose_mat = get1d();
                const double ux = get1d();
                const double uz = get1d();
                const Vector3f center((float)(a + 0.9f * ux), 0.2f, (float)(c + 0.9f * uz));
                if (!((center - Vector3f(4, 0.2f, 0)).length() > 0.9f)) continue;
                if (choose_mat < 0.5)
                {
                    const double r0 = get1d(), r1 = get1d(), g0 = get1d(), g1 = get1d(), b0 = get1d(), b1 = get1d();
                    list.push_back(new sphere(center, 0.2f, new lambertian(new constant_texture(Vector3f(r0 * r1, g0 * g1, b0 * b1)))));
                }
                else if (choose_mat < 0.75)
                {
                    const double r = get1d(), g = get1d(), bl = get1d();
                    list.push_back(new sphere(center, 0.2f, new metal(Vector3f(0.5 * (1 + r), 0.5 * (1 + g), 0.5 * (1 + bl)))));
                }
                else
                    list.push_back(new sphere(center, 0.2f, new dielectric(1.5, new constant_texture(Vector3f(1.0, 1.0, 1.0)),
                                                                           new constant_texture(Vector3f(1.0, 1.0, 1.0)))));
            }
        cam = camera(Vector3f(-15, 2, 1), Vector3f(0, 0, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.1f, 10.0);
    }
    else if (name.rfind("soupfile:", 0) == 0)
    {
        // raw float32 triangle soup: n x 9 floats (v0,v1,v2), one lambertian mesh
        std::ifstream is(name.substr(9), std::ios::binary);
        if (!is.is_open()) throw std::runtime_error("cannot open soup file");
        is.seekg(0, std::ios::end);
        const size_t bytes = is.tellg();
        is.seekg(0);
        const int nt = int(bytes / 36);
        std::vector<float> raw(size_t(nt) * 9);
        is.read((char *)raw.data(), bytes);
        Vector3f *v = new Vector3f[size_t(nt) * 3];
        Vector3f *n = new Vector3f[size_t(nt) * 3];
        Vector2f *uv = new Vector2f[size_t(nt) * 3];
        std::vector<int> idx(size_t(nt) * 3);
        for (size_t i = 0; i < size_t(nt) * 3; ++i)
        {
            v[i] = Vector3f(raw[3 * i], raw[3 * i + 1], raw[3 * i + 2]);
            n[i] = Vector3f(0, 0, 1);
            uv[i] = Vector2f(0, 0);
            idx[i] = int(i);
        }
        b.meshes.push_back(std::make_shared<triangle_mesh>(nt, nt * 3, v, idx.data(), n, uv,
                           std::make_unique<lambertian>(new constant_texture(Vector3f(0.5, 0.5, 0.5))), true, "soup", true));
        push_triangles(b.meshes, b.keep, lights);
        for (auto &t : b.keep) list.push_back(t.get());
        cam = camera(Vector3f(0.5, 0.5, 3.0), Vector3f(0.5, 0.5, 0.5), Vector3f(0, 1, 0), 30.0, aspect, 0.0, 10.0);
        env = Vector3f(1, 1, 1);
    }
    else
        throw std::runtime_error("unknown scene " + name);

    if (sky) env = Vector3f(0.5, 0.6, 0.8);
    b.creation_order = list;
    hitable **arr = new hitable *[list.size()];
    for (size_t i = 0; i < list.size(); ++i) arr[i] = list[i];
    b.scene.reset(new Scene(parallel_bvh_node::create_bvh(arr, (int)list.size(), 0.0, 0.0),
                            hdr ? hdr_env : new environment_map(std::make_unique<constant_texture>(env)), cam, lights));
    return b;
}

// ------------------------------------------------------------------------------------------
// Ray counting wrapper (closest: t_max == FLT_MAX, shadow: t_max < 2), SURVEY §8d
// ------------------------------------------------------------------------------------------
static std::atomic<uint64_t> g_closest{0}, g_shadow{0};
static thread_local uint64_t tl_closest = 0, tl_shadow = 0;
struct counting_world : public hitable
{
    counting_world(hitable *inner_) : inner(inner_) {}
    bool hit(const ray &r, double t_min, double t_max, hit_record &rec) const override
    {
        if (t_max < 2.0) ++tl_shadow; else ++tl_closest;
        return inner->hit(r, t_min, t_max, rec);
    }
    bool bounding_box(double t0, double t1, aabb &b) const override { return inner->bounding_box(t0, t1, b); }
    std::unique_ptr<hitable> inner;
};
static void flush_counts()
{
    g_closest += tl_closest; g_shadow += tl_shadow; tl_closest = tl_shadow = 0;
}

static double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ------------------------------------------------------------------------------------------
static int cmd_trace(int argc, char **argv)
{
    if (argc < 7) return 2;
    const std::string scene_name = argv[2];
    const double aspect = std::atof(argv[3]);
    const bool shadow = std::string(argv[4]) == "shadow";
    const double t0 = now_s();
    built_scene &b = build_scene(scene_name, aspect);
    const double t_build = now_s() - t0;
    std::map<const hitable *, int> id_of;
    for (size_t i = 0; i < b.creation_order.size(); ++i) id_of[b.creation_order[i]] = (int)i;

    std::ifstream is(argv[5], std::ios::binary);
    is.seekg(0, std::ios::end);
    const size_t n = size_t(is.tellg()) / 24;
    is.seekg(0);
    std::vector<float> rays(n * 6);
    is.read((char *)rays.data(), n * 24);

    std::vector<int32_t> prim(n);
    std::vector<double> tt(n), uu(n), vv(n);
    const double t1 = now_s();
    for (size_t i = 0; i < n; ++i)
    {
        const ray r(Vector3f(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]),
                    Vector3f(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]));
        hit_record rec;
        // path.cpp:10 / path.cpp:50
        const bool h = shadow ? b.scene->world->hit(r, EPSILON, 1 - SHADOW_EPSILON, rec)
                              : b.scene->world->hit(r, EPSILON, FLT_MAX, rec);
        prim[i] = h ? id_of[rec.obj] : -1;
        tt[i] = h ? rec.t : 0.0;
        uu[i] = h ? rec.uv.x : 0.0;
        vv[i] = h ? rec.uv.y : 0.0;
    }
    const double t_trace = now_s() - t1;
    std::ofstream os(argv[6], std::ios::binary | std::ios::trunc);
    os.write((char *)prim.data(), n * 4);
    os.write((char *)tt.data(), n * 8);
    os.write((char *)uu.data(), n * 8);
    os.write((char *)vv.data(), n * 8);
    std::printf("{\"cmd\": \"trace\", \"scene\": \"%s\", \"n_rays\": %zu, \"n_prims\": %zu, \"build_s\": %.6f, "
                "\"trace_s\": %.6f, \"threads\": 1, \"bvh_nodes\": %u}\n",
                scene_name.c_str(), n, b.creation_order.size(), t_build, t_trace, (unsigned)Scene::num_bvh_nodes);
    return 0;
}

static int cmd_render(int argc, char **argv)
{
    if (argc < 7) return 2;
    const std::string scene_name = argv[2];
    const int nx = std::atoi(argv[3]), ny = std::atoi(argv[4]);
    const uint_fast64_t ns = std::atoll(argv[5]);
    const std::string out = argv[6];
    int threads = (int)std::thread::hardware_concurrency();
    uint32_t seed = 1;
    int y0 = 0, y1 = ny, ystep = 1;
    std::string integrator_name = "path"; // path::Li, ao::Li (ao.cpp:4-28), normals_renderer::Li (debug_renderer.h:8-17), ass_1_renderer::Li (assignment_1.cpp:4-38)
    for (int i = 7; i < argc; ++i)
    {
        const std::string a = argv[i];
        if (a == "--threads") threads = std::atoi(argv[++i]);
        else if (a == "--seed") seed = (uint32_t)std::atoll(argv[++i]);
        else if (a == "--rows") { y0 = std::atoi(argv[i + 1]); y1 = std::atoi(argv[i + 2]); i += 2; }
        else if (a == "--ystep") ystep = std::atoi(argv[++i]);
        else if (a == "--integrator") integrator_name = argv[++i];
    }
    if (integrator_name != "path" && integrator_name != "ao" && integrator_name != "normals" && integrator_name != "direct")
    {
        std::fprintf(stderr, "unknown integrator %s\n", integrator_name.c_str());
        return 2;
    }
    const double tb = now_s();
    built_scene &b = build_scene(scene_name, double(nx) / double(ny));
    const double t_build = now_s() - tb;
    Scene *scene = b.scene.get();
    scene->world.reset(new counting_world(scene->world.release()));
    const bool has_lights = scene->lights.list_size > 0;
    (void)has_lights;
    viewer film(nx, ny, ns, 3);
    path integrator;
    ao ao_integrator;
    normals_renderer normals_integrator;
    ass_1_renderer direct_integrator;
    const int which = integrator_name == "ao" ? 1 : (integrator_name == "normals" ? 2 : (integrator_name == "direct" ? 3 : 0));
    std::atomic<uint64_t> nonfinite{0};

    tf::Taskflow tf(threads > 0 ? threads : 1);
    // path.cpp:120-147 with the full row range of ao.h:17
    tf.parallel_for(y1 - 1, y0 - 1, -ystep, [&, seed](int y)
    {
#ifdef FRT_REF_CTR_RNG
        sampler random_sampler;
#else
        sampler random_sampler(int(seed) * 7919 + y * 39);
#endif
        for (int x = 0; x < nx; x++)
        {
            Vector3f col(0.0, 0.0, 0.0);
            for (uint_fast64_t s = 0; s < ns; s++)
            {
#ifdef FRT_REF_CTR_RNG
                random_sampler.begin_sample(seed, uint32_t(y * nx + x), uint32_t(s), has_lights);
#endif
                double u = double(x + random_sampler.get1d()) / double(nx);
                double v = double(y + random_sampler.get1d()) / double(ny);
                ray r = scene->cam.get_ray(u, v, random_sampler.get2d());
                hit_record hrec;
                const Vector3f sample = which == 1 ? ao_integrator.Li(r, scene, 0, hrec, 0.0, random_sampler)
                                        : which == 2 ? normals_integrator.Li(r, scene, 0, hrec, 0.0, random_sampler)
                                        : which == 3 ? direct_integrator.Li(r, scene, 0, hrec, 0.0, random_sampler)
                                        : integrator.Li(r, scene, 0, hrec, 0.0, random_sampler);
                if (!(std::isfinite(sample[0]) && std::isfinite(sample[1]) && std::isfinite(sample[2])))
                {
                    ++nonfinite;
                    continue;
                }
                col += sample;
            }
            film.add_sample(Vector2i(x, y), col);
        }
        flush_counts();
    });
    const double t1 = now_s();
    tf.dispatch().get();
    const double secs = now_s() - t1;

    // image.h:89-118 (PFM, rows in buffer order = bottom row first), without the $HOME prefix
    {
        std::ofstream os(out, std::ios::binary | std::ios::trunc);
        os << "PF\n" << nx << " " << ny << "\n-1\n";
        std::vector<float> row(size_t(nx) * 3);
        for (int y = 0; y < ny; ++y)
        {
            for (int i = 0; i < nx * 3; ++i) row[i] = (float)film.fout_image[size_t(y) * nx * 3 + i];
            os.write((char *)row.data(), row.size() * 4);
        }
    }
    const int rows = (y1 - y0 + ystep - 1) / ystep;
    const double samples = double(rows) * nx * double(ns);
    const double rays = double(g_closest.load() + g_shadow.load());
    std::printf("{\"cmd\": \"render\", \"scene\": \"%s\", \"nx\": %d, \"ny\": %d, \"spp\": %llu, \"rows\": %d, "
                "\"threads\": %d, \"seconds\": %.6f, \"build_s\": %.6f, \"samples\": %.0f, \"samples_per_s\": %.3f, "
                "\"closest_rays\": %llu, \"shadow_rays\": %llu, \"mrays_per_s\": %.6f, \"nonfinite\": %llu, "
                "\"rng\": \"%s\", \"node_visits\": %u, \"tri_tests\": %u}\n",
                scene_name.c_str(), nx, ny, (unsigned long long)ns, rows, threads, secs, t_build, samples,
                samples / secs, (unsigned long long)g_closest.load(), (unsigned long long)g_shadow.load(),
                rays / secs / 1e6, (unsigned long long)nonfinite.load(),
#ifdef FRT_REF_CTR_RNG
                "philox",
#else
                "mt19937",
#endif
                (unsigned)Scene::num_rays_traversed, (unsigned)Scene::num_ray_tri_intersections);
    return 0;
}

// Scene dump: lets the tests check that the product's own loader / scene builders feed the GPU
// exactly the geometry and materials the reference driver renders.
static int cmd_dump(int argc, char **argv)
{
    if (argc < 5) return 2;
    built_scene &b = build_scene(argv[2], std::atof(argv[3]));
    std::ofstream os(argv[4], std::ios::binary | std::ios::trunc);
    auto wd = [&](double x) { os.write((char *)&x, 8); };
    auto wv = [&](const Vector3f &v) { wd(v[0]); wd(v[1]); wd(v[2]); };
    const camera &c = b.scene->cam;
    wd(3.0); // version tag
    wv(c.origin); wv(c.lower_left_corner); wv(c.horizontal); wv(c.vertical); wv(c.u); wv(c.v); wv(c.w); wd(c.lens_radius);
    wd((double)b.creation_order.size());
    wd((double)b.scene->lights.list_size);
    for (hitable *h : b.creation_order)
    {
        material *m = nullptr;
        if (auto *t = dynamic_cast<triangle *>(h))
        {
            wd(0.0);
            const Vector3f &v0 = t->mesh->vertices[t->V[0]];
            wv(v0); wv(v0 + t->edge1); wv(v0 + t->edge2);
            wv(t->mesh->normals[t->V[0]]); wv(t->mesh->normals[t->V[1]]); wv(t->mesh->normals[t->V[2]]);
            wd(t->inv_area); wd(t->mesh->use_geometry_normals ? 1.0 : 0.0);
            m = t->mat_ptr;
        }
        else if (auto *s = dynamic_cast<sphere *>(h))
        {
            wd(1.0);
            wv(s->center); wv(Vector3f(s->radius, 0, 0)); wv(Vector3f(0, 0, 0));
            wv(Vector3f(0, 0, 0)); wv(Vector3f(0, 0, 0)); wv(Vector3f(0, 0, 0));
            wd(0.0); wd(0.0);
            m = s->mat_ptr;
        }
        hit_record dummy;
        dummy.u = dummy.v = 0;
        double kind = -1, p[12] = {0};
        if (auto *l = dynamic_cast<lambertian *>(m)) { kind = 0; Vector3f a = l->albedo->value(dummy); p[0] = a[0]; p[1] = a[1]; p[2] = a[2]; }
        else if (auto *ph = dynamic_cast<modified_phong *>(m))
        {
            kind = 1; Vector3f a = ph->diffuse_reflectance->value(dummy), s = ph->specular_reflectance->value(dummy);
            p[0] = a[0]; p[1] = a[1]; p[2] = a[2]; p[3] = s[0]; p[4] = s[1]; p[5] = s[2]; p[6] = ph->specular_exponent;
        }
        else if (auto *d = dynamic_cast<dielectric *>(m))
        {
            kind = 2; Vector3f s = d->specular_reflectance->value(dummy);
            p[0] = s[0]; p[1] = s[1]; p[2] = s[2]; p[6] = d->ref_idx;
        }
        else if (auto *rc = dynamic_cast<rough_conductor *>(m))
        {
            kind = 3; Vector3f s = rc->specular_reflectance->value(dummy);
            p[0] = s[0]; p[1] = s[1]; p[2] = s[2]; p[3] = rc->eta[0]; p[4] = rc->eta[1]; p[5] = rc->eta[2];
            p[6] = rc->alphaU; p[7] = rc->k[0]; p[8] = rc->k[1]; p[9] = rc->k[2];
            p[10] = rc->distribution_type == microfacet_distributions::ggx ? 0 : 1;
        }
        else if (auto *mt = dynamic_cast<metal *>(m)) { kind = 4; p[0] = mt->albedo[0]; p[1] = mt->albedo[1]; p[2] = mt->albedo[2]; }
        else if (auto *dl = dynamic_cast<diffuse_light *>(m)) { kind = 5; Vector3f e = dl->emit->value(dummy); p[0] = e[0]; p[1] = e[1]; p[2] = e[2]; }
        wd(kind);
        for (double x : p) wd(x);
    }
    std::printf("{\"cmd\": \"dump\", \"n_prims\": %zu, \"n_lights\": %d, \"bvh_nodes\": %u, \"bvh_leaves\": %u}\n",
                b.creation_order.size(), b.scene->lights.list_size, (unsigned)Scene::num_bvh_nodes,
                (unsigned)Scene::num_bvh_leaf_nodes);
    return 0;
}

// BSDF unit vectors: each input record is 24 doubles
//   kind, p[0..10], n[3], wi[3], s[3], wo_probe[3]
// and each output record 12 doubles (layout of frt_oracle_bsdf, oracle/frt_oracle.h), computed by the
// reference's own material::scatter / pdf::generate / pdf::value / eval_bsdf.
static int cmd_bsdf(int argc, char **argv)
{
    if (argc < 4) return 2;
    std::ifstream is(argv[2], std::ios::binary);
    is.seekg(0, std::ios::end);
    const size_t n = size_t(is.tellg()) / (24 * 8);
    is.seekg(0);
    std::vector<double> in(n * 24), out(n * 12, 0.0);
    is.read((char *)in.data(), n * 24 * 8);
    for (size_t i = 0; i < n; ++i)
    {
        const double *q = &in[i * 24];
        const int kind = int(q[0]);
        const double *p = q + 1;
        std::unique_ptr<material> m;
        if (kind == 0) m.reset(new lambertian(new constant_texture(Vector3f(p[0], p[1], p[2]))));
        else if (kind == 1) m.reset(new modified_phong(new constant_texture(Vector3f(p[0], p[1], p[2])), new constant_texture(Vector3f(p[3], p[4], p[5])), p[6]));
        else if (kind == 2) m.reset(new dielectric(p[6], new constant_texture(Vector3f(p[0], p[1], p[2])), new constant_texture(Vector3f(1, 1, 1))));
        else if (kind == 3) m.reset(new rough_conductor(p[6], 1.0, Vector3f(p[3], p[4], p[5]), Vector3f(p[7], p[8], p[9]), new constant_texture(Vector3f(p[0], p[1], p[2])), p[10] == 0 ? "ggx" : "beckmann"));
        else if (kind == 4) m.reset(new metal(Vector3f(p[0], p[1], p[2])));
        else continue;
        hit_record hrec;
        hrec.t = 1.0; hrec.u = hrec.v = 0.0;
        hrec.p = Vector3f(0, 0, 0);
        hrec.normal = Vector3f(q[12], q[13], q[14]);
        hrec.wi = Vector3f(q[15], q[16], q[17]);
        hrec.mat_ptr = m.get();
        const ray r_in(hrec.wi, -hrec.wi);
        scatter_record srec(hrec);
        if (!m->scatter(r_in, hrec, srec, Vector3f(q[18], q[19], q[20]))) continue;
        Vector3f dir = srec.is_specular ? srec.specular_ray.direction()
                                        : srec.pdf_ptr->generate(Vector2f(q[18], q[19]), srec);
        const Vector3f probe(q[21], q[22], q[23]);
        double *o = &out[i * 12];
        o[0] = dir[0]; o[1] = dir[1]; o[2] = dir[2];
        o[3] = srec.sampled_pdf;
        o[4] = srec.pdf_ptr->value(hrec, dir);
        const Vector3f f = m->eval_bsdf(r_in, hrec, dir), fp = m->eval_bsdf(r_in, hrec, probe);
        o[5] = f[0]; o[6] = f[1]; o[7] = f[2];
        o[8] = srec.pdf_ptr->value(hrec, probe);
        o[9] = fp[0]; o[10] = fp[1]; o[11] = fp[2];
    }
    std::ofstream os(argv[3], std::ios::binary | std::ios::trunc);
    os.write((char *)out.data(), out.size() * 8);
    std::printf("{\"cmd\": \"bsdf\", \"n\": %zu}\n", n);
    return 0;
}

#ifdef FRT_DROPIN
// frt_dropin flatten <scene> <aspect> <max_leaf> <out.bin>: the frt_scene_desc arrays the flattener produces when it
// walks the REFERENCE's Scene (real parallel_bvh_node / triangle / material / texture objects), one after the other:
// header of 16 uint32 (counts), then nodes, tris, tri_normals, spheres, materials, lights, camera, env, textures,
// tex_texels, tri_uv.  tests/test_host.py compares them byte for byte with the host layer's own.
static int cmd_flatten(int argc, char **argv)
{
    if (argc < 6) return 2;
    built_scene &b = build_scene(argv[2], std::atof(argv[3]));
    std::unordered_map<const hitable *, int> ids;
    for (size_t i = 0; i < b.creation_order.size(); ++i) ids[b.creation_order[i]] = (int)i;
    frt_flattener fl(&ids, std::atoi(argv[4]));
    const frt_flat_scene flat = fl.flatten(b.scene.get());
    const frt_scene_desc d = flat.desc();
    std::ofstream os(argv[5], std::ios::binary | std::ios::trunc);
    const uint32_t head[16] = {d.n_nodes, d.n_tris, d.n_spheres, d.n_materials, d.n_lights, d.tri_normals ? 1u : 0u, d.env_image ? 1u : 0u,
                               d.env_nx, d.env_ny, d.n_textures, (uint32_t)d.n_texels, d.tri_uv ? 1u : 0u, 0, 0, 0, 0};
    os.write((const char *)head, sizeof head);
    os.write((const char *)d.nodes, sizeof(frt_bvh4_node) * (size_t)d.n_nodes);
    os.write((const char *)d.tris, sizeof(frt_float4) * 3 * (size_t)d.n_tris);
    if (d.tri_normals) os.write((const char *)d.tri_normals, sizeof(frt_float4) * 3 * (size_t)d.n_tris);
    os.write((const char *)d.spheres, sizeof(frt_sphere) * (size_t)d.n_spheres);
    os.write((const char *)d.materials, sizeof(frt_material) * (size_t)d.n_materials);
    os.write((const char *)d.lights, sizeof(frt_light) * (size_t)d.n_lights);
    os.write((const char *)&d.camera, sizeof d.camera);
    os.write((const char *)d.env_color, sizeof d.env_color);
    os.write((const char *)&d.ao_t_max, sizeof d.ao_t_max);
    os.write((const char *)d.textures, sizeof(frt_texture) * (size_t)d.n_textures);
    os.write((const char *)d.tex_texels, sizeof(float) * 3 * (size_t)d.n_texels);
    if (d.tri_uv) os.write((const char *)d.tri_uv, sizeof(float) * 6 * (size_t)d.n_tris);
    std::printf("{\"cmd\": \"flatten\", \"n_nodes\": %u, \"n_tris\": %u, \"n_spheres\": %u, \"n_textures\": %u}\n", d.n_nodes, d.n_tris,
                d.n_spheres, d.n_textures);
    return 0;
}

// frt_dropin gpurender <scene> <w> <h> <spp> <out.pfm> [--seed S] [--gpus N] [--integrator path|ao|normals|direct]:
// the reference's renderer<> (integrator.h:11-47) instantiated with gpu_path_integrator on the reference's Scene.
static int cmd_gpurender(int argc, char **argv)
{
    if (argc < 7) return 2;
    const int nx = std::atoi(argv[3]), ny = std::atoi(argv[4]);
    const uint_fast64_t ns = std::atoll(argv[5]);
    uint32_t seed = 1, integ = FRT_INTEGRATOR_PATH;
    int gpus = 1;
    for (int i = 7; i < argc; ++i)
    {
        const std::string a = argv[i];
        if (a == "--seed") seed = (uint32_t)std::atoll(argv[++i]);
        else if (a == "--gpus") gpus = std::atoi(argv[++i]);
        else if (a == "--integrator")
        {
            const std::string v = argv[++i];
            integ = v == "ao" ? FRT_INTEGRATOR_AO : v == "normals" ? FRT_INTEGRATOR_NORMALS : v == "direct" ? FRT_INTEGRATOR_DIRECT : FRT_INTEGRATOR_PATH;
        }
    }
    built_scene &b = build_scene(argv[2], double(nx) / double(ny));
    std::unordered_map<const hitable *, int> ids;
    for (size_t i = 0; i < b.creation_order.size(); ++i) ids[b.creation_order[i]] = (int)i;
    viewer film(nx, ny, ns, 3);
    renderer<gpu_path_integrator> render;
    render.seed = seed; render.n_gpus = gpus; render.integrator = integ; render.prim_ids = &ids;
    std::streambuf *keep = std::cout.rdbuf(std::cerr.rdbuf()); // the reference's timing block goes to stderr: stdout is one JSON line
    const double secs = render.Render(b.scene.get(), film);
    std::cout.rdbuf(keep);
    {
        std::ofstream os(argv[6], std::ios::binary | std::ios::trunc);
        os << "PF\n" << nx << " " << ny << "\n-1\n";
        std::vector<float> row(size_t(nx) * 3);
        for (int y = 0; y < ny; ++y)
        {
            for (int i = 0; i < nx * 3; ++i) row[i] = (float)film.fout_image[size_t(y) * nx * 3 + i];
            os.write((char *)row.data(), row.size() * 4);
        }
    }
    unsigned long long ldr_sum = 0;
    for (size_t i = 0; i < size_t(nx) * ny * 3; ++i) ldr_sum += film.out_image[i];
    std::printf("{\"cmd\": \"gpurender\", \"seconds\": %.6f, \"samples\": %llu, \"closest_rays\": %llu, \"shadow_rays\": %llu, "
                "\"ldr_sum\": %llu}\n", secs, (unsigned long long)render.stats.samples, (unsigned long long)render.stats.closest_rays,
                (unsigned long long)render.stats.shadow_rays, ldr_sum);
    gpu_path_integrator::release_cached_context();
    return 0;
}
#endif

int main(int argc, char **argv)
{
    if (const char *a = std::getenv("FRT_ASSETS")) g_assets = a;
    if (argc < 2) { std::fprintf(stderr, "usage: frt_ref trace|render|dump ...\n"); return 2; }
    const std::string cmd = argv[1];
    try
    {
        // The reference double-owns its primitives (parallel_bvh.h:134,148 vs triangle.cpp:17), so
        // nothing is destroyed: leave through _Exit once the output is flushed.
        int rc = 2;
        if (cmd == "trace") rc = cmd_trace(argc, argv);
        else if (cmd == "render") rc = cmd_render(argc, argv);
        else if (cmd == "dump") rc = cmd_dump(argc, argv);
        else if (cmd == "bsdf") rc = cmd_bsdf(argc, argv);
#ifdef FRT_DROPIN
        else if (cmd == "flatten") rc = cmd_flatten(argc, argv);
        else if (cmd == "gpurender") rc = cmd_gpurender(argc, argv);
#endif
        std::fflush(stdout);
        std::_Exit(rc);
    }
    catch (const std::exception &e)
    {
        std::fprintf(stderr, "frt_ref: %s\n", e.what());
        return 1;
    }
    return 2;
}
