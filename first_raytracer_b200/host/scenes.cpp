// Scene builders of the north-star configurations, written against the reference's own API
// the way main.cpp writes them (cornell_box_obj main.cpp:222-252, veach_mis main.cpp:281-314,
// furnace_test_scene's teapot main.cpp:198-212).  Deviations are the ones BASELINE/SURVEY §8c
// fix: veach plates are rough_conductor with the alphas of veach_mi/mi.xml and the root is a
// parallel_bvh_node; the unshipped Mitsuba matpreview mesh is replaced by cube/teapot.obj.
#include "scenes.h"
#include <random>
#include <cstdlib>
#include <stdexcept>
#include "mesh_loader.h"
#include "parallel_bvh.h"

std::atomic<unsigned int> Scene::num_rays_traversed{0};
std::atomic<unsigned int> Scene::num_ray_tri_intersections{0};
std::atomic<unsigned int> Scene::num_ray_box_intersections{0};
std::atomic<unsigned int> Scene::num_bvh_nodes{0};
std::atomic<unsigned int> Scene::num_bvh_leaf_nodes{0};

namespace {

std::string asset(const std::string &rel)
{
    const char *root = std::getenv("FRT_ASSETS");
    return std::string(root ? root : "assets") + "/" + rel;
}

// Remember creation order before create_bvh reorders the array (parallel_bvh.h:93-98).
Scene *finish(std::vector<hitable *> &list, const camera &cam, const Vector3f &env, const std::vector<hitable *> &lights,
              frt_scene_ids *ids)
{
    if (ids)
    {
        ids->id.clear();
        for (size_t i = 0; i < list.size(); ++i) ids->id[list[i]] = (int)i;
    }
    return new Scene(parallel_bvh_node::create_bvh(list.data(), (int)list.size(), 0.0, 0.0),
                     new environment_map(std::make_unique<constant_texture>(env)), cam, lights);
}

std::shared_ptr<triangle_mesh> quad(const Vector3f &a, const Vector3f &b, const Vector3f &c, const Vector3f &d,
                                    std::unique_ptr<material> mat, const char *name)
{
    Vector3f *v = new Vector3f[4]{a, b, c, d};
    const Vector3f nn = unit_vector(cross(b - a, c - a));
    Vector3f *n = new Vector3f[4]{nn, nn, nn, nn};
    Vector2f *uv = new Vector2f[4]{Vector2f(0, 0), Vector2f(1, 0), Vector2f(1, 1), Vector2f(0, 1)};
    const int idx[6] = {0, 1, 2, 0, 2, 3};
    return std::make_shared<triangle_mesh>(2, 4, v, idx, n, uv, std::move(mat), true, name, true);
}

void add_mesh(const std::shared_ptr<triangle_mesh> &mesh, std::vector<hitable *> &list, std::vector<hitable *> &lights)
{
    for (int i = 0; i < mesh->nTriangles; ++i)
    {
        list.push_back(new triangle(mesh, i));
        if (dynamic_cast<diffuse_light *>(mesh->mat.get())) lights.push_back(list.back());
    }
}

} // namespace

Scene *cornell_box_obj(const double &aspect, const char *variant, frt_scene_ids *ids)
{
    std::vector<hitable *> list, lights;
    const bool own_file = variant && (std::string(variant) == "Glossy" || std::string(variant) == "Sphere");
    const std::string file = std::string("CornellBox/CornellBox-") + (own_file ? variant : "Original") + ".obj";
    // Use geometry normals (main.cpp:229); "Ao" = cornell_box_ao (main.cpp:254-280): the same file with its (generated,
    // smooth) vertex normals and a white environment
    const std::string var = variant ? variant : "Original";
    auto meshes = mesh_loader::load_obj(asset(file), var != "Ao");
    // CornellBox-Glossy.mtl as shipped has no `light` entry, so that scene renders black; give the
    // `light` group the emitter of CornellBox-Original.mtl:81-88 (the oracle driver does the same).
    for (auto &m : meshes)
        if (m->name == "light" && !dynamic_cast<diffuse_light *>(m->mat.get()))
            m->mat.reset(new diffuse_light(new constant_texture(Vector3f(17, 12, 4))));
    // "Metal": CornellBox-Original with the tall box as a `metal` mirror (material.h:111-130; as shipped only
    // random_scene uses that class); "Lens": CornellBox-Original through a thin lens (camera.h:10-35)
    if (var == "Metal")
        for (auto &m : meshes)
            if (m->name == "tallBox") m->mat.reset(new metal(Vector3f(0.8, 0.7, 0.6)));
    for (auto &m : meshes) add_mesh(m, list, lights);
    camera cam = var == "Lens" ? camera(Vector3f(0, 1, 3.9f), Vector3f(0, 1, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.2, 4.0)
                               : camera(Vector3f(0, 1, 3.9f), Vector3f(0, 1, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.0, 10.0);
    return finish(list, cam, var == "Ao" ? Vector3f(1.0, 1.0, 1.0) : Vector3f(0.0, 0.0, 0.0), lights, ids);
}

Scene *veach_mis(const double &aspect, frt_scene_ids *ids, const char *distribution)
{
    std::vector<hitable *> list, lights;
    // veach_mi/veach_mi.obj numbers each object's faces from 1 again; OBJ indices are global, so as
    // shipped every plate collapses onto the floor's first two triangles (coincident geometry whose
    // winner depends on traversal order).  The per-object files carry the intended geometry
    // (SURVEY §8c(iii)); floor colour = Kd of `floor` in veach_mi/veach_mi.mtl.
    std::vector<std::shared_ptr<triangle_mesh>> meshes;
    const double alpha[4] = {0.005, 0.02, 0.05, 0.1}; // veach_mi/mi.xml:84,95,106,117
    for (int p = -1; p < 4; ++p)
    {
        const std::string file = p < 0 ? "veach_mi/floor.obj" : "veach_mi/plate" + std::to_string(p + 1) + ".obj";
        for (auto &m : mesh_loader::load_obj(asset(file), false))
        {
            if (p < 0)
                m->mat.reset(new lambertian(new constant_texture(FromSrgb(Vector3f(0.4, 0.4, 0.4)))));
            else
                m->mat.reset(new rough_conductor(alpha[p], 1.0, Vector3f(1.65746, 0.880369, 0.521229),
                                                 Vector3f(9.22387, 6.26952, 4.837), new constant_texture(Vector3f(1, 1, 1)), distribution));
            m->name = p < 0 ? "floor" : "plate" + std::to_string(p + 1);
            meshes.push_back(m);
        }
    }
    for (auto &m : meshes) add_mesh(m, list, lights);
    // main.cpp:292-302: the world and the light list hold separate sphere objects
    const Vector3f c[5] = {Vector3f(10, 10, 4), Vector3f(-1.25f, 0, 0), Vector3f(-3.75f, 0, 0), Vector3f(1.25f, 0, 0), Vector3f(3.75f, 0, 0)};
    const double r[5] = {0.5, 0.1f, 0.03333f, 0.3f, 0.9f};
    const double e[5] = {800, 100, 901.803f, 11.1111f, 1.23457};
    for (int i = 0; i < 5; ++i)
    {
        list.push_back(new sphere(c[i], r[i], new diffuse_light(new constant_texture(Vector3f(e[i], e[i], e[i])))));
        lights.push_back(new sphere(c[i], r[i], new diffuse_light(new constant_texture(Vector3f(e[i], e[i], e[i])))));
    }
    camera cam(Vector3f(0, 2, 15), Vector3f(0, -2, 2.5), Vector3f(0, 1, 0), 28.0, aspect, 0.0, 50.0);
    return finish(list, cam, Vector3f(0.0, 0.0, 0.0), lights, ids);
}

Scene *teapot_scene(const double &aspect, frt_scene_ids *ids)
{
    std::vector<hitable *> list, lights;
    auto meshes = mesh_loader::load_obj(asset("cube/teapot.obj"), false);
    meshes.push_back(quad(Vector3f(-150, 0, -150), Vector3f(-150, 0, 150), Vector3f(150, 0, 150), Vector3f(150, 0, -150),
                          std::make_unique<lambertian>(new constant_texture(Vector3f(0.4, 0.4, 0.4))), "floor"));
    meshes.push_back(quad(Vector3f(-40, 160, -40), Vector3f(40, 160, -40), Vector3f(40, 160, 40), Vector3f(-40, 160, 40),
                          std::make_unique<diffuse_light>(new constant_texture(Vector3f(30, 30, 30))), "light"));
    for (auto &m : meshes) add_mesh(m, list, lights);
    camera cam(Vector3f(0, 125, 175.0), Vector3f(0, 50, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.0, 300.0);
    return finish(list, cam, Vector3f(0.05, 0.06, 0.08), lights, ids);
}

// furnace_test_scene (main.cpp:177-221): the teapot through create_triangle_mesh(file, toWorld = identity, bsdf, lights)
// with a dielectric(1.5) on every mesh, NO lights (pick_sample gives -1 and path.cpp:40 skips next-event estimation),
// camera main.cpp:207-212.  Its environment is data/ennis.hdr, which the reference repository does not ship:
// "furnace+hdr" uses the shipped data/test.hdr (the file main.cpp:170 loads).
Scene *furnace_test_scene(const double &aspect, frt_scene_ids *ids)
{
    std::vector<hitable *> list, lights;
    Matrix4x4 I;
    I.set_identity();
    auto meshes = mesh_loader::load_obj(asset("cube/teapot.obj"), false);
    Matrix4x4 inv;
    if (!I.invert(inv)) throw std::runtime_error("Cannot compute inverse. Matrix isn't singular.");
    for (auto &m : meshes)
    {
        m->mat.reset(new dielectric(1.5, new constant_texture(Vector3f(1.0, 1.0, 1.0)), new constant_texture(Vector3f(1.0, 1.0, 1.0))));
        for (int j = 0; j < m->nVertices; ++j)
        {
            m->vertices[j] = I * m->vertices[j];
            m->normals[j] = inv.normal_transform(m->normals[j]);
        }
        add_mesh(m, list, lights);
    }
    camera cam(Vector3f(0, 125, 175.0), Vector3f(0, 50, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.0, 300.0);
    return finish(list, cam, Vector3f(0.0, 0.0, 0.0), lights, ids);
}

// A reduced random_scene (main.cpp:55-113): the 1000-radius ground sphere under a checker_texture, the sphere light,
// small spheres of the three materials random_scene draws (lambertian / metal / dielectric) on a fixed grid, plus
// checker_texture (texture.h:30-49) and image_texture (texture.h:51-112) on spheres (get_sphere_uv hitable.h:15-21)
// and on triangles (uv interpolation triangle.h:104-113).  The oracle driver builds the same scene from the
// reference's own classes (oracle/ref_driver/ref_main.cpp "textured").
Scene *textured_scene(const double &aspect, frt_scene_ids *ids)
{
    std::vector<hitable *> list, lights;
    texture *checker = new checker_texture(new constant_texture(Vector3f(0.2f, 0.3f, 0.1f)),
                                           new constant_texture(Vector3f(0.99f, 0.99f, 0.99f)), 1, 1);
    auto tiles = [&]() { return new image_texture(std::make_unique<image>(asset("data/tiles.bmp"), formats::STBI_BMP)); };
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
    std::vector<std::shared_ptr<triangle_mesh>> meshes;
    meshes.push_back(quad(Vector3f(5, 0, -4), Vector3f(5, 0, 4), Vector3f(5, 4, 4), Vector3f(5, 4, -4),
                          std::make_unique<lambertian>(tiles()), "picture"));
    meshes.push_back(quad(Vector3f(-4, 0, -5), Vector3f(5, 0, -5), Vector3f(5, 4, -5), Vector3f(-4, 4, -5),
                          std::make_unique<lambertian>(new checker_texture(new constant_texture(Vector3f(0.9, 0.8, 0.2)), tiles(), 3, 2)),
                          "checkerwall"));
    for (auto &m : meshes) add_mesh(m, list, lights);
    camera cam(Vector3f(-9, 3, 5), Vector3f(0.5, 0.6, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.1f, 10.0);
    return finish(list, cam, Vector3f(0.15, 0.18, 0.24), lights, ids);
}

// random_scene, main.cpp:55-113, as written: 22 x 22 cells, one 0.2-radius sphere per cell unless it would touch the
// spot at (4, 0.2, 0); lambertian / metal / dielectric by a uniform draw; the checkered 1000-radius ground sphere,
// the sphere light, the thin-lens camera (aperture 0.1).  The reference seeds its sampler with time(0)
// (sampler.h:14), so every run of it places other spheres: this is the instance of std::mt19937(1), with the
// draws taken one per statement (the reference leaves their order to the compiler: constructor arguments).
// Centres are rounded to float32 (the C ABI carries spheres as float32).  Its environment, data/grace.hdr, is not
// in the repository: "random_scene+hdr" uses the shipped data/test.hdr.
Scene *random_scene(const double &aspect, frt_scene_ids *ids)
{
    std::vector<hitable *> list, lights;
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
    camera cam(Vector3f(-15, 2, 1), Vector3f(0, 0, 0), Vector3f(0, 1, 0), 40.0, aspect, 0.1f, 10.0);
    return finish(list, cam, Vector3f(0.0, 0.0, 0.0), lights, ids);
}

// splitmix64 stream, centre ~ U[0,1)^3, two edges ~ U[-l,l]^3
static inline uint64_t splitmix64(uint64_t &s)
{
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline float unit_float(uint64_t &s) { return (float)(splitmix64(s) >> 40) * (1.0f / 16777216.0f); }

void soup_triangles(size_t n, uint64_t seed, float edge, std::vector<float> &out)
{
    out.resize(n * 9);
    uint64_t s = seed;
    for (size_t i = 0; i < n; ++i)
    {
        float c[3], e1[3], e2[3];
        for (int k = 0; k < 3; ++k) c[k] = unit_float(s);
        for (int k = 0; k < 3; ++k) e1[k] = (2.0f * unit_float(s) - 1.0f) * edge;
        for (int k = 0; k < 3; ++k) e2[k] = (2.0f * unit_float(s) - 1.0f) * edge;
        for (int k = 0; k < 3; ++k)
        {
            out[9 * i + k] = c[k];
            out[9 * i + 3 + k] = c[k] + e1[k];
            out[9 * i + 6 + k] = c[k] + e2[k];
        }
    }
}

Scene *soup_scene(const double &aspect, const float *tri9, size_t n, frt_scene_ids *ids)
{
    std::vector<hitable *> list, lights;
    Vector3f *v = new Vector3f[n * 3];
    Vector3f *nn = new Vector3f[n * 3];
    Vector2f *uv = new Vector2f[n * 3];
    std::vector<int> idx(n * 3);
    for (size_t i = 0; i < n * 3; ++i)
    {
        v[i] = Vector3f(tri9[3 * i], tri9[3 * i + 1], tri9[3 * i + 2]);
        nn[i] = Vector3f(0, 0, 1);
        idx[i] = (int)i;
    }
    auto mesh = std::make_shared<triangle_mesh>((int)n, (int)(n * 3), v, idx.data(), nn, uv,
                                                std::make_unique<lambertian>(new constant_texture(Vector3f(0.5, 0.5, 0.5))), true, "soup", true);
    add_mesh(mesh, list, lights);
    camera cam(Vector3f(0.5, 0.5, 3.0), Vector3f(0.5, 0.5, 0.5), Vector3f(0, 1, 0), 30.0, aspect, 0.0, 10.0);
    return finish(list, cam, Vector3f(1, 1, 1), lights, ids);
}

Scene *make_scene(const std::string &name, double aspect, frt_scene_ids *ids)
{
    // "<scene>+sky": the same scene under a constant sky-blue environment (the shipped scenes are lit by
    // emitters under a black environment, where ao::Li is identically zero); the oracle driver does the same
    const std::string sky = "+sky";
    if (name.size() > sky.size() && name.compare(name.size() - sky.size(), sky.size(), sky) == 0)
    {
        Scene *s = make_scene(name.substr(0, name.size() - sky.size()), aspect, ids);
        s->env_map.reset(new environment_map(std::make_unique<constant_texture>(Vector3f(0.5, 0.6, 0.8))));
        return s;
    }
    // "<scene>+hdr": the same scene under the HDR lat-long environment image main.cpp:170 loads (data/test.hdr,
    // environment_map(filename) material.h:209-213)
    const std::string hdr = "+hdr";
    if (name.size() > hdr.size() && name.compare(name.size() - hdr.size(), hdr.size(), hdr) == 0)
    {
        Scene *s = make_scene(name.substr(0, name.size() - hdr.size()), aspect, ids);
        s->env_map.reset(new environment_map(asset("data/test.hdr")));
        return s;
    }
    if (name == "cornell" || name == "cornell_box_obj") return cornell_box_obj(aspect, "Original", ids);
    if (name == "cornell_glossy") return cornell_box_obj(aspect, "Glossy", ids);
    if (name == "cornell_sphere") return cornell_box_obj(aspect, "Sphere", ids);
    if (name == "cornell_metal") return cornell_box_obj(aspect, "Metal", ids);
    if (name == "cornell_lens") return cornell_box_obj(aspect, "Lens", ids);
    if (name == "cornell_ao" || name == "cornell_box_ao") return cornell_box_obj(aspect, "Ao", ids);
    if (name == "veach" || name == "veach_mis") return veach_mis(aspect, ids, "ggx");
    if (name == "veach_beckmann") return veach_mis(aspect, ids, "beckmann");
    if (name == "textured") return textured_scene(aspect, ids);
    if (name == "random_scene") return random_scene(aspect, ids);
    if (name == "teapot") return teapot_scene(aspect, ids);
    if (name == "furnace" || name == "furnace_test_scene") return furnace_test_scene(aspect, ids);
    if (name.rfind("soup:", 0) == 0)
    {
        const size_t n = (size_t)std::atoll(name.c_str() + 5);
        std::vector<float> t;
        soup_triangles(n, 0x5EED0001ull, 0.004f, t);
        return soup_scene(aspect, t.data(), n, ids);
    }
    throw std::runtime_error("unknown scene '" + name + "'");
}
