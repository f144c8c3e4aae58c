// frt_main — the reference's main() (main.cpp:483-525) with the GPU integrator selectable:
//   frt_main [--ns N] [--scene cornell|veach|teapot|soup:N] [--res WxH] [--integrator gpu_path]
//            [--gpus N] [--seed S] [--out NAME] [--png FILE] [--checkpoint FILE] [--resume FILE]
// --checkpoint writes the film sums after the render; --resume starts from such a file and renders only the samples
// it does not hold yet (same seed, same resolution), so an interrupted --ns 512 job can be finished as --ns 1024.
// Defaults follow the reference: 512x512, ns = 100 (main.cpp:485-487).  The only integrator in
// this product is gpu_path (renderer<gpu_path_integrator>); asking for another one is an error
// rather than a silent CPU path.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include "gpu_path_integrator.h"
#include "integrator.h"
#include "png_writer.h"
#include "scenes.h"

namespace
{
// Film checkpoint, version 2: besides the film size and the samples it holds, everything that decides WHICH estimator
// those samples came from — scene (FNV-1a hash of its name), integrator, depth cap — so that --resume cannot add
// samples of another scene / integrator onto saved sums.
struct ckpt_header { char magic[8]; uint32_t nx, ny, spp_done, seed; uint32_t version, integrator, max_depth, scene_hash; };

uint32_t fnv1a(const std::string &s)
{
    uint32_t h = 2166136261u;
    for (unsigned char c : s) { h ^= c; h *= 16777619u; }
    return h;
}

void write_checkpoint(const std::string &path, int nx, int ny, uint32_t spp_done, uint32_t seed, uint32_t integrator, uint32_t max_depth,
                      const std::string &scene, const std::vector<float> &sums)
{
    ckpt_header h{{'F', 'R', 'T', 'C', 'K', 'P', 'T', '2'}, (uint32_t)nx, (uint32_t)ny, spp_done, seed, 2u, integrator, max_depth, fnv1a(scene)};
    std::ofstream f(path, std::ios::binary);
    f.write((const char *)&h, sizeof h);
    f.write((const char *)sums.data(), (std::streamsize)(sums.size() * sizeof(float)));
    if (!f) throw std::runtime_error("cannot write checkpoint " + path);
}

ckpt_header read_checkpoint(const std::string &path, int nx, int ny, uint32_t integrator, uint32_t max_depth, const std::string &scene,
                            std::vector<float> &sums)
{
    ckpt_header h{};
    std::ifstream f(path, std::ios::binary);
    f.read((char *)&h, sizeof h);
    if (!f || std::memcmp(h.magic, "FRTCKPT2", 8) != 0 || h.version != 2u) throw std::runtime_error("not a version-2 film checkpoint: " + path);
    if ((int)h.nx != nx || (int)h.ny != ny) throw std::runtime_error("checkpoint resolution differs from --res");
    if (h.scene_hash != fnv1a(scene)) throw std::runtime_error("checkpoint was rendered from another --scene");
    if (h.integrator != integrator) throw std::runtime_error("checkpoint was rendered by another --integrator");
    if (h.max_depth != max_depth) throw std::runtime_error("checkpoint was rendered with another depth cap");
    sums.resize((size_t)nx * ny * 3);
    f.read((char *)sums.data(), (std::streamsize)(sums.size() * sizeof(float)));
    if (!f) throw std::runtime_error("truncated checkpoint " + path);
    return h;
}
} // namespace

int main(int argc, const char **argv)
{
    int nx = 512, ny = 512;
    uint_fast64_t ns = 100;
    constexpr int comp = 3;
    std::string scene_name = "cornell", integrator_name = "gpu_path", out = "out_test", checkpoint, resume, png;
    int gpus = 1;
    uint32_t seed = 1;
    for (int i = 1; i < argc; ++i)
    {
        const std::string a = argv[i];
        auto next = [&]() -> const char * { if (i + 1 >= argc) { std::cerr << "missing value after " << a << "\n"; std::exit(2); } return argv[++i]; };
        if (a == "--ns") ns = std::atoll(next());
        else if (a == "--scene") scene_name = next();
        else if (a == "--integrator") integrator_name = next();
        else if (a == "--gpus") gpus = std::atoi(next());
        else if (a == "--seed") seed = (uint32_t)std::atoll(next());
        else if (a == "--out") out = next();
        else if (a == "--checkpoint") checkpoint = next();
        else if (a == "--png") png = next();
        else if (a == "--resume") resume = next();
        else if (a == "--res") { const char *r = next(); if (std::sscanf(r, "%dx%d", &nx, &ny) != 2) { std::cerr << "bad --res\n"; return 2; } }
        else { std::cerr << "unknown argument " << a << "\n"; return 2; }
    }
    uint32_t li = FRT_INTEGRATOR_PATH;
    if (integrator_name == "gpu_ao") li = FRT_INTEGRATOR_AO;
    else if (integrator_name == "gpu_normals") li = FRT_INTEGRATOR_NORMALS;
    else if (integrator_name == "gpu_ass_1") li = FRT_INTEGRATOR_DIRECT;
    else if (integrator_name != "gpu_path")
    {
        std::cerr << "integrator '" << integrator_name << "' is not part of this build (gpu_path, gpu_ao, gpu_normals, gpu_ass_1)\n";
        return 2;
    }
    viewer film_viewer(nx, ny, ns, comp);
    std::cout << "Resolution: " << nx << "x" << ny << std::endl;
    std::cout << "Setting number of samples to " << ns << std::endl;
    try
    {
        const auto t11 = std::chrono::high_resolution_clock::now();
        std::unique_ptr<Scene> scene(make_scene(scene_name, double(nx) / double(ny)));
        const auto t22 = std::chrono::high_resolution_clock::now();
        std::cout << "\nBVH construction took me " << std::chrono::duration<double>(t22 - t11).count() << " seconds.";

        renderer<gpu_path_integrator> render;
        render.n_gpus = gpus;
        render.seed = seed;
        render.integrator = li;
        std::vector<float> resume_sums, final_sums;
        if (!resume.empty())
        {
            const ckpt_header h = read_checkpoint(resume, nx, ny, li, render.max_depth, scene_name, resume_sums);
            if (h.spp_done > ns) throw std::runtime_error("checkpoint already holds more samples than --ns");
            render.seed = h.seed;
            render.spp_begin = h.spp_done;
            render.resume_sums = &resume_sums;
            std::cout << "\nResuming at sample " << h.spp_done << " of " << ns << " (seed " << h.seed << ")";
        }
        if (!checkpoint.empty()) render.film_sums_out = &final_sums;
        const double render_time = render.Render(scene.get(), film_viewer);
        const frt_stats &st = render.stats;
        std::cout << " GPU: " << st.samples << " samples, " << st.closest_rays << " closest + " << st.shadow_rays
                  << " shadow rays in " << st.render_ms << " ms device time = "
                  << (st.closest_rays + st.shadow_rays) / (st.render_ms * 1e3) << " Mrays/s" << std::endl;
        if (!checkpoint.empty()) write_checkpoint(checkpoint, nx, ny, (uint32_t)ns, render.seed, li, render.max_depth, scene_name, final_sums);
        // what the reference's window shows (viewer::out_image, viewer.cpp:117-125), as a file
        if (!png.empty() && !frt_png::write_rgb8(png, film_viewer.out_image.get(), nx, ny)) throw std::runtime_error("cannot write " + png);
        film_viewer.save_and_destroy(out + "_" + std::to_string(render_time) + "s");
        gpu_path_integrator::release_cached_context();
    }
    catch (const std::exception &e)
    {
        std::cerr << "frt_main: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
