// Headless film with the reference viewer's interface (viewer.h:9-26).  The GLFW/GLEW preview
// window of viewer.cpp:18-107 is unavailable on this box (no graphics stack): init() returns
// nullptr and background_thread() just waits for the render to finish.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <future>
#include <memory>
#include <string>
#include "geometry.h"

struct GLFWwindow;
typedef unsigned char GLubyte;
typedef double GLdouble;

struct viewer
{
    viewer(int nx_, int ny_, uint_fast64_t ns_, int num_channels_)
        : to_exit(false), nx(nx_), ny(ny_), ns(ns_), num_channels(num_channels_), window(nullptr)
    {
        const size_t n = (size_t)nx * ny * num_channels + 64; // viewer.cpp:12-13 slack
        out_image.reset(new GLubyte[n]());
        fout_image.reset(new GLdouble[n]());
    }
    GLFWwindow *init() { return nullptr; }
    void background_thread(const std::shared_future<void> &future, GLFWwindow *) { future.wait(); to_exit = true; }
    // viewer.cpp:109-132: `sample` is the SUM over ns samples; row 0 is the bottom row
    void add_sample(const Vector2i &pixel, Vector3f &sample)
    {
        sample /= double(ns);
        const size_t index = ((size_t)pixel[1] * nx + pixel[0]) * num_channels;
        for (int c = 0; c < 3; ++c)
        {
            out_image[index + c] = (GLubyte) int(std::pow(1 - std::exp(-sample[c]), 1 / 2.2) * 255 + .5);
            fout_image[index + c] = sample[c];
        }
    }
    // image.h:89-118 PFM layout (rows in buffer order, little endian, scale -1); the reference
    // prepends $HOME/ on Linux (image.h:94-98) — here the name is used as given.
    void save_and_destroy(const std::string &filename)
    {
        std::printf("Saving PFM...\n");
        FILE *f = std::fopen((filename + ".pfm").c_str(), "wb");
        if (!f) return;
        std::fprintf(f, "PF\n%d %d\n-1\n", nx, ny);
        for (size_t i = 0; i < (size_t)nx * ny * num_channels; ++i)
        {
            const float v = (float)fout_image[i];
            std::fwrite(&v, sizeof v, 1, f);
        }
        std::fclose(f);
    }

    bool to_exit;
    const int nx, ny;
    const uint_fast64_t ns;
    const int num_channels;
    std::unique_ptr<GLubyte[]> out_image;
    std::unique_ptr<GLdouble[]> fout_image;
    GLFWwindow *window;
};
