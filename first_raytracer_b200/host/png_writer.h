// Minimal PNG writer for the 8-bit tone-mapped film (viewer::out_image).  The reference shows that buffer in a GLFW
// window (viewer.cpp:18-107), which does not exist on a headless box: here it goes to a file instead.  Self-contained
// (the reference's image::save_image, image.cpp:24-58, goes through stb_image_write): 8-bit RGB, no filter, zlib
// "stored" blocks — nothing to decompress wrongly, at 3 bytes per pixel.  Row 0 of the film is the BOTTOM row
// (viewer.cpp:111 indexes y upwards), PNG rows run top-down, hence the flip (image.cpp:30 flips on write too).
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace frt_png {

inline uint32_t crc32(const unsigned char *p, size_t n, uint32_t crc = 0)
{
    static uint32_t table[256];
    static bool ready = false;
    if (!ready)
    {
        for (uint32_t i = 0; i < 256; ++i)
        {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return ~crc;
}

inline void put32(std::vector<unsigned char> &v, uint32_t x)
{
    v.push_back((unsigned char)(x >> 24)); v.push_back((unsigned char)(x >> 16)); v.push_back((unsigned char)(x >> 8)); v.push_back((unsigned char)x);
}

inline void chunk(std::vector<unsigned char> &out, const char tag[4], const std::vector<unsigned char> &data)
{
    put32(out, (uint32_t)data.size());
    const size_t start = out.size();
    out.insert(out.end(), tag, tag + 4);
    out.insert(out.end(), data.begin(), data.end());
    put32(out, crc32(out.data() + start, out.size() - start));
}

// rgb: ny rows of nx RGB pixels, row 0 = bottom.  Returns false when the file cannot be written.
inline bool write_rgb8(const std::string &path, const unsigned char *rgb, int nx, int ny)
{
    if (nx <= 0 || ny <= 0 || !rgb) return false;
    std::vector<unsigned char> raw;
    raw.reserve((size_t)ny * (3 * (size_t)nx + 1));
    for (int y = ny - 1; y >= 0; --y)
    {
        raw.push_back(0); // filter type: none
        raw.insert(raw.end(), rgb + (size_t)y * nx * 3, rgb + (size_t)(y + 1) * nx * 3);
    }
    std::vector<unsigned char> z;
    z.push_back(0x78); z.push_back(0x01); // zlib header, no compression
    uint32_t a = 1, b = 0;                // Adler-32 of the raw data
    for (size_t pos = 0; pos < raw.size();)
    {
        const size_t n = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + n == raw.size() ? 1 : 0); // BFINAL, BTYPE = 00 (stored)
        z.push_back((unsigned char)(n & 0xFF)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 0xFF)); z.push_back((unsigned char)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        for (size_t i = pos; i < pos + n; ++i) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
        pos += n;
    }
    put32(z, (b << 16) | a);
    std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<unsigned char> ihdr;
    put32(ihdr, (uint32_t)nx); put32(ihdr, (uint32_t)ny);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0); // 8 bit, RGB, deflate, no filter, no interlace
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    std::fclose(f);
    return ok;
}

} // namespace frt_png
