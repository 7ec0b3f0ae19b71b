// image_out.cpp — see image_out.hpp.
#include "image_out.hpp"

#include <zlib.h>

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace rtbhost {

namespace {

struct File {
    FILE* f;
    explicit File(const std::string& path) : f(std::fopen(path.c_str(), "wb")) {
        if (!f) throw std::runtime_error("cannot open " + path + " for writing");
    }
    ~File() { if (f) std::fclose(f); }
    void write(const void* p, size_t n) {
        if (n && std::fwrite(p, 1, n, f) != n) throw std::runtime_error("short write");
    }
};

void be32(uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; }

void chunk(File& out, const char tag[4], const uint8_t* data, size_t n) {
    uint8_t hdr[8];
    be32(hdr, (uint32_t)n);
    std::memcpy(hdr + 4, tag, 4);
    out.write(hdr, 8);
    out.write(data, n);
    uLong crc = crc32(0L, reinterpret_cast<const Bytef*>(tag), 4);
    if (n) crc = crc32(crc, data, (uInt)n);
    uint8_t c[4];
    be32(c, (uint32_t)crc);
    out.write(c, 4);
}

}  // namespace

void write_ppm(const std::string& path, const uint8_t* rgb8, int width, int height) {
    File out(path);
    char hdr[64];
    int n = std::snprintf(hdr, sizeof hdr, "P6\n%d %d\n255\n", width, height);
    out.write(hdr, (size_t)n);
    out.write(rgb8, (size_t)width * height * 3);
}

void write_png(const std::string& path, const uint8_t* rgb8, int width, int height) {
    const size_t row = (size_t)width * 3;
    std::vector<uint8_t> raw((row + 1) * (size_t)height);
    for (int y = 0; y < height; y++) {
        raw[(row + 1) * y] = 0;   // filter type 0 (None)
        std::memcpy(&raw[(row + 1) * y + 1], rgb8 + row * y, row);
    }
    uLongf bound = compressBound((uLong)raw.size());
    std::vector<uint8_t> z(bound);
    if (compress2(z.data(), &bound, raw.data(), (uLong)raw.size(), 6) != Z_OK) throw std::runtime_error("zlib compress2 failed");
    File out(path);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    out.write(sig, 8);
    uint8_t ihdr[13];
    be32(ihdr, (uint32_t)width); be32(ihdr + 4, (uint32_t)height);
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;   // 8-bit, colour type 2 (RGB)
    chunk(out, "IHDR", ihdr, 13);
    chunk(out, "IDAT", z.data(), bound);
    chunk(out, "IEND", nullptr, 0);
}

void save_image(const std::string& path, const uint8_t* rgb8, int width, int height) {
    const size_t n = path.size();
    if (n >= 4 && path.compare(n - 4, 4, ".ppm") == 0) write_ppm(path, rgb8, width, height);
    else write_png(path, rgb8, width, height);
}

}  // namespace rtbhost
