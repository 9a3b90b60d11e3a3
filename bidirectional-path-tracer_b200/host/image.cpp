// Image output: colors_to_rgb8 and write_png of the reference's src/image/wpng.h:14-89, with the PNG
// container written directly on zlib (libpng is not in this image): 8-bit RGB, no interlace, filter 0.
#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <filesystem>
#include <stdexcept>

#include "scene.h"

void colors_to_rgb8(const std::vector<color>& fb, int W, int H, int samples_per_pixel, std::vector<uint8_t>& out) {
    out.resize(size_t(W) * size_t(H) * 3);
    const double scale = 1.0 / std::max(1, samples_per_pixel);                  // wpng.h:18
    auto clampd = [](double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); };
    for (size_t i = 0; i < size_t(W) * size_t(H); ++i)
        for (int c = 0; c < 3; ++c)                                             // gamma 2, clamp [0,0.999], 256x truncate
            out[3 * i + c] = static_cast<uint8_t>(256.0 * std::sqrt(clampd(fb[i][c] * scale, 0.0, 0.999)));
}

namespace {
void put32(std::vector<uint8_t>& v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back(uint8_t(x >> s)); }
void chunk(std::FILE* fp, const char type[4], const std::vector<uint8_t>& data) {
    std::vector<uint8_t> head;
    put32(head, uint32_t(data.size()));
    std::fwrite(head.data(), 1, 4, fp);
    std::fwrite(type, 1, 4, fp);
    if (!data.empty()) std::fwrite(data.data(), 1, data.size(), fp);
    uLong crc = crc32(0L, reinterpret_cast<const Bytef*>(type), 4);
    if (!data.empty()) crc = crc32(crc, data.data(), uInt(data.size()));
    std::vector<uint8_t> tail;
    put32(tail, uint32_t(crc));
    std::fwrite(tail.data(), 1, 4, fp);
}
}  // namespace

void write_png_path(const std::string& path, const uint8_t* rgb, int W, int H) {
    if (!rgb || W <= 0 || H <= 0) throw std::runtime_error("write_png: bad image");
    std::FILE* fp = std::fopen(path.c_str(), "wb");
    if (!fp) throw std::runtime_error("write_png: cannot open output file");
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::fwrite(sig, 1, 8, fp);
    std::vector<uint8_t> ihdr;
    put32(ihdr, uint32_t(W)); put32(ihdr, uint32_t(H));
    ihdr.insert(ihdr.end(), {8, 2, 0, 0, 0});   // bit depth 8, colour type RGB, deflate, filter base, no interlace
    chunk(fp, "IHDR", ihdr);
    std::vector<uint8_t> raw(size_t(H) * (size_t(W) * 3 + 1));
    for (int j = 0; j < H; ++j) {
        raw[size_t(j) * (size_t(W) * 3 + 1)] = 0;   // filter type None
        std::copy(rgb + size_t(j) * W * 3, rgb + size_t(j + 1) * W * 3, raw.begin() + size_t(j) * (size_t(W) * 3 + 1) + 1);
    }
    uLongf clen = compressBound(uLong(raw.size()));
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), uLong(raw.size()), 6) != Z_OK) { std::fclose(fp); throw std::runtime_error("write_png: deflate failed"); }
    comp.resize(clen);
    chunk(fp, "IDAT", comp);
    chunk(fp, "IEND", {});
    std::fclose(fp);
}

void write_png(const std::string& file_name, const std::vector<uint8_t>& rgb, int W, int H) {   // wpng.h:38-89
    if (rgb.size() != size_t(W) * size_t(H) * 3) throw std::runtime_error("write_png: RGB buffer size mismatch");
    std::filesystem::create_directories("output");                                              // wpng.h:45-49
    write_png_path((std::filesystem::path("output") / file_name).string(), rgb.data(), W, H);
}
