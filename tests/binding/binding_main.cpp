// TEST INFRASTRUCTURE.  The drop-in proof: the reference's REAL headers (included from where they lie, nothing copied)
// + tests/binding/bdpt_binding.h + libbdpt.so in one translation unit, as a maintainer of the reference would build it.
// Built by oracle/Makefile (target _ref/binding_demo) only where /root/reference exists; the GPU box runs the
// prebuilt binary.  The scene comes from a small binary file the test writes, and is rebuilt here with the
// reference's OWN classes (triangle, triangle_collection, lambertian / metal / dielectric / diffuse_light, camera), so
// what is exercised is exactly the path reference object graph -> bdpt_binding.h -> C-ABI -> PNG.
//
// usage: binding_demo <scene.bin> <out.png> [seed]
//   scene.bin: u64 n_tris, u32 n_mats, u32 pad | n_tris x 9 f64 | n_tris x u32 | n_mats x {u32 type, u32 pad, 3 f64 rgb, f64 param}
//              | camera: f64 aspect, i32 width, i32 spp, i32 depth, i32 pad, 3 f64 background, f64 vfov, 3 f64 from, 3 f64 at, 3 f64 vup
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <filesystem>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#define private public          // read access to triangle::v0.. and the material parameters (see bdpt_binding.h)
#include "main.h"
#include "bvh.h"
#include "camera.h"
#include "hittable.h"
#include "material.h"
#include "triangle.h"
#include "texture.h"
#include "wpng.h"
#undef private

// libpng is not in this image (oracle/shim/png.h aborts): the PNG goes through libbdpt's own writer, same bytes in
#define BDPT_BINDING_WRITE_PNG(name, rgb, w, h) \
    do { if (bdpt_write_png((name).c_str(), (rgb).data(), (w), (h))) throw std::runtime_error(bdpt_host_last_error()); } while (0)
#include "bdpt_binding.h"

namespace {
template <class T> T rd(std::ifstream& f) { T v; f.read(reinterpret_cast<char*>(&v), sizeof v); if (!f) throw std::runtime_error("scene.bin: truncated"); return v; }
}

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: %s scene.bin out.png [seed]\n", argv[0]); return 2; }
    try {
        std::ifstream f(argv[1], std::ios::binary);
        if (!f) throw std::runtime_error("cannot open scene.bin");
        const uint64_t n_tris = rd<uint64_t>(f);
        const uint32_t n_mats = rd<uint32_t>(f);
        rd<uint32_t>(f);
        std::vector<double> v(n_tris * 9);
        std::vector<uint32_t> mid(n_tris);
        f.read(reinterpret_cast<char*>(v.data()), v.size() * 8);
        f.read(reinterpret_cast<char*>(mid.data()), mid.size() * 4);
        std::vector<shared_ptr<material>> mats;
        for (uint32_t k = 0; k < n_mats; ++k) {
            const uint32_t type = rd<uint32_t>(f);
            rd<uint32_t>(f);
            const double r = rd<double>(f), g = rd<double>(f), b = rd<double>(f), p = rd<double>(f);
            switch (type) {   // the loader's four classes, scene_loader.h:101-170
                case 0: mats.push_back(make_shared<lambertian>(color(r, g, b))); break;
                case 1: mats.push_back(make_shared<metal>(color(r, g, b), p)); break;
                case 2: mats.push_back(make_shared<dielectric>(p)); break;
                default: mats.push_back(make_shared<diffuse_light>(color(r, g, b))); break;
            }
        }
        triangle_collection world;
        for (uint64_t i = 0; i < n_tris; ++i) {
            const double* q = v.data() + i * 9;
            world.add(triangle(point3(q[0], q[1], q[2]), point3(q[3], q[4], q[5]), point3(q[6], q[7], q[8]), mats.at(mid[i])));
        }
        camera cam;
        cam.aspect_ratio = rd<double>(f);
        cam.image_width = rd<int32_t>(f);
        cam.samples_per_pixel = rd<int32_t>(f);
        cam.max_depth = rd<int32_t>(f);
        rd<int32_t>(f);
        { const double r = rd<double>(f), g = rd<double>(f), b = rd<double>(f); cam.background = color(r, g, b); }
        cam.vfov = rd<double>(f);
        { const double x = rd<double>(f), y = rd<double>(f), z = rd<double>(f); cam.lookfrom = point3(x, y, z); }
        { const double x = rd<double>(f), y = rd<double>(f), z = rd<double>(f); cam.lookat = point3(x, y, z); }
        { const double x = rd<double>(f), y = rd<double>(f), z = rd<double>(f); cam.vup = vec3(x, y, z); }
        cam.defocus_angle = 0;
        cam.file_name = argv[2];
        const uint64_t seed = argc > 3 ? std::strtoull(argv[3], nullptr, 10) : 0;
        bdpt_render_scene(cam, world, 0, seed);
        std::printf("binding_demo: %llu triangles, %u materials -> %s\n", (unsigned long long)n_tris, n_mats, argv[2]);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "binding_demo: %s\n", e.what());
        return 1;
    }
    return 0;
}
