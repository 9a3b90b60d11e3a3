// bdpt_binding.h -- the ONE file a maintainer of the reference adds to src/ to render through libbdpt.so.
//
// Include it after the reference's own headers (main.h, camera.h, triangle.h, material.h, texture.h, wpng.h) and
// call bdpt_render_scene(scene.cam, scene.world) where main.cpp:68 calls scene.cam.render(*world_bvh, lights);
// link -lbdpt.  It flattens the reference's own triangle_collection into the soup bdpt_scene_upload takes
// (prim ID = insertion order), maps the four material classes the loader can build
// (scene_loader.h:101-170) onto bdpt_material, copies the 13 public camera fields (camera.h:26-41), renders, and
// hands the 8-bit image to write_png (wpng.h:38).
//
// Access: triangle keeps v0 / v1 / v2 private (triangle.h:123-132) and the material classes keep their parameters
// private (material.h:65,87,120,145).  The binding needs read access to them: either compile the TU that includes the
// reference headers with `-Dprivate=public` (what tests/binding/binding_main.cpp and oracle/ref_harness.cpp do), or add
// `friend struct bdpt_access;` to those five classes.  Nothing else in the reference changes.
//
// This file is compiled and run against the reference's real headers by tests/test_binding_gpu.py (the binary is
// built by oracle/Makefile where /root/reference exists): its PNG equals, byte for byte, the one the ctypes binding
// produces for the same soup, camera and seed.
#pragma once
#include <bdpt.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#ifndef BDPT_BINDING_WRITE_PNG   // the reference's own writer by default (wpng.h:38: creates ./output/, throws on failure)
#define BDPT_BINDING_WRITE_PNG(name, rgb, w, h) write_png((name).c_str(), (rgb), (w), (h))
#endif

struct bdpt_access {
    static void rgb(const color& c, double out[3]) { out[0] = c.x(); out[1] = c.y(); out[2] = c.z(); }

    // material.h:42-147 -> bdpt_material.  Textures other than solid_color cannot come out of the reference's loader;
    // they are sampled at (0, 0, origin), which is exact for solid_color (texture.h:14-26).
    static bdpt_material material_of(const shared_ptr<material>& m) {
        bdpt_material o{};
        if (auto l = std::dynamic_pointer_cast<diffuse_light>(m)) {
            o.type = BDPT_MAT_LIGHT;
            rgb(l->tex->value(0, 0, point3(0, 0, 0)), o.rgb);
        } else if (auto d = std::dynamic_pointer_cast<dielectric>(m)) {
            o.type = BDPT_MAT_DIELECTRIC;
            o.param = d->refraction_index;
        } else if (auto k = std::dynamic_pointer_cast<metal>(m)) {
            o.type = BDPT_MAT_METAL;
            rgb(k->albedo, o.rgb);
            o.param = k->fuzz;
        } else if (auto a = std::dynamic_pointer_cast<lambertian>(m)) {
            o.type = BDPT_MAT_LAMBERTIAN;
            rgb(a->tex->value(0, 0, point3(0, 0, 0)), o.rgb);
        } else {
            throw std::runtime_error("bdpt_binding: material class outside the loader's four (scene_loader.h:101-170)");
        }
        return o;
    }

    // triangle_collection (triangle.h:135-230) -> n x 9 doubles in insertion order + material table
    static void flatten(const triangle_collection& world, std::vector<double>& verts, std::vector<uint32_t>& mat_id,
                        std::vector<bdpt_material>& mats) {
        std::map<const material*, uint32_t> seen;
        verts.reserve(world.tris.size() * 9);
        mat_id.reserve(world.tris.size());
        for (const triangle& t : world.tris) {
            const point3* p[3] = {&t.v0, &t.v1, &t.v2};   // the points as the loader stored them: v0 + edge re-rounds
            for (const point3* q : p) { verts.push_back(q->x()); verts.push_back(q->y()); verts.push_back(q->z()); }
            auto it = seen.find(t.mat.get());
            if (it == seen.end()) {
                it = seen.emplace(t.mat.get(), uint32_t(mats.size())).first;
                mats.push_back(material_of(t.mat));
            }
            mat_id.push_back(it->second);
        }
    }

    static bdpt_camera camera_of(const camera& cam) {   // camera.h:26-41
        bdpt_camera c{};
        c.aspect_ratio = cam.aspect_ratio; c.image_width = cam.image_width;
        c.samples_per_pixel = cam.samples_per_pixel; c.max_depth = cam.max_depth;
        c.vfov = cam.vfov; c.defocus_angle = cam.defocus_angle; c.focus_dist = cam.focus_dist;
        for (int k = 0; k < 3; ++k) {
            c.background[k] = cam.background[k]; c.lookfrom[k] = cam.lookfrom[k];
            c.lookat[k] = cam.lookat[k]; c.vup[k] = cam.vup[k];
        }
        return c;
    }
};

// camera::render(world, lights) (camera.h:43) on the GPU.  `lights` is implied: the triangles whose material is a
// diffuse_light, in insertion order (scene_loader.h:190-202), every triangle when there is none (main.cpp:67).
inline void bdpt_render_scene(const camera& cam, const triangle_collection& world, int device = 0, uint64_t seed = 0,
                              int bvh = BDPT_BVH_FAST) {
    std::vector<double> verts;
    std::vector<uint32_t> mat_id;
    std::vector<bdpt_material> mats;
    bdpt_access::flatten(world, verts, mat_id, mats);
    bdpt_ctx* ctx = nullptr;
    if (bdpt_ctx_create(device, &ctx)) throw std::runtime_error(std::string("bdpt: ") + bdpt_last_error(nullptr));
    const bdpt_camera c = bdpt_access::camera_of(cam);
    bdpt_render_opts o{};
    o.seed = seed;
    o.bvh = bvh;
    int W = 0, H = 0;
    std::vector<uint8_t> rgb;
    auto fail = [&](const char* what) {
        const std::string e = std::string(what) + ": " + bdpt_last_error(ctx);
        bdpt_ctx_destroy(ctx);
        throw std::runtime_error(e);
    };
    if (bdpt_scene_upload(ctx, verts.data(), mat_id.data(), mat_id.size(), mats.data(), uint32_t(mats.size()))) fail("bdpt_scene_upload");
    if (bdpt_render(ctx, &c, &o)) fail("bdpt_render");
    if (bdpt_film_size(ctx, &W, &H)) fail("bdpt_film_size");   // image_height = int(image_width / aspect_ratio), camera.h:161
    rgb.resize(size_t(W) * H * 3);
    const int n = int(std::sqrt(double(cam.samples_per_pixel)));   // sqrt_spp, camera.h:164
    if (bdpt_film_to_rgb8(ctx, std::max(1, n * n), rgb.data())) fail("bdpt_film_to_rgb8");   // camera.h:139-140 + colors_to_rgb8, wpng.h:14-35
    bdpt_ctx_destroy(ctx);
    BDPT_BINDING_WRITE_PNG(cam.file_name, rgb, W, H);
}
