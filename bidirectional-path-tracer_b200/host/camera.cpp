// camera::render (src/camera.h:43-145): the frame loop, re-hosted on libbdpt.  The worker threads and the
// shared-mt19937 sample loop of the reference are replaced by one host thread per GPU driving the
// wavefront kernels; strata are sharded across GPUs and the film sums are added on the host
// (the NCCL path lives in the torch.distributed harness of bench.py; this CLI has no NCCL dependency).
#include <chrono>
#include <iostream>
#include <stdexcept>
#include <thread>

#include "scene.h"

namespace {
struct guard {
    bdpt_ctx* c = nullptr;
    ~guard() { if (c) bdpt_ctx_destroy(c); }
};
void check(int rc, bdpt_ctx* ctx, const char* what) {
    if (rc) throw std::runtime_error(std::string(what) + " failed: " + bdpt_last_error(ctx));
}
}  // namespace

void camera::render(const hittable& world, const hittable& lights) {
    (void)lights;
    const std::vector<triangle>* prims = nullptr;
    if (auto b = dynamic_cast<const bvh_node*>(&world)) prims = &b->prims;
    else if (auto t = dynamic_cast<const triangle_collection*>(&world)) prims = &t->tris;
    if (!prims) throw std::runtime_error("camera::render: world must be a bvh_node or a triangle_collection");
    bool from_image = options.scene_image != nullptr;
    const void* image = options.scene_image;
    uint64_t image_bytes = options.scene_image_bytes;
    std::vector<double> verts; std::vector<uint32_t> mat_id; std::vector<bdpt_material> mats;
    if (!from_image) flatten_triangles(*prims, verts, mat_id, mats);
    // several GPUs: flatten and build the two trees ONCE into a scene image and hand every context the same block,
    // instead of letting each worker's bdpt_scene_upload rebuild them side by side on the same host cores
    struct image_owner { void* p = nullptr; ~image_owner() { if (p) bdpt_scene_image_free(p); } } built;
    if (!from_image && options.gpus > 1) {
        if (bdpt_scene_image_build(verts.data(), mat_id.data(), mat_id.size(), mats.data(), uint32_t(mats.size()), 0, &built.p, &image_bytes))
            throw std::runtime_error(std::string("bdpt_scene_image_build failed: ") + bdpt_scene_image_last_error());
        image = built.p;
        from_image = true;
    }

    bdpt_camera c{};
    c.aspect_ratio = aspect_ratio; c.image_width = image_width; c.samples_per_pixel = samples_per_pixel;
    c.max_depth = max_depth; c.vfov = vfov; c.defocus_angle = defocus_angle; c.focus_dist = focus_dist;
    for (int k = 0; k < 3; ++k) { c.background[k] = background[k]; c.lookfrom[k] = lookfrom[k]; c.lookat[k] = lookat[k]; c.vup[k] = vup[k]; }
    int H = int(image_width / aspect_ratio);   // camera.h:161-162
    H = H < 1 ? 1 : H;
    const int W = image_width;
    const int G = options.gpus < 1 ? 1 : options.gpus;

    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::vector<double>> films(G);
    std::vector<std::string> errors(G);
    std::vector<bdpt_stats_t> stats(G);
    std::vector<double> ms(G, 0.0);
    auto worker = [&](int g) {
        try {
            guard ctx;
            if (bdpt_ctx_create(g, &ctx.c)) throw std::runtime_error(std::string("bdpt_ctx_create: ") + bdpt_last_error(nullptr));
            if (from_image) check(bdpt_scene_upload_image(ctx.c, image, image_bytes, 0), ctx.c, "bdpt_scene_upload_image");
            else check(bdpt_scene_upload(ctx.c, verts.data(), mat_id.data(), mat_id.size(), mats.data(), uint32_t(mats.size())), ctx.c, "bdpt_scene_upload");
            check(bdpt_film_clear(ctx.c, W, H), ctx.c, "bdpt_film_clear");
            bdpt_render_opts o{};
            o.seed = options.seed; o.stratum_begin = g; o.stratum_step = G; o.bvh = options.bvh;
            check(bdpt_render(ctx.c, &c, &o), ctx.c, "bdpt_render");
            films[g].resize(size_t(W) * H * 3);
            check(bdpt_film_download(ctx.c, films[g].data()), ctx.c, "bdpt_film_download");
            bdpt_stats(ctx.c, &stats[g]);
            bdpt_last_device_ms(ctx.c, &ms[g]);
        } catch (const std::exception& e) { errors[g] = e.what(); }
    };
    std::vector<std::thread> pool;
    for (int g = 1; g < G; ++g) pool.emplace_back(worker, g);
    worker(0);
    for (auto& th : pool) th.join();
    for (int g = 0; g < G; ++g) if (!errors[g].empty()) throw std::runtime_error("GPU " + std::to_string(g) + ": " + errors[g]);

    render_report rep;
    rep.width = W; rep.height = H;
    rep.film = films[0];
    for (int g = 1; g < G; ++g) for (size_t i = 0; i < rep.film.size(); ++i) rep.film[i] += films[g][i];
    for (int g = 0; g < G; ++g) {
        rep.samples += stats[g].samples; rep.rays_closest += stats[g].rays_closest; rep.rays_shadow += stats[g].rays_shadow;
        rep.launches += stats[g].kernel_launches; rep.device_ms = std::max(rep.device_ms, ms[g]);
    }
    rep.wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    if (options.write_file) {
        const int n = int(std::sqrt(double(samples_per_pixel)));
        std::vector<color> fb(size_t(W) * H);
        for (size_t i = 0; i < fb.size(); ++i) fb[i] = color(rep.film[3 * i], rep.film[3 * i + 1], rep.film[3 * i + 2]);
        std::vector<uint8_t> rgb;
        colors_to_rgb8(fb, W, H, n * n, rgb);          // camera.h:139-141
        write_png(file_name, rgb, W, H);               // camera.h:142
    }
    if (!options.quiet) {
        std::clog << "Rendered " << W << "x" << H << " @ " << rep.samples / double(size_t(W) * H) << " spp on " << G << " GPU(s): "
                  << rep.device_ms << " ms device, " << rep.wall_s << " s wall, "
                  << rep.samples / (rep.device_ms * 1e3) << " Msamples/s, "
                  << (rep.rays_closest + rep.rays_shadow) / (rep.device_ms * 1e3) << " Mrays/s ("
                  << rep.rays_closest << " closest + " << rep.rays_shadow << " shadow), " << rep.launches << " kernel launches\n";
    }
    last_report = std::move(rep);
}
