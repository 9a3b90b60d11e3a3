// camera::render (src/camera.h:43-145): the frame loop, re-hosted on libbdpt.  The worker threads and the
// shared-mt19937 sample loop of the reference are replaced by one host thread per GPU driving the
// wavefront kernels; strata are sharded across GPUs and the per-GPU film sums are added by ONE NCCL
// sum-reduce over NVLink onto GPU 0 (bdpt_film_reduce), which alone is read back.  The progress bar (camera.h:67-109)
// and the BvhStats block (stats.h:34-50, camera.h:144) keep the reference's format on std::clog.
#include <atomic>
#include <chrono>
#include <cmath>
#include <iomanip>
#include <iostream>
#include <map>
#include <stdexcept>
#include <thread>

#include "scene.h"

namespace {
struct ctx_set {
    std::vector<bdpt_ctx*> c;
    ~ctx_set() { for (auto* x : c) if (x) bdpt_ctx_destroy(x); }
};
void check(int rc, bdpt_ctx* ctx, const char* what) {
    if (rc) throw std::runtime_error(std::string(what) + " failed: " + bdpt_last_error(ctx));
}
}  // namespace

// hittable.h:81-84: the angle's sine and cosine, once
rotate_y::rotate_y(shared_ptr<hittable> obj, double angle_degrees) : object(std::move(obj)) {
    const double radians = angle_degrees * 3.1415926535897932385 / 180.0;   // degrees_to_radians, main.h:24-26
    sin_theta = std::sin(radians);
    cos_theta = std::cos(radians);
}

void collect_triangles(const hittable& h, std::vector<triangle>& out, std::vector<constant_medium>* media, std::vector<sphere>* spheres) {
    if (auto t = dynamic_cast<const triangle*>(&h)) { out.push_back(*t); return; }
    if (auto c = dynamic_cast<const triangle_collection*>(&h)) { out.insert(out.end(), c->tris.begin(), c->tris.end()); return; }
    if (auto b = dynamic_cast<const bvh_node*>(&h)) {
        out.insert(out.end(), b->prims.begin(), b->prims.end());
        if (!b->media.empty()) {
            if (!media) throw std::runtime_error("collect_triangles: a constant_medium where only triangles are allowed");
            media->insert(media->end(), b->media.begin(), b->media.end());
        }
        if (!b->spheres.empty()) {
            if (!spheres) throw std::runtime_error("collect_triangles: an analytic sphere where only triangles are allowed");
            spheres->insert(spheres->end(), b->spheres.begin(), b->spheres.end());
        }
        return;
    }
    if (auto l = dynamic_cast<const hittable_list*>(&h)) { for (const auto& o : l->objects) collect_triangles(*o, out, media, spheres); return; }
    if (auto sp = dynamic_cast<const sphere*>(&h)) {
        if (!spheres) throw std::runtime_error("collect_triangles: an analytic sphere where only triangles are allowed");
        spheres->push_back(*sp);
        return;
    }
    if (auto cm = dynamic_cast<const constant_medium*>(&h)) {
        if (!media) throw std::runtime_error("collect_triangles: a constant_medium where only triangles are allowed");
        media->push_back(*cm);
        return;
    }
    const size_t first = out.size(), first_medium = media ? media->size() : 0, first_sphere = spheres ? spheres->size() : 0;
    // a transform around a medium moves its boundary: the medium's boundary is replaced by the moved triangles
    auto move_media = [&](auto&& wrap) {
        if (!media) return;
        for (size_t k = first_medium; k < media->size(); ++k) {
            auto moved = make_shared<triangle_collection>();
            collect_triangles(*wrap((*media)[k].boundary), moved->tris);
            (*media)[k].boundary = moved;
        }
    };
    if (auto tr = dynamic_cast<const translate*>(&h)) {
        collect_triangles(*tr->object, out, media, spheres);
        for (size_t i = first; i < out.size(); ++i)
            for (point3* p : {&out[i].v0, &out[i].v1, &out[i].v2})   // hittable.h:64: rec.p += offset
                *p = point3(p->x() + tr->offset.x(), p->y() + tr->offset.y(), p->z() + tr->offset.z());
        for (size_t i = first_sphere; spheres && i < spheres->size(); ++i) {
            point3& c = (*spheres)[i].center;
            c = point3(c.x() + tr->offset.x(), c.y() + tr->offset.y(), c.z() + tr->offset.z());
        }
        move_media([&](const shared_ptr<hittable>& b) { return make_shared<translate>(b, tr->offset); });
        return;
    }
    if (auto ro = dynamic_cast<const rotate_y*>(&h)) {
        collect_triangles(*ro->object, out, media, spheres);
        const double c = ro->cos_theta, s = ro->sin_theta;
        for (size_t i = first; i < out.size(); ++i)
            for (point3* p : {&out[i].v0, &out[i].v1, &out[i].v2})   // object -> world, hittable.h:134-138
                *p = point3((c * p->x()) + (s * p->z()), p->y(), (-s * p->x()) + (c * p->z()));
        for (size_t i = first_sphere; spheres && i < spheres->size(); ++i) {
            point3& q = (*spheres)[i].center;
            q = point3((c * q.x()) + (s * q.z()), q.y(), (-s * q.x()) + (c * q.z()));
        }
        move_media([&](const shared_ptr<hittable>& b) { auto r = make_shared<rotate_y>(b, 0.0); r->sin_theta = s; r->cos_theta = c; return r; });
        return;
    }
    throw std::runtime_error("collect_triangles: unknown hittable (triangle, triangle_collection, hittable_list, bvh_node, translate, rotate_y)");
}

void flatten_spheres(const std::vector<sphere>& spheres, const std::vector<triangle>& tris, std::vector<bdpt_material>& mats,
                     std::vector<bdpt_sphere>& out) {
    out.clear();
    if (spheres.empty()) return;
    // the table's order is first use over the triangles (flatten_triangles): rebuild its index, then look up / append
    std::map<const material*, uint32_t> index;
    for (const triangle& t : tris) index.emplace(t.mat.get(), uint32_t(index.size()));
    if (index.size() != mats.size()) throw std::runtime_error("flatten_spheres: the material table does not belong to these triangles");
    for (const sphere& sp : spheres) {
        if (!sp.mat || sp.mat->phase) throw std::runtime_error("sphere: needs a surface material");
        if (!(sp.radius > 0.0)) throw std::runtime_error("sphere: radius must be > 0");
        auto it = index.find(sp.mat.get());
        if (it == index.end()) {
            bdpt_material m{};
            m.type = uint32_t(sp.mat->kind);
            m.rgb[0] = sp.mat->rgb.x(); m.rgb[1] = sp.mat->rgb.y(); m.rgb[2] = sp.mat->rgb.z();
            m.param = sp.mat->param;
            it = index.emplace(sp.mat.get(), uint32_t(mats.size())).first;
            mats.push_back(m);
        }
        bdpt_sphere d{};
        d.center[0] = sp.center.x(); d.center[1] = sp.center.y(); d.center[2] = sp.center.z();
        d.radius = sp.radius; d.mat_id = it->second;
        out.push_back(d);
    }
}

void flatten_media(const std::vector<constant_medium>& media, std::vector<bdpt_medium>& out, std::vector<double>& boundary_verts) {
    out.clear(); boundary_verts.clear();
    for (const constant_medium& m : media) {
        std::vector<triangle> tris;
        collect_triangles(*m.boundary, tris);
        bdpt_medium d{};
        d.density = m.density;
        d.albedo[0] = m.albedo.x(); d.albedo[1] = m.albedo.y(); d.albedo[2] = m.albedo.z();
        d.first_tri = boundary_verts.size() / 9; d.n_tris = tris.size();
        for (const triangle& t : tris)
            for (const point3* p : {&t.v0, &t.v1, &t.v2}) { boundary_verts.push_back(p->x()); boundary_verts.push_back(p->y()); boundary_verts.push_back(p->z()); }
        out.push_back(d);
    }
}

void print_bvh_stats(std::ostream& os, const render_report& r) {   // stats.h:34-50, same lines
    os << "\nBVH Stats:\n";
    os << "  Rays traced:         " << r.rays_closest << "\n";
    os << "  BVH nodes built:     " << r.bvh_nodes_built << "\n";
    os << "  AABB tests:          " << r.aabb_tests << " (hits: " << r.aabb_hits << ")\n";
    os << "  BVH node visits:     " << r.bvh_node_visits << "\n";
    os << "  Triangle tests:      " << r.triangle_tests << " (hits: " << r.triangle_hits << ")\n";
}

void camera::render(const hittable& world, const hittable& lights) {
    (void)lights;
    const std::vector<triangle>* prims = nullptr;
    std::vector<triangle> collected;
    std::vector<constant_medium> media = options.media;
    std::vector<sphere> spheres = options.spheres;
    if (auto b = dynamic_cast<const bvh_node*>(&world)) {
        prims = &b->prims;
        media.insert(media.end(), b->media.begin(), b->media.end());
        spheres.insert(spheres.end(), b->spheres.begin(), b->spheres.end());
    } else if (auto t = dynamic_cast<const triangle_collection*>(&world)) prims = &t->tris;
    else { collect_triangles(world, collected, &media, &spheres); prims = &collected; }   // hittable_list, translate, rotate_y, constant_medium, sphere
    std::vector<bdpt_medium> media_desc;
    std::vector<double> media_verts;
    flatten_media(media, media_desc, media_verts);
    bool from_image = options.scene_image != nullptr;
    const void* image = options.scene_image;
    uint64_t image_bytes = options.scene_image_bytes;
    std::vector<double> verts; std::vector<uint32_t> mat_id; std::vector<bdpt_material> mats;
    std::vector<bdpt_texture> textures; std::vector<unsigned char> tex_data;
    std::vector<bdpt_sphere> sphere_desc;
    if (!from_image) {
        flatten_triangles(*prims, verts, mat_id, mats);
        flatten_textures(*prims, mats, textures, tex_data);
        flatten_spheres(spheres, *prims, mats, sphere_desc);   // may append the spheres' materials to the table
    } else if (!spheres.empty())
        throw std::runtime_error("camera::render: analytic spheres need the scene's material table, which a scene image does not expose; render without the image");
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

    using clock = std::chrono::steady_clock;
    const auto t0 = clock::now();
    ctx_set ctxs;
    ctxs.c.assign(G, nullptr);
    for (int g = 0; g < G; ++g)
        if (bdpt_ctx_create(g, &ctxs.c[g])) throw std::runtime_error(std::string("bdpt_ctx_create: ") + bdpt_last_error(nullptr));
    std::vector<std::string> errors(G);
    std::vector<bdpt_stats_t> stats(G);
    std::vector<double> ms(G, 0.0);
    std::atomic<int> rendering{G};
    auto worker = [&](int g) {
        try {
            bdpt_ctx* ctx = ctxs.c[g];
            if (from_image) check(bdpt_scene_upload_image(ctx, image, image_bytes, 0), ctx, "bdpt_scene_upload_image");
            else check(bdpt_scene_upload(ctx, verts.data(), mat_id.data(), mat_id.size(), mats.data(), uint32_t(mats.size())), ctx, "bdpt_scene_upload");
            if (!textures.empty())
                check(bdpt_scene_set_textures(ctx, textures.data(), uint32_t(textures.size()), tex_data.data(), tex_data.size()), ctx, "bdpt_scene_set_textures");
            if (!sphere_desc.empty())
                check(bdpt_scene_set_spheres(ctx, sphere_desc.data(), uint32_t(sphere_desc.size())), ctx, "bdpt_scene_set_spheres");
            if (!media_desc.empty())
                check(bdpt_scene_set_media(ctx, media_desc.data(), uint32_t(media_desc.size()), media_verts.data(), media_verts.size() / 9), ctx, "bdpt_scene_set_media");
            check(bdpt_film_clear(ctx, W, H), ctx, "bdpt_film_clear");
            bdpt_render_opts o{};
            o.seed = options.seed; o.stratum_begin = g; o.stratum_step = G; o.bvh = options.bvh;
            o.flags = options.stats ? BDPT_RENDER_STATS : 0;
            o.weighting = options.weighting;
            check(bdpt_render(ctx, &c, &o), ctx, "bdpt_render");
            bdpt_stats(ctx, &stats[g]);
            bdpt_last_device_ms(ctx, &ms[g]);
        } catch (const std::exception& e) { errors[g] = e.what(); }
        rendering.fetch_sub(1);
    };
    // Progress bar, camera.h:67-109: the reference counts finished rows; here the samples handed to the GPUs, in rows
    const int barWidth = progress_bar_length;
    const int total_rows = H;
    auto put_time = [](std::ostream& os, double seconds) {
        const int rem = int(std::ceil(seconds)), hrs = rem / 3600, mins = (rem % 3600) / 60, secs = rem % 60;
        os << (hrs > 0 ? (std::to_string(hrs) + ":") : "") << std::setfill('0') << std::setw(2) << mins << ':' << std::setw(2) << secs << std::setfill(' ');
    };
    std::thread reporter([&] {
        if (options.quiet || barWidth <= 0) return;
        while (rendering.load() > 0) {
            uint64_t done_s = 0, total_s = 0;
            for (int g = 0; g < G; ++g) { uint64_t d = 0, t = 0; bdpt_render_progress(ctxs.c[g], &d, &t); done_s += d; total_s += t; }
            const double progress = total_s ? double(done_s) / double(total_s) : 0.0;
            const int done = int(progress * total_rows), filled = int(progress * barWidth + 0.5);
            const double dt = std::chrono::duration<double>(clock::now() - t0).count();
            const double rps = dt > 0.0 ? done / dt : 0.0, sec_left = rps > 0.0 ? (total_rows - done) / rps : 0.0;
            std::clog << '\r' << '[' << std::string(filled, '#') << std::string(barWidth - filled, ' ') << "] " << std::setw(3) << int(progress * 100)
                      << "% | " << done << "/" << total_rows << " | ETA: ";
            put_time(std::clog, sec_left);
            std::clog << std::flush;
            for (int k = 0; k < 10 && rendering.load() > 0; ++k) std::this_thread::sleep_for(std::chrono::milliseconds(100));
        }
        std::clog << "\r[" << std::string(barWidth, '#') << "] 100% | " << total_rows << "/" << total_rows << " | Runtime: ";
        put_time(std::clog, std::chrono::duration<double>(clock::now() - t0).count());
        std::clog << "\n";
    });
    std::vector<std::thread> pool;
    for (int g = 1; g < G; ++g) pool.emplace_back(worker, g);
    worker(0);
    for (auto& th : pool) th.join();
    reporter.join();
    for (int g = 0; g < G; ++g) if (!errors[g].empty()) throw std::runtime_error("GPU " + std::to_string(g) + ": " + errors[g]);

    render_report rep;
    rep.width = W; rep.height = H;
    if (G > 1) {   // the one exchange step of the frame: NCCL sum-reduce of the films onto GPU 0
        check(bdpt_film_reduce(ctxs.c.data(), G, 0), ctxs.c[0], "bdpt_film_reduce");
        bdpt_last_device_ms(ctxs.c[0], &rep.film_reduce_ms);
    }
    rep.film.resize(size_t(W) * H * 3);
    check(bdpt_film_download(ctxs.c[0], rep.film.data()), ctxs.c[0], "bdpt_film_download");
    for (int g = 0; g < G; ++g) {
        rep.samples += stats[g].samples; rep.rays_closest += stats[g].rays_closest; rep.rays_shadow += stats[g].rays_shadow;
        rep.launches += stats[g].kernel_launches; rep.device_ms = std::max(rep.device_ms, ms[g]);
        rep.aabb_tests += stats[g].aabb_tests; rep.aabb_hits += stats[g].aabb_hits; rep.triangle_hits += stats[g].triangle_hits;
        rep.bvh_node_visits += stats[g].nodes_closest + stats[g].nodes_shadow;
        rep.triangle_tests += stats[g].tris_closest + stats[g].tris_shadow;
    }
    rep.bvh_nodes_built = stats[0].bvh_nodes_built;
    rep.wall_s = std::chrono::duration<double>(clock::now() - t0).count();

    if (options.write_file) {
        const int n = int(std::sqrt(double(samples_per_pixel)));
        std::vector<uint8_t> rgb(size_t(W) * H * 3);
        check(bdpt_film_to_rgb8(ctxs.c[0], n * n, rgb.data()), ctxs.c[0], "bdpt_film_to_rgb8");   // camera.h:139-141 (wpng.h:14-35 on the device film)
        write_png(file_name, rgb, W, H);               // camera.h:142
    }
    if (!options.quiet) {
        print_bvh_stats(std::clog, rep);               // camera.h:144
        if (!options.stats) std::clog << "  (box / node / triangle counters need --stats; rays traced = closest-hit queries, camera.h:334)\n";
        std::clog << "Rendered " << W << "x" << H << " @ " << rep.samples / double(size_t(W) * H) << " spp on " << G << " GPU(s): "
                  << rep.device_ms << " ms device, " << rep.wall_s << " s wall, "
                  << rep.samples / (rep.device_ms * 1e3) << " Msamples/s, "
                  << (rep.rays_closest + rep.rays_shadow) / (rep.device_ms * 1e3) << " Mrays/s ("
                  << rep.rays_closest << " closest + " << rep.rays_shadow << " shadow), " << rep.launches << " kernel launches";
        if (G > 1) std::clog << ", film reduce (NCCL) " << rep.film_reduce_ms << " ms";
        std::clog << "\n";
    }
    last_report = std::move(rep);
}
