// Entry point: `bdpt scene.yaml` exactly as the reference's src/main.cpp:62-78, plus overrides for the
// settings BASELINE.json's configs change (the YAMLs say 800x800 / 1280x720).
#include <cstring>
#include <iostream>

#include "scene.h"

int main(int argc, char** argv) {
    if (argc < 2) {
        std::cerr << "usage: bdpt scene.yaml [--res WxH] [--spp N] [--depth N] [--gpus N] [--bvh ref|fast] [--mis] [--seed S] [--out name.png] [--cache file.bdptimg] [--stats] [--quiet]\n";
        return 2;
    }
    // --cache FILE: keep the scene image (flattened soup + both BVHs) beside the scene; a later run whose YAML and
    // OBJ files are unchanged maps it and skips OBJ parsing, flattening and the tree builds
    const char* cache = nullptr;
    for (int i = 2; i + 1 < argc; ++i) if (!std::strcmp(argv[i], "--cache")) cache = argv[i + 1];
    bdpt_scene_file* cached = nullptr;
    struct closer { bdpt_scene_file*& f; ~closer() { if (f) bdpt_scene_file_free(f); } } close_cached{cached};
    try {
        scene_load_result scene;
        if (cache) {
            const void* image = nullptr;
            uint64_t bytes = 0;
            int hit = 0;
            if (bdpt_scene_file_open_cached(argv[1], cache, &cached, &image, &bytes, &hit)) throw std::runtime_error(bdpt_host_last_error());
            scene.cam = load_camera_from_yaml(argv[1]);
            scene.cam.options.scene_image = image;
            scene.cam.options.scene_image_bytes = bytes;
            std::clog << "Scene image " << cache << (hit ? ": cache hit, " : ": built and written, ") << bytes / 1e6 << " MB\n";
        } else scene = load_scene_from_yaml(argv[1]);
        for (int i = 2; i < argc; ++i) {
            auto next = [&](const char* flag) -> const char* {
                if (i + 1 >= argc) throw std::runtime_error(std::string("missing value for ") + flag);
                return argv[++i];
            };
            if (!std::strcmp(argv[i], "--res")) {
                int w = 0, h = 0;
                if (std::sscanf(next("--res"), "%dx%d", &w, &h) != 2 || w <= 0 || h <= 0) throw std::runtime_error("--res wants WxH");
                scene.cam.image_width = w; scene.cam.aspect_ratio = double(w) / double(h);
            } else if (!std::strcmp(argv[i], "--spp")) scene.cam.samples_per_pixel = std::atoi(next("--spp"));
            else if (!std::strcmp(argv[i], "--depth")) scene.cam.max_depth = std::atoi(next("--depth"));
            else if (!std::strcmp(argv[i], "--gpus")) scene.cam.options.gpus = std::atoi(next("--gpus"));
            else if (!std::strcmp(argv[i], "--seed")) scene.cam.options.seed = std::strtoull(next("--seed"), nullptr, 0);
            else if (!std::strcmp(argv[i], "--out")) scene.cam.file_name = next("--out");
            else if (!std::strcmp(argv[i], "--cache")) next("--cache");
            else if (!std::strcmp(argv[i], "--stats")) scene.cam.options.stats = true;     // count the reference's BvhStats (slower)
            else if (!std::strcmp(argv[i], "--quiet")) scene.cam.options.quiet = true;
            else if (!std::strcmp(argv[i], "--mis")) scene.cam.options.weighting = BDPT_WEIGHT_MIS;   // the unbiased estimator; default: the reference's sum
            else if (!std::strcmp(argv[i], "--bvh")) scene.cam.options.bvh = std::strcmp(next("--bvh"), "ref") ? BDPT_BVH_FAST : BDPT_BVH_REFERENCE;
            else throw std::runtime_error(std::string("unknown flag ") + argv[i]);
        }
        auto world_bvh = make_shared<bvh_node>(scene.world.to_hittable_list());           // main.cpp:66
        const triangle_collection& lights = scene.lights.empty() ? scene.world : scene.lights;   // main.cpp:67
        scene.cam.render(*world_bvh, lights);                                             // main.cpp:68
    } catch (const std::exception& e) {
        std::cerr << "Failed to load scene: " << e.what() << "\n";                        // main.cpp:72-75
        return 1;
    }
    return 0;
}
