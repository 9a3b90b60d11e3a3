// C-ABI access to the host front-end (include/bdpt.h "scene files" section): lets a non-C++ caller
// (tests, bench.py) use the same YAML/OBJ loader and PNG writer as the CLI.
#include <cstring>
#include <string>

#include "scene.h"

struct bdpt_scene_file {
    scene_load_result scene;
    std::vector<double> verts;
    std::vector<uint32_t> mat_id;
    std::vector<bdpt_material> mats;
};

static thread_local std::string g_host_err;

extern "C" {

const char* bdpt_host_last_error(void) { return g_host_err.c_str(); }

int bdpt_scene_file_load(const char* yaml_path, bdpt_scene_file** out) {
    if (!yaml_path || !out) return BDPT_ERR_ARG;
    *out = nullptr;
    try {
        auto* f = new bdpt_scene_file{load_scene_from_yaml(yaml_path), {}, {}, {}};
        flatten_triangles(f->scene.world.tris, f->verts, f->mat_id, f->mats);
        *out = f;
    } catch (const std::exception& e) {
        g_host_err = e.what();
        return BDPT_ERR_ARG;
    }
    return BDPT_OK;
}

void bdpt_scene_file_free(bdpt_scene_file* f) { delete f; }

int bdpt_scene_file_info(const bdpt_scene_file* f, uint64_t* n_tris, uint32_t* n_mats, uint64_t* n_lights,
                         bdpt_camera* cam, char* output_name, uint64_t output_cap) {
    if (!f) return BDPT_ERR_ARG;
    if (n_tris) *n_tris = f->scene.world.size();
    if (n_mats) *n_mats = uint32_t(f->mats.size());
    if (n_lights) *n_lights = f->scene.lights.size();
    if (cam) {
        const camera& c = f->scene.cam;
        std::memset(cam, 0, sizeof *cam);
        cam->aspect_ratio = c.aspect_ratio; cam->image_width = c.image_width; cam->samples_per_pixel = c.samples_per_pixel;
        cam->max_depth = c.max_depth; cam->vfov = c.vfov; cam->defocus_angle = c.defocus_angle; cam->focus_dist = c.focus_dist;
        for (int k = 0; k < 3; ++k) {
            cam->background[k] = c.background[k]; cam->lookfrom[k] = c.lookfrom[k];
            cam->lookat[k] = c.lookat[k]; cam->vup[k] = c.vup[k];
        }
    }
    if (output_name && output_cap) {
        std::strncpy(output_name, f->scene.cam.file_name.c_str(), output_cap - 1);
        output_name[output_cap - 1] = '\0';
    }
    return BDPT_OK;
}

int bdpt_scene_file_copy(const bdpt_scene_file* f, double* verts, uint32_t* mat_id, bdpt_material* mats) {
    if (!f) return BDPT_ERR_ARG;
    if (verts) std::memcpy(verts, f->verts.data(), f->verts.size() * sizeof(double));
    if (mat_id) std::memcpy(mat_id, f->mat_id.data(), f->mat_id.size() * sizeof(uint32_t));
    if (mats) std::memcpy(mats, f->mats.data(), f->mats.size() * sizeof(bdpt_material));
    return BDPT_OK;
}

int bdpt_write_png(const char* path, const uint8_t* rgb, int width, int height) {
    try { write_png_path(path, rgb, width, height); } catch (const std::exception& e) { g_host_err = e.what(); return BDPT_ERR_ARG; }
    return BDPT_OK;
}

}  // extern "C"
