// C-ABI access to the host front-end (include/bdpt.h "scene files" section): lets a non-C++ caller
// (tests, bench.py) use the same YAML/OBJ loader and PNG writer as the CLI.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#include "scene.h"

struct bdpt_scene_file {
    scene_load_result scene;
    std::vector<double> verts;
    std::vector<uint32_t> mat_id;
    std::vector<bdpt_material> mats;
    std::vector<bdpt_texture> textures;
    std::vector<unsigned char> tex_data;
    std::vector<bdpt_sphere> spheres;
    // scene-image cache (bdpt_scene_file_open_cached)
    const void* image = nullptr;
    uint64_t image_bytes = 0;
    bool image_mapped = false;     // else allocated by bdpt_scene_image_build
    bool from_cache = false;
    bdpt_scene_image_info_t image_info{};
    ~bdpt_scene_file() {
        if (!image) return;
        if (image_mapped) bdpt_scene_image_unmap(image, image_bytes);
        else bdpt_scene_image_free(const_cast<void*>(image));
    }
};

static thread_local std::string g_host_err;

extern "C" {

const char* bdpt_host_last_error(void) { return g_host_err.c_str(); }

int bdpt_scene_file_load(const char* yaml_path, bdpt_scene_file** out) {
    if (!yaml_path || !out) return BDPT_ERR_ARG;
    *out = nullptr;
    try {
        auto* f = new bdpt_scene_file;
        f->scene = load_scene_from_yaml(yaml_path);
        flatten_triangles(f->scene.world.tris, f->verts, f->mat_id, f->mats);
        flatten_textures(f->scene.world.tris, f->mats, f->textures, f->tex_data);
        flatten_spheres(f->scene.cam.options.spheres, f->scene.world.tris, f->mats, f->spheres);   // may append their materials
        *out = f;
    } catch (const std::exception& e) {
        g_host_err = e.what();
        return BDPT_ERR_ARG;
    }
    return BDPT_OK;
}

void bdpt_scene_file_free(bdpt_scene_file* f) { delete f; }

int bdpt_scene_file_textures(const bdpt_scene_file* f, uint32_t* n_tex, bdpt_texture* tex, uint64_t* data_bytes, void* data) {
    if (!f) return BDPT_ERR_ARG;
    if (n_tex) *n_tex = uint32_t(f->textures.size());
    if (data_bytes) *data_bytes = f->tex_data.size();
    if (tex && !f->textures.empty()) std::memcpy(tex, f->textures.data(), f->textures.size() * sizeof(bdpt_texture));
    if (data && !f->tex_data.empty()) std::memcpy(data, f->tex_data.data(), f->tex_data.size());
    return BDPT_OK;
}

int bdpt_scene_file_spheres(const bdpt_scene_file* f, uint32_t* n_spheres, bdpt_sphere* spheres) {
    if (!f) return BDPT_ERR_ARG;
    if (n_spheres) *n_spheres = uint32_t(f->spheres.size());
    if (spheres && !f->spheres.empty()) std::memcpy(spheres, f->spheres.data(), f->spheres.size() * sizeof(bdpt_sphere));
    return BDPT_OK;
}

int bdpt_scene_file_media(const bdpt_scene_file* f, uint32_t* n_media, bdpt_medium* media, uint64_t* n_boundary_tris, double* boundary_verts) {
    if (!f) return BDPT_ERR_ARG;
    try {
        std::vector<bdpt_medium> d;
        std::vector<double> v;
        flatten_media(f->scene.cam.options.media, d, v);
        if (n_media) *n_media = uint32_t(d.size());
        if (n_boundary_tris) *n_boundary_tris = v.size() / 9;
        if (media && !d.empty()) std::memcpy(media, d.data(), d.size() * sizeof(bdpt_medium));
        if (boundary_verts && !v.empty()) std::memcpy(boundary_verts, v.data(), v.size() * sizeof(double));
    } catch (const std::exception& e) { g_host_err = e.what(); return BDPT_ERR_ARG; }
    return BDPT_OK;
}

int bdpt_scene_file_info(const bdpt_scene_file* f, uint64_t* n_tris, uint32_t* n_mats, uint64_t* n_lights,
                         bdpt_camera* cam, char* output_name, uint64_t output_cap) {
    if (!f) return BDPT_ERR_ARG;
    if (n_tris) *n_tris = f->from_cache ? f->image_info.n_tris : f->scene.world.size();
    if (n_mats) *n_mats = f->from_cache ? f->image_info.n_mats : uint32_t(f->mats.size());
    if (n_lights) *n_lights = f->from_cache ? f->image_info.n_lights : f->scene.lights.size();
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
    if (f->from_cache) { g_host_err = "scene was opened from its image cache: there is no triangle soup to copy"; return BDPT_ERR_STATE; }
    if (verts) std::memcpy(verts, f->verts.data(), f->verts.size() * sizeof(double));
    if (mat_id) std::memcpy(mat_id, f->mat_id.data(), f->mat_id.size() * sizeof(uint32_t));
    if (mats) std::memcpy(mats, f->mats.data(), f->mats.size() * sizeof(bdpt_material));
    return BDPT_OK;
}

int bdpt_write_png(const char* path, const uint8_t* rgb, int width, int height) {
    try { write_png_path(path, rgb, width, height); } catch (const std::exception& e) { g_host_err = e.what(); return BDPT_ERR_ARG; }
    return BDPT_OK;
}

int bdpt_scene_image_save(const char* path, const void* image, uint64_t bytes) {
    if (!path || !image || !bytes) return BDPT_ERR_ARG;
    // A temporary file of this call's own: several ranks or processes that miss the same cache at once must not
    // truncate each other's half-written file (a reader maps with MAP_POPULATE and would take a SIGBUS).  The last
    // rename wins; every one of them publishes a complete image of the same sources.
    static std::atomic<unsigned> serial{0};
    const std::string tmp = std::string(path) + "." + std::to_string(long(::getpid())) + "." + std::to_string(serial.fetch_add(1)) + ".tmp";
    const int fd = ::open(tmp.c_str(), O_WRONLY | O_CREAT | O_EXCL, 0644);
    if (fd < 0) { g_host_err = "cannot write " + tmp; return BDPT_ERR_ARG; }
    const unsigned char* p = static_cast<const unsigned char*>(image);
    uint64_t left = bytes;
    bool ok = true;
    while (left && ok) {
        const ssize_t w = ::write(fd, p, size_t(std::min<uint64_t>(left, uint64_t(1) << 30)));
        if (w <= 0) ok = false; else { p += w; left -= uint64_t(w); }
    }
    if (ok && ::fsync(fd) != 0) ok = false;
    if (::close(fd) != 0) ok = false;
    if (!ok || std::rename(tmp.c_str(), path) != 0) {
        std::remove(tmp.c_str());
        g_host_err = std::string("cannot write ") + path;
        return BDPT_ERR_ARG;
    }
    return BDPT_OK;
}

int bdpt_scene_image_map(const char* path, const void** image, uint64_t* bytes) {
    if (!path || !image || !bytes) return BDPT_ERR_ARG;
    *image = nullptr; *bytes = 0;
    const int fd = ::open(path, O_RDONLY);
    if (fd < 0) { g_host_err = std::string("cannot open ") + path; return BDPT_ERR_ARG; }
    struct stat st;
    if (::fstat(fd, &st) != 0 || st.st_size <= 0) { ::close(fd); g_host_err = std::string("cannot stat ") + path; return BDPT_ERR_ARG; }
    void* p = ::mmap(nullptr, size_t(st.st_size), PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
    ::close(fd);
    if (p == MAP_FAILED) { g_host_err = std::string("cannot map ") + path; return BDPT_ERR_NOMEM; }
    *image = p; *bytes = uint64_t(st.st_size);
    return BDPT_OK;
}

void bdpt_scene_image_unmap(const void* image, uint64_t bytes) {
    if (image && bytes) ::munmap(const_cast<void*>(image), size_t(bytes));
}

int bdpt_scene_source_key(const char* yaml_path, uint64_t* key) {
    if (!yaml_path || !key) return BDPT_ERR_ARG;
    try { *key = scene_source_key(yaml_path); } catch (const std::exception& e) { g_host_err = e.what(); return BDPT_ERR_ARG; }
    return BDPT_OK;
}

int bdpt_scene_file_open_cached(const char* yaml_path, const char* cache_path, bdpt_scene_file** out, const void** image,
                                uint64_t* image_bytes, int* hit) {
    if (!yaml_path || !cache_path || !out || !image || !image_bytes) return BDPT_ERR_ARG;
    *out = nullptr; *image = nullptr; *image_bytes = 0;
    if (hit) *hit = 0;
    bdpt_scene_file* f = nullptr;
    try {
        const uint64_t key = scene_source_key(yaml_path);
        f = new bdpt_scene_file;
        const void* img = nullptr;
        uint64_t n = 0;
        if (bdpt_scene_image_map(cache_path, &img, &n) == BDPT_OK) {
            // a stale, foreign or damaged file is not an error: the cache is rebuilt below
            if (bdpt_scene_image_info(img, n, 1, &f->image_info) == BDPT_OK && f->image_info.source_key == key) {
                f->image = img; f->image_bytes = n; f->image_mapped = true; f->from_cache = true;
                f->scene.cam = load_camera_from_yaml(yaml_path);
            } else bdpt_scene_image_unmap(img, n);
        }
        if (!f->from_cache) {
            f->scene = load_scene_from_yaml(yaml_path);
            flatten_triangles(f->scene.world.tris, f->verts, f->mat_id, f->mats);
            void* built = nullptr;
            if (bdpt_scene_image_build(f->verts.data(), f->mat_id.data(), f->mat_id.size(), f->mats.data(), uint32_t(f->mats.size()), key,
                                       &built, &n) != BDPT_OK)
                throw std::runtime_error(bdpt_scene_image_last_error());
            f->image = built; f->image_bytes = n; f->image_mapped = false;
            bdpt_scene_image_info(built, n, 0, &f->image_info);
            if (bdpt_scene_image_save(cache_path, built, n) != BDPT_OK)
                std::fprintf(stderr, "bdpt: scene image not cached: %s\n", g_host_err.c_str());   // rendering goes on without a cache file
        }
    } catch (const std::exception& e) {
        delete f;
        g_host_err = e.what();
        return BDPT_ERR_ARG;
    }
    *out = f; *image = f->image; *image_bytes = f->image_bytes;
    if (hit) *hit = f->from_cache ? 1 : 0;
    return BDPT_OK;
}

}  // extern "C"
