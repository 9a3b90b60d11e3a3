// Host-side object model of the drop-in surface: the names a user of the reference touches in
// main.cpp (src/main.cpp:62-78) — vec3/point3/color, material classes, triangle, triangle_collection,
// hittable_list, bvh_node, camera, scene_load_result — kept source-compatible, but as plain descriptors:
// shading, intersection and the BVH live on the GPU behind include/bdpt.h.
//   reference                                   here
//   material + 4 subclasses (material.h:16-147)  `material` with a kind tag; same constructors
//   triangle (triangle.h:19-133)                 3 points + material handle
//   triangle_collection (triangle.h:135-230)     std::vector<triangle>, add/size/to_hittable_list
//   bvh_node(hittable_list) (bvh.h:13)           records the primitive list; the trees are built at upload
//   camera (camera.h:23-477)                     same public fields; render() drives libbdpt
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <ostream>
#include <string>
#include <vector>

#include "../../include/bdpt.h"

using std::make_shared;
using std::shared_ptr;

class vec3 {
  public:
    double e[3];
    vec3() : e{0, 0, 0} {}
    vec3(double x, double y, double z) : e{x, y, z} {}
    double x() const { return e[0]; }
    double y() const { return e[1]; }
    double z() const { return e[2]; }
    double operator[](int i) const { return e[i]; }
    double& operator[](int i) { return e[i]; }
    double length_squared() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    double length() const { return std::sqrt(length_squared()); }
};
using point3 = vec3;
using color = vec3;

// texture.h:7-87 as descriptors (the arithmetic of value() runs on the GPU: csrc/texture.cuh).  noise_texture draws its
// perlin tables on the host exactly as perlin.h:6-14,68-75 does, from a default-seeded mt19937 of its own standing in for
// main.h:28-32's (so the first noise texture of a scene gets the tables the reference would give it); image_texture reads
// binary PPM (P6) files and applies rtw_image's float round trip (rtw_stb_image.h:45-60,90-112 over stb's gamma 2.2).
class texture {
  public:
    enum kind_t { Solid = BDPT_TEX_SOLID, Checker = BDPT_TEX_CHECKER, Noise = BDPT_TEX_NOISE, Image = BDPT_TEX_IMAGE };
    virtual ~texture() = default;
    kind_t kind = Solid;
    color a, b;                       // solid / checker even, checker odd
    double scale = 1.0;
    int width = 0, height = 0;        // image
    std::vector<unsigned char> data;  // noise: 768 doubles + 768 int32; image: width * height * 3 bytes
};
class solid_color : public texture {
  public:
    explicit solid_color(const color& albedo) { kind = Solid; a = albedo; }
    solid_color(double r, double g, double bl) : solid_color(color(r, g, bl)) {}
};
class checker_texture : public texture {   // texture.h:28-49 (the two-colour constructor, :34-35)
  public:
    checker_texture(double scale_, const color& c1, const color& c2) { kind = Checker; scale = scale_; a = c1; b = c2; }
};
class noise_texture : public texture {     // texture.h:75-85
  public:
    explicit noise_texture(double scale_);
};
class image_texture : public texture {     // texture.h:51-73
  public:
    explicit image_texture(const char* filename);
};

class material {
  public:
    enum kind_t { Lambertian = BDPT_MAT_LAMBERTIAN, Metal = BDPT_MAT_METAL, Dielectric = BDPT_MAT_DIELECTRIC,
                  DiffuseLight = BDPT_MAT_LIGHT };
    virtual ~material() = default;
    kind_t kind;
    color rgb;       // albedo or emission
    double param;    // fuzz or refraction index
    shared_ptr<texture> tex;   // lambertian albedo / light emission through a texture (material.h:45,136), else null
    bool phase = false;        // isotropic: only meaningful inside a constant_medium (flatten_triangles refuses it on a triangle)
    virtual bool is_delta() const { return kind == Metal || kind == Dielectric; }   // material.h:38,85,118
  protected:
    material(kind_t k, const color& c, double p) : kind(k), rgb(c), param(p) {}
};
class lambertian : public material {   // material.h:42-67
  public:
    explicit lambertian(const color& albedo) : material(Lambertian, albedo, 0.0) {}
    explicit lambertian(shared_ptr<texture> t) : material(Lambertian, color(0.5, 0.5, 0.5), 0.0) { tex = std::move(t); }   // material.h:45
};
class metal : public material {        // material.h:69-90 (fuzz clamped to <= 1, :71)
  public:
    metal(const color& albedo, double fuzz) : material(Metal, albedo, fuzz < 1 ? fuzz : 1) {}
};
class dielectric : public material {   // material.h:92-131
  public:
    explicit dielectric(double refraction_index) : material(Dielectric, color(1, 1, 1), refraction_index) {}
};
class diffuse_light : public material {  // material.h:133-147
  public:
    explicit diffuse_light(const color& emit) : material(DiffuseLight, emit, 0.0) {}
    explicit diffuse_light(shared_ptr<texture> t) : material(DiffuseLight, color(1, 1, 1), 0.0) { tex = std::move(t); }   // material.h:135
};

class isotropic : public material {      // material.h:149-171: the phase function of a constant_medium; not a surface material here
  public:
    explicit isotropic(const color& albedo) : material(Lambertian, albedo, 0.0) { phase = true; }
};

class hittable {
  public:
    virtual ~hittable() = default;
};

class triangle : public hittable {
  public:
    triangle(const point3& a, const point3& b, const point3& c, shared_ptr<material> m) : v0(a), v1(b), v2(c), mat(std::move(m)) {}
    point3 v0, v1, v2;
    shared_ptr<material> mat;
};

class hittable_list : public hittable {
  public:
    std::vector<shared_ptr<hittable>> objects;
    void add(shared_ptr<hittable> o) { objects.push_back(std::move(o)); }
    void clear() { objects.clear(); }
};

class triangle_collection : public hittable {
  public:
    std::vector<triangle> tris;
    void add(const triangle& t) { tris.push_back(t); }
    void clear() { tris.clear(); }
    size_t size() const { return tris.size(); }
    bool empty() const { return tris.empty(); }
    hittable_list to_hittable_list() const {   // triangle.h:191
        hittable_list l;
        l.objects.reserve(tris.size());
        for (const auto& t : tris) l.add(make_shared<triangle>(t));
        return l;
    }
};

// Instancing (hittable.h:46-157).  The reference moves the RAY into the object's frame at every hit query; here the
// wrapped object's triangles are moved into the world frame ONCE, when the scene is flattened (collect_triangles): the
// device walks one tree of world-space triangles and pays nothing per ray.  Same geometry, not the same rounding: hit
// distances agree with the reference's ray-transform form to ~1e-12 relative, prim IDs except on knife edges
// (tests/test_instancing.py).  translate(rotate_y(obj, deg), offset) is the usual order.
class translate : public hittable {      // hittable.h:46-77
  public:
    translate(shared_ptr<hittable> obj, const vec3& off) : object(std::move(obj)), offset(off) {}
    shared_ptr<hittable> object;
    vec3 offset;
};
class rotate_y : public hittable {       // hittable.h:79-157 (angle in degrees, counter-clockwise seen from +y)
  public:
    rotate_y(shared_ptr<hittable> obj, double angle_degrees);
    shared_ptr<hittable> object;
    double sin_theta, cos_theta;
};
// constant_medium (materials/volumes/constant_medium.h:8-61): a closed (convex) boundary, a density, the colour of its
// isotropic phase function.  On the device: bdpt_scene_set_media (include/bdpt.h) -- the boundary's triangles are NOT
// world prims.  Put one into the world list (hittable_list / bvh_node) like the reference does, or into
// camera::options.media (what the YAML loader's `media:` section does).
class constant_medium : public hittable {
  public:
    constant_medium(shared_ptr<hittable> boundary_, double density_, const color& albedo_)
        : boundary(std::move(boundary_)), density(density_), albedo(albedo_) {}
    shared_ptr<hittable> boundary;
    double density;
    color albedo;
};
// An analytic sphere prim (include/bdpt.h: bdpt_scene_set_spheres).  NOT a reference class: the reference's loader tessellates its
// `Sphere` surfaces into 960 triangles (scene_loader.h:212-242), and so does this one unless the surface says `analytic: true`.
class sphere : public hittable {
  public:
    sphere(const point3& c, double r, shared_ptr<material> m) : center(c), radius(r), mat(std::move(m)) {}
    point3 center;
    double radius;
    shared_ptr<material> mat;
};
// the triangles of any object graph (triangle, triangle_collection, hittable_list, bvh_node, translate, rotate_y), in
// traversal order (= prim ID), in world space; throws on a hittable it does not know.  constant_medium and sphere objects met on
// the way go to `media` / `spheres` (with the transforms above them applied) when those are given, else they are an error.
void collect_triangles(const hittable& h, std::vector<class triangle>& out, std::vector<constant_medium>* media = nullptr,
                       std::vector<sphere>* spheres = nullptr);
// bdpt_scene_set_spheres' argument: the spheres' materials are looked up in (or appended to) the table flatten_triangles made of `tris`
void flatten_spheres(const std::vector<sphere>& spheres, const std::vector<class triangle>& tris, std::vector<bdpt_material>& mats,
                     std::vector<bdpt_sphere>& out);
// bdpt_scene_set_media's arguments from a list of media (boundaries flattened with collect_triangles)
void flatten_media(const std::vector<constant_medium>& media, std::vector<bdpt_medium>& out, std::vector<double>& boundary_verts);

// bvh_node(hittable_list) (bvh.h:13-18): here it only records the primitives, in list order (= prim ID);
// both trees (reference topology and SAH) are built inside bdpt_scene_upload.
class bvh_node : public hittable {
  public:
    explicit bvh_node(hittable_list list) {
        prims.reserve(list.objects.size());
        for (auto& o : list.objects) collect_triangles(*o, prims, &media, &spheres);   // triangles as they are; instances baked (above)
    }
    std::vector<triangle> prims;
    std::vector<constant_medium> media;   // constant_medium objects of the list
    std::vector<sphere> spheres;          // analytic sphere objects of the list
};

struct render_options {         // overrides the CLI adds on top of the YAML
    int gpus = 1;               // strata are sharded over this many devices
    int bvh = BDPT_BVH_FAST;    // both give the reference's hits; REFERENCE also walks its topology
    int weighting = BDPT_WEIGHT_REF_SUM;   // REF_SUM: the reference's estimator (camera.h:294-323); BDPT_WEIGHT_MIS: balance-heuristic weights + light-tracing splats
    uint64_t seed = 0;
    bool quiet = false;
    bool stats = false;         // count the reference's BvhStats (stats.h:8-16) during the render (BDPT_RENDER_STATS: slower)
    bool write_file = true;     // write ./output/<file_name> like camera.h:139-142
    // a prebuilt scene image (include/bdpt.h, bdpt_scene_image_build / _map): when set, render() uploads it instead
    // of flattening `world` and building the trees, and `world` may be empty
    const void* scene_image = nullptr;
    uint64_t scene_image_bytes = 0;
    // participating media beside the world (the YAML loader's `media:` section puts them here; a constant_medium inside the
    // world list is found by render() itself)
    std::vector<constant_medium> media;
    // analytic sphere prims beside the world (`analytic: true` on a YAML Sphere surface); a sphere inside the world list is found
    // by render() itself
    std::vector<sphere> spheres;
};

struct render_report {
    double device_ms = 0, wall_s = 0;
    uint64_t samples = 0, rays_closest = 0, rays_shadow = 0, launches = 0;
    // stats.h:8-16, summed over the GPUs (all but rays_traced / bvh_nodes_built are 0 unless options.stats)
    uint64_t aabb_tests = 0, aabb_hits = 0, bvh_node_visits = 0, triangle_tests = 0, triangle_hits = 0, bvh_nodes_built = 0;
    double film_reduce_ms = 0;  // several GPUs: the NCCL sum-reduce of the film onto GPU 0
    int width = 0, height = 0;
    std::vector<double> film;   // W*H*3 fp64 SUM over strata (camera.h:55)
};

class camera {
  public:
    // camera.h:26-41
    double aspect_ratio = 1.0;
    int image_width = 100;
    int samples_per_pixel = 50;
    int max_depth = 10;
    color background = color(0, 0, 0);
    double vfov = 90;
    point3 lookfrom = point3(0, 0, 0);
    point3 lookat = point3(0, 0, -1);
    vec3 vup = vec3(0, 1, 0);
    double defocus_angle = 0;
    double focus_dist = 10;
    int progress_bar_length = 30;   // camera.h:39: width of the progress bar on std::clog
    std::string file_name = "image.png";

    render_options options;     // not in the reference: device-side knobs
    render_report last_report;

    // camera.h:43. `world` may be a bvh_node (as main.cpp:66 builds) or a triangle_collection; `lights`
    // is accepted for signature compatibility — the device derives the light list from the materials, in
    // world order, which is what scene_loader.h:190-202 / main.cpp:67 produce.
    void render(const hittable& world, const hittable& lights);
};

struct scene_load_result {      // scene_loader.h:22-26
    camera cam;
    triangle_collection world;
    triangle_collection lights;
};

scene_load_result load_scene_from_yaml(const std::string& path);   // scene_loader.h:480-523
// Scene-image cache support (include/bdpt.h; not in the reference): the camera section alone, and a
// key over the bytes of the YAML and of every OBJ file its `object` surfaces name, in scene order.
camera load_camera_from_yaml(const std::string& path);
uint64_t scene_source_key(const std::string& path);

// stats.h:34-50: the block camera::render prints on std::clog when it is done (camera.h:144)
void print_bvh_stats(std::ostream& os, const render_report& r);

// image output, wpng.h:14-89
void colors_to_rgb8(const std::vector<color>& framebuffer, int image_width, int image_height, int samples_per_pixel,
                    std::vector<uint8_t>& out_rgb);
void write_png(const std::string& file_name, const std::vector<uint8_t>& rgb, int image_width, int image_height);
void write_png_path(const std::string& path, const uint8_t* rgb, int image_width, int image_height);

// flatten a primitive list to the C-ABI's soup (prim ID = list order); material table dedups by pointer
void flatten_triangles(const std::vector<triangle>& tris, std::vector<double>& verts, std::vector<uint32_t>& mat_id,
                       std::vector<bdpt_material>& mats);
// the texture table of a flattened scene for bdpt_scene_set_textures; sets bdpt_material.tex of `mats` (same order as
// flatten_triangles made them)
void flatten_textures(const std::vector<triangle>& tris, std::vector<bdpt_material>& mats, std::vector<bdpt_texture>& tex,
                      std::vector<unsigned char>& data);
