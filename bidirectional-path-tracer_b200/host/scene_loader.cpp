// YAML + OBJ scene front-end: behaviour of the reference's src/scene/scene_loader.h (schema in SURVEY
// Appendix B), written against yaml_lite instead of yaml-cpp.  Not on the timed path.
#include <algorithm>
#include <random>
#include <array>
#include <charconv>
#include <future>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>
#include <unordered_map>

#include "scene.h"
#include "yaml_lite.h"

namespace {

using yaml_lite::node;
using material_map = std::unordered_map<std::string, shared_ptr<material>>;

// scalar helpers with the reference's "default on anything odd" behaviour (scene_loader.h:32-72)
std::string to_string(const node& n, const std::string& def = {}) { return n.is_scalar() ? n.scalar : def; }
double to_double(const node& n, double def = 0.0) { double v; return n.as_double(v) ? v : def; }
int to_int(const node& n, int def = 0) { int v; return n.as_int(v) ? v : def; }
std::vector<double> to_double_list(const node& n) {
    std::vector<double> out;
    if (!n.is_sequence()) return out;
    out.reserve(n.seq.size());
    for (const auto& e : n.seq) out.push_back(to_double(e, 0.0));
    return out;
}
color read_color(const node& n, const color& fallback) {            // scene_loader.h:74-79
    const auto v = to_double_list(n);
    return v.size() < 3 ? fallback : color(v[0], v[1], v[2]);
}
color read_color_scaled(const node& n, const color& fallback) {     // scene_loader.h:81-92: 0-255 -> 0-1
    const auto v = to_double_list(n);
    if (v.size() < 3) return fallback;
    double r = v[0], g = v[1], b = v[2];
    const double maxc = std::max({std::fabs(r), std::fabs(g), std::fabs(b)});
    if (maxc > 1.0 && maxc <= 255.0) { const double s = 1.0 / 255.0; r *= s; g *= s; b *= s; }
    return color(r, g, b);
}

// Extension of the reference's schema (its loader only ever builds solid colours): a `texture:` map inside a
// lambertian or light material -- {type: checker, scale: s, even: [r,g,b], odd: [r,g,b]} | {type: noise, scale: s} |
// {type: image, file: picture.ppm (relative to the YAML)} -- builds the reference's checker_texture / noise_texture /
// image_texture (texture.h:28-87) and the material's texture constructor (material.h:45,135).
std::string g_scene_dir;
void reset_scene_rng();   // a scene load starts at the head of the random stream, like a fresh run of the reference
shared_ptr<texture> build_texture(const node& n, bool emission) {
    if (!n.is_map()) return nullptr;
    const std::string type = to_string(n["type"], "");
    // colours follow the material they sit in: albedo 0-255 -> 0-1 (scene_loader.h:81-92), emission linear HDR (:123-125)
    auto col = [&](const node& c, const color& fallback) { return emission ? read_color(c, fallback) : read_color_scaled(c, fallback); };
    if (type == "checker")
        return make_shared<checker_texture>(to_double(n["scale"], 1.0), col(n["even"], color(1, 1, 1)), col(n["odd"], color(0, 0, 0)));
    if (type == "noise") return make_shared<noise_texture>(to_double(n["scale"], 1.0));
    if (type == "image") {
        std::string file = to_string(n["file"], "");
        if (!file.empty() && file[0] != '/' && !g_scene_dir.empty()) file = g_scene_dir + "/" + file;
        return make_shared<image_texture>(file.c_str());
    }
    if (type == "solid") return make_shared<solid_color>(col(n["color"], color(0, 0, 0)));
    throw std::runtime_error("unknown texture type '" + type + "'");
}

shared_ptr<material> build_material(const node& n) {                 // scene_loader.h:101-170
    if (!n.is_map()) throw std::runtime_error("Material must be a mapping");
    const color black(0, 0, 0);
    const std::string type = to_string(n["type"], "");
    if (!type.empty() && n["texture"].defined()) {
        if (type == "lambertian") return make_shared<lambertian>(build_texture(n["texture"], false));
        if (type == "light" || type == "diffuse_light") return make_shared<diffuse_light>(build_texture(n["texture"], true));
    }
    if (!type.empty()) {
        const color c = read_color_scaled(n["color"], read_color_scaled(n["albedo"],
                        read_color_scaled(n["base_color"], read_color_scaled(n["base_colour"], black))));
        if (type == "light" || type == "diffuse_light")
            return make_shared<diffuse_light>(read_color(n["emission"], black));   // linear HDR, never /255
        if (type == "lambertian") return make_shared<lambertian>(c);
        if (type == "metal") return make_shared<metal>(c, std::clamp(to_double(n["roughness"], 0.0), 0.0, 1.0));
        if (type == "dielectric" || type == "glass") {
            const double ior = to_double(n["ior"], 1.5);
            return make_shared<dielectric>(ior > 0.0 ? ior : 1.5);
        }
    }
    // legacy PBR-ish keys (scene_loader.h:138-169)
    color base = read_color_scaled(n["base_color"], black);
    if (n["base_colour"].defined()) base = read_color_scaled(n["base_colour"], base);
    color emission = read_color_scaled(n["emission"], black);
    if (emission.length_squared() > 0.0) {
        const double maxc = std::max({std::fabs(emission.x()), std::fabs(emission.y()), std::fabs(emission.z())});
        if (maxc > 50.0) { const double s = 50.0 / maxc; emission = color(emission.x() * s, emission.y() * s, emission.z() * s); }
        return make_shared<diffuse_light>(emission);
    }
    double transmission = to_double(n["transmission"], 0.0);
    if (transmission == 0.0) transmission = to_double(n["spec_trans"], 0.0);
    const double ior = to_double(n["ior"], 1.5);
    if (transmission > 0.0) return make_shared<dielectric>(ior > 0.0 ? ior : 1.5);
    const double metallic = to_double(n["metallic"], 0.0);
    const double roughness = std::clamp(to_double(n["roughness"], 0.0), 0.0, 1.0);
    if (metallic > 0.5) return make_shared<metal>(base, roughness);
    return make_shared<lambertian>(base);
}

material_map load_materials(const node& n) {                         // scene_loader.h:173-188
    material_map out;
    if (!n.is_map()) return out;
    for (const auto& kv : n.map) {
        try { out[kv.first] = build_material(kv.second); } catch (...) { /* bad entries are skipped */ }
    }
    return out;
}

struct sink {                                                        // scene_loader.h:190-202
    triangle_collection& world;
    triangle_collection& lights;
    void add(const point3& a, const point3& b, const point3& c, const shared_ptr<material>& m) {
        triangle t(a, b, c, m);
        world.add(t);
        if (m->kind == material::DiffuseLight) lights.add(t);
    }
};

shared_ptr<material> resolve_material(const node& m, const material_map& mats) {   // scene_loader.h:318-329,411-422
    shared_ptr<material> mat;
    if (m.is_scalar()) { auto it = mats.find(m.scalar); if (it != mats.end()) mat = it->second; }
    else if (m.is_map()) mat = build_material(m);
    if (!mat) mat = make_shared<lambertian>(color(0.8, 0.8, 0.8));
    return mat;
}

void load_tri_mesh(const node& mesh, sink& out) {                    // scene_loader.h:244-272
    const node& data = mesh["data"];
    if (!data.is_map()) throw std::runtime_error("Mesh missing data field");
    if (!data["vertices"].is_sequence()) throw std::runtime_error("Missing vertices");
    const auto v = to_double_list(data["vertices"]);
    if (v.size() % 9 != 0) throw std::runtime_error("Vertices length not a multiple of 9");
    if (!mesh["material"].defined()) throw std::runtime_error("Missing material field");
    const auto mat = build_material(mesh["material"]);
    for (size_t i = 0; i < v.size(); i += 9)
        out.add(point3(v[i], v[i + 1], v[i + 2]), point3(v[i + 3], v[i + 4], v[i + 5]), point3(v[i + 6], v[i + 7], v[i + 8]), mat);
}

point3 sphere_point(double theta, double phi, const point3& c, double r) {   // scene_loader.h:204-210
    const double st = std::sin(theta);
    const double x = r * st * std::cos(phi), y = r * std::cos(theta), z = r * st * std::sin(phi);
    return point3(c.x() + x, c.y() + y, c.z() + z);
}

// `analytic: true` (extension; SURVEY 8 "sphere row"): the sphere becomes ONE exact prim (bdpt_scene_set_spheres) instead of the
// reference's 960 triangles; it lands in `analytic` (cam.options.spheres)
void load_sphere(const node& mesh, sink& out, std::vector<sphere>& analytic) {   // scene_loader.h:212-242,274-294
    if (!mesh["material"].defined()) throw std::runtime_error("Missing material field");
    const auto mat = build_material(mesh["material"]);
    const node& data = mesh["data"];
    if (!data.is_map()) throw std::runtime_error("Missing data field");
    const auto cv = to_double_list(data["center"]);
    const point3 center = cv.size() < 3 ? point3(0, 0, 0) : point3(cv[0], cv[1], cv[2]);
    const double radius = to_double(data["radius"], 0.0);
    if (radius <= 0.0) throw std::runtime_error("Missing or invalid radius field");
    const std::string an = to_string(mesh["analytic"], "");
    if (an == "true" || an == "True" || an == "yes" || an == "1") {
        if (mat->kind == material::DiffuseLight) throw std::runtime_error("an analytic sphere cannot be a light (lights are sampled over triangles): drop `analytic`");
        analytic.emplace_back(center, radius, mat);
        return;
    }
    const int lat_steps = 16, lon_steps = 32;   // 960 triangles
    const double pi = 3.14159265358979323846;
    for (int lat = 0; lat < lat_steps; ++lat) {
        const double t0 = pi * (double(lat) / lat_steps), t1 = pi * (double(lat + 1) / lat_steps);
        for (int lon = 0; lon < lon_steps; ++lon) {
            const double p0 = 2.0 * pi * (double(lon) / lon_steps), p1 = 2.0 * pi * (double(lon + 1) / lon_steps);
            const point3 p00 = sphere_point(t0, p0, center, radius), p01 = sphere_point(t0, p1, center, radius);
            const point3 p10 = sphere_point(t1, p0, center, radius), p11 = sphere_point(t1, p1, center, radius);
            if (lat > 0) out.add(p00, p10, p11, mat);
            if (lat < lat_steps - 1) out.add(p00, p11, p01, mat);
        }
    }
}

void load_indexed_mesh(const node& mesh, sink& out, const material_map& mats) {   // scene_loader.h:297-343
    const node &vn = mesh["vertices"], &tn = mesh["triangles"];
    if (!vn.is_sequence()) throw std::runtime_error("Indexed mesh missing vertices");
    if (!tn.is_sequence()) throw std::runtime_error("Indexed mesh missing triangles");
    std::vector<point3> verts;
    verts.reserve(vn.seq.size());
    for (const auto& v : vn.seq) {
        const auto vals = to_double_list(v);
        if (vals.size() < 3) continue;
        verts.emplace_back(vals[0], vals[1], vals[2]);
    }
    const auto mat = resolve_material(mesh["material"], mats);
    for (const auto& tri : tn.seq) {
        std::vector<int> idx;
        if (tri.is_sequence()) for (const auto& t : tri.seq) idx.push_back(to_int(t, 0));
        if (idx.size() < 3) continue;
        out.add(verts.at(size_t(idx[0])), verts.at(size_t(idx[1])), verts.at(size_t(idx[2])), mat);   // 0-based; .at throws
    }
}

// OBJ: 'v' and 'f' lines only, 1-based and negative indices, fan triangulation (scene_loader.h:346-397).
// Tokenised by hand with std::from_chars (correctly rounded like the reference's stream extraction, a third of
// strtod's cost): the reference's istringstream-per-line parse is the slow part at 10^7 faces.
// A file is parsed into a plain vertex-triple list so that the files of a scene can be read side by side.
using tri_points = std::array<point3, 3>;

// strtod's contract on [p, eol): skips blanks, returns false when no number starts there
inline bool parse_double(const char*& p, const char* eol, double& out) {
    while (p < eol && (*p == ' ' || *p == '\t')) ++p;
    const char* q = p;
    if (q < eol && *q == '+') ++q;
    const auto r = std::from_chars(q, eol, out);
    if (r.ec == std::errc()) { p = r.ptr; return true; }
    if (r.ec == std::errc::result_out_of_range || (q < eol && (*q == 'i' || *q == 'I' || *q == 'n' || *q == 'N'))) {
        // overflow / underflow / inf / nan spellings: leave those to strtod (the line is NUL- or newline-terminated)
        char* e;
        out = std::strtod(p, &e);
        if (e != p && e <= eol) { p = e; return true; }
    }
    return false;
}

std::vector<tri_points> parse_obj_file(const std::filesystem::path& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("Failed to open OBJ file: " + path.string());
    std::string buf((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    std::vector<point3> verts;
    verts.reserve(buf.size() / 64);
    std::vector<tri_points> tris;
    tris.reserve(buf.size() / 48);
    std::vector<int> fidx;
    const char* p = buf.c_str();
    const char* end = p + buf.size();
    while (p < end) {
        const char* eol = static_cast<const char*>(std::memchr(p, '\n', size_t(end - p)));
        if (!eol) eol = end;
        const char* q = p;
        while (q < eol && (*q == ' ' || *q == '\t')) ++q;
        const bool sep = q + 1 < eol && (q[1] == ' ' || q[1] == '\t');
        if (q < eol && *q == 'v' && sep) {
            const char* t = q + 1;
            double x, y, z;
            if (parse_double(t, eol, x) && parse_double(t, eol, y) && parse_double(t, eol, z)) verts.emplace_back(x, y, z);
        } else if (q < eol && *q == 'f' && sep) {
            fidx.clear();
            const char* t = q + 1;
            while (t < eol) {
                while (t < eol && (*t == ' ' || *t == '\t' || *t == '\r')) ++t;
                if (t >= eol) break;
                long vi = 0;                                  // "vi", "vi/vt", "vi/vt/vn", "vi//vn": leading int
                const char* d = (*t == '+') ? t + 1 : t;
                const auto r = std::from_chars(d, eol, vi);
                if (r.ec == std::errc()) { fidx.push_back(vi > 0 ? int(vi - 1) : int(verts.size()) + int(vi)); t = r.ptr; }
                else ++t;
                while (t < eol && *t != ' ' && *t != '\t') ++t;   // skip the /vt/vn tail
            }
            for (size_t k = 2; k < fidx.size(); ++k)
                tris.push_back(tri_points{verts.at(size_t(fidx[0])), verts.at(size_t(fidx[k - 1])), verts.at(size_t(fidx[k]))});
        }
        p = eol + 1;
    }
    return tris;
}

using obj_jobs = std::unordered_map<const node*, std::future<std::vector<tri_points>>>;

std::filesystem::path object_path(const node& n, const std::filesystem::path& dir) {
    const std::string rel = to_string(n["file"]);
    if (rel.empty()) throw std::runtime_error("Object missing file field");
    return dir / rel;
}

void load_object(const node& n, const std::filesystem::path& dir, sink& out, const material_map& mats, obj_jobs& jobs) {   // :399-425
    const std::filesystem::path path = object_path(n, dir);
    auto it = jobs.find(&n);
    const std::vector<tri_points> tris = it != jobs.end() ? it->second.get() : parse_obj_file(path);   // get() rethrows the parse error here, in scene order
    const auto mat = resolve_material(n["material"], mats);
    for (const tri_points& t : tris) out.add(t[0], t[1], t[2], mat);
}

void load_camera(const node& n, camera& cam) {                       // scene_loader.h:427-476
    if (!n.is_map()) throw std::runtime_error("Camera section must be a mapping");
    const node& res = n["resolution"];
    if (!res.is_sequence() || res.size() < 2) throw std::runtime_error("Camera missing resolution");
    const auto rv = to_double_list(res);
    const int width = int(rv[0]), height = int(rv[1]);
    if (width <= 0 || height <= 0) throw std::runtime_error("Resolution values must be positive");
    cam.image_width = width;
    cam.aspect_ratio = double(width) / double(height);
    double vfov = cam.vfov;
    if (n["vfov"].defined() || n["fov"].defined()) vfov = to_double(n["vfov"], to_double(n["fov"], vfov));
    cam.vfov = std::clamp(vfov, 1.0, 179.0);
    cam.focus_dist = to_double(n["focus_distance"], cam.focus_dist);
    cam.defocus_angle = 0.0;                                          // aperture settings are ignored (:463)
    cam.lookfrom = read_color(n["location"], cam.lookfrom);
    cam.lookat = read_color(n["look_at"], cam.lookat);
    cam.vup = read_color(n["up"], cam.vup);
    cam.background = read_color(n["background"], cam.background);
    cam.samples_per_pixel = to_int(n["samples_per_pixel"], cam.samples_per_pixel);
    cam.max_depth = to_int(n["max_depth"], cam.max_depth);
    const std::string output = to_string(n["output"]);
    if (!output.empty()) cam.file_name = output;
}

}  // namespace

namespace {
void load_media(const node& root, camera& cam) {
    // Extension of the reference's schema (it can construct no constant_medium, constant_medium.h:8-61):
    //   media:
    //     - density: 0.01
    //       color: [230, 230, 230]                  (0-255 or 0-1, like a material's)
    //       box: [[x0, y0, z0], [x1, y1, z1]]       an axis-aligned box as the boundary, or
    //       vertices: [...] / triangles: [...]      a closed convex mesh, as a `mesh` surface spells it
    //       transform: {rotate_y: deg, translate: [x, y, z]}
    // They land in cam.options.media: camera::render hands them to bdpt_scene_set_media beside the world.
    const node& media = root["media"];
    if (media.is_sequence()) {
        for (const auto& md : media.seq) {
            if (!md.is_map()) throw std::runtime_error("media entries must be mappings");
            const double density = to_double(md["density"], 0.0);
            if (!(density > 0.0)) throw std::runtime_error("medium: density must be > 0");
            auto boundary = make_shared<triangle_collection>();
            auto skin = make_shared<lambertian>(color(0, 0, 0));
            const auto box = md["box"];
            if (box.is_sequence()) {
                if (box.seq.size() != 2) throw std::runtime_error("medium: box must be [[x0, y0, z0], [x1, y1, z1]]");
                const auto lo = to_double_list(box.seq[0]), hi = to_double_list(box.seq[1]);
                if (lo.size() != 3 || hi.size() != 3) throw std::runtime_error("medium: box must be [[x0, y0, z0], [x1, y1, z1]]");
                point3 c[8];
                for (int i = 0; i < 8; ++i) c[i] = point3(i & 1 ? hi[0] : lo[0], i & 2 ? hi[1] : lo[1], i & 4 ? hi[2] : lo[2]);
                static const int faces[6][4] = {{0, 1, 3, 2}, {4, 6, 7, 5}, {0, 4, 5, 1}, {2, 3, 7, 6}, {0, 2, 6, 4}, {1, 5, 7, 3}};
                for (const auto& q : faces) {
                    boundary->add(triangle(c[q[0]], c[q[1]], c[q[2]], skin));
                    boundary->add(triangle(c[q[0]], c[q[2]], c[q[3]], skin));
                }
            } else {
                triangle_collection none;
                sink bs{*boundary, none};
                material_map no_mats;
                load_indexed_mesh(md, bs, no_mats);
            }
            if (boundary->empty()) throw std::runtime_error("medium: empty boundary (give `box` or `vertices` + `triangles`)");
            shared_ptr<hittable> obj = boundary;
            const node& tf = md["transform"];
            if (tf.is_map()) {
                const auto off = to_double_list(tf["translate"]);
                if (tf["translate"].defined() && off.size() != 3) throw std::runtime_error("transform.translate must be [x, y, z]");
                if (tf["rotate_y"].defined()) obj = make_shared<rotate_y>(obj, to_double(tf["rotate_y"], 0.0));
                if (off.size() == 3) obj = make_shared<translate>(obj, vec3(off[0], off[1], off[2]));
            }
            cam.options.media.emplace_back(obj, density, read_color_scaled(md["color"], color(1, 1, 1)));
        }
    } else if (media.defined()) throw std::runtime_error("media must be a sequence");
}
}  // namespace

scene_load_result load_scene_from_yaml(const std::string& path) {    // scene_loader.h:480-523
    const node root = yaml_lite::load_file(path);
    if (!root.is_map()) throw std::runtime_error("Scene root must be a mapping");
    scene_load_result result;
    load_camera(root["camera"], result.cam);
    g_scene_dir = std::filesystem::path(path).parent_path().string();
    reset_scene_rng();
    const material_map mats = load_materials(root["materials"]);
    const node* surfaces = &root["surfaces"];
    if (!surfaces->defined()) surfaces = &root["scene"];
    if (!surfaces->is_sequence()) throw std::runtime_error("Scene/surfaces field missing or not a sequence");
    const std::filesystem::path dir = std::filesystem::path(path).parent_path();
    sink out{result.world, result.lights};
    // OBJ files are read and tokenised concurrently; their triangles still enter the world in scene order
    obj_jobs jobs;
    for (const auto& mesh : surfaces->seq)
        if (mesh.is_map() && to_string(mesh["type"]) == "object" && !to_string(mesh["file"]).empty())
            jobs.emplace(&mesh, std::async(std::launch::async, parse_obj_file, object_path(mesh, dir)));
    for (const auto& mesh : surfaces->seq) {
        if (!mesh.is_map()) throw std::runtime_error("Scene entries must be mappings");
        const std::string type = to_string(mesh["type"]);
        if (type.empty()) throw std::runtime_error("Mesh missing type field");
        const size_t w0 = result.world.size(), l0 = result.lights.size(), s0 = result.cam.options.spheres.size();
        if (type == "TriMesh") load_tri_mesh(mesh, out);
        else if (type == "Sphere") load_sphere(mesh, out, result.cam.options.spheres);
        else if (type == "mesh") load_indexed_mesh(mesh, out, mats);
        else if (type == "object") load_object(mesh, dir, out, mats, jobs);
        else std::cerr << "Unknown mesh type: " << type << "\n";
        // Extension of the reference's schema (it can construct no translate / rotate_y, hittable.h:46-157):
        //   transform: {rotate_y: <degrees>, translate: [x, y, z]}   = translate(rotate_y(surface, deg), offset)
        // applied to the triangles this surface just added (world and light list alike), baked like collect_triangles does.
        const node& tf = mesh["transform"];
        if (tf.is_map()) {
            const double deg = to_double(tf["rotate_y"], 0.0);
            const auto off = to_double_list(tf["translate"]);
            if (tf["translate"].defined() && off.size() != 3) throw std::runtime_error("transform.translate must be [x, y, z]");
            auto bake = [&](std::vector<triangle>& tris, size_t first) {
                if (first >= tris.size()) return;
                auto span = make_shared<triangle_collection>();
                span->tris.assign(tris.begin() + long(first), tris.end());
                shared_ptr<hittable> obj = span;
                if (tf["rotate_y"].defined()) obj = make_shared<rotate_y>(obj, deg);
                if (off.size() == 3) obj = make_shared<translate>(obj, vec3(off[0], off[1], off[2]));
                std::vector<triangle> moved;
                collect_triangles(*obj, moved);
                std::copy(moved.begin(), moved.end(), tris.begin() + long(first));
            };
            bake(result.world.tris, w0);
            bake(result.lights.tris, l0);
            for (size_t i = s0; i < result.cam.options.spheres.size(); ++i) {   // an analytic sphere of this surface: its centre moves
                sphere& sp = result.cam.options.spheres[i];
                shared_ptr<hittable> obj = make_shared<sphere>(sp);
                if (tf["rotate_y"].defined()) obj = make_shared<rotate_y>(obj, deg);
                if (off.size() == 3) obj = make_shared<translate>(obj, vec3(off[0], off[1], off[2]));
                std::vector<triangle> none;
                std::vector<sphere> moved;
                collect_triangles(*obj, none, nullptr, &moved);
                sp = moved.at(0);
            }
        } else if (tf.defined()) throw std::runtime_error("transform must be a mapping {rotate_y: deg, translate: [x, y, z]}");
    }
    load_media(root, result.cam);
    std::cout << "Triangles: " << result.world.size() << "\n";
    return result;
}

camera load_camera_from_yaml(const std::string& path) {
    const node root = yaml_lite::load_file(path);
    if (!root.is_map()) throw std::runtime_error("Scene root must be a mapping");
    camera cam;
    load_camera(root["camera"], cam);
    load_media(root, cam);   // the scene-image cache holds the world only: media are re-read with the camera
    const node* surfaces = &root["surfaces"];
    if (!surfaces->defined()) surfaces = &root["scene"];
    if (surfaces->is_sequence())
        for (const auto& mesh : surfaces->seq)
            if (mesh.is_map() && mesh["analytic"].defined())
                throw std::runtime_error("analytic spheres need the scene's material table, which a scene image does not expose: render this scene without --cache");
    return cam;
}

namespace {
// word-wise multiply-xor mix (not cryptographic): detects edited, replaced or truncated source files
uint64_t mix_bytes(uint64_t h, const char* p, size_t n) {
    h ^= n * 0x9E3779B97F4A7C15ull;
    uint64_t lane[4] = {h, h ^ 0xC2B2AE3D27D4EB4Full, h ^ 0x165667B19E3779F9ull, h ^ 0x27D4EB2F165667C5ull};
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        uint64_t w[4];
        std::memcpy(w, p + i, 32);
        for (int k = 0; k < 4; ++k) { lane[k] = (lane[k] ^ w[k]) * 0x100000001B3ull; lane[k] = (lane[k] << 29) | (lane[k] >> 35); }
    }
    for (; i < n; ++i) { lane[0] = (lane[0] ^ uint64_t(static_cast<unsigned char>(p[i]))) * 0x100000001B3ull; lane[0] = (lane[0] << 29) | (lane[0] >> 35); }
    uint64_t r = n;
    for (int k = 0; k < 4; ++k) { r = (r ^ lane[k]) * 0xFF51AFD7ED558CCDull; r ^= r >> 33; }
    return r;
}
std::string slurp(const std::filesystem::path& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("Failed to open file: " + path.string());
    return std::string((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
}
}  // namespace

uint64_t scene_source_key(const std::string& path) {
    const std::string yaml = slurp(path);
    uint64_t key = mix_bytes(0x42445054ull, yaml.data(), yaml.size());
    const node root = yaml_lite::load_file(path);
    if (!root.is_map()) throw std::runtime_error("Scene root must be a mapping");
    const node* surfaces = &root["surfaces"];
    if (!surfaces->defined()) surfaces = &root["scene"];
    if (!surfaces->is_sequence()) throw std::runtime_error("Scene/surfaces field missing or not a sequence");
    const std::filesystem::path dir = std::filesystem::path(path).parent_path();
    std::vector<std::future<uint64_t>> jobs;
    for (const auto& mesh : surfaces->seq)
        if (mesh.is_map() && to_string(mesh["type"]) == "object" && !to_string(mesh["file"]).empty()) {
            const std::filesystem::path obj = object_path(mesh, dir);
            jobs.push_back(std::async(std::launch::async, [obj] { const std::string b = slurp(obj); return mix_bytes(1, b.data(), b.size()); }));
        }
    for (auto& j : jobs) { const uint64_t h = j.get(); key = mix_bytes(key, reinterpret_cast<const char*>(&h), sizeof h); }
    return key ? key : 1;   // 0 means "no key" in an image header
}

// ---------------------------------------------------------------- textures (texture.h, perlin.h, rtw_stb_image.h)
namespace {
struct host_rng {   // main.h:28-37: one default-seeded mt19937 + uniform_real_distribution
    std::mt19937 gen;
    std::uniform_real_distribution<double> dist{0.0, 1.0};
    double next() { return dist(gen); }
    double next(double mn, double mx) { return mn + (mx - mn) * next(); }
    int next_int(int mn, int mx) { return int(next(mn, mx + 1)); }   // main.h:39-42
};
host_rng& scene_rng() { static host_rng r; return r; }
void reset_scene_rng() { scene_rng() = host_rng(); }
}  // namespace

noise_texture::noise_texture(double scale_) {
    kind = Noise; scale = scale_;
    host_rng& r = scene_rng();
    data.resize(256 * 3 * sizeof(double) + 3 * 256 * sizeof(int32_t));
    double* randvec = reinterpret_cast<double*>(data.data());
    int32_t* perm = reinterpret_cast<int32_t*>(data.data() + 256 * 3 * sizeof(double));
    for (int i = 0; i < 256; ++i) {   // perlin.h:7-9: unit_vector(vec3::random(-1, 1)); g++ evaluates the arguments z, y, x
        const double z = r.next(-1, 1), y = r.next(-1, 1), x = r.next(-1, 1);
        const double inv = 1 / std::sqrt(x * x + y * y + z * z);      // vec3.h:114-116: v / length = (1/length) * v
        randvec[3 * i] = inv * x; randvec[3 * i + 1] = inv * y; randvec[3 * i + 2] = inv * z;
    }
    for (int t = 0; t < 3; ++t) {     // perlin.h:11-13,68-84
        int32_t* p = perm + 256 * t;
        for (int i = 0; i < 256; ++i) p[i] = i;
        for (int i = 255; i > 0; --i) std::swap(p[i], p[r.next_int(0, i)]);
    }
}

image_texture::image_texture(const char* filename) {
    kind = Image;
    std::ifstream f(filename, std::ios::binary);
    std::string magic;
    int maxv = 0;
    if (!(f >> magic >> width >> height >> maxv) || magic != "P6" || maxv != 255 || width <= 0 || height <= 0) {
        std::cerr << "ERROR: Could not load image file '" << filename << "' (binary PPM, P6, maxval 255).\n";   // rtw_stb_image.h:36
        width = height = 0;
        return;
    }
    f.get();
    std::vector<unsigned char> raw(size_t(width) * height * 3);
    f.read(reinterpret_cast<char*>(raw.data()), std::streamsize(raw.size()));
    data.resize(raw.size());
    for (size_t i = 0; i < raw.size(); ++i) {
        // stb_image's stbi__ldr_to_hdr: (float)(pow(byte / 255.0f, 2.2f) * 1.0f) in float; then rtw_stb_image.h:90-97
        const float value = std::pow(float(raw[i]) / 255.0f, 2.2f) * 1.0f;
        data[i] = value <= 0.0 ? 0 : (1.0 <= value ? 255 : static_cast<unsigned char>(256.0 * value));
    }
}

void flatten_textures(const std::vector<triangle>& tris, std::vector<bdpt_material>& mats, std::vector<bdpt_texture>& tex,
                      std::vector<unsigned char>& data) {
    tex.clear(); data.clear();
    std::unordered_map<const material*, uint32_t> mat_index;
    std::unordered_map<const texture*, uint32_t> tex_index;
    for (const triangle& t : tris) {
        if (mat_index.count(t.mat.get())) continue;
        const uint32_t mi = uint32_t(mat_index.size());
        mat_index.emplace(t.mat.get(), mi);
        if (!t.mat->tex || mi >= mats.size()) continue;
        const texture* tx = t.mat->tex.get();
        auto it = tex_index.find(tx);
        if (it == tex_index.end()) {
            bdpt_texture d{};
            d.type = uint32_t(tx->kind); d.scale = tx->scale; d.width = tx->width; d.height = tx->height;
            for (int k = 0; k < 3; ++k) { d.a[k] = tx->a[k]; d.b[k] = tx->b[k]; }
            while (data.size() % 8) data.push_back(0);
            d.data_offset = data.size();
            data.insert(data.end(), tx->data.begin(), tx->data.end());
            it = tex_index.emplace(tx, uint32_t(tex.size())).first;
            tex.push_back(d);
        }
        mats[mi].tex = it->second + 1;
    }
}

void flatten_triangles(const std::vector<triangle>& tris, std::vector<double>& verts, std::vector<uint32_t>& mat_id,
                       std::vector<bdpt_material>& mats) {
    verts.resize(tris.size() * 9);
    mat_id.resize(tris.size());
    mats.clear();
    std::unordered_map<const material*, uint32_t> index;
    for (size_t i = 0; i < tris.size(); ++i) {
        const triangle& t = tris[i];
        const point3* p[3] = {&t.v0, &t.v1, &t.v2};
        for (int k = 0; k < 3; ++k) for (int c = 0; c < 3; ++c) verts[9 * i + 3 * k + c] = (*p[k])[c];
        auto it = index.find(t.mat.get());
        if (it == index.end()) {
            if (t.mat->phase) throw std::runtime_error("isotropic is the phase function of a constant_medium, not a surface material");
            bdpt_material m{};
            m.type = uint32_t(t.mat->kind);
            m.rgb[0] = t.mat->rgb.x(); m.rgb[1] = t.mat->rgb.y(); m.rgb[2] = t.mat->rgb.z();
            m.param = t.mat->param;
            it = index.emplace(t.mat.get(), uint32_t(mats.size())).first;
            mats.push_back(m);
        }
        mat_id[i] = it->second;
    }
}
