// TEST INFRASTRUCTURE ONLY — never linked, imported or executed by the product path.
//
// CPU restatement ("port") of the reference's BDPT hot path, written from the reference's behaviour
// with flat arrays instead of its shared_ptr/virtual object graph.  Every function cites the
// reference file:line it follows (paths relative to /root/reference/src).  The arithmetic keeps the
// reference's exact operation order in fp64 (e.g. v/t is (1/t)*v, vec3.h:40-42,98-100) so that in
// ORC_RNG_MT19937 mode a single-threaded render is BIT-IDENTICAL to the reference's own headers
// (pinned by tests/test_oracle_vs_reference.py here and by tests/golden/ on the GPU box).
// ORC_RNG_PHILOX mode swaps only the random stream for the counter-based generator the CUDA kernels
// use, which makes GPU-vs-port whole-render comparison possible sample by sample.
//
// Language: C++17, not plain C, because bit parity needs libstdc++'s std::sort (BVH leaf order for
// equal keys, bvh.h:41) and std::mt19937 + std::uniform_real_distribution (main.h:28-32).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <random>
#include <thread>
#include <vector>

#define ORC_PREFIX port_
#include "oracle_api.h"

namespace {

const double kInf = std::numeric_limits<double>::infinity();
const double kPi = 3.1415926535897932385;  // main.h:20

// ---------------------------------------------------------------- vec3.h:7-159
struct V3 { double x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator*(double t, V3 v) { return {t * v.x, t * v.y, t * v.z}; }       // vec3.h:90-92
inline V3 operator*(V3 v, double t) { return t * v; }                             // vec3.h:94-96
inline V3 operator/(V3 v, double t) { return (1 / t) * v; }                       // vec3.h:98-100
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }       // vec3.h:102-106
inline V3 cross(V3 u, V3 v) {                                                     // vec3.h:108-112
    return {u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x};
}
inline double len2(V3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }            // vec3.h:43-45
inline double len(V3 v) { return std::sqrt(len2(v)); }
inline V3 unit(V3 v) { return v / len(v); }                                       // vec3.h:114-116
inline double comp(const V3& v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

// ---------------------------------------------------------------- RNG (main.h:28-42)
// Philox4x32-10, key = (seed_lo ^ stream, seed_hi), counter = (sample_lo, sample_hi, block, 0x5EED).
// One block yields two doubles in [0,1): (hi:lo >> 11) * 2^-53.  Identical code in csrc/rng.cuh.
struct Philox {
    static inline void round(uint32_t c[4], uint32_t k0, uint32_t k1) {
        const uint64_t p0 = uint64_t(0xD2511F53u) * c[0];
        const uint64_t p1 = uint64_t(0xCD9E8D57u) * c[2];
        uint32_t n0 = uint32_t(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = uint32_t(p1);
        uint32_t n2 = uint32_t(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = uint32_t(p0);
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    }
    static inline void block(uint64_t seed, uint32_t stream, uint64_t sample, uint32_t blk, uint32_t out[4]) {
        uint32_t k0 = uint32_t(seed) ^ stream, k1 = uint32_t(seed >> 32);
        uint32_t c[4] = {uint32_t(sample), uint32_t(sample >> 32), blk, 0x5EEDu};
        for (int i = 0; i < 10; ++i) {
            round(c, k0, k1);
            k0 += 0x9E3779B9u;
            k1 += 0xBB67AE85u;
        }
        out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
    }
};

// Deterministic sin/cos and x^5 used ONLY in ORC_RNG_PHILOX mode (never in the reference-pinned
// MT19937 mode, which keeps libm's std::sin/std::cos/std::pow for bit parity with the reference).
// libm and CUDA's math library differ in the last ulp, and camera::visible()'s far-endpoint test is a
// rounding knife edge, so a per-sample GPU-vs-port comparison needs transcendental
// functions that are plain IEEE add/mul sequences: Cody-Waite reduction by pi/2 + the classic degree-13/12
// minimax kernels on [-pi/4, pi/4] (error < 1 ulp).  csrc/render_math.cuh states the same algorithm.
inline void det_sincos(double phi, double& s_out, double& c_out) {
    const int k = int(phi * 0.63661977236758134308 + 0.5);               // nearest multiple of pi/2
    const double y = (phi - k * 1.57079632673412561417e+00) - k * 6.07710050650619224932e-11;
    const double z = y * y;
    const double s = y + y * z * (-1.66666666666666324348e-01 + z * (8.33333333332248946124e-03 + z * (-1.98412698298579493134e-04 +
                     z * (2.75573137070700676789e-06 + z * (-2.50507602534068634195e-08 + z * 1.58969099521155010221e-10)))));
    const double c = 1.0 - 0.5 * z + z * z * (4.16666666666666019037e-02 + z * (-1.38888888888741095749e-03 + z * (2.48015872894767294178e-05 +
                     z * (-2.75573143513906633035e-07 + z * (2.08757232129817482790e-09 + z * -1.13596475577881948265e-11)))));
    switch (k & 3) {
        case 0: s_out = s; c_out = c; break;
        case 1: s_out = c; c_out = -s; break;
        case 2: s_out = -s; c_out = -c; break;
        default: s_out = -c; c_out = s; break;
    }
}
inline double det_pow5(double x) { const double x2 = x * x; const double x4 = x2 * x2; return x4 * x; }

// Deterministic natural logarithm for ORC_RNG_PHILOX mode (constant_medium.h:44 draws log(random_double()) inside hit()):
// fdlibm's __ieee754_log as a fixed sequence of IEEE +, -, *, / on the argument's mantissa and exponent, so the CUDA
// kernels (csrc/render_math.cuh: det_log, same constants, same order, -fmad=false) and this port agree bit for bit.
// x in [0, 1): 0 -> -inf; the smallest non-zero draw is 2^-53, a normal number.
inline double det_log(double x) {
    if (!(x > 0.0)) return -std::numeric_limits<double>::infinity();
    uint64_t bits;
    std::memcpy(&bits, &x, 8);
    int e = int((bits >> 52) & 0x7FF) - 1023;
    bits = (bits & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
    double m;
    std::memcpy(&m, &bits, 8);
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    const double f = m - 1.0;
    const double s_ = f / (2.0 + f);
    const double z = s_ * s_, w = z * z;
    const double t1 = w * (3.999999999940941908e-01 + w * (2.222219843214978396e-01 + w * 1.531383769920937332e-01));
    const double t2 = z * (6.666666666666735130e-01 + w * (2.857142874366239149e-01 + w * (1.818357216161805012e-01 + w * 1.479819860511658591e-01)));
    const double R = t2 + t1;
    const double hfsq = 0.5 * f * f;
    const double dk = double(e);
    return dk * 6.93147180369123816490e-01 - ((hfsq - (s_ * (hfsq + R) + dk * 1.90821492927058770002e-10)) - f);
}

struct Rng {
    bool det = false;  // PHILOX mode: deterministic sin/cos/pow5 (see above)
    // mode MT: all streams alias one engine, exactly the reference's single global stream
    std::mt19937* mt = nullptr;
    std::uniform_real_distribution<double>* dist = nullptr;
    // mode PHILOX
    uint64_t seed = 0, sample = 0;
    uint32_t stream = 0, draws = 0;
    double next() {
        if (mt) return (*dist)(*mt);
        uint32_t o[4];
        Philox::block(seed, stream, sample, draws >> 1, o);
        const uint32_t lo = (draws & 1) ? o[2] : o[0], hi = (draws & 1) ? o[3] : o[1];
        ++draws;
        return double(((uint64_t(hi) << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
    }
    double next(double mn, double mx) { return mn + (mx - mn) * next(); }  // main.h:34-37
};

// vec3.h:128-135 — g++ evaluates the three constructor arguments of vec3::random(min,max)
// (vec3.h:58-60) in the order pinned by tests/test_oracle_vs_reference.py; see kRandomVecOrder.
inline V3 random_unit_vector(Rng& r) {
    for (;;) {
        // argument evaluation order of `vec3(random_double(..), random_double(..), random_double(..))`
        // with this toolchain is last-to-first (z, y, x); verified bit-exactly against the reference.
        double z = r.next(-1, 1);
        double y = r.next(-1, 1);
        double x = r.next(-1, 1);
        V3 p{x, y, z};
        double lensq = len2(p);
        if (1e-160 < lensq && lensq <= 1) return p / std::sqrt(lensq);
    }
}
inline V3 reflect(V3 v, V3 n) { return v - 2 * dot(v, n) * n; }  // vec3.h:138-140
inline V3 refract(V3 uv, V3 n, double etai_over_etat) {          // vec3.h:142-147
    double cos_theta = std::fmin(dot(-uv, n), 1.0);
    V3 perp = etai_over_etat * (uv + cos_theta * n);
    V3 par = -std::sqrt(std::fabs(1.0 - len2(perp))) * n;
    return perp + par;
}
inline V3 random_cosine_direction(Rng& r) {  // vec3.h:149-159
    double r1 = r.next(), r2 = r.next();
    double phi = 2 * kPi * r1;
    double cs = 0, sn = 0;
    if (r.det) det_sincos(phi, sn, cs); else { cs = std::cos(phi); sn = std::sin(phi); }
    double x = cs * std::sqrt(r2);
    double y = sn * std::sqrt(r2);
    double z = std::sqrt(1 - r2);
    return {x, y, z};
}
struct Onb {  // onb.h:4-24
    V3 a0, a1, a2;
    explicit Onb(V3 n) {
        a2 = unit(n);
        V3 a = (std::fabs(a2.x) > 0.9) ? V3{0, 1, 0} : V3{1, 0, 0};
        a1 = unit(cross(a2, a));
        a0 = cross(a2, a1);
    }
    V3 transform(V3 v) const { return (v.x * a0) + (v.y * a1) + (v.z * a2); }
};

// ---------------------------------------------------------------- interval.h / aabb.h
struct Ival {
    double mn, mx;
    double size() const { return mx - mn; }
};
inline Ival merge(Ival a, Ival b) { return {a.mn <= b.mn ? a.mn : b.mn, a.mx >= b.mx ? a.mx : b.mx}; }  // interval.h:12-16
struct Box {
    Ival ax[3];
    void pad() {  // aabb.h:81-88 (expand: interval.h:36-39)
        const double delta = 0.0001;
        for (int a = 0; a < 3; ++a)
            if (ax[a].size() < delta) { double p = delta / 2; ax[a] = {ax[a].mn - p, ax[a].mx + p}; }
    }
    int longest_axis() const {  // aabb.h:68-75 (ties -> later axis)
        if (ax[0].size() > ax[1].size()) return ax[0].size() > ax[2].size() ? 0 : 2;
        return ax[1].size() > ax[2].size() ? 1 : 2;
    }
};
inline Box merge(const Box& a, const Box& b) {  // aabb.h:27-33
    Box r;
    for (int i = 0; i < 3; ++i) r.ax[i] = merge(a.ax[i], b.ax[i]);
    r.pad();
    return r;
}
inline Box empty_box() {  // aabb.h:92 (padding an empty interval leaves (+inf,-inf))
    Box r;
    for (int i = 0; i < 3; ++i) r.ax[i] = {kInf, -kInf};
    r.pad();
    return r;
}

struct Stats { uint64_t rays = 0, node_visits = 0, aabb_tests = 0, aabb_hits = 0, tri_tests = 0, tri_hits = 0, shadow = 0; };
thread_local Stats tl_stats;
std::atomic<uint64_t> g_stats[8];
void flush_stats() {
    g_stats[0] += tl_stats.rays; g_stats[1] += tl_stats.node_visits; g_stats[2] += tl_stats.aabb_tests;
    g_stats[3] += tl_stats.aabb_hits; g_stats[4] += tl_stats.tri_tests; g_stats[5] += tl_stats.tri_hits;
    g_stats[7] += tl_stats.shadow;
    tl_stats = Stats();
}

inline bool box_hit(const Box& b, V3 o, V3 d, double tmin, double tmax) {  // aabb.h:41-66
    ++tl_stats.aabb_tests;
    for (int a = 0; a < 3; ++a) {
        const double adinv = 1.0 / comp(d, a);
        double t0 = (b.ax[a].mn - comp(o, a)) * adinv;
        double t1 = (b.ax[a].mx - comp(o, a)) * adinv;
        if (t0 < t1) {
            if (t0 > tmin) tmin = t0;
            if (t1 < tmax) tmax = t1;
        } else {
            if (t1 > tmin) tmin = t1;
            if (t0 < tmax) tmax = t0;
        }
        if (tmax <= tmin) return false;
    }
    ++tl_stats.aabb_hits;
    return true;
}

// ---------------------------------------------------------------- triangle.h:19-133
struct Tri {
    V3 v0, e1, e2, normal;
    double area;
    Box box;
    uint32_t mat;
};
inline Tri make_tri(V3 v0, V3 v1, V3 v2, uint32_t mat) {  // triangle.h:21-39
    Tri t;
    t.v0 = v0; t.e1 = v1 - v0; t.e2 = v2 - v0;
    V3 n = cross(t.e1, t.e2);
    t.normal = unit(n);
    t.area = 0.5 * len(n);
    t.mat = mat;
    double mnx = std::fmin(v0.x, std::fmin(v1.x, v2.x)), mxx = std::fmax(v0.x, std::fmax(v1.x, v2.x));
    double mny = std::fmin(v0.y, std::fmin(v1.y, v2.y)), mxy = std::fmax(v0.y, std::fmax(v1.y, v2.y));
    double mnz = std::fmin(v0.z, std::fmin(v1.z, v2.z)), mxz = std::fmax(v0.z, std::fmax(v1.z, v2.z));
    t.box.ax[0] = {mnx, mxx}; t.box.ax[1] = {mny, mxy}; t.box.ax[2] = {mnz, mxz};  // aabb.h:18-25, unpadded
    return t;
}
struct Hit {  // hittable.h:10-27 plus the prim ID the reference lacks (F6)
    V3 p, normal;
    double t, u, v;
    bool front;
    uint32_t prim;
};
inline bool tri_hit(const Tri& tr, uint32_t id, V3 o, V3 d, double tmin, double tmax, Hit& rec) {  // triangle.h:41-74
    ++tl_stats.tri_tests;
    V3 pvec = cross(d, tr.e2);
    double det = dot(tr.e1, pvec);
    if (std::fabs(det) < 1e-8) return false;
    double inv_det = 1.0 / det;
    V3 tvec = o - tr.v0;
    double u = dot(tvec, pvec) * inv_det;
    if (u < 0.0 || u > 1.0) return false;
    V3 qvec = cross(tvec, tr.e1);
    double v = dot(d, qvec) * inv_det;
    if (v < 0.0 || u + v > 1.0) return false;
    double t = dot(tr.e2, qvec) * inv_det;
    if (!(tmin <= t && t <= tmax)) return false;  // interval.h:22-24, inclusive
    rec.t = t;
    rec.p = o + t * d;  // ray.h:21-23
    rec.u = u; rec.v = v;
    rec.front = dot(d, tr.normal) < 0;  // hittable.h:20-26
    rec.normal = rec.front ? tr.normal : -tr.normal;
    rec.prim = id;
    ++tl_stats.tri_hits;
    return true;
}

// ---------------------------------------------------------------- bvh.h:13-86
struct Node {
    Box box;
    int32_t left, right;  // internal: node indices; leaf: triangle (prim) IDs (left==right for span 1)
    bool leaf;
};

// constant_medium (materials/volumes/constant_medium.h:8-61) with its isotropic phase function (material.h:149-171): a closed
// boundary mesh, kept OUTSIDE the world soup (its triangles have no prim ID and no material), a density and a colour.
constexpr uint32_t kMediumPrim = 0x20000000u;   // Hit.prim of a scatter point inside medium k: kMediumPrim | k
struct Medium {
    double neg_inv_density;
    V3 albedo;
    std::vector<Tri> boundary;
};

// Analytic sphere (an extension: the reference only tessellates its spheres, scene_loader.h:212-242; SURVEY 8 "sphere row").  A world
// prim with a material, kept in a side list and asked after the triangle walk, in list order, with the bound found so far.
constexpr uint32_t kSpherePrim = 0x10000000u;   // Hit.prim of a hit on sphere k: kSpherePrim | k
struct Sphere { V3 c; double r; uint32_t mat; };
// The classic quadratic in the reference's own style: inclusive interval like triangle.h:68 (interval.h:22-24), outward normal
// (p - c) / r face-forwarded like hittable.h:20-26; u = v = 0.  Nearer root first, then the farther one.
inline bool sphere_hit(const Sphere& sp, uint32_t id, V3 o, V3 d, double tmin, double tmax, Hit& rec) {
    V3 oc = sp.c - o;
    double a = len2(d);
    double h = dot(d, oc);
    double c = len2(oc) - sp.r * sp.r;
    double discriminant = h * h - a * c;
    if (discriminant < 0) return false;
    double sqrtd = std::sqrt(discriminant);
    double root = (h - sqrtd) / a;
    if (!(tmin <= root && root <= tmax)) {
        root = (h + sqrtd) / a;
        if (!(tmin <= root && root <= tmax)) return false;
    }
    rec.t = root;
    rec.p = o + root * d;
    V3 outward = (rec.p - sp.c) / sp.r;
    rec.front = dot(d, outward) < 0;
    rec.normal = rec.front ? outward : -outward;
    rec.u = 0; rec.v = 0;
    rec.prim = kSpherePrim | id;
    return true;
}

struct Scene {
    std::vector<Sphere> spheres;
    std::vector<Medium> media;
    std::vector<Tri> tris;
    std::vector<orc_material> mats;
    std::vector<orc_texture> tex;
    std::vector<unsigned char> tex_data;

    // perlin::noise / turb, perlin.h:16-49,77-91
    double perlin_noise(const unsigned char* tables, V3 p) const {
        const double* randvec = reinterpret_cast<const double*>(tables);
        const int32_t* perm = reinterpret_cast<const int32_t*>(tables + 256 * 3 * sizeof(double));
        double u = p.x - std::floor(p.x), v = p.y - std::floor(p.y), w = p.z - std::floor(p.z);
        int i = int(std::floor(p.x)), j = int(std::floor(p.y)), k = int(std::floor(p.z));
        double uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
        double accum = 0.0;
        for (int di = 0; di < 2; di++)
            for (int dj = 0; dj < 2; dj++)
                for (int dk = 0; dk < 2; dk++) {
                    int idx = perm[(i + di) & 255] ^ perm[256 + ((j + dj) & 255)] ^ perm[512 + ((k + dk) & 255)];
                    V3 c{randvec[3 * idx], randvec[3 * idx + 1], randvec[3 * idx + 2]};
                    V3 weight_v{u - di, v - dj, w - dk};
                    accum += (di * uu + (1 - di) * (1 - uu)) * (dj * vv + (1 - dj) * (1 - vv)) * (dk * ww + (1 - dk) * (1 - ww)) * dot(c, weight_v);
                }
        return accum;
    }
    double perlin_turb(const unsigned char* tables, V3 p, int depth) const {
        double accum = 0.0, weight = 1.0;
        V3 tp = p;
        for (int i = 0; i < depth; i++) {
            accum += weight * perlin_noise(tables, tp);
            weight *= 0.5;
            tp = {tp.x * 2, tp.y * 2, tp.z * 2};
        }
        return std::fabs(accum);
    }
    // texture::value(u, v, p), texture.h:14-87
    V3 colour(const orc_material& m, double u, double v, V3 p) const {
        if (m.tex == 0 || m.tex > tex.size()) return {m.rgb[0], m.rgb[1], m.rgb[2]};
        const orc_texture& t = tex[m.tex - 1];
        if (t.type == ORC_TEX_CHECKER) {
            double inv_scale = 1.0 / t.scale;
            int xi = int(std::floor(inv_scale * p.x)), yi = int(std::floor(inv_scale * p.y)), zi = int(std::floor(inv_scale * p.z));
            return ((xi + yi + zi) % 2 == 0) ? V3{t.a[0], t.a[1], t.a[2]} : V3{t.b[0], t.b[1], t.b[2]};
        }
        if (t.type == ORC_TEX_NOISE)
            return V3{.5, .5, .5} * (1 + std::sin(t.scale * p.z + 10 * perlin_turb(tex_data.data() + t.data_offset, p, 7)));
        if (t.type == ORC_TEX_IMAGE) {
            if (t.height <= 0) return {0, 1, 1};
            u = u < 0 ? 0 : (u > 1 ? 1 : u);
            v = 1.0 - (v < 0 ? 0 : (v > 1 ? 1 : v));
            int i = int(u * t.width), j = int(v * t.height);
            i = i < 0 ? 0 : (i < t.width ? i : t.width - 1);
            j = j < 0 ? 0 : (j < t.height ? j : t.height - 1);
            const unsigned char* px = tex_data.data() + t.data_offset + (size_t(j) * t.width + i) * 3;
            double color_scale = 1.0 / 255.0;
            return {color_scale * px[0], color_scale * px[1], color_scale * px[2]};
        }
        return {t.a[0], t.a[1], t.a[2]};
    }
    std::vector<uint32_t> lights;  // prim IDs in reference light-list order (scene_loader.h:190-202; main.cpp:67)
    std::vector<Node> nodes;
    std::vector<uint32_t> order;   // builder's working array == DFS leaf order when finished
    int32_t root = -1;

    int32_t build(size_t start, size_t end) {  // bvh.h:20-48
        int32_t me = int32_t(nodes.size());
        nodes.push_back(Node());
        Box bb = empty_box();
        for (size_t i = start; i < end; ++i) bb = merge(bb, tris[order[i]].box);
        int axis = bb.longest_axis();
        size_t span = end - start;
        Node n;
        n.box = bb;
        if (span == 1) {
            n.leaf = true; n.left = n.right = int32_t(order[start]);
        } else if (span == 2) {
            n.leaf = true; n.left = int32_t(order[start]); n.right = int32_t(order[start + 1]);
        } else {
            // bvh.h:41,68-74: std::sort (unstable introsort) on the box minimum along `axis`
            std::sort(order.begin() + start, order.begin() + end, [&](uint32_t a, uint32_t b) {
                return tris[a].box.ax[axis].mn < tris[b].box.ax[axis].mn;
            });
            size_t mid = start + span / 2;
            n.leaf = false;
            n.left = build(start, mid);
            n.right = build(mid, end);
        }
        nodes[me] = n;
        return me;
    }

    bool hit_node(int32_t ni, V3 o, V3 d, double tmin, double tmax, Hit& rec) const {  // bvh.h:50-59
        ++tl_stats.node_visits;
        const Node& n = nodes[ni];
        if (!box_hit(n.box, o, d, tmin, tmax)) return false;
        bool hl, hr;
        if (n.leaf) {
            hl = tri_hit(tris[n.left], uint32_t(n.left), o, d, tmin, tmax, rec);
            hr = tri_hit(tris[n.right], uint32_t(n.right), o, d, tmin, hl ? rec.t : tmax, rec);
        } else {
            hl = hit_node(n.left, o, d, tmin, tmax, rec);
            hr = hit_node(n.right, o, d, tmin, hl ? rec.t : tmax, rec);
        }
        return hl || hr;
    }
    bool hit(V3 o, V3 d, double tmin, double tmax, Hit& rec) const {
        ++tl_stats.rays;
        return hit_node(root, o, d, tmin, tmax, rec);
    }

    // boundary->hit(r, ray_t): the boundary as a list of triangles in order (triangle_collection::hit, triangle.h:150-164);
    // only t is used by the medium
    static bool boundary_hit(const Medium& m, V3 o, V3 d, double tmin, double tmax, double& t_out) {
        bool any = false;
        double closest = tmax;
        Hit tmp;
        for (const Tri& tr : m.boundary)
            if (tri_hit(tr, 0, o, d, tmin, closest, tmp)) { any = true; closest = tmp.t; }
        t_out = closest;
        return any;
    }
    // constant_medium::hit, constant_medium.h:22-58.  ONE draw, taken only when the ray's part inside the boundary
    // overlaps ray_t.
    static bool medium_hit(const Medium& m, V3 o, V3 d, double tmin, double tmax, Rng& rng, double& t_out) {
        double t1, t2;
        if (!boundary_hit(m, o, d, -kInf, kInf, t1)) return false;
        if (!boundary_hit(m, o, d, t1 + 0.0001, kInf, t2)) return false;
        if (t1 < tmin) t1 = tmin;
        if (t2 > tmax) t2 = tmax;
        if (t1 >= t2) return false;
        if (t1 < 0) t1 = 0;
        const double ray_length = len(d);
        const double distance_inside_boundary = (t2 - t1) * ray_length;
        const double xi = rng.next();
        const double hit_distance = m.neg_inv_density * (rng.det ? det_log(xi) : std::log(xi));
        if (hit_distance > distance_inside_boundary) return false;
        t_out = t1 + hit_distance / ray_length;
        return true;
    }
    // The part of [tmin, tmax] a ray spends inside medium m, as constant_medium::hit measures it (the same two boundary
    // hits and clamps as medium_hit above, without the draw): exp(-density * this) is the probability that medium_hit
    // does NOT scatter the ray before tmax -- the transmittance the MIS estimator weighs with.
    static double medium_overlap(const Medium& m, V3 o, V3 d, double tmin, double tmax) {
        double t1, t2;
        if (!boundary_hit(m, o, d, -kInf, kInf, t1)) return 0.0;
        if (!boundary_hit(m, o, d, t1 + 0.0001, kInf, t2)) return 0.0;
        if (t1 < tmin) t1 = tmin;
        if (t2 > tmax) t2 = tmax;
        if (t1 >= t2) return 0.0;
        if (t1 < 0) t1 = 0;
        return (t2 - t1) * len(d);
    }
    // transmittance of all media along o + t d, t in [tmin, tmax]
    double transmittance(V3 o, V3 d, double tmin, double tmax) const {
        double tau = 0.0;
        for (const Medium& m : media) tau += medium_overlap(m, o, d, tmin, tmax) / -m.neg_inv_density;
        return tau > 0.0 ? std::exp(-tau) : 1.0;
    }
    // world.hit with media in the world list.  The reference reaches a medium whenever its BVH walk gets to it, with the
    // bound found so far; here the triangles go first and every medium then sees [tmin, best]: one draw per medium either
    // way, and the same event decides (scatter before the nearer of exit and best), so the outcome has the same
    // distribution whatever the visiting order.  `key`: ORC_RNG_PHILOX draw index of medium 0 for this query (media take
    // key, key + 1, ...; the path streams pass their running counter and get it back advanced by the draws taken).
    uint32_t mat_of(uint32_t prim) const { return (prim & kSpherePrim) ? spheres[prim & ~kSpherePrim].mat : tris[prim].mat; }
    bool extras() const { return !media.empty() || !spheres.empty(); }
    // any sphere within [tmin, tmax]?  (the shadow query's share of the side list)
    bool spheres_occlude(V3 o, V3 d, double tmin, double tmax) const {
        Hit tmp;
        for (size_t k = 0; k < spheres.size(); ++k) if (sphere_hit(spheres[k], uint32_t(k), o, d, tmin, tmax, tmp)) return true;
        return false;
    }
    // triangles, then the sphere prims (no media: they draw random numbers) -- world.hit of the MIS estimator
    bool hit_prims(V3 o, V3 d, double tmin, double tmax, Hit& rec) const {
        bool found = hit(o, d, tmin, tmax, rec);
        double best = found ? rec.t : tmax;
        for (size_t k = 0; k < spheres.size(); ++k)
            if (sphere_hit(spheres[k], uint32_t(k), o, d, tmin, best, rec)) { found = true; best = rec.t; }
        return found;
    }
    bool hit_world(V3 o, V3 d, double tmin, double tmax, Hit& rec, Rng& rng) const {
        bool found = hit(o, d, tmin, tmax, rec);
        double best = found ? rec.t : tmax;
        for (size_t k = 0; k < spheres.size(); ++k)
            if (sphere_hit(spheres[k], uint32_t(k), o, d, tmin, best, rec)) { found = true; best = rec.t; }
        for (size_t k = 0; k < media.size(); ++k) {
            double t;
            if (medium_hit(media[k], o, d, tmin, best, rng, t)) {
                found = true; best = t;
                rec.t = t; rec.p = o + t * d;
                rec.normal = V3{1, 0, 0}; rec.front = true;   // constant_medium.h:52-53: arbitrary
                rec.u = 0; rec.v = 0;
                rec.prim = kMediumPrim | uint32_t(k);
            }
        }
        return found;
    }
};

// ---------------------------------------------------------------- camera.h
struct Cam {  // camera.h:148-197
    orc_camera c;
    int H, sqrt_spp;
    double recip_sqrt_spp;
    V3 center, pixel00, du, dv, background;
    V3 fwd, ul;            // MIS / light tracing only: viewing direction, upper-left corner of the viewport
    double plane_dist_px;  // distance of the image plane in units of one pixel's side
    explicit Cam(const orc_camera& in) : c(in) {
        H = int(c.image_width / c.aspect_ratio);
        H = (H < 1) ? 1 : H;
        sqrt_spp = int(std::sqrt(c.samples_per_pixel));
        recip_sqrt_spp = 1.0 / sqrt_spp;
        V3 lookfrom{c.lookfrom[0], c.lookfrom[1], c.lookfrom[2]}, lookat{c.lookat[0], c.lookat[1], c.lookat[2]};
        V3 vup{c.vup[0], c.vup[1], c.vup[2]};
        background = {c.background[0], c.background[1], c.background[2]};
        center = lookfrom;
        double theta = c.vfov * kPi / 180.0;  // main.h:24-26
        double h = std::tan(theta / 2);
        double vh = 2 * h * c.focus_dist;
        double vw = vh * (double(c.image_width) / H);
        V3 w = unit(lookfrom - lookat);
        V3 u = unit(cross(vup, w));
        V3 v = cross(w, u);
        V3 vu = vw * u;
        V3 vv = vh * -v;
        du = vu / c.image_width;
        dv = vv / H;
        ul = center - (c.focus_dist * w) - vu / 2 - vv / 2;
        pixel00 = ul + 0.5 * (du + dv);
        fwd = -w;
        plane_dist_px = c.focus_dist / len(du);
    }
    void get_ray(int i, int j, int s_i, int s_j, Rng& r, V3& o, V3& d) const {  // camera.h:199-223
        double px = ((s_i + r.next()) * recip_sqrt_spp) - 0.5;
        double py = ((s_j + r.next()) * recip_sqrt_spp) - 0.5;
        V3 ps = pixel00 + ((i + px) * du) + ((j + py) * dv);
        o = center;  // defocus_angle <= 0 (scene_loader.h:463)
        d = ps - o;
        (void)r.next();  // ray_time, camera.h:210 (drawn, unused downstream)
    }
};

struct Vertex {  // camera.h:236-243
    V3 p, normal, throughput, wi, emission, albedo;
    uint32_t mat;
    bool delta, is_light;
    bool iso = false;   // a scatter point inside a constant_medium: the material is its isotropic phase function
    int index = 0;      // position on its subpath (camera: bounce; light: 0 = the emitter, then 1 + bounce)
};

struct Integrator {
    const Scene& sc;
    const Cam& cam;
    Integrator(const Scene& s, const Cam& c) : sc(s), cam(c) {}

    // material.h:133-147
    V3 emitted(const orc_material& m, bool front, double u = 0, double v = 0, V3 p = V3{0, 0, 0}) const {
        if (m.type != ORC_MAT_LIGHT || !front) return {0, 0, 0};
        return sc.colour(m, u, v, p);
    }
    // material.h:60-63; everything but lambertian evaluates to zero (material.h:35-37)
    V3 eval_bsdf(const orc_material& m, V3 albedo) const {   // albedo = tex->value(rec.u, rec.v, rec.p) of that vertex
        if (m.type != ORC_MAT_LAMBERTIAN) return {0, 0, 0};
        return albedo / kPi;
    }
    V3 eval_bsdf(const Vertex& v) const {
        if (v.iso) return v.albedo / (4 * kPi);               // isotropic::evaluate_bsdf, material.h:166-168
        return eval_bsdf(sc.mats[v.mat], v.albedo);
    }

    void trace_path(V3 o, V3 d, int depth, std::vector<Vertex>& path, V3 T, V3* bg, Rng& rng, int index0 = 0) const {  // camera.h:325-370
        for (int bounce = 0; bounce < depth; ++bounce) {
            Hit rec;
            if (!(sc.extras() ? sc.hit_world(o, d, 0.001, kInf, rec, rng) : sc.hit(o, d, 0.001, kInf, rec))) {
                if (bg) *bg = *bg + T * cam.background;
                break;
            }
            if (rec.prim & kMediumPrim) {   // scattered inside a medium: isotropic (material.h:149-171)
                const Medium& md = sc.media[rec.prim & ~kMediumPrim];
                Vertex vx;
                vx.p = rec.p; vx.normal = rec.normal; vx.mat = 0;
                vx.throughput = T;
                vx.wi = unit(-d);
                vx.emission = V3{0, 0, 0};
                vx.albedo = md.albedo;
                vx.delta = false; vx.is_light = false; vx.iso = true; vx.index = index0 + bounce;
                path.push_back(vx);
                // scatter: attenuation = albedo, sphere_pdf (pdf.h:18-29): generate() = random_unit_vector, value = 1/(4 pi);
                // scattering_pdf = 1/(4 pi) (material.h:161-164); camera.h:361-368
                V3 sd = random_unit_vector(rng);
                double pdf_value = 1 / (4 * kPi);
                double spdf = 1 / (4 * kPi);
                T = T * md.albedo * spdf / pdf_value;
                o = rec.p; d = sd;
                continue;
            }
            const orc_material& m = sc.mats[sc.mat_of(rec.prim)];
            Vertex vx;
            vx.index = index0 + bounce;
            vx.p = rec.p; vx.normal = rec.normal; vx.mat = sc.mat_of(rec.prim);
            vx.throughput = T;
            vx.wi = unit(-d);
            vx.emission = emitted(m, rec.front, rec.u, rec.v, rec.p);
            vx.albedo = (m.type == ORC_MAT_LAMBERTIAN) ? sc.colour(m, rec.u, rec.v, rec.p) : V3{0, 0, 0};
            vx.delta = (m.type == ORC_MAT_METAL || m.type == ORC_MAT_DIELECTRIC);
            vx.is_light = (m.type == ORC_MAT_LIGHT);
            path.push_back(vx);

            if (m.type == ORC_MAT_LIGHT) break;  // material.h:22-24: scatter() false
            V3 att = (m.type == ORC_MAT_LAMBERTIAN) ? vx.albedo : V3{m.rgb[0], m.rgb[1], m.rgb[2]};
            if (m.type == ORC_MAT_METAL) {  // material.h:73-83
                V3 refl = reflect(d, rec.normal);
                refl = unit(refl) + (m.param * random_unit_vector(rng));
                T = T * att;
                o = rec.p; d = refl;
                continue;
            }
            if (m.type == ORC_MAT_DIELECTRIC) {  // material.h:96-116
                double ri = rec.front ? (1.0 / m.param) : m.param;
                V3 ud = unit(d);
                double cos_theta = std::fmin(dot(-ud, rec.normal), 1.0);
                double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
                bool cannot_refract = ri * sin_theta > 1.0;
                V3 dir;
                bool refl = cannot_refract;
                if (!refl) {  // material.h:125-130 Schlick
                    double r0 = (1 - ri) / (1 + ri);
                    r0 = r0 * r0;
                    double rf = r0 + (1 - r0) * (rng.det ? det_pow5(1 - cos_theta) : std::pow((1 - cos_theta), 5));
                    refl = rf > rng.next();
                }
                dir = refl ? reflect(ud, rec.normal) : refract(ud, rec.normal, ri);
                T = T * V3{1.0, 1.0, 1.0};
                o = rec.p; d = dir;
                continue;
            }
            // lambertian: material.h:47-58, pdf.h:32-47, camera.h:361-368
            Onb uvw(rec.normal);
            V3 sd = uvw.transform(random_cosine_direction(rng));
            double pdf_value = std::fmax(0, dot(unit(sd), uvw.a2) / kPi);
            if (pdf_value <= 0) break;
            double ct = dot(rec.normal, unit(sd));
            double spdf = ct < 0 ? 0 : ct / kPi;
            T = T * att * spdf / pdf_value;
            o = rec.p; d = sd;
        }
    }

    bool build_light_path(int depth, std::vector<Vertex>& path, Rng& rng) const {  // camera.h:372-418
        if (depth <= 0 || sc.lights.empty()) return false;
        // triangle.h:199-224
        double total = 0.0;
        for (uint32_t id : sc.lights) total += sc.tris[id].area;
        if (total <= 0.0) return false;
        double pick = rng.next(0.0, total);
        const Tri* chosen = &sc.tris[sc.lights.back()];
        double accum = 0.0;
        for (uint32_t id : sc.lights) {
            accum += sc.tris[id].area;
            if (pick <= accum) { chosen = &sc.tris[id]; break; }
        }
        double u = rng.next(), v = rng.next();  // triangle.h:107-119
        if (u + v > 1.0) { u = 1.0 - u; v = 1.0 - v; }
        V3 pos = chosen->v0 + (u * chosen->e1) + (v * chosen->e2);
        V3 n = chosen->normal;
        double pdf = 1.0 / total;
        const orc_material& m = sc.mats[chosen->mat];
        V3 emission = emitted(m, true, 0, 0, pos);   // camera.h:389-392: u = v = 0
        if (len2(emission) <= 0) return false;

        Vertex ev;
        ev.p = pos; ev.normal = n; ev.mat = chosen->mat;
        ev.throughput = V3{1.0, 1.0, 1.0} / std::max(pdf, 1e-8);
        ev.wi = n; ev.emission = emission; ev.albedo = V3{0, 0, 0}; ev.delta = false; ev.is_light = true;
        path.push_back(ev);

        V3 dir = Onb(n).transform(random_cosine_direction(rng));
        V3 du = unit(dir);
        double cos_theta = std::max(0.0, dot(n, du));
        if (cos_theta <= 0) return true;
        double pdf_dir = std::max(cos_theta / kPi, 1e-8);
        V3 T = ev.throughput * emission * (cos_theta / pdf_dir);
        trace_path(pos + (0.001 * n), du, depth - 1, path, T, nullptr, rng, 1);
        return true;
    }

    bool visible(V3 a, V3 b) const {  // camera.h:425-438
        V3 dir = b - a;
        double dist = len(dir);
        if (dist <= 0) return false;
        V3 du = dir / dist;
        double max_t = dist - 0.001;
        if (max_t <= 0) return false;
        Hit tmp;
        ++tl_stats.shadow;
        if (sc.spheres_occlude(a + (0.001 * du), du, 0.001, max_t)) return false;   // the sphere prims of the side list (none: a no-op)
        return !sc.hit(a + (0.001 * du), du, 0.001, max_t, tmp);
    }
    // the same with media in the world: a medium between the two points scatters the shadow ray with the probability
    // constant_medium::hit gives it (one draw per medium crossed).  ORC_RNG_PHILOX: the draw of medium k for the pair
    // (camera vertex ci, light vertex li) is draw ((ci * 128 + li) * 16 + k) of the sample's shadow stream (stream 2).
    bool visible_media(V3 a, V3 b, Rng& shadow_rng, int ci, int li) const {
        V3 dir = b - a;
        double dist = len(dir);
        if (dist <= 0) return false;
        V3 du = dir / dist;
        double max_t = dist - 0.001;
        if (max_t <= 0) return false;
        ++tl_stats.shadow;
        const V3 so = a + (0.001 * du);
        for (size_t k = 0; k < sc.media.size(); ++k) {
            if (!shadow_rng.mt) shadow_rng.draws = uint32_t((ci * 128 + li) * 16 + int(k));
            double t;
            if (Scene::medium_hit(sc.media[k], so, du, 0.001, max_t, shadow_rng, t)) return false;
        }
        if (sc.spheres_occlude(so, du, 0.001, max_t)) return false;
        Hit tmp;
        return !sc.hit(so, du, 0.001, max_t, tmp);
    }

    V3 connect(const Vertex& cv, const Vertex& lv, Rng* shadow_rng = nullptr) const {  // camera.h:440-475
        if (cv.delta || lv.delta) return {0, 0, 0};
        V3 dir = lv.p - cv.p;
        double dist2 = len2(dir);
        if (dist2 <= 0) return {0, 0, 0};
        V3 du = unit(dir);
        double cos_cam = std::fabs(dot(cv.normal, du));
        double cos_light = std::fabs(dot(lv.normal, -du));
        if (cos_cam <= 0 || cos_light <= 0) return {0, 0, 0};
        if (!(sc.extras() ? visible_media(cv.p, lv.p, *shadow_rng, cv.index, lv.index) : visible(cv.p, lv.p))) return {0, 0, 0};
        V3 f_cam = eval_bsdf(cv);
        if (len2(f_cam) <= 0) return {0, 0, 0};
        V3 f_light = lv.is_light ? lv.emission : eval_bsdf(lv);
        if (len2(f_light) <= 0) return {0, 0, 0};
        V3 contrib = cv.throughput * f_cam;
        contrib = contrib * lv.throughput * f_light;
        contrib = contrib * (cos_cam * cos_light) / dist2;
        return contrib;
    }

    // ------------------------------------------------------------ weighting = MIS (not in the reference: SURVEY F3, 8f-3)
    // The same two random walks as above (same draws, same vertices), combined with balance-heuristic weights over
    // every way of building the path: camera path hits the light (s = 0), connection to the light subpath's emitter
    // point (s = 1), vertex connections (s >= 2) and light tracing onto the film (t = 1, a SPLAT into another pixel).
    // Weights are kept in the recursive form of Georgiev's vertex-connection formulation (two running quantities
    // per subpath, dVCM and dVC), which makes every connection O(1).  Differences from the reference's unweighted
    // sum that an unbiased estimator needs: a lambertian surface reflects only to the side the path arrived on
    // (material.h:47-52 samples about the face-forwarded normal; the reference's connection uses |cos| on both
    // sides), emitters are one-sided absorbers on both walks (material.h:140, camera.h:409), a full path holds at
    // most max_depth surface vertices (what path_trace_color's depth counter allows, camera.h:255-292), and the
    // shadow segment stops short of the far surface instead of on it.  Metal stays a delta vertex even
    // when fuzzy, as material.h:85 declares it.
    // ORC_MIS_ONLY=<mask> (test hook): bit 0 s=0, 1 s=1, 2 s>=2, 3 t=1 -- the chosen strategies with weight 1, the rest
    // with weight 0, so each family of strategies can be checked on its own against the reference's path tracer
    static int mis_only() { static const int m = [] { const char* e = std::getenv("ORC_MIS_ONLY"); return e ? std::atoi(e) : 0; }(); return m; }
    static double pick_w(double w, int bit) { const int m = mis_only(); return m == 0 ? w : ((m >> bit) & 1 ? 1.0 : 0.0); }

    struct MisVertex {
        V3 p, n, T, albedo;        // position, normal facing the arriving path, throughput up to here
        double dVCM, dVC, cos_in;
        int len;                   // surface vertices of the subpath up to and including this one
        bool medium = false;       // a scatter point inside a medium: isotropic phase function, no cosine
        // pdf per solid angle of scattering back towards the previous vertex, the BSDF value, and the cosine of a direction
        double rev_pdf() const { return medium ? 1 / (4 * kPi) : cos_in / kPi; }
        V3 f() const { return medium ? albedo / (4 * kPi) : albedo / kPi; }
        double cosine(V3 du) const { return medium ? 1.0 : dot(n, du); }
        double dir_pdf(double c) const { return medium ? 1 / (4 * kPi) : c / kPi; }
    };
    // transmittance of the media along the segment visible_mis tests (1 without media)
    double tr_mis(V3 a, V3 b) const {
        if (sc.media.empty()) return 1.0;
        V3 dir = b - a;
        double dist = len(dir);
        V3 du = dir / dist;
        double max_t = dist - 0.001 - 1e-6 * (1.0 + dist);
        return sc.transmittance(a + (0.001 * du), du, 0.001, max_t);
    }

    bool visible_mis(V3 a, V3 b) const {
        V3 dir = b - a;
        double dist = len(dir);
        if (dist <= 0) return false;
        V3 du = dir / dist;
        double max_t = dist - 0.001 - 1e-6 * (1.0 + dist);   // stops short of b's own surface
        if (max_t <= 0.001) return false;
        Hit tmp;
        ++tl_stats.shadow;
        if (sc.spheres_occlude(a + (0.001 * du), du, 0.001, max_t)) return false;   // the sphere prims of the side list
        return !sc.hit(a + (0.001 * du), du, 0.001, max_t, tmp);
    }

    double light_area() const {
        double total = 0.0;
        for (uint32_t id : sc.lights) total += sc.tris[id].area;
        return total;
    }

    // one subpath; `verts` receives its connectable (lambertian) vertices.  camera: adds the s = 0 terms to *result.
    void walk_mis(V3 o, V3 d, int depth, int len0, V3 T, double dVCM, double dVC, bool camera, double pdfA_light,
                  std::vector<MisVertex>& verts, V3* result, Rng& rng) const {
        int length = len0;   // surface vertices already on the subpath (light subpath: the emitter = 1)
        for (int bounce = 0; bounce < depth; ++bounce) {
            Hit rec;
            if (!(sc.media.empty() ? sc.hit_prims(o, d, 0.001, kInf, rec) : sc.hit_world(o, d, 0.001, kInf, rec, rng))) {
                if (camera) *result = *result + T * cam.background;
                break;
            }
            ++length;
            const double dist2 = rec.t * rec.t * len2(d);
            // media: the pdf of getting this far holds the transmittance of the segment (and, for a scatter point, the density,
            // which every strategy that makes this vertex shares): the segment's 1/dist^2 becomes Tr/dist^2 in every pdf
            const double tr = sc.media.empty() ? 1.0 : sc.transmittance(o, d, 0.001, rec.t);
            if (rec.prim & kMediumPrim) {   // isotropic scatter point: connectable, no cosine, 1/(4 pi) both ways
                const Medium& md = sc.media[rec.prim & ~kMediumPrim];
                dVCM *= dist2 / tr;
                MisVertex v;
                v.p = rec.p; v.n = rec.normal; v.T = T; v.albedo = md.albedo; v.dVCM = dVCM; v.dVC = dVC; v.cos_in = 1.0; v.len = length;
                v.medium = true;
                verts.push_back(v);
                V3 sd = random_unit_vector(rng);
                const double dirPdfW = 1 / (4 * kPi), revPdfW = 1 / (4 * kPi);
                T = T * md.albedo * revPdfW / dirPdfW;   // attenuation * scattering_pdf / pdf, as trace_path writes it
                dVC = (1.0 / dirPdfW) * (dVC * revPdfW + dVCM);
                dVCM = 1.0 / dirPdfW;
                o = rec.p; d = sd;
                continue;
            }
            const orc_material& m = sc.mats[sc.mat_of(rec.prim)];
            const double cos_in = std::fabs(dot(rec.normal, unit(-d)));
            dVCM *= sc.media.empty() ? dist2 : dist2 / tr;
            dVCM /= cos_in; dVC /= cos_in;
            if (m.type == ORC_MAT_LIGHT) {           // absorbs; a camera path that arrives on the emitting side sees it
                if (camera && rec.front) {
                    V3 Le = sc.colour(m, rec.u, rec.v, rec.p);
                    double w = 1.0;
                    if (length > 1) {
                        const double directPdfA = pdfA_light, emissionPdfW = pdfA_light * cos_in / kPi;
                        w = pick_w(1.0 / (1.0 + directPdfA * dVCM + emissionPdfW * dVC), 0);
                    }
                    *result = *result + w * (T * Le);
                }
                break;
            }
            V3 att = (m.type == ORC_MAT_LAMBERTIAN) ? sc.colour(m, rec.u, rec.v, rec.p) : V3{m.rgb[0], m.rgb[1], m.rgb[2]};
            if (m.type == ORC_MAT_METAL || m.type == ORC_MAT_DIELECTRIC) {
                V3 dir;
                if (m.type == ORC_MAT_METAL) {
                    V3 refl = reflect(d, rec.normal);
                    dir = unit(refl) + (m.param * random_unit_vector(rng));
                    T = T * att;
                } else {
                    double ri = rec.front ? (1.0 / m.param) : m.param;
                    V3 ud = unit(d);
                    double cos_theta = std::fmin(dot(-ud, rec.normal), 1.0);
                    double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
                    bool refl = ri * sin_theta > 1.0;
                    if (!refl) {
                        double r0 = (1 - ri) / (1 + ri);
                        r0 = r0 * r0;
                        double rf = r0 + (1 - r0) * (rng.det ? det_pow5(1 - cos_theta) : std::pow((1 - cos_theta), 5));
                        refl = rf > rng.next();
                    }
                    dir = refl ? reflect(ud, rec.normal) : refract(ud, rec.normal, ri);
                    T = T * V3{1.0, 1.0, 1.0};
                }
                const double cos_out = std::fabs(dot(rec.normal, unit(dir)));
                dVCM = 0.0; dVC *= cos_out;          // a delta vertex cannot be connected to
                o = rec.p; d = dir;
                continue;
            }
            MisVertex v;
            v.p = rec.p; v.n = rec.normal; v.T = T; v.albedo = att; v.dVCM = dVCM; v.dVC = dVC; v.cos_in = cos_in; v.len = length;
            verts.push_back(v);
            Onb uvw(rec.normal);
            V3 sd = uvw.transform(random_cosine_direction(rng));
            double pdf_value = std::fmax(0, dot(unit(sd), uvw.a2) / kPi);
            if (pdf_value <= 0) break;
            double ct = dot(rec.normal, unit(sd));
            double spdf = ct < 0 ? 0 : ct / kPi;
            T = T * att * spdf / pdf_value;
            const double dirPdfW = pdf_value, revPdfW = cos_in / kPi;
            dVC = (ct / dirPdfW) * (dVC * revPdfW + dVCM);
            dVCM = 1.0 / dirPdfW;
            o = rec.p; d = sd;
        }
    }

    // one sample: returns what lands in its own pixel; light-tracing splats go through `splat(pixel, colour)`
    template <class Splat>
    V3 sample_mis(V3 o, V3 d, int depth, Rng& cam_rng, Rng& light_rng, int W, int H, Splat splat) const {
        const double n_light = double(W) * double(H);   // light subpaths per pass over the frame: every one may splat anywhere
        V3 result{0, 0, 0};
        std::vector<MisVertex> cp, lp;
        const double total = light_area();
        const double pdfA = total > 0 ? 1.0 / total : 0.0;
        {   // camera subpath: the pdf of a primary ray per unit solid angle, for an image plane measured in pixels
            const double c = dot(unit(d), cam.fwd);
            const double cameraPdfW = cam.plane_dist_px * cam.plane_dist_px / (c * c * c);
            walk_mis(o, d, depth, 0, {1, 1, 1}, n_light / cameraPdfW, 0.0, true, pdfA, cp, &result, cam_rng);
        }
        // light subpath: build_light_path's emitter and first direction (camera.h:372-414), same draws
        if (depth <= 0 || sc.lights.empty() || total <= 0.0) return result;
        double pick = light_rng.next(0.0, total);
        const Tri* chosen = &sc.tris[sc.lights.back()];
        double accum = 0.0;
        for (uint32_t id : sc.lights) {
            accum += sc.tris[id].area;
            if (pick <= accum) { chosen = &sc.tris[id]; break; }
        }
        double u = light_rng.next(), v = light_rng.next();
        if (u + v > 1.0) { u = 1.0 - u; v = 1.0 - v; }
        const V3 y0 = chosen->v0 + (u * chosen->e1) + (v * chosen->e2), n0 = chosen->normal;
        const V3 Le = emitted(sc.mats[chosen->mat], true, 0, 0, y0);
        if (len2(Le) <= 0) return result;
        {
            V3 dir = Onb(n0).transform(random_cosine_direction(light_rng));
            V3 du = unit(dir);
            double cos_theta = std::max(0.0, dot(n0, du));
            if (cos_theta > 0) {
                const double emissionPdfW = pdfA * std::max(cos_theta / kPi, 1e-8);
                V3 T = (V3{1.0, 1.0, 1.0} / std::max(pdfA, 1e-8)) * Le * (cos_theta / std::max(cos_theta / kPi, 1e-8));
                walk_mis(y0 + (0.001 * n0), du, depth - 1, 1, T, pdfA / emissionPdfW, cos_theta / emissionPdfW, false, pdfA, lp, nullptr, light_rng);
            }
        }
        for (const MisVertex& c : cp) {
            const V3 fc = c.f();
            // s = 1: the emitter point itself
            if (c.len + 1 <= depth) {
                V3 dir = y0 - c.p;
                double dist2 = len2(dir);
                if (dist2 > 0) {
                    V3 du = unit(dir);
                    const double cosToLight = c.cosine(du), cosAtLight = dot(n0, -du);
                    if (cosToLight > 0 && cosAtLight > 0 && visible_mis(c.p, y0)) {
                        const double tr = tr_mis(c.p, y0);
                        const double directPdfW = pdfA * dist2 / cosAtLight, emissionPdfW = pdfA * cosAtLight / kPi;
                        const double wLight = (c.dir_pdf(cosToLight) * tr) / directPdfW;
                        const double wCamera = (emissionPdfW * cosToLight * tr / (directPdfW * cosAtLight)) * (c.dVCM + c.dVC * c.rev_pdf());
                        const double w = pick_w(1.0 / (wLight + 1.0 + wCamera), 1);
                        result = result + (w * cosToLight * tr / directPdfW) * (c.T * fc * Le);
                    }
                }
            }
            // s >= 2
            for (const MisVertex& l : lp) {
                if (c.len + l.len > depth) continue;
                V3 dir = l.p - c.p;
                double dist2 = len2(dir);
                if (dist2 <= 0) continue;
                V3 du = unit(dir);
                const double cosCamera = c.cosine(du), cosLight = l.cosine(-du);
                if (cosCamera <= 0 || cosLight <= 0) continue;
                if (!visible_mis(c.p, l.p)) continue;
                const double tr = tr_mis(c.p, l.p);
                const double camDirPdfA = c.dir_pdf(cosCamera) * cosLight * tr / dist2, lightDirPdfA = l.dir_pdf(cosLight) * cosCamera * tr / dist2;
                const double wLight = camDirPdfA * (l.dVCM + l.dVC * l.rev_pdf());
                const double wCamera = lightDirPdfA * (c.dVCM + c.dVC * c.rev_pdf());
                const double w = pick_w(1.0 / (wLight + 1.0 + wCamera), 2);
                result = result + (w * cosLight * cosCamera * tr / dist2) * (c.T * fc * l.T * l.f());
            }
        }
        // t = 1: every light-subpath vertex seen from the camera lands on the film, wherever that is
        for (const MisVertex& l : lp) {
            V3 dir = cam.center - l.p;
            double dist2 = len2(dir);
            if (dist2 <= 0) continue;
            V3 du = unit(dir);
            const double cosToCamera = l.cosine(du), cosAtCamera = dot(cam.fwd, -du);
            if (cosToCamera <= 0 || cosAtCamera <= 0) continue;
            // where the line camera -> vertex crosses the viewport, in pixel coordinates from its upper-left corner
            V3 P = cam.center + (cam.c.focus_dist / cosAtCamera) * (-du);
            V3 rel = P - cam.ul;
            const double xc = dot(rel, cam.du) / len2(cam.du), yc = dot(rel, cam.dv) / len2(cam.dv);
            if (!(xc >= 0 && yc >= 0 && xc < W && yc < H)) continue;
            if (!visible_mis(l.p, cam.center)) continue;
            const double ipcd = cam.plane_dist_px / cosAtCamera;
            const double imageToSurface = (ipcd * ipcd / cosAtCamera) * cosToCamera * tr_mis(l.p, cam.center) / dist2;
            const double wLight = (imageToSurface / n_light) * (l.dVCM + l.dVC * l.rev_pdf());
            const double w = pick_w(1.0 / (wLight + 1.0), 3);
            splat(int(yc) * W + int(xc), (w * imageToSurface / n_light) * (l.T * l.f()));
        }
        return result;
    }

    V3 sample(V3 o, V3 d, int depth, Rng& cam_rng, Rng& light_rng) const {  // camera.h:294-323
        std::vector<Vertex> cp, lp;
        cp.reserve(depth);
        V3 bg{0, 0, 0};
        trace_path(o, d, depth, cp, {1, 1, 1}, &bg, cam_rng);
        V3 result = bg;
        for (const auto& vx : cp)
            if (!vx.delta && len2(vx.emission) > 0) result = result + vx.throughput * vx.emission;
        lp.reserve(depth);
        if (!build_light_path(depth, lp, light_rng)) return result;
        Rng shadow = cam_rng;   // media only: the sample's shadow stream (MT mode: the one global stream, as the reference)
        if (!shadow.mt) { shadow.stream = 2; shadow.draws = 0; }
        for (const auto& cv : cp)
            for (const auto& lv : lp) result = result + connect(cv, lv, &shadow);
        return result;
    }
};

template <class F>
void parallel_for(uint64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) { f(0, n); flush_stats(); return; }
    std::vector<std::thread> pool;
    uint64_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        uint64_t a = std::min<uint64_t>(n, t * chunk), b = std::min<uint64_t>(n, a + chunk);
        if (a < b) pool.emplace_back([=] { f(a, b); flush_stats(); });
    }
    for (auto& th : pool) th.join();
}

}  // namespace

extern "C" {

void* port_scene_create(const double* v, const uint32_t* mat_id, uint64_t n, const orc_material* mats, uint32_t n_mats) {
    auto* s = new Scene;
    s->mats.assign(mats, mats + n_mats);
    s->tris.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        const double* p = v + 9 * i;
        s->tris.push_back(make_tri({p[0], p[1], p[2]}, {p[3], p[4], p[5]}, {p[6], p[7], p[8]}, mat_id[i]));
        if (mats[mat_id[i]].type == ORC_MAT_LIGHT) s->lights.push_back(uint32_t(i));
    }
    if (s->lights.empty())  // main.cpp:67
        for (uint64_t i = 0; i < n; ++i) s->lights.push_back(uint32_t(i));
    s->order.resize(n);
    for (uint64_t i = 0; i < n; ++i) s->order[i] = uint32_t(i);
    s->nodes.reserve(size_t(1.4 * n) + 4);
    s->root = s->build(0, n);
    g_stats[6] = s->nodes.size();
    return s;
}
void* port_scene_create_tex(const double* v, const uint32_t* mat_id, uint64_t n, const orc_material* mats, uint32_t n_mats,
                            const orc_texture* tex, uint32_t n_tex, const void* data, uint64_t data_bytes) {
    auto* s = static_cast<Scene*>(port_scene_create(v, mat_id, n, mats, n_mats));
    s->tex.assign(tex, tex + n_tex);
    const unsigned char* d = static_cast<const unsigned char*>(data);
    s->tex_data.assign(d, d + data_bytes);
    return s;
}
void port_texture_value(void* p, uint32_t k, uint64_t n, const double* uvp, double* rgb) {
    auto* s = static_cast<Scene*>(p);
    orc_material m{};
    m.type = ORC_MAT_LAMBERTIAN; m.tex = k + 1;
    for (uint64_t i = 0; i < n; ++i) {
        V3 c = s->colour(m, uvp[5 * i], uvp[5 * i + 1], {uvp[5 * i + 2], uvp[5 * i + 3], uvp[5 * i + 4]});
        rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    }
}
void port_scene_destroy(void* p) { delete static_cast<Scene*>(p); }

// media: n x {density, r, g, b, first boundary triangle, boundary triangle count} as doubles; boundary: triangles x 9 doubles
void port_scene_set_media(void* p, const double* media6, uint32_t n_media, const double* boundary) {
    auto* s = static_cast<Scene*>(p);
    s->media.clear();
    for (uint32_t k = 0; k < n_media; ++k) {
        const double* m = media6 + 6 * k;
        Medium md;
        md.neg_inv_density = -1 / m[0];   // constant_medium.h:12
        md.albedo = V3{m[1], m[2], m[3]};
        const uint64_t first = uint64_t(m[4]), cnt = uint64_t(m[5]);
        for (uint64_t i = first; i < first + cnt; ++i) {
            const double* q = boundary + 9 * i;
            md.boundary.push_back(make_tri({q[0], q[1], q[2]}, {q[3], q[4], q[5]}, {q[6], q[7], q[8]}, 0));
        }
        s->media.push_back(md);
    }
}
double port_det_log(double x) { return det_log(x); }
// spheres: n x {cx, cy, cz, radius, material index} as doubles
void port_scene_set_spheres(void* p, const double* spheres5, uint32_t n) {
    auto* s = static_cast<Scene*>(p);
    s->spheres.clear();
    for (uint32_t k = 0; k < n; ++k) {
        const double* q = spheres5 + 5 * k;
        s->spheres.push_back(Sphere{V3{q[0], q[1], q[2]}, q[3], uint32_t(q[4])});
    }
}
uint64_t port_bvh_node_count(void* p) { return static_cast<Scene*>(p)->nodes.size(); }
void port_bvh_leaf_order(void* p, uint32_t* order) {
    auto* s = static_cast<Scene*>(p);
    std::copy(s->order.begin(), s->order.end(), order);
}

void port_trace(void* p, const double* rays, uint64_t n, uint32_t* prim, double* t, double* u, double* v, int nthreads) {
    auto* s = static_cast<Scene*>(p);
    parallel_for(n, nthreads, [=](uint64_t a, uint64_t b) {
        for (uint64_t i = a; i < b; ++i) {
            const double* q = rays + 8 * i;
            Hit rec;
            bool h = s->hit({q[0], q[1], q[2]}, {q[3], q[4], q[5]}, q[6], q[7], rec);
            double best = h ? rec.t : q[7];   // the sphere prims of the side list, as hit_world asks them (media draw numbers: not here)
            for (size_t k = 0; k < s->spheres.size(); ++k)
                if (sphere_hit(s->spheres[k], uint32_t(k), {q[0], q[1], q[2]}, {q[3], q[4], q[5]}, q[6], best, rec)) { h = true; best = rec.t; }
            if (prim) prim[i] = h ? rec.prim : 0xFFFFFFFFu;
            if (t) t[i] = h ? rec.t : 0.0;
            if (u) u[i] = h ? rec.u : 0.0;
            if (v) v[i] = h ? rec.v : 0.0;
        }
    });
}

void port_visible(void* p, const double* a, const double* b, uint64_t n, uint8_t* out, int nthreads) {
    auto* s = static_cast<Scene*>(p);
    orc_camera oc;
    std::memset(&oc, 0, sizeof oc);
    oc.aspect_ratio = 1; oc.image_width = 1; oc.samples_per_pixel = 1; oc.vfov = 90; oc.focus_dist = 10;
    oc.lookat[2] = -1; oc.vup[1] = 1;
    Cam cam(oc);
    Integrator integ(*s, cam);
    parallel_for(n, nthreads, [&, a, b, out](uint64_t i0, uint64_t i1) {
        for (uint64_t i = i0; i < i1; ++i)
            out[i] = integ.visible({a[3 * i], a[3 * i + 1], a[3 * i + 2]}, {b[3 * i], b[3 * i + 1], b[3 * i + 2]}) ? 1 : 0;
    });
}

int port_render_ex(void* p, const orc_camera* c, int rng_mode, uint64_t seed, int row0, int row1, double* film, int nthreads,
                   int weighting) {
    auto* s = static_cast<Scene*>(p);
    Cam cam(*c);
    Integrator integ(*s, cam);
    const int W = c->image_width, H = cam.H, n = cam.sqrt_spp;
    row0 = std::max(0, row0);
    row1 = std::min(H, row1);
    if (row1 <= row0) return H;
    if (rng_mode == ORC_RNG_MT19937) nthreads = 1;  // one global stream: main.h:28-32
    std::mt19937 mt;  // default seed, as the reference's static generator (main.h:28-32) ...
    if (seed) mt.seed(uint32_t(seed));  // ... unless the caller asks for an independent stream (convergence tests)
    std::uniform_real_distribution<double> dist(0.0, 1.0);
    std::mutex splat_mu;   // MIS: light-tracing splats land in other threads' rows
    auto splat = [&](int pix, V3 cval) {
        std::lock_guard<std::mutex> lock(splat_mu);
        film[3 * size_t(pix)] += cval.x; film[3 * size_t(pix) + 1] += cval.y; film[3 * size_t(pix) + 2] += cval.z;
    };
    parallel_for(uint64_t(row1 - row0), nthreads, [&](uint64_t a, uint64_t b) {
        Rng cr, lr;
        if (rng_mode == ORC_RNG_MT19937) { cr.mt = lr.mt = &mt; cr.dist = lr.dist = &dist; }
        for (int j = row0 + int(a); j < row0 + int(b); ++j)
            for (int i = 0; i < W; ++i) {  // camera.h:116-125
                V3 px{0, 0, 0};
                for (int s_j = 0; s_j < n; ++s_j)
                    for (int s_i = 0; s_i < n; ++s_i) {
                        if (rng_mode != ORC_RNG_MT19937) {
                            uint64_t sample = (uint64_t(j) * W + i) * uint64_t(n * n) + uint64_t(s_j * n + s_i);
                            cr = Rng(); lr = Rng();
                            cr.seed = lr.seed = seed; cr.sample = lr.sample = sample;
                            cr.stream = 0; lr.stream = 1;
                            cr.det = lr.det = true;
                        }
                        V3 o, d;
                        cam.get_ray(i, j, s_i, s_j, cr, o, d);
                        px = px + (weighting == ORC_WEIGHT_MIS ? integ.sample_mis(o, d, c->max_depth, cr, lr, W, H, splat)
                                                              : integ.sample(o, d, c->max_depth, cr, lr));
                    }
                if (weighting == ORC_WEIGHT_MIS) splat(j * W + i, px);   // the film already holds splats: add, under the lock
                else {
                    double* out = film + 3 * (size_t(j) * W + i);
                    out[0] = px.x; out[1] = px.y; out[2] = px.z;
                }
            }
    });
    return H;
}

int port_render(void* p, const orc_camera* c, int rng_mode, uint64_t seed, int row0, int row1, double* film, int nthreads) {
    return port_render_ex(p, c, rng_mode, seed, row0, row1, film, nthreads, ORC_WEIGHT_REF_SUM);
}

void port_stats_get(orc_stats* o) {
    flush_stats();
    o->rays = g_stats[0]; o->node_visits = g_stats[1]; o->aabb_tests = g_stats[2]; o->aabb_hits = g_stats[3];
    o->tri_tests = g_stats[4]; o->tri_hits = g_stats[5]; o->nodes_built = g_stats[6]; o->shadow_rays = g_stats[7];
}
void port_stats_reset(void) {
    tl_stats = Stats();
    for (int i = 0; i < 8; ++i) if (i != 6) g_stats[i] = 0;
}

// test hook: the PHILOX-mode deterministic sin/cos (accuracy is pinned in tests/test_oracle.py)
void port_det_sincos(double phi, double* s, double* c) { det_sincos(phi, *s, *c); }

void port_colors_to_rgb8(const double* film, int W, int H, int spp, uint8_t* out) {  // wpng.h:14-35
    const double scale = 1.0 / std::max(1, spp);
    auto clampd = [](double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); };
    for (size_t i = 0; i < size_t(W) * H; ++i)
        for (int c = 0; c < 3; ++c) {
            double x = film[3 * i + c] * scale;
            out[3 * i + c] = static_cast<uint8_t>(256.0 * std::sqrt(clampd(x, 0.0, 0.999)));
        }
}

}  // extern "C"
