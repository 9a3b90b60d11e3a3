// GPU-resident scene: flattened BVHs, triangle records, material table, light CDF.
// Layout notes (DESIGN.md "Data layout in HBM"): every record is a multiple of 16 bytes and is read
// with 16-byte vector loads.
#pragma once
#include "dmath.cuh"

namespace bdpt {

// ---- reference-topology BVH (replaces src/acceleration/bvh.h + aabb.h), fp64 boxes ------------
// 64 B = 4 x 16-byte loads. Internal: left/right are node indices. Leaf (span 1 or 2, bvh.h:35-39):
// left/right are triangle (prim) IDs; span 1 has left == right and the triangle is tested twice,
// exactly as the reference does.
struct __align__(16) RefNode {
    double lo_x, hi_x, lo_y, hi_y, lo_z, hi_z;  // aabb.h:8 x,y,z intervals (padded per aabb.h:81-88)
    int32_t left, right;
    uint32_t leaf;                              // 1 = children are triangles
    uint32_t pad;
};
static_assert(sizeof(RefNode) == 64, "RefNode must be 64 bytes");

// fp64 triangle for intersection: triangle.h:21-31 (v0, edge1, edge2), 80 B = 5 x 16-byte loads.
struct __align__(16) TriGeom {
    double v0x, v0y, v0z, e1x, e1y, e1z, e2x, e2y, e2z;
    uint32_t mat;
    uint32_t pad;
};
static_assert(sizeof(TriGeom) == 80, "TriGeom must be 80 bytes");

// shading side-table: unit geometric normal (triangle.h:32) + material, 32 B
struct __align__(16) TriShade {
    double nx, ny, nz;
    uint32_t mat;
    uint32_t pad;
};
static_assert(sizeof(TriShade) == 32, "TriShade must be 32 bytes");

// material table entry (material.h:42-147), 48 B
struct __align__(16) Mat {
    double r, g, b;
    double param;
    uint32_t type;
    uint32_t tex;      // 0: (r, g, b) is the colour; k > 0: texture k - 1
    uint32_t pad[2];
};
static_assert(sizeof(Mat) == 48, "Mat must be 48 bytes");

// ---- fast BVH: binned-SAH binary tree, both child boxes in the parent, fp32 conservative ------
// 64 B. child >= 0: node index. child < 0: leaf, ~child = (first << 3) | (count - 1), first = index into
// the leaf-ordered prim list, 1..8 triangles.
struct __align__(16) FastNode {
    float l_lo[3], l_hi[3];
    float r_lo[3], r_hi[3];
    int32_t left, right;
    uint32_t pad[2];
};
static_assert(sizeof(FastNode) == 64, "FastNode must be 64 bytes");

// ---- 4-wide form of the fast BVH (collapsed from the binary tree, fast_bvh.cu): the walk of a ray is a chain of
// dependent node fetches, and a 4-wide node halves the length of that chain.  128 B = 8 x 16-byte loads, boxes
// plane-major so one load brings the same plane of all four children.  child: >= 0 wide node index, < 0 leaf
// code as in FastNode, kWideEmpty: no child in this slot.
constexpr int32_t kWideEmpty = INT32_MIN;
constexpr int kWideStackLevels = 64;   // 4-wide walk: up to three deferred children per level (depth <= 20)
struct __align__(16) WideNode {
    float lo_x[4], lo_y[4], lo_z[4], hi_x[4], hi_y[4], hi_z[4];
    int32_t child[4];
    uint32_t pad[4];
};
static_assert(sizeof(WideNode) == 128, "WideNode must be 128 bytes");

// texture table entry (texture.h:14-87), 80 B
struct __align__(16) Tex {
    double ar, ag, ab, br, bg, bb;   // solid / checker even; checker odd
    double scale;
    uint64_t data;                   // byte offset into tex_data
    uint32_t type;
    int32_t width, height;
    uint32_t pad;
};
static_assert(sizeof(Tex) == 80, "Tex must be 80 bytes");

// constant_medium (constant_medium.h:8-61) + isotropic (material.h:149-171): bdpt_scene_set_media
// node of the small tree over a LARGE boundary mesh (more than kMedTreeMin triangles; a box's 12 are a plain loop): boxes in fp64, padded;
// b >= 0: children a, b; b < 0: leaf of -b triangles from a (relative to the medium's first triangle)
struct __align__(16) MedNode { double lo[3], hi[3]; int32_t a, b; int32_t pad[2]; };
static_assert(sizeof(MedNode) == 64, "MedNode must be 64 bytes");
constexpr uint32_t kMedTreeMin = 16;
struct __align__(16) MediumDev {
    double neg_inv_density;          // constant_medium.h:12
    double r, g, b;                  // isotropic's colour
    double lo[3], hi[3];             // bounds of the boundary mesh: a ray that misses them has no boundary hit
    uint32_t first, count;           // boundary triangles in med_tris
    const MedNode* nodes;            // tree over them (device pointer), null for small boundaries
};
static_assert(sizeof(MediumDev) == 96, "MediumDev must be 96 bytes");
// analytic sphere of bdpt_scene_set_spheres: a world prim of a side list (the reference only tessellates)
struct __align__(16) SphereDev { double cx, cy, cz, r; uint32_t mat; uint32_t pad[3]; };
static_assert(sizeof(SphereDev) == 48, "SphereDev must be 48 bytes");
constexpr uint32_t kSpherePrim = 0x10000000u;   // hit / vertex prim of a hit on sphere k: kSpherePrim | k
constexpr uint32_t kMediumPrim = 0x20000000u;   // hit / vertex prim of a scatter point inside medium k: kMediumPrim | k (prim IDs stay below 2^28)
#ifdef __CUDACC__
// sphere prims (oracle/port.cpp: sphere_hit), shared by the render kernels and the batch queries.
// The quadratic in the reference's own style: inclusive interval like triangle.h:68, nearer root first, then the farther one.
__device__ __forceinline__ bool sphere_hit_t(const SphereDev& sp, const D3& o, const D3& d, double tmin, double tmax, double& t_out) {
    const D3 oc = mk(sp.cx, sp.cy, sp.cz) - o;
    const double a = len2(d);
    const double h = dot(d, oc);
    const double c = len2(oc) - sp.r * sp.r;
    const double discriminant = h * h - a * c;
    if (discriminant < 0) return false;
    const double sqrtd = sqrt(discriminant);
    double root = (h - sqrtd) / a;
    if (!(tmin <= root && root <= tmax)) {
        root = (h + sqrtd) / a;
        if (!(tmin <= root && root <= tmax)) return false;
    }
    t_out = root;
    return true;
}
// outward unit normal of sphere k at p: (p - c) / r
__device__ __forceinline__ D3 sphere_normal(const SphereDev& sp, const D3& p) { return (p - mk(sp.cx, sp.cy, sp.cz)) / sp.r; }
#endif

struct SceneView {
    const WideNode* wide_nodes;   // may be null (device-built tree): the binary walk is used
    uint32_t n_wide_nodes;
    uint32_t use_wide;
    const RefNode* ref_nodes;
    const FastNode* fast_nodes;
    const uint32_t* fast_prims;   // leaf-ordered prim IDs for the fast BVH
    const TriGeom* fast_tris;     // the triangles in that order (copy): leaf k's triangles are fast_tris[first .. first+cnt)
    const struct Tri32* tris32;   // the same triangles in fp32 for the walk's pre-filter (traverse_w32.cuh), built at upload
    float vmax32;                 // >= the largest |coordinate| of any vertex (pre-filter error scale)
    const uint32_t* dfs_rank;     // prim ID -> last position in the reference DFS leaf order (tie rule)
    const int32_t* ref_parent;    // reference node -> parent node (-1 for the root)
    const int32_t* prim_leaf;     // prim ID -> reference leaf node holding it
    const double* chain_box;      // prim ID -> 6 doubles, intersection of its reference chain's boxes
    const TriGeom* tris;
    const TriShade* shade;
    const Mat* mats;
    const Tex* tex;               // texture table (bdpt_scene_set_textures), null when the scene has none
    const unsigned char* tex_data;
    uint32_t n_tex;
    const MediumDev* media;       // participating media (bdpt_scene_set_media), null when the scene has none
    const TriGeom* med_tris;      // their boundary meshes
    uint32_t n_media;
    const SphereDev* spheres;     // analytic spheres (bdpt_scene_set_spheres), null when the scene has none
    uint32_t n_spheres;
    const uint32_t* lights;       // prim IDs, reference light-list order
    const double* light_area;     // running sums of light triangle areas in list order (triangle.h:206-212)
    uint32_t n_lights;
    uint32_t n_mats;
    uint32_t n_tris;
    uint32_t n_ref_nodes;
    uint32_t n_fast_nodes;
    double light_total_area;      // summed in list order like triangle.h:200-203
    D3 center;                    // centre of the scene bounds
    double extent;                // largest side of the scene bounds (tie-band scale, traverse_fast.cuh)
};

}  // namespace bdpt
