// Reference-exact BVH traversal (replaces bvh_node::hit, src/acceleration/bvh.h:50-59, aabb::hit,
// aabb.h:41-66, and triangle::hit, src/objects/primatives/triangle.h:41-74).
//
// The reference recursion visits left then right whenever the node's box passes, and hands the right
// child the interval [tmin, hit_left ? rec.t : tmax].  By induction that upper bound is always the
// best t found so far, so the recursion equals this iterative form: pop a node, test its box against
// [tmin, best]; leaf -> test its two triangles in order; internal -> push right, then left.  The box is
// tested at POP time (not push time) so the right subtree sees the bound left behind by the left one.
// A triangle is accepted when tmin <= t <= best (inclusive, interval.h:22-24), so among equal-t hits
// the LAST one in DFS order wins, exactly as in the reference.  Compiled with -fmad=false.
#pragma once
#include "scene.cuh"

namespace bdpt {

constexpr int kRefStackLevels = 48;  // deferred children only: one per tree level

struct RayD {
    D3 o, d;
    double tmin, tmax;
};

struct HitD {
    double t, u, v;
    uint32_t prim;
};

// nodes = boxes fetched (the reference tree: one aabb test each, bvh.h:51 + aabb.h:42; the fast tree: two child boxes
// each), box_hits = boxes passed (aabb.h:64), tris = triangle tests (triangle.h:42), tri_hits = accepted (triangle.h:72)
struct TravStats { uint32_t nodes, tris, box_hits, tri_hits; };

__device__ __forceinline__ RefNode load_ref_node(const RefNode* p) {
    // 4 x LDG.128 through the read-only path
    const double2* q = reinterpret_cast<const double2*>(p);
    double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    int4 w = __ldg(reinterpret_cast<const int4*>(q + 3));
    RefNode n;
    n.lo_x = a.x; n.hi_x = a.y; n.lo_y = b.x; n.hi_y = b.y; n.lo_z = c.x; n.hi_z = c.y;
    n.left = w.x; n.right = w.y; n.leaf = uint32_t(w.z); n.pad = 0;
    return n;
}

// aabb.h:41-66 with the per-axis 1.0/dir hoisted (same IEEE quotient every time)
__device__ __forceinline__ bool slab(double lo, double hi, double o, double adinv, double& tmin, double& tmax) {
    double t0 = (lo - o) * adinv;
    double t1 = (hi - o) * adinv;
    if (t0 < t1) {
        if (t0 > tmin) tmin = t0;
        if (t1 < tmax) tmax = t1;
    } else {
        if (t1 > tmin) tmin = t1;
        if (t0 < tmax) tmax = t0;
    }
    return !(tmax <= tmin);
}

__device__ __forceinline__ bool box_hit(const RefNode& n, const D3& o, const D3& adinv, double tmin, double tmax) {
    if (!slab(n.lo_x, n.hi_x, o.x, adinv.x, tmin, tmax)) return false;
    if (!slab(n.lo_y, n.hi_y, o.y, adinv.y, tmin, tmax)) return false;
    return slab(n.lo_z, n.hi_z, o.z, adinv.z, tmin, tmax);
}

// triangle.h:41-67: everything up to the interval check. True when the ray's line meets the triangle
// (|det| >= 1e-8, u in [0,1], v >= 0, u+v <= 1); t is then the reference's t, bit for bit.
__device__ __forceinline__ bool tri_candidate(const TriGeom* tp, const D3& o, const D3& d, double& t_out,
                                              double& u_out, double& v_out) {
    const double2* q = reinterpret_cast<const double2*>(tp);
    double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), e = __ldg(q + 3);
    double e2z = __ldg(reinterpret_cast<const double*>(q + 4));
    D3 v0 = mk(a.x, a.y, b.x), e1 = mk(b.y, c.x, c.y), e2 = mk(e.x, e.y, e2z);
    D3 pvec = cross(d, e2);
    double det = dot(e1, pvec);
    if (fabs(det) < 1e-8) return false;
    double inv_det = 1.0 / det;
    D3 tvec = o - v0;
    double u = dot(tvec, pvec) * inv_det;
    if (u < 0.0 || u > 1.0) return false;
    D3 qvec = cross(tvec, e1);
    double v = dot(d, qvec) * inv_det;
    if (v < 0.0 || u + v > 1.0) return false;
    t_out = dot(e2, qvec) * inv_det;
    u_out = u; v_out = v;
    return true;
}

// triangle.h:41-74. Returns true and updates (t,u,v) when tmin <= t <= tmax (interval.h:22-24).
__device__ __forceinline__ bool tri_hit(const TriGeom* tp, const D3& o, const D3& d, double tmin, double tmax,
                                        double& t_out, double& u_out, double& v_out) {
    double t, u, v;
    if (!tri_candidate(tp, o, d, t, u, v)) return false;
    if (!(tmin <= t && t <= tmax)) return false;
    t_out = t; u_out = u; v_out = v;
    return true;
}

// stack: this thread's column in a [levels][blockDim] shared array; `stride` = blockDim.
template <bool ANY, bool STATS>
__device__ __forceinline__ bool traverse_ref(const SceneView& sc, const RayD& r, HitD& hit, int32_t* stack,
                                             int stride, TravStats* st) {
    const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
    double best = r.tmax;
    bool found = false;
    int sp = 0;
    int32_t cur = 0;  // root
    uint32_t nn = 0, nt = 0, nb = 0, nh = 0;
    for (;;) {
        RefNode n = load_ref_node(sc.ref_nodes + cur);
        if (STATS) ++nn;
        bool descend = false;
        if (box_hit(n, r.o, adinv, r.tmin, best)) {
            if (STATS) ++nb;
            if (n.leaf) {
                if (STATS) nt += 2;
                if (tri_hit(sc.tris + n.left, r.o, r.d, r.tmin, best, hit.t, hit.u, hit.v)) {
                    best = hit.t; hit.prim = uint32_t(n.left); found = true;
                    if (STATS) ++nh;
                    if (ANY) break;
                }
                if (tri_hit(sc.tris + n.right, r.o, r.d, r.tmin, best, hit.t, hit.u, hit.v)) {
                    best = hit.t; hit.prim = uint32_t(n.right); found = true;
                    if (STATS) ++nh;
                    if (ANY) break;
                }
            } else {
                stack[sp * stride] = n.right;
                ++sp;
                cur = n.left;
                descend = true;
            }
        }
        if (!descend) {
            if (sp == 0) break;
            --sp;
            cur = stack[sp * stride];
        }
    }
    if (STATS && st) { st->nodes += nn; st->tris += nt; st->box_hits += nb; st->tri_hits += nh; }
    return found;
}

}  // namespace bdpt
