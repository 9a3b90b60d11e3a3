// Fast traversal with reference-identical answers.
//
// Topology: binned-SAH binary tree (fast_bvh.cu), both child boxes stored in the parent as fp32 values
// rounded OUTWARD from the fp64 triangle bounds (one 64-byte node fetch tests two children).  The slab
// arithmetic itself runs in fp64 on the widened box values: B200 issues fp64 at half the fp32 rate, the
// loop is latency/bandwidth bound, and it keeps culling strictly conservative (the >= 1 fp32 ulp of
// outward rounding dwarfs the fp64 rounding error of the slab test).
// Triangles are decided by the same fp64 Moller-Trumbore as the reference (triangle.h:41-74).  The
// reference's answer is "the LAST triangle in its DFS leaf order among those with the minimum t"
// (bvh.h:55-56 + inclusive interval.h:22-24); here nodes are visited near-to-far, so that rule is
// applied explicitly: accept t < best, or t == best with a higher reference DFS rank.
#pragma once
#include "traverse_ref.cuh"

namespace bdpt {

// The inline fallback to the reference traversal reuses the caller's StackEntry array as int32 slots.  A thread-private
// stack (stride 1: every caller today) is contiguous, so all 2 x levels int32 slots are usable -- the reference tree may
// be up to kRefStackLevels - 1 deep, more than the StackEntry count; a column of a [levels][blockDim] array can only
// use the first word of every entry.
__device__ __forceinline__ int ref_stride(int stride) { return stride == 1 ? 1 : 2 * stride; }


// One slab of a child box in scalar type T (float on the normal path, double for rays whose origin is so
// far outside the scene that fp32 would eat the outward padding).  NaN (0 * inf, origin exactly on a slab
// plane of a zero-direction axis) is dropped by fmin/fmax, which keeps the slab: conservative.
template <class T>
__device__ __forceinline__ void slab_fast(T lo, T hi, T o, T inv, T& tn, T& tf) {
    const T t0 = (lo - o) * inv, t1 = (hi - o) * inv;
    tn = fmax(tn, fmin(t0, t1));
    tf = fmin(tf, fmax(t0, t1));
}

enum { TR_MISS = 0, TR_HIT = 1, TR_DEFER = 2, TR_SLOW = 3 };

// Does the reference's traversal REACH triangle `prim` with the untouched bound [tmin, tmax], i.e. does
// every box on its root-to-leaf chain pass aabb::hit?  One test of the chain's intersection box accepts;
// only when that fails is the chain walked (host_scene.cu: build_ref_bvh).
__device__ __forceinline__ bool reference_reaches(const SceneView& sc, uint32_t prim, const RayD& r, const D3& adinv,
                                                  uint32_t& walk) {
    const double2* q = reinterpret_cast<const double2*>(sc.chain_box + 6 * size_t(prim));
    const double2 bx = __ldg(q), by = __ldg(q + 1), bz = __ldg(q + 2);
    RefNode cb;
    cb.lo_x = bx.x; cb.hi_x = bx.y; cb.lo_y = by.x; cb.hi_y = by.y; cb.lo_z = bz.x; cb.hi_z = bz.y;
    ++walk;
    if (box_hit(cb, r.o, adinv, r.tmin, r.tmax)) return true;
    int32_t node = __ldg(sc.prim_leaf + prim);
    while (node >= 0) {
        const RefNode nd = load_ref_node(sc.ref_nodes + node);
        ++walk;
        if (!box_hit(nd, r.o, adinv, r.tmin, r.tmax)) return false;
        node = __ldg(sc.ref_parent + node);
    }
    return true;
}

// Shadow rays of connect_vertices end ON a surface: max_t = dist - 0.001 lands on the far vertex's own
// triangle, which therefore occludes most segments.  Test that triangle before walking
// anything: if its t is inside [tmin, tmax] and the reference reaches it, the answer is "occluded".
// Anything else (outside the interval, not reached, zero direction component) decides nothing.
__device__ __forceinline__ bool far_endpoint_occludes(const SceneView& sc, const RayD& r, uint32_t prim, TravStats* st) {
    if (r.d.x == 0.0 || r.d.y == 0.0 || r.d.z == 0.0) return false;
    double t, u, v;
    if (st) ++st->tris;
    if (!tri_candidate(sc.tris + prim, r.o, r.d, t, u, v)) return false;
    if (!(r.tmin <= t && t <= r.tmax)) return false;
    const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
    uint32_t walk = 0;
    const bool reached = reference_reaches(sc, prim, r, adinv, walk);
    if (st) st->nodes += walk;
    return reached;
}

// traversal stack entry: deferred child + its box-entry distance (fp32, a lower bound), so that a
// child made unreachable by a closer hit found in the meantime is dropped at pop time without a fetch
typedef unsigned long long StackEntry;
__device__ __forceinline__ StackEntry pack_entry(int32_t node, float t) {
    return (StackEntry(uint32_t(node)) << 32) | StackEntry(__float_as_uint(t));
}

// Where the reference's answer is NOT simply "closest triangle" — and how this kernel reproduces it:
//  (1) A zero direction component: the reference's slab test yields 0*inf = NaN when the origin lies on
//      that slab's plane and its compare chain (aabb.h:49-61) then rejects the node (a ray inside the
//      floor plane misses the whole scene at the root).
//  (2) Knife edges: bvh_node::hit hands a subtree the interval [tmin, best] and aabb::hit rejects it when
//      tmax <= tmin (aabb.h:61).  A triangle whose t equals the current bound to within rounding (a
//      coplanar duplicate surface, the far endpoint of a visible() segment at dist-0.001, a hit at tmin)
//      can sit in a node whose slab-computed entry is >= that bound, and is then never tested.  Which of
//      two tied triangles wins therefore depends on the reference topology and on fp64 rounding.
//  Both depend on the reference tree, so such rays are re-run through the reference-exact traversal.
//  Detection: every triangle the ray's line meets reports its exact t; if it lies within a band of
//  tmin, of tmax, or of another candidate's t, the ray is flagged.  The band (1e-10 of the scene extent
//  in t units) is far above fp64 rounding (~1e-13) and far below the outward box padding (2^-16 of the
//  extent), so a flagged-band triangle can never have been culled by THIS tree's boxes.
// Returns TR_HIT / TR_MISS, or TR_DEFER when DEFER is set and the ray needs the reference-exact traversal
// (the caller queues it for the exact kernel); with DEFER unset that traversal runs inline.
// `stack` is this thread's column of a [levels][stride] shared array of StackEntry.
// `limit` > 0 bounds the walk: a ray that has touched more than `limit` nodes + triangles gives up with
// TR_SLOW (nothing decided) so that the caller can re-run it, unbounded, among its own kind — a launch's
// duration is then no longer the latency chain of its single worst ray (a segment running inside a
// tessellated plane touches ~1000 nodes against a mean of ~25).
template <bool ANY, bool STATS, bool DEFER, class T>
__device__ __forceinline__ int traverse_fast_t(const SceneView& sc, const RayD& r, HitD& hit, StackEntry* stack,
                                               int stride, TravStats* st, uint32_t limit) {
    const T ox = T(r.o.x), oy = T(r.o.y), oz = T(r.o.z);
    const T ix = T(1.0 / r.d.x), iy = T(1.0 / r.d.y), iz = T(1.0 / r.d.z);
    const double inv_max = fmin(fabs(1.0 / r.d.x), fmin(fabs(1.0 / r.d.y), fabs(1.0 / r.d.z)));
    const double band_abs = 1e-10 * sc.extent * inv_max;
    // bounds in T, rounded outward
    const T tmin_t = sizeof(T) == 4 ? T(__double2float_rd(r.tmin)) : T(r.tmin);
    double best = r.tmax;
    T best_t = sizeof(T) == 4 ? T(__double2float_ru(best)) : T(best);
    uint32_t best_rank = 0;
    bool found = false, flagged = false;
    int sp = 0;
    int32_t cur = 0;
    uint32_t nn = 0, nt = 0;
    uint32_t work = 0;
    constexpr int kMaxEdge = 4;
    uint32_t edge_prim[kMaxEdge];
    int n_edge = 0;
    bool done = false;
    while (!done) {
        // ---- inner nodes: descend until a leaf (or the stack runs dry)
        while (cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
            const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            const int4 w = __ldg(reinterpret_cast<const int4*>(q + 3));
            if (STATS) ++nn;
            ++work;
            // a = l_lo.xyz l_hi.x | b = l_hi.yz r_lo.xy | c = r_lo.z r_hi.xyz
            T ln = tmin_t, lf = best_t, rn = tmin_t, rf = best_t;
            slab_fast<T>(T(a.x), T(a.w), ox, ix, ln, lf);
            slab_fast<T>(T(a.y), T(b.x), oy, iy, ln, lf);
            slab_fast<T>(T(a.z), T(b.y), oz, iz, ln, lf);
            slab_fast<T>(T(b.z), T(c.y), ox, ix, rn, rf);
            slab_fast<T>(T(b.w), T(c.z), oy, iy, rn, rf);
            slab_fast<T>(T(c.x), T(c.w), oz, iz, rn, rf);
            const bool hl = ln <= lf, hr = rn <= rf;
            int32_t lc = w.x, rc = w.y;
            if (hl && hr) {
                float far_t = float(rn);
                if (rn < ln) { const int32_t t = lc; lc = rc; rc = t; far_t = float(ln); }
                stack[sp * stride] = pack_entry(rc, far_t);
                ++sp;
                cur = lc;
            } else if (hl) cur = lc;
            else if (hr) cur = rc;
            else {
                // pop, dropping children that a closer hit has put out of reach
                for (;;) {
                    if (sp == 0) { done = true; break; }
                    --sp;
                    const StackEntry e = stack[sp * stride];
                    if (T(__uint_as_float(uint32_t(e))) <= best_t) { cur = int32_t(uint32_t(e >> 32)); break; }
                }
                if (done) break;
            }
        }
        if (done) break;
        // ---- leaf: cur = ~(first << 3 | count-1)
        {
            const uint32_t code = uint32_t(~cur);
            const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
            work += cnt;
            if (limit && work > limit) {
                if (STATS && st) { st->nodes += nn; st->tris += nt; }
                return TR_SLOW;
            }
            for (uint32_t k = 0; k < cnt; ++k) {
                const uint32_t prim = __ldg(sc.fast_prims + first + k);
                if (STATS) ++nt;
                double t, u, v;
                if (!tri_candidate(sc.tris + prim, r.o, r.d, t, u, v)) continue;
                const double band = band_abs + 1e-10 * fabs(t);
                const bool edge = fabs(t - r.tmin) <= band || fabs(t - r.tmax) <= band;
                if (ANY) {
                    if (edge) {   // knife edge: decided after the walk, by the reference's own box chain
                        if (r.tmin <= t && t <= r.tmax) {
                            if (n_edge < kMaxEdge) edge_prim[n_edge++] = prim; else flagged = true;
                        }
                        continue;
                    }
                    if (r.tmin <= t && t <= r.tmax) {                  // certain occluder
                        hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; found = true; done = true; break;
                    }
                } else {
                    if (found && fabs(t - best) <= band) flagged = true;   // tie with the current best
                    if (edge) flagged = true;
                    if (r.tmin <= t && t <= best) {
                        const uint32_t rank = __ldg(sc.dfs_rank + prim);
                        if (!found || t < best || rank > best_rank) {
                            best = t; best_rank = rank; found = true;
                            best_t = sizeof(T) == 4 ? T(__double2float_ru(best)) : T(best);
                            hit.t = t; hit.u = u; hit.v = v; hit.prim = prim;
                        }
                    }
                }
            }
            if (done) break;
            for (;;) {
                if (sp == 0) { done = true; break; }
                --sp;
                const StackEntry e = stack[sp * stride];
                if (T(__uint_as_float(uint32_t(e))) <= best_t) { cur = int32_t(uint32_t(e >> 32)); break; }
            }
        }
    }
    if (STATS && st) { st->nodes += nn; st->tris += nt; }
    if (ANY) {
        if (found) return TR_HIT;
        if (!flagged) {
            // No certain occluder.  The reference reports a hit iff some knife-edge triangle that passes its
            // (bit-identical) interval test is actually REACHED, i.e. every box on its root-to-leaf chain in
            // the reference tree passes aabb::hit with the untouched bound [tmin, tmax] (the bound only
            // shrinks after a hit, and then the answer is already "occluded").
            if (n_edge == 0) return TR_MISS;
            const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
            uint32_t walk = 0;
            for (int k = 0; k < n_edge && !found; ++k) {
                if (reference_reaches(sc, edge_prim[k], r, adinv, walk)) { hit.prim = edge_prim[k]; hit.t = r.tmax; hit.u = 0.0; hit.v = 0.0; found = true; }
            }
            if (STATS && st) st->nodes += walk;
            return found ? TR_HIT : TR_MISS;
        }
    } else if (!flagged) {
        return found ? TR_HIT : TR_MISS;
    }
    if (DEFER) return TR_DEFER;
    return traverse_ref<ANY, STATS>(sc, r, hit, reinterpret_cast<int32_t*>(stack), ref_stride(stride), st) ? TR_HIT : TR_MISS;
}

// Entry point: picks the arithmetic of the box tests.
//  - a zero direction component -> reference-exact traversal (case 1 above)
//  - origin (or the scene itself) farther than 8 scene extents from the scene centre / the world origin -> fp64 slabs (fp32 would eat the padding)
//  - otherwise fp32 slabs: outward padding of 2^-16 extent covers the <= 5e-6 extent fp32 error
template <bool ANY, bool STATS, bool DEFER>
__device__ __forceinline__ int traverse_fast(const SceneView& sc, const RayD& r, HitD& hit, StackEntry* stack,
                                             int stride, TravStats* st, uint32_t limit = 0) {
    if (r.d.x == 0.0 || r.d.y == 0.0 || r.d.z == 0.0) {
        if (DEFER) return TR_DEFER;
        return traverse_ref<ANY, STATS>(sc, r, hit, reinterpret_cast<int32_t*>(stack), ref_stride(stride), st) ? TR_HIT : TR_MISS;
    }
    const double far = 8.0 * sc.extent;
    const bool near_scene = fabs(r.o.x - sc.center.x) <= far && fabs(r.o.y - sc.center.y) <= far && fabs(r.o.z - sc.center.z) <= far &&
                            fabs(sc.center.x) <= far && fabs(sc.center.y) <= far && fabs(sc.center.z) <= far;
    if (near_scene) return traverse_fast_t<ANY, STATS, DEFER, float>(sc, r, hit, stack, stride, st, limit);
    return traverse_fast_t<ANY, STATS, DEFER, double>(sc, r, hit, stack, stride, st, limit);
}

}  // namespace bdpt
