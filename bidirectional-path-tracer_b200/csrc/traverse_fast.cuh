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

__device__ __forceinline__ bool slab_fast(double lo, double hi, double o, double adinv, double& tmin, double& tmax) {
    double t0 = (lo - o) * adinv;
    double t1 = (hi - o) * adinv;
    // NaN (0 * inf) only when the origin lies exactly on a slab plane of a zero-direction axis:
    // fmin/fmax drop the NaN, which keeps the slab (conservative).
    double tn = fmin(t0, t1), tf = fmax(t0, t1);
    tmin = fmax(tmin, tn);
    tmax = fmin(tmax, tf);
    return tmin <= tmax;
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
//  in t units) is far above fp64 rounding (~1e-13) and far below the outward box padding (2^-20 of the
//  extent), so a flagged-band triangle can never have been culled by THIS tree's boxes.
template <bool ANY, bool STATS>
__device__ __forceinline__ bool traverse_fast(const SceneView& sc, const RayD& r, HitD& hit, int32_t* stack,
                                              int stride, TravStats* st) {
    if (r.d.x == 0.0 || r.d.y == 0.0 || r.d.z == 0.0) return traverse_ref<ANY, STATS>(sc, r, hit, stack, stride, st);
    const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
    const double band_abs = 1e-10 * sc.extent * fmin(fabs(adinv.x), fmin(fabs(adinv.y), fabs(adinv.z)));
    double best = r.tmax;
    uint32_t best_rank = 0;
    bool found = false, flagged = false;
    int sp = 0;
    int32_t cur = 0;
    uint32_t nn = 0, nt = 0;
    constexpr int kMaxEdge = 4;
    uint32_t edge_prim[kMaxEdge];
    int n_edge = 0;
    for (;;) {
        if (cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
            float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            int4 w = __ldg(reinterpret_cast<const int4*>(q + 3));
            if (STATS) ++nn;
            // a = l_lo.xyz l_hi.x | b = l_hi.yz r_lo.xy | c = r_lo.z r_hi.xyz
            double ln = r.tmin, lf = best, rn = r.tmin, rf = best;
            bool hl = slab_fast(a.x, a.w, r.o.x, adinv.x, ln, lf) && slab_fast(a.y, b.x, r.o.y, adinv.y, ln, lf) &&
                      slab_fast(a.z, b.y, r.o.z, adinv.z, ln, lf);
            bool hr = slab_fast(b.z, c.y, r.o.x, adinv.x, rn, rf) && slab_fast(b.w, c.z, r.o.y, adinv.y, rn, rf) &&
                      slab_fast(c.x, c.w, r.o.z, adinv.z, rn, rf);
            int32_t lc = w.x, rc = w.y;
            if (hl && hr) {
                if (rn < ln) { int32_t t = lc; lc = rc; rc = t; }
                stack[sp * stride] = rc;
                ++sp;
                cur = lc;
                continue;
            } else if (hl) { cur = lc; continue; }
            else if (hr) { cur = rc; continue; }
        } else {
            // leaf: cur = ~(first << 3 | count-1)
            const uint32_t code = uint32_t(~cur);
            const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
            bool done = false;
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
                            hit.t = t; hit.u = u; hit.v = v; hit.prim = prim;
                        }
                    }
                }
            }
            if (done) break;
        }
        if (sp == 0) break;
        --sp;
        cur = stack[sp * stride];
    }
    if (STATS && st) { st->nodes += nn; st->tris += nt; }
    const uint32_t nn_walk = nn;
    if (ANY) {
        if (found) return true;
        if (flagged) return traverse_ref<ANY, STATS>(sc, r, hit, stack, stride, st);
        // No certain occluder.  The reference reports a hit iff some knife-edge triangle that passes its
        // (bit-identical) interval test is actually REACHED, i.e. every box on its root-to-leaf chain in
        // the reference tree passes aabb::hit with the untouched bound [tmin, tmax] (the bound only
        // shrinks after a hit, and then the answer is already "occluded").
        for (int k = 0; k < n_edge; ++k) {
            int32_t node = __ldg(sc.prim_leaf + edge_prim[k]);
            bool reached = true;
            while (node >= 0) {
                RefNode nd = load_ref_node(sc.ref_nodes + node);
                if (STATS) ++nn;
                if (!box_hit(nd, r.o, adinv, r.tmin, r.tmax)) { reached = false; break; }
                node = __ldg(sc.ref_parent + node);
            }
            if (reached) { hit.prim = edge_prim[k]; hit.t = r.tmax; hit.u = 0.0; hit.v = 0.0; found = true; break; }
        }
        if (STATS && st) st->nodes += nn - nn_walk;
        return found;
    }
    if (flagged) return traverse_ref<ANY, STATS>(sc, r, hit, stack, stride, st);
    return found;
}

}  // namespace bdpt
