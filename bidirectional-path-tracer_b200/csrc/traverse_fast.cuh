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

template <bool ANY, bool STATS>
__device__ __forceinline__ bool traverse_fast(const SceneView& sc, const RayD& r, HitD& hit, int32_t* stack,
                                              int stride, TravStats* st) {
    const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
    double best = r.tmax;
    uint32_t best_rank = 0;
    bool found = false;
    int sp = 0;
    int32_t cur = 0;
    uint32_t nn = 0, nt = 0;
    if (sc.n_fast_nodes == 0) return false;
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
            // leaf children carry their triangle count in the low bits of w.z: lcnt | rcnt << 16
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
            for (uint32_t k = 0; k < cnt; ++k) {
                const uint32_t prim = __ldg(sc.fast_prims + first + k);
                if (STATS) ++nt;
                double t, u, v;
                if (tri_hit(sc.tris + prim, r.o, r.d, r.tmin, best, t, u, v)) {
                    if (ANY) { hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; found = true; break; }
                    const uint32_t rank = __ldg(sc.dfs_rank + prim);
                    if (!found || t < best || rank > best_rank) {
                        best = t; best_rank = rank; found = true;
                        hit.t = t; hit.u = u; hit.v = v; hit.prim = prim;
                    }
                }
            }
            if (ANY && found) break;
        }
        if (sp == 0) break;
        --sp;
        cur = stack[sp * stride];
    }
    if (STATS && st) { st->nodes += nn; st->tris += nt; }
    return found;
}

}  // namespace bdpt
