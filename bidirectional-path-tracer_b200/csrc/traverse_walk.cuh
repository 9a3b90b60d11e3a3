// Resumable form of the fast traversal (traverse_fast.cuh), for the persistent wavefront kernels.
//
// A warp of a while-while traversal runs as long as its slowest ray; with a mean of ~25 steps and single
// rays of ~1000 (a segment inside a tessellated plane) most lanes idle most of the time (ncu: 12 of 32
// lanes active per instruction in k_trace_paths, 5 of 32 in k_shadow).  Here the walk keeps its whole
// state in a struct, hands control back ("yields") after a time slice of a few leaf visits, and is
// re-entered after the lanes whose ray is decided have fetched new rays — Aila & Laine's persistent
// threads with dynamic fetch.  The arithmetic, the tie/knife-edge flagging and therefore the
// answers are those of traverse_fast_t<.., float>; rays that need fp64 slabs or the reference tree are
// not started here at all (walk_eligible) and go to the deferred list.
#pragma once
#include "traverse_fast.cuh"

namespace bdpt {

enum { WALK_YIELD = -1 };

// fp32 slabs are valid for this ray (see traverse_fast): no zero direction component, origin and scene
// within 8 extents of each other and of the world origin
__device__ __forceinline__ bool walk_eligible(const SceneView& sc, const RayD& r) {
    if (r.d.x == 0.0 || r.d.y == 0.0 || r.d.z == 0.0) return false;
    const double far = 8.0 * sc.extent;
    return fabs(r.o.x - sc.center.x) <= far && fabs(r.o.y - sc.center.y) <= far && fabs(r.o.z - sc.center.z) <= far &&
           fabs(sc.center.x) <= far && fabs(sc.center.y) <= far && fabs(sc.center.z) <= far;
}

template <bool ANY, bool STATS>
struct Walk {
    // ray (fp64 for the triangle tests, fp32 for the slabs)
    RayD r;
    float ox, oy, oz, ix, iy, iz, tmin_t, best_t;
    double best, band_abs;
    // answer so far
    HitD hit;
    uint32_t best_rank;
    bool found, flagged;
    // walk position
    int sp;
    int32_t cur;
    // ANY: knife-edge candidates decided after the walk
    uint32_t edge_prim[4];
    int n_edge;
    TravStats ts;

    __device__ __forceinline__ void init(const SceneView& sc, const RayD& ray) {
        r = ray;
        ox = float(r.o.x); oy = float(r.o.y); oz = float(r.o.z);
        ix = float(1.0 / r.d.x); iy = float(1.0 / r.d.y); iz = float(1.0 / r.d.z);
        const double inv_max = fmin(fabs(1.0 / r.d.x), fmin(fabs(1.0 / r.d.y), fabs(1.0 / r.d.z)));
        band_abs = 1e-10 * sc.extent * inv_max;
        tmin_t = __double2float_rd(r.tmin);
        best = r.tmax;
        best_t = __double2float_ru(best);
        best_rank = 0;
        found = false; flagged = false;
        sp = 0; cur = 0;
        n_edge = 0;
        ts.nodes = 0; ts.tris = 0;
    }

    // pops until a child still within reach is found; false when the stack is empty
    __device__ __forceinline__ bool pop(const StackEntry* stack) {
        while (sp > 0) {
            --sp;
            const StackEntry e = stack[sp];
            if (__uint_as_float(uint32_t(e)) <= best_t) { cur = int32_t(uint32_t(e >> 32)); return true; }
        }
        return false;
    }

    // Walks until the ray is decided (returns TR_HIT / TR_MISS / TR_DEFER) or until it has visited `rounds`
    // leaves (returns WALK_YIELD; call again to continue; rounds <= 0: no limit).  The time slice is what
    // brings the warp back together: __activemask() inside the walk only sees the lanes that happen to be
    // scheduled together, not the lanes that still have work.
    __device__ __forceinline__ int run(const SceneView& sc, StackEntry* stack, int rounds);
    // the decision once the walk has ended (the tail of traverse_fast_t)
    __device__ __forceinline__ int finish(const SceneView& sc);
};

template <bool ANY, bool STATS>
__device__ __forceinline__ int Walk<ANY, STATS>::run(const SceneView& sc, StackEntry* stack, int rounds) {
    for (;;) {
        // ---- inner nodes: descend until a leaf (or the stack runs dry)
        while (cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
            const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            const int4 w = __ldg(reinterpret_cast<const int4*>(q + 3));
            if (STATS) ++ts.nodes;
            // a = l_lo.xyz l_hi.x | b = l_hi.yz r_lo.xy | c = r_lo.z r_hi.xyz
            float ln = tmin_t, lf = best_t, rn = tmin_t, rf = best_t;
            slab_fast<float>(a.x, a.w, ox, ix, ln, lf);
            slab_fast<float>(a.y, b.x, oy, iy, ln, lf);
            slab_fast<float>(a.z, b.y, oz, iz, ln, lf);
            slab_fast<float>(b.z, c.y, ox, ix, rn, rf);
            slab_fast<float>(b.w, c.z, oy, iy, rn, rf);
            slab_fast<float>(c.x, c.w, oz, iz, rn, rf);
            const bool hl = ln <= lf, hr = rn <= rf;
            int32_t lc = w.x, rc = w.y;
            if (hl && hr) {
                float far_t = rn;
                if (rn < ln) { const int32_t t = lc; lc = rc; rc = t; far_t = ln; }
                stack[sp] = pack_entry(rc, far_t);
                ++sp;
                cur = lc;
            } else if (hl) cur = lc;
            else if (hr) cur = rc;
            else if (!pop(stack)) return finish(sc);
        }
        // ---- leaf: cur = ~(first << 3 | count-1)
        const uint32_t code = uint32_t(~cur);
        const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
        for (uint32_t k = 0; k < cnt; ++k) {
            const uint32_t prim = __ldg(sc.fast_prims + first + k);
            if (STATS) ++ts.tris;
            double t, u, v;
            if (!tri_candidate(sc.tris + prim, r.o, r.d, t, u, v)) continue;
            const double band = band_abs + 1e-10 * fabs(t);
            const bool edge = fabs(t - r.tmin) <= band || fabs(t - r.tmax) <= band;
            if (ANY) {
                if (edge) {   // knife edge: decided after the walk, by the reference's own box chain
                    if (r.tmin <= t && t <= r.tmax) {
                        if (n_edge < 4) edge_prim[n_edge++] = prim; else flagged = true;
                    }
                    continue;
                }
                if (r.tmin <= t && t <= r.tmax) {                  // certain occluder
                    hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; found = true;
                    return TR_HIT;
                }
            } else {
                if (found && fabs(t - best) <= band) flagged = true;   // tie with the current best
                if (edge) flagged = true;
                if (r.tmin <= t && t <= best) {
                    const uint32_t rank = __ldg(sc.dfs_rank + prim);
                    if (!found || t < best || rank > best_rank) {
                        best = t; best_rank = rank; found = true;
                        best_t = __double2float_ru(best);
                        hit.t = t; hit.u = u; hit.v = v; hit.prim = prim;
                    }
                }
            }
        }
        if (!pop(stack)) return finish(sc);
        if (--rounds == 0) return WALK_YIELD;
    }
}

template <bool ANY, bool STATS>
__device__ __forceinline__ int Walk<ANY, STATS>::finish(const SceneView& sc) {
    if (ANY) {
        if (!flagged) {
            // No certain occluder: a knife-edge triangle occludes iff the reference's traversal reaches it
            // (traverse_fast.cuh).
            if (n_edge == 0) return TR_MISS;
            const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
            uint32_t walk = 0;
            for (int k = 0; k < n_edge; ++k) {
                if (reference_reaches(sc, edge_prim[k], r, adinv, walk)) {
                    hit.prim = edge_prim[k]; hit.t = r.tmax; hit.u = 0.0; hit.v = 0.0; found = true;
                    break;
                }
            }
            if (STATS) ts.nodes += walk;
            return found ? TR_HIT : TR_MISS;
        }
    } else if (!flagged) {
        return found ? TR_HIT : TR_MISS;
    }
    return TR_DEFER;
}

}  // namespace bdpt
