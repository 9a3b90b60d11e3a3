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
#include "prefilter32.cuh"
#include "traverse_fast.cuh"

namespace bdpt {

enum { WALK_YIELD = -1 };

// -DBDPT_PREFILTER=1: leaf_tests rejects triangles with the conservative fp32 pre-filter (prefilter32.cuh) before the
// reference's fp64 test (most triangles of a visited leaf are misses; ~35 full-rate fp32 instructions against ~80 half-rate
// fp64 ones).  Parity-green (the filter only rejects what the exact test rejects, or what lies beyond the bound / before
// tmin by far more than the flagging bands), but MEASURED SLOWER and therefore off: bench step 21.85 -> 23.3 ms, shadow
// stage +12 %, config-2-style batch 2,604 -> 2,486 Mrays/s (profiles/r2_rejected_experiments.txt).  The walk waits on
// dependent loads, not on issue slots: the filter's 48-byte record is one more load in front of the 80-byte one.
#ifndef BDPT_PREFILTER
#define BDPT_PREFILTER 0
#endif

// fp32 slabs are valid for this ray (see traverse_fast): no zero direction component, origin and scene
// within 8 extents of each other and of the world origin
__device__ __forceinline__ bool walk_eligible(const SceneView& sc, const RayD& r) {
    if (r.d.x == 0.0 || r.d.y == 0.0 || r.d.z == 0.0) return false;
    const double far = 8.0 * sc.extent;
    return fabs(r.o.x - sc.center.x) <= far && fabs(r.o.y - sc.center.y) <= far && fabs(r.o.z - sc.center.z) <= far &&
           fabs(sc.center.x) <= far && fabs(sc.center.y) <= far && fabs(sc.center.z) <= far;
}

// One slab in fp32 with explicit FMAs (this file is compiled with -fmad=false for the fp64 parity code):
// t = plane * inv - o * inv.  Against (plane - o) * inv this moves the result by a few ulps of |o * inv|,
// which stays far inside the outward padding of the boxes (2^-16 of the scene extent, fast_bvh.cu) for the
// rays that are walk_eligible (origin within 8 extents).
__device__ __forceinline__ void slab_fma(float lo, float hi, float noi, float inv, float& tn, float& tf) {
    const float t0 = __fmaf_rn(lo, inv, noi), t1 = __fmaf_rn(hi, inv, noi);
    tn = fmaxf(tn, fminf(t0, t1));
    tf = fminf(tf, fmaxf(t0, t1));
}

template <bool ANY, bool STATS>
struct Walk {
    // ray (fp64 for the triangle tests, fp32 for the slabs)
    RayD r;
    float ox, oy, oz, ix, iy, iz, tmin_t, best_t;   // ox.. hold -o * inv: one FMA per slab plane
    float cull_hi;     // upper end of the part of the ray this walk is responsible for (split walks), else FLT_MAX
    uint32_t work;     // nodes + triangles touched so far
#if BDPT_PREFILTER
    float kD, kU, kT;  // error scales of the fp32 pre-filter for this ray (make_ray32)
#endif
    double best, band_abs;
    // answer so far
    HitD hit;
    bool found, flagged;
    // walk position
    int sp;
    int32_t cur;
    int32_t pend;      // run_wide: a leaf whose triangle tests are postponed (0: none; leaf codes are negative)
    // ANY: knife-edge candidates decided after the walk
    uint32_t edge_prim[4];
    int n_edge;
    TravStats ts;

    // [cull_lo, cull_hi] (fp32, conservative): only boxes overlapping this part of the ray are opened.  Hits are
    // still accepted over the ray's whole interval, so a split walk returns a valid hit of the full ray and the
    // closest one whenever it lies in its part.
    __device__ __forceinline__ void init(const SceneView& sc, const RayD& ray, float cull_lo = 0.f,
                                         float cull_hi_ = 3.402823466e+38f) {
        r = ray;
        ix = float(1.0 / r.d.x); iy = float(1.0 / r.d.y); iz = float(1.0 / r.d.z);
        ox = -float(r.o.x) * ix; oy = -float(r.o.y) * iy; oz = -float(r.o.z) * iz;
        const double inv_max = fmin(fabs(1.0 / r.d.x), fmin(fabs(1.0 / r.d.y), fabs(1.0 / r.d.z)));
        band_abs = 1e-10 * sc.extent * inv_max;
        tmin_t = fmaxf(__double2float_rd(r.tmin), cull_lo);
        cull_hi = cull_hi_;
        best = r.tmax;
        best_t = fminf(__double2float_ru(best), cull_hi);
#if BDPT_PREFILTER
        {
            const float fx = float(r.d.x), fy = float(r.d.y), fz = float(r.d.z);
            const float dn = __double2float_ru(sqrt(double(fx) * fx + double(fy) * fy + double(fz) * fz) * 1.000001);
            const float rho = fmaxf(fabsf(float(r.o.x)), fmaxf(fabsf(float(r.o.y)), fabsf(float(r.o.z)))) * 1.000001f + sc.vmax32;
            kD = kFilterK * dn; kU = kFilterK * rho * dn; kT = kFilterK * rho;
        }
#endif
        work = 0;
        found = false; flagged = false;
        sp = 0; cur = 0; pend = 0;
        n_edge = 0;
        ts.nodes = 0; ts.tris = 0; ts.box_hits = 0; ts.tri_hits = 0;
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
    // leaves (returns WALK_YIELD; call again to continue; rounds <= 0: no limit), or TR_SLOW once it has touched
    // more than `limit` nodes + triangles (limit 0: never).  The time slice is what
    // brings the warp back together: __activemask() inside the walk only sees the lanes that happen to be
    // scheduled together, not the lanes that still have work.
    __device__ __forceinline__ int run(const SceneView& sc, StackEntry* stack, int rounds, uint32_t limit = 0);
    // The same over the 4-wide nodes (sc.wide_nodes): same leaves, same candidates' tests, up to three pushes per node.
    __device__ __forceinline__ int run4(const SceneView& sc, StackEntry* stack, int rounds, uint32_t limit = 0);
    // The same walk for a whole warp in lock step (all 32 lanes must call it; `busy` = this lane has a ray):
    // every trip each walking lane opens ONE node, a lane that reaches a leaf postpones its triangle tests and
    // keeps walking, and the tests run for all postponed leaves at once when at least `t_leaf` lanes have one
    // or at most `t_walk` lanes can still walk.  Returns after `trips` trips, or earlier when no lane is busy
    // or (min_idle > 0) at least `min_idle` lanes are idle.  Per lane: WALK_YIELD (call again) or TR_*.
    __device__ __forceinline__ int run_wide(const SceneView& sc, StackEntry* stack, bool busy, int trips, int t_leaf,
                                            int t_walk, int min_idle, uint32_t limit = 0);
    // triangle tests of one leaf; true: ANY walk found a certain occluder (hit is set)
    __device__ __forceinline__ bool leaf_tests(const SceneView& sc, uint32_t first, uint32_t cnt);
    // the decision once the walk has ended (the tail of traverse_fast_t)
    __device__ __forceinline__ int finish(const SceneView& sc);
};

template <bool ANY, bool STATS>
__device__ __forceinline__ int Walk<ANY, STATS>::run(const SceneView& sc, StackEntry* stack, int rounds, uint32_t limit) {
    for (;;) {
        // ---- inner nodes: descend until a leaf (or the stack runs dry)
        while (cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
            const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            const int4 w = __ldg(reinterpret_cast<const int4*>(q + 3));
            if (STATS) ++ts.nodes;
            ++work;
            // a = l_lo.xyz l_hi.x | b = l_hi.yz r_lo.xy | c = r_lo.z r_hi.xyz
            float ln = tmin_t, lf = best_t, rn = tmin_t, rf = best_t;
            slab_fma(a.x, a.w, ox, ix, ln, lf);
            slab_fma(a.y, b.x, oy, iy, ln, lf);
            slab_fma(a.z, b.y, oz, iz, ln, lf);
            slab_fma(b.z, c.y, ox, ix, rn, rf);
            slab_fma(b.w, c.z, oy, iy, rn, rf);
            slab_fma(c.x, c.w, oz, iz, rn, rf);
            const bool hl = ln <= lf, hr = rn <= rf;
            if (STATS) ts.box_hits += uint32_t(hl) + uint32_t(hr);
            int32_t lc = w.x, rc = w.y;
            if (hl && hr) {
                float far_t = rn;
                if (rn < ln) { const int32_t t = lc; lc = rc; rc = t; far_t = ln; }
                stack[sp] = pack_entry(rc, far_t);
                ++sp;
                cur = lc;
#ifdef BDPT_PREFETCH_FAR
                if (rc >= 0) {   // the deferred child is popped later: start its 64 bytes on their way to L1
                    const char* pf = reinterpret_cast<const char*>(sc.fast_nodes + rc);
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + 32));
                }
#endif
            } else if (hl) cur = lc;
            else if (hr) cur = rc;
            else if (!pop(stack)) return finish(sc);
        }
        // ---- leaf: cur = ~(first << 3 | count-1)
        const uint32_t code = uint32_t(~cur);
        const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
        work += cnt;
        if (limit && work > limit) return TR_SLOW;   // nothing decided: the caller parks the ray for a split walk
        if (leaf_tests(sc, first, cnt)) return TR_HIT;
        if (!pop(stack)) return finish(sc);
        if (--rounds == 0) return WALK_YIELD;
    }
}

// compare-exchange of (entry distance, child) pairs: the smaller distance ends in a
__device__ __forceinline__ void cswap(float& da, int32_t& ca, float& db, int32_t& cb) {
    const bool sw = db < da;
    const float d0 = sw ? db : da, d1 = sw ? da : db;
    const int32_t c0 = sw ? cb : ca, c1 = sw ? ca : cb;
    da = d0; db = d1; ca = c0; cb = c1;
}

template <bool ANY, bool STATS>
__device__ __forceinline__ int Walk<ANY, STATS>::run4(const SceneView& sc, StackEntry* stack, int rounds, uint32_t limit) {
    const float kMissT = 3.402823466e+38f;
    for (;;) {
        while (cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(sc.wide_nodes + cur);
            const float4 lx = __ldg(q), ly = __ldg(q + 1), lz = __ldg(q + 2), hx = __ldg(q + 3), hy = __ldg(q + 4), hz = __ldg(q + 5);
            const int4 ch = __ldg(reinterpret_cast<const int4*>(q + 6));
            if (STATS) ++ts.nodes;
            ++work;
            float d0, d1, d2, d3;
            int32_t c0 = ch.x, c1 = ch.y, c2 = ch.z, c3 = ch.w;
#define BDPT_WIDE_CHILD(D, C, LX, HX, LY, HY, LZ, HZ)                              \
            {                                                                      \
                float tn = tmin_t, tf = best_t;                                    \
                slab_fma(LX, HX, ox, ix, tn, tf);                                  \
                slab_fma(LY, HY, oy, iy, tn, tf);                                  \
                slab_fma(LZ, HZ, oz, iz, tn, tf);                                  \
                D = (tn <= tf && C != kWideEmpty) ? tn : kMissT;                   \
            }
            BDPT_WIDE_CHILD(d0, c0, lx.x, hx.x, ly.x, hy.x, lz.x, hz.x)
            BDPT_WIDE_CHILD(d1, c1, lx.y, hx.y, ly.y, hy.y, lz.y, hz.y)
            BDPT_WIDE_CHILD(d2, c2, lx.z, hx.z, ly.z, hy.z, lz.z, hz.z)
            BDPT_WIDE_CHILD(d3, c3, lx.w, hx.w, ly.w, hy.w, lz.w, hz.w)
#undef BDPT_WIDE_CHILD
            // sort by entry distance (misses carry FLT_MAX and end up last)
            cswap(d0, c0, d1, c1); cswap(d2, c2, d3, c3); cswap(d0, c0, d2, c2); cswap(d1, c1, d3, c3); cswap(d1, c1, d2, c2);
            if (d0 == kMissT) { if (!pop(stack)) return finish(sc); continue; }
            // nearest child next; the others wait on the stack, farthest pushed first
            if (d3 != kMissT) { stack[sp] = pack_entry(c3, d3); ++sp; }
            if (d2 != kMissT) { stack[sp] = pack_entry(c2, d2); ++sp; }
            if (d1 != kMissT) { stack[sp] = pack_entry(c1, d1); ++sp; }
            cur = c0;
        }
        const uint32_t code = uint32_t(~cur);
        const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
        work += cnt;
        if (limit && work > limit) return TR_SLOW;
        if (leaf_tests(sc, first, cnt)) return TR_HIT;
        if (!pop(stack)) return finish(sc);
        if (--rounds == 0) return WALK_YIELD;
    }
}

template <bool ANY, bool STATS>
__device__ __forceinline__ bool Walk<ANY, STATS>::leaf_tests(const SceneView& sc, uint32_t first, uint32_t cnt) {
#if BDPT_PREFILTER
    const float fox = float(r.o.x), foy = float(r.o.y), foz = float(r.o.z), fdx = float(r.d.x), fdy = float(r.d.y), fdz = float(r.d.z);
    const float f_tmin = __double2float_rd(r.tmin), f_best = __double2float_ru(ANY ? r.tmax : best);   // the RAY's interval, not a split walk's part
#endif
    for (uint32_t k = 0; k < cnt; ++k) {
        if (STATS) ++ts.tris;
#if BDPT_PREFILTER
        if (sc.tris32) {
            const Tri32* T = sc.tris32 + first + k;
            const float4 q0 = __ldg(&T->q0), q1 = __ldg(&T->q1), q2 = __ldg(&T->q2);
            if (prefilter_rejects(fox, foy, foz, fdx, fdy, fdz, f_tmin, f_best, kD, kU, kT, q0, q1, q2)) continue;
        }
#endif
        double t, u, v;
        if (!tri_candidate(sc.fast_tris + first + k, r.o, r.d, t, u, v)) continue;
        const uint32_t prim = __ldg(sc.fast_prims + first + k);   // only candidates need their id
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
                if (STATS) ++ts.tri_hits;
                return true;
            }
        } else {
            if (found && fabs(t - best) <= band) flagged = true;   // tie with the current best
            if (edge) flagged = true;
            // t == best (the reference's DFS order decides) cannot be settled here and need not be: an exact
            // tie is inside the band above, so the ray is flagged and re-traced through the reference tree
            if (r.tmin <= t && t <= best) {
                if (!found || t < best) {
                    if (STATS) ++ts.tri_hits;
                    best = t; found = true;
                    best_t = fminf(__double2float_ru(best), cull_hi);
                    hit.u = u; hit.v = v; hit.prim = prim;   // hit.t: `best`, copied by finish()
                }
            }
        }
    }
    return false;
}

constexpr int32_t kWalkDry = INT32_MIN;   // cur: the stack has run dry (never a leaf code: first < 2^28)

template <bool ANY, bool STATS>
__device__ __forceinline__ int Walk<ANY, STATS>::run_wide(const SceneView& sc, StackEntry* stack, bool busy, int trips,
                                                          int t_leaf, int t_walk, int min_idle, uint32_t limit) {
    int result = WALK_YIELD;
    for (int trip = 0; trip < trips; ++trip) {
        // ---- one node per walking lane
        if (busy && cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
            const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            const int4 w = __ldg(reinterpret_cast<const int4*>(q + 3));
            if (STATS) ++ts.nodes;
            ++work;
            float ln = tmin_t, lf = best_t, rn = tmin_t, rf = best_t;
            slab_fma(a.x, a.w, ox, ix, ln, lf);
            slab_fma(a.y, b.x, oy, iy, ln, lf);
            slab_fma(a.z, b.y, oz, iz, ln, lf);
            slab_fma(b.z, c.y, ox, ix, rn, rf);
            slab_fma(b.w, c.z, oy, iy, rn, rf);
            slab_fma(c.x, c.w, oz, iz, rn, rf);
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
            else if (!pop(stack)) cur = kWalkDry;
        }
        // ---- a leaf moves into the postponed slot when that is free, and the lane walks on
        if (busy && cur < 0 && cur != kWalkDry && pend == 0) {
            pend = cur;
            if (!pop(stack)) cur = kWalkDry;
        }
        // ---- triangle tests for every postponed leaf at once
        const uint32_t m_pend = __ballot_sync(0xFFFFFFFFu, busy && pend != 0);
        const uint32_t m_walk = __ballot_sync(0xFFFFFFFFu, busy && cur >= 0);
        if (m_pend && (__popc(m_pend) >= t_leaf || __popc(m_walk) <= t_walk)) {
            if (busy && pend != 0) {
                const uint32_t code = uint32_t(~pend);
                const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
                pend = 0;
                work += cnt;
                if (limit && work > limit) { result = TR_SLOW; busy = false; }
                else if (leaf_tests(sc, first, cnt)) { result = TR_HIT; busy = false; }
                // a closer hit lowers best_t: pop() and the slab test drop what no longer matters as it comes up
            }
        }
        if (busy && cur == kWalkDry && pend == 0) { result = finish(sc); busy = false; }
        const uint32_t m_busy = __ballot_sync(0xFFFFFFFFu, busy);
        if (!m_busy || (min_idle > 0 && 32 - __popc(m_busy) >= min_idle)) break;
    }
    return result;
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
        hit.t = best;
        return found ? TR_HIT : TR_MISS;
    }
    return TR_DEFER;
}

// One ray walked by a whole warp: lane j opens only the boxes that overlap the j-th of 32 equal parts of the
// ray's passage through the scene, so a ray that would touch ~1000 nodes on its own (a segment running
// inside a tessellated plane) costs each lane ~1/32 of that plus the top of the tree.  Every lane still
// accepts hits over the whole interval, so the answer is: ANY — occluded if any lane found an occluder;
// closest — the smallest t over the lanes, deferred to the reference tree when two different triangles tie
// within the band, exactly as within one walk.  Must be called by all 32 lanes with the same ray (which must
// be walk_eligible).  The result is valid in every lane.
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
    return v;
}

template <bool ANY, bool STATS>
__device__ __forceinline__ int split_walk(const SceneView& sc, const RayD& r, HitD& hit, StackEntry* stack, TravStats& ts) {
    const int lane = threadIdx.x & 31;
    // the ray's passage through a cube around the scene (wider than every padded box of the fast tree)
    const double h = 0.5005 * sc.extent + 1e-300;
    double t0 = r.tmin, t1 = r.tmax;
    {
        const double cx[3] = {sc.center.x, sc.center.y, sc.center.z}, o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double inv = 1.0 / d[k];
            const double a = (cx[k] - h - o[k]) * inv, b = (cx[k] + h - o[k]) * inv;
            t0 = fmax(t0, fmin(a, b));
            t1 = fmin(t1, fmax(a, b));
        }
    }
    if (!(t0 <= t1)) return TR_MISS;   // never enters the scene: no box can be opened, no candidate exists
    const double step = (t1 - t0) * (1.0 / 32.0);
    const float lo = __double2float_rd(t0 + step * lane), hi = lane == 31 ? __double2float_ru(t1) : __double2float_ru(t0 + step * (lane + 1));
    Walk<ANY, STATS> W;
    W.init(sc, r, lo, hi);
    const int rc = W.run(sc, stack, 0, 0);
    if (STATS) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { W.ts.nodes += __shfl_xor_sync(0xFFFFFFFFu, W.ts.nodes, o); W.ts.tris += __shfl_xor_sync(0xFFFFFFFFu, W.ts.tris, o); }
        ts.nodes += W.ts.nodes; ts.tris += W.ts.tris;
    }
    if (ANY) {
        const uint32_t hits = __ballot_sync(0xFFFFFFFFu, rc == TR_HIT);
        if (hits) {
            const int src = __ffs(hits) - 1;
            hit.prim = __shfl_sync(0xFFFFFFFFu, W.hit.prim, src);
            hit.t = __shfl_sync(0xFFFFFFFFu, W.hit.t, src); hit.u = 0.0; hit.v = 0.0;
            return TR_HIT;
        }
        return __any_sync(0xFFFFFFFFu, rc == TR_DEFER) ? TR_DEFER : TR_MISS;
    }
    if (__any_sync(0xFFFFFFFFu, rc == TR_DEFER)) return TR_DEFER;
    const double inf = __longlong_as_double(0x7FF0000000000000LL);
    const double mine = W.found ? W.best : inf;
    const double tmin_all = warp_min(mine);
    if (!(tmin_all < inf)) return TR_MISS;
    const uint32_t winners = __ballot_sync(0xFFFFFFFFu, W.found && mine == tmin_all);
    const int src = __ffs(winners) - 1;
    const uint32_t prim = __shfl_sync(0xFFFFFFFFu, W.hit.prim, src);
    const double band = W.band_abs + 1e-10 * fabs(tmin_all);
    if (__any_sync(0xFFFFFFFFu, W.found && W.hit.prim != prim && fabs(mine - tmin_all) <= band)) return TR_DEFER;
    hit.prim = prim;
    hit.t = tmin_all;
    hit.u = __shfl_sync(0xFFFFFFFFu, W.hit.u, src);
    hit.v = __shfl_sync(0xFFFFFFFFu, W.hit.v, src);
    return TR_HIT;
}

}  // namespace bdpt
