// Warp-synchronous fp32 walk with an exact fp64 finish ("walk32") — the bulk traversal of round 2.
//
// Why: the per-lane while-while walk of traverse_walk.cuh carries the ray in fp64, tests every triangle of every
// visited leaf with the reference's fp64 Moller-Trumbore and lets the 32 lanes of a warp drift apart; ncu put it at
// 8.4 (closest) / 4.7 (shadow) active lanes per instruction and 43 % issue-active at a 64-register cap
// (profiles/r1_ncu_full_k_trace_fast_final.csv).  Here
//   * the walk itself is fp32 only: ray, slabs AND a conservative fp32 Moller-Trumbore PRE-FILTER with rigorous error
//     bounds.  A triangle the filter rejects is one the reference's fp64 test (triangle.h:41-74) rejects too, with a
//     margin of ~1e-5 of the ray length against an fp64 noise of ~1e-13; what it cannot reject becomes a CANDIDATE
//     (a leaf-order index in a 4-entry list);
//   * all 32 lanes of a warp run the same loop trips: a trip is a few "pop if needed, then open ONE node" steps; a
//     lane that reaches a leaf waits there, and the triangle pre-filter runs for all waiting lanes at once as soon as
//     `t_leaf` lanes wait (or few lanes still walk) — so both phases execute with many lanes instead of 8 and 5;
//   * the reference's fp64 test, the tie / knife-edge flagging and the DFS-order rule of traverse_fast.cuh run ONCE per
//     ray over its candidates (typically one), at the warp's meeting point, for all finished lanes together.
// The answers are those of traverse_fast_t / Walk: the candidate set always contains every triangle the exact test
// could accept within [tmin, best] plus everything within the flagging bands, and the decision logic over that set
// (decide() below) restates Walk::leaf_tests + Walk::finish over a list.
//
// Error bounds of the pre-filter (delta = 2^-24, round to nearest; inputs are the fp64 values rounded to fp32, so
// each carries a relative error <= delta; rho >= |o|_inf + |v0|_inf, |T| <= sqrt(3) rho):
//   pvec = d x e2      |dP|_inf <= 4 delta |d||e2|
//   det  = e1 . pvec   |dDet|   <= 11 delta |d||e1||e2|
//   U    = T . pvec    |dU|     <= 21 delta rho |d||e2|        (T = o - v0, |dT_i| <= 2.01 delta rho)
//   V    = d . (T x e1)|dV|     <= 24 delta rho |d||e1|
//   TN   = e2 . (T x e1) |dTN|  <= 24 delta rho |e1||e2|
// The filter uses K = 2^-17 = 128 delta in place of 11 / 21 / 24 delta: >= 5x slack, which also covers the rounding of
// the comparisons themselves and the reference's own fp64 rounding by ten orders of magnitude.
#pragma once
#include <cmath>
#include <cstring>

#include "prefilter32.cuh"
#include "traverse_walk.cuh"

namespace bdpt {

enum { WALK32_FLUSH = -2 };                  // run(): the candidate list is full; call decide(last = false) and continue
constexpr int32_t kWalkPop = INT32_MIN + 1;  // cur: take the next stack entry (never a leaf code: first < 2^28 - 1)
constexpr int kMaxCand = 4;

#ifdef __CUDACC__
template <bool ANY, bool STATS>
struct Walk32 {
    Ray32 R;                     // origin, direction, interval images and error scales (pre-filter)
    float ix, iy, iz, nox, noy, noz;   // slabs: 1/d (the fp64 quotient, rounded) and -o/d for one FMA per plane
    float best_t;                // the walk's current upper bound (R.best_t is the ray's own tmax image)
    int32_t cur;
    int sp;
    uint32_t c0, c1, c2, c3;     // candidates: leaf-order indices for the exact test (scalars: an indexed array would
    int ncand;                   //   put the whole walk state in local memory)
    uint32_t skip;               // ANY: prim that is known not to occlude (the far endpoint's triangle), else BDPT_MISS
    bool flagged;                // an exact test saw a tie / knife edge: the reference tree decides
    bool occluded;               // ANY: a certain occluder was found (c0 = its leaf-order index)
    TravStats ts;

    __device__ __forceinline__ void push(uint32_t k) {
        if (ncand == 0) c0 = k; else if (ncand == 1) c1 = k; else if (ncand == 2) c2 = k; else c3 = k;
        ++ncand;
    }
    __device__ __forceinline__ uint32_t get(int i) const { return i == 0 ? c0 : (i == 1 ? c1 : (i == 2 ? c2 : c3)); }

    __device__ __forceinline__ void init(const SceneView& sc, const RayD& r, uint32_t skip_prim = 0xFFFFFFFFu) {
        const double o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
        R = make_ray32(o, d, r.tmin, r.tmax, sc.vmax32);
        ix = float(1.0 / r.d.x); iy = float(1.0 / r.d.y); iz = float(1.0 / r.d.z);
        nox = -R.ox * ix; noy = -R.oy * iy; noz = -R.oz * iz;
        best_t = R.best_t;
        cur = 0; sp = 0; ncand = 0; c0 = c1 = c2 = c3 = 0; skip = skip_prim;
        flagged = false; occluded = false;
        ts.nodes = 0; ts.tris = 0; ts.box_hits = 0; ts.tri_hits = 0;
    }

    __device__ __forceinline__ int filter(const Tri32* __restrict__ T, float& t_hi, uint32_t& prim) const {
        const float4 q0 = __ldg(&T->q0), q1 = __ldg(&T->q1), q2 = __ldg(&T->q2);
        prim = __float_as_uint(q2.w);
        return filter32<ANY>(R, best_t, q0, q1, q2, t_hi);
    }

    // One time slice for the whole warp (all 32 lanes must call; busy = this lane holds a ray that is not finished).
    // A trip = [the pre-filter for the lanes waiting at a leaf, when enough of them wait] + up to `nstep` node steps
    // per lane with no vote in between (a lane that reaches a leaf sits out the rest of the trip) + one vote.
    // Returns per lane: WALK_YIELD (call again), WALK32_FLUSH (call decide(last = false), then again) or TR_HIT (the
    // walk is over: call decide(last = true)).  TR_HIT is only a "done" marker here.
    __device__ __forceinline__ int run(const SceneView& sc, StackEntry* stack, bool busy, int trips, int nstep, int t_leaf, int t_node,
                                       int min_idle) {
        bool stalled = false;   // candidate list full: wait for the caller's flush
        const uint32_t m_busy = __ballot_sync(0xFFFFFFFFu, busy);
        for (int trip = 0; trip < trips; ++trip) {
            const bool live = busy && !stalled && cur != kWalkDry;
            const bool at_leaf = live && cur < 0 && cur != kWalkPop;
            const uint32_t m_leaf = __ballot_sync(0xFFFFFFFFu, at_leaf);
            // ---- leaf phase: the pre-filter for every waiting lane at once
            if (m_leaf) {
                const uint32_t m_walk = __ballot_sync(0xFFFFFFFFu, live && !at_leaf);
                if (__popc(m_leaf) >= t_leaf || __popc(m_walk) <= t_node) {
                    if (at_leaf) {
                        const uint32_t code = uint32_t(~cur);
                        const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;   // builders make leaves of 1 or 2 triangles
                        const bool room = ncand + int(cnt) <= kMaxCand;
                        if (!room) stalled = true;
                        const Tri32* T = sc.tris32 + first;
                        bool done = false;
#pragma unroll
                        for (uint32_t k = 0; k < 2; ++k) {
                            __syncwarp(m_leaf);   // the waiting lanes test their k-th triangles together
                            if (room && k < cnt && !done) {
                                if (STATS) ++ts.tris;
                                float t_hi = 0.f;
                                uint32_t prim;
                                const int f = filter(T + k, t_hi, prim);
                                if (f != 0 && !(ANY && prim == skip)) {
                                    if (ANY && f == 2) { occluded = true; c0 = first + k; done = true; }
                                    else {
                                        push(first + k);
                                        if (!ANY && f == 2) best_t = fminf(best_t, t_hi);
                                    }
                                }
                            }
                        }
                        if (room) cur = done ? kWalkDry : kWalkPop;
                    }
                }
            }
            __syncwarp();
            // ---- node steps: every lane takes every step (predicated), so the warp is whole at each node fetch
#pragma unroll 1
            for (int s = 0; s < nstep; ++s) {
                const bool act = busy && !stalled && (cur >= 0 || cur == kWalkPop);
                if (!__any_sync(0xFFFFFFFFu, act)) break;
                if (act && cur == kWalkPop) {   // one stack entry per step; an entry a closer hit has put out of reach costs a step
                    if (sp == 0) cur = kWalkDry;
                    else {
                        --sp;
                        const StackEntry e = stack[sp];
                        if (__uint_as_float(uint32_t(e)) <= best_t) cur = int32_t(uint32_t(e >> 32));
                    }
                }
                __syncwarp();
                if (act && cur >= 0) {
                    const float4* q = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
                    const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
                    const int2 w = __ldg(reinterpret_cast<const int2*>(q + 3));
                    if (STATS) ++ts.nodes;
                    float ln = R.tmin_t, lf = best_t, rn = R.tmin_t, rf = best_t;
                    slab_fma(a.x, a.w, nox, ix, ln, lf);
                    slab_fma(a.y, b.x, noy, iy, ln, lf);
                    slab_fma(a.z, b.y, noz, iz, ln, lf);
                    slab_fma(b.z, c.y, nox, ix, rn, rf);
                    slab_fma(b.w, c.z, noy, iy, rn, rf);
                    slab_fma(c.x, c.w, noz, iz, rn, rf);
                    const bool hl = ln <= lf, hr = rn <= rf;
                    if (STATS) ts.box_hits += uint32_t(hl) + uint32_t(hr);
                    const bool swap = rn < ln;
                    if (hl && hr) { stack[sp] = pack_entry(swap ? w.x : w.y, swap ? ln : rn); ++sp; }
                    cur = (hl && hr) ? (swap ? w.y : w.x) : (hl ? w.x : (hr ? w.y : kWalkPop));
                }
                __syncwarp();
            }
            const uint32_t m_done = __ballot_sync(0xFFFFFFFFu, busy && (stalled || cur == kWalkDry));
            if (m_done == m_busy || (min_idle > 0 && __popc(m_done) + (32 - __popc(m_busy)) >= min_idle)) break;
        }
        if (!busy) return WALK_YIELD;
        if (cur == kWalkDry) return TR_HIT;
        return stalled ? WALK32_FLUSH : WALK_YIELD;
    }

    // The exact decision over the candidate list: Walk::leaf_tests + Walk::finish of traverse_walk.cuh, restated for
    // a list.  `last` = the walk is over (finish) / the list is full (flush: keep the winner, go on walking).
    // Returns TR_HIT / TR_MISS / TR_DEFER when last, else WALK_YIELD.
    __device__ __forceinline__ int decide(const SceneView& sc, const RayD& r, HitD& hit, bool last) {
        const double inv_max = fmin(fabs(1.0 / r.d.x), fmin(fabs(1.0 / r.d.y), fabs(1.0 / r.d.z)));
        const double band_abs = 1e-10 * sc.extent * inv_max;
        if (ANY) {
            if (occluded) { hit.prim = __ldg(sc.fast_prims + c0); hit.t = r.tmax; hit.u = 0.0; hit.v = 0.0; return TR_HIT; }
#pragma unroll
            for (int k = 0; k < kMaxCand; ++k) {
                if (k >= ncand) break;
                const uint32_t ck = get(k);
                if (STATS) ++ts.tris;
                double t, u, v;
                if (!tri_candidate(sc.fast_tris + ck, r.o, r.d, t, u, v)) continue;
                if (!(r.tmin <= t && t <= r.tmax)) continue;
                const uint32_t prim = __ldg(sc.fast_prims + ck);
                if (prim == skip) continue;
                const double band = band_abs + 1e-10 * fabs(t);
                const bool edge = fabs(t - r.tmin) <= band || fabs(t - r.tmax) <= band;
                bool occ = true;
                if (edge) {   // knife edge: occludes iff the reference's own box chain reaches it (traverse_fast.cuh)
                    const D3 adinv = mk(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
                    uint32_t walk = 0;
                    occ = reference_reaches(sc, prim, r, adinv, walk);
                    if (STATS) ts.nodes += walk;
                }
                if (occ) { occluded = true; c0 = ck; hit.prim = prim; hit.t = t; hit.u = u; hit.v = v; cur = kWalkDry; return TR_HIT; }
            }
            ncand = 0;
            return last ? TR_MISS : WALK_YIELD;
        } else {
            double best = r.tmax;
            bool found = false;
            uint32_t best_k = 0;
#pragma unroll
            for (int k = 0; k < kMaxCand; ++k) {
                if (k >= ncand) break;
                const uint32_t ck = get(k);
                if (STATS) ++ts.tris;
                double t, u, v;
                if (!tri_candidate(sc.fast_tris + ck, r.o, r.d, t, u, v)) continue;
                const double band = band_abs + 1e-10 * fabs(t);
                if (found && fabs(t - best) <= band) flagged = true;                          // tie with the current best
                if (fabs(t - r.tmin) <= band || fabs(t - r.tmax) <= band) flagged = true;     // knife edge
                if (r.tmin <= t && t <= best && (!found || t < best)) {
                    if (STATS && last) ++ts.tri_hits;
                    best = t; found = true; best_k = ck;
                    hit.t = t; hit.u = u; hit.v = v;
                }
            }
            if (!last) {
                ncand = 0;
                if (found) { c0 = best_k; ncand = 1; best_t = fminf(best_t, __double2float_ru(best)); }
                return WALK_YIELD;
            }
            if (flagged) return TR_DEFER;
            if (!found) return TR_MISS;
            hit.prim = __ldg(sc.fast_prims + best_k);
            return TR_HIT;
        }
    }
};
#endif  // __CUDACC__

}  // namespace bdpt
