// Wavefront bidirectional path tracer: camera::render's sample loop (src/camera.h:111-129) and
// bidirectional_color (camera.h:294-323) restructured as queue-driven kernels.
//
//   K0 k_generate      get_ray (camera.h:199-223) + build_light_path's emitter (camera.h:372-414)
//   K1 k_trace_paths   world.hit for every live subpath (camera.h:336) -> per-material queues
//   K2 k_shade         one launch over the per-material queues (material.h:42-147): vertex push, emission,
//                      scatter, throughput update (camera.h:342-368); survivors -> next ray queue
//   K3 k_connect       connect_vertices minus visibility (camera.h:440-475) -> shadow-ray queue
//   K4 k_shadow        visible() (camera.h:425-438) + film accumulation (camera.h:121-124)
//
// Path regeneration: a wavefront has B sample SLOTS.  A slot whose two subpaths have both ended is
// connected (K3/K4), freed, and refilled with the next sample by K0 in the following iteration, so every
// iteration traces a full queue instead of the thinning tail of one fixed batch.
//
// All arithmetic is fp64 in the reference's operation order and this file is compiled with -fmad=false,
// so a sample is computed exactly as the CPU oracle's counter-based mode computes it; only the order in
// which contributions are summed into a pixel differs (fp64 atomics).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "ctx.h"
#include "render_math.cuh"
#include "texture.cuh"
#include "traverse_fast.cuh"
#include "traverse_walk.cuh"
#include "traverse_w32.cuh"

namespace bdpt {

namespace {

constexpr int kBlock = 64;
constexpr uint32_t kLightFlag = 0x80000000u;
// Front blocks give every parked ray a warp of its own (lane 0): these rays are latency chains of ~1000
// dependent steps, and 32 of them in one warp would serialise on each other's loop phases (measured: 7x
// slower than alone).  Issue slots are not the constraint — a few hundred such rays per launch.
constexpr int kFrontPerBlock = kBlock / 32;
constexpr uint32_t kSlowFlag = 0x80000000u;   // q_exact entry: re-run the FAST traversal without a work limit (else: reference traversal)

// device counters (uint32 each)
enum { C_ACTIVE0 = 0, C_ACTIVE1 = 1, C_MAT0 = 2 /* ..5 */, C_SHADOW = 6, C_MAXNC = 7, C_MAXNL = 8, C_SUMNC = 9,
       C_SUMNL = 10, C_OVERFLOW = 11, C_PAIRS_LO = 12, C_PAIRS_HI = 13 /* 64-bit, 8-byte aligned */, C_CURSOR = 14 /* dynamic-fetch cursor of the trace launch */, C_EXACT_SHADOW = 15,
       C_DONE0 = 16, C_DONE1 = 17, C_SHADOW_TOTAL = 18 /* ..19: 64-bit, shadow rays of the whole worker run */, C_SHADOW_LIVE = 20, C_EXACT0 = 21, C_EXACT1 = 22,
       C_MAT4 = 23 /* fifth material queue: scatter points inside a medium (isotropic) */,
       C_SLOW_S0 = 24, C_SLOW_S1 = 25 /* parked shadow segments, ping-pong */, C_COUNT = 28 };

struct Frame {           // camera.h:148-197 derived values + this call's sample selection
    D3 center, pixel00, du, dv, background;
    double recip_sqrt_spp;
    int W, H, n;         // n = sqrt_spp
    int depth;
    int row0, rows;      // selected rows
    int s_begin, s_step; // selected strata: s_begin + k*s_step
    uint64_t seed;
    int64_t npix;        // rows * W
    // BDPT_WEIGHT_MIS only: viewing direction, upper-left corner of the viewport, |du|^2, |dv|^2, distance of the image
    // plane in pixel sides, focus distance, light subpaths per pass over the frame (= W * H)
    D3 fwd, ul;
    double du2, dv2, plane_px, focus, n_light;
};

// Subpath state and path vertices are gathered through queues of arbitrary ids, so they are stored as
// records (AoS): one thread touches whole 32-byte sectors of its own record instead of one 8-byte word in
// each of a dozen arrays.
struct __align__(32) PathRec {   // 96 B = 3 sectors
    double ox, oy, oz, dx, dy, dz;   // ray of the next world.hit (sectors 0-1: all the trace kernel reads)
    double tx, ty, tz;               // throughput
    double hit_t;
    uint32_t draws, hit_prim;
    int32_t remaining;
    uint32_t pad;
};
static_assert(sizeof(PathRec) == 96, "PathRec must be 96 bytes");
struct __align__(32) VertRec {   // 64 B = 2 sectors
    double px, py, pz;
    double ax, ay, az;               // camera vertex: throughput * f (camera.h:471); light vertex: throughput
    uint32_t prim;                   // light vertex: bit 31 = is_light (camera.h:348)
    uint32_t pad[3];
};
static_assert(sizeof(VertRec) == 64, "VertRec must be 64 bytes");

// BDPT_WEIGHT_MIS: the two running quantities of the recursive balance-heuristic weights (Georgiev's vertex-connection
// form: every connection then costs O(1)), per subpath and, frozen at the moment of the hit, per stored vertex.
struct __align__(16) PathMis { double dVCM, dVC; };
struct __align__(32) VertMis { double dVCM, dVC, cos_in; uint32_t len, pad; };   // len = surface vertices of the subpath up to this one
static_assert(sizeof(VertMis) == 32, "VertMis must be 32 bytes");
constexpr uint32_t kBackFlag = 0x40000000u;   // VertRec.prim: the path arrived on the back side (the facing normal is -n)

struct __align__(32) SlowShadow {   // a parked shadow segment: ray, finished contribution, pixel; 96 B
    double ox, oy, oz, dx, dy, dz, tmax, cx, cy, cz;
    uint32_t pix, kind;              // kind 1: needs the reference traversal
    uint32_t pad[2];
};
static_assert(sizeof(SlowShadow) == 96, "SlowShadow must be 96 bytes");

// Shadow queue: SoA, written at consecutive slots (warp-aggregated append) and read by index.
struct SQ {
    double *ox, *oy, *oz, *dx, *dy, *dz, *tmax, *cx, *cy, *cz;   // segment and its finished contribution
    uint32_t *pix, *far;             // pixel; prim of the far (light-side) vertex
    uint32_t* live;                  // segments the far-endpoint test did not settle (indices into this queue)
    uint32_t* cnt;                   // SQ_*
    // participating media only (null otherwise): what k_media_segments needs to settle the media's share of a segment
    unsigned long long* mkey;        // REF_SUM: (sample id << 20) | (camera vertex index * 128 + light vertex index): the draws' key
    double* msum;                    // MIS: wLight + wCamera of the segment's weight with the transmittance taken out
};
constexpr uint32_t kDeadSeg = 0xFFFFFFFEu;   // SQ.far of a segment a medium scattered: the shadow stage skips it
enum { SQ_N = 0, SQ_LIVE = 1, SQ_OVERFLOW = 2, SQ_CURSOR = 3 /* dynamic-fetch cursor of stage 2 */, SQ_COUNT = 8 };

struct WF {              // wavefront workspace, all device pointers
    int64_t B;           // batch capacity in samples
    int depth;
    int64_t scap;        // shadow queue capacity
    // subpath state, index = path id (camera subpath of sample b: b; light subpath: B + b)
    PathRec* path;
    // per slot: the sample it holds (index into this call's sample list) and how many of its subpaths run
    int64_t* sample;
    uint32_t* live;
    uint32_t* q_slots[2];            // slots whose subpaths have both ended (ping-pong: done list -> free list)
    uint32_t *q_active[2], *q_mat[5];   // q_mat[4]: isotropic (scatter points inside media)
    uint32_t* q_exact[2];            // path rays for the next launch's front blocks (reference traversal or unbounded re-run), ping-pong
    SlowShadow* ss[2];               // parked shadow segments, ping-pong
    int64_t sscap;
    uint32_t* counters;
    // vertex pools, index = b * depth + slot (a sample's vertices are contiguous)
    VertRec *cv, *lv;
    uint32_t *ncam, *nlight;
    SQ sq[2];                        // shadow queues, ping-pong: the shadow stage of one iteration runs on its own
                                     // stream while the main stream already traces the next
    int textured;                    // the scene has textures: lcol exists
    D3* lcol;                        // beside lv: the light vertex's colour as connect_vertices needs it (emission / albedo at that point)
    int media;                       // the scene has participating media: the shadow queues carry mkey / msum
    int mis;                         // BDPT_WEIGHT_MIS: the three arrays below exist
    PathMis* pmis;                   // per path id
    VertMis *cvm, *lvm;              // beside cv / lv
    void* base;
    size_t bytes;
};

// one atomic per warp: every thread of the warp must call this (pred = whether it appends)
__device__ __forceinline__ uint32_t warp_append(uint32_t* counter, bool pred) {
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (mask) {
        const int leader = __ffs(mask) - 1;
        if (lane == leader) base = atomicAdd(counter, uint32_t(__popc(mask)));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
    }
    return base + uint32_t(__popc(mask & ((1u << lane) - 1u)));
}

// one atomic per block of NT threads: every thread of the block must call this.  The streaming kernels
// (generate, shade) append from every warp to the same counter; ncu showed them waiting on that one address
// (29 % of k_generate's stall samples on the warp leaders' atomicAdd), so their warps pool their counts first.
constexpr int kWide = 256;
constexpr int kGen = 128;    // k_generate: 84 registers, 5 blocks of 128 stay resident where 2 of 256 would
template <int NT>
__device__ __forceinline__ uint32_t block_append(uint32_t* counter, bool pred) {
    __shared__ uint32_t s_off[NT / 32];
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_off[wid] = uint32_t(__popc(mask));
    __syncthreads();
    if (wid == 0) {
        const uint32_t c = lane < NT / 32 ? s_off[lane] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < NT / 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, NT / 32 - 1);
        uint32_t base = 0;
        if (lane == 0 && total) base = atomicAdd(counter, total);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (lane < NT / 32) s_off[lane] = base + incl - c;
    }
    __syncthreads();
    const uint32_t r = s_off[wid] + uint32_t(__popc(mask & ((1u << lane) - 1u)));
    __syncthreads();   // s_off is free again for the next call
    return r;
}

__device__ __forceinline__ void film_add(double* film, uint32_t pix, D3 c) {
    atomicAdd(film + 3 * size_t(pix), c.x);
    atomicAdd(film + 3 * size_t(pix) + 1, c.y);
    atomicAdd(film + 3 * size_t(pix) + 2, c.z);
}

// 16-byte vector accessors of the records
__device__ __forceinline__ void store_path(PathRec* r, D3 o, D3 d, D3 T, uint32_t draws, int32_t remaining) {
    double2* q = reinterpret_cast<double2*>(r);
    q[0] = make_double2(o.x, o.y); q[1] = make_double2(o.z, d.x); q[2] = make_double2(d.y, d.z);
    q[3] = make_double2(T.x, T.y); q[4] = make_double2(T.z, 0.0);
    reinterpret_cast<int4*>(r)[5] = make_int4(int(draws), -1, remaining, 0);
}
// index: the vertex's position on its subpath (camera: bounce; light: 0 = the emitter, then 1 + bounce) -- keys the shadow
// stream's draws when the scene has media
__device__ __forceinline__ void store_vertex(VertRec* r, D3 p, D3 a, uint32_t prim, uint32_t index = 0) {
    double2* q = reinterpret_cast<double2*>(r);
    q[0] = make_double2(p.x, p.y); q[1] = make_double2(p.z, a.x); q[2] = make_double2(a.y, a.z);
    reinterpret_cast<int4*>(r)[3] = make_int4(int(prim), int(index), 0, 0);
}
__device__ __forceinline__ void load_vertex(const VertRec* r, D3& p, D3& a, uint32_t& prim, uint32_t* index = nullptr) {
    const double2* q = reinterpret_cast<const double2*>(r);
    const double2 q0 = q[0], q1 = q[1], q2 = q[2];
    const int4 tail = reinterpret_cast<const int4*>(r)[3];
    prim = uint32_t(tail.x);
    if (index) *index = uint32_t(tail.y);
    p = mk(q0.x, q0.y, q1.x); a = mk(q1.y, q2.x, q2.y);
}

// A subpath of slot b has ended.  The thread that ends the slot's LAST live subpath appends the slot to the
// done list; connection and refill happen in later launches, so the vertices written by either subpath
// are visible by then.  Must be called by every thread of the warp (ended = false for bystanders).
__device__ __forceinline__ void retire(const WF& w, bool ended, uint32_t b, int dn) {
    bool last = false;
    if (ended) last = atomicSub(&w.live[b], 1u) == 1u;
    const uint32_t slot = warp_append(w.counters + C_DONE0 + dn, last);
    if (last) w.q_slots[dn][slot] = b;
}

// the same for kernels that append per block (every thread of the block must call it)
template <int NT>
__device__ __forceinline__ void retire_block(const WF& w, bool ended, uint32_t b, int dn) {
    bool last = false;
    if (ended) last = atomicSub(&w.live[b], 1u) == 1u;
    const uint32_t slot = block_append<NT>(w.counters + C_DONE0 + dn, last);
    if (last) w.q_slots[dn][slot] = b;
}

// linear sample index of this call -> pixel, stratum and the global sample id that keys the RNG
struct SampleId { int i, j, s_i, s_j; uint32_t pix; uint64_t sid; };
__device__ __forceinline__ SampleId decode_sample(const Frame& f, int64_t L) {
    SampleId s;
    const int64_t k = L / f.npix, p = L % f.npix;
    s.j = f.row0 + int(p / f.W);
    s.i = int(p % f.W);
    const int stratum = f.s_begin + int(k) * f.s_step;
    s.s_j = stratum / f.n;
    s.s_i = stratum % f.n;
    s.pix = uint32_t(s.j) * uint32_t(f.W) + uint32_t(s.i);
    s.sid = uint64_t(s.pix) * uint64_t(f.n * f.n) + uint64_t(stratum);
    return s;
}

// ------------------------------------------------------------------ K0
// Fills `ngen` free slots (free_list, or slots 0..ngen-1 on a worker's first iteration) with samples
// gen_base .. gen_base+ngen-1 of this call; consecutive threads take consecutive samples = neighbouring
// pixels, so the primary rays of a refill are coherent.
template <bool MIS, bool TEX>
__global__ void __launch_bounds__(kGen) k_generate(SceneView sc, Frame f, WF w, const uint32_t* __restrict__ free_list,
                                                   int64_t gen_base, int64_t ngen, int cur, int dn) {
    const int64_t gi = int64_t(blockIdx.x) * kGen + threadIdx.x;
    const bool valid = gi < ngen;
    bool cam_live = false, light_live = false;
    int64_t b = 0;
    if (valid) {
        b = free_list ? int64_t(free_list[gi]) : gi;
        const SampleId s = decode_sample(f, gen_base + gi);
        w.sample[b] = gen_base + gi;
        // camera subpath: get_ray, camera.h:199-223 (defocus off: scene_loader.h:463)
        Rng rc;
        rc.init(f.seed, 0u, s.sid, 0u);
        const double px = ((s.s_i + rc.next()) * f.recip_sqrt_spp) - 0.5;
        const double py = ((s.s_j + rc.next()) * f.recip_sqrt_spp) - 0.5;
        const D3 ps = f.pixel00 + ((s.i + px) * f.du) + ((s.j + py) * f.dv);
        const D3 d = ps - f.center;
        (void)rc.next();  // ray time (camera.h:210): drawn, never used
        store_path(w.path + b, f.center, d, mk(1.0, 1.0, 1.0), rc.draws, f.depth);
        if (MIS) {   // pdf of this primary ray per unit solid angle, for an image plane measured in pixels
            const double c = dot(unit(d), f.fwd);
            PathMis pm;
            pm.dVCM = f.n_light / (f.plane_px * f.plane_px / (c * c * c)); pm.dVC = 0.0;
            w.pmis[b] = pm;
        }
        w.ncam[b] = 0;
        cam_live = f.depth > 0;

        // light subpath: build_light_path, camera.h:372-418
        uint32_t nl = 0;
        const int64_t lp = w.B + b;
        if (f.depth > 0 && sc.n_lights > 0 && sc.light_total_area > 0.0) {
            Rng rl;
            rl.init(f.seed, 1u, s.sid, 0u);
            // triangle_collection::sample_surface, triangle.h:199-224: linear scan `pick <= accum`
            const double pick = rl.next(0.0, sc.light_total_area);
            uint32_t lo = 0, hi = sc.n_lights - 1;   // light_area holds the running sums (same order of adds)
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (pick <= __ldg(sc.light_area + mid)) hi = mid; else lo = mid + 1;
            }
            const uint32_t prim = __ldg(sc.lights + lo);
            double u = rl.next(), v = rl.next();      // triangle::sample, triangle.h:107-119
            if (u + v > 1.0) { u = 1.0 - u; v = 1.0 - v; }
            const TriGeom tg = sc.tris[prim];
            const D3 pos = mk(tg.v0x, tg.v0y, tg.v0z) + (u * mk(tg.e1x, tg.e1y, tg.e1z)) + (v * mk(tg.e2x, tg.e2y, tg.e2z));
            const TriShade sh = sc.shade[prim];
            const D3 n = mk(sh.nx, sh.ny, sh.nz);
            const double pdf = 1.0 / sc.light_total_area;
            const Mat m = sc.mats[sh.mat];
            // emitted(front = true, u = v = 0, p), camera.h:386-392
            const D3 emission = (m.type != BDPT_MAT_LIGHT) ? mk(0, 0, 0) : (TEX ? mat_colour(sc, m, pos, [](double& tu, double& tv) { tu = 0.0; tv = 0.0; }) : mk(m.r, m.g, m.b));
            if (len2(emission) > 0) {
                const D3 T0 = mk(1.0, 1.0, 1.0) / fmax(pdf, 1e-8);
                store_vertex(w.lv + size_t(b) * size_t(w.depth), pos, T0, prim | kLightFlag);
                if (TEX) w.lcol[size_t(b) * size_t(w.depth)] = emission;
                nl = 1;
                const D3 dir = Onb(n).transform(random_cosine_direction(rl));
                const D3 du = unit(dir);
                const double cos_theta = fmax(0.0, dot(n, du));
                if (cos_theta > 0 && f.depth - 1 > 0) {
                    const double pdf_dir = fmax(cos_theta / kPi, 1e-8);
                    const D3 T = T0 * emission * (cos_theta / pdf_dir);
                    const D3 o = pos + (0.001 * n);
                    store_path(w.path + lp, o, du, T, rl.draws, f.depth - 1);
                    if (MIS) {
                        const double emissionPdfW = pdf * pdf_dir;
                        PathMis pm;
                        pm.dVCM = pdf / emissionPdfW; pm.dVC = cos_theta / emissionPdfW;
                        w.pmis[lp] = pm;
                    }
                    light_live = true;
                }
            }
        }
        w.nlight[b] = nl;
        w.live[b] = uint32_t(cam_live) + uint32_t(light_live);
    }
    uint32_t slot = block_append<kGen>(w.counters + C_ACTIVE0 + cur, cam_live);
    if (cam_live) w.q_active[cur][slot] = uint32_t(b);
    slot = block_append<kGen>(w.counters + C_ACTIVE0 + cur, light_live);
    if (light_live) w.q_active[cur][slot] = uint32_t(w.B + b);
    const bool dead = valid && !cam_live && !light_live;   // nothing to trace: straight to the done list
    slot = block_append<kGen>(w.counters + C_DONE0 + dn, dead);
    if (dead) w.q_slots[dn][slot] = uint32_t(b);
}

// ------------------------------------------------------------------ K1
// world.hit over the live subpaths (camera.h:336: interval [0.001, inf)); a miss ends the subpath and,
// for a camera subpath, adds throughput * background (camera.h:337-340).  Hits are binned by material
// class so each shading kernel runs divergence-free over its own queue.
// result of one closest-hit query: record the hit and pick the material queue, or finish the subpath
// ------------------------------------------------------------------ participating media (constant_medium.h:8-61)
// boundary->hit(r, [tmin, tmax]) over the boundary triangles in order (triangle_collection::hit, triangle.h:150-164): only t matters
__device__ __forceinline__ bool boundary_hit(const TriGeom* __restrict__ med_tris, const MediumDev& m, const D3& o, const D3& d, double tmin,
                                             double tmax, double& t_out) {
    bool any = false;
    double closest = tmax;
    for (uint32_t i = 0; i < m.count; ++i) {
        double t, u, v;
        if (tri_candidate(med_tris + m.first + i, o, d, t, u, v) && tmin <= t && t <= closest) { any = true; closest = t; }
    }
    t_out = closest;
    return any;
}
// The same query through the tree over a large boundary (MedNode, built by bdpt_scene_set_media): the smallest candidate t inside
// [tmin, tmax], found by visiting only the leaves whose padded boxes the line crosses before the best t so far -- the same number
// the loop over all triangles returns.
__device__ __noinline__ bool boundary_hit_tree(const TriGeom* __restrict__ tris, const MedNode* __restrict__ nodes, D3 o, D3 d, double tmin, double tmax,
                                               double* t_out) {
    const double oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    double inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) inv[k] = 1.0 / dd[k];
    int32_t stack[24];   // <= 4096 triangles, median splits, leaves of <= 4: depth <= 11
    int sp = 0;
    int32_t cur = 0;
    bool any = false;
    double closest = tmax;
    for (;;) {
        const MedNode* nd = nodes + cur;
        double tn = tmin, tf = closest;
        bool hit = true;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double lo = nd->lo[k], hi = nd->hi[k];
            if (dd[k] == 0.0) { if (oo[k] < lo || oo[k] > hi) hit = false; }
            else {
                const double a = (lo - oo[k]) * inv[k], b = (hi - oo[k]) * inv[k];
                tn = fmax(tn, fmin(a, b)); tf = fmin(tf, fmax(a, b));
            }
        }
        if (hit && !(tn > tf)) {
            const int32_t a = nd->a, b = nd->b;
            if (b < 0) {
                for (int32_t i = 0; i < -b; ++i) {
                    double t, u, v;
                    if (tri_candidate(tris + a + i, o, d, t, u, v) && tmin <= t && t <= closest) { any = true; closest = t; }
                }
            } else {
                stack[sp++] = b;
                cur = a;
                continue;
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    *t_out = closest;
    return any;
}
// Both boundary hits from ONE walk of a large boundary's tree: every candidate t along the whole line is kept (a closed mesh gives a
// handful), then the two queries are answered from them as for a small boundary.  false: more than `cap` candidates -- ask twice.
__device__ __noinline__ bool boundary_candidates_tree(const TriGeom* __restrict__ tris, const MedNode* __restrict__ nodes, D3 o, D3 d, double* ts, uint32_t cap,
                                                      uint32_t* n_out) {
    const double oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    double inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) inv[k] = 1.0 / dd[k];
    int32_t stack[24];
    int sp = 0;
    int32_t cur = 0;
    uint32_t n = 0;
    for (;;) {
        const MedNode* nd = nodes + cur;
        double tn = -__longlong_as_double(0x7FF0000000000000LL), tf = __longlong_as_double(0x7FF0000000000000LL);
        bool hit = true;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double lo = nd->lo[k], hi = nd->hi[k];
            if (dd[k] == 0.0) { if (oo[k] < lo || oo[k] > hi) hit = false; }
            else {
                const double a = (lo - oo[k]) * inv[k], b = (hi - oo[k]) * inv[k];
                tn = fmax(tn, fmin(a, b)); tf = fmin(tf, fmax(a, b));
            }
        }
        if (hit && !(tn > tf)) {
            const int32_t a = nd->a, b = nd->b;
            if (b < 0) {
                for (int32_t i = 0; i < -b; ++i) {
                    double t, u, v;
                    if (tri_candidate(tris + a + i, o, d, t, u, v) && t == t) {
                        if (n == cap) return false;
                        ts[n++] = t;
                    }
                }
            } else {
                stack[sp++] = b;
                cur = a;
                continue;
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    *n_out = n;
    return true;
}
// The two boundary hits constant_medium::hit asks for (constant_medium.h:25-29): the nearest over the whole line, then the nearest from
// 0.0001 behind it.  Small boundaries (a box is 12 triangles) are met ONCE: the candidates' t are kept and the second query reads them
// back -- the same comparisons on the same numbers as two loops over the triangles.
__device__ __forceinline__ bool boundary_pair(const TriGeom* __restrict__ med_tris, const MediumDev& m, const D3& o, const D3& d, double& t1, double& t2) {
    const double inf = __longlong_as_double(0x7FF0000000000000LL);
    constexpr uint32_t kKeep = 16;
    double ts[kKeep];
    uint32_t n = 0;
    if (m.nodes) {
        if (!boundary_candidates_tree(med_tris + m.first, m.nodes, o, d, ts, kKeep, &n))
            return boundary_hit_tree(med_tris + m.first, m.nodes, o, d, -inf, inf, &t1) && boundary_hit_tree(med_tris + m.first, m.nodes, o, d, t1 + 0.0001, inf, &t2);
    } else if (m.count > kKeep) {
        return boundary_hit(med_tris, m, o, d, -inf, inf, t1) && boundary_hit(med_tris, m, o, d, t1 + 0.0001, inf, t2);
    } else {
        for (uint32_t i = 0; i < m.count; ++i) {
            double t, u, v;
            if (tri_candidate(med_tris + m.first + i, o, d, t, u, v) && t == t) ts[n++] = t;
        }
    }
    double c1 = inf;
    bool any = false;
    for (uint32_t i = 0; i < n; ++i)
        if (ts[i] <= c1) { any = true; c1 = ts[i]; }
    if (!any) return false;
    t1 = c1;
    const double from = t1 + 0.0001;
    double c2 = inf;
    any = false;
    for (uint32_t i = 0; i < n; ++i)
        if (from <= ts[i] && ts[i] <= c2) { any = true; c2 = ts[i]; }
    t2 = c2;
    return any;
}
// false: the part [tmin, tmax] of the ray certainly has nothing in common with the (padded) bounds of the boundary mesh.  Both boundary
// hits of constant_medium::hit, if they exist, lie inside the line's slab interval [tn, tf]; after its clamps to [tmin, tmax] the medium
// then sees t1 >= t2 and answers "no hit" WITHOUT a draw -- so skipping the boundary loops here changes nothing.
__device__ __forceinline__ bool line_meets_bounds(const MediumDev& m, const D3& o, const D3& d, double tmin, double tmax) {
    const double oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    double tn = -__longlong_as_double(0x7FF0000000000000LL), tf = __longlong_as_double(0x7FF0000000000000LL);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (dd[k] == 0.0) { if (oo[k] < m.lo[k] || oo[k] > m.hi[k]) return false; }
        else {
            const double a = (m.lo[k] - oo[k]) / dd[k], b = (m.hi[k] - oo[k]) / dd[k];
            tn = fmax(tn, fmin(a, b)); tf = fmin(tf, fmax(a, b));
        }
    }
    return !(tn > tf) && !(tf < tmin) && !(tn > tmax);
}
// constant_medium::hit, constant_medium.h:22-58 (oracle/port.cpp: medium_hit).  ONE draw, only when the part of the ray inside
// the boundary overlaps [tmin, tmax].
__device__ __forceinline__ bool medium_hit(const TriGeom* __restrict__ med_tris, const MediumDev& m, const D3& o, const D3& d, double tmin,
                                           double tmax, Rng& rng, double& t_out) {
    if (!line_meets_bounds(m, o, d, tmin, tmax)) return false;
    const double inf = __longlong_as_double(0x7FF0000000000000LL);
    double t1, t2;
    if (!boundary_pair(med_tris, m, o, d, t1, t2)) return false;
    if (t1 < tmin) t1 = tmin;
    if (t2 > tmax) t2 = tmax;
    if (t1 >= t2) return false;
    if (t1 < 0) t1 = 0;
    const double ray_length = len(d);
    const double distance_inside_boundary = (t2 - t1) * ray_length;
    const double hit_distance = m.neg_inv_density * det_log(rng.next());
    if (hit_distance > distance_inside_boundary) return false;
    t_out = t1 + hit_distance / ray_length;
    return true;
}

// The part of [tmin, tmax] the ray spends inside medium m as constant_medium::hit measures it (medium_hit without the draw);
// oracle/port.cpp: medium_overlap.
__device__ __forceinline__ double medium_overlap(const TriGeom* __restrict__ med_tris, const MediumDev& m, const D3& o, const D3& d, double tmin,
                                                 double tmax) {
    if (!line_meets_bounds(m, o, d, tmin, tmax)) return 0.0;
    const double inf = __longlong_as_double(0x7FF0000000000000LL);
    double t1, t2;
    if (!boundary_pair(med_tris, m, o, d, t1, t2)) return 0.0;
    if (t1 < tmin) t1 = tmin;
    if (t2 > tmax) t2 = tmax;
    if (t1 >= t2) return 0.0;
    if (t1 < 0) t1 = 0;
    return (t2 - t1) * len(d);
}
// Transmittance of all media along o + t d, t in [tmin, tmax] (oracle/port.cpp: transmittance): what BDPT_WEIGHT_MIS weighs with.
// Rare and heavy (a loop over every boundary triangle): kept out of line.
__device__ __noinline__ double media_transmittance(const MediumDev* __restrict__ media, const TriGeom* __restrict__ med_tris, uint32_t n_media,
                                                   D3 o, D3 d, double tmin, double tmax) {
    double tau = 0.0;
    for (uint32_t k = 0; k < n_media; ++k) {
        const MediumDev md = media[k];
        tau += medium_overlap(med_tris, md, o, d, tmin, tmax) / -md.neg_inv_density;
    }
    return tau > 0.0 ? exp(-tau) : 1.0;
}

constexpr int kClsIsotropic = 4;   // material queue of the scatter points inside media (after BDPT_MAT_* 0..3)

// analytic spheres: sphere_hit_t / sphere_normal live in scene.cuh (bdpt_trace asks the side list too)
// material index and geometric normal of a stored vertex's prim (triangle or, in the flavours that know them, sphere)
template <bool EX>
__device__ __forceinline__ uint32_t prim_mat(const SceneView& sc, uint32_t prim) {
    if (EX && (prim & kSpherePrim)) return sc.spheres[prim & ~kSpherePrim].mat;
    return sc.shade[prim].mat;
}

// world.hit's media for one path ray.  Out of line on purpose: inlined into finish_trace its fp64 triangle loops raise the
// register pressure of every traversal kernel (k_trace_fast: 150 -> 538 bytes of spills), scenes without media included.
struct MediaOut { double t; uint32_t prim, found; };
__device__ __noinline__ void spheres_after_triangles(const SphereDev* __restrict__ spheres, uint32_t n_spheres, const PathRec* path, MediaOut* io) {
    const double2* q = reinterpret_cast<const double2*>(path);
    const double2 q0 = q[0], q1 = q[1], q2 = q[2];
    const D3 o = mk(q0.x, q0.y, q1.x), d = mk(q1.y, q2.x, q2.y);
    double best = io->found ? io->t : __longlong_as_double(0x7FF0000000000000LL);
    for (uint32_t k = 0; k < n_spheres; ++k) {   // in list order, against the bound so far
        double ts;
        if (sphere_hit_t(spheres[k], o, d, 0.001, best, ts)) { io->found = 1u; best = ts; io->t = ts; io->prim = kSpherePrim | k; }
    }
}
__device__ __noinline__ void media_after_triangles(const MediumDev* __restrict__ media, const TriGeom* __restrict__ med_tris, uint32_t n_media,
                                                   PathRec* path, uint64_t seed, uint32_t stream, uint64_t sid, MediaOut* io) {
    const double2* q = reinterpret_cast<const double2*>(path);
    const double2 q0 = q[0], q1 = q[1], q2 = q[2];
    const D3 o = mk(q0.x, q0.y, q1.x), d = mk(q1.y, q2.x, q2.y);
    Rng rng;
    rng.init(seed, stream, sid, path->draws);
    const uint32_t d0 = rng.draws;
    double best = io->found ? io->t : __longlong_as_double(0x7FF0000000000000LL);
    for (uint32_t k = 0; k < n_media; ++k) {
        double tm;
        if (medium_hit(med_tris, media[k], o, d, 0.001, best, rng, tm)) { io->found = 1u; best = tm; io->t = tm; io->prim = kMediumPrim | k; }
    }
    if (rng.draws != d0) path->draws = rng.draws;
}

__device__ __forceinline__ RayD load_path_ray(const WF& w, uint32_t pid) {
    RayD r;
    const double2* q = reinterpret_cast<const double2*>(w.path + pid);
    const double2 q0 = q[0], q1 = q[1], q2 = q[2];
    r.o = mk(q0.x, q0.y, q1.x);
    r.d = mk(q1.y, q2.x, q2.y);
    r.tmin = 0.001;
    r.tmax = __longlong_as_double(0x7FF0000000000000LL);
    return r;
}
template <bool MEDIA = false>
__device__ __forceinline__ int finish_trace(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid, bool found,
                                            const HitD& h, double* film) {
    uint32_t prim = h.prim;
    double t = h.t;
    if (MEDIA && (sc.n_media | sc.n_spheres)) {   // the side lists of the world (sphere prims, media), asked after the triangles with the bound found so far (port: hit_world)
        const bool is_cam = pid < uint32_t(w.B);
        const uint32_t b = is_cam ? pid : pid - uint32_t(w.B);
        MediaOut mo;
        mo.found = found ? 1u : 0u; mo.prim = prim; mo.t = t;
        // the sphere prims of the side list first (their own small call: a scene of spheres only never pays the media's big one)
        if (sc.n_spheres) { spheres_after_triangles(sc.spheres, sc.n_spheres, w.path + pid, &mo); found = mo.found != 0; prim = mo.prim; t = mo.t; }
        if (sc.n_media) {
            media_after_triangles(sc.media, sc.med_tris, sc.n_media, w.path + pid, f.seed, is_cam ? 0u : 1u, decode_sample(f, w.sample[b]).sid, &mo);
            found = mo.found != 0; prim = mo.prim; t = mo.t;
        }
    }
    if (found) {
        w.path[pid].hit_prim = prim;
        w.path[pid].hit_t = t;
        if (MEDIA && (prim & kMediumPrim)) return kClsIsotropic;
        if (MEDIA && (prim & kSpherePrim)) return int(sc.mats[sc.spheres[prim & ~kSpherePrim].mat].type);
        return int(sc.mats[__ldg(&sc.shade[prim].mat)].type);
    }
    if (pid < uint32_t(w.B)) {
        const PathRec& pr = w.path[pid];
        const D3 T = mk(pr.tx, pr.ty, pr.tz);
        const D3 c = T * f.background;
        if (c.x != 0.0 || c.y != 0.0 || c.z != 0.0) film_add(film, decode_sample(f, w.sample[pid]).pix, c);
    }
    return -1;
}
template <bool MEDIA = false>
__device__ __forceinline__ void bin_by_material(const WF& w, int cls, uint32_t pid) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t slot = warp_append(w.counters + C_MAT0 + k, cls == k);
        if (cls == k) w.q_mat[k][slot] = pid;
    }
    if (MEDIA && __any_sync(0xFFFFFFFFu, cls == kClsIsotropic)) {   // scenes with media only
        const uint32_t slot = warp_append(w.counters + C_MAT4, cls == kClsIsotropic);
        if (cls == kClsIsotropic) w.q_mat[kClsIsotropic][slot] = pid;
    }
}

// Traversal stacks live in local memory (L1-cached, lane-interleaved by the hardware): a shared-memory stack
// of tree-depth levels cost 24.5 KB per block, capped residency at 6 blocks per SM and shrank L1.
#ifndef BDPT_WIDE_WALK
#define BDPT_WIDE_WALK 0   // 1: bulk rays take the warp-wide lock-step walk with postponed leaves (Walk::run_wide): parity-green, measured 25 % slower
#endif
#ifndef BDPT_TRAV_BLOCKS
#define BDPT_TRAV_BLOCKS 8
#endif
constexpr int kTravBlocksPerSM = BDPT_TRAV_BLOCKS * 128 / kBlock;   // residency target of the traversal kernels (register cap = 65536 / (128 * this))
constexpr int kStackLevels = 32;              // StackEntry slots (fast tree, depth <= 23) = 64 int32 slots (reference tree)
#define BDPT_DECLARE_STACK StackEntry l_stack[kStackLevels]

// Blocks [0, n_def_blocks) re-trace the rays the PREVIOUS iteration's launch deferred (ties, knife edges,
// zero direction components: traverse_fast.cuh) through the reference-exact traversal; there are a handful
// per launch but each walks ~1000 nodes of the reference tree, so instead of a launch of their own (whose
// duration would be one thread's latency chain) they ride at the front of the next bulk launch and their
// paths simply lag one iteration behind — path regeneration makes that free.
template <int BVH, bool STATS, bool MEDIA = false>
__global__ void __launch_bounds__(kBlock, kTravBlocksPerSM) k_trace_paths(SceneView sc, Frame f, WF w, const uint32_t* __restrict__ queue,
                                                       const uint32_t* __restrict__ n_ptr, int dn, int ex, uint32_t n_def_blocks,
                                                       uint32_t limit, double* film, unsigned long long* stats) {
    BDPT_DECLARE_STACK;
    int cls = -1;
    uint32_t pid = 0, defer_tag = 0;
    bool defer = false, ended = false;
    if (blockIdx.x < n_def_blocks) {
        const uint32_t idx = blockIdx.x * kFrontPerBlock + (threadIdx.x >> 5);
        if ((threadIdx.x & 31) == 0 && idx < w.counters[C_EXACT0 + (ex ^ 1)]) {
            const uint32_t e = w.q_exact[ex ^ 1][idx];
            pid = e & ~kSlowFlag;
            const RayD r = load_path_ray(w, pid);
            HitD h;
            TravStats ts{0, 0};
            bool found;
            if (BVH != BDPT_BVH_REFERENCE && (e & kSlowFlag))
                found = traverse_fast<false, STATS, false>(sc, r, h, l_stack, 1, &ts) == TR_HIT;
            else
                found = traverse_ref<false, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts);
            if (STATS) add_trav<0, BVH == BDPT_BVH_REFERENCE>(stats, ts);
            cls = finish_trace<MEDIA>(sc, f, w, pid, found, h, film);
            ended = cls < 0;
        }
    } else {
        const uint32_t idx = (blockIdx.x - n_def_blocks) * kBlock + threadIdx.x;
        if (idx < *n_ptr) {
            pid = queue[idx];
            const RayD r = load_path_ray(w, pid);
            HitD h;
            bool found;
            TravStats ts{0, 0};
            if (BVH == BDPT_BVH_REFERENCE) {
                found = traverse_ref<false, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts);
            } else {
                const int rc = traverse_fast<false, STATS, true>(sc, r, h, l_stack, 1, &ts, limit);
                found = rc == TR_HIT;
                defer = rc == TR_DEFER || rc == TR_SLOW;
                defer_tag = rc == TR_SLOW ? kSlowFlag : 0u;
            }
            if (STATS) add_trav<0, BVH == BDPT_BVH_REFERENCE>(stats, ts);
            if (!defer) { cls = finish_trace<MEDIA>(sc, f, w, pid, found, h, film); ended = cls < 0; }
        }
    }
    if (BVH != BDPT_BVH_REFERENCE) {   // rays that need the reference tree: compacted for the next launch's front blocks
        const uint32_t slot = warp_append(w.counters + C_EXACT0 + ex, defer);
        if (defer) w.q_exact[ex][slot] = pid | defer_tag;
    }
    bin_by_material<MEDIA>(w, cls, pid);
    retire(w, ended, pid < uint32_t(w.B) ? pid : pid - uint32_t(w.B), dn);
}

// ------------------------------------------------------------------ K1, fast flavour: persistent warps with dynamic fetch
// Every warp keeps pulling rays from the queue (one atomic per refill) into the lanes whose ray is decided,
// and its walk (traverse_walk.cuh) hands control back as soon as fewer than `threshold` lanes are still
// walking.  All queue appends happen at the top of the loop, where the warp is converged.
// Blocks [0, n_def_blocks) are the front blocks described above k_trace_paths.
// a refill is worth its atomic and its ray set-up only when at least `min_idle` lanes are idle (or all are)
__device__ __forceinline__ uint32_t warp_fetch(uint32_t* cursor, bool need, uint32_t n, bool& exhausted, int min_idle) {
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, need);
    if (!mask || exhausted) return 0xFFFFFFFFu;
    if (mask != 0xFFFFFFFFu && __popc(mask) < min_idle) return 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(cursor, uint32_t(__popc(mask)));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (base + uint32_t(__popc(mask)) >= n) exhausted = true;   // nothing left after this batch
    const uint32_t my = base + uint32_t(__popc(mask & ((1u << lane) - 1u)));
    return (need && my < n) ? my : 0xFFFFFFFFu;
}

template <bool STATS, bool MEDIA = false>
__global__ void __launch_bounds__(kBlock, kTravBlocksPerSM) k_trace_fast(SceneView sc, Frame f, WF w, const uint32_t* __restrict__ queue,
                                                                        const uint32_t* __restrict__ n_ptr, int dn, int ex,
                                                                        uint32_t n_def_blocks, int threshold, uint32_t limit, double* film,
                                                                        unsigned long long* stats) {
    BDPT_DECLARE_STACK;
    if (blockIdx.x < n_def_blocks) {
        int cls = -1;
        uint32_t pid = 0;
        bool ended = false;
        const uint32_t idx = blockIdx.x * kFrontPerBlock + (threadIdx.x >> 5);
        const bool lane0 = (threadIdx.x & 31) == 0;
        if (idx < w.counters[C_EXACT0 + (ex ^ 1)]) {   // warp-uniform
            const uint32_t e = w.q_exact[ex ^ 1][idx];
            pid = e & ~kSlowFlag;
            const RayD r = load_path_ray(w, pid);
            HitD h;
            TravStats ts{0, 0};
            // over the work limit: the whole warp walks it, split 32 ways; anything else (or a tie found on the
            // way): the reference traversal on lane 0
            int rc = TR_DEFER;
            if (e & kSlowFlag) rc = split_walk<false, STATS>(sc, r, h, l_stack, ts);
            if (lane0) {
            const bool found = rc == TR_DEFER ? traverse_ref<false, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts) : rc == TR_HIT;
            if (STATS) add_trav<0, false>(stats, ts);
            cls = finish_trace<MEDIA>(sc, f, w, pid, found, h, film);
            ended = cls < 0;
            }
        }
        bin_by_material<MEDIA>(w, cls, pid);
        retire(w, ended, pid < uint32_t(w.B) ? pid : pid - uint32_t(w.B), dn);
        return;
    }
    const uint32_t n = *n_ptr;
    Walk<false, STATS> W;
    bool have = false, exhausted = false;
    int result = WALK_YIELD;   // != WALK_YIELD: this lane's ray `pid` is decided and waits for the epilogue
    uint32_t pid = 0;
    for (;;) {
        __syncwarp();
        // ---- epilogue of the lanes whose ray was decided in the last round
        const bool fin = result != WALK_YIELD;
        if (__any_sync(0xFFFFFFFFu, fin)) {
            int cls = -1;
            bool ended = false, defer = false;
            if (fin) {
                if (STATS && result != TR_DEFER && result != TR_SLOW) add_trav<0, false>(stats, W.ts);
                if (result == TR_DEFER || result == TR_SLOW) defer = true;
                else { cls = finish_trace<MEDIA>(sc, f, w, pid, result == TR_HIT, W.hit, film); ended = cls < 0; }
            }
            const uint32_t slot = warp_append(w.counters + C_EXACT0 + ex, defer);
            if (defer) w.q_exact[ex][slot] = pid | (result == TR_SLOW ? kSlowFlag : 0u);
            bin_by_material<MEDIA>(w, cls, pid);
            retire(w, ended, pid < uint32_t(w.B) ? pid : pid - uint32_t(w.B), dn);
            result = WALK_YIELD;
        }
        // ---- refill the idle lanes
        const uint32_t my = warp_fetch(w.counters + C_CURSOR, !have, n, exhausted, (threshold >> 8) & 0xFF);
        if (my != 0xFFFFFFFFu) {
            pid = queue[my];
            const RayD r = load_path_ray(w, pid);
            if (walk_eligible(sc, r)) { W.init(sc, r); have = true; }
            else result = TR_DEFER;   // fp64 slabs or the reference tree: the next launch's front blocks
        }
        if (!__any_sync(0xFFFFFFFFu, have || result != WALK_YIELD)) break;
        // ---- walk
#if BDPT_WIDE_WALK
        {
            const int rc = W.run_wide(sc, l_stack, have, threshold & 0xFF, (threshold >> 16) & 0xFF, (threshold >> 24) & 0xFF,
                                      exhausted ? 0 : ((threshold >> 8) & 0xFF), limit);
            if (have && rc != WALK_YIELD) { result = rc; have = false; }
        }
#else
        if (have) {
            const int rc = W.run(sc, l_stack, exhausted ? 0 : (threshold & 0xFF), limit);
            if (rc != WALK_YIELD) { result = rc; have = false; }
        }
#endif
    }
}


// ------------------------------------------------------------------ K1, round 2: warp-synchronous fp32 walk + exact fp64 finish
// (traverse_w32.cuh).  Same persistent-warp frame as k_trace_fast — dynamic fetch, front blocks for the rays
// that need the reference tree — but the 32 lanes of a warp take every loop trip together, the walk carries no fp64
// state (the ray is re-read from its path record for the exact tests at the meeting point), and the register
// budget buys more resident warps.
#ifndef BDPT_W32_BLOCKS
#define BDPT_W32_BLOCKS 16
#endif
constexpr int kW32BlocksPerSM = BDPT_W32_BLOCKS;   // blocks of 64 threads per SM (register cap = 65536 / (64 * this))

// knobs: bits 0-5 trips per slice | 6-9 node steps per trip | 10-15 idle lanes that end a slice (refill) | 16-23 waiting
// leaves that start the pre-filter | 24-31 walkers at or below which it starts anyway
struct W32Knobs { int trips, nstep, min_idle, t_leaf, t_node; };
__device__ __forceinline__ W32Knobs unpack_knobs(int k) {
    W32Knobs r;
    r.trips = k & 0x3F; r.nstep = (k >> 6) & 0xF; r.min_idle = (k >> 10) & 0x3F; r.t_leaf = (k >> 16) & 0xFF; r.t_node = (k >> 24) & 0xFF;
    return r;
}

template <bool STATS>
__global__ void __launch_bounds__(kBlock, kW32BlocksPerSM) k_trace_w32(SceneView sc, Frame f, WF w, const uint32_t* __restrict__ queue,
                                                                       const uint32_t* __restrict__ n_ptr, int dn, int ex,
                                                                       uint32_t n_def_blocks, int knobs, double* film,
                                                                       unsigned long long* stats) {
    BDPT_DECLARE_STACK;
    if (blockIdx.x < n_def_blocks) {   // rays the previous launch handed to the reference tree, one per warp (lane 0)
        int cls = -1;
        uint32_t pid = 0;
        bool ended = false;
        const uint32_t idx = blockIdx.x * kFrontPerBlock + (threadIdx.x >> 5);
        if ((threadIdx.x & 31) == 0 && idx < w.counters[C_EXACT0 + (ex ^ 1)]) {
            pid = w.q_exact[ex ^ 1][idx] & ~kSlowFlag;
            const RayD r = load_path_ray(w, pid);
            HitD h;
            TravStats ts{0, 0};
            const bool found = traverse_ref<false, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts);
            if (STATS) add_trav<0, true>(stats, ts);
            cls = finish_trace(sc, f, w, pid, found, h, film);
            ended = cls < 0;
        }
        bin_by_material(w, cls, pid);
        retire(w, ended, pid < uint32_t(w.B) ? pid : pid - uint32_t(w.B), dn);
        return;
    }
    const uint32_t n = *n_ptr;
    const W32Knobs kn = unpack_knobs(knobs);
    Walk32<false, STATS> W;
    W.cur = kWalkDry;
    bool have = false, exhausted = false;
    int result = WALK_YIELD;   // TR_HIT: walk over, decide | WALK32_FLUSH | TR_DEFER: not walkable here
    uint32_t pid = 0;
    for (;;) {
        __syncwarp();
        if (result == WALK32_FLUSH) {   // candidate list full: the exact test keeps the winner, the walk goes on
            HitD h;
            W.decide(sc, load_path_ray(w, pid), h, false);
            result = WALK_YIELD;
        }
        const bool fin = result != WALK_YIELD;
        if (__any_sync(0xFFFFFFFFu, fin)) {
            int cls = -1;
            bool ended = false, defer = false;
            if (fin) {
                int rc = result;
                HitD h;
                if (rc == TR_HIT) rc = W.decide(sc, load_path_ray(w, pid), h, true);
                if (rc == TR_DEFER) defer = true;
                else {
                    if (STATS) add_trav<0, false>(stats, W.ts);
                    cls = finish_trace(sc, f, w, pid, rc == TR_HIT, h, film);
                    ended = cls < 0;
                }
                have = false;
                result = WALK_YIELD;
            }
            const uint32_t slot = warp_append(w.counters + C_EXACT0 + ex, defer);
            if (defer) w.q_exact[ex][slot] = pid;
            bin_by_material(w, cls, pid);
            retire(w, ended, pid < uint32_t(w.B) ? pid : pid - uint32_t(w.B), dn);
        }
        const uint32_t my = warp_fetch(w.counters + C_CURSOR, !have, n, exhausted, kn.min_idle);
        if (my != 0xFFFFFFFFu) {
            pid = queue[my];
            const RayD r = load_path_ray(w, pid);
            if (walk_eligible(sc, r)) { W.init(sc, r); have = true; }
            else result = TR_DEFER;   // fp64 slabs or the reference tree: the next launch's front blocks
        }
        if (!__any_sync(0xFFFFFFFFu, have || result != WALK_YIELD)) break;
        const int rc = W.run(sc, l_stack, have, kn.trips, kn.nstep, kn.t_leaf, kn.t_node, exhausted ? 0 : kn.min_idle);
        if (have) result = rc;
    }
}

// hit point and face-forwarded geometric normal (ray.h:21-23, hittable.h:20-26)
struct Surf { D3 o, d, T, p, n; double t; bool front; uint32_t prim, draws; int32_t remaining; };
template <bool EX = false>
__device__ __forceinline__ Surf load_surf(const SceneView& sc, const WF& w, uint32_t pid) {
    Surf s;
    const double2* q = reinterpret_cast<const double2*>(w.path + pid);
    const double2 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3], q4 = q[4];
    const int4 q5 = reinterpret_cast<const int4*>(w.path + pid)[5];
    s.o = mk(q0.x, q0.y, q1.x);
    s.d = mk(q1.y, q2.x, q2.y);
    s.T = mk(q3.x, q3.y, q4.x);
    s.prim = uint32_t(q5.y);
    s.draws = uint32_t(q5.x);
    s.remaining = q5.z;
    const double t = q4.y;
    s.t = t;
    s.p = s.o + t * s.d;
    D3 n;
    if (EX && (s.prim & kSpherePrim)) n = sphere_normal(sc.spheres[s.prim & ~kSpherePrim], s.p);
    else { const TriShade sh = sc.shade[s.prim]; n = mk(sh.nx, sh.ny, sh.nz); }
    s.front = dot(s.d, n) < 0;
    s.n = s.front ? n : -n;
    return s;
}

// (u, v) of the hit as triangle::hit left them in the record (triangle.h:66-67): the same arithmetic on the same inputs
struct HitUV {
    const SceneView& sc; const Surf& s;
    __device__ __forceinline__ void operator()(double& u, double& v) const {
        double t;
        if ((s.prim & kSpherePrim) || !tri_candidate(sc.tris + s.prim, s.o, s.d, t, u, v)) { u = 0.0; v = 0.0; }   // a sphere prim: u = v = 0
    }
};

// BDPT_WEIGHT_MIS: the subpath's running weight quantities as they stand ON this vertex (the hit just found):
// the segment's area conversion and the arrival cosine go in
struct MisAtHit { double dVCM, dVC, cos_in; uint32_t len; };
template <bool EX = false>
__device__ __forceinline__ MisAtHit mis_arrive(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid, const Surf& s) {
    MisAtHit m;
    const PathMis pm = w.pmis[pid];
    m.cos_in = fabs(dot(s.n, unit(-s.d)));
    double dist2 = s.t * s.t * len2(s.d);
    // media: the pdf of getting this far holds the segment's transmittance -- 1/dist^2 becomes Tr/dist^2 (port: walk_mis)
    if (EX && sc.n_media) dist2 = dist2 / media_transmittance(sc.media, sc.med_tris, sc.n_media, s.o, s.d, 0.001, s.t);
    m.dVCM = pm.dVCM * dist2 / m.cos_in;
    m.dVC = pm.dVC / m.cos_in;
    m.len = uint32_t(f.depth - s.remaining + 1);   // camera: remaining starts at depth; light: at depth - 1 behind the emitter
    return m;
}

__device__ __forceinline__ void store_ray(const WF& w, uint32_t pid, D3 o, D3 d, D3 T, uint32_t draws, int32_t remaining) {
    store_path(w.path + pid, o, d, T, draws, remaining);
}

// ------------------------------------------------------------------ K2: diffuse_light (material.h:133-147)
// scatter() is false: the subpath ends here.  Camera subpath: bidirectional_color adds throughput *
// emission for this non-delta vertex (camera.h:305-309); it never connects (evaluate_bsdf = 0).
// Light subpath: the vertex is connectable with f_light = emission when hit from the front (camera.h:463).
template <bool MIS, bool TEX, bool EX = false>
__device__ __forceinline__ bool shade_light(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid, double* film) {
    {
        const bool is_cam = pid < uint32_t(w.B);
        const uint32_t b = is_cam ? pid : pid - uint32_t(w.B);
        const Surf s = load_surf<EX>(sc, w, pid);
        const Mat m = sc.mats[prim_mat<EX>(sc, s.prim)];
        const D3 Le = TEX ? mat_colour(sc, m, s.p, HitUV{sc, s}) : mk(m.r, m.g, m.b);
        if (s.front && len2(Le) > 0) {   // emitted() is one-sided (material.h:140)
            if (is_cam) {
                double wgt = 1.0;
                if (MIS) {   // s = 0: weighed against every way the light subpath could have produced this path
                    const MisAtHit h = mis_arrive<EX>(sc, f, w, pid, s);
                    if (h.len > 1) {
                        const double pdfA = 1.0 / sc.light_total_area;
                        wgt = 1.0 / (1.0 + pdfA * h.dVCM + (pdfA * h.cos_in / kPi) * h.dVC);
                    }
                }
                film_add(film, decode_sample(f, w.sample[b]).pix, MIS ? wgt * (s.T * Le) : s.T * Le);
            } else if (!MIS) {   // MIS: an emitter absorbs; a light-subpath vertex on it connects to nothing
                const uint32_t slot = w.nlight[b];
                store_vertex(w.lv + size_t(b) * size_t(w.depth) + slot, s.p, s.T, s.prim | kLightFlag, uint32_t(f.depth - s.remaining));
                if (TEX) w.lcol[size_t(b) * size_t(w.depth) + slot] = Le;
                w.nlight[b] = slot + 1;
            }
        }
    }
    return false;   // scatter() is false: the subpath always ends on a light
}

// ------------------------------------------------------------------ K2: lambertian (material.h:42-67)
template <bool MIS, bool TEX, bool EX = false>
__device__ __forceinline__ bool shade_lambert(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid) {
    bool live = false;
    {
        const Surf s = load_surf<EX>(sc, w, pid);
        MisAtHit mh{0.0, 0.0, 0.0, 0u};
        if (MIS) mh = mis_arrive<EX>(sc, f, w, pid, s);
        const uint32_t side = (MIS && !s.front) ? kBackFlag : 0u;
        const Mat m = sc.mats[prim_mat<EX>(sc, s.prim)];
        const D3 albedo = TEX ? mat_colour(sc, m, s.p, HitUV{sc, s}) : mk(m.r, m.g, m.b);   // tex->value(rec.u, rec.v, rec.p), material.h:47,61
        const bool is_cam = pid < uint32_t(w.B);
        const uint32_t b = is_cam ? pid : pid - uint32_t(w.B);
        // vertex push (camera.h:342-349) — only vertices that can contribute: evaluate_bsdf = albedo/pi
        // (material.h:60-63); an all-zero BSDF makes connect_vertices return 0 (camera.h:460,468)
        const D3 fr = albedo / kPi;
        if (len2(fr) > 0) {
            if (is_cam) {
                const uint32_t slot = w.ncam[b];
                const D3 tf = s.T * fr;   // camera.h:471: cam_v.throughput * f_cam
                store_vertex(w.cv + size_t(b) * size_t(w.depth) + slot, s.p, tf, s.prim | side, uint32_t(f.depth - s.remaining));
                if (MIS) { VertMis vm; vm.dVCM = mh.dVCM; vm.dVC = mh.dVC; vm.cos_in = mh.cos_in; vm.len = mh.len; vm.pad = 0; w.cvm[size_t(b) * size_t(w.depth) + slot] = vm; }
                w.ncam[b] = slot + 1;
            } else {
                const uint32_t slot = w.nlight[b];
                store_vertex(w.lv + size_t(b) * size_t(w.depth) + slot, s.p, s.T, s.prim | side, uint32_t(f.depth - s.remaining));
                if (TEX) w.lcol[size_t(b) * size_t(w.depth) + slot] = albedo;
                if (MIS) { VertMis vm; vm.dVCM = mh.dVCM; vm.dVC = mh.dVC; vm.cos_in = mh.cos_in; vm.len = mh.len; vm.pad = 0; w.lvm[size_t(b) * size_t(w.depth) + slot] = vm; }
                w.nlight[b] = slot + 1;
            }
        }
        const int32_t remaining = s.remaining - 1;
        if (remaining > 0) {
            // scatter: cosine_pdf about the face-forwarded normal (material.h:47-52, pdf.h:32-47, onb.h:6-11)
            Rng rng;
            rng.init(f.seed, is_cam ? 0u : 1u, decode_sample(f, w.sample[b]).sid, s.draws);
            const Onb uvw(s.n);
            const D3 sd = uvw.transform(random_cosine_direction(rng));
            const double pdf_value = fmax(0.0, dot(unit(sd), uvw.a2) / kPi);   // pdf.h:37-40
            if (pdf_value > 0) {                                               // camera.h:362-363
                const double ct = dot(s.n, unit(sd));                          // material.h:55-58
                const double spdf = ct < 0 ? 0 : ct / kPi;
                const D3 T = s.T * albedo * spdf / pdf_value;                  // camera.h:366
                store_ray(w, pid, s.p, sd, T, rng.draws, remaining);           // camera.h:367: ray(rec.p, dir)
                if (MIS) {   // the sampled direction's pdf forwards, the arrival direction's pdf backwards
                    PathMis pm;
                    pm.dVC = (ct / pdf_value) * (mh.dVC * (mh.cos_in / kPi) + mh.dVCM);
                    pm.dVCM = 1.0 / pdf_value;
                    w.pmis[pid] = pm;
                }
                live = true;
            }
        }
    }
    return live;
}

// ------------------------------------------------------------------ K2: isotropic (material.h:149-171), the phase function of a
// constant_medium (constant_medium.h:14,54).  The vertex is a point INSIDE the medium: no triangle, the reference's arbitrary normal
// (1,0,0) and front_face = true (constant_medium.h:52-53).  Not delta: evaluate_bsdf = albedo / (4 pi) (material.h:166-168), so the
// vertex is stored and connected; scatter: attenuation = albedo, sphere_pdf (pdf.h:18-29): direction = random_unit_vector(), pdf =
// scattering_pdf = 1 / (4 pi) (material.h:154-164; camera.h:361-368).  BDPT_WEIGHT_MIS: a connectable vertex without a cosine, 1 / (4 pi) in
// both directions (oracle/port.cpp: walk_mis).
template <bool MIS>
__device__ __forceinline__ bool shade_isotropic(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid) {
    const double2* q = reinterpret_cast<const double2*>(w.path + pid);
    const double2 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3], q4 = q[4];
    const int4 q5 = reinterpret_cast<const int4*>(w.path + pid)[5];
    const D3 o = mk(q0.x, q0.y, q1.x), d = mk(q1.y, q2.x, q2.y), T = mk(q3.x, q3.y, q4.x);
    const D3 p = o + q4.y * d;   // rec.p = r.at(rec.t), constant_medium.h:50
    const uint32_t prim = uint32_t(q5.y), draws = uint32_t(q5.x);
    const int32_t remaining_in = q5.z;
    const MediumDev md = sc.media[prim & ~kMediumPrim];
    const D3 albedo = mk(md.r, md.g, md.b);
    const bool is_cam = pid < uint32_t(w.B);
    const uint32_t b = is_cam ? pid : pid - uint32_t(w.B);
    const uint32_t index = uint32_t(f.depth - remaining_in);
    const D3 fr = albedo / (4 * kPi);
    VertMis vm{0.0, 0.0, 1.0, 0u, 0u};
    if (MIS) {   // the running weight quantities as they stand on this vertex: no arrival cosine, the segment's transmittance
        const PathMis pm = w.pmis[pid];
        const double dist2 = q4.y * q4.y * len2(d);
        vm.dVCM = pm.dVCM * (dist2 / media_transmittance(sc.media, sc.med_tris, sc.n_media, o, d, 0.001, q4.y));
        vm.dVC = pm.dVC;
        vm.len = uint32_t(f.depth - remaining_in + 1);
    }
    if (len2(fr) > 0) {
        if (is_cam) {
            const uint32_t slot = w.ncam[b];
            store_vertex(w.cv + size_t(b) * size_t(w.depth) + slot, p, T * fr, prim, index);
            if (MIS) w.cvm[size_t(b) * size_t(w.depth) + slot] = vm;
            w.ncam[b] = slot + 1;
        } else {
            const uint32_t slot = w.nlight[b];
            store_vertex(w.lv + size_t(b) * size_t(w.depth) + slot, p, T, prim, index);
            if (w.lcol) w.lcol[size_t(b) * size_t(w.depth) + slot] = albedo;
            if (MIS) w.lvm[size_t(b) * size_t(w.depth) + slot] = vm;
            w.nlight[b] = slot + 1;
        }
    }
    const int32_t remaining = remaining_in - 1;
    if (remaining <= 0) return false;
    Rng rng;
    rng.init(f.seed, is_cam ? 0u : 1u, decode_sample(f, w.sample[b]).sid, draws);
    const D3 sd = random_unit_vector(rng);
    const double pdf_value = 1 / (4 * kPi);
    const double spdf = 1 / (4 * kPi);
    const D3 T2 = T * albedo * spdf / pdf_value;   // camera.h:366
    store_ray(w, pid, p, sd, T2, rng.draws, remaining);
    if (MIS) {
        PathMis pm;
        pm.dVC = (1.0 / pdf_value) * (vm.dVC * spdf + vm.dVCM);
        pm.dVCM = 1.0 / pdf_value;
        w.pmis[pid] = pm;
    }
    return true;
}

// ------------------------------------------------------------------ K2: metal (material.h:69-90)
// Always a delta vertex, even when fuzzy (material.h:85): never stored, never connected.  The fuzzed
// direction is not rejected when it points below the surface (material.h:74-82).
template <bool MIS, bool EX = false>
__device__ __forceinline__ bool shade_metal(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid) {
    bool live = false;
    {
        const int32_t remaining = w.path[pid].remaining - 1;
        if (remaining > 0) {
            const Surf s = load_surf<EX>(sc, w, pid);
            const Mat m = sc.mats[prim_mat<EX>(sc, s.prim)];
            const bool is_cam = pid < uint32_t(w.B);
            const uint32_t b = is_cam ? pid : pid - uint32_t(w.B);
            Rng rng;
            rng.init(f.seed, is_cam ? 0u : 1u, decode_sample(f, w.sample[b]).sid, s.draws);
            D3 refl = reflect(s.d, s.n);
            refl = unit(refl) + (m.param * random_unit_vector(rng));
            const D3 T = s.T * mk(m.r, m.g, m.b);                              // camera.h:355
            store_ray(w, pid, s.p, refl, T, rng.draws, remaining);
            if (MIS) {   // a delta vertex (material.h:85): nothing connects to it
                const MisAtHit mh = mis_arrive<EX>(sc, f, w, pid, s);
                PathMis pm;
                pm.dVCM = 0.0; pm.dVC = mh.dVC * fabs(dot(s.n, unit(refl)));
                w.pmis[pid] = pm;
            }
            live = true;
        }
    }
    return live;
}

// ------------------------------------------------------------------ K2: dielectric (material.h:92-131)
template <bool MIS, bool EX = false>
__device__ __forceinline__ bool shade_dielectric(const SceneView& sc, const Frame& f, const WF& w, uint32_t pid) {
    bool live = false;
    {
        const int32_t remaining = w.path[pid].remaining - 1;
        if (remaining > 0) {
            const Surf s = load_surf<EX>(sc, w, pid);
            const Mat m = sc.mats[prim_mat<EX>(sc, s.prim)];
            const bool is_cam = pid < uint32_t(w.B);
            const uint32_t b = is_cam ? pid : pid - uint32_t(w.B);
            Rng rng;
            rng.init(f.seed, is_cam ? 0u : 1u, decode_sample(f, w.sample[b]).sid, s.draws);
            const double ri = s.front ? (1.0 / m.param) : m.param;
            const D3 ud = unit(s.d);
            const double cos_theta = fmin(dot(-ud, s.n), 1.0);
            const double sin_theta = sqrt(1.0 - cos_theta * cos_theta);
            bool refl = ri * sin_theta > 1.0;                                   // total internal reflection
            if (!refl) {                                                        // Schlick, material.h:125-130
                double r0 = (1 - ri) / (1 + ri);
                r0 = r0 * r0;
                const double rf = r0 + (1 - r0) * det_pow5(1 - cos_theta);
                refl = rf > rng.next();
            }
            const D3 dir = refl ? reflect(ud, s.n) : refract(ud, s.n, ri);
            const D3 T = s.T * mk(1.0, 1.0, 1.0);
            store_ray(w, pid, s.p, dir, T, rng.draws, remaining);
            if (MIS) {
                const MisAtHit mh = mis_arrive<EX>(sc, f, w, pid, s);
                PathMis pm;
                pm.dVCM = 0.0; pm.dVC = mh.dVC * fabs(dot(s.n, unit(dir)));
                w.pmis[pid] = pm;
            }
            live = true;
        }
    }
    return live;
}

// ------------------------------------------------------------------ K2: one launch over the four material queues
// Thread idx walks the concatenation light | lambertian | metal | dielectric of the queues k_trace_* filled, so
// a warp straddles at most one class boundary; the survivors' queue append and the retirement of ended
// subpaths are shared.
#ifndef BDPT_SHADE_BLOCKS
#define BDPT_SHADE_BLOCKS 3
#endif
template <bool MIS, bool TEX, bool MEDIA = false>
__global__ void __launch_bounds__(kWide, BDPT_SHADE_BLOCKS) k_shade(SceneView sc, Frame f, WF w, int nxt, int dn, double* film) {
    const uint32_t idx = blockIdx.x * kWide + threadIdx.x;
    const uint32_t n_li = w.counters[C_MAT0 + BDPT_MAT_LIGHT], n_la = w.counters[C_MAT0 + BDPT_MAT_LAMBERTIAN],
                   n_me = w.counters[C_MAT0 + BDPT_MAT_METAL], n_di = w.counters[C_MAT0 + BDPT_MAT_DIELECTRIC];
    const uint32_t n_iso = MEDIA ? w.counters[C_MAT4] : 0u;
    const bool valid = idx < n_li + n_la + n_me + n_di + n_iso;
    bool live = false;
    uint32_t pid = 0;
    if (valid) {
        if (idx < n_li) { pid = w.q_mat[BDPT_MAT_LIGHT][idx]; live = shade_light<MIS, TEX, MEDIA>(sc, f, w, pid, film); }
        else if (idx < n_li + n_la) { pid = w.q_mat[BDPT_MAT_LAMBERTIAN][idx - n_li]; live = shade_lambert<MIS, TEX, MEDIA>(sc, f, w, pid); }
        else if (idx < n_li + n_la + n_me) { pid = w.q_mat[BDPT_MAT_METAL][idx - n_li - n_la]; live = shade_metal<MIS, MEDIA>(sc, f, w, pid); }
        else if (idx < n_li + n_la + n_me + n_di) { pid = w.q_mat[BDPT_MAT_DIELECTRIC][idx - n_li - n_la - n_me]; live = shade_dielectric<MIS, MEDIA>(sc, f, w, pid); }
        else if (MEDIA) { pid = w.q_mat[kClsIsotropic][idx - n_li - n_la - n_me - n_di]; live = shade_isotropic<MIS>(sc, f, w, pid); }
    }
    const uint32_t slot = block_append<kWide>(w.counters + C_ACTIVE0 + nxt, live);
    if (live) w.q_active[nxt][slot] = pid;
    retire_block<kWide>(w, valid && !live, pid < uint32_t(w.B) ? pid : pid - uint32_t(w.B), dn);
}

// counters of an iteration's outputs back to zero (ACTIVE[cur] keeps the survivors of the last one)
__global__ void k_reset_counters(uint32_t* __restrict__ c, int nxt, int dn, int ex) {
    const int t = threadIdx.x;
    if (t >= C_MAT0 && t <= C_EXACT_SHADOW) c[t] = 0;
    if (t == 0) { c[C_ACTIVE0 + nxt] = 0; c[C_MAT4] = 0; c[C_DONE0 + dn] = 0; c[C_EXACT0 + ex] = 0; }
}

// ------------------------------------------------------------------ vertex-count reduction over the done list
__global__ void __launch_bounds__(256) k_counts(WF w, int dn) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    uint32_t nc = 0, nl = 0;
    if (i < w.counters[C_DONE0 + dn]) { const uint32_t b = w.q_slots[dn][i]; nc = w.ncam[b]; nl = w.nlight[b]; }
    const uint32_t mc = __reduce_max_sync(0xFFFFFFFFu, nc), ml = __reduce_max_sync(0xFFFFFFFFu, nl);
    const uint32_t sc_ = __reduce_add_sync(0xFFFFFFFFu, nc), sl = __reduce_add_sync(0xFFFFFFFFu, nl);
    const uint32_t pairs = __reduce_add_sync(0xFFFFFFFFu, nc * nl + (w.mis ? nl : 0u));   // <= 32 * (depth^2 + depth); MIS: + the splats
    if ((threadIdx.x & 31) == 0) {
        if (pairs) atomicAdd(reinterpret_cast<unsigned long long*>(w.counters + C_PAIRS_LO), (unsigned long long)pairs);
        if (mc) atomicMax(w.counters + C_MAXNC, mc);
        if (ml) atomicMax(w.counters + C_MAXNL, ml);
        if (sc_) atomicAdd(w.counters + C_SUMNC, sc_);
        if (sl) atomicAdd(w.counters + C_SUMNL, sl);
    }
}

// ------------------------------------------------------------------ K3
// connect_vertices (camera.h:440-475) for camera-vertex slots [k0, k1) of every finished sample (the done
// list) against all of the sample's light vertices.  Everything except visible() is evaluated here; pairs that survive the
// early-outs become shadow rays carrying their finished contribution.
template <bool MEDIA>
__global__ void __launch_bounds__(kBlock) k_connect(SceneView sc, Frame f, WF w, SQ q, const uint32_t* __restrict__ done, int64_t nb,
                                                   int k0, int k1) {
    const int64_t idx = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int64_t total = int64_t(k1 - k0) * nb;
    const bool valid = idx < total;
    const int k = valid ? k0 + int(idx / nb) : 0;
    const int64_t b = valid ? int64_t(done[idx % nb]) : 0;
    uint32_t nl = 0;
    D3 cp = mk(0, 0, 0), ctf = mk(0, 0, 0), cn = mk(0, 0, 0);
    uint32_t pix = 0, cindex = 0;
    uint64_t sid = 0;
    if (valid && uint32_t(k) < w.ncam[b]) {
        nl = w.nlight[b];
        uint32_t cprim;
        load_vertex(w.cv + size_t(b) * size_t(w.depth) + size_t(k), cp, ctf, cprim, &cindex);
        if (MEDIA && (cprim & kMediumPrim)) cn = mk(1.0, 0.0, 0.0);   // a scatter point inside a medium: constant_medium.h:52
        else if (MEDIA && (cprim & kSpherePrim)) cn = sphere_normal(sc.spheres[cprim & ~kSpherePrim], cp);
        else {
            const TriShade sh = sc.shade[cprim];
            cn = mk(sh.nx, sh.ny, sh.nz);   // |dot| below makes the face flip irrelevant (camera.h:451-452)
        }
        const SampleId sm = decode_sample(f, w.sample[b]);
        pix = sm.pix; sid = sm.sid;
    }
    const uint32_t nl_max = __reduce_max_sync(0xFFFFFFFFu, nl);
    for (uint32_t m = 0; m < nl_max; ++m) {
        bool emit = false;
        D3 so = mk(0, 0, 0), sd = mk(0, 0, 0), contrib = mk(0, 0, 0);
        double max_t = 0;
        uint32_t far_prim = 0;
        unsigned long long mkey = 0;
        if (m < nl) {
            D3 lp, lT;
            uint32_t lprim, lindex;
            load_vertex(w.lv + size_t(b) * size_t(w.depth) + size_t(m), lp, lT, lprim, &lindex);
            const D3 dir = lp - cp;
            const double dist2 = len2(dir);
            if (dist2 > 0) {
                const D3 du = unit(dir);
                far_prim = lprim & ~kLightFlag;
                const bool l_medium = MEDIA && (far_prim & kMediumPrim) != 0;
                TriShade lsh;
                lsh.nx = 1.0; lsh.ny = 0.0; lsh.nz = 0.0; lsh.mat = 0; lsh.pad = 0;
                const bool l_sphere = MEDIA && (far_prim & kSpherePrim) != 0;
                if (l_sphere) {
                    const SphereDev sp = sc.spheres[far_prim & ~kSpherePrim];
                    const D3 sn = sphere_normal(sp, lp);
                    lsh.nx = sn.x; lsh.ny = sn.y; lsh.nz = sn.z; lsh.mat = sp.mat;
                } else if (!l_medium) lsh = sc.shade[far_prim];
                const D3 ln = mk(lsh.nx, lsh.ny, lsh.nz);
                const double cos_cam = fabs(dot(cn, du));
                const double cos_light = fabs(dot(ln, -du));
                if (cos_cam > 0 && cos_light > 0) {
                    // visible() prologue, camera.h:425-436 (its own dist, from sqrt of the same length_squared)
                    const double dist = len(dir);
                    const D3 vdu = dir / dist;
                    max_t = dist - 0.001;
                    if (dist > 0 && max_t > 0) {
                        D3 lrgb;
                        if (w.lcol) lrgb = w.lcol[size_t(b) * size_t(w.depth) + size_t(m)];
                        else if (l_medium) { const MediumDev md = sc.media[far_prim & ~kMediumPrim]; lrgb = mk(md.r, md.g, md.b); }
                        else { const Mat lm = sc.mats[lsh.mat]; lrgb = mk(lm.r, lm.g, lm.b); }
                        const D3 f_light = (lprim & kLightFlag) ? lrgb : (l_medium ? lrgb / (4 * kPi) : lrgb / kPi);   // camera.h:463-466; material.h:166-168
                        if (len2(f_light) > 0) {
                            contrib = ctf * lT * f_light;                               // camera.h:471-472
                            contrib = contrib * (cos_cam * cos_light) / dist2;          // camera.h:473
                            so = cp + (0.001 * vdu);
                            sd = vdu;
                            emit = true;
                            if (l_medium || l_sphere) far_prim = BDPT_MISS;   // no far-endpoint triangle to test first
                            // the far vertex's own triangle settles ~80 % of the segments (k_shadow_pre does the same for the
                            // survivors' queue): asked here first, before the side lists' loops over boundary triangles
                            if (MEDIA && far_prim != BDPT_MISS) {
                                RayD sr;
                                sr.o = so; sr.d = sd; sr.tmin = 0.001; sr.tmax = max_t;
                                if (far_endpoint_occludes(sc, sr, far_prim, nullptr)) emit = false;
                            }
                            // the sphere prims of the side list occlude like any prim (oracle/port.cpp: spheres_occlude)
                            for (uint32_t sk = 0; MEDIA && sk < sc.n_spheres && emit; ++sk) {
                                double ts;
                                if (sphere_hit_t(sc.spheres[sk], so, sd, 0.001, max_t, ts)) emit = false;
                            }
                            // media between the two vertices scatter the shadow ray (world.hit inside visible(), camera.h:437): one draw
                            // of the sample's shadow stream per (pair, medium) -- oracle/port.cpp: visible_media.  Asked by
                            // k_media_segments over the queued segments, where every lane has one (here ~3 of 32 lanes would).
                            if (MEDIA && q.mkey) mkey = ((unsigned long long)sid << 20) | (unsigned long long)(cindex * 128u + lindex);
                        }
                    }
                }
            }
        }
        const uint32_t slot = warp_append(q.cnt + SQ_N, emit);
        if (emit) {
            if (slot < uint32_t(w.scap)) {
                q.ox[slot] = so.x; q.oy[slot] = so.y; q.oz[slot] = so.z;
                q.dx[slot] = sd.x; q.dy[slot] = sd.y; q.dz[slot] = sd.z;
                q.tmax[slot] = max_t;
                q.cx[slot] = contrib.x; q.cy[slot] = contrib.y; q.cz[slot] = contrib.z;
                q.pix[slot] = pix;
                q.far[slot] = far_prim;
                if (MEDIA && q.mkey) q.mkey[slot] = mkey;
            } else {
                q.cnt[SQ_OVERFLOW] = 1;
            }
        }
    }
}

// ------------------------------------------------------------------ K3, BDPT_WEIGHT_MIS
// The same pairing as k_connect, with balance-heuristic weights in O(1) per pair from the vertices' frozen dVCM / dVC
// (oracle/port.cpp sample_mis states the same arithmetic): light vertex 0 is the emitter point (s = 1), the others are
// surface vertices (s >= 2); a lambertian surface reflects only to the side the path arrived on; a full path holds at
// most `depth` surface vertices.  The thread of camera slot 0 also emits the sample's light-tracing SPLATS (t = 1): each
// light-subpath surface vertex, seen from the pinhole, lands in whatever pixel it projects to.  All of it goes through
// the shadow queue (the segment stops short of the far surface: SURVEY F4 is a property of the reference's estimator,
// not of this one) and reaches the film as atomic adds.
template <bool EX>
__device__ __forceinline__ void push_segment(const SceneView& sc, const WF& w, const SQ& q, bool emit, D3 from, D3 to_dir_unit, double dist, D3 contrib, uint32_t pix,
                                             double msum = 0.0) {
    const double max_t = dist - 0.001 - 1e-6 * (1.0 + dist);
    emit = emit && max_t > 0.001;
    if (EX && emit) {   // the sphere prims of the side list occlude like any prim (oracle/port.cpp: visible_mis)
        const D3 so = from + (0.001 * to_dir_unit);
        for (uint32_t sk = 0; sk < sc.n_spheres && emit; ++sk) {
            double ts;
            if (sphere_hit_t(sc.spheres[sk], so, to_dir_unit, 0.001, max_t, ts)) emit = false;
        }
    }
    const uint32_t slot = warp_append(q.cnt + SQ_N, emit);
    if (emit) {
        if (slot < uint32_t(w.scap)) {
            const D3 so = from + (0.001 * to_dir_unit);
            q.ox[slot] = so.x; q.oy[slot] = so.y; q.oz[slot] = so.z;
            q.dx[slot] = to_dir_unit.x; q.dy[slot] = to_dir_unit.y; q.dz[slot] = to_dir_unit.z;
            q.tmax[slot] = max_t;
            q.cx[slot] = contrib.x; q.cy[slot] = contrib.y; q.cz[slot] = contrib.z;
            q.pix[slot] = pix;
            q.far[slot] = BDPT_MISS;   // no far-endpoint shortcut: the far surface is outside the segment by construction
            if (EX && q.msum) q.msum[slot] = msum;
        } else {
            q.cnt[SQ_OVERFLOW] = 1;
        }
    }
}

template <bool EX>
__global__ void __launch_bounds__(kBlock) k_connect_mis(SceneView sc, Frame f, WF w, SQ q, const uint32_t* __restrict__ done, int64_t nb,
                                                       int k0, int k1) {
    const int64_t idx = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int64_t total = int64_t(k1 - k0) * nb;
    const bool valid = idx < total;
    const int k = valid ? k0 + int(idx / nb) : 0;
    const int64_t b = valid ? int64_t(done[idx % nb]) : 0;
    const uint32_t nl = valid ? w.nlight[b] : 0u;
    const bool have_cam = valid && uint32_t(k) < w.ncam[b];
    D3 cp = mk(0, 0, 0), ctf = mk(0, 0, 0), cn = mk(0, 0, 0);
    VertMis cm{0.0, 0.0, 0.0, 0u, 0u};
    uint32_t pix = 0;
    bool c_med = false;
    if (have_cam) {
        uint32_t cprim;
        load_vertex(w.cv + size_t(b) * size_t(w.depth) + size_t(k), cp, ctf, cprim);
        const uint32_t cid = cprim & ~(kBackFlag | kLightFlag);
        c_med = EX && (cid & kMediumPrim) != 0;   // a scatter point inside a medium: no cosine, 1 / (4 pi) both ways
        if (c_med) cn = mk(1.0, 0.0, 0.0);
        else if (EX && (cid & kSpherePrim)) cn = sphere_normal(sc.spheres[cid & ~kSpherePrim], cp);
        else { const TriShade sh = sc.shade[cid]; cn = mk(sh.nx, sh.ny, sh.nz); }
        if (cprim & kBackFlag) cn = -cn;
        cm = w.cvm[size_t(b) * size_t(w.depth) + size_t(k)];
        pix = decode_sample(f, w.sample[b]).pix;
    }
    const bool splatter = valid && k == 0;   // one thread per sample also walks the light vertices for the t = 1 strategy
    const uint32_t nl_max = __reduce_max_sync(0xFFFFFFFFu, (have_cam || splatter) ? nl : 0u);
    const double pdfA = 1.0 / sc.light_total_area;
    for (uint32_t m = 0; m < nl_max; ++m) {
        bool emit = false, emit_splat = false;
        D3 dirc = mk(0, 0, 0), dirs = mk(0, 0, 0), contrib = mk(0, 0, 0), contrib_s = mk(0, 0, 0), lp = mk(0, 0, 0);
        double distc = 0, dists = 0;
        // media: the segment's transmittance Tr is found by k_media_segments over the queue (all lanes busy there); every weight is
        // 1 / (S Tr + 1) and every contribution C Tr / (S Tr + 1) with S and C free of Tr -- S travels in q.msum, C as the contribution
        double msum = 0, msum_s = 0;
        uint32_t spix = 0;
        if (m < nl && (have_cam || splatter)) {
            D3 lT;
            uint32_t lprim;
            load_vertex(w.lv + size_t(b) * size_t(w.depth) + size_t(m), lp, lT, lprim);
            const uint32_t lid = lprim & ~(kBackFlag | kLightFlag);
            const bool l_med = EX && (lid & kMediumPrim) != 0;
            D3 ln = mk(1.0, 0.0, 0.0);
            uint32_t lmat = 0;
            if (l_med) {}
            else if (EX && (lid & kSpherePrim)) { const SphereDev sp = sc.spheres[lid & ~kSpherePrim]; ln = sphere_normal(sp, lp); lmat = sp.mat; }
            else { const TriShade lsh = sc.shade[lid]; ln = mk(lsh.nx, lsh.ny, lsh.nz); lmat = lsh.mat; }
            if (lprim & kBackFlag) ln = -ln;
            D3 lrgb;
            if (w.lcol) lrgb = w.lcol[size_t(b) * size_t(w.depth) + size_t(m)];
            else if (l_med) { const MediumDev md = sc.media[lid & ~kMediumPrim]; lrgb = mk(md.r, md.g, md.b); }
            else { const Mat lm = sc.mats[lmat]; lrgb = mk(lm.r, lm.g, lm.b); }
            const D3 lf = l_med ? lrgb / (4 * kPi) : lrgb / kPi;                       // BSDF value of the light vertex
            const bool media = EX && sc.n_media != 0;
            if (have_cam) {
                const D3 dir = lp - cp;
                const double dist2 = len2(dir);
                if (dist2 > 0) {
                    const D3 du = unit(dir);
                    const double cosCamera = c_med ? 1.0 : dot(cn, du), cosLight = l_med ? 1.0 : dot(ln, -du);
                    if (cosCamera > 0 && cosLight > 0) {
                        dirc = du; distc = len(dir);
                        const double tr = 1.0;
                        const double c_dir = c_med ? 1 / (4 * kPi) : cosCamera / kPi;       // pdf of the camera vertex scattering this way
                        const double c_rev = c_med ? 1 / (4 * kPi) : cm.cos_in / kPi;       // ... and back towards its predecessor
                        if (lprim & kLightFlag) {   // s = 1: the emitter point (ctf = T * f of the camera vertex)
                            if (cm.len + 1u <= uint32_t(f.depth)) {
                                const double directPdfW = pdfA * dist2 / cosLight, emissionPdfW = pdfA * cosLight / kPi;
                                const double wLight = (c_dir * tr) / directPdfW;
                                const double wCamera = (emissionPdfW * cosCamera * tr / (directPdfW * cosLight)) * (cm.dVCM + cm.dVC * c_rev);
                                const double wgt = media ? 1.0 : 1.0 / (wLight + 1.0 + wCamera);
                                msum = wLight + wCamera;
                                contrib = (wgt * cosCamera * tr / directPdfW) * (ctf * lrgb);
                                emit = true;
                            }
                        } else {
                            const VertMis vm = w.lvm[size_t(b) * size_t(w.depth) + size_t(m)];
                            if (cm.len + vm.len <= uint32_t(f.depth)) {
                                const double l_dir = l_med ? 1 / (4 * kPi) : cosLight / kPi, l_rev = l_med ? 1 / (4 * kPi) : vm.cos_in / kPi;
                                const double camDirPdfA = c_dir * cosLight * tr / dist2, lightDirPdfA = l_dir * cosCamera * tr / dist2;
                                const double wLight = camDirPdfA * (vm.dVCM + vm.dVC * l_rev);
                                const double wCamera = lightDirPdfA * (cm.dVCM + cm.dVC * c_rev);
                                const double wgt = media ? 1.0 : 1.0 / (wLight + 1.0 + wCamera);
                                msum = wLight + wCamera;
                                contrib = (wgt * cosLight * cosCamera * tr / dist2) * (ctf * lT * lf);
                                emit = true;
                            }
                        }
                    }
                }
            }
            if (splatter && !(lprim & kLightFlag)) {   // t = 1
                const D3 dir = f.center - lp;
                const double dist2 = len2(dir);
                if (dist2 > 0) {
                    const D3 du = unit(dir);
                    const double cosToCamera = l_med ? 1.0 : dot(ln, du), cosAtCamera = dot(f.fwd, -du);
                    if (cosToCamera > 0 && cosAtCamera > 0) {
                        const D3 P = f.center + (f.focus / cosAtCamera) * (-du);
                        const D3 rel = P - f.ul;
                        const double xc = dot(rel, f.du) / f.du2, yc = dot(rel, f.dv) / f.dv2;
                        if (xc >= 0 && yc >= 0 && xc < f.W && yc < f.H) {
                            const VertMis vm = w.lvm[size_t(b) * size_t(w.depth) + size_t(m)];
                            const double ipcd = f.plane_px / cosAtCamera;
                            dirs = du; dists = len(dir);
                            const double tr = 1.0;
                            const double imageToSurface = (ipcd * ipcd / cosAtCamera) * cosToCamera * tr / dist2;
                            const double wLight = (imageToSurface / f.n_light) * (vm.dVCM + vm.dVC * (l_med ? 1 / (4 * kPi) : vm.cos_in / kPi));
                            const double wgt = media ? 1.0 : 1.0 / (wLight + 1.0);
                            msum_s = wLight;
                            contrib_s = (wgt * imageToSurface / f.n_light) * (lT * lf);
                            spix = uint32_t(int(yc)) * uint32_t(f.W) + uint32_t(int(xc));
                            emit_splat = true;
                        }
                    }
                }
            }
        }
        push_segment<EX>(sc, w, q, emit, cp, dirc, distc, contrib, pix, msum);
        push_segment<EX>(sc, w, q, emit_splat, lp, dirs, dists, contrib_s, spix, msum_s);
    }
}

// ------------------------------------------------------------------ K3b, participating media only
// The media's share of every queued segment, one segment per thread (in the connect kernels a handful of lanes per warp hold a
// segment at any one time: ncu showed 3 of 32 busy in the boundary loops).  REF_SUM: a medium between the two vertices scatters the
// shadow ray with the probability constant_medium::hit gives it, one draw of the sample's shadow stream per (pair, medium)
// (oracle/port.cpp: visible_media); a scattered segment is marked dead.  MIS: the analytic transmittance finishes weight and
// contribution (oracle/port.cpp: tr_mis).
template <bool MIS>
__global__ void __launch_bounds__(kBlock) k_media_segments(SceneView sc, Frame f, WF w, SQ q) {
    const uint32_t idx = blockIdx.x * kBlock + threadIdx.x;
    if (idx >= q.cnt[SQ_N] || idx >= uint32_t(w.scap)) return;
    const D3 so = mk(q.ox[idx], q.oy[idx], q.oz[idx]), sd = mk(q.dx[idx], q.dy[idx], q.dz[idx]);
    const double max_t = q.tmax[idx];
    if (MIS) {
        const double tr = media_transmittance(sc.media, sc.med_tris, sc.n_media, so, sd, 0.001, max_t);
        const double wt = tr / (q.msum[idx] * tr + 1.0);
        q.cx[idx] *= wt; q.cy[idx] *= wt; q.cz[idx] *= wt;
    } else {
        const unsigned long long key = q.mkey[idx];
        const uint64_t sid = uint64_t(key >> 20);
        const uint32_t pair = uint32_t(key & 0xFFFFFu);
        for (uint32_t k = 0; k < sc.n_media; ++k) {
            Rng srng;
            srng.init(f.seed, 2u, sid, pair * 16u + k);
            double tm;
            if (medium_hit(sc.med_tris, sc.media[k], so, sd, 0.001, max_t, srng, tm)) { q.far[idx] = kDeadSeg; return; }
        }
    }
}

// ------------------------------------------------------------------ K4
// visible() (camera.h:437: world.hit over [0.001, dist-0.001], inclusive ends — SURVEY F4) and the
// film accumulation of every unoccluded connection.
__device__ __forceinline__ RayD load_shadow_ray(const SQ& q, uint32_t idx) {
    RayD r;
    r.o = mk(q.ox[idx], q.oy[idx], q.oz[idx]);
    r.d = mk(q.dx[idx], q.dy[idx], q.dz[idx]);
    r.tmin = 0.001;
    r.tmax = q.tmax[idx];
    return r;
}

// Stage 1 (fast flavour): the far-endpoint test settles ~80 % of the segments with one triangle
// and one box; the rest are compacted into q_shadow_live so that stage 2's warps are full of rays that
// really walk the tree.
template <bool STATS>
__global__ void __launch_bounds__(kBlock) k_shadow_pre(SceneView sc, WF w, SQ q, unsigned long long* stats) {
    const uint32_t idx = blockIdx.x * kBlock + threadIdx.x;
    if (idx == 0) atomicAdd(reinterpret_cast<unsigned long long*>(w.counters + C_SHADOW_TOTAL), (unsigned long long)q.cnt[SQ_N]);
    bool live = false;
    if (idx < q.cnt[SQ_N]) {
        const RayD r = load_shadow_ray(q, idx);
        TravStats ts{0, 0};
        const uint32_t far = q.far[idx];   // BDPT_MISS (MIS segments): there is no far surface inside the segment to test first
        live = far != kDeadSeg && (far == BDPT_MISS || !far_endpoint_occludes(sc, r, far, STATS ? &ts : nullptr));
        if (STATS) add_trav<1, false>(stats, ts);
    }
    const uint32_t slot = warp_append(q.cnt + SQ_LIVE, live);
    if (live) q.live[slot] = idx;
}

// Stage 2.  Blocks [0, n_slow_blocks) finish the segments that an EARLIER launch gave up on — over the work
// limit (re-run unbounded) or in need of the reference tree (ties, knife edges) — from their own copies of
// the payload; the other blocks walk this round's live segments under the work limit and park what they
// cannot settle in ss[sx] for the next launch.
template <int BVH, bool STATS>
__global__ void __launch_bounds__(kBlock, kTravBlocksPerSM) k_shadow(SceneView sc, WF w, SQ q, int sx, uint32_t n_slow_blocks, uint32_t limit, double* film,
                                                  unsigned long long* stats) {
    BDPT_DECLARE_STACK;
    if (BVH == BDPT_BVH_REFERENCE) {
        const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
        if (i == 0) atomicAdd(reinterpret_cast<unsigned long long*>(w.counters + C_SHADOW_TOTAL), (unsigned long long)q.cnt[SQ_N]);
        if (i < q.cnt[SQ_N] && q.far[i] != kDeadSeg) {
            const RayD r = load_shadow_ray(q, i);
            HitD h;
            TravStats ts{0, 0};
            const bool found = traverse_ref<true, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts);
            if (STATS) add_trav<1, true>(stats, ts);
            if (!found) film_add(film, q.pix[i], mk(q.cx[i], q.cy[i], q.cz[i]));
        }
        return;
    }
    if (blockIdx.x < n_slow_blocks) {   // one parked segment per warp at a time, warp-stride over the device-side count
        const uint32_t n = min(w.counters[C_SLOW_S0 + (sx ^ 1)], uint32_t(w.sscap));
        if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(&stats[ST_DEFER_SHADOW], (unsigned long long)n);
        if ((threadIdx.x & 31) != 0) return;
        for (uint32_t i = blockIdx.x * kFrontPerBlock + (threadIdx.x >> 5); i < n; i += n_slow_blocks * kFrontPerBlock) {
            const SlowShadow* e = w.ss[sx ^ 1] + i;
            const double2* pq = reinterpret_cast<const double2*>(e);
            const double2 q0 = pq[0], q1 = pq[1], q2 = pq[2], q3 = pq[3], q4 = pq[4];
            const int4 q5 = reinterpret_cast<const int4*>(e)[5];
            RayD r;
            r.o = mk(q0.x, q0.y, q1.x); r.d = mk(q1.y, q2.x, q2.y); r.tmin = 0.001; r.tmax = q3.x;
            HitD h;
            TravStats ts{0, 0};
            bool found;
            if (q5.y) found = traverse_ref<true, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts);
            else found = traverse_fast<true, STATS, false>(sc, r, h, l_stack, 1, &ts) == TR_HIT;
            if (STATS) add_trav<1, false>(stats, ts);
            if (!found) film_add(film, uint32_t(q5.x), mk(q3.y, q4.x, q4.y));
        }
        return;
    }
    const uint32_t i = (blockIdx.x - n_slow_blocks) * kBlock + threadIdx.x;
    bool park = false;
    int kind = 0;
    uint32_t idx = 0;
    RayD r;
    if (i < q.cnt[SQ_LIVE]) {
        idx = q.live[i];
        r = load_shadow_ray(q, idx);
        HitD h;
        TravStats ts{0, 0};
        const int rc = traverse_fast<true, STATS, true>(sc, r, h, l_stack, 1, &ts, limit);
        park = rc == TR_DEFER || rc == TR_SLOW;
        kind = rc == TR_DEFER ? 1 : 0;
        if (STATS) add_trav<1, false>(stats, ts);
        if (rc == TR_MISS) film_add(film, q.pix[idx], mk(q.cx[idx], q.cy[idx], q.cz[idx]));
    }
    const uint32_t slot = warp_append(w.counters + C_SLOW_S0 + sx, park);
    if (park) {
        const D3 c = mk(q.cx[idx], q.cy[idx], q.cz[idx]);
        const uint32_t pix = q.pix[idx];
        if (slot < uint32_t(w.sscap)) {
            SlowShadow* e = w.ss[sx] + slot;
            double2* pq = reinterpret_cast<double2*>(e);
            pq[0] = make_double2(r.o.x, r.o.y); pq[1] = make_double2(r.o.z, r.d.x); pq[2] = make_double2(r.d.y, r.d.z);
            pq[3] = make_double2(r.tmax, c.x); pq[4] = make_double2(c.y, c.z);
            reinterpret_cast<int4*>(e)[5] = make_int4(int(pix), kind, 0, 0);
        } else {   // parking lot full: settle it here, unbounded
            HitD h;
            TravStats ts{0, 0};
            const bool found = traverse_fast<true, STATS, false>(sc, r, h, l_stack, 1, &ts) == TR_HIT;
            if (STATS) add_trav<1, false>(stats, ts);
            if (!found) film_add(film, pix, c);
        }
    }
}

// ------------------------------------------------------------------ K4 stage 2, fast flavour: persistent warps with dynamic fetch
// Same structure as k_trace_fast over this round's live segments (q.live).  A segment the fp32 walk cannot
// settle (tie / knife edge / fp64 slabs needed) is parked in ss[sx] with its payload and finished by the
// front blocks of a later launch through the reference traversal.
template <bool STATS>
__global__ void __launch_bounds__(kBlock, kTravBlocksPerSM) k_shadow_fast(SceneView sc, WF w, SQ q, int sx, uint32_t n_slow_blocks,
                                                                         int threshold, uint32_t limit, double* film,
                                                                         unsigned long long* stats) {
    BDPT_DECLARE_STACK;
    if (blockIdx.x < n_slow_blocks) {   // one parked segment per warp at a time, warp-stride over the device-side count
        const uint32_t n = min(w.counters[C_SLOW_S0 + (sx ^ 1)], uint32_t(w.sscap));
        if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(&stats[ST_DEFER_SHADOW], (unsigned long long)n);
        const bool lane0 = (threadIdx.x & 31) == 0;
        for (uint32_t i = blockIdx.x * kFrontPerBlock + (threadIdx.x >> 5); i < n; i += n_slow_blocks * kFrontPerBlock) {
            const SlowShadow* e = w.ss[sx ^ 1] + i;
            const double2* pq = reinterpret_cast<const double2*>(e);
            const double2 q0 = pq[0], q1 = pq[1], q2 = pq[2], q3 = pq[3], q4 = pq[4];
            const int4 q5 = reinterpret_cast<const int4*>(e)[5];
            RayD r;
            r.o = mk(q0.x, q0.y, q1.x); r.d = mk(q1.y, q2.x, q2.y); r.tmin = 0.001; r.tmax = q3.x;
            HitD h;
            TravStats ts{0, 0};
            int rc = TR_DEFER;   // kind 1: the reference traversal; kind 0 (over the work limit): split over the warp
            if (q5.y == 0) rc = split_walk<true, STATS>(sc, r, h, l_stack, ts);
            if (!lane0) continue;
            const bool found = rc == TR_DEFER ? traverse_ref<true, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts) : rc == TR_HIT;
            if (STATS) add_trav<1, false>(stats, ts);
            if (!found) film_add(film, uint32_t(q5.x), mk(q3.y, q4.x, q4.y));
        }
        return;
    }
    const uint32_t n = q.cnt[SQ_LIVE];
    {
        // Few segments survive k_shadow_pre, and their walks differ by two orders of magnitude in length: if every
        // resident warp took 32 of them at once, nothing would be left to refill the lanes that finish early and
        // each warp would idle behind its longest ray.  Only as many warps stay as leave `over` rounds of refills.
        const uint32_t over = (uint32_t(threshold) >> 16) & 0xFFu;
        const uint32_t warp_id = ((blockIdx.x - n_slow_blocks) * kBlock + threadIdx.x) >> 5;
        if (over > 1u && uint64_t(warp_id) * 32u * over >= n) return;
    }
    Walk<true, STATS> W;
    bool have = false, exhausted = false;
    int result = WALK_YIELD;
    uint32_t idx = 0;
    for (;;) {
        __syncwarp();
        const bool fin = result != WALK_YIELD;
        if (__any_sync(0xFFFFFFFFu, fin)) {
            const bool park = fin && (result == TR_DEFER || result == TR_SLOW);
            if (fin && STATS && !park) add_trav<1, false>(stats, W.ts);
            if (fin && result == TR_MISS) film_add(film, q.pix[idx], mk(q.cx[idx], q.cy[idx], q.cz[idx]));
            const uint32_t slot = warp_append(w.counters + C_SLOW_S0 + sx, park);
            if (park) {
                const RayD r = load_shadow_ray(q, idx);
                const D3 c = mk(q.cx[idx], q.cy[idx], q.cz[idx]);
                const uint32_t pix = q.pix[idx];
                if (slot < uint32_t(w.sscap)) {
                    SlowShadow* e = w.ss[sx] + slot;
                    double2* pq = reinterpret_cast<double2*>(e);
                    pq[0] = make_double2(r.o.x, r.o.y); pq[1] = make_double2(r.o.z, r.d.x); pq[2] = make_double2(r.d.y, r.d.z);
                    pq[3] = make_double2(r.tmax, c.x); pq[4] = make_double2(c.y, c.z);
                    reinterpret_cast<int4*>(e)[5] = make_int4(int(pix), result == TR_DEFER ? 1 : 0, 0, 0);
                } else {   // parking lot full: settle it here
                    HitD h;
                    const bool found = traverse_ref<true, false>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, nullptr);
                    if (!found) film_add(film, pix, c);
                }
            }
            result = WALK_YIELD;
        }
        const uint32_t my = warp_fetch(q.cnt + SQ_CURSOR, !have, n, exhausted, (threshold >> 8) & 0xFF);
        if (my != 0xFFFFFFFFu) {
            idx = q.live[my];
            const RayD r = load_shadow_ray(q, idx);
            if (walk_eligible(sc, r)) { W.init(sc, r); have = true; }
            else result = TR_DEFER;
        }
        if (!__any_sync(0xFFFFFFFFu, have || result != WALK_YIELD)) break;
#if BDPT_WIDE_WALK
        {
            const int rc = W.run_wide(sc, l_stack, have, threshold & 0xFF, (threshold >> 16) & 0xFF, (threshold >> 24) & 0xFF,
                                      exhausted ? 0 : ((threshold >> 8) & 0xFF), limit);
            if (have && rc != WALK_YIELD) { result = rc; have = false; }
        }
#else
        if (have) {
            const int rc = W.run(sc, l_stack, exhausted ? 0 : (threshold & 0xFF), limit);
            if (rc != WALK_YIELD) { result = rc; have = false; }
        }
#endif
    }
}


// ------------------------------------------------------------------ K4 stage 2, round 2: the shadow walk on traverse_w32.cuh
// Any-hit: a triangle the fp32 filter finds certainly inside the segment ends the walk at once, without any fp64
// work; only what the filter cannot decide (a graze, a hit within ~1e-5 of an end of the segment) reaches the
// exact test.  The far endpoint's own triangle, which k_shadow_pre has already cleared, is skipped by ID.
template <bool STATS>
__global__ void __launch_bounds__(kBlock, kW32BlocksPerSM) k_shadow_w32(SceneView sc, WF w, SQ q, int sx, uint32_t n_slow_blocks, int knobs,
                                                                        double* film, unsigned long long* stats) {
    BDPT_DECLARE_STACK;
    if (blockIdx.x < n_slow_blocks) {   // parked segments (they need the reference tree): one per warp at a time
        const uint32_t n = min(w.counters[C_SLOW_S0 + (sx ^ 1)], uint32_t(w.sscap));
        if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(&stats[ST_DEFER_SHADOW], (unsigned long long)n);
        if ((threadIdx.x & 31) != 0) return;
        for (uint32_t i = blockIdx.x * kFrontPerBlock + (threadIdx.x >> 5); i < n; i += n_slow_blocks * kFrontPerBlock) {
            const SlowShadow* e = w.ss[sx ^ 1] + i;
            const double2* pq = reinterpret_cast<const double2*>(e);
            const double2 q0 = pq[0], q1 = pq[1], q2 = pq[2], q3 = pq[3], q4 = pq[4];
            const int4 q5 = reinterpret_cast<const int4*>(e)[5];
            RayD r;
            r.o = mk(q0.x, q0.y, q1.x); r.d = mk(q1.y, q2.x, q2.y); r.tmin = 0.001; r.tmax = q3.x;
            HitD h;
            TravStats ts{0, 0};
            const bool found = traverse_ref<true, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &ts);
            if (STATS) add_trav<1, true>(stats, ts);
            if (!found) film_add(film, uint32_t(q5.x), mk(q3.y, q4.x, q4.y));
        }
        return;
    }
    const uint32_t n = q.cnt[SQ_LIVE];
    const W32Knobs kn = unpack_knobs(knobs);
    Walk32<true, STATS> W;
    W.cur = kWalkDry;
    bool have = false, exhausted = false;
    int result = WALK_YIELD;
    uint32_t idx = 0;
    for (;;) {
        __syncwarp();
        if (result == WALK32_FLUSH) {
            HitD h;
            W.decide(sc, load_shadow_ray(q, idx), h, false);
            result = WALK_YIELD;
        }
        const bool fin = result != WALK_YIELD;
        if (__any_sync(0xFFFFFFFFu, fin)) {
            bool park = false;
            if (fin) {
                int rc = result;
                HitD h;
                if (rc == TR_HIT) rc = W.occluded ? TR_HIT : W.decide(sc, load_shadow_ray(q, idx), h, true);
                park = rc == TR_DEFER;
                if (STATS && !park) add_trav<1, false>(stats, W.ts);
                if (rc == TR_MISS) film_add(film, q.pix[idx], mk(q.cx[idx], q.cy[idx], q.cz[idx]));
                have = false;
                result = WALK_YIELD;
            }
            const uint32_t slot = warp_append(w.counters + C_SLOW_S0 + sx, park);
            if (park) {
                const RayD r = load_shadow_ray(q, idx);
                const D3 c = mk(q.cx[idx], q.cy[idx], q.cz[idx]);
                const uint32_t pix = q.pix[idx];
                if (slot < uint32_t(w.sscap)) {
                    SlowShadow* e = w.ss[sx] + slot;
                    double2* pq = reinterpret_cast<double2*>(e);
                    pq[0] = make_double2(r.o.x, r.o.y); pq[1] = make_double2(r.o.z, r.d.x); pq[2] = make_double2(r.d.y, r.d.z);
                    pq[3] = make_double2(r.tmax, c.x); pq[4] = make_double2(c.y, c.z);
                    reinterpret_cast<int4*>(e)[5] = make_int4(int(pix), 1, 0, 0);
                } else {   // parking lot full: settle it here
                    HitD h;
                    const bool found = traverse_ref<true, false>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, nullptr);
                    if (!found) film_add(film, pix, c);
                }
            }
        }
        const uint32_t my = warp_fetch(q.cnt + SQ_CURSOR, !have, n, exhausted, kn.min_idle);
        if (my != 0xFFFFFFFFu) {
            idx = q.live[my];
            const RayD r = load_shadow_ray(q, idx);
            if (walk_eligible(sc, r)) { W.init(sc, r, q.far[idx]); have = true; }
            else result = TR_DEFER;
        }
        if (!__any_sync(0xFFFFFFFFu, have || result != WALK_YIELD)) break;
        const int rc = W.run(sc, l_stack, have, kn.trips, kn.nstep, kn.t_leaf, kn.t_node, exhausted ? 0 : kn.min_idle);
        if (have) result = rc;
    }
}

}  // namespace

// ------------------------------------------------------------------ host side
namespace {

struct Carver {
    char* p; size_t off = 0;
    template <class T> T* take(size_t n) {
        off = (off + 255) & ~size_t(255);
        T* r = p ? reinterpret_cast<T*>(p + off) : nullptr;
        off += n * sizeof(T);
        return r;
    }
};

void carve(WF& w, char* base) {
    Carver c{base};
    const size_t P = size_t(2 * w.B), V = size_t(w.depth) * size_t(w.B), B = size_t(w.B), S = size_t(w.scap);
    w.path = c.take<PathRec>(P);
    w.sample = c.take<int64_t>(B); w.live = c.take<uint32_t>(B);
    for (int k = 0; k < 2; ++k) w.q_slots[k] = c.take<uint32_t>(B);
    for (int k = 0; k < 2; ++k) w.q_active[k] = c.take<uint32_t>(P);
    for (int k = 0; k < 5; ++k) w.q_mat[k] = c.take<uint32_t>(P);
    for (int k = 0; k < 2; ++k) w.q_exact[k] = c.take<uint32_t>(P);
    for (int k = 0; k < 2; ++k) w.ss[k] = c.take<SlowShadow>(size_t(w.sscap));
    w.counters = c.take<uint32_t>(C_COUNT);
    w.cv = c.take<VertRec>(V); w.lv = c.take<VertRec>(V);
    w.ncam = c.take<uint32_t>(B); w.nlight = c.take<uint32_t>(B);
    if (w.textured) w.lcol = c.take<D3>(V);
    if (w.mis) { w.pmis = c.take<PathMis>(P); w.cvm = c.take<VertMis>(V); w.lvm = c.take<VertMis>(V); }
    for (int k = 0; k < 2; ++k) {
        SQ& q = w.sq[k];
        double** sq[] = {&q.ox, &q.oy, &q.oz, &q.dx, &q.dy, &q.dz, &q.tmax, &q.cx, &q.cy, &q.cz};
        for (double** d : sq) *d = c.take<double>(S);
        q.pix = c.take<uint32_t>(S); q.far = c.take<uint32_t>(S); q.live = c.take<uint32_t>(S);
        q.cnt = c.take<uint32_t>(SQ_COUNT);
        q.mkey = (w.media && !w.mis) ? c.take<unsigned long long>(S) : nullptr;
        q.msum = (w.media && w.mis) ? c.take<double>(S) : nullptr;
    }
    w.bytes = c.off + 256;
}

// camera::initialize, camera.h:160-197 (host fp64, no FMA contraction: -ffp-contract=off)
Frame make_frame(const bdpt_camera& c, int H) {
    Frame f{};
    f.W = c.image_width; f.H = H;
    f.n = int(std::sqrt(double(c.samples_per_pixel)));
    f.recip_sqrt_spp = 1.0 / f.n;
    f.depth = c.max_depth;
    const D3 lookfrom = mk(c.lookfrom[0], c.lookfrom[1], c.lookfrom[2]), lookat = mk(c.lookat[0], c.lookat[1], c.lookat[2]);
    const D3 vup = mk(c.vup[0], c.vup[1], c.vup[2]);
    f.background = mk(c.background[0], c.background[1], c.background[2]);
    f.center = lookfrom;
    const double theta = c.vfov * 3.1415926535897932385 / 180.0;   // main.h:24-26
    const double h = std::tan(theta / 2);
    const double vh = 2 * h * c.focus_dist;
    const double vw = vh * (double(c.image_width) / H);
    const D3 w = unit(lookfrom - lookat);
    const D3 u = unit(cross(vup, w));
    const D3 v = cross(w, u);
    const D3 vu = vw * u;
    const D3 vv = vh * -v;
    f.du = vu / c.image_width;
    f.dv = vv / H;
    const D3 ul = f.center - (c.focus_dist * w) - vu / 2 - vv / 2;
    f.pixel00 = ul + 0.5 * (f.du + f.dv);
    f.fwd = -w; f.ul = ul;
    f.du2 = len2(f.du); f.dv2 = len2(f.dv);
    f.plane_px = c.focus_dist / len(f.du);
    f.focus = c.focus_dist;
    f.n_light = double(f.W) * double(f.H);
    return f;
}

// Per-kernel-class device time: a CUDA event pair around every launch group on the launching stream,
// read back whenever the stream is known to be idle.  (bench.py's roofline uses the trace/shadow classes.)
struct KernelTimers {
    struct Pair { cudaEvent_t a, b; int cls; };
    std::vector<Pair> pool;
    size_t used = 0;
    int open = -1;
    void begin(int cls, cudaStream_t st) {
        if (used == pool.size()) {
            Pair p{};
            cudaEventCreate(&p.a); cudaEventCreate(&p.b);
            pool.push_back(p);
        }
        pool[used].cls = cls;
        cudaEventRecord(pool[used].a, st);
        open = int(used);
    }
    void end(int, cudaStream_t st) { cudaEventRecord(pool[size_t(open)].b, st); ++used; }
    void collect(double* ms, uint64_t* launches) {
        for (size_t i = 0; i < used; ++i) {
            float t = 0.f;
            if (cudaEventSynchronize(pool[i].b) == cudaSuccess && cudaEventElapsedTime(&t, pool[i].a, pool[i].b) == cudaSuccess) {
                ms[pool[i].cls] += t; launches[pool[i].cls] += 1;
            }
        }
        used = 0;
    }
    ~KernelTimers() { for (auto& p : pool) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); } }
};
KernelTimers* ctx_timers(bdpt_ctx* ctx) {
    if (!ctx->timers) ctx->timers = new KernelTimers();
    return static_cast<KernelTimers*>(ctx->timers);
}

inline unsigned grid_for(int64_t n, int block) { return unsigned(std::max<int64_t>(1, (n + block - 1) / block)); }

}  // namespace

namespace {

// One wavefront in flight: its own stream, workspace, timers and tallies.  bdpt_render keeps several
// (opts.streams) so that the long tail of one wavefront — late bounces with few live rays, whose launch
// time is the latency of the slowest ray — overlaps the bulk of another.
struct Worker {
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;      // the shadow stage (k_shadow_pre, k_shadow) of each iteration
    cudaEvent_t done = nullptr;
    cudaEvent_t ev_connect = nullptr, ev_shadow[2] = {nullptr, nullptr};
    WF wf{};
    KernelTimers tm;
    uint32_t* hc = nullptr;              // pinned read-back of the device counters
    uint64_t rays_closest = 0, rays_shadow = 0, cam_verts = 0, light_verts = 0, launches = 0, samples = 0;
    uint64_t deferred_closest = 0, deferred_shadow = 0;
    double kernel_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint64_t kernel_launches[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::string err;
    int rc = BDPT_OK;
    ~Worker() {
        if (wf.base) cudaFree(wf.base);
        if (hc) cudaFreeHost(hc);
        if (done) cudaEventDestroy(done);
        if (ev_connect) cudaEventDestroy(ev_connect);
        for (auto& e : ev_shadow) if (e) cudaEventDestroy(e);
        if (stream2) cudaStreamDestroy(stream2);
        if (stream) cudaStreamDestroy(stream);
    }
};
struct Workers { std::vector<Worker*> w; ~Workers() { for (auto* x : w) delete x; } };

#define WK_CUDA(call)                                                                   \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) { wk.err = std::string(#call) + ": " + cudaGetErrorString(e__); wk.rc = BDPT_ERR_CUDA; return; } \
    } while (0)

struct Job {
    SceneView sc; Frame f; bdpt_render_opts o;
    int64_t total; double* film; unsigned long long* d_stats;
    size_t smem; unsigned exact_grid; int device; int n_workers;
    int64_t B;                       // sample slots per wavefront (every worker's workspace holds at least this many)
    std::atomic<int64_t>* next;      // next sample of this call not yet handed to a worker
    std::atomic<int64_t>* progress;  // the context's public copy of it (bdpt_render_progress)
    uint32_t limit, limit_shadow;    // work limits of the bulk traversal launches (traverse_walk.cuh); 0 = none
    unsigned resident_blocks;        // blocks of a traversal kernel the device keeps resident (persistent grids)
    int threshold;                   // dynamic fetch: a walk yields when fewer lanes than this are still walking
    bool mis;                        // BDPT_WEIGHT_MIS
    bool tex;                        // the scene has textures
    bool media;                      // the scene has participating media: the MEDIA flavours of trace / shade / connect run
    uint64_t shadow_batch;           // pairs to collect in a shadow queue before its shadow stage is launched (0: every iteration)
    bool walk32;                     // bulk traversal = the warp-synchronous fp32 walk (traverse_w32.cuh); false: round 1's per-lane walk
    int knobs32;                     // its time-slice / refill / leaf-batch parameters (unpack_knobs)
    unsigned resident32;             // blocks of k_trace_w32 / k_shadow_w32 the device keeps resident
};

// One worker: a wavefront of B slots, refilled from the call's shared sample counter until it runs dry.
// Iteration = refill free slots -> trace every live subpath one bounce -> shade -> (host reads the counter
// block: the only sync) -> connect + shadow the samples that finished -> their slots become the free list.
void run_worker(const Job& J, Worker& wk, int /*id*/) {
    cudaSetDevice(J.device);
    const SceneView& sc = J.sc; const Frame& f = J.f; const bdpt_render_opts& o = J.o;
    WF& wf = wk.wf;
    cudaStream_t st = wk.stream;
    KernelTimers& tm = wk.tm;
    uint32_t* hc = wk.hc;
    const bool want_stats = (o.flags & BDPT_RENDER_STATS) != 0;
    const size_t smem = J.smem;
    WK_CUDA(cudaMemsetAsync(wf.counters, 0, C_COUNT * sizeof(uint32_t), st));
    WK_CUDA(cudaEventRecord(wk.ev_connect, st));              // the side stream starts after the counter reset
    WK_CUDA(cudaStreamWaitEvent(wk.stream2, wk.ev_connect, 0));
    int cur = 0, dn = 0, ex = 0;
    int64_t n_active = 0, n_free = J.B;
    uint32_t n_def = 0;              // rays deferred by the previous iteration's trace launch
    int sx = 0;                      // parking lot the next shadow launch fills; it drains ss[sx ^ 1]
    int sb = 0;                      // shadow queue the next connect round writes
    bool first = true;
    cudaStream_t st2 = (o.flags & BDPT_RENDER_SERIAL) ? wk.stream : wk.stream2;
    constexpr unsigned kFrontBlocks = 148;    // 296 warps drain the parked segments, one segment per warp at a time
    // the shadow stage of one connect round, on the side stream: front blocks drain what earlier launches
    // parked, the rest walk this round's live segments
    auto launch_shadow = [&](const SQ& q, uint32_t ns_bound) {
        if (wf.media && ns_bound) {   // the media's share of the queued segments first (REF_SUM: kills, MIS: weighs)
            tm.begin(KT_CONNECT, st2);
            if (J.mis) k_media_segments<true><<<grid_for(ns_bound, kBlock), kBlock, 0, st2>>>(sc, f, wf, q);
            else k_media_segments<false><<<grid_for(ns_bound, kBlock), kBlock, 0, st2>>>(sc, f, wf, q);
            tm.end(KT_CONNECT, st2);
            wk.launches += 1;
        }
        tm.begin(KT_SHADOW, st2);
        if (o.bvh == BDPT_BVH_REFERENCE) {
            if (want_stats) k_shadow<BDPT_BVH_REFERENCE, true><<<grid_for(ns_bound, kBlock), kBlock, smem, st2>>>(sc, wf, q, sx, 0u, 0u, J.film, J.d_stats);
            else k_shadow<BDPT_BVH_REFERENCE, false><<<grid_for(ns_bound, kBlock), kBlock, smem, st2>>>(sc, wf, q, sx, 0u, 0u, J.film, J.d_stats);
        } else {
            cudaMemsetAsync(wf.counters + C_SLOW_S0 + sx, 0, sizeof(uint32_t), st2);
            if (ns_bound) {
                if (want_stats) k_shadow_pre<true><<<grid_for(ns_bound, kBlock), kBlock, 0, st2>>>(sc, wf, q, J.d_stats);
                else k_shadow_pre<false><<<grid_for(ns_bound, kBlock), kBlock, 0, st2>>>(sc, wf, q, J.d_stats);
                wk.launches += 1;
            }
            // persistent: no more blocks than stay resident; the warps pull their segments from q.live
            const unsigned gm = ns_bound ? std::min(grid_for(ns_bound, kBlock), J.walk32 ? J.resident32 : J.resident_blocks) : 0u;
            if (J.walk32) {
                if (want_stats) k_shadow_w32<true><<<kFrontBlocks + gm, kBlock, 0, st2>>>(sc, wf, q, sx, kFrontBlocks, J.knobs32, J.film, J.d_stats);
                else k_shadow_w32<false><<<kFrontBlocks + gm, kBlock, 0, st2>>>(sc, wf, q, sx, kFrontBlocks, J.knobs32, J.film, J.d_stats);
            } else if (want_stats) k_shadow_fast<true><<<kFrontBlocks + gm, kBlock, 0, st2>>>(sc, wf, q, sx, kFrontBlocks, J.threshold, J.limit_shadow, J.film, J.d_stats);
            else k_shadow_fast<false><<<kFrontBlocks + gm, kBlock, 0, st2>>>(sc, wf, q, sx, kFrontBlocks, J.threshold, J.limit_shadow, J.film, J.d_stats);
            sx ^= 1;
        }
        tm.end(KT_SHADOW, st2);
        wk.launches += 1;
    };
    bool acc_open = false;           // the queue wf.sq[sb] holds segments of earlier iterations that no shadow launch has taken yet
    uint64_t acc_bound = 0;          // upper bound of how many
    auto flush_shadow_batch = [&]() {
        if (!acc_open) return;
        acc_open = false;
        const uint32_t ns_bound = uint32_t(std::min<uint64_t>(acc_bound, uint64_t(wf.scap)));
        acc_bound = 0;
        if (!ns_bound) return;
        cudaError_t e = cudaEventRecord(wk.ev_connect, st);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st2, wk.ev_connect, 0);
        if (e == cudaSuccess) { launch_shadow(wf.sq[sb], ns_bound); e = cudaEventRecord(wk.ev_shadow[sb], st2); }
        if (e != cudaSuccess) { wk.err = std::string("shadow batch: ") + cudaGetErrorString(e); wk.rc = BDPT_ERR_CUDA; return; }
        sb ^= 1;
    };
    for (;;) {
        int64_t ngen = 0, gen_base = 0;
        if (n_free > 0) {
            gen_base = J.next->fetch_add(n_free);
            J.progress->store(std::min<int64_t>(J.total, gen_base + n_free), std::memory_order_relaxed);
            ngen = std::max<int64_t>(0, std::min<int64_t>(n_free, J.total - gen_base));
        }
        if (ngen == 0 && n_active == 0 && n_def == 0) break;
        const int nxt = cur ^ 1;
        // counters of this iteration's outputs (ACTIVE[cur] keeps the survivors of the last one)
        k_reset_counters<<<1, 32, 0, st>>>(wf.counters, nxt, dn, ex);   // one launch instead of five memsets
        wk.launches += 1;
        if (ngen > 0) {
            tm.begin(KT_GENERATE, st);
            {
                const uint32_t* fl = first ? nullptr : wf.q_slots[dn ^ 1];
                const unsigned gg = grid_for(ngen, kGen);
                if (J.mis) { if (J.tex) k_generate<true, true><<<gg, kGen, 0, st>>>(sc, f, wf, fl, gen_base, ngen, cur, dn); else k_generate<true, false><<<gg, kGen, 0, st>>>(sc, f, wf, fl, gen_base, ngen, cur, dn); }
                else { if (J.tex) k_generate<false, true><<<gg, kGen, 0, st>>>(sc, f, wf, fl, gen_base, ngen, cur, dn); else k_generate<false, false><<<gg, kGen, 0, st>>>(sc, f, wf, fl, gen_base, ngen, cur, dn); }
            }
            tm.end(KT_GENERATE, st);
            wk.launches += 1;
            wk.samples += uint64_t(ngen);
        }
        first = false;
        const unsigned g = grid_for(n_active + 2 * ngen, kBlock);   // upper bound; the kernels read the device count
        const unsigned gd = n_def ? grid_for(n_def, kFrontPerBlock) : 0u;
        tm.begin(KT_TRACE, st);
#define BDPT_TP(B, S) do { if (J.media) k_trace_paths<B, S, true><<<g + gd, kBlock, smem, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.limit, J.film, J.d_stats); \
                            else k_trace_paths<B, S, false><<<g + gd, kBlock, smem, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.limit, J.film, J.d_stats); } while (0)
        if (o.bvh == BDPT_BVH_REFERENCE) { if (want_stats) BDPT_TP(BDPT_BVH_REFERENCE, true); else BDPT_TP(BDPT_BVH_REFERENCE, false); }
        else {
            const unsigned gp = std::min(g, J.walk32 ? J.resident32 : J.resident_blocks);
            if (J.walk32) {
                if (want_stats) k_trace_w32<true><<<gd + gp, kBlock, 0, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.knobs32, J.film, J.d_stats);
                else k_trace_w32<false><<<gd + gp, kBlock, 0, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.knobs32, J.film, J.d_stats);
            } else if (J.media) {
                if (want_stats) k_trace_fast<true, true><<<gd + gp, kBlock, 0, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.threshold, J.limit, J.film, J.d_stats);
                else k_trace_fast<false, true><<<gd + gp, kBlock, 0, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.threshold, J.limit, J.film, J.d_stats);
            } else if (want_stats) k_trace_fast<true><<<gd + gp, kBlock, 0, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.threshold, J.limit, J.film, J.d_stats);
            else k_trace_fast<false><<<gd + gp, kBlock, 0, st>>>(sc, f, wf, wf.q_active[cur], wf.counters + C_ACTIVE0 + cur, dn, ex, gd, J.threshold, J.limit, J.film, J.d_stats);
        }
#undef BDPT_TP
        tm.end(KT_TRACE, st);
        tm.begin(KT_SHADE, st);
        const unsigned gsh = grid_for(n_active + 2 * ngen + int64_t(n_def), kWide);   // the lagged rays of the front blocks are shaded too
        if (J.mis && J.media) { if (J.tex) k_shade<true, true, true><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); else k_shade<true, false, true><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); }
        else if (J.mis) { if (J.tex) k_shade<true, true><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); else k_shade<true, false><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); }
        else if (J.media) { if (J.tex) k_shade<false, true, true><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); else k_shade<false, false, true><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); }
        else { if (J.tex) k_shade<false, true><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); else k_shade<false, false><<<gsh, kWide, 0, st>>>(sc, f, wf, nxt, dn, J.film); }
        k_counts<<<grid_for(J.B, 256), 256, 0, st>>>(wf, dn);
        tm.end(KT_SHADE, st);
        wk.launches += 3;
        WK_CUDA(cudaMemcpyAsync(hc, wf.counters, C_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        WK_CUDA(cudaStreamSynchronize(st));
        wk.rays_closest += hc[C_ACTIVE0 + cur];
        n_def = hc[C_EXACT0 + ex];
        wk.deferred_closest += n_def;
        ex ^= 1;
        n_active = hc[C_ACTIVE0 + nxt];
        const int64_t n_done = hc[C_DONE0 + dn];
        const int max_nl = int(hc[C_MAXNL]);
        const int max_nc = J.mis ? std::max(1, int(hc[C_MAXNC])) : int(hc[C_MAXNC]);   // MIS: camera slot 0's thread also emits the splats
        wk.cam_verts += hc[C_SUMNC]; wk.light_verts += hc[C_SUMNL];
        // connections of the samples that finished in this iteration
        if (n_done > 0 && max_nc > 0 && max_nl > 0) {
            // every (camera vertex, light vertex) pair is at most one shadow ray: if they all fit, one round
            const uint64_t pairs = uint64_t(hc[C_PAIRS_LO]) | (uint64_t(hc[C_PAIRS_HI]) << 32);
            const int slots_per_round = pairs <= uint64_t(wf.scap) ? max_nc : int(std::max<int64_t>(1, wf.scap / (n_done * int64_t(max_nl)) - (J.mis ? 1 : 0)));
            const bool one_round = slots_per_round >= max_nc;
            // Batched shadow stage (J.shadow_batch > 0): the segments of SEVERAL iterations are appended to one queue and walked by
            // one launch.  A shadow launch of one iteration is less than one wave of the machine, so its duration is the latency chain
            // of its longest segment whatever it holds: fewer, fuller launches cost the same each.  The pair count (an upper bound
            // of the segments) decides when the batch is flushed; a multi-round iteration flushes and takes the unbatched path.
            if (acc_open && (!one_round || acc_bound + pairs > uint64_t(wf.scap))) flush_shadow_batch();
            for (int k0 = 0; k0 < max_nc; k0 += slots_per_round) {
                const int k1 = std::min(max_nc, k0 + slots_per_round);
                const SQ& q = wf.sq[sb];
                if (!acc_open) {
                    WK_CUDA(cudaStreamWaitEvent(st, wk.ev_shadow[sb], 0));   // the shadow stage that last read this queue
                    WK_CUDA(cudaMemsetAsync(q.cnt, 0, SQ_COUNT * sizeof(uint32_t), st));
                }
                tm.begin(KT_CONNECT, st);
                if (J.mis && J.media) k_connect_mis<true><<<grid_for(int64_t(k1 - k0) * n_done, kBlock), kBlock, 0, st>>>(sc, f, wf, q, wf.q_slots[dn], n_done, k0, k1);
                else if (J.mis) k_connect_mis<false><<<grid_for(int64_t(k1 - k0) * n_done, kBlock), kBlock, 0, st>>>(sc, f, wf, q, wf.q_slots[dn], n_done, k0, k1);
                else if (J.media) k_connect<true><<<grid_for(int64_t(k1 - k0) * n_done, kBlock), kBlock, 0, st>>>(sc, f, wf, q, wf.q_slots[dn], n_done, k0, k1);
                else k_connect<false><<<grid_for(int64_t(k1 - k0) * n_done, kBlock), kBlock, 0, st>>>(sc, f, wf, q, wf.q_slots[dn], n_done, k0, k1);
                tm.end(KT_CONNECT, st);
                wk.launches += 1;
                if (one_round && J.shadow_batch > 0) {   // keep filling this queue; flush when enough pairs have gone in
                    acc_open = true;
                    acc_bound += pairs;
                    if (acc_bound >= J.shadow_batch) flush_shadow_batch();
                    if (wk.rc != BDPT_OK) return;
                    continue;
                }
                // one round: size the shadow launch by the pair count (an upper bound) and skip the read-back
                uint32_t ns_bound = uint32_t(std::min<uint64_t>(pairs, uint64_t(wf.scap)));
                if (slots_per_round < max_nc) {
                    uint32_t* hq = hc + C_COUNT;
                    WK_CUDA(cudaMemcpyAsync(hq, q.cnt, SQ_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                    WK_CUDA(cudaStreamSynchronize(st));
                    if (hq[SQ_OVERFLOW]) { wk.err = "shadow queue overflow"; wk.rc = BDPT_ERR_STATE; return; }
                    ns_bound = hq[SQ_N];
                }
                if (ns_bound) {
                    WK_CUDA(cudaEventRecord(wk.ev_connect, st));
                    WK_CUDA(cudaStreamWaitEvent(st2, wk.ev_connect, 0));
                    launch_shadow(q, ns_bound);
                    WK_CUDA(cudaEventRecord(wk.ev_shadow[sb], st2));
                    sb ^= 1;
                }
            }
        }
        n_free = n_done;
        dn ^= 1;
        cur = nxt;
    }
    flush_shadow_batch();
    if (wk.rc != BDPT_OK) return;
    // segments still parked by the last shadow launch: one more launch of front blocks only (they never park)
    if (o.bvh != BDPT_BVH_REFERENCE) launch_shadow(wf.sq[0], 0);
    WK_CUDA(cudaStreamSynchronize(st2));
    WK_CUDA(cudaMemcpyAsync(hc, wf.counters, C_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    WK_CUDA(cudaStreamSynchronize(st));
    wk.rays_shadow += uint64_t(hc[C_SHADOW_TOTAL]) | (uint64_t(hc[C_SHADOW_TOTAL + 1]) << 32);
    tm.collect(wk.kernel_ms, wk.kernel_launches);
    WK_CUDA(cudaEventRecord(wk.done, st));
}

}  // namespace

void render_free(bdpt_ctx* ctx) {
    if (ctx->timers) { delete static_cast<KernelTimers*>(ctx->timers); ctx->timers = nullptr; }
    if (ctx->wf) { delete static_cast<Workers*>(ctx->wf); ctx->wf = nullptr; }
}

int render_impl(bdpt_ctx* ctx, const bdpt_camera* cam, const bdpt_render_opts* opts_in) {
    bdpt_render_opts o{};
    if (opts_in) o = *opts_in;
    if (cam->image_width <= 0 || cam->samples_per_pixel <= 0 || cam->aspect_ratio <= 0 || cam->max_depth < 0) {
        ctx->err = "bad camera"; return BDPT_ERR_ARG;
    }
    if (cam->defocus_angle > 0) { ctx->err = "defocus blur is off in the reference's loader (scene_loader.h:463)"; return BDPT_ERR_ARG; }
    if (o.bvh != BDPT_BVH_REFERENCE && o.bvh != BDPT_BVH_FAST) { ctx->err = "unknown bvh flavour"; return BDPT_ERR_ARG; }
    if (o.weighting != BDPT_WEIGHT_REF_SUM && o.weighting != BDPT_WEIGHT_MIS) { ctx->err = "unknown weighting"; return BDPT_ERR_ARG; }
    const bool mis = o.weighting == BDPT_WEIGHT_MIS;
    const bool textured = ctx->view.n_tex > 0;
    const bool has_media = ctx->view.n_media > 0;   // the shadow queues then carry what k_media_segments needs
    if (ctx->view.n_spheres > 0 && ctx->spheres_max_mat >= ctx->view.n_mats) { ctx->err = "a sphere names a material outside the scene's table"; return BDPT_ERR_ARG; }
    int H = int(cam->image_width / cam->aspect_ratio);   // camera.h:161-162
    H = H < 1 ? 1 : H;
    if (!ctx->film || ctx->film_w != cam->image_width || ctx->film_h != H) {
        int rc = bdpt_film_clear(ctx, cam->image_width, H);
        if (rc) return rc;
    }
    Frame f = make_frame(*cam, H);
    if (f.n < 1) { ctx->err = "samples_per_pixel < 1"; return BDPT_ERR_ARG; }
    f.seed = o.seed;
    const int n2 = f.n * f.n;
    int s_begin = std::max(0, o.stratum_begin), s_end = o.stratum_end <= 0 ? n2 : std::min(n2, o.stratum_end);
    const int s_step = o.stratum_step <= 0 ? 1 : o.stratum_step;
    const int row0 = std::max(0, o.row_begin), row1 = o.row_end <= 0 ? H : std::min(H, o.row_end);
    const int64_t n_strata = s_end > s_begin ? (s_end - s_begin + s_step - 1) / s_step : 0;
    f.row0 = row0; f.rows = std::max(0, row1 - row0);
    f.s_begin = s_begin; f.s_step = s_step;
    f.npix = int64_t(f.rows) * f.W;
    const int64_t total = n_strata * f.npix;
    ctx->last_ms = 0.0;
    if (total <= 0 || f.depth <= 0) {
        // depth 0: trace_path pushes nothing and adds no background (camera.h:333) -> black
        ctx->samples += uint64_t(std::max<int64_t>(0, total));
        return BDPT_OK;
    }

    // wavefront size and how many wavefronts are kept in flight
    int64_t B = o.batch_samples > 0 ? o.batch_samples : (int64_t(1) << 21);
    B = std::min<int64_t>(B, int64_t(1) << 30);
    int S = o.streams > 0 ? o.streams : 4;
    S = int(std::min<int64_t>(std::min(S, 16), std::max<int64_t>(1, total / 8192)));
    B = std::min<int64_t>(B, (total + S - 1) / S);   // every worker gets slots to fill
    if (o.batch_samples <= 0) {
        // the default takes at most half of the free device memory: a slot costs ~5 KB at depth 16 and ~25 KB at
        // depth 80 (vertex pools and shadow queues grow with the depth)
        WF probe{};
        probe.B = 65536; probe.depth = f.depth; probe.mis = mis ? 1 : 0; probe.textured = textured ? 1 : 0; probe.media = has_media ? 1 : 0;
        probe.scap = std::max<int64_t>(probe.B * 8, int64_t(f.depth) * probe.B * (mis ? 2 : 1));
        probe.sscap = std::max<int64_t>(4096, probe.scap / 8);
        carve(probe, nullptr);
        const double per_slot = double(probe.bytes) / double(probe.B);
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
            size_t held = 0;   // workspaces this context already holds can be reused or are freed before growing
            if (ctx->wf) for (auto* wk : static_cast<Workers*>(ctx->wf)->w) held += wk->wf.base ? wk->wf.bytes : 0;
            const int64_t cap = int64_t(0.5 * double(free_b + held) / (per_slot * S));
            B = std::max<int64_t>(4096, std::min<int64_t>(B, cap));
        }
    }
    Workers* ws = static_cast<Workers*>(ctx->wf);
    if (!ws) { ws = new Workers(); ctx->wf = ws; }
    while (int(ws->w.size()) < S) ws->w.push_back(new Worker());
    for (int i = 0; i < S; ++i) {
        Worker& wk = *ws->w[i];
        if (!wk.stream) {
            BDPT_CUDA(ctx, cudaStreamCreateWithFlags(&wk.stream, cudaStreamNonBlocking));
            BDPT_CUDA(ctx, cudaEventCreateWithFlags(&wk.done, cudaEventDisableTiming));
            BDPT_CUDA(ctx, cudaStreamCreateWithFlags(&wk.stream2, cudaStreamNonBlocking));
            BDPT_CUDA(ctx, cudaEventCreateWithFlags(&wk.ev_connect, cudaEventDisableTiming));
            for (auto& e : wk.ev_shadow) BDPT_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            BDPT_CUDA(ctx, cudaMallocHost(&wk.hc, (C_COUNT + SQ_COUNT) * sizeof(uint32_t)));
        }
        if (!wk.wf.base || wk.wf.B < B || wk.wf.depth < f.depth || wk.wf.mis != (mis ? 1 : 0) || wk.wf.textured != (textured ? 1 : 0) ||
            wk.wf.media != (has_media ? 1 : 0)) {
            if (wk.wf.base) { BDPT_CUDA(ctx, cudaStreamSynchronize(wk.stream)); cudaFree(wk.wf.base); }
            std::memset(&wk.wf, 0, sizeof(WF));
            wk.wf.B = B; wk.wf.depth = f.depth; wk.wf.mis = mis ? 1 : 0; wk.wf.textured = textured ? 1 : 0; wk.wf.media = has_media ? 1 : 0;
            wk.wf.scap = std::max<int64_t>(B * 8, int64_t(f.depth) * B * (mis ? 2 : 1));   // >= one camera slot against every light vertex (MIS: + the splats)
            wk.wf.scap = std::min<int64_t>(wk.wf.scap, int64_t(0xFFFFFFF0u));
            wk.wf.sscap = std::max<int64_t>(4096, wk.wf.scap / 8);
            carve(wk.wf, nullptr);
            void* base = nullptr;
            cudaError_t e = cudaMalloc(&base, wk.wf.bytes);
            if (e != cudaSuccess) { ctx->err = std::string("wavefront workspace: ") + cudaGetErrorString(e); wk.wf.base = nullptr; return BDPT_ERR_NOMEM; }
            wk.wf.base = base;
            carve(wk.wf, static_cast<char*>(base));
        }
        wk.rays_closest = wk.rays_shadow = wk.cam_verts = wk.light_verts = wk.launches = wk.samples = 0;
        wk.deferred_closest = wk.deferred_shadow = 0;
        for (int k = 0; k < 8; ++k) { wk.kernel_ms[k] = 0.0; wk.kernel_launches[k] = 0; }
        wk.err.clear(); wk.rc = BDPT_OK;
    }

    Job J;
    J.sc = ctx->view; J.f = f; J.o = o; J.total = total; J.film = ctx->film; J.d_stats = ctx->d_stats;
    // Shadow segments are collected over several iterations before their shadow stage is launched (run_worker): measured on the
    // bench step 21.85 -> 20.9 ms (380 -> 397 Msamples/s) from 4 M pairs per batch on, flat beyond (profiles/r2_shadow_batch_sweep.txt)
    J.shadow_batch = 4000000;
    if (const char* e = std::getenv("BDPT_SHADOW_BATCH")) J.shadow_batch = std::strtoull(e, nullptr, 10);
    J.mis = mis; J.tex = textured; J.media = ctx->view.n_media > 0 || ctx->view.n_spheres > 0;
    if (J.media && std::getenv("BDPT_WALK32") && std::atoi(std::getenv("BDPT_WALK32")) != 0) { ctx->err = "BDPT_WALK32 does not support participating media or analytic spheres"; return BDPT_ERR_ARG; }
    // traversal stack: one StackEntry per tree level per thread (deferred children only)
    const size_t levels = size_t(std::max(ctx->host.fast_depth, ctx->host.ref_depth)) + 2;
    if (ctx->host.fast_depth + 2 > uint32_t(kStackLevels) || ctx->host.ref_depth + 2 > uint32_t(2 * kStackLevels)) {
        ctx->err = "BVH deeper than the traversal stack"; return BDPT_ERR_ARG;
    }
    (void)levels;
    J.smem = 0;
    J.exact_grid = 148 * 2;
    J.device = ctx->device; J.n_workers = S; J.B = B;
    // Work limit: rays over this many nodes + triangles are parked and re-run split over a warp (traverse_walk.cuh:
    // split_walk).  Off by default: on the named scenes ~16 % of the unoccluded shadow segments exceed 160 steps
    // (mean 115), so parking them was measured slower; BDPT_WORK_LIMIT=n turns it on.
    J.limit = 0;
    if (const char* e = std::getenv("BDPT_WORK_LIMIT")) J.limit = uint32_t(std::strtoul(e, nullptr, 10));
    J.limit_shadow = J.limit;
    if (const char* e = std::getenv("BDPT_WORK_LIMIT_SHADOW")) J.limit_shadow = uint32_t(std::strtoul(e, nullptr, 10));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    J.resident_blocks = unsigned(sms * kTravBlocksPerSM);
#if BDPT_WIDE_WALK
    int slice = 16, min_idle = 8;   // trips per call of the lock-step walk and idle lanes that trigger a refill
#else
    int slice = 4, min_idle = 8;    // walk time slice (leaf visits) and idle lanes that trigger a refill
#endif
    int t_leaf = 16, t_walk = 8;    // lock-step walk: postponed leaves that start the triangle tests / walkers below which they start anyway
#if !BDPT_WIDE_WALK
    t_leaf = 1;                     // per-lane walk: these bits hold the shadow kernel's refill rounds instead (k_shadow_fast; 1 = off: 2 and 4 measured within noise, 8 slower)
    if (const char* e = std::getenv("BDPT_SHADOW_OVER")) t_leaf = std::atoi(e);
#endif
    if (const char* e = std::getenv("BDPT_WALK_SLICE")) slice = std::atoi(e);
    if (const char* e = std::getenv("BDPT_MIN_IDLE")) min_idle = std::atoi(e);
    if (const char* e = std::getenv("BDPT_T_LEAF")) t_leaf = std::atoi(e);
    if (const char* e = std::getenv("BDPT_T_WALK")) t_walk = std::atoi(e);
    J.threshold = int(uint32_t(slice & 0xFF) | (uint32_t(min_idle & 0xFF) << 8) | (uint32_t(t_leaf & 0xFF) << 16) | (uint32_t(t_walk & 0xFF) << 24));
    {
        // BDPT_WALK32=1: the warp-synchronous fp32 walk (traverse_w32.cuh).  Parity-green and at 15.3 instead of 8.4
        // active lanes per instruction, but measured SLOWER (31.9 vs 21.7 ms per bench step): the walk is bound by
        // dependent-load latency, and a warp in lock step waits for the slowest of its lanes at every node.
        // The default therefore stays the per-lane walk of traverse_walk.cuh.
        J.walk32 = false;
        if (const char* e = std::getenv("BDPT_WALK32")) J.walk32 = std::atoi(e) != 0;
        int trips = 16, nstep = 4, idle = 8, tleaf = 8, tnode = 4;
        if (const char* e = std::getenv("BDPT_W32_TRIPS")) trips = std::atoi(e);
        if (const char* e = std::getenv("BDPT_W32_NSTEP")) nstep = std::atoi(e);
        if (const char* e = std::getenv("BDPT_W32_IDLE")) idle = std::atoi(e);
        if (const char* e = std::getenv("BDPT_W32_TLEAF")) tleaf = std::atoi(e);
        if (const char* e = std::getenv("BDPT_W32_TNODE")) tnode = std::atoi(e);
        J.knobs32 = int(uint32_t(trips & 0x3F) | (uint32_t(nstep & 0xF) << 6) | (uint32_t(idle & 0x3F) << 10) | (uint32_t(tleaf & 0xFF) << 16) | (uint32_t(tnode & 0xFF) << 24));
        int per_sm = kW32BlocksPerSM;   // what the registers ptxas really used allow
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace_w32<false>, kBlock, 0) != cudaSuccess || per_sm < 1) per_sm = kW32BlocksPerSM;
        J.resident32 = unsigned(sms * per_sm);
    }
    std::atomic<int64_t> next_sample{0};
    J.next = &next_sample;
    J.progress = &ctx->progress_done;
    ctx->progress_done.store(0, std::memory_order_relaxed);
    ctx->progress_total.store(total, std::memory_order_relaxed);

    // the film clear (and anything else queued on the context stream) happens-before the workers
    BDPT_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    for (int i = 0; i < S; ++i) BDPT_CUDA(ctx, cudaStreamWaitEvent(ws->w[i]->stream, ctx->ev0, 0));
    std::vector<std::thread> pool;
    for (int i = 1; i < S; ++i) pool.emplace_back([&, i] { run_worker(J, *ws->w[i], i); });
    run_worker(J, *ws->w[0], 0);
    for (auto& th : pool) th.join();
    for (int i = 0; i < S; ++i) {
        Worker& wk = *ws->w[i];
        if (wk.rc != BDPT_OK) { ctx->err = wk.err; cudaDeviceSynchronize(); return wk.rc; }
        BDPT_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, wk.done, 0));
    }
    BDPT_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    BDPT_CUDA(ctx, cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    // tallies go into the device counter block so bdpt_stats reads one place
    unsigned long long cur_stats[ST_COUNT];
    BDPT_CUDA(ctx, cudaMemcpy(cur_stats, ctx->d_stats, sizeof cur_stats, cudaMemcpyDeviceToHost));
    for (int i = 0; i < S; ++i) {
        const Worker& wk = *ws->w[i];
        cur_stats[ST_RAYS_CLOSEST] += wk.rays_closest; cur_stats[ST_RAYS_SHADOW] += wk.rays_shadow;
        cur_stats[ST_CAM_VERTS] += wk.cam_verts; cur_stats[ST_LIGHT_VERTS] += wk.light_verts;
        cur_stats[ST_DEFER_CLOSEST] += wk.deferred_closest; cur_stats[ST_DEFER_SHADOW] += wk.deferred_shadow;
        ctx->launches += wk.launches; ctx->samples += wk.samples;
        for (int k = 0; k < 8; ++k) { ctx->kernel_ms[k] += wk.kernel_ms[k]; ctx->kernel_launches[k] += wk.kernel_launches[k]; }
    }
    BDPT_CUDA(ctx, cudaMemcpy(ctx->d_stats, cur_stats, sizeof cur_stats, cudaMemcpyHostToDevice));
    return BDPT_OK;
}

}  // namespace bdpt
