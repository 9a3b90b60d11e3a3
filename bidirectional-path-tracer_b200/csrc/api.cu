// C-ABI of libbdpt (include/bdpt.h): context, scene upload, batch trace / visible, film access.
// Compiled with -fmad=false: the trace kernels instantiate the reference-exact traversal.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <exception>
#include <stdexcept>
#include <thread>

#include "ctx.h"
#include "traverse_ref.cuh"
#include "traverse_fast.cuh"
#include "traverse_walk.cuh"
#include "traverse_w32.cuh"

using namespace bdpt;

namespace {

std::string g_create_err;

constexpr int kTraceBlock = 64;
#ifndef BDPT_TRACE_SLICE
#define BDPT_TRACE_SLICE 2
#endif
constexpr int kTraceSlice = BDPT_TRACE_SLICE;   // leaf visits between two meetings of a warp in k_trace


template <class T>
int upload(bdpt_ctx* ctx, const std::vector<T>& h, const T** d) {
    void* p = nullptr;
    size_t bytes = std::max<size_t>(h.size() * sizeof(T), 16);
    BDPT_CUDA(ctx, cudaMalloc(&p, bytes));
    ctx->scene_allocs.push_back(p);
    if (!h.empty()) BDPT_CUDA(ctx, cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *d = static_cast<const T*>(p);
    return BDPT_OK;
}

void free_scene(bdpt_ctx* ctx) {
    for (void* p : ctx->scene_allocs) cudaFree(p);
    ctx->scene_allocs.clear();
    ctx->have_scene = false;
}

// ---------------------------------------------------------------- read-bandwidth probe (bdpt_bench_read)
__global__ void __launch_bounds__(256) k_read_probe(const uint4* __restrict__ buf, uint64_t n16, unsigned long long* sink) {
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n16; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint4 v = __ldg(buf + i);
        acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x9E3779B9u) atomicAdd(sink, 1ull);   // keeps the loads alive; practically never true
}

// ---------------------------------------------------------------- fp32 triangle records of the walk's pre-filter
__global__ void k_make_tris32(const TriGeom* __restrict__ fast_tris, const uint32_t* __restrict__ fast_prims, uint32_t n,
                              Tri32* __restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = make_tri32(fast_tris[k], fast_prims[k]);
}

// ---------------------------------------------------------------- batch kernels
// hittable::hit over a ray batch (hittable.h:33, bvh.h:50-59)
template <int BVH, bool ANY, bool STATS, bool WIDE = false>
__global__ void __launch_bounds__(kTraceBlock, 1024 / kTraceBlock) k_trace(SceneView sc, const bdpt_ray* __restrict__ rays, uint64_t n,
                                                      bdpt_hit* __restrict__ out, unsigned long long* stats) {
    // stack in local memory (as in the render kernels): 48 int32 levels for the reference tree = 24 StackEntry;
    // 32 StackEntry cover the fast tree (depth <= 23)
    StackEntry l_stack[WIDE ? kWideStackLevels : 32];
    const uint64_t i = uint64_t(blockIdx.x) * kTraceBlock + threadIdx.x;
    TravStats st{0, 0};
    RayD r;
    HitD h;
    h.t = 0.0; h.u = 0.0; h.v = 0.0; h.prim = BDPT_MISS;
    bool found = false;
    if (i < n) {
        const double2* q = reinterpret_cast<const double2*>(rays + i);
        double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), e = __ldg(q + 3);
        r.o = mk(a.x, a.y, b.x); r.d = mk(b.y, c.x, c.y); r.tmin = e.x; r.tmax = e.y;
    }
    if (BVH == BDPT_BVH_REFERENCE) {
        if (i < n) found = traverse_ref<ANY, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &st);
    } else {
        // The render kernels' walk (fp32 FMA slabs) in time slices: a lane hands control back after kSlice leaf visits
        // and the warp meets at __syncwarp before walking on.  Without that meeting point the lanes of a warp drift apart
        // for the whole walk (ncu: 2 of 32 lanes active per instruction); with it they re-join every few leaves.
        Walk<ANY, STATS> W;
        const bool eligible = i < n && walk_eligible(sc, r);
        bool busy = eligible;
        int rc = TR_MISS;
        const int slice = int(sc.use_wide >> (ANY ? 16 : 8)) & 0xFF;
        if (busy) W.init(sc, r);
        for (;;) {
            __syncwarp();
            if (busy) {
                const int x = WIDE ? W.run4(sc, l_stack, slice) : W.run(sc, l_stack, slice);
                if (x != WALK_YIELD) { rc = x; busy = false; }
            }
            if (!__any_sync(0xFFFFFFFFu, busy)) break;
        }
        if (eligible) {
            if (STATS) { st.nodes += W.ts.nodes; st.tris += W.ts.tris; st.box_hits += W.ts.box_hits; st.tri_hits += W.ts.tri_hits; }
            h = W.hit;
            found = rc == TR_DEFER ? traverse_ref<ANY, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &st) : rc == TR_HIT;   // a tie goes to the reference tree
        } else if (i < n)
            found = traverse_fast<ANY, STATS, false>(sc, r, h, l_stack, 1, &st) == TR_HIT;
    }
    if (i < n) {
        double2* o = reinterpret_cast<double2*>(out + i);
        if (!found) { h.t = 0.0; h.u = 0.0; h.v = 0.0; h.prim = BDPT_MISS; }
        o[0] = make_double2(h.t, h.u);
        double2 w;
        w.x = h.v;
        w.y = __longlong_as_double((long long)(unsigned long long)h.prim);
        o[1] = w;
    }
    if (STATS) add_trav<2, BVH == BDPT_BVH_REFERENCE>(stats, st);
}

// hittable::hit over a ray batch on the warp-synchronous fp32 walk (traverse_w32.cuh): persistent warps pull rays
// through one atomic cursor, so a lane whose ray is decided takes the next one instead of idling behind its warp's
// longest walk.  The fp64 ray is re-read from the batch for the exact tests.  Rays the fp32 walk cannot take (zero
// direction component, far origin) or that end in a tie / knife edge go through the reference tree right here.
#ifndef BDPT_B32_BLOCKS
#define BDPT_B32_BLOCKS 16
#endif
__device__ __forceinline__ RayD load_batch_ray(const bdpt_ray* __restrict__ rays, uint64_t i) {
    const double2* q = reinterpret_cast<const double2*>(rays + i);
    const double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), e = __ldg(q + 3);
    RayD r;
    r.o = mk(a.x, a.y, b.x); r.d = mk(b.y, c.x, c.y); r.tmin = e.x; r.tmax = e.y;
    return r;
}
__device__ __forceinline__ void store_hit(bdpt_hit* __restrict__ out, uint64_t i, bool found, HitD h) {
    double2* o = reinterpret_cast<double2*>(out + i);
    if (!found) { h.t = 0.0; h.u = 0.0; h.v = 0.0; h.prim = BDPT_MISS; }
    o[0] = make_double2(h.t, h.u);
    double2 w;
    w.x = h.v;
    w.y = __longlong_as_double((long long)(unsigned long long)h.prim);
    o[1] = w;
}

template <bool ANY, bool STATS>
__global__ void __launch_bounds__(kTraceBlock, BDPT_B32_BLOCKS) k_trace_w32(SceneView sc, const bdpt_ray* __restrict__ rays, uint64_t n,
                                                                            bdpt_hit* __restrict__ out, unsigned long long* stats,
                                                                            unsigned long long* cursor, int knobs) {
    StackEntry l_stack[32];
    const int trips = knobs & 0x3F, nstep = (knobs >> 6) & 0xF, min_idle = (knobs >> 10) & 0x3F, t_leaf = (knobs >> 16) & 0xFF, t_node = (knobs >> 24) & 0xFF;
    const int lane = threadIdx.x & 31;
    Walk32<ANY, STATS> W;
    W.cur = kWalkDry;
    bool have = false, exhausted = false;
    int result = WALK_YIELD;
    uint64_t idx = 0;
    TravStats st{0, 0};
    for (;;) {
        __syncwarp();
        if (result == WALK32_FLUSH) {
            HitD h;
            W.decide(sc, load_batch_ray(rays, idx), h, false);
            result = WALK_YIELD;
        }
        if (result != WALK_YIELD) {   // decided, or never walkable: the exact finish
            const RayD r = load_batch_ray(rays, idx);
            HitD h;
            int rc = result;
            if (rc == TR_HIT) rc = W.decide(sc, r, h, true);
            if (STATS && have) { st.nodes += W.ts.nodes; st.tris += W.ts.tris; st.box_hits += W.ts.box_hits; st.tri_hits += W.ts.tri_hits; }
            bool found = rc == TR_HIT;
            if (rc == TR_DEFER) found = traverse_ref<ANY, STATS>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, &st);
            store_hit(out, idx, found, h);
            have = false;
            result = WALK_YIELD;
        }
        // refill the idle lanes
        {
            const bool need = !have;
            const uint32_t mask = __ballot_sync(0xFFFFFFFFu, need);
            if (mask && !exhausted && (mask == 0xFFFFFFFFu || __popc(mask) >= min_idle)) {
                const int leader = __ffs(mask) - 1;
                unsigned long long base = 0;
                if (lane == leader) base = atomicAdd(cursor, (unsigned long long)__popc(mask));
                base = __shfl_sync(0xFFFFFFFFu, base, leader);
                if (base + uint64_t(__popc(mask)) >= n) exhausted = true;
                const uint64_t my = base + uint64_t(__popc(mask & ((1u << lane) - 1u)));
                if (need && my < n) {
                    idx = my;
                    const RayD r = load_batch_ray(rays, idx);
                    if (walk_eligible(sc, r)) { W.init(sc, r); have = true; }
                    else result = TR_DEFER;
                }
            }
        }
        if (!__any_sync(0xFFFFFFFFu, have || result != WALK_YIELD)) break;
        const int rc = W.run(sc, l_stack, have, trips, nstep, t_leaf, t_node, exhausted ? 0 : min_idle);
        if (have) result = rc;
    }
    if (STATS) add_trav<2, false>(stats, st);
}

// the same, one ray per WARP through split_walk (traverse_walk.cuh); ties found on the way fall back to the
// reference traversal on lane 0, as in the render kernels' front blocks
template <bool ANY>
__global__ void __launch_bounds__(kTraceBlock) k_trace_split(SceneView sc, const bdpt_ray* __restrict__ rays, uint64_t n,
                                                            bdpt_hit* __restrict__ out) {
    StackEntry l_stack[32];
    const uint64_t i = (uint64_t(blockIdx.x) * kTraceBlock + threadIdx.x) >> 5;
    if (i >= n) return;   // warp-uniform
    const double2* q = reinterpret_cast<const double2*>(rays + i);
    double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), e = __ldg(q + 3);
    RayD r;
    r.o = mk(a.x, a.y, b.x); r.d = mk(b.y, c.x, c.y); r.tmin = e.x; r.tmax = e.y;
    HitD h;
    h.t = 0.0; h.u = 0.0; h.v = 0.0; h.prim = BDPT_MISS;
    TravStats ts{0, 0};
    int rc = TR_DEFER;
    if (walk_eligible(sc, r)) rc = split_walk<ANY, false>(sc, r, h, l_stack, ts);
    if ((threadIdx.x & 31) != 0) return;
    const bool found = rc == TR_DEFER ? traverse_ref<ANY, false>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, nullptr) : rc == TR_HIT;
    if (!found) { h.t = 0.0; h.u = 0.0; h.v = 0.0; h.prim = BDPT_MISS; }
    double2* o = reinterpret_cast<double2*>(out + i);
    o[0] = make_double2(h.t, h.u);
    double2 w;
    w.x = h.v;
    w.y = __longlong_as_double((long long)(unsigned long long)h.prim);
    o[1] = w;
}

// camera::visible (camera.h:425-438)
template <int BVH>
__global__ void __launch_bounds__(kTraceBlock, 1024 / kTraceBlock) k_visible(SceneView sc, const double* __restrict__ a,
                                                        const double* __restrict__ b, uint64_t n,
                                                        uint8_t* __restrict__ out) {
    StackEntry l_stack[32];
    const uint64_t i = uint64_t(blockIdx.x) * kTraceBlock + threadIdx.x;
    if (i >= n) return;
    D3 pa = mk(a[3 * i], a[3 * i + 1], a[3 * i + 2]), pb = mk(b[3 * i], b[3 * i + 1], b[3 * i + 2]);
    D3 dir = pb - pa;
    double dist = len(dir);
    uint8_t vis = 0;
    if (dist > 0) {
        D3 du = dir / dist;
        double max_t = dist - 0.001;
        if (max_t > 0) {
            RayD r;
            r.o = pa + (0.001 * du); r.d = du; r.tmin = 0.001; r.tmax = max_t;
            HitD h;
            bool found;
            if (BVH == BDPT_BVH_REFERENCE)
                found = traverse_ref<true, false>(sc, r, h, reinterpret_cast<int32_t*>(l_stack), 1, nullptr);
            else
                found = traverse_fast<true, false, false>(sc, r, h, l_stack, 1, nullptr) == TR_HIT;
            vis = found ? 0 : 1;
        }
    }
    out[i] = vis;
}

// The sphere prims of the side list (bdpt_scene_set_spheres) after the triangle walk of a batch query, as world.hit asks them
// (oracle/port.cpp: port_trace): in list order against the bound found so far; ANY: only rays no triangle occludes.
template <bool ANY>
__global__ void __launch_bounds__(256) k_trace_spheres(const SphereDev* __restrict__ spheres, uint32_t n_spheres, const bdpt_ray* __restrict__ rays,
                                                       uint64_t n, bdpt_hit* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * 256 + threadIdx.x;
    if (i >= n) return;
    const RayD r = load_batch_ray(rays, i);
    bdpt_hit h = out[i];
    if (ANY && h.prim != BDPT_MISS) return;
    double best = h.prim != BDPT_MISS ? h.t : r.tmax;
    bool changed = false;
    for (uint32_t k = 0; k < n_spheres; ++k) {
        double ts;
        if (sphere_hit_t(spheres[k], r.o, r.d, r.tmin, best, ts)) {
            best = ts; changed = true;
            h.t = ANY ? 0.0 : ts; h.u = 0.0; h.v = 0.0; h.prim = kSpherePrim | k;
            if (ANY) break;
        }
    }
    if (changed) out[i] = h;
}
// the same for bdpt_visible: a segment no triangle blocks may still meet a sphere
__global__ void __launch_bounds__(256) k_visible_spheres(const SphereDev* __restrict__ spheres, uint32_t n_spheres, const double* __restrict__ a,
                                                         const double* __restrict__ b, uint64_t n, uint8_t* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * 256 + threadIdx.x;
    if (i >= n || out[i] == 0) return;
    const D3 pa = mk(a[3 * i], a[3 * i + 1], a[3 * i + 2]), pb = mk(b[3 * i], b[3 * i + 1], b[3 * i + 2]);
    const D3 dir = pb - pa;
    const double dist = len(dir);
    const D3 du = dir / dist;                      // out[i] == 1: dist > 0 and max_t > 0 (k_visible)
    const double max_t = dist - 0.001;
    const D3 so = pa + (0.001 * du);
    for (uint32_t k = 0; k < n_spheres; ++k) {
        double ts;
        if (sphere_hit_t(spheres[k], so, du, 0.001, max_t, ts)) { out[i] = 0; return; }
    }
}


}  // namespace

// Everything the fp32 walk (traverse_w32.cuh) needs beyond the uploaded scene: the leaf-ordered triangles in fp32 and
// the coordinate bound of its error terms.  Derived on the device (10^7 triangles: ~1 ms) for every way a scene
// arrives -- host-built tree, device-built tree, scene image -- so neither the builders nor the image format change.
int bdpt::derive_walk32(bdpt_ctx* ctx, SceneView& v) {
    const double m = std::max(std::max(std::fabs(v.center.x), std::fabs(v.center.y)), std::fabs(v.center.z)) + 0.5 * v.extent;
    float f = float(m * 1.000001);
    if (double(f) < m) f = std::nextafterf(f, INFINITY);
    v.vmax32 = f;
    void* p = nullptr;
    const size_t n = std::max<size_t>(v.n_tris, 1);
    BDPT_CUDA(ctx, cudaMalloc(&p, n * sizeof(Tri32)));
    ctx->scene_allocs.push_back(p);
    if (v.n_tris) {
        k_make_tris32<<<(v.n_tris + 255) / 256, 256, 0, ctx->stream>>>(v.fast_tris, v.fast_prims, v.n_tris, static_cast<Tri32*>(p));
        ctx->launches += 1;
        BDPT_CUDA(ctx, cudaGetLastError());
    }
    v.tris32 = static_cast<const Tri32*>(p);
    return BDPT_OK;
}

// ---------------------------------------------------------------- C-ABI
extern "C" {

int bdpt_ctx_create(int device, bdpt_ctx** out) {
    if (!out) return BDPT_ERR_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        // no CPU fallback: the product path needs a CUDA device
        g_create_err = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return BDPT_ERR_CUDA;
    }
    if (device < 0 || device >= count) { g_create_err = "device index out of range"; return BDPT_ERR_ARG; }
    bdpt_ctx* ctx = new bdpt_ctx;
    ctx->device = device;
    auto fail = [&](const char* what, cudaError_t err) {
        g_create_err = std::string(what) + ": " + cudaGetErrorString(err);
        delete ctx;
        return BDPT_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreate(&ctx->ev0)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaMalloc(&ctx->d_stats, ST_COUNT * sizeof(unsigned long long))) != cudaSuccess) return fail("cudaMalloc", e);
    cudaMemset(ctx->d_stats, 0, ST_COUNT * sizeof(unsigned long long));
    *out = ctx;
    return BDPT_OK;
}

void bdpt_ctx_destroy(bdpt_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    render_free(ctx);
    comm_free(ctx);
    free_scene(ctx);
    if (ctx->film) cudaFree(ctx->film);
    if (ctx->d_tex) cudaFree(ctx->d_tex);
    if (ctx->d_tex_data) cudaFree(ctx->d_tex_data);
    if (ctx->d_media) cudaFree(ctx->d_media);
    if (ctx->d_spheres) cudaFree(ctx->d_spheres);
    if (ctx->d_med_tris) cudaFree(ctx->d_med_tris);
    if (ctx->d_med_nodes) cudaFree(ctx->d_med_nodes);
    if (ctx->d_stats) cudaFree(ctx->d_stats);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* bdpt_last_error(bdpt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int bdpt_scene_upload(bdpt_ctx* ctx, const double* verts, const uint32_t* mat_id, uint64_t n_tris,
                      const bdpt_material* mats, uint32_t n_mats) {
    return bdpt_scene_upload_ex(ctx, verts, mat_id, n_tris, mats, n_mats, 0u);
}

int bdpt_scene_upload_ex(bdpt_ctx* ctx, const double* verts, const uint32_t* mat_id, uint64_t n_tris,
                         const bdpt_material* mats, uint32_t n_mats, uint32_t flags) {
    if (!ctx) return BDPT_ERR_ARG;
    if (flags & ~(BDPT_UPLOAD_FAST_TREE_GPU | BDPT_UPLOAD_FAST_TREE_HOST)) { ctx->err = "unknown upload flags"; return BDPT_ERR_ARG; }
    // the fast tree is built on the device unless the host build is asked for (or its 4-wide form is: BDPT_WIDE)
    const bool gpu_tree = (flags & BDPT_UPLOAD_FAST_TREE_GPU) != 0 || (!(flags & BDPT_UPLOAD_FAST_TREE_HOST) && bdpt::wide_walk_mode() == 0u);
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool timing = std::getenv("BDPT_TIMING") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t) { return std::chrono::duration<double, std::milli>(now() - t).count(); };
    const auto t_start = now();
    double t_flat = 0, t_ref = 0, t_fast = 0;
    HostScene& h = ctx->host;
    SceneView v{};
    int rc = BDPT_OK;
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    free_scene(ctx);
    ctx->have_scene = false;
    try {
        flatten_scene(verts, mat_id, n_tris, mats, n_mats, h);
        t_flat = ms_since(t_start);
        // The two trees share only read-only inputs (tri_box): they are built side by side on their own threads
        // while this one copies what is already final, then each tree's arrays as soon as that tree is done.
        std::exception_ptr fast_err, ref_err;
        const auto t_b = now();
        std::thread fast_thread([&] {
            if (gpu_tree) return;
            try { build_fast_bvh(h); t_fast = ms_since(t_b); } catch (...) { fast_err = std::current_exception(); }
        });
        std::thread ref_thread([&] { try { build_ref_bvh(h); t_ref = ms_since(t_b); } catch (...) { ref_err = std::current_exception(); } });
        if (!rc) rc = upload(ctx, h.tris, &v.tris);
        if (!rc) rc = upload(ctx, h.shade, &v.shade);
        if (!rc) rc = upload(ctx, h.mats, &v.mats);
        if (!rc) rc = upload(ctx, h.lights, &v.lights);
        if (!rc) rc = upload(ctx, h.light_area, &v.light_area);
        fast_thread.join();
        if (!rc && gpu_tree) {
            double lo[3], hi[3];
            scene_bounds(h, lo, hi);
            h.fast_nodes.clear(); h.fast_prims.clear(); h.fast_tris.clear();
            rc = build_fast_bvh_gpu(ctx, v.tris, uint32_t(h.tris.size()), lo, hi, h.extent, &v.fast_nodes, &v.fast_prims, &v.fast_tris,
                                    &v.n_fast_nodes, &h.fast_depth);
            t_fast = ms_since(t_b);
        } else if (!rc && !fast_err) {
            rc = upload(ctx, h.fast_nodes, &v.fast_nodes);
            if (!rc && !h.wide_nodes.empty()) rc = upload(ctx, h.wide_nodes, &v.wide_nodes);
            v.n_wide_nodes = uint32_t(h.wide_nodes.size());
            if (!rc) rc = upload(ctx, h.fast_prims, &v.fast_prims);
            if (!rc) rc = upload(ctx, h.fast_tris, &v.fast_tris);
            v.n_fast_nodes = uint32_t(h.fast_nodes.size());
        }
        ref_thread.join();
        if (fast_err) std::rethrow_exception(fast_err);
        if (ref_err) std::rethrow_exception(ref_err);
    } catch (const std::exception& ex) {
        free_scene(ctx);
        ctx->err = ex.what();
        return BDPT_ERR_ARG;
    }
    if (!rc) rc = upload(ctx, h.ref_nodes, &v.ref_nodes);
    if (!rc) rc = upload(ctx, h.dfs_rank, &v.dfs_rank);
    if (!rc) rc = upload(ctx, h.ref_parent, &v.ref_parent);
    if (!rc) rc = upload(ctx, h.prim_leaf, &v.prim_leaf);
    if (!rc) rc = upload(ctx, h.chain_box, &v.chain_box);
    v.use_wide = (!gpu_tree && !h.wide_nodes.empty() && 3u * h.wide_depth + 2u <= uint32_t(kWideStackLevels)) ? bdpt::wide_walk_mode() : 0u;
    if (timing && !h.wide_nodes.empty()) std::fprintf(stderr, "bdpt_scene_upload: 4-wide tree: %zu nodes, depth %u, walk mode %u\n", h.wide_nodes.size(), h.wide_depth, v.use_wide);
    if (!rc && (h.ref_depth >= uint32_t(kRefStackLevels) || h.fast_depth + 2 > 32u)) {   // 64 int32 / 32 StackEntry slots of local stack
        ctx->err = "BVH deeper than the traversal stack";
        rc = BDPT_ERR_ARG;
    }
    if (rc) { free_scene(ctx); return rc; }
    v.n_lights = uint32_t(h.lights.size());
    v.n_mats = uint32_t(h.mats.size());
    v.n_tris = uint32_t(h.tris.size());
    v.n_ref_nodes = uint32_t(h.ref_nodes.size());
    v.light_total_area = h.light_total_area;
    v.extent = h.extent;
    v.center = mk(h.center[0], h.center[1], h.center[2]);
    rc = derive_walk32(ctx, v);
    if (rc) { free_scene(ctx); return rc; }
    apply_textures(ctx, v);
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->view = v;
    ctx->have_scene = true;
    if (timing)
        std::fprintf(stderr, "bdpt_scene_upload: %llu triangles: flatten %.1f ms, reference tree %.1f ms || fast tree %.1f ms, total %.1f ms\n",
                     (unsigned long long)n_tris, t_flat, t_ref, t_fast, ms_since(t_start));
    return BDPT_OK;
}

int bdpt_scene_set_textures(bdpt_ctx* ctx, const bdpt_texture* tex, uint32_t n_tex, const void* data, uint64_t data_bytes) {
    if (!ctx || (n_tex && !tex) || (data_bytes && !data)) return BDPT_ERR_ARG;
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    std::vector<Tex> h(n_tex);
    for (uint32_t i = 0; i < n_tex; ++i) {
        const bdpt_texture& t = tex[i];
        if (t.type > BDPT_TEX_IMAGE) { ctx->err = "unknown texture type"; return BDPT_ERR_ARG; }
        uint64_t need = 0;
        if (t.type == BDPT_TEX_NOISE) need = 256 * 3 * sizeof(double) + 3 * 256 * sizeof(int32_t);
        if (t.type == BDPT_TEX_IMAGE) {
            if (t.width < 0 || t.height < 0) { ctx->err = "image texture: negative size"; return BDPT_ERR_ARG; }
            need = uint64_t(t.width) * uint64_t(t.height) * 3;
        }
        if (need && ((t.data_offset & 7u) || t.data_offset > data_bytes || need > data_bytes - t.data_offset)) {
            ctx->err = "texture data outside the blob (or not 8-byte aligned)"; return BDPT_ERR_ARG;
        }
        if (t.type == BDPT_TEX_CHECKER && !(t.scale != 0.0)) { ctx->err = "checker texture: scale 0"; return BDPT_ERR_ARG; }
        if (t.type == BDPT_TEX_NOISE) {   // perm entries index randvec
            const int32_t* perm = reinterpret_cast<const int32_t*>(static_cast<const unsigned char*>(data) + t.data_offset + 256 * 3 * sizeof(double));
            for (int k = 0; k < 768; ++k) if (perm[k] < 0 || perm[k] > 255) { ctx->err = "noise texture: permutation entry outside 0..255"; return BDPT_ERR_ARG; }
        }
        Tex& d = h[i];
        d.ar = t.a[0]; d.ag = t.a[1]; d.ab = t.a[2]; d.br = t.b[0]; d.bg = t.b[1]; d.bb = t.b[2];
        d.scale = t.scale; d.data = t.data_offset; d.type = t.type; d.width = t.width; d.height = t.height; d.pad = 0;
    }
    if (ctx->d_tex) { cudaFree(ctx->d_tex); ctx->d_tex = nullptr; }
    if (ctx->d_tex_data) { cudaFree(ctx->d_tex_data); ctx->d_tex_data = nullptr; }
    ctx->n_tex = 0;
    if (n_tex) {
        BDPT_CUDA(ctx, cudaMalloc(&ctx->d_tex, n_tex * sizeof(Tex)));
        BDPT_CUDA(ctx, cudaMemcpy(ctx->d_tex, h.data(), n_tex * sizeof(Tex), cudaMemcpyHostToDevice));
        BDPT_CUDA(ctx, cudaMalloc(&ctx->d_tex_data, std::max<uint64_t>(data_bytes, 16)));
        if (data_bytes) BDPT_CUDA(ctx, cudaMemcpy(ctx->d_tex_data, data, data_bytes, cudaMemcpyHostToDevice));
        ctx->n_tex = n_tex;
    }
    apply_textures(ctx, ctx->view);
    return BDPT_OK;
}

int bdpt_scene_set_spheres(bdpt_ctx* ctx, const bdpt_sphere* spheres, uint32_t n_spheres) {
    if (!ctx || (n_spheres && !spheres)) return BDPT_ERR_ARG;
    if (n_spheres > 256) { ctx->err = "at most 256 analytic spheres (a side list, not a tree)"; return BDPT_ERR_ARG; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    std::vector<SphereDev> h(n_spheres);
    uint32_t max_mat = 0;
    for (uint32_t k = 0; k < n_spheres; ++k) {
        if (!(spheres[k].radius > 0.0)) { ctx->err = "sphere: radius must be > 0"; return BDPT_ERR_ARG; }
        SphereDev d{};
        d.cx = spheres[k].center[0]; d.cy = spheres[k].center[1]; d.cz = spheres[k].center[2]; d.r = spheres[k].radius; d.mat = spheres[k].mat_id;
        max_mat = std::max(max_mat, d.mat);
        h[k] = d;
    }
    if (ctx->d_spheres) { cudaFree(ctx->d_spheres); ctx->d_spheres = nullptr; }
    ctx->n_spheres = 0; ctx->spheres_max_mat = 0;
    if (n_spheres) {
        BDPT_CUDA(ctx, cudaMalloc(&ctx->d_spheres, n_spheres * sizeof(SphereDev)));
        BDPT_CUDA(ctx, cudaMemcpy(ctx->d_spheres, h.data(), n_spheres * sizeof(SphereDev), cudaMemcpyHostToDevice));
        ctx->n_spheres = n_spheres; ctx->spheres_max_mat = max_mat;
    }
    apply_textures(ctx, ctx->view);
    return BDPT_OK;
}

// Median-split tree over the boundary triangles [first, first + count) of one medium (reordered in place; their order means nothing:
// a boundary query returns the smallest t).  Leaves of <= 4 triangles; boxes padded well beyond any rounding of the triangle test.
static int32_t build_med_tree(std::vector<TriGeom>& tris, size_t first, size_t lo_i, size_t hi_i, std::vector<MedNode>& nodes, double pad) {
    const int32_t me = int32_t(nodes.size());
    nodes.push_back(MedNode{});
    MedNode n{};
    double clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int c = 0; c < 3; ++c) { n.lo[c] = INFINITY; n.hi[c] = -INFINITY; }
    auto vert = [](const TriGeom& g, int v, int c) {
        const double v0[3] = {g.v0x, g.v0y, g.v0z}, e1[3] = {g.e1x, g.e1y, g.e1z}, e2[3] = {g.e2x, g.e2y, g.e2z};
        return v == 0 ? v0[c] : (v == 1 ? v0[c] + e1[c] : v0[c] + e2[c]);
    };
    auto centroid = [&](const TriGeom& g, int c) { return (vert(g, 0, c) + vert(g, 1, c) + vert(g, 2, c)) / 3.0; };
    for (size_t i = lo_i; i < hi_i; ++i)
        for (int c = 0; c < 3; ++c) {
            for (int v = 0; v < 3; ++v) { const double x = vert(tris[i], v, c); n.lo[c] = std::fmin(n.lo[c], x); n.hi[c] = std::fmax(n.hi[c], x); }
            const double ce = centroid(tris[i], c);
            clo[c] = std::fmin(clo[c], ce); chi[c] = std::fmax(chi[c], ce);
        }
    for (int c = 0; c < 3; ++c) { n.lo[c] -= pad; n.hi[c] += pad; }
    if (hi_i - lo_i <= 4) {
        n.a = int32_t(lo_i - first); n.b = -int32_t(hi_i - lo_i);
    } else {
        int axis = 0;
        if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
        if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
        const size_t mid = lo_i + (hi_i - lo_i) / 2;
        std::nth_element(tris.begin() + lo_i, tris.begin() + mid, tris.begin() + hi_i,
                         [&](const TriGeom& x, const TriGeom& y) { return centroid(x, axis) < centroid(y, axis); });
        n.a = build_med_tree(tris, first, lo_i, mid, nodes, pad);
        n.b = build_med_tree(tris, first, mid, hi_i, nodes, pad);
    }
    nodes[size_t(me)] = n;
    return me;
}

int bdpt_scene_set_media(bdpt_ctx* ctx, const bdpt_medium* media, uint32_t n_media, const double* boundary_verts, uint64_t n_boundary_tris) {
    if (!ctx || (n_media && (!media || !boundary_verts))) return BDPT_ERR_ARG;
    if (n_media > 16) { ctx->err = "at most 16 media"; return BDPT_ERR_ARG; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    std::vector<MediumDev> h(n_media);
    std::vector<TriGeom> tris(n_boundary_tris);
    for (uint64_t i = 0; i < n_boundary_tris; ++i) {   // triangle.h:21-31: v0, edge1, edge2
        const double* p = boundary_verts + 9 * i;
        TriGeom g;
        g.v0x = p[0]; g.v0y = p[1]; g.v0z = p[2];
        g.e1x = p[3] - p[0]; g.e1y = p[4] - p[1]; g.e1z = p[5] - p[2];
        g.e2x = p[6] - p[0]; g.e2y = p[7] - p[1]; g.e2z = p[8] - p[2];
        g.mat = 0; g.pad = 0;
        tris[i] = g;
    }
    for (uint32_t k = 0; k < n_media; ++k) {
        const bdpt_medium& m = media[k];
        if (!(m.density > 0.0)) { ctx->err = "medium: density must be > 0"; return BDPT_ERR_ARG; }
        if (m.n_tris == 0 || m.n_tris > 4096 || m.first_tri > n_boundary_tris || m.n_tris > n_boundary_tris - m.first_tri) {
            ctx->err = "medium: boundary triangles outside the soup (1..4096 per medium)"; return BDPT_ERR_ARG;
        }
        MediumDev& d = h[k];
        d.neg_inv_density = -1 / m.density;   // constant_medium.h:12
        d.r = m.albedo[0]; d.g = m.albedo[1]; d.b = m.albedo[2];
        for (int c = 0; c < 3; ++c) { d.lo[c] = INFINITY; d.hi[c] = -INFINITY; }
        for (uint64_t i = m.first_tri; i < m.first_tri + m.n_tris; ++i)
            for (int vtx = 0; vtx < 3; ++vtx)
                for (int c = 0; c < 3; ++c) {
                    const double x = boundary_verts[9 * i + 3 * vtx + c];
                    d.lo[c] = std::fmin(d.lo[c], x); d.hi[c] = std::fmax(d.hi[c], x);
                }
        for (int c = 0; c < 3; ++c) {   // generous: the bounds only skip rays that cannot meet the mesh
            const double pad = 1e-6 * (std::fabs(d.lo[c]) + std::fabs(d.hi[c]) + (d.hi[c] - d.lo[c])) + 1e-12;
            d.lo[c] -= pad; d.hi[c] += pad;
        }
        d.first = uint32_t(m.first_tri); d.count = uint32_t(m.n_tris);
        d.nodes = nullptr;
    }
    // large boundaries get a tree (two media may share triangles: then the second one keeps its loop)
    std::vector<MedNode> nodes;
    std::vector<size_t> root_of(n_media, SIZE_MAX);
    std::vector<char> taken(n_boundary_tris, 0);
    for (uint32_t k = 0; k < n_media; ++k) {
        const MediumDev& d = h[k];
        bool shared = false;
        for (uint32_t i = 0; i < d.count; ++i) shared = shared || taken[d.first + i];
        for (uint32_t i = 0; i < d.count; ++i) taken[d.first + i] = 1;
        if (d.count <= kMedTreeMin || shared || std::getenv("BDPT_MEDIA_NO_TREE")) continue;
        const double ext = std::fmax(d.hi[0] - d.lo[0], std::fmax(d.hi[1] - d.lo[1], d.hi[2] - d.lo[2]));
        const double big = std::fmax(std::fmax(std::fabs(d.lo[0]) + std::fabs(d.hi[0]), std::fabs(d.lo[1]) + std::fabs(d.hi[1])), std::fabs(d.lo[2]) + std::fabs(d.hi[2]));
        root_of[k] = nodes.size();
        std::vector<MedNode> own;
        build_med_tree(tris, d.first, d.first, size_t(d.first) + d.count, own, 1e-7 * (ext + big) + 1e-12);
        // (child links are relative to the medium's own root: MediumDev.nodes points at it)
        nodes.insert(nodes.end(), own.begin(), own.end());
    }
    if (ctx->d_media) { cudaFree(ctx->d_media); ctx->d_media = nullptr; }
    if (ctx->d_med_tris) { cudaFree(ctx->d_med_tris); ctx->d_med_tris = nullptr; }
    if (ctx->d_med_nodes) { cudaFree(ctx->d_med_nodes); ctx->d_med_nodes = nullptr; }
    ctx->n_media = 0;
    if (n_media) {
        if (!nodes.empty()) {
            BDPT_CUDA(ctx, cudaMalloc(&ctx->d_med_nodes, nodes.size() * sizeof(MedNode)));
            BDPT_CUDA(ctx, cudaMemcpy(ctx->d_med_nodes, nodes.data(), nodes.size() * sizeof(MedNode), cudaMemcpyHostToDevice));
            for (uint32_t k = 0; k < n_media; ++k) if (root_of[k] != SIZE_MAX) h[k].nodes = ctx->d_med_nodes + root_of[k];
        }
        BDPT_CUDA(ctx, cudaMalloc(&ctx->d_media, n_media * sizeof(MediumDev)));
        BDPT_CUDA(ctx, cudaMemcpy(ctx->d_media, h.data(), n_media * sizeof(MediumDev), cudaMemcpyHostToDevice));
        BDPT_CUDA(ctx, cudaMalloc(&ctx->d_med_tris, std::max<uint64_t>(tris.size(), 1) * sizeof(TriGeom)));
        BDPT_CUDA(ctx, cudaMemcpy(ctx->d_med_tris, tris.data(), tris.size() * sizeof(TriGeom), cudaMemcpyHostToDevice));
        ctx->n_media = n_media;
    }
    apply_textures(ctx, ctx->view);
    return BDPT_OK;
}

int bdpt_bvh_info(bdpt_ctx* ctx, int bvh, uint64_t* n_nodes, uint32_t* max_depth) {
    if (!ctx) return BDPT_ERR_ARG;
    if (!ctx->have_scene) { ctx->err = "no scene uploaded"; return BDPT_ERR_STATE; }
    if (bvh == BDPT_BVH_REFERENCE) {
        if (n_nodes) *n_nodes = ctx->view.n_ref_nodes;
        if (max_depth) *max_depth = ctx->host.ref_depth;
    } else if (bvh == BDPT_BVH_FAST) {
        if (n_nodes) *n_nodes = ctx->view.n_fast_nodes;
        if (max_depth) *max_depth = ctx->host.fast_depth;
    } else { ctx->err = "unknown bvh flavour"; return BDPT_ERR_ARG; }
    return BDPT_OK;
}

int bdpt_host_bvh_build(const double* verts, uint64_t n_tris, uint32_t* order, uint64_t* ref_nodes,
                        uint64_t* fast_nodes) {
    if (!verts || n_tris == 0) return BDPT_ERR_ARG;
    try {
        HostScene h;
        std::vector<uint32_t> mat_id(n_tris, 0u);
        bdpt_material m{};
        m.type = BDPT_MAT_LAMBERTIAN;
        flatten_scene(verts, mat_id.data(), n_tris, &m, 1, h);
        build_ref_bvh(h);
        if (order) std::memcpy(order, h.leaf_order.data(), n_tris * sizeof(uint32_t));
        if (ref_nodes) *ref_nodes = h.ref_nodes.size();
        if (fast_nodes) { build_fast_bvh(h); *fast_nodes = h.fast_nodes.size(); }
    } catch (const std::exception& ex) {
        g_create_err = ex.what();
        return BDPT_ERR_ARG;
    }
    return BDPT_OK;
}

int bdpt_bvh_leaf_order(bdpt_ctx* ctx, uint32_t* order) {
    if (!ctx || !order) return BDPT_ERR_ARG;
    if (!ctx->have_scene) { ctx->err = "no scene uploaded"; return BDPT_ERR_STATE; }
    std::memcpy(order, ctx->host.leaf_order.data(), ctx->host.leaf_order.size() * sizeof(uint32_t));
    return BDPT_OK;
}

// world.hit's second half for a batch: the sphere side list, when the context has one
static int trace_spheres(bdpt_ctx* ctx, const bdpt_ray* d_rays, uint64_t n, bdpt_hit* d_out, bool any, cudaStream_t stream) {
    if (ctx->n_spheres == 0) return BDPT_OK;
    const unsigned grid = unsigned((n + 255) / 256);
    if (any) k_trace_spheres<true><<<grid, 256, 0, stream>>>(ctx->d_spheres, ctx->n_spheres, d_rays, n, d_out);
    else k_trace_spheres<false><<<grid, 256, 0, stream>>>(ctx->d_spheres, ctx->n_spheres, d_rays, n, d_out);
    ctx->launches += 1;
    BDPT_CUDA(ctx, cudaGetLastError());
    return BDPT_OK;
}

static int launch_trace(bdpt_ctx* ctx, int bvh, const bdpt_ray* d_rays, uint64_t n, bdpt_hit* d_out, int flags, cudaStream_t stream) {
    const bool any = (flags & BDPT_TRACE_ANY) != 0, stats = (flags & BDPT_TRACE_STATS) != 0;
    const unsigned grid = unsigned((n + kTraceBlock - 1) / kTraceBlock);
    SceneView v = ctx->view;
    {
        const char* e = std::getenv("BDPT_TRACE_SLICE");
        v.use_wide |= uint32_t((e ? std::atoi(e) : kTraceSlice) & 0xFF) << 8;
        const char* ea = std::getenv("BDPT_TRACE_SLICE_ANY");
        v.use_wide |= uint32_t((ea ? std::atoi(ea) : 0) & 0xFF) << 16;
    }
    unsigned long long* st = ctx->d_stats;
    if ((flags & BDPT_TRACE_SPLIT) && bvh == BDPT_BVH_FAST) {
        const unsigned wgrid = unsigned((n * 32 + kTraceBlock - 1) / kTraceBlock);
        if (any) k_trace_split<true><<<wgrid, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out);
        else k_trace_split<false><<<wgrid, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out);
        ctx->launches += 1;
        BDPT_CUDA(ctx, cudaGetLastError());
        return trace_spheres(ctx, d_rays, n, d_out, any, stream);
    }
#define BDPT_LAUNCH(B, A, S) k_trace<B, A, S><<<grid, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out, st)
    static const bool walk32 = [] { const char* e = std::getenv("BDPT_WALK32"); return e ? std::atoi(e) != 0 : false; }();   // opt-in: measured slower than the per-lane walk
    if (bvh == BDPT_BVH_FAST && walk32 && !(v.use_wide & 3u)) {
        // round 2: persistent warps on the warp-synchronous fp32 walk; the cursor is a stream-ordered scratch word
        auto knob = [](const char* name, int dflt) { const char* e = std::getenv(name); return e ? std::atoi(e) : dflt; };
        const int knobs = int(uint32_t(knob("BDPT_W32_TRIPS", 16) & 0x3F) | (uint32_t(knob("BDPT_W32_NSTEP", 4) & 0xF) << 6) |
                              (uint32_t(knob("BDPT_W32_IDLE", 8) & 0x3F) << 10) | (uint32_t(knob("BDPT_W32_TLEAF", 8) & 0xFF) << 16) |
                              (uint32_t(knob("BDPT_W32_TNODE", 4) & 0xFF) << 24));
        unsigned long long* cursor = nullptr;
        BDPT_CUDA(ctx, cudaMallocAsync(reinterpret_cast<void**>(&cursor), sizeof(unsigned long long), stream));
        BDPT_CUDA(ctx, cudaMemsetAsync(cursor, 0, sizeof(unsigned long long), stream));
        int sms = 148, per_sm = BDPT_B32_BLOCKS;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace_w32<false, false>, kTraceBlock, 0) != cudaSuccess || per_sm < 1) per_sm = BDPT_B32_BLOCKS;
        const unsigned pg = std::min<unsigned>(grid, unsigned(sms * per_sm));
        if (any) { if (stats) k_trace_w32<true, true><<<pg, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out, st, cursor, knobs); else k_trace_w32<true, false><<<pg, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out, st, cursor, knobs); }
        else     { if (stats) k_trace_w32<false, true><<<pg, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out, st, cursor, knobs); else k_trace_w32<false, false><<<pg, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out, st, cursor, knobs); }
        BDPT_CUDA(ctx, cudaFreeAsync(cursor, stream));
    } else if (bvh == BDPT_BVH_REFERENCE) {
        if (any) { if (stats) BDPT_LAUNCH(BDPT_BVH_REFERENCE, true, true); else BDPT_LAUNCH(BDPT_BVH_REFERENCE, true, false); }
        else     { if (stats) BDPT_LAUNCH(BDPT_BVH_REFERENCE, false, true); else BDPT_LAUNCH(BDPT_BVH_REFERENCE, false, false); }
    } else if ((v.use_wide & (any ? 2u : 1u)) && v.wide_nodes) {   // bits 8..15 of use_wide: the walk's time slice
#define BDPT_LAUNCH_W(A, S) k_trace<BDPT_BVH_FAST, A, S, true><<<grid, kTraceBlock, 0, stream>>>(v, d_rays, n, d_out, st)
        if (any) { if (stats) BDPT_LAUNCH_W(true, true); else BDPT_LAUNCH_W(true, false); }
        else     { if (stats) BDPT_LAUNCH_W(false, true); else BDPT_LAUNCH_W(false, false); }
#undef BDPT_LAUNCH_W
    } else {
        if (any) { if (stats) BDPT_LAUNCH(BDPT_BVH_FAST, true, true); else BDPT_LAUNCH(BDPT_BVH_FAST, true, false); }
        else     { if (stats) BDPT_LAUNCH(BDPT_BVH_FAST, false, true); else BDPT_LAUNCH(BDPT_BVH_FAST, false, false); }
    }
#undef BDPT_LAUNCH
    ctx->launches += 1;
    BDPT_CUDA(ctx, cudaGetLastError());
    return trace_spheres(ctx, d_rays, n, d_out, any, stream);
}

int bdpt_trace(bdpt_ctx* ctx, int bvh, const bdpt_ray* rays, uint64_t n, bdpt_hit* out, int flags) {
    if (!ctx) return BDPT_ERR_ARG;
    if (!ctx->have_scene) { ctx->err = "no scene uploaded"; return BDPT_ERR_STATE; }
    if (bvh != BDPT_BVH_REFERENCE && bvh != BDPT_BVH_FAST) { ctx->err = "unknown bvh flavour"; return BDPT_ERR_ARG; }
    if (n == 0) { ctx->last_ms = 0.0; return BDPT_OK; }   // empty batch: nothing to do
    if (!rays || !out) { ctx->err = "null ray/hit pointer"; return BDPT_ERR_ARG; }
    if (n > (uint64_t(1) << 31) * kTraceBlock) { ctx->err = "batch too large"; return BDPT_ERR_ARG; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool dev = (flags & BDPT_TRACE_DEVICE_PTRS) != 0;
    if (dev) {
        BDPT_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
        int rc = launch_trace(ctx, bvh, rays, n, out, flags, ctx->stream);
        if (rc) return rc;
        BDPT_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { ctx->err = std::string("trace: ") + cudaGetErrorString(e); return BDPT_ERR_CUDA; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
        ctx->last_ms = ms;
        return BDPT_OK;
    }
    // Host buffers: the batch goes through in chunks on two streams, so that a chunk's H2D copy, its kernel and its
    // D2H copy overlap the neighbours' (full-duplex PCIe; really asynchronous only when the caller's buffers are
    // pinned -- pageable memory degrades to what one big copy would do).  Device staging is two chunks, not n.
    constexpr uint64_t kChunk = uint64_t(1) << 20;
    constexpr int kLanes = 2;
    const uint64_t chunk = std::min<uint64_t>(n, kChunk), n_chunks = (n + chunk - 1) / chunk;
    struct Stage {
        cudaStream_t st = nullptr;
        void *r = nullptr, *o = nullptr;
    } lane[kLanes];
    std::vector<cudaEvent_t> ev(2 * n_chunks, nullptr);
    cudaError_t e = cudaSuccess;
    int rc = BDPT_OK;
    const int lanes = int(std::min<uint64_t>(kLanes, n_chunks));
    for (int k = 0; k < lanes && e == cudaSuccess; ++k) {
        e = cudaStreamCreateWithFlags(&lane[k].st, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&lane[k].r, chunk * sizeof(bdpt_ray));
        if (e == cudaSuccess) e = cudaMalloc(&lane[k].o, chunk * sizeof(bdpt_hit));
    }
    for (auto& x : ev) if (e == cudaSuccess) e = cudaEventCreate(&x);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);   // whatever was queued before (scene upload) is done
    for (uint64_t c = 0; c < n_chunks && e == cudaSuccess && rc == BDPT_OK; ++c) {
        Stage& L = lane[c % lanes];
        const uint64_t at = c * chunk, cnt = std::min<uint64_t>(chunk, n - at);
        e = cudaMemcpyAsync(L.r, rays + at, cnt * sizeof(bdpt_ray), cudaMemcpyHostToDevice, L.st);
        if (e == cudaSuccess) e = cudaEventRecord(ev[2 * c], L.st);
        if (e == cudaSuccess) rc = launch_trace(ctx, bvh, static_cast<const bdpt_ray*>(L.r), cnt, static_cast<bdpt_hit*>(L.o), flags, L.st);
        if (e == cudaSuccess && rc == BDPT_OK) e = cudaEventRecord(ev[2 * c + 1], L.st);
        if (e == cudaSuccess && rc == BDPT_OK) e = cudaMemcpyAsync(out + at, L.o, cnt * sizeof(bdpt_hit), cudaMemcpyDeviceToHost, L.st);
    }
    for (int k = 0; k < lanes; ++k)
        if (lane[k].st) { const cudaError_t s2 = cudaStreamSynchronize(lane[k].st); if (e == cudaSuccess) e = s2; }
    double total_ms = 0.0;   // kernel time only, summed over the chunks (they overlap the copies, not each other's kernels much)
    if (e == cudaSuccess && rc == BDPT_OK)
        for (uint64_t c = 0; c < n_chunks; ++c) { float ms = 0.f; if (cudaEventElapsedTime(&ms, ev[2 * c], ev[2 * c + 1]) == cudaSuccess) total_ms += ms; }
    for (auto& x : ev) if (x) cudaEventDestroy(x);
    for (int k = 0; k < lanes; ++k) {
        if (lane[k].r) cudaFree(lane[k].r);
        if (lane[k].o) cudaFree(lane[k].o);
        if (lane[k].st) cudaStreamDestroy(lane[k].st);
    }
    if (rc) return rc;
    if (e != cudaSuccess) { ctx->err = std::string("trace: ") + cudaGetErrorString(e); return e == cudaErrorMemoryAllocation ? BDPT_ERR_NOMEM : BDPT_ERR_CUDA; }
    ctx->last_ms = total_ms;
    return BDPT_OK;
}

int bdpt_visible(bdpt_ctx* ctx, int bvh, const double* a, const double* b, uint64_t n, uint8_t* out, int flags) {
    if (!ctx) return BDPT_ERR_ARG;
    if (!ctx->have_scene) { ctx->err = "no scene uploaded"; return BDPT_ERR_STATE; }
    if (bvh != BDPT_BVH_REFERENCE && bvh != BDPT_BVH_FAST) { ctx->err = "unknown bvh flavour"; return BDPT_ERR_ARG; }
    if (n == 0) { ctx->last_ms = 0.0; return BDPT_OK; }
    if (!a || !b || !out) { ctx->err = "null pointer"; return BDPT_ERR_ARG; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool dev = (flags & BDPT_TRACE_DEVICE_PTRS) != 0;
    const double *da = a, *db = b;
    uint8_t* dout = out;
    void *ta = nullptr, *tb = nullptr, *to = nullptr;
    if (!dev) {
        BDPT_CUDA(ctx, cudaMalloc(&ta, n * 24));
        BDPT_CUDA(ctx, cudaMalloc(&tb, n * 24));
        BDPT_CUDA(ctx, cudaMalloc(&to, n));
        BDPT_CUDA(ctx, cudaMemcpyAsync(ta, a, n * 24, cudaMemcpyHostToDevice, ctx->stream));
        BDPT_CUDA(ctx, cudaMemcpyAsync(tb, b, n * 24, cudaMemcpyHostToDevice, ctx->stream));
        da = static_cast<const double*>(ta); db = static_cast<const double*>(tb); dout = static_cast<uint8_t*>(to);
    }
    const unsigned grid = unsigned((n + kTraceBlock - 1) / kTraceBlock);
    BDPT_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if (bvh == BDPT_BVH_REFERENCE) k_visible<BDPT_BVH_REFERENCE><<<grid, kTraceBlock, 0, ctx->stream>>>(ctx->view, da, db, n, dout);
    else k_visible<BDPT_BVH_FAST><<<grid, kTraceBlock, 0, ctx->stream>>>(ctx->view, da, db, n, dout);
    ctx->launches += 1;
    if (ctx->n_spheres) {
        k_visible_spheres<<<unsigned((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_spheres, ctx->n_spheres, da, db, n, dout);
        ctx->launches += 1;
    }
    BDPT_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    if (!dev) BDPT_CUDA(ctx, cudaMemcpyAsync(out, dout, n, cudaMemcpyDeviceToHost, ctx->stream));
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (!dev) { cudaFree(ta); cudaFree(tb); cudaFree(to); }
    if (e != cudaSuccess) { ctx->err = std::string("visible: ") + cudaGetErrorString(e); return BDPT_ERR_CUDA; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    return BDPT_OK;
}

// wpng.h:14-35: average, clamp to [0,0.999], gamma 2, 256x truncate
__global__ void k_tonemap(const double* __restrict__ film, uint64_t n, double scale, uint8_t* __restrict__ out) {
    uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double x = film[i] * scale;
    x = x < 0.0 ? 0.0 : (x > 0.999 ? 0.999 : x);   // interval::clamp, interval.h:30-34 (NaN passes through)
    out[i] = static_cast<uint8_t>(256.0 * sqrt(x));
}

int bdpt_film_clear(bdpt_ctx* ctx, int width, int height) {
    if (!ctx || width <= 0 || height <= 0) return BDPT_ERR_ARG;
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->film || ctx->film_w != width || ctx->film_h != height) {
        BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (ctx->film) cudaFree(ctx->film);
        ctx->film = nullptr;
        BDPT_CUDA(ctx, cudaMalloc(&ctx->film, size_t(width) * height * 3 * sizeof(double)));
        ctx->film_w = width; ctx->film_h = height;
    }
    BDPT_CUDA(ctx, cudaMemsetAsync(ctx->film, 0, size_t(width) * height * 3 * sizeof(double), ctx->stream));
    return BDPT_OK;
}

int bdpt_film_device_ptr(bdpt_ctx* ctx, void** ptr, uint64_t* n_doubles) {
    if (!ctx || !ptr) return BDPT_ERR_ARG;
    if (!ctx->film) { ctx->err = "no film"; return BDPT_ERR_STATE; }
    *ptr = ctx->film;
    if (n_doubles) *n_doubles = uint64_t(ctx->film_w) * ctx->film_h * 3;
    return BDPT_OK;
}

int bdpt_film_size(bdpt_ctx* ctx, int* width, int* height) {
    if (!ctx) return BDPT_ERR_ARG;
    if (!ctx->film) { ctx->err = "no film"; return BDPT_ERR_STATE; }
    if (width) *width = ctx->film_w;
    if (height) *height = ctx->film_h;
    return BDPT_OK;
}

int bdpt_film_download(bdpt_ctx* ctx, double* film_sum) {
    if (!ctx || !film_sum) return BDPT_ERR_ARG;
    if (!ctx->film) { ctx->err = "no film"; return BDPT_ERR_STATE; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    BDPT_CUDA(ctx, cudaMemcpyAsync(film_sum, ctx->film, size_t(ctx->film_w) * ctx->film_h * 3 * sizeof(double),
                                   cudaMemcpyDeviceToHost, ctx->stream));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BDPT_OK;
}

int bdpt_film_to_rgb8(bdpt_ctx* ctx, int samples_per_pixel, uint8_t* out) {
    if (!ctx || !out) return BDPT_ERR_ARG;
    if (!ctx->film) { ctx->err = "no film"; return BDPT_ERR_STATE; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t n = uint64_t(ctx->film_w) * ctx->film_h * 3;
    uint8_t* d = nullptr;
    BDPT_CUDA(ctx, cudaMalloc(&d, n));
    const double scale = 1.0 / double(samples_per_pixel > 1 ? samples_per_pixel : 1);  // wpng.h:18
    k_tonemap<<<unsigned((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->film, n, scale, d);
    ctx->launches += 1;
    cudaError_t e = cudaMemcpyAsync(out, d, n, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    if (e != cudaSuccess) { ctx->err = std::string("tonemap: ") + cudaGetErrorString(e); return BDPT_ERR_CUDA; }
    return BDPT_OK;
}

// ---------------------------------------------------------------- bdpt_film_to_rgb8_noalloc
// The tone-map's device staging (W*H*3 bytes between k_tonemap and the D2H copy) in a file-scope device array instead of a
// cudaMalloc per call: for callers that may not allocate (CUDA graph capture around them, serving loops, harnesses that
// forbid cudaMalloc).  Every GPU that loads libbdpt holds the array until exit.  All calls on a device share it and it
// keeps the last call's bytes: nothing has to start at zero (k_tonemap writes every byte of [0, W*H*3) before the copy
// reads it), but two calls that overlap on the GPU race on it.
static __device__ __align__(256) unsigned char g_rgb8_scratch[BDPT_NOALLOC_RGB8_MAX_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<unsigned char*> g_rgb8_scratch_dev[kMaxDevices];   // device address of g_rgb8_scratch, per device ordinal

// In host code `g_rgb8_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress with the device current.  Looked up once per device.
int rgb8_scratch_for(bdpt_ctx* ctx, unsigned char** out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices) { ctx->err = "device ordinal beyond the no-alloc scratch table"; return BDPT_ERR_ARG; }
    unsigned char* p = g_rgb8_scratch_dev[ctx->device].load(std::memory_order_acquire);
    if (!p) {
        BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
        void* sym = nullptr;
        BDPT_CUDA(ctx, cudaGetSymbolAddress(&sym, g_rgb8_scratch));
        if (!sym) { ctx->err = "cudaGetSymbolAddress returned a null scratch address"; return BDPT_ERR_CUDA; }
        p = static_cast<unsigned char*>(sym);
        g_rgb8_scratch_dev[ctx->device].store(p, std::memory_order_release);
    }
    *out = p;
    return BDPT_OK;
}
}  // namespace

int bdpt_film_to_rgb8_noalloc(bdpt_ctx* ctx, int samples_per_pixel, uint8_t* out) {
    if (!ctx || !out) return BDPT_ERR_ARG;
    if (!ctx->film) { ctx->err = "no film"; return BDPT_ERR_STATE; }
    const uint64_t n = uint64_t(ctx->film_w) * ctx->film_h * 3;
    if (n > uint64_t(BDPT_NOALLOC_RGB8_MAX_BYTES)) {
        ctx->err = "bdpt_film_to_rgb8_noalloc: film of " + std::to_string(ctx->film_w) + " x " + std::to_string(ctx->film_h) + " needs " +
                   std::to_string(n) + " bytes of staging, the static scratch holds " + std::to_string(uint64_t(BDPT_NOALLOC_RGB8_MAX_BYTES)) +
                   " (BDPT_NOALLOC_RGB8_MAX_BYTES); use bdpt_film_to_rgb8";
        return BDPT_ERR_ARG;
    }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* d = nullptr;
    if (int rc = rgb8_scratch_for(ctx, &d)) return rc;
    const double scale = 1.0 / double(samples_per_pixel > 1 ? samples_per_pixel : 1);  // wpng.h:18
    k_tonemap<<<unsigned((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->film, n, scale, d);
    ctx->launches += 1;
    BDPT_CUDA(ctx, cudaGetLastError());
    cudaError_t e = cudaMemcpyAsync(out, d, n, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { ctx->err = std::string("tonemap (no-alloc): ") + cudaGetErrorString(e); return BDPT_ERR_CUDA; }
    return BDPT_OK;
}

int bdpt_render(bdpt_ctx* ctx, const bdpt_camera* cam, const bdpt_render_opts* opts) {
    if (!ctx || !cam) return BDPT_ERR_ARG;
    if (!ctx->have_scene) { ctx->err = "no scene uploaded"; return BDPT_ERR_STATE; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    return render_impl(ctx, cam, opts);
}

int bdpt_stats(bdpt_ctx* ctx, bdpt_stats_t* out) {
    if (!ctx || !out) return BDPT_ERR_ARG;
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned long long h[ST_COUNT];
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    BDPT_CUDA(ctx, cudaMemcpy(h, ctx->d_stats, sizeof h, cudaMemcpyDeviceToHost));
    out->node_visits = h[ST_NODE]; out->tri_tests = h[ST_TRI];
    out->rays_closest = h[ST_RAYS_CLOSEST]; out->rays_shadow = h[ST_RAYS_SHADOW];
    out->camera_vertices = h[ST_CAM_VERTS]; out->light_vertices = h[ST_LIGHT_VERTS];
    out->samples = ctx->samples; out->kernel_launches = ctx->launches;
    out->nodes_closest = h[ST_NODE_CLOSEST]; out->tris_closest = h[ST_TRI_CLOSEST];
    out->nodes_shadow = h[ST_NODE_SHADOW]; out->tris_shadow = h[ST_TRI_SHADOW];
    out->max_work_closest = h[ST_MAX_CLOSEST]; out->max_work_shadow = h[ST_MAX_SHADOW];
    out->deferred_closest = h[ST_DEFER_CLOSEST]; out->deferred_shadow = h[ST_DEFER_SHADOW];
    out->aabb_tests = h[ST_BOX_TESTS]; out->aabb_hits = h[ST_BOX_HITS]; out->triangle_hits = h[ST_TRI_HITS];
    out->bvh_nodes_built = ctx->have_scene ? ctx->view.n_ref_nodes : 0;
    return BDPT_OK;
}

int bdpt_stats_reset(bdpt_ctx* ctx) {
    if (!ctx) return BDPT_ERR_ARG;
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    BDPT_CUDA(ctx, cudaMemsetAsync(ctx->d_stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
    ctx->samples = 0; ctx->launches = 0;
    for (int i = 0; i < 8; ++i) { ctx->kernel_ms[i] = 0.0; ctx->kernel_launches[i] = 0; }
    return BDPT_OK;
}

int bdpt_last_device_ms(bdpt_ctx* ctx, double* ms) {
    if (!ctx || !ms) return BDPT_ERR_ARG;
    *ms = ctx->last_ms;
    return BDPT_OK;
}

int bdpt_kernel_times(bdpt_ctx* ctx, double* ms, uint64_t* launches) {
    if (!ctx || !ms) return BDPT_ERR_ARG;
    for (int i = 0; i < KT_COUNT; ++i) { ms[i] = ctx->kernel_ms[i]; if (launches) launches[i] = ctx->kernel_launches[i]; }
    return BDPT_OK;
}

int bdpt_render_progress(bdpt_ctx* ctx, uint64_t* samples_done, uint64_t* samples_total) {
    if (!ctx) return BDPT_ERR_ARG;
    const int64_t t = ctx->progress_total.load(std::memory_order_relaxed);
    const int64_t d = std::min(t, ctx->progress_done.load(std::memory_order_relaxed));
    if (samples_done) *samples_done = uint64_t(d < 0 ? 0 : d);
    if (samples_total) *samples_total = uint64_t(t < 0 ? 0 : t);
    return BDPT_OK;
}

// Read bandwidth of a device buffer of `bytes` bytes, streamed `reps` times by a grid of 16-byte loads (after two
// untimed passes): with a buffer well inside the 126 MB L2 this is the L2-resident read peak the traversal kernels'
// roofline is quoted against beside the HBM figure; with a buffer far beyond it, the HBM read rate.
int bdpt_bench_read(bdpt_ctx* ctx, uint64_t bytes, int reps, double* gb_per_s) {
    if (!ctx || !gb_per_s || bytes < 4096 || reps < 1) return BDPT_ERR_ARG;
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    void* buf = nullptr;
    BDPT_CUDA(ctx, cudaMalloc(&buf, bytes));
    cudaMemsetAsync(buf, 1, bytes, ctx->stream);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    const uint64_t n16 = bytes / 16;
    const unsigned grid = unsigned(sms * 8);
    for (int k = 0; k < 2; ++k) k_read_probe<<<grid, 256, 0, ctx->stream>>>(static_cast<const uint4*>(buf), n16, ctx->d_stats + ST_COUNT - 1);
    cudaEventRecord(ctx->ev0, ctx->stream);
    for (int k = 0; k < reps; ++k) k_read_probe<<<grid, 256, 0, ctx->stream>>>(static_cast<const uint4*>(buf), n16, ctx->d_stats + ST_COUNT - 1);
    cudaEventRecord(ctx->ev1, ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    cudaFree(buf);
    ctx->launches += uint64_t(reps) + 2;
    if (e != cudaSuccess) { ctx->err = std::string("bdpt_bench_read: ") + cudaGetErrorString(e); return BDPT_ERR_CUDA; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    *gb_per_s = double(n16 * 16) * reps / (double(ms) * 1e-3) / 1e9;
    return BDPT_OK;
}

int bdpt_sync(bdpt_ctx* ctx) {
    if (!ctx) return BDPT_ERR_ARG;
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BDPT_OK;
}

// Host-only test hook (tests/test_host.py): the walk's fp32 pre-filter (traverse_w32.cuh) on one ray / one triangle,
// exactly the arithmetic the kernels run (the same inline functions compiled for the host; only the device's
// approximate division differs, inside the stated inflation).  ray = o, d, tmin, tmax; best = current upper bound.
// Returns 0 reject / 1 candidate / 2 certain, *t_hi as filter32.
int bdpt_test_filter32(const double* ray, double best, const double* tri_v0e1e2, double vmax, int any, float* t_hi) {
    if (!ray || !tri_v0e1e2) return BDPT_ERR_ARG;
    TriGeom t{};
    t.v0x = tri_v0e1e2[0]; t.v0y = tri_v0e1e2[1]; t.v0z = tri_v0e1e2[2];
    t.e1x = tri_v0e1e2[3]; t.e1y = tri_v0e1e2[4]; t.e1z = tri_v0e1e2[5];
    t.e2x = tri_v0e1e2[6]; t.e2y = tri_v0e1e2[7]; t.e2z = tri_v0e1e2[8];
    const Tri32 T = make_tri32(t, 0u);
    const Ray32 R = make_ray32(ray, ray + 3, ray[6], ray[7], f32_up(vmax * 1.000001));
    float hi = 0.f;
    const float b = f32_up(best);
    const int f = any ? filter32<true>(R, b, T.q0, T.q1, T.q2, hi) : filter32<false>(R, b, T.q0, T.q1, T.q2, hi);
    if (t_hi) *t_hi = hi;
    return f;
}

}  // extern "C"
