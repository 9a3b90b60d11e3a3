// Fast-tree builder on the GPU (replaces the host build of the acceleration tree, bvh.h:13-48, for
// big scenes).  It produces the same structure the walk reads (FastNode: both child boxes in the parent, fp32
// rounded outward; leaves of <= 2 triangles in a leaf-ordered triangle copy), so every kernel and every parity
// property is unchanged: the tree only decides which triangles the exact fp64 test sees.
//
//   1. k_morton      63-bit Morton code of each triangle's box centre within the scene bounds
//   2. cub radix sort (code, prim id)
//   3. k_gather      triangles copied into sorted order (fast_tris)
//   4. k_split       top-down, one launch per level: a range splits where the highest differing code bit
//                    flips (spatial median, LBVH), clamped so that no child holds more than 2^(kMaxLevel - level)
//                    triangles: the depth is bounded by construction (the walk's stack has 32 slots), whatever
//                    the distribution.  Ranges of <= 2 triangles are leaves.
//   5. k_boxes       bottom-up, one launch per level: child boxes from leaves' triangles or from the children's
//                    own boxes (fp64), padded and rounded outward like the host builder (fast_bvh.cu).
// Node ids are handed out per level by an atomic counter, so levels are contiguous in memory (the top of the
// tree shares cache lines); ids within a level vary from run to run, results do not.
//
// Round 2, SAH at every level (default; BDPT_GPU_TREE_SAH=0 keeps the Morton splits above the bottoms, BDPT_GPU_TREE_LBVH=1 gives
// the plain tree above): ranges of more than kSahLeafMax triangles are split by binned SAH with an in-place, stable partition
// (k_sah_big: one thread block per node, 16 bins on each axis, the host builder's algorithm one level at a time; the Morton sort
// stays as the initial order, which makes the bins of neighbouring triangles coincide and the partition's writes coalesce), and
// a range of at most kSahLeafMax triangles is not split by its Morton codes but by the exact surface-area heuristic -- its triangles sorted by centroid on each axis in
// turn, every position swept, the cheapest kept and the range's ids re-ordered accordingly (one thread per node, everything
// in local memory).  Most of a ray's node visits are in the last levels above the surface it hits, which is where Z-order
// splits of a few dozen triangles of a surface are poor (long overlapping boxes); the spatial-median top stays.  A variant
// with a binned-SAH TOP over Morton clusters (HLBVH's scheme) measured no better than the plain tree and was dropped
// (profiles/r2_rejected_experiments.txt).
#include "ctx.h"

#include <cub/device/device_radix_sort.cuh>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace bdpt {
namespace {

constexpr int kMaxLevel = 27;   // internal nodes live on levels 0..kMaxLevel: at most 28 stack entries per ray
constexpr int kBuildBlock = 256;

__device__ __forceinline__ void grow_tri(const TriGeom& t, double lo[3], double hi[3]) {
    const double p[3][3] = {{t.v0x, t.v0y, t.v0z}, {t.v0x + t.e1x, t.v0y + t.e1y, t.v0z + t.e1z}, {t.v0x + t.e2x, t.v0y + t.e2y, t.v0z + t.e2z}};
#pragma unroll
    for (int v = 0; v < 3; ++v)
#pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] = fmin(lo[k], p[v][k]); hi[k] = fmax(hi[k], p[v][k]); }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x) {   // 21 bits -> every third bit
    x &= 0x1FFFFFull;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(kBuildBlock) k_morton(const TriGeom* __restrict__ tris, uint32_t n, double lx, double ly, double lz,
                                                        double sx, double sy, double sz, unsigned long long* __restrict__ keys,
                                                        uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * kBuildBlock + threadIdx.x;
    if (i >= n) return;
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    grow_tri(tris[i], lo, hi);
    const double c[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
    const double q[3] = {(c[0] - lx) * sx, (c[1] - ly) * sy, (c[2] - lz) * sz};   // [0, 2^21)
    unsigned long long code = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double v = fmin(fmax(q[k], 0.0), 2097151.0);
        code |= spread21((unsigned long long)v) << (2 - k);
    }
    keys[i] = code;
    ids[i] = i;
}

__global__ void __launch_bounds__(kBuildBlock) k_gather(const TriGeom* __restrict__ tris, const uint32_t* __restrict__ prims, uint32_t n,
                                                        TriGeom* __restrict__ out) {
    const uint32_t i = blockIdx.x * kBuildBlock + threadIdx.x;
    if (i >= n) return;
    const int4* src = reinterpret_cast<const int4*>(tris + prims[i]);
    int4* dst = reinterpret_cast<int4*>(out + i);
#pragma unroll
    for (int k = 0; k < 5; ++k) dst[k] = src[k];
}

__device__ __forceinline__ int32_t leaf_code(uint32_t first, uint32_t cnt) { return int32_t(~((first << 3) | (cnt - 1u))); }

// fp32 box of every triangle, by prim id (outward-rounded; only the split heuristic reads them)
struct __align__(16) PBox { float lx, ly, lz, p0, hx, hy, hz, p1; };
static_assert(sizeof(PBox) == 32, "PBox must be 32 bytes");
__global__ void __launch_bounds__(kBuildBlock) k_prim_boxes(const TriGeom* __restrict__ tris, uint32_t n, PBox* __restrict__ pbox) {
    const uint32_t i = blockIdx.x * kBuildBlock + threadIdx.x;
    if (i >= n) return;
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    grow_tri(tris[i], lo, hi);
    pbox[i] = PBox{__double2float_rd(lo[0]), __double2float_rd(lo[1]), __double2float_rd(lo[2]), 0.f,
                   __double2float_ru(hi[0]), __double2float_ru(hi[1]), __double2float_ru(hi[2]), 0.f};
}

constexpr int kSahLeafMax = 32;   // ranges up to this many triangles are split by the exact SAH sweep (local arrays of this size)

// Exact SAH split of ids[b, e) (cnt <= kSahLeafMax): returns the split position (b < mid < e) and leaves ids re-ordered so that
// [b, mid) and [mid, e) are the two sides.  Cost of a split after k triangles of an axis-sorted order: A(L) k + A(R) (cnt - k).
__device__ __forceinline__ uint32_t sah_split_small(uint32_t* __restrict__ ids, const PBox* __restrict__ pbox, uint32_t b, uint32_t cnt) {
    uint32_t id[kSahLeafMax];
    float lo[kSahLeafMax][3], hi[kSahLeafMax][3], key[kSahLeafMax];
    uint8_t ord[kSahLeafMax], best_ord[kSahLeafMax];
    for (uint32_t k = 0; k < cnt; ++k) {
        id[k] = ids[b + k];
        const float4 a = __ldg(reinterpret_cast<const float4*>(pbox + id[k])), c = __ldg(reinterpret_cast<const float4*>(pbox + id[k]) + 1);
        lo[k][0] = a.x; lo[k][1] = a.y; lo[k][2] = a.z; hi[k][0] = c.x; hi[k][1] = c.y; hi[k][2] = c.z;
        best_ord[k] = uint8_t(k);
    }
    float best = INFINITY;
    uint32_t best_k = cnt / 2;
    float rarea[kSahLeafMax];
    for (int axis = 0; axis < 3; ++axis) {
        for (uint32_t k = 0; k < cnt; ++k) { ord[k] = uint8_t(k); key[k] = lo[k][axis] + hi[k][axis]; }
        for (uint32_t k = 1; k < cnt; ++k) {   // insertion sort of ord by centroid (ties: by position, deterministic)
            const uint8_t o = ord[k];
            const float v = key[o];
            uint32_t j = k;
            while (j > 0 && key[ord[j - 1]] > v) { ord[j] = ord[j - 1]; --j; }
            ord[j] = o;
        }
        float l3[3] = {INFINITY, INFINITY, INFINITY}, h3[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t k = cnt; k-- > 1;) {   // suffix areas: rarea[k] = area of ord[k..cnt)
            const uint8_t o = ord[k];
            for (int c = 0; c < 3; ++c) { l3[c] = fminf(l3[c], lo[o][c]); h3[c] = fmaxf(h3[c], hi[o][c]); }
            const float dx = h3[0] - l3[0], dy = h3[1] - l3[1], dz = h3[2] - l3[2];
            rarea[k] = dx * dy + dy * dz + dz * dx;
        }
        for (int c = 0; c < 3; ++c) { l3[c] = INFINITY; h3[c] = -INFINITY; }
        for (uint32_t k = 1; k < cnt; ++k) {   // split after k triangles
            const uint8_t o = ord[k - 1];
            for (int c = 0; c < 3; ++c) { l3[c] = fminf(l3[c], lo[o][c]); h3[c] = fmaxf(h3[c], hi[o][c]); }
            const float dx = h3[0] - l3[0], dy = h3[1] - l3[1], dz = h3[2] - l3[2];
            const float cost = (dx * dy + dy * dz + dz * dx) * float(k) + rarea[k] * float(cnt - k);
            if (cost < best) {
                best = cost; best_k = k;
                for (uint32_t q = 0; q < cnt; ++q) best_ord[q] = ord[q];
            }
        }
    }
    for (uint32_t k = 0; k < cnt; ++k) ids[b + k] = id[best_ord[k]];
    return b + best_k;
}

// nodes [first_node, first_node + n_level) of one level split their ranges; children that stay internal get ids
// from `counter` (all of them on the next level)
__global__ void __launch_bounds__(kBuildBlock) k_split(const unsigned long long* __restrict__ keys, uint32_t* __restrict__ nb,
                                                       uint32_t* __restrict__ ne, FastNode* __restrict__ nodes, uint32_t* counter,
                                                       uint32_t first_node, uint32_t n_level, int level,
                                                       uint32_t* __restrict__ ids, const PBox* __restrict__ pbox, uint32_t sah_leaf, int sah_all) {
    const uint32_t j = blockIdx.x * kBuildBlock + threadIdx.x;
    if (j >= n_level) return;
    const uint32_t i = first_node + j;
    const uint32_t b = nb[i], e = ne[i], cnt = e - b;
    if (sah_all && cnt > sah_leaf) return;   // k_sah_big's node
    int32_t link[2];
    if (cnt == 1) {   // a one-triangle scene: the root needs two children
        link[0] = link[1] = leaf_code(b, 1);
    } else {
        const unsigned long long kf = keys[b], kl = keys[e - 1];
        uint32_t mid;
        if (cnt <= sah_leaf) mid = sah_split_small(ids, pbox, b, cnt);   // (keys are not read again below this node)
        else if (kf == kl) mid = b + cnt / 2;
        else {
            const int p = __clzll((long long)(kf ^ kl));
            uint32_t lo = b, hi = e - 1;   // keys[lo] shares more than p leading bits with kf, keys[hi] exactly p
            while (hi - lo > 1) {
                const uint32_t m = lo + (hi - lo) / 2;
                if (__clzll((long long)(kf ^ keys[m])) > p) lo = m; else hi = m;
            }
            mid = hi;
        }
        const uint32_t cap = 1u << (kMaxLevel - level);   // cnt <= 2 * cap by induction (host checks the root)
        if (e - mid > cap) mid = e - cap;
        if (mid - b > cap) mid = b + cap;
        const uint32_t cb[2] = {b, mid}, ce[2] = {mid, e};
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const uint32_t c = ce[s] - cb[s];
            if (c <= 2) link[s] = leaf_code(cb[s], c);
            else {
                const uint32_t id = atomicAdd(counter, 1u);
                nb[id] = cb[s]; ne[id] = ce[s];
                link[s] = int32_t(id);
            }
        }
    }
    nodes[i].left = link[0];
    nodes[i].right = link[1];
    nodes[i].pad[0] = 0; nodes[i].pad[1] = 0;
}


// ---------------------------------------------------------------- binned SAH for the ranges above kSahLeafMax (BDPT_GPU_TREE_SAH=1)
// One thread block per node: centroid bounds, 16 bins on each of the three axes in shared memory, the cheapest of the 45
// planes (A(L) |L| + A(R) |R|), the range's ids partitioned in place through ids_tmp.  The algorithm of the host builder
// (fast_bvh.cu) one level at a time; a node's cost is one block streaming its range three times, so the top levels of a
// 10^7-triangle scene are slow (one block per huge node) -- what this variant is for is tree QUALITY on the device.
constexpr int kSahBlock = 128;
constexpr int kBins = 16;
__device__ __forceinline__ void atomic_min_f(float* a, float v) {
    if (v >= 0.f) atomicMin(reinterpret_cast<int*>(a), __float_as_int(v)); else atomicMax(reinterpret_cast<unsigned*>(a), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float* a, float v) {
    if (v >= 0.f) atomicMax(reinterpret_cast<int*>(a), __float_as_int(v)); else atomicMin(reinterpret_cast<unsigned*>(a), __float_as_uint(v));
}
__device__ __forceinline__ void load_pbox(const PBox* __restrict__ pbox, uint32_t id, float lo[3], float hi[3]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(pbox + id)), c = __ldg(reinterpret_cast<const float4*>(pbox + id) + 1);
    lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = c.x; hi[1] = c.y; hi[2] = c.z;
}

__device__ __forceinline__ void load_box_plain(const PBox* pb, uint32_t k, float lo[3], float hi[3]) {   // pcur is written by this kernel: no read-only path
    const float4 a = reinterpret_cast<const float4*>(pb + k)[0], c = reinterpret_cast<const float4*>(pb + k)[1];
    lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = c.x; hi[1] = c.y; hi[2] = c.z;
}
// boxes into position order (pcur[k] = pbox[ids[k]]): the big-node passes then read their ranges sequentially
__global__ void __launch_bounds__(kBuildBlock) k_gather_boxes(const PBox* __restrict__ pbox, const uint32_t* __restrict__ ids, uint32_t n, PBox* __restrict__ pcur) {
    const uint32_t i = blockIdx.x * kBuildBlock + threadIdx.x;
    if (i >= n) return;
    const float4* s4 = reinterpret_cast<const float4*>(pbox + ids[i]);
    float4* d4 = reinterpret_cast<float4*>(pcur + i);
    d4[0] = __ldg(s4); d4[1] = __ldg(s4 + 1);
}

// NT threads per node; nodes with cnt in (lo_cnt, hi_cnt] are this launch's (two launches per level: big blocks for the few huge
// ranges at the top, small blocks for the many ranges further down).  Each thread walks a CONTIGUOUS chunk of the range and keeps
// a running (bin, count, box) per axis that is flushed to the shared bins only when the bin changes: the range is in Morton order
// (the partition below is stable), neighbours fall into the same bin, and the shared-memory atomics stop being 32-way conflicts.
template <int NT>
__global__ void __launch_bounds__(NT) k_sah_big(PBox* __restrict__ pcur, PBox* __restrict__ ptmp, uint32_t* __restrict__ ids, uint32_t* __restrict__ ids_tmp,
                                                uint32_t* __restrict__ nb, uint32_t* __restrict__ ne, FastNode* __restrict__ nodes,
                                                uint32_t* counter, uint32_t first_node, int level, uint32_t lo_cnt, uint32_t hi_cnt) {
    const uint32_t i = first_node + blockIdx.x;
    const uint32_t b = nb[i], e = ne[i], cnt = e - b;
    if (cnt <= lo_cnt || cnt > hi_cnt) return;
    __shared__ float s_clo[3], s_chi[3];
    __shared__ unsigned s_cnt[3][kBins];
    __shared__ float s_lo[3][kBins][3], s_hi[3][kBins][3];
    __shared__ float s_best_cost[3];
    __shared__ int s_best_plane[3];
    __shared__ int s_axis, s_plane;
    __shared__ unsigned s_scan[NT];
    __shared__ unsigned s_nleft;
    const int t = threadIdx.x;
    if (t < 3) { s_clo[t] = INFINITY; s_chi[t] = -INFINITY; }
    for (int k = t; k < 3 * kBins; k += NT) {
        const int a = k / kBins, bi = k % kBins;
        s_cnt[a][bi] = 0;
        for (int c = 0; c < 3; ++c) { s_lo[a][bi][c] = INFINITY; s_hi[a][bi][c] = -INFINITY; }
    }
    __syncthreads();
    const uint32_t per = (cnt + NT - 1) / NT;
    const uint32_t c0 = min(cnt, uint32_t(t) * per), c1 = min(cnt, c0 + per);
    {   // centroid bounds (twice the centroid: lo + hi)
        float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t k = c0; k < c1; ++k) {
            float lo[3], hi[3];
            load_box_plain(pcur, b + k, lo, hi);
            for (int a = 0; a < 3; ++a) { const float c2 = lo[a] + hi[a]; mn[a] = fminf(mn[a], c2); mx[a] = fmaxf(mx[a], c2); }
        }
        for (int a = 0; a < 3; ++a) { if (mn[a] <= mx[a]) { atomic_min_f(&s_clo[a], mn[a]); atomic_max_f(&s_chi[a], mx[a]); } }
    }
    __syncthreads();
    float scale[3];
    for (int a = 0; a < 3; ++a) { const float ext = s_chi[a] - s_clo[a]; scale[a] = ext > 0.f ? float(kBins) * 0.999999f / ext : 0.f; }
    {   // binning with run-length accumulation per axis
        int cur[3] = {-1, -1, -1};
        unsigned rc[3] = {0, 0, 0};
        float rl[3][3], rh[3][3];
        auto flush = [&](int a) {
            if (cur[a] < 0 || rc[a] == 0) return;
            atomicAdd(&s_cnt[a][cur[a]], rc[a]);
            for (int c = 0; c < 3; ++c) { atomic_min_f(&s_lo[a][cur[a]][c], rl[a][c]); atomic_max_f(&s_hi[a][cur[a]][c], rh[a][c]); }
        };
        for (uint32_t k = c0; k < c1; ++k) {
            float lo[3], hi[3];
            load_box_plain(pcur, b + k, lo, hi);
            for (int a = 0; a < 3; ++a) {
                const int bi = min(kBins - 1, max(0, int(((lo[a] + hi[a]) - s_clo[a]) * scale[a])));
                if (bi != cur[a]) {
                    flush(a);
                    cur[a] = bi; rc[a] = 0;
                    for (int c = 0; c < 3; ++c) { rl[a][c] = INFINITY; rh[a][c] = -INFINITY; }
                }
                ++rc[a];
                for (int c = 0; c < 3; ++c) { rl[a][c] = fminf(rl[a][c], lo[c]); rh[a][c] = fmaxf(rh[a][c], hi[c]); }
            }
        }
        for (int a = 0; a < 3; ++a) flush(a);
    }
    __syncthreads();
    if (t < 3) {   // one axis per thread: suffix boxes, then the sweep
        float best = INFINITY;
        int plane = -1;
        float rarea[kBins];
        unsigned rn[kBins];
        float l3[3] = {INFINITY, INFINITY, INFINITY}, h3[3] = {-INFINITY, -INFINITY, -INFINITY};
        unsigned acc = 0;
        for (int k = kBins - 1; k >= 0; --k) {
            acc += s_cnt[t][k];
            for (int c = 0; c < 3; ++c) { l3[c] = fminf(l3[c], s_lo[t][k][c]); h3[c] = fmaxf(h3[c], s_hi[t][k][c]); }
            rn[k] = acc;
            const float dx = h3[0] - l3[0], dy = h3[1] - l3[1], dz = h3[2] - l3[2];
            rarea[k] = acc ? dx * dy + dy * dz + dz * dx : 0.f;
        }
        for (int c = 0; c < 3; ++c) { l3[c] = INFINITY; h3[c] = -INFINITY; }
        acc = 0;
        for (int k = 0; k + 1 < kBins; ++k) {   // plane between bins k and k + 1
            acc += s_cnt[t][k];
            for (int c = 0; c < 3; ++c) { l3[c] = fminf(l3[c], s_lo[t][k][c]); h3[c] = fmaxf(h3[c], s_hi[t][k][c]); }
            const unsigned nr = rn[k + 1];
            if (acc == 0 || nr == 0) continue;
            const float dx = h3[0] - l3[0], dy = h3[1] - l3[1], dz = h3[2] - l3[2];
            const float cost = (dx * dy + dy * dz + dz * dx) * float(acc) + rarea[k + 1] * float(nr);
            if (cost < best) { best = cost; plane = k + 1; }
        }
        s_best_cost[t] = best; s_best_plane[t] = plane;
    }
    __syncthreads();
    if (t == 0) {
        int axis = -1;
        float best = INFINITY;
        for (int a = 0; a < 3; ++a) if (s_best_plane[a] > 0 && s_best_cost[a] < best) { best = s_best_cost[a]; axis = a; }
        s_axis = axis; s_plane = axis >= 0 ? s_best_plane[axis] : 0;
        if (axis >= 0) {   // depth bound: no child may hold more than 2^(kMaxLevel - level) triangles (as in split_node)
            const uint32_t cap = 1u << (kMaxLevel - level);
            unsigned nl = 0;
            for (int k = 0; k < s_plane; ++k) nl += s_cnt[axis][k];
            if (nl > cap || cnt - nl > cap) s_axis = -1;
        }
    }
    __syncthreads();
    const int axis = s_axis, plane = s_plane;
    uint32_t n_left;
    if (axis < 0) n_left = cnt / 2;   // no plane (coincident centroids) or the depth bound: split the list where it stands
    else {
        unsigned mine = 0;
        for (uint32_t k = c0; k < c1; ++k) {
            float lo[3], hi[3];
            load_box_plain(pcur, b + k, lo, hi);
            const int bi = min(kBins - 1, max(0, int(((lo[axis] + hi[axis]) - s_clo[axis]) * scale[axis])));
            mine += bi < plane ? 1u : 0u;
        }
        s_scan[t] = mine;
        __syncthreads();
        if (t == 0) { unsigned run = 0; for (int k = 0; k < NT; ++k) { const unsigned v = s_scan[k]; s_scan[k] = run; run += v; } s_nleft = run; }
        __syncthreads();
        n_left = s_nleft;
        uint32_t lpos = s_scan[t], rpos = n_left + (c0 - s_scan[t]);
        for (uint32_t k = c0; k < c1; ++k) {   // stable: both sides keep their relative (Morton) order
            const float4 q0 = reinterpret_cast<const float4*>(pcur + b + k)[0], q1 = reinterpret_cast<const float4*>(pcur + b + k)[1];
            const int bi = min(kBins - 1, max(0, int(((axis == 0 ? q0.x + q1.x : (axis == 1 ? q0.y + q1.y : q0.z + q1.z)) - s_clo[axis]) * scale[axis])));
            const uint32_t dst = b + (bi < plane ? lpos++ : rpos++);
            ids_tmp[dst] = ids[b + k];
            reinterpret_cast<float4*>(ptmp + dst)[0] = q0; reinterpret_cast<float4*>(ptmp + dst)[1] = q1;
        }
        __syncthreads();
        for (uint32_t k = b + t; k < e; k += NT) {
            ids[k] = ids_tmp[k];
            reinterpret_cast<float4*>(pcur + k)[0] = reinterpret_cast<const float4*>(ptmp + k)[0];
            reinterpret_cast<float4*>(pcur + k)[1] = reinterpret_cast<const float4*>(ptmp + k)[1];
        }
    }
    if (t == 0) {
        const uint32_t cb[2] = {b, b + n_left}, ce[2] = {b + n_left, e};
        int32_t link[2];
        for (int s = 0; s < 2; ++s) {
            const uint32_t c = ce[s] - cb[s];
            if (c <= 2) link[s] = leaf_code(cb[s], c);
            else {
                const uint32_t id = atomicAdd(counter, 1u);
                nb[id] = cb[s]; ne[id] = ce[s];
                link[s] = int32_t(id);
            }
        }
        nodes[i].left = link[0]; nodes[i].right = link[1];
        nodes[i].pad[0] = 0; nodes[i].pad[1] = 0;
    }
}

__global__ void __launch_bounds__(kBuildBlock) k_boxes(const TriGeom* __restrict__ ftris, FastNode* __restrict__ nodes, double* __restrict__ own,
                                                       uint32_t first_node, uint32_t n_level, double eps) {
    const uint32_t j = blockIdx.x * kBuildBlock + threadIdx.x;
    if (j >= n_level) return;
    const uint32_t i = first_node + j;
    const int32_t link[2] = {nodes[i].left, nodes[i].right};
    double lo[2][3], hi[2][3];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        for (int k = 0; k < 3; ++k) { lo[s][k] = INFINITY; hi[s][k] = -INFINITY; }
        if (link[s] < 0) {
            const uint32_t code = uint32_t(~link[s]);
            const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
            for (uint32_t k = 0; k < cnt; ++k) grow_tri(ftris[first + k], lo[s], hi[s]);
        } else {
            const double* o = own + 6 * size_t(link[s]);
            for (int k = 0; k < 3; ++k) { lo[s][k] = o[k]; hi[s][k] = o[3 + k]; }
        }
    }
    FastNode& nd = nodes[i];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        nd.l_lo[k] = __double2float_rd(lo[0][k] - eps); nd.l_hi[k] = __double2float_ru(hi[0][k] + eps);
        nd.r_lo[k] = __double2float_rd(lo[1][k] - eps); nd.r_hi[k] = __double2float_ru(hi[1][k] + eps);
    }
    double* o = own + 6 * size_t(i);
#pragma unroll
    for (int k = 0; k < 3; ++k) { o[k] = fmin(lo[0][k], lo[1][k]); o[3 + k] = fmax(hi[0][k], hi[1][k]); }
}

struct Scratch {
    std::vector<void*> ptrs;
    ~Scratch() { for (void* p : ptrs) cudaFree(p); }
    template <class T> cudaError_t take(T** p, size_t n) {
        void* q = nullptr;
        const cudaError_t e = cudaMalloc(&q, std::max<size_t>(n * sizeof(T), 16));
        if (e == cudaSuccess) { ptrs.push_back(q); *p = static_cast<T*>(q); }
        return e;
    }
};

}  // namespace

#define GB_CUDA(call)                                                                             \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess) {                                                                  \
            ctx->err = std::string("gpu bvh build: " #call ": ") + cudaGetErrorString(e_);       \
            return e_ == cudaErrorMemoryAllocation ? BDPT_ERR_NOMEM : BDPT_ERR_CUDA;              \
        }                                                                                         \
    } while (0)

// d_tris: the scene's triangles in prim order, already on the device (queued on ctx->stream).  lo/hi: scene
// bounds, extent: their largest side.  On success the three arrays are owned by ctx->scene_allocs.
int build_fast_bvh_gpu(bdpt_ctx* ctx, const TriGeom* d_tris, uint32_t n, const double lo[3], const double hi[3], double extent,
                       const FastNode** out_nodes, const uint32_t** out_prims, const TriGeom** out_tris, uint32_t* out_n_nodes,
                       uint32_t* out_depth) {
    if (n == 0 || n >= (1u << 28)) { ctx->err = "gpu bvh build: triangle count out of range"; return BDPT_ERR_ARG; }
    const bool timing = std::getenv("BDPT_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    cudaStream_t st = ctx->stream;
    Scratch tmp;
    unsigned long long *keys = nullptr, *keys_sorted = nullptr;
    uint32_t *ids = nullptr, *nb = nullptr, *ne = nullptr, *counter = nullptr;
    double* own = nullptr;
    void* sort_tmp = nullptr;
    GB_CUDA(tmp.take(&keys, n));
    GB_CUDA(tmp.take(&keys_sorted, n));
    GB_CUDA(tmp.take(&ids, n));
    GB_CUDA(tmp.take(&nb, n));
    GB_CUDA(tmp.take(&ne, n));
    GB_CUDA(tmp.take(&counter, 1));
    GB_CUDA(tmp.take(&own, 6 * size_t(n)));
    // results (kept): at most n - 1 internal nodes
    FastNode* nodes = nullptr;
    uint32_t* prims = nullptr;
    TriGeom* ftris = nullptr;
    {
        void* p = nullptr;
        GB_CUDA(cudaMalloc(&p, size_t(n) * sizeof(FastNode)));  ctx->scene_allocs.push_back(p); nodes = static_cast<FastNode*>(p);
        GB_CUDA(cudaMalloc(&p, size_t(n) * sizeof(uint32_t)));  ctx->scene_allocs.push_back(p); prims = static_cast<uint32_t*>(p);
        GB_CUDA(cudaMalloc(&p, size_t(n) * sizeof(TriGeom)));   ctx->scene_allocs.push_back(p); ftris = static_cast<TriGeom*>(p);
    }
    const unsigned grid_n = (n + kBuildBlock - 1) / kBuildBlock;
    double sc[3];
    for (int k = 0; k < 3; ++k) sc[k] = hi[k] > lo[k] ? 2097152.0 / (hi[k] - lo[k]) : 0.0;
    k_morton<<<grid_n, kBuildBlock, 0, st>>>(d_tris, n, lo[0], lo[1], lo[2], sc[0], sc[1], sc[2], keys, ids);
    size_t sort_bytes = 0;
    GB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys, keys_sorted, ids, prims, int(n), 0, 63, st));
    GB_CUDA(tmp.take(reinterpret_cast<char**>(&sort_tmp), sort_bytes));
    GB_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, keys, keys_sorted, ids, prims, int(n), 0, 63, st));
    // exact-SAH bottoms: ranges of <= sah_leaf triangles are split by the surface-area sweep over fp32 triangle boxes
    uint32_t sah_leaf = kSahLeafMax;
    if (const char* e = std::getenv("BDPT_GPU_TREE_LBVH")) if (std::atoi(e) != 0) sah_leaf = 0;
    if (const char* e = std::getenv("BDPT_GPU_TREE_SAH_LEAF")) sah_leaf = uint32_t(std::max(0, std::min(kSahLeafMax, std::atoi(e))));
    PBox* pbox = nullptr;
    GB_CUDA(tmp.take(&pbox, n));
    k_prim_boxes<<<grid_n, kBuildBlock, 0, st>>>(d_tris, n, pbox);
    int sah_all = sah_leaf >= 2 ? 1 : 0;   // binned SAH above the exact-sweep ranges as well (k_sah_big); BDPT_GPU_TREE_SAH=0: Morton splits there
    if (const char* e = std::getenv("BDPT_GPU_TREE_SAH")) sah_all = std::atoi(e) != 0 && sah_leaf >= 2 ? 1 : 0;
    uint32_t* ids_tmp = nullptr;
    PBox *pcur = nullptr, *ptmp = nullptr;
    if (sah_all) {
        GB_CUDA(tmp.take(&ids_tmp, n));
        GB_CUDA(tmp.take(&pcur, n));
        GB_CUDA(tmp.take(&ptmp, n));
        k_gather_boxes<<<grid_n, kBuildBlock, 0, st>>>(pbox, prims, n, pcur);
    }
    // ---- top-down
    const uint32_t root_range[2] = {0u, n};
    const uint32_t one = 1u;
    GB_CUDA(cudaMemcpyAsync(nb, &root_range[0], 4, cudaMemcpyHostToDevice, st));
    GB_CUDA(cudaMemcpyAsync(ne, &root_range[1], 4, cudaMemcpyHostToDevice, st));
    GB_CUDA(cudaMemcpyAsync(counter, &one, 4, cudaMemcpyHostToDevice, st));
    if ((uint64_t(n) + 1) / 2 > (1ull << kMaxLevel)) { ctx->err = "gpu bvh build: too many triangles for the depth bound"; return BDPT_ERR_ARG; }
    std::vector<uint32_t> level_first{0u};
    uint32_t total = 1;
    for (int level = 0;; ++level) {
        const uint32_t first = level_first.back(), cnt = total - first;
        if (cnt == 0) { level_first.pop_back(); break; }
        if (level > kMaxLevel) { ctx->err = "gpu bvh build: depth bound violated"; return BDPT_ERR_CUDA; }
        if (sah_all) {
            constexpr uint32_t kHuge = 8192;   // ranges above this get 1024 threads, the others 128
            if (level <= 12) k_sah_big<1024><<<cnt, 1024, 0, st>>>(pcur, ptmp, prims, ids_tmp, nb, ne, nodes, counter, first, level, kHuge, 0xFFFFFFFFu);
            k_sah_big<kSahBlock><<<cnt, kSahBlock, 0, st>>>(pcur, ptmp, prims, ids_tmp, nb, ne, nodes, counter, first, level, sah_leaf, level <= 12 ? kHuge : 0xFFFFFFFFu);
        }
        k_split<<<(cnt + kBuildBlock - 1) / kBuildBlock, kBuildBlock, 0, st>>>(keys_sorted, nb, ne, nodes, counter, first, cnt, level, prims, pbox, sah_leaf, sah_all);
        uint32_t now = 0;
        GB_CUDA(cudaMemcpyAsync(&now, counter, 4, cudaMemcpyDeviceToHost, st));
        GB_CUDA(cudaStreamSynchronize(st));
        level_first.push_back(total);
        total = now;
    }
    // the triangles in their final order (the SAH bottoms re-ordered the ids inside their ranges)
    k_gather<<<grid_n, kBuildBlock, 0, st>>>(d_tris, prims, n, ftris);
    // ---- bottom-up
    const double eps = std::max(extent, 1e-30) * (1.0 / 65536.0);   // the host builder's outward padding
    for (size_t l = level_first.size(); l-- > 0;) {
        const uint32_t first = level_first[l];
        const uint32_t cnt = (l + 1 < level_first.size() ? level_first[l + 1] : total) - first;
        k_boxes<<<(cnt + kBuildBlock - 1) / kBuildBlock, kBuildBlock, 0, st>>>(ftris, nodes, own, first, cnt, eps);
    }
    GB_CUDA(cudaGetLastError());
    GB_CUDA(cudaStreamSynchronize(st));
    *out_nodes = nodes; *out_prims = prims; *out_tris = ftris;
    *out_n_nodes = total;
    *out_depth = uint32_t(level_first.size()) + 1;   // internal levels + the leaves below them
    if (timing)
        std::fprintf(stderr, "  fast tree on the GPU: %u nodes, %zu levels, %.1f ms\n", total, level_first.size(),
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    return BDPT_OK;
}

}  // namespace bdpt
