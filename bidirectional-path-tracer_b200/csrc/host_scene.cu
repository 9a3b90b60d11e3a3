// Host-side scene flattening and the reference-topology BVH builder.
// Replaces: triangle ctor (src/objects/primatives/triangle.h:21-39), triangle_collection light list
// (triangle.h:135-230, scene_loader.h:190-202), bvh_node ctor (src/acceleration/bvh.h:20-48) and the
// aabb merge / pad rules (src/acceleration/aabb.h:27-33,81-88, src/core/interval.h:12-16,36-39).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <limits>
#include <map>
#include <stdexcept>
#include <thread>

#include "host_scene.h"

namespace bdpt {

namespace {

struct Box6 { double lo[3], hi[3]; };

inline void pad_axis(double& lo, double& hi) {  // aabb.h:81-88 via interval.h:36-39
    const double delta = 0.0001;
    if (hi - lo < delta) { const double p = delta / 2; lo = lo - p; hi = hi + p; }
}

inline Box6 merge_pad(const Box6& a, const double* b /* lo xyz hi xyz */) {  // aabb.h:27-33
    Box6 r;
    for (int k = 0; k < 3; ++k) {
        r.lo[k] = a.lo[k] <= b[k] ? a.lo[k] : b[k];
        r.hi[k] = a.hi[k] >= b[3 + k] ? a.hi[k] : b[3 + k];
        pad_axis(r.lo[k], r.hi[k]);
    }
    return r;
}

inline int longest_axis(const Box6& b) {  // aabb.h:68-75 (ties go to the later axis)
    const double sx = b.hi[0] - b.lo[0], sy = b.hi[1] - b.lo[1], sz = b.hi[2] - b.lo[2];
    if (sx > sy) return sx > sz ? 0 : 2;
    return sy > sz ? 1 : 2;
}

// Node count of the reference tree over a span: f(1) = f(2) = 1, f(n) = 1 + f(n/2) + f(n - n/2) (bvh.h:35-46).
// It depends on the span's size only, so every node's preorder index is known before anything is built and
// subtrees can be built by different threads straight into their final places.
struct NodeCount {
    std::map<size_t, size_t> memo;
    size_t operator()(size_t n) {
        if (n <= 2) return 1;
        auto it = memo.find(n);
        if (it != memo.end()) return it->second;
        const size_t v = 1 + (*this)(n / 2) + (*this)(n - n / 2);
        memo[n] = v;
        return v;
    }
};

// std::sort with its top quicksort levels spread over threads.  It calls libstdc++'s own building blocks in the
// order std::__sort does (bits/stl_algo.h: __introsort_loop = partition around the median-of-three pivot, recurse
// into the right part, loop on the left; then one __final_insertion_sort over everything).  The two parts of a
// partition are disjoint and sorted independently of each other, so handing the right part to another thread
// changes nothing about the moves made: the permutation is exactly std::sort's, ties included.
template <class It, class Cmp>
void introsort_loop_forked(It first, It last, long depth_limit, Cmp comp, int forks) {
    std::vector<std::thread> pool;
    while (last - first > 16) {   // _S_threshold
        if (depth_limit == 0) { std::__partial_sort(first, last, last, comp); break; }
        --depth_limit;
        It cut = std::__unguarded_partition_pivot(first, last, comp);
        if (forks > 0 && last - cut > 4096) {
            --forks;
            pool.emplace_back([=] { introsort_loop_forked(cut, last, depth_limit, comp, forks); });
        } else {
            std::__introsort_loop(cut, last, depth_limit, comp);
        }
        last = cut;
    }
    for (auto& th : pool) th.join();
}

template <class It, class Less>
void sort_like_std(It first, It last, Less less, int forks) {
    if (first == last) return;
    auto comp = __gnu_cxx::__ops::__iter_comp_iter(less);
    introsort_loop_forked(first, last, long(std::__lg(last - first)) * 2, comp, forks);
    std::__final_insertion_sort(first, last, comp);
}

struct RefBuilder {
    HostScene& s;
    std::vector<uint32_t>& order;
    std::atomic<uint32_t> max_depth{0};
    std::map<size_t, size_t> counts;   // filled before the build, read-only during it
    explicit RefBuilder(HostScene& sc) : s(sc), order(sc.leaf_order) {}

    size_t count(size_t n) const { return n <= 2 ? 1 : counts.at(n); }

    void build(size_t start, size_t end, uint32_t depth, size_t me, int fork_levels) {
        uint32_t seen = max_depth.load(std::memory_order_relaxed);
        while (depth > seen && !max_depth.compare_exchange_weak(seen, depth, std::memory_order_relaxed)) {}
        const double inf = std::numeric_limits<double>::infinity();
        Box6 bb;  // aabb::empty (aabb.h:92): padding (+inf,-inf) leaves it unchanged (size = -inf < delta
                  // gives (+inf - p, -inf + p) = (+inf, -inf))
        for (int k = 0; k < 3; ++k) { bb.lo[k] = inf; bb.hi[k] = -inf; }
        for (size_t i = start; i < end; ++i) bb = merge_pad(bb, &s.tri_box[6 * size_t(order[i])]);
        const int axis = longest_axis(bb);
        const size_t span = end - start;
        RefNode n;
        n.lo_x = bb.lo[0]; n.hi_x = bb.hi[0]; n.lo_y = bb.lo[1]; n.hi_y = bb.hi[1];
        n.lo_z = bb.lo[2]; n.hi_z = bb.hi[2];
        n.pad = 0;
        if (span == 1) {
            n.leaf = 1; n.left = n.right = int32_t(order[start]);
        } else if (span == 2) {
            n.leaf = 1; n.left = int32_t(order[start]); n.right = int32_t(order[start + 1]);
        } else {
            // bvh.h:41,68-74: libstdc++ std::sort (unstable) on the box minimum along `axis`; the same
            // library, comparator results and initial order give the reference's permutation.
            // Sorted as (key, prim) pairs: std::sort's moves depend only on the comparison outcomes, so the
            // permutation equals that of sorting the pointers, without a cache miss per comparison.
            const double* tb = s.tri_box.data();
            struct KeyId { double key; uint32_t id; };
            KeyId few[128];                 // most nodes are small: no heap allocation for them (92 k nodes for 180 k triangles)
            std::vector<KeyId> many;
            KeyId* tmp = few;
            if (span > 128) { many.resize(span); tmp = many.data(); }
            for (size_t i = 0; i < span; ++i) { const uint32_t p = order[start + i]; tmp[i] = KeyId{tb[6 * size_t(p) + axis], p}; }
            auto by_key = [](const KeyId& a, const KeyId& b) { return a.key < b.key; };
            if (fork_levels > 0 && span > 32768) sort_like_std(tmp, tmp + span, by_key, fork_levels);
            else std::sort(tmp, tmp + span, by_key);   // the same introsort over pointers as over vector iterators: the same permutation
            for (size_t i = 0; i < span; ++i) order[start + i] = tmp[i].id;
            const size_t mid = start + span / 2;
            const size_t left = me + 1, right = me + 1 + count(mid - start);   // preorder
            n.leaf = 0;
            n.left = int32_t(left);
            n.right = int32_t(right);
            if (fork_levels > 0 && span > 4096) {
                std::thread th([=] { build(start, mid, depth + 1, left, fork_levels - 1); });
                build(mid, end, depth + 1, right, fork_levels - 1);
                th.join();
            } else {
                build(start, mid, depth + 1, left, 0);
                build(mid, end, depth + 1, right, 0);
            }
        }
        s.ref_nodes[me] = n;
    }
};

}  // namespace

void flatten_scene(const double* verts, const uint32_t* mat_id, uint64_t n, const bdpt_material* mats,
                   uint32_t n_mats, HostScene& out) {
    if (!verts || !mat_id || !mats || n == 0 || n_mats == 0) throw std::runtime_error("empty scene");
    if (n >= 0x7FFFFFFFull) throw std::runtime_error("too many triangles");
    out = HostScene();
    out.mats.resize(n_mats);
    for (uint32_t i = 0; i < n_mats; ++i) {
        if (mats[i].type > BDPT_MAT_LIGHT) throw std::runtime_error("unknown material type");
        Mat m;
        m.r = mats[i].rgb[0]; m.g = mats[i].rgb[1]; m.b = mats[i].rgb[2];
        m.param = mats[i].param;
        if (mats[i].type == BDPT_MAT_METAL && m.param > 1.0) m.param = 1.0;  // material.h:71
        m.type = mats[i].type;
        m.tex = mats[i].tex;
        m.pad[0] = m.pad[1] = 0;
        out.mats[i] = m;
    }
    out.tris.resize(n); out.shade.resize(n); out.area.resize(n); out.tri_box.resize(6 * n);
    // triangles are independent: flattened on a few threads (the light list, which is ordered, afterwards)
    std::atomic<bool> bad_mat{false};
    auto flatten_range = [&](uint64_t i0, uint64_t i1) {
        for (uint64_t i = i0; i < i1; ++i) {
            if (mat_id[i] >= n_mats) { bad_mat = true; return; }
            const double* p = verts + 9 * i;
            const D3 v0 = mk(p[0], p[1], p[2]), v1 = mk(p[3], p[4], p[5]), v2 = mk(p[6], p[7], p[8]);
            const D3 e1 = v1 - v0, e2 = v2 - v0;
            const D3 nn = cross(e1, e2);
            const D3 un = unit(nn);                      // triangle.h:27 (NaN for degenerate triangles, as there)
            TriGeom g;
            g.v0x = v0.x; g.v0y = v0.y; g.v0z = v0.z; g.e1x = e1.x; g.e1y = e1.y; g.e1z = e1.z;
            g.e2x = e2.x; g.e2y = e2.y; g.e2z = e2.z; g.mat = mat_id[i]; g.pad = 0;
            out.tris[i] = g;
            TriShade sh;
            sh.nx = un.x; sh.ny = un.y; sh.nz = un.z; sh.mat = mat_id[i]; sh.pad = 0;
            out.shade[i] = sh;
            out.area[i] = 0.5 * len(nn);
            double* b = &out.tri_box[6 * i];
            b[0] = std::fmin(v0.x, std::fmin(v1.x, v2.x)); b[3] = std::fmax(v0.x, std::fmax(v1.x, v2.x));
            b[1] = std::fmin(v0.y, std::fmin(v1.y, v2.y)); b[4] = std::fmax(v0.y, std::fmax(v1.y, v2.y));
            b[2] = std::fmin(v0.z, std::fmin(v1.z, v2.z)); b[5] = std::fmax(v0.z, std::fmax(v1.z, v2.z));
        }
    };
    {
        const unsigned nt = n < 65536 ? 1u : std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
        const uint64_t per = (n + nt - 1) / nt;
        std::vector<std::thread> pool;
        for (unsigned k = 1; k < nt; ++k) pool.emplace_back(flatten_range, std::min(n, k * per), std::min(n, (k + 1) * per));
        flatten_range(0, std::min(n, per));
        for (auto& th : pool) th.join();
    }
    if (bad_mat) throw std::runtime_error("material index out of range");
    for (uint64_t i = 0; i < n; ++i)
        if (out.mats[mat_id[i]].type == BDPT_MAT_LIGHT) out.lights.push_back(uint32_t(i));
    if (out.lights.empty())  // main.cpp:67: no lights -> the whole world is the light list
        for (uint64_t i = 0; i < n; ++i) out.lights.push_back(uint32_t(i));
    out.light_area.resize(out.lights.size());
    double total = 0.0;  // triangle.h:200-203 sums in list order
    for (size_t k = 0; k < out.lights.size(); ++k) {
        total += out.area[out.lights[k]];
        out.light_area[k] = total;   // running sum = the `accum` of the reference's linear scan
    }
    out.light_total_area = total;
}

void build_ref_bvh(HostScene& s) {
    const size_t n = s.tris.size();
    s.leaf_order.resize(n);
    for (size_t i = 0; i < n; ++i) s.leaf_order[i] = uint32_t(i);
    RefBuilder b(s);
    NodeCount nc;
    s.ref_nodes.assign(nc(n), RefNode());
    b.counts = nc.memo;
    const bool timing = std::getenv("BDPT_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (timing) std::fprintf(stderr, "  reference tree: %s at %.1f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    };
    static const int fork_levels = [] { const char* e = std::getenv("BDPT_REF_FORKS"); return e ? std::atoi(e) : 4; }();
    b.build(0, n, 0, 0, fork_levels);   // up to 2^fork_levels subtrees in flight (4: 16)
    lap("nodes built");
    s.ref_depth = b.max_depth.load() + 1;
    s.dfs_rank.resize(n);
    for (size_t k = 0; k < n; ++k) s.dfs_rank[s.leaf_order[k]] = uint32_t(k);
    s.ref_parent.assign(s.ref_nodes.size(), -1);
    s.prim_leaf.assign(n, -1);
    for (size_t i = 0; i < s.ref_nodes.size(); ++i) {
        const RefNode& nd = s.ref_nodes[i];
        if (nd.leaf) { s.prim_leaf[nd.left] = int32_t(i); s.prim_leaf[nd.right] = int32_t(i); }
        else { s.ref_parent[nd.left] = int32_t(i); s.ref_parent[nd.right] = int32_t(i); }
    }
    // Intersection of the chain's boxes, top-down (preorder: a parent precedes its children).  A ray that
    // passes aabb::hit on this box passes it on every box of the chain (each slab bound of the intersection
    // is the tightest of the chain's, and IEEE rounding is monotone), so traversal can accept "the reference
    // reaches this triangle" with one box test; a failure proves nothing (child boxes are not nested in
    // their parents where pad_to_minimums, aabb.h:81-88, widened a thin child) and falls back to the chain.
    lap("rank and parent tables");
    s.chain_box.assign(6 * n, 0.0);
    struct Chain {
        HostScene& s;
        bool fork;
        // node i under the running intersection `up` of its ancestors' boxes; the right subtrees of the first
        // four levels go to their own threads
        void go(int32_t i, const double* up, int depth) {
            const RefNode& nd = s.ref_nodes[size_t(i)];
            double a[6] = {nd.lo_x, nd.hi_x, nd.lo_y, nd.hi_y, nd.lo_z, nd.hi_z};
            if (up)
                for (int k = 0; k < 6; k += 2) { a[k] = std::fmax(a[k], up[k]); a[k + 1] = std::fmin(a[k + 1], up[k + 1]); }
            if (nd.leaf) {
                std::copy(a, a + 6, &s.chain_box[6 * size_t(nd.left)]);
                std::copy(a, a + 6, &s.chain_box[6 * size_t(nd.right)]);
                return;
            }
            if (fork && depth < 4) {
                std::thread right([&] { go(nd.right, a, depth + 1); });
                go(nd.left, a, depth + 1);
                right.join();
            } else {
                go(nd.left, a, depth + 1);
                go(nd.right, a, depth + 1);
            }
        }
    };
    Chain{s, n >= 65536}.go(0, nullptr, 0);
    lap("chain boxes");
}

}  // namespace bdpt
