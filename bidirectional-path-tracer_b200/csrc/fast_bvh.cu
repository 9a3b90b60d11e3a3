// Binned-SAH builder for the fast BVH (host).  Not a reference behaviour: the reference only has the
// median split of src/acceleration/bvh.h:20-48 (kept bit-exactly in host_scene.cu); this tree exists so
// that traversal visits fewer nodes while returning the same hits (traverse_fast.cuh).
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <thread>
#include <vector>
#include <cmath>
#include <limits>

#include "host_scene.h"

namespace bdpt {

namespace {

constexpr int kBins = 16;
constexpr uint32_t kMaxLeaf = 2;
constexpr double kTravCost = 1.0, kTriCost = 1.5;

struct B3 {
    double lo[3], hi[3];
    void reset() {
        for (int k = 0; k < 3; ++k) { lo[k] = std::numeric_limits<double>::infinity(); hi[k] = -lo[k]; }
    }
    void grow(const double* b) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b[k]); hi[k] = std::max(hi[k], b[3 + k]); }
    }
    void grow(const B3& o) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], o.lo[k]); hi[k] = std::max(hi[k], o.hi[k]); }
    }
    double half_area() const {
        double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0)) return 0.0;
        return dx * dy + dy * dz + dz * dx;
    }
};

inline float down32(double v) {
    float f = float(v);
    if (double(f) > v) f = std::nextafterf(f, -INFINITY);
    return f;
}
inline float up32(double v) {
    float f = float(v);
    if (double(f) < v) f = std::nextafterf(f, INFINITY);
    return f;
}

struct Task { uint32_t begin, end, depth; int32_t node; int side; };

// fn(chunk_begin, chunk_end, chunk_index) over [b, e) on up to `threads` threads (the big nodes at the top of the
// tree are built by one task: their linear passes are what is worth splitting)
template <class F>
void par_chunks(uint32_t b, uint32_t e, unsigned threads, F fn) {
    const uint32_t n = e - b;
    if (threads <= 1 || n < 32768) { fn(b, e, 0u); return; }
    std::vector<std::thread> pool;
    const uint32_t per = (n + threads - 1) / threads;
    for (unsigned k = 1; k < threads; ++k) {
        const uint32_t cb = b + std::min(n, k * per), ce = b + std::min(n, (k + 1) * per);
        if (cb < ce) pool.emplace_back([=] { fn(cb, ce, k); });
    }
    fn(b, b + std::min(n, per), 0u);
    for (auto& th : pool) th.join();
}
constexpr unsigned kTopThreads = 8;

}  // namespace

// bounds of all triangles; sets s.extent (largest side) and s.center
void scene_bounds(HostScene& s, double lo[3], double hi[3]) {
    const uint32_t n = uint32_t(s.tris.size());
    const double* tb = s.tri_box.data();
    B3 scene;
    scene.reset();
    for (uint32_t i = 0; i < n; ++i) scene.grow(tb + 6 * size_t(i));
    double ext = 0.0;
    for (int k = 0; k < 3; ++k) ext = std::max(ext, scene.hi[k] - scene.lo[k]);
    s.extent = ext;
    for (int k = 0; k < 3; ++k) { s.center[k] = 0.5 * (scene.lo[k] + scene.hi[k]); lo[k] = scene.lo[k]; hi[k] = scene.hi[k]; }
}

void build_fast_bvh(HostScene& s) {
    const uint32_t n = uint32_t(s.tris.size());
    const double* tb = s.tri_box.data();
    s.fast_prims.resize(n);
    for (uint32_t i = 0; i < n; ++i) s.fast_prims[i] = i;
    s.fast_nodes.clear();
    s.fast_nodes.reserve(n / 2 + 4);
    double slo[3], shi[3];
    scene_bounds(s, slo, shi);
    const double ext = s.extent;
    // outward padding: far above fp64 slab/triangle rounding noise, far below feature size
    const double eps = std::max(ext, 1e-30) * (1.0 / 65536.0);

    auto range_box = [&](uint32_t b, uint32_t e) {
        B3 r; r.reset();
        for (uint32_t i = b; i < e; ++i) r.grow(tb + 6 * size_t(s.fast_prims[i]));
        return r;
    };
    auto store_box = [&](FastNode& nd, int side, const B3& b) {
        float* lo = side ? nd.r_lo : nd.l_lo;
        float* hi = side ? nd.r_hi : nd.l_hi;
        for (int k = 0; k < 3; ++k) { lo[k] = down32(b.lo[k] - eps); hi[k] = up32(b.hi[k] + eps); }
    };
    auto leaf_code = [](uint32_t first, uint32_t cnt) { return int32_t(~((first << 3) | (cnt - 1))); };

    // One (sub)tree: tasks are expanded depth-first into `nodes` (node 0 = the task seeded with node -1).
    // `stop_at` > 0 makes it breadth-first instead and stops as soon as that many tasks are pending, which is
    // how the top of the tree is split into independent subtrees for the worker threads.
    auto grow = [&](std::vector<FastNode>& nodes, std::deque<Task>& todo, uint32_t& max_depth, size_t stop_at) {
    while (!todo.empty() && (stop_at == 0 || todo.size() < stop_at)) {
        Task t;
        if (stop_at) { t = todo.front(); todo.pop_front(); } else { t = todo.back(); todo.pop_back(); }
        const uint32_t cnt = t.end - t.begin;
        max_depth = std::max(max_depth, t.depth);
        // leaf?  (never for the root task: it must produce node 0)
        if (t.node >= 0 && cnt <= kMaxLeaf) {
            // decided by the parent (it stored the box); just encode
            FastNode& p = nodes[t.node];
            (t.side ? p.right : p.left) = leaf_code(t.begin, cnt);
            continue;
        }
        int32_t me;
        if (t.node < 0) me = 0;
        else {
            me = int32_t(nodes.size());
            nodes.push_back(FastNode());
            FastNode& p = nodes[t.node];
            (t.side ? p.right : p.left) = me;
        }
        uint32_t mid = t.begin;
        if (cnt == 1) {
            // degenerate root: duplicate the single triangle as both children
            FastNode nd{};
            const B3 bb = range_box(t.begin, t.end);
            store_box(nd, 0, bb); store_box(nd, 1, bb);
            nd.left = leaf_code(t.begin, 1); nd.right = leaf_code(t.begin, 1);
            nodes[me] = nd;
            continue;
        }
        // centroid bounds
        double clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        const bool fork = stop_at && cnt >= 32768;   // only the single-threaded top phase may fork
        if (!fork) {
            for (uint32_t i = t.begin; i < t.end; ++i) {
                const double* b = tb + 6 * size_t(s.fast_prims[i]);
                for (int k = 0; k < 3; ++k) {
                    double c = 0.5 * (b[k] + b[3 + k]);
                    clo[k] = std::min(clo[k], c); chi[k] = std::max(chi[k], c);
                }
            }
        } else {
            struct alignas(64) CB { double lo[3], hi[3]; } part[kTopThreads];
            for (auto& x : part) for (int k = 0; k < 3; ++k) { x.lo[k] = INFINITY; x.hi[k] = -INFINITY; }
            par_chunks(t.begin, t.end, kTopThreads, [&](uint32_t cb, uint32_t ce, unsigned q) {
                double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
                for (uint32_t i = cb; i < ce; ++i) {
                    const double* b = tb + 6 * size_t(s.fast_prims[i]);
                    for (int k = 0; k < 3; ++k) {
                        double c = 0.5 * (b[k] + b[3 + k]);
                        lo[k] = std::min(lo[k], c); hi[k] = std::max(hi[k], c);
                    }
                }
                for (int k = 0; k < 3; ++k) { part[q].lo[k] = lo[k]; part[q].hi[k] = hi[k]; }
            });
            for (auto& x : part) for (int k = 0; k < 3; ++k) { clo[k] = std::min(clo[k], x.lo[k]); chi[k] = std::max(chi[k], x.hi[k]); }
        }
        int axis = 0;
        for (int k = 1; k < 3; ++k) if (chi[k] - clo[k] > chi[axis] - clo[axis]) axis = k;
        const double cext = chi[axis] - clo[axis];
        bool split_ok = false;
        B3 lbox, rbox;   // child boxes of a SAH split: unions of its bins (what range_box would recompute)
        if (cext > 0 && t.depth < 22 && cnt > 2) {
            // binned SAH over all three axes
            double best_cost = std::numeric_limits<double>::infinity();
            int best_axis = -1, best_bin = -1, best_nb = kBins;
            B3 best_bins[kBins];
            // few triangles need few bins: the sweeps below cost per bin, and most nodes of a tree are small
            const int nb = int(std::min<uint32_t>(kBins, std::max<uint32_t>(4u, 2u * cnt)));
            best_nb = nb;
            // one pass over the triangles fills the bins of all three axes
            struct alignas(64) PB { B3 b[3][kBins]; uint32_t c[3][kBins]; };
            double sc3[3];
            for (int ax = 0; ax < 3; ++ax) { const double e = chi[ax] - clo[ax]; sc3[ax] = e > 0 ? nb / e : 0.0; }
            auto fill = [&](PB& pb, uint32_t cb, uint32_t ce) {
                for (int ax = 0; ax < 3; ++ax) for (int k = 0; k < nb; ++k) { pb.b[ax][k].reset(); pb.c[ax][k] = 0; }
                for (uint32_t i = cb; i < ce; ++i) {
                    const double* bx = tb + 6 * size_t(s.fast_prims[i]);
                    for (int ax = 0; ax < 3; ++ax) {
                        int bi = int((0.5 * (bx[ax] + bx[3 + ax]) - clo[ax]) * sc3[ax]);
                        bi = std::min(std::max(bi, 0), nb - 1);
                        pb.b[ax][bi].grow(bx); ++pb.c[ax][bi];
                    }
                }
            };
            PB all;
            if (!fork) fill(all, t.begin, t.end);
            else {
                std::vector<PB> part(kTopThreads);
                for (auto& x : part) fill(x, 0, 0);
                par_chunks(t.begin, t.end, kTopThreads, [&](uint32_t cb, uint32_t ce, unsigned q) { fill(part[q], cb, ce); });
                fill(all, 0, 0);
                for (auto& x : part)
                    for (int ax = 0; ax < 3; ++ax)
                        for (int k = 0; k < nb; ++k) { all.b[ax][k].grow(x.b[ax][k]); all.c[ax][k] += x.c[ax][k]; }
            }
            for (int ax = 0; ax < 3; ++ax) {
                if (!(chi[ax] - clo[ax] > 0)) continue;
                const B3* bins = all.b[ax]; const uint32_t* bc = all.c[ax];
                double ra[kBins]; uint32_t rc[kBins];
                B3 acc; acc.reset(); uint32_t c = 0;
                for (int k = nb - 1; k > 0; --k) { acc.grow(bins[k]); c += bc[k]; ra[k] = acc.half_area(); rc[k] = c; }
                acc.reset(); c = 0;
                for (int k = 0; k < nb - 1; ++k) {
                    acc.grow(bins[k]); c += bc[k];
                    if (c == 0 || rc[k + 1] == 0) continue;
                    double cost = acc.half_area() * c + ra[k + 1] * rc[k + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = k; }
                }
            }
            if (best_axis >= 0) for (int k = 0; k < nb; ++k) best_bins[k] = all.b[best_axis][k];
            if (best_axis >= 0) {
                const double sc = sc3[best_axis];
                const double lo = clo[best_axis];
                auto it = std::partition(s.fast_prims.begin() + t.begin, s.fast_prims.begin() + t.end, [&](uint32_t p) {
                    const double* b = tb + 6 * size_t(p);
                    int bi = int((0.5 * (b[best_axis] + b[3 + best_axis]) - lo) * sc);
                    bi = std::min(std::max(bi, 0), best_nb - 1);
                    return bi <= best_bin;
                });
                mid = uint32_t(it - s.fast_prims.begin());
                split_ok = mid > t.begin && mid < t.end;
                lbox.reset(); rbox.reset();
                for (int k = 0; k < best_nb; ++k) (k <= best_bin ? lbox : rbox).grow(best_bins[k]);
            }
        }
        if (!split_ok) {
            mid = t.begin + cnt / 2;
            std::nth_element(s.fast_prims.begin() + t.begin, s.fast_prims.begin() + mid, s.fast_prims.begin() + t.end,
                             [&](uint32_t a, uint32_t b) {
                                 return tb[6 * size_t(a) + axis] + tb[6 * size_t(a) + 3 + axis] <
                                        tb[6 * size_t(b) + axis] + tb[6 * size_t(b) + 3 + axis];
                             });
        }
        FastNode nd{};
        if (!split_ok) { lbox = range_box(t.begin, mid); rbox = range_box(mid, t.end); }
        store_box(nd, 0, lbox);
        store_box(nd, 1, rbox);
        nd.left = nd.right = 0;
        nodes[me] = nd;
        // children: push right first so the left subtree is laid out next (preorder, near the parent)
        todo.push_back(Task{mid, t.end, t.depth + 1, me, 1});
        todo.push_back(Task{t.begin, mid, t.depth + 1, me, 0});
    }
    };

    const bool timing = std::getenv("BDPT_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (timing) std::fprintf(stderr, "  fast tree: %s at %.1f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    };
    // top of the tree on this thread until there are enough independent subtrees ...
    uint32_t max_depth = 0;
    std::deque<Task> top;
    s.fast_nodes.push_back(FastNode());   // root: always an internal node so the kernel never special-cases
    top.push_back(Task{0, n, 0, -1, 0});
    grow(s.fast_nodes, top, max_depth, 32);
    lap("top levels done");
    // ... then one subtree per task, on worker threads, each into its own node array (prim ranges are disjoint)
    std::vector<Task> subs(top.begin(), top.end());
    std::sort(subs.begin(), subs.end(), [](const Task& a, const Task& b) { return a.end - a.begin > b.end - b.begin; });
    std::vector<std::vector<FastNode>> local(subs.size());
    std::vector<uint32_t> local_depth(subs.size(), 0);
    std::atomic<size_t> next{0};
    auto worker = [&] {
        for (size_t i = next.fetch_add(1); i < subs.size(); i = next.fetch_add(1)) {
            const Task& t = subs[i];
            if (t.end - t.begin <= kMaxLeaf) continue;   // a leaf of the top part: encoded below
            std::deque<Task> todo;
            local[i].push_back(FastNode());
            todo.push_back(Task{t.begin, t.end, t.depth, -1, 0});
            grow(local[i], todo, local_depth[i], 0);
        }
    };
    const unsigned n_threads = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    std::vector<std::thread> pool;
    for (unsigned k = 1; k < n_threads && k < subs.size(); ++k) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
    lap("subtrees done");
    // splice: a subtree's internal links move by its base index; leaf codes (negative) are global already
    std::vector<int32_t> base(subs.size(), 0);
    {
        size_t total_nodes = s.fast_nodes.size();
        for (size_t i = 0; i < subs.size(); ++i) { base[i] = int32_t(total_nodes); total_nodes += local[i].size(); }
        s.fast_nodes.resize(total_nodes);
    }
    for (size_t i = 0; i < subs.size(); ++i) {
        const Task& t = subs[i];
        FastNode& p = s.fast_nodes[t.node];
        max_depth = std::max(max_depth, std::max(t.depth, local_depth[i]));
        (t.side ? p.right : p.left) = local[i].empty() ? leaf_code(t.begin, t.end - t.begin) : base[i];
    }
    std::atomic<size_t> next_copy{0};
    auto copier = [&] {
        for (size_t i = next_copy.fetch_add(1); i < subs.size(); i = next_copy.fetch_add(1)) {
            FastNode* dst = s.fast_nodes.data() + base[i];
            for (FastNode nd : local[i]) {
                if (nd.left >= 0) nd.left += base[i];
                if (nd.right >= 0) nd.right += base[i];
                *dst++ = nd;
            }
        }
    };
    pool.clear();
    for (unsigned k = 1; k < n_threads && k < subs.size(); ++k) pool.emplace_back(copier);
    copier();
    for (auto& th : pool) th.join();
    (void)kTravCost; (void)kTriCost;
    s.fast_depth = max_depth + 1;
    lap("spliced");
    s.fast_tris.resize(n);
    par_chunks(0, n, n_threads, [&](uint32_t cb, uint32_t ce, unsigned) {
        for (uint32_t k = cb; k < ce; ++k) s.fast_tris[k] = s.tris[s.fast_prims[k]];
    });
    // the 4-wide form is an option of bdpt_trace (BDPT_WIDE, ctx.h): built only when asked for
    s.wide_nodes.clear();
    s.wide_depth = 0;
    if (const char* e = std::getenv("BDPT_WIDE")) {
        if (std::atoi(e) & 3) { collapse_wide(s); lap("collapsed to 4-wide"); }
    }
}

// Binary -> 4-wide: a wide node starts from a binary node's two children and, while it has a free slot, replaces
// the internal child of largest surface area by that child's two children.  Boxes are copied, never recomputed,
// so the wide tree is exactly as conservative as the binary one; leaves are the same leaf codes.
void collapse_wide(HostScene& s) {
    struct Slot { int32_t link; float lo[3], hi[3]; };
    auto slots_of = [&](int32_t b, Slot out[2]) {
        const FastNode& nd = s.fast_nodes[size_t(b)];
        out[0].link = nd.left; out[1].link = nd.right;
        for (int k = 0; k < 3; ++k) { out[0].lo[k] = nd.l_lo[k]; out[0].hi[k] = nd.l_hi[k]; out[1].lo[k] = nd.r_lo[k]; out[1].hi[k] = nd.r_hi[k]; }
    };
    auto area = [](const Slot& c) {
        const double dx = double(c.hi[0]) - c.lo[0], dy = double(c.hi[1]) - c.lo[1], dz = double(c.hi[2]) - c.lo[2];
        return dx * dy + dy * dz + dz * dx;
    };
    struct Item { int32_t bin, wide; uint32_t depth; };
    s.wide_nodes.clear();
    s.wide_nodes.reserve(s.fast_nodes.size() / 2 + 4);
    s.wide_nodes.push_back(WideNode());
    std::vector<Item> todo{Item{0, 0, 1}};
    uint32_t max_depth = 1;
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        max_depth = std::max(max_depth, it.depth);
        Slot c[4];
        int n = 2;
        slots_of(it.bin, c);
        while (n < 4) {
            int best = -1;
            double best_area = -1.0;
            for (int k = 0; k < n; ++k)
                if (c[k].link >= 0 && area(c[k]) > best_area) { best_area = area(c[k]); best = k; }
            if (best < 0) break;
            Slot two[2];
            slots_of(c[best].link, two);
            c[best] = two[0];
            c[n++] = two[1];
        }
        WideNode w;
        std::memset(&w, 0, sizeof w);
        int32_t kids[4];
        for (int k = 0; k < 4; ++k) {
            if (k < n) {
                w.lo_x[k] = c[k].lo[0]; w.lo_y[k] = c[k].lo[1]; w.lo_z[k] = c[k].lo[2];
                w.hi_x[k] = c[k].hi[0]; w.hi_y[k] = c[k].hi[1]; w.hi_z[k] = c[k].hi[2];
                if (c[k].link >= 0) {
                    kids[k] = int32_t(s.wide_nodes.size());
                    s.wide_nodes.push_back(WideNode());
                    w.child[k] = kids[k];
                } else { kids[k] = -1; w.child[k] = c[k].link; }
            } else { kids[k] = -1; w.child[k] = kWideEmpty; }
        }
        s.wide_nodes[size_t(it.wide)] = w;
        for (int k = n - 1; k >= 0; --k)
            if (kids[k] >= 0) todo.push_back(Item{c[k].link, kids[k], it.depth + 1});
    }
    s.wide_depth = max_depth;
}

}  // namespace bdpt
