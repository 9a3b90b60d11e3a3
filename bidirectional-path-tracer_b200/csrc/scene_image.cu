// Scene image: the flattened soup and both BVHs of a scene as ONE relocatable block of bytes, laid
// out exactly as the kernels read them.  Building it (flatten + both host tree builds, the reference's
// bvh_node ctor bvh.h:13-48 among them) is the expensive part of bdpt_scene_upload; an image is built once,
// kept by the caller (memory, or a file through bdpt_scene_image_save / _map), and uploaded with ONE
// cudaMalloc and ONE host-to-device copy.  Every section starts on a 256-byte boundary, so the device
// pointers are base + offset and keep the 16-byte alignment the vector loads need.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <stdexcept>
#include <thread>

#include "ctx.h"
#include "traverse_ref.cuh"

using namespace bdpt;

namespace {

enum Section {
    SEC_TRIS = 0, SEC_SHADE, SEC_MATS, SEC_LIGHTS, SEC_LIGHT_AREA, SEC_FAST_NODES, SEC_WIDE_NODES, SEC_FAST_PRIMS, SEC_FAST_TRIS,
    SEC_REF_NODES, SEC_DFS_RANK, SEC_REF_PARENT, SEC_PRIM_LEAF, SEC_CHAIN_BOX,
    SEC_LEAF_ORDER,   // host only (bdpt_bvh_leaf_order): last, so the device copy stops before it
    SEC_COUNT
};

constexpr uint32_t kVersion = 2;   // 2: + the 4-wide nodes
constexpr uint64_t kAlign = 256;
constexpr char kMagic[8] = {'B', 'D', 'P', 'T', 'I', 'M', 'G', '1'};

struct ImageHeader {
    char magic[8];
    uint32_t version, header_bytes;
    uint64_t total_bytes;
    uint64_t payload_hash;     // image_hash over [header_bytes, total_bytes)
    uint64_t source_key;       // the caller's key of what the image was built from (0 = none)
    uint64_t layout_sig;       // record sizes this library was compiled with
    uint32_t n_tris, n_mats, n_lights, n_ref_nodes, n_fast_nodes, ref_depth, fast_depth, n_wide_nodes;
    uint32_t wide_depth, reserved;
    double light_total_area, extent, center[3];
    struct { uint64_t off, bytes; } sec[SEC_COUNT];
    unsigned char pad[512 - (8 + 8 + 8 * 4 + 4 * 10 + 8 * 5 + 16 * SEC_COUNT)];
};
static_assert(sizeof(ImageHeader) == 512, "ImageHeader must be 512 bytes");

constexpr uint64_t layout_sig() {
    return uint64_t(sizeof(RefNode)) | uint64_t(sizeof(FastNode)) << 8 | uint64_t(sizeof(TriGeom)) << 16 |
           uint64_t(sizeof(TriShade)) << 24 | uint64_t(sizeof(Mat)) << 32 | uint64_t(sizeof(WideNode) & 0xFF) << 56 | uint64_t(SEC_COUNT) << 40 | uint64_t(kVersion) << 48;
}

// Four independent multiply-xor lanes over 64-bit words (the payload is a multiple of 256 bytes): a few GB/s
// on one core, enough to notice a torn or truncated file; it is not a cryptographic digest.
uint64_t image_hash(const unsigned char* p, uint64_t bytes) {
    uint64_t h[4] = {0x9E3779B97F4A7C15ull, 0xC2B2AE3D27D4EB4Full, 0x165667B19E3779F9ull, 0x27D4EB2F165667C5ull};
    const uint64_t words = bytes / 8;
    uint64_t w[4];
    uint64_t i = 0;
    for (; i + 4 <= words; i += 4) {
        std::memcpy(w, p + 8 * i, 32);
        for (int k = 0; k < 4; ++k) {
            h[k] = (h[k] ^ w[k]) * 0x100000001B3ull;
            h[k] = (h[k] << 29) | (h[k] >> 35);
        }
    }
    for (; i < words; ++i) {
        std::memcpy(w, p + 8 * i, 8);
        h[0] = (h[0] ^ w[0]) * 0x100000001B3ull;
        h[0] = (h[0] << 29) | (h[0] >> 35);
    }
    uint64_t r = bytes;
    for (int k = 0; k < 4; ++k) r = (r ^ h[k]) * 0xFF51AFD7ED558CCDull, r ^= r >> 33;
    return r;
}

std::string g_image_err;

// expected bytes of each section given the header's counts
void expected_bytes(const ImageHeader& h, uint64_t out[SEC_COUNT]) {
    const uint64_t n = h.n_tris;
    out[SEC_TRIS] = n * sizeof(TriGeom);
    out[SEC_SHADE] = n * sizeof(TriShade);
    out[SEC_MATS] = uint64_t(h.n_mats) * sizeof(Mat);
    out[SEC_LIGHTS] = uint64_t(h.n_lights) * sizeof(uint32_t);
    out[SEC_LIGHT_AREA] = uint64_t(h.n_lights) * sizeof(double);
    out[SEC_FAST_NODES] = uint64_t(h.n_fast_nodes) * sizeof(FastNode);
    out[SEC_WIDE_NODES] = uint64_t(h.n_wide_nodes) * sizeof(WideNode);
    out[SEC_FAST_PRIMS] = n * sizeof(uint32_t);
    out[SEC_FAST_TRIS] = n * sizeof(TriGeom);
    out[SEC_REF_NODES] = uint64_t(h.n_ref_nodes) * sizeof(RefNode);
    out[SEC_DFS_RANK] = n * sizeof(uint32_t);
    out[SEC_REF_PARENT] = uint64_t(h.n_ref_nodes) * sizeof(int32_t);
    out[SEC_PRIM_LEAF] = n * sizeof(int32_t);
    out[SEC_CHAIN_BOX] = n * 6 * sizeof(double);
    out[SEC_LEAF_ORDER] = n * sizeof(uint32_t);
}

// Header and section-table checks; with `verify` also the payload hash.  Returns nullptr and sets `why` on failure.
const ImageHeader* check_image(const void* image, uint64_t bytes, bool verify, std::string& why) {
    if (!image || bytes < sizeof(ImageHeader)) { why = "scene image: shorter than its header"; return nullptr; }
    if (reinterpret_cast<uintptr_t>(image) % 8) { why = "scene image: buffer must be 8-byte aligned"; return nullptr; }
    const ImageHeader* h = static_cast<const ImageHeader*>(image);
    if (std::memcmp(h->magic, kMagic, 8)) { why = "scene image: bad magic"; return nullptr; }
    if (h->version != kVersion || h->header_bytes != sizeof(ImageHeader)) { why = "scene image: unknown version"; return nullptr; }
    if (h->layout_sig != layout_sig()) { why = "scene image: built by a library with different record layouts"; return nullptr; }
    if (h->total_bytes != bytes) { why = "scene image: truncated (header says " + std::to_string(h->total_bytes) + " bytes, got " + std::to_string(bytes) + ")"; return nullptr; }
    if (h->n_tris == 0 || h->n_mats == 0 || h->n_lights == 0 || h->n_ref_nodes == 0 || h->n_fast_nodes == 0) { why = "scene image: empty scene"; return nullptr; }
    uint64_t want[SEC_COUNT];
    expected_bytes(*h, want);
    uint64_t at = sizeof(ImageHeader);
    for (int s = 0; s < SEC_COUNT; ++s) {
        if (h->sec[s].off != at || h->sec[s].bytes != want[s]) { why = "scene image: section table does not match the counts"; return nullptr; }
        at = (at + want[s] + kAlign - 1) / kAlign * kAlign;
    }
    if (at != bytes) { why = "scene image: size does not match the section table"; return nullptr; }
    if (h->ref_depth >= uint32_t(kRefStackLevels) || h->fast_depth + 2 > 32u) { why = "BVH deeper than the traversal stack"; return nullptr; }
    if (verify && image_hash(static_cast<const unsigned char*>(image) + sizeof(ImageHeader), bytes - sizeof(ImageHeader)) != h->payload_hash) {
        why = "scene image: payload hash mismatch (corrupt file)";
        return nullptr;
    }
    return h;
}

// One linear pass over the payload: every index a kernel follows must stay inside its array.  The header check above
// only ties the section SIZES to the counts; a bit-rotted or crafted file that passes it would otherwise send the walk
// to wild addresses (a device fault an operator has to clear).  ~1 ms per million triangles.
template <class T> const T* sec(const ImageHeader* h, int s) { return reinterpret_cast<const T*>(reinterpret_cast<const unsigned char*>(h) + h->sec[s].off); }
bool validate_payload(const ImageHeader* h, std::string& why) {
    const uint64_t n = h->n_tris;
    auto bad = [&](const char* what) { why = std::string("scene image: ") + what + " out of range (corrupt payload)"; return false; };
    auto leaf_ok = [&](int32_t c) { const uint32_t code = uint32_t(~c); return uint64_t(code >> 3) + (code & 7u) + 1u <= n; };
    const FastNode* fn = sec<FastNode>(h, SEC_FAST_NODES);
    for (uint32_t i = 0; i < h->n_fast_nodes; ++i)
        for (int32_t c : {fn[i].left, fn[i].right})
            if (c >= 0 ? uint32_t(c) >= h->n_fast_nodes : !leaf_ok(c)) return bad("fast-tree link");
    const WideNode* wn = sec<WideNode>(h, SEC_WIDE_NODES);
    for (uint32_t i = 0; i < h->n_wide_nodes; ++i)
        for (int32_t c : wn[i].child)
            if (c != kWideEmpty && (c >= 0 ? uint32_t(c) >= h->n_wide_nodes : !leaf_ok(c))) return bad("wide-tree link");
    const RefNode* rn = sec<RefNode>(h, SEC_REF_NODES);
    const int32_t* parent = sec<int32_t>(h, SEC_REF_PARENT);
    for (uint32_t i = 0; i < h->n_ref_nodes; ++i) {
        const uint64_t lim = rn[i].leaf ? n : h->n_ref_nodes;
        if (rn[i].left < 0 || rn[i].right < 0 || uint64_t(rn[i].left) >= lim || uint64_t(rn[i].right) >= lim) return bad("reference-tree link");
        if (parent[i] < -1 || parent[i] >= int64_t(h->n_ref_nodes)) return bad("reference-tree parent");
    }
    const TriGeom* tris = sec<TriGeom>(h, SEC_TRIS);
    const TriGeom* ftris = sec<TriGeom>(h, SEC_FAST_TRIS);
    const TriShade* shade = sec<TriShade>(h, SEC_SHADE);
    const uint32_t* fprims = sec<uint32_t>(h, SEC_FAST_PRIMS);
    const uint32_t* rank = sec<uint32_t>(h, SEC_DFS_RANK);
    const int32_t* pleaf = sec<int32_t>(h, SEC_PRIM_LEAF);
    const uint32_t* order = sec<uint32_t>(h, SEC_LEAF_ORDER);
    for (uint64_t i = 0; i < n; ++i) {
        if (tris[i].mat >= h->n_mats || ftris[i].mat >= h->n_mats || shade[i].mat >= h->n_mats) return bad("material id");
        if (fprims[i] >= n || order[i] >= n) return bad("prim id");
        if (rank[i] >= 2 * n) return bad("DFS rank");   // a span-1 leaf lists its triangle twice (bvh.h:35-39)
        if (pleaf[i] < 0 || uint32_t(pleaf[i]) >= h->n_ref_nodes) return bad("prim leaf");
    }
    const uint32_t* lights = sec<uint32_t>(h, SEC_LIGHTS);
    for (uint32_t i = 0; i < h->n_lights; ++i) if (lights[i] >= n) return bad("light id");
    const Mat* mats = sec<Mat>(h, SEC_MATS);
    for (uint32_t i = 0; i < h->n_mats; ++i) if (mats[i].type > BDPT_MAT_LIGHT) return bad("material class");
    return true;
}

template <class T>
void put(unsigned char* base, const ImageHeader& h, int s, const std::vector<T>& v) {
    if (v.size() * sizeof(T) != h.sec[s].bytes) throw std::runtime_error("scene image: internal section size mismatch");
    if (!v.empty()) std::memcpy(base + h.sec[s].off, v.data(), v.size() * sizeof(T));
}

}  // namespace

extern "C" {

int bdpt_scene_image_build(const double* verts, const uint32_t* mat_id, uint64_t n_tris, const bdpt_material* mats,
                           uint32_t n_mats, uint64_t source_key, void** image, uint64_t* bytes) {
    if (!image || !bytes) return BDPT_ERR_ARG;
    *image = nullptr;
    *bytes = 0;
    try {
        HostScene hs;
        flatten_scene(verts, mat_id, n_tris, mats, n_mats, hs);
        std::exception_ptr fast_err;
        std::thread fast_thread([&] { try { build_fast_bvh(hs); } catch (...) { fast_err = std::current_exception(); } });
        try { build_ref_bvh(hs); } catch (...) { fast_thread.join(); throw; }
        fast_thread.join();
        if (fast_err) std::rethrow_exception(fast_err);
        if (hs.ref_depth >= uint32_t(kRefStackLevels) || hs.fast_depth + 2 > 32u) throw std::runtime_error("BVH deeper than the traversal stack");

        ImageHeader h;
        std::memset(&h, 0, sizeof h);
        std::memcpy(h.magic, kMagic, 8);
        h.version = kVersion; h.header_bytes = sizeof(ImageHeader);
        h.source_key = source_key; h.layout_sig = layout_sig();
        h.n_tris = uint32_t(hs.tris.size()); h.n_mats = uint32_t(hs.mats.size()); h.n_lights = uint32_t(hs.lights.size());
        h.n_ref_nodes = uint32_t(hs.ref_nodes.size()); h.n_fast_nodes = uint32_t(hs.fast_nodes.size());
        h.ref_depth = hs.ref_depth; h.fast_depth = hs.fast_depth;
        h.n_wide_nodes = uint32_t(hs.wide_nodes.size()); h.wide_depth = hs.wide_depth;
        h.light_total_area = hs.light_total_area; h.extent = hs.extent;
        for (int k = 0; k < 3; ++k) h.center[k] = hs.center[k];
        uint64_t want[SEC_COUNT];
        expected_bytes(h, want);
        uint64_t at = sizeof(ImageHeader);
        for (int s = 0; s < SEC_COUNT; ++s) {
            h.sec[s].off = at; h.sec[s].bytes = want[s];
            at = (at + want[s] + kAlign - 1) / kAlign * kAlign;
        }
        h.total_bytes = at;
        unsigned char* base = static_cast<unsigned char*>(std::aligned_alloc(4096, (at + 4095) / 4096 * 4096));
        if (!base) return BDPT_ERR_NOMEM;
        std::memset(base, 0, at);   // the alignment gaps are part of the hash
        try {
            put(base, h, SEC_TRIS, hs.tris); put(base, h, SEC_SHADE, hs.shade); put(base, h, SEC_MATS, hs.mats);
            put(base, h, SEC_LIGHTS, hs.lights); put(base, h, SEC_LIGHT_AREA, hs.light_area);
            put(base, h, SEC_FAST_NODES, hs.fast_nodes); put(base, h, SEC_WIDE_NODES, hs.wide_nodes); put(base, h, SEC_FAST_PRIMS, hs.fast_prims); put(base, h, SEC_FAST_TRIS, hs.fast_tris);
            put(base, h, SEC_REF_NODES, hs.ref_nodes); put(base, h, SEC_DFS_RANK, hs.dfs_rank); put(base, h, SEC_REF_PARENT, hs.ref_parent);
            put(base, h, SEC_PRIM_LEAF, hs.prim_leaf); put(base, h, SEC_CHAIN_BOX, hs.chain_box); put(base, h, SEC_LEAF_ORDER, hs.leaf_order);
        } catch (...) { std::free(base); throw; }
        h.payload_hash = image_hash(base + sizeof(ImageHeader), at - sizeof(ImageHeader));
        std::memcpy(base, &h, sizeof h);
        *image = base;
        *bytes = at;
    } catch (const std::exception& ex) {
        g_image_err = ex.what();
        return BDPT_ERR_ARG;
    }
    return BDPT_OK;
}

void bdpt_scene_image_free(void* image) { std::free(image); }

const char* bdpt_scene_image_last_error(void) { return g_image_err.c_str(); }

int bdpt_scene_image_info(const void* image, uint64_t bytes, int verify, bdpt_scene_image_info_t* out) {
    std::string why;
    const ImageHeader* h = check_image(image, bytes, (verify & 1) != 0, why);
    if (h && (verify & 2) && !validate_payload(h, why)) h = nullptr;
    if (!h) { g_image_err = why; return BDPT_ERR_ARG; }
    if (out) {
        out->n_tris = h->n_tris; out->n_mats = h->n_mats; out->n_lights = h->n_lights;
        out->n_ref_nodes = h->n_ref_nodes; out->n_fast_nodes = h->n_fast_nodes;
        out->ref_depth = h->ref_depth; out->fast_depth = h->fast_depth;
        out->source_key = h->source_key; out->payload_hash = h->payload_hash;
        out->device_bytes = h->sec[SEC_LEAF_ORDER].off - sizeof(ImageHeader);
    }
    return BDPT_OK;
}

int bdpt_scene_upload_image(bdpt_ctx* ctx, const void* image, uint64_t bytes, uint32_t flags) {
    if (!ctx) return BDPT_ERR_ARG;
    if (flags & ~(BDPT_IMAGE_VERIFY | BDPT_IMAGE_TRUSTED)) { ctx->err = "unknown image upload flags"; return BDPT_ERR_ARG; }
    const bool timing = std::getenv("BDPT_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    std::string why;
    const ImageHeader* h = check_image(image, bytes, (flags & BDPT_IMAGE_VERIFY) != 0, why);
    if (!h) { ctx->err = why; return BDPT_ERR_ARG; }
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    // The new block is allocated and its copy started BEFORE the resident scene is touched: the payload's range check
    // runs on the host while the copy is in flight, and a refused image leaves the old scene in place.
    const unsigned char* src = static_cast<const unsigned char*>(image);
    const uint64_t dev_bytes = h->sec[SEC_LEAF_ORDER].off - sizeof(ImageHeader);
    unsigned char* d = nullptr;
    BDPT_CUDA(ctx, cudaMalloc(&d, dev_bytes));
    cudaError_t e = cudaMemcpyAsync(d, src + sizeof(ImageHeader), dev_bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
        cudaFree(d);
        ctx->err = std::string("cudaMemcpyAsync(scene image): ") + cudaGetErrorString(e);
        return BDPT_ERR_CUDA;
    }
    if (!(flags & BDPT_IMAGE_TRUSTED) && !validate_payload(h, why)) {   // every link and index must stay inside its array
        cudaStreamSynchronize(ctx->stream);
        cudaFree(d);
        ctx->err = why;
        return BDPT_ERR_ARG;
    }
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // the copy, and whatever still read the old scene
    for (void* p : ctx->scene_allocs) cudaFree(p);
    ctx->scene_allocs.clear();
    ctx->have_scene = false;
    ctx->scene_allocs.push_back(d);
    auto at = [&](int s) { return d + (h->sec[s].off - sizeof(ImageHeader)); };
    SceneView v{};
    v.tris = reinterpret_cast<const TriGeom*>(at(SEC_TRIS));
    v.shade = reinterpret_cast<const TriShade*>(at(SEC_SHADE));
    v.mats = reinterpret_cast<const Mat*>(at(SEC_MATS));
    v.lights = reinterpret_cast<const uint32_t*>(at(SEC_LIGHTS));
    v.light_area = reinterpret_cast<const double*>(at(SEC_LIGHT_AREA));
    v.fast_nodes = reinterpret_cast<const FastNode*>(at(SEC_FAST_NODES));
    v.wide_nodes = h->n_wide_nodes ? reinterpret_cast<const WideNode*>(at(SEC_WIDE_NODES)) : nullptr;   // optional section
    v.n_wide_nodes = h->n_wide_nodes;
    v.use_wide = (h->n_wide_nodes && 3u * h->wide_depth + 2u <= uint32_t(kWideStackLevels)) ? wide_walk_mode() : 0u;
    v.fast_prims = reinterpret_cast<const uint32_t*>(at(SEC_FAST_PRIMS));
    v.fast_tris = reinterpret_cast<const TriGeom*>(at(SEC_FAST_TRIS));
    v.ref_nodes = reinterpret_cast<const RefNode*>(at(SEC_REF_NODES));
    v.dfs_rank = reinterpret_cast<const uint32_t*>(at(SEC_DFS_RANK));
    v.ref_parent = reinterpret_cast<const int32_t*>(at(SEC_REF_PARENT));
    v.prim_leaf = reinterpret_cast<const int32_t*>(at(SEC_PRIM_LEAF));
    v.chain_box = reinterpret_cast<const double*>(at(SEC_CHAIN_BOX));
    v.n_lights = h->n_lights; v.n_mats = h->n_mats; v.n_tris = h->n_tris; v.n_ref_nodes = h->n_ref_nodes; v.n_fast_nodes = h->n_fast_nodes;
    v.light_total_area = h->light_total_area; v.extent = h->extent;
    v.center = mk(h->center[0], h->center[1], h->center[2]);
    apply_textures(ctx, v);
    if (int rc = derive_walk32(ctx, v)) { for (void* p : ctx->scene_allocs) cudaFree(p); ctx->scene_allocs.clear(); return rc; }

    // what the other calls read from the host copy: the tree depths (workspace sizing) and the DFS leaf order
    HostScene& hs = ctx->host;
    hs = HostScene{};
    hs.ref_depth = h->ref_depth; hs.fast_depth = h->fast_depth; hs.wide_depth = h->wide_depth;
    hs.extent = h->extent; hs.light_total_area = h->light_total_area;
    for (int k = 0; k < 3; ++k) hs.center[k] = h->center[k];
    const uint32_t* lo = reinterpret_cast<const uint32_t*>(src + h->sec[SEC_LEAF_ORDER].off);
    hs.leaf_order.assign(lo, lo + h->n_tris);

    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // the caller may free or unmap the image on return
    ctx->view = v;
    ctx->have_scene = true;
    if (timing)
        std::fprintf(stderr, "bdpt_scene_upload_image: %u triangles, %.1f MB to the device in %.1f ms\n", h->n_tris, dev_bytes / 1e6,
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    return BDPT_OK;
}

}  // extern "C"
