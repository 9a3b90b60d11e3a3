// libbdpt context: one CUDA device, one stream, device-resident scene + film.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "host_scene.h"
#include "traverse_ref.cuh"

struct bdpt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    double last_ms = 0.0;

    bdpt::HostScene host;
    bool have_scene = false;
    bdpt::SceneView view{};
    std::vector<void*> scene_allocs;

    // film (camera.h:55): fp64 RGB sum
    double* film = nullptr;
    int film_w = 0, film_h = 0;

    // counters
    unsigned long long* d_stats = nullptr;  // [16] device counters
    uint64_t launches = 0;
    uint64_t samples = 0;

    // texture table of bdpt_scene_set_textures: belongs to the context, survives scene uploads
    bdpt::Tex* d_tex = nullptr;
    unsigned char* d_tex_data = nullptr;
    uint32_t n_tex = 0;
    // media of bdpt_scene_set_media: likewise
    bdpt::MediumDev* d_media = nullptr;
    bdpt::TriGeom* d_med_tris = nullptr;
    bdpt::MedNode* d_med_nodes = nullptr;    // trees over the large boundary meshes, all media back to back
    uint32_t n_media = 0;
    bdpt::SphereDev* d_spheres = nullptr;   // bdpt_scene_set_spheres
    uint32_t n_spheres = 0;
    uint32_t spheres_max_mat = 0;           // largest material index a sphere names (checked against the scene at render time)

    // NCCL communicator of bdpt_comm_init (film_reduce.cu); ncclComm_t
    void* comm = nullptr;
    int comm_rank = 0, comm_size = 1;
    // samples handed to the wavefronts by the bdpt_render call in flight (bdpt_render_progress reads them from any thread)
    std::atomic<int64_t> progress_done{0}, progress_total{0};

    // wavefront workspace (render.cu)
    void* wf = nullptr;
    void* timers = nullptr;
    double kernel_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};       // per kernel class, KT_*
    uint64_t kernel_launches[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

namespace bdpt {

#define BDPT_CUDA(ctx, call)                                                              \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);             \
            return BDPT_ERR_CUDA;                                                         \
        }                                                                                 \
    } while (0)

// device counter slots
enum { ST_NODE = 0, ST_TRI = 1, ST_RAYS_CLOSEST = 2, ST_RAYS_SHADOW = 3, ST_CAM_VERTS = 4, ST_LIGHT_VERTS = 5,
       ST_NODE_CLOSEST = 6, ST_TRI_CLOSEST = 7, ST_NODE_SHADOW = 8, ST_TRI_SHADOW = 9, ST_MAX_CLOSEST = 10, ST_MAX_SHADOW = 11, ST_DEFER_CLOSEST = 12, ST_DEFER_SHADOW = 13,
       ST_BOX_TESTS = 14, ST_BOX_HITS = 15, ST_TRI_HITS = 16, ST_COUNT = 24 };
// the reference's seven BvhStats counters (stats.h:8-16) need the box / hit tallies beside nodes and triangles
#ifdef __CUDACC__
template <int KIND /* 0 closest, 1 shadow, 2 batch trace */, bool REF_TREE>
__device__ __forceinline__ void add_trav(unsigned long long* stats, const TravStats& ts) {
    const int sn = KIND == 0 ? ST_NODE_CLOSEST : (KIND == 1 ? ST_NODE_SHADOW : ST_NODE), st = KIND == 0 ? ST_TRI_CLOSEST : (KIND == 1 ? ST_TRI_SHADOW : ST_TRI);
    atomicAdd(&stats[sn], (unsigned long long)ts.nodes);
    atomicAdd(&stats[st], (unsigned long long)ts.tris);
    if (KIND < 2) atomicMax(&stats[KIND == 0 ? ST_MAX_CLOSEST : ST_MAX_SHADOW], (unsigned long long)(ts.nodes + ts.tris));
    atomicAdd(&stats[ST_BOX_TESTS], (unsigned long long)ts.nodes * (REF_TREE ? 1ull : 2ull));
    atomicAdd(&stats[ST_BOX_HITS], (unsigned long long)ts.box_hits);
    atomicAdd(&stats[ST_TRI_HITS], (unsigned long long)ts.tri_hits);
}
#endif
// kernel classes of the wavefront (bdpt_kernel_times)
enum { KT_GENERATE = 0, KT_TRACE = 1, KT_SHADE = 2, KT_CONNECT = 3, KT_SHADOW = 4, KT_TRACE_EXACT = 5, KT_SHADOW_EXACT = 6, KT_COUNT = 7 };

// BDPT_WIDE: bdpt_trace walks the 4-wide nodes for closest-hit (bit 0) / any-hit (bit 1) queries.  Default 0: once
// k_trace's walk had its warp meeting point the binary walk measured 3-4 % faster than the 4-wide one.
inline uint32_t wide_walk_mode() {
    const char* e = std::getenv("BDPT_WIDE");
    return e ? uint32_t(std::atoi(e)) & 3u : 0u;
}

int render_impl(bdpt_ctx* ctx, const bdpt_camera* cam, const bdpt_render_opts* opts);
// api.cu: the fp32 triangle records + error scale of the warp-synchronous walk (traverse_w32.cuh), queued on ctx->stream
int derive_walk32(bdpt_ctx* ctx, SceneView& v);
void comm_free(bdpt_ctx* ctx);   // film_reduce.cu
inline void apply_textures(bdpt_ctx* ctx, SceneView& v) {
    v.tex = ctx->d_tex; v.tex_data = ctx->d_tex_data; v.n_tex = ctx->n_tex;
    v.media = ctx->d_media; v.med_tris = ctx->d_med_tris; v.n_media = ctx->n_media;
    v.spheres = ctx->d_spheres; v.n_spheres = ctx->n_spheres;
}
void render_free(bdpt_ctx* ctx);
// gpu_bvh.cu: the fast tree built on the device from triangles already uploaded (queued on ctx->stream)
int build_fast_bvh_gpu(bdpt_ctx* ctx, const TriGeom* d_tris, uint32_t n, const double lo[3], const double hi[3], double extent,
                       const FastNode** out_nodes, const uint32_t** out_prims, const TriGeom** out_tris, uint32_t* out_n_nodes,
                       uint32_t* out_depth);

}  // namespace bdpt
