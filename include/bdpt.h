/* libbdpt — C-ABI of the B200-native BDPT hot path.
 *
 * The reference (teehee567/Bidirectional-Path-Tracer) has no FFI layer; its de-facto operator
 * boundary is three C++ calls in one TU (paths relative to /root/reference/src):
 *   camera::render(world, lights)                 camera.h:43      -> bdpt_render_*
 *   hittable::hit(ray, interval, hit_record&)     hittable.h:33    -> bdpt_trace / bdpt_visible
 *   bvh_node(hittable_list)                       bvh.h:13-48      -> bdpt_scene_upload (+ BVH build)
 * Everything here is plain pointers and sizes; no torch / C++ types cross the boundary.
 * All calls return 0 on success, a negative bdpt_status otherwise; bdpt_last_error() gives the text.
 * A context is bound to one CUDA device and is not thread-safe; contexts are independent.
 * There is NO CPU fallback: without a CUDA device bdpt_ctx_create fails with BDPT_ERR_CUDA.
 */
#ifndef BDPT_H
#define BDPT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bdpt_ctx bdpt_ctx;

typedef enum {
    BDPT_OK = 0,
    BDPT_ERR_ARG = -1,     /* bad argument (null pointer, size 0, unknown enum) */
    BDPT_ERR_CUDA = -2,    /* CUDA runtime error, text in bdpt_last_error */
    BDPT_ERR_STATE = -3,   /* call order (e.g. trace before scene upload) */
    BDPT_ERR_NOMEM = -4
} bdpt_status;

/* material classes reachable from the reference's loader (scene_loader.h:101-170, material.h:42-147) */
enum { BDPT_MAT_LAMBERTIAN = 0, BDPT_MAT_METAL = 1, BDPT_MAT_DIELECTRIC = 2, BDPT_MAT_LIGHT = 3 };

typedef struct {
    uint32_t type;    /* BDPT_MAT_* */
    uint32_t tex;     /* 0: the colour is rgb (solid_color, texture.h:14-26); k > 0: texture k-1 of bdpt_scene_set_textures
                         gives the lambertian albedo / the light's emission (material.h:45,136) */
    double   rgb[3];  /* albedo (lambertian, metal) or emission (light); unused for dielectric */
    double   param;   /* metal: fuzz, clamped to <=1 as material.h:71; dielectric: refraction index */
} bdpt_material;

/* Textures (texture.h:28-87, perlin.h).  The reference's loader can construct none of them
 * (scene_loader.h:101-170 only ever makes solid colours), so they are an extension of the scene surface; the arithmetic
 * is the reference's: texture::value(u, v, p) with (u, v) = the triangle's barycentric hit coordinates
 * (triangle.h:66-67) and p the hit point.  Checker colours are solid colours (texture.h:34-35). */
enum { BDPT_TEX_SOLID = 0, BDPT_TEX_CHECKER = 1, BDPT_TEX_NOISE = 2, BDPT_TEX_IMAGE = 3 };
typedef struct {
    uint32_t type;         /* BDPT_TEX_* */
    uint32_t pad;
    double   a[3];         /* solid: the colour; checker: the even colour */
    double   b[3];         /* checker: the odd colour */
    double   scale;        /* checker: cell size (texture.h:31); noise: frequency of the sine (texture.h:80) */
    uint64_t data_offset;  /* into the data blob, 8-byte aligned.  noise: perlin's tables (perlin.h:62-66): 256 x 3 doubles
                              randvec, then perm_x, perm_y, perm_z as 3 x 256 int32.  image: width * height * 3 bytes, rows top to bottom */
    int32_t  width, height;/* image */
} bdpt_texture;

/* The public camera fields of camera.h:26-41 that shape the image. */
typedef struct {
    double  aspect_ratio;
    int32_t image_width;
    int32_t samples_per_pixel;   /* sqrt_spp = int(sqrt(spp)) strata per axis, camera.h:164 */
    int32_t max_depth;
    int32_t pad;
    double  background[3];
    double  vfov;
    double  lookfrom[3];
    double  lookat[3];
    double  vup[3];
    double  defocus_angle;       /* must be <= 0 (scene_loader.h:463 forces it off) */
    double  focus_dist;
} bdpt_camera;

/* BVH flavours */
enum {
    BDPT_BVH_REFERENCE = 0,  /* the reference's median-split topology (bvh.h:20-48), fp64 boxes:
                                traversal results are bit-identical to bvh_node::hit */
    BDPT_BVH_FAST      = 1   /* binned-SAH topology, fp32 conservative boxes, ordered traversal;
                                triangles are still decided in fp64 with the reference's tie rule */
};

/* trace flags */
enum {
    BDPT_TRACE_CLOSEST = 0,
    BDPT_TRACE_ANY     = 1,   /* shadow query: prim = any occluder or MISS, t undefined */
    BDPT_TRACE_DEVICE_PTRS = 16, /* rays/out pointers are device memory of the ctx's device */
    BDPT_TRACE_STATS = 32,       /* count node visits / triangle tests into bdpt_stats (slower) */
    BDPT_TRACE_SPLIT = 64        /* BDPT_BVH_FAST only: every ray is walked by a whole warp, split 32 ways along its length
                                    (the kernel path bdpt_render uses for rays over its work limit); same answers */
};

#define BDPT_MISS 0xFFFFFFFFu

/* n x 8 doubles: ox,oy,oz, dx,dy,dz, tmin,tmax (ray.h:6-29 + interval.h:4-24, inclusive ends) */
typedef struct { double o[3], d[3], tmin, tmax; } bdpt_ray;
/* hit_record (hittable.h:10-27) reduced to what cannot be recomputed, plus the prim ID it lacks */
typedef struct { double t, u, v; uint32_t prim; uint32_t pad; } bdpt_hit;

typedef struct {
    uint64_t rays_closest;   /* world.hit queries from trace_path (camera.h:336) */
    uint64_t rays_shadow;    /* world.hit queries from visible() (camera.h:437) */
    uint64_t node_visits;    /* stats build only (BDPT_STATS), else 0 */
    uint64_t tri_tests;      /* stats build only */
    uint64_t samples;        /* (pixel, stratum) samples rendered */
    uint64_t kernel_launches;/* our kernels launched since ctx creation / last reset */
    uint64_t camera_vertices;
    uint64_t light_vertices;
    /* bdpt_render with BDPT_RENDER_STATS: traversal work split by ray kind */
    uint64_t nodes_closest, tris_closest, nodes_shadow, tris_shadow;
    /* ... and the most expensive single ray of each kind (nodes + triangles it touched): the latency floor of a launch */
    uint64_t max_work_closest, max_work_shadow;
    /* rays the fast traversal handed to the reference-exact kernels (ties, knife edges, zero direction components) */
    uint64_t deferred_closest, deferred_shadow;
    /* The rest of the reference's BvhStats (stats.h:8-16), counted with node_visits / tri_tests (BDPT_TRACE_STATS,
     * BDPT_RENDER_STATS).  Through BDPT_BVH_REFERENCE they are the reference's own numbers for closest-hit queries
     * (aabb_tests == bvh_node_visits there: bvh.h:51-52); through BDPT_BVH_FAST a node holds both child boxes, so
     * aabb_tests = 2 x nodes fetched and aabb_hits = child boxes passed.  bvh_nodes_built = nodes of the reference
     * topology (stats.h:15; bvh.h:14). */
    uint64_t aabb_tests, aabb_hits, triangle_hits, bvh_nodes_built;
} bdpt_stats_t;

/* Device time per kernel class of bdpt_render since the last bdpt_stats_reset, from CUDA event pairs on
 * the context's stream around each launch (ms[i], launches[i]; i = BDPT_KT_*). */
enum { BDPT_KT_GENERATE = 0, BDPT_KT_TRACE = 1, BDPT_KT_SHADE = 2, BDPT_KT_CONNECT = 3, BDPT_KT_SHADOW = 4,
       BDPT_KT_TRACE_EXACT = 5, BDPT_KT_SHADOW_EXACT = 6 /* the deferred rays' reference-topology kernels */, BDPT_KT_COUNT = 7 };

typedef struct {
    uint64_t seed;
    int32_t  stratum_begin;  /* strata [begin,end) of the sqrt_spp^2 per pixel, row-major s_j*n+s_i */
    int32_t  stratum_end;    /* <=0 means all */
    int32_t  stratum_step;   /* rank g of G renders begin=g, step=G; 0/1 = contiguous */
    int32_t  bvh;            /* BDPT_BVH_* used for traversal */
    int32_t  row_begin;      /* rows [row_begin,row_end) ; row_end<=0 means image height */
    int32_t  row_end;
    int64_t  batch_samples;  /* wavefront size; <=0 picks a default */
    int32_t  flags;          /* BDPT_RENDER_* */
    int32_t  streams;        /* wavefronts kept in flight on separate CUDA streams; <=0 picks a default (4) */
    int32_t  weighting;      /* BDPT_WEIGHT_*: how the sampling strategies of a path are combined */
    int32_t  pad;
} bdpt_render_opts;
/* BDPT_WEIGHT_REF_SUM (default, the parity gate): the reference's estimator, bidirectional_color camera.h:294-323 --
 *   emission at every camera vertex plus every camera x light vertex connection, all with weight 1.
 * BDPT_WEIGHT_MIS: the same two random walks per sample combined with balance-heuristic weights over all
 *   strategies -- camera path hits a light (s = 0), connection to the emitter point (s = 1), vertex connections
 *   (s >= 2) and light tracing (t = 1), whose contributions are SPLATS: atomic adds into whatever pixel the light
 *   vertex projects to.  Unbiased: its expectation is that of the reference's own path_trace_color
 *   (camera.h:255-292), at most max_depth surface vertices per path; not a reference behaviour. */
enum { BDPT_WEIGHT_REF_SUM = 0, BDPT_WEIGHT_MIS = 1 };
enum { BDPT_RENDER_STATS = 1,  /* count node visits / triangle tests of the traversal kernels (slower) */
       BDPT_RENDER_SERIAL = 2  /* keep every kernel of a wavefront on ONE stream (the shadow stage normally overlaps the
                                  next iteration on a side stream): per-kernel event times are then each kernel's own */ };

int  bdpt_ctx_create(int device, bdpt_ctx** out);
void bdpt_ctx_destroy(bdpt_ctx* ctx);
const char* bdpt_last_error(bdpt_ctx* ctx);   /* ctx may be NULL: last creation error */

/* Triangle soup in scene.world insertion order (prim ID = index); verts is n_tris x 9
 * doubles (v0,v1,v2: triangle.h:21). Lights are the triangles whose material is BDPT_MAT_LIGHT, in
 * insertion order (scene_loader.h:190-202); if there are none every triangle is a light (main.cpp:67).
 * Builds both BVH flavours (the reference topology on the host, the fast tree on the device) and uploads them. */
int bdpt_scene_upload(bdpt_ctx* ctx, const double* verts, const uint32_t* mat_id, uint64_t n_tris,
                      const bdpt_material* mats, uint32_t n_mats);
/* The same with options.  Where the BDPT_BVH_FAST tree is built (results are identical either way: the tree only selects
 * candidates for the exact test):
 *   default (0) / BDPT_UPLOAD_FAST_TREE_GPU   on the device (csrc/gpu_bvh.cu): Morton sort, then binned SAH with in-place
 *       partition at every level down to ranges of 32 triangles and the exact SAH sweep below -- the host builder's algorithm
 *       and its tree quality (20.2 nodes per closest ray on the bench scene for both), 11 ms for 180 k triangles against 35-45;
 *       the reference-topology tree (bvh.h:20-48) is built on the host meanwhile
 *   BDPT_UPLOAD_FAST_TREE_HOST               the host's threaded binned-SAH build (also what bdpt_scene_image_build uses, which
 *       needs no device, and what carries the optional 4-wide tree of BDPT_WIDE) */
#define BDPT_UPLOAD_FAST_TREE_GPU 1u
#define BDPT_UPLOAD_FAST_TREE_HOST 2u
int bdpt_scene_upload_ex(bdpt_ctx* ctx, const double* verts, const uint32_t* mat_id, uint64_t n_tris,
                         const bdpt_material* mats, uint32_t n_mats, uint32_t flags);
/* The texture table bdpt_material.tex refers to, with its data blob (copied; may be freed on return).  Belongs to the
 * context, not to the scene: call it before or after bdpt_scene_upload / bdpt_scene_upload_image, again to replace it,
 * with n_tex = 0 to drop it. */
int bdpt_scene_set_textures(bdpt_ctx* ctx, const bdpt_texture* tex, uint32_t n_tex, const void* data, uint64_t data_bytes);
/* Participating media: constant_medium (materials/volumes/constant_medium.h:8-61) with its isotropic phase function
 * (material.h:149-171); SURVEY 8f-4.  Not reachable from the reference's loader either, so an extension of the scene surface
 * like the textures.  A medium is a closed boundary mesh (a convex one for the reference's two-hit logic to mean "inside"),
 * a density and a colour.  The boundary triangles are NOT world prims: they have no prim ID, no material and occlude nothing.
 * bdpt_render then makes world.hit what it is in the reference with the media in the world list: after the triangle walk
 * every medium is asked with the interval [tmin, nearest hit so far] (constant_medium.h:22-58: two boundary hits, clamp,
 * ONE random draw, scatter distance -1/density * log(xi)); a scatter point is a path vertex with the isotropic material
 * (uniform sphere scatter, f = albedo / 4 pi, the reference's arbitrary normal (1,0,0)); visible() lets a medium
 * between two vertices scatter the shadow ray with the same probability.  BDPT_WEIGHT_MIS: a scatter point is a
 * connectable vertex (no cosine, 1 / (4 pi) both ways) and every pdf and contribution holds the analytic transmittance of its
 * segment.  bdpt_trace / bdpt_visible stay deterministic queries (media draw random numbers).  Limits: 16 media, 4096 boundary triangles each
 * (a boundary of more than 16 gets a small tree of its own; a box's 12 are a loop).
 * Belongs to the context like the texture table: before or after the scene upload, again to replace, n_media = 0 to drop. */
typedef struct {
    double   density;        /* > 0 */
    double   albedo[3];      /* isotropic's solid colour */
    uint64_t first_tri;      /* boundary triangles [first_tri, first_tri + n_tris) of boundary_verts */
    uint64_t n_tris;
} bdpt_medium;
int bdpt_scene_set_media(bdpt_ctx* ctx, const bdpt_medium* media, uint32_t n_media, const double* boundary_verts /* n x 9 */,
                         uint64_t n_boundary_tris);
/* Analytic spheres (north_star: "ray-triangle/sphere intersection").  NOT a reference behaviour: the reference tessellates every
 * sphere into 960 triangles (scene_loader.h:212-242,274-294) and that stays the default of the loader; this is the opt-in exact
 * primitive.  A sphere is a world prim with a material of the scene's table (lambertian, metal or dielectric; an emitter must be
 * triangles: lights are sampled by area over triangles, triangle.h:199-224), kept in a side list of at most 256 that every
 * world.hit asks after the triangle walk, in list order, with the bound found so far (closest hit), and that every visible()
 * query asks for any hit.  Intersection: the quadratic in fp64, nearer root first, inclusive interval like triangle.h:68, outward
 * normal (p - c) / r face-forwarded like hittable.h:20-26.  Either weighting.  bdpt_trace / bdpt_visible ask the sphere list too while one is set (world queries: a hit on sphere k reports
 * prim = 0x10000000 | k, u = v = 0); media draw random numbers and stay out of the batch queries.  Belongs to the context like textures and media; mat_id is checked against the uploaded scene when rendering. */
typedef struct { double center[3]; double radius; uint32_t mat_id; uint32_t pad; } bdpt_sphere;
int bdpt_scene_set_spheres(bdpt_ctx* ctx, const bdpt_sphere* spheres, uint32_t n_spheres);
int bdpt_bvh_info(bdpt_ctx* ctx, int bvh, uint64_t* n_nodes, uint32_t* max_depth);

/* ---- scene image: everything bdpt_scene_upload derives from the soup -- the flattened triangle
 * records, the light list, the reference-topology tree (bvh.h:13-48) with its tie-rule side tables and the
 * binned-SAH tree -- as ONE relocatable block in the kernels' own layout.  Build it once (host only, no CUDA
 * device needed), keep it in memory or in a file, upload it with one allocation and one host-to-device copy:
 * the reference rebuilds its BVH on every run (main.cpp:66).  Rendering from an image is bit-identical to
 * rendering from bdpt_scene_upload of the same soup.  The image is tied to this library's record layouts
 * (a mismatch is refused) and is not portable across endianness. */
typedef struct {
    uint64_t n_tris, n_ref_nodes, n_fast_nodes;
    uint32_t n_mats, n_lights, ref_depth, fast_depth;
    uint64_t source_key;     /* as passed to bdpt_scene_image_build */
    uint64_t payload_hash;
    uint64_t device_bytes;   /* what bdpt_scene_upload_image copies to the device */
} bdpt_scene_image_info_t;
/* *image is allocated by the library (4096-byte aligned); release with bdpt_scene_image_free. */
int  bdpt_scene_image_build(const double* verts, const uint32_t* mat_id, uint64_t n_tris, const bdpt_material* mats,
                            uint32_t n_mats, uint64_t source_key, void** image, uint64_t* bytes);
void bdpt_scene_image_free(void* image);
/* Header / section-table checks (always); verify & 1: the payload hash; verify & 2: the range check of every link and
 * index in the payload (what bdpt_scene_upload_image runs unless told BDPT_IMAGE_TRUSTED); out may be NULL. */
int  bdpt_scene_image_info(const void* image, uint64_t bytes, int verify, bdpt_scene_image_info_t* out);
const char* bdpt_scene_image_last_error(void);
#define BDPT_IMAGE_VERIFY 1u   /* check the payload hash before copying (one pass over the image on the host); the file
                                  front-end (bdpt_scene_file_open_cached, bdpt --cache) checks it when it opens the file */
#define BDPT_IMAGE_TRUSTED 2u  /* skip the range check of the payload's links and indices (tree links, prim / material / light
                                  ids), which otherwise runs on the host while the copy is in flight (~1 ms per 30 k triangles):
                                  only for an image this process has just built itself */
/* image: host memory, 8-byte aligned (pinned memory makes the copy faster); may be freed on return. */
int  bdpt_scene_upload_image(bdpt_ctx* ctx, const void* image, uint64_t bytes, uint32_t flags);
/* DFS leaf order of the BDPT_BVH_REFERENCE build: order[k] = prim ID (n_tris entries) */
int bdpt_bvh_leaf_order(bdpt_ctx* ctx, uint32_t* order);

/* Host-only (no CUDA device needed): build the BDPT_BVH_REFERENCE topology for a triangle soup and
 * return its DFS leaf order, node count and, when fast_nodes != NULL, the fast BVH's node count.
 * Lets the host builder be checked against the reference without a GPU. */
int bdpt_host_bvh_build(const double* verts, uint64_t n_tris, uint32_t* order /* n_tris */,
                        uint64_t* ref_nodes, uint64_t* fast_nodes);

/* hittable::hit over a batch (hittable.h:33 / bvh.h:50-59). */
int bdpt_trace(bdpt_ctx* ctx, int bvh, const bdpt_ray* rays, uint64_t n, bdpt_hit* out, int flags);
/* camera::visible (camera.h:425-438) over n segments a->b (n x 3 doubles each); out[i] = 1 visible. */
int bdpt_visible(bdpt_ctx* ctx, int bvh, const double* a, const double* b, uint64_t n, uint8_t* out,
                 int flags);

/* camera::render's sample loop (camera.h:111-129): adds every selected (pixel, stratum) sample of
 * bidirectional_color (camera.h:294-323) into the context's device film (fp64 RGB SUM, camera.h:55).
 * The film persists across calls until bdpt_film_clear. */
int bdpt_render(bdpt_ctx* ctx, const bdpt_camera* cam, const bdpt_render_opts* opts);
int bdpt_film_clear(bdpt_ctx* ctx, int width, int height);
/* device pointer of the film (W*H*3 doubles) so a caller can reduce it across GPUs in place */
int bdpt_film_device_ptr(bdpt_ctx* ctx, void** ptr, uint64_t* n_doubles);
int bdpt_film_download(bdpt_ctx* ctx, double* film_sum /* W*H*3 */);
/* Dimensions of the context's film.  bdpt_render sizes the film by the reference's own rule, image_height =
 * max(1, int(image_width / aspect_ratio)) (camera.h:161-162), so this is the one source of truth for the buffer a
 * caller hands to bdpt_film_download / bdpt_film_to_rgb8. */
int bdpt_film_size(bdpt_ctx* ctx, int* width, int* height);
/* wpng.h:14-35 on the device film: average, clamp [0,0.999], sqrt, 256x truncate. out: W*H*3 bytes (host) */
int bdpt_film_to_rgb8(bdpt_ctx* ctx, int samples_per_pixel, uint8_t* out);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as bdpt_film_to_rgb8, byte
 * for byte, but the W*H*3 bytes of device staging between the tone-map kernel and the copy to `out` live in a static
 * device array of BDPT_NOALLOC_RGB8_MAX_BYTES (32 MiB: up to 3840 x 2160; held by every GPU that loads the library
 * until exit) instead of a cudaMalloc/cudaFree pair per call: the call allocates nothing on the device or the host.
 * A film larger than the limit is refused with BDPT_ERR_ARG and a message naming both sizes.  All calls on a device
 * share the array, which keeps the previous call's bytes (none of them is read: every byte is written before the copy);
 * calls that overlap on the GPU -- two contexts of one device from two threads -- race on it: serialise them. */
#define BDPT_NOALLOC_RGB8_MAX_BYTES (32u << 20)
int bdpt_film_to_rgb8_noalloc(bdpt_ctx* ctx, int samples_per_pixel, uint8_t* out);

/* ---- the film reduce: the one exchange step of a multi-GPU frame.  Every rank renders its strata
 * (bdpt_render_opts.stratum_begin / stratum_step) into its own film; ONE ncclReduce(sum, fp64) over NVLink then adds
 * the films in place onto `root` -- what camera.h:117-124 does with `pixel_color +=` on one framebuffer.
 * libnccl.so.2 is opened on first use (a process that already holds an NCCL, e.g. torch's, shares it).
 *   one process, n devices:   bdpt_film_reduce(ctxs, n, root)           (communicators are created once per device list)
 *   one process per GPU:      id from bdpt_comm_unique_id on one rank, carried to the others by the launcher;
 *                             bdpt_comm_init on every rank, once; bdpt_film_reduce_comm(ctx, root) per frame.
 * Both are blocking; bdpt_last_device_ms then gives the device time of the reduce on the calling / root context. */
int bdpt_film_reduce(bdpt_ctx** ctxs, int n, int root);
int bdpt_comm_unique_id(void* id128 /* 128 bytes out */);
int bdpt_comm_init(bdpt_ctx* ctx, int nranks, int rank, const void* id128);
int bdpt_film_reduce_comm(bdpt_ctx* ctx, int root);
/* The communicator of bdpt_comm_init belongs to the process (one per device) and outlives the context that joined:
 * another context on the same device takes it over with bdpt_comm_attach instead of a new rendezvous. */
int bdpt_comm_attach(bdpt_ctx* ctx);
/* Samples of the bdpt_render call in flight already handed to the GPU, out of how many: callable from another
 * thread while bdpt_render runs (the progress bar of camera::render, camera.h:67-109). */
int bdpt_render_progress(bdpt_ctx* ctx, uint64_t* samples_done, uint64_t* samples_total);

int bdpt_stats(bdpt_ctx* ctx, bdpt_stats_t* out);
int bdpt_stats_reset(bdpt_ctx* ctx);
/* GPU time (ms, CUDA events on the ctx stream) of the last bdpt_trace / bdpt_visible / bdpt_render
 * device work, excluding host<->device copies. */
int bdpt_last_device_ms(bdpt_ctx* ctx, double* ms);
int bdpt_sync(bdpt_ctx* ctx);
/* Measurement aid: read bandwidth (GB/s) of a `bytes`-byte device buffer streamed `reps` times with 16-byte loads.
 * bytes well inside the L2 (e.g. 48 MiB) = the L2-resident read peak; bytes far beyond it = the HBM read rate. */
int bdpt_bench_read(bdpt_ctx* ctx, uint64_t bytes, int reps, double* gb_per_s);
int bdpt_kernel_times(bdpt_ctx* ctx, double* ms /* BDPT_KT_COUNT */, uint64_t* launches /* BDPT_KT_COUNT */);

/* ---- scene files: the host front-end behind load_scene_from_yaml (scene_loader.h:480-523) and
 * write_png (wpng.h:38-89), for callers that are not C++. Host only; errors via bdpt_host_last_error. */
typedef struct bdpt_scene_file bdpt_scene_file;
int  bdpt_scene_file_load(const char* yaml_path, bdpt_scene_file** out);
void bdpt_scene_file_free(bdpt_scene_file* f);
int  bdpt_scene_file_info(const bdpt_scene_file* f, uint64_t* n_tris, uint32_t* n_mats, uint64_t* n_lights,
                          bdpt_camera* cam, char* output_name, uint64_t output_cap);
/* copies the flattened soup: verts n_tris*9, mat_id n_tris, mats n_mats (any may be NULL) */
int  bdpt_scene_file_copy(const bdpt_scene_file* f, double* verts, uint32_t* mat_id, bdpt_material* mats);
/* The scene's texture table for bdpt_scene_set_textures (a `texture:` map in a lambertian / light material: checker, noise,
 * image -- an extension of the reference's schema, host/scene_loader.cpp).  Call with tex = data = NULL for the sizes. */
int  bdpt_scene_file_textures(const bdpt_scene_file* f, uint32_t* n_tex, bdpt_texture* tex, uint64_t* data_bytes, void* data);
/* The scene's participating media for bdpt_scene_set_media (a `media:` section in the YAML -- an extension of the
 * reference's schema, host/scene_loader.cpp).  Call with media = boundary_verts = NULL for the sizes. */
int  bdpt_scene_file_media(const bdpt_scene_file* f, uint32_t* n_media, bdpt_medium* media, uint64_t* n_boundary_tris, double* boundary_verts);
/* The scene's analytic spheres for bdpt_scene_set_spheres (`analytic: true` on a YAML Sphere surface: an extension; without it the
 * surface is tessellated like the reference does).  Their materials are part of the table bdpt_scene_file_copy returns (appended
 * after the triangles' materials).  Call with spheres = NULL for the count. */
int  bdpt_scene_file_spheres(const bdpt_scene_file* f, uint32_t* n_spheres, bdpt_sphere* spheres);
int  bdpt_write_png(const char* path, const uint8_t* rgb, int width, int height);
/* Scene-image files: save writes to a temporary file of its own beside `path` ("<path>.<pid>.<n>.tmp"), syncs it and
 * renames it over `path`, so concurrent savers (ranks missing the same cache) never write into each other's file; map gives a read-only mapping of the file
 * (page cache backed: nothing is copied until the upload reads it). */
int  bdpt_scene_image_save(const char* path, const void* image, uint64_t bytes);
int  bdpt_scene_image_map(const char* path, const void** image, uint64_t* bytes);
void bdpt_scene_image_unmap(const void* image, uint64_t bytes);
/* Key of a YAML scene's sources: a hash of the YAML bytes and of every OBJ file its `object` surfaces name, in
 * scene order -- what bdpt_scene_file_open_cached compares with an image's source_key. */
int  bdpt_scene_source_key(const char* yaml_path, uint64_t* key);
/* load_scene_from_yaml with a scene-image cache beside it: when cache_path holds an image whose source_key
 * matches, only the YAML is parsed (camera, output name) and the image is mapped -- no OBJ parsing, no
 * flattening, no BVH build; otherwise the scene is loaded and flattened as bdpt_scene_file_load does, the
 * image is built and written to cache_path.  *hit = 1 when the cache was used.  The image stays valid until
 * bdpt_scene_file_free; bdpt_scene_file_copy on a cache hit fails (there is no soup). */
int  bdpt_scene_file_open_cached(const char* yaml_path, const char* cache_path, bdpt_scene_file** out,
                                 const void** image, uint64_t* image_bytes, int* hit);
const char* bdpt_host_last_error(void);

/* Test hook, host only: the fp32 pre-filter of the warp-synchronous walk (csrc/traverse_w32.cuh) on one ray
 * (o, d, tmin, tmax) and one triangle (v0, e1, e2); best = the walk's current upper bound, vmax >= the largest
 * |coordinate| of the scene.  0 = rejected, 1 = candidate for the exact fp64 test, 2 = certainly accepted. */
int  bdpt_test_filter32(const double* ray, double best, const double* tri_v0e1e2, double vmax, int any, float* t_hi);

#ifdef __cplusplus
}
#endif
#endif
