/* TEST INFRASTRUCTURE ONLY — never linked, imported or executed by the product path.
 *
 * One C interface, implemented twice:
 *   - oracle/port.cpp          (prefix port_): our CPU restatement of the reference algorithm
 *   - oracle/ref_harness.cpp   (prefix ref_):  the reference's OWN headers, compiled where they lie
 *                                              under /root/reference, driven through this interface
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * either library.
 *
 * Conventions: triangles are a soup of N x 9 doubles (v0,v1,v2) in scene.world insertion order;
 * prim ID := that insertion index.  Rays are n x 8 doubles (ox,oy,oz,dx,dy,dz,tmin,tmax).
 * A miss is prim 0xFFFFFFFF.  Films are W*H*3 doubles holding the per-pixel SUM over strata
 * (reference camera.h:55,117-124), not the average.
 */
#ifndef BDPT_ORACLE_API_H
#define BDPT_ORACLE_API_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_MAT_LAMBERTIAN = 0, ORC_MAT_METAL = 1, ORC_MAT_DIELECTRIC = 2, ORC_MAT_LIGHT = 3 };

typedef struct {
    uint32_t type;    /* ORC_MAT_* */
    uint32_t tex;     /* 0: rgb is the colour; k > 0: texture k-1 of scene_create_tex (lambertian albedo / light emission) */
    double   rgb[3];  /* albedo (lambertian, metal) or emission (light) */
    double   param;   /* metal: fuzz (already clamped to <=1), dielectric: refraction index */
} orc_material;

/* the 13 public camera fields of reference camera.h:26-41 that affect the image */
typedef struct {
    double aspect_ratio;
    int32_t image_width;
    int32_t samples_per_pixel;
    int32_t max_depth;
    int32_t pad;
    double background[3];
    double vfov;
    double lookfrom[3];
    double lookat[3];
    double vup[3];
    double defocus_angle;   /* must be <= 0: the YAML loader forces it off (scene_loader.h:463) */
    double focus_dist;
} orc_camera;

/* texture.h:14-87 / perlin.h (not reachable from the reference's loader: SURVEY F8, 8f-4) */
enum { ORC_TEX_SOLID = 0, ORC_TEX_CHECKER = 1, ORC_TEX_NOISE = 2, ORC_TEX_IMAGE = 3 };
typedef struct {
    uint32_t type, pad;
    double   a[3], b[3];   /* solid colour / checker even; checker odd */
    double   scale;
    uint64_t data_offset;  /* noise: 768 doubles randvec + 768 int32 perm_x|y|z; image: width*height*3 bytes as the
                              reference's rtw_image holds them (after its float round trip) */
    int32_t  width, height;
} orc_texture;

enum { ORC_RNG_MT19937 = 0,   /* the reference's global default-seeded mt19937 stream (main.h:28-32) */
       ORC_RNG_PHILOX  = 1 }; /* counter-based stream keyed (seed, pixel, stratum, subpath); port only  */

typedef struct {
    uint64_t rays;          /* world.hit queries */
    uint64_t node_visits;   /* bvh_node::hit entries (stats.h bvh_node_visits) */
    uint64_t aabb_tests;
    uint64_t aabb_hits;
    uint64_t tri_tests;
    uint64_t tri_hits;
    uint64_t nodes_built;
    uint64_t shadow_rays;   /* visible() queries among `rays` */
} orc_stats;

#ifndef ORC_PREFIX
#define ORC_PREFIX port_
#endif
#define ORC_CAT2(a, b) a##b
#define ORC_CAT(a, b) ORC_CAT2(a, b)
#define ORC(name) ORC_CAT(ORC_PREFIX, name)

void*    ORC(scene_create)(const double* verts, const uint32_t* mat_id, uint64_t n_tris,
                           const orc_material* mats, uint32_t n_mats);
/* The same with a texture table.  port: `data` holds every table / image as described above.  ref: builds the
 * reference's own checker_texture / noise_texture / image_texture objects; a noise texture makes its OWN perlin tables
 * (perlin.h:6-14, from the global RNG) and an image texture is loaded by the reference's rtw_image from a PPM file the
 * harness writes from `data` (raw RGB there) -- ref_texture_data then exports what the reference ended up with, so
 * that the port and the CUDA path can be given bit-identical tables. */
void*    ORC(scene_create_tex)(const double* verts, const uint32_t* mat_id, uint64_t n_tris, const orc_material* mats,
                               uint32_t n_mats, const orc_texture* tex, uint32_t n_tex, const void* data, uint64_t data_bytes);
/* texture k evaluated at n points: uvp = n x 5 doubles (u, v, px, py, pz) -> rgb = n x 3 doubles (texture.h value()) */
void     ORC(texture_value)(void* scene, uint32_t tex_index, uint64_t n, const double* uvp, double* rgb);
int      ref_texture_data(void* scene, uint32_t tex_index, void* out /* noise: 6144 + 3072 bytes; image: w*h*3 */);
/* constant_medium (materials/volumes/constant_medium.h:8-61) + isotropic (material.h:149-171), SURVEY 8f-4: n_media x 6
 * doubles {density, r, g, b, first boundary triangle, boundary triangle count} and the boundary soup (triangles x 9
 * doubles, closed meshes; NOT world prims).  port: media are tested after the triangle walk, in list order (oracle/port.cpp
 * hit_world); ref: real constant_medium objects around a bvh_node of the boundary, added to the world list. */
void     ORC(scene_set_media)(void* scene, const double* media6, uint32_t n_media, const double* boundary_verts);
double   port_det_log(double x);
/* port only: analytic spheres (an extension; the reference tessellates): n x 5 doubles {cx, cy, cz, radius, material index}, world
 * prims of a side list asked after the triangle walk (oracle/port.cpp: sphere_hit, hit_world). */
void     port_scene_set_spheres(void* scene, const double* spheres5, uint32_t n);
void     ORC(scene_destroy)(void* scene);
uint64_t ORC(bvh_node_count)(void* scene);
/* DFS leaf order: order[k] = prim ID of the k-th triangle in the builder's final array. */
void     ORC(bvh_leaf_order)(void* scene, uint32_t* order);
/* closest hit; outputs may be NULL. nthreads<=1 runs serially. */
void     ORC(trace)(void* scene, const double* rays, uint64_t n, uint32_t* prim, double* t,
                    double* u, double* v, int nthreads);
/* camera.h:425-438 on n segments a->b (each n x 3 doubles); out[i]=1 when visible */
void     ORC(visible)(void* scene, const double* a, const double* b, uint64_t n, uint8_t* out,
                      int nthreads);
/* camera.h:43-145 minus threads/PNG: rows [row0,row1), every stratum; film is W*H*3 (only those rows
 * written). rng_mode MT19937 requires nthreads==1 for a reproducible stream. Returns image height. */
int      ORC(render)(void* scene, const orc_camera* cam, int rng_mode, uint64_t seed, int row0,
                     int row1, double* film, int nthreads);
/* port only: the same with the estimator chosen.  ORC_WEIGHT_REF_SUM = the reference's unweighted sum
 * (camera.h:294-323); ORC_WEIGHT_MIS = balance-heuristic BDPT with light-tracing splats (film must be zeroed by
 * the caller: splats ADD into pixels outside [row0,row1)).
 * ref only: render_pt = the reference's own unbiased path tracer, camera::path_trace_color (camera.h:255-292), over the
 * same sample loop -- the yardstick the MIS estimator is validated against. */
enum { ORC_WEIGHT_REF_SUM = 0, ORC_WEIGHT_MIS = 1 };
int      port_render_ex(void* scene, const orc_camera* cam, int rng_mode, uint64_t seed, int row0, int row1, double* film,
                        int nthreads, int weighting);
int      ref_render_pt(void* scene, const orc_camera* cam, uint64_t seed, int row0, int row1, double* film);
void     ORC(stats_get)(orc_stats* out);
void     ORC(stats_reset)(void);
/* wpng.h:14-35 */
void     ORC(colors_to_rgb8)(const double* film, int W, int H, int samples_per_pixel, uint8_t* out);

#ifdef __cplusplus
}
#endif
#endif
