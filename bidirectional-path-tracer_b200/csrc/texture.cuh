// texture::value(u, v, p) (src/materials/textures/texture.h:14-87, perlin.h:4-93) for the shading kernels: solid colour,
// checker, Perlin-noise marble and image lookup, in the reference's operation order (fp64, no FMA contraction: this
// is compiled with -fmad=false like everything that must round as the CPU build does).  Not reachable from the
// reference's own loader: the scene surface is extended by bdpt_scene_set_textures.
#pragma once
#include "scene.cuh"

namespace bdpt {

// perlin::noise, perlin.h:16-34 + perlin_interp :77-91.  tables: 256 x 3 doubles randvec, then perm_x, perm_y, perm_z.
__device__ __forceinline__ double perlin_noise(const unsigned char* __restrict__ tables, D3 p) {
    const double* randvec = reinterpret_cast<const double*>(tables);
    const int* perm = reinterpret_cast<const int*>(tables + 256 * 3 * sizeof(double));
    const double fx = floor(p.x), fy = floor(p.y), fz = floor(p.z);
    const double u = p.x - fx, v = p.y - fy, w = p.z - fz;
    const int i = int(fx), j = int(fy), k = int(fz);
    const double uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
    double accum = 0.0;
#pragma unroll
    for (int di = 0; di < 2; ++di)
#pragma unroll
        for (int dj = 0; dj < 2; ++dj)
#pragma unroll
            for (int dk = 0; dk < 2; ++dk) {
                const int idx = __ldg(perm + ((i + di) & 255)) ^ __ldg(perm + 256 + ((j + dj) & 255)) ^ __ldg(perm + 512 + ((k + dk) & 255));
                const D3 c = mk(__ldg(randvec + 3 * idx), __ldg(randvec + 3 * idx + 1), __ldg(randvec + 3 * idx + 2));
                const D3 weight_v = mk(u - di, v - dj, w - dk);
                accum += (di * uu + (1 - di) * (1 - uu)) * (dj * vv + (1 - dj) * (1 - vv)) * (dk * ww + (1 - dk) * (1 - ww)) * dot(c, weight_v);
            }
    return accum;
}

// perlin::turb, perlin.h:36-49
__device__ __forceinline__ double perlin_turb(const unsigned char* __restrict__ tables, D3 p, int depth) {
    double accum = 0.0, weight = 1.0;
    D3 tp = p;
    for (int i = 0; i < depth; ++i) {
        accum += weight * perlin_noise(tables, tp);
        weight *= 0.5;
        tp = mk(tp.x * 2, tp.y * 2, tp.z * 2);   // vec3::operator*=(double), vec3.h:30-35
    }
    return fabs(accum);
}

__device__ __forceinline__ D3 tex_value(const SceneView& sc, uint32_t tex_id, double u, double v, D3 p) {
    const Tex t = sc.tex[tex_id];
    if (t.type == BDPT_TEX_CHECKER) {   // texture.h:37-45
        const double inv_scale = 1.0 / t.scale;
        const int xi = int(floor(inv_scale * p.x)), yi = int(floor(inv_scale * p.y)), zi = int(floor(inv_scale * p.z));
        return ((xi + yi + zi) % 2 == 0) ? mk(t.ar, t.ag, t.ab) : mk(t.br, t.bg, t.bb);
    }
    if (t.type == BDPT_TEX_NOISE)       // texture.h:79-81
        return mk(.5, .5, .5) * (1 + sin(t.scale * p.z + 10 * perlin_turb(sc.tex_data + t.data, p, 7)));
    if (t.type == BDPT_TEX_IMAGE) {     // texture.h:57-71, rtw_stb_image.h:63-88
        if (t.height <= 0) return mk(0, 1, 1);
        u = u < 0.0 ? 0.0 : (u > 1.0 ? 1.0 : u);                  // interval(0,1).clamp
        v = 1.0 - (v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v));
        int i = int(u * t.width), j = int(v * t.height);
        i = i < 0 ? 0 : (i < t.width ? i : t.width - 1);
        j = j < 0 ? 0 : (j < t.height ? j : t.height - 1);
        const unsigned char* px = sc.tex_data + t.data + (size_t(j) * size_t(t.width) + size_t(i)) * 3;
        const double color_scale = 1.0 / 255.0;
        return mk(color_scale * px[0], color_scale * px[1], color_scale * px[2]);
    }
    return mk(t.ar, t.ag, t.ab);        // solid_color, texture.h:14-26
}

// The colour of material m at a hit: its own rgb, or its texture at (u, v, p).  (u, v) are needed by image textures
// only and cost a second Moller-Trumbore: `uv` computes them on demand, with the very arithmetic of the hit test.
template <class UV>
__device__ __forceinline__ D3 mat_colour(const SceneView& sc, const Mat& m, D3 p, UV uv) {
    if (m.tex == 0u || sc.tex == nullptr || m.tex > sc.n_tex) return mk(m.r, m.g, m.b);
    double u = 0.0, v = 0.0;
    if (sc.tex[m.tex - 1u].type == BDPT_TEX_IMAGE) uv(u, v);
    return tex_value(sc, m.tex - 1u, u, v, p);
}

}  // namespace bdpt
