// The fp32 pre-filter of the triangle tests: plain arithmetic shared by the warp-synchronous walk (traverse_w32.cuh),
// the per-lane walk (traverse_walk.cuh: Walk::leaf_tests rejects with it before the fp64 test) and the host
// (bdpt_test_filter32, tests/test_host.py hammers it against the exact test without a GPU).
#pragma once
#include <cmath>
#include <cstring>

#include "scene.cuh"

namespace bdpt {

constexpr float kFilterK = 7.62939453125e-06f;   // 2^-17

// One triangle of the fast tree's leaf order in fp32 for the pre-filter: 48 B = 3 x LDG.128.
//   q0 = v0.xyz, e1.x | q1 = e1.yz, e2.xy | q2 = e2.z, |e1| (rounded up), |e2| (rounded up), prim ID (bits)
// Built on the device at upload from fast_tris / fast_prims (derive_walk32, api.cu).
struct __align__(16) Tri32 { float4 q0, q1, q2; };
static_assert(sizeof(Tri32) == 48, "Tri32 must be 48 bytes");

// ---- the parts of the walk that are plain arithmetic, shared with the host so that a CPU test can hammer the
// pre-filter against the exact fp64 test without a GPU (bdpt_test_filter32, tests/test_host.py)
BDPT_HD float f32_down(double v) {
#ifdef __CUDA_ARCH__
    return __double2float_rd(v);
#else
    float f = float(v);
    if (double(f) > v) f = std::nextafterf(f, -INFINITY);
    return f;
#endif
}
BDPT_HD float f32_up(double v) {
#ifdef __CUDA_ARCH__
    return __double2float_ru(v);
#else
    float f = float(v);
    if (double(f) < v) f = std::nextafterf(f, INFINITY);
    return f;
#endif
}
BDPT_HD float fma32(float a, float b, float c) {
#ifdef __CUDA_ARCH__
    return __fmaf_rn(a, b, c);
#else
    return std::fmaf(a, b, c);
#endif
}
BDPT_HD float div32(float a, float b) {   // device: the 2-ulp approximate quotient; its error is inside the 1e-6 inflation below
#ifdef __CUDA_ARCH__
    return __fdividef(a, b);
#else
    return a / b;
#endif
}

BDPT_HD Tri32 make_tri32(const TriGeom& t, uint32_t prim) {
    Tri32 r;
    r.q0.x = float(t.v0x); r.q0.y = float(t.v0y); r.q0.z = float(t.v0z); r.q0.w = float(t.e1x);
    r.q1.x = float(t.e1y); r.q1.y = float(t.e1z); r.q1.z = float(t.e2x); r.q1.w = float(t.e2y);
    const double n1 = sqrt(t.e1x * t.e1x + t.e1y * t.e1y + t.e1z * t.e1z), n2 = sqrt(t.e2x * t.e2x + t.e2y * t.e2y + t.e2z * t.e2z);
    r.q2.x = float(t.e2z); r.q2.y = f32_up(n1 * 1.0000001); r.q2.z = f32_up(n2 * 1.0000001);
#ifdef __CUDA_ARCH__
    r.q2.w = __uint_as_float(prim);
#else
    std::memcpy(&r.q2.w, &prim, 4);
#endif
    return r;
}

// the fp32 image of a ray: what the pre-filter reads
struct Ray32 {
    float ox, oy, oz, dx, dy, dz;
    float tmin_t, best_t;        // conservative fp32 images of [tmin, tmax]: rounded down / up
    float tmin_hi, tmax_lo;      // tmin rounded up, tmax rounded down: "certainly inside the interval"
    float kD, kU, kT;            // pre-filter error scales: K |d|, K rho |d|, K rho
};
BDPT_HD Ray32 make_ray32(const double o[3], const double d[3], double tmin, double tmax, float vmax32) {
    Ray32 r;
    r.ox = float(o[0]); r.oy = float(o[1]); r.oz = float(o[2]);
    r.dx = float(d[0]); r.dy = float(d[1]); r.dz = float(d[2]);
    r.tmin_t = f32_down(tmin); r.tmin_hi = f32_up(tmin);
    r.best_t = f32_up(tmax); r.tmax_lo = f32_down(tmax);
    const float dn = f32_up(sqrt(double(r.dx) * r.dx + double(r.dy) * r.dy + double(r.dz) * r.dz) * 1.000001);
    const float rho = fmaxf(fabsf(r.ox), fmaxf(fabsf(r.oy), fabsf(r.oz))) * 1.000001f + vmax32;
    r.kD = kFilterK * dn; r.kU = kFilterK * rho * dn; r.kT = kFilterK * rho;
    return r;
}

// The conservative pre-filter.  0: the reference's test rejects this triangle (or it cannot matter: beyond the
// current bound / before tmin by more than every flagging band).  1: candidate.  2: candidate that the reference
// certainly accepts, with t <= t_hi (closest) / certainly inside (tmin, tmax) (ANY).
template <bool ANY>
BDPT_HD int filter32(const Ray32& R, float best_t, const float4& q0, const float4& q1, const float4& q2, float& t_hi) {
    const float e1x = q0.w, e1y = q1.x, e1z = q1.y, e2x = q1.z, e2y = q1.w, e2z = q2.x, n1 = q2.y, n2 = q2.z;
    const float px = fma32(R.dy, e2z, -(R.dz * e2y)), py = fma32(R.dz, e2x, -(R.dx * e2z)), pz = fma32(R.dx, e2y, -(R.dy * e2x));
    const float det = fma32(e1x, px, fma32(e1y, py, e1z * pz));
    const float tx = R.ox - q0.x, ty = R.oy - q0.y, tz = R.oz - q0.z;
    const float U = fma32(tx, px, fma32(ty, py, tz * pz));
    const float qx = fma32(ty, e1z, -(tz * e1y)), qy = fma32(tz, e1x, -(tx * e1z)), qz = fma32(tx, e1y, -(ty * e1x));
    const float V = fma32(R.dx, qx, fma32(R.dy, qy, R.dz * qz));
    const float TN = fma32(e2x, qx, fma32(e2y, qy, e2z * qz));
    const float m12 = n1 * n2;
    const float eD = fmaxf(R.kD * m12, 1e-30f), eU = R.kU * n2, eV = R.kU * n1, eT = R.kT * m12;
    const float adet = fabsf(det);
    // one exit: the lanes of a warp take this function together, an early return would split them for the rest of it
    const bool sign_known = adet > eD;   // else (grazing, sliver, NaN): only the exact test can tell
    const bool neg = det < 0.f;
    const float Us = neg ? -U : U, Vs = neg ? -V : V, Ts = neg ? -TN : TN;
    const float eS = eD + eU + eV;
    const bool outside = Us < -eU || Vs < -eV || Us + Vs > adet + eS;      // u < 0, v < 0 or u + v > 1, beyond doubt
    const bool beyond = Ts - eT > best_t * (adet + eD);                     // t > best, beyond doubt (best_t = inf: never)
    const bool before = Ts + eT < R.tmin_t * (adet - eD);                   // t < tmin, beyond doubt
    int res = (sign_known && (outside || beyond || before)) ? 0 : 1;
    // certainly inside the triangle, |det| certainly >= 1e-8 (triangle.h:47), t certainly >= tmin?
    if (res == 1 && sign_known && Us > eU && Vs > eV && Us + Vs < adet - eS && adet - eD > 2e-8f) {
        const float lo = div32(Ts - eT, adet + eD) * 0.999999f, hi = div32(Ts + eT, adet - eD) * 1.000001f;
        if (lo >= R.tmin_hi && (!ANY || hi <= R.tmax_lo)) { t_hi = hi; res = 2; }
    }
    return res;
}

// Reject-only form for the per-lane walk: true when filter32 would return 0 -- the reference's fp64 test rejects this
// triangle, or its t lies beyond `best_t` / before `tmin_t` by more than every flagging band (the margins are ~1e-5 of
// the ray length against bands of 1e-10).  (ox..dz): the ray rounded to fp32; kD, kU, kT: make_ray32's error scales.
BDPT_HD bool prefilter_rejects(float ox, float oy, float oz, float dx, float dy, float dz, float tmin_t, float best_t, float kD,
                               float kU, float kT, const float4& q0, const float4& q1, const float4& q2) {
    const float e1x = q0.w, e1y = q1.x, e1z = q1.y, e2x = q1.z, e2y = q1.w, e2z = q2.x, n1 = q2.y, n2 = q2.z;
    const float px = fma32(dy, e2z, -(dz * e2y)), py = fma32(dz, e2x, -(dx * e2z)), pz = fma32(dx, e2y, -(dy * e2x));
    const float det = fma32(e1x, px, fma32(e1y, py, e1z * pz));
    const float tx = ox - q0.x, ty = oy - q0.y, tz = oz - q0.z;
    const float U = fma32(tx, px, fma32(ty, py, tz * pz));
    const float qx = fma32(ty, e1z, -(tz * e1y)), qy = fma32(tz, e1x, -(tx * e1z)), qz = fma32(tx, e1y, -(ty * e1x));
    const float V = fma32(dx, qx, fma32(dy, qy, dz * qz));
    const float TN = fma32(e2x, qx, fma32(e2y, qy, e2z * qz));
    const float m12 = n1 * n2;
    const float eD = fmaxf(kD * m12, 1e-30f), eU = kU * n2, eV = kU * n1, eT = kT * m12;
    const float adet = fabsf(det);
    const bool sign_known = adet > eD;
    const bool neg = det < 0.f;
    const float Us = neg ? -U : U, Vs = neg ? -V : V, Ts = neg ? -TN : TN;
    const float eS = eD + eU + eV;
    const bool outside = Us < -eU || Vs < -eV || Us + Vs > adet + eS;
    const bool beyond = Ts - eT > best_t * (adet + eD);
    const bool before = Ts + eT < tmin_t * (adet - eD);
    return sign_known && (outside || beyond || before);
}

}  // namespace bdpt
