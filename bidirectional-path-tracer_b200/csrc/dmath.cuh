// fp64 vector math for the reference-exact kernels.
// Operation ORDER follows the reference's vec3 (src/core/vec3.h:7-159) so that, compiled with
// -fmad=false (no FMA contraction: the reference's flagless x86-64 build has none), every IEEE add,
// mul, div and sqrt rounds exactly as on the CPU.  E.g. v/t is (1/t)*v (vec3.h:40-42,98-100).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define BDPT_HD __host__ __device__ __forceinline__

namespace bdpt {

struct D3 { double x, y, z; };

BDPT_HD D3 mk(double x, double y, double z) { D3 r; r.x = x; r.y = y; r.z = z; return r; }
BDPT_HD D3 operator+(D3 a, D3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
BDPT_HD D3 operator-(D3 a, D3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
BDPT_HD D3 operator-(D3 a) { return mk(-a.x, -a.y, -a.z); }
BDPT_HD D3 operator*(D3 a, D3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
BDPT_HD D3 operator*(double t, D3 v) { return mk(t * v.x, t * v.y, t * v.z); }   // vec3.h:90-92
BDPT_HD D3 operator*(D3 v, double t) { return t * v; }                           // vec3.h:94-96
BDPT_HD D3 operator/(D3 v, double t) { return (1 / t) * v; }                     // vec3.h:98-100
BDPT_HD double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }     // vec3.h:102-106
BDPT_HD D3 cross(D3 u, D3 v) {                                                   // vec3.h:108-112
    return mk(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
BDPT_HD double len2(D3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }          // vec3.h:43-45
BDPT_HD double len(D3 v) { return sqrt(len2(v)); }
BDPT_HD D3 unit(D3 v) { return v / len(v); }                                     // vec3.h:114-116
BDPT_HD D3 reflect(D3 v, D3 n) { return v - 2 * dot(v, n) * n; }                 // vec3.h:138-140
BDPT_HD D3 refract(D3 uv, D3 n, double etai_over_etat) {                         // vec3.h:142-147
    double cos_theta = fmin(dot(-uv, n), 1.0);
    D3 perp = etai_over_etat * (uv + cos_theta * n);
    D3 par = -sqrt(fabs(1.0 - len2(perp))) * n;
    return perp + par;
}

// onb.h:4-24
struct Onb {
    D3 a0, a1, a2;
    BDPT_HD explicit Onb(D3 n) {
        a2 = unit(n);
        D3 a = (fabs(a2.x) > 0.9) ? mk(0, 1, 0) : mk(1, 0, 0);
        a1 = unit(cross(a2, a));
        a0 = cross(a2, a1);
    }
    BDPT_HD D3 transform(D3 v) const { return (v.x * a0) + (v.y * a1) + (v.z * a2); }
};

}  // namespace bdpt
