// Random numbers and sampling helpers of the integrator kernels.
// Replaces random_double()/random_double(min,max) (src/main.h:28-37: one global mt19937 shared by all
// threads) with a counter-based generator, and restates random_unit_vector /
// random_cosine_direction (src/core/vec3.h:128-159) on top of it.
#pragma once
#include "dmath.cuh"

namespace bdpt {

// Philox4x32-10. key = (seed_lo ^ stream, seed_hi), counter = (sample_lo, sample_hi, block, 0x5EED).
// One block yields two doubles in [0,1): ((hi:lo) >> 11) * 2^-53.  stream 0 = camera subpath,
// 1 = light subpath; `sample` is the global (pixel, stratum) index, so the image does not depend on how
// samples are sharded over batches or GPUs.
__device__ __forceinline__ void philox_block(uint64_t seed, uint32_t stream, uint64_t sample, uint32_t blk,
                                             uint32_t out[4]) {
    uint32_t k0 = uint32_t(seed) ^ stream, k1 = uint32_t(seed >> 32);
    uint32_t c0 = uint32_t(sample), c1 = uint32_t(sample >> 32), c2 = blk, c3 = 0x5EEDu;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        const uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

struct Rng {
    uint64_t seed, sample;
    uint32_t stream, draws;
    uint32_t cache[2];
    bool have;
    __device__ __forceinline__ void init(uint64_t seed_, uint32_t stream_, uint64_t sample_, uint32_t draws_) {
        seed = seed_; stream = stream_; sample = sample_; draws = draws_; have = false;
    }
    __device__ __forceinline__ double next() {
        uint32_t lo, hi;
        if ((draws & 1u) && have) { lo = cache[0]; hi = cache[1]; have = false; }
        else {
            uint32_t o[4];
            philox_block(seed, stream, sample, draws >> 1, o);
            if (draws & 1u) { lo = o[2]; hi = o[3]; }
            else { lo = o[0]; hi = o[1]; cache[0] = o[2]; cache[1] = o[3]; have = true; }
        }
        ++draws;
        return double(((uint64_t(hi) << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
    }
    __device__ __forceinline__ double next(double mn, double mx) { return mn + (mx - mn) * next(); }  // main.h:34-37
};

// sin/cos as a fixed sequence of IEEE adds and multiplies (no FMA: -fmad=false), so that the result does
// not depend on which math library evaluates it: Cody-Waite reduction by pi/2 followed by the classic
// degree-13 / degree-12 minimax kernels on [-pi/4, pi/4]; error below 1 ulp.  phi in [0, 2*pi].
// Why: camera::visible()'s far endpoint is a rounding knife edge, so one ulp of difference
// in a sampled direction flips connection decisions; with this the kernels and the CPU oracle's
// counter-based mode agree sample by sample.
__device__ __forceinline__ void det_sincos(double phi, double& s_out, double& c_out) {
    const int k = int(phi * 0.63661977236758134308 + 0.5);
    const double y = (phi - k * 1.57079632673412561417e+00) - k * 6.07710050650619224932e-11;
    const double z = y * y;
    const double s = y + y * z * (-1.66666666666666324348e-01 + z * (8.33333333332248946124e-03 + z * (-1.98412698298579493134e-04 +
                     z * (2.75573137070700676789e-06 + z * (-2.50507602534068634195e-08 + z * 1.58969099521155010221e-10)))));
    const double c = 1.0 - 0.5 * z + z * z * (4.16666666666666019037e-02 + z * (-1.38888888888741095749e-03 + z * (2.48015872894767294178e-05 +
                     z * (-2.75573143513906633035e-07 + z * (2.08757232129817482790e-09 + z * -1.13596475577881948265e-11)))));
    switch (k & 3) {
        case 0: s_out = s; c_out = c; break;
        case 1: s_out = c; c_out = -s; break;
        case 2: s_out = -s; c_out = -c; break;
        default: s_out = -c; c_out = s; break;
    }
}
// log as a fixed sequence of IEEE +, -, *, / (fdlibm's __ieee754_log on mantissa and exponent; -fmad=false): the draw inside
// constant_medium::hit (constant_medium.h:44) must come out bit for bit as in the oracle port (oracle/port.cpp: det_log, the same
// constants in the same order), for the same reason as det_sincos.  x in [0, 1): 0 -> -inf; no subnormals (draws are k * 2^-53).
__device__ __forceinline__ double det_log(double x) {
    if (!(x > 0.0)) return -__longlong_as_double(0x7FF0000000000000LL);
    unsigned long long bits = (unsigned long long)__double_as_longlong(x);
    int e = int((bits >> 52) & 0x7FFull) - 1023;
    bits = (bits & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
    double m = __longlong_as_double((long long)bits);
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    const double f = m - 1.0;
    const double s_ = f / (2.0 + f);
    const double z = s_ * s_, w = z * z;
    const double t1 = w * (3.999999999940941908e-01 + w * (2.222219843214978396e-01 + w * 1.531383769920937332e-01));
    const double t2 = z * (6.666666666666735130e-01 + w * (2.857142874366239149e-01 + w * (1.818357216161805012e-01 + w * 1.479819860511658591e-01)));
    const double R = t2 + t1;
    const double hfsq = 0.5 * f * f;
    const double dk = double(e);
    return dk * 6.93147180369123816490e-01 - ((hfsq - (s_ * (hfsq + R) + dk * 1.90821492927058770002e-10)) - f);
}
__device__ __forceinline__ double det_pow5(double x) { const double x2 = x * x; const double x4 = x2 * x2; return x4 * x; }

constexpr double kPi = 3.1415926535897932385;  // main.h:20

// vec3.h:128-135: rejection sampling in the cube; the reference's argument evaluation order draws z, y, x.
__device__ __forceinline__ D3 random_unit_vector(Rng& r) {
    for (;;) {
        const double z = r.next(-1, 1);
        const double y = r.next(-1, 1);
        const double x = r.next(-1, 1);
        const D3 p = mk(x, y, z);
        const double lensq = len2(p);
        if (1e-160 < lensq && lensq <= 1) return p / sqrt(lensq);
    }
}

// vec3.h:149-159
__device__ __forceinline__ D3 random_cosine_direction(Rng& r) {
    const double r1 = r.next(), r2 = r.next();
    const double phi = 2 * kPi * r1;
    double sn, cs;
    det_sincos(phi, sn, cs);
    const double x = cs * sqrt(r2);
    const double y = sn * sqrt(r2);
    const double z = sqrt(1 - r2);
    return mk(x, y, z);
}

}  // namespace bdpt
