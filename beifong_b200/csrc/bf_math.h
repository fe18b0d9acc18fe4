// beifong_b200 — fp32 arithmetic shared by the host scene code and the sm_100a kernels.
//
// The reference (JacobMackay/beifong, a Mitsuba 2 fork) computes everything on this path in fp32
// through Enoki's scalar wrappers (mitsuba.conf: "float": "float").  To make the CUDA path agree
// with the reference's scalar CPU variant ray by ray, every operation here is an IEEE fp32
// operation in a fixed order: the translation unit is compiled with `-fmad=false` (device) and
// without FMA contraction (host), and a fused multiply-add is issued only where the reference
// writes fmadd/fmsub/fnmadd or where Enoki's dot()/cross() use an FMA chain.  No fast-math
// intrinsics are used on the parity path.
//
// Reference: include/mitsuba/core/math.h:17-131, vector.h:116-135, frame.h:17-118,
// transform.h:85-160, warp.h:54-90,268-283,325-349, random.h:74-116 (+ enoki/random.h PCG32).
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define BF_HD __host__ __device__ __forceinline__
#define BF_D __device__ __forceinline__
#else
#define BF_HD inline
#define BF_D inline
#endif

namespace bf {

// include/mitsuba/core/math.h:17-41
#define BF_PI 3.14159265358979323846f
#define BF_TWO_PI 6.28318530717958647692f
#define BF_INV_PI 0.31830988618379067154f
#define BF_INV_SQRT_PI 0.56418958354775628695f
#define BF_EPS 5.9604644775390625e-08f                 /* numeric_limits<float>::epsilon() / 2 */
#define BF_RAY_EPS (5.9604644775390625e-08f * 1500.f)  /* math.h:37 */
#define BF_SHADOW_EPS (5.9604644775390625e-08f * 1500.f * 10.f)

BF_HD float bf_inf() {
#if defined(__CUDA_ARCH__)
    return __int_as_float(0x7f800000);
#else
    return INFINITY;
#endif
}
BF_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
BF_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

BF_HD float fmadd(float a, float b, float c) { return fmaf(a, b, c); }
BF_HD float fmsub(float a, float b, float c) { return fmaf(a, b, -c); }
BF_HD float fnmadd(float a, float b, float c) { return fmaf(-a, b, c); }
BF_HD float rcp(float x) { return 1.f / x; }
BF_HD float rsqrt_(float x) { return 1.f / sqrtf(x); }
BF_HD float sqr(float x) { return x * x; }
BF_HD float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.f)); }
BF_HD float sign_(float x) { return u2f(0x3f800000u | (f2u(x) & 0x80000000u)); }
BF_HD float mulsign(float a, float b) { return u2f(f2u(a) ^ (f2u(b) & 0x80000000u)); }
BF_HD float mulsign_neg(float a, float b) { return u2f(f2u(a) ^ ((~f2u(b)) & 0x80000000u)); }
BF_HD float lerp_(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }
BF_HD float clamp_(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

struct V3 {
    float x, y, z;
    BF_HD V3() : x(0.f), y(0.f), z(0.f) {}
    BF_HD V3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
struct V2 {
    float x, y;
    BF_HD V2() : x(0.f), y(0.f) {}
    BF_HD V2(float a, float b) : x(a), y(b) {}
};
BF_HD V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
BF_HD V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
BF_HD V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
BF_HD V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
BF_HD V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
BF_HD V3 operator/(V3 a, float s) { float r = rcp(s); return V3(a.x * r, a.y * r, a.z * r); }
// Enoki's dot_ is an FMA chain over the lanes; cross() is shuffle + fmsub.
BF_HD float dot(V3 a, V3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
BF_HD float abs_dot(V3 a, V3 b) { return fabsf(dot(a, b)); }
BF_HD float squared_norm(V3 a) { return dot(a, a); }
BF_HD float norm(V3 a) { return sqrtf(dot(a, a)); }
BF_HD V3 normalize(V3 a) { return a * rsqrt_(squared_norm(a)); }
BF_HD V3 cross(V3 a, V3 b) {
    return V3(fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x));
}
BF_HD V3 fmadd3(V3 a, float s, V3 c) { return V3(fmadd(a.x, s, c.x), fmadd(a.y, s, c.y), fmadd(a.z, s, c.z)); }
BF_HD float hmax_abs(V3 p) { return fmaxf(fmaxf(fabsf(p.x), fabsf(p.y)), fabsf(p.z)); }

// "Jacob functions", include/mitsuba/core/math.h:68-131
BF_HD float jabs(float x) { return x >= 0.f ? x : -x; }
BF_HD float sinc(float x) { return jabs(x) > BF_EPS ? sinf(x) / x : 1.f; }
BF_HD float tri(float x) { return jabs(x) < 0.5f ? 1.f - 2.f * jabs(x) : 0.f; }
BF_HD float rect(float x) { return jabs(x) < 0.5f ? 1.f : 0.f; }
// math.h:108-123 is a literal subtraction loop, not fmodf; the rounding of every step is kept.
// The loop is bounded like the oracle's (the reference would spin for ever on such input).
BF_HD float fmodulo(float a, float b) {
    float result = jabs(a);
    const float ab = jabs(b);
    uint32_t guard = 0;
    while (result - ab >= BF_EPS) {
        result -= ab;
        if (++guard > (1u << 26)) break;
    }
    result = (a < 0.f) ? ab - result : result;
    result += (b < 0.f) ? b : 0.f;
    return result;
}
BF_HD float wchirp(float t, float f, float w, float a) {
    return 2 * a * a * w * tri(t / w) * sinc(BF_TWO_PI * f * w * tri(t / w));
}

// include/mitsuba/core/vector.h:116-135
BF_HD void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = sign_(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    s = V3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = V3(b, sign + sqr(n.y) * a, -n.y);
}
struct Frame {
    V3 s, t, n;
    BF_HD Frame() {}
    BF_HD explicit Frame(V3 v) : n(v) { coordinate_system(v, s, t); }
    BF_HD Frame(V3 s_, V3 t_, V3 n_) : s(s_), t(t_), n(n_) {}
    BF_HD V3 to_local(V3 v) const { return V3(dot(v, s), dot(v, t), dot(v, n)); }
    BF_HD V3 to_world(V3 v) const { return s * v.x + t * v.y + n * v.z; }
};

// Row-major 4x4, include/mitsuba/core/transform.h:85-160.
struct M4 { float m[4][4]; };
BF_HD V3 xf_point_affine(const M4 &M, V3 p) {  // transform.h:91-99
    float r[3];
#pragma unroll
    for (int j = 0; j < 3; ++j)
        r[j] = fmadd(M.m[j][2], p.z, fmadd(M.m[j][1], p.y, fmadd(M.m[j][0], p.x, M.m[j][3])));
    return V3(r[0], r[1], r[2]);
}
BF_HD V3 xf_point(const M4 &M, V3 p) {         // transform.h:105-112 (projective: true division)
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
        r[j] = fmadd(M.m[j][2], p.z, fmadd(M.m[j][1], p.y, fmadd(M.m[j][0], p.x, M.m[j][3])));
    float iw = rcp(r[3]);
    return V3(r[0] * iw, r[1] * iw, r[2] * iw);
}
BF_HD V3 xf_vector(const M4 &M, V3 v) {        // transform.h:119-127
    float r[3];
#pragma unroll
    for (int j = 0; j < 3; ++j)
        r[j] = fmadd(M.m[j][2], v.z, fmadd(M.m[j][1], v.y, M.m[j][0] * v.x));
    return V3(r[0], r[1], r[2]);
}
BF_HD V3 xf_normal(const M4 &Minv, V3 n) {     // transform.h:134-142 (inverse transpose)
    float r[3];
#pragma unroll
    for (int j = 0; j < 3; ++j)
        r[j] = fmadd(Minv.m[2][j], n.z, fmadd(Minv.m[1][j], n.y, Minv.m[0][j] * n.x));
    return V3(r[0], r[1], r[2]);
}

// PCG32: public O'Neill generator as wrapped by enoki/random.h (submodule absent from the
// reference tree; pinned commit 9e51ab8).  TEA: include/mitsuba/core/random.h:74-116.
#define BF_PCG32_DEFAULT_STATE 0x853c49e6748fea9bULL
#define BF_PCG32_DEFAULT_STREAM 0xda3e39cb94b95bdbULL
#define BF_PCG32_MULT 0x5851f42d4c957f2dULL
struct PCG32 {
    uint64_t state, inc;
    BF_HD uint32_t next_u32() {
        uint64_t old = state;
        state = old * BF_PCG32_MULT + inc;
        uint32_t xorshifted = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    BF_HD void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u; next_u32(); state += initstate; next_u32();
    }
    BF_HD float next_float32() { return u2f((next_u32() >> 9) | 0x3f800000u) - 1.f; }
};
// src/librender/sampler.cpp:89-92 instantiates sample_tea_64 with UInt64 arguments.
BF_HD uint64_t sample_tea_64_wide(uint64_t v0, uint64_t v1) {
    uint64_t sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v0 + (v1 << 32);
}

// include/mitsuba/core/warp.h:54-90, 325-349, 268-283
BF_HD V2 square_to_uniform_disk_concentric(V2 sample) {
    float x = fmsub(2.f, sample.x, 1.f), y = fmsub(2.f, sample.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), quadrant_1_or_3 = fabsf(x) < fabsf(y);
    float r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    float phi = .25f * BF_PI * rp / r;
    if (quadrant_1_or_3) phi = .5f * BF_PI - phi;
    if (is_zero) phi = 0.f;
    float s = sinf(phi), c = cosf(phi);
    return V2(r * c, r * s);
}
BF_HD V3 square_to_cosine_hemisphere(V2 sample) {
    V2 p = square_to_uniform_disk_concentric(sample);
    float sq = fmadd(p.y, p.y, p.x * p.x);
    return V3(p.x, p.y, safe_sqrt(1.f - sq));
}
BF_HD float square_to_cosine_hemisphere_pdf(V3 v) { return BF_INV_PI * v.z; }

// "Approximating the erfinv function", M. Giles — the single-precision polynomial Enoki uses.
BF_HD float erfinv_(float x) {
    float w = -logf((1.f - x) * (1.f + x)), p;
    if (w < 5.f) {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = fmadd(p, w, 3.43273939e-07f); p = fmadd(p, w, -3.5233877e-06f);
        p = fmadd(p, w, -4.39150654e-06f); p = fmadd(p, w, 0.00021858087f);
        p = fmadd(p, w, -0.00125372503f); p = fmadd(p, w, -0.00417768164f);
        p = fmadd(p, w, 0.246640727f); p = fmadd(p, w, 1.50140941f);
    } else {
        w = sqrtf(w) - 3.f;
        p = -0.000200214257f;
        p = fmadd(p, w, 0.000100950558f); p = fmadd(p, w, 0.00134934322f);
        p = fmadd(p, w, -0.00367342844f); p = fmadd(p, w, 0.00573950773f);
        p = fmadd(p, w, -0.0076224613f); p = fmadd(p, w, 0.00943887047f);
        p = fmadd(p, w, 1.00167406f); p = fmadd(p, w, 2.83297682f);
    }
    return p * x;
}

}  // namespace bf
