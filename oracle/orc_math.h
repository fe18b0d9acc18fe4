// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Scalar fp32 restatement of the arithmetic helpers the reference's radar hot path uses.
// Every function cites the reference file:line it follows (paths under /root/reference).
//
// Third-party dependency absent from the reference tree: mitsuba-renderer/enoki @
// 9e51ab8680dd354bc6a5c06eb03653ff73871bae (ext/enoki is an empty submodule).  In the
// reference's scalar variants Enoki's elementary operations are thin wrappers over IEEE
// fp32 arithmetic; this header fixes ONE canonical arithmetic for them, compiled with
// -ffp-contract=off so that an FMA happens exactly where the reference writes
// fmadd/fmsub/fnmadd:
//   rcp(x) = 1/x, rsqrt(x) = 1/sqrt(x), safe_sqrt(x) = sqrt(max(x,0)),
//   dot(a,b) = fma(a.z,b.z, fma(a.y,b.y, a.x*b.x))   (Enoki's generic dot_ fma chain),
//   cross(a,b)_x = fmsub(a.y,b.z, a.z*b.y)           (Enoki's shuffle/fmsub cross),
//   v / s = v * rcp(s)                                (Enoki's array / scalar),
//   normalize(v) = v * rsqrt(dot(v,v)),
//   sin/cos/acos/exp/log/erf = libm; erfinv = Giles' single-precision polynomial
//   (the published algorithm Enoki's special.h implements).
// PCG32 is the public O'Neill generator exactly as enoki/random.h wraps it.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <algorithm>

namespace orc {

// ---- constants: include/mitsuba/core/math.h:17-41 ------------------------------------------
constexpr float Pi = 3.14159265358979323846f;
constexpr float TwoPi = 6.28318530717958647692f;
constexpr float InvPi = 0.31830988618379067154f;
constexpr float InvTwoPi = 0.15915494309189533577f;
constexpr float InvSqrtPi = 0.56418958354775628695f;
constexpr float Infinity = std::numeric_limits<float>::infinity();
constexpr float Epsilon = std::numeric_limits<float>::epsilon() / 2;  // math.h:36
constexpr float RayEpsilon = Epsilon * 1500;                            // math.h:37
constexpr float ShadowEpsilon = RayEpsilon * 10;                        // math.h:38

inline float fmadd(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float fmsub(float a, float b, float c) { return std::fmaf(a, b, -c); }
inline float fnmadd(float a, float b, float c) { return std::fmaf(-a, b, c); }
inline float rcp(float x) { return 1.f / x; }
inline float rsqrt(float x) { return 1.f / std::sqrt(x); }
inline float sqr(float x) { return x * x; }
inline float safe_sqrt(float x) { return std::sqrt(std::max(x, 0.f)); }
inline float sign_(float x) { return std::copysign(1.f, x); }             // enoki::sign
inline float mulsign(float a, float b) {                                  // a * sign(b), bitwise
    uint32_t ua, ub; std::memcpy(&ua, &a, 4); std::memcpy(&ub, &b, 4);
    ua ^= (ub & 0x80000000u); float r; std::memcpy(&r, &ua, 4); return r;
}
inline float mulsign_neg(float a, float b) {                              // a * -sign(b)
    uint32_t ua, ub; std::memcpy(&ua, &a, 4); std::memcpy(&ub, &b, 4);
    ua ^= ((~ub) & 0x80000000u); float r; std::memcpy(&r, &ua, 4); return r;
}
inline float lerp_(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }
inline float clamp_(float x, float lo, float hi) { return std::min(std::max(x, lo), hi); }

struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    explicit V3(float s) : x(s), y(s), z(s) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, float s) { float r = rcp(s); return {a.x * r, a.y * r, a.z * r}; }
inline float dot(V3 a, V3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
inline float abs_dot(V3 a, V3 b) { return std::fabs(dot(a, b)); }
inline float squared_norm(V3 a) { return dot(a, a); }
inline float norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a * rsqrt(squared_norm(a)); }
inline V3 cross(V3 a, V3 b) {
    return {fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x)};
}
inline V3 fmadd3(V3 a, float s, V3 c) { return {fmadd(a.x, s, c.x), fmadd(a.y, s, c.y), fmadd(a.z, s, c.z)}; }
inline float hmax_abs(V3 p) { return std::max(std::max(std::fabs(p.x), std::fabs(p.y)), std::fabs(p.z)); }

struct V2 { float x, y; V2() : x(0), y(0) {} V2(float a, float b) : x(a), y(b) {} };

// ---- "Jacob functions": include/mitsuba/core/math.h:68-131 ----------------------------------
inline float jabs(float x) { return x >= 0.f ? x : -x; }                        // math.h:70-72
inline float sinc(float x) { return jabs(x) > Epsilon ? std::sin(x) / x : 1.f; } // math.h:76-80
inline float tri(float x) { return jabs(x) < 0.5f ? 1.f - 2.f * jabs(x) : 0.f; } // math.h:84-86
inline float rect(float x) { return jabs(x) < 0.5f ? 1.f : 0.f; }                // math.h:90-92
// math.h:108-123: a literal subtraction loop (NOT fmodf).  `all(x)` on a scalar float is
// `(bool) x`, so the guarded subtraction fires whenever the loop condition holds.
inline float fmodulo(float a, float b) {
    float result = jabs(a);
    const float ab = jabs(b);
    uint64_t guard = 0;
    while (result - ab >= Epsilon) {
        result -= ((result - ab) != 0.f ? 1.f : 0.f) >= Epsilon ? ab : 0.f;
        if (++guard > (1ull << 26)) break;  // the reference would spin; callers never get here
    }
    result = (a < 0.f) ? ab - result : result;
    result += (b < 0.f) ? b : 0.f;
    return result;
}
// math.h:127-131
inline float wchirp(float t, float f, float w, float a) {
    return 2 * a * a * w * tri(t / w) * sinc(TwoPi * f * w * tri(t / w));
}

// ---- coordinate_system: include/mitsuba/core/vector.h:116-135 --------------------------------
inline void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = sign_(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    s = V3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = V3(b, sign + sqr(n.y) * a, -n.y);
}

// ---- Frame: include/mitsuba/core/frame.h:17-43 -----------------------------------------------
struct Frame {
    V3 s, t, n;
    Frame() {}
    explicit Frame(V3 v) : n(v) { coordinate_system(v, s, t); }
    Frame(V3 s_, V3 t_, V3 n_) : s(s_), t(t_), n(n_) {}
    V3 to_local(V3 v) const { return V3(dot(v, s), dot(v, t), dot(v, n)); }
    V3 to_world(V3 v) const { return s * v.x + t * v.y + n * v.z; }
};

// ---- Transform4f: include/mitsuba/core/transform.h:85-160 (row-major m[r][c]) ----------------
struct M4 {
    float m[4][4];
    static M4 identity() { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = i == j; return r; }
    static M4 from_rows16(const float *p) { M4 r; std::memcpy(r.m, p, 64); return r; }
};
inline M4 mul(const M4 &a, const M4 &b) {
    M4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) {
        double s = 0; for (int k = 0; k < 4; ++k) s += (double) a.m[i][k] * b.m[k][j];
        r.m[i][j] = (float) s;
    }
    return r;
}
// General 4x4 inverse (scene-load time only; the reference uses enoki::inverse).
inline M4 inverse(const M4 &a) {
    double inv[16], m[16];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[i * 4 + j] = a.m[i][j];
    inv[0] = m[5]*m[10]*m[15]-m[5]*m[11]*m[14]-m[9]*m[6]*m[15]+m[9]*m[7]*m[14]+m[13]*m[6]*m[11]-m[13]*m[7]*m[10];
    inv[4] = -m[4]*m[10]*m[15]+m[4]*m[11]*m[14]+m[8]*m[6]*m[15]-m[8]*m[7]*m[14]-m[12]*m[6]*m[11]+m[12]*m[7]*m[10];
    inv[8] = m[4]*m[9]*m[15]-m[4]*m[11]*m[13]-m[8]*m[5]*m[15]+m[8]*m[7]*m[13]+m[12]*m[5]*m[11]-m[12]*m[7]*m[9];
    inv[12] = -m[4]*m[9]*m[14]+m[4]*m[10]*m[13]+m[8]*m[5]*m[14]-m[8]*m[6]*m[13]-m[12]*m[5]*m[10]+m[12]*m[6]*m[9];
    inv[1] = -m[1]*m[10]*m[15]+m[1]*m[11]*m[14]+m[9]*m[2]*m[15]-m[9]*m[3]*m[14]-m[13]*m[2]*m[11]+m[13]*m[3]*m[10];
    inv[5] = m[0]*m[10]*m[15]-m[0]*m[11]*m[14]-m[8]*m[2]*m[15]+m[8]*m[3]*m[14]+m[12]*m[2]*m[11]-m[12]*m[3]*m[10];
    inv[9] = -m[0]*m[9]*m[15]+m[0]*m[11]*m[13]+m[8]*m[1]*m[15]-m[8]*m[3]*m[13]-m[12]*m[1]*m[11]+m[12]*m[3]*m[9];
    inv[13] = m[0]*m[9]*m[14]-m[0]*m[10]*m[13]-m[8]*m[1]*m[14]+m[8]*m[2]*m[13]+m[12]*m[1]*m[10]-m[12]*m[2]*m[9];
    inv[2] = m[1]*m[6]*m[15]-m[1]*m[7]*m[14]-m[5]*m[2]*m[15]+m[5]*m[3]*m[14]+m[13]*m[2]*m[7]-m[13]*m[3]*m[6];
    inv[6] = -m[0]*m[6]*m[15]+m[0]*m[7]*m[14]+m[4]*m[2]*m[15]-m[4]*m[3]*m[14]-m[12]*m[2]*m[7]+m[12]*m[3]*m[6];
    inv[10] = m[0]*m[5]*m[15]-m[0]*m[7]*m[13]-m[4]*m[1]*m[15]+m[4]*m[3]*m[13]+m[12]*m[1]*m[7]-m[12]*m[3]*m[5];
    inv[14] = -m[0]*m[5]*m[14]+m[0]*m[6]*m[13]+m[4]*m[1]*m[14]-m[4]*m[2]*m[13]-m[12]*m[1]*m[6]+m[12]*m[2]*m[5];
    inv[3] = -m[1]*m[6]*m[11]+m[1]*m[7]*m[10]+m[5]*m[2]*m[11]-m[5]*m[3]*m[10]-m[9]*m[2]*m[7]+m[9]*m[3]*m[6];
    inv[7] = m[0]*m[6]*m[11]-m[0]*m[7]*m[10]-m[4]*m[2]*m[11]+m[4]*m[3]*m[10]+m[8]*m[2]*m[7]-m[8]*m[3]*m[6];
    inv[11] = -m[0]*m[5]*m[11]+m[0]*m[7]*m[9]+m[4]*m[1]*m[11]-m[4]*m[3]*m[9]-m[8]*m[1]*m[7]+m[8]*m[3]*m[5];
    inv[15] = m[0]*m[5]*m[10]-m[0]*m[6]*m[9]-m[4]*m[1]*m[10]+m[4]*m[2]*m[9]+m[8]*m[1]*m[6]-m[8]*m[2]*m[5];
    double det = m[0]*inv[0]+m[1]*inv[4]+m[2]*inv[8]+m[3]*inv[12];
    M4 r; double id = 1.0 / det;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = (float) (inv[i * 4 + j] * id);
    return r;
}
inline M4 transpose(const M4 &a) { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i]; return r; }
// transform.h:91-99 transform_affine(Point): result = col3; result = fmadd(col_i, arg_i, result)
inline V3 xf_point_affine(const M4 &M, V3 p) {
    float r[3];
    for (int j = 0; j < 3; ++j)
        r[j] = fmadd(M.m[j][2], p.z, fmadd(M.m[j][1], p.y, fmadd(M.m[j][0], p.x, M.m[j][3])));
    return V3(r[0], r[1], r[2]);
}
// transform.h:105-112 operator*(Point): same + perspective divide (true division)
inline V3 xf_point(const M4 &M, V3 p) {
    float r[4];
    for (int j = 0; j < 4; ++j)
        r[j] = fmadd(M.m[j][2], p.z, fmadd(M.m[j][1], p.y, fmadd(M.m[j][0], p.x, M.m[j][3])));
    float iw = rcp(r[3]);
    return V3(r[0] * iw, r[1] * iw, r[2] * iw);
}
// transform.h:119-127 operator*(Vector): result = col0 * x; fmadd for y, z
inline V3 xf_vector(const M4 &M, V3 v) {
    float r[3];
    for (int j = 0; j < 3; ++j)
        r[j] = fmadd(M.m[j][2], v.z, fmadd(M.m[j][1], v.y, M.m[j][0] * v.x));
    return V3(r[0], r[1], r[2]);
}
// transform.h:134-142 operator*(Normal): uses inverse_transpose (pass M = inverse; columns of
// inverse_transpose are rows of the inverse)
inline V3 xf_normal(const M4 &Minv, V3 n) {
    float r[3];
    for (int j = 0; j < 3; ++j)
        r[j] = fmadd(Minv.m[2][j], n.z, fmadd(Minv.m[1][j], n.y, Minv.m[0][j] * n.x));
    return V3(r[0], r[1], r[2]);
}

// ---- PCG32 (enoki/random.h; public O'Neill generator) and TEA (core/random.h:74-116) ----------
constexpr uint64_t PCG32_DEFAULT_STATE = 0x853c49e6748fea9bULL;
constexpr uint64_t PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdbULL;
constexpr uint64_t PCG32_MULT = 0x5851f42d4c957f2dULL;
struct PCG32 {
    uint64_t state, inc;
    PCG32(uint64_t initstate = PCG32_DEFAULT_STATE, uint64_t initseq = PCG32_DEFAULT_STREAM) { seed(initstate, initseq); }
    void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u; next_u32(); state += initstate; next_u32();
    }
    uint32_t next_u32() {
        uint64_t old = state;
        state = old * PCG32_MULT + inc;
        uint32_t xorshifted = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    float next_float32() {
        uint32_t u = (next_u32() >> 9) | 0x3f800000u; float f; std::memcpy(&f, &u, 4); return f - 1.f;
    }
};
inline uint32_t sample_tea_32(uint32_t v0, uint32_t v1, int rounds = 4) {   // random.h:74-85
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v1;
}
inline uint64_t sample_tea_64(uint32_t v0, uint32_t v1, int rounds = 4) {   // random.h:104-116
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return (uint64_t) v0 + ((uint64_t) v1 << 32);
}
// sampler.cpp:89-92 instantiates sample_tea_64 with UInt64 arguments (the wavefront index is an
// arange<UInt64>), i.e. the same rounds carried out in 64-bit arithmetic without 32-bit wrap.
inline uint64_t sample_tea_64_wide(uint64_t v0, uint64_t v1, int rounds = 4) {
    uint64_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v0 + (v1 << 32);
}
inline float sample_tea_float32(uint32_t v0, uint32_t v1, int rounds = 4) { // random.h:131-134
    uint32_t u = (sample_tea_32(v0, v1, rounds) >> 9) | 0x3f800000u; float f; std::memcpy(&f, &u, 4); return f - 1.f;
}
inline double sample_tea_float64(uint32_t v0, uint32_t v1, int rounds = 4) {
    uint64_t u = (sample_tea_64(v0, v1, rounds) >> 12) | 0x3ff0000000000000ull; double d; std::memcpy(&d, &u, 8); return d - 1.0;
}

// ---- warps: include/mitsuba/core/warp.h:54-90, 325-349 ----------------------------------------
inline V2 square_to_uniform_disk_concentric(V2 sample) {
    float x = fmsub(2.f, sample.x, 1.f), y = fmsub(2.f, sample.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), quadrant_1_or_3 = std::fabs(x) < std::fabs(y);
    float r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    float phi = .25f * Pi * rp / r;
    if (quadrant_1_or_3) phi = .5f * Pi - phi;
    if (is_zero) phi = 0.f;
    float s = std::sin(phi), c = std::cos(phi);
    return V2(r * c, r * s);
}
inline V3 square_to_cosine_hemisphere(V2 sample) {
    V2 p = square_to_uniform_disk_concentric(sample);
    float sq = fmadd(p.y, p.y, p.x * p.x);  // squared_norm of a 2-vector (dot_ fma chain)
    float z = safe_sqrt(1.f - sq);
    return V3(p.x, p.y, z);
}
inline float square_to_cosine_hemisphere_pdf(V3 v) { return InvPi * v.z; }
inline V3 square_to_uniform_cone(V2 sample, float cos_cutoff) {             // warp.h:268-283
    float one_minus = 1.f - cos_cutoff;
    V2 p = square_to_uniform_disk_concentric(sample);
    float pn = fmadd(p.y, p.y, p.x * p.x);
    float z = cos_cutoff + one_minus * (1.f - pn);
    float f = safe_sqrt(one_minus * (2.f - one_minus * pn));
    return V3(p.x * f, p.y * f, z);
}

// ---- erfinv: "Approximating the erfinv function", M. Giles (single precision) ------------------
inline float erfinv(float x) {
    float w = -std::log((1.f - x) * (1.f + x)), p;
    if (w < 5.f) {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = fmadd(p, w, 3.43273939e-07f); p = fmadd(p, w, -3.5233877e-06f);
        p = fmadd(p, w, -4.39150654e-06f); p = fmadd(p, w, 0.00021858087f);
        p = fmadd(p, w, -0.00125372503f); p = fmadd(p, w, -0.00417768164f);
        p = fmadd(p, w, 0.246640727f); p = fmadd(p, w, 1.50140941f);
    } else {
        w = std::sqrt(w) - 3.f;
        p = -0.000200214257f;
        p = fmadd(p, w, 0.000100950558f); p = fmadd(p, w, 0.00134934322f);
        p = fmadd(p, w, -0.00367342844f); p = fmadd(p, w, 0.00573950773f);
        p = fmadd(p, w, -0.0076224613f); p = fmadd(p, w, 0.00943887047f);
        p = fmadd(p, w, 1.00167406f); p = fmadd(p, w, 2.83297682f);
    }
    return p * x;
}

}  // namespace orc
