// beifong_b200 — device functions of the radar path: surface interaction, BSDFs, signal model,
// transmitters, receivers.  Each function names the reference code it replaces (paths under
// /root/reference).  fp32 everywhere, FMA only where the reference has one (see bf_math.h); the
// few `double` expressions are the ones the reference promotes through a double literal
// (`wavelengths*1e-9`, `(MAX-MIN)/2*1e-9`).
#pragma once
#include "bf_scene.h"

// Large device functions are kept out of line: one copy each keeps the shade kernel inside the
// instruction cache (fully inlined it was 326 KB of SASS, ncu: 38 % of stall samples "no instruction").
#define BF_NI __device__ __noinline__

namespace bf {

// atomicAdd with a returned value, issued by ONE elected lane: written as PTX so that the compiler does not wrap it in its
// own warp aggregation (elect + SHFL broadcast of the result right behind the atomic, which makes the issuing warp wait
// out the ~1 us round trip on the spot; the callers consume the result one loop iteration later instead).
__device__ __forceinline__ uint32_t atomic_add_async(uint32_t *p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}

template <int S> struct Spec {
    float v[S];
    BF_D Spec() {
#pragma unroll
        for (int k = 0; k < S; ++k) v[k] = 0.f;
    }
    BF_D explicit Spec(float s) {
#pragma unroll
        for (int k = 0; k < S; ++k) v[k] = s;
    }
};
template <int S> BF_D float hmax_spec(const Spec<S> &s) {
    float m = s.v[0];
#pragma unroll
    for (int k = 1; k < S; ++k) m = fmaxf(m, s.v[k]);
    return m;
}
// enoki hsum of a 4-vector: (v0+v1)+(v2+v3)
template <int S> BF_D float hsum_spec(const Spec<S> &s) {
    if constexpr (S == 1) return s.v[0];
    else return (s.v[0] + s.v[1]) + (s.v[2] + s.v[3]);
}

// `MTS_C*rcp(wavelengths*1e-9)`: the double literal promotes the whole expression
// (integrator.cpp:1592,1603,1619; wignertransmitter.cpp:233).
BF_D double freq_from_wl_d(float c, float wl_nm) { return (double) c * (1.0 / ((double) wl_nm * 1e-9)); }
BF_D float freq_from_wl(float c, float wl_nm) { return (float) freq_from_wl_d(c, wl_nm); }

// Surface interaction: the subset of SurfaceInteraction3f (interaction.h:33-172) consumed downstream.
struct SI {
    float t;
    V3 p, n, wi;
    Frame sh;
    uint32_t shape, prim;
    BF_D bool is_valid() const { return t != bf_inf(); }
    BF_D V3 to_world(V3 v) const { return sh.to_world(v); }
    BF_D V3 to_local(V3 v) const { return sh.to_local(v); }
};
struct PositionSample { V3 p, n; float time, pdf; bool delta; };
struct DirectionSample { V3 p, n, d; float time, pdf, dist; bool delta; int object; };

// ---- spectra: src/spectra/uniform.cpp:27-40, distr_1d.h:387-403 (regular), :664-687 (irregular) ---
template <int S> BF_NI float spectrum_eval_1(const DevScene &sc, uint32_t id, float wl) {
    const DevSpectrum t = sc.spectra[id];
    const float *lam = sc.spec_pool + t.offset, *val = lam + t.n;
    switch (t.kind) {
        case BF_SPEC_UNIFORM:
            if (S > 1) return (wl >= sc.C.lambda_min_nm && wl <= sc.C.lambda_max_nm) ? val[0] : 0.f;
            return val[0];
        case BF_SPEC_REGULAR: {
            bool active = wl >= lam[0] && wl <= lam[t.n - 1];
            float x = (wl - lam[0]) * t.inv_interval;
            long long xi = (x == x) ? (long long) fminf(fmaxf(x, -1.f), 4.0e9f) : 0;
            long long hi = (long long) t.n - 2;
            uint32_t index = (uint32_t) (xi < 0 ? 0 : (xi > hi ? hi : xi));
            float y0 = active ? val[index] : 0.f, y1 = active ? val[index + 1] : 0.f;
            float w1 = x - (float) index, w0 = 1.f - w1;
            return fmadd(w0, y0, w1 * y1);
        }
        case BF_SPEC_IRREGULAR: {
            bool active = wl >= lam[0] && wl <= lam[t.n - 1];
            uint32_t lo = 0, hi = t.n;
            while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (active && lam[mid] < wl) lo = mid + 1; else hi = mid; }
            uint32_t m = lo < t.n - 1u ? lo : t.n - 1u;
            uint32_t index = (m > 1u ? m : 1u) - 1u;
            float x0 = lam[index], x1 = lam[index + 1], y0 = val[index], y1 = val[index + 1];
            float x = (wl - x0) / (x1 - x0);
            return active ? fmadd(x, y1 - y0, y0) : 0.f;
        }
        default:
            return 0.f;
    }
}
template <int S> BF_D Spec<S> spectrum_eval(const DevScene &sc, uint32_t id, const Spec<S> &wl) {
    Spec<S> r;
#pragma unroll
    for (int k = 0; k < S; ++k) r.v[k] = spectrum_eval_1<S>(sc, id, wl.v[k]);
    return r;
}

// ---- Rectangle: src/shapes/rectangle.cpp:111-125, 132-220, 229-298 ---------------------------------
BF_D PositionSample rect_sample_position(const DevRect &r, float time, V2 sample) {
    PositionSample ps;
    ps.p = xf_point_affine(r.to_world, V3(sample.x * 2.f - 1.f, sample.y * 2.f - 1.f, 0.f));
    ps.n = r.fn; ps.pdf = r.inv_area; ps.time = time; ps.delta = false;
    return ps;
}
// rectangle.cpp:132-220.  `m_to_object * ds.p/2` is the projective operator* followed by /2.
// inv_wl = 1.0 / (wavelength0 * 1e-9) in double (the literal promotes the expression): callers that need the same
// quotient for the frequency (`MTS_C*rcp(wavelengths*1e-9)`) evaluate the division once and pass it in.
BF_D double inv_wavelength_d(float wavelength0) { return 1.0 / ((double) wavelength0 * 1e-9); }
BF_D float rect_sample_wigner_impl(const DevRect &r, V3 ds_p, V3 ds_d, double inv_wl) {
    V3 r_hat = xf_point(r.to_object, ds_p) / 2.f;
    V3 tv(fmadd(r.ns.z, ds_d.z, fmadd(r.ns.y, ds_d.y, r.ns.x * ds_d.x)),
          fmadd(r.nt.z, ds_d.z, fmadd(r.nt.y, ds_d.y, r.nt.x * ds_d.x)),
          fmadd(r.nn.z, ds_d.z, fmadd(r.nn.y, ds_d.y, r.nn.x * ds_d.x)));
    float nu_x = (float) ((double) tv.x * inv_wl), nu_y = (float) ((double) tv.y * inv_wl);
    float gain = 4 * tri(r_hat.x) * tri(r_hat.y) *
                 sinc(BF_TWO_PI * nu_x * r.wid_x * tri(r_hat.x)) *
                 sinc(BF_TWO_PI * nu_y * r.wid_y * tri(r_hat.y));
    return gain;
}
BF_NI float rect_sample_wigner(const DevRect &r, V3 ds_p, V3 ds_d, float wavelength0) {
    return rect_sample_wigner_impl(r, ds_p, ds_d, inv_wavelength_d(wavelength0));
}
// Phased arrays: sample_wigner (phasedtransmitter.cpp:275-290 == phasedreceiver.cpp:289-304) with W_rect_2D (:261-273):
// the real part of sum_i W_rect(r_hat_i, nu_hat_i) exp(2 pi i nu_hat_i . r'_i) psi'_i over the n^2 virtual elements
// whose footprint contains the point.  Complex arithmetic spelled out as std::complex<float> evaluates it.
BF_NI float phased_sample_wigner(const DevScene &sc, const DevEndpoint &ep, V3 ds_p, V3 ds_d, float wavelength0) {
    float w_re = 0.f;
    const double inv_wl = 1.0 / ((double) wavelength0 * 1e-9);
    const DevVElem *ve = sc.velems + ep.velem_offset;
    for (uint32_t i = 0; i < ep.n_velems; ++i) {
        const DevVElem &e = ve[i];
        V3 r_hat = xf_point(e.to_object, ds_p) / 2.f;
        if (!(jabs(r_hat.x) <= 0.5f && jabs(r_hat.y) <= 0.5f)) continue;
        V3 tv(fmadd(e.ns.z, ds_d.z, fmadd(e.ns.y, ds_d.y, e.ns.x * ds_d.x)),
              fmadd(e.nt.z, ds_d.z, fmadd(e.nt.y, ds_d.y, e.nt.x * ds_d.x)),
              fmadd(e.nn.z, ds_d.z, fmadd(e.nn.y, ds_d.y, e.nn.x * ds_d.x)));
        V3 nu_hat((float) ((double) tv.x * inv_wl), (float) ((double) tv.y * inv_wl), (float) ((double) tv.z * inv_wl));
        float g = 4 * ep.wid_x * ep.wid_y * tri(r_hat.x) * tri(r_hat.y) *
                  sinc(BF_TWO_PI * nu_hat.x * ep.wid_x * tri(r_hat.x)) *
                  sinc(BF_TWO_PI * nu_hat.y * ep.wid_y * tri(r_hat.y));
        float th = BF_TWO_PI * dot(nu_hat, e.r_dash);
        float sn, cs;
        sincosf(th, &sn, &cs);
        float a_re = g * cs, a_im = g * sn;
        w_re += a_re * e.psi_re - a_im * e.psi_im;
    }
    return w_re;
}
// Shape::sample_direction / pdf_direction, src/librender/shape.cpp:323-354
BF_D DirectionSample shape_sample_direction(const DevRect &r, V3 it_p, float it_time, V2 sample) {
    PositionSample ps = rect_sample_position(r, it_time, sample);
    DirectionSample ds;
    ds.p = ps.p; ds.n = ps.n; ds.time = ps.time; ds.pdf = ps.pdf; ds.delta = false;
    ds.d = ds.p - it_p;
    float dist_squared = squared_norm(ds.d);
    ds.dist = sqrtf(dist_squared);
    ds.d = ds.d / ds.dist;
    float dp = abs_dot(ds.d, ds.n);
    ds.pdf *= (dp != 0.f) ? dist_squared / dp : 0.f;
    ds.object = 0;
    return ds;
}
BF_D float shape_pdf_direction(const DevRect &r, const DirectionSample &ds) {
    float pdf = r.inv_area, dp = abs_dot(ds.d, ds.n);
    pdf *= (dp != 0.f) ? (ds.dist * ds.dist) / dp : 0.f;
    return pdf;
}

// ---- surface interaction ------------------------------------------------------------------------------
// Mesh::compute_surface_interaction (src/librender/mesh.cpp:452-548), Rectangle::
// compute_surface_interaction (rectangle.cpp:265-298), initialize_sh_frame + wi
// (interaction.h:159-162, 613-644).
BF_D SI compute_si_impl(const DevScene &sc, V3 ray_o, V3 ray_d, float4 hit) {
    SI si;
    si.t = hit.x;
    if (hit.x == bf_inf()) {
        si.wi = -ray_d; si.shape = BF_NONE; si.prim = 0;
        return si;
    }
    uint32_t slot = __float_as_uint(hit.w);
    V3 dp_du;
    if (slot & 0x80000000u) {
        // rectangle.cpp:265-298; the shading frame of a rectangle is a per-rectangle constant (DevRect::sh_s / sh_t)
        const DevRect &r = sc.rects[slot & 0x7fffffffu];
        si.p = fmadd3(ray_d, hit.x, ray_o);
        si.n = r.fn; si.sh.n = r.fn; si.sh.s = r.sh_s; si.sh.t = r.sh_t;
        si.shape = r.shape; si.prim = 0;
        si.wi = si.to_local(-ray_d);
        return si;
    } else {
        float4 a = __ldg(&sc.tris[3 * (size_t) slot]), b = __ldg(&sc.tris[3 * (size_t) slot + 1]), c = __ldg(&sc.tris[3 * (size_t) slot + 2]);
        V3 p0(a.x, a.y, a.z), p1(b.x, b.y, b.z), p2(c.x, c.y, c.z);
        si.shape = __float_as_uint(a.w); si.prim = __float_as_uint(b.w);
        float b1 = hit.y, b2 = hit.z, b0 = 1.f - b1 - b2;
        V3 dp0 = p1 - p0, dp1 = p2 - p0;
        si.p = p0 * b0 + p1 * b1 + p2 * b2;
        si.n = normalize(cross(dp0, dp1));
        V3 dp_dv;
        coordinate_system(si.n, dp_du, dp_dv);
        if (sc.shapes[si.shape].has_normals) {
            float4 na = __ldg(&sc.tri_nrm[3 * (size_t) slot]), nb = __ldg(&sc.tri_nrm[3 * (size_t) slot + 1]), nc = __ldg(&sc.tri_nrm[3 * (size_t) slot + 2]);
            V3 n0(na.x, na.y, na.z), n1(nb.x, nb.y, nb.z), n2(nc.x, nc.y, nc.z);
            si.sh.n = normalize(n0 * b0 + n1 * b1 + n2 * b2);
        } else {
            si.sh.n = si.n;
        }
    }
    float dn = dot(si.sh.n, dp_du);
    si.sh.s = normalize(V3(fnmadd(si.sh.n.x, dn, dp_du.x), fnmadd(si.sh.n.y, dn, dp_du.y), fnmadd(si.sh.n.z, dn, dp_du.z)));
    si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = si.to_local(-ray_d);
    return si;
}

BF_NI SI compute_si(const DevScene &sc, V3 ray_o, V3 ray_d, float4 hit) { return compute_si_impl(sc, ray_o, ray_d, hit); }

// Shape::doppler (src/librender/shape.cpp:375-389): 2 * dot(si.wi, m_velocity * Point3f(si.to_local(si.p))) / MTS_C * si.wavelengths.
// Reproduced as written: the velocity transform is applied (projectively) to the hit position expressed in the shading frame.
template <int S> BF_NI Spec<S> shape_doppler(const DevScene &sc, const SI &si, const Spec<S> &wl) {
    const V3 pl = si.to_local(si.p);
    const V3 vp = xf_point(sc.velocities[si.shape], pl);
    const float f = 2 * dot(si.wi, vp) / sc.C.c;
    Spec<S> r;
#pragma unroll
    for (int k = 0; k < S; ++k) r.v[k] = f * wl.v[k];
    return r;
}

// ---- microfacet distribution: include/mitsuba/render/microfacet.h -------------------------------------
struct Microfacet {
    int type; float au, av; bool sample_visible;
    BF_D Microfacet(int type_, float au_, float av_, bool sv) : type(type_), au(fmaxf(au_, 1e-4f)), av(fmaxf(av_, 1e-4f)), sample_visible(sv) {}
    BF_D float eval(V3 m) const {                                           // :179-196
        float alpha_uv = au * av, cos_theta = m.z, cos_theta_2 = sqr(cos_theta), result;
        if (type == BF_MICROFACET_BECKMANN)
            result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / cos_theta_2) / (BF_PI * alpha_uv * sqr(cos_theta_2));
        else
            result = rcp(BF_PI * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * cos_theta > 1e-20f ? result : 0.f;
    }
    BF_D float smith_g1(V3 v, V3 m) const {                                 // :330-358
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == BF_MICROFACET_BECKMANN) {
            float a = rsqrt_(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    BF_D float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    BF_D float pdf(V3 wi, V3 m) const {                                     // :206-215
        float result = eval(m);
        if (sample_visible) result *= smith_g1(wi, m) * abs_dot(wi, m) / wi.z;
        else result *= m.z;
        return result;
    }
    BF_D V2 sample_visible_11(float cos_theta_i, V2 sample) const {         // :361-409
        if (type == BF_MICROFACET_BECKMANN) {
            float tan_theta_i = safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = erff(cot_theta_i);
            sample.x = fmaxf(fminf(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = fmaxf(fminf(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sample.x)));
            sample.x *= 1.f + maxval + BF_INV_SQRT_PI * tan_theta_i * expf(-sqr(cot_theta_i));
#pragma unroll 1
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_(x), value = 1.f + x + BF_INV_SQRT_PI * tan_theta_i * expf(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return V2(erfinv_(x), erfinv_(fmsub(2.f, sample.y, 1.f)));
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = .5f * (1.f + cos_theta_i);
            p.y = lerp_(safe_sqrt(1.f - sqr(p.x)), p.y, s);
            float x = p.x, y = p.y, z = safe_sqrt(1.f - fmadd(p.y, p.y, p.x * p.x));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = rcp(fmadd(sin_theta_i, y, cos_theta_i * z));
            return V2(fmsub(cos_theta_i, y, sin_theta_i * z) * nrm, x * nrm);
        }
    }
    BF_D void sample(V3 wi, V2 sample, V3 &m_out, float &pdf_out) const {   // :217-322
        if (!sample_visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2, pdf;
            if (au == av) {
                float ang = (2.f * BF_PI) * sample.y; sin_phi = sinf(ang); cos_phi = cosf(ang);
                alpha_2 = au * au;
            } else {
                float ratio = av / au, tmp = ratio * tanf((2.f * BF_PI) * sample.y);
                cos_phi = rsqrt_(fmadd(tmp, tmp, 1.f));
                cos_phi = mulsign(cos_phi, fabsf(sample.y - .5f) - .25f);
                sin_phi = cos_phi * tmp;
                alpha_2 = rcp(sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == BF_MICROFACET_BECKMANN) {
                cos_theta = rsqrt_(fnmadd(alpha_2, logf(1.f - sample.x), 1.f));
                cos_theta_2 = sqr(cos_theta);
                float cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf = (1.f - sample.x) / (BF_PI * au * av * cos_theta_3);
            } else {
                float tan_theta_m_2 = alpha_2 * sample.x / (1.f - sample.x);
                cos_theta = rsqrt_(1.f + tan_theta_m_2);
                cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf = rcp(BF_PI * au * av * cos_theta_3 * sqr(temp));
            }
            float sin_theta = sqrtf(1.f - cos_theta_2);
            m_out = V3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta); pdf_out = pdf;
        } else {
            V3 wi_p = normalize(V3(au * wi.x, av * wi.y, wi.z));
            // Frame::sincos_phi (frame.h:107-118)
            float sin_theta_2 = fnmadd(wi_p.z, wi_p.z, 1.f), inv_sin_theta = rsqrt_(sin_theta_2);
            float rx = wi_p.x * inv_sin_theta, ry = wi_p.y * inv_sin_theta;
            float cos_phi, sin_phi;
            if (fabsf(sin_theta_2) <= 4.f * BF_EPS) { cos_phi = 1.f; sin_phi = 0.f; }
            else { cos_phi = clamp_(rx, -1.f, 1.f); sin_phi = clamp_(ry, -1.f, 1.f); }
            float cos_theta = wi_p.z;
            V2 slope = sample_visible_11(cos_theta, sample);
            slope = V2(fmsub(cos_phi, slope.x, sin_phi * slope.y) * au, fmadd(sin_phi, slope.x, cos_phi * slope.y) * av);
            V3 m = normalize(V3(-slope.x, -slope.y, 1));
            float pdf = eval(m) * smith_g1(wi, m) * abs_dot(wi, m) / wi.z;
            m_out = m; pdf_out = pdf;
        }
    }
};
// include/mitsuba/render/fresnel.h:92-116
BF_NI float fresnel_conductor(float cos_theta_i, float eta_r, float eta_i) {
    float cos_theta_i_2 = cos_theta_i * cos_theta_i, sin_theta_i_2 = 1.f - cos_theta_i_2, sin_theta_i_4 = sin_theta_i_2 * sin_theta_i_2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - sin_theta_i_2,
          a_2_pb_2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r),
          a = safe_sqrt(.5f * (a_2_pb_2 + temp_1));
    float term_1 = a_2_pb_2 + cos_theta_i_2, term_2 = 2.f * cos_theta_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a_2_pb_2 * cos_theta_i_2 + sin_theta_i_4, term_4 = term_2 * sin_theta_i_2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return .5f * (r_s + r_p);
}
BF_D V3 reflect(V3 wi, V3 m) {  // fresnel.h:278-280
    float k = 2.f * dot(wi, m);
    return V3(fmsub(m.x, k, wi.x), fmsub(m.y, k, wi.y), fmsub(m.z, k, wi.z));
}

// ---- BSDFs: src/bsdfs/diffuse.cpp:78-135, roughconductor.cpp:196-394, conductor.cpp, twosided.cpp:94-181
struct BSDFSample { V3 wo; float pdf, eta; bool delta; };

BF_D bool bsdf_is_smooth(const bf_bsdf_desc &b) { return b.kind == BF_BSDF_DIFFUSE || b.kind == BF_BSDF_ROUGHCONDUCTOR; }

// BK >= 0: the BSDF kind is known at compile time (material-class shade kernels); -1: read from the descriptor.
template <int S, int BK> BF_D Spec<S> onesided_eval_impl(const DevScene &sc, const bf_bsdf_desc &b, const Spec<S> &wl, V3 wi, V3 wo) {
    Spec<S> r;
    const uint32_t kind = BK >= 0 ? (uint32_t) BK : b.kind;
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (kind == BF_BSDF_DIFFUSE) {
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return r;
        Spec<S> refl = spectrum_eval<S>(sc, b.reflectance, wl);
#pragma unroll
        for (int k = 0; k < S; ++k) r.v[k] = refl.v[k] * BF_INV_PI * cos_theta_o;
        return r;
    } else if (kind == BF_BSDF_ROUGHCONDUCTOR) {
        bool active = cos_theta_i > 0.f && cos_theta_o > 0.f;
        if (!active) return r;
        V3 H = normalize(wo + wi);
        Microfacet distr((int) b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        float D = distr.eval(H);
        active &= D != 0.f;
        float G = distr.G(wi, wo, H);
        float result = D * G / (4.f * wi.z);
        Spec<S> eta = spectrum_eval<S>(sc, b.eta, wl), kk = spectrum_eval<S>(sc, b.k, wl);
        float cosh_ = dot(wi, H);
        Spec<S> sr(1.f);
        if (b.specular_reflectance != BF_NONE) sr = spectrum_eval<S>(sc, b.specular_reflectance, wl);
#pragma unroll
        for (int k = 0; k < S; ++k) {
            float F = fresnel_conductor(cosh_, eta.v[k], kk.v[k]);
            float res = result;
            if (b.specular_reflectance != BF_NONE) res *= sr.v[k];
            r.v[k] = active ? F * res : 0.f;
        }
        return r;
    }
    return r;
}
template <int BK> BF_D float onesided_pdf_impl(const bf_bsdf_desc &b, V3 wi, V3 wo) {
    const uint32_t kind = BK >= 0 ? (uint32_t) BK : b.kind;
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (kind == BF_BSDF_DIFFUSE) {
        float pdf = square_to_cosine_hemisphere_pdf(wo);
        return (cos_theta_i > 0.f && cos_theta_o > 0.f) ? pdf : 0.f;
    } else if (kind == BF_BSDF_ROUGHCONDUCTOR) {
        V3 m = normalize(wo + wi);
        bool active = cos_theta_i > 0.f && cos_theta_o > 0.f && dot(wi, m) > 0.f && dot(wo, m) > 0.f;
        if (!active) return 0.f;
        Microfacet distr((int) b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        float result;
        if (b.sample_visible) result = distr.eval(m) * distr.smith_g1(wi, m) / (4.f * cos_theta_i);
        else result = distr.pdf(wi, m) / (4.f * dot(wo, m));
        return result;
    }
    return 0.f;
}
template <int S, int BK> BF_D Spec<S> onesided_sample_impl(const DevScene &sc, const bf_bsdf_desc &b, const Spec<S> &wl, V3 wi, V2 sample2, BSDFSample &bs) {
    Spec<S> r;
    const uint32_t kind = BK >= 0 ? (uint32_t) BK : b.kind;
    bs.wo = V3(0.f, 0.f, 0.f); bs.pdf = 0.f; bs.eta = 0.f; bs.delta = false;
    float cos_theta_i = wi.z;
    if (kind == BF_BSDF_DIFFUSE) {
        if (!(cos_theta_i > 0.f)) return r;
        bs.wo = square_to_cosine_hemisphere(sample2);
        bs.pdf = square_to_cosine_hemisphere_pdf(bs.wo);
        bs.eta = 1.f;
        if (bs.pdf > 0.f) r = spectrum_eval<S>(sc, b.reflectance, wl);
        return r;
    } else if (kind == BF_BSDF_ROUGHCONDUCTOR) {
        if (!(cos_theta_i > 0.f)) return r;
        Microfacet distr((int) b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        V3 m; distr.sample(wi, sample2, m, bs.pdf);
        bs.wo = reflect(wi, m);
        bs.eta = 1.f;
        bool active = bs.pdf != 0.f && bs.wo.z > 0.f;
        float weight;
        if (b.sample_visible) weight = distr.smith_g1(bs.wo, m);
        else weight = distr.G(wi, bs.wo, m) * dot(wi, m) / (cos_theta_i * m.z);
        bs.pdf /= 4.f * dot(bs.wo, m);
        Spec<S> eta = spectrum_eval<S>(sc, b.eta, wl), kk = spectrum_eval<S>(sc, b.k, wl);
        Spec<S> sr(1.f);
        if (b.specular_reflectance != BF_NONE) sr = spectrum_eval<S>(sc, b.specular_reflectance, wl);
        float cosm = dot(wi, m);
#pragma unroll
        for (int k = 0; k < S; ++k) {
            float F = fresnel_conductor(cosm, eta.v[k], kk.v[k]);
            float w = weight;
            if (b.specular_reflectance != BF_NONE) w *= sr.v[k];
            r.v[k] = active ? F * w : 0.f;
        }
        return r;
    } else if (kind == BF_BSDF_CONDUCTOR) {
        if (!(cos_theta_i > 0.f)) return r;
        bs.wo = V3(-wi.x, -wi.y, wi.z); bs.eta = 1.f; bs.pdf = 1.f; bs.delta = true;
        Spec<S> eta = spectrum_eval<S>(sc, b.eta, wl), kk = spectrum_eval<S>(sc, b.k, wl);
        Spec<S> sr(1.f);
        if (b.specular_reflectance != BF_NONE) sr = spectrum_eval<S>(sc, b.specular_reflectance, wl);
#pragma unroll
        for (int k = 0; k < S; ++k) {
            float v = (b.specular_reflectance != BF_NONE) ? sr.v[k] : 1.f;
            r.v[k] = v * fresnel_conductor(cos_theta_i, eta.v[k], kk.v[k]);
        }
        return r;
    }
    return r;
}
// out-of-line generic versions (one copy each keeps the generic shade kernel's code small)
template <int S> BF_NI Spec<S> onesided_eval(const DevScene &sc, const bf_bsdf_desc &b, const Spec<S> &wl, V3 wi, V3 wo) { return onesided_eval_impl<S, -1>(sc, b, wl, wi, wo); }
BF_NI float onesided_pdf(const bf_bsdf_desc &b, V3 wi, V3 wo) { return onesided_pdf_impl<-1>(b, wi, wo); }
template <int S> BF_NI Spec<S> onesided_sample(const DevScene &sc, const bf_bsdf_desc &b, const Spec<S> &wl, V3 wi, V2 sample2, BSDFSample &bs) { return onesided_sample_impl<S, -1>(sc, b, wl, wi, sample2, bs); }
// twosided.cpp:94-181: strict inequalities, cos_theta == 0 gives a zero result
// Front / back selection of twosided: the branch taken decides (descriptor, wi, wo); one call site each.
BF_D const bf_bsdf_desc *twosided_select(const DevScene &sc, const bf_bsdf_desc &b, V3 &wi, bool &flipped, bool &dead) {
    flipped = false; dead = false;
    if (!b.two_sided) return &b;
    if (wi.z > 0.f) return &b;
    if (wi.z < 0.f) {
        flipped = true;
        wi = V3(wi.x, wi.y, wi.z * -1.f);
        return b.back_bsdf != BF_NONE ? &sc.bsdfs[b.back_bsdf] : &b;
    }
    dead = true;   // cos_theta == 0 (or NaN): zero result
    return &b;
}
template <int S, int BK = -1> BF_D Spec<S> bsdf_eval(const DevScene &sc, uint32_t id, const Spec<S> &wl, V3 wi, V3 wo) {
    bool flipped, dead;
    const bf_bsdf_desc *b = twosided_select(sc, sc.bsdfs[id], wi, flipped, dead);
    if (dead) return Spec<S>();
    if (flipped) wo = V3(wo.x, wo.y, wo.z * -1.f);
    if constexpr (BK >= 0) return onesided_eval_impl<S, BK>(sc, *b, wl, wi, wo);
    else return onesided_eval<S>(sc, *b, wl, wi, wo);
}
template <int BK = -1> BF_D float bsdf_pdf(const DevScene &sc, uint32_t id, V3 wi, V3 wo) {
    bool flipped, dead;
    const bf_bsdf_desc *b = twosided_select(sc, sc.bsdfs[id], wi, flipped, dead);
    if (dead) return 0.f;
    if (flipped) wo = V3(wo.x, wo.y, wo.z * -1.f);
    if constexpr (BK >= 0) return onesided_pdf_impl<BK>(*b, wi, wo);
    else return onesided_pdf(*b, wi, wo);
}
template <int S, int BK = -1> BF_D Spec<S> bsdf_sample(const DevScene &sc, uint32_t id, const Spec<S> &wl, V3 wi, V2 s2, BSDFSample &bs) {
    bool flipped, dead;
    const bf_bsdf_desc *b = twosided_select(sc, sc.bsdfs[id], wi, flipped, dead);
    if (dead) { bs.wo = V3(0.f, 0.f, 0.f); bs.pdf = 0.f; bs.eta = 0.f; bs.delta = false; return Spec<S>(); }
    Spec<S> r;
    if constexpr (BK >= 0) r = onesided_sample_impl<S, BK>(sc, *b, wl, wi, s2, bs);
    else r = onesided_sample<S>(sc, *b, wl, wi, s2, bs);
    if (flipped) bs.wo.z *= -1.f;
    return r;
}

// twosided(diffuse) / diffuse as the material-class-0 shade kernel uses it: select() is twosided.cpp's front / back choice for
// this lane's wi, texture() the reflectance lookup at the lane's current wavelengths (evaluated once; a second call is a
// no-op), eval / pdf / sample are diffuse.cpp:78-135 behind twosided.cpp:94-181 on those two.
template <int S> struct DiffuseLane {
    const bf_bsdf_desc *b;
    V3 wi;
    bool flipped, dead, have_refl;
    Spec<S> refl;
    BF_D void select(const DevScene &sc, uint32_t id, V3 wi_) {
        wi = wi_; have_refl = false;
        b = twosided_select(sc, sc.bsdfs[id], wi, flipped, dead);
    }
    BF_D void texture(const DevScene &sc, const Spec<S> &wl) {
        if (have_refl) return;
        have_refl = true;
        if (!dead) refl = spectrum_eval<S>(sc, b->reflectance, wl);
    }
    BF_D Spec<S> eval(V3 wo) const {
        Spec<S> r;
        if (dead) return r;
        if (flipped) wo = V3(wo.x, wo.y, wo.z * -1.f);
        if (!(wi.z > 0.f && wo.z > 0.f)) return r;
#pragma unroll
        for (int k = 0; k < S; ++k) r.v[k] = refl.v[k] * BF_INV_PI * wo.z;
        return r;
    }
    BF_D float pdf(V3 wo) const {
        if (dead) return 0.f;
        if (flipped) wo = V3(wo.x, wo.y, wo.z * -1.f);
        float p = square_to_cosine_hemisphere_pdf(wo);
        return (wi.z > 0.f && wo.z > 0.f) ? p : 0.f;
    }
    BF_D Spec<S> sample(V2 sample2, BSDFSample &bs) const {
        Spec<S> r;
        bs.wo = V3(0.f, 0.f, 0.f); bs.pdf = 0.f; bs.eta = 0.f; bs.delta = false;
        if (dead || !(wi.z > 0.f)) return r;
        bs.wo = square_to_cosine_hemisphere(sample2);
        bs.pdf = square_to_cosine_hemisphere_pdf(bs.wo);
        bs.eta = 1.f;
        if (bs.pdf > 0.f) r = refl;
        if (flipped) bs.wo.z *= -1.f;
        return r;
    }
};

// ---- signal model: wignertransmitter.cpp:111-188 (== phasedtransmitter.cpp:176-253, wignerreceiver.cpp:116-201)
// rcp(repfreq), t_ext / 2 and f_ext / t_ext are evaluated once per scene (DevEndpoint::sig_*), with the reference's operations.
BF_D float eval_signal_impl(const DevEndpoint &ep, float time, float frequency0) {
    const bf_signal_desc &sg = ep.signal;
    if (sg.signaltype == BF_SIG_LINFMCW) {
        float t = fmodulo(time, ep.sig_period);
        float ti = ep.sig_ti;
        float fi = sg.f_centre + ep.sig_slope * (t - ti);
        return rect((t - ti) / sg.t_ext) > 0.f ? wchirp(t - ti, frequency0 - fi, sg.t_ext, sg.amplitude) : 0.f;
    } else if (sg.signaltype == BF_SIG_PULSE) {
        float t = fmodulo(time, ep.sig_period);
        float ti = ep.sig_ti;
        float fi = sg.f_centre;
        return rect((t - ti) / sg.t_ext) > 0.f ? wchirp(t - ti, frequency0 - fi, sg.t_ext, sg.amplitude) : 0.f;
    }
    return sg.amplitude * sg.amplitude;
}
BF_D float sample_delta_frequency_impl(const DevEndpoint &ep, float time) {
    const bf_signal_desc &sg = ep.signal;
    if (sg.signaltype == BF_SIG_LINFMCW) {
        float t = fmodulo(time, ep.sig_period);
        float ti = ep.sig_ti;
        return sg.f_centre + ep.sig_slope * (t - ti);
    }
    return sg.f_centre;
}
BF_NI float eval_signal_ni(const DevEndpoint &ep, float time, float frequency0) { return eval_signal_impl(ep, time, frequency0); }
BF_NI float sample_delta_frequency_ni(const DevEndpoint &ep, float time) { return sample_delta_frequency_impl(ep, time); }
// INL: the endpoint is a kernel-parameter copy in a specialised kernel (inline, constant-bank operands); otherwise one out-of-line copy
template <bool INL> BF_D float eval_signal(const DevEndpoint &ep, float time, float frequency0) {
    if constexpr (INL) return eval_signal_impl(ep, time, frequency0); else return eval_signal_ni(ep, time, frequency0);
}
template <bool INL> BF_D float sample_delta_frequency(const DevEndpoint &ep, float time) {
    if constexpr (INL) return sample_delta_frequency_impl(ep, time); else return sample_delta_frequency_ni(ep, time);
}
template <int S> BF_D Spec<S> sample_shifted(float sample) {                    // math.h:494-517
    Spec<S> r;
#pragma unroll
    for (int k = 0; k < S; ++k) { float v = sample + (float) k / (float) S; if (v > 1.f) v -= 1.f; r.v[k] = v; }
    return r;
}

// ---- transmitters / emitters ------------------------------------------------------------------------------
// `wl` is the path's current wavelength state: the reference mutates si.wavelengths through a
// const_cast (wignertransmitter.cpp:220-221,439-440) and the last write decides the frequency bin.
// TXK >= 0 (specialised shade kernels): the scene has ONE transmitter of that kind; its endpoint and rectangle are read
// from the copies in the kernel parameter block.  TXK = -1: everything is looked up in the scene tables.
template <int TXK> BF_D const DevEndpoint &tx_ref(const DevScene &sc, uint32_t i) {
    if constexpr (TXK >= 0) return sc.tx0_ep; else return sc.txs[i];
}
template <int TXK> BF_D const DevRect &tx_rect_ref(const DevScene &sc, const DevEndpoint &tx) {
    if constexpr (TXK >= 0) return sc.tx0_rect; else return sc.rects[sc.shapes[tx.shape].index];
}
template <int S, int TXK> BF_D Spec<S> tx_eval_impl(const DevScene &sc, const DevEndpoint &tx, const SI &si, float si_time, Spec<S> &wl) {
    Spec<S> r;
    constexpr bool INL = TXK >= 0;
    const uint32_t kind = TXK >= 0 ? (uint32_t) TXK : tx.kind;
    if (kind == BF_TX_AREA) {                                   // areatransmitter.cpp:65-74
        if (!(si.wi.z > 0.f)) return r;
        Spec<S> rad = spectrum_eval<S>(sc, tx.texture, wl);
        float area = tx_rect_ref<TXK>(sc, tx).area;
#pragma unroll
        for (int k = 0; k < S; ++k) r.v[k] = rad.v[k] * area;
        return r;
    } else if (kind == BF_TX_AREALIGHT) {                       // src/emitters/area.cpp:63-72
        if (!(si.wi.z > 0.f)) return r;
        return spectrum_eval<S>(sc, tx.texture, wl);
    } else if (kind == BF_TX_WIGNER || kind == BF_TX_PHASED) {  // wignertransmitter.cpp:193-271, phasedtransmitter.cpp:296-381
        float signal_power;
        if (tx.signal.resample_freq) {
            float f = sample_delta_frequency<INL>(tx, si_time);
            signal_power = 1.f;
            float w = sc.C.c * rcp(f) * 1e9f;
#pragma unroll
            for (int k = 0; k < S; ++k) wl.v[k] = w;
        } else {
            signal_power = eval_signal<INL>(tx, si_time, freq_from_wl(sc.C.c, wl.v[0]));
        }
        Spec<S> geom = spectrum_eval<S>(sc, tx.texture, wl);
        const DevRect &rc = tx_rect_ref<TXK>(sc, tx);
        if (kind == BF_TX_WIGNER) {
            // the reference reads an uninitialised ds.d here; defined as the zero vector (oracle.cpp tx_eval)
            float w = rect_sample_wigner(rc, si.p, V3(0.f, 0.f, 0.f), wl.v[0]);
            if (!(si.wi.z > 0.f)) return r;
#pragma unroll
            for (int k = 0; k < S; ++k) r.v[k] = signal_power * tx.signal.gain * (geom.v[k] * w) * BF_TWO_PI;
        } else {
            float inv_area = rcp(rc.area);
            float w = phased_sample_wigner(sc, tx, si.p, V3(0.f, 0.f, 0.f), wl.v[0]);
            if (!(si.wi.z > 0.f)) return r;
#pragma unroll
            for (int k = 0; k < S; ++k) r.v[k] = signal_power * tx.signal.gain * (geom.v[k] * inv_area * w);
        }
        return r;
    }
    return r;  // spot: eval = 0 (spot.cpp:166)
}
template <int S, int TXK> BF_D Spec<S> tx_sample_direction_impl(const DevScene &sc, const DevEndpoint &tx, V3 it_p, float it_time, Spec<S> &wl, V2 sample, DirectionSample &ds) {
    Spec<S> out;
    constexpr bool INL = TXK >= 0;
    const uint32_t kind = TXK >= 0 ? (uint32_t) TXK : tx.kind;
    if (kind == BF_TX_SPOT) {                                   // spot.cpp:97-115, 141-164
        ds.p = V3(tx.to_world.m[0][3], tx.to_world.m[1][3], tx.to_world.m[2][3]);
        ds.n = V3(0.f, 0.f, 0.f); ds.pdf = 1.f; ds.time = it_time; ds.delta = true; ds.object = 0;
        ds.d = ds.p - it_p; ds.dist = norm(ds.d);
        float inv_dist = rcp(ds.dist);
        ds.d = ds.d * inv_dist;
        V3 local_d = xf_vector(tx.to_object, -ds.d);
        Spec<S> result = spectrum_eval<S>(sc, tx.texture, wl);
        V3 ld = normalize(local_d);
        float cos_theta = ld.z;
#pragma unroll
        for (int k = 0; k < S; ++k) {
            float beam = cos_theta >= tx.cos_beam ? result.v[k] : result.v[k] * ((tx.cutoff - acosf(cos_theta)) * tx.inv_transition);
            float v = cos_theta <= tx.cos_cutoff ? 0.f : beam;
            out.v[k] = v * (inv_dist * inv_dist);
        }
        return out;
    }
    const DevRect &rc = tx_rect_ref<TXK>(sc, tx);
    ds = shape_sample_direction(rc, it_p, it_time, sample);
    bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
    if (kind == BF_TX_AREA || kind == BF_TX_AREALIGHT) {     // areatransmitter.cpp:131-182, area.cpp
        Spec<S> rad = spectrum_eval<S>(sc, tx.texture, wl);
#pragma unroll
        for (int k = 0; k < S; ++k) out.v[k] = active ? rad.v[k] / ds.pdf : 0.f;
        return out;
    }
    // wignertransmitter.cpp:373-534
    Spec<S> tex = spectrum_eval<S>(sc, tx.texture, wl);
    Spec<S> geom;
#pragma unroll
    for (int k = 0; k < S; ++k) geom.v[k] = tex.v[k] / ds.pdf;
    if ((double) ds.dist > 5e-7) ds.time += -ds.dist / sc.C.c;
    float signal_power;
    if (tx.signal.resample_freq) {
        float f = sample_delta_frequency<INL>(tx, ds.time);
        signal_power = 1.f;
        float w = sc.C.c * rcp(f) * 1e9f;
#pragma unroll
        for (int k = 0; k < S; ++k) wl.v[k] = w;
    } else {
        signal_power = eval_signal<INL>(tx, ds.time, freq_from_wl(sc.C.c, wl.v[0]));
    }
    if (kind == BF_TX_PHASED) {                                 // phasedtransmitter.cpp:596-654
        float w = phased_sample_wigner(sc, tx, ds.p, ds.d * -1.f, wl.v[0]);
#pragma unroll
        for (int k = 0; k < S; ++k) geom.v[k] *= w;
        ds.pdf *= w;
        ds.pdf = sqrtf(ds.pdf * ds.pdf);
#pragma unroll
        for (int k = 0; k < S; ++k) out.v[k] = active ? signal_power * tx.signal.gain * geom.v[k] * 1.f : 0.f;
        return out;
    }
    float w;
    if constexpr (INL) w = rect_sample_wigner_impl(rc, ds.p, ds.d * -1.f, inv_wavelength_d(wl.v[0]));
    else w = rect_sample_wigner(rc, ds.p, ds.d * -1.f, wl.v[0]);
#pragma unroll
    for (int k = 0; k < S; ++k) geom.v[k] *= w;
    ds.pdf *= w;
    float extents = rc.two_pi_inv_area;                         // rcp(area) * 2 pi
#pragma unroll
    for (int k = 0; k < S; ++k) out.v[k] = active ? signal_power * tx.signal.gain * geom.v[k] * extents : 0.f;
    return out;
}
template <int TXK> BF_D float tx_pdf_direction_impl(const DevScene &sc, const DevEndpoint &tx, float wl0, const DirectionSample &ds) {
    const uint32_t kind = TXK >= 0 ? (uint32_t) TXK : tx.kind;
    if (kind == BF_TX_SPOT) return 0.f;
    const DevRect &rc = tx_rect_ref<TXK>(sc, tx);
    float dp = dot(ds.d, ds.n);
    bool active = dp < 0.f;
    float value = shape_pdf_direction(rc, ds);
    if (kind == BF_TX_WIGNER)                                   // wignertransmitter.cpp:540-577
        value *= rect_sample_wigner(rc, ds.p, ds.d * -1.f, wl0);
    else if (kind == BF_TX_PHASED) {                            // phasedtransmitter.cpp:660-702
        value *= phased_sample_wigner(sc, tx, ds.p, ds.d * -1.f, wl0);
        value = sqrtf(value * value);
    }
    return active ? value : 0.f;
}

// Out-of-line generic versions.  tx_eval and tx_pdf_direction only run for a path that HITS a transmitter's aperture (a few
// millimetres wide): they are always the out-of-line generic copies, which keeps them out of the specialised kernels' hot code.
template <int S> BF_NI Spec<S> tx_eval_generic(const DevScene &sc, const DevEndpoint &tx, const SI &si, float si_time, Spec<S> &wl) { return tx_eval_impl<S, -1>(sc, tx, si, si_time, wl); }
template <int S> BF_NI Spec<S> tx_sample_direction_generic(const DevScene &sc, const DevEndpoint &tx, V3 it_p, float it_time, Spec<S> &wl, V2 sample, DirectionSample &ds) { return tx_sample_direction_impl<S, -1>(sc, tx, it_p, it_time, wl, sample, ds); }
BF_NI float tx_pdf_direction_generic(const DevScene &sc, const DevEndpoint &tx, float wl0, const DirectionSample &ds) { return tx_pdf_direction_impl<-1>(sc, tx, wl0, ds); }
template <int S> BF_D Spec<S> tx_eval(const DevScene &sc, const DevEndpoint &tx, const SI &si, float si_time, Spec<S> &wl) { return tx_eval_generic<S>(sc, tx, si, si_time, wl); }
template <int S, int TXK = -1> BF_D Spec<S> tx_sample_direction(const DevScene &sc, uint32_t index, V3 it_p, float it_time, Spec<S> &wl, V2 sample, DirectionSample &ds) {
    if constexpr (TXK >= 0) return tx_sample_direction_impl<S, TXK>(sc, sc.tx0_ep, it_p, it_time, wl, sample, ds);
    else return tx_sample_direction_generic<S>(sc, sc.txs[index], it_p, it_time, wl, sample, ds);
}
BF_D float tx_pdf_direction(const DevScene &sc, const DevEndpoint &tx, float wl0, const DirectionSample &ds) { return tx_pdf_direction_generic(sc, tx, wl0, ds); }

// ---- receivers / sensors -------------------------------------------------------------------------------------
struct RayGen { V3 o, d; float mint, maxt, time; };
// spectrum.h:310-314, 360-371
template <int S> BF_D void sample_wavelength(const DevScene &sc, float sample, Spec<S> &wl, Spec<S> &weight) {
    if constexpr (S == 1) { wl = Spec<S>(__int_as_float(0x7fc00000)); weight = Spec<S>(1.f); }
    else {
        Spec<S> s = sample_shifted<S>(sample);
        float range = sc.C.cie_max_nm - sc.C.cie_min_nm;
#pragma unroll
        for (int k = 0; k < S; ++k) { wl.v[k] = s.v[k] * range + sc.C.cie_min_nm; weight.v[k] = range; }
    }
}
// omnidirectional.cpp:71-107, fluxmeter.cpp:63-84, wignerreceiver.cpp:150-269, perspective.cpp:166-196
// RXK >= 0: the receiver kind is a compile-time constant (specialised primary kernels); -1: read from the scene.
// The receiver and its rectangle are the copies in the kernel parameter block (sc.rx_ep, sc.rx_rect); the frame around
// the sampled position's normal (= the rectangle's normal) is the rectangle's precomputed coordinate_system(fn).
// inv_wl: 1.0 / (wl[0] * 1e-9) in double, shared with the caller's `MTS_C*rcp(wavelengths*1e-9)` (integrator.cpp:1603).
template <int S, int RXK = -1> BF_D void rx_sample_ray(const DevScene &sc, float time, float sample1, V2 sample2, V2 sample3, RayGen &ray, Spec<S> &wl, Spec<S> &weight, double &inv_wl) {
    const DevEndpoint &d = sc.rx_ep;
    constexpr bool INL = RXK >= 0;
    const uint32_t kind = RXK >= 0 ? (uint32_t) RXK : d.kind;
    ray.mint = BF_RAY_EPS; ray.maxt = bf_inf(); ray.time = time;
    if (kind == BF_RX_PERSPECTIVE) {
        sample_wavelength<S>(sc, sample1, wl, weight);
        V3 near_p = xf_point(d.sample_to_camera, V3(sample2.x, sample2.y, 0.f));
        V3 dd = normalize(near_p);
        float inv_z = rcp(dd.z);
        ray.mint = d.near_clip * inv_z; ray.maxt = d.far_clip * inv_z;
        ray.o = xf_point_affine(d.to_world, V3(0.f, 0.f, 0.f));
        ray.d = xf_vector(d.to_world, dd);
        inv_wl = inv_wavelength_d(wl.v[0]);
        return;
    }
    const DevRect &r = sc.rx_rect;
    const Frame fr(r.cs_s, r.cs_t, r.fn);                          // Frame3f(ps.n), ps.n = m_frame.n
    if (kind == BF_RX_OMNIDIRECTIONAL || kind == BF_RX_FLUXMETER) {
        PositionSample ps = rect_sample_position(r, time, sample2);
        V3 local = square_to_cosine_hemisphere(sample3);
        Spec<S> ww; sample_wavelength<S>(sc, sample1, wl, ww);
        ray.o = ps.p; ray.d = fr.to_world(local);
        float k = kind == BF_RX_OMNIDIRECTIONAL ? r.area : BF_PI;
#pragma unroll
        for (int i = 0; i < S; ++i) weight.v[i] = ww.v[i] * k;
        inv_wl = inv_wavelength_d(wl.v[0]);
        return;
    }
    Spec<S> freq, power;
    const bf_signal_desc &sg = d.signal;
    if (d.receive_type == BF_RECV_RAW || d.receive_type == BF_RECV_RAW_RESAMPLE || !sg.sig_is_delta) {
        Spec<S> fs = sample_shifted<S>(sample1);
#pragma unroll
        for (int k = 0; k < S; ++k) freq.v[k] = fs.v[k] * sg.f_ext + (sg.f_centre - sg.f_ext / 2);
        if (d.receive_type == BF_RECV_RAW || d.receive_type == BF_RECV_RAW_RESAMPLE) power = Spec<S>(1.f);
        else power = Spec<S>(eval_signal<false>(d, time, freq.v[0]));
    } else {
        freq = Spec<S>(sample_delta_frequency<INL>(d, time)); power = Spec<S>(1.f);
    }
#pragma unroll
    for (int k = 0; k < S; ++k) wl.v[k] = sc.C.c * rcp(freq.v[k]) * 1e9f;
    inv_wl = inv_wavelength_d(wl.v[0]);
    PositionSample ps = rect_sample_position(r, time, sample2);
    V3 dloc = square_to_cosine_hemisphere(sample3);
    float geom_gain;
    if (kind == BF_RX_PHASED) {                                  // phasedreceiver.cpp:313-375 (:353)
        float w = phased_sample_wigner(sc, d, ps.p, dloc, wl.v[0]);
        float dn = dot(dloc, ps.n);                                // local direction against the world normal, as written
        geom_gain = w * ps.pdf * (1 - dn * dn * dn * dn);
    } else {
        float w;
        if constexpr (INL) w = rect_sample_wigner_impl(r, ps.p, dloc, inv_wl);
        else w = rect_sample_wigner(r, ps.p, dloc, wl.v[0]);
        geom_gain = w * ps.pdf;
    }
    float extents0 = r.area_pi;                                    // area * pi
    ray.o = ps.p; ray.d = fr.to_world(dloc);
#pragma unroll
    for (int k = 0; k < S; ++k) {
        float extents = extents0;
        if (!sg.sig_is_delta) extents *= sc.C.c * rcp(sg.f_ext) * 1e9f;
        weight.v[k] = power.v[k] * sg.gain * geom_gain * extents;
    }
}

BF_D float mis_weight(float pdf_a, float pdf_b) {                     // pathtimefrequency.cpp:472-477
    pdf_a *= pdf_a; pdf_b *= pdf_b;
    return pdf_a > 0.f ? pdf_a / (pdf_a + pdf_b) : 0.f;
}

}  // namespace bf
