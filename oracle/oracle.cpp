// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// Scalar CPU restatement of the reference's radar path-integration hot path
// (JacobMackay/beifong, paths below are under /root/reference).  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
// library; the product (beifong_b200/) never does.
//
// PARITY PINNING: the reference cannot be compiled here (ext/enoki is an empty submodule, every
// header needs it).  What is pinned against the reference's own tests: TEA (test_random.py:9-27),
// warps (test_warp.py:68-75,142-149), Möller–Trumbore / rectangle conventions
// (test_mesh.py:257-298, test_rectangle.py:37-62), accel == brute force (test_kdtrees.py:25-91),
// diffuse / microfacet closed forms (test_diffuse.py:16-39, test_microfacet.py), and the vendored
// Embree 3.8.0 built from the reference sources (oracle/_ref/libembree3.so) for first hits.
// The radar-specific arithmetic (receive_sample, pathtimefrequency, wigner endpoints,
// SignalBlock::put) has NO test or golden vector in the reference: for those rows
// this oracle is "parity unpinned" and is anchored on the call sites it cites line by line.
//
// Arithmetic conventions: see orc_math.h.  Spectrum width S (1 mono / 4 spectral) is a run-time
// constant here; arrays are float[4].
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include <thread>
#include <dlfcn.h>
#include "orc_math.h"
#include "../include/beifong_b200/flat_scene.h"

namespace orc {

// Expressions such as `MTS_C*rcp(wavelengths*1e-9)` contain a double literal: Enoki follows the C++
// promotion rules, so the product, the reciprocal and the multiplication by MTS_C are carried out
// in double and rounded to float once, on assignment (integrator.cpp:1592,1603,1619;
// wignertransmitter.cpp:233; rectangle.cpp:160).
static inline double freq_from_wl_d(float c, float wl_nm) { return (double) c * (1.0 / ((double) wl_nm * 1e-9)); }
static inline float freq_from_wl(float c, float wl_nm) { return (float) freq_from_wl_d(c, wl_nm); }

struct Spec {
    float v[4];
    Spec() { v[0] = v[1] = v[2] = v[3] = 0.f; }
    explicit Spec(float s) { v[0] = v[1] = v[2] = v[3] = s; }
};
static int S_ = 1;  // set per scene before use (thread-safe because each Scene copies it)

struct Ray {
    V3 o, d;
    float mint = RayEpsilon, maxt = Infinity, time = 0.f, phase = 0.f;  // ray.h:30-38
    Spec wl;
    V3 at(float t) const { return fmadd3(d, t, o); }                  // ray.h:96
};

// ---- records (include/mitsuba/render/records.h:20-208, interaction.h:33-172) -----------------
struct PositionSample { V3 p, n; V2 uv; float time = 0, pdf = 0; bool delta = false; int object = -1; };
struct DirectionSample : PositionSample { V3 d; float dist = 0; };
struct SI {
    float t = Infinity, time = 0, phase = 0;
    Spec wl;
    V3 p;
    int shape = -1;
    V2 uv;
    V3 n;
    Frame sh;
    V3 dp_du, dp_dv, wi;
    uint32_t prim = 0;
    bool is_valid() const { return t != Infinity; }
    V3 to_world(V3 v) const { return sh.to_world(v); }
    V3 to_local(V3 v) const { return sh.to_local(v); }
};
struct PI { float t = Infinity; V2 uv; uint32_t prim = 0; int shape = -1; };

struct SpectrumTex { uint32_t kind = 0; std::vector<float> lambda, value; float inv_interval = 0; };
struct MeshData { std::vector<float> pos; std::vector<uint32_t> idx; std::vector<float> nrm; uint32_t nv = 0, nf = 0;
                  std::vector<float> pos_rest, nrm_rest; /* as imported: BF_TAG_MESH_MOTION moves the mesh from this pose */ };
struct RectData { M4 to_world, to_object; Frame frame; float inv_area = 0, area = 0; };
struct ShapeRec { int type = 0; /*0 mesh 1 rect*/ int index = 0; uint32_t bsdf = 0; int tx = -1; bool is_rx = false; };
// Virtual element (i, j) of a phased array: phasedtransmitter.cpp:119-167 == phasedreceiver.cpp:115-173
struct VElem { M4 to_object; V3 ns, nt, nn; V3 r_dash; float psi_re = 1.f, psi_im = 0.f; };
struct TxRec { int shape = -1; bf_endpoint_desc d; M4 to_world, to_object; float inv_transition = 0, cos_cutoff = 0, cos_beam = 0, cutoff = 0;
               std::vector<VElem> velems; };

// ---- minimal BVH over triangles (test-side accelerator; equality with brute force is tested) ---
struct BvhNode { float lo[3], hi[3]; uint32_t left, count; /* count>0: leaf, left = first */ };
struct TriRef { uint32_t shape, prim; };

struct Scene;
typedef void *RTCDevice; typedef void *RTCScene; typedef void *RTCGeometry;
struct EmbreeApi {
    void *lib = nullptr;
    RTCDevice (*rtcNewDevice)(const char *) = nullptr;
    RTCScene (*rtcNewScene)(RTCDevice) = nullptr;
    void (*rtcSetSceneFlags)(RTCScene, int) = nullptr;
    RTCGeometry (*rtcNewGeometry)(RTCDevice, int) = nullptr;
    void (*rtcSetSharedGeometryBuffer)(RTCGeometry, int, unsigned, int, const void *, size_t, size_t, size_t) = nullptr;
    void (*rtcCommitGeometry)(RTCGeometry) = nullptr;
    unsigned (*rtcAttachGeometry)(RTCScene, RTCGeometry) = nullptr;
    void (*rtcReleaseGeometry)(RTCGeometry) = nullptr;
    void (*rtcCommitScene)(RTCScene) = nullptr;
    void (*rtcIntersect1)(RTCScene, void *, void *) = nullptr;
    void (*rtcOccluded1)(RTCScene, void *, void *) = nullptr;
    void (*rtcReleaseScene)(RTCScene) = nullptr;
    void (*rtcReleaseDevice)(RTCDevice) = nullptr;
};
// Embree 3 ABI structs (public rtcore_ray.h layout)
struct alignas(16) RTCRay_ { float org_x, org_y, org_z, tnear, dir_x, dir_y, dir_z, time, tfar; unsigned mask, id, flags; };
struct alignas(16) RTCHit_ { float Ng_x, Ng_y, Ng_z, u, v; unsigned primID, geomID, instID[1]; };
struct alignas(16) RTCRayHit_ { RTCRay_ ray; RTCHit_ hit; };
struct RTCIntersectContext_ { int flags; void *filter; unsigned instID[1]; };

struct Scene {
    bf_constants C;
    int S = 1;
    std::vector<SpectrumTex> spectra;
    std::vector<bf_bsdf_desc> bsdfs;
    std::vector<MeshData> meshes;
    std::vector<RectData> rects;
    std::vector<ShapeRec> shapes;
    std::vector<M4> velocities;      // per shape, identity by default (shape.cpp:39-42)
    std::vector<TxRec> txs;
    int rx_shape = -1; bf_endpoint_desc rx; bool has_rx = false;
    M4 rx_to_world;
    std::vector<VElem> rx_velems;
    bf_adc_desc adc; bf_integrator_desc integ;
    // accel
    int accel = 0;  // 0 brute force, 1 own BVH, 2 embree
    std::vector<BvhNode> nodes; std::vector<TriRef> tris;
    EmbreeApi em; RTCDevice em_dev = nullptr; RTCScene em_scene = nullptr;
    std::vector<int> em_geom_shape;
    std::string err;
    float wl_half_band_m() const { return (C.lambda_max_nm - C.lambda_min_nm) / 2 * 1e-9f; }
};

// ---- spectra: uniform.cpp:27-40, distr_1d.h:387-403 (regular), :664-687 (irregular) -------------
static float spectrum_eval_1(const Scene &sc, const SpectrumTex &t, float wl) {
    switch (t.kind) {
        case BF_SPEC_UNIFORM:
            if (sc.S > 1) return (wl >= sc.C.lambda_min_nm && wl <= sc.C.lambda_max_nm) ? t.value[0] : 0.f;
            return t.value[0];
        case BF_SPEC_REGULAR: {
            bool active = wl >= t.lambda.front() && wl <= t.lambda.back();
            float x = (wl - t.lambda.front()) * t.inv_interval;
            uint32_t n = (uint32_t) t.value.size();
            int64_t xi = (x == x) ? (int64_t) std::min(std::max(x, -1.f), 4.0e9f) : 0;
            uint32_t index = (uint32_t) std::min<int64_t>(std::max<int64_t>(xi, 0), (int64_t) n - 2);
            float y0 = active ? t.value[index] : 0.f, y1 = active ? t.value[index + 1] : 0.f;
            float w1 = x - (float) index, w0 = 1.f - w1;
            return fmadd(w0, y0, w1 * y1);
        }
        case BF_SPEC_IRREGULAR: {
            bool active = wl >= t.lambda.front() && wl <= t.lambda.back();
            uint32_t n = (uint32_t) t.lambda.size();
            uint32_t lo = 0, hi = n;  // enoki::binary_search: first index with !(node < x)
            while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (active && t.lambda[mid] < wl) lo = mid + 1; else hi = mid; }
            uint32_t index = std::max(std::min(lo, n - 1u), 1u) - 1u;
            float x0 = t.lambda[index], x1 = t.lambda[index + 1], y0 = t.value[index], y1 = t.value[index + 1];
            float x = (wl - x0) / (x1 - x0);
            return active ? fmadd(x, y1 - y0, y0) : 0.f;
        }
        default:  // BF_SPEC_ZERO_OUTSIDE_VISIBLE (Texture::D65 default outside 300..830 nm)
            return 0.f;
    }
}
static Spec spectrum_eval(const Scene &sc, uint32_t id, const Spec &wl) {
    Spec r;
    const SpectrumTex &t = sc.spectra[id];
    for (int k = 0; k < sc.S; ++k) r.v[k] = spectrum_eval_1(sc, t, wl.v[k]);
    return r;
}
static float spectrum_eval_scalar(const Scene &sc, uint32_t id) {  // Texture::eval_1 of a uniform
    return sc.spectra[id].value[0];
}

// ---- Mesh: include/mitsuba/render/mesh.h:194-226, src/librender/mesh.cpp:452-548 ----------------
static inline V3 mesh_vertex(const MeshData &m, uint32_t i) { return V3(m.pos[3 * i], m.pos[3 * i + 1], m.pos[3 * i + 2]); }
static inline V3 mesh_normal(const MeshData &m, uint32_t i) { return V3(m.nrm[3 * i], m.nrm[3 * i + 1], m.nrm[3 * i + 2]); }
static inline bool ray_intersect_triangle(const MeshData &m, uint32_t index, const Ray &ray, float &t_out, float &u_out, float &v_out) {
    uint32_t i0 = m.idx[3 * index], i1 = m.idx[3 * index + 1], i2 = m.idx[3 * index + 2];
    V3 p0 = mesh_vertex(m, i0), p1 = mesh_vertex(m, i1), p2 = mesh_vertex(m, i2);
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = cross(ray.d, e2);
    float inv_det = rcp(dot(e1, pvec));
    V3 tvec = ray.o - p0;
    float u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    float v = dot(ray.d, qvec) * inv_det;
    active &= v >= 0.f && u + v <= 1.f;
    float t = dot(e2, qvec) * inv_det;
    active &= t >= ray.mint && t <= ray.maxt;
    t_out = t; u_out = u; v_out = v;
    return active;
}
static SI mesh_compute_si(const MeshData &m, const Ray &, const PI &pi) {
    float b1 = pi.uv.x, b2 = pi.uv.y, b0 = 1.f - b1 - b2;
    uint32_t i0 = m.idx[3 * pi.prim], i1 = m.idx[3 * pi.prim + 1], i2 = m.idx[3 * pi.prim + 2];
    V3 p0 = mesh_vertex(m, i0), p1 = mesh_vertex(m, i1), p2 = mesh_vertex(m, i2);
    V3 dp0 = p1 - p0, dp1 = p2 - p0;
    SI si;
    si.t = pi.t;
    si.p = p0 * b0 + p1 * b1 + p2 * b2;
    si.n = normalize(cross(dp0, dp1));
    si.uv = V2(b1, b2);
    coordinate_system(si.n, si.dp_du, si.dp_dv);
    if (!m.nrm.empty()) {
        V3 n0 = mesh_normal(m, i0), n1 = mesh_normal(m, i1), n2 = mesh_normal(m, i2);
        si.sh.n = normalize(n0 * b0 + n1 * b1 + n2 * b2);
    } else {
        si.sh.n = si.n;
    }
    return si;
}

// ---- Rectangle: src/shapes/rectangle.cpp:83-95, 111-125, 132-220, 229-298 -----------------------
static RectData make_rect(const float *to_world16) {
    RectData r;
    r.to_world = M4::from_rows16(to_world16);
    r.to_object = inverse(r.to_world);
    V3 dp_du = xf_vector(r.to_world, V3(2, 0, 0)), dp_dv = xf_vector(r.to_world, V3(0, 2, 0));
    V3 normal = normalize(xf_normal(r.to_object, V3(0, 0, 1)));
    r.frame = Frame(dp_du, dp_dv, normal);
    r.area = norm(cross(r.frame.s, r.frame.t));
    r.inv_area = rcp(r.area);
    return r;
}
static inline bool rect_intersect(const RectData &r, const Ray &ray_, float &t_out, float &lx, float &ly) {
    V3 o = xf_point_affine(r.to_object, ray_.o), d = xf_vector(r.to_object, ray_.d);
    float t = -o.z * rcp(d.z);                       // ray.d_rcp.z of the transformed ray
    V3 local = fmadd3(d, t, o);
    t_out = t; lx = local.x; ly = local.y;
    return t >= ray_.mint && t <= ray_.maxt && std::fabs(local.x) <= 1.f && std::fabs(local.y) <= 1.f;
}
static SI rect_compute_si(const RectData &r, const Ray &ray, const PI &pi) {
    SI si;
    si.t = pi.t;
    si.p = ray.at(pi.t);
    si.n = r.frame.n; si.sh.n = r.frame.n;
    si.dp_du = r.frame.s; si.dp_dv = r.frame.t;
    si.uv = V2(fmadd(pi.uv.x, .5f, .5f), fmadd(pi.uv.y, .5f, .5f));
    return si;
}
static PositionSample rect_sample_position(const RectData &r, float time, V2 sample) {
    PositionSample ps;
    ps.p = xf_point_affine(r.to_world, V3(sample.x * 2.f - 1.f, sample.y * 2.f - 1.f, 0.f));
    ps.n = r.frame.n; ps.pdf = r.inv_area; ps.uv = sample; ps.time = time; ps.delta = false;
    return ps;
}
// rectangle.cpp:132-220.  NOTE `m_to_object * ds.p/2` is the projective operator* followed by /2.
static float rect_sample_wigner(const RectData &r, V3 ds_p, V3 ds_d, float wavelength0) {
    float wid_x = norm(r.frame.s), wid_y = norm(r.frame.t);
    V3 r_hat = xf_point(r.to_object, ds_p) / 2.f;
    V3 fs = normalize(r.frame.s), ft = normalize(r.frame.t), fn = normalize(r.frame.n);
    // Transform4f::from_frame: rows = (s, t, n); transform_affine(Vector) = column fma chain
    double inv_wl = 1.0 / ((double) wavelength0 * 1e-9);
    V3 tv(fmadd(fs.z, ds_d.z, fmadd(fs.y, ds_d.y, fs.x * ds_d.x)),
          fmadd(ft.z, ds_d.z, fmadd(ft.y, ds_d.y, ft.x * ds_d.x)),
          fmadd(fn.z, ds_d.z, fmadd(fn.y, ds_d.y, fn.x * ds_d.x)));
    V3 nu_hat((float) ((double) tv.x * inv_wl), (float) ((double) tv.y * inv_wl), (float) ((double) tv.z * inv_wl));
    float gain = 4 * tri(r_hat.x) * tri(r_hat.y) *
                 sinc(TwoPi * nu_hat.x * wid_x * tri(r_hat.x)) *
                 sinc(TwoPi * nu_hat.y * wid_y * tri(r_hat.y));
    return gain;
}

// ---- phased arrays: phasedtransmitter.cpp:104-167, 261-290 == phasedreceiver.cpp:115-173, 275-304 ----
// Load-time geometry of the n x n virtual elements (double arithmetic rounded to fp32 once, as for every
// load-time matrix of this oracle).  Transform::inverse() of a product is the product of the explicit
// inverses (transform.h:53-61,161-171): to_object = S(1/w) T(-r_v) array_loc^-1.
static std::vector<VElem> make_velems(const bf_array_desc &a, const bf_constants &C) {
    std::vector<VElem> out;
    int n = (int) a.n_elems;
    if (n <= 0) return out;
    std::vector<V3> locs;
    for (int i = 0; i < n; ++i) {
        // array_centre - elem_spacing*elem_axis*(i - n/2.f + 0.5)   (even)   |   (i - (n - 1.f)/2.f)   (odd)
        double k = (n % 2 == 0) ? ((double) ((float) i - ((float) n / 2.f)) + 0.5) : (double) ((float) i - ((float) n - 1.f) / 2.f);
        V3 l;
        l.x = (float) (0.0 - (double) (a.elem_spacing[0] * a.elem_axis[0]) * k);
        l.y = (float) (0.0 - (double) (a.elem_spacing[1] * a.elem_axis[1]) * k);
        l.z = (float) (0.0 - (double) (a.elem_spacing[2] * a.elem_axis[2]) * k);
        locs.push_back(l);
    }
    V3 steer(std::sin(a.steering_vector[0]), std::sin(a.steering_vector[1]), std::sin(a.steering_vector[2]));
    M4 A = M4::from_rows16(a.array_loc), Ainv = inverse(A);
    float k_psi = (float) (1.0 / ((double) (C.lambda_max_nm - C.lambda_min_nm) * 1e-9 / 2));
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) {
        VElem e;
        V3 r_v = (locs[i] + locs[j]) / 2.f;
        e.r_dash = locs[i] - locs[j];
        float sx = a.elem_dims[0] / 2, sy = a.elem_dims[1] / 2, sz = a.elem_dims[2];
        M4 TS = M4::identity(), TSinv = M4::identity();
        TS.m[0][0] = sx; TS.m[1][1] = sy; TS.m[2][2] = sz; TS.m[0][3] = r_v.x; TS.m[1][3] = r_v.y; TS.m[2][3] = r_v.z;
        TSinv.m[0][0] = rcp(sx); TSinv.m[1][1] = rcp(sy); TSinv.m[2][2] = rcp(sz);
        TSinv.m[0][3] = -r_v.x * rcp(sx); TSinv.m[1][3] = -r_v.y * rcp(sy); TSinv.m[2][3] = -r_v.z * rcp(sz);
        M4 to_world = mul(A, TS);
        e.to_object = mul(TSinv, Ainv);
        V3 dp_du = xf_vector(to_world, V3(2, 0, 0)), dp_dv = xf_vector(to_world, V3(0, 2, 0));
        V3 normal = normalize(xf_normal(e.to_object, V3(0, 0, 1)));
        e.ns = normalize(dp_du); e.nt = normalize(dp_dv); e.nn = normalize(normal);
        // exp(i * k_psi * dot(array_centre - r_dash, sin(steering)))
        float arg = k_psi * dot(V3(0, 0, 0) - e.r_dash, steer);
        e.psi_re = std::cos(arg); e.psi_im = std::sin(arg);
        out.push_back(e);
    }
    return out;
}
// sample_wigner (phasedtransmitter.cpp:275-290): real part of the phasor sum over the virtual elements
static float phased_sample_wigner(const std::vector<VElem> &ve, const float wid[3], V3 ds_p, V3 ds_d, float wavelength0) {
    float w_re = 0.f, w_im = 0.f;
    double inv_wl = 1.0 / ((double) wavelength0 * 1e-9);
    for (const VElem &e : ve) {
        V3 r_hat = xf_point(e.to_object, ds_p) / 2.f;
        if (!(jabs(r_hat.x) <= 0.5f && jabs(r_hat.y) <= 0.5f)) continue;
        V3 tv(fmadd(e.ns.z, ds_d.z, fmadd(e.ns.y, ds_d.y, e.ns.x * ds_d.x)),
              fmadd(e.nt.z, ds_d.z, fmadd(e.nt.y, ds_d.y, e.nt.x * ds_d.x)),
              fmadd(e.nn.z, ds_d.z, fmadd(e.nn.y, ds_d.y, e.nn.x * ds_d.x)));
        V3 nu_hat((float) ((double) tv.x * inv_wl), (float) ((double) tv.y * inv_wl), (float) ((double) tv.z * inv_wl));
        // W_rect_2D (:261-273)
        float g = 4 * wid[0] * wid[1] * tri(r_hat.x) * tri(r_hat.y) *
                  sinc(TwoPi * nu_hat.x * wid[0] * tri(r_hat.x)) *
                  sinc(TwoPi * nu_hat.y * wid[1] * tri(r_hat.y));
        float th = TwoPi * dot(nu_hat, e.r_dash);
        float a_re = g * std::cos(th), a_im = g * std::sin(th);
        w_re += a_re * e.psi_re - a_im * e.psi_im;
        w_im += a_re * e.psi_im + a_im * e.psi_re;
    }
    (void) w_im;
    return w_re;
}

// ---- Shape base: src/librender/shape.cpp:323-354 ------------------------------------------------
static DirectionSample shape_sample_direction(const Scene &sc, int shape, V3 it_p, float it_time, V2 sample) {
    const RectData &r = sc.rects[sc.shapes[shape].index];
    PositionSample ps = rect_sample_position(r, it_time, sample);
    DirectionSample ds; static_cast<PositionSample &>(ds) = ps;
    ds.d = ds.p - it_p;
    float dist_squared = squared_norm(ds.d);
    ds.dist = std::sqrt(dist_squared);
    ds.d = ds.d / ds.dist;
    float dp = abs_dot(ds.d, ds.n);
    ds.pdf *= (dp != 0.f) ? dist_squared / dp : 0.f;
    ds.object = shape;
    return ds;
}
static float shape_pdf_direction(const Scene &sc, int shape, const DirectionSample &ds) {
    const RectData &r = sc.rects[sc.shapes[shape].index];
    float pdf = r.inv_area, dp = abs_dot(ds.d, ds.n);
    pdf *= (dp != 0.f) ? (ds.dist * ds.dist) / dp : 0.f;
    return pdf;
}

// ---- scene queries: src/librender/scene.cpp:129-178, scene_native.inl / scene_embree.inl ---------
struct Counters { uint32_t n_closest = 0, n_shadow = 0; };

static void bvh_build(Scene &sc);
static bool slab(const BvhNode &n, const Ray &ray, const float *idir, float tmax) {
    float t0 = ray.mint, t1 = tmax;
    const float o[3] = {ray.o.x, ray.o.y, ray.o.z};
    for (int a = 0; a < 3; ++a) {
        float tn = (n.lo[a] - o[a]) * idir[a], tf = (n.hi[a] - o[a]) * idir[a];
        if (tn > tf) std::swap(tn, tf);
        // NaN-safe, conservative (boxes are padded at build time)
        t0 = tn > t0 ? tn : t0; t1 = tf < t1 ? tf : t1;
    }
    return t0 <= t1 * 1.0000004f + 1e-30f;
}
static inline bool better(float t, uint32_t shape, uint32_t prim, const PI &best) {
    // strict "<" in scene/prim order == brute-force loop order (lowest (shape, prim) wins ties)
    if (t < best.t) return true;
    if (t == best.t && best.shape >= 0 && (shape < (uint32_t) best.shape || (shape == (uint32_t) best.shape && prim < best.prim))) return true;
    return false;
}
static PI scene_intersect_preliminary(const Scene &sc, const Ray &ray, Counters *cnt, int accel_override = -1) {
    if (cnt) cnt->n_closest++;
    PI best;
    int accel = accel_override >= 0 ? accel_override : sc.accel;
    // analytic rectangles are always tested directly
    for (size_t s = 0; s < sc.shapes.size(); ++s) {
        if (sc.shapes[s].type != 1) continue;
        float t, lx, ly;
        if (rect_intersect(sc.rects[sc.shapes[s].index], ray, t, lx, ly) && better(t, (uint32_t) s, 0, best)) {
            best.t = t; best.uv = V2(lx, ly); best.prim = 0; best.shape = (int) s;
        }
    }
    if (accel == 0) {
        for (size_t s = 0; s < sc.shapes.size(); ++s) {
            if (sc.shapes[s].type != 0) continue;
            const MeshData &m = sc.meshes[sc.shapes[s].index];
            for (uint32_t f = 0; f < m.nf; ++f) {
                float t, u, v;
                if (ray_intersect_triangle(m, f, ray, t, u, v) && better(t, (uint32_t) s, f, best)) {
                    best.t = t; best.uv = V2(u, v); best.prim = f; best.shape = (int) s;
                }
            }
        }
    } else if (accel == 1) {
        if (sc.nodes.empty()) return best;
        float idir[3] = {1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z};
        uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
        while (sp) {
            const BvhNode &n = sc.nodes[stack[--sp]];
            if (!slab(n, ray, idir, std::min(ray.maxt, best.t))) continue;
            if (n.count) {
                for (uint32_t i = 0; i < n.count; ++i) {
                    const TriRef &tr = sc.tris[n.left + i];
                    const MeshData &m = sc.meshes[sc.shapes[tr.shape].index];
                    float t, u, v;
                    if (ray_intersect_triangle(m, tr.prim, ray, t, u, v) && better(t, tr.shape, tr.prim, best)) {
                        best.t = t; best.uv = V2(u, v); best.prim = tr.prim; best.shape = (int) tr.shape;
                    }
                }
            } else { stack[sp++] = n.left; stack[sp++] = n.left + 1; }
        }
    } else {
        // scene_embree.inl:54-67,147-263: rtcIntersect1 with tnear=mint, tfar=maxt, mask=0, flags=0
        RTCIntersectContext_ ctx; ctx.flags = 0; ctx.filter = nullptr; ctx.instID[0] = 0xffffffffu;
        RTCRayHit_ rh;
        rh.ray.org_x = ray.o.x; rh.ray.org_y = ray.o.y; rh.ray.org_z = ray.o.z; rh.ray.tnear = ray.mint;
        rh.ray.dir_x = ray.d.x; rh.ray.dir_y = ray.d.y; rh.ray.dir_z = ray.d.z; rh.ray.time = 0;
        rh.ray.tfar = std::min(ray.maxt, best.t); rh.ray.mask = 0; rh.ray.id = 0; rh.ray.flags = 0;
        rh.hit.geomID = 0xffffffffu; rh.hit.primID = 0xffffffffu; rh.hit.instID[0] = 0xffffffffu;
        sc.em.rtcIntersect1(sc.em_scene, &ctx, &rh);
        if (rh.hit.geomID != 0xffffffffu && rh.ray.tfar < best.t) {
            best.t = rh.ray.tfar; best.uv = V2(rh.hit.u, rh.hit.v); best.prim = rh.hit.primID;
            best.shape = sc.em_geom_shape[rh.hit.geomID];
        }
    }
    return best;
}
// interaction.h:613-644 + Scene::ray_intersect (scene.cpp:129-137)
static SI compute_si(const Scene &sc, const Ray &ray, const PI &pi) {
    SI si;
    if (pi.t == Infinity) {
        // invalid: only t = inf and wi = -d are observable downstream
        si.t = Infinity; si.time = ray.time; si.wl = ray.wl; si.wi = -ray.d;
        return si;
    }
    const ShapeRec &sh = sc.shapes[pi.shape];
    si = sh.type == 0 ? mesh_compute_si(sc.meshes[sh.index], ray, pi) : rect_compute_si(sc.rects[sh.index], ray, pi);
    si.prim = pi.prim; si.shape = pi.shape; si.time = ray.time; si.wl = ray.wl;
    // initialize_sh_frame (interaction.h:159-162)
    float dn = dot(si.sh.n, si.dp_du);
    si.sh.s = normalize(V3(fnmadd(si.sh.n.x, dn, si.dp_du.x), fnmadd(si.sh.n.y, dn, si.dp_du.y), fnmadd(si.sh.n.z, dn, si.dp_du.z)));
    si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = si.to_local(-ray.d);
    return si;
}
static SI scene_ray_intersect(const Scene &sc, const Ray &ray, Counters *cnt, PI *pi_out = nullptr) {
    PI pi = scene_intersect_preliminary(sc, ray, cnt);
    if (pi_out) *pi_out = pi;
    return compute_si(sc, ray, pi);
}
static bool scene_ray_test(const Scene &sc, const Ray &ray, Counters *cnt, int accel_override = -1) {
    if (cnt) cnt->n_shadow++;
    int accel = accel_override >= 0 ? accel_override : sc.accel;
    for (size_t s = 0; s < sc.shapes.size(); ++s) {
        if (sc.shapes[s].type != 1) continue;
        float t, lx, ly;
        if (rect_intersect(sc.rects[sc.shapes[s].index], ray, t, lx, ly)) return true;
    }
    if (accel == 0) {
        for (size_t s = 0; s < sc.shapes.size(); ++s) {
            if (sc.shapes[s].type != 0) continue;
            const MeshData &m = sc.meshes[sc.shapes[s].index];
            for (uint32_t f = 0; f < m.nf; ++f) { float t, u, v; if (ray_intersect_triangle(m, f, ray, t, u, v)) return true; }
        }
        return false;
    } else if (accel == 1) {
        if (sc.nodes.empty()) return false;
        float idir[3] = {1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z};
        uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
        while (sp) {
            const BvhNode &n = sc.nodes[stack[--sp]];
            if (!slab(n, ray, idir, ray.maxt)) continue;
            if (n.count) {
                for (uint32_t i = 0; i < n.count; ++i) {
                    const TriRef &tr = sc.tris[n.left + i];
                    float t, u, v;
                    if (ray_intersect_triangle(sc.meshes[sc.shapes[tr.shape].index], tr.prim, ray, t, u, v)) return true;
                }
            } else { stack[sp++] = n.left; stack[sp++] = n.left + 1; }
        }
        return false;
    } else {
        // scene_embree.inl:272-286
        RTCIntersectContext_ ctx; ctx.flags = 0; ctx.filter = nullptr; ctx.instID[0] = 0xffffffffu;
        RTCRay_ r;
        r.org_x = ray.o.x; r.org_y = ray.o.y; r.org_z = ray.o.z; r.tnear = ray.mint;
        r.dir_x = ray.d.x; r.dir_y = ray.d.y; r.dir_z = ray.d.z; r.time = 0; r.tfar = ray.maxt;
        r.mask = 0; r.id = 0; r.flags = 0;
        sc.em.rtcOccluded1(sc.em_scene, &ctx, &r);
        return r.tfar != ray.maxt;
    }
}

// ---- own BVH build: median split on the largest centroid axis, padded boxes ---------------------
static void tri_bounds(const Scene &sc, const TriRef &tr, float *lo, float *hi) {
    const MeshData &m = sc.meshes[sc.shapes[tr.shape].index];
    for (int a = 0; a < 3; ++a) { lo[a] = Infinity; hi[a] = -Infinity; }
    for (int k = 0; k < 3; ++k) {
        uint32_t vi = m.idx[3 * tr.prim + k];
        for (int a = 0; a < 3; ++a) { float x = m.pos[3 * vi + a]; lo[a] = std::min(lo[a], x); hi[a] = std::max(hi[a], x); }
    }
}
static void bvh_build_rec(Scene &sc, uint32_t node, uint32_t first, uint32_t count, std::vector<float> &cent) {
    BvhNode &n0 = sc.nodes[node];
    for (int a = 0; a < 3; ++a) { n0.lo[a] = Infinity; n0.hi[a] = -Infinity; }
    float clo[3] = {Infinity, Infinity, Infinity}, chi[3] = {-Infinity, -Infinity, -Infinity};
    for (uint32_t i = first; i < first + count; ++i) {
        float lo[3], hi[3]; tri_bounds(sc, sc.tris[i], lo, hi);
        for (int a = 0; a < 3; ++a) {
            n0.lo[a] = std::min(n0.lo[a], lo[a]); n0.hi[a] = std::max(n0.hi[a], hi[a]);
            float c = 0.5f * (lo[a] + hi[a]); clo[a] = std::min(clo[a], c); chi[a] = std::max(chi[a], c);
        }
    }
    for (int a = 0; a < 3; ++a) {
        float pad = 1e-5f * std::max(std::max(std::fabs(n0.lo[a]), std::fabs(n0.hi[a])), 1e-3f) + 1e-7f;
        n0.lo[a] -= pad; n0.hi[a] += pad;
    }
    if (count <= 4) { n0.left = first; n0.count = count; return; }
    int axis = 0; float ext = chi[0] - clo[0];
    for (int a = 1; a < 3; ++a) if (chi[a] - clo[a] > ext) { ext = chi[a] - clo[a]; axis = a; }
    uint32_t mid = first + count / 2;
    if (ext > 0) {
        std::nth_element(sc.tris.begin() + first, sc.tris.begin() + mid, sc.tris.begin() + first + count,
            [&](const TriRef &a, const TriRef &b) {
                float la[3], ha[3], lb[3], hb[3]; tri_bounds(sc, a, la, ha); tri_bounds(sc, b, lb, hb);
                return la[axis] + ha[axis] < lb[axis] + hb[axis];
            });
    }
    uint32_t left = (uint32_t) sc.nodes.size();
    sc.nodes.push_back(BvhNode()); sc.nodes.push_back(BvhNode());
    sc.nodes[node].left = left; sc.nodes[node].count = 0;
    bvh_build_rec(sc, left, first, mid - first, cent);
    bvh_build_rec(sc, left + 1, mid, first + count - mid, cent);
}
static void bvh_build(Scene &sc) {
    sc.nodes.clear(); sc.tris.clear();
    for (size_t s = 0; s < sc.shapes.size(); ++s)
        if (sc.shapes[s].type == 0)
            for (uint32_t f = 0; f < sc.meshes[sc.shapes[s].index].nf; ++f) sc.tris.push_back({(uint32_t) s, f});
    if (sc.tris.empty()) return;
    sc.nodes.reserve(sc.tris.size());
    sc.nodes.push_back(BvhNode());
    std::vector<float> cent;
    bvh_build_rec(sc, 0, 0, (uint32_t) sc.tris.size(), cent);
}

static bool embree_attach(Scene &sc, const char *path) {
    EmbreeApi &e = sc.em;
    e.lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!e.lib) { sc.err = std::string("dlopen failed: ") + dlerror(); return false; }
#define LD(name) *(void **) (&e.name) = dlsym(e.lib, #name); if (!e.name) { sc.err = "missing symbol " #name; return false; }
    LD(rtcNewDevice) LD(rtcNewScene) LD(rtcSetSceneFlags) LD(rtcNewGeometry) LD(rtcSetSharedGeometryBuffer)
    LD(rtcCommitGeometry) LD(rtcAttachGeometry) LD(rtcReleaseGeometry) LD(rtcCommitScene) LD(rtcIntersect1)
    LD(rtcOccluded1) LD(rtcReleaseScene) LD(rtcReleaseDevice)
#undef LD
    sc.em_dev = e.rtcNewDevice("threads=1");                   // scene_embree.inl:24-39
    sc.em_scene = e.rtcNewScene(sc.em_dev);
    e.rtcSetSceneFlags(sc.em_scene, 1 /*RTC_SCENE_FLAG_DYNAMIC*/);
    sc.em_geom_shape.clear();
    for (size_t s = 0; s < sc.shapes.size(); ++s) {
        if (sc.shapes[s].type != 0) continue;
        MeshData &m = sc.meshes[sc.shapes[s].index];
        // mesh.cpp:789-805: shared FLOAT3 vertex buffer (stride 12) and UINT3 index buffer.  Embree
        // may read 16 bytes at the last vertex: pad the host arrays.
        m.pos.reserve(m.pos.size() + 4);
        RTCGeometry g = e.rtcNewGeometry(sc.em_dev, 0 /*RTC_GEOMETRY_TYPE_TRIANGLE*/);
        e.rtcSetSharedGeometryBuffer(g, 1 /*VERTEX*/, 0, 0x9003 /*FLOAT3*/, m.pos.data(), 0, 12, m.nv);
        e.rtcSetSharedGeometryBuffer(g, 0 /*INDEX*/, 0, 0x5003 /*UINT3*/, m.idx.data(), 0, 12, m.nf);
        e.rtcCommitGeometry(g);
        unsigned id = e.rtcAttachGeometry(sc.em_scene, g);
        e.rtcReleaseGeometry(g);
        if (sc.em_geom_shape.size() <= id) sc.em_geom_shape.resize(id + 1, -1);
        sc.em_geom_shape[id] = (int) s;
    }
    e.rtcCommitScene(sc.em_scene);
    return true;
}

// ---- BSDFs ---------------------------------------------------------------------------------------
struct BSDFSample { V3 wo; float pdf = 0, eta = 1; bool delta = false; };
// include/mitsuba/render/microfacet.h (line references in the function bodies)
struct Microfacet {
    int type; float au, av; bool sample_visible;
    Microfacet(int type_, float au_, float av_, bool sv) : type(type_), au(std::max(au_, 1e-4f)), av(std::max(av_, 1e-4f)), sample_visible(sv) {}
    float eval(V3 m) const {                                               // :179-196
        float alpha_uv = au * av, cos_theta = m.z, cos_theta_2 = sqr(cos_theta), result;
        if (type == BF_MICROFACET_BECKMANN)
            result = std::exp(-(sqr(m.x / au) + sqr(m.y / av)) / cos_theta_2) / (Pi * alpha_uv * sqr(cos_theta_2));
        else
            result = rcp(Pi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * cos_theta > 1e-20f ? result : 0.f;
    }
    float smith_g1(V3 v, V3 m) const {                                     // :330-358
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == BF_MICROFACET_BECKMANN) {
            float a = rsqrt(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = 2.f / (1.f + std::sqrt(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    float pdf(V3 wi, V3 m) const {                                         // :206-215
        float result = eval(m);
        if (sample_visible) result *= smith_g1(wi, m) * abs_dot(wi, m) / wi.z;
        else result *= m.z;
        return result;
    }
    V2 sample_visible_11(float cos_theta_i, V2 sample) const {             // :361-409
        if (type == BF_MICROFACET_BECKMANN) {
            float tan_theta_i = safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = std::erf(cot_theta_i);
            sample.x = std::max(std::min(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = std::max(std::min(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * std::erf(std::sqrt(-std::log(sample.x)));
            sample.x *= 1.f + maxval + InvSqrtPi * tan_theta_i * std::exp(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv(x), value = 1.f + x + InvSqrtPi * tan_theta_i * std::exp(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return V2(erfinv(x), erfinv(fmsub(2.f, sample.y, 1.f)));
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
    void sample(V3 wi, V2 sample, V3 &m_out, float &pdf_out) const {       // :217-322
        if (!sample_visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2, pdf;
            if (au == av) {
                float ang = (2.f * Pi) * sample.y; sin_phi = std::sin(ang); cos_phi = std::cos(ang);
                alpha_2 = au * au;
            } else {
                float ratio = av / au, tmp = ratio * std::tan((2.f * Pi) * sample.y);
                cos_phi = rsqrt(fmadd(tmp, tmp, 1.f));
                cos_phi = mulsign(cos_phi, std::fabs(sample.y - .5f) - .25f);
                sin_phi = cos_phi * tmp;
                alpha_2 = rcp(sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == BF_MICROFACET_BECKMANN) {
                cos_theta = rsqrt(fnmadd(alpha_2, std::log(1.f - sample.x), 1.f));
                cos_theta_2 = sqr(cos_theta);
                float cos_theta_3 = std::max(cos_theta_2 * cos_theta, 1e-20f);
                pdf = (1.f - sample.x) / (Pi * au * av * cos_theta_3);
            } else {
                float tan_theta_m_2 = alpha_2 * sample.x / (1.f - sample.x);
                cos_theta = rsqrt(1.f + tan_theta_m_2);
                cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, cos_theta_3 = std::max(cos_theta_2 * cos_theta, 1e-20f);
                pdf = rcp(Pi * au * av * cos_theta_3 * sqr(temp));
            }
            float sin_theta = std::sqrt(1.f - cos_theta_2);
            m_out = V3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta); pdf_out = pdf;
        } else {
            V3 wi_p = normalize(V3(au * wi.x, av * wi.y, wi.z));
            // Frame::sincos_phi (frame.h:107-118)
            float sin_theta_2 = fnmadd(wi_p.z, wi_p.z, 1.f), inv_sin_theta = rsqrt(sin_theta_2);
            float rx = wi_p.x * inv_sin_theta, ry = wi_p.y * inv_sin_theta;
            float cos_phi, sin_phi;
            if (std::fabs(sin_theta_2) <= 4.f * Epsilon) { cos_phi = 1.f; sin_phi = 0.f; }
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
static float fresnel_conductor(float cos_theta_i, float eta_r, float eta_i) {
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
static inline V3 reflect(V3 wi, V3 m) {  // fresnel.h:278-280: fmsub(m, 2 dot(wi, m), wi)
    float k = 2.f * dot(wi, m);
    return V3(fmsub(m.x, k, wi.x), fmsub(m.y, k, wi.y), fmsub(m.z, k, wi.z));
}

static bool bsdf_is_smooth(const Scene &sc, uint32_t id) {  // has_flag(flags, BSDFFlags::Smooth)
    uint32_t k = sc.bsdfs[id].kind;
    return k == BF_BSDF_DIFFUSE || k == BF_BSDF_ROUGHCONDUCTOR;
}
// one-sided kernels ------------------------------------------------------------------------------
static Spec onesided_eval(const Scene &sc, const bf_bsdf_desc &b, const SI &si, V3 wi, V3 wo) {
    Spec r;
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (b.kind == BF_BSDF_DIFFUSE) {                               // diffuse.cpp:109-124
        bool active = cos_theta_i > 0.f && cos_theta_o > 0.f;
        if (!active) return r;
        Spec refl = spectrum_eval(sc, b.reflectance, si.wl);
        for (int k = 0; k < sc.S; ++k) r.v[k] = refl.v[k] * InvPi * cos_theta_o;
        return r;
    } else if (b.kind == BF_BSDF_ROUGHCONDUCTOR) {                 // roughconductor.cpp:277-344
        bool active = cos_theta_i > 0.f && cos_theta_o > 0.f;
        if (!active) return r;
        V3 H = normalize(wo + wi);
        Microfacet distr((int) b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        float D = distr.eval(H);
        active &= D != 0.f;
        float G = distr.G(wi, wo, H);
        float result = D * G / (4.f * wi.z);
        Spec eta = spectrum_eval(sc, b.eta, si.wl), kk = spectrum_eval(sc, b.k, si.wl);
        float cosh = dot(wi, H);
        Spec sr(1.f);
        if (b.specular_reflectance != BF_NONE) sr = spectrum_eval(sc, b.specular_reflectance, si.wl);
        for (int k = 0; k < sc.S; ++k) {
            float F = fresnel_conductor(cosh, eta.v[k], kk.v[k]);
            float res = result;
            if (b.specular_reflectance != BF_NONE) res *= sr.v[k];
            r.v[k] = active ? F * res : 0.f;
        }
        return r;
    }
    return r;  // smooth conductor / null: eval = 0 (delta lobes)
}
static float onesided_pdf(const Scene &, const bf_bsdf_desc &b, const SI &, V3 wi, V3 wo) {
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (b.kind == BF_BSDF_DIFFUSE) {                               // diffuse.cpp:126-135
        float pdf = square_to_cosine_hemisphere_pdf(wo);
        return (cos_theta_i > 0.f && cos_theta_o > 0.f) ? pdf : 0.f;
    } else if (b.kind == BF_BSDF_ROUGHCONDUCTOR) {                 // roughconductor.cpp:346-381
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
static Spec onesided_sample(const Scene &sc, const bf_bsdf_desc &b, const SI &si, V3 wi, float, V2 sample2, BSDFSample &bs) {
    Spec r;
    bs = BSDFSample(); bs.wo = V3(0, 0, 0); bs.pdf = 0.f; bs.eta = 0.f;  // zero<BSDFSample3f>()
    float cos_theta_i = wi.z;
    if (b.kind == BF_BSDF_DIFFUSE) {                               // diffuse.cpp:78-107
        if (!(cos_theta_i > 0.f)) return r;
        bs.wo = square_to_cosine_hemisphere(sample2);
        bs.pdf = square_to_cosine_hemisphere_pdf(bs.wo);
        bs.eta = 1.f;
        Spec value = spectrum_eval(sc, b.reflectance, si.wl);
        if (bs.pdf > 0.f) r = value;
        return r;
    } else if (b.kind == BF_BSDF_ROUGHCONDUCTOR) {                 // roughconductor.cpp:196-275
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
        Spec eta = spectrum_eval(sc, b.eta, si.wl), kk = spectrum_eval(sc, b.k, si.wl);
        Spec sr(1.f);
        if (b.specular_reflectance != BF_NONE) sr = spectrum_eval(sc, b.specular_reflectance, si.wl);
        float cosm = dot(wi, m);
        for (int k = 0; k < sc.S; ++k) {
            float F = fresnel_conductor(cosm, eta.v[k], kk.v[k]);
            float w = weight;
            if (b.specular_reflectance != BF_NONE) w *= sr.v[k];
            r.v[k] = active ? F * w : 0.f;
        }
        return r;
    } else if (b.kind == BF_BSDF_CONDUCTOR) {                      // conductor.cpp sample()
        if (!(cos_theta_i > 0.f)) return r;
        bs.wo = V3(-wi.x, -wi.y, wi.z); bs.eta = 1.f; bs.pdf = 1.f; bs.delta = true;
        Spec eta = spectrum_eval(sc, b.eta, si.wl), kk = spectrum_eval(sc, b.k, si.wl);
        Spec sr(1.f);
        if (b.specular_reflectance != BF_NONE) sr = spectrum_eval(sc, b.specular_reflectance, si.wl);
        for (int k = 0; k < sc.S; ++k) {
            float v = (b.specular_reflectance != BF_NONE) ? sr.v[k] : 1.f;
            r.v[k] = v * fresnel_conductor(cos_theta_i, eta.v[k], kk.v[k]);
        }
        return r;
    }
    return r;
}
// twosided.cpp:94-181 (front/back select; strict inequalities: cos_theta == 0 -> zero result)
static Spec bsdf_eval(const Scene &sc, uint32_t id, const SI &si, V3 wo) {
    const bf_bsdf_desc &b = sc.bsdfs[id];
    if (!b.two_sided) return onesided_eval(sc, b, si, si.wi, wo);
    Spec r;
    if (si.wi.z > 0.f) r = onesided_eval(sc, b, si, si.wi, wo);
    if (si.wi.z < 0.f) {
        const bf_bsdf_desc &bb = b.back_bsdf != BF_NONE ? sc.bsdfs[b.back_bsdf] : b;
        r = onesided_eval(sc, bb, si, V3(si.wi.x, si.wi.y, si.wi.z * -1.f), V3(wo.x, wo.y, wo.z * -1.f));
    }
    return r;
}
static float bsdf_pdf(const Scene &sc, uint32_t id, const SI &si, V3 wo) {
    const bf_bsdf_desc &b = sc.bsdfs[id];
    if (!b.two_sided) return onesided_pdf(sc, b, si, si.wi, wo);
    float r = 0.f;
    if (si.wi.z > 0.f) r = onesided_pdf(sc, b, si, si.wi, wo);
    if (si.wi.z < 0.f) {
        const bf_bsdf_desc &bb = b.back_bsdf != BF_NONE ? sc.bsdfs[b.back_bsdf] : b;
        r = onesided_pdf(sc, bb, si, V3(si.wi.x, si.wi.y, si.wi.z * -1.f), V3(wo.x, wo.y, wo.z * -1.f));
    }
    return r;
}
static Spec bsdf_sample(const Scene &sc, uint32_t id, const SI &si, float s1, V2 s2, BSDFSample &bs) {
    const bf_bsdf_desc &b = sc.bsdfs[id];
    if (!b.two_sided) return onesided_sample(sc, b, si, si.wi, s1, s2, bs);
    Spec r; bs = BSDFSample(); bs.eta = 0.f;
    if (si.wi.z > 0.f) r = onesided_sample(sc, b, si, si.wi, s1, s2, bs);
    if (si.wi.z < 0.f) {
        const bf_bsdf_desc &bb = b.back_bsdf != BF_NONE ? sc.bsdfs[b.back_bsdf] : b;
        r = onesided_sample(sc, bb, si, V3(si.wi.x, si.wi.y, si.wi.z * -1.f), s1, s2, bs);
        bs.wo.z *= -1.f;
    }
    return r;
}

// ---- signal model shared by wigner / phased endpoints -------------------------------------------
// wignertransmitter.cpp:111-146 (== phasedtransmitter.cpp:176-210, wignerreceiver.cpp:116-146).
// Only frequency[0] is used; the value is broadcast over the S channels.
static float eval_signal(const bf_signal_desc &sg, float time, float frequency0) {
    if (sg.signaltype == BF_SIG_LINFMCW) {
        float t = fmodulo(time, rcp(sg.repfreq));
        float ti = 0 + sg.t_ext / 2;
        float fi = sg.f_centre + (sg.f_ext / sg.t_ext) * (t - ti);
        return rect((t - ti) / sg.t_ext) > 0.f ? wchirp(t - ti, frequency0 - fi, sg.t_ext, sg.amplitude) : 0.f;
    } else if (sg.signaltype == BF_SIG_PULSE) {
        float t = fmodulo(time, rcp(sg.repfreq));
        float ti = 0 + sg.t_ext / 2;
        float fi = sg.f_centre;
        return rect((t - ti) / sg.t_ext) > 0.f ? wchirp(t - ti, frequency0 - fi, sg.t_ext, sg.amplitude) : 0.f;
    }
    return sg.amplitude * sg.amplitude;
}
// wignertransmitter.cpp:152-168.  "pulse"/other leave `frequencies` uninitialised in the reference;
// the oracle defines that case as f_centre.
static float sample_delta_frequency(const bf_signal_desc &sg, float time) {
    if (sg.signaltype == BF_SIG_LINFMCW) {
        float t = fmodulo(time, rcp(sg.repfreq));
        float ti = 0 + sg.t_ext / 2;
        return sg.f_centre + (sg.f_ext / sg.t_ext) * (t - ti);
    }
    return sg.f_centre;
}
static inline Spec sample_shifted(float sample, int S) {                   // math.h:494-517
    Spec r;
    for (int k = 0; k < S; ++k) { float v = sample + (float) k / (float) S; if (v > 1.f) v -= 1.f; r.v[k] = v; }
    return r;
}
// transmitter sample_frequency (wignertransmitter.cpp:174-188): returns frequencies + signal power
static void tx_sample_frequency(const Scene &sc, const bf_signal_desc &sg, float time, float u, Spec &freq, Spec &power) {
    if (sg.sig_is_delta) {
        float f = sample_delta_frequency(sg, time);
        freq = Spec(f); power = Spec(1.f);
    } else {
        Spec fs = sample_shifted(u, sc.S);
        for (int k = 0; k < sc.S; ++k) freq.v[k] = fs.v[k] * sg.f_ext + (sg.f_centre - sg.f_ext / 2);
        power = Spec(eval_signal(sg, time, freq.v[0]));
    }
}

// ---- transmitters ----------------------------------------------------------------------------------
// Returns the spectrum; mutates si.wl / si.phase exactly where the reference const_casts.
static Spec tx_eval(const Scene &sc, const TxRec &tx, SI &si) {
    const bf_endpoint_desc &d = tx.d;
    Spec r;
    if (d.kind == BF_TX_AREA) {                                    // areatransmitter.cpp:65-74
        if (!(si.wi.z > 0.f)) return r;
        Spec rad = spectrum_eval(sc, d.texture, si.wl);
        float area = sc.rects[sc.shapes[tx.shape].index].area;
        for (int k = 0; k < sc.S; ++k) r.v[k] = rad.v[k] * area;
        return r;
    } else if (d.kind == BF_TX_AREALIGHT) {                        // src/emitters/area.cpp:63-72
        if (!(si.wi.z > 0.f)) return r;
        return spectrum_eval(sc, d.texture, si.wl);
    } else if (d.kind == BF_TX_WIGNER) {                           // wignertransmitter.cpp:193-271
        float signal_power;
        if (d.signal.resample_freq) {
            float f = sample_delta_frequency(d.signal, si.time);
            signal_power = 1.f;
            for (int k = 0; k < sc.S; ++k) si.wl.v[k] = sc.C.c * rcp(f) * 1e9f;
            si.phase = 0.f;
        } else {
            signal_power = eval_signal(d.signal, si.time, freq_from_wl(sc.C.c, si.wl.v[0]));
            si.phase = 0.f;
        }
        Spec geom = spectrum_eval(sc, d.texture, si.wl);
        // `DirectionSample3f ds(si); ds.d *= -1` reads an uninitialised ds.d in the reference;
        // the oracle defines it as the zero vector (sinc terms = 1).
        float w = rect_sample_wigner(sc.rects[sc.shapes[tx.shape].index], si.p, V3(0, 0, 0), si.wl.v[0]);
        if (!(si.wi.z > 0.f)) return r;
        for (int k = 0; k < sc.S; ++k) r.v[k] = signal_power * d.signal.gain * (geom.v[k] * w) * TwoPi;
        return r;
    }
    else if (d.kind == BF_TX_PHASED) {                           // phasedtransmitter.cpp:296-381
        float signal_power;
        if (d.signal.resample_freq) {
            float f = sample_delta_frequency(d.signal, si.time);
            signal_power = 1.f;
            for (int k = 0; k < sc.S; ++k) si.wl.v[k] = sc.C.c * rcp(f) * 1e9f;
            si.phase = 0.f;
        } else {
            signal_power = eval_signal(d.signal, si.time, freq_from_wl(sc.C.c, si.wl.v[0]));
            si.phase = 0.f;
        }
        Spec geom = spectrum_eval(sc, d.texture, si.wl);
        float inv_area = rcp(sc.rects[sc.shapes[tx.shape].index].area);
        // ds.d uninitialised in the reference as for the wigner transmitter: defined as the zero vector
        float w = phased_sample_wigner(tx.velems, d.array.elem_dims, si.p, V3(0, 0, 0), si.wl.v[0]);
        if (!(si.wi.z > 0.f)) return r;
        for (int k = 0; k < sc.S; ++k) r.v[k] = signal_power * d.signal.gain * (geom.v[k] * inv_area * w);
        return r;
    }
    return r;  // spot: eval = 0 (spot.cpp:166)
}
struct TxDirResult { DirectionSample ds; Spec spec; };
// `it` is the interaction at the scattering point; wl / phase are mutated like the reference does.
static TxDirResult tx_sample_direction(const Scene &sc, const TxRec &tx, SI &it, V2 sample) {
    const bf_endpoint_desc &d = tx.d;
    TxDirResult out;
    DirectionSample &ds = out.ds;
    if (d.kind == BF_TX_SPOT) {                                    // spot.cpp:141-164
        ds.p = V3(tx.to_world.m[0][3], tx.to_world.m[1][3], tx.to_world.m[2][3]);
        ds.n = V3(0, 0, 0); ds.uv = V2(0, 0); ds.pdf = 1.f; ds.time = it.time; ds.delta = true;
        ds.d = ds.p - it.p; ds.dist = norm(ds.d);
        float inv_dist = rcp(ds.dist);
        ds.d = ds.d * inv_dist;
        V3 local_d = xf_vector(tx.to_object, -ds.d);
        // falloff_curve (spot.cpp:97-115)
        Spec result = spectrum_eval(sc, d.texture, it.wl);
        V3 ld = normalize(local_d);
        float cos_theta = ld.z;
        for (int k = 0; k < sc.S; ++k) {
            float beam = cos_theta >= tx.cos_beam ? result.v[k] : result.v[k] * ((tx.cutoff - std::acos(cos_theta)) * tx.inv_transition);
            float v = cos_theta <= tx.cos_cutoff ? 0.f : beam;
            out.spec.v[k] = v * (inv_dist * inv_dist);
        }
        return out;
    }
    // area-like: Shape::sample_direction, then facing test
    ds = shape_sample_direction(sc, tx.shape, it.p, it.time, sample);
    bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
    if (d.kind == BF_TX_AREA || d.kind == BF_TX_AREALIGHT) {       // areatransmitter.cpp:131-182, area.cpp
        Spec rad = spectrum_eval(sc, d.texture, it.wl);
        for (int k = 0; k < sc.S; ++k) out.spec.v[k] = active ? rad.v[k] / ds.pdf : 0.f;
        return out;
    }
    // wignertransmitter.cpp:373-534
    Spec tex = spectrum_eval(sc, d.texture, it.wl);
    Spec geom; for (int k = 0; k < sc.S; ++k) geom.v[k] = tex.v[k] / ds.pdf;
    if ((double) ds.dist > 5e-7) ds.time += -ds.dist / sc.C.c;
    float signal_power;
    if (d.signal.resample_freq) {
        float f = sample_delta_frequency(d.signal, ds.time);
        signal_power = 1.f;
        for (int k = 0; k < sc.S; ++k) it.wl.v[k] = sc.C.c * rcp(f) * 1e9f;
        it.phase = 0.f;
    } else {
        signal_power = eval_signal(d.signal, ds.time, freq_from_wl(sc.C.c, it.wl.v[0]));
        it.phase = 0.f;
    }
    if (d.kind == BF_TX_PHASED) {                                  // phasedtransmitter.cpp:596-654
        float w = phased_sample_wigner(tx.velems, d.array.elem_dims, ds.p, ds.d * -1.f, it.wl.v[0]);
        for (int k = 0; k < sc.S; ++k) geom.v[k] *= w;
        ds.pdf *= w;
        ds.pdf = std::sqrt(ds.pdf * ds.pdf);
        for (int k = 0; k < sc.S; ++k) out.spec.v[k] = active ? signal_power * d.signal.gain * geom.v[k] * 1.f : 0.f;
        return out;
    }
    float w = rect_sample_wigner(sc.rects[sc.shapes[tx.shape].index], ds.p, ds.d * -1.f, it.wl.v[0]);
    for (int k = 0; k < sc.S; ++k) geom.v[k] *= w;
    ds.pdf *= w;
    float extents = rcp(sc.rects[sc.shapes[tx.shape].index].area) * TwoPi;
    for (int k = 0; k < sc.S; ++k) out.spec.v[k] = active ? signal_power * d.signal.gain * geom.v[k] * extents : 0.f;
    return out;
}
static float tx_pdf_direction(const Scene &sc, const TxRec &tx, const SI &it, const DirectionSample &ds) {
    const bf_endpoint_desc &d = tx.d;
    if (d.kind == BF_TX_SPOT) return 0.f;
    float dp = dot(ds.d, ds.n);
    bool active = dp < 0.f;
    float value = shape_pdf_direction(sc, tx.shape, ds);
    if (d.kind == BF_TX_WIGNER) {                                  // wignertransmitter.cpp:540-577
        float w = rect_sample_wigner(sc.rects[sc.shapes[tx.shape].index], ds.p, ds.d * -1.f, it.wl.v[0]);
        value *= w;
    } else if (d.kind == BF_TX_PHASED) {                           // phasedtransmitter.cpp:660-702
        value *= phased_sample_wigner(tx.velems, d.array.elem_dims, ds.p, ds.d * -1.f, it.wl.v[0]);
        value = std::sqrt(value * value);
    }
    return active ? value : 0.f;
}

// scene.cpp:249-299
static TxDirResult scene_sample_transmitter_direction(const Scene &sc, SI &ref, V2 sample, Counters *cnt) {
    TxDirResult r;
    if (sc.txs.empty()) { r.ds.pdf = 0.f; return r; }
    if (sc.txs.size() == 1) {
        r = tx_sample_direction(sc, sc.txs[0], ref, sample);
        r.ds.object = 0;
    } else {
        float n = (float) sc.txs.size();
        float transmitter_pdf = 1.f / n;
        uint32_t index = std::min((uint32_t) (sample.x * n), (uint32_t) sc.txs.size() - 1);
        sample.x = (sample.x - index * transmitter_pdf) * n;
        r = tx_sample_direction(sc, sc.txs[index], ref, sample);
        r.ds.object = (int) index;
        r.ds.pdf *= transmitter_pdf;
        for (int k = 0; k < sc.S; ++k) r.spec.v[k] *= rcp(transmitter_pdf);
    }
    bool active = r.ds.pdf != 0.f;
    if (active) {
        Ray ray; ray.o = ref.p; ray.d = r.ds.d;
        ray.mint = RayEpsilon * (1.f + hmax_abs(ref.p));
        ray.maxt = r.ds.dist * (1.f - ShadowEpsilon);
        ray.time = ref.time; ray.wl = ref.wl;
        if (scene_ray_test(sc, ray, cnt)) r.spec = Spec(0.f);
    }
    return r;
}
static float scene_pdf_transmitter_direction(const Scene &sc, const SI &ref, const DirectionSample &ds) {
    if (sc.txs.size() == 1) return tx_pdf_direction(sc, sc.txs[0], ref, ds);
    return tx_pdf_direction(sc, sc.txs[ds.object], ref, ds) * (1.f / (float) sc.txs.size());
}

// ---- receivers / sensors ----------------------------------------------------------------------------
struct Sampler {
    PCG32 rng;
    float next_1d() { return rng.next_float32(); }
    V2 next_2d() { float a = next_1d(), b = next_1d(); return V2(a, b); }
};
// spectrum.h:310-314, 360-371
static void sample_wavelength(const Scene &sc, float sample, Spec &wl, Spec &weight) {
    if (sc.S == 1) { wl = Spec(std::numeric_limits<float>::quiet_NaN()); weight = Spec(1.f); return; }
    Spec s = sample_shifted(sample, sc.S);
    float range = sc.C.cie_max_nm - sc.C.cie_min_nm;
    for (int k = 0; k < sc.S; ++k) { wl.v[k] = s.v[k] * range + sc.C.cie_min_nm; weight.v[k] = range; }
}
static bool rx_needs_aperture_sample(const Scene &sc) { return sc.rx.kind != BF_RX_PERSPECTIVE; }
static void rx_sample_ray(const Scene &sc, float time, float sample1, V2 sample2, V2 sample3, Ray &ray, Spec &weight) {
    const bf_endpoint_desc &d = sc.rx;
    if (d.kind == BF_RX_PERSPECTIVE) {                             // perspective.cpp:166-196
        Spec wl, ww; sample_wavelength(sc, sample1, wl, ww);
        ray.time = time; ray.wl = wl;
        M4 s2c = M4::from_rows16(d.sample_to_camera);
        V3 near_p = xf_point(s2c, V3(sample2.x, sample2.y, 0.f));
        V3 dd = normalize(near_p);
        float inv_z = rcp(dd.z);
        ray.mint = d.near_clip * inv_z; ray.maxt = d.far_clip * inv_z;
        ray.o = xf_point_affine(sc.rx_to_world, V3(0, 0, 0));
        ray.d = xf_vector(sc.rx_to_world, dd);
        weight = ww;
        return;
    }
    const RectData &r = sc.rects[sc.shapes[sc.rx_shape].index];
    if (d.kind == BF_RX_OMNIDIRECTIONAL || d.kind == BF_RX_FLUXMETER) {
        // omnidirectional.cpp:71-107 / fluxmeter.cpp:63-84
        PositionSample ps = rect_sample_position(r, time, sample2);
        V3 local = square_to_cosine_hemisphere(sample3);
        Spec wl, ww; sample_wavelength(sc, sample1, wl, ww);
        ray = Ray(); ray.o = ps.p; ray.d = Frame(ps.n).to_world(local); ray.time = time; ray.wl = wl;
        float k = d.kind == BF_RX_OMNIDIRECTIONAL ? r.area : Pi;
        for (int i = 0; i < sc.S; ++i) weight.v[i] = ww.v[i] * k;
        return;
    }
    // wignerreceiver.cpp:150-269
    Spec freq, power;
    const bf_signal_desc &sg = d.signal;
    if (d.receive_type == BF_RECV_RAW || d.receive_type == BF_RECV_RAW_RESAMPLE) {
        Spec fs = sample_shifted(sample1, sc.S);
        for (int k = 0; k < sc.S; ++k) freq.v[k] = fs.v[k] * sg.f_ext + (sg.f_centre - sg.f_ext / 2);
        power = Spec(1.f);
    } else {
        if (sg.sig_is_delta) { freq = Spec(sample_delta_frequency(sg, time)); power = Spec(1.f); }
        else {
            Spec fs = sample_shifted(sample1, sc.S);
            for (int k = 0; k < sc.S; ++k) freq.v[k] = fs.v[k] * sg.f_ext + (sg.f_centre - sg.f_ext / 2);
            power = Spec(eval_signal(sg, time, freq.v[0]));
        }
    }
    Spec wavelength; for (int k = 0; k < sc.S; ++k) wavelength.v[k] = sc.C.c * rcp(freq.v[k]) * 1e9f;
    PositionSample ps = rect_sample_position(r, time, sample2);
    Frame fr(ps.n);
    V3 dloc = square_to_cosine_hemisphere(sample3);
    float geom_gain;
    if (d.kind == BF_RX_PHASED) {                                  // phasedreceiver.cpp:313-375 (:353)
        float w = phased_sample_wigner(sc.rx_velems, d.array.elem_dims, ps.p, dloc, wavelength.v[0]);
        float dn = dot(dloc, ps.n);                                // local direction against the world normal, as written
        geom_gain = w * ps.pdf * (1 - dn * dn * dn * dn);
    } else {
        float w = rect_sample_wigner(r, ps.p, dloc, wavelength.v[0]);
        geom_gain = w * ps.pdf;
    }
    float extents0 = r.area * Pi;
    ray = Ray(); ray.o = ps.p; ray.d = fr.to_world(dloc); ray.time = time; ray.phase = 0.f; ray.wl = wavelength;
    for (int k = 0; k < sc.S; ++k) {
        float extents = extents0;
        if (!sg.sig_is_delta) extents *= sc.C.c * rcp(sg.f_ext) * 1e9f;
        weight.v[k] = power.v[k] * sg.gain * geom_gain * extents;
    }
}

// ---- integrators ----------------------------------------------------------------------------------
static inline float mis_weight(float pdf_a, float pdf_b) {             // pathtimefrequency.cpp:472-477
    pdf_a *= pdf_a; pdf_b *= pdf_b;
    return pdf_a > 0.f ? pdf_a / (pdf_a + pdf_b) : 0.f;
}
static inline float hmax_spec(const Spec &s, int S) { float m = s.v[0]; for (int k = 1; k < S; ++k) m = std::max(m, s.v[k]); return m; }
static inline float hsum_spec(const Spec &s, int S) {                  // enoki hsum of a 4-vector = (v0+v1)+(v2+v3)
    if (S == 1) return s.v[0];
    return (s.v[0] + s.v[1]) + (s.v[2] + s.v[3]);
}
static inline void ray_update_state(const Scene &sc, Ray &ray, float t) {  // ray.h:89-93
    ray.time += t / sc.C.c;
    // `(MAX-MIN)/2*1e-9`: the double literal promotes the quotient and the += to double
    double denom = (double) ((sc.C.lambda_max_nm - sc.C.lambda_min_nm) / 2) * 1e-9;
    ray.phase = (float) ((double) ray.phase + (double) (TwoPi * t) / denom);
}
static inline Ray spawn_ray(const SI &si, V3 d) {                       // interaction.h:61-64
    Ray r; r.o = si.p; r.d = d; r.mint = (1.f + hmax_abs(si.p)) * RayEpsilon; r.maxt = Infinity;
    r.time = si.time; r.wl = si.wl; r.phase = 0.f; return r;
}
struct PathOut {
    Spec result; bool valid = false; float scalar = 0.f;  // scalar: pathlength (gen 2) / pathtime (gen 1)
    PI first; uint32_t depth = 0;
};
// gen 3: pathtimefrequency.cpp:103-460.  `ray_` is mutated where the reference const_casts.
static PathOut path_time_frequency(const Scene &sc, Sampler &sampler, Ray &ray_, Counters *cnt) {
    PathOut out;
    Ray ray = ray_;
    float emission_weight = 1.f;
    Spec throughput(1.f), result(0.f);
    bool active = true;
    SI si = scene_ray_intersect(sc, ray, cnt, &out.first);
    bool valid_ray = si.is_valid();
    int transmitter = si.is_valid() ? sc.shapes[si.shape].tx : -1;
    if (si.is_valid()) { ray_update_state(sc, ray, -si.t); si.time = ray.time; }
    int depth = 1;
    for (;; ++depth) {
        if (transmitter >= 0) {
            Spec e = tx_eval(sc, sc.txs[transmitter], si);
            if (active) for (int k = 0; k < sc.S; ++k) result.v[k] += emission_weight * throughput.v[k] * e.v[k];
        }
        active &= si.is_valid();
        if (depth > sc.integ.rr_depth) {
            float q = std::min(hmax_spec(throughput, sc.S) * 1.f, .95f);
            active &= sampler.next_1d() < q;
            float iq = rcp(q);
            for (int k = 0; k < sc.S; ++k) throughput.v[k] *= iq;
        }
        if ((uint32_t) depth >= (uint32_t) sc.integ.max_depth || !active) break;
        uint32_t bsdf = sc.shapes[si.shape].bsdf;
        bool active_e = active && bsdf_is_smooth(sc, bsdf);
        if (active_e) {
            TxDirResult tr = scene_sample_transmitter_direction(sc, si, sampler.next_2d(), cnt);
            if (sc.integ.doppler) {
                // pathtimefrequency.cpp:252-256 (commented out in the reference; restored behind bf_integrator_desc::doppler):
                // wavelengths += si.shape->doppler(si), Shape::doppler = 2*dot(si.wi, m_velocity*Point3f(si.to_local(si.p)))/MTS_C*si.wavelengths (shape.cpp:375-389)
                M4 vel; memset(&vel, 0, sizeof(vel)); for (int i = 0; i < 4; ++i) vel.m[i][i] = 1.f;
                if (si.shape < sc.velocities.size()) vel = sc.velocities[si.shape];
                const V3 vp = xf_point(vel, si.to_local(si.p));
                const float f = 2 * dot(si.wi, vp) / sc.C.c;
                for (int k = 0; k < sc.S; ++k) si.wl.v[k] += f * si.wl.v[k];
            }
            active_e &= tr.ds.pdf != 0.f;
            V3 wo = si.to_local(tr.ds.d);
            Spec bsdf_val = active_e ? bsdf_eval(sc, bsdf, si, wo) : Spec(0.f);
            float bsdf_pdf_ = active_e ? bsdf_pdf(sc, bsdf, si, wo) : 0.f;
            float mis = tr.ds.delta ? 1.f : mis_weight(tr.ds.pdf, bsdf_pdf_);
            if (active_e) for (int k = 0; k < sc.S; ++k) result.v[k] += mis * throughput.v[k] * bsdf_val.v[k] * tr.spec.v[k];
        }
        float s1 = sampler.next_1d(); V2 s2 = sampler.next_2d();
        BSDFSample bs; Spec bsdf_val = bsdf_sample(sc, bsdf, si, s1, s2, bs);
        bool any_nz = false;
        for (int k = 0; k < sc.S; ++k) { throughput.v[k] = throughput.v[k] * bsdf_val.v[k]; any_nz |= throughput.v[k] != 0.f; }
        active &= any_nz;
        if (!active) break;
        ray = spawn_ray(si, si.to_world(bs.wo));
        SI si_bsdf = scene_ray_intersect(sc, ray, cnt);
        if (active) { ray_update_state(sc, ray, -si_bsdf.t); si_bsdf.time = ray.time; }
        transmitter = si_bsdf.is_valid() ? sc.shapes[si_bsdf.shape].tx : -1;
        if (transmitter >= 0) {
            // DirectionSample3f ds(si_bsdf, si): records.h:168-174
            DirectionSample ds; ds.p = si_bsdf.p; ds.n = si_bsdf.sh.n; ds.uv = si_bsdf.uv; ds.time = si_bsdf.time;
            ds.d = si_bsdf.p - si.p; ds.dist = norm(ds.d); ds.d = ds.d / ds.dist; ds.object = transmitter;
            float transmitter_pdf = !bs.delta ? scene_pdf_transmitter_direction(sc, si, ds) : 0.f;
            emission_weight = mis_weight(bs.pdf, transmitter_pdf);
        }
        si = si_bsdf;
    }
    if (valid_ray) { ray_.time = ray.time; ray_.wl = si.wl; ray_.phase += ray.phase; }
    out.result = result; out.valid = valid_ray; out.depth = (uint32_t) depth;
    return out;
}
// gen 2 / gen 1: pathlength.cpp:114-325 and pathtime.cpp:114-286 share the loop; they differ in the
// scalar accumulated (pathlength.cpp:146,161,209,307 vs pathtime.cpp:140,228).
static PathOut path_emitters(const Scene &sc, Sampler &sampler, const Ray &ray_, Counters *cnt, int gen) {
    PathOut out;
    Ray ray = ray_;
    float emission_weight = 1.f;
    Spec throughput(1.f), result(0.f);
    bool active = true;
    float scalar = 0.f;
    SI si = scene_ray_intersect(sc, ray, cnt, &out.first);
    bool valid_ray = si.is_valid();
    int emitter = si.is_valid() ? sc.shapes[si.shape].tx : -1;
    if (gen == BF_GEN_RANGE) scalar += valid_ray ? si.t : 0.f;
    else scalar = si.is_valid() ? si.t / 3.0e8f : 0.f;
    int depth = 1;
    for (;; ++depth) {
        if (emitter >= 0) {
            Spec e = tx_eval(sc, sc.txs[emitter], si);
            if (active) for (int k = 0; k < sc.S; ++k) result.v[k] += emission_weight * throughput.v[k] * e.v[k];
            if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;
        }
        active &= si.is_valid();
        if (depth > sc.integ.rr_depth) {
            float q = std::min(hmax_spec(throughput, sc.S) * 1.f, .95f);
            active &= sampler.next_1d() < q;
            float iq = rcp(q);
            for (int k = 0; k < sc.S; ++k) throughput.v[k] *= iq;
        }
        if ((uint32_t) depth >= (uint32_t) sc.integ.max_depth || !active) break;
        uint32_t bsdf = sc.shapes[si.shape].bsdf;
        bool active_e = active && bsdf_is_smooth(sc, bsdf);
        if (active_e) {
            TxDirResult tr = scene_sample_transmitter_direction(sc, si, sampler.next_2d(), cnt);
            active_e &= tr.ds.pdf != 0.f;
            V3 wo = si.to_local(tr.ds.d);
            Spec bsdf_val = active_e ? bsdf_eval(sc, bsdf, si, wo) : Spec(0.f);
            float bsdf_pdf_ = active_e ? bsdf_pdf(sc, bsdf, si, wo) : 0.f;
            float mis = tr.ds.delta ? 1.f : mis_weight(tr.ds.pdf, bsdf_pdf_);
            if (active_e) for (int k = 0; k < sc.S; ++k) result.v[k] += mis * throughput.v[k] * bsdf_val.v[k] * tr.spec.v[k];
            if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;   // pathlength.cpp:209 (inside `if (any(active_e))`)
        }
        float s1 = sampler.next_1d(); V2 s2 = sampler.next_2d();
        BSDFSample bs; Spec bsdf_val = bsdf_sample(sc, bsdf, si, s1, s2, bs);
        bool any_nz = false;
        for (int k = 0; k < sc.S; ++k) { throughput.v[k] = throughput.v[k] * bsdf_val.v[k]; any_nz |= throughput.v[k] != 0.f; }
        active &= any_nz;
        if (!active) break;
        ray = spawn_ray(si, si.to_world(bs.wo));
        SI si_bsdf = scene_ray_intersect(sc, ray, cnt);
        emitter = si_bsdf.is_valid() ? sc.shapes[si_bsdf.shape].tx : -1;
        if (emitter >= 0) {
            DirectionSample ds; ds.p = si_bsdf.p; ds.n = si_bsdf.sh.n; ds.uv = si_bsdf.uv; ds.time = si_bsdf.time;
            ds.d = si_bsdf.p - si.p; ds.dist = norm(ds.d); ds.d = ds.d / ds.dist; ds.object = emitter;
            float emitter_pdf = !bs.delta ? scene_pdf_transmitter_direction(sc, si, ds) : 0.f;
            emission_weight = mis_weight(bs.pdf, emitter_pdf);
        }
        si = si_bsdf;
        if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;
        else scalar += si.is_valid() ? si.t / 3.0e8f : 0.f;
    }
    out.result = result; out.valid = valid_ray; out.scalar = scalar; out.depth = (uint32_t) depth;
    return out;
}

// ---- sample loop + film ----------------------------------------------------------------------------
#define ORC_MAX_CHANNELS 1024
static uint32_t n_channels(const Scene &sc) {
    if (sc.integ.gen == BF_GEN_PHASE) return 3 + sc.integ.phase_bins;
    if (sc.integ.gen == BF_GEN_RANGE) return 5 + sc.integ.range_bins;
    return 5 + 50 * 3;
}
static void film_size(const Scene &sc, uint32_t &w, uint32_t &h) {
    if (sc.integ.gen == BF_GEN_PHASE) { w = sc.adc.window_t_bins; h = sc.adc.window_f_bins; }
    else { w = sc.rx.film_width ? sc.rx.film_width : 1; h = sc.rx.film_height ? sc.rx.film_height : 1; }
}
// SignalBlock::put single-bin branch (signalblock.cpp:162-169); ImageBlock::put is its twin.
static inline void block_put(float *data, uint32_t C, int off_x, int off_y, uint32_t size_x, uint32_t size_y, V2 pos_, const float *value) {
    float px = pos_.x - ((float) off_x - 0.f + .5f), py = pos_.y - ((float) off_y - 0.f + .5f);
    float cx = std::ceil(px - .5f), cy = std::ceil(py - .5f);
    // ceil2int: NaN / out-of-range convert to INT_MIN on x86 and are rejected by the unsigned test
    if (!(cx >= 0.f && cy >= 0.f && cx < (float) size_x && cy < (float) size_y)) return;
    uint32_t lx = (uint32_t) cx, ly = (uint32_t) cy;
    size_t offset = (size_t) C * ((size_t) ly * size_x + lx);
    for (uint32_t k = 0; k < C; ++k) data[offset + k] += value[k];
}
// Discretised reconstruction filter: ReconstructionFilter::init_discretization (src/libcore/rfilter.cpp:9-20) over
// GaussianFilter::eval (src/rfilters/gaussian.cpp:37-49; radius = 4 stddev) or TentFilter::eval (src/rfilters/tent.cpp:30-38).
struct Filter { float values[32]; float radius, scale; uint32_t n, border; bool wide; };
static Filter make_filter(const bf_adc_desc &a) {
    Filter f;
    std::memset(&f, 0, sizeof(f));
    f.radius = a.filter_radius;
    f.wide = a.filter_radius > .5f + RayEpsilon;                      // signalblock.cpp:113
    if (!f.wide) return f;
    float m_stddev = a.filter_radius / 4.f;
    float m_alpha = -1.f / (2.f * m_stddev * m_stddev);
    float m_bias = std::exp(m_alpha * (f.radius * f.radius));
    float m_inv_radius = 1.f / f.radius;
    for (size_t i = 0; i < 31; ++i) {
        float x = (f.radius * i) / 31, v = 0.f;
        if (a.filter_kind == 1) v = std::max(0.f, std::exp(m_alpha * (x * x)) - m_bias);
        else if (a.filter_kind == 2) v = std::max(0.f, 1.f - std::abs(x * m_inv_radius));
        else if (a.filter_kind == 3 || a.filter_kind == 4) {
            // MitchellNetravaliFilter::eval (src/rfilters/mitchell.cpp:52-64, B = C = 1/3); CatmullRomFilter::eval
            // (src/rfilters/catmullrom.cpp:33-45, B = 0, C = 0.5)
            float m_b = a.filter_kind == 3 ? 1.f / 3.f : 0.f, m_c = a.filter_kind == 3 ? 1.f / 3.f : .5f;
            float x2 = x * x, x3 = x2 * x;
            float result = (1.f / 6.f) * (x < 1 ? (12.f - 9.f * m_b - 6.f * m_c) * x3 + (-18.f + 12.f * m_b + 6.f * m_c) * x2 + (6.f - 2.f * m_b)
                                                : (-m_b - 6.f * m_c) * x3 + (6.f * m_b + 30.f * m_c) * x2 + (-12.f * m_b - 48.f * m_c) * x + (8.f * m_b + 24.f * m_c));
            v = x < 2.f ? result : 0.f;
        } else if (a.filter_kind == 5) {
            // LanczosSincFilter::eval (src/rfilters/lanczos.cpp:47-55), radius = lobes
            float x1 = 3.14159265358979323846f * x, x2 = x1 / f.radius, result = (std::sin(x1) * std::sin(x2)) / (x1 * x2);
            v = x < 5.9604644775390625e-08f ? 1.f : (x > f.radius ? 0.f : result);
        }
        f.values[i] = v;
    }
    f.values[31] = 0;
    f.scale = 31 / f.radius;
    f.border = (uint32_t) (int) std::ceil(f.radius - .5f - 2.f * RayEpsilon);
    f.n = (uint32_t) (int) std::ceil((f.radius - 2.f * RayEpsilon) * 2.f);     // signalblock.cpp:121
    return f;
}
static inline float eval_discretized(const Filter &f, float x) {       // include/mitsuba/core/rfilter.h:62-65
    int index = std::min((int) std::abs(x * f.scale), 31);
    return f.values[std::max(index, 0)];
}
// SignalBlock::put, wide branch (signalblock.cpp:113-161), into a block with a border of f.border bins, followed by the crop
// of SignalBlock::put(block) into the border-less ADC storage (signalblock.cpp:49-77, hdradc.cpp:166-176).
static inline void block_put_wide(float *data, uint32_t C, int off_x, int off_y, uint32_t size_x, uint32_t size_y, const Filter &f, V2 pos_, const float *value) {
    const int border = (int) f.border;
    const int bsx = (int) size_x + 2 * border, bsy = (int) size_y + 2 * border;
    float px = pos_.x - ((float) (off_x - border) + .5f), py = pos_.y - ((float) (off_y - border) + .5f);
    if (!(std::abs(px) < 1e9f && std::abs(py) < 1e9f)) return;      // non-finite positions index out of bounds in the reference
    int lox = std::max((int) std::ceil(px - f.radius), 0), loy = std::max((int) std::ceil(py - f.radius), 0);
    uint32_t hix = (uint32_t) std::min((int) std::floor(px + f.radius), bsx - 1), hiy = (uint32_t) std::min((int) std::floor(py + f.radius), bsy - 1);
    float basex = (float) lox - px, basey = (float) loy - py;
    float wx[20], wy[20];
    for (uint32_t i = 0; i < f.n; ++i) { wx[i] = eval_discretized(f, basex + (float) i); wy[i] = eval_discretized(f, basey + (float) i); }
    for (uint32_t yr = 0; yr < f.n; ++yr) {
        uint32_t y = (uint32_t) loy + yr;
        bool enabled = y <= hiy;
        for (uint32_t xr = 0; xr < f.n; ++xr) {
            uint32_t x = (uint32_t) lox + xr;
            float weight = wy[yr] * wx[xr];
            enabled &= x <= hix;
            int tx = (int) x - border, ty = (int) y - border;      // crop to the storage
            if (!enabled || tx < 0 || ty < 0 || tx >= (int) size_x || ty >= (int) size_y) continue;
            size_t offset = (size_t) C * ((size_t) ty * size_x + (size_t) tx);
            for (uint32_t k = 0; k < C; ++k) data[offset + k] += value[k] * weight;
        }
    }
}
static void seed_sample(const Scene &sc, Sampler &s, uint64_t seed, uint64_t idx) {
    (void) sc;
    s.rng.seed(sample_tea_64_wide(seed, idx), sample_tea_64_wide(idx, seed));  // sampler.cpp:89-92
}
// One iteration of receive_sample (integrator.cpp:1538-1667) or render_sample (:255-310).
static void one_sample(const Scene &sc, Sampler &sampler, float *block, bf_sample_record *rec) {
    Counters cnt;
    uint32_t C = n_channels(sc);
    float aovs[ORC_MAX_CHANNELS];
    for (uint32_t k = 0; k < C; ++k) aovs[k] = 0.f;
    uint32_t W, H; film_size(sc, W, H);
    if (sc.integ.gen == BF_GEN_PHASE) {
        V2 position_sample = sampler.next_2d();
        V2 aperture_sample(.5f, .5f);
        if (rx_needs_aperture_sample(sc)) aperture_sample = sampler.next_2d();
        float time = sc.rx.adc_sampling_start;
        if (sc.rx.adc_sampling_time > 0.f) time += sampler.next_1d() * sc.rx.adc_sampling_time; else time = 0.f;
        float wavelength_sample = sampler.next_1d();
        Ray ray; Spec ray_weight;
        rx_sample_ray(sc, time, wavelength_sample, position_sample, aperture_sample, ray, ray_weight);
        PathOut res; V2 tf(0.f, 0.f);
        bool ran = true;
        if (sc.rx.receive_type == BF_RECV_MIX_RESAMPLE) {
            float f_rx0 = freq_from_wl(sc.C.c, ray.wl.v[0]);
            res = path_time_frequency(sc, sampler, ray, &cnt);
            tf.x = time - sc.rx.adc_sampling_start;
            tf.y = (float) std::fabs(freq_from_wl_d(sc.C.c, ray.wl.v[0]) - (double) f_rx0);
        } else if (sc.rx.receive_type == BF_RECV_RAW || sc.rx.receive_type == BF_RECV_RAW_RESAMPLE) {
            res = path_time_frequency(sc, sampler, ray, &cnt);
            tf.x = time - sc.rx.adc_sampling_start;
            tf.y = freq_from_wl(sc.C.c, ray.wl.v[0]);
        } else { ran = false; }
        // phase.cpp:85-138 (phase_bins == 0: bare pathtimefrequency, no extra channels)
        float hs = hsum_spec(res.result, sc.S);
        float phase = std::fmod(ray.phase, TwoPi);
        phase += phase < 0.f ? TwoPi : 0.f;
        if (ran) {
            float bin_width = TwoPi / (float) std::max(1u, sc.integ.phase_bins);
            for (uint32_t k = 0; k < sc.integ.phase_bins; ++k) {
                float centre = (float) (bin_width * (k + 0.5));
                aovs[3 + k] = rect((phase - centre) / bin_width) > 0.f ? (res.valid ? hs : 0.f) : 0.f;
            }
        }
        tf.x *= (float) sc.adc.t_bins / sc.adc.t_bandwidth;
        tf.y *= (float) sc.adc.f_bins / sc.adc.f_bandwidth;
        Spec wres; for (int k = 0; k < sc.S; ++k) wres.v[k] = std::fabs(ray_weight.v[k]) * res.result.v[k];
        aovs[0] = res.valid ? hsum_spec(wres, sc.S) : 0.f;
        aovs[1] = res.valid ? 1.f : 0.f;
        aovs[2] = 1.f;
        if (block) {
            const Filter filt = make_filter(sc.adc);
            if (filt.wide) block_put_wide(block, C, (int) sc.adc.window_offset_t, (int) sc.adc.window_offset_f, W, H, filt, tf, aovs);
            else block_put(block, C, (int) sc.adc.window_offset_t, (int) sc.adc.window_offset_f, W, H, tf, aovs);
        }
        if (rec) {
            rec->first_shape = res.first.shape >= 0 ? (uint32_t) res.first.shape : BF_NONE;
            rec->first_prim = res.first.shape >= 0 ? res.first.prim : BF_NONE; rec->first_t = res.first.t;
            rec->valid = res.valid; rec->tf_t = tf.x; rec->tf_f = tf.y; rec->value = aovs[0]; rec->phase = ray.phase;
            rec->n_closest = cnt.n_closest; rec->n_shadow = cnt.n_shadow; rec->wavelength0 = ray.wl.v[0]; rec->depth = res.depth;
        }
    } else {
        // render_sample (integrator.cpp:255-310), 1x1 film: pos = (0,0)
        V2 position_sample = sampler.next_2d();
        V2 aperture_sample(.5f, .5f);
        if (rx_needs_aperture_sample(sc)) aperture_sample = sampler.next_2d();
        float time = sc.rx.shutter_open;
        if (sc.rx.shutter_open_time > 0.f) time += sampler.next_1d() * sc.rx.shutter_open_time;
        float wavelength_sample = sampler.next_1d();
        V2 adjusted((position_sample.x - 0.f) / (float) W, (position_sample.y - 0.f) / (float) H);
        Ray ray; Spec ray_weight;
        rx_sample_ray(sc, time, wavelength_sample, adjusted, aperture_sample, ray, ray_weight);
        PathOut res = path_emitters(sc, sampler, ray, &cnt, (int) sc.integ.gen);
        // wrappers write their bins from the UNWEIGHTED spectrum (range.cpp:141-161, time.cpp:118-153)
        if (sc.integ.gen == BF_GEN_RANGE) {
            for (uint32_t i = 0; i < sc.integ.range_bins; ++i) {
                float lo = (float) i * sc.integ.dr, hi = (float) i * sc.integ.dr + sc.integ.dr;
                aovs[5 + i] = (res.scalar >= lo && res.scalar < hi) ? res.result.v[0] : 0.f;
            }
        } else {
            float dt = 0.5e-9f;
            for (int i = 0; i < 50; ++i) {
                float lo = (float) i * dt, hi = (float) i * dt + dt;
                float x = (res.scalar >= lo && res.scalar < hi) ? res.result.v[0] : 0.f;
                aovs[5 + 3 * i] = x; aovs[5 + 3 * i + 1] = x; aovs[5 + 3 * i + 2] = x;
            }
        }
        Spec wres; for (int k = 0; k < sc.S; ++k) wres.v[k] = ray_weight.v[k] * res.result.v[k];
        // mono: xyz = spec.x broadcast; spectral: CIE matching functions vanish at radar wavelengths
        float x = sc.S == 1 ? wres.v[0] : 0.f;
        aovs[0] = x; aovs[1] = x; aovs[2] = x; aovs[3] = res.valid ? 1.f : 0.f; aovs[4] = 1.f;
        if (block) block_put(block, C, 0, 0, W, H, position_sample, aovs);
        if (rec) {
            rec->first_shape = res.first.shape >= 0 ? (uint32_t) res.first.shape : BF_NONE;
            rec->first_prim = res.first.shape >= 0 ? res.first.prim : BF_NONE; rec->first_t = res.first.t;
            rec->valid = res.valid; rec->tf_t = position_sample.x; rec->tf_f = position_sample.y; rec->value = res.result.v[0];
            rec->phase = res.scalar; rec->n_closest = cnt.n_closest; rec->n_shadow = cnt.n_shadow; rec->wavelength0 = ray.wl.v[0];
            rec->depth = res.depth;
        }
    }
}

static void run_range(const Scene &sc, uint64_t seed, uint64_t first, uint64_t n, float *block, bf_sample_record *recs, int rng_mode, uint64_t *queries) {
    Sampler sampler;
    if (rng_mode == 1) sampler.rng.seed(seed + PCG32_DEFAULT_STATE, PCG32_DEFAULT_STREAM);  // independent.cpp:59-64
    uint64_t q = 0;
    bf_sample_record tmp;
    for (uint64_t i = 0; i < n; ++i) {
        if (rng_mode == 0) seed_sample(sc, sampler, seed, first + i);
        bf_sample_record *r = recs ? recs + i : (queries ? &tmp : nullptr);
        one_sample(sc, sampler, block, r);
        if (r) q += r->n_closest + r->n_shadow;
    }
    if (queries) *queries = q;
}

}  // namespace orc

// =====================================================================================================
// C API
// =====================================================================================================
using namespace orc;
extern "C" {

struct orc_scene { Scene sc; };

const char *orc_last_error(orc_scene *s) { return s ? s->sc.err.c_str() : "null scene"; }

// Rebuild a scene from the flat blob written by bf_scene_export.
orc_scene *orc_scene_import(const void *blob, size_t bytes) {
    const uint8_t *p = (const uint8_t *) blob;
    if (bytes < sizeof(bf_flat_header)) return nullptr;
    const bf_flat_header *h = (const bf_flat_header *) p;
    if (h->magic != BF_FLAT_MAGIC) return nullptr;
    orc_scene *os = new orc_scene();
    Scene &sc = os->sc;
    sc.C = h->consts; sc.S = (int) h->consts.spectrum_width;
    size_t off = sizeof(bf_flat_header);
    while (off + sizeof(bf_flat_record) <= bytes) {
        const bf_flat_record *r = (const bf_flat_record *) (p + off);
        const uint8_t *pl = p + off + sizeof(bf_flat_record);
        switch (r->tag) {
            case BF_TAG_SPECTRUM: {
                SpectrumTex t; const uint32_t *u = (const uint32_t *) pl; t.kind = u[0]; uint32_t n = u[1];
                const float *f = (const float *) (pl + 8);
                t.lambda.assign(f, f + n); t.value.assign(f + n, f + 2 * n);
                if (t.kind == BF_SPEC_REGULAR && n >= 2) {
                    double interval = ((double) t.lambda.back() - (double) t.lambda.front()) / (double) (n - 1);
                    t.inv_interval = (float) (1.0 / interval);
                }
                sc.spectra.push_back(t); break;
            }
            case BF_TAG_BSDF: sc.bsdfs.push_back(*(const bf_bsdf_desc *) pl); break;
            case BF_TAG_MESH: {
                const uint32_t *u = (const uint32_t *) pl; MeshData m; m.nv = u[0]; m.nf = u[1];
                uint32_t bsdf = u[2], has_n = u[3];
                const float *pos = (const float *) (pl + 16); const uint32_t *idx = (const uint32_t *) (pos + 3 * (size_t) m.nv);
                m.pos.assign(pos, pos + 3 * (size_t) m.nv); m.idx.assign(idx, idx + 3 * (size_t) m.nf);
                if (has_n) { const float *nr = (const float *) (idx + 3 * (size_t) m.nf); m.nrm.assign(nr, nr + 3 * (size_t) m.nv); }
                ShapeRec s; s.type = 0; s.index = (int) sc.meshes.size(); s.bsdf = bsdf;
                sc.meshes.push_back(std::move(m)); sc.shapes.push_back(s); break;
            }
            case BF_TAG_RECT: {
                const float *f = (const float *) pl; uint32_t bsdf = *(const uint32_t *) (pl + 64);
                ShapeRec s; s.type = 1; s.index = (int) sc.rects.size(); s.bsdf = bsdf;
                sc.rects.push_back(make_rect(f)); sc.shapes.push_back(s); break;
            }
            case BF_TAG_TX: {
                TxRec t; t.shape = (int) *(const uint32_t *) pl; if (*(const uint32_t *) pl == BF_NONE) t.shape = -1;
                t.d = *(const bf_endpoint_desc *) (pl + 8);
                t.to_world = M4::from_rows16(t.d.to_world); t.to_object = inverse(t.to_world);
                if (t.d.kind == BF_TX_SPOT) {                       // spot.cpp:85-93
                    float cutoff = t.d.cutoff_angle_deg * (Pi / 180.f), beam = t.d.beam_width_deg * (Pi / 180.f);
                    t.cutoff = cutoff; t.inv_transition = 1.0f / (cutoff - beam);
                    t.cos_cutoff = std::cos(cutoff); t.cos_beam = std::cos(beam);
                }
                if (t.d.kind == BF_TX_PHASED) t.velems = make_velems(t.d.array, sc.C);
                if (t.shape >= 0) sc.shapes[t.shape].tx = (int) sc.txs.size();
                sc.txs.push_back(t); break;
            }
            case BF_TAG_RX: {
                uint32_t sh = *(const uint32_t *) pl; sc.rx_shape = sh == BF_NONE ? -1 : (int) sh;
                sc.rx = *(const bf_endpoint_desc *) (pl + 8); sc.has_rx = true;
                sc.rx_to_world = M4::from_rows16(sc.rx.to_world);
                if (sc.rx.kind == BF_RX_PHASED) sc.rx_velems = make_velems(sc.rx.array, sc.C);
                break;
            }
            case BF_TAG_ADC: sc.adc = *(const bf_adc_desc *) pl; break;
            case BF_TAG_INTEGRATOR: sc.integ = *(const bf_integrator_desc *) pl; break;
            case BF_TAG_MESH_MOTION: {
                // bf_scene_update_transform on a mesh: world = motion * rest (Transform::transform_affine, transform.h:91-99);
                // normals by the inverse transpose, re-normalised (mesh.cpp:117-124)
                uint32_t shape = *(const uint32_t *) pl;
                if (shape >= sc.shapes.size() || sc.shapes[shape].type != 0) break;
                MeshData &m = sc.meshes[sc.shapes[shape].index];
                if (m.pos_rest.empty()) { m.pos_rest = m.pos; m.nrm_rest = m.nrm; }
                M4 M = M4::from_rows16((const float *) (pl + 8)), Minv = inverse(M);
                for (uint32_t v = 0; v < m.nv; ++v) {
                    V3 p = xf_point_affine(M, V3(m.pos_rest[3 * v], m.pos_rest[3 * v + 1], m.pos_rest[3 * v + 2]));
                    m.pos[3 * v] = p.x; m.pos[3 * v + 1] = p.y; m.pos[3 * v + 2] = p.z;
                    if (!m.nrm_rest.empty()) {
                        V3 n = xf_normal(Minv, V3(m.nrm_rest[3 * v], m.nrm_rest[3 * v + 1], m.nrm_rest[3 * v + 2]));
                        if (n.x != 0.f || n.y != 0.f || n.z != 0.f) n = normalize(n);
                        m.nrm[3 * v] = n.x; m.nrm[3 * v + 1] = n.y; m.nrm[3 * v + 2] = n.z;
                    }
                }
                break;
            }
            case BF_TAG_VELOCITY: {
                uint32_t shape = *(const uint32_t *) pl;
                if (sc.velocities.size() < sc.shapes.size()) { M4 id; memset(&id, 0, sizeof(id)); for (int i = 0; i < 4; ++i) id.m[i][i] = 1.f; sc.velocities.resize(sc.shapes.size(), id); }
                if (shape < sc.velocities.size()) memcpy(sc.velocities[shape].m, pl + 8, 64);
                break;
            }
            default: break;
        }
        off += sizeof(bf_flat_record) + ((r->payload_bytes + 7) & ~(size_t) 7);
    }
    // the receiver's record may precede the record of the shape it sits on (a re-posed receiver is rewritten in place)
    if (sc.rx_shape >= 0 && (size_t) sc.rx_shape < sc.shapes.size()) sc.shapes[sc.rx_shape].is_rx = true;
    return os;
}
void orc_scene_destroy(orc_scene *s) {
    if (!s) return;
    if (s->sc.em_scene) { s->sc.em.rtcReleaseScene(s->sc.em_scene); s->sc.em.rtcReleaseDevice(s->sc.em_dev); }
    delete s;
}
// accel: 0 brute force, 1 oracle BVH, 2 Embree (embree_path = oracle/_ref/libembree3.so)
int orc_scene_set_accel(orc_scene *s, int accel, const char *embree_path) {
    Scene &sc = s->sc;
    if (accel == 1 && sc.nodes.empty()) bvh_build(sc);
    if (accel == 2 && !sc.em_scene) { if (!embree_path || !embree_attach(sc, embree_path)) return -1; }
    sc.accel = accel;
    return 0;
}
int orc_profile_size(orc_scene *s, uint32_t *w, uint32_t *h, uint32_t *c) {
    film_size(s->sc, *w, *h); *c = n_channels(s->sc); return 0;
}
// rng_mode 0: per-sample streams (GPU-comparable); 1: reference-faithful single stream
int orc_receive(orc_scene *s, uint64_t seed, uint64_t first, uint64_t n, float *profile, int accumulate, int rng_mode, int n_threads, uint64_t *queries) {
    const Scene &sc = s->sc;
    uint32_t W, H; film_size(sc, W, H); uint32_t C = n_channels(sc);
    size_t nf = (size_t) W * H * C;
    if (!accumulate) std::fill(profile, profile + nf, 0.f);
    if (n_threads <= 1 || rng_mode == 1) { run_range(sc, seed, first, n, profile, nullptr, rng_mode, queries); return 0; }
    std::vector<std::vector<float>> priv(n_threads, std::vector<float>(nf, 0.f));
    std::vector<uint64_t> q(n_threads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; ++t) {
        uint64_t a = first + n * t / n_threads, b = first + n * (t + 1) / n_threads;
        th.emplace_back([&, t, a, b]() { run_range(sc, seed, a, b - a, priv[t].data(), nullptr, 0, &q[t]); });
    }
    for (auto &t : th) t.join();
    uint64_t qq = 0;
    for (int t = 0; t < n_threads; ++t) { for (size_t i = 0; i < nf; ++i) profile[i] += priv[t][i]; qq += q[t]; }
    if (queries) *queries = qq;
    return 0;
}
int orc_sample_records(orc_scene *s, uint64_t seed, uint64_t first, uint64_t n, bf_sample_record *recs, int rng_mode) {
    run_range(s->sc, seed, first, n, nullptr, recs, rng_mode, nullptr); return 0;
}
int orc_trace_first_hits(orc_scene *s, const float *o, const float *d, const float *mint, const float *maxt, uint64_t n,
                         uint32_t *shape, uint32_t *prim, float *t, float *u, float *v, int accel) {
    for (uint64_t i = 0; i < n; ++i) {
        Ray r; r.o = V3(o[3 * i], o[3 * i + 1], o[3 * i + 2]); r.d = V3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        r.mint = mint ? mint[i] : RayEpsilon; r.maxt = maxt ? maxt[i] : Infinity;
        PI pi = scene_intersect_preliminary(s->sc, r, nullptr, accel);
        shape[i] = pi.shape >= 0 ? (uint32_t) pi.shape : BF_NONE; prim[i] = pi.shape >= 0 ? pi.prim : BF_NONE;
        t[i] = pi.t; if (u) u[i] = pi.uv.x; if (v) v[i] = pi.uv.y;
    }
    return 0;
}
int orc_trace_occluded(orc_scene *s, const float *o, const float *d, const float *mint, const float *maxt, uint64_t n, uint8_t *hit, int accel) {
    for (uint64_t i = 0; i < n; ++i) {
        Ray r; r.o = V3(o[3 * i], o[3 * i + 1], o[3 * i + 2]); r.d = V3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        r.mint = mint ? mint[i] : RayEpsilon; r.maxt = maxt ? maxt[i] : Infinity;
        hit[i] = scene_ray_test(s->sc, r, nullptr, accel) ? 1 : 0;
    }
    return 0;
}
// surface interaction at a first hit (test_mesh.py:257-298 known answers): out = p(3) n(3) uv(2) dp_du(3) wi(3) sh_n(3)
int orc_surface_interaction(orc_scene *s, const float *o, const float *d, float *out17, float *t_out, uint32_t *prim_out) {
    Ray r; r.o = V3(o[0], o[1], o[2]); r.d = V3(d[0], d[1], d[2]);
    PI pi; SI si = scene_ray_intersect(s->sc, r, nullptr, &pi);
    *t_out = si.t; *prim_out = pi.prim;
    if (!si.is_valid()) return 1;
    float v[17] = {si.p.x, si.p.y, si.p.z, si.n.x, si.n.y, si.n.z, si.uv.x, si.uv.y, si.dp_du.x, si.dp_du.y, si.dp_du.z,
                   si.wi.x, si.wi.y, si.wi.z, si.sh.n.x, si.sh.n.y, si.sh.n.z};
    std::memcpy(out17, v, sizeof(v));
    return 0;
}

// ---- unit-level hooks for the known-answer tests ----------------------------------------------------
uint32_t orc_pcg32_next(uint64_t *state, uint64_t *inc) { PCG32 r; r.state = *state; r.inc = *inc; uint32_t v = r.next_u32(); *state = r.state; return v; }
void orc_pcg32_seed(uint64_t initstate, uint64_t initseq, uint64_t *state, uint64_t *inc) { PCG32 r(initstate, initseq); *state = r.state; *inc = r.inc; }
float orc_pcg32_next_float(uint64_t *state, uint64_t *inc) { PCG32 r; r.state = *state; r.inc = *inc; float v = r.next_float32(); *state = r.state; return v; }
uint32_t orc_tea32(uint32_t a, uint32_t b, int rounds) { return sample_tea_32(a, b, rounds); }
uint64_t orc_tea64(uint32_t a, uint32_t b, int rounds) { return sample_tea_64(a, b, rounds); }
uint64_t orc_tea64_wide(uint64_t a, uint64_t b, int rounds) { return sample_tea_64_wide(a, b, rounds); }
float orc_tea_float32(uint32_t a, uint32_t b, int rounds) { return sample_tea_float32(a, b, rounds); }
double orc_tea_float64(uint32_t a, uint32_t b, int rounds) { return sample_tea_float64(a, b, rounds); }
void orc_square_to_uniform_disk_concentric(float x, float y, float *out) { V2 r = square_to_uniform_disk_concentric(V2(x, y)); out[0] = r.x; out[1] = r.y; }
void orc_square_to_cosine_hemisphere(float x, float y, float *out) { V3 r = square_to_cosine_hemisphere(V2(x, y)); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
void orc_coordinate_system(const float *n, float *s, float *t) { V3 a, b; coordinate_system(V3(n[0], n[1], n[2]), a, b); s[0] = a.x; s[1] = a.y; s[2] = a.z; t[0] = b.x; t[1] = b.y; t[2] = b.z; }
float orc_sinc(float x) { return sinc(x); }
float orc_tri(float x) { return tri(x); }
float orc_rect(float x) { return rect(x); }
float orc_fmodulo(float a, float b) { return fmodulo(a, b); }
float orc_wchirp(float t, float f, float w, float a) { return wchirp(t, f, w, a); }
float orc_erfinv(float x) { return erfinv(x); }
float orc_fresnel_conductor(float c, float er, float ei) { return fresnel_conductor(c, er, ei); }
float orc_microfacet_eval(int type, float au, float av, const float *m) { return Microfacet(type, au, av, true).eval(V3(m[0], m[1], m[2])); }
float orc_microfacet_smith_g1(int type, float au, float av, const float *v, const float *m) { return Microfacet(type, au, av, true).smith_g1(V3(v[0], v[1], v[2]), V3(m[0], m[1], m[2])); }
void orc_microfacet_sample(int type, float au, float av, int visible, const float *wi, float u1, float u2, float *m_out, float *pdf_out) {
    V3 m; Microfacet(type, au, av, visible != 0).sample(V3(wi[0], wi[1], wi[2]), V2(u1, u2), m, *pdf_out); m_out[0] = m.x; m_out[1] = m.y; m_out[2] = m.z;
}
// BSDF hooks on a scene's BSDF table: wi / wo in the local frame, wavelengths from `wl`
static SI hook_si(orc_scene *s, const float *wi, const float *wl) { SI si; si.t = 1.f; si.wi = V3(wi[0], wi[1], wi[2]); for (int k = 0; k < s->sc.S; ++k) si.wl.v[k] = wl[k]; return si; }
void orc_bsdf_eval(orc_scene *s, uint32_t id, const float *wi, const float *wo, const float *wl, float *out) { Spec r = bsdf_eval(s->sc, id, hook_si(s, wi, wl), V3(wo[0], wo[1], wo[2])); std::memcpy(out, r.v, 16); }
float orc_bsdf_pdf(orc_scene *s, uint32_t id, const float *wi, const float *wo, const float *wl) { return bsdf_pdf(s->sc, id, hook_si(s, wi, wl), V3(wo[0], wo[1], wo[2])); }
void orc_bsdf_sample(orc_scene *s, uint32_t id, const float *wi, float u1, float u2, float u3, const float *wl, float *wo, float *pdf, float *weight) {
    BSDFSample bs; Spec r = bsdf_sample(s->sc, id, hook_si(s, wi, wl), u1, V2(u2, u3), bs); wo[0] = bs.wo.x; wo[1] = bs.wo.y; wo[2] = bs.wo.z; *pdf = bs.pdf; std::memcpy(weight, r.v, 16);
}
float orc_rect_sample_wigner(orc_scene *s, uint32_t shape, const float *p, const float *d, float wavelength_nm) {
    return rect_sample_wigner(s->sc.rects[s->sc.shapes[shape].index], V3(p[0], p[1], p[2]), V3(d[0], d[1], d[2]), wavelength_nm);
}
// which: -1 = the receiver's array, >= 0 = transmitter index
float orc_phased_sample_wigner(orc_scene *s, int which, const float *p, const float *d, float wavelength_nm) {
    const std::vector<VElem> &ve = which < 0 ? s->sc.rx_velems : s->sc.txs[which].velems;
    const float *wid = which < 0 ? s->sc.rx.array.elem_dims : s->sc.txs[which].d.array.elem_dims;
    return phased_sample_wigner(ve, wid, V3(p[0], p[1], p[2]), V3(d[0], d[1], d[2]), wavelength_nm);
}
float orc_rect_area(orc_scene *s, uint32_t shape) { return s->sc.rects[s->sc.shapes[shape].index].area; }
float orc_eval_signal(const bf_signal_desc *sg, float time, float f0) { return eval_signal(*sg, time, f0); }
float orc_sample_delta_frequency(const bf_signal_desc *sg, float time) { return sample_delta_frequency(*sg, time); }
float orc_spectrum_eval(orc_scene *s, uint32_t id, float wl) { return spectrum_eval_1(s->sc, s->sc.spectra[id], wl); }
// SignalBlock::put bin edges: returns flat bin index or -1
int orc_block_bin(uint32_t size_x, uint32_t size_y, int off_x, int off_y, float px, float py) {
    std::vector<float> data((size_t) size_x * size_y, 0.f); float one = 1.f;
    block_put(data.data(), 1, off_x, off_y, size_x, size_y, V2(px, py), &one);
    for (size_t i = 0; i < data.size(); ++i) if (data[i] != 0.f) return (int) i;
    return -1;
}

// SignalBlock::put with a wide filter: one sample of value 1 at (px, py) into a size_x x size_y single-channel storage
void orc_block_splat(uint32_t size_x, uint32_t size_y, int off_x, int off_y, uint32_t filter_kind, float filter_radius, float px, float py, float *out) {
    bf_adc_desc a; std::memset(&a, 0, sizeof(a)); a.filter_kind = filter_kind; a.filter_radius = filter_radius;
    const Filter f = make_filter(a);
    float one = 1.f;
    if (f.wide) block_put_wide(out, 1, off_x, off_y, size_x, size_y, f, V2(px, py), &one);
    else block_put(out, 1, off_x, off_y, size_x, size_y, V2(px, py), &one);
}

}  // extern "C"
