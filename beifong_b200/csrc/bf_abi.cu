// beifong_b200 — C ABI (include/beifong_b200/abi.h): host scene object, flattening, upload, the
// wavefront driver loop.  There is no CPU fallback behind this file: bf_scene_commit fails with
// BF_ERR_CUDA when no device is usable and every compute entry point requires a committed scene.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>      // header-only; ranges are no-ops unless a profiler injects itself

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "bf_wave.h"

using namespace bf;

namespace bf {

M4 m4_inverse(const M4 &a) {
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

// Rectangle::Rectangle / update (src/shapes/rectangle.cpp:75-95)
DevRect make_rect(const float *to_world16, uint32_t shape) {
    DevRect r;
    memset(&r, 0, sizeof(r));
    memcpy(r.to_world.m, to_world16, 64);
    r.to_object = m4_inverse(r.to_world);
    V3 dp_du = xf_vector(r.to_world, V3(2.f, 0.f, 0.f)), dp_dv = xf_vector(r.to_world, V3(0.f, 2.f, 0.f));
    V3 normal = normalize(xf_normal(r.to_object, V3(0.f, 0.f, 1.f)));
    r.fs = dp_du; r.ft = dp_dv; r.fn = normal;
    r.area = norm(cross(r.fs, r.ft));
    r.inv_area = rcp(r.area);
    r.wid_x = norm(r.fs); r.wid_y = norm(r.ft);
    r.ns = normalize(r.fs); r.nt = normalize(r.ft); r.nn = normalize(r.fn);
    r.shape = shape;
    coordinate_system(r.fn, r.cs_s, r.cs_t);
    {   // interaction.h:159-162 with sh_frame.n = n = fn and dp_du = fs
        const float dn = dot(r.fn, r.fs);
        r.sh_s = normalize(V3(fnmadd(r.fn.x, dn, r.fs.x), fnmadd(r.fn.y, dn, r.fs.y), fnmadd(r.fn.z, dn, r.fs.z)));
        r.sh_t = cross(r.fn, r.sh_s);
    }
    r.two_pi_inv_area = rcp(r.area) * BF_TWO_PI;
    r.area_pi = r.area * BF_PI;
    // padded world-space box of the four corners: margin 1e-4 of the largest coordinate / extent involved, far above the
    // rounding of the object-space test it guards (a few 1e-7 relative)
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, mag = 1e-3f;
    for (int c = 0; c < 4; ++c) {
        V3 p = xf_point_affine(r.to_world, V3((c & 1) ? 1.f : -1.f, (c & 2) ? 1.f : -1.f, 0.f));
        const float v[3] = {p.x, p.y, p.z};
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], v[k]); hi[k] = std::max(hi[k], v[k]); mag = std::max(mag, std::fabs(v[k])); }
    }
    for (int k = 0; k < 3; ++k) { r.box_lo[k] = lo[k] - 1e-4f * mag; r.box_hi[k] = hi[k] + 1e-4f * mag; }
    return r;
}

static M4 m4_identity() { M4 r; memset(&r, 0, sizeof(r)); for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f; return r; }
static M4 m4_mul(const M4 &a, const M4 &b) {
    M4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) {
        double acc = 0;
        for (int k = 0; k < 4; ++k) acc += (double) a.m[i][k] * b.m[k][j];
        r.m[i][j] = (float) acc;
    }
    return r;
}
// Phased-array virtual elements, appended to `out` (phasedtransmitter.cpp:104-167).  to_world = array_loc *
// translate(r_v) * scale(w.x/2, w.y/2, w.z); Transform::inverse() of the product is the product of the explicit
// inverses (transform.h:53-61,161-171).  Load-time matrices are evaluated in double and rounded once.
void make_velems(const bf_array_desc &a, const bf_constants &C, std::vector<DevVElem> &out) {
    const int n = (int) a.n_elems;
    std::vector<V3> locs;
    for (int i = 0; i < n; ++i) {
        double k = (n % 2 == 0) ? ((double) ((float) i - ((float) n / 2.f)) + 0.5) : (double) ((float) i - ((float) n - 1.f) / 2.f);
        locs.push_back(V3((float) (0.0 - (double) (a.elem_spacing[0] * a.elem_axis[0]) * k),
                          (float) (0.0 - (double) (a.elem_spacing[1] * a.elem_axis[1]) * k),
                          (float) (0.0 - (double) (a.elem_spacing[2] * a.elem_axis[2]) * k)));
    }
    const V3 steer(sinf(a.steering_vector[0]), sinf(a.steering_vector[1]), sinf(a.steering_vector[2]));
    M4 A; memcpy(A.m, a.array_loc, 64);
    const M4 Ainv = m4_inverse(A);
    const float k_psi = (float) (1.0 / ((double) (C.lambda_max_nm - C.lambda_min_nm) * 1e-9 / 2));
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) {
        DevVElem e;
        memset(&e, 0, sizeof(e));
        const V3 r_v = (locs[i] + locs[j]) / 2.f;
        e.r_dash = locs[i] - locs[j];
        const float sx = a.elem_dims[0] / 2, sy = a.elem_dims[1] / 2, sz = a.elem_dims[2];
        M4 TS, TSinv;
        memset(&TS, 0, sizeof(TS)); memset(&TSinv, 0, sizeof(TSinv));
        TS.m[0][0] = sx; TS.m[1][1] = sy; TS.m[2][2] = sz; TS.m[3][3] = 1.f; TS.m[0][3] = r_v.x; TS.m[1][3] = r_v.y; TS.m[2][3] = r_v.z;
        TSinv.m[0][0] = rcp(sx); TSinv.m[1][1] = rcp(sy); TSinv.m[2][2] = rcp(sz); TSinv.m[3][3] = 1.f;
        TSinv.m[0][3] = -r_v.x * rcp(sx); TSinv.m[1][3] = -r_v.y * rcp(sy); TSinv.m[2][3] = -r_v.z * rcp(sz);
        const M4 to_world = m4_mul(A, TS);
        e.to_object = m4_mul(TSinv, Ainv);
        const V3 dp_du = xf_vector(to_world, V3(2.f, 0.f, 0.f)), dp_dv = xf_vector(to_world, V3(0.f, 2.f, 0.f));
        const V3 normal = normalize(xf_normal(e.to_object, V3(0.f, 0.f, 1.f)));
        e.ns = normalize(dp_du); e.nt = normalize(dp_dv); e.nn = normalize(normal);
        const float arg = k_psi * dot(V3(0.f, 0.f, 0.f) - e.r_dash, steer);
        e.psi_re = cosf(arg); e.psi_im = sinf(arg);
        out.push_back(e);
    }
}

}  // namespace bf

struct HostShape { uint32_t type, index, bsdf; int32_t tx; };
struct HostTx { int32_t shape; bf_endpoint_desc d; };

template <typename T> struct DevBuf {
    T *p = nullptr; size_t n = 0;
    cudaError_t alloc(size_t count) {
        if (count <= n && p) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; n = 0;
        cudaError_t e = cudaMalloc((void **) &p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) n = count;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

struct WaveSlot {
    Wave wave; uint32_t capacity = 0; std::vector<void *> allocs;
    cudaStream_t stream = nullptr; bool own_stream = false; cudaEvent_t done = nullptr;
    uint32_t *h_counts = nullptr;     // pinned read-back of queue counts
    uint32_t *path_stash = nullptr; uint32_t stash_cap = 0;   // k_path's parking areas (path_stash_bytes), allocated on first use
    DevRect *d_rects = nullptr; size_t n_rects = 0; DevEndpoint *d_rx = nullptr;   // bf_receive_batch: this slot's copy of the tables an item re-poses
};

// Launch-shape knobs.  Defaults come from the environment ONCE, when the scene is created (BF_WAVE_LANES, BF_WAVE_STREAMS,
// BF_FUSED_PRIMARY, BF_PRIMARY_DEFER, BF_TRACE_DEFER, BF_TRACE_ROUNDS); bf_set_tunable changes them per scene afterwards.
// None of them changes a result: they select between launch sequences that are tested bit-identical per sample.
struct Tunables {
    int64_t wave_lanes = 0;       // 0: chosen from the request size
    int64_t wave_streams = 8;
    int64_t fused_primary = 1;
    int64_t primary_defer = -1;   // -1: bf_scene_commit's estimate
    int64_t trace_defer = -1;
    int64_t trace_rounds = -1;    // batches of 32 rays per active traversal warp (0: all warps of the grid, -1: by the request, see receive_into)
    int64_t use_graph = 1;        // small calls: replay the wave's launch sequence from a CUDA graph
    int64_t mega_batch = 1;       // bf_receive_batch: items that k_mega can run go out as one launch
    int64_t path_fused = 1;       // scenes with a deferred primary: k_path integrates the rectangle-only paths in registers
    int64_t path_stash_max = 65536;   // k_path: most entries of one CTA's parking area (path_stash_cap); 17 (S = 1) / 23 (S = 4) words x 148 CTAs each per wave slot
    int64_t l2_persist_mb = 0;    // > 0: an L2 access-policy window (persisting) over the first MB of the breadth-first BVH node array, i.e. its top levels
    int64_t mega_max = 262144;    // requests up to this many samples, and every request on a scene without meshes (-1: never), run as ONE k_mega launch
};
static int64_t env_i64(const char *name, int64_t dflt, int64_t lo, int64_t hi) {
    const char *e = getenv(name);
    if (!e || !*e) return dflt;
    long long v = atoll(e);
    if (v < lo) v = lo;
    if (v > hi) v = hi;
    return (int64_t) v;
}
static Tunables tunables_from_env() {
    Tunables t;
    t.wave_lanes = env_i64("BF_WAVE_LANES", 0, 0, 1ll << 28);
    t.wave_streams = env_i64("BF_WAVE_STREAMS", 8, 1, 16);
    t.fused_primary = env_i64("BF_FUSED_PRIMARY", 1, 0, 1);
    t.primary_defer = env_i64("BF_PRIMARY_DEFER", -1, -1, 1);
    t.trace_defer = env_i64("BF_TRACE_DEFER", -1, -1, 1);
    t.trace_rounds = env_i64("BF_TRACE_ROUNDS", -1, -1, 1024);
    t.use_graph = env_i64("BF_USE_GRAPH", 1, 0, 1);
    t.mega_batch = env_i64("BF_MEGA_BATCH", 1, 0, 1);
    t.path_fused = env_i64("BF_PATH_FUSED", 1, 0, 1);
    t.mega_max = env_i64("BF_MEGA_MAX", 262144, -1, 1ll << 28);
    t.l2_persist_mb = env_i64("BF_L2_PERSIST_MB", 0, 0, 4096);
    t.path_stash_max = env_i64("BF_PATH_STASH_MAX", 65536, 1024, 1 << 20);
    return t;
}

struct bf_scene {
    bf_constants C;
    Tunables tune;
    std::vector<HostSpectrum> spectra;
    std::vector<bf_bsdf_desc> bsdfs;
    std::vector<HostMesh> meshes;
    std::vector<DevRect> rects;
    std::vector<size_t> rect_blob_off;   // offset of each rectangle's record in the export blob (bf_scene_update_rect rewrites it)
    std::vector<HostShape> shapes;
    std::vector<M4> velocities;      // per shape (identity unless bf_scene_set_shape_velocity)
    std::vector<HostTx> txs;
    int32_t rx_shape = -1; bf_endpoint_desc rx; bool has_rx = false; size_t rx_blob_off = (size_t) -1;
    bf_adc_desc adc; bool has_adc = false;
    bf_integrator_desc integ; bool has_integ = false;
    std::vector<uint8_t> blob;        // recorded construction calls (flat_scene.h)
    std::string err;

    // device side
    bool committed = false, tables_dirty = false;
    bool tables_light = false;       // tables_dirty, but only rectangle poses and / or a non-phased receiver descriptor changed since the last upload
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;        // the stream every entry point enqueues on: own_stream, or the caller's (bf_scene_set_stream)
    cudaStream_t own_stream = nullptr;
    size_t l2_window_max = 0; bool l2_window_touched = false;     // tune.l2_persist_mb
    DevScene dev;
    DevBuf<uint4> d_nodes; DevBuf<float4> d_tris, d_tri_nrm;
    DevBuf<float4> d_tris_rest, d_nrm_rest; DevBuf<float> d_node_box; bool rest_saved = false;    // bf_scene_update_transform (refit)
    std::vector<uint32_t> bvh_level_first; float bvh_pad_abs = 0.f;
    DevBuf<DevRect> d_rects; DevBuf<DevShape> d_shapes; DevBuf<bf_bsdf_desc> d_bsdfs;
    DevBuf<DevSpectrum> d_spectra; DevBuf<float> d_spec_pool; DevBuf<DevEndpoint> d_txs, d_rx; DevBuf<DevVElem> d_velems; DevBuf<M4> d_velocities;
    DevBuf<float> d_profile; size_t profile_floats = 0;
    DevBuf<float> d_batch; uint32_t batch_items = 0;   // bf_receive_batch: one profile per item
    // ... small items in one k_mega_batch launch: per-item tables [MegaItem | DevEndpoint | DevRect x n_rects] x items + one cursor each
    DevBuf<uint8_t> d_items; uint8_t *h_items = nullptr; size_t h_items_bytes = 0; cudaEvent_t items_ev = nullptr;
    bool geometry_moved = false;     // the device geometry differs from the pinned image (bf_scene_reupload would undo a refit)
    bool primary_defer = false;      // fused primary: traverse the rays that enter the mesh bounds in a separate dense launch
    uint32_t auto_rounds = 8;        // k_trace batches per warp for the request being issued (tune.trace_rounds < 0)
    bool path_defer = false;         // same, at the (higher) share of rays below which k_path + dense traversal wins over k_primary
    DevFilter filter; bool wide_filter = false;    // discretised reconstruction filter when wider than one bin
    DevBuf<uint32_t> d_developed;    // bf_profile_develop output (fp32 or fp16 components)
    // image of the geometry tables in pinned host memory (re-upload without a rebuild)
    void *h_geom = nullptr; size_t h_geom_bytes = 0, h_nodes_bytes = 0, h_tris_bytes = 0, h_nrm_bytes = 0;
    void *h_tables = nullptr; size_t h_tables_bytes = 0;   // pinned staging of the small tables (upload_tables)
    // ring of pinned staging slots for the light uploads (rectangle poses + receiver descriptor): a slot is rewritten only
    // after the event recorded behind its copy has completed, so per-frame re-posing never synchronises the stream
    static const int RING = 32;
    void *h_ring = nullptr; size_t ring_slot_bytes = 0; int ring_next = 0; cudaEvent_t ring_ev[RING] = {};
    // wave
    std::vector<WaveSlot> slots;      // slot 0 runs on `stream`
    DevBuf<unsigned long long> d_totals;
    cudaEvent_t ev_fork = nullptr;
    void *nccl_comm = nullptr;        // bf_multi_init
    // stats / timing
    bf_stats stats;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    std::vector<cudaEvent_t> ev_trace; size_t ev_trace_used = 0; std::vector<uint8_t> ev_kind;
    bool time_trace = false;
};

static void multi_destroy_comm(void *comm);
namespace { struct NvtxRange { explicit NvtxRange(const char *name) { nvtxRangePushA(name); } ~NvtxRange() { nvtxRangePop(); } }; }
static thread_local std::string g_create_error;

static int fail(bf_scene *s, int code, const std::string &msg) {
    if (s) s->err = msg; else g_create_error = msg;
    return code;
}
#define BF_CUDA(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return fail(s, BF_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));           \
    } while (0)

static void record(bf_scene *s, uint32_t tag, const void *a, size_t na, const void *b = nullptr, size_t nb = 0,
                   const void *c = nullptr, size_t nc = 0, const void *d = nullptr, size_t nd = 0) {
    bf_flat_record r; r.tag = tag; r.reserved = 0; r.payload_bytes = na + nb + nc + nd;
    size_t off = s->blob.size();
    size_t padded = (r.payload_bytes + 7) & ~(size_t) 7;
    s->blob.resize(off + sizeof(r) + padded, 0);
    memcpy(s->blob.data() + off, &r, sizeof(r));
    uint8_t *p = s->blob.data() + off + sizeof(r);
    if (na) memcpy(p, a, na); p += na;
    if (nb) memcpy(p, b, nb); p += nb;
    if (nc) memcpy(p, c, nc); p += nc;
    if (nd) memcpy(p, d, nd);
}

// ReconstructionFilter::init_discretization (rfilter.cpp:9-20) for the ADC's filter; GaussianFilter::eval (gaussian.cpp:47-49),
// TentFilter::eval (tent.cpp:36-38).  Host arithmetic, fp32 as in the reference's scalar variants.
static void build_filter(bf_scene *s) {
    const bf_adc_desc &a = s->adc;
    s->wide_filter = s->has_adc && s->integ.gen == BF_GEN_PHASE && a.filter_radius > 0.5f + BF_RAY_EPS;
    if (!s->wide_filter) return;
    DevFilter &f = s->filter;
    const float radius = a.filter_radius;
    const float stddev = radius / 4.f, alpha = -1.f / (2.f * stddev * stddev), bias = std::exp(alpha * (radius * radius));
    // Mitchell-Netravali polynomial (mitchell.cpp:52-64; catmullrom.cpp:33-45 is the same with B = 0, C = 0.5)
    auto cubic = [](float x, float B, float C) {
        const float x2 = x * x, x3 = x2 * x;
        const float r = (1.f / 6.f) * (x < 1.f ? (12.f - 9.f * B - 6.f * C) * x3 + (-18.f + 12.f * B + 6.f * C) * x2 + (6.f - 2.f * B)
                                               : (-B - 6.f * C) * x3 + (6.f * B + 30.f * C) * x2 + (-12.f * B - 48.f * C) * x + (8.f * B + 24.f * C));
        return x < 2.f ? r : 0.f;
    };
    for (int i = 0; i < 31; ++i) {
        const float x = (radius * (float) i) / 31.f;
        float v = 0.f;
        switch (a.filter_kind) {
            case 1: v = std::max(0.f, std::exp(alpha * (x * x)) - bias); break;
            case 2: v = std::max(0.f, 1.f - std::fabs(x * (1.f / radius))); break;
            case 3: v = cubic(x, 1.f / 3.f, 1.f / 3.f); break;
            case 4: v = cubic(x, 0.f, .5f); break;
            case 5: {   // lanczos.cpp:47-55
                const float x1 = 3.14159265358979323846f * x, x2 = x1 / radius;
                v = x < 5.9604644775390625e-08f ? 1.f : (x > radius ? 0.f : (std::sin(x1) * std::sin(x2)) / (x1 * x2));
                break;
            }
        }
        f.values[i] = v;
    }
    f.values[31] = 0.f;
    f.radius = radius; f.scale = 31.f / radius;
    f.n = (uint32_t) (int) std::ceil((radius - 2.f * BF_RAY_EPS) * 2.f);
    f.border = (uint32_t) (int) std::ceil(radius - .5f - 2.f * BF_RAY_EPS);
}

// Which primary kernel: the share of receiver rays (cosine-weighted about the receiver rectangle's normal, from its centre)
// that enter the padded bounds of the meshes, on a 64 x 64 stratified grid.  Below a quarter, traversal is left to a
// separate dense launch (k_primary<DEFER>); results are the same either way.  Scenes k_path can run switch at one half:
// measured on a B200 (tools/bounds_fraction.py), C3 at 0.32 gains 15 % from k_path + dense k_trace, C5 at 1.0 loses 4 %.
static void estimate_primary_defer(bf_scene *s) {
    s->primary_defer = false; s->path_defer = false;
    if (!s->has_rx || s->rx_shape < 0 || (size_t) s->rx_shape >= s->shapes.size() || s->shapes[s->rx_shape].type != 1) return;
    const DevRect &r = s->rects[s->shapes[s->rx_shape].index];
    const float *lo = s->dev.mesh_lo, *hi = s->dev.mesh_hi;
    if (!(lo[0] <= hi[0])) return;
    const double o[3] = {r.to_world.m[0][3], r.to_world.m[1][3], r.to_world.m[2][3]};
    const double fs[3] = {r.ns.x, r.ns.y, r.ns.z}, ft[3] = {r.nt.x, r.nt.y, r.nt.z}, fn[3] = {r.nn.x, r.nn.y, r.nn.z};
    int inside = 0;
    const int N = 64;
    for (int iy = 0; iy < N; ++iy)
        for (int ix = 0; ix < N; ++ix) {
            // cosine-weighted direction: uniform disk (polar form is fine for an estimate) lifted to the hemisphere
            const double u = (ix + 0.5) / N, v = (iy + 0.5) / N, rad = std::sqrt(u), phi = 6.283185307179586 * v;
            const double x = rad * std::cos(phi), y = rad * std::sin(phi), z = std::sqrt(std::max(0.0, 1.0 - u));
            double t0 = 0.0, t1 = 1e30;
            for (int a = 0; a < 3; ++a) {
                const double d = x * fs[a] + y * ft[a] + z * fn[a];
                if (std::fabs(d) < 1e-12) { if (o[a] < lo[a] || o[a] > hi[a]) { t1 = -1.0; break; } continue; }
                const double ta = (lo[a] - o[a]) / d, tb = (hi[a] - o[a]) / d;
                t0 = std::max(t0, std::min(ta, tb)); t1 = std::min(t1, std::max(ta, tb));
            }
            inside += t0 <= t1;
        }
    s->stats.mesh_bounds_fraction = (double) inside / (double) (N * N);
    s->primary_defer = inside * 4 < N * N;
    s->path_defer = inside * 2 < N * N;
}

static uint32_t n_channels(const bf_scene *s) {
    if (s->integ.gen == BF_GEN_PHASE) return 3 + s->integ.phase_bins;
    if (s->integ.gen == BF_GEN_RANGE) return 5 + s->integ.range_bins;
    return 5 + 50 * 3;
}
static void film_size(const bf_scene *s, uint32_t &w, uint32_t &h) {
    if (s->integ.gen == BF_GEN_PHASE) { w = s->adc.window_t_bins; h = s->adc.window_f_bins; }
    else { w = s->rx.film_width ? s->rx.film_width : 1; h = s->rx.film_height ? s->rx.film_height : 1; }
}

static DevEndpoint make_endpoint(const bf_endpoint_desc &d, int32_t shape) {
    DevEndpoint e;
    memset(&e, 0, sizeof(e));
    e.kind = d.kind; e.texture = d.texture; e.receive_type = d.receive_type; e.shape = shape;
    e.signal = d.signal;
    e.sig_period = rcp(d.signal.repfreq); e.sig_ti = 0 + d.signal.t_ext / 2; e.sig_slope = d.signal.f_ext / d.signal.t_ext;
    e.adc_sampling_start = d.adc_sampling_start; e.adc_sampling_time = d.adc_sampling_time;
    e.near_clip = d.near_clip; e.far_clip = d.far_clip;
    e.shutter_open = d.shutter_open; e.shutter_open_time = d.shutter_open_time;
    e.film_width = d.film_width; e.film_height = d.film_height;
    memcpy(e.to_world.m, d.to_world, 64);
    bool identity_ok = false;
    for (int i = 0; i < 16; ++i) if (d.to_world[i] != 0.f) identity_ok = true;
    if (!identity_ok) { for (int i = 0; i < 4; ++i) e.to_world.m[i][i] = 1.f; }
    e.to_object = m4_inverse(e.to_world);
    memcpy(e.sample_to_camera.m, d.sample_to_camera, 64);
    // spot.cpp:85-93
    float cutoff = d.cutoff_angle_deg * (BF_PI / 180.f), beam = d.beam_width_deg * (BF_PI / 180.f);
    e.cutoff = cutoff;
    e.inv_transition = 1.0f / (cutoff - beam);
    e.cos_cutoff = cosf(cutoff);
    e.cos_beam = cosf(beam);
    return e;
}

extern "C" {

uint32_t bf_abi_version(void) { return BF_ABI_VERSION; }

int bf_scene_create(const bf_constants *consts, bf_scene **out) {
    if (!consts || !out) return fail(nullptr, BF_ERR_INVALID, "bf_scene_create: null argument");
    if (consts->spectrum_width != 1 && consts->spectrum_width != 4)
        return fail(nullptr, BF_ERR_UNSUPPORTED, "bf_scene_create: spectrum_width must be 1 (mono) or 4 (spectral)");
    if (!(consts->c > 0.f)) return fail(nullptr, BF_ERR_INVALID, "bf_scene_create: propagation speed must be positive");
    bf_scene *s = new bf_scene();
    s->C = *consts;
    s->tune = tunables_from_env();
    memset(&s->rx, 0, sizeof(s->rx)); memset(&s->adc, 0, sizeof(s->adc)); memset(&s->integ, 0, sizeof(s->integ));
    memset(&s->stats, 0, sizeof(s->stats)); memset(&s->dev, 0, sizeof(s->dev));
    *out = s;
    return BF_OK;
}

static void free_wave(bf_scene *s);
static void apply_l2_persistence(bf_scene *s);

void bf_scene_destroy(bf_scene *s) {
    if (!s) return;
    if (s->committed) {
        cudaSetDevice(s->device);
        if (s->stream) cudaStreamSynchronize(s->stream);
        if (s->nccl_comm) multi_destroy_comm(s->nccl_comm);
        free_wave(s);
        s->d_nodes.release(); s->d_tris.release(); s->d_tri_nrm.release(); s->d_tris_rest.release(); s->d_nrm_rest.release(); s->d_node_box.release(); s->d_rects.release(); s->d_shapes.release();
        s->d_bsdfs.release(); s->d_spectra.release(); s->d_spec_pool.release(); s->d_txs.release(); s->d_rx.release(); s->d_velems.release(); s->d_velocities.release();
        s->d_profile.release(); s->d_batch.release(); s->d_items.release();
        if (s->h_items) { cudaFreeHost(s->h_items); s->h_items = nullptr; s->h_items_bytes = 0; }
        if (s->items_ev) { cudaEventDestroy(s->items_ev); s->items_ev = nullptr; } s->d_developed.release(); s->d_totals.release();
        if (s->h_geom) cudaFreeHost(s->h_geom);
        if (s->h_tables) cudaFreeHost(s->h_tables);
        if (s->h_ring) cudaFreeHost(s->h_ring);
        for (cudaEvent_t e : s->ring_ev) if (e) cudaEventDestroy(e);
        if (s->ev_fork) cudaEventDestroy(s->ev_fork);
        if (s->ev_begin) cudaEventDestroy(s->ev_begin);
        if (s->ev_end) cudaEventDestroy(s->ev_end);
        for (cudaEvent_t e : s->ev_trace) cudaEventDestroy(e);
        if (s->own_stream) cudaStreamDestroy(s->own_stream);
    }
    delete s;
}

const char *bf_last_error(const bf_scene *s) { return s ? s->err.c_str() : g_create_error.c_str(); }

int bf_scene_add_spectrum(bf_scene *s, uint32_t kind, uint32_t n, const float *lambda_nm, const float *values, uint32_t *spectrum_id) {
    if (!s || !values || n == 0) return fail(s, BF_ERR_INVALID, "bf_scene_add_spectrum: null / empty");
    if (kind > BF_SPEC_ZERO_OUTSIDE_VISIBLE) return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_add_spectrum: unknown kind");
    if ((kind == BF_SPEC_REGULAR || kind == BF_SPEC_IRREGULAR) && (n < 2 || !lambda_nm))
        return fail(s, BF_ERR_INVALID, "bf_scene_add_spectrum: tabulated spectra need >= 2 nodes with wavelengths");
    HostSpectrum t; t.kind = kind;
    t.value.assign(values, values + n);
    if (lambda_nm) t.lambda.assign(lambda_nm, lambda_nm + n); else t.lambda.assign(n, 0.f);
    if (kind == BF_SPEC_IRREGULAR)
        for (uint32_t i = 1; i < n; ++i)
            if (!(t.lambda[i] > t.lambda[i - 1])) return fail(s, BF_ERR_INVALID, "bf_scene_add_spectrum: wavelengths must increase");
    uint32_t hdr[2] = {kind, n};
    record(s, BF_TAG_SPECTRUM, hdr, 8, t.lambda.data(), 4 * (size_t) n, t.value.data(), 4 * (size_t) n);
    if (spectrum_id) *spectrum_id = (uint32_t) s->spectra.size();
    s->spectra.push_back(std::move(t));
    s->tables_dirty = true; s->tables_light = false;
    return BF_OK;
}

int bf_scene_add_bsdf(bf_scene *s, const bf_bsdf_desc *desc, uint32_t *bsdf_id) {
    if (!s || !desc) return fail(s, BF_ERR_INVALID, "bf_scene_add_bsdf: null");
    if (desc->kind > BF_BSDF_NULL_BLACK) return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_add_bsdf: BSDF plugin outside the radar path (supported: diffuse, roughconductor, conductor, twosided)");
    auto spec_ok = [&](uint32_t id) { return id < s->spectra.size(); };
    if (desc->kind == BF_BSDF_DIFFUSE && !spec_ok(desc->reflectance)) return fail(s, BF_ERR_INVALID, "bf_scene_add_bsdf: reflectance spectrum id out of range");
    if ((desc->kind == BF_BSDF_ROUGHCONDUCTOR || desc->kind == BF_BSDF_CONDUCTOR) && (!spec_ok(desc->eta) || !spec_ok(desc->k)))
        return fail(s, BF_ERR_INVALID, "bf_scene_add_bsdf: eta / k spectrum id out of range");
    if (desc->specular_reflectance != BF_NONE && !spec_ok(desc->specular_reflectance)) return fail(s, BF_ERR_INVALID, "bf_scene_add_bsdf: specular_reflectance id out of range");
    if (desc->back_bsdf != BF_NONE && desc->back_bsdf >= s->bsdfs.size()) return fail(s, BF_ERR_INVALID, "bf_scene_add_bsdf: back_bsdf id out of range");
    record(s, BF_TAG_BSDF, desc, sizeof(*desc));
    if (bsdf_id) *bsdf_id = (uint32_t) s->bsdfs.size();
    s->bsdfs.push_back(*desc);
    s->tables_dirty = true; s->tables_light = false;
    return BF_OK;
}

int bf_scene_add_mesh(bf_scene *s, const float *positions, uint32_t n_vertices, const uint32_t *faces, uint32_t n_faces,
                      const float *normals, uint32_t bsdf_id, uint32_t *shape_id) {
    if (!s || !positions || !faces) return fail(s, BF_ERR_INVALID, "bf_scene_add_mesh: null");
    if (bsdf_id >= s->bsdfs.size()) return fail(s, BF_ERR_INVALID, "bf_scene_add_mesh: bsdf id out of range");
    if (s->committed) return fail(s, BF_ERR_STATE, "bf_scene_add_mesh: scene already committed");
    for (size_t i = 0; i < 3 * (size_t) n_faces; ++i)
        if (faces[i] >= n_vertices) return fail(s, BF_ERR_INVALID, "bf_scene_add_mesh: face index out of range");
    HostMesh m; m.nv = n_vertices; m.nf = n_faces;
    m.pos.assign(positions, positions + 3 * (size_t) n_vertices);
    m.idx.assign(faces, faces + 3 * (size_t) n_faces);
    if (normals) m.nrm.assign(normals, normals + 3 * (size_t) n_vertices);
    for (int a = 0; a < 3; ++a) { m.rest_lo[a] = INFINITY; m.rest_hi[a] = -INFINITY; }
    for (uint32_t v = 0; v < n_vertices; ++v)
        for (int a = 0; a < 3; ++a) { m.rest_lo[a] = std::min(m.rest_lo[a], positions[3 * (size_t) v + a]); m.rest_hi[a] = std::max(m.rest_hi[a], positions[3 * (size_t) v + a]); }
    uint32_t hdr[4] = {n_vertices, n_faces, bsdf_id, normals ? 1u : 0u};
    record(s, BF_TAG_MESH, hdr, 16, m.pos.data(), 4 * m.pos.size(), m.idx.data(), 4 * m.idx.size(), normals ? m.nrm.data() : nullptr, normals ? 4 * m.nrm.size() : 0);
    HostShape sh; sh.type = 0; sh.index = (uint32_t) s->meshes.size(); sh.bsdf = bsdf_id; sh.tx = -1;
    if (shape_id) *shape_id = (uint32_t) s->shapes.size();
    s->meshes.push_back(std::move(m));
    s->shapes.push_back(sh);
    return BF_OK;
}

int bf_scene_add_rect(bf_scene *s, const float to_world[16], uint32_t bsdf_id, uint32_t *shape_id) {
    if (!s || !to_world) return fail(s, BF_ERR_INVALID, "bf_scene_add_rect: null");
    if (bsdf_id >= s->bsdfs.size()) return fail(s, BF_ERR_INVALID, "bf_scene_add_rect: bsdf id out of range");
    if (s->committed) return fail(s, BF_ERR_STATE, "bf_scene_add_rect: scene already committed (use bf_scene_update_rect)");
    if (s->rects.size() >= BF_MAX_RECTS) return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_add_rect: more than 64 analytic rectangles");
    uint32_t tail[2] = {bsdf_id, 0};
    s->rect_blob_off.push_back(s->blob.size());
    record(s, BF_TAG_RECT, to_world, 64, tail, 8);
    HostShape sh; sh.type = 1; sh.index = (uint32_t) s->rects.size(); sh.bsdf = bsdf_id; sh.tx = -1;
    uint32_t id = (uint32_t) s->shapes.size();
    s->rects.push_back(make_rect(to_world, id));
    s->shapes.push_back(sh);
    if (shape_id) *shape_id = id;
    return BF_OK;
}

int bf_scene_set_shape_velocity(bf_scene *s, uint32_t shape_id, const float velocity[16]) {
    if (!s || !velocity) return BF_ERR_INVALID;
    if (shape_id >= s->shapes.size()) return fail(s, BF_ERR_INVALID, "bf_scene_set_shape_velocity: shape id out of range");
    s->velocities.resize(s->shapes.size(), m4_identity());
    memcpy(s->velocities[shape_id].m, velocity, 64);
    uint32_t hdr[2] = {shape_id, 0};
    record(s, BF_TAG_VELOCITY, hdr, 8, velocity, 64);
    s->tables_dirty = true; s->tables_light = false;
    return BF_OK;
}

static int check_endpoint(bf_scene *s, uint32_t shape_id, const bf_endpoint_desc *d, bool is_tx) {
    if (!d) return fail(s, BF_ERR_INVALID, "endpoint: null descriptor");
    bool free_standing = is_tx ? d->kind == BF_TX_SPOT : d->kind == BF_RX_PERSPECTIVE;
    if ((is_tx && d->kind == BF_TX_PHASED) || (!is_tx && d->kind == BF_RX_PHASED)) {
        if (d->array.n_elems < 1 || d->array.n_elems > BF_MAX_ARRAY_ELEMS) return fail(s, BF_ERR_INVALID, "phased endpoint: n_elems must be in [1, 64]");
        if (d->array.elem_dims[0] == 0.f || d->array.elem_dims[1] == 0.f || d->array.elem_dims[2] == 0.f)
            return fail(s, BF_ERR_INVALID, "phased endpoint: elem_dims has a zero component (scale() of the virtual element is singular)");
    }
    if (is_tx && d->kind > BF_TX_AREALIGHT) return fail(s, BF_ERR_UNSUPPORTED, "unknown transmitter kind");
    if (!is_tx && d->kind > BF_RX_PERSPECTIVE) return fail(s, BF_ERR_UNSUPPORTED, "unknown receiver kind");
    if (!free_standing) {
        if (shape_id >= s->shapes.size()) return fail(s, BF_ERR_INVALID, "endpoint: shape id out of range");
        if (s->shapes[shape_id].type != 1) return fail(s, BF_ERR_UNSUPPORTED, "endpoint: only rectangle apertures are supported (every radar scene of the reference uses rectangles)");
    }
    if (is_tx && d->texture >= s->spectra.size()) return fail(s, BF_ERR_INVALID, "endpoint: texture spectrum id out of range");
    return BF_OK;
}

int bf_scene_add_transmitter(bf_scene *s, uint32_t shape_id, const bf_endpoint_desc *desc, uint32_t *transmitter_id) {
    if (!s) return BF_ERR_INVALID;
    int rc = check_endpoint(s, shape_id, desc, true);
    if (rc) return rc;
    HostTx t; t.shape = desc->kind == BF_TX_SPOT ? -1 : (int32_t) shape_id; t.d = *desc;
    uint32_t hdr[2] = {t.shape < 0 ? BF_NONE : shape_id, 0};
    record(s, BF_TAG_TX, hdr, 8, desc, sizeof(*desc));
    if (t.shape >= 0) s->shapes[shape_id].tx = (int32_t) s->txs.size();
    if (transmitter_id) *transmitter_id = (uint32_t) s->txs.size();
    s->txs.push_back(t);
    s->tables_dirty = true; s->tables_light = false;
    return BF_OK;
}

int bf_scene_set_receiver(bf_scene *s, uint32_t shape_id, const bf_endpoint_desc *desc) {
    if (!s) return BF_ERR_INVALID;
    int rc = check_endpoint(s, shape_id, desc, false);
    if (rc) return rc;
    if (s->committed && s->integ.gen != BF_GEN_PHASE) {
        // gen 1/2: the film size comes from the sensor; the device profile was sized at commit (a larger film would scatter out of bounds)
        const uint32_t w0 = s->rx.film_width ? s->rx.film_width : 1, h0 = s->rx.film_height ? s->rx.film_height : 1;
        const uint32_t w1 = desc->film_width ? desc->film_width : 1, h1 = desc->film_height ? desc->film_height : 1;
        if (w0 != w1 || h0 != h1) return fail(s, BF_ERR_STATE, "bf_scene_set_receiver: the film size cannot change after bf_scene_commit");
    }
    const bool light = s->committed && s->has_rx && s->rx.kind == desc->kind && desc->kind != BF_RX_PHASED && s->rx.texture == desc->texture;
    s->rx = *desc; s->has_rx = true;
    s->rx_shape = desc->kind == BF_RX_PERSPECTIVE ? -1 : (int32_t) shape_id;
    uint32_t hdr[2] = {s->rx_shape < 0 ? BF_NONE : shape_id, 0};
    if (s->rx_blob_off != (size_t) -1) {   // re-posed receiver: rewrite its record in place, the export stays one RX record long
        uint8_t *p = s->blob.data() + s->rx_blob_off + sizeof(bf_flat_record);
        memcpy(p, hdr, 8); memcpy(p + 8, desc, sizeof(*desc));
    } else {
        s->rx_blob_off = s->blob.size();
        record(s, BF_TAG_RX, hdr, 8, desc, sizeof(*desc));
    }
    if (light) { if (!s->tables_dirty) s->tables_light = true; } else s->tables_light = false;
    s->tables_dirty = true;
    return BF_OK;
}

int bf_scene_set_adc(bf_scene *s, const bf_adc_desc *desc) {
    if (!s || !desc) return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: null");
    if (desc->t_bins == 0 || desc->f_bins == 0 || desc->window_t_bins == 0 || desc->window_f_bins == 0)
        return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: zero-sized ADC");
    if (!(desc->t_bandwidth > 0.f) || !(desc->f_bandwidth > 0.f)) return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: bandwidths must be positive");
    // Receivers warn when the reconstruction filter is wider than one bin (phasedreceiver.cpp:59-62);
    // the accelerated path implements the single-bin branch of SignalBlock::put (signalblock.cpp:162-169).
    // Wider filters take the n x n splat (signalblock.cpp:113-161): gaussian (radius = 4 stddev, gaussian.cpp:37-45), tent
    // (radius 1, tent.cpp:30-34), mitchell (B = C = 1/3) and catmullrom (radius 2), lanczos (radius = lobes).
    if (desc->filter_radius > 0.5f + BF_RAY_EPS) {
        if (desc->filter_kind == 1) { if (!(desc->filter_radius <= 8.f)) return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: gaussian filter radius (4 stddev) must be <= 8 bins"); }
        else if (desc->filter_kind == 2) { if (desc->filter_radius != 1.f) return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: the tent filter has radius 1"); }
        else if (desc->filter_kind == 3 || desc->filter_kind == 4) { if (desc->filter_radius != 2.f) return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: the mitchell / catmullrom filters have radius 2"); }
        else if (desc->filter_kind == 5) { if (!(desc->filter_radius >= 1.f && desc->filter_radius <= 8.f && desc->filter_radius == std::floor(desc->filter_radius))) return fail(s, BF_ERR_INVALID, "bf_scene_set_adc: lanczos radius = lobes, an integer in [1, 8]"); }
        else return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_set_adc: unknown wide reconstruction filter kind (1 gaussian, 2 tent, 3 mitchell, 4 catmullrom, 5 lanczos)");
    }
    if (s->committed) return fail(s, BF_ERR_STATE, "bf_scene_set_adc: scene already committed");
    s->adc = *desc; s->has_adc = true;
    record(s, BF_TAG_ADC, desc, sizeof(*desc));
    return BF_OK;
}

int bf_scene_set_integrator(bf_scene *s, const bf_integrator_desc *desc) {
    if (!s || !desc) return fail(s, BF_ERR_INVALID, "bf_scene_set_integrator: null");
    if (desc->gen < BF_GEN_TIME || desc->gen > BF_GEN_PHASE) return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_set_integrator: gen must be 1 (time), 2 (range) or 3 (phase)");
    if (desc->rng_mode != 0) return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_set_integrator: only per-sample PCG32 streams (rng_mode 0) run on the GPU");
    if (desc->max_depth == 0 || desc->max_depth < -1) return fail(s, BF_ERR_INVALID, "bf_scene_set_integrator: max_depth must be -1 or > 0 (integrator.cpp:1713-1720)");
    if (desc->gen == BF_GEN_RANGE && !(desc->dr > 0.f)) return fail(s, BF_ERR_INVALID, "bf_scene_set_integrator: range integrator needs dr > 0");
    if (s->committed) return fail(s, BF_ERR_STATE, "bf_scene_set_integrator: scene already committed");
    s->integ = *desc; s->has_integ = true;
    record(s, BF_TAG_INTEGRATOR, desc, sizeof(*desc));
    return BF_OK;
}

int bf_profile_size(const bf_scene *s, uint32_t *width, uint32_t *height, uint32_t *channels) {
    if (!s || !s->has_integ || !s->has_rx) return BF_ERR_STATE;
    if (s->integ.gen == BF_GEN_PHASE && !s->has_adc) return BF_ERR_STATE;
    uint32_t w, h; film_size(s, w, h);
    if (width) *width = w;
    if (height) *height = h;
    if (channels) *channels = n_channels(s);
    return BF_OK;
}
int bf_profile_channels(const bf_scene *s, uint32_t *channels) { return bf_profile_size(s, nullptr, nullptr, channels); }

int bf_scene_export(const bf_scene *s, void *dst, size_t capacity, size_t *required) {
    if (!s) return BF_ERR_INVALID;
    size_t total = sizeof(bf_flat_header) + s->blob.size();
    if (required) *required = total;
    if (!dst) return BF_OK;
    if (capacity < total) return BF_ERR_INVALID;
    bf_flat_header h; h.magic = BF_FLAT_MAGIC; h.abi_version = BF_ABI_VERSION; h.total_bytes = total; h.consts = s->C;
    memcpy(dst, &h, sizeof(h));
    if (!s->blob.empty()) memcpy((uint8_t *) dst + sizeof(h), s->blob.data(), s->blob.size());
    return BF_OK;
}

// ---- upload ---------------------------------------------------------------------------------------------
static int upload_tables(bf_scene *s) {
    // spectra
    std::vector<DevSpectrum> ds; std::vector<float> pool;
    for (const HostSpectrum &t : s->spectra) {
        DevSpectrum d; d.kind = t.kind; d.n = (uint32_t) t.value.size(); d.offset = (uint32_t) pool.size(); d.inv_interval = 0.f;
        if (t.kind == BF_SPEC_REGULAR && d.n >= 2) {
            double interval = ((double) t.lambda.back() - (double) t.lambda.front()) / (double) (d.n - 1);
            d.inv_interval = (float) (1.0 / interval);
        }
        pool.insert(pool.end(), t.lambda.begin(), t.lambda.end());
        pool.insert(pool.end(), t.value.begin(), t.value.end());
        ds.push_back(d);
    }
    std::vector<DevShape> sh(s->shapes.size());
    int32_t cls1_kind = -2;   // -2: no class-1 shape seen yet
    for (size_t i = 0; i < sh.size(); ++i) {
        memset(&sh[i], 0, sizeof(DevShape));
        sh[i].type = s->shapes[i].type; sh[i].index = s->shapes[i].index; sh[i].bsdf = s->shapes[i].bsdf; sh[i].tx = s->shapes[i].tx;
        sh[i].has_normals = s->shapes[i].type == 0 && !s->meshes[s->shapes[i].index].nrm.empty();
        const bf_bsdf_desc &bd = s->bsdfs[s->shapes[i].bsdf];
        const uint32_t back_kind = (bd.two_sided && bd.back_bsdf != BF_NONE) ? s->bsdfs[bd.back_bsdf].kind : bd.kind;
        sh[i].mat_class = (bd.kind == BF_BSDF_DIFFUSE && back_kind == BF_BSDF_DIFFUSE) ? 0u : 1u;
        if (sh[i].mat_class == 1u) {
            const int32_t k = (bd.kind == back_kind) ? (int32_t) bd.kind : -1;
            cls1_kind = cls1_kind == -2 ? k : (cls1_kind == k ? k : -1);
        }
    }
    std::vector<DevEndpoint> txs;
    std::vector<DevVElem> velems;
    auto with_array = [&](DevEndpoint e, const bf_endpoint_desc &d, bool phased) {
        if (phased) {
            e.velem_offset = (uint32_t) velems.size();
            make_velems(d.array, s->C, velems);
            e.n_velems = (uint32_t) velems.size() - e.velem_offset;
            e.wid_x = d.array.elem_dims[0]; e.wid_y = d.array.elem_dims[1];
        }
        return e;
    };
    for (const HostTx &t : s->txs) txs.push_back(with_array(make_endpoint(t.d, t.shape), t.d, t.d.kind == BF_TX_PHASED));
    DevEndpoint rx = with_array(make_endpoint(s->rx, s->rx_shape), s->rx, s->has_rx && s->rx.kind == BF_RX_PHASED);

    BF_CUDA(s->d_spectra.alloc(ds.size())); BF_CUDA(s->d_spec_pool.alloc(pool.size()));
    BF_CUDA(s->d_shapes.alloc(sh.size())); BF_CUDA(s->d_bsdfs.alloc(s->bsdfs.size()));
    BF_CUDA(s->d_rects.alloc(s->rects.size())); BF_CUDA(s->d_txs.alloc(txs.size())); BF_CUDA(s->d_rx.alloc(1));
    BF_CUDA(s->d_velems.alloc(velems.size()));
    s->velocities.resize(s->shapes.size(), m4_identity());
    BF_CUDA(s->d_velocities.alloc(s->velocities.size()));
    // The tables are packed into ONE pinned staging buffer and copied with cudaMemcpyAsync on the scene's stream, so the
    // copies are ordered before every kernel of the next bf_receive (the wave streams fork from that stream).  The
    // synchronize keeps the staging buffer from being rewritten while an earlier upload is still in flight (every wave
    // stream has been joined back into s->stream by the end of bf_receive).
    BF_CUDA(cudaStreamSynchronize(s->stream));
    struct Part { void *dst; const void *src; size_t bytes; };
    const Part parts[] = {
        {s->d_velems.p, velems.data(), velems.size() * sizeof(DevVElem)}, {s->d_velocities.p, s->velocities.data(), s->velocities.size() * sizeof(M4)}, {s->d_spectra.p, ds.data(), ds.size() * sizeof(DevSpectrum)},
        {s->d_spec_pool.p, pool.data(), pool.size() * 4}, {s->d_shapes.p, sh.data(), sh.size() * sizeof(DevShape)},
        {s->d_bsdfs.p, s->bsdfs.data(), s->bsdfs.size() * sizeof(bf_bsdf_desc)}, {s->d_rects.p, s->rects.data(), s->rects.size() * sizeof(DevRect)},
        {s->d_txs.p, txs.data(), txs.size() * sizeof(DevEndpoint)}, {s->d_rx.p, &rx, sizeof(rx)}};
    size_t total = 0;
    for (const Part &pt : parts) total += (pt.bytes + 15) & ~(size_t) 15;
    if (total > s->h_tables_bytes) {
        if (s->h_tables) cudaFreeHost(s->h_tables);
        if (s->h_ring) cudaFreeHost(s->h_ring);
        for (cudaEvent_t e : s->ring_ev) if (e) cudaEventDestroy(e);
        s->h_tables = nullptr; s->h_tables_bytes = 0;
        BF_CUDA(cudaMallocHost(&s->h_tables, total * 2));
        s->h_tables_bytes = total * 2;
    }
    size_t off = 0;
    for (const Part &pt : parts) {
        if (pt.bytes == 0) continue;
        memcpy((uint8_t *) s->h_tables + off, pt.src, pt.bytes);
        BF_CUDA(cudaMemcpyAsync(pt.dst, (uint8_t *) s->h_tables + off, pt.bytes, cudaMemcpyHostToDevice, s->stream));
        off += (pt.bytes + 15) & ~(size_t) 15;
    }

    DevScene &d = s->dev;
    d.C = s->C; d.S = s->C.spectrum_width;
    d.n_rects = (uint32_t) s->rects.size(); d.n_shapes = (uint32_t) sh.size(); d.n_txs = (uint32_t) txs.size();
    d.rects = s->d_rects.p; d.shapes = s->d_shapes.p; d.bsdfs = s->d_bsdfs.p; d.spectra = s->d_spectra.p;
    d.spec_pool = s->d_spec_pool.p; d.txs = s->d_txs.p; d.rx = s->d_rx.p; d.velems = s->d_velems.p; d.velocities = s->d_velocities.p;
    d.adc = s->adc; d.integ = s->integ;
    build_filter(s);
    estimate_primary_defer(s);
    film_size(s, d.film_w, d.film_h); d.channels = n_channels(s);
    d.rx_needs_aperture = s->rx.kind != BF_RX_PERSPECTIVE;   // Sensor::needs_aperture_sample
    d.hint_cls1_bsdf = cls1_kind < 0 ? -1 : cls1_kind;
    d.hint_rx_kind = (int32_t) s->rx.kind;
    d.hint_tx_kind = txs.empty() ? -1 : (int32_t) txs[0].kind;
    for (const DevEndpoint &e : txs) if ((int32_t) e.kind != d.hint_tx_kind) d.hint_tx_kind = -1;
    d.phase_denom = (double) ((s->C.lambda_max_nm - s->C.lambda_min_nm) / 2) * 1e-9;
    d.adc_scale_t = (float) s->adc.t_bins / s->adc.t_bandwidth; d.adc_scale_f = (float) s->adc.f_bins / s->adc.f_bandwidth;
    d.rx_ep = rx;
    memset(&d.rx_rect, 0, sizeof(DevRect)); memset(&d.tx0_ep, 0, sizeof(DevEndpoint)); memset(&d.tx0_rect, 0, sizeof(DevRect));
    if (s->rx_shape >= 0) d.rx_rect = s->rects[s->shapes[s->rx_shape].index];
    if (!txs.empty()) {
        d.tx0_ep = txs[0];
        if (txs[0].shape >= 0) d.tx0_rect = s->rects[s->shapes[txs[0].shape].index];
    }
    s->tables_dirty = false; s->tables_light = false;
    return BF_OK;
}

// Light upload: only rectangle poses and / or the (non-phased) receiver descriptor changed since the last upload (the
// per-frame re-posing of C4 / C5, animated_trans_rad.py:305-355).  Re-stages the rectangle table and the receiver endpoint
// through the pinned ring and refreshes the copies in the kernel parameter block; nothing else of the scene is touched.
static int upload_tables_light(bf_scene *s, WaveSlot *into = nullptr) {
    DevScene &d = s->dev;
    const size_t rect_bytes = s->rects.size() * sizeof(DevRect);
    const size_t need = ((rect_bytes + 15) & ~(size_t) 15) + sizeof(DevEndpoint);
    if (!s->h_ring || need > s->ring_slot_bytes) {
        BF_CUDA(cudaDeviceSynchronize());
        if (s->h_ring) cudaFreeHost(s->h_ring);
        s->h_ring = nullptr;
        BF_CUDA(cudaMallocHost(&s->h_ring, need * 2 * bf_scene::RING));
        s->ring_slot_bytes = need * 2;
        for (int i = 0; i < bf_scene::RING; ++i) if (!s->ring_ev[i]) BF_CUDA(cudaEventCreateWithFlags(&s->ring_ev[i], cudaEventDisableTiming));
    }
    // destination: the scene's tables on the scene's stream, or (bf_receive_batch) a wave slot's private copies on its stream
    DevRect *dst_rects = s->d_rects.p; DevEndpoint *dst_rx = s->d_rx.p; cudaStream_t stream = s->stream;
    if (into) {
        if (into->n_rects < s->rects.size() || !into->d_rx) {
            BF_CUDA(cudaStreamSynchronize(into->stream));
            if (into->d_rects) { cudaFree(into->d_rects); into->allocs.erase(std::remove(into->allocs.begin(), into->allocs.end(), (void *) into->d_rects), into->allocs.end()); }
            if (into->d_rx) { cudaFree(into->d_rx); into->allocs.erase(std::remove(into->allocs.begin(), into->allocs.end(), (void *) into->d_rx), into->allocs.end()); }
            BF_CUDA(cudaMalloc((void **) &into->d_rects, std::max<size_t>(rect_bytes, 16))); into->allocs.push_back(into->d_rects);
            BF_CUDA(cudaMalloc((void **) &into->d_rx, sizeof(DevEndpoint))); into->allocs.push_back(into->d_rx);
            into->n_rects = s->rects.size();
        }
        dst_rects = into->d_rects; dst_rx = into->d_rx; stream = into->stream;
    }
    const int slot = s->ring_next; s->ring_next = (slot + 1) % bf_scene::RING;
    BF_CUDA(cudaEventSynchronize(s->ring_ev[slot]));      // completed long ago unless RING uploads are in flight
    uint8_t *h = (uint8_t *) s->h_ring + (size_t) slot * s->ring_slot_bytes;
    DevEndpoint rx = make_endpoint(s->rx, s->rx_shape);
    if (rect_bytes) {
        memcpy(h, s->rects.data(), rect_bytes);
        BF_CUDA(cudaMemcpyAsync(dst_rects, h, rect_bytes, cudaMemcpyHostToDevice, stream));
    }
    memcpy(h + ((rect_bytes + 15) & ~(size_t) 15), &rx, sizeof(rx));
    BF_CUDA(cudaMemcpyAsync(dst_rx, h + ((rect_bytes + 15) & ~(size_t) 15), sizeof(rx), cudaMemcpyHostToDevice, stream));
    BF_CUDA(cudaEventRecord(s->ring_ev[slot], stream));
    if (d.n_nodes != 0) estimate_primary_defer(s);
    d.rx_needs_aperture = s->rx.kind != BF_RX_PERSPECTIVE;
    d.rx_ep = rx;
    memset(&d.rx_rect, 0, sizeof(DevRect));
    if (s->rx_shape >= 0) d.rx_rect = s->rects[s->shapes[s->rx_shape].index];
    if (d.n_txs != 0 && d.tx0_ep.shape >= 0) d.tx0_rect = s->rects[s->shapes[d.tx0_ep.shape].index];
    // a slot-private upload leaves the scene's own device tables stale: they are refreshed before the next plain bf_receive
    s->tables_dirty = into != nullptr; s->tables_light = into != nullptr;
    return BF_OK;
}

// Copy the pinned image of the geometry tables (nodes, triangles, normals) to the device.
static int upload_geometry(bf_scene *s) {
    const uint8_t *h = (const uint8_t *) s->h_geom;
    if (s->h_nodes_bytes) BF_CUDA(cudaMemcpyAsync(s->d_nodes.p, h, s->h_nodes_bytes, cudaMemcpyHostToDevice, s->stream));
    if (s->h_tris_bytes) BF_CUDA(cudaMemcpyAsync(s->d_tris.p, h + s->h_nodes_bytes, s->h_tris_bytes, cudaMemcpyHostToDevice, s->stream));
    if (s->h_nrm_bytes) BF_CUDA(cudaMemcpyAsync(s->d_tri_nrm.p, h + s->h_nodes_bytes + s->h_tris_bytes, s->h_nrm_bytes, cudaMemcpyHostToDevice, s->stream));
    return BF_OK;
}

int bf_scene_commit(bf_scene *s, int device) {
    if (!s) return BF_ERR_INVALID;
    NvtxRange range("bf_scene_commit (SAH build + upload)");
    if (!s->has_rx) return fail(s, BF_ERR_STATE, "bf_scene_commit: no receiver / sensor set");
    if (!s->has_integ) return fail(s, BF_ERR_STATE, "bf_scene_commit: no integrator set");
    if (s->integ.gen == BF_GEN_PHASE && !s->has_adc) return fail(s, BF_ERR_STATE, "bf_scene_commit: phase integrator needs an ADC");
    if (s->committed) return fail(s, BF_ERR_STATE, "bf_scene_commit: already committed");
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return fail(s, BF_ERR_CUDA, std::string("bf_scene_commit: no CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback");
    if (device < 0 || device >= n_dev) return fail(s, BF_ERR_INVALID, "bf_scene_commit: device index out of range");
    BF_CUDA(cudaSetDevice(device));
    s->device = device;
    cudaDeviceProp prop;
    BF_CUDA(cudaGetDeviceProperties(&prop, device));
    s->sm_count = prop.multiProcessorCount;
    BF_CUDA(cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking));
    s->stream = s->own_stream;
    BF_CUDA(cudaEventCreate(&s->ev_begin)); BF_CUDA(cudaEventCreate(&s->ev_end));

    // ---- BVH over all mesh triangles, shape order ----
    auto t0 = std::chrono::steady_clock::now();
    std::vector<BvhTriIn> tris;
    float scene_abs = 0.f;
    for (size_t si = 0; si < s->shapes.size(); ++si) {
        if (s->shapes[si].type != 0) continue;
        const HostMesh &m = s->meshes[s->shapes[si].index];
        for (uint32_t f = 0; f < m.nf; ++f) {
            BvhTriIn t; t.shape = (uint32_t) si; t.prim = f;
            for (int k = 0; k < 3; ++k) {
                uint32_t vi = m.idx[3 * (size_t) f + k];
                float *dst = k == 0 ? t.p0 : (k == 1 ? t.p1 : t.p2);
                V3 p(m.pos[3 * (size_t) vi], m.pos[3 * (size_t) vi + 1], m.pos[3 * (size_t) vi + 2]);
                if (m.moved) p = xf_point_affine(m.motion, p);          // bf_scene_update_transform before the commit
                dst[0] = p.x; dst[1] = p.y; dst[2] = p.z;
                for (int a = 0; a < 3; ++a) scene_abs = std::max(scene_abs, std::fabs(dst[a]));
            }
            tris.push_back(t);
        }
    }
    for (const DevRect &r : s->rects)
        for (int a = 0; a < 3; ++a)
            scene_abs = std::max(scene_abs, std::fabs(r.to_world.m[a][3]) + std::fabs(r.to_world.m[a][0]) + std::fabs(r.to_world.m[a][1]));
    {   // bounds of the meshes, padded like the rectangles' boxes (1e-4 of the largest coordinate)
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, mag = 1e-3f;
        for (const BvhTriIn &t : tris)
            for (int a = 0; a < 3; ++a) {
                lo[a] = std::min(lo[a], std::min(t.p0[a], std::min(t.p1[a], t.p2[a])));
                hi[a] = std::max(hi[a], std::max(t.p0[a], std::max(t.p1[a], t.p2[a])));
            }
        for (int a = 0; a < 3; ++a) if (!tris.empty()) mag = std::max(mag, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
        for (int a = 0; a < 3; ++a) { s->dev.mesh_lo[a] = lo[a] - 1e-4f * mag; s->dev.mesh_hi[a] = hi[a] + 1e-4f * mag; }
    }
    WideBvh bvh;
    build_wide_bvh(tris, 4e-6f * std::max(scene_abs, 1e-3f), bvh);
    if (2 * bvh.stats.depth + 2 > BF_TRACE_STACK_DEPTH)   // node groups and parked triangle groups share the stack
        return fail(s, BF_ERR_UNSUPPORTED, "bf_scene_commit: BVH deeper than the traversal stack");
    bool any_normals = false;
    for (const HostMesh &m : s->meshes) any_normals |= !m.nrm.empty();
    s->h_nodes_bytes = bvh.nodes.size() * 4; s->h_tris_bytes = bvh.tris.size() * 4;
    s->h_nrm_bytes = any_normals ? bvh.tris.size() * 4 : 0;
    s->h_geom_bytes = s->h_nodes_bytes + s->h_tris_bytes + s->h_nrm_bytes;
    BF_CUDA(cudaMallocHost(&s->h_geom, std::max<size_t>(s->h_geom_bytes, 16)));
    uint8_t *hg = (uint8_t *) s->h_geom;
    if (s->h_nodes_bytes) memcpy(hg, bvh.nodes.data(), s->h_nodes_bytes);
    if (s->h_tris_bytes) memcpy(hg + s->h_nodes_bytes, bvh.tris.data(), s->h_tris_bytes);
    if (any_normals) {
        float *nr = (float *) (hg + s->h_nodes_bytes + s->h_tris_bytes);
        for (size_t slot = 0; slot < bvh.tri_src.size(); ++slot) {
            const BvhTriIn &t = tris[bvh.tri_src[slot]];
            const HostMesh &m = s->meshes[s->shapes[t.shape].index];
            for (int k = 0; k < 3; ++k) {
                float *dst = nr + 12 * slot + 4 * k;
                if (m.nrm.empty()) { dst[0] = dst[1] = dst[2] = 0.f; }
                else {
                    uint32_t vi = m.idx[3 * (size_t) t.prim + k];
                    V3 n(m.nrm[3 * (size_t) vi], m.nrm[3 * (size_t) vi + 1], m.nrm[3 * (size_t) vi + 2]);
                    if (m.moved) { n = xf_normal(m4_inverse(m.motion), n); if (n.x != 0.f || n.y != 0.f || n.z != 0.f) n = normalize(n); }
                    dst[0] = n.x; dst[1] = n.y; dst[2] = n.z;
                }
                dst[3] = 0.f;
            }
        }
    }
    s->stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    s->stats.bvh_nodes = bvh.stats.n_nodes; s->stats.bvh_tris = bvh.stats.n_tris; s->stats.bvh_depth = bvh.stats.depth;
    s->bvh_level_first = bvh.level_first; s->bvh_pad_abs = bvh.pad_abs;
    s->stats.n_rects = (uint32_t) s->rects.size();

    auto t1 = std::chrono::steady_clock::now();
    BF_CUDA(s->d_nodes.alloc(bvh.nodes.size() / 4)); BF_CUDA(s->d_tris.alloc(bvh.tris.size() / 4));
    if (any_normals) BF_CUDA(s->d_tri_nrm.alloc(bvh.tris.size() / 4));
    int rc = upload_geometry(s);
    if (rc) return rc;
    rc = upload_tables(s);
    if (rc) return rc;
    s->dev.nodes = s->d_nodes.p; s->dev.tris = s->d_tris.p; s->dev.tri_nrm = any_normals ? s->d_tri_nrm.p : nullptr;
    s->dev.n_nodes = bvh.stats.n_nodes; s->dev.n_tris = bvh.stats.n_tris;

    s->profile_floats = (size_t) s->dev.film_w * s->dev.film_h * s->dev.channels;
    BF_CUDA(s->d_profile.alloc(s->profile_floats));
    BF_CUDA(cudaMemsetAsync(s->d_profile.p, 0, s->profile_floats * 4, s->stream));
    BF_CUDA(s->d_totals.alloc(4));
    BF_CUDA(cudaMemsetAsync(s->d_totals.p, 0, 32, s->stream));
    BF_CUDA(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
    BF_CUDA(cudaStreamSynchronize(s->stream));
    s->stats.upload_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
    s->committed = true;
    apply_l2_persistence(s);      // BF_L2_PERSIST_MB / bf_set_tunable("l2_persist_mb"): off unless asked for
    return BF_OK;
}

int bf_scene_reupload(bf_scene *s) {
    if (!s || !s->committed) return fail(s, BF_ERR_STATE, "bf_scene_reupload: scene not committed");
    BF_CUDA(cudaSetDevice(s->device));
    int rc = upload_geometry(s);
    if (rc) return rc;
    if (s->geometry_moved) {      // the pinned image is the committed pose: re-apply the motions set since
        LaunchCfg cfg; memset(&cfg, 0, sizeof(cfg)); cfg.stream = s->stream; cfg.sm_count = s->sm_count;
        for (size_t si = 0; si < s->shapes.size(); ++si) {
            if (s->shapes[si].type != 0 || !s->meshes[s->shapes[si].index].moved) continue;
            const HostMesh &m = s->meshes[s->shapes[si].index];
            launch_refit_tris(s->d_tris_rest.p, s->dev.tri_nrm ? s->d_nrm_rest.p : nullptr, s->d_tris.p, s->d_tri_nrm.p, s->dev.n_tris, (uint32_t) si, m.motion,
                              m4_inverse(m.motion), cfg);
        }
        launch_refit_nodes(s->d_nodes.p, s->d_tris.p, s->d_node_box.p, s->bvh_level_first.data(), (uint32_t) s->bvh_level_first.size() - 1, s->bvh_pad_abs, cfg);
    }
    return upload_tables(s);
}

int bf_scene_device_bytes(const bf_scene *s, size_t *geometry_bytes) {
    if (!s || !s->committed) return BF_ERR_STATE;
    if (geometry_bytes) *geometry_bytes = s->h_geom_bytes;
    return BF_OK;
}

int bf_scene_update_rect(bf_scene *s, uint32_t shape_id, const float to_world[16]) {
    if (!s || !to_world) return BF_ERR_INVALID;
    if (shape_id >= s->shapes.size() || s->shapes[shape_id].type != 1) return fail(s, BF_ERR_INVALID, "bf_scene_update_rect: not a rectangle");
    s->rects[s->shapes[shape_id].index] = make_rect(to_world, shape_id);
    memcpy(s->blob.data() + s->rect_blob_off[s->shapes[shape_id].index] + sizeof(bf_flat_record), to_world, 64);   // keep the export current
    if (!s->tables_dirty) s->tables_light = true;
    s->tables_dirty = true;
    return BF_OK;
}

// Padded world-space bounds of all meshes from the (moved) rest-pose boxes: conservative, host only
static void refresh_mesh_bounds(bf_scene *s) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, mag = 1e-3f;
    bool any = false;
    for (const HostMesh &m : s->meshes) {
        if (m.nf == 0) continue;
        any = true;
        for (int c = 0; c < 8; ++c) {
            V3 p((c & 1) ? m.rest_hi[0] : m.rest_lo[0], (c & 2) ? m.rest_hi[1] : m.rest_lo[1], (c & 4) ? m.rest_hi[2] : m.rest_lo[2]);
            if (m.moved) p = xf_point_affine(m.motion, p);
            const float v[3] = {p.x, p.y, p.z};
            for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], v[a]); hi[a] = std::max(hi[a], v[a]); }
        }
    }
    for (int a = 0; a < 3; ++a) if (any) mag = std::max(mag, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
    for (int a = 0; a < 3; ++a) { s->dev.mesh_lo[a] = lo[a] - 2e-4f * mag; s->dev.mesh_hi[a] = hi[a] + 2e-4f * mag; }
}

int bf_scene_update_transform(bf_scene *s, uint32_t shape_id, const float motion[16]) {
    NvtxRange range("bf_scene_update_transform (refit)");
    if (!s || !motion) return BF_ERR_INVALID;
    if (shape_id >= s->shapes.size()) return fail(s, BF_ERR_INVALID, "bf_scene_update_transform: shape id out of range");
    if (s->shapes[shape_id].type == 1) return bf_scene_update_rect(s, shape_id, motion);
    HostMesh &m = s->meshes[s->shapes[shape_id].index];
    memcpy(m.motion.m, motion, 64);
    m.moved = true;
    uint32_t hdr[2] = {shape_id, 0};
    record(s, BF_TAG_MESH_MOTION, hdr, 8, motion, 64);
    if (!s->committed) return BF_OK;            // the moved vertices go into the build
    BF_CUDA(cudaSetDevice(s->device));
    const uint32_t n_tris = s->dev.n_tris;
    if (n_tris == 0) return BF_OK;
    if (!s->rest_saved) {
        // first motion after the commit: the rest-pose triangle (and normal) records in leaf order, rebuilt from the vertices as
        // given to bf_scene_add_mesh (the committed table may hold a pose set before the commit); (shape, prim) of every slot
        // come from the pinned image of the committed geometry
        const float *ht = (const float *) ((const uint8_t *) s->h_geom + s->h_nodes_bytes);
        std::vector<float> rest(12 * (size_t) n_tris), rest_n(s->dev.tri_nrm ? 12 * (size_t) n_tris : 0);
        for (size_t slot = 0; slot < n_tris; ++slot) {
            uint32_t shape, prim; memcpy(&shape, ht + 12 * slot + 3, 4); memcpy(&prim, ht + 12 * slot + 7, 4);
            const HostMesh &hm = s->meshes[s->shapes[shape].index];
            for (int k = 0; k < 3; ++k) {
                const uint32_t vi = hm.idx[3 * (size_t) prim + k];
                float *dst = rest.data() + 12 * slot + 4 * k;
                dst[0] = hm.pos[3 * (size_t) vi]; dst[1] = hm.pos[3 * (size_t) vi + 1]; dst[2] = hm.pos[3 * (size_t) vi + 2];
                memcpy(dst + 3, ht + 12 * slot + 4 * k + 3, 4);
                if (!rest_n.empty()) {
                    float *dn = rest_n.data() + 12 * slot + 4 * k;
                    if (hm.nrm.empty()) { dn[0] = dn[1] = dn[2] = 0.f; }
                    else { dn[0] = hm.nrm[3 * (size_t) vi]; dn[1] = hm.nrm[3 * (size_t) vi + 1]; dn[2] = hm.nrm[3 * (size_t) vi + 2]; }
                    dn[3] = 0.f;
                }
            }
        }
        BF_CUDA(s->d_tris_rest.alloc(3 * (size_t) n_tris));
        BF_CUDA(cudaMemcpy(s->d_tris_rest.p, rest.data(), 48 * (size_t) n_tris, cudaMemcpyHostToDevice));
        if (!rest_n.empty()) {
            BF_CUDA(s->d_nrm_rest.alloc(3 * (size_t) n_tris));
            BF_CUDA(cudaMemcpy(s->d_nrm_rest.p, rest_n.data(), 48 * (size_t) n_tris, cudaMemcpyHostToDevice));
        }
        BF_CUDA(s->d_node_box.alloc(6 * (size_t) s->dev.n_nodes));
        s->rest_saved = true;
    }
    LaunchCfg cfg; memset(&cfg, 0, sizeof(cfg)); cfg.stream = s->stream; cfg.sm_count = s->sm_count;
    uint64_t launches = 0;
    launches += launch_refit_tris(s->d_tris_rest.p, s->dev.tri_nrm ? s->d_nrm_rest.p : nullptr, s->d_tris.p, s->d_tri_nrm.p, n_tris, shape_id, m.motion,
                                  m4_inverse(m.motion), cfg);
    launches += launch_refit_nodes(s->d_nodes.p, s->d_tris.p, s->d_node_box.p, s->bvh_level_first.data(), (uint32_t) s->bvh_level_first.size() - 1,
                                   s->bvh_pad_abs, cfg);
    s->stats.kernel_launches += launches;
    BF_CUDA(cudaGetLastError());
    refresh_mesh_bounds(s);
    if (s->dev.n_nodes != 0 && s->has_rx) estimate_primary_defer(s);
    s->geometry_moved = true;
    return BF_OK;
}

// ---- wave slots --------------------------------------------------------------------------------------------
// A call to bf_receive is cut into waves of BF_WAVE_LANES lanes; BF_WAVE_STREAMS waves are in flight at a
// time, each on its own stream with its own state buffers.  Every traversal launch ends with a latency-bound
// tail (a handful of rays with ~100 sequential node visits); with several waves in flight the tail of one
// wave's launch overlaps the bulk of another's.

static void free_slot(WaveSlot &sl, bool keep_stream) {
    for (void *p : sl.allocs) cudaFree(p);
    sl.allocs.clear();
    memset(&sl.wave, 0, sizeof(sl.wave));
    sl.capacity = 0; sl.path_stash = nullptr; sl.d_rects = nullptr; sl.n_rects = 0; sl.d_rx = nullptr;
    if (!keep_stream) {
        if (sl.stream && sl.own_stream) cudaStreamDestroy(sl.stream);
        if (sl.done) cudaEventDestroy(sl.done);
        if (sl.h_counts) cudaFreeHost(sl.h_counts);
        sl.stream = nullptr; sl.done = nullptr; sl.h_counts = nullptr;
    }
}
static void free_wave(bf_scene *s) {
    for (WaveSlot &sl : s->slots) free_slot(sl, false);
    s->slots.clear();
}

// L2 persistence for the top of the BVH: the node array is in breadth-first order, so a window over its
// first bytes covers the levels every ray visits.  Kernels launched on `st` afterwards inherit the window.
static void apply_l2_window(bf_scene *s, cudaStream_t st) {
    if (!s->l2_window_touched && s->tune.l2_persist_mb <= 0) return;
    cudaStreamAttrValue a;
    memset(&a, 0, sizeof(a));
    size_t bytes = std::min<size_t>((size_t) s->tune.l2_persist_mb << 20, s->dev.n_nodes * (size_t) 80);
    bytes = std::min<size_t>(bytes, s->l2_window_max);
    a.accessPolicyWindow.base_ptr = bytes ? (void *) s->d_nodes.p : nullptr;
    a.accessPolicyWindow.num_bytes = bytes;
    a.accessPolicyWindow.hitRatio = bytes ? 1.f : 0.f;
    a.accessPolicyWindow.hitProp = bytes ? cudaAccessPropertyPersisting : cudaAccessPropertyNormal;
    a.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
    if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &a) != cudaSuccess) cudaGetLastError();
    s->l2_window_touched = true;
}
static void apply_l2_persistence(bf_scene *s) {
    if (!s->committed || (!s->l2_window_touched && s->tune.l2_persist_mb <= 0)) return;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, s->device) != cudaSuccess) { cudaGetLastError(); return; }
    s->l2_window_max = (size_t) std::max(prop.accessPolicyMaxWindowSize, 0);
    const size_t want = std::min<size_t>((size_t) s->tune.l2_persist_mb << 20, (size_t) std::max(prop.persistingL2CacheMaxSize, 0));
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) != cudaSuccess) cudaGetLastError();
    if (want == 0) cudaCtxResetPersistingL2Cache();
    apply_l2_window(s, s->stream);
    for (WaveSlot &sl : s->slots) if (sl.own_stream) apply_l2_window(s, sl.stream);
}

static int ensure_slot(bf_scene *s, size_t index, uint32_t lanes, bool records) {
    while (s->slots.size() <= index) {
        WaveSlot sl;
        memset(&sl.wave, 0, sizeof(sl.wave));
        if (s->slots.empty()) { sl.stream = s->stream; sl.own_stream = false; }
        else { BF_CUDA(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking)); sl.own_stream = true; apply_l2_window(s, sl.stream); }
        BF_CUDA(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
        BF_CUDA(cudaMallocHost((void **) &sl.h_counts, 64));
        s->slots.push_back(sl);
    }
    WaveSlot &sl = s->slots[index];
    if (sl.capacity >= lanes && (!records || sl.wave.records)) return BF_OK;
    BF_CUDA(cudaStreamSynchronize(sl.stream));
    free_slot(sl, true);
    auto A = [&](void **p, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) sl.allocs.push_back(*p);
        return e;
    };
    Wave &w = sl.wave;
    size_t n = lanes;
    BF_CUDA(A((void **) &w.ray_o, n * 16)); BF_CUDA(A((void **) &w.ray_d, n * 16)); BF_CUDA(A((void **) &w.sh_d, n * 16));
    BF_CUDA(A((void **) &w.hit, n * 16)); BF_CUDA(A((void **) &w.occl, n * 4)); BF_CUDA(A((void **) &w.rng, n * 16));
    BF_CUDA(A((void **) &w.x1, n * 16)); BF_CUDA(A((void **) &w.x2, n * 16));
    int n_spec = s->C.spectrum_width == 1 ? 1 : 5;
    for (int k = 0; k < n_spec; ++k) BF_CUDA(A((void **) &w.spec[k], n * 16));
    BF_CUDA(A((void **) &w.flags, n * 4));
    BF_CUDA(A((void **) &w.queue[0], n * 4)); BF_CUDA(A((void **) &w.queue_h, n * 4)); BF_CUDA(A((void **) &w.queue_h2, n * 4)); BF_CUDA(A((void **) &w.queue_s, n * 4)); BF_CUDA(A((void **) &w.queue_d, n * 4));
    BF_CUDA(A((void **) &w.counters, (size_t) BF_CTR_STRIDE * (BF_MAX_BOUNCES + 2) * 4));
    if (records) { BF_CUDA(A((void **) &w.records, n * sizeof(bf_sample_record))); BF_CUDA(A((void **) &w.sample_of, n * 4)); }
    w.capacity = lanes;
    w.totals = s->d_totals.p;
    sl.capacity = lanes;
    return BF_OK;
}

static cudaEvent_t next_trace_event(bf_scene *s, int kind) {
    if (s->ev_trace_used == s->ev_trace.size()) {
        cudaEvent_t e; cudaEventCreate(&e); s->ev_trace.push_back(e); s->ev_kind.push_back(0);
    }
    s->ev_kind[s->ev_trace_used] = (uint8_t) kind;
    return s->ev_trace[s->ev_trace_used++];
}

// One wave on one slot: lanes [0, n) <-> sample indices [first, first + n)
static int run_wave(bf_scene *s, WaveSlot &sl, uint64_t seed, uint64_t first, uint32_t n, float *profile, bool with_records,
                    const DevScene *scene_override = nullptr) {
    const DevScene &sc = scene_override ? *scene_override : s->dev;
    NvtxRange range("bf wave");
    Wave w = sl.wave;
    if (!with_records) { w.records = nullptr; w.sample_of = nullptr; }
    // BF_FUSED_PRIMARY=0 selects the unfused raygen + trace pair (lanes = samples), kept for A/B measurements
    // wide reconstruction filters: every sample (misses too) is splatted by k_finalize_wide, so lanes = samples
    const bool fused = s->tune.fused_primary != 0 && !s->wide_filter;
    if (!fused) w.sample_of = nullptr;
    LaunchCfg cfg; cfg.stream = sl.stream; cfg.sm_count = s->sm_count; cfg.trace_rounds = s->tune.trace_rounds >= 0 ? (uint32_t) s->tune.trace_rounds : s->auto_rounds;
    uint32_t *ctr = w.counters;
    const bool unbounded = sc.integ.max_depth < 0 || sc.integ.max_depth > 64;
    const uint32_t max_depth = sc.integ.max_depth < 0 ? BF_MAX_BOUNCES : std::min<uint32_t>((uint32_t) sc.integ.max_depth, BF_MAX_BOUNCES);
    const uint32_t ctr_blocks = unbounded ? BF_MAX_BOUNCES + 2 : max_depth + 2;
    const bool mega = fused && s->tune.mega_max >= 0 && (sc.n_nodes == 0 || (int64_t) n <= s->tune.mega_max);
    BF_CUDA(cudaMemsetAsync(ctr, 0, (size_t) BF_CTR_STRIDE * (mega ? 1 : ctr_blocks) * 4, sl.stream));
    uint64_t launches = 0;
    const bool timing = s->time_trace;
    auto timed = [&](int kind, auto &&fn) {
        if (timing) { cudaEventRecord(next_trace_event(s, kind), sl.stream); }
        launches += fn();
        if (timing) { cudaEventRecord(next_trace_event(s, kind), sl.stream); }
    };
    bool path = false;
    if (mega) {
        // scenes without meshes, and small requests on any scene: the whole path of a sample in one thread, one launch
        timed(BF_K_PRIMARY, [&] { return launch_mega(sc, w, seed, first, n, ctr + 2, profile, cfg); });
        s->stats.kernel_launches += launches;
        BF_CUDA(cudaGetLastError());
        return BF_OK;
    }
    if (fused) {
        w.n_lanes = ctr + 6;
        // few receiver rays reach the meshes (bf_scene_commit's estimate): k_primary leaves those to a dense k_trace launch
        const bool path_ok = s->tune.path_fused != 0 && path_kernel_available(sc);
        const bool defer = (s->tune.primary_defer < 0 ? s->primary_defer || (s->path_defer && path_ok) : s->tune.primary_defer != 0) && sc.n_nodes != 0;
        // ... and the paths that stay clear of the meshes altogether are integrated by k_path without touching the wave arrays
        path = defer && path_ok;
        if (path) {
            if (!sl.path_stash) {
                sl.stash_cap = path_stash_cap(sl.capacity, s->sm_count, (uint32_t) s->tune.path_stash_max);
                BF_CUDA(cudaMalloc((void **) &sl.path_stash, path_stash_bytes(sc, s->sm_count, sl.stash_cap)));
                sl.allocs.push_back(sl.path_stash);
            }
            timed(BF_K_PRIMARY, [&] { return launch_path(sc, w, seed, first, n, ctr, profile, w.queue_d, sl.path_stash, sl.stash_cap, cfg); });
            timed(BF_K_TRACE_CLOSEST, [&] { return launch_trace_closest(sc, w, w.queue_d, ctr + 0, 0, 1, ctr + 4, ctr + 3, cfg, false); });
        } else {
            timed(BF_K_PRIMARY, [&] { return launch_primary(sc, w, seed, first, n, ctr + 4, ctr + 2, profile, defer ? w.queue[0] : nullptr, ctr + 0, cfg); });
            if (defer) timed(BF_K_TRACE_CLOSEST, [&] { return launch_trace_closest(sc, w, w.queue[0], ctr + 0, 0, 1, ctr + 4, ctr + 3, cfg, false); });
        }
    } else {
        w.n_lanes = nullptr;
        timed(BF_K_RAYGEN, [&] { return launch_raygen(sc, w, seed, first, n, cfg); });
        timed(BF_K_TRACE_CLOSEST, [&] { return launch_trace_closest(sc, w, nullptr, nullptr, n, 1, ctr + 4, ctr + 2, cfg); });
    }
    // k_path has already sorted the continuation rays of depth 1: what reaches queue[0] mostly needs the BVH
    const bool trace_defer = !path && (s->tune.trace_defer < 0 ? s->primary_defer : s->tune.trace_defer != 0) && sc.n_nodes != 0;
    for (uint32_t depth = 1; depth <= max_depth; ++depth) {
        uint32_t *blk = ctr + (size_t) BF_CTR_STRIDE * depth, *prev = ctr + (size_t) BF_CTR_STRIDE * (depth - 1);
        timed(BF_K_SHADE, [&] { return launch_shade(sc, w, depth, prev + 4, n, w.queue[0], blk, cfg); });
        if (depth >= max_depth) break;
        timed(BF_K_TRACE_SHADOW, [&] { return launch_trace_shadow(sc, w, w.queue_s, blk + 1, blk + 3, cfg); });
        if (trace_defer) {
            // the shadow launch above has consumed queue_s: it carries the deferred lanes to the dense follow-up launch
            timed(BF_K_TRACE_CLOSEST, [&] { return launch_trace_closest(sc, w, w.queue[0], blk + 0, 0, depth + 1, blk + 4, blk + 2, cfg, true, w.queue_s, blk + 6); });
            timed(BF_K_TRACE_CLOSEST, [&] { return launch_trace_closest(sc, w, w.queue_s, blk + 6, 0, depth + 1, blk + 4, blk + 7, cfg, false); });
        } else {
            timed(BF_K_TRACE_CLOSEST, [&] { return launch_trace_closest(sc, w, w.queue[0], blk + 0, 0, depth + 1, blk + 4, blk + 2, cfg); });
        }
        if (unbounded && depth >= 8) {
            BF_CUDA(cudaMemcpyAsync(sl.h_counts, blk, 8, cudaMemcpyDeviceToHost, sl.stream));
            BF_CUDA(cudaStreamSynchronize(sl.stream));
            if (sl.h_counts[0] == 0) break;
        }
    }
    timed(BF_K_FINALIZE, [&] { return launch_finalize(sc, w, n, profile, s->wide_filter ? &s->filter : nullptr, cfg); });
    s->stats.kernel_launches += launches;
    BF_CUDA(cudaGetLastError());
    return BF_OK;
}

static int receive_into(bf_scene *s, uint64_t seed, uint64_t first_sample, uint64_t n_samples, int accumulate, float *profile) {
    NvtxRange range("bf_receive");
    BF_CUDA(cudaSetDevice(s->device));
    if (s->tables_dirty) { int rc = (s->tables_light && s->committed) ? upload_tables_light(s) : upload_tables(s); if (rc) return rc; }
    // Wave size: BF_WAVE_LANES when set; otherwise the request is cut into as many waves as there are slots, between 2^22 and
    // 2^25 samples each (larger waves amortise the traversal tails: profiles/r01d_tune_wave.txt)
    const uint64_t streams = (uint64_t) s->tune.wave_streams;
    uint64_t want_lanes = (uint64_t) std::max<int64_t>(s->tune.wave_lanes, 1024);
    if (s->tune.wave_lanes == 0) {
        const uint64_t per_slot = (n_samples + streams - 1) / streams;
        want_lanes = std::min<uint64_t>(std::max<uint64_t>((per_slot + 1023) / 1024 * 1024, 1u << 22), 1u << 25);
    }
    const uint32_t lanes = (uint32_t) std::min<uint64_t>(std::max<uint64_t>(n_samples, 1024), want_lanes);
    const uint64_t n_waves = (n_samples + lanes - 1) / lanes;
    // per-kernel timing needs the launches serialised on one stream
    const size_t n_slots = s->time_trace ? 1 : (size_t) std::min<uint64_t>(streams, std::max<uint64_t>(n_waves, 1));
    for (size_t k = 0; k < n_slots; ++k) { int rc = ensure_slot(s, k, lanes, false); if (rc) return rc; }
    // A traversal launch trades its tail against its latency: few warps with many batches each pay the per-warp tail (the
    // last rays of a warp run at a few lanes of 32) once per many rays but leave the SMs to the launch's slowest warp.  A
    // request of several waves in flight fills those gaps with the other waves' launches, so it takes 48 batches per warp
    // (C2 +1.5 %, C3 +7.6 % over 8: tools/ab_env4.sh, profiles/r02u); a request that is one wave keeps 8 for its latency.
    s->auto_rounds = n_waves >= 4 ? 48u : 8u;
    s->ev_trace_used = 0;
    BF_CUDA(cudaEventRecord(s->ev_begin, s->stream));
    if (!accumulate) BF_CUDA(cudaMemsetAsync(profile, 0, s->profile_floats * 4, s->stream));
    if (n_slots > 1) {
        BF_CUDA(cudaEventRecord(s->ev_fork, s->stream));
        for (size_t k = 1; k < n_slots; ++k) BF_CUDA(cudaStreamWaitEvent(s->slots[k].stream, s->ev_fork, 0));
    }
    uint64_t wave_index = 0;
    for (uint64_t done = 0; done < n_samples; done += lanes, ++wave_index) {
        uint32_t n = (uint32_t) std::min<uint64_t>(lanes, n_samples - done);
        int rc = run_wave(s, s->slots[wave_index % n_slots], seed, first_sample + done, n, profile, false);
        if (rc) return rc;
    }
    for (size_t k = 1; k < n_slots; ++k) {
        BF_CUDA(cudaEventRecord(s->slots[k].done, s->slots[k].stream));
        BF_CUDA(cudaStreamWaitEvent(s->stream, s->slots[k].done, 0));
    }
    BF_CUDA(cudaEventRecord(s->ev_end, s->stream));
    s->stats.samples += n_samples;
    return BF_OK;
}

int bf_receive(bf_scene *s, uint64_t seed, uint64_t first_sample, uint64_t n_samples, int accumulate) {
    if (!s) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_receive: scene not committed");
    return receive_into(s, seed, first_sample, n_samples, accumulate, s->d_profile.p);
}

// Many receive() calls as ONE submission (C4: the frames of a moving-target sequence; C5: the elements of a receive array --
// the reference holds std::vector<ref<Receiver>>, scene.h:197, and calls receive() per receiver): item i re-poses the listed
// rectangles and / or the receiver, then integrates n_samples into ITS OWN profile, slot i of a device-side batch buffer.
// Nothing synchronises between items (light table uploads through the pinned ring); one copy brings all profiles back.
// Every item's tables into one pinned block, one upload, one memset of the profiles and cursors, one k_mega_batch launch.
static int receive_batch_one_launch(bf_scene *s, uint32_t n_items, const bf_batch_item *items, uint32_t n_samples) {
    if (s->tables_dirty && !s->tables_light) { int rc = upload_tables(s); if (rc) return rc; }     // a heavy change pending from before the batch
    const size_t n_rects = s->rects.size();
    const size_t rect_bytes = (n_rects * sizeof(DevRect) + 15) & ~(size_t) 15;
    const size_t ep_bytes = (sizeof(DevEndpoint) + 15) & ~(size_t) 15, item_bytes = (sizeof(MegaItem) + 15) & ~(size_t) 15;
    const size_t off_ep = item_bytes * n_items, off_rects = off_ep + ep_bytes * n_items;
    const size_t off_cursors = off_rects + rect_bytes * n_items, total = off_cursors + 4 * (size_t) n_items;
    if (s->h_items_bytes < total || s->d_items.n < total) {
        BF_CUDA(cudaStreamSynchronize(s->stream));
        if (s->h_items) cudaFreeHost(s->h_items);
        s->h_items = nullptr; s->h_items_bytes = 0;
        BF_CUDA(cudaMallocHost((void **) &s->h_items, total));
        s->h_items_bytes = total;
        BF_CUDA(s->d_items.alloc(total));
    }
    if (!s->items_ev) BF_CUDA(cudaEventCreateWithFlags(&s->items_ev, cudaEventDisableTiming));
    BF_CUDA(cudaEventSynchronize(s->items_ev));          // the previous batch's upload has left the pinned block
    uint8_t *h = s->h_items, *d = s->d_items.p;
    for (uint32_t i = 0; i < n_items; ++i) {
        const bf_batch_item &it = items[i];
        for (uint32_t k = 0; k < it.n_rect_updates; ++k) {
            int rc = bf_scene_update_rect(s, it.rect_shape_ids[k], it.rect_to_world + 16 * (size_t) k);
            if (rc) return rc;
        }
        if (it.receiver) { int rc = bf_scene_set_receiver(s, it.receiver_shape, it.receiver); if (rc) return rc; }
        if (s->tables_dirty && !s->tables_light)
            return fail(s, BF_ERR_UNSUPPORTED, "bf_receive_batch: an item may re-pose rectangles and replace the receiver by one of the same kind only");
        if ((s->rx.kind != BF_RX_PERSPECTIVE) != (s->dev.rx_needs_aperture != 0))
            return fail(s, BF_ERR_UNSUPPORTED, "bf_receive_batch: an item may re-pose rectangles and replace the receiver by one of the same kind only");
        MegaItem m;
        memset((void *) &m, 0, sizeof(m));
        m.seed = it.seed; m.first_sample = it.first_sample;
        m.profile = s->d_batch.p + (size_t) i * s->profile_floats;
        m.rects = (const DevRect *) (d + off_rects + rect_bytes * i);
        m.rx = (const DevEndpoint *) (d + off_ep + ep_bytes * i);
        m.rx_ep = make_endpoint(s->rx, s->rx_shape);
        if (s->rx_shape >= 0) m.rx_rect = s->rects[s->shapes[s->rx_shape].index];
        if (s->dev.n_txs != 0 && s->dev.tx0_ep.shape >= 0) m.tx0_rect = s->rects[s->shapes[s->dev.tx0_ep.shape].index];
        memcpy(h + item_bytes * i, &m, sizeof(m));
        memcpy(h + off_ep + ep_bytes * i, &m.rx_ep, sizeof(DevEndpoint));
        if (n_rects) memcpy(h + off_rects + rect_bytes * i, s->rects.data(), n_rects * sizeof(DevRect));
    }
    s->ev_trace_used = 0;
    BF_CUDA(cudaEventRecord(s->ev_begin, s->stream));
    BF_CUDA(cudaMemcpyAsync(d, h, off_cursors, cudaMemcpyHostToDevice, s->stream));
    BF_CUDA(cudaEventRecord(s->items_ev, s->stream));
    BF_CUDA(cudaMemsetAsync(d + off_cursors, 0, 4 * (size_t) n_items, s->stream));
    BF_CUDA(cudaMemsetAsync(s->d_batch.p, 0, (size_t) n_items * s->profile_floats * 4, s->stream));
    LaunchCfg cfg; memset(&cfg, 0, sizeof(cfg)); cfg.stream = s->stream; cfg.sm_count = s->sm_count;
    if (s->time_trace) cudaEventRecord(next_trace_event(s, BF_K_PRIMARY), s->stream);
    s->stats.kernel_launches += launch_mega_batch(s->dev, (const MegaItem *) d, n_items, n_samples, (uint32_t *) (d + off_cursors), s->d_totals.p, cfg);
    if (s->time_trace) cudaEventRecord(next_trace_event(s, BF_K_PRIMARY), s->stream);
    BF_CUDA(cudaGetLastError());
    BF_CUDA(cudaEventRecord(s->ev_end, s->stream));
    s->stats.samples += (uint64_t) n_items * n_samples;
    // the scene's own device tables hold none of the items' poses: refreshed (light) before the next plain bf_receive
    s->tables_dirty = true; s->tables_light = true;
    return BF_OK;
}

int bf_receive_batch(bf_scene *s, uint32_t n_items, const bf_batch_item *items, uint64_t n_samples, float *profiles, size_t n_floats) {
    NvtxRange range("bf_receive_batch");
    if (!s || (n_items && !items)) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_receive_batch: scene not committed");
    if (profiles && n_floats < (size_t) n_items * s->profile_floats) return fail(s, BF_ERR_INVALID, "bf_receive_batch: destination too small");
    BF_CUDA(cudaSetDevice(s->device));
    if (s->d_batch.n < (size_t) n_items * s->profile_floats) BF_CUDA(cudaStreamSynchronize(s->stream));
    BF_CUDA(s->d_batch.alloc((size_t) n_items * s->profile_floats));
    // Items that fit one wave run concurrently: item i goes to wave slot i % slots, on that slot's stream, with the slot's OWN
    // copy of the tables the item re-poses (rectangles, receiver), so neighbouring items never wait for each other.  Larger
    // items (several waves each) already fill the machine and run one after the other on the scene's tables.
    const bool concurrent = n_samples <= (1u << 22) && !s->time_trace && s->tune.wave_streams > 1;
    // Items small enough for k_mega (no meshes, or <= mega_max samples) go out as ONE launch over (samples, items)
    const bool one_launch = n_items > 1 && s->tune.mega_batch != 0 && s->tune.fused_primary != 0 && !s->wide_filter &&
        s->tune.mega_max >= 0 && n_samples > 0 && n_samples <= (1u << 22) && (s->dev.n_nodes == 0 || (int64_t) n_samples <= s->tune.mega_max) &&
        (uint64_t) n_items <= 65535u;
    if (one_launch) {
        int rc = receive_batch_one_launch(s, n_items, items, (uint32_t) n_samples);
        if (rc) return rc;
    } else if (!concurrent) {
        for (uint32_t i = 0; i < n_items; ++i) {
            const bf_batch_item &it = items[i];
            for (uint32_t k = 0; k < it.n_rect_updates; ++k) {
                int rc = bf_scene_update_rect(s, it.rect_shape_ids[k], it.rect_to_world + 16 * (size_t) k);
                if (rc) return rc;
            }
            if (it.receiver) { int rc = bf_scene_set_receiver(s, it.receiver_shape, it.receiver); if (rc) return rc; }
            int rc = receive_into(s, it.seed, it.first_sample, n_samples, 0, s->d_batch.p + (size_t) i * s->profile_floats);
            if (rc) return rc;
        }
    } else {
        if (s->tables_dirty && !s->tables_light) { int rc = upload_tables(s); if (rc) return rc; }     // a heavy change pending from before the batch
        const size_t n_slots = (size_t) std::min<int64_t>(s->tune.wave_streams, std::max<uint32_t>(n_items, 1u));
        const uint32_t lanes = (uint32_t) std::max<uint64_t>(n_samples, 1024);
        for (size_t k = 0; k < n_slots; ++k) { int rc = ensure_slot(s, k, lanes, false); if (rc) return rc; }
        s->auto_rounds = 8u;        // items of <= 2^22 samples: measured no better at 16, worse at 32 (C5, tools/ab_env4.sh)
        s->ev_trace_used = 0;
        BF_CUDA(cudaEventRecord(s->ev_begin, s->stream));
        BF_CUDA(cudaEventRecord(s->ev_fork, s->stream));
        for (size_t k = 1; k < n_slots; ++k) BF_CUDA(cudaStreamWaitEvent(s->slots[k].stream, s->ev_fork, 0));
        for (uint32_t i = 0; i < n_items; ++i) {
            const bf_batch_item &it = items[i];
            for (uint32_t k = 0; k < it.n_rect_updates; ++k) {
                int rc = bf_scene_update_rect(s, it.rect_shape_ids[k], it.rect_to_world + 16 * (size_t) k);
                if (rc) return rc;
            }
            if (it.receiver) { int rc = bf_scene_set_receiver(s, it.receiver_shape, it.receiver); if (rc) return rc; }
            if (s->tables_dirty && !s->tables_light)
                return fail(s, BF_ERR_UNSUPPORTED, "bf_receive_batch: an item may re-pose rectangles and replace the receiver by one of the same kind only");
            WaveSlot &sl = s->slots[i % n_slots];
            int rc = upload_tables_light(s, &sl);
            if (rc) return rc;
            DevScene sc = s->dev;
            sc.rects = sl.d_rects; sc.rx = sl.d_rx;
            float *profile = s->d_batch.p + (size_t) i * s->profile_floats;
            BF_CUDA(cudaMemsetAsync(profile, 0, s->profile_floats * 4, sl.stream));
            rc = run_wave(s, sl, it.seed, it.first_sample, (uint32_t) n_samples, profile, false, &sc);
            if (rc) return rc;
            s->stats.samples += n_samples;
        }
        for (size_t k = 1; k < n_slots; ++k) {
            BF_CUDA(cudaEventRecord(s->slots[k].done, s->slots[k].stream));
            BF_CUDA(cudaStreamWaitEvent(s->stream, s->slots[k].done, 0));
        }
        BF_CUDA(cudaEventRecord(s->ev_end, s->stream));
    }
    s->batch_items = n_items;
    if (profiles && n_items) {
        BF_CUDA(cudaMemcpyAsync(profiles, s->d_batch.p, (size_t) n_items * s->profile_floats * 4, cudaMemcpyDeviceToHost, s->stream));
        BF_CUDA(cudaStreamSynchronize(s->stream));
    }
    return BF_OK;
}
int bf_batch_device_ptr(bf_scene *s, void **dptr, size_t *n_floats, void **cuda_stream) {
    if (!s || !s->committed) return BF_ERR_STATE;
    if (dptr) *dptr = s->d_batch.p;
    if (n_floats) *n_floats = (size_t) s->batch_items * s->profile_floats;
    if (cuda_stream) *cuda_stream = (void *) s->stream;
    return BF_OK;
}

// ---- multi-GPU: one ncclReduce of the per-rank profiles -----------------------------------------
// NCCL is bound at run time so that the library has no link-time dependency on it.
namespace {
struct NcclId { char internal[128]; };
struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(void **, int, NcclId, int) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, void *, cudaStream_t) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string error;
};
NcclApi &nccl_api() {
    static NcclApi api;
    if (api.handle || !api.error.empty()) return api;
    const char *names[] = {getenv("BF_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        if (!n || !*n) continue;
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
        api.error = dlerror();
    }
    if (!api.handle) { if (api.error.empty()) api.error = "libnccl.so.2 not found"; return api; }
    api.GetUniqueId = (decltype(api.GetUniqueId)) dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank)) dlsym(api.handle, "ncclCommInitRank");
    api.Reduce = (decltype(api.Reduce)) dlsym(api.handle, "ncclReduce");
    api.CommDestroy = (decltype(api.CommDestroy)) dlsym(api.handle, "ncclCommDestroy");
    api.GetErrorString = (decltype(api.GetErrorString)) dlsym(api.handle, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.Reduce || !api.CommDestroy) { api.error = "libnccl: missing symbols"; api.handle = nullptr; }
    else api.error.clear();
    return api;
}
std::string nccl_err(const NcclApi &api, int rc) { return api.GetErrorString ? api.GetErrorString(rc) : ("nccl error " + std::to_string(rc)); }
}  // namespace

}  // extern "C"
static void multi_destroy_comm(void *comm) { if (nccl_api().handle) nccl_api().CommDestroy(comm); }
extern "C" {
int bf_multi_unique_id(void *id, size_t capacity) {
    if (!id || capacity < BF_MULTI_ID_BYTES) return BF_ERR_INVALID;
    NcclApi &api = nccl_api();
    if (!api.handle) return fail(nullptr, BF_ERR_UNSUPPORTED, "bf_multi_unique_id: NCCL not available: " + api.error);
    NcclId nid;
    int rc = api.GetUniqueId(&nid);
    if (rc != 0) return fail(nullptr, BF_ERR_CUDA, "ncclGetUniqueId: " + nccl_err(api, rc));
    memcpy(id, &nid, sizeof(nid));
    return BF_OK;
}
int bf_multi_init(bf_scene *s, int rank, int world, const void *id, size_t id_bytes) {
    if (!s || !id || id_bytes < BF_MULTI_ID_BYTES || world < 1 || rank < 0 || rank >= world) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_multi_init: scene not committed");
    NcclApi &api = nccl_api();
    if (!api.handle) return fail(s, BF_ERR_UNSUPPORTED, "bf_multi_init: NCCL not available: " + api.error);
    BF_CUDA(cudaSetDevice(s->device));
    if (s->nccl_comm) { api.CommDestroy(s->nccl_comm); s->nccl_comm = nullptr; }
    NcclId nid; memcpy(&nid, id, sizeof(nid));
    int rc = api.CommInitRank(&s->nccl_comm, world, nid, rank);
    if (rc != 0) { s->nccl_comm = nullptr; return fail(s, BF_ERR_CUDA, "ncclCommInitRank: " + nccl_err(api, rc)); }
    return BF_OK;
}
int bf_multi_reduce_comm(bf_scene *s, void *nccl_comm, int root) {
    if (!s || !nccl_comm) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_multi_reduce: scene not committed");
    NcclApi &api = nccl_api();
    if (!api.handle) return fail(s, BF_ERR_UNSUPPORTED, "bf_multi_reduce: NCCL not available: " + api.error);
    BF_CUDA(cudaSetDevice(s->device));
    int rc = api.Reduce(s->d_profile.p, s->d_profile.p, s->profile_floats, /* ncclFloat32 */ 7, /* ncclSum */ 0, root, nccl_comm, s->stream);
    if (rc != 0) return fail(s, BF_ERR_CUDA, "ncclReduce: " + nccl_err(api, rc));
    return BF_OK;
}
int bf_multi_reduce(bf_scene *s, int root) {
    if (!s) return BF_ERR_INVALID;
    if (!s->nccl_comm) return fail(s, BF_ERR_STATE, "bf_multi_reduce: bf_multi_init has not been called");
    return bf_multi_reduce_comm(s, s->nccl_comm, root);
}

int bf_scene_set_stream(bf_scene *s, void *cuda_stream) {
    if (!s) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_scene_set_stream: the scene is not committed");
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(cudaStreamSynchronize(s->stream));     // nothing of this scene is left on the stream being given up
    cudaStream_t next = cuda_stream ? (cudaStream_t) cuda_stream : s->own_stream;
    if (cuda_stream) {
        // a handle that is not a stream of this device's context fails here, not inside a later launch
        const cudaError_t e = cudaStreamQuery(next);
        if (e != cudaSuccess && e != cudaErrorNotReady) { cudaGetLastError(); return fail(s, BF_ERR_INVALID, std::string("bf_scene_set_stream: not a usable stream (") + cudaGetErrorString(e) + ")"); }
        int dev = s->device;
        if (cudaStreamGetDevice(next, &dev) != cudaSuccess) cudaGetLastError();
        else if (dev != s->device) return fail(s, BF_ERR_INVALID, "bf_scene_set_stream: the stream belongs to another device than the scene");
    }
    if (s->l2_window_touched && s->stream != s->own_stream) {      // leave no access-policy window behind on a stream that is not ours
        cudaStreamAttrValue a;
        memset(&a, 0, sizeof(a));
        if (cudaStreamSetAttribute(s->stream, cudaStreamAttributeAccessPolicyWindow, &a) != cudaSuccess) cudaGetLastError();
    }
    s->stream = next;
    apply_l2_window(s, next);
    if (!s->slots.empty() && !s->slots[0].own_stream) s->slots[0].stream = next;      // wave slot 0 runs on the scene's stream
    return BF_OK;
}

int bf_synchronize(bf_scene *s) {
    if (!s || !s->committed) return BF_ERR_STATE;
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(cudaStreamSynchronize(s->stream));
    return BF_OK;
}

int bf_profile_read(bf_scene *s, float *dst, size_t n_floats) {
    if (!s || !dst) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_profile_read: scene not committed");
    if (n_floats < s->profile_floats) return fail(s, BF_ERR_INVALID, "bf_profile_read: destination too small");
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(cudaMemcpyAsync(dst, s->d_profile.p, s->profile_floats * 4, cudaMemcpyDeviceToHost, s->stream));
    BF_CUDA(cudaStreamSynchronize(s->stream));
    return BF_OK;
}

int bf_profile_develop(bf_scene *s, uint32_t component_format, void *dst, size_t dst_bytes) {
    if (!s || !dst) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_profile_develop: scene not committed");
    if (component_format != BF_COMPONENT_FLOAT32 && component_format != BF_COMPONENT_FLOAT16)
        return fail(s, BF_ERR_UNSUPPORTED, "bf_profile_develop: component_format must be float32 or float16");
    if (s->integ.gen != BF_GEN_PHASE)
        return fail(s, BF_ERR_UNSUPPORTED, "bf_profile_develop: only the HDRADC storage (Y, A, W + AOVs) of the phase integrator is developed; "
                                           "gen 1/2 films (X, Y, Z, A, W: HDRFilm's XYZ -> RGB develop, hdrfilm.cpp:304-307) are outside the path");
    const uint32_t C = s->dev.channels;
    if (C <= 3) return fail(s, BF_ERR_UNSUPPORTED, "bf_profile_develop: a storage without AOV channels has no developed form in the reference");
    const size_t n_bins = (size_t) s->dev.film_w * s->dev.film_h, esz = component_format == BF_COMPONENT_FLOAT16 ? 2 : 4;
    const size_t bytes = n_bins * (C - 1) * esz;
    if (dst_bytes < bytes) return fail(s, BF_ERR_INVALID, "bf_profile_develop: destination too small");
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(s->d_developed.alloc((bytes + 3) / 4));
    LaunchCfg cfg{s->stream, s->sm_count, 0};
    launch_develop(s->d_profile.p, s->d_developed.p, (uint32_t) n_bins, C, esz == 2, cfg);
    BF_CUDA(cudaGetLastError());
    BF_CUDA(cudaMemcpyAsync(dst, s->d_developed.p, bytes, cudaMemcpyDeviceToHost, s->stream));
    BF_CUDA(cudaStreamSynchronize(s->stream));
    return BF_OK;
}

int bf_profile_power_db(bf_scene *s, uint32_t channel, double sample_count, int mode, float *dst, size_t n_floats) {
    if (!s || !dst) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_profile_power_db: scene not committed");
    const uint32_t C = s->dev.channels, bins = (uint32_t) (s->profile_floats / C);
    if (channel >= C || !(sample_count > 0.0) || (mode != 0 && mode != 1)) return fail(s, BF_ERR_INVALID, "bf_profile_power_db: channel / sample_count / mode");
    if (n_floats < bins) return fail(s, BF_ERR_INVALID, "bf_profile_power_db: destination too small");
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(s->d_developed.alloc(bins));
    LaunchCfg cfg; memset(&cfg, 0, sizeof(cfg)); cfg.stream = s->stream; cfg.sm_count = s->sm_count;
    s->stats.kernel_launches += launch_power_db(s->d_profile.p, (float *) s->d_developed.p, bins, C, channel, sample_count, mode, cfg);
    BF_CUDA(cudaMemcpyAsync(dst, s->d_developed.p, (size_t) bins * 4, cudaMemcpyDeviceToHost, s->stream));
    BF_CUDA(cudaStreamSynchronize(s->stream));
    return BF_OK;
}

int bf_profile_device_ptr(bf_scene *s, void **dptr, size_t *n_floats, void **cuda_stream) {
    if (!s || !s->committed) return BF_ERR_STATE;
    if (dptr) *dptr = s->d_profile.p;
    if (n_floats) *n_floats = s->profile_floats;
    if (cuda_stream) *cuda_stream = (void *) s->stream;
    return BF_OK;
}

static int trace_rays_common(bf_scene *s, const float *origins, const float *directions, const float *mint, const float *maxt,
                             uint64_t n_rays, uint32_t *shape, uint32_t *prim, float *t, float *u, float *v, uint8_t *occl, uint32_t *steps = nullptr) {
    if (!s || !origins || !directions) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "trace: scene not committed");
    if (n_rays == 0) return BF_OK;
    if (n_rays > (1ull << 28)) return fail(s, BF_ERR_INVALID, "trace: too many rays in one call");
    BF_CUDA(cudaSetDevice(s->device));
    if (s->tables_dirty) { int rc = (s->tables_light && s->committed) ? upload_tables_light(s) : upload_tables(s); if (rc) return rc; }
    size_t n = (size_t) n_rays;
    std::vector<float4> ho(n), hd(n);
    for (size_t i = 0; i < n; ++i) {
        ho[i] = make_float4(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2], mint ? mint[i] : BF_RAY_EPS);
        hd[i] = make_float4(directions[3 * i], directions[3 * i + 1], directions[3 * i + 2], maxt ? maxt[i] : INFINITY);
    }
    DevBuf<float4> d_o, d_d, d_hit; DevBuf<uint32_t> d_a, d_b, d_cursor;
    auto cleanup = [&]() { d_o.release(); d_d.release(); d_hit.release(); d_a.release(); d_b.release(); d_cursor.release(); };
    cudaError_t e = cudaSuccess;
    if ((e = d_o.alloc(n)) || (e = d_d.alloc(n)) || (e = d_hit.alloc(n)) || (e = d_a.alloc(n)) || (e = d_b.alloc(n)) || (e = d_cursor.alloc(1))) {
        cleanup();
        return fail(s, BF_ERR_CUDA, std::string("trace: ") + cudaGetErrorString(e));
    }
    LaunchCfg cfg; cfg.stream = s->stream; cfg.sm_count = s->sm_count; cfg.trace_rounds = s->tune.trace_rounds >= 0 ? (uint32_t) s->tune.trace_rounds : s->auto_rounds;
    cudaMemcpyAsync(d_o.p, ho.data(), n * 16, cudaMemcpyHostToDevice, s->stream);
    cudaMemcpyAsync(d_d.p, hd.data(), n * 16, cudaMemcpyHostToDevice, s->stream);
    cudaMemsetAsync(d_cursor.p, 0, 4, s->stream);
    if (!occl) {
        s->stats.kernel_launches += launch_trace_rays(s->dev, d_o.p, d_d.p, (uint32_t) n, d_hit.p, steps ? d_b.p : nullptr, d_cursor.p, cfg);
        if (steps) cudaMemcpyAsync(steps, d_b.p, n * 4, cudaMemcpyDeviceToHost, s->stream);
        s->stats.kernel_launches += launch_resolve_hits(s->dev, d_hit.p, (uint32_t) n, d_a.p, d_b.p, cfg);
        std::vector<float4> hh(n);
        cudaMemcpyAsync(hh.data(), d_hit.p, n * 16, cudaMemcpyDeviceToHost, s->stream);
        if (shape) cudaMemcpyAsync(shape, d_a.p, n * 4, cudaMemcpyDeviceToHost, s->stream);
        if (prim) cudaMemcpyAsync(prim, d_b.p, n * 4, cudaMemcpyDeviceToHost, s->stream);
        e = cudaStreamSynchronize(s->stream);
        for (size_t i = 0; i < n && e == cudaSuccess; ++i) {
            if (t) t[i] = hh[i].x;
            if (u) u[i] = hh[i].y;
            if (v) v[i] = hh[i].z;
        }
    } else {
        s->stats.kernel_launches += launch_trace_rays(s->dev, d_o.p, d_d.p, (uint32_t) n, nullptr, d_a.p, d_cursor.p, cfg);
        std::vector<uint32_t> hh(n);
        cudaMemcpyAsync(hh.data(), d_a.p, n * 4, cudaMemcpyDeviceToHost, s->stream);
        e = cudaStreamSynchronize(s->stream);
        for (size_t i = 0; i < n && e == cudaSuccess; ++i) occl[i] = (uint8_t) hh[i];
    }
    if (e == cudaSuccess) e = cudaGetLastError();
    cleanup();
    if (e != cudaSuccess) return fail(s, BF_ERR_CUDA, std::string("trace: ") + cudaGetErrorString(e));
    return BF_OK;
}

int bf_trace_first_hits(bf_scene *s, const float *origins, const float *directions, const float *mint, const float *maxt,
                        uint64_t n_rays, uint32_t *shape, uint32_t *prim, float *t, float *u, float *v) {
    return trace_rays_common(s, origins, directions, mint, maxt, n_rays, shape, prim, t, u, v, nullptr);
}
// No allocation on this path: no DevBuf, no std::vector, no table upload (see abi.h).  The scratch array, its
// cudaGetSymbolAddress look-up and the kernels that use it are in bf_kernels.cu (static_trace_first_hits).
int bf_trace_first_hits_static(bf_scene *s, const float *origins, const float *directions, const float *mint, const float *maxt,
                               uint64_t n_rays, uint32_t *shape, uint32_t *prim, float *t, float *u, float *v) {
    if (!s || !origins || !directions) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_trace_first_hits_static: scene not committed");
    if (n_rays > BF_STATIC_TRACE_MAX_RAYS)
        return fail(s, BF_ERR_INVALID, "bf_trace_first_hits_static: at most 65536 rays per call (BF_STATIC_TRACE_MAX_RAYS, the size of the static scratch array)");
    if (s->device >= BF_STATIC_TRACE_MAX_DEVICES)
        return fail(s, BF_ERR_UNSUPPORTED, "bf_trace_first_hits_static: device ordinal beyond BF_STATIC_TRACE_MAX_DEVICES");
    if (s->tables_dirty)
        return fail(s, BF_ERR_STATE, "bf_trace_first_hits_static: scene edits pending; apply them with bf_receive or bf_trace_first_hits first (that may allocate)");
    if (n_rays == 0) return BF_OK;
    BF_CUDA(cudaSetDevice(s->device));
    LaunchCfg cfg; memset(&cfg, 0, sizeof(cfg)); cfg.stream = s->stream; cfg.sm_count = s->sm_count; cfg.trace_rounds = 8;
    int launches = 0;
    cudaError_t e = static_trace_first_hits(s->dev, s->device, origins, directions, mint, maxt, (uint32_t) n_rays, shape, prim, t, u, v, cfg, &launches);
    s->stats.kernel_launches += (uint64_t) launches;
    const cudaError_t e2 = cudaStreamSynchronize(s->stream);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) return fail(s, BF_ERR_CUDA, std::string("bf_trace_first_hits_static: ") + cudaGetErrorString(e));
    return BF_OK;
}
int bf_debug_trace_steps(bf_scene *s, const float *origins, const float *directions, const float *mint, const float *maxt,
                         uint64_t n_rays, uint32_t *steps, float *t) {
    if (!steps) return BF_ERR_INVALID;
    return trace_rays_common(s, origins, directions, mint, maxt, n_rays, nullptr, nullptr, t, nullptr, nullptr, nullptr, steps);
}
int bf_trace_occluded(bf_scene *s, const float *origins, const float *directions, const float *mint, const float *maxt,
                      uint64_t n_rays, uint8_t *hit) {
    if (!hit) return BF_ERR_INVALID;
    return trace_rays_common(s, origins, directions, mint, maxt, n_rays, nullptr, nullptr, nullptr, nullptr, nullptr, hit);
}

int bf_debug_sample_records(bf_scene *s, uint64_t seed, uint64_t first_sample, uint64_t n_samples, bf_sample_record *records) {
    if (!s || !records) return BF_ERR_INVALID;
    if (!s->committed) return fail(s, BF_ERR_STATE, "bf_debug_sample_records: scene not committed");
    if (n_samples == 0) return BF_OK;
    if (n_samples > (1u << 24)) return fail(s, BF_ERR_INVALID, "bf_debug_sample_records: at most 2^24 samples per call");
    BF_CUDA(cudaSetDevice(s->device));
    if (s->tables_dirty) { int rc = (s->tables_light && s->committed) ? upload_tables_light(s) : upload_tables(s); if (rc) return rc; }
    int rc = ensure_slot(s, 0, (uint32_t) std::max<uint64_t>(n_samples, 1024), true);
    if (rc) return rc;
    DevBuf<float> scratch;
    BF_CUDA(scratch.alloc(s->profile_floats));
    BF_CUDA(cudaMemsetAsync(scratch.p, 0, s->profile_floats * 4, s->stream));
    s->ev_trace_used = 0;
    rc = run_wave(s, s->slots[0], seed, first_sample, (uint32_t) n_samples, scratch.p, true);
    if (rc == BF_OK) {
        cudaError_t e = cudaMemcpyAsync(records, s->slots[0].wave.records, n_samples * sizeof(bf_sample_record), cudaMemcpyDeviceToHost, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = fail(s, BF_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaStreamSynchronize(s->stream);
    scratch.release();
    return rc;
}

int bf_get_stats(const bf_scene *s_, bf_stats *stats) {
    bf_scene *s = const_cast<bf_scene *>(s_);
    if (!s || !stats) return BF_ERR_INVALID;
    if (s->committed) {
        BF_CUDA(cudaSetDevice(s->device));
        unsigned long long t[4] = {0, 0, 0, 0};
        BF_CUDA(cudaStreamSynchronize(s->stream));
        BF_CUDA(cudaMemcpy(t, s->d_totals.p, 32, cudaMemcpyDeviceToHost));
        s->stats.closest_queries = t[0]; s->stats.shadow_queries = t[1];
        s->stats.path_closest_queries = t[2]; s->stats.path_shadow_queries = t[3];
    }
    *stats = s->stats;
    return BF_OK;
}

int bf_set_tunable(bf_scene *s, const char *name, int64_t value) {
    if (!s || !name) return BF_ERR_INVALID;
    struct { const char *n; int64_t *p; int64_t lo, hi; } table[] = {
        {"wave_lanes", &s->tune.wave_lanes, 0, 1ll << 28}, {"wave_streams", &s->tune.wave_streams, 1, 16},
        {"fused_primary", &s->tune.fused_primary, 0, 1}, {"primary_defer", &s->tune.primary_defer, -1, 1},
        {"trace_defer", &s->tune.trace_defer, -1, 1}, {"trace_rounds", &s->tune.trace_rounds, -1, 1024},
        {"use_graph", &s->tune.use_graph, 0, 1}, {"mega_batch", &s->tune.mega_batch, 0, 1},
        {"path_fused", &s->tune.path_fused, 0, 1}, {"mega_max", &s->tune.mega_max, -1, 1ll << 28},
        {"l2_persist_mb", &s->tune.l2_persist_mb, 0, 4096}, {"path_stash_max", &s->tune.path_stash_max, 1024, 1 << 20}};
    for (auto &t : table)
        if (strcmp(t.n, name) == 0) {
            if (value < t.lo || value > t.hi) return fail(s, BF_ERR_INVALID, std::string("bf_set_tunable: value out of range for ") + name);
            if (t.p == &s->tune.wave_lanes && value != 0 && value < 1024) value = 1024;
            *t.p = value;
            if (t.p == &s->tune.l2_persist_mb) apply_l2_persistence(s);
            return BF_OK;
        }
    return fail(s, BF_ERR_INVALID, std::string("bf_set_tunable: unknown tunable ") + name);
}

int bf_set_trace_timing(bf_scene *s, int enabled) {
    if (!s) return BF_ERR_INVALID;
    s->time_trace = enabled != 0;
    return BF_OK;
}

int bf_last_receive_ms(bf_scene *s, float *total_ms, float *traverse_ms) {
    if (!s || !s->committed) return BF_ERR_STATE;
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(cudaEventSynchronize(s->ev_end));
    float ms = 0.f;
    BF_CUDA(cudaEventElapsedTime(&ms, s->ev_begin, s->ev_end));
    if (total_ms) *total_ms = ms;
    if (traverse_ms) {
        float acc = 0.f;
        for (size_t i = 0; i + 1 < s->ev_trace_used; i += 2) {
            float m = 0.f;
            if (s->ev_kind[i] != BF_K_TRACE_CLOSEST && s->ev_kind[i] != BF_K_TRACE_SHADOW && s->ev_kind[i] != BF_K_PRIMARY) continue;
            if (cudaEventElapsedTime(&m, s->ev_trace[i], s->ev_trace[i + 1]) == cudaSuccess) acc += m;
        }
        *traverse_ms = s->time_trace ? acc : -1.f;
    }
    return BF_OK;
}

int bf_last_launch_list(bf_scene *s, float *ms, uint32_t *kind, uint32_t capacity, uint32_t *n_out) {
    if (!s || !s->committed || !ms || !kind || !n_out) return BF_ERR_STATE;
    if (!s->time_trace) return fail(s, BF_ERR_STATE, "bf_last_launch_list: enable bf_set_trace_timing first");
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(cudaEventSynchronize(s->ev_end));
    uint32_t n = 0;
    for (size_t i = 0; i + 1 < s->ev_trace_used && n < capacity; i += 2, ++n) {
        BF_CUDA(cudaEventElapsedTime(&ms[n], s->ev_trace[i], s->ev_trace[i + 1]));
        kind[n] = s->ev_kind[i];
    }
    *n_out = n;
    return BF_OK;
}

int bf_last_kernel_ms(bf_scene *s, float *ms, uint32_t *launches) {
    if (!s || !s->committed || !ms || !launches) return BF_ERR_STATE;
    if (!s->time_trace) return fail(s, BF_ERR_STATE, "bf_last_kernel_ms: enable bf_set_trace_timing first");
    BF_CUDA(cudaSetDevice(s->device));
    BF_CUDA(cudaEventSynchronize(s->ev_end));
    for (int k = 0; k < BF_K_COUNT; ++k) { ms[k] = 0.f; launches[k] = 0; }
    for (size_t i = 0; i + 1 < s->ev_trace_used; i += 2) {
        float m = 0.f;
        BF_CUDA(cudaEventElapsedTime(&m, s->ev_trace[i], s->ev_trace[i + 1]));
        ms[s->ev_kind[i]] += m; launches[s->ev_kind[i]]++;
    }
    return BF_OK;
}

}  // extern "C"
