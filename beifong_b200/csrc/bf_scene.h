// beifong_b200 — flattened scene: the POD tables the kernels read and the host object behind the
// C ABI handle.  Mirrors what Scene::Scene (src/librender/scene.cpp:22-120) and
// Shape::Shape (src/librender/shape.cpp:38-99) gather from the plugin tree.
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include <vector_types.h>

#include "../../include/beifong_b200/abi.h"
#include "../../include/beifong_b200/flat_scene.h"
#include "bf_math.h"
#include "bf_bvh.h"

namespace bf {

#define BF_MAX_RECTS 64u

// Rectangle (src/shapes/rectangle.cpp:83-95): everything derived from to_world at load time.
struct DevRect {
    M4 to_world, to_object;
    V3 fs, ft, fn;          // m_frame: dp_du, dp_dv (not normalised), normal
    V3 ns, nt, nn;          // normalize(frame.s/t/n), used by sample_wigner (rectangle.cpp:152-158)
    float area, inv_area;   // |s x t|
    float wid_x, wid_y;     // |s|, |t|
    uint32_t shape;         // scene shape id
    uint32_t pad;
    float box_lo[3], box_hi[3];   // world-space bounding box, padded: the traversal kernel's conservative pre-test
    // Per-rectangle constants the kernels would otherwise recompute for every sample (same fp32 operations, done once on the host):
    V3 cs_s, cs_t;          // coordinate_system(fn): the frame receivers build around the sampled position's normal
    V3 sh_s, sh_t;          // initialize_sh_frame for a hit on this rectangle (interaction.h:159-162): Gram-Schmidt of dp_du against fn
    float two_pi_inv_area;  // rcp(area) * 2 pi (wignertransmitter.cpp:528)
    float area_pi;          // area * pi        (wignerreceiver.cpp:262)
};
struct alignas(16) DevShape {   // 32 bytes: (type, index, bsdf, tx) is one 128-bit load
    uint32_t type;          // 0 mesh, 1 rectangle
    uint32_t index;         // mesh index / rectangle index
    uint32_t bsdf;
    int32_t tx;             // transmitter / emitter id attached to the shape, -1 none
    uint32_t has_normals;   // vertex normals present (Mesh::has_vertex_normals)
    uint32_t mat_class;     // 0: diffuse BSDF, 1: anything else (shade-queue split, bf_trace.cuh)
    uint32_t pad[2];
};
// Spectrum texture (src/spectra/uniform.cpp, regular.cpp, irregular.cpp)
struct DevSpectrum {
    uint32_t kind, n, offset;   // offset into the float pool: lambda[n] then value[n]
    float inv_interval;         // regular: 1 / ((hi - lo) / (n - 1))
};
// Virtual element (i, j) of a phased array (phasedtransmitter.cpp:119-167 == phasedreceiver.cpp:115-173):
// world -> element-local transform, the normalised element frame, r' = loc_i - loc_j and the steering phasor psi'.
struct DevVElem {
    M4 to_object;
    V3 ns, nt, nn;
    V3 r_dash;
    float psi_re, psi_im;
};
struct DevEndpoint {
    uint32_t kind, texture, receive_type;
    int32_t shape;              // parent shape id, -1 if free standing
    uint32_t velem_offset, n_velems;   // phased: slice of DevScene::velems (n_elems^2 entries)
    float wid_x, wid_y;         // phased: elem_dims.x / .y
    bf_signal_desc signal;
    float sig_period, sig_ti, sig_slope;   // rcp(repfreq), 0 + t_ext / 2, f_ext / t_ext: scene constants of eval_signal / sample_delta_frequency
    float adc_sampling_start, adc_sampling_time;
    float cutoff, inv_transition, cos_cutoff, cos_beam;   // spot.cpp:85-93
    float near_clip, far_clip, shutter_open, shutter_open_time;
    uint32_t film_width, film_height;
    M4 to_world, to_object, sample_to_camera;
};

// Passed to every kernel by value (constant bank): pointers to the tables + small constants.
// Discretised reconstruction filter (ReconstructionFilter::init_discretization, src/libcore/rfilter.cpp:9-20): the
// MTS_FILTER_RESOLUTION + 1 = 32 table values, built on the host.  Only read by the wide-filter finalize kernel.
struct DevFilter {
    float values[32];
    float radius, scale;      // m_radius, m_scale_factor = 31 / radius
    uint32_t n;               // footprint width: ceil((radius - 2 RayEpsilon) * 2) (signalblock.cpp:121)
    uint32_t border;          // m_border_size = ceil(radius - 0.5 - 2 RayEpsilon)
};

struct DevScene {
    bf_constants C;
    uint32_t S;
    uint32_t n_rects, n_shapes, n_txs, n_nodes, n_tris;
    const uint4 *nodes;        // 5 per node
    const float4 *tris;        // 3 per triangle (leaf order)
    const float4 *tri_nrm;     // 3 per triangle (vertex normals, leaf order) or null
    const DevRect *rects;
    const DevShape *shapes;
    const bf_bsdf_desc *bsdfs;
    const DevSpectrum *spectra;
    const float *spec_pool;
    const DevEndpoint *txs;
    const DevEndpoint *rx;
    const DevVElem *velems;
    const M4 *velocities;      // per shape: the `velocity` transform (shape.cpp:39-42); read only when integ.doppler is set
    bf_adc_desc adc;
    bf_integrator_desc integ;
    uint32_t film_w, film_h, channels;
    uint32_t rx_needs_aperture;
    // what the whole scene has in common, for the specialised shade kernels: the kind shared by every transmitter and by
    // every BSDF (front and back) of material class 1, or -1
    int32_t hint_tx_kind, hint_cls1_bsdf, hint_rx_kind;
    float mesh_lo[3], mesh_hi[3];   // padded world-space bounds of all mesh triangles: rays that miss them skip the traversal
    double phase_denom;        // (lambda_max - lambda_min) / 2 * 1e-9 evaluated like ray.h:92
    float adc_scale_t, adc_scale_f;   // (float) t_bins / t_bandwidth, (float) f_bins / f_bandwidth (integrator.cpp:1639)
    // Copies of the receiver, of the first transmitter and of their rectangles INSIDE the kernel parameter block: the
    // specialised kernels (one receiver / one transmitter of a known kind) read them as constant-bank operands instead of
    // chasing rx -> shape -> rect pointers through L1 for every sample.
    DevEndpoint rx_ep, tx0_ep;
    DevRect rx_rect, tx0_rect;
};

struct HostMesh {
    std::vector<float> pos;          // as given to bf_scene_add_mesh: the rest pose bf_scene_update_transform moves
    std::vector<uint32_t> idx;
    std::vector<float> nrm;
    uint32_t nv = 0, nf = 0;
    bool moved = false; M4 motion;   // bf_scene_update_transform: world position = motion * rest position
    float rest_lo[3], rest_hi[3];    // bounds of the rest pose
};
struct HostSpectrum {
    uint32_t kind = 0;
    std::vector<float> lambda, value;
};

DevRect make_rect(const float *to_world16, uint32_t shape);
void make_velems(const bf_array_desc &a, const bf_constants &C, std::vector<DevVElem> &out);
M4 m4_inverse(const M4 &a);

}  // namespace bf
