// beifong_b200 — wavefront state (SoA in HBM) and the kernel launchers.
//
// One lane = one radar path sample (one iteration of SamplingIntegrator::receive_sample,
// src/librender/integrator.cpp:1538-1667, or render_sample, :255-310).  A wave of lanes is carried
// through   raygen -> trace(closest) -> [ shade -> trace(shadow) + trace(closest) ]* -> bin scatter
// with lane ids compacted into queues between stages.  All arrays are indexed by lane id, 16-byte
// records so that every access is one 128-bit load/store.
#pragma once
#include <cuda_runtime.h>
#include "bf_scene.h"

namespace bf {

enum : uint32_t { LANE_ACTIVE = 1u, LANE_VALID = 2u, LANE_BS_DELTA = 4u, LANE_PENDING = 8u };

#define BF_MAX_BOUNCES 1024u
#define BF_TRACE_SM_STACK 8        // traversal stack entries kept in shared memory per thread
#define BF_TRACE_LOCAL_STACK 56    // deeper entries spill to local memory
#define BF_TRACE_STACK_DEPTH (BF_TRACE_SM_STACK + BF_TRACE_LOCAL_STACK)
#define BF_CTR_STRIDE 8u     // per-depth counter block: [0] next-closest count, [1] shadow count,
                             // [2] closest work cursor, [3] shadow work cursor, [4], [5] hit counts of that closest trace per material class

struct Wave {
    uint32_t capacity;
    float4 *ray_o;       // o.xyz, mint         (shared by the continuation and the shadow ray)
    float4 *ray_d;       // d.xyz, maxt         continuation ray
    float4 *sh_d;        // d.xyz, maxt         shadow ray
    float4 *hit;         // t, u, v, slot       closest-hit result (slot: leaf-order triangle index,
                         //                     0x80000000|rect index for rectangles)
    uint32_t *occl;      // any-hit result
    ulonglong2 *rng;     // PCG32 state, inc
    float4 *x1;          // time, phase, bs.pdf, scalar (gen 2 path length / gen 1 path time)
    float4 *x2;          // gen 3: sample time, f_rx ; gen 1/2: film position ; .w = ray weight (S = 1)
    float4 *spec[5];     // S = 1: [0] = (throughput, result, pending, wavelength)
                         // S = 4: throughput, result, pending, wavelengths, ray weight
    uint32_t *flags;
    uint32_t *queue[2];  // [0] lane-id queue of continuation rays (shade -> trace), [1] unused
    uint32_t *queue_h;   // lane-id queue of lanes whose continuation ray hit a diffuse surface (trace -> shade)
    uint32_t *queue_h2;  // ... any other material: a shading warp runs one BSDF class
    uint32_t *queue_s;   // lane-id queue for shadow rays
    uint32_t *queue_d;   // k_path: lane-id queue of receiver rays that enter the mesh bounds
    uint32_t *counters;  // BF_CTR_STRIDE * (BF_MAX_BOUNCES + 2)
    unsigned long long *totals;  // [0] closest queries traced, [1] shadow queries traced
    bf_sample_record *records;   // optional debug records, indexed by sample (= lane unless sample_of is set)
    uint32_t *sample_of;         // fused primary kernel: lane -> sample offset in the wave (lanes are compacted hits)
    uint32_t *n_lanes;           // fused primary kernel: device-side count of compacted lanes
};

struct LaunchCfg {
    cudaStream_t stream;
    int sm_count;
    uint32_t trace_rounds;   // k_trace: batches of 32 rays per active warp (0 = every warp of the grid takes part)
};

// Kernel launchers (bf_kernels.cu).  Return the number of kernels launched.
// raygen + first closest-hit query + compaction of the hits into lanes [0, *w.n_lanes); misses are binned on the spot
// defer_queue != nullptr: rays that enter the mesh bounds are not traversed here but queued (lane ids) for a
// launch_trace_closest(depth 1) over defer_queue / defer_count
int launch_primary(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, uint32_t *hit_count,
                   uint32_t *work_cursor, float *profile, uint32_t *defer_queue, uint32_t *defer_count, const LaunchCfg &cfg);
// k_path: raygen + the rectangle-only paths start to end in registers; everything else is spilled into the wave arrays
// (counter layout: ctr + 0 defer_queue length, + 2 work cursor, + 4 / + 5 shade(1) class queues, ctr + BF_CTR_STRIDE + 0 / + 1
// the continuation / shadow queues that shade(1) appends to).  path_kernel_available: the scene has an instantiation.
bool path_kernel_available(const DevScene &sc);
uint32_t path_stash_cap(uint32_t lanes, int sm_count, uint32_t cap_max);     // entries per CTA for a wave slot of `lanes` samples
size_t path_stash_bytes(const DevScene &sc, int sm_count, uint32_t cap);     // per wave slot: the CTAs' parking areas
int launch_path(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, uint32_t *ctr, float *profile,
                uint32_t *defer_queue, uint32_t *stash, uint32_t stash_cap, const LaunchCfg &cfg);
// k_mega: one thread carries one sample from the receiver ray to its bin (scenes without meshes; small requests)
int launch_mega(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, uint32_t *work_cursor, float *profile,
                const LaunchCfg &cfg);
int launch_raygen(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, const LaunchCfg &cfg);
// queue == nullptr: lanes [0, n_direct); else lane ids from queue[0 .. *count)
// depth = the shade iteration this trace feeds; lanes that hit are appended to w.queue_h / *hit_count
int launch_trace_closest(const DevScene &sc, const Wave &w, const uint32_t *queue, const uint32_t *count, uint32_t n_direct,
                         uint32_t depth, uint32_t *hit_count, uint32_t *work_cursor, const LaunchCfg &cfg, bool count_queries = true,
                         uint32_t *defer_queue = nullptr, uint32_t *defer_count = nullptr);
int launch_trace_shadow(const DevScene &sc, const Wave &w, const uint32_t *queue, const uint32_t *count, uint32_t *work_cursor,
                        const LaunchCfg &cfg);
// shades the lanes of w.queue_h[0 .. hit_counts[0]) and w.queue_h2[0 .. hit_counts[1]); n_bound >= their sum
int launch_shade(const DevScene &sc, const Wave &w, uint32_t depth, const uint32_t *hit_counts, uint32_t n_bound,
                 uint32_t *next_queue, uint32_t *ctr_block, const LaunchCfg &cfg);
// filter == nullptr: box filter (one bin per sample); otherwise the separable n x n splat of SignalBlock::put
int launch_finalize(const DevScene &sc, const Wave &w, uint32_t n, float *profile, const DevFilter *filter, const LaunchCfg &cfg);
// Moving meshes (bf_scene_update_transform): re-pose the triangles of one shape from their rest pose, then refit the wide BVH level by level
int launch_refit_tris(const float4 *rest, const float4 *rest_nrm, float4 *tris, float4 *nrm, uint32_t n_tris, uint32_t shape, const M4 &M, const M4 &Minv,
                      const LaunchCfg &cfg);
int launch_refit_nodes(uint4 *nodes, const float4 *tris, float *node_box, const uint32_t *level_first, uint32_t n_levels, float pad_abs, const LaunchCfg &cfg);
int launch_power_db(const float *raw, float *out, uint32_t n_bins, uint32_t channels, uint32_t ch, double sample_count, int mode, const LaunchCfg &cfg);
int launch_develop(const float *raw, void *out, uint32_t n_bins, uint32_t channels, bool half_out, const LaunchCfg &cfg);
// One item of a batched k_mega launch (bf_receive_batch): what differs from the committed scene
struct MegaItem {
    uint64_t seed, first_sample;
    float *profile;
    const DevRect *rects;      // this item's rectangle table and receiver (device)
    const DevEndpoint *rx;
    DevEndpoint rx_ep;         // ... and the copies DevScene carries by value
    DevRect rx_rect, tx0_rect;
};
// all items in one launch (grid.y = item); cursors: one zeroed word per item
int launch_mega_batch(const DevScene &sc, const MegaItem *items, uint32_t n_items, uint32_t n, uint32_t *cursors, unsigned long long *totals,
                      const LaunchCfg &cfg);
// Stand-alone ray queries (Scene::ray_intersect_preliminary / ray_test for caller-provided rays)
int launch_trace_rays(const DevScene &sc, const float4 *ray_o, const float4 *ray_d, uint32_t n, float4 *hit, uint32_t *occl,
                      uint32_t *work_cursor, const LaunchCfg &cfg);
// resolve (slot -> shape, prim) for hits produced by launch_trace_rays
int launch_resolve_hits(const DevScene &sc, const float4 *hit, uint32_t n, uint32_t *shape, uint32_t *prim, const LaunchCfg &cfg);
// the same query with all temporaries in bf_kernels.cu's static device array (bf_trace_first_hits_static): no allocation
cudaError_t static_trace_first_hits(const DevScene &sc, int device, const float *origins, const float *directions, const float *mint,
                                    const float *maxt, uint32_t n, uint32_t *shape, uint32_t *prim, float *t, float *u, float *v,
                                    const LaunchCfg &cfg, int *launches);

}  // namespace bf
