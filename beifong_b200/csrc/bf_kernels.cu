// beifong_b200 — sm_100a wavefront kernels of the radar path:
//   k_raygen    receiver ray generation       (integrator.cpp:1538-1569 / :255-279, receivers/*.cpp)
//   k_trace     8-wide compressed-BVH traversal + Möller-Trumbore + analytic rectangles
//               (Scene::ray_intersect_preliminary / ray_test, scene.cpp:129-178; mesh.h:194-226;
//                rectangle.cpp:229-263)
//   k_shade     surface interaction, emission, russian roulette, transmitter sampling (NEE) + MIS,
//               BSDF sampling, next-ray spawn, queue compaction
//               (pathtimefrequency.cpp:103-460, pathlength.cpp:114-325, pathtime.cpp:114-286)
//   k_finalize  phase/range/time binning + range-bin scatter
//               (phase.cpp:85-138, range.cpp:141-161, time.cpp:118-153, integrator.cpp:1588-1665,
//                signalblock.cpp:162-169)
// Compiled with -fmad=false: see bf_math.h for the arithmetic contract.
#include <cuda_fp16.h>
#include <mutex>
#include "bf_wave.h"
#include "bf_device.cuh"
#include "bf_trace.cuh"

namespace bf {

// =========================================================================================================
// helpers
// =========================================================================================================
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// warp-aggregated queue push: one atomic per warp, lane ids written in lane order
__device__ __forceinline__ void queue_push(bool pred, uint32_t value, uint32_t *queue, uint32_t *counter, unsigned long long *total) {
    uint32_t mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0) return;
    uint32_t leader = __ffs(mask) - 1, base = 0;
    if (lane_id() == leader) {
        base = atomicAdd(counter, (uint32_t) __popc(mask));
        if (total) atomicAdd(total, (unsigned long long) __popc(mask));
    }
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) queue[base + __popc(mask & ((1u << lane_id()) - 1u))] = value;
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// Deferred pair of warp-aggregated queue pushes: reserve() issues the two slot reservations (one atomic per queue per
// warp) and complete(), called one loop iteration later, reads their results and writes the lane ids, so the ~1 us
// round trip of a returning global atomic overlaps the next lane's shading.
struct DeferredPush2 {
    uint32_t base_a = 0, base_b = 0;   // atomic results, valid in the leader lanes
    uint32_t meta = 0;                 // bit 0 / 1: this lane pushes to a / b; bits 8..12 / 16..20: rank; bits 24..31 unused
    uint32_t value = 0;
    __device__ __forceinline__ void reserve(bool pred_a, uint32_t *counter_a, bool pred_b, uint32_t *counter_b, uint32_t v) {
        const uint32_t mask_a = __ballot_sync(0xffffffffu, pred_a), mask_b = __ballot_sync(0xffffffffu, pred_b);
        const uint32_t lane = lane_id(), lt = (1u << lane) - 1u;
        if (mask_a && lane == (uint32_t) (__ffs(mask_a) - 1)) base_a = atomic_add_async(counter_a, (uint32_t) __popc(mask_a));
        if (mask_b && lane == (uint32_t) (__ffs(mask_b) - 1)) base_b = atomic_add_async(counter_b, (uint32_t) __popc(mask_b));
        meta = (pred_a ? 1u : 0u) | (pred_b ? 2u : 0u) | ((uint32_t) __popc(mask_a & lt) << 8) | ((uint32_t) __popc(mask_b & lt) << 16);
        value = v;
    }
    __device__ __forceinline__ void complete(uint32_t *queue_a, uint32_t *queue_b) {
        const uint32_t mask_a = __ballot_sync(0xffffffffu, meta & 1u), mask_b = __ballot_sync(0xffffffffu, meta & 2u);
        if (mask_a) {
            const uint32_t base = __shfl_sync(0xffffffffu, base_a, __ffs(mask_a) - 1);
            if (meta & 1u) queue_a[base + ((meta >> 8) & 31u)] = value;
        }
        if (mask_b) {
            const uint32_t base = __shfl_sync(0xffffffffu, base_b, __ffs(mask_b) - 1);
            if (meta & 2u) queue_b[base + ((meta >> 16) & 31u)] = value;
        }
        meta = 0;
    }
};

__device__ __forceinline__ bf_sample_record *rec_of(const Wave &w, uint32_t l) { return w.records + (w.sample_of ? w.sample_of[l] : l); }

template <int S> struct SpecIO;
template <> struct SpecIO<1> {
    static __device__ __forceinline__ void load(const Wave &w, uint32_t l, Spec<1> &thr, Spec<1> &res, Spec<1> &pend, Spec<1> &wl) {
        float4 a = w.spec[0][l]; thr.v[0] = a.x; res.v[0] = a.y; pend.v[0] = a.z; wl.v[0] = a.w;
    }
    static __device__ __forceinline__ void store(const Wave &w, uint32_t l, const Spec<1> &thr, const Spec<1> &res, const Spec<1> &pend, const Spec<1> &wl) {
        w.spec[0][l] = make_float4(thr.v[0], res.v[0], pend.v[0], wl.v[0]);
    }
    static __device__ __forceinline__ Spec<1> load_weight(const Wave &w, uint32_t l, float4 x2) { Spec<1> r; r.v[0] = x2.w; return r; }
    static __device__ __forceinline__ void store_weight(const Wave &, uint32_t, const Spec<1> &wg, float4 &x2) { x2.w = wg.v[0]; }
};
template <> struct SpecIO<4> {
    static __device__ __forceinline__ Spec<4> ld(const float4 *p, uint32_t l) { float4 a = p[l]; Spec<4> r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; return r; }
    static __device__ __forceinline__ void st(float4 *p, uint32_t l, const Spec<4> &s) { p[l] = make_float4(s.v[0], s.v[1], s.v[2], s.v[3]); }
    static __device__ __forceinline__ void load(const Wave &w, uint32_t l, Spec<4> &thr, Spec<4> &res, Spec<4> &pend, Spec<4> &wl) {
        thr = ld(w.spec[0], l); res = ld(w.spec[1], l); pend = ld(w.spec[2], l); wl = ld(w.spec[3], l);
    }
    static __device__ __forceinline__ void store(const Wave &w, uint32_t l, const Spec<4> &thr, const Spec<4> &res, const Spec<4> &pend, const Spec<4> &wl) {
        st(w.spec[0], l, thr); st(w.spec[1], l, res); st(w.spec[2], l, pend); st(w.spec[3], l, wl);
    }
    static __device__ __forceinline__ Spec<4> load_weight(const Wave &w, uint32_t l, float4) { return ld(w.spec[4], l); }
    static __device__ __forceinline__ void store_weight(const Wave &w, uint32_t l, const Spec<4> &wg, float4 &) { st(w.spec[4], l, wg); }
};

// =========================================================================================================
// raygen
// =========================================================================================================
// One out-of-line copy of the TEA mixer (called twice per sample): halves its share of the primary kernel's code footprint.
__device__ __noinline__ uint64_t tea64_shared(uint64_t v0, uint64_t v1) { return sample_tea_64_wide(v0, v1); }

// What raygen produces for one sample: the receiver ray and the per-sample state of the path.
template <int S> struct PrimaryState {
    RayGen ray;
    PCG32 rng;
    Spec<S> wl, weight;
    float time;          // ray time
    float4 x2;           // gen 3: (sample time, f_rx, 0, weight[S == 1]); gen 1/2: (film position, 0, weight)
    double inv_wl;       // 1.0 / (wl[0] * 1e-9): the quotient behind f_rx, kept for the bin position of a sample that misses
};

// Head of receive_sample (integrator.cpp:1544-1569) / render_sample (:263-279) for sample `idx`.
template <int S, int GEN = -1, int RXK = -1>
__device__ __forceinline__ void raygen_sample(const DevScene &sc, uint64_t seed, uint64_t idx, PrimaryState<S> &ps) {
    PCG32 rng;
    rng.seed(tea64_shared(seed, idx), tea64_shared(idx, seed));   // sampler.cpp:89-92
    const DevEndpoint &rx = sc.rx_ep;
    V2 position_sample; position_sample.x = rng.next_float32(); position_sample.y = rng.next_float32();
    V2 aperture_sample(.5f, .5f);
    if (sc.rx_needs_aperture) { aperture_sample.x = rng.next_float32(); aperture_sample.y = rng.next_float32(); }
    float time;
    float4 x2 = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((GEN >= 0 ? (uint32_t) GEN : sc.integ.gen) == BF_GEN_PHASE) {
        // integrator.cpp:1544-1569
        time = rx.adc_sampling_start;
        if (rx.adc_sampling_time > 0.f) time += rng.next_float32() * rx.adc_sampling_time; else time = 0.f;
        float wavelength_sample = rng.next_float32();
        rx_sample_ray<S, RXK>(sc, time, wavelength_sample, position_sample, aperture_sample, ps.ray, ps.wl, ps.weight, ps.inv_wl);
        x2.x = time;
        x2.y = (float) ((double) sc.C.c * ps.inv_wl);   // f_rx of mix_resample: MTS_C*rcp(wavelengths*1e-9) (integrator.cpp:1603)
    } else {
        // integrator.cpp:263-279 (1x1 film, pos = 0)
        time = rx.shutter_open;
        if (rx.shutter_open_time > 0.f) time += rng.next_float32() * rx.shutter_open_time;
        float wavelength_sample = rng.next_float32();
        V2 adjusted((position_sample.x - 0.f) / (float) sc.film_w, (position_sample.y - 0.f) / (float) sc.film_h);
        rx_sample_ray<S, RXK>(sc, time, wavelength_sample, adjusted, aperture_sample, ps.ray, ps.wl, ps.weight, ps.inv_wl);
        x2.x = position_sample.x; x2.y = position_sample.y;
    }
    if (S == 1) x2.w = ps.weight.v[0];
    ps.rng = rng; ps.time = time; ps.x2 = x2;
}

// Writes the state of a freshly generated sample into lane l.
template <int S>
__device__ __forceinline__ void store_primary(const Wave &w, uint32_t l, const PrimaryState<S> &ps) {
    float4 x2 = ps.x2;
    SpecIO<S>::store_weight(w, l, ps.weight, x2);
    w.ray_o[l] = make_float4(ps.ray.o.x, ps.ray.o.y, ps.ray.o.z, ps.ray.mint);
    w.ray_d[l] = make_float4(ps.ray.d.x, ps.ray.d.y, ps.ray.d.z, ps.ray.maxt);
    w.rng[l] = make_ulonglong2(ps.rng.state, ps.rng.inc);
    w.x1[l] = make_float4(ps.time, 0.f, 0.f, 0.f);
    w.x2[l] = x2;
    SpecIO<S>::store(w, l, Spec<S>(1.f), Spec<S>(0.f), Spec<S>(0.f), ps.wl);
    w.flags[l] = LANE_ACTIVE;
}
__device__ __forceinline__ void init_record(bf_sample_record *rp, float wavelength0) {
    bf_sample_record r;
    r.first_shape = BF_NONE; r.first_prim = BF_NONE; r.first_t = bf_inf(); r.valid = 0; r.tf_t = 0.f; r.tf_f = 0.f;
    r.value = 0.f; r.phase = 0.f; r.n_closest = 1; r.n_shadow = 0; r.wavelength0 = wavelength0; r.depth = 0;
    *rp = r;
}

template <int S>
__global__ void __launch_bounds__(256) k_raygen(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, uint64_t seed, uint64_t first_sample, uint32_t n) {
    uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n) return;
    PrimaryState<S> ps;
    raygen_sample<S>(sc, seed, first_sample + l, ps);
    store_primary<S>(w, l, ps);
    if (w.records) init_record(w.records + l, ps.wl.v[0]);
}

// Where a sample lands on the film.  gen 3: (time, frequency) -> fractional ADC bin (integrator.cpp:1588-1639), `ran` is
// false for receive types the reference leaves unhandled; gen 1/2: the film position sample.
// f_d = MTS_C*rcp(wavelengths[0]*1e-9) in double for the sample's final wavelength (gen 3; unused otherwise).
__device__ __forceinline__ void sample_position(const DevScene &sc, float4 x2, double f_d, float &posx, float &posy, bool &ran, int &off_x, int &off_y) {
    ran = true; off_x = 0; off_y = 0;
    if (sc.integ.gen == BF_GEN_PHASE) {
        const DevEndpoint &rx = sc.rx_ep;
        float tfx = 0.f, tfy = 0.f;
        if (rx.receive_type == BF_RECV_MIX_RESAMPLE) {
            tfx = x2.x - rx.adc_sampling_start;
            tfy = (float) fabs(f_d - (double) x2.y);
        } else if (rx.receive_type == BF_RECV_RAW || rx.receive_type == BF_RECV_RAW_RESAMPLE) {
            tfx = x2.x - rx.adc_sampling_start;
            tfy = (float) f_d;
        } else {
            ran = false;
        }
        tfx *= sc.adc_scale_t;      // (float) t_bins / t_bandwidth
        tfy *= sc.adc_scale_f;
        posx = tfx; posy = tfy;
        off_x = (int) sc.adc.window_offset_t; off_y = (int) sc.adc.window_offset_f;
    } else {
        posx = x2.x; posy = x2.y;
    }
}
// SignalBlock::put, box filter of radius 0.5 (signalblock.cpp:162-169): destination pixel or -1
__device__ __forceinline__ int pixel_of(const DevScene &sc, float posx, float posy, int off_x, int off_y) {
    float px = posx - ((float) off_x - 0.f + .5f), py = posy - ((float) off_y - 0.f + .5f);
    float cx = ceilf(px - .5f), cy = ceilf(py - .5f);
    if (cx >= 0.f && cy >= 0.f && cx < (float) sc.film_w && cy < (float) sc.film_h) return (int) ((uint32_t) cy * sc.film_w + (uint32_t) cx);
    return -1;
}

// =========================================================================================================
// primary: raygen + first closest-hit query + compaction, fused
// =========================================================================================================
// One persistent kernel generates the receiver rays and traverses them (the loop of k_trace<false>, bf_trace.cuh).  A ray
// that leaves the scene ends its sample on the spot: all it contributes is a count in the weight channel of its bin
// (SignalBlock::put of a zero value), so nothing of it is ever written to HBM.  A ray that hits gets the next free lane:
// the wave state holds the surviving samples only, densely packed, and every later kernel reads and writes it coalesced.
#define PRIM_MISS_PIXELS 2048

struct PrimaryIO {
    uint64_t seed, first_sample;
    uint32_t n;
    uint32_t *work_cursor;
    uint32_t *hit_count;         // [2]: lengths of the two material-class queues
    uint32_t *n_lanes;           // number of lanes handed out
    unsigned long long *total;   // closest queries traced
    float *profile;
    uint32_t *defer_queue, *defer_count;   // DEFER: lanes whose ray enters the mesh bounds, traversed by the k_trace launch that follows
};

#ifndef BF_PRIMARY_CULL
#define BF_PRIMARY_CULL false
#endif
#ifndef BF_PRIMARY_MIN_BLOCKS
#define BF_PRIMARY_MIN_BLOCKS 8
#endif
// DEFER (scenes where few receiver rays reach the meshes): no traversal here at all.  A ray that misses the padded mesh
// bounds within [mint, nearest rectangle hit] is decided by the rectangles alone and committed as before; a ray that enters
// them gets its lane straight away and is queued for a k_trace<closest> launch over just those lanes (dense warps), which
// records the hit and feeds the shade queues, or leaves the lane a miss for k_finalize to count.  Every warp then generates
// 32 samples per round, and the traversal code is out of this kernel's instruction footprint.
template <int S, int GEN, int RXK, bool DEFER>
__global__ void __launch_bounds__(TRACE_THREADS, BF_PRIMARY_MIN_BLOCKS) k_primary(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, const __grid_constant__ PrimaryIO io) {
    __shared__ float s_rect[RECT_SM_FLOATS * BF_MAX_RECTS];
    __shared__ uint2 s_stack[TRACE_SM_STACK * TRACE_THREADS];
    __shared__ uint32_t s_miss[PRIM_MISS_PIXELS];
    uint2 l_stack[TRACE_LOCAL_STACK];
    const uint32_t n_pixels = sc.film_w * sc.film_h, C = sc.channels;
    const bool miss_smem = n_pixels <= PRIM_MISS_PIXELS;
    const uint32_t w_channel = (sc.integ.gen == BF_GEN_PHASE ? 3u : 5u) - 1u;
    if (miss_smem) for (uint32_t i = threadIdx.x; i < n_pixels; i += blockDim.x) s_miss[i] = 0u;
    stage_rects(sc, s_rect);
    const uint32_t n = io.n;
    if (io.total && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(io.total, (unsigned long long) n);
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;

    // work distribution as in k_trace: a static first batch, then chunks claimed one ahead
    const uint32_t n_warps = gridDim.x * (TRACE_THREADS / 32);
    const uint32_t static_total = n_warps * 32u;
    const bool dynamic = n > static_total;
    uint32_t chunk = 32;
    while (dynamic && chunk < 256 && chunk * 16 * n_warps <= n - static_total) chunk *= 2;
    uint32_t c_next = (blockIdx.x * (TRACE_THREADS / 32) + (threadIdx.x >> 5)) * 32u, c_end = c_next + 32u;   // warp-uniform
    uint32_t claim = 0xffffffffu;
    if (dynamic && lane == 0) claim = static_total + atomic_add_async(io.work_cursor, chunk);
    bool out_of_work = c_next >= n;

    enum { EMPTY = 0, BUSY = 1, FINISHED = 2, EXHAUSTED = 3 };
    int state = EMPTY;
    uint32_t smp = 0;            // sample offset in the wave
    bool deferred = false;       // DEFER: the ray needs the BVH
    TraceLane L;
    PrimaryState<S> ps;
    while (true) {
        // ---- commit finished samples (warp converged here) ----
        const bool fin = state == FINISHED;
        const bool hit = fin && (L.best.slot != BF_NONE || (DEFER && deferred));      // = gets a lane
        if (fin && !hit) {
            float posx, posy; bool ran; int off_x, off_y;
            sample_position(sc, ps.x2, (double) sc.C.c * ps.inv_wl, posx, posy, ran, off_x, off_y);
            const int pixel = pixel_of(sc, posx, posy, off_x, off_y);
            if (pixel >= 0) {
                if (miss_smem) atomicAdd(&s_miss[pixel], 1u);
                else atomicAdd(io.profile + (size_t) C * pixel + w_channel, 1.f);
            }
            if (w.records) {
                bf_sample_record *r = w.records + smp;
                init_record(r, ps.wl.v[0]);
                r->depth = 1; r->tf_t = posx; r->tf_f = posy;
            }
        }
        const uint32_t hits = __ballot_sync(0xffffffffu, hit);
        if (hits) {
            const bool shade_now = hit && !(DEFER && deferred);
            const uint32_t cls = shade_now ? sc.shapes[L.best.shape].mat_class : 0u;
            const uint32_t m0 = __ballot_sync(0xffffffffu, shade_now && cls == 0u), m1 = __ballot_sync(0xffffffffu, shade_now && cls != 0u);
            const uint32_t md = DEFER ? hits & ~(m0 | m1) : 0u;
            // the reservations are in flight together (lane slots, the two class queues, the traversal queue), one wait
            uint32_t base = 0;
            if (lane == 0) base = atomic_add_async(io.n_lanes, (uint32_t) __popc(hits));
            else if (lane == 1 && m0) base = atomic_add_async(io.hit_count + 0, (uint32_t) __popc(m0));
            else if (lane == 2 && m1) base = atomic_add_async(io.hit_count + 1, (uint32_t) __popc(m1));
            else if (DEFER && lane == 3 && md) base = atomic_add_async(io.defer_count, (uint32_t) __popc(md));
            const uint32_t base_l = __shfl_sync(0xffffffffu, base, 0), base_0 = __shfl_sync(0xffffffffu, base, 1), base_1 = __shfl_sync(0xffffffffu, base, 2);
            const uint32_t base_d = DEFER ? __shfl_sync(0xffffffffu, base, 3) : 0u;
            if (hit) {
                const uint32_t dl = base_l + __popc(hits & lt_mask);
                store_primary<S>(w, dl, ps);
                if (w.sample_of) w.sample_of[dl] = smp;
                if (!shade_now) {
                    io.defer_queue[base_d + __popc(md & lt_mask)] = dl;
                } else {
                    w.hit[dl] = make_float4(L.best.t, L.best.u, L.best.v, __uint_as_float(L.best.slot));
                    if (cls == 0u) w.queue_h[base_0 + __popc(m0 & lt_mask)] = dl;
                    else w.queue_h2[base_1 + __popc(m1 & lt_mask)] = dl;
                }
                if (w.records) init_record(w.records + smp, ps.wl.v[0]);
            }
        }
        if (fin) state = EMPTY;
        // ---- refill empty lanes from the open chunk, running over into the chunk claimed ahead ----
        uint32_t want = __ballot_sync(0xffffffffu, state == EMPTY);
        if (want) {
            const uint32_t cnt = __popc(want), rank = __popc(want & lt_mask);
            const uint32_t avail = out_of_work ? 0u : c_end - c_next;
            uint32_t i = rank < avail ? c_next + rank : 0xffffffffu;
            if (!out_of_work && cnt > avail) {
                const uint32_t nxt = __shfl_sync(0xffffffffu, claim, 0);
                if (nxt < n) {
                    if (lane == 0) claim = static_total + atomic_add_async(io.work_cursor, chunk);
                    if (rank >= avail) i = nxt + (rank - avail);
                    c_next = nxt + (cnt - avail); c_end = nxt + chunk;
                } else {
                    out_of_work = true;
                }
            } else if (!out_of_work) {
                c_next += cnt;
            }
            if (!out_of_work && c_next >= n) out_of_work = true;
            if (state == EMPTY) {
                if (i < n) {
                    smp = i;
                    raygen_sample<S, GEN, RXK>(sc, io.seed, io.first_sample + i, ps);
                    L.o = ps.ray.o; L.d = ps.ray.d; L.mint = ps.ray.mint; L.maxt = ps.ray.maxt;
                    if (DEFER) { deferred = trace_begin<false, true>(sc, s_rect, L); state = FINISHED; }
                    else state = trace_begin<false, BF_PRIMARY_CULL>(sc, s_rect, L) ? BUSY : FINISHED;
                } else {
                    state = EXHAUSTED;
                }
            }
        }
        const uint32_t busy = __ballot_sync(0xffffffffu, state == BUSY);
        const uint32_t finm = __ballot_sync(0xffffffffu, state == FINISHED);
        if (busy == 0 && finm == 0) break;
        if (DEFER || busy == 0) continue;
        const int min_busy = out_of_work ? 1 : TRACE_REFILL;
        if (BF_PRIMARY_CULL && !out_of_work && finm != 0 && __popc(busy) < min_busy) continue;
        do {
            if (state == BUSY) {
                if (!trace_step<false>(sc, s_stack, l_stack, L)) state = FINISHED;
            }
        } while (__popc(__ballot_sync(0xffffffffu, state == BUSY)) >= min_busy);
    }
    if (miss_smem) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n_pixels; i += blockDim.x) {
            const uint32_t v = s_miss[i];
            if (v) atomicAdd(io.profile + (size_t) C * i + w_channel, (float) v);
        }
    }
}


__global__ void k_resolve_hits(const __grid_constant__ DevScene sc, const float4 *__restrict__ hit, uint32_t n, uint32_t *shape, uint32_t *prim) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 h = hit[i];
    uint32_t slot = __float_as_uint(h.w);
    if (h.x == bf_inf() || slot == BF_NONE) { shape[i] = BF_NONE; prim[i] = BF_NONE; return; }
    if (slot & 0x80000000u) { shape[i] = sc.rects[slot & 0x7fffffffu].shape; prim[i] = 0; return; }
    shape[i] = __float_as_uint(__ldg(&sc.tris[3 * (size_t) slot]).w);
    prim[i] = __float_as_uint(__ldg(&sc.tris[3 * (size_t) slot + 1]).w);
}

// =========================================================================================================
// shade
// =========================================================================================================
// Ray::update_state (ray.h:89-93): time += t / MTS_C; phase += 2 pi t / ((MAX-MIN)/2*1e-9) (double)
__device__ __forceinline__ void update_state(const DevScene &sc, float &time, float &phase, float t) {
    time += t / sc.C.c;
    phase = (float) ((double) phase + (double) (BF_TWO_PI * t) / sc.phase_denom);
}

// Register-resident state of one lane between two shade steps (what the wave arrays hold in HBM).
template <int S> struct LaneRegs {
    uint32_t flags;
    V3 o, d;                 // the ray whose closest hit is `hit`
    float4 hit;              // t, u, v, slot
    PCG32 rng;
    float time, phase, bs_pdf, scalar;   // x1
    Spec<S> thr, res, pend, wl;
};
// What one shade step hands to the ray queries that follow it.
struct ShadeOut {
    bool push_next, push_shadow;
    V3 origin; float mint;   // Interaction::spawn_ray: shared by the continuation and the shadow ray
    V3 new_d;                // continuation ray direction (maxt = inf)
    V3 sh_d; float sh_maxt;  // shadow ray
};

// One iteration of the integrator loop for one lane: the tail of the previous iteration (time / phase update, emitter MIS)
// and the body (emission, russian roulette, NEE, BSDF sampling) of PathTimeFrequencyIntegrator::sample
// (pathtimefrequency.cpp:154-433) and its PathLength / PathTime twins.  Shared by k_shade (state in the wave arrays) and
// k_path (state in registers).
// GEN / TXK / BK >= 0: the integrator generation, the transmitter kind and the BSDF kind are compile-time constants (the
// device functions are then inlined with their other branches folded away, which keeps the executed code inside the 32 KB
// instruction cache and the intermediate structs in registers); -1: read from the scene (out-of-line generic functions).
// TX_HIT = false: the caller guarantees that the surface hit carries no transmitter (k_path parks only such hits).
// DOPPLER = false: the kernel is never launched for a scene with the integrator's doppler flag set (the specialised instantiations).
template <int S, int GEN, int TXK, int BK, bool TX_HIT = true, bool DOPPLER = true>
__device__ __forceinline__ void shade_core(const DevScene &sc, uint32_t depth, LaneRegs<S> &R, bf_sample_record *rec, ShadeOut &out) {
    const uint32_t gen = GEN >= 0 ? (uint32_t) GEN : sc.integ.gen;
    bool push_next = false, push_shadow = false;
    uint32_t flags = R.flags;
    Spec<S> &thr = R.thr, &res = R.res, &pend = R.pend, &wl = R.wl;
    PCG32 &rng = R.rng;
    const V3 o = R.o, d = R.d;
    SI si = (GEN >= 0) ? compute_si_impl(sc, o, d, R.hit) : compute_si(sc, o, d, R.hit);
    float time = R.time, phase = R.phase, bs_pdf = R.bs_pdf, scalar = R.scalar;
    float emission_weight = 1.f;
    // (type, index, bsdf, tx) of the shape hit: one 128-bit load
    const uint4 shp = si.is_valid() ? __ldg(reinterpret_cast<const uint4 *>(sc.shapes + si.shape)) : make_uint4(0u, 0u, 0u, 0xffffffffu);
    const int transmitter = TX_HIT ? (int) shp.w : -1;

    if (depth == 1) {
        if (si.is_valid()) flags |= LANE_VALID;
        if (gen == BF_GEN_PHASE) { if (si.is_valid()) update_state(sc, time, phase, -si.t); }
        else if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;
        else scalar = si.is_valid() ? si.t / 3.0e8f : 0.f;
        if (rec) {
            rec->first_shape = si.is_valid() ? si.shape : BF_NONE;
            rec->first_prim = si.is_valid() ? si.prim : BF_NONE;
            rec->first_t = si.t;
        }
    } else {
        // tail of the previous loop iteration (pathtimefrequency.cpp:368-399)
        if (gen == BF_GEN_PHASE) { phase = 0.f; update_state(sc, time, phase, -si.t); }
        if (transmitter >= 0) {
            DirectionSample ds;   // DirectionSample3f(si_bsdf, si), records.h:168-174
            ds.p = si.p; ds.n = si.sh.n; ds.time = time; ds.delta = false; ds.pdf = 0.f;
            ds.d = si.p - o; ds.dist = norm(ds.d); ds.d = ds.d / ds.dist; ds.object = transmitter;
            float tx_pdf = 0.f;
            if (!(flags & LANE_BS_DELTA)) {
                tx_pdf = tx_pdf_direction(sc, sc.txs[transmitter], wl.v[0], ds);
                if (sc.n_txs > 1) tx_pdf *= (1.f / (float) sc.n_txs);
            }
            emission_weight = mis_weight(bs_pdf, tx_pdf);
        }
        if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;
        else if (gen == BF_GEN_TIME) scalar += si.is_valid() ? si.t / 3.0e8f : 0.f;
    }

    // ---- loop body (pathtimefrequency.cpp:154-364) ----
    bool active = true;
    if (transmitter >= 0) {
        Spec<S> e = tx_eval<S>(sc, sc.txs[transmitter], si, time, wl);
#pragma unroll
        for (int k = 0; k < S; ++k) res.v[k] += emission_weight * thr.v[k] * e.v[k];
        if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;      // pathlength.cpp:161
    }
    active &= si.is_valid();
    if ((int) depth > sc.integ.rr_depth) {
        float q = fminf(hmax_spec<S>(thr) * 1.f, .95f);
        active &= rng.next_float32() < q;
        float iq = rcp(q);
#pragma unroll
        for (int k = 0; k < S; ++k) thr.v[k] *= iq;
    }
    bool go_on = !(depth >= (uint32_t) sc.integ.max_depth || !active);
    V3 new_d(0.f, 0.f, 0.f);
    out.sh_d = V3(0.f, 0.f, 0.f); out.sh_maxt = 0.f;
    if (go_on) {
        const uint32_t bsdf = shp.z;
        bool active_e = (BK == BF_BSDF_DIFFUSE || BK == BF_BSDF_ROUGHCONDUCTOR) ? true : bsdf_is_smooth(sc.bsdfs[bsdf]);
        // Diffuse class: the twosided front / back choice and the reflectance texture are evaluated ONCE per lane (the
        // reference's eval(), pdf() and sample() each repeat them on the same wi and, after NEE, the same wavelengths)
        DiffuseLane<S> dl;
        if constexpr (BK == BF_BSDF_DIFFUSE) dl.select(sc, bsdf, si.wi);
        if (active_e) {
            // Scene::sample_transmitter_direction (scene.cpp:249-299)
            V2 sample; sample.x = rng.next_float32(); sample.y = rng.next_float32();
            DirectionSample ds; Spec<S> tspec;
            ds.pdf = 0.f; ds.delta = false; ds.dist = 0.f;
            if (TXK >= 0 || sc.n_txs == 1) {        // TXK >= 0: launched for single-transmitter scenes only
                tspec = tx_sample_direction<S, TXK>(sc, 0u, si.p, time, wl, sample, ds);
            } else if (sc.n_txs > 1) {
                float nt = (float) sc.n_txs, transmitter_pdf = 1.f / nt;
                uint32_t index = min((uint32_t) (sample.x * nt), sc.n_txs - 1u);
                sample.x = (sample.x - index * transmitter_pdf) * nt;
                tspec = tx_sample_direction<S, -1>(sc, index, si.p, time, wl, sample, ds);
                ds.pdf *= transmitter_pdf;
                float inv = rcp(transmitter_pdf);
#pragma unroll
                for (int k = 0; k < S; ++k) tspec.v[k] *= inv;
            }
            if (DOPPLER && gen == BF_GEN_PHASE && sc.integ.doppler) {
                // pathtimefrequency.cpp:252-256, commented out in the reference ("Took doppler out to test"), behind the flag
                const Spec<S> shift = shape_doppler<S>(sc, si, wl);
#pragma unroll
                for (int k = 0; k < S; ++k) wl.v[k] += shift.v[k];
            }
            active_e &= ds.pdf != 0.f;
            if (active_e) {
                V3 wo = si.to_local(ds.d);
                Spec<S> bsdf_val; float bpdf;
                if constexpr (BK == BF_BSDF_DIFFUSE) {
                    dl.texture(sc, wl);
                    bsdf_val = dl.eval(wo); bpdf = dl.pdf(wo);
                } else {
                    bsdf_val = bsdf_eval<S, BK>(sc, bsdf, wl, si.wi, wo);
                    bpdf = bsdf_pdf<BK>(sc, bsdf, si.wi, wo);
                }
                float mis = ds.delta ? 1.f : mis_weight(ds.pdf, bpdf);
                bool nonzero = false;
#pragma unroll
                for (int k = 0; k < S; ++k) { pend.v[k] = mis * thr.v[k] * bsdf_val.v[k] * tspec.v[k]; nonzero |= pend.v[k] != 0.f; }
                if (rec) rec->n_shadow++;
                if (nonzero) {   // a zero contribution needs no visibility test
                    flags |= LANE_PENDING;
                    push_shadow = true;
                    out.sh_d = ds.d; out.sh_maxt = ds.dist * (1.f - BF_SHADOW_EPS);
                }
            }
            if (gen == BF_GEN_RANGE) scalar += si.is_valid() ? si.t : 0.f;  // pathlength.cpp:209
        }
        // BSDF sampling (pathtimefrequency.cpp:346-364)
        float s1 = rng.next_float32(); (void) s1;
        V2 s2; s2.x = rng.next_float32(); s2.y = rng.next_float32();
        BSDFSample bs;
        Spec<S> bsdf_val;
        if constexpr (BK == BF_BSDF_DIFFUSE) { dl.texture(sc, wl); bsdf_val = dl.sample(s2, bs); }
        else bsdf_val = bsdf_sample<S, BK>(sc, bsdf, wl, si.wi, s2, bs);
        bool any_nz = false;
#pragma unroll
        for (int k = 0; k < S; ++k) { thr.v[k] = thr.v[k] * bsdf_val.v[k]; any_nz |= thr.v[k] != 0.f; }
        active &= any_nz;
        if (active) {
            new_d = si.to_world(bs.wo);
            bs_pdf = bs.pdf;
            flags = bs.delta ? (flags | LANE_BS_DELTA) : (flags & ~LANE_BS_DELTA);
            push_next = true;
        }
        go_on = active;
    }
    // Interaction::spawn_ray (interaction.h:61-64); the shadow ray shares origin and mint
    out.origin = si.p; out.mint = (1.f + hmax_abs(si.p)) * BF_RAY_EPS;
    out.new_d = new_d;
    out.push_next = push_next; out.push_shadow = push_shadow;
    if (!go_on) flags &= ~LANE_ACTIVE;
    R.flags = flags;
    R.time = time; R.phase = phase; R.bs_pdf = bs_pdf; R.scalar = scalar;
    if (rec) {
        rec->depth = depth;
        if (push_next) rec->n_closest++;
    }
}

#ifndef BF_SHADE_MIN_BLOCKS
#define BF_SHADE_MIN_BLOCKS 5
#endif
// QSEL: 0 / 1 = shade the lanes of one material-class queue, 2 = both.  GEN / TXK / BK: see shade_core.
template <int S, int GEN, int TXK, int BK, int QSEL>
__global__ void __launch_bounds__(128, BF_SHADE_MIN_BLOCKS) k_shade(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, uint32_t depth,
                                               const uint32_t *__restrict__ hit_counts, uint32_t *__restrict__ next_queue, uint32_t *ctr) {
    // two input queues, one per material class; the second starts on a warp boundary of the virtual index space
    const uint32_t n_a = QSEL == 1 ? 0u : hit_counts[0], n_b = QSEL == 0 ? 0u : hit_counts[1];
    const uint32_t n_a_round = (n_a + 31u) & ~31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = n_a_round + ((n_b + 31u) & ~31u);   // keep warps converged for the ballot in queue_push
    // lane id of virtual index i, or BF_NONE in the padding between / after the two queues
    auto lane_of = [&](uint32_t i) -> uint32_t {
        const bool second = i >= n_a_round;
        const uint32_t j = second ? i - n_a_round : i;
        if (i >= n_round || j >= (second ? n_b : n_a)) return BF_NONE;
        return second ? w.queue_h2[j] : w.queue_h[j];
    };
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t l_cur = lane_of(i), l_next = lane_of(i + stride);
    DeferredPush2 dpush;
    for (; i < n_round; i += stride) {
        bool push_next = false, push_shadow = false;
        const uint32_t l = l_cur;
        // queue entries are read two iterations ahead; the records of the next lane are pulled into L2 while this one
        // is shaded (queue entry -> records -> triangle is a chain of dependent DRAM round trips otherwise)
        const uint32_t l_after = lane_of(i + 2u * stride);
        if (l_next != BF_NONE) {
            prefetch_l2(w.flags + l_next); prefetch_l2(w.ray_o + l_next); prefetch_l2(w.ray_d + l_next); prefetch_l2(w.hit + l_next);
            prefetch_l2(w.x1 + l_next); prefetch_l2(w.rng + l_next);
#pragma unroll
            for (int k = 0; k < (S == 1 ? 1 : 4); ++k) prefetch_l2(w.spec[k] + l_next);
        }
        if (l != BF_NONE) {
            LaneRegs<S> R;
            R.flags = w.flags[l];
            const float4 ro = w.ray_o[l], rd = w.ray_d[l];
            R.hit = w.hit[l];
            const float4 x1 = w.x1[l];
            { ulonglong2 r = w.rng[l]; R.rng.state = r.x; R.rng.inc = r.y; }
            SpecIO<S>::load(w, l, R.thr, R.res, R.pend, R.wl);
            if (R.flags & LANE_PENDING) {
                if (!w.occl[l]) {
#pragma unroll
                    for (int k = 0; k < S; ++k) R.res.v[k] += R.pend.v[k];
                }
                R.flags &= ~LANE_PENDING;
            }
            R.o = V3(ro.x, ro.y, ro.z); R.d = V3(rd.x, rd.y, rd.z);
            R.time = x1.x; R.phase = x1.y; R.bs_pdf = x1.z; R.scalar = x1.w;
            ShadeOut so;
            shade_core<S, GEN, TXK, BK, true, (GEN < 0)>(sc, depth, R, w.records ? rec_of(w, l) : nullptr, so);      // doppler scenes run the generic kernel
            push_next = so.push_next; push_shadow = so.push_shadow;
            if (push_next || push_shadow) {
                w.ray_o[l] = make_float4(so.origin.x, so.origin.y, so.origin.z, so.mint);
                if (push_next) w.ray_d[l] = make_float4(so.new_d.x, so.new_d.y, so.new_d.z, bf_inf());
                if (push_shadow) w.sh_d[l] = make_float4(so.sh_d.x, so.sh_d.y, so.sh_d.z, so.sh_maxt);
            }
            w.flags[l] = R.flags;
            w.x1[l] = make_float4(R.time, R.phase, R.bs_pdf, R.scalar);
            SpecIO<S>::store(w, l, R.thr, R.res, R.pend, R.wl);
            w.rng[l] = make_ulonglong2(R.rng.state, R.rng.inc);
        }
        dpush.complete(w.queue_s, next_queue);       // the previous iteration's pushes
        dpush.reserve(push_shadow, ctr + 1, push_next, ctr + 0, l);
        l_cur = l_next; l_next = l_after;
    }
    dpush.complete(w.queue_s, next_queue);
}

// =========================================================================================================
// finalize: wrappers' binning + SignalBlock::put / ImageBlock::put (single-bin box filter)
// =========================================================================================================
#define FIN_THREADS 256
#define FIN_SMEM_FLOATS 8192

// fmodf(x, 2 pi) with fmodf's exact result, through one double division instead of the library's subtract-and-shift
// loop (|x / 2 pi| reaches 1e5 for the carrier phase): q = trunc(|x| / y) is off by at most one, q * y is exact in
// double (27 + 24 bits), and the corrected remainder is a float-representable number, so no step rounds.
__device__ __forceinline__ float fmod_two_pi(float x) {
    const float ax = fabsf(x);
    if (!(ax < 1e8f)) return fmodf(x, BF_TWO_PI);      // inf / nan / huge: library path
    const double y = (double) BF_TWO_PI;
    const double q = trunc((double) ax / y);
    double r = fma(-q, y, (double) ax);
    if (r < 0.0) r += y; else if (r >= y) r -= y;
    return copysignf((float) r, x);
}

// Warp-aggregated float scatter: lanes with key >= 0 and val != 0 add val to dst[key .. key + rep).  Lanes that
// share a key are summed in registers first (one pass over the set bits of the participation ballot), and the
// lowest lane of each key issues the atomics.  Used when the lanes of a warp land in the same pixel (1x1 films).
__device__ __forceinline__ void agg_add(float *dst, int key, float val, uint32_t rep) {
    const bool mine = key >= 0 && val != 0.f;
    uint32_t todo = __ballot_sync(0xffffffffu, mine);
    if (todo == 0) return;
    float acc = 0.f;
    bool first = true;
    const uint32_t lane = lane_id();
    while (todo) {
        const int i = __ffs(todo) - 1;
        todo &= todo - 1;
        const int k = __shfl_sync(0xffffffffu, key, i);
        const float v = __shfl_sync(0xffffffffu, val, i);
        if (k == key) { acc += v; if ((uint32_t) i < lane) first = false; }
    }
    if (mine && first)
        for (uint32_t r = 0; r < rep; ++r) atomicAdd(dst + key + r, acc);
}

// What one sample adds to the profile (the tail of receive_sample / render_sample).
struct Splat {
    int pixel;          // destination pixel or -1 (box filter)
    float posx, posy;   // fractional sample position and the block offset (wide filters)
    int off_x, off_y;
    float val_a;        // weighted value: channel Y (gen 3) or X = Y = Z (gen 1/2)
    bool cnt_valid;     // alpha channel
    int var_ch;         // the single non-zero wrapper channel of this sample, or -1
    float var_val;
    uint32_t var_rep;   // gen 1 writes the value to 3 consecutive channels
};

// res: the path's radiance with the pending NEE term already resolved; phase / scalar: x1.y / x1.w of the lane
template <int S>
__device__ __forceinline__ Splat finalize_core(const DevScene &sc, uint32_t flags, float x1_phase, float x1_scalar, float4 x2, Spec<S> res,
                                               const Spec<S> &wl, const Spec<S> &weight, bf_sample_record *rec) {
    Splat sp; sp.pixel = -1; sp.val_a = 0.f; sp.cnt_valid = false; sp.var_ch = -1; sp.var_val = 0.f; sp.var_rep = 1;
    const uint32_t gen = sc.integ.gen;
    const bool valid = flags & LANE_VALID;
    float posx, posy;
    int off_x = 0, off_y = 0;
    bool ran;
    sample_position(sc, x2, gen == BF_GEN_PHASE ? freq_from_wl_d(sc.C.c, wl.v[0]) : 0.0, posx, posy, ran, off_x, off_y);
    if (gen == BF_GEN_PHASE) {
        float wl0 = wl.v[0];
        float phase = valid ? x1_phase : 0.f;
        if (!ran) {
#pragma unroll
            for (int k = 0; k < S; ++k) res.v[k] = 0.f;
            phase = 0.f;
        }
        // phase.cpp:117-124; a zero value writes nothing, so the phase is only reduced where one is written
        float hs = hsum_spec<S>(res);
        float pv = (valid && ran) ? hs : 0.f;
        if (ran && sc.integ.phase_bins > 0 && pv != 0.f) {
            float ph = fmod_two_pi(phase);
            ph += ph < 0.f ? BF_TWO_PI : 0.f;
            float bin_width = BF_TWO_PI / (float) sc.integ.phase_bins;
            for (uint32_t k = 0; k < sc.integ.phase_bins; ++k) {
                float centre = (float) ((double) bin_width * ((double) k + 0.5));
                if (rect((ph - centre) / bin_width) > 0.f) { sp.var_ch = 3 + (int) k; sp.var_val = pv; }
            }
        }
        Spec<S> wres;
#pragma unroll
        for (int k = 0; k < S; ++k) wres.v[k] = fabsf(weight.v[k]) * res.v[k];
        sp.val_a = (valid && ran) ? hsum_spec<S>(wres) : 0.f;
        sp.cnt_valid = valid && ran;
        if (rec) {
            rec->valid = (valid && ran) ? 1u : 0u; rec->tf_t = posx; rec->tf_f = posy; rec->value = sp.val_a;
            rec->phase = phase; rec->wavelength0 = wl0;
        }
    } else {
        // range.cpp:141-161 / time.cpp:118-153: bins from the UNWEIGHTED spectrum, first channel
        float scalar = x1_scalar, r0 = res.v[0];
        if (gen == BF_GEN_RANGE) {
            for (uint32_t i = 0; i < sc.integ.range_bins; ++i) {
                float lo = (float) i * sc.integ.dr, hi = (float) i * sc.integ.dr + sc.integ.dr;
                if (scalar >= lo && scalar < hi && r0 != 0.f) { sp.var_ch = 5 + (int) i; sp.var_val = r0; }
            }
        } else {
            const float dt = 0.5e-9f;
            for (int i = 0; i < 50; ++i) {
                float lo = (float) i * dt, hi = (float) i * dt + dt;
                if (scalar >= lo && scalar < hi && r0 != 0.f) { sp.var_ch = 5 + 3 * i; sp.var_val = r0; sp.var_rep = 3; }
            }
        }
        Spec<S> wres;
#pragma unroll
        for (int k = 0; k < S; ++k) wres.v[k] = weight.v[k] * res.v[k];
        // mono: xyz = value broadcast; spectral: the CIE matching functions vanish at radar wavelengths
        float x = S == 1 ? wres.v[0] : 0.f;
        sp.val_a = x; sp.cnt_valid = valid;
        if (rec) {
            rec->valid = valid ? 1u : 0u; rec->tf_t = posx; rec->tf_f = posy; rec->value = r0; rec->phase = scalar; rec->wavelength0 = wl.v[0];
        }
    }
    sp.pixel = pixel_of(sc, posx, posy, off_x, off_y);
    sp.posx = posx; sp.posy = posy; sp.off_x = off_x; sp.off_y = off_y;
    return sp;
}

template <int S>
__device__ __forceinline__ Splat finalize_lane(const DevScene &sc, const Wave &w, uint32_t l) {
    uint32_t flags = w.flags[l];
    float4 x1 = w.x1[l], x2 = w.x2[l];
    Spec<S> thr, res, pend, wl;
    SpecIO<S>::load(w, l, thr, res, pend, wl);
    if ((flags & LANE_PENDING) && !w.occl[l]) {
#pragma unroll
        for (int k = 0; k < S; ++k) res.v[k] += pend.v[k];
    }
    Spec<S> weight = SpecIO<S>::load_weight(w, l, x2);
    return finalize_core<S>(sc, flags, x1.y, x1.w, x2, res, wl, weight, w.records ? rec_of(w, l) : nullptr);
}

// SMEM: the whole profile fits a per-CTA shared-memory histogram (range profiles): value channels are fp32 (CAS-loop
// shared atomics), the two count channels (alpha, weight) are accumulated as integers with native shared atomics and
// converted at the flush, which is one global atomic per non-zero entry.  Otherwise samples go to HBM directly, as one
// 128-bit reduction per sample when a pixel is four channels wide.
// SignalBlock::put of the samples of one converged warp (box filter).  When all the lanes that land in the film land in ONE
// bin (1 x 1 films: every warp) their counts and values are summed in registers first; otherwise one set of atomics per lane.
// SMEM: dst is a shared-memory copy of the profile whose count channels hold integers.
template <bool SMEM>
__device__ __forceinline__ void scatter_warp(const Splat &sp, float *dst, uint32_t C, uint32_t n_fixed) {
    const int pixel = sp.pixel;
    const uint32_t in_film = __ballot_sync(0xffffffffu, pixel >= 0);
    if (in_film == 0) return;
    const int pixel0 = __shfl_sync(0xffffffffu, pixel, __ffs(in_film) - 1);
    const bool one_pixel = __all_sync(0xffffffffu, pixel < 0 || pixel == pixel0) && (in_film & (in_film - 1));
    if (one_pixel) {
        const uint32_t nv = __popc(__ballot_sync(0xffffffffu, pixel >= 0 && sp.cnt_valid));
        if (lane_id() == (uint32_t) (__ffs(in_film) - 1)) {
            float *px = dst + (size_t) C * pixel;
            if (SMEM) {
                atomicAdd(reinterpret_cast<uint32_t *>(px + n_fixed - 1), (uint32_t) __popc(in_film));
                if (nv) atomicAdd(reinterpret_cast<uint32_t *>(px + n_fixed - 2), nv);
            } else {
                atomicAdd(px + n_fixed - 1, (float) __popc(in_film));
                if (nv) atomicAdd(px + n_fixed - 2, (float) nv);
            }
        }
        agg_add(dst, pixel >= 0 ? pixel * (int) C : -1, sp.val_a, n_fixed - 2);   // gen 3: Y; gen 1/2: X = Y = Z
        agg_add(dst, (pixel >= 0 && sp.var_ch >= 0) ? pixel * (int) C + sp.var_ch : -1, sp.var_val, sp.var_rep);
        return;
    }
    if (pixel < 0) return;
    float *px = dst + (size_t) C * pixel;
    if (SMEM) {
        atomicAdd(reinterpret_cast<uint32_t *>(px + n_fixed - 1), 1u);
        if (sp.cnt_valid) atomicAdd(reinterpret_cast<uint32_t *>(px + n_fixed - 2), 1u);
        if (sp.val_a != 0.f)
            for (uint32_t r = 0; r < n_fixed - 2; ++r) atomicAdd(px + r, sp.val_a);
        if (sp.var_ch >= 0 && sp.var_val != 0.f)
            for (uint32_t r = 0; r < sp.var_rep; ++r) atomicAdd(px + sp.var_ch + r, sp.var_val);
    } else if (C == 4 && n_fixed == 3) {
        // gen 3, one phase bin: (Y, A, W, P0) is one aligned 16-byte record
        atomicAdd(reinterpret_cast<float4 *>(px), make_float4(sp.val_a, sp.cnt_valid ? 1.f : 0.f, 1.f, sp.var_ch >= 0 ? sp.var_val : 0.f));
    } else {
        atomicAdd(px + n_fixed - 1, 1.f);
        if (sp.cnt_valid) atomicAdd(px + n_fixed - 2, 1.f);
        if (sp.val_a != 0.f)
            for (uint32_t r = 0; r < n_fixed - 2; ++r) atomicAdd(px + r, sp.val_a);
        if (sp.var_ch >= 0 && sp.var_val != 0.f)
            for (uint32_t r = 0; r < sp.var_rep; ++r) atomicAdd(px + sp.var_ch + r, sp.var_val);
    }
}
// flush of a shared-memory profile copy: one global atomic per non-zero entry
__device__ __forceinline__ void flush_hist(const float *s_hist, uint32_t total_floats, uint32_t C, uint32_t n_fixed, float *profile) {
    for (uint32_t i = threadIdx.x; i < total_floats; i += blockDim.x) {
        const uint32_t ch = i % C;
        const bool is_count = ch == n_fixed - 1 || ch == n_fixed - 2;
        const float v = is_count ? (float) __float_as_uint(s_hist[i]) : s_hist[i];
        if (v != 0.f) atomicAdd(profile + i, v);
    }
}

template <int S, bool SMEM>
__global__ void __launch_bounds__(FIN_THREADS) k_finalize(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, uint32_t n,
                                                         float *__restrict__ profile) {
    __shared__ float s_hist[SMEM ? FIN_SMEM_FLOATS : 1];
    if (w.n_lanes) n = min(n, *w.n_lanes);      // compacted wave: the lanes handed out by k_primary
    const uint32_t C = sc.channels;
    const uint32_t total_floats = C * sc.film_w * sc.film_h;
    if (SMEM) {
        for (uint32_t i = threadIdx.x; i < total_floats; i += blockDim.x) s_hist[i] = 0.f;
        __syncthreads();
    }
    const uint32_t n_fixed = sc.integ.gen == BF_GEN_PHASE ? 3u : 5u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n + 31u) & ~31u;
    for (uint32_t l = blockIdx.x * blockDim.x + threadIdx.x; l < n_round; l += stride) {
        Splat sp; sp.pixel = -1; sp.val_a = 0.f; sp.cnt_valid = false; sp.var_ch = -1; sp.var_val = 0.f; sp.var_rep = 1;
        if (l < n) sp = finalize_lane<S>(sc, w, l);
        scatter_warp<SMEM>(sp, SMEM ? s_hist : profile, C, n_fixed);
    }
    if (SMEM) {
        __syncthreads();
        flush_hist(s_hist, total_floats, C, n_fixed, profile);
    }
}

// =========================================================================================================
// mega: one sample = one thread, start to end
// =========================================================================================================
// The whole path of a sample in one thread: receiver ray, then per depth a closest-hit query (rectangles + BVH), one shade step
// (shade_core, the generic instantiation), the shadow query, until the path ends; then the binning (finalize_core).  Nothing
// is written to HBM but the profile.  Two uses, both latency-bound, where one launch beats the ~17 of the wavefront:
//   * scenes without meshes (C1 trans_rad.xml, C4 animated_trans_rad.py: three rectangles), at any size;
//   * small receive() calls on any scene (the reference's drivers use 1 000 - 65 536 samples per call).
// Same arithmetic as the wavefront sequence (records bit-identical: tests/test_gpu_parity.py::test_mega_kernel_equals_wavefront).
#define MEGA_THREADS 128
struct MegaIO {
    uint64_t seed, first_sample;
    uint32_t n;
    uint32_t *work_cursor;
    unsigned long long *totals;
    float *profile;
    uint32_t hist_floats;
};
template <int S>
__device__ __forceinline__ void mega_body(const DevScene &sc, bf_sample_record *records, const MegaIO &io, uint32_t *mega_smem) {
    float *s_rect = reinterpret_cast<float *>(mega_smem);
    uint2 *s_stack = reinterpret_cast<uint2 *>(mega_smem + RECT_SM_FLOATS * BF_MAX_RECTS);
    float *s_hist = reinterpret_cast<float *>(s_stack + TRACE_SM_STACK * MEGA_THREADS);
    uint2 l_stack[TRACE_LOCAL_STACK];
    static_assert(MEGA_THREADS == TRACE_THREADS, "trace_step indexes the shared-memory stack with TRACE_THREADS");
    const uint32_t C = sc.channels, n_fixed = sc.integ.gen == BF_GEN_PHASE ? 3u : 5u;
    const bool hist = io.hist_floats != 0u;
    if (hist) for (uint32_t i = threadIdx.x; i < io.hist_floats; i += blockDim.x) s_hist[i] = 0.f;
    stage_rects(sc, s_rect);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gen = sc.integ.gen;
    uint32_t q_closest = 0, q_shadow = 0;
    while (true) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(io.work_cursor, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= io.n) break;
        const uint32_t i = base + lane;
        const bool have = i < io.n;
        PrimaryState<S> ps;
        LaneRegs<S> R;
        bf_sample_record *rec = (records && have) ? records + i : nullptr;
        V3 ro(0.f, 0.f, 0.f), rd(0.f, 0.f, 1.f);
        float mint = 0.f, maxt = 0.f;
        bool alive = have;
        if (have) {
            raygen_sample<S>(sc, io.seed, io.first_sample + i, ps);
            ro = ps.ray.o; rd = ps.ray.d; mint = ps.ray.mint; maxt = ps.ray.maxt;
            R.flags = LANE_ACTIVE; R.rng = ps.rng;
            R.thr = Spec<S>(1.f); R.res = Spec<S>(0.f); R.pend = Spec<S>(0.f); R.wl = ps.wl;
            R.time = ps.time; R.phase = 0.f; R.bs_pdf = 0.f; R.scalar = 0.f;
            if (rec) init_record(rec, ps.wl.v[0]);
        }
        for (uint32_t depth = 1; alive; ++depth) {
            TraceLane L; L.o = ro; L.d = rd; L.mint = mint; L.maxt = maxt;
            if (trace_begin<false>(sc, s_rect, L)) { while (trace_step<false>(sc, s_stack, l_stack, L)) { } }
            ++q_closest;
            if (L.best.slot == BF_NONE) {
                // the miss branch of the integrator loop (pathtimefrequency.cpp:368-371 with t = inf): see k_trace
                if (depth > 1 && gen == BF_GEN_PHASE) { R.time = -bf_inf(); R.phase = -bf_inf(); }
                if (rec) rec->depth = depth;
                break;
            }
            R.o = ro; R.d = rd; R.hit = make_float4(L.best.t, L.best.u, L.best.v, __uint_as_float(L.best.slot));
            ShadeOut so;
            shade_core<S, -1, -1, -1>(sc, depth, R, rec, so);
            if (so.push_shadow) {
                TraceLane T; T.o = so.origin; T.d = so.sh_d; T.mint = so.mint; T.maxt = so.sh_maxt;
                if (trace_begin<true>(sc, s_rect, T)) { while (trace_step<true>(sc, s_stack, l_stack, T)) { } }
                ++q_shadow;
                if (!T.occluded) {
#pragma unroll
                    for (int k = 0; k < S; ++k) R.res.v[k] += R.pend.v[k];
                }
                R.flags &= ~LANE_PENDING;
            }
            if (!so.push_next || depth >= BF_MAX_BOUNCES) break;
            ro = so.origin; rd = so.new_d; mint = so.mint; maxt = bf_inf();
        }
        Splat sp; sp.pixel = -1; sp.val_a = 0.f; sp.cnt_valid = false; sp.var_ch = -1; sp.var_val = 0.f; sp.var_rep = 1;
        if (have) {
            float4 x2 = ps.x2;
            sp = finalize_core<S>(sc, R.flags, R.phase, R.scalar, x2, R.res, R.wl, ps.weight, rec);
        }
        __syncwarp();
        if (hist) scatter_warp<true>(sp, s_hist, C, n_fixed); else scatter_warp<false>(sp, io.profile, C, n_fixed);
    }
    {
        uint32_t qc = q_closest, qs = q_shadow;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { qc += __shfl_xor_sync(0xffffffffu, qc, o); qs += __shfl_xor_sync(0xffffffffu, qs, o); }
        if (lane == 0 && io.totals) {
            if (qc) atomicAdd(io.totals + 0, (unsigned long long) qc);
            if (qs) atomicAdd(io.totals + 1, (unsigned long long) qs);
        }
    }
    if (hist) {
        __syncthreads();
        flush_hist(s_hist, io.hist_floats, C, n_fixed, io.profile);
    }
}
template <int S>
__global__ void __launch_bounds__(MEGA_THREADS) k_mega(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, const __grid_constant__ MegaIO io) {
    extern __shared__ __align__(16) uint32_t mega_smem[];
    mega_body<S>(sc, w.records, io, mega_smem);
}

// bf_receive_batch of small items (C4: 256 frames of 1 000 samples; the phased sweep on a small scene) in ONE launch:
// blockIdx.y is the item.  An item differs from the committed scene in its rectangle table, its receiver and the copies of
// both inside DevScene (MegaItem); the CTA builds the item's DevScene in shared memory -- the parameter block with those
// members replaced -- and runs the k_mega body on it.  Reading the scene from shared memory instead of the constant bank
// costs instructions; launching, uploading and zeroing per item cost 20 us each, ten times the work of such an item.
template <int S>
__global__ void __launch_bounds__(MEGA_THREADS) k_mega_batch(const __grid_constant__ DevScene sc0, const MegaItem *__restrict__ items, uint32_t n,
                                                             uint32_t *cursors, unsigned long long *totals, uint32_t hist_floats) {
    extern __shared__ __align__(16) uint32_t mega_smem[];
    __shared__ __align__(16) uint32_t s_scene[(sizeof(DevScene) + 3) / 4];
    DevScene &sc = *reinterpret_cast<DevScene *>(s_scene);
    const MegaItem &it = items[blockIdx.y];
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&sc0);
        for (uint32_t i = threadIdx.x; i < sizeof(DevScene) / 4; i += blockDim.x) s_scene[i] = src[i];
        __syncthreads();
        auto patch = [&](void *dst, const void *from, uint32_t bytes) {
            uint32_t *d = static_cast<uint32_t *>(dst); const uint32_t *f = static_cast<const uint32_t *>(from);
            for (uint32_t i = threadIdx.x; i < bytes / 4; i += blockDim.x) d[i] = f[i];
        };
        patch(&sc.rx_ep, &it.rx_ep, sizeof(DevEndpoint));
        patch(&sc.rx_rect, &it.rx_rect, sizeof(DevRect));
        patch(&sc.tx0_rect, &it.tx0_rect, sizeof(DevRect));
        if (threadIdx.x == 0) { sc.rects = it.rects; sc.rx = it.rx; }
        __syncthreads();
    }
    MegaIO io;
    io.seed = it.seed; io.first_sample = it.first_sample; io.n = n; io.work_cursor = cursors + blockIdx.y; io.totals = totals;
    io.profile = it.profile; io.hist_floats = hist_floats;
    mega_body<S>(sc, nullptr, io, mega_smem);
}

// =========================================================================================================
// path: raygen + the paths that never come near a mesh, start to end in registers
// =========================================================================================================
// Scenes whose receiver rays mostly miss the mesh bounds (bf_scene_commit's estimate; C2: 98 %) spend their samples on three
// kinds of path: the receiver ray leaves the scene (half of them), it hits a diffuse rectangle (the ground) and both the
// shadow ray and the continuation ray are decided by the rectangles alone, or -- a few percent -- some ray enters the mesh
// bounds.  k_path integrates the first two kinds completely: receiver ray, rectangle tests, one shade step (shade_core,
// the code k_shade runs), rectangle / mesh-bounds tests of the two spawned rays, binning (finalize_core, the code k_finalize
// runs) into a per-CTA shared-memory profile.  Nothing of such a sample touches HBM.  Lanes of the third kind are SPILLED
// into the wave arrays at the stage they have reached, exactly in the form the wavefront kernels expect there:
//   receiver ray enters the mesh bounds          -> store_primary + defer queue       (k_trace<closest> -> k_shade(1) ...)
//   rectangle hit that is not plain diffuse      -> store_primary + hit + class queue  (k_shade(1) ...)
//   shadow / continuation ray still undecided    -> the state k_shade(1) would have stored + shadow / next queues
// The two halves of the work -- generating receiver rays (raygen_sample + rectangle tests) and shading the rectangle hits
// (shade_core + pre-tests + finalize_core) -- are ~16 KB and ~30 KB of executed code.  Run side by side on one SM they
// thrash its 32 KB instruction cache (measured: 40 % of the stall samples "no instruction", profiles/r02c).  So the kernel
// runs ONE CTA of 32 warps per SM and the whole CTA alternates between two long phases, separated by __syncthreads():
//   G  every warp generates rounds of 32 samples; misses are binned, spills are stored, plain rectangle hits are PARKED
//      in the CTA's stash (a word-major array in global memory, 0.5 MB per CTA: it lives in L2), until the stash is full;
//   S  every warp takes batches of 32 parked hits -- full warps, whatever the hit rate -- and carries them to the end.
// One phase lasts ~100 us, so the instruction cache is refilled a few thousand times per second instead of per round.
#ifndef PATH_THREADS
#define PATH_THREADS 1024    // 32 warps at 64 registers: measured 10.3 Gsamples/s against 9.8 (768 x 80) and 8.7 (512 x 108), profiles/r02e
#endif
#define PATH_WARPS (PATH_THREADS / 32)
#ifndef BF_PATH_STASH_CAP
#define BF_PATH_STASH_CAP 8192u
#endif
template <int S> struct PathStash {
    // words per parked hit: o(3) d(3) t slot | rng state(2) inc(2) | time f_rx sample | wl[S] weight[S]
    static constexpr int WL = 15, WEIGHT = 15 + S, WORDS = 15 + 2 * S;
};
struct PathIO {
    uint64_t seed, first_sample;
    uint32_t n;
    uint32_t *work_cursor;
    uint32_t *n_lanes;                        // lanes handed out to spilled samples
    unsigned long long *totals;               // [0] closest, [1] shadow queries (all kernels); [2], [3] the ones decided here
    float *profile;
    uint32_t *defer_queue, *defer_count;      // receiver rays that enter the mesh bounds
    uint32_t *hit_count;                      // [2]: class queues of shade(1) (w.queue_h / w.queue_h2)
    uint32_t *next_queue, *next_count;        // continuation rays after shade(1) that are not decided here
    uint32_t *shadow_count;                   // shadow rays after shade(1) that enter the mesh bounds (w.queue_s)
    uint32_t hist_floats;                     // > 0: the profile fits a shared-memory copy of that many floats
    uint32_t *stash;                          // gridDim.x x WORDS x stash_cap words
    uint32_t stash_cap;
};

// SignalBlock::put of one finished sample (box filter), as k_finalize does it per lane
template <bool SMEM>
__device__ __forceinline__ void scatter_splat(const Splat &sp, float *dst, uint32_t C, uint32_t n_fixed) {
    if (sp.pixel < 0) return;
    float *px = dst + (size_t) C * sp.pixel;
    if (SMEM) {
        atomicAdd(reinterpret_cast<uint32_t *>(px + n_fixed - 1), 1u);
        if (sp.cnt_valid) atomicAdd(reinterpret_cast<uint32_t *>(px + n_fixed - 2), 1u);
        if (sp.val_a != 0.f)
            for (uint32_t r = 0; r < n_fixed - 2; ++r) atomicAdd(px + r, sp.val_a);
        if (sp.var_ch >= 0 && sp.var_val != 0.f)
            for (uint32_t r = 0; r < sp.var_rep; ++r) atomicAdd(px + sp.var_ch + r, sp.var_val);
    } else if (C == 4 && n_fixed == 3) {
        atomicAdd(reinterpret_cast<float4 *>(px), make_float4(sp.val_a, sp.cnt_valid ? 1.f : 0.f, 1.f, sp.var_ch >= 0 ? sp.var_val : 0.f));
    } else {
        atomicAdd(px + n_fixed - 1, 1.f);
        if (sp.cnt_valid) atomicAdd(px + n_fixed - 2, 1.f);
        if (sp.val_a != 0.f)
            for (uint32_t r = 0; r < n_fixed - 2; ++r) atomicAdd(px + r, sp.val_a);
        if (sp.var_ch >= 0 && sp.var_val != 0.f)
            for (uint32_t r = 0; r < sp.var_rep; ++r) atomicAdd(px + sp.var_ch + r, sp.var_val);
    }
}

#ifndef PATH_BOUND_THREADS
#define PATH_BOUND_THREADS PATH_THREADS      // register budget = 65536 / this
#endif
template <int S, int RXK, int TXK>
__global__ void __launch_bounds__(PATH_BOUND_THREADS, 1) k_path(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, const __grid_constant__ PathIO io) {
    extern __shared__ __align__(16) uint32_t path_smem[];
    float *s_rect = reinterpret_cast<float *>(path_smem);                        // RECT_SM_FLOATS * BF_MAX_RECTS
    uint32_t *s_plain = path_smem + RECT_SM_FLOATS * BF_MAX_RECTS;               // per rectangle: 1 = diffuse class, no transmitter attached
    uint32_t *s_ctl = s_plain + BF_MAX_RECTS;                                    // [0] parked hits, [1] hits taken
    float *s_hist = reinterpret_cast<float *>(s_ctl + 4);
    constexpr uint32_t C_FIXED = 3u;                                             // gen 3: Y, A, W
    const uint32_t C = sc.channels;
    const bool hist = io.hist_floats != 0u;
    if (hist) for (uint32_t i = threadIdx.x; i < io.hist_floats; i += blockDim.x) s_hist[i] = 0.f;
    for (uint32_t r = threadIdx.x; r < sc.n_rects; r += blockDim.x) {
        const DevShape sh = sc.shapes[sc.rects[r].shape];
        s_plain[r] = (sh.mat_class == 0u && sh.tx < 0) ? 1u : 0u;
    }
    if (threadIdx.x < 4) s_ctl[threadIdx.x] = 0u;
    stage_rects(sc, s_rect);      // ends with __syncthreads()
    const uint32_t n = io.n;
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    const uint32_t cap = io.stash_cap;
    uint32_t *st = io.stash + (size_t) blockIdx.x * PathStash<S>::WORDS * cap;    // word k of entry e: st[k * cap + e]
    float *dst = hist ? s_hist : io.profile;
    volatile uint32_t *v_ctl = s_ctl;

    // work distribution: a static first chunk per warp, then chunks claimed one ahead from the global cursor
    const uint32_t n_warps = gridDim.x * PATH_WARPS;
    uint32_t chunk = 32;
    while (chunk < 512 && (uint64_t) chunk * 16u * n_warps <= n) chunk *= 2;
    const uint32_t static_total = n_warps * chunk;
    // warp ranks interleave the CTAs, so that a request smaller than the grid still spreads over all SMs
    uint32_t c_next = ((threadIdx.x >> 5) * gridDim.x + blockIdx.x) * chunk, c_end = c_next + chunk;   // warp-uniform
    uint32_t claim = 0xffffffffu;
    const bool dynamic = n > static_total;
    if (dynamic && lane == 0) claim = static_total + atomic_add_async(io.work_cursor, chunk);
    bool out_of_work = c_next >= n;

    uint32_t q_closest = 0, q_shadow = 0;   // ray queries decided in this thread
    while (true) {
        // ======================= phase G: generate until the stash is full =======================
        while (!out_of_work && v_ctl[0] + PATH_THREADS <= cap) {
            const uint32_t i = c_next + lane;
            c_next += 32u;
            if (c_next >= c_end) {          // chunk used up: open the one claimed ahead
                const uint32_t nxt = dynamic ? __shfl_sync(0xffffffffu, claim, 0) : 0xffffffffu;
                if (nxt < n) {
                    if (lane == 0) claim = static_total + atomic_add_async(io.work_cursor, chunk);
                    c_next = nxt; c_end = nxt + chunk;
                } else {
                    out_of_work = true;
                }
            }
            if (c_next >= n) out_of_work = true;
            enum { K_NONE = 0, K_MISS = 1, K_PARK = 2, K_DEFER = 3, K_SHADE1 = 4 };
            int kind = K_NONE;
            PrimaryState<S> ps;
            TraceLane L;
            if (i < n) {
                raygen_sample<S, BF_GEN_PHASE, RXK>(sc, io.seed, io.first_sample + i, ps);
                L.o = ps.ray.o; L.d = ps.ray.d; L.mint = ps.ray.mint; L.maxt = ps.ray.maxt;
                ++q_closest;
                if (trace_begin<false, true>(sc, s_rect, L)) kind = K_DEFER;
                else if (L.best.slot == BF_NONE) kind = K_MISS;
                else kind = s_plain[L.best.slot & 0x7fffffffu] ? K_PARK : K_SHADE1;
                if (w.records) init_record(w.records + i, ps.wl.v[0]);
            }
            if (kind == K_MISS) {
                // SignalBlock::put of a zero value: one count in the weight channel of the sample's bin
                float posx, posy; bool ran; int off_x, off_y;
                sample_position(sc, ps.x2, (double) sc.C.c * ps.inv_wl, posx, posy, ran, off_x, off_y);
                const int pixel = pixel_of(sc, posx, posy, off_x, off_y);
                if (pixel >= 0) {
                    if (hist) atomicAdd(reinterpret_cast<uint32_t *>(s_hist + (size_t) C * pixel + C_FIXED - 1u), 1u);
                    else atomicAdd(io.profile + (size_t) C * pixel + C_FIXED - 1u, 1.f);
                }
                if (w.records) { bf_sample_record *r = w.records + i; r->depth = 1; r->tf_t = posx; r->tf_f = posy; }
            }
            // spills before the shade step: the lane gets the state k_primary would have given it
            const uint32_t m_spill = __ballot_sync(0xffffffffu, kind >= K_DEFER);
            if (m_spill) {
                const uint32_t cls = kind == K_SHADE1 ? sc.shapes[L.best.shape].mat_class : 0u;
                const uint32_t m0 = __ballot_sync(0xffffffffu, kind == K_SHADE1 && cls == 0u), m1 = __ballot_sync(0xffffffffu, kind == K_SHADE1 && cls != 0u);
                const uint32_t md = m_spill & ~(m0 | m1);
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(io.n_lanes, (uint32_t) __popc(m_spill));
                else if (lane == 1 && m0) base = atomicAdd(io.hit_count + 0, (uint32_t) __popc(m0));
                else if (lane == 2 && m1) base = atomicAdd(io.hit_count + 1, (uint32_t) __popc(m1));
                else if (lane == 3 && md) base = atomicAdd(io.defer_count, (uint32_t) __popc(md));
                const uint32_t base_l = __shfl_sync(0xffffffffu, base, 0), base_0 = __shfl_sync(0xffffffffu, base, 1),
                               base_1 = __shfl_sync(0xffffffffu, base, 2), base_d = __shfl_sync(0xffffffffu, base, 3);
                if (kind >= K_DEFER) {
                    const uint32_t dl = base_l + __popc(m_spill & lt_mask);
                    store_primary<S>(w, dl, ps);
                    if (w.sample_of) w.sample_of[dl] = i;
                    if (kind == K_DEFER) {
                        io.defer_queue[base_d + __popc(md & lt_mask)] = dl;     // (the query stays counted here, as in k_primary)
                    } else {
                        w.hit[dl] = make_float4(L.best.t, L.best.u, L.best.v, __uint_as_float(L.best.slot));
                        if (cls == 0u) w.queue_h[base_0 + __popc(m0 & lt_mask)] = dl;
                        else w.queue_h2[base_1 + __popc(m1 & lt_mask)] = dl;
                    }
                }
            }
            // park the plain rectangle hits: one shared-memory reservation per warp
            const uint32_t m_park = __ballot_sync(0xffffffffu, kind == K_PARK);
            if (m_park) {
                uint32_t pbase = 0;
                if (lane == 0) pbase = atomicAdd(s_ctl, (uint32_t) __popc(m_park));
                pbase = __shfl_sync(0xffffffffu, pbase, 0);
                if (kind == K_PARK) {
                    uint32_t *e = st + pbase + __popc(m_park & lt_mask);
                    e[0 * cap] = __float_as_uint(L.o.x); e[1 * cap] = __float_as_uint(L.o.y); e[2 * cap] = __float_as_uint(L.o.z);
                    e[3 * cap] = __float_as_uint(L.d.x); e[4 * cap] = __float_as_uint(L.d.y); e[5 * cap] = __float_as_uint(L.d.z);
                    e[6 * cap] = __float_as_uint(L.best.t); e[7 * cap] = L.best.slot;
                    e[8 * cap] = (uint32_t) ps.rng.state; e[9 * cap] = (uint32_t) (ps.rng.state >> 32);
                    e[10 * cap] = (uint32_t) ps.rng.inc; e[11 * cap] = (uint32_t) (ps.rng.inc >> 32);
                    e[12 * cap] = __float_as_uint(ps.time); e[13 * cap] = __float_as_uint(ps.x2.y); e[14 * cap] = i;
#pragma unroll
                    for (int k = 0; k < S; ++k) {
                        e[(PathStash<S>::WL + k) * cap] = __float_as_uint(ps.wl.v[k]);
                        e[(PathStash<S>::WEIGHT + k) * cap] = __float_as_uint(ps.weight.v[k]);
                    }
                }
            }
        }
        const int all_out = __syncthreads_and(out_of_work ? 1 : 0);      // parked entries are visible to the whole CTA
        const uint32_t total = v_ctl[0];
        // ======================= phase S: shade the parked hits, 32 at a time =======================
        while (true) {
            uint32_t first = 0;
            if (lane == 0) first = atomicAdd(s_ctl + 1, 32u);
            first = __shfl_sync(0xffffffffu, first, 0);
            if (first >= total) break;
            const bool act = first + lane < total;
            LaneRegs<S> R;
            Spec<S> weight;
            float f_rx = 0.f, t_sample = 0.f;
            uint32_t smp = 0;
            if (act) {
                const uint32_t *e = st + first + lane;
                R.o = V3(__uint_as_float(e[0 * cap]), __uint_as_float(e[1 * cap]), __uint_as_float(e[2 * cap]));
                R.d = V3(__uint_as_float(e[3 * cap]), __uint_as_float(e[4 * cap]), __uint_as_float(e[5 * cap]));
                R.hit = make_float4(__uint_as_float(e[6 * cap]), 0.f, 0.f, __uint_as_float(e[7 * cap]));
                R.rng.state = (uint64_t) e[8 * cap] | ((uint64_t) e[9 * cap] << 32);
                R.rng.inc = (uint64_t) e[10 * cap] | ((uint64_t) e[11 * cap] << 32);
                t_sample = __uint_as_float(e[12 * cap]); f_rx = __uint_as_float(e[13 * cap]); smp = e[14 * cap];
#pragma unroll
                for (int k = 0; k < S; ++k) {
                    R.wl.v[k] = __uint_as_float(e[(PathStash<S>::WL + k) * cap]);
                    weight.v[k] = __uint_as_float(e[(PathStash<S>::WEIGHT + k) * cap]);
                }
            }
            bool sh_spill = false, nx_spill = false;
            ShadeOut so; so.push_next = false; so.push_shadow = false;
            float4 x2 = make_float4(t_sample, f_rx, 0.f, S == 1 ? weight.v[0] : 0.f);
            if (act) {
                R.flags = LANE_ACTIVE;
                R.thr = Spec<S>(1.f); R.res = Spec<S>(0.f); R.pend = Spec<S>(0.f);
                R.time = t_sample; R.phase = 0.f; R.bs_pdf = 0.f; R.scalar = 0.f;
                bf_sample_record *rec = w.records ? w.records + smp : nullptr;
                shade_core<S, BF_GEN_PHASE, TXK, BF_BSDF_DIFFUSE, false, false>(sc, 1u, R, rec, so);
                if (so.push_shadow) {
                    TraceLane T; T.o = so.origin; T.d = so.sh_d; T.mint = so.mint; T.maxt = so.sh_maxt;
                    if (trace_begin<true>(sc, s_rect, T)) {
                        sh_spill = true;
                    } else {
                        ++q_shadow;
                        if (!T.occluded) {
#pragma unroll
                            for (int k = 0; k < S; ++k) R.res.v[k] += R.pend.v[k];
                        }
                        R.flags &= ~LANE_PENDING;
                    }
                }
                if (so.push_next) {
                    TraceLane T; T.o = so.origin; T.d = so.new_d; T.mint = so.mint; T.maxt = bf_inf();
                    const bool need = trace_begin<false, true>(sc, s_rect, T);
                    if (need || T.best.slot != BF_NONE) {
                        nx_spill = true;         // k_trace<closest> redoes the query for this lane
                    } else {
                        // the continuation ray leaves the scene: the miss branch of k_trace (pathtimefrequency.cpp:368-371, t = inf)
                        ++q_closest;
                        R.time = -bf_inf(); R.phase = -bf_inf();
                        if (rec) rec->depth = 2u;
                    }
                }
                if (!sh_spill && !nx_spill) {
                    const Splat sp = finalize_core<S>(sc, R.flags, R.phase, R.scalar, x2, R.res, R.wl, weight, rec);
                    if (hist) scatter_splat<true>(sp, dst, C, C_FIXED); else scatter_splat<false>(sp, dst, C, C_FIXED);
                }
            }
            // spills after the shade step: what k_shade(1) would have left in the wave arrays
            const uint32_t m_spill = __ballot_sync(0xffffffffu, sh_spill || nx_spill);
            if (m_spill) {
                const uint32_t ms = __ballot_sync(0xffffffffu, sh_spill), mn = __ballot_sync(0xffffffffu, nx_spill);
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(io.n_lanes, (uint32_t) __popc(m_spill));
                else if (lane == 1 && ms) base = atomicAdd(io.shadow_count, (uint32_t) __popc(ms));
                else if (lane == 2 && mn) base = atomicAdd(io.next_count, (uint32_t) __popc(mn));
                const uint32_t base_l = __shfl_sync(0xffffffffu, base, 0), base_s = __shfl_sync(0xffffffffu, base, 1), base_n = __shfl_sync(0xffffffffu, base, 2);
                if (sh_spill || nx_spill) {
                    const uint32_t dl = base_l + __popc(m_spill & lt_mask);
                    w.ray_o[dl] = make_float4(so.origin.x, so.origin.y, so.origin.z, so.mint);
                    if (nx_spill) w.ray_d[dl] = make_float4(so.new_d.x, so.new_d.y, so.new_d.z, bf_inf());
                    if (sh_spill) w.sh_d[dl] = make_float4(so.sh_d.x, so.sh_d.y, so.sh_d.z, so.sh_maxt);
                    w.flags[dl] = R.flags;
                    w.x1[dl] = make_float4(R.time, R.phase, R.bs_pdf, R.scalar);
                    float4 x2s = x2;
                    SpecIO<S>::store_weight(w, dl, weight, x2s);
                    w.x2[dl] = x2s;
                    SpecIO<S>::store(w, dl, R.thr, R.res, R.pend, R.wl);
                    w.rng[dl] = make_ulonglong2(R.rng.state, R.rng.inc);
                    if (w.sample_of) w.sample_of[dl] = smp;
                    if (sh_spill) w.queue_s[base_s + __popc(ms & lt_mask)] = dl;
                    if (nx_spill) io.next_queue[base_n + __popc(mn & lt_mask)] = dl;
                }
            }
        }
        if (all_out) break;
        __syncthreads();                     // every warp has left phase S
        if (threadIdx.x == 0) { s_ctl[0] = 0u; s_ctl[1] = 0u; }
        __syncthreads();
    }
    // ray-query totals of this warp
    {
        uint32_t qc = q_closest, qs = q_shadow;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { qc += __shfl_xor_sync(0xffffffffu, qc, o); qs += __shfl_xor_sync(0xffffffffu, qs, o); }
        if (lane == 0 && io.totals) {
            if (qc) { atomicAdd(io.totals + 0, (unsigned long long) qc); atomicAdd(io.totals + 2, (unsigned long long) qc); }
            if (qs) { atomicAdd(io.totals + 1, (unsigned long long) qs); atomicAdd(io.totals + 3, (unsigned long long) qs); }
        }
    }
    if (hist) {
        __syncthreads();
        flush_hist(s_hist, io.hist_floats, C, C_FIXED, io.profile);
    }
}

// SignalBlock::put, filters wider than one bin (signalblock.cpp:113-161): the sample is spread over an n x n footprint with
// separable weights looked up in the discretised filter.  The reference splats into a block with a border of `border` bins
// and crops it when the block is put into the ADC storage (signalblock.cpp:49-77): here the crop is the bounds test on the
// destination bin.  Float-to-int conversions saturate (the reference's are undefined for non-finite positions).
__device__ __forceinline__ float filter_weight(const DevFilter &f, float x) {
    const int index = min((int) fabsf(x * f.scale), 31);          // rfilter.h:62-65
    return f.values[max(index, 0)];
}
template <int S>
__global__ void __launch_bounds__(FIN_THREADS) k_finalize_wide(const __grid_constant__ DevScene sc, const __grid_constant__ Wave w, uint32_t n,
                                                              float *__restrict__ profile, const __grid_constant__ DevFilter f) {
    if (w.n_lanes) n = min(n, *w.n_lanes);
    const uint32_t C = sc.channels;
    const uint32_t n_fixed = sc.integ.gen == BF_GEN_PHASE ? 3u : 5u;
    const int border = (int) f.border, size_x = (int) sc.film_w + 2 * border, size_y = (int) sc.film_h + 2 * border;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t l = blockIdx.x * blockDim.x + threadIdx.x; l < n; l += stride) {
        const Splat sp = finalize_lane<S>(sc, w, l);
        const float px = sp.posx - ((float) (sp.off_x - border) + .5f), py = sp.posy - ((float) (sp.off_y - border) + .5f);
        if (!(fabsf(px) < 1e9f && fabsf(py) < 1e9f)) continue;
        const int lox = max((int) ceilf(px - f.radius), 0), loy = max((int) ceilf(py - f.radius), 0);
        const uint32_t hix = (uint32_t) min((int) floorf(px + f.radius), size_x - 1), hiy = (uint32_t) min((int) floorf(py + f.radius), size_y - 1);
        const float basex = (float) lox - px, basey = (float) loy - py;
        for (uint32_t yr = 0; yr < f.n; ++yr) {
            const uint32_t y = (uint32_t) loy + yr;
            bool enabled = y <= hiy;
            const float wy = filter_weight(f, basey + (float) yr);
            for (uint32_t xr = 0; xr < f.n; ++xr) {
                const uint32_t x = (uint32_t) lox + xr;
                const float weight = wy * filter_weight(f, basex + (float) xr);
                enabled &= x <= hix;
                const int tx = (int) x - border, ty = (int) y - border;
                if (!enabled || weight == 0.f || tx < 0 || ty < 0 || tx >= (int) sc.film_w || ty >= (int) sc.film_h) continue;
                float *dst = profile + (size_t) C * ((size_t) ty * sc.film_w + (size_t) tx);
                atomicAdd(dst + n_fixed - 1, weight);                                   // W: value 1
                if (sp.cnt_valid) atomicAdd(dst + n_fixed - 2, weight);                 // A: value 1
                if (sp.val_a != 0.f)
                    for (uint32_t r = 0; r < n_fixed - 2; ++r) atomicAdd(dst + r, sp.val_a * weight);
                if (sp.var_ch >= 0 && sp.var_val != 0.f)
                    for (uint32_t r = 0; r < sp.var_rep; ++r) atomicAdd(dst + sp.var_ch + r, sp.var_val * weight);
            }
        }
    }
}

// =========================================================================================================
// launchers
// =========================================================================================================
#ifndef BF_TRACE_GRID_BLOCKS
#define BF_TRACE_GRID_BLOCKS TRACE_MIN_BLOCKS
#endif
static int trace_grid(const LaunchCfg &cfg) { return cfg.sm_count * BF_TRACE_GRID_BLOCKS; }

int launch_raygen(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, const LaunchCfg &cfg) {
    if (n == 0) return 0;
    dim3 grid((n + 255) / 256), block(256);
    if (sc.S == 1) k_raygen<1><<<grid, block, 0, cfg.stream>>>(sc, w, seed, first_sample, n);
    else k_raygen<4><<<grid, block, 0, cfg.stream>>>(sc, w, seed, first_sample, n);
    return 1;
}
static TraceIO trace_io(const Wave &w, const float4 *dirs, const uint32_t *queue, const uint32_t *count, uint32_t n_direct,
                        uint32_t *work_cursor, const LaunchCfg &cfg) {
    TraceIO io;
    memset(&io, 0, sizeof(io));
    io.ray_o = w.ray_o; io.ray_d = dirs; io.queue = queue; io.count = count; io.n_direct = n_direct; io.work_cursor = work_cursor;
    io.rounds = cfg.trace_rounds; io.sm_count = (uint32_t) cfg.sm_count;
    return io;
}
int launch_trace_closest(const DevScene &sc, const Wave &w, const uint32_t *queue, const uint32_t *count, uint32_t n_direct,
                         uint32_t depth, uint32_t *hit_count, uint32_t *work_cursor, const LaunchCfg &cfg, bool count_queries,
                         uint32_t *defer_queue, uint32_t *defer_count) {
    TraceIO io = trace_io(w, w.ray_d, queue, count, n_direct, work_cursor, cfg);
    io.hit = w.hit; io.hit_queue[0] = w.queue_h; io.hit_queue[1] = w.queue_h2; io.hit_count[0] = hit_count; io.hit_count[1] = hit_count + 1; io.x1 = w.x1; io.records = w.records; io.sample_of = w.sample_of; io.depth = depth;
    io.total = count_queries ? w.totals + 0 : nullptr;     // the deferred half of a query is counted by the launch that deferred it
    io.hit_queue[2] = defer_queue; io.hit_count[2] = defer_count;
    if (defer_queue) k_trace<false, true><<<trace_grid(cfg), TRACE_THREADS, 0, cfg.stream>>>(sc, io);
    else k_trace<false><<<trace_grid(cfg), TRACE_THREADS, 0, cfg.stream>>>(sc, io);
    return 1;
}
int launch_trace_shadow(const DevScene &sc, const Wave &w, const uint32_t *queue, const uint32_t *count, uint32_t *work_cursor,
                        const LaunchCfg &cfg) {
    TraceIO io = trace_io(w, w.sh_d, queue, count, 0, work_cursor, cfg);
    io.occl = w.occl; io.total = w.totals + 1;
    k_trace<true><<<trace_grid(cfg), TRACE_THREADS, 0, cfg.stream>>>(sc, io);
    return 1;
}
template <int S>
static int launch_shade_s(const DevScene &sc, const Wave &w, uint32_t depth, const uint32_t *hit_counts, int grid, uint32_t *next_queue,
                          uint32_t *ctr_block, const LaunchCfg &cfg) {
    // Specialised kernels for the radar configurations (gen-3 integrator, wigner or area transmitters): one launch per material class.
#define BF_SHADE(T_, B_, Q_) k_shade<S, BF_GEN_PHASE, T_, B_, Q_><<<grid, 128, 0, cfg.stream>>>(sc, w, depth, hit_counts, next_queue, ctr_block)
#define BF_SHADE_T(T_) do { \
        BF_SHADE(T_, BF_BSDF_DIFFUSE, 0); \
        if (sc.hint_cls1_bsdf == BF_BSDF_ROUGHCONDUCTOR) BF_SHADE(T_, BF_BSDF_ROUGHCONDUCTOR, 1); else BF_SHADE(T_, -1, 1); } while (0)
    if (sc.integ.gen == BF_GEN_PHASE && sc.n_txs == 1 && !sc.integ.doppler && sc.hint_tx_kind == BF_TX_WIGNER) { BF_SHADE_T(BF_TX_WIGNER); return 2; }
    if (sc.integ.gen == BF_GEN_PHASE && sc.n_txs == 1 && !sc.integ.doppler && sc.hint_tx_kind == BF_TX_AREA) { BF_SHADE_T(BF_TX_AREA); return 2; }
#undef BF_SHADE_T
#undef BF_SHADE
    k_shade<S, -1, -1, -1, 2><<<grid, 128, 0, cfg.stream>>>(sc, w, depth, hit_counts, next_queue, ctr_block);
    return 1;
}
int launch_shade(const DevScene &sc, const Wave &w, uint32_t depth, const uint32_t *hit_counts, uint32_t n_bound,
                 uint32_t *next_queue, uint32_t *ctr_block, const LaunchCfg &cfg) {
    int grid = (int) min((uint32_t) (cfg.sm_count * 8), (n_bound + 64u + 127u) / 128u);   // n_bound = upper bound of the queue lengths
    if (grid < 1) grid = 1;
    return sc.S == 1 ? launch_shade_s<1>(sc, w, depth, hit_counts, grid, next_queue, ctr_block, cfg)
                     : launch_shade_s<4>(sc, w, depth, hit_counts, grid, next_queue, ctr_block, cfg);
}
int launch_finalize(const DevScene &sc, const Wave &w, uint32_t n, float *profile, const DevFilter *filter, const LaunchCfg &cfg) {
    if (n == 0) return 0;
    int grid = (int) min((uint32_t) (cfg.sm_count * 4), (n + FIN_THREADS - 1) / FIN_THREADS);
    if (filter) {
        if (sc.S == 1) k_finalize_wide<1><<<grid, FIN_THREADS, 0, cfg.stream>>>(sc, w, n, profile, *filter);
        else k_finalize_wide<4><<<grid, FIN_THREADS, 0, cfg.stream>>>(sc, w, n, profile, *filter);
        return 1;
    }
    const bool smem = sc.channels * sc.film_w * sc.film_h <= FIN_SMEM_FLOATS;
    if (sc.S == 1) { if (smem) k_finalize<1, true><<<grid, FIN_THREADS, 0, cfg.stream>>>(sc, w, n, profile); else k_finalize<1, false><<<grid, FIN_THREADS, 0, cfg.stream>>>(sc, w, n, profile); }
    else { if (smem) k_finalize<4, true><<<grid, FIN_THREADS, 0, cfg.stream>>>(sc, w, n, profile); else k_finalize<4, false><<<grid, FIN_THREADS, 0, cfg.stream>>>(sc, w, n, profile); }
    return 1;
}
int launch_primary(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, uint32_t *hit_count,
                   uint32_t *work_cursor, float *profile, uint32_t *defer_queue, uint32_t *defer_count, const LaunchCfg &cfg) {
    if (n == 0) return 0;
    PrimaryIO io;
    io.seed = seed; io.first_sample = first_sample; io.n = n; io.work_cursor = work_cursor; io.hit_count = hit_count; io.n_lanes = w.n_lanes;
    io.total = w.totals + 0; io.profile = profile; io.defer_queue = defer_queue; io.defer_count = defer_count;
    const bool defer = defer_queue != nullptr;
    const int grid = cfg.sm_count * BF_PRIMARY_MIN_BLOCKS;
    // specialised instantiations: gen-3 integrator with a wigner or an omnidirectional receiver; generic otherwise
#define BF_PRIMARY(S_, G_, R_) do { if (defer) k_primary<S_, G_, R_, true><<<grid, TRACE_THREADS, 0, cfg.stream>>>(sc, w, io); \
                                     else k_primary<S_, G_, R_, false><<<grid, TRACE_THREADS, 0, cfg.stream>>>(sc, w, io); } while (0)
#define BF_PRIMARY_S(S_) do { \
        if (sc.integ.gen == BF_GEN_PHASE && sc.hint_rx_kind == BF_RX_WIGNER) BF_PRIMARY(S_, BF_GEN_PHASE, BF_RX_WIGNER); \
        else if (sc.integ.gen == BF_GEN_PHASE && sc.hint_rx_kind == BF_RX_OMNIDIRECTIONAL) BF_PRIMARY(S_, BF_GEN_PHASE, BF_RX_OMNIDIRECTIONAL); \
        else BF_PRIMARY(S_, -1, -1); } while (0)
    if (sc.S == 1) BF_PRIMARY_S(1); else BF_PRIMARY_S(4);
#undef BF_PRIMARY_S
#undef BF_PRIMARY
    return 1;
}
// Dynamic shared memory of k_path: rectangles, per-rectangle flags, phase counters, the profile copy
static size_t path_smem_bytes(uint32_t hist_floats) {
    return ((size_t) RECT_SM_FLOATS * BF_MAX_RECTS + BF_MAX_RECTS + 4 + hist_floats) * 4;
}
bool path_kernel_available(const DevScene &sc) {
    return sc.integ.gen == BF_GEN_PHASE && !sc.integ.doppler && sc.n_txs == 1 && (sc.hint_tx_kind == BF_TX_WIGNER || sc.hint_tx_kind == BF_TX_AREA) &&
           (sc.hint_rx_kind == BF_RX_WIGNER || sc.hint_rx_kind == BF_RX_OMNIDIRECTIONAL) && sc.n_nodes != 0;
}
// Entries of one CTA's parking area.  A CTA alternates a generate and a shade phase each time the area fills up, and every switch
// costs an instruction-cache refill, two CTA-wide barriers and the ragged ends of both phases: measured on C2 (2^25-sample waves,
// 226 000 samples per CTA of which ~45 % are parked) 8 192 entries 1.055e10 samples/s, 16 384 1.069e10, 32 768 1.085e10, 65 536
// 1.092e10 (profiles/r02z).  So the area holds what one CTA can park in a wave -- no more than the samples it generates -- between
// BF_PATH_STASH_CAP and `cap_max` (bf_set_tunable "path_stash_max") entries; a full area only ends the phase early, it never overflows.
uint32_t path_stash_cap(uint32_t lanes, int sm_count, uint32_t cap_max) {
    const uint64_t per_cta = ((uint64_t) lanes + (uint64_t) sm_count - 1u) / (uint64_t) std::max(sm_count, 1);
    const uint64_t cap = (per_cta + PATH_THREADS + 1023u) / 1024u * 1024u;
    return (uint32_t) std::min<uint64_t>(std::max<uint64_t>(cap, BF_PATH_STASH_CAP), std::max<uint64_t>(cap_max, BF_PATH_STASH_CAP));
}
size_t path_stash_bytes(const DevScene &sc, int sm_count, uint32_t cap) {
    return (size_t) sm_count * (sc.S == 1 ? PathStash<1>::WORDS : PathStash<4>::WORDS) * cap * 4;
}
int launch_path(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, uint32_t *ctr, float *profile,
                uint32_t *defer_queue, uint32_t *stash, uint32_t stash_cap, const LaunchCfg &cfg) {
    if (n == 0) return 0;
    PathIO io;
    io.seed = seed; io.first_sample = first_sample; io.n = n;
    io.work_cursor = ctr + 2; io.n_lanes = w.n_lanes; io.totals = w.totals; io.profile = profile;
    io.defer_queue = defer_queue; io.defer_count = ctr + 0; io.hit_count = ctr + 4;
    uint32_t *blk1 = ctr + BF_CTR_STRIDE;
    io.next_queue = w.queue[0]; io.next_count = blk1 + 0; io.shadow_count = blk1 + 1;
    const uint32_t floats = sc.channels * sc.film_w * sc.film_h;
    io.hist_floats = floats <= FIN_SMEM_FLOATS ? floats : 0u;
    io.stash = stash; io.stash_cap = stash_cap;
    const size_t smem = path_smem_bytes(io.hist_floats);
    const int grid = cfg.sm_count;       // one CTA per SM: the CTA-wide phases are what keeps the instruction cache warm
#define BF_PATH(S_, R_, T_) k_path<S_, R_, T_><<<grid, PATH_THREADS, smem, cfg.stream>>>(sc, w, io)
#define BF_PATH_RT(S_) do { \
        if (sc.hint_rx_kind == BF_RX_WIGNER && sc.hint_tx_kind == BF_TX_WIGNER) BF_PATH(S_, BF_RX_WIGNER, BF_TX_WIGNER); \
        else if (sc.hint_rx_kind == BF_RX_WIGNER) BF_PATH(S_, BF_RX_WIGNER, BF_TX_AREA); \
        else if (sc.hint_tx_kind == BF_TX_WIGNER) BF_PATH(S_, BF_RX_OMNIDIRECTIONAL, BF_TX_WIGNER); \
        else BF_PATH(S_, BF_RX_OMNIDIRECTIONAL, BF_TX_AREA); } while (0)
    if (sc.S == 1) BF_PATH_RT(1); else BF_PATH_RT(4);
#undef BF_PATH_RT
#undef BF_PATH
    return 1;
}

int launch_mega(const DevScene &sc, const Wave &w, uint64_t seed, uint64_t first_sample, uint32_t n, uint32_t *work_cursor, float *profile,
                const LaunchCfg &cfg) {
    if (n == 0) return 0;
    MegaIO io;
    io.seed = seed; io.first_sample = first_sample; io.n = n; io.work_cursor = work_cursor; io.totals = w.totals; io.profile = profile;
    const uint32_t floats = sc.channels * sc.film_w * sc.film_h;
    io.hist_floats = floats <= 4096u ? floats : 0u;
    const size_t smem = ((size_t) RECT_SM_FLOATS * BF_MAX_RECTS + 2 * TRACE_SM_STACK * MEGA_THREADS + io.hist_floats) * 4;
    // one warp per 32 samples up to a full machine; small requests get a small grid (and a small flush)
    const int grid = (int) std::min<uint32_t>((uint32_t) cfg.sm_count * 4u, (n + MEGA_THREADS - 1) / MEGA_THREADS);
    if (sc.S == 1) k_mega<1><<<grid, MEGA_THREADS, smem, cfg.stream>>>(sc, w, io);
    else k_mega<4><<<grid, MEGA_THREADS, smem, cfg.stream>>>(sc, w, io);
    return 1;
}
#ifndef MEGA_BATCH_CTAS_PER_SM
#define MEGA_BATCH_CTAS_PER_SM 8u
#endif
int launch_mega_batch(const DevScene &sc, const MegaItem *items, uint32_t n_items, uint32_t n, uint32_t *cursors, unsigned long long *totals,
                      const LaunchCfg &cfg) {
    if (n == 0 || n_items == 0) return 0;
    const uint32_t floats = sc.channels * sc.film_w * sc.film_h;
    const uint32_t hist_floats = floats <= 4096u ? floats : 0u;
    const size_t smem = ((size_t) RECT_SM_FLOATS * BF_MAX_RECTS + 2 * TRACE_SM_STACK * MEGA_THREADS + hist_floats) * 4;
    // per item: one warp per 32 samples, capped so that the whole batch is a few CTAs per SM
    const uint32_t per_item_cap = std::max<uint32_t>(((uint32_t) cfg.sm_count * MEGA_BATCH_CTAS_PER_SM + n_items - 1) / n_items, 1u);
    const dim3 grid(std::min<uint32_t>(per_item_cap, (n + MEGA_THREADS - 1) / MEGA_THREADS), n_items);
    if (sc.S == 1) k_mega_batch<1><<<grid, MEGA_THREADS, smem, cfg.stream>>>(sc, items, n, cursors, totals, hist_floats);
    else k_mega_batch<4><<<grid, MEGA_THREADS, smem, cfg.stream>>>(sc, items, n, cursors, totals, hist_floats);
    return 1;
}
int launch_trace_rays(const DevScene &sc, const float4 *ray_o, const float4 *ray_d, uint32_t n, float4 *hit, uint32_t *occl,
                      uint32_t *work_cursor, const LaunchCfg &cfg) {
    if (n == 0) return 0;
    TraceIO io;
    memset(&io, 0, sizeof(io));
    io.ray_o = ray_o; io.ray_d = ray_d; io.n_direct = n; io.work_cursor = work_cursor; io.hit = hit; io.occl = hit ? nullptr : occl;
    io.rounds = cfg.trace_rounds; io.sm_count = (uint32_t) cfg.sm_count;
    io.steps = hit ? occl : nullptr;   // closest queries: `occl`, when given, receives the per-ray step count
    if (hit) k_trace<false><<<trace_grid(cfg), TRACE_THREADS, 0, cfg.stream>>>(sc, io);
    else k_trace<true><<<trace_grid(cfg), TRACE_THREADS, 0, cfg.stream>>>(sc, io);
    return 1;
}
int launch_resolve_hits(const DevScene &sc, const float4 *hit, uint32_t n, uint32_t *shape, uint32_t *prim, const LaunchCfg &cfg) {
    if (n == 0) return 0;
    k_resolve_hits<<<(n + 255) / 256, 256, 0, cfg.stream>>>(sc, hit, n, shape, prim);
    return 1;
}

// =========================================================================================================
// refit: moving meshes without a rebuild (bf_scene_update_transform)
// =========================================================================================================
// A mesh that moves rigidly between frames (the reference rebuilds its scene per frame,
// animated_trans_rad.py:289-311) keeps the TOPOLOGY of the 8-wide BVH and gets new boxes:
//   k_refit_tris    the rest-pose triangle records of the moved shape are transformed with the same fp32 fmadd chain as
//                   Transform::transform_affine (transform.h:91-99), so the device triangles equal those of a scene built on
//                   the moved vertices bit for bit (and with them t, u, v of every hit);
//   k_refit_level   one thread per node, one launch per tree level, deepest first: child boxes from the triangles (leaf children)
//                   or from the boxes the level below has just written (inner children), then the node's own quantisation
//                   frame and its 8-bit child corners, rounded outwards exactly as bf_bvh.cpp does.
__global__ void __launch_bounds__(256) k_refit_tris(const float4 *__restrict__ rest, const float4 *__restrict__ rest_nrm, float4 *__restrict__ tris,
                                                    float4 *__restrict__ nrm, uint32_t n_tris, uint32_t shape, M4 M, M4 Minv) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_tris) return;
    const float4 a = rest[3 * (size_t) slot];
    if (__float_as_uint(a.w) != shape) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 v = k == 0 ? a : rest[3 * (size_t) slot + k];
        const V3 p = xf_point_affine(M, V3(v.x, v.y, v.z));
        tris[3 * (size_t) slot + k] = make_float4(p.x, p.y, p.z, v.w);
        if (rest_nrm) {
            const float4 nv = rest_nrm[3 * (size_t) slot + k];
            V3 n = xf_normal(Minv, V3(nv.x, nv.y, nv.z));
            if (n.x != 0.f || n.y != 0.f || n.z != 0.f) n = normalize(n);
            nrm[3 * (size_t) slot + k] = make_float4(n.x, n.y, n.z, 0.f);
        }
    }
}

__global__ void __launch_bounds__(128) k_refit_level(uint4 *__restrict__ nodes, const float4 *__restrict__ tris, float *__restrict__ node_box,
                                                     uint32_t first, uint32_t count, float pad_abs) {
    const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= first + count) return;
    uint4 *np = nodes + 5 * (size_t) i;
    const uint4 n0 = np[0], n1 = np[1];
    const uint32_t imask = n0.w >> 24, child_base = n1.x, tri_base = n1.y;
    float clo[8][3], chi[8][3];
    bool used[8];
    float lo[3] = {bf_inf(), bf_inf(), bf_inf()}, hi[3] = {-bf_inf(), -bf_inf(), -bf_inf()};
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const uint32_t m = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
        used[s] = m != 0u;
        if (!used[s]) continue;
        const uint32_t low5 = m & 31u;
        if (low5 >= 24u) {         // inner child: the box the level below has written
            const uint32_t c = child_base + __popc(imask & ((1u << s) - 1u));
#pragma unroll
            for (int a = 0; a < 3; ++a) { clo[s][a] = node_box[6 * (size_t) c + a]; chi[s][a] = node_box[6 * (size_t) c + 3 + a]; }
        } else {                   // leaf child: 1 - 3 triangles, boxes padded as in build_wide_bvh
            const uint32_t cnt = __popc(m >> 5);
#pragma unroll
            for (int a = 0; a < 3; ++a) { clo[s][a] = bf_inf(); chi[s][a] = -bf_inf(); }
            for (uint32_t t = 0; t < cnt; ++t) {
                const size_t slot = (size_t) tri_base + low5 + t;
                const float4 p0 = tris[3 * slot], p1 = tris[3 * slot + 1], p2 = tris[3 * slot + 2];
                const float v[3][3] = {{p0.x, p0.y, p0.z}, {p1.x, p1.y, p1.z}, {p2.x, p2.y, p2.z}};
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const float l = fminf(v[0][a], fminf(v[1][a], v[2][a])), h = fmaxf(v[0][a], fmaxf(v[1][a], v[2][a]));
                    const float pad = pad_abs + 4e-7f * fmaxf(fabsf(l), fabsf(h));
                    clo[s][a] = fminf(clo[s][a], l - pad); chi[s][a] = fmaxf(chi[s][a], h + pad);
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], clo[s][a]); hi[a] = fmaxf(hi[a], chi[s][a]); }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) { node_box[6 * (size_t) i + a] = lo[a]; node_box[6 * (size_t) i + 3 + a] = hi[a]; }
    // quantisation frame: origin = lo, per-axis power-of-two step with 255 steps covering the extent (bf_bvh.cpp)
    uint32_t ebyte[3]; float scale[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const double ext = (double) hi[a] - (double) lo[a];
        int e = ext > 0.0 ? (int) ceil(log2(ext / 255.0)) : -100;
        e = max(-120, min(120, e));
        while (ldexp(255.0, e) < ext * (1.0 + 1e-6)) ++e;
        ebyte[a] = (uint32_t) (e + 127);
        scale[a] = ldexpf(1.0f, e);
    }
    uint32_t q[6][2] = {{0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}};      // qlo.x, qlo.y, qlo.z, qhi.x, qhi.y, qhi.z
#pragma unroll
    for (int s = 0; s < 8; ++s) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            int ql = 255, qh = 0;          // empty slot: inverted box
            if (used[s]) {
                const double lo_q = floor(((double) clo[s][a] - (double) lo[a]) / (double) scale[a]);
                const double hi_q = ceil(((double) chi[s][a] - (double) lo[a]) / (double) scale[a]);
                ql = (int) fmax(0.0, fmin(255.0, lo_q)); qh = (int) fmax(0.0, fmin(255.0, hi_q));
                while (ql > 0 && lo[a] + (float) ql * scale[a] > clo[s][a]) --ql;
                while (qh < 255 && lo[a] + (float) qh * scale[a] < chi[s][a]) ++qh;
            }
            q[a][s >> 2] |= (uint32_t) ql << (8 * (s & 3));
            q[3 + a][s >> 2] |= (uint32_t) qh << (8 * (s & 3));
        }
    }
    np[0] = make_uint4(__float_as_uint(lo[0]), __float_as_uint(lo[1]), __float_as_uint(lo[2]), ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24));
    np[2] = make_uint4(q[0][0], q[0][1], q[1][0], q[1][1]);
    np[3] = make_uint4(q[2][0], q[2][1], q[3][0], q[3][1]);
    np[4] = make_uint4(q[4][0], q[4][1], q[5][0], q[5][1]);
}

int launch_refit_tris(const float4 *rest, const float4 *rest_nrm, float4 *tris, float4 *nrm, uint32_t n_tris, uint32_t shape, const M4 &M, const M4 &Minv,
                      const LaunchCfg &cfg) {
    if (n_tris == 0) return 0;
    k_refit_tris<<<(n_tris + 255) / 256, 256, 0, cfg.stream>>>(rest, rest_nrm, tris, nrm, n_tris, shape, M, Minv);
    return 1;
}
int launch_refit_nodes(uint4 *nodes, const float4 *tris, float *node_box, const uint32_t *level_first, uint32_t n_levels, float pad_abs, const LaunchCfg &cfg) {
    int launches = 0;
    for (uint32_t l = n_levels; l-- > 0;) {
        const uint32_t first = level_first[l], count = level_first[l + 1] - first;
        if (count == 0) continue;
        k_refit_level<<<(count + 127) / 128, 128, 0, cfg.stream>>>(nodes, tris, node_box, first, count, pad_abs);
        ++launches;
    }
    return launches;
}

// =========================================================================================================
// develop: HDRADC::bitmap(raw = false) (hdradc.cpp:219-275) -- the step after the path
// =========================================================================================================
// The reference flags channel 2 of the storage as the Struct weight field and converts to a target without it:
// every remaining channel becomes value * (1.f / weight) (struct.cpp:1761-1766, 1810-1811), stored in the ADC's
// component format.  An empty bin (weight 0) therefore develops to 0 * inf = NaN, as it does in the reference.
// One thread per output element; consecutive threads read consecutive floats of the raw profile (one skipped per
// bin), the weight comes from the same lines.  HBM-bound: 4 C bytes read and 4 (C - 1) or 2 (C - 1) written per bin.
template <typename OUT>
__global__ void __launch_bounds__(256) k_develop(const float *__restrict__ raw, OUT *__restrict__ out, uint32_t n_bins, uint32_t C) {
    const uint32_t Co = C - 1u;
    const uint64_t total = (uint64_t) n_bins * Co, stride = (uint64_t) gridDim.x * blockDim.x;
    for (uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; j < total; j += stride) {
        const uint32_t bin = (uint32_t) (j / Co), c = (uint32_t) (j - (uint64_t) bin * Co);
        const float *px = raw + (size_t) bin * C;
        const float inv_weight = 1.f / px[2];
        const float v = px[c < 2u ? c : c + 1u] * inv_weight;
        if constexpr (sizeof(OUT) == 2) out[j] = __float2half_rn(v);
        else out[j] = v;
    }
}

// Receive.ipynb cells 26-27 and 31: what the reference's scripts do with bitmap(raw=True) on the host, one value per bin --
//   mode 0:  10 * log10(v * (v > 0) / sample_count + eps)       (cells 26-27; numpy promotes to float64, eps = 2^-52)
//   mode 1:  10 * log10(|v| / sample_count + eps)               (cell 31, with sample_count = 1 there)
// evaluated in double like numpy does, one thread per bin, reading channel `ch` of the raw profile.
__global__ void __launch_bounds__(256) k_power_db(const float *__restrict__ raw, float *__restrict__ out, uint32_t n_bins, uint32_t C, uint32_t ch,
                                                  double inv_count, int mode) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_bins; i += stride) {
        const double v = (double) raw[(size_t) i * C + ch];
        const double lin = (mode == 0 ? (v > 0.0 ? v : 0.0) : fabs(v)) * inv_count + 2.220446049250313e-16;
        out[i] = (float) (10.0 * log10(lin));
    }
}
int launch_power_db(const float *raw, float *out, uint32_t n_bins, uint32_t channels, uint32_t ch, double sample_count, int mode, const LaunchCfg &cfg) {
    if (n_bins == 0) return 0;
    const int grid = (int) std::min<uint32_t>((n_bins + 255) / 256, (uint32_t) cfg.sm_count * 8);
    k_power_db<<<grid, 256, 0, cfg.stream>>>(raw, out, n_bins, channels, ch, 1.0 / sample_count, mode);
    return 1;
}

int launch_develop(const float *raw, void *out, uint32_t n_bins, uint32_t channels, bool half_out, const LaunchCfg &cfg) {
    if (n_bins == 0 || channels < 2) return 0;
    const uint64_t total = (uint64_t) n_bins * (channels - 1u);
    const int grid = (int) std::min<uint64_t>((total + 255) / 256, (uint64_t) cfg.sm_count * 8);
    if (half_out) k_develop<__half><<<grid, 256, 0, cfg.stream>>>(raw, (__half *) out, n_bins, channels);
    else k_develop<float><<<grid, 256, 0, cfg.stream>>>(raw, (float *) out, n_bins, channels);
    return 1;
}

// =========================================================================================================
// bf_trace_first_hits_static: closest-hit queries for callers that may not allocate device memory
// =========================================================================================================
// The stand-alone ray query (launch_trace_rays + launch_resolve_hits above) with every device-side temporary -- the staged
// caller arrays, the packed rays, k_trace's work cursor and hit records, the resolved ids -- in ONE file-scope array instead
// of cudaMalloc'ed buffers.  Every GPU that loads the library holds the array (6.25 MiB + 256 B) until the process exits; every
// call on a device shares it and finds the previous call's contents in it, so the one word that must start at zero (the work
// cursor) is cleared per call, and two calls that overlap on a GPU (two scenes on two streams) race.
// On the host `g_static_trace` is a stand-in symbol, not a device address: host code gets the address from
// cudaGetSymbolAddress, once per device with that device current (static_trace_base); device code names the array directly.
#define ST_N BF_STATIC_TRACE_MAX_RAYS
enum : uint32_t {              // word offsets into g_static_trace; each block is sized for ST_N rays
    ST_CURSOR = 0, ST_HDR = 64,
    ST_IN_O = ST_HDR, ST_IN_D = ST_IN_O + 3u * ST_N, ST_IN_MINT = ST_IN_D + 3u * ST_N, ST_IN_MAXT = ST_IN_MINT + ST_N,
    ST_RO = ST_IN_MAXT + ST_N, ST_RD = ST_RO + 4u * ST_N, ST_HIT = ST_RD + 4u * ST_N,
    ST_SHAPE = ST_HIT + 4u * ST_N, ST_PRIM = ST_SHAPE + ST_N, ST_T = ST_PRIM + ST_N, ST_U = ST_T + ST_N, ST_V = ST_U + ST_N,
    ST_WORDS = ST_V + ST_N
};
static_assert(ST_N % 64 == 0 && ST_RO % 4 == 0, "float4 blocks of the static scratch must stay 16-byte aligned");
static __device__ __align__(256) uint32_t g_static_trace[ST_WORDS];

// (origin xyz, mint) / (direction xyz, maxt) records from the staged caller arrays; the defaults are trace_rays_common's
__global__ void __launch_bounds__(256) k_static_pack(uint32_t n, int has_mint, int has_maxt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *o = reinterpret_cast<const float *>(g_static_trace + ST_IN_O), *d = reinterpret_cast<const float *>(g_static_trace + ST_IN_D);
    const float mint = has_mint ? __uint_as_float(g_static_trace[ST_IN_MINT + i]) : BF_RAY_EPS;
    const float maxt = has_maxt ? __uint_as_float(g_static_trace[ST_IN_MAXT + i]) : bf_inf();
    reinterpret_cast<float4 *>(g_static_trace + ST_RO)[i] = make_float4(o[3 * i], o[3 * i + 1], o[3 * i + 2], mint);
    reinterpret_cast<float4 *>(g_static_trace + ST_RD)[i] = make_float4(d[3 * i], d[3 * i + 1], d[3 * i + 2], maxt);
}
// k_resolve_hits on the static hit records, plus the planar t / u / v the caller asked for
__global__ void __launch_bounds__(256) k_static_resolve(const __grid_constant__ DevScene sc, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 h = reinterpret_cast<const float4 *>(g_static_trace + ST_HIT)[i];
    const uint32_t slot = __float_as_uint(h.w);
    uint32_t shape = BF_NONE, prim = BF_NONE;
    if (!(h.x == bf_inf() || slot == BF_NONE)) {
        if (slot & 0x80000000u) { shape = sc.rects[slot & 0x7fffffffu].shape; prim = 0; }
        else { shape = __float_as_uint(__ldg(&sc.tris[3 * (size_t) slot]).w); prim = __float_as_uint(__ldg(&sc.tris[3 * (size_t) slot + 1]).w); }
    }
    g_static_trace[ST_SHAPE + i] = shape; g_static_trace[ST_PRIM + i] = prim;
    g_static_trace[ST_T + i] = __float_as_uint(h.x); g_static_trace[ST_U + i] = __float_as_uint(h.y); g_static_trace[ST_V + i] = __float_as_uint(h.z);
}

// Device address of g_static_trace on the CURRENT device (`device` is its ordinal): looked up once per device.
static cudaError_t static_trace_base(int device, uint32_t **base) {
    static std::mutex lock;
    static uint32_t *known[BF_STATIC_TRACE_MAX_DEVICES] = {};
    if (device < 0 || device >= BF_STATIC_TRACE_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> guard(lock);
    if (!known[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_trace);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        known[device] = static_cast<uint32_t *>(p);
    }
    *base = known[device];
    return cudaSuccess;
}

// Allocates nothing (host or device).  The caller arrays may be host or device memory (cudaMemcpyDefault); the copies back
// are queued on cfg.stream and the caller synchronises it.
cudaError_t static_trace_first_hits(const DevScene &sc, int device, const float *origins, const float *directions, const float *mint,
                                    const float *maxt, uint32_t n, uint32_t *shape, uint32_t *prim, float *t, float *u, float *v,
                                    const LaunchCfg &cfg, int *launches) {
    *launches = 0;
    if (n == 0) return cudaSuccess;
    if (n > ST_N) return cudaErrorInvalidValue;
    uint32_t *base = nullptr;
    cudaError_t e = static_trace_base(device, &base);
    if (e != cudaSuccess) return e;
    cudaStream_t st = cfg.stream;
    if ((e = cudaMemsetAsync(base + ST_CURSOR, 0, ST_HDR * 4, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(base + ST_IN_O, origins, (size_t) n * 12, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(base + ST_IN_D, directions, (size_t) n * 12, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if (mint && (e = cudaMemcpyAsync(base + ST_IN_MINT, mint, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if (maxt && (e = cudaMemcpyAsync(base + ST_IN_MAXT, maxt, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    k_static_pack<<<(n + 255) / 256, 256, 0, st>>>(n, mint != nullptr, maxt != nullptr);
    TraceIO io;
    memset(&io, 0, sizeof(io));
    io.ray_o = reinterpret_cast<const float4 *>(base + ST_RO); io.ray_d = reinterpret_cast<const float4 *>(base + ST_RD);
    io.n_direct = n; io.work_cursor = base + ST_CURSOR; io.hit = reinterpret_cast<float4 *>(base + ST_HIT);
    io.rounds = cfg.trace_rounds; io.sm_count = (uint32_t) cfg.sm_count;
    k_trace<false><<<trace_grid(cfg), TRACE_THREADS, 0, st>>>(sc, io);
    k_static_resolve<<<(n + 255) / 256, 256, 0, st>>>(sc, n);
    *launches = 3;
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (shape && (e = cudaMemcpyAsync(shape, base + ST_SHAPE, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if (prim && (e = cudaMemcpyAsync(prim, base + ST_PRIM, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if (t && (e = cudaMemcpyAsync(t, base + ST_T, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if (u && (e = cudaMemcpyAsync(u, base + ST_U, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    if (v && (e = cudaMemcpyAsync(v, base + ST_V, (size_t) n * 4, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    return cudaSuccess;
}

}  // namespace bf
