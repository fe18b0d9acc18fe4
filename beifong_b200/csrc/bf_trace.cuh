// beifong_b200 — ray queries: 8-wide compressed-BVH traversal + Möller-Trumbore + analytic rectangles.
// Replaces Scene::ray_intersect_preliminary / Scene::ray_test (src/librender/scene.cpp:129-178) and the
// accelerators behind them (kdtree.h:2078-2171,2361-2390; scene_embree.inl:147-312; scene_optix.inl).
//
// Execution model: one persistent grid; every warp keeps its 32 lanes busy by claiming new rays from a
// global cursor whenever fewer than TRACE_REFILL lanes are still traversing (lanes of a warp finish at
// very different times: half of the receiver rays leave the scene after the root test).  A lane's
// traversal state lives in registers, its stack in shared memory (first BF_TRACE_SM_STACK entries) and
// local memory (the rest).  Results are committed at the refill points, where the warp is converged,
// so queue pushes are one atomic per warp.
#pragma once
#include "bf_wave.h"
#include "bf_device.cuh"

namespace bf {

#define TRACE_THREADS 128
#define TRACE_SM_STACK BF_TRACE_SM_STACK
#define TRACE_LOCAL_STACK BF_TRACE_LOCAL_STACK
#ifndef TRACE_REFILL
#define TRACE_REFILL 20
#endif
#ifndef TRACE_MIN_BLOCKS
#define TRACE_MIN_BLOCKS 8
#endif
#ifndef TRACE_TRI_QUEUE
#define TRACE_TRI_QUEUE 1    // k_trace: leaf triangle groups a lane can queue (besides the one it is testing) for the warp's triangle phase (0: test them right behind the node step)
#endif
#ifndef TRACE_TRI_LANES
#define TRACE_TRI_LANES 12   // the triangle phase starts when this many lanes hold queued triangles (and runs down to half of it)
#endif
#ifndef TRACE_TRI_MIN
#define TRACE_TRI_MIN 0      // > 0: triangle tests are postponed when fewer lanes than this are testing.  Measured slower on C2 (8 / 16 / 24: trace_closest 1.93 / 1.88 / 1.92 ms per 64 M samples against 1.69 ms without, profiles/r02e), so off
#endif

__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x) {
    uint32_t r;
    asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(r) : "r"(x));
    return r;
}
__device__ __forceinline__ uint32_t byte_of(uint32_t x, int i) { return (x >> (8 * i)) & 0xffu; }
// byte j of x as the float 32768 + byte: the byte lands in mantissa bits 8..15 of 0x47000000 (one PRMT
// instead of a quarter-rate I2F); the 32768 is folded into the additive term of the slab FMA.
__device__ __forceinline__ float magic_byte(uint32_t x, int j) {
    return __uint_as_float(__byte_perm(x, 0x47000000u, 0x7504u | ((uint32_t) j << 4)));
}

// Ray state is read once per query: keep it out of L1 so that the cache holds BVH nodes and triangles.
__device__ __forceinline__ float4 ld_stream(const float4 *p) {
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ld_stream(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.global.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

struct HitAcc {
    float t; float u, v; uint32_t slot; uint32_t shape, prim;
};

// Möller-Trumbore exactly as Mesh::ray_intersect_triangle (mesh.h:194-226): no backface culling,
// no determinant epsilon, u in [0,1], v >= 0, u+v <= 1, t in [mint, maxt].
__device__ __forceinline__ bool tri_test(const float4 *__restrict__ tris, uint32_t slot, V3 o, V3 d, float mint, float maxt,
                                         float &t_out, float &u_out, float &v_out, uint32_t &shape, uint32_t &prim) {
    float4 a = __ldg(&tris[3 * (size_t) slot]), b = __ldg(&tris[3 * (size_t) slot + 1]), c = __ldg(&tris[3 * (size_t) slot + 2]);
    V3 p0(a.x, a.y, a.z), p1(b.x, b.y, b.z), p2(c.x, c.y, c.z);
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = cross(d, e2);
    float inv_det = rcp(dot(e1, pvec));
    V3 tvec = o - p0;
    float u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    float v = dot(d, qvec) * inv_det;
    active &= v >= 0.f && u + v <= 1.f;
    float t = dot(e2, qvec) * inv_det;
    active &= t >= mint && t <= maxt;
    t_out = t; u_out = u; v_out = v;
    shape = __float_as_uint(a.w); prim = __float_as_uint(b.w);
    return active;
}

// Per-lane traversal state
struct TraceLane {
    V3 o, d;
    float mint, maxt;
    float idx, idy, idz;
    uint32_t octinv;
    HitAcc best;
    uint2 ngroup;
    int sp;
    bool occluded;
};

// Ray setup + analytic rectangles (rectangle.cpp:229-263).  Returns false when the query is already
// decided (any-hit occluded by a rectangle, or no BVH).
#define RECT_SM_FLOATS 19   // per rectangle in shared memory: 12 to_object, 1 shape id, 6 padded world box

template <bool ANY_HIT, bool CULL = true>
__device__ __forceinline__ bool trace_begin(const DevScene &sc, const float *s_rect, TraceLane &L) {
    L.best.t = L.maxt; L.best.u = 0.f; L.best.v = 0.f; L.best.slot = BF_NONE; L.best.shape = BF_NONE; L.best.prim = BF_NONE;
    L.occluded = false;
    const V3 o = L.o, d = L.d;
    const float tiny = 1e-30f;
    L.idx = __fdividef(1.f, fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x));
    L.idy = __fdividef(1.f, fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y));
    L.idz = __fdividef(1.f, fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z));
    const bool quick = L.mint > 0.f;
    const float t_lo = L.mint * 0.999f, t_hi = L.maxt * 1.001f;
    for (uint32_t r = 0; r < sc.n_rects; ++r) {
        const float *m = s_rect + 12 * r;
        float oz = fmadd(m[10], o.z, fmadd(m[9], o.y, fmadd(m[8], o.x, m[11])));
        float dz = fmadd(m[10], d.z, fmadd(m[9], d.y, m[8] * d.x));
        // Conservative rejections, so that a warp skips the exact test of a rectangle none of its rays can reach:
        // (1) t = -oz / dz is negative, or smaller than mint in magnitude (rays leaving the rectangle they start on);
        // (2) the ray misses the rectangle's padded world-space box.
        if (quick && (oz * dz > 0.f || fabsf(oz) < 0.999f * L.mint * fabsf(dz))) continue;
        {
            const float *bx = s_rect + 13 * BF_MAX_RECTS + 6 * r;
            float x0 = (bx[0] - o.x) * L.idx, x1 = (bx[3] - o.x) * L.idx;
            float y0 = (bx[1] - o.y) * L.idy, y1 = (bx[4] - o.y) * L.idy;
            float z0 = (bx[2] - o.z) * L.idz, z1 = (bx[5] - o.z) * L.idz;
            float t_in = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), t_lo));
            float t_out = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t_hi));
            if (t_in > t_out * 1.001f) continue;
        }
        float ox = fmadd(m[2], o.z, fmadd(m[1], o.y, fmadd(m[0], o.x, m[3])));
        float oy = fmadd(m[6], o.z, fmadd(m[5], o.y, fmadd(m[4], o.x, m[7])));
        float dx = fmadd(m[2], d.z, fmadd(m[1], d.y, m[0] * d.x));
        float dy = fmadd(m[6], d.z, fmadd(m[5], d.y, m[4] * d.x));
        float t = -oz * rcp(dz);
        float lx = fmadd(dx, t, ox), ly = fmadd(dy, t, oy);
        bool hit = t >= L.mint && t <= L.maxt && fabsf(lx) <= 1.f && fabsf(ly) <= 1.f;
        if (!hit) continue;
        if (ANY_HIT) { L.occluded = true; return false; }
        uint32_t shape = __float_as_uint(s_rect[12 * BF_MAX_RECTS + r]);
        if (t < L.best.t || (t == L.best.t && shape < L.best.shape)) {
            L.best.t = t; L.best.u = lx; L.best.v = ly; L.best.slot = 0x80000000u | r; L.best.shape = shape; L.best.prim = 0;
        }
    }
    if (sc.n_nodes == 0) return false;
    if (CULL) {   // Rays that miss the padded bounds of the meshes within [mint, closest hit so far] skip the traversal altogether
        // (C2: the car fills a few percent of the receiver's hemisphere; most shadow rays never come near it).
        float x0 = (sc.mesh_lo[0] - o.x) * L.idx, x1 = (sc.mesh_hi[0] - o.x) * L.idx;
        float y0 = (sc.mesh_lo[1] - o.y) * L.idy, y1 = (sc.mesh_hi[1] - o.y) * L.idy;
        float z0 = (sc.mesh_lo[2] - o.z) * L.idz, z1 = (sc.mesh_hi[2] - o.z) * L.idz;
        float t_in = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), t_lo));
        float t_out = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), L.best.t * 1.001f));
        if (t_in > t_out * 1.001f) return false;
    }
    L.octinv = (d.x < 0.f ? 0u : 4u) | (d.y < 0.f ? 0u : 2u) | (d.z < 0.f ? 0u : 1u);
    L.ngroup = make_uint2(0u, 0x80000000u);
    L.sp = 0;
    return true;
}

// One traversal step: pop one child of the current node group, test its 8 quantised child boxes, test
// the triangles of the leaves hit.  Returns false when the traversal is finished.
// Closest hit: lowest t, ties broken by the lowest (shape, prim) — the order a brute-force loop over the
// scene's shapes visits them in (ray_intersect_naive, scene_native.inl:43-61).
// The node half of a traversal step: pops one child of the current node group, tests its 8 quantised child boxes.  On return
// L.ngroup holds the inner children hit and `tgroup` the leaf triangles hit (base slot, bit mask).  Precondition: L.ngroup has a node.
__device__ __forceinline__ void node_step(const DevScene &sc, uint2 *s_stack, uint2 *l_stack, TraceLane &L, uint2 &tgroup) {
    const uint32_t tid = threadIdx.x;
    {
        const uint32_t hits_imask = L.ngroup.y;
        const uint32_t child_bit_index = 31u - __clz(hits_imask);
        const uint32_t child_node_base = L.ngroup.x;
        L.ngroup.y &= ~(1u << child_bit_index);
        if (L.ngroup.y > 0x00ffffffu) {
            if (L.sp < TRACE_SM_STACK) s_stack[L.sp * TRACE_THREADS + tid] = L.ngroup;
            else l_stack[L.sp - TRACE_SM_STACK] = L.ngroup;
            ++L.sp;
        }
        const uint32_t slot_index = (child_bit_index - 24u) ^ L.octinv;
        const uint32_t relative_index = __popc(hits_imask & ~(0xffffffffu << slot_index));
        const uint32_t node_index = child_node_base + relative_index;

        const uint4 *np = sc.nodes + 5 * (size_t) node_index;
        const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);

        const float px = __uint_as_float(n0.x), py = __uint_as_float(n0.y), pz = __uint_as_float(n0.z);
        const uint32_t ex = n0.w & 0xffu, ey = (n0.w >> 8) & 0xffu, ez = (n0.w >> 16) & 0xffu, imask = n0.w >> 24;
        L.ngroup.x = n1.x;
        tgroup.x = n1.y;
        uint32_t hitmask = 0;

        // t = (32768 + q) * a + (b - 32768 a); the folded constant costs at most |a| / 512 of accuracy, the
        // +-|a| / 128 slack below keeps the test conservative (a = one quantisation step in t)
        const float ax = __uint_as_float(ex << 23) * L.idx, ay = __uint_as_float(ey << 23) * L.idy, az = __uint_as_float(ez << 23) * L.idz;
        const float bx = fmaf(-32768.f, ax, (px - L.o.x) * L.idx), by = fmaf(-32768.f, ay, (py - L.o.y) * L.idy),
                    bz = fmaf(-32768.f, az, (pz - L.o.z) * L.idz);
        const float sx = fabsf(ax) * 0.0078125f, sy = fabsf(ay) * 0.0078125f, sz = fabsf(az) * 0.0078125f;
        const float bxl = bx - sx, bxh = bx + sx, byl = by - sy, byh = by + sy, bzl = bz - sz, bzh = bz + sz;
        const float tmin_ray = L.mint * (1.f - 4e-7f);
        const float tmax_ray = L.best.t * (1.f + 4e-7f);
        const uint32_t octinv4 = L.octinv * 0x01010101u;
        const bool nx = L.d.x < 0.f, ny = L.d.y < 0.f, nz = L.d.z < 0.f;

#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const uint32_t meta4 = half ? n1.w : n1.z;
            const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
            const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
            const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1f1f1f1fu;
            const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
            const uint32_t qlx = half ? n2.y : n2.x, qly = half ? n2.w : n2.z, qlz = half ? n3.y : n3.x;
            const uint32_t qhx = half ? n3.w : n3.z, qhy = half ? n4.y : n4.x, qhz = half ? n4.w : n4.z;
            const uint32_t x_min = nx ? qhx : qlx, x_max = nx ? qlx : qhx;
            const uint32_t y_min = ny ? qhy : qly, y_max = ny ? qly : qhy;
            const uint32_t z_min = nz ? qhz : qlz, z_max = nz ? qlz : qhz;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float tminx = fmaf(magic_byte(x_min, j), ax, bxl), tmaxx = fmaf(magic_byte(x_max, j), ax, bxh);
                float tminy = fmaf(magic_byte(y_min, j), ay, byl), tmaxy = fmaf(magic_byte(y_max, j), ay, byh);
                float tminz = fmaf(magic_byte(z_min, j), az, bzl), tmaxz = fmaf(magic_byte(z_max, j), az, bzh);
                float cmin = fmaxf(fmaxf(tminx, tminy), fmaxf(tminz, tmin_ray));
                float cmax = fminf(fminf(tmaxx, tmaxy), fminf(tmaxz, tmax_ray));
                if (cmin <= cmax)
                    hitmask |= byte_of(child_bits4, j) << byte_of(bit_index4, j);
            }
        }
        L.ngroup.y = (hitmask & 0xff000000u) | imask;
        tgroup.y = hitmask & 0x00ffffffu;
    }

}

template <bool ANY_HIT>
__device__ __forceinline__ bool trace_step(const DevScene &sc, uint2 *s_stack, uint2 *l_stack, TraceLane &L) {
    const uint32_t tid = threadIdx.x;
    uint2 tgroup = make_uint2(0u, 0u);
    if (L.ngroup.y <= 0x00ffffffu) {
        // the entry popped last was a parked triangle group (see below)
        tgroup = L.ngroup;
        L.ngroup = make_uint2(0u, 0u);
    } else {
        node_step(sc, s_stack, l_stack, L, tgroup);
    }

    // Triangle tests: the lanes of a warp have very different numbers of them after a node step.  Once fewer than
    // TRACE_TRI_MIN lanes are still testing, a lane that has node work left parks its remaining triangles on the stack
    // (an entry with only the low 24 bits set) and goes on with the nodes; the parked group is tested when it is popped,
    // with whatever lanes are in their triangle phase then (Ylitie, Karras, Laine 2017, "postponing").  Results do not
    // depend on the order: closest-hit ties are broken by (shape, prim).
    while (tgroup.y != 0) {
        if (TRACE_TRI_MIN > 0 && L.ngroup.y > 0x00ffffffu && __popc(__activemask()) < TRACE_TRI_MIN) break;
        const uint32_t i = 31u - __clz(tgroup.y);
        tgroup.y &= ~(1u << i);
        const uint32_t slot = tgroup.x + i;
        float t, u, v; uint32_t shape, prim;
        if (tri_test(sc.tris, slot, L.o, L.d, L.mint, L.maxt, t, u, v, shape, prim)) {
            if (ANY_HIT) { L.occluded = true; return false; }
            if (t < L.best.t || (t == L.best.t && (shape < L.best.shape || (shape == L.best.shape && prim < L.best.prim)))) {
                L.best.t = t; L.best.u = u; L.best.v = v; L.best.slot = slot; L.best.shape = shape; L.best.prim = prim;
            }
        }
    }

    if (tgroup.y != 0) {      // parked (only reached with node work left in L.ngroup)
        if (L.sp < TRACE_SM_STACK) s_stack[L.sp * TRACE_THREADS + tid] = tgroup;
        else l_stack[L.sp - TRACE_SM_STACK] = tgroup;
        ++L.sp;
    }
    if (L.ngroup.y <= 0x00ffffffu) {
        if (L.sp == 0) return false;
        --L.sp;
        L.ngroup = L.sp < TRACE_SM_STACK ? s_stack[L.sp * TRACE_THREADS + tid] : l_stack[L.sp - TRACE_SM_STACK];
    }
    return true;
}

__device__ __forceinline__ void stage_rects(const DevScene &sc, float *s_rect) {
    for (uint32_t i = threadIdx.x; i < sc.n_rects * 12u; i += blockDim.x) {
        uint32_t r = i / 12u, k = i % 12u;
        s_rect[i] = sc.rects[r].to_object.m[k / 4u][k % 4u];
    }
    for (uint32_t r = threadIdx.x; r < sc.n_rects; r += blockDim.x)
        s_rect[12 * BF_MAX_RECTS + r] = __uint_as_float(sc.rects[r].shape);
    for (uint32_t i = threadIdx.x; i < sc.n_rects * 6u; i += blockDim.x) {
        uint32_t r = i / 6u, k = i % 6u;
        s_rect[13 * BF_MAX_RECTS + i] = k < 3u ? sc.rects[r].box_lo[k] : sc.rects[r].box_hi[k - 3u];
    }
    __syncthreads();
}

// What a trace launch reads and writes.
struct TraceIO {
    const float4 *ray_o, *ray_d;     // (o, mint), (d, maxt) indexed by lane id
    const uint32_t *queue;           // lane ids, or null for lanes [0, n_direct)
    const uint32_t *count;           // device-side queue length
    uint32_t n_direct;
    float4 *hit;                     // closest: (t, u, v, slot)
    uint32_t *occl;                  // any-hit: 0 / 1
    // closest: lanes whose ray hit something, split by the material class of the surface hit (the shade kernel's
    // input: class 0 = diffuse, class 1 = everything else), so that a shading warp runs one BSDF.  May be null.
    uint32_t *hit_queue[3], *hit_count[3];   // [2] (DEFER): lanes whose ray enters the mesh bounds, for the dense follow-up launch
    float4 *x1;                      // closest, gen 3, depth > 1: a missed continuation ray leaves time = phase = -inf
    bf_sample_record *records;
    const uint32_t *sample_of;       // lane -> record index (null: identity)
    uint32_t depth;                  // the shade iteration this trace feeds
    uint32_t *steps;                 // optional debug output: traversal steps per ray
    uint32_t *work_cursor;
    unsigned long long *total;
    uint32_t rounds;                 // > 0: a launch uses only as many warps as give each of them ~`rounds` batches of 32 rays
    uint32_t sm_count;
};

// Work distribution: a warp claims chunks of consecutive rays from the global cursor and always has the NEXT chunk's
// claim in flight (the atomic is issued when the previous chunk is opened and its result is first read when that
// chunk runs dry), so the ~1 us round trip of a returning global atomic is off the critical path.  The same is
// done for hit-queue pushes: the slot reservation is issued at one commit point and consumed at the next.
// DEFER (closest queries of scenes where few rays reach the meshes): this launch decides the rays that miss the mesh bounds
// from the rectangles alone and queues the others (hit_queue[2]) for a second, plain launch that traverses them with dense
// warps -- the split k_primary<DEFER> makes for the receiver rays.
template <bool ANY_HIT, bool DEFER = false>
__global__ void __launch_bounds__(TRACE_THREADS, TRACE_MIN_BLOCKS) k_trace(const __grid_constant__ DevScene sc, const __grid_constant__ TraceIO io) {
    __shared__ float s_rect[RECT_SM_FLOATS * BF_MAX_RECTS];
    __shared__ uint2 s_stack[TRACE_SM_STACK * TRACE_THREADS];
    // closest-hit queries queue their triangle tests (measured on C3: 1 537 us against 1 675 us, 16.7 against 13.9 threads per
    // instruction, profiles/r02n); any-hit queries test on the spot (an early hit ends the ray: queueing made them slower)
    constexpr int TQ = ANY_HIT ? 0 : TRACE_TRI_QUEUE;
    __shared__ uint2 s_tq[TQ > 0 ? TQ * TRACE_THREADS : 1];         // per-lane queue of leaf triangle groups awaiting their tests
    uint2 l_stack[TRACE_LOCAL_STACK];
    stage_rects(sc, s_rect);
    const uint32_t n = io.queue ? *io.count : io.n_direct;
    if (io.total && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(io.total, (unsigned long long) n);
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    const bool push_hits = !ANY_HIT && io.hit_queue[0] != nullptr;

    // A launch is bound by its slowest rays: a warp that holds one batch of 32 rays runs every traversal step until the
    // last of them is done (mean 16 steps, 99.99th percentile 66), at 5-9 active lanes of 32 (profiles/r01h).  So a launch
    // of n rays uses only as many warps as give each of them ~`rounds` batches, refilled at TRACE_REFILL busy lanes from
    // the global cursor: the tail is paid once per warp instead of once per batch.  The active warps are spread over the
    // SMs first (block b runs on SM b % sm_count while the whole grid is resident), then over the SM's sub-partitions.
    const uint32_t n_warps_grid = gridDim.x * (TRACE_THREADS / 32);
    uint32_t n_warps = n_warps_grid, warp_rank = blockIdx.x * (TRACE_THREADS / 32) + (threadIdx.x >> 5);
    if (io.rounds && io.sm_count && gridDim.x % io.sm_count == 0) {
        const uint32_t floor_warps = min(n_warps_grid, io.sm_count * (TRACE_THREADS / 32));     // tiny launches: latency first
        n_warps = min(n_warps_grid, max((n + 32u * io.rounds - 1u) / (32u * io.rounds), floor_warps));
        warp_rank = ((blockIdx.x / io.sm_count) * (TRACE_THREADS / 32) + (threadIdx.x >> 5)) * io.sm_count + blockIdx.x % io.sm_count;
        if (warp_rank >= n_warps) return;
    }
    // Every warp starts on a static batch of 32 rays (no atomic, a small launch is spread over the whole grid); the
    // rest of the queue is handed out dynamically in chunks sized for >= ~8 claims per warp.
    const uint32_t static_total = n_warps * 32u;
    const bool dynamic = n > static_total;
    uint32_t chunk = 32;
    while (dynamic && chunk < 256 && chunk * 16 * n_warps <= n - static_total) chunk *= 2;
    // a launch smaller than one static batch per warp (the deferred-lane launches, deep bounces) is cut evenly over all
    // the warps instead: such a launch is bound by its slowest rays, and fewer rays per warp means more of them run at once
    const uint32_t quota = dynamic ? 32u : max((n + n_warps - 1u) / n_warps, 1u);
    uint32_t c_next = warp_rank * quota, c_end = c_next + quota;   // warp-uniform
    uint32_t claim = 0xffffffffu;                   // lane 0: start of the chunk claimed ahead
    if (dynamic && lane == 0) claim = static_total + atomic_add_async(io.work_cursor, chunk);
    bool out_of_work = c_next >= n;

    // deferred hit-queue push
    uint32_t pend_base = 0, pend_meta = 0;          // meta: bit 31 valid, bits 29..30 queue (class 0 / 1, 2 = deferred)
    constexpr int N_QUEUES = DEFER ? 3 : 2;
    bool deferred = false;

    enum { EMPTY = 0, BUSY = 1, FINISHED = 2, EXHAUSTED = 3 };
    int state = EMPTY;
    uint32_t l = 0, n_steps = 0, pend_l = 0;
    uint32_t qcnt = 0;                      // triangle groups queued in s_tq for this lane
    uint2 tpend = make_uint2(0u, 0u);       // the triangle group this lane is testing (base slot, mask of the triangles left)
    TraceLane L;
    while (true) {
        // ---- complete the push reserved at the previous commit point ----
        if (push_hits) {
            uint32_t any = __ballot_sync(0xffffffffu, pend_meta >> 31);
            if (any) {
#pragma unroll
                for (int c = 0; c < N_QUEUES; ++c) {
                    uint32_t m = __ballot_sync(0xffffffffu, (pend_meta >> 31) && ((pend_meta >> 29) & 3u) == (uint32_t) c);
                    if (m) {
                        uint32_t base = __shfl_sync(0xffffffffu, pend_base, __ffs(m) - 1);
                        if ((m >> lane) & 1u) io.hit_queue[c][base + __popc(m & lt_mask)] = pend_l;
                    }
                }
                pend_meta = 0;
            }
        }
        // ---- commit finished lanes (warp converged here) ----
        bool push = false;
        uint32_t cls = 0;
        if (state == FINISHED) {
            if (io.steps) io.steps[l] = n_steps;
            if (ANY_HIT) {
                io.occl[l] = L.occluded ? 1u : 0u;
            } else if (DEFER && deferred) {
                push = true; cls = 2;                // nothing decided yet: the follow-up launch redoes the query for this lane
            } else {
                const bool hit = L.best.slot != BF_NONE;
                io.hit[l] = make_float4(hit ? L.best.t : bf_inf(), L.best.u, L.best.v, __uint_as_float(L.best.slot));
                push = hit && push_hits;
                if (push) cls = sc.shapes[L.best.shape].mat_class;
                if (!hit && push_hits) {
                    // the miss branch of the integrator loop (pathtimefrequency.cpp:368-371 with t = inf, then
                    // `active &= si.is_valid()` ends the path): nothing else of the shade step is observable
                    if (io.depth > 1 && sc.integ.gen == BF_GEN_PHASE) {
                        float4 x = io.x1[l]; x.x = -bf_inf(); x.y = -bf_inf(); io.x1[l] = x;
                    }
                    if (io.records) io.records[io.sample_of ? io.sample_of[l] : l].depth = io.depth;
                }
            }
            state = EMPTY;
        }
        if (push_hits) {
#pragma unroll
            for (int c = 0; c < N_QUEUES; ++c) {
                uint32_t m = __ballot_sync(0xffffffffu, push && cls == (uint32_t) c);
                if (m) {
                    if (lane == (uint32_t) (__ffs(m) - 1)) pend_base = atomic_add_async(io.hit_count[c], (uint32_t) __popc(m));   // consumed next round
                    if ((m >> lane) & 1u) { pend_meta = 0x80000000u | ((uint32_t) c << 29); pend_l = l; }
                }
            }
        }
        // ---- refill empty lanes from the open chunk, running over into the chunk claimed ahead ----
        uint32_t want = __ballot_sync(0xffffffffu, state == EMPTY);
        if (want) {
            const uint32_t cnt = __popc(want), rank = __popc(want & lt_mask);
            const uint32_t avail = out_of_work ? 0u : c_end - c_next;
            uint32_t i = rank < avail ? c_next + rank : 0xffffffffu;
            if (!out_of_work && cnt > avail) {      // warp-uniform: open the next chunk (chunk >= 32 >= cnt - avail)
                const uint32_t nxt = __shfl_sync(0xffffffffu, claim, 0);
                if (nxt < n) {
                    if (lane == 0) claim = static_total + atomic_add_async(io.work_cursor, chunk);
                    if (rank >= avail) i = nxt + (rank - avail);
                    c_next = nxt + (cnt - avail); c_end = nxt + chunk;
                } else {
                    out_of_work = true;             // the cursor is past the end: so is every later claim
                }
            } else if (!out_of_work) {
                c_next += cnt;
            }
            if (!out_of_work && c_next >= n) out_of_work = true;
            if (state == EMPTY) {
                if (i < n) {
                    l = io.queue ? ld_stream(io.queue + i) : i;
                    float4 ro = ld_stream(io.ray_o + l), rd = ld_stream(io.ray_d + l);
                    L.o = V3(ro.x, ro.y, ro.z); L.d = V3(rd.x, rd.y, rd.z); L.mint = ro.w; L.maxt = rd.w;
                    n_steps = 0; qcnt = 0u; tpend = make_uint2(0u, 0u);
                    const bool traverse = trace_begin<ANY_HIT>(sc, s_rect, L);
                    if (DEFER) { deferred = traverse; state = FINISHED; }
                    else state = traverse ? BUSY : FINISHED;
                } else {
                    state = EXHAUSTED;              // only reached once the cursor is past the end
                }
            }
        }
        uint32_t busy = __ballot_sync(0xffffffffu, state == BUSY);
        uint32_t fin = __ballot_sync(0xffffffffu, state == FINISHED);
        if (busy == 0 && fin == 0) {
            if (__ballot_sync(0xffffffffu, pend_meta >> 31) == 0) break;
            continue;                                // one more round to flush the deferred push
        }
        if (busy == 0) continue;
        const bool drained = out_of_work;
        const int min_busy = drained ? 1 : TRACE_REFILL;
        // Most rays are decided in trace_begin (rectangles, mesh bounds): while rays remain, keep refilling until enough
        // lanes actually traverse, so that a node step is never run for a handful of lanes
        if (!drained && fin != 0 && __popc(busy) < min_busy) continue;
        // ---- traverse until too few lanes are left ----
        if (TQ == 0) {
            do {
                if (state == BUSY) {
                    ++n_steps;
                    if (!trace_step<ANY_HIT>(sc, s_stack, l_stack, L)) state = FINISHED;
                }
            } while (__popc(__ballot_sync(0xffffffffu, state == BUSY)) >= min_busy);
            continue;
        }
        // Two phases per iteration.  NODE: every busy lane does one node step; the leaf triangles it finds are not tested on the
        // spot (after a node step ~2 of 32 lanes have one: profiles/r02e) but kept as a (base slot, bit mask) group: the lane's
        // current group in registers, up to TQ more in its shared-memory queue.  TRIANGLE: when enough lanes hold a group (or
        // nothing else is left to do) those lanes test one triangle each per pass, for as long as enough of them have some; what
        // is left stays queued.  Results do not depend on the order of the tests: closest-hit ties are broken by (shape, prim).
        do {
            if (state == BUSY && qcnt < (uint32_t) TQ) {
                if (L.ngroup.y <= 0x00ffffffu && L.sp > 0) {
                    --L.sp;
                    L.ngroup = L.sp < TRACE_SM_STACK ? s_stack[L.sp * TRACE_THREADS + threadIdx.x] : l_stack[L.sp - TRACE_SM_STACK];
                }
                if (L.ngroup.y > 0x00ffffffu) {
                    uint2 tg;
                    ++n_steps; node_step(sc, s_stack, l_stack, L, tg);
                    if (tg.y != 0u) {
                        if (tpend.y == 0u) tpend = tg;
                        else { s_tq[qcnt * TRACE_THREADS + threadIdx.x] = tg; ++qcnt; }
                    }
                }
            }
            const uint32_t have_tri = __ballot_sync(0xffffffffu, state == BUSY && tpend.y != 0u);
            const uint32_t can_node = __ballot_sync(0xffffffffu, state == BUSY && qcnt < (uint32_t) TQ && (L.ngroup.y > 0x00ffffffu || L.sp > 0));
            if (have_tri != 0u && (__popc(have_tri) >= TRACE_TRI_LANES || can_node == 0u)) {
                uint32_t m = have_tri;
                do {
                    if (state == BUSY && tpend.y != 0u) {
                        const uint32_t i = 31u - __clz(tpend.y);
                        tpend.y &= ~(1u << i);
                        const uint32_t slot = tpend.x + i;
                        if (tpend.y == 0u && qcnt > 0u) { --qcnt; tpend = s_tq[qcnt * TRACE_THREADS + threadIdx.x]; }
                        float t, u, v; uint32_t shape, prim;
                        if (tri_test(sc.tris, slot, L.o, L.d, L.mint, L.maxt, t, u, v, shape, prim)) {
                            if (ANY_HIT) { L.occluded = true; state = FINISHED; qcnt = 0u; tpend.y = 0u; }
                            else if (t < L.best.t || (t == L.best.t && (shape < L.best.shape || (shape == L.best.shape && prim < L.best.prim)))) {
                                L.best.t = t; L.best.u = u; L.best.v = v; L.best.slot = slot; L.best.shape = shape; L.best.prim = prim;
                            }
                        }
                    }
                    m = __ballot_sync(0xffffffffu, state == BUSY && tpend.y != 0u);
                } while (m != 0u && (__popc(m) >= (TRACE_TRI_LANES + 1) / 2 || can_node == 0u));
            }
            // nothing left anywhere for this lane: the traversal is finished
            if (state == BUSY && tpend.y == 0u && L.ngroup.y <= 0x00ffffffu && L.sp == 0) state = FINISHED;
        } while (__popc(__ballot_sync(0xffffffffu, state == BUSY)) >= min_busy);
    }
}

}  // namespace bf
