// beifong_b200 — host-side SAH BVH build and compression into 80-byte 8-wide nodes.
//
// Replaces the reference's CPU acceleration structures on the radar path:
//   Scene::accel_init_cpu  src/librender/scene_native.inl:3-10 (SAH kd-tree, kdtree.h:1891-2005)
//                          src/librender/scene_embree.inl:24-39 (Embree BVH)
//   Scene::accel_init_gpu  src/librender/scene_optix.inl:236-300 (OptiX GAS)
// Only triangles go into the BVH; analytic rectangles (src/shapes/rectangle.cpp) are tested
// directly by the traversal kernel.
//
// Node layout (5 x 16 B, read by the kernel as five 128-bit loads):
//   q0: p.x p.y p.z | e.x e.y e.z imask        origin of the quantisation grid, per-axis power-of-two
//                                              scale (biased exponent byte), internal-child mask
//   q1: child_base | tri_base | meta[0..3] | meta[4..7]
//   q2: qlo.x[0..7] | qlo.y[0..7]              8-bit child box corners, lo rounded down, hi rounded up
//   q3: qlo.z[0..7] | qhi.x[0..7]
//   q4: qhi.y[0..7] | qhi.z[0..7]
// meta[i]: 0 = empty slot; internal child: 0b001'11000 + i; leaf child: unary(count)<<5 | first
// triangle offset from tri_base (count <= 3, at most 24 triangles per node).
// Triangle record (3 x 16 B): (p0.xyz, shape) (p1.xyz, prim) (p2.xyz, flags): the full-precision
// world-space vertices the reference's Möller-Trumbore test (include/mitsuba/render/mesh.h:194-226)
// is evaluated on, so first hits agree bit for bit with the scalar CPU variant.
#pragma once
#include <stdint.h>
#include <vector>

namespace bf {

struct BvhTriIn {
    float p0[3], p1[3], p2[3];
    uint32_t shape, prim;
};

struct BvhStats {
    uint32_t n_nodes = 0, n_tris = 0, depth = 0;
    double sah_cost = 0.0;
    double build_ms = 0.0;
};

struct WideBvh {
    std::vector<uint32_t> nodes;     // 20 u32 per node
    std::vector<float> tris;         // 12 floats per triangle, leaf order
    std::vector<uint32_t> tri_src;   // leaf slot -> index into the input triangle array
    std::vector<uint32_t> level_first;   // nodes are in breadth-first order: index of the first node of every depth, + the node count
    float pad_abs = 0.f;
    BvhStats stats;
};

// pad_abs: every triangle's box is grown by this absolute margin (world units) before the build
// so that the conservative slab test never rejects a triangle its exact fp32 intersection test
// would accept.
void build_wide_bvh(const std::vector<BvhTriIn> &tris, float pad_abs, WideBvh &out);

}  // namespace bf
