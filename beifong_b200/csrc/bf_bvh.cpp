// beifong_b200 — binned-SAH binary BVH build, collapse to 8-wide, 80-byte node compression.
// See bf_bvh.h for the layout and the reference interfaces this replaces.
#include "bf_bvh.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <thread>

namespace bf {
namespace {

constexpr int kBins = 16;
constexpr float kCostNode = 1.0f;   // relative cost of one wide-node visit
constexpr float kCostTri = 0.3f;    // relative cost of one triangle test
constexpr uint32_t kLeafMax = 3;    // triangles per leaf child (unary count in 3 bits)

// Share of a wide-node visit charged to one more binary level when a group of <= 3 triangles decides between staying one leaf
// child and being split further (build-time experiment knob BF_BVH_SPLIT_SHARE; results never depend on the tree shape).
inline float split_node_share() {
    static const float v = [] { const char *e = std::getenv("BF_BVH_SPLIT_SHARE"); float x = e ? (float) std::atof(e) : 0.25f; return x > 0.f ? x : 0.25f; }();
    return v;
}

struct Ref {
    float lo[3], hi[3];
    uint32_t id;
};
struct Node2 {
    float lo[3], hi[3];
    uint32_t left, right;   // children (internal)
    uint32_t first, count;  // refs range (leaf: count > 0)
};

inline float half_area(const float *lo, const float *hi) {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

struct Builder {
    std::vector<Ref> refs;
    std::vector<Node2> nodes;
    std::atomic<uint32_t> next_node{0};
    std::atomic<int> live_threads{0};
    int max_threads = 1;

    void bounds(uint32_t first, uint32_t count, float *lo, float *hi, float *clo, float *chi) const {
        for (int a = 0; a < 3; ++a) {
            lo[a] = clo[a] = std::numeric_limits<float>::infinity();
            hi[a] = chi[a] = -std::numeric_limits<float>::infinity();
        }
        for (uint32_t i = first; i < first + count; ++i) {
            const Ref &r = refs[i];
            for (int a = 0; a < 3; ++a) {
                lo[a] = std::min(lo[a], r.lo[a]);
                hi[a] = std::max(hi[a], r.hi[a]);
                float c = 0.5f * (r.lo[a] + r.hi[a]);
                clo[a] = std::min(clo[a], c);
                chi[a] = std::max(chi[a], c);
            }
        }
    }

    void build(uint32_t node, uint32_t first, uint32_t count, int depth) {
        float lo[3], hi[3], clo[3], chi[3];
        bounds(first, count, lo, hi, clo, chi);
        Node2 &n = nodes[node];
        std::memcpy(n.lo, lo, 12);
        std::memcpy(n.hi, hi, 12);
        n.left = n.right = 0;
        n.first = first;
        n.count = count;
        if (count == 1) return;

        // binned SAH over the three axes
        float best_cost = std::numeric_limits<float>::infinity();
        int best_axis = -1, best_split = -1;
        for (int a = 0; a < 3; ++a) {
            float ext = chi[a] - clo[a];
            if (!(ext > 0.f)) continue;
            float blo[kBins][3], bhi[kBins][3];
            uint32_t bcnt[kBins];
            for (int b = 0; b < kBins; ++b) {
                bcnt[b] = 0;
                for (int k = 0; k < 3; ++k) {
                    blo[b][k] = std::numeric_limits<float>::infinity();
                    bhi[b][k] = -std::numeric_limits<float>::infinity();
                }
            }
            float scale = (float) kBins / ext;
            for (uint32_t i = first; i < first + count; ++i) {
                const Ref &r = refs[i];
                float c = 0.5f * (r.lo[a] + r.hi[a]);
                int b = std::min(kBins - 1, std::max(0, (int) ((c - clo[a]) * scale)));
                bcnt[b]++;
                for (int k = 0; k < 3; ++k) {
                    blo[b][k] = std::min(blo[b][k], r.lo[k]);
                    bhi[b][k] = std::max(bhi[b][k], r.hi[k]);
                }
            }
            float rarea[kBins];
            uint32_t rcnt[kBins];
            float alo[3], ahi[3];
            for (int k = 0; k < 3; ++k) { alo[k] = std::numeric_limits<float>::infinity(); ahi[k] = -alo[k]; }
            uint32_t acc = 0;
            for (int b = kBins - 1; b > 0; --b) {
                acc += bcnt[b];
                for (int k = 0; k < 3; ++k) { alo[k] = std::min(alo[k], blo[b][k]); ahi[k] = std::max(ahi[k], bhi[b][k]); }
                rcnt[b] = acc;
                rarea[b] = acc ? half_area(alo, ahi) : 0.f;
            }
            for (int k = 0; k < 3; ++k) { alo[k] = std::numeric_limits<float>::infinity(); ahi[k] = -alo[k]; }
            acc = 0;
            for (int b = 0; b < kBins - 1; ++b) {
                acc += bcnt[b];
                for (int k = 0; k < 3; ++k) { alo[k] = std::min(alo[k], blo[b][k]); ahi[k] = std::max(ahi[k], bhi[b][k]); }
                if (acc == 0 || rcnt[b + 1] == 0) continue;
                float cost = half_area(alo, ahi) * (float) acc + rarea[b + 1] * (float) rcnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = b; }
            }
        }

        float area = half_area(lo, hi);
        if (count <= kLeafMax) {
            float leaf_cost = kCostTri * area * (float) count;
            float split_cost = kCostNode * area * split_node_share() + kCostTri * best_cost;  // a binary level is a fraction of a wide level
            if (best_axis < 0 || leaf_cost <= split_cost) return;  // stay a leaf
        }

        uint32_t mid;
        if (best_axis >= 0) {
            float ext = chi[best_axis] - clo[best_axis];
            float scale = (float) kBins / ext;
            float c0 = clo[best_axis];
            int a = best_axis, s = best_split;
            Ref *b = refs.data() + first, *e = b + count;
            Ref *m = std::partition(b, e, [=](const Ref &r) {
                float c = 0.5f * (r.lo[a] + r.hi[a]);
                int bin = std::min(kBins - 1, std::max(0, (int) ((c - c0) * scale)));
                return bin <= s;
            });
            mid = first + (uint32_t) (m - b);
        } else {
            mid = first + count / 2;  // all centroids coincide
        }
        if (mid == first || mid == first + count) mid = first + count / 2;

        uint32_t l = next_node.fetch_add(2);
        n.left = l;
        n.right = l + 1;
        n.count = 0;
        uint32_t lc = mid - first, rc = first + count - mid;
        if (count > 65536 && live_threads.load() < max_threads) {
            live_threads.fetch_add(1);
            std::thread t([this, l, first, lc, depth]() { build(l, first, lc, depth + 1); live_threads.fetch_sub(1); });
            build(l + 1, mid, rc, depth + 1);
            t.join();
        } else {
            build(l, first, lc, depth + 1);
            build(l + 1, mid, rc, depth + 1);
        }
    }
};

inline uint32_t f2u_(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

}  // namespace

void build_wide_bvh(const std::vector<BvhTriIn> &tris, float pad_abs, WideBvh &out) {
    auto t0 = std::chrono::steady_clock::now();
    out.nodes.clear();
    out.tris.clear();
    out.tri_src.clear();
    out.level_first.clear();
    out.stats = BvhStats();
    const uint32_t N = (uint32_t) tris.size();
    if (N == 0) return;

    Builder B;
    B.refs.resize(N);
    for (uint32_t i = 0; i < N; ++i) {
        const BvhTriIn &t = tris[i];
        Ref &r = B.refs[i];
        for (int a = 0; a < 3; ++a) {
            float lo = std::min(t.p0[a], std::min(t.p1[a], t.p2[a]));
            float hi = std::max(t.p0[a], std::max(t.p1[a], t.p2[a]));
            // absolute pad + a relative one for coordinates much larger than the scene scale hint
            float pad = pad_abs + 4e-7f * std::max(std::fabs(lo), std::fabs(hi));
            r.lo[a] = lo - pad;
            r.hi[a] = hi + pad;
        }
        r.id = i;
    }
    B.nodes.resize(2 * (size_t) N);
    B.next_node = 1;
    B.max_threads = (int) std::max(1u, std::min(16u, std::thread::hardware_concurrency())) - 1;
    B.build(0, 0, N, 0);

    // ---- collapse to 8-wide, breadth first so that the top of the tree is contiguous -----------------
    struct Item { uint32_t n2; uint32_t wide; uint32_t depth; };
    std::deque<Item> queue;
    out.nodes.assign(20, 0u);
    queue.push_back({0u, 0u, 1u});
    out.tris.reserve(12 * (size_t) N);
    out.tri_src.reserve(N);
    double sah = 0.0;
    const float root_area = std::max(half_area(B.nodes[0].lo, B.nodes[0].hi), 1e-30f);
    uint32_t max_depth = 0;

    while (!queue.empty()) {
        Item it = queue.front();
        queue.pop_front();
        if (it.depth > max_depth) out.level_first.push_back(it.wide);      // breadth first: depths never decrease along the queue
        max_depth = std::max(max_depth, it.depth);
        const Node2 &root = B.nodes[it.n2];
        uint32_t child[8];
        int k = 0;
        if (root.count > 0) {
            child[k++] = it.n2;  // degenerate: the whole tree is one leaf
        } else {
            child[k++] = root.left;
            child[k++] = root.right;
            while (k < 8) {
                int best = -1;
                float best_a = -1.f;
                for (int i = 0; i < k; ++i) {
                    const Node2 &c = B.nodes[child[i]];
                    if (c.count > 0) continue;
                    float a = half_area(c.lo, c.hi);
                    if (a > best_a) { best_a = a; best = i; }
                }
                if (best < 0) break;
                const Node2 &c = B.nodes[child[best]];
                child[best] = c.left;
                child[k++] = c.right;
            }
        }
        sah += kCostNode * half_area(root.lo, root.hi) / root_area;

        // node frame
        float p[3], hi[3];
        for (int a = 0; a < 3; ++a) { p[a] = root.lo[a]; hi[a] = root.hi[a]; }
        int ebyte[3];
        float scale[3];
        for (int a = 0; a < 3; ++a) {
            double ext = (double) hi[a] - (double) p[a];
            int e = ext > 0.0 ? (int) std::ceil(std::log2(ext / 255.0)) : -100;
            e = std::max(-120, std::min(120, e));
            // make sure ceil((hi - p) / 2^e) <= 255 also after fp32 rounding of p + q * 2^e
            while (std::ldexp(255.0, e) < ext * (1.0 + 1e-6)) ++e;
            ebyte[a] = e + 127;
            scale[a] = std::ldexp(1.0f, e);
        }

        // slot assignment: slot s prefers children lying towards ((s&4)?+:-, (s&2)?+:-, (s&1)?+:-)
        float cc[8][3];
        for (int i = 0; i < k; ++i)
            for (int a = 0; a < 3; ++a)
                cc[i][a] = 0.5f * (B.nodes[child[i]].lo[a] + B.nodes[child[i]].hi[a]) - 0.5f * (p[a] + hi[a]);
        int slot_of[8];
        int child_in_slot[8];
        for (int s = 0; s < 8; ++s) child_in_slot[s] = -1;
        for (int i = 0; i < 8; ++i) slot_of[i] = -1;
        for (int round = 0; round < k; ++round) {
            float best = -std::numeric_limits<float>::infinity();
            int bi = -1, bs = -1;
            for (int i = 0; i < k; ++i) {
                if (slot_of[i] >= 0) continue;
                for (int s = 0; s < 8; ++s) {
                    if (child_in_slot[s] >= 0) continue;
                    float sc = ((s & 4) ? cc[i][0] : -cc[i][0]) + ((s & 2) ? cc[i][1] : -cc[i][1]) + ((s & 1) ? cc[i][2] : -cc[i][2]);
                    if (sc > best) { best = sc; bi = i; bs = s; }
                }
            }
            slot_of[bi] = bs;
            child_in_slot[bs] = bi;
        }

        uint32_t imask = 0;
        uint8_t meta[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint8_t qlo[3][8], qhi[3][8];
        std::memset(qlo, 0, sizeof(qlo));
        std::memset(qhi, 0, sizeof(qhi));
        uint32_t child_base = (uint32_t) (out.nodes.size() / 20);
        uint32_t tri_base = (uint32_t) (out.tris.size() / 12);
        uint32_t n_internal = 0, tri_off = 0;
        for (int s = 0; s < 8; ++s) {
            int ci = child_in_slot[s];
            if (ci < 0) {
                // empty slot: inverted box, never hit; meta = 0 contributes no bits anyway
                for (int a = 0; a < 3; ++a) { qlo[a][s] = 255; qhi[a][s] = 0; }
                continue;
            }
            const Node2 &c = B.nodes[child[ci]];
            for (int a = 0; a < 3; ++a) {
                double lo_q = std::floor(((double) c.lo[a] - (double) p[a]) / (double) scale[a]);
                double hi_q = std::ceil(((double) c.hi[a] - (double) p[a]) / (double) scale[a]);
                int ql = (int) std::max(0.0, std::min(255.0, lo_q));
                int qh = (int) std::max(0.0, std::min(255.0, hi_q));
                while (ql > 0 && p[a] + (float) ql * scale[a] > c.lo[a]) --ql;
                while (qh < 255 && p[a] + (float) qh * scale[a] < c.hi[a]) ++qh;
                qlo[a][s] = (uint8_t) ql;
                qhi[a][s] = (uint8_t) qh;
            }
            if (c.count == 0) {
                imask |= 1u << s;
                meta[s] = (uint8_t) ((1u << 5) | (24u + (uint32_t) s));
                ++n_internal;
            } else {
                uint32_t unary = c.count == 1 ? 1u : (c.count == 2 ? 3u : 7u);
                meta[s] = (uint8_t) ((unary << 5) | tri_off);
                for (uint32_t i = 0; i < c.count; ++i) {
                    uint32_t src = B.refs[c.first + i].id;
                    const BvhTriIn &t = tris[src];
                    float rec[12] = {t.p0[0], t.p0[1], t.p0[2], 0.f, t.p1[0], t.p1[1], t.p1[2], 0.f, t.p2[0], t.p2[1], t.p2[2], 0.f};
                    std::memcpy(&rec[3], &t.shape, 4);
                    std::memcpy(&rec[7], &t.prim, 4);
                    out.tris.insert(out.tris.end(), rec, rec + 12);
                    out.tri_src.push_back(src);
                }
                tri_off += c.count;
                sah += kCostTri * (double) c.count * half_area(c.lo, c.hi) / root_area;
            }
        }
        // reserve the internal children contiguously, in slot order
        out.nodes.resize(out.nodes.size() + 20 * (size_t) n_internal, 0u);
        uint32_t rel = 0;
        for (int s = 0; s < 8; ++s) {
            int ci = child_in_slot[s];
            if (ci < 0 || B.nodes[child[ci]].count != 0) continue;
            queue.push_back({child[ci], child_base + rel, it.depth + 1});
            ++rel;
        }

        uint32_t *w = out.nodes.data() + 20 * (size_t) it.wide;
        w[0] = f2u_(p[0]);
        w[1] = f2u_(p[1]);
        w[2] = f2u_(p[2]);
        w[3] = (uint32_t) ebyte[0] | ((uint32_t) ebyte[1] << 8) | ((uint32_t) ebyte[2] << 16) | (imask << 24);
        w[4] = child_base;
        w[5] = tri_base;
        auto pack4 = [](const uint8_t *b) { return (uint32_t) b[0] | ((uint32_t) b[1] << 8) | ((uint32_t) b[2] << 16) | ((uint32_t) b[3] << 24); };
        w[6] = pack4(meta);
        w[7] = pack4(meta + 4);
        w[8] = pack4(qlo[0]);  w[9] = pack4(qlo[0] + 4);
        w[10] = pack4(qlo[1]); w[11] = pack4(qlo[1] + 4);
        w[12] = pack4(qlo[2]); w[13] = pack4(qlo[2] + 4);
        w[14] = pack4(qhi[0]); w[15] = pack4(qhi[0] + 4);
        w[16] = pack4(qhi[1]); w[17] = pack4(qhi[1] + 4);
        w[18] = pack4(qhi[2]); w[19] = pack4(qhi[2] + 4);
    }

    out.level_first.push_back((uint32_t) (out.nodes.size() / 20));
    out.pad_abs = pad_abs;
    out.stats.n_nodes = (uint32_t) (out.nodes.size() / 20);
    out.stats.n_tris = (uint32_t) (out.tris.size() / 12);
    out.stats.depth = max_depth;
    out.stats.sah_cost = sah;
    out.stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace bf
