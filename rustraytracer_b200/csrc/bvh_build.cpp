// Host BVH build + flatten for the GPU traversal kernels.
//
// Replaces BvhNode::new / handle_two (reference hittable.rs:637-752): the reference builds a
// median-split tree on a random axis with one primitive per leaf.  The reference's *hit set* is
// tree-independent (hit(prim) = f64 leaf-box test ∧ f64 primitive test), so
// the GPU side is free to use a better tree: binned SAH (16 bins), up to 4 primitives per leaf,
// f32 boxes rounded outward from the f64 primitive boxes.  Exactness is restored at the leaves
// by the f64 tests in traverse.cuh.
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <future>
#include <thread>
#include "rt_internal.h"

namespace rt {
namespace {

struct FBox { float mn[3], mx[3]; };

inline float round_down(double x) { float f = (float)x; if ((double)f > x) f = std::nextafterf(f, -INFINITY); return f; }
inline float round_up(double x) { float f = (float)x; if ((double)f < x) f = std::nextafterf(f, INFINITY); return f; }
inline void grow(FBox& a, const FBox& b) { for (int k = 0; k < 3; k++) { a.mn[k] = std::min(a.mn[k], b.mn[k]); a.mx[k] = std::max(a.mx[k], b.mx[k]); } }
inline FBox empty_box() { FBox b; for (int k = 0; k < 3; k++) { b.mn[k] = FLT_MAX; b.mx[k] = -FLT_MAX; } return b; }
inline float half_area(const FBox& b) {
    float dx = b.mx[0] - b.mn[0], dy = b.mx[1] - b.mn[1], dz = b.mx[2] - b.mn[2];
    if (dx < 0 || dy < 0 || dz < 0) return 0.f;
    return dx * dy + dy * dz + dz * dx;
}

struct TmpNode { FBox box; int32_t left, right; uint32_t first, count; };  // left<0 ⇒ leaf over order[first,first+count)

struct Builder {
    const std::vector<FBox>& pb;
    std::vector<uint32_t>& idx;
    std::vector<TmpNode> nodes;
    std::atomic<uint32_t> n_nodes{0};
    static constexpr int BINS = 16;
    uint32_t LEAF_MAX = 4;            // hard cap on primitives per leaf (the leaf word has room for 4); RT_SAH_MAXLEAF overrides (A/B runs)
    float c_isect = 1.5f;      // SAH: cost of one f64 leaf test relative to one node step (RT_SAH_ISECT overrides, for tuning)
    Builder(const std::vector<FBox>& b, std::vector<uint32_t>& i) : pb(b), idx(i) { nodes.resize(std::max<size_t>(2 * b.size(), 2)); }

    uint32_t alloc() { return n_nodes.fetch_add(1); }

    void build(uint32_t node, uint32_t first, uint32_t count, int par_depth) {
        FBox bounds = empty_box(), cb = empty_box();
        for (uint32_t i = first; i < first + count; i++) {
            const FBox& b = pb[idx[i]];
            grow(bounds, b);
            for (int k = 0; k < 3; k++) { float c = 0.5f * (b.mn[k] + b.mx[k]); cb.mn[k] = std::min(cb.mn[k], c); cb.mx[k] = std::max(cb.mx[k], c); }
        }
        TmpNode& nd = nodes[node];
        nd.box = bounds; nd.first = first; nd.count = count; nd.left = nd.right = -1;
        if (count <= 1) return;
        // binned SAH on the widest centroid axis (try all three axes)
        float best_cost = FLT_MAX; int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; axis++) {
            float lo = cb.mn[axis], ext = cb.mx[axis] - lo;
            if (!(ext > 0.f)) continue;
            float scale = (float)BINS / ext;
            FBox bb[BINS]; uint32_t bc[BINS];
            for (int b = 0; b < BINS; b++) { bb[b] = empty_box(); bc[b] = 0; }
            for (uint32_t i = first; i < first + count; i++) {
                const FBox& b = pb[idx[i]];
                int bin = std::min(BINS - 1, std::max(0, (int)((0.5f * (b.mn[axis] + b.mx[axis]) - lo) * scale)));
                grow(bb[bin], b); bc[bin]++;
            }
            float right_area[BINS]; uint32_t right_cnt[BINS];
            FBox acc = empty_box(); uint32_t c = 0;
            for (int b = BINS - 1; b > 0; b--) { grow(acc, bb[b]); c += bc[b]; right_area[b] = half_area(acc); right_cnt[b] = c; }
            acc = empty_box(); c = 0;
            for (int b = 0; b < BINS - 1; b++) {
                grow(acc, bb[b]); c += bc[b];
                if (c == 0 || right_cnt[b + 1] == 0) continue;
                float cost = half_area(acc) * (float)c + right_area[b + 1] * (float)right_cnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        uint32_t mid;
        if (best_axis < 0) {
            if (count <= LEAF_MAX) return;
            mid = first + count / 2;                       // all centroids coincide: split by index
        } else {
            // leaf cost: count * C_isect ; split cost: C_trav + SAH.  C_isect/C_trav = 1.5 (f64 leaf test)
            float ha = half_area(bounds);
            float split_cost = 1.0f + c_isect * best_cost / std::max(ha, 1e-30f);
            if (count <= LEAF_MAX && c_isect * (float)count <= split_cost) return;
            float lo = cb.mn[best_axis], scale = (float)BINS / (cb.mx[best_axis] - lo);
            auto it = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t p) {
                const FBox& b = pb[p];
                int bin = std::min(BINS - 1, std::max(0, (int)((0.5f * (b.mn[best_axis] + b.mx[best_axis]) - lo) * scale)));
                return bin <= best_bin;
            });
            mid = (uint32_t)(it - idx.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        uint32_t l = alloc(), r = alloc();
        nodes[node].left = (int32_t)l; nodes[node].right = (int32_t)r;
        uint32_t lc = mid - first, rc = first + count - mid;
        if (par_depth > 0 && count > 65536) {
            auto fut = std::async(std::launch::async, [=]() { build(l, first, lc, par_depth - 1); });
            build(r, mid, rc, par_depth - 1);
            fut.get();
        } else {
            build(l, first, lc, 0);
            build(r, mid, rc, 0);
        }
    }
};

inline int32_t encode_leaf(uint32_t first_slot, uint32_t count) { return ~(int32_t)((first_slot << 2) | (count - 1)); }

}  // namespace

// Conservativeness of the f32 boxes (the same rule in bvh_device.cu:k_leaves).  The traversal's f32 slab test (traverse.cuh:node4_test) must never
// reject a child whose f64 leaf test (hittable.rs:494-508 on the primitive's own f64 box) would pass.  Its error budget is carried on the RAY side:
//   * planes: f64 -> f32 rounded outward here, so the f32 box contains the f64 box;
//   * origin: o32 = fl(o), |o32 - o| <= 2^-24 |o|  -> make_rayf's `pad` (4.8e-7 max|o| per unit of |1/d|, 8x the bound);
//   * q = fl(o32 * idx): 2^-24 |q|                 -> + 2.4e-7 |q| (4x);
//   * idx = fl(1 / fl(d)): relative 2^-23, one rounding in the fma: 2^-24 -> t0 -= 4.8e-7 |t0|, t1 += 4.8e-7 |t1| (2.7x).
// That already covers every rounding in the test; the relative pad of 2^-22 of the coordinate magnitude added here on top is a second,
// independent margin (it loosens a box by 0.25 ppm and was measured to cost nothing), kept identical in both builders so that they cull alike.
rt_status build_bvh(const std::vector<Box64>& boxes, BuiltBvh& out, int threads) {
    out.nodes.clear(); out.order.clear(); out.max_stack = 0;
    size_t n = boxes.size();
    if (n == 0) return RT_OK;
    if (n >= (1ull << 29)) { set_error("too many primitives for the leaf encoding (2^29)"); return RT_ERR_INVALID; }
    std::vector<FBox> pb(n);
    for (size_t i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) {
            // outward rounding plus a relative pad of 2^-22 of the coordinate magnitude (covers the f32
            // rounding of the slab arithmetic against the box's own coordinates; the ray-origin part is
            // padded per ray in the kernel)
            double lo = boxes[i].mn[k], hi = boxes[i].mx[k];
            double pad = (std::fabs(lo) + std::fabs(hi)) * 2.4e-7;
            pb[i].mn[k] = round_down(lo - pad); pb[i].mx[k] = round_up(hi + pad);
        }
    std::vector<uint32_t> idx(n);
    for (size_t i = 0; i < n; i++) idx[i] = (uint32_t)i;
    Builder b(pb, idx);
    if (const char* e = getenv("RT_SAH_ISECT")) { float v = (float)atof(e); if (v > 0.f) b.c_isect = v; }
    if (const char* e = getenv("RT_SAH_MAXLEAF")) { int v = atoi(e); if (v >= 1 && v <= 4) b.LEAF_MAX = (uint32_t)v; }
    uint32_t root = b.alloc();
    int par_depth = 0; while ((1 << par_depth) < threads) par_depth++;
    b.build(root, 0, (uint32_t)n, par_depth + 1);
    uint32_t total = b.n_nodes.load();
#if RT_BVH_WIDTH == 4
    // collapse to arity 4: a BVH4 node takes a BVH2 inner node's two children and keeps replacing the inner child with the
    // largest surface area by its own two children until it has four (or only leaves remain); leaves are folded into the
    // parent's child words.  Breadth-first numbering, root = 0.
    (void)total;
    auto set_child = [&](DNode4& d, int k, const FBox& bx, int32_t word) {
        d.mnx[k] = bx.mn[0]; d.mxx[k] = bx.mx[0]; d.mny[k] = bx.mn[1]; d.mxy[k] = bx.mx[1]; d.mnz[k] = bx.mn[2]; d.mxz[k] = bx.mx[2]; d.child[k] = word;
    };
    FBox never; for (int k = 0; k < 3; k++) { never.mn[k] = FLT_MAX; never.mx[k] = -FLT_MAX; }
    if (b.nodes[root].left < 0) {
        DNode4 d; memset(&d, 0, sizeof(d));
        set_child(d, 0, b.nodes[root].box, encode_leaf(0, b.nodes[root].count));
        for (int k = 1; k < 4; k++) set_child(d, k, never, EMPTY_CHILD);
        out.nodes.push_back(d);
    } else {
        std::vector<uint32_t> queue;            // BVH2 inner node behind every BVH4 node, in BVH4 index order
        queue.push_back(root);
        out.nodes.reserve(n / 2 + 16);
        for (size_t qi = 0; qi < queue.size(); qi++) {
            const TmpNode& t = b.nodes[queue[qi]];
            uint32_t kids[4] = {(uint32_t)t.left, (uint32_t)t.right, 0, 0}; int nk = 2;
            while (nk < 4) {
                int pick = -1; float best = -1.f;
                for (int k = 0; k < nk; k++) if (b.nodes[kids[k]].left >= 0) { float a = half_area(b.nodes[kids[k]].box); if (a > best) { best = a; pick = k; } }
                if (pick < 0) break;
                const TmpNode& c = b.nodes[kids[pick]];
                kids[pick] = (uint32_t)c.left; kids[nk++] = (uint32_t)c.right;
            }
            DNode4 d; memset(&d, 0, sizeof(d));
            for (int k = 0; k < 4; k++) {
                if (k >= nk) { set_child(d, k, never, EMPTY_CHILD); continue; }
                const TmpNode& c = b.nodes[kids[k]];
                if (c.left >= 0) { set_child(d, k, c.box, (int32_t)queue.size()); queue.push_back(kids[k]); }
                else set_child(d, k, c.box, encode_leaf(c.first, c.count));
            }
            out.nodes.push_back(d);
        }
        // worst-case traversal stack: descending into one child leaves the node's other children on the stack
        std::vector<int> need(out.nodes.size(), 0);
        for (size_t i = out.nodes.size(); i-- > 0;) {
            int kids_n = 0, deepest = 0;
            for (int k = 0; k < 4; k++) {
                int32_t c = out.nodes[i].child[k];
                if (c == EMPTY_CHILD) continue;
                kids_n++;
                if (c >= 0) deepest = std::max(deepest, need[c]);
            }
            need[i] = kids_n - 1 + deepest;
        }
        out.max_stack = need[0] + 1;            // + the sentinel
    }
#else
    // flatten: internal nodes only; leaves are folded into the parent's child words
    std::vector<int32_t> remap(total, -1);
    uint32_t n_internal = 0;
    for (uint32_t i = 0; i < total; i++) if (b.nodes[i].left >= 0) remap[i] = (int32_t)n_internal++;
    auto child_word = [&](uint32_t c) -> int32_t { const TmpNode& t = b.nodes[c]; return t.left >= 0 ? remap[c] : encode_leaf(t.first, t.count); };
    auto put_child = [&](DNode& d, int side, const FBox& bx) {
        int o = side * 4;
        d.v[o + 0] = bx.mn[0]; d.v[o + 1] = bx.mx[0]; d.v[o + 2] = bx.mn[1]; d.v[o + 3] = bx.mx[1];
        d.v[8 + side * 2] = bx.mn[2]; d.v[9 + side * 2] = bx.mx[2];
    };
    if (n_internal == 0) {
        // the whole scene is one leaf: synthesise a root whose right child can never be hit
        DNode d; memset(&d, 0, sizeof(d));
        put_child(d, 0, b.nodes[root].box);
        FBox never; for (int k = 0; k < 3; k++) { never.mn[k] = FLT_MAX; never.mx[k] = -FLT_MAX; }
        put_child(d, 1, never);
        d.left = encode_leaf(0, b.nodes[root].count); d.right = encode_leaf(0, 1);
        out.nodes.push_back(d);
    } else {
        out.nodes.resize(n_internal);
        // root must be node 0: alloc order guarantees root == 0 and remap[0] == 0
        for (uint32_t i = 0; i < total; i++) {
            const TmpNode& t = b.nodes[i];
            if (t.left < 0) continue;
            DNode d; memset(&d, 0, sizeof(d));
            put_child(d, 0, b.nodes[t.left].box); put_child(d, 1, b.nodes[t.right].box);
            d.left = child_word((uint32_t)t.left); d.right = child_word((uint32_t)t.right);
            out.nodes[remap[i]] = d;
        }
    }
#endif
    out.order = std::move(idx);
    return RT_OK;
}

}  // namespace rt
