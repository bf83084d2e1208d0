// Device BVH build: LBVH over 63-bit Morton codes, collapsed to the traversal kernels' BVH4 on the GPU.
//
// Replaces BvhNode::new / handle_two (reference hittable.rs:637-752: a recursive median split on a random axis, one primitive per
// leaf, single-threaded) for scenes where the time to the first pixel matters more than the last 20-30 % of traversal speed: the
// host binned-SAH build (bvh_build.cpp) takes 0.7 s per million triangles, this one a few milliseconds.  The hit set does not
// depend on the tree (hit(prim) = f64 leaf box AND f64 primitive test, decided in traverse.cuh), so both builders give
// bit-identical renders; only the number of node steps differs.
//
// Tree quality (round 2, profiles/lbvh_quality_r02.json).  Round 1's trees walked 12-28 % slower than the host binned-SAH tree.  Two things were tried:
//   * LEAF SIZE — what closed the gap.  The collapse used to turn every subtree of <= 4 primitives into one leaf word; every slot of a visited leaf costs an
//     exact f64 test, the most expensive code of the walk (at ~8 of 32 lanes).  With at most 2 primitives per leaf (1 from a million primitives up, where
//     the tree does not compete for L2 any more) the radix tree is within 3-8 % of the SAH tree on the 1 M / 10 M soups and 7.7 % on the dragon's extend time.
//   * TOPOLOGY — PLOC (parallel locally-ordered clustering, Meister & Bittner 2018: every cluster of the Morton-ordered cluster array looks for the neighbour
//     within PLOC_RADIUS positions whose union box with it has the smallest surface area, mutually nearest pairs merge, the array is compacted, until one
//     cluster is left).  Built, parity-green, and MEASURED NO BETTER here: equal on the dragon (extend 32.24 vs 32.25 ms), 20-25 % slower on the soups,
//     ~100 rounds and 2x the build time.  It stays selectable with RT_LBVH=ploc; the default is the Karras radix tree.
// Pipeline (all on the scene's device, default stream, n = primitives):
//   k_centroid_bounds   bounds of the box centroids (block reduce + 6 ordered-u64 atomics)
//   k_morton            63-bit Morton key of every centroid (21 bits per axis), value = primitive index
//   cub radix sort      (key, value) pairs — library plumbing, the only non-hand-written step
//   k_leaves            sorted leaf j: outward-rounded f32 box, and the leaf-ordered copy of the 80-byte slot records
//   k_radix_tree        Karras 2012: one thread per internal node finds its key range and split (ties broken by index)
//   k_fit               bottom-up box fit, second arrival at a node does the union
//   k_collapse          level by level from the root: a BVH4 node takes a binary node's two children and opens the inner child with
//                       the largest surface area until it has four; subtrees of <= 4 primitives become leaf words over a contiguous
//                       slot range (a radix-tree node always covers a contiguous range of the sorted order)
//   k_stack_need        bottom-up worst-case traversal stack, the same bound the host build checks
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cstring>
#include <string>
#include <cfloat>
#include <cstdio>
#include <vector>
#include "rt_internal.h"
#include "scene.h"

namespace rt {
namespace {

#define CKB(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { set_error(std::string("device BVH build: ") + cudaGetErrorString(e_)); cleanup(); return RT_ERR_CUDA; } } while (0)

__device__ __forceinline__ unsigned long long enc_f64(double x) {      // order-preserving map double -> u64
    long long b = __double_as_longlong(x);
    return b >= 0 ? (unsigned long long)b ^ 0x8000000000000000ull : ~(unsigned long long)b;
}
__device__ __forceinline__ double dec_f64(unsigned long long u) {
    long long b = (u & 0x8000000000000000ull) ? (long long)(u ^ 0x8000000000000000ull) : (long long)~u;
    return __longlong_as_double(b);
}
__global__ void k_init_bounds(unsigned long long* b) { if (threadIdx.x < 3) b[threadIdx.x] = ~0ull; else if (threadIdx.x < 6) b[threadIdx.x] = 0ull; }
__global__ void __launch_bounds__(256) k_centroid_bounds(const Box64* __restrict__ boxes, uint32_t n, unsigned long long* __restrict__ bounds) {
    double mn[3] = {1e308, 1e308, 1e308}, mx[3] = {-1e308, -1e308, -1e308};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int k = 0; k < 3; k++) { double c = 0.5 * boxes[i].mn[k] + 0.5 * boxes[i].mx[k]; mn[k] = fmin(mn[k], c); mx[k] = fmax(mx[k], c); }
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o > 0; o >>= 1) { mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o)); mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o)); }
        if ((threadIdx.x & 31) == 0) { atomicMin(bounds + k, enc_f64(mn[k])); atomicMax(bounds + 3 + k, enc_f64(mx[k])); }
    }
}
__device__ __forceinline__ unsigned long long spread21(unsigned long long x) {       // abc -> a00b00c
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull; x = (x | x << 16) & 0x1f0000ff0000ffull; x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull; x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
__global__ void __launch_bounds__(256) k_morton(const Box64* __restrict__ boxes, uint32_t n, const unsigned long long* __restrict__ bounds,
                                                unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals) {
    double lo[3], inv[3];
    for (int k = 0; k < 3; k++) { lo[k] = dec_f64(bounds[k]); double ext = dec_f64(bounds[3 + k]) - lo[k]; inv[k] = ext > 0. ? 2097152. / ext : 0.; }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned long long q[3];
        for (int k = 0; k < 3; k++) { double c = 0.5 * boxes[i].mn[k] + 0.5 * boxes[i].mx[k]; double f = (c - lo[k]) * inv[k]; q[k] = (unsigned long long)fmin(fmax(f, 0.), 2097151.); }
        keys[i] = spread21(q[0]) << 2 | spread21(q[1]) << 1 | spread21(q[2]);
        vals[i] = i;
    }
}

struct FB { float mn[3], mx[3]; };
// node ids: internal i in [0, n-1) ; leaf j -> (n-1) + j
__global__ void __launch_bounds__(256) k_leaves(const Box64* __restrict__ boxes, const uint32_t* __restrict__ order, uint32_t n, FB* __restrict__ fb) {
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const uint32_t p = order[j];
        FB b;
        // the same rule as the host build (bvh_build.cpp:leaf_box_f32): relative pad of 2^-22 of the coordinate magnitude, then outward rounding
        for (int k = 0; k < 3; k++) {
            const double lo = boxes[p].mn[k], hi = boxes[p].mx[k], pad = (fabs(lo) + fabs(hi)) * 2.4e-7;
            b.mn[k] = __double2float_rd(lo - pad); b.mx[k] = __double2float_ru(hi + pad);
        }
        fb[(size_t)(n - 1) + j] = b;
    }
}
// leaf slot `pos` of the finished tree holds sorted leaf leaf_at[pos], i.e. primitive order[leaf_at[pos]]
__global__ void __launch_bounds__(256) k_permute_slots(const DSlot* __restrict__ slots_in, const uint32_t* __restrict__ order, const uint32_t* __restrict__ leaf_at, uint32_t n,
                                                       DSlot* __restrict__ slots_out) {
    for (uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x; pos < n; pos += gridDim.x * blockDim.x) {
        const double2* s = reinterpret_cast<const double2*>(slots_in + order[leaf_at[pos]]); double2* d = reinterpret_cast<double2*>(slots_out + pos);
        d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = s[3]; d[4] = s[4];
    }
}
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, uint32_t n, int i, int j) {
    if (j < 0 || j >= (int)n) return -1;
    unsigned long long x = keys[i] ^ keys[j];
    return x ? __clzll((long long)x) : 64 + __clz(i ^ j);
}
__global__ void __launch_bounds__(256) k_radix_tree(const unsigned long long* __restrict__ keys, uint32_t n, int32_t* __restrict__ left, int32_t* __restrict__ right,
                                                    int32_t* __restrict__ parent, uint32_t* __restrict__ first, uint32_t* __restrict__ count, uint32_t* __restrict__ n_big, uint32_t leaf_max) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int)n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1) if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) { if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t; if (t == 1) break; }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int32_t lc = lo == gamma ? (int32_t)(n - 1) + gamma : gamma, rc = hi == gamma + 1 ? (int32_t)(n - 1) + gamma + 1 : gamma + 1;
    left[i] = lc; right[i] = rc; parent[lc] = i; parent[rc] = i;
    first[i] = (uint32_t)lo; count[i] = (uint32_t)(hi - lo + 1);
    if ((uint32_t)(hi - lo + 1) > leaf_max) atomicAdd(n_big, 1u);
    if (i == 0) parent[0] = -1;
}
__global__ void __launch_bounds__(256) k_fit(uint32_t n, const int32_t* __restrict__ left, const int32_t* __restrict__ right, const int32_t* __restrict__ parent,
                                             FB* fb, uint32_t* __restrict__ flags) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int32_t node = parent[(size_t)(n - 1) + j];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(flags + node, 1u) == 0u) return;       // first arrival: the sibling subtree is not done yet
        __threadfence();
        const volatile FB* a = fb + left[node]; const volatile FB* b = fb + right[node];
        FB u;
        for (int k = 0; k < 3; k++) { u.mn[k] = fminf(a->mn[k], b->mn[k]); u.mx[k] = fmaxf(a->mx[k], b->mx[k]); }
        for (int k = 0; k < 3; k++) { ((volatile FB*)fb)[node].mn[k] = u.mn[k]; ((volatile FB*)fb)[node].mx[k] = u.mx[k]; }
        node = parent[node];
    }
}
// ---- PLOC.  Node ids: internal i in [0, n-1), leaf j -> (n-1) + j; the cluster array C holds node ids in Morton order of their first leaf.
constexpr int PLOC_RADIUS = 16;
constexpr int PLOC_BLOCK = 256;
__device__ __forceinline__ float union_area(const FB& a, const FB& b) {
    const float dx = fmaxf(a.mx[0], b.mx[0]) - fminf(a.mn[0], b.mn[0]), dy = fmaxf(a.mx[1], b.mx[1]) - fminf(a.mn[1], b.mn[1]), dz = fmaxf(a.mx[2], b.mx[2]) - fminf(a.mn[2], b.mn[2]);
    return dx * dy + dy * dz + dz * dx;
}
__global__ void k_ploc_init(uint32_t n, uint32_t* __restrict__ C) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) C[i] = (n - 1) + i;
}
// nearest neighbour of every cluster within PLOC_RADIUS positions: smallest union area, ties to the smaller position (the cost is symmetric and
// computed by the same expression for (i, j) and (j, i), so mutual pairs are found consistently).  The boxes of the block's window go through shared memory.
__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_nn(const uint32_t* __restrict__ C, uint32_t m, const FB* __restrict__ fb, uint32_t* __restrict__ nn) {
    __shared__ FB tile[PLOC_BLOCK + 2 * PLOC_RADIUS];
    const long long base = (long long)blockIdx.x * PLOC_BLOCK - PLOC_RADIUS;
    for (int t = threadIdx.x; t < PLOC_BLOCK + 2 * PLOC_RADIUS; t += PLOC_BLOCK) {
        const long long g = base + t;
        if (g >= 0 && g < (long long)m) tile[t] = fb[C[g]];
    }
    __syncthreads();
    const long long i = (long long)blockIdx.x * PLOC_BLOCK + threadIdx.x;
    if (i >= (long long)m) return;
    const FB me = tile[threadIdx.x + PLOC_RADIUS];
    float best = FLT_MAX; long long bj = -1;
    for (int o = -PLOC_RADIUS; o <= PLOC_RADIUS; o++) {
        const long long j = i + o;
        if (o == 0 || j < 0 || j >= (long long)m) continue;
        const float a = union_area(me, tile[threadIdx.x + PLOC_RADIUS + o]);
        if (a < best) { best = a; bj = j; }              // ascending j: the first minimum is the one with the smaller position
    }
    nn[i] = (uint32_t)bj;
}
// mutual nearest neighbours merge into a new internal node (ids are handed out downwards from n-2, so that the last merge — the root — gets id 0);
// the lower position keeps the merged cluster, the upper one is dropped
__global__ void __launch_bounds__(256) k_ploc_merge(const uint32_t* __restrict__ C, uint32_t m, const uint32_t* __restrict__ nn, FB* __restrict__ fb, int32_t* __restrict__ left,
                                                    int32_t* __restrict__ right, uint32_t* __restrict__ count, uint32_t n, uint32_t* __restrict__ next_id, uint32_t* __restrict__ n_big,
                                                    uint32_t* __restrict__ C_tmp, uint32_t* __restrict__ keep, uint32_t leaf_max) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t j = nn[i];
    if (nn[j] != i) { C_tmp[i] = C[i]; keep[i] = 1u; return; }
    if (i > j) { keep[i] = 0u; return; }
    const uint32_t a = C[i], b = C[j], leaf0 = n - 1;
    const uint32_t id = atomicSub(next_id, 1u) - 1u;
    const FB ba = fb[a], bb = fb[b];
    FB u;
    for (int k = 0; k < 3; k++) { u.mn[k] = fminf(ba.mn[k], bb.mn[k]); u.mx[k] = fmaxf(ba.mx[k], bb.mx[k]); }
    fb[id] = u; left[id] = (int32_t)a; right[id] = (int32_t)b;
    const uint32_t cnt = (a >= leaf0 ? 1u : count[a]) + (b >= leaf0 ? 1u : count[b]);
    count[id] = cnt;
    if (cnt > leaf_max) atomicAdd(n_big, 1u);
    C_tmp[i] = id; keep[i] = 1u;
}
__global__ void __launch_bounds__(256) k_ploc_compact(const uint32_t* __restrict__ C_tmp, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos, uint32_t m, uint32_t* __restrict__ C_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m && keep[i]) C_out[pos[i]] = C_tmp[i];
}
// no mutual pair found in a round (cannot happen with a strict order on the costs; kept as a guard so that the loop always ends): pair up neighbours
__global__ void __launch_bounds__(256) k_ploc_force_nn(uint32_t m, uint32_t* __restrict__ nn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) nn[i] = (i ^ 1u) < m ? (i ^ 1u) : i - 1u;
}

__device__ __forceinline__ float half_area(const FB& b) {
    const float dx = b.mx[0] - b.mn[0], dy = b.mx[1] - b.mn[1], dz = b.mx[2] - b.mn[2];
    return dx * dy + dy * dz + dz * dx;
}
// q items: (binary internal node, BVH4 index, first leaf position of the node's subtree, -).  The leaves of a subtree occupy [first, first + count) of
// the final slot order: first[left] = first[node], first[right] = first[node] + count[left] — for the radix tree that reproduces the Morton order,
// for PLOC (whose clusters are not contiguous in it) it defines the order, written to leaf_at[] by the subtrees that become leaf words.
__global__ void __launch_bounds__(128) k_collapse(uint32_t n, const int32_t* __restrict__ left, const int32_t* __restrict__ right, const uint32_t* __restrict__ count,
                                                  const FB* __restrict__ fb, const uint4* __restrict__ q_in, uint32_t n_in, uint4* __restrict__ q_out, uint32_t* __restrict__ n_out,
                                                  uint32_t* __restrict__ n_nodes, DNode4* __restrict__ nodes, uint32_t* __restrict__ leaf_at, uint32_t leaf_max) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_in) return;
    const uint4 item = q_in[t];
    const int32_t leaf0 = (int32_t)(n - 1);
    auto cnt = [&](int32_t id) -> uint32_t { return id >= leaf0 ? 1u : count[id]; };
    int32_t kids[4] = {left[item.x], right[item.x], 0, 0}; int nk = 2;
    uint32_t kfirst[4] = {item.z, item.z + cnt(left[item.x]), 0u, 0u};
    while (nk < 4) {
        int pick = -1; float best = -1.f;
        for (int k = 0; k < nk; k++) if (cnt(kids[k]) > leaf_max) { const float a = half_area(fb[kids[k]]); if (a > best) { best = a; pick = k; } }
        if (pick < 0) break;
        const int32_t c = kids[pick];
        kids[pick] = left[c];                                   // keeps kfirst[pick]
        kids[nk] = right[c]; kfirst[nk] = kfirst[pick] + cnt(left[c]); nk++;
    }
    DNode4 d;
    for (int k = 0; k < 4; k++) {
        d.pad[k] = 0;
        if (k >= nk) { d.mnx[k] = d.mny[k] = d.mnz[k] = FLT_MAX; d.mxx[k] = d.mxy[k] = d.mxz[k] = -FLT_MAX; d.child[k] = EMPTY_CHILD; continue; }
        const int32_t c = kids[k]; const FB b = fb[c];
        d.mnx[k] = b.mn[0]; d.mxx[k] = b.mx[0]; d.mny[k] = b.mn[1]; d.mxy[k] = b.mx[1]; d.mnz[k] = b.mn[2]; d.mxz[k] = b.mx[2];
        const uint32_t c_n = cnt(c);
        if (c_n > leaf_max) {
            const uint32_t idx = atomicAdd(n_nodes, 1u);
            d.child[k] = (int32_t)idx;
            q_out[atomicAdd(n_out, 1u)] = make_uint4((uint32_t)c, idx, kfirst[k], 0u);
        } else {
            d.child[k] = (int32_t)~((kfirst[k] << 2) | (c_n - 1u));
            int32_t st[4]; int sp = 0; st[sp++] = c; uint32_t pos = kfirst[k];        // a subtree of <= 4 leaves: depth-first, left before right
            while (sp > 0) {
                const int32_t x = st[--sp];
                if (x >= leaf0) leaf_at[pos++] = (uint32_t)(x - leaf0);
                else { st[sp++] = right[x]; st[sp++] = left[x]; }
            }
        }
    }
    nodes[item.y] = d;
}
__global__ void __launch_bounds__(256) k_stack_need(const DNode4* __restrict__ nodes, uint32_t begin, uint32_t end, int32_t* __restrict__ need) {
    const uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    int kids = 0, deepest = 0;
    for (int k = 0; k < 4; k++) { const int32_t c = nodes[i].child[k]; if (c == EMPTY_CHILD) continue; kids++; if (c >= 0) deepest = max(deepest, need[c]); }
    need[i] = kids - 1 + deepest;
}

}  // namespace

// d_boxes: f64 primitive boxes, d_slots_in: slot records in primitive order.  On success *nodes_out / *slots_out are device
// allocations owned by the caller (n_nodes BVH4 nodes, n leaf-ordered slots).
rt_status build_bvh_device(const Box64* d_boxes, const DSlot* d_slots_in, uint32_t n, DNode4** nodes_out, uint32_t* n_nodes_out, DSlot** slots_out, int* max_stack_out, float* build_ms) {
    std::vector<void*> tmp;
    DNode4* nodes_big = nullptr; DSlot* slots = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    auto cleanup = [&]() { for (void* p : tmp) cudaFree(p); if (nodes_big) cudaFree(nodes_big); if (slots) cudaFree(slots); if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); };
    auto alloc = [&](size_t bytes) -> void* { void* p = nullptr; if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) return nullptr; tmp.push_back(p); return p; };
    if (n < 8) { set_error("device BVH build: needs at least 8 primitives"); return RT_ERR_INVALID; }
    if (n >= (1u << 29)) { set_error("device BVH build: too many primitives for the leaf encoding (2^29)"); return RT_ERR_INVALID; }   // leaf word ~((first << 2) | count-1) must stay negative
    CKB(cudaEventCreate(&e0)); CKB(cudaEventCreate(&e1));
    CKB(cudaEventRecord(e0, 0));
    const int grid = 148 * 8;
    unsigned long long* bounds = (unsigned long long*)alloc(6 * 8);
    unsigned long long *keys = (unsigned long long*)alloc((size_t)n * 8), *keys2 = (unsigned long long*)alloc((size_t)n * 8);
    uint32_t *vals = (uint32_t*)alloc((size_t)n * 4), *order = (uint32_t*)alloc((size_t)n * 4);
    const size_t n_ids = 2 * (size_t)n - 1;
    FB* fb = (FB*)alloc(n_ids * sizeof(FB));
    int32_t *left = (int32_t*)alloc((size_t)n * 4), *right = (int32_t*)alloc((size_t)n * 4), *parent = (int32_t*)alloc(n_ids * 4);
    uint32_t *first = (uint32_t*)alloc((size_t)n * 4), *count = (uint32_t*)alloc((size_t)n * 4), *flags = (uint32_t*)alloc((size_t)n * 4);
    uint32_t* ctr = (uint32_t*)alloc(4 * 4);               // [0] nodes with > 4 primitives, [1] BVH4 nodes, [2] next queue size, [3] PLOC: next internal node id
    if (!bounds || !keys || !keys2 || !vals || !order || !fb || !left || !right || !parent || !first || !count || !flags || !ctr) { set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
    if (cudaMalloc((void**)&slots, (size_t)n * sizeof(DSlot)) != cudaSuccess) { slots = nullptr; set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
    k_init_bounds<<<1, 32>>>(bounds);
    k_centroid_bounds<<<grid, 256>>>(d_boxes, n, bounds);
    k_morton<<<grid, 256>>>(d_boxes, n, bounds, keys, vals);
    size_t sort_bytes = 0;
    CKB(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys, keys2, vals, order, (int)n, 0, 63));
    void* sort_tmp = alloc(sort_bytes);
    if (!sort_tmp) { set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
    CKB(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, keys, keys2, vals, order, (int)n, 0, 63));
    k_leaves<<<grid, 256>>>(d_boxes, order, n, fb);
    CKB(cudaMemsetAsync(ctr, 0, 16));
    // largest subtree that becomes one leaf word (1..4 slots): see "Tree quality" at the top of this file.  RT_LBVH_LEAF overrides (A/B runs).
    uint32_t leaf_max = n >= 1000000u ? 1u : 2u;
    if (const char* e = getenv("RT_LBVH_LEAF")) { const int v = atoi(e); if (v >= 1 && v <= 4) leaf_max = (uint32_t)v; }
    const char* topo = getenv("RT_LBVH");
    const bool ploc = topo && !strcmp(topo, "ploc");
    int ploc_rounds = 0;
    if (!ploc) {
        CKB(cudaMemsetAsync(flags, 0, (size_t)n * 4));
        k_radix_tree<<<(n - 1 + 255) / 256, 256>>>(keys2, n, left, right, parent, first, count, ctr, leaf_max);
        k_fit<<<(n + 255) / 256, 256>>>(n, left, right, parent, fb, flags);
    } else {
        // scratch: the sort's key / value buffers are free again (keys: 8n bytes = two u32 arrays)
        uint32_t *Ca = reinterpret_cast<uint32_t*>(keys), *Cb = Ca + n, *nn = vals, *keep = flags, *pos = first;
        uint32_t* Ctmp = (uint32_t*)alloc((size_t)n * 4);
        size_t scan_bytes = 0;
        CKB(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, keep, pos, (int)n));
        void* scan_tmp = alloc(scan_bytes);
        if (!Ctmp || !scan_tmp) { set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
        const uint32_t start_id = n - 1;
        CKB(cudaMemcpy(ctr + 3, &start_id, 4, cudaMemcpyHostToDevice));            // ctr[3]: next internal id, handed out downwards
        k_ploc_init<<<grid, 256>>>(n, Ca);
        uint32_t m = n;
        bool force = false;
        while (m > 1) {
            const uint32_t g = (m + 255) / 256;
            if (force) k_ploc_force_nn<<<g, 256>>>(m, nn);
            else k_ploc_nn<<<(m + PLOC_BLOCK - 1) / PLOC_BLOCK, PLOC_BLOCK>>>(Ca, m, fb, nn);
            k_ploc_merge<<<g, 256>>>(Ca, m, nn, fb, left, right, count, n, ctr + 3, ctr, Ctmp, keep, leaf_max);
            CKB(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, keep, pos, (int)m));
            k_ploc_compact<<<g, 256>>>(Ctmp, keep, pos, m, Cb);
            uint32_t last[2];
            CKB(cudaMemcpy(&last[0], pos + (m - 1), 4, cudaMemcpyDeviceToHost));
            CKB(cudaMemcpy(&last[1], keep + (m - 1), 4, cudaMemcpyDeviceToHost));
            const uint32_t m2 = last[0] + last[1];
            if (m2 > m || m2 == 0) { set_error("device BVH build: PLOC compaction went wrong"); cleanup(); return RT_ERR_CUDA; }
            force = m2 == m;                                    // no mutual pair this round: pair up neighbours next time (guard, see k_ploc_force_nn)
            m = m2; std::swap(Ca, Cb);
            if (++ploc_rounds > 4096) { set_error("device BVH build: PLOC did not converge"); cleanup(); return RT_ERR_CUDA; }
        }
        uint32_t left_over = 0;
        CKB(cudaMemcpy(&left_over, ctr + 3, 4, cudaMemcpyDeviceToHost));
        if (left_over != 0) { set_error("device BVH build: PLOC made " + std::to_string(n - 1 - left_over) + " of " + std::to_string(n - 1) + " merges"); cleanup(); return RT_ERR_CUDA; }
    }
    if (getenv("RT_TRACE_BUILD")) fprintf(stderr, "bvh: device topology %s%s\n", ploc ? "PLOC, rounds " : "radix tree", ploc ? std::to_string(ploc_rounds).c_str() : "");
    uint32_t n_big = 0;
    CKB(cudaMemcpy(&n_big, ctr, 4, cudaMemcpyDeviceToHost));
    CKB(cudaGetLastError());
    // every BVH4 node stands for a distinct binary node with more than 4 primitives
    if (cudaMalloc((void**)&nodes_big, (size_t)n_big * sizeof(DNode4)) != cudaSuccess) { nodes_big = nullptr; set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
    uint4 *qa = (uint4*)alloc((size_t)n_big * 16), *qb = (uint4*)alloc((size_t)n_big * 16);
    int32_t* need = (int32_t*)alloc((size_t)n_big * 4);
    uint32_t* leaf_at = (uint32_t*)alloc((size_t)n * 4);
    if (!qa || !qb || !need || !leaf_at) { set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
    const uint4 root = make_uint4(0u, 0u, 0u, 0u); const uint32_t one = 1;
    CKB(cudaMemcpy(qa, &root, 16, cudaMemcpyHostToDevice));
    CKB(cudaMemcpy(ctr + 1, &one, 4, cudaMemcpyHostToDevice));
    std::vector<uint32_t> level_begin; uint32_t n_in = 1, made = 1;
    while (n_in > 0) {
        level_begin.push_back(made - n_in);
        CKB(cudaMemsetAsync(ctr + 2, 0, 4));
        k_collapse<<<(n_in + 127) / 128, 128>>>(n, left, right, count, fb, qa, n_in, qb, ctr + 2, ctr + 1, nodes_big, leaf_at, leaf_max);
        uint32_t c2[2];
        CKB(cudaMemcpy(c2, ctr + 1, 8, cudaMemcpyDeviceToHost));
        made = c2[0]; n_in = c2[1];
        std::swap(qa, qb);
        if (level_begin.size() > 4096) { set_error("device BVH build: tree too deep"); cleanup(); return RT_ERR_INVALID; }
    }
    level_begin.push_back(made);
    for (size_t l = level_begin.size() - 1; l-- > 0;) {
        const uint32_t b = level_begin[l], e = level_begin[l + 1];
        if (e > b) k_stack_need<<<(e - b + 255) / 256, 256>>>(nodes_big, b, e, need);
    }
    k_permute_slots<<<grid, 256>>>(d_slots_in, order, leaf_at, n, slots);
    int32_t need0 = 0;
    CKB(cudaMemcpy(&need0, need, 4, cudaMemcpyDeviceToHost));
    CKB(cudaGetLastError());
    DNode4* nodes = nullptr;
    if (cudaMalloc((void**)&nodes, (size_t)made * sizeof(DNode4)) != cudaSuccess) { set_error("device BVH build: out of device memory"); cleanup(); return RT_ERR_CUDA; }
    if (cudaMemcpy(nodes, nodes_big, (size_t)made * sizeof(DNode4), cudaMemcpyDeviceToDevice) != cudaSuccess) { cudaFree(nodes); set_error("device BVH build: copy failed"); cleanup(); return RT_ERR_CUDA; }
    cudaEventRecord(e1, 0); cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    if (build_ms) *build_ms = ms;
    *nodes_out = nodes; *n_nodes_out = made; *slots_out = slots; *max_stack_out = need0 + 1;
    slots = nullptr;                                         // ownership moves to the caller
    cleanup();
    return RT_OK;
}

}  // namespace rt
