// Internal host/device layout of librtcore (not part of the ABI).
//
// Data layout in HBM:
//   nodes   : BVH2, 64 B/node, four 16-byte words (Aila–Laine child-box packing), f32 boxes rounded
//             outward from the reference's f64 primitive boxes;
//   slots   : leaf-ordered primitive slots, 80 B each = five 16-byte words: nine f64 vertex
//             coordinates (the reference's own f64 values — the exact test runs on them),
//             prim id, kind;
//   prims   : 16 B/prim record (type/flip/mesh flags, material, light, payload offset);
//   points / normals / uvs / tri_index : the meshes, concatenated, f64 (read by shade only);
//   shapes  : f64 payload of spheres and rects; xforms: forward + inverse 4x4 per transform.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <string>
#include <thread>
#include <utility>
#include <vector>
#include <sys/mman.h>
#include "../../include/rtcore.h"

namespace rt {

constexpr double SMALL = 0.001;          // consts.rs:32
constexpr double INF = 1e308;            // consts.rs:30
constexpr double PI = 3.14159265358979;  // consts.rs:31

// Host-side bulk arrays (mesh tables, primitive runs: gigabytes at 50 M triangles).  HVec is a std::vector whose resize() leaves trivially constructible
// elements uninitialised, so that the pages are first touched by the THREADS that fill them (par_for) instead of being zeroed by one thread first.
// Blocks of 4 MB and more are 2 MB-aligned and marked MADV_HUGEPAGE: a cold 10 M-triangle ingest is otherwise bound by ~400 K first-touch page faults
// (measured here: 1.06 s cold against 0.17 s into already-mapped memory, 8 threads), with transparent huge pages in `madvise` mode it is ~800.
template <class T> struct NoInitAlloc : std::allocator<T> {
    template <class U> struct rebind { typedef NoInitAlloc<U> other; };
    NoInitAlloc() = default;
    template <class U> NoInitAlloc(const NoInitAlloc<U>&) {}
    T* allocate(size_t n) {
        if (n > SIZE_MAX / sizeof(T)) throw std::bad_alloc();
        const size_t bytes = n * sizeof(T);
        if (bytes >= ((size_t)4 << 20)) {
            void* p = nullptr;
            if (posix_memalign(&p, (size_t)2 << 20, bytes) != 0 || !p) throw std::bad_alloc();
            madvise(p, bytes, MADV_HUGEPAGE);                  // advisory: failure leaves ordinary pages
            return (T*)p;
        }
        T* p = (T*)malloc(bytes ? bytes : 1);
        if (!p) throw std::bad_alloc();
        return p;
    }
    void deallocate(T* p, size_t) { free(p); }
    template <class U, class... A> void construct(U* p, A&&... a) {
        if constexpr (sizeof...(A) == 0) ::new ((void*)p) U; else ::new ((void*)p) U(std::forward<A>(a)...);
    }
};
template <class T> using HVec = std::vector<T, NoInitAlloc<T>>;
// host threads used by the ingest / build code: all of them, or RT_HOST_THREADS (measurement knob: ingest on 1 thread vs all)
inline size_t host_threads() {
    if (const char* e = getenv("RT_HOST_THREADS")) { const long v = atol(e); if (v > 0) return (size_t)v; }
    return std::max<size_t>(1, std::thread::hardware_concurrency());
}
// fn(begin, end) over [0, n) on the host threads (one chunk each); runs inline for small n
template <class F> inline void par_for(size_t n, size_t min_per_thread, F fn) {
    size_t t = std::max<size_t>(1, std::min<size_t>(host_threads(), n / std::max<size_t>(min_per_thread, 1)));
    if (t <= 1) { fn((size_t)0, n); return; }
    std::vector<std::thread> pool;
    const size_t chunk = (n + t - 1) / t;
    for (size_t k = 0; k < t; k++) { const size_t b = std::min(n, k * chunk), e = std::min(n, b + chunk); if (e > b) pool.emplace_back([=]() { fn(b, e); }); }
    for (auto& th : pool) th.join();
}
template <class T> inline void par_assign(HVec<T>& dst, const T* src, size_t n) {
    dst.resize(n);
    T* d = dst.data();
    par_for(n, (size_t)1 << 18, [=](size_t b, size_t e) { memcpy(d + b, src + b, (e - b) * sizeof(T)); });
}

struct HostMesh {
    HVec<double> p, n, uv;               // xyz / xyz / uv
    HVec<uint32_t> ind;
};

struct DNode { float v[12]; int32_t left, right, pad0, pad1; };   // 64 B
static_assert(sizeof(DNode) == 64, "node must be 64 bytes");
// v[0..3] = l.minx,l.maxx,l.miny,l.maxy ; v[4..7] = r.minx,r.maxx,r.miny,r.maxy ; v[8..11] = l.minz,l.maxz,r.minz,r.maxz

// BVH4 (default): the binned-SAH BVH2 collapsed to arity 4 on the host — half the dependent node fetches per ray, which is what
// bounds the latency-bound traversal (ncu r01: ~950 cycles per BVH2 step per lane, long-scoreboard 45 % of stalls).  128 B/node,
// eight 16-byte words: six planes of four child coordinates, four child words (>= 0 node, < 0 leaf, EMPTY_CHILD = unused slot
// whose box is inverted so that it can never be hit), one pad word.
struct DNode4 { float mnx[4], mxx[4], mny[4], mxy[4], mnz[4], mxz[4]; int32_t child[4]; int32_t pad[4]; };
static_assert(sizeof(DNode4) == 128, "BVH4 node must be 128 bytes");
constexpr int32_t EMPTY_CHILD = 0x7fffffff;
#ifndef RT_BVH_WIDTH
#define RT_BVH_WIDTH 4
#endif
#if RT_BVH_WIDTH == 4
typedef DNode4 DNodeT;
#else
typedef DNode DNodeT;
#endif

struct DSlot { double v[9]; uint32_t prim; uint32_t kind; };      // 80 B ; kind 0 = triangle, 1 = other
static_assert(sizeof(DSlot) == 80, "slot must be 80 bytes");

struct DPrim { uint32_t info, mat, light, a; };                   // 16 B
// info: bits 0-2 type, bit 3 flip_face, bit 4 mesh has normals, bit 5 mesh has uvs, bits 6.. mesh index
constexpr uint32_t PI_FLIP = 8u, PI_HAS_N = 16u, PI_HAS_UV = 32u;

struct DMesh { uint64_t p_base, n_base, uv_base; };               // element (vertex) offsets

struct DXform { double fwd[16], inv[16]; };

// Diffuse: prim/two_sided/color/area.  Infinite (light.rs:69-79): text_id, to_world/to_obj = xforms[xform] (RT_NONE =
// identity), and the Distribution2D of make_infinite_light (light.rs:608-638, distribution.rs:98-148) as flat device tables:
//   dist_f    [nu*nv]      the luminance*sin(theta) image; row v's `func` is the SLICE f[v .. v+nu) (distribution.rs:120)
//   dist_cdf  [nv*(nu+1)]  row v's normalised cdf
//   row_int   [nv]         row v's func_int  (= the marginal's `func`)
//   marg_cdf  [nv+1]       the marginal's normalised cdf ;  marg_int = the marginal's func_int
// eq_class: smallest index of a light field-equal to this one (`#[derive(PartialEq)] enum Light`, light.rs:28; used by integrator.rs:621)
struct DLight {
    uint32_t prim, two_sided, type, text_id; double color[3]; double area;
    uint32_t xform, nu, nv, eq_class; const double *dist_f, *dist_cdf, *row_int, *marg_cdf; double marg_int;
    // delta lights (light.rs:522-582): Point / Spot position (to_world x origin) or the Distant direction; the Spot cosines
    double pos[3]; double cos_width, cos_falloff;
};
constexpr double DISTANT_WORLD_RADIUS = 1e10;   // light.rs:579
// texel storage of Hdr / Image textures: one u32 per texel.  Hdr: the RGBE8 quadruple AFTER the re-quantisation of
// material.rs:578 (r | g<<8 | b<<16 | e<<24); Image: r | g<<8 | b<<16.
struct DImage { uint64_t offset; uint32_t w, h; };
constexpr double WORLD_RADIUS = 10000.0;   // light.rs:636

constexpr uint32_t MAX_DIRECT = 8;
// everything the kernels read; all pointers are device pointers
struct DScene {
    const DNodeT* nodes; const DSlot* slots; const DPrim* prims; const DMesh* meshes;
    const float* tri32;                       // 48 B per leaf slot: the f32 copy the prefilter reads (traverse.cuh:tri_prefilter): v0.xyz v1.x | v1.yz v2.xy | v2.z, prim, kind, pmax
    const double* points; const double* normals; const double* uvs; const uint32_t* tri_index;
    const double* shapes; const DXform* xforms;
    const rt_texture* textures; const rt_material* materials; const DLight* lights;
    const rt_disney* disneys;                 // Material::Disney records, indexed by the material's tex[0]
    const uint8_t* prim_bin;                  // shading bin of every primitive (material type * 2 + is_triangle), built at commit: one gather in ExtendJob::finish instead of prims -> materials
    const DImage* images; const uint32_t* texels;
    const double* perlin;                     // PERLIN_STRIDE doubles per Texture::Perlin: scale, perm_x[256], perm_y[256], perm_z[256], 256 gradient vectors
    const double* mat_alpha;                  // 2 per material: the microfacet alphas of Plastic / Glass / solid-textured Metal after the optional
                                              // roughness remap (microfacet.rs:442-446: an f64 log per shaded vertex otherwise); NaN = evaluate on the device
    uint32_t n_nodes, n_prims, n_lights, n_slots;
    uint32_t has_infinite, has_perlin;        // any Light::Infinite in `lights` (misses then emit, light.rs:499-513); any Texture::Perlin
    uint32_t has_disney, pad1_;               // any Material::Disney (f64 shade kernels only)
    uint32_t has_delta, n_direct;             // any Point / Spot / Distant light (f64 shade kernels only); direct shapes (below)
    // the scene's few parametric shapes when they are kept OUT of the tree and tested once per ray before the walk (traverse.cuh:direct_shapes)
    uint32_t direct[MAX_DIRECT];
    float direct_box[MAX_DIRECT][6];          // their leaf boxes rounded outward to f32 (min xyz, max xyz): a conservative cull in front of the exact f64 test
};

constexpr size_t PERLIN_STRIDE = 1 + 3 * 256 + 3 * 256;
struct Box64 { double mn[3], mx[3]; };

// host-side primitive record kept by rt_scene (compact; triangles of a mesh are a run)
struct HostPrim { uint8_t type, flip; uint32_t mat, light, mesh, ind, xform; uint32_t shape; };

void set_error(const std::string& msg);

// bvh_build.cpp: binned-SAH BVH2 over conservative f32 boxes.  Returns nodes (root = 0) and the
// leaf-ordered primitive permutation.  max 4 prims per leaf.
struct BuiltBvh { std::vector<DNodeT> nodes; std::vector<uint32_t> order; int max_stack = 0; };
constexpr int TRAVERSAL_STACK = 64;         // per-thread stack entries of the traversal kernels (traverse.cuh)
rt_status build_bvh(const std::vector<Box64>& boxes, BuiltBvh& out, int threads);

}  // namespace rt
