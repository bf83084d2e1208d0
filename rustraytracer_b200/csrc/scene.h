// rt_scene: host tables + committed device image.
#pragma once
#include <atomic>
#include "rt_internal.h"

namespace rt { constexpr uint32_t N_BATCH_COUNTERS = 256; }
// the film: (r, g, b, n) running sums as 4 doubles per pixel on `device`; `owned` = allocated by rt_film_create (else caller memory)
struct rt_film { double* d = nullptr; uint32_t w = 0, h = 0; bool owned = false; int device = 0; };
struct rt_scene {
    // host tables (mirror of `Objects`, reference geometry.rs:13-21)
    std::vector<rt::HostMesh> meshes;
    rt::HVec<rt::HostPrim> prims;
    std::vector<double> shapes;              // 12 per shape: v[0..4], xform (-1 none), f64 leaf box mn[3], mx[3]
    std::vector<rt::DXform> xforms;
    std::vector<rt_texture> textures;
    std::vector<rt_material> materials;
    std::vector<rt_disney> disneys;          // rt_scene_add_disney
    std::vector<rt_light> lights;
    std::vector<rt::DImage> images;          // Hdr / Image texel blocks (rt_scene_add_image_*)
    std::vector<uint32_t> texels;
    std::vector<double> perlin;              // PERLIN_STRIDE doubles per rt_scene_add_perlin
    // committed
    bool committed = false;
    int device = 0;
    rt::DScene d{};                          // device pointers
    uint32_t* d_counter = nullptr;           // device work counters of the persistent batch kernels (ring of N_BATCH_COUNTERS)
    std::atomic<uint32_t> counter_next{0};   // next ring slot: concurrent rt_closest_device calls on different streams get different counters
    std::vector<void*> allocations;          // every cudaMalloc of this scene
    uint64_t bytes_nodes = 0, bytes_slots = 0, bytes_tables = 0;
    uint32_t build_mode = 0;                 // RT_BUILD_*
    rt_build_stats build_stats{};
    uint32_t bin_mask = 0;                   // shading bins (material type * 2 + is_triangle) that occur in the scene
    // wavefront state (render.cu)
    void* wavefront = nullptr;
    uint32_t wavefront_size = 1u << 24;   // paths in flight (≈ 480 B of queue state each: 8 GB); measured r01, dragon 128 spp: 4 Mi 1931, 8 Mi 2044, 16 Mi 2105, 32 Mi 2150 Mrays/s
    std::atomic<int> cancel{0};
    rt_counters counters{};
};

namespace rt {
// scene_host.cpp
Box64 prim_box(const rt_scene& s, const HostPrim& p);
double prim_area(const rt_scene& s, const HostPrim& p);
bool invert4(const double m[16], double out[16]);
// what commit hands to the device uploader (device.cu)
struct HostImage {
    std::vector<DNodeT> nodes; HVec<DSlot> slots; HVec<DPrim> prims; std::vector<DMesh> meshes;      // slots / prims: filled in full by the flatten threads (HVec: not zeroed first)
    std::vector<double> shapes;              // points / normals / uvs / indices are uploaded from rt_scene::meshes directly
    std::vector<DLight> lights;
    // Infinite lights: flat Distribution2D tables, concatenated; DLight pointers hold OFFSETS (in doubles) until upload
    std::vector<double> dist;
    std::vector<double> mat_alpha;           // 2 per material, NaN = not constant
    // RT_BUILD_DEVICE_LBVH: `nodes` and `slots` are empty.  The uploader computes the f64 boxes and the slot records of the triangles on
    // the device from the uploaded tables (device.cu:k_fill_tri_slots), scatters the few non-triangle ones prepared here (`other_*`,
    // slot kind 1), builds the tree where it will be walked (bvh_device.cu) and permutes the slots into leaf order there
    bool device_build = false; std::vector<uint32_t> other_idx; std::vector<Box64> other_boxes;
    // direct shapes (traverse.cuh:direct_shapes): the scene's parametric shapes when there are at most MAX_DIRECT of them next to triangles — kept out
    // of the tree (host build: not handed to the builder; device build: a dead slot with an inverted box) and tested once per ray before the walk
    std::vector<uint32_t> direct;
    size_t n_slots = 0;
    int max_stack = 0; double build_ms = 0.;
};
// bvh_device.cu
rt_status build_bvh_device(const Box64* d_boxes, const DSlot* d_slots_in, uint32_t n, DNode4** nodes_out, uint32_t* n_nodes_out, DSlot** slots_out, int* max_stack_out, float* build_ms);
rt_status build_host_image(const rt_scene& s, HostImage& img, bool device_build);
rt_status upload_scene(rt_scene& s, HostImage& img);         // device.cu
void free_scene_device(rt_scene& s);                          // device.cu
void free_wavefront(rt_scene& s);                             // render.cu
void render_prefilter_stats(unsigned long long out[4]);       // render.cu (RT_PREFILTER_CHECK builds)
}
