// Device half of the C ABI, part 1: scene upload and the ray-batch entry points
// (rt_intersect*, rt_closest_device, rt_occluded*).  sm_100a only; there is no CPU fallback —
// every entry point fails with RT_ERR_CUDA when no device is usable.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include "scene.h"
#include "traverse.cuh"

namespace rt {

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); return RT_ERR_CUDA; } \
    } while (0)

// Host -> device copy of one big table (the mesh arrays and the primitive table are gigabytes at 50 M triangles).  cudaMemcpy from pageable memory
// stages through the driver's own small pinned buffer on ONE host thread (measured: ~6 GB in ~0.8 s of the 1.07 s commit); here the host threads copy
// 32 MB chunks into two pinned buffers in turn (par_for) while the DMA of the previous chunk runs (cudaMemcpyAsync on a copy stream, one event per
// buffer), so the copy runs at the slower of host memcpy bandwidth and PCIe instead of their sum.  Small tables take the plain path.
static rt_status h2d(void* dst, const void* src, size_t bytes) {
    constexpr size_t CHUNK = (size_t)32 << 20;
    static std::mutex mu;
    static void* pinned[2] = {nullptr, nullptr};
    static bool pinned_failed = false;
    if (bytes < 2 * CHUNK || getenv("RT_H2D_PLAIN")) { CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice)); return RT_OK; }
    std::lock_guard<std::mutex> lock(mu);
    if (!pinned[0] && !pinned_failed) {
        if (cudaHostAlloc(&pinned[0], CHUNK, cudaHostAllocPortable) != cudaSuccess || cudaHostAlloc(&pinned[1], CHUNK, cudaHostAllocPortable) != cudaSuccess) {
            cudaGetLastError(); if (pinned[0]) cudaFreeHost(pinned[0]); pinned[0] = pinned[1] = nullptr; pinned_failed = true;
        }
    }
    if (pinned_failed) { CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice)); return RT_OK; }
    cudaStream_t st = nullptr; cudaEvent_t ev[2] = {nullptr, nullptr};
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaError_t e = cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming);
    size_t k = 0;
    for (size_t off = 0; off < bytes && e == cudaSuccess; off += CHUNK, k++) {
        const size_t n = std::min(CHUNK, bytes - off);
        const int b = (int)(k & 1);
        if (k >= 2) e = cudaEventSynchronize(ev[b]);                  // the DMA that last read this buffer is done
        if (e != cudaSuccess) break;
        char* d = (char*)pinned[b]; const char* sp = (const char*)src + off;
        par_for(n, (size_t)1 << 20, [=](size_t b0, size_t e0) { memcpy(d + b0, sp + b0, e0 - b0); });
        e = cudaMemcpyAsync((char*)dst + off, pinned[b], n, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaEventRecord(ev[b], st);
    }
    const cudaError_t e2 = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = e2;
    if (ev[0]) cudaEventDestroy(ev[0]);
    if (ev[1]) cudaEventDestroy(ev[1]);
    cudaStreamDestroy(st);
    if (e != cudaSuccess) { set_error(std::string("host -> device copy: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}

template <class T, class A> static rt_status up(rt_scene& s, const std::vector<T, A>& h, const T*& dptr, uint64_t& bytes) {
    dptr = nullptr;
    size_t n = h.size() * sizeof(T);
    void* p = nullptr;
    CK(cudaMalloc(&p, n ? n : 16));
    s.allocations.push_back(p);
    if (n) { if (rt_status st = h2d(p, h.data(), n)) return st; }
    dptr = (const T*)p; bytes += n;
    return RT_OK;
}

// The mesh tables (points, normals, uvs, indices of every mesh back to back; DMesh holds the element offsets): copied mesh by mesh
// straight into one device allocation — no concatenated host copy (4.8 GB of it at 50 M triangles).
template <class T, class F> static rt_status up_meshes(rt_scene& s, F field, const T*& dptr, uint64_t& bytes) {
    dptr = nullptr;
    size_t total = 0;
    for (const HostMesh& m : s.meshes) total += field(m).size();
    void* p = nullptr;
    CK(cudaMalloc(&p, total ? total * sizeof(T) : 16));
    s.allocations.push_back(p);
    size_t off = 0;
    for (const HostMesh& m : s.meshes) {
        const HVec<T>& v = field(m);
        if (!v.empty()) { if (rt_status st = h2d((T*)p + off, v.data(), v.size() * sizeof(T))) return st; }
        off += v.size();
    }
    dptr = (const T*)p; bytes += total * sizeof(T);
    return RT_OK;
}

// RT_BUILD_DEVICE_LBVH: the f64 box (scene_host.cpp:prim_box, triangle arm) and the 80-byte slot record (build_host_image:fill_slots)
// of every triangle, computed from the uploaded tables instead of on the host — at 50 M triangles the host versions and their 6.4 GB
// of uploads were most of the commit.  Same arithmetic (no FMA contraction in this library), so the records are the host's bit for bit.
__global__ void __launch_bounds__(256) k_fill_tri_slots(uint32_t n, const DPrim* __restrict__ prims, const DMesh* __restrict__ meshes, const double* __restrict__ points,
                                                        const double* __restrict__ uvs, const uint32_t* __restrict__ tri_index, Box64* __restrict__ boxes, DSlot* __restrict__ slots) {
    for (uint64_t i64 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i64 < n; i64 += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)i64;
        const DPrim p = prims[i];
        if ((p.info & 7u) != RT_PRIM_TRIANGLE) continue;
        const DMesh m = meshes[p.info >> 6];
        const uint32_t id[3] = {tri_index[p.a], tri_index[p.a + 1], tri_index[p.a + 2]};
        DSlot sl;
        for (int v = 0; v < 3; v++) for (int k = 0; k < 3; k++) sl.v[3 * v + k] = points[3 * (m.p_base + id[v]) + k];
        sl.prim = i; sl.kind = 0;
        Box64 b;
        for (int k = 0; k < 3; k++) { b.mn[k] = fmin(sl.v[k], fmin(sl.v[3 + k], sl.v[6 + k])); b.mx[k] = fmax(sl.v[k], fmax(sl.v[3 + k], sl.v[6 + k])); }
        if (p.info & PI_HAS_UV) {
            // hittable.rs:366-378: uv determinant below 1e-7 AND an exactly zero geometric normal: the triangle always returns None (kind 2)
            const double* u0 = uvs + 2 * (m.uv_base + id[0]); const double* u1 = uvs + 2 * (m.uv_base + id[1]); const double* u2 = uvs + 2 * (m.uv_base + id[2]);
            const double d02x = u0[0] - u2[0], d02y = u0[1] - u2[1], d12x = u1[0] - u2[0], d12y = u1[1] - u2[1];
            const double determinant = d02x * d12y - d02y * d12x;
            if (fabs(determinant) < SMALL / 10000.) {
                const double a[3] = {sl.v[6] - sl.v[0], sl.v[7] - sl.v[1], sl.v[8] - sl.v[2]}, e[3] = {sl.v[3] - sl.v[0], sl.v[4] - sl.v[1], sl.v[5] - sl.v[2]};
                const double c[3] = {a[1] * e[2] - a[2] * e[1], a[2] * e[0] - a[0] * e[2], a[0] * e[1] - a[1] * e[0]};
                if (c[0] * c[0] + c[1] * c[1] + c[2] * c[2] == 0.) sl.kind = 2;
            }
        }
        boxes[i] = b; slots[i] = sl;
    }
}
// the non-triangle primitives' boxes come from the host (transforms, padding: scene_host.cpp:prim_box); their slot is {prim, kind 1}
__global__ void k_scatter_other_slots(uint32_t n, const uint32_t* __restrict__ idx, const Box64* __restrict__ ob, Box64* __restrict__ boxes, DSlot* __restrict__ slots) {
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const uint32_t i = idx[j];
        DSlot sl; for (int k = 0; k < 9; k++) sl.v[k] = 0.;
        sl.prim = i; sl.kind = ob[j].mn[0] > ob[j].mx[0] ? 2u : 1u;      // inverted box = a direct shape: dead slot (scene_host.cpp)
        boxes[i] = ob[j]; slots[i] = sl;
    }
}

// The prefilter's 48-byte f32 record of every leaf slot (traverse.cuh:tri_prefilter), derived on the device from the leaf-ordered f64 slots of
// either builder: vertices rounded to nearest f32, the primitive id and slot kind, and pmax = max |coordinate| of the f64 vertices rounded UP
// (the error bound's |p| <= pmax must hold for the f64 values).
__global__ void __launch_bounds__(256) k_make_tri32(uint32_t n, const DSlot* __restrict__ slots, float4* __restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const double2* q = reinterpret_cast<const double2*>(slots + i);
        const double2 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3], q4 = q[4];
        const uint2 tag = *reinterpret_cast<const uint2*>(&q4.y);
        const double v[9] = {q0.x, q0.y, q1.x, q1.y, q2.x, q2.y, q3.x, q3.y, q4.x};
        double m = 0.;
        for (int k = 0; k < 9; k++) m = fmax(m, fabs(v[k]));
        out[3 * i + 0] = make_float4((float)v[0], (float)v[1], (float)v[2], (float)v[3]);
        out[3 * i + 1] = make_float4((float)v[4], (float)v[5], (float)v[6], (float)v[7]);
        out[3 * i + 2] = make_float4((float)v[8], __uint_as_float(tag.x), __uint_as_float(tag.y), __double2float_ru(m));
    }
}

rt_status upload_scene(rt_scene& s, HostImage& img) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { set_error("no CUDA device: librtcore has no CPU path"); return RT_ERR_CUDA; }
    CK(cudaGetDevice(&s.device));
    DScene d{}; uint64_t b = 0;
    s.bytes_nodes = s.bytes_slots = s.bytes_tables = 0;
    rt_status st;
    uint32_t n_nodes = (uint32_t)img.nodes.size();
    if ((st = up(s, img.prims, d.prims, b))) return st;
    if ((st = up(s, img.meshes, d.meshes, b))) return st;
    if ((st = up_meshes<double>(s, [](const HostMesh& m) -> const HVec<double>& { return m.p; }, d.points, b))) return st;
    if ((st = up_meshes<double>(s, [](const HostMesh& m) -> const HVec<double>& { return m.n; }, d.normals, b))) return st;
    if ((st = up_meshes<double>(s, [](const HostMesh& m) -> const HVec<double>& { return m.uv; }, d.uvs, b))) return st;
    if ((st = up_meshes<uint32_t>(s, [](const HostMesh& m) -> const HVec<uint32_t>& { return m.ind; }, d.tri_index, b))) return st;
    if (img.device_build) {
#if RT_BVH_WIDTH == 4
        // RT_BUILD_DEVICE_LBVH: boxes and primitive-ordered slots go up as temporaries, the tree is built where it will be walked
        void *tb = nullptr, *ts = nullptr;
        const uint32_t n = (uint32_t)img.n_slots;
        CK(cudaMalloc(&tb, (size_t)n * sizeof(Box64)));
        if (cudaMalloc(&ts, (size_t)n * sizeof(DSlot)) != cudaSuccess) { cudaFree(tb); set_error("upload: out of device memory"); return RT_ERR_CUDA; }
        k_fill_tri_slots<<<148 * 8, 256>>>(n, d.prims, d.meshes, d.points, d.uvs, d.tri_index, (Box64*)tb, (DSlot*)ts);
        if (!img.other_idx.empty()) {
            const uint32_t no = (uint32_t)img.other_idx.size();
            void *oi = nullptr, *ob = nullptr;
            if (cudaMalloc(&oi, (size_t)no * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&ob, (size_t)no * sizeof(Box64)) != cudaSuccess) {
                cudaFree(oi); cudaFree(tb); cudaFree(ts); set_error("upload: out of device memory"); return RT_ERR_CUDA;
            }
            cudaMemcpy(oi, img.other_idx.data(), (size_t)no * sizeof(uint32_t), cudaMemcpyHostToDevice);
            cudaMemcpy(ob, img.other_boxes.data(), (size_t)no * sizeof(Box64), cudaMemcpyHostToDevice);
            k_scatter_other_slots<<<(no + 255) / 256, 256>>>(no, (const uint32_t*)oi, (const Box64*)ob, (Box64*)tb, (DSlot*)ts);
            cudaDeviceSynchronize(); cudaFree(oi); cudaFree(ob);
        }
        if (cudaDeviceSynchronize() != cudaSuccess || cudaGetLastError() != cudaSuccess) { cudaFree(tb); cudaFree(ts); set_error("upload: slot kernels failed"); return RT_ERR_CUDA; }
        DNode4* nodes = nullptr; DSlot* slots = nullptr; float ms = 0.f;
        st = build_bvh_device((const Box64*)tb, (const DSlot*)ts, n, &nodes, &n_nodes, &slots, &img.max_stack, &ms);
        cudaFree(tb); cudaFree(ts);
        if (st != RT_OK) return st;
        s.allocations.push_back(nodes); s.allocations.push_back(slots);
        d.nodes = nodes; d.slots = slots; img.build_ms = ms;
        s.bytes_nodes = (uint64_t)n_nodes * sizeof(DNode4); s.bytes_slots = (uint64_t)n * sizeof(DSlot);
        if (getenv("RT_TRACE_BUILD")) fprintf(stderr, "bvh: device LBVH, %u nodes, %u slots, worst-case stack %d, %.2f ms\n", n_nodes, n, img.max_stack, ms);
#else
        set_error("device BVH build needs RT_BVH_WIDTH == 4"); return RT_ERR_INVALID;
#endif
    } else {
        if ((st = up(s, img.nodes, d.nodes, s.bytes_nodes))) return st;
        if ((st = up(s, img.slots, d.slots, s.bytes_slots))) return st;
    }
#if RT_PREFILTER
    {   // f32 prefilter records, leaf order
        void* tp = nullptr;
        CK(cudaMalloc(&tp, std::max<size_t>(img.n_slots, 1) * 3 * sizeof(float4)));
        s.allocations.push_back(tp);
        if (img.n_slots) {
            k_make_tri32<<<148 * 8, 256>>>((uint32_t)img.n_slots, d.slots, (float4*)tp);
            CK(cudaGetLastError()); CK(cudaDeviceSynchronize());
        }
        d.tri32 = (const float*)tp; s.bytes_slots += (uint64_t)img.n_slots * 3 * sizeof(float4);
    }
#endif
    if ((st = up(s, img.shapes, d.shapes, b))) return st;
    if ((st = up(s, s.xforms, d.xforms, b))) return st;
    if ((st = up(s, s.textures, d.textures, b))) return st;
    if ((st = up(s, s.materials, d.materials, b))) return st;
    if ((st = up(s, s.disneys, d.disneys, b))) return st;
    {   // per-primitive shading bin
        HVec<uint8_t> bins; bins.resize(img.prims.size());
        uint8_t* bp = bins.data(); const DPrim* pp = img.prims.data(); const rt_material* mp = s.materials.data();
        par_for(bins.size(), (size_t)1 << 18, [=](size_t b0, size_t e0) {
            for (size_t i = b0; i < e0; i++) bp[i] = (uint8_t)(mp[pp[i].mat].type * 2u + ((pp[i].info & 7u) == RT_PRIM_TRIANGLE ? 1u : 0u));
        });
        if ((st = up(s, bins, d.prim_bin, b))) return st;
    }
    d.has_disney = 0;
    for (auto& m : s.materials) if (m.type == RT_MAT_DISNEY) d.has_disney = 1;
    if ((st = up(s, img.mat_alpha, d.mat_alpha, b))) return st;
    if ((st = up(s, s.images, d.images, b))) return st;
    if ((st = up(s, s.texels, d.texels, b))) return st;
    if ((st = up(s, s.perlin, d.perlin, b))) return st;
    d.has_perlin = 0;
    for (auto& t : s.textures) if (t.type == RT_TEX_PERLIN) d.has_perlin = 1;
    // Infinite lights: upload the flat Distribution2D tables and turn the offsets kept in DLight into device pointers
    const double* d_dist = nullptr;
    if ((st = up(s, img.dist, d_dist, b))) return st;
    std::vector<DLight> lights = img.lights;
    d.has_delta = 0;
    for (auto& l : lights) {
        if (l.type == RT_LIGHT_POINT || l.type == RT_LIGHT_SPOT || l.type == RT_LIGHT_DISTANT) d.has_delta = 1;
        if (l.type != RT_LIGHT_INFINITE) continue;
        d.has_infinite = 1;
        l.dist_f = d_dist + ((uintptr_t)l.dist_f - 1); l.dist_cdf = d_dist + ((uintptr_t)l.dist_cdf - 1);
        l.row_int = d_dist + ((uintptr_t)l.row_int - 1); l.marg_cdf = d_dist + ((uintptr_t)l.marg_cdf - 1);
    }
    if ((st = up(s, lights, d.lights, b))) return st;
    s.bytes_tables = b;
    // work counters of the persistent batch kernels: a ring, one per call in flight, so that calls on different streams never share one
    { void* cp = nullptr; CK(cudaMalloc(&cp, N_BATCH_COUNTERS * sizeof(uint32_t))); s.allocations.push_back(cp); s.d_counter = (uint32_t*)cp; }
    d.n_direct = (uint32_t)std::min<size_t>(img.direct.size(), MAX_DIRECT);
    for (uint32_t k = 0; k < d.n_direct; k++) {
        d.direct[k] = img.direct[k];
        // the shape's f64 leaf box (shapes[.. + 6 .. 11]) padded and rounded outward like every box of the tree (bvh_build.cpp)
        const double* bx = &img.shapes[img.prims[img.direct[k]].a + 6];
        for (int a = 0; a < 3; a++) {
            const double lo = bx[a], hi = bx[3 + a], pad = (std::fabs(lo) + std::fabs(hi)) * 2.4e-7;
            d.direct_box[k][a] = std::nextafterf((float)(lo - pad), -INFINITY); d.direct_box[k][3 + a] = std::nextafterf((float)(hi + pad), INFINITY);
        }
    }
    d.n_nodes = n_nodes; d.n_prims = (uint32_t)img.prims.size();
    d.n_lights = (uint32_t)img.lights.size(); d.n_slots = (uint32_t)img.n_slots;
    s.d = d;
    return RT_OK;
}

void free_scene_device(rt_scene& s) {
    for (void* p : s.allocations) cudaFree(p);
    s.allocations.clear(); s.d = DScene{}; s.d_counter = nullptr; s.committed = false;
}

// ---------------------------------------------------------------- kernels
__global__ void __launch_bounds__(128) k_intersect(const __grid_constant__ DScene sc, uint64_t n, const rt_ray* __restrict__ rays, double tmin, double tmax, rt_hit* __restrict__ hits) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const double2* rp = reinterpret_cast<const double2*>(rays + i);
        double2 a = __ldg(rp), b = __ldg(rp + 1), c = __ldg(rp + 2);
        D3 o = mk3(a.x, a.y, b.x), d = mk3(b.y, c.x, c.y);
        Hit h = trace_closest(sc, o, d, tmin, tmax);
        rt_hit out; memset(&out, 0, sizeof(out));
        out.prim_index = RT_NONE;
        if (h.prim != RT_NONE) {
            Rec r;
            if (prim_record(sc, h.prim, o, d, tmin, tmax, true, r)) {
                out.t = r.t; out.p[0] = r.p.x; out.p[1] = r.p.y; out.p[2] = r.p.z;
                out.n[0] = r.n.x; out.n[1] = r.n.y; out.n[2] = r.n.z;
                out.ns[0] = r.sn.x; out.ns[1] = r.sn.y; out.ns[2] = r.sn.z;
                out.uv[0] = r.uv.x; out.uv[1] = r.uv.y; out.prim_index = h.prim; out.front = r.front ? 1u : 0u;
            }
        }
        hits[i] = out;
    }
}

__global__ void __launch_bounds__(128) k_closest(const __grid_constant__ DScene sc, uint64_t n, const rt_ray* __restrict__ rays, double tmin, double tmax,
                                                 uint32_t* __restrict__ prim, double* __restrict__ t) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const double2* rp = reinterpret_cast<const double2*>(rays + i);
        double2 a = __ldg(rp), b = __ldg(rp + 1), c = __ldg(rp + 2);
        Hit h = trace_closest(sc, mk3(a.x, a.y, b.x), mk3(b.y, c.x, c.y), tmin, tmax);
        prim[i] = h.prim; t[i] = h.t;
    }
}

// The ray-cast sweep entry (rt_closest_device) through the persistent scheduler of traverse.cuh: lanes whose ray has finished
// fetch the next one from a global counter instead of idling until the slowest ray of the warp is done — the soup's
// incoherent rays differ by an order of magnitude in traversal length.
struct ClosestJob {
    static constexpr bool shared_leaf_ray = false;      // traverse.cuh:trace_jobs
    const DScene* sc; const rt_ray* rays; double tmin_, tmax_; uint32_t* prim; double* t;
    __device__ __forceinline__ double tmin(const TState&) const { return tmin_; }
    __device__ __forceinline__ double tmax(const TState&) const { return tmax_; }
    __device__ __forceinline__ void ray(uint32_t i, D3& o, D3& d) const {
        const double2* rp = reinterpret_cast<const double2*>(rays + i);
        const double2 a = __ldg(rp), b = __ldg(rp + 1), c = __ldg(rp + 2);
        o = mk3(a.x, a.y, b.x); d = mk3(b.y, c.x, c.y);
    }
    __device__ __forceinline__ bool load(uint32_t i, TState& st, int32_t* stack) {
        D3 o, d; ray(i, o, d);
        tstate_init(*sc, st, stack, o, d, tmin_, tmax_, false, 0., 0u);
        return true;
    }
    __device__ __forceinline__ void finish(uint32_t i, const TState& st) { prim[i] = st.best.prim; t[i] = st.best.t; }
};
__global__ void __launch_bounds__(128, 5) k_closest_persistent(const __grid_constant__ DScene sc, ClosestJob job, uint32_t n, uint32_t* counter) {
    job.sc = &sc;
    trace_jobs(sc, job, n, counter);
}

__global__ void __launch_bounds__(128) k_occluded(const __grid_constant__ DScene sc, uint64_t n, const rt_shadow_query* __restrict__ q, uint8_t* __restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        rt_shadow_query s = q[i];
        out[i] = unoccluded(sc, mk3(s.p0[0], s.p0[1], s.p0[2]), mk3(s.p1[0], s.p1[1], s.p1[2]), s.p1_prim_index, s.infinite_light != 0) ? 1 : 0;
    }
}

static int grid_for(uint64_t n, int block) {
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint64_t want = (n + block - 1) / block;
    uint64_t cap = (uint64_t)sms * 64;                 // a multiple of the SM count; grid-stride covers the rest
    return (int)std::max<uint64_t>(1, std::min(want, cap));
}

static rt_status check_scene(rt_scene* s, const char* who) {
    if (!s) { set_error(std::string(who) + ": null scene"); return RT_ERR_INVALID; }
    if (!s->committed) { set_error(std::string(who) + ": scene not committed"); return RT_ERR_STATE; }
    CK(cudaSetDevice(s->device));
    return RT_OK;
}

}  // namespace rt

using namespace rt;

extern "C" {

int32_t rt_device_count(void) { int c = 0; if (cudaGetDeviceCount(&c) != cudaSuccess) return 0; return c; }
rt_status rt_set_device(int32_t device) { CK(cudaSetDevice(device)); return RT_OK; }

// RT_PREFILTER_CHECK builds: {triangle slot tests, skipped by the f32 prefilter, exact hits, violations} summed over every traversal kernel
// since the library was loaded; zeros otherwise.  A violation = the prefilter skipped a slot whose exact hit the traversal would have used.
rt_status rt_debug_prefilter_stats(uint64_t out[4]) {
    if (!out) { set_error("rt_debug_prefilter_stats: null argument"); return RT_ERR_INVALID; }
    unsigned long long acc[4] = {0, 0, 0, 0};
    cudaDeviceSynchronize();
    prefilter_stats_add(acc); render_prefilter_stats(acc);
    for (int k = 0; k < 4; k++) out[k] = acc[k];
    return RT_OK;
}

rt_status rt_intersect_device(rt_scene* s, uint64_t n, const rt_ray* d_rays, double tmin, double tmax, rt_hit* d_hits, void* stream) {
    rt_status st = check_scene(s, "rt_intersect_device"); if (st) return st;
    if (n == 0) return RT_OK;
    if (!d_rays || !d_hits) { set_error("rt_intersect_device: null buffer"); return RT_ERR_INVALID; }
    k_intersect<<<grid_for(n, 128), 128, 0, (cudaStream_t)stream>>>(s->d, n, d_rays, tmin, tmax, d_hits);
    CK(cudaGetLastError());
    return RT_OK;
}
rt_status rt_closest_device(rt_scene* s, uint64_t n, const rt_ray* d_rays, double tmin, double tmax, uint32_t* d_prim, double* d_t, void* stream) {
    rt_status st = check_scene(s, "rt_closest_device"); if (st) return st;
    if (n == 0) return RT_OK;
    if (!d_rays || !d_prim || !d_t) { set_error("rt_closest_device: null buffer"); return RT_ERR_INVALID; }
    if (!s->d_counter) { k_closest<<<grid_for(n, 128), 128, 0, (cudaStream_t)stream>>>(s->d, n, d_rays, tmin, tmax, d_prim, d_t); CK(cudaGetLastError()); return RT_OK; }
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
    const uint64_t CHUNK = 1ull << 31;
    for (uint64_t base = 0; base < n; base += CHUNK) {
        uint32_t m = (uint32_t)std::min<uint64_t>(CHUNK, n - base);
        uint32_t* counter = s->d_counter + (s->counter_next.fetch_add(1u) % N_BATCH_COUNTERS);
        CK(cudaMemsetAsync(counter, 0, sizeof(uint32_t), (cudaStream_t)stream));
        ClosestJob job{nullptr, d_rays + base, tmin, tmax, d_prim + base, d_t + base};
        k_closest_persistent<<<sms * 5, 128, 0, (cudaStream_t)stream>>>(s->d, job, m, counter);
    }
    CK(cudaGetLastError());
    return RT_OK;
}
rt_status rt_occluded_device(rt_scene* s, uint64_t n, const rt_shadow_query* d_q, uint8_t* d_out, void* stream) {
    rt_status st = check_scene(s, "rt_occluded_device"); if (st) return st;
    if (n == 0) return RT_OK;
    if (!d_q || !d_out) { set_error("rt_occluded_device: null buffer"); return RT_ERR_INVALID; }
    k_occluded<<<grid_for(n, 128), 128, 0, (cudaStream_t)stream>>>(s->d, n, d_q, d_out);
    CK(cudaGetLastError());
    return RT_OK;
}

rt_status rt_intersect_range(rt_scene* s, uint64_t n, const rt_ray* rays, double tmin, double tmax, rt_hit* hits) {
    rt_status st = check_scene(s, "rt_intersect"); if (st) return st;
    if (n == 0) return RT_OK;
    if (!rays || !hits) { set_error("rt_intersect: null buffer"); return RT_ERR_INVALID; }
    rt_ray* d_r = nullptr; rt_hit* d_h = nullptr;
    CK(cudaMalloc(&d_r, n * sizeof(rt_ray)));
    if (cudaMalloc(&d_h, n * sizeof(rt_hit)) != cudaSuccess) { cudaFree(d_r); set_error("rt_intersect: out of device memory"); return RT_ERR_CUDA; }
    cudaError_t e = cudaMemcpy(d_r, rays, n * sizeof(rt_ray), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) { st = rt_intersect_device(s, n, d_r, tmin, tmax, d_h, nullptr); if (st == RT_OK) e = cudaMemcpy(hits, d_h, n * sizeof(rt_hit), cudaMemcpyDeviceToHost); }
    cudaFree(d_r); cudaFree(d_h);
    if (st) return st;
    if (e != cudaSuccess) { set_error(std::string("rt_intersect: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}
// geometry::get_intersection (geometry.rs:238-240): (SMALL, INFINITY)
rt_status rt_intersect(rt_scene* s, uint64_t n, const rt_ray* rays, rt_hit* hits) { return rt_intersect_range(s, n, rays, SMALL, INF, hits); }

rt_status rt_occluded(rt_scene* s, uint64_t n, const rt_shadow_query* q, uint8_t* out) {
    rt_status st = check_scene(s, "rt_occluded"); if (st) return st;
    if (n == 0) return RT_OK;
    if (!q || !out) { set_error("rt_occluded: null buffer"); return RT_ERR_INVALID; }
    rt_shadow_query* d_q = nullptr; uint8_t* d_o = nullptr;
    CK(cudaMalloc(&d_q, n * sizeof(rt_shadow_query)));
    if (cudaMalloc(&d_o, n) != cudaSuccess) { cudaFree(d_q); set_error("rt_occluded: out of device memory"); return RT_ERR_CUDA; }
    cudaError_t e = cudaMemcpy(d_q, q, n * sizeof(rt_shadow_query), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) { st = rt_occluded_device(s, n, d_q, d_o, nullptr); if (st == RT_OK) e = cudaMemcpy(out, d_o, n, cudaMemcpyDeviceToHost); }
    cudaFree(d_q); cudaFree(d_o);
    if (st) return st;
    if (e != cudaSuccess) { set_error(std::string("rt_occluded: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}

}  // extern "C"
