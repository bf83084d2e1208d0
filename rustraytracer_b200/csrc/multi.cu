// Device half of the C ABI, part 3: several GPUs of one box.
//
// The reference spreads a render over NUM_THREADS = 6 CPU threads pulling 16x16 tiles from a queue and merging into one Mutex-guarded film
// (render::tile_multithread, render.rs:13-159).  Here the scene is REPLICATED on every GPU, the work is split by SAMPLE RANGE — GPU g renders
// samples [b + g S / G, b + (g+1) S / G) of every pixel; the RNG is keyed by (pixel, global sample index), so the union is bit-for-bit the
// single-GPU render of S samples up to the order of the f64 additions — and the (r,g,b,n) films are summed with ONE ncclAllReduce(f64, sum) over
// NVLink / NVSwitch.  There is no per-ray exchange, hence no fused compute+collective kernel: the reduce moves 32 MiB (1024^2) or 128 MiB (2048^2)
// once per render next to a render of 0.1 - 100 s.
// Two ways to form the communicator, both behind the ABI:
//   * one process, one host thread per GPU: rt_comm_create_local(devices) = ncclCommInitAll, rt_render_multi(...) runs the threads;
//   * one process per GPU (torchrun, MPI): rank 0 calls rt_comm_unique_id, every rank rt_comm_create_rank(id, world, rank) = ncclCommInitRank,
//     then rt_render on its own sample range and rt_film_allreduce.
// NCCL is bound at run time (dlopen of libnccl.so.2, the copy already in the process if there is one — e.g. torch's): librtcore.so has no
// link-time dependency on it and still loads where NCCL is absent; the multi-GPU entry points then fail with RT_ERR_STATE.
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#include "scene.h"
#include "film.cuh"

namespace rt {

// rt_film_reduce_resolve_rgb8: the films of up to 16 GPUs summed and resolved by one kernel, launched on every GPU that holds a film, each over its
// own slice of the pixels (the shape of a reduce-scatter + gather, without NCCL).  p[k] are the N accumulators — one local, the others read in place over
// NVLink / NVSwitch with ordinary loads (peer access): 32 bytes per pixel and film, each read once, fully coalesced (two 16-byte loads per lane) —
// and the 3 bytes per pixel are stored straight into the root GPU's image buffer (peer stores).  Every GPU pulls (N-1)/N of ONE film's bytes, so no
// link carries more than any other (the first version had the root pull N-1 whole films: 4.2 ms against NCCL's 3.4 at 8 GPUs, 2048^2;
// profiles/reduce_resolve_r02.json).  The sum runs in the order given, so its bits do not depend on how a ring or a tree associates the additions.
constexpr int REDUCE_MAX_FILMS = 16;
struct FilmPtrs { const double2* p[REDUCE_MAX_FILMS]; };
__global__ void __launch_bounds__(256) k_reduce_resolve(const __grid_constant__ FilmPtrs f, int n_films, uint64_t begin, uint64_t end, uint8_t* __restrict__ rgb) {
    for (uint64_t i = begin + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < end; i += (uint64_t)gridDim.x * blockDim.x) {
        double2 rg = __ldcs(f.p[0] + 2 * i), bn = __ldcs(f.p[0] + 2 * i + 1);                  // streamed: every byte is used once
        for (int k = 1; k < n_films; k++) {
            const double2 a = __ldcs(f.p[k] + 2 * i), b = __ldcs(f.p[k] + 2 * i + 1);
            rg.x += a.x; rg.y += a.y; bn.x += b.x; bn.y += b.y;
        }
        resolve_pixel(rg.x, rg.y, bn.x, bn.y, rgb + 3 * i);
    }
}

namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false; std::string why;
};
const NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, []() {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);       // the copy already loaded (torch bundles one)
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { api.why = std::string("NCCL not found: ") + dlerror(); return; }
        bool all = true;
        auto bind = [&](auto& fn, const char* name) { fn = reinterpret_cast<std::remove_reference_t<decltype(fn)>>(dlsym(h, name)); if (!fn) { all = false; api.why = std::string("NCCL symbol missing: ") + name; } };
        bind(api.GetUniqueId, "ncclGetUniqueId"); bind(api.CommInitRank, "ncclCommInitRank"); bind(api.CommInitAll, "ncclCommInitAll"); bind(api.CommDestroy, "ncclCommDestroy");
        bind(api.AllReduce, "ncclAllReduce"); bind(api.GroupStart, "ncclGroupStart"); bind(api.GroupEnd, "ncclGroupEnd"); bind(api.GetErrorString, "ncclGetErrorString");
        api.ok = all;
    });
    return api;
}
bool nccl_ready(const char* who) {
    if (nccl().ok) return true;
    set_error(std::string(who) + ": " + nccl().why);
    return false;
}
#define NCK(call)                                                                                                   \
    do {                                                                                                            \
        ncclResult_t r_ = (call);                                                                                   \
        if (r_ != ncclSuccess) { set_error(std::string(#call) + ": " + nccl().GetErrorString(r_)); return RT_ERR_CUDA; } \
    } while (0)
#define CK(call)                                                                                                    \
    do {                                                                                                            \
        cudaError_t e_ = (call);                                                                                    \
        if (e_ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); return RT_ERR_CUDA; } \
    } while (0)

}  // namespace
}  // namespace rt

// local: comms[i] lives on devices[i] of this process; rank form: one comm on `devices[0]`, `world` ranks in total
struct rt_comm {
    std::vector<ncclComm_t> comms; std::vector<int> devices; std::vector<cudaStream_t> streams;
    uint32_t world = 0; bool local = false;
};

using namespace rt;

extern "C" {

rt_status rt_comm_create_local(const int32_t* devices, uint32_t n, rt_comm** out) {
    if (!devices || !n || !out) { set_error("rt_comm_create_local: bad argument"); return RT_ERR_INVALID; }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { set_error("rt_comm_create_local: no CUDA device: librtcore has no CPU path"); return RT_ERR_CUDA; }
    for (uint32_t i = 0; i < n; i++) {
        if (devices[i] < 0 || devices[i] >= count) { set_error("rt_comm_create_local: device index out of range"); return RT_ERR_INVALID; }
        for (uint32_t j = 0; j < i; j++) if (devices[j] == devices[i]) { set_error("rt_comm_create_local: a device is listed twice"); return RT_ERR_INVALID; }
    }
    if (!nccl_ready("rt_comm_create_local")) return RT_ERR_STATE;
    rt_comm* c = new rt_comm(); c->local = true; c->world = n;
    c->devices.assign(devices, devices + n); c->comms.resize(n); c->streams.resize(n, nullptr);
    ncclResult_t r = nccl().CommInitAll(c->comms.data(), (int)n, c->devices.data());
    if (r != ncclSuccess) { set_error(std::string("ncclCommInitAll: ") + nccl().GetErrorString(r)); delete c; return RT_ERR_CUDA; }
    int prev = 0; cudaGetDevice(&prev);
    for (uint32_t i = 0; i < n; i++) { cudaSetDevice(c->devices[i]); cudaStreamCreateWithFlags(&c->streams[i], cudaStreamNonBlocking); }
    cudaSetDevice(prev);
    *out = c; return RT_OK;
}
rt_status rt_comm_unique_id(uint8_t id_out[128]) {
    if (!id_out) { set_error("rt_comm_unique_id: null argument"); return RT_ERR_INVALID; }
    if (!nccl_ready("rt_comm_unique_id")) return RT_ERR_STATE;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NCK(nccl().GetUniqueId(&id));
    memcpy(id_out, &id, 128);
    return RT_OK;
}
rt_status rt_comm_create_rank(const uint8_t id_in[128], uint32_t world, uint32_t rank, rt_comm** out) {
    if (!id_in || !out || !world || rank >= world) { set_error("rt_comm_create_rank: bad argument"); return RT_ERR_INVALID; }
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("rt_comm_create_rank: no CUDA device: librtcore has no CPU path"); return RT_ERR_CUDA; }
    if (!nccl_ready("rt_comm_create_rank")) return RT_ERR_STATE;
    ncclUniqueId id; memcpy(&id, id_in, 128);
    rt_comm* c = new rt_comm(); c->local = false; c->world = world; c->devices.push_back(dev); c->comms.resize(1); c->streams.resize(1, nullptr);
    ncclResult_t r = nccl().CommInitRank(&c->comms[0], (int)world, id, (int)rank);
    if (r != ncclSuccess) { set_error(std::string("ncclCommInitRank: ") + nccl().GetErrorString(r)); delete c; return RT_ERR_CUDA; }
    cudaStreamCreateWithFlags(&c->streams[0], cudaStreamNonBlocking);
    *out = c; return RT_OK;
}
rt_status rt_comm_destroy(rt_comm* c) {
    if (!c) return RT_OK;
    int prev = 0; cudaGetDevice(&prev);
    for (size_t i = 0; i < c->comms.size(); i++) {
        cudaSetDevice(c->devices[i]);
        if (c->streams[i]) cudaStreamDestroy(c->streams[i]);
        if (c->comms[i] && nccl().ok) nccl().CommDestroy(c->comms[i]);
    }
    cudaSetDevice(prev);
    delete c; return RT_OK;
}
rt_status rt_comm_size(const rt_comm* c, uint32_t* world, uint32_t* n_local) {
    if (!c) { set_error("rt_comm_size: null communicator"); return RT_ERR_INVALID; }
    if (world) *world = c->world;
    if (n_local) *n_local = (uint32_t)c->comms.size();
    return RT_OK;
}

// In-place sum of the (r,g,b,n) films over all ranks: ncclAllReduce(f64, sum), 4 * w * h values.  films[i] must live on the communicator's i-th
// local device.  Blocking: returns when every local film holds the total.
rt_status rt_film_allreduce(rt_comm* c, rt_film* const* films, uint32_t n_films) {
    if (!c || !films || n_films != c->comms.size()) { set_error("rt_film_allreduce: needs one film per local device of the communicator"); return RT_ERR_INVALID; }
    for (uint32_t i = 0; i < n_films; i++) {
        if (!films[i] || films[i]->w != films[0]->w || films[i]->h != films[0]->h) { set_error("rt_film_allreduce: films differ in size"); return RT_ERR_INVALID; }
        if (films[i]->device != c->devices[i]) { set_error("rt_film_allreduce: film " + std::to_string(i) + " is not on the communicator's device"); return RT_ERR_INVALID; }
    }
    int prev = 0; cudaGetDevice(&prev);
    const size_t count = (size_t)films[0]->w * films[0]->h * 4;
    NCK(nccl().GroupStart());
    for (uint32_t i = 0; i < n_films; i++) {
        cudaSetDevice(c->devices[i]);
        ncclResult_t r = nccl().AllReduce(films[i]->d, films[i]->d, count, ncclFloat64, ncclSum, c->comms[i], c->streams[i]);
        if (r != ncclSuccess) { nccl().GroupEnd(); cudaSetDevice(prev); set_error(std::string("ncclAllReduce: ") + nccl().GetErrorString(r)); return RT_ERR_CUDA; }
    }
    NCK(nccl().GroupEnd());
    rt_status st = RT_OK;
    for (uint32_t i = 0; i < n_films; i++) {
        cudaSetDevice(c->devices[i]);
        cudaError_t e = cudaStreamSynchronize(c->streams[i]);
        if (e != cudaSuccess) { set_error(std::string("rt_film_allreduce: ") + cudaGetErrorString(e)); st = RT_ERR_CUDA; }
    }
    cudaSetDevice(prev);
    return st;
}

// A second scene with the same host tables, committed on `device` (the scene is replicated).  Build mode and wavefront size carry over.
rt_status rt_scene_replicate(const rt_scene* src, int32_t device, rt_scene** out) {
    if (!src || !out) { set_error("rt_scene_replicate: null argument"); return RT_ERR_INVALID; }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { set_error("rt_scene_replicate: no CUDA device: librtcore has no CPU path"); return RT_ERR_CUDA; }
    if (device < 0 || device >= count) { set_error("rt_scene_replicate: device index out of range"); return RT_ERR_INVALID; }
    rt_scene* s = new rt_scene();
    s->meshes = src->meshes; s->prims = src->prims; s->shapes = src->shapes; s->xforms = src->xforms; s->textures = src->textures;
    s->materials = src->materials; s->disneys = src->disneys; s->lights = src->lights; s->images = src->images; s->texels = src->texels; s->perlin = src->perlin;
    s->build_mode = src->build_mode; s->wavefront_size = src->wavefront_size;
    int prev = 0; cudaGetDevice(&prev);
    rt_status st = RT_ERR_CUDA;
    if (cudaSetDevice(device) == cudaSuccess) st = rt_scene_commit(s); else set_error("rt_scene_replicate: cudaSetDevice failed");
    cudaSetDevice(prev);
    if (st != RT_OK) { rt_scene_destroy(s); return st; }
    *out = s; return RT_OK;
}

// render::tile_multithread (render.rs:13-159) on the n GPUs of a local communicator: scenes[g] / films[g] live on its g-th device; one host thread
// per GPU renders its share of desc's sample range into films[g] (which it ADDS to, like rt_render), then the films are summed in place.
// Films must therefore enter with the same content on every GPU only if that content is meant to be counted n times: clear them first.
rt_status rt_render_multi(rt_comm* c, rt_scene* const* scenes, rt_film* const* films, uint32_t n, const rt_camera* cam,
                          const rt_integrator_desc* integ, const rt_film_desc* desc) {
    if (!c || !scenes || !films || !cam || !integ || !desc) { set_error("rt_render_multi: null argument"); return RT_ERR_INVALID; }
    if (!c->local || n != c->comms.size()) { set_error("rt_render_multi: needs a local communicator (rt_comm_create_local) and one scene + film per device of it"); return RT_ERR_INVALID; }
    if (desc->sample_end < desc->sample_begin) { set_error("rt_render_multi: bad sample range"); return RT_ERR_INVALID; }
    for (uint32_t g = 0; g < n; g++) {
        if (!scenes[g] || !films[g]) { set_error("rt_render_multi: null scene or film"); return RT_ERR_INVALID; }
        if (!scenes[g]->committed) { set_error("rt_render_multi: scene " + std::to_string(g) + " is not committed"); return RT_ERR_STATE; }
        if (scenes[g]->device != c->devices[g] || films[g]->device != c->devices[g]) { set_error("rt_render_multi: scene / film " + std::to_string(g) + " is not on the communicator's device"); return RT_ERR_INVALID; }
    }
    const uint64_t total = desc->sample_end - desc->sample_begin;
    std::vector<rt_status> status(n, RT_OK); std::vector<std::string> msg(n);
    std::vector<std::thread> pool;
    for (uint32_t g = 0; g < n; g++) pool.emplace_back([&, g]() {
        rt_film_desc d = *desc;
        d.sample_begin = desc->sample_begin + (uint32_t)(total * g / n);
        d.sample_end = desc->sample_begin + (uint32_t)(total * (g + 1) / n);
        status[g] = cudaSetDevice(c->devices[g]) == cudaSuccess ? rt_render(scenes[g], cam, integ, &d, films[g]) : RT_ERR_CUDA;
        if (status[g] != RT_OK) msg[g] = rt_last_error_string();
    });
    for (auto& t : pool) t.join();
    for (uint32_t g = 0; g < n; g++) if (status[g] != RT_OK) { set_error("rt_render_multi: device " + std::to_string(c->devices[g]) + ": " + msg[g]); return status[g]; }
    return rt_film_allreduce(c, films, n);
}

// The N films of one process's GPUs -> one RGB8 image, without NCCL and without a reduced film in memory (k_reduce_resolve above).  What follows
// rt_render on several GPUs when only the picture is wanted (render.rs:117-153: merge the workers' tiles, then save); rt_film_allreduce stays for
// callers that go on accumulating.  The films are not modified.
rt_status rt_film_reduce_resolve_rgb8(rt_film* const* films, uint32_t n_films, uint8_t* host_rgb) {
    if (!films || !host_rgb || n_films == 0) { set_error("rt_film_reduce_resolve_rgb8: null argument"); return RT_ERR_INVALID; }
    if (n_films > (uint32_t)REDUCE_MAX_FILMS) { set_error("rt_film_reduce_resolve_rgb8: at most " + std::to_string(REDUCE_MAX_FILMS) + " films"); return RT_ERR_INVALID; }
    for (uint32_t i = 0; i < n_films; i++) {
        if (!films[i]) { set_error("rt_film_reduce_resolve_rgb8: null film"); return RT_ERR_INVALID; }
        if (films[i]->w != films[0]->w || films[i]->h != films[0]->h) { set_error("rt_film_reduce_resolve_rgb8: films differ in size"); return RT_ERR_INVALID; }
    }
    const int root = films[0]->device;
    int prev = 0; cudaGetDevice(&prev);
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    FilmPtrs fp{};
    std::vector<int> devs;                                         // the distinct GPUs, root first: each one takes a slice of the pixels
    for (uint32_t i = 0; i < n_films; i++) {
        if (reinterpret_cast<uintptr_t>(films[i]->d) & 15u) { set_error("rt_film_reduce_resolve_rgb8: film memory must be 16-byte aligned"); return RT_ERR_INVALID; }
        fp.p[i] = reinterpret_cast<const double2*>(films[i]->d);
        if (std::find(devs.begin(), devs.end(), films[i]->device) == devs.end()) devs.push_back(films[i]->device);
    }
    for (int a : devs) for (int b : devs) {
        if (a == b) continue;
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, a, b));
        if (!can) { set_error("rt_film_reduce_resolve_rgb8: device " + std::to_string(a) + " has no peer access to device " + std::to_string(b) + " (use rt_film_allreduce + rt_film_resolve_rgb8)"); return RT_ERR_STATE; }
        CK(cudaSetDevice(a));
        const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) { set_error(std::string("rt_film_reduce_resolve_rgb8: cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    }
    CK(cudaSetDevice(root));
    const uint64_t n = (uint64_t)films[0]->w * films[0]->h;
    uint8_t* d = nullptr;
    CK(cudaMalloc((void**)&d, n * 3));
    const uint64_t nd = devs.size(), slice = (n + nd - 1) / nd;
    cudaError_t e = cudaSuccess;
    for (uint64_t g = 0; g < nd && e == cudaSuccess; g++) {
        const uint64_t b0 = std::min(n, g * slice), e0 = std::min(n, b0 + slice);
        if (e0 == b0) continue;
        if ((e = cudaSetDevice(devs[g])) != cudaSuccess) break;
        int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, devs[g]);
        k_reduce_resolve<<<(unsigned)std::min<uint64_t>((e0 - b0 + 255) / 256, (uint64_t)sms * 8), 256, 0, cudaStreamPerThread>>>(fp, (int)n_films, b0, e0, d);
        e = cudaGetLastError();
    }
    for (uint64_t g = 0; g < nd; g++) {                            // every launched kernel is waited for, error or not, before the buffer is freed
        cudaSetDevice(devs[g]);
        const cudaError_t w = cudaStreamSynchronize(cudaStreamPerThread);
        if (e == cudaSuccess) e = w;
    }
    cudaSetDevice(root);
    if (e == cudaSuccess) e = cudaMemcpy(host_rgb, d, n * 3, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) { set_error(std::string("rt_film_reduce_resolve_rgb8: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}

}  // extern "C"

