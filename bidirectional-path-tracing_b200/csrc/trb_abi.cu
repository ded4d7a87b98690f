// trb_abi.cu -- device half of include/trb.h: context, scene upload, wave scheduling, parity entry points.
// One context = one GPU.  All launches go to the context's stream; nothing here falls back to the CPU.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/trb.h"
#include "host/tr_host.h"
#include "kernels/bdpt_kernels.cuh"
#include "kernels/pt_kernels.cuh"
#include "kernels/di_kernels.cuh"
#include "kernels/bvh_build.cuh"
#include "kernels/trace_coop.cuh"
#include "kernels/trace_wide.cuh"

using namespace trbk;

namespace {

thread_local std::string g_create_err;

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {
        if (count <= n && p) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) n = count; else p = nullptr;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    size_t bytes() const { return n * sizeof(T); }
};

#define TRB_SCENE_ARRAYS 12
enum LaunchKind { LK_CLOSEST = 0, LK_ANY = 1, LK_LOGIC = 2, LK_N = 3 };
#define TRB_NSTATS 16

}  // namespace

struct trb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;  // shadow-ray traversal of stage d overlaps the closest-hit traversal of stage d+1
    std::string err;
    cudaDeviceProp prop{};
    // scene
    bool has_scene = false;
    bool built_on_device = false;   // the acceleration structure and the triangle records came from kernels/bvh_build.cuh
    int trace_blocks_per_sm = TRB_TRACE_MINBLOCKS;   // persistent traversal grid = SMs x this ($TRB_TRACE_GRID_BLOCKS at trb_create: experiments with room left for other kernels)
    float near_inv_max = 0.f;       // TraceArgs::inv_max of the uploaded scene (kernels/trace.cuh: near_cull)
    bool force_binary = false;      // TRB_TRAVERSAL=binary at trb_create: every ray on the reference's binary tree (A/B measurements, tests of k_trace_coop)
    trb::FlatScene flat;
    DevBuf<trb::DevNode> d_nodes;      // the reference's binary tree (degenerate rays, k_pt_check)
    DevBuf<trb::DevNode4> d_wnodes;   // the 4-wide upper tree (every other ray)
    DevBuf<trb::DevTri> d_tris;
    DevBuf<trb::DevShade> d_shade;
    DevBuf<trb::DevMaterial> d_mats;
    DevBuf<trb::DevEmitter> d_emitters;
    DevBuf<trb::DevEmitterTri> d_etris;
    DevBuf<float> d_cdf;
    DevBuf<float4> d_esphere;
    DevBuf<trb::DevTexInfo> d_tex_info;   // bitmap textures: empty for an untextured scene
    DevBuf<trb::DevTriUV> d_tri_uv;
    DevBuf<float> d_texels;
    SceneDev sc{};
    uint8_t* pinned = nullptr;  // flattened scene staged in pinned host memory (trb_scene_refresh)
    size_t pinned_bytes = 0;
    // wave buffers
    DevBuf<RayT> cam_ray, ray, sray, sray2;   // two shadow queues: one is traced while the next is being filled
    DevBuf<HitT> hit, hit0;
    DevBuf<float4> st_t, st_v;
    DevBuf<RayT> ray_b;               // BDPT: second set of the queue-ordered walk records (WaveParams: ray_in / ray_out ...)
    DevBuf<float4> st_t_b, sv_a, sv_b;
    DevBuf<LightVertex> lv;
    DevBuf<int32_t> nlv, q_a, q_b, q_conn;
    DevBuf<uint32_t> pt_ctr;             // path tracer: one counter per queue / traversal launch
    DevBuf<int32_t> defer_q[2];          // slots of the rays k_trace_wide leaves to the binary tree: [0] main stream (closest hits), [1] shadow stream
    DevBuf<EyeVertex> ev;
    DevBuf<float4> lv_tex, ev_tex;       // textured scenes only: (triangle, u, v) of the stored vertices
    DevBuf<Pending> spend, spend2;
    DevBuf<uint32_t> counters;           // per stage CTR_N counters
    DevBuf<uint32_t> fetch_ctr;          // queue cursor of the stand-alone trace entry points
    DevBuf<unsigned long long> d_stats;  // TRB_NSTATS words: [0] closest rays, [1] any rays, [2..5] wide node visits / tri tests (closest, any), [6,7] re-traces,
                                         // [8..11] binary-tree node visits / tri tests of the deferred rays (closest, any), [12,13] deferred rays
    DevBuf<float> fb;                    // frame buffer W*H*3
    // scratch for the parity entry points
    DevBuf<RayT> io_rays;
    DevBuf<HitT> io_hits;
    DevBuf<uint8_t> io_occ;
    // timing
    std::vector<cudaEvent_t> events, events2;   // per stream
    std::vector<int> event_kind, event_kind2;
    size_t n_events = 0, n_events2 = 0;
    std::vector<cudaEvent_t> dep_a, dep_b;       // untimed dependency events between the two streams
    bool timing = false;
    uint64_t launches = 0;
    uint64_t wave_used = 0;      // samples per wavefront of the last render call

    int fail(int code, const std::string& m) { err = m; return code; }
    int cuda_fail(cudaError_t e, const char* what) {
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return TRB_ERR_CUDA;
    }
    void mark(int kind, bool second = false) {  // kind < 0: a begin marker on the second stream, not a launch
        if (kind >= 0) ++launches;
        if (!timing) return;
        std::vector<cudaEvent_t>& evs = second ? events2 : events;
        std::vector<int>& kinds = second ? event_kind2 : event_kind;
        size_t& n = second ? n_events2 : n_events;
        if (n == evs.size()) { cudaEvent_t ev; cudaEventCreate(&ev); evs.push_back(ev); kinds.push_back(0); }
        kinds[n] = kind;
        cudaEventRecord(evs[n++], second ? stream2 : stream);
    }
};

#define TRB_CUDA(ctx, call)                                      \
    do {                                                         \
        cudaError_t e__ = (call);                                \
        if (e__ != cudaSuccess) return (ctx)->cuda_fail(e__, #call); \
    } while (0)

namespace {

inline int grid_for(size_t n, int sm_count) {
    size_t blocks = (n + TRB_BLOCK - 1) / TRB_BLOCK;
    size_t cap = size_t(sm_count) * 32;  // grid-stride kernels: a multiple of the SM count
    return int(std::max<size_t>(1, std::min(blocks, cap)));
}
inline int grid_exact(size_t n) { return int(std::max<size_t>(1, (n + TRB_BLOCK - 1) / TRB_BLOCK)); }


// `ctr`: the launch's own TRACE_CTR_WORDS counters (ray-queue cursor, deferred count, cursor of the deferred launch), zeroed by the
// caller on the stream.  k_trace_wide traces every ray but those with a zero / non-finite direction component, which it appends to
// defer_q; k_trace_coop then traces those on the reference's binary tree (a grid of warps that find an empty queue and leave).
int ensure_defer(trb_context* c, int which, size_t n) { TRB_CUDA(c, c->defer_q[which].alloc(n)); return TRB_OK; }

void launch_closest(trb_context* c, const RayT* rays, const int32_t* q, const uint32_t* n_ptr, uint32_t n_fixed, HitT* hits,
                    uint32_t* ctr, unsigned long long* visit) {
    TraceArgs a{};
    a.bvh = BvhView{c->sc.nodes, c->sc.tris, c->sc.root_ref};
    a.wnodes = reinterpret_cast<const F4*>(c->d_wnodes.p); a.wroot = c->flat.n_wnodes ? 0 : c->sc.root_ref;
    a.n_nodes = uint32_t(c->flat.n_nodes); a.n_wnodes = uint32_t(c->flat.n_wnodes); a.n_tris = uint32_t(c->flat.n_tris);
    a.rays = rays; a.q = q; a.n_ptr = n_ptr; a.n_fixed = n_fixed; a.cap = 0xffffffffu; a.hits = hits; a.fetch = ctr;
    a.inv_max = c->near_inv_max;
    a.defer_q = c->defer_q[0].p; a.defer_n = ctr + 1; a.defer_cap = uint32_t(std::min<size_t>(c->defer_q[0].n, 0xffffffffu));
    a.visit = visit; a.visit_exh = visit ? c->d_stats.p + 6 : nullptr;
    const int grid = c->prop.multiProcessorCount * c->trace_blocks_per_sm;
    if (c->force_binary) { k_trace_coop<false><<<grid, TRB_TRACE_BLOCK, 0, c->stream>>>(a); return; }
    k_trace_wide<false><<<grid, TRB_TRACE_BLOCK, 0, c->stream>>>(a);
    a.q = c->defer_q[0].p; a.n_ptr = ctr + 1; a.n_fixed = 0u; a.cap = uint32_t(std::min<size_t>(c->defer_q[0].n, 0xffffffffu)); a.fetch = ctr + 2;
    if (visit) a.visit = c->d_stats.p + 8;
    k_trace_coop<false><<<grid, TRB_TRACE_BLOCK, 0, c->stream>>>(a);
}

void launch_any(trb_context* c, const RayT* rays, const uint32_t* n_ptr, uint32_t n_fixed, uint32_t cap, uint8_t* occ_out,
                const Pending* pend, float* accum, uint32_t* ctr, unsigned long long* visit, cudaStream_t stream = nullptr) {
    const int which = (stream && stream != c->stream) ? 1 : 0;
    if (!stream) stream = c->stream;
    TraceArgs a{};
    a.bvh = BvhView{c->sc.nodes, c->sc.tris, c->sc.root_ref};
    a.wnodes = reinterpret_cast<const F4*>(c->d_wnodes.p); a.wroot = c->flat.n_wnodes ? 0 : c->sc.root_ref;
    a.n_nodes = uint32_t(c->flat.n_nodes); a.n_wnodes = uint32_t(c->flat.n_wnodes); a.n_tris = uint32_t(c->flat.n_tris);
    a.rays = rays; a.q = nullptr; a.n_ptr = n_ptr; a.n_fixed = n_fixed; a.cap = cap; a.occ_out = occ_out; a.pend = pend; a.accum = accum; a.fetch = ctr;
    a.inv_max = c->near_inv_max;
    a.defer_q = c->defer_q[which].p; a.defer_n = ctr + 1; a.defer_cap = uint32_t(std::min<size_t>(c->defer_q[which].n, 0xffffffffu));
    a.visit = visit; a.visit_exh = visit ? c->d_stats.p + 7 : nullptr;
    const int grid = c->prop.multiProcessorCount * c->trace_blocks_per_sm;
    if (c->force_binary) { k_trace_coop<true><<<grid, TRB_TRACE_BLOCK, 0, stream>>>(a); return; }
    k_trace_wide<true><<<grid, TRB_TRACE_BLOCK, 0, stream>>>(a);
    a.q = c->defer_q[which].p; a.n_ptr = ctr + 1; a.n_fixed = 0u; a.cap = uint32_t(std::min<size_t>(c->defer_q[which].n, 0xffffffffu)); a.fetch = ctr + 2;
    if (visit) a.visit = c->d_stats.p + 10;
    k_trace_coop<true><<<grid, TRB_TRACE_BLOCK, 0, stream>>>(a);
}

// depth cap and vertices per subpath of a render description (bdpt.h:68,188: `depth < rrDepth`, depth starts at 1)
inline bool rr_enabled(const trb_render_desc& d) { return (d.flags & TRB_FLAG_RUSSIAN_ROULETTE) != 0; }
inline int max_depth_of(const trb_render_desc& d) {
    if (!rr_enabled(d)) return d.rr_depth;
    return d.max_depth > 0 ? d.max_depth : std::max(TRB_RR_DEFAULT_MAX_DEPTH, d.rr_depth);
}
inline int max_vertices_of(const trb_render_desc& d) { return std::max(max_depth_of(d) - 1, 0); }

// Samples per wavefront when the caller leaves it to the library (trb_render_desc.wave_samples == 0).  Bigger waves amortise the tails of
// the persistent kernels and - on scenes larger than the L2 - put more rays in flight per node (measured, B200, profiles/r2u_wave.log:
// 10 M triangles 75.2 / 79.3 / 81.5 Msamples/s at 2^22 / 2^23 / 2^24; caustic 117.0 / 117.8 / 118.2).  Memory is what a 180 GB part has:
// the largest power of two <= 2^24 whose wave buffers (bytes per slot below: what ensure_wave allocates) fit in a quarter of the
// device memory that is free now (buffers the context already holds count as free), never below 2^20.
// `shadow`: BDPT (two shadow-ray queues of K + 1 entries per slot, the second set of walk records); false = path tracer / direct integrator.
int64_t default_wave_samples(trb_context* c, int K, bool shadow = true) {
    const size_t kk = size_t(std::max(K, 1));
    const size_t per_slot = (shadow ? 308 + 100 * (kk + 1) : 228) + 68 * kk + (c->sc.has_tex ? 16 * kk + 16 : 0);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return int64_t(1) << 22;
    const size_t held = (c->lv.bytes() + c->sray.bytes() + c->sray2.bytes() + c->spend.bytes() + c->spend2.bytes() + c->ray.bytes() + c->cam_ray.bytes());
    const size_t budget = (free_b + held) / 4;
    int shift = 24;
    while (shift > 20 && (per_slot << shift) > budget) --shift;
    return int64_t(1) << shift;
}

int ensure_wave(trb_context* c, size_t P, int K, bool shadow = true) {
    size_t kk = size_t(std::max(K, 1));
    size_t shadow_cap = shadow ? P * (kk + 1) : 1;  // (the path tracer traces closest hits only)
    TRB_CUDA(c, c->cam_ray.alloc(P)); TRB_CUDA(c, c->ray.alloc(P));
    TRB_CUDA(c, c->hit.alloc(P)); TRB_CUDA(c, c->hit0.alloc(P));
    TRB_CUDA(c, c->st_t.alloc(P)); TRB_CUDA(c, c->st_v.alloc(P));
    if (shadow) { TRB_CUDA(c, c->ray_b.alloc(P)); TRB_CUDA(c, c->st_t_b.alloc(P)); TRB_CUDA(c, c->sv_a.alloc(P)); TRB_CUDA(c, c->sv_b.alloc(P)); }   // (BDPT only)
    TRB_CUDA(c, c->lv.alloc(P * kk)); TRB_CUDA(c, c->nlv.alloc(P));
    TRB_CUDA(c, c->q_a.alloc(P)); TRB_CUDA(c, c->q_b.alloc(P)); TRB_CUDA(c, c->q_conn.alloc(P * kk)); TRB_CUDA(c, c->ev.alloc(P));  // q_conn: (slot, light vertex) pairs
    TRB_CUDA(c, c->sray.alloc(shadow_cap)); TRB_CUDA(c, c->spend.alloc(shadow_cap));
    TRB_CUDA(c, c->sray2.alloc(shadow_cap)); TRB_CUDA(c, c->spend2.alloc(shadow_cap));
    if (c->sc.has_tex) { TRB_CUDA(c, c->lv_tex.alloc(P * kk)); TRB_CUDA(c, c->ev_tex.alloc(P)); }
    TRB_CUDA(c, c->counters.alloc(size_t(2 * kk + 8) * CTR_N));
    TRB_CUDA(c, c->defer_q[0].alloc(P)); TRB_CUDA(c, c->defer_q[1].alloc(shadow_cap));
    TRB_CUDA(c, c->d_stats.alloc(TRB_NSTATS));
    return TRB_OK;
}

__global__ void k_fold_stats(const uint32_t* __restrict__ counters, int n_stage, int light_stages, int K, uint32_t P,
                             unsigned long long* __restrict__ stats, uint32_t shadow_cap) {
    // stage 0 = light_init (its QOUT feeds light depth 1), stages 1..K = light bounces, K+1 = eye_init, K+2.. = eye bounces
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long closest = P, any = 0;
    for (int s = 0; s < n_stage; ++s) {
        uint32_t qo = counters[s * CTR_N + CTR_QOUT], sh = min(counters[s * CTR_N + CTR_SHADOW], shadow_cap);
        any += sh;
        bool is_eye_init = (s == K + 1);
        if (!is_eye_init) closest += qo;  // every queued slot is traced once more; eye depth 1 reuses the primary hit
    }
    unsigned long long dc = 0, da = 0;   // rays k_trace_wide left to the binary tree (word [1] of every launch's counters)
    for (int s = 0; s < n_stage; ++s) { dc += counters[s * CTR_N + CTR_FETCH_CLOSEST + 1]; da += counters[s * CTR_N + CTR_FETCH_ANY + 1]; }
    stats[0] += closest; stats[1] += any; stats[12] += dc; stats[13] += da;
}

}  // namespace

extern "C" {

int trb_create(int device, trb_context** out) {
    if (!out) return TRB_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0 || device < 0 || device >= n) {
        g_create_err = std::string("no usable CUDA device (the BDPT path has no CPU fallback): ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
        // surface through the host error channel too
        return TRB_ERR_NO_DEVICE;
    }
    auto* c = new trb_context();
    c->device = device;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&c->prop, device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess) {
        g_create_err = "cudaSetDevice / stream creation failed";
        delete c;
        return TRB_ERR_NO_DEVICE;
    }
    if (c->prop.major < 10) {
        g_create_err = "device is not sm_100 class (library is built for sm_100a only)";
        cudaStreamDestroy(c->stream);
        cudaStreamDestroy(c->stream2);
        delete c;
        return TRB_ERR_NO_DEVICE;
    }
    if (const char* e = std::getenv("TRB_TRAVERSAL")) c->force_binary = std::string(e) == "binary";
    if (const char* e = std::getenv("TRB_TRACE_GRID_BLOCKS")) { const int v = std::atoi(e); if (v >= 1 && v <= TRB_TRACE_MINBLOCKS) c->trace_blocks_per_sm = v; }
    *out = c;
    return TRB_OK;
}

void trb_destroy(trb_context* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->d_nodes.release(); c->d_wnodes.release(); c->defer_q[0].release(); c->defer_q[1].release(); c->d_tris.release(); c->d_shade.release(); c->d_mats.release(); c->d_emitters.release();
    c->d_etris.release(); c->d_cdf.release(); c->d_esphere.release(); c->pt_ctr.release();
    c->d_tex_info.release(); c->d_tri_uv.release(); c->d_texels.release(); c->lv_tex.release(); c->ev_tex.release();
    c->cam_ray.release(); c->ray.release(); c->sray.release(); c->sray2.release(); c->spend2.release(); c->hit.release(); c->hit0.release(); c->st_t.release();
    c->ray_b.release(); c->st_t_b.release(); c->sv_a.release(); c->sv_b.release();
    c->st_v.release(); c->lv.release(); c->nlv.release(); c->q_a.release(); c->q_b.release(); c->q_conn.release(); c->ev.release(); c->spend.release();
    c->counters.release(); c->fetch_ctr.release(); c->d_stats.release(); c->fb.release(); c->io_rays.release(); c->io_hits.release(); c->io_occ.release();
    if (c->pinned) cudaFreeHost(c->pinned);
    for (auto ev : c->events) cudaEventDestroy(ev);
    for (auto ev : c->events2) cudaEventDestroy(ev);
    for (auto ev : c->dep_a) cudaEventDestroy(ev);
    for (auto ev : c->dep_b) cudaEventDestroy(ev);
    cudaStreamSynchronize(c->stream2);
    cudaStreamDestroy(c->stream2);
    cudaStreamDestroy(c->stream);
    delete c;
}

const char* trb_last_error(const trb_context* c) { return c ? c->err.c_str() : g_create_err.c_str(); }

const char* trb_build_info(void) {
#ifdef TRB_COUNTERS
    return "trb sm_100a fmad=off exact_transcendentals=" "1" " counters=on";
#else
    return "trb sm_100a fmad=off exact_transcendentals=" "1" " counters=off";
#endif
}

}  // extern "C"

namespace {

void scene_array_sizes(const trb::FlatScene& f, size_t sizes[TRB_SCENE_ARRAYS]) {
    const size_t v[TRB_SCENE_ARRAYS] = {f.n_nodes * sizeof(trb::DevNode), f.n_tris * sizeof(trb::DevTri), f.n_tris * sizeof(trb::DevShade),
                                        f.materials.size() * sizeof(trb::DevMaterial), f.emitters.size() * sizeof(trb::DevEmitter),
                                        f.emitter_tris.size() * sizeof(trb::DevEmitterTri), f.emitter_cdf.size() * sizeof(float),
                                        f.emitter_sphere.size() * sizeof(float), f.tex_info.size() * sizeof(trb::DevTexInfo),
                                        f.n_tri_uv * sizeof(trb::DevTriUV), f.texels.size() * sizeof(float), f.n_wnodes * sizeof(trb::DevNode4)};
    for (int k = 0; k < TRB_SCENE_ARRAYS; ++k) sizes[k] = v[k];
}

// The flattened scene in ONE pinned blob, so that trb_scene_refresh is plain async copies from pinned memory.  Arrays the host
// built are copied from their vectors; arrays the device built (kernels/bvh_build.cuh) are read back from the device.
int stage_scene(trb_context* c) {
    const trb::FlatScene& f = c->flat;
    size_t sizes[TRB_SCENE_ARRAYS];
    scene_array_sizes(f, sizes);
    const void* srcs[TRB_SCENE_ARRAYS] = {f.nodes.data(), f.tris.data(), f.shade.data(), f.materials.data(), f.emitters.data(), f.emitter_tris.data(),
                                          f.emitter_cdf.data(), f.emitter_sphere.data(), f.tex_info.data(), f.tri_uv.data(), f.texels.data(), f.wnodes.data()};
    const bool on_host[TRB_SCENE_ARRAYS] = {!f.nodes.empty(), !f.tris.empty(), !f.shade.empty(), true, true, true, true, true, true, !f.tri_uv.empty(), true, !f.wnodes.empty()};
    const void* devs[TRB_SCENE_ARRAYS] = {c->d_nodes.p, c->d_tris.p, c->d_shade.p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, c->d_tri_uv.p, nullptr, c->d_wnodes.p};
    size_t total = 0;
    for (size_t b : sizes) total += (b + 255) & ~size_t(255);
    if (c->pinned) { cudaFreeHost(c->pinned); c->pinned = nullptr; }
    TRB_CUDA(c, cudaMallocHost(reinterpret_cast<void**>(&c->pinned), std::max<size_t>(total, 256)));
    c->pinned_bytes = total;
    size_t off = 0;
    std::vector<std::thread> th;   // the staging copy (1.8 GB for 10 M triangles) on a few host threads
    for (int k = 0; k < TRB_SCENE_ARRAYS; ++k) {
        if (sizes[k] && !on_host[k]) {
            TRB_CUDA(c, cudaMemcpyAsync(c->pinned + off, devs[k], sizes[k], cudaMemcpyDeviceToHost, c->stream));
        } else {
            const size_t piece = size_t(64) << 20;
            for (size_t o = 0; o < sizes[k]; o += piece) {
                uint8_t* dst = c->pinned + off + o; const char* src = static_cast<const char*>(srcs[k]) + o; const size_t len = std::min(piece, sizes[k] - o);
                if (sizes[k] <= piece || th.size() >= 8) std::memcpy(dst, src, len);
                else th.emplace_back([=]() { std::memcpy(dst, src, len); });
            }
        }
        off += (sizes[k] + 255) & ~size_t(255);
    }
    for (auto& t : th) t.join();
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

// BVH::build + the per-triangle records on the device (kernels/bvh_build.cuh): same arrays as trb::flatten_geometry
int build_geometry_on_device(trb_context* c, const trb_scene_desc& s) {
    trb::FlatScene& f = c->flat;
    const bool timing = std::getenv("TRB_BUILD_TIMING") != nullptr;
    auto t_prev = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        cudaStreamSynchronize(c->stream);
        auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[trb build] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
        t_prev = now;
    };
    const uint32_t n = uint32_t(s.n_tris);
    DevBuf<float> d_pos, d_nrm, d_tc;
    DevBuf<int32_t> d_tv, d_tn, d_tm, d_ts, d_sf, d_se, d_tt, ord_a, ord_b, range_of;
    DevBuf<float4> tmin, tmax, tcen;
    DevBuf<uint32_t> is_l, pref, r_pos, last_r, counter;
    DevBuf<BuildNode> nodes;
    DevBuf<uint8_t> cub_tmp;
    auto release = [&]() {
        d_pos.release(); d_nrm.release(); d_tc.release(); d_tv.release(); d_tn.release(); d_tm.release(); d_ts.release(); d_sf.release(); d_se.release();
        d_tt.release(); ord_a.release(); ord_b.release(); range_of.release(); tmin.release(); tmax.release(); tcen.release(); is_l.release(); pref.release(); r_pos.release(); last_r.release();
        counter.release(); nodes.release(); cub_tmp.release();
    };
    struct Guard { decltype(release)& r; ~Guard() { r(); } } guard{release};
    auto up = [&](auto& buf, const auto* src, size_t count) -> cudaError_t {
        cudaError_t e = buf.alloc(count);
        if (e == cudaSuccess && count) e = cudaMemcpyAsync(buf.p, src, count * sizeof(*src), cudaMemcpyHostToDevice, c->stream);
        return e;
    };
    TRB_CUDA(c, up(d_pos, s.positions, 3 * s.n_positions)); TRB_CUDA(c, up(d_nrm, s.normals, 3 * s.n_normals));
    TRB_CUDA(c, up(d_tv, s.tri_v, 3 * size_t(n))); TRB_CUDA(c, up(d_tn, s.tri_n, 3 * size_t(n)));
    TRB_CUDA(c, up(d_tm, s.tri_mat, size_t(n))); TRB_CUDA(c, up(d_ts, s.tri_shape, size_t(n)));
    TRB_CUDA(c, up(d_sf, f.shape_first.data(), f.shape_first.size())); TRB_CUDA(c, up(d_se, f.shape_to_emitter.data(), f.shape_to_emitter.size()));
    if (f.textured) { TRB_CUDA(c, up(d_tc, s.texcoords, 2 * s.n_texcoords)); TRB_CUDA(c, up(d_tt, s.tri_t, 3 * size_t(n))); }
    TRB_CUDA(c, cudaMemcpyAsync(c->d_mats.p, f.materials.data(), f.materials.size() * sizeof(trb::DevMaterial), cudaMemcpyHostToDevice, c->stream));
    lap("raw arrays to the device");
    TRB_CUDA(c, ord_a.alloc(n)); TRB_CUDA(c, ord_b.alloc(n)); TRB_CUDA(c, range_of.alloc(n));
    TRB_CUDA(c, tmin.alloc(n)); TRB_CUDA(c, tmax.alloc(n)); TRB_CUDA(c, tcen.alloc(n));
    TRB_CUDA(c, is_l.alloc(size_t(n) + 1)); TRB_CUDA(c, pref.alloc(size_t(n) + 1)); TRB_CUDA(c, counter.alloc(1));
    TRB_CUDA(c, r_pos.alloc(size_t(n) + 1)); TRB_CUDA(c, last_r.alloc(size_t(n) + 1));
    const size_t cap = 2 * size_t(n) + 2;   // a leaf holds >= 1 triangle and the tree is binary
    TRB_CUDA(c, nodes.alloc(cap));
    size_t tmp_bytes = 0;
    size_t tmp_bytes2 = 0;
    TRB_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, is_l.p, pref.p, int(n + 1), c->stream));
    TRB_CUDA(c, cub::DeviceScan::InclusiveScan(nullptr, tmp_bytes2, r_pos.p, last_r.p, cub::Max(), int(n + 1), c->stream));
    tmp_bytes = std::max(tmp_bytes, tmp_bytes2);
    TRB_CUDA(c, cub_tmp.alloc(tmp_bytes));
    lap("scratch allocation");
    BuildScene bs{d_pos.p, d_nrm.p, d_tv.p, d_tn.p, d_tm.p, d_ts.p, d_sf.p, d_se.p, c->d_mats.p, d_tc.p, d_tt.p, n};
    const int B = 256, gN = int((size_t(n) + B - 1) / B), gN1 = int((size_t(n) + 1 + B - 1) / B);
    auto gK = [&](int32_t count) { return std::max(1, (count + B - 1) / B); };
    k_bb_prep<<<gN, B, 0, c->stream>>>(bs, tmin.p, tmax.p, tcen.p, ord_a.p, range_of.p);
    k_bb_root<<<1, 1, 0, c->stream>>>(nodes.p, n, counter.p);
    std::vector<std::pair<int32_t, int32_t>> levels;   // node id ranges in creation order
    int32_t lb = 0, le = 1;
    int32_t* ord = ord_a.p; int32_t* ord_next = ord_b.p;
    double kind_ms[6] = {0, 0, 0, 0, 0, 0};
    auto klap = [&](int kind) {   // TRB_BUILD_TIMING only: per-kernel-kind time summed over the levels (serialises the loop)
        if (!timing) return;
        cudaStreamSynchronize(c->stream);
        auto now = std::chrono::steady_clock::now();
        kind_ms[kind] += std::chrono::duration<double, std::milli>(now - t_prev).count();
        t_prev = now;
    };
    while (le > lb) {
        levels.push_back({lb, le});
        if (levels.size() > size_t(TRB_STACK)) return c->fail(TRB_ERR_UNSUPPORTED, "BVH deeper than the traversal stack");
        k_bb_bounds<<<gN, B, 0, c->stream>>>(nodes.p, lb, ord, range_of.p, tmin.p, tmax.p, tcen.p, n); klap(0);
        k_bb_split<<<gK(le - lb), B, 0, c->stream>>>(nodes.p, lb, le); klap(1);
        k_bb_flags<<<gN1, B, 0, c->stream>>>(nodes.p, lb, ord, range_of.p, tcen.p, is_l.p, r_pos.p, n); klap(2);
        TRB_CUDA(c, cub::DeviceScan::ExclusiveSum(cub_tmp.p, tmp_bytes, is_l.p, pref.p, int(n + 1), c->stream));
        TRB_CUDA(c, cub::DeviceScan::InclusiveScan(cub_tmp.p, tmp_bytes, r_pos.p, last_r.p, cub::Max(), int(n + 1), c->stream)); klap(3);
        k_bb_children<<<gK(le - lb), B, 0, c->stream>>>(nodes.p, lb, le, pref.p, counter.p); klap(4);
        k_bb_permute<<<gN, B, 0, c->stream>>>(nodes.p, lb, ord, ord_next, range_of.p, is_l.p, pref.p, last_r.p, n); klap(5);
        std::swap(ord, ord_next);
        uint32_t total = 0;
        TRB_CUDA(c, cudaMemcpyAsync(&total, counter.p, sizeof(total), cudaMemcpyDeviceToHost, c->stream));
        TRB_CUDA(c, cudaStreamSynchronize(c->stream));
        if (size_t(total) > cap) return c->fail(TRB_ERR_CUDA, "device BVH build: node capacity exceeded");
        lb = le; le = int32_t(total);
    }
    const int32_t total_nodes = le;
    if (timing) std::fprintf(stderr, "[trb build] levels: bounds %.1f split %.1f flags %.1f scan %.1f children %.1f permute %.1f ms\n", kind_ms[0], kind_ms[1],
                             kind_ms[2], kind_ms[3], kind_ms[4], kind_ms[5]);
    lap("levels (tail)");
    for (size_t l = levels.size(); l-- > 0;) k_bb_sizes<<<gK(levels[l].second - levels[l].first), B, 0, c->stream>>>(nodes.p, levels[l].first, levels[l].second);
    for (size_t l = 0; l < levels.size(); ++l) k_bb_number<<<gK(levels[l].second - levels[l].first), B, 0, c->stream>>>(nodes.p, levels[l].first, levels[l].second);
    BuildNode root;
    TRB_CUDA(c, cudaMemcpyAsync(&root, nodes.p, sizeof(root), cudaMemcpyDeviceToHost, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    f.n_nodes = size_t(root.sub_inner); f.n_tris = n; f.n_tri_uv = f.textured ? n : 0;
    f.root_ref = root.left >= 0 ? 0 : ~int32_t((0u << 3) | n);
    f.max_depth = int32_t(levels.size()) - 1;
    for (int k = 0; k < 3; ++k) { f.scene_min[k] = root.bmin[k]; f.scene_max[k] = root.bmax[k]; }
    TRB_CUDA(c, c->d_nodes.alloc(f.n_nodes)); TRB_CUDA(c, c->d_tris.alloc(n)); TRB_CUDA(c, c->d_shade.alloc(n)); TRB_CUDA(c, c->d_tri_uv.alloc(f.n_tri_uv));
    k_bb_emit_nodes<<<gK(total_nodes), B, 0, c->stream>>>(nodes.p, total_nodes, c->d_nodes.p);
    k_bb_emit_tris<<<gN, B, 0, c->stream>>>(bs, ord, c->d_tris.p, c->d_shade.p, f.textured ? c->d_tri_uv.p : nullptr);
    TRB_CUDA(c, cudaGetLastError());
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    lap("numbering + emit");
    f.n_wnodes = 0; f.wide_depth = 0; f.wnodes.clear();
    if (root.left >= 0 && !f.wide_collapse) {
        // the SAH upper tree: the leaf list comes down (32 B per leaf), the tree is built on the host cores (trb::build_sah_wide - the same
        // call the host build makes on the same list), the 128 B nodes go up.  f.wnodes stays on the host (staging for trb_scene_refresh).
        const size_t n_leaves = size_t(root.sub_nodes - root.sub_inner);
        DevBuf<trb::LeafBox> d_leaves;
        struct LGuard { DevBuf<trb::LeafBox>& b; ~LGuard() { b.release(); } } lguard{d_leaves};
        TRB_CUDA(c, d_leaves.alloc(n_leaves));
        k_bb_emit_leaves<<<gK(total_nodes), B, 0, c->stream>>>(nodes.p, total_nodes, d_leaves.p);
        std::vector<trb::LeafBox> leaves(n_leaves);
        TRB_CUDA(c, cudaMemcpyAsync(leaves.data(), d_leaves.p, n_leaves * sizeof(trb::LeafBox), cudaMemcpyDeviceToHost, c->stream));
        TRB_CUDA(c, cudaStreamSynchronize(c->stream));
        lap("leaf list to the host");
        trb::build_sah_wide(leaves.data(), n_leaves, f.wnodes, f.wide_depth);
        lap("SAH upper tree (host cores)");
        f.n_wnodes = f.wnodes.size();
        TRB_CUDA(c, c->d_wnodes.alloc(f.n_wnodes));
        TRB_CUDA(c, cudaMemcpyAsync(c->d_wnodes.p, f.wnodes.data(), f.n_wnodes * sizeof(trb::DevNode4), cudaMemcpyHostToDevice, c->stream));
        TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    } else if (root.left >= 0) {
        // TRB_UPLOAD_WIDE_COLLAPSE: the collapse of the reference's own tree (k_wide_*: the device twin of trb::build_wide_nodes), one level
        // per pass; scratch of the build is re-used
        DevBuf<trb::DevNode4> wtmp;
        struct WGuard { DevBuf<trb::DevNode4>& b; ~WGuard() { b.release(); } } wguard{wtmp};
        TRB_CUDA(c, wtmp.alloc(f.n_nodes));   // upper bound: every wide node takes at least one binary inner node
        int32_t* lvl = range_of.p; int32_t* lvl_next = ord_next;
        int4* slots = reinterpret_cast<int4*>(tmin.p);
        TRB_CUDA(c, cudaMemsetAsync(lvl, 0, sizeof(int32_t), c->stream));   // the root is node 0 in creation order
        uint32_t base = 0; int32_t n_level = 1;
        while (n_level > 0) {
            if (size_t(base) + size_t(n_level) > f.n_nodes) return c->fail(TRB_ERR_CUDA, "device BVH build: wide node capacity exceeded");
            k_wide_expand<<<gK(n_level + 1), B, 0, c->stream>>>(nodes.p, lvl, n_level, slots, is_l.p);
            TRB_CUDA(c, cub::DeviceScan::ExclusiveSum(cub_tmp.p, tmp_bytes, is_l.p, pref.p, n_level + 1, c->stream));
            k_wide_emit<<<gK(n_level), B, 0, c->stream>>>(nodes.p, n_level, slots, pref.p, base, wtmp.p, lvl_next);
            uint32_t next_n = 0;
            TRB_CUDA(c, cudaMemcpyAsync(&next_n, pref.p + n_level, sizeof(next_n), cudaMemcpyDeviceToHost, c->stream));
            TRB_CUDA(c, cudaStreamSynchronize(c->stream));
            base += uint32_t(n_level); std::swap(lvl, lvl_next); n_level = int32_t(next_n); ++f.wide_depth;
        }
        f.n_wnodes = base;
        TRB_CUDA(c, c->d_wnodes.alloc(base));
        TRB_CUDA(c, cudaMemcpyAsync(c->d_wnodes.p, wtmp.p, size_t(base) * sizeof(trb::DevNode4), cudaMemcpyDeviceToDevice, c->stream));
        TRB_CUDA(c, cudaGetLastError());
        TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    lap("wide tree");
    if (timing) std::fprintf(stderr, "[trb build] %d levels, %d nodes (%zu inner)\n", int(levels.size()), total_nodes, f.n_nodes);
    return TRB_OK;
}

}  // namespace

extern "C" {

int trb_scene_upload(trb_context* c, const trb_scene_desc* s) { return trb_scene_upload_ex(c, s, TRB_UPLOAD_AUTO); }

int trb_scene_upload_ex(trb_context* c, const trb_scene_desc* s, uint32_t flags) {
    if (!c || !s) return TRB_ERR_INVALID;
    cudaSetDevice(c->device);
    std::string err;
    trb::FlatScene& f = c->flat;
    c->has_scene = false;
    auto code = [&]() { return err.find("illum 5") != std::string::npos ? TRB_ERR_UNSUPPORTED : TRB_ERR_INVALID; };
    const auto t_up0 = std::chrono::steady_clock::now();
    if (!trb::flatten_tables(*s, f, err)) return c->fail(code(), err);
    if (std::getenv("TRB_BUILD_TIMING")) std::fprintf(stderr, "[trb build] %-28s %8.2f ms\n", "host tables + validation", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_up0).count());
    if (f.emitters.empty()) return c->fail(TRB_ERR_INVALID, "scene has no emitter (the reference divides by emitters.size())");
    if (s->n_tris == 0) return c->fail(TRB_ERR_INVALID, "empty scene");
    if (s->n_tris >= (size_t(1) << 27)) return c->fail(TRB_ERR_UNSUPPORTED, "more than 2^27 triangles");   // (owner lane << 27 | position) queue words
    f.wide_collapse = (flags & TRB_UPLOAD_WIDE_COLLAPSE) != 0;
    flags &= ~uint32_t(TRB_UPLOAD_WIDE_COLLAPSE);
    if (flags > TRB_UPLOAD_DEVICE_BUILD) return c->fail(TRB_ERR_INVALID, "trb_scene_upload_ex: unknown flags");
    const bool on_device = flags == TRB_UPLOAD_DEVICE_BUILD || (flags == TRB_UPLOAD_AUTO && s->n_tris >= TRB_DEVICE_BUILD_MIN_TRIS);
    TRB_CUDA(c, c->d_mats.alloc(f.materials.size()));
    TRB_CUDA(c, c->d_emitters.alloc(f.emitters.size())); TRB_CUDA(c, c->d_etris.alloc(f.emitter_tris.size()));
    TRB_CUDA(c, c->d_cdf.alloc(f.emitter_cdf.size())); TRB_CUDA(c, c->d_esphere.alloc(f.emitter_sphere.size() / 4));
    TRB_CUDA(c, c->d_tex_info.alloc(f.tex_info.size())); TRB_CUDA(c, c->d_texels.alloc(f.texels.size()));
    if (c->pinned) { cudaFreeHost(c->pinned); c->pinned = nullptr; }
    if (on_device) {
        int rc = build_geometry_on_device(c, *s);
        if (rc) return rc;
    } else {
        if (!trb::flatten_geometry(*s, f, err)) return c->fail(code(), err);
        TRB_CUDA(c, c->d_nodes.alloc(f.n_nodes)); TRB_CUDA(c, c->d_tris.alloc(f.n_tris));
        TRB_CUDA(c, c->d_shade.alloc(f.n_tris)); TRB_CUDA(c, c->d_tri_uv.alloc(f.n_tri_uv)); TRB_CUDA(c, c->d_wnodes.alloc(f.n_wnodes));
    }
    if (3 * f.wide_depth + 4 > 3 * TRB_STACK) return c->fail(TRB_ERR_UNSUPPORTED, "wide BVH deeper than the traversal stack");
    if (f.max_depth + 2 > TRB_STACK) return c->fail(TRB_ERR_UNSUPPORTED, "BVH deeper than the traversal stack");
    c->built_on_device = on_device;
    c->sc.nodes = reinterpret_cast<const F4*>(c->d_nodes.p);
    c->sc.tris = reinterpret_cast<const F4*>(c->d_tris.p);
    c->sc.shade = c->d_shade.p; c->sc.mats = c->d_mats.p; c->sc.emitters = c->d_emitters.p; c->sc.etris = c->d_etris.p;
    c->sc.tri_uv = reinterpret_cast<const float4*>(c->d_tri_uv.p); c->sc.tex_info = reinterpret_cast<const int4*>(c->d_tex_info.p);
    c->sc.texels = c->d_texels.p; c->sc.has_tex = f.tex_info.empty() ? 0 : 1;
    c->sc.cdf = c->d_cdf.p; c->sc.esphere = c->d_esphere.p; c->sc.root_ref = f.root_ref; c->sc.n_emitters = int32_t(f.emitters.size());
    {   // near_cull (kernels/trace.cuh) is tied to the size of the scene's coordinates; TRB_NEAR_CULL=0 at upload keeps the old rule (A/B runs)
        float amax = 0.f;
        for (int k = 0; k < 3; ++k) amax = std::max(amax, std::max(std::fabs(f.scene_min[k]), std::fabs(f.scene_max[k])));
        const char* e = std::getenv("TRB_NEAR_CULL");
        c->near_inv_max = (e && std::string(e) == "0") ? 0.f : (amax > 0.f && std::isfinite(amax) ? 400.f / amax : 0.f);
    }
    c->has_scene = true;
    if (on_device) {
        // the tables go up directly; the pinned staging blob (geometry read back from the device) is made by the first trb_scene_refresh
        size_t sizes[TRB_SCENE_ARRAYS];
        scene_array_sizes(f, sizes);
        void* dsts[TRB_SCENE_ARRAYS] = {nullptr, nullptr, nullptr, c->d_mats.p, c->d_emitters.p, c->d_etris.p, c->d_cdf.p, c->d_esphere.p, c->d_tex_info.p, nullptr, c->d_texels.p, nullptr};
        const void* srcs[TRB_SCENE_ARRAYS] = {nullptr, nullptr, nullptr, f.materials.data(), f.emitters.data(), f.emitter_tris.data(), f.emitter_cdf.data(),
                                              f.emitter_sphere.data(), f.tex_info.data(), nullptr, f.texels.data(), nullptr};
        for (int k = 0; k < TRB_SCENE_ARRAYS; ++k)
            if (dsts[k] && sizes[k]) TRB_CUDA(c, cudaMemcpyAsync(dsts[k], srcs[k], sizes[k], cudaMemcpyHostToDevice, c->stream));
        TRB_CUDA(c, cudaStreamSynchronize(c->stream));
        return TRB_OK;
    }
    int rc = stage_scene(c);
    if (rc) return rc;
    return trb_scene_refresh(c);
}

int trb_scene_refresh(trb_context* c) {
    if (!c) return TRB_ERR_INVALID;
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    cudaSetDevice(c->device);
    if (!c->pinned) { int rc = stage_scene(c); if (rc) return rc; }
    void* dsts[TRB_SCENE_ARRAYS] = {c->d_nodes.p, c->d_tris.p, c->d_shade.p, c->d_mats.p, c->d_emitters.p, c->d_etris.p, c->d_cdf.p, c->d_esphere.p,
                                    c->d_tex_info.p, c->d_tri_uv.p, c->d_texels.p, c->d_wnodes.p};
    size_t sizes[TRB_SCENE_ARRAYS];
    scene_array_sizes(c->flat, sizes);
    size_t off = 0;
    for (int k = 0; k < TRB_SCENE_ARRAYS; ++k) {
        if (sizes[k]) TRB_CUDA(c, cudaMemcpyAsync(dsts[k], c->pinned + off, sizes[k], cudaMemcpyHostToDevice, c->stream));
        off += (sizes[k] + 255) & ~size_t(255);
    }
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

/* Test hook: the device scene's geometry arrays read back (nodes, triangles, shading records) so that a device-built scene can be
 * compared with a host-built one. which: 0 binary nodes (64 B), 1 triangles (48 B), 2 shading records (96 B), 3 texture coordinates (32 B), 4 wide nodes (128 B). */
int trb_scene_download(trb_context* c, int which, void* out, size_t capacity_bytes, size_t* count_out) {
    if (!c) return TRB_ERR_INVALID;
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    cudaSetDevice(c->device);
    const trb::FlatScene& f = c->flat;
    const void* src[5] = {c->d_nodes.p, c->d_tris.p, c->d_shade.p, c->d_tri_uv.p, c->d_wnodes.p};
    const size_t cnt[5] = {f.n_nodes, f.n_tris, f.n_tris, f.n_tri_uv, f.n_wnodes};
    const size_t rec[5] = {sizeof(trb::DevNode), sizeof(trb::DevTri), sizeof(trb::DevShade), sizeof(trb::DevTriUV), sizeof(trb::DevNode4)};
    if (which < 0 || which > 4) return c->fail(TRB_ERR_INVALID, "trb_scene_download: which must be 0..4");
    if (count_out) *count_out = cnt[which];
    if (!out) return TRB_OK;
    if (capacity_bytes < cnt[which] * rec[which]) return c->fail(TRB_ERR_INVALID, "trb_scene_download: buffer too small");
    if (cnt[which]) TRB_CUDA(c, cudaMemcpyAsync(out, src[which], cnt[which] * rec[which], cudaMemcpyDeviceToHost, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_scene_built_on_device(const trb_context* c) { return (c && c->has_scene && c->built_on_device) ? 1 : 0; }

size_t trb_scene_device_bytes(const trb_context* c) { return (c && c->has_scene) ? c->flat.bytes() : 0; }

}  // extern "C"

// ---- wave runner -------------------------------------------------------------------------------------------
namespace {

struct WaveRun {
    const trb_render_desc* desc;
    float* d_accum;            // frame buffer or parity accumulation buffer
    int32_t* d_splat_pix;      // parity only
    const RayT* d_inject_rays; // parity only
    const float* d_inject_rng;
    int inject_stride;
    bool parity;
};

int run_wave(trb_context* c, const WaveRun& w, const trb::CameraDev& cam, int64_t wave_base, int P, int spp_begin, int spp_count) {
    const trb_render_desc& d = *w.desc;
    const int K = max_vertices_of(d);
    const int kk = std::max(K, 1);
    const int sm = c->prop.multiProcessorCount;
    WaveParams p;
    std::memset(&p, 0, sizeof(p));
    p.sc = c->sc; p.cam = cam;
    p.P = P; p.K = kk; p.rr_depth = d.rr_depth; p.rr_on = rr_enabled(d) ? 1 : 0; p.max_depth = max_depth_of(d); p.spp = d.spp; p.spp_begin = spp_begin; p.spp_count = spp_count;
    p.wave_base = wave_base; p.seed = d.seed; p.flags = d.flags;
    p.inject_rays = w.d_inject_rays; p.inject_rng = w.d_inject_rng; p.inject_stride = w.inject_stride; p.parity = w.parity ? 1 : 0;
    p.accum = w.d_accum; p.splat_pix = w.d_splat_pix;
    p.cam_ray = c->cam_ray.p; p.ray = c->ray.p; p.hit = c->hit.p; p.hit0 = c->hit0.p; p.st_t = c->st_t.p; p.st_v = c->st_v.p;
    p.lv = c->lv.p; p.nlv = c->nlv.p; p.ev = c->ev.p; p.lv_tex = c->lv_tex.p; p.ev_tex = c->ev_tex.p; p.q_conn = c->q_conn.p; p.sray = c->sray.p; p.spend = c->spend.p;
    p.shadow_cap = uint32_t(std::min<size_t>(c->sray.n, 0xffffffffu));
    const int n_stage = 2 * kk + 4;
    TRB_CUDA(c, cudaMemsetAsync(c->counters.p, 0, size_t(n_stage) * CTR_N * sizeof(uint32_t), c->stream));
    unsigned long long* visit = nullptr;
#ifdef TRB_COUNTERS
    visit = c->d_stats.p + 2;
#endif
    int32_t* qa = c->q_a.p; int32_t* qb = c->q_b.p;
    // the two sets of queue-ordered walk records: a stage reads the set its predecessor wrote
    auto use_sets = [&](bool in_a) {
        p.q_in = in_a ? qa : qb;                       p.q_out = in_a ? qb : qa;
        p.ray_in = in_a ? c->ray.p : c->ray_b.p;       p.ray_out = in_a ? c->ray_b.p : c->ray.p;
        p.stt_in = in_a ? c->st_t.p : c->st_t_b.p;     p.stt_out = in_a ? c->st_t_b.p : c->st_t.p;
        p.sv_in = in_a ? c->sv_a.p : c->sv_b.p;        p.sv_out = in_a ? c->sv_b.p : c->sv_a.p;
    };
    auto stage_ctr = [&](int s) { return c->counters.p + size_t(s) * CTR_N; };
    const int gP = grid_exact(size_t(P));
    const int gB = int((size_t(P) + TRB_SORT_BLOCK - 1) / TRB_SORT_BLOCK);

    // Shadow rays of stage d are traced on the second stream while the first goes on with the closest-hit traversal
    // and the bounce of stage d+1 (the tail of one persistent kernel is filled by the other).  Two shadow queues
    // alternate; a bounce may refill a queue only after the traversal that read it has finished.
    const size_t n_dep = size_t(n_stage) + 2;
    while (c->dep_a.size() < n_dep) { cudaEvent_t e; cudaEventCreateWithFlags(&e, cudaEventDisableTiming); c->dep_a.push_back(e); }
    while (c->dep_b.size() < n_dep) { cudaEvent_t e; cudaEventCreateWithFlags(&e, cudaEventDisableTiming); c->dep_b.push_back(e); }
    int n_any = 0;                       // shadow launches issued so far in this wave
    int last_any_of_buf[2] = {-1, -1};   // index (into dep_b) of the last traversal that read each queue
    auto shadow_buffers = [&]() {        // before a bounce kernel: pick the queue and wait until it is free again
        const int b = n_any & 1;
        p.sray = b ? c->sray2.p : c->sray.p; p.spend = b ? c->spend2.p : c->spend.p;
        if (last_any_of_buf[b] >= 0) cudaStreamWaitEvent(c->stream, c->dep_b[last_any_of_buf[b]], 0);
    };
    const bool overlap = (d.flags & TRB_FLAG_NO_OVERLAP) == 0;
    auto closest = [&](const RayT* rays, const int32_t* q, const uint32_t* n_ptr, uint32_t n_fixed, uint32_t* fetch) {
        launch_closest(c, rays, q, n_ptr, n_fixed, p.hit, fetch, visit); c->mark(LK_CLOSEST);
    };
    auto trace_shadow = [&]() {          // after the bounce kernel(s) of a stage
        const int b = n_any & 1;
        if (!overlap) {
            launch_any(c, p.sray, p.counters + CTR_SHADOW, 0u, p.shadow_cap, nullptr, p.spend, p.accum, p.counters + CTR_FETCH_ANY,
                       visit ? visit + 2 : nullptr, c->stream);
            c->mark(LK_ANY);
            ++n_any;
            return;
        }
        cudaEventRecord(c->dep_a[n_any], c->stream);
        cudaStreamWaitEvent(c->stream2, c->dep_a[n_any], 0);
        c->mark(-1, true);
        launch_any(c, p.sray, p.counters + CTR_SHADOW, 0u, p.shadow_cap, nullptr, p.spend, p.accum, p.counters + CTR_FETCH_ANY,
                   visit ? visit + 2 : nullptr, c->stream2);
        c->mark(LK_ANY, true);
        cudaEventRecord(c->dep_b[n_any], c->stream2);
        last_any_of_buf[b] = n_any;
        ++n_any;
    };

    k_gen_camera<<<gP, TRB_BLOCK, 0, c->stream>>>(p); c->mark(LK_LOGIC);
    launch_closest(c, p.cam_ray, nullptr, nullptr, uint32_t(P), p.hit, stage_ctr(2 * kk + 2) + CTR_FETCH_CLOSEST, visit); c->mark(LK_CLOSEST);  // (camera rays are coherent as generated)
    // light subpath
    int stage = 0;
    p.counters = stage_ctr(stage); use_sets(false);   // writes set A
    k_light_init<<<gP, TRB_BLOCK, 0, c->stream>>>(p); c->mark(LK_LOGIC);
    const bool light_only = (d.flags & TRB_FLAG_LIGHT_TRACING) != 0, eye_only = (d.flags & TRB_FLAG_PATH_TRACING) != 0;   // bdpt.h:16-17, 229-237
    for (int depth = 1; depth <= (eye_only ? 0 : K); ++depth) {
        const uint32_t* n_in = stage_ctr(stage) + CTR_QOUT;
        use_sets(stage % 2 == 0);
        ++stage;
        p.counters = stage_ctr(stage);
        closest(p.ray_in, nullptr, n_in, 0u, p.counters + CTR_FETCH_CLOSEST);   // rays and hits by queue position
        shadow_buffers();
        k_light_bounce<<<gB, TRB_SORT_BLOCK, 0, c->stream>>>(p, depth, n_in); c->mark(LK_LOGIC);
        trace_shadow();
    }
    // eye subpath
    stage = kk + 1;
    p.counters = stage_ctr(stage); use_sets(false);
    k_eye_init<<<gP, TRB_BLOCK, 0, c->stream>>>(p); c->mark(LK_LOGIC);
    for (int depth = 1; depth <= (light_only ? 1 : K); ++depth) {   // (LIGHT_TRACING: one pass that adds the primary hit's emission)
        const uint32_t* n_in = stage_ctr(stage) + CTR_QOUT;
        const bool even = ((stage - (kk + 1)) % 2 == 0);
        use_sets(even);
        ++stage;
        p.counters = stage_ctr(stage);
        if (depth > 1) closest(p.ray_in, nullptr, n_in, 0u, p.counters + CTR_FETCH_CLOSEST);
        shadow_buffers();
        k_eye_bounce<<<gB, TRB_SORT_BLOCK, 0, c->stream>>>(p, depth, n_in);
        if (c->sc.has_tex) k_connect_vertices<true><<<grid_for(size_t(P) * size_t(kk), sm), TRB_BLOCK, 0, c->stream>>>(p, p.counters + CTR_CONN);
        else k_connect_vertices<false><<<grid_for(size_t(P) * size_t(kk), sm), TRB_BLOCK, 0, c->stream>>>(p, p.counters + CTR_CONN);
        c->mark(LK_LOGIC);
        ++c->launches;
        trace_shadow();
    }
    for (int b = 0; b < 2; ++b)  // join: the wave is complete on the first stream
        if (last_any_of_buf[b] >= 0) cudaStreamWaitEvent(c->stream, c->dep_b[last_any_of_buf[b]], 0);
    k_fold_stats<<<1, 32, 0, c->stream>>>(c->counters.p, n_stage, kk, kk, uint32_t(P), c->d_stats.p, p.shadow_cap); c->mark(LK_LOGIC);
    TRB_CUDA(c, cudaGetLastError());
    return TRB_OK;
}

int check_desc(trb_context* c, const trb_render_desc* d) {
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    if (!d || d->width <= 0 || d->height <= 0 || d->spp <= 0 || d->rr_depth < 1) return c->fail(TRB_ERR_INVALID, "bad render description");
    if (d->rr_depth > 64) return c->fail(TRB_ERR_INVALID, "rr_depth > 64 is not supported");
    if ((d->flags & TRB_FLAG_LIGHT_TRACING) && (d->flags & TRB_FLAG_PATH_TRACING))
        return c->fail(TRB_ERR_INVALID, "TRB_FLAG_LIGHT_TRACING and TRB_FLAG_PATH_TRACING exclude each other");
    if (rr_enabled(*d) && (max_depth_of(*d) < d->rr_depth || max_depth_of(*d) > 64))
        return c->fail(TRB_ERR_INVALID, "Russian roulette: max_depth must lie in [rr_depth, 64]");
    return TRB_OK;
}

void collect_stats(trb_context* c, trb_stats* st, uint64_t samples) {
    if (!st) return;
    std::memset(st, 0, sizeof(*st));
    unsigned long long h[TRB_NSTATS] = {0};
    cudaMemcpyAsync(h, c->d_stats.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream);
    cudaStreamSynchronize(c->stream);
    st->samples = samples; st->rays_closest = h[0]; st->rays_any = h[1]; st->node_visits = h[2]; st->tri_tests = h[3]; st->node_visits_any = h[4]; st->tri_tests_any = h[5]; st->retraces = h[6]; st->retraces_any = h[7];
    st->node_visits_bin = h[8]; st->tri_tests_bin = h[9]; st->node_visits_bin_any = h[10]; st->tri_tests_bin_any = h[11]; st->rays_deferred = h[12]; st->rays_deferred_any = h[13];
    st->launches = c->launches; st->wave_samples = c->wave_used;
    double ms[LK_N] = {0, 0, 0};
    // kernel time summed per kind on each stream; the two streams overlap, so the kinds can add up to more than ms_total
    for (size_t i = 1; i < c->n_events; ++i) {
        float t = 0.f;
        cudaEventElapsedTime(&t, c->events[i - 1], c->events[i]);
        ms[c->event_kind[i]] += t;
    }
    for (size_t i = 1; i < c->n_events2; ++i) {
        if (c->event_kind2[i] < 0) continue;  // a begin marker
        float t = 0.f;
        cudaEventElapsedTime(&t, c->events2[i - 1], c->events2[i]);
        ms[c->event_kind2[i]] += t;
    }
    if (c->n_events >= 2) { float t = 0.f; cudaEventElapsedTime(&t, c->events[0], c->events[c->n_events - 1]); st->ms_total = t; }
    st->ms_trace_closest = ms[LK_CLOSEST]; st->ms_trace_any = ms[LK_ANY]; st->ms_logic = ms[LK_LOGIC];
}

int render_impl(trb_context* c, const trb_render_desc* d, int accumulate, float* d_rgb, trb_stats* stats) {
    int rc = check_desc(c, d);
    if (rc) return rc;
    cudaSetDevice(c->device);
    const int64_t npix = int64_t(d->width) * d->height;
    const int spp_begin = d->spp_begin;
    const int spp_end = d->spp_end > 0 ? d->spp_end : d->spp;
    if (spp_begin < 0 || spp_end > d->spp || spp_begin > spp_end) return c->fail(TRB_ERR_INVALID, "bad sample range");
    const int spp_count = spp_end - spp_begin;
    const int K = std::max(max_vertices_of(*d), 1);
    int64_t P = d->wave_samples > 0 ? d->wave_samples : default_wave_samples(c, K);
    const int64_t total = npix * spp_count;
    P = std::max<int64_t>(1, std::min(P, std::max<int64_t>(total, 1)));
    if (P > (int64_t(1) << 25)) return c->fail(TRB_ERR_INVALID, "wave_samples must not exceed 2^25");  // (slot << 6 | vertex) pair words
    rc = ensure_wave(c, size_t(P), K);
    if (rc) return rc;
    c->wave_used = uint64_t(P);
    trb::CameraDev cam;
    trb::make_camera(*d, cam);
    c->timing = stats != nullptr; c->n_events = 0; c->n_events2 = 0; c->launches = 0;
    TRB_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, TRB_NSTATS * sizeof(unsigned long long), c->stream));
    c->mark(LK_LOGIC);
    if (!accumulate) TRB_CUDA(c, cudaMemsetAsync(d_rgb, 0, size_t(npix) * 3 * sizeof(float), c->stream));
    WaveRun w{d, d_rgb, nullptr, nullptr, nullptr, 0, false};
    for (int64_t base = 0; base < total; base += P) {
        int n = int(std::min<int64_t>(P, total - base));
        rc = run_wave(c, w, cam, base, n, spp_begin, spp_count);
        if (rc) return rc;
    }
    c->mark(LK_LOGIC);
    if (stats) collect_stats(c, stats, uint64_t(total));
    return TRB_OK;
}


// ---- PathTracerIntegrator on the same wave buffers (kernels/pt_kernels.cuh; path.h:15-262) ------------------------------
__global__ void k_read_draws(const float4* __restrict__ st_v, int n, int32_t* __restrict__ out);  // defined with the parity entry points

// rays traced by a path-tracer wave = P primary rays + the sizes of all traced queues
__global__ void k_pt_fold_stats(const uint32_t* __restrict__ ctr, uint32_t n_ray, uint32_t P, unsigned long long* __restrict__ stats) {
    __shared__ unsigned long long part[256];
    unsigned long long s = 0;
    for (uint32_t i = threadIdx.x; i < n_ray; i += blockDim.x) s += ctr[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = P;
        for (int i = 0; i < 256; ++i) t += part[i];
        stats[0] += t;
    }
}

int check_path_desc(trb_context* c, const trb_render_desc* d, const trb_path_desc* pd, PtParams* out) {
    int rc = check_desc(c, d);
    if (rc) return rc;
    if (!pd) return c->fail(TRB_ERR_INVALID, "no path description");
    if (pd->emitter_samples < 0 || pd->emitter_samples > 16 || pd->bsdf_samples < 0 || pd->bsdf_samples > 16)
        return c->fail(TRB_ERR_INVALID, "emitter_samples / bsdf_samples must lie in [0, 16]");
    if (pd->max_depth < -1 || pd->max_depth > 256) return c->fail(TRB_ERR_INVALID, "path max_depth must lie in [-1, 256]");
    const int cap = d->max_depth > 0 ? d->max_depth : TRB_PT_DEFAULT_DEPTH_CAP;
    if (cap > 256) return c->fail(TRB_ERR_INVALID, "depth cap (trb_render_desc.max_depth) must be <= 256");
    out->is_explicit = pd->is_explicit ? 1 : 0; out->max_depth = pd->max_depth; out->rr_depth = d->rr_depth; out->rr_prob = pd->rr_prob;
    out->emitter_samples = pd->emitter_samples; out->bsdf_samples = pd->bsdf_samples;
    // stack levels: a bounded recursion (maxDepth >= 0) enters maxDepth calls and makes one more that only draws its roulette
    // number (path.h:70-74), the roulette-driven one gets the cap
    out->cap = pd->max_depth >= 0 ? pd->max_depth + 1 : cap;
    return TRB_OK;
}

int run_pt_wave(trb_context* c, const WaveRun& w, const PtParams& pt, const trb::CameraDev& cam, int64_t wave_base, int P, int spp_begin, int spp_count) {
    const trb_render_desc& d = *w.desc;
    WaveParams p;
    std::memset(&p, 0, sizeof(p));
    p.sc = c->sc; p.cam = cam;
    p.P = P; p.K = pt.cap; p.rr_depth = d.rr_depth; p.max_depth = pt.cap; p.spp = d.spp; p.spp_begin = spp_begin; p.spp_count = spp_count;
    p.wave_base = wave_base; p.seed = d.seed; p.flags = d.flags;
    p.inject_rays = w.d_inject_rays; p.inject_rng = w.d_inject_rng; p.inject_stride = w.inject_stride; p.parity = w.parity ? 1 : 0;
    p.accum = w.d_accum;
    p.cam_ray = c->cam_ray.p; p.ray = c->ray.p; p.hit = c->hit.p; p.hit0 = c->hit0.p; p.st_t = c->st_t.p; p.st_v = c->st_v.p;
    p.lv = c->lv.p; p.nlv = c->nlv.p; p.ev = c->ev.p; p.lv_tex = c->lv_tex.p; p.ev_tex = c->ev_tex.p;
    const int D = pt.cap, M = pt.emitter_samples, NB = pt.bsdf_samples;
    // counters: [0, n_ray) sizes of the traced queues (their sum + P = rays traced), then the other queues, then one fetch cursor per traversal launch
    const size_t n_ray = size_t(D) * size_t(1 + M + NB) + size_t(pt.max_depth >= 0 ? pt.max_depth : 0) + 8;
    const size_t n_other = size_t(D) + 8, n_fetch = TRACE_CTR_WORDS * (n_ray + 8);
    TRB_CUDA(c, c->pt_ctr.alloc(n_ray + n_other + n_fetch));
    TRB_CUDA(c, cudaMemsetAsync(c->pt_ctr.p, 0, (n_ray + n_other + n_fetch) * sizeof(uint32_t), c->stream));
    size_t i_ray = 0, i_other = n_ray, i_fetch = n_ray + n_other;
    auto ray_ctr = [&]() { return c->pt_ctr.p + i_ray++; };
    auto other_ctr = [&]() { return c->pt_ctr.p + i_other++; };
    auto fetch_ctr = [&]() { uint32_t* f = c->pt_ctr.p + i_fetch; i_fetch += TRACE_CTR_WORDS; return f; };
    unsigned long long* visit = nullptr;
#ifdef TRB_COUNTERS
    visit = c->d_stats.p + 2;
#endif
    const int gP = grid_exact(size_t(P));
    k_gen_camera<<<gP, TRB_BLOCK, 0, c->stream>>>(p); c->mark(LK_LOGIC);
    launch_closest(c, p.cam_ray, nullptr, nullptr, uint32_t(P), p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
    int32_t* q_cur = c->q_a.p; int32_t* q_next = c->q_b.p;
    uint32_t* n_cur = other_ctr();
    k_pt_init<<<gP, TRB_BLOCK, 0, c->stream>>>(p, q_cur, n_cur); c->mark(LK_LOGIC);
    if (!pt.is_explicit) {
        for (int depth = 0; depth <= std::max(pt.max_depth, 0); ++depth) {
            uint32_t* n_out = ray_ctr();
            k_pt_implicit<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, q_cur, n_cur, q_next, n_out); c->mark(LK_LOGIC);
            launch_closest(c, p.ray, q_next, n_out, 0u, p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
            std::swap(q_cur, q_next); n_cur = n_out;
        }
    } else {
        // forward: descend, one stack level per depth
        for (int depth = 0; depth < D; ++depth) {
            uint32_t* n_try = ray_ctr();
            uint32_t* n_next = other_ctr();
            k_pt_begin<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, q_cur, n_cur, c->q_conn.p, n_try); c->mark(LK_LOGIC);
            launch_closest(c, p.ray, c->q_conn.p, n_try, 0u, p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
            k_pt_check<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, c->q_conn.p, n_try, q_next, n_next, c->d_stats.p); c->mark(LK_LOGIC);
            std::swap(q_cur, q_next); n_cur = n_next;
        }
        // backward: direct illumination and the combination, deepest level first
        for (int depth = D - 1; depth >= 0; --depth) {
            for (int e = 0; e < M; ++e) {
                uint32_t* n_try = ray_ctr();
                k_pt_emitter_gen<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, e, c->q_conn.p, n_try); c->mark(LK_LOGIC);
                launch_closest(c, p.ray, c->q_conn.p, n_try, 0u, p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
                k_pt_emitter_resolve<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, c->q_conn.p, n_try); c->mark(LK_LOGIC);
            }
            for (int b = 0; b < NB; ++b) {
                uint32_t* n_try = ray_ctr();
                k_pt_bsdf_gen<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, c->q_conn.p, n_try); c->mark(LK_LOGIC);
                launch_closest(c, p.ray, c->q_conn.p, n_try, 0u, p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
                k_pt_bsdf_resolve<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth, c->q_conn.p, n_try); c->mark(LK_LOGIC);
            }
            k_pt_combine<<<gP, TRB_BLOCK, 0, c->stream>>>(p, pt, depth); c->mark(LK_LOGIC);
        }
    }
    k_pt_fold_stats<<<1, 256, 0, c->stream>>>(c->pt_ctr.p, uint32_t(i_ray), uint32_t(P), c->d_stats.p); c->mark(LK_LOGIC);
    TRB_CUDA(c, cudaGetLastError());
    return TRB_OK;
}

int render_path_impl(trb_context* c, const trb_render_desc* d, const trb_path_desc* pd, int accumulate, float* d_rgb, trb_stats* stats) {
    PtParams pt;
    int rc = check_path_desc(c, d, pd, &pt);
    if (rc) return rc;
    cudaSetDevice(c->device);
    const int64_t npix = int64_t(d->width) * d->height;
    const int spp_begin = d->spp_begin;
    const int spp_end = d->spp_end > 0 ? d->spp_end : d->spp;
    if (spp_begin < 0 || spp_end > d->spp || spp_begin > spp_end) return c->fail(TRB_ERR_INVALID, "bad sample range");
    const int spp_count = spp_end - spp_begin;
    // (B200, profiles/r2_node_phase_experiments.md r7: 2^21 / 2^22 / 2^23 samples per wave -> 64.8 / 89.1 / 111.6 Msamples/s on the caustic scene: a path's
    // 64 depth levels are 64 launches per stage whatever the wave holds, and deep levels hold few rays)
    int64_t P = d->wave_samples > 0 ? d->wave_samples : default_wave_samples(c, pt.is_explicit ? pt.cap : 1, /*shadow=*/false);
    const int64_t total = npix * spp_count;
    P = std::max<int64_t>(1, std::min(P, std::max<int64_t>(total, 1)));
    rc = ensure_wave(c, size_t(P), pt.is_explicit ? pt.cap : 1, /*shadow=*/false);
    if (rc) return rc;
    c->wave_used = uint64_t(P);
    trb::CameraDev cam;
    trb::make_camera(*d, cam);
    c->timing = stats != nullptr; c->n_events = 0; c->n_events2 = 0; c->launches = 0;
    TRB_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, TRB_NSTATS * sizeof(unsigned long long), c->stream));
    c->mark(LK_LOGIC);
    if (!accumulate) TRB_CUDA(c, cudaMemsetAsync(d_rgb, 0, size_t(npix) * 3 * sizeof(float), c->stream));
    WaveRun w{d, d_rgb, nullptr, nullptr, nullptr, 0, false};
    for (int64_t base = 0; base < total; base += P) {
        int n = int(std::min<int64_t>(P, total - base));
        rc = run_pt_wave(c, w, pt, cam, base, n, spp_begin, spp_count);
        if (rc) return rc;
    }
    c->mark(LK_LOGIC);
    if (stats) collect_stats(c, stats, uint64_t(total));
    return TRB_OK;
}

// ---- DirectIntegrator on the same wave buffers (kernels/di_kernels.cuh; direct.h:74-467) ----------------------------------
int check_direct_desc(trb_context* c, const trb_render_desc* d, const trb_direct_desc* dd, DiParams* out) {
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    if (!d || d->width <= 0 || d->height <= 0 || d->spp <= 0) return c->fail(TRB_ERR_INVALID, "bad render description");
    if (!dd) return c->fail(TRB_ERR_INVALID, "no direct-integrator description");
    if (dd->strategy < TRB_DI_MIS || dd->strategy > TRB_DI_BSDF) return c->fail(TRB_ERR_INVALID, "Error: wrong strategy");  // direct.h:460
    if (dd->emitter_samples < 0 || dd->emitter_samples > 64 || dd->bsdf_samples < 0 || dd->bsdf_samples > 64)
        return c->fail(TRB_ERR_INVALID, "emitter_samples / bsdf_samples must lie in [0, 64]");
    const bool emitter_loop = dd->strategy == TRB_DI_AREA || dd->strategy == TRB_DI_SOLID_ANGLE || dd->strategy == TRB_DI_COSINE_HEMISPHERE;
    if ((emitter_loop && dd->emitter_samples == 0) || (dd->strategy == TRB_DI_BSDF && dd->bsdf_samples == 0))
        return c->fail(TRB_ERR_INVALID, "this strategy divides by its sample count (direct.h:192,229,263,332): it must be >= 1");
    out->strategy = dd->strategy; out->emitter_samples = dd->emitter_samples; out->bsdf_samples = dd->bsdf_samples;
    return TRB_OK;
}

int run_di_wave(trb_context* c, const WaveRun& w, const DiParams& dp, const trb::CameraDev& cam, int64_t wave_base, int P, int spp_begin, int spp_count) {
    const trb_render_desc& d = *w.desc;
    WaveParams p;
    std::memset(&p, 0, sizeof(p));
    p.sc = c->sc; p.cam = cam;
    p.P = P; p.K = 1; p.rr_depth = d.rr_depth; p.max_depth = 1; p.spp = d.spp; p.spp_begin = spp_begin; p.spp_count = spp_count;
    p.wave_base = wave_base; p.seed = d.seed; p.flags = d.flags;
    p.inject_rays = w.d_inject_rays; p.inject_rng = w.d_inject_rng; p.inject_stride = w.inject_stride; p.parity = w.parity ? 1 : 0;
    p.accum = w.d_accum;
    p.cam_ray = c->cam_ray.p; p.ray = c->ray.p; p.hit = c->hit.p; p.hit0 = c->hit0.p; p.st_t = c->st_t.p; p.st_v = c->st_v.p;
    p.lv = c->lv.p; p.nlv = c->nlv.p; p.ev = c->ev.p; p.lv_tex = c->lv_tex.p; p.ev_tex = c->ev_tex.p;
    const int M = dp.strategy == DI_BSDF ? 0 : dp.emitter_samples;
    const int NB = (dp.strategy == DI_MIS || dp.strategy == DI_BSDF) ? dp.bsdf_samples : 0;
    // counters: [0, n_ray) sizes of the traced queues (their sum + P = rays traced), then one fetch cursor per traversal launch
    const size_t n_ray = size_t(M + NB) + 8, n_fetch = TRACE_CTR_WORDS * (n_ray + 8);
    TRB_CUDA(c, c->pt_ctr.alloc(n_ray + n_fetch));
    TRB_CUDA(c, cudaMemsetAsync(c->pt_ctr.p, 0, (n_ray + n_fetch) * sizeof(uint32_t), c->stream));
    size_t i_ray = 0, i_fetch = n_ray;
    unsigned long long* visit = nullptr;
#ifdef TRB_COUNTERS
    visit = c->d_stats.p + 2;
#endif
    const int gP = grid_exact(size_t(P));
    k_gen_camera<<<gP, TRB_BLOCK, 0, c->stream>>>(p); c->mark(LK_LOGIC);
    auto fetch_ctr = [&]() { uint32_t* f = c->pt_ctr.p + i_fetch; i_fetch += TRACE_CTR_WORDS; return f; };
    launch_closest(c, p.cam_ray, nullptr, nullptr, uint32_t(P), p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
    k_di_init<<<gP, TRB_BLOCK, 0, c->stream>>>(p); c->mark(LK_LOGIC);
    for (int phase = 0; phase < 2; ++phase)
        for (int i = 0; i < (phase == 0 ? M : NB); ++i) {
            uint32_t* n_try = c->pt_ctr.p + i_ray++;
            k_di_gen<<<gP, TRB_BLOCK, 0, c->stream>>>(p, dp, phase, c->q_conn.p, n_try); c->mark(LK_LOGIC);
            launch_closest(c, p.ray, c->q_conn.p, n_try, 0u, p.hit, fetch_ctr(), visit); c->mark(LK_CLOSEST);
            k_di_resolve<<<gP, TRB_BLOCK, 0, c->stream>>>(p, dp, phase, c->q_conn.p, n_try); c->mark(LK_LOGIC);
        }
    k_di_combine<<<gP, TRB_BLOCK, 0, c->stream>>>(p, dp); c->mark(LK_LOGIC);
    k_pt_fold_stats<<<1, 256, 0, c->stream>>>(c->pt_ctr.p, uint32_t(i_ray), uint32_t(P), c->d_stats.p); c->mark(LK_LOGIC);
    TRB_CUDA(c, cudaGetLastError());
    return TRB_OK;
}

int render_direct_impl(trb_context* c, const trb_render_desc* d, const trb_direct_desc* dd, int accumulate, float* d_rgb, trb_stats* stats) {
    DiParams dp;
    int rc = check_direct_desc(c, d, dd, &dp);
    if (rc) return rc;
    cudaSetDevice(c->device);
    const int64_t npix = int64_t(d->width) * d->height;
    const int spp_begin = d->spp_begin;
    const int spp_end = d->spp_end > 0 ? d->spp_end : d->spp;
    if (spp_begin < 0 || spp_end > d->spp || spp_begin > spp_end) return c->fail(TRB_ERR_INVALID, "bad sample range");
    const int spp_count = spp_end - spp_begin;
    int64_t P = d->wave_samples > 0 ? d->wave_samples : default_wave_samples(c, 1, /*shadow=*/false);
    const int64_t total = npix * spp_count;
    P = std::max<int64_t>(1, std::min(P, std::max<int64_t>(total, 1)));
    if (P > (int64_t(1) << 25)) return c->fail(TRB_ERR_INVALID, "wave_samples must not exceed 2^25");
    rc = ensure_wave(c, size_t(P), 1, /*shadow=*/false);
    if (rc) return rc;
    c->wave_used = uint64_t(P);
    trb::CameraDev cam;
    trb::make_camera(*d, cam);
    c->timing = stats != nullptr; c->n_events = 0; c->n_events2 = 0; c->launches = 0;
    TRB_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, TRB_NSTATS * sizeof(unsigned long long), c->stream));
    c->mark(LK_LOGIC);
    if (!accumulate) TRB_CUDA(c, cudaMemsetAsync(d_rgb, 0, size_t(npix) * 3 * sizeof(float), c->stream));
    WaveRun w{d, d_rgb, nullptr, nullptr, nullptr, 0, false};
    for (int64_t base = 0; base < total; base += P) {
        int n = int(std::min<int64_t>(P, total - base));
        rc = run_di_wave(c, w, dp, cam, base, n, spp_begin, spp_count);
        if (rc) return rc;
    }
    c->mark(LK_LOGIC);
    if (stats) collect_stats(c, stats, uint64_t(total));
    return TRB_OK;
}

}  // namespace

extern "C" {

int trb_framebuffer_upload(trb_context* c, const float* rgb, int width, int height) {
    if (!c || !rgb || width <= 0 || height <= 0) return TRB_ERR_INVALID;
    cudaSetDevice(c->device);
    const size_t nf = size_t(width) * height * 3;
    TRB_CUDA(c, c->fb.alloc(nf));
    TRB_CUDA(c, cudaMemcpyAsync(c->fb.p, rgb, nf * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_render_direct_device(trb_context* c, const trb_render_desc* d, const trb_direct_desc* dd, int accumulate, float* d_rgb, trb_stats* stats) {
    if (!c || !d_rgb) return TRB_ERR_INVALID;
    int rc = render_direct_impl(c, d, dd, accumulate, d_rgb, stats);
    if (rc) return rc;
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_render_direct(trb_context* c, const trb_render_desc* d, const trb_direct_desc* dd, int accumulate, float* rgb_out, trb_stats* stats) {
    if (!c || !d) return TRB_ERR_INVALID;
    if (d->width <= 0 || d->height <= 0) return c->fail(TRB_ERR_INVALID, "bad film size");
    cudaSetDevice(c->device);
    size_t nf = size_t(d->width) * d->height * 3;
    bool fresh = c->fb.n < nf;
    TRB_CUDA(c, c->fb.alloc(nf));
    if (fresh && accumulate) TRB_CUDA(c, cudaMemsetAsync(c->fb.p, 0, nf * sizeof(float), c->stream));
    int rc = render_direct_impl(c, d, dd, accumulate, c->fb.p, stats);
    if (rc) return rc;
    if (rgb_out) TRB_CUDA(c, cudaMemcpyAsync(rgb_out, c->fb.p, nf * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

/* DirectIntegrator::render on n given camera rays with injected random numbers (n x stride, consumed in the reference's order).
 * li_out: n x 3; draws_out (optional): numbers consumed per sample. */
int trb_direct_paths(trb_context* c, const trb_render_desc* d, const trb_direct_desc* dd, const trb_ray* cam_rays, const float* rng, size_t n,
                     size_t stride, float* li_out, int32_t* draws_out) {
    if (!c || !cam_rays || !rng || !li_out) return TRB_ERR_INVALID;
    DiParams dp;
    int rc = check_direct_desc(c, d, dd, &dp);
    if (rc) return rc;
    if (n == 0) return TRB_OK;
    if (n > (size_t(1) << 24)) return c->fail(TRB_ERR_INVALID, "trb_direct_paths: at most 2^24 samples per call");
    if (stride < 1) return c->fail(TRB_ERR_INVALID, "trb_direct_paths: rng stride must be >= 1");
    cudaSetDevice(c->device);
    rc = ensure_wave(c, n, 1, /*shadow=*/false);
    if (rc) return rc;
    DevBuf<RayT> d_rays; DevBuf<float> d_rng, d_acc; DevBuf<int32_t> d_draws;
    TRB_CUDA(c, d_rays.alloc(n)); TRB_CUDA(c, d_rng.alloc(n * stride)); TRB_CUDA(c, d_acc.alloc(n * 3)); TRB_CUDA(c, d_draws.alloc(n));
    TRB_CUDA(c, cudaMemcpyAsync(d_rays.p, cam_rays, n * sizeof(RayT), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaMemcpyAsync(d_rng.p, rng, n * stride * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaMemsetAsync(d_acc.p, 0, d_acc.bytes(), c->stream));
    TRB_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, TRB_NSTATS * sizeof(unsigned long long), c->stream));
    trb::CameraDev cam;
    trb::make_camera(*d, cam);
    c->timing = false; c->n_events = 0; c->n_events2 = 0; c->launches = 0;
    WaveRun w{d, d_acc.p, nullptr, d_rays.p, d_rng.p, int(stride), true};
    rc = run_di_wave(c, w, dp, cam, 0, int(n), 0, d->spp);
    if (rc == TRB_OK) {
        k_read_draws<<<grid_exact(n), TRB_BLOCK, 0, c->stream>>>(c->st_v.p, int(n), d_draws.p);
        std::vector<int32_t> draws(n);
        cudaError_t e = cudaMemcpyAsync(li_out, d_acc.p, d_acc.bytes(), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(draws.data(), d_draws.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = c->cuda_fail(e, "trb_direct_paths download");
        else if (draws_out) std::copy(draws.begin(), draws.end(), draws_out);
    }
    d_rays.release(); d_rng.release(); d_acc.release(); d_draws.release();
    return rc;
}

int trb_render_path_device(trb_context* c, const trb_render_desc* d, const trb_path_desc* pd, int accumulate, float* d_rgb, trb_stats* stats) {
    if (!c || !d_rgb) return TRB_ERR_INVALID;
    int rc = render_path_impl(c, d, pd, accumulate, d_rgb, stats);
    if (rc) return rc;
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_render_path(trb_context* c, const trb_render_desc* d, const trb_path_desc* pd, int accumulate, float* rgb_out, trb_stats* stats) {
    if (!c || !d) return TRB_ERR_INVALID;
    if (d->width <= 0 || d->height <= 0) return c->fail(TRB_ERR_INVALID, "bad film size");
    cudaSetDevice(c->device);
    size_t nf = size_t(d->width) * d->height * 3;
    bool fresh = c->fb.n < nf;
    TRB_CUDA(c, c->fb.alloc(nf));
    if (fresh && accumulate) TRB_CUDA(c, cudaMemsetAsync(c->fb.p, 0, nf * sizeof(float), c->stream));
    int rc = render_path_impl(c, d, pd, accumulate, c->fb.p, stats);
    if (rc) return rc;
    if (rgb_out) TRB_CUDA(c, cudaMemcpyAsync(rgb_out, c->fb.p, nf * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

/* PathTracerIntegrator::render on n given camera rays with injected random numbers (n x stride, consumed in the reference's
 * depth-first order).  li_out: n x 3; draws_out (optional): numbers consumed per path. */
int trb_path_paths(trb_context* c, const trb_render_desc* d, const trb_path_desc* pd, const trb_ray* cam_rays, const float* rng, size_t n,
                   size_t stride, float* li_out, int32_t* draws_out) {
    if (!c || !cam_rays || !rng || !li_out) return TRB_ERR_INVALID;
    PtParams pt;
    int rc = check_path_desc(c, d, pd, &pt);
    if (rc) return rc;
    if (n == 0) return TRB_OK;
    if (n > (size_t(1) << 24)) return c->fail(TRB_ERR_INVALID, "trb_path_paths: at most 2^24 paths per call");
    cudaSetDevice(c->device);
    if (stride < 8) return c->fail(TRB_ERR_INVALID, "trb_path_paths: rng stride must be >= 8");
    rc = ensure_wave(c, n, pt.is_explicit ? pt.cap : 1, /*shadow=*/false);
    if (rc) return rc;
    DevBuf<RayT> d_rays; DevBuf<float> d_rng, d_acc; DevBuf<int32_t> d_draws;
    TRB_CUDA(c, d_rays.alloc(n)); TRB_CUDA(c, d_rng.alloc(n * stride)); TRB_CUDA(c, d_acc.alloc(n * 3)); TRB_CUDA(c, d_draws.alloc(n));
    TRB_CUDA(c, cudaMemcpyAsync(d_rays.p, cam_rays, n * sizeof(RayT), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaMemcpyAsync(d_rng.p, rng, n * stride * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaMemsetAsync(d_acc.p, 0, d_acc.bytes(), c->stream));
    TRB_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, TRB_NSTATS * sizeof(unsigned long long), c->stream));
    trb::CameraDev cam;
    trb::make_camera(*d, cam);
    c->timing = false; c->n_events = 0; c->n_events2 = 0; c->launches = 0;
    WaveRun w{d, d_acc.p, nullptr, d_rays.p, d_rng.p, int(stride), true};
    rc = run_pt_wave(c, w, pt, cam, 0, int(n), 0, d->spp);
    if (rc == TRB_OK) {
        k_read_draws<<<grid_exact(n), TRB_BLOCK, 0, c->stream>>>(c->st_v.p, int(n), d_draws.p);
        std::vector<int32_t> draws(n);
        cudaError_t e = cudaMemcpyAsync(li_out, d_acc.p, d_acc.bytes(), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(draws.data(), d_draws.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = c->cuda_fail(e, "trb_path_paths download");
        else if (draws_out) std::copy(draws.begin(), draws.end(), draws_out);
    }
    d_rays.release(); d_rng.release(); d_acc.release(); d_draws.release();
    return rc;
}

}  // extern "C"

namespace {
}  // namespace

extern "C" {

int trb_render_bdpt_device(trb_context* c, const trb_render_desc* d, int accumulate, float* d_rgb, trb_stats* stats) {
    if (!c || !d_rgb) return TRB_ERR_INVALID;
    int rc = render_impl(c, d, accumulate, d_rgb, stats);
    if (rc) return rc;
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_render_bdpt(trb_context* c, const trb_render_desc* d, int accumulate, float* rgb_out, trb_stats* stats) {
    if (!c || !d) return TRB_ERR_INVALID;
    if (d->width <= 0 || d->height <= 0) return c->fail(TRB_ERR_INVALID, "bad film size");
    cudaSetDevice(c->device);
    size_t nf = size_t(d->width) * d->height * 3;
    bool fresh = c->fb.n < nf;
    TRB_CUDA(c, c->fb.alloc(nf));
    if (fresh && accumulate) TRB_CUDA(c, cudaMemsetAsync(c->fb.p, 0, nf * sizeof(float), c->stream));
    int rc = render_impl(c, d, accumulate, c->fb.p, stats);
    if (rc) return rc;
    if (rgb_out) TRB_CUDA(c, cudaMemcpyAsync(rgb_out, c->fb.p, nf * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

}  // extern "C"

// ---- one frame on several GPUs of this process: sample-index sharding + ONE ncclReduce over NVLink ------------------
// (bench.py does the same across processes with torch.distributed).  NCCL is loaded at run time by its SONAME
// - the library has no link-time dependency on it, and a process that already holds torch's copy gets that one.
#include <dlfcn.h>
#include <mutex>
namespace {
struct Nccl {
    void* lib = nullptr;
    int (*CommInitAll)(void** comms, int n, const int* devs) = nullptr;
    int (*CommDestroy)(void* comm) = nullptr;
    int (*Reduce)(const void* send, void* recv, size_t count, int dtype, int op, int root, void* comm, cudaStream_t st) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string err;
    bool load() {
        if (lib) return true;
        // TRB_NCCL_LIB, when set, is the only library tried (an explicit override; a bad path fails instead of silently taking another copy)
        const char* over = std::getenv("TRB_NCCL_LIB");
        const char* names[] = {over && *over ? over : "libnccl.so.2", over && *over ? nullptr : "libnccl.so"};
        for (const char* nm : names) { if (nm && (lib = dlopen(nm, RTLD_NOW | RTLD_LOCAL))) break; }
        if (!lib) {
            const char* e = dlerror();   // ONE call: dlerror() clears the message it returns
            err = std::string("cannot load ") + (over && *over ? over : "libnccl.so.2") + ": " + (e ? e : "?");
            return false;
        }
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(dlsym(lib, "ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
        Reduce = reinterpret_cast<decltype(Reduce)>(dlsym(lib, "ncclReduce"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(dlsym(lib, "ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(dlsym(lib, "ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
        if (!CommInitAll || !CommDestroy || !Reduce || !GroupStart || !GroupEnd) { err = "libnccl.so.2 lacks a required symbol"; dlclose(lib); lib = nullptr; return false; }
        return true;
    }
    std::vector<int> devs;      // communicators are kept for the device list they were made for
    std::vector<void*> comms;
    bool comms_for(const std::vector<int>& want) {
        if (want == devs && !comms.empty()) return true;
        for (void* cm : comms) CommDestroy(cm);
        comms.assign(want.size(), nullptr); devs = want;
        int rc = CommInitAll(comms.data(), int(want.size()), want.data());
        if (rc != 0) { err = std::string("ncclCommInitAll: ") + (GetErrorString ? GetErrorString(rc) : "error"); comms.clear(); devs.clear(); return false; }
        return true;
    }
};
Nccl g_nccl;
std::mutex g_nccl_mutex;
}  // namespace

namespace {
// fn(context, description of its sample range, device frame buffer, stats or null) -> status
template <class RenderOne>
int render_multi(trb_context* const* ctxs, int n, const trb_render_desc* d, float* rgb_out, trb_stats* stats, RenderOne fn) {
    trb_context* c0 = ctxs[0];
    if (d->width <= 0 || d->height <= 0 || d->spp < n) return c0->fail(TRB_ERR_INVALID, "trb_render_bdpt_multi: needs spp >= number of devices");
    const int s0 = d->spp_begin, s1 = d->spp_end > 0 ? d->spp_end : d->spp;
    // every device gets a non-empty slice of [s0, s1): an empty one would reach render_impl as spp_end == 0, which means "up to spp"
    if (s0 < 0 || s1 > d->spp || s1 - s0 < n) return c0->fail(TRB_ERR_INVALID, "trb_render_*_multi: the sample range must hold at least one sample per device");
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (!g_nccl.load()) return c0->fail(TRB_ERR_UNSUPPORTED, g_nccl.err);
    std::vector<int> devs;
    for (int g = 0; g < n; ++g) { if (!ctxs[g]) return TRB_ERR_INVALID; devs.push_back(ctxs[g]->device); }
    for (int g = 0; g < n; ++g) for (int h = g + 1; h < n; ++h) if (devs[g] == devs[h]) return c0->fail(TRB_ERR_INVALID, "trb_render_bdpt_multi: one context per device");
    if (!g_nccl.comms_for(devs)) return c0->fail(TRB_ERR_CUDA, g_nccl.err);
    const size_t nf = size_t(d->width) * d->height * 3;
    std::vector<int> rc(size_t(n), TRB_OK);
    std::vector<std::thread> th;
    for (int g = 0; g < n; ++g) {
        th.emplace_back([&, g]() {   // one host thread per device: each context has its own streams
            trb_context* c = ctxs[g];
            cudaSetDevice(c->device);
            if (c->fb.alloc(nf) != cudaSuccess) { rc[g] = c->fail(TRB_ERR_CUDA, "frame buffer allocation"); return; }
            trb_render_desc part = *d;
            part.spp_begin = s0 + int(int64_t(s1 - s0) * g / n);
            part.spp_end = s0 + int(int64_t(s1 - s0) * (g + 1) / n);
            rc[g] = fn(c, &part, c->fb.p, g == 0 ? stats : nullptr);
        });
    }
    for (auto& t : th) t.join();
    for (int g = 0; g < n; ++g) if (rc[g] != TRB_OK) return rc[g];
    // contributions are normalised by the total spp: the partial frame buffers are plain summands -> one reduce onto device 0
    int nrc = g_nccl.GroupStart();
    for (int g = 0; g < n && nrc == 0; ++g) {
        cudaSetDevice(ctxs[g]->device);
        nrc = g_nccl.Reduce(ctxs[g]->fb.p, ctxs[g]->fb.p, nf, /*ncclFloat*/ 7, /*ncclSum*/ 0, 0, g_nccl.comms[g], ctxs[g]->stream);
    }
    const int erc = g_nccl.GroupEnd();
    if (nrc == 0) nrc = erc;
    if (nrc != 0) return c0->fail(TRB_ERR_CUDA, std::string("ncclReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc) : "error"));
    for (int g = 0; g < n; ++g) { cudaSetDevice(ctxs[g]->device); TRB_CUDA(ctxs[g], cudaStreamSynchronize(ctxs[g]->stream)); }
    cudaSetDevice(c0->device);
    if (rgb_out) TRB_CUDA(c0, cudaMemcpyAsync(rgb_out, c0->fb.p, nf * sizeof(float), cudaMemcpyDeviceToHost, c0->stream));
    TRB_CUDA(c0, cudaStreamSynchronize(c0->stream));
    return TRB_OK;
}
}  // namespace

extern "C" int trb_render_bdpt_multi(trb_context* const* ctxs, int n, const trb_render_desc* d, float* rgb_out, trb_stats* stats) {
    if (!ctxs || n < 1 || !d || !ctxs[0]) return TRB_ERR_INVALID;
    if (n == 1) return trb_render_bdpt(ctxs[0], d, 0, rgb_out, stats);
    return render_multi(ctxs, n, d, rgb_out, stats,
                        [](trb_context* c, const trb_render_desc* part, float* fb, trb_stats* st) { return render_impl(c, part, 0, fb, st); });
}

extern "C" int trb_render_path_multi(trb_context* const* ctxs, int n, const trb_render_desc* d, const trb_path_desc* pd, float* rgb_out, trb_stats* stats) {
    if (!ctxs || n < 1 || !d || !pd || !ctxs[0]) return TRB_ERR_INVALID;
    if (n == 1) return trb_render_path(ctxs[0], d, pd, 0, rgb_out, stats);
    return render_multi(ctxs, n, d, rgb_out, stats,
                        [pd](trb_context* c, const trb_render_desc* part, float* fb, trb_stats* st) { return render_path_impl(c, part, pd, 0, fb, st); });
}

extern "C" int trb_render_direct_multi(trb_context* const* ctxs, int n, const trb_render_desc* d, const trb_direct_desc* dd, float* rgb_out, trb_stats* stats) {
    if (!ctxs || n < 1 || !d || !dd || !ctxs[0]) return TRB_ERR_INVALID;
    if (n == 1) return trb_render_direct(ctxs[0], d, dd, 0, rgb_out, stats);
    return render_multi(ctxs, n, d, rgb_out, stats,
                        [dd](trb_context* c, const trb_render_desc* part, float* fb, trb_stats* st) { return render_direct_impl(c, part, dd, 0, fb, st); });
}

// ---- parity entry points -------------------------------------------------------------------------------------
namespace {

// hit record -> trb_hit with the range semantics of AcceleratorBVH::intersect (accel.h:131-133,169-170)
__global__ void k_export_hits(SceneDev sc, const RayT* __restrict__ rays, const HitT* __restrict__ hits, size_t n, trb_hit* __restrict__ out) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    HitT h = hits[i];
    RayT r = rays[i];
    trb_hit o;
    o.t = h.pos >= 0 ? h.t : FLT_MAX; o.u = 0.f; o.v = 0.f; o.hit = 0; o.shape = -1; o.prim = -1; o.tri = -1; o._pad = 0;
    if (hit_valid(h, r)) {
        const trb::DevShade& s = sc.shade[h.pos];
        o.hit = 1; o.u = h.u; o.v = h.v; o.shape = s.shape; o.prim = s.prim; o.tri = s.tri;
    }
    out[i] = o;
}

__global__ void k_bsdf_batch(SceneDev sc, const trb_bsdf_query* __restrict__ q, size_t n, int n_mats, trb_bsdf_out* __restrict__ out) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    trb_bsdf_query x = q[i];
    trb_bsdf_out o;
    memset(&o, 0, sizeof(o));
    if (x.mat >= 0 && x.mat < n_mats) {
        Mat m = load_mat(sc.mats, x.mat);
        V3 wo = mk(x.wo[0], x.wo[1], x.wo[2]), wi = mk(x.wi[0], x.wi[1], x.wi[2]);
        V3 e = bsdf_eval(m, wo, wi);
        o.eval[0] = e.x; o.eval[1] = e.y; o.eval[2] = e.z;
        o.pdf = bsdf_pdf(m, wo, wi);
        V3 swi; float spdf;
        V3 sv = bsdf_sample(m, wo, x.u[0], x.u[1], &swi, &spdf);
        o.sample_val[0] = sv.x; o.sample_val[1] = sv.y; o.sample_val[2] = sv.z; o.sample_pdf = spdf;
        o.sample_wi[0] = swi.x; o.sample_wi[1] = swi.y; o.sample_wi[2] = swi.z; o.type = m.type;
    }
    out[i] = o;
}

__global__ void k_read_draws(const float4* __restrict__ st_v, int n, int32_t* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = int32_t(__float_as_uint(st_v[i].z));
}

int trace_device(trb_context* c, bool any, const RayT* d_rays, size_t n, HitT* d_hits, uint8_t* d_occ, float* ms_out) {
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    if (n > 0xffffffffull) return c->fail(TRB_ERR_INVALID, "too many rays for one call");
    cudaSetDevice(c->device);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms_out) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, c->stream); }
    unsigned long long* visit = nullptr;
#ifdef TRB_COUNTERS
    TRB_CUDA(c, c->d_stats.alloc(TRB_NSTATS));
    visit = c->d_stats.p + 2;
#endif
    if (n) {
        TRB_CUDA(c, c->fetch_ctr.alloc(TRACE_CTR_WORDS));
        TRB_CUDA(c, cudaMemsetAsync(c->fetch_ctr.p, 0, TRACE_CTR_WORDS * sizeof(uint32_t), c->stream));
        TRB_CUDA(c, c->defer_q[0].alloc(n));
        if (any) launch_any(c, d_rays, nullptr, uint32_t(n), uint32_t(n), d_occ, nullptr, nullptr, c->fetch_ctr.p, visit ? visit + 2 : nullptr);
        else launch_closest(c, d_rays, nullptr, nullptr, uint32_t(n), d_hits, c->fetch_ctr.p, visit);
    }
    ++c->launches;
    if (ms_out) { cudaEventRecord(e1, c->stream); }
    TRB_CUDA(c, cudaGetLastError());
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (ms_out) { cudaEventElapsedTime(ms_out, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1); }
    return TRB_OK;
}

}  // namespace

extern "C" {

int trb_trace_closest(trb_context* c, const trb_ray* rays, size_t n, trb_hit* out) {
    if (!c || (n && (!rays || !out))) return TRB_ERR_INVALID;
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    cudaSetDevice(c->device);
    static_assert(sizeof(trb_ray) == sizeof(RayT) && sizeof(trb_hit) == 32, "ABI layout");
    TRB_CUDA(c, c->io_rays.alloc(n)); TRB_CUDA(c, c->io_hits.alloc(n));
    DevBuf<trb_hit> d_out;
    TRB_CUDA(c, d_out.alloc(n));
    TRB_CUDA(c, cudaMemcpyAsync(c->io_rays.p, rays, n * sizeof(RayT), cudaMemcpyHostToDevice, c->stream));
    int rc = trace_device(c, false, c->io_rays.p, n, c->io_hits.p, nullptr, nullptr);
    if (rc == TRB_OK && n) {
        k_export_hits<<<grid_exact(n), TRB_BLOCK, 0, c->stream>>>(c->sc, c->io_rays.p, c->io_hits.p, n, d_out.p);
        cudaError_t e = cudaMemcpyAsync(out, d_out.p, n * sizeof(trb_hit), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = c->cuda_fail(e, "trb_trace_closest download");
    }
    d_out.release();
    return rc;
}

int trb_trace_any(trb_context* c, const trb_ray* rays, size_t n, uint8_t* out) {
    if (!c || (n && (!rays || !out))) return TRB_ERR_INVALID;
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    cudaSetDevice(c->device);
    TRB_CUDA(c, c->io_rays.alloc(n)); TRB_CUDA(c, c->io_occ.alloc(n));
    TRB_CUDA(c, cudaMemcpyAsync(c->io_rays.p, rays, n * sizeof(RayT), cudaMemcpyHostToDevice, c->stream));
    int rc = trace_device(c, true, c->io_rays.p, n, nullptr, c->io_occ.p, nullptr);
    if (rc) return rc;
    TRB_CUDA(c, cudaMemcpyAsync(out, c->io_occ.p, n, cudaMemcpyDeviceToHost, c->stream));
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_trace_closest_device(trb_context* c, const trb_ray* d_rays, size_t n, trb_hit* d_out, float* ms_out) {
    if (!c || !d_rays || !d_out) return TRB_ERR_INVALID;
    cudaSetDevice(c->device);
    TRB_CUDA(c, c->io_hits.alloc(n));
    int rc = trace_device(c, false, reinterpret_cast<const RayT*>(d_rays), n, c->io_hits.p, nullptr, ms_out);
    if (rc) return rc;
    if (n) k_export_hits<<<grid_exact(n), TRB_BLOCK, 0, c->stream>>>(c->sc, reinterpret_cast<const RayT*>(d_rays), c->io_hits.p, n, d_out);
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}

int trb_trace_any_device(trb_context* c, const trb_ray* d_rays, size_t n, uint8_t* d_out, float* ms_out) {
    if (!c || !d_rays || !d_out) return TRB_ERR_INVALID;
    return trace_device(c, true, reinterpret_cast<const RayT*>(d_rays), n, nullptr, d_out, ms_out);
}

int trb_bsdf_batch(trb_context* c, const trb_bsdf_query* q, size_t n, trb_bsdf_out* out) {
    if (!c || (n && (!q || !out))) return TRB_ERR_INVALID;
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    cudaSetDevice(c->device);
    DevBuf<trb_bsdf_query> dq; DevBuf<trb_bsdf_out> dout;
    TRB_CUDA(c, dq.alloc(n)); TRB_CUDA(c, dout.alloc(n));
    cudaError_t e = cudaMemcpyAsync(dq.p, q, n * sizeof(trb_bsdf_query), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && n) {
        k_bsdf_batch<<<grid_exact(n), TRB_BLOCK, 0, c->stream>>>(c->sc, dq.p, n, int(c->flat.materials.size()), dout.p);
        ++c->launches;
        e = cudaMemcpyAsync(out, dout.p, n * sizeof(trb_bsdf_out), cudaMemcpyDeviceToHost, c->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    dq.release(); dout.release();
    if (e != cudaSuccess) return c->cuda_fail(e, "trb_bsdf_batch");
    return TRB_OK;
}

int trb_bdpt_paths(trb_context* c, const trb_render_desc* d, const trb_ray* cam_rays, const float* rng, size_t n, size_t stride,
                   float* li_out, int32_t* splat_pix, float* splat_rgb, int32_t* draws_out) {
    if (!c || !cam_rays || !rng || !li_out || !splat_pix || !splat_rgb) return TRB_ERR_INVALID;
    int rc = check_desc(c, d);
    if (rc) return rc;
    if (n == 0) return TRB_OK;
    if (n > (size_t(1) << 24)) return c->fail(TRB_ERR_INVALID, "trb_bdpt_paths: at most 2^24 paths per call");
    cudaSetDevice(c->device);
    const int K = std::max(max_vertices_of(*d), 1);  // rr_depth-1, or max_depth-1 under Russian roulette
    // worst case draws per path: 2 + 6 + 3K (light) + 7K (eye) + 2
    if (stride < size_t(10 + 10 * K)) return c->fail(TRB_ERR_INVALID, "trb_bdpt_paths: rng stride must be >= 10 + 10*K, K = max vertices per subpath");
    rc = ensure_wave(c, n, K);
    if (rc) return rc;
    DevBuf<RayT> d_rays; DevBuf<float> d_rng, d_acc; DevBuf<int32_t> d_pix, d_draws;
    TRB_CUDA(c, d_rays.alloc(n)); TRB_CUDA(c, d_rng.alloc(n * stride)); TRB_CUDA(c, d_acc.alloc(n * size_t(K + 1) * 3));
    TRB_CUDA(c, d_pix.alloc(n * size_t(K))); TRB_CUDA(c, d_draws.alloc(n));
    TRB_CUDA(c, cudaMemcpyAsync(d_rays.p, cam_rays, n * sizeof(RayT), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaMemcpyAsync(d_rng.p, rng, n * stride * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    TRB_CUDA(c, cudaMemsetAsync(d_acc.p, 0, d_acc.bytes(), c->stream));
    TRB_CUDA(c, cudaMemsetAsync(d_pix.p, 0xff, d_pix.bytes(), c->stream));
    TRB_CUDA(c, cudaMemsetAsync(c->d_stats.p, 0, TRB_NSTATS * sizeof(unsigned long long), c->stream));
    trb::CameraDev cam;
    trb::make_camera(*d, cam);
    c->timing = false; c->n_events = 0; c->n_events2 = 0; c->launches = 0;
    WaveRun w{d, d_acc.p, d_pix.p, d_rays.p, d_rng.p, int(stride), true};
    rc = run_wave(c, w, cam, 0, int(n), 0, d->spp);
    if (rc == TRB_OK) {
        k_read_draws<<<grid_exact(n), TRB_BLOCK, 0, c->stream>>>(c->st_v.p, int(n), d_draws.p);
        std::vector<float> acc(n * size_t(K + 1) * 3);
        std::vector<int32_t> pix(n * size_t(K)), draws(n);
        cudaError_t e = cudaMemcpyAsync(acc.data(), d_acc.p, d_acc.bytes(), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(pix.data(), d_pix.p, d_pix.bytes(), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(draws.data(), d_draws.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = c->cuda_fail(e, "trb_bdpt_paths download");
        else {
            for (size_t i = 0; i < n; ++i) {
                const float* a = acc.data() + i * size_t(K + 1) * 3;
                li_out[3 * i] = a[0]; li_out[3 * i + 1] = a[1]; li_out[3 * i + 2] = a[2];
                for (int k = 0; k < K; ++k) {
                    const float* s = a + 3 * (k + 1);
                    bool any = s[0] != 0.f || s[1] != 0.f || s[2] != 0.f;  // an occluded splat leaves no trace in the reference either
                    splat_pix[i * K + k] = any ? pix[i * K + k] : -1;
                    for (int ch = 0; ch < 3; ++ch) splat_rgb[(i * K + k) * 3 + ch] = s[ch];
                }
                if (draws_out) draws_out[i] = draws[i];
            }
        }
    }
    d_rays.release(); d_rng.release(); d_acc.release(); d_pix.release(); d_draws.release();
    return rc;
}

}  // extern "C"

// ---- an entry point that allocates nothing (no counterpart in the reference) --------
// Some callers cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  This is
// trb_trace_any_device with the scratch the traversal needs - the launch's three queue counters (bdpt_common.cuh: TRACE_CTR_WORDS) and
// the queue of ray slots k_trace_wide hands to k_trace_coop (rays with a zero / non-finite direction component, at most one entry per
// ray) - in a file-scope device array instead of cudaMalloc'ed buffers of the context.  Every GPU that loads the library holds the
// array (256 B + 4 B x TRB_NOALLOC_MAX_RAYS = 4 MiB + 256 B) until exit.  All calls on a device share it and it keeps the last call's
// contents: the counters are cleared on the stream before every launch (the queue needs no clearing: only entries below the freshly
// counted length are read); two calls that overlap on one GPU (two contexts of one device from two threads) race on it.
static __device__ __align__(256) unsigned char g_noalloc_scratch[256 + sizeof(int32_t) * size_t(TRB_NOALLOC_MAX_RAYS)];

namespace {
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
std::mutex g_noalloc_mutex;
void* g_noalloc_ptr[64] = {nullptr};
int noalloc_scratch(trb_context* c, unsigned char** out) {
    if (c->device < 0 || c->device >= 64) return c->fail(TRB_ERR_UNSUPPORTED, "trb_trace_any_device_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[c->device]) {
        void* p = nullptr;
        TRB_CUDA(c, cudaSetDevice(c->device));
        TRB_CUDA(c, cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return c->fail(TRB_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the scratch array");
        g_noalloc_ptr[c->device] = p;
    }
    *out = static_cast<unsigned char*>(g_noalloc_ptr[c->device]);
    return TRB_OK;
}
}  // namespace

extern "C" int trb_trace_any_device_noalloc(trb_context* c, const trb_ray* d_rays, size_t n, uint8_t* d_out) {
    if (!c) return TRB_ERR_INVALID;
    if (n && (!d_rays || !d_out)) return c->fail(TRB_ERR_INVALID, "trb_trace_any_device_noalloc: null ray or output pointer");
    if (!c->has_scene) return c->fail(TRB_ERR_NO_SCENE, "no scene uploaded");
    if (n > size_t(TRB_NOALLOC_MAX_RAYS))
        return c->fail(TRB_ERR_INVALID, "trb_trace_any_device_noalloc: " + std::to_string(n) + " rays exceed the fixed scratch of TRB_NOALLOC_MAX_RAYS = " +
                                            std::to_string(size_t(TRB_NOALLOC_MAX_RAYS)) + " rays per call; split the batch or use trb_trace_any_device");
    if (n == 0) return TRB_OK;
    unsigned char* scratch = nullptr;
    int rc = noalloc_scratch(c, &scratch);
    if (rc) return rc;
    TRB_CUDA(c, cudaSetDevice(c->device));
    uint32_t* ctr = reinterpret_cast<uint32_t*>(scratch);               // [0] ray cursor, [1] deferred count, [2] cursor of the deferred launch
    int32_t* defer_q = reinterpret_cast<int32_t*>(scratch + 256);
    TRB_CUDA(c, cudaMemsetAsync(ctr, 0, TRACE_CTR_WORDS * sizeof(uint32_t), c->stream));
    TraceArgs a{};
    a.bvh = BvhView{c->sc.nodes, c->sc.tris, c->sc.root_ref};
    a.wnodes = reinterpret_cast<const F4*>(c->d_wnodes.p); a.wroot = c->flat.n_wnodes ? 0 : c->sc.root_ref;
    a.n_nodes = uint32_t(c->flat.n_nodes); a.n_wnodes = uint32_t(c->flat.n_wnodes); a.n_tris = uint32_t(c->flat.n_tris);
    a.rays = reinterpret_cast<const RayT*>(d_rays); a.q = nullptr; a.n_ptr = nullptr; a.n_fixed = uint32_t(n); a.cap = uint32_t(n);
    a.occ_out = d_out; a.pend = nullptr; a.accum = nullptr; a.fetch = ctr;
    a.defer_q = defer_q; a.defer_n = ctr + 1; a.defer_cap = uint32_t(TRB_NOALLOC_MAX_RAYS);
    a.visit = nullptr; a.visit_exh = nullptr;
    const int grid = c->prop.multiProcessorCount * c->trace_blocks_per_sm;
    k_trace_wide<true><<<grid, TRB_TRACE_BLOCK, 0, c->stream>>>(a);
    a.q = defer_q; a.n_ptr = ctr + 1; a.n_fixed = 0u; a.cap = uint32_t(TRB_NOALLOC_MAX_RAYS); a.fetch = ctr + 2;
    k_trace_coop<true><<<grid, TRB_TRACE_BLOCK, 0, c->stream>>>(a);
    c->launches += 2;
    TRB_CUDA(c, cudaGetLastError());
    TRB_CUDA(c, cudaStreamSynchronize(c->stream));
    return TRB_OK;
}
