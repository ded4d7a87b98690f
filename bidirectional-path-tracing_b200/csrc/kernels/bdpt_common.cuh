// bdpt_common.cuh -- wave state, RNG and the scene services shared by the BDPT wavefront kernels.
#pragma once
#include "../host/tr_host.h"
#include "bsdf.cuh"
#include "trace.cuh"
#include "vec.cuh"

namespace trbk {

struct SceneDev {
    const F4* nodes;
    const F4* tris;
    const trb::DevShade* shade;
    const trb::DevMaterial* mats;
    const trb::DevEmitter* emitters;
    const trb::DevEmitterTri* etris;
    const float* cdf;
    const float4* tri_uv;   // 2 per triangle (DevTriUV), build order; textured scenes only
    const int4* tex_info;   // DevTexInfo per texture
    const float* texels;    // RGB
    int32_t has_tex;        // 0: every reflectance is a constant (no shipped scene has map_Kd / map_Ks)
    const float4* esphere;  // per emitter: shape centre xyz, radius (renderer.cpp:349-358; DirectIntegrator only)
    int32_t root_ref;
    int32_t n_emitters;
};

// One shadow-ray work item: the ray plus the contribution that lands at `target` if it is unoccluded.
struct alignas(16) Pending {
    float r, g, b;
    uint32_t target;  // float index / 3 into the accumulation buffer
};

// Light vertex as stored for connectVertices: 64 B (PathVertex, bdpt.h:24-35, is ~176 B AoS)
struct alignas(16) LightVertex {
    float px, py, pz, vcm;
    float nx, ny, nz, vc;
    float wox, woy, woz;
    int32_t mat;
    float tx, ty, tz, rr;   // rr = PathVertex::rr (bdpt.h:31-32): 1, or 0.5 under Russian roulette
};

// Eye vertex handed from k_eye_bounce to k_connect_vertices (same 64 B shape as LightVertex; `target` is the
// accumulation slot of the sample: its pixel, or its Li slot in parity mode)
struct alignas(16) EyeVertex {
    float px, py, pz, vcm;
    float nx, ny, nz, vc;
    float wox, woy, woz;
    int32_t mat;             // | EV_RR_HALF when the vertex's continuation probability is 0.5
    float tx, ty, tz;
    uint32_t target;
};
enum : int32_t { EV_RR_HALF = 1 << 30 };

struct WaveParams {
    SceneDev sc;
    trb::CameraDev cam;
    // wave geometry
    int32_t P;            // sample slots in this wave
    int32_t K;            // max vertices per subpath: rr_depth - 1, or max_depth - 1 under Russian roulette
    int32_t rr_depth;
    int32_t rr_on;        // TRB_FLAG_RUSSIAN_ROULETTE: the reference's loop test with NO_RR 0 (bdpt.h:68,129,188,201)
    int32_t max_depth;    // hard cap on `depth` (== rr_depth when rr_on is 0)
    int32_t spp;          // total spp (normaliser)
    int32_t spp_begin;    // first sample index of this call
    int32_t spp_count;    // samples per pixel rendered by this call
    int64_t wave_base;    // first local sample number of the wave: local = s_local * (W*H) + pixel
    uint32_t seed;
    uint32_t flags;
    // parity mode (trb_bdpt_paths): rays and random numbers come from the caller
    const RayT* inject_rays;
    const float* inject_rng;
    int32_t inject_stride;
    int32_t parity;
    // accumulation target: frame buffer (3 floats/pixel) or, in parity mode, (K+1) RGB slots per path
    float* accum;
    int32_t* splat_pix;   // parity mode: pixel index per (path, light depth-1)
    // wave buffers
    RayT* cam_ray;        // P
    RayT* ray;            // P   current walk ray (path tracer / direct integrator: per slot)
    HitT* hit;            // P   current walk hit (BDPT: entry k belongs to queue entry k, see below; the others: per slot)
    HitT* hit0;           // P   primary hit (eye walk depth 1 reuses it: bdpt.h:70 re-traces the same ray)
    float4* st_t;         // P   throughput.xyz, vcm (path tracer / direct integrator: per slot)
    float4* st_v;         // P   vc, flags(bits), rng_ctr(bits), unused - per slot; BDPT: the value a walk ENDED with (the eye walk
                          //     and the draw counts of the parity entry point pick it up there)
    // BDPT: the state of a live walk travels in QUEUE ORDER - entry k of (ray_in, hit, stt_in, sv_in) belongs to slot q_in[k], and a
    // bounce writes the walks it continues to entry idx of the *_out arrays, idx being their place in q_out.  Every read and write
    // of a bounce kernel is then coalesced whatever slots are still alive, and the closest-hit traversal reads its rays and writes
    // its hits by queue position, with no index list in between.
    const RayT* ray_in;  RayT* ray_out;      // P each
    const float4* stt_in; float4* stt_out;   // P each: throughput.xyz, vcm
    const float4* sv_in;  float4* sv_out;    // P each: vc, flags, rng_ctr, unused
    LightVertex* lv;      // K * P.  BDPT: vertex j of slot i at [i*K + j] (slot-major: the <= K pairs that connect one eye vertex read one
                          // contiguous <= 64 K-byte run; depth-major put them P * 64 B = 1 GB apart at 2^24-sample waves).  The path tracer keeps its
                          // vertex stack depth-major ([d*P + i]: one kernel works on one depth of every slot)
    int32_t* nlv;         // P
    int32_t* q_in;        // P   live slots entering this bounce
    int32_t* q_out;       // P   live slots leaving it
    EyeVertex* ev;        // P   eye vertex of the current bounce (connecting slots only)
    float4* lv_tex;       // K * P (indexed like lv), textured scenes only: (triangle position bits, u, v, -) of the stored vertex, for BitmapTexture3f::eval
    float4* ev_tex;       // P
    int32_t* q_conn;      // P*K (slot << 6 | light vertex) pairs to connect this bounce, dense (path tracer: a P-slot queue)
    uint32_t* counters;   // [0]=q_out count, [1]=shadow count, [2]=closest rays traced, [3]=any rays traced
    RayT* sray;           // shadow capacity
    Pending* spend;       // shadow capacity
    uint32_t shadow_cap;
};

// a traversal launch owns three consecutive words: [0] ray-queue cursor of k_trace_wide, [1] number of rays it deferred to the binary tree,
// [2] cursor of the k_trace_coop launch that traces those
enum : uint32_t { CTR_QOUT = 0, CTR_SHADOW = 1, CTR_CONN = 2, CTR_SPARE = 3, CTR_FETCH_CLOSEST = 4, CTR_FETCH_ANY = 8, CTR_N = 12 };
enum : uint32_t { TRACE_CTR_WORDS = 3 };
enum : uint32_t { SLOT_PURE_SPECULAR = 1u };
#define TRB_EST_MASK (TRB_FLAG_LIGHT_TRACING | TRB_FLAG_PATH_TRACING)   // the reference's estimator A/B switches (bdpt.h:16-17)

// ---- random numbers ----------------------------------------------------------------------------------
// Production: counter-based, keyed by (seed, pixel, sample index, dimension) - independent of wave size and GPU count.  The key
// is 64 bits (two independent 32-bit hashes of the triple): configs[3] renders 3840 x 2160 x 4096 = 3.4e10 samples, eight times
// the 2^32 streams a 32-bit key can name, and a shared stream would repeat jitter, light and eye subpaths at two pixels.
// Parity: the caller's floats, consumed in the reference's order (Q17).
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
struct Rng {
    const float* inj;  // parity mode pointer to this path's row, or nullptr
    uint32_t key;      // hash of (seed, pixel, sample)
    uint32_t key2;     // a second, independent hash of the same triple
    uint32_t ctr;
    uint32_t lim;      // parity mode: floats in the row; past its end every draw reads 0.99 (ends any roulette)
    __device__ __forceinline__ float next() {
        uint32_t c = ctr++;
        if (inj) return c < lim ? inj[c] : 0.99f;
        uint32_t h = mix32(key2 ^ mix32(key ^ mix32(c * 0x9e3779b9u + 0x85ebca6bu)));
        return float(h >> 8) * (1.0f / 16777216.0f);
    }
};
__device__ __forceinline__ Rng make_rng(const WaveParams& p, int slot, uint32_t pixel, uint32_t sample, uint32_t ctr) {
    Rng r;
    r.inj = p.inject_rng ? p.inject_rng + size_t(slot) * p.inject_stride : nullptr;
    r.key = mix32(p.seed ^ mix32(pixel * 0x9e3779b9u + 0x7f4a7c15u) ^ mix32(sample * 0x85ebca6bu + 0xc2b2ae35u));
    r.key2 = mix32(mix32(p.seed + 0x632be5abu) + pixel * 0xcc9e2d51u) ^ mix32(sample * 0x1b873593u + 0x68e31da4u);
    r.ctr = ctr;
    r.lim = uint32_t(p.inject_stride);
    return r;
}

// local sample number -> (pixel, sample index)
__device__ __forceinline__ void slot_ids(const WaveParams& p, int slot, uint32_t* pixel, uint32_t* sample) {
    if (p.parity) { *pixel = 0; *sample = uint32_t(slot); return; }
    int64_t g = p.wave_base + slot;
    int64_t np = int64_t(p.cam.width) * p.cam.height;
    *pixel = uint32_t(g % np);
    *sample = uint32_t(p.spp_begin + g / np);
}

// warp-aggregated queue push (ballot + popc + one atomic per warp): returns the slot in the queue
__device__ __forceinline__ uint32_t queue_push(uint32_t* counter, bool want) {
    unsigned mask = __ballot_sync(0xffffffffu, want);
    if (mask == 0) return 0;
    int leader = __ffs(mask) - 1;
    int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, uint32_t(__popc(mask)));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + uint32_t(__popc(mask & ((1u << lane) - 1u)));
}

// ---- SurfaceInteraction from a hit record (accel.h:133-166) --------------------------------------------
struct Surf {
    V3 p, wo, n;   // position, local outgoing direction, shading normal (frameNs.n)
    Frame f;       // frameNs
    float t;
    int32_t mat, shape, emitter;
};

__device__ __forceinline__ bool hit_valid(const HitT& h, const RayT& r) {
    return h.pos >= 0 && h.t <= r.max_t && h.t >= r.min_t;  // accel.h:131-133
}

__device__ __forceinline__ Surf make_surf(const SceneDev& sc, const HitT& h, const RayT& r) {
    const float4* s = reinterpret_cast<const float4*>(sc.shade + h.pos);
    float4 a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2), d = __ldg(s + 3), e = __ldg(s + 4), f = __ldg(s + 5);
    Surf o;
    o.p = bary(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), h.u, h.v);
    o.n = normalize(bary(mk(d.x, d.y, d.z), mk(e.x, e.y, e.z), mk(f.x, f.y, f.z), h.u, h.v));
    o.f = make_frame(o.n);
    o.wo = to_local(o.f, mk(-r.dx, -r.dy, -r.dz));
    o.t = h.t;
    o.shape = __float_as_int(a.w); o.mat = __float_as_int(c.w); o.emitter = __float_as_int(d.w);
    return o;
}

// ---- bitmap textures (BitmapTexture3f::eval, core.h:569-585): nearest texel at the wrapped interpolated coordinate -------
__device__ __forceinline__ V3 tex_eval(const SceneDev& sc, int tex, int pos, float u, float v) {
    const float4 a = __ldg(sc.tri_uv + 2 * size_t(pos)), b = __ldg(sc.tri_uv + 2 * size_t(pos) + 1);
    const int4 ti = __ldg(sc.tex_info + tex);
    const float k = 1 - u - v;
    float sx = ((a.x * k + a.z * u) + b.x * v) + 1.0f, sy = ((a.y * k + a.w * u) + b.y * v) + 1.0f;
    sx = sx - floorf(sx); sy = sy - floorf(sy);
    const int px = min(max(int(sx * float(ti.y)), 0), ti.y - 1), py = min(max(int(sy * float(ti.z)), 0), ti.z - 1);
    const float* t = sc.texels + 3 * (size_t(uint32_t(ti.x)) + size_t(ti.y) * py + px);
    return mk(__ldg(t), __ldg(t + 1), __ldg(t + 2));
}
// the material as the BSDF sees it at a hit: Kd / Ks replaced by their bitmaps where the material has them
__device__ __forceinline__ void apply_tex(const SceneDev& sc, Mat& m, int pos, float u, float v) {
    if (m.tex & 0xffff) m.kd = tex_eval(sc, (m.tex & 0xffff) - 1, pos, u, v);
    if (m.tex >> 16) m.ks = tex_eval(sc, (m.tex >> 16) - 1, pos, u, v);
}
__device__ __forceinline__ Mat load_mat_at(const SceneDev& sc, int id, const HitT& h) {
    Mat m = load_mat(sc.mats, id);
    if (sc.has_tex && m.tex) apply_tex(sc, m, h.pos, h.u, h.v);
    return m;
}
__device__ __forceinline__ Mat load_mat_stored(const SceneDev& sc, int id, const float4* rec) {  // rec: the vertex's entry in lv_tex / ev_tex
    Mat m = load_mat(sc.mats, id);
    if (sc.has_tex && m.tex) { const float4 t = *rec; apply_tex(sc, m, __float_as_int(t.x), t.y, t.z); }
    return m;
}
__device__ __forceinline__ float4 tex_record(const HitT& h) { return make_float4(__int_as_float(h.pos), h.u, h.v, 0.f); }

// ---- emitter services (integrator.cpp:46-100, math.h:99-111,229-234) -----------------------------------
struct EmitterSample {
    V3 pos, n, radiance;
    float sel_pdf, area_pdf;
    int id;   // emitter index (Integrator::selectEmitter)
};

__device__ __forceinline__ EmitterSample sample_emitter(const SceneDev& sc, Rng& rng) {
    EmitterSample es;
    float s0 = rng.next();
    uint32_t ne = uint32_t(sc.n_emitters);
    uint32_t id = uint32_t(s0 * float(ne));
    id = min(id, ne - 1u);
    es.sel_pdf = 1.f / float(ne);
    es.id = int(id);
    const float4* ep = reinterpret_cast<const float4*>(sc.emitters + id);
    float4 e0 = __ldg(ep), e1 = __ldg(ep + 1);
    es.radiance = mk(e0.x, e0.y, e0.z);
    float area = e0.w;
    int first_face = __float_as_int(e1.y), n_faces = __float_as_int(e1.z), cdf_off = __float_as_int(e1.w);
    // Distribution1D::sample: upper_bound over n_faces+1 entries, then clamp(index-1, 0, n_faces-1)
    float s1 = rng.next();
    const float* cdf = sc.cdf + cdf_off;
    int lo = 0, hi = n_faces + 1;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (!(s1 < __ldg(cdf + mid))) lo = mid + 1; else hi = mid; }
    int prim = max(0, min(lo - 1, n_faces - 1));
    float ux = rng.next(), uy = rng.next();
    float su = sqrtf(1.f - ux);
    float bu = 1 - su, bv = su * uy;  // squareToUniformTriangle
    const float4* tp = reinterpret_cast<const float4*>(sc.etris + first_face + prim);
    float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
    V3 v0 = mk(q0.x, q0.y, q0.z), v1 = mk(q0.w, q1.x, q1.y), v2 = mk(q1.z, q1.w, q2.x);
    V3 n0 = mk(q2.y, q2.z, q2.w), n1 = mk(q3.x, q3.y, q3.z), n2 = mk(q3.w, q4.x, q4.y);
    es.pos = bary(v0, v1, v2, bu, bv);
    es.n = normalize(bary(n0, n1, n2, bu, bv));
    es.area_pdf = 1.f / area;
    return es;
}

// visibilityQuery's ray (bdpt.h:498-505): Ray(start, (end-start)/dist, Epsilon, dist - 0.00001f)
__device__ __forceinline__ RayT visibility_ray(V3 start, V3 end) {
    V3 dir = end - start;
    float dist = sqrtf(dot(dir, dir));
    dir = dir / dist;
    RayT r;
    r.ox = start.x; r.oy = start.y; r.oz = start.z; r.dx = dir.x; r.dy = dir.y; r.dz = dir.z;
    r.min_t = 1e-8f; r.max_t = dist - 0.00001f;
    return r;
}

__device__ __forceinline__ void push_shadow(const WaveParams& p, bool want, const RayT& r, V3 c, uint32_t target) {
    uint32_t idx = queue_push(p.counters + CTR_SHADOW, want);
    if (want && idx < p.shadow_cap) {
        p.sray[idx] = r;
        Pending pd; pd.r = c.x; pd.g = c.y; pd.b = c.z; pd.target = target;
        p.spend[idx] = pd;
    }
}

}  // namespace trbk
