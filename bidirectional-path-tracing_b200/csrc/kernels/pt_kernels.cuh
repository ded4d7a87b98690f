// pt_kernels.cuh -- the reference's unidirectional path tracer (PathTracerIntegrator, src/integrators/path.h:15-262) on the
// BDPT wavefront's buffers and traversal kernels.  The author uses it as the independent estimator BDPT is checked against
// (data/a5/bonus_bdpt/tinyrender/cbox_bdpt_path.toml); SURVEY 8(f) N2.
//
// recursiveExplicit (path.h:68-206) is depth-first: a vertex draws its roulette number and its indirect direction(s), the
// WHOLE deeper recursion runs, and only then the vertex draws its direct-illumination numbers.  The wavefront keeps that
// order of random numbers with two passes over a per-slot stack of vertex records:
//   forward, depth 0,1,2..:  k_pt_begin (roulette, first BSDF sample) -> trace_closest -> k_pt_check (emitter hit: the re-sampling
//                            loop of path.h:88-108 in place; else descend or stop)
//   backward, deepest first: m x [k_pt_emitter_gen, trace_closest, k_pt_emitter_resolve] + nb x [k_pt_bsdf_gen, trace_closest,
//                            k_pt_bsdf_resolve] + k_pt_combine  (Lr = direct + Li * f cos / pdf / n / cumRR, then the 1/rrProb scale)
// All rays are closest-hit queries (path.h:136: the emitter test compares the hit's shapeID).  The reference recurses without
// bound when maxDepth = -1; the device stops descending at `pt_cap` vertices (trb_render_desc.max_depth, default 64).
#pragma once
#include "bdpt_common.cuh"

namespace trbk {

struct PtParams {
    int32_t is_explicit, max_depth, rr_depth, cap;
    float rr_prob;
    int32_t emitter_samples, bsdf_samples;
};

// vertex record of the stack (LightVertex layout): p|pdf, n|nSamples bits, wo|mat+flags, f cos|unused
enum : int32_t { PT_HAS_CHILD = 1 << 30, PT_RR_SCALED = 1 << 29, PT_MAT_MASK = (1 << 29) - 1 };

struct PtVertex { V3 p, n, wo; Frame f; int mat; };
__device__ __forceinline__ PtVertex pt_load_vertex(const WaveParams& p, int d, int slot, float4* c_out, float4* d_out, float* pdf, uint32_t* ns) {
    const float4* lp = reinterpret_cast<const float4*>(p.lv + size_t(d) * p.P + slot);
    float4 a = lp[0], b = lp[1], c = lp[2];
    if (d_out) *d_out = lp[3];
    if (c_out) *c_out = c;
    if (pdf) *pdf = a.w;
    if (ns) *ns = __float_as_uint(b.w);
    PtVertex v;
    v.p = mk(a.x, a.y, a.z); v.n = mk(b.x, b.y, b.z); v.wo = mk(c.x, c.y, c.z); v.mat = __float_as_int(c.w) & PT_MAT_MASK;
    v.f = make_frame(v.n);
    return v;
}
__device__ __forceinline__ RayT pt_ray(V3 o, V3 d) {  // Ray(hit.p, wiW, Epsilon, infinity) (path.h:43,99,135,172)
    RayT r; r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.min_t = 1e-8f; r.max_t = __int_as_float(0x7f800000);
    return r;
}
// getEmission(visibilityInteraction): a miss leaves the value-initialised SurfaceInteraction, i.e. material 0 (core.h:173-180)
__device__ __forceinline__ V3 pt_vis_emission(const WaveParams& p, bool valid, int mat) {
    const float4* mp = reinterpret_cast<const float4*>(p.sc.mats + (valid ? mat : 0));
    float4 c = __ldg(mp + 2);
    return mk(c.x, c.y, c.z);
}
__device__ __forceinline__ void pt_accum(const WaveParams& p, int slot, V3 c) {
    uint32_t pixel, sample;
    slot_ids(p, slot, &pixel, &sample);
    const uint32_t target = p.parity ? uint32_t(slot) : pixel;
    c = c * (p.parity ? 1.f : (1.f / p.spp));  // renderer.cpp:201
    float* a = p.accum + size_t(target) * 3;
    atomicAdd(a, c.x); atomicAdd(a + 1, c.y); atomicAdd(a + 2, c.z);
}

// ---- PathTracerIntegrator::render (path.h:239-251): primary hit, emitter seen directly ------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_init(const __grid_constant__ WaveParams p, int32_t* __restrict__ q_out, uint32_t* __restrict__ n_out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false;
    if (i < p.P) {
        RayT cr = p.cam_ray[i];
        HitT h = p.hit[i];
        p.nlv[i] = 0;
        p.st_t[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (hit_valid(h, cr)) {
            const float4* s = reinterpret_cast<const float4*>(p.sc.shade + h.pos);
            const int mat = __float_as_int(__ldg(s + 2).w);
            V3 ke = pt_vis_emission(p, true, mat);
            if (!is_zero(ke)) pt_accum(p, i, ke);
            else { go = true; p.ray[i] = cr; }
        }
    }
    uint32_t idx = queue_push(n_out, go);
    if (go) q_out[idx] = i;
}

// ---- implicit path tracing (path.h:35-66), one bounce; throughput in st_t ------------------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_implicit(const __grid_constant__ WaveParams p, PtParams pt, int depth, const int32_t* __restrict__ q_in,
                                                           const uint32_t* __restrict__ n_in, int32_t* __restrict__ q_out, uint32_t* __restrict__ n_out) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false; int slot = 0;
    if (k < *n_in) {
        slot = q_in[k];
        RayT r = p.ray[slot];
        HitT h = p.hit[slot];
        float4 st = p.st_t[slot];
        V3 T = depth == 0 ? mk(1.f) : mk(st.x, st.y, st.z);
        bool valid = hit_valid(h, r);
        // depth 0 arrives with the primary hit; deeper, `h` is the hit of the direction sampled at depth-1
        bool shade = true;
        if (depth > 0) {
            if (!valid) shade = false;  // `return v3f(0.f)`
            else {
                Surf v = make_surf(p.sc, h, r);
                V3 ke = pt_vis_emission(p, true, v.mat);
                if (!is_zero(ke)) {
                    if (dot(v.n, mk(-r.dx, -r.dy, -r.dz)) > 0.f) pt_accum(p, slot, ke * T);  // path.h:54-56
                    shade = false;
                }
            }
        }
        if (shade && depth < pt.max_depth) {
            Surf s = make_surf(p.sc, h, r);
            Mat m = load_mat_at(p.sc, s.mat, h);
            uint32_t pixel, sample;
            slot_ids(p, slot, &pixel, &sample);
            float4 sv = p.st_v[slot];
            Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
            float ux = rng.next(), uy = rng.next();
            V3 wi; float pdf;
            V3 fcos = bsdf_sample(m, s.wo, ux, uy, &wi, &pdf);
            T = T * fcos * (1.f / pdf);
            p.ray[slot] = pt_ray(s.p, to_world(s.f, wi));
            p.st_t[slot] = make_float4(T.x, T.y, T.z, 0.f);
            sv.z = __uint_as_float(rng.ctr);
            p.st_v[slot] = sv;
            go = true;
        }
    }
    uint32_t idx = queue_push(n_out, go);
    if (go) q_out[idx] = slot;
}

// ---- recursiveExplicit, head of the call at `depth` (path.h:68-101): roulette, vertex record, first indirect direction ----
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_begin(const __grid_constant__ WaveParams p, PtParams pt, int depth, const int32_t* __restrict__ q_in,
                                                        const uint32_t* __restrict__ n_in, int32_t* __restrict__ q_try, uint32_t* __restrict__ n_try) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false; int slot = 0;
    if (k < *n_in) {
        slot = q_in[k];
        RayT r = p.ray[slot];
        HitT h = p.hit[slot];
        uint32_t pixel, sample;
        slot_ids(p, slot, &pixel, &sample);
        float4 sv = p.st_v[slot];
        Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
        const float rrSample = rng.next();
        if (depth < pt.max_depth || (pt.max_depth == -1 && (depth < pt.rr_depth || rrSample < pt.rr_prob))) {
            Surf s = make_surf(p.sc, h, r);
            Mat m = load_mat_at(p.sc, s.mat, h);
            float ux = rng.next(), uy = rng.next();
            V3 wi; float pdf;
            V3 fcos = bsdf_sample(m, s.wo, ux, uy, &wi, &pdf);
            const bool scaled = pt.max_depth == -1 && !(depth < pt.rr_depth) && (rrSample < pt.rr_prob);  // path.h:199-201
            LightVertex v;
            v.px = s.p.x; v.py = s.p.y; v.pz = s.p.z; v.vcm = pdf;
            v.nx = s.n.x; v.ny = s.n.y; v.nz = s.n.z; v.vc = __uint_as_float(1u);
            v.wox = s.wo.x; v.woy = s.wo.y; v.woz = s.wo.z; v.mat = s.mat | (scaled ? PT_RR_SCALED : 0);
            v.tx = fcos.x; v.ty = fcos.y; v.tz = fcos.z; v.rr = 0.f;
            p.lv[size_t(depth) * p.P + slot] = v;
            if (p.sc.has_tex) p.lv_tex[size_t(depth) * p.P + slot] = tex_record(h);
            p.nlv[slot] = depth + 1;
            p.ray[slot] = pt_ray(s.p, to_world(s.f, wi));
            go = true;
        }
        sv.z = __uint_as_float(rng.ctr);
        p.st_v[slot] = sv;
    }
    uint32_t idx = queue_push(n_try, go);
    if (go) q_try[idx] = slot;
}

// ---- after tracing the indirect direction (path.h:88-118): an emitter (or, on a miss, material 0 being emissive) makes the
// reference draw again with probability 0.95 - for a mirror that is the same direction over and over, 20 times on average.
// Those few slots run the whole `do .. while` here, with the scalar traversal of trace.cuh (same hit semantics as the warp
// kernel), so that the order of random numbers stays the reference's; then: descend or stop. ------------------------------------
#define TRB_PT_MAX_RETRIES 4096
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_check(const __grid_constant__ WaveParams p, PtParams pt, int depth,
                                                        const int32_t* __restrict__ q_try, const uint32_t* __restrict__ n_try,
                                                        int32_t* __restrict__ q_next, uint32_t* __restrict__ n_next,
                                                        unsigned long long* __restrict__ extra_rays) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool descend = false; int slot = 0;
    if (k < *n_try) {
        slot = q_try[k];
        RayT r = p.ray[slot];
        HitT h = p.hit[slot];
        bool valid = hit_valid(h, r);
        int vmat = 0;
        if (valid) vmat = __float_as_int(__ldg(reinterpret_cast<const float4*>(p.sc.shade + h.pos) + 2).w);
        V3 ke = pt_vis_emission(p, valid, vmat);
        LightVertex* vp = p.lv + size_t(depth) * p.P + slot;
        if (!is_zero(ke)) {
            uint32_t pixel, sample;
            slot_ids(p, slot, &pixel, &sample);
            float4 sv = p.st_v[slot];
            Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
            float pdf0; uint32_t ns;
            PtVertex v = pt_load_vertex(p, depth, slot, nullptr, nullptr, &pdf0, &ns);
            Mat m = load_mat_stored(p.sc, v.mat, p.lv_tex + size_t(depth) * p.P + slot);
            const BvhView bvh = bvh_of(p.sc);
            uint32_t tries = 0, traced = 0;
            while (!is_zero(ke) && rng.next() < 0.95f && tries < TRB_PT_MAX_RETRIES) {
                float ux = rng.next(), uy = rng.next();
                V3 wi; float pdf;
                V3 fcos = bsdf_sample(m, v.wo, ux, uy, &wi, &pdf);
                const RayT r2 = pt_ray(v.p, to_world(v.f, wi));
                // a mirror re-samples the very same direction (glass: one of two): the same ray has the same hit
                const bool same = r2.dx == r.dx && r2.dy == r.dy && r2.dz == r.dz;
                r = r2;
                if (!same) { bool occ; trace_ray<false>(bvh, r, &h, &occ, nullptr, nullptr); ++traced; }
                ++ns; ++tries;
                vp->vcm = pdf; vp->tx = fcos.x; vp->ty = fcos.y; vp->tz = fcos.z;
                valid = hit_valid(h, r);
                vmat = valid ? __float_as_int(__ldg(reinterpret_cast<const float4*>(p.sc.shade + h.pos) + 2).w) : 0;
                ke = pt_vis_emission(p, valid, vmat);
            }
            if (tries) {
                vp->vc = __uint_as_float(ns);
                p.ray[slot] = r; p.hit[slot] = h;
                if (extra_rays && traced) atomicAdd(extra_rays, (unsigned long long)traced);
            }
            sv.z = __uint_as_float(rng.ctr);
            p.st_v[slot] = sv;
        }
        if (is_zero(ke) && valid && depth + 1 < pt.cap) {
            const V3 fcos = mk(vp->tx, vp->ty, vp->tz);
            if (!is_zero(fcos)) { vp->mat |= PT_HAS_CHILD; descend = true; }  // path.h:110-117
        }
    }
    uint32_t in = queue_push(n_next, descend);
    if (descend) q_next[in] = slot;
}

// ---- direct illumination by emitter sampling (path.h:121-153), sample e of vertex `depth` ------------------------------
// stash in ev[slot]: (wiLocal, areaToSolidAngle), (emitterPdf, emitterAreaPdf, emitter id, -), emitterEstimator, bsdfEstimator
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_emitter_gen(const __grid_constant__ WaveParams p, PtParams pt, int depth, int e,
                                                              int32_t* __restrict__ q_try, uint32_t* __restrict__ n_try) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false;
    if (slot < p.P && p.nlv[slot] > depth) {
        PtVertex v = pt_load_vertex(p, depth, slot, nullptr, nullptr, nullptr, nullptr);
        uint32_t pixel, sample;
        slot_ids(p, slot, &pixel, &sample);
        float4 sv = p.st_v[slot];
        Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
        float4* st = reinterpret_cast<float4*>(p.ev + slot);
        if (e == 0) { st[2] = make_float4(0.f, 0.f, 0.f, 0.f); st[3] = make_float4(0.f, 0.f, 0.f, 0.f); }
        EmitterSample es = sample_emitter(p.sc, rng);  // selectEmitter + sampleEmitterPosition
        V3 wiW = normalize(es.pos - v.p);
        V3 wiL = to_local(v.f, wiW);
        V3 dlt = es.pos - v.p;
        float d2 = dot(dlt, dlt);
        float cosOut = dot(-wiW, es.n);
        if (cosOut > 0.f && wiL.z > 0.f) {
            st[0] = make_float4(wiL.x, wiL.y, wiL.z, cosOut * (1.f / d2));
            st[1] = make_float4(es.sel_pdf, es.area_pdf, __int_as_float(es.id), 0.f);
            p.ray[slot] = pt_ray(v.p, wiW);
            go = true;
        }
        sv.z = __uint_as_float(rng.ctr);
        p.st_v[slot] = sv;
    }
    uint32_t idx = queue_push(n_try, go);
    if (go) q_try[idx] = slot;
}

__global__ void __launch_bounds__(TRB_BLOCK) k_pt_emitter_resolve(const __grid_constant__ WaveParams p, PtParams pt, int depth,
                                                                  const int32_t* __restrict__ q_try, const uint32_t* __restrict__ n_try) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= *n_try) return;
    const int slot = q_try[k];
    RayT r = p.ray[slot];
    HitT h = p.hit[slot];
    if (!hit_valid(h, r)) return;
    const float4* s = reinterpret_cast<const float4*>(p.sc.shade + h.pos);
    const int vmat = __float_as_int(__ldg(s + 2).w), vem = __float_as_int(__ldg(s + 3).w);
    float4* st = reinterpret_cast<float4*>(p.ev + slot);
    float4 a = st[0], b = st[1];
    if (vem != __float_as_int(b.z)) return;  // visibilityInteraction.shapeID == emitter.shapeID
    PtVertex v = pt_load_vertex(p, depth, slot, nullptr, nullptr, nullptr, nullptr);
    Mat m = load_mat_stored(p.sc, v.mat, p.lv_tex + size_t(depth) * p.P + slot);
    V3 Li = pt_vis_emission(p, true, vmat);
    V3 wiL = mk(a.x, a.y, a.z);
    float a2s = a.w, emitterPdf = b.x, emitterAreaPdf = b.y;
    const Lobe lobe = glossy_lobe(m, v.wo, wiL);   // one powf for the pdf and the value (bsdf.cuh)
    float bsdfPdf = bsdf_pdf(m, v.wo, wiL, lobe);
    float f = float(pt.emitter_samples) * (emitterAreaPdf * emitterPdf * (1.f / a2s)), g = float(pt.bsdf_samples) * bsdfPdf;
    float weight = f / (f + g);
    V3 c = weight * Li * bsdf_eval(m, v.wo, wiL, lobe) * (1.f / emitterAreaPdf) * (1.f / emitterPdf) * a2s;
    float4 acc = st[2];
    st[2] = make_float4(acc.x + c.x, acc.y + c.y, acc.z + c.z, 0.f);
}

// ---- direct illumination by BSDF sampling (path.h:156-190), sample b of vertex `depth` ----------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_bsdf_gen(const __grid_constant__ WaveParams p, PtParams pt, int depth,
                                                           int32_t* __restrict__ q_try, uint32_t* __restrict__ n_try) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false;
    if (slot < p.P && p.nlv[slot] > depth) {
        PtVertex v = pt_load_vertex(p, depth, slot, nullptr, nullptr, nullptr, nullptr);
        Mat m = load_mat_stored(p.sc, v.mat, p.lv_tex + size_t(depth) * p.P + slot);
        uint32_t pixel, sample;
        slot_ids(p, slot, &pixel, &sample);
        float4 sv = p.st_v[slot];
        Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
        float ux = rng.next(), uy = rng.next();
        V3 wi; float pdf;
        V3 fcos = bsdf_sample(m, v.wo, ux, uy, &wi, &pdf);
        if (!is_zero(fcos)) {
            float4* st = reinterpret_cast<float4*>(p.ev + slot);
            st[0] = make_float4(fcos.x, fcos.y, fcos.z, pdf);
            p.ray[slot] = pt_ray(v.p, to_world(v.f, wi));
            go = true;
        }
        sv.z = __uint_as_float(rng.ctr);
        p.st_v[slot] = sv;
    }
    uint32_t idx = queue_push(n_try, go);
    if (go) q_try[idx] = slot;
}

__global__ void __launch_bounds__(TRB_BLOCK) k_pt_bsdf_resolve(const __grid_constant__ WaveParams p, PtParams pt, int depth,
                                                               const int32_t* __restrict__ q_try, const uint32_t* __restrict__ n_try) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= *n_try) return;
    const int slot = q_try[k];
    RayT r = p.ray[slot];
    HitT h = p.hit[slot];
    if (!hit_valid(h, r)) return;
    Surf vis = make_surf(p.sc, h, r);
    V3 Li = pt_vis_emission(p, true, vis.mat);
    if (is_zero(Li)) return;
    if (vis.emitter < 0) return;  // (an emissive face on a shape that is no emitter: getEmitterByID would be out of range in the reference)
    const float4* ep = reinterpret_cast<const float4*>(p.sc.emitters + vis.emitter);
    float emitterPdf = 1.f / float(p.sc.n_emitters), emitterAreaPdf = 1.f / __ldg(ep).w;
    PtVertex v = pt_load_vertex(p, depth, slot, nullptr, nullptr, nullptr, nullptr);
    V3 dlt = v.p - vis.p;   // glm::distance2(visibilityInteraction.p, hit.p)
    float d2 = dot(dlt, dlt);
    const F4* tp = p.sc.tris + 3 * size_t(h.pos);
    F4 q1 = ld4(tp + 1), q2 = ld4(tp + 2);
    V3 ng = normalize(cross(mk(q1.x, q1.y, q1.z), mk(q2.x, q2.y, q2.z)));  // frameNg.n (accel.h:161)
    float cosOut = dot(mk(-r.dx, -r.dy, -r.dz), ng);
    if (!(cosOut > 0.f)) return;
    float4* st = reinterpret_cast<float4*>(p.ev + slot);
    float4 a = st[0];
    float a2s = cosOut * (1.f / d2);
    float f = float(pt.bsdf_samples) * a.w, g = float(pt.emitter_samples) * (emitterPdf * emitterAreaPdf * (1.f / a2s));
    float weight = f / (f + g);
    V3 c = weight * Li * mk(a.x, a.y, a.z) * (1.f / a.w);
    float4 acc = st[3];
    st[3] = make_float4(acc.x + c.x, acc.y + c.y, acc.z + c.z, 0.f);
}

// ---- tail of the call at `depth` (path.h:153,190-203): Lr = direct + indirect, roulette scale; at depth 0: the sample's value ----
__global__ void __launch_bounds__(TRB_BLOCK) k_pt_combine(const __grid_constant__ WaveParams p, PtParams pt, int depth) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= p.P || p.nlv[slot] <= depth) return;
    float4 c4, d4; float pdf; uint32_t ns;
    pt_load_vertex(p, depth, slot, &c4, &d4, &pdf, &ns);
    const int flags = __float_as_int(c4.w);
    const float4* st = reinterpret_cast<const float4*>(p.ev + slot);
    float4 ee = st[2], be = st[3];
    V3 emitterEstimator = pt.emitter_samples == 0 ? mk(0.f) : mk(ee.x, ee.y, ee.z) / float(pt.emitter_samples);
    V3 bsdfEstimator = pt.bsdf_samples == 0 ? mk(0.f) : mk(be.x, be.y, be.z) / float(pt.bsdf_samples);
    V3 direct = emitterEstimator + bsdfEstimator;
    V3 indirect = mk(0.f);
    if (flags & PT_HAS_CHILD) {
        float4 lr = p.st_t[slot];
        const float cum = ns > 1u ? 0.95f : 1.f;
        indirect = mk(lr.x, lr.y, lr.z) * mk(d4.x, d4.y, d4.z) * (1.f / pdf) * (1.f / float(ns)) * (1.f / cum);
    }
    V3 Lr = direct + indirect;
    if (flags & PT_RR_SCALED) Lr = Lr * (1.f / pt.rr_prob);
    if (depth == 0) pt_accum(p, slot, Lr);
    else p.st_t[slot] = make_float4(Lr.x, Lr.y, Lr.z, 0.f);
}

}  // namespace trbk
