// bdpt_kernels.cuh -- the BDPT wavefront: one launch per stage over a wave of P samples.
// Replaces BDPTIntegrator (src/integrators/bdpt.h:37-518) and the offline pixel loop of
// Renderer::render (src/core/renderer.cpp:130-214).  Stage order per wave (host side: trb_abi.cu):
//   gen_camera -> trace_closest -> light_init -> K x [trace_closest, light_bounce, trace_any+resolve]
//   -> eye_init -> K x [(trace_closest), eye_bounce, trace_any+resolve]
// Live samples are compacted into queues with warp-ballot pushes; every shadow ray carries its pending
// contribution, so one resolve kernel serves light-tracing splats, NEE and vertex connections alike.
#pragma once
#include "bdpt_common.cuh"

namespace trbk {

#define TRB_BLOCK 128

__device__ __forceinline__ BvhView bvh_of(const SceneDev& sc) { return BvhView{sc.nodes, sc.tris, sc.root_ref}; }

// (the traversal kernels live in trace_wide.cuh / trace_coop.cuh)

// ---- camera rays (renderer.cpp:161-192) ------------------------------------------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_gen_camera(const __grid_constant__ WaveParams p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.P) return;
    uint32_t ctr = 0;
    RayT r;
    if (p.parity) {
        r = p.inject_rays[i];
    } else {
        uint32_t pixel, sample;
        slot_ids(p, i, &pixel, &sample);
        Rng rng = make_rng(p, i, pixel, sample, 0);
        int j = int(pixel % uint32_t(p.cam.width)), row = int(pixel / uint32_t(p.cam.width));
        float y = (1.f - ((float)row + 0.5f) * p.cam.inv_h) * 2.f - 1.f;
        float x = (((float)j + 0.5f) * p.cam.inv_w) * 2.f - 1.f;
        float px, py;
        if (p.spp == 1 || (p.flags & TRB_FLAG_NO_JITTER)) {
            px = x * p.cam.tan_half_fov * p.cam.aspect; py = y * p.cam.tan_half_fov;
        } else {
            float rx = rng.next(), ry = rng.next();
            rx -= 0.5f; ry -= 0.5f;
            rx = rx * p.cam.inv_w; ry = ry * p.cam.inv_h;
            px = (x + rx) * p.cam.tan_half_fov * p.cam.aspect; py = (y + ry) * p.cam.tan_half_fov;
        }
        float4 d4 = mat_mul(p.cam.c2w, make_float4(px, py, -1.f, 0.f));
        V3 d = normalize(mk(d4.x, d4.y, d4.z));
        r.ox = p.cam.eye[0]; r.oy = p.cam.eye[1]; r.oz = p.cam.eye[2];
        r.dx = d.x; r.dy = d.y; r.dz = d.z; r.min_t = 1.f; r.max_t = 1000.f;
        ctr = rng.ctr;
    }
    p.cam_ray[i] = r;
    p.st_v[i] = make_float4(0.f, __uint_as_float(0u), __uint_as_float(ctr), 0.f);
}

// ---- Russian roulette (bdpt.h:68,129,188,201 with NO_RR 0; the shipped build has NO_RR 1 => rr == 1) --------
// Continuation probability stored with the vertex reached at `depth`, from the throughput ON ARRIVAL.
__device__ __forceinline__ float rr_prob(const WaveParams& p, int depth, V3 T) {
    if (!p.rr_on) return 1.f;
    return (depth + 1) < p.rr_depth ? 1.f : (luminance(T) < 0.01f ? 0.5f : 1.f);
}
// The loop test `depth < rrDepth || (sampler.next() < rr && !NO_RR)` evaluated for the NEXT iteration: past rrDepth it
// always draws one number (Q3), and continues only under Russian roulette (and below the device's hard cap).
__device__ __forceinline__ bool rr_continue(const WaveParams& p, int depth_next, float rr, Rng& rng) {
    if (depth_next < p.rr_depth) return true;
    float u = rng.next();
    return p.rr_on && u < rr && depth_next < p.max_depth;
}

// ---- connectToCamera (bdpt.h:295-371) + splatToImagePlane (bdpt.h:485-496) ---------------------------------
__device__ __forceinline__ void connect_to_camera(const WaveParams& p, bool active, int slot, int depth, const Surf& s, const Mat& m,
                                                  V3 T, float vcm, float vc, float rr) {
    bool want = false;
    RayT sr; V3 radiance = mk(0.f); uint32_t target = 0;
    if (active) {
        V3 eye = mk(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
        V3 fwd = mk(p.cam.forward[0], p.cam.forward[1], p.cam.forward[2]);
        V3 e2l = s.p - eye;
        float invD2 = 1.f / length2(e2l);
        e2l = e2l * sqrtf(invD2);
        float4 uv = mat_mul(p.cam.w2c, make_float4(s.p.x, s.p.y, s.p.z, 1.f));
        uv = mat_mul(p.cam.clip, uv);
        uv = make_float4(uv.x / uv.w, uv.y / uv.w, uv.z / uv.w, uv.w / uv.w);
        uv = mat_mul(p.cam.ndc2screen, uv);
        int xp = __float2int_rz(uv.x), yp = __float2int_rz(uv.y);
        bool ok = !(xp < 0 || yp < 0 || xp >= p.cam.width || yp >= p.cam.height);
        float cosCamera = dot(fwd, e2l);
        ok = ok && !(cosCamera <= 0.f);
        if (ok) {
            V3 wi = to_local(s.f, -e2l);
            const Lobe lobe = glossy_lobe(m, s.wo, wi);   // one powf for the value and the reverse pdf (bsdf.cuh)
            V3 fcos = bsdf_eval(m, s.wo, wi, lobe);
            if (!(is_zero(fcos) || wi.z <= 0.f)) {
                sr = visibility_ray(eye, s.p);
                float ipd = p.cam.vnear_dist / cosCamera;
                float imgToSolid = ipd * ipd / cosCamera;
                float solidToArea = wi.z * invD2;
                float imgToSurf = imgToSolid * solidToArea;
                float surfToImg = 1.f / imgToSurf;
                int nlight = p.cam.width * p.cam.height;
                radiance = T * (fcos * (1.f / wi.z)) * (1.f / surfToImg) * (1.f / nlight) * (1.f / p.spp);
                float revPdfA = 1.f * imgToSurf;
                float prevRev = bsdf_pdf(m, wi, s.wo, lobe) * rr;  // wi/wo swapped (bdpt.h:342)
                float lightWeight = (revPdfA / nlight) * (vcm + prevRev * vc);
                float mis = 1.f / (lightWeight + 1.f + 0.f);
                if (!(p.flags & TRB_EST_MASK)) radiance = radiance * mis;   // bdpt.h:355-357: unweighted under LIGHT_TRACING / PATH_TRACING
                int pixel = yp * p.cam.width + xp;
                if (p.parity) {
                    target = uint32_t(slot) * uint32_t(p.K + 1) + uint32_t(depth);
                    p.splat_pix[size_t(slot) * p.K + (depth - 1)] = pixel;
                } else {
                    target = uint32_t(pixel);
                }
                want = true;
            }
        }
    }
    push_shadow(p, want, sr, radiance, target);
}

// ---- ContinuePathRandomWalk (bdpt.h:243-291); returns false when the walk stops ------------------------------
__device__ __forceinline__ bool continue_walk(const Surf& s, const Mat& m, Rng& rng, float rr, V3* T, float* vc, float* vcm, RayT* next, V3* wi_out) {
    float ux = rng.next(), uy = rng.next();
    V3 wi; float pdf;
    Lobe lobe;
    V3 fcos = bsdf_sample(m, s.wo, ux, uy, &wi, &pdf, &lobe);
    pdf *= rr;
    *wi_out = wi;
    float absCosOut = fabsf(wi.z);
    if (is_zero(fcos)) return false;
    *T = *T * (fcos * (1.f / pdf));
    bool delta = is_delta(m);
    float prevRev = delta ? pdf : bsdf_pdf(m, wi, s.wo, lobe) * rr;
    if (delta) {
        *vc = (absCosOut / pdf) * (prevRev * *vc);
        *vcm = 0.f;
    } else {
        *vc = (absCosOut / pdf) * (*vcm + prevRev * *vc);
        *vcm = 1.f / pdf;
    }
    V3 d = to_world(s.f, wi);
    next->ox = s.p.x; next->oy = s.p.y; next->oz = s.p.z; next->dx = d.x; next->dy = d.y; next->dz = d.z;
    next->min_t = 1e-8f; next->max_t = FLT_MAX;
    return true;
}

// ---- light subpath start (bdpt.h:158-186) -------------------------------------------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_light_init(const __grid_constant__ WaveParams p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool in = i < p.P;
    bool go = false;
    RayT lr_out; float4 st_out = make_float4(0.f, 0.f, 0.f, 0.f), sv = st_out;
    lr_out.ox = lr_out.oy = lr_out.oz = lr_out.dx = lr_out.dy = lr_out.dz = lr_out.min_t = lr_out.max_t = 0.f;
    if (in) {
        RayT cr = p.cam_ray[i];
        HitT h = p.hit[i];
        p.hit0[i] = h;
        p.nlv[i] = 0;
        sv = p.st_v[i];
        uint32_t ctr = __float_as_uint(sv.z);
        uint32_t flags = 0;
        if (hit_valid(h, cr)) flags |= 2u;   // bdpt.h:225: no subpaths at all when the primary ray misses
        if ((flags & 2u) && !(p.flags & TRB_FLAG_PATH_TRACING)) {   // (PATH_TRACING, bdpt.h:232-233: no light walk, no draws)
            uint32_t pixel, sample;
            slot_ids(p, i, &pixel, &sample);
            Rng rng = make_rng(p, i, pixel, sample, ctr);
            EmitterSample es = sample_emitter(p.sc, rng);
            float ux = rng.next(), uy = rng.next();
            V3 ed = sq_to_uniform_hemisphere(ux, uy);
            float emissionPdf = TRB_INV_TWOPI * es.area_pdf;
            float areaPdf = es.area_pdf * es.sel_pdf;
            emissionPdf *= es.sel_pdf;
            Frame lf = make_frame(es.n);
            V3 wd = to_world(lf, ed);
            V3 T = ed.z * es.radiance * (1.f / emissionPdf);
            float vc = ed.z * (1.f / emissionPdf);
            float vcm = areaPdf / emissionPdf;
            if (!(ed.z <= 0.f)) {
                if (rr_continue(p, 1, 1.f, rng)) {  // the loop test of bdpt.h:188 at depth 1 (draws when rrDepth <= 1, Q3)
                    go = true;
                    RayT lr; lr.ox = es.pos.x; lr.oy = es.pos.y; lr.oz = es.pos.z; lr.dx = wd.x; lr.dy = wd.y; lr.dz = wd.z;
                    lr.min_t = 1e-8f; lr.max_t = FLT_MAX;
                    lr_out = lr;
                    st_out = make_float4(T.x, T.y, T.z, vcm);
                    sv.x = vc;
                }
            }
            ctr = rng.ctr;
        }
        sv.y = __uint_as_float(flags); sv.z = __uint_as_float(ctr);
        p.st_v[i] = sv;
    }
    uint32_t idx = queue_push(p.counters + CTR_QOUT, go);
    if (go) { p.q_out[idx] = i; p.ray_out[idx] = lr_out; p.stt_out[idx] = st_out; p.sv_out[idx] = sv; }
}

// ---- launch shapes of the bounce kernels ---------------------------------------------------------------------
#ifndef TRB_SORT_BLOCK
#define TRB_SORT_BLOCK 256
#endif
// Launch bounds of the bounce / connect kernels.  The bounce kernels wait on the triangle and material their hit names (ncu r1o: long scoreboard,
// 23-35 % occupancy), so resident warps count for more than a few spilled registers: capping k_eye_bounce for 3 blocks per SM
// (92 -> 80 registers) and k_light_bounce for 4 (78 -> 64) takes the logic share of a caustic step from 35.2 to 31.6 ms.  Tighter
// caps (eye 64, light 48, connect 72 / 64) and no cap at all (94 / 86 / 94) are slower (profiles/r1o_ncu_logic_summary.md).
#ifndef TRB_LB_MINB
#define TRB_LB_MINB 4
#endif
#ifndef TRB_EB_MINB
#define TRB_EB_MINB 3
#endif
#define TRB_LB_BOUNDS __launch_bounds__(TRB_SORT_BLOCK, TRB_LB_MINB)
#define TRB_EB_BOUNDS __launch_bounds__(TRB_SORT_BLOCK, TRB_EB_MINB)
#ifndef TRB_CV_COMPACT
#define TRB_CV_COMPACT 1
#endif
#ifdef TRB_CV_MINB
#define TRB_CV_BOUNDS __launch_bounds__(TRB_BLOCK, TRB_CV_MINB)
#else
#define TRB_CV_BOUNDS __launch_bounds__(TRB_BLOCK)
#endif
// ---- one light bounce (body of the loop at bdpt.h:188-216) -------------------------------------------------
__global__ void TRB_LB_BOUNDS k_light_bounce(const __grid_constant__ WaveParams p, int depth, const uint32_t* __restrict__ n_in) {
    uint32_t n = *n_in;
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x * blockDim.x >= n) return;  // block-uniform
    const bool in = k < n;
    int slot = 0;
    bool alive = false, connect = false, go = false;
    Surf s; Mat m; V3 T = mk(0.f); float vcm = 0.f, vc = 0.f;
    float4 sv = make_float4(0.f, 0.f, 0.f, 0.f);
    RayT r;
    if (in) {
        slot = p.q_in[k];        // the walk's records are read by queue position: four coalesced streams
        r = p.ray_in[k];
        HitT h = p.hit[k];
        sv = p.sv_in[k];
        const float4 st = p.stt_in[k];
        if (hit_valid(h, r)) {
            alive = true;
            s = make_surf(p.sc, h, r);
            m = load_mat_at(p.sc, s.mat, h);
            T = mk(st.x, st.y, st.z); vcm = st.w; vc = sv.x;
            const float distSquared = s.t * s.t;
            const float absCosIn = fabsf(s.wo.z);
            vcm *= (distSquared / absCosIn);
            vc *= (1.f / absCosIn);
            connect = !is_delta(m);
        }
    }
    const float rr = rr_prob(p, depth, T);
    connect_to_camera(p, connect, slot, depth, s, m, T, vcm, vc, rr);
    if (alive) {
        uint32_t pixel, sample;
        slot_ids(p, slot, &pixel, &sample);
        Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
        V3 Tn = T; float vcn = vc, vcmn = vcm; RayT nr; V3 wi;
        if (continue_walk(s, m, rng, rr, &Tn, &vcn, &vcmn, &nr, &wi)) {
            if (connect) {  // non-delta vertices whose continuation succeeded are kept (bdpt.h:211-215, Q11)
                int j = p.nlv[slot];
                LightVertex lv;
                lv.px = s.p.x; lv.py = s.p.y; lv.pz = s.p.z; lv.vcm = vcm;
                lv.nx = s.n.x; lv.ny = s.n.y; lv.nz = s.n.z; lv.vc = vc;
                lv.wox = s.wo.x; lv.woy = s.wo.y; lv.woz = s.wo.z; lv.mat = s.mat;
                lv.tx = T.x; lv.ty = T.y; lv.tz = T.z; lv.rr = rr;
                p.lv[size_t(slot) * p.K + j] = lv;   // slot-major: the pairs of one eye vertex read one contiguous run (bdpt_common.cuh)
                if (p.sc.has_tex) p.lv_tex[size_t(slot) * p.K + j] = tex_record(p.hit[k]);
                p.nlv[slot] = j + 1;
            }
            if (rr_continue(p, depth + 1, rr, rng)) {
                go = true;
                r = nr;
                T = Tn; vcm = vcmn;
                sv.x = vcn;
            }
        }
        sv.z = __uint_as_float(rng.ctr);
    }
    uint32_t idx = queue_push(p.counters + CTR_QOUT, go);
    if (go) { p.q_out[idx] = slot; p.ray_out[idx] = r; p.stt_out[idx] = make_float4(T.x, T.y, T.z, vcm); p.sv_out[idx] = sv; }
    else if (in) p.st_v[slot] = sv;   // the walk ends here: what the eye walk (flags, draw count) finds under the slot
}

// ---- eye subpath start (bdpt.h:46-66) ------------------------------------------------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_eye_init(const __grid_constant__ WaveParams p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false;
    RayT r_out; float4 st_out = make_float4(0.f, 0.f, 0.f, 0.f), sv = st_out;
    r_out.ox = r_out.oy = r_out.oz = r_out.dx = r_out.dy = r_out.dz = r_out.min_t = r_out.max_t = 0.f;
    if (i < p.P) {
        sv = p.st_v[i];
        uint32_t flags = __float_as_uint(sv.y);
        if (flags & 2u) {
            RayT cr = p.cam_ray[i];
            V3 fwd = mk(p.cam.forward[0], p.cam.forward[1], p.cam.forward[2]);
            float cosCamera = dot(fwd, mk(cr.dx, cr.dy, cr.dz));
            float ipd = p.cam.vnear_dist / cosCamera;
            float imgToSolid = ipd * ipd / cosCamera;
            float t1Pdf = 1.f * imgToSolid;
            float vcm = float(p.cam.width * p.cam.height) * (1.f / t1Pdf);
            bool cont = 1 < p.rr_depth || (p.flags & TRB_FLAG_LIGHT_TRACING);
            if (!cont) {  // the loop test of bdpt.h:68 draws (Q3); with rr == 1 the draw always passes
                sv.z = __uint_as_float(__float_as_uint(sv.z) + 1u);
                cont = p.rr_on && 1 < p.max_depth;
            }
            if (cont) {
                go = true;
                r_out = cr;
                st_out = make_float4(1.f, 1.f, 1.f, vcm);
                sv.x = 0.f;
                sv.y = __uint_as_float(flags | SLOT_PURE_SPECULAR);
            }
            p.st_v[i] = sv;
        }
    }
    uint32_t idx = queue_push(p.counters + CTR_QOUT, go);
    if (go) { p.q_out[idx] = i; p.ray_out[idx] = r_out; p.hit[idx] = p.hit0[i]; p.stt_out[idx] = st_out; p.sv_out[idx] = sv; }
}

// ---- one eye bounce (body of the loop at bdpt.h:68-154) --------------------------------------------------------
__global__ void TRB_EB_BOUNDS k_eye_bounce(const __grid_constant__ WaveParams p, int depth, const uint32_t* __restrict__ n_in) {
    uint32_t n = *n_in;
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x * blockDim.x >= n) return;  // block-uniform
    const bool in = k < n;
    int slot = 0;
    bool alive = false, connect = false, go = false;
    Surf s; Mat m; V3 T = mk(0.f); float vcm = 0.f, vc = 0.f;
    float4 sv = make_float4(0.f, 0.f, 0.f, 0.f);
    RayT nr_out;
    nr_out.ox = nr_out.oy = nr_out.oz = nr_out.dx = nr_out.dy = nr_out.dz = nr_out.min_t = nr_out.max_t = 0.f;
    uint32_t pixel = 0, sample = 0, li_target = 0;
    const float li_scale = p.parity ? 1.f : (1.f / p.spp);  // renderer.cpp:201: averageRadiance * (1.f / spp)
    Rng rng; rng.inj = nullptr; rng.key = 0; rng.key2 = 0; rng.ctr = 0; rng.lim = 0;
    if (in) {
        slot = p.q_in[k];        // the walk's records are read by queue position: four coalesced streams
        RayT r = p.ray_in[k];
        HitT h = p.hit[k];
        sv = p.sv_in[k];
        const float4 st = p.stt_in[k];
        slot_ids(p, slot, &pixel, &sample);
        li_target = p.parity ? uint32_t(slot) * uint32_t(p.K + 1) : pixel;
        rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
        if (hit_valid(h, r)) {
            alive = true;
            s = make_surf(p.sc, h, r);
            m = load_mat_at(p.sc, s.mat, h);
            T = mk(st.x, st.y, st.z); vcm = st.w; vc = sv.x;
            const float distSquared = s.t * s.t;
            const float absCosIn = fabsf(s.wo.z);
            vcm *= (distSquared / absCosIn);
            vc *= (1.f / absCosIn);
            if (m.emissive) {  // bdpt.h:79-125: emitter hit ends the walk
                bool pure = (__float_as_uint(sv.y) & SLOT_PURE_SPECULAR) != 0;
                V3 c;
                if (depth > 1) {
                    const float4* ep = reinterpret_cast<const float4*>(p.sc.emitters + s.emitter);
                    float4 e0 = __ldg(ep);
                    float emitterPdf = 1.f / float(p.sc.n_emitters);
                    c = mk(e0.x, e0.y, e0.z) * T;
                    float posPdfA = 1.f / (e0.w * emitterPdf);
                    float dirPdfW = TRB_INV_TWOPI;
                    float cameraWeight = posPdfA * vcm + (posPdfA * dirPdfW) * vc;
                    float mis = 1.f / (1.f + cameraWeight);
                    if (!pure && !(p.flags & TRB_EST_MASK)) c = c * mis;                 // bdpt.h:95-100
                    if ((p.flags & TRB_FLAG_PATH_TRACING) && !pure) c = mk(0.f, 0.f, 0.f);   // bdpt.h:110-113: NEE already counted this light path
                } else {
                    c = m.ke;
                }
                c = c * li_scale;
                float* a = p.accum + size_t(li_target) * 3;
                atomicAdd(a, c.x); atomicAdd(a + 1, c.y); atomicAdd(a + 2, c.z);
                alive = false;
            } else {
                connect = !is_delta(m);
            }
            if (p.flags & TRB_FLAG_LIGHT_TRACING) { alive = false; connect = false; }   // bdpt.h:229-231: Li = getEmission(hit), no eye walk
        }
    }
    const float rr = rr_prob(p, depth, T);
    // ---- connectToLight (bdpt.h:374-430) ----
    {
        bool want = false; RayT sr; V3 Li = mk(0.f);
        if (connect) {
            EmitterSample es = sample_emitter(p.sc, rng);
            V3 l2e = s.p - es.pos;
            float d2 = length2(l2e);
            l2e = l2e * (1.f / sqrtf(d2));
            V3 wi = to_local(s.f, -l2e);
            float cosAtLight = dot(es.n, l2e);
            float cosAtEye = wi.z;
            if (!(cosAtLight <= 0.f || cosAtEye <= 0.f)) {
                float connA = es.sel_pdf * es.area_pdf;
                float connW = connA * d2 / cosAtLight;
                float dirPdfW = TRB_INV_TWOPI;
                const Lobe lobe = glossy_lobe(m, s.wo, wi);
                Li = bsdf_eval(m, s.wo, wi, lobe) * (1.f / connW) * T * es.radiance;
                if (!is_zero(Li)) {
                    sr = visibility_ray(s.p, es.pos);
                    float lightRevW = bsdf_pdf(m, s.wo, wi, lobe) * rr;
                    float lightWeight = lightRevW / connW;
                    float eyePrevRevW = bsdf_pdf(m, wi, s.wo, lobe) * rr;
                    float eyeCurRevA = cosAtEye * (1.f / d2) * dirPdfW;
                    float eyeWeight = eyeCurRevA * (vcm + eyePrevRevW * vc);
                    float mis = 1.f / (lightWeight + 1.f + eyeWeight);
                    if (!(p.flags & TRB_EST_MASK)) Li = Li * mis;   // bdpt.h:426-428
                    Li = Li * li_scale;
                    want = true;
                }
            }
        }
        push_shadow(p, want, sr, Li, li_target);
    }
    // ---- connectVertices (bdpt.h:434-483): handed to k_connect_vertices as a DENSE list of (slot, light vertex) pairs ----
    // A slot has 0..K stored light vertices (3-4 on average): one thread per (slot, j < K) left most lanes of the pair kernel
    // without work.  Each connecting lane reserves nlv entries (warp scan + one atomic per warp) and writes them in order.
    {
        const int nl = (connect && !(p.flags & TRB_EST_MASK)) ? p.nlv[slot] : 0;   // bdpt.h:145-149
        int inc = nl;
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        const int total = __shfl_sync(0xffffffffu, inc, 31);
        if (total) {
            uint32_t base = 0;
            if (lane == 31) base = atomicAdd(p.counters + CTR_CONN, uint32_t(total));
            base = __shfl_sync(0xffffffffu, base, 31) + uint32_t(inc - nl);
            for (int j = 0; j < nl; ++j) p.q_conn[base + uint32_t(j)] = (slot << 6) | j;
        }
        if (nl > 0) {
            EyeVertex e;
            e.px = s.p.x; e.py = s.p.y; e.pz = s.p.z; e.vcm = vcm;
            e.nx = s.n.x; e.ny = s.n.y; e.nz = s.n.z; e.vc = vc;
            e.wox = s.wo.x; e.woy = s.wo.y; e.woz = s.wo.z; e.mat = s.mat | (rr < 1.f ? EV_RR_HALF : 0);
            e.tx = T.x; e.ty = T.y; e.tz = T.z; e.target = li_target;
            p.ev[slot] = e;
            if (p.sc.has_tex) p.ev_tex[slot] = tex_record(p.hit[k]);
        }
    }
    if (in) {
        if (alive) {
            if (connect) sv.y = __uint_as_float(__float_as_uint(sv.y) & ~SLOT_PURE_SPECULAR);
            V3 Tn = T; float vcn = vc, vcmn = vcm; RayT nr; V3 wi;
            if (continue_walk(s, m, rng, rr, &Tn, &vcn, &vcmn, &nr, &wi)) {
                if (rr_continue(p, depth + 1, rr, rng)) {
                    go = true;
                    nr_out = nr;
                    T = Tn; vcm = vcmn;
                    sv.x = vcn;
                }
            }
        }
        sv.z = __uint_as_float(rng.ctr);
    }
    uint32_t idx = queue_push(p.counters + CTR_QOUT, go);
    if (go) { p.q_out[idx] = slot; p.ray_out[idx] = nr_out; p.stt_out[idx] = make_float4(T.x, T.y, T.z, vcm); p.sv_out[idx] = sv; }
    else if (in) p.st_v[slot] = sv;   // the walk ends here: the draw count the parity entry point reads
}

// ---- connectVertices (bdpt.h:434-483): one thread per (eye vertex, stored light vertex) pair of the dense list k_eye_bounce wrote.
// Replaces the per-thread loop over the sample's light vertices (profiles/r1e: 7 active threads per instruction in k_eye_bounce). ----
// TEX: the scene has bitmap textures (their lookup is compiled out of the untextured instance, which every shipped scene runs)
// the two cosines of a connection, exactly as cv_evaluate forms them: they decide whether the pair is evaluated at all (bdpt.h:447-449)
__device__ __forceinline__ bool cv_cosines_pass(const WaveParams& p, int pair) {
    const int slot = pair >> 6;
    const uint32_t j = uint32_t(pair & 63);
    const float4* ep = reinterpret_cast<const float4*>(p.ev + slot);
    const float4 ea = ep[0], eb = ep[1];
    const float4* lp = reinterpret_cast<const float4*>(p.lv + size_t(slot) * p.K + j);
    const float4 a = lp[0], b = lp[1];
    V3 l2e = mk(ea.x, ea.y, ea.z) - mk(a.x, a.y, a.z);
    const float invD2 = 1.f / length2(l2e);
    l2e = l2e * sqrtf(invD2);
    const float cosL = to_local(make_frame(mk(b.x, b.y, b.z)), l2e).z, cosE = to_local(make_frame(mk(eb.x, eb.y, eb.z)), -l2e).z;
    return !(cosL <= 0.f || cosE <= 0.f);
}
// one (eye vertex, light vertex) pair: the contribution and its shadow ray; false when there is nothing to trace
template <bool TEX>
__device__ __forceinline__ bool cv_evaluate(const WaveParams& p, int pair, float li_scale, RayT* sr, V3* Li_out, uint32_t* target) {
    const int slot = pair >> 6;
    const uint32_t j = uint32_t(pair & 63);
    const float4* ep = reinterpret_cast<const float4*>(p.ev + slot);
    float4 ea = ep[0], eb = ep[1], ec = ep[2], ed = ep[3];
    const float4* lp = reinterpret_cast<const float4*>(p.lv + size_t(slot) * p.K + j);
    float4 a = lp[0], b = lp[1], c = lp[2], d = lp[3];
    *target = __float_as_uint(ed.w);
    V3 epnt = mk(ea.x, ea.y, ea.z), lpnt = mk(a.x, a.y, a.z);
    V3 l2e = epnt - lpnt;
    float invD2 = 1.f / length2(l2e);
    l2e = l2e * sqrtf(invD2);
    Frame lf = make_frame(mk(b.x, b.y, b.z));
    Frame ef = make_frame(mk(eb.x, eb.y, eb.z));
    V3 l_wo = mk(c.x, c.y, c.z), e_wo = mk(ec.x, ec.y, ec.z);
    V3 l_wi = to_local(lf, l2e);
    V3 e_wi = to_local(ef, -l2e);
    float cosL = l_wi.z, cosE = e_wi.z;
    if (cosL <= 0.f || cosE <= 0.f) return false;
    Mat lm = load_mat(p.sc.mats, __float_as_int(c.w));
    if (TEX && lm.tex) { const float4 t = p.lv_tex[size_t(slot) * p.K + j]; apply_tex(p.sc, lm, __float_as_int(t.x), t.y, t.z); }
    const int emat = __float_as_int(ec.w);
    const float lrr = d.w, err = (emat & EV_RR_HALF) ? 0.5f : 1.f;  // PathVertex::rr of the two vertices
    Mat m = load_mat(p.sc.mats, emat & ~EV_RR_HALF);
    if (TEX && m.tex) { const float4 t = p.ev_tex[slot]; apply_tex(p.sc, m, __float_as_int(t.x), t.y, t.z); }
    const Lobe ll = glossy_lobe(lm, l_wo, l_wi), le = glossy_lobe(m, e_wo, e_wi);   // two powf instead of six
    V3 Li = bsdf_eval(lm, l_wo, l_wi, ll) * bsdf_eval(m, e_wo, e_wi, le);
    Li = Li * (mk(d.x, d.y, d.z) * mk(ed.x, ed.y, ed.z) * invD2);
    float eyeRevW = bsdf_pdf(lm, l_wo, l_wi, ll) * lrr;
    float lightPrevRevW = bsdf_pdf(lm, l_wi, l_wo, ll) * lrr;
    float lightRevW = bsdf_pdf(m, e_wo, e_wi, le) * err;
    float eyePrevRevW = bsdf_pdf(m, e_wi, e_wo, le) * err;
    float lightRevA = lightRevW * (cosL) * invD2;
    float eyeRevA = eyeRevW * (cosE) * invD2;
    float lightWeight = lightRevA * (a.w + lightPrevRevW * b.w);
    float eyeWeight = eyeRevA * (ea.w + eyePrevRevW * eb.w);
    float mis = 1.f / (lightWeight + 1.f + eyeWeight);
    Li = Li * mis;
    Li = Li * li_scale;
    *Li_out = Li;
    if (is_zero(Li)) return false;
    *sr = visibility_ray(epnt, lpnt);
    return true;
}

template <bool TEX>
__global__ void TRB_CV_BOUNDS k_connect_vertices(const __grid_constant__ WaveParams p, const uint32_t* __restrict__ n_pairs_ptr) {
    const unsigned long long total = *n_pairs_ptr;
    const unsigned long long total_w = (total + 31ull) & ~31ull;  // whole warps take part in the ballots
    const float li_scale = p.parity ? 1.f : (1.f / p.spp);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
#if TRB_CV_COMPACT
    // Half the pairs leave at the cosine test and the rest cost ten times as much: a warp tests 32 pairs, keeps the survivors in a
    // shared-memory list and evaluates them 32 at a time with every lane busy (ncu before: 14 active threads per instruction).
    __shared__ int s_list[TRB_BLOCK / 32][64];
    int* list = s_list[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    int cnt = 0;   // warp-uniform
    for (; t < total_w; t += stride) {
        int pair = 0; bool pass = false;
        if (t < total) { pair = p.q_conn[t]; pass = cv_cosines_pass(p, pair); }   // (the pairs of one eye vertex are neighbours: one broadcast read of its record)
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (pass) list[cnt + __popc(m & lt)] = pair;
        cnt += __popc(m);
        __syncwarp();
        if (cnt >= 32) {
            cnt -= 32;
            RayT sr; V3 Li = mk(0.f); uint32_t target = 0;
            const bool want = cv_evaluate<TEX>(p, list[cnt + lane], li_scale, &sr, &Li, &target);
            push_shadow(p, want, sr, Li, target);
            __syncwarp();
        }
    }
    {
        RayT sr; V3 Li = mk(0.f); uint32_t target = 0;
        const bool want = lane < cnt && cv_evaluate<TEX>(p, list[lane], li_scale, &sr, &Li, &target);
        push_shadow(p, want, sr, Li, target);
    }
#else
    for (; t < total_w; t += stride) {
        bool want = false; RayT sr; V3 Li = mk(0.f); uint32_t target = 0;
        if (t < total) want = cv_evaluate<TEX>(p, p.q_conn[t], li_scale, &sr, &Li, &target);
        push_shadow(p, want, sr, Li, target);
    }
#endif
}

}  // namespace trbk
