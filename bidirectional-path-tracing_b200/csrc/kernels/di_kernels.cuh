// di_kernels.cuh -- the reference's DirectIntegrator (src/integrators/direct.h:16-467, `type = "direct"`) on the wavefront's
// buffers and traversal kernel: one-bounce direct illumination with the five sampling strategies of direct.h:449-462,
// written by the reference for SPHERE emitters (centre / radius = Scene::getShapeCenter / getShapeRadius,
// renderer.cpp:349-358; the host keeps them per emitter in SceneDev::esphere).  SURVEY 8(f) N4.
//
//   k_di_init     primary hit: emitter seen directly -> its radiance (direct.h:151-154); else the shading vertex is kept
//   per sample i of the emitter loop, then of the BSDF loop (the reference's order of random numbers):
//     k_di_gen     draws the sample's numbers, builds the ray, stashes what the estimator needs
//     k_trace_coop closest hit (every query of direct.h is scene.bvh->intersect)
//     k_di_resolve adds the sample's term to the running estimator (same order of additions as the reference's loop)
//   k_di_combine  Lr = estimator / n (direct.h:192,229,263,332) or the two MIS estimators (direct.h:404,439,442) -> frame buffer
// M_PI is platform.h:51's FLOAT constant, so `4 * M_PI * r * r` is fp32; raySphereIntersect (direct.h:16-69) is fp64 as written.
#pragma once
#include "pt_kernels.cuh"

namespace trbk {

struct DiParams {
    int32_t strategy;  // TRB_DI_*
    int32_t emitter_samples, bsdf_samples;
};
enum : int32_t { DI_MIS = 0, DI_AREA = 1, DI_SOLID_ANGLE = 2, DI_COSINE_HEMISPHERE = 3, DI_BSDF = 4 };

__device__ __forceinline__ V3 sq_to_uniform_sphere(float ux, float uy) {  // math.h:119-127
    float phi = ux * TRB_PI * 2.0f;
    float cosTheta = 1.f - (2.f * uy);
    float sinTheta = sqrtf(fmaxf(1.f - cosTheta * cosTheta, 0.f));
    float sn, cs; t_sincosf(phi, &sn, &cs);
    return mk(sinTheta * cs, sinTheta * sn, cosTheta);
}

// direct.h:123-124 / :427-428: cosine of the half-angle the sphere subtends from p
__device__ __forceinline__ float di_cos_theta_max(V3 p, V3 center, float radius) {
    float sinThetaMaxSquared = radius * radius / length2(center - p);
    return sqrtf(fmaxf(0.f, 1.f - sinThetaMaxSquared));
}

// raySphereIntersect + quadratic (direct.h:16-69), fp64 as the reference writes them
__device__ __forceinline__ bool di_ray_sphere(const RayT& r, V3 center, float radius) {
    V3 o = mk(r.ox, r.oy, r.oz) - center, d = mk(r.dx, r.dy, r.dz);
    double c = double(dot(o, o) - (radius * radius));
    double b = double(dot(o, d)) * 2.0;
    double a = double(dot(d, d));
    double disc = b * b - 4 * a * c;
    double t0, t1;
    if (disc > 0) {
        double sq = sqrt(disc), inv2a = 1 / (2 * a);
        t0 = (-b + sq) * inv2a; t1 = (-b - sq) * inv2a;
    } else if (disc == 0) {
        t0 = (-b + sqrt(disc)) / (2 * a); t1 = t0;
    } else return false;
    const double inf = __longlong_as_double(0x7ff0000000000000LL), mn = double(r.min_t), mx = double(r.max_t);
    t0 = t0 < 0 ? inf : t0; t1 = t1 < 0 ? inf : t1;
    t0 = (t0 > mn && t0 < mx) ? t0 : inf;
    t1 = (t1 > mn && t1 < mx) ? t1 : inf;
    double closest = fmin(t0, t1);
    return closest > mn && closest < mx;
}

__device__ __forceinline__ float4 di_sphere(const SceneDev& sc, int emitter) { return __ldg(sc.esphere + emitter); }
__device__ __forceinline__ V3 di_radiance(const SceneDev& sc, int emitter) {
    float4 e0 = __ldg(reinterpret_cast<const float4*>(sc.emitters + emitter));
    return mk(e0.x, e0.y, e0.z);
}
__device__ __forceinline__ RayT di_ray(V3 o, V3 d, float max_t) {
    RayT r; r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.min_t = 1e-8f; r.max_t = max_t;
    return r;
}
__device__ __forceinline__ float di_balance(float nf, float fPdf, float ng, float gPdf) { float f = nf * fPdf, g = ng * gPdf; return f / (f + g); }  // direct.h:81-84

// ---- every strategy's head (direct.h:147-154): primary hit; an emitter seen directly is its radiance --------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_di_init(const __grid_constant__ WaveParams p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.P) return;
    RayT cr = p.cam_ray[i];
    HitT h = p.hit[i];
    int n = 0;
    if (hit_valid(h, cr)) {
        Surf s = make_surf(p.sc, h, cr);
        V3 ke = pt_vis_emission(p, true, s.mat);
        if (!is_zero(ke)) pt_accum(p, i, ke);
        else {
            LightVertex v;
            v.px = s.p.x; v.py = s.p.y; v.pz = s.p.z; v.vcm = 0.f;
            v.nx = s.n.x; v.ny = s.n.y; v.nz = s.n.z; v.vc = 0.f;
            v.wox = s.wo.x; v.woy = s.wo.y; v.woz = s.wo.z; v.mat = s.mat;
            v.tx = 0.f; v.ty = 0.f; v.tz = 0.f; v.rr = 0.f;
            p.lv[i] = v;
            if (p.sc.has_tex) p.lv_tex[i] = tex_record(h);
            float4* st = reinterpret_cast<float4*>(p.ev + i);
            st[2] = make_float4(0.f, 0.f, 0.f, 0.f); st[3] = make_float4(0.f, 0.f, 0.f, 0.f);
            n = 1;
        }
    }
    p.nlv[i] = n;
}

// ---- one sample of the emitter loop (phase 0) or of the BSDF loop (phase 1): numbers, ray, stash -----------------------
// stash in ev[slot]: [0] = (wi local | f cos, sampled pdf), [1] = (emitterPdf, emitter id, areaToSolidAngle, -),
//                    [2] = running emitter-loop estimator, [3] = running BSDF-loop estimator
__global__ void __launch_bounds__(TRB_BLOCK) k_di_gen(const __grid_constant__ WaveParams p, DiParams dp, int phase,
                                                      int32_t* __restrict__ q_try, uint32_t* __restrict__ n_try) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    bool go = false;
    if (slot < p.P && p.nlv[slot] > 0) {
        PtVertex v = pt_load_vertex(p, 0, slot, nullptr, nullptr, nullptr, nullptr);
        uint32_t pixel, sample;
        slot_ids(p, slot, &pixel, &sample);
        float4 sv = p.st_v[slot];
        Rng rng = make_rng(p, slot, pixel, sample, __float_as_uint(sv.z));
        float4* st = reinterpret_cast<float4*>(p.ev + slot);
        const float inf = __int_as_float(0x7f800000);
        if (phase == 1) {  // direct.h:251-256, :409-414: the ray is traced whatever the sample's value
            Mat m = load_mat_stored(p.sc, v.mat, p.lv_tex + slot);
            float ux = rng.next(), uy = rng.next();
            V3 wi; float pdf;
            V3 fcos = bsdf_sample(m, v.wo, ux, uy, &wi, &pdf);
            st[0] = make_float4(fcos.x, fcos.y, fcos.z, pdf);
            p.ray[slot] = di_ray(v.p, to_world(v.f, wi), inf);
            go = true;
        } else if (dp.strategy == DI_COSINE_HEMISPHERE) {  // direct.h:215-218
            float ux = rng.next(), uy = rng.next();
            V3 local = sq_to_cosine_hemisphere(ux, uy);
            V3 world = normalize(to_world(v.f, local));
            st[0] = make_float4(local.x, local.y, local.z, 0.f);
            p.ray[slot] = di_ray(v.p, world, inf);
            go = true;
        } else {
            // selectEmitter (integrator.cpp:46-51), the emitter's sphere
            float s0 = rng.next();
            uint32_t ne = uint32_t(p.sc.n_emitters);
            uint32_t id = min(uint32_t(s0 * float(ne)), ne - 1u);
            float emitterPdf = 1.f / float(ne);
            float4 sp = di_sphere(p.sc, int(id));
            V3 center = mk(sp.x, sp.y, sp.z);
            float radius = sp.w;
            float ux = rng.next(), uy = rng.next();
            if (dp.strategy == DI_AREA) {  // sampleSphereByArea (direct.h:96-109) + :171-178
                V3 ne_ = sq_to_uniform_sphere(ux, uy);
                V3 pos = ne_ * radius + center;
                V3 wiW = normalize(pos - v.p);
                float outPdf = 1.f / (4 * TRB_PI * radius * radius);
                float d2 = length2(v.p - pos);
                float cosOut = dot(-wiW, ne_);
                V3 wiL = to_local(v.f, wiW);
                if (!(cosOut <= 0.f || wiL.z <= 0.f)) {
                    st[0] = make_float4(wiL.x, wiL.y, wiL.z, outPdf);
                    st[1] = make_float4(emitterPdf, __int_as_float(int(id)), cosOut * (1.f / d2), 0.f);
                    p.ray[slot] = di_ray(v.p, wiW, sqrtf(d2) - 1e-8f);
                    go = true;
                }
            } else {  // sampleSphereBySolidAngle (direct.h:111-141) + :297-304 / :370-377
                V3 centerDir = normalize(center - v.p);
                Frame fr = make_frame(centerDir);
                float cosThetaMax = di_cos_theta_max(v.p, center, radius);
                float cosTheta = (1.f - ux) + (ux * cosThetaMax);
                float phi = uy * TRB_PI * 2.0f;
                float sinTheta = sqrtf(fmaxf(1.f - (cosTheta * cosTheta), 0.f));
                float sn, cs; t_sincosf(phi, &sn, &cs);
                V3 wiW = to_world(fr, mk(sinTheta * cs, sinTheta * sn, cosTheta));
                float outPdf = TRB_INV_TWOPI * (1.f / (1.f - cosThetaMax));
                V3 wiL = to_local(v.f, wiW);
                if (!(wiL.z <= 0.f)) {
                    st[0] = make_float4(wiL.x, wiL.y, wiL.z, outPdf);
                    st[1] = make_float4(emitterPdf, __int_as_float(int(id)), 0.f, 0.f);
                    V3 dc = center - v.p;
                    p.ray[slot] = di_ray(v.p, wiW, dp.strategy == DI_MIS ? inf : sqrtf(dot(dc, dc)) + 1e-8f);
                    go = true;
                }
            }
        }
        sv.z = __uint_as_float(rng.ctr);
        p.st_v[slot] = sv;
    }
    uint32_t idx = queue_push(n_try, go);
    if (go) q_try[idx] = slot;
}

// ---- the sample's term (direct.h:182-190, :222-227, :257-261, :309-330, :380-402, :415-437) -------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_di_resolve(const __grid_constant__ WaveParams p, DiParams dp, int phase,
                                                          const int32_t* __restrict__ q_try, const uint32_t* __restrict__ n_try) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= *n_try) return;
    const int slot = q_try[k];
    RayT r = p.ray[slot];
    HitT h = p.hit[slot];
    const bool valid = hit_valid(h, r);
    int vmat = 0, vem = -1;
    if (valid) {
        const float4* s = reinterpret_cast<const float4*>(p.sc.shade + h.pos);
        vmat = __float_as_int(__ldg(s + 2).w); vem = __float_as_int(__ldg(s + 3).w);
    }
    float4* st = reinterpret_cast<float4*>(p.ev + slot);
    const float4 a = st[0], b = st[1];
    PtVertex v = pt_load_vertex(p, 0, slot, nullptr, nullptr, nullptr, nullptr);
    Mat m = load_mat_stored(p.sc, v.mat, p.lv_tex + slot);
    V3 c = mk(0.f);
    int which = 2;
    if (phase == 1) {
        which = 3;
        if (!valid) return;
        V3 ke = pt_vis_emission(p, true, vmat);
        V3 fcos = mk(a.x, a.y, a.z);
        if (dp.strategy == DI_BSDF) {
            c = ke * fcos * float(1.0 / double(a.w));  // `(1.0 / pdf)`: a double quotient, then glm's float scalar
        } else {
            if (is_zero(ke) || vem < 0) return;
            float4 sp = di_sphere(p.sc, vem);
            float cosThetaMax = di_cos_theta_max(v.p, mk(sp.x, sp.y, sp.z), sp.w);
            float emitterSolidAnglePdf = TRB_INV_TWOPI * (1.f / (1.f - cosThetaMax));
            emitterSolidAnglePdf *= 1.f / float(p.sc.n_emitters);
            float weight = di_balance(float(dp.bsdf_samples), a.w, float(dp.emitter_samples), emitterSolidAnglePdf);
            c = ke * fcos * weight * (1.f / a.w);
        }
    } else if (dp.strategy == DI_COSINE_HEMISPHERE) {
        if (!valid) return;
        V3 local = mk(a.x, a.y, a.z);
        c = pt_vis_emission(p, true, vmat) * bsdf_eval(m, v.wo, local) * (1.0f / cosine_hemisphere_pdf(local));
    } else if (dp.strategy == DI_AREA) {
        if (valid) return;  // `if (!intersects2)`: nothing between the point and the sampled position
        const int id = __float_as_int(b.y);
        c = di_radiance(p.sc, id) * bsdf_eval(m, v.wo, mk(a.x, a.y, a.z)) * (1.f / a.w) * (1.f / b.x) * b.z;
    } else {
        const int id = __float_as_int(b.y);
        bool contributes;
        if (valid) contributes = vem == id;  // emitter.shapeID == visibilityInteraction.shapeID
        else { float4 sp = di_sphere(p.sc, id); contributes = di_ray_sphere(r, mk(sp.x, sp.y, sp.z), sp.w); }
        if (!contributes) return;
        V3 wiL = mk(a.x, a.y, a.z);
        if (dp.strategy == DI_MIS) {
            const Lobe lobe = glossy_lobe(m, v.wo, wiL);
            float bsdfPdf = bsdf_pdf(m, v.wo, wiL, lobe);
            float weight = di_balance(float(dp.emitter_samples), a.w * b.x, float(dp.bsdf_samples), bsdfPdf);
            c = di_radiance(p.sc, id) * bsdf_eval(m, v.wo, wiL, lobe) * weight * (1.f / a.w) * (1.f / b.x);
        } else {
            c = di_radiance(p.sc, id) * bsdf_eval(m, v.wo, wiL) * (1.f / a.w) * (1.f / b.x);
        }
    }
    float4 acc = st[which];
    st[which] = make_float4(acc.x + c.x, acc.y + c.y, acc.z + c.z, 0.f);
}

// ---- the sample's value (direct.h:192,229,263,332; MIS :404,439,442) ---------------------------------------------------
__global__ void __launch_bounds__(TRB_BLOCK) k_di_combine(const __grid_constant__ WaveParams p, DiParams dp) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= p.P || p.nlv[slot] <= 0) return;
    const float4* st = reinterpret_cast<const float4*>(p.ev + slot);
    float4 ee = st[2], be = st[3];
    V3 Lr;
    if (dp.strategy == DI_MIS) {
        V3 e = dp.emitter_samples == 0 ? mk(0.f) : mk(ee.x, ee.y, ee.z) / float(dp.emitter_samples);
        V3 b = dp.bsdf_samples == 0 ? mk(0.f) : mk(be.x, be.y, be.z) / float(dp.bsdf_samples);
        Lr = mk(0.f) + (e + b);
    } else if (dp.strategy == DI_BSDF) {
        Lr = mk(be.x, be.y, be.z) / float(dp.bsdf_samples);
    } else {
        Lr = mk(ee.x, ee.y, ee.z) / float(dp.emitter_samples);
    }
    pt_accum(p, slot, Lr);
}

}  // namespace trbk
