// bsdf.cuh -- BSDF::{eval,pdf,sample} for the five material classes of the reference, in the local
// shading frame, returning f*cos(theta_i) exactly as the reference does.
//   DiffuseBSDF  src/bsdfs/diffuse.h:35-61      PhongBSDF   src/bsdfs/phong.h:61-105
//   MixtureBSDF  src/bsdfs/mixture.h:60-151     MirrorBSDF  src/bsdfs/perfectmirror.h:33-59
//   GlassBSDF    src/bsdfs/glass.h:40-108       Warp::*     src/core/math.h:118-268
// Operation order follows the C++ expressions so that, with -fmad=false, only the transcendental
// functions can differ from the x86-64 build.  Where a transcendental decides a DIRECTION (sampling: sin/cos of the
// azimuth, the pow of the Phong-lobe polar angle) it goes through fp64 (TRB_EXACT_TRANSCENDENTALS) so that the fp32
// result is the correctly rounded one glibc returns in all but ~1e-8 of the cases - a path through the glass sphere
// amplifies a 1-ulp direction error by orders of magnitude.  Where it only scales a VALUE (the Phong lobe in
// eval / pdf: six pow per vertex connection) CUDA's fp32 powf (<= 4 ulp, 5e-7 relative, against the 1e-5 parity
// tolerance) is used: the fp64 pow was 52 % of k_eye_bounce's instructions (profiles/r1e_ncu_lines_k_eye_bounce.txt).
#pragma once
#include "../host/tr_host.h"
#include "vec.cuh"

#ifndef TRB_EXACT_TRANSCENDENTALS
#define TRB_EXACT_TRANSCENDENTALS 1
#endif

namespace trbk {

#if TRB_EXACT_TRANSCENDENTALS
__device__ __forceinline__ float t_sinf(float x) { return (float)sin((double)x); }
__device__ __forceinline__ float t_cosf(float x) { return (float)cos((double)x); }
__device__ __forceinline__ float t_powf(float x, float y) { return (float)pow((double)x, (double)y); }
__device__ __forceinline__ void t_sincosf(float x, float* s, float* c) { double ds, dc; sincos((double)x, &ds, &dc); *s = (float)ds; *c = (float)dc; }
__device__ __forceinline__ float v_powf(float x, float y) { return powf(x, y); }  // value-only uses
#else
__device__ __forceinline__ float v_powf(float x, float y) { return powf(x, y); }
__device__ __forceinline__ float t_sinf(float x) { return sinf(x); }
__device__ __forceinline__ float t_cosf(float x) { return cosf(x); }
__device__ __forceinline__ float t_powf(float x, float y) { return powf(x, y); }
__device__ __forceinline__ void t_sincosf(float x, float* s, float* c) { sincosf(x, s, c); }
#endif

struct Mat {
    V3 kd, ks, ke, tf;
    float ns, ni, scale, sw;
    int cls;
    uint32_t type;
    bool emissive;
    int tex;   // DevMaterial::tex
};

__device__ __forceinline__ Mat load_mat(const trb::DevMaterial* mats, int id) {
    const float4* p = reinterpret_cast<const float4*>(mats + id);
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4);
    Mat m;
    m.kd = mk(a.x, a.y, a.z); m.cls = __float_as_int(a.w);
    m.ks = mk(b.x, b.y, b.z); m.type = __float_as_uint(b.w);
    m.ke = mk(c.x, c.y, c.z); m.ns = c.w;
    m.tf = mk(d.x, d.y, d.z); m.ni = d.w;
    m.scale = e.x; m.sw = e.y; m.emissive = __float_as_int(e.z) != 0; m.tex = __float_as_int(e.w);
    return m;
}

__device__ __forceinline__ bool is_delta(const Mat& m) { return (m.type & (0x1u | 0x20u | 0x40u)) != 0; }  // BSDF::EDelta

__device__ __forceinline__ V3 reflect_z(V3 d) { return mk(-d.x, -d.y, d.z); }

// ---- Warp (math.h) ---------------------------------------------------------------------------------
__device__ __forceinline__ V3 sq_to_uniform_hemisphere(float ux, float uy) {  // math.h:136-144
    float phi = ux * TRB_PI * 2.0f;
    float cosTheta = uy;
    float sinTheta = sqrtf(fmaxf(1.f - (cosTheta * cosTheta), 0.f));
    float sn, cs; t_sincosf(phi, &sn, &cs);  // one shared range reduction
    return mk(sinTheta * cs, sinTheta * sn, cosTheta);
}

__device__ __forceinline__ V3 sq_to_cosine_hemisphere(float ux, float uy) {  // math.h:153-192
    float rx = (2.f * ux) - 1.f, ry = (2.f * uy) - 1.f;
    float dx = 0.f, dy = 0.f;
    if (!(rx == 0 && ry == 0)) {
        float phi, radius;
        if ((rx * rx) > (ry * ry)) { radius = rx; phi = (TRB_PI * 0.25f) * (ry * (1.f / rx)); }
        else { radius = ry; phi = (TRB_PI * 0.5f) - ((TRB_PI * 0.25f) * (rx * (1.f / ry))); }
        float sn, cs; t_sincosf(phi, &sn, &cs);
        dx = radius * cs; dy = radius * sn;
    }
    float z = 1.0f - (dx * dx + dy * dy);
    z = fmaxf(z, 0.f);
    z = sqrtf(z);
    return mk(dx, dy, z);
}

__device__ __forceinline__ float cosine_hemisphere_pdf(V3 v) { return v.z >= 0.f ? v.z * TRB_INV_PI : 0.f; }  // math.h:194-208

__device__ __forceinline__ V3 sq_to_phong_lobe(float ux, float uy, float exponent) {  // math.h:210-219
    float cosTheta = t_powf(ux, 1.f / (exponent + 2));
    float sinTheta = sqrtf(fmaxf(1.f - (cosTheta * cosTheta), 0.f));
    float phi = uy * 2.f * TRB_PI;
    float sn, cs; t_sincosf(phi, &sn, &cs);
    return mk(sinTheta * cs, sinTheta * sn, cosTheta);
}

__device__ __forceinline__ float phong_lobe_pdf(V3 v, float exponent) {  // math.h:221-227
    return v.z >= 0.f ? (exponent + 2) * TRB_INV_TWOPI * v_powf(v.z, exponent) : 0.f;
}

// ---- eval / pdf / sample ---------------------------------------------------------------------------
// The Phong lobe of a glossy material sees (wo, wi) only through c = dot(wi, reflect_z(wo)).  Reflecting wi instead gives the same float:
// the products commute and are summed in the same order, and nothing is contracted (-fmad=false).  So the eval of phong.h:61-74, the pdf of
// phong.h:76-87 for (wo, wi) and that pdf for (wi, wo) - what every connection of bdpt.h asks of a vertex - share ONE powf(c, ns); its
// value is bit for bit what each of them computed on its own (ncu, round 2: the three inlined powf were 40 % of k_connect_vertices'
// warp instructions at 6 active lanes).
struct Lobe { float c, pw; };   // c = dot(wi, reflect_z(wo)); pw = powf(c, ns) when c >= 0
__device__ __forceinline__ Lobe glossy_lobe(const Mat& m, V3 wo, V3 wi) {
    Lobe l; l.c = 0.f; l.pw = 0.f;
    if (m.cls == trb::MAT_PHONG || m.cls == trb::MAT_MIXTURE) {
        l.c = dot(wi, reflect_z(wo));
        l.pw = l.c >= 0.f ? v_powf(l.c, m.ns) : 0.f;
    }
    return l;
}
// powf(min(max(c, 0), 1), e) from the lobe's powf(c, e): the clamp only matters outside [0, 1] (and for a NaN, which fmaxf drops)
__device__ __forceinline__ float lobe_pow_clamped(const Lobe& l, float e) {
    if (l.c >= 0.f && l.c <= 1.f) return l.pw;
    if (l.c > 1.f) return 1.f;                                      // powf(1, e)
    return e > 0.f ? 0.f : (e == 0.f ? 1.f : v_powf(0.f, e));      // powf(0, e)
}

__device__ __forceinline__ V3 glossy_eval(const Mat& m, V3 wo, V3 wi, const Lobe& l) {  // phong.h:61-74 == mixture.h:60-73
    V3 val = mk(0.f);
    if (wi.z >= 0.f && wo.z >= 0.f) {
        val = val + m.kd * TRB_INV_PI;
        float e = m.ns;
        val = val + m.ks * (e + 2) * TRB_INV_TWOPI * lobe_pow_clamped(l, e);
        val = val * m.scale;
        val = val * wi.z;
    }
    return val;
}

__device__ __forceinline__ V3 bsdf_eval(const Mat& m, V3 wo, V3 wi, const Lobe& l) {
    switch (m.cls) {
        case trb::MAT_DIFFUSE:
            if (wi.z >= 0.f && wo.z >= 0.f) return m.kd * TRB_INV_PI * wi.z;  // diffuse.h:35-43
            return mk(0.f);
        case trb::MAT_PHONG:
        case trb::MAT_MIXTURE: return glossy_eval(m, wo, wi, l);
        default: return mk(0.f);  // delta lobes: perfectmirror.h:33-38, glass.h:55-58
    }
}

// phong.h:76-87: Frame(reflect(wo)).toLocal(wi).z is the lobe's c
__device__ __forceinline__ float phong_pdf(const Mat& m, const Lobe& l) { return l.c >= 0.f ? (m.ns + 2) * TRB_INV_TWOPI * l.pw : 0.f; }

// `l`: the lobe of (wo, wi) or of (wi, wo) - the same
__device__ __forceinline__ float bsdf_pdf(const Mat& m, V3 wo, V3 wi, const Lobe& l) {
    switch (m.cls) {
        case trb::MAT_DIFFUSE: return cosine_hemisphere_pdf(wi);  // diffuse.h:45-50
        case trb::MAT_PHONG: return phong_pdf(m, l);
        case trb::MAT_MIXTURE: {  // mixture.h:75-93
            float pp = phong_pdf(m, l);
            float pd = cosine_hemisphere_pdf(wi);
            return (pp * m.sw) + (pd * (1.f - m.sw));
        }
        default: return 0.f;
    }
}
// one-off queries
__device__ __forceinline__ V3 bsdf_eval(const Mat& m, V3 wo, V3 wi) { return bsdf_eval(m, wo, wi, glossy_lobe(m, wo, wi)); }
__device__ __forceinline__ float bsdf_pdf(const Mat& m, V3 wo, V3 wi) { return bsdf_pdf(m, wo, wi, glossy_lobe(m, wo, wi)); }

__device__ __forceinline__ float fresnel_dielectric(float eta_i, float eta_t, float cos_i, float cos_t) {  // glass.h:40-53
    float eta = eta_i / eta_t;
    float sin2_t = eta * eta * (fmaxf(0.f, 1.f - cos_i * cos_i));
    if (sin2_t >= 1.f) return 1.f;
    float rPar = ((eta_t * cos_i) - (eta_i * cos_t)) / ((eta_t * cos_i) + (eta_i * cos_t));
    float rPer = ((eta_i * cos_i) - (eta_t * cos_t)) / ((eta_i * cos_i) + (eta_t * cos_t));
    return (rPar * rPar + rPer * rPer) * 0.5f;
}

// BSDF::sample: writes wi and pdf, returns f*cos; `lobe_out`: the lobe of (wo, wi), for the caller's reverse pdf
__device__ __forceinline__ V3 bsdf_sample(const Mat& m, V3 wo, float ux, float uy, V3* wi, float* pdf, Lobe* lobe_out = nullptr) {
    Lobe lobe; lobe.c = 0.f; lobe.pw = 0.f;
    V3 ret;
    switch (m.cls) {
        case trb::MAT_DIFFUSE: {  // diffuse.h:52-61
            *wi = sq_to_cosine_hemisphere(ux, uy);
            *pdf = cosine_hemisphere_pdf(*wi);
            ret = bsdf_eval(m, wo, *wi, lobe);
            break;
        }
        case trb::MAT_PHONG: {  // phong.h:89-105
            Frame rf = make_frame(reflect_z(wo));
            V3 s = sq_to_phong_lobe(ux, uy, m.ns);
            *pdf = phong_lobe_pdf(s, m.ns);
            *wi = to_world(rf, s);
            lobe = glossy_lobe(m, wo, *wi);
            ret = glossy_eval(m, wo, *wi, lobe);
            break;
        }
        case trb::MAT_MIXTURE: {  // mixture.h:95-151
            if (ux < m.sw) {
                Frame rf = make_frame(reflect_z(wo));
                V3 s = sq_to_phong_lobe(ux / m.sw, uy, m.ns);
                *wi = to_world(rf, s);
            } else {
                *wi = sq_to_cosine_hemisphere((ux - m.sw) / (1.f - m.sw), uy);
            }
            lobe = glossy_lobe(m, wo, *wi);
            ret = glossy_eval(m, wo, *wi, lobe);
            *pdf = bsdf_pdf(m, wo, *wi, lobe);
            break;
        }
        case trb::MAT_MIRROR: {  // perfectmirror.h:48-59
            *pdf = 1.f;
            *wi = reflect_z(wo);
            ret = mk(1.f);
            break;
        }
        default: {  // glass.h:60-108
            *pdf = 1.f;
            bool entering = wo.z > 0.f;
            float eta_i = 1.f, eta_t = m.ni;
            if (!entering) { float s = eta_i; eta_i = eta_t; eta_t = s; }
            float eta = eta_i / eta_t;
            float sin2_i = fmaxf(0.f, 1.f - wo.z * wo.z);
            float sin2_t = eta * eta * sin2_i;
            float cos_t = sqrtf(fmaxf(0.f, 1.f - sin2_t));
            cos_t = entering ? -cos_t : cos_t;
            float fr = fresnel_dielectric(eta_i, eta_t, fabsf(wo.z), fabsf(cos_t));
            if (ux < fr) { *wi = reflect_z(wo); ret = mk(1.f); }
            else { *wi = mk(eta * -wo.x, eta * -wo.y, cos_t); ret = m.tf; }
            break;
        }
    }
    if (lobe_out) *lobe_out = lobe;
    return ret;
}

}  // namespace trbk
