/* oracle/bdpt_oracle.cpp -- TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or executed from the
 * product (bidirectional-path-tracing_b200/); only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
 * leg may use it, and only as the checker / the reported CPU baseline.
 *
 * Plain scalar C++ restatement of the reference's BDPT hot path (JackMinn/Bidirectional-Path-Tracing), one
 * function per reference function, same control flow (per-sample while loops, exhaustive stack traversal),
 * each citing the file:line it follows.  PARITY PINNED: tests/test_oracle.py checks this file against golden
 * vectors generated from the UNMODIFIED reference compiled in oracle/_ref (tools/gen_golden.py): BVH primitive
 * order, emitters, closest hits and occlusion (bit-exact), BSDF eval/pdf/sample, the Sampler stream
 * (bit-exact) and whole BDPT samples (Li + splats) under identical seeds.
 *
 * Build: make -C oracle port   (g++ -O2 -ffp-contract=off; x86-64 baseline has no FMA)
 */
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <random>
#include <thread>
#include <vector>

namespace {

/* ---- src/core/platform.h:48-66 ------------------------------------------------------------------------ */
const float kPi = 3.14159265358979323846f, kInvPi = 0.31830988618379067154f, kInvTwoPi = 0.15915494309189533577f;
const float kEps = 1e-8f;

struct v3 { float x, y, z; };
inline v3 V(float a, float b, float c) { return v3{a, b, c}; }
inline v3 V(float a) { return v3{a, a, a}; }
inline v3 operator+(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
inline v3 operator-(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
inline v3 operator-(v3 a) { return V(-a.x, -a.y, -a.z); }
inline v3 operator*(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
inline v3 operator*(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
inline v3 operator*(float s, v3 a) { return V(s * a.x, s * a.y, s * a.z); }
inline v3 operator/(v3 a, float s) { return V(a.x / s, a.y / s, a.z / s); }
inline bool operator==(v3 a, v3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(v3 a, v3 b) { return !(a == b); }
inline float at(const v3& a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }
/* glm 0.9.9: dot(vec3) = (x+y)+z; cross; normalize = v * inversesqrt(dot); length = sqrt(dot) */
inline float dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline v3 cross(v3 x, v3 y) { return V(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
inline float length2(v3 a) { return dot(a, a); }
inline float length(v3 a) { return std::sqrt(dot(a, a)); }
inline v3 normalize(v3 a) { return a * (1.f / std::sqrt(dot(a, a))); }
inline v3 vmin(v3 a, v3 b) { return V(b.x < a.x ? b.x : a.x, b.y < a.y ? b.y : a.y, b.z < a.z ? b.z : a.z); }
inline v3 vmax(v3 a, v3 b) { return V(a.x < b.x ? b.x : a.x, a.y < b.y ? b.y : a.y, a.z < b.z ? b.z : a.z); }

/* src/core/math.h:19-22 */
inline v3 barycentric(v3 a, v3 b, v3 c, float u, float v) { return (a * (1 - u - v) + b * u) + c * v; }
/* src/core/math.h:56-58 */
inline float getLuminance(v3 rgb) { return dot(rgb, V(0.212671f, 0.715160f, 0.072169f)); }

/* src/core/math.h:42-51 + src/core/core.h:152-167 */
struct Frame {
    v3 s, t, n;
    Frame() : s(V(0)), t(V(0)), n(V(0)) {}
    explicit Frame(v3 a) : n(a) {
        if (std::abs(a.x) > std::abs(a.y)) {
            float invLen = 1.f / std::sqrt(a.x * a.x + a.z * a.z);
            t = V(a.z * invLen, 0.f, -a.x * invLen);
        } else {
            float invLen = 1.f / std::sqrt(a.y * a.y + a.z * a.z);
            t = V(0.f, a.z * invLen, -a.y * invLen);
        }
        s = cross(t, a);
    }
    v3 toLocal(v3 v) const { return V(dot(v, s), dot(v, t), dot(v, n)); }
    v3 toWorld(v3 v) const { return (s * v.x + t * v.y) + n * v.z; }
};

/* src/core/core.h:117-122 */
struct Ray {
    v3 o, d;
    float min_t, max_t;
    Ray(v3 co, v3 cd, float mn = kEps, float mx = std::numeric_limits<float>::max()) : o(co), d(cd), min_t(mn), max_t(mx) {}
};

/* src/core/core.h:173-180 */
struct SurfaceInteraction {
    v3 p = V(0), wo = V(0), wi = V(0);
    float t = 0, u = 0, v = 0;
    size_t shapeID = 0, primID = 0;
    Frame frameNs;
    int matID = 0;
};

/* src/core/math.h:63-76: std::mt19937 + std::uniform_real_distribution<float>(0,1) as libstdc++ implements it:
 * generate_canonical<float,24> takes ONE 32-bit draw: float(u32) / 2^32, clamped below 1 */
struct Sampler {
    std::mt19937 g;
    const float* injected = nullptr;  /* parity mode: numbers supplied by the caller */
    size_t pos = 0;
    explicit Sampler(int seed) : g(seed) {}
    float next() {
        if (injected) return injected[pos++];
        ++pos;
        float sum = float(g() - std::mt19937::min());
        float ret = sum / 4294967296.0f;
        if (ret >= 1.0f) ret = std::nextafter(1.0f, 0.0f);
        return ret * (1.f - 0.f) + 0.f;
    }
    void next2D(float* a, float* b) { *a = next(); *b = next(); }
};

/* ---- Warp (src/core/math.h:118-268) ---------------------------------------------------------------------- */
inline v3 squareToUniformHemisphere(float sx, float sy) {
    float phi = sx * kPi * 2.0f;
    float cosTheta = sy;
    float sinTheta = sqrtf(float(std::fmax(1.f - (cosTheta * cosTheta), 0)));
    return V(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}
inline void squareToUniformDiskConcentric(float sx, float sy, float* ox, float* oy) {
    float phi, radius;
    float rx = (2.f * sx) - 1.f, ry = (2.f * sy) - 1.f;
    if (rx == 0 && ry == 0) { *ox = 0.f; *oy = 0.f; return; }
    if ((rx * rx) > (ry * ry)) { radius = rx; phi = (kPi * 0.25f) * (ry * (1.f / rx)); }
    else { radius = ry; phi = (kPi * 0.5f) - ((kPi * 0.25f) * (rx * (1.f / ry))); }
    *ox = radius * cosf(phi); *oy = radius * sinf(phi);
}
inline v3 squareToCosineHemisphere(float sx, float sy) {
    float dx, dy;
    squareToUniformDiskConcentric(sx, sy, &dx, &dy);
    float z = 1.0f - (dx * dx + dy * dy);
    z = std::fmax(z, 0.f);
    z = sqrtf(z);
    return V(dx, dy, z);
}
inline float squareToCosineHemispherePdf(v3 v) { return v.z >= 0.f ? v.z * kInvPi : 0.f; }
inline v3 squareToPhongLobe(float sx, float sy, float exponent) {
    float cosTheta = powf(sx, 1.f / (exponent + 2));
    float sinTheta = sqrtf(float(std::fmax(1.f - (cosTheta * cosTheta), 0)));
    float phi = sy * 2.f * kPi;
    return V(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}
inline float squareToPhongLobePdf(v3 v, float exponent) { return v.z >= 0.f ? (exponent + 2) * kInvTwoPi * powf(v.z, exponent) : 0.f; }

/* ---- scene --------------------------------------------------------------------------------------------- */
struct Material { v3 kd, ks, ke, tf; float ns, ni; int illum; };

enum { EDiffuseReflection = 0x2, EGlossyReflection = 0x8, EDeltaReflection = 0x20, EDeltaTransmission = 0x40, ENull = 0x1 };
const unsigned EDelta = ENull | EDeltaReflection | EDeltaTransmission;

/* BSDF classes: diffuse.h, phong.h, mixture.h, perfectmirror.h, glass.h; selection renderer.cpp:259-271 */
struct BSDF {
    int kind = 0;  /* 0 diffuse(7) 1 phong(else) 2 mixture(8) 3 mirror(3) 4 glass(6) */
    v3 kd = V(0), ks = V(0), emission = V(0), tf = V(0);
    float exponent = 0, ior = 1, scale = 1, specularSamplingWeight = 0;
    unsigned combinedType = 0;

    static v3 reflect(v3 d) { return V(-d.x, -d.y, d.z); }

    v3 evalGlossy(const SurfaceInteraction& i) const { /* phong.h:61-74, mixture.h:60-73 */
        v3 val = V(0.f);
        if (i.wi.z >= 0.f && i.wo.z >= 0.f) {
            val = val + kd * kInvPi;
            float cosTheta = std::fmin(std::fmax(dot(i.wi, reflect(i.wo)), 0.f), 1.f);
            val = val + ks * (exponent + 2) * kInvTwoPi * powf(cosTheta, exponent);
            val = val * scale;
            val = val * i.wi.z;
        }
        return val;
    }
    v3 eval(const SurfaceInteraction& i) const {
        switch (kind) {
            case 0: /* diffuse.h:35-43 */
                if (i.wi.z >= 0.f && i.wo.z >= 0.f) return kd * kInvPi * i.wi.z;
                return V(0.f);
            case 1: case 2: return evalGlossy(i);
            default: return V(0.f); /* perfectmirror.h:33-38, glass.h:55-58 */
        }
    }
    float pdfPhong(const SurfaceInteraction& i) const { /* phong.h:76-87 */
        Frame reflectionSpace(reflect(i.wo));
        return squareToPhongLobePdf(reflectionSpace.toLocal(i.wi), exponent);
    }
    float pdf(const SurfaceInteraction& i) const {
        switch (kind) {
            case 0: return squareToCosineHemispherePdf(i.wi); /* diffuse.h:45-50 */
            case 1: return pdfPhong(i);
            case 2: { /* mixture.h:75-93 */
                float pdfP = pdfPhong(i), pdfD = squareToCosineHemispherePdf(i.wi);
                return (pdfP * specularSamplingWeight) + (pdfD * (1.f - specularSamplingWeight));
            }
            default: return 0.f;
        }
    }
    static float FresnelDielectric(float eta_i, float eta_t, float cos_i, float cos_t) { /* glass.h:40-53 */
        float eta = eta_i / eta_t;
        float sin2_t = eta * eta * (std::fmax(0.f, 1.f - cos_i * cos_i));
        if (sin2_t >= 1.f) return 1.f;
        float rParallel = ((eta_t * cos_i) - (eta_i * cos_t)) / ((eta_t * cos_i) + (eta_i * cos_t));
        float rPerpendicular = ((eta_i * cos_i) - (eta_t * cos_t)) / ((eta_i * cos_i) + (eta_t * cos_t));
        return (rParallel * rParallel + rPerpendicular * rPerpendicular) * 0.5f;
    }
    v3 sample(SurfaceInteraction& i, float sx, float sy, float* pdfOut) const {
        switch (kind) {
            case 0: { /* diffuse.h:52-61 */
                i.wi = squareToCosineHemisphere(sx, sy);
                *pdfOut = pdf(i);
                return eval(i);
            }
            case 1: { /* phong.h:89-105 */
                Frame reflectionSpace(reflect(i.wo));
                v3 s = squareToPhongLobe(sx, sy, exponent);
                *pdfOut = squareToPhongLobePdf(s, exponent);
                i.wi = reflectionSpace.toWorld(s);
                return eval(i);
            }
            case 2: { /* mixture.h:95-151 */
                v3 val;
                if (sx < specularSamplingWeight) {
                    Frame reflectionSpace(reflect(i.wo));
                    v3 s = squareToPhongLobe(sx / specularSamplingWeight, sy, exponent);
                    i.wi = reflectionSpace.toWorld(s);
                    val = eval(i);
                } else {
                    i.wi = squareToCosineHemisphere((sx - specularSamplingWeight) / (1.f - specularSamplingWeight), sy);
                    val = eval(i);
                }
                *pdfOut = pdf(i);
                return val;
            }
            case 3: { /* perfectmirror.h:48-59 */
                *pdfOut = 1.f;
                i.wi = reflect(i.wo);
                return V(1.f);
            }
            default: { /* glass.h:60-108 */
                *pdfOut = 1;
                bool isEntering = i.wo.z > 0.f;
                float eta_i = 1.f, eta_t = ior;
                if (!isEntering) std::swap(eta_i, eta_t);
                float eta = eta_i / eta_t;
                float sin2_i = std::fmax(0.f, 1.f - i.wo.z * i.wo.z);
                float sin2_t = eta * eta * sin2_i;
                float cos_t = sqrtf(std::fmax(0.f, 1.f - sin2_t));
                cos_t = isEntering ? -cos_t : cos_t;
                float fresnel = FresnelDielectric(eta_i, eta_t, std::fabs(i.wo.z), std::fabs(cos_t));
                if (sx < fresnel) { i.wi = reflect(i.wo); return V(1.f); }
                i.wi = V(eta * -i.wo.x, eta * -i.wo.y, cos_t);
                return tf;
            }
        }
    }
};

BSDF makeBSDF(const Material& m) {
    BSDF b;
    b.kd = m.kd; b.ks = m.ks; b.emission = m.ke; b.tf = m.tf; b.exponent = m.ns; b.ior = m.ni;
    if (m.illum == 7) { b.kind = 0; b.combinedType = EDiffuseReflection; }
    else if (m.illum == 3) { b.kind = 3; b.combinedType = EDeltaReflection; }
    else if (m.illum == 6) { b.kind = 4; b.combinedType = EDeltaReflection | EDeltaTransmission; }
    else {
        b.kind = (m.illum == 8) ? 2 : 1;
        b.combinedType = EGlossyReflection | EDiffuseReflection;
        v3 maxValue = m.ks + m.kd; /* phong.h:41-47 */
        float actualMax = std::max(std::max(maxValue.x, maxValue.y), maxValue.z);
        b.scale = actualMax > 1.0f ? 0.99f * (1.0f / actualMax) : 1.0f;
        float dAvg = getLuminance(m.kd * b.scale), sAvg = getLuminance(m.ks * b.scale);
        b.specularSamplingWeight = sAvg / (dAvg + sAvg);
    }
    return b;
}

}  // namespace
