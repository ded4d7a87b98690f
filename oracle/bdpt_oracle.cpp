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
    v3 ng = V(0);  /* frameNg.n (accel.h:161); only path.h's bsdf-sampling estimator reads it */
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
struct Material { v3 kd, ks, ke, tf; float ns, ni; int illum; int tex = 0; /* (1 + map_Kd texture) | (1 + map_Ks texture) << 16 */ };

/* Tex + BitmapTexture3f (core.h:417-497, 525-585): decoded images (gamma-expanded, vertically flipped), attrib.texcoords and the
 * per-corner texcoord indices, shape-major like the other triangle arrays */
struct TexSet {
    struct Tex { int w = 0, h = 0; std::vector<float> cs; };
    std::vector<Tex> textures;
    std::vector<float> texcoords;   /* 2 per vt */
    std::vector<int> tri_t;         /* 3 per triangle */
    const std::vector<int>* shapeFirst = nullptr;
    v3 getAverage(int t) const { /* :546-552 */
        const Tex& x = textures[t];
        v3 s = V(0);
        for (size_t i = 0; i < x.cs.size() / 3; i++) s = s + V(x.cs[i * 3 + 0], x.cs[i * 3 + 1], x.cs[i * 3 + 2]);
        float scale = float(1.0 / (x.cs.size() / 3.0));
        return s * scale;
    }
    v3 getMax(int t) const { /* :561-567, from -FLT_MIN as written */
        const Tex& x = textures[t];
        const float m = -std::numeric_limits<float>::min();
        v3 s = V(m, m, m);
        for (size_t i = 0; i < x.cs.size() / 3; i++) s = vmax(s, V(x.cs[i * 3 + 0], x.cs[i * 3 + 1], x.cs[i * 3 + 2]));
        return s;
    }
    v3 eval(int t, const SurfaceInteraction& hit) const { /* :569-585 */
        const Tex& x = textures[t];
        const int tri = (*shapeFirst)[hit.shapeID] + int(hit.primID);
        const int i0 = tri_t[3 * tri], i1 = tri_t[3 * tri + 1], i2 = tri_t[3 * tri + 2];
        const float ax = texcoords[2 * i0], ay = texcoords[2 * i0 + 1], bx = texcoords[2 * i1], by = texcoords[2 * i1 + 1];
        const float cx = texcoords[2 * i2], cy = texcoords[2 * i2 + 1];
        const float k = 1 - hit.u - hit.v;  /* barycentric(): a * (1 - u - v) + b * u + c * v */
        float sx = ((ax * k + bx * hit.u) + cx * hit.v) + 1.0f, sy = ((ay * k + by * hit.u) + cy * hit.v) + 1.0f;
        sx = sx - std::floor(sx); sy = sy - std::floor(sy);
        const int px = std::min(std::max(int(sx * x.w), 0), x.w - 1), py = std::min(std::max(int(sy * x.h), 0), x.h - 1);
        const int i = x.w * py + px;
        return V(x.cs[i * 3 + 0], x.cs[i * 3 + 1], x.cs[i * 3 + 2]);
    }
};

enum { EDiffuseReflection = 0x2, EGlossyReflection = 0x8, EDeltaReflection = 0x20, EDeltaTransmission = 0x40, ENull = 0x1 };
const unsigned EDelta = ENull | EDeltaReflection | EDeltaTransmission;

/* BSDF classes: diffuse.h, phong.h, mixture.h, perfectmirror.h, glass.h; selection renderer.cpp:259-271 */
struct BSDF {
    int kind = 0;  /* 0 diffuse(7) 1 phong(else) 2 mixture(8) 3 mirror(3) 4 glass(6) */
    v3 kd = V(0), ks = V(0), emission = V(0), tf = V(0);
    float exponent = 0, ior = 1, scale = 1, specularSamplingWeight = 0;
    unsigned combinedType = 0;
    int kdTex = -1, ksTex = -1;     /* diffuse_texname / specular_texname */
    const TexSet* ts = nullptr;
    v3 kdAt(const SurfaceInteraction& i) const { return kdTex >= 0 ? ts->eval(kdTex, i) : kd; }  /* Texture::eval */
    v3 ksAt(const SurfaceInteraction& i) const { return ksTex >= 0 ? ts->eval(ksTex, i) : ks; }

    static v3 reflect(v3 d) { return V(-d.x, -d.y, d.z); }

    v3 evalGlossy(const SurfaceInteraction& i) const { /* phong.h:61-74, mixture.h:60-73 */
        v3 val = V(0.f);
        if (i.wi.z >= 0.f && i.wo.z >= 0.f) {
            val = val + kdAt(i) * kInvPi;
            float cosTheta = std::fmin(std::fmax(dot(i.wi, reflect(i.wo)), 0.f), 1.f);
            val = val + ksAt(i) * (exponent + 2) * kInvTwoPi * powf(cosTheta, exponent);
            val = val * scale;
            val = val * i.wi.z;
        }
        return val;
    }
    v3 eval(const SurfaceInteraction& i) const {
        switch (kind) {
            case 0: /* diffuse.h:35-43 */
                if (i.wi.z >= 0.f && i.wo.z >= 0.f) return kdAt(i) * kInvPi * i.wi.z;
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

BSDF makeBSDF(const Material& m, const TexSet* ts = nullptr) {
    BSDF b;
    b.kd = m.kd; b.ks = m.ks; b.emission = m.ke; b.tf = m.tf; b.exponent = m.ns; b.ior = m.ni;
    b.ts = ts; b.kdTex = (m.tex & 0xffff) - 1; b.ksTex = ((m.tex >> 16) & 0xffff) - 1;
    if (m.illum == 7) b.ksTex = -1;                      /* DiffuseBSDF reads map_Kd only (diffuse.h:23-26) */
    if (m.illum == 3 || m.illum == 6) b.kdTex = b.ksTex = -1;
    if (m.illum == 7) { b.kind = 0; b.combinedType = EDiffuseReflection; }
    else if (m.illum == 3) { b.kind = 3; b.combinedType = EDeltaReflection; }
    else if (m.illum == 6) { b.kind = 4; b.combinedType = EDeltaReflection | EDeltaTransmission; }
    else {
        b.kind = (m.illum == 8) ? 2 : 1;
        b.combinedType = EGlossyReflection | EDiffuseReflection;
        /* phong.h:41-47: Texture::getMax / getAverage - the constant itself, or the bitmap's */
        v3 maxValue = (b.ksTex >= 0 ? ts->getMax(b.ksTex) : m.ks) + (b.kdTex >= 0 ? ts->getMax(b.kdTex) : m.kd);
        float actualMax = std::max(std::max(maxValue.x, maxValue.y), maxValue.z);
        b.scale = actualMax > 1.0f ? 0.99f * (1.0f / actualMax) : 1.0f;
        float dAvg = getLuminance((b.kdTex >= 0 ? ts->getAverage(b.kdTex) : m.kd) * b.scale);
        float sAvg = getLuminance((b.ksTex >= 0 ? ts->getAverage(b.ksTex) : m.ks) * b.scale);
        b.specularSamplingWeight = sAvg / (dAvg + sAvg);
    }
    return b;
}

}  // namespace

namespace {

/* ---- externals/bvh.h -------------------------------------------------------------------------------------- */
struct BBox { /* bvh.h:28-91 */
    v3 min, max, extent;
    BBox() : min(V(0)), max(V(0)), extent(V(0)) {}
    explicit BBox(v3 p) : min(p), max(p) { extent = max - min; }
    void expandToInclude(v3 p) { min = vmin(min, p); max = vmax(max, p); extent = max - min; }
    void expandToInclude(const BBox& b) { min = vmin(min, b.min); max = vmax(max, b.max); extent = max - min; }
    uint32_t maxDimension() const { /* bvh.h:82-87, including the y/z comparison as written */
        uint32_t result = 0;
        if (extent.y > extent.x) result = 1;
        if (extent.z > extent.y) result = 2;
        return result;
    }
    bool intersect(const Ray& r) const { /* bvh.h:33-69; never reports distances */
        float tmin = (min.x - r.o.x) / r.d.x, tmax = (max.x - r.o.x) / r.d.x;
        if (tmin > tmax) std::swap(tmin, tmax);
        float tymin = (min.y - r.o.y) / r.d.y, tymax = (max.y - r.o.y) / r.d.y;
        if (tymin > tymax) std::swap(tymin, tymax);
        if ((tmin > tymax) || (tymin > tmax)) return false;
        if (tymin > tmin) tmin = tymin;
        if (tymax < tmax) tmax = tymax;
        float tzmin = (min.z - r.o.z) / r.d.z, tzmax = (max.z - r.o.z) / r.d.z;
        if (tzmin > tzmax) std::swap(tzmin, tzmax);
        if ((tmin > tzmax) || (tzmin > tmax)) return false;
        return true;
    }
};

struct FlatNode { BBox bbox; uint32_t start, nPrims, rightOffset; }; /* bvh.h:102-105 */

struct IntersectionInfo { float t = 0, u = 0, v = 0; int object = -1; /* index into prims, -1 = nullptr */ };

struct Emitter { /* core.h:324-332 */
    size_t shapeID; float area; v3 radiance; std::vector<float> cdf;
};

struct Scene {
    std::vector<v3> vertices, normals;
    std::vector<int> tri_v, tri_n, tri_mat, tri_shape;  /* shape-major */
    std::vector<int> shapeFirst;                        /* first global triangle of each shape */
    std::vector<Material> materials;
    TexSet tex;
    std::vector<BSDF> bsdfs;
    std::vector<Emitter> emitters;
    std::vector<int> prims;        /* build_prims: global triangle ids, reordered by the build */
    std::vector<FlatNode> flatTree;
    /* camera / film (Config) */
    v3 cam_o = V(0, 0.8f, 3.8f), cam_at = V(0, 0.8f, 0), cam_up = V(0, 1, 0);
    float fov = 30.f;
    int width = 800, height = 600, spp = 1, rrDepth = 5;
    /* PathTracerIntegrator settings (main.cpp:94-102 defaults) */
    bool ptExplicit = true; int ptMaxDepth = -1; float ptRRProb = 0.95f; int ptEmitterSamples = 1, ptBsdfSamples = 0;
    int ptDepthCap = 0;  /* > 0: the device's hard cap on the recursion depth (the reference has none); 0 = none */
    /* DirectIntegrator settings (main.cpp:88-93): 0 mis, 1 area, 2 solidAngle, 3 cosineHemisphere, 4 bsdf */
    int diStrategy = 0, diEmitterSamples = 1, diBsdfSamples = 1;
    std::vector<v3> shapesCenter; std::vector<float> shapesMaxX;  /* renderer.cpp:294-304, :349-358 */
    bool rrOn = false;   /* false = NO_RR 1, the shipped build */
    int rrMaxDepth = 0;  /* 0 = no cap (the reference) */

    v3 vert(int t, int k) const { return vertices[tri_v[3 * t + k]]; }
    v3 norm(int t, int k) const { return normals[tri_n[3 * t + k]]; }

    /* accel.h:71-106 */
    BBox triBBox(int t) const { BBox b(vert(t, 0)); b.expandToInclude(vert(t, 1)); b.expandToInclude(vert(t, 2)); return b; }
    v3 triCentroid(int t) const { return (vert(t, 0) + vert(t, 1) + vert(t, 2)) / 3.0f; }

    /* BVH::build, bvh.h:147-247 (explicit stack, right pushed before left, in-place partition) */
    void buildBVH() {
        const uint32_t Untouched = 0xffffffff, TouchedTwice = 0xfffffffd, leafSize = 4;
        prims.resize(tri_mat.size());
        for (size_t i = 0; i < prims.size(); i++) prims[i] = int(i); /* accel.h:115-120: shape-major creation order */
        struct Entry { uint32_t parent, start, end; };
        std::vector<Entry> todo(128);
        uint32_t stackptr = 0, nNodes = 0;
        todo[stackptr] = {0xfffffffc, 0, uint32_t(prims.size())};
        stackptr++;
        std::vector<FlatNode> buildnodes;
        buildnodes.reserve(prims.size() * 2);
        while (stackptr > 0) {
            Entry bnode = todo[--stackptr];
            uint32_t start = bnode.start, end = bnode.end, nPrims = end - start;
            nNodes++;
            FlatNode node;
            node.start = start; node.nPrims = nPrims; node.rightOffset = Untouched;
            BBox bb = triBBox(prims[start]);
            BBox bc(triCentroid(prims[start]));
            for (uint32_t p = start + 1; p < end; ++p) {
                bb.expandToInclude(triBBox(prims[p]));
                bc.expandToInclude(triCentroid(prims[p]));
            }
            node.bbox = bb;
            if (nPrims <= leafSize) node.rightOffset = 0;
            buildnodes.push_back(node);
            if (bnode.parent != 0xfffffffc) {
                buildnodes[bnode.parent].rightOffset--;
                if (buildnodes[bnode.parent].rightOffset == TouchedTwice) buildnodes[bnode.parent].rightOffset = nNodes - 1 - bnode.parent;
            }
            if (node.rightOffset == 0) continue;
            uint32_t split_dim = bc.maxDimension();
            float split_coord = .5f * (at(bc.min, split_dim) + at(bc.max, split_dim));
            uint32_t mid = start;
            for (uint32_t i = start; i < end; ++i) {
                if (at(triCentroid(prims[i]), split_dim) < split_coord) { std::swap(prims[i], prims[mid]); ++mid; }
            }
            if (mid == start || mid == end) mid = start + (end - start) / 2;
            if (stackptr + 2 > todo.size()) todo.resize(todo.size() * 2);
            todo[stackptr++] = {nNodes - 1, mid, end};
            todo[stackptr++] = {nNodes - 1, start, mid};
        }
        flatTree = buildnodes;
    }

    /* rayTriangleIntersect, core.h:379-400 */
    static bool rayTriangleIntersect(const Ray& r, v3 v0, v3 v1, v3 v2, float& t, float& u, float& v) {
        v3 v0v1 = v1 - v0, v0v2 = v2 - v0;
        v3 pvec = cross(r.d, v0v2);
        float det = dot(v0v1, pvec);
        if (std::fabs(det) < kEps) return false;
        float invDet = 1 / det;
        v3 tvec = r.o - v0;
        u = dot(tvec, pvec) * invDet;
        if (u < 0 || u > 1) return false;
        v3 qvec = cross(tvec, v0v1);
        v = dot(r.d, qvec) * invDet;
        if (v < 0 || u + v > 1) return false;
        t = dot(v0v2, qvec) * invDet;
        return true;
    }
    /* AcceleratorBVH::BVHNode::getIntersection, accel.h:27-52 */
    bool primIntersect(int prim, const Ray& ray, IntersectionInfo* I) const {
        int tr = prims[prim];
        float t, u, v;
        if (rayTriangleIntersect(ray, vert(tr, 0), vert(tr, 1), vert(tr, 2), t, u, v)) {
            if (t > 1e-3) { I->t = t; I->u = u; I->v = v; I->object = prim; return true; }
        }
        return false;
    }
    /* BVH::getIntersection, bvh.h:259-352: bbhits stay 0 because BBox::intersect never writes them */
    bool getIntersection(const Ray& ray, IntersectionInfo* intersection, bool occlusion) const {
        intersection->t = ray.max_t;
        intersection->object = -1;
        float bbhits[4] = {};
        int32_t closer, other;
        struct Trav { uint32_t i; float mint; };
        Trav todo[64];
        int32_t stackptr = 0;
        todo[stackptr].i = 0;
        todo[stackptr].mint = ray.min_t;
        while (stackptr >= 0) {
            int ni = todo[stackptr].i;
            float near = todo[stackptr].mint;
            stackptr--;
            const FlatNode& node = flatTree[ni];
            if (near > intersection->t) continue;
            if (node.rightOffset == 0) {
                for (uint32_t o = 0; o < node.nPrims; ++o) {
                    IntersectionInfo current;
                    bool hit = primIntersect(int(node.start + o), ray, &current);
                    if (hit) {
                        if (occlusion && current.t <= ray.max_t && current.t >= ray.min_t) return true;
                        if (current.t < intersection->t) *intersection = current;
                    }
                }
            } else {
                bool hitc0 = flatTree[ni + 1].bbox.intersect(ray);
                bool hitc1 = flatTree[ni + node.rightOffset].bbox.intersect(ray);
                if (hitc0 && hitc1) {
                    closer = ni + 1;
                    other = ni + node.rightOffset;
                    if (bbhits[2] < bbhits[0]) { std::swap(bbhits[0], bbhits[2]); std::swap(bbhits[1], bbhits[3]); std::swap(closer, other); }
                    todo[++stackptr] = {uint32_t(other), bbhits[2]};
                    todo[++stackptr] = {uint32_t(closer), bbhits[0]};
                } else if (hitc0) {
                    todo[++stackptr] = {uint32_t(ni + 1), bbhits[0]};
                } else if (hitc1) {
                    todo[++stackptr] = {uint32_t(ni + node.rightOffset), bbhits[2]};
                }
            }
        }
        return intersection->object != -1;
    }
    /* AcceleratorBVH::intersect, accel.h:125-172 */
    bool intersect(const Ray& ray, SurfaceInteraction& info) const {
        IntersectionInfo iInfo;
        if (getIntersection(ray, &iInfo, false)) {
            info.t = iInfo.t;
            if (iInfo.t <= ray.max_t && iInfo.t >= ray.min_t) {
                int tr = prims[iInfo.object];
                v3 v0 = vert(tr, 0), v1 = vert(tr, 1), v2 = vert(tr, 2), n0 = norm(tr, 0), n1 = norm(tr, 1), n2 = norm(tr, 2);
                info.shapeID = size_t(tri_shape[tr]);
                info.primID = size_t(tr - shapeFirst[tri_shape[tr]]);
                info.t = iInfo.t; info.u = iInfo.u; info.v = iInfo.v;
                info.p = barycentric(v0, v1, v2, iInfo.u, iInfo.v);
                info.ng = normalize(cross(v1 - v0, v2 - v0));
                info.frameNs = Frame(normalize(barycentric(n0, n1, n2, info.u, info.v)));
                info.wo = info.frameNs.toLocal(-ray.d);
                info.matID = tri_mat[tr];
                return true;
            }
            return false;
        }
        info.t = std::numeric_limits<float>::max();
        return false;
    }

    /* Scene::load tail: BSDFs (renderer.cpp:259-271), emitters (:279-305), getShapeArea (:317-339) */
    void finish() {
        bsdfs.clear();
        tex.shapeFirst = &shapeFirst;
        for (auto& m : materials) bsdfs.push_back(makeBSDF(m, &tex));
        size_t nShapes = tri_shape.empty() ? 0 : size_t(tri_shape.back()) + 1;
        shapeFirst.assign(nShapes + 1, 0);
        for (size_t t = 0; t < tri_shape.size(); t++) shapeFirst[tri_shape[t] + 1] = int(t + 1);
        for (size_t s = 1; s <= nShapes; s++) shapeFirst[s] = std::max(shapeFirst[s], shapeFirst[s - 1]);
        emitters.clear();
        for (size_t s = 0; s < nShapes; s++) {
            int first = shapeFirst[s], last = shapeFirst[s + 1];
            if (first == last) continue;
            const BSDF& b = bsdfs[tri_mat[first]];
            if (!(length2(b.emission) > 0.f)) continue; /* BSDF::isEmissive, core.h:306-308 */
            Emitter e;
            e.shapeID = s; e.radiance = b.emission; e.cdf.push_back(0.f);
            for (int t = first; t < last; t++) {
                v3 e1 = vert(t, 1) - vert(t, 0), e2 = vert(t, 2) - vert(t, 0), e3 = cross(e1, e2);
                e.cdf.push_back(e.cdf.back() + 0.5f * std::sqrt(e3.x * e3.x + e3.y * e3.y + e3.z * e3.z));
            }
            e.area = e.cdf.back();
            float sum = e.cdf.back(); /* Distribution1D::normalize, math.h:94-101 */
            for (float& c : e.cdf) c /= sum;
            emitters.push_back(e);
        }
        /* Scene::load's tail (renderer.cpp:294-304): centre = running sum over mesh.indices / their number; the AABB's max.x */
        shapesCenter.assign(nShapes, V(0.f)); shapesMaxX.assign(nShapes, -std::numeric_limits<float>::infinity());
        for (size_t sh = 0; sh < nShapes; sh++) {
            v3 c = V(0.f);
            for (int t = shapeFirst[sh]; t < shapeFirst[sh + 1]; t++)
                for (int k = 0; k < 3; k++) { v3 p = vert(t, k); c = c + p; shapesMaxX[sh] = std::max(shapesMaxX[sh], p.x); }
            shapesCenter[sh] = c / float(3 * (shapeFirst[sh + 1] - shapeFirst[sh]));
        }
        buildBVH();
    }
    v3 getShapeCenter(size_t shapeID) const { return shapesCenter[shapeID]; }
    float getShapeRadius(size_t shapeID) const { return shapesMaxX[shapeID] - shapesCenter[shapeID].x; } /* renderer.cpp:349-353 */
};

}  // namespace

namespace {

/* ---- camera matrices with glm's operation order (renderer.cpp:143-153, bdpt.h:485-490) --------------------- */
struct M4 { float m[4][4]; };  /* m[col][row] */
inline M4 ident() { M4 r; std::memset(&r, 0, sizeof(r)); for (int i = 0; i < 4; i++) r.m[i][i] = 1.f; return r; }
inline void mulv(const M4& M, const float v[4], float o[4]) { /* (m0*x + m1*y) + (m2*z + m3*w) */
    for (int r = 0; r < 4; r++) o[r] = (M.m[0][r] * v[0] + M.m[1][r] * v[1]) + (M.m[2][r] * v[2] + M.m[3][r] * v[3]);
}
inline M4 mulm(const M4& A, const M4& B) {
    M4 R;
    for (int j = 0; j < 4; j++) for (int r = 0; r < 4; r++)
        R.m[j][r] = ((A.m[0][r] * B.m[j][0] + A.m[1][r] * B.m[j][1]) + A.m[2][r] * B.m[j][2]) + A.m[3][r] * B.m[j][3];
    return R;
}
inline M4 lookAt(v3 eye, v3 center, v3 up) {
    v3 f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f);
    M4 R = ident();
    R.m[0][0] = s.x; R.m[1][0] = s.y; R.m[2][0] = s.z;
    R.m[0][1] = u.x; R.m[1][1] = u.y; R.m[2][1] = u.z;
    R.m[0][2] = -f.x; R.m[1][2] = -f.y; R.m[2][2] = -f.z;
    R.m[3][0] = -dot(s, eye); R.m[3][1] = -dot(u, eye); R.m[3][2] = dot(f, eye);
    return R;
}
inline M4 perspective(float fovy, float aspect, float zn, float zf) {
    float th = std::tan(fovy / 2.f);
    M4 R; std::memset(&R, 0, sizeof(R));
    R.m[0][0] = 1.f / (aspect * th); R.m[1][1] = 1.f / th; R.m[2][2] = -(zf + zn) / (zf - zn); R.m[2][3] = -1.f;
    R.m[3][2] = -(2.f * zf * zn) / (zf - zn);
    return R;
}
inline M4 scaleM(float x, float y, float z) { M4 R = ident(); R.m[0][0] *= x; R.m[1][1] *= y; R.m[2][2] *= z; return R; }
inline M4 translateM(float x, float y, float z) {
    M4 I = ident(), R = I;
    for (int r = 0; r < 4; r++) R.m[3][r] = ((I.m[0][r] * x + I.m[1][r] * y) + I.m[2][r] * z) + I.m[3][r];
    return R;
}
M4 inverse(const M4& M) { /* glm::inverse(mat4): cofactors, then one reciprocal determinant */
    auto m = [&](int c, int r) { return M.m[c][r]; };
    float c00 = m(2,2)*m(3,3) - m(3,2)*m(2,3), c02 = m(1,2)*m(3,3) - m(3,2)*m(1,3), c03 = m(1,2)*m(2,3) - m(2,2)*m(1,3);
    float c04 = m(2,1)*m(3,3) - m(3,1)*m(2,3), c06 = m(1,1)*m(3,3) - m(3,1)*m(1,3), c07 = m(1,1)*m(2,3) - m(2,1)*m(1,3);
    float c08 = m(2,1)*m(3,2) - m(3,1)*m(2,2), c10 = m(1,1)*m(3,2) - m(3,1)*m(1,2), c11 = m(1,1)*m(2,2) - m(2,1)*m(1,2);
    float c12 = m(2,0)*m(3,3) - m(3,0)*m(2,3), c14 = m(1,0)*m(3,3) - m(3,0)*m(1,3), c15 = m(1,0)*m(2,3) - m(2,0)*m(1,3);
    float c16 = m(2,0)*m(3,2) - m(3,0)*m(2,2), c18 = m(1,0)*m(3,2) - m(3,0)*m(1,2), c19 = m(1,0)*m(2,2) - m(2,0)*m(1,2);
    float c20 = m(2,0)*m(3,1) - m(3,0)*m(2,1), c22 = m(1,0)*m(3,1) - m(3,0)*m(1,1), c23 = m(1,0)*m(2,1) - m(2,0)*m(1,1);
    float F0[4] = {c00,c00,c02,c03}, F1[4] = {c04,c04,c06,c07}, F2[4] = {c08,c08,c10,c11}, F3[4] = {c12,c12,c14,c15}, F4[4] = {c16,c16,c18,c19}, F5[4] = {c20,c20,c22,c23};
    float V0[4] = {m(1,0),m(0,0),m(0,0),m(0,0)}, V1[4] = {m(1,1),m(0,1),m(0,1),m(0,1)}, V2[4] = {m(1,2),m(0,2),m(0,2),m(0,2)}, V3[4] = {m(1,3),m(0,3),m(0,3),m(0,3)};
    float SA[4] = {1,-1,1,-1}, SB[4] = {-1,1,-1,1};
    M4 I;
    for (int k = 0; k < 4; k++) {
        I.m[0][k] = ((V1[k]*F0[k] - V2[k]*F1[k]) + V3[k]*F2[k]) * SA[k];
        I.m[1][k] = ((V0[k]*F0[k] - V2[k]*F3[k]) + V3[k]*F4[k]) * SB[k];
        I.m[2][k] = ((V0[k]*F1[k] - V1[k]*F3[k]) + V3[k]*F5[k]) * SA[k];
        I.m[3][k] = ((V0[k]*F2[k] - V1[k]*F4[k]) + V2[k]*F5[k]) * SB[k];
    }
    float det = (m(0,0)*I.m[0][0] + m(0,1)*I.m[1][0]) + (m(0,2)*I.m[2][0] + m(0,3)*I.m[3][0]);
    float inv = 1.f / det;
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) I.m[c][r] *= inv;
    return I;
}

struct PathVertex { SurfaceInteraction hit; v3 throughput; float vcm, vc, rr; }; /* bdpt.h:24-35 */

/* BDPTIntegrator (bdpt.h:37-518) with LIGHT_TRACING=0, PATH_TRACING=0, NO_RR=1 as the file is shipped (NO_RR=0 selectable, see m_noRR) */
struct BDPT {
    const Scene& scene;
    std::vector<v3>* rgb;  /* RenderBuffer */
    int m_rrDepth;
    /* NO_RR is a #define (bdpt.h:18, shipped as 1); here a run-time switch so that the reference's Russian-roulette code
     * (NO_RR 0) can be checked too.  m_maxDepth > 0 adds the device's hard depth cap, which the reference lacks (bdpt.h:66-67). */
    bool m_noRR;
    int m_maxDepth;
    bool belowCap(int depth) const { return m_maxDepth <= 0 || depth < m_maxDepth; }
    explicit BDPT(const Scene& s, std::vector<v3>* fb) : scene(s), rgb(fb), m_rrDepth(s.rrDepth), m_noRR(!s.rrOn), m_maxDepth(s.rrMaxDepth) {}
    float deg2radFov() const { return kPi / 180.f * scene.fov; }

    /* integrator.cpp:35-100 */
    const BSDF* getBSDF(const SurfaceInteraction& h) const { return &scene.bsdfs[scene.tri_mat[scene.shapeFirst[h.shapeID] + int(h.primID)]]; }
    v3 getEmission(const SurfaceInteraction& h) const { return scene.materials[h.matID].ke; }
    size_t selectEmitter(float sample, float& pdf) const {
        size_t id = size_t(sample * scene.emitters.size());
        id = std::min(id, scene.emitters.size() - 1);
        pdf = 1.f / scene.emitters.size();
        return id;
    }
    size_t getEmitterIDByShapeID(size_t shapeID) const {
        for (size_t i = 0; i < scene.emitters.size(); i++) if (scene.emitters[i].shapeID == shapeID) return i;
        return scene.emitters.size();
    }
    void sampleEmitterPosition(Sampler& sampler, const Emitter& em, v3& n, v3& pos, float& pdf) const {
        float s = sampler.next();
        auto it = std::upper_bound(em.cdf.begin(), em.cdf.end(), s); /* Distribution1D::sample, math.h:107-111 */
        int primID = std::min(std::max(int(it - em.cdf.begin()) - 1, 0), int(em.cdf.size()) - 2);
        float sx, sy; sampler.next2D(&sx, &sy);
        float u = std::sqrt(1.f - sx); /* squareToUniformTriangle, math.h:229-234 */
        float bu = 1 - u, bv = u * sy;
        int t = scene.shapeFirst[em.shapeID] + primID;
        pos = barycentric(scene.vert(t, 0), scene.vert(t, 1), scene.vert(t, 2), bu, bv);
        n = normalize(barycentric(scene.norm(t, 0), scene.norm(t, 1), scene.norm(t, 2), bu, bv));
        pdf = 1.f / em.area;
    }

    bool visibilityQuery(v3 start, v3 end) const { /* bdpt.h:498-506 */
        v3 dir = end - start;
        float dist = length(dir);
        dir = dir / dist;
        IntersectionInfo info;
        Ray ray(start, dir, kEps, dist - 0.00001f);
        return scene.getIntersection(ray, &info, true);
    }
    void splatToImagePlane(v3 position, int& xPixel, int& yPixel) const { /* bdpt.h:485-496 */
        float aspectRatio = (float)scene.width / scene.height;
        M4 worldToCamera = lookAt(scene.cam_o, scene.cam_at, scene.cam_up);
        M4 cameraToClip = perspective(deg2radFov(), aspectRatio, 1.f, 1000.f);
        M4 NDCToScreen = mulm(mulm(scaleM(float(scene.width), float(scene.height), 1.0f), scaleM(0.5f, -0.5f, 1.f)), translateM(1.f, -1.f, 0.f));
        float uv[4] = {position.x, position.y, position.z, 1.f}, t[4];
        mulv(worldToCamera, uv, t); mulv(cameraToClip, t, uv);
        float w = uv[3];
        for (int k = 0; k < 4; k++) uv[k] /= w;
        mulv(NDCToScreen, uv, t);
        xPixel = static_cast<int>(t[0]);
        yPixel = static_cast<int>(t[1]);
    }

    bool ContinuePathRandomWalk(SurfaceInteraction& hit, Sampler& sampler, float rrProbability, PathVertex& pathVertex, v3& throughput,
                                int& depth, float& vc, float& vcm, Ray& wi) const { /* bdpt.h:243-291 */
        const BSDF* bsdf = getBSDF(hit);
        bool isDeltaBsdf = (bsdf->combinedType & EDelta);
        float bsdfPdf_w, sx, sy;
        sampler.next2D(&sx, &sy);
        v3 bsdfCosTheta = bsdf->sample(hit, sx, sy, &bsdfPdf_w);
        bsdfPdf_w *= rrProbability;
        pathVertex.hit.wi = hit.wi;
        float absCosOut = std::fabs(hit.wi.z);
        if (bsdfCosTheta == V(0.f)) return false;
        throughput = throughput * (bsdfCosTheta * (1.f / bsdfPdf_w));
        depth++;
        SurfaceInteraction reverseHit(hit);
        std::swap(reverseHit.wi, reverseHit.wo);
        float previousVertexReversePdf_w = isDeltaBsdf ? bsdfPdf_w : bsdf->pdf(reverseHit) * rrProbability;
        if (isDeltaBsdf) {
            vc = (absCosOut / bsdfPdf_w) * (previousVertexReversePdf_w * vc);
            vcm = 0.f;
        } else {
            vc = (absCosOut / bsdfPdf_w) * (vcm + previousVertexReversePdf_w * vc);
            vcm = 1.f / bsdfPdf_w;
        }
        wi = Ray(hit.p, hit.frameNs.toWorld(hit.wi));
        return true;
    }

    void connectToCamera(const PathVertex& lightVertex) const { /* bdpt.h:295-371 */
        v3 cameraForward = normalize(scene.cam_at - scene.cam_o);
        v3 eyeToLightVert = lightVertex.hit.p - scene.cam_o;
        float invDistanceSquared = 1.f / length2(eyeToLightVert);
        eyeToLightVert = eyeToLightVert * std::sqrt(invDistanceSquared);
        int xPixel, yPixel;
        splatToImagePlane(lightVertex.hit.p, xPixel, yPixel);
        if (xPixel < 0 || yPixel < 0 || xPixel >= scene.width || yPixel >= scene.height) return;
        float cosCamera = dot(cameraForward, eyeToLightVert);
        if (cosCamera <= 0.f) return;
        SurfaceInteraction ei(lightVertex.hit);
        ei.wi = ei.frameNs.toLocal(-eyeToLightVert);
        v3 bsdfCosTheta = getBSDF(ei)->eval(ei);
        if (bsdfCosTheta == V(0.f) || ei.wi.z <= 0.f) return;
        if (visibilityQuery(scene.cam_o, lightVertex.hit.p)) return;
        float virtualNearPlaneDistance = (1.f / std::tan(deg2radFov() * 0.5f)) * scene.height * 0.5f;
        float imagePointToCameraDist = virtualNearPlaneDistance / cosCamera;
        float imageAreaToCameraSolidAngle = imagePointToCameraDist * imagePointToCameraDist / cosCamera;
        float cameraSolidAngleToSurfaceArea = ei.wi.z * invDistanceSquared;
        float imageAreaToSurfaceArea = imageAreaToCameraSolidAngle * cameraSolidAngleToSurfaceArea;
        float surfaceAreaToImageArea = 1.f / imageAreaToSurfaceArea;
        int nlightSubpath = scene.width * scene.height;
        v3 radiance = lightVertex.throughput * (bsdfCosTheta * (1.f / ei.wi.z)) * (1.f / surfaceAreaToImageArea) * (1.f / nlightSubpath) * (1.f / scene.spp);
        float reversePdf_a = 1.f * imageAreaToSurfaceArea;
        SurfaceInteraction reverseHit(ei);
        std::swap(reverseHit.wi, reverseHit.wo);
        float previousVertexReversePdf_w = getBSDF(reverseHit)->pdf(reverseHit) * lightVertex.rr;
        float lightWeight = (reversePdf_a / nlightSubpath) * (lightVertex.vcm + previousVertexReversePdf_w * lightVertex.vc);
        float eyeWeight = 0.f;
        float misWeight = 1.f / (lightWeight + 1.f + eyeWeight);
        radiance = radiance * misWeight;
        int pixel = yPixel * scene.width + xPixel;
        (*rgb)[pixel] = (*rgb)[pixel] + radiance; /* under the per-pixel lock in the reference, bdpt.h:363-370 */
    }

    v3 connectToLight(const PathVertex& eyeVertex, Sampler& sampler) const { /* bdpt.h:374-430 */
        float emitterPdf;
        size_t emitterID = selectEmitter(sampler.next(), emitterPdf);
        const Emitter& emitter = scene.emitters[emitterID];
        v3 emitterNormal, emitterPosition;
        float emitterPositionPdf;
        sampleEmitterPosition(sampler, emitter, emitterNormal, emitterPosition, emitterPositionPdf);
        v3 l2e = eyeVertex.hit.p - emitterPosition;
        float d2 = length2(l2e);
        l2e = l2e * (1.f / std::sqrt(d2));
        SurfaceInteraction ei(eyeVertex.hit);
        ei.wi = ei.frameNs.toLocal(-l2e);
        float cosAtLight = dot(emitterNormal, l2e), cosAtEyeVertex = ei.wi.z;
        if (cosAtLight <= 0.f || cosAtEyeVertex <= 0.f) return V(0.f);
        float emitterConnectPdf_a = emitterPdf * emitterPositionPdf;
        float emitterConnectPdf_w = emitterConnectPdf_a * d2 / cosAtLight;
        float emitterDirectionPdf_w = kInvTwoPi;
        v3 Li = getBSDF(ei)->eval(ei) * (1.f / emitterConnectPdf_w) * eyeVertex.throughput * emitter.radiance;
        if (Li == V(0.f)) return V(0.f);
        if (visibilityQuery(eyeVertex.hit.p, emitterPosition)) return V(0.f);
        float lightPathReversePdf_w = getBSDF(ei)->pdf(ei) * eyeVertex.rr;
        float lightWeight = lightPathReversePdf_w / emitterConnectPdf_w;
        SurfaceInteraction ri(ei);
        std::swap(ri.wi, ri.wo);
        float eyePathPreviousVertexReversePdf_w = getBSDF(ri)->pdf(ri) * eyeVertex.rr;
        float eyePathCurrentVertexReversePdf_a = cosAtEyeVertex * (1.f / d2) * emitterDirectionPdf_w;
        float eyeWeight = eyePathCurrentVertexReversePdf_a * (eyeVertex.vcm + eyePathPreviousVertexReversePdf_w * eyeVertex.vc);
        float misWeight = 1.f / (lightWeight + 1.f + eyeWeight);
        return Li * misWeight;
    }

    v3 connectVertices(const PathVertex& lightVertex, const PathVertex& eyeVertex) const { /* bdpt.h:434-483 */
        v3 l2e = eyeVertex.hit.p - lightVertex.hit.p;
        float invDistanceSquared = 1.f / length2(l2e);
        l2e = l2e * std::sqrt(invDistanceSquared);
        SurfaceInteraction lh(lightVertex.hit);
        lh.wi = lh.frameNs.toLocal(l2e);
        SurfaceInteraction eh(eyeVertex.hit);
        eh.wi = eh.frameNs.toLocal(-l2e);
        float cosAtLightVertex = lh.wi.z, cosAtEyeVertex = eh.wi.z;
        if (cosAtLightVertex <= 0.f || cosAtEyeVertex <= 0.f) return V(0.f);
        if (visibilityQuery(eyeVertex.hit.p, lightVertex.hit.p)) return V(0.f);
        v3 Li = getBSDF(lh)->eval(lh) * getBSDF(eh)->eval(eh);
        Li = Li * (lightVertex.throughput * eyeVertex.throughput * invDistanceSquared);
        float eyePathReversePdf_w = getBSDF(lh)->pdf(lh) * lightVertex.rr;
        std::swap(lh.wi, lh.wo);
        float lightPathPreviousVertexReversePdf_w = getBSDF(lh)->pdf(lh) * lightVertex.rr;
        float lightPathReversePdf_w = getBSDF(eh)->pdf(eh) * eyeVertex.rr;
        eh.wo = eh.wi;
        eh.wi = eyeVertex.hit.wo;
        float eyePathPreviousVertexReversePdf_w = getBSDF(eh)->pdf(eh) * eyeVertex.rr;
        float lightPathReversePdf_a = lightPathReversePdf_w * (cosAtLightVertex) * invDistanceSquared;
        float eyePathReversePdf_a = eyePathReversePdf_w * (cosAtEyeVertex) * invDistanceSquared;
        float lightWeight = lightPathReversePdf_a * (lightVertex.vcm + lightPathPreviousVertexReversePdf_w * lightVertex.vc);
        float eyeWeight = eyePathReversePdf_a * (eyeVertex.vcm + eyePathPreviousVertexReversePdf_w * eyeVertex.vc);
        float misWeight = 1.f / (lightWeight + 1.f + eyeWeight);
        return Li * misWeight;
    }

    void lightSubpathWalk(Sampler& sampler, std::vector<PathVertex>& lightSubpath) const { /* bdpt.h:158-217 */
        float emitterPdf, emitterAreaPdf, emitterEmissionPdf;
        v3 normalOut, positionOut;
        size_t id = selectEmitter(sampler.next(), emitterPdf);
        const Emitter& emitter = scene.emitters[id];
        sampleEmitterPosition(sampler, emitter, normalOut, positionOut, emitterAreaPdf);
        float sx, sy; sampler.next2D(&sx, &sy);
        v3 emitterDirection = squareToUniformHemisphere(sx, sy);
        emitterEmissionPdf = kInvTwoPi * emitterAreaPdf;
        emitterAreaPdf *= emitterPdf;
        emitterEmissionPdf *= emitterPdf;
        Frame lightFrame(normalOut);
        Ray wi(positionOut, lightFrame.toWorld(emitterDirection));
        v3 throughput = emitterDirection.z * emitter.radiance * (1.f / emitterEmissionPdf);
        float vc = emitterDirection.z * (1.f / emitterEmissionPdf);
        float vcm = emitterAreaPdf / emitterEmissionPdf;
        if (emitterDirection.z <= 0.f) return;
        int depth = 1;
        float rrProbability = 1.f;
        while (depth < m_rrDepth || (sampler.next() < rrProbability && !m_noRR && belowCap(depth))) {
            SurfaceInteraction hit;
            if (!scene.intersect(wi, hit)) break;
            const float distSquared = hit.t * hit.t, absCosIn = std::fabs(hit.wo.z);
            vcm *= (distSquared / absCosIn);
            vc *= (1.f / absCosIn);
            rrProbability = (depth + 1) < m_rrDepth ? 1.f : (getLuminance(throughput) < 0.01f ? 0.5f : 1.f); /* bdpt.h:129,201 */
            if (m_noRR) rrProbability = 1.f;                                                                  /* bdpt.h:130-132,202-204 */
            unsigned hitBSDFType = getBSDF(hit)->combinedType;
            PathVertex lightVertex{hit, throughput, vcm, vc, rrProbability};
            if (!(hitBSDFType & EDelta)) connectToCamera(lightVertex);
            if (!ContinuePathRandomWalk(hit, sampler, rrProbability, lightVertex, throughput, depth, vc, vcm, wi)) break;
            if (!(hitBSDFType & EDelta)) lightSubpath.push_back(lightVertex);
        }
    }

    void eyeSubpathWalk(const std::vector<PathVertex>& lightSubpath, const Ray& ray, Sampler& sampler, v3& Li) const { /* bdpt.h:46-155 */
        bool isPathPureSpecular = true;
        v3 cameraForward = normalize(scene.cam_at - scene.cam_o);
        float cosCamera = dot(cameraForward, ray.d);
        float virtualNearPlaneDistance = (1.f / std::tan(deg2radFov() * 0.5f)) * scene.height * 0.5f;
        float imagePointToCameraDist = virtualNearPlaneDistance / cosCamera;
        float imageAreaToCameraSolidAngle = imagePointToCameraDist * imagePointToCameraDist / cosCamera;
        float t1Pdf = 1.f * imageAreaToCameraSolidAngle;
        Ray wi(ray);
        v3 throughput = V(1.f);
        float vc = 0.f;
        float vcm = scene.width * scene.height * (1.f / t1Pdf);
        int depth = 1;
        float rrProbability = 1.f;
        while (depth < m_rrDepth || (sampler.next() < rrProbability && !m_noRR && belowCap(depth))) {
            SurfaceInteraction hit;
            if (!scene.intersect(wi, hit)) break;
            const float distSquared = hit.t * hit.t, absCosIn = std::fabs(hit.wo.z);
            vcm *= (distSquared / absCosIn);
            vc *= (1.f / absCosIn);
            if (getEmission(hit) != V(0.f)) {
                const Emitter& emitter = scene.emitters[getEmitterIDByShapeID(hit.shapeID)];
                float emitterPdf = 1.f / scene.emitters.size();
                if (depth > 1) {
                    v3 contribution = emitter.radiance * throughput;
                    float emitterPositionPdf_a = 1.f / (emitter.area * emitterPdf);
                    float emitterDirectionPdf_w = kInvTwoPi;
                    float cameraWeight = emitterPositionPdf_a * vcm + (emitterPositionPdf_a * emitterDirectionPdf_w) * vc;
                    float misWeight = 1.f / (1.f + cameraWeight);
                    if (!isPathPureSpecular) contribution = contribution * misWeight;
                    Li = Li + contribution;
                } else if (depth == 1) {
                    Li = Li + getEmission(hit);
                }
                break;
            }
            rrProbability = (depth + 1) < m_rrDepth ? 1.f : (getLuminance(throughput) < 0.01f ? 0.5f : 1.f); /* bdpt.h:129,201 */
            if (m_noRR) rrProbability = 1.f;                                                                  /* bdpt.h:130-132,202-204 */
            PathVertex eyeVertex{hit, throughput, vcm, vc, rrProbability};
            unsigned hitBSDFType = getBSDF(hit)->combinedType;
            if (!(hitBSDFType & EDelta)) {
                isPathPureSpecular = false;
                Li = Li + connectToLight(eyeVertex, sampler);
                for (size_t i = 0; i < lightSubpath.size(); i++) Li = Li + connectVertices(lightSubpath[i], eyeVertex);
            }
            if (!ContinuePathRandomWalk(hit, sampler, rrProbability, eyeVertex, throughput, depth, vc, vcm, wi)) break;
        }
    }

    v3 render(const Ray& ray, Sampler& sampler) const { /* bdpt.h:219-241 */
        v3 Li = V(0.f);
        SurfaceInteraction hit;
        if (scene.intersect(ray, hit)) {
            std::vector<PathVertex> lightSubpath;
            lightSubpathWalk(sampler, lightSubpath);
            eyeSubpathWalk(lightSubpath, ray, sampler, Li);
        }
        return Li;
    }
};


/* PathTracerIntegrator (src/integrators/path.h:15-262): the unidirectional estimator the author cross-checks BDPT with
 * (data/a5/bonus_bdpt/tinyrender/cbox_bdpt_path.toml).  Integrator services are BDPT's (same base class). */
struct PathTracer {
    const Scene& scene;
    BDPT svc;
    bool m_isExplicit; int m_maxDepth, m_rrDepth; float m_rrProb; size_t m_emitterSamples, m_bsdfSamples; int m_cap;
    explicit PathTracer(const Scene& s) : scene(s), svc(s, nullptr), m_isExplicit(s.ptExplicit), m_maxDepth(s.ptMaxDepth), m_rrDepth(s.rrDepth),
        m_rrProb(s.ptRRProb), m_emitterSamples(size_t(s.ptEmitterSamples)), m_bsdfSamples(size_t(s.ptBsdfSamples)), m_cap(s.ptDepthCap) {}
    static float balanceHeuristic(float nf, float fPdf, float ng, float gPdf) { float f = nf * fPdf, g = ng * gPdf; return f / (f + g); } /* :30-33 */
    static Ray mkRay(v3 o, v3 d) { return Ray(o, d, kEps, std::numeric_limits<float>::infinity()); }

    v3 recursiveImplicit(Sampler& sampler, SurfaceInteraction& hit, int depth) const { /* path.h:35-66 */
        if (depth < m_maxDepth) {
            float sx, sy; sampler.next2D(&sx, &sy);
            float pdf;
            v3 brdfCosTheta = svc.getBSDF(hit)->sample(hit, sx, sy, &pdf);
            SurfaceInteraction vis;
            v3 wiW = hit.frameNs.toWorld(hit.wi);
            bool intersects = scene.intersect(mkRay(hit.p, wiW), vis);
            if (intersects) {
                if (svc.getEmission(vis) == V(0.f)) {
                    v3 Li = recursiveImplicit(sampler, vis, depth + 1);
                    return Li * brdfCosTheta * float(1.0 / pdf);
                } else {
                    v3 Li = svc.getEmission(vis);
                    Li = dot(vis.frameNs.n, -wiW) > 0.f ? Li : V(0.f);
                    return Li * brdfCosTheta * float(1.0 / pdf);
                }
            }
            return V(0.f);
        }
        return V(0.f);
    }

    v3 recursiveExplicit(Sampler& sampler, SurfaceInteraction& hit, int depth) const { /* path.h:68-206 */
        v3 Lr = V(0.f);
        float rrSample = sampler.next();
        v3 indirectEstimator = V(0.f), directEstimator = V(0.f);
        if (depth < m_maxDepth || (m_maxDepth == -1 && (depth < m_rrDepth || rrSample < m_rrProb))) {
            {
                unsigned int nSamples = 0;
                float indRRProb = 0.95f, cumRRProb = 1.f;
                SurfaceInteraction vis;
                bool intersects = false;
                float pdf;
                v3 brdfCosTheta;
                do {
                    vis = SurfaceInteraction();
                    float sx, sy; sampler.next2D(&sx, &sy);
                    brdfCosTheta = svc.getBSDF(hit)->sample(hit, sx, sy, &pdf);
                    v3 wiW = hit.frameNs.toWorld(hit.wi);
                    intersects = scene.intersect(mkRay(hit.p, wiW), vis);
                    nSamples++;
                } while (svc.getEmission(vis) != V(0.f) && sampler.next() < indRRProb);
                cumRRProb = nSamples > 1 ? indRRProb : 1.f;
                if (intersects && brdfCosTheta != V(0.f)) {
                    if (svc.getEmission(vis) == V(0.f) && brdfCosTheta != V(0.f) && (m_cap <= 0 || depth + 1 < m_cap)) {
                        v3 Li = recursiveExplicit(sampler, vis, depth + 1);
                        indirectEstimator = Li * brdfCosTheta * (1.f / pdf) * (1.f / nSamples) * (1.f / cumRRProb);
                    }
                }
            }
            {
                v3 emitterEstimator = V(0.f), bsdfEstimator = V(0.f);
                for (unsigned int i = 0; i < m_emitterSamples; ++i) {
                    float emitterPdf, emitterAreaPdf;
                    v3 normalOut, positionOut;
                    size_t id = svc.selectEmitter(sampler.next(), emitterPdf);
                    const Emitter& emitter = scene.emitters[id];
                    svc.sampleEmitterPosition(sampler, emitter, normalOut, positionOut, emitterAreaPdf);
                    v3 wiW = normalize(positionOut - hit.p);
                    v3 wiLocal = hit.frameNs.toLocal(wiW);
                    float lightDistanceSquared = length2(positionOut - hit.p); /* glm::distance2(a, b) = length2(b - a) */
                    float cosOutgoing = dot(-wiW, normalOut);
                    if (cosOutgoing > 0.f && wiLocal.z > 0.f) {
                        SurfaceInteraction vis;
                        bool intersects = scene.intersect(mkRay(hit.p, wiW), vis);
                        if (intersects) {
                            if (vis.shapeID == emitter.shapeID) {
                                v3 Li = svc.getEmission(vis);
                                hit.wi = wiLocal;
                                float areaToSolidAngle = cosOutgoing * (1.f / lightDistanceSquared);
                                float bsdfPdf = svc.getBSDF(hit)->pdf(hit);
                                float weight = balanceHeuristic(float(m_emitterSamples), emitterAreaPdf * emitterPdf * (1.f / areaToSolidAngle), float(m_bsdfSamples), bsdfPdf);
                                emitterEstimator = emitterEstimator + weight * Li * svc.getBSDF(hit)->eval(hit) * (1.f / emitterAreaPdf) * (1.f / emitterPdf) * areaToSolidAngle;
                            }
                        }
                    }
                }
                emitterEstimator = m_emitterSamples == 0 ? V(0.f) : (emitterEstimator / float(m_emitterSamples));
                for (unsigned int i = 0; i < m_bsdfSamples; ++i) {
                    float sx, sy; sampler.next2D(&sx, &sy);
                    float bsdfPdf;
                    v3 brdfCosTheta = svc.getBSDF(hit)->sample(hit, sx, sy, &bsdfPdf);
                    if (brdfCosTheta != V(0.f)) {
                        SurfaceInteraction vis;
                        v3 wiW = hit.frameNs.toWorld(hit.wi);
                        bool intersects = scene.intersect(mkRay(hit.p, wiW), vis);
                        if (intersects) {
                            v3 Li = svc.getEmission(vis);
                            if (Li == V(0.f)) continue;
                            const Emitter& emitter = scene.emitters[svc.getEmitterIDByShapeID(vis.shapeID)];
                            float emitterPdf = 1.f / scene.emitters.size();
                            float emitterAreaPdf = 1.f / emitter.area;
                            float lightDistanceSquared = length2(hit.p - vis.p);
                            float cosOutgoing = dot(-wiW, vis.ng);
                            if (cosOutgoing > 0.f) {
                                float areaToSolidAngle = cosOutgoing * (1.f / lightDistanceSquared);
                                float weight = balanceHeuristic(float(m_bsdfSamples), bsdfPdf, float(m_emitterSamples), emitterPdf * emitterAreaPdf * (1.f / areaToSolidAngle));
                                bsdfEstimator = bsdfEstimator + weight * Li * brdfCosTheta * (1.f / bsdfPdf);
                            }
                        }
                    }
                }
                bsdfEstimator = m_bsdfSamples == 0 ? V(0.f) : (bsdfEstimator / float(m_bsdfSamples));
                directEstimator = emitterEstimator + bsdfEstimator;
            }
            Lr = directEstimator + indirectEstimator;
            if (m_maxDepth == -1 && !(depth < m_rrDepth) && (rrSample < m_rrProb)) Lr = Lr * (1.f / m_rrProb);
        }
        return Lr;
    }

    v3 render(const Ray& ray, Sampler& sampler) const { /* path.h:208-251 */
        SurfaceInteraction hit;
        if (scene.intersect(ray, hit)) {
            if (svc.getEmission(hit) != V(0.f)) return svc.getEmission(hit);
            return m_isExplicit ? recursiveExplicit(sampler, hit, 0) : recursiveImplicit(sampler, hit, 0);
        }
        return V(0.f);
    }
};


/* DirectIntegrator (src/integrators/direct.h:16-467, `type = "direct"`): one-bounce direct illumination with five sampling
 * strategies, written for sphere emitters (centre = shape centre, radius = AABB max.x - centre.x, renderer.cpp:349-358).
 * SURVEY 8(f) N4.  M_PI is platform.h:51's FLOAT constant, so `4 * M_PI * r * r` is fp32. */
inline v3 squareToUniformSphere(float sx, float sy) { /* math.h:119-127 */
    float phi = sx * kPi * 2.0f;
    float cosTheta = 1.f - (2.f * sy);
    float sinTheta = sqrtf(float(std::fmax(1.f - cosTheta * cosTheta, 0)));
    return V(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}
inline bool quadratic(double a, double b, double c, double& t0, double& t1) { /* direct.h:16-35 */
    double discriminant = b * b - 4 * a * c;
    if (discriminant > 0) {
        double sqrtDiscriminant = std::sqrt(discriminant), inv2a = 1 / (2 * a);
        t0 = (-b + sqrtDiscriminant) * inv2a; t1 = (-b - sqrtDiscriminant) * inv2a;
        return true;
    } else if (discriminant == 0) {
        t0 = (-b + std::sqrt(discriminant)) / (2 * a); t1 = t0;
        return true;
    }
    return false;
}
inline bool raySphereIntersect(const Ray& ray_, v3 sphereCenter, float sphereRadius, float& outT) { /* direct.h:37-69 */
    v3 newOrigin = ray_.o - sphereCenter;
    double c = dot(newOrigin, newOrigin) - (sphereRadius * sphereRadius);
    double b = dot(newOrigin, ray_.d) * 2.0;
    double a = dot(ray_.d, ray_.d);
    double r1, r2;
    r1 = ray_.min_t;
    bool intersects = quadratic(a, b, c, r1, r2);
    if (!intersects) return false; /* (r2 is read uninitialised in the reference on this branch; the result is false either way) */
    const double inf = std::numeric_limits<double>::infinity();
    r1 = r1 < 0 ? inf : r1; r2 = r2 < 0 ? inf : r2;
    r1 = (r1 > ray_.min_t && r1 < ray_.max_t) ? r1 : inf;
    r2 = (r2 > ray_.min_t && r2 < ray_.max_t) ? r2 : inf;
    double closestT = std::fmin(r1, r2);
    if (closestT > ray_.min_t && closestT < ray_.max_t) { outT = float(closestT); return true; }
    return false;
}

struct Direct {
    const Scene& scene;
    BDPT svc;
    size_t m_emitterSamples, m_bsdfSamples; int m_strategy;
    explicit Direct(const Scene& s) : scene(s), svc(s, nullptr), m_emitterSamples(size_t(s.diEmitterSamples)), m_bsdfSamples(size_t(s.diBsdfSamples)),
        m_strategy(s.diStrategy) {}
    static float balanceHeuristic(float nf, float fPdf, float ng, float gPdf) { float f = nf * fPdf, g = ng * gPdf; return f / (f + g); } /* :81-84 */
    static const float inf() { return std::numeric_limits<float>::infinity(); }

    void sampleSphereByArea(float sx, float sy, v3 pShading, v3 emitterCenter, float emitterRadius, v3& pos, v3& ne, v3& wiW, float& pdf) const { /* :96-109 */
        ne = squareToUniformSphere(sx, sy);
        pos = ne * emitterRadius + emitterCenter;
        wiW = normalize(pos - pShading);
        pdf = 1.f / (4 * kPi * emitterRadius * emitterRadius);
    }
    static float coneCosThetaMax(v3 pShading, v3 emitterCenter, float emitterRadius) { /* :123-124, :427-428 */
        float sinThetaMaxSquared = emitterRadius * emitterRadius / length2(emitterCenter - pShading);
        return sqrtf(std::fmax(0.f, 1.f - sinThetaMaxSquared));
    }
    void sampleSphereBySolidAngle(float sx, float sy, v3 pShading, v3 emitterCenter, float emitterRadius, v3& wiW, float& pdf) const { /* :111-141 */
        v3 centerDir = normalize(emitterCenter - pShading);
        Frame fr(centerDir);
        float cosThetaMax = coneCosThetaMax(pShading, emitterCenter, emitterRadius);
        float cosTheta = (1.f - sx) + (sx * cosThetaMax);
        float phi = sy * kPi * 2.0f;
        float sinTheta = sqrtf(float(std::fmax(1.f - (cosTheta * cosTheta), 0)));
        v3 sampleDir = V(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
        wiW = fr.toWorld(sampleDir);
        pdf = kInvTwoPi * (1.f / (1.f - cosThetaMax));
    }

    v3 renderArea(SurfaceInteraction& hit, Sampler& sampler) const { /* :143-197 */
        v3 Lr = V(0.f);
        for (unsigned int i = 0; i < m_emitterSamples; ++i) {
            float emitterPdf;
            size_t id = svc.selectEmitter(sampler.next(), emitterPdf);
            const Emitter& emitter = scene.emitters[id];
            v3 emitterCenter = scene.getShapeCenter(emitter.shapeID);
            float emitterRadius = scene.getShapeRadius(emitter.shapeID);
            float sx, sy; sampler.next2D(&sx, &sy);
            v3 outPos, outNe, outwiW; float outPdf;
            sampleSphereByArea(sx, sy, hit.p, emitterCenter, emitterRadius, outPos, outNe, outwiW, outPdf);
            float lightDistanceSquared = length2(hit.p - outPos);
            float cosOutgoing = dot(-outwiW, outNe);
            v3 wiLocal = hit.frameNs.toLocal(outwiW);
            if (cosOutgoing <= 0.f || wiLocal.z <= 0.f) continue;
            Ray r(hit.p, outwiW, kEps, std::sqrt(lightDistanceSquared) - kEps);
            SurfaceInteraction vis;
            if (!scene.intersect(r, vis)) {
                hit.wi = wiLocal;
                float areaToSolidAngle = cosOutgoing * (1.f / lightDistanceSquared);
                Lr = Lr + emitter.radiance * svc.getBSDF(hit)->eval(hit) * (1.f / outPdf) * (1.f / emitterPdf) * areaToSolidAngle;
            }
        }
        return Lr / float(m_emitterSamples);
    }
    v3 renderCosineHemisphere(SurfaceInteraction& hit, Sampler& sampler) const { /* :200-234 */
        v3 Lr = V(0.f);
        for (unsigned int i = 0; i < m_emitterSamples; ++i) {
            float sx, sy; sampler.next2D(&sx, &sy);
            v3 localRayDir = squareToCosineHemisphere(sx, sy);
            v3 worldRayDir = normalize(hit.frameNs.toWorld(localRayDir));
            SurfaceInteraction vis;
            if (scene.intersect(Ray(hit.p, worldRayDir, kEps, inf()), vis)) {
                v3 emittedRadiance = svc.getEmission(vis);
                hit.wi = localRayDir;
                Lr = Lr + emittedRadiance * svc.getBSDF(hit)->eval(hit) * (1.0f / squareToCosineHemispherePdf(localRayDir));
            }
        }
        return Lr / float(m_emitterSamples);
    }
    v3 renderBSDF(SurfaceInteraction& hit, Sampler& sampler) const { /* :236-268 */
        v3 Lr = V(0.f);
        for (unsigned int i = 0; i < m_bsdfSamples; ++i) {
            float sx, sy; sampler.next2D(&sx, &sy);
            float pdf;
            v3 brdfCosTheta = svc.getBSDF(hit)->sample(hit, sx, sy, &pdf);
            SurfaceInteraction vis;
            if (scene.intersect(Ray(hit.p, hit.frameNs.toWorld(hit.wi), kEps, inf()), vis)) {
                v3 emittedRadiance = svc.getEmission(vis);
                Lr = Lr + emittedRadiance * brdfCosTheta * float(1.0 / pdf);
            }
        }
        return Lr / float(m_bsdfSamples);
    }
    /* the emitter loop shared by renderSolidAngle (:284-331, weighted = false, ray up to the centre distance + Epsilon) and
     * renderMIS (:357-403, weighted = true, unbounded ray) */
    v3 solidAngleLoop(SurfaceInteraction& hit, Sampler& sampler, bool weighted) const {
        v3 Lr = V(0.f);
        for (unsigned int i = 0; i < m_emitterSamples; ++i) {
            float emitterPdf;
            size_t id = svc.selectEmitter(sampler.next(), emitterPdf);
            const Emitter& emitter = scene.emitters[id];
            v3 emitterCenter = scene.getShapeCenter(emitter.shapeID);
            float emitterRadius = scene.getShapeRadius(emitter.shapeID);
            float sx, sy; sampler.next2D(&sx, &sy);
            v3 outwiW; float outPdf;
            sampleSphereBySolidAngle(sx, sy, hit.p, emitterCenter, emitterRadius, outwiW, outPdf);
            v3 wiLocal = hit.frameNs.toLocal(outwiW);
            if (wiLocal.z <= 0.f) continue;
            Ray r(hit.p, outwiW, kEps, weighted ? inf() : length(emitterCenter - hit.p) + kEps);
            SurfaceInteraction vis;
            bool contributes;
            if (scene.intersect(r, vis)) contributes = emitter.shapeID == vis.shapeID;
            else { float outT; contributes = raySphereIntersect(r, emitterCenter, emitterRadius, outT); }
            if (contributes) {
                hit.wi = wiLocal;
                if (weighted) {
                    float bsdfPdf = svc.getBSDF(hit)->pdf(hit);
                    float weight = balanceHeuristic(float(m_emitterSamples), outPdf * emitterPdf, float(m_bsdfSamples), bsdfPdf);
                    Lr = Lr + emitter.radiance * svc.getBSDF(hit)->eval(hit) * weight * (1.f / outPdf) * (1.f / emitterPdf);
                } else {
                    Lr = Lr + emitter.radiance * svc.getBSDF(hit)->eval(hit) * (1.f / outPdf) * (1.f / emitterPdf);
                }
            }
        }
        return Lr;
    }
    v3 renderSolidAngle(SurfaceInteraction& hit, Sampler& sampler) const { return solidAngleLoop(hit, sampler, false) / float(m_emitterSamples); }
    v3 renderMIS(SurfaceInteraction& hit, Sampler& sampler) const { /* :340-447 */
        v3 emitterEstimator = solidAngleLoop(hit, sampler, true), bsdfEstimator = V(0.f);
        emitterEstimator = m_emitterSamples == 0 ? V(0.f) : (emitterEstimator / float(m_emitterSamples));
        for (unsigned int i = 0; i < m_bsdfSamples; ++i) {
            float sx, sy; sampler.next2D(&sx, &sy);
            float outPdf;
            v3 brdfCosTheta = svc.getBSDF(hit)->sample(hit, sx, sy, &outPdf);
            SurfaceInteraction vis;
            if (scene.intersect(Ray(hit.p, hit.frameNs.toWorld(hit.wi), kEps, inf()), vis)) {
                v3 emittedRadiance = svc.getEmission(vis);
                if (emittedRadiance == V(0.f)) continue;
                const Emitter& emitter = scene.emitters[svc.getEmitterIDByShapeID(vis.shapeID)];
                float cosThetaMax = coneCosThetaMax(hit.p, scene.getShapeCenter(emitter.shapeID), scene.getShapeRadius(emitter.shapeID));
                float emitterSolidAnglePdf = kInvTwoPi * (1.f / (1.f - cosThetaMax));
                emitterSolidAnglePdf *= 1.f / scene.emitters.size(); /* getEmitterPdf */
                float weight = balanceHeuristic(float(m_bsdfSamples), outPdf, float(m_emitterSamples), emitterSolidAnglePdf);
                bsdfEstimator = bsdfEstimator + emittedRadiance * brdfCosTheta * weight * (1.f / outPdf);
            }
        }
        bsdfEstimator = m_bsdfSamples == 0 ? V(0.f) : (bsdfEstimator / float(m_bsdfSamples));
        return V(0.f) + (emitterEstimator + bsdfEstimator);
    }
    v3 render(const Ray& ray, Sampler& sampler) const { /* :449-462; every strategy starts the same way (:147-154) */
        SurfaceInteraction hit;
        if (!scene.intersect(ray, hit)) return V(0.f);
        v3 emittedRadiance = svc.getEmission(hit);
        if (emittedRadiance != V(0.f)) return emittedRadiance;
        switch (m_strategy) {
            case 0: return renderMIS(hit, sampler);
            case 1: return renderArea(hit, sampler);
            case 2: return renderSolidAngle(hit, sampler);
            case 3: return renderCosineHemisphere(hit, sampler);
            default: return renderBSDF(hit, sampler);
        }
    }
};

}  // namespace

/* ---- C interface (ctypes: oracle/oracle_port.py) ------------------------------------------------------------ */
extern "C" {

void* orc_create(const float* positions, size_t npos, const float* normals, size_t nnrm, const int* tri_v, const int* tri_n,
                 const int* tri_mat, const int* tri_shape, size_t ntris, const float* mats16, size_t nmat,
                 const float* texcoords, size_t ntexcoords, const int* tri_t, const float* const* tex_rgb, const int* tex_wh, size_t ntex) {
    Scene* s = new Scene();
    for (size_t i = 0; i < npos; i++) s->vertices.push_back(V(positions[3 * i], positions[3 * i + 1], positions[3 * i + 2]));
    for (size_t i = 0; i < nnrm; i++) s->normals.push_back(V(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]));
    s->tri_v.assign(tri_v, tri_v + 3 * ntris); s->tri_n.assign(tri_n, tri_n + 3 * ntris);
    s->tri_mat.assign(tri_mat, tri_mat + ntris); s->tri_shape.assign(tri_shape, tri_shape + ntris);
    for (size_t i = 0; i < nmat; i++) {
        const float* m = mats16 + 16 * i;
        s->materials.push_back(Material{V(m[0], m[1], m[2]), V(m[3], m[4], m[5]), V(m[6], m[7], m[8]), V(m[9], m[10], m[11]), m[12], m[13], int(m[14]), ntex ? int(m[15]) : 0});
    }
    if (ntex) {
        s->tex.texcoords.assign(texcoords, texcoords + 2 * ntexcoords);
        s->tex.tri_t.assign(tri_t, tri_t + 3 * ntris);
        for (size_t t = 0; t < ntex; t++) {
            TexSet::Tex x; x.w = tex_wh[2 * t]; x.h = tex_wh[2 * t + 1];
            x.cs.assign(tex_rgb[t], tex_rgb[t] + size_t(x.w) * x.h * 3);
            s->tex.textures.push_back(std::move(x));
        }
    }
    s->finish();
    return s;
}
void orc_destroy(void* h) { delete static_cast<Scene*>(h); }

void orc_set_config(void* h, const float* eye, const float* at, const float* up, float fov, int width, int height, int spp, int rrDepth) {
    Scene* s = static_cast<Scene*>(h);
    s->cam_o = V(eye[0], eye[1], eye[2]); s->cam_at = V(at[0], at[1], at[2]); s->cam_up = V(up[0], up[1], up[2]);
    s->fov = fov; s->width = width; s->height = height; s->spp = spp; s->rrDepth = rrDepth;
}

void orc_set_rr(void* h, int rr_on, int max_depth) {
    Scene* s = static_cast<Scene*>(h);
    s->rrOn = rr_on != 0; s->rrMaxDepth = max_depth;
}

void orc_set_path(void* h, int is_explicit, int max_depth, float rr_prob, int emitter_samples, int bsdf_samples, int depth_cap) {
    Scene* s = static_cast<Scene*>(h);
    s->ptExplicit = is_explicit != 0; s->ptMaxDepth = max_depth; s->ptRRProb = rr_prob; s->ptEmitterSamples = emitter_samples;
    s->ptBsdfSamples = bsdf_samples; s->ptDepthCap = depth_cap;
}

/* PathTracerIntegrator::render on given camera rays; random numbers from Sampler(seed) or injected rows */
void orc_path_paths(void* h, const float* rays8, const int* seeds, const float* injected, int64_t stride, int64_t n, float* Li, int* nDraws) {
    Scene* s = static_cast<Scene*>(h);
    PathTracer pt(*s);
    for (int64_t i = 0; i < n; i++) {
        Sampler sm(seeds ? seeds[i] : 0);
        if (!seeds) sm.injected = injected + i * stride;
        const float* r = rays8 + 8 * i;
        v3 L = pt.render(Ray(V(r[0], r[1], r[2]), V(r[3], r[4], r[5]), r[6], r[7]), sm);
        Li[3 * i] = L.x; Li[3 * i + 1] = L.y; Li[3 * i + 2] = L.z;
        if (nDraws) nDraws[i] = int(sm.pos);
    }
}

void orc_set_direct(void* h, int strategy, int emitter_samples, int bsdf_samples) {
    Scene* s = static_cast<Scene*>(h);
    s->diStrategy = strategy; s->diEmitterSamples = emitter_samples; s->diBsdfSamples = bsdf_samples;
}

/* DirectIntegrator::render on given camera rays; random numbers from Sampler(seed) or injected rows */
void orc_direct_paths(void* h, const float* rays8, const int* seeds, const float* injected, int64_t stride, int64_t n, float* Li, int* nDraws) {
    Scene* s = static_cast<Scene*>(h);
    Direct di(*s);
    for (int64_t i = 0; i < n; i++) {
        Sampler sm(seeds ? seeds[i] : 0);
        if (!seeds) sm.injected = injected + i * stride;
        const float* r = rays8 + 8 * i;
        v3 L = di.render(Ray(V(r[0], r[1], r[2]), V(r[3], r[4], r[5]), r[6], r[7]), sm);
        Li[3 * i] = L.x; Li[3 * i + 1] = L.y; Li[3 * i + 2] = L.z;
        if (nDraws) nDraws[i] = int(sm.pos);
    }
}

/* per shape: centre xyz + radius (Scene::getShapeCenter / getShapeRadius) */
void orc_shape_spheres(void* h, float* out4) {
    Scene* s = static_cast<Scene*>(h);
    for (size_t i = 0; i < s->shapesCenter.size(); i++) {
        out4[4 * i] = s->shapesCenter[i].x; out4[4 * i + 1] = s->shapesCenter[i].y; out4[4 * i + 2] = s->shapesCenter[i].z;
        out4[4 * i + 3] = s->getShapeRadius(i);
    }
}

void orc_bvh_order(void* h, int* order) { Scene* s = static_cast<Scene*>(h); std::copy(s->prims.begin(), s->prims.end(), order); }
int orc_bvh_node_count(void* h) { return int(static_cast<Scene*>(h)->flatTree.size()); }

int orc_emitters(void* h, int* shape, float* area, float* radiance, float* cdf_concat, int max_e) {
    Scene* s = static_cast<Scene*>(h);
    size_t off = 0;
    for (size_t i = 0; i < s->emitters.size() && int(i) < max_e; i++) {
        shape[i] = int(s->emitters[i].shapeID); area[i] = s->emitters[i].area;
        radiance[3 * i] = s->emitters[i].radiance.x; radiance[3 * i + 1] = s->emitters[i].radiance.y; radiance[3 * i + 2] = s->emitters[i].radiance.z;
        for (float c : s->emitters[i].cdf) cdf_concat[off++] = c;
    }
    return int(s->emitters.size());
}

void orc_camera_matrices(void* h, float* out64) {
    Scene* s = static_cast<Scene*>(h);
    M4 w2c = lookAt(s->cam_o, s->cam_at, s->cam_up), c2w = inverse(w2c);
    M4 clip = perspective(kPi / 180.f * s->fov, (float)s->width / s->height, 1.f, 1000.f);
    M4 n2s = mulm(mulm(scaleM(float(s->width), float(s->height), 1.0f), scaleM(0.5f, -0.5f, 1.f)), translateM(1.f, -1.f, 0.f));
    std::memcpy(out64, w2c.m, 64); std::memcpy(out64 + 16, c2w.m, 64); std::memcpy(out64 + 32, clip.m, 64); std::memcpy(out64 + 48, n2s.m, 64);
}

static Ray mkRay(const float* r) { return Ray(V(r[0], r[1], r[2]), V(r[3], r[4], r[5]), r[6], r[7]); }

/* ids: n x 3 (hit, shapeID, primID); tuv: n x 3 */
void orc_intersect(void* h, const float* rays8, int64_t n, int* ids, float* tuv) {
    Scene* s = static_cast<Scene*>(h);
    for (int64_t i = 0; i < n; i++) {
        SurfaceInteraction si;
        bool hit = s->intersect(mkRay(rays8 + 8 * i), si);
        ids[3 * i] = hit; ids[3 * i + 1] = hit ? int(si.shapeID) : -1; ids[3 * i + 2] = hit ? int(si.primID) : -1;
        tuv[3 * i] = si.t; tuv[3 * i + 1] = hit ? si.u : 0.f; tuv[3 * i + 2] = hit ? si.v : 0.f;
    }
}
void orc_occluded(void* h, const float* rays8, int64_t n, uint8_t* out) {
    Scene* s = static_cast<Scene*>(h);
    for (int64_t i = 0; i < n; i++) { IntersectionInfo info; out[i] = s->getIntersection(mkRay(rays8 + 8 * i), &info, true) ? 1 : 0; }
}
/* q: n x 8 (wo, wi, u); out: n x 12 (eval3, pdf, sample3, sample_pdf, sample_wi3, type) */
void orc_bsdf(void* h, int mat, const float* q, int64_t n, float* out) {
    Scene* s = static_cast<Scene*>(h);
    const BSDF& b = s->bsdfs[mat];
    for (int64_t i = 0; i < n; i++) {
        const float* x = q + 8 * i; float* o = out + 12 * i;
        SurfaceInteraction si; si.wo = V(x[0], x[1], x[2]); si.wi = V(x[3], x[4], x[5]);
        v3 e = b.eval(si); o[0] = e.x; o[1] = e.y; o[2] = e.z; o[3] = b.pdf(si);
        SurfaceInteraction ss; ss.wo = si.wo;
        float pdf = 0.f; v3 sv = b.sample(ss, x[6], x[7], &pdf);
        o[4] = sv.x; o[5] = sv.y; o[6] = sv.z; o[7] = pdf; o[8] = ss.wi.x; o[9] = ss.wi.y; o[10] = ss.wi.z; o[11] = float(b.combinedType);
    }
}
void orc_sampler_stream(int seed, int64_t n, float* out) { Sampler s(seed); for (int64_t i = 0; i < n; i++) out[i] = s.next(); }

/* BDPTIntegrator::render per path. seeds != NULL: private Sampler(seed); else injected numbers (n x stride).
 * Splats are returned by scanning the cleared frame buffer exactly like oracle/ref_harness.cpp does. */
void orc_bdpt_paths(void* h, const float* rays8, const int* seeds, const float* injected, int64_t stride, int64_t n, int maxSplats,
                    float* Li, int* splat_pix, float* splat_rgb, int* nDraws) {
    Scene* s = static_cast<Scene*>(h);
    std::vector<v3> fb(size_t(s->width) * s->height, V(0.f));
    BDPT bdpt(*s, &fb);
    for (int64_t i = 0; i < n; i++) {
        std::fill(fb.begin(), fb.end(), V(0.f));
        Sampler sm(seeds ? seeds[i] : 0);
        if (!seeds) sm.injected = injected + i * stride;
        v3 L = bdpt.render(mkRay(rays8 + 8 * i), sm);
        Li[3 * i] = L.x; Li[3 * i + 1] = L.y; Li[3 * i + 2] = L.z;
        int k = 0;
        for (int m = 0; m < maxSplats; m++) splat_pix[i * maxSplats + m] = -1;
        for (size_t p = 0; p < fb.size(); p++) {
            if (fb[p].x != 0.f || fb[p].y != 0.f || fb[p].z != 0.f) {
                if (k < maxSplats) {
                    splat_pix[i * maxSplats + k] = int(p);
                    splat_rgb[(i * maxSplats + k) * 3] = fb[p].x; splat_rgb[(i * maxSplats + k) * 3 + 1] = fb[p].y; splat_rgb[(i * maxSplats + k) * 3 + 2] = fb[p].z;
                }
                k++;
            }
        }
        if (nDraws) nDraws[i] = int(sm.pos);
    }
}

/* Renderer::render offline branch (renderer.cpp:130-214) over `threads` static pixel chunks (parallelfor.h:25-66).
 * Unlike the reference, every chunk owns its Sampler (seed + chunk) and its splat buffer, so the run is race-free
 * and deterministic; the buffers are summed at the end.  Returns seconds spent rendering. */
double orc_render(void* h, float* rgb_out, int threads, int seed) {
    Scene* s = static_cast<Scene*>(h);
    const int W = s->width, H = s->height, np = W * H;
    threads = std::max(1, threads);
    M4 worldToCamera = lookAt(s->cam_o, s->cam_at, s->cam_up), cameraToWorld = inverse(worldToCamera);
    const float invWidth = 1.f / W, invHeight = 1.f / H;
    const float angle = std::tan(kPi / 180.f * s->fov * 0.5f), aspectRatio = (float)W / H;
    std::vector<std::vector<v3>> fbs(threads, std::vector<v3>(size_t(np), V(0.f)));
    auto t0 = std::chrono::high_resolution_clock::now();
    auto work = [&](int tid) {
        BDPT bdpt(*s, &fbs[tid]);
        Sampler sampler(seed + tid);
        int start = int(int64_t(np) * tid / threads), end = int(int64_t(np) * (tid + 1) / threads);
        for (int pixel = start; pixel < end; ++pixel) {
            int j = pixel % W, i = pixel / W;
            float y = (1.f - ((float)i + 0.5f) * invHeight) * 2.f - 1.f;
            float x = (((float)j + 0.5f) * invWidth) * 2.f - 1.f;
            v3 averageRadiance = V(0.f);
            for (int sidx = 0; sidx < s->spp; sidx++) {
                float px, py;
                if (s->spp == 1) { px = x * angle * aspectRatio; py = y * angle; }
                else {
                    float rx, ry; sampler.next2D(&rx, &ry);
                    rx -= 0.5f; ry -= 0.5f; rx = rx * invWidth; ry = ry * invHeight;
                    px = (x + rx) * angle * aspectRatio; py = (y + ry) * angle;
                }
                float ip[4] = {px, py, -1.f, 0.f}, d4[4];
                mulv(cameraToWorld, ip, d4);
                Ray ray(s->cam_o, normalize(V(d4[0], d4[1], d4[2])), 1.f, 1000.f);
                averageRadiance = averageRadiance + bdpt.render(ray, sampler);
            }
            fbs[tid][pixel] = fbs[tid][pixel] + averageRadiance * (1.f / s->spp);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work, t);
    work(0);
    for (auto& t : pool) t.join();
    auto t1 = std::chrono::high_resolution_clock::now();
    if (rgb_out) {
        for (int p = 0; p < np; p++) {
            v3 a = V(0.f);
            for (int t = 0; t < threads; t++) a = a + fbs[t][p];
            rgb_out[3 * p] = a.x; rgb_out[3 * p + 1] = a.y; rgb_out[3 * p + 2] = a.z;
        }
    }
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
