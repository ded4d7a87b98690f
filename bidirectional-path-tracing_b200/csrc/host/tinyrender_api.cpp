// tinyrender_api.cpp -- TinyRender's C++ host API (include/tinyrender/tinyrender.h) on top of the C ABI.
// Host logic only: every computation goes through trb_* (include/trb.h) to the GPU.
#include "../../../include/tinyrender/tinyrender.h"

#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <mutex>
#include <thread>

#include "tr_host.h"

namespace TinyRender {

namespace {
trb_render_desc make_desc(const Config& c) {
    trb_render_desc d;
    std::memset(&d, 0, sizeof(d));
    d.eye[0] = c.camera.o.x; d.eye[1] = c.camera.o.y; d.eye[2] = c.camera.o.z;
    d.at[0] = c.camera.at.x; d.at[1] = c.camera.at.y; d.at[2] = c.camera.at.z;
    d.up[0] = c.camera.up.x; d.up[1] = c.camera.up.y; d.up[2] = c.camera.up.z;
    d.fov_deg = c.camera.fov;
    d.width = c.width; d.height = c.height; d.spp = c.spp;
    d.rr_depth = c.integratorSettings.bdpt.rrDepth;
    d.seed = c.seed; d.wave_samples = c.waveSamples;
    if (c.russianRoulette) d.flags |= TRB_FLAG_RUSSIAN_ROULETTE;
    d.max_depth = c.maxDepth;
    return d;
}
trb_ray to_ray(const Ray& r) {
    trb_ray o;
    o.o[0] = r.o.x; o.o[1] = r.o.y; o.o[2] = r.o.z; o.d[0] = r.d.x; o.d[1] = r.d.y; o.d[2] = r.d.z;
    o.min_t = r.min_t; o.max_t = r.max_t;
    return o;
}
}  // namespace

// ---- loadTOML (main.cpp:22-116) ---------------------------------------------------------------------------
bool loadTOML(Config& config, const std::string& inputFile) {
    trb::TomlConfig t;
    std::string err;
    if (!trb::load_toml(inputFile, t, err)) throw std::runtime_error(err);
    config.tomlFile = inputFile;
    config.objFile = t.obj_path;
    config.camera.o = v3f(t.desc.eye[0], t.desc.eye[1], t.desc.eye[2]);
    config.camera.at = v3f(t.desc.at[0], t.desc.at[1], t.desc.at[2]);
    config.camera.up = v3f(t.desc.up[0], t.desc.up[1], t.desc.up[2]);
    config.camera.fov = t.desc.fov_deg;
    config.width = t.desc.width; config.height = t.desc.height; config.spp = t.desc.spp;
    config.seed = t.desc.seed; config.waveSamples = t.desc.wave_samples; config.gpus = t.gpus;
    config.russianRoulette = (t.desc.flags & TRB_FLAG_RUSSIAN_ROULETTE) != 0; config.maxDepth = t.desc.max_depth;
    config.checkpointFile = t.checkpoint; config.previewPng = t.preview; config.checkpointEvery = t.checkpoint_every;
    static const char* known[] = {"normal", "simple", "ao", "ro", "direct", "path", "bdpt"};
    static const EIntegrator ids[] = {ENormalIntegrator, ESimpleIntegrator, EAOIntegrator, EROIntegrator, EDirectIntegrator,
                                      EPathTracerIntegrator, EBDPTIntegrator};
    bool found = false;
    for (int k = 0; k < 7; ++k) if (t.type == known[k]) { config.integrator = ids[k]; found = true; }
    if (!found && !t.realtime) throw std::runtime_error("Invalid integrator type");  // main.cpp:109
    config.integratorSettings.bdpt.rrDepth = t.desc.rr_depth;
    auto& pt = config.integratorSettings.pt;
    pt.isExplicit = t.path.is_explicit != 0; pt.maxDepth = t.path.max_depth; pt.rrDepth = t.desc.rr_depth; pt.rrProb = t.path.rr_prob;
    pt.emitterSamples = size_t(std::max(t.path.emitter_samples, 0)); pt.bsdfSamples = size_t(std::max(t.path.bsdf_samples, 0));
    auto& di = config.integratorSettings.di;  // main.cpp:88-93
    di.emitterSamples = size_t(std::max(t.direct.emitter_samples, 0)); di.bsdfSamples = size_t(std::max(t.direct.bsdf_samples, 0));
    static const char* strategies[] = {"mis", "area", "solidAngle", "cosineHemisphere", "bsdf"};
    di.samplingStrategy = (t.direct.strategy >= 0 && t.direct.strategy <= 4) ? strategies[t.direct.strategy] : "emitter";
    return t.realtime;
}

// ---- Scene (renderer.cpp:235-315) -------------------------------------------------------------------------
Scene::Scene(const Config& c) : config(c) {}

Scene::~Scene() {
    for (trb_context* c : ctx) if (c) trb_destroy(c);
    if (host) trb_host_scene_free(host);
}

const trb_scene_desc* Scene::desc() const { return host ? &sdesc : nullptr; }

trb_context* Scene::context(int device) const {
    if (device < 0) return nullptr;
    if (size_t(device) >= ctx.size()) ctx.resize(size_t(device) + 1, nullptr);
    if (!ctx[device]) {
        trb_context* c = nullptr;
        if (trb_create(device, &c) != TRB_OK) { err = trb_last_error(nullptr); return nullptr; }
        if (host && trb_scene_upload(c, &sdesc) != TRB_OK) { err = trb_last_error(c); trb_destroy(c); return nullptr; }
        ctx[device] = c;
    }
    return ctx[device];
}

bool Scene::load(bool isRealTime) {
    if (isRealTime) { err = "the realtime (OpenGL) passes are not available: no graphics stack on this platform"; std::printf("%s\n", err.c_str()); return false; }
    if (trb_host_load_obj(config.objFile.c_str(), &host) != TRB_OK) { err = trb_host_last_error(); std::printf("%s\n", err.c_str()); return false; }
    trb_host_scene_desc(host, &sdesc);
    nShapes = sdesc.n_shapes; nTriangles = sdesc.n_tris;
    // BSDF class by illum (renderer.cpp:259-271)
    bsdfs.clear();
    for (size_t i = 0; i < sdesc.n_materials; ++i) {
        const trb_material& m = sdesc.materials[i];
        std::unique_ptr<BSDF> b;
        switch (m.illum) {
            case 7: b.reset(new DiffuseBSDF(*this, i)); b->combinedType = BSDF::EDiffuseReflection; break;
            case 3: b.reset(new MirrorBSDF(*this, i)); b->combinedType = BSDF::EDeltaReflection; break;
            case 6: b.reset(new GlassBSDF(*this, i)); b->combinedType = BSDF::EDeltaReflection | BSDF::EDeltaTransmission; break;
            case 8: b.reset(new MixtureBSDF(*this, i)); b->combinedType = BSDF::EGlossyReflection | BSDF::EDiffuseReflection; break;
            case 5: err = "illum 5: the reference builds no BSDF for it (renderer.cpp:259-270)"; std::printf("%s\n", err.c_str()); return false;
            default: b.reset(new PhongBSDF(*this, i)); b->combinedType = BSDF::EGlossyReflection | BSDF::EDiffuseReflection; break;
        }
        b->emissive = m.ke[0] != 0.f || m.ke[1] != 0.f || m.ke[2] != 0.f;
        bsdfs.push_back(std::move(b));
    }
    // emitters (renderer.cpp:279-305)
    {
        std::vector<int32_t> shape(64);
        std::vector<float> area(64), rad(64 * 3), cdf(sdesc.n_tris + 128);
        int n = trb_host_emitters(&sdesc, shape.data(), area.data(), rad.data(), 64, cdf.data(), cdf.size());
        if (n < 0) { err = trb_host_last_error(); std::printf("%s\n", err.c_str()); return false; }
        emitters.clear();
        size_t off = 0;
        std::vector<size_t> faces(sdesc.n_shapes, 0);
        for (size_t t = 0; t < sdesc.n_tris; ++t) ++faces[sdesc.tri_shape[t]];
        for (int e = 0; e < n; ++e) {
            Emitter em;
            em.shapeID = size_t(shape[e]); em.area = area[e]; em.radiance = v3f(rad[3 * e], rad[3 * e + 1], rad[3 * e + 2]);
            size_t nf = faces[em.shapeID] + 1;
            em.faceAreaCdf.assign(cdf.begin() + off, cdf.begin() + off + nf);
            off += nf;
            emitters.push_back(std::move(em));
        }
    }
    for (int k = 0; k < 3; ++k) { aabb.min[k] = 3.4e38f; aabb.max[k] = -3.4e38f; }
    for (size_t v = 0; v < sdesc.n_positions; ++v)
        for (int k = 0; k < 3; ++k) { aabb.min[k] = std::fmin(aabb.min[k], sdesc.positions[3 * v + k]); aabb.max[k] = std::fmax(aabb.max[k], sdesc.positions[3 * v + k]); }
    {   // the WorldData view: per-shape centre and radius as Scene::load derives them (renderer.cpp:294-304)
        std::vector<float> sph(4 * sdesc.n_shapes);
        if (trb_host_shape_spheres(&sdesc, sph.data(), sdesc.n_shapes) != TRB_OK) { err = trb_host_last_error(); std::printf("%s\n", err.c_str()); return false; }
        worldData.owner = this; worldData.flat = &sdesc;
        worldData.shapesCenter.clear(); worldData.shapesRadius.clear();
        for (size_t k = 0; k < sdesc.n_shapes; ++k) { worldData.shapesCenter.push_back(v3f(sph[4 * k], sph[4 * k + 1], sph[4 * k + 2])); worldData.shapesRadius.push_back(sph[4 * k + 3]); }
        for (auto& b : bsdfs) b->emission = v3f(sdesc.materials[b->matID].ke[0], sdesc.materials[b->matID].ke[1], sdesc.materials[b->matID].ke[2]);
    }
    std::printf("Found %zu shapes, %zu triangles, %zu emitters\n", nShapes, nTriangles, emitters.size());
    // AcceleratorBVH::build (accel.h:115-123) + device upload: fails loudly without a GPU
    if (!context(0)) { std::printf("%s\n", err.c_str()); return false; }
    bvh.reset(new AcceleratorBVH(*this));
    return true;
}

// ---- AcceleratorBVH (accel.h:113-172) -----------------------------------------------------------------------
bool AcceleratorBVH::build() { return scene.context(0) != nullptr; }

bool AcceleratorBVH::intersectBatch(const trb_ray* rays, size_t n, trb_hit* out) const {
    trb_context* c = scene.context(0);
    return c && trb_trace_closest(c, rays, n, out) == TRB_OK;
}
bool AcceleratorBVH::occludedBatch(const trb_ray* rays, size_t n, uint8_t* out) const {
    trb_context* c = scene.context(0);
    return c && trb_trace_any(c, rays, n, out) == TRB_OK;
}
bool AcceleratorBVH::occluded(const Ray& ray) const {
    trb_ray r = to_ray(ray);
    uint8_t o = 0;
    return occludedBatch(&r, 1, &o) && o != 0;
}

namespace {
// host copies of the few fp32 expressions that fill SurfaceInteraction after a hit (accel.h:136-166, math.h:19-22,42-51)
struct F3 { float x, y, z; };
inline F3 bary3(const float* a, const float* b, const float* c, float u, float v) {
    float w = 1 - u - v;
    return {(a[0] * w + b[0] * u) + c[0] * v, (a[1] * w + b[1] * u) + c[1] * v, (a[2] * w + b[2] * u) + c[2] * v};
}
inline float dot3(F3 a, F3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline F3 cross3(F3 x, F3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
}  // namespace

bool AcceleratorBVH::intersect(const Ray& ray, SurfaceInteraction& info, bool) const {
    trb_ray r = to_ray(ray);
    trb_hit h;
    if (!intersectBatch(&r, 1, &h)) return false;
    info.t = h.t;
    if (!h.hit) return false;
    const trb_scene_desc& s = *scene.desc();
    const int32_t* tv = s.tri_v + 3 * size_t(h.tri);
    const int32_t* tn = s.tri_n + 3 * size_t(h.tri);
    F3 p = bary3(s.positions + 3 * tv[0], s.positions + 3 * tv[1], s.positions + 3 * tv[2], h.u, h.v);
    F3 n = bary3(s.normals + 3 * tn[0], s.normals + 3 * tn[1], s.normals + 3 * tn[2], h.u, h.v);
    float inv = 1.f / std::sqrt(dot3(n, n));
    n = {n.x * inv, n.y * inv, n.z * inv};
    F3 t;  // coordinateSystem (math.h:42-51)
    if (std::fabs(n.x) > std::fabs(n.y)) { float il = 1.f / std::sqrt(n.x * n.x + n.z * n.z); t = {n.z * il, 0.f, -n.x * il}; }
    else { float il = 1.f / std::sqrt(n.y * n.y + n.z * n.z); t = {0.f, n.z * il, -n.y * il}; }
    F3 sdir = cross3(t, n);
    F3 md = {-ray.d.x, -ray.d.y, -ray.d.z};
    info.p = v3f(p.x, p.y, p.z);
    info.wo = v3f(dot3(md, sdir), dot3(md, t), dot3(md, n));
    info.u = h.u; info.v = h.v;
    info.shapeID = size_t(h.shape); info.primID = size_t(h.prim);
    info.matID = s.tri_mat[h.tri];
    return true;
}

// ---- BSDF (core.h:256-318) ------------------------------------------------------------------------------------
// the reference's constructor (core.h:299-305): emission = materials[matID].emission
BSDF::BSDF(const WorldData& d, const Config& c, size_t id) : worldData(d), config(c), scene(*d.owner), matID(id) {
    if (!d.owner || !d.flat) throw std::runtime_error("BSDF: the WorldData view is not bound to a loaded Scene");
    if (id >= d.flat->n_materials) throw std::out_of_range("BSDF: material index out of range");
    const trb_material& m = d.flat->materials[id];
    emission = v3f(m.ke[0], m.ke[1], m.ke[2]);
    emissive = m.ke[0] != 0.f || m.ke[1] != 0.f || m.ke[2] != 0.f;
}
BSDF::BSDF(const Scene& s, size_t id) : worldData(s.worldData), config(s.config), scene(s), matID(id) {}

float Scene::getShapeRadius(size_t shapeID) const { return worldData.shapesRadius.at(shapeID); }
v3f Scene::getShapeCenter(size_t shapeID) const { return worldData.shapesCenter.at(shapeID); }
float Scene::getShapeArea(size_t shapeID) const {   // the running fp32 sum Distribution1D::add keeps (math.h:88-91), face order
    if (shapeID >= sdesc.n_shapes) throw std::out_of_range("Scene::getShapeArea: shape index out of range");
    float sum = 0.f;
    for (size_t t = 0; t < sdesc.n_tris; ++t) {
        if (size_t(sdesc.tri_shape[t]) != shapeID) continue;
        const float* a = sdesc.positions + 3 * size_t(sdesc.tri_v[3 * t]);
        const float* b = sdesc.positions + 3 * size_t(sdesc.tri_v[3 * t + 1]);
        const float* c = sdesc.positions + 3 * size_t(sdesc.tri_v[3 * t + 2]);
        const float e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        const float cx = e1[1] * e2[2] - e2[1] * e1[2], cy = e1[2] * e2[0] - e2[2] * e1[0], cz = e1[0] * e2[1] - e2[0] * e1[1];   // glm::cross order
        sum += 0.5f * std::sqrt(cx * cx + cy * cy + cz * cz);
    }
    return sum;
}

size_t Integrator::getEmitterIDByShapeID(size_t shapeID) const {
    for (size_t e = 0; e < scene.emitters.size(); ++e) if (scene.emitters[e].shapeID == shapeID) return e;
    throw std::out_of_range("Integrator::getEmitterIDByShapeID: the shape is not an emitter");
}
const BSDF* Integrator::getBSDF(const SurfaceInteraction& hit) const {   // material of face primID of shape shapeID
    const trb_scene_desc* d = scene.desc();
    size_t first = 0;
    while (first < d->n_tris && size_t(d->tri_shape[first]) < hit.shapeID) ++first;   // triangles are shape-major
    const size_t t = first + hit.primID;
    if (t >= d->n_tris || size_t(d->tri_shape[t]) != hit.shapeID) throw std::out_of_range("Integrator::getBSDF: no such face");
    return scene.bsdfs.at(size_t(d->tri_mat[t])).get();
}
v3f Integrator::getEmission(const SurfaceInteraction& hit) const {
    const trb_scene_desc* d = scene.desc();
    if (hit.matID < 0 || size_t(hit.matID) >= d->n_materials) throw std::out_of_range("Integrator::getEmission: material index out of range");
    const trb_material& m = d->materials[hit.matID];
    return v3f(m.ke[0], m.ke[1], m.ke[2]);
}
size_t Integrator::selectEmitter(float sample, float& pdf) const {
    size_t id = size_t(sample * scene.emitters.size());
    id = std::min(id, scene.emitters.size() - 1);
    pdf = 1.f / scene.emitters.size();
    return id;
}

bool BSDF::evalBatch(const trb_bsdf_query* q, size_t n, trb_bsdf_out* out) const {
    trb_context* c = scene.context(0);
    return c && trb_bsdf_batch(c, q, n, out) == TRB_OK;
}
namespace {
trb_bsdf_out bsdf_one(const BSDF& b, const SurfaceInteraction& i, const v2f& u) {
    trb_bsdf_query q;
    std::memset(&q, 0, sizeof(q));
    q.wo[0] = i.wo.x; q.wo[1] = i.wo.y; q.wo[2] = i.wo.z; q.wi[0] = i.wi.x; q.wi[1] = i.wi.y; q.wi[2] = i.wi.z;
    q.u[0] = u.x; q.u[1] = u.y; q.mat = int32_t(b.matID);
    trb_bsdf_out o;
    std::memset(&o, 0, sizeof(o));
    if (!b.evalBatch(&q, 1, &o)) throw std::runtime_error("BSDF: device evaluation failed (no CPU fallback)");
    return o;
}
}  // namespace
v3f BSDF::eval(const SurfaceInteraction& i) const { trb_bsdf_out o = bsdf_one(*this, i, v2f()); return v3f(o.eval[0], o.eval[1], o.eval[2]); }
float BSDF::pdf(const SurfaceInteraction& i) const { return bsdf_one(*this, i, v2f()).pdf; }
v3f BSDF::sample(SurfaceInteraction& i, const v2f& s, float* pdf) const {
    trb_bsdf_out o = bsdf_one(*this, i, s);
    i.wi = v3f(o.sample_wi[0], o.sample_wi[1], o.sample_wi[2]);
    if (pdf) *pdf = o.sample_pdf;
    return v3f(o.sample_val[0], o.sample_val[1], o.sample_val[2]);
}

// ---- Integrator (integrator.cpp:16-30) ----------------------------------------------------------------------
bool Integrator::init() {
    rgb.reset(new RenderBuffer(scene.config.width, scene.config.height));
    rgb->clear();
    return true;
}

bool Integrator::save() {
    std::string out = scene.config.tomlFile;
    size_t dot = out.find_last_of('.');
    size_t slash = out.find_last_of('/');
    if (dot != std::string::npos && (slash == std::string::npos || dot > slash)) out = out.substr(0, dot);
    out += ".exr";
    static_assert(sizeof(v3f) == 12, "RenderBuffer must be packed RGB floats");
    if (trb_host_save_exr(out.c_str(), reinterpret_cast<const float*>(rgb->data.get()), rgb->width, rgb->height) != TRB_OK) {
        std::printf("%s\n", trb_host_last_error());
        return false;
    }
    std::printf("Saved EXR image to %s\n", out.c_str());
    return true;
}

// ---- BDPTIntegrator (bdpt.h:37-518) ---------------------------------------------------------------------------
BDPTIntegrator::BDPTIntegrator(const Scene& s) : Integrator(s) {
    m_rrDepth = s.config.integratorSettings.bdpt.rrDepth;  // bdpt.h:38-43
    m_rrProb = s.config.integratorSettings.bdpt.rrProb;
}

v3f BDPTIntegrator::render(const Ray& ray, Sampler& sampler) const {
    trb_context* c = scene.context(0);
    if (!c) throw std::runtime_error("BDPTIntegrator::render: no device context");
    trb_render_desc d = make_desc(scene.config);
    const int K = std::max(m_rrDepth - 1, 1);
    const size_t stride = size_t(10 + 10 * K);
    // the path consumes a prefix of the sampler's stream: feed a copy's numbers, then advance the real sampler
    Sampler ahead = sampler;
    std::vector<float> rng(stride);
    for (float& f : rng) f = ahead.next();
    trb_ray r = to_ray(ray);
    float li[3] = {0, 0, 0};
    std::vector<int32_t> pix(size_t(K), -1);
    std::vector<float> srgb(size_t(K) * 3, 0.f);
    int32_t draws = 0;
    if (trb_bdpt_paths(c, &d, &r, rng.data(), 1, stride, li, pix.data(), srgb.data(), &draws) != TRB_OK)
        throw std::runtime_error(std::string("BDPTIntegrator::render: ") + trb_last_error(c));
    for (int k = 0; k < draws; ++k) sampler.next();
    if (rgb) {  // light-tracing splats land in the frame buffer as a side effect (bdpt.h:360-370); the reference takes a per-pixel lock there
        static std::mutex splat_lock;   // one lock for the few splats of a sample: render() may be called from several host threads
        std::lock_guard<std::mutex> guard(splat_lock);
        for (int k = 0; k < K; ++k)
            if (pix[k] >= 0 && pix[k] < rgb->width * rgb->height) rgb->data[pix[k]] += v3f(srgb[3 * k], srgb[3 * k + 1], srgb[3 * k + 2]);
    }
    return v3f(li[0], li[1], li[2]);
}

// Whole frame on G devices of this process: sample-index sharding (replicated scene, disjoint RNG streams; contributions are
// normalised by the total spp, so partial frame buffers are plain summands), one host thread per device, host-side sum.
template <class RenderCall, class MultiCall>
static bool render_frame_sharded(const Scene& scene, const trb_render_desc& base, RenderBuffer& rgb, trb_stats* stats, RenderCall call, MultiCall multi) {
    const Config& cfg = scene.config;
    const int G = std::max(1, std::min(cfg.gpus, cfg.spp));
    const size_t nf = size_t(cfg.width) * cfg.height * 3;
    float* out = reinterpret_cast<float*>(rgb.data.get());
    if (G == 1) {
        trb_context* c = scene.context(0);
        if (!c) { std::printf("%s\n", scene.error().c_str()); return false; }
        if (call(c, base, out, stats) != TRB_OK) { std::printf("%s\n", trb_last_error(c)); return false; }
        return true;
    }
    {   // the library's own multi-GPU call: one NCCL reduce of the partial frame buffers over NVLink
        std::vector<trb_context*> cs;
        cs.assign(size_t(G), nullptr);
        for (int g = 0; g < G; ++g) { cs[g] = scene.context(g); if (!cs[g]) { std::printf("%s\n", scene.error().c_str()); return false; } }
        const int rc = multi(cs.data(), G, base, out, stats);
        if (rc == TRB_OK) return true;
        if (rc != TRB_ERR_UNSUPPORTED) { std::printf("%s\n", trb_last_error(cs[0])); return false; }
        std::printf("NCCL is not available (%s): summing the partial frame buffers on the host\n", trb_last_error(cs[0]));
    }
    std::vector<std::vector<float>> part;
    part.assign(size_t(G), std::vector<float>(nf));
    std::vector<int> rc;
    rc.assign(size_t(G), 0);
    std::vector<trb_context*> cs;
    cs.assign(size_t(G), nullptr);
    for (int g = 0; g < G; ++g) { cs[g] = scene.context(g); if (!cs[g]) { std::printf("%s\n", scene.error().c_str()); return false; } }
    std::vector<std::thread> th;
    for (int g = 0; g < G; ++g) {
        th.emplace_back([&, g]() {
            trb_render_desc d = base;
            d.spp_begin = int32_t(int64_t(cfg.spp) * g / G);
            d.spp_end = int32_t(int64_t(cfg.spp) * (g + 1) / G);
            rc[g] = call(cs[g], d, part[g].data(), g == 0 ? stats : nullptr);
        });
    }
    for (auto& t : th) t.join();
    for (int g = 0; g < G; ++g) if (rc[g] != TRB_OK) { std::printf("%s\n", trb_last_error(cs[g])); return false; }
    for (size_t i = 0; i < nf; ++i) { float s = 0.f; for (int g = 0; g < G; ++g) s += part[g][i]; out[i] = s; }
    return true;
}

bool BDPTIntegrator::renderFrame() {
    return render_frame_sharded(scene, make_desc(scene.config), *rgb, &lastStats,
                                [](trb_context* c, const trb_render_desc& d, float* out, trb_stats* st) { return trb_render_bdpt(c, &d, 0, out, st); },
                                [](trb_context* const* cs, int n, const trb_render_desc& d, float* out, trb_stats* st) { return trb_render_bdpt_multi(cs, n, &d, out, st); });
}

bool BDPTIntegrator::renderRange(int sppBegin, int sppEnd, bool accumulate) {
    trb_context* c = scene.context(0);
    if (!c) { std::printf("%s\n", scene.error().c_str()); return false; }
    trb_render_desc d = make_desc(scene.config);
    d.spp_begin = sppBegin; d.spp_end = sppEnd;
    if (trb_render_bdpt(c, &d, accumulate ? 1 : 0, reinterpret_cast<float*>(rgb->data.get()), &lastStats) != TRB_OK) { std::printf("%s\n", trb_last_error(c)); return false; }
    return true;
}

void BDPTIntegrator::describe(trb_checkpoint_info& info) const {
    std::memset(&info, 0, sizeof(info));
    info.desc = make_desc(scene.config);
    info.integrator = 0;
}

// ---- PathTracerIntegrator (path.h:15-262) ---------------------------------------------------------------------
PathTracerIntegrator::PathTracerIntegrator(const Scene& s) : Integrator(s) {
    const auto& pt = s.config.integratorSettings.pt;  // path.h:16-22
    m_isExplicit = pt.isExplicit; m_maxDepth = pt.maxDepth; m_rrDepth = pt.rrDepth; m_rrProb = pt.rrProb;
    m_emitterSamples = pt.emitterSamples; m_bsdfSamples = pt.bsdfSamples;
    std::cout << "Samples per pixel are: " << s.config.spp << std::endl;  // path.h:24-27
    std::cout << "DI emitter samples are: " << m_emitterSamples << std::endl;
    std::cout << "DI bsdf samples are: " << m_bsdfSamples << std::endl;
    std::cout << "Max depth is: " << m_maxDepth << std::endl;
}

trb_path_desc PathTracerIntegrator::pathDesc() const {
    trb_path_desc p;
    p.is_explicit = m_isExplicit ? 1 : 0; p.max_depth = m_maxDepth; p.rr_prob = m_rrProb;
    p.emitter_samples = int32_t(m_emitterSamples); p.bsdf_samples = int32_t(m_bsdfSamples);
    return p;
}

v3f PathTracerIntegrator::render(const Ray& ray, Sampler& sampler) const {
    trb_context* c = scene.context(0);
    if (!c) throw std::runtime_error("PathTracerIntegrator::render: no device context");
    trb_render_desc d = make_desc(scene.config);
    d.rr_depth = m_rrDepth;
    const trb_path_desc pd = pathDesc();
    // the walk consumes a prefix of the sampler's stream: feed a copy's numbers, then advance the real sampler
    const size_t stride = 4096;
    Sampler ahead = sampler;
    std::vector<float> rng(stride);
    for (float& f : rng) f = ahead.next();
    trb_ray r = to_ray(ray);
    float li[3] = {0, 0, 0};
    int32_t draws = 0;
    if (trb_path_paths(c, &d, &pd, &r, rng.data(), 1, stride, li, &draws) != TRB_OK)
        throw std::runtime_error(std::string("PathTracerIntegrator::render: ") + trb_last_error(c));
    for (int k = 0; k < draws; ++k) sampler.next();
    return v3f(li[0], li[1], li[2]);
}

bool PathTracerIntegrator::renderFrame() {
    trb_render_desc base = make_desc(scene.config);
    base.rr_depth = m_rrDepth;
    const trb_path_desc pd = pathDesc();
    return render_frame_sharded(scene, base, *rgb, &lastStats,
                                [pd](trb_context* c, const trb_render_desc& d, float* out, trb_stats* st) { return trb_render_path(c, &d, &pd, 0, out, st); },
                                [pd](trb_context* const* cs, int n, const trb_render_desc& d, float* out, trb_stats* st) { return trb_render_path_multi(cs, n, &d, &pd, out, st); });
}

bool PathTracerIntegrator::renderRange(int sppBegin, int sppEnd, bool accumulate) {
    trb_context* c = scene.context(0);
    if (!c) { std::printf("%s\n", scene.error().c_str()); return false; }
    trb_render_desc d = make_desc(scene.config);
    d.rr_depth = m_rrDepth; d.spp_begin = sppBegin; d.spp_end = sppEnd;
    const trb_path_desc pd = pathDesc();
    if (trb_render_path(c, &d, &pd, accumulate ? 1 : 0, reinterpret_cast<float*>(rgb->data.get()), &lastStats) != TRB_OK) { std::printf("%s\n", trb_last_error(c)); return false; }
    return true;
}

void PathTracerIntegrator::describe(trb_checkpoint_info& info) const {
    std::memset(&info, 0, sizeof(info));
    info.desc = make_desc(scene.config);
    info.desc.rr_depth = m_rrDepth;
    info.path = pathDesc();
    info.integrator = 1;
}

// ---- DirectIntegrator (direct.h:74-467) -------------------------------------------------------------------------
DirectIntegrator::DirectIntegrator(const Scene& s) : Integrator(s) {
    const auto& di = s.config.integratorSettings.di;  // direct.h:75-79
    m_emitterSamples = di.emitterSamples; m_bsdfSamples = di.bsdfSamples; m_samplingStrategy = di.samplingStrategy;
}

trb_direct_desc DirectIntegrator::directDesc() const {
    trb_direct_desc d;
    d.strategy = m_samplingStrategy == "mis" ? TRB_DI_MIS : m_samplingStrategy == "area" ? TRB_DI_AREA
               : m_samplingStrategy == "solidAngle" ? TRB_DI_SOLID_ANGLE : m_samplingStrategy == "cosineHemisphere" ? TRB_DI_COSINE_HEMISPHERE
               : m_samplingStrategy == "bsdf" ? TRB_DI_BSDF : TRB_DI_INVALID;
    d.emitter_samples = int32_t(m_emitterSamples); d.bsdf_samples = int32_t(m_bsdfSamples); d.reserved = 0;
    return d;
}

v3f DirectIntegrator::render(const Ray& ray, Sampler& sampler) const {
    trb_context* c = scene.context(0);
    if (!c) throw std::runtime_error("DirectIntegrator::render: no device context");
    const trb_render_desc d = make_desc(scene.config);
    const trb_direct_desc dd = directDesc();
    if (dd.strategy == TRB_DI_INVALID) throw std::runtime_error("Error: wrong strategy");  // direct.h:460
    // the sample consumes a prefix of the sampler's stream: feed a copy's numbers, then advance the real sampler
    const size_t stride = 3 * m_emitterSamples + 2 * m_bsdfSamples + 4;
    Sampler ahead = sampler;
    std::vector<float> rng(stride);
    for (float& f : rng) f = ahead.next();
    trb_ray r = to_ray(ray);
    float li[3] = {0, 0, 0};
    int32_t draws = 0;
    if (trb_direct_paths(c, &d, &dd, &r, rng.data(), 1, stride, li, &draws) != TRB_OK)
        throw std::runtime_error(std::string("DirectIntegrator::render: ") + trb_last_error(c));
    for (int k = 0; k < draws; ++k) sampler.next();
    return v3f(li[0], li[1], li[2]);
}

bool DirectIntegrator::renderFrame() {
    const trb_direct_desc dd = directDesc();
    if (dd.strategy == TRB_DI_INVALID) { std::cout << "Error: wrong strategy" << std::endl; return false; }  // direct.h:460 (the reference then exits)
    return render_frame_sharded(scene, make_desc(scene.config), *rgb, &lastStats,
                                [dd](trb_context* c, const trb_render_desc& d, float* out, trb_stats* st) { return trb_render_direct(c, &d, &dd, 0, out, st); },
                                [dd](trb_context* const* cs, int n, const trb_render_desc& d, float* out, trb_stats* st) { return trb_render_direct_multi(cs, n, &d, &dd, out, st); });
}

bool DirectIntegrator::renderRange(int sppBegin, int sppEnd, bool accumulate) {
    trb_context* c = scene.context(0);
    if (!c) { std::printf("%s\n", scene.error().c_str()); return false; }
    trb_render_desc d = make_desc(scene.config);
    d.spp_begin = sppBegin; d.spp_end = sppEnd;
    const trb_direct_desc dd = directDesc();
    if (dd.strategy == TRB_DI_INVALID) { std::cout << "Error: wrong strategy" << std::endl; return false; }
    if (trb_render_direct(c, &d, &dd, accumulate ? 1 : 0, reinterpret_cast<float*>(rgb->data.get()), &lastStats) != TRB_OK) { std::printf("%s\n", trb_last_error(c)); return false; }
    return true;
}

void DirectIntegrator::describe(trb_checkpoint_info& info) const {
    std::memset(&info, 0, sizeof(info));
    info.desc = make_desc(scene.config);
    const trb_direct_desc dd = directDesc();
    info.path.is_explicit = dd.strategy; info.path.emitter_samples = dd.emitter_samples; info.path.bsdf_samples = dd.bsdf_samples;  // trb.h: integrator 2
    info.integrator = 2;
}

// ---- Renderer (renderer.cpp:53-227) ---------------------------------------------------------------------------
Renderer::Renderer(const Config& config) : scene(config) {}

bool Renderer::init(bool isRealTime, bool nogui_) {
    realTime = isRealTime; nogui = nogui_;
    if (!scene.load(isRealTime)) return false;
    if (scene.config.integrator == EBDPTIntegrator) integrator.reset(new BDPTIntegrator(scene));
    else if (scene.config.integrator == EPathTracerIntegrator) integrator.reset(new PathTracerIntegrator(scene));
    else if (scene.config.integrator == EDirectIntegrator) integrator.reset(new DirectIntegrator(scene));
    else {
        std::printf("Only the BDPT, path and direct integrators are hosted on the B200 path (renderer.type = \"bdpt\" | \"path\" | \"direct\")\n");
        return false;
    }
    return integrator->init();
}

void Renderer::render() {
    if (!integrator) return;
    const Config& cfg = scene.config;
    if (cfg.checkpointEvery <= 0 || cfg.checkpointFile.empty()) {
        if (!integrator->renderFrame()) throw std::runtime_error("Renderer::render: device render failed");
        return;
    }
    // the frame in slices of checkpointEvery samples per pixel; after each slice the running sum is saved, and a
    // run that finds a checkpoint of this very render (same camera, film, spp, depth, seed, integrator, scene) continues from it.
    trb_checkpoint_info me;
    integrator->describe(me);
    const trb_scene_desc* sd = scene.desc();
    me.scene_hash = trb_host_scene_hash(sd);   // content of every scene array: an edited OBJ / MTL / texture starts a new render
    if (cfg.gpus > 1) std::printf("Checkpointed renders run on one device (gpus = %d ignored): each slice is rendered by context 0\n", cfg.gpus);
    float* out = reinterpret_cast<float*>(integrator->rgb->data.get());
    const size_t nf = size_t(cfg.width) * cfg.height * 3;
    int done = 0;
    trb_checkpoint_info old;
    if (trb_host_load_checkpoint(cfg.checkpointFile.c_str(), &old, nullptr, 0) == TRB_OK && trb_host_checkpoint_matches(&old, &me) &&
        old.spp_done > 0 && old.spp_done <= cfg.spp && trb_host_load_checkpoint(cfg.checkpointFile.c_str(), &old, out, nf) == TRB_OK) {
        trb_context* c = scene.context(0);
        if (!c || trb_framebuffer_upload(c, out, cfg.width, cfg.height) != TRB_OK) throw std::runtime_error("Renderer::render: cannot restore the checkpoint");
        done = old.spp_done;
        std::printf("Resuming from %s: %d of %d samples per pixel done\n", cfg.checkpointFile.c_str(), done, cfg.spp);
    }
    int written = 0;
    while (done < cfg.spp) {
        if (cfg.stopAfterCheckpoints > 0 && written >= cfg.stopAfterCheckpoints) { std::printf("Stopped after %d checkpoint(s)\n", written); return; }
        const int end = std::min(done + cfg.checkpointEvery, cfg.spp);
        if (!integrator->renderRange(done, end, done > 0)) throw std::runtime_error("Renderer::render: device render failed");
        done = end;
        me.spp_done = done;
        if (trb_host_save_checkpoint(cfg.checkpointFile.c_str(), &me, out) != TRB_OK) std::printf("%s\n", trb_host_last_error());
        else { ++written; std::printf("Checkpoint: %d of %d samples per pixel\n", done, cfg.spp); }
        if (!cfg.previewPng.empty()) {
            // the buffer is normalised by the TOTAL spp: scale the preview up to what is there so far
            if (trb_host_save_png(cfg.previewPng.c_str(), out, cfg.width, cfg.height, float(cfg.spp) / float(done)) != TRB_OK) std::printf("%s\n", trb_host_last_error());
        }
    }
}

void Renderer::cleanUp() {
    if (integrator) integrator->cleanUp();
}

}  // namespace TinyRender
