// tinyrender.h -- TinyRender's C++ host API for the offline BDPT path, re-hosted on the B200 library.
//
// Same names, argument meaning and error behaviour as the reference's in-process operator API
// (JackMinn/Bidirectional-Path-Tracing):
//   Config, Camera               src/core/core.h:183-248
//   loadTOML                     src/main.cpp:22-116       (returns the `realtime` flag; throws std::runtime_error
//                                                           on an invalid integrator type)
//   Ray, SurfaceInteraction      src/core/core.h:117-122, 173-180
//   RenderBuffer                 src/core/core.h:128-146
//   Sampler                      src/core/math.h:63-76     (std::mt19937 + uniform_real_distribution<float>)
//   BSDF + the five subclasses   src/core/core.h:256-318, src/bsdfs/*.h
//   AcceleratorBVH               src/core/accel.h:113-172
//   Scene                        src/core/core.h:352-374, src/core/renderer.cpp:235-315
//   Integrator, BDPTIntegrator   src/core/integrator.h:23-70, src/integrators/bdpt.h:37-518
//   PathTracerIntegrator         src/integrators/path.h:15-262
//   DirectIntegrator             src/integrators/direct.h:74-467
//   Renderer                     src/core/renderer.h:20-33, src/core/renderer.cpp:53-227
// Two additive changes: Integrator gains `virtual bool renderFrame()` (the whole offline pixel loop of
// Renderer::render as ONE call; BDPTIntegrator implements it with trb_render_bdpt) and the single-item virtuals
// have batch twins.  Everything computes on the GPU through the C ABI of include/trb.h; there is no CPU path:
// Scene::load returns false (and says why) when no sm_100 device is usable.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "../trb.h"

namespace TinyRender {

struct v2f { float x = 0, y = 0; v2f() = default; v2f(float a, float b) : x(a), y(b) {} };
struct v3f {
    float x = 0, y = 0, z = 0;
    v3f() = default;
    explicit v3f(float s) : x(s), y(s), z(s) {}
    v3f(float a, float b, float c) : x(a), y(b), z(c) {}
    v3f& operator+=(const v3f& o) { x += o.x; y += o.y; z += o.z; return *this; }
    v3f operator+(const v3f& o) const { return {x + o.x, y + o.y, z + o.z}; }
    v3f operator*(float s) const { return {x * s, y * s, z * s}; }
    bool operator==(const v3f& o) const { return x == o.x && y == o.y && z == o.z; }
};

enum EIntegrator { ENormalIntegrator, ESimpleIntegrator, EAOIntegrator, EROIntegrator, EDirectIntegrator, EPathTracerIntegrator, EBDPTIntegrator };

struct Camera {  // core.h:183-190
    v3f o, at, up;
    float fov = 30.f;
};

struct Config {  // core.h:195-248 (the fields the offline BDPT path reads) + optional B200 keys
    EIntegrator integrator = EBDPTIntegrator;
    Camera camera;
    std::string objFile, tomlFile;
    int width = 0, height = 0, spp = 1;
    struct IntegratorConfig {
        struct bdpt_s { int rrDepth = 5; float rrProb = 0.f; } bdpt;  // main.cpp:103-107 (rrProb is read and never used)
        struct pt_s {  // main.cpp:94-102, core.h:229-236
            bool isExplicit = true; int maxDepth = -1; int rrDepth = 5; float rrProb = 0.95f;
            size_t emitterSamples = 1, bsdfSamples = 0;
        } pt;
        struct di_s {  // main.cpp:88-93: the reference's own default strategy name, "emitter", is one DirectIntegrator::render rejects
            size_t emitterSamples = 1, bsdfSamples = 1; std::string samplingStrategy = "emitter";
        } di;
    } integratorSettings;
    // additive [renderer] keys (defaults reproduce the reference run)
    int gpus = 1;                 // devices used by Renderer::render, sharded by sample index
    uint32_t seed = 260450963u;   // renderer.cpp:155
    int64_t waveSamples = 0;      // samples per wavefront, 0 = library default
    bool russianRoulette = false; // bdpt.h:18 ships NO_RR 1; true runs the reference's own NO_RR 0 code (TRB_FLAG_RUSSIAN_ROULETTE)
    int maxDepth = 0;             // depth cap under russianRoulette, 0 = TRB_RR_DEFAULT_MAX_DEPTH
    // restartable renders: every checkpointEvery samples per pixel the running frame buffer goes to checkpointFile (and, tone-
    // mapped, to previewPng); a run that finds a checkpoint of the same render continues from it
    std::string checkpointFile, previewPng;
    int checkpointEvery = 0;
    int stopAfterCheckpoints = 0; // > 0: leave Renderer::render() after that many checkpoints, as an interrupted run would (tests; CLI: $TRB_STOP_AFTER_CHECKPOINTS)
};

// main.cpp:22-116. Returns renderer.realtime. Throws std::runtime_error("Invalid integrator type") for an unknown
// renderer.type, and std::runtime_error for a file that cannot be read or parsed.
bool loadTOML(Config& config, const std::string& inputFile);

struct Ray {  // core.h:117-122
    v3f o, d;
    float min_t, max_t;
    Ray(const v3f& o_, const v3f& d_, float tmin = 1e-8f, float tmax = 3.402823466e+38f) : o(o_), d(d_), min_t(tmin), max_t(tmax) {}
};

struct SurfaceInteraction {  // core.h:173-180 (frames are rebuilt on the device from the shading normal)
    v3f p, wo, wi;
    float t = 0, u = 0, v = 0;
    size_t shapeID = 0, primID = 0;
    int matID = 0;
};

struct RenderBuffer {  // core.h:128-146
    int width, height;
    std::unique_ptr<v3f[]> data;
    RenderBuffer(int w, int h) : width(w), height(h), data(new v3f[size_t(w) * h]) {}
    void clear() { for (size_t i = 0; i < size_t(width) * height; ++i) data[i] = v3f(0.f); }
};

struct Sampler {  // math.h:63-76
    std::mt19937 g;
    std::uniform_real_distribution<float> d;
    explicit Sampler(int seed) : g(seed), d(0.f, 1.f) {}
    float next() { return d(g); }
    v2f next2D() { float a = d(g); float b = d(g); return {a, b}; }
};

struct Scene;

// WorldData (core.h:151-160 holds tinyobj's attrib / shapes / materials): here a thin VIEW over the flat arrays of the loaded
// scene (what tinyobj::LoadObj yields, shape-major; trb_scene_desc), plus the per-shape centres and radii Scene::load derives
// (renderer.cpp:294-304).  It exists so that code written against the reference's signatures - BSDF(const WorldData&, const
// Config&, size_t), scene.worldData - compiles; the arrays are owned by the Scene.
struct WorldData {
    const Scene* owner = nullptr;
    const trb_scene_desc* flat = nullptr;        // attrib.vertices / normals / texcoords, indices, material ids, materials
    std::vector<v3f> shapesCenter;               // renderer.cpp:294-301
    std::vector<float> shapesRadius;             // shapesAABOX[s].max.x - shapesCenter[s].x (renderer.cpp:349-353)
    size_t nShapes() const { return flat ? flat->n_shapes : 0; }
    size_t nMaterials() const { return flat ? flat->n_materials : 0; }
};

// BSDF (core.h:256-318). eval/sample return f * cos(theta_i); directions are in the local shading frame.
struct BSDF {
    enum EBSDFType { EUnknown = 0x0, EDiffuseReflection = 0x2, EGlossyReflection = 0x8, EDeltaReflection = 0x20, EDeltaTransmission = 0x40,
                     EDelta = 0x1 | EDeltaReflection | EDeltaTransmission };
    const WorldData& worldData;   // core.h:296-298
    const Config& config;
    v3f emission;                 // core.h:299, :303: materials[matID].emission
    const Scene& scene;
    size_t matID;
    unsigned combinedType = EUnknown;
    bool emissive = false;
    BSDF(const WorldData& d, const Config& c, size_t matID);   // the reference's constructor (core.h:299-305)
    BSDF(const Scene& s, size_t id);
    virtual ~BSDF() = default;
    virtual v3f eval(const SurfaceInteraction& i) const;
    virtual float pdf(const SurfaceInteraction& i) const;
    virtual v3f sample(SurfaceInteraction& i, const v2f& sample, float* pdf) const;  // writes i.wi
    unsigned getType() const { return combinedType; }
    bool isEmissive() const { return emissive; }
    virtual std::string toString() const = 0;
    // batch twin of the three virtuals (one device launch): q/out as in trb.h
    bool evalBatch(const trb_bsdf_query* q, size_t n, trb_bsdf_out* out) const;
};
struct DiffuseBSDF : BSDF { using BSDF::BSDF; std::string toString() const override { return "Diffuse"; } };       // bsdfs/diffuse.h
struct PhongBSDF : BSDF { using BSDF::BSDF; std::string toString() const override { return "Phong"; } };           // bsdfs/phong.h
struct MixtureBSDF : BSDF { using BSDF::BSDF; std::string toString() const override { return "Mixture"; } };       // bsdfs/mixture.h
struct MirrorBSDF : BSDF { using BSDF::BSDF; std::string toString() const override { return "Mirror"; } };         // bsdfs/perfectmirror.h
struct GlassBSDF : BSDF { using BSDF::BSDF; std::string toString() const override { return "Glass"; } };           // bsdfs/glass.h

struct Emitter {  // core.h:324-332
    size_t shapeID = 0;
    float area = 0.f;
    v3f radiance;
    std::vector<float> faceAreaCdf;  // Distribution1D::cdf, normalised
};

// AcceleratorBVH (accel.h:113-172): same hit semantics (closest hit, then the [min_t,max_t] test on it only).
struct AcceleratorBVH {
    const Scene& scene;
    explicit AcceleratorBVH(const Scene& s) : scene(s) {}
    bool build();  // the reference-order BVH is built and uploaded by Scene::load; kept for source compatibility
    bool intersect(const Ray& ray, SurfaceInteraction& info, bool occlusion = false) const;  // `occlusion` ignored, as in the reference
    bool occluded(const Ray& ray) const;  // bvh->getIntersection(ray, &info, true) as used by bdpt.h:505
    // batch twins
    bool intersectBatch(const trb_ray* rays, size_t n, trb_hit* out) const;
    bool occludedBatch(const trb_ray* rays, size_t n, uint8_t* out) const;
};

struct Scene {  // core.h:352-374
    const Config& config;  // a reference, as in the reference: the config must outlive the scene
    std::vector<std::unique_ptr<BSDF>> bsdfs;
    std::vector<Emitter> emitters;
    std::unique_ptr<AcceleratorBVH> bvh;
    WorldData worldData;          // core.h:355 (a view, see WorldData)
    struct { float min[3], max[3]; } aabb{};
    size_t nShapes = 0, nTriangles = 0;
    float getShapeArea(size_t shapeID) const;     // renderer.cpp:317-339: sum of the shape's face areas (fp32, face order)
    float getShapeRadius(size_t shapeID) const;   // renderer.cpp:349-353
    v3f getShapeCenter(size_t shapeID) const;     // renderer.cpp:355-358
    size_t getFirstLight() const { return emitters.empty() ? 0 : emitters[0].shapeID; }   // renderer.cpp:360-362

    explicit Scene(const Config& config);
    ~Scene();
    bool load(bool isRealTime);  // renderer.cpp:235-315; false on I/O errors, unsupported materials, or no usable GPU
    trb_context* context(int device = 0) const;  // device 0 is created by load(); others on demand
    const trb_scene_desc* desc() const;
    const std::string& error() const { return err; }

private:
    friend struct Renderer;
    trb_host_scene* host = nullptr;
    trb_scene_desc sdesc{};
    mutable std::vector<trb_context*> ctx;
    mutable std::string err;
};

struct Integrator {  // integrator.h:23-70
    const Scene& scene;
    std::unique_ptr<RenderBuffer> rgb;
    explicit Integrator(const Scene& s) : scene(s) {}
    virtual ~Integrator() = default;
    virtual bool init();                                        // integrator.cpp:16-20: allocate + clear rgb
    virtual v3f render(const Ray&, Sampler&) const = 0;         // one sample
    virtual bool renderFrame() { return false; }                // ADDED: the offline pixel loop as one device call
    // ADDED: samples [sppBegin, sppEnd) of every pixel on device 0, onto the context's frame buffer when `accumulate`; rgb = the sum so far
    virtual bool renderRange(int /*sppBegin*/, int /*sppEnd*/, bool /*accumulate*/) { return false; }
    virtual void describe(trb_checkpoint_info& info) const { std::memset(&info, 0, sizeof(info)); }
    virtual void cleanUp() { save(); }                          // integrator.cpp:22-24
    bool save();                                                // writes <toml stem>.exr (integrator.cpp:26-30, utils.h:95-156)
    // emitter / BSDF getters (integrator.h:37-60, integrator.cpp:32-63): host evaluators over the tables the device uses
    const Emitter& getEmitterByID(int emitterID) const { return scene.emitters[size_t(emitterID)]; }
    size_t getEmitterIDByShapeID(size_t shapeID) const;         // throws std::out_of_range where the reference asserts
    float getEmitterPdf(const Emitter&) const { return 1.f / scene.emitters.size(); }   // integrator.cpp:60-62: the selection probability, as written
    const BSDF* getBSDF(const SurfaceInteraction& hit) const;   // by (shapeID, primID), integrator.cpp:35-39
    v3f getEmission(const SurfaceInteraction& hit) const;       // materials[hit.matID].emission, integrator.cpp:41-44
    size_t selectEmitter(float sample, float& pdf) const;       // integrator.cpp:46-51
};

struct BDPTIntegrator : Integrator {  // bdpt.h:37-518
    int m_rrDepth;
    float m_rrProb;
    mutable trb_stats lastStats{};
    explicit BDPTIntegrator(const Scene& s);
    // One BDPT sample on the device with the sampler's own number stream, consumed in the reference's order; the
    // light-tracing splats are added into rgb->data as the reference does (bdpt.h:360-370). Returns Li.
    v3f render(const Ray& ray, Sampler& sampler) const override;
    bool renderFrame() override;
    bool renderRange(int sppBegin, int sppEnd, bool accumulate) override;
    void describe(trb_checkpoint_info& info) const override;
};

struct PathTracerIntegrator : Integrator {  // path.h:15-262, the estimator the reference cross-checks BDPT with
    bool m_isExplicit; int m_maxDepth, m_rrDepth; float m_rrProb; size_t m_emitterSamples, m_bsdfSamples;
    mutable trb_stats lastStats{};
    explicit PathTracerIntegrator(const Scene& s);
    // One sample on the device with the sampler's own number stream, consumed in the reference's (depth-first) order.
    v3f render(const Ray& ray, Sampler& sampler) const override;
    bool renderFrame() override;
    bool renderRange(int sppBegin, int sppEnd, bool accumulate) override;
    void describe(trb_checkpoint_info& info) const override;
    trb_path_desc pathDesc() const;
};

struct DirectIntegrator : Integrator {  // direct.h:74-467: one-bounce direct illumination, five sampling strategies
    size_t m_emitterSamples, m_bsdfSamples;
    std::string m_samplingStrategy;  // "mis" | "area" | "solidAngle" | "cosineHemisphere" | "bsdf"
    mutable trb_stats lastStats{};
    explicit DirectIntegrator(const Scene& s);
    // One sample on the device with the sampler's own number stream, consumed in the reference's order.  An unknown strategy
    // throws std::runtime_error("Error: wrong strategy") where the reference prints that and exits (direct.h:460-461).
    v3f render(const Ray& ray, Sampler& sampler) const override;
    bool renderFrame() override;
    bool renderRange(int sppBegin, int sppEnd, bool accumulate) override;
    void describe(trb_checkpoint_info& info) const override;
    trb_direct_desc directDesc() const;
};

struct Renderer {  // renderer.h:20-33
    std::unique_ptr<Integrator> integrator;
    Scene scene;
    bool realTime = false, nogui = true;
    explicit Renderer(const Config& config);
    bool init(bool isRealTime, bool nogui);  // renderer.cpp:55-110 (offline branch; isRealTime => false: no graphics stack)
    void render();                           // renderer.cpp:112-215 (offline branch) => integrator->renderFrame(), or the
                                             // checkpointed loop over integrator->renderRange() (Config::checkpointEvery)
    void cleanUp();                          // renderer.cpp:221-227
};

}  // namespace TinyRender
