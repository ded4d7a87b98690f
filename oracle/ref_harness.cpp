/* oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Thin extern "C" batch interface over the UNMODIFIED reference (compiled from /root/reference by
 * oracle/Makefile into oracle/_ref/libtr_ref.so).  It is used to
 *   (1) generate the golden vectors committed under tests/golden/ (tools/gen_golden.py),
 *   (2) validate the plain-C++ restatement in oracle/bdpt_oracle.cpp,
 *   (3) time the reference's own CPU renderer (bench.py --impl reference / cpu_baseline).
 * Every entry point calls the reference's own functions; nothing is re-implemented here.
 *
 * Reference entry points exercised (file:line under /root/reference):
 *   loadTOML                               src/main.cpp:22-116
 *   Renderer::{init,render}                src/core/renderer.cpp:55-110,112-215
 *   AcceleratorBVH::intersect              src/core/accel.h:125-172
 *   BVH::getIntersection(occlusion)        externals/bvh.h:259-352
 *   BSDF::{eval,pdf,sample}                src/bsdfs/*.h
 *   BDPTIntegrator::{render,visibilityQuery}  src/integrators/bdpt.h:219-241,498-514
 *   Sampler                                src/core/math.h:63-76
 */
#include <core/core.h>
#include <core/accel.h>
#include <core/renderer.h>
#include <bsdfs/diffuse.h>
#include <bsdfs/phong.h>
#include <bsdfs/mixture.h>
#include <bsdfs/perfectmirror.h>
#include <bsdfs/glass.h>
#include <integrators/bdpt.h>
#include <chrono>
#include <cstdint>

bool loadTOML(TinyRender::Config& config, const std::string& inputFile); /* src/main.cpp:22 */
extern std::unique_ptr<std::mutex[]> g_FrameBufferLocks;                  /* src/main.cpp:17 */

using namespace TinyRender;

namespace {
struct Handle {
    Config config;
    std::unique_ptr<Renderer> renderer;
    const BDPTIntegrator* bdpt = nullptr;
    std::vector<int> triBase;  // first global triangle id of each shape (shape-major numbering)
    size_t nTris = 0;
};

/* The reference's BVH keeps its flat tree public but its primitive list private; the harness gets the
 * build-reordered primitive order from AcceleratorBVH::objects, which BVH::build permutes in place
 * (externals/bvh.h:147-247 works on the pointer it was given, accel.h:121). */
inline Ray mkRay(const float* r) { return Ray(v3f(r[0], r[1], r[2]), v3f(r[3], r[4], r[5]), r[6], r[7]); }
}  // namespace

extern "C" {

void* refh_open(const char* toml) {
    auto* h = new Handle();
    try {
        bool rt = loadTOML(h->config, toml);
        if (rt) { delete h; return nullptr; }
        h->renderer.reset(new Renderer(h->config));
        if (!h->renderer->init(false, true)) { delete h; return nullptr; }
    } catch (std::exception const& e) {
        std::cerr << "refh_open: " << e.what() << std::endl;
        delete h;
        return nullptr;
    }
    size_t fb = size_t(h->config.width) * h->config.height;
    g_FrameBufferLocks.reset(new std::mutex[fb]); /* as src/main.cpp:137-143 */
    h->bdpt = dynamic_cast<const BDPTIntegrator*>(h->renderer->integrator.get());
    const auto& shapes = h->renderer->scene.worldData.shapes;
    int base = 0;
    for (auto& s : shapes) { h->triBase.push_back(base); base += int(s.mesh.indices.size() / 3); }
    h->nTris = base;
    return h;
}

void refh_close(void* hp) { delete static_cast<Handle*>(hp); }

/* counts: [nShapes, nMaterials, nVertexFloats, nNormalFloats, nTris, nEmitters, width, height, spp, rrDepth] */
void refh_counts(void* hp, int64_t* out) {
    auto* h = static_cast<Handle*>(hp);
    const auto& wd = h->renderer->scene.worldData;
    out[0] = wd.shapes.size(); out[1] = wd.materials.size();
    out[2] = wd.attrib.vertices.size(); out[3] = wd.attrib.normals.size();
    out[4] = h->nTris; out[5] = h->renderer->scene.emitters.size();
    out[6] = h->config.width; out[7] = h->config.height; out[8] = h->config.spp;
    out[9] = h->config.integratorSettings.pt.rrDepth;
}

/* Dump of what tinyobj produced, to validate the product's own OBJ/MTL loader. */
void refh_scene_dump(void* hp, float* vertices, float* normals, int* tri_v, int* tri_n, int* tri_mat,
                     int* tri_shape) {
    auto* h = static_cast<Handle*>(hp);
    const auto& wd = h->renderer->scene.worldData;
    std::copy(wd.attrib.vertices.begin(), wd.attrib.vertices.end(), vertices);
    std::copy(wd.attrib.normals.begin(), wd.attrib.normals.end(), normals);
    size_t t = 0;
    for (size_t j = 0; j < wd.shapes.size(); j++) {
        const auto& m = wd.shapes[j].mesh;
        for (size_t i = 0; i < m.indices.size(); i += 3, t++) {
            for (int k = 0; k < 3; k++) {
                tri_v[3 * t + k] = m.indices[i + k].vertex_index;
                tri_n[3 * t + k] = m.indices[i + k].normal_index;
            }
            tri_mat[t] = m.material_ids[i / 3];
            tri_shape[t] = int(j);
        }
    }
}

/* materials: per material 16 floats: Kd(3) Ks(3) Ke(3) Tf(3) Ns Ni illum pad */
void refh_materials(void* hp, float* out) {
    auto* h = static_cast<Handle*>(hp);
    const auto& mats = h->renderer->scene.worldData.materials;
    for (size_t i = 0; i < mats.size(); i++) {
        float* o = out + 16 * i;
        for (int k = 0; k < 3; k++) {
            o[k] = mats[i].diffuse[k]; o[3 + k] = mats[i].specular[k];
            o[6 + k] = mats[i].emission[k]; o[9 + k] = mats[i].transmittance[k];
        }
        o[12] = mats[i].shininess; o[13] = mats[i].ior; o[14] = float(mats[i].illum); o[15] = 0.f;
    }
}

/* emitters: per emitter [shapeID, area, radiance(3)] as 5 floats; cdf arrays concatenated after a size table */
void refh_emitters(void* hp, float* out5, int* cdf_sizes) {
    auto* h = static_cast<Handle*>(hp);
    const auto& em = h->renderer->scene.emitters;
    for (size_t i = 0; i < em.size(); i++) {
        out5[5 * i + 0] = float(em[i].shapeID); out5[5 * i + 1] = em[i].area;
        out5[5 * i + 2] = em[i].radiance.x; out5[5 * i + 3] = em[i].radiance.y; out5[5 * i + 4] = em[i].radiance.z;
        cdf_sizes[i] = int(em[i].faceAreaDistribution.cdf.size());
    }
}
void refh_emitter_cdf(void* hp, int e, float* cdf) {
    auto* h = static_cast<Handle*>(hp);
    const auto& c = h->renderer->scene.emitters[e].faceAreaDistribution.cdf;
    std::copy(c.begin(), c.end(), cdf);
}

/* order[pos] = global (shape-major) triangle id sitting at position pos of the build-reordered list. */
void refh_bvh_order(void* hp, int* order) {
    auto* h = static_cast<Handle*>(hp);
    const auto& objs = h->renderer->scene.bvh->objects;
    for (size_t p = 0; p < objs.size(); p++) {
        auto* n = static_cast<const AcceleratorBVH::BVHNode*>(objs[p]);
        order[p] = h->triBase[n->shapeID] + int(n->faceID / 3);
    }
}

/* Flat tree: returns node count when arrays are null. bbox = 6 floats (min,max); meta = start,nPrims,rightOffset */
int refh_bvh_nodes(void* hp, float* bbox, uint32_t* meta, int maxNodes) {
    auto* h = static_cast<Handle*>(hp);
    const BVHFlatNode* ft = h->renderer->scene.bvh->bvh->flatTree;
    /* nNodes is private: walk the pre-order array (left = i+1, right = i+rightOffset) to find its extent */
    int n = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int i = stack.back(); stack.pop_back();
        n = std::max(n, i + 1);
        if (ft[i].rightOffset != 0) { stack.push_back(i + 1); stack.push_back(i + int(ft[i].rightOffset)); }
    }
    if (!bbox) return n;
    for (int i = 0; i < n && i < maxNodes; i++) {
        for (int k = 0; k < 3; k++) { bbox[6 * i + k] = ft[i].bbox.min[k]; bbox[6 * i + 3 + k] = ft[i].bbox.max[k]; }
        meta[3 * i] = ft[i].start; meta[3 * i + 1] = ft[i].nPrims; meta[3 * i + 2] = ft[i].rightOffset;
    }
    return n;
}

/* Level 1: closest hit through AcceleratorBVH::intersect. rays: n x 8 floats (o,d,min_t,max_t).
 * ids: n x 3 (hit?1:0, shapeID, primID); tuv: n x 3; geo: n x 12 (p, wo, frameNs.n, frameNs.s) */
void refh_intersect(void* hp, const float* rays, int64_t n, int* ids, float* tuv, float* geo) {
    auto* h = static_cast<Handle*>(hp);
    const auto& bvh = *h->renderer->scene.bvh;
    for (int64_t i = 0; i < n; i++) {
        Ray r = mkRay(rays + 8 * i);
        SurfaceInteraction si = SurfaceInteraction();
        bool hit = bvh.intersect(r, si);
        ids[3 * i] = hit ? 1 : 0;
        ids[3 * i + 1] = hit ? int(si.shapeID) : -1;
        ids[3 * i + 2] = hit ? int(si.primID) : -1;
        tuv[3 * i] = si.t; tuv[3 * i + 1] = hit ? si.u : 0.f; tuv[3 * i + 2] = hit ? si.v : 0.f;
        if (geo) {
            float* g = geo + 12 * i;
            if (hit) {
                for (int k = 0; k < 3; k++) {
                    g[k] = si.p[k]; g[3 + k] = si.wo[k]; g[6 + k] = si.frameNs.n[k]; g[9 + k] = si.frameNs.s[k];
                }
            } else {
                for (int k = 0; k < 12; k++) g[k] = 0.f;
            }
        }
    }
}

/* Level 1: raw occlusion query, exactly the call made at bdpt.h:505. */
void refh_occluded(void* hp, const float* rays, int64_t n, uint8_t* out) {
    auto* h = static_cast<Handle*>(hp);
    const auto& bvh = *h->renderer->scene.bvh->bvh;
    for (int64_t i = 0; i < n; i++) {
        Ray r = mkRay(rays + 8 * i);
        IntersectionInfo info = IntersectionInfo();
        out[i] = bvh.getIntersection(r, &info, true) ? 1 : 0;
    }
}

/* BDPTIntegrator::visibilityQuery(start,end): segs n x 6 */
void refh_visibility(void* hp, const float* segs, int64_t n, uint8_t* out) {
    auto* h = static_cast<Handle*>(hp);
    for (int64_t i = 0; i < n; i++) {
        const float* s = segs + 6 * i;
        out[i] = h->bdpt->visibilityQuery(v3f(s[0], s[1], s[2]), v3f(s[3], s[4], s[5])) ? 1 : 0;
    }
}

/* Level 2: BSDF batch. q: n x 8 floats (wo(3), wi(3), u(2)); out: n x 12 floats
 * (eval(3), pdf, sample_value(3), sample_pdf, sample_wi(3), type). eval/pdf use (wo,wi); sample uses (wo,u). */
void refh_bsdf(void* hp, int matID, const float* q, int64_t n, float* out) {
    auto* h = static_cast<Handle*>(hp);
    const BSDF* b = h->renderer->scene.bsdfs[matID].get();
    for (int64_t i = 0; i < n; i++) {
        const float* x = q + 8 * i;
        float* o = out + 12 * i;
        SurfaceInteraction si = SurfaceInteraction();
        si.matID = matID;
        si.wo = v3f(x[0], x[1], x[2]);
        si.wi = v3f(x[3], x[4], x[5]);
        v3f e = b->eval(si);
        o[0] = e.x; o[1] = e.y; o[2] = e.z;
        o[3] = b->pdf(si);
        SurfaceInteraction ss = SurfaceInteraction();
        ss.matID = matID;
        ss.wo = si.wo;
        float pdf = 0.f;
        v3f sv = b->sample(ss, v2f(x[6], x[7]), &pdf);
        o[4] = sv.x; o[5] = sv.y; o[6] = sv.z; o[7] = pdf;
        o[8] = ss.wi.x; o[9] = ss.wi.y; o[10] = ss.wi.z;
        o[11] = float(b->getType());
    }
}

/* The reference Sampler's float stream for a seed (src/core/math.h:63-76). */
void refh_sampler_stream(int seed, int64_t n, float* out) {
    Sampler s(seed);
    for (int64_t i = 0; i < n; i++) out[i] = s.next();
}

/* Level 2: full BDPT samples, single-threaded, one private Sampler(seed) per path.
 * rays: n x 8; seeds: n; Li: n x 3; splats: up to maxSplats per path as (pixel, r, g, b) in
 * splat_pix[n*maxSplats] (-1 = none) / splat_rgb[n*maxSplats*3]; nDraws[n] = sampler draws consumed
 * (measured by replaying the seed until the engine state matches).
 * The light-tracing splats are recovered as the non-zero pixels of the (pre-cleared) RenderBuffer, which
 * is the only place the reference records them (bdpt.h:360-370). */
void refh_bdpt_paths(void* hp, const float* rays, const int* seeds, int64_t n, int maxSplats, float* Li,
                     int* splat_pix, float* splat_rgb, int* nDraws) {
    auto* h = static_cast<Handle*>(hp);
    RenderBuffer& rb = *h->renderer->integrator->rgb;
    const int np = rb.width * rb.height;
    for (int64_t i = 0; i < n; i++) {
        rb.clear();
        Ray r = mkRay(rays + 8 * i);
        Sampler s(seeds[i]);
        v3f L = h->bdpt->render(r, s);
        Li[3 * i] = L.x; Li[3 * i + 1] = L.y; Li[3 * i + 2] = L.z;
        int k = 0;
        for (int m = 0; m < maxSplats; m++) splat_pix[i * maxSplats + m] = -1;
        for (int p = 0; p < np; p++) {
            const v3f& c = rb.data[p];
            if (c.x != 0.f || c.y != 0.f || c.z != 0.f) {
                if (k < maxSplats) {
                    splat_pix[i * maxSplats + k] = p;
                    splat_rgb[(i * maxSplats + k) * 3 + 0] = c.x;
                    splat_rgb[(i * maxSplats + k) * 3 + 1] = c.y;
                    splat_rgb[(i * maxSplats + k) * 3 + 2] = c.z;
                }
                k++;
            }
        }
        if (nDraws) {
            Sampler replay(seeds[i]);
            int c = 0;
            while (!(replay.g == s.g) && c < 100000) { replay.next(); c++; }
            nDraws[i] = c;
        }
    }
    rb.clear();
}

/* The four camera matrices, built with the reference's glm by the expressions of renderer.cpp:143-153 /
 * bdpt.h:485-490: worldToCamera, cameraToWorld, cameraToClip, NDCToScreen (column-major). */
void refh_camera_matrices(void* hp, float* out64) {
    auto* h = static_cast<Handle*>(hp);
    const Config& c = h->config;
    float aspectRatio = (float)c.width / c.height;
    auto worldToCamera = glm::lookAt(c.camera.o, c.camera.at, c.camera.up);
    auto cameraToWorld = glm::inverse(worldToCamera);
    auto cameraToClip = glm::perspective(deg2rad * c.camera.fov, aspectRatio, 1.f, 1000.f);
    auto NDCToScreen = glm::scale(glm::mat4(1.f), v3f(c.width, c.height, 1.0f)) * glm::scale(glm::mat4(1.f), v3f(0.5f, -0.5f, 1.f)) * glm::translate(glm::mat4(1.f), v3f(1.f, -1.f, 0.f));
    std::memcpy(out64, glm::value_ptr(worldToCamera), 64);
    std::memcpy(out64 + 16, glm::value_ptr(cameraToWorld), 64);
    std::memcpy(out64 + 32, glm::value_ptr(cameraToClip), 64);
    std::memcpy(out64 + 48, glm::value_ptr(NDCToScreen), 64);
}

/* Whole samples through ANY offline integrator of the reference (Integrator::render is the virtual the pixel loop calls,
 * renderer.cpp:194): used for PathTracerIntegrator (src/integrators/path.h:15-262), which returns everything in Li. */
void refh_integrator_paths(void* hp, const float* rays, const int* seeds, int64_t n, float* Li, int* nDraws) {
    auto* h = static_cast<Handle*>(hp);
    const Integrator* in = h->renderer->integrator.get();
    for (int64_t i = 0; i < n; i++) {
        Ray r = mkRay(rays + 8 * i);
        Sampler s(seeds[i]);
        v3f L = in->render(r, s);
        Li[3 * i] = L.x; Li[3 * i + 1] = L.y; Li[3 * i + 2] = L.z;
        if (nDraws) {
            Sampler replay(seeds[i]);
            int c = 0;
            while (!(replay.g == s.g) && c < 1000000) { replay.next(); c++; }
            nDraws[i] = c;
        }
    }
}

/* Scene::getShapeCenter / getShapeRadius (renderer.cpp:349-358) per shape: centre xyz + radius - DirectIntegrator's sphere emitters */
void refh_shape_spheres(void* hp, float* out4) {
    auto* h = static_cast<Handle*>(hp);
    const Scene& sc = h->renderer->scene;
    for (size_t i = 0; i < sc.worldData.shapes.size(); i++) {
        v3f c = sc.getShapeCenter(i);
        out4[4 * i] = c.x; out4[4 * i + 1] = c.y; out4[4 * i + 2] = c.z; out4[4 * i + 3] = sc.getShapeRadius(i);
    }
}

/* Whole-frame render through Renderer::render (the reference's own threaded pixel loop).
 * Returns the seconds spent inside render(); rgb_out = W*H*3 floats. */
double refh_render(void* hp, float* rgb_out) {
    auto* h = static_cast<Handle*>(hp);
    RenderBuffer& rb = *h->renderer->integrator->rgb;
    rb.clear();
    auto t0 = std::chrono::high_resolution_clock::now();
    h->renderer->render();
    auto t1 = std::chrono::high_resolution_clock::now();
    if (rgb_out) {
        const int np = rb.width * rb.height;
        for (int p = 0; p < np; p++) { rgb_out[3 * p] = rb.data[p].x; rgb_out[3 * p + 1] = rb.data[p].y; rgb_out[3 * p + 2] = rb.data[p].z; }
    }
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
