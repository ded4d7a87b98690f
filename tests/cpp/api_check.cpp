// tests/cpp/api_check.cpp -- drives TinyRender's C++ host API (include/tinyrender/tinyrender.h) the way the
// reference's own code does: loadTOML -> Renderer::init -> per-sample integrator->render(ray, sampler) with a private
// Sampler(seed) per path (what oracle/ref_harness.cpp does with the unmodified reference), then scene.bvh->intersect
// and BSDF::eval/pdf/sample on single items.
//   api_check <scene.toml> <in.bin> <out.bin>
// in.bin : int32 n, then n x {float ray[8], int32 seed};  out.bin: n x {float li[3], int32 hit, float t, float wo_z}
#include <cstdio>
#include <cstring>
#include <vector>

#include "tinyrender/tinyrender.h"

using namespace TinyRender;

int main(int argc, char** argv) {
    if (argc != 4) return 2;
    Config config;
    bool rt = loadTOML(config, argv[1]);
    Renderer renderer(config);
    if (!renderer.init(rt, true)) return 3;
    {   // the reference's own signatures (core.h:299-305, 352-374; integrator.h:37-60) on the loaded scene
        const Scene& sc = renderer.scene;
        struct MyDiffuse : BSDF {   // a BSDF subclass written against the reference's constructor
            MyDiffuse(const WorldData& d, const Config& c, size_t matID) : BSDF(d, c, matID) { combinedType = EDiffuseReflection; }
            std::string toString() const override { return "MyDiffuse"; }
        };
        MyDiffuse mine(sc.worldData, sc.config, 0);
        const Integrator& in = *renderer.integrator;
        if (&mine.worldData != &sc.worldData || &mine.config != &config || !(mine.emission == sc.bsdfs[0]->emission)) return 6;
        if (sc.emitters.empty()) return 6;
        const Emitter& em = in.getEmitterByID(0);
        if (in.getEmitterIDByShapeID(em.shapeID) != 0 || in.getEmitterPdf(em) != 1.f / sc.emitters.size()) return 6;
        float pdf = 0.f;
        if (in.selectEmitter(0.999999f, pdf) != sc.emitters.size() - 1 || pdf != 1.f / sc.emitters.size()) return 6;
        if (sc.getShapeArea(em.shapeID) != em.area || !(sc.getShapeRadius(em.shapeID) > 0.f) || sc.getFirstLight() != em.shapeID) return 6;
        SurfaceInteraction lightFace; lightFace.shapeID = em.shapeID; lightFace.primID = 0;
        const BSDF* lb = in.getBSDF(lightFace);
        lightFace.matID = int(lb->matID);
        if (!lb->isEmissive() || !(in.getEmission(lightFace) == em.radiance)) return 6;
        v3f c = sc.getShapeCenter(em.shapeID);
        std::printf("api_check: emitter shape %zu area %g centre (%g %g %g) radius %g\n", em.shapeID, double(em.area), double(c.x), double(c.y), double(c.z),
                    double(sc.getShapeRadius(em.shapeID)));
    }
    FILE* f = std::fopen(argv[2], "rb");
    if (!f) return 4;
    int32_t n = 0;
    if (std::fread(&n, 4, 1, f) != 1) return 4;
    std::vector<float> out;
    for (int i = 0; i < n; ++i) {
        float r[8]; int32_t seed;
        if (std::fread(r, 4, 8, f) != 8 || std::fread(&seed, 4, 1, f) != 1) return 4;
        Ray ray(v3f(r[0], r[1], r[2]), v3f(r[3], r[4], r[5]), r[6], r[7]);
        Sampler sampler(seed);
        v3f li = renderer.integrator->render(ray, sampler);
        SurfaceInteraction si;
        bool hit = renderer.scene.bvh->intersect(ray, si);
        out.push_back(li.x); out.push_back(li.y); out.push_back(li.z);
        int32_t h = hit ? 1 : 0; float hf; std::memcpy(&hf, &h, 4);
        out.push_back(hf); out.push_back(si.t); out.push_back(hit ? si.wo.z : 0.f);
        if (hit && i < 8) {  // single-item BSDF calls must agree with themselves through both entry points
            const BSDF* b = renderer.scene.bsdfs[size_t(si.matID)].get();
            float pdf = 0.f;
            v3f val = b->sample(si, v2f(0.3f, 0.6f), &pdf);
            v3f ev = b->eval(si);
            float p2 = b->pdf(si);
            bool delta = (b->getType() & BSDF::EDelta) != 0;
            if (!delta && (!(val == ev) || pdf != p2)) { std::fprintf(stderr, "BSDF single-item mismatch on %s\n", b->toString().c_str()); return 5; }
        }
    }
    std::fclose(f);
    f = std::fopen(argv[3], "wb");
    std::fwrite(out.data(), 4, out.size(), f);
    std::fclose(f);
    std::printf("api_check: %d paths, splat energy in rgb: %g\n", n, double(renderer.integrator->rgb->data[0].x));
    return 0;
}
