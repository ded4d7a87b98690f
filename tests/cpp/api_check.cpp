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
