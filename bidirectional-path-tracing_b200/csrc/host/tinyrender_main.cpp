// tinyrender_main.cpp -- the reference's CLI (src/main.cpp:121-181) on the B200 host API:
//     tinyrender_b200 <scene.toml> [nogui]
// loadTOML -> Renderer{init, render, cleanUp}; prints "Render took: ... seconds." and writes <toml stem>.exr.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../../include/tinyrender/tinyrender.h"

static void run(const std::string& file, bool nogui) {
    TinyRender::Config config;
    bool isRealTime;
    try {
        isRealTime = TinyRender::loadTOML(config, file);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "Error while parsing scene file: %s\n", e.what());
        std::exit(EXIT_FAILURE);
    }
    if (const char* stop = std::getenv("TRB_STOP_AFTER_CHECKPOINTS")) config.stopAfterCheckpoints = std::atoi(stop);
    TinyRender::Renderer renderer(config);
    if (!renderer.init(isRealTime, nogui)) {
        std::fprintf(stderr, "Renderer initialisation failed\n");
        std::exit(EXIT_FAILURE);
    }
    auto t0 = std::chrono::steady_clock::now();
    renderer.render();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("Render took: %f seconds.\n", sec);
    const trb_stats* st = nullptr;
    if (auto* b = dynamic_cast<TinyRender::BDPTIntegrator*>(renderer.integrator.get())) st = &b->lastStats;
    else if (auto* pt = dynamic_cast<TinyRender::PathTracerIntegrator*>(renderer.integrator.get())) st = &pt->lastStats;
    else if (auto* di = dynamic_cast<TinyRender::DirectIntegrator*>(renderer.integrator.get())) st = &di->lastStats;
    if (st && st->samples)
        std::printf("device: %.3f ms, %.2f Msamples/s, %.1f rays/sample, %llu launches\n", st->ms_total,
                    double(st->samples) / (st->ms_total * 1e3), double(st->rays_closest + st->rays_any) / double(st->samples),
                    (unsigned long long)st->launches);
    renderer.cleanUp();
}

int main(int argc, char* argv[]) {
    if (argc != 2 && argc != 3) {
        std::fprintf(stderr, "Syntax: %s <scene.toml> [nogui]\n", argv[0]);
        return EXIT_FAILURE;
    }
    bool nogui = argc == 3 && std::strcmp(argv[2], "nogui") == 0;
    run(argv[1], nogui);
    return EXIT_SUCCESS;
}
