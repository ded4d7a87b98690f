// oracle/ref_b200_binding.h -- INTEGRATION.md section B, as code: the stub a maintainer of the reference adds to
// src/integrators/bdpt.h so that BDPTIntegrator routes the whole offline frame through libtrb.so (include/trb.h).
//
// `make -C oracle ref_b200` compiles the UNMODIFIED reference sources with three edits applied to copies in the build directory
// (never in /root/reference, never in this repo):
//   1. src/core/integrator.h   + `virtual bool renderFrame() { return false; }`            (Integrator, integrator.h:23-70)
//   2. src/core/renderer.cpp   + `if (integrator->renderFrame()) return;` at the top of the offline branch (renderer.cpp:130-140)
//   3. src/integrators/bdpt.h  + `bool renderFrame() override;` in BDPTIntegrator and this file included at its end (bdpt.h:219-241)
// and links the result with libtrb.so -> oracle/_ref/tinyrender_ref_b200: the reference's own main(), loadTOML, tinyobj loader,
// Scene::load and saveEXR around the B200 render.  tests/test_gpu_image.py::test_patched_reference_renders_on_the_device runs it.
//
// This file is included INSIDE the reference's TR_NAMESPACE (after struct BDPTIntegrator), so the reference's types are in scope.
// It is test infrastructure (it exists to prove the drop-in claim); the product does not include it.
#pragma once
extern "C" {
#include "trb.h"
}
#include <cstdlib>
#include <cstring>

inline bool BDPTIntegrator::renderFrame() {
    static trb_context* ctx = nullptr;
    const WorldData& wd = scene.worldData;
    if (!ctx) {
        if (trb_create(0, &ctx) != TRB_OK) {   // no sm_100 device: keep the reference's CPU loop
            std::cerr << "trb_create: " << trb_last_error(nullptr) << " - falling back to the CPU pixel loop" << std::endl;
            return false;
        }
        // shape-major triangles, the order AcceleratorBVH::build creates its objects in (accel.h:116-120)
        std::vector<int32_t> tv, tn, tm, ts;
        for (size_t s = 0; s < wd.shapes.size(); ++s) {
            const tinyobj::mesh_t& m = wd.shapes[s].mesh;
            for (size_t f = 0; f < m.indices.size() / 3; ++f) {
                for (int k = 0; k < 3; ++k) { tv.push_back(m.indices[3 * f + k].vertex_index); tn.push_back(m.indices[3 * f + k].normal_index); }
                tm.push_back(m.material_ids[f]); ts.push_back(int32_t(s));
            }
        }
        std::vector<trb_material> mats(wd.materials.size());
        for (size_t i = 0; i < mats.size(); ++i) {
            const tinyobj::material_t& t = wd.materials[i];
            std::memset(&mats[i], 0, sizeof(trb_material));
            for (int k = 0; k < 3; ++k) {
                mats[i].kd[k] = t.diffuse[k]; mats[i].ks[k] = t.specular[k]; mats[i].ke[k] = t.emission[k]; mats[i].tf[k] = t.transmittance[k];
            }
            mats[i].ns = t.shininess; mats[i].ni = t.ior; mats[i].illum = t.illum;
        }
        trb_scene_desc sd;
        std::memset(&sd, 0, sizeof(sd));
        sd.positions = wd.attrib.vertices.data(); sd.n_positions = wd.attrib.vertices.size() / 3;
        sd.normals = wd.attrib.normals.data(); sd.n_normals = wd.attrib.normals.size() / 3;
        sd.tri_v = tv.data(); sd.tri_n = tn.data(); sd.tri_mat = tm.data(); sd.tri_shape = ts.data(); sd.n_tris = tm.size();
        sd.materials = mats.data(); sd.n_materials = mats.size(); sd.n_shapes = wd.shapes.size();
        if (trb_scene_upload(ctx, &sd) != TRB_OK) { std::cerr << "trb_scene_upload: " << trb_last_error(ctx) << std::endl; return false; }
    }
    const Config& c = scene.config;
    trb_render_desc d;
    std::memset(&d, 0, sizeof(d));
    for (int k = 0; k < 3; ++k) { d.eye[k] = c.camera.o[k]; d.at[k] = c.camera.at[k]; d.up[k] = c.camera.up[k]; }
    d.fov_deg = c.camera.fov; d.width = c.width; d.height = c.height; d.spp = c.spp;
    d.rr_depth = m_rrDepth;
    d.seed = 260450963u;
    if (const char* s = std::getenv("TRB_SEED")) d.seed = uint32_t(std::strtoul(s, nullptr, 10));
    trb_stats st;
    if (trb_render_bdpt(ctx, &d, 0, &rgb->data[0].x, &st) != TRB_OK) {   // RenderBuffer is packed v3f, row 0 = top (core.h:128-146)
        std::cerr << "trb_render_bdpt: " << trb_last_error(ctx) << std::endl;
        return false;
    }
    std::cout << "B200 path: " << st.samples << " samples in " << st.ms_total << " ms on the device, " << st.launches << " kernel launches" << std::endl;
    return true;
}
