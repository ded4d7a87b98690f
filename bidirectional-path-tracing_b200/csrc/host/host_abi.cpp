// host_abi.cpp -- the trb_host_* half of include/trb.h: everything that runs without a GPU.
#include <cstring>
#include <string>

#include "tr_host.h"

struct trb_host_scene {
    trb::HostScene s;
};

namespace {
thread_local std::string g_host_err;
int fail(int code, const std::string& msg) { g_host_err = msg; return code; }
}  // namespace

extern "C" {

const char* trb_host_last_error(void) { return g_host_err.c_str(); }

int trb_host_load_obj(const char* obj_path, trb_host_scene** out) {
    if (!obj_path || !out) return fail(TRB_ERR_INVALID, "trb_host_load_obj: null argument");
    auto* h = new trb_host_scene();
    std::string err;
    if (!trb::load_obj(obj_path, h->s, err)) {
        delete h;
        *out = nullptr;
        return fail(err.find("not supported") != std::string::npos ? TRB_ERR_UNSUPPORTED : TRB_ERR_IO, err);
    }
    *out = h;
    return TRB_OK;
}

int trb_host_scene_desc(const trb_host_scene* h, trb_scene_desc* out) {
    if (!h || !out) return fail(TRB_ERR_INVALID, "trb_host_scene_desc: null argument");
    *out = h->s.desc();
    return TRB_OK;
}

int trb_host_shape_name(const trb_host_scene* h, size_t shape, char* buf, size_t buflen) {
    if (!h || !buf || buflen == 0 || shape >= h->s.shape_names.size()) return fail(TRB_ERR_INVALID, "trb_host_shape_name: bad argument");
    std::strncpy(buf, h->s.shape_names[shape].c_str(), buflen - 1);
    buf[buflen - 1] = '\0';
    return TRB_OK;
}

void trb_host_scene_free(trb_host_scene* h) { delete h; }

int trb_host_load_toml(const char* toml_path, trb_render_desc* out, char* obj_path_out, size_t obj_len, char* type_out,
                       size_t type_len, int* realtime_out) {
    if (!toml_path || !out) return fail(TRB_ERR_INVALID, "trb_host_load_toml: null argument");
    trb::TomlConfig cfg;
    std::string err;
    if (!trb::load_toml(toml_path, cfg, err)) return fail(TRB_ERR_IO, err);
    *out = cfg.desc;
    if (obj_path_out && obj_len) { std::strncpy(obj_path_out, cfg.obj_path.c_str(), obj_len - 1); obj_path_out[obj_len - 1] = '\0'; }
    if (type_out && type_len) { std::strncpy(type_out, cfg.type.c_str(), type_len - 1); type_out[type_len - 1] = '\0'; }
    if (realtime_out) *realtime_out = cfg.realtime ? 1 : 0;
    return TRB_OK;
}

int trb_host_load_toml_path(const char* toml_path, trb_path_desc* out) {
    if (!toml_path || !out) return fail(TRB_ERR_INVALID, "trb_host_load_toml_path: null argument");
    trb::TomlConfig cfg;
    std::string err;
    if (!trb::load_toml(toml_path, cfg, err)) return fail(TRB_ERR_IO, err);
    *out = cfg.path;
    return TRB_OK;
}

int trb_host_load_toml_direct(const char* toml_path, trb_direct_desc* out) {
    if (!toml_path || !out) return fail(TRB_ERR_INVALID, "trb_host_load_toml_direct: null argument");
    trb::TomlConfig cfg;
    std::string err;
    if (!trb::load_toml(toml_path, cfg, err)) return fail(TRB_ERR_IO, err);
    *out = cfg.direct;
    return TRB_OK;
}

int trb_host_save_checkpoint(const char* path, const trb_checkpoint_info* info, const float* rgb) {
    if (!path || !info) return fail(TRB_ERR_INVALID, "trb_host_save_checkpoint: null argument");
    std::string err;
    if (!trb::save_checkpoint(path, *info, rgb, err)) return fail(TRB_ERR_IO, err);
    return TRB_OK;
}

int trb_host_load_checkpoint(const char* path, trb_checkpoint_info* info_out, float* rgb_out, size_t capacity_floats) {
    if (!path || !info_out) return fail(TRB_ERR_INVALID, "trb_host_load_checkpoint: null argument");
    std::string err;
    if (!trb::load_checkpoint(path, *info_out, rgb_out, capacity_floats, err)) return fail(TRB_ERR_IO, err);
    return TRB_OK;
}

uint64_t trb_host_scene_hash(const trb_scene_desc* s) {
    if (!s) return 1;
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t bytes) {
        const unsigned char* b = static_cast<const unsigned char*>(p);
        // 8 bytes per step (FNV-1a over 64-bit words, tail by bytes): the geometry of a 10 M-triangle scene is ~0.5 GB
        size_t i = 0;
        for (; i + 8 <= bytes; i += 8) { uint64_t w; std::memcpy(&w, b + i, 8); h = (h ^ w) * 1099511628211ull; }
        for (; i < bytes; ++i) h = (h ^ b[i]) * 1099511628211ull;
    };
    const uint64_t counts[8] = {s->n_positions, s->n_normals, s->n_tris, s->n_materials, s->n_shapes, s->n_texcoords, s->n_textures, 0};
    mix(counts, sizeof(counts));
    if (s->positions) mix(s->positions, s->n_positions * 3 * sizeof(float));
    if (s->normals) mix(s->normals, s->n_normals * 3 * sizeof(float));
    if (s->tri_v) mix(s->tri_v, s->n_tris * 3 * sizeof(int32_t));
    if (s->tri_n) mix(s->tri_n, s->n_tris * 3 * sizeof(int32_t));
    if (s->tri_mat) mix(s->tri_mat, s->n_tris * sizeof(int32_t));
    if (s->tri_shape) mix(s->tri_shape, s->n_tris * sizeof(int32_t));
    if (s->materials) mix(s->materials, s->n_materials * sizeof(trb_material));
    if (s->texcoords) mix(s->texcoords, s->n_texcoords * 2 * sizeof(float));
    if (s->tri_t) mix(s->tri_t, s->n_tris * 3 * sizeof(int32_t));
    for (size_t t = 0; s->textures && t < s->n_textures; ++t) {
        const int32_t wh[2] = {s->textures[t].width, s->textures[t].height};
        mix(wh, sizeof(wh));
        if (s->textures[t].rgb && wh[0] > 0 && wh[1] > 0) mix(s->textures[t].rgb, size_t(wh[0]) * wh[1] * 3 * sizeof(float));
    }
    return h ? h : 1;
}

int trb_host_checkpoint_matches(const trb_checkpoint_info* a, const trb_checkpoint_info* b) {
    return (a && b && trb::checkpoint_matches(*a, *b)) ? 1 : 0;
}

int trb_host_save_png(const char* path, const float* rgb, int width, int height, float exposure) {
    if (!path) return fail(TRB_ERR_INVALID, "trb_host_save_png: null path");
    std::string err;
    if (!trb::save_png_srgb(path, rgb, width, height, exposure, err)) return fail(TRB_ERR_IO, err);
    return TRB_OK;
}

int trb_host_save_exr(const char* path, const float* rgb, int width, int height) {
    if (!path) return fail(TRB_ERR_INVALID, "trb_host_save_exr: null path");
    std::string err;
    if (!trb::save_exr_bgr_half(path, rgb, width, height, err)) return fail(TRB_ERR_IO, err);
    return TRB_OK;
}

int trb_host_bvh_order(const trb_scene_desc* s, int32_t* order_out, size_t n) {
    if (!s || !order_out || n != s->n_tris) return fail(TRB_ERR_INVALID, "trb_host_bvh_order: bad argument");
    trb::RefTree t;
    std::string err;
    if (!trb::build_ref_tree(*s, t, err)) return fail(TRB_ERR_INVALID, err);
    std::memcpy(order_out, t.order.data(), n * sizeof(int32_t));
    return TRB_OK;
}

int trb_host_emitters(const trb_scene_desc* s, int32_t* shape_out, float* area_out, float* radiance_out, size_t max_emitters,
                      float* cdf_out, size_t cdf_capacity) {
    if (!s) return fail(TRB_ERR_INVALID, "trb_host_emitters: null scene");
    std::vector<trb::DevMaterial> mats;
    std::vector<trb::DevEmitter> em;
    std::vector<trb::DevEmitterTri> et;
    std::vector<float> cdf;
    std::vector<int32_t> s2e;
    std::string err;
    if (!trb::derive_materials(*s, mats, err)) return fail(TRB_ERR_UNSUPPORTED, err);
    if (!trb::derive_emitters(*s, mats, em, et, cdf, s2e, err)) return fail(TRB_ERR_UNSUPPORTED, err);
    for (size_t i = 0; i < em.size() && i < max_emitters; ++i) {
        if (shape_out) shape_out[i] = em[i].shape;
        if (area_out) area_out[i] = em[i].area;
        if (radiance_out) for (int k = 0; k < 3; ++k) radiance_out[3 * i + k] = em[i].radiance[k];
    }
    if (cdf_out) std::memcpy(cdf_out, cdf.data(), std::min(cdf.size(), cdf_capacity) * sizeof(float));
    return int(em.size());
}

int trb_host_shape_spheres(const trb_scene_desc* s, float* out4, size_t n_shapes) {
    if (!s || !out4) return fail(TRB_ERR_INVALID, "trb_host_shape_spheres: null argument");
    if (n_shapes < s->n_shapes) return fail(TRB_ERR_INVALID, "trb_host_shape_spheres: buffer too small");
    for (size_t t = 0; t < s->n_tris; ++t)
        for (int k = 0; k < 3; ++k)
            if (s->tri_v[3 * t + k] < 0 || size_t(s->tri_v[3 * t + k]) >= s->n_positions) return fail(TRB_ERR_INVALID, "vertex index out of range");
    std::vector<float> sph;
    trb::shape_spheres(*s, sph);
    std::memcpy(out4, sph.data(), sph.size() * sizeof(float));
    return TRB_OK;
}

int trb_host_camera_matrices(const trb_render_desc* d, float* out64) {
    if (!d || !out64) return fail(TRB_ERR_INVALID, "trb_host_camera_matrices: null argument");
    trb::CameraDev c;
    trb::make_camera(*d, c);
    std::memcpy(out64, c.w2c, 64);
    std::memcpy(out64 + 16, c.c2w, 64);
    std::memcpy(out64 + 32, c.clip, 64);
    std::memcpy(out64 + 48, c.ndc2screen, 64);
    return TRB_OK;
}

}  // extern "C"
