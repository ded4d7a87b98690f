// obj_loader.cpp -- Wavefront OBJ/MTL reader producing exactly what the reference's loader produces
// for Scene::load (src/core/renderer.cpp:249: tinyobj::LoadObj(..., triangulate = true), tinyobj v1.2.0).
//
// Bit-compatibility matters: vertex floats, triangle order and per-face material ids feed the
// reference-order BVH build and the bit-exact intersection parity.  Three behaviours of the reference's
// parser are therefore restated deliberately:
//   * decimal -> double conversion accumulates digit by digit (integer part *10, fraction digit * 10^-k)
//     and is then narrowed to float (tiny_obj_loader.h:525-648) - not strtod;
//   * polygons with > 3 vertices are ear-clipped in the projection plane picked from the first
//     non-degenerate corner (tiny_obj_loader.h:1043-1239) - not fanned;
//   * `o` and `g` start a new shape, `usemtl` only switches the material of following faces
//     (tiny_obj_loader.h:1861-1990); shapes without faces are dropped.
// tests/test_host.py (test_reference_loader_parity_if_available, test_obj_loader_semantics) and tools/gen_golden.py pin this file
// against the reference's tinyobj output for every shipped scene.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

#include "tr_host.h"

namespace trb {
namespace {

inline bool is_space(char c) { return c == ' ' || c == '\t'; }
inline bool is_digit(char c) { return c >= '0' && c <= '9'; }

// digit-accumulating decimal parser (see header comment); returns false when [s,e) is not a number
bool parse_double_compat(const char* s, const char* e, double* out) {
    if (s >= e) return false;
    double m = 0.0;
    int exp10 = 0;
    bool neg = false, eneg = false;
    const char* c = s;
    if (*c == '+' || *c == '-') { neg = (*c == '-'); ++c; }
    else if (!is_digit(*c)) return false;
    int nread = 0;
    while (c != e && is_digit(*c)) { m *= 10; m += int(*c - '0'); ++c; ++nread; }
    if (nread == 0) return false;
    if (c != e) {
        bool go_exp = false;
        if (*c == '.') {
            ++c;
            int k = 1;
            static const double lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
            while (c != e && is_digit(*c)) {
                m += int(*c - '0') * (k < 8 ? lut[k] : std::pow(10.0, -k));
                ++k; ++c;
            }
            go_exp = (c != e);
        } else if (*c == 'e' || *c == 'E') {
            go_exp = true;
        }
        if (go_exp && (*c == 'e' || *c == 'E')) {
            ++c;
            if (c != e && (*c == '+' || *c == '-')) { eneg = (*c == '-'); ++c; }
            else if (c == e || !is_digit(*c)) return false;
            int n = 0;
            while (c != e && is_digit(*c)) { exp10 = exp10 * 10 + int(*c - '0'); ++c; ++n; }
            if (eneg) exp10 = -exp10;
            if (n == 0) return false;
        }
    }
    double v = exp10 ? std::ldexp(m * std::pow(5.0, exp10), exp10) : m;
    *out = neg ? -v : v;
    return true;
}

struct Cursor {
    const char* p;
    void skip() { while (is_space(*p)) ++p; }
    float real(double dflt = 0.0) {
        skip();
        const char* e = p + std::strcspn(p, " \t\r");
        double v = dflt;
        parse_double_compat(p, e, &v);
        p = e;
        return float(v);
    }
    int integer() {
        skip();
        const char* e = p + std::strcspn(p, " \t\r");
        int v = std::atoi(p);
        p = e;
        return v;
    }
    bool at_eol() const { return *p == '\0' || *p == '\r' || *p == '\n'; }
};

struct Corner { int v = -1, vt = -1, vn = -1; };

// OBJ indices are 1-based; negative ones count from the end; 0 is invalid
bool fix_index(int idx, int n, int* out) {
    if (idx > 0) { *out = idx - 1; return true; }
    if (idx == 0) return false;
    *out = n + idx;
    return true;
}

bool parse_corner(Cursor& c, int nv, int nvn, int nvt, Corner* out) {
    Corner r;
    if (!fix_index(std::atoi(c.p), nv, &r.v)) return false;
    c.p += std::strcspn(c.p, "/ \t\r");
    if (*c.p != '/') { *out = r; return true; }
    ++c.p;
    if (*c.p == '/') {  // v//vn
        ++c.p;
        if (!fix_index(std::atoi(c.p), nvn, &r.vn)) return false;
        c.p += std::strcspn(c.p, "/ \t\r");
        *out = r;
        return true;
    }
    if (!fix_index(std::atoi(c.p), nvt, &r.vt)) return false;  // v/vt
    c.p += std::strcspn(c.p, "/ \t\r");
    if (*c.p != '/') { *out = r; return true; }
    ++c.p;  // v/vt/vn
    if (!fix_index(std::atoi(c.p), nvn, &r.vn)) return false;
    c.p += std::strcspn(c.p, "/ \t\r");
    *out = r;
    return true;
}

// crossing-number point-in-polygon test on a triangle (the classic pnpoly formulation)
bool point_in_tri(const float* x, const float* y, float tx, float ty) {
    bool inside = false;
    for (int i = 0, j = 2; i < 3; j = i++) {
        if (((y[i] > ty) != (y[j] > ty)) && (tx < (x[j] - x[i]) * (ty - y[i]) / (y[j] - y[i]) + x[i])) inside = !inside;
    }
    return inside;
}

struct ShapeAcc {
    std::string name;
    std::vector<Corner> corners;  // 3 per triangle
    std::vector<int> mats;        // per triangle
};

// ear clipping of one polygon, appended to `shape` (see header comment for the reference behaviour)
void emit_face(const std::vector<Corner>& face, int material, const std::vector<float>& v, ShapeAcc& shape) {
    size_t np = face.size();
    if (np < 3) return;
    auto push_tri = [&](const Corner& a, const Corner& b, const Corner& c) {
        shape.corners.push_back(a); shape.corners.push_back(b); shape.corners.push_back(c);
        shape.mats.push_back(material);
    };
    // projection axes from the first corner whose cross product is non-degenerate
    size_t ax0 = 1, ax1 = 2;
    for (size_t k = 0; k < np; ++k) {
        size_t a = size_t(face[k % np].v), b = size_t(face[(k + 1) % np].v), c = size_t(face[(k + 2) % np].v);
        if (3 * a + 2 >= v.size() || 3 * b + 2 >= v.size() || 3 * c + 2 >= v.size()) continue;
        float e0x = v[3 * b] - v[3 * a], e0y = v[3 * b + 1] - v[3 * a + 1], e0z = v[3 * b + 2] - v[3 * a + 2];
        float e1x = v[3 * c] - v[3 * b], e1y = v[3 * c + 1] - v[3 * b + 1], e1z = v[3 * c + 2] - v[3 * b + 2];
        float cx = std::fabs(e0y * e1z - e0z * e1y);
        float cy = std::fabs(e0z * e1x - e0x * e1z);
        float cz = std::fabs(e0x * e1y - e0y * e1x);
        const float eps = std::numeric_limits<float>::epsilon();
        if (cx > eps || cy > eps || cz > eps) {
            if (!(cx > cy && cx > cz)) {
                ax0 = 0;
                if (cz > cx && cz > cy) ax1 = 1;
            }
            break;
        }
    }
    float area = 0.f;
    for (size_t k = 0; k < np; ++k) {
        size_t a = size_t(face[k % np].v), b = size_t(face[(k + 1) % np].v);
        if (3 * a + ax0 >= v.size() || 3 * a + ax1 >= v.size() || 3 * b + ax0 >= v.size() || 3 * b + ax1 >= v.size()) continue;
        float ax = v[3 * a + ax0], ay = v[3 * a + ax1], bx = v[3 * b + ax0], by = v[3 * b + ax1];
        area += (ax * by - ay * bx) * 0.5f;
    }
    std::vector<Corner> rem = face;
    size_t guess = 0;
    int rounds = 10;
    while (rem.size() > 3 && rounds > 0) {
        np = rem.size();
        if (guess >= np) { rounds -= 1; guess -= np; }
        Corner ind[3];
        float px[3], py[3];
        for (size_t k = 0; k < 3; k++) {
            ind[k] = rem[(guess + k) % np];
            size_t vi = size_t(ind[k].v);
            if (3 * vi + ax0 >= v.size() || 3 * vi + ax1 >= v.size()) { px[k] = 0.f; py[k] = 0.f; }
            else { px[k] = v[3 * vi + ax0]; py[k] = v[3 * vi + ax1]; }
        }
        float e0x = px[1] - px[0], e0y = py[1] - py[0], e1x = px[2] - px[1], e1y = py[2] - py[1];
        float cr = e0x * e1y - e0y * e1x;
        if (cr * area < 0.0f) { guess += 1; continue; }  // reflex corner
        bool overlap = false;
        for (size_t o = 3; o < np; ++o) {
            size_t idx = (guess + o) % np;
            if (idx >= rem.size()) continue;
            size_t ovi = size_t(rem[idx].v);
            if (3 * ovi + ax0 >= v.size() || 3 * ovi + ax1 >= v.size()) continue;
            if (point_in_tri(px, py, v[3 * ovi + ax0], v[3 * ovi + ax1])) { overlap = true; break; }
        }
        if (overlap) { guess += 1; continue; }
        push_tri(ind[0], ind[1], ind[2]);
        rem.erase(rem.begin() + long((guess + 1) % np));
    }
    if (rem.size() == 3) push_tri(rem[0], rem[1], rem[2]);
}

void default_material(trb_material& m) {
    std::memset(&m, 0, sizeof(m));
    m.ns = 1.f;  // tinyobj InitMaterial: shininess 1, ior 1, illum 0
    m.ni = 1.f;
    m.illum = 0;
}

// tinyobj's texture-name parse (tiny_obj_loader.h ParseTextureNameAndOption): options start with '-', the last plain token is the name
std::string texture_name(const char* t) {
    std::istringstream ss{std::string(t)};
    std::string w, name;
    while (ss >> w) {
        if (w[0] == '-') {  // every option of the MTL spec takes one value except -o/-s/-t (up to three) and -mm (two)
            int take = (w == "-o" || w == "-s" || w == "-t") ? 3 : (w == "-mm" ? 2 : 1);
            for (int k = 0; k < take; ++k) {
                std::streampos pos = ss.tellg();
                std::string v;
                if (!(ss >> v)) break;
                char* end = nullptr;
                std::strtod(v.c_str(), &end);
                const bool numeric = end && *end == '\0';
                if (take > 1 && k > 0 && !numeric) { ss.clear(); ss.seekg(pos); break; }  // optional trailing values
            }
            continue;
        }
        name = w;
    }
    return name;
}

bool load_mtl(const std::string& path, std::vector<trb_material>& mats, std::vector<std::string>& names,
              std::map<std::string, int>& index, std::vector<std::string>& tex_files, std::string& err) {
    std::ifstream in(path);
    if (!in) { err += "cannot open material file " + path + "\n"; return false; }
    trb_material cur; default_material(cur);
    std::string cur_name;
    std::string line;
    auto flush = [&]() {
        if (!cur_name.empty()) { index.insert({cur_name, int(mats.size())}); mats.push_back(cur); names.push_back(cur_name); }
    };
    while (std::getline(in, line)) {
        size_t last = line.find_last_not_of(" \t\r\n");
        if (last == std::string::npos) continue;
        line.resize(last + 1);
        Cursor c{line.c_str()};
        c.skip();
        const char* t = c.p;
        if (*t == '\0' || *t == '#') continue;
        auto key3 = [&](const char* k) { return t[0] == k[0] && t[1] == k[1] && is_space(t[2]); };
        if (!std::strncmp(t, "newmtl", 6) && is_space(t[6])) {
            flush();
            default_material(cur);
            cur_name = std::string(t + 7);
            continue;
        }
        if (key3("Kd")) { c.p = t + 2; cur.kd[0] = c.real(); cur.kd[1] = c.real(); cur.kd[2] = c.real(); continue; }
        if (key3("Ks")) { c.p = t + 2; cur.ks[0] = c.real(); cur.ks[1] = c.real(); cur.ks[2] = c.real(); continue; }
        if (key3("Ke")) { c.p = t + 2; cur.ke[0] = c.real(); cur.ke[1] = c.real(); cur.ke[2] = c.real(); continue; }
        if (key3("Kt") || key3("Tf")) { c.p = t + 2; cur.tf[0] = c.real(); cur.tf[1] = c.real(); cur.tf[2] = c.real(); continue; }
        if (key3("Ni")) { c.p = t + 2; cur.ni = c.real(); continue; }
        if (key3("Ns")) { c.p = t + 2; cur.ns = c.real(); continue; }
        if (!std::strncmp(t, "illum", 5) && is_space(t[5])) { c.p = t + 6; cur.illum = c.integer(); continue; }
        if ((!std::strncmp(t, "map_Kd", 6) || !std::strncmp(t, "map_Ks", 6)) && is_space(t[6])) {  // diffuse_texname / specular_texname
            const std::string fn = texture_name(t + 7);
            if (!fn.empty()) {
                size_t id = 0;
                while (id < tex_files.size() && tex_files[id] != fn) ++id;
                if (id == tex_files.size()) tex_files.push_back(fn);
                if (id >= 0x7fff) { err += "too many textures\n"; return false; }
                const int32_t code = int32_t(id) + 1;
                if (t[5] == 'd') cur.tex = (cur.tex & ~0xffff) | code;
                else cur.tex = (cur.tex & 0xffff) | (code << 16);
            }
            continue;
        }
        // Ka, d, Tr, Pr, ... : parsed by the reference but never read on the BDPT path
    }
    flush();
    return true;
}

}  // namespace

trb_scene_desc HostScene::desc() const {
    trb_scene_desc d;
    d.positions = positions.data(); d.n_positions = positions.size() / 3;
    d.normals = normals.data(); d.n_normals = normals.size() / 3;
    d.tri_v = tri_v.data(); d.tri_n = tri_n.data(); d.tri_mat = tri_mat.data(); d.tri_shape = tri_shape.data();
    d.n_tris = tri_mat.size();
    d.materials = materials.data(); d.n_materials = materials.size();
    d.n_shapes = n_shapes;
    texture_views.clear();
    for (const Texture& t : textures) texture_views.push_back(trb_texture{t.rgb.data(), t.w, t.h});
    d.texcoords = texcoords.data(); d.n_texcoords = texcoords.size() / 2;
    d.tri_t = tri_t.empty() ? nullptr : tri_t.data();
    d.textures = texture_views.data(); d.n_textures = texture_views.size();
    return d;
}

void load_ppm_texture(const std::string& path, HostScene::Texture& out) {
    out.file = path;
    auto pink = [&]() { out.w = 1; out.h = 1; out.rgb = {1.f, 0.f, 1.f}; };  // Tex::pink
    // Tex::load always reads `<stem>.ppm` (core.h:490-496)
    std::string ppm = path;
    {
        const size_t slash = ppm.find_last_of("/\\"), dot = ppm.find_last_of('.');
        if (dot != std::string::npos && (slash == std::string::npos || dot > slash + 1)) ppm.resize(dot);
        ppm += ".ppm";
    }
    FILE* f = std::fopen(ppm.c_str(), "rb");
    if (!f) { pink(); return; }
    int w = 0, h = 0, e = 0;
    if (std::fscanf(f, "%*s %d %d %d%*c", &w, &h, &e) < 3 || w <= 0 || h <= 0 || e <= 0) { std::fclose(f); pink(); return; }
    const size_t sz = size_t(w) * h * 3;
    std::vector<uint8_t> ct(sz, 0);
    if (std::fread(ct.data(), 1, sz, f) == 0) { std::fclose(f); pink(); return; }
    std::fclose(f);
    float lut[256];
    for (int b = 0; b < 256; ++b) lut[b] = std::pow(float(b) / float(e), 2.2f);  // Tex::pf, float overload
    out.w = w; out.h = h;
    out.rgb.resize(sz);
    for (size_t i = 0; i < sz; ++i) {  // Tex::fl: the vertically flipped image
        const size_t j = i / 3, x = j % size_t(w), y = j / size_t(w);
        out.rgb[i] = lut[ct[3 * ((size_t(h) - y - 1) * size_t(w) + x) + i % 3]];
    }
}

bool load_obj(const std::string& obj_path, HostScene& out, std::string& err) {
    std::ifstream in(obj_path);
    if (!in) { err = "cannot open " + obj_path; return false; }
    std::string base;
    {
        size_t slash = obj_path.find_last_of("/\\");
        base = slash == std::string::npos ? std::string() : obj_path.substr(0, slash + 1);
    }
    out = HostScene();
    std::vector<float>& vt = out.texcoords;
    std::vector<std::string> tex_files;
    std::map<std::string, int> mat_index;
    std::vector<ShapeAcc> shapes;
    ShapeAcc shape;
    std::vector<std::vector<Corner>> group;  // faces waiting for their material to be closed
    int material = -1;
    std::string name;

    auto flush_group = [&]() {
        for (auto& f : group) emit_face(f, material, out.positions, shape);
        bool had = !group.empty();
        group.clear();
        return had;
    };

    std::string line;
    while (std::getline(in, line)) {
        while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
        if (line.empty()) continue;
        Cursor c{line.c_str()};
        c.skip();
        const char* t = c.p;
        if (*t == '\0' || *t == '#') continue;
        if (t[0] == 'v' && is_space(t[1])) {
            c.p = t + 2;
            out.positions.push_back(c.real()); out.positions.push_back(c.real()); out.positions.push_back(c.real());
            continue;
        }
        if (t[0] == 'v' && t[1] == 'n' && is_space(t[2])) {
            c.p = t + 3;
            out.normals.push_back(c.real()); out.normals.push_back(c.real()); out.normals.push_back(c.real());
            continue;
        }
        if (t[0] == 'v' && t[1] == 't' && is_space(t[2])) {
            c.p = t + 3;
            vt.push_back(c.real()); vt.push_back(c.real());
            continue;
        }
        if (t[0] == 'f' && is_space(t[1])) {
            c.p = t + 2;
            c.skip();
            std::vector<Corner> face;
            while (!c.at_eol()) {
                Corner k;
                if (!parse_corner(c, int(out.positions.size() / 3), int(out.normals.size() / 3), int(vt.size() / 2), &k)) {
                    err = "failed to parse `f' line (zero index) in " + obj_path;
                    return false;
                }
                face.push_back(k);
                c.p += std::strspn(c.p, " \t\r");
            }
            group.push_back(std::move(face));
            continue;
        }
        if (!std::strncmp(t, "usemtl", 6) && is_space(t[6])) {
            std::string nm(t + 7);
            auto it = mat_index.find(nm);
            int id = it == mat_index.end() ? -1 : it->second;
            if (id != material) { flush_group(); material = id; }
            continue;
        }
        if (!std::strncmp(t, "mtllib", 6) && is_space(t[6])) {
            std::istringstream ss(std::string(t + 7));
            std::string fn;
            bool found = false;
            while (!found && std::getline(ss, fn, ' ')) {
                if (fn.empty()) continue;
                std::string e2;
                if (load_mtl(base + fn, out.materials, out.material_names, mat_index, tex_files, e2)) found = true;
                else if (e2.find("not supported") != std::string::npos) { err = e2; return false; }
            }
            if (!found) err += "WARN: failed to load material file(s)\n";
            continue;
        }
        if ((t[0] == 'g' || t[0] == 'o') && is_space(t[1])) {
            bool had = flush_group();
            // `g` keeps a shape that already holds faces; `o` keeps it when this flush produced some
            bool keep = (t[0] == 'g') ? !shape.corners.empty() : had;
            if (keep) { shape.name = name; shapes.push_back(shape); }
            shape = ShapeAcc();
            if (t[0] == 'o') {
                name = std::string(t + 2);
            } else {
                std::istringstream ss(std::string(t + 2));
                std::vector<std::string> names;
                std::string w;
                while (ss >> w) names.push_back(w);
                // the group's first name (the reference's parser re-reads the keyword itself as token 0
                // and keeps token 1, tiny_obj_loader.h:1939-1955)
                name = names.size() > 0 ? names[0] : std::string();
            }
            continue;
        }
        // `s`, `t`, anything else: irrelevant to the BDPT path
    }
    bool had = flush_group();
    if (had || !shape.corners.empty()) { shape.name = name; shapes.push_back(shape); }

    for (size_t j = 0; j < shapes.size(); ++j) {
        const ShapeAcc& s = shapes[j];
        out.shape_names.push_back(s.name);
        for (size_t f = 0; f < s.mats.size(); ++f) {
            for (int k = 0; k < 3; ++k) {
                out.tri_v.push_back(s.corners[3 * f + k].v);
                out.tri_n.push_back(s.corners[3 * f + k].vn);
                out.tri_t.push_back(s.corners[3 * f + k].vt);
            }
            out.tri_mat.push_back(s.mats[f]);
            out.tri_shape.push_back(int32_t(j));
        }
    }
    out.n_shapes = shapes.size();
    // BitmapTexture3f (core.h:528-543): a relative texture name is resolved against the OBJ file's directory
    for (const std::string& fn : tex_files) {
        HostScene::Texture t;
        load_ppm_texture((!fn.empty() && fn[0] == '/') ? fn : base + fn, t);
        out.textures.push_back(std::move(t));
    }
    if (out.tri_mat.empty()) { err = "no faces in " + obj_path; return false; }
    return true;
}

}  // namespace trb
