// toml_exr.cpp -- the two file formats on either side of the hot path.
//   load_toml : the TOML scene file of the reference (loadTOML, src/main.cpp:22-116): tables [input],
//               [camera], [film], [renderer]; same keys, same defaults, same error for a bad `type`.
//   save_exr_bgr_half : saveEXR (src/core/utils.h:95-156): scan-line OpenEXR, no compression, channels
//               B, G, R stored as half, row 0 = top, no colour transform.
// Both are small self-contained implementations (the reference uses cpptoml and tinyexr).
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

#include "tr_host.h"

namespace trb {
namespace {

struct TomlValue {
    enum Kind { STR, INT, FLT, BOOL, ARR } kind = STR;
    std::string s;
    long long i = 0;
    double f = 0.0;
    bool b = false;
    std::vector<TomlValue> arr;
};
using TomlTable = std::map<std::string, TomlValue>;

std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos) return "";
    size_t b = s.find_last_not_of(" \t\r\n");
    return s.substr(a, b - a + 1);
}

// strips a trailing comment that is not inside a string
std::string strip_comment(const std::string& s) {
    bool in_str = false;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] == '"' && (i == 0 || s[i - 1] != '\\')) in_str = !in_str;
        if (s[i] == '#' && !in_str) return s.substr(0, i);
    }
    return s;
}

bool parse_scalar(const std::string& t, TomlValue& v) {
    if (t.empty()) return false;
    if (t[0] == '"') {
        if (t.size() < 2 || t.back() != '"') return false;
        v.kind = TomlValue::STR;
        v.s.clear();
        for (size_t i = 1; i + 1 < t.size(); ++i) {
            if (t[i] == '\\' && i + 2 < t.size()) { ++i; v.s += (t[i] == 'n' ? '\n' : t[i] == 't' ? '\t' : t[i]); }
            else v.s += t[i];
        }
        return true;
    }
    if (t == "true" || t == "false") { v.kind = TomlValue::BOOL; v.b = (t == "true"); return true; }
    bool is_float = t.find_first_of(".eE") != std::string::npos;
    std::string u;
    for (char c : t) if (c != '_') u += c;
    char* end = nullptr;
    errno = 0;
    if (is_float) { v.kind = TomlValue::FLT; v.f = std::strtod(u.c_str(), &end); }
    else { v.kind = TomlValue::INT; v.i = std::strtoll(u.c_str(), &end, 10); v.f = double(v.i); }
    return end && *end == '\0' && end != u.c_str();
}

bool parse_value(const std::string& t, TomlValue& v) {
    if (!t.empty() && t[0] == '[') {
        if (t.back() != ']') return false;
        v.kind = TomlValue::ARR;
        std::string inner = t.substr(1, t.size() - 2);
        std::stringstream ss(inner);
        std::string item;
        while (std::getline(ss, item, ',')) {
            item = trim(item);
            if (item.empty()) continue;
            TomlValue e;
            if (!parse_scalar(item, e)) return false;
            v.arr.push_back(e);
        }
        return true;
    }
    return parse_scalar(t, v);
}

bool parse_toml(const std::string& path, std::map<std::string, TomlTable>& out, std::string& err) {
    std::ifstream in(path);
    if (!in) { err = path + " could not be opened for parsing"; return false; }
    std::string line, table;
    int ln = 0;
    while (std::getline(in, line)) {
        ++ln;
        line = trim(strip_comment(line));
        if (line.empty()) continue;
        if (line[0] == '[') {
            if (line.back() != ']') { err = "bad table header at line " + std::to_string(ln); return false; }
            table = trim(line.substr(1, line.size() - 2));
            out[table];
            continue;
        }
        size_t eq = line.find('=');
        if (eq == std::string::npos) { err = "expected key = value at line " + std::to_string(ln); return false; }
        std::string key = trim(line.substr(0, eq));
        std::string val = trim(line.substr(eq + 1));
        // multi-line arrays
        while (!val.empty() && val[0] == '[' && val.back() != ']' && std::getline(in, line)) { ++ln; val += " " + trim(strip_comment(line)); val = trim(val); }
        TomlValue v;
        if (!parse_value(val, v)) { err = "bad value at line " + std::to_string(ln); return false; }
        out[table][key] = v;
    }
    return true;
}

const TomlValue* find(const TomlTable& t, const char* k) { auto it = t.find(k); return it == t.end() ? nullptr : &it->second; }

// cpptoml get_as<double>: accepts a float (and converts an integer)
double get_double(const TomlTable& t, const char* k, double dflt) {
    const TomlValue* v = find(t, k);
    if (!v) return dflt;
    if (v->kind == TomlValue::FLT) return v->f;
    if (v->kind == TomlValue::INT) return double(v->i);
    return dflt;
}
int get_int(const TomlTable& t, const char* k, int dflt) {
    const TomlValue* v = find(t, k);
    return (v && v->kind == TomlValue::INT) ? int(v->i) : dflt;
}
bool get_bool(const TomlTable& t, const char* k, bool dflt) {
    const TomlValue* v = find(t, k);
    return (v && v->kind == TomlValue::BOOL) ? v->b : dflt;
}
std::string get_str(const TomlTable& t, const char* k, const std::string& dflt) {
    const TomlValue* v = find(t, k);
    return (v && v->kind == TomlValue::STR) ? v->s : dflt;
}
// cpptoml get_array_of<double>: empty (=> default) unless EVERY element is a float
void get_vec3(const TomlTable& t, const char* k, const double (&dflt)[3], float* out) {
    const TomlValue* v = find(t, k);
    bool ok = v && v->kind == TomlValue::ARR && v->arr.size() >= 3;
    if (ok) for (auto& e : v->arr) if (e.kind != TomlValue::FLT) ok = false;
    for (int i = 0; i < 3; ++i) out[i] = float(ok ? v->arr[i].f : dflt[i]);
}

}  // namespace

bool load_toml(const std::string& path, TomlConfig& out, std::string& err) {
    std::map<std::string, TomlTable> doc;
    if (!parse_toml(path, doc, err)) return false;
    for (const char* req : {"input", "camera", "film", "renderer"}) {
        if (!doc.count(req)) { err = std::string("missing table [") + req + "]"; return false; }
    }
    std::memset(&out.desc, 0, sizeof(out.desc));
    const TomlTable& input = doc["input"];
    const TomlValue* obj = find(input, "objfile");
    if (!obj || obj->kind != TomlValue::STR) { err = "input.objfile missing"; return false; }
    // relative paths are resolved against the toml's directory (renderer.cpp:236-241)
    std::string objfile = obj->s;
    for (char& c : objfile) if (c == '\\') c = '/';
    if (!objfile.empty() && objfile[0] != '/') {
        size_t slash = path.find_last_of('/');
        objfile = (slash == std::string::npos ? std::string() : path.substr(0, slash + 1)) + objfile;
    }
    out.obj_path = objfile;

    const TomlTable& cam = doc["camera"];
    out.desc.fov_deg = float(get_double(cam, "fov", 30.));
    get_vec3(cam, "eye", {1., 1., 0.}, out.desc.eye);
    get_vec3(cam, "at", {0., 0., 0.}, out.desc.at);
    get_vec3(cam, "up", {0., 1., 0.}, out.desc.up);
    const TomlTable& film = doc["film"];
    out.desc.width = get_int(film, "width", 768);
    out.desc.height = get_int(film, "height", 576);
    const TomlTable& ren = doc["renderer"];
    out.realtime = get_bool(ren, "realtime", false);
    out.type = get_str(ren, "type", "normal");
    out.desc.rr_depth = get_int(ren, "rrDepth", 5);   // main.cpp:105
    out.desc.spp = get_int(ren, "spp", 1);            // main.cpp:112
    out.desc.spp_begin = 0; out.desc.spp_end = 0;
    out.desc.seed = uint32_t(get_int(ren, "seed", 260450963));  // renderer.cpp:155 (extension key)
    out.desc.flags = 0;
    // extension keys: the reference's compiled-out Russian roulette (bdpt.h:18 NO_RR) and the device's depth cap
    if (get_bool(ren, "russianRoulette", false)) out.desc.flags |= TRB_FLAG_RUSSIAN_ROULETTE;
    out.desc.max_depth = get_int(ren, "maxDepth", 0);
    out.desc.reserved = 0;
    out.desc.wave_samples = get_int(ren, "waveSamples", 0);  // extension keys: samples per wavefront, devices
    out.gpus = std::max(1, get_int(ren, "gpus", 1));
    // PathTracerIntegrator (main.cpp:94-102); rrDepth is shared with the BDPT branch (:98,105)
    out.path.is_explicit = get_bool(ren, "isExplicit", true) ? 1 : 0;
    out.path.max_depth = get_int(ren, "maxDepth", -1);
    out.path.rr_prob = float(get_double(ren, "rrProb", 0.95));
    out.path.emitter_samples = get_int(ren, "emitterSamples", 1);
    out.path.bsdf_samples = get_int(ren, "bsdfSamples", 0);
    // DirectIntegrator (main.cpp:88-93); its bsdfSamples default is 1 where the path tracer's is 0, and its samplingStrategy default,
    // "emitter", is a name DirectIntegrator::render does not know (direct.h:449-461)
    out.direct.emitter_samples = get_int(ren, "emitterSamples", 1);
    out.direct.bsdf_samples = get_int(ren, "bsdfSamples", 1);
    out.direct.reserved = 0;
    {
        const std::string st = get_str(ren, "samplingStrategy", "emitter");
        out.direct.strategy = st == "mis" ? TRB_DI_MIS : st == "area" ? TRB_DI_AREA : st == "solidAngle" ? TRB_DI_SOLID_ANGLE
                            : st == "cosineHemisphere" ? TRB_DI_COSINE_HEMISPHERE : st == "bsdf" ? TRB_DI_BSDF : TRB_DI_INVALID;
    }
    // restartable renders: file names relative to the toml's directory, like objfile
    auto resolve = [&](std::string f) {
        if (!f.empty() && f[0] != '/') { size_t slash = path.find_last_of('/'); f = (slash == std::string::npos ? std::string() : path.substr(0, slash + 1)) + f; }
        return f;
    };
    out.checkpoint = resolve(get_str(ren, "checkpoint", ""));
    out.preview = resolve(get_str(ren, "preview", ""));
    out.checkpoint_every = get_int(ren, "checkpointEvery", 0);
    return true;
}

// ---- EXR -------------------------------------------------------------------------------------------
namespace {
uint16_t float_to_half(float f) {
    uint32_t x; std::memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t e = (x >> 23) & 0xffu, m = x & 0x7fffffu;
    if (e == 255) return uint16_t(sign | 0x7c00u | (m ? 0x200u : 0u));
    if (e == 0) return uint16_t(sign);
    int ne = int(e) - 127 + 15;
    if (ne >= 31) return uint16_t(sign | 0x7c00u);
    if (ne <= 0) {
        if (14 - ne > 24) return uint16_t(sign);
        uint32_t mant = m | 0x800000u;
        uint32_t h = mant >> (14 - ne);
        if ((mant >> (13 - ne)) & 1u) ++h;
        return uint16_t(sign | h);
    }
    uint32_t h = (uint32_t(ne) << 10) | (m >> 13);
    if (m & 0x1000u) ++h;  // round half up on the dropped bit, may carry into the exponent
    return uint16_t(sign | h);
}
void put(std::vector<uint8_t>& b, const void* p, size_t n) { const uint8_t* c = static_cast<const uint8_t*>(p); b.insert(b.end(), c, c + n); }
void put_str(std::vector<uint8_t>& b, const char* s) { put(b, s, std::strlen(s) + 1); }
void put_i32(std::vector<uint8_t>& b, int32_t v) { put(b, &v, 4); }
void put_attr(std::vector<uint8_t>& b, const char* name, const char* type, const void* data, int32_t size) {
    put_str(b, name); put_str(b, type); put_i32(b, size); put(b, data, size_t(size));
}
}  // namespace

bool save_exr_bgr_half(const std::string& path, const float* rgb, int w, int h, std::string& err) {
    if (!rgb || w <= 0 || h <= 0) { err = "save_exr: bad arguments"; return false; }
    std::vector<uint8_t> hd;
    const uint8_t magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};
    put(hd, magic, 8);
    std::vector<uint8_t> ch;
    for (const char* name : {"B", "G", "R"}) {  // alphabetical = the (A)BGR order of utils.h:124-129
        put_str(ch, name);
        put_i32(ch, 1);  // HALF
        const uint8_t lin[4] = {0, 0, 0, 0};
        put(ch, lin, 4);
        put_i32(ch, 1); put_i32(ch, 1);
    }
    ch.push_back(0);
    put_attr(hd, "channels", "chlist", ch.data(), int32_t(ch.size()));
    const uint8_t comp = 0;  // NO_COMPRESSION (tinyexr's InitEXRHeader default, as the reference leaves it)
    put_attr(hd, "compression", "compression", &comp, 1);
    const int32_t win[4] = {0, 0, w - 1, h - 1};
    put_attr(hd, "dataWindow", "box2i", win, 16);
    put_attr(hd, "displayWindow", "box2i", win, 16);
    const uint8_t lo = 0;
    put_attr(hd, "lineOrder", "lineOrder", &lo, 1);
    const float par = 1.f;
    put_attr(hd, "pixelAspectRatio", "float", &par, 4);
    const float swc[2] = {0.f, 0.f};
    put_attr(hd, "screenWindowCenter", "v2f", swc, 8);
    const float sww = 1.f;
    put_attr(hd, "screenWindowWidth", "float", &sww, 4);
    hd.push_back(0);

    const size_t line_bytes = size_t(w) * 3 * 2;
    const uint64_t table_start = hd.size();
    const uint64_t data_start = table_start + uint64_t(h) * 8;
    std::vector<uint8_t> out;
    out.reserve(size_t(data_start) + size_t(h) * (8 + line_bytes));
    out = hd;
    for (int y = 0; y < h; ++y) { uint64_t off = data_start + uint64_t(y) * (8 + line_bytes); put(out, &off, 8); }
    std::vector<uint16_t> line(size_t(w) * 3);
    for (int y = 0; y < h; ++y) {
        put_i32(out, y);
        put_i32(out, int32_t(line_bytes));
        const float* row = rgb + size_t(y) * w * 3;
        for (int x = 0; x < w; ++x) {
            line[x] = float_to_half(row[3 * x + 2]);                  // B
            line[size_t(w) + x] = float_to_half(row[3 * x + 1]);      // G
            line[2 * size_t(w) + x] = float_to_half(row[3 * x + 0]);  // R
        }
        put(out, line.data(), line_bytes);
    }
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { err = "cannot write " + path; return false; }
    size_t nw = std::fwrite(out.data(), 1, out.size(), f);
    std::fclose(f);
    if (nw != out.size()) { err = "short write to " + path; return false; }
    return true;
}

}  // namespace trb
