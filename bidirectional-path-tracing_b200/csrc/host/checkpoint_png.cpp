// checkpoint_png.cpp -- SURVEY 8(f) N3: restartable renders and a quick-look image.
//   * checkpoint = the running-sum frame buffer + how many samples per pixel it holds + the render description that made it.
//     Contributions are normalised by the TOTAL spp (bdpt.h:330-332, renderer.cpp:201) and RNG streams are keyed by
//     (seed, pixel, sample index), so rendering [spp_done, spp) with `accumulate` on top of a restored buffer gives the
//     image an uninterrupted run gives (up to the order of float atomics).
//   * PNG = 8-bit sRGB of the linear buffer (the reference only has this transfer curve in its realtime blit,
//     renderpass.cpp:397); zlib "stored" blocks, so no compression library is needed.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "tr_host.h"

namespace trb {
namespace {
const char kMagic[8] = {'T', 'R', 'B', 'C', 'K', 'P', 'T', '1'};
uint64_t fnv1a(const void* p, size_t n, uint64_t h = 1469598103934665603ull) {
    const unsigned char* b = static_cast<const unsigned char*>(p);
    for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}
uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(const unsigned char* p, size_t n, uint32_t c = 0xffffffffu) {
    if (!crc_ready) {
        for (uint32_t i = 0; i < 256; ++i) { uint32_t v = i; for (int k = 0; k < 8; ++k) v = (v & 1) ? 0xedb88320u ^ (v >> 1) : v >> 1; crc_table[i] = v; }
        crc_ready = true;
    }
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xff] ^ (c >> 8);
    return c;
}
void be32(std::vector<unsigned char>& o, uint32_t v) { o.push_back(v >> 24); o.push_back(v >> 16); o.push_back(v >> 8); o.push_back(v); }
void chunk(std::vector<unsigned char>& png, const char* type, const std::vector<unsigned char>& data) {
    be32(png, uint32_t(data.size()));
    size_t start = png.size();
    png.insert(png.end(), type, type + 4);
    png.insert(png.end(), data.begin(), data.end());
    be32(png, crc32(png.data() + start, png.size() - start) ^ 0xffffffffu);
}
}  // namespace

bool save_checkpoint(const std::string& path, const trb_checkpoint_info& info, const float* rgb, std::string& err) {
    if (!rgb || info.desc.width <= 0 || info.desc.height <= 0) { err = "save_checkpoint: bad argument"; return false; }
    const size_t n = size_t(info.desc.width) * info.desc.height * 3;
    const std::string tmp = path + ".tmp";   // write-then-rename: an interrupted save never destroys the previous checkpoint
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) { err = "cannot write " + tmp; return false; }
    uint64_t sum = fnv1a(&info, sizeof(info));
    sum = fnv1a(rgb, n * sizeof(float), sum);
    bool ok = std::fwrite(kMagic, 1, 8, f) == 8 && std::fwrite(&info, sizeof(info), 1, f) == 1 && std::fwrite(rgb, sizeof(float), n, f) == n &&
              std::fwrite(&sum, sizeof(sum), 1, f) == 1;
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { err = "short write to " + tmp; std::remove(tmp.c_str()); return false; }
    if (std::rename(tmp.c_str(), path.c_str()) != 0) { err = "cannot rename " + tmp; return false; }
    return true;
}

bool load_checkpoint(const std::string& path, trb_checkpoint_info& info, float* rgb, size_t capacity, std::string& err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot read " + path; return false; }
    char magic[8];
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, kMagic, 8) == 0 && std::fread(&info, sizeof(info), 1, f) == 1;
    if (!ok || info.desc.width <= 0 || info.desc.height <= 0) { std::fclose(f); err = path + ": not a checkpoint"; return false; }
    const size_t n = size_t(info.desc.width) * info.desc.height * 3;
    if (!rgb) { std::fclose(f); return true; }   // header only
    if (capacity < n) { std::fclose(f); err = "load_checkpoint: buffer too small"; return false; }
    uint64_t sum = 0;
    ok = std::fread(rgb, sizeof(float), n, f) == n && std::fread(&sum, sizeof(sum), 1, f) == 1;
    std::fclose(f);
    if (!ok) { err = path + ": truncated checkpoint"; return false; }
    uint64_t want = fnv1a(&info, sizeof(info));
    want = fnv1a(rgb, n * sizeof(float), want);
    if (want != sum) { err = path + ": checksum mismatch"; return false; }
    return true;
}

bool save_png_srgb(const std::string& path, const float* rgb, int w, int h, float exposure, std::string& err) {
    if (!rgb || w <= 0 || h <= 0) { err = "save_png: bad argument"; return false; }
    std::vector<unsigned char> raw;
    raw.reserve(size_t(h) * (size_t(w) * 3 + 1));
    for (int y = 0; y < h; ++y) {
        raw.push_back(0);  // filter: none
        for (int x = 0; x < w * 3; ++x) {
            float v = rgb[size_t(y) * w * 3 + x] * exposure;
            v = !(v > 0.f) ? 0.f : (v > 1.f ? 1.f : v);
            float s = v <= 0.0031308f ? 12.92f * v : 1.055f * std::pow(v, 1.f / 2.4f) - 0.055f;
            raw.push_back(static_cast<unsigned char>(s * 255.f + 0.5f));
        }
    }
    std::vector<unsigned char> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (unsigned char c : raw) { a = (a + c) % 65521u; b = (b + a) % 65521u; }
    for (size_t off = 0; off < raw.size(); off += 65535) {
        size_t len = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + len == raw.size() ? 1 : 0);
        z.push_back(len & 0xff); z.push_back(len >> 8); z.push_back(~len & 0xff); z.push_back((~len >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + len);
    }
    be32(z, (b << 16) | a);
    std::vector<unsigned char> png = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<unsigned char> ihdr;
    be32(ihdr, uint32_t(w)); be32(ihdr, uint32_t(h));
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(png, "IHDR", ihdr);
    chunk(png, "IDAT", z);
    chunk(png, "IEND", {});
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { err = "cannot write " + path; return false; }
    bool ok = std::fwrite(png.data(), 1, png.size(), f) == png.size();
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) err = "short write to " + path;
    return ok;
}

// Same render?  Everything that shapes the image must agree; the sample range, wave size and stream flags may differ.
bool checkpoint_matches(const trb_checkpoint_info& a, const trb_checkpoint_info& b) {
    const trb_render_desc &x = a.desc, &y = b.desc;
    const uint32_t image_flags = TRB_FLAG_NO_JITTER | TRB_FLAG_RUSSIAN_ROULETTE | TRB_FLAG_LIGHT_TRACING | TRB_FLAG_PATH_TRACING;
    bool same = std::memcmp(x.eye, y.eye, sizeof(x.eye)) == 0 && std::memcmp(x.at, y.at, sizeof(x.at)) == 0 && std::memcmp(x.up, y.up, sizeof(x.up)) == 0 &&
                x.fov_deg == y.fov_deg && x.width == y.width && x.height == y.height && x.spp == y.spp && x.rr_depth == y.rr_depth && x.seed == y.seed &&
                (x.flags & image_flags) == (y.flags & image_flags) && x.max_depth == y.max_depth && a.integrator == b.integrator && a.scene_hash == b.scene_hash;
    if (same && (a.integrator == 1 || a.integrator == 2)) same = std::memcmp(&a.path, &b.path, sizeof(a.path)) == 0;
    return same;
}
}  // namespace trb
