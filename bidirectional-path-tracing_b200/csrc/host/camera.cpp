// camera.cpp -- camera matrices with the same operation order as the glm 0.9.9 calls the reference makes
// (src/core/renderer.cpp:143-153, src/integrators/bdpt.h:49-55,485-490), so that camera rays and splat
// pixel coordinates are reproduced to the last bit.  Column-major 4x4, m[col*4+row].
#include <cmath>
#include <cstring>

#include "tr_host.h"

namespace trb {
namespace {

struct M4 { float m[16]; float& at(int c, int r) { return m[c * 4 + r]; } float at(int c, int r) const { return m[c * 4 + r]; } };

M4 identity() { M4 r; std::memset(&r, 0, sizeof(r)); r.at(0, 0) = r.at(1, 1) = r.at(2, 2) = r.at(3, 3) = 1.f; return r; }

inline float dot3(const float* a, const float* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
inline void cross3(const float* x, const float* y, float* o) {
    o[0] = x[1] * y[2] - y[1] * x[2];
    o[1] = x[2] * y[0] - y[2] * x[0];
    o[2] = x[0] * y[1] - y[0] * x[1];
}
inline void normalize3(float* v) {  // v * (1 / sqrt(dot(v,v)))
    float inv = 1.f / std::sqrt(dot3(v, v));
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}

// column j of A*B = ((A0*b0 + A1*b1) + A2*b2) + A3*b3
M4 mul(const M4& A, const M4& B) {
    M4 R;
    for (int j = 0; j < 4; ++j)
        for (int r = 0; r < 4; ++r)
            R.at(j, r) = ((A.at(0, r) * B.at(j, 0) + A.at(1, r) * B.at(j, 1)) + A.at(2, r) * B.at(j, 2)) + A.at(3, r) * B.at(j, 3);
    return R;
}

M4 look_at_rh(const float* eye, const float* center, const float* up) {
    float f[3] = {center[0] - eye[0], center[1] - eye[1], center[2] - eye[2]};
    normalize3(f);
    float s[3]; cross3(f, up, s); normalize3(s);
    float u[3]; cross3(s, f, u);
    M4 R = identity();
    R.at(0, 0) = s[0]; R.at(1, 0) = s[1]; R.at(2, 0) = s[2];
    R.at(0, 1) = u[0]; R.at(1, 1) = u[1]; R.at(2, 1) = u[2];
    R.at(0, 2) = -f[0]; R.at(1, 2) = -f[1]; R.at(2, 2) = -f[2];
    R.at(3, 0) = -dot3(s, eye); R.at(3, 1) = -dot3(u, eye); R.at(3, 2) = dot3(f, eye);
    return R;
}

M4 perspective_rh_no(float fovy, float aspect, float zn, float zf) {
    const float th = std::tan(fovy / 2.f);
    M4 R; std::memset(&R, 0, sizeof(R));
    R.at(0, 0) = 1.f / (aspect * th);
    R.at(1, 1) = 1.f / th;
    R.at(2, 2) = -(zf + zn) / (zf - zn);
    R.at(2, 3) = -1.f;
    R.at(3, 2) = -(2.f * zf * zn) / (zf - zn);
    return R;
}

M4 scale_of(const M4& m, float x, float y, float z) {
    M4 R;
    const float v[3] = {x, y, z};
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 4; ++r) R.at(c, r) = m.at(c, r) * v[c];
    for (int r = 0; r < 4; ++r) R.at(3, r) = m.at(3, r);
    return R;
}

M4 translate_of(const M4& m, float x, float y, float z) {
    M4 R = m;
    for (int r = 0; r < 4; ++r) R.at(3, r) = ((m.at(0, r) * x + m.at(1, r) * y) + m.at(2, r) * z) + m.at(3, r);
    return R;
}

// cofactor expansion in the order of glm::inverse for mat4
M4 inverse(const M4& M) {
    auto m = [&](int c, int r) { return M.at(c, r); };
    float c00 = m(2, 2) * m(3, 3) - m(3, 2) * m(2, 3), c02 = m(1, 2) * m(3, 3) - m(3, 2) * m(1, 3), c03 = m(1, 2) * m(2, 3) - m(2, 2) * m(1, 3);
    float c04 = m(2, 1) * m(3, 3) - m(3, 1) * m(2, 3), c06 = m(1, 1) * m(3, 3) - m(3, 1) * m(1, 3), c07 = m(1, 1) * m(2, 3) - m(2, 1) * m(1, 3);
    float c08 = m(2, 1) * m(3, 2) - m(3, 1) * m(2, 2), c10 = m(1, 1) * m(3, 2) - m(3, 1) * m(1, 2), c11 = m(1, 1) * m(2, 2) - m(2, 1) * m(1, 2);
    float c12 = m(2, 0) * m(3, 3) - m(3, 0) * m(2, 3), c14 = m(1, 0) * m(3, 3) - m(3, 0) * m(1, 3), c15 = m(1, 0) * m(2, 3) - m(2, 0) * m(1, 3);
    float c16 = m(2, 0) * m(3, 2) - m(3, 0) * m(2, 2), c18 = m(1, 0) * m(3, 2) - m(3, 0) * m(1, 2), c19 = m(1, 0) * m(2, 2) - m(2, 0) * m(1, 2);
    float c20 = m(2, 0) * m(3, 1) - m(3, 0) * m(2, 1), c22 = m(1, 0) * m(3, 1) - m(3, 0) * m(1, 1), c23 = m(1, 0) * m(2, 1) - m(2, 0) * m(1, 1);
    const float F0[4] = {c00, c00, c02, c03}, F1[4] = {c04, c04, c06, c07}, F2[4] = {c08, c08, c10, c11};
    const float F3[4] = {c12, c12, c14, c15}, F4[4] = {c16, c16, c18, c19}, F5[4] = {c20, c20, c22, c23};
    const float V0[4] = {m(1, 0), m(0, 0), m(0, 0), m(0, 0)}, V1[4] = {m(1, 1), m(0, 1), m(0, 1), m(0, 1)};
    const float V2[4] = {m(1, 2), m(0, 2), m(0, 2), m(0, 2)}, V3[4] = {m(1, 3), m(0, 3), m(0, 3), m(0, 3)};
    const float SA[4] = {+1, -1, +1, -1}, SB[4] = {-1, +1, -1, +1};
    M4 I;
    for (int k = 0; k < 4; ++k) {
        I.at(0, k) = ((V1[k] * F0[k] - V2[k] * F1[k]) + V3[k] * F2[k]) * SA[k];
        I.at(1, k) = ((V0[k] * F0[k] - V2[k] * F3[k]) + V3[k] * F4[k]) * SB[k];
        I.at(2, k) = ((V0[k] * F1[k] - V1[k] * F3[k]) + V3[k] * F5[k]) * SA[k];
        I.at(3, k) = ((V0[k] * F2[k] - V1[k] * F4[k]) + V2[k] * F5[k]) * SB[k];
    }
    const float d0 = m(0, 0) * I.at(0, 0), d1 = m(0, 1) * I.at(1, 0), d2 = m(0, 2) * I.at(2, 0), d3 = m(0, 3) * I.at(3, 0);
    const float det = (d0 + d1) + (d2 + d3);
    const float inv = 1.f / det;
    for (int k = 0; k < 16; ++k) I.m[k] *= inv;
    return I;
}

}  // namespace

void make_camera(const trb_render_desc& d, CameraDev& out) {
    std::memset(&out, 0, sizeof(out));
    const float kPi = 3.14159265358979323846f;  // platform.h:51 (M_PI redefined as float)
    const float W = float(d.width), H = float(d.height);
    for (int k = 0; k < 3; ++k) out.eye[k] = d.eye[k];
    out.tan_half_fov = std::tan(kPi / 180.f * d.fov_deg * 0.5f);  // renderer.cpp:149 (deg2rad expands to M_PI / 180.f)
    out.aspect = W / H;                                            // renderer.cpp:150
    out.forward[0] = d.at[0] - d.eye[0]; out.forward[1] = d.at[1] - d.eye[1]; out.forward[2] = d.at[2] - d.eye[2];
    normalize3(out.forward);                                       // bdpt.h:49
    M4 w2c = look_at_rh(d.eye, d.at, d.up);
    M4 c2w = inverse(w2c);
    M4 clip = perspective_rh_no(kPi / 180.f * d.fov_deg, out.aspect, 1.f, 1000.f);
    M4 n2s = mul(mul(scale_of(identity(), W, H, 1.0f), scale_of(identity(), 0.5f, -0.5f, 1.f)), translate_of(identity(), 1.f, -1.f, 0.f));
    std::memcpy(out.w2c, w2c.m, sizeof(out.w2c));
    std::memcpy(out.c2w, c2w.m, sizeof(out.c2w));
    std::memcpy(out.clip, clip.m, sizeof(out.clip));
    std::memcpy(out.ndc2screen, n2s.m, sizeof(out.ndc2screen));
    out.inv_w = 1.f / W; out.inv_h = 1.f / H;
    out.vnear_dist = (1.f / std::tan(kPi / 180.f * d.fov_deg * 0.5f)) * H * 0.5f;  // bdpt.h:52
    out.width = d.width; out.height = d.height;
}

}  // namespace trb
