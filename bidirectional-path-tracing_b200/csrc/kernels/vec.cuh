// vec.cuh -- fp32 vector helpers whose operation ORDER matches the glm 0.9.9 calls made by the reference
// (dot = (x+y)+z, cross as glm::cross, normalize = v * (1/sqrt(dot))), so that together with -fmad=false the
// device reproduces the reference's un-contracted x86-64 arithmetic bit for bit wherever no
// transcendental function is involved.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

namespace trbk {

struct V3 {
    float x, y, z;
};

__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 mk(float s) { return V3{s, s, s}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 x, V3 y) {
    return mk(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
__device__ __forceinline__ float length2(V3 a) { return dot(a, a); }
__device__ __forceinline__ V3 normalize(V3 v) { return v * (1.f / sqrtf(dot(v, v))); }
__device__ __forceinline__ bool is_zero(V3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }  // v == v3f(0.f)
__device__ __forceinline__ float luminance(V3 c) { return dot(c, mk(0.212671f, 0.715160f, 0.072169f)); }  // math.h:56-58

// Frame (core.h:152-167) built by coordinateSystem (math.h:42-51): s = cross(t, n)
struct Frame {
    V3 s, t, n;
};
__device__ __forceinline__ Frame make_frame(V3 a) {
    Frame f;
    f.n = a;
    if (fabsf(a.x) > fabsf(a.y)) {
        float invLen = 1.f / sqrtf(a.x * a.x + a.z * a.z);
        f.t = mk(a.z * invLen, 0.f, -a.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(a.y * a.y + a.z * a.z);
        f.t = mk(0.f, a.z * invLen, -a.y * invLen);
    }
    f.s = cross(f.t, a);
    return f;
}
__device__ __forceinline__ V3 to_local(const Frame& f, V3 v) { return mk(dot(v, f.s), dot(v, f.t), dot(v, f.n)); }
__device__ __forceinline__ V3 to_world(const Frame& f, V3 v) { return (f.s * v.x + f.t * v.y) + f.n * v.z; }

// barycentric (math.h:19-22): a*(1-u-v) + b*u + c*v
__device__ __forceinline__ V3 bary(V3 a, V3 b, V3 c, float u, float v) { return (a * (1 - u - v) + b * u) + c * v; }

// column-major mat4 * vec4 in glm's order: (m0*x + m1*y) + (m2*z + m3*w)
__device__ __forceinline__ float4 mat_mul(const float* m, float4 v) {
    float4 r;
    r.x = (m[0] * v.x + m[4] * v.y) + (m[8] * v.z + m[12] * v.w);
    r.y = (m[1] * v.x + m[5] * v.y) + (m[9] * v.z + m[13] * v.w);
    r.z = (m[2] * v.x + m[6] * v.y) + (m[10] * v.z + m[14] * v.w);
    r.w = (m[3] * v.x + m[7] * v.y) + (m[11] * v.z + m[15] * v.w);
    return r;
}

#define TRB_PI 3.14159265358979323846f
#define TRB_INV_PI 0.31830988618379067154f
#define TRB_INV_TWOPI 0.15915494309189533577f

}  // namespace trbk
