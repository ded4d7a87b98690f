// trace.cuh -- BVH traversal with the reference's exact hit semantics.
//
// Replaces BVH::getIntersection + BBox::intersect (externals/bvh.h:259-352, 33-69),
// AcceleratorBVH::BVHNode::getIntersection + rayTriangleIntersect (src/core/accel.h:27-52,
// src/core/core.h:379-400) and the range test of AcceleratorBVH::intersect (accel.h:131-133).
//
// Reference semantics: BBox::intersect never reports distances, so the reference visits
// EVERY node whose box the infinite line passes (decided by six IEEE divisions) and answers
//   closest : argmin (t, position in the build-reordered primitive list) over triangles with
//             (double)t > 1e-3 and t < ray.max_t, all of whose ancestors' boxes passed;
//   any-hit : exists such a triangle with t <= max_t (bvh.h:299-307 + :351).
// This traversal returns exactly that, but ordered and pruned:
//   * box decision: (b - o) * (1/d) is within 2 ulp of the reference's (b - o) / d; when the slab
//     overlap is decided by less than ~5 ulp the six divisions are redone exactly as bvh.h:33-69
//     (rare slow path), so the SET of visited subtrees is the reference's;
//   * pruning by entry distance uses a pad far above Moller-Trumbore's error for non-degenerate
//     configurations; a ray that meets an ill-conditioned candidate (grazing / origin in the triangle's
//     plane, where the reference's t is rounding noise) is re-traced exhaustively, i.e. with the
//     reference's own visiting rule;
//   * ties in t keep the lower list position (bvh.h:305 strict '<', first found wins).
// The functions are __host__ __device__ so that tools/host_trace_check.cpp can run the identical code
// against the compiled reference on the CPU box (development aid only; the product runs it on the GPU).
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define TRB_HD __host__ __device__ __forceinline__
#define TRB_HD_NOINLINE __host__ __device__ __noinline__
#else
#define TRB_HD inline
#define TRB_HD_NOINLINE inline
#endif

namespace trbk {

struct F4 { float x, y, z, w; };

struct RayT {
    float ox, oy, oz, dx, dy, dz, min_t, max_t;
};

struct HitT {
    float t, u, v;
    int32_t pos;  // position in the reordered triangle list, -1 = none
};

struct BvhView {
    const F4* nodes;  // 4 x F4 per node (DevNode)
    const F4* tris;   // 3 x F4 per triangle (DevTri)
    int32_t root_ref;
};

#ifndef TRB_STACK
#define TRB_STACK 64
#endif

TRB_HD int32_t f2i(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int32_t i; memcpy(&i, &f, 4); return i;
#endif
}

#if defined(__CUDA_ARCH__)
#define TRB_LDG4(p) __ldg(reinterpret_cast<const float4*>(p))
TRB_HD F4 ld4(const F4* p) { float4 v = TRB_LDG4(p); return F4{v.x, v.y, v.z, v.w}; }
TRB_HD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
TRB_HD float frcp(float a) { return __frcp_rn(a); }  // RN(1/a): the same value as fdiv(1.f, a), cheaper sequence
#else
TRB_HD F4 ld4(const F4* p) { return *p; }
TRB_HD float fdiv(float a, float b) { return a / b; }
TRB_HD float frcp(float a) { return 1.f / a; }
#endif

// BBox::intersect verbatim in behaviour (bvh.h:33-69), returning also the clipped interval for ordering.
// NaN comparisons are false exactly as in the reference.  Everything by value: pointer arguments to a
// non-inlined function would force the caller's boxes into local memory.
struct BoxR { float enter, exit_; bool hit; };
TRB_HD_NOINLINE BoxR box_exact(float mnx, float mny, float mnz, float mxx, float mxy, float mxz,
                               float ox, float oy, float oz, float dx, float dy, float dz) {
    BoxR o; o.enter = 0.f; o.exit_ = 0.f; o.hit = false;
    float tmin = fdiv(mnx - ox, dx), tmax = fdiv(mxx - ox, dx);
    if (tmin > tmax) { float s = tmin; tmin = tmax; tmax = s; }
    float tymin = fdiv(mny - oy, dy), tymax = fdiv(mxy - oy, dy);
    if (tymin > tymax) { float s = tymin; tymin = tymax; tymax = s; }
    if ((tmin > tymax) || (tymin > tmax)) return o;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = fdiv(mnz - oz, dz), tzmax = fdiv(mxz - oz, dz);
    if (tzmin > tzmax) { float s = tzmin; tzmin = tzmax; tzmax = s; }
    if ((tmin > tzmax) || (tzmin > tmax)) return o;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    o.enter = (tmin == tmin) ? tmin : -INFINITY;   // NaN -> never prune
    o.exit_ = (tmax == tmax) ? tmax : INFINITY;
    o.hit = true;
    return o;
}

struct RayPre {
    float ix, iy, iz;  // 1/d, IEEE
    bool degenerate;   // a zero/non-finite direction component: every box goes through box_exact
};

TRB_HD RayPre ray_pre(const RayT& r) {
    RayPre p;
    p.ix = frcp(r.dx); p.iy = frcp(r.dy); p.iz = frcp(r.dz);
    float s = (p.ix - p.ix) + (p.iy - p.iy) + (p.iz - p.iz);  // NaN unless all finite
    p.degenerate = !(s == 0.f);
    return p;
}

// A box the ray LEAVES before this parameter cannot own an accepted hit: a hit point lies in its triangle, hence in the box, and the
// reference accepts only t > 1e-3 (accel.h:43).  Half of that threshold is left as margin for the rounding of the exit parameter
// (<= a few ulp(coordinate) x |1/d|) and of t itself, and only for rays none of whose direction components is small against the
// scene's coordinates (inv_max = 400 / largest |coordinate|: both errors stay below 1e-4); any other ray keeps the old rule, a box wholly
// behind the origin by more than the threshold.  What it buys: a secondary ray starts ON a surface, and when that surface is an
// axis-aligned wall (flat leaf box: exit ~ 0) its own leaf - 2-4 certain rejections - is no longer tested.
TRB_HD float near_cull(const RayPre& p, float inv_max) {
    const float m = fmaxf(fabsf(p.ix), fmaxf(fabsf(p.iy), fabsf(p.iz)));
    return m <= inv_max ? 5e-4f : -0.0010011f;
}

// Decides one child box the way the reference does; enter/exit are the (approximate) line parameters.
// The reference misses iff some CROSS-axis pair has tmin_i > tmax_j (bvh.h:33-69); the same-axis pair never
// decides.  Three tiers: (1) box_fast: the clipped interval [en, ex] is clearly non-empty / clearly empty -
// straight-line code, no branch, so that the two child tests of a node stay converged; (2) box_slow: the
// interval is empty-or-touching only because one axis binds both ends (flat boxes: every axis-aligned wall of
// the Cornell box) - then the six cross-axis gaps decide; (3) a cross-axis gap is within a few ulp, or the
// direction has a zero / non-finite component: the reference's own six IEEE divisions are redone (box_exact).
struct BoxF { float en, ex, nx, ny, nz, xx, xy, xz, tol; bool pass, undecided; };
TRB_HD BoxF box_fast(float mnx, float mny, float mnz, float mxx, float mxy, float mxz, const RayT& r, const RayPre& p) {
    BoxF f;
    float ax = (mnx - r.ox) * p.ix, bx = (mxx - r.ox) * p.ix;
    float ay = (mny - r.oy) * p.iy, by = (mxy - r.oy) * p.iy;
    float az = (mnz - r.oz) * p.iz, bz = (mxz - r.oz) * p.iz;
    f.nx = fminf(ax, bx); f.ny = fminf(ay, by); f.nz = fminf(az, bz);
    f.xx = fmaxf(ax, bx); f.xy = fmaxf(ay, by); f.xz = fmaxf(az, bz);
    f.en = fmaxf(fmaxf(f.nx, f.ny), f.nz);
    f.ex = fminf(fminf(f.xx, f.xy), f.xz);
    float gap = f.ex - f.en;
    f.tol = 6e-7f * fmaxf(fabsf(f.en), fabsf(f.ex)) + 1e-30f;
    f.pass = gap > f.tol;
    f.undecided = p.degenerate || !(f.pass || gap < -f.tol);
    return f;
}
// tiers 2 and 3 for an undecided box; updates en/ex when the exact test supplies them
TRB_HD bool box_slow(float mnx, float mny, float mnz, float mxx, float mxy, float mxz, const RayT& r, const RayPre& p, BoxF& f) {
    if (!p.degenerate) {
        float g = fminf(fminf(fminf(f.xy - f.nx, f.xz - f.nx), fminf(f.xx - f.ny, f.xz - f.ny)), fminf(f.xx - f.nz, f.xy - f.nz));
        float tol = f.tol + f.tol;
        if (g > tol) return true;
        if (g < -tol) return false;
    }
    BoxR e = box_exact(mnx, mny, mnz, mxx, mxy, mxz, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz);
    f.en = e.enter; f.ex = e.exit_;
    return e.hit;
}
TRB_HD bool box_test(const float* bmin, const float* bmax, const RayT& r, const RayPre& p, float* enter, float* exit_) {
    BoxF f = box_fast(bmin[0], bmin[1], bmin[2], bmax[0], bmax[1], bmax[2], r, p);
    bool hit = f.pass;
    if (f.undecided) hit = box_slow(bmin[0], bmin[1], bmin[2], bmax[0], bmax[1], bmax[2], r, p, f);
    *enter = f.en; *exit_ = f.ex;
    return hit;
}

// rayTriangleIntersect (core.h:379-400) + the t > 1e-3 acceptance of accel.h:43, same operation order.
// `ill` is set when the candidate is numerically ill-conditioned (see file header).
// The IEEE division 1/det is only paid by candidates that survive sign/magnitude pre-rejects which are strictly
// weaker than the reference's own tests on u and v (a pre-reject fires only where uu < 0, uu > 1, vv < 0 or
// uu + vv > 1 is certain for ANY rounding of the two multiplications), so accepted hits and their (t,u,v) are
// bit-identical to computing invDet first.
TRB_HD bool tri_test(const F4& q0, const F4& q1, const F4& q2, const RayT& r, float* t, float* u, float* v, bool* ill) {
    // pvec = cross(d, v0v2)
    float px = r.dy * q2.z - q2.y * r.dz, py = r.dz * q2.x - q2.z * r.dx, pz = r.dx * q2.y - q2.x * r.dy;
    float det = (q1.x * px + q1.y * py) + q1.z * pz;
    float adet = fabsf(det);
    if (adet < 1e-8f) return false;
    float tx = r.ox - q0.x, ty = r.oy - q0.y, tz = r.oz - q0.z;
    float un = (tx * px + ty * py) + tz * pz;
    const bool guard = adet <= 1e10f;           // keeps un * (1/det) away from underflow in the sign argument
    float s_un = det < 0.f ? -un : un;          // un * sign(det), exact
    if (guard && (s_un <= -1e-30f || s_un > adet * 1.000001f)) return false;  // uu < 0 or uu > 1 for certain
    // qvec = cross(tvec, v0v1)
    float qx = ty * q1.z - q1.y * tz, qy = tz * q1.x - q1.z * tx, qz = tx * q1.y - q1.x * ty;
    float vn = (r.dx * qx + r.dy * qy) + r.dz * qz;
    float s_vn = det < 0.f ? -vn : vn;
    if (guard && (s_vn <= -1e-30f || (s_un + s_vn) > adet * 1.000002f)) return false;  // vv < 0 or uu + vv > 1 for certain
    float invDet = frcp(det);
    float uu = un * invDet;
    if (uu < 0 || uu > 1) return false;
    float vv = vn * invDet;
    if (vv < 0 || uu + vv > 1) return false;
    float tt = ((q2.x * qx + q2.y * qy) + q2.z * qz) * invDet;
    // conditioning: |det| against |e1||e2||d| (grazing); q1.w = 1e-6f * |e1|^2 * |e2|^2 from the host build
    float dd = (r.dx * r.dx + r.dy * r.dy) + r.dz * r.dz;
    if (det * det < q1.w * dd) *ill = true;
    if (!(tt >= 0.001f)) return false;  // (double)t > 1e-3  <=>  t >= 0.001f for fp32 t
    *t = tt; *u = uu; *v = vv;
    return true;
}

// One traversal. ANY: occlusion query.  PRUNE=false visits what the reference visits (no distance test).
// Returns true when an ill-conditioned candidate was met (caller re-runs with PRUNE=false).
template <bool ANY, bool PRUNE>
TRB_HD bool traverse(const BvhView& bvh, const RayT& r, const RayPre& pre, HitT* hit, bool* occluded,
                     uint32_t* n_nodes, uint32_t* n_tris) {
    int32_t stack[TRB_STACK];
    int sp = 0;
    int32_t cur = bvh.root_ref;
    bool ill = false;
    float best = r.max_t;  // bvh.h:263: intersection->t = ray.max_t
    HitT h; h.t = r.max_t; h.u = 0.f; h.v = 0.f; h.pos = -1;
    *occluded = false;
    // bvh.h:287: the root is skipped when ray.min_t > ray.max_t
    if (r.min_t > r.max_t) { *hit = h; return false; }
    // One structured loop: every iteration is "inner node step" or "leaf step", both end by setting `cur`
    // (-1 = done), and the branches re-join at the bottom of the iteration.  No `continue`/`break` out of
    // divergent code: with independent thread scheduling those leave lanes permanently split.
    while (cur != -1) {
        if (cur >= 0) {
            const F4* n = bvh.nodes + 4 * size_t(cur);
            F4 a = ld4(n), b = ld4(n + 1), c = ld4(n + 2), d = ld4(n + 3);
            if (n_nodes) ++*n_nodes;
            const float c0min[3] = {a.x, a.y, a.z}, c0max[3] = {a.w, b.x, b.y};
            const float c1min[3] = {b.z, b.w, c.x}, c1max[3] = {c.y, c.z, c.w};
            float e0, x0, e1, x1;
            bool h0 = box_test(c0min, c0max, r, pre, &e0, &x0);
            bool h1 = box_test(c1min, c1max, r, pre, &e1, &x1);
            if (PRUNE) {
                float lim = best + 1e-3f * (1.f + fabsf(best));
                h0 = h0 && !(e0 > lim) && !(x0 < -1e-3f * (1.f + fabsf(x0)));
                h1 = h1 && !(e1 > lim) && !(x1 < -1e-3f * (1.f + fabsf(x1)));
            }
            int32_t r0 = f2i(d.x), r1 = f2i(d.y);
            // reference order on equal footing: left first (bvh.h:317-336); nearer first when pruning
            bool swap = PRUNE && h0 && h1 && (e1 < e0);
            int32_t first = swap ? r1 : r0, second = swap ? r0 : r1;
            bool hf = swap ? h1 : h0, hs = swap ? h0 : h1;
            if (hf && hs) { stack[sp++] = second; cur = first; }
            else if (hf) cur = first;
            else if (hs) cur = second;
            else cur = sp ? stack[--sp] : -1;
        } else {
            uint32_t code = uint32_t(~cur);
            uint32_t first = code >> 3, cnt = code & 7u;
            bool done = false;
            for (uint32_t k = 0; k < cnt; ++k) {
                const F4* tp = bvh.tris + 3 * size_t(first + k);
                F4 q0 = ld4(tp), q1 = ld4(tp + 1), q2 = ld4(tp + 2);
                if (n_tris) ++*n_tris;
                float t, u, v;
                if (tri_test(q0, q1, q2, r, &t, &u, &v, &ill)) {
                    if (ANY) {
                        // bvh.h:299-307: immediate return inside [min_t,max_t]; otherwise recorded when
                        // t < max_t, which also makes the final `object != nullptr` true
                        if ((t <= r.max_t && t >= r.min_t) || t < r.max_t) { *occluded = true; done = true; }
                    } else {
                        int32_t pos = int32_t(first + k);
                        // strict '<' in list order == (t, pos) lexicographic minimum for any visiting order
                        if (t < h.t || (t == h.t && h.pos >= 0 && pos < h.pos)) { h.t = t; h.u = u; h.v = v; h.pos = pos; best = t; }
                    }
                }
            }
            cur = (done || sp == 0) ? -1 : stack[--sp];
        }
    }
    if (ANY && *occluded) ill = false;
    *hit = h;
    return ill;
}

// Full query with the exhaustive re-trace for ill-conditioned rays.
template <bool ANY>
TRB_HD void trace_ray(const BvhView& bvh, const RayT& r, HitT* hit, bool* occluded, uint32_t* n_nodes, uint32_t* n_tris) {
    RayPre pre = ray_pre(r);
    bool ill = traverse<ANY, true>(bvh, r, pre, hit, occluded, n_nodes, n_tris);
    if (ill && !(ANY && *occluded)) {
        traverse<ANY, false>(bvh, r, pre, hit, occluded, n_nodes, n_tris);
    }
}

}  // namespace trbk
