// trace_wide.cuh -- traversal of the 4-wide upper tree over the reference's leaves (host/tr_host.h: DevNode4).
//
// What is found is what trace.cuh finds (and what BVH::getIntersection, externals/bvh.h:259-352, finds):
//   closest : the (t, list position) minimum over the triangles of every leaf whose box passes BBox::intersect
//             (bvh.h:33-69) - for a ray whose direction has three finite non-zero components "the leaf box passes" is the
//             same set as "all ancestor boxes pass" (every box is the exact union of its children, fl((b - o) / d) is
//             monotone in b; the cross-axis comparisons that decide the slab test can only get easier for a larger box);
//   any-hit : some such triangle with t <= max_t.
// Consequences used here:
//   * an INNER box only has to be conservative: it is rejected when the straight-line slab test is decisive
//     (`gap < -tol`, which implies a cross-axis violation by more than the rounding of o + t d), and entered otherwise -
//     no exact six-division test above the leaves;
//   * a LEAF box is decided exactly as the reference decides it (box_fast / box_slow / box_exact of trace.cuh);
//   * the shape of the upper tree is free: four children per node halves the number of dependent fetches per ray.
// Rays with a zero / non-finite direction component (RayPre::degenerate) do NOT come here: NaN comparisons make an
// ancestor's test fail where the leaf's would pass; they are traced on the reference's own binary tree (trace.cuh /
// trace_coop.cuh: k_trace_coop).
//
// traverse_wide() below is the scalar statement of the rule (host-compilable: tools/host_trace_check.cpp checks it
// against the compiled reference on a box without a GPU); k_trace_wide is the production kernel.
#pragma once
#include "trace.cuh"

namespace trbk {

#ifndef TRB_WIDE_EMPTY
#define TRB_WIDE_EMPTY INT32_MIN
#endif

struct WideView {
    const F4* wnodes;   // 8 x F4 per node (DevNode4: two 64-byte halves)
    const F4* tris;     // 3 x F4 per triangle (DevTri)
    int32_t root_ref;   // 0 (wide node 0) or a leaf code when the scene is a single leaf
    float inv_max;      // near_cull (trace.cuh): 400 / the scene's largest |coordinate|; 0 = the old rule for every ray
};

struct WideCounters { uint32_t nodes, boxes, tris, max_sp; };

// One child of a wide node: leaf children get the reference's decision, inner children a conservative one.
TRB_HD bool wide_child(float mnx, float mny, float mnz, float mxx, float mxy, float mxz, int32_t ref, const RayT& r, const RayPre& p,
                       float* en, float* ex) {
    BoxF f = box_fast(mnx, mny, mnz, mxx, mxy, mxz, r, p);
    bool hit = f.pass;
    if (f.undecided) hit = ref >= 0 ? true : box_slow(mnx, mny, mnz, mxx, mxy, mxz, r, p, f);
    *en = f.en; *ex = f.ex;
    return hit && ref != TRB_WIDE_EMPTY;
}

// ANY: occlusion query.  PRUNE=false visits every leaf the reference reaches.  Returns true when an ill-conditioned
// candidate was met (the caller re-runs with PRUNE=false, as trace_ray does).  The ray must not be degenerate.
template <bool ANY, bool PRUNE>
TRB_HD bool traverse_wide(const WideView& bvh, const RayT& r, const RayPre& pre, HitT* hit, bool* occluded, WideCounters* cnt) {
    int32_t stack[3 * TRB_STACK];
    int sp = 0;
    int32_t cur = bvh.root_ref;
    bool ill = false;
    float best = r.max_t;  // bvh.h:263
    HitT h; h.t = r.max_t; h.u = 0.f; h.v = 0.f; h.pos = -1;
    *occluded = false;
    if (r.min_t > r.max_t) { *hit = h; return false; }   // bvh.h:287
    const int32_t NONE = 0x7fffffff;
    const float ncull = near_cull(pre, bvh.inv_max);
    while (cur != NONE) {
        if (cur >= 0) {
            const F4* n = bvh.wnodes + 8 * size_t(cur);   // two 64-byte halves: {c0min c0max c1min c1max ref0 ref1 pad pad}
            if (cnt) { ++cnt->nodes; }
            float L[4][6]; int32_t ref[4];
            for (int hf = 0; hf < 2; ++hf) {
                const F4 a = ld4(n + 4 * hf), b = ld4(n + 4 * hf + 1), c = ld4(n + 4 * hf + 2), d = ld4(n + 4 * hf + 3);
                const float c0[6] = {a.x, a.y, a.z, a.w, b.x, b.y}, c1[6] = {b.z, b.w, c.x, c.y, c.z, c.w};
                for (int k = 0; k < 6; ++k) { L[2 * hf][k] = c0[k]; L[2 * hf + 1][k] = c1[k]; }
                ref[2 * hf] = f2i(d.x); ref[2 * hf + 1] = f2i(d.y);
            }
            float en[4]; bool ok[4];
            const float lim = best + 1e-3f * (1.f + fabsf(best));
            for (int k = 0; k < 4; ++k) {
                float ex;
                if (cnt && ref[k] != TRB_WIDE_EMPTY) ++cnt->boxes;
                ok[k] = wide_child(L[k][0], L[k][1], L[k][2], L[k][3], L[k][4], L[k][5], ref[k], r, pre, &en[k], &ex);
                if (PRUNE) ok[k] = ok[k] && !(en[k] > lim) && !(ex < ncull);
            }
            // onto the stack in slot order, slot 0 on top; when pruning a closest-hit ray the nearest child goes on top of that
            // (what k_trace_wide does: a full sort by entry distance saves 3 % of the triangle tests and costs more in instructions)
            int ord[4] = {3, 2, 1, 0};
            if (PRUNE && !ANY) {
                int nb = -1;
                for (int k = 0; k < 4; ++k) if (ok[k] && (nb < 0 || en[k] < en[nb])) nb = k;
                if (nb >= 0) { int j = 0; for (int k = 3; k >= 0; --k) if (k != nb) ord[j++] = k; ord[3] = nb; }
            }
            for (int a = 0; a < 4; ++a) if (ok[ord[a]]) stack[sp++] = ref[ord[a]];
            if (cnt && uint32_t(sp) > cnt->max_sp) cnt->max_sp = uint32_t(sp);
            cur = sp ? stack[--sp] : NONE;
        } else {
            const uint32_t code = uint32_t(~cur);
            const uint32_t first = code >> 3, n_tri = code & 7u;
            bool done = false;
            for (uint32_t k = 0; k < n_tri; ++k) {
                const F4* tp = bvh.tris + 3 * size_t(first + k);
                F4 q0 = ld4(tp), q1 = ld4(tp + 1), q2 = ld4(tp + 2);
                if (cnt) ++cnt->tris;
                float t, u, v;
                if (tri_test(q0, q1, q2, r, &t, &u, &v, &ill)) {
                    if (ANY) {
                        if ((t <= r.max_t && t >= r.min_t) || t < r.max_t) { *occluded = true; done = true; }   // bvh.h:299-307 + :351
                    } else {
                        const int32_t pos = int32_t(first + k);
                        if (t < h.t || (t == h.t && h.pos >= 0 && pos < h.pos)) { h.t = t; h.u = u; h.v = v; h.pos = pos; best = t; }
                    }
                }
            }
            cur = (done || sp == 0) ? NONE : stack[--sp];
        }
    }
    if (ANY && *occluded) ill = false;
    *hit = h;
    return ill;
}

// Full query: wide tree for ordinary rays, the reference's binary tree for degenerate ones, exhaustive re-trace of
// ill-conditioned rays.
template <bool ANY>
TRB_HD void trace_ray_wide(const WideView& wide, const BvhView& bin, const RayT& r, HitT* hit, bool* occluded, WideCounters* cnt) {
    RayPre pre = ray_pre(r);
    if (pre.degenerate) { trace_ray<ANY>(bin, r, hit, occluded, nullptr, nullptr); return; }
    bool ill = traverse_wide<ANY, true>(wide, r, pre, hit, occluded, cnt);
    if (ill && !(ANY && *occluded)) traverse_wide<ANY, false>(wide, r, pre, hit, occluded, cnt);
}

}  // namespace trbk

#if defined(__CUDACC__)
#include "trace_coop.cuh"

#ifndef TRB_SSTACK
#define TRB_SSTACK 12         // stack entries per lane kept in shared memory ([entry][thread]: conflict-free); deeper entries spill to local memory
#endif
#define TRB_LSTACK (3 * TRB_STACK - TRB_SSTACK)   // a visit pushes at most 3; the upload checks 3 * wide_depth + 4 <= 3 * TRB_STACK

namespace trbk {

struct F8 { float a, b, c, d, e, f, g, h; };
// 256-bit read-only global load (sm_100: LDG.E.256): half the L1 wavefronts of two LDG.128 when every lane reads its own line
__device__ __forceinline__ F8 ld8(const float* p) {
    F8 v;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v.a), "=f"(v.b), "=f"(v.c), "=f"(v.d), "=f"(v.e), "=f"(v.f), "=f"(v.g), "=f"(v.h) : "l"(p));
    return v;
}

__device__ __forceinline__ F8 ld8_2x4(const float* p) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    return F8{a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
}
#ifndef TRB_WIDE_LD256
#define TRB_WIDE_LD256 1
#endif
#ifndef TRB_WIDE_PREFETCH
#define TRB_WIDE_PREFETCH 0      // 1: prefetch the next node of the lane at the end of the node phase (measured: see profiles/r2_node_phase_experiments.md)
#endif

// The production traversal kernel: persistent warps, lane refill, per-warp triangle queue (trace_coop.cuh) over the 4-wide tree.
// Per node visit a lane loads 4 x 32 B, tests four boxes in straight-line code and
//   * queues the triangles of up to two hit leaves (one range push, <= 8 entries: the queue invariant of trace_coop.cuh) and parks
//     further hit leaves on its stack as leaf references;
//   * continues with the nearest hit inner child (first in slot order for occlusion / exhaustive rays) and stacks the others.
// A ray with a zero / non-finite direction component is not traced here: its slot is appended to `defer_q` for k_trace_coop.
template <bool ANY>
__global__ void __launch_bounds__(TRB_TRACE_BLOCK, TRB_TRACE_MINBLOCKS) k_trace_wide(const TraceArgs a) {
    __shared__ CoopWarp sh_all[TRB_TRACE_BLOCK / 32];
    __shared__ int32_t sstack[TRB_SSTACK][TRB_TRACE_BLOCK];
    CoopWarp& sh = sh_all[threadIdx.x >> 5];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int32_t NONE = 0x7fffffff;
    uint32_t n = a.n_ptr ? *a.n_ptr : a.n_fixed;
    if (a.n_ptr) n = min(n, a.cap);
    bool more = true;       // warp-uniform: the ray queue may still hold rays
    uint32_t head = 0;      // warp-uniform read cursor of the triangle queue

    int32_t cur = NONE;     // wide node (>= 0), parked leaf reference (< 0) or NONE
    int sp = 0;
    bool have = false, exh = false;
    RayT r; RayPre pre;
    float ncull = -0.0010011f;   // near_cull(pre) of the lane's ray
    int32_t slot = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.min_t = r.max_t = 0.f;
    pre.ix = pre.iy = pre.iz = 0.f; pre.degenerate = false;
    int32_t lstack[TRB_LSTACK];
#ifdef TRB_COUNTERS
    uint32_t c_nodes = 0, c_tris = 0, c_exh = 0;
#endif
    sh.outst[lane] = 0; sh.flags[lane] = 0; sh.best[lane] = 0ull;
    if (lane == 0) sh.tail = 0;
    __syncwarp();

    auto push = [&](int32_t x) {
        TRB_CHECK(sp < TRB_SSTACK + TRB_LSTACK);
        if (sp < TRB_SSTACK) sstack[sp][threadIdx.x] = x; else lstack[sp - TRB_SSTACK] = x;
        ++sp;
    };
    auto pop = [&]() -> int32_t {
        if (sp == 0) return NONE;
        --sp;
        return sp < TRB_SSTACK ? sstack[sp][threadIdx.x] : lstack[sp - TRB_SSTACK];
    };
    // this lane's ray must test the triangles of leaf code ca, then of cb (cb = 0: none); <= 8 entries in all
    auto push_leaves = [&](uint32_t ca, uint32_t cb) {
        const uint32_t na = ca & 7u, cnt = na + (cb & 7u);
        TRB_CHECK(na >= 1u && na <= 4u && (cb & 7u) <= 4u && (ca >> 3) + na <= a.n_tris && (cb >> 3) + (cb & 7u) <= a.n_tris);
        const uint32_t base = atomicAdd(&sh.tail, cnt);
        const uint32_t ea = (uint32_t(lane) << 27) | (ca >> 3), eb = ((uint32_t(lane) << 27) | (cb >> 3)) - na;
#pragma unroll
        for (uint32_t k = 0; k < 4u; ++k)   // straight-line predicated stores: a leaf holds <= 4 triangles
            if (k < cnt) sh.queue[(base + k) & (TRB_COOP_Q - 1u)] = (k < na ? ea : eb) + k;
        if (cnt > 4u) {
#pragma unroll
            for (uint32_t k = 4; k < 8u; ++k)
                if (k < cnt) sh.queue[(base + k) & (TRB_COOP_Q - 1u)] = (k < na ? ea : eb) + k;
        }
        sh.outst[lane] += int32_t(cnt);   // only the owner adds; testers subtract atomically between __syncwarp()s
    };
    auto start_ray = [&]() {
        sh.best[lane] = (unsigned long long)__float_as_uint(r.max_t) << 32;   // bvh.h:263: intersection->t = ray.max_t
        sh.flags[lane] = 0; sp = 0; cur = NONE;
        if (!(r.min_t > r.max_t)) cur = a.wroot;                               // bvh.h:287
    };

    while (true) {
        const bool idle = cur == NONE && (!have || sh.outst[lane] == 0);
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask == FULL || (__popc(idle_mask) >= TRB_FETCH_T && more)) {
            bool want = idle, fin = false, occ = false;
            if (idle && have) {
                const uint32_t fl = sh.flags[lane];
                occ = (fl & 2u) != 0;
                if ((fl & 1u) && !exh && !(ANY && occ)) {
                    exh = true;   // ill-conditioned candidate met: visit every leaf the reference reaches (no distance pruning)
                    start_ray();
                    want = false;
#ifdef TRB_COUNTERS
                    ++c_exh;
#endif
                } else {
                    have = false;
                    fin = true;
                }
            }
            // A refill's round trips - the finished shadow ray's pending contribution, the queue cursor - are started before anything
            // waits on either (the compiler keeps a load behind an atomic it cannot prove unrelated); the next ray's load follows.
            Pending pd; pd.r = pd.g = pd.b = 0.f; pd.target = 0u;
            const bool resolve = ANY && fin && !occ && a.pend != nullptr;
            if (resolve) pd = a.pend[slot];
            unsigned wm = 0u; uint32_t cnt = 0u, base = 0u;
            if (more) {
                wm = __ballot_sync(FULL, want);
                cnt = uint32_t(__popc(wm));
                if (lane == 0 && cnt) base = atomicAdd(a.fetch, cnt);
            }
            if (fin) {
                if (ANY) {
                    if (a.occ_out) a.occ_out[slot] = occ ? 1 : 0;
                    if (resolve) {
                        float* dst = a.accum + size_t(pd.target) * 3;
                        atomicAdd(dst, pd.r); atomicAdd(dst + 1, pd.g); atomicAdd(dst + 2, pd.b);
                    }
                } else {
                    const unsigned long long b = sh.best[lane];
                    HitT h; h.t = r.max_t; h.u = 0.f; h.v = 0.f; h.pos = -1;
                    if (uint32_t(b >> 32) != __float_as_uint(r.max_t)) {
                        // the winner's (u, v) were left in shared memory by the lane whose test won (triangle phase): the values of
                        // the very test that won, so bit-identical to the reference's
                        const float2 uv = sh.uv[lane];
                        h.t = __uint_as_float(uint32_t(b >> 32)); h.u = uv.x; h.v = uv.y; h.pos = int32_t(uint32_t(b));
                    }
                    a.hits[slot] = h;
                }
            }
            if (more) {
                base = __shfl_sync(FULL, base, 0);
                const uint32_t idx = base + uint32_t(__popc(wm & lt));
                if (want && idx < n) {
                    slot = a.q ? a.q[idx] : int32_t(idx);
                    const float4* rp = reinterpret_cast<const float4*>(a.rays + slot);   // (plain loads: the rays were written by the previous kernel)
                    float4 r0 = rp[0], r1 = rp[1];
                    r.ox = r0.x; r.oy = r0.y; r.oz = r0.z; r.dx = r0.w; r.dy = r1.x; r.dz = r1.y; r.min_t = r1.z; r.max_t = r1.w;
                    pre = ray_pre(r);
                    ncull = near_cull(pre, a.inv_max);
                    if (pre.degenerate) { const uint32_t at = atomicAdd(a.defer_n, 1u); TRB_CHECK(at < a.defer_cap); a.defer_q[at] = slot; }   // for the reference's own tree (k_trace_coop)
                    else { have = true; exh = false; start_ray(); }
                }
                more = cnt == 0 || (base + cnt < n);
            }
            __syncwarp();
            if (__ballot_sync(FULL, have) == 0 && !more) break;
        }
        // ---- node phase ----
        if (cur != NONE) {
            const bool stop = ANY && (sh.flags[lane] & 2u) != 0;   // already occluded: nothing left to find
            if (cur < 0) {          // a leaf parked on the stack (its box passed when its parent was visited)
                if (stop) sp = 0; else push_leaves(uint32_t(~cur), 0u);
                cur = pop();
            } else {
                TRB_CHECK(uint32_t(cur) < a.n_wnodes);
                const float* np = reinterpret_cast<const float*>(a.wnodes + 8 * size_t(cur));
#if TRB_WIDE_LD256
                const F8 h0a = ld8(np), h0b = ld8(np + 8), h1a = ld8(np + 16), h1b = ld8(np + 24);   // 4 x LDG.256
#else
                const F8 h0a = ld8_2x4(np), h0b = ld8_2x4(np + 8), h1a = ld8_2x4(np + 16), h1b = ld8_2x4(np + 24);   // 8 x LDG.128
#endif
#ifdef TRB_COUNTERS
                ++c_nodes;
#endif
                float lim = 0.f;
                if (!exh) {
                    const float best = ANY ? r.max_t : __uint_as_float(uint32_t(sh.best[lane] >> 32));
                    lim = best + 1e-3f * (1.f + fabsf(best));
                }
                // One child: an inner child is entered unless the straight-line test rejects it; a leaf child gets the reference's
                // decision - its slow tiers run right here (flat boxes - every wall of a Cornell box - take them on every visit).
                // Then the distance cull (beyond the best hit / wholly behind the origin), unless the ray is in exhaustive mode.
                uint32_t ca = 0u, cb = 0u; int nl = 0;        // leaf codes queued by this visit
                int32_t nref = -1; float nen = 0.f;           // the inner child to continue with
                int32_t o0 = -1, o1 = -1, o2 = -1;            // the other hit inner children, slot order
#define TRB_WCHILD(mnx, mny, mnz, mxx, mxy, mxz, rf) { \
                    const int32_t ref = (rf); \
                    BoxF f = box_fast(mnx, mny, mnz, mxx, mxy, mxz, r, pre); \
                    bool hit = f.pass || f.undecided; \
                    if (f.undecided && ref < 0) hit = box_slow(mnx, mny, mnz, mxx, mxy, mxz, r, pre, f); \
                    hit = hit && ref != TRB_WIDE_EMPTY && !stop && (exh || (!(f.en > lim) && !(f.ex < ncull))); \
                    if (hit && ref < 0) { if (nl == 0) ca = uint32_t(~ref); else if (nl == 1) cb = uint32_t(~ref); else push(ref); ++nl; } \
                    if (hit && ref >= 0) { \
                        const bool take = nref < 0 || (!ANY && !exh && f.en < nen); \
                        const int32_t other = take ? nref : ref; \
                        if (take) { nref = ref; nen = f.en; } \
                        if (other >= 0) { if (o0 < 0) o0 = other; else if (o1 < 0) o1 = other; else o2 = other; } \
                    } }
                TRB_WCHILD(h0a.a, h0a.b, h0a.c, h0a.d, h0a.e, h0a.f, f2i(h0b.e))
                TRB_WCHILD(h0a.g, h0a.h, h0b.a, h0b.b, h0b.c, h0b.d, f2i(h0b.f))
                TRB_WCHILD(h1a.a, h1a.b, h1a.c, h1a.d, h1a.e, h1a.f, f2i(h1b.e))
                TRB_WCHILD(h1a.g, h1a.h, h1b.a, h1b.b, h1b.c, h1b.d, f2i(h1b.f))
#undef TRB_WCHILD
                if (nl > 0) push_leaves(ca, cb);
                if (o2 >= 0) push(o2);
                if (o1 >= 0) push(o1);
                if (o0 >= 0) push(o0);
                if (stop) sp = 0;
                cur = nref >= 0 ? nref : pop();
            }
#if TRB_WIDE_PREFETCH
            // the node this lane visits next is known now, a triangle phase and perhaps a refill away: one line, certain to be used
            if (cur >= 0 && cur != NONE) asm volatile("prefetch.global.L1 [%0];" :: "l"(a.wnodes + 8 * size_t(cur)));
#endif
        }
        __syncwarp();
        // ---- triangle phases: full warps of queued tests; drain when no lane is traversing or the queue is filling up ----
        uint32_t count = sh.tail - head;
        TRB_CHECK(count <= TRB_COOP_Q);   // the per-warp queue never wraps onto unread entries
        const bool trav = __ballot_sync(FULL, cur != NONE) != 0;
        while (count >= TRB_COOP_DRAIN || (count > 0u && (!trav || count > TRB_COOP_Q - 256u))) {
            const uint32_t nb_ = min(count, 32u);
            const bool act = uint32_t(lane) < nb_;
            const uint32_t e = act ? sh.queue[(head + uint32_t(lane)) & (TRB_COOP_Q - 1u)] : (uint32_t(lane) << 27);
            const int owner = int(e >> 27);
            const uint32_t pos = e & 0x07ffffffu;
            RayT o;   // the owner's ray, from the owner's registers
            o.ox = __shfl_sync(FULL, r.ox, owner); o.oy = __shfl_sync(FULL, r.oy, owner); o.oz = __shfl_sync(FULL, r.oz, owner);
            o.dx = __shfl_sync(FULL, r.dx, owner); o.dy = __shfl_sync(FULL, r.dy, owner); o.dz = __shfl_sync(FULL, r.dz, owner);
            o.max_t = __shfl_sync(FULL, r.max_t, owner); o.min_t = 0.f;
            float t = 0.f, u = 0.f, v = 0.f;
            bool ill = false, thit = false;
            unsigned long long mine = 0ull;
            if (act) {
                TRB_CHECK(pos < a.n_tris);
                const F4* tp = a.bvh.tris + 3 * size_t(pos);
                F4 q0 = ld4(tp), q1 = ld4(tp + 1), q2 = ld4(tp + 2);
#ifdef TRB_COUNTERS
                ++c_tris;
#endif
                thit = tri_test(q0, q1, q2, o, &t, &u, &v, &ill);
            }
            if (thit) {
                if (ANY) {
                    // bvh.h:299-307 + :351: occluded iff some accepted t <= max_t (min_t <= max_t holds for a traced ray)
                    if (t <= o.max_t) { atomicOr(&sh.flags[owner], 2u); ill = false; }
                } else {
                    // (t, list position) lexicographic minimum == strict '<' in list order for any visiting order
                    mine = ((unsigned long long)__float_as_uint(t) << 32) | pos;
                    atomicMin(&sh.best[owner], mine);
                }
            }
            if (ill) atomicOr(&sh.flags[owner], 1u);
            // outstanding-test counts: one subtraction per RUN of entries of the same owner (a leaf's triangles are adjacent in the
            // queue), not one per lane - every lane hitting the same word serialises in the shared-memory atomic unit
            const int prev_owner = __shfl_up_sync(FULL, owner, 1);
            const bool run_head = act && (lane == 0 || prev_owner != owner);
            const unsigned heads = __ballot_sync(FULL, run_head);
            if (run_head) {
                const unsigned rest = lane == 31 ? 0u : (heads >> (lane + 1));
                const int len = rest ? __ffs(int(rest)) : int(nb_) - lane;
                { const int before = atomicSub(&sh.outst[owner], len); TRB_CHECK(len >= 1 && before >= len); (void)before; }
            }
            head += nb_; count -= nb_;
            __syncwarp();
            // every merge of this phase is in: the one lane whose packed value is the owner's minimum now leaves its (u, v) there
            // (list positions are unique, so at most one lane matches; a later phase can only replace it after its own __syncwarp)
            if (!ANY && thit && sh.best[owner] == mine) sh.uv[owner] = make_float2(u, v);
        }
        if (!ANY) __syncwarp();   // the (u, v) of the last phase are in place before a completed ray reads its own
    }
#ifdef TRB_COUNTERS
    if (a.visit) {
        atomicAdd(a.visit, (unsigned long long)c_nodes); atomicAdd(a.visit + 1, (unsigned long long)c_tris);
        if (a.visit_exh) atomicAdd(a.visit_exh, (unsigned long long)c_exh);
    }
#endif
}

}  // namespace trbk
#endif  // __CUDACC__
