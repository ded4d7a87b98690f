// trace_warp.cuh -- the production traversal kernels: persistent warps over a ray queue.
//
// Same hit semantics as the scalar traverse<>() of trace.cuh (which stays as the readable statement of the
// rule and as the host-compilable check); what changes is how a warp spends its issue slots.  The scalar
// loop measured 5.4 (closest) / 8.3 (any) active threads per issued instruction on B200
// (profiles/r1a_ncu_lines_*): lanes finish at different times and sit in different states (inner node /
// leaf), so the two code paths serialise inside every iteration.  Here:
//   * persistent warps REFILL idle lanes from the ray queue whenever TRB_FETCH_T lanes are idle (warp ballot
//     + popc ranks, one atomicAdd per refill), so the loop stays populated and the tail is balanced;
//   * a lane does not switch state at a leaf: the leaf's triangle range is POSTPONED into a two-entry pending
//     list (registers) and the lane keeps descending.  Every iteration is "node phase" then "triangle phase":
//     each lane with a current node tests its two child boxes, each lane with pending triangles tests ONE
//     triangle, so most lanes are active in both phases instead of half the warp idling in each;
//     sibling leaves are adjacent in the reference's primitive order (pre-order build), so one node adds at
//     most one contiguous range;
//   * the exhaustive re-trace of an ill-conditioned ray is a per-lane flag (prune off) on the same code path:
//     the lane restarts, the other 31 keep going.
// The closest hit is the lexicographic minimum of (t, list position) over the triangles reached, which does
// not depend on the order of testing, so postponing leaves changes no result (only how early pruning bites).
// Results: closest -> HitT per ray; any -> occlusion byte, or the pending contribution resolved into the
// frame buffer with atomics (replaces g_FrameBufferLocks, bdpt.h:360-370).
#pragma once
#include "bdpt_common.cuh"

#ifndef TRB_FETCH_T
#define TRB_FETCH_T 8        // refill when this many lanes are idle
#endif
#ifndef TRB_TRACE_BLOCK
#define TRB_TRACE_BLOCK 128
#endif
#ifndef TRB_TRACE_MINBLOCKS
#define TRB_TRACE_MINBLOCKS 8  // resident blocks per SM the register budget is capped for (8 x 128 threads x 64 registers)
#endif

#ifndef TRB_HOIST
#define TRB_HOIST 1          // issue the triangle fetch of this iteration BEFORE the node phase (two fetches in flight per lane)
#endif

namespace trbk {

struct TraceArgs {
    BvhView bvh;
    const RayT* rays;
    const int32_t* q;        // closest: optional indirection (ray/hit slot of queue entry k)
    const uint32_t* n_ptr;   // device-side count (queue length) or nullptr
    uint32_t n_fixed;
    uint32_t cap;            // any: clamp of *n_ptr (shadow queue capacity)
    HitT* hits;              // closest
    uint8_t* occ_out;        // any (ABI)
    const Pending* pend;     // any (wavefront): contribution per shadow ray
    float* accum;
    uint32_t* fetch;         // queue cursor of THIS launch (zeroed by the host before the launch)
    unsigned long long* visit;      // TRB_COUNTERS builds: [0] nodes visited, [1] triangles tested
    unsigned long long* visit_exh;  // TRB_COUNTERS builds: exhaustive re-traces
};

template <bool ANY>
__global__ void __launch_bounds__(TRB_TRACE_BLOCK, TRB_TRACE_MINBLOCKS) k_trace_warp(const TraceArgs a) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    uint32_t n = a.n_ptr ? *a.n_ptr : a.n_fixed;
    if (ANY && a.n_ptr) n = min(n, a.cap);
    bool more = true;  // warp-uniform: the queue may still hold rays

    // lane state
    int32_t cur = -1;          // current inner node or -1
    int sp = 0;
    uint32_t p0 = 0, p1 = 0;   // pending triangle ranges: (first << 4) | count, count <= 8
    int npend = 0;
    bool have = false, exh = false, ill = false, occ = false;
    RayT r; RayPre pre; HitT h;
    int32_t slot = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.min_t = r.max_t = 0.f;
    pre.ix = pre.iy = pre.iz = 0.f; pre.degenerate = false;
    h.t = 0.f; h.u = 0.f; h.v = 0.f; h.pos = -1;
    int32_t stack[TRB_STACK];
#ifdef TRB_COUNTERS
    uint32_t c_nodes = 0, c_tris = 0, c_exh = 0;
#endif
    auto start_ray = [&]() {   // (re)start the traversal of the lane's ray
        h.t = r.max_t; h.u = 0.f; h.v = 0.f; h.pos = -1;  // bvh.h:263: intersection->t = ray.max_t
        ill = false; occ = false; sp = 0; npend = 0; cur = -1;
        if (!(r.min_t > r.max_t)) {                         // bvh.h:287: the root is skipped when min_t > max_t
            if (a.bvh.root_ref >= 0) cur = a.bvh.root_ref;
            else { const uint32_t code = uint32_t(~a.bvh.root_ref); p0 = ((code >> 3) << 4) | (code & 7u); npend = 1; }
        }
    };

    while (true) {
        const bool idle = (cur < 0 && npend == 0);
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask == FULL || (__popc(idle_mask) >= TRB_FETCH_T && more)) {
            bool want = idle;
            if (idle && have) {
                if (ill && !exh && !(ANY && occ)) {
                    // ill-conditioned candidate met: visit what the reference visits (no distance pruning)
                    exh = true;
                    start_ray();
                    want = false;
#ifdef TRB_COUNTERS
                    ++c_exh;
#endif
                } else {
                    have = false;
                    if (ANY) {
                        if (a.occ_out) a.occ_out[slot] = occ ? 1 : 0;
                        if (a.pend && !occ) {
                            Pending pd = a.pend[slot];
                            float* dst = a.accum + size_t(pd.target) * 3;
                            atomicAdd(dst, pd.r); atomicAdd(dst + 1, pd.g); atomicAdd(dst + 2, pd.b);
                        }
                    } else {
                        a.hits[slot] = h;
                    }
                }
            }
            if (more) {
                const unsigned wm = __ballot_sync(FULL, want);
                const uint32_t cnt = uint32_t(__popc(wm));
                uint32_t base = 0;
                if (lane == 0 && cnt) base = atomicAdd(a.fetch, cnt);
                base = __shfl_sync(FULL, base, 0);
                const uint32_t idx = base + uint32_t(__popc(wm & lt));
                if (want && idx < n) {
                    slot = a.q ? a.q[idx] : int32_t(idx);
                    const float4* rp = reinterpret_cast<const float4*>(a.rays + slot);
                    float4 r0 = rp[0], r1 = rp[1];
                    r.ox = r0.x; r.oy = r0.y; r.oz = r0.z; r.dx = r0.w; r.dy = r1.x; r.dz = r1.y; r.min_t = r1.z; r.max_t = r1.w;
                    pre = ray_pre(r);
                    have = true; exh = false;
                    start_ray();
                }
                more = cnt == 0 || (base + cnt < n);
            }
            if (__ballot_sync(FULL, have) == 0) break;
            continue;
        }
#if TRB_HOIST
        // the triangle tested in this iteration is known now: fetch it first so that its latency overlaps the node phase
        const bool do_tri = npend > 0;
        const uint32_t tfirst = p0 >> 4, tcnt = p0 & 15u;
        F4 q0, q1, q2;
        if (do_tri) {
            const F4* tp = a.bvh.tris + 3 * size_t(tfirst);
            q0 = ld4(tp); q1 = ld4(tp + 1); q2 = ld4(tp + 2);
        }
#endif
        // ---- node phase: two child boxes of the current inner node ----
        if (cur >= 0 && npend < 2) {
            const F4* np = a.bvh.nodes + 4 * size_t(cur);
            F4 na = ld4(np), nb = ld4(np + 1), nc = ld4(np + 2), nd = ld4(np + 3);
#ifdef TRB_COUNTERS
            ++c_nodes;
#endif
            // both child tests as straight-line code; the rare undecided box takes the slow tiers afterwards
            BoxF f0 = box_fast(na.x, na.y, na.z, na.w, nb.x, nb.y, r, pre);
            BoxF f1 = box_fast(nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, r, pre);
            bool h0 = f0.pass, h1 = f1.pass;
            if (f0.undecided || f1.undecided) {
                if (f0.undecided) h0 = box_slow(na.x, na.y, na.z, na.w, nb.x, nb.y, r, pre, f0);
                if (f1.undecided) h1 = box_slow(nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, r, pre, f1);
            }
            const float e0 = f0.en, x0 = f0.ex, e1 = f1.en, x1 = f1.ex;
            if (!exh) {
                const float best = ANY ? r.max_t : h.t;
                const float lim = best + 1e-3f * (1.f + fabsf(best));
                h0 = h0 && !(e0 > lim) && !(x0 < -1e-3f * (1.f + fabsf(x0)));
                h1 = h1 && !(e1 > lim) && !(x1 < -1e-3f * (1.f + fabsf(x1)));
            }
            const int32_t r0 = f2i(nd.x), r1 = f2i(nd.y);
            // leaves are postponed: one contiguous range (the right sibling's triangles follow the left's)
            const bool l0 = h0 && r0 < 0, l1 = h1 && r1 < 0;
            if (l0 || l1) {
                const uint32_t k0 = uint32_t(~r0), k1 = uint32_t(~r1);
                const uint32_t first = l0 ? (k0 >> 3) : (k1 >> 3);
                const uint32_t cnt = (l0 ? (k0 & 7u) : 0u) + (l1 ? (k1 & 7u) : 0u);
                const uint32_t e = (first << 4) | cnt;
                if (npend == 0) p0 = e; else p1 = e;
                ++npend;
            }
            const bool i0 = h0 && r0 >= 0, i1 = h1 && r1 >= 0;
            if (i0 && i1) {
                const bool swap = !exh && (e1 < e0);  // nearer child first when pruning
                stack[sp++] = swap ? r0 : r1;
                cur = swap ? r1 : r0;
            } else if (i0) cur = r0;
            else if (i1) cur = r1;
            else cur = sp ? stack[--sp] : -1;
        }
        // ---- triangle phase: one postponed triangle ----
#if TRB_HOIST
        if (do_tri) {
            const uint32_t first = tfirst, cnt = tcnt;
#else
        if (npend > 0) {
            const uint32_t first = p0 >> 4, cnt = p0 & 15u;
            const F4* tp = a.bvh.tris + 3 * size_t(first);
            F4 q0 = ld4(tp), q1 = ld4(tp + 1), q2 = ld4(tp + 2);
#endif
#ifdef TRB_COUNTERS
            ++c_tris;
#endif
            float t, u, v;
            bool done = false;
            if (tri_test(q0, q1, q2, r, &t, &u, &v, &ill)) {
                if (ANY) {
                    // bvh.h:299-307 + :351: inside [min_t,max_t] returns at once; t < max_t makes `object` non-null
                    if ((t <= r.max_t && t >= r.min_t) || t < r.max_t) { occ = true; done = true; ill = false; }
                } else {
                    const int32_t pos = int32_t(first);
                    // strict '<' in list order == (t, pos) lexicographic minimum for any visiting order
                    if (t < h.t || (t == h.t && h.pos >= 0 && pos < h.pos)) { h.t = t; h.u = u; h.v = v; h.pos = pos; }
                }
            }
            if (done) { cur = -1; sp = 0; npend = 0; }
            else if (cnt > 1) p0 = ((first + 1) << 4) | (cnt - 1);
            else { p0 = p1; --npend; }
        }
    }
#ifdef TRB_COUNTERS
    if (a.visit) {
        atomicAdd(a.visit, (unsigned long long)c_nodes); atomicAdd(a.visit + 1, (unsigned long long)c_tris);
        if (a.visit_exh) atomicAdd(a.visit_exh, (unsigned long long)c_exh);
    }
#endif
}

}  // namespace trbk
