// trace_coop.cuh -- persistent-warp traversal of the reference's BINARY tree (64 B DevNode), warp-cooperative triangle tests.
//
// Since round 2 the production rays go through the 4-wide upper tree (trace_wide.cuh: k_trace_wide); this kernel traces what
// that one sets aside - rays with a zero / non-finite direction component, for which the reference's NaN behaviour in
// BBox::intersect (bvh.h:33-69) makes "all ancestors pass" differ from "the leaf passes" - on the reference's own topology.
// It is also the shared machinery of both kernels (TraceArgs, the per-warp triangle queue):
//   * persistent warps REFILL idle lanes from the ray queue whenever TRB_FETCH_T lanes are idle (warp ballot + popc ranks, one
//     atomicAdd per refill), so the loop stays populated and the tail is balanced;
//   * node phase: every lane with a current node tests its two child boxes; a reached leaf (or pair of sibling leaves = one
//     contiguous range) is pushed triangle by triangle into the warp's queue in shared memory as (owner lane, triangle);
//   * triangle phase: runs whenever the queue holds >= 32 entries (or nothing else is left to do); lane i tests entry i for
//     its OWNER's ray - read from the owner's registers with warp shuffles - and merges the result into the owner's record in
//     shared memory: atomicMin on the packed (t bits << 32 | list position) for closest hits, which IS the reference's rule
//     (min t, ties -> first in the build-reordered list, bvh.h:305), an OR of flags for occlusion / ill-conditioning.  (u, v)
//     of the winner are recomputed once at completion by the same arithmetic, hence bit-identical;
//   * the exhaustive re-trace of an ill-conditioned ray is a per-lane flag (prune off) on the same code path.
// Results: closest -> HitT per ray; any -> occlusion byte, or the pending contribution resolved into the frame buffer with
// atomics (replaces g_FrameBufferLocks, bdpt.h:360-370).
// Measured and removed in round 2 (notes kept under profiles/): per-lane postponed leaves (r1d), top levels of the tree
// staged in shared memory (r1n_top_levels_shared.md: 3-7 % slower), sorting the ray queue (r1f_tuning_variants.md).
#pragma once
#include "bdpt_common.cuh"

#ifndef TRB_FETCH_T
#define TRB_FETCH_T 16       // refill when this many lanes are idle (B200, SAH upper tree: 4 / 8 / 12 / 16 / 20 / 24 / 28 -> caustic 129.9 / 135.9 / 139.5 / 141.8 / 140.9 / 137.9 / 122.1 Msamples/s, profiles/r2_node_phase_experiments.md)
#endif
#ifndef TRB_TRACE_BLOCK
#define TRB_TRACE_BLOCK 128
#endif
#ifndef TRB_TRACE_MINBLOCKS
#define TRB_TRACE_MINBLOCKS 8  // resident blocks per SM the register budget is capped for (8 x 128 threads x 64 registers)
#endif
#ifndef TRB_COOP_DRAIN
#define TRB_COOP_DRAIN 32u    // run a triangle phase when this many tests are queued
#endif
#ifndef TRB_COOP_Q
#define TRB_COOP_Q 512u       // queue entries per warp (power of two); a node phase adds at most 8 x 32
#endif

// -DTRB_CHECKS (build.py --variant checks): a device assert on every index the traversal kernels form - node, triangle, stack, per-warp
// queue occupancy, deferred-ray queue, outstanding-test counts.  compute-sanitizer is closed on this pool, so this variant run over
// tools/sanitize_small.py (tests/test_gpu_build.py::test_checked_build_runs_clean) is the memory-safety evidence of the traversal.
#ifdef TRB_CHECKS
#include <cassert>
#define TRB_CHECK(x) assert(x)
#else
#define TRB_CHECK(x) ((void)0)
#endif

namespace trbk {

struct TraceArgs {
    BvhView bvh;             // the reference's binary tree (k_trace_coop)
    const F4* wnodes;        // the 4-wide upper tree (k_trace_wide); wroot = 0 or a leaf code
    int32_t wroot;
    const RayT* rays;
    const int32_t* q;        // optional indirection (ray / result slot of queue entry k)
    const uint32_t* n_ptr;   // device-side count (queue length) or nullptr
    uint32_t n_fixed;
    uint32_t cap;            // clamp of *n_ptr (queue capacity)
    HitT* hits;              // closest
    uint8_t* occ_out;        // any (ABI)
    const Pending* pend;     // any (wavefront): contribution per shadow ray
    float* accum;
    uint32_t* fetch;         // queue cursor of THIS launch (zeroed by the host before the launch)
    int32_t* defer_q;        // k_trace_wide: slots of the rays it leaves to k_trace_coop, and their count
    uint32_t* defer_n;
    uint32_t n_nodes, n_wnodes, n_tris, defer_cap;   // array lengths (TRB_CHECKS builds assert against them)
    float inv_max;           // near_cull (trace.cuh): 400 / the scene's largest |coordinate|
    unsigned long long* visit;      // TRB_COUNTERS builds: [0] nodes visited, [1] triangles tested
    unsigned long long* visit_exh;  // TRB_COUNTERS builds: exhaustive re-traces
};

struct CoopWarp {
    unsigned long long best[32];   // closest: (float bits of t << 32) | list position; init (bits(max_t) << 32)
    int32_t outst[32];             // triangles queued for this lane's ray and not yet tested
    uint32_t flags[32];            // bit 0: ill-conditioned candidate met, bit 1: occluded
    uint32_t tail;                 // queue write cursor (monotone; entries live at index & (Q-1))
    uint32_t pad[3];
    float2 uv[32];                 // k_trace_wide, closest: (u, v) of the current winner, written by the lane whose test won
    uint32_t queue[TRB_COOP_Q];    // (owner lane << 27) | triangle position
};

template <bool ANY>
__global__ void __launch_bounds__(TRB_TRACE_BLOCK, TRB_TRACE_MINBLOCKS) k_trace_coop(const TraceArgs a) {
    __shared__ CoopWarp sh_all[TRB_TRACE_BLOCK / 32];
    CoopWarp& sh = sh_all[threadIdx.x >> 5];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    uint32_t n = a.n_ptr ? *a.n_ptr : a.n_fixed;
    if (a.n_ptr) n = min(n, a.cap);
    bool more = true;       // warp-uniform: the ray queue may still hold rays
    uint32_t head = 0;      // warp-uniform read cursor of the triangle queue

    int32_t cur = -1;       // current inner node or -1
    int sp = 0;
    bool have = false, exh = false;
    RayT r; RayPre pre;
    int32_t slot = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.min_t = r.max_t = 0.f;
    pre.ix = pre.iy = pre.iz = 0.f; pre.degenerate = false;
    int32_t stack[TRB_STACK];
#ifdef TRB_COUNTERS
    uint32_t c_nodes = 0, c_tris = 0, c_exh = 0;
#endif
    sh.outst[lane] = 0; sh.flags[lane] = 0; sh.best[lane] = 0ull;
    if (lane == 0) sh.tail = 0;
    __syncwarp();

    auto push_range = [&](uint32_t first, uint32_t cnt) {   // this lane's ray must test triangles [first, first+cnt)
        TRB_CHECK(cnt >= 1u && cnt <= 8u && first + cnt <= a.n_tris);
        const uint32_t base = atomicAdd(&sh.tail, cnt);
        const uint32_t e0 = (uint32_t(lane) << 27) | first;
#pragma unroll
        for (uint32_t k = 0; k < 4u; ++k)   // straight-line predicated stores: a leaf holds <= 4 triangles ...
            if (k < cnt) sh.queue[(base + k) & (TRB_COOP_Q - 1u)] = e0 + k;
        if (cnt > 4u) {                     // ... only two sibling leaves hit together give more (cnt <= 8)
#pragma unroll
            for (uint32_t k = 4; k < 8u; ++k)
                if (k < cnt) sh.queue[(base + k) & (TRB_COOP_Q - 1u)] = e0 + k;
        }
        sh.outst[lane] += int32_t(cnt);   // only the owner adds; testers subtract atomically between __syncwarp()s
    };
    auto start_ray = [&]() {
        sh.best[lane] = (unsigned long long)__float_as_uint(r.max_t) << 32;   // bvh.h:263: intersection->t = ray.max_t
        sh.flags[lane] = 0; sp = 0; cur = -1;
        if (!(r.min_t > r.max_t)) {                                            // bvh.h:287
            if (a.bvh.root_ref >= 0) cur = a.bvh.root_ref;
            else { const uint32_t code = uint32_t(~a.bvh.root_ref); push_range(code >> 3, code & 7u); }
        }
    };

    while (true) {
        const bool idle = cur < 0 && (!have || sh.outst[lane] == 0);
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask == FULL || (__popc(idle_mask) >= TRB_FETCH_T && more)) {
            bool want = idle;
            if (idle && have) {
                const uint32_t fl = sh.flags[lane];
                const bool occ = (fl & 2u) != 0;
                if ((fl & 1u) && !exh && !(ANY && occ)) {
                    exh = true;   // ill-conditioned candidate met: visit what the reference visits (no distance pruning)
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
                        const unsigned long long b = sh.best[lane];
                        HitT h; h.t = r.max_t; h.u = 0.f; h.v = 0.f; h.pos = -1;
                        if (uint32_t(b >> 32) != __float_as_uint(r.max_t)) {
                            // the winner's (t, u, v): the same arithmetic on the same inputs as when it won
                            const uint32_t pos = uint32_t(b);
                            const F4* tp = a.bvh.tris + 3 * size_t(pos);
                            F4 q0 = ld4(tp), q1 = ld4(tp + 1), q2 = ld4(tp + 2);
                            bool ill2 = false; float t, u, v;
                            if (tri_test(q0, q1, q2, r, &t, &u, &v, &ill2)) { h.t = t; h.u = u; h.v = v; h.pos = int32_t(pos); }
                        }
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
            __syncwarp();
            if (__ballot_sync(FULL, have) == 0) break;
        }
        // ---- node phase ----
        if (cur >= 0) {
            TRB_CHECK(uint32_t(cur) < a.n_nodes);
            const F4* np = a.bvh.nodes + 4 * size_t(cur);
            const F4 na = ld4(np), nb = ld4(np + 1), nc = ld4(np + 2), nd = ld4(np + 3);
#ifdef TRB_COUNTERS
            ++c_nodes;
#endif
            BoxF f0 = box_fast(na.x, na.y, na.z, na.w, nb.x, nb.y, r, pre);
            BoxF f1 = box_fast(nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, r, pre);
            bool h0 = f0.pass, h1 = f1.pass;
            if (f0.undecided || f1.undecided) {
                if (f0.undecided) h0 = box_slow(na.x, na.y, na.z, na.w, nb.x, nb.y, r, pre, f0);
                if (f1.undecided) h1 = box_slow(nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, r, pre, f1);
            }
            const float e0 = f0.en, x0 = f0.ex, e1 = f1.en, x1 = f1.ex;
            const bool stop = ANY && (sh.flags[lane] & 2u) != 0;   // already occluded: nothing left to find
            if (!exh) {
                const float best = ANY ? r.max_t : __uint_as_float(uint32_t(sh.best[lane] >> 32));
                const float lim = best + 1e-3f * (1.f + fabsf(best));
                // a box wholly behind the origin: exit < -1e-3 (1 + |exit|)  <=>  exit < -1e-3 / (1 - 1e-3)
                h0 = h0 && !(e0 > lim) && !(x0 < -0.0010011f);
                h1 = h1 && !(e1 > lim) && !(x1 < -0.0010011f);
            }
            const int32_t r0 = f2i(nd.x), r1 = f2i(nd.y);
            const bool l0 = h0 && r0 < 0, l1 = h1 && r1 < 0;
            if ((l0 || l1) && !stop) {   // sibling leaves are adjacent in the primitive order: one contiguous range
                const uint32_t k0 = uint32_t(~r0), k1 = uint32_t(~r1);
                push_range(l0 ? (k0 >> 3) : (k1 >> 3), (l0 ? (k0 & 7u) : 0u) + (l1 ? (k1 & 7u) : 0u));
            }
            const bool i0 = h0 && r0 >= 0, i1 = h1 && r1 >= 0;
            if (stop) { cur = -1; sp = 0; }
            else if (i0 && i1) {
                const bool swap = !exh && (e1 < e0);  // nearer child first when pruning
                TRB_CHECK(sp < TRB_STACK);
                stack[sp++] = swap ? r0 : r1;
                cur = swap ? r1 : r0;
            } else if (i0) cur = r0;
            else if (i1) cur = r1;
            else cur = sp ? stack[--sp] : -1;
        }
        __syncwarp();
        // ---- triangle phases: full warps of queued tests; drain when no lane is traversing or the queue is filling up ----
        uint32_t count = sh.tail - head;
        TRB_CHECK(count <= TRB_COOP_Q);   // the per-warp queue never wraps onto unread entries
        const bool trav = __ballot_sync(FULL, cur >= 0) != 0;
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
            if (act) {
                TRB_CHECK(pos < a.n_tris);
                const F4* tp = a.bvh.tris + 3 * size_t(pos);
                F4 q0 = ld4(tp), q1 = ld4(tp + 1), q2 = ld4(tp + 2);
#ifdef TRB_COUNTERS
                ++c_tris;
#endif
                float t, u, v;
                bool ill = false;
                const bool hit = tri_test(q0, q1, q2, o, &t, &u, &v, &ill);
                if (hit) {
                    if (ANY) {
                        // bvh.h:299-307 + :351: occluded iff some accepted t <= max_t (min_t <= max_t holds for a traced ray)
                        if (t <= o.max_t) { atomicOr(&sh.flags[owner], 2u); ill = false; }
                    } else {
                        // (t, list position) lexicographic minimum == strict '<' in list order for any visiting order
                        atomicMin(&sh.best[owner], ((unsigned long long)__float_as_uint(t) << 32) | pos);
                    }
                }
                if (ill) atomicOr(&sh.flags[owner], 1u);
                { const int before = atomicSub(&sh.outst[owner], 1); TRB_CHECK(before >= 1); (void)before; }
            }
            head += nb_; count -= nb_;
            __syncwarp();
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
