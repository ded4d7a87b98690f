// tools/host_trace_check.cpp -- DEVELOPMENT AID, not part of the product or of the test-suite's product path.
// Compiles the __host__ __device__ traversal of csrc/kernels/trace.cuh for the CPU so that its hit semantics
// can be compared with the compiled reference on a box without a GPU (this container).  Build:
//   g++ -std=c++17 -O2 -fPIC -shared -ffp-contract=off tools/host_trace_check.cpp \
//       bidirectional-path-tracing_b200/build/{obj_loader,scene_build,camera,toml_exr,host_abi}.o -o /tmp/libhc.so
#include <cstring>
#include <string>
#include "../bidirectional-path-tracing_b200/csrc/host/tr_host.h"
#include "../bidirectional-path-tracing_b200/csrc/kernels/trace.cuh"

// TraceArgs::inv_max as trb_scene_upload sets it (kernels/trace.cuh: near_cull); HC_NEAR_CULL=0 = the old rule
static float hc_inv_max(const trb::FlatScene& fs) {
    if (const char* e = std::getenv("HC_NEAR_CULL")) if (e[0] == '0') return 0.f;
    float amax = 0.f;
    for (int k = 0; k < 3; ++k) amax = std::max(amax, std::max(std::fabs(fs.scene_min[k]), std::fabs(fs.scene_max[k])));
    return amax > 0.f ? 400.f / amax : 0.f;
}

extern "C" int hc_trace(const trb_scene_desc* s, const float* rays8, long n, int any, float* tuv, int* tri_out,
                        unsigned char* occ, unsigned long long* counters, int force_exhaustive) {
    trb::FlatScene fs;
    std::string err;
    if (!trb::flatten_scene(*s, fs, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
    trbk::BvhView bv{reinterpret_cast<const trbk::F4*>(fs.nodes.data()), reinterpret_cast<const trbk::F4*>(fs.tris.data()), fs.root_ref};
    unsigned long long nn = 0, nt = 0, nill = 0;
    for (long i = 0; i < n; ++i) {
        trbk::RayT r; memcpy(&r, rays8 + 8 * i, 32);
        trbk::HitT h; bool oc = false; uint32_t a = 0, b = 0;
        trbk::RayPre pre = trbk::ray_pre(r);
        if (force_exhaustive == 1) {
            if (any) trbk::traverse<true, false>(bv, r, pre, &h, &oc, &a, &b); else trbk::traverse<false, false>(bv, r, pre, &h, &oc, &a, &b);
        } else {
            bool ill = any ? trbk::traverse<true, true>(bv, r, pre, &h, &oc, &a, &b) : trbk::traverse<false, true>(bv, r, pre, &h, &oc, &a, &b);
            if (ill && !(any && oc) && force_exhaustive != 2) {
                ++nill;
                if (any) trbk::traverse<true, false>(bv, r, pre, &h, &oc, &a, &b); else trbk::traverse<false, false>(bv, r, pre, &h, &oc, &a, &b);
            }
        }
        nn += a; nt += b;
        if (any) occ[i] = oc;
        else {
            tuv[3 * i] = h.t; tuv[3 * i + 1] = h.u; tuv[3 * i + 2] = h.v;
            tri_out[i] = h.pos >= 0 ? fs.tris[h.pos].tri : -1;
        }
    }
    counters[0] = nn; counters[1] = nt; counters[2] = nill;
    return 0;
}

#include "../bidirectional-path-tracing_b200/csrc/kernels/trace_wide.cuh"

// the 4-wide upper tree (trace_wide.cuh) on the same rays; counters: nodes, boxes, tris, max stack, degenerate rays
extern "C" int hc_trace_wide(const trb_scene_desc* s, const float* rays8, long n, int any, float* tuv, int* tri_out,
                             unsigned char* occ, unsigned long long* counters) {
    static trb::FlatScene fs;            // cached across calls on the same scene pointer (development aid)
    static const void* cached = nullptr;
    std::string err;
    if (cached != s->positions) {
        if (!trb::flatten_scene(*s, fs, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        cached = s->positions;
    }
    trbk::BvhView bv{reinterpret_cast<const trbk::F4*>(fs.nodes.data()), reinterpret_cast<const trbk::F4*>(fs.tris.data()), fs.root_ref};
    trbk::WideView wv{reinterpret_cast<const trbk::F4*>(fs.wnodes.data()), reinterpret_cast<const trbk::F4*>(fs.tris.data()), fs.wnodes.empty() ? fs.root_ref : 0, hc_inv_max(fs)};
    trbk::WideCounters c{0, 0, 0, 0};
    unsigned long long nn = 0, nb = 0, nt = 0, ndeg = 0;
    for (long i = 0; i < n; ++i) {
        trbk::RayT r; memcpy(&r, rays8 + 8 * i, 32);
        trbk::HitT h; bool oc = false;
        if (trbk::ray_pre(r).degenerate) ++ndeg;
        c.nodes = c.boxes = c.tris = 0;
        if (any) trbk::trace_ray_wide<true>(wv, bv, r, &h, &oc, &c); else trbk::trace_ray_wide<false>(wv, bv, r, &h, &oc, &c);
        nn += c.nodes; nb += c.boxes; nt += c.tris;
        if (any) occ[i] = oc;
        else {
            tuv[3 * i] = h.t; tuv[3 * i + 1] = h.u; tuv[3 * i + 2] = h.v;
            tri_out[i] = h.pos >= 0 ? fs.tris[h.pos].tri : -1;
        }
    }
    counters[0] = nn; counters[1] = nb; counters[2] = nt; counters[3] = c.max_sp; counters[4] = ndeg;
    counters[5] = fs.wnodes.size(); counters[6] = fs.nodes.size(); counters[7] = (unsigned long long)fs.wide_depth;
    return 0;
}

// The claim the wide tree rests on (tr_host.h: DevNode4), checked by brute force against the reference's own box test restated
// in trace.cuh (box_exact = bvh.h:33-69): over EVERY leaf of the reference's tree, "the leaf box passes" must imply "every ancestor
// box passes" for a ray whose direction has three finite non-zero components.  out[0] non-degenerate rays, out[1] (ray, leaf)
// pairs whose leaf box passes, out[2] of those with a failing ancestor (must be 0), out[3] degenerate rays (excluded),
// out[4] the same violation count among the degenerate rays (informational: why they keep the binary tree).
#include <thread>
extern "C" int hc_leaf_implies_ancestors(const trb_scene_desc* s, const float* rays8, long n, unsigned long long* out) {
    trb::RefTree tree; std::string err;
    if (!trb::build_ref_tree(*s, tree, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
    const size_t nn = tree.nodes.size();
    std::vector<int32_t> parent(nn, -1);
    for (size_t i = 0; i < nn; ++i) if (tree.nodes[i].left >= 0) { parent[tree.nodes[i].left] = int32_t(i); parent[tree.nodes[i].right] = int32_t(i); }
    const unsigned T = std::max(1u, std::min(std::thread::hardware_concurrency(), 16u));
    std::vector<std::vector<unsigned long long>> part(T, std::vector<unsigned long long>(5, 0));
    auto work = [&](unsigned tid) {
        std::vector<unsigned char> pass(nn);
        auto& c = part[tid];
        for (long i = tid; i < n; i += T) {
            trbk::RayT r; memcpy(&r, rays8 + 8 * i, 32);
            const bool deg = trbk::ray_pre(r).degenerate;
            ++c[deg ? 3 : 0];
            for (size_t k = 1; k < nn; ++k) {   // the root's own box is never tested (bvh.h:276)
                const auto& b = tree.nodes[k];
                pass[k] = trbk::box_exact(b.bmin[0], b.bmin[1], b.bmin[2], b.bmax[0], b.bmax[1], b.bmax[2], r.ox, r.oy, r.oz, r.dx, r.dy, r.dz).hit;
            }
            pass[0] = 1;
            for (size_t k = 1; k < nn; ++k) {
                if (tree.nodes[k].left >= 0 || !pass[k]) continue;
                if (!deg) ++c[1];
                bool anc = true;
                for (int32_t a = parent[k]; a > 0; a = parent[a]) anc = anc && pass[a];
                if (!anc) ++c[deg ? 4 : 2];
            }
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < T; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& t : th) t.join();
    for (int k = 0; k < 5; ++k) { out[k] = 0; for (unsigned t = 0; t < T; ++t) out[k] += part[t][k]; }
    return 0;
}

// the host-built wide nodes (structure tests)
extern "C" int hc_wide_nodes(const trb_scene_desc* s, void* out, size_t* n) {
    trb::FlatScene fs; std::string err;
    if (!trb::flatten_scene(*s, fs, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
    if (out) memcpy(out, fs.wnodes.data(), std::min(*n, fs.wnodes.size()) * sizeof(trb::DevNode4));
    *n = fs.wnodes.size();
    return 0;
}

// ---- experiment: a binned-SAH upper tree over the reference's LEAVES (instead of the collapse of its midpoint tree) ----------------
// Same leaves, same leaf boxes, same primitive order; only the inner nodes differ.  Returns the same outputs as hc_trace_wide so that the
// hits can be compared and the node visits counted.  (tools/exp_sah_upper.py)
#include <algorithm>
#include <numeric>
namespace {
struct LeafRef { float bmin[3], bmax[3], cen[3]; uint32_t start, count; };
float area3(const float* mn, const float* mx) { return trb::wide_box_area(mn, mx); }
int32_t sah_build(std::vector<LeafRef>& L, size_t lo, size_t hi, trb::RefTree& out) {
    trb::RefTree::Node nd;
    for (int k = 0; k < 3; ++k) { nd.bmin[k] = 3.4e38f; nd.bmax[k] = -3.4e38f; }
    float cmn[3] = {3.4e38f, 3.4e38f, 3.4e38f}, cmx[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (size_t i = lo; i < hi; ++i)
        for (int k = 0; k < 3; ++k) {
            nd.bmin[k] = std::min(nd.bmin[k], L[i].bmin[k]); nd.bmax[k] = std::max(nd.bmax[k], L[i].bmax[k]);
            cmn[k] = std::min(cmn[k], L[i].cen[k]); cmx[k] = std::max(cmx[k], L[i].cen[k]);
        }
    nd.left = nd.right = -1; nd.start = L[lo].start; nd.count = L[lo].count;
    const int32_t me = int32_t(out.nodes.size());
    out.nodes.push_back(nd);
    if (hi - lo == 1) return me;
    // binned SAH, 16 bins per axis
    const int NB = 16;
    float best = 3.4e38f; int baxis = -1, bsplit = -1;
    for (int ax = 0; ax < 3; ++ax) {
        const float ext = cmx[ax] - cmn[ax];
        if (!(ext > 0.f)) continue;
        struct Bin { float mn[3], mx[3]; int n; } bins[NB];
        for (auto& b : bins) { b.n = 0; for (int k = 0; k < 3; ++k) { b.mn[k] = 3.4e38f; b.mx[k] = -3.4e38f; } }
        for (size_t i = lo; i < hi; ++i) {
            int b = std::min(NB - 1, int((L[i].cen[ax] - cmn[ax]) / ext * NB));
            bins[b].n++;
            for (int k = 0; k < 3; ++k) { bins[b].mn[k] = std::min(bins[b].mn[k], L[i].bmin[k]); bins[b].mx[k] = std::max(bins[b].mx[k], L[i].bmax[k]); }
        }
        float la[NB], ra[NB]; int ln[NB], rn[NB];
        float mn[3] = {3.4e38f, 3.4e38f, 3.4e38f}, mx[3] = {-3.4e38f, -3.4e38f, -3.4e38f}; int n = 0;
        for (int b = 0; b < NB; ++b) { n += bins[b].n; for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], bins[b].mn[k]); mx[k] = std::max(mx[k], bins[b].mx[k]); } la[b] = n ? area3(mn, mx) : 0.f; ln[b] = n; }
        for (int k = 0; k < 3; ++k) { mn[k] = 3.4e38f; mx[k] = -3.4e38f; } n = 0;
        for (int b = NB - 1; b >= 0; --b) { n += bins[b].n; for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], bins[b].mn[k]); mx[k] = std::max(mx[k], bins[b].mx[k]); } ra[b] = n ? area3(mn, mx) : 0.f; rn[b] = n; }
        for (int b = 0; b + 1 < NB; ++b) {
            if (!ln[b] || !rn[b + 1]) continue;
            const float cost = la[b] * ln[b] + ra[b + 1] * rn[b + 1];
            if (cost < best) { best = cost; baxis = ax; bsplit = b; }
        }
    }
    size_t mid;
    if (baxis < 0) mid = lo + (hi - lo) / 2;
    else {
        const float ext = cmx[baxis] - cmn[baxis];
        auto it = std::partition(L.begin() + lo, L.begin() + hi, [&](const LeafRef& l) { return std::min(NB - 1, int((l.cen[baxis] - cmn[baxis]) / ext * NB)) <= bsplit; });
        mid = size_t(it - L.begin());
        if (mid == lo || mid == hi) mid = lo + (hi - lo) / 2;
    }
    const int32_t l = sah_build(L, lo, mid, out);
    const int32_t r = sah_build(L, mid, hi, out);
    out.nodes[me].left = l; out.nodes[me].right = r;
    return me;
}
}  // namespace

extern "C" int hc_trace_wide_sah(const trb_scene_desc* s, const float* rays8, long n, int any, float* tuv, int* tri_out, unsigned char* occ,
                                 unsigned long long* counters) {
    static trb::FlatScene fs; static std::vector<trb::DevNode4> wn; static const void* cached = nullptr;
    std::string err;
    if (cached != s->positions) {
        if (!trb::flatten_scene(*s, fs, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        trb::RefTree ref;
        if (!trb::build_ref_tree(*s, ref, err)) return -1;
        std::vector<LeafRef> L;
        for (const auto& nd : ref.nodes)
            if (nd.left < 0) { LeafRef l; for (int k = 0; k < 3; ++k) { l.bmin[k] = nd.bmin[k]; l.bmax[k] = nd.bmax[k]; l.cen[k] = 0.5f * (nd.bmin[k] + nd.bmax[k]); } l.start = nd.start; l.count = nd.count; L.push_back(l); }
        trb::RefTree sah;
        sah_build(L, 0, L.size(), sah);
        int32_t depth = 0;
        trb::build_wide_nodes(sah, wn, depth);
        counters[7] = (unsigned long long)depth;
        cached = s->positions;
    }
    trbk::BvhView bv{reinterpret_cast<const trbk::F4*>(fs.nodes.data()), reinterpret_cast<const trbk::F4*>(fs.tris.data()), fs.root_ref};
    trbk::WideView wv{reinterpret_cast<const trbk::F4*>(wn.data()), reinterpret_cast<const trbk::F4*>(fs.tris.data()), wn.empty() ? fs.root_ref : 0, hc_inv_max(fs)};
    trbk::WideCounters c{0, 0, 0, 0};
    unsigned long long nn = 0, nb = 0, nt = 0;
    for (long i = 0; i < n; ++i) {
        trbk::RayT r; memcpy(&r, rays8 + 8 * i, 32);
        trbk::HitT h; bool oc = false;
        c.nodes = c.boxes = c.tris = 0;
        if (any) trbk::trace_ray_wide<true>(wv, bv, r, &h, &oc, &c); else trbk::trace_ray_wide<false>(wv, bv, r, &h, &oc, &c);
        nn += c.nodes; nb += c.boxes; nt += c.tris;
        if (any) occ[i] = oc; else { tuv[3 * i] = h.t; tuv[3 * i + 1] = h.u; tuv[3 * i + 2] = h.v; tri_out[i] = h.pos >= 0 ? fs.tris[h.pos].tri : -1; }
    }
    counters[0] = nn; counters[1] = nb; counters[2] = nt; counters[3] = c.max_sp; counters[5] = wn.size();
    return 0;
}
