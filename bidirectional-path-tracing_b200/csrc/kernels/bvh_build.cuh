// bvh_build.cuh -- the reference's BVH build (externals/bvh.h:147-247) on the device, with the reference's EXACT result:
// the same tree, the same build-reordered primitive list (which is the tie-break rank of the hit search, bvh.h:305) and the
// same pre-order node numbering as the host build of csrc/host/scene_build.cpp.  SURVEY 8(f) N1.
//
// The reference builds depth-first with an in-place partition (bvh.h:207-213):
//     mid = start;  for i in [start, end):  if centroid(a[i])[dim] < split:  swap(a[i], a[mid]); ++mid
// whose swap sequence defines the order of the right-hand part.  Sibling ranges are disjoint, so the tree can be built one
// LEVEL at a time, every range of a level in parallel, provided each range ends up in the order the sequential loop leaves:
//   * the elements that go left keep their relative order (each is appended at `mid`);
//   * the block of right-hand elements seen so far, B = a[mid, i), behaves as a queue: a right-hand element is appended at
//     its tail, a left-hand element moves B's FRONT element to the tail.  With f = the first right-hand position and
//     j = i - f, the sequence of tail insertions is  P[j] = a[f + j]  if a[f + j] goes right,  else  P[h(j)],
//     h(j) = number of left-hand elements in (f, f + j)  (< j: the queue head at that moment), and the right-hand part is
//     P[H .. end - f), H = number of left-hand elements after f.  h comes from ONE prefix sum of the left/right flags over the
//     whole array; a destination resolves its source by following h until it lands on a right-hand element.  One step moves
//     the position back by r = the number of right-hand elements before it, which is constant inside a RUN of left-hand
//     elements, so the steps of a run are taken at once (the run's start comes from a running maximum of the right-hand
//     positions, a second scan): a mesh split across its rows gives runs of a whole row - thousands of single steps, a handful
//     of run jumps.
// Per level: bounds of every range (warp-aggregated float atomics; min / max are order-independent), split decision, flags,
// prefix sum (CUB), children, permutation.  Then subtree sizes bottom-up, pre-order numbers top-down, the 64 B two-child
// nodes and the per-triangle records gathered in the final order.  All arithmetic as the host build does it (fp32, no FMA).
#pragma once
#include <cub/device/device_scan.cuh>

#include "../host/tr_host.h"

namespace trbk {

struct alignas(16) BuildNode {
    uint32_t start, end;
    int32_t parent, left, right;   // ids in creation (level) order; left < 0: leaf
    int32_t dim; float split;
    uint32_t first_r;              // position of the first element that goes right
    uint32_t mid;
    int32_t permute;               // 0: leaf, or the median-index fallback of bvh.h:216-218 (order untouched)
    float bmin[3], bmax[3], cmin[3], cmax[3];
    int32_t sub_nodes, sub_inner;  // subtree sizes (all nodes / inner nodes)
    int32_t pre, inner_pre;        // pre-order number among all nodes / among inner nodes
    int32_t pad[2];
};
static_assert(sizeof(BuildNode) == 112, "BuildNode");

struct BuildScene {   // raw scene arrays on the device (what trb_scene_desc points at on the host)
    const float* positions; const float* normals;
    const int32_t* tri_v; const int32_t* tri_n; const int32_t* tri_mat; const int32_t* tri_shape;
    const int32_t* shape_first; const int32_t* shape_to_emitter;
    const trb::DevMaterial* mats;
    const float* texcoords; const int32_t* tri_t;
    uint32_t n;
};

// float min / max through integer atomics: exact, order-independent (a -0 / +0 tie may keep either zero; they compare equal).
// The value is read first (from L2; a stale read only costs a needless atomic, min / max are monotone): after the first few
// elements of a range almost none moves its bounds, so almost no atomic is issued.
__device__ __forceinline__ void atomic_min_f(float* a, float v) {
    if (!(v < __ldcg(a))) return;
    if (v >= 0.f) atomicMin(reinterpret_cast<int*>(a), __float_as_int(v)); else atomicMax(reinterpret_cast<unsigned*>(a), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float* a, float v) {
    if (!(v > __ldcg(a))) return;
    if (v >= 0.f) atomicMax(reinterpret_cast<int*>(a), __float_as_int(v)); else atomicMin(reinterpret_cast<unsigned*>(a), __float_as_uint(v));
}

// accel.h:71-106: box by min / max over the three vertices, centroid = ((v0 + v1) + v2) / 3.0f
__global__ void k_bb_prep(BuildScene s, float4* __restrict__ tmin, float4* __restrict__ tmax, float4* __restrict__ tcen,
                          int32_t* __restrict__ ord, int32_t* __restrict__ range_of) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= s.n) return;
    const float* a = s.positions + 3 * size_t(s.tri_v[3 * size_t(t)]);
    const float* b = s.positions + 3 * size_t(s.tri_v[3 * size_t(t) + 1]);
    const float* c = s.positions + 3 * size_t(s.tri_v[3 * size_t(t) + 2]);
    float mn[3], mx[3], ce[3];
    for (int k = 0; k < 3; ++k) {
        mn[k] = fminf(fminf(a[k], b[k]), c[k]); mx[k] = fmaxf(fmaxf(a[k], b[k]), c[k]);
        ce[k] = __fdiv_rn((a[k] + b[k]) + c[k], 3.0f);
    }
    tmin[t] = make_float4(mn[0], mn[1], mn[2], 0.f); tmax[t] = make_float4(mx[0], mx[1], mx[2], 0.f); tcen[t] = make_float4(ce[0], ce[1], ce[2], 0.f);
    ord[t] = int32_t(t);   // objects are created shape-major (accel.h:116-120)
    range_of[t] = 0;
}

__device__ __forceinline__ void init_node(BuildNode& nd, uint32_t start, uint32_t end, int32_t parent) {
    const float inf = __int_as_float(0x7f800000);
    nd.start = start; nd.end = end; nd.parent = parent; nd.left = -1; nd.right = -1; nd.dim = 0; nd.split = 0.f;
    nd.first_r = 0xffffffffu; nd.mid = start; nd.permute = 0;
    for (int k = 0; k < 3; ++k) { nd.bmin[k] = inf; nd.bmax[k] = -inf; nd.cmin[k] = inf; nd.cmax[k] = -inf; }
    nd.sub_nodes = 1; nd.sub_inner = 0; nd.pre = 0; nd.inner_pre = 0; nd.pad[0] = nd.pad[1] = 0;
}
__global__ void k_bb_root(BuildNode* nodes, uint32_t n, uint32_t* counter) { init_node(nodes[0], 0u, n, -1); *counter = 1u; }

// bounds of every range of the level (bvh.h:171-178): triangle boxes and centroids
__global__ void k_bb_bounds(BuildNode* __restrict__ nodes, int32_t lb, const int32_t* __restrict__ ord, const int32_t* __restrict__ range_of,
                            const float4* __restrict__ tmin, const float4* __restrict__ tmax, const float4* __restrict__ tcen, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < n;
    const int32_t node = in ? range_of[i] : -1;
    const bool act = in && node >= lb;
    float v[12];
    if (act) {
        const int32_t t = ord[i];
        const float4 a = tmin[t], b = tmax[t], c = tcen[t];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = b.x; v[4] = b.y; v[5] = b.z; v[6] = c.x; v[7] = c.y; v[8] = c.z; v[9] = c.x; v[10] = c.y; v[11] = c.z;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, act ? node : -1);
    if (peers == 0xffffffffu && act) {   // the whole warp lies in one range (the rule at the top of the tree): reduce, then 12 atomics
#pragma unroll
        for (int k = 0; k < 12; ++k) {
            const bool is_min = k < 3 || (k >= 6 && k < 9);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) { const float o = __shfl_xor_sync(0xffffffffu, v[k], d); v[k] = is_min ? fminf(v[k], o) : fmaxf(v[k], o); }
        }
        if ((threadIdx.x & 31) != 0) return;
    } else if (!act) return;
    BuildNode& nd = nodes[node];
    for (int k = 0; k < 3; ++k) {
        atomic_min_f(&nd.bmin[k], v[k]); atomic_max_f(&nd.bmax[k], v[3 + k]);
        atomic_min_f(&nd.cmin[k], v[6 + k]); atomic_max_f(&nd.cmax[k], v[9 + k]);
    }
}

// leaf test (leafSize 4, bvh.h:122,183), split axis with maxDimension's y/z quirk (bvh.h:82-87), midpoint (bvh.h:203)
__global__ void k_bb_split(BuildNode* __restrict__ nodes, int32_t lb, int32_t le) {
    int32_t k = lb + int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (k >= le) return;
    BuildNode& nd = nodes[k];
    if (nd.end - nd.start <= 4u) { nd.dim = -1; return; }
    const float ex = nd.cmax[0] - nd.cmin[0], ey = nd.cmax[1] - nd.cmin[1], ez = nd.cmax[2] - nd.cmin[2];
    int dim = 0;
    if (ey > ex) dim = 1;
    if (ez > ey) dim = 2;
    nd.dim = dim;
    nd.split = .5f * (nd.cmin[dim] + nd.cmax[dim]);
}

__global__ void k_bb_flags(BuildNode* __restrict__ nodes, int32_t lb, const int32_t* __restrict__ ord, const int32_t* __restrict__ range_of,
                           const float4* __restrict__ tcen, uint32_t* __restrict__ is_l, uint32_t* __restrict__ r_pos, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    uint32_t l = 0;
    if (i < n) {
        const int32_t node = range_of[i];
        if (node >= lb) {
            const int dim = nodes[node].dim;
            if (dim >= 0) {
                const float4 c = tcen[ord[i]];
                const float x = dim == 0 ? c.x : (dim == 1 ? c.y : c.z);
                l = x < nodes[node].split ? 1u : 0u;
                if (!l && i < __ldcg(&nodes[node].first_r)) atomicMin(&nodes[node].first_r, i);
            }
        }
    }
    is_l[i] = l;   // entry n is a zero so that the exclusive sum has n + 1 entries
    r_pos[i] = l ? 0u : i;   // running maximum = the nearest right-hand (or inactive) position at or before i
}

// mid (with the median-index fallback of bvh.h:216-218) and the two children (bvh.h:221-227)
__global__ void k_bb_children(BuildNode* __restrict__ nodes, int32_t lb, int32_t le, const uint32_t* __restrict__ pref, uint32_t* __restrict__ counter) {
    int32_t k = lb + int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (k >= le) return;
    BuildNode& nd = nodes[k];
    if (nd.dim < 0) return;
    const uint32_t cnt = nd.end - nd.start, nl = pref[nd.end] - pref[nd.start];
    if (nl == 0u || nl == cnt) { nd.mid = nd.start + cnt / 2u; nd.permute = 0; }
    else { nd.mid = nd.start + nl; nd.permute = 1; }
    const uint32_t base = atomicAdd(counter, 2u);
    nd.left = int32_t(base); nd.right = int32_t(base + 1u);
    init_node(nodes[base], nd.start, nd.mid, k);
    init_node(nodes[base + 1u], nd.mid, nd.end, k);
}

// the order the reference's swap loop leaves (see the file header), and the range each position belongs to on the next level
__global__ void k_bb_permute(const BuildNode* __restrict__ nodes, int32_t lb, const int32_t* __restrict__ ord, int32_t* __restrict__ ord_next,
                             int32_t* __restrict__ range_of, const uint32_t* __restrict__ is_l, const uint32_t* __restrict__ pref,
                             const uint32_t* __restrict__ last_r, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t node = range_of[i];
    if (node < lb || nodes[node].dim < 0) { ord_next[i] = ord[i]; return; }
    const BuildNode& nd = nodes[node];
    range_of[i] = i < nd.mid ? nd.left : nd.right;
    if (!nd.permute) { ord_next[i] = ord[i]; return; }
    if (is_l[i]) ord_next[nd.start + (pref[i] - pref[nd.start])] = ord[i];   // stable on the left
    if (i >= nd.mid) {                                                       // destination i of the right-hand part
        const uint32_t f = nd.first_r, pf = pref[f];
        const uint32_t H = pref[nd.end] - pf;
        uint32_t pos = f + H + (i - nd.mid);
        while (is_l[pos]) {
            const uint32_t r = (pos - f) - (pref[pos] - pf);   // right-hand elements in [f, pos): one step is pos -= r
            const uint32_t b = last_r[pos] + 1u;               // first position of this run of left-hand elements
            pos -= ((pos - b) / r + 1u) * r;                   // all the steps that stay inside the run, and the one that leaves it
        }
        ord_next[i] = ord[pos];
    }
}

// subtree sizes, one level at a time from the deepest
__global__ void k_bb_sizes(BuildNode* __restrict__ nodes, int32_t lb, int32_t le) {
    int32_t k = lb + int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (k >= le) return;
    BuildNode& nd = nodes[k];
    if (nd.left < 0) { nd.sub_nodes = 1; nd.sub_inner = 0; return; }
    nd.sub_nodes = 1 + nodes[nd.left].sub_nodes + nodes[nd.right].sub_nodes;
    nd.sub_inner = 1 + nodes[nd.left].sub_inner + nodes[nd.right].sub_inner;
}
// pre-order numbers (the reference's flat array: left child = i + 1, right child = i + rightOffset), one level at a time from the root
__global__ void k_bb_number(BuildNode* __restrict__ nodes, int32_t lb, int32_t le) {
    int32_t k = lb + int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (k >= le) return;
    const BuildNode& nd = nodes[k];
    if (nd.left < 0) return;
    BuildNode& l = nodes[nd.left]; BuildNode& r = nodes[nd.right];
    l.pre = nd.pre + 1; l.inner_pre = nd.inner_pre + 1;
    r.pre = nd.pre + 1 + l.sub_nodes; r.inner_pre = nd.inner_pre + 1 + l.sub_inner;
}

// the 64 B two-child node of the traversal kernels (tr_host.h: DevNode), inner nodes only, numbered in pre-order
__global__ void k_bb_emit_nodes(const BuildNode* __restrict__ nodes, int32_t total, trb::DevNode* __restrict__ out) {
    int32_t k = int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (k >= total) return;
    const BuildNode& nd = nodes[k];
    if (nd.left < 0) return;
    const BuildNode& l = nodes[nd.left]; const BuildNode& r = nodes[nd.right];
    trb::DevNode d;
    for (int c = 0; c < 3; ++c) { d.c0min[c] = l.bmin[c]; d.c0max[c] = l.bmax[c]; d.c1min[c] = r.bmin[c]; d.c1max[c] = r.bmax[c]; }
    d.ref0 = l.left >= 0 ? l.inner_pre : ~int32_t((l.start << 3) | (l.end - l.start));
    d.ref1 = r.left >= 0 ? r.inner_pre : ~int32_t((r.start << 3) | (r.end - r.start));
    d.pad0 = 0; d.pad1 = 0;
    out[nd.inner_pre] = d;
}

// the leaves of the finished tree in primitive order (a leaf's ordinal = its pre-order number minus the inner nodes before it): the
// input of the SAH upper tree, which is built on the host cores (trb::build_sah_wide)
__global__ void k_bb_emit_leaves(const BuildNode* __restrict__ nodes, int32_t total, trb::LeafBox* __restrict__ out) {
    int32_t k = int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (k >= total) return;
    const BuildNode& nd = nodes[k];
    if (nd.left >= 0) return;
    trb::LeafBox l;
    for (int c = 0; c < 3; ++c) { l.bmin[c] = nd.bmin[c]; l.bmax[c] = nd.bmax[c]; }
    l.start = nd.start; l.count = nd.end - nd.start;
    out[nd.pre - nd.inner_pre] = l;
}

// ---- the 4-wide upper tree (tr_host.h: DevNode4), level by level: the device twin of trb::build_wide_nodes ----
// children of every wide node of a level: start from the binary node's two children and open, while fewer than four, the inner
// child of largest surface area (ties: the first); an opened child is replaced in place by its two children
__global__ void k_wide_expand(const BuildNode* __restrict__ nodes, const int32_t* __restrict__ level, int32_t n_level, int4* __restrict__ slots,
                              uint32_t* __restrict__ n_inner) {
    int32_t i = int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (i > n_level) return;
    if (i == n_level) { n_inner[i] = 0u; return; }   // the exclusive sum has n_level + 1 entries
    const BuildNode& b = nodes[level[i]];
    int32_t sl[4] = {b.left, b.right, -1, -1};
    int n = 2;
    while (n < 4) {
        int best = -1; float best_a = 0.f;
        for (int k = 0; k < n; ++k) {
            const BuildNode& c = nodes[sl[k]];
            if (c.left < 0) continue;
            const float a = trb::wide_box_area(c.bmin, c.bmax);
            if (best < 0 || a > best_a) { best = k; best_a = a; }
        }
        if (best < 0) break;
        const BuildNode& c = nodes[sl[best]];
        for (int k = n; k > best + 1; --k) sl[k] = sl[k - 1];
        sl[best] = c.left; sl[best + 1] = c.right;
        ++n;
    }
    uint32_t cnt = 0;
    for (int k = 0; k < 4; ++k) if (sl[k] >= 0 && nodes[sl[k]].left >= 0) ++cnt;
    slots[i] = make_int4(sl[0], sl[1], sl[2], sl[3]);
    n_inner[i] = cnt;
}
// the node records of the level and the binary nodes of the next one: inner children are numbered in (parent order, slot order)
__global__ void k_wide_emit(const BuildNode* __restrict__ nodes, int32_t n_level, const int4* __restrict__ slots, const uint32_t* __restrict__ first_inner,
                            uint32_t base, trb::DevNode4* __restrict__ out, int32_t* __restrict__ next_level) {
    int32_t i = int32_t(blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= n_level) return;
    const int4 s4 = slots[i];
    const int32_t sl[4] = {s4.x, s4.y, s4.z, s4.w};
    const uint32_t next_base = base + uint32_t(n_level);
    uint32_t at = first_inner[i];
    trb::DevNode4 w;
    for (int k = 0; k < 4; ++k) {
        trb::DevNode4::Half& hf = w.half[k >> 1];
        float* mn = (k & 1) ? hf.c1min : hf.c0min; float* mx = (k & 1) ? hf.c1max : hf.c0max;
        int32_t& ref = (k & 1) ? hf.ref1 : hf.ref0;
        hf.pad0 = 0; hf.pad1 = 0;
        if (sl[k] < 0) { for (int a = 0; a < 3; ++a) { mn[a] = 0.f; mx[a] = 0.f; } ref = TRB_WIDE_EMPTY; continue; }
        const BuildNode& c = nodes[sl[k]];
        for (int a = 0; a < 3; ++a) { mn[a] = c.bmin[a]; mx[a] = c.bmax[a]; }
        if (c.left >= 0) { next_level[at] = sl[k]; ref = int32_t(next_base + at); ++at; }
        else ref = ~int32_t((c.start << 3) | (c.end - c.start));
    }
    out[base + uint32_t(i)] = w;
}

// the per-triangle records in the final primitive order (what flatten_geometry gathers on the host)
__global__ void k_bb_emit_tris(BuildScene s, const int32_t* __restrict__ ord, trb::DevTri* __restrict__ tris, trb::DevShade* __restrict__ shade,
                               trb::DevTriUV* __restrict__ tri_uv) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= s.n) return;
    const int32_t t = ord[p];
    const float* a = s.positions + 3 * size_t(s.tri_v[3 * size_t(t)]);
    const float* b = s.positions + 3 * size_t(s.tri_v[3 * size_t(t) + 1]);
    const float* c = s.positions + 3 * size_t(s.tri_v[3 * size_t(t) + 2]);
    const float* n0 = s.normals + 3 * size_t(s.tri_n[3 * size_t(t)]);
    const float* n1 = s.normals + 3 * size_t(s.tri_n[3 * size_t(t) + 1]);
    const float* n2 = s.normals + 3 * size_t(s.tri_n[3 * size_t(t) + 2]);
    const int32_t mat = s.tri_mat[t], shape = s.tri_shape[t];
    trb::DevTri dt;
    for (int k = 0; k < 3; ++k) { dt.v0[k] = a[k]; dt.e1[k] = b[k] - a[k]; dt.e2[k] = c[k] - a[k]; }
    dt.tri = t;
    const float e1 = (dt.e1[0] * dt.e1[0] + dt.e1[1] * dt.e1[1]) + dt.e1[2] * dt.e1[2];
    const float e2 = (dt.e2[0] * dt.e2[0] + dt.e2[1] * dt.e2[1]) + dt.e2[2] * dt.e2[2];
    dt.illk = 1e-6f * e1 * e2;
    const trb::DevMaterial& m = s.mats[mat];
    dt.flags = m.cls | (m.emissive ? 8 : 0);
    tris[p] = dt;
    trb::DevShade ds;
    for (int k = 0; k < 3; ++k) { ds.v0[k] = a[k]; ds.v1[k] = b[k]; ds.v2[k] = c[k]; ds.n0[k] = n0[k]; ds.n1[k] = n1[k]; ds.n2[k] = n2[k]; }
    ds.shape = shape; ds.prim = t - s.shape_first[shape]; ds.mat = mat; ds.emitter = s.shape_to_emitter[shape]; ds.tri = t; ds.pad = 0;
    shade[p] = ds;
    if (tri_uv) {
        trb::DevTriUV uv = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}};
        if (m.tex) {
            for (int k = 0; k < 2; ++k) {
                uv.st0[k] = s.texcoords[2 * size_t(s.tri_t[3 * size_t(t)]) + k]; uv.st1[k] = s.texcoords[2 * size_t(s.tri_t[3 * size_t(t) + 1]) + k];
                uv.st2[k] = s.texcoords[2 * size_t(s.tri_t[3 * size_t(t) + 2]) + k];
            }
        }
        tri_uv[p] = uv;
    }
}

}  // namespace trbk
