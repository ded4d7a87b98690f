// scene_build.cpp -- the tail of Scene::load on the host (reference: src/core/renderer.cpp:259-339,
// src/core/accel.h:115-123, externals/bvh.h:147-247) and the flattening into the device layout.
//
// Why the reference's BVH *topology* is kept (DESIGN.md "BVH"): the reference's traversal is exhaustive
// along the infinite line and its answer is "the min-t triangle among those whose ancestors' boxes all
// pass BBox::intersect; ties -> first in the build-reordered list" (bvh.h:259-352).  Bit-exact
// hit ids need (a) that primitive order and (b) the same set of boxes, so the device tree has the same
// partition; what changes is the node layout (children boxes in the parent, 64 B, float4 loads) and the
// traversal (ordered, pruned, stack in registers/shared memory).
#include <algorithm>
#include <array>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <utility>
#include <vector>

#include "tr_host.h"

namespace trb {

namespace {
struct V3 { float x, y, z; };
inline V3 vtx(const trb_scene_desc& s, int32_t i) { return {s.positions[3 * i], s.positions[3 * i + 1], s.positions[3 * i + 2]}; }
inline V3 nrm(const trb_scene_desc& s, int32_t i) { return {s.normals[3 * i], s.normals[3 * i + 1], s.normals[3 * i + 2]}; }
inline float lum(float r, float g, float b) { return (r * 0.212671f + g * 0.715160f) + b * 0.072169f; }  // math.h:56-58 (glm dot order)
}  // namespace

size_t FlatScene::bytes() const {
    return n_nodes * sizeof(DevNode) + n_wnodes * sizeof(DevNode4) + n_tris * sizeof(DevTri) + n_tris * sizeof(DevShade) +
           materials.size() * sizeof(DevMaterial) + emitters.size() * sizeof(DevEmitter) +
           emitter_tris.size() * sizeof(DevEmitterTri) + emitter_cdf.size() * sizeof(float) + emitter_sphere.size() * sizeof(float) +
           tex_info.size() * sizeof(DevTexInfo) + n_tri_uv * sizeof(DevTriUV) + texels.size() * sizeof(float);
}

void texture_max_average(const trb_texture& t, float mx[3], float avg[3]) {
    const size_t n = size_t(t.width) * t.height;
    float s[3] = {0.f, 0.f, 0.f};
    for (int k = 0; k < 3; ++k) mx[k] = -std::numeric_limits<float>::min();  // as written (core.h:554-556)
    for (size_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) { s[k] += t.rgb[3 * i + k]; mx[k] = std::max(mx[k], t.rgb[3 * i + k]); }
    const float scale = float(1.0 / (double(3 * n) / 3.0));
    for (int k = 0; k < 3; ++k) avg[k] = s[k] * scale;
}

// BSDF constructors: diffuse.h:20-32, phong.h:25-52, mixture.h:25-52, perfectmirror.h:18-26, glass.h:20-31;
// class selection by illum: renderer.cpp:259-271.
bool derive_materials(const trb_scene_desc& s, std::vector<DevMaterial>& out, std::string& err) {
    out.resize(s.n_materials);
    for (size_t i = 0; i < s.n_materials; ++i) {
        const trb_material& m = s.materials[i];
        DevMaterial d;
        std::memset(&d, 0, sizeof(d));
        {
            const int32_t kd_t = m.tex & 0xffff, ks_t = (m.tex >> 16) & 0xffff;
            if (size_t(kd_t) > s.n_textures || size_t(ks_t) > s.n_textures || (m.tex && !s.textures)) { err = "material " + std::to_string(i) + ": texture index out of range"; return false; }
            for (int32_t t : {kd_t, ks_t})
                if (t && (!s.textures[t - 1].rgb || s.textures[t - 1].width <= 0 || s.textures[t - 1].height <= 0)) { err = "empty texture"; return false; }
        }
        for (int k = 0; k < 3; ++k) { d.kd[k] = m.kd[k]; d.ks[k] = m.ks[k]; d.ke[k] = m.ke[k]; d.tf[k] = m.tf[k]; }
        d.ns = m.ns; d.ni = m.ni;
        d.emissive = (m.ke[0] != 0.f || m.ke[1] != 0.f || m.ke[2] != 0.f) ? 1 : 0;
        // which bitmaps the BSDF class reads: Diffuse map_Kd (diffuse.h:23-26), Phong / Mixture both (phong.h:28-36), mirror and glass none
        d.tex = m.illum == 7 ? (m.tex & 0xffff) : (m.illum == 3 || m.illum == 6) ? 0 : m.tex;
        d.scale = 1.f; d.spec_weight = 0.f;
        const uint32_t EDiffuseReflection = 0x2, EGlossyReflection = 0x8, EDeltaReflection = 0x20, EDeltaTransmission = 0x40;
        switch (m.illum) {
            case 7: d.cls = MAT_DIFFUSE; d.type = EDiffuseReflection; break;
            case 3: d.cls = MAT_MIRROR; d.type = EDeltaReflection; break;
            case 6: d.cls = MAT_GLASS; d.type = EDeltaReflection | EDeltaTransmission; break;
            case 5:
                err = "material " + std::to_string(i) + ": illum 5 has no BSDF in the reference (renderer.cpp:259-270)";
                return false;
            default: {
                d.cls = (m.illum == 8) ? MAT_MIXTURE : MAT_PHONG;
                d.type = EGlossyReflection | EDiffuseReflection;
                // energy-conservation scale and lobe-selection weight (phong.h:41-47)
                // with a bitmap the constructor takes the texture's maximum and average (Texture::getMax / getAverage)
                float kd_max[3] = {m.kd[0], m.kd[1], m.kd[2]}, kd_avg[3] = {m.kd[0], m.kd[1], m.kd[2]};
                float ks_max[3] = {m.ks[0], m.ks[1], m.ks[2]}, ks_avg[3] = {m.ks[0], m.ks[1], m.ks[2]};
                if (m.tex & 0xffff) texture_max_average(s.textures[(m.tex & 0xffff) - 1], kd_max, kd_avg);
                if (m.tex >> 16) texture_max_average(s.textures[(m.tex >> 16) - 1], ks_max, ks_avg);
                float mx = ks_max[0] + kd_max[0], my = ks_max[1] + kd_max[1], mz = ks_max[2] + kd_max[2];
                float actual = std::max(std::max(mx, my), mz);
                d.scale = actual > 1.0f ? 0.99f * (1.0f / actual) : 1.0f;
                float dAvg = lum(kd_avg[0] * d.scale, kd_avg[1] * d.scale, kd_avg[2] * d.scale);
                float sAvg = lum(ks_avg[0] * d.scale, ks_avg[1] * d.scale, ks_avg[2] * d.scale);
                d.spec_weight = sAvg / (dAvg + sAvg);
            }
        }
        out[i] = d;
    }
    return true;
}

// Emitters: one per shape whose FIRST face material is emissive (renderer.cpp:279-291); face-area CDF and
// total area as Scene::getShapeArea / Distribution1D do (renderer.cpp:317-339, math.h:81-112).
bool derive_emitters(const trb_scene_desc& s, const std::vector<DevMaterial>& mats, std::vector<DevEmitter>& em,
                     std::vector<DevEmitterTri>& etris, std::vector<float>& cdf, std::vector<int32_t>& shape_to_emitter,
                     std::string& err) {
    em.clear(); etris.clear(); cdf.clear();
    shape_to_emitter.assign(s.n_shapes, -1);
    size_t t = 0;
    for (size_t sh = 0; sh < s.n_shapes; ++sh) {
        size_t first = t;
        while (t < s.n_tris && size_t(s.tri_shape[t]) == sh) ++t;
        size_t nf = t - first;
        if (nf == 0) { err = "shape without faces"; return false; }
        int32_t m0 = s.tri_mat[first];
        if (m0 < 0 || size_t(m0) >= s.n_materials) { err = "face without a material (usemtl missing)"; return false; }
        if (!mats[m0].emissive) continue;
        DevEmitter e;
        std::memset(&e, 0, sizeof(e));
        e.shape = int32_t(sh);
        e.first_face = int32_t(etris.size());
        e.n_faces = int32_t(nf);
        e.cdf_offset = int32_t(cdf.size());
        for (int k = 0; k < 3; ++k) e.radiance[k] = mats[m0].ke[k];
        size_t c0 = cdf.size();
        cdf.push_back(0.f);
        for (size_t f = first; f < t; ++f) {
            V3 a = vtx(s, s.tri_v[3 * f]), b = vtx(s, s.tri_v[3 * f + 1]), c = vtx(s, s.tri_v[3 * f + 2]);
            V3 e1{b.x - a.x, b.y - a.y, b.z - a.z}, e2{c.x - a.x, c.y - a.y, c.z - a.z};
            V3 cr{e1.y * e2.z - e2.y * e1.z, e1.z * e2.x - e2.z * e1.x, e1.x * e2.y - e2.x * e1.y};  // glm::cross order
            float area = 0.5f * std::sqrt(cr.x * cr.x + cr.y * cr.y + cr.z * cr.z);
            cdf.push_back(cdf.back() + area);
            DevEmitterTri et;
            std::memset(&et, 0, sizeof(et));
            V3 n0 = nrm(s, s.tri_n[3 * f]), n1 = nrm(s, s.tri_n[3 * f + 1]), n2 = nrm(s, s.tri_n[3 * f + 2]);
            et.v0[0] = a.x; et.v0[1] = a.y; et.v0[2] = a.z; et.v1[0] = b.x; et.v1[1] = b.y; et.v1[2] = b.z;
            et.v2[0] = c.x; et.v2[1] = c.y; et.v2[2] = c.z;
            et.n0[0] = n0.x; et.n0[1] = n0.y; et.n0[2] = n0.z; et.n1[0] = n1.x; et.n1[1] = n1.y; et.n1[2] = n1.z;
            et.n2[0] = n2.x; et.n2[1] = n2.y; et.n2[2] = n2.z;
            etris.push_back(et);
        }
        e.area = cdf.back();
        float sum = cdf.back();
        for (size_t k = c0; k < cdf.size(); ++k) cdf[k] /= sum;  // Distribution1D::normalize divides every entry
        shape_to_emitter[sh] = int32_t(em.size());
        em.push_back(e);
    }
    // The reference looks the emitter up by shape whenever a face's material is emissive (bdpt.h:79-81,
    // integrator.cpp:53-58) and walks off the end of the list when the shape is not an emitter.
    for (size_t f = 0; f < s.n_tris; ++f) {
        int32_t m = s.tri_mat[f];
        if (m < 0 || size_t(m) >= s.n_materials) { err = "face without a material (usemtl missing)"; return false; }
        if (mats[m].emissive && shape_to_emitter[s.tri_shape[f]] < 0) {
            err = "emissive face in a shape whose first face is not emissive: undefined in the reference";
            return false;
        }
    }
    return true;
}

void shape_spheres(const trb_scene_desc& s, std::vector<float>& out4) {
    out4.assign(4 * s.n_shapes, 0.f);
    size_t t = 0;
    for (size_t sh = 0; sh < s.n_shapes; ++sh) {
        float cx = 0.f, cy = 0.f, cz = 0.f, mx = -std::numeric_limits<float>::infinity();
        size_t first = t;
        for (; t < s.n_tris && size_t(s.tri_shape[t]) == sh; ++t)
            for (int k = 0; k < 3; ++k) {
                V3 p = vtx(s, s.tri_v[3 * t + k]);
                cx += p.x; cy += p.y; cz += p.z;
                mx = std::max(mx, p.x);
            }
        const float n = float(3 * (t - first));
        cx /= n; cy /= n; cz /= n;
        out4[4 * sh] = cx; out4[4 * sh + 1] = cy; out4[4 * sh + 2] = cz; out4[4 * sh + 3] = mx - cx;
    }
}

// ---- reference-order BVH build ---------------------------------------------------------------------
// Host threads for the build: the arithmetic and the resulting order are those of the sequential algorithm (every thread
// works on a disjoint primitive range); only the wall time changes (10 M triangles: 1.9 s -> see DESIGN section 12).
static unsigned build_threads() {
    if (const char* e = std::getenv("TRB_BUILD_THREADS")) { int v = std::atoi(e); if (v >= 1) return unsigned(std::min(v, 64)); }
    return std::max(1u, std::min(std::thread::hardware_concurrency(), 16u));
}
template <class F>
static void parallel_for(size_t n, size_t grain, F fn) {  // fn(begin, end)
    const unsigned T = build_threads();
    if (T <= 1 || n < 2 * grain) { fn(size_t(0), n); return; }
    const size_t chunks = std::min<size_t>(T, (n + grain - 1) / grain);
    std::vector<std::thread> th;
    for (size_t c = 1; c < chunks; ++c) th.emplace_back([=]() { fn(n * c / chunks, n * (c + 1) / chunks); });
    fn(size_t(0), n / chunks);
    for (auto& t : th) t.join();
}

bool build_ref_tree(const trb_scene_desc& s, RefTree& out, std::string& err) {
    const size_t n = s.n_tris;
    if (n == 0) { err = "empty scene"; return false; }
    // per-triangle bounds and centroid, computed as AcceleratorBVH::BVHNode::{getBBox,getCentroid} do
    // (accel.h:71-106): box by min/max over the vertices, centroid = ((v0 + v1) + v2) / 3.0f
    std::vector<float> bmin(3 * n), bmax(3 * n), cen(3 * n);
    parallel_for(n, 1 << 16, [&](size_t t0, size_t t1) {
        for (size_t t = t0; t < t1; ++t) {
            V3 a = vtx(s, s.tri_v[3 * t]), b = vtx(s, s.tri_v[3 * t + 1]), c = vtx(s, s.tri_v[3 * t + 2]);
            const float pa[3] = {a.x, a.y, a.z}, pb[3] = {b.x, b.y, b.z}, pc[3] = {c.x, c.y, c.z};
            for (int k = 0; k < 3; ++k) {
                bmin[3 * t + k] = std::min(std::min(pa[k], pb[k]), pc[k]);
                bmax[3 * t + k] = std::max(std::max(pa[k], pb[k]), pc[k]);
                cen[3 * t + k] = ((pa[k] + pb[k]) + pc[k]) / 3.0f;
            }
        }
    });
    out.order.resize(n);
    for (size_t t = 0; t < n; ++t) out.order[t] = int32_t(t);  // objects are created shape-major, accel.h:116-120
    std::vector<int32_t>& ord = out.order;

    struct Todo { uint32_t start, end; int32_t parent; bool is_right; };
    // One step of the reference's loop (bvh.h:159-246): bound the range, emit the node, split it.  `nodes` is the array the
    // node goes into (the whole tree, or a subtree built on its own); children are pushed right first so that the left one
    // is built next: pre-order, as the reference's flat array.
    auto step = [&](const Todo& td, std::vector<RefTree::Node>& nodes, std::vector<Todo>& stack) {
        const uint32_t start = td.start, end = td.end, cnt = end - start;
        RefTree::Node node;
        node.start = start; node.count = cnt; node.left = -1; node.right = -1;
        float cmin[3], cmax[3];
        for (int k = 0; k < 3; ++k) {
            node.bmin[k] = bmin[3 * ord[start] + k]; node.bmax[k] = bmax[3 * ord[start] + k];
            cmin[k] = cmax[k] = cen[3 * ord[start] + k];
        }
        for (uint32_t p = start + 1; p < end; ++p) {
            const int32_t t = ord[p];
            for (int k = 0; k < 3; ++k) {
                node.bmin[k] = std::min(node.bmin[k], bmin[3 * t + k]); node.bmax[k] = std::max(node.bmax[k], bmax[3 * t + k]);
                cmin[k] = std::min(cmin[k], cen[3 * t + k]); cmax[k] = std::max(cmax[k], cen[3 * t + k]);
            }
        }
        const int32_t me = int32_t(nodes.size());
        nodes.push_back(node);
        if (td.parent >= 0) {
            if (td.is_right) nodes[td.parent].right = me; else nodes[td.parent].left = me;
        }
        if (cnt <= 4) return;  // leafSize = 4 (bvh.h:122)
        // split axis: BBox::maxDimension with its y/z comparison quirk (bvh.h:82-87)
        const float ex = cmax[0] - cmin[0], ey = cmax[1] - cmin[1], ez = cmax[2] - cmin[2];
        int dim = 0;
        if (ey > ex) dim = 1;
        if (ez > ey) dim = 2;
        const float split = .5f * (cmin[dim] + cmax[dim]);
        // in-place partition; the swap sequence defines the final primitive order (bvh.h:207-213)
        uint32_t mid = start;
        for (uint32_t i = start; i < end; ++i) {
            if (cen[3 * ord[i] + dim] < split) { std::swap(ord[i], ord[mid]); ++mid; }
        }
        if (mid == start || mid == end) mid = start + (end - start) / 2;
        stack.push_back({mid, end, me, true});
        stack.push_back({start, mid, me, false});
    };

    const unsigned T = build_threads();
    const uint32_t defer_below = (T > 1 && n >= (size_t(1) << 17)) ? uint32_t(n / (size_t(T) * 8)) : 0u;  // 0: plain sequential build
    // top of the tree, sequentially and in pre-order; a range smaller than `defer_below` is set aside as a subtree task and
    // leaves a placeholder in the pre-order sequence
    struct Item { int32_t top_index; int32_t task; };  // one of the two is >= 0
    struct Task { Todo td; std::vector<RefTree::Node> nodes; };
    std::vector<RefTree::Node> top;
    std::vector<Item> seq;
    std::vector<Task> tasks;
    std::vector<std::pair<int32_t, bool>> task_parent;  // (top parent, is_right) per task
    {
        std::vector<Todo> stack;
        stack.push_back({0u, uint32_t(n), -1, false});
        while (!stack.empty()) {
            Todo td = stack.back();
            stack.pop_back();
            if (defer_below && td.end - td.start < defer_below && td.parent >= 0) {
                seq.push_back({-1, int32_t(tasks.size())});
                task_parent.push_back({td.parent, td.is_right});
                Task t; t.td = {td.start, td.end, -1, false};
                tasks.push_back(std::move(t));
                continue;
            }
            seq.push_back({int32_t(top.size()), -1});
            step(td, top, stack);
        }
    }
    // the set-aside subtrees, in parallel (disjoint primitive ranges), each into its own node array with local indices
    if (!tasks.empty()) {
        std::atomic<size_t> next{0};
        auto worker = [&]() {
            std::vector<Todo> stack;
            for (size_t i = next.fetch_add(1); i < tasks.size(); i = next.fetch_add(1)) {
                Task& t = tasks[i];
                t.nodes.reserve(2 * size_t(t.td.end - t.td.start) / 3 + 4);
                stack.clear();
                stack.push_back(t.td);
                while (!stack.empty()) { Todo td = stack.back(); stack.pop_back(); step(td, t.nodes, stack); }
            }
        };
        std::vector<std::thread> th;
        for (unsigned k = 1; k < T; ++k) th.emplace_back(worker);
        worker();
        for (auto& t : th) t.join();
    }
    // stitch: final index = position in the pre-order sequence, a subtree counting for all of its nodes
    std::vector<int32_t> top_final(top.size(), 0), task_base(tasks.size(), 0);
    size_t total = 0;
    for (const Item& it : seq) {
        if (it.top_index >= 0) { top_final[it.top_index] = int32_t(total); total += 1; }
        else { task_base[it.task] = int32_t(total); total += tasks[it.task].nodes.size(); }
    }
    out.nodes.assign(total, RefTree::Node{});
    for (size_t i = 0; i < top.size(); ++i) {
        RefTree::Node nd = top[i];
        if (nd.left >= 0) nd.left = top_final[nd.left];
        if (nd.right >= 0) nd.right = top_final[nd.right];
        out.nodes[top_final[i]] = nd;
    }
    for (size_t k = 0; k < tasks.size(); ++k) {
        const int32_t base = task_base[k];
        const std::vector<RefTree::Node>& sub = tasks[k].nodes;
        for (size_t i = 0; i < sub.size(); ++i) {
            RefTree::Node nd = sub[i];
            if (nd.left >= 0) { nd.left += base; nd.right += base; }
            out.nodes[size_t(base) + i] = nd;
        }
        RefTree::Node& par = out.nodes[top_final[task_parent[k].first]];
        if (task_parent[k].second) par.right = base; else par.left = base;
    }
    return true;
}

// The 4-wide upper tree (tr_host.h: DevNode4), level by level.  A wide node starts from a binary inner node's two children and
// opens, while it has fewer than four, the inner child of largest surface area (ties: the first); an opened child is replaced
// in place by its two children, so the children keep the reference's left-to-right order.  Inner children are numbered in
// (parent order, slot order) on the next level - the same rule the device build follows with a prefix sum.
void build_wide_nodes(const RefTree& tree, std::vector<DevNode4>& out, int32_t& depth) {
    out.clear(); depth = 0;
    if (tree.nodes.empty() || tree.nodes[0].left < 0) return;   // a single leaf: the root reference is the leaf itself
    std::vector<int32_t> level{0}, next;
    out.reserve(tree.nodes.size() / 3 + 16);   // (a wide node stands for about two inner binary nodes: no re-allocation on the way in the common case)
    while (!level.empty()) {
        const size_t base = out.size();
        out.resize(base + level.size());
        // children of every node of the level (disjoint outputs), then the numbering of the inner ones
        std::vector<int32_t> slots(4 * level.size(), -1);
        std::vector<uint8_t> n_inner_of(level.size(), 0);
        parallel_for(level.size(), 1 << 14, [&](size_t i0, size_t i1) {
            for (size_t i = i0; i < i1; ++i) {
                int32_t* sl = &slots[4 * i];
                const RefTree::Node& b = tree.nodes[level[i]];
                sl[0] = b.left; sl[1] = b.right;
                int n = 2;
                while (n < 4) {
                    int best = -1; float best_a = 0.f;
                    for (int k = 0; k < n; ++k) {
                        const RefTree::Node& c = tree.nodes[sl[k]];
                        if (c.left < 0) continue;
                        const float a = wide_box_area(c.bmin, c.bmax);
                        if (best < 0 || a > best_a) { best = k; best_a = a; }
                    }
                    if (best < 0) break;
                    const RefTree::Node& c = tree.nodes[sl[best]];
                    for (int k = n; k > best + 1; --k) sl[k] = sl[k - 1];
                    sl[best] = c.left; sl[best + 1] = c.right;
                    ++n;
                }
                uint8_t ci = 0;
                for (int k = 0; k < n; ++k) if (tree.nodes[sl[k]].left >= 0) ++ci;
                n_inner_of[i] = ci;
            }
        });
        next.clear();
        std::vector<uint32_t> first_inner(level.size() + 1, 0);
        for (size_t i = 0; i < level.size(); ++i) first_inner[i + 1] = first_inner[i] + n_inner_of[i];
        next.resize(first_inner[level.size()]);
        const size_t next_base = base + level.size();
        parallel_for(level.size(), 1 << 14, [&](size_t i0, size_t i1) {
            for (size_t i = i0; i < i1; ++i) {
                DevNode4& w = out[base + i];
                uint32_t at = first_inner[i];
                for (int k = 0; k < 4; ++k) {
                    DevNode4::Half& hf = w.half[k >> 1];
                    float* mn = (k & 1) ? hf.c1min : hf.c0min; float* mx = (k & 1) ? hf.c1max : hf.c0max;
                    int32_t& ref = (k & 1) ? hf.ref1 : hf.ref0;
                    hf.pad0 = hf.pad1 = 0;
                    const int32_t ci = slots[4 * i + k];
                    if (ci < 0) { for (int a = 0; a < 3; ++a) mn[a] = mx[a] = 0.f; ref = TRB_WIDE_EMPTY; continue; }
                    const RefTree::Node& c = tree.nodes[ci];
                    for (int a = 0; a < 3; ++a) { mn[a] = c.bmin[a]; mx[a] = c.bmax[a]; }
                    if (c.left >= 0) { next[at] = ci; ref = int32_t(next_base + at); ++at; }
                    else ref = ~int32_t((c.start << 3) | c.count);
                }
            }
        });
        level.swap(next);
        ++depth;
    }
}

// ---- the SAH upper tree over the reference's leaves (tr_host.h: build_sah_wide) ------------------------------------------------------
namespace {
struct SahBin { float mn[3], mx[3]; uint32_t n; };
inline void bin_clear(SahBin& b) { b.n = 0; for (int k = 0; k < 3; ++k) { b.mn[k] = 3.4e38f; b.mx[k] = -3.4e38f; } }
inline void bin_add(SahBin& b, const LeafBox& l) { ++b.n; for (int k = 0; k < 3; ++k) { b.mn[k] = std::min(b.mn[k], l.bmin[k]); b.mx[k] = std::max(b.mx[k], l.bmax[k]); } }
inline void bin_merge(SahBin& b, const SahBin& o) { b.n += o.n; for (int k = 0; k < 3; ++k) { b.mn[k] = std::min(b.mn[k], o.mn[k]); b.mx[k] = std::max(b.mx[k], o.mx[k]); } }
constexpr int SAH_BINS = 16;
// nb = bins of this node: min(16, leaves in the range) - most nodes of the tree hold 2-4 leaves, and clearing and sweeping 48 bins for
// each of them was most of the build time
inline int sah_bin_of(const LeafBox& l, int ax, float cmn, float ext, int nb) {
    const float c = 0.5f * (l.bmin[ax] + l.bmax[ax]);
    return std::min(nb - 1, int((c - cmn) / ext * float(nb)));
}
}  // namespace

// a fixed set of threads that runs f(0) .. f(T-1) once per run(): the top of the SAH tree makes hundreds of short passes over one range
namespace {
struct BuildPool {
    unsigned T;
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv, done_cv;
    std::function<void(unsigned)> job;
    uint64_t gen = 0;
    unsigned pending = 0;
    bool stop = false;
    explicit BuildPool(unsigned t) : T(t) { for (unsigned k = 1; k < T; ++k) th.emplace_back([this, k]() { loop(k); }); }
    void loop(unsigned k) {
        uint64_t seen = 0;
        for (;;) {
            std::function<void(unsigned)> j;
            { std::unique_lock<std::mutex> l(m); cv.wait(l, [&]() { return stop || gen != seen; }); if (stop) return; seen = gen; j = job; }
            j(k);
            { std::lock_guard<std::mutex> l(m); if (--pending == 0) done_cv.notify_one(); }
        }
    }
    void run(const std::function<void(unsigned)>& f) {
        if (T == 1) { f(0); return; }
        { std::lock_guard<std::mutex> l(m); job = f; pending = T - 1; ++gen; }
        cv.notify_all();
        f(0);
        std::unique_lock<std::mutex> l(m);
        done_cv.wait(l, [&]() { return pending == 0; });
    }
    ~BuildPool() { { std::lock_guard<std::mutex> l(m); stop = true; } cv.notify_all(); for (auto& x : th) x.join(); }
};
}  // namespace

void build_sah_wide(const LeafBox* L_in, size_t n, std::vector<DevNode4>& out, int32_t& depth) {
    out.clear(); depth = 0;
    const auto t_begin = std::chrono::steady_clock::now();
    if (n < 2) return;   // a single leaf: the root reference is the leaf itself
    // The leaf records themselves are partitioned (32 B moves), not an index: every range is contiguous memory, and a subtree task
    // works on a private copy that fits the caches.  The top of the tree partitions a range from one array into the other (`in_scr`
    // says where a range lives), all threads on one range; two passes per node: bins, then scatter + the children's bounds.
    std::vector<LeafBox> all(L_in, L_in + n);
    std::vector<LeafBox> scratch;
    const unsigned T = build_threads();
    BuildPool pool(T);
    struct Bounds { float bmin[3], bmax[3], cmn[3], cmx[3]; };
    struct Todo { uint32_t lo, hi; int32_t parent; bool is_right; bool in_scr; bool has_b; Bounds b; };
    auto bounds_clear = [](Bounds& b) { for (int k = 0; k < 3; ++k) { b.bmin[k] = 3.4e38f; b.bmax[k] = -3.4e38f; b.cmn[k] = 3.4e38f; b.cmx[k] = -3.4e38f; } };
    auto bounds_add = [](Bounds& b, const LeafBox& l) {
        for (int k = 0; k < 3; ++k) {
            b.bmin[k] = std::min(b.bmin[k], l.bmin[k]); b.bmax[k] = std::max(b.bmax[k], l.bmax[k]);
            const float c = 0.5f * (l.bmin[k] + l.bmax[k]);
            b.cmn[k] = std::min(b.cmn[k], c); b.cmx[k] = std::max(b.cmx[k], c);
        }
    };
    auto bounds_merge = [](Bounds& b, const Bounds& o) {
        for (int k = 0; k < 3; ++k) {
            b.bmin[k] = std::min(b.bmin[k], o.bmin[k]); b.bmax[k] = std::max(b.bmax[k], o.bmax[k]);
            b.cmn[k] = std::min(b.cmn[k], o.cmn[k]); b.cmx[k] = std::max(b.cmx[k], o.cmx[k]);
        }
    };
    auto bin_slice = [](const LeafBox* L, uint32_t lo, uint32_t hi, const float cmn[3], const float ext[3], int nbins, SahBin (*bins)[SAH_BINS]) {
        for (int ax = 0; ax < 3; ++ax) for (int b = 0; b < nbins; ++b) bin_clear(bins[ax][b]);
        for (uint32_t i = lo; i < hi; ++i)
            for (int ax = 0; ax < 3; ++ax) if (ext[ax] > 0.f) bin_add(bins[ax][sah_bin_of(L[i], ax, cmn[ax], ext[ax], nbins)], L[i]);
    };
    // best split of a binned range (ties: first axis, first bin); false when every centroid coincides
    auto choose_split = [](SahBin (*bins)[SAH_BINS], int nbins, const float ext[3], int* baxis, int* bsplit) {
        float best = 3.4e38f; *baxis = -1; *bsplit = -1;
        for (int ax = 0; ax < 3; ++ax) {
            if (!(ext[ax] > 0.f)) continue;
            float ra[SAH_BINS]; uint32_t rn[SAH_BINS];
            SahBin acc; bin_clear(acc);
            for (int b = nbins - 1; b >= 0; --b) { bin_merge(acc, bins[ax][b]); rn[b] = acc.n; ra[b] = acc.n ? wide_box_area(acc.mn, acc.mx) : 0.f; }
            bin_clear(acc);
            for (int b = 0; b + 1 < nbins; ++b) {
                bin_merge(acc, bins[ax][b]);
                if (!acc.n || !rn[b + 1]) continue;
                const float cost = wide_box_area(acc.mn, acc.mx) * float(acc.n) + ra[b + 1] * float(rn[b + 1]);
                if (cost < best) { best = cost; *baxis = ax; *bsplit = b; }
            }
        }
        return *baxis >= 0;
    };
    auto new_node = [](std::vector<RefTree::Node>& nodes, const Todo& td, const Bounds& nb, const LeafBox& first) {
        RefTree::Node node;
        node.left = -1; node.right = -1; node.start = first.start; node.count = first.count;
        for (int k = 0; k < 3; ++k) { node.bmin[k] = nb.bmin[k]; node.bmax[k] = nb.bmax[k]; }
        const int32_t me = int32_t(nodes.size());
        nodes.push_back(node);
        if (td.parent >= 0) { if (td.is_right) nodes[td.parent].right = me; else nodes[td.parent].left = me; }
        return me;
    };
    // one node over L[lo, hi) on the calling thread: bounds, best split, stable partition in place
    auto step = [&](LeafBox* L, const Todo& td, std::vector<RefTree::Node>& nodes, std::vector<Todo>& stack) {
        const uint32_t lo = td.lo, hi = td.hi, cnt = hi - lo;
        Bounds nb;
        if (td.has_b) nb = td.b; else { bounds_clear(nb); for (uint32_t i = lo; i < hi; ++i) bounds_add(nb, L[i]); }
        const int32_t me = new_node(nodes, td, nb, L[lo]);
        if (cnt == 1) return;
        const float* cmn = nb.cmn;
        const float ext[3] = {nb.cmx[0] - nb.cmn[0], nb.cmx[1] - nb.cmn[1], nb.cmx[2] - nb.cmn[2]};
        SahBin bins[3][SAH_BINS];
        const int nbins = int(std::min<uint32_t>(SAH_BINS, cnt));
        bin_slice(L, lo, hi, cmn, ext, nbins, bins);
        int baxis, bsplit;
        uint32_t mid = lo + cnt / 2;   // coincident centroids: split the list in the middle, order untouched
        if (choose_split(bins, nbins, ext, &baxis, &bsplit)) {
            const float c0 = cmn[baxis], e0 = ext[baxis];
            auto goes_left = [&](const LeafBox& l) { return sah_bin_of(l, baxis, c0, e0, nbins) <= bsplit; };
            uint32_t m;
            if (cnt <= 64) {   // small range: stable partition through a stack buffer (std::stable_partition allocates one per call)
                LeafBox right[64]; uint32_t nr = 0, w = lo;
                for (uint32_t i = lo; i < hi; ++i) { if (goes_left(L[i])) L[w++] = L[i]; else right[nr++] = L[i]; }
                for (uint32_t i = 0; i < nr; ++i) L[w + i] = right[i];
                m = w;
            } else {
                m = uint32_t(std::stable_partition(L + lo, L + hi, goes_left) - L);
            }
            if (m > lo && m < hi) mid = m;
        }
        Todo r{mid, hi, me, true, td.in_scr, false, Bounds()}, l{lo, mid, me, false, td.in_scr, false, Bounds()};
        stack.push_back(r);
        stack.push_back(l);
    };
    // the same node with all threads on the range: per-slice bins (their counts give every slice's share of the left side, so the
    // partition needs no counting pass), then one pass that scatters the range into the OTHER array, stable, and gathers the bounds of
    // both children on the way.  min / max / count reductions and slice-ordered offsets: the result does not depend on T.
    auto step_wide = [&](const Todo& td, std::vector<RefTree::Node>& nodes, std::vector<Todo>& stack) {
        const uint32_t lo = td.lo, hi = td.hi, cnt = hi - lo;
        const LeafBox* src = td.in_scr ? scratch.data() : all.data();
        LeafBox* dst = td.in_scr ? all.data() : scratch.data();
        auto slice = [&](unsigned t, uint32_t* a, uint32_t* b) { *a = lo + uint32_t(uint64_t(cnt) * t / T); *b = lo + uint32_t(uint64_t(cnt) * (t + 1) / T); };
        Bounds nb;
        if (td.has_b) nb = td.b;
        else {
            std::vector<Bounds> part(T);
            pool.run([&](unsigned t) { uint32_t a, b; slice(t, &a, &b); bounds_clear(part[t]); for (uint32_t i = a; i < b; ++i) bounds_add(part[t], src[i]); });
            bounds_clear(nb);
            for (unsigned t = 0; t < T; ++t) bounds_merge(nb, part[t]);
        }
        const int32_t me = new_node(nodes, td, nb, src[lo]);
        const float* cmn = nb.cmn;
        const float ext[3] = {nb.cmx[0] - nb.cmn[0], nb.cmx[1] - nb.cmn[1], nb.cmx[2] - nb.cmn[2]};
        const int nbins = SAH_BINS;
        std::vector<std::array<std::array<SahBin, SAH_BINS>, 3>> part(T);
        pool.run([&](unsigned t) { uint32_t a, b; slice(t, &a, &b); bin_slice(src, a, b, cmn, ext, nbins, reinterpret_cast<SahBin(*)[SAH_BINS]>(part[t].data())); });
        SahBin bins[3][SAH_BINS];
        for (int ax = 0; ax < 3; ++ax) for (int b = 0; b < nbins; ++b) { bin_clear(bins[ax][b]); for (unsigned t = 0; t < T; ++t) bin_merge(bins[ax][b], part[t][ax][b]); }
        int baxis, bsplit;
        if (!choose_split(bins, nbins, ext, &baxis, &bsplit)) {   // coincident centroids: the middle of the list, nothing moves
            const uint32_t mid = lo + cnt / 2;
            Todo r{mid, hi, me, true, td.in_scr, false, Bounds()}, l{lo, mid, me, false, td.in_scr, false, Bounds()};
            stack.push_back(r); stack.push_back(l);
            return;
        }
        std::vector<uint32_t> nl(T + 1, 0);
        for (unsigned t = 0; t < T; ++t) { uint32_t c = 0; for (int b = 0; b <= bsplit; ++b) c += part[t][baxis][b].n; nl[t + 1] = nl[t] + c; }
        const uint32_t n_left = nl[T];   // (0 < n_left < cnt: choose_split only takes splits with leaves on both sides)
        const float c0 = cmn[baxis], e0 = ext[baxis];
        std::vector<Bounds> bl(T), br(T);
        pool.run([&](unsigned t) {
            uint32_t a, b; slice(t, &a, &b);
            uint32_t wl = lo + nl[t], wr = lo + n_left + ((a - lo) - nl[t]);
            bounds_clear(bl[t]); bounds_clear(br[t]);
            for (uint32_t i = a; i < b; ++i) {
                if (sah_bin_of(src[i], baxis, c0, e0, nbins) <= bsplit) { dst[wl++] = src[i]; bounds_add(bl[t], src[i]); }
                else { dst[wr++] = src[i]; bounds_add(br[t], src[i]); }
            }
        });
        Todo r{lo + n_left, hi, me, true, !td.in_scr, true, Bounds()}, l{lo, lo + n_left, me, false, !td.in_scr, true, Bounds()};
        bounds_clear(r.b); bounds_clear(l.b);
        for (unsigned t = 0; t < T; ++t) { bounds_merge(l.b, bl[t]); bounds_merge(r.b, br[t]); }
        stack.push_back(r); stack.push_back(l);
    };
    // top of the tree with all threads on one range, ranges below `defer_below` leaves as subtree tasks
    const uint32_t defer_below = (T > 1 && n >= (size_t(1) << 15)) ? uint32_t(n / (size_t(T) * 8)) : 0u;
    struct Task { Todo td; int32_t parent; bool is_right; std::vector<RefTree::Node> nodes; };
    RefTree sah;
    std::vector<Task> tasks;
    {
        if (defer_below) scratch.resize(n);
        std::vector<Todo> stack;
        stack.push_back(Todo{0u, uint32_t(n), -1, false, false, false, Bounds()});
        while (!stack.empty()) {
            Todo td = stack.back(); stack.pop_back();
            if (defer_below && td.hi - td.lo < defer_below && td.parent >= 0) {
                Task t; t.td = td; t.parent = td.parent; t.is_right = td.is_right;
                tasks.push_back(std::move(t));
                continue;
            }
            if (defer_below && td.hi - td.lo >= 4096u) step_wide(td, sah.nodes, stack);
            else step(td.in_scr ? scratch.data() : all.data(), td, sah.nodes, stack);
        }
    }
    const auto t_top = std::chrono::steady_clock::now();
    auto t_tasks = t_top;
    if (!tasks.empty()) {   // disjoint ranges: every task copies its leaves and builds its subtree from the copy
        std::atomic<size_t> next{0};
        pool.run([&](unsigned) {
            std::vector<Todo> stack;
            std::vector<LeafBox> local;
            for (size_t i = next.fetch_add(1); i < tasks.size(); i = next.fetch_add(1)) {
                Task& t = tasks[i];
                const LeafBox* from = t.td.in_scr ? scratch.data() : all.data();
                local.assign(from + t.td.lo, from + t.td.hi);
                t.nodes.reserve(2 * local.size());
                stack.clear();
                Todo root{0u, uint32_t(local.size()), -1, false, false, t.td.has_b, t.td.b};
                stack.push_back(root);
                while (!stack.empty()) { Todo td = stack.back(); stack.pop_back(); step(local.data(), td, t.nodes, stack); }
            }
        });
        t_tasks = std::chrono::steady_clock::now();
        // stitch in task order (the order the top pass met them: deterministic): bases by a prefix sum, copies on all threads
        std::vector<size_t> base(tasks.size() + 1, sah.nodes.size());
        for (size_t k = 0; k < tasks.size(); ++k) base[k + 1] = base[k] + tasks[k].nodes.size();
        sah.nodes.resize(base[tasks.size()]);
        std::atomic<size_t> nxt{0};
        pool.run([&](unsigned) {
            for (size_t k = nxt.fetch_add(1); k < tasks.size(); k = nxt.fetch_add(1)) {
                const int32_t b = int32_t(base[k]);
                RefTree::Node* dst = sah.nodes.data() + base[k];
                for (RefTree::Node nd : tasks[k].nodes) { if (nd.left >= 0) { nd.left += b; nd.right += b; } *dst++ = nd; }
                std::vector<RefTree::Node>().swap(tasks[k].nodes);
            }
        });
        for (size_t k = 0; k < tasks.size(); ++k) {
            RefTree::Node& par = sah.nodes[tasks[k].parent];
            if (tasks[k].is_right) par.right = int32_t(base[k]); else par.left = int32_t(base[k]);
        }
    }
    const auto t_tree = std::chrono::steady_clock::now();
    build_wide_nodes(sah, out, depth);
    if (std::getenv("TRB_BUILD_TIMING"))
        std::fprintf(stderr, "[trb build] SAH upper tree: %zu leaves, %zu binary nodes in %.1f ms (top %.1f, %zu subtree tasks on %u threads %.1f, stitch %.1f), collapse to %zu wide nodes in %.1f ms\n", n,
                     sah.nodes.size(), std::chrono::duration<double, std::milli>(t_tree - t_begin).count(), std::chrono::duration<double, std::milli>(t_top - t_begin).count(), tasks.size(), T,
                     std::chrono::duration<double, std::milli>(t_tasks - t_top).count(), std::chrono::duration<double, std::milli>(t_tree - t_tasks).count(), out.size(),
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_tree).count());
}

bool flatten_scene(const trb_scene_desc& s, FlatScene& out, std::string& err) {
    return flatten_tables(s, out, err) && flatten_geometry(s, out, err);
}

// Everything but the acceleration structure and the per-triangle records: validation, BSDF parameters, textures, emitters.
bool flatten_tables(const trb_scene_desc& s, FlatScene& out, std::string& err) {
    if (!s.positions || !s.normals || !s.tri_v || !s.tri_n || !s.tri_mat || !s.tri_shape || !s.materials) {
        err = "null array in scene description"; return false;
    }
    std::atomic<int> bad{0};  // 1 vertex index, 2 normal index, 3 shape order
    parallel_for(s.n_tris, 1 << 16, [&](size_t t0, size_t t1) {
        for (size_t t = t0; t < t1; ++t) {
            for (int k = 0; k < 3; ++k) {
                if (s.tri_v[3 * t + k] < 0 || size_t(s.tri_v[3 * t + k]) >= s.n_positions) bad.store(1);
                else if (s.tri_n[3 * t + k] < 0 || size_t(s.tri_n[3 * t + k]) >= s.n_normals) { int z = 0; bad.compare_exchange_strong(z, 2); }
            }
            if (s.tri_shape[t] < 0 || size_t(s.tri_shape[t]) >= s.n_shapes || (t > 0 && s.tri_shape[t] < s.tri_shape[t - 1])) { int z = 0; bad.compare_exchange_strong(z, 3); }
        }
    });
    if (bad.load() == 1) { err = "vertex index out of range"; return false; }
    if (bad.load() == 2) { err = "normal index out of range (the reference needs vn on every face)"; return false; }
    if (bad.load() == 3) { err = "tri_shape must be non-decreasing and < n_shapes"; return false; }
    if (!derive_materials(s, out.materials, err)) return false;
    // bitmap textures: table, texels, per-triangle coordinates (filled with the triangles below)
    out.tex_info.clear(); out.texels.clear(); out.tri_uv.clear();
    bool textured = false;
    for (const DevMaterial& m : out.materials) textured = textured || m.tex != 0;
    if (textured) {
        for (size_t t = 0; t < s.n_textures; ++t) {
            const trb_texture& x = s.textures[t];
            if (!x.rgb || x.width <= 0 || x.height <= 0) { err = "empty texture"; return false; }
            if (out.texels.size() / 3 + size_t(x.width) * x.height > 0xffffffffu) { err = "textures too large"; return false; }
            out.tex_info.push_back(DevTexInfo{uint32_t(out.texels.size() / 3), x.width, x.height, 0});
            out.texels.insert(out.texels.end(), x.rgb, x.rgb + size_t(x.width) * x.height * 3);
        }
        if (!s.tri_t || !s.texcoords) { err = "textured material without texture coordinates"; return false; }
        for (size_t t = 0; t < s.n_tris; ++t) {
            if (!out.materials[s.tri_mat[t]].tex) continue;
            for (int k = 0; k < 3; ++k)  // the reference indexes attrib.texcoords with -1 here (core.h:572-577): undefined
                if (s.tri_t[3 * t + k] < 0 || size_t(s.tri_t[3 * t + k]) >= s.n_texcoords) { err = "face of a textured material without vt: undefined in the reference"; return false; }
        }
    }
    std::vector<int32_t>& shape_to_emitter = out.shape_to_emitter;
    if (!derive_emitters(s, out.materials, out.emitters, out.emitter_tris, out.emitter_cdf, shape_to_emitter, err)) return false;
    out.shape_first.assign(s.n_shapes + 1, 0);
    for (size_t t = 0; t < s.n_tris; ++t) out.shape_first[s.tri_shape[t] + 1] = int32_t(t + 1);
    for (size_t k = 1; k <= s.n_shapes; ++k) out.shape_first[k] = std::max(out.shape_first[k], out.shape_first[k - 1]);
    out.textured = textured;
    out.nodes.clear(); out.wnodes.clear(); out.tris.clear(); out.shade.clear(); out.order.clear();
    out.n_nodes = out.n_tris = out.n_tri_uv = out.n_wnodes = 0; out.wide_depth = 0;
    {
        std::vector<float> sph;
        shape_spheres(s, sph);
        out.emitter_sphere.clear();
        for (const DevEmitter& e : out.emitters) out.emitter_sphere.insert(out.emitter_sphere.end(), sph.begin() + 4 * e.shape, sph.begin() + 4 * e.shape + 4);
    }

    return true;
}

// The reference-order BVH on the host cores and the per-triangle records in its primitive order (the device build of
// kernels/bvh_build.cuh produces the same arrays on the GPU).
bool flatten_geometry(const trb_scene_desc& s, FlatScene& out, std::string& err) {
    const bool textured = out.textured;
    const std::vector<int32_t>& shape_to_emitter = out.shape_to_emitter;
    const std::vector<int32_t>& shape_first = out.shape_first;
    RefTree tree;
    if (!build_ref_tree(s, tree, err)) return false;
    out.order = tree.order;
    const size_t n = s.n_tris;

    out.tris.resize(n);
    out.shade.resize(n);
    if (textured) out.tri_uv.assign(n, DevTriUV{{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}});
    parallel_for(n, 1 << 16, [&](size_t p0, size_t p1) {
    for (size_t p = p0; p < p1; ++p) {
        const int32_t t = tree.order[p];
        V3 a = vtx(s, s.tri_v[3 * t]), b = vtx(s, s.tri_v[3 * t + 1]), c = vtx(s, s.tri_v[3 * t + 2]);
        V3 n0 = nrm(s, s.tri_n[3 * t]), n1 = nrm(s, s.tri_n[3 * t + 1]), n2 = nrm(s, s.tri_n[3 * t + 2]);
        const int32_t mat = s.tri_mat[t];
        DevTri& dt = out.tris[p];
        dt.v0[0] = a.x; dt.v0[1] = a.y; dt.v0[2] = a.z;
        dt.e1[0] = b.x - a.x; dt.e1[1] = b.y - a.y; dt.e1[2] = b.z - a.z;
        dt.e2[0] = c.x - a.x; dt.e2[1] = c.y - a.y; dt.e2[2] = c.z - a.z;
        dt.tri = t;
        {   // same operation order as the device evaluated it: ((1e-6f * e1.e1) * e2.e2), each dot as (x*x + y*y) + z*z
            const float e1 = (dt.e1[0] * dt.e1[0] + dt.e1[1] * dt.e1[1]) + dt.e1[2] * dt.e1[2];
            const float e2 = (dt.e2[0] * dt.e2[0] + dt.e2[1] * dt.e2[1]) + dt.e2[2] * dt.e2[2];
            dt.illk = 1e-6f * e1 * e2;
        }
        dt.flags = out.materials[mat].cls | (out.materials[mat].emissive ? 8 : 0);
        DevShade& ds = out.shade[p];
        ds.v0[0] = a.x; ds.v0[1] = a.y; ds.v0[2] = a.z; ds.v1[0] = b.x; ds.v1[1] = b.y; ds.v1[2] = b.z;
        ds.v2[0] = c.x; ds.v2[1] = c.y; ds.v2[2] = c.z;
        ds.n0[0] = n0.x; ds.n0[1] = n0.y; ds.n0[2] = n0.z; ds.n1[0] = n1.x; ds.n1[1] = n1.y; ds.n1[2] = n1.z;
        ds.n2[0] = n2.x; ds.n2[1] = n2.y; ds.n2[2] = n2.z;
        ds.shape = s.tri_shape[t]; ds.prim = t - shape_first[s.tri_shape[t]]; ds.mat = mat;
        ds.emitter = shape_to_emitter[s.tri_shape[t]]; ds.tri = t; ds.pad = 0;
        if (textured && out.materials[mat].tex) {
            DevTriUV& uv = out.tri_uv[p];
            for (int k = 0; k < 2; ++k) {
                uv.st0[k] = s.texcoords[2 * s.tri_t[3 * t] + k]; uv.st1[k] = s.texcoords[2 * s.tri_t[3 * t + 1] + k];
                uv.st2[k] = s.texcoords[2 * s.tri_t[3 * t + 2] + k];
            }
        }
    }
    });

    // inner nodes only, in the reference's pre-order (left child adjacent); leaves become negative child references
    std::vector<int32_t> inner_index(tree.nodes.size(), -1);
    int32_t n_inner = 0;
    for (size_t i = 0; i < tree.nodes.size(); ++i) if (tree.nodes[i].left >= 0) inner_index[i] = n_inner++;
    auto make_ref = [&](int32_t ni) -> int32_t {
        const RefTree::Node& nd = tree.nodes[ni];
        if (nd.left >= 0) return inner_index[ni];
        return ~int32_t((nd.start << 3) | nd.count);
    };
    out.nodes.resize(size_t(n_inner));
    std::atomic<bool> apart{false};
    parallel_for(tree.nodes.size(), 1 << 15, [&](size_t i0, size_t i1) {
        for (size_t i = i0; i < i1; ++i) {
            const RefTree::Node& nd = tree.nodes[i];
            if (nd.left < 0) continue;
            DevNode& d = out.nodes[inner_index[i]];
            const RefTree::Node& l = tree.nodes[nd.left];
            const RefTree::Node& r = tree.nodes[nd.right];
            for (int k = 0; k < 3; ++k) { d.c0min[k] = l.bmin[k]; d.c0max[k] = l.bmax[k]; d.c1min[k] = r.bmin[k]; d.c1max[k] = r.bmax[k]; }
            d.ref0 = make_ref(nd.left); d.ref1 = make_ref(nd.right); d.pad0 = d.pad1 = 0;
            // the traversal kernel postpones sibling leaves as ONE contiguous triangle range (trace_coop.cuh)
            if (l.left < 0 && r.left < 0 && r.start != l.start + l.count) apart.store(true);
        }
    });
    if (apart.load()) { err = "sibling leaves are not adjacent in the primitive order"; return false; }
    out.root_ref = make_ref(0);
    for (int k = 0; k < 3; ++k) { out.scene_min[k] = tree.nodes[0].bmin[k]; out.scene_max[k] = tree.nodes[0].bmax[k]; }
    // depth (pre-order array: children always follow their parent)
    std::vector<int32_t> depth(tree.nodes.size(), 0);
    int32_t md = 0;
    for (size_t i = 0; i < tree.nodes.size(); ++i) {
        if (tree.nodes[i].left >= 0) { depth[tree.nodes[i].left] = depth[i] + 1; depth[tree.nodes[i].right] = depth[i] + 1; }
        md = std::max(md, depth[i]);
    }
    out.max_depth = md;
    {   // the upper tree the traversal walks: SAH over the reference's leaves (primitive order = the reference's pre-order)
        std::vector<LeafBox> leaves;
        for (const RefTree::Node& nd : tree.nodes)
            if (nd.left < 0) { LeafBox l; for (int k = 0; k < 3; ++k) { l.bmin[k] = nd.bmin[k]; l.bmax[k] = nd.bmax[k]; } l.start = nd.start; l.count = nd.count; leaves.push_back(l); }
        if (out.wide_collapse) build_wide_nodes(tree, out.wnodes, out.wide_depth);
        else build_sah_wide(leaves.data(), leaves.size(), out.wnodes, out.wide_depth);
    }
    out.n_nodes = out.nodes.size(); out.n_tris = out.tris.size(); out.n_tri_uv = out.tri_uv.size(); out.n_wnodes = out.wnodes.size();
    return true;
}

}  // namespace trb
