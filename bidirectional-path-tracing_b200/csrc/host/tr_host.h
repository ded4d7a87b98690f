// tr_host.h -- host-side (CPU, no CUDA) half of the product: scene loading, the reference-order BVH
// build, flattening into the device layout, camera matrices, TOML and EXR I/O.
// This is the part of TinyRender's Scene::load / loadTOML / saveEXR that stays on the host
// (reference: src/core/renderer.cpp:235-339, src/main.cpp:22-116, src/core/utils.h:95-156).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../../include/trb.h"

namespace trb {

// ---- what tinyobj::LoadObj(triangulate=true) yields, flattened shape-major -----------------------
struct HostScene {
    std::vector<float> positions, normals;
    std::vector<int32_t> tri_v, tri_n, tri_mat, tri_shape;
    std::vector<trb_material> materials;
    std::vector<std::string> shape_names, material_names;
    size_t n_shapes = 0;
    // bitmap textures (core.h:417-497, 525-585): attrib.texcoords, per-corner vt indices, the decoded images
    std::vector<float> texcoords;
    std::vector<int32_t> tri_t;
    struct Texture { std::string file; int32_t w = 0, h = 0; std::vector<float> rgb; };
    std::vector<Texture> textures;
    mutable std::vector<trb_texture> texture_views;  // what desc() points at
    trb_scene_desc desc() const;
};

// Tex::load (core.h:456-497): `<stem>.ppm` next to the name given, binary P6, gamma-expanded and vertically flipped;
// a file that cannot be read becomes the reference's 1x1 pink texel.
void load_ppm_texture(const std::string& path, HostScene::Texture& out);
// BitmapTexture3f::{getMax,getAverage} (core.h:546-566)
void texture_max_average(const trb_texture& t, float mx[3], float avg[3]);

bool load_obj(const std::string& obj_path, HostScene& out, std::string& err);

// ---- device layouts (shared with the CUDA side; all 16-byte aligned, float4-vectorised) ----------
// 64-byte 2-wide node: the two CHILD boxes live in the parent, so one 64 B fetch (two 32 B sectors,
// 4x LDG.128) decides both children.  Same topology as the reference's flat tree
// (externals/bvh.h:102-105, 48 B/node holding its OWN box) - see DESIGN.md for why topology is kept.
struct alignas(16) DevNode {
    float c0min[3], c0max[3];  // child 0 (reference: node ni+1)
    float c1min[3], c1max[3];  // child 1 (reference: node ni+rightOffset)
    int32_t ref0, ref1;        // >=0: inner node index; <0: leaf, ~ref = (first << 3) | count
    int32_t pad0, pad1;
};
static_assert(sizeof(DevNode) == 64, "node must be 64 B");

// 128-byte 4-wide node of the UPPER tree the traversal kernels walk (kernels/trace_wide.cuh).
// The reference's answer is "min (t, list position) over the triangles ALL of whose ancestor boxes pass BBox::intersect"
// (bvh.h:259-352).  Every box of the reference's tree is the exact min/max union of its children's (bvh.h:171-182) and
// fl((b - o) / d) is monotone in b, so for a ray whose direction has three finite non-zero components a LEAF box that
// passes implies that all its ancestors pass: the candidate set is decided by the leaf boxes alone, and any tree over the
// reference's leaves whose inner boxes are unions of leaf boxes finds it.  This one is the reference's own binary tree
// collapsed to four children per node (the child of largest surface area is opened first): half the dependent fetches per
// ray.  Leaves, their exact boxes and the primitive order (= the tie-break rank) are the reference's.  Rays with a zero or
// non-finite direction component (NaN poisoning in the slab test, where the equivalence does not hold) are traced on the
// binary DevNode tree.  Laid out as two 64-byte halves of two children each (the DevNode layout twice), 32-byte aligned
// pieces: 4 x LDG.256 per node (sm_100 has 256-bit global loads), and a half is consumed as soon as its two loads land.
#define TRB_WIDE_EMPTY INT32_MIN   // an unused child slot (never a leaf code: first < 2^27)
struct alignas(32) DevNode4 {
    struct Half {
        float c0min[3], c0max[3];
        float c1min[3], c1max[3];
        int32_t ref0, ref1;   // >= 0: wide node index; < 0: leaf, ~ref = (first << 3) | count; TRB_WIDE_EMPTY: unused slot
        int32_t pad0, pad1;
    } half[2];                // children 0,1 and 2,3 in the reference's left-to-right order
};
static_assert(sizeof(DevNode4) == 128, "wide node must be 128 B");

// 48-byte intersection record, stored in the reference's build-reordered primitive order so that the
// array position IS the tie-break rank (first found wins, bvh.h:305).
struct alignas(16) DevTri {
    float v0[3];
    int32_t tri;    // global shape-major triangle id
    float e1[3];    // v1 - v0 (one fp32 subtraction, identical to core.h:386)
    float illk;     // 1e-6f * |e1|^2 * |e2|^2: the conditioning bound of tri_test (trace.cuh), so that it is not recomputed per test
    float e2[3];    // v2 - v0
    int32_t flags;  // bits 0-2: material class; bit 3: emissive material
};
static_assert(sizeof(DevTri) == 48, "triangle record must be 48 B");

// 96-byte shading record (same order as DevTri): what AcceleratorBVH::intersect gathers after the
// closest hit (accel.h:136-166).
struct alignas(16) DevShade {
    float v0[3]; int32_t shape;
    float v1[3]; int32_t prim;
    float v2[3]; int32_t mat;
    float n0[3]; int32_t emitter;  // emitter index of the owning shape or -1
    float n1[3]; int32_t tri;
    float n2[3]; int32_t pad;
};
static_assert(sizeof(DevShade) == 96, "shading record must be 96 B");

enum MatClass : int32_t { MAT_DIFFUSE = 0, MAT_PHONG = 1, MAT_MIXTURE = 2, MAT_MIRROR = 3, MAT_GLASS = 4, MAT_NCLASS = 5 };

// BSDF parameters derived exactly as the reference's BSDF constructors do.
struct alignas(16) DevMaterial {
    float kd[3]; int32_t cls;
    float ks[3]; uint32_t type;       // BSDF::combinedType
    float ke[3]; float ns;            // emission; Phong exponent
    float tf[3]; float ni;            // glass transmittance; IOR
    float scale;                      // phong.h:41-43
    float spec_weight;                // specularSamplingWeight, phong.h:45-47
    int32_t emissive;                 // any(ke != 0)
    int32_t tex;                      // (1 + map_Kd texture) | (1 + map_Ks texture) << 16 as the BSDF class reads them; 0 = constants
};
static_assert(sizeof(DevMaterial) == 80, "material record must be 80 B");

struct alignas(16) DevEmitter {
    float radiance[3]; float area;
    int32_t shape, first_face, n_faces, cdf_offset;  // first_face indexes emitter_tris; cdf has n_faces+1 entries
};
static_assert(sizeof(DevEmitter) == 32, "emitter record must be 32 B");

// emitter faces in the shape's own face order (what sampleEmitterPosition indexes, integrator.cpp:73-100)
struct alignas(16) DevEmitterTri {
    float v0[3], v1[3], v2[3];
    float n0[3], n1[3], n2[3];
    float pad[2];
};
static_assert(sizeof(DevEmitterTri) == 80, "emitter triangle must be 80 B");

// bitmap textures (BitmapTexture3f, core.h:525-585): table entry + per-triangle texture coordinates (build order, like DevTri)
struct alignas(16) DevTexInfo { uint32_t first_texel; int32_t w, h, pad; };
struct alignas(16) DevTriUV { float st0[2], st1[2], st2[2], pad[2]; };
static_assert(sizeof(DevTexInfo) == 16 && sizeof(DevTriUV) == 32, "texture records");

struct FlatScene {
    std::vector<DevNode> nodes;
    std::vector<DevNode4> wnodes;       // the 4-wide upper tree (level order); empty after a device build
    std::vector<DevTri> tris;
    std::vector<DevShade> shade;
    std::vector<DevMaterial> materials;
    std::vector<DevEmitter> emitters;
    std::vector<DevEmitterTri> emitter_tris;
    std::vector<float> emitter_cdf;
    std::vector<DevTexInfo> tex_info;   // empty for an untextured scene (then tri_uv and texels are empty too)
    std::vector<DevTriUV> tri_uv;
    std::vector<float> texels;          // RGB, all textures back to back
    std::vector<float> emitter_sphere;  // per emitter: Scene::getShapeCenter xyz, getShapeRadius (renderer.cpp:349-358) - DirectIntegrator's sphere lights
    std::vector<int32_t> order;  // order[pos] = global tri id (host build only)
    // counts of the device arrays (nodes / tris = shade / tri_uv): the vectors above are empty after a device build
    size_t n_nodes = 0, n_tris = 0, n_tri_uv = 0, n_wnodes = 0;
    int32_t wide_depth = 0;      // levels of the wide tree
    bool wide_collapse = false;  // TRB_UPLOAD_WIDE_COLLAPSE: the upper tree is the collapse of the reference's own tree (built where the BVH is built) instead of the SAH tree
    bool textured = false;
    std::vector<int32_t> shape_to_emitter, shape_first;  // per shape: emitter index or -1; first global triangle (n_shapes + 1)
    int32_t root_ref = 0;        // root as a child reference (leaf when the scene has <= 4 triangles)
    int32_t max_depth = 0;       // tree depth (traversal stack bound)
    float scene_min[3], scene_max[3];
    size_t bytes() const;
};

// Reference-order BVH build (externals/bvh.h:147-247).  order[pos] = global triangle id.
struct RefTree {
    struct Node { float bmin[3], bmax[3]; uint32_t start, count; int32_t left, right; Node() {} };  // left<0: leaf.  (The empty constructor keeps vector::resize from zero-filling millions of nodes on one thread before the build threads write them.)
    std::vector<Node> nodes;  // pre-order, node 0 = root
    std::vector<int32_t> order;
};
bool build_ref_tree(const trb_scene_desc& s, RefTree& out, std::string& err);
// One leaf of the reference's tree: its exact box and its primitive range.  The list of leaves in primitive order (increasing `start`
// = the reference's pre-order) is ALL the upper tree depends on; both upload paths (host build, device build) hand over the same list.
struct alignas(16) LeafBox { float bmin[3]; uint32_t start; float bmax[3]; uint32_t count; };
static_assert(sizeof(LeafBox) == 32, "leaf record must be 32 B");
// The upper tree the traversal kernels walk (round 2): a binned-SAH binary tree over the reference's leaves (16 bins per axis, cost =
// area x leaves), built on the host cores - deterministic whatever the thread count - then collapsed to 4-wide nodes by
// build_wide_nodes.  Against the collapse of the reference's own midpoint tree it cuts the node visits per ray by 23 % (Cornell scenes)
// to 37 % (tessellated scene) and the slab tests by 32-39 %, hits unchanged (profiles/r2_sah_upper_experiment.txt).
void build_sah_wide(const LeafBox* leaves, size_t n_leaves, std::vector<DevNode4>& out, int32_t& depth);
// a tree collapsed to 4-wide nodes in level order (same result as kernels/bvh_build.cuh: k_wide_*)
void build_wide_nodes(const RefTree& tree, std::vector<DevNode4>& out, int32_t& depth);
// surface-area measure that orders the collapse: ((dx*dy) + (dy*dz)) + (dz*dx), fp32, no contraction (host and device alike)
#if defined(__CUDACC__)
__host__ __device__
#endif
inline float wide_box_area(const float mn[3], const float mx[3]) {
    const float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
    return (dx * dy + dy * dz) + dz * dx;
}

bool derive_materials(const trb_scene_desc& s, std::vector<DevMaterial>& out, std::string& err);
bool derive_emitters(const trb_scene_desc& s, const std::vector<DevMaterial>& mats, std::vector<DevEmitter>& em,
                     std::vector<DevEmitterTri>& tris, std::vector<float>& cdf, std::vector<int32_t>& shape_to_emitter,
                     std::string& err);
// Scene::load's tail (renderer.cpp:294-304): per shape the mean of the vertices its indices name (running fp32 sum in index
// order) and radius = AABB max.x - centre.x; out4 = n_shapes x (cx, cy, cz, r)
void shape_spheres(const trb_scene_desc& s, std::vector<float>& out4);
bool flatten_scene(const trb_scene_desc& s, FlatScene& out, std::string& err);      // = tables + geometry
bool flatten_tables(const trb_scene_desc& s, FlatScene& out, std::string& err);
bool flatten_geometry(const trb_scene_desc& s, FlatScene& out, std::string& err);

// ---- camera (glm-order arithmetic; renderer.cpp:143-153, bdpt.h:49-55,485-496) --------------------
struct CameraDev {
    float eye[3]; float tan_half_fov;     // angle = tanf(deg2rad*fov*0.5f)
    float forward[3]; float aspect;       // normalize(at - o); (float)W / H
    float c2w[16];                        // cameraToWorld, column-major
    float w2c[16];                        // worldToCamera
    float clip[16];                       // cameraToClip = perspective(deg2rad*fov, aspect, 1, 1000)
    float ndc2screen[16];                 // scale(W,H,1)*scale(.5,-.5,1)*translate(1,-1,0)
    float inv_w, inv_h;                   // 1.f/W, 1.f/H
    float vnear_dist;                     // (1/tanf(deg2rad*fov*.5f)) * H * .5f   (bdpt.h:52)
    int32_t width, height;
    int32_t pad[3];
};
void make_camera(const trb_render_desc& d, CameraDev& out);

// ---- TOML / EXR ---------------------------------------------------------------------------------
struct TomlConfig {
    trb_render_desc desc;
    trb_path_desc path;  // [renderer] keys of type = "path" (main.cpp:94-102)
    trb_direct_desc direct;  // [renderer] keys of type = "direct" (main.cpp:88-93)
    std::string obj_path, type;
    bool realtime = false;
    int gpus = 1;
    // restartable renders (extension keys): renderer.checkpoint = "file", renderer.checkpointEvery = spp, renderer.preview = "file.png"
    std::string checkpoint, preview;
    int checkpoint_every = 0;
};
bool load_toml(const std::string& path, TomlConfig& out, std::string& err);
bool save_checkpoint(const std::string& path, const trb_checkpoint_info& info, const float* rgb, std::string& err);
bool load_checkpoint(const std::string& path, trb_checkpoint_info& info, float* rgb, size_t capacity, std::string& err);
bool checkpoint_matches(const trb_checkpoint_info& a, const trb_checkpoint_info& b);
bool save_png_srgb(const std::string& path, const float* rgb, int w, int h, float exposure, std::string& err);
bool save_exr_bgr_half(const std::string& path, const float* rgb, int w, int h, std::string& err);

}  // namespace trb
