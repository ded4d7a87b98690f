/* trb.h -- C ABI of the B200-native BDPT hot path ("TinyRender on B200").
 *
 * This is the drop-in boundary underneath TinyRender's C++ host API (the headers under include/tinyrender).  The
 * reference (JackMinn/Bidirectional-Path-Tracing) has no FFI layer: its operator API is in-process C++
 * (Scene / Integrator / BSDF / Renderer).  Each entry point below cites the reference interface it
 * replaces (file:line under the reference root).
 *
 * Conventions: extern "C"; plain pointers and sizes; every function returns 0 (TRB_OK) or a negative
 * trb_status; the message is available from trb_last_error().  No exception crosses the boundary.  The
 * caller owns every host buffer it passes; the context owns all device memory.  A context is bound to
 * ONE device and is not thread-safe (one caller thread per context).  There is no CPU fallback: every
 * compute entry point fails with TRB_ERR_NO_DEVICE when no sm_100 device is usable.
 */
#ifndef TRB_H
#define TRB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum trb_status {
    TRB_OK = 0,
    TRB_ERR_INVALID = -1,   /* bad argument / inconsistent description            */
    TRB_ERR_NO_DEVICE = -2, /* no usable CUDA device (the product has no CPU path) */
    TRB_ERR_CUDA = -3,      /* CUDA runtime error; see trb_last_error              */
    TRB_ERR_NO_SCENE = -4,  /* compute call before trb_scene_upload                */
    TRB_ERR_IO = -5,        /* file could not be read / written / parsed           */
    TRB_ERR_UNSUPPORTED = -6 /* e.g. illum 5 (reference builds no BSDF, renderer.cpp:259-270) */
} trb_status;

typedef struct trb_context trb_context;       /* device side: one GPU, its scene, its wave buffers */
typedef struct trb_host_scene trb_host_scene; /* host side: what Scene::load produces              */

/* ---- plain-old-data mirrors of the reference's types ------------------------------------------- */

/* Ray (src/core/core.h:117-122): origin, direction, [min_t, max_t]. 32 bytes. */
typedef struct trb_ray {
    float o[3];
    float d[3];
    float min_t, max_t;
} trb_ray;

/* Result of AcceleratorBVH::intersect (src/core/accel.h:125-172): hit flag, SurfaceInteraction::{t,u,v,
 * shapeID,primID}.  `tri` is the shape-major global triangle number (sum of earlier shapes' face counts +
 * primID).  On a miss shape = prim = tri = -1 and t follows the reference: FLT_MAX when nothing was hit,
 * the (out-of-range) closest t when the closest hit failed the [min_t,max_t] test (accel.h:131-133,170). */
typedef struct trb_hit {
    float t, u, v;
    int32_t hit;
    int32_t shape, prim, tri;
    int32_t _pad;
} trb_hit;

/* tinyobj::material_t fields the path reads (diffuse.h:20-32, phong.h:25-52, glass.h:20-31, renderer.cpp:229-231). */
typedef struct trb_material {
    float kd[3];  /* Kd  */
    float ks[3];  /* Ks  */
    float ke[3];  /* Ke  : emission                                    */
    float tf[3];  /* Tf/Kt: glass transmittance                        */
    float ns;     /* Ns  : Phong exponent                              */
    float ni;     /* Ni  : index of refraction                         */
    int32_t illum; /* 7 diffuse, 3 mirror, 6 glass, 8 mixture, 5 unsupported, else phong (renderer.cpp:259-271) */
    int32_t tex;   /* bitmap textures (diffuse_texname / specular_texname, diffuse.h:23-26, phong.h:28-36):
                      (1 + index of the map_Kd texture) | (1 + index of the map_Ks texture) << 16; 0 = the constants above */
} trb_material;

/* Tex (core.h:417-497) as BitmapTexture3f holds it: width*height RGB floats, gamma-expanded (pow(byte / maxval, 2.2f)) and
 * vertically flipped, i.e. row 0 is the LAST row of the PPM file. */
typedef struct trb_texture {
    const float* rgb;
    int32_t width, height;
} trb_texture;

/* What tinyobj::LoadObj(triangulate=true) + Scene::load hold (src/core/renderer.cpp:235-315), flattened.
 * Triangles are listed shape-major in file order: tri_shape[] is non-decreasing. */
typedef struct trb_scene_desc {
    const float* positions;   /* attrib.vertices, 3 floats per vertex  */
    size_t n_positions;       /* number of vertices                    */
    const float* normals;     /* attrib.normals, 3 floats per normal   */
    size_t n_normals;
    const int32_t* tri_v;     /* 3 vertex indices per triangle         */
    const int32_t* tri_n;     /* 3 normal indices per triangle         */
    const int32_t* tri_mat;   /* mesh.material_ids                     */
    const int32_t* tri_shape; /* owning shape                          */
    size_t n_tris;
    const trb_material* materials;
    size_t n_materials;
    size_t n_shapes;
    /* bitmap textures; all zero / NULL for an untextured scene */
    const float* texcoords;   /* attrib.texcoords, 2 floats per vt               */
    size_t n_texcoords;
    const int32_t* tri_t;     /* 3 texcoord indices per triangle, -1 = none; may be NULL when n_textures == 0 */
    const trb_texture* textures;
    size_t n_textures;
} trb_scene_desc;

/* Config (src/core/core.h:195-248) restricted to what BDPT reads, plus the sample range that makes the
 * call shardable by sample index.  Contributions are always normalised by the TOTAL spp, so partial
 * frame buffers of disjoint [spp_begin, spp_end) ranges add up to the full render. */
typedef struct trb_render_desc {
    float eye[3], at[3], up[3];
    float fov_deg;
    int32_t width, height;
    int32_t spp;       /* total samples per pixel (the normaliser, bdpt.h:332, renderer.cpp:201)       */
    int32_t rr_depth;  /* renderer.rrDepth: subpaths have at most rr_depth-1 vertices (bdpt.h:68,188)  */
    int32_t spp_begin; /* first sample index rendered by this call (0 for a whole render)              */
    int32_t spp_end;   /* one past the last; 0 means spp                                                */
    uint32_t seed;     /* key of the counter-based RNG; streams are keyed by (seed, pixel, sample, dim) */
    uint32_t flags;    /* TRB_FLAG_*                                                                    */
    int64_t wave_samples; /* samples per wavefront launch; 0 = library default: up to 2^24, sized to a quarter of the free device memory */
    int32_t max_depth; /* TRB_FLAG_RUSSIAN_ROULETTE only: hard cap on the walk depth (the reference has none,
                          bdpt.h:66-67); 0 = TRB_RR_DEFAULT_MAX_DEPTH; rr_depth <= max_depth <= 64            */
    int32_t reserved;
} trb_render_desc;

#define TRB_FLAG_NONE 0u
#define TRB_FLAG_NO_JITTER 1u /* force the spp==1 pixel-centre path of renderer.cpp:169-180 */
#define TRB_FLAG_NO_OVERLAP 2u /* one stream: every kernel runs alone (per-kernel timing for the roofline line); same results */
/* Russian roulette as the reference writes it but compiles out (`#define NO_RR 1`, bdpt.h:18): past rr_depth a walk
 * continues with probability rr = (luminance(throughput) < 0.01 ? 0.5 : 1) (bdpt.h:68,129,188,201), pdfs and reverse
 * pdfs are scaled by the vertex's rr (bdpt.h:250,272,342,410,417,461-472).  Off (the default) is the shipped
 * behaviour: rr_depth is a hard cap.  renderer.rrProb is read but never used by the reference (bdpt.h:40). */
#define TRB_FLAG_RUSSIAN_ROULETTE 4u
#define TRB_RR_DEFAULT_MAX_DEPTH 32
/* The reference's estimator A/B switches (`#define LIGHT_TRACING 0`, `#define PATH_TRACING 0`, bdpt.h:16-17), which its author flips to
 * check the bidirectional estimator against its two halves.  LIGHT_TRACING: only the light walk runs; splats carry no MIS weight and Li is
 * the emission of the primary hit (bdpt.h:229-231, 355-357).  PATH_TRACING: only the eye walk runs; no vertex connections, next-event
 * estimation and emitter hits carry no MIS weight, and an emitter hit counts only on a purely specular path (bdpt.h:95-118, 145-149,
 * 232-233, 426-428).  Setting both is refused (the reference's #if / #elif would take LIGHT_TRACING with PATH_TRACING's weights). */
#define TRB_FLAG_LIGHT_TRACING 16u
#define TRB_FLAG_PATH_TRACING 32u
/* Accepted and ignored.  (Round 1 had an experiment behind it: a counting sort of each ray queue before traversal; the traversal got
 * 10-15 % faster and the sort cost more than that - profiles/r1f_tuning_variants.md - so the code was removed in round 2.) */
#define TRB_FLAG_SORT_RAYS 8u

/* PathTracerIntegrator settings (Config::integratorSettings.pt, core.h:229-236; TOML defaults main.cpp:94-102).
 * rrDepth comes from trb_render_desc.rr_depth.  The reference recurses without bound when max_depth == -1; the device
 * keeps at most trb_render_desc.max_depth vertices per path (0 = TRB_PT_DEFAULT_DEPTH_CAP, <= 256). */
typedef struct trb_path_desc {
    int32_t is_explicit;     /* renderer.isExplicit (default true): recursiveExplicit, else recursiveImplicit      */
    int32_t max_depth;       /* renderer.maxDepth (default -1 = Russian roulette past rrDepth with rr_prob)         */
    float rr_prob;           /* renderer.rrProb (default 0.95)                                                      */
    int32_t emitter_samples; /* renderer.emitterSamples (default 1), 0..16                                          */
    int32_t bsdf_samples;    /* renderer.bsdfSamples (default 0), 0..16                                             */
} trb_path_desc;
#define TRB_PT_DEFAULT_DEPTH_CAP 64

/* DirectIntegrator settings (Config::integratorSettings.di, core.h; TOML keys main.cpp:88-93).  The reference exits with
 * "Error: wrong strategy" for any other name (direct.h:460-461) - its own TOML default, "emitter", included. */
typedef struct trb_direct_desc {
    int32_t strategy;        /* renderer.samplingStrategy: TRB_DI_*                                                   */
    int32_t emitter_samples; /* renderer.emitterSamples (default 1), 0..64; >= 1 unless strategy is mis or bsdf        */
    int32_t bsdf_samples;    /* renderer.bsdfSamples (default 1), 0..64; >= 1 for strategy bsdf                        */
    int32_t reserved;
} trb_direct_desc;
#define TRB_DI_MIS 0               /* "mis"              direct.h:340-447 */
#define TRB_DI_AREA 1              /* "area"             direct.h:143-197 */
#define TRB_DI_SOLID_ANGLE 2       /* "solidAngle"       direct.h:271-338 */
#define TRB_DI_COSINE_HEMISPHERE 3 /* "cosineHemisphere" direct.h:200-234 */
#define TRB_DI_BSDF 4              /* "bsdf"             direct.h:236-268 */
#define TRB_DI_INVALID (-1)        /* any other name                      */

typedef struct trb_stats {
    double ms_total;          /* device time of the whole call, CUDA events            */
    double ms_trace_closest;  /* summed over launches                                   */
    double ms_trace_any;
    double ms_logic;          /* generation / bounce / connection kernels               */
    uint64_t samples;         /* Integrator::render calls                               */
    uint64_t rays_closest;    /* BVH closest-hit queries traced                         */
    uint64_t rays_any;        /* occlusion queries traced                               */
    uint64_t launches;        /* kernels launched by this call                          */
    uint64_t node_visits;     /* closest-hit rays; only when built with TRB_COUNTERS (else 0) */
    uint64_t tri_tests;
    uint64_t node_visits_any; /* occlusion rays; TRB_COUNTERS builds only                */
    uint64_t tri_tests_any;
    uint64_t retraces;        /* closest-hit rays re-traced without distance pruning (ill-conditioned candidate met); TRB_COUNTERS only */
    uint64_t retraces_any;    /* same for occlusion rays */
    /* node_visits / node_visits_any count 128-byte nodes of the 4-wide upper tree (k_trace_wide).  Rays with a zero / non-finite
     * direction component are traced on the reference's binary tree instead (k_trace_coop, 64-byte nodes): */
    uint64_t rays_deferred;      /* closest-hit rays traced on the binary tree (always counted)               */
    uint64_t rays_deferred_any;  /* occlusion rays traced on the binary tree (always counted)                 */
    uint64_t node_visits_bin;    /* their 64-byte node visits and triangle tests; TRB_COUNTERS builds only     */
    uint64_t tri_tests_bin;
    uint64_t node_visits_bin_any;
    uint64_t tri_tests_bin_any;
    uint64_t wave_samples;       /* samples per wavefront this call used (trb_render_desc.wave_samples, or the library's choice) */
} trb_stats;

/* BSDF query in the local shading frame (SurfaceInteraction::{wo,wi}, core.h:173-180) */
typedef struct trb_bsdf_query {
    float wo[3];
    float wi[3];
    float u[2];     /* the v2f handed to BSDF::sample */
    int32_t mat;
    int32_t _pad;
} trb_bsdf_query;

typedef struct trb_bsdf_out {
    float eval[3];       /* BSDF::eval  (f * cos)          */
    float pdf;           /* BSDF::pdf                       */
    float sample_val[3]; /* BSDF::sample return value       */
    float sample_pdf;    /* *pdf written by BSDF::sample    */
    float sample_wi[3];  /* i.wi written by BSDF::sample    */
    uint32_t type;       /* BSDF::getType()                 */
} trb_bsdf_out;

/* ---- host side (replaces tinyobj::LoadObj + Scene::load, renderer.cpp:235-315; loadTOML, main.cpp:22-116) */
int trb_host_load_obj(const char* obj_path, trb_host_scene** out);
int trb_host_scene_desc(const trb_host_scene*, trb_scene_desc* out); /* pointers stay owned by the host scene */
int trb_host_shape_name(const trb_host_scene*, size_t shape, char* buf, size_t buflen);
void trb_host_scene_free(trb_host_scene*);
/* Parses the reference's TOML scene file. obj_path_out receives input.objfile resolved against the
 * toml's directory (renderer.cpp:236-241). type_out receives renderer.type. */
int trb_host_load_toml(const char* toml_path, trb_render_desc* out, char* obj_path_out, size_t obj_len,
                       char* type_out, size_t type_len, int* realtime_out);
/* the `type = "path"` keys of the same file (main.cpp:94-102: isExplicit, maxDepth, rrProb, emitterSamples, bsdfSamples) */
int trb_host_load_toml_path(const char* toml_path, struct trb_path_desc* out);
/* the `type = "direct"` keys of the same file (main.cpp:88-93: emitterSamples, bsdfSamples, samplingStrategy); an unknown
 * strategy name is reported as TRB_DI_INVALID, which every render entry point refuses */
int trb_host_load_toml_direct(const char* toml_path, struct trb_direct_desc* out);
/* restartable renders.  A checkpoint holds the running-sum frame buffer, the number of samples per pixel in it
 * and the description of the render; trb_host_checkpoint_matches says whether a description continues it.  Resume = restore the
 * buffer (trb_framebuffer_upload), then render [spp_done, spp) with accumulate = 1.  The save is write-then-rename. */
typedef struct trb_checkpoint_info {
    trb_render_desc desc;
    trb_path_desc path;     /* integrator == 1; integrator == 2 keeps its trb_direct_desc here: is_explicit = strategy,
                               emitter_samples, bsdf_samples as named, the rest 0 */
    int32_t integrator;     /* 0 = bdpt, 1 = path, 2 = direct */
    int32_t spp_done;
    uint64_t scene_hash;    /* caller's fingerprint of the scene (0 = unchecked) */
} trb_checkpoint_info;
int trb_host_save_checkpoint(const char* path, const trb_checkpoint_info* info, const float* rgb);
/* rgb_out may be null to read the header only */
int trb_host_load_checkpoint(const char* path, trb_checkpoint_info* info_out, float* rgb_out, size_t capacity_floats);
int trb_host_checkpoint_matches(const trb_checkpoint_info* a, const trb_checkpoint_info* b);
/* Fingerprint of a scene for trb_checkpoint_info::scene_hash: FNV-1a over the CONTENT of every array of the description (positions,
 * normals, indices, material and shape ids, material records, texture coordinates and texels) - an edited vertex, Kd / Ke / illum
 * or texture changes it, so a checkpoint of the old scene is not continued on the new one.  Never 0. */
uint64_t trb_host_scene_hash(const trb_scene_desc* scene);
/* quick-look image: 8-bit sRGB PNG of exposure * rgb (row 0 = top) */
int trb_host_save_png(const char* path, const float* rgb, int width, int height, float exposure);
/* saveEXR (src/core/utils.h:95-156): channels B,G,R stored as half, row 0 = top. */
int trb_host_save_exr(const char* path, const float* rgb, int width, int height);
const char* trb_host_last_error(void);

/* Host-side derived data, exposed so tests can pin them against the reference (no GPU needed):
 * the build-reordered primitive order of BVH::build (externals/bvh.h:147-247): order[pos] = global tri id */
int trb_host_bvh_order(const trb_scene_desc*, int32_t* order_out, size_t n);
/* emitters as Scene::load builds them (renderer.cpp:279-305): returns count; per emitter shape id, area,
 * radiance(3); cdf_out receives the concatenated normalised face-area CDFs (n_faces+1 each) if non-null */
int trb_host_emitters(const trb_scene_desc*, int32_t* shape_out, float* area_out, float* radiance_out,
                      size_t max_emitters, float* cdf_out, size_t cdf_capacity);
/* Scene::getShapeCenter / getShapeRadius (renderer.cpp:294-304, 349-358) per shape: out4 = n_shapes x (cx, cy, cz, radius);
 * DirectIntegrator treats every emitter as the sphere so described */
int trb_host_shape_spheres(const trb_scene_desc*, float* out4, size_t n_shapes);
/* camera matrices as glm builds them (renderer.cpp:143-153, bdpt.h:485-490): 4 column-major mat4:
 * worldToCamera, cameraToWorld, cameraToClip, NDCToScreen */
int trb_host_camera_matrices(const trb_render_desc*, float* out64);

/* ---- device side ------------------------------------------------------------------------------- */
int trb_create(int device, trb_context** out);
void trb_destroy(trb_context*);
const char* trb_last_error(const trb_context*);

/* Scene::load's tail + AcceleratorBVH::build (renderer.cpp:259-312, accel.h:115-123): derives BSDF
 * parameters, emitters and the BVH on the host and uploads the flattened scene. */
int trb_scene_upload(trb_context*, const trb_scene_desc*);   /* = trb_scene_upload_ex(.., TRB_UPLOAD_AUTO) */
/* The same with the place of BVH::build chosen.  The device build (csrc/kernels/bvh_build.cuh) produces the
 * reference's exact tree and primitive order one level at a time, then gathers the triangle records on the device; the host
 * build does it on the host cores.  Both give the same device scene (tests compare them through trb_scene_download). */
#define TRB_UPLOAD_AUTO 0u          /* device build from TRB_DEVICE_BUILD_MIN_TRIS triangles on, host build below */
#define TRB_UPLOAD_HOST_BUILD 1u
#define TRB_UPLOAD_DEVICE_BUILD 2u
/* OR-able: the 4-wide upper tree is the collapse of the reference's own (midpoint-split) tree, made where the BVH is built, instead of
 * the SAH tree over the reference's leaves that the host cores build (the default: 23-37 % fewer node visits per ray; it adds about
 * 0.1 s per million triangles to the upload).  Same hits either way. */
#define TRB_UPLOAD_WIDE_COLLAPSE 0x100u
#define TRB_DEVICE_BUILD_MIN_TRIS 65536u
int trb_scene_upload_ex(trb_context*, const trb_scene_desc*, uint32_t flags);
int trb_scene_built_on_device(const trb_context*);
/* Test hook: geometry arrays of the device scene read back. which: 0 nodes (64 B records), 1 triangles (48 B), 2 shading records
 * (96 B), 3 texture coordinates (32 B).  out may be NULL to query the count. */
int trb_scene_download(trb_context*, int which, void* out, size_t capacity_bytes, size_t* count_out);
/* Copies the already flattened scene (kept in pinned host memory by trb_scene_upload) to the device again:
 * the per-frame host->device transfer of an end-to-end render without redoing the host-side BVH build. */
int trb_scene_refresh(trb_context*);
/* bytes of the flattened device scene (nodes + triangles + shading records + tables) */
size_t trb_scene_device_bytes(const trb_context*);

/* Renderer::render offline branch + BDPTIntegrator (renderer.cpp:130-214, bdpt.h:37-518).
 * rgb_out: host, width*height*3 floats, row 0 = top (RenderBuffer, core.h:128-146); may be NULL.
 * The device frame buffer is cleared first unless accumulate != 0. */
int trb_render_bdpt(trb_context*, const trb_render_desc*, int accumulate, float* rgb_out, trb_stats* stats);
/* Same, result left in a caller-provided DEVICE buffer (width*height*3 floats) - used for the multi-GPU
 * NCCL reduce and the device-resident throughput measurement. */
int trb_render_bdpt_device(trb_context*, const trb_render_desc*, int accumulate, float* d_rgb, trb_stats* stats);

/* One frame on n GPUs from ONE process: context g (one per device) renders the sample indices
 * [g*spp/n, (g+1)*spp/n) of every pixel, then ONE ncclReduce(sum) over NVLink brings the partial frame buffers together on
 * ctxs[0]'s device; rgb_out (host) receives the sum.  Every context must hold the same scene.  libnccl.so.2 is loaded at run
 * time ($TRB_NCCL_LIB overrides the name); TRB_ERR_UNSUPPORTED when it is absent.  Replaces the T-thread pixel loop +
 * g_FrameBufferLocks of renderer.cpp:130-214 / main.cpp:137-143 at node scale. */
int trb_render_bdpt_multi(trb_context* const* ctxs, int n, const trb_render_desc*, float* rgb_out, trb_stats* stats);
int trb_render_path_multi(trb_context* const* ctxs, int n, const trb_render_desc*, const trb_path_desc*, float* rgb_out, trb_stats* stats);

/* Restore the context's frame buffer (the one trb_render_*(accumulate = 1, host buffer out) adds to), e.g. from a checkpoint. */
int trb_framebuffer_upload(trb_context*, const float* rgb, int width, int height);

/* The reference's unidirectional path tracer (src/integrators/path.h:15-262, `type = "path"`) on the same wavefront:
 * the independent estimator the reference checks BDPT against (cbox_bdpt_path.toml).  Same conventions as trb_render_bdpt. */
int trb_render_path(trb_context*, const trb_render_desc*, const trb_path_desc*, int accumulate, float* rgb_out, trb_stats* out);
int trb_render_path_device(trb_context*, const trb_render_desc*, const trb_path_desc*, int accumulate, float* d_rgb, trb_stats* out);
/* PathTracerIntegrator::render on n given camera rays with injected random numbers (n x stride floats, consumed in the
 * reference's depth-first order); li_out n x 3; draws_out optional.  A path that needs more than `stride` numbers reads
 * 0.99 past the end of its row (which ends every roulette); draws_out then exceeds stride. */
int trb_path_paths(trb_context*, const trb_render_desc*, const trb_path_desc*, const trb_ray* cam_rays, const float* rng, size_t n,
                   size_t stride, float* li_out, int32_t* draws_out);

/* The reference's DirectIntegrator (src/integrators/direct.h:74-467, `type = "direct"`): one-bounce direct
 * illumination, five sampling strategies, every emitter treated as the sphere (shape centre, AABB max.x - centre.x).  Same
 * conventions as trb_render_bdpt; rr_depth is not read. */
int trb_render_direct(trb_context*, const trb_render_desc*, const trb_direct_desc*, int accumulate, float* rgb_out, trb_stats* out);
int trb_render_direct_device(trb_context*, const trb_render_desc*, const trb_direct_desc*, int accumulate, float* d_rgb, trb_stats* out);
int trb_render_direct_multi(trb_context* const* ctxs, int n, const trb_render_desc*, const trb_direct_desc*, float* rgb_out, trb_stats* stats);
/* DirectIntegrator::render on n given camera rays with injected random numbers (n x stride floats, consumed in the
 * reference's order: the emitter loop's (1 +) 2 numbers per sample, then the BSDF loop's 2); li_out n x 3; draws_out optional. */
int trb_direct_paths(trb_context*, const trb_render_desc*, const trb_direct_desc*, const trb_ray* cam_rays, const float* rng, size_t n,
                     size_t stride, float* li_out, int32_t* draws_out);

/* Parity entry points (north_star levels 1 and 2). Host buffers in, host buffers out. */
int trb_trace_closest(trb_context*, const trb_ray* rays, size_t n, trb_hit* out);   /* accel.h:125-172 */
int trb_trace_any(trb_context*, const trb_ray* rays, size_t n, uint8_t* out);       /* bvh.h:259-352, occlusion=true (bdpt.h:505) */
int trb_bsdf_batch(trb_context*, const trb_bsdf_query* q, size_t n, trb_bsdf_out* out); /* BSDF::{eval,pdf,sample,getType} */
/* BDPTIntegrator::render on n given camera rays with injected random numbers: rng is n x stride floats,
 * consumed per path in the reference's order.  li_out: n x 3.  Light-tracing splats:
 * splat_pix n x K pixel indices (-1 = none), splat_rgb n x K x 3, slot = light depth-1, with K = max(rr_depth-1, 1)
 * (max(max_depth-1, 1) under TRB_FLAG_RUSSIAN_ROULETTE); stride >= 10 + 10 K.
 * draws_out (optional): random numbers consumed per path. */
int trb_bdpt_paths(trb_context*, const trb_render_desc*, const trb_ray* cam_rays, const float* rng, size_t n,
                   size_t stride, float* li_out, int32_t* splat_pix, float* splat_rgb, int32_t* draws_out);

/* Device-resident variants used by the benchmark (pointers are DEVICE pointers). */
int trb_trace_closest_device(trb_context*, const trb_ray* d_rays, size_t n, trb_hit* d_out, float* ms_out);
int trb_trace_any_device(trb_context*, const trb_ray* d_rays, size_t n, uint8_t* d_out, float* ms_out);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same query and results as trb_trace_any_device (d_rays and
 * d_out are DEVICE pointers the caller owns; d_out[i] = 1 when ray i is occluded), but the call allocates NOTHING: the traversal's
 * scratch (queue counters + the queue of rays handed to the binary tree) lives in a fixed device array inside the library, which
 * every GPU that loads the library holds until exit (4 MiB + 256 B).  n > TRB_NOALLOC_MAX_RAYS is refused with TRB_ERR_INVALID.
 * All calls on a device share that array: calls that overlap on one GPU race; calls on one context (one stream) do not. */
#define TRB_NOALLOC_MAX_RAYS (1u << 20)
int trb_trace_any_device_noalloc(trb_context*, const trb_ray* d_rays, size_t n, uint8_t* d_out);

/* library build info: "sm_100a;fmad=off;..." */
const char* trb_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* TRB_H */
