// tools/host_trace_check.cpp -- DEVELOPMENT AID, not part of the product or of the test-suite's product path.
// Compiles the __host__ __device__ traversal of csrc/kernels/trace.cuh for the CPU so that its hit semantics
// can be compared with the compiled reference on a box without a GPU (this container).  Build:
//   g++ -std=c++17 -O2 -fPIC -shared -ffp-contract=off tools/host_trace_check.cpp \
//       bidirectional-path-tracing_b200/build/{obj_loader,scene_build,camera,toml_exr,host_abi}.o -o /tmp/libhc.so
#include <cstring>
#include <string>
#include "../bidirectional-path-tracing_b200/csrc/host/tr_host.h"
#include "../bidirectional-path-tracing_b200/csrc/kernels/trace.cuh"

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
