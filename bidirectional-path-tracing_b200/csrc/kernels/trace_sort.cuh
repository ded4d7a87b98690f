// trace_sort.cuh -- optional re-ordering of a ray queue before traversal (TRB_FLAG_SORT_RAYS; an experiment, off by default:
// B200, serialised pass - traversal 29.0 -> 25.7 / 34.4 -> 31.1 ms (tess 10 M, closest / any) and 51.6 -> 43.9 / 64.5 -> 58.2 ms
// (caustic), but this three-kernel sort costs 28 / 110 ms on the same passes; it would have to run in < 0.1 ms per queue to pay).
// Secondary rays of a wave are incoherent; on a scene whose acceleration structure exceeds L2 (configs[4]: 0.8 GB) every warp
// then walks 32 unrelated root-to-leaf chains (ncu r1k: L1 hit 69 %, L2 hit 52 %, no eligible warp 38 % of the time).
// A counting sort of the QUEUE (ray records stay where they are) by an 18-bit key - Morton code of the origin on a 16^3 grid
// over the scene box, then 2 bits per direction component - puts rays that start in the same cell and head the same way
// next to each other; the persistent warps fetch contiguous chunks, so lanes of a warp share nodes.  Results do not depend on
// the order (hits are per ray; frame-buffer sums are atomic either way).
#pragma once
#include "bdpt_common.cuh"

namespace trbk {

#define TRB_SORT_BINS (1u << 18)

struct SortArgs {
    const RayT* rays;
    const int32_t* q_in;      // slots, or nullptr for the identity
    const uint32_t* n_ptr;    // device count, or nullptr -> n_fixed
    uint32_t n_fixed, cap;
    float mnx, mny, mnz, sx, sy, sz;   // scene box -> [0,16)^3
    uint32_t* keys;           // per queue entry
    uint32_t* hist;           // TRB_SORT_BINS, zeroed by the caller
    int32_t* q_out;
};

__device__ __forceinline__ uint32_t spread4(uint32_t v) {  // 4 bits -> every third bit
    return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6);
}
__device__ __forceinline__ uint32_t dir2(float d) { return d < -0.5f ? 0u : (d < 0.f ? 1u : (d < 0.5f ? 2u : 3u)); }
__device__ __forceinline__ uint32_t sort_key(const SortArgs& a, const RayT& r) {
    const uint32_t cx = uint32_t(fminf(fmaxf((r.ox - a.mnx) * a.sx, 0.f), 15.f));
    const uint32_t cy = uint32_t(fminf(fmaxf((r.oy - a.mny) * a.sy, 0.f), 15.f));
    const uint32_t cz = uint32_t(fminf(fmaxf((r.oz - a.mnz) * a.sz, 0.f), 15.f));
    const uint32_t morton = spread4(cx) | (spread4(cy) << 1) | (spread4(cz) << 2);
    return (morton << 6) | (dir2(r.dx) << 4) | (dir2(r.dy) << 2) | dir2(r.dz);
}
__device__ __forceinline__ uint32_t sort_n(const SortArgs& a) { return min(a.n_ptr ? *a.n_ptr : a.n_fixed, a.cap); }

__global__ void __launch_bounds__(256) k_sort_count(const SortArgs a) {
    const uint32_t n = sort_n(a);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int32_t slot = a.q_in ? a.q_in[i] : int32_t(i);
        const float4* rp = reinterpret_cast<const float4*>(a.rays + slot);
        const float4 r0 = rp[0], r1 = rp[1];
        RayT r; r.ox = r0.x; r.oy = r0.y; r.oz = r0.z; r.dx = r0.w; r.dy = r1.x; r.dz = r1.y; r.min_t = r1.z; r.max_t = r1.w;
        const uint32_t key = sort_key(a, r);
        a.keys[i] = key;
        atomicAdd(a.hist + key, 1u);
    }
}

// exclusive scan of the histogram in place: one block, 1024 threads x 256 bins each
__global__ void __launch_bounds__(1024) k_sort_scan(uint32_t* __restrict__ hist) {
    __shared__ uint32_t part[1024];
    const uint32_t per = TRB_SORT_BINS / 1024u, first = threadIdx.x * per;
    uint32_t s = 0;
    for (uint32_t k = 0; k < per; ++k) s += hist[first + k];
    part[threadIdx.x] = s;
    __syncthreads();
    for (uint32_t d = 1; d < 1024u; d <<= 1) {   // Hillis-Steele inclusive scan
        uint32_t v = threadIdx.x >= d ? part[threadIdx.x - d] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - s;
    for (uint32_t k = 0; k < per; ++k) { const uint32_t c = hist[first + k]; hist[first + k] = run; run += c; }
}

__global__ void __launch_bounds__(256) k_sort_scatter(const SortArgs a) {
    const uint32_t n = sort_n(a);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int32_t slot = a.q_in ? a.q_in[i] : int32_t(i);
        const uint32_t pos = atomicAdd(a.hist + a.keys[i], 1u);
        a.q_out[pos] = slot;
    }
}

}  // namespace trbk
