// slb_fast.cuh — speedups.advance_board for boards with 4 <= H, W <= 32 in the reference's uint16
// layout: one warp per board, cells -> bit-planes in registers on the way in, k steps on chip,
// planes -> cells on the way out (one HBM round trip for k steps).
#pragma once
#include "slb_resident.cuh"

namespace slb {

constexpr int FAST_WARPS = 4;                 // warps (= boards) per CTA
constexpr int FAST_STAGE_U16 = 32 * 34;       // per-warp staging area (H <= 32 rows, stride <= 34)

// all boards 8-byte aligned and the shared image equals the memory image
__host__ __device__ inline bool fast_vec_ok(int h, int w, const void *a, const void *b) {
    return row_stride(w) == w && ((size_t)h * w * 2) % 8 == 0 && ((uintptr_t)a % 8) == 0 &&
           ((uintptr_t)b % 8) == 0;
}

// ---------------------------------------------------------------------------------------------
// A1: advance_board, k steps on chip.  grid = ceil(N / FAST_WARPS).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FAST_WARPS * 32)
advance_fast_kernel(const uint16_t *__restrict__ in, uint16_t *__restrict__ out, int n, int h, int w,
                    const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                    uint64_t *counters, int k_steps, bool vec) {
    __shared__ __align__(16) uint16_t stage[FAST_WARPS][FAST_STAGE_U16];
    const int warp = threadIdx.x >> 5, i = blockIdx.x * FAST_WARPS + warp;
    if (i >= n) return;
    const Geo g = make_geo(h, w);
    const size_t cells = (size_t)h * w;
    uint32_t P[16];
    load_planes(in + i * cells, stage[warp], h, w, vec, P);
    uint64_t counter = counters ? counters[i] : 0;
    const float p = spawn_prob[i];
    bool ch;
    for (int k = 0; k < k_steps; k++) counter = physics_planes(P, g, p, seed, env_id0 + i, counter, ch);
    store_planes(out + i * cells, stage[warp], h, w, vec, P);
    if (counters && (threadIdx.x & 31) == 0) counters[i] = counter;
}

inline void launch_advance_fast(const uint16_t *in, uint16_t *out, int n, int h, int w, const float *p,
                                uint64_t seed, uint64_t env0, uint64_t *ctr, int k, cudaStream_t st) {
    int grid = (n + FAST_WARPS - 1) / FAST_WARPS;
    advance_fast_kernel<<<grid, FAST_WARPS * 32, 0, st>>>(in, out, n, h, w, p, seed, env0, ctr, k,
                                                          fast_vec_ok(h, w, in, out));
}

}  // namespace slb

#include "slb_fast_side.cuh"
