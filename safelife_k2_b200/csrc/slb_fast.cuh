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
template <int HW>  // HW = 26: the benchmark board size at compile time; 0: runtime shape
__global__ void __launch_bounds__(FAST_WARPS * 32)
advance_fast_kernel(const uint16_t *__restrict__ in, uint16_t *__restrict__ out, int n, int h_, int w_,
                    const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                    uint64_t *counters, int k_steps, bool vec) {
    __shared__ __align__(16) uint16_t stage[FAST_WARPS][FAST_STAGE_U16];
    const int warp = threadIdx.x >> 5, i = blockIdx.x * FAST_WARPS + warp;
    if (i >= n) return;
    const int h = HW ? HW : h_, w = HW ? HW : w_;
    const Geo g = make_geo(h, w);
    const size_t cells = (size_t)h * w;
    uint32_t P[16];
    load_planes<HW>(in + i * cells, stage[warp], h, w, vec, P);
    uint64_t counter = counters ? counters[i] : 0;
    const float p = spawn_prob[i];
    uint32_t touched = 0;  // this lane's cells that flipped (single step), or all ones once anything changed
    if (k_steps == 1) {
        PhysDelta dl;
        bool quiet;
        counter = physics_compute(P, g, p, seed, env_id0 + i, counter, dl, quiet);
        if (__any_sync(FULL, dl.flip != 0)) {
            physics_apply(P, dl);
            touched = dl.flip;
        }
    } else {
        bool ch, any_ch = false;
        for (int k = 0; k < k_steps; k++) {
            counter = physics_planes(P, g, p, seed, env_id0 + i, counter, ch);
            any_ch |= ch;
        }
        if (any_ch) touched = 0xffffffffu;  // several steps: take the full back-transposition
    }
    const int lane = threadIdx.x & 31;
    const int most = __reduce_max_sync(FULL, __popc(touched));  // busiest row
    if (most > 8) {
        store_planes<HW>(out + i * cells, stage[warp], h, w, vec, P);
    } else {
        // Few cells changed (or none): the staging area still holds the input image.  A cell that flipped
        // lost every bit and, where it was born, got alive / destructible / colour bits only (physics_apply),
        // so its final value is read off five planes and patched into the image, which is then copied out —
        // no back-transposition of the planes.
        const int S = row_stride(w);
        uint16_t *sm = stage[warp];
        uint32_t t = touched;
        while (t) {
            const int x = __ffs(t) - 1;
            t &= t - 1;
            const uint32_t v = ((P[0] >> x) & 1u) | (((P[3] >> x) & 1u) << 3) | (((P[9] >> x) & 1u) << 9) |
                               (((P[10] >> x) & 1u) << 10) | (((P[11] >> x) & 1u) << 11);
            sm[lane * S + x] = (uint16_t)v;
        }
        __syncwarp();
        uint16_t *dst = out + i * cells;
        if (vec) {
            const uint2 *s2 = reinterpret_cast<const uint2 *>(sm);
            uint2 *d2 = reinterpret_cast<uint2 *>(dst);
            for (int j = lane; j < ((int)cells >> 2); j += 32) d2[j] = s2[j];
        } else {
            for (int e = lane; e < (int)cells; e += 32) {
                int r = e / w, c = e - r * w;
                dst[e] = sm[r * S + c];
            }
        }
    }
    if (counters && (threadIdx.x & 31) == 0) counters[i] = counter;
}

inline void launch_advance_fast(const uint16_t *in, uint16_t *out, int n, int h, int w, const float *p,
                                uint64_t seed, uint64_t env0, uint64_t *ctr, int k, cudaStream_t st) {
    int grid = (n + FAST_WARPS - 1) / FAST_WARPS;
    if (h == 26 && w == 26)
        advance_fast_kernel<26><<<grid, FAST_WARPS * 32, 0, st>>>(in, out, n, h, w, p, seed, env0, ctr, k,
                                                                  fast_vec_ok(h, w, in, out));
    else
        advance_fast_kernel<0><<<grid, FAST_WARPS * 32, 0, st>>>(in, out, n, h, w, p, seed, env0, ctr, k,
                                                                 fast_vec_ok(h, w, in, out));
}

}  // namespace slb

#include "slb_fast_side.cuh"
