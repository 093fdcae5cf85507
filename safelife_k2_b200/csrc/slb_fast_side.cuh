// slb_fast_side.cuh — side_effect_score rollouts on bit-planes (side_effects.py:59-141): one warp
// per finished episode; both boards (inaction b0, action b1) live in registers; the per-cell-type
// count maps are bit-sliced vertical counters in shared memory (one ripple add per matched key and
// sample), transposed to the uint16 [2,K,H,W] output layout once at the end.
#pragma once
#include "slb_planes_env.cuh"

namespace slb {

constexpr int SE_FAST_MAX_KEYS = 64;

struct SideSmem {  // per-warp carve-up of dynamic shared memory (in 32-bit words)
    int kcap, nplanes;  // kcap: number of cell types the shared counters have room for
    __host__ __device__ size_t words() const {
        return (size_t)2 * kcap * nplanes * 32  // counters [branch][key][plane][lane]
               + (size_t)kcap * 16              // per-key plane selectors (0 / ~0)
               + 32 * 34 / 2;                   // uint16 staging area for load/store
    }
};

// _add_cell_distribution (side_effects.py:59-86) for one board.
__device__ __forceinline__ void se_accumulate_planes(const uint32_t P[16], const Geo &g, int branch,
                                                     int kmax, int nplanes, int &nkeys, bool &overflow,
                                                     uint32_t *cnt, uint32_t *sel) {
    const int lane = threadIdx.x & 31;
    // unchanging = (cell & (frozen|destructible|movable)) == frozen  -> zeroed (:61-62);
    // the destructible bit is dropped from every cell (:62); agent cells and zeros are skipped (:68-70)
    const uint32_t keepm = ~(P[4] & ~P[3] & ~P[2] & ~P[15]);
    uint32_t Q[16];
    uint32_t nz = 0;
#pragma unroll
    for (int b = 0; b < 16; b++) {
        Q[b] = (b == 3) ? 0u : (P[b] & keepm);
        nz |= Q[b];
    }
    uint32_t todo = nz & ~Q[1];  // cells still to be attributed to a key
    if (!__any_sync(FULL, todo != 0)) return;
    for (int k = 0;; k++) {
        if (k == nkeys) {
            // discover the next cell type: first unattributed cell in raster order
            unsigned lanes = __ballot_sync(FULL, todo != 0);
            if (!lanes) break;
            if (nkeys == kmax) { overflow = true; break; }
            const int src = __ffs(lanes) - 1;
            const int x = __ffs(__shfl_sync(FULL, todo, src)) - 1;
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 16; b++) v |= ((Q[b] >> x) & 1u) << b;
            v = __shfl_sync(FULL, v, src);
            if (lane < 16) sel[k * 16 + lane] = ((v >> lane) & 1u) ? 0xffffffffu : 0u;
            for (int j = 0; j < 2 * nplanes; j++) {  // zero both branches' counters of the new key
                int br = j / nplanes, p = j - br * nplanes;
                cnt[((size_t)(br * kmax + k) * nplanes + p) * 32 + lane] = 0;
            }
            nkeys++;
            __syncwarp();
        }
        const uint4 *s4 = reinterpret_cast<const uint4 *>(sel + k * 16);
        uint32_t diff = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint4 s = s4[q];
            diff |= (Q[4 * q] ^ s.x) | (Q[4 * q + 1] ^ s.y) | (Q[4 * q + 2] ^ s.z) | (Q[4 * q + 3] ^ s.w);
        }
        uint32_t carry = todo & ~diff;
        todo &= diff;
        uint32_t *c = cnt + ((size_t)(branch * kmax + k) * nplanes) * 32 + lane;
        for (int p = 0; p < nplanes && __any_sync(FULL, carry != 0); p++) {
            uint32_t v = c[p * 32];
            c[p * 32] = v ^ carry;
            carry &= v;
        }
        if (!__any_sync(FULL, todo != 0)) break;
    }
}

__global__ void __launch_bounds__(128)
side_effect_fast_kernel(const uint16_t *__restrict__ board0, const uint16_t *__restrict__ board1, int n,
                        int h, int w, const int32_t *__restrict__ num_steps, int num_samples,
                        const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                        uint64_t *counters, int kmax, int kcap, int retry, int nplanes, uint16_t *keys,
                        uint16_t *counts, int32_t *nkeys_out, bool vec) {
    extern __shared__ __align__(16) uint32_t se_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + warp;
    if (i >= n) return;
    // kcap < kmax is the first pass: most episodes see a handful of cell types, and counters for a
    // handful fit 4 warps per CTA and 2 CTAs per SM.  An episode that needs more reports -1 and leaves
    // its draw counter alone; the retry pass (full capacity, few warps per SM) redoes only those.
    if (retry && nkeys_out[i] != -1) return;
    SideSmem lay{kcap, nplanes};
    uint32_t *base = se_smem + warp * lay.words();
    uint32_t *cnt = base;
    uint32_t *sel = cnt + (size_t)2 * kcap * nplanes * 32;
    uint16_t *stage = reinterpret_cast<uint16_t *>(sel + (size_t)kcap * 16);
    const Geo g = make_geo(h, w);
    const size_t cells = (size_t)h * w;
    uint32_t A[16], C[16];
    load_planes(board0 + i * cells, stage, h, w, vec, A);
    load_planes(board1 + i * cells, stage, h, w, vec, C);
    const uint64_t env = env_id0 + (uint64_t)i;
    const float thr = spawn_prob[i];
    bool ch;
    uint64_t counter = counters ? counters[i] : 0;
    const int catch_up = num_steps[i];
    for (int t = 0; t < catch_up; t++) counter = physics_planes(A, g, thr, seed, env, counter, ch);  // :135-136
    int nkeys = 0;
    bool overflow = false;
    for (int t = 0; t < num_samples && !overflow; t++) {  // :137-141, the inaction board draws first
        counter = physics_planes(A, g, thr, seed, env, counter, ch);
        counter = physics_planes(C, g, thr, seed, env, counter, ch);
        se_accumulate_planes(A, g, 0, kcap, nplanes, nkeys, overflow, cnt, sel);
        se_accumulate_planes(C, g, 1, kcap, nplanes, nkeys, overflow, cnt, sel);
    }
    if (overflow) {  // warp-uniform
        if (lane == 0) nkeys_out[i] = -1;
        return;
    }
    if (counters && lane == 0) counters[i] = counter;
    __syncwarp();
    // keys: ctype, with the destructible flag added back for life-like cells and spawners (:72-75)
    for (int k = lane; k < kmax; k += 32) {
        uint32_t key = 0;
        if (k < nkeys) {
            for (int b = 0; b < 16; b++) key |= (sel[k * 16 + b] & 1u) << b;
            uint32_t bt = key & ~(uint32_t)COLORS;
            if (bt == ALIVE || bt == (FROZEN | SPAWNING)) key |= DESTRUCTIBLE;
        }
        keys[(size_t)i * kmax + k] = (uint16_t)key;
    }
    if (lane == 0) nkeys_out[i] = nkeys;
    // counters -> uint16 maps [2][kmax][H][W] (same transposition as planes -> cells)
    const bool vec_out = fast_vec_ok(h, w, counts, counts);
    for (int br = 0; br < 2; br++)
        for (int k = 0; k < nkeys; k++) {
            uint32_t V[16];
#pragma unroll
            for (int p = 0; p < 16; p++)
                V[p] = (p < nplanes) ? cnt[((size_t)(br * kcap + k) * nplanes + p) * 32 + lane] : 0u;
            store_planes(counts + (((size_t)i * 2 + br) * kmax + k) * cells, stage, h, w, vec_out, V);
        }
}

// returns false when the counters do not fit in shared memory (caller takes the generic kernel)
inline bool launch_side_effect_fast(const uint16_t *b0, const uint16_t *b1, int n, int h, int w,
                                    const int32_t *num_steps, int num_samples, const float *p,
                                    uint64_t seed, uint64_t env0, uint64_t *ctr, int kmax,
                                    uint16_t *keys, uint16_t *counts, int32_t *nkeys, size_t smem_limit,
                                    cudaStream_t st) {
    int nplanes = 1;
    while ((1 << nplanes) <= num_samples) nplanes++;
    const bool vec = fast_vec_ok(h, w, b0, b1);
    auto pass = [&](int kcap, int retry) -> bool {
        SideSmem lay{kcap, nplanes};
        size_t per_warp = lay.words() * 4;
        if (per_warp % 16) return false;  // keep each warp's region 16-byte aligned
        int warps = 4;
        while (warps > 1 && warps * per_warp > smem_limit) warps >>= 1;
        if (warps * per_warp > smem_limit) return false;
        size_t bytes = warps * per_warp;
        if (bytes > 48 * 1024)
            cudaFuncSetAttribute(side_effect_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        int grid = (n + warps - 1) / warps;
        side_effect_fast_kernel<<<grid, warps * 32, bytes, st>>>(b0, b1, n, h, w, num_steps, num_samples, p, seed,
                                                              env0, ctr, kmax, kcap, retry, nplanes, keys,
                                                              counts, nkeys, vec);
        return true;
    };
    const int kfirst = 8;
    if (kmax <= kfirst) return pass(kmax, 0);
    // make sure the full-capacity pass can run before committing to the two-pass scheme
    {
        SideSmem lay{kmax, nplanes};
        if ((lay.words() * 4) % 16 || lay.words() * 4 > smem_limit) return false;
    }
    return pass(kfirst, 0) && pass(kmax, 1);
}

}  // namespace slb
