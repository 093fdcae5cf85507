// slb_fast_side.cuh — side_effect_score rollouts on bit-planes (side_effects.py:59-141): one warp
// per finished episode; both boards (inaction b0, action b1) live in registers; the per-cell-type
// count maps are bit-sliced vertical counters (one weighted add per matched key and run of identical
// board states), transposed to the uint16 [2,K,H,W] output layout once at the end.
//
// Counter storage ("split" mode, whenever a key's output slice is large enough): bits 0-3 of every
// counter plus one PENDING plane worth 16 live in shared memory; bits >= 4 live in the episode's own
// slice of the output array (`counts[i][branch][key]` is H*W uint16 and only receives its final
// contents at the very end).  An add of weight < 16 is then five shared loads, a fixed carry chain and
// five stores — no votes, no data-dependent loop — with the carry out of bit 3 parked in the pending
// plane; the pending planes of a branch are folded into the high bits before the weights added since
// the last fold could exceed 15 (so a cell's nibble overflows at most once in between).  A warp needs
// 2*kcap*5*128 B of shared memory instead of 2*kcap*nplanes*128 B: twice the resident episodes per SM.
//
// A board state is counted once per RUN of identical states: physics_compute says what would flip,
// and only when something does is the old state added with the run length as its weight.  A board
// that reached a draw-free fixed point (`quiet`) is not stepped again; when both boards have, the
// remaining samples are added to the two run lengths and the loop ends.
#pragma once
#include "slb_planes_env.cuh"

namespace slb {

constexpr int SE_FAST_MAX_KEYS = 64;
constexpr int SE_SPLIT_LO = 5;  // shared planes in split mode: bits 0-3 and the pending-16 plane

struct SideSmem {  // per-warp carve-up of dynamic shared memory (in 32-bit words)
    int kcap, lo;  // kcap: number of cell types the counters have room for; lo: planes in shared memory
    __host__ __device__ size_t words() const {
        return (size_t)2 * kcap * lo * 32  // counters [branch][key][plane][lane]
               + (size_t)kcap * 16         // per-key plane selectors (0 / ~0)
               + 32 * 34 / 2;              // uint16 staging area for load/store
    }
};

struct SideCtr {  // where one episode's counters live
    uint32_t *cnt, *sel;  // shared: low planes, selectors
    uint16_t *counts;     // the episode's [2][kmax][cells] output slice (high planes while running)
    int kcap, kmax, lo, nplanes;
    bool split;
    bool rare;  // a countable cell of the episode has a bit outside selector slots 0-7
    size_t cells;
    // high planes of (branch, key): the 4-byte aligned start of that key's output slice
    __device__ __forceinline__ uint32_t *hi(int br, int k) const {
        uintptr_t a = reinterpret_cast<uintptr_t>(counts + ((size_t)br * kmax + k) * cells);
        return reinterpret_cast<uint32_t *>((a + 3) & ~(uintptr_t)3) + (threadIdx.x & 31);
    }
    __device__ __forceinline__ uint32_t *low(int br, int k) const {
        return cnt + ((size_t)(br * kcap + k) * lo) * 32 + (threadIdx.x & 31);
    }
};

// split mode, weight < 16, pending plane known to have room
__device__ __forceinline__ void se_add_fast(uint32_t *c, uint32_t m, uint32_t w) {
    const uint32_t v0 = c[0], v1 = c[32], v2 = c[64], v3 = c[96], v4 = c[128];
    if (w == 1u) {  // the usual case on a board that changes every step
        uint32_t cy = v0 & m;
        c[0] = v0 ^ m;
        c[32] = v1 ^ cy; cy &= v1;
        c[64] = v2 ^ cy; cy &= v2;
        c[96] = v3 ^ cy; cy &= v3;
        c[128] = v4 | cy;
        return;
    }
    const uint32_t x0 = (w & 1u) ? m : 0u, x1 = (w & 2u) ? m : 0u, x2 = (w & 4u) ? m : 0u, x3 = (w & 8u) ? m : 0u;
    uint32_t cy = v0 & x0;
    c[0] = v0 ^ x0;
    c[32] = v1 ^ x1 ^ cy; cy = maj3(v1, x1, cy);
    c[64] = v2 ^ x2 ^ cy; cy = maj3(v2, x2, cy);
    c[96] = v3 ^ x3 ^ cy; cy = maj3(v3, x3, cy);
    c[128] = v4 | cy;
}

// Selector slot -> cell bit.  Slots 0-7 hold the bits benchmark boards use after the masking of
// _add_cell_distribution (alive, frozen, spawning, colours, pushable, pullable), slots 8-13 the rest;
// the agent and destructible bits (slots 14, 15) never take part in a match.
constexpr uint64_t SE_SLOT_BITS = 0x31EDC865F2BA9740ull;  // nibble s = cell bit of slot s
__host__ __device__ __forceinline__ int se_slot_bit(int s) { return (int)((SE_SLOT_BITS >> (4 * s)) & 15u); }

// ripple add of weight * m over `np` planes at c (stride 32 words), starting with carry `cy` and the
// weight bits `wb`; stops once the weight's bits and the carries ran out.  Returns the carry out.
__device__ __forceinline__ uint32_t se_add_ripple(uint32_t *c, int np, uint32_t m, uint32_t &wb, uint32_t cy) {
    for (int p = 0; p < np; p++, wb >>= 1) {
        if (wb == 0 && !__any_sync(FULL, cy != 0)) break;
        const uint32_t x = (wb & 1u) ? m : 0u, v = c[p * 32];
        c[p * 32] = v ^ x ^ cy;
        cy = maj3(v, x, cy);
    }
    return cy;
}

// split mode: fold the pending planes of one branch into the high bits
__device__ __forceinline__ void se_fold_pending(const SideCtr &sc, int br, int nkeys) {
    const int nhi = sc.nplanes - 4;
    for (int k = 0; k < nkeys; k++) {
        uint32_t *lw = sc.low(br, k);
        uint32_t cy = lw[128];
        if (!__any_sync(FULL, cy != 0)) continue;
        lw[128] = 0;
        uint32_t *hi = sc.hi(br, k);
        uint32_t v[8];
#pragma unroll
        for (int p = 0; p < 8; p++) v[p] = (p < nhi) ? hi[p * 32] : 0u;  // all loads in flight together
#pragma unroll
        for (int p = 0; p < 8; p++)
            if (p < nhi) {
                hi[p * 32] = v[p] ^ cy;
                cy &= v[p];
            }
        uint32_t none = 0;
        if (nhi > 8) se_add_ripple(hi + 8 * 32, nhi - 8, 0u, none, cy);
    }
}

// _add_cell_distribution (side_effects.py:59-86) for one board, `weight` times (weight >= 1, warp-uniform).
// wsum: weights added to this branch since its pending planes were last folded.
__device__ __forceinline__ void se_accumulate_planes(const uint32_t P[16], int branch, uint32_t weight,
                                                     const SideCtr &sc, int &nkeys, bool &overflow,
                                                     uint32_t &wsum) {
    const int lane = threadIdx.x & 31;
    // unchanging = (cell & (frozen|destructible|movable)) == frozen  -> zeroed (:61-62);
    // the destructible bit is dropped from every cell (:62); agent cells and zeros are skipped (:68-70).
    // On the cells of `todo` the masked value equals the cell itself (minus destructible, agent clear),
    // so keys are matched against the planes P directly.
    const uint32_t keepm = ~(P[4] & ~P[3] & ~P[2] & ~P[15]);
    const uint32_t nz = or3(or3(P[0], P[2], P[4]), or3(P[5], P[6], P[7]), or3(P[8], P[9], P[10])) |
                        or3(or3(P[11], P[12], P[13]), P[14], P[15]);
    uint32_t todo = nz & keepm & ~P[1];  // cells still to be attributed to a key
    if (!__any_sync(FULL, todo != 0)) return;
    const bool fast = sc.split && weight < 16;
    if (sc.split) {
        if (!fast || wsum + weight > 15) {
            se_fold_pending(sc, branch, nkeys);
            wsum = 0;
        }
        if (fast) wsum += weight;
    }
    for (int k = 0;; k++) {
        if (k == nkeys) {
            // discover the next cell type: first unattributed cell in raster order
            unsigned lanes = __ballot_sync(FULL, todo != 0);
            if (!lanes) break;
            if (nkeys == sc.kcap) { overflow = true; break; }
            const int src = __ffs(lanes) - 1;
            const int x = __ffs(__shfl_sync(FULL, todo, src)) - 1;
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 16; b++)
                if (b != 1 && b != 3) v |= ((P[b] >> x) & 1u) << b;
            v = __shfl_sync(FULL, v, src);
            if (lane < 16) sc.sel[k * 16 + lane] = ((v >> se_slot_bit(lane)) & 1u) ? 0xffffffffu : 0u;
            for (int br = 0; br < 2; br++) {  // zero both branches' counters of the new key
                uint32_t *lw = sc.low(br, k);
                for (int p = 0; p < sc.lo; p++) lw[p * 32] = 0;
                if (sc.split) {
                    uint32_t *hi = sc.hi(br, k);
                    for (int p = 0; p < sc.nplanes - 4; p++) hi[p * 32] = 0;
                }
            }
            nkeys++;
            __syncwarp();
        }
        const uint4 *s4 = reinterpret_cast<const uint4 *>(sc.sel + k * 16);
        const uint4 sa = s4[0], sb = s4[1];
        uint32_t diff = (P[0] ^ sa.x) | (P[4] ^ sa.y) | (P[7] ^ sa.z) | (P[9] ^ sa.w) |
                        (P[10] ^ sb.x) | (P[11] ^ sb.y) | (P[2] ^ sb.z) | (P[15] ^ sb.w);
        if (sc.rare) {  // some cell of the episode carries one of the other bits
            const uint4 sc4 = s4[2], sd = s4[3];
            diff |= (P[5] ^ sc4.x) | (P[6] ^ sc4.y) | (P[8] ^ sc4.z) | (P[12] ^ sc4.w) | (P[13] ^ sd.x) | (P[14] ^ sd.y);
        }
        const uint32_t m = todo & ~diff;
        todo &= diff;
        uint32_t *c = sc.low(branch, k);
        if (fast) {
            se_add_fast(c, m, weight);
        } else {
            uint32_t wb = weight;
            if (sc.split) {  // pending planes were folded above: nibble here, the rest in the output slice
                const uint32_t cy = se_add_ripple(c, 4, m, wb, 0u);
                if (wb != 0 || __any_sync(FULL, cy != 0)) se_add_ripple(sc.hi(branch, k), sc.nplanes - 4, m, wb, cy);
            } else {
                se_add_ripple(c, sc.lo, m, wb, 0u);
            }
        }
        if (!__any_sync(FULL, todo != 0)) break;
    }
}

__global__ void __launch_bounds__(128, 4)
side_effect_fast_kernel(const uint16_t *__restrict__ board0, const uint16_t *__restrict__ board1, int n,
                        int h, int w, const int32_t *__restrict__ num_steps, int num_samples,
                        const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                        uint64_t *counters, int kmax, int kcap, int retry, int nplanes, int lo,
                        uint16_t *keys, uint16_t *counts, int32_t *nkeys_out, bool vec) {
    extern __shared__ __align__(16) uint32_t se_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + warp;
    if (i >= n) return;
    // kcap < kmax is the first pass: most episodes see a handful of cell types, and counters for a
    // handful fit 4 warps per CTA and 4 CTAs per SM.  An episode that needs more reports -1 and leaves
    // its draw counter alone; the retry pass (full capacity, few warps per SM) redoes only those.
    if (retry && nkeys_out[i] != -1) return;
    SideSmem lay{kcap, lo};
    uint32_t *base = se_smem + warp * lay.words();
    const Geo g = make_geo(h, w);
    const size_t cells = (size_t)h * w;
    SideCtr sc;
    sc.cnt = base;
    sc.sel = base + (size_t)2 * kcap * lo * 32;
    sc.counts = counts + (size_t)i * 2 * kmax * cells;
    sc.kcap = kcap; sc.kmax = kmax; sc.lo = lo; sc.nplanes = nplanes; sc.cells = cells;
    sc.split = lo < nplanes;
    uint16_t *stage = reinterpret_cast<uint16_t *>(sc.sel + (size_t)kcap * 16);
    uint32_t A[16], C[16];
    load_planes(board0 + i * cells, stage, h, w, vec, A);
    load_planes(board1 + i * cells, stage, h, w, vec, C);
    {   // cells that _add_cell_distribution zeroes are frozen and stay as they are for good; the step
        // creates alive / destructible / colour bits only: the other bits of countable cells never grow
        // (the player itself — frozen, preserving, inhibiting — is skipped by the count and never changes)
        const uint32_t ra = (A[5] | A[6] | A[8] | A[12] | A[13] | A[14]) & ~(A[4] & ~A[3] & ~A[2] & ~A[15]) & ~(A[1] & A[4]);
        const uint32_t rc = (C[5] | C[6] | C[8] | C[12] | C[13] | C[14]) & ~(C[4] & ~C[3] & ~C[2] & ~C[15]) & ~(C[1] & C[4]);
        sc.rare = __any_sync(FULL, (ra | rc) != 0);
    }
    const uint64_t env = env_id0 + (uint64_t)i;
    const float thr = spawn_prob[i];
    uint64_t counter = counters ? counters[i] : 0;
    const int catch_up = num_steps[i];
    bool fixA = false, fixC = false;  // the board is at a draw-free fixed point (warp-uniform)
    PhysDelta dl;
    for (int t = 0; t < catch_up && !fixA; t++) {  // :135-136
        counter = physics_compute(A, g, thr, seed, env, counter, dl, fixA);
        if (__any_sync(FULL, dl.flip != 0)) physics_apply(A, dl);
    }
    int nkeys = 0;
    bool overflow = false;
    uint32_t runA = 0, runC = 0;  // samples the current state of the board still has to be counted for
    uint32_t wsumA = 0, wsumC = 0;
    // :137-141, the inaction board draws first; iteration t == num_samples only counts the last runs
    for (int t = 0; t <= num_samples && !overflow; t++) {
        const bool last = t == num_samples;
        bool flip = false;
        if (!last && !fixA) {
            counter = physics_compute(A, g, thr, seed, env, counter, dl, fixA);
            flip = __any_sync(FULL, dl.flip != 0);
        }
        if ((flip || last) && runA) {
            se_accumulate_planes(A, 0, runA, sc, nkeys, overflow, wsumA);
            runA = 0;
        }
        if (flip) physics_apply(A, dl);
        flip = false;
        if (!last && !fixC) {
            counter = physics_compute(C, g, thr, seed, env, counter, dl, fixC);
            flip = __any_sync(FULL, dl.flip != 0);
        }
        if ((flip || last) && runC && !overflow) {
            se_accumulate_planes(C, 1, runC, sc, nkeys, overflow, wsumC);
            runC = 0;
        }
        if (flip) physics_apply(C, dl);
        if (!last) {
            runA++;
            runC++;
            if (fixA && fixC) {  // nothing changes any more and no draw is consumed
                runA += (uint32_t)(num_samples - 1 - t);
                runC += (uint32_t)(num_samples - 1 - t);
                t = num_samples - 1;
            }
        }
    }
    if (overflow) {  // warp-uniform
        if (lane == 0) nkeys_out[i] = -1;
        return;
    }
    if (sc.split) {
        se_fold_pending(sc, 0, nkeys);
        se_fold_pending(sc, 1, nkeys);
    }
    if (counters && lane == 0) counters[i] = counter;
    __syncwarp();
    // keys: ctype, with the destructible flag added back for life-like cells and spawners (:72-75)
    for (int k = lane; k < kmax; k += 32) {
        uint32_t key = 0;
        if (k < nkeys) {
            for (int b = 0; b < 16; b++) key |= (sc.sel[k * 16 + b] & 1u) << se_slot_bit(b);
            uint32_t bt = key & ~(uint32_t)COLORS;
            if (bt == ALIVE || bt == (FROZEN | SPAWNING)) key |= DESTRUCTIBLE;
        }
        keys[(size_t)i * kmax + k] = (uint16_t)key;
    }
    if (lane == 0) nkeys_out[i] = nkeys;
    // counters -> uint16 maps [2][kmax][H][W] (same transposition as planes -> cells); every lane reads
    // all of a key's planes (the high ones from the output slice itself) before store_planes' first
    // __syncwarp, and only after it is the slice overwritten with the final map
    const bool vec_out = fast_vec_ok(h, w, counts, counts);
    const int nlow = sc.split ? 4 : lo;
    for (int br = 0; br < 2; br++)
        for (int k = 0; k < nkeys; k++) {
            uint32_t V[16];
            const uint32_t *lw = sc.low(br, k), *hi = sc.hi(br, k);
#pragma unroll
            for (int p = 0; p < 16; p++)
                V[p] = (p < nlow) ? lw[p * 32] : (p < nplanes) ? hi[(p - nlow) * 32] : 0u;
            store_planes(sc.counts + ((size_t)br * kmax + k) * cells, stage, h, w, vec_out, V);
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
    // the high counter planes fit in a key's own output slice (H*W uint16) when it is large enough
    int lo = nplanes;
    if (nplanes > SE_SPLIT_LO && (size_t)h * w * 2 >= (size_t)(nplanes - 4) * 128 + 4) lo = SE_SPLIT_LO;
    const bool vec = fast_vec_ok(h, w, b0, b1);
    auto pass = [&](int kcap, int retry) -> bool {
        SideSmem lay{kcap, lo};
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
                                                              env0, ctr, kmax, kcap, retry, nplanes, lo, keys,
                                                              counts, nkeys, vec);
        return true;
    };
    const int kfirst = 8;
    if (kmax <= kfirst) return pass(kmax, 0);
    // make sure the full-capacity pass can run before committing to the two-pass scheme
    {
        SideSmem lay{kmax, lo};
        if ((lay.words() * 4) % 16 || lay.words() * 4 > smem_limit) return false;
    }
    return pass(kfirst, 0) && pass(kmax, 1);
}

}  // namespace slb
