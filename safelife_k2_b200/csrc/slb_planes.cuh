// slb_planes.cuh — bit-sliced SafeLife on one warp per board (H <= 32 rows across lanes, W <= 32
// cells across the bits of one 32-bit word, 16 bit-planes per lane in registers).
//
// A board row r lives in lane r as P[b] (b = cell bit 0..15): bit c of P[b] = bit b of cell (r, c).
// Bits >= W of every plane are kept zero; lanes >= H hold zeros.  Vertical neighbours come from
// warp shuffles (toroidal: lane (r +- 1) mod H), horizontal ones from rotations within W bits.
// Semantics follow the reference exactly; parity is tested bit for bit.
#pragma once
#include "slb_common.cuh"

namespace slb {

constexpr unsigned FULL = 0xffffffffu;

struct Geo {
    int h, w;
    int up, dn;      // source lanes of the rows above / below this lane (toroidal)
    uint32_t wmask;  // low W bits
    uint32_t hib;    // 1 << (W - 1): multiplying by it moves column 0 to column W - 1
    bool row;        // lane < h
};

__device__ __forceinline__ Geo make_geo(int h, int w) {
    Geo g;
    int lane = threadIdx.x & 31;
    g.h = h; g.w = w;
    g.row = lane < h;
    g.up = g.row ? (lane == 0 ? h - 1 : lane - 1) : lane;
    g.dn = g.row ? (lane == h - 1 ? 0 : lane + 1) : lane;
    g.wmask = (w >= 32) ? 0xffffffffu : ((1u << w) - 1u);
    g.hib = 1u << (w - 1);
    return g;
}

// value of the LEFT neighbour (cell c-1) at bit c / of the RIGHT neighbour (cell c+1) at bit c
__device__ __forceinline__ uint32_t from_left(uint32_t x, const Geo &g) {
    return ((x << 1) | (x >> (g.w - 1))) & g.wmask;
}
__device__ __forceinline__ uint32_t from_right(uint32_t x, const Geo &g) {
    return ((x >> 1) | (x << (g.w - 1))) & g.wmask;
}

// The same neighbours WITHOUT the final mask: columns 0 .. W-1 are right, bits >= W hold rubbish.  For
// values that are only combined column by column and ANDed with a clean plane before they are stored or
// rotated again (every summary of the step below): a shift plus a multiply-add instead of two shifts and
// a masked OR, and the multiply-add issues on the FMA pipe, which the bit-sliced code leaves idle.
__device__ __forceinline__ uint32_t from_left_raw(uint32_t x, const Geo &g) {
    return x * 2u + (x >> (g.w - 1));          // bit 0 of 2x is clear: the sum is an OR
}
__device__ __forceinline__ uint32_t from_right_raw(uint32_t x, const Geo &g) {
    return x * g.hib + (x >> 1);               // x * hib: column 0 -> column W-1, the rest above it
}

__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) { return a ^ b ^ c; }
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) {
    return (a & b) | (c & (a | b));
}
__device__ __forceinline__ uint32_t or3(uint32_t a, uint32_t b, uint32_t c) { return a | b | c; }

// ---------------------------------------------------------------------------------------------
// 32 cells x 16 bits  <->  16 planes x 32 bits, in registers.
// x[j] (j < 16) = cell(2j) | cell(2j+1) << 16 (the row as it lies in memory, zero padded).
// Step 1 (PRMT): y[j] = cell(j) | cell(j+16) << 16.  Step 2: a 16x16 bit transpose carried out on
// both 16-bit halves at once (block-swap stages 8, 4, 2, 1; stage 8 is a byte permute).
// The transpose is an involution, so the inverse runs the same stages and then undoes step 1.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void transpose16_halves(uint32_t y[16]) {
#pragma unroll
    for (int i = 0; i < 8; i++) {  // stage 8: rows i <-> i+8, swap bytes
        uint32_t a = y[i], b = y[i + 8];
        y[i] = __byte_perm(a, b, 0x6240);      // a0 b0 a2 b2
        y[i + 8] = __byte_perm(a, b, 0x7351);  // a1 b1 a3 b3
    }
#pragma unroll
    for (int s = 4; s >= 1; s >>= 1) {
        const uint32_t m = (s == 4) ? 0x0f0f0f0fu : (s == 2) ? 0x33333333u : 0x55555555u;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (i & s) continue;
            uint32_t a = y[i], b = y[i + s];
            y[i] = (a & m) | ((b << s) & ~m);
            y[i + s] = ((a >> s) & m) | (b & ~m);
        }
    }
}

__device__ __forceinline__ void cells_to_planes(const uint32_t x[16], uint32_t P[16]) {
#pragma unroll
    for (int j = 0; j < 16; j++)
        P[j] = __byte_perm(x[j >> 1], x[8 + (j >> 1)], (j & 1) ? 0x7632 : 0x5410);
    transpose16_halves(P);
}

__device__ __forceinline__ void planes_to_cells(const uint32_t P[16], uint32_t x[16]) {
    uint32_t y[16];
#pragma unroll
    for (int j = 0; j < 16; j++) y[j] = P[j];
    transpose16_halves(y);
#pragma unroll
    for (int j = 0; j < 8; j++) {
        // y[2j] = cell(2j) | cell(2j+16)<<16, y[2j+1] = cell(2j+1) | cell(2j+17)<<16
        x[j] = __byte_perm(y[2 * j], y[2 * j + 1], 0x5410);      // cell(2j) | cell(2j+1)<<16
        x[8 + j] = __byte_perm(y[2 * j], y[2 * j + 1], 0x7632);  // cell(2j+16) | cell(2j+17)<<16
    }
}

// ---------------------------------------------------------------------------------------------
// Staging between the uint16 [H,W] layout in global memory and per-lane planes.  `sm` is this
// warp's shared staging area of H * S uint16 (S = padded row stride, see row_stride()).
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ int row_stride(int w) {
    int s = (w + 1) & ~1;           // even
    if (((s >> 1) & 1) == 0) s += 2;  // S/2 odd -> conflict-free 32-bit row reads across lanes
    return s;
}

// the planes of the board whose cells lie in shared memory at `sm` (row stride row_stride(w)); lane = row
template <int HW = 0>
__device__ __forceinline__ void planes_from_image(const uint16_t *sm, int h_, int w_, uint32_t P[16]) {
    const int h = HW ? HW : h_, w = HW ? HW : w_;
    const int lane = threadIdx.x & 31, S = row_stride(w);
    uint32_t x[16];
    const uint32_t *row = reinterpret_cast<const uint32_t *>(sm) + lane * (S >> 1);
#pragma unroll
    for (int j = 0; j < 16; j++) {
        uint32_t v = 0;
        if (lane < h && 2 * j < w) {
            v = row[j];
            if (2 * j + 1 >= w) v &= 0xffffu;
        }
        x[j] = v;
    }
    cells_to_planes(x, P);
}

// the inverse: the planes written as cells into the shared image (callers sync the warp around it)
template <int HW = 0>
__device__ __forceinline__ void image_from_planes(uint16_t *sm, int h_, int w_, const uint32_t P[16]) {
    const int h = HW ? HW : h_, w = HW ? HW : w_;
    const int lane = threadIdx.x & 31, S = row_stride(w);
    uint32_t x[16];
    planes_to_cells(P, x);
    uint32_t *row = reinterpret_cast<uint32_t *>(sm) + lane * (S >> 1);
#pragma unroll
    for (int j = 0; j < 16; j++)
        if (lane < h && 2 * j < w) row[j] = x[j];  // a padding half-word may be written: never read back
}

// vec: all boards are 8-byte aligned and S == w (then the smem image equals the memory image)
// HW > 0: the board is HW x HW, known at compile time (row predicates and strides fold away).
template <int HW = 0>
__device__ __forceinline__ void load_planes(const uint16_t *__restrict__ src, uint16_t *sm, int h_,
                                            int w_, bool vec, uint32_t P[16]) {
    const int h = HW ? HW : h_, w = HW ? HW : w_;
    const int lane = threadIdx.x & 31, cells = h * w, S = row_stride(w);
    __syncwarp();
    if (vec) {
        const uint2 *s2 = reinterpret_cast<const uint2 *>(src);
        uint2 *d2 = reinterpret_cast<uint2 *>(sm);
        for (int j = lane; j < (cells >> 2); j += 32) d2[j] = s2[j];
    } else {
        for (int e = lane; e < cells; e += 32) {
            int r = e / w, c = e - r * w;
            sm[r * S + c] = src[e];
        }
    }
    __syncwarp();
    planes_from_image<HW>(sm, h, w, P);
}

template <int HW = 0>
__device__ __forceinline__ void store_planes(uint16_t *__restrict__ dst, uint16_t *sm, int h_, int w_,
                                             bool vec, const uint32_t P[16]) {
    const int h = HW ? HW : h_, w = HW ? HW : w_;
    const int lane = threadIdx.x & 31, cells = h * w, S = row_stride(w);
    uint32_t x[16];
    planes_to_cells(P, x);
    __syncwarp();
    uint32_t *row = reinterpret_cast<uint32_t *>(sm) + lane * (S >> 1);
#pragma unroll
    for (int j = 0; j < 16; j++)
        if (lane < h && 2 * j < w) row[j] = x[j];  // a padding half-word may be written: never read back
    __syncwarp();
    if (vec) {
        const uint2 *s2 = reinterpret_cast<const uint2 *>(sm);
        uint2 *d2 = reinterpret_cast<uint2 *>(dst);
        for (int j = lane; j < (cells >> 2); j += 32) d2[j] = s2[j];
    } else {
        for (int e = lane; e < cells; e += 32) {
            int r = e / w, c = e - r * w;
            dst[e] = sm[r * S + c];
        }
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// One advance_board step on the planes (advance_board.c:34-120, closed form of Appendix A).
// Returns the env's updated draw counter; `changed` (warp-uniform) says whether any cell changed.
//
// The evaluation is lazy but exact: the alive count n of every 3x3 is always computed; the flag
// summaries (preserving / inhibiting / spawning), the draws and the colour inheritance are only
// evaluated when some cell of this board can actually need them (all tests are warp votes, i.e.
// per board).  A board on which nothing dies, nothing can be born and no spawner exists costs the
// count and three votes.
// ---------------------------------------------------------------------------------------------
// (box_or / box_two: columns 0 .. W-1 only — bits >= W of the result are rubbish, see from_left_raw)
__device__ __forceinline__ uint32_t box_or(uint32_t x, const Geo &g) {  // OR over the 3x3, centre included
    const uint32_t v = or3(x, __shfl_sync(FULL, x, g.up), __shfl_sync(FULL, x, g.dn));
    return or3(from_left_raw(v, g), v, from_right_raw(v, g));
}

// ">= 2 of the alive cells in the 3x3 carry the flag" (pairwise AND accumulation of
// advance_board.c:18,28 in closed form); src = alive & flag
__device__ __forceinline__ uint32_t box_two(uint32_t src, const Geo &g) {
    const uint32_t u = __shfl_sync(FULL, src, g.up), d = __shfl_sync(FULL, src, g.dn);
    const uint32_t f0 = xor3(src, u, d), f1 = maj3(src, u, d);  // column count, 2 bits
    const uint32_t hi = or3(from_left_raw(f1, g), f1, from_right_raw(f1, g));
    return hi | maj3(from_left_raw(f0, g), f0, from_right_raw(f0, g));
}

// The step is split into "compute what flips" and "apply it" so that a caller can still use the old
// board between the two (the side-effect kernel counts a board once per run of identical states).
struct PhysDelta {
    uint32_t flip;               // cells that die or are born
    uint32_t born, nD, nR, nG, nB;  // births and the destructible / colour bits they inherit
};

// `quiet` (warp-uniform): nothing can die or be born and there is no spawner — a fixed point that
// consumes no draws, now and at every later step.
__device__ __forceinline__ uint64_t physics_compute(const uint32_t P[16], const Geo &g, float spawn_prob,
                                                    uint64_t seed, uint64_t env, uint64_t counter,
                                                    PhysDelta &dl, bool &quiet) {
    dl.flip = dl.born = dl.nD = dl.nR = dl.nG = dl.nB = 0;
    quiet = false;
    const uint32_t a = P[0], fz = P[4], sp = P[7];
    // ---- alive count of the 3x3 (centre included): n = s0 + 2 n1 + 4 n2 (+ 8), n <= 9 ----
    uint32_t is3, is4;
    {
        const uint32_t u = __shfl_sync(FULL, a, g.up), d = __shfl_sync(FULL, a, g.dn);
        const uint32_t v0 = xor3(a, u, d), v1 = maj3(a, u, d);
        const uint32_t l0 = from_left_raw(v0, g), r0 = from_right_raw(v0, g);
        const uint32_t l1 = from_left_raw(v1, g), r1 = from_right_raw(v1, g);
        const uint32_t s0 = xor3(l0, v0, r0), c0 = maj3(l0, v0, r0);
        const uint32_t s1 = xor3(l1, v1, r1), c1 = maj3(l1, v1, r1);
        const uint32_t n1 = c0 ^ s1, k = c0 & s1, n2 = c1 ^ k;
        is3 = s0 & n1 & ~n2;
        is4 = ~s0 & ~n1 & n2;  // n2 set implies the 8s bit (c1 & k) is clear
    }
    const uint32_t die0 = a & ~(fz | is3 | is4);       // dies unless preserved
    const uint32_t open0 = ~(a | fz) & g.wmask;        // dead and not frozen
    const uint32_t birth0 = open0 & is3;               // born unless inhibited
    if (!__any_sync(FULL, (die0 | birth0 | sp) != 0)) {  // fixed point, no spawner
        quiet = true;
        return counter;
    }
    // ---- flag summaries: taken from every cell of the 3x3, alive or not ----
    uint32_t PR = 0, IH = 0;
    if (__any_sync(FULL, (P[5] | P[6]) != 0)) {
        PR = box_or(P[5], g);
        IH = box_or(P[6], g);
    }
    const uint32_t die = die0 & ~PR;
    const uint32_t open = open0 & ~IH;
    const uint32_t birth = open & is3;
    const bool any_sp = __any_sync(FULL, sp != 0);
    uint32_t spawn = 0;
    if (any_sp) {
        const uint32_t elig = open & ~is3 & box_or(sp, g);  // the cells that consume a draw (:110)
        if (__any_sync(FULL, elig != 0)) {
            // raster-order rank of every eligible cell: exclusive prefix over rows, then in the row.
            // The draw of rank r depends on r alone, so lane j evaluates the draws of ranks j, j + 32, ...
            // (one Philox evaluation per 32 draws instead of one per eligible cell of the busiest row);
            // a vote packs the 32 outcomes and every row picks the bits of its own ranks.
            const uint64_t thr = spawn_threshold(spawn_prob);
            int cnt = __popc(elig), incl = cnt;
            const int lane = threadIdx.x & 31;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += t;
            }
            const int total = __shfl_sync(FULL, incl, 31);
            for (int c0 = 0; c0 < total; c0 += 32) {  // warp-uniform
                const bool hit = (c0 + lane < total) && stream_u53(seed, env, counter + (uint64_t)(c0 + lane)) < thr;
                const uint32_t outcomes = __ballot_sync(FULL, hit);
                uint32_t e = elig;
                int r = incl - cnt - c0;  // rank of this row's first eligible cell, relative to the chunk
                while (e) {
                    const uint32_t bit = e & (0u - e);
                    e ^= bit;
                    if ((unsigned)r < 32u && ((outcomes >> r) & 1u)) spawn |= bit;
                    r++;
                }
            }
            counter += (uint64_t)total;
        }
    }
    const uint32_t born = birth | spawn;
    uint32_t nD = 0, nR = 0, nG = 0, nB = 0;
    if (__any_sync(FULL, born != 0)) {
        // inherited bits: present in >= 2 alive cells of the 3x3; D2 = destructible | exit (:44-46)
        nD = (birth & box_two(a & (P[3] | P[8]), g)) | spawn;  // spawned cells are always destructible
        uint32_t tR = box_two(a & P[9], g), tG = box_two(a & P[10], g), tB = box_two(a & P[11], g);
        if (any_sp) {  // colour of ANY spawning cell in the 3x3, alive or not (:19)
            tR |= box_or(sp & P[9], g);
            tG |= box_or(sp & P[10], g);
            tB |= box_or(sp & P[11], g);
        }
        nR = born & tR; nG = born & tG; nB = born & tB;
    }
    dl.flip = die | born;
    dl.born = born; dl.nD = nD; dl.nR = nR; dl.nG = nG; dl.nB = nB;
    return counter;
}

__device__ __forceinline__ void physics_apply(uint32_t P[16], const PhysDelta &dl) {
    const uint32_t keep = ~dl.flip;
#pragma unroll
    for (int b = 0; b < 16; b++) P[b] &= keep;  // killed cells lose everything; births discard residue
    P[0] |= dl.born;
    P[3] |= dl.nD;
    P[9] |= dl.nR;
    P[10] |= dl.nG;
    P[11] |= dl.nB;
}

__device__ __forceinline__ uint64_t physics_planes(uint32_t P[16], const Geo &g, float spawn_prob,
                                                   uint64_t seed, uint64_t env, uint64_t counter,
                                                   bool &changed) {
    PhysDelta dl;
    bool quiet;
    counter = physics_compute(P, g, spawn_prob, seed, env, counter, dl, quiet);
    changed = !quiet && __any_sync(FULL, dl.flip != 0);
    if (changed) physics_apply(P, dl);
    return counter;
}

}  // namespace slb
