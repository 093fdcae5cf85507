// slb_common.cuh — shared device helpers for libsafelife_b200 (sm_100a).
//
// Cell bit layout: safelife/safelife_game.py:74-98 and speedups_src/constants.h:4-25 of the
// reference; re-declared here as the CUDA-side contract.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/safelife_b200.h"

namespace slb {

enum : uint32_t {
    ALIVE = 1u << 0,
    AGENT = 1u << 1,
    PUSHABLE = 1u << 2,
    DESTRUCTIBLE = 1u << 3,
    FROZEN = 1u << 4,
    PRESERVING = 1u << 5,
    INHIBITING = 1u << 6,
    SPAWNING = 1u << 7,
    EXIT = 1u << 8,
    RED = 1u << 9,
    GREEN = 1u << 10,
    BLUE = 1u << 11,
    PULLABLE = 1u << 15,
    COLORS = 7u << 9,
    MOVABLE = PUSHABLE | PULLABLE,
};

// ---------------------------------------------------------------------------------------------
// Counter-based spawn stream (replaces speedups_src/random.c): Philox4x32-10.
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t stream_u53(uint64_t seed, uint64_t env, uint64_t idx) {
    uint32_t c0 = (uint32_t)idx, c1 = (uint32_t)(idx >> 32);
    uint32_t c2 = (uint32_t)env, c3 = (uint32_t)(env >> 32);
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; r++) {
#ifdef __CUDA_ARCH__
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
#else
        uint64_t p0 = 0xD2511F53ull * c0, p1 = 0xCD9E8D57ull * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    // numpy random_sample construction: (a >> 5) * 2^26 + (b >> 6)
    return ((uint64_t)(c0 >> 5) << 26) | (uint64_t)(c1 >> 6);
}

// `random_float() < spawn_prob` (advance_board.c:110: double < (double)float) as an integer test:
// u53 / 2^53 < p  <=>  u53 < ceil(p * 2^53)  (p * 2^53 is exact in double).
__host__ __device__ __forceinline__ uint64_t spawn_threshold(float p) {
    double t = (double)p * 9007199254740992.0;
    if (!(t > 0.0)) return 0;  // also NaN
    if (t >= 9007199254740992.0) return 1ull << 53;
    uint64_t f = (uint64_t)t;
    return ((double)f < t) ? f + 1 : f;
}

// ---------------------------------------------------------------------------------------------
// point_table (safelife_game.py:554-565), rows = goal colour, cols = cell colour (k r g y b m c w)
// ---------------------------------------------------------------------------------------------
__device__ __constant__ int8_t c_point_table[64] = {
    0, -1, 0, 0, 0, 0, 0, 0,      //
    -3, 3, -3, 0, -3, 0, -3, -3,  //
    0, -3, 5, 0, 0, 0, 3, 0,      //
    -3, 0, 0, 3, 0, 0, 0, 0,      //
    3, -3, 3, 0, 5, 3, 3, 3,      //
    -3, 3, -3, 0, -3, 5, -3, -3,  //
    3, -3, 3, 0, 3, 0, 5, 3,      //
    0, -1, 0, 0, 0, 0, 0, 0,
};

// per-cell contributions to current_points / performance_ratio (safelife_game.py:590-631)
struct CellScore {
    int points;    // point_table[g][c] * alive
    int current;   // sign(point_table)[g][c] * (alive & not(frozen & immovable))
    int possible;  // max_c sign(point_table)[g][c] = goal colour not in {black, white}
};

__device__ __forceinline__ CellScore cell_score(uint32_t b, uint32_t g) {
    CellScore s;
    int gc = (g >> 9) & 7, cc = (b >> 9) & 7;
    int pt = (b & ALIVE) ? (int)c_point_table[gc * 8 + cc] : 0;
    s.points = pt;
    bool m = (b & ALIVE) && ((b & (FROZEN | MOVABLE)) != FROZEN);
    s.current = m ? ((pt > 0) - (pt < 0)) : 0;
    s.possible = (gc != 0 && gc != 7) ? 1 : 0;
    return s;
}

// can_exit (safelife_game.py:522-526): int >= float64 product
__device__ __forceinline__ int can_exit_rule(double min_perf, int completed, int possible) {
    if (min_perf < 0.0) return 1;
    return (double)completed >= min_perf * (double)possible;
}

__device__ __forceinline__ int warp_sum(int v) { return __reduce_add_sync(0xffffffffu, v); }

}  // namespace slb
