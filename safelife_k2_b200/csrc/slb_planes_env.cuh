// slb_planes_env.cuh — execute_action, scores and exit colours on bit-planes (one warp per env).
#pragma once
#include "slb_planes.cuh"

namespace slb {

__device__ __forceinline__ uint32_t bit_at(uint32_t plane, int lane_y, int x) {
    return (__shfl_sync(FULL, plane, lane_y) >> x) & 1u;
}

// ---------------------------------------------------------------------------------------------
// execute_action on planes (safelife_game.py:347-393, move_agent :308-345).
// Requires H >= 4 and W >= 4 so that the agent cell A, front F, two-ahead F2 and behind B are four
// distinct cells (smaller boards alias F2 and B and take the literal generic path instead).
// All reads see the pre-action board, which is what the reference reads too: the pull test on B
// happens after the moves, but no move writes B.  Warp-uniform control flow (one env per warp).
// Returns the action reward; updates ax, ay, orientation, game_over.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int action_planes(uint32_t P[16], const Geo &g, int &ax, int &ay,
                                             int &orientation, int &game_over, int action,
                                             int can_exit) {
    if (game_over || action < 1 || action > 8) return 0;
    const int lane = threadIdx.x & 31;
    const int dir = (action - 1) & 3;
    orientation = dir;
    const int fx = (dir == 1) - (dir == 3), fy = (dir == 2) - (dir == 0);
    const int h = g.h, w = g.w;
    const int xF = (ax + fx + w) % w, yF = (ay + fy + h) % h;
    uint32_t nz = 0;
#pragma unroll
    for (int b = 0; b < 16; b++) nz |= P[b];
    const uint32_t f_nz = bit_at(nz, yF, xF);
    if (action >= 5) {  // TOGGLE
        if (!f_nz) {
            // empty -> life | player colour (alive | destructible | colour of the agent cell)
            const uint32_t m = (lane == yF) ? (1u << xF) : 0u;
            P[0] |= m;
            P[3] |= m;
#pragma unroll
            for (int b = 9; b <= 11; b++)
                if (bit_at(P[b], ay, ax)) P[b] |= m;
        } else if (bit_at(P[3], yF, xF)) {  // destructible -> empty
            const uint32_t k = (lane == yF) ? ~(1u << xF) : 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 16; b++) P[b] &= k;
        }
        return 0;
    }
    // MOVE
    const int xB = (ax - fx + w) % w, yB = (ay - fy + h) % h;
    const int xF2 = (ax + 2 * fx + 2 * w) % w, yF2 = (ay + 2 * fy + 2 * h) % h;
    bool moved = false, push = false;
    int reward = 0;
    if (!f_nz) {
        moved = true;
    } else if (bit_at(P[8], yF, xF) && can_exit) {
        game_over = 1;
        reward = 1;
    } else if (bit_at(P[2], yF, xF)) {
        if (!bit_at(nz, yF2, xF2)) {
            moved = true;
            push = true;  // the pushed cell moves on to F2
        } else if (bit_at(P[8], yF2, xF2)) {
            moved = true;  // pushed out of the level: F is simply overwritten by the agent
        }
    }
    if (!moved) return reward;
    const bool pull = bit_at(P[15], yB, xB) != 0;
    // Every written cell takes the value of its predecessor along the direction of motion:
    //   F <- A always; F2 <- F when pushing; A <- B when pulling (else A <- 0; B <- 0 when pulling).
    uint32_t set_m = 0, clr_m = 0;  // this lane's destination bits that receive / are cleared
    if (lane == yF) { set_m |= 1u << xF; }
    if (push && lane == yF2) { set_m |= 1u << xF2; }
    if (lane == ay) { if (pull) set_m |= 1u << ax; else clr_m |= 1u << ax; }
    if (pull && lane == yB) { clr_m |= 1u << xB; }
    const uint32_t keep = ~(set_m | clr_m);
    if (fy == 0) {
        // horizontal: predecessor is the cell one step back along x in the same row
#pragma unroll
        for (int b = 0; b < 16; b++) {
            uint32_t pred = (fx > 0) ? from_left(P[b], g) : from_right(P[b], g);
            P[b] = (P[b] & keep) | (pred & set_m);
        }
    } else {
        // vertical: predecessor is the same column in the row one step back along y
        const int src = g.row ? ((fy > 0) ? g.up : g.dn) : lane;
#pragma unroll
        for (int b = 0; b < 16; b++) {
            uint32_t pred = __shfl_sync(FULL, P[b], src);
            P[b] = (P[b] & keep) | (pred & set_m);
        }
    }
    ax = xF;
    ay = yF;
    return reward;
}

// ---------------------------------------------------------------------------------------------
// current_points / performance_ratio on planes (safelife_game.py:554-565, 590-631).
// Returns this lane's row sums; the caller warp-reduces.
// ---------------------------------------------------------------------------------------------
struct RowScore { int points, current, possible; };

__device__ __forceinline__ RowScore score_planes(const uint32_t B[16], const uint32_t G[16],
                                                 const Geo &g) {
    const uint32_t c0 = B[9], c1 = B[10], c2 = B[11];  // cell colour bits r, g, b
    const uint32_t g0 = G[9], g1 = G[10], g2 = G[11];  // goal colour bits
    // minterms: index = r + 2 g + 4 b  ->  k r g y b m c w
    const uint32_t Ck = ~(c0 | c1 | c2), Cr = c0 & ~c1 & ~c2, Cg = ~c0 & c1 & ~c2, Cy = c0 & c1 & ~c2;
    const uint32_t Cb = ~c0 & ~c1 & c2, Cm = c0 & ~c1 & c2, Cc = ~c0 & c1 & c2, Cw = c0 & c1 & c2;
    const uint32_t Gk = ~(g0 | g1 | g2), Gr = g0 & ~g1 & ~g2, Gg = ~g0 & g1 & ~g2, Gy = g0 & g1 & ~g2;
    const uint32_t Gb = ~g0 & ~g1 & g2, Gm = g0 & ~g1 & g2, Gc = ~g0 & g1 & g2, Gw = g0 & g1 & g2;
    // point_table value classes
    const uint32_t s5 = (Gg & Cg) | (Gb & Cb) | (Gm & Cm) | (Gc & Cc);
    const uint32_t s3 = (Gr & Cr) | (Gg & Cc) | (Gy & Cy) | (Gm & Cr) |
                        (Gb & (Ck | Cg | Cm | Cc | Cw)) | (Gc & (Ck | Cg | Cb | Cw));
    const uint32_t sm1 = (Gk | Gw) & Cr;
    const uint32_t sm3 = ((Gr | Gm) & (Ck | Cg | Cb | Cc | Cw)) | ((Gg | Gb | Gc) & Cr) | (Gy & Ck);
    const uint32_t a = B[0] & g.wmask;
    // mask of performance_ratio: alive and not (frozen and immovable) (:616-620)
    const uint32_t m = a & (~B[4] | B[2] | B[15]);
    RowScore r;
    r.points = 5 * __popc(a & s5) + 3 * __popc(a & s3) - __popc(a & sm1) - 3 * __popc(a & sm3);
    r.current = __popc(m & (s5 | s3)) - __popc(m & (sm1 | sm3));
    r.possible = __popc(~(Gk | Gw) & g.wmask);  // goal colour neither black nor white
    return r;
}

// update_exit_colors (safelife_game.py:531-537): the level's initial exit cells (row mask e0) are
// overwritten with frozen|exit (272) or frozen|exit|red (784).
__device__ __forceinline__ void paint_exits(uint32_t P[16], uint32_t e0, int can_exit) {
    if (!__any_sync(FULL, e0 != 0)) return;
    const uint32_t k = ~e0;
#pragma unroll
    for (int b = 0; b < 16; b++) P[b] &= k;
    P[4] |= e0;
    P[8] |= e0;
    if (can_exit) P[9] |= e0;
}

}  // namespace slb
