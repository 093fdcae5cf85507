// slb_planes_env.cuh — execute_action, scores and exit colours on bit-planes (one warp per env).
#pragma once
#include "slb_planes.cuh"

namespace slb {

// ---------------------------------------------------------------------------------------------
// execute_action on planes (safelife_game.py:347-393, move_agent :308-345).
// Requires H >= 4 and W >= 4 so that the agent cell A, front F, two-ahead F2 and behind B are four
// distinct cells (smaller boards alias F2 and B and take the literal generic path instead).
// All reads see the pre-action board, which is what the reference reads too: the pull test on B
// happens after the moves, but no move writes B.  Warp-uniform control flow (one env per warp).
// Returns the action reward; updates ax, ay, orientation, game_over; `touched` (warp-uniform) says
// whether the board was written.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t bit_of(uint32_t plane, int x) { return (plane >> x) & 1u; }

__device__ __forceinline__ int action_planes(uint32_t P[16], const Geo &g, int &ax, int &ay,
                                             int &orientation, int &game_over, int action,
                                             int can_exit, bool &touched) {
    touched = false;
    if (game_over || action < 1 || action > 8) return 0;
    const int lane = threadIdx.x & 31;
    const int dir = (action - 1) & 3;
    orientation = dir;
    // forward vector: up (0,-1), right (1,0), down (0,1), left (-1,0)
    const int fx = (dir & 1) ? 2 - dir : 0, fy = (dir & 1) ? 0 : dir - 1;
    const int h = g.h, w = g.w;
    int xF = ax + fx, yF = ay + fy;
    xF = xF < 0 ? w - 1 : (xF >= w ? 0 : xF);
    yF = yF < 0 ? h - 1 : (yF >= h ? 0 : yF);
    uint32_t nz = 0;
#pragma unroll
    for (int b = 0; b < 16; b++) nz |= P[b];
    // front cell summary, extracted by its row's lane: bit0 non-empty, 1 exit, 2 pushable, 3 destructible
    uint32_t sF = bit_of(nz, xF) | (bit_of(P[8], xF) << 1) | (bit_of(P[2], xF) << 2) | (bit_of(P[3], xF) << 3);
    sF = __shfl_sync(FULL, sF, yF);
    if (action >= 5) {  // TOGGLE
        if (!(sF & 1u)) {
            // empty -> life | player colour (alive | destructible | colour of the agent cell)
            uint32_t cA = bit_of(P[9], ax) | (bit_of(P[10], ax) << 1) | (bit_of(P[11], ax) << 2);
            cA = __shfl_sync(FULL, cA, ay);
            const uint32_t m = (lane == yF) ? (1u << xF) : 0u;
            P[0] |= m;
            P[3] |= m;
            if (cA & 1u) P[9] |= m;
            if (cA & 2u) P[10] |= m;
            if (cA & 4u) P[11] |= m;
            touched = true;
        } else if (sF & 8u) {  // destructible -> empty
            const uint32_t k = (lane == yF) ? ~(1u << xF) : 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 16; b++) P[b] &= k;
            touched = true;
        }
        return 0;
    }
    // MOVE
    bool moved = false, push = false;
    int reward = 0;
    int xF2 = xF, yF2 = yF;
    if (!(sF & 1u)) {
        moved = true;
    } else if ((sF & 2u) && can_exit) {
        game_over = 1;
        reward = 1;
    } else if (sF & 4u) {
        xF2 = xF + fx; yF2 = yF + fy;
        xF2 = xF2 < 0 ? w - 1 : (xF2 >= w ? 0 : xF2);
        yF2 = yF2 < 0 ? h - 1 : (yF2 >= h ? 0 : yF2);
        uint32_t s2 = bit_of(nz, xF2) | (bit_of(P[8], xF2) << 1);
        s2 = __shfl_sync(FULL, s2, yF2);
        if (!(s2 & 1u)) {
            moved = true;
            push = true;  // the pushed cell moves on to F2
        } else if (s2 & 2u) {
            moved = true;  // pushed out of the level: F is simply overwritten by the agent
        }
    }
    if (!moved) return reward;
    touched = true;
    int xB = ax - fx, yB = ay - fy;
    xB = xB < 0 ? w - 1 : (xB >= w ? 0 : xB);
    yB = yB < 0 ? h - 1 : (yB >= h ? 0 : yB);
    const bool pull = (__shfl_sync(FULL, P[15], yB) >> xB) & 1u;
    // Every moving cell advances one step along the direction of motion: A -> F always, F -> F2 when
    // pushing, B -> A when pulling.  src_m: this lane's cells that move; all of them are vacated,
    // and each destination receives its predecessor (a pushed-out cell's old place F receives A).
    uint32_t src_m = 0;
    if (lane == ay) src_m |= 1u << ax;
    if (push && lane == yF) src_m |= 1u << xF;
    if (pull && lane == yB) src_m |= 1u << xB;
    uint32_t dst_m;  // this lane's cells that receive a value
    if (fy == 0) {
        // horizontal: every moving cell lies in row ay; one step along x, toroidal within W bits
        const int sh = fx > 0 ? 1 : 31;  // rotate-left amount in a 32-bit word
        const int fix = fx > 0 ? w : 32 - w;  // where a bit that left the W-bit field reappears
        const bool wrap = (w < 32) && (fx > 0 ? (src_m >> (w - 1)) & 1u : src_m & 1u);
        const bool any_wrap = __any_sync(FULL, wrap);
        dst_m = __funnelshift_l(src_m, src_m, sh);
        if (any_wrap) dst_m = (dst_m & g.wmask) | ((dst_m & ~g.wmask) >> fix);
        const uint32_t keep = ~(src_m | dst_m);
        if (!any_wrap) {
#pragma unroll
            for (int b = 0; b < 16; b++) {
                const uint32_t t = P[b] & src_m;
                P[b] = (P[b] & keep) | __funnelshift_l(t, t, sh);
            }
        } else {
#pragma unroll
            for (int b = 0; b < 16; b++) {
                uint32_t t = P[b] & src_m;
                t = __funnelshift_l(t, t, sh);
                t = (t & g.wmask) | ((t & ~g.wmask) >> fix);
                P[b] = (P[b] & keep) | t;
            }
        }
    } else {
        // vertical: same column, the value comes from the row one step back along y
        const int from = g.row ? ((fy > 0) ? g.up : g.dn) : lane;
        dst_m = __shfl_sync(FULL, src_m, from);
        const uint32_t keep = ~(src_m | dst_m);
#pragma unroll
        for (int b = 0; b < 16; b++) {
            const uint32_t t = __shfl_sync(FULL, P[b] & src_m, from);
            P[b] = (P[b] & keep) | t;
        }
    }
    ax = xF;
    ay = yF;
    return reward;
}

// The same, also reporting whether the action changed what the score reads.  The score only looks
// at alive cells; an action creates, destroys or moves whole cells, and every such change of an
// alive cell flips the alive plane somewhere (the agent is not alive, and a moving cell only ever
// lands on an empty cell, on the agent's old place or on a cell that is overwritten), so comparing
// the alive plane before and after is exact.
__device__ __forceinline__ int action_planes(uint32_t P[16], const Geo &g, int &ax, int &ay,
                                             int &orientation, int &game_over, int action,
                                             int can_exit, bool &touched, bool &life_touched) {
    const uint32_t a0 = P[0];
    const int reward = action_planes(P, g, ax, ay, orientation, game_over, action, can_exit, touched);
    life_touched = touched && __any_sync(FULL, a0 != P[0]);
    return reward;
}

// ---------------------------------------------------------------------------------------------
// current_points / performance_ratio on planes (safelife_game.py:554-565, 590-631).
// Returns this lane's row sums; the caller warp-reduces.
// ---------------------------------------------------------------------------------------------
struct RowScore { int points, current, possible; };

// f(a, b, c) for every bit position, f given by its truth table: bit (4a + 2b + c) of TT
template <unsigned TT>
__device__ __forceinline__ uint32_t lut3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(TT));
    return r;
}

__device__ __forceinline__ RowScore score_planes(const uint32_t B[16], const uint32_t G[16],
                                                 const Geo &g) {
    const uint32_t c0 = B[9], c1 = B[10], c2 = B[11];  // cell colour bits r, g, b
    const uint32_t g0 = G[9], g1 = G[10], g2 = G[11];  // goal colour bits
    // Colour index = r + 2 g + 4 b: k 0, r 1, g 2, y 3, b 4, m 5, c 6, w 7.  Every value class of the
    // point table is an OR over goal colours of (goal colour in a set) & (cell colour in a set); a
    // set of colours is a function of three planes, i.e. ONE lop3 whose truth table is the set's
    // bit mask (29 logic operations for all four classes).
#define SLB_CELL_IN(mask) lut3<(mask)>(c2, c1, c0)
#define SLB_GOAL_IN(mask) lut3<(mask)>(g2, g1, g0)
    const uint32_t Ck = SLB_CELL_IN(0x01), Cr = SLB_CELL_IN(0x02), Cy = SLB_CELL_IN(0x08), Cc = SLB_CELL_IN(0x40);
    const uint32_t Gkw = SLB_GOAL_IN(0x81);  // black or white
    const uint32_t Grm = SLB_GOAL_IN(0x22);  // red or magenta
    // +5: cell colour == goal colour and goal in {g, b, m, c}
    const uint32_t ne = (c0 ^ g0) | (c1 ^ g1) | (c2 ^ g2);
    const uint32_t s5 = ~ne & SLB_GOAL_IN(0x74);
    // +3: (r|m, r) (g, c) (y, y) (b, {k g m c w}) (c, {k g b w})
    const uint32_t Gy = SLB_GOAL_IN(0x08);
    const uint32_t s3 = (Grm & Cr) | (SLB_GOAL_IN(0x04) & Cc) | (Gy & Cy) | (SLB_GOAL_IN(0x10) & SLB_CELL_IN(0xE5)) |
                        (SLB_GOAL_IN(0x40) & SLB_CELL_IN(0x95));
    // -1: (k|w, r)
    const uint32_t sm1 = Gkw & Cr;
    // -3: (r|m, {k g b c w}) (g|b|c, r) (y, k)
    const uint32_t sm3 = (Grm & SLB_CELL_IN(0xD5)) | (SLB_GOAL_IN(0x54) & Cr) | (Gy & Ck);
#undef SLB_CELL_IN
#undef SLB_GOAL_IN
    const uint32_t a = B[0] & g.wmask;
    // mask of performance_ratio: alive and not (frozen and immovable) (:616-620)
    const uint32_t m = a & (~B[4] | B[2] | B[15]);
    RowScore r;
    r.points = 5 * __popc(a & s5) + 3 * __popc(a & s3) - __popc(a & sm1) - 3 * __popc(a & sm3);
    r.current = __popc(m & (s5 | s3)) - __popc(m & (sm1 | sm3));
    r.possible = __popc(~Gkw & g.wmask);  // goal colour neither black nor white
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

// ---------------------------------------------------------------------------------------------
// SimpleSideEffectPenalty's per-step count (env_wrappers.py:319-346) for one word of cells:
// B = current board, S = the level's initial board, G = goals, e0 = initial exit cells.  The player
// bits (agent, destructible, frozen, preserving, inhibiting = 122) are ignored on both boards, exit
// cells count as unchanged, and a cell is no side effect if it is unchanged, or was red life and no
// longer is, or lies on a blue goal and now holds non-red life.  Returns the mask of side-effect cells.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t side_effect_word(const uint32_t B[16], const uint32_t S[16],
                                                     const uint32_t G[16], uint32_t e0, uint32_t wmask) {
    uint32_t diff = 0;
#pragma unroll
    for (int b = 0; b < 16; b++) {
        if (b == 1 || b == 3 || b == 4 || b == 5 || b == 6) continue;  // CellTypes.player
        diff |= B[b] ^ S[b];
    }
    diff &= ~e0;                                   // board[exit_locs] = start_board[exit_locs]
    const uint32_t b0 = (B[0] & ~e0) | (S[0] & e0), b9 = (B[9] & ~e0) | (S[9] & e0);
    const uint32_t start_red = S[0] & S[9], end_red = b0 & b9;
    const uint32_t goal_cell = G[11] & ~G[9] & ~G[10];
    const uint32_t end_alive = b0 & ~b9;
    const uint32_t non_effects = ~diff | (start_red & ~end_red) | (goal_cell & end_alive);
    return ~non_effects & wmask;
}

// the same for one uint16 cell
__device__ __forceinline__ int side_effect_cell(uint32_t b, uint32_t s, uint32_t g, bool is_exit) {
    const uint32_t player = AGENT | DESTRUCTIBLE | FROZEN | PRESERVING | INHIBITING;
    b &= ~player; s &= ~player;
    if (is_exit) b = s;
    const uint32_t red_life = ALIVE | RED;
    const bool start_red = (s & red_life) == red_life, end_red = (b & red_life) == red_life;
    const bool goal_cell = (g & COLORS) == BLUE, end_alive = (b & red_life) == ALIVE;
    return !((b == s) || (start_red && !end_red) || (goal_cell && end_alive));
}

}  // namespace slb
