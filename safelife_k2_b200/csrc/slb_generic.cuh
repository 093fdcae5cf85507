// slb_generic.cuh — shape-generic kernels (any H, W >= 3): one CTA per board/env, one cell per
// thread per pass, uint16 word-per-cell.  These are the correctness-first path and the fallback
// for shapes without a bit-sliced fast path (slb_fast.cuh).  Semantics: SURVEY.md Appendix A/B.
#pragma once
#include "slb_common.cuh"

namespace slb {

constexpr int GEN_THREADS = 256;

// ---------------------------------------------------------------------------------------------
// One cell of advance_board (advance_board.c:34-120 in closed form).  `src` is the whole board
// (shared or global memory).  Returns the new value when no draw is needed; when the cell is
// spawn-eligible (dead, not frozen, not inhibited, n != 3, spawner in the 3x3 — the only place a
// draw is consumed, advance_board.c:110) sets eligible and returns the SPAWNED value; the caller
// picks it or the old value after drawing.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cell_rule(const uint16_t *src, int h, int w, int y, int x,
                                              bool &eligible) {
    int ym = (y == 0) ? h - 1 : y - 1, yp = (y == h - 1) ? 0 : y + 1;
    int xm = (x == 0) ? w - 1 : x - 1, xp = (x == w - 1) ? 0 : x + 1;
    const int rows[3] = {ym * w, y * w, yp * w};
    const int cols[3] = {xm, x, xp};
    uint32_t n = 0, flags = 0, once = 0, twice = 0, spc = 0;
#pragma unroll
    for (int j = 0; j < 3; j++) {
#pragma unroll
        for (int i = 0; i < 3; i++) {
            uint32_t v = src[rows[j] + cols[i]];
            flags |= v;
            if (v & SPAWNING) spc |= v;
            if (v & ALIVE) {
                n++;
                // bit 8 <- destructible|exit (advance_board.c:44-46), bits 9-11 colours
                uint32_t f = (v & (EXIT | COLORS)) | ((v & DESTRUCTIBLE) << 5);
                twice |= once & f;
                once |= f;
            }
        }
    }
    uint32_t self = src[rows[1] + x];
    uint32_t inherit = (twice | spc) & COLORS;
    eligible = false;
    if (self & ALIVE) {
        bool keep = (self & FROZEN) || (flags & PRESERVING) || n == 3 || n == 4;
        return keep ? self : 0u;
    }
    if ((self & FROZEN) || (flags & INHIBITING)) return self;
    if (n == 3) return ALIVE | inherit | ((twice & EXIT) ? (uint32_t)DESTRUCTIBLE : 0u);
    if (flags & SPAWNING) {
        eligible = true;
        return ALIVE | DESTRUCTIBLE | inherit;
    }
    return self;
}

// Block-wide exclusive prefix count of a predicate over threads in thread-index order, plus the
// block total.  `warp_tot` is shared scratch of at least blockDim/32 ints.  Two __syncthreads.
__device__ __forceinline__ int block_rank(bool pred, int *warp_tot, int &total) {
    unsigned ballot = __ballot_sync(0xffffffffu, pred);
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int r = __popc(ballot & ((1u << lane) - 1u));
    if (lane == 0) warp_tot[wid] = __popc(ballot);
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5, before = 0, tot = 0;
    for (int k = 0; k < nw; k++) {
        int t = warp_tot[k];
        if (k < wid) before += t;
        tot += t;
    }
    __syncthreads();
    total = tot;
    return r + before;
}

// One advance_board pass for the CTA's board.  src: old board (shared or global, not written),
// dst: global output.  `counter` is the env's running draw counter (same value in every thread);
// returns the updated counter.  If `score_goals` is set this pass is the GOALS pass of an env step:
// for every cell the freshly written BOARD value (board_new[c], written by this same thread in the
// previous pass) and the new goals value feed the score accumulators.
template <bool SCORE>
__device__ __forceinline__ uint64_t advance_pass(const uint16_t *src, uint16_t *dst, int h, int w,
                                                 uint64_t thr, uint64_t seed, uint64_t env,
                                                 uint64_t counter, int *warp_tot,
                                                 const uint16_t *board_new, int &points,
                                                 int &current, int &possible) {
    const int cells = h * w;
    for (int base = 0; base < cells; base += blockDim.x) {
        int c = base + threadIdx.x;
        bool eligible = false;
        uint32_t nv = 0, self = 0;
        if (c < cells) {
            int y = c / w, x = c - y * w;
            nv = cell_rule(src, h, w, y, x, eligible);
            self = src[c];
        }
        int total;
        int rank = block_rank(eligible, warp_tot, total);
        if (eligible) {
            uint64_t u = stream_u53(seed, env, counter + (uint64_t)rank);
            if (!(u < thr)) nv = self;
        }
        counter += (uint64_t)total;
        if (c < cells) {
            dst[c] = (uint16_t)nv;
            if (SCORE) {
                CellScore s = cell_score(board_new[c], nv);
                points += s.points;
                current += s.current;
                possible += s.possible;
            }
        }
    }
    return counter;
}

// Block reduction of three ints; result valid in every thread.  scratch: 3 * 32 ints.
__device__ __forceinline__ void block_sum3(int &a, int &b, int &c, int *scratch) {
    a = warp_sum(a);
    b = warp_sum(b);
    c = warp_sum(c);
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) {
        scratch[wid] = a;
        scratch[32 + wid] = b;
        scratch[64 + wid] = c;
    }
    __syncthreads();
    int ra = 0, rb = 0, rc = 0;
    for (int k = 0; k < nw; k++) {
        ra += scratch[k];
        rb += scratch[32 + k];
        rc += scratch[64 + k];
    }
    a = ra; b = rb; c = rc;
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// execute_action on a word-per-cell board (safelife_game.py:347-393, move_agent :308-345), literal
// read/write order so that wrap aliasing (W == 3: two-ahead == behind) behaves identically.
// Run by ONE thread.  Returns the action reward.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int wrap_add(int a, int d, int m) {
    int r = (a + d) % m;
    return r < 0 ? r + m : r;
}

__device__ inline int execute_action_cells(uint16_t *b, int h, int w, int &ax, int &ay,
                                           int &orientation, int &game_over, int action,
                                           int can_exit) {
    if (game_over || action < 1 || action > 8) return 0;
    int dir = (action - 1) & 3;
    orientation = dir;
    // forward vector: up (0,-1), right (1,0), down (0,1), left (-1,0)  (relative_loc :294-306)
    int fx = (dir == 1) - (dir == 3), fy = (dir == 2) - (dir == 0);
    int x0 = ax, y0 = ay;
    int x1 = wrap_add(x0, fx, w), y1 = wrap_add(y0, fy, h);
    int reward = 0;
    if (action <= 4) {
        int x2 = wrap_add(x0, -fx, w), y2 = wrap_add(y0, -fy, h);
        bool moved = false;
        uint32_t front = b[y1 * w + x1];
        if (front == 0) {
            b[y1 * w + x1] = b[y0 * w + x0];
            b[y0 * w + x0] = 0;
            moved = true;
        } else if ((front & EXIT) && can_exit) {
            game_over = 1;
            reward = 1;
        } else if (front & PUSHABLE) {
            int x3 = wrap_add(x0, 2 * fx, w), y3 = wrap_add(y0, 2 * fy, h);
            uint32_t far = b[y3 * w + x3];
            if (far == 0) {
                b[y3 * w + x3] = (uint16_t)front;
                b[y1 * w + x1] = b[y0 * w + x0];
                b[y0 * w + x0] = 0;
                moved = true;
            } else if (far & EXIT) {
                b[y1 * w + x1] = b[y0 * w + x0];
                b[y0 * w + x0] = 0;
                moved = true;
            }
        }
        if (moved) {
            ax = x1;
            ay = y1;
            if (!(x0 == x1 && y0 == y1) && (b[y2 * w + x2] & PULLABLE)) {
                b[y0 * w + x0] = b[y2 * w + x2];
                b[y2 * w + x2] = 0;
            }
        }
    } else {
        uint32_t pc = b[y0 * w + x0] & COLORS;
        uint32_t t = b[y1 * w + x1];
        if (t == 0)
            b[y1 * w + x1] = (uint16_t)(ALIVE | DESTRUCTIBLE | pc);
        else if (t & DESTRUCTIBLE)
            b[y1 * w + x1] = 0;
    }
    return reward;
}

}  // namespace slb
