// slb_generic_kernels.cuh — __global__ kernels of the shape-generic path (see slb_generic.cuh).
#pragma once
#include "slb_generic.cuh"

namespace slb {

// ---------------------------------------------------------------------------------------------
// A1: advance_board for N boards, out of place.  grid = N, block = GEN_THREADS.
// USE_SMEM: the old board is staged in shared memory (h*w*2 bytes dynamic).
// ---------------------------------------------------------------------------------------------
template <bool USE_SMEM>
__global__ void __launch_bounds__(GEN_THREADS)
advance_generic_kernel(const uint16_t *__restrict__ in, uint16_t *__restrict__ out, int h, int w,
                       const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                       uint64_t *counters) {
    extern __shared__ uint16_t sm_board[];
    __shared__ int warp_tot[GEN_THREADS / 32];
    const int i = blockIdx.x, cells = h * w;
    const uint16_t *src = in + (size_t)i * cells;
    if (USE_SMEM) {
        for (int c = threadIdx.x; c < cells; c += blockDim.x) sm_board[c] = src[c];
        __syncthreads();
        src = sm_board;
    }
    uint64_t counter = counters ? counters[i] : 0;
    int d0 = 0, d1 = 0, d2 = 0;
    counter = advance_pass<false>(src, out + (size_t)i * cells, h, w, spawn_threshold(spawn_prob[i]),
                                  seed, env_id0 + (uint64_t)i, counter, warp_tot, nullptr, d0, d1, d2);
    if (counters && threadIdx.x == 0) counters[i] = counter;
}

// ---------------------------------------------------------------------------------------------
// pool prepare: baseline score + exit mask per level.  grid = P.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GEN_THREADS) pool_prepare_kernel(slb_pool pool) {
    __shared__ int scratch[96];
    const int l = blockIdx.x, h = pool.h, w = pool.w, cells = h * w, wpr = pool.words_per_row;
    const uint16_t *b = pool.board + (size_t)l * cells, *g = pool.goals + (size_t)l * cells;
    uint32_t *em = pool.exit_mask + (size_t)l * h * wpr;
    for (int k = threadIdx.x; k < h * wpr; k += blockDim.x) {
        int y = k / wpr, x0 = (k - y * wpr) * 32;
        uint32_t m = 0;
        for (int j = 0; j < 32 && x0 + j < w; j++)
            if (b[y * w + x0 + j] & EXIT) m |= 1u << j;
        em[k] = m;
    }
    int pts = 0, cur = 0, pos = 0;
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        CellScore s = cell_score(b[c], g[c]);
        cur += s.current;
    }
    block_sum3(pts, cur, pos, scratch);
    if (threadIdx.x == 0) {
        pool.baseline[l] = cur;
        slb_level_consts rec = {};
        rec.spawn_prob = pool.spawn_prob[l];
        rec.baseline = cur;
        rec.flags = 0u;  // res_level_flags_kernel adds SLB_LEVEL_GOALS_STATIC on fast-path shapes
        rec.min_performance = pool.min_performance[l];
        reinterpret_cast<slb_level_consts *>(pool.level_consts)[l] = rec;
    }
}

// CTA-wide reset of one env from the pool (safelife_env.py:188-198).  Every thread must call.
__device__ inline void reset_env_cta(const slb_pool &pool, int level, uint16_t *board,
                                     uint16_t *goals, slb_env_state *st, int *scratch) {
    const int h = pool.h, w = pool.w, cells = h * w;
    const uint16_t *b0 = pool.board + (size_t)level * cells, *g0 = pool.goals + (size_t)level * cells;
    int pts = 0, cur = 0, pos = 0;
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        uint32_t b = b0[c], g = g0[c];
        CellScore s = cell_score(b, g);
        // _old_game_value is taken AFTER update_exit_colors (safelife_env.py:191-192): cells at the
        // exit locations are repainted to the (dead) exit value and score nothing
        pts += (b & EXIT) ? 0 : s.points;
        cur += s.current;
        pos += s.possible;
        goals[c] = (uint16_t)g;
        board[c] = (uint16_t)b;
    }
    block_sum3(pts, cur, pos, scratch);
    const int baseline = pool.baseline[level];
    const int completed = cur - baseline, possible = pos - baseline;
    const int ce = can_exit_rule(pool.min_performance[level], completed, possible);
    // update_exit_colors (safelife_game.py:531-537)
    const uint16_t exit_val = (uint16_t)(FROZEN | EXIT | (ce ? RED : 0u));
    for (int c = threadIdx.x; c < cells; c += blockDim.x)
        if (b0[c] & EXIT) board[c] = exit_val;
    if (threadIdx.x == 0) {
        st->agent_x = pool.agent_x[level];
        st->agent_y = pool.agent_y[level];
        st->orientation = pool.orientation[level];
        st->game_over = 0;
        st->num_steps = 0;
        st->level_id = level;
        st->episode_length = 0;
        st->episode_completed = 0;
        st->old_points = pts;
        st->can_exit = ce;
        st->completed = completed;
        st->possible = possible;
        st->episode_reward = 0;
        st->min_performance = pool.min_performance[level];  // deserialize (safelife_game.py:205)
        st->flags = 0;
        // rng_counter deliberately untouched: the stream keeps running across episodes
    }
    __syncthreads();
}

__global__ void __launch_bounds__(GEN_THREADS)
reset_generic_kernel(slb_pool pool, const int32_t *__restrict__ level_ids,
                     const uint8_t *__restrict__ mask, uint16_t *board, uint16_t *goals,
                     slb_env_state *state) {
    __shared__ int scratch[96];
    const int i = blockIdx.x;
    if (mask && !mask[i]) return;
    const size_t cells = (size_t)pool.h * pool.w;
    reset_env_cta(pool, level_ids[i], board + i * cells, goals + i * cells, state + i, scratch);
}

// ---------------------------------------------------------------------------------------------
// A7/A8 stand-alone: one thread per env.
// ---------------------------------------------------------------------------------------------
__global__ void execute_action_kernel(uint16_t *board, int n, int h, int w, slb_env_state *state,
                                      const uint8_t *__restrict__ actions, int32_t *reward) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    slb_env_state *s = state + i;
    int ax = s->agent_x, ay = s->agent_y, o = s->orientation, go = s->game_over;
    int r = execute_action_cells(board + (size_t)i * h * w, h, w, ax, ay, o, go, actions[i],
                                 s->can_exit);
    s->agent_x = ax; s->agent_y = ay; s->orientation = o; s->game_over = go;
    if (reward) reward[i] = r;
}

// ---------------------------------------------------------------------------------------------
// A9-A11 stand-alone: grid = N.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GEN_THREADS)
points_generic_kernel(slb_pool pool, const uint16_t *__restrict__ board,
                      const uint16_t *__restrict__ goals, const slb_env_state *__restrict__ state,
                      int32_t *points, int32_t *completed, int32_t *possible, uint8_t *can_exit) {
    __shared__ int scratch[96];
    const int i = blockIdx.x, cells = pool.h * pool.w;
    const uint16_t *b = board + (size_t)i * cells, *g = goals + (size_t)i * cells;
    int pts = 0, cur = 0, pos = 0;
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        CellScore s = cell_score(b[c], g[c]);
        pts += s.points; cur += s.current; pos += s.possible;
    }
    block_sum3(pts, cur, pos, scratch);
    if (threadIdx.x == 0) {
        const int level = state[i].level_id, baseline = pool.baseline[level];
        if (points) points[i] = pts;
        if (completed) completed[i] = cur - baseline;
        if (possible) possible[i] = pos - baseline;
        if (can_exit)
            can_exit[i] = (uint8_t)can_exit_rule(state[i].min_performance, cur - baseline, pos - baseline);
    }
}

// ---------------------------------------------------------------------------------------------
// update_exit_colors stand-alone (safelife_game.py:531-537): can_exit() on the current board, then
// repaint the initial exit locations; refreshes state.can_exit/completed/possible.  grid = N.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GEN_THREADS)
exit_colors_generic_kernel(slb_pool pool, uint16_t *board, const uint16_t *__restrict__ goals,
                           slb_env_state *state) {
    __shared__ int scratch[96];
    const int i = blockIdx.x, cells = pool.h * pool.w;
    uint16_t *b = board + (size_t)i * cells;
    const uint16_t *g = goals + (size_t)i * cells;
    int pts = 0, cur = 0, pos = 0;
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        CellScore s = cell_score(b[c], g[c]);
        cur += s.current; pos += s.possible;
    }
    block_sum3(pts, cur, pos, scratch);
    const int level = state[i].level_id, baseline = pool.baseline[level];
    const int ce = can_exit_rule(state[i].min_performance, cur - baseline, pos - baseline);
    const uint16_t *b0 = pool.board + (size_t)level * cells;
    const uint16_t exit_val = (uint16_t)(FROZEN | EXIT | (ce ? RED : 0u));
    for (int c = threadIdx.x; c < cells; c += blockDim.x)
        if (b0[c] & EXIT) b[c] = exit_val;
    if (threadIdx.x == 0) {
        state[i].can_exit = ce;
        state[i].completed = cur - baseline;
        state[i].possible = pos - baseline;
        state[i].flags &= ~SLB_ENV_RESCORE;
    }
}

}  // namespace slb
