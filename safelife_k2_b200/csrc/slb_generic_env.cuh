// slb_generic_env.cuh — shape-generic fused env step, side-effect rollout, stats and observation
// kernels.  One CTA per env / episode.
#pragma once
#include "slb_generic_kernels.cuh"
#include "slb_planes_env.cuh"

namespace slb {

// ---------------------------------------------------------------------------------------------
// A12: SafeLifeEnv.step (safelife_env.py:157-186), fused.  grid = N, dynamic smem = h*w*2.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GEN_THREADS)
env_step_generic_kernel(slb_pool pool, uint16_t *board, uint16_t *goals, slb_env_state *state,
                        const uint8_t *__restrict__ actions, int time_limit, uint64_t seed,
                        uint64_t env_id0, int32_t *reward_out, uint8_t *done_out,
                        int auto_reset_stride) {
    extern __shared__ uint16_t sm_cells[];
    __shared__ int warp_tot[GEN_THREADS / 32];
    __shared__ int scratch[96];
    __shared__ int sh_act[5];  // ax, ay, orientation, game_over, action reward

    const int i = blockIdx.x, h = pool.h, w = pool.w, cells = h * w;
    uint16_t *b = board + (size_t)i * cells, *g = goals + (size_t)i * cells;
    slb_env_state *st = state + i;
    const int level = st->level_id;
    const uint64_t env = env_id0 + (uint64_t)i;
    const uint64_t thr = spawn_threshold(pool.spawn_prob[level]);
    uint64_t counter = st->rng_counter;

    for (int c = threadIdx.x; c < cells; c += blockDim.x) sm_cells[c] = b[c];
    __syncthreads();
    if (threadIdx.x == 0) {
        int ax = st->agent_x, ay = st->agent_y, o = st->orientation, go = st->game_over;
        int r = execute_action_cells(sm_cells, h, w, ax, ay, o, go, actions[i], st->can_exit);
        sh_act[0] = ax; sh_act[1] = ay; sh_act[2] = o; sh_act[3] = go; sh_act[4] = r;
    }
    __syncthreads();
    int pts = 0, cur = 0, pos = 0;
    // board first, then goals: the RNG order of SafeLifeGame.advance_board (safelife_game.py:657-660)
    counter = advance_pass<false>(sm_cells, b, h, w, thr, seed, env, counter, warp_tot, nullptr,
                                  pts, cur, pos);
    __syncthreads();
    for (int c = threadIdx.x; c < cells; c += blockDim.x) sm_cells[c] = g[c];
    __syncthreads();
    counter = advance_pass<true>(sm_cells, g, h, w, thr, seed, env, counter, warp_tot, b, pts, cur,
                                 pos);
    block_sum3(pts, cur, pos, scratch);

    const int baseline = pool.baseline[level];
    const int completed = cur - baseline, possible = pos - baseline;
    const int ce = can_exit_rule(st->min_performance, completed, possible);
    // update_exit_colors: exit locations of the INITIAL board (safelife_game.py:528-537)
    const uint16_t *b0 = pool.board + (size_t)level * cells;
    const uint16_t exit_val = (uint16_t)(FROZEN | EXIT | (ce ? RED : 0u));
    for (int c = threadIdx.x; c < cells; c += blockDim.x)
        if (b0[c] & EXIT) b[c] = exit_val;

    __shared__ int sh_done;
    if (threadIdx.x == 0) {
        const int reward = sh_act[4] + pts - st->old_points;
        const int ep_len = st->episode_length + 1;
        const int times_up = ep_len > time_limit;
        const int done = times_up || sh_act[3];
        st->agent_x = sh_act[0]; st->agent_y = sh_act[1]; st->orientation = sh_act[2];
        st->game_over = sh_act[3];
        st->num_steps += 1;
        st->episode_length = ep_len;
        st->episode_completed = done;
        st->old_points = pts;
        st->can_exit = ce;
        st->completed = completed;
        st->possible = possible;
        st->episode_reward += reward;
        st->rng_counter = counter;
        st->flags = 0;  // this kernel re-scores and repaints every step
        if (reward_out) reward_out[i] = reward;
        if (done_out) done_out[i] = (uint8_t)done;
        sh_done = done;
    }
    __syncthreads();
    if (auto_reset_stride > 0 && sh_done) {
        __threadfence_block();
        reset_env_cta(pool, (level + auto_reset_stride) % pool.n_levels, b, g, st, scratch);
    }
}

// ---------------------------------------------------------------------------------------------
// SafeLifeGame.advance_board alone (safelife_game.py:657-660) on uint16-resident state.  grid = N,
// dynamic smem = h*w*2.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(GEN_THREADS)
game_advance_generic_kernel(slb_pool pool, uint16_t *board, uint16_t *goals, slb_env_state *state,
                            uint64_t seed, uint64_t env_id0) {
    extern __shared__ uint16_t sm_cells[];
    __shared__ int warp_tot[GEN_THREADS / 32];
    const int i = blockIdx.x, h = pool.h, w = pool.w, cells = h * w;
    uint16_t *b = board + (size_t)i * cells, *g = goals + (size_t)i * cells;
    slb_env_state *st = state + i;
    const uint64_t env = env_id0 + (uint64_t)i;
    const uint64_t thr = spawn_threshold(pool.spawn_prob[st->level_id]);
    uint64_t counter = st->rng_counter;
    int d0 = 0, d1 = 0, d2 = 0;
    for (int c = threadIdx.x; c < cells; c += blockDim.x) sm_cells[c] = b[c];
    __syncthreads();
    counter = advance_pass<false>(sm_cells, b, h, w, thr, seed, env, counter, warp_tot, nullptr, d0, d1, d2);
    __syncthreads();
    for (int c = threadIdx.x; c < cells; c += blockDim.x) sm_cells[c] = g[c];
    __syncthreads();
    counter = advance_pass<false>(sm_cells, g, h, w, thr, seed, env, counter, warp_tot, nullptr, d0, d1, d2);
    if (threadIdx.x == 0) {
        st->rng_counter = counter;
        st->num_steps += 1;
    }
}

// ---------------------------------------------------------------------------------------------
// A13/A14: side_effect_score up to the EMD (side_effects.py:59-141).  grid = N episodes.
// buf: 4 boards of h*w cells (b0 ping/pong, b1 ping/pong) in shared memory (USE_SMEM) or in the
// caller's scratch.  Key table per episode in shared memory; count maps in global memory.
// ---------------------------------------------------------------------------------------------
constexpr int SE_MAX_KEYS = 64;

__device__ __forceinline__ void se_accumulate(const uint16_t *bd, int cells, int kmax,
                                              uint32_t *sh_keys, int *sh_overflow,
                                              uint16_t *counts) {
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        uint32_t v = bd[c];
        bool unchanging = (v & (FROZEN | DESTRUCTIBLE | MOVABLE)) == FROZEN;  // side_effects.py:61
        uint32_t ctype = unchanging ? 0u : (v & ~(uint32_t)DESTRUCTIBLE);     // :62
        if (!ctype || (ctype & AGENT)) continue;                              // :68-70
        uint32_t key = ctype, base = ctype & ~(uint32_t)COLORS;
        if (base == ALIVE || base == (FROZEN | SPAWNING)) key |= DESTRUCTIBLE;  // :72-75
        int slot = -1;
        for (int k = 0; k < kmax; k++) {
            uint32_t cur = ((volatile uint32_t *)sh_keys)[k];
            if (cur == 0) {  // empty slot: try to claim it; if somebody else won, look at what they put
                uint32_t prev = atomicCAS(&sh_keys[k], 0u, key);
                cur = prev ? prev : key;
            }
            if (cur == key) { slot = k; break; }
        }
        if (slot < 0) { *sh_overflow = 1; continue; }
        counts[(size_t)slot * cells + c] += 1;  // each cell is owned by one thread
    }
}

template <bool USE_SMEM>
__global__ void __launch_bounds__(GEN_THREADS)
side_effect_generic_kernel(const uint16_t *__restrict__ board0, const uint16_t *__restrict__ board1,
                           int h, int w, const int32_t *__restrict__ num_steps, int num_samples,
                           const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                           uint64_t *counters, int kmax, uint16_t *keys, uint16_t *counts,
                           int32_t *nkeys, uint16_t *scratch_g) {
    extern __shared__ uint16_t sm_buf[];
    __shared__ int warp_tot[GEN_THREADS / 32];
    __shared__ uint32_t sh_keys[SE_MAX_KEYS];
    __shared__ int sh_overflow;
    const int i = blockIdx.x, cells = h * w;
    uint16_t *buf = USE_SMEM ? sm_buf : scratch_g + (size_t)i * 4 * cells;
    uint16_t *a0 = buf, *a1 = buf + cells, *c0 = buf + 2 * cells, *c1 = buf + 3 * cells;
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        a0[c] = board0[(size_t)i * cells + c];
        c0[c] = board1[(size_t)i * cells + c];
    }
    if (threadIdx.x < SE_MAX_KEYS) sh_keys[threadIdx.x] = 0;
    if (threadIdx.x == 0) sh_overflow = 0;
    __syncthreads();
    const uint64_t env = env_id0 + (uint64_t)i, thr = spawn_threshold(spawn_prob[i]);
    uint64_t counter = counters ? counters[i] : 0;
    int d0 = 0, d1 = 0, d2 = 0;
    uint16_t *cnt0 = counts + (size_t)i * 2 * kmax * cells, *cnt1 = cnt0 + (size_t)kmax * cells;
    const int catch_up = num_steps[i];
    for (int t = 0; t < catch_up; t++) {  // side_effects.py:135-136
        counter = advance_pass<false>(a0, a1, h, w, thr, seed, env, counter, warp_tot, nullptr, d0, d1, d2);
        __syncthreads();
        uint16_t *tmp = a0; a0 = a1; a1 = tmp;
    }
    for (int t = 0; t < num_samples; t++) {  // :137-141, inaction board draws first
        counter = advance_pass<false>(a0, a1, h, w, thr, seed, env, counter, warp_tot, nullptr, d0, d1, d2);
        counter = advance_pass<false>(c0, c1, h, w, thr, seed, env, counter, warp_tot, nullptr, d0, d1, d2);
        __syncthreads();
        uint16_t *tmp = a0; a0 = a1; a1 = tmp;
        tmp = c0; c0 = c1; c1 = tmp;
        se_accumulate(a0, cells, kmax, sh_keys, &sh_overflow, cnt0);
        se_accumulate(c0, cells, kmax, sh_keys, &sh_overflow, cnt1);
        __syncthreads();
    }
    if (counters && threadIdx.x == 0) counters[i] = counter;
    if (threadIdx.x < kmax) keys[(size_t)i * kmax + threadIdx.x] = (uint16_t)sh_keys[threadIdx.x];
    if (threadIdx.x == 0) {
        int nk = 0;
        for (int k = 0; k < kmax; k++) nk += sh_keys[k] != 0;
        nkeys[i] = sh_overflow ? -1 : nk;
    }
}

// SimpleSideEffectPenalty's count of side-effect cells per env (env_wrappers.py:319-346); grid = N
__global__ void __launch_bounds__(GEN_THREADS)
penalty_generic_kernel(slb_pool pool, const uint16_t *__restrict__ board,
                       const uint16_t *__restrict__ goals, const slb_env_state *__restrict__ state,
                       int32_t *count) {
    __shared__ int scratch[96];
    const int i = blockIdx.x, cells = pool.h * pool.w, level = state[i].level_id;
    const uint16_t *b = board + (size_t)i * cells, *g = goals + (size_t)i * cells;
    const uint16_t *s = pool.board + (size_t)level * cells;
    int c = 0, d1 = 0, d2 = 0;
    for (int k = threadIdx.x; k < cells; k += blockDim.x) c += side_effect_cell(b[k], s[k], g[k], (s[k] & EXIT) != 0);
    block_sum3(c, d1, d2, scratch);
    if (threadIdx.x == 0) count[i] = c;
}

// ---------------------------------------------------------------------------------------------
// statistics for the all-reduce.  Single CTA loops over envs (n is small next to a step).
// ---------------------------------------------------------------------------------------------
// SafeLifeEnv.global_counter (safelife_env.py:169-173) around a step kernel that does not count itself:
// who was done before the step, then (after it) num_steps += fresh, episodes_completed += done & fresh.
__global__ void __launch_bounds__(256)
gc_pre_kernel(int n, const slb_env_state *__restrict__ state, uint8_t *already) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) already[i] = (uint8_t)(state[i].episode_completed != 0);
}

__global__ void __launch_bounds__(256)
gc_post_kernel(int n, const uint8_t *__restrict__ done, const uint8_t *__restrict__ already,
               unsigned long long *gc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool fresh = i < n && !already[i];
    const unsigned f = __ballot_sync(0xffffffffu, fresh), d = __ballot_sync(0xffffffffu, fresh && done[i]);
    if ((threadIdx.x & 31) == 0) {
        const int slot = (i >> 5) & 63;
        if (f) atomicAdd(gc + 2 * slot, (unsigned long long)__popc(f));
        if (d) atomicAdd(gc + 2 * slot + 1, (unsigned long long)__popc(d));
    }
}

// `game.min_performance = value` from outside (env_wrappers.py:316): can_exit() re-evaluated from the
// env's last performance_ratio(), exits repainted by the next step (SLB_ENV_RESCORE).
__global__ void __launch_bounds__(256)
set_min_performance_kernel(int n, slb_env_state *state, const uint8_t *__restrict__ mask, double value) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || (mask && !mask[i])) return;
    slb_env_state *s = state + i;
    s->min_performance = value;
    s->can_exit = can_exit_rule(value, s->completed, s->possible);
    s->flags |= SLB_ENV_RESCORE;
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
stats_kernel(int n, const slb_env_state *__restrict__ state, const int32_t *__restrict__ reward,
             const uint8_t *__restrict__ done, unsigned long long *stats) {
    long long acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const slb_env_state s = state[i];
        acc[0] += reward ? reward[i] : 0;
        const int d = done ? done[i] : s.episode_completed;
        acc[1] += d;
        if (d) {
            acc[2] += s.episode_length;
            acc[3] += s.episode_reward;
            acc[4] += s.completed;
            acc[5] += s.possible;
        }
        acc[6] += 1;
        acc[7] += s.game_over;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        long long v = acc[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v != 0) atomicAdd(stats + k, (unsigned long long)v);
    }
}

// ---------------------------------------------------------------------------------------------
// get_obs (safelife_env.py:125-155) + recenter_view (helper_utils.py:41-74).  grid = N.
// out[i][vy][vx] = (board + ((goals & colours, white removed) << 3)) cropped toroidally around the
// agent; exit cells of the INITIAL board that fall outside the view are clamped onto its rim, in
// raster order of the exit cells (later ones overwrite earlier ones, like the fancy assignment).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint16_t obs_word(uint32_t b, uint32_t g, int remove_white) {
    g &= COLORS;
    if (remove_white && g == COLORS) g = 0;
    return (uint16_t)(b + (g << 3));  // uint16 add, wraps like numpy
}

__device__ __forceinline__ int pymod_i(int a, int m) {
    int r = a % m;
    return r < 0 ? r + m : r;
}

__global__ void __launch_bounds__(GEN_THREADS)
get_obs_kernel(slb_pool pool, const uint16_t *__restrict__ board, const uint16_t *__restrict__ goals,
               const slb_env_state *__restrict__ state, int vh, int vw, int remove_white,
               uint16_t *obs) {
    const int i = blockIdx.x, h = pool.h, w = pool.w, cells = h * w;
    const uint16_t *b = board + (size_t)i * cells, *g = goals + (size_t)i * cells;
    const int x0 = state[i].agent_x, y0 = state[i].agent_y, level = state[i].level_id;
    const int x1 = x0 - vw / 2, y1 = y0 - vh / 2;
    uint16_t *o = obs + (size_t)i * vh * vw;
    for (int k = threadIdx.x; k < vh * vw; k += blockDim.x) {
        int vy = k / vw, vx = k - vy * vw;
        int c = pymod_i(y1 + vy, h) * w + pymod_i(x1 + vx, w);
        o[k] = obs_word(b[c], g[c], remove_white);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t *em = pool.exit_mask + (size_t)level * h * pool.words_per_row;
        for (int y = 0; y < h; y++)
            for (int wd = 0; wd < pool.words_per_row; wd++) {
                uint32_t m = em[y * pool.words_per_row + wd];
                while (m) {
                    int x = wd * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    int jy = pymod_i(y - y0 + h / 2, h) - h / 2;
                    int jx = pymod_i(x - x0 + w / 2, w) - w / 2;
                    jy = min(max(jy + vh / 2, 0), vh - 1);
                    jx = min(max(jx + vw / 2, 0), vw - 1);
                    o[jy * vw + jx] = obs_word(b[y * w + x], g[y * w + x], remove_white);
                }
            }
    }
}

}  // namespace slb
