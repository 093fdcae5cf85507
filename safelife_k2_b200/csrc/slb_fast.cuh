// slb_fast.cuh — bit-sliced fast path for boards with 4 <= H, W <= 32 (the 26x26 benchmark levels):
// one warp per board / env, state in registers as bit-planes, k steps per HBM round trip.
#pragma once
#include "slb_planes_env.cuh"

namespace slb {

constexpr int FAST_WARPS = 4;                 // warps (= envs) per CTA
constexpr int FAST_STAGE_U16 = 32 * 34;       // per-warp staging area (H <= 32 rows, stride <= 34)

inline bool fast_shape(int h, int w) { return h >= 4 && w >= 4 && h <= 32 && w <= 32; }

// all boards 8-byte aligned and the shared image equals the memory image
__host__ __device__ inline bool fast_vec_ok(int h, int w, const void *a, const void *b) {
    return row_stride(w) == w && ((size_t)h * w * 2) % 8 == 0 && ((uintptr_t)a % 8) == 0 &&
           ((uintptr_t)b % 8) == 0;
}

// ---------------------------------------------------------------------------------------------
// A1: advance_board, k steps on chip.  grid = ceil(N / FAST_WARPS).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FAST_WARPS * 32)
advance_fast_kernel(const uint16_t *__restrict__ in, uint16_t *__restrict__ out, int n, int h, int w,
                    const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                    uint64_t *counters, int k_steps, bool vec) {
    __shared__ __align__(16) uint16_t stage[FAST_WARPS][FAST_STAGE_U16];
    const int warp = threadIdx.x >> 5, i = blockIdx.x * FAST_WARPS + warp;
    if (i >= n) return;
    const Geo g = make_geo(h, w);
    const size_t cells = (size_t)h * w;
    uint32_t P[16];
    load_planes(in + i * cells, stage[warp], h, w, vec, P);
    uint64_t counter = counters ? counters[i] : 0;
    const uint64_t thr = spawn_threshold(spawn_prob[i]);
    for (int k = 0; k < k_steps; k++) counter = physics_planes(P, g, thr, seed, env_id0 + i, counter);
    store_planes(out + i * cells, stage[warp], h, w, vec, P);
    if (counters && (threadIdx.x & 31) == 0) counters[i] = counter;
}

inline void launch_advance_fast(const uint16_t *in, uint16_t *out, int n, int h, int w, const float *p,
                                uint64_t seed, uint64_t env0, uint64_t *ctr, int k, cudaStream_t st) {
    int grid = (n + FAST_WARPS - 1) / FAST_WARPS;
    advance_fast_kernel<<<grid, FAST_WARPS * 32, 0, st>>>(in, out, n, h, w, p, seed, env0, ctr, k,
                                                          fast_vec_ok(h, w, in, out));
}

// ---------------------------------------------------------------------------------------------
// Env state in registers (warp-uniform scalars)
// ---------------------------------------------------------------------------------------------
struct EnvRegs {
    int ax, ay, orientation, game_over, num_steps, level, ep_len, ep_done;
    int old_points, can_exit, completed, possible;
    long long ep_reward;
    uint64_t counter;
};

__device__ __forceinline__ EnvRegs load_env(const slb_env_state *s) {
    EnvRegs e;
    e.ax = s->agent_x; e.ay = s->agent_y; e.orientation = s->orientation; e.game_over = s->game_over;
    e.num_steps = s->num_steps; e.level = s->level_id; e.ep_len = s->episode_length;
    e.ep_done = s->episode_completed; e.old_points = s->old_points; e.can_exit = s->can_exit;
    e.completed = s->completed; e.possible = s->possible; e.ep_reward = s->episode_reward;
    e.counter = s->rng_counter;
    return e;
}

__device__ __forceinline__ void store_env(slb_env_state *s, const EnvRegs &e) {
    s->agent_x = e.ax; s->agent_y = e.ay; s->orientation = e.orientation; s->game_over = e.game_over;
    s->num_steps = e.num_steps; s->level_id = e.level; s->episode_length = e.ep_len;
    s->episode_completed = e.ep_done; s->old_points = e.old_points; s->can_exit = e.can_exit;
    s->completed = e.completed; s->possible = e.possible; s->episode_reward = e.ep_reward;
    s->rng_counter = e.counter;
}

// reset from the pool on planes (safelife_env.py:188-198); every lane of the warp calls
__device__ __forceinline__ void reset_env_planes(const slb_pool &pool, int level, uint16_t *sm,
                                                 const Geo &g, uint32_t B[16], uint32_t G[16],
                                                 uint32_t &e0, EnvRegs &e) {
    const int lane = threadIdx.x & 31;
    const size_t cells = (size_t)g.h * g.w;
    const bool vec = fast_vec_ok(g.h, g.w, pool.board, pool.goals);
    load_planes(pool.board + level * cells, sm, g.h, g.w, vec, B);
    load_planes(pool.goals + level * cells, sm, g.h, g.w, vec, G);
    e0 = g.row ? pool.exit_mask[(size_t)level * g.h + lane] : 0u;
    RowScore rs = score_planes(B, G, g);
    const int cur = warp_sum(rs.current), pos = warp_sum(rs.possible);
    const int baseline = pool.baseline[level];
    e.completed = cur - baseline;
    e.possible = pos - baseline;
    e.can_exit = can_exit_rule(pool.min_performance[level], e.completed, e.possible);
    paint_exits(B, e0, e.can_exit);
    // _old_game_value is taken after update_exit_colors (safelife_env.py:191-192)
    rs = score_planes(B, G, g);
    e.old_points = warp_sum(rs.points);
    e.ax = pool.agent_x[level]; e.ay = pool.agent_y[level]; e.orientation = pool.orientation[level];
    e.game_over = 0; e.num_steps = 0; e.level = level; e.ep_len = 0; e.ep_done = 0; e.ep_reward = 0;
}

// ---------------------------------------------------------------------------------------------
// A12 fused: T env steps per launch, board + goals resident in registers between steps.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FAST_WARPS * 32)
env_rollout_fast_kernel(slb_pool pool, int n, uint16_t *board, uint16_t *goals, slb_env_state *state,
                        const uint8_t *__restrict__ actions, int t_steps, int time_limit,
                        uint64_t seed, uint64_t env_id0, int32_t *reward_out, uint8_t *done_out,
                        int auto_reset_stride, bool vec) {
    __shared__ __align__(16) uint16_t stage[FAST_WARPS][FAST_STAGE_U16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * FAST_WARPS + warp;
    if (i >= n) return;
    const Geo g = make_geo(pool.h, pool.w);
    const size_t cells = (size_t)pool.h * pool.w;
    uint16_t *sm = stage[warp];
    uint32_t B[16], G[16];
    load_planes(board + i * cells, sm, g.h, g.w, vec, B);
    load_planes(goals + i * cells, sm, g.h, g.w, vec, G);
    EnvRegs e = load_env(state + i);
    const uint64_t env = env_id0 + (uint64_t)i;
    uint64_t thr = spawn_threshold(pool.spawn_prob[e.level]);
    uint32_t e0 = g.row ? pool.exit_mask[(size_t)e.level * g.h + lane] : 0u;
    int baseline = pool.baseline[e.level];
    double min_perf = pool.min_performance[e.level];

    for (int t = 0; t < t_steps; t++) {
        const int action = actions[(size_t)t * n + i];
        int reward = action_planes(B, g, e.ax, e.ay, e.orientation, e.game_over, action, e.can_exit);
        // board first, then goals: the order the stream is consumed in (safelife_game.py:657-660)
        e.num_steps += 1;
        e.counter = physics_planes(B, g, thr, seed, env, e.counter);
        e.counter = physics_planes(G, g, thr, seed, env, e.counter);
        RowScore rs = score_planes(B, G, g);
        const int pts = warp_sum(rs.points), cur = warp_sum(rs.current), pos = warp_sum(rs.possible);
        reward += pts - e.old_points;
        e.old_points = pts;
        e.ep_len += 1;
        e.ep_reward += reward;
        e.completed = cur - baseline;
        e.possible = pos - baseline;
        e.can_exit = can_exit_rule(min_perf, e.completed, e.possible);
        paint_exits(B, e0, e.can_exit);
        const int done = (e.ep_len > time_limit) || e.game_over;
        e.ep_done = done;
        if (lane == 0) {
            if (reward_out) reward_out[(size_t)t * n + i] = reward;
            if (done_out) done_out[(size_t)t * n + i] = (uint8_t)done;
        }
        if (done && auto_reset_stride > 0) {
            reset_env_planes(pool, (e.level + auto_reset_stride) % pool.n_levels, sm, g, B, G, e0, e);
            thr = spawn_threshold(pool.spawn_prob[e.level]);
            baseline = pool.baseline[e.level];
            min_perf = pool.min_performance[e.level];
        }
    }
    store_planes(board + i * cells, sm, g.h, g.w, vec, B);
    store_planes(goals + i * cells, sm, g.h, g.w, vec, G);
    if (lane == 0) store_env(state + i, e);
}

inline void launch_env_rollout_fast(const slb_pool &pool, int n, uint16_t *board, uint16_t *goals,
                                    slb_env_state *state, const uint8_t *actions, int t_steps,
                                    int time_limit, uint64_t seed, uint64_t env0, int32_t *reward,
                                    uint8_t *done, int stride, cudaStream_t st) {
    int grid = (n + FAST_WARPS - 1) / FAST_WARPS;
    env_rollout_fast_kernel<<<grid, FAST_WARPS * 32, 0, st>>>(
        pool, n, board, goals, state, actions, t_steps, time_limit, seed, env0, reward, done, stride,
        fast_vec_ok(pool.h, pool.w, board, goals));
}

}  // namespace slb

#include "slb_fast_side.cuh"
