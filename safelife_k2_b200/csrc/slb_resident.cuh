// slb_resident.cuh — the RESIDENT env state of boards with 4 <= H, W <= 32 (the 26x26 benchmark
// levels) and the kernels that step it: one warp per env, board and goals as 16 bit-planes in
// registers (lane = row, bit = column), loaded from and stored to HBM *already bit-sliced*, so a
// step has no cell<->plane transposition and no shared-memory staging at all.
//
// Resident layout of one board (64*H bytes, see slb_resident_bytes): 4 plane groups x H rows x
// uint4; group g, row r is the uint4 at index g*H + r and holds the row's words of 4 planes:
//     group 0: alive(0) frozen(4) red(9) green(10)        group 2: agent(1) preserving(5) inhibiting(6) exit(8)
//     group 1: blue(11) destructible(3) pushable(2) pullable(15)   group 3: spawning(7) bits 12, 13, 14
// A warp reads a group with one fully coalesced 16-byte load per lane.  Groups 2 and 3 are only
// read / written when some board of the pool (or an imported board) has one of their bits: the
// pool keeps an OR of every cell it has seen (pool.plane_bits); a group outside it is all-zero in
// memory (buffers start zeroed) and stays so, because no rule creates a bit that no cell has,
// except births (bits 0, 3, 9-11: groups 0/1, always present) and exit painting (bits 4, 8, 9 —
// only where an exit cell, which has bit 8, exists).
//
// The uint16 [N,H,W] cell layout of the reference stays the format of every API boundary
// (slb_resident_import / slb_resident_export, slb_advance_board, the level pool, observations).
#pragma once
#include <cstdlib>

#include "slb_planes_env.cuh"

namespace slb {

#ifndef SLB_RES_WARPS
#define SLB_RES_WARPS 2
#endif
#ifndef SLB_RES_SLOTS
#define SLB_RES_SLOTS 28  // resident warps per SM the step kernels are compiled for (register budget)
#endif
constexpr int RES_WARPS = SLB_RES_WARPS;  // warps (= envs) per CTA
// the host-step kernel keeps four warps per CTA whatever the others use: its last CTA copies the whole
// reward / done block to the pinned host mirror, and half the CTAs means half the tickets
constexpr int RES_HOST_WARPS = 4;

inline bool fast_shape(int h, int w) { return h >= 4 && w >= 4 && h <= 32 && w <= 32; }

__host__ __device__ inline size_t resident_words(int h) { return (size_t)16 * h; }

// which groups a plane-bit mask touches (bit g set = group g present); groups 0/1 always
__host__ __device__ inline unsigned groups_of(uint32_t plane_bits) {
    unsigned gm = 3u;
    if (plane_bits & ((1u << 1) | (1u << 5) | (1u << 6) | (1u << 8))) gm |= 4u;
    if (plane_bits & ((1u << 7) | (7u << 12))) gm |= 8u;
    return gm;
}

__device__ __forceinline__ void load_res(const uint32_t *__restrict__ base, const Geo &g,
                                         unsigned groups, uint32_t P[16]) {
    const int lane = threadIdx.x & 31;
    const uint4 *b4 = reinterpret_cast<const uint4 *>(base);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    const uint4 q0 = g.row ? b4[lane] : z;
    const uint4 q1 = g.row ? b4[g.h + lane] : z;
    const uint4 q2 = (g.row && (groups & 4u)) ? b4[2 * g.h + lane] : z;
    const uint4 q3 = (g.row && (groups & 8u)) ? b4[3 * g.h + lane] : z;
    P[0] = q0.x; P[4] = q0.y; P[9] = q0.z; P[10] = q0.w;
    P[11] = q1.x; P[3] = q1.y; P[2] = q1.z; P[15] = q1.w;
    P[1] = q2.x; P[5] = q2.y; P[6] = q2.z; P[8] = q2.w;
    P[7] = q3.x; P[12] = q3.y; P[13] = q3.z; P[14] = q3.w;
}

__device__ __forceinline__ void store_res(uint32_t *__restrict__ base, const Geo &g, unsigned groups,
                                          const uint32_t P[16]) {
    const int lane = threadIdx.x & 31;
    if (!g.row) return;
    uint4 *b4 = reinterpret_cast<uint4 *>(base);
    b4[lane] = make_uint4(P[0], P[4], P[9], P[10]);
    b4[g.h + lane] = make_uint4(P[11], P[3], P[2], P[15]);
    if (groups & 4u) b4[2 * g.h + lane] = make_uint4(P[1], P[5], P[6], P[8]);
    if (groups & 8u) b4[3 * g.h + lane] = make_uint4(P[7], P[12], P[13], P[14]);
}

// ---------------------------------------------------------------------------------------------
// Env scalars in registers (warp-uniform).
// ---------------------------------------------------------------------------------------------
struct EnvRegs {
    int ax, ay, orientation, game_over, num_steps, level, ep_len, ep_done;
    int old_points, can_exit, completed, possible;
    long long ep_reward;
    uint64_t counter;
    double min_perf;
    uint32_t flags;
};

// The 80-byte slb_env_state record moves as 5 x 16 B.
__device__ __forceinline__ EnvRegs load_env(const slb_env_state *s) {
    const uint4 *q = reinterpret_cast<const uint4 *>(s);
    const uint4 a = q[0], b = q[1], c = q[2], d = q[3], f = q[4];
    EnvRegs e;
    e.ax = (int)a.x; e.ay = (int)a.y; e.orientation = (int)a.z; e.game_over = (int)a.w;
    e.num_steps = (int)b.x; e.level = (int)b.y; e.ep_len = (int)b.z; e.ep_done = (int)b.w;
    e.old_points = (int)c.x; e.can_exit = (int)c.y; e.completed = (int)c.z; e.possible = (int)c.w;
    e.ep_reward = (long long)(((uint64_t)d.y << 32) | d.x);
    e.counter = ((uint64_t)d.w << 32) | d.z;
    e.min_perf = __hiloint2double((int)f.y, (int)f.x);
    e.flags = f.z;
    return e;
}

__device__ __forceinline__ void store_env(slb_env_state *s, const EnvRegs &e) {
    uint4 *q = reinterpret_cast<uint4 *>(s);
    q[0] = make_uint4((uint32_t)e.ax, (uint32_t)e.ay, (uint32_t)e.orientation, (uint32_t)e.game_over);
    q[1] = make_uint4((uint32_t)e.num_steps, (uint32_t)e.level, (uint32_t)e.ep_len, (uint32_t)e.ep_done);
    q[2] = make_uint4((uint32_t)e.old_points, (uint32_t)e.can_exit, (uint32_t)e.completed, (uint32_t)e.possible);
    q[3] = make_uint4((uint32_t)(uint64_t)e.ep_reward, (uint32_t)((uint64_t)e.ep_reward >> 32),
                      (uint32_t)e.counter, (uint32_t)(e.counter >> 32));
    q[4] = make_uint4((uint32_t)__double2loint(e.min_perf), (uint32_t)__double2hiint(e.min_perf), e.flags, 0u);
}

struct LevelConsts {
    float spawn_prob;
    int baseline;
    uint32_t flags;
    uint32_t e0;  // this lane's row of the level's initial exit cells
};

// pool.level_consts: one 32-byte slb_level_consts record per level; a step needs its first 16 bytes
// (min_performance, the second half, is copied into the env's state by the reset)
__device__ __forceinline__ LevelConsts load_level(const slb_pool &pool, int level, const Geo &g) {
    LevelConsts c;
    const int lane = threadIdx.x & 31;
    const uint4 a = reinterpret_cast<const uint4 *>(pool.level_consts)[2 * level];
    c.spawn_prob = __uint_as_float(a.x);
    c.baseline = (int)a.y;
    c.flags = pool.external_writes ? 0u : a.z;
    c.e0 = g.row ? pool.exit_mask[level * g.h + lane] : 0u;
    return c;
}

__device__ __forceinline__ double load_level_min_perf(const slb_pool &pool, int level) {
    const uint2 b = *reinterpret_cast<const uint2 *>(reinterpret_cast<const uint4 *>(pool.level_consts) + 2 * level + 1);
    return __hiloint2double((int)b.y, (int)b.x);
}

// score -> completed / possible / can_exit (safelife_game.py:522-526, 601-631), then repaint the
// exits (:531-537).  Returns current_points of the board BEFORE repainting (exit cells are dead
// either way, so the value is the same after it).
__device__ __forceinline__ int score_and_paint(uint32_t B[16], const uint32_t G[16], const Geo &g,
                                               const LevelConsts &lc, EnvRegs &e, bool paint) {
    const RowScore rs = score_planes(B, G, g);
    const int pts = warp_sum(rs.points), cur = warp_sum(rs.current), pos = warp_sum(rs.possible);
    e.completed = cur - lc.baseline;
    e.possible = pos - lc.baseline;
    e.can_exit = can_exit_rule(e.min_perf, e.completed, e.possible);
    if (paint) {
        paint_exits(B, lc.e0, e.can_exit);
        e.flags &= ~SLB_ENV_RESCORE;
    }
    return pts;
}

// reset from the pool's resident level copies (safelife_env.py:188-198); every lane calls
__device__ __forceinline__ void reset_env_res(const slb_pool &pool, int level, const Geo &g,
                                              unsigned gb, unsigned gg, uint32_t B[16], uint32_t G[16],
                                              LevelConsts &lc, EnvRegs &e) {
    const size_t words = resident_words(g.h);
    load_res((const uint32_t *)pool.board_res + level * words, g, gb, B);
    load_res((const uint32_t *)pool.goals_res + level * words, g, gg, G);
    lc = load_level(pool, level, g);
    e.min_perf = load_level_min_perf(pool, level);  // deserialize restores the level's value (safelife_game.py:205)
    e.flags = 0;
    score_and_paint(B, G, g, lc, e, true);
    // _old_game_value is taken AFTER update_exit_colors (safelife_env.py:190-192): re-score, an
    // exit cell that was alive in the level record is dead once painted
    e.old_points = warp_sum(score_planes(B, G, g).points);
    e.ax = pool.agent_x[level]; e.ay = pool.agent_y[level]; e.orientation = pool.orientation[level];
    e.game_over = 0; e.num_steps = 0; e.level = level; e.ep_len = 0; e.ep_done = 0; e.ep_reward = 0;
    // e.counter deliberately untouched: the stream keeps running across episodes
}

// One env step on the planes in registers (safelife_env.py:157-186): action, physics of board then
// goals, reward, exit colours, counters, done.  Shared by the rollout kernel
// and the fused step + observation kernel.  Returns the step's reward; `done` as the env reports it.
// `gc` (may be NULL): SafeLifeEnv.global_counter on the device (safelife_env.py:169-173), GC_SLOTS pairs
// {num_steps, episodes_completed}; env i adds to pair i % GC_SLOTS so that the adds of one launch spread over
// GC_SLOTS addresses (fire-and-forget reductions; the host sums the pairs when it reads the counter).
constexpr int GC_SLOTS = 64;

template <unsigned GB, unsigned GG>
__device__ __forceinline__ int env_step_once(const slb_pool &pool, const Geo &g, uint32_t B[16], uint32_t G[16],
                                             LevelConsts &lc, EnvRegs &e, int action, uint64_t seed, uint64_t env,
                                             int time_limit, bool &g_dirty, int &done,
                                             unsigned long long *gc = nullptr, int slot = 0) {
    bool touched, life_touched, ch_b, ch_g = false;
    const int already = e.ep_done;
    int reward = action_planes(B, g, e.ax, e.ay, e.orientation, e.game_over, action, e.can_exit, touched,
                               life_touched);
    // board first, then goals: the order the stream is consumed in (safelife_game.py:657-660)
    e.num_steps += 1;
    e.counter = physics_planes(B, g, lc.spawn_prob, seed, env, e.counter, ch_b);
    if (!(lc.flags & SLB_LEVEL_GOALS_STATIC))  // a static goals board is its own successor
        e.counter = physics_planes(G, g, lc.spawn_prob, seed, env, e.counter, ch_g);
    g_dirty |= ch_g;
    // The score reads alive cells (their colour, frozen / movable bits) and goal colours.  When
    // the action moved no living cell and neither board changed, points, completed, possible
    // and can_exit are what they were, and the exit cells (frozen, never overwritten by an
    // action) still carry the colour painted for that can_exit.
    if (life_touched || ch_b || ch_g || pool.external_writes || (e.flags & SLB_ENV_RESCORE)) {
        const int pts = score_and_paint(B, G, g, lc, e, true);
        reward += pts - e.old_points;
        e.old_points = pts;
    }
    e.ep_len += 1;
    e.ep_reward += reward;
    done = (e.ep_len > time_limit) || e.game_over;
    e.ep_done = done;
    if (gc && !already && (threadIdx.x & 31) == 0) {  // only until the first done (safelife_env.py:169)
        atomicAdd(gc + 2 * (slot & (GC_SLOTS - 1)), 1ull);
        if (done) atomicAdd(gc + 2 * (slot & (GC_SLOTS - 1)) + 1, 1ull);
    }
    return reward;
}

// ---------------------------------------------------------------------------------------------
// A12 fused: T env steps per launch (T = 1 with a policy in the loop); board + goals stay in
// registers between steps.  grid = ceil(N / RES_WARPS).
// ---------------------------------------------------------------------------------------------
// GB / GG: compile-time plane-group masks of boards / goals (supersets of pool.plane_groups):
// planes of absent groups are constant zero, so their registers and all work on them fold away.
// HW: board height = width when known at compile time (26: the benchmark levels), 0 = run time.
// ONE: exactly one step per launch (the policy-in-the-loop case): no step loop, no action look-ahead.
template <unsigned GB, unsigned GG, int HW, bool ONE>
__global__ void __launch_bounds__(RES_WARPS * 32, SLB_RES_SLOTS / RES_WARPS)
res_env_rollout_kernel(slb_pool pool, int n, uint32_t *board, uint32_t *goals, slb_env_state *state,
                       const uint8_t *__restrict__ actions, int t_steps, int time_limit,
                       uint64_t seed, uint64_t env_id0, int32_t *reward_out, uint8_t *done_out,
                       int auto_reset_stride, int l2_prefetch, unsigned long long *gc) {
    const int lane = threadIdx.x & 31;
    int i = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    // A warp owns env i, i + W, i + 2 W, ... (W = warps of the grid).  By default the grid is one CTA per
    // RES_WARPS envs (the loop runs once); with SLB_RES_ROUNDS > 1 the launcher may give a warp several envs,
    // the next env's lines travelling from HBM to L2 while the current one is computed (see the launcher for
    // what that measured).
    const int stride = gridDim.x * RES_WARPS;
    // Programmatic dependent launch: let the next kernel on the stream begin launching now (its CTAs
    // take the SM slots ours free up); wait for the previous kernel's writes before the first read.
    asm volatile("griddepcontrol.launch_dependents;");
    if (i >= n) return;
    const Geo g = make_geo(HW ? HW : pool.h, HW ? HW : pool.w);
    const int nb = (__popc(GB) * 16 * g.h + 127) >> 7, ng = (__popc(GG) * 16 * g.h + 127) >> 7;
    if (l2_prefetch) {
        // While the previous kernel on the stream drains, pull this env's lines from HBM into L2
        // (one 128-byte line per lane).  L2 is the point of coherence, so a line prefetched before
        // the previous kernel's last writes land still reads back current after the wait below.
        const uint32_t off = (uint32_t)i * (16u * (uint32_t)g.h);
        if (lane < nb) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(board + off) + lane * 128));
        if (lane < ng) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(goals + off) + lane * 128));
        if (lane >= 30) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(state + i) + (lane - 30) * 64));
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for (; i < n; i += stride) {
        const uint32_t off = (uint32_t)i * (16u * (uint32_t)g.h);  // resident words per board: 16 * H
        uint32_t *bp = board + off, *gp = goals + off;
        if (i + stride < n) {  // the next env of this warp: HBM -> L2 while this one is stepped
            const uint32_t offn = (uint32_t)(i + stride) * (16u * (uint32_t)g.h);
            if (lane < nb) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(board + offn) + lane * 128));
            if (lane < ng) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(goals + offn) + lane * 128));
            if (lane >= 30) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(state + i + stride) + (lane - 30) * 64));
        }
        uint32_t B[16], G[16];
        EnvRegs e = load_env(state + i);
        int action = actions[i];
        load_res(bp, g, GB, B);
        load_res(gp, g, GG, G);
        LevelConsts lc = load_level(pool, e.level, g);
        const uint64_t env = env_id0 + (uint64_t)i;
        bool g_dirty = false;
        const int steps = ONE ? 1 : t_steps;

        for (int t = 0; t < steps; t++) {
            int next_action = 0;
            if (!ONE && t + 1 < steps) next_action = actions[(size_t)(t + 1) * n + i];
            int done;
            const int reward = env_step_once<GB, GG>(pool, g, B, G, lc, e, action, seed, env, time_limit, g_dirty, done, gc, i);
            if (lane == 0) {
                const size_t o = (size_t)t * n + i;
                if (reward_out) reward_out[o] = reward;
                if (done_out) done_out[o] = (uint8_t)done;
            }
            if (done && auto_reset_stride > 0) {
                reset_env_res(pool, (e.level + auto_reset_stride) % pool.n_levels, g, GB, GG, B, G, lc, e);
                g_dirty = true;
            }
            action = next_action;
        }
        store_res(bp, g, GB, B);
        if (g_dirty) store_res(gp, g, GG, G);
        if (lane == 0) store_env(state + i, e);
    }
}

// picks the smallest instantiated superset of the pool's plane groups
inline void launch_res_env_rollout(const slb_pool &pool, int n, uint32_t *board, uint32_t *goals,
                                   slb_env_state *state, const uint8_t *actions, int t_steps,
                                   int time_limit, uint64_t seed, uint64_t env0, int32_t *reward,
                                   uint8_t *done, int stride, cudaStream_t st,
                                   unsigned long long *gc = nullptr) {
    int grid = (n + RES_WARPS - 1) / RES_WARPS;
    // SLB_RES_ROUNDS = R > 1 (default 1 = off): a batch of up to R resident waves is launched as ONE balanced
    // wave of CTAs whose warps take several envs in turn (see the kernel).  Measured on B200 (round 2,
    // profiles/r02): +7 % on append-still at 8 192 envs (uniform cost per env), but -10 % on the mixed archives
    // at 8 192 / 16 384 envs and on prune-spawn at 16 384 — a static share per warp loses the hardware CTA
    // scheduler's dynamic balancing when the cost per env varies (spawners, oscillators) — hence off.
    static const int max_rounds = []() { const char *v = getenv("SLB_RES_ROUNDS"); return v ? atoi(v) : 1; }();
    static const int wave = []() {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        return sms * (SLB_RES_SLOTS / RES_WARPS);
    }();
    if (grid > wave && grid <= max_rounds * wave) {
        const int rounds = (grid + wave - 1) / wave;
        grid = (grid + rounds - 1) / rounds;
    }
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    // launched with programmatic stream serialization (see the griddepcontrol pair in the kernel)
    static const bool pdl = []() { const char *v = getenv("SLB_PDL"); return !v || v[0] != '0'; }();
    static const bool pf = []() { const char *v = getenv("SLB_L2_PREFETCH"); return !v || v[0] != '0'; }();
    const int l2_prefetch = pdl && pf;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(RES_WARPS * 32);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
#define SLB_ROLL3(GBV, GGV, HWV, ONEV)                                                               \
    cudaLaunchKernelEx(&cfg, res_env_rollout_kernel<GBV, GGV, HWV, ONEV>, pool, n, board, goals,      \
                       state, actions, t_steps, time_limit, seed, env0, reward, done, stride,         \
                       l2_prefetch, gc)
#define SLB_ROLL2(GBV, GGV, HWV)                                                                     \
    do {                                                                                             \
        if (t_steps == 1) SLB_ROLL3(GBV, GGV, HWV, true);                                            \
        else SLB_ROLL3(GBV, GGV, HWV, false);                                                        \
    } while (0)
#define SLB_ROLL(GBV, GGV)                                                                           \
    do {                                                                                             \
        if (pool.h == 26 && pool.w == 26) SLB_ROLL2(GBV, GGV, 26);                                   \
        else SLB_ROLL2(GBV, GGV, 0);                                                                 \
    } while (0)
    // SLB_RES_CHUNK = C > 0 (one-step launches only): a batch of more than C envs goes out as ceil(n / C)
    // launches of at most C envs each (C = -1: one resident wave), back to back on the stream with
    // programmatic serialization — every launch is then the one-wave case whose loads the previous launch's
    // tail hides (L2 prefetch ahead of the dependency wait), instead of two or three lock-step waves that
    // each pay the whole load -> compute -> store latency chain.  An env's results do not depend on how the
    // batch is cut (its stream is keyed by env0 + index).
    static const int chunk_cfg = []() { const char *v = getenv("SLB_RES_CHUNK"); return v ? atoi(v) : 0; }();
    const int chunk = (t_steps == 1 && chunk_cfg != 0) ? (chunk_cfg < 0 ? wave * RES_WARPS : chunk_cfg) : 0;
    const int n_all = n;
    uint32_t *const board0 = board, *const goals0 = goals;
    slb_env_state *const state0 = state;
    const uint8_t *const actions0 = actions;
    int32_t *const reward0 = reward;
    uint8_t *const done0 = done;
    const uint64_t env00 = env0;
    for (int lo = 0; lo < n_all; lo += (chunk > 0 ? chunk : n_all)) {
        n = (chunk > 0 && n_all - lo > chunk) ? chunk : n_all - lo;
        const size_t woff = (size_t)lo * 16u * (size_t)pool.h;
        board = board0 + woff; goals = goals0 + woff; state = state0 + lo; actions = actions0 + lo;
        reward = reward0 ? reward0 + lo : nullptr; done = done0 ? done0 + lo : nullptr;
        env0 = env00 + (uint64_t)lo;
        if (chunk > 0) cfg.gridDim = dim3((n + RES_WARPS - 1) / RES_WARPS);
        if (!(gg & ~3u) && !(gb & ~7u)) SLB_ROLL(7u, 3u);        // no spawners anywhere, plain goals
        else if (!(gg & ~3u)) SLB_ROLL(15u, 3u);                 // spawners on the boards, plain goals
        else SLB_ROLL(15u, 15u);
    }
#undef SLB_ROLL3
#undef SLB_ROLL2
#undef SLB_ROLL
}

// ---------------------------------------------------------------------------------------------
// The host-buffer step without a DMA copy on either side (slb_env_step_host*, n <= 4096): the actions
// travel in the launch parameters (4 KB by value), the step writes reward / done to a device staging
// block, and the last CTA to finish (ticket counter) copies the whole block to the caller's mapped
// pinned mirror with 16-byte stores, fences system-wide and publishes the step's sequence number
// there; the host waits on that word instead of on the stream.
// ---------------------------------------------------------------------------------------------
constexpr int HOST_STEP_MAX_ENVS = 4096;
struct ActBlock {
    uint8_t a[HOST_STEP_MAX_ENVS];
};
struct HostTail {
    int32_t *d_reward;   // device staging [reward 4n | done n], 16-byte aligned, padded to 16
    uint8_t *d_done;
    unsigned *counter;   // device, zero between launches
    uint8_t *h_block;    // mapped pinned host mirror of the staging block
    uint32_t *h_seq;     // mapped pinned host word
    uint32_t seq;
};

template <unsigned GB, unsigned GG, int HW>
__global__ void __launch_bounds__(RES_HOST_WARPS * 32, SLB_RES_SLOTS / RES_HOST_WARPS)
res_env_step_host_kernel(slb_pool pool, int n, uint32_t *board, uint32_t *goals, slb_env_state *state,
                         const __grid_constant__ ActBlock acts, int time_limit, uint64_t seed,
                         uint64_t env_id0, int auto_reset_stride, HostTail ht) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_HOST_WARPS + (threadIdx.x >> 5);
    if (i < n) {
        const Geo g = make_geo(HW ? HW : pool.h, HW ? HW : pool.w);
        const uint32_t off = (uint32_t)i * (16u * (uint32_t)g.h);
        uint32_t *bp = board + off, *gp = goals + off;
        uint32_t B[16], G[16];
        EnvRegs e = load_env(state + i);
        const int action = acts.a[i];
        load_res(bp, g, GB, B);
        load_res(gp, g, GG, G);
        LevelConsts lc = load_level(pool, e.level, g);
        const uint64_t env = env_id0 + (uint64_t)i;
        bool g_dirty = false;
        int done;
        const int reward = env_step_once<GB, GG>(pool, g, B, G, lc, e, action, seed, env, time_limit, g_dirty, done);
        if (lane == 0) {
            ht.d_reward[i] = reward;
            ht.d_done[i] = (uint8_t)done;
        }
        if (done && auto_reset_stride > 0) {
            reset_env_res(pool, (e.level + auto_reset_stride) % pool.n_levels, g, GB, GG, B, G, lc, e);
            g_dirty = true;
        }
        store_res(bp, g, GB, B);
        if (g_dirty) store_res(gp, g, GG, G);
        if (lane == 0) store_env(state + i, e);
    }
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(ht.counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        const int nvec = (5 * n + 15) >> 4;
        const uint4 *src = reinterpret_cast<const uint4 *>(ht.d_reward);
        uint4 *dst = reinterpret_cast<uint4 *>(ht.h_block);
        // (four loads in flight per thread: the copy is this CTA's alone and every load is an L2 round trip)
        const int T = blockDim.x;
        for (int j = threadIdx.x; j < nvec; j += 4 * T) {
            uint4 v[4];
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (j + q * T < nvec) v[q] = __ldcg(src + j + q * T);
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (j + q * T < nvec) dst[j + q * T] = v[q];
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            *ht.counter = 0;
            __threadfence_system();  // release: every thread's stores (fenced above) before the sequence number
            *reinterpret_cast<volatile uint32_t *>(ht.h_seq) = ht.seq;
        }
    }
}

inline void launch_res_env_step_host(const slb_pool &pool, int n, uint32_t *board, uint32_t *goals,
                                     slb_env_state *state, const ActBlock &acts, int time_limit,
                                     uint64_t seed, uint64_t env0, int stride, const HostTail &ht,
                                     cudaStream_t st) {
    const int grid = (n + RES_HOST_WARPS - 1) / RES_HOST_WARPS;
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
#define SLB_HS2(GBV, GGV, HWV)                                                                          \
    res_env_step_host_kernel<GBV, GGV, HWV><<<grid, RES_HOST_WARPS * 32, 0, st>>>(                           \
        pool, n, board, goals, state, acts, time_limit, seed, env0, stride, ht)
#define SLB_HS(GBV, GGV)                                                                                \
    do {                                                                                                \
        if (pool.h == 26 && pool.w == 26) SLB_HS2(GBV, GGV, 26);                                        \
        else SLB_HS2(GBV, GGV, 0);                                                                      \
    } while (0)
    if (!(gg & ~3u) && !(gb & ~7u)) SLB_HS(7u, 3u);
    else if (!(gg & ~3u)) SLB_HS(15u, 3u);
    else SLB_HS(15u, 15u);
#undef SLB_HS2
#undef SLB_HS
}

// pool prepare, after the resident copies exist: which levels have goals that never change
// (a fixed point of advance_board without spawners; safelife_game.py:657-660 advances goals too)
__global__ void __launch_bounds__(RES_WARPS * 32)
res_level_flags_kernel(slb_pool pool) {
    const int lane = threadIdx.x & 31;
    const int l = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    if (l >= pool.n_levels) return;
    const Geo g = make_geo(pool.h, pool.w);
    uint32_t G[16];
    load_res((const uint32_t *)pool.goals_res + l * resident_words(g.h), g, 15u, G);
    const bool spawners = __any_sync(FULL, G[7] != 0);
    bool changed;
    physics_planes(G, g, 0.0f, 0, 0, 0, changed);
    if (lane == 0 && !spawners && !changed)
        reinterpret_cast<slb_level_consts *>(pool.level_consts)[l].flags |= SLB_LEVEL_GOALS_STATIC;
}

// ---------------------------------------------------------------------------------------------
// The single stages as their own launches (the SafeLifeGame methods call them one by one).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RES_WARPS * 32)
res_reset_kernel(slb_pool pool, int n, const int32_t *__restrict__ level_ids,
                 const uint8_t *__restrict__ mask, uint32_t *board, uint32_t *goals,
                 slb_env_state *state) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    if (i >= n || (mask && !mask[i])) return;
    const Geo g = make_geo(pool.h, pool.w);
    const size_t words = resident_words(g.h);
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    uint32_t B[16], G[16];
    LevelConsts lc;
    EnvRegs e = load_env(state + i);
    reset_env_res(pool, level_ids[i], g, gb, gg, B, G, lc, e);
    store_res(board + i * words, g, gb, B);
    store_res(goals + i * words, g, gg, G);
    if (lane == 0) store_env(state + i, e);
}

// execute_action alone (safelife_game.py:347-393)
__global__ void __launch_bounds__(RES_WARPS * 32)
res_action_kernel(slb_pool pool, int n, uint32_t *board, slb_env_state *state,
                  const uint8_t *__restrict__ actions, int32_t *reward_out) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    if (i >= n) return;
    const Geo g = make_geo(pool.h, pool.w);
    const unsigned gb = pool.plane_groups[0];
    uint32_t *bp = board + i * resident_words(g.h);
    uint32_t B[16];
    load_res(bp, g, gb, B);
    EnvRegs e = load_env(state + i);
    bool touched;
    const int reward = action_planes(B, g, e.ax, e.ay, e.orientation, e.game_over, actions[i], e.can_exit, touched);
    if (touched) store_res(bp, g, gb, B);
    if (lane == 0) {
        store_env(state + i, e);
        if (reward_out) reward_out[i] = reward;
    }
}

// SafeLifeGame.advance_board alone (safelife_game.py:657-660): num_steps += 1, board then goals
__global__ void __launch_bounds__(RES_WARPS * 32)
res_advance_kernel(slb_pool pool, int n, uint32_t *board, uint32_t *goals, slb_env_state *state,
                   uint64_t seed, uint64_t env_id0) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    if (i >= n) return;
    const Geo g = make_geo(pool.h, pool.w);
    const size_t words = resident_words(g.h);
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    uint32_t P[16];
    EnvRegs e = load_env(state + i);
    const float p = pool.spawn_prob[e.level];
    const uint64_t env = env_id0 + (uint64_t)i;
    bool ch;
    load_res(board + i * words, g, gb, P);
    e.counter = physics_planes(P, g, p, seed, env, e.counter, ch);
    if (ch) store_res(board + i * words, g, gb, P);
    load_res(goals + i * words, g, gg, P);
    e.counter = physics_planes(P, g, p, seed, env, e.counter, ch);
    if (ch) store_res(goals + i * words, g, gg, P);
    e.num_steps += 1;
    if (lane == 0) store_env(state + i, e);
}

// current_points / performance_ratio / can_exit (read-only), or update_exit_colors (PAINT)
template <bool PAINT>
__global__ void __launch_bounds__(RES_WARPS * 32)
res_score_kernel(slb_pool pool, int n, uint32_t *board, const uint32_t *goals, slb_env_state *state,
                 int32_t *points, int32_t *completed, int32_t *possible, uint8_t *can_exit) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    if (i >= n) return;
    const Geo g = make_geo(pool.h, pool.w);
    const size_t words = resident_words(g.h);
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    uint32_t B[16], G[16];
    load_res(board + i * words, g, gb, B);
    load_res(goals + i * words, g, gg, G);
    EnvRegs e = load_env(state + i);
    const LevelConsts lc = load_level(pool, e.level, g);
    const int pts = score_and_paint(B, G, g, lc, e, PAINT);
    if (PAINT) {
        store_res(board + i * words, g, gb, B);
        if (lane == 0) store_env(state + i, e);
    } else if (lane == 0) {
        if (points) points[i] = pts;
        if (completed) completed[i] = e.completed;
        if (possible) possible[i] = e.possible;
        if (can_exit) can_exit[i] = (uint8_t)e.can_exit;
    }
}

// SimpleSideEffectPenalty's count of side-effect cells per env (env_wrappers.py:319-346)
__global__ void __launch_bounds__(RES_WARPS * 32)
res_penalty_kernel(slb_pool pool, int n, const uint32_t *__restrict__ board,
                   const uint32_t *__restrict__ goals, const slb_env_state *__restrict__ state,
                   int32_t *count) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + (threadIdx.x >> 5);
    if (i >= n) return;
    const Geo g = make_geo(pool.h, pool.w);
    const size_t words = resident_words(g.h);
    const int level = state[i].level_id;
    uint32_t B[16], S[16], G[16];
    load_res(board + i * words, g, pool.plane_groups[0], B);
    load_res((const uint32_t *)pool.board_res + level * words, g, 15u, S);
    load_res(goals + i * words, g, 3u, G);
    const uint32_t e0 = g.row ? pool.exit_mask[(size_t)level * g.h + lane] : 0u;
    const int c = warp_sum(__popc(side_effect_word(B, S, G, e0, g.row ? g.wmask : 0u)));
    if (lane == 0) count[i] = c;
}

// ---------------------------------------------------------------------------------------------
// cells <-> resident planes (API boundary).  The import ORs every cell it sees into *plane_bits.
// ---------------------------------------------------------------------------------------------
constexpr int RES_STAGE_U16 = 32 * 34;  // per-warp staging area (H <= 32 rows, stride <= 34)

__global__ void __launch_bounds__(RES_WARPS * 32)
res_import_kernel(const uint16_t *__restrict__ cells, uint32_t *planes, int n, int h, int w,
                  uint32_t *plane_bits, bool vec) {
    __shared__ __align__(16) uint16_t stage[RES_WARPS][RES_STAGE_U16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + warp;
    if (i >= n) return;
    const Geo g = make_geo(h, w);
    uint32_t P[16];
    load_planes(cells + (size_t)i * h * w, stage[warp], h, w, vec, P);
    store_res(planes + i * resident_words(h), g, 15u, P);
    if (plane_bits) {
        uint32_t bits = 0;
#pragma unroll
        for (int b = 0; b < 16; b++) bits |= (P[b] != 0u) ? (1u << b) : 0u;
        bits = __reduce_or_sync(FULL, bits);
        if (lane == 0 && (bits & ~*plane_bits)) atomicOr(plane_bits, bits);
    }
}

__global__ void __launch_bounds__(RES_WARPS * 32)
res_export_kernel(const uint32_t *__restrict__ planes, uint16_t *cells, int n, int h, int w, bool vec) {
    __shared__ __align__(16) uint16_t stage[RES_WARPS][RES_STAGE_U16];
    const int warp = threadIdx.x >> 5;
    const int i = blockIdx.x * RES_WARPS + warp;
    if (i >= n) return;
    const Geo g = make_geo(h, w);
    uint32_t P[16];
    load_res(planes + i * resident_words(h), g, 15u, P);
    store_planes(cells + (size_t)i * h * w, stage[warp], h, w, vec, P);
}

// ---------------------------------------------------------------------------------------------
// get_obs (safelife_env.py:125-155) + recenter_view (helper_utils.py:41-74) from resident planes.
// out[i][vy][vx] = (board + ((goals & colours, white removed) << 3)) cropped toroidally around the
// agent; exit cells of the INITIAL board that fall outside the view are clamped onto its rim, in
// raster order of the exit cells (later ones overwrite earlier ones, like the fancy assignment).
// The combined board is formed on the planes, transposed once into a per-warp shared image of
// uint16 cells, and gathered from there (any view size, also views larger than the board).
// ---------------------------------------------------------------------------------------------
// python's a % m for the view arithmetic, where a lies within a few multiples of m around [0, m)
// (agent position +- view extent): one conditional add / subtract; the integer division is the rare path
__device__ __forceinline__ int pymod_w(int a, int m) {
    if (a < 0) a += m;
    else if (a >= m) a -= m;
    if ((unsigned)a >= (unsigned)m) {  // only views larger than the board get here
        a %= m;
        if (a < 0) a += m;
    }
    return a;
}

struct ObsChannels { uint8_t ch[16]; };

// the low 8 bits of `bits` as eight uint16 values 0 / 1 (element j = bit j), one 16-byte store
__device__ __forceinline__ uint4 spread8(uint32_t bits) {
    const uint32_t u = (bits & 0xffu) * 0x8001u;  // bit j at position j and at position j + 15
    return make_uint4(u & 0x00010001u, (u >> 2) & 0x00010001u, (u >> 4) & 0x00010001u, (u >> 6) & 0x00010001u);
}

template <bool IDENT>
__device__ __forceinline__ uint32_t obs_remap(uint32_t w, const ObsChannels &t, int c_n) {
    if (IDENT) return w & ((1u << c_n) - 1u);
    uint32_t r = 0;
#pragma unroll
    for (int c = 0; c < 16; c++)
        if (c < c_n) r |= ((w >> t.ch[c]) & 1u) << c;
    return r;
}

constexpr int RES_VIEW_U16 = 1089;  // largest view the fused channel output stages (33 x 33)

// Observation of safelife_env.py:125-155 for one env per warp.  c_n == 0: packed words, obs is
// uint16 [N, vh, vw].  c_n > 0 (vh * vw <= RES_VIEW_U16): the channel split is fused in, obs is
// uint16 [N, vh, vw, c_n] and is written with 16-byte stores (obs must be 16-byte aligned).
// obs_emit: the part after the planes are in registers (O = board planes, overwritten; G = goals), shared
// by the stand-alone kernel and the fused step + observation kernel.  x0, y0: agent; e0: this lane's row of
// the level's initial exit cells; sm / view: the warp's staging areas.
template <bool CHANNELS, bool IDENT>
__device__ __forceinline__ void obs_emit(uint32_t O[16], const uint32_t G[16], const Geo &g, int i, int x0, int y0,
                                         uint32_t e0, int vh, int vw, int remove_white, const ObsChannels &tab,
                                         int c_n, uint16_t *obs, uint16_t *sm, uint16_t *view) {
    const int lane = threadIdx.x & 31;
    const int h = g.h, w = g.w, S = row_stride(w);
    {
        uint32_t cr = G[9], cg = G[10], cb = G[11];
        if (remove_white) {
            const uint32_t nw = ~(cr & cg & cb);
            cr &= nw; cg &= nw; cb &= nw;
        }
        // uint16 add of (goal colours << 3) onto the board word: ripple through bits 12..15
        uint32_t c = O[12] & cr;
        O[12] ^= cr;
        uint32_t t = O[13] ^ cg;
        const uint32_t c2 = (O[13] & cg) | (t & c);
        O[13] = t ^ c;
        t = O[14] ^ cb;
        c = (O[14] & cb) | (t & c2);
        O[14] = t ^ c2;
        O[15] ^= c;
    }
    {
        uint32_t x[16];
        planes_to_cells(O, x);
        __syncwarp();
        uint32_t *row = reinterpret_cast<uint32_t *>(sm) + lane * (S >> 1);
#pragma unroll
        for (int j = 0; j < 16; j++)
            if (lane < h && 2 * j < w) row[j] = x[j];
        __syncwarp();
    }
    const int x1 = x0 - vw / 2, y1 = y0 - vh / 2;
    const int V = vh * vw;
    // the view goes to global memory (packed words) or to shared memory in channel order
    uint16_t *o = CHANNELS ? view : obs + (size_t)i * V;
    // lane = view column (blocks of 32 for wider views); source column once per block, source row
    // stepped with a wrap per view row
    for (int vx0 = 0; vx0 < vw; vx0 += 32) {
        const int vx = vx0 + lane;
        const int sx = pymod_w(x1 + vx, w);
        int sy = pymod_w(y1, h);
        uint16_t *orow = o + vx;
        for (int vy = 0; vy < vh; vy++) {
            if (vx < vw) {
                const uint32_t v = sm[sy * S + sx];
                *orow = CHANNELS ? (uint16_t)obs_remap<IDENT>(v, tab, c_n) : (uint16_t)v;
            }
            orow += vw;
            if (++sy == h) sy = 0;
        }
    }
    __syncwarp();
    // move_to_perimeter: every initial exit cell, raster order
    unsigned rows = __ballot_sync(FULL, e0 != 0u);
    while (rows) {
        const int y = __ffs(rows) - 1;
        rows &= rows - 1;
        uint32_t m = __shfl_sync(FULL, e0, y);
        while (m) {
            const int x = __ffs(m) - 1;
            m &= m - 1;
            int jy = pymod_w(y - y0 + h / 2, h) - h / 2;
            int jx = pymod_w(x - x0 + w / 2, w) - w / 2;
            jy = min(max(jy + vh / 2, 0), vh - 1);
            jx = min(max(jx + vw / 2, 0), vw - 1);
            if (lane == 0) {
                const uint32_t v = sm[y * S + x];
                o[jy * vw + jx] = CHANNELS ? (uint16_t)obs_remap<IDENT>(v, tab, c_n) : (uint16_t)v;
            }
        }
    }
    if (!CHANNELS) return;
    __syncwarp();
    // channel split: element e of this env (e = m * c_n + c) is bit c of o[m].  The env's T = V * c_n
    // output elements start at global element E0 = i * T, which is only 2-byte aligned in general:
    // single stores up to the first 16-byte boundary, 16-byte stores, single stores for the rest.
    const uint32_t T = (uint32_t)V * (uint32_t)c_n;
    const size_t E0 = (size_t)i * T;
    const uint32_t head = min((uint32_t)((8 - (E0 & 7)) & 7), T);  // elements before the boundary
    const uint32_t chunks = (T - head) >> 3;
    const uint32_t tail0 = head + (chunks << 3);
    uint16_t *out = obs + E0;
    for (uint32_t e = lane; e < head + (T - tail0); e += 32) {
        const uint32_t l = e < head ? e : tail0 + (e - head);
        const uint32_t m = l / (uint32_t)c_n;
        out[l] = (uint16_t)((o[m] >> (l - m * c_n)) & 1u);
    }
    uint4 *out4 = reinterpret_cast<uint4 *>(out + head);
    uint32_t q0 = 0;
    if (IDENT && c_n == 15) {
        // The default channels (bits 0..14 in order): 8 cells are exactly 120 elements = 15 chunks.  Phase 1:
        // a lane takes the 8 cells of group gq (one 16-byte shared load) plus the next one, packs their
        // 15-bit words into a 128-bit string, shifts it by `head` (chunks are aligned to global 16-byte
        // boundaries, cells to the env's first element) and parks the 15 bytes in a 16-byte slot of the
        // staging area (the cell image there is dead by now).  Phase 2: lanes take consecutive chunks —
        // one byte each — so that the 16-byte global stores of a warp are contiguous.
        const uint32_t groups = (T - head) / 120u;
        uint4 *slots = reinterpret_cast<uint4 *>(sm);
        __syncwarp();
        for (uint32_t gq = lane; gq < groups; gq += 32) {
            const uint4 wa = *reinterpret_cast<const uint4 *>(o + 8 * gq);
            const uint32_t w0 = wa.x & 0xffffu, w1 = wa.x >> 16, w2 = wa.y & 0xffffu, w3 = wa.y >> 16;
            const uint32_t w4 = wa.z & 0xffffu, w5 = wa.z >> 16, w6 = wa.w & 0xffffu, w7 = wa.w >> 16;
            const uint32_t w8 = o[8 * gq + 8];
            const uint32_t p0 = w0 | (w1 << 15) | (w2 << 30);
            const uint32_t p1 = (w2 >> 2) | (w3 << 13) | (w4 << 28);
            const uint32_t p2 = (w4 >> 4) | (w5 << 11) | (w6 << 26);
            const uint32_t p3 = (w6 >> 6) | (w7 << 9) | (w8 << 24);
            slots[gq] = make_uint4(__funnelshift_r(p0, p1, head), __funnelshift_r(p1, p2, head),
                                   __funnelshift_r(p2, p3, head), p3 >> head);
        }
        __syncwarp();
        q0 = 15u * groups;
        const uint8_t *bytes = reinterpret_cast<const uint8_t *>(sm);
        uint32_t gg = (uint32_t)lane / 15u, jj = (uint32_t)lane - 15u * gg;  // chunk q = 15 gg + jj
        for (uint32_t q = lane; q < q0; q += 32) {
            out4[q] = spread8(bytes[gg * 16 + jj]);
            gg += 2; jj += 2;  // 32 = 2 * 15 + 2
            if (jj >= 15) { jj -= 15; gg++; }
        }
    }
    // chunk q starts at element head + 8 q: (cell m, channel c0) advance by 256 elements per iteration
    const uint32_t l0 = head + ((q0 + (uint32_t)lane) << 3);
    uint32_t m = l0 / (uint32_t)c_n;
    int c0 = (int)(l0 - m * c_n);
    const uint32_t dm = 256u / (uint32_t)c_n;
    const int dc = (int)(256u - dm * c_n);
    for (uint32_t q = q0 + lane; q < chunks; q += 32) {
        uint32_t bits = (uint32_t)o[m] >> c0, mm = m;
        for (int have = c_n - c0; have < 8; have += c_n) bits |= (uint32_t)o[++mm] << have;
        out4[q] = spread8(bits);
        m += dm;
        c0 += dc;
        if (c0 >= c_n) { c0 -= c_n; m++; }
    }
}

constexpr int RES_VIEW_ROW = 1104;  // RES_VIEW_U16 + 1 cells, rows 16-byte aligned

template <bool CHANNELS, bool IDENT>
__global__ void __launch_bounds__(RES_WARPS * 32)
res_obs_kernel(slb_pool pool, int n, const uint32_t *__restrict__ board,
               const uint32_t *__restrict__ goals, const slb_env_state *__restrict__ state, int vh,
               int vw, int remove_white, ObsChannels tab, int c_n, uint16_t *obs) {
    __shared__ __align__(16) uint16_t stage[RES_WARPS][RES_STAGE_U16];
    __shared__ __align__(16) uint16_t view_sm[CHANNELS ? RES_WARPS : 1][CHANNELS ? RES_VIEW_ROW : 8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * RES_WARPS + warp;
    if (i >= n) return;
    const Geo g = make_geo(pool.h, pool.w);
    const size_t words = resident_words(g.h);
    uint32_t O[16], G[16];
    load_res(board + i * words, g, pool.plane_groups[0], O);
    load_res(goals + i * words, g, 3u, G);  // only the colour planes (groups 0/1) are needed
    const slb_env_state *st = state + i;
    const int x0 = st->agent_x, y0 = st->agent_y, level = st->level_id;
    const uint32_t e0 = g.row ? pool.exit_mask[(size_t)level * g.h + lane] : 0u;
    obs_emit<CHANNELS, IDENT>(O, G, g, i, x0, y0, e0, vh, vw, remove_white, tab, c_n, obs, stage[warp],
                              view_sm[CHANNELS ? warp : 0]);
}

// ---------------------------------------------------------------------------------------------
// The policy-in-the-loop pair in one launch: one env step (the ONE path of res_env_rollout_kernel)
// followed by the observation of the state it produced (after an in-kernel auto-reset: of the new
// episode), straight from the planes in registers.  Saves the second launch and the observation
// kernel's reload of board and goals.
// ---------------------------------------------------------------------------------------------
struct ObsArgs {
    int vh, vw, remove_white, c_n;
    ObsChannels tab;
    uint16_t *obs;
};

template <unsigned GB, unsigned GG, int HW, bool CHANNELS, bool IDENT>
__global__ void __launch_bounds__(RES_WARPS * 32, SLB_RES_SLOTS / RES_WARPS)
res_env_step_obs_kernel(slb_pool pool, int n, uint32_t *board, uint32_t *goals, slb_env_state *state,
                        const uint8_t *__restrict__ actions, int time_limit, uint64_t seed, uint64_t env_id0,
                        int32_t *reward_out, uint8_t *done_out, int auto_reset_stride, int l2_prefetch,
                        ObsArgs oa, unsigned long long *gc) {
    __shared__ __align__(16) uint16_t stage[RES_WARPS][RES_STAGE_U16];
    __shared__ __align__(16) uint16_t view_sm[CHANNELS ? RES_WARPS : 1][CHANNELS ? RES_VIEW_ROW : 8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i = blockIdx.x * RES_WARPS + warp;
    asm volatile("griddepcontrol.launch_dependents;");
    if (i >= n) return;
    const Geo g = make_geo(HW ? HW : pool.h, HW ? HW : pool.w);
    const uint32_t off = (uint32_t)i * (16u * (uint32_t)g.h);
    uint32_t *bp = board + off, *gp = goals + off;
    uint32_t B[16], G[16];
    if (l2_prefetch) {
        const int nb = (__popc(GB) * 16 * g.h + 127) >> 7, ng = (__popc(GG) * 16 * g.h + 127) >> 7;
        if (lane < nb) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)bp + lane * 128));
        if (lane < ng) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)gp + lane * 128));
        if (lane >= 30) asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)(state + i) + (lane - 30) * 64));
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    EnvRegs e = load_env(state + i);
    const int action = actions[i];
    load_res(bp, g, GB, B);
    load_res(gp, g, GG, G);
    LevelConsts lc = load_level(pool, e.level, g);
    const uint64_t env = env_id0 + (uint64_t)i;
    bool g_dirty = false;
    int done;
    const int reward = env_step_once<GB, GG>(pool, g, B, G, lc, e, action, seed, env, time_limit, g_dirty, done, gc, i);
    if (lane == 0) {
        if (reward_out) reward_out[i] = reward;
        if (done_out) done_out[i] = (uint8_t)done;
    }
    if (done && auto_reset_stride > 0) {
        reset_env_res(pool, (e.level + auto_reset_stride) % pool.n_levels, g, GB, GG, B, G, lc, e);
        g_dirty = true;
    }
    store_res(bp, g, GB, B);
    if (g_dirty) store_res(gp, g, GG, G);
    if (lane == 0) store_env(state + i, e);
    obs_emit<CHANNELS, IDENT>(B, G, g, i, e.ax, e.ay, lc.e0, oa.vh, oa.vw, oa.remove_white, oa.tab, oa.c_n, oa.obs,
                              stage[warp], view_sm[CHANNELS ? warp : 0]);
}

inline cudaError_t launch_res_env_step_obs(const slb_pool &pool, int n, uint32_t *board, uint32_t *goals,
                                           slb_env_state *state, const uint8_t *actions, int time_limit,
                                           uint64_t seed, uint64_t env0, int32_t *reward, uint8_t *done, int stride,
                                           const ObsArgs &oa, bool ident, cudaStream_t st,
                                           unsigned long long *gc = nullptr) {
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    static const bool pdl = []() { const char *v = getenv("SLB_PDL"); return !v || v[0] != '0'; }();
    static const bool pf = []() { const char *v = getenv("SLB_L2_PREFETCH"); return !v || v[0] != '0'; }();
    const int l2_prefetch = pdl && pf;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((n + RES_WARPS - 1) / RES_WARPS);
    cfg.blockDim = dim3(RES_WARPS * 32);
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const bool chan = oa.c_n > 0;
#define SLB_SO3(GBV, GGV, HWV, CH, ID)                                                                        \
    return cudaLaunchKernelEx(&cfg, res_env_step_obs_kernel<GBV, GGV, HWV, CH, ID>, pool, n, board, goals,     \
                              state, actions, time_limit, seed, env0, reward, done, stride, l2_prefetch, oa, gc)
#define SLB_SO2(GBV, GGV, HWV)                                                                                \
    do {                                                                                                      \
        if (!chan) SLB_SO3(GBV, GGV, HWV, false, true);                                                       \
        else if (ident) SLB_SO3(GBV, GGV, HWV, true, true);                                                   \
        else SLB_SO3(GBV, GGV, HWV, true, false);                                                             \
    } while (0)
#define SLB_SO(GBV, GGV)                                                                                      \
    do {                                                                                                      \
        if (pool.h == 26 && pool.w == 26) SLB_SO2(GBV, GGV, 26);                                              \
        else SLB_SO2(GBV, GGV, 0);                                                                            \
    } while (0)
    if (!(gg & ~3u) && !(gb & ~7u)) SLB_SO(7u, 3u);
    else if (!(gg & ~3u)) SLB_SO(15u, 3u);
    else SLB_SO(15u, 15u);
#undef SLB_SO3
#undef SLB_SO2
#undef SLB_SO
}

// packed observation words [M] -> boolean channels [M, C] uint16 (safelife_env.py:152-154):
// out[m*C + c] = (words[m] >> channels[c]) & 1.  HBM-write bound (2*C bytes out per 2 bytes in): every
// thread produces 8 consecutive outputs as one 16-byte store.  A cell's C output bits are first
// gathered into the low bits of one word ("channel order"; the identity when channels = 0..C-1), so
// the 8 outputs are 8 consecutive bits of the concatenation of successive cells.
template <bool IDENT>
__global__ void __launch_bounds__(256)
obs_channels_kernel(const uint16_t *__restrict__ words, size_t m_words, ObsChannels t, int c_n, uint16_t *out) {
    const size_t total = m_words * (size_t)c_n;  // uint16 elements
    const size_t chunks = total >> 3;
    uint4 *out4 = reinterpret_cast<uint4 *>(out);
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < chunks; q += (size_t)gridDim.x * blockDim.x) {
        const size_t e0 = q << 3;
        size_t m;
        int c0;
        if (total <= 0xffffffffull) {
            const uint32_t mm = (uint32_t)e0 / (uint32_t)c_n;
            m = mm;
            c0 = (int)((uint32_t)e0 - mm * (uint32_t)c_n);
        } else {
            m = e0 / (size_t)c_n;
            c0 = (int)(e0 - m * (size_t)c_n);
        }
        uint32_t bits = obs_remap<IDENT>(words[m], t, c_n) >> c0;
        for (int have = c_n - c0; have < 8; have += c_n) {
            m++;  // (m < m_words: the chunk lies entirely inside the output)
            bits |= obs_remap<IDENT>(words[m], t, c_n) << have;
        }
        out4[q] = spread8(bits);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // the last total % 8 elements
        for (size_t e = chunks << 3; e < total; e++) {
            const size_t m = e / (size_t)c_n;
            out[e] = (uint16_t)((words[m] >> t.ch[e - m * (size_t)c_n]) & 1u);
        }
    }
}

}  // namespace slb
