/*
 * safelife_b200.h — C ABI of libsafelife_b200.so: the B200 (sm_100a) batched SafeLife hot path.
 *
 * This is the drop-in boundary.  It replaces, for N lockstep boards on one GPU, what the reference
 * binds through its CPython extension `safelife.speedups` (speedups_src/module.c:256-274) plus the
 * Python rule code that sits directly on top of it on the hot path.  Every entry point is
 * `extern "C"`, takes plain pointers/sizes (no torch types), launches asynchronously on the
 * caller's CUDA stream (`stream` is a cudaStream_t passed as void*) and returns 0 on success or a
 * negative slb_status; slb_last_error() gives the message (thread-local).  There is no CPU
 * fallback anywhere behind this ABI.
 *
 * Pointers named d_* are DEVICE pointers; h_* are HOST pointers (the *_host entry points copy
 * in/out themselves and are synchronous).
 *
 * Cell encoding (uint16, safelife/safelife_game.py:74-98, speedups_src/constants.h:4-25):
 *   bit 0 alive, 1 agent, 2 pushable, 3 destructible, 4 frozen, 5 preserving, 6 inhibiting,
 *   7 spawning, 8 exit, 9/10/11 colour r/g/b, 15 pullable (12-14 unused, passed through).
 *
 * Spawn randomness (replaces speedups_src/random.c:14-57): a counter-based stream.  Draw j of
 * env e under seed s is Philox4x32-10(counter={j_lo,j_hi,e_lo,e_hi}, key={s_lo,s_hi});
 * u53 = (x0>>5)<<26 | (x1>>6); the reference's `random_float() < spawn_prob`
 * (advance_board.c:110, double < (double)float) is evaluated exactly as u53 < ceil(p_f32 * 2^53).
 * Draws are consumed in raster order by eligible cells only, board before goals, exactly like the
 * reference; the per-env draw counter is explicit state.
 */
#ifndef SAFELIFE_B200_H
#define SAFELIFE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SLB_ABI_VERSION 5

typedef enum {
    SLB_OK = 0,
    SLB_ERR_INVALID = -1,     /* bad shape / argument (reference: SystemError/ValueError, module.c:29-32) */
    SLB_ERR_CUDA = -2,        /* CUDA runtime error; message has cudaGetErrorString */
    SLB_ERR_UNSUPPORTED = -3, /* size outside what the kernels support */
    SLB_ERR_OVERFLOW = -4,    /* side-effect key table overflow */
} slb_status;

const char *slb_last_error(void);
int slb_abi_version(void);
/* number of kernels launched by this library in this process so far (bench `gpu_launches`) */
uint64_t slb_launch_count(void);

/* Per-env scalar state (replaces the GameState/SafeLifeEnv attributes listed in SURVEY.md A6/A12:
 * safelife_game.py:155-168, safelife_env.py:162-175,190-195).  80 bytes, one per env, 16-byte aligned. */
typedef struct {
    int32_t agent_x, agent_y;     /* GameState.agent_loc (x, y) */
    int32_t orientation;          /* 0 up, 1 right, 2 down, 3 left */
    int32_t game_over;            /* 0 / 1 */
    int32_t num_steps;            /* SafeLifeGame.num_steps */
    int32_t level_id;             /* index into the level pool */
    int32_t episode_length;       /* SafeLifeEnv.episode_length */
    int32_t episode_completed;    /* SafeLifeEnv.episode_completed */
    int32_t old_points;           /* SafeLifeEnv._old_game_value */
    int32_t can_exit;             /* can_exit() on the current board (kept fresh by reset/step) */
    int32_t completed, possible;  /* last performance_ratio() */
    int64_t episode_reward;       /* SafeLifeEnv.episode_reward */
    uint64_t rng_counter;         /* next draw index of this env's stream */
    double min_performance;       /* GameState.min_performance (safelife_game.py:166): the level's value at
                                     reset; a wrapper may overwrite it (slb_env_set_min_performance) */
    uint32_t flags;               /* SLB_ENV_RESCORE */
    uint32_t reserved;
} slb_env_state;
/* the next step re-evaluates score, can_exit and the exit colours even if nothing on the board changed
 * (set when something the cached values depend on was changed from outside; cleared by that step) */
#define SLB_ENV_RESCORE 1u

/* Level pool on the device (replaces _init_data of each game + the per-level scalars,
 * safelife_game.py:184-212; file_finder.py:78-98 record layout). */
typedef struct {
    int32_t n_levels, h, w, words_per_row; /* words_per_row = (w + 31) / 32 */
    const uint16_t *board;          /* [P,H,W] */
    const uint16_t *goals;          /* [P,H,W] */
    const int32_t *agent_x;         /* [P] */
    const int32_t *agent_y;         /* [P] */
    const int32_t *orientation;     /* [P] */
    const float *spawn_prob;        /* [P] python float narrowed to C float (module.c:22-24) */
    const double *min_performance;  /* [P] */
    int32_t *baseline;              /* [P]   written by slb_pool_prepare */
    uint32_t *exit_mask;            /* [P,H,words_per_row] written by slb_pool_prepare */
    void *board_res;                /* [P] boards in the RESIDENT format, written by slb_pool_prepare */
    void *goals_res;                /* [P] goals  in the RESIDENT format, written by slb_pool_prepare */
    uint32_t *plane_bits;           /* [2] device: OR of every board / goals cell seen (pool + imports);
                                       zeroed by the caller, maintained by slb_pool_prepare / _import */
    uint32_t plane_groups[2];       /* host copy as plane-group masks (bit g = group g in use) for boards /
                                       goals; written by slb_pool_prepare and slb_resident_import */
    void *level_consts;             /* [P] 32-byte slb_level_consts records, written by slb_pool_prepare */
    uint32_t external_writes;       /* host flag, set by slb_resident_import into an env buffer and by the
                                       single-stage calls (slb_execute_action, slb_game_advance,
                                       slb_update_exit_colors): some env's board / goals were written
                                       outside the fused step and reset, so the fused step may no longer
                                       assume that an env's goals equal its level's or that the score kept
                                       in slb_env_state matches the board; it then recomputes both every
                                       step (always exact either way) */
    uint32_t reserved;
} slb_pool;

/* What the step kernels read per level, packed so that it arrives with two 16-byte loads. */
typedef struct {
    float spawn_prob;
    int32_t baseline;
    uint32_t flags;                 /* SLB_LEVEL_GOALS_STATIC */
    uint32_t pad0;
    double min_performance;
    uint64_t pad1;
} slb_level_consts;
/* the level's goals board is a fixed point of advance_board and holds no spawner: it never changes
 * (no action touches goals), so an env that plays the level need not advance it */
#define SLB_LEVEL_GOALS_STATIC 1u

/* ---- RESIDENT env state ------------------------------------------------------------------------
 * Between steps an env's board and goals live on the device in the layout the kernels compute in,
 * not in the reference's uint16 [H,W] layout:
 *   - 4 <= H, W <= 32 (the 26x26 benchmark levels): bit-sliced, 64*H bytes per board — 4 plane
 *     groups x H rows x 16 bytes; group g of row r holds that row's 32-bit column masks of planes
 *     {alive 0, frozen 4, red 9, green 10}, {blue 11, destructible 3, pushable 2, pullable 15},
 *     {agent 1, preserving 5, inhibiting 6, exit 8}, {spawning 7, bits 12, 13, 14} for g = 0..3.
 *     Groups 2/3 are only touched when some cell of the pool has one of their bits (plane_bits).
 *   - any other shape >= 3x3: the uint16 [H,W] cells themselves (2*H*W bytes per board).
 * Resident buffers are caller-allocated, N * slb_resident_bytes(h, w) bytes, 16-byte aligned and
 * ZERO-INITIALISED; slb_resident_import / _export convert from / to uint16 [N,H,W] cells (the
 * format of every other boundary: level pool, speedups.advance_board, side effects, observations).
 */
size_t slb_resident_bytes(int h, int w);
/* d_cells [N,H,W] uint16 -> d_res.  `pool` (may be NULL for a conversion that no env kernel will
 * read) is the pool whose envs the boards join, which = 0 for boards, 1 for goals: the cells' bits
 * are ORed into pool->plane_bits[which] and, after a stream synchronisation, pool->plane_groups. */
int slb_resident_import(const uint16_t *d_cells, void *d_res, int n, int h, int w,
                        slb_pool *pool, int which, void *stream);
int slb_resident_export(const void *d_res, uint16_t *d_cells, int n, int h, int w, void *stream);

/* slb_pool is passed to the kernels BY VALUE at launch: a CUDA graph into which the caller captured one
 * of the calls below must be captured again after anything that changes the struct (slb_pool_prepare,
 * slb_resident_import with a pool, the single-stage calls that set external_writes).  Preparing a pool
 * again invalidates the envs that play it until they are reset. */

/* Fills pool->baseline (performance_ratio's baseline_score, safelife_game.py:622),
 * pool->exit_mask (update_exit_locs, :528-529), the resident copies of the levels, plane_bits and
 * (synchronising the stream once) plane_groups. */
int slb_pool_prepare(slb_pool *pool, void *stream);

/* ---- A1-A4: speedups.advance_board for N boards (module.c:19-44 -> advance_board.c:34-120) ----
 * d_in/d_out: [N,H,W] uint16, d_in != d_out.  d_spawn_prob: [N] float.  Env ids are env_id0 + i.
 * d_counters: [N] draw counters, read and advanced (NULL: start at 0, not written back).
 * k_steps >= 1 steps are applied; d_in is never written.  d_scratch [N,H,W] is required only when
 * k_steps > 1 and the board shape has no on-chip fast path (may be NULL otherwise). */
int slb_advance_board(const uint16_t *d_in, uint16_t *d_out, int n, int h, int w,
                      const float *d_spawn_prob, uint64_t seed, uint64_t env_id0,
                      uint64_t *d_counters, int k_steps, uint16_t *d_scratch, void *stream);

/* Same contract with HOST buffers (what `speedups.advance_board(ndarray, p)` binds): copies in,
 * steps, copies out, synchronises.  h_counters may be NULL. */
int slb_advance_board_host(const uint16_t *h_in, uint16_t *h_out, int n, int h, int w,
                           const float *h_spawn_prob, uint64_t seed, uint64_t env_id0,
                           uint64_t *h_counters, int k_steps);

/* In every function below d_board / d_goals are RESIDENT buffers of N envs (see above).
 *
 * ---- A17: reset (safelife_env.py:188-198; revert/deserialize safelife_game.py:196-212,229-234) --
 * For every env i with d_mask == NULL or d_mask[i] != 0: load level d_level_ids[i]. */
int slb_env_reset(const slb_pool *pool, int n, const int32_t *d_level_ids, const uint8_t *d_mask,
                  void *d_board, void *d_goals, slb_env_state *d_state, void *stream);

/* `game.min_performance = value` for the envs with d_mask == NULL or d_mask[i] != 0 (what the reference's
 * SimpleSideEffectPenalty.reset does after env.reset(), env_wrappers.py:313-317): stores the value, re-evaluates
 * can_exit() (safelife_game.py:522-526) from the env's last performance_ratio() so that the next
 * execute_action sees it, and marks the env so that its next step repaints the exits (the reference's
 * update_exit_colors runs every step).  The board is not touched: exit colours change at the next step. */
int slb_env_set_min_performance(int n, slb_env_state *d_state, const uint8_t *d_mask, double value,
                                void *stream);

/* ---- A7/A8: execute_action (safelife_game.py:347-393, move_agent :308-345) -----------------------
 * d_actions: [N] uint8, 0 NULL, 1-4 MOVE U/R/D/L, 5-8 TOGGLE U/R/D/L (safelife_env.py:61-71).
 * Uses and updates d_state (agent_x/y, orientation, game_over); can_exit is read from
 * d_state[i].can_exit.  d_reward: [N] int32 action reward (0 / +1). */
int slb_execute_action(slb_pool *pool, int n, void *d_board, slb_env_state *d_state,
                       const uint8_t *d_actions, int32_t *d_reward, void *stream);

/* ---- A4: SafeLifeGame.advance_board (safelife_game.py:657-660) on resident state: num_steps += 1,
 * board then goals (the order the env's stream is consumed in), spawn_prob of each env's level. */
int slb_game_advance(slb_pool *pool, int n, void *d_board, void *d_goals,
                     slb_env_state *d_state, uint64_t seed, uint64_t env_id0, void *stream);

/* ---- A9-A11: current_points, performance_ratio, can_exit (safelife_game.py:590-631, 522-526) ----
 * Any output pointer may be NULL.  Does not modify state. */
int slb_points_and_performance(const slb_pool *pool, int n, const void *d_board,
                               const void *d_goals, const slb_env_state *d_state,
                               int32_t *d_points, int32_t *d_completed, int32_t *d_possible,
                               uint8_t *d_can_exit, void *stream);

/* update_exit_colors (safelife_game.py:531-537): evaluates can_exit() on the current board/goals,
 * repaints the level's initial exit cells (272 / 784) and refreshes d_state[i].can_exit,
 * .completed, .possible. */
int slb_update_exit_colors(slb_pool *pool, int n, void *d_board, const void *d_goals,
                           slb_env_state *d_state, void *stream);

/* ---- A12: one SafeLifeEnv.step for N envs, fused (safelife_env.py:157-186) ----------------------
 * execute_action -> advance board -> advance goals -> current_points -> reward -> counters ->
 * update_exit_colors -> done.  d_reward [N] int32, d_done [N] uint8 (either may be NULL).
 * auto_reset_stride > 0: an env that finishes this step is reset (after its outputs are written)
 * to level (level_id + auto_reset_stride) % n_levels. */
int slb_env_step(const slb_pool *pool, int n, void *d_board, void *d_goals,
                 slb_env_state *d_state, const uint8_t *d_actions, int time_limit, uint64_t seed,
                 uint64_t env_id0, int32_t *d_reward, uint8_t *d_done, int auto_reset_stride,
                 void *stream);

/* T fused env-steps with pre-staged actions: d_actions [T,N] uint8, d_reward [T,N] int32,
 * d_done [T,N] uint8 (outputs may be NULL).  Board state stays on chip between steps when the
 * shape has a fast path; otherwise this loops slb_env_step. */
int slb_env_rollout(const slb_pool *pool, int n, void *d_board, void *d_goals,
                    slb_env_state *d_state, const uint8_t *d_actions, int t_steps, int time_limit,
                    uint64_t seed, uint64_t env_id0, int32_t *d_reward, uint8_t *d_done,
                    int auto_reset_stride, void *stream);

/* Host-buffer step (the call a CPU-side policy makes): actions in, reward/done out, state stays
 * resident on the device.  Synchronous. */
int slb_env_step_host(const slb_pool *pool, int n, void *d_board, void *d_goals,
                      slb_env_state *d_state, const uint8_t *h_actions, int time_limit,
                      uint64_t seed, uint64_t env_id0, int32_t *h_reward, uint8_t *h_done,
                      int auto_reset_stride, void *stream);

/* The same without the final wait: H2D copy, step and D2H copy are queued on `stream` (give steps that
 * should overlap their own streams; at most 16 streams with a step in flight per calling thread).  h_actions is consumed
 * before the call returns; h_reward / h_done are valid only after slb_stream_wait(stream), called from
 * the SAME host thread, returns (a plain cudaStreamSynchronize is not enough: on a non-default stream
 * the step goes through a page-locked staging block inside the library and slb_stream_wait hands the
 * results over; the host buffers need not be page-locked).  On a non-default stream, batches of up to 4096
 * envs on boards of 4..32 rows / columns run as ONE kernel launch without any DMA copy: the actions travel
 * by value in the launch parameters, and the step's last CTA writes reward / done to the page-locked block
 * and publishes a sequence number that slb_stream_wait spins on (SLB_HOST_DIRECT=0 turns this off).
 * Other cases run {H2D copy, step, D2H copy} as one cached CUDA graph (SLB_HOST_GRAPH=0: plain calls). */
int slb_env_step_host_async(const slb_pool *pool, int n, void *d_board, void *d_goals,
                            slb_env_state *d_state, const uint8_t *h_actions, int time_limit,
                            uint64_t seed, uint64_t env_id0, int32_t *h_reward, uint8_t *h_done,
                            int auto_reset_stride, void *stream);
/* Low-latency wait for the slb_env_step_host_async call in flight on `stream` (spins on the step's
 * published sequence number, or on cudaStreamQuery where the step ran as copies), then delivers its
 * outputs; without such a call it waits for everything queued on the stream. */
int slb_stream_wait(void *stream);

/* A host stepper binds every argument of slb_env_step_host_async that does not change from step to
 * step, so that the per-step call crosses the FFI with 4 arguments (a ctypes call costs ~0.25 us per
 * argument).  `pool` must outlive the stepper.  step = slb_env_step_host_async, wait =
 * slb_stream_wait on the bound stream (same thread rule). */
typedef struct slb_stepper slb_stepper;
int slb_stepper_create(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                       int time_limit, uint64_t seed, uint64_t env_id0, int auto_reset_stride,
                       void *stream, slb_stepper **out);
int slb_stepper_step(slb_stepper *s, const uint8_t *h_actions, int32_t *h_reward, uint8_t *h_done);
int slb_stepper_wait(slb_stepper *s);
void slb_stepper_destroy(slb_stepper *s);
/* Optional: the caller's outputs are ONE page-locked block [int32 reward x n | uint8 done x n] (16-byte aligned,
 * n a multiple of 16, e.g. a torch pin_memory() tensor) that stays allocated while it is bound: steps called with
 * exactly these pointers then have their last CTA write the outputs straight into that block — the copy out of
 * the library's own page-locked mirror is skipped (about 1 us of host time per 4 096-env step).  Steps called with
 * other pointers work as before.  NULL, NULL unbinds.  SLB_ERR_INVALID if the block is not page-locked or not of
 * that shape. */
int slb_stepper_bind_outputs(slb_stepper *s, int32_t *h_reward, uint8_t *h_done);

/* ---- A13/A14: side_effect_score up to the EMD (side_effects.py:59-141) ---------------------------
 * For episode i: b0 = d_board0[i] advanced d_num_steps[i] times, then num_samples paired steps of
 * b0 (inaction, draws first) and b1 = d_board1[i] (action); per-cell-type count maps.
 * d_keys [N,kmax] uint16 (0 = empty slot), d_counts [N,2,kmax,H,W] uint16 (must be zeroed by the
 * caller), d_nkeys [N] int32 (-1 on key-table overflow).  d_counters [N] in/out (may be NULL).
 * d_scratch: [N,4,H,W] uint16 workspace, required only when 4*H*W*2 bytes exceed shared memory. */
int slb_side_effect_counts(const uint16_t *d_board0, const uint16_t *d_board1, int n, int h, int w,
                           const int32_t *d_num_steps, int num_samples, const float *d_spawn_prob,
                           uint64_t seed, uint64_t env_id0, uint64_t *d_counters, int kmax,
                           uint16_t *d_keys, uint16_t *d_counts, int32_t *d_nkeys,
                           uint16_t *d_scratch, void *stream);

/* ---- "next" row 4: the EMD stage of side_effect_score for a whole batch (side_effects.py:12-56) ---------
 * The reference hands every (episode, cell type) pair of density maps to third-party pyemd.emd
 * (pyemd 0.5.1, side_effects.py:56; not available here: the value is "parity unpinned" against pyemd and checked
 * against two independent CPU solvers).  Here: problem p = i * kmax + slot takes the two uint16 count maps
 * d_counts[i, 0 / 1, slot] of slb_side_effect_counts (densities = counts / num_samples), the changed cells
 * (|a - b| > 1e-3 max|a - b|), and computes EMD-hat as pyemd.emd does (Pele & Werman's emd_hat_gd_metric: cancel
 * min(a, b) inside every cell, exact min-cost transport of the rest, extra_mass_penalty per unit of surplus).
 * Slots with d_keys[p] == 0 are skipped.
 * slb_emd_sizes: d_n_src[p] / d_n_snk[p] = number of supplying / demanding cells of problem p (0 for skipped
 * slots) — what the caller sizes the solver with.
 * slb_emd_solve: solves the problems d_order[0 .. n_problems) (largest first balances best), one warp each;
 * d_emd[p] receives the value.  d_lut [H,W] doubles: ground distance of two cells whose coordinates differ by
 * (|dy|, |dx|) after the wrap (made on the host exactly as side_effects.py:50-55); wrap_x / wrap_y as there.
 * vmax >= max(n_src + n_snk), ft_stride >= max(n_src * n_snk) over the problems; d_scratch: scratch_doubles
 * doubles, at least 4 * ft_stride (the grid is sized to what fits); d_queue: one zeroed uint32; d_status: one
 * zeroed int32, set to 1 if a problem did not converge (its d_emd is -1).  SLB_ERR_UNSUPPORTED when vmax
 * exceeds the solver's shared memory (about 1800 changed cells). */
int slb_emd_sizes(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                  int num_samples, int32_t *d_n_src, int32_t *d_n_snk, void *stream);
int slb_emd_solve(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                  int num_samples, const double *d_lut, int wrap_x, int wrap_y, double extra_mass_penalty,
                  const int32_t *d_order, int n_problems, int vmax, size_t ft_stride, double *d_scratch,
                  size_t scratch_doubles, uint32_t *d_queue, double *d_emd, int32_t *d_status, void *stream);

/* slb_emd_solve for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as slb_emd_solve, without d_scratch / scratch_doubles / d_queue: the
 * flow matrices and the queue word live in a static device array of SLB_EMD_STATIC_SCRATCH_DOUBLES doubles
 * (16 MiB) that every GPU which loads the library holds until exit.  The call allocates and frees nothing (no
 * cudaMalloc / cudaFree, no allocating library routine); it clears the queue word on `stream` itself.
 * Limit: ft_stride * 4 <= SLB_EMD_STATIC_SCRATCH_DOUBLES, i.e. n_src * n_snk <= 524 288 per problem (every
 * 26 x 26 density map fits: <= 338 * 338); beyond that SLB_ERR_UNSUPPORTED with a message naming the limit —
 * use slb_emd_solve.  All calls on a device share the array: two calls that overlap on the GPU (different
 * streams, or different host threads / processes' streams of one context) race; calls on one stream are safe. */
#define SLB_EMD_STATIC_SCRATCH_DOUBLES 2097152
int slb_emd_solve_static(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                         int num_samples, const double *d_lut, int wrap_x, int wrap_y,
                         double extra_mass_penalty, const int32_t *d_order, int n_problems, int vmax,
                         size_t ft_stride, double *d_emd, int32_t *d_status, void *stream);

/* ---- statistics vector for the multi-GPU all-reduce -----------------------------
 * d_stats[8] int64 += { sum reward, n done, sum episode_length of done envs, sum episode_reward of
 * done envs, sum completed of done, sum possible of done, n envs stepped, n game_over }. */
int slb_accumulate_stats(int n, const slb_env_state *d_state, const int32_t *d_reward,
                         const uint8_t *d_done, int64_t *d_stats, void *stream);

/* ---- "next" row 1: observation (safelife_env.py:125-155, helper_utils.py:41-74) ------------------
 * d_obs: [N,vh,vw] uint16 words (bit b = channel b).  Channel split is a view-level bit test. */
int slb_get_obs(const slb_pool *pool, int n, const void *d_board, const void *d_goals,
                const slb_env_state *d_state, int vh, int vw, int remove_white_goals,
                uint16_t *d_obs, void *stream);

/* ---- "next" row 2: SimpleSideEffectPenalty.step's count (env_wrappers.py:319-346) ------------------
 * d_count[i] = number of cells of env i that differ from the level's initial board, ignoring the
 * player's attribute bits, exit cells, red life that went away and non-red life on blue goals. */
int slb_side_effect_penalty(const slb_pool *pool, int n, const void *d_board, const void *d_goals,
                            const slb_env_state *d_state, int32_t *d_count, void *stream);

/* get_obs with the channel split fused in (safelife_env.py:125-155 in one call): d_out uint16
 * [N, vh, vw, n_channels], 16-byte aligned; h_channels as for slb_obs_channels.  Boards of 4..32 rows /
 * columns with vh * vw <= 1089 (33 x 33) run as ONE kernel; any other shape goes through slb_get_obs +
 * slb_obs_channels and needs d_words_scratch (uint16 [N, vh, vw]; SLB_ERR_UNSUPPORTED if it is NULL). */
int slb_get_obs_channels(const slb_pool *pool, int n, const void *d_board, const void *d_goals,
                         const slb_env_state *d_state, int vh, int vw, int remove_white_goals,
                         const uint8_t *h_channels, int n_channels, uint16_t *d_words_scratch,
                         uint16_t *d_out, void *stream);

/* SafeLifeEnv.step up to and including get_obs (safelife_env.py:157-186 + 125-155) in ONE launch, for boards
 * of 4..32 rows / columns: slb_env_step, then the observation of the state it produced (of the new episode
 * where an env was auto-reset).  n_channels = 0: packed words uint16 [N, vh, vw]; else the channel split
 * uint16 [N, vh, vw, n_channels] (vh * vw <= 1089, d_obs 16-byte aligned).  Other shapes:
 * SLB_ERR_UNSUPPORTED — call slb_env_step and slb_get_obs(_channels). */
int slb_env_step_obs(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                     const uint8_t *d_actions, int time_limit, uint64_t seed, uint64_t env_id0,
                     int32_t *d_reward, uint8_t *d_done, int auto_reset_stride, int vh, int vw,
                     int remove_white_goals, const uint8_t *h_channels, int n_channels, uint16_t *d_obs,
                     void *stream);

/* SafeLifeEnv.step for a policy on the GPU (safelife_env.py:157-186) with everything that does not change
 * from step to step bound once: step = slb_env_step_obs (vh > 0 and d_obs given; SLB_ERR_UNSUPPORTED where
 * that call is) or slb_env_step.  Rewards go to the d_reward given to the step, or to the one given at
 * creation when that is NULL.  h_channels is copied.  `pool` must outlive the stepper.
 * d_global_counter (may be NULL): SafeLifeEnv.global_counter kept on the device (safelife_env.py:169-173) —
 * 64 pairs of uint64 {num_steps, episodes_completed}, zeroed by the caller; every step adds, for each env that
 * was not already done, 1 to num_steps and its `done` to episodes_completed, env i into pair i % 64 (the
 * caller sums the pairs when it reads the counter; no host synchronisation per step). */
typedef struct slb_dev_stepper slb_dev_stepper;
int slb_dev_stepper_create(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                           int time_limit, uint64_t seed, uint64_t env_id0, int auto_reset_stride,
                           int32_t *d_reward, int vh, int vw, int remove_white_goals,
                           const uint8_t *h_channels, int n_channels, uint64_t *d_global_counter,
                           slb_dev_stepper **out);
int slb_dev_stepper_step(slb_dev_stepper *s, const uint8_t *d_actions, int32_t *d_reward, uint8_t *d_done,
                         uint16_t *d_obs, void *stream);
void slb_dev_stepper_destroy(slb_dev_stepper *s);

/* The channel split of safelife_env.py:152-154 for M packed words: d_out [M, n_channels] uint16,
 * d_out[m, c] = (d_words[m] >> h_channels[c]) & 1.  h_channels: n_channels <= 16 bit indices in HOST
 * memory (read before the call returns); d_out 16-byte aligned. */
int slb_obs_channels(const uint16_t *d_words, size_t m, const uint8_t *h_channels, int n_channels,
                     uint16_t *d_out, void *stream);

/* ---- measurement aid (the integer roofline is measured on the box, not assumed) ----
 * Launches a kernel of `ctas` x 256 threads in which every thread issues `iters` x 8 independent
 * dependent-chain integer instructions: kind 0 = LOP3 only (the ALU pipe the bit-sliced stencil runs on),
 * kind 1 = LOP3 and IMAD alternating (ALU + FMA pipes, the issue-rate ceiling).  The caller times the
 * launch with events; *h_warp_insts (host, may be NULL) receives the warp instructions it issues
 * (ctas * 8 warps * iters * 8, loop overhead not counted). */
int slb_microbench_int(int kind, int ctas, int iters, uint32_t *d_sink, uint64_t *h_warp_insts, void *stream);

#ifdef __cplusplus
}
#endif
#endif
