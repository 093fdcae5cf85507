/*
 * slo_oracle.h — CPU ORACLE for the SafeLife hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (safelife_k2_b200 + libsafelife_b200.so) never links, imports
 * or calls it; the product path fails loudly when its CUDA library is missing.
 *
 * Every function is a plain-C restatement (written from the closed-form semantics, not copied)
 * of the reference function cited next to it.  Parity of this oracle is PINNED: see
 * tests/test_oracle_vs_reference.py (runs the compiled reference itself, in the build container)
 * and tests/golden/ (vectors generated from the reference, script committed).
 * Exception: the earth-mover-distance value of side_effect_score (third-party pyemd 0.5.1,
 * absent from the reference tree) — "parity unpinned" for that one number, see DESIGN.md.
 */
#ifndef SLO_ORACLE_H
#define SLO_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Counter-based spawn stream (the build's choice): draw j of env e under seed s is
 * Philox4x32-10(counter = {j_lo, j_hi, e_lo, e_hi}, key = {s_lo, s_hi}); the double is built the
 * way numpy's legacy random_sample builds it: ((x0 >> 5) * 2^26 + (x1 >> 6)) / 2^53. */
typedef struct {
    uint64_t seed;
    uint64_t env;
    uint64_t counter; /* next draw index; advanced by every draw */
} slo_stream;

void slo_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
uint64_t slo_stream_u53(uint64_t seed, uint64_t env, uint64_t index);
double slo_stream_next(slo_stream *st);
/* fill n doubles of env's stream starting at draw index `start` (for the Python feeder) */
void slo_stream_fill(uint64_t seed, uint64_t env, uint64_t start, int64_t n, double *out);

/* speedups_src/advance_board.c:34-120 (Appendix A of SURVEY.md). `st` may be NULL when the
 * board has no spawners (a draw with st == NULL aborts). */
void slo_advance_board(const uint16_t *in, uint16_t *out, int h, int w, float spawn_prob,
                       slo_stream *st);

/* Per-env scalar state, same field order as the product's slb_env_state (include/safelife_b200.h). */
typedef struct {
    int32_t agent_x, agent_y, orientation, game_over;
    int32_t num_steps, level_id, episode_length, episode_completed;
    int32_t old_points, can_exit, completed, possible;
    int64_t episode_reward;
    uint64_t rng_counter;
    double min_performance; /* GameState.min_performance: the level's value after a reset, writable */
    uint32_t flags, reserved; /* the product's bookkeeping; always 0 here */
} slo_env_state; /* 80 bytes */

/* safelife_game.py:347-393 for the nine env actions of safelife_env.py:61-71 (0 NULL, 1-4 MOVE
 * U/R/D/L, 5-8 TOGGLE U/R/D/L); move_agent :308-345.  `can_exit` is the value can_exit() would
 * return on this board.  Returns the action reward (0 or +1). */
int slo_execute_action(uint16_t *board, int h, int w, slo_env_state *s, int action, int can_exit);

/* safelife_game.py:590-599 */
int64_t slo_current_points(const uint16_t *board, const uint16_t *goals, int h, int w);

/* safelife_game.py:601-631 with unit_rewards=True.  `baseline` = the same masked sum on the
 * level's initial board/goals (slo_baseline_score). */
int64_t slo_baseline_score(const uint16_t *board0, const uint16_t *goals0, int h, int w);
void slo_performance_ratio(const uint16_t *board, const uint16_t *goals, int h, int w,
                           int64_t baseline, int64_t *completed, int64_t *possible);
/* safelife_game.py:522-526 */
int slo_can_exit(double min_performance, int64_t completed, int64_t possible);
/* safelife_game.py:528-537: exit locations come from the level's INITIAL board */
void slo_update_exit_colors(uint16_t *board, const uint16_t *board0, int h, int w, int can_exit);

/* Level pool record set (structure-of-arrays over P levels). */
typedef struct {
    int32_t n_levels, h, w, pad;
    const uint16_t *board;     /* [P,H,W] */
    const uint16_t *goals;     /* [P,H,W] */
    const int32_t *agent_x;    /* [P] */
    const int32_t *agent_y;    /* [P] */
    const int32_t *orientation;/* [P] */
    const float *spawn_prob;   /* [P] python float narrowed to C float, module.c:22-24 */
    const double *min_performance; /* [P] */
} slo_pool;

/* safelife_env.py:188-198 (reset) + safelife_game.py:196-212,229-234 (revert/deserialize) */
void slo_env_reset(const slo_pool *pool, int level_id, uint16_t *board, uint16_t *goals,
                   slo_env_state *s);

/* safelife_env.py:157-186: one SafeLifeEnv.step.  Returns reward; *done_out = episode_completed;
 * *times_up_out = times_up. */
int64_t slo_env_step(const slo_pool *pool, uint16_t *board, uint16_t *goals, slo_env_state *s,
                     int action, int time_limit, uint64_t seed, uint64_t env_id,
                     int *done_out, int *times_up_out);

/* N envs (serial here: this image has no libgomp; callers thread over slices, ctypes drops the GIL).  Arrays are [N,...]. env ids = env_id0 + i.
 * use_ref_advance: if non-NULL, a function pointer with the reference's advance_board signature
 * plus a stream setter is used for the physics instead of slo_advance_board (bench "reference" arm). */
typedef void (*slo_ref_advance_fn)(uint16_t *b1, uint16_t *b2, int nrow, int ncol, float spawn_prob);
typedef void (*slo_ref_stream_set_fn)(uint64_t seed, uint64_t env, uint64_t counter);
typedef uint64_t (*slo_ref_stream_get_fn)(void);
void slo_set_reference_advance(slo_ref_advance_fn adv, slo_ref_stream_set_fn set,
                               slo_ref_stream_get_fn get);

void slo_env_reset_batch(const slo_pool *pool, int n, const int32_t *level_ids, uint16_t *boards,
                         uint16_t *goals, slo_env_state *states);
void slo_env_step_batch(const slo_pool *pool, int n, uint16_t *boards, uint16_t *goals,
                        slo_env_state *states, const int32_t *actions, int time_limit,
                        uint64_t seed, uint64_t env_id0, int64_t *rewards, uint8_t *dones,
                        int nthreads);
void slo_advance_board_batch(int n, int h, int w, const uint16_t *in, uint16_t *out,
                             const float *spawn_prob, uint64_t seed, uint64_t env_id0,
                             uint64_t *counters, int nthreads);

/* side_effects.py:59-161 up to (not including) the EMD: catch-up of the initial board by
 * num_steps, then num_samples paired steps (inaction board draws first), per-cell-type count
 * maps.  keys[kmax] (sorted ascending), counts[2][kmax][h][w] (branch 0 = inaction, 1 = action),
 * returns number of keys, or -1 if more than kmax distinct keys appear.  The stream continues
 * from *counter. */
int slo_side_effect_counts(const uint16_t *board0, const uint16_t *board1, int h, int w,
                           int num_steps, int num_samples, float spawn_prob, uint64_t seed,
                           uint64_t env_id, uint64_t *counter, int kmax, uint16_t *keys,
                           uint16_t *counts);

#ifdef __cplusplus
}
#endif
#endif
