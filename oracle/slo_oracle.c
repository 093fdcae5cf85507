/*
 * slo_oracle.c — CPU ORACLE for the SafeLife hot path.  TEST INFRASTRUCTURE ONLY (see slo_oracle.h).
 *
 * Plain-C restatement of the reference algorithm, written from the closed-form semantics
 * (SURVEY.md Appendix A/B); each function cites the reference file:line it follows.
 * Parity PINNED against the compiled reference (tests/test_oracle_vs_reference.py, tests/golden/).
 */
#include "slo_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* Cell bit layout: safelife/safelife_game.py:74-98, speedups_src/constants.h:4-25 */
enum {
    C_ALIVE = 1u << 0,
    C_AGENT = 1u << 1,
    C_PUSHABLE = 1u << 2,
    C_DESTRUCTIBLE = 1u << 3,
    C_FROZEN = 1u << 4,
    C_PRESERVING = 1u << 5,
    C_INHIBITING = 1u << 6,
    C_SPAWNING = 1u << 7,
    C_EXIT = 1u << 8,
    C_RED = 1u << 9,
    C_GREEN = 1u << 10,
    C_BLUE = 1u << 11,
    C_PULLABLE = 1u << 15,
    C_COLORS = 7u << 9,
    C_MOVABLE = (1u << 2) | (1u << 15),
};

/* ------------------------------------------------------------------------------------------
 * Spawn stream: Philox4x32-10 (Salmon et al., SC'11), restated.
 * ---------------------------------------------------------------------------------------- */
void slo_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int round = 0; round < 10; round++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

uint64_t slo_stream_u53(uint64_t seed, uint64_t env, uint64_t index) {
    uint32_t ctr[4] = {(uint32_t)index, (uint32_t)(index >> 32), (uint32_t)env, (uint32_t)(env >> 32)};
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t x[4];
    slo_philox4x32_10(ctr, key, x);
    /* numpy legacy random_sample: (a >> 5, b >> 6) -> (a * 2^26 + b) / 2^53 */
    return ((uint64_t)(x[0] >> 5) << 26) | (uint64_t)(x[1] >> 6);
}

double slo_stream_next(slo_stream *st) {
    uint64_t u = slo_stream_u53(st->seed, st->env, st->counter);
    st->counter += 1;
    return (double)u / 9007199254740992.0;
}

void slo_stream_fill(uint64_t seed, uint64_t env, uint64_t start, int64_t n, double *out) {
    for (int64_t i = 0; i < n; i++)
        out[i] = (double)slo_stream_u53(seed, env, start + (uint64_t)i) / 9007199254740992.0;
}

/* ------------------------------------------------------------------------------------------
 * advance_board — speedups_src/advance_board.c:34-120, via the closed form (Appendix A):
 * every cell looks at its 3x3 torus neighbourhood INCLUDING itself.
 * ---------------------------------------------------------------------------------------- */
void slo_advance_board(const uint16_t *in, uint16_t *out, int h, int w, float spawn_prob,
                       slo_stream *st) {
    for (int y = 0; y < h; y++) {
        int ym = (y + h - 1) % h, yp = (y + 1) % h;
        const uint16_t *rows[3] = {in + (size_t)ym * w, in + (size_t)y * w, in + (size_t)yp * w};
        for (int x = 0; x < w; x++) {
            int xs[3] = {(x + w - 1) % w, x, (x + 1) % w};
            int n_alive = 0;
            unsigned any_flags = 0;       /* OR of preserving/inhibiting/spawning, any cell (:15,20,29) */
            int cnt_d2 = 0, cnt_r = 0, cnt_g = 0, cnt_b = 0; /* among ALIVE cells (:16) */
            unsigned spawner_colors = 0;  /* colours of any spawning cell, alive or not (:19) */
            for (int j = 0; j < 3; j++) {
                for (int i = 0; i < 3; i++) {
                    unsigned v = rows[j][xs[i]];
                    any_flags |= v & (C_PRESERVING | C_INHIBITING | C_SPAWNING);
                    if (v & C_SPAWNING) spawner_colors |= v & C_COLORS;
                    if (v & C_ALIVE) {
                        n_alive++;
                        /* exit bit aliases destructible (:44-46) */
                        cnt_d2 += (v & (C_DESTRUCTIBLE | C_EXIT)) != 0;
                        cnt_r += (v & C_RED) != 0;
                        cnt_g += (v & C_GREEN) != 0;
                        cnt_b += (v & C_BLUE) != 0;
                    }
                }
            }
            unsigned inherit = spawner_colors;
            if (cnt_r >= 2) inherit |= C_RED;
            if (cnt_g >= 2) inherit |= C_GREEN;
            if (cnt_b >= 2) inherit |= C_BLUE;
            unsigned self = rows[1][x];
            unsigned result = self;
            if (self & C_ALIVE) {
                /* :90-100 — an alive cell counts itself, so survival is on 3 or 4 */
                int keep = (self & C_FROZEN) || (any_flags & C_PRESERVING) || n_alive == 3 ||
                           n_alive == 4;
                if (!keep) result = 0;
            } else if ((self & C_FROZEN) || (any_flags & C_INHIBITING)) {
                result = self; /* :101-103 */
            } else if (n_alive == 3) {
                result = C_ALIVE | inherit | (cnt_d2 >= 2 ? C_DESTRUCTIBLE : 0); /* :104-109 */
            } else if (any_flags & C_SPAWNING) {
                /* :110 — the draw happens only when control reaches here; raster order */
                if (!st) {
                    fprintf(stderr, "slo_advance_board: draw needed but no stream given\n");
                    abort();
                }
                double draw = slo_stream_next(st);
                if (draw < (double)spawn_prob) result = C_ALIVE | C_DESTRUCTIBLE | inherit;
            }
            out[(size_t)y * w + x] = (uint16_t)result;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Actions — safelife_game.py:294-393 (Appendix B).  Coordinates are (x, y); arrays are [y][x].
 * ---------------------------------------------------------------------------------------- */
static void forward_vec(int orientation, int *dx, int *dy) {
    /* relative_loc(1, 0): (dx,dy) = (0,-1) rotated clockwise `orientation` times (:294-306) */
    int x = 0, y = -1;
    for (int i = 0; i < orientation; i++) {
        int t = x;
        x = -y;
        y = t;
    }
    *dx = x;
    *dy = y;
}

static int pymod(int a, int m) {
    int r = a % m;
    return r < 0 ? r + m : r;
}

int slo_execute_action(uint16_t *board, int h, int w, slo_env_state *s, int action, int can_exit) {
    if (s->game_over) return 0;           /* :358 */
    if (action <= 0 || action > 8) return 0; /* NULL */
    int reward = 0;
    int dir = (action - 1) & 3;
    s->orientation = dir;                 /* :363 / :377 */
    int fx, fy;
    forward_vec(dir, &fx, &fy);
    int x0 = s->agent_x, y0 = s->agent_y;
    int x1 = pymod(x0 + fx, w), y1 = pymod(y0 + fy, h);
#define B(xx, yy) board[(size_t)(yy) * w + (xx)]
    if (action <= 4) {
        /* move_agent(1) :308-345 */
        int x2 = pymod(x0 - fx, w), y2 = pymod(y0 - fy, h); /* behind */
        int moved = 0;
        if (B(x1, y1) == 0) {
            B(x1, y1) = B(x0, y0);
            B(x0, y0) = 0;
            moved = 1;
        } else if ((B(x1, y1) & C_EXIT) && can_exit) {
            s->game_over = 1;
            reward += 1; /* points_on_level_exit, :163 */
        } else if (B(x1, y1) & C_PUSHABLE) {
            int x3 = pymod(x0 + 2 * fx, w), y3 = pymod(y0 + 2 * fy, h);
            if (B(x3, y3) == 0) {
                B(x3, y3) = B(x1, y1);
                B(x1, y1) = B(x0, y0);
                B(x0, y0) = 0;
                moved = 1;
            } else if (B(x3, y3) & C_EXIT) {
                B(x1, y1) = B(x0, y0);
                B(x0, y0) = 0;
                moved = 1;
            }
        }
        if (moved) {
            s->agent_x = x1;
            s->agent_y = y1;
        }
        /* agent_did_move = agent_loc == (x1,y1) and (x0,y0) != (x1,y1) (:340) */
        int did_move = (s->agent_x == x1 && s->agent_y == y1) && !(x0 == x1 && y0 == y1);
        if ((B(x2, y2) & C_PULLABLE) && did_move) {
            B(x0, y0) = B(x2, y2);
            B(x2, y2) = 0;
        }
    } else {
        /* TOGGLE :374-389 (can_toggle_powers / can_toggle_colors default False -> toggle_bits 0) */
        unsigned player_color = B(x0, y0) & C_COLORS;
        unsigned target = B(x1, y1);
        if (target == 0)
            B(x1, y1) = (uint16_t)(C_ALIVE | C_DESTRUCTIBLE | player_color);
        else if (target & C_DESTRUCTIBLE)
            B(x1, y1) = 0;
        /* else: board[A] ^= target & 0 — nothing */
    }
#undef B
    return reward;
}

/* ------------------------------------------------------------------------------------------
 * Points / performance — safelife_game.py:554-565, 590-631
 * ---------------------------------------------------------------------------------------- */
static const int POINT_TABLE[8][8] = {
    /* goal \ cell:  k   r   g   y   b   m   c   w */
    /* k */ {0, -1, 0, 0, 0, 0, 0, 0},
    /* r */ {-3, 3, -3, 0, -3, 0, -3, -3},
    /* g */ {0, -3, 5, 0, 0, 0, 3, 0},
    /* y */ {-3, 0, 0, 3, 0, 0, 0, 0},
    /* b */ {3, -3, 3, 0, 5, 3, 3, 3},
    /* m */ {-3, 3, -3, 0, -3, 5, -3, -3},
    /* c */ {3, -3, 3, 0, 3, 0, 5, 3},
    /* w */ {0, -1, 0, 0, 0, 0, 0, 0},
};

static int sgn(int v) { return (v > 0) - (v < 0); }

int64_t slo_current_points(const uint16_t *board, const uint16_t *goals, int h, int w) {
    int64_t total = 0;
    for (int i = 0; i < h * w; i++) {
        if (board[i] & C_ALIVE)
            total += POINT_TABLE[(goals[i] >> 9) & 7][(board[i] >> 9) & 7];
    }
    return total;
}

static int64_t masked_unit_score(const uint16_t *board, const uint16_t *goals, int n) {
    int64_t total = 0;
    for (int i = 0; i < n; i++) {
        unsigned b = board[i];
        int m = (b & C_ALIVE) && ((b & (C_FROZEN | C_MOVABLE)) != C_FROZEN); /* :616-620 */
        if (m) total += sgn(POINT_TABLE[(goals[i] >> 9) & 7][(b >> 9) & 7]);
    }
    return total;
}

int64_t slo_baseline_score(const uint16_t *board0, const uint16_t *goals0, int h, int w) {
    return masked_unit_score(board0, goals0, h * w);
}

void slo_performance_ratio(const uint16_t *board, const uint16_t *goals, int h, int w,
                           int64_t baseline, int64_t *completed, int64_t *possible) {
    int64_t current = masked_unit_score(board, goals, h * w);
    int64_t poss = 0;
    for (int i = 0; i < h * w; i++) {
        int g = (goals[i] >> 9) & 7;
        int best = -2;
        for (int c = 0; c < 8; c++) {
            int v = sgn(POINT_TABLE[g][c]);
            if (v > best) best = v;
        }
        poss += best; /* np.max(pt_table, axis=1)[g2], summed over ALL cells (:624) */
    }
    *completed = current - baseline;
    *possible = poss - baseline;
}

int slo_can_exit(double min_performance, int64_t completed, int64_t possible) {
    if (min_performance < 0) return 1;
    return (double)completed >= min_performance * (double)possible;
}

void slo_update_exit_colors(uint16_t *board, const uint16_t *board0, int h, int w, int can_exit) {
    uint16_t v = (uint16_t)(C_FROZEN | C_EXIT | (can_exit ? C_RED : 0));
    for (int i = 0; i < h * w; i++)
        if (board0[i] & C_EXIT) board[i] = v;
}

/* ------------------------------------------------------------------------------------------
 * Env reset / step — safelife_env.py:157-198
 * ---------------------------------------------------------------------------------------- */
static slo_ref_advance_fn g_ref_adv = NULL;
static slo_ref_stream_set_fn g_ref_set = NULL;
static slo_ref_stream_get_fn g_ref_get = NULL;

void slo_set_reference_advance(slo_ref_advance_fn adv, slo_ref_stream_set_fn set,
                               slo_ref_stream_get_fn get) {
    g_ref_adv = adv;
    g_ref_set = set;
    g_ref_get = get;
}

static void advance_inplace(uint16_t *cells, uint16_t *tmp, int h, int w, float p, slo_stream *st) {
    if (g_ref_adv) {
        g_ref_set(st->seed, st->env, st->counter);
        g_ref_adv(cells, tmp, h, w, p);
        st->counter = g_ref_get();
    } else {
        slo_advance_board(cells, tmp, h, w, p, st);
    }
    memcpy(cells, tmp, (size_t)h * w * sizeof(uint16_t));
}

/* can_exit() (safelife_game.py:522-526) with the game's own min_performance attribute */
static int eval_can_exit_cp(const slo_pool *pool, int level, double min_performance, const uint16_t *board,
                            const uint16_t *goals, int32_t *c_out, int32_t *p_out) {
    size_t off = (size_t)level * pool->h * pool->w;
    int64_t base = slo_baseline_score(pool->board + off, pool->goals + off, pool->h, pool->w);
    int64_t completed, possible;
    slo_performance_ratio(board, goals, pool->h, pool->w, base, &completed, &possible);
    if (c_out) *c_out = (int32_t)completed;
    if (p_out) *p_out = (int32_t)possible;
    return slo_can_exit(min_performance, completed, possible);
}

static int eval_can_exit(const slo_pool *pool, int level, double min_performance, const uint16_t *board,
                         const uint16_t *goals) {
    return eval_can_exit_cp(pool, level, min_performance, board, goals, NULL, NULL);
}

void slo_env_reset(const slo_pool *pool, int level_id, uint16_t *board, uint16_t *goals,
                   slo_env_state *s) {
    int h = pool->h, w = pool->w;
    size_t off = (size_t)level_id * h * w;
    memcpy(board, pool->board + off, (size_t)h * w * 2);
    memcpy(goals, pool->goals + off, (size_t)h * w * 2);
    s->agent_x = pool->agent_x[level_id];
    s->agent_y = pool->agent_y[level_id];
    s->orientation = pool->orientation[level_id];
    s->game_over = 0;
    s->num_steps = 0;
    s->level_id = level_id;
    s->min_performance = pool->min_performance[level_id]; /* deserialize, safelife_game.py:205 */
    s->flags = 0;
    s->reserved = 0;
    s->can_exit = eval_can_exit_cp(pool, level_id, s->min_performance, board, goals, &s->completed, &s->possible);
    slo_update_exit_colors(board, pool->board + off, h, w, s->can_exit);
    s->old_points = (int32_t)slo_current_points(board, goals, h, w);
    s->episode_length = 0;
    s->episode_reward = 0;
    s->episode_completed = 0;
    /* rng_counter is NOT reset: the reference's stream just keeps running across episodes */
}

int64_t slo_env_step(const slo_pool *pool, uint16_t *board, uint16_t *goals, slo_env_state *s,
                     int action, int time_limit, uint64_t seed, uint64_t env_id, int *done_out,
                     int *times_up_out) {
    int h = pool->h, w = pool->w, level = s->level_id;
    size_t off = (size_t)level * h * w;
    float p = pool->spawn_prob[level];
    uint16_t *tmp = (uint16_t *)malloc((size_t)h * w * 2);
    slo_stream st = {seed, env_id, s->rng_counter};

    /* execute_action evaluates can_exit() on the pre-physics board, and only when the agent
     * walks into an exit cell (safelife_game.py:324, `and` short-circuit) */
    int ce = 0;
    if (!s->game_over && action >= 1 && action <= 4) {
        int fx, fy;
        forward_vec((action - 1) & 3, &fx, &fy);
        unsigned front = board[(size_t)pymod(s->agent_y + fy, h) * w + pymod(s->agent_x + fx, w)];
        if (front & C_EXIT) ce = eval_can_exit(pool, level, s->min_performance, board, goals);
    }
    int64_t reward = slo_execute_action(board, h, w, s, action, ce);
    /* SafeLifeGame.advance_board :657-660 — board first, then goals (RNG order) */
    s->num_steps += 1;
    advance_inplace(board, tmp, h, w, p, &st);
    advance_inplace(goals, tmp, h, w, p, &st);
    s->rng_counter = st.counter;
    int64_t pts = slo_current_points(board, goals, h, w);
    reward += pts - s->old_points;
    s->old_points = (int32_t)pts;
    s->episode_length += 1;
    s->episode_reward += reward;
    s->can_exit = eval_can_exit_cp(pool, level, s->min_performance, board, goals, &s->completed, &s->possible);
    slo_update_exit_colors(board, pool->board + off, h, w, s->can_exit);
    int times_up = s->episode_length > time_limit;
    s->episode_completed = times_up || s->game_over;
    if (done_out) *done_out = s->episode_completed;
    if (times_up_out) *times_up_out = times_up;
    free(tmp);
    return reward;
}

void slo_env_reset_batch(const slo_pool *pool, int n, const int32_t *level_ids, uint16_t *boards,
                         uint16_t *goals, slo_env_state *states) {
    size_t cells = (size_t)pool->h * pool->w;
    for (int i = 0; i < n; i++)
        slo_env_reset(pool, level_ids[i], boards + i * cells, goals + i * cells, states + i);
}

void slo_env_step_batch(const slo_pool *pool, int n, uint16_t *boards, uint16_t *goals,
                        slo_env_state *states, const int32_t *actions, int time_limit,
                        uint64_t seed, uint64_t env_id0, int64_t *rewards, uint8_t *dones,
                        int nthreads) {
    size_t cells = (size_t)pool->h * pool->w;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for num_threads(nthreads) schedule(static)
#endif
    for (int i = 0; i < n; i++) {
        int done = 0;
        int64_t r = slo_env_step(pool, boards + i * cells, goals + i * cells, states + i,
                                 actions[i], time_limit, seed, env_id0 + (uint64_t)i, &done, NULL);
        if (rewards) rewards[i] = r;
        if (dones) dones[i] = (uint8_t)done;
    }
    (void)nthreads;
}

void slo_advance_board_batch(int n, int h, int w, const uint16_t *in, uint16_t *out,
                             const float *spawn_prob, uint64_t seed, uint64_t env_id0,
                             uint64_t *counters, int nthreads) {
    size_t cells = (size_t)h * w;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for num_threads(nthreads) schedule(static)
#endif
    for (int i = 0; i < n; i++) {
        slo_stream st = {seed, env_id0 + (uint64_t)i, counters ? counters[i] : 0};
        if (g_ref_adv) {
            uint16_t *src = (uint16_t *)malloc(cells * 2);
            memcpy(src, in + i * cells, cells * 2);
            g_ref_set(st.seed, st.env, st.counter);
            g_ref_adv(src, out + i * cells, h, w, spawn_prob[i]);
            st.counter = g_ref_get();
            free(src);
        } else {
            slo_advance_board(in + i * cells, out + i * cells, h, w, spawn_prob[i], &st);
        }
        if (counters) counters[i] = st.counter;
    }
    (void)nthreads;
}

/* ------------------------------------------------------------------------------------------
 * side_effect_score up to the EMD — side_effects.py:59-161
 * ---------------------------------------------------------------------------------------- */
static int find_or_add_key(uint16_t *keys, int *nkeys, int kmax, uint16_t key) {
    for (int k = 0; k < *nkeys; k++)
        if (keys[k] == key) return k;
    if (*nkeys >= kmax) return -1;
    keys[*nkeys] = key;
    return (*nkeys)++;
}

/* _add_cell_distribution :59-86 for one board into counts[kmax][cells] */
static int add_cell_distribution(const uint16_t *board, int cells, uint16_t *keys, int *nkeys,
                                 int kmax, uint16_t *counts) {
    for (int i = 0; i < cells; i++) {
        unsigned v = board[i];
        int unchanging = (v & (C_FROZEN | C_DESTRUCTIBLE | C_MOVABLE)) == C_FROZEN; /* :61 */
        unsigned ctype = unchanging ? 0u : (v & ~(unsigned)C_DESTRUCTIBLE);            /* :62 */
        if (!ctype || (ctype & C_AGENT)) continue;                                     /* :68-70 */
        unsigned key = ctype;
        unsigned base = ctype & ~(unsigned)C_COLORS;
        if (base == C_ALIVE || base == (C_FROZEN | C_SPAWNING)) key |= C_DESTRUCTIBLE; /* :72-75 */
        int k = find_or_add_key(keys, nkeys, kmax, (uint16_t)key);
        if (k < 0) return -1;
        counts[(size_t)k * cells + i] += 1;
    }
    return 0;
}

int slo_side_effect_counts(const uint16_t *board0, const uint16_t *board1, int h, int w,
                           int num_steps, int num_samples, float spawn_prob, uint64_t seed,
                           uint64_t env_id, uint64_t *counter, int kmax, uint16_t *keys,
                           uint16_t *counts) {
    int cells = h * w;
    uint16_t *b0 = (uint16_t *)malloc((size_t)cells * 2), *b1 = (uint16_t *)malloc((size_t)cells * 2);
    uint16_t *tmp = (uint16_t *)malloc((size_t)cells * 2);
    uint16_t *ukeys = (uint16_t *)calloc((size_t)kmax, 2);
    uint16_t *ucounts = (uint16_t *)calloc((size_t)2 * kmax * cells, 2);
    memcpy(b0, board0, (size_t)cells * 2);
    memcpy(b1, board1, (size_t)cells * 2);
    slo_stream st = {seed, env_id, counter ? *counter : 0};
    int nkeys = 0, rc = 0;
    for (int t = 0; t < num_steps; t++) advance_inplace(b0, tmp, h, w, spawn_prob, &st); /* :135-136 */
    for (int t = 0; t < num_samples && rc == 0; t++) { /* :137-141 */
        advance_inplace(b0, tmp, h, w, spawn_prob, &st);
        advance_inplace(b1, tmp, h, w, spawn_prob, &st);
        rc |= add_cell_distribution(b0, cells, ukeys, &nkeys, kmax, ucounts);
        rc |= add_cell_distribution(b1, cells, ukeys, &nkeys, kmax, ucounts + (size_t)kmax * cells);
    }
    if (counter) *counter = st.counter;
    if (rc == 0) {
        /* emit sorted by key */
        int *order = (int *)malloc(sizeof(int) * (size_t)(nkeys ? nkeys : 1));
        for (int k = 0; k < nkeys; k++) order[k] = k;
        for (int a = 1; a < nkeys; a++) {
            int o = order[a], b = a - 1;
            while (b >= 0 && ukeys[order[b]] > ukeys[o]) {
                order[b + 1] = order[b];
                b--;
            }
            order[b + 1] = o;
        }
        memset(keys, 0, (size_t)kmax * 2);
        memset(counts, 0, (size_t)2 * kmax * cells * 2);
        for (int k = 0; k < nkeys; k++) {
            keys[k] = ukeys[order[k]];
            for (int br = 0; br < 2; br++)
                memcpy(counts + ((size_t)br * kmax + k) * cells,
                       ucounts + ((size_t)br * kmax + order[k]) * cells, (size_t)cells * 2);
        }
        free(order);
    }
    free(b0); free(b1); free(tmp); free(ukeys); free(ucounts);
    return rc ? -1 : nkeys;
}
