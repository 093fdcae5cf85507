// slb_big_rows.cuh — advance_board for big resident boards (slb_big.cuh layout) as a REGISTER SLIDING WINDOW:
// no shared-memory snapshot of the board at all.
//
// Boards whose rows are WPR = 2, 4 or 8 words (W = 64, 128, 256): a CTA of 256 threads owns one env.  Lane
// l of a warp owns word-column j = l % WPR of sub-strip q = l / WPR; warp w, sub-strip q is strip
// s = w * (32 / WPR) + q of 8 * 32 / WPR strips, R = H / strips consecutive rows each.  A thread walks down
// its strip with the 11 neighbourhood source planes of rows r-1, r, r+1 in registers: every row of the board
// is loaded from HBM exactly once per step, straight into the registers the rule is evaluated on (4 coalesced
// 16-byte loads per row and thread; a warp reads whole 128-byte row segments), the vertical 3-row sums need
// no memory access, and the left / right neighbour words are the neighbouring lanes' reduced words (segmented
// shuffles = the torus).  In-place update: a thread loads the old rows just above and below its strip BEFORE
// the CTA barrier that precedes the first store; rows inside the strip are only ever written by their owner,
// after it has read them.  A row's lines are requested (L1 prefetch) two rows ahead of its evaluation.
//
// Spawn draws need the raster-order rank of every eligible cell of the board: the eligible words and, for
// items that have any, the spawn colours are parked in shared memory (6 words per item), scanned CTA-wide
// and applied by a fix-up pass, as in slb_big.cuh.
#pragma once
#include "slb_big.cuh"

namespace slb {

constexpr int ROWS_THREADS = 256;

inline bool rows_shape(int h, int w) {
    if (w != 64 && w != 128 && w != 256) return false;
    const int strips = (ROWS_THREADS / 32) * (32 / (w / 32));
    return h % strips == 0 && h / strips >= 2;
}

// 6 words per item for the spawn draws + the 11 source planes of every thread's bottom halo row
__host__ __device__ inline size_t rows_smem_bytes(int h, int w) {
    return ((size_t)6 * h * (w / 32) + (size_t)11 * ROWS_THREADS) * 4;
}

struct RowsSmem {
    uint32_t *elig, *rank, *tr, *tg, *tb, *draws;  // [items] each
    uint32_t *halo;                                // [11][ROWS_THREADS]
    int *warp_tot;
};

// one row of one word-column: the raw planes (for the write-back) and the 11 source planes
struct RowRegs {
    uint32_t P[16];
    uint32_t s[11];
};

template <unsigned GROUPS>
__device__ __forceinline__ void row_issue(const uint32_t *planes, int items, int it, uint4 q[4]) {
    const uint4 *b4 = reinterpret_cast<const uint4 *>(planes);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    q[0] = b4[it];
    q[1] = b4[items + it];
    q[2] = (GROUPS & 4u) ? b4[2 * items + it] : z;
    q[3] = (GROUPS & 8u) ? b4[3 * items + it] : z;
}

// request a row's lines into L1 ahead of its loads (no registers are tied up while it travels)
template <unsigned GROUPS>
__device__ __forceinline__ void row_prefetch(const uint32_t *planes, int items, int it) {
    const uint4 *b4 = reinterpret_cast<const uint4 *>(planes);
    asm volatile("prefetch.global.L1 [%0];" ::"l"(b4 + it));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(b4 + items + it));
    if (GROUPS & 4u) asm volatile("prefetch.global.L1 [%0];" ::"l"(b4 + 2 * items + it));
    if (GROUPS & 8u) asm volatile("prefetch.global.L1 [%0];" ::"l"(b4 + 3 * items + it));
}

__device__ __forceinline__ void row_fill(RowRegs &r, const uint4 q[4]) {
    uint32_t *P = r.P;
    P[0] = q[0].x; P[4] = q[0].y; P[9] = q[0].z; P[10] = q[0].w;
    P[11] = q[1].x; P[3] = q[1].y; P[2] = q[1].z; P[15] = q[1].w;
    P[1] = q[2].x; P[5] = q[2].y; P[6] = q[2].z; P[8] = q[2].w;
    P[7] = q[3].x; P[12] = q[3].y; P[13] = q[3].z; P[14] = q[3].w;
    const uint32_t a = P[0], sp = P[7];
    r.s[0] = a; r.s[1] = P[5]; r.s[2] = P[6]; r.s[3] = sp;
    r.s[4] = a & (P[3] | P[8]);  // D2 = destructible | exit (advance_board.c:44-46)
    r.s[5] = a & P[9]; r.s[6] = a & P[10]; r.s[7] = a & P[11];
    r.s[8] = sp & P[9]; r.s[9] = sp & P[10]; r.s[10] = sp & P[11];
}

// left / right neighbour lanes inside this lane's row segment
struct RowLanes { int ll, lr; };

template <int K>
__device__ __forceinline__ uint32_t rw_or(const RowRegs &u, const RowRegs &m, const RowRegs &d, const RowLanes &L) {
    const uint32_t c = or3(u.s[K], m.s[K], d.s[K]);
    const uint32_t l = __shfl_sync(0xffffffffu, c, L.ll), r = __shfl_sync(0xffffffffu, c, L.lr);
    return or3(__funnelshift_l(l, c, 1), c, __funnelshift_r(c, r, 1));
}

template <int K>
__device__ __forceinline__ ColSums rw_cols(const RowRegs &u, const RowRegs &m, const RowRegs &d, const RowLanes &L) {
    const uint32_t s_m = xor3(u.s[K], m.s[K], d.s[K]), c_m = maj3(u.s[K], m.s[K], d.s[K]);
    const uint32_t s_l = __shfl_sync(0xffffffffu, s_m, L.ll), s_r = __shfl_sync(0xffffffffu, s_m, L.lr);
    const uint32_t c_l = __shfl_sync(0xffffffffu, c_m, L.ll), c_r = __shfl_sync(0xffffffffu, c_m, L.lr);
    ColSums c;
    c.sl = __funnelshift_l(s_l, s_m, 1); c.sm = s_m; c.sr = __funnelshift_r(s_m, s_r, 1);
    c.cl = __funnelshift_l(c_l, c_m, 1); c.cm = c_m; c.cr = __funnelshift_r(c_m, c_r, 1);
    return c;
}

template <int K>
__device__ __forceinline__ uint32_t rw_two(const RowRegs &u, const RowRegs &m, const RowRegs &d, const RowLanes &L) {
    const ColSums c = rw_cols<K>(u, m, d, L);
    return or3(c.cl, c.cm, c.cr) | maj3(c.sl, c.sm, c.sr);
}

// The rule for row `mid` (rows up / dn are its vertical neighbours), advance_board.c:89-119 in the closed
// form of SURVEY Appendix A with the lazy-but-exact evaluation of big_rule_items<true>: every lane of the
// warp calls; summaries that no cell of the warp's rows can need are decided by votes.  Writes the row back
// if any cell flipped; returns the row's eligible-for-spawning word (parked in shared memory by the caller).
template <unsigned GROUPS>
__device__ __forceinline__ uint32_t rows_rule(uint32_t *planes, int items, int it, const RowRegs &u, RowRegs &m,
                                              const RowRegs &d, const RowLanes &L, const RowsSmem &sm) {
    const uint32_t a = m.P[0], fz = m.P[4];
    uint32_t is3, is4;
    {
        const ColSums c = rw_cols<0>(u, m, d, L);
        const uint32_t s0 = xor3(c.sl, c.sm, c.sr), c0 = maj3(c.sl, c.sm, c.sr);
        const uint32_t s1 = xor3(c.cl, c.cm, c.cr), c1 = maj3(c.cl, c.cm, c.cr);
        const uint32_t n1 = c0 ^ s1, k = c0 & s1, n2 = c1 ^ k;  // n = s0 + 2 n1 + 4 n2 (+ 8)
        is3 = s0 & n1 & ~n2;
        is4 = ~s0 & ~n1 & n2;
    }
    const uint32_t die0 = a & ~(fz | is3 | is4);
    const uint32_t open0 = ~(a | fz);
    uint32_t PR = 0, IH = 0;
    if ((GROUPS & 4u) && __any_sync(0xffffffffu, (u.s[1] | u.s[2] | m.s[1] | m.s[2] | d.s[1] | d.s[2]) != 0u)) {
        PR = rw_or<1>(u, m, d, L);
        IH = rw_or<2>(u, m, d, L);
    }
    const uint32_t die = die0 & ~PR;
    const uint32_t open = open0 & ~IH;
    const uint32_t birth = open & is3;
    const bool want_s = (GROUPS & 8u) && __any_sync(0xffffffffu, (u.s[3] | m.s[3] | d.s[3]) != 0u);
    uint32_t elig = 0;
    if (want_s) elig = open & ~is3 & rw_or<3>(u, m, d, L);
    const uint32_t flip = die | birth;
    const bool any_birth = __any_sync(0xffffffffu, birth != 0u);
    const bool any_elig = want_s && __any_sync(0xffffffffu, elig != 0u);
    uint32_t nD = 0, nR = 0, nG = 0, nB = 0;
    if (any_birth || any_elig) {
        uint32_t tR = rw_two<5>(u, m, d, L);
        uint32_t tG = rw_two<6>(u, m, d, L);
        uint32_t tB = rw_two<7>(u, m, d, L);
        if (want_s) {
            tR |= rw_or<8>(u, m, d, L);
            tG |= rw_or<9>(u, m, d, L);
            tB |= rw_or<10>(u, m, d, L);
        }
        if (any_birth) nD = birth & rw_two<4>(u, m, d, L);
        nR = birth & tR; nG = birth & tG; nB = birth & tB;
        if (any_elig && elig) {  // the colours a spawned cell would take (advance_board.c:110-117)
            sm.tr[it] = tR; sm.tg[it] = tG; sm.tb[it] = tB;
        }
    }
    if (flip) {
        const uint32_t keep = ~flip;
        const bool g2 = (GROUPS & 4u) && (m.P[1] | m.P[5] | m.P[6] | m.P[8]) != 0u;
        const bool g3 = (GROUPS & 8u) && (m.P[7] | m.P[12] | m.P[13] | m.P[14]) != 0u;
        uint4 *b4 = reinterpret_cast<uint4 *>(planes);
        b4[it] = make_uint4((m.P[0] & keep) | birth, m.P[4] & keep, (m.P[9] & keep) | nR, (m.P[10] & keep) | nG);
        b4[items + it] = make_uint4((m.P[11] & keep) | nB, (m.P[3] & keep) | nD, m.P[2] & keep, m.P[15] & keep);
        // a group that is all zero in this item stays zero (a flip only clears bits in it)
        if (g2) b4[2 * items + it] = make_uint4(m.P[1] & keep, m.P[5] & keep, m.P[6] & keep, m.P[8] & keep);
        if (g3) b4[3 * items + it] = make_uint4(m.P[7] & keep, m.P[12] & keep, m.P[13] & keep, m.P[14] & keep);
    }
    return elig;
}

// One advance_board step of the CTA's board, in place (advance_board.c:34-120).  Every thread of the CTA
// must call.  Returns the env's updated draw counter (same in all threads).
template <unsigned GROUPS, int WPR>
__device__ __forceinline__ uint64_t rows_physics(uint32_t *planes, int h, const RowsSmem &sm, float spawn_prob,
                                                 uint64_t seed, uint64_t env, uint64_t counter) {
    constexpr int SUB = 32 / WPR;                       // sub-strips per warp
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = lane & (WPR - 1), q = lane / WPR;
    const int strips = (ROWS_THREADS / 32) * SUB, R = h / strips;
    const int r0 = (warp * SUB + q) * R;
    const int items = h * WPR;
    RowLanes L;
    L.ll = (lane & ~(WPR - 1)) | ((lane - 1) & (WPR - 1));
    L.lr = (lane & ~(WPR - 1)) | ((lane + 1) & (WPR - 1));
    RowRegs A, B, C;
    uint32_t *halo = sm.halo + threadIdx.x;  // this thread's column of the halo block
    {
        uint4 qa[4], qb[4], qc[4], qh[4];
        row_issue<GROUPS>(planes, items, ((r0 == 0) ? h - 1 : r0 - 1) * WPR + j, qa);
        row_issue<GROUPS>(planes, items, r0 * WPR + j, qb);
        row_issue<GROUPS>(planes, items, (r0 + 1) * WPR + j, qc);
        row_issue<GROUPS>(planes, items, ((r0 + R == h) ? 0 : r0 + R) * WPR + j, qh);
        row_fill(A, qa); row_fill(B, qb); row_fill(C, qc);
        RowRegs t;
        row_fill(t, qh);
#pragma unroll
        for (int k = 0; k < 11; k++) halo[k * ROWS_THREADS] = t.s[k];
    }
    __syncthreads();  // every halo row has been read: stores may begin
    uint32_t seen = 0;
    // rows r0 + k: (up, mid, dn) rotate through (A, B, C); the row two below is prefetched at the top of a
    // body and loaded at its end, into the registers of the row that just left the window
#define SLB_ROW_BODY(UP, MID, DN, K)                                                                     \
    if ((K) < R) {                                                                                       \
        const int r_ = r0 + (K);                                                                         \
        const bool own_ = (K) + 2 < R;                                                                   \
        if (own_) row_prefetch<GROUPS>(planes, items, (r_ + 2) * WPR + j);                               \
        const int it_ = r_ * WPR + j;                                                                    \
        const uint32_t e_ = rows_rule<GROUPS>(planes, items, it_, UP, MID, DN, L, sm);                   \
        if (GROUPS & 8u) { sm.elig[it_] = e_; seen |= e_; }                                              \
        if (own_) {                                                                                      \
            uint4 nq_[4];                                                                                \
            row_issue<GROUPS>(planes, items, (r_ + 2) * WPR + j, nq_);                                   \
            row_fill(UP, nq_);                                                                           \
        } else {                                                                                           \
            _Pragma("unroll") for (int k_ = 0; k_ < 11; k_++) UP.s[k_] = halo[k_ * ROWS_THREADS];                       \
        }                                                                                                \
    }
    for (int k = 0; k < R; k += 3) {
        SLB_ROW_BODY(A, B, C, k)
        SLB_ROW_BODY(B, C, A, k + 1)
        SLB_ROW_BODY(C, A, B, k + 2)
    }
#undef SLB_ROW_BODY
    if (!(GROUPS & 8u)) return counter;
    if (!__syncthreads_or(seen != 0u)) return counter;  // uniform; also orders the parked words
    // ---- spawn draws: ranks by a CTA-wide scan, one Philox evaluation per lane, outcomes packed by votes ----
    const int T = blockDim.x;
    int total = 0;
    for (int base = 0; base < items; base += T) {  // uniform trip count
        const int it = base + threadIdx.x;
        const int cnt = (it < items) ? __popc(sm.elig[it]) : 0;
        int chunk_total;
        const int rank = big_scan(cnt, sm.warp_tot, chunk_total);
        if (it < items) sm.rank[it] = (uint32_t)(total + rank);
        total += chunk_total;
    }
    {
        const uint64_t thr = spawn_threshold(spawn_prob);
        const int nwarps = T >> 5;
        for (int w0 = warp; w0 * 32 < total; w0 += nwarps) {
            const int dd = w0 * 32 + lane;
            const bool hit = dd < total && stream_u53(seed, env, counter + (uint64_t)dd) < thr;
            const uint32_t word = __ballot_sync(FULL, hit);
            if (lane == 0) sm.draws[w0] = word;
        }
    }
    __syncthreads();
    for (int it = threadIdx.x; it < items; it += T) {
        uint32_t e = sm.elig[it];
        if (!e) continue;
        const uint32_t rank = sm.rank[it];
        const int w0 = (int)(rank >> 5), nw = (total + 31) >> 5;
        const uint32_t lo = sm.draws[w0];
        const uint32_t hi = (w0 + 1 < nw) ? sm.draws[w0 + 1] : 0u;
        uint32_t bits = __funnelshift_r(lo, hi, rank & 31u);  // this item's outcomes, first eligible cell in bit 0
        uint32_t spawn = 0;
        while (e) {
            const uint32_t bit = e & (0u - e);
            e ^= bit;
            if (bits & 1u) spawn |= bit;
            bits >>= 1;
        }
        if (!spawn) continue;
        const uint32_t tR = sm.tr[it], tG = sm.tg[it], tB = sm.tb[it];
        uint4 *b4 = reinterpret_cast<uint4 *>(planes);
        const uint32_t keep = ~spawn;  // spawned cells were dead: their residue bits are discarded
        uint4 v = b4[it];
        b4[it] = make_uint4((v.x & keep) | spawn, v.y & keep, (v.z & keep) | (spawn & tR), (v.w & keep) | (spawn & tG));
        v = b4[items + it];
        b4[items + it] = make_uint4((v.x & keep) | (spawn & tB), (v.y & keep) | spawn, v.z & keep, v.w & keep);
        if (GROUPS & 4u) {
            v = b4[2 * items + it];
            if (v.x | v.y | v.z | v.w) b4[2 * items + it] = make_uint4(v.x & keep, v.y & keep, v.z & keep, v.w & keep);
        }
        v = b4[3 * items + it];
        if (v.x | v.y | v.z | v.w) b4[3 * items + it] = make_uint4(v.x & keep, v.y & keep, v.z & keep, v.w & keep);
    }
    __syncthreads();
    return counter + (uint64_t)total;
}

// The env kernel body for big boards: one CTA per env; `ops` selects the stages.  WPR > 0: the board's rows are
// WPR words and the CTA has ROWS_THREADS threads — advance_board runs as the register sliding window above;
// WPR == 0: any big shape, shared-memory snapshot (slb_big.cuh).
template <int WPR>
__device__ __forceinline__ void big_env_body(const slb_pool &pool, const BigArgs &a, uint32_t *big_smem) {
    __shared__ int scratch[96];
    __shared__ int warp_tot[32];
    __shared__ int sh_act[5];
    const int i = blockIdx.x;
    const BigGeo g = big_geo(pool.h, pool.w);
    BigSmem sm;
    RowsSmem rs;
    if (WPR == 0) {
        sm.src = big_smem;
        sm.elig = big_smem + (size_t)BIG_SRC * g.items;
        sm.rowf = sm.elig + g.items;
        sm.warp_tot = warp_tot;
    } else {
        rs.elig = big_smem; rs.rank = rs.elig + g.items; rs.tr = rs.rank + g.items; rs.tg = rs.tr + g.items;
        rs.tb = rs.tg + g.items; rs.draws = rs.tb + g.items; rs.halo = rs.draws + g.items;
        rs.warp_tot = warp_tot;
    }
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    const size_t words = (size_t)16 * g.items;
    uint32_t *board = a.board + i * words, *goals = a.goals ? a.goals + i * words : nullptr;
    slb_env_state *st = a.state + i;
    const int ops = a.ops;

    if (ops & BIG_RESET) {
        if (a.mask && !a.mask[i]) return;
        big_reset(pool, a.level_ids[i], g, gb, gg, board, goals, st, scratch);
        return;
    }
    const int level = st->level_id;
    const uint64_t env = a.env_id0 + (uint64_t)i;
    int act_reward = 0;
    if (ops & BIG_ACTION) {
        if (threadIdx.x < 32) {
            int ax = st->agent_x, ay = st->agent_y, o = st->orientation, go = st->game_over;
            const int r = big_action(board, g, gb, pool.w, ax, ay, o, go, a.actions[i], st->can_exit);
            if (threadIdx.x == 0) { sh_act[0] = ax; sh_act[1] = ay; sh_act[2] = o; sh_act[3] = go; sh_act[4] = r; }
        }
        __syncthreads();
        act_reward = sh_act[4];
        if (threadIdx.x == 0) {
            st->agent_x = sh_act[0]; st->agent_y = sh_act[1]; st->orientation = sh_act[2]; st->game_over = sh_act[3];
            if (!(ops & BIG_BOOK) && a.reward_out) a.reward_out[i] = act_reward;
        }
    }
    uint64_t counter = st->rng_counter;
    if (ops & BIG_PHYSICS) {
        const float p = pool.spawn_prob[level];
        // board first, then goals: the order the stream is consumed in (safelife_game.py:657-660)
        if (WPR == 0) {
            counter = big_physics(board, g, gb, sm, p, a.seed, env, counter);
            __syncthreads();
            counter = big_physics(goals, g, gg, sm, p, a.seed, env, counter);
        } else {
            constexpr int W_ = WPR ? WPR : 2;
            if (gb & 8u) counter = rows_physics<15u, W_>(board, g.h, rs, p, a.seed, env, counter);
            else if (gb & 4u) counter = rows_physics<7u, W_>(board, g.h, rs, p, a.seed, env, counter);
            else counter = rows_physics<3u, W_>(board, g.h, rs, p, a.seed, env, counter);
            __syncthreads();
            if (gg & 8u) counter = rows_physics<15u, W_>(goals, g.h, rs, p, a.seed, env, counter);
            else if (gg & 4u) counter = rows_physics<7u, W_>(goals, g.h, rs, p, a.seed, env, counter);
            else counter = rows_physics<3u, W_>(goals, g.h, rs, p, a.seed, env, counter);
        }
        __syncthreads();
        if (threadIdx.x == 0) { st->rng_counter = counter; st->num_steps += 1; }
    }
    if (!(ops & (BIG_BOOK | BIG_POINTS | BIG_PAINT))) return;
    const uint32_t *e0 = pool.exit_mask + (size_t)level * g.items;
    int pts, cur, pos;
    big_score(board, goals, g, e0, false, scratch, pts, cur, pos);
    const int baseline = pool.baseline[level];
    const int completed = cur - baseline, possible = pos - baseline;
    const int ce = can_exit_rule(st->min_performance, completed, possible);
    if (ops & BIG_POINTS) {
        if (threadIdx.x == 0) {
            if (a.points) a.points[i] = pts;
            if (a.completed) a.completed[i] = completed;
            if (a.possible) a.possible[i] = possible;
            if (a.can_exit) a.can_exit[i] = (uint8_t)ce;
        }
        return;
    }
    big_paint(board, g, gb, e0, ce);
    __shared__ int sh_done;
    if (threadIdx.x == 0) {
        st->can_exit = ce; st->completed = completed; st->possible = possible;
        st->flags &= ~SLB_ENV_RESCORE;
        if (ops & BIG_BOOK) {
            const int reward = act_reward + pts - st->old_points;
            const int ep_len = st->episode_length + 1;
            const int done = (ep_len > a.time_limit) || st->game_over;
            st->episode_length = ep_len;
            st->episode_completed = done;
            st->old_points = pts;
            st->episode_reward += reward;
            if (a.reward_out) a.reward_out[i] = reward;
            if (a.done_out) a.done_out[i] = (uint8_t)done;
            sh_done = done;
        }
    }
    if (!(ops & BIG_BOOK)) return;
    __syncthreads();
    if (a.auto_reset_stride > 0 && sh_done)
        big_reset(pool, (level + a.auto_reset_stride) % pool.n_levels, g, gb, gg, board, goals, st, scratch);
}


__global__ void __launch_bounds__(BIG_THREADS, 2)
big_env_kernel(slb_pool pool, BigArgs a) {
    extern __shared__ __align__(16) uint32_t big_smem[];
    big_env_body<0>(pool, a, big_smem);
}

template <int WPR>
__global__ void __launch_bounds__(ROWS_THREADS, 2)
big_rows_env_kernel(slb_pool pool, BigArgs a) {
    extern __shared__ __align__(16) uint32_t big_smem[];
    big_env_body<WPR>(pool, a, big_smem);
}

}  // namespace slb
