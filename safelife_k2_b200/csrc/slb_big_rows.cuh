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

#ifndef SLB_ROWS_THREADS
#define SLB_ROWS_THREADS 256
#endif
constexpr int ROWS_THREADS = SLB_ROWS_THREADS;  // (a compile-time knob for A/B builds: 128 = four CTAs' worth of registers per SM)

inline bool rows_shape(int h, int w) {
    if (w != 64 && w != 128 && w != 256) return false;
    const int strips = (ROWS_THREADS / 32) * (32 / (w / 32));
    return h % strips == 0 && h / strips >= 2;
}

// 6 words per item for the spawn draws + the 11 source planes of every thread's bottom halo row
// (the all-stages kernel; its rank array and draw words are per item)
__host__ __device__ inline size_t rows_smem_bytes(int h, int w) {
    return ((size_t)6 * h * (w / 32) + (size_t)12 * ROWS_THREADS + (size_t)16 * ROWS_THREADS) * 4;
}
// The step kernel: four words per item (eligible word, three spawn colours), the halo block, and
// ONE region that is the threads' staging slots while the window slides (four 16-byte slots per thread) and the
// packed draw outcomes afterwards (one bit per eligible cell: at most one word per item) — 60 KB at 256 x 256.
__host__ __device__ inline size_t rows_step_smem_bytes(int h, int w) {
    const size_t items = (size_t)h * (w / 32), stage = (size_t)16 * ROWS_THREADS;
    return ((size_t)4 * items + (size_t)12 * ROWS_THREADS + (stage > items ? stage : items)) * 4;
}

struct RowsSmem {
    uint32_t *elig, *rank, *tr, *tg, *tb, *draws;  // [items] each
    uint32_t *halo;                                // [11][ROWS_THREADS]
    uint32_t *stage;                               // [4][ROWS_THREADS] uint4 (16-byte aligned)
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

// The same four loads as asynchronous copies into the thread's staging slots (one 16-byte slot per plane group,
// [group][thread]: conflict-free): no register is tied up while the row travels; row_take_async waits for
// the thread's own copies and reads the slots.
template <unsigned GROUPS>
__device__ __forceinline__ void row_issue_async(const uint32_t *planes, int items, int it, uint4 *stage) {
    const uint4 *b4 = reinterpret_cast<const uint4 *>(planes);
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(stage);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s0), "l"(b4 + it) : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s0 + 16u * ROWS_THREADS), "l"(b4 + items + it) : "memory");
    if (GROUPS & 4u)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s0 + 32u * ROWS_THREADS), "l"(b4 + 2 * items + it) : "memory");
    if (GROUPS & 8u)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s0 + 48u * ROWS_THREADS), "l"(b4 + 3 * items + it) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
}

template <unsigned GROUPS>
__device__ __forceinline__ void row_take_async(const uint4 *stage, uint4 q[4]) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    q[0] = stage[0];
    q[1] = stage[ROWS_THREADS];
    q[2] = (GROUPS & 4u) ? stage[2 * ROWS_THREADS] : z;
    q[3] = (GROUPS & 8u) ? stage[3 * ROWS_THREADS] : z;
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
// OUT: col[0..2] receive the row's NEW r / g / b planes (what the fused score of the step kernel reads).
template <unsigned GROUPS, bool OUT = false>
__device__ __forceinline__ uint32_t rows_rule(uint32_t *planes, int items, int it, const RowRegs &u, RowRegs &m,
                                              const RowRegs &d, const RowLanes &L, const RowsSmem &sm,
                                              uint32_t *col = nullptr) {
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
    if (OUT) {  // births are flips: without a flip the planes are the old ones
        col[0] = (m.P[9] & ~flip) | nR; col[1] = (m.P[10] & ~flip) | nG; col[2] = (m.P[11] & ~flip) | nB;
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
__global__ void __launch_bounds__(ROWS_THREADS, 512 / ROWS_THREADS)
big_rows_env_kernel(slb_pool pool, BigArgs a) {
    extern __shared__ __align__(16) uint32_t big_smem[];
    big_env_body<WPR>(pool, a, big_smem);
}


// =====================================================================================================
// The fused step (execute_action + advance_board x2 + score + exit colours + bookkeeping) for the sliding-
// window shapes, restructured around what profiles/r02/b_bigrows_ncu.txt showed: the kernel above issues on
// 29 % of the cycles with 16 resident warps per SM — it waits on memory — so this variant removes exposed
// latencies instead of instructions.  OPT bits (compile time; SLB_BIG_OPT picks an instantiation):
//   RO_PREFETCH  every thread requests its share of the env's board / goals lines into L2 before anything
//                else happens (state load, action, barriers): the first rows and every later pass hit L2
//   RO_HOIST     a row's four 16-byte loads are issued at the TOP of the body two rows ahead and land while
//                the rule of the current row is evaluated (the kernel above issues them at the end of a body
//                and uses them at once; it only prefetches into L1)
//   RO_FIXB      the spawn fix-up reads the planes of four items before it modifies and stores any of them
//                (the loop above is load -> store -> load ...: one L2 round trip per item)
//   RO_SCORE     when goals hold no spawner, current_points / performance_ratio are accumulated inside the
//                goals' sliding window (new goal colours are in registers; the board's two plane groups are
//                requested at the top of the body) instead of a separate pass that reloads both boards
//   RO_PDL       programmatic dependent launch: see the griddepcontrol pair in rows_step_body
// Always: slb_env_state lives in shared memory for the whole step (one 80-byte load, one store), the level's
// exit mask is parked in shared memory while the physics runs, spawn ranks come from ONE CTA-wide scan (a
// thread owns items / 256 consecutive items) instead of one scan per 256 items.
// =====================================================================================================
enum : int { RO_PREFETCH = 1, RO_HOIST = 2, RO_FIXB = 4, RO_SCORE = 8, RO_NEW = 16, RO_PDL = 32 };

struct RowsAcc { int pts, cur, pos; };

__device__ __forceinline__ void rows_score_add(RowsAcc &acc, const uint4 &b0, const uint4 &b1, uint32_t g9,
                                               uint32_t g10, uint32_t g11) {
    Geo full;
    full.h = 0; full.w = 32; full.up = full.dn = 0; full.wmask = 0xffffffffu; full.hib = 0x80000000u; full.row = true;
    uint32_t B[16], G[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { B[k] = 0u; G[k] = 0u; }
    B[0] = b0.x; B[4] = b0.y; B[9] = b0.z; B[10] = b0.w;
    B[11] = b1.x; B[3] = b1.y; B[2] = b1.z; B[15] = b1.w;
    G[9] = g9; G[10] = g10; G[11] = g11;
    const RowScore rs = score_planes(B, G, full);
    acc.pts += rs.points; acc.cur += rs.current; acc.pos += rs.possible;
}

// One advance_board step of the CTA's board, in place.  SCORE (goals only, GROUPS without spawners): `other`
// = the env's board, already final; the row's score terms are added to acc.
template <unsigned GROUPS, int WPR, int OPT, bool SCORE>
__device__ __forceinline__ uint64_t rows_physics2(uint32_t *planes, int h, const RowsSmem &sm, float spawn_prob,
                                                  uint64_t seed, uint64_t env, uint64_t counter,
                                                  const uint32_t *other, RowsAcc &acc) {
    static_assert(!SCORE || !(GROUPS & 8u), "the fused score needs goals without spawners");
    constexpr int SUB = 32 / WPR;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = lane & (WPR - 1), q = lane / WPR;
    const int strips = (ROWS_THREADS / 32) * SUB, R = h / strips;
    const int r0 = (warp * SUB + q) * R;
    const int items = h * WPR;
    RowLanes L;
    L.ll = (lane & ~(WPR - 1)) | ((lane - 1) & (WPR - 1));
    L.lr = (lane & ~(WPR - 1)) | ((lane + 1) & (WPR - 1));
    RowRegs A, B, C;
    uint32_t *halo = sm.halo + threadIdx.x;
    {
        uint4 qa[4], qb[4], qc[4], qh[4];
        row_issue<GROUPS>(planes, items, ((r0 == 0) ? h - 1 : r0 - 1) * WPR + j, qa);
        row_issue<GROUPS>(planes, items, r0 * WPR + j, qb);
        row_issue<GROUPS>(planes, items, (r0 + 1) * WPR + j, qc);
        row_issue<GROUPS>(planes, items, ((r0 + R == h) ? 0 : r0 + R) * WPR + j, qh);
        if (R > 2) row_prefetch<GROUPS>(planes, items, (r0 + 2) * WPR + j);
        row_fill(A, qa); row_fill(B, qb); row_fill(C, qc);
        RowRegs t;
        row_fill(t, qh);
#pragma unroll
        for (int k = 0; k < 11; k++) halo[k * ROWS_THREADS] = t.s[k];
    }
    __syncthreads();  // every halo row has been read: stores may begin
    uint32_t seen = 0;
    const uint4 *o4 = reinterpret_cast<const uint4 *>(other);
    uint4 *stage = reinterpret_cast<uint4 *>(sm.stage) + threadIdx.x;  // this thread's slot of each group's block
#define SLB_ROW_BODY2(UP, MID, DN, K)                                                                    \
    if ((K) < R) {                                                                                       \
        const int r_ = r0 + (K);                                                                         \
        const bool own_ = (K) + 2 < R;                                                                   \
        const int it_ = r_ * WPR + j;                                                                    \
        uint4 nq_[4];                                                                                    \
        if (OPT & RO_HOIST) {                                                                            \
            if (own_) row_issue_async<GROUPS>(planes, items, (r_ + 2) * WPR + j, stage);                 \
            if ((K) + 3 < R) row_prefetch<GROUPS>(planes, items, (r_ + 3) * WPR + j);                    \
        } else if (own_) {                                                                               \
            row_prefetch<GROUPS>(planes, items, (r_ + 2) * WPR + j);                                     \
        }                                                                                                \
        uint4 ob0_, ob1_;                                                                                \
        if (SCORE) { ob0_ = o4[it_]; ob1_ = o4[items + it_]; }                                           \
        uint32_t col_[3];                                                                                \
        const uint32_t e_ = rows_rule<GROUPS, SCORE>(planes, items, it_, UP, MID, DN, L, sm, col_);      \
        if (GROUPS & 8u) { sm.elig[it_] = e_; seen |= e_; }                                              \
        if (SCORE) rows_score_add(acc, ob0_, ob1_, col_[0], col_[1], col_[2]);                           \
        if (own_) {                                                                                      \
            if (OPT & RO_HOIST) row_take_async<GROUPS>(stage, nq_);                                      \
            else row_issue<GROUPS>(planes, items, (r_ + 2) * WPR + j, nq_);                              \
            row_fill(UP, nq_);                                                                           \
        } else {                                                                                         \
            _Pragma("unroll") for (int k_ = 0; k_ < 11; k_++) UP.s[k_] = halo[k_ * ROWS_THREADS];        \
        }                                                                                                \
    }
    for (int k = 0; k < R; k += 3) {
        SLB_ROW_BODY2(A, B, C, k)
        SLB_ROW_BODY2(B, C, A, k + 1)
        SLB_ROW_BODY2(C, A, B, k + 2)
    }
#undef SLB_ROW_BODY2
    if (!(GROUPS & 8u)) return counter;
    if (!__syncthreads_or(seen != 0u)) return counter;  // uniform; also orders the parked words
    // ---- spawn draws: ONE CTA-wide scan (thread t owns the `per` consecutive items from t * per on, in
    // raster order), one Philox evaluation per lane packed by votes, fix-up by the same owner ----
    const int per = items / ROWS_THREADS, i0 = threadIdx.x * per;
    int cnt = 0;
    for (int k = 0; k < per; k++) cnt += __popc(sm.elig[i0 + k]);
    int total;
    int rank = big_scan(cnt, sm.warp_tot, total);
    {
        const uint64_t thr = spawn_threshold(spawn_prob);
        const int nwarps = ROWS_THREADS >> 5;
        for (int w0 = warp; w0 * 32 < total; w0 += nwarps) {
            const int dd = w0 * 32 + lane;
            const bool hit = dd < total && stream_u53(seed, env, counter + (uint64_t)dd) < thr;
            const uint32_t word = __ballot_sync(FULL, hit);
            if (lane == 0) sm.draws[w0] = word;
        }
    }
    __syncthreads();
    uint4 *b4 = reinterpret_cast<uint4 *>(planes);
    const int nw = (total + 31) >> 5;
    constexpr int BATCH = (OPT & RO_FIXB) ? 4 : 1;
    for (int k0 = 0; k0 < per; k0 += BATCH) {
        uint32_t sp[BATCH];
#pragma unroll
        for (int u = 0; u < BATCH; u++) {
            sp[u] = 0u;
            if (k0 + u >= per) continue;
            uint32_t e = sm.elig[i0 + k0 + u];
            if (!e) continue;
            const int w0 = rank >> 5;
            const uint32_t lo = sm.draws[w0];
            const uint32_t hi = (w0 + 1 < nw) ? sm.draws[w0 + 1] : 0u;
            uint32_t bits = __funnelshift_r(lo, hi, (uint32_t)rank & 31u);  // first eligible cell in bit 0
            rank += __popc(e);
            uint32_t spawn = 0;
            while (e) {
                const uint32_t bit = e & (0u - e);
                e ^= bit;
                if (bits & 1u) spawn |= bit;
                bits >>= 1;
            }
            sp[u] = spawn;
        }
        uint4 v0[BATCH], v1[BATCH], v2[BATCH], v3[BATCH];
#pragma unroll
        for (int u = 0; u < BATCH; u++) {
            if (!sp[u]) continue;
            const int it = i0 + k0 + u;
            v0[u] = b4[it];
            v1[u] = b4[items + it];
            if (GROUPS & 4u) v2[u] = b4[2 * items + it];
            v3[u] = b4[3 * items + it];
        }
#pragma unroll
        for (int u = 0; u < BATCH; u++) {
            const uint32_t spawn = sp[u];
            if (!spawn) continue;
            const int it = i0 + k0 + u;
            const uint32_t tR = sm.tr[it], tG = sm.tg[it], tB = sm.tb[it];
            const uint32_t keep = ~spawn;  // spawned cells were dead: their residue bits are discarded
            uint4 v = v0[u];
            b4[it] = make_uint4((v.x & keep) | spawn, v.y & keep, (v.z & keep) | (spawn & tR), (v.w & keep) | (spawn & tG));
            v = v1[u];
            b4[items + it] = make_uint4((v.x & keep) | (spawn & tB), (v.y & keep) | spawn, v.z & keep, v.w & keep);
            if (GROUPS & 4u) {
                v = v2[u];
                if (v.x | v.y | v.z | v.w) b4[2 * items + it] = make_uint4(v.x & keep, v.y & keep, v.z & keep, v.w & keep);
            }
            v = v3[u];
            if (v.x | v.y | v.z | v.w) b4[3 * items + it] = make_uint4(v.x & keep, v.y & keep, v.z & keep, v.w & keep);
        }
    }
    __syncthreads();
    return counter + (uint64_t)total;
}

// the separate score pass (goals with spawners, or RO_SCORE off): two items' loads in flight per thread
__device__ __forceinline__ void rows_score_pass(const uint32_t *board, const uint32_t *goals, int items, RowsAcc &acc) {
    const uint4 *b4 = reinterpret_cast<const uint4 *>(board), *g4 = reinterpret_cast<const uint4 *>(goals);
    for (int it = threadIdx.x; it < items; it += 2 * ROWS_THREADS) {
        const int i2 = it + ROWS_THREADS;
        const bool two = i2 < items;
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        const uint4 a0 = b4[it], a1 = b4[items + it], ga0 = g4[it], ga1 = g4[items + it];
        const uint4 c0 = two ? b4[i2] : z, c1 = two ? b4[items + i2] : z;
        const uint4 gc0 = two ? g4[i2] : z, gc1 = two ? g4[items + i2] : z;
        rows_score_add(acc, a0, a1, ga0.z, ga0.w, ga1.x);
        if (two) rows_score_add(acc, c0, c1, gc0.z, gc0.w, gc1.x);
    }
}

template <int WPR, int OPT>
__device__ __forceinline__ void rows_step_body(const slb_pool &pool, const BigArgs &a, uint32_t *big_smem) {
    __shared__ int scratch[96];
    __shared__ int warp_tot[32];
    __shared__ __align__(16) slb_env_state shst;
    __shared__ int sh_act, sh_baseline, sh_done, sh_reward;
    __shared__ float sh_prob;
    const int i = blockIdx.x;
    const BigGeo g = big_geo(pool.h, pool.w);
    const int items = g.items;
    RowsSmem rs;
    // (the single scan keeps ranks in registers: no rank array here)
    rs.elig = big_smem; rs.rank = nullptr; rs.tr = rs.elig + items; rs.tg = rs.tr + items;
    rs.tb = rs.tg + items; rs.halo = rs.tb + items;
    rs.stage = rs.halo + 12 * ROWS_THREADS;  // (11 halo planes, padded to keep the slots 16-byte aligned)
    rs.draws = rs.stage;  // the draws are evaluated after the window's last staged row has been taken
    rs.warp_tot = warp_tot;
    const unsigned gb = pool.plane_groups[0], gg = pool.plane_groups[1];
    const size_t words = (size_t)16 * items;
    uint32_t *board = a.board + i * words, *goals = a.goals + i * words;
    slb_env_state *st = a.state + i;
    // RO_PDL (the launch carries the programmatic-serialization attribute): the next launch on the stream may
    // take SM slots as ours free up; its CTAs request their env's lines into L2 and then wait for this grid to
    // complete before their first read (L2 is the point of coherence: a prefetched line reads back current)
    if (OPT & RO_PDL) asm volatile("griddepcontrol.launch_dependents;");
    if (OPT & RO_PREFETCH) {  // 128-byte lines of the groups that exist: items / 8 lines per group
        const int lpg = items >> 3;
        for (int k = threadIdx.x; k < 4 * lpg; k += ROWS_THREADS) {
            const unsigned grp = 1u << (k / lpg);
            if (gb & grp) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(board) + (size_t)k * 128));
            if (gg & grp) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(goals) + (size_t)k * 128));
        }
    }
    if (OPT & RO_PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x < 5)
        reinterpret_cast<uint4 *>(&shst)[threadIdx.x] = reinterpret_cast<const uint4 *>(st)[threadIdx.x];
    if (threadIdx.x == 32) sh_act = a.actions[i];
    __syncthreads();
    const int level = shst.level_id;
    const uint64_t env = a.env_id0 + (uint64_t)i;
    const int per = items / ROWS_THREADS, i0 = threadIdx.x * per;
    // which of this thread's items hold exit cells (a level has a handful): looked up now, while nothing waits
    // for it, so that the repaint at the end of the step touches the exit mask only for those
    const uint32_t *e0 = pool.exit_mask + (size_t)level * items;
    uint32_t exit_items = 0;
    if (per <= 32) {
        for (int k = 0; k < per; k++) exit_items |= (e0[i0 + k] != 0u) ? (1u << k) : 0u;
    } else {
        exit_items = 0xffffffffu;  // (more than 32 items per thread: every item is looked at again)
    }
    if (threadIdx.x == 32) { sh_prob = pool.spawn_prob[level]; sh_baseline = pool.baseline[level]; }
    if (threadIdx.x < 32) {
        int ax = shst.agent_x, ay = shst.agent_y, o = shst.orientation, go = shst.game_over;
        const int r = big_action(board, g, gb, pool.w, ax, ay, o, go, sh_act, shst.can_exit);
        __syncwarp();
        if (threadIdx.x == 0) {
            shst.agent_x = ax; shst.agent_y = ay; shst.orientation = o; shst.game_over = go;
            sh_reward = r;
        }
    }
    __syncthreads();
    uint64_t counter = shst.rng_counter;
    const float p = sh_prob;
    RowsAcc acc = {0, 0, 0};
    // board first, then goals: the order the stream is consumed in (safelife_game.py:657-660)
    if (gb & 8u) counter = rows_physics2<15u, WPR, OPT, false>(board, g.h, rs, p, a.seed, env, counter, nullptr, acc);
    else if (gb & 4u) counter = rows_physics2<7u, WPR, OPT, false>(board, g.h, rs, p, a.seed, env, counter, nullptr, acc);
    else counter = rows_physics2<3u, WPR, OPT, false>(board, g.h, rs, p, a.seed, env, counter, nullptr, acc);
    __syncthreads();
    bool scored = false;
    if (gg & 8u) counter = rows_physics2<15u, WPR, OPT, false>(goals, g.h, rs, p, a.seed, env, counter, nullptr, acc);
    else if (OPT & RO_SCORE) {
        scored = true;
        if (gg & 4u) counter = rows_physics2<7u, WPR, OPT, true>(goals, g.h, rs, p, a.seed, env, counter, board, acc);
        else counter = rows_physics2<3u, WPR, OPT, true>(goals, g.h, rs, p, a.seed, env, counter, board, acc);
    } else {
        if (gg & 4u) counter = rows_physics2<7u, WPR, OPT, false>(goals, g.h, rs, p, a.seed, env, counter, nullptr, acc);
        else counter = rows_physics2<3u, WPR, OPT, false>(goals, g.h, rs, p, a.seed, env, counter, nullptr, acc);
    }
    if (!scored) {
        __syncthreads();
        rows_score_pass(board, goals, items, acc);
    }
    int pts = acc.pts, cur = acc.cur, pos = acc.pos;
    big_sum3(pts, cur, pos, scratch);  // (its barriers also order the score's reads before the repaint)
    const int completed = cur - sh_baseline, possible = pos - sh_baseline;
    const int ce = can_exit_rule(shst.min_performance, completed, possible);
    // update_exit_colors (safelife_game.py:531-537)
    for (int k = 0; exit_items != 0u && k < per; k++) {
        if (per <= 32 && !((exit_items >> k) & 1u)) continue;
        const uint32_t m = __ldg(e0 + i0 + k);
        if (!m) continue;
        uint32_t P[16];
        big_load(board, g, i0 + k, gb, P);
#pragma unroll
        for (int b = 0; b < 16; b++) P[b] &= ~m;
        P[4] |= m; P[8] |= m;
        if (ce) P[9] |= m;
        big_store(board, g, i0 + k, gb, P);
    }
    if (threadIdx.x == 0) {
        shst.rng_counter = counter; shst.num_steps += 1;
        shst.can_exit = ce; shst.completed = completed; shst.possible = possible;
        shst.flags &= ~SLB_ENV_RESCORE;
        const int reward = sh_reward + pts - shst.old_points;
        const int ep_len = shst.episode_length + 1;
        const int done = (ep_len > a.time_limit) || shst.game_over;
        shst.episode_length = ep_len;
        shst.episode_completed = done;
        shst.old_points = pts;
        shst.episode_reward += reward;
        if (a.reward_out) a.reward_out[i] = reward;
        if (a.done_out) a.done_out[i] = (uint8_t)done;
        sh_done = done;
    }
    __syncthreads();
    if (threadIdx.x < 5)
        reinterpret_cast<uint4 *>(st)[threadIdx.x] = reinterpret_cast<const uint4 *>(&shst)[threadIdx.x];
    if (a.auto_reset_stride > 0 && sh_done) {
        __syncthreads();
        big_reset(pool, (level + a.auto_reset_stride) % pool.n_levels, g, gb, gg, board, goals, st, scratch);
    }
}

template <int WPR, int OPT>
__global__ void __launch_bounds__(ROWS_THREADS, 512 / ROWS_THREADS)
big_rows_step_kernel(slb_pool pool, BigArgs a) {
    extern __shared__ __align__(16) uint32_t big_smem[];
    rows_step_body<WPR, OPT>(pool, a, big_smem);
}

}  // namespace slb
