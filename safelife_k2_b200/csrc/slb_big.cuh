// slb_big.cuh — resident bit-plane kernels for boards larger than one warp's 32x32 (e.g. the 256x256
// scaling configuration): W a multiple of 32, one CTA per env.
//
// Resident layout of one board (64*H*(W/32) bytes, same as the small layout with rows widened to
// W/32 words): 4 plane groups x H rows x (W/32) words x uint4; "item" = one 32-cell word of a row,
// item index = row * wpr + word, so consecutive threads read consecutive uint4.
//
// One advance_board step: phase A writes the 11 neighbourhood SOURCE planes of the whole board into
// shared memory (alive; preserving, inhibiting, spawning; alive & {D2, r, g, b}; spawning & {r, g,
// b}); after one barrier every item reads the 3x3 words around it, aligns the left/right neighbours
// with one funnel shift each and reduces the nine words with carry-save adders.  Results go
// straight back to the resident planes in global memory (an item's planes are only ever read and
// written by its own thread; neighbours come from the shared snapshot).  Spawn draws need the
// raster-order rank of every eligible cell of the board: eligible words are parked in shared
// memory, scanned CTA-wide, and a fix-up pass applies the successful draws.
#pragma once
#include "slb_resident.cuh"

namespace slb {

constexpr int BIG_THREADS = 512;
constexpr int BIG_SRC = 11;  // source planes in shared memory
constexpr size_t BIG_SMEM_LIMIT = 200 * 1024;

__host__ __device__ inline size_t big_smem_bytes(int h, int w) {
    const size_t items = (size_t)h * (w / 32);
    return (BIG_SRC + 1) * items * 4 + (size_t)h * 4;  // sources + eligible words + row flags
}

inline bool big_shape(int h, int w) {
    return !fast_shape(h, w) && h >= 3 && w >= 32 && (w % 32) == 0 && big_smem_bytes(h, w) <= BIG_SMEM_LIMIT;
}

struct BigGeo {
    int h, wpr, items;
    int wshift;  // log2(wpr) when wpr is a power of two, else -1
};

__device__ __forceinline__ BigGeo big_geo(int h, int w) {
    BigGeo g;
    g.h = h; g.wpr = w / 32; g.items = h * g.wpr;
    g.wshift = (g.wpr & (g.wpr - 1)) ? -1 : __ffs(g.wpr) - 1;
    return g;
}

__device__ __forceinline__ void big_row_word(const BigGeo &g, int item, int &r, int &j) {
    if (g.wshift >= 0) { r = item >> g.wshift; j = item & (g.wpr - 1); }
    else { r = item / g.wpr; j = item - r * g.wpr; }
}

__device__ __forceinline__ void big_load(const uint32_t *__restrict__ base, const BigGeo &g, int item,
                                         unsigned groups, uint32_t P[16]) {
    const uint4 *b4 = reinterpret_cast<const uint4 *>(base);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    const uint4 q0 = b4[item];
    const uint4 q1 = b4[g.items + item];
    const uint4 q2 = (groups & 4u) ? b4[2 * g.items + item] : z;
    const uint4 q3 = (groups & 8u) ? b4[3 * g.items + item] : z;
    P[0] = q0.x; P[4] = q0.y; P[9] = q0.z; P[10] = q0.w;
    P[11] = q1.x; P[3] = q1.y; P[2] = q1.z; P[15] = q1.w;
    P[1] = q2.x; P[5] = q2.y; P[6] = q2.z; P[8] = q2.w;
    P[7] = q3.x; P[12] = q3.y; P[13] = q3.z; P[14] = q3.w;
}

__device__ __forceinline__ void big_store(uint32_t *__restrict__ base, const BigGeo &g, int item,
                                          unsigned groups, const uint32_t P[16]) {
    uint4 *b4 = reinterpret_cast<uint4 *>(base);
    b4[item] = make_uint4(P[0], P[4], P[9], P[10]);
    b4[g.items + item] = make_uint4(P[11], P[3], P[2], P[15]);
    if (groups & 4u) b4[2 * g.items + item] = make_uint4(P[1], P[5], P[6], P[8]);
    if (groups & 8u) b4[3 * g.items + item] = make_uint4(P[7], P[12], P[13], P[14]);
}

// The shared snapshot is item-major: word k of item i at src[i * BIG_SRC + k] (BIG_SRC is odd, so
// consecutive items hit different banks); the nine neighbour items are addressed once per item and
// every source plane is an immediate offset from them.
struct BigNbr {
    const uint32_t *ul, *um, *ur, *ml, *mm, *mr, *dl, *dm, *dr;
};

__device__ __forceinline__ BigNbr big_nbr(const uint32_t *src, const BigGeo &g, int r, int j) {
    const int ru = (r == 0) ? g.h - 1 : r - 1, rd = (r == g.h - 1) ? 0 : r + 1;
    const int jl = (j == 0) ? g.wpr - 1 : j - 1, jr = (j == g.wpr - 1) ? 0 : j + 1;
    const int up = ru * g.wpr, mid = r * g.wpr, dn = rd * g.wpr;
    BigNbr nb;
    nb.ul = src + (up + jl) * BIG_SRC; nb.um = src + (up + j) * BIG_SRC; nb.ur = src + (up + jr) * BIG_SRC;
    nb.ml = src + (mid + jl) * BIG_SRC; nb.mm = src + (mid + j) * BIG_SRC; nb.mr = src + (mid + jr) * BIG_SRC;
    nb.dl = src + (dn + jl) * BIG_SRC; nb.dm = src + (dn + j) * BIG_SRC; nb.dr = src + (dn + jr) * BIG_SRC;
    return nb;
}

// The 3x3 summaries of source plane K around an item.  The three rows are reduced FIRST, separately for
// the left, own and right word (that is word-parallel and needs no alignment); only the reduced words
// are then aligned to the item's columns, so a summary costs 2 or 4 funnel shifts instead of 6.

// OR over the 3x3
template <int K>
__device__ __forceinline__ uint32_t nine_or(const BigNbr &nb) {
    const uint32_t l = or3(nb.ul[K], nb.ml[K], nb.dl[K]);
    const uint32_t m = or3(nb.um[K], nb.mm[K], nb.dm[K]);
    const uint32_t r = or3(nb.ur[K], nb.mr[K], nb.dr[K]);
    return or3(__funnelshift_l(l, m, 1), m, __funnelshift_r(m, r, 1));
}

// per-column 3-row sums (s + 2 c), aligned: left neighbour column, own column, right neighbour column
struct ColSums { uint32_t sl, sm, sr, cl, cm, cr; };

template <int K>
__device__ __forceinline__ ColSums col_sums(const BigNbr &nb) {
    uint32_t u = nb.ul[K], m = nb.ml[K], d = nb.dl[K];
    const uint32_t s_l = xor3(u, m, d), c_l = maj3(u, m, d);
    u = nb.um[K]; m = nb.mm[K]; d = nb.dm[K];
    const uint32_t s_m = xor3(u, m, d), c_m = maj3(u, m, d);
    u = nb.ur[K]; m = nb.mr[K]; d = nb.dr[K];
    const uint32_t s_r = xor3(u, m, d), c_r = maj3(u, m, d);
    ColSums c;
    c.sl = __funnelshift_l(s_l, s_m, 1); c.sm = s_m; c.sr = __funnelshift_r(s_m, s_r, 1);
    c.cl = __funnelshift_l(c_l, c_m, 1); c.cm = c_m; c.cr = __funnelshift_r(c_m, c_r, 1);
    return c;
}

// >= 2 of the nine: some column holds two, or two columns hold an odd number
template <int K, class NB>
__device__ __forceinline__ uint32_t nine_two(const NB &nb) {
    const ColSums c = col_sums<K>(nb);
    return or3(c.cl, c.cm, c.cr) | maj3(c.sl, c.sm, c.sr);
}

// The same summaries with the left / right words taken from the neighbouring LANES instead of shared
// memory: when a row is a power-of-two number of words (<= 32) and the item loop advances by a multiple
// of 32, the words of a row sit in adjacent lanes of one warp, so every lane reduces only its own
// column of three words (3 loads instead of 9, a third of the row-reduction arithmetic) and fetches its
// neighbours' reduced words with two shuffles each (segmented: lane +-1 wraps inside the row's lanes,
// which is the torus).  All 32 lanes must call.
struct BigNbrW {
    const uint32_t *um, *mm, *dm;
    int ll, lr;  // lanes holding the word to the left / right in the same row
};

__device__ __forceinline__ BigNbrW big_nbr_w(const uint32_t *src, const BigGeo &g, int r, int j) {
    const int ru = (r == 0) ? g.h - 1 : r - 1, rd = (r == g.h - 1) ? 0 : r + 1;
    BigNbrW nb;
    nb.um = src + (ru * g.wpr + j) * BIG_SRC;
    nb.mm = src + (r * g.wpr + j) * BIG_SRC;
    nb.dm = src + (rd * g.wpr + j) * BIG_SRC;
    const int lane = threadIdx.x & 31, seg = lane & ~(g.wpr - 1);
    nb.ll = seg | ((lane - 1) & (g.wpr - 1));
    nb.lr = seg | ((lane + 1) & (g.wpr - 1));
    return nb;
}

template <int K>
__device__ __forceinline__ uint32_t nine_or(const BigNbrW &nb) {
    const uint32_t m = or3(nb.um[K], nb.mm[K], nb.dm[K]);
    const uint32_t l = __shfl_sync(0xffffffffu, m, nb.ll), r = __shfl_sync(0xffffffffu, m, nb.lr);
    return or3(__funnelshift_l(l, m, 1), m, __funnelshift_r(m, r, 1));
}

template <int K>
__device__ __forceinline__ ColSums col_sums(const BigNbrW &nb) {
    const uint32_t u = nb.um[K], m = nb.mm[K], d = nb.dm[K];
    const uint32_t s_m = xor3(u, m, d), c_m = maj3(u, m, d);
    const uint32_t s_l = __shfl_sync(0xffffffffu, s_m, nb.ll), s_r = __shfl_sync(0xffffffffu, s_m, nb.lr);
    const uint32_t c_l = __shfl_sync(0xffffffffu, c_m, nb.ll), c_r = __shfl_sync(0xffffffffu, c_m, nb.lr);
    ColSums c;
    c.sl = __funnelshift_l(s_l, s_m, 1); c.sm = s_m; c.sr = __funnelshift_r(s_m, s_r, 1);
    c.cl = __funnelshift_l(c_l, c_m, 1); c.cm = c_m; c.cr = __funnelshift_r(c_m, c_r, 1);
    return c;
}

// n == 3 / n == 4 of the nine (the adder tree of physics_planes)
template <int K, class NB>
__device__ __forceinline__ void nine_count(const NB &nb, uint32_t &is3, uint32_t &is4) {
    const ColSums c = col_sums<K>(nb);
    const uint32_t s0 = xor3(c.sl, c.sm, c.sr), c0 = maj3(c.sl, c.sm, c.sr);
    const uint32_t s1 = xor3(c.cl, c.cm, c.cr), c1 = maj3(c.cl, c.cm, c.cr);
    const uint32_t n1 = c0 ^ s1, k = c0 & s1, n2 = c1 ^ k;  // n = s0 + 2 n1 + 4 n2 (+ 8)
    is3 = s0 & n1 & ~n2;
    is4 = ~s0 & ~n1 & n2;  // n2 set implies the 8s bit (c1 & k) is clear
}

// CTA-wide exclusive prefix sum of one int per thread (thread order); total returned to all.
__device__ __forceinline__ int big_scan(int v, int *warp_tot, int &total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    int before = 0, tot = 0;
    for (int k = 0; k < nw; k++) {
        const int t = warp_tot[k];
        if (k < wid) before += t;
        tot += t;
    }
    __syncthreads();
    total = tot;
    return before + incl - v;
}

struct BigSmem {
    uint32_t *src;    // [BIG_SRC][items]
    uint32_t *elig;   // [items]
    uint32_t *rowf;   // [h]: bit 0 row has preserving|inhibiting, bit 1 row has spawning
    int *warp_tot;    // [BIG_THREADS / 32]
};

template <bool WARP> struct BigNbrSel;
template <> struct BigNbrSel<false> {
    typedef BigNbr type;
    static __device__ __forceinline__ BigNbr make(const uint32_t *src, const BigGeo &g, int r, int j) { return big_nbr(src, g, r, j); }
};
template <> struct BigNbrSel<true> {
    typedef BigNbrW type;
    static __device__ __forceinline__ BigNbrW make(const uint32_t *src, const BigGeo &g, int r, int j) { return big_nbr_w(src, g, r, j); }
};

// Phase B of big_physics: the rule for every item of the board.  WARP: left / right words come from the
// neighbouring lanes (BigNbrW); the lazily evaluated summaries are then decided per warp (votes), not
// per item, because every lane has to take part in the shuffles.
template <bool WARP>
__device__ __forceinline__ void big_rule_items(uint32_t *planes, const BigGeo &g, unsigned groups,
                                               const BigSmem &sm, int any7) {
    const int T = blockDim.x;
    for (int it = threadIdx.x; it < g.items; it += T) {
        int r, j;
        big_row_word(g, it, r, j);
        typename BigNbrSel<WARP>::type nb = BigNbrSel<WARP>::make(sm.src, g, r, j);
        const uint32_t rf = sm.rowf[(r == 0) ? g.h - 1 : r - 1] | sm.rowf[r] | sm.rowf[(r == g.h - 1) ? 0 : r + 1];
        const bool want_f = WARP ? __any_sync(0xffffffffu, (rf & 1u) != 0u) : (rf & 1u) != 0u;
        const bool want_s = WARP ? __any_sync(0xffffffffu, (rf & 2u) != 0u) : (rf & 2u) != 0u;
        uint32_t P[16];
        const unsigned live = groups & (3u | sm.elig[it]);  // phase A left the item's group flags here
        big_load(planes, g, it, live, P);
        const uint32_t a = P[0], fz = P[4];
        uint32_t is3, is4;
        nine_count<0>(nb, is3, is4);
        const uint32_t die0 = a & ~(fz | is3 | is4);
        const uint32_t open0 = ~(a | fz);
        uint32_t PR = 0, IH = 0;
        if (want_f) {
            PR = nine_or<1>(nb);
            IH = nine_or<2>(nb);
        }
        const uint32_t die = die0 & ~PR;
        const uint32_t open = open0 & ~IH;
        const uint32_t birth = open & is3;
        uint32_t elig = 0;
        if (want_s) elig = open & ~is3 & nine_or<3>(nb);
        if (any7) sm.elig[it] = elig;
        const uint32_t flip = die | birth;
        uint32_t nD = 0, nR = 0, nG = 0, nB = 0;
        if (WARP ? __any_sync(0xffffffffu, birth != 0u) : birth != 0u) {
            nD = birth & nine_two<4>(nb);
            uint32_t tR = nine_two<5>(nb);
            uint32_t tG = nine_two<6>(nb);
            uint32_t tB = nine_two<7>(nb);
            if (want_s) {
                tR |= nine_or<8>(nb);
                tG |= nine_or<9>(nb);
                tB |= nine_or<10>(nb);
            }
            nR = birth & tR; nG = birth & tG; nB = birth & tB;
        }
        if (flip) {
            const uint32_t keep = ~flip;
#pragma unroll
            for (int b = 0; b < 16; b++) P[b] &= keep;
            P[0] |= birth; P[3] |= nD; P[9] |= nR; P[10] |= nG; P[11] |= nB;
            big_store(planes, g, it, live, P);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// One advance_board step of the CTA's board, in place on its resident planes (advance_board.c:34-120,
// closed form of SURVEY Appendix A; same lazy-but-exact evaluation as physics_planes, per item).
// Every thread of the CTA must call.  Returns the env's updated draw counter (same in all threads).
// ---------------------------------------------------------------------------------------------
__device__ inline uint64_t big_physics(uint32_t *planes, const BigGeo &g, unsigned groups,
                                       const BigSmem &sm, float spawn_prob, uint64_t seed, uint64_t env,
                                       uint64_t counter) {
    const int T = blockDim.x;
    // ---- phase A: snapshot of the source planes ----
    for (int r = threadIdx.x; r < g.h; r += T) sm.rowf[r] = 0u;
    __syncthreads();
    int has7 = 0;
    for (int it = threadIdx.x; it < g.items; it += T) {
        uint32_t P[16];
        big_load(planes, g, it, groups, P);
        const uint32_t a = P[0], sp = P[7];
        uint32_t *s = sm.src + it * BIG_SRC;
        s[0] = a;
        s[1] = P[5];
        s[2] = P[6];
        s[3] = sp;
        s[4] = a & (P[3] | P[8]);  // D2 = destructible | exit (advance_board.c:44-46)
        s[5] = a & P[9];
        s[6] = a & P[10];
        s[7] = a & P[11];
        s[8] = sp & P[9];
        s[9] = sp & P[10];
        s[10] = sp & P[11];
        // which of the two rarely populated plane groups hold anything in this item: phase B neither
        // loads nor stores a group that is all zero here (no rule can create a bit in it)
        sm.elig[it] = ((P[1] | P[5] | P[6] | P[8]) ? 4u : 0u) | ((sp | P[12] | P[13] | P[14]) ? 8u : 0u);
        const uint32_t f = ((P[5] | P[6]) ? 1u : 0u) | (sp ? 2u : 0u);
        if (f) {
            int r, j;
            big_row_word(g, it, r, j);
            atomicOr(&sm.rowf[r], f);
        }
        has7 |= sp != 0u;
    }
    const int any7 = __syncthreads_or(has7);  // also the barrier between the phases
    // ---- phase B: rule per item ----
    // rows that are a power-of-two number of words, whole warps of items: neighbours through the warp
    if (g.wshift >= 0 && g.wpr <= 32 && (g.items & 31) == 0 && (T & 31) == 0)
        big_rule_items<true>(planes, g, groups, sm, any7);
    else
        big_rule_items<false>(planes, g, groups, sm, any7);
    if (!any7) return counter;  // uniform
    // ---- spawn draws ----
    // (1) raster-order rank of every eligible cell: CTA-wide scan over the items' popcounts; the ranks
    //     are parked in source slot 0 of each item (the count / flag sources are dead after phase B);
    // (2) the draws counter .. counter + total - 1 are evaluated by ALL threads, one draw per lane, a
    //     warp vote packing 32 outcomes into one word of a bit stream (source slot 1 of item w);
    // (3) every item with eligible cells pulls its popcount bits out of the stream and applies them.
    __syncthreads();
    int total = 0;
    for (int base = 0; base < g.items; base += T) {  // uniform trip count
        const int it = base + threadIdx.x;
        const int cnt = (it < g.items) ? __popc(sm.elig[it]) : 0;
        int chunk_total;
        const int rank = big_scan(cnt, sm.warp_tot, chunk_total);
        if (it < g.items) sm.src[it * BIG_SRC] = (uint32_t)(total + rank);
        total += chunk_total;
    }
    if (total == 0) return counter;  // uniform
    {
        const uint64_t thr = spawn_threshold(spawn_prob);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = T >> 5;
        for (int w0 = warp; w0 * 32 < total; w0 += nwarps) {
            const int d = w0 * 32 + lane;
            const bool hit = d < total && stream_u53(seed, env, counter + (uint64_t)d) < thr;
            const uint32_t word = __ballot_sync(FULL, hit);
            if (lane == 0) sm.src[w0 * BIG_SRC + 1] = word;
        }
    }
    __syncthreads();
    for (int it = threadIdx.x; it < g.items; it += T) {
        uint32_t e = sm.elig[it];
        if (!e) continue;
        const uint32_t rank = sm.src[it * BIG_SRC];
        const int w0 = (int)(rank >> 5), nw = (total + 31) >> 5;
        const uint32_t lo = sm.src[w0 * BIG_SRC + 1];
        const uint32_t hi = (w0 + 1 < nw) ? sm.src[(w0 + 1) * BIG_SRC + 1] : 0u;
        uint32_t bits = __funnelshift_r(lo, hi, rank & 31u);  // this item's outcomes, first eligible cell in bit 0
        uint32_t spawn = 0;
        while (e) {
            const uint32_t bit = e & (0u - e);
            e ^= bit;
            if (bits & 1u) spawn |= bit;
            bits >>= 1;
        }
        if (!spawn) continue;
        int r, j;
        big_row_word(g, it, r, j);
        const BigNbr nb = big_nbr(sm.src, g, r, j);
        const uint32_t tR = nine_two<5>(nb) | nine_or<8>(nb);
        const uint32_t tG = nine_two<6>(nb) | nine_or<9>(nb);
        const uint32_t tB = nine_two<7>(nb) | nine_or<10>(nb);
        uint32_t P[16];
        big_load(planes, g, it, groups, P);
        const uint32_t keep = ~spawn;  // spawned cells were dead: their residue bits are discarded
#pragma unroll
        for (int b = 0; b < 16; b++) P[b] &= keep;
        P[0] |= spawn; P[3] |= spawn;  // always destructible (advance_board.c:110-114)
        P[9] |= spawn & tR; P[10] |= spawn & tG; P[11] |= spawn & tB;
        big_store(planes, g, it, groups, P);
    }
    __syncthreads();
    return counter + (uint64_t)total;
}

// ---------------------------------------------------------------------------------------------
// Single cells of a resident big board, for execute_action: lane p < 16 owns plane p, a warp vote
// assembles / scatters the uint16 cell value.  Called by all 32 lanes of ONE warp.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int plane_word(int p, int item, int items) {
    const int idx = (int)((0x7FED432BCA915680ull >> (4 * p)) & 15ull);  // plane -> group * 4 + slot
    return ((idx >> 2) * items + item) * 4 + (idx & 3);
}

__device__ __forceinline__ uint32_t big_cell_read(const uint32_t *planes, const BigGeo &g, int x, int y) {
    const int lane = threadIdx.x & 31;
    uint32_t bit = 0;
    if (lane < 16) bit = (planes[plane_word(lane, y * g.wpr + (x >> 5), g.items)] >> (x & 31)) & 1u;
    return __ballot_sync(FULL, bit != 0u) & 0xffffu;
}

__device__ __forceinline__ void big_cell_write(uint32_t *planes, const BigGeo &g, unsigned groups, int x,
                                               int y, uint32_t v) {
    const int lane = threadIdx.x & 31;
    if (lane < 16) {
        const int idx = (int)((0x7FED432BCA915680ull >> (4 * lane)) & 15ull);
        if ((groups >> (idx >> 2)) & 1u) {
            uint32_t *w = planes + plane_word(lane, y * g.wpr + (x >> 5), g.items);
            const uint32_t m = 1u << (x & 31);
            *w = (*w & ~m) | (((v >> lane) & 1u) ? m : 0u);
        }
    }
    __syncwarp();
}

// execute_action (safelife_game.py:347-393, move_agent :308-345), literal read / write order like
// execute_action_cells, on single cells of the resident planes.  One warp.
__device__ inline int big_action(uint32_t *b, const BigGeo &g, unsigned groups, int w, int &ax, int &ay,
                                 int &orientation, int &game_over, int action, int can_exit) {
    if (game_over || action < 1 || action > 8) return 0;
    const int h = g.h;
    const int dir = (action - 1) & 3;
    orientation = dir;
    const int fx = (dir & 1) ? 2 - dir : 0, fy = (dir & 1) ? 0 : dir - 1;
    const int x0 = ax, y0 = ay;
    const int x1 = (x0 + fx + w) % w, y1 = (y0 + fy + h) % h;
    int reward = 0;
    if (action <= 4) {
        const int x2 = (x0 - fx + w) % w, y2 = (y0 - fy + h) % h;
        bool moved = false;
        const uint32_t front = big_cell_read(b, g, x1, y1);
        if (front == 0) {
            big_cell_write(b, g, groups, x1, y1, big_cell_read(b, g, x0, y0));
            big_cell_write(b, g, groups, x0, y0, 0);
            moved = true;
        } else if ((front & EXIT) && can_exit) {
            game_over = 1;
            reward = 1;
        } else if (front & PUSHABLE) {
            const int x3 = (x0 + 2 * fx + 2 * w) % w, y3 = (y0 + 2 * fy + 2 * h) % h;
            const uint32_t far = big_cell_read(b, g, x3, y3);
            if (far == 0) {
                big_cell_write(b, g, groups, x3, y3, front);
                big_cell_write(b, g, groups, x1, y1, big_cell_read(b, g, x0, y0));
                big_cell_write(b, g, groups, x0, y0, 0);
                moved = true;
            } else if (far & EXIT) {
                big_cell_write(b, g, groups, x1, y1, big_cell_read(b, g, x0, y0));
                big_cell_write(b, g, groups, x0, y0, 0);
                moved = true;
            }
        }
        if (moved) {
            ax = x1;
            ay = y1;
            const uint32_t behind = big_cell_read(b, g, x2, y2);
            if (behind & PULLABLE) {
                big_cell_write(b, g, groups, x0, y0, behind);
                big_cell_write(b, g, groups, x2, y2, 0);
            }
        }
    } else {
        const uint32_t pc = big_cell_read(b, g, x0, y0) & COLORS;
        const uint32_t t = big_cell_read(b, g, x1, y1);
        if (t == 0)
            big_cell_write(b, g, groups, x1, y1, ALIVE | DESTRUCTIBLE | pc);
        else if (t & DESTRUCTIBLE)
            big_cell_write(b, g, groups, x1, y1, 0);
    }
    return reward;
}

// ---------------------------------------------------------------------------------------------
// score / paint / reset passes over the CTA's env.  Every thread of the CTA must call.
// ---------------------------------------------------------------------------------------------
// block reduction of three ints; scratch: 96 ints
__device__ __forceinline__ void big_sum3(int &a, int &b, int &c, int *scratch) {
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) { scratch[wid] = a; scratch[32 + wid] = b; scratch[64 + wid] = c; }
    __syncthreads();
    int ra = 0, rb = 0, rc = 0;
    for (int k = 0; k < nw; k++) { ra += scratch[k]; rb += scratch[32 + k]; rc += scratch[64 + k]; }
    a = ra; b = rb; c = rc;
    __syncthreads();
}

// current_points / performance_ratio sums (safelife_game.py:590-631); skip_exits: cells of the exit
// mask e0 count as dead (what they are once painted)
__device__ inline void big_score(const uint32_t *board, const uint32_t *goals, const BigGeo &g,
                                 const uint32_t *e0, bool skip_exits, int *scratch, int &pts, int &cur,
                                 int &pos) {
    Geo full;
    full.h = g.h; full.w = 32; full.up = full.dn = 0; full.wmask = 0xffffffffu; full.hib = 0x80000000u; full.row = true;
    pts = cur = pos = 0;
    for (int it = threadIdx.x; it < g.items; it += blockDim.x) {
        uint32_t B[16], G[16];
        big_load(board, g, it, 3u, B);  // every plane the score reads lives in groups 0 / 1
        big_load(goals, g, it, 3u, G);
        if (skip_exits) B[0] &= ~e0[it];
        const RowScore rs = score_planes(B, G, full);
        pts += rs.points; cur += rs.current; pos += rs.possible;
    }
    big_sum3(pts, cur, pos, scratch);
}

// update_exit_colors' repaint (safelife_game.py:531-537)
__device__ inline void big_paint(uint32_t *board, const BigGeo &g, unsigned groups, const uint32_t *e0,
                                 int can_exit) {
    for (int it = threadIdx.x; it < g.items; it += blockDim.x) {
        const uint32_t m = e0[it];
        if (!m) continue;
        uint32_t P[16];
        big_load(board, g, it, groups, P);
#pragma unroll
        for (int b = 0; b < 16; b++) P[b] &= ~m;
        P[4] |= m; P[8] |= m;
        if (can_exit) P[9] |= m;
        big_store(board, g, it, groups, P);
    }
}

// ---------------------------------------------------------------------------------------------
// The env kernel for big boards: one CTA per env; `ops` selects the stages (the fused step is
// ACTION | PHYSICS | BOOK; the single stages serve SafeLifeGame's method-by-method API).
// ---------------------------------------------------------------------------------------------
enum : int { BIG_RESET = 1, BIG_ACTION = 2, BIG_PHYSICS = 4, BIG_BOOK = 8, BIG_POINTS = 16, BIG_PAINT = 32 };

struct BigArgs {
    uint32_t *board, *goals;
    slb_env_state *state;
    const uint8_t *actions;
    const int32_t *level_ids;
    const uint8_t *mask;
    int32_t *reward_out;
    uint8_t *done_out;
    int32_t *points, *completed, *possible;
    uint8_t *can_exit;
    uint64_t seed, env_id0;
    int n, time_limit, auto_reset_stride, ops;
};

// load level `level` into the env (safelife_env.py:188-198); fills the warp-0-owned scalars in *st
__device__ inline void big_reset(const slb_pool &pool, int level, const BigGeo &g, unsigned gb, unsigned gg,
                                 uint32_t *board, uint32_t *goals, slb_env_state *st, int *scratch) {
    const size_t words = (size_t)16 * g.items;
    const uint4 *b0 = reinterpret_cast<const uint4 *>((const uint32_t *)pool.board_res + level * words);
    const uint4 *g0 = reinterpret_cast<const uint4 *>((const uint32_t *)pool.goals_res + level * words);
    uint4 *b4 = reinterpret_cast<uint4 *>(board), *g4 = reinterpret_cast<uint4 *>(goals);
    for (int k = threadIdx.x; k < 4 * g.items; k += blockDim.x) {
        const unsigned grp = 1u << (k / g.items);
        if (gb & grp) b4[k] = b0[k];
        if (gg & grp) g4[k] = g0[k];
    }
    __syncthreads();
    const uint32_t *e0 = pool.exit_mask + (size_t)level * g.items;
    int pts, cur, pos;
    big_score(board, goals, g, e0, true, scratch, pts, cur, pos);  // points as they are once painted
    const int baseline = pool.baseline[level];
    const int completed = cur - baseline, possible = pos - baseline;
    const int ce = can_exit_rule(pool.min_performance[level], completed, possible);
    big_paint(board, g, gb, e0, ce);
    if (threadIdx.x == 0) {
        st->agent_x = pool.agent_x[level]; st->agent_y = pool.agent_y[level];
        st->orientation = pool.orientation[level];
        st->game_over = 0; st->num_steps = 0; st->level_id = level; st->episode_length = 0;
        st->episode_completed = 0; st->old_points = pts; st->can_exit = ce; st->completed = completed;
        st->possible = possible; st->episode_reward = 0;
        st->min_performance = pool.min_performance[level];  // deserialize (safelife_game.py:205)
        st->flags = 0;
        // rng_counter deliberately untouched: the stream keeps running across episodes
    }
    __syncthreads();
}

// (big_env_kernel itself lives in slb_big_rows.cuh, next to the sliding-window physics it can use)

// SimpleSideEffectPenalty's count of side-effect cells per env (env_wrappers.py:319-346); grid = N
__global__ void __launch_bounds__(256)
big_penalty_kernel(slb_pool pool, const uint32_t *__restrict__ board, const uint32_t *__restrict__ goals,
                   const slb_env_state *__restrict__ state, int32_t *count) {
    __shared__ int scratch[96];
    const BigGeo g = big_geo(pool.h, pool.w);
    const int i = blockIdx.x, level = state[i].level_id;
    const size_t words = (size_t)16 * g.items;
    const uint32_t *e0 = pool.exit_mask + (size_t)level * g.items;
    int c = 0, d1 = 0, d2 = 0;
    for (int it = threadIdx.x; it < g.items; it += blockDim.x) {
        uint32_t B[16], S[16], G[16];
        big_load(board + i * words, g, it, pool.plane_groups[0], B);
        big_load((const uint32_t *)pool.board_res + level * words, g, it, 15u, S);
        big_load(goals + i * words, g, it, 3u, G);
        c += __popc(side_effect_word(B, S, G, e0[it], 0xffffffffu));
    }
    big_sum3(c, d1, d2, scratch);
    if (threadIdx.x == 0) count[i] = c;
}

// ---------------------------------------------------------------------------------------------
// cells <-> resident planes for big boards (API boundary); grid = N, one item per thread per pass.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
big_import_kernel(const uint16_t *__restrict__ cells, uint32_t *planes, int h, int w, uint32_t *plane_bits) {
    const BigGeo g = big_geo(h, w);
    const uint4 *src = reinterpret_cast<const uint4 *>(cells + (size_t)blockIdx.x * h * w);  // 64 B per item
    uint32_t *dst = planes + (size_t)blockIdx.x * 16 * g.items;
    uint32_t bits = 0;
    for (int it = threadIdx.x; it < g.items; it += blockDim.x) {
        uint32_t x[16], P[16];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 v = src[it * 4 + q];
            x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
        }
        cells_to_planes(x, P);
        big_store(dst, g, it, 15u, P);
#pragma unroll
        for (int b = 0; b < 16; b++) bits |= (P[b] != 0u) ? (1u << b) : 0u;
    }
    if (plane_bits) {
        bits = __reduce_or_sync(FULL, bits);
        if ((threadIdx.x & 31) == 0 && (bits & ~*plane_bits)) atomicOr(plane_bits, bits);
    }
}

__global__ void __launch_bounds__(256)
big_export_kernel(const uint32_t *__restrict__ planes, uint16_t *cells, int h, int w) {
    const BigGeo g = big_geo(h, w);
    const uint32_t *src = planes + (size_t)blockIdx.x * 16 * g.items;
    uint4 *dst = reinterpret_cast<uint4 *>(cells + (size_t)blockIdx.x * h * w);
    for (int it = threadIdx.x; it < g.items; it += blockDim.x) {
        uint32_t x[16], P[16];
        big_load(src, g, it, 15u, P);
        planes_to_cells(P, x);
#pragma unroll
        for (int q = 0; q < 4; q++) dst[it * 4 + q] = make_uint4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
    }
}

// get_obs from resident big boards (safelife_env.py:125-155, helper_utils.py:41-74): one thread per
// view cell gathers its 16 plane bits.  grid = N.
__device__ __forceinline__ uint32_t big_cell_value(const uint32_t *planes, const BigGeo &g, int x, int y) {
    const int item = y * g.wpr + (x >> 5), sh = x & 31;
    uint32_t v = 0;
#pragma unroll
    for (int p = 0; p < 16; p++) v |= ((planes[plane_word(p, item, g.items)] >> sh) & 1u) << p;
    return v;
}

__global__ void __launch_bounds__(256)
big_obs_kernel(slb_pool pool, const uint32_t *__restrict__ board, const uint32_t *__restrict__ goals,
               const slb_env_state *__restrict__ state, int vh, int vw, int remove_white, uint16_t *obs) {
    const BigGeo g = big_geo(pool.h, pool.w);
    const int i = blockIdx.x, h = pool.h, w = pool.w;
    const uint32_t *b = board + (size_t)i * 16 * g.items, *gl = goals + (size_t)i * 16 * g.items;
    const int x0 = state[i].agent_x, y0 = state[i].agent_y, level = state[i].level_id;
    const int x1 = x0 - vw / 2, y1 = y0 - vh / 2;
    uint16_t *o = obs + (size_t)i * vh * vw;
    for (int k = threadIdx.x; k < vh * vw; k += blockDim.x) {
        const int vy = k / vw, vx = k - vy * vw;
        const int y = pymod_w(y1 + vy, h), x = pymod_w(x1 + vx, w);
        uint32_t gv = big_cell_value(gl, g, x, y) & COLORS;
        if (remove_white && gv == COLORS) gv = 0;
        o[k] = (uint16_t)(big_cell_value(b, g, x, y) + (gv << 3));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t *em = pool.exit_mask + (size_t)level * g.items;
        for (int it = 0; it < g.items; it++) {
            uint32_t m = em[it];
            while (m) {
                const int y = it / g.wpr, x = (it - y * g.wpr) * 32 + __ffs(m) - 1;  // (rare) exit cells only
                m &= m - 1;
                int jy = pymod_w(y - y0 + h / 2, h) - h / 2;
                int jx = pymod_w(x - x0 + w / 2, w) - w / 2;
                jy = min(max(jy + vh / 2, 0), vh - 1);
                jx = min(max(jx + vw / 2, 0), vw - 1);
                uint32_t gv = big_cell_value(gl, g, x, y) & COLORS;
                if (remove_white && gv == COLORS) gv = 0;
                o[jy * vw + jx] = (uint16_t)(big_cell_value(b, g, x, y) + (gv << 3));
            }
        }
    }
}

}  // namespace slb
