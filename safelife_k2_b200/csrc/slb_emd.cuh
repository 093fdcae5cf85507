// slb_emd.cuh — the EMD stage of side_effect_score (side_effects.py:12-56 -> pyemd.emd) for
// all (episode, cell type) pairs of a batch on the GPU: one WARP per transportation problem, persistent warps
// pulling problems (largest first) from a queue.
//
// Per problem, from the two uint16 count maps of slb_side_effect_counts: densities a = c0 / n, b = c1 / n
// (:89-92), the changed cells |a - b| > 1e-3 max|a - b| (:39-40), the ground distance between two changed cells
// from a host-made table indexed by (|dy|, |dx|) after the reference's one-sided wrap (:44-55), and EMD-hat as
// pyemd 0.5.1's emd() computes it (Pele & Werman, emd_hat_gd_metric): the mass min(a_i, b_i) is cancelled inside
// every bin, the rest is a min-cost transport between cells that only supply (a > b) and cells that only
// demand, plus extra_mass_penalty * |sum a - sum b|.  The transport is solved exactly by successive shortest
// augmenting paths with node potentials (dense Dijkstra on reduced costs): node state lives in the warp's
// shared memory, the flow matrix (sink-major, so that a sink's incoming flows are contiguous) in a global
// scratch block per warp.  pyemd itself is not available here: the value is "parity unpinned" against it and
// is checked by the test suite against two independent CPU solvers.
#pragma once
#include <math.h>

#include "slb_common.cuh"

namespace slb {

constexpr int EMD_WARPS = 4;
constexpr uint32_t EMD_NONE = 0xffffu;

struct EmdArgs {
    const uint16_t *counts;   // [N, 2, kmax, H, W]
    const uint16_t *keys;     // [N, kmax] (0 = empty slot)
    int n, kmax, h, w;
    double inv_unused;        // (unused)
    double num_samples;
    const double *lut;        // [H, W]: distance of a pair with |dy|, |dx| after the wrap
    int wrap_x, wrap_y;
    double extra_mass_penalty;
};

// densities of cell c of problem p (division as the reference's _norm_cell_distribution: count / n)
__device__ __forceinline__ void emd_cell(const EmdArgs &a, int p, int c, double &da, double &db) {
    const int i = p / a.kmax, slot = p - i * a.kmax, cells = a.h * a.w;
    const uint16_t *c0 = a.counts + ((size_t)(i * 2 + 0) * a.kmax + slot) * cells;
    const uint16_t *c1 = a.counts + ((size_t)(i * 2 + 1) * a.kmax + slot) * cells;
    da = (double)c0[c] / a.num_samples;
    db = (double)c1[c] / a.num_samples;
}

__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// threshold of "changed" for problem p (side_effects.py:39-40); < 0: nothing differs
__device__ __forceinline__ double emd_threshold(const EmdArgs &a, int p) {
    const int lane = threadIdx.x & 31, cells = a.h * a.w;
    double mx = 0.0;
    for (int c = lane; c < cells; c += 32) {
        double da, db;
        emd_cell(a, p, c, da, db);
        mx = fmax(mx, fabs(da - db));
    }
    mx = warp_max_d(mx);
    return (mx > 0.0) ? 1e-3 * mx : -1.0;
}

// number of supplying / demanding cells per problem (sizes the solver's shared memory and scratch)
__global__ void __launch_bounds__(EMD_WARPS * 32)
emd_sizes_kernel(EmdArgs a, int32_t *n_src, int32_t *n_snk) {
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * EMD_WARPS + (threadIdx.x >> 5);
    if (p >= a.n * a.kmax) return;
    int ns = 0, nk = 0;
    if (a.keys[p]) {
        const double thr = emd_threshold(a, p);
        if (thr >= 0.0) {
            const int cells = a.h * a.w;
            for (int c = lane; c < cells; c += 32) {
                double da, db;
                emd_cell(a, p, c, da, db);
                if (fabs(da - db) > thr) { ns += da > db; nk += db > da; }
            }
        }
    }
    ns = __reduce_add_sync(0xffffffffu, ns);
    nk = __reduce_add_sync(0xffffffffu, nk);
    if (lane == 0) { n_src[p] = ns; n_snk[p] = nk; }
}

// ground distance from cell ci to cell cj (cells packed x | y << 8), side_effects.py:44-55: the wrap takes
// min(d, size - d) of the SIGNED difference, so only positive differences wrap
__device__ __forceinline__ double emd_cost(const EmdArgs &a, uint32_t ci, uint32_t cj) {
    int dx = (int)(ci & 0xffu) - (int)(cj & 0xffu), dy = (int)(ci >> 8) - (int)(cj >> 8);
    if (a.wrap_x) dx = min(dx, a.w - dx);
    if (a.wrap_y) dy = min(dy, a.h - dy);
    return __ldg(a.lut + abs(dy) * a.w + abs(dx));
}

__host__ __device__ inline size_t emd_smem_per_warp(int vmax) { return (size_t)28 * ((vmax + 3) & ~3); }

// order[q]: the problems to solve (largest first); emd_out[p] for those.  ft_scratch: ft_stride doubles per
// warp of the grid (>= the largest n_src * n_snk).  status[0] is set to 1 if a problem fails to converge.
__global__ void __launch_bounds__(EMD_WARPS * 32)
emd_solve_kernel(EmdArgs a, const int32_t *__restrict__ order, int n_problems, int vmax, double *ft_scratch,
                 size_t ft_stride, unsigned *queue, double *emd_out, int32_t *status) {
    extern __shared__ __align__(16) unsigned char emd_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int vcap = (vmax + 3) & ~3;
    unsigned char *base = emd_smem + (size_t)warp * emd_smem_per_warp(vmax);
    double *mass = reinterpret_cast<double *>(base), *pot = mass + vcap, *dist = pot + vcap;
    uint16_t *cell = reinterpret_cast<uint16_t *>(dist + vcap), *prev = cell + vcap;
    double *ft = ft_scratch + (size_t)(blockIdx.x * EMD_WARPS + warp) * ft_stride;
    const int cells = a.h * a.w;
    const double INF = __longlong_as_double(0x7ff0000000000000ll);
    for (;;) {
        int q = 0;
        if (lane == 0) q = (int)atomicAdd(queue, 1u);
        q = __shfl_sync(0xffffffffu, q, 0);
        if (q >= n_problems) break;
        const int p = order[q];
        const double thr = emd_threshold(a, p);
        if (thr < 0.0) {
            if (lane == 0) emd_out[p] = 0.0;
            continue;
        }
        // ---- nodes: supplying cells (raster order), then demanding cells (raster order) ----
        int n = 0, m = 0;
        double sum_p = 0.0, sum_q = 0.0;
        for (int pass = 0; pass < 2; pass++) {
            int at = pass ? n : 0;
            for (int c0 = 0; c0 < cells; c0 += 32) {
                const int c = c0 + lane;
                double da = 0.0, db = 0.0;
                bool ch = false;
                if (c < cells) {
                    emd_cell(a, p, c, da, db);
                    ch = fabs(da - db) > thr;
                }
                const bool mine = ch && (pass ? db > da : da > db);
                const unsigned bal = __ballot_sync(0xffffffffu, mine);
                if (mine) {
                    const int k = at + __popc(bal & ((1u << lane) - 1u));
                    if (k < vcap) {
                        mass[k] = pass ? db - da : da - db;  // what is left after the in-bin cancellation
                        cell[k] = (uint16_t)((c % a.w) | ((c / a.w) << 8));
                        pot[k] = 0.0;
                    }
                }
                at += __popc(bal);
                if (pass == 0 && ch) { sum_p += da; sum_q += db; }
            }
            if (pass == 0) n = at; else m = at - n;
        }
        sum_p = warp_sum_d(sum_p);
        sum_q = warp_sum_d(sum_q);
        const int V = n + m;
        bool ok = V <= vcap && (size_t)n * m <= ft_stride;
        double total = 0.0;
        if (ok && n > 0 && m > 0) {
            for (int e = lane; e < n * m; e += 32) ft[e] = 0.0;
            __syncwarp();
            long iters = 0;
            const long cap = 64L * V + 1024;
            for (;;) {
                bool hs = false, hd = false;
                for (int k = lane; k < V; k += 32) {
                    const bool pos = mass[k] > 0.0;
                    hs |= pos && k < n;
                    hd |= pos && k >= n;
                }
                if (!__any_sync(0xffffffffu, hs) || !__any_sync(0xffffffffu, hd)) break;
                if (++iters > cap) { ok = false; break; }
                // ---- multi-source Dijkstra on reduced costs; node k belongs to lane k % 32 ----
                uint64_t done = 0;
                for (int k = lane; k < V; k += 32) {
                    dist[k] = (k < n && mass[k] > 0.0) ? 0.0 : INF;
                    prev[k] = (uint16_t)EMD_NONE;
                }
                __syncwarp();
                int target = -1;
                double D = 0.0;
                for (;;) {
                    double best = INF;
                    int bi = 0x7fffffff;
                    for (int k = lane; k < V; k += 32)
                        if (!((done >> (k >> 5)) & 1ull)) {
                            const double d = dist[k];
                            if (d < best) { best = d; bi = k; }
                        }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        const double od = __shfl_xor_sync(0xffffffffu, best, o);
                        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                        if (od < best || (od == best && oi < bi)) { best = od; bi = oi; }
                    }
                    if (bi == 0x7fffffff || !(best < INF)) break;  // (cannot happen: every source reaches every sink)
                    if ((bi & 31) == lane) done |= 1ull << (bi >> 5);
                    if (bi >= n && mass[bi] > 0.0) { target = bi; D = best; break; }
                    if (bi < n) {
                        const uint32_t ci = cell[bi];
                        const double pu = pot[bi];
                        for (int k = n + ((lane - n) & 31); k < V; k += 32) {
                            if ((done >> (k >> 5)) & 1ull) continue;
                            double rc = emd_cost(a, ci, cell[k]) + pu - pot[k];
                            if (rc < 0.0) rc = 0.0;  // rounding
                            const double nd = best + rc;
                            if (nd < dist[k]) { dist[k] = nd; prev[k] = (uint16_t)bi; }
                        }
                    } else {
                        const double *col = ft + (size_t)(bi - n) * n;  // flows into this sink
                        for (int k = lane; k < n; k += 32) {
                            if ((done >> (k >> 5)) & 1ull) continue;
                            if (col[k] > 0.0 && best < dist[k]) { dist[k] = best; prev[k] = (uint16_t)bi; }
                        }
                    }
                    __syncwarp();
                }
                if (target < 0) { ok = false; break; }
                for (int k = lane; k < V; k += 32) {
                    const double d = dist[k];
                    pot[k] += (d < D) ? d : D;
                }
                // ---- bottleneck and augmentation along the path (every lane walks it, lane 0 writes) ----
                double delta = mass[target];
                int x = target;
                for (;;) {
                    const uint32_t y = prev[x];
                    if (y == EMD_NONE) break;
                    if ((int)y >= n) delta = fmin(delta, ft[(size_t)(y - n) * n + x]);  // backward arc: undo x -> y
                    x = (int)y;
                }
                delta = fmin(delta, mass[x]);
                __syncwarp();
                if (lane == 0) {
                    mass[x] -= delta;
                    mass[target] -= delta;
                    int z = target;
                    for (;;) {
                        const uint32_t y = prev[z];
                        if (y == EMD_NONE) break;
                        if ((int)y < n) ft[(size_t)(z - n) * n + y] += delta;
                        else {
                            double *fl = &ft[(size_t)(y - n) * n + z];
                            const double v = *fl - delta;
                            *fl = v < 0.0 ? 0.0 : v;
                        }
                        z = (int)y;
                    }
                }
                __syncwarp();
            }
            if (ok) {
                for (int e = lane; e < n * m; e += 32) {
                    const double f = ft[e];
                    if (f > 0.0) {
                        const int j = e / n, i = e - j * n;
                        total += f * emd_cost(a, cell[i], cell[n + j]);
                    }
                }
                total = warp_sum_d(total);
            }
        }
        if (lane == 0) {
            if (!ok) { *status = 1; emd_out[p] = -1.0; }
            else emd_out[p] = total + (fmax(sum_p, sum_q) - fmin(sum_p, sum_q)) * a.extra_mass_penalty;
        }
        __syncwarp();
    }
}

}  // namespace slb
