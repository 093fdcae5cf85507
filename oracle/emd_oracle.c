/*
 * emd_oracle.c — TEST INFRASTRUCTURE (oracle): an exact solver of the transportation problem behind
 * side_effects.earth_mover_distance (/root/reference/safelife/side_effects.py:12-56), independent of the
 * scipy LP and of the CUDA kernel it checks.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may use anything under oracle/.
 *
 * The reference hands the changed cells to third-party pyemd 0.5.1 (requirements.txt:1; call site
 * side_effects.py:56: pyemd.emd(a[changed], b[changed], dist, extra_mass_penalty)), which is neither
 * vendored under /root/reference nor installed here: PARITY UNPINNED against pyemd itself.  What is
 * restated here is the published algorithm pyemd wraps — Pele & Werman's EMD-hat ("Fast and robust earth
 * mover's distances", ICCV 2009), `emd_hat_gd_metric`:
 *
 *   EMD-hat(P, Q, C) = min_F sum_ij F_ij C_ij + |sum P - sum Q| * extra_mass_penalty
 *   subject to F >= 0, sum_j F_ij <= P_i, sum_i F_ij <= Q_j, sum_ij F_ij = min(sum P, sum Q),
 *
 * where the `_gd_metric` entry point first cancels min(P_i, Q_i) inside every bin (a zero-cost flow when
 * C_ii = 0; optimal when C is a metric) and solves the problem that is left: bins that only supply against
 * bins that only demand.  slo_emd_transport solves that remaining problem exactly by successive shortest
 * augmenting paths with node potentials (Dijkstra on reduced costs, dense); the reduction, the choice of the
 * changed cells and the ground distance are in oracle/emd.py.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* min sum f_ij c_ij, f >= 0, row sums <= supply, column sums <= demand, total = min(sum supply, sum demand).
 * cost: n x m row-major, >= 0.  flow (may be NULL): n x m out.  Returns the optimal cost, or -1 on a
 * failure to converge (iteration cap). */
double slo_emd_transport(int n, int m, const double *supply, const double *demand, const double *cost,
                         double *flow) {
    if (n <= 0 || m <= 0) return 0.0;
    const int V = n + m;
    double *f = (double *)calloc((size_t)n * m, sizeof(double));
    double *rs = (double *)malloc(sizeof(double) * n), *rd = (double *)malloc(sizeof(double) * m);
    double *pot = (double *)calloc(V, sizeof(double)), *dist = (double *)malloc(sizeof(double) * V);
    int *prev = (int *)malloc(sizeof(int) * V);
    char *done = (char *)malloc(V);
    memcpy(rs, supply, sizeof(double) * n);
    memcpy(rd, demand, sizeof(double) * m);
    long iters = 0;
    const long cap = 64L * (long)(n + m) + 1024;
    int ok = 1;
    for (;;) {
        int have_src = 0, have_dst = 0;
        for (int i = 0; i < n; i++) have_src |= rs[i] > 0.0;
        for (int j = 0; j < m; j++) have_dst |= rd[j] > 0.0;
        if (!have_src || !have_dst) break;
        if (++iters > cap) { ok = 0; break; }
        /* multi-source Dijkstra over the residual graph; nodes 0..n-1 sources, n..n+m-1 sinks.
         * arc i -> j: reduced cost c_ij + pot[i] - pot[n+j] (>= 0); arc j -> i where f_ij > 0: reduced cost 0 */
        for (int x = 0; x < V; x++) { dist[x] = INFINITY; prev[x] = -1; done[x] = 0; }
        for (int i = 0; i < n; i++) if (rs[i] > 0.0) dist[i] = 0.0;
        int target = -1;
        for (;;) {
            int x = -1;
            double best = INFINITY;
            for (int y = 0; y < V; y++) if (!done[y] && dist[y] < best) { best = dist[y]; x = y; }
            if (x < 0) break;
            done[x] = 1;
            if (x >= n && rd[x - n] > 0.0) { target = x; break; }
            if (x < n) {
                const double *c = cost + (size_t)x * m;
                for (int j = 0; j < m; j++) {
                    if (done[n + j]) continue;
                    double rc = c[j] + pot[x] - pot[n + j];
                    if (rc < 0.0) rc = 0.0; /* rounding */
                    if (best + rc < dist[n + j]) { dist[n + j] = best + rc; prev[n + j] = x; }
                }
            } else {
                const int j = x - n;
                for (int i = 0; i < n; i++) {
                    if (done[i] || !(f[(size_t)i * m + j] > 0.0)) continue;
                    if (best < dist[i]) { dist[i] = best; prev[i] = x; }
                }
            }
        }
        if (target < 0) { ok = 0; break; } /* cannot happen: every source reaches every sink */
        const double D = dist[target];
        for (int x = 0; x < V; x++) pot[x] += (dist[x] < D) ? dist[x] : D;
        /* bottleneck along the path */
        double delta = rd[target - n];
        int x = target;
        while (prev[x] >= 0) {
            int y = prev[x];
            if (y >= n) { /* backward arc y(sink) -> x(source): limited by the flow x -> y */
                double fl = f[(size_t)x * m + (y - n)];
                if (fl < delta) delta = fl;
            }
            x = y;
        }
        if (rs[x] < delta) delta = rs[x];
        /* augment */
        rs[x] -= delta;
        rd[target - n] -= delta;
        x = target;
        while (prev[x] >= 0) {
            int y = prev[x];
            if (y < n) f[(size_t)y * m + (x - n)] += delta;      /* forward source y -> sink x */
            else {
                double *fl = &f[(size_t)x * m + (y - n)];          /* backward: undo flow x -> y */
                *fl -= delta;
                if (*fl < 0.0) *fl = 0.0;
            }
            x = y;
        }
    }
    double total = 0.0;
    for (int i = 0; i < n; i++)
        for (int j = 0; j < m; j++) total += f[(size_t)i * m + j] * cost[(size_t)i * m + j];
    if (flow) memcpy(flow, f, sizeof(double) * (size_t)n * m);
    free(f); free(rs); free(rd); free(pot); free(dist); free(prev); free(done);
    return ok ? total : -1.0;
}
