/* Oracle (TEST INFRASTRUCTURE): rectangular linear-sum-assignment, restating the algorithm of the
 * third-party dependency the reference calls at ConditionalDETR/models/matcher.py:86 —
 * scipy.optimize.linear_sum_assignment (scipy 1.18.1 in this image; unpinned in the reference's
 * requirements).  Algorithm: D. F. Crouse, "On implementing 2D rectangular assignment algorithms",
 * IEEE T-AES 52(4), 2016 — shortest augmenting paths with dual variables, no initialisation phase,
 * tall matrices solved on the transpose.  Tie-breaking follows the scan order of the `remaining`
 * list (filled in reverse, swap-remove), preferring an unassigned column among equal minima.
 * Pinned against scipy itself in tests/test_lsap_oracle.py (random, integer-tie and constant matrices).
 *
 * int fod_oracle_lsap(nr, nc, cost[nr*nc] row-major double, row_out, col_out) -> number of pairs,
 *   -1 infeasible, -2 invalid (NaN / -inf) entries.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static int64_t shortest_path(int64_t nc, const double* cost, const double* u, const double* v, int64_t* pred,
                             const int64_t* row_of_col, double* dist, int64_t start, unsigned char* in_sr,
                             unsigned char* in_sc, int64_t* todo, double* min_out) {
    double lo = 0.0;
    int64_t n_todo = nc, i = start, sink = -1;
    for (int64_t t = 0; t < nc; ++t) todo[t] = nc - t - 1;
    for (int64_t j = 0; j < nc; ++j) { in_sc[j] = 0; dist[j] = INFINITY; }
    while (sink < 0) {
        int64_t best = -1;
        double best_val = INFINITY;
        in_sr[i] = 1;
        for (int64_t t = 0; t < n_todo; ++t) {
            const int64_t j = todo[t];
            const double r = lo + cost[i * nc + j] - u[i] - v[j];
            if (r < dist[j]) { pred[j] = i; dist[j] = r; }
            if (dist[j] < best_val || (dist[j] == best_val && row_of_col[j] < 0)) { best_val = dist[j]; best = t; }
        }
        lo = best_val;
        if (lo == INFINITY) return -1;
        const int64_t j = todo[best];
        if (row_of_col[j] < 0) sink = j; else i = row_of_col[j];
        in_sc[j] = 1;
        todo[best] = todo[--n_todo];
    }
    *min_out = lo;
    return sink;
}

static int cmp_pair(const void* a, const void* b) {
    const int64_t x = ((const int64_t*)a)[0], y = ((const int64_t*)b)[0];
    return (x > y) - (x < y);
}

int fod_oracle_lsap(int64_t nr, int64_t nc, const double* cost_in, int64_t* row_out, int64_t* col_out) {
    if (nr == 0 || nc == 0) return 0;
    const int transposed = nc < nr;
    double* cost = (double*)malloc(sizeof(double) * nr * nc);
    if (transposed) {
        for (int64_t i = 0; i < nr; ++i)
            for (int64_t j = 0; j < nc; ++j) cost[j * nr + i] = cost_in[i * nc + j];
        int64_t t = nr; nr = nc; nc = t;
    } else {
        for (int64_t i = 0; i < nr * nc; ++i) cost[i] = cost_in[i];
    }
    for (int64_t i = 0; i < nr * nc; ++i)
        if (cost[i] != cost[i] || cost[i] == -INFINITY) { free(cost); return -2; }
    double* u = (double*)calloc(nr, sizeof(double));
    double* v = (double*)calloc(nc, sizeof(double));
    double* dist = (double*)malloc(sizeof(double) * nc);
    int64_t* pred = (int64_t*)malloc(sizeof(int64_t) * nc);
    int64_t* col_of_row = (int64_t*)malloc(sizeof(int64_t) * nr);
    int64_t* row_of_col = (int64_t*)malloc(sizeof(int64_t) * nc);
    int64_t* todo = (int64_t*)malloc(sizeof(int64_t) * nc);
    unsigned char* in_sr = (unsigned char*)malloc(nr);
    unsigned char* in_sc = (unsigned char*)malloc(nc);
    for (int64_t i = 0; i < nr; ++i) col_of_row[i] = -1;
    for (int64_t j = 0; j < nc; ++j) { row_of_col[j] = -1; pred[j] = -1; }
    int rc = 0;
    for (int64_t cur = 0; cur < nr; ++cur) {
        double lo;
        for (int64_t i = 0; i < nr; ++i) in_sr[i] = 0;
        const int64_t sink = shortest_path(nc, cost, u, v, pred, row_of_col, dist, cur, in_sr, in_sc, todo, &lo);
        if (sink < 0) { rc = -1; break; }
        u[cur] += lo;
        for (int64_t i = 0; i < nr; ++i)
            if (in_sr[i] && i != cur) u[i] += lo - dist[col_of_row[i]];
        for (int64_t j = 0; j < nc; ++j)
            if (in_sc[j]) v[j] -= lo - dist[j];
        int64_t j = sink;
        for (;;) {
            const int64_t i = pred[j];
            row_of_col[j] = i;
            const int64_t prev = col_of_row[i];
            col_of_row[i] = j;
            j = prev;
            if (i == cur) break;
        }
    }
    if (rc == 0) {
        if (transposed) {
            int64_t* pairs = (int64_t*)malloc(sizeof(int64_t) * 2 * nr);
            for (int64_t i = 0; i < nr; ++i) { pairs[2 * i] = col_of_row[i]; pairs[2 * i + 1] = i; }
            qsort(pairs, nr, 2 * sizeof(int64_t), cmp_pair);
            for (int64_t i = 0; i < nr; ++i) { row_out[i] = pairs[2 * i]; col_out[i] = pairs[2 * i + 1]; }
            free(pairs);
        } else {
            for (int64_t i = 0; i < nr; ++i) { row_out[i] = i; col_out[i] = col_of_row[i]; }
        }
        rc = (int)nr;
    }
    free(cost); free(u); free(v); free(dist); free(pred); free(col_of_row); free(row_of_col); free(todo);
    free(in_sr); free(in_sc);
    return rc;
}
