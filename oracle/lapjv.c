/*
 * oracle/lapjv.c -- TEST INFRASTRUCTURE ONLY (CPU oracle; never on the product path).
 *
 * CPU restatement of the dense Jonker-Volgenant linear assignment solver that the
 * reference calls as `lap.lapjv` (PyPI `lapx`, pinned 0.5.11.post1 in
 * /root/reference/uv.lock:1789-1795; dependency line pyproject.toml:121).  The lapx
 * source is NOT under /root/reference and is not installed in this image, so this
 * file restates the published algorithm (R. Jonker, A. Volgenant, "A shortest
 * augmenting path algorithm for dense and sparse linear assignment problems",
 * Computing 38, 1987) in the four-phase structure lapjv implementations use:
 * column reduction + reduction transfer, augmenting row reduction (twice),
 * shortest-augmenting-path augmentation with the TODO/SCAN/READY column partition.
 *
 * PARITY STATUS: "parity unpinned" for tie-breaking -- with exact cost ties the
 * optimum is not unique and lapx's scan order decides; that order cannot be checked
 * here.  The optimum VALUE is pinned against scipy.optimize.linear_sum_assignment
 * (tests/test_oracle_lap.py) and the extension semantics against the reference's
 * call sites:
 *   boxmot/utils/matching.py:54-55        lapjv(cost, extend_cost=True, cost_limit=thresh)
 *   boxmot/utils/association.py:50-51,109 lapjv(cost, extend_cost=True[, cost_limit])
 *   boxmot/trackers/boosttrack/assoc.py:113  lapjv(-cost, extend_cost=True)
 *
 * The CUDA solver (boxmot_b200/csrc/lapjv.cuh) follows the same scan order, so the
 * two agree bit-for-bit on every input, ties included.
 */
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define LARGE 1000000.0

typedef double cost_t;

/* Implicit extended matrix: rows [0,nr) x cols [0,nc) hold `c` (row-major, ld = nc);
 * the two off-diagonal blocks hold `pad`; the bottom-right block holds 0.
 * When `extended` is 0 the matrix is the square nr x nr `c` itself. */
typedef struct {
    const cost_t *c;
    int nr, nc, n, extended;
    cost_t pad;
    long scans; /* instrumentation: number of full row/column sweeps */
    int skip_mode;      /* 1: model of the GPU shortcut (see scan_cols) */
    int hm_set; cost_t hmax;
} mat_t;

static inline cost_t at(const mat_t *m, int i, int j)
{
    if (!m->extended) return m->c[(size_t)i * m->nc + j];
    if (i < m->nr) return j < m->nc ? m->c[(size_t)i * m->nc + j] : m->pad;
    return j < m->nc ? m->pad : 0.0;
}

/* Phase 1: column reduction and reduction transfer. */
static int ccrrt(mat_t *m, int *free_rows, int *x, int *y, cost_t *v)
{
    const int n = m->n;
    int n_free = 0;
    char *unique = (char *)malloc((size_t)n);
    for (int i = 0; i < n; i++) { x[i] = -1; v[i] = LARGE; y[i] = 0; }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            const cost_t c = at(m, i, j);
            if (c < v[j]) { v[j] = c; y[j] = i; }
        }
    m->scans += n;
    memset(unique, 1, (size_t)n);
    for (int j = n - 1; j >= 0; j--) {
        const int i = y[j];
        if (x[i] < 0) x[i] = j;
        else { unique[i] = 0; y[j] = -1; }
    }
    for (int i = 0; i < n; i++) {
        if (x[i] < 0) free_rows[n_free++] = i;
        else if (unique[i]) {
            const int j = x[i];
            cost_t mn = LARGE;
            for (int j2 = 0; j2 < n; j2++) {
                if (j2 == j) continue;
                const cost_t c = at(m, i, j2) - v[j2];
                if (c < mn) mn = c;
            }
            m->scans++;
            v[j] -= mn;
        }
    }
    free(unique);
    return n_free;
}

/* Phase 2: augmenting row reduction. */
static int carr(mat_t *m, int n_free, int *free_rows, int *x, int *y, cost_t *v)
{
    const int n = m->n;
    int current = 0, new_free = 0;
    long rr_cnt = 0;
    while (current < n_free) {
        int i0, j1, j2;
        cost_t v1, v2, v1_new;
        int v1_lowers;
        rr_cnt++;
        const int fi = free_rows[current++];
        j1 = 0; v1 = at(m, fi, 0) - v[0];
        j2 = -1; v2 = LARGE;
        for (int j = 1; j < n; j++) {
            const cost_t c = at(m, fi, j) - v[j];
            if (c < v2) {
                if (c >= v1) { v2 = c; j2 = j; }
                else { v2 = v1; v1 = c; j2 = j1; j1 = j; }
            }
        }
        m->scans++;
        i0 = y[j1];
        v1_new = v[j1] - (v2 - v1);
        v1_lowers = v1_new < v[j1];
        if (rr_cnt < (long)current * n) {
            if (v1_lowers) v[j1] = v1_new;
            else if (i0 >= 0 && j2 >= 0) { j1 = j2; i0 = y[j2]; }
            if (i0 >= 0) {
                if (v1_lowers) free_rows[--current] = i0;
                else free_rows[new_free++] = i0;
            }
        } else {
            if (i0 >= 0) free_rows[new_free++] = i0;
        }
        x[fi] = j1;
        y[j1] = fi;
    }
    return new_free;
}

/* Move the columns with minimal d from TODO to the front (SCAN list). */
static int find_min_cols(int n, int lo, const cost_t *d, int *cols)
{
    int hi = lo + 1;
    cost_t mind = d[cols[lo]];
    for (int k = hi; k < n; k++) {
        const int j = cols[k];
        if (d[j] <= mind) {
            if (d[j] < mind) { hi = lo; mind = d[j]; }
            cols[k] = cols[hi];
            cols[hi++] = j;
        }
    }
    return hi;
}

/* Relax TODO columns through every SCAN column; return a free column reached at
 * the current minimum, or -1. */
static int scan_cols(mat_t *m, int *plo, int *phi, cost_t *d, int *cols, int *pred,
                     const int *y, const cost_t *v)
{
    const int n = m->n;
    int lo = *plo, hi = *phi;
    while (lo != hi) {
        int j = cols[lo++];
        const int i = y[j];
        const cost_t mind = d[j];
        const cost_t h = at(m, i, j) - v[j] - mind;
        if (m->skip_mode && m->extended && i >= m->nr) {
            /* All dummy rows hold the same entries, so a sweep from a dummy row with
             * h <= an earlier dummy sweep's h (same path) cannot lower any d[j]:
             * (c - v[j]) - h is monotone in h.  The CUDA solver skips such sweeps;
             * this branch is its CPU model (results must equal the plain algorithm). */
            if (m->hm_set && h <= m->hmax) continue;
            m->hmax = h; m->hm_set = 1;
        }
        m->scans++;
        for (int k = hi; k < n; k++) {
            j = cols[k];
            const cost_t cred = at(m, i, j) - v[j] - h;
            if (cred < d[j]) {
                d[j] = cred;
                pred[j] = i;
                if (cred == mind) {
                    if (y[j] < 0) return j;
                    cols[k] = cols[hi];
                    cols[hi++] = j;
                }
            }
        }
    }
    *plo = lo; *phi = hi;
    return -1;
}

/* One modified-Dijkstra shortest augmenting path from free row `start`. */
static int find_path(mat_t *m, int start, const int *y, cost_t *v, int *pred,
                     int *cols, cost_t *d)
{
    const int n = m->n;
    int lo = 0, hi = 0, final_j = -1, n_ready = 0;
    m->hm_set = 0;
    for (int j = 0; j < n; j++) {
        cols[j] = j; pred[j] = start; d[j] = at(m, start, j) - v[j];
    }
    m->scans++;
    while (final_j == -1) {
        if (lo == hi) {
            n_ready = lo;
            hi = find_min_cols(n, lo, d, cols);
            m->scans++;
            for (int k = lo; k < hi; k++) {
                const int j = cols[k];
                if (y[j] < 0) final_j = j;
            }
        }
        if (final_j == -1)
            final_j = scan_cols(m, &lo, &hi, d, cols, pred, y, v);
    }
    {
        const cost_t mind = d[cols[lo]];
        for (int k = 0; k < n_ready; k++) {
            const int j = cols[k];
            v[j] += d[j] - mind;
        }
    }
    return final_j;
}

/* Phase 3: augmentation for the remaining free rows. */
static void augment(mat_t *m, int n_free, const int *free_rows, int *x, int *y, cost_t *v)
{
    const int n = m->n;
    int *pred = (int *)malloc(sizeof(int) * (size_t)n);
    int *cols = (int *)malloc(sizeof(int) * (size_t)n);
    cost_t *d = (cost_t *)malloc(sizeof(cost_t) * (size_t)n);
    for (int f = 0; f < n_free; f++) {
        const int fi = free_rows[f];
        int i = -1;
        int j = find_path(m, fi, y, v, pred, cols, d);
        while (i != fi) {
            i = pred[j];
            y[j] = i;
            const int t = x[i]; x[i] = j; j = t;
        }
    }
    free(pred); free(cols); free(d);
}

static void solve_square(mat_t *m, int *x, int *y)
{
    const int n = m->n;
    int *free_rows = (int *)malloc(sizeof(int) * (size_t)n);
    cost_t *v = (cost_t *)malloc(sizeof(cost_t) * (size_t)n);
    int ret = ccrrt(m, free_rows, x, y, v);
    for (int it = 0; ret > 0 && it < 2; it++)
        ret = carr(m, ret, free_rows, x, y, v);
    if (ret > 0) augment(m, ret, free_rows, x, y, v);
    free(v); free(free_rows);
}

/*
 * oracle_lapjv: the `lap.lapjv(cost, extend_cost, cost_limit)` contract.
 *   cost        row-major nr x nc doubles
 *   extend      non-zero: pad to (nr+nc)^2 (off-diagonal blocks cost_limit/2 when
 *               cost_limit is finite, else max(cost)+1; zero dummy block)
 *   cost_limit  >= DBL_MAX / inf means "no limit"; a finite limit implies extension
 *   x[nr], y[nc] receive the assignment (-1 = unassigned)
 * Returns 0, or -1 for a non-square matrix without extension.  *scans (optional)
 * receives the number of row sweeps performed (used to size the GPU work estimate).
 */
static int lapjv_impl(const double *cost, int nr, int nc, int extend, double cost_limit,
                      int *x, int *y, long *scans, int skip_mode)
{
    mat_t m;
    m.skip_mode = skip_mode; m.hm_set = 0; m.hmax = 0.0;
    const int limited = cost_limit < DBL_MAX;
    m.c = cost; m.nr = nr; m.nc = nc; m.scans = 0;
    if (scans) *scans = 0;
    if (nr == 0 || nc == 0) {
        for (int i = 0; i < nr; i++) x[i] = -1;
        for (int j = 0; j < nc; j++) y[j] = -1;
        return 0;
    }
    if (extend || limited) {
        m.extended = 1; m.n = nr + nc;
        if (limited) m.pad = cost_limit / 2.0;
        else {
            cost_t mx = cost[0];
            for (size_t k = 1; k < (size_t)nr * nc; k++) if (cost[k] > mx) mx = cost[k];
            m.pad = mx + 1.0;
        }
    } else {
        if (nr != nc) return -1;
        m.extended = 0; m.n = nr; m.pad = 0.0;
    }
    int *xe = (int *)malloc(sizeof(int) * (size_t)m.n);
    int *ye = (int *)malloc(sizeof(int) * (size_t)m.n);
    solve_square(&m, xe, ye);
    for (int i = 0; i < nr; i++) x[i] = (m.extended && xe[i] >= nc) ? -1 : xe[i];
    for (int j = 0; j < nc; j++) y[j] = (m.extended && ye[j] >= nr) ? -1 : ye[j];
    if (scans) *scans = m.scans;
    free(xe); free(ye);
    return 0;
}

int oracle_lapjv(const double *cost, int nr, int nc, int extend, double cost_limit,
                 int *x, int *y, long *scans)
{
    return lapjv_impl(cost, nr, nc, extend, cost_limit, x, y, scans, 0);
}

/* CPU model of the CUDA solver's exact shortcut (tests assert it equals oracle_lapjv). */
int oracle_lapjv_skip(const double *cost, int nr, int nc, int extend, double cost_limit,
                      int *x, int *y, long *scans)
{
    return lapjv_impl(cost, nr, nc, extend, cost_limit, x, y, scans, 1);
}
