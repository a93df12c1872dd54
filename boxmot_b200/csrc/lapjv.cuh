// Warp-cooperative dense Jonker-Volgenant assignment solver (device code, sm_100a).
//
// Replaces the reference's calls into lap.lapjv (lapx 0.5.11.post1):
//   boxmot/utils/matching.py:54-55, boxmot/utils/association.py:50-51,109,
//   boxmot/trackers/boosttrack/assoc.py:113
// with the (nr+nc)^2 extension of `extend_cost=True` / `cost_limit` kept IMPLICIT (the
// pad and zero blocks are constants, never stored).
//
// One warp solves one problem.  All per-problem vectors (v, d, x, y, pred, cols,
// free_rows) live in shared memory; the cost block is reached through a generic
// pointer (shared memory when it fits, otherwise an L2-resident global scratch).
// The sequential structure of lapjv is preserved step for step -- every minimum is
// resolved to the lowest (value, index) exactly as the strict `<` scans of the scalar
// code do, and the TODO/SCAN/READY column permutation is updated in the same order --
// so the assignment equals oracle/lapjv.c on every input, exact ties included.
// The one liberty taken is provably result-neutral: a SCAN column owned by a dummy
// row is not swept when an earlier dummy-row sweep of the same path had h >= its h
// (all dummy rows hold identical entries, so the sweep cannot lower any d[j]);
// oracle/lapjv.c:oracle_lapjv_skip is the CPU model of exactly this rule.
#pragma once
#include "common.cuh"
#include <math_constants.h>

namespace bm {

#define BM_LAP_LARGE 1000000.0

struct LapMat {
    const double* c;  // nr x nc block, row-major, leading dimension ld
    int ld;
    int nr, nc, n;
    int extended;     // 1: implicit (nr+nc)^2 extension; 0: square nr x nr
    double pad;       // value of the two off-diagonal blocks
};

struct LapWork {      // n entries each, shared memory
    double* v;
    double* d;
    int* x;
    int* y;
    int* pred;
    int* cols;
    int* free_rows;
    int* aux;
};

__device__ __forceinline__ double lap_at(const LapMat& m, int i, int j) {
    if (!m.extended) return m.c[(size_t)i * m.ld + j];
    if (i < m.nr) return j < m.nc ? m.c[(size_t)i * m.ld + j] : m.pad;
    return j < m.nc ? m.pad : 0.0;
}

struct LapKey {
    double c;
    int j;
};
__device__ __forceinline__ bool lap_less(const LapKey& a, const LapKey& b) {
    return a.c < b.c || (a.c == b.c && a.j < b.j);
}

// Warp minimum of a double (all lanes get the result).
__device__ __forceinline__ double warp_min_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double t = shfl_xor_d(v, o);
        v = t < v ? t : v;
    }
    return v;
}

// Phase 1: column reduction + reduction transfer.  Returns the number of free rows.
static __device__ int lap_ccrrt(const LapMat& m, LapWork& w) {
    const int n = m.n, lane = lane_id();
    for (int j = lane; j < n; j += 32) {
        double vm = BM_LAP_LARGE;
        int yi = 0;
        if (!m.extended) {
            for (int i = 0; i < n; i++) {
                double c = m.c[(size_t)i * m.ld + j];
                if (c < vm) { vm = c; yi = i; }
            }
        } else if (j < m.nc) {
            for (int i = 0; i < m.nr; i++) {
                double c = m.c[(size_t)i * m.ld + j];
                if (c < vm) { vm = c; yi = i; }
            }
            if (m.pad < vm) { vm = m.pad; yi = m.nr; }   // first dummy row
        } else {
            if (m.pad < vm) { vm = m.pad; yi = 0; }      // real rows all hold pad
            if (0.0 < vm) { vm = 0.0; yi = m.nr; }       // first dummy row holds 0
        }
        w.v[j] = vm;
        w.y[j] = yi;
        w.x[j] = -1;
        w.aux[j] = 0;
    }
    __syncwarp();
    // Reverse-order claim: row i keeps the LAST (largest) column that chose it.
    for (int j = lane; j < n; j += 32) {
        atomicMax(&w.x[w.y[j]], j);
        atomicAdd(&w.aux[w.y[j]], 1);
    }
    __syncwarp();
    for (int j = lane; j < n; j += 32) {
        int i = w.y[j];
        if (w.x[i] != j) w.y[j] = -1;
    }
    __syncwarp();
    int n_free = 0;
    for (int i = 0; i < n; i++) {
        const int xi = w.x[i];
        if (xi < 0) {
            if (lane == 0) w.free_rows[n_free] = i;
            n_free++;
        } else if (w.aux[i] == 1) {
            double mn = BM_LAP_LARGE;
            for (int j2 = lane; j2 < n; j2 += 32) {
                if (j2 == xi) continue;
                double c = lap_at(m, i, j2) - w.v[j2];
                if (c < mn) mn = c;
            }
            mn = warp_min_d(mn);
            if (lane == 0) w.v[xi] -= mn;
            __syncwarp();
        }
    }
    __syncwarp();
    return n_free;
}

// Phase 2: augmenting row reduction.
static __device__ int lap_carr(const LapMat& m, LapWork& w, int n_free) {
    const int n = m.n, lane = lane_id();
    int current = 0, new_free = 0;
    long long rr_cnt = 0;
    const long long guard = 4LL * n * n + 64;
    while (current < n_free && rr_cnt < guard) {
        rr_cnt++;
        const int fi = w.free_rows[current++];
        LapKey b1 = {CUDART_INF, 0x7fffffff}, b2 = {CUDART_INF, 0x7fffffff};
        for (int j = lane; j < n; j += 32) {
            LapKey k = {lap_at(m, fi, j) - w.v[j], j};
            if (lap_less(k, b1)) { b2 = b1; b1 = k; }
            else if (lap_less(k, b2)) { b2 = k; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            LapKey o1 = {shfl_xor_d(b1.c, o), __shfl_xor_sync(BM_FULL, b1.j, o)};
            LapKey o2 = {shfl_xor_d(b2.c, o), __shfl_xor_sync(BM_FULL, b2.j, o)};
            LapKey n1, n2;
            if (lap_less(b1, o1)) { n1 = b1; n2 = lap_less(b2, o1) ? b2 : o1; }
            else { n1 = o1; n2 = lap_less(o2, b1) ? o2 : b1; }
            b1 = n1; b2 = n2;
        }
        int j1 = b1.j, j2;
        double v1 = b1.c, v2;
        if (b2.c < BM_LAP_LARGE) { v2 = b2.c; j2 = b2.j; } else { v2 = BM_LAP_LARGE; j2 = -1; }
        int i0 = w.y[j1];
        const double vj1 = w.v[j1];
        const double v1_new = vj1 - (v2 - v1);
        const bool lowers = v1_new < vj1;
        __syncwarp();
        if (rr_cnt < (long long)current * n) {
            if (lowers) { if (lane == 0) w.v[j1] = v1_new; }
            else if (i0 >= 0 && j2 >= 0) { j1 = j2; i0 = w.y[j2]; }
            if (i0 >= 0) {
                if (lowers) { --current; if (lane == 0) w.free_rows[current] = i0; }
                else { if (lane == 0) w.free_rows[new_free] = i0; new_free++; }
            }
        } else if (i0 >= 0) {
            if (lane == 0) w.free_rows[new_free] = i0;
            new_free++;
        }
        if (lane == 0) { w.x[fi] = j1; w.y[j1] = fi; }
        __syncwarp();
    }
    return new_free;
}

// Phase 3 helper: one shortest augmenting path from free row `start` (modified
// Dijkstra over the TODO/SCAN/READY partition of `cols`).  Returns the free column
// that ends the path, or -1 when the iteration guard trips (non-finite costs).
static __device__ int lap_find_path(const LapMat& m, LapWork& w, int start) {
    const int n = m.n, lane = lane_id();
    for (int j = lane; j < n; j += 32) {
        w.cols[j] = j;
        w.pred[j] = start;
        w.d[j] = lap_at(m, start, j) - w.v[j];
    }
    __syncwarp();
    int lo = 0, hi = 0, n_ready = 0, final_j = -1;
    bool hm_set = false;
    double hmax = 0.0;
    int guard = 4 * n + 16;
    while (final_j < 0 && --guard > 0) {
        if (lo == hi) {
            n_ready = lo;
            // ---- move the columns of minimal d to the front of TODO (new SCAN list)
            double mind = w.d[w.cols[lo]];
            hi = lo + 1;
            for (int base = lo + 1; base < n; base += 32) {
                const int k = base + lane;
                const bool ok = k < n;
                const int j = ok ? w.cols[k] : 0;
                const double dj = ok ? w.d[j] : CUDART_INF;
                unsigned open = BM_FULL;   // lanes not yet passed by the scalar scan
                for (;;) {
                    const unsigned cand = __ballot_sync(BM_FULL, ok && dj <= mind) & open;
                    if (!cand) break;
                    const int e = __ffs(cand) - 1;
                    const double dje = shfl_d(dj, e);
                    const int je = __shfl_sync(BM_FULL, j, e);
                    if (dje < mind) { hi = lo; mind = dje; }
                    if (lane == 0) { w.cols[base + e] = w.cols[hi]; w.cols[hi] = je; }
                    hi++;
                    open = e == 31 ? 0u : (BM_FULL << (e + 1));
                }
                __syncwarp();
            }
            // ---- a free column already at the minimum ends the path (last one wins)
            for (int base = lo; base < hi; base += 32) {
                const int k = base + lane;
                const bool ok = k < hi;
                const int j = ok ? w.cols[k] : 0;
                const unsigned fb = __ballot_sync(BM_FULL, ok && w.y[j] < 0);
                if (fb) final_j = __shfl_sync(BM_FULL, j, 31 - __clz(fb));
            }
        }
        if (final_j < 0) {
            int l = lo, h = hi;
            while (l != h && final_j < 0) {
                // ---- pick the next SCAN column whose sweep can have an effect
                const int ks = l + lane;
                const bool oks = ks < h;
                const int sj0 = oks ? w.cols[ks] : 0;
                const int si0 = oks ? w.y[sj0] : 0;
                const double sm0 = oks ? w.d[sj0] : 0.0;
                const double sh0 = oks ? (lap_at(m, si0, sj0) - w.v[sj0]) - sm0 : 0.0;
                const bool dummy0 = m.extended && si0 >= m.nr;
                const unsigned need =
                    __ballot_sync(BM_FULL, oks && !(dummy0 && hm_set && sh0 <= hmax));
                if (!need) { l = (l + 32 < h) ? l + 32 : h; continue; }
                const int e0 = __ffs(need) - 1;
                const int si = __shfl_sync(BM_FULL, si0, e0);
                const double smind = shfl_d(sm0, e0);
                const double sh = shfl_d(sh0, e0);
                if (__shfl_sync(BM_FULL, (int)dummy0, e0)) { hmax = sh; hm_set = true; }
                l += e0 + 1;
                // ---- relax every TODO column through row si
                for (int base = h; base < n && final_j < 0; base += 32) {
                    const int k = base + lane;
                    const bool ok = k < n;
                    const int jj = ok ? w.cols[k] : 0;
                    bool ev = false;
                    if (ok) {
                        const double cred = (lap_at(m, si, jj) - w.v[jj]) - sh;
                        if (cred < w.d[jj]) {
                            w.d[jj] = cred;
                            w.pred[jj] = si;
                            ev = cred == smind;
                        }
                    }
                    unsigned eb = __ballot_sync(BM_FULL, ev);
                    while (eb) {
                        const int e2 = __ffs(eb) - 1;
                        eb &= eb - 1;
                        const int je = __shfl_sync(BM_FULL, jj, e2);
                        if (w.y[je] < 0) { final_j = je; break; }
                        if (lane == 0) { w.cols[base + e2] = w.cols[h]; w.cols[h] = je; }
                        h++;
                    }
                    __syncwarp();
                }
            }
            if (final_j < 0) { lo = l; hi = h; }
        }
    }
    if (final_j < 0) return -1;
    const double mind = w.d[w.cols[lo]];
    for (int k = lane; k < n_ready; k += 32) {
        const int j = w.cols[k];
        w.v[j] += w.d[j] - mind;
    }
    __syncwarp();
    return final_j;
}

// Solve the assignment problem; on return w.x[i] (i < n) / w.y[j] (j < n) hold the
// extended assignment.  Returns 0, or 1 when a guard tripped (non-finite costs).
static __device__ int lap_solve_warp(const LapMat& m, LapWork& w) {
    const int lane = lane_id();
    int n_free = lap_ccrrt(m, w);
    for (int it = 0; n_free > 0 && it < 2; it++) n_free = lap_carr(m, w, n_free);
    int bad = 0;
    for (int f = 0; f < n_free; f++) {
        const int start = w.free_rows[f];
        const int fj = lap_find_path(m, w, start);
        if (fj < 0) { bad = 1; break; }
        if (lane == 0) {
            int i = -1, j = fj, hops = 0;
            while (i != start && hops++ <= m.n) {
                i = w.pred[j];
                w.y[j] = i;
                const int t = w.x[i];
                w.x[i] = j;
                j = t;
            }
        }
        __syncwarp();
    }
    return bad;
}

// Shared-memory bytes lap_solve_warp needs for an n-column problem (8-byte aligned).
__host__ __device__ inline size_t lap_work_bytes(int n) {
    return (size_t)n * (2 * sizeof(double) + 6 * sizeof(int));
}

__device__ inline LapWork lap_carve(void* smem, int n) {
    LapWork w;
    w.v = reinterpret_cast<double*>(smem);
    w.d = w.v + n;
    w.x = reinterpret_cast<int*>(w.d + n);
    w.y = w.x + n;
    w.pred = w.y + n;
    w.cols = w.pred + n;
    w.free_rows = w.cols + n;
    w.aux = w.free_rows + n;
    return w;
}

}  // namespace bm
