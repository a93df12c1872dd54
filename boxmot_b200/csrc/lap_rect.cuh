// Warp-cooperative exact solver for the rectangular assignment problem with a "null"
// option -- the engine's fast path for every assignment the reference sends to
// lap.lapjv(cost, extend_cost=True[, cost_limit]) (boxmot/utils/association.py:50-51,109,
// boxmot/utils/matching.py:54-55, boxmot/trackers/boosttrack/assoc.py:113).
//
// lapjv's (nr+nc)^2 extension (real block = cost, off-diagonal blocks = pad, dummy block
// = 0) is equivalent to: every real row either takes a real column at c[i][j] or stays
// unmatched, and each matched pair saves 2*pad.  So the optimum of the extended problem
// restricted to the real block is the optimum of
//        min  sum_{matched (i,j)} (c[i][j] - null_cost),      null_cost = 2*pad
// (= cost_limit when a limit is given, 2*(max(c)+1) otherwise).  That problem is solved
// here directly by successive shortest augmenting paths over the REAL rows only
// (Crouse-style row/column duals, all columns start free with v = 0), with the null
// option as an always-free sink of reduced cost null_cost - u[i].  No dummy rows are
// ever augmented, which is what makes it ~100x shorter than emulating lapjv on the
// extension (csrc/lapjv.cuh keeps that step-for-step emulation for the cases where
// lapjv's internal tie-breaking is observable, see ocsort_step.cu).
//
// Exactness: the result is an optimal assignment in fp64.  When the optimum is unique
// it equals lapjv's x/y (tests/test_gpu_units.py); when it is not, any optimal solution
// may be returned.
#pragma once
#include "common.cuh"
#include <math_constants.h>

namespace bm {

struct RectWork {     // shared memory, carved by rect_carve
    double *u, *v, *spc, *scd;   // [nr] [nc] [nc] [nc]
    int *x, *y, *path, *sc, *fr; // [nr] [nc] [nc] [nc] [nr]
};

__host__ __device__ inline size_t rect_work_bytes(int nr, int nc) {
    return (size_t)nr * 8 + (size_t)nc * 24 + ((size_t)nr * 2 + (size_t)nc * 3) * 4 + 8;
}

__device__ inline RectWork rect_carve(void* smem, int nr, int nc) {
    RectWork w;
    w.u = reinterpret_cast<double*>(smem);
    w.v = w.u + nr;
    w.spc = w.v + nc;
    w.scd = w.spc + nc;
    w.x = reinterpret_cast<int*>(w.scd + nc);
    w.y = w.x + nr;
    w.path = w.y + nc;
    w.sc = w.path + nc;
    w.fr = w.sc + nc;
    return w;
}

// (value, index) arg-min over the warp; ties go to the lower index.
__device__ __forceinline__ void warp_argmin(double& s, int& j) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double so = shfl_xor_d(s, o);
        const int jo = __shfl_xor_sync(BM_FULL, j, o);
        if (so < s || (so == s && jo < j)) { s = so; j = jo; }
    }
}

__device__ __forceinline__ double warp_min_rect(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, shfl_xor_d(v, o));
    return v;
}

// Cost accessor over a stored row-major matrix.
struct DenseCost {
    static constexpr bool kBounded = false, kStaged = false;
    const double* c; int ld;
    __device__ __forceinline__ double operator()(int i, int j) const { return c[(size_t)i * ld + j]; }
    __device__ __forceinline__ double lower(int i, int j, bool* exact) const { *exact = true; return (*this)(i, j); }
};

// Solve with one warp.  cost(i, j): entry accessor (a stored matrix, or a functor that
// evaluates the entry on demand -- the solver only touches the rows it sweeps).
// A cost functor with kStaged = true gets cost.stage(i, spc) called (by the whole warp,
// converged) before every sweep of row i -- it may evaluate the row into a buffer with the
// help of other warps, skipping columns with spc[j] == -inf -- and must be given rowmin.
// rowmin/rowarg: optional per-row minimum and its column (computed by the caller while
// the costs were generated); pass nullptr to have the warp compute them.
// On return w.x[i] = column of row i or -1, w.y[j] = row of column j or -1.
// Returns 0, or 1 when a guard tripped (NaN costs).
template <typename CostFn>
static __device__ int rect_solve_warp(const CostFn& cost, int nr, int nc, double null_cost,
                                      RectWork& w, const double* rowmin, const int* rowarg) {
    const int lane = lane_id();
    for (int j = lane; j < nc; j += 32) { w.v[j] = 0.0; w.y[j] = 0x7fffffff; }
    __syncwarp();
    // ---- row reduction: u[i] = min_j c[i][j] (capped by the null option), greedy claim
    if (rowmin) {
        for (int i = lane; i < nr; i += 32) {
            double m = rowmin[i];
            int a = rowarg[i];
            if (!(m <= null_cost)) { m = null_cost; a = -1; }
            w.u[i] = m;
            w.x[i] = a;
        }
    } else if constexpr (CostFn::kStaged) {
        return 1;            // staged cost functors must come with row minima
    } else {
        for (int i = 0; i < nr; i++) {
            double m = CUDART_INF;
            int a = 0x7fffffff;
            for (int j = lane; j < nc; j += 32) {
                const double cc = cost(i, j);
                if (cc < m) { m = cc; a = j; }
            }
            warp_argmin(m, a);
            if (!(m <= null_cost)) { m = null_cost; a = -1; }
            if (lane == 0) { w.u[i] = m; w.x[i] = a; }
        }
    }
    __syncwarp();
    for (int i = lane; i < nr; i += 32) { const int a = w.x[i]; if (a >= 0) atomicMin(&w.y[a], i); }
    __syncwarp();
    int n_free = 0;
    for (int base = 0; base < nr; base += 32) {
        const int i = base + lane;
        bool fr = false;
        if (i < nr) {
            // a < 0: the row's cheapest option is the null sink (reduced cost 0 with u = null_cost);
            // it is assigned there for good -- it owns no column, so no later path runs through it
            const int a = w.x[i];
            if (a >= 0 && w.y[a] != i) { fr = true; w.x[i] = -1; }
        }
        const unsigned b = __ballot_sync(BM_FULL, fr);
        if (fr) w.fr[n_free + __popc(b & ((1u << lane) - 1))] = i;
        n_free += __popc(b);
    }
    for (int j = lane; j < nc; j += 32) if (w.y[j] == 0x7fffffff) w.y[j] = -1;
    __syncwarp();

    // ---- one shortest augmenting path per free row
    int bad = 0;
    for (int f = 0; f < n_free && !bad; f++) {
        const int r0 = w.fr[f];
        for (int j = lane; j < nc; j += 32) w.spc[j] = CUDART_INF;
        __syncwarp();
        double minval = 0.0, best_null = CUDART_INF, bfree = CUDART_INF;
        int i = r0, null_row = -1, nsc = 0, sink = -2;
        int guard = nc + 2;
        while (sink == -2) {
            if (--guard < 0) { bad = 1; break; }
            const double ui = w.u[i];
            const double cn = minval + (null_cost - ui);
            if (cn < best_null) { best_null = cn; null_row = i; }
            double lowest = CUDART_INF;
            int jl = 0x7fffffff;
            if constexpr (CostFn::kStaged) cost.stage(i, w.spc);
            if constexpr (CostFn::kBounded) {
                // Expensive entries: cost.lower(i, j) is a cheap rigorous lower bound.  `bfree` is the
                // exact distance of the best free column (or the null option) found so far, an upper
                // bound of the final path length; an entry whose bound cannot beat min(spc[j], bfree)
                // can neither lie on the shortest path nor be selected before the search ends, and
                // the duals only read the distances of SCANNED columns, so it is never evaluated.
                double lbf = CUDART_INF;
                int jf = 0x7fffffff;
                for (int j = lane; j < nc; j += 32) {          // pass A: most promising free column
                    if (w.spc[j] == -CUDART_INF || w.y[j] >= 0) continue;
                    bool ex;
                    const double r = minval + ((cost.lower(i, j, &ex) - ui) - w.v[j]);
                    if (r < lbf) { lbf = r; jf = j; }
                }
                warp_argmin(lbf, jf);
                if (jf != 0x7fffffff && lane == (jf & 31)) {
                    const double r = minval + ((cost(i, jf) - ui) - w.v[jf]);
                    if (r < w.spc[jf]) { w.spc[jf] = r; w.path[jf] = i; }
                }
                __syncwarp();
                if (jf != 0x7fffffff) bfree = fmin(bfree, w.spc[jf]);
                bfree = fmin(bfree, best_null);
                double nb = CUDART_INF;                         // best free distance after this sweep
                for (int j = lane; j < nc; j += 32) {          // pass B
                    double s = w.spc[j];
                    if (s == -CUDART_INF) continue;
                    bool ex;
                    double r = minval + ((cost.lower(i, j, &ex) - ui) - w.v[j]);
                    if (r < fmin(s, bfree) || (ex && r < s)) {
                        if (!ex) r = minval + ((cost(i, j) - ui) - w.v[j]);
                        if (r < s) { s = r; w.spc[j] = r; w.path[j] = i; }
                    }
                    if (s < lowest) { lowest = s; jl = j; }
                    if (w.y[j] < 0) nb = fmin(nb, s);
                }
                nb = warp_min_rect(nb);
                bfree = fmin(bfree, nb);
            } else
            for (int j = lane; j < nc; j += 32) {
                double s = w.spc[j];
                if (s == -CUDART_INF) continue;          // already scanned
                const double r = minval + ((cost(i, j) - ui) - w.v[j]);
                if (r < s) { s = r; w.spc[j] = r; w.path[j] = i; }
                if (s < lowest) { lowest = s; jl = j; }
            }
            warp_argmin(lowest, jl);
            if (!(lowest < best_null)) { sink = -1; break; }      // null option ends the path
            minval = lowest;
            const int yj = w.y[jl];
            __syncwarp();
            if (lane == 0) { w.sc[nsc] = jl; w.scd[nsc] = lowest; w.spc[jl] = -CUDART_INF; }
            nsc++;
            __syncwarp();
            if (yj < 0) sink = jl; else i = yj;
        }
        if (bad) break;
        const double fin = sink == -1 ? best_null : minval;
        // dual updates (rows first: they read y[] before the path is flipped)
        for (int k = lane; k < nsc; k += 32) {
            const int j = w.sc[k];
            const double dk = fin - w.scd[k];
            const int yi = w.y[j];
            if (yi >= 0) w.u[yi] += dk;
            w.v[j] -= dk;
        }
        if (lane == 0) w.u[r0] += fin;
        __syncwarp();
        if (lane == 0) {
            int hops = 0;
            if (sink == -1) {
                int ii = null_row, j = w.x[ii];
                w.x[ii] = -1;
                while (ii != r0 && hops++ <= nr) {
                    ii = w.path[j];
                    w.y[j] = ii;
                    const int t = w.x[ii];
                    w.x[ii] = j;
                    j = t;
                }
            } else {
                int j = sink, ii;
                do {
                    ii = w.path[j];
                    w.y[j] = ii;
                    const int t = w.x[ii];
                    w.x[ii] = j;
                    j = t;
                } while (ii != r0 && hops++ <= nr);
            }
        }
        __syncwarp();
    }
    return bad;
}

}  // namespace bm
