// Stand-alone kernels behind the unit-testable C-ABI entry points (function-level
// parity tests call the same device code the fused engines use).
#include "../../include/boxmot_b200.h"
#include "iou.cuh"
#include "kf_xysr.cuh"
#include "kf_xyah.cuh"
#include "cosine_tc.cuh"
#include "strongsort_units.cuh"
#include <algorithm>
#include "lapjv.cuh"
#include "lap_rect.cuh"
#include <cstdio>
#include <limits>
#include <mutex>

namespace bm {

// One warp per problem, LAP_WPB problems per CTA.  The cost block is read in place
// from global memory (L2-resident at these sizes).
constexpr int LAP_WPB = 4;

template <int RECT>
__global__ void __launch_bounds__(LAP_WPB * 32)
lap_batch_kernel(const double* __restrict__ cost, int n_problems, int nr_max, int nc_max,
                 const int* __restrict__ nr_, const int* __restrict__ nc_, double cost_limit,
                 int limited, int* __restrict__ x_out, int* __restrict__ y_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int wid = warp_id(), lane = lane_id();
    const int b = blockIdx.x * LAP_WPB + wid;
    if (b >= n_problems) return;
    const int nr = nr_[b], nc = nc_[b];
    int* x = x_out + (size_t)b * nr_max;
    int* y = y_out + (size_t)b * nc_max;
    if (nr <= 0 || nc <= 0) {
        for (int i = lane; i < nr_max; i += 32) x[i] = -1;
        for (int j = lane; j < nc_max; j += 32) y[j] = -1;
        return;
    }
    const double* c = cost + (size_t)b * nr_max * nc_max;
    double pad;
    if (limited) pad = cost_limit / 2.0;
    else {
        double mx = -CUDART_INF;
        for (int e = lane; e < nr * nc; e += 32) mx = fmax(mx, c[(size_t)(e / nc) * nc_max + (e % nc)]);
        for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, shfl_xor_d(mx, o));
        pad = mx + 1.0;
    }
    const size_t wbytes = (lap_work_bytes(nr_max + nc_max) + 15) / 16 * 16;
    if (RECT) {   // fast exact solver on the real rows (csrc/lap_rect.cuh)
        RectWork rw = rect_carve(smem_raw + (size_t)wid * wbytes, nr, nc);
        rect_solve_warp(DenseCost{c, nc_max}, nr, nc, 2.0 * pad, rw, nullptr, nullptr);
        for (int i = lane; i < nr_max; i += 32) x[i] = i < nr ? rw.x[i] : -1;
        for (int j = lane; j < nc_max; j += 32) y[j] = j < nc ? rw.y[j] : -1;
        return;
    }
    LapWork w = lap_carve(smem_raw + (size_t)wid * wbytes, nr + nc);
    LapMat m = {c, nc_max, nr, nc, nr + nc, 1, pad};
    lap_solve_warp(m, w);
    for (int i = lane; i < nr_max; i += 32) x[i] = (i < nr && w.x[i] < nc) ? w.x[i] : -1;
    for (int j = lane; j < nc_max; j += 32) y[j] = (j < nc && w.y[j] < nr) ? w.y[j] : -1;
}

__global__ void iou_matrix_kernel(int func, const double* __restrict__ a, int na,
                                  const double* __restrict__ b, int nb, double diag,
                                  double* __restrict__ out) {
    const size_t total = (size_t)na * nb;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total;
         e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / nb), j = (int)(e % nb);
        Box A = {a[i * 4 + 0], a[i * 4 + 1], a[i * 4 + 2], a[i * 4 + 3]};
        Box B = {b[j * 4 + 0], b[j * 4 + 1], b[j * 4 + 2], b[j * 4 + 3]};
        out[e] = asso_pair(func, A, B, diag);
    }
}

__global__ void xysr_predict_kernel(double* x, double* P, int n, XysrConst kc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double xs[7], Ps[49];
    for (int c = 0; c < 7; c++) xs[c] = x[(size_t)i * 7 + c];
    for (int c = 0; c < 49; c++) Ps[c] = P[(size_t)i * 49 + c];
    xysr_predict(xs, Ps, kc);
    for (int c = 0; c < 7; c++) x[(size_t)i * 7 + c] = xs[c];
    for (int c = 0; c < 49; c++) P[(size_t)i * 49 + c] = Ps[c];
}

__global__ void xysr_update_kernel(double* x, double* P, const double* z, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double xs[7], Ps[49], zs[4];
    for (int c = 0; c < 7; c++) xs[c] = x[(size_t)i * 7 + c];
    for (int c = 0; c < 49; c++) Ps[c] = P[(size_t)i * 49 + c];
    for (int c = 0; c < 4; c++) zs[c] = z[(size_t)i * 4 + c];
    xysr_update(xs, Ps, zs);
    for (int c = 0; c < 7; c++) x[(size_t)i * 7 + c] = xs[c];
    for (int c = 0; c < 49; c++) P[(size_t)i * 49 + c] = Ps[c];
}

__global__ void xysrb_predict_kernel(double* x, double* Pb, int n, XysrConst kc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double xs[7], Ps[13];
    for (int c = 0; c < 7; c++) xs[c] = x[(size_t)i * 7 + c];
    for (int c = 0; c < 13; c++) Ps[c] = Pb[(size_t)i * 13 + c];
    xysrb_predict(xs, Ps, kc);
    for (int c = 0; c < 7; c++) x[(size_t)i * 7 + c] = xs[c];
    for (int c = 0; c < 13; c++) Pb[(size_t)i * 13 + c] = Ps[c];
}

__global__ void xysrb_update_kernel(double* x, double* Pb, const double* z, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double xs[7], Ps[13], zs[4];
    for (int c = 0; c < 7; c++) xs[c] = x[(size_t)i * 7 + c];
    for (int c = 0; c < 13; c++) Ps[c] = Pb[(size_t)i * 13 + c];
    for (int c = 0; c < 4; c++) zs[c] = z[(size_t)i * 4 + c];
    xysrb_update(xs, Ps, zs);
    for (int c = 0; c < 7; c++) x[(size_t)i * 7 + c] = xs[c];
    for (int c = 0; c < 13; c++) Pb[(size_t)i * 13 + c] = Ps[c];
}


// ---- constant-velocity box filters (XYAH / XYWH), dense [n][8][8] covariance at the ABI
__device__ __forceinline__ void cv8_load(const double* mean, const double* cov, int i, double* x, double* P) {
    for (int c = 0; c < 8; c++) x[c] = mean[(size_t)i * 8 + c];
    const double* C = cov + (size_t)i * 64;
    for (int c = 0; c < 4; c++) {
        P[c] = C[c * 8 + c]; P[4 + c] = C[c * 8 + 4 + c]; P[8 + c] = C[(4 + c) * 8 + c]; P[12 + c] = C[(4 + c) * 8 + 4 + c];
    }
}
__device__ __forceinline__ void cv8_store(double* mean, double* cov, int i, const double* x, const double* P) {
    for (int c = 0; c < 8; c++) mean[(size_t)i * 8 + c] = x[c];
    xyah_expand(P, cov + (size_t)i * 64);
}

// op: 0 initiate (z -> mean, cov), 1 predict, 2 update (z, conf), 3 multi_predict
template <int KIND>
__global__ void cv8_kernel(int op, double* mean, double* cov, const double* z, const double* conf, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double x[8], P[XA_P], zs[4];
    if (z) for (int c = 0; c < 4; c++) zs[c] = z[(size_t)i * 4 + c];
    if (op == 0) {
        cv8_initiate<KIND>(zs, x, P);
    } else {
        cv8_load(mean, cov, i, x, P);
        if (op == 1) cv8_predict<KIND>(x, P);
        else if (op == 3) cv8_predict<KIND, true>(x, P);
        else cv8_update<KIND>(x, P, zs, conf ? conf[i] : 0.0);
    }
    cv8_store(mean, cov, i, x, P);
}

// out[t][d] = gating_distance(track t, measurement d)
template <int KIND>
__global__ void cv8_gating_kernel(const double* mean, const double* cov, int nt, const double* meas, int nd,
                                  int only_position, int gaussian, double* out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nt * nd) return;
    const int t = e / nd, d = e % nd;
    double x[8], P[XA_P], zs[4];
    cv8_load(mean, cov, t, x, P);
    for (int c = 0; c < 4; c++) zs[c] = meas[(size_t)d * 4 + c];
    out[e] = cv8_gating<KIND>(x, P, zs, only_position != 0, gaussian != 0);
}


// ---- StrongSORT cost post-processing (strongsort/sort/linear_assignment.py, tracker.py), one thread per
// (track, detection) entry; the per-entry arithmetic lives in strongsort_units.cuh (shared with the engine).
__global__ void ss_gated_metric_kernel(double* __restrict__ cost, const double* __restrict__ mean,
                                       const double* __restrict__ cov, const double* __restrict__ trk, int nt,
                                       const double* __restrict__ meas, int nd, double mc_lambda, double id_weight) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nt * nd) return;
    const int t = e / nd, d = e % nd;
    double x[8], P[XA_P], z[4];
    cv8_load(mean, cov, t, x, P);
    for (int c = 0; c < 4; c++) z[c] = meas[(size_t)d * 4 + c];
    cost[e] = ss_gated_entry(cost[e], x, P, trk + (size_t)t * SS_FIELDS, z, mc_lambda, id_weight);
}

// _enhance_cost_matrix (linear_assignment.py:251-273); det [nd][3] = quality_score, conf, cls
__global__ void ss_enhance_cost_kernel(double* __restrict__ cost, const double* __restrict__ trk, int nt,
                                       const double* __restrict__ det, int nd) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nt * nd) return;
    const double* b = det + (size_t)(e % nd) * 3;
    cost[e] = ss_enhance_entry(cost[e], trk + (size_t)(e / nd) * SS_FIELDS, b[0], b[1], b[2]);
}

// iou_cost (iou_matching.py:49-87): rows of tracks with time_since_update > 1 are INFTY_COST
__global__ void ss_iou_cost_kernel(const double* __restrict__ mean, const double* __restrict__ trk, int nt,
                                   const double* __restrict__ tlwh, int nd, double* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nt * nd) return;
    const int t = e / nd, d = e % nd;
    if (trk[(size_t)t * SS_FIELDS + SS_TSU] > 1.0) { out[e] = SS_INFTY_COST; return; }
    out[e] = ss_iou_entry(mean + (size_t)t * 8, tlwh + (size_t)d * 4);
}


// Feature EMA of StrongSORT's Track.update (strongsort/sort/track.py:238-268, adaptive alpha :319-335),
// one warp per updated track: fp64 arithmetic on fp32 storage.
__device__ __forceinline__ double warp_sum_d(double v) {
    for (int o = 16; o > 0; o >>= 1) v += shfl_xor_d(v, o);
    return v;
}
__global__ void ss_feature_update_kernel(const float* __restrict__ last, const int* __restrict__ has_last,
                                         const float* __restrict__ feat, const double* __restrict__ conf, int n, int E,
                                         double base_alpha, double* __restrict__ ac, double* __restrict__ sim_out,
                                         float* __restrict__ out) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n) return;
    const float* f = feat + (size_t)w * E;
    const float* l = last + (size_t)w * E;
    float* o = out + (size_t)w * E;
    double f2 = 0.0;
    for (int i = lane; i < E; i += 32) f2 += (double)f[i] * (double)f[i];
    const double fn = sqrt(warp_sum_d(f2)) + 1e-8;                       // :240
    if (!has_last[w]) {                                                  // :261-263 first feature
        for (int i = lane; i < E; i += 32) o[i] = (float)((double)f[i] / fn);
        if (lane == 0 && sim_out) sim_out[w] = CUDART_NAN;
        return;
    }
    double dot = 0.0, n2 = 0.0, l2 = 0.0;
    for (int i = lane; i < E; i += 32) {
        const double a = (double)f[i] / fn, b = (double)l[i];
        dot += a * b; n2 += a * a; l2 += b * b;
    }
    dot = warp_sum_d(dot); n2 = warp_sum_d(n2); l2 = warp_sum_d(l2);
    const double sim = dot / (sqrt(n2) * sqrt(l2) + 1e-8);               // :245-247
    const double c = conf[w];
    const double cf = c > 0.7 ? 1.0 : (c > 0.3 ? 0.5 : 0.2);             // :324
    const double af = sim > 0.7 ? 1.0 : (sim > 0.4 ? 0.7 : 0.4);         // :327
    const double alpha = fmin(fmax(base_alpha * cf * af, 0.1), 0.95);    // :330-333
    double s2 = 0.0;
    for (int i = lane; i < E; i += 32) {
        const double v = alpha * (double)l[i] + (1.0 - alpha) * ((double)f[i] / fn);   // :254-256
        s2 += v * v;
    }
    const double sn = sqrt(warp_sum_d(s2)) + 1e-8;                       // :257
    for (int i = lane; i < E; i += 32)
        o[i] = (float)((alpha * (double)l[i] + (1.0 - alpha) * ((double)f[i] / fn)) / sn);
    if (lane == 0) {
        ac[w] = 0.9 * ac[w] + 0.1 * sim;                                 // :248
        if (sim_out) sim_out[w] = sim;
    }
}


// StrongSORT Track.predict / update / mark_missed / camera_update (strongsort/sort/track.py:173-317) on a
// flat per-track table, one thread per track (strongsort_units.cuh).
__global__ void ss_track_kernel(int op, double* __restrict__ mean, double* __restrict__ cov, double* __restrict__ trk,
                                const double* __restrict__ det, int n, int n_init, int max_age) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double* a = trk + (size_t)i * ST_FIELDS;
    if (op == 2) { ss_track_mark_missed(a, max_age); return; }
    if (op == 3) { ss_track_camera(mean + (size_t)i * 8, a, det + (size_t)i * 7); return; }
    double x[8], P[XA_P];
    cv8_load(mean, cov, i, x, P);
    if (op == 0) ss_track_predict(x, P, a);
    else ss_track_update(x, P, a, det + (size_t)i * 7, n_init);
    cv8_store(mean, cov, i, x, P);
}


// ---- embedding distance (matching.py:230-321) as a stand-alone batched contraction
__global__ void cosine_prep_kernel(const float* a, int na, const float* b, int nb, int E, int* idx, float* norms) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= na + nb) return;
    const float* src = w < na ? a + (size_t)w * E : b + (size_t)(w - na) * E;
    float ss = 0.f;
    for (int e = lane; e < E; e += 32) ss += src[e] * src[e];
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(BM_FULL, ss, o);
    if (lane == 0) { norms[w] = sqrtf(ss); idx[w] = w < na ? w : w - na; }
}

struct UnitCosine {
    CosineProblem p;
    __device__ CosineProblem operator()(int) const { return p; }
};

// ---- StrongSORT appearance metric (strongsort/sort/linear_assignment.py:382-413, :468-497,
// :595-618): cost[t][d] = min over the track's stored samples of 1 - clip(<f / (|f| + 1e-8),
// g / (|g| + 1e-8)>, -1, 1); INFTY_COST = 1e5 for a track without samples.
__global__ void nn_prep_kernel(const float* bank, const int* nb, int T, int Bcap, const float* feats, int D, int E,
                               int* idx, float* norms, float* out) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nbank = T * Bcap;
    if (w < nbank + D) {
        bool live = true;
        if (w < nbank) live = (w % Bcap) < nb[w / Bcap];
        if (live) {
            const float* src = w < nbank ? bank + (size_t)w * E : feats + (size_t)(w - nbank) * E;
            float ss = 0.f;
            for (int e = lane; e < E; e += 32) ss += src[e] * src[e];
            for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(BM_FULL, ss, o);
            if (lane == 0) norms[w] = sqrtf(ss) + 1e-8f;
        }
    }
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int nidx = Bcap > D ? Bcap : D;
    if (g < nidx) idx[g] = g;
    if (g < T * D && nb[g / D] <= 0) out[g] = 1e5f;
}

struct NnCosine {
    const float *bank, *feats, *norms;
    const int *nb, *idx;
    int T, Bcap, D, E;
    float* out;
    __device__ CosineProblem operator()(int t) const {
        CosineProblem p;
        p.A = feats; p.ra = idx; p.na = norms + (size_t)T * Bcap; p.nA = D;
        p.B = bank + (size_t)t * Bcap * E; p.rb = idx; p.nb = norms + (size_t)t * Bcap; p.nB = nb[t];
        p.E = E; p.out = out + (size_t)t * D; p.out_ld = 0; p.raw = 2;
        return p;
    }
};

// ---- StrongSORT min_cost_matching (strongsort/sort/linear_assignment.py:14-87) on a given cost
// matrix: entries > max_distance become max_distance + 1e-5 (:67), rectangular linear sum
// assignment (:70, scipy), pairs whose cost is still > max_distance are rejected (:82).
struct ClampCost {
    static constexpr bool kBounded = false, kStaged = false;
    const double* c; int ld; double maxd; int transposed;
    __device__ __forceinline__ double operator()(int i, int j) const {
        const double v = transposed ? c[(size_t)j * ld + i] : c[(size_t)i * ld + j];
        return v > maxd ? maxd + 1e-5 : v;
    }
    __device__ __forceinline__ double lower(int i, int j, bool* exact) const { *exact = true; return (*this)(i, j); }
};

__global__ void __launch_bounds__(LAP_WPB * 32)
min_cost_matching_kernel(const double* __restrict__ cost, int n_problems, int nr_max, int nc_max,
                         const int* __restrict__ nr_, const int* __restrict__ nc_, double max_distance,
                         int* __restrict__ x_out, int* __restrict__ y_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int wid = warp_id(), lane = lane_id();
    const int b = blockIdx.x * LAP_WPB + wid;
    if (b >= n_problems) return;
    const int nr = nr_[b], nc = nc_[b];
    int* x = x_out + (size_t)b * nr_max;
    int* y = y_out + (size_t)b * nc_max;
    for (int i = lane; i < nr_max; i += 32) x[i] = -1;
    for (int j = lane; j < nc_max; j += 32) y[j] = -1;
    if (nr <= 0 || nc <= 0) return;
    __syncwarp();
    const double* c = cost + (size_t)b * nr_max * nc_max;
    const size_t wbytes = (rect_work_bytes(nr_max > nc_max ? nr_max : nc_max, nr_max > nc_max ? nr_max : nc_max) + 15) / 16 * 16;
    const bool tr = nr > nc;                 // every row of the smaller side is assigned
    const int n1 = tr ? nc : nr, n2 = tr ? nr : nc;
    RectWork rw = rect_carve(smem_raw + (size_t)wid * wbytes, n1, n2);
    const ClampCost cc = {c, nc_max, max_distance, tr ? 1 : 0};
    rect_solve_warp(cc, n1, n2, CUDART_INF, rw, nullptr, nullptr);
    for (int i = lane; i < n1; i += 32) {
        const int j = rw.x[i];
        if (j < 0 || cc(i, j) > max_distance) continue;
        const int row = tr ? j : i, col = tr ? i : j;
        x[row] = col;
        y[col] = row;
    }
}

// reference form: one thread per output, fp32 FFMA dot product
__global__ void cosine_ffma_kernel(CosineProblem p) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.nA * p.nB) return;
    const int i = e / p.nB, j = e % p.nB;
    const float* x = p.A + (size_t)p.ra[i] * p.E;
    const float* y = p.B + (size_t)p.rb[j] * p.E;
    float acc = 0.f;
    for (int k = 0; k < p.E; k++) acc = fmaf(x[k], y[k], acc);
    const double cs = (double)acc / ((double)p.na[i] * (double)p.nb[j]);
    p.out[(size_t)p.ra[i] * p.out_ld + j] = (float)fmax(0.0, 1.0 - cs);
}

}  // namespace bm

static int unit_fail(const char* what, cudaError_t e) {
    fprintf(stderr, "boxmot_b200: %s: %s\n", what, cudaGetErrorString(e));
    return -2;
}

extern "C" {

static int lap_batch_launch(int rect, const double* d_cost, int n_problems, int nr_max, int nc_max,
                            const int* d_nr, const int* d_nc, double cost_limit, int* d_x, int* d_y,
                            void* stream) {
    if (n_problems <= 0) return 0;
    const int limited = cost_limit < std::numeric_limits<double>::max();
    const size_t wbytes = (bm::lap_work_bytes(nr_max + nc_max) + 15) / 16 * 16;
    const size_t smem = wbytes * bm::LAP_WPB;
    auto kern = rect ? bm::lap_batch_kernel<1> : bm::lap_batch_kernel<0>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return unit_fail("lap smem", e);
    const int grid = (n_problems + bm::LAP_WPB - 1) / bm::LAP_WPB;
    kern<<<grid, bm::LAP_WPB * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        d_cost, n_problems, nr_max, nc_max, d_nr, d_nc, cost_limit, limited, d_x, d_y);
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("lap_batch_kernel", e);
}

int bm_lap_solve_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                       const int* d_nc, double cost_limit, int* d_x, int* d_y, void* stream) {
    return lap_batch_launch(0, d_cost, n_problems, nr_max, nc_max, d_nr, d_nc, cost_limit, d_x, d_y, stream);
}

int bm_lap_rect_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                      const int* d_nc, double cost_limit, int* d_x, int* d_y, void* stream) {
    return lap_batch_launch(1, d_cost, n_problems, nr_max, nc_max, d_nr, d_nc, cost_limit, d_x, d_y, stream);
}

int bm_iou_matrix(int asso_func, const double* d_a, int na, const double* d_b, int nb, double img_w,
                  double img_h, double* d_out, void* stream) {
    if (na <= 0 || nb <= 0) return 0;
    const size_t total = (size_t)na * nb;
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    bm::iou_matrix_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        asso_func, d_a, na, d_b, nb, sqrt(img_w * img_w + img_h * img_h), d_out);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("iou_matrix_kernel", e);
}

int bm_xysr_predict(double* d_x, double* d_P, int n, double q_xy, double q_s, void* stream) {
    if (n <= 0) return 0;
    bm::XysrConst kc = {q_xy, q_s};
    bm::xysr_predict_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_x, d_P, n, kc);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("xysr_predict_kernel", e);
}

int bm_xysr_update(double* d_x, double* d_P, const double* d_z, int n, void* stream) {
    if (n <= 0) return 0;
    bm::xysr_update_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_x, d_P, d_z, n);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("xysr_update_kernel", e);
}

int bm_xysrb_predict(double* d_x, double* d_Pb, int n, double q_xy, double q_s, void* stream) {
    if (n <= 0) return 0;
    bm::XysrConst kc = {q_xy, q_s};
    bm::xysrb_predict_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_x, d_Pb, n, kc);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("xysrb_predict_kernel", e);
}

int bm_xysrb_update(double* d_x, double* d_Pb, const double* d_z, int n, void* stream) {
    if (n <= 0) return 0;
    bm::xysrb_update_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_x, d_Pb, d_z, n);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("xysrb_update_kernel", e);
}

int bm_kf_cv8(int kind, int op, double* d_mean, double* d_cov, const double* d_z, const double* d_conf, int n,
              void* stream) {
    if (n <= 0) return 0;
    if ((kind != 0 && kind != 1) || op < 0 || op > 3 || !d_mean || !d_cov || ((op == 0 || op == 2) && !d_z)) return -1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (kind == 0) bm::cv8_kernel<bm::CV8_XYAH><<<(n + 127) / 128, 128, 0, st>>>(op, d_mean, d_cov, d_z, d_conf, n);
    else bm::cv8_kernel<bm::CV8_XYWH><<<(n + 127) / 128, 128, 0, st>>>(op, d_mean, d_cov, d_z, d_conf, n);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("cv8_kernel", e);
}

int bm_kf_cv8_gating(int kind, const double* d_mean, const double* d_cov, int n_tracks, const double* d_meas,
                     int n_meas, int only_position, int gaussian, double* d_out, void* stream) {
    if (n_tracks <= 0 || n_meas <= 0) return 0;
    if ((kind != 0 && kind != 1) || !d_mean || !d_cov || !d_meas || !d_out) return -1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int n = n_tracks * n_meas;
    if (kind == 0)
        bm::cv8_gating_kernel<bm::CV8_XYAH><<<(n + 127) / 128, 128, 0, st>>>(d_mean, d_cov, n_tracks, d_meas, n_meas,
                                                                            only_position, gaussian, d_out);
    else
        bm::cv8_gating_kernel<bm::CV8_XYWH><<<(n + 127) / 128, 128, 0, st>>>(d_mean, d_cov, n_tracks, d_meas, n_meas,
                                                                            only_position, gaussian, d_out);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("cv8_gating_kernel", e);
}

int bm_strongsort_gated_metric(double* d_cost, const double* d_mean, const double* d_cov, const double* d_trk,
                               int n_tracks, const double* d_meas, int n_dets, double mc_lambda, double id_weight,
                               void* stream) {
    const int n = n_tracks * n_dets;
    if (n <= 0) return 0;
    if (!d_cost || !d_mean || !d_cov || !d_trk || !d_meas) return -1;
    bm::ss_gated_metric_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        d_cost, d_mean, d_cov, d_trk, n_tracks, d_meas, n_dets, mc_lambda, id_weight);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("ss_gated_metric_kernel", e);
}

int bm_strongsort_enhance_cost(double* d_cost, const double* d_trk, int n_tracks, const double* d_det, int n_dets,
                               void* stream) {
    const int n = n_tracks * n_dets;
    if (n <= 0) return 0;
    if (!d_cost || !d_trk || !d_det) return -1;
    bm::ss_enhance_cost_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_cost, d_trk, n_tracks,
                                                                                              d_det, n_dets);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("ss_enhance_cost_kernel", e);
}

int bm_strongsort_iou_cost(const double* d_mean, const double* d_trk, int n_tracks, const double* d_tlwh, int n_dets,
                           double* d_out, void* stream) {
    const int n = n_tracks * n_dets;
    if (n <= 0) return 0;
    if (!d_mean || !d_trk || !d_tlwh || !d_out) return -1;
    bm::ss_iou_cost_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_mean, d_trk, n_tracks,
                                                                                          d_tlwh, n_dets, d_out);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("ss_iou_cost_kernel", e);
}

int bm_strongsort_feature_update(const float* d_last, const int* d_has_last, const float* d_feat, const double* d_conf,
                                 int n, int emb_dim, double ema_alpha, double* d_appearance_consistency,
                                 double* d_similarity, float* d_out, void* stream) {
    if (n <= 0) return 0;
    if (!d_last || !d_has_last || !d_feat || !d_conf || !d_appearance_consistency || !d_out || emb_dim <= 0) return -1;
    bm::ss_feature_update_kernel<<<(n + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        d_last, d_has_last, d_feat, d_conf, n, emb_dim, ema_alpha, d_appearance_consistency, d_similarity, d_out);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("ss_feature_update_kernel", e);
}

int bm_strongsort_track(int op, double* d_mean, double* d_cov, double* d_trk, const double* d_det, int n, int n_init,
                        int max_age, void* stream) {
    if (n <= 0) return 0;
    if (op < 0 || op > 3 || !d_trk || (op != 2 && !d_mean) || (op < 2 && !d_cov) || ((op == 1 || op == 3) && !d_det)) return -1;
    bm::ss_track_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(op, d_mean, d_cov, d_trk, d_det, n,
                                                                                       n_init, max_age);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("ss_track_kernel", e);
}

int bm_nn_cosine_distance(const float* d_bank, const int* d_nb, int n_tracks, int bank_cap, const float* d_feats,
                           int n_feats, int emb_dim, float* d_out, void* d_work, void* stream) {
    if (n_tracks <= 0 || n_feats <= 0) return 0;
    if (!d_bank || !d_nb || !d_feats || !d_out || !d_work || bank_cap <= 0 || emb_dim <= 0) return -1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nidx = bank_cap > n_feats ? bank_cap : n_feats;
    int* idx = static_cast<int*>(d_work);
    float* norms = reinterpret_cast<float*>(idx + nidx);
    const long long warps = (long long)n_tracks * bank_cap + n_feats;
    const long long thr = std::max<long long>(warps * 32, std::max<long long>((long long)n_tracks * n_feats, nidx));
    bm::nn_prep_kernel<<<(unsigned)((thr + 127) / 128), 128, 0, st>>>(d_bank, d_nb, n_tracks, bank_cap, d_feats, n_feats,
                                                                     emb_dim, idx, norms, d_out);
    cudaError_t e = cudaFuncSetAttribute(bm::cosine_tc_kernel<bm::NnCosine>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)bm::cosine_tc_smem_bytes());
    if (e != cudaSuccess) return unit_fail("cosine_tc_kernel<NnCosine> attribute", e);
    bm::NnCosine fn = {d_bank, d_feats, norms, d_nb, idx, n_tracks, bank_cap, n_feats, emb_dim, d_out};
    bm::cosine_tc_kernel<bm::NnCosine><<<dim3((n_feats + bm::TC_M - 1) / bm::TC_M, n_tracks), bm::TC_NT,
                                         bm::cosine_tc_smem_bytes(), st>>>(fn);
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("nn cosine kernel", e);
}

int bm_min_cost_matching_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                               const int* d_nc, double max_distance, int* d_x, int* d_y, void* stream) {
    if (n_problems <= 0) return 0;
    if (!d_cost || !d_nr || !d_nc || !d_x || !d_y || nr_max <= 0 || nc_max <= 0) return -1;
    const int nm = nr_max > nc_max ? nr_max : nc_max;
    const size_t wbytes = (bm::rect_work_bytes(nm, nm) + 15) / 16 * 16;
    const size_t smem = wbytes * bm::LAP_WPB;
    cudaError_t e = cudaFuncSetAttribute(bm::min_cost_matching_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return unit_fail("min_cost_matching smem", e);
    const int grid = (n_problems + bm::LAP_WPB - 1) / bm::LAP_WPB;
    bm::min_cost_matching_kernel<<<grid, bm::LAP_WPB * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        d_cost, n_problems, nr_max, nc_max, d_nr, d_nc, max_distance, d_x, d_y);
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("min_cost_matching_kernel", e);
}

int bm_cosine_distance(const float* d_a, int na, const float* d_b, int nb, int emb_dim, float* d_out, void* d_work,
                       int mode, void* stream) {
    if (na <= 0 || nb <= 0) return 0;
    if (!d_a || !d_b || !d_out || !d_work || emb_dim <= 0 || (mode != 0 && mode != 1)) return -1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int* idx = static_cast<int*>(d_work);
    float* norms = reinterpret_cast<float*>(idx + na + nb);
    bm::cosine_prep_kernel<<<(na + nb + 3) / 4, 128, 0, st>>>(d_a, na, d_b, nb, emb_dim, idx, norms);
    bm::CosineProblem p;
    p.A = d_a; p.ra = idx; p.na = norms; p.nA = na;
    p.B = d_b; p.rb = idx + na; p.nb = norms + na; p.nB = nb;
    p.E = emb_dim; p.out = d_out; p.out_ld = nb;
    if (mode == 1) {
        bm::cosine_ffma_kernel<<<(na * nb + 127) / 128, 128, 0, st>>>(p);
    } else {
        cudaError_t e = cudaFuncSetAttribute(bm::cosine_tc_kernel<bm::UnitCosine>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bm::cosine_tc_smem_bytes());
        if (e != cudaSuccess) return unit_fail("cosine_tc_kernel attribute", e);
        bm::cosine_tc_kernel<bm::UnitCosine><<<dim3((na + bm::TC_M - 1) / bm::TC_M, 1), bm::TC_NT,
                                               bm::cosine_tc_smem_bytes(), st>>>(bm::UnitCosine{p});
    }
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("cosine kernel", e);
}

}  // extern "C"

// ---- bm_pack_rows: the engines' padded output rows as one dense ragged buffer ---------------
// Stream s owns rows [offsets[s], offsets[s + 1]) of `rows`; both destinations may be mapped pinned
// HOST memory, in which case the kernel's stores are the device-to-host transfer (64 bytes per row
// that exists instead of a copy of the whole padded array).
namespace bm {
__global__ void __launch_bounds__(256)
pack_rows_kernel(const double* __restrict__ out, const int* __restrict__ nout, int S, int out_rows,
                 double* __restrict__ rows, int* __restrict__ offsets) {
    __shared__ int s_part[8];
    __shared__ int s_off;
    const int s = blockIdx.x, tid = threadIdx.x;
    int acc = 0;
    for (int i = tid; i < s; i += blockDim.x) acc += nout[i];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(BM_FULL, acc, o);
    if ((tid & 31) == 0) s_part[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        int t = 0;
        for (int i = 0; i < 8; i++) t += s_part[i];
        s_off = t;
        offsets[s] = t;
        if (s == S - 1) offsets[S] = t + nout[s];
    }
    __syncthreads();
    const int off = s_off, n = nout[s];
    const double2* src = reinterpret_cast<const double2*>(out + (size_t)s * out_rows * 8);
    double2* dst = reinterpret_cast<double2*>(rows + (size_t)off * 8);
    for (int i = tid; i < n * 4; i += blockDim.x) dst[i] = src[i];
}
}  // namespace bm

extern "C" int bm_pack_rows(const double* d_out, const int* d_nout, int n_streams, int out_rows, double* rows_dst,
                            int* offsets_dst, void* stream) {
    if (n_streams <= 0) return 0;
    if (!d_out || !d_nout || !rows_dst || !offsets_dst || out_rows <= 0) return -1;
    bm::pack_rows_kernel<<<n_streams, 256, 0, static_cast<cudaStream_t>(stream)>>>(d_out, d_nout, n_streams, out_rows,
                                                                                   rows_dst, offsets_dst);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("pack_rows_kernel", e);
}

// ---- bm_pack_mot_rows: convert_to_mot_format (boxmot/engine/utils.py:101-146, ndarray branch) for every
// stream of a frame, fused with the ragged packing of bm_pack_rows: 9 columns per row =
// frame, id, left, top, width, height (rounded half-to-even like np.round, as int32), 1, class, confidence.
namespace bm {
__global__ void __launch_bounds__(256)
pack_mot_rows_kernel(const double* __restrict__ out, const int* __restrict__ nout, int S, int out_rows, int frame_idx,
                     double* __restrict__ rows, int* __restrict__ offsets) {
    __shared__ int s_part[8];
    __shared__ int s_off;
    const int s = blockIdx.x, tid = threadIdx.x;
    int acc = 0;
    for (int i = tid; i < s; i += blockDim.x) acc += nout[i];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(BM_FULL, acc, o);
    if ((tid & 31) == 0) s_part[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        int t = 0;
        for (int i = 0; i < 8; i++) t += s_part[i];
        s_off = t;
        offsets[s] = t;
        if (s == S - 1) offsets[S] = t + nout[s];
    }
    __syncthreads();
    const int n = nout[s];
    for (int i = tid; i < n; i += blockDim.x) {
        const double* r = out + ((size_t)s * out_rows + i) * 8;
        double* d = rows + (size_t)(s_off + i) * 9;
        const double w = r[2] - r[0], h = r[3] - r[1];          // ops.xyxy2ltwh
        d[0] = (double)frame_idx;
        d[1] = (double)(int)r[4];
        d[2] = (double)(int)rint(r[0]); d[3] = (double)(int)rint(r[1]);
        d[4] = (double)(int)rint(w); d[5] = (double)(int)rint(h);
        d[6] = 1.0;
        d[7] = (double)(int)r[6];
        d[8] = r[5];
    }
}
}  // namespace bm

extern "C" int bm_pack_mot_rows(const double* d_out, const int* d_nout, int n_streams, int out_rows, int frame_idx,
                                double* rows_dst, int* offsets_dst, void* stream) {
    if (n_streams <= 0) return 0;
    if (!d_out || !d_nout || !rows_dst || !offsets_dst || out_rows <= 0) return -1;
    bm::pack_mot_rows_kernel<<<n_streams, 256, 0, static_cast<cudaStream_t>(stream)>>>(d_out, d_nout, n_streams, out_rows,
                                                                                       frame_idx, rows_dst, offsets_dst);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : unit_fail("pack_mot_rows_kernel", e);
}

// ---- bm_cosine_distance_noalloc: bm_cosine_distance for callers that cannot allocate -------
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The row-index / norm
// scratch that bm_cosine_distance takes from the caller lives in ONE file-scope device array;
// every GPU that loads the library holds these 128 KiB until the process exits.  In HOST code
// the array's name is only a stand-in for the symbol, not a device address: the address handed
// to the kernels comes from cudaGetSymbolAddress, taken once per device with that device current.
#define BM_COSINE_NOALLOC_SCRATCH_BYTES (8 * BM_COSINE_NOALLOC_MAX_ROWS)
static __device__ __align__(256) unsigned char g_cosine_noalloc_scratch[BM_COSINE_NOALLOC_SCRATCH_BYTES];

namespace bm { void set_last_error(const char* msg); }

namespace {
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kNoallocMaxDevices] = {};   // device address of the scratch array, per device

int noalloc_fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    char msg[256];
    if (e != cudaSuccess) snprintf(msg, sizeof(msg), "bm_cosine_distance_noalloc: %s: %s", what, cudaGetErrorString(e));
    else snprintf(msg, sizeof(msg), "bm_cosine_distance_noalloc: %s", what);
    bm::set_last_error(msg);
    return code;
}

// Once per device, with that device current: resolve the symbol and opt the tensor-core kernel
// into its dynamic shared memory.  Neither call allocates.
int noalloc_setup(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(-2, "cudaGetDevice", e);
    if (dev < 0 || dev >= kNoallocMaxDevices) return noalloc_fail(-1, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_cosine_noalloc_scratch);
        if (e != cudaSuccess || !ptr) return noalloc_fail(-2, "cudaGetSymbolAddress(g_cosine_noalloc_scratch)", e);
        e = cudaFuncSetAttribute(bm::cosine_tc_kernel<bm::UnitCosine>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)bm::cosine_tc_smem_bytes());
        if (e != cudaSuccess) return noalloc_fail(-2, "cosine_tc_kernel attribute", e);
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}
}  // namespace

extern "C" int bm_cosine_distance_noalloc(const float* d_a, int na, const float* d_b, int nb, int emb_dim, float* d_out,
                                          int mode, void* stream) {
    if (na <= 0 || nb <= 0) return 0;
    if (!d_a || !d_b || !d_out || emb_dim <= 0 || (mode != 0 && mode != 1)) return noalloc_fail(-1, "bad argument");
    if ((long long)na + (long long)nb > BM_COSINE_NOALLOC_MAX_ROWS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "na + nb = %lld exceeds BM_COSINE_NOALLOC_MAX_ROWS = %d (static scratch of %d bytes)",
                 (long long)na + nb, BM_COSINE_NOALLOC_MAX_ROWS, BM_COSINE_NOALLOC_SCRATCH_BYTES);
        return noalloc_fail(-4, msg);
    }
    void* scratch = nullptr;
    const int rc = noalloc_setup(&scratch);
    if (rc != 0) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // Layout as bm_cosine_distance's d_work: (na + nb) row indices, then (na + nb) norms.  The
    // prep kernel rewrites every entry the contraction reads, so nothing of the previous call's
    // contents survives into this one and no clearing is needed.
    int* idx = static_cast<int*>(scratch);
    float* norms = reinterpret_cast<float*>(idx + na + nb);
    bm::cosine_prep_kernel<<<(na + nb + 3) / 4, 128, 0, st>>>(d_a, na, d_b, nb, emb_dim, idx, norms);
    bm::CosineProblem p;
    p.A = d_a; p.ra = idx; p.na = norms; p.nA = na;
    p.B = d_b; p.rb = idx + na; p.nb = norms + na; p.nB = nb;
    p.E = emb_dim; p.out = d_out; p.out_ld = nb;
    if (mode == 1)
        bm::cosine_ffma_kernel<<<(na * nb + 127) / 128, 128, 0, st>>>(p);
    else
        bm::cosine_tc_kernel<bm::UnitCosine><<<dim3((na + bm::TC_M - 1) / bm::TC_M, 1), bm::TC_NT,
                                               bm::cosine_tc_smem_bytes(), st>>>(bm::UnitCosine{p});
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : noalloc_fail(-2, "kernel launch", e);
}
