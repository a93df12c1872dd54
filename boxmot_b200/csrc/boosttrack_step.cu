// BoostTrack(++) frame step: five launches (six with ReID) advance every stream of the engine
// by one frame with no host synchronisation in between.  Mirrors BoostTrack.update
// (boxmot/trackers/boosttrack/boosttrack.py:221-336), its boosts (:372-456) and
// boosttrack/assoc.py:9-200.
//
//   B1 bt_predict_kernel   thread per live track     camera_update (:81-103), KalmanFilter.predict
//                          (HBM streaming)           (kalmanfilter.py:84-117), counters, get_confidence
//                                                    (:66-70), association operands by list position
//   B2 bt_boost_kernel     CTA per stream            DLO boost (:414-456: soft-BIoU, Mahalanobis column
//                          (all detections x         softmax, shape similarity), DUO boost (:372-412),
//                          tracks, on the fly)       confidence gate (:258-262)
//      cosine_tc_kernel    tensor cores              dets_embs @ tracker_embs.T (:272-277), raw dot products
//   B3 bt_assoc_kernel     CTA per stream            score matrix of assoc.py:156-200 (materialised: the
//                                                    column softmax needs whole columns), fast path or exact
//                                                    LAP (:106-114), IoU / embedding gate (:117-153)
//   B3u bt_update_kernel   thread per match          Kalman update of the matches (:72-79)
//   B4 bt_finalize_kernel  CTA per stream            births in
//                                                    unmatched order (:311-315), output rows in list order
//                                                    with the shape filter (:317-346), removal (:329)
//   B5 bt_emb_kernel       warp per pair             update_emb (:116-118) / raw copy for births
#include "boosttrack_engine.cuh"
#include "cosine_tc.cuh"
#include <cstdio>
#include <cstdlib>

namespace bm {

constexpr int B1_NT = 128;
// Threads of the per-stream CTAs (boost and association kernels): 256, or 512 when the per-stream
// tables leave room for only one CTA per SM (crowd caps) -- the same shared memory then feeds 16
// warps instead of 8.  The kernels stride by blockDim.x.
constexpr int BT_NT_MAX = 512;
#define BT_NT ((int)blockDim.x)
#define BT_NW ((int)blockDim.x >> 5)
constexpr int BM_ERR_LAPBAD = 8;

// ============================================================================ B1
__device__ __forceinline__ void bt_load(const double* base, int T, int slot, double* x, double* Pb) {
#pragma unroll
    for (int c = 0; c < 8; c++) x[c] = base[c * T + slot];
#pragma unroll
    for (int c = 0; c < 16; c++) Pb[c] = base[(8 + c) * T + slot];
}
__device__ __forceinline__ void bt_store(double* base, int T, int slot, const double* x, const double* Pb) {
#pragma unroll
    for (int c = 0; c < 8; c++) base[c * T + slot] = x[c];
#pragma unroll
    for (int c = 0; c < 16; c++) base[(8 + c) * T + slot] = Pb[c];
}

__global__ void __launch_bounds__(B1_NT)
bt_predict_kernel(BoostState st, BoostParams pr, const double* __restrict__ warps) {
    const int chunks = (st.Tcap + B1_NT - 1) / B1_NT;
    const int s = blockIdx.x / chunks;
    const int p = (blockIdx.x % chunks) * B1_NT + threadIdx.x;
    BoostView v(st, s);
    const int T = st.Tcap;
    const int nT = v.hdr[BTH_NLIST];
    if (p == 0) st.n_updates[s] += (unsigned long long)nT;
    if (p >= nT) return;
    const int slot = v.order[p];
    double x[8], Pb[16];
    bt_load(v.xP, T, slot, x, Pb);
    if (warps) {
        // camera_update: both corners through the 2x3 warp, state rebuilt from the warped box
        const double* w = warps + (size_t)s * 6;
        double b[4];
        bt_x_to_box(x, b);
        const double x1 = (w[0] * b[0] + w[1] * b[1]) + w[2] * 1.0, y1 = (w[3] * b[0] + w[4] * b[1]) + w[5] * 1.0;
        const double x2 = (w[0] * b[2] + w[1] * b[3]) + w[2] * 1.0, y2 = (w[3] * b[2] + w[4] * b[3]) + w[5] * 1.0;
        const double ww = x2 - x1, hh = y2 - y1;
        x[0] = x1 + ww / 2; x[1] = y1 + hh / 2; x[2] = hh; x[3] = ww / hh;
    }
    bt_kf_predict(x, Pb);
    bt_store(v.xP, T, slot, x, Pb);
    const int age = v.I(BTI_AGE, slot) + 1;
    v.I(BTI_AGE, slot) = age;
    const int tsu0 = v.I(BTI_TSU, slot);
    if (tsu0 > 0) v.I(BTI_STREAK, slot) = 0;
    const int tsu = tsu0 + 1;
    v.I(BTI_TSU, slot) = tsu;
    // get_confidence: 0.9^(7 - age) while age < 7, else 0.9^(time_since_update - 1)
    const int k = age < 7 ? 7 - age : tsu - 1;
    const double conf = st.powtab[min(k, st.npow - 1)];
    double* tp = st.tp + (size_t)s * TP_COUNT * T;
    double b[4];
    bt_x_to_box(x, b);
#pragma unroll
    for (int c = 0; c < 4; c++) {
        tp[(TP_X1 + c) * T + p] = b[c];
        tp[(TP_Z0 + c) * T + p] = x[c];
        tp[(TP_S0 + c) * T + p] = 1.0 / Pb[4 * c];      // np.reciprocal(diag(P[:4,:4]))
    }
    tp[TP_CONF * T + p] = conf;
    tp[TP_AGE1 * T + p] = (double)(tsu - 1);
}

// ============================================================================ shared helpers
struct BtSmem {
    double* tcol;        // [TP_COUNT][Tcap] per-column operands
    double* colsum;      // [Tcap]
    double* dz;          // [Dcap][4] detections as (x, y, h, r)
    double* dcf;         // [Dcap]
    double* rowa;        // [Dcap] row maxima / minima
    double* rowb;        // [Dcap]
    float4* dbox;        // [Dcap]
    int *di, *flag, *rarg, *rowcnt, *hit_t, *d2t, *um_d;   // [Dcap]
    int *colcnt, *t2d, *slot_of;                            // [Tcap]
    int* misc;           // [16]
    void* work;          // solver work area
    float* tf;           // [3][Tcap] fp32 screening table (centre x, reach for the boost pass, reach otherwise)
    unsigned short* cand;   // [BT_NW][Tcap] per-warp list of the columns that passed the screen
};

__host__ __device__ inline size_t bt_work_bytes(int Tcap, int Dcap) {
    const size_t b = rect_work_bytes(Dcap, Tcap), c = lap_work_bytes(2 * Dcap);
    return ((b > c ? b : c) + 15) / 16 * 16;
}
__host__ __device__ inline size_t bt_smem_bytes(int Tcap, int Dcap, int threads) {
    const size_t t = Tcap, d = Dcap;
    return 8 * (TP_COUNT * t + t + 4 * d + 3 * d) + 16 * d + 4 * (7 * d + 3 * t + 16) + 12 * t + 2 * (threads / 32) * t + bt_work_bytes(Tcap, Dcap) + 32;
}
inline int bt_threads(int Tcap, int Dcap) {
    const size_t one_per_sm = 113 * 1024, limit = 227 * 1024;
    static const int forced = [] { const char* e = getenv("BM_BT_THREADS"); return e ? atoi(e) : 0; }();
    if ((forced == 256 || forced == BT_NT_MAX) && bt_smem_bytes(Tcap, Dcap, forced) <= limit) return forced;
    return bt_smem_bytes(Tcap, Dcap, 256) > one_per_sm && bt_smem_bytes(Tcap, Dcap, BT_NT_MAX) <= limit ? BT_NT_MAX : 256;
}
__device__ inline BtSmem bt_carve(unsigned char* base, int Tcap, int Dcap) {
    BtSmem m;
    m.dbox = reinterpret_cast<float4*>(base);                       // 16-byte aligned first
    double* q = reinterpret_cast<double*>(base + 16 * (size_t)Dcap);
    m.tcol = q; q += (size_t)TP_COUNT * Tcap;
    m.colsum = q; q += Tcap;
    m.dz = q; q += 4 * (size_t)Dcap;
    m.dcf = q; q += Dcap;
    m.rowa = q; q += Dcap;
    m.rowb = q; q += Dcap;
    m.work = q; q = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(q) + bt_work_bytes(Tcap, Dcap));
    int* k = reinterpret_cast<int*>(q);
    m.di = k; k += Dcap; m.flag = k; k += Dcap; m.rarg = k; k += Dcap; m.rowcnt = k; k += Dcap;
    m.hit_t = k; k += Dcap; m.d2t = k; k += Dcap; m.um_d = k; k += Dcap;
    m.colcnt = k; k += Tcap; m.t2d = k; k += Tcap; m.slot_of = k; k += Tcap;
    m.misc = k; k += 16;
    m.tf = reinterpret_cast<float*>(k);
    m.cand = reinterpret_cast<unsigned short*>(m.tf + 3 * (size_t)Tcap);
    return m;
}

__device__ __forceinline__ Box bt_trk_box(const double* tcol, int T, int t) {
    return Box{tcol[TP_X1 * T + t], tcol[TP_Y1 * T + t], tcol[TP_X2 * T + t], tcol[TP_Y2 * T + t]};
}

// get_mh_dist_matrix (boosttrack.py:356-369): sum over (x, y, h, r) of (z - x)^2 / P_ii
__device__ __forceinline__ double bt_mh(const double* z, const double* tcol, int T, int t) {
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const double dlt = z[c] - tcol[(TP_Z0 + c) * T + t];
        const double q = (dlt * dlt) * tcol[(TP_S0 + c) * T + t];
        acc = c == 0 ? q : acc + q;
    }
    return acc;
}

// shape_similarity_v1 / v2 (assoc.py:15-34); v1 divides BOTH terms by the larger width
__device__ __forceinline__ double bt_shape(const Box& d, const Box& t, int corr) {
    const double dw = d.x2 - d.x1, dh = d.y2 - d.y1, tw = t.x2 - t.x1, th = t.y2 - t.y1;
    const double den2 = corr ? fmax(dh, th) : fmax(dw, tw);
    return exp(-(fabs(dw - tw) / fmax(dw, tw) + fabs(dh - th) / den2));
}

// soft_biou_batch (assoc.py:69-103): both boxes grown by (1 - track confidence) * k
__device__ __forceinline__ double bt_soft_biou(const Box& a, const Box& b, double conf) {
    const double k1 = 0.25, k2 = 0.5, g = 1 - conf;
    const double ax1 = a.x1 - (a.x2 - a.x1) * g * k1, bx1 = b.x1 - (b.x2 - b.x1) * g * k2;
    const double ay1 = a.y1 - (a.y2 - a.y1) * g * k1, by1 = b.y1 - (b.y2 - b.y1) * g * k2;
    const double ax2 = a.x2 + (a.x2 - a.x1) * g * k1, bx2 = b.x2 + (b.x2 - b.x1) * g * k2;
    const double ay2 = a.y2 + (a.y2 - a.y1) * g * k1, by2 = b.y2 + (b.y2 - b.y1) * g * k2;
    const double w = fmax(0.0, fmin(ax2, bx2) - fmax(ax1, bx1));
    const double h = fmax(0.0, fmin(ay2, by2) - fmax(ay1, by1));
    const double wh = w * h;
    const double den = (ax2 - ax1) * (ay2 - ay1) + (bx2 - bx1) * (by2 - by1) - wh;
    if (wh == 0.0 && den > 0.0) return 0.0;          // 0 / den, without the division
    return wh / den;
}

// iou_batch (assoc.py:50-66) with the same shortcut for disjoint boxes
__device__ __forceinline__ double bt_iou(const Box& a, const Box& b) {
    const double w = fmax(0.0, fmin(a.x2, b.x2) - fmax(a.x1, b.x1));
    const double h = fmax(0.0, fmin(a.y2, b.y2) - fmax(a.y1, b.y1));
    const double wh = w * h;
    const double den = (a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1) - wh;
    if (wh == 0.0 && den > 0.0) return 0.0;
    return wh / den;
}

// x^1.5 as x * sqrt(x) carried in double-double, i.e. correctly rounded except in rare
// half-way cases -- the same contract as the libm pow behind numpy's `max_s ** 1.5`.
__device__ __forceinline__ double bt_pow15(double x) {
    if (!(x > 0.0)) return x == 0.0 ? 0.0 : CUDART_NAN;
    const double s = sqrt(x);
    const double e = fma(-s, s, x) / (2.0 * s);          // sqrt(x) = s + e
    const double hi = x * s;
    const double lo = fma(x, s, -hi) + x * e;
    return hi + lo;
}

// ---------------------------------------------------------------------------------------
// fp32 screening.  Most detection x track pairs are far apart IN X: their boxes (even after
// soft-BIoU's growth) are disjoint and their Mahalanobis distance is beyond the 13.2767 limit
// (mh >= dx^2 / P00), so every term of the reference's matrices is a STRUCTURAL value for them
// (IoU = soft-BIoU = 0 exactly, Mahalanobis similarity = 0, exp(limit - min(mh, limit)) = 1).
// One float compare per pair -- |cx_d - cx_t| > reach_t + reach_d, with margins far above the fp32
// rounding error -- decides which pairs are surely in that state; everything else is evaluated
// exactly in fp64.  Degenerate / non-finite boxes are never screened out (reach = +inf).
//   tf[0][t] = centre x, tf[1][t] = max(2 half-widths, Mahalanobis radius)   (boost pass: the track
//   box grows by <= 0.5 of its width per side), tf[2][t] = max(half-width, Mahalanobis radius).
enum { TF_CX = 0, TF_RB, TF_RA, TF_COUNT };
__device__ inline void bt_stage_floats(const BtSmem& sm, int T, int nT) {
    for (int t = threadIdx.x; t < nT; t += BT_NT) {
        const double x1 = sm.tcol[TP_X1 * T + t], x2 = sm.tcol[TP_X2 * T + t], s0 = sm.tcol[TP_S0 * T + t];
        const double hgt = sm.tcol[TP_Y2 * T + t] - sm.tcol[TP_Y1 * T + t];
        const bool ok = x2 > x1 && (x2 - x1) < 1e30 && s0 > 0 && s0 < 1e30 && fabs(x1) < 1e30 && hgt > 0 && hgt < 1e30;
        const float hw = (float)((x2 - x1) * 0.5) * 1.0001f;
        const float rmh = sqrtf((float)BT_MH_LIMIT * 1.02f / ((float)s0 * 0.999f)) * 1.001f + 0.5f;
        sm.tf[TF_CX * T + t] = (float)((x1 + x2) * 0.5);
        sm.tf[TF_RB * T + t] = ok ? fmaxf(2.f * hw, rmh) : CUDART_INF_F;
        sm.tf[TF_RA * T + t] = ok ? fmaxf(hw, rmh) : CUDART_INF_F;
    }
}
// far in x: grow = 1.5 for the boost pass (the detection grows by <= 0.25 of its width per side), 1 otherwise
__device__ __forceinline__ bool bt_far_x(const float4 det, float grow, const float* tf, int which, int T, int t) {
    const float dcx = 0.5f * (det.x + det.z), dhw = 0.5f * (det.z - det.x) * 1.0001f;
    const float cx = tf[TF_CX * T + t];
    const float margin = 0.05f + 1e-5f * (fabsf(dcx) + fabsf(cx));
    return dhw > 0.f && det.w > det.y && fabsf(dcx - cx) > tf[which * T + t] + grow * dhw + margin;   // false for NaN / inf operands
}

// Stage the per-column operands of the nT listed tracks (and MhDist_similarity's column sums
// over the detections listed in sm.dz[0..n)) into shared memory.
__device__ inline void bt_stage_columns(const BoostState& st, int s, const BtSmem& sm, int nT) {
    const int T = st.Tcap;
    const double* tp = st.tp + (size_t)s * TP_COUNT * T;
    for (int i = threadIdx.x; i < TP_COUNT * nT; i += BT_NT) {
        const int f = i / nT, t = i % nT;
        sm.tcol[f * T + t] = tp[f * T + t];
    }
}
// np.exp(limit - min(mh, limit)).sum(0) per column: a thread owns a column and adds the rows in order
// (1.0 for a screened-out or masked pair: exp(0) == 1 exactly).
__device__ inline void bt_column_sums(const BtSmem& sm, int T, int n, int nT, double* /*scratch*/) {
    for (int t = threadIdx.x; t < nT; t += BT_NT) {
        double acc = 0.0;
        for (int d = 0; d < n; d++) {
            double e = 1.0;
            if (!bt_far_x(sm.dbox[d], 1.f, sm.tf, TF_RA, T, t)) {
                const double mh = bt_mh(sm.dz + 4 * d, sm.tcol, T, t);
                if (!(mh > BT_MH_LIMIT)) e = exp(BT_MH_LIMIT - mh);
            }
            acc = d == 0 ? e : acc + e;
        }
        sm.colsum[t] = acc;
    }
}
__device__ __forceinline__ double bt_mh_sim(double mh, double colsum) {
    if (mh > BT_MH_LIMIT) return 0.0;
    return exp(BT_MH_LIMIT - mh) / colsum;
}

// ============================================================================ B2
template <int MAXNT>
__global__ void __launch_bounds__(MAXNT)
bt_boost_kernel(BoostState st, BoostParams pr, const float* __restrict__ dets, const int* __restrict__ ndets) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int T = st.Tcap, Dcap = st.Dcap;
    BtSmem sm = bt_carve(smem_raw, T, Dcap);
    BoostView v(st, s);
    const int nT = v.hdr[BTH_NLIST];
    int nD = ndets[s];
    if (nD > Dcap) { if (tid == 0) atomicOr(&v.hdr[BTH_ERR], BM_ERR_DET_OVERFLOW); nD = Dcap; }
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    for (int d = tid; d < nD; d += BT_NT) {
        const float4 b = make_float4(dsrc[d * 6 + 0], dsrc[d * 6 + 1], dsrc[d * 6 + 2], dsrc[d * 6 + 3]);
        sm.dbox[d] = b;
        const double bb[4] = {(double)b.x, (double)b.y, (double)b.z, (double)b.w};
        bt_box_to_z(bb, sm.dz + 4 * d);
        sm.dcf[d] = (double)dsrc[d * 6 + 4];
        sm.rowa[d] = -CUDART_INF;      // max_s
        sm.rowb[d] = CUDART_INF;       // min Mahalanobis distance
        sm.flag[d] = 0;                // use_vt hit
    }
    bt_stage_columns(st, s, sm, nT);
    __syncthreads();
    bt_stage_floats(sm, T, nT);
    __syncthreads();
    const bool boosts = nT > 0 && nD > 0 && (pr.use_dlo || pr.use_duo);
    if (boosts) {
        const bool rich = pr.use_dlo && pr.rich_s;
        if (rich) { bt_column_sums(sm, T, nD, nT, st.cost + (size_t)s * st.cost_stride); __syncthreads(); }
        for (int d = wid; d < nD; d += BT_NW) {
            const float4 bf = sm.dbox[d];
            const Box a = {(double)bf.x, (double)bf.y, (double)bf.z, (double)bf.w};
            double mx = -CUDART_INF, mn = CUDART_INF;
            int hit = 0, deferred = 0;
            // Phase 1 evaluates the pairs that can reach a large S; a pair that is masked by the
            // Mahalanobis limit AND has disjoint grown boxes has S = shape / 3 <= 1/3 exactly (shape
            // similarity is exp of a non-positive number) and cannot pass the use_vt test (>= 0.8):
            // its exp is only evaluated (phase 2) when the row maximum is still below 1/3.
            unsigned short* cand = sm.cand + (size_t)wid * T;
            int nc = 0;
            for (int base = 0; base < nT; base += 32) {
                const int t = base + lane;
                const bool near = t < nT && !bt_far_x(bf, 1.5f, sm.tf, TF_RB, T, t);
                if (t < nT && !near) deferred = 1;
                const unsigned bal = __ballot_sync(BM_FULL, near);
                if (near) cand[nc + __popc(bal & ((1u << lane) - 1))] = (unsigned short)t;
                nc += __popc(bal);
            }
            __syncwarp();
            for (int i = lane; i < nc; i += 32) {
                const int t = cand[i];
                const double mh = bt_mh(sm.dz + 4 * d, sm.tcol, T, t);
                mn = fmin(mn, mh);
                if (!pr.use_dlo) continue;
                const Box b = bt_trk_box(sm.tcol, T, t);
                double S;
                if (rich) {
                    const double sb = bt_soft_biou(a, b, sm.tcol[TP_CONF * T + t]);
                    if (mh > BT_MH_LIMIT && sb == 0.0) { deferred = 1; continue; }
                    const double ms = bt_mh_sim(mh, sm.colsum[t]);
                    const double sh = bt_shape(a, b, pr.s_corr);
                    S = ((ms + sh) + sb) / 3;
                } else {
                    S = bt_iou(a, b);
                }
                mx = fmax(mx, S);
                if (S > fmax(0.95 - sm.tcol[TP_AGE1 * T + t], 0.8)) hit = 1;
            }
            __syncwarp();
            for (int o = 16; o > 0; o >>= 1) {
                mx = fmax(mx, shfl_xor_d(mx, o));
                mn = fmin(mn, shfl_xor_d(mn, o));
                hit |= __shfl_xor_sync(BM_FULL, hit, o);
                deferred |= __shfl_xor_sync(BM_FULL, deferred, o);
            }
            if (deferred && !(mx >= 1.0 / 3.0)) {
                for (int t = lane; t < nT; t += 32) {
                    const double mh = bt_mh(sm.dz + 4 * d, sm.tcol, T, t);
                    const Box b = bt_trk_box(sm.tcol, T, t);
                    if (!(mh > BT_MH_LIMIT && bt_soft_biou(a, b, sm.tcol[TP_CONF * T + t]) == 0.0)) continue;
                    mx = fmax(mx, ((0.0 + bt_shape(a, b, pr.s_corr)) + 0.0) / 3);
                }
                for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, shfl_xor_d(mx, o));
            }
            if (lane == 0) { sm.rowa[d] = mx; sm.rowb[d] = mn; sm.flag[d] = hit; }
        }
        __syncthreads();
        if (pr.use_dlo) {
            for (int d = tid; d < nD; d += BT_NT) {
                double c = sm.dcf[d];
                const double ms = sm.rowa[d];
                if (!pr.use_sb && !pr.use_vt) {
                    c = fmax(c, ms * pr.dlo_coef);
                } else {
                    if (pr.use_sb) c = fmax(c, 0.65 * c + (1 - 0.65) * bt_pow15(ms));
                    if (pr.use_vt && sm.flag[d]) c = fmax(c, pr.det_thresh + 1e-5);
                }
                sm.dcf[d] = c;
            }
            __syncthreads();
        }
        if (pr.use_duo) {
            // candidates: far from every track and still below the threshold; a candidate is
            // boosted unless an overlapping (IoU > 0.3) candidate has a larger confidence
            for (int d = tid; d < nD; d += BT_NT) sm.flag[d] = (sm.rowb[d] > BT_MH_LIMIT && sm.dcf[d] < pr.det_thresh) ? 1 : 0;
            __syncthreads();
            for (int d = tid; d < nD; d += BT_NT) {
                if (!sm.flag[d]) continue;
                const float4 bf = sm.dbox[d];
                const Box a = {(double)bf.x, (double)bf.y, (double)bf.z, (double)bf.w};
                const double c = sm.dcf[d];
                bool keep = true;
                for (int j = 0; j < nD && keep; j++) {
                    if (j == d || !sm.flag[j]) continue;
                    const float4 q = sm.dbox[j];
                    const Box b = {(double)q.x, (double)q.y, (double)q.z, (double)q.w};
                    if (iou_pair(a, b) > 0.3 && sm.dcf[j] > c) keep = false;
                }
                if (keep) sm.flag[d] = 2;
            }
            __syncthreads();
            for (int d = tid; d < nD; d += BT_NT) if (sm.flag[d] == 2) sm.dcf[d] = pr.det_thresh + 1e-4;
            __syncthreads();
        }
    }
    double* dconf = st.dconf + (size_t)s * Dcap;
    for (int d = tid; d < nD; d += BT_NT) dconf[d] = sm.dcf[d];
    if (wid == 0) {
        int* hs = st.hsel + (size_t)s * Dcap;
        const int n = warp_compact(nD, hs, 0, [&](int d) { return sm.dcf[d] >= pr.det_thresh; }, [&](int d) { return d; });
        if (lane == 0) v.hdr[BTH_ND] = n;
    }
}

// ============================================================================ B3
template <int MAXNT>
__global__ void __launch_bounds__(MAXNT)
bt_assoc_kernel(BoostState st, BoostParams pr, const float* __restrict__ dets) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int T = st.Tcap, Dcap = st.Dcap;
    BtSmem sm = bt_carve(smem_raw, T, Dcap);
    BoostView v(st, s);
    const int nT = v.hdr[BTH_NLIST], nD = v.hdr[BTH_ND];
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    const int* hs = st.hsel + (size_t)s * Dcap;
    const double* dconf = st.dconf + (size_t)s * Dcap;
    double* cost = st.cost + (size_t)s * st.cost_stride;
    const float* sim = st.embsim + (size_t)s * Dcap * T;
    int* upd_slot = st.upd + (size_t)s * 2 * Dcap;
    int* upd_det = upd_slot + Dcap;
    __shared__ double s_cmx[BT_NT_MAX / 32];
    if (tid < 16) sm.misc[tid] = 0;
    for (int d = tid; d < nD; d += BT_NT) {
        const int i = hs[d];
        const float4 b = make_float4(dsrc[i * 6 + 0], dsrc[i * 6 + 1], dsrc[i * 6 + 2], dsrc[i * 6 + 3]);
        sm.dbox[d] = b;
        const double bb[4] = {(double)b.x, (double)b.y, (double)b.z, (double)b.w};
        bt_box_to_z(bb, sm.dz + 4 * d);
        sm.dcf[d] = dconf[i];
        sm.di[d] = i;
        sm.d2t[d] = -1; sm.rowcnt[d] = 0; sm.hit_t[d] = -1;
    }
    for (int t = tid; t < nT; t += BT_NT) { sm.colcnt[t] = 0; sm.t2d[t] = -1; sm.slot_of[t] = v.order[t]; }
    bt_stage_columns(st, s, sm, nT);
    __syncthreads();
    bt_stage_floats(sm, T, nT);
    __syncthreads();
    if (nD > 0 && nT > 0) {
        bt_column_sums(sm, T, nD, nT, cost);
        __syncthreads();
        const double l_emb = (((1 + pr.l_iou) + pr.l_shape) + pr.l_mhd) * 1.5;
        double cmx = -CUDART_INF;
        for (int d = wid; d < nD; d += BT_NW) {
            const float4 bf = sm.dbox[d];
            const Box a = {(double)bf.x, (double)bf.y, (double)bf.z, (double)bf.w};
            const double dc = sm.dcf[d];
            const float* srow = sim + (size_t)sm.di[d] * T;
            double best = CUDART_INF;
            int barg = 0x7fffffff, hits = 0, hit = -1;
            unsigned short* cand = sm.cand + (size_t)wid * T;
            int nc = 0;
            auto account = [&](int t, double c) {
                cost[(size_t)d * nT + t] = -c;
                if (c > pr.iou_thr) { hits++; hit = t; atomicAdd(&sm.colcnt[t], 1); }
                if (-c < best || (-c == best && t < barg)) { best = -c; barg = t; }
                cmx = fmax(cmx, -c);
            };
            for (int base = 0; base < nT; base += 32) {
                const int t = base + lane;
                const bool near = t < nT && !(pr.iou_thr > 0 && bt_far_x(bf, 1.f, sm.tf, TF_RA, T, t));
                if (t < nT && !near) {
                    // IoU = 0, confidence term = 0 (IoU < thr), Mahalanobis similarity = 0, shape term x 0
                    double c = 0.0;
                    if (pr.with_reid) c = c + l_emb * (double)srow[t];
                    account(t, c);
                }
                const unsigned bal = __ballot_sync(BM_FULL, near);
                if (near) cand[nc + __popc(bal & ((1u << lane) - 1))] = (unsigned short)t;
                nc += __popc(bal);
            }
            __syncwarp();
            for (int i = lane; i < nc; i += 32) {
                const int t = cand[i];
                const Box b = bt_trk_box(sm.tcol, T, t);
                const double iou = bt_iou(a, b);
                double cf = dc * sm.tcol[TP_CONF * T + t];
                if (iou < pr.iou_thr) cf = 0.0;
                double c = iou;
                c = c + (pr.l_iou * cf) * iou;
                const double mh = bt_mh(sm.dz + 4 * d, sm.tcol, T, t);
                c = c + pr.l_mhd * bt_mh_sim(mh, sm.colsum[t]);
                // conf == 0 (IoU below the threshold): (lambda * 0) * shape == 0 for the finite shape term
                if (cf != 0.0) c = c + (pr.l_shape * cf) * bt_shape(a, b, pr.s_corr);
                if (pr.with_reid) c = c + l_emb * (double)srow[t];
                account(t, c);
            }
            __syncwarp();
            warp_argmin(best, barg);
            for (int o = 16; o > 0; o >>= 1) {
                hits += __shfl_xor_sync(BM_FULL, hits, o);
                hit = max(hit, __shfl_xor_sync(BM_FULL, hit, o));
            }
            if (lane == 0) { sm.rowa[d] = best; sm.rarg[d] = barg; sm.rowcnt[d] = hits; sm.hit_t[d] = hit; }
        }
        for (int o = 16; o > 0; o >>= 1) cmx = fmax(cmx, shfl_xor_d(cmx, o));
        if (lane == 0) s_cmx[wid] = cmx;
        __syncthreads();
        if (wid == 0) {
            int rmax = 0, cmax_ = 0;
            for (int d = lane; d < nD; d += 32) rmax = max(rmax, sm.rowcnt[d]);
            for (int t = lane; t < nT; t += 32) cmax_ = max(cmax_, sm.colcnt[t]);
            for (int o = 16; o > 0; o >>= 1) {
                rmax = max(rmax, __shfl_xor_sync(BM_FULL, rmax, o));
                cmax_ = max(cmax_, __shfl_xor_sync(BM_FULL, cmax_, o));
            }
            const bool fast = rmax == 1 && cmax_ == 1;       // match(): `a` is a partial permutation
            if (fast) {
                for (int d = lane; d < nD; d += 32) sm.d2t[d] = sm.rowcnt[d] == 1 ? sm.hit_t[d] : -1;
            } else {
                int bad;
                int* xsol;
                if (nD <= nT) {
                    // every detection is assigned in every optimum of lapjv's extension (the padding
                    // costs more than any real entry): the fast exact solver applies
                    RectWork w = rect_carve(sm.work, nD, nT);
                    bad = rect_solve_warp(DenseCost{cost, nT}, nD, nT, CUDART_INF, w, sm.rowa, sm.rarg);
                    xsol = w.x;
                } else {
                    double pad = s_cmx[0];
                    for (int i = 1; i < BT_NW; i++) pad = fmax(pad, s_cmx[i]);
                    LapMat m = {cost, nT, nD, nT, nD + nT, 1, pad + 1.0};
                    LapWork w = lap_carve(sm.work, nD + nT);
                    bad = lap_solve_warp(m, w);
                    xsol = w.x;
                }
                if (bad && lane == 0) atomicOr(&v.hdr[BTH_ERR], BM_ERR_LAPBAD);
                for (int d = lane; d < nD; d += 32) { const int t = xsol[d]; sm.d2t[d] = (t >= 0 && t < nT) ? t : -1; }
                if (lane == 0) atomicAdd(&v.hdr[BTH_LAPS], 1);
            }
            __syncwarp();
            // linear_assignment's gate (assoc.py:141-150): IoU >= thr, or IoU >= thr / 2 with an
            // embedding similarity >= 0.75
            for (int d = lane; d < nD; d += 32) {
                const int t = sm.d2t[d];
                int stt = 0;
                if (t >= 0) {
                    const float4 bf = sm.dbox[d];
                    const Box a = {(double)bf.x, (double)bf.y, (double)bf.z, (double)bf.w};
                    const double iou = iou_pair(a, bt_trk_box(sm.tcol, T, t));
                    bool ok = iou >= pr.iou_thr;
                    if (!ok && pr.with_reid) ok = iou >= pr.iou_thr / 2 && (double)sim[(size_t)sm.di[d] * T + t] >= 0.75;
                    stt = ok ? 1 : 2;
                }
                sm.rowcnt[d] = stt;
            }
        }
        __syncthreads();
    }
    if (wid == 0) {
        // unmatched detections: never assigned (ascending), then gate-rejected pairs in row order
        int nud = warp_compact(nD, sm.um_d, 0, [&](int d) { return sm.rowcnt[d] == 0; }, [&](int d) { return d; });
        nud = warp_compact(nD, sm.um_d, nud, [&](int d) { return sm.rowcnt[d] == 2; }, [&](int d) { return d; });
        const int nup = warp_compact(nD, upd_slot, 0, [&](int d) { return sm.rowcnt[d] == 1; },
                                     [&](int d) { return sm.slot_of[sm.d2t[d]]; });
        warp_compact(nD, upd_det, 0, [&](int d) { return sm.rowcnt[d] == 1; }, [&](int d) { return sm.di[d]; });
        int* born = st.born + (size_t)s * Dcap;
        for (int r = lane; r < nud; r += 32) born[r] = sm.di[sm.um_d[r]];
        if (lane == 0) {
            v.hdr[BTH_NUPD] = nup;
            v.hdr[BTH_NBORN] = nud;
            v.hdr[BTH_MATCH] += nup;
        }
    }
}

// ============================================================================ B4
__device__ __forceinline__ int bt_block_excl_scan(int val, int* total, int* ws /*[33]*/) {
    const int lane = lane_id(), wid = warp_id();
    int inc = val;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(BM_FULL, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) ws[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int nw = (blockDim.x + 31) / 32;
        int w = lane < nw ? ws[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(BM_FULL, w, o);
            if (lane >= o) w += t;
        }
        if (lane < nw) ws[lane] = w;
        if (lane == 31) ws[32] = w;
    }
    __syncthreads();
    const int base = wid > 0 ? ws[wid - 1] : 0;
    *total = ws[32];
    __syncthreads();
    return base + inc - val;
}

// KalmanBoxTracker.update (boosttrack.py:72-79) for the frame's matches, one thread per pair.
constexpr int B3U_NT = 64;
__global__ void __launch_bounds__(B3U_NT)
bt_update_kernel(BoostState st, const float* __restrict__ dets) {
    const int chunks = (st.Dcap + B3U_NT - 1) / B3U_NT;
    const int s = blockIdx.x / chunks;
    const int k = (blockIdx.x % chunks) * B3U_NT + threadIdx.x;
    BoostView v(st, s);
    if (k >= v.hdr[BTH_NUPD]) return;
    const int T = st.Tcap, Dcap = st.Dcap;
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    const int* us = st.upd + (size_t)s * 2 * Dcap;
    const int sl = us[k], di = us[Dcap + k];
    const double b[4] = {(double)dsrc[di * 6 + 0], (double)dsrc[di * 6 + 1], (double)dsrc[di * 6 + 2], (double)dsrc[di * 6 + 3]};
    double z[4], x[8], Pb[16];
    bt_box_to_z(b, z);
    bt_load(v.xP, T, sl, x, Pb);
    bt_kf_update(x, Pb, z);
    bt_store(v.xP, T, sl, x, Pb);
    v.I(BTI_TSU, sl) = 0;
    v.I(BTI_STREAK, sl) += 1;
    v.I(BTI_DETIND, sl) = di;
    v.confcls[0 * T + sl] = st.dconf[(size_t)s * Dcap + di];
    v.confcls[1 * T + sl] = (double)dsrc[di * 6 + 5];
}

// One CTA per stream, one thread per slot / list position (blockDim = Tcap rounded up to a warp).
__global__ void __launch_bounds__(1024)
bt_finalize_kernel(BoostState st, BoostParams pr, const float* __restrict__ dets, double* __restrict__ out,
                   int* __restrict__ nout) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int* inuse = reinterpret_cast<int*>(smem_raw);     // [Tcap]
    int* bslot = inuse + st.Tcap;                       // [Tcap]
    __shared__ int ws[33];
    const int s = blockIdx.x, tid = threadIdx.x;
    const int T = st.Tcap, Dcap = st.Dcap;
    BoostView v(st, s);
    const int nT = v.hdr[BTH_NLIST];
    int born = v.hdr[BTH_NBORN];
    const int id0 = v.hdr[BTH_NEXT_ID];
    const int frame = v.hdr[BTH_FRAME] + 1;
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    const double* dconf = st.dconf + (size_t)s * Dcap;
    if (tid < T) inuse[tid] = 0;
    __syncthreads();
    int slot = -1;
    if (tid < nT) { slot = v.order[tid]; inuse[slot] = 1; }
    __syncthreads();
    // births take the free slots in ascending order (ids in unmatched-detection order)
    const int isfree = (tid < T && !inuse[tid]) ? 1 : 0;
    int nfree;
    const int frank = bt_block_excl_scan(isfree, &nfree, ws);
    if (born > nfree) {
        if (tid == 0) atomicOr(&v.hdr[BTH_ERR], BM_ERR_TRACK_OVERFLOW);
        born = nfree;
    }
    if (isfree && frank < born) bslot[frank] = tid;
    __syncthreads();
    if (tid < born) {
        const int di = st.born[(size_t)s * Dcap + tid], sl = bslot[tid];
        const double b[4] = {(double)dsrc[di * 6 + 0], (double)dsrc[di * 6 + 1], (double)dsrc[di * 6 + 2], (double)dsrc[di * 6 + 3]};
        double z[4];
        bt_box_to_z(b, z);
        for (int c = 0; c < 4; c++) {
            v.xP[c * T + sl] = z[c];
            v.xP[(4 + c) * T + sl] = 0.0;
            v.xP[(8 + 4 * c) * T + sl] = 10.0; v.xP[(9 + 4 * c) * T + sl] = 0.0;
            v.xP[(10 + 4 * c) * T + sl] = 0.0; v.xP[(11 + 4 * c) * T + sl] = 10000.0;
        }
        v.I(BTI_ID, sl) = id0 + tid + 1;              // KalmanBoxTracker.count is incremented first (:53-56)
        v.I(BTI_TSU, sl) = 0; v.I(BTI_STREAK, sl) = 0; v.I(BTI_AGE, sl) = 0; v.I(BTI_DETIND, sl) = di;
        v.confcls[0 * T + sl] = dconf[di];
        v.confcls[1 * T + sl] = (double)dsrc[di * 6 + 5];
        if (st.bornslot) st.bornslot[(size_t)s * Dcap + tid] = sl;
    }
    if (tid >= nT && tid < nT + born) slot = bslot[tid - nT];
    __syncthreads();
    // output rows in list order with filter_outputs (:338-346); dead tracks leave the list (:329)
    const int nAll = nT + born;
    int emit = 0, keep = 0;
    double b[4];
    if (tid < nAll) {
        const int tsu = v.I(BTI_TSU, slot);
        emit = (tsu < 1 && (v.I(BTI_STREAK, slot) >= pr.min_hits || frame <= pr.min_hits)) ? 1 : 0;
        keep = tsu > pr.max_age ? 0 : 1;
        if (emit) {
            double x[4];
#pragma unroll
            for (int c = 0; c < 4; c++) x[c] = v.xP[c * T + slot];
            bt_x_to_box(x, b);
            const double w = b[2] - b[0], h = b[3] - b[1];
            if (!(w / h <= pr.ar_thresh && w * h > pr.min_box_area)) emit = 0;
        }
    }
    int n_emit, n_keep;
    const int erank = bt_block_excl_scan(emit, &n_emit, ws);
    const int krank = bt_block_excl_scan(keep, &n_keep, ws);
    if (emit) {
        double* o = out + ((size_t)s * min(T, Dcap) + erank) * 8;
        o[0] = b[0]; o[1] = b[1]; o[2] = b[2]; o[3] = b[3];
        o[4] = (double)v.I(BTI_ID, slot);
        o[5] = v.confcls[0 * T + slot];
        o[6] = v.confcls[1 * T + slot];
        o[7] = (double)v.I(BTI_DETIND, slot);
    }
    if (keep) v.order[krank] = slot;
    if (tid == 0) {
        nout[s] = n_emit;
        v.hdr[BTH_NBORN] = born;
        v.hdr[BTH_NLIST] = n_keep;
        v.hdr[BTH_NEXT_ID] = id0 + born;
        v.hdr[BTH_FRAME] = frame;
        v.hdr[BTH_BIRTHS] += born;
    }
}

// ============================================================================ B5
// update_emb (boosttrack.py:116-118) with alpha = 0.95 + 0.05 (1 - trust), trust from the BOOSTED
// confidence (:299-305), and the raw copy for births (:311-315).  One warp per pair.
constexpr int B5_NT = 256;
__global__ void __launch_bounds__(B5_NT)
bt_emb_kernel(BoostState st, BoostParams pr, const float* __restrict__ embs) {
    const int s = blockIdx.x, lane = lane_id();
    const int T = st.Tcap, Dcap = st.Dcap, E = pr.E;
    const int* hdr = st.hdr + (size_t)s * BTH_COUNT;
    const int nupd = hdr[BTH_NUPD], nborn = hdr[BTH_NBORN];
    const int k = blockIdx.y * (B5_NT / 32) + warp_id();
    if (k >= nupd + nborn) return;
    int slot, di;
    const bool birth = k >= nupd;
    if (!birth) {
        const int* us = st.upd + (size_t)s * 2 * Dcap;
        slot = us[k]; di = us[Dcap + k];
    } else {
        slot = st.bornslot[(size_t)s * Dcap + (k - nupd)];
        di = st.born[(size_t)s * Dcap + (k - nupd)];
    }
    float* dst = st.emb + ((size_t)s * T + slot) * E;
    const float* src = embs + ((size_t)s * Dcap + di) * E;
    const bool vec = (E & 3) == 0;
    if (birth) {
        if (vec) for (int e = lane * 4; e < E; e += 128) *reinterpret_cast<float4*>(dst + e) = __ldg(reinterpret_cast<const float4*>(src + e));
        else for (int e = lane; e < E; e += 32) dst[e] = src[e];
        return;
    }
    const double conf = st.dconf[(size_t)s * Dcap + di];
    const double trust = (conf - pr.det_thresh) / (1 - pr.det_thresh);
    const double af = 0.95;
    const double alpha = af + (1 - af) * (1 - trust);
    double ss = 0.0;
    for (int e = lane; e < E; e += 32) {
        const double t = alpha * (double)dst[e] + (1 - alpha) * (double)src[e];
        ss += t * t;
    }
    for (int o = 16; o > 0; o >>= 1) ss += shfl_xor_d(ss, o);
    const double nrm = sqrt(ss);
    for (int e = lane; e < E; e += 32) {
        const double t = alpha * (double)dst[e] + (1 - alpha) * (double)src[e];
        dst[e] = (float)(t / nrm);
    }
}

struct BoostDot {
    BoostState st;
    const float* embs;
    int E;
    __device__ CosineProblem operator()(int s) const {
        const int* hdr = st.hdr + (size_t)s * BTH_COUNT;
        CosineProblem p;
        p.A = embs + (size_t)s * st.Dcap * E; p.ra = st.hsel + (size_t)s * st.Dcap; p.na = nullptr; p.nA = hdr[BTH_ND];
        p.B = st.emb + (size_t)s * st.Tcap * E; p.rb = st.order + (size_t)s * st.Tcap; p.nb = nullptr; p.nB = hdr[BTH_NLIST];
        p.E = E;
        p.out = st.embsim + (size_t)s * st.Dcap * st.Tcap; p.out_ld = st.Tcap;
        p.raw = 1;
        return p;
    }
};

size_t boosttrack_smem(const BoostState& st) { return bt_smem_bytes(st.Tcap, st.Dcap, bt_threads(st.Tcap, st.Dcap)); }

cudaError_t boosttrack_configure(const BoostState&) {
    cudaError_t e = smem_optin(bt_boost_kernel<256>);
    if (e != cudaSuccess) return e;
    e = smem_optin(bt_assoc_kernel<256>);
    if (e != cudaSuccess) return e;
    e = smem_optin(bt_boost_kernel<BT_NT_MAX>);
    if (e != cudaSuccess) return e;
    e = smem_optin(bt_assoc_kernel<BT_NT_MAX>);
    if (e != cudaSuccess) return e;
    return smem_optin(cosine_tc_kernel<BoostDot>);
}

static cudaError_t bt_dbg(const char* name, cudaStream_t stream) {
    static const bool on = [] { const char* e = getenv("BM_DEBUG_SYNC"); return e && e[0] == '1'; }();
    if (!on) return cudaSuccess;
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) fprintf(stderr, "boxmot_b200: kernel %s failed: %s\n", name, cudaGetErrorString(e));
    return e;
}

// ev: optional 7 events around the 6 launches
cudaError_t boosttrack_step_launch(const BoostState& st, const BoostParams& pr, const float* dets, const int* ndets,
                                   const float* embs, const double* warps, double* out, int* nout,
                                   cudaStream_t stream, cudaEvent_t* ev) {
    const int c1 = (st.Tcap + B1_NT - 1) / B1_NT;
    const size_t sm = boosttrack_smem(st);
    cudaError_t e;
#define BM_EV(k) do { if (ev && (e = cudaEventRecord(ev[k], stream)) != cudaSuccess) return e; } while (0)
    BM_EV(0);
    bt_predict_kernel<<<st.S * c1, B1_NT, 0, stream>>>(st, pr, warps);
    if ((e = bt_dbg("bt_predict_kernel", stream)) != cudaSuccess) return e;
    BM_EV(1);
    const int nt = bt_threads(st.Tcap, st.Dcap);
    if (nt == 256) bt_boost_kernel<256><<<st.S, 256, sm, stream>>>(st, pr, dets, ndets);
    else bt_boost_kernel<BT_NT_MAX><<<st.S, nt, sm, stream>>>(st, pr, dets, ndets);
    if ((e = bt_dbg("bt_boost_kernel", stream)) != cudaSuccess) return e;
    BM_EV(2);
    if (pr.with_reid) {
        cosine_tc_kernel<BoostDot><<<dim3((st.Dcap + TC_M - 1) / TC_M, st.S), TC_NT, cosine_tc_smem_bytes(), stream>>>(
            BoostDot{st, embs, pr.E});
        if ((e = bt_dbg("cosine_tc_kernel<BoostDot>", stream)) != cudaSuccess) return e;
    }
    BM_EV(3);
    if (nt == 256) bt_assoc_kernel<256><<<st.S, 256, sm, stream>>>(st, pr, dets);
    else bt_assoc_kernel<BT_NT_MAX><<<st.S, nt, sm, stream>>>(st, pr, dets);
    if ((e = bt_dbg("bt_assoc_kernel", stream)) != cudaSuccess) return e;
    bt_update_kernel<<<st.S * ((st.Dcap + B3U_NT - 1) / B3U_NT), B3U_NT, 0, stream>>>(st, dets);
    if ((e = bt_dbg("bt_update_kernel", stream)) != cudaSuccess) return e;
    BM_EV(4);
    const int nt4 = (st.Tcap + 31) / 32 * 32;
    bt_finalize_kernel<<<st.S, nt4, (size_t)st.Tcap * 8, stream>>>(st, pr, dets, out, nout);
    if ((e = bt_dbg("bt_finalize_kernel", stream)) != cudaSuccess) return e;
    BM_EV(5);
    if (pr.with_reid) {
        bt_emb_kernel<<<dim3(st.S, (st.Dcap + B5_NT / 32 - 1) / (B5_NT / 32)), B5_NT, 0, stream>>>(st, pr, embs);
        if ((e = bt_dbg("bt_emb_kernel", stream)) != cudaSuccess) return e;
    }
    BM_EV(6);
#undef BM_EV
    return cudaGetLastError();
}

}  // namespace bm
