// Device-resident, stream-batched BoostTrack engine: state layout and per-track helpers.
//
// One engine holds S independent streams (one reference `BoostTrack` instance each,
// boxmot/trackers/boosttrack/boosttrack.py:154-336).  Track state is SoA in HBM,
// [stream][field][slot] with `slot` fastest (coalesced per-field loads by a warp of tracks);
// the reference's `self.trackers` list ORDER (= output row order, cost-matrix column order)
// is the per-stream `order` array (list position -> slot).
#pragma once
#include "common.cuh"
#include "iou.cuh"
#include "lapjv.cuh"
#include "lap_rect.cuh"

namespace bm {

enum { BTI_ID = 0, BTI_TSU, BTI_STREAK, BTI_AGE, BTI_DETIND, BTI_COUNT };
enum { BTH_NLIST = 0, BTH_NEXT_ID, BTH_FRAME, BTH_ERR, BTH_LAPS, BTH_MATCH, BTH_BIRTHS, BTH_NUPD, BTH_NBORN, BTH_ND,
       BTH_COUNT = 12 };
// per-list-position operands written by the predict kernel
enum { TP_X1 = 0, TP_Y1, TP_X2, TP_Y2, TP_CONF, TP_Z0, TP_Z1, TP_Z2, TP_Z3, TP_S0, TP_S1, TP_S2, TP_S3, TP_AGE1, TP_COUNT };
constexpr int BTX_N = 24;            // doubles of Kalman state per track: x[8] + 4 blocks {P[b][b], P[b][b+4], P[b+4][b], P[b+4][b+4]}
constexpr int BT_CL = 32;            // listed (not screened-out) rows per column before a column is evaluated in full
__host__ __device__ inline size_t bt_list_doubles(int Tcap) { return (size_t)Tcap * BT_CL + ((size_t)Tcap * BT_CL + 3) / 4; }
constexpr double BT_MH_LIMIT = 13.2767;   // assoc.py:38, boosttrack.py:377

struct BoostParams {
    double det_thresh, iou_thr, min_box_area, ar_thresh, l_iou, l_mhd, l_shape, dlo_coef;
    int max_age, min_hits, use_dlo, use_duo, s_corr, rich_s, use_sb, use_vt, with_reid, E;
};

struct BoostState {
    int S, Tcap, Dcap, npow;
    double* xP;          // [S][24][Tcap]
    double* confcls;     // [S][2][Tcap]
    int* ints;           // [S][BTI_COUNT][Tcap]
    int* order;          // [S][Tcap]
    int* hdr;            // [S][BTH_COUNT]
    unsigned long long* n_updates;   // [S]
    double* tp;          // [S][TP_COUNT][Tcap] predicted boxes / confidence / Mahalanobis operands by list position
    double* dconf;       // [S][Dcap]  detection confidence after the DLO / DUO boosts
    int* hsel;           // [S][Dcap]  detection rows with boosted conf >= det_thresh, ascending
    double* cost;        // [S][cost_stride] association score matrix (L2-resident); also the column lists' scratch
    size_t cost_stride;  // doubles per stream: max(Dcap*Tcap, column-list scratch)
    int* upd;            // [S][2][Dcap] (slot, detection row) of the frame's matches
    int* born;           // [S][Dcap]  detection rows that start a track, in id order
    int* bornslot;       // [S][Dcap]
    float* emb;          // [S][Tcap][E]   (with_reid)
    float* embsim;       // [S][Dcap][Tcap] dets_embs @ tracker_embs.T by (detection row, list position)
    const double* powtab;   // [npow] 0.9^k evaluated by the host's libm (KalmanBoxTracker.get_confidence)
};

// Streams [s0, s0 + n) as an engine of their own (every array is stream-major); see ocs_substate.
inline BoostState bt_substate(const BoostState& st, int s0, int n, int E) {
    BoostState g = st;
    const size_t o = (size_t)s0, T = st.Tcap, D = st.Dcap;
    g.S = n;
    g.xP += o * BTX_N * T;  g.confcls += o * 2 * T;  g.ints += o * BTI_COUNT * T;  g.order += o * T;
    g.hdr += o * BTH_COUNT;  g.n_updates += o;  g.tp += o * TP_COUNT * T;
    g.dconf += o * D;  g.hsel += o * D;  g.cost += o * st.cost_stride;
    g.upd += o * 2 * D;  g.born += o * D;
    if (st.bornslot) g.bornslot += o * D;
    if (st.emb) g.emb += o * T * E;
    if (st.embsim) g.embsim += o * D * T;
    return g;
}

struct BoostView {
    double *xP, *confcls;
    int *ints, *order, *hdr;
    int Tcap;
    __device__ BoostView(const BoostState& st, int s) : Tcap(st.Tcap) {
        const size_t T = st.Tcap;
        xP = st.xP + (size_t)s * BTX_N * T;
        confcls = st.confcls + (size_t)s * 2 * T;
        ints = st.ints + (size_t)s * BTI_COUNT * T;
        order = st.order + (size_t)s * T;
        hdr = st.hdr + (size_t)s * BTH_COUNT;
    }
    __device__ __forceinline__ int& I(int f, int slot) const { return ints[f * Tcap + slot]; }
};

// convert_bbox_to_z (boosttrack.py:19-28): [x, y, h, r = w / (h + 1e-6)]
__device__ __forceinline__ void bt_box_to_z(const double* b, double* z) {
    const double w = b[2] - b[0], h = b[3] - b[1];
    z[0] = b[0] + w / 2.0;
    z[1] = b[1] + h / 2.0;
    z[2] = h;
    z[3] = w / (h + 1e-6);
}

// convert_x_to_bbox (boosttrack.py:31-42)
__device__ __forceinline__ void bt_x_to_box(const double* x, double* b) {
    const double h = x[2], r = x[3];
    const double w = r <= 0 ? 0.0 : r * h;
    b[0] = x[0] - w / 2.0;
    b[1] = x[1] - h / 2.0;
    b[2] = x[0] + w / 2.0;
    b[3] = x[1] + h / 2.0;
}

// KalmanFilter.predict (kalmanfilter.py:84-117): x <- F x, P <- F (P F^T) + Q with
// Q = diag(1,1,1,1,.01,.01,.01,.01); the covariance only couples (i, i+4), so the four 2x2
// blocks are evaluated with the dense code's own term order.
__device__ __forceinline__ void bt_kf_predict(double* x, double* Pb) {
#pragma unroll
    for (int b = 0; b < 4; b++) {
        x[b] = x[b] + x[b + 4];
        const double p = Pb[4 * b], c1 = Pb[4 * b + 1], c2 = Pb[4 * b + 2], v = Pb[4 * b + 3];
        Pb[4 * b] = ((p + c1) + (c2 + v)) + 1.0;
        Pb[4 * b + 1] = c1 + v;
        Pb[4 * b + 2] = c2 + v;
        Pb[4 * b + 3] = v + 0.01;
    }
}

// KalmanFilter.update (kalmanfilter.py:119-157): S = H P H^T + R is diagonal here, so the
// Cholesky solve is two divisions by sqrt(S_b) per entry.  R = diag(1, 1, 10, 0.01).
__device__ __forceinline__ void bt_kf_update(double* x, double* Pb, const double* z) {
    const double R[4] = {1.0, 1.0, 10.0, 0.01};
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const double p = Pb[4 * b], c1 = Pb[4 * b + 1], c2 = Pb[4 * b + 2], v = Pb[4 * b + 3];
        const double Sb = p + R[b];
        const double l = sqrt(Sb);
        const double k0 = (p / l) / l, k1 = (c2 / l) / l;
        const double y = z[b] - x[b];
        x[b] = x[b] + y * k0;
        x[b + 4] = x[b + 4] + y * k1;
        Pb[4 * b] = p - k0 * (Sb * k0);
        Pb[4 * b + 1] = c1 - k0 * (Sb * k1);
        Pb[4 * b + 2] = c2 - k1 * (Sb * k0);
        Pb[4 * b + 3] = v - k1 * (Sb * k1);
    }
}

__host__ __device__ inline void bt_expand(const double* Pb, double* P /*64*/) {
    for (int i = 0; i < 64; i++) P[i] = 0.0;
    for (int b = 0; b < 4; b++) {
        P[b * 8 + b] = Pb[4 * b];
        P[b * 8 + b + 4] = Pb[4 * b + 1];
        P[(b + 4) * 8 + b] = Pb[4 * b + 2];
        P[(b + 4) * 8 + b + 4] = Pb[4 * b + 3];
    }
}

size_t boosttrack_smem(const BoostState& st);
cudaError_t boosttrack_configure(const BoostState& st);
cudaError_t boosttrack_step_launch(const BoostState& st, const BoostParams& pr, const float* dets, const int* ndets,
                                   const float* embs, const double* warps, double* out, int* nout,
                                   cudaStream_t stream, cudaEvent_t* ev = nullptr);

}  // namespace bm
