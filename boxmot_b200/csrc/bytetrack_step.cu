// ByteTrack frame step: ONE launch advances every stream by one frame; one CTA of two warps
// per stream runs ByteTrack.update (boxmot/trackers/bytetrack/bytetrack.py:158-302) end to
// end -- score split, pool build + multi_predict, the three thresholded assignments
// (high-score, low-score, unconfirmed), births, list bookkeeping, duplicate removal and the
// output rows -- with every list kept in the reference's order.
//
// Costs: iou_distance (boxmot/utils/matching.py:184-227) and fuse_score (:488-560) in fp64
// with the reference's operation order; assignment: lap.lapjv(cost, extend_cost=True,
// cost_limit=thresh) (:54-55) via the exact rectangular solver with the null option
// (csrc/lap_rect.cuh; unmatched lists are index-sorted in the reference, so only the optimal
// matching itself is observable).  Kalman filter: csrc/kf_xyah.cuh.
#include "bytetrack_engine.cuh"
#include "lap_rect.cuh"
#include "cosine_tc.cuh"
#include <cstdio>
#include <cstdlib>

namespace bm {

constexpr int BT_NT = 64;
enum { TF_MATCHED = 1, TF_NEWLOST = 2, TF_REFIND = 4, TF_DUP = 8, TF_USED = 16, TF_AGED = 32 };
enum { BM_NH = 0, BM_NLO, BM_NTR, BM_NP, BM_NU, BM_NUPD, BM_NR, BM_NREST, BM_NBORN, BM_NNL, BM_NRF, BM_NA2, BM_NL2,
       BM_BAD, BM_N1, BM_N2, BM_NFU, BM_MISC = 24 };

struct ByteSmem {
    float4 *hbox, *lbox;      // [Dcap] raw high / low score detections
    float* hconf;             // [Dcap]
    void* work;               // solver work area
    int *hind, *lind, *dflag, *rest, *rflag, *born;  // [Dcap]
    int *pool, *unc, *rtr, *xs, *nlost, *refind, *lista, *listb, *tflag, *upd_slot, *upd_det, *rarg;   // [Tcap]
    double* rmin;             // [Tcap] row minima of the cost matrix being built
    int* misc;
};

__host__ __device__ inline size_t byte_smem_bytes(int Tcap, int Dcap) {
    size_t b = 2 * 16 * (size_t)Dcap + (rect_work_bytes(Tcap, Dcap) + 15) / 16 * 16;
    b += 8 * (size_t)Tcap + 4 * ((size_t)Dcap + 6 * (size_t)Dcap + 12 * (size_t)Tcap + BM_MISC);
    return (b + 15) / 16 * 16;
}

__device__ inline ByteSmem byte_carve(unsigned char* base, int Tcap, int Dcap) {
    ByteSmem m;
    m.hbox = reinterpret_cast<float4*>(base); base += 16 * (size_t)Dcap;
    m.lbox = reinterpret_cast<float4*>(base); base += 16 * (size_t)Dcap;
    m.work = base; base += (rect_work_bytes(Tcap, Dcap) + 15) / 16 * 16;
    m.rmin = reinterpret_cast<double*>(base); base += 8 * (size_t)Tcap;
    m.hconf = reinterpret_cast<float*>(base); base += 4 * (size_t)Dcap;
    int* q = reinterpret_cast<int*>(base);
    m.hind = q; q += Dcap; m.lind = q; q += Dcap; m.dflag = q; q += Dcap; m.rest = q; q += Dcap; m.rflag = q; q += Dcap; m.born = q; q += Dcap;
    m.pool = q; q += Tcap; m.unc = q; q += Tcap; m.rtr = q; q += Tcap; m.xs = q; q += Tcap; m.nlost = q; q += Tcap;
    m.refind = q; q += Tcap; m.lista = q; q += Tcap; m.listb = q; q += Tcap; m.tflag = q; q += Tcap;
    m.upd_slot = q; q += Tcap; m.upd_det = q; q += Tcap; m.rarg = q; q += Tcap;
    m.misc = q;
    return m;
}

// STrack.__init__ conversions (bytetrack.py:17-21; boxmot/utils/ops.py:10-23, :46-60, :90-103)
struct DetGeom { double xc, yc, w, h; };
__device__ __forceinline__ DetGeom det_xywh(const float4 r) {
    const double x1 = r.x, y1 = r.y, x2 = r.z, y2 = r.w;
    return DetGeom{(x1 + x2) / 2, (y1 + y2) / 2, x2 - x1, y2 - y1};
}
__device__ __forceinline__ Box det_xyxy(const float4 r) {           // STrack.xyxy with mean = None (:105-116)
    const DetGeom g = det_xywh(r);
    return Box{g.xc - g.w / 2, g.yc - g.h / 2, g.xc + g.w / 2, g.yc + g.h / 2};
}
__device__ __forceinline__ void det_xyah(const float4 r, double* z) {
    const DetGeom g = det_xywh(r);
    const double tx = g.xc - g.w / 2, ty = g.yc - g.h / 2;          // xywh2tlwh
    z[0] = tx + g.w / 2; z[1] = ty + g.h / 2; z[2] = g.w / g.h; z[3] = g.h;   // tlwh2xyah
}

// enhanced_fuse_score(cost, dets, 0.5, True) on one entry (matching.py:520-544)
__device__ __forceinline__ double fuse_entry(double cost, double conf) {
    const double sim = 1.0 - cost;
    const double weight = conf > 0.7 ? conf * 1.2 : conf;
    const double mask = conf >= 0.5 ? 1.0 : 0.0;
    double fused = 1.0 - (sim * weight) * mask;
    if (conf < 0.5) fused = fused * 2.0;
    return fused;
}


// Measurement of a detection and box of a mean for the two filters:
//   XYAH (ByteTrack): tlwh2xyah(xywh2tlwh(xyxy2xywh(det)))  (bytetrack.py:17-21), box :105-116
//   XYWH (BoT-SORT):  xyxy2xywh(det)                         (botsort_track.py:15), box :155-159
template <int KIND>
__device__ __forceinline__ void det_measurement(const float4 r, double* z) {
    if (KIND == CV8_XYAH) { det_xyah(r, z); return; }
    const DetGeom g = det_xywh(r);
    z[0] = g.xc; z[1] = g.yc; z[2] = g.w; z[3] = g.h;
}
template <int KIND>
__device__ __forceinline__ Box mean_to_xyxy(const double* x) {
    if (KIND == CV8_XYAH) return xyah_to_xyxy(x);
    return Box{x[0] - x[2] / 2, x[1] - x[3] / 2, x[0] + x[2] / 2, x[1] + x[3] / 2};
}

// BoT-SORT cost entry (botsort.py:223-234 first round, :311-326 unconfirmed round)
__device__ __forceinline__ double bot_cost(double iou_d, double conf, bool fuse, const ByteParams& pr,
                                           const float* embd, int idx) {
    const bool far = iou_d > pr.proximity_thresh;
    if (fuse) iou_d = fuse_entry(iou_d, conf);
    if (!pr.with_reid) return iou_d;
    double e = (double)embd[idx] / 2.0;
    if (e > pr.appearance_thresh) e = 1.0;
    if (far) e = 1.0;
    return fmin(iou_d, e);
}

// STrack.multi_gmc (botsort_track.py:92-104) for a warp without rotation / shear: kron(I4, R)
// with diagonal R scales the (x, w, vx, vw) and (y, h, vy, vh) components separately and keeps
// the covariance in block form.  Warps with rotation / shear go through the dense slow path
// (botd_predict_gmc): the callers switch the track over first.
__device__ __forceinline__ void bot_gmc(double* x, double* P, const double* w6, int* hdr) {
    const double r00 = w6[0], r01 = w6[1], tx = w6[2], r10 = w6[3], r11 = w6[4], ty = w6[5];
    if (r01 != 0.0 || r10 != 0.0) { atomicOr(&hdr[BH_ERR], BM_ERR_WARP); return; }   // not reached
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double r = (i & 1) ? r11 : r00;
        x[i] = r * x[i]; x[4 + i] = r * x[4 + i];
        P[i] = (r * P[i]) * r; P[4 + i] = (r * P[4 + i]) * r; P[8 + i] = (r * P[8 + i]) * r; P[12 + i] = (r * P[12 + i]) * r;
    }
    x[0] += tx; x[1] += ty;
}

// STrack.update_cls (botsort_track.py:51-64) after `self.cls = new_track.cls`: returns the
// track's class.  hist: [2*BOT_HIST][Tcap] doubles (class ids, then vote sums) of one stream.
__device__ __forceinline__ double bot_vote(double* hist, int Tcap, int slot, int& nh, double cls, double conf, int* hdr) {
    double out = cls, best = 0.0;
    bool found = false;
    for (int i = 0; i < nh; i++) {
        const double c = hist[i * Tcap + slot];
        double sum = hist[(BOT_HIST + i) * Tcap + slot];
        if (cls == c) { sum += conf; hist[(BOT_HIST + i) * Tcap + slot] = sum; found = true; }
        if (sum > best) { best = sum; out = c; }
    }
    if (!found) {
        if (nh < BOT_HIST) {
            hist[nh * Tcap + slot] = cls;
            hist[(BOT_HIST + nh) * Tcap + slot] = conf;
            nh++;
        } else {
            atomicOr(&hdr[BH_ERR], BM_ERR_CLS_HIST);
        }
        out = cls;
    }
    return out;
}

struct ByteView {
    double *xP, *confcls, *tbox, *dP;
    int *ints, *act, *lost, *hdr;
    int Tcap;
    __device__ ByteView(const ByteState& st, int s) : Tcap(st.Tcap) {
        const size_t T = st.Tcap;
        xP = st.xP + (size_t)s * XA_N * T;
        dP = st.dP ? st.dP + (size_t)s * 64 * T : nullptr;
        confcls = st.confcls + (size_t)s * 2 * T;
        tbox = st.tbox + (size_t)s * 4 * T;
        ints = st.ints + (size_t)s * BI_COUNT * T;
        act = st.act + (size_t)s * T;
        lost = st.lost + (size_t)s * T;
        hdr = st.hdr + (size_t)s * BH_COUNT;
    }
    __device__ __forceinline__ int& I(int f, int slot) const { return ints[f * Tcap + slot]; }
    __device__ __forceinline__ Box box(int slot) const {
        return Box{tbox[slot], tbox[Tcap + slot], tbox[2 * Tcap + slot], tbox[3 * Tcap + slot]};
    }
    __device__ __forceinline__ void set_box(int slot, const Box& b) const {
        tbox[slot] = b.x1; tbox[Tcap + slot] = b.y1; tbox[2 * Tcap + slot] = b.x2; tbox[3 * Tcap + slot] = b.y2;
    }
    __device__ __forceinline__ void load(int slot, double* x, double* P) const {
#pragma unroll
        for (int c = 0; c < 8; c++) x[c] = xP[c * Tcap + slot];
#pragma unroll
        for (int c = 0; c < XA_P; c++) P[c] = xP[(8 + c) * Tcap + slot];
    }
    __device__ __forceinline__ void store(int slot, const double* x, const double* P) const {
#pragma unroll
        for (int c = 0; c < 8; c++) xP[c * Tcap + slot] = x[c];
#pragma unroll
        for (int c = 0; c < XA_P; c++) xP[(8 + c) * Tcap + slot] = P[c];
    }
};

// ---- dense 8x8 slow path of the XYWH filter (BoT-SORT tracks that saw a camera-motion warp with
// rotation / shear: kron(I4, R) couples x<->y, w<->h and their velocities, botsort_track.py:91-104).
// The reference's own dense expressions (base_kalman_filter.py:111-155, xywh_kf.py:29-66); kept out of line
// so that the register budget of the step kernel's common path is untouched.
__device__ __forceinline__ void botd_load(const ByteView& v, int slot, double* x, double* P) {
    for (int c = 0; c < 8; c++) x[c] = v.xP[c * v.Tcap + slot];
    for (int c = 0; c < 64; c++) P[c] = v.dP[(size_t)c * v.Tcap + slot];
}
__device__ __forceinline__ void botd_store(const ByteView& v, int slot, const double* x, const double* P) {
    for (int c = 0; c < 8; c++) v.xP[c * v.Tcap + slot] = x[c];
    for (int c = 0; c < 64; c++) v.dP[(size_t)c * v.Tcap + slot] = P[c];
}
__device__ __noinline__ void bot_make_dense(const ByteView& v, int slot) {
    double pb[XA_P], P[64];
    for (int c = 0; c < XA_P; c++) pb[c] = v.xP[(8 + c) * v.Tcap + slot];
    xyah_expand(pb, P);
    for (int c = 0; c < 64; c++) v.dP[(size_t)c * v.Tcap + slot] = P[c];
    v.I(BI_DENSE, slot) = 1;
}
// multi_predict (optional) followed by multi_gmc with any 2x3 warp (optional), on a dense track
__device__ __noinline__ void botd_predict_gmc(const ByteView& v, int slot, bool zero_size_vel, bool predict, const double* w6) {
    double x[8], P[64];
    botd_load(v, slot, x, P);
    if (predict) {
        if (zero_size_vel) { x[6] = 0.0; x[7] = 0.0; }
        double sp[4], sv[4];
        cv8_std<CV8_XYWH>(x, 1.0, 1.0, 1e-2, 1e-5, sp, sv);
        for (int i = 0; i < 4; i++) x[i] = x[i] + x[4 + i];
        for (int i = 0; i < 4; i++)                      // F P
            for (int j = 0; j < 8; j++) P[i * 8 + j] = P[i * 8 + j] + P[(i + 4) * 8 + j];
        for (int i = 0; i < 8; i++)                      // (F P) F^T
            for (int j = 0; j < 4; j++) P[i * 8 + j] = P[i * 8 + j] + P[i * 8 + j + 4];
        for (int i = 0; i < 4; i++) { P[i * 9] = P[i * 9] + sp[i] * sp[i]; P[(4 + i) * 9] = P[(4 + i) * 9] + sv[i] * sv[i]; }
    }
    if (w6) {
        const double r[4] = {w6[0], w6[1], w6[3], w6[4]};
        for (int b = 0; b < 4; b++) {                    // mean = R8x8 . mean (+ t on the first pair)
            const double a = x[2 * b], c = x[2 * b + 1];
            x[2 * b] = r[0] * a + r[1] * c;
            x[2 * b + 1] = r[2] * a + r[3] * c;
        }
        x[0] += w6[2]; x[1] += w6[5];
        double T[64];
        for (int bi = 0; bi < 4; bi++)                   // T = R8x8 . P
            for (int j = 0; j < 8; j++) {
                const double a = P[(2 * bi) * 8 + j], c = P[(2 * bi + 1) * 8 + j];
                T[(2 * bi) * 8 + j] = r[0] * a + r[1] * c;
                T[(2 * bi + 1) * 8 + j] = r[2] * a + r[3] * c;
            }
        for (int i = 0; i < 8; i++)                      // P = T . R8x8^T
            for (int bj = 0; bj < 4; bj++) {
                const double a = T[i * 8 + 2 * bj], c = T[i * 8 + 2 * bj + 1];
                P[i * 8 + 2 * bj] = a * r[0] + c * r[1];
                P[i * 8 + 2 * bj + 1] = a * r[2] + c * r[3];
            }
    }
    botd_store(v, slot, x, P);
}
// KalmanFilterXYWH.update (base_kalman_filter.py:129-155) with a dense covariance
__device__ __noinline__ void botd_update(const ByteView& v, int slot, const double* z) {
    double x[8], P[64];
    botd_load(v, slot, x, P);
    double sp[4], sv[4];
    cv8_std<CV8_XYWH>(x, 1.0, 1.0, 1e-1, 0.0, sp, sv);
    double S[16], L[16];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) S[i * 4 + j] = P[i * 8 + j] + (i == j ? sp[i] * sp[i] : 0.0);
    for (int i = 0; i < 16; i++) L[i] = 0.0;
    for (int j = 0; j < 4; j++) {                        // cho_factor (lower)
        double d = S[j * 4 + j];
        for (int k = 0; k < j; k++) d -= L[j * 4 + k] * L[j * 4 + k];
        d = sqrt(d);
        L[j * 4 + j] = d;
        for (int i = j + 1; i < 4; i++) {
            double a = S[i * 4 + j];
            for (int k = 0; k < j; k++) a -= L[i * 4 + k] * L[j * 4 + k];
            L[i * 4 + j] = a / d;
        }
    }
    double K[32];                                        // K = P H^T S^-1: cho_solve per row of P[:, :4]
    for (int r = 0; r < 8; r++) {
        double b[4];
        for (int i = 0; i < 4; i++) b[i] = P[r * 8 + i];
        for (int i = 0; i < 4; i++) { for (int k = 0; k < i; k++) b[i] -= L[i * 4 + k] * b[k]; b[i] = b[i] / L[i * 4 + i]; }
        for (int i = 3; i >= 0; i--) { for (int k = i + 1; k < 4; k++) b[i] -= L[k * 4 + i] * b[k]; b[i] = b[i] / L[i * 4 + i]; }
        for (int i = 0; i < 4; i++) K[r * 4 + i] = b[i];
    }
    double inn[4];
    for (int i = 0; i < 4; i++) inn[i] = z[i] - x[i];
    for (int r = 0; r < 8; r++) {
        double acc = 0.0;
        for (int i = 0; i < 4; i++) acc += inn[i] * K[r * 4 + i];
        x[r] = x[r] + acc;
    }
    double SK[32];                                       // S K^T (4 x 8); multi_dot evaluates K (S K^T)
    for (int i = 0; i < 4; i++)
        for (int r = 0; r < 8; r++) {
            double acc = 0.0;
            for (int k = 0; k < 4; k++) acc += S[i * 4 + k] * K[r * 4 + k];
            SK[i * 8 + r] = acc;
        }
    for (int a = 0; a < 8; a++)
        for (int b = 0; b < 8; b++) {
            double acc = 0.0;
            for (int k = 0; k < 4; k++) acc += K[a * 4 + k] * SK[k * 8 + b];
            P[a * 8 + b] = P[a * 8 + b] - acc;
        }
    botd_store(v, slot, x, P);
}

// iou_pair with the division skipped when the boxes do not overlap: max(0, w) * max(0, h) is
// then exactly 0 and 0 / (area_a + area_b) is exactly +0.0 whenever that sum is positive
// (anything else -- degenerate or non-finite boxes -- takes the full expression).
__device__ __forceinline__ double iou_sparse(const Box& a, const Box& b) {
    const double w = fmin(a.x2, b.x2) - fmax(a.x1, b.x1), h = fmin(a.y2, b.y2) - fmax(a.y1, b.y1);
    if (!(w > 0.0 && h > 0.0) && (a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1) > 0.0) return 0.0;
    return iou_pair(a, b);
}

// Cost matrix [nr][nc] of one association round, built a row per warp: entry (r, c) =
// f(1 - iou(row box, detection box), detection position); the exact row minimum and its
// lowest column go to rmin / rarg for the solver's row reduction.  dsel maps the column to the
// detection position (nullptr = identity).
template <typename CostOf>
__device__ __forceinline__ void byte_cost_rows(const ByteView& v, const int* rowslot, int nr, const float4* dbox,
                                               const int* dsel, int nc, double* cost, double* rmin, int* rarg, CostOf f) {
    const int lane = lane_id();
    for (int r = warp_id(); r < nr; r += BT_NT / 32) {
        const Box a = v.box(rowslot[r]);
        double m = CUDART_INF;
        int arg = 0x7fffffff;
        for (int c = lane; c < nc; c += 32) {
            const int d = dsel ? dsel[c] : c;
            const double cc = f(1.0 - iou_sparse(a, det_xyxy(dbox[d])), d);
            cost[(size_t)r * nc + c] = cc;
            if (cc < m) { m = cc; arg = c; }
        }
        warp_argmin(m, arg);
        if (lane == 0) { rmin[r] = m; rarg[r] = arg; }
    }
}

// One thresholded assignment (enhanced_linear_assignment, matching.py:30-108) by warp 0 on the
// cost matrix in `cost` [nr][nc]: xs[r] = column or -1 (row unmatched, x < 0) or -2 (assigned
// above the threshold: in neither list, as in the reference); colflag[c] |= 1 when y[c] >= 0.
__device__ inline void byte_assign(const double* cost, int nr, int nc, double thresh, ByteSmem& sm, int* xs,
                                   int* colflag, int* hdr) {
    const int lane = lane_id();
    if (nr == 0 || nc == 0) {
        for (int r = lane; r < nr; r += 32) xs[r] = -1;
        __syncwarp();
        return;
    }
    RectWork w = rect_carve(sm.work, nr, nc);
    const int bad = rect_solve_warp(DenseCost{cost, nc}, nr, nc, thresh, w, sm.rmin, sm.rarg);
    if (bad && lane == 0) sm.misc[BM_BAD] = 1;
    for (int r = lane; r < nr; r += 32) {
        const int c = w.x[r];
        int v = -1;
        if (c >= 0) { v = cost[(size_t)r * nc + c] <= thresh ? c : -2; colflag[c] |= 1; }
        xs[r] = v;
    }
    if (lane == 0) atomicAdd(&hdr[BH_LAPS], 1);
    __syncwarp();
}

// MODE 0 = ByteTrack (XYAH filter, fuse_score in the first round), MODE 1 = BoT-SORT
// (boxmot/trackers/botsort/botsort.py:94-411: XYWH filter, camera-motion warp, IoU / embedding
// fusion with proximity and appearance gates, class voting, new_track_thresh; the embedding
// distances come from bot_embed_kernel, the feature EMA is applied by bot_ema_kernel).
// <= 146 registers: seven CTAs per SM keep all 1024 streams of the bench in one wave on 148 SMs
enum { MODE_BYTE = 0, MODE_BOT = 1 };
template <int MODE>
__global__ void __launch_bounds__(BT_NT, 7)
byte_step_kernel(ByteState st, ByteParams pr, const float* __restrict__ dets, const int* __restrict__ ndets,
                 const double* __restrict__ warps, double* __restrict__ out, int* __restrict__ nout) {
    constexpr int KIND = MODE == MODE_BOT ? CV8_XYWH : CV8_XYAH;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    ByteSmem sm = byte_carve(smem_raw, Tcap, Dcap);
    ByteView v(st, s);
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    double* cost = st.scratch + (size_t)s * Tcap * Dcap;
    const int nA = v.hdr[BH_NACT], nL = v.hdr[BH_NLOST], id0 = v.hdr[BH_NEXT_ID];
    const int frame = v.hdr[BH_FRAME] + 1;
    int nD_in = ndets[s];
    if (nD_in > Dcap) { if (tid == 0) atomicOr(&v.hdr[BH_ERR], BM_ERR_DET_OVERFLOW); nD_in = Dcap; }
    for (int i = tid; i < Tcap; i += BT_NT) sm.tflag[i] = 0;
    for (int i = tid; i < Dcap; i += BT_NT) { sm.dflag[i] = 0; sm.rflag[i] = 0; }
    if (tid < BM_MISC) sm.misc[tid] = 0;
    __syncthreads();

    // ---- score split (bytetrack.py:171-181) by warp 0; pool / unconfirmed lists (:188-198) by warp 1
    if (wid == 0) {
        auto conf = [&](int i) { return (double)dsrc[i * 6 + 4]; };
        const int nh = warp_compact(nD_in, sm.hind, 0, [&](int i) { return conf(i) > pr.track_thresh; },
                                    [&](int i) { return i; });
        const int nl = warp_compact(nD_in, sm.lind, 0,
                                    [&](int i) { return conf(i) > pr.min_conf && conf(i) < pr.track_thresh; },
                                    [&](int i) { return i; });
        if (lane == 0) { sm.misc[BM_NH] = nh; sm.misc[BM_NLO] = nl; }
    } else {
        const int ntr = warp_compact(nA, sm.pool, 0, [&](int i) { return v.I(BI_ACTIVATED, v.act[i]) != 0; },
                                     [&](int i) { return v.act[i]; });
        const int nu = warp_compact(nA, sm.unc, 0, [&](int i) { return v.I(BI_ACTIVATED, v.act[i]) == 0; },
                                    [&](int i) { return v.act[i]; });
        for (int i = lane; i < nL; i += 32) sm.pool[ntr + i] = v.lost[i];      // joint_stracks: ids are disjoint
        if (lane == 0) { sm.misc[BM_NTR] = ntr; sm.misc[BM_NP] = ntr + nL; sm.misc[BM_NU] = nu; }
    }
    __syncthreads();
    const int nH = sm.misc[BM_NH], nLo = sm.misc[BM_NLO], nP = sm.misc[BM_NP], nU = sm.misc[BM_NU];
    for (int d = tid; d < nH; d += BT_NT) {
        const float* r = dsrc + sm.hind[d] * 6;
        sm.hbox[d] = make_float4(r[0], r[1], r[2], r[3]);
        sm.hconf[d] = r[4];
    }
    for (int d = tid; d < nLo; d += BT_NT) {
        const float* r = dsrc + sm.lind[d] * 6;
        sm.lbox[d] = make_float4(r[0], r[1], r[2], r[3]);
    }
    // ---- STrack.multi_predict on the pool (:41-54); unconfirmed tracks are NOT predicted
    const double* w6 = (MODE == MODE_BOT && warps) ? warps + (size_t)s * 6 : nullptr;
    const bool rot = w6 && (w6[1] != 0.0 || w6[3] != 0.0);
    for (int p = tid; p < nP; p += BT_NT) {
        const int slot = sm.pool[p];
        if (MODE == MODE_BOT && v.dP && (rot || v.I(BI_DENSE, slot))) {      // dense covariance (slow path)
            if (!v.I(BI_DENSE, slot)) bot_make_dense(v, slot);
            botd_predict_gmc(v, slot, v.I(BI_STATE, slot) != BT_TRACKED, true, w6);
            double xm[8];
            for (int c = 0; c < 8; c++) xm[c] = v.xP[c * Tcap + slot];
            v.set_box(slot, mean_to_xyxy<KIND>(xm));
            sm.tflag[slot] |= TF_USED;
            continue;
        }
        double x[8], P[XA_P];
        v.load(slot, x, P);
        if (v.I(BI_STATE, slot) != BT_TRACKED) {
            x[7] = 0.0;
            if (MODE == MODE_BOT) x[6] = 0.0;       // botsort_track.py:83-85 resets both size velocities
        }
        cv8_predict<KIND, true>(x, P);       // STrack.multi_predict
        if (MODE == MODE_BOT && warps) bot_gmc(x, P, warps + (size_t)s * 6, v.hdr);
        v.store(slot, x, P);
        v.set_box(slot, mean_to_xyxy<KIND>(x));
        sm.tflag[slot] |= TF_USED;
    }
    if (st.pc_act) {      // per_class: the other classes' active tracks keep their slots
        const int cur = st.pc_meta[(size_t)s * 2];
        for (int c = 0; c < st.pc_C; c++) {
            if (c == cur) continue;
            const int m = st.pc_n[(size_t)s * st.pc_C + c];
            const int* l = st.pc_act + ((size_t)s * st.pc_C + c) * Tcap;
            for (int i = tid; i < m; i += BT_NT) sm.tflag[l[i]] |= TF_USED;
        }
    }
    for (int p = tid; p < nU; p += BT_NT) {          // unconfirmed tracks are not predicted, only warped
        const int slot = sm.unc[p];
        sm.tflag[slot] |= TF_USED;
        if (MODE == MODE_BOT && warps) {
            if (v.dP && (rot || v.I(BI_DENSE, slot))) {
                if (!v.I(BI_DENSE, slot)) bot_make_dense(v, slot);
                botd_predict_gmc(v, slot, false, false, w6);
                double xm[8];
                for (int c = 0; c < 8; c++) xm[c] = v.xP[c * Tcap + slot];
                v.set_box(slot, mean_to_xyxy<KIND>(xm));
                continue;
            }
            double x[8], P[XA_P];
            v.load(slot, x, P);
            bot_gmc(x, P, warps + (size_t)s * 6, v.hdr);
            v.store(slot, x, P);
            v.set_box(slot, mean_to_xyxy<KIND>(x));
        }
    }
    if (tid == 0) st.n_updates[s] += (unsigned long long)nP;
    __syncthreads();

    // ---- first association: pool x high-score detections, fuse_score, thresh = match_thresh (:199-219)
    const float* embd = MODE == MODE_BOT ? st.embd + (size_t)s * Tcap * Dcap : nullptr;
    if (MODE == MODE_BYTE) {
        byte_cost_rows(v, sm.pool, nP, sm.hbox, nullptr, nH, cost, sm.rmin, sm.rarg,
                       [&](double iou_d, int d) { return fuse_entry(iou_d, (double)sm.hconf[d]); });
    } else {
        // the row's slot is needed for the embedding lookup: one row per warp, so it is warp-uniform
        for (int r = wid; r < nP; r += BT_NT / 32) {
            const int slot = sm.pool[r];
            const Box a = v.box(slot);
            double m = CUDART_INF;
            int arg = 0x7fffffff;
            for (int c = lane; c < nH; c += 32) {
                const double cc = bot_cost(1.0 - iou_sparse(a, det_xyxy(sm.hbox[c])), (double)sm.hconf[c],
                                           pr.fuse_first != 0, pr, embd, slot * Dcap + c);
                cost[(size_t)r * nH + c] = cc;
                if (cc < m) { m = cc; arg = c; }
            }
            warp_argmin(m, arg);
            if (lane == 0) { sm.rmin[r] = m; sm.rarg[r] = arg; }
        }
    }
    __syncthreads();
    if (wid == 0) {
        byte_assign(cost, nP, nH, pr.match_thresh, sm, sm.xs, sm.dflag, v.hdr);
        const int n1 = warp_compact(nP, sm.upd_slot, 0, [&](int t) { return sm.xs[t] >= 0; }, [&](int t) { return sm.pool[t]; });
        warp_compact(nP, sm.upd_det, 0, [&](int t) { return sm.xs[t] >= 0; }, [&](int t) { return sm.hind[sm.xs[t]]; });
        // re_activate (:68-80): matched pool tracks that were not Tracked, in pool order
        const int nrf = warp_compact(nP, sm.refind, 0,
                                     [&](int t) { return sm.xs[t] >= 0 && v.I(BI_STATE, sm.pool[t]) != BT_TRACKED; },
                                     [&](int t) { return sm.pool[t]; });
        for (int t = lane; t < nP; t += 32) if (sm.xs[t] >= 0) sm.tflag[sm.pool[t]] |= TF_MATCHED;
        for (int k = lane; k < nrf; k += 32) sm.tflag[sm.refind[k]] |= TF_REFIND;
        // r_tracked_stracks (:231-235): unmatched pool rows (x < 0) whose state is Tracked
        const int nr = warp_compact(nP, sm.rtr, 0,
                                    [&](int t) { return sm.xs[t] == -1 && v.I(BI_STATE, sm.pool[t]) == BT_TRACKED; },
                                    [&](int t) { return sm.pool[t]; });
        // detections left for the unconfirmed round: y < 0 (:252)
        const int nrest = warp_compact(nH, sm.rest, 0, [&](int d) { return sm.dflag[d] == 0; }, [&](int d) { return d; });
        if (lane == 0) { sm.misc[BM_NUPD] = n1; sm.misc[BM_N1] = n1; sm.misc[BM_NRF] = nrf; sm.misc[BM_NR] = nr; sm.misc[BM_NREST] = nrest; }
    }
    __syncthreads();

    // ---- second association: still-tracked leftovers x low-score detections, thresh 0.5 (:221-250)
    const int nR = sm.misc[BM_NR], nRest = sm.misc[BM_NREST];
    byte_cost_rows(v, sm.rtr, nR, sm.lbox, nullptr, nLo, cost, sm.rmin, sm.rarg, [&](double iou_d, int) { return iou_d; });
    __syncthreads();
    if (wid == 0) {
        byte_assign(cost, nR, nLo, 0.5, sm, sm.xs, sm.born /*column flags unused*/, v.hdr);
        const int n0 = sm.misc[BM_NUPD];
        const int n1 = warp_compact(nR, sm.upd_slot, n0, [&](int t) { return sm.xs[t] >= 0; }, [&](int t) { return sm.rtr[t]; });
        warp_compact(nR, sm.upd_det, n0, [&](int t) { return sm.xs[t] >= 0; }, [&](int t) { return sm.lind[sm.xs[t]]; });
        for (int t = lane; t < nR; t += 32) if (sm.xs[t] >= 0) sm.tflag[sm.rtr[t]] |= TF_MATCHED;
        // mark_lost (:246-250), in u_track order
        const int nnl = warp_compact(nR, sm.nlost, 0, [&](int t) { return sm.xs[t] == -1; }, [&](int t) { return sm.rtr[t]; });
        for (int k = lane; k < nnl; k += 32) { const int slot = sm.nlost[k]; sm.tflag[slot] |= TF_NEWLOST; v.I(BI_STATE, slot) = BT_LOST; }
        if (lane == 0) { sm.misc[BM_NUPD] = n1; sm.misc[BM_N2] = n1; sm.misc[BM_NNL] = nnl; }
    }
    __syncthreads();

    // ---- unconfirmed tracks x remaining high-score detections, fuse_score, thresh 0.7 (:252-264)
    if (MODE == MODE_BYTE) {
        byte_cost_rows(v, sm.unc, nU, sm.hbox, sm.rest, nRest, cost, sm.rmin, sm.rarg,
                       [&](double iou_d, int d) { return fuse_entry(iou_d, (double)sm.hconf[d]); });
    } else {
        for (int r = wid; r < nU; r += BT_NT / 32) {
            const int slot = sm.unc[r];
            const Box a = v.box(slot);
            double m = CUDART_INF;
            int arg = 0x7fffffff;
            for (int c = lane; c < nRest; c += 32) {
                const int d = sm.rest[c];
                const double cc = bot_cost(1.0 - iou_sparse(a, det_xyxy(sm.hbox[d])), (double)sm.hconf[d], true, pr, embd,
                                           slot * Dcap + d);
                cost[(size_t)r * nRest + c] = cc;
                if (cc < m) { m = cc; arg = c; }
            }
            warp_argmin(m, arg);
            if (lane == 0) { sm.rmin[r] = m; sm.rarg[r] = arg; }
        }
    }
    __syncthreads();
    if (wid == 0) {
        byte_assign(cost, nU, nRest, 0.7, sm, sm.xs, sm.rflag, v.hdr);
        const int n0 = sm.misc[BM_NUPD];
        const int n1 = warp_compact(nU, sm.upd_slot, n0, [&](int t) { return sm.xs[t] >= 0; }, [&](int t) { return sm.unc[t]; });
        warp_compact(nU, sm.upd_det, n0, [&](int t) { return sm.xs[t] >= 0; }, [&](int t) { return sm.hind[sm.rest[sm.xs[t]]]; });
        for (int t = lane; t < nU; t += 32) {
            if (sm.xs[t] >= 0) sm.tflag[sm.unc[t]] |= TF_MATCHED;
            else if (sm.xs[t] == -1) v.I(BI_STATE, sm.unc[t]) = BT_REMOVED;      // mark_removed (:261-264)
        }
        // births (:266-272): leftover detections in index order, conf >= det_thresh
        const int nb = warp_compact(nRest, sm.born, 0,
                                    [&](int k) { return sm.rflag[k] == 0 && !((double)sm.hconf[sm.rest[k]] < pr.det_thresh); },
                                    [&](int k) { return sm.rest[k]; });
        // free slots = slots in neither list at the start of the frame
        const int nfree = warp_compact(Tcap, sm.rtr, 0, [&](int t) { return (sm.tflag[t] & TF_USED) == 0; }, [&](int t) { return t; });
        int nbc = nb;
        if (nb > nfree) { if (lane == 0) atomicOr(&v.hdr[BH_ERR], BM_ERR_TRACK_OVERFLOW); nbc = nfree; }
        if (lane == 0) { sm.misc[BM_NUPD] = n1; sm.misc[BM_NBORN] = nbc; }
    }
    __syncthreads();

    // ---- STrack.update / re_activate (:68-103) for every matched pair; activate (:56-66) for births
    const int nUpd = sm.misc[BM_NUPD], nBorn = sm.misc[BM_NBORN];
    for (int k = tid; k < nUpd + nBorn; k += BT_NT) {
        double x[8], P[XA_P], z[4];
        int slot, drow;
        if (k < nUpd) {
            slot = sm.upd_slot[k]; drow = sm.upd_det[k];
        } else {
            slot = sm.rtr[k - nUpd]; drow = sm.hind[sm.born[k - nUpd]];
        }
        const float* r = dsrc + drow * 6;
        det_measurement<KIND>(make_float4(r[0], r[1], r[2], r[3]), z);
        const bool dense = MODE == MODE_BOT && v.dP && k < nUpd && v.I(BI_DENSE, slot);
        if (dense) {
            botd_update(v, slot, z);
            for (int c = 0; c < 8; c++) x[c] = v.xP[c * Tcap + slot];
            v.I(BI_ACTIVATED, slot) = 1;
        } else if (k < nUpd) {
            v.load(slot, x, P);
            cv8_update<KIND>(x, P, z, 0.0);
            v.I(BI_ACTIVATED, slot) = 1;
        } else {
            cv8_initiate<KIND>(z, x, P);
            if (MODE == MODE_BOT) v.I(BI_DENSE, slot) = 0;
            v.I(BI_ID, slot) = id0 + (k - nUpd) + 1;
            v.I(BI_ACTIVATED, slot) = frame == 1 ? 1 : 0;
            v.I(BI_START, slot) = frame;
            v.I(BI_INREMOVED, slot) = 0;
            v.I(BI_NHIST, slot) = 0;
        }
        if (!dense) v.store(slot, x, P);
        v.set_box(slot, mean_to_xyxy<KIND>(x));
        v.I(BI_STATE, slot) = BT_TRACKED;
        v.I(BI_FRAME, slot) = frame;
        v.I(BI_DETIND, slot) = drow;
        v.confcls[slot] = (double)r[4];
        double cls = (double)r[5];
        if (MODE == MODE_BOT) {
            cls = bot_vote(st.chist + (size_t)s * 2 * BOT_HIST * Tcap, Tcap, slot, v.I(BI_NHIST, slot), cls, (double)r[4],
                           v.hdr);
            // feature EMA (botsort_track.py:40-49): first-round and unconfirmed matches, births; the
            // low-score detections of the second round carry no feature
            if (pr.with_reid && (k < sm.misc[BM_N1] || k >= sm.misc[BM_N2])) {
                const int q = atomicAdd(&sm.misc[BM_NFU], 1);
                int* fu = st.fupd + (size_t)s * 2 * Dcap;
                fu[q] = slot;
                fu[Dcap + q] = drow | (k >= nUpd ? 1 << 30 : 0);
            }
        }
        v.confcls[Tcap + slot] = cls;
    }
    if (MODE == MODE_BOT) {
        __syncthreads();
        if (tid == 0) v.hdr[BH_NFUPD] = sm.misc[BM_NFU];
    }
    __syncthreads();
    // ---- lost tracks age out (:273-276); the flag only takes effect one frame later (:283-285)
    for (int i = tid; i < nL; i += BT_NT) {
        const int slot = v.lost[i];
        if (frame - v.I(BI_FRAME, slot) > pr.max_time_lost) { v.I(BI_STATE, slot) = BT_REMOVED; sm.tflag[slot] |= TF_AGED; }
    }
    __syncthreads();

    // ---- new lists (:278-285)
    if (wid == 0) {
        int na = warp_compact(nA, sm.lista, 0, [&](int i) { return v.I(BI_STATE, v.act[i]) == BT_TRACKED; },
                              [&](int i) { return v.act[i]; });
        for (int k = lane; k < nBorn; k += 32) sm.lista[na + k] = sm.rtr[k];
        na += nBorn;
        const int nrf = sm.misc[BM_NRF];
        if (na + nrf > Tcap) { if (lane == 0) atomicOr(&v.hdr[BH_ERR], BM_ERR_TRACK_OVERFLOW); }
        for (int k = lane; k < nrf && na + k < Tcap; k += 32) sm.lista[na + k] = sm.refind[k];
        na = min(Tcap, na + nrf);
        int nl = warp_compact(nL, sm.listb, 0,
                              [&](int i) { const int sl = v.lost[i]; return !(sm.tflag[sl] & TF_REFIND) && !v.I(BI_INREMOVED, sl); },
                              [&](int i) { return v.lost[i]; });
        nl = warp_compact(sm.misc[BM_NNL], sm.listb, nl, [&](int k) { return !v.I(BI_INREMOVED, sm.nlost[k]); },
                          [&](int k) { return sm.nlost[k]; });
        __syncwarp();
        for (int i = lane; i < nL; i += 32) { const int sl = v.lost[i]; if (sm.tflag[sl] & TF_AGED) v.I(BI_INREMOVED, sl) = 1; }
        if (lane == 0) { sm.misc[BM_NA2] = na; sm.misc[BM_NL2] = nl; }
    }
    __syncthreads();
    // ---- remove_duplicate_stracks (:332-346): iou_distance < 0.15 between an active and a lost track
    const int nA2 = sm.misc[BM_NA2], nL2 = sm.misc[BM_NL2];
    for (int e = tid; e < nA2 * nL2; e += BT_NT) {
        const int a = sm.lista[e / nL2], b = sm.listb[e % nL2];
        if (1.0 - iou_sparse(v.box(a), v.box(b)) < 0.15) {
            const int tp = v.I(BI_FRAME, a) - v.I(BI_START, a), tq = v.I(BI_FRAME, b) - v.I(BI_START, b);
            atomicOr(&sm.tflag[tp > tq ? b : a], TF_DUP);
        }
    }
    __syncthreads();
    if (wid == 0) {
        const int na = warp_compact(nA2, v.act, 0, [&](int i) { return !(sm.tflag[sm.lista[i]] & TF_DUP); },
                                    [&](int i) { return sm.lista[i]; });
        const int nl = warp_compact(nL2, v.lost, 0, [&](int i) { return !(sm.tflag[sm.listb[i]] & TF_DUP); },
                                    [&](int i) { return sm.listb[i]; });
        // output rows (:287-301): activated tracks in list order, box from the Kalman mean
        const int nrows = warp_compact(na, sm.xs, 0, [&](int i) { return v.I(BI_ACTIVATED, v.act[i]) != 0; },
                                       [&](int i) { return v.act[i]; });
        const int out_rows = min(Tcap, Dcap);
        if (nrows > out_rows && lane == 0) atomicOr(&v.hdr[BH_ERR], BM_ERR_TRACK_OVERFLOW);
        if (lane == 0) {
            v.hdr[BH_NACT] = na; v.hdr[BH_NLOST] = nl; v.hdr[BH_FRAME] = frame; v.hdr[BH_NEXT_ID] = id0 + nBorn;
            v.hdr[BH_MATCH] += nUpd; v.hdr[BH_BIRTHS] += nBorn;
            if (sm.misc[BM_BAD]) atomicOr(&v.hdr[BH_ERR], 8);
            nout[s] = min(nrows, out_rows);
            sm.misc[BM_NR] = min(nrows, out_rows);
        }
    }
    __syncthreads();
    const int nrows = sm.misc[BM_NR];
    double* orow = out + (size_t)s * min(Tcap, Dcap) * 8;
    for (int k = tid; k < nrows; k += BT_NT) {
        const int slot = sm.xs[k];
        const Box b = v.box(slot);
        double* o = orow + (size_t)k * 8;
        o[0] = b.x1; o[1] = b.y1; o[2] = b.x2; o[3] = b.y2;
        o[4] = (double)v.I(BI_ID, slot);
        o[5] = v.confcls[slot]; o[6] = v.confcls[Tcap + slot];
        o[7] = (double)v.I(BI_DETIND, slot);
    }
}

// ============================================================================ BoT-SORT side kernels
// bot_embed_kernel: one CTA per stream, before the step.
//   (1) high-score split (same predicate and order as the step kernel),
//   (2) detection features as STrack.__init__ leaves them (botsort_track.py:36-48: normalised
//       twice, because smooth_feat aliases curr_feat on the first update_features call), fp32,
//   (3) embedding_distance (boxmot/utils/matching.py:230-321) of every live track (active and
//       lost lists) against every high-score detection: max(0, 1 - t.d / (|t| |d|)), a
//       [live x E] . [E x nH] contraction with fp32 FFMA tiles, stored as float by slot.
constexpr int BE_NT = 256, BE_TILE = 64, BE_K = 32;

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(BM_FULL, v, o);
    return v;
}

__global__ void __launch_bounds__(BE_NT)
bot_embed_kernel(ByteState st, ByteParams pr, const float* __restrict__ dets, const int* __restrict__ ndets,
                 const float* __restrict__ embs) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap, E = pr.emb_dim;
    float* At = reinterpret_cast<float*>(smem_raw);            // [BE_TILE][BE_K + 1]
    float* Bt = At + BE_TILE * (BE_K + 1);                     // [BE_TILE][BE_K + 1]
    float* tn = Bt + BE_TILE * (BE_K + 1);                     // [Tcap] |smooth_feat| by row
    float* dn = tn + Tcap;                                     // [Dcap] |feature| by high-score position
    int* rows = reinterpret_cast<int*>(dn + Dcap);             // [Tcap] live slots
    int* hind = rows + Tcap;                                   // [Dcap]
    __shared__ int s_nh;
    ByteView v(st, s);
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    const int nA = v.hdr[BH_NACT], nL = v.hdr[BH_NLOST], nLive = nA + nL;
    int nD_in = ndets[s];
    if (nD_in > Dcap) nD_in = Dcap;
    if (wid == 0) {
        const int nh = warp_compact(nD_in, hind, 0, [&](int i) { return (double)dsrc[i * 6 + 4] > pr.track_thresh; },
                                    [&](int i) { return i; });
        if (lane == 0) s_nh = nh;
    }
    for (int i = tid; i < nLive; i += BE_NT) rows[i] = i < nA ? v.act[i] : v.lost[i - nA];
    __syncthreads();
    const int nH = s_nh;
    int* g_live = st.elive + (size_t)s * Tcap;
    int* g_hind = st.ehind + (size_t)s * Dcap;
    for (int i = tid; i < nLive; i += BE_NT) g_live[i] = rows[i];
    for (int i = tid; i < nH; i += BE_NT) g_hind[i] = hind[i];
    if (tid == 0) { v.hdr[BH_NLIVE] = nLive; v.hdr[BH_NH] = nH; }
    const float* esrc = embs + (size_t)s * Dcap * E;
    float* dfeat = st.dfeat + (size_t)s * Dcap * E;
    const float* feat = st.feat + (size_t)s * Tcap * E;
    for (int d = wid; d < nH; d += BE_NT / 32) {
        const float* src = esrc + (size_t)hind[d] * E;
        float* dst = dfeat + (size_t)hind[d] * E;
        float ss = 0.f;
        for (int e = lane; e < E; e += 32) { const float x = src[e]; ss += x * x; }
        const float n1 = sqrtf(warp_sum_f(ss));
        ss = 0.f;
        for (int e = lane; e < E; e += 32) { const float x = src[e] / n1; ss += x * x; }
        const float n2 = sqrtf(warp_sum_f(ss));
        ss = 0.f;
        for (int e = lane; e < E; e += 32) { const float x = (src[e] / n1) / n2; dst[e] = x; ss += x * x; }
        ss = warp_sum_f(ss);
        if (lane == 0) { dn[d] = sqrtf(ss); st.edn[(size_t)s * Dcap + d] = dn[d]; }
    }
    for (int i = wid; i < nLive; i += BE_NT / 32) {
        const float* src = feat + (size_t)rows[i] * E;
        float ss = 0.f;
        for (int e = lane; e < E; e += 32) { const float x = src[e]; ss += x * x; }
        ss = warp_sum_f(ss);
        if (lane == 0) { tn[i] = sqrtf(ss); st.etn[(size_t)s * Tcap + i] = tn[i]; }
    }
    if (st.tensor_cores) return;          // the contraction runs on the tensor cores (cosine_tc_kernel)
    __syncthreads();
    float* embd = st.embd + (size_t)s * Tcap * Dcap;
    const int ty = tid / 16, tx = tid % 16;
    for (int r0 = 0; r0 < nLive; r0 += BE_TILE) {
        for (int c0 = 0; c0 < nH; c0 += BE_TILE) {
            float acc[4][4] = {};
            for (int k0 = 0; k0 < E; k0 += BE_K) {
                for (int q = tid; q < BE_TILE * BE_K; q += BE_NT) {
                    const int i = q / BE_K, kk = q % BE_K;
                    const bool kin = k0 + kk < E;
                    At[i * (BE_K + 1) + kk] = (r0 + i < nLive && kin) ? feat[(size_t)rows[r0 + i] * E + k0 + kk] : 0.f;
                    Bt[i * (BE_K + 1) + kk] = (c0 + i < nH && kin) ? dfeat[(size_t)hind[c0 + i] * E + k0 + kk] : 0.f;
                }
                __syncthreads();
#pragma unroll 8
                for (int kk = 0; kk < BE_K; kk++) {
                    float a[4], b[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) { a[i] = At[(ty * 4 + i) * (BE_K + 1) + kk]; b[i] = Bt[(tx * 4 + i) * (BE_K + 1) + kk]; }
#pragma unroll
                    for (int i = 0; i < 4; i++)
#pragma unroll
                        for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
                }
                __syncthreads();
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int r = r0 + ty * 4 + i, c = c0 + tx * 4 + j;
                    if (r < nLive && c < nH) {
                        const double cs = (double)acc[i][j] / ((double)tn[r] * (double)dn[c]);
                        embd[(size_t)rows[r] * Dcap + c] = (float)fmax(0.0, 1.0 - cs);
                    }
                }
        }
    }
}

static size_t bot_embed_smem(const ByteState& st) {
    return sizeof(float) * (2 * BE_TILE * (BE_K + 1) + st.Tcap + st.Dcap) + sizeof(int) * (st.Tcap + st.Dcap);
}

// bot_ema_kernel: STrack.update_features (botsort_track.py:40-49) for the frame's feature
// updates, one warp per (slot, detection): births copy the detection feature; matches blend
// smooth = 0.9 smooth + 0.1 (feat / |feat|) and renormalise, all in float32 as the reference.
__global__ void __launch_bounds__(256)
bot_ema_kernel(ByteState st, ByteParams pr) {
    const int s = blockIdx.x, lane = lane_id();
    const int q = blockIdx.y * 8 + warp_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap, E = pr.emb_dim;
    if (q >= st.hdr[(size_t)s * BH_COUNT + BH_NFUPD]) return;
    const int* fu = st.fupd + (size_t)s * 2 * Dcap;
    const int slot = fu[q], rw = fu[Dcap + q], r = rw & ((1 << 30) - 1);
    const float* src = st.dfeat + ((size_t)s * Dcap + r) * E;
    float* dst = st.feat + ((size_t)s * Tcap + slot) * E;
    if (rw >> 30) {
        for (int e = lane; e < E; e += 32) dst[e] = src[e];
        return;
    }
    float ss = 0.f;
    for (int e = lane; e < E; e += 32) { const float x = src[e]; ss += x * x; }
    const float n1 = sqrtf(warp_sum_f(ss));
    ss = 0.f;
    for (int e = lane; e < E; e += 32) { const float m = 0.9f * dst[e] + 0.1f * (src[e] / n1); ss += m * m; }
    const float n2 = sqrtf(warp_sum_f(ss));
    for (int e = lane; e < E; e += 32) dst[e] = (0.9f * dst[e] + 0.1f * (src[e] / n1)) / n2;
}

// One contraction per stream for cosine_tc_kernel
struct BotCosine {
    ByteState st;
    int E;
    __device__ CosineProblem operator()(int s) const {
        const int* hdr = st.hdr + (size_t)s * BH_COUNT;
        CosineProblem p;
        p.A = st.feat + (size_t)s * st.Tcap * E; p.ra = st.elive + (size_t)s * st.Tcap;
        p.na = st.etn + (size_t)s * st.Tcap; p.nA = hdr[BH_NLIVE];
        p.B = st.dfeat + (size_t)s * st.Dcap * E; p.rb = st.ehind + (size_t)s * st.Dcap;
        p.nb = st.edn + (size_t)s * st.Dcap; p.nB = hdr[BH_NH];
        p.E = E;
        p.out = st.embd + (size_t)s * st.Tcap * st.Dcap; p.out_ld = st.Dcap;
        return p;
    }
};

size_t bytetrack_step_smem(const ByteState& st) { return byte_smem_bytes(st.Tcap, st.Dcap); }

cudaError_t bytetrack_step_configure(size_t) {
    cudaError_t e = smem_optin(byte_step_kernel<MODE_BYTE>);
    if (e != cudaSuccess) return e;
    e = smem_optin(byte_step_kernel<MODE_BOT>);
    if (e != cudaSuccess) return e;
    e = smem_optin(bot_embed_kernel);
    if (e != cudaSuccess) return e;
    return smem_optin(cosine_tc_kernel<BotCosine>);
}

cudaError_t bytetrack_step_launch(const ByteState& st, const ByteParams& pr, const float* dets, const int* ndets,
                                  double* out, int* nout, cudaStream_t stream) {
    byte_step_kernel<MODE_BYTE><<<st.S, BT_NT, bytetrack_step_smem(st), stream>>>(st, pr, dets, ndets, nullptr, out, nout);
    return cudaGetLastError();
}

cudaError_t botsort_step_launch(const ByteState& st, const ByteParams& pr, const float* dets, const int* ndets,
                                const float* embs, const double* warps, double* out, int* nout, cudaStream_t stream) {
    if (pr.with_reid) {
        const size_t sm = bot_embed_smem(st);
        bot_embed_kernel<<<st.S, BE_NT, sm, stream>>>(st, pr, dets, ndets, embs);
        if (st.tensor_cores) {
            cosine_tc_kernel<BotCosine><<<dim3((st.Tcap + TC_M - 1) / TC_M, st.S), TC_NT, cosine_tc_smem_bytes(), stream>>>(
                BotCosine{st, pr.emb_dim});
        }
    }
    byte_step_kernel<MODE_BOT><<<st.S, BT_NT, bytetrack_step_smem(st), stream>>>(st, pr, dets, ndets, warps, out, nout);
    if (pr.with_reid) bot_ema_kernel<<<dim3(st.S, (st.Dcap + 7) / 8), 256, 0, stream>>>(st, pr);
    return cudaGetLastError();
}

}  // namespace bm
