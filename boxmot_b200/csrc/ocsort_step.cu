// OC-SORT frame step: FOUR launches advance every stream of the engine by one frame,
// with no host synchronisation in between (all lifecycle decisions are made on the device).
// Mirrors OcSort.update (boxmot/trackers/ocsort/ocsort.py:246-439) and enhanced_associate
// (boxmot/utils/association.py:377-533 with repairs R1-R3).
//
//   K1 ocs_predict_kernel   thread per live track        Kalman predict (ocsort.py:281-288,
//                           (HBM-streaming, SoA state)   :173-186) + association operands
//                                                        and the confidence split (:265-275)
//   K2 ocs_assoc_kernel     CTA (2 warps) per stream     exact row minima of the IoU + OCM cost
//                           (latency bound)              (association.py:427-461; fp32-screened,
//                                                        matrix never materialised), permutation
//                                                        fast path or exact LAP with on-demand
//                                                        cost rows, BYTE (:330-356), OCR (:358-386)
//   K3 ocs_update_kernel    thread per matched track     Kalman update + ORU (:321-324,
//                           (HBM-streaming)              xysr_kf.py:211-291)
//   K4 ocs_finalize_kernel  CTA per stream               misses, births, output rows, dead-track
//                                                        removal (:388-439)
//
// Splitting the step keeps the register-heavy fp64 Kalman code out of the association CTA,
// whose occupancy would otherwise be 2 CTAs/SM, and lets the state-streaming phases run with
// a thread per track over the whole GPU.  (A single fused launch with one 64-thread CTA per
// stream was measured 10-20 % slower at every stream count from 1 to 1024: 64 threads per
// stream cannot keep enough loads in flight for the predict / update / finalize phases.)
#include "ocsort_engine.cuh"
#include "lap_rect.cuh"
#include "cosine_tc.cuh"
#include <cstdio>
#include <cstdlib>

namespace bm {

constexpr int K1_NT = 128;
constexpr int K2_NT = 64;
constexpr int K3_NT = 64;
static_assert(K2_NT == 64, "OcmCostStaged's named barrier and the helper warp assume two warps per association CTA");

// ============================================================================ K1
// Blocks [0, S*chunks) predict; the trailing ceil(S/4) blocks split each stream's detections
// by confidence (ocsort.py:265-275; one warp per stream, stable order) for the kernels that
// follow.  dets: [S][Dcap][6] float32 (x1,y1,x2,y2,conf,cls) -- float32 IS the reference's
// precision: BaseTracker.setup_decorator round-trips every ndarray through float32
// (boxmot/trackers/basetracker.py:122-128).
// Confidence split of one stream by ONE warp (ocsort.py:265-275; stable order).
__device__ __forceinline__ void ocs_split_warp(const OcsState& st, const OcsParams& pr, int s,
                                               const float* __restrict__ dets, const int* __restrict__ ndets) {
    const int lane = lane_id(), Dcap = st.Dcap;
    int* hdr = st.hdr + (size_t)s * OH_COUNT;
    int nD_in = ndets[s];
    if (nD_in > Dcap) { if (lane == 0) atomicOr(&hdr[OH_ERR], BM_ERR_DET_OVERFLOW); nD_in = Dcap; }
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    int* hs = st.hsel + (size_t)s * 2 * Dcap;
    int nh = 0, ns = 0;
    for (int base = 0; base < nD_in; base += 128) {          // 4 independent loads in flight
        float cf[4];
#pragma unroll
        for (int k = 0; k < 4; k++) { const int i = base + 32 * k + lane; cf[k] = i < nD_in ? dsrc[i * 6 + 4] : -1.f; }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = base + 32 * k + lane;
            const double conf = (double)cf[k];
            const bool hi = i < nD_in && conf > pr.det_thresh;
            const bool se = pr.use_byte && i < nD_in && conf > pr.min_conf && conf < pr.det_thresh;
            const unsigned bh = __ballot_sync(BM_FULL, hi), bs = __ballot_sync(BM_FULL, se);
            const unsigned lt = (1u << lane) - 1;
            if (hi) hs[nh + __popc(bh & lt)] = i;
            if (se) hs[Dcap + ns + __popc(bs & lt)] = i;
            nh += __popc(bh); ns += __popc(bs);
        }
    }
    if (lane == 0) { hdr[OH_ND] = nh; hdr[OH_NS] = ns; }
}

// Predict of the track at list position p (ocsort.py:173-186, xysr_kf.py:137-175) and its
// association operands.
template <bool DENSE>
__device__ __forceinline__ void ocs_predict_track(const OcsState& st, const OcsParams& pr, const OcsView& v, int s, int p) {
    const int Tcap = st.Tcap;
    const XysrConst kc = {pr.q_xy, pr.q_s};
    const int slot = v.order[p];
    double x[7];
    if (DENSE && v.dP && (v.I(OI_FLAGS, slot) & OF_DENSE)) {
        ocs_dense_predict(v, slot, kc, x);
    } else {
        double P[13];
        load_xP(v.xP, Tcap, slot, x, P);
        if (x[6] + x[2] <= 0) x[6] *= 0.0;
        xysrb_predict(x, P, kc);
        store_xP(v.xP, Tcap, slot, x, P);
    }
    const int age = v.I(OI_AGE, slot) + 1;
    v.I(OI_AGE, slot) = age;
    const int tsu = v.I(OI_TSU, slot);
    if (tsu > 0) v.I(OI_STREAK, slot) = 0;
    v.I(OI_TSU, slot) = tsu + 1;
    double b[4];
    x_to_xyxy(x, b);      // a NaN box makes K2 drop the track (ocsort.py:284-288)
    double* ab = st.abuf + (size_t)s * AB_COUNT * Tcap;
#pragma unroll
    for (int c = 0; c < 4; c++) ab[(AB_X1 + c) * Tcap + p] = b[c];
    ab[AB_VY * Tcap + p] = v.vel[0 * Tcap + slot];
    ab[AB_VX * Tcap + p] = v.vel[1 * Tcap + slot];
    double kb[4];
    const bool ok = ocs_lookback(v, slot, age, pr.delta_t, kb);
    // placeholder observation is [-1,-1,-1,-1,-1] -> centre (-1,-1)
    ab[AB_KCX * Tcap + p] = ok ? (kb[0] + kb[2]) / 2.0 : -1.0;
    ab[AB_KCY * Tcap + p] = ok ? (kb[1] + kb[3]) / 2.0 : -1.0;
    ab[AB_KVALID * Tcap + p] = ok ? 1.0 : 0.0;
}

// DeepOCSORT CMC correction of one track, BEFORE its predict (deepocsort.py:191-208 +
// xysr_kf.py:111-135), for any 2x3 warp [m | t]:
//   * boxes: both corners through p -> m p + t;
//   * the reference keeps ONE numpy array as `last_observation`, `observations[age_last]` and
//     the detection row, so the newest observation is warped once through `last_observation`
//     (when its sum is > 0) and once more through the `observations` window when
//     age - age_last <= delta_t;
//   * the Kalman mean / covariance and, while frozen, the frozen copy are warped (x[:2], x[4:6],
//     P[:2,:2], P[4:6,4:6] -- not the cross blocks); the ORU anchor (the LIVE history's last
//     measurement, xysr_kf.py:184-192) is not.
// A warp without rotation / shear (m diagonal) keeps the covariance's block structure and is applied to
// the 13-entry block form with the reference's own products.  Any other warp couples x and y: the track
// switches to the dense 7x7 covariance (OF_DENSE, side array dP) for the rest of its life and every
// later predict / update / ORU replay of it runs the dense arithmetic (kf_xysr.cuh xysr_*).
__device__ __forceinline__ void docs_warp_box(double* b0, double* b1, double* b2, double* b3, const double* w6) {
    const double x1 = *b0, y1 = *b1, x2 = *b2, y2 = *b3;
    *b0 = (w6[0] * x1 + w6[1] * y1) + w6[2];
    *b1 = (w6[3] * x1 + w6[4] * y1) + w6[5];
    *b2 = (w6[0] * x2 + w6[1] * y2) + w6[2];
    *b3 = (w6[3] * x2 + w6[4] * y2) + w6[5];
}
// P[o:o+2, o:o+2] <- (m @ P2) @ m.T on a dense 7x7
__device__ __forceinline__ void docs_warp_block(double* P, int o, const double* w6) {
    const double p00 = P[o * 7 + o], p01 = P[o * 7 + o + 1], p10 = P[(o + 1) * 7 + o], p11 = P[(o + 1) * 7 + o + 1];
    const double m00 = w6[0], m01 = w6[1], m10 = w6[3], m11 = w6[4];
    const double t00 = m00 * p00 + m01 * p10, t01 = m00 * p01 + m01 * p11;
    const double t10 = m10 * p00 + m11 * p10, t11 = m10 * p01 + m11 * p11;
    P[o * 7 + o] = t00 * m00 + t01 * m01;
    P[o * 7 + o + 1] = t00 * m10 + t01 * m11;
    P[(o + 1) * 7 + o] = t10 * m00 + t11 * m01;
    P[(o + 1) * 7 + o + 1] = t10 * m10 + t11 * m11;
}
static __device__ __noinline__ void docs_affine_dense(const OcsView& v, int pass, int slot, const double* w6) {
    const int T = v.Tcap;
    double* b = pass ? v.sxP : v.xP;
    const double x0 = b[0 * T + slot], x1 = b[1 * T + slot], x4 = b[4 * T + slot], x5 = b[5 * T + slot];
    b[0 * T + slot] = (w6[0] * x0 + w6[1] * x1) + w6[2];
    b[1 * T + slot] = (w6[3] * x0 + w6[4] * x1) + w6[5];
    b[4 * T + slot] = w6[0] * x4 + w6[1] * x5;
    b[5 * T + slot] = w6[3] * x4 + w6[4] * x5;
    double P[49];
    load_dense(v, pass, slot, P);
    docs_warp_block(P, 0, w6);
    docs_warp_block(P, 4, w6);
    store_dense(v, pass, slot, P);
}
__device__ __forceinline__ void docs_affine_track(const OcsView& v, int slot, const double* w6, int delta_t) {
    const int T = v.Tcap;
    const double a = w6[0], tx = w6[2], d = w6[4], ty = w6[5];
    const bool rot = w6[1] != 0.0 || w6[3] != 0.0;
    if (rot && !v.dP) { atomicOr(&v.hdr[OH_ERR], 32); return; }
    const int age = v.I(OI_AGE, slot);
    if (v.I(OI_HITS, slot) > 0) {
        double lo[5];
#pragma unroll
        for (int c = 0; c < 5; c++) lo[c] = v.lastobs[c * T + slot];
        const double lsum = (((lo[0] + lo[1]) + lo[2]) + lo[3]) + lo[4];
        int times = lsum > 0 ? 1 : 0;
        if (v.I(OI_TSU, slot) <= delta_t) times++;          // age_last = age - time_since_update
        for (int k = 0; k < times; k++) {
            if (rot) docs_warp_box(&lo[0], &lo[1], &lo[2], &lo[3], w6);
            else {
                lo[0] = a * lo[0] + tx; lo[2] = a * lo[2] + tx;
                lo[1] = d * lo[1] + ty; lo[3] = d * lo[3] + ty;
            }
        }
#pragma unroll
        for (int c = 0; c < 4; c++) v.lastobs[c * T + slot] = lo[c];
        // the rest of the window; the newest observation's ring copy follows last_observation
        for (int r = 0; r < v.R; r++) {
            const int ra = v.ring_age[r * T + slot];
            if (ra < 0 || ra < age - delta_t) continue;
            if (ra == age - v.I(OI_TSU, slot)) {
#pragma unroll
                for (int c = 0; c < 4; c++) v.ring[(r * 4 + c) * T + slot] = lo[c];
            } else {
                double* q = v.ring + (size_t)(r * 4) * T + slot;
                if (rot) docs_warp_box(&q[0], &q[T], &q[2 * T], &q[3 * T], w6);
                else {
                    q[0] = a * q[0] + tx; q[2 * T] = a * q[2 * T] + tx;
                    q[T] = d * q[T] + ty; q[3 * T] = d * q[3 * T] + ty;
                }
            }
        }
    }
    int flags = v.I(OI_FLAGS, slot);
    if (rot && !(flags & OF_DENSE)) { ocs_make_dense(v, slot); flags |= OF_DENSE; }
    for (int pass = 0; pass < 2; pass++) {
        if (pass == 1 && !(!(flags & OF_OBSERVED) && (flags & OF_SAVED))) break;
        double* b = pass ? v.sxP : v.xP;
        if (flags & OF_DENSE) { docs_affine_dense(v, pass, slot, w6); continue; }
        b[0 * T + slot] = a * b[0 * T + slot] + tx;
        b[1 * T + slot] = d * b[1 * T + slot] + ty;
        b[4 * T + slot] = a * b[4 * T + slot];
        b[5 * T + slot] = d * b[5 * T + slot];
        // Pb = {P[b][b], P[b][b+4], P[b+4][b], P[b+4][b+4]}: only the two diagonal 2x2 blocks
        // P[:2,:2] and P[4:6,4:6] are warped by the reference (m P m^T), not the cross terms
        b[(7 + 0) * T + slot] = (a * b[(7 + 0) * T + slot]) * a;
        b[(7 + 3) * T + slot] = (a * b[(7 + 3) * T + slot]) * a;
        b[(7 + 4) * T + slot] = (d * b[(7 + 4) * T + slot]) * d;
        b[(7 + 7) * T + slot] = (d * b[(7 + 7) * T + slot]) * d;
    }
}

template <bool DEEP>
__global__ void __launch_bounds__(K1_NT)
ocs_predict_kernel(OcsState st, OcsParams pr, const float* __restrict__ dets, const int* __restrict__ ndets,
                   const double* __restrict__ warps) {
    const int chunks = (st.Tcap + K1_NT - 1) / K1_NT;
    if ((int)blockIdx.x >= st.S * chunks) {
        const int s = ((int)blockIdx.x - st.S * chunks) * (K1_NT / 32) + warp_id();
        if (s < st.S) ocs_split_warp(st, pr, s, dets, ndets);
        return;
    }
    const int s = blockIdx.x / chunks;
    const int p = (blockIdx.x % chunks) * K1_NT + threadIdx.x;
    OcsView v(st, s);
    const int nT = v.hdr[OH_NLIST];
    if (p == 0) st.n_updates[s] += (unsigned long long)nT;
    if (p < nT) {
        if (DEEP && warps) docs_affine_track(v, v.order[p], warps + (size_t)s * 6, pr.delta_t);
        ocs_predict_track<DEEP>(st, pr, v, s, p);
    }
}

// ============================================================================ K2
// One CTA of two warps per stream; ~23 KB of shared memory and <= 128 registers keep eight
// CTAs resident per SM, i.e. ALL streams of the headline workload in a single wave: the
// kernel is a chain of short dependent phases, so throughput comes from co-residency.
struct OcsSmem {
    float4 *dbox, *sbox;                 // [Dcap] high-score / BYTE detections (inputs are float32: exact)
    double* rmin;                        // [Dcap]
    void* work;                          // screening tables, later the solver's work area
    int *dind, *sind, *d2t, *um_d, *rowcnt, *hit_t, *rarg;   // [Dcap]
    int *slot_of, *t2d, *um_t, *colcnt;                     // [Tcap]
    int *misc, *cand;
};

enum { M_NT = 0, M_NUD, M_NUT, M_FAST, M_ANYBAD, M_NUPD, M_LAPBAD, M_ROW, M_COUNT = 8 };

// fp32 screening tables, built once per stream-frame (they alias the solver's work area,
// which is idle while the row summaries are computed).
//   boxf[t]  the track box rounded OUTWARDS to float (x1,y1 down; x2,y2 up): if the float
//            boxes do not overlap, the fp64 boxes do not either.  A degenerate / non-finite
//            box is stored as (-inf,-inf,+inf,+inf) so it is always evaluated exactly.
//   kc[t]    centre of the observation delta_t frames back as (x_hi, x_lo, y_hi, y_lo) float
//            pairs: (cx_hi - x_hi) + (cx_lo - x_lo) is the fp64 difference to ~1e-7 relative
//   mvf[t]   (vx, vy) as floats, (0,0) when the OCM term is exactly 0 for this track
//   cs       [warps][Tcap] per-row screening cosines
struct OcmScreen {
    float4* boxf;
    float4* kc;
    float2* mvf;
    float* cs;
};
__host__ __device__ inline size_t ocm_screen_bytes(int Tcap, int warps) {
    return (size_t)Tcap * (sizeof(float4) + sizeof(float4) + sizeof(float2)) + (size_t)warps * Tcap * sizeof(float);
}
__host__ __device__ inline size_t ocs_work_bytes(int Tcap, int Dcap) {
    // the lapjv emulation only runs with more detections than tracks: n = nD + nT < 2 * Dcap
    const size_t a = ocm_screen_bytes(Tcap, K2_NT / 32), b = rect_work_bytes(Dcap, Tcap), c = lap_work_bytes(2 * Dcap);
    const size_t m = a > b ? (a > c ? a : c) : (b > c ? b : c);
    return (m + 15) / 16 * 16;
}

// per-warp candidate scratch: [Tcap + 64] packed candidates (ocm_scan_rows) or
// LANE_CAND results (8 B) + packed candidates (4 B) (ocm_scan_rows_lane)
constexpr int LANE_CAND = 128;
__host__ __device__ inline size_t ocs_cand_ints(int Tcap) {
    const size_t a = (size_t)Tcap + 64, b = 3 * LANE_CAND;
    return ((a > b ? a : b) + 3) / 4 * 4;
}
__host__ __device__ inline size_t ocs_smem_bytes(int Tcap, int Dcap, int use_byte) {
    const size_t d = (size_t)Dcap, t = (size_t)Tcap;
    size_t b = 16 * d * (use_byte ? 2 : 1) + 8 * d + ocs_work_bytes(Tcap, Dcap);
    b += 4 * ((use_byte ? 7 : 6) * d + 4 * t + M_COUNT + ocs_cand_ints(Tcap) * (K2_NT / 32));
    return (b + 15) / 16 * 16;
}

__device__ inline OcsSmem ocs_carve(unsigned char* base, int Tcap, int Dcap, int use_byte) {
    OcsSmem m;
    m.dbox = reinterpret_cast<float4*>(base); base += 16 * (size_t)Dcap;
    if (use_byte) { m.sbox = reinterpret_cast<float4*>(base); base += 16 * (size_t)Dcap; } else m.sbox = nullptr;
    m.work = base; base += ocs_work_bytes(Tcap, Dcap);
    m.cand = reinterpret_cast<int*>(base); base += 4 * ocs_cand_ints(Tcap) * (K2_NT / 32);
    m.rmin = reinterpret_cast<double*>(base); base += 8 * (size_t)Dcap;
    int* q = reinterpret_cast<int*>(base);
    m.dind = q; q += Dcap;
    if (use_byte) { m.sind = q; q += Dcap; } else m.sind = nullptr;
    m.d2t = q; q += Dcap;
    m.um_d = q; q += Dcap;
    m.rowcnt = q; q += Dcap;
    m.hit_t = q; q += Dcap;
    m.rarg = q; q += Dcap;
    m.slot_of = q; q += Tcap;
    m.t2d = q; q += Tcap;
    m.um_t = q; q += Tcap;
    m.colcnt = q; q += Tcap;
    m.misc = q;
    return m;
}

// arccos with the three exact landmarks pinned (the reference's structural zeros in
// the OCM cost come from arccos(0) == pi/2 exactly).
__device__ __forceinline__ double acos_pinned(double c) {
    if (c == 0.0) return CUDART_PI / 2.0;
    if (c >= 1.0) return 0.0;
    if (c <= -1.0) return CUDART_PI;
    return acos(c);
}

// ---------------------------------------------------------------------------------------
// First-round cost, evaluated ON DEMAND.  cost(d, t) = -(iou + OCM) exactly as
// enhanced_associate builds it (association.py:427-461), det-major.  The assignment solver
// needs (a) every row's minimum and (b) the full rows it sweeps while augmenting -- on
// tracking data that is ~2 rows per frame -- so the matrix is never materialised.  The
// fp64 operands stay where the predict kernel wrote them (L2-resident global memory); only
// the fp32 screening tables live in shared memory.
struct OcmCost {
    // bounded sweeps (lower()) measured slower than plain exact sweeps on B200: two extra
    // passes of L2 loads per sweep cost more than the ~140 exact entries they avoid
    static constexpr bool kBounded = false, kStaged = false;
    const float4* dbox;                // high-score detections (shared memory)
    const double* tb; int Tcap;        // operands per track column [AB_COUNT][Tcap] (global, L2)
    int func; double diag, inertia;

    // velocity-direction term; exactly 0 for a track that was never observed (valid = 0)
    // or has no velocity yet (vx = vy = 0 -> cos = 0 -> arccos = pi/2 exactly)
    __device__ __forceinline__ double motion(double cxd, double cyd, int t) const {
        // all operands are fetched up front (one L2 round trip instead of three dependent ones)
        const double vx = tb[AB_VX * Tcap + t], vy = tb[AB_VY * Tcap + t], valid = tb[AB_KVALID * Tcap + t];
        const double kx = tb[AB_KCX * Tcap + t], ky = tb[AB_KCY * Tcap + t];
        if (valid == 0.0 || (vx == 0.0 && vy == 0.0)) return 0.0;
        // speed_direction_batch(dets, previous_obs)  (association.py:10-20)
        double dx = cxd - kx, dy = cyd - ky;
        const double norm = sqrt(dx * dx + dy * dy) + 1e-6;
        dx = dx / norm; dy = dy / norm;
        double cs = vx * dx + vy * dy;
        cs = fmin(1.0, fmax(-1.0, cs));
        const double ang = (CUDART_PI / 2.0 - fabs(acos_pinned(cs))) / CUDART_PI;
        return (1.0 * ang) * inertia;
    }
    __device__ __forceinline__ Box det(int d) const {
        const float4 v = dbox[d];
        return Box{(double)v.x, (double)v.y, (double)v.z, (double)v.w};
    }
    __device__ __forceinline__ Box trk(int t) const {
        return Box{tb[AB_X1 * Tcap + t], tb[AB_Y1 * Tcap + t], tb[AB_X2 * Tcap + t], tb[AB_Y2 * Tcap + t]};
    }
    __device__ __forceinline__ double entry(const Box& a, double cxd, double cyd, int t, double* iou_out) const {
        const Box b = trk(t);
        const double m = motion(cxd, cyd, t);        // issues its loads before the IoU arithmetic
        const double iou = asso_pair(func, a, b, diag);
        *iou_out = iou;
        return -(iou + m);
    }
    __device__ __forceinline__ double operator()(int d, int t) const {
        const Box a = det(d);
        double iou;
        return entry(a, (a.x1 + a.x2) / 2.0, (a.y1 + a.y2) / 2.0, t, &iou);
    }
    // Cheap rigorous lower bound of cost(d, t) for the solver's sweeps (*exact = true when the
    // returned value IS the entry).  asso_func = iou only: boxes that do not overlap give
    // cost = -motion exactly, and motion <= inertia * asin(cos + eps) / pi with cos taken in
    // fp32 (error << eps); everything else is evaluated exactly.
    __device__ __forceinline__ double lower(int d, int t, bool* exact) const {
        const Box a = det(d), b = trk(t);
        const double w = fmin(a.x2, b.x2) - fmax(a.x1, b.x1), h = fmin(a.y2, b.y2) - fmax(a.y1, b.y1);
        const double den = (a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1);
        const double cxd = (a.x1 + a.x2) / 2.0, cyd = (a.y1 + a.y2) / 2.0;
        if (func != ASSO_IOU || !((w <= 0.0 || h <= 0.0) && den > 0.0)) {
            double iou;
            *exact = true;
            return entry(a, cxd, cyd, t, &iou);
        }
        const double vx = tb[AB_VX * Tcap + t], vy = tb[AB_VY * Tcap + t];
        if (tb[AB_KVALID * Tcap + t] == 0.0 || (vx == 0.0 && vy == 0.0)) { *exact = true; return -0.0; }
        *exact = false;
        const float dx = (float)(cxd - tb[AB_KCX * Tcap + t]), dy = (float)(cyd - tb[AB_KCY * Tcap + t]);
        const float n2 = dx * dx + dy * dy;
        const float nrm = n2 * rsqrtf(fmaxf(n2, 1e-30f));
        const float cs = __fdividef((float)vx * dx + (float)vy * dy, nrm + 1e-6f);
        const double ub = (double)asinf(fminf(1.f, cs + 2e-5f)) + 4e-6;  // >= asin(true cos)
        return -(ub * (1.0 / CUDART_PI) * inertia + 1e-12);
    }
};

// The same cost for the solver's sweeps with BOTH warps of the CTA evaluating the swept row
// into a shared-memory buffer (the fp64 entry is a long dependent chain, so a sweep's time is
// its chunk count): warp 0 runs the solver and calls stage(); warp 1 sits in helper() until
// release().  Named barrier 1 (64 threads) separates "row published" / "row evaluated".
struct OcmCostStaged {
    static constexpr bool kBounded = false, kStaged = true;
    OcmCost base;
    double* rowc;          // [nc]
    int* cur;              // row being swept, -1 = solver finished
    const double* spc;     // solver's distance array (column scanned <=> -inf)
    int nc;
    static __device__ __forceinline__ void bar() { asm volatile("bar.sync 1, 64;" ::: "memory"); }
    __device__ __forceinline__ void fill(int i, int w) const {
        const Box a = base.det(i);
        const double cx = (a.x1 + a.x2) / 2.0, cy = (a.y1 + a.y2) / 2.0;
        for (int j = 32 * w + lane_id(); j < nc; j += 64) {
            if (spc[j] == -CUDART_INF) continue;
            double iou;
            rowc[j] = base.entry(a, cx, cy, j, &iou);
        }
    }
    __device__ __forceinline__ void stage(int i, const double*) const {
        __syncwarp();
        if (lane_id() == 0) *cur = i;
        bar();
        fill(i, 0);
        bar();
    }
    __device__ __forceinline__ void release() const {
        __syncwarp();
        if (lane_id() == 0) *cur = -1;
        bar();
    }
    __device__ __forceinline__ void helper() const {
        for (;;) {
            bar();
            const int i = *(volatile int*)cur;
            if (i < 0) break;
            fill(i, 1);
            bar();
        }
    }
    __device__ __forceinline__ double operator()(int, int j) const { return rowc[j]; }
    __device__ __forceinline__ double lower(int i, int j, bool* exact) const { *exact = true; return rowc[j]; }
};

// Returns true when the predicted box is NaN (the reference drops such a track).
__device__ inline bool ocm_screen_fill(const OcmCost& c, const OcmScreen& sc, int t) {
    const Box b = c.trk(t);
    const bool ok = (b.x2 - b.x1) * (b.y2 - b.y1) > 0.0 && b.x2 > b.x1;
    sc.boxf[t] = ok ? make_float4(__double2float_rd(b.x1), __double2float_rd(b.y1),
                                  __double2float_ru(b.x2), __double2float_ru(b.y2))
                    : make_float4(-CUDART_INF_F, -CUDART_INF_F, CUDART_INF_F, CUDART_INF_F);
    const double vx = c.tb[AB_VX * c.Tcap + t], vy = c.tb[AB_VY * c.Tcap + t];
    const bool mv = c.tb[AB_KVALID * c.Tcap + t] != 0.0 && (vx != 0.0 || vy != 0.0);
    float fx = mv ? (float)vx : 0.f, fy = mv ? (float)vy : 0.f;
    if (mv && fx == 0.f && fy == 0.f) fx = 1e-30f;      // keep "moving" distinguishable
    sc.mvf[t] = make_float2(fx, fy);
    const double kx = c.tb[AB_KCX * c.Tcap + t], ky = c.tb[AB_KCY * c.Tcap + t];
    const float kxh = (float)kx, kyh = (float)ky;
    sc.kc[t] = make_float4(kxh, (float)(kx - (double)kxh), kyh, (float)(ky - (double)kyh));
    return b.x1 != b.x1 || b.y1 != b.y1 || b.x2 != b.x2 || b.y2 != b.y2;
}
// fp32 cosine between the track's velocity (mv) and the direction from its look-back
// centre (kc, hi/lo split) to the detection centre (association.py:10-20, :437-438).
__device__ __forceinline__ float ocm_screen_cos(const float4 kc, const float2 mv, float cxh, float cxl,
                                                float cyh, float cyl) {
    const float dx = (cxh - kc.x) + (cxl - kc.y), dy = (cyh - kc.z) + (cyl - kc.w);
    const float n2 = fmaf(dy, dy, dx * dx);
    const float nrm = n2 * rsqrtf(fmaxf(n2, 1e-30f));
    return __fdividef(fmaf(mv.y, dy, mv.x * dx), nrm + 1e-6f);
}
constexpr float OCM_BAND = 1e-4f;   // >> error of the fp32 cosine (~1e-6), << spacing of candidates
constexpr float OCM_NOCAND = -4.f;

// Row summaries for the solver and the permutation fast path: exact minimum / argmin of
// every row, number of IoU hits (> thr) and the last hit column.  Each warp scans its rows
// (d = warp, warp + W, ...) one after the other.
// With asso_func = iou an entry whose boxes do not overlap is exactly -(0 + motion), and
// motion is a monotone function of the OCM cosine, so only (i) entries whose outward-rounded
// float boxes overlap and (ii) the remaining entries whose fp32 cosine is within OCM_BAND of
// the row's best are evaluated in fp64 (the rest provably cannot be the minimum); entries
// whose motion is exactly 0 compete as -0.0.  Other association functions, and detections
// with a degenerate box, evaluate every entry.  The exact evaluations are batched ACROSS rows
// (one lane per candidate, segmented arg-min per row), flushed at row boundaries.
// Outputs (shared memory, pre-set by the caller to +inf / INT_MAX / 0 / -1): rmin, rarg,
// rowcnt, hit_t; colcnt accumulates the per-column hits; zero_t is scratch [nD].
__device__ inline void ocm_scan_rows(const OcmCost& c, const OcmScreen& sc, int nD, int nT, double iou_thr,
                                     int* cand /*[Tcap + 64] per warp*/, int* colcnt, double* rmin, int* rarg,
                                     int* rowcnt, int* hit_t, int* zero_ts) {
    const int lane = lane_id(), wid = warp_id(), NW = blockDim.x >> 5;
    float* csrow = sc.cs + (size_t)wid * c.Tcap;
    int ncand = 0;
    auto flush = [&]() {
        __syncwarp();
        for (int k0 = 0; k0 < ncand; k0 += 32) {
            const int k = k0 + lane;
            const bool valid = k < ncand;
            const int pk = valid ? cand[k] : 0;
            const int d = pk >> 16;
            int t = pk & 0xffff;
            double cc = CUDART_INF;
            if (valid) {
                const Box a = c.det(d);
                double iou;
                cc = c.entry(a, (a.x1 + a.x2) / 2.0, (a.y1 + a.y2) / 2.0, t, &iou);
                if (iou > iou_thr) { atomicAdd(&rowcnt[d], 1); atomicMax(&hit_t[d], t); atomicAdd(&colcnt[t], 1); }
            }
            const int row = valid ? d : -1 - lane;           // rows are contiguous in the buffer
            const int prev = __shfl_up_sync(BM_FULL, row, 1);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const double occ = __shfl_down_sync(BM_FULL, cc, o);
                const int ot = __shfl_down_sync(BM_FULL, t, o), orow = __shfl_down_sync(BM_FULL, row, o);
                if (lane + o < 32 && orow == row && (occ < cc || (occ == cc && ot < t))) { cc = occ; t = ot; }
            }
            if (valid && (lane == 0 || prev != row)) {       // segment head merges into the row state
                const double cur = rmin[d];
                if (cc < cur || (cc == cur && t < rarg[d])) { rmin[d] = cc; rarg[d] = t; }
            }
            __syncwarp();
        }
        ncand = 0;
    };
    auto push = [&](bool pred, int d, int t) {
        const unsigned b = __ballot_sync(BM_FULL, pred);
        if (pred) cand[ncand + __popc(b & ((1u << lane) - 1))] = (d << 16) | t;
        ncand += __popc(b);
    };
    for (int d = wid; d < nD; d += NW) {
        const float4 af = c.dbox[d];
        const double cxd = ((double)af.x + (double)af.z) / 2.0, cyd = ((double)af.y + (double)af.w) / 2.0;
        const float cxh = (float)cxd, cyh = (float)cyd;
        const float cxl = (float)(cxd - (double)cxh), cyl = (float)(cyd - (double)cyh);
        const bool sparse = c.func == ASSO_IOU &&
                            ((double)af.z - (double)af.x) * ((double)af.w - (double)af.y) > 0.0 && af.z > af.x;
        int zero_t = 0x7fffffff;
        float m32 = -CUDART_INF_F;
        for (int base = 0; base < nT; base += 128) {         // 4 independent chunks in flight
            bool ex[4];
            float csv[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int t = base + 32 * u + lane;
                ex[u] = false;
                csv[u] = OCM_NOCAND;
                if (t < nT) {
                    ex[u] = true;
                    if (sparse) {
                        const float4 b = sc.boxf[t];
                        ex[u] = af.x < b.z && b.x < af.z && af.y < b.w && b.y < af.w;
                        if (!ex[u]) {
                            const float2 mv = sc.mvf[t];
                            if (mv.x == 0.f && mv.y == 0.f) zero_t = min(zero_t, t);
                            else {
                                csv[u] = ocm_screen_cos(sc.kc[t], mv, cxh, cxl, cyh, cyl);
                                m32 = fmaxf(m32, csv[u]);
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int t = base + 32 * u + lane;
                if (base + 32 * u < nT) {                    // warp-uniform
                    if (t < nT) csrow[t] = csv[u];
                    push(ex[u], d, t);
                }
            }
        }
        if (sparse) {
            for (int o = 16; o > 0; o >>= 1) {
                m32 = fmaxf(m32, __shfl_xor_sync(BM_FULL, m32, o));
                zero_t = min(zero_t, __shfl_xor_sync(BM_FULL, zero_t, o));
            }
            const float lim = m32 - OCM_BAND;
            __syncwarp();
            for (int base = 0; base < nT; base += 32) {
                const int t = base + lane;
                const float v = t < nT ? csrow[t] : OCM_NOCAND;
                push(v >= lim && v != OCM_NOCAND, d, t);
            }
        }
        if (lane == 0) zero_ts[d] = zero_t;
        if (ncand >= 32) flush();
    }
    flush();
    // zero-motion, zero-IoU entries hold -(0.0 + 0.0) = -0.0
    for (int d = wid + lane * NW; d < nD; d += 32 * NW) {
        const int zt = zero_ts[d];
        const double rm = rmin[d];
        if (zt != 0x7fffffff && (-0.0 < rm || (-0.0 == rm && zt < rarg[d]))) { rmin[d] = -0.0; rarg[d] = zt; }
    }
}

// asso_func = iou: the same row summaries with ONE LANE PER ROW.  A lane walks all tracks of
// its detection (the screening tables are read as shared-memory broadcasts, so the walk has
// no shuffles, ballots or list traffic and pipelines freely) and keeps in registers
//   * the tracks whose outward-rounded float box overlaps the detection (up to OVL_CAP
//     10-bit columns packed in two 64-bit words),
//   * a 6-deep shift register of the columns whose fp32 cosine came within OCM_BAND of the
//     RUNNING maximum (a superset of those within the band of the final maximum; the fp32
//     value is recomputed for the survivors),
//   * the first column whose entry is exactly -0.0.
// Afterwards every lane evaluates its few candidates in fp64 with the reference's operation
// order.  Rows that overflow a register list, lose a possibly-in-band column out of the shift
// register, or have a degenerate detection box are redone by the whole warp with every entry
// evaluated exactly.  Results are identical to ocm_scan_rows (exact minima, lowest column on
// ties).
constexpr int OVL_CAP = 12;
// The fp32 screen of one row (detection box af) against all nT tracks: one lane, no cross-lane traffic.
// Phase 1 lists the columns whose outward-rounded float box overlaps the detection and bounds their
// IoU in fp32.  The OCM term lies in [-|inertia|/2, |inertia|/2] and a disjoint pair has IoU = 0
// exactly, so once some column's IoU exceeds |inertia| (+0.01: ~1e3 x the fp32 error of the bound)
// no disjoint column can hold the row minimum: such a "strong" row needs no phase 2.
__device__ __forceinline__ bool ocm_lane_overlaps(const float4 af, const OcmScreen& sc, int nT, float strong_thr,
                                                  ScanRow& r) {
    bool dense = !(((double)af.z - (double)af.x) * ((double)af.w - (double)af.y) > 0.0 && af.z > af.x);
    const float aa = (af.z - af.x) * (af.w - af.y);
    unsigned long long ov0 = 0, ov1 = 0;
    int nov = 0;
    float best = 0.f;
#pragma unroll 4
    for (int t = 0; t < nT; t++) {
        const float4 b = sc.boxf[t];
        if (af.x < b.z && b.x < af.z && af.y < b.w && b.y < af.w) {
            if (nov < 6) ov0 |= (unsigned long long)t << (10 * nov);
            else if (nov < OVL_CAP) ov1 |= (unsigned long long)t << (10 * (nov - 6));
            nov++;
            const float inter = (fminf(af.z, b.z) - fmaxf(af.x, b.x)) * (fminf(af.w, b.w) - fmaxf(af.y, b.y));
            best = fmaxf(best, __fdividef(inter, aa + (b.z - b.x) * (b.w - b.y) - inter));
        }
    }
    if (nov > OVL_CAP) dense = true;
    r.ov0 = ov0; r.ov1 = ov1; r.bf = 0; r.nov = nov; r.nbf = 0; r.zero_t = 0x7fffffff; r.dense = dense ? 1 : 0;
    return best > strong_thr;
}
// Phase 2 (rows without a strong overlap): the disjoint columns whose fp32 OCM cosine is within
// OCM_BAND of the row's best, and the lowest zero-motion column.
__device__ __forceinline__ void ocm_lane_band(const float4 af, const OcmScreen& sc, int nT, ScanRow& r) {
    const double cxd = ((double)af.x + (double)af.z) / 2.0, cyd = ((double)af.y + (double)af.w) / 2.0;
    const float cxh = (float)cxd, cyh = (float)cyd;
    const float cxl = (float)(cxd - (double)cxh), cyl = (float)(cyd - (double)cyh);
    unsigned long long bl = 0;
    int nb = 0, zero_t = 0x7fffffff;
    float runmax = -CUDART_INF_F;
#pragma unroll 4
    for (int t = 0; t < nT; t++) {
        const float4 b = sc.boxf[t];
        const float2 mv = sc.mvf[t];
        const float4 kc = sc.kc[t];
        if (af.x < b.z && b.x < af.z && af.y < b.w && b.y < af.w) {
            // listed by phase 1
        } else if (mv.x == 0.f && mv.y == 0.f) {
            zero_t = min(zero_t, t);
        } else {
            const float cs = ocm_screen_cos(kc, mv, cxh, cxl, cyh, cyl);
            if (cs >= runmax - OCM_BAND) {
                bl = (bl << 10) | (unsigned long long)t;
                nb++;
                runmax = fmaxf(runmax, cs);
            }
        }
    }
    const float lim = runmax - OCM_BAND;
    const int nbk = min(nb, 6);
    if (nb > 6) {
        // the oldest retained column bounds (within one band) everything shifted out before it
        const int t5 = (int)((bl >> 50) & 1023u);
        if (ocm_screen_cos(sc.kc[t5], sc.mvf[t5], cxh, cxl, cyh, cyl) >= lim - OCM_BAND) r.dense = 1;
    }
    // survivors of the band filter (fp32 value against the FINAL maximum)
    unsigned long long bf = 0;
    int nbf = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        if (k < nbk) {
            const int t = (int)((bl >> (10 * k)) & 1023u);
            if (ocm_screen_cos(sc.kc[t], sc.mvf[t], cxh, cxl, cyh, cyl) >= lim) {
                bf |= (unsigned long long)t << (10 * nbf);
                nbf++;
            }
        }
    }
    r.bf = bf; r.nbf = nbf; r.zero_t = zero_t;
}
__device__ __forceinline__ float ocm_strong_thr(double inertia) { return (float)fabs(inertia) + 0.01f; }
// Both phases in ONE walk over the tracks, for the association CTA's own screen: there the 32 rows of
// a warp are walked together, so a second, divergent walk for the few weak rows would cost every lane.
__device__ __forceinline__ ScanRow ocm_lane_screen(const float4 af, const OcmScreen& sc, int nT) {
    const double cxd = ((double)af.x + (double)af.z) / 2.0, cyd = ((double)af.y + (double)af.w) / 2.0;
    const float cxh = (float)cxd, cyh = (float)cyd;
    const float cxl = (float)(cxd - (double)cxh), cyl = (float)(cyd - (double)cyh);
    ScanRow r;
    r.dense = !(((double)af.z - (double)af.x) * ((double)af.w - (double)af.y) > 0.0 && af.z > af.x) ? 1 : 0;
    unsigned long long ov0 = 0, ov1 = 0, bl = 0;
    int nov = 0, nb = 0, zero_t = 0x7fffffff;
    float runmax = -CUDART_INF_F;
#pragma unroll 4
    for (int t = 0; t < nT; t++) {
        const float4 b = sc.boxf[t];
        const float2 mv = sc.mvf[t];
        const float4 kc = sc.kc[t];
        if (af.x < b.z && b.x < af.z && af.y < b.w && b.y < af.w) {
            if (nov < 6) ov0 |= (unsigned long long)t << (10 * nov);
            else if (nov < OVL_CAP) ov1 |= (unsigned long long)t << (10 * (nov - 6));
            nov++;
        } else if (mv.x == 0.f && mv.y == 0.f) {
            zero_t = min(zero_t, t);
        } else {
            const float cs = ocm_screen_cos(kc, mv, cxh, cxl, cyh, cyl);
            if (cs >= runmax - OCM_BAND) {
                bl = (bl << 10) | (unsigned long long)t;
                nb++;
                runmax = fmaxf(runmax, cs);
            }
        }
    }
    if (nov > OVL_CAP) r.dense = 1;
    const float lim = runmax - OCM_BAND;
    const int nbk = min(nb, 6);
    if (nb > 6) {
        const int t5 = (int)((bl >> 50) & 1023u);
        if (ocm_screen_cos(sc.kc[t5], sc.mvf[t5], cxh, cxl, cyh, cyl) >= lim - OCM_BAND) r.dense = 1;
    }
    unsigned long long bf = 0;
    int nbf = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        if (k < nbk) {
            const int t = (int)((bl >> (10 * k)) & 1023u);
            if (ocm_screen_cos(sc.kc[t], sc.mvf[t], cxh, cxl, cyh, cyl) >= lim) {
                bf |= (unsigned long long)t << (10 * nbf);
                nbf++;
            }
        }
    }
    r.ov0 = ov0; r.ov1 = ov1; r.bf = bf; r.nov = nov; r.nbf = nbf; r.zero_t = zero_t;
    return r;
}

// pre: rows screened by ocs_screen_kernel (nullptr = screen here, with the CTA's own tables)
__device__ inline void ocm_scan_rows_lane(const OcmCost& c, const OcmScreen& sc, int nD, int nT, double iou_thr,
                                          int* colcnt, double* rmin, int* rarg, int* rowcnt, int* hit_t,
                                          int* scratch /*[3 * LANE_CAND] per warp, 8-byte aligned*/,
                                          const ScanRow* __restrict__ pre = nullptr) {
    const int lane = lane_id(), wid = warp_id(), NW = blockDim.x >> 5;
    double* cres = reinterpret_cast<double*>(scratch);
    int* clist = scratch + 2 * LANE_CAND;
    for (int r0 = 32 * wid; r0 < nD; r0 += 32 * NW) {
        const int d = r0 + lane;
        const bool live = d < nD;
        const ScanRow sr = pre ? pre[live ? d : r0] : ocm_lane_screen(c.dbox[live ? d : r0], sc, nT);
        const unsigned long long ov0 = sr.ov0, ov1 = sr.ov1, bf = sr.bf;
        const int nov = sr.nov, nbf = sr.nbf, zero_t = sr.zero_t;
        const bool dense = sr.dense != 0;
        // ---- exact evaluation: the warp's candidates are pooled in shared memory so that all
        // 32 lanes evaluate (the fp64 entry is a long dependent chain), then every lane folds
        // the results of its own row; LANE_CAND candidates per round
        const int ntot = (live && !dense) ? nov + nbf : 0;
        double best = CUDART_INF;
        int bt = 0x7fffffff, cnt = 0, ht = -1, done = 0;
        for (;;) {
            const int rem = ntot - done;
            int incl = rem;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int up = __shfl_up_sync(BM_FULL, incl, o);
                if (lane >= o) incl += up;
            }
            const int total = __shfl_sync(BM_FULL, incl, 31);
            if (total == 0) break;
            const int excl = incl - rem;
            const int take = max(0, min(rem, LANE_CAND - excl));
            for (int j = 0; j < take; j++) {
                const int q = done + j;
                const unsigned long long w = q < 6 ? ov0 >> (10 * q) : q < nov ? ov1 >> (10 * (q - 6)) : 0;
                const int t = q < nov ? (int)(w & 1023u) : (int)((bf >> (10 * (q - nov))) & 1023u);
                clist[excl + j] = (lane << 10) | t;
            }
            __syncwarp();
            const int nsub = min(total, LANE_CAND);
            for (int k = lane; k < nsub; k += 32) {
                const int pk = clist[k], t = pk & 1023;
                const Box a = c.det(r0 + (pk >> 10));
                double iou;
                cres[k] = c.entry(a, (a.x1 + a.x2) / 2.0, (a.y1 + a.y2) / 2.0, t, &iou);
                if (iou > iou_thr) { atomicAdd(&colcnt[t], 1); clist[k] = pk | (1 << 20); }
            }
            __syncwarp();
            for (int j = 0; j < take; j++) {
                const int pk = clist[excl + j], t = pk & 1023;
                const double cc = cres[excl + j];
                if (pk & (1 << 20)) { cnt++; ht = max(ht, t); }
                if (cc < best || (cc == best && t < bt)) { best = cc; bt = t; }
            }
            done += take;
            __syncwarp();
        }
        // zero-motion, zero-IoU entries hold -(0.0 + 0.0) = -0.0
        if (zero_t != 0x7fffffff && (-0.0 < best || (-0.0 == best && zero_t < bt))) { best = -0.0; bt = zero_t; }
        if (live && !dense) { rmin[d] = best; rarg[d] = bt; rowcnt[d] = cnt; hit_t[d] = ht; }
        // ---- rare rows: every entry exactly, by the whole warp
        unsigned dm = __ballot_sync(BM_FULL, live && dense);
        while (dm) {
            const int src = __ffs(dm) - 1;
            dm &= dm - 1;
            const int dd = r0 + src;
            const Box a = c.det(dd);
            const double ax = (a.x1 + a.x2) / 2.0, ay = (a.y1 + a.y2) / 2.0;
            double m = CUDART_INF;
            int mt = 0x7fffffff, hc = 0, hmax = -1;
            for (int t = lane; t < nT; t += 32) {
                double iou;
                const double cc = c.entry(a, ax, ay, t, &iou);
                if (iou > iou_thr) { hc++; hmax = max(hmax, t); atomicAdd(&colcnt[t], 1); }
                if (cc < m) { m = cc; mt = t; }
            }
            warp_argmin(m, mt);
            for (int o = 16; o > 0; o >>= 1) {
                hc += __shfl_xor_sync(BM_FULL, hc, o);
                hmax = max(hmax, __shfl_xor_sync(BM_FULL, hmax, o));
            }
            if (lane == 0) { rmin[dd] = m; rarg[dd] = mt; rowcnt[dd] = hc; hit_t[dd] = hmax; }
        }
    }
}

// ---------------------------------------------------------------------------------------
// DeepOCSORT first round (association.py:427-474): total = (iou + OCM) + emb, with
//   emb = dets_embs @ trk_embs.T (tensor-core contraction, st.embsim), emb[iou <= 0] = 0,
//   then compute_aw_max_metric (:320-374): every row and column with >= 2 positive entries
//   scales its weight by 1 - max(second / max - aw_param, 0) / (1 - aw_param); or emb * w
//   with aw_off.
// The weights need every entry of the row AND of the column, so this round materialises the
// matrix (stream-private, L2-resident scratch): pass A a warp per row (base cost, masked
// embedding, row top-2, IoU hit counts), pass B a thread per column (column top-2), pass C a
// warp per row (final cost -(total), row minima for the solver, matrix maximum for lapjv's
// padding).  Column index = list position; `remap` is set after a NaN-track removal.
template <int NT>
__device__ inline void deep_build_cost(const OcsState& st, const OcsParams& pr, int s, const OcmCost& ocm,
                                       const int* dind, int nD, int nT, bool remap, int* colcnt, double* rmin,
                                       int* rarg, int* rowcnt, int* hit_t, double* s_cmx) {
    const int tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    double* cost = st.scratch + (size_t)s * st.scratch_elems;       // [nD][nT]
    double* embm = cost + (size_t)Tcap * Dcap;                      // [nD][nT]
    double* roww = st.wbuf + (size_t)s * (Dcap + Tcap);
    double* colw = roww + Dcap;
    const float* sim = st.embsim + (size_t)s * Dcap * Tcap;
    const int* epos = st.epos + (size_t)s * Tcap;
    const bool aw = pr.emb_on && !pr.aw_off;
    for (int d = wid; d < nD; d += NT / 32) {
        const Box a = ocm.det(d);
        const double cx = (a.x1 + a.x2) / 2.0, cy = (a.y1 + a.y2) / 2.0;
        const float* srow = sim + (size_t)dind[d] * Tcap;
        double m1 = 0.0, m2 = 0.0;
        int npos = 0, hits = 0, hit = -1;
        for (int t = lane; t < nT; t += 32) {
            double iou;
            const double c = ocm.entry(a, cx, cy, t, &iou);
            double e = 0.0;
            if (pr.emb_on) {
                e = (double)srow[remap ? epos[t] : t];
                if (iou <= 0) e = 0.0;
            }
            cost[(size_t)d * nT + t] = c;
            embm[(size_t)d * nT + t] = e;
            if (e > 0) { npos++; if (e > m1) { m2 = m1; m1 = e; } else if (e > m2) m2 = e; }
            if (iou > pr.iou_thr) { hits++; hit = t; atomicAdd(&colcnt[t], 1); }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const double b1 = shfl_xor_d(m1, o), b2 = shfl_xor_d(m2, o);
            m2 = fmax(fmin(m1, b1), fmax(m2, b2));
            m1 = fmax(m1, b1);
            npos += __shfl_xor_sync(BM_FULL, npos, o);
            hits += __shfl_xor_sync(BM_FULL, hits, o);
            hit = max(hit, __shfl_xor_sync(BM_FULL, hit, o));
        }
        if (lane == 0) {
            double w = 1.0;
            if (aw && npos >= 2) w = 1.0 - fmax(m2 / m1 - pr.aw_param, 0.0) / (1.0 - pr.aw_param);
            roww[d] = w;
            rowcnt[d] = hits;
            hit_t[d] = hit;
        }
    }
    __syncthreads();
    if (aw) {
        for (int t = tid; t < nT; t += NT) {
            double m1 = 0.0, m2 = 0.0;
            int npos = 0;
            for (int d = 0; d < nD; d++) {
                const double e = embm[(size_t)d * nT + t];
                if (e > 0) { npos++; if (e > m1) { m2 = m1; m1 = e; } else if (e > m2) m2 = e; }
            }
            colw[t] = npos >= 2 ? 1.0 - fmax(m2 / m1 - pr.aw_param, 0.0) / (1.0 - pr.aw_param) : 1.0;
        }
        __syncthreads();
    }
    double cmx = -CUDART_INF;
    for (int d = wid; d < nD; d += NT / 32) {
        const double rw = roww[d];
        double best = CUDART_INF;
        int barg = 0x7fffffff;
        for (int t = lane; t < nT; t += 32) {
            double c = cost[(size_t)d * nT + t];
            if (pr.emb_on) {
                const double e = embm[(size_t)d * nT + t];
                // w_emb = full(w); w[row] *= row_weight; w[:, col] *= col_weight; w * emb
                const double w = aw ? (pr.w_emb * rw) * colw[t] : pr.w_emb;
                c = -((-c) + w * e);
                cost[(size_t)d * nT + t] = c;
            }
            cmx = fmax(cmx, c);
            if (c < best) { best = c; barg = t; }
        }
        warp_argmin(best, barg);
        if (lane == 0) { rmin[d] = best; rarg[d] = barg; }
    }
    for (int o = 16; o > 0; o >>= 1) cmx = fmax(cmx, shfl_xor_d(cmx, o));
    if (lane == 0) s_cmx[wid] = cmx;
    __syncthreads();
}

// The same build as its own launch with four times the threads of the association CTA (the pass
// is a flat loop over detection x track pairs of long fp64 chains: it wants occupancy, not the
// association kernel's shared-memory working set).  Streams whose prediction produced a NaN box
// are left to the association kernel, which removes the track first (OH_DEEP = 0).
constexpr int KD_NT = 256;
__global__ void __launch_bounds__(KD_NT, 4)
deep_cost_kernel(OcsState st, OcsParams pr, const float* __restrict__ dets) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x;
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    int* hdr = st.hdr + (size_t)s * OH_COUNT;
    const int nD = hdr[OH_ND], nT = hdr[OH_NLIST];
    float4* dbox = reinterpret_cast<float4*>(smem_raw);
    double* rmin = reinterpret_cast<double*>(dbox + Dcap);
    int* dind = reinterpret_cast<int*>(rmin + Dcap);
    int* rarg = dind + Dcap;
    int* rowcnt = rarg + Dcap;
    int* hit_t = rowcnt + Dcap;
    int* colcnt = hit_t + Dcap;
    __shared__ double s_cmx[KD_NT / 32];
    __shared__ int s_bad;
    if (tid == 0) s_bad = 0;
    __syncthreads();
    const double* ab = st.abuf + (size_t)s * AB_COUNT * Tcap;
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    const int* hs = st.hsel + (size_t)s * 2 * Dcap;
    for (int d = tid; d < nD; d += KD_NT) {
        const int i = hs[d];
        dbox[d] = make_float4(dsrc[i * 6 + 0], dsrc[i * 6 + 1], dsrc[i * 6 + 2], dsrc[i * 6 + 3]);
        dind[d] = i;
    }
    for (int p = tid; p < nT; p += KD_NT) {
        colcnt[p] = 0;
        bool nan = false;
        for (int f = 0; f < 4; f++) nan |= ab[f * Tcap + p] != ab[f * Tcap + p];
        if (nan) s_bad = 1;
    }
    __syncthreads();
    if (s_bad || nD <= 0 || nT <= 0) {
        if (tid == 0) hdr[OH_DEEP] = 0;
        return;
    }
    const OcmCost ocm = {dbox, ab, Tcap, pr.asso_func, pr.diag, pr.inertia};
    deep_build_cost<KD_NT>(st, pr, s, ocm, dind, nD, nT, false, colcnt, rmin, rarg, rowcnt, hit_t, s_cmx);
    double* od = st.deep_d + (size_t)s * (Dcap + 1);
    int* oi = st.deep_i + (size_t)s * (3 * Dcap + Tcap);
    for (int d = tid; d < nD; d += KD_NT) {
        od[d] = rmin[d];
        oi[d] = rarg[d]; oi[Dcap + d] = rowcnt[d]; oi[2 * Dcap + d] = hit_t[d];
    }
    for (int p = tid; p < nT; p += KD_NT) oi[3 * Dcap + p] = colcnt[p];
    if (tid == 0) {
        double cmx = s_cmx[0];
        for (int i = 1; i < KD_NT / 32; i++) cmx = fmax(cmx, s_cmx[i]);
        od[Dcap] = cmx;
        hdr[OH_DEEP] = 1;
    }
}
static size_t deep_cost_smem(const OcsState& st) { return (size_t)st.Dcap * (16 + 8 + 16) + (size_t)st.Tcap * 4; }

// Second-round association shared by BYTE and OCR: rows = the nA detections listed in
// `alist` (indices into a detection set `abox`), cols = the nB list positions in
// `blist` with boxes fetched by `bfetch`.  lapjv on -iou, accept pairs with iou >= thr
// (ocsort.py:336-353, :363-380).  Writes pair_b[r] = column index or -1.  Returns
// whether the block ran (max iou > thr and no NaN).  The matrix lives in the stream's
// L2-resident scratch.
template <typename BFetch>
__device__ bool ocs_second_round(const OcsState& st, const OcsParams& pr, int s, OcsSmem& sm, int Tcap, int Dcap,
                                 const float4* abox, const int* alist, int nA,
                                 const int* blist, int nB, BFetch bfetch, int* pair_b) {
    const int tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const size_t need = (size_t)nA * nB;
    double* cost = st.scratch + (size_t)s * st.scratch_elems;
    if (need > st.scratch_elems) {
        if (tid == 0) atomicOr(&st.hdr[(size_t)s * OH_COUNT + OH_ERR], BM_ERR_SCRATCH);
        for (int r = tid; r < nA; r += K2_NT) pair_b[r] = -1;
        __syncthreads();
        return false;
    }
    double mx = -CUDART_INF;   // max of iou_left
    double cmx = -CUDART_INF;  // max of -iou_left
    int nanf = 0;
    for (size_t e = tid; e < need; e += K2_NT) {
        const int r = (int)(e / nB), c = (int)(e % nB);
        const float4 av = abox[alist[r]];
        const Box a = {(double)av.x, (double)av.y, (double)av.z, (double)av.w};
        const Box b = bfetch(blist[c]);
        const double iou = asso_pair(pr.asso_func, a, b, pr.diag);
        if (iou != iou) nanf = 1;
        mx = fmax(mx, iou);
        const double cc = -iou;
        cmx = fmax(cmx, cc);
        cost[e] = cc;
    }
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, shfl_xor_d(mx, o));
        cmx = fmax(cmx, shfl_xor_d(cmx, o));
        nanf |= __shfl_xor_sync(BM_FULL, nanf, o);
    }
    __shared__ double s_mx[K2_NT / 32], s_cmx[K2_NT / 32];
    __shared__ int s_nan[K2_NT / 32];
    if (lane == 0) { s_mx[wid] = mx; s_cmx[wid] = cmx; s_nan[wid] = nanf; }
    __syncthreads();
    mx = s_mx[0]; cmx = s_cmx[0]; nanf = s_nan[0];
    for (int i = 1; i < K2_NT / 32; i++) { mx = fmax(mx, s_mx[i]); cmx = fmax(cmx, s_cmx[i]); nanf |= s_nan[i]; }
    __syncthreads();
    const bool run = !nanf && mx > pr.iou_thr;
    if (!run) {
        for (int r = tid; r < nA; r += K2_NT) pair_b[r] = -1;
        __syncthreads();
        return false;
    }
    if (wid == 0) {
        // Only pairs with iou >= thr survive, and the caller re-sorts what is left: any
        // optimal solution gives the same result, so the fast exact solver is used.
        RectWork w = rect_carve(sm.work, nA, nB);
        const int bad = rect_solve_warp(DenseCost{cost, nB}, nA, nB, 2.0 * (cmx + 1.0), w, nullptr, nullptr);
        if (bad && lane == 0) sm.misc[M_LAPBAD] = 1;
        for (int r = lane; r < nA; r += 32) {
            int c = w.x[r];
            // `if iou_left[m0, m1] < thr: continue`
            if (c >= 0 && -cost[(size_t)r * nB + c] < pr.iou_thr) c = -1;
            pair_b[r] = c;
        }
        if (lane == 0) atomicAdd(&st.hdr[(size_t)s * OH_COUNT + OH_LAPS], 1);
    }
    __syncthreads();
    return true;
}

struct FetchPred {     // predicted boxes (global, written by the predict kernel)
    const double* tb; int Tcap;
    __device__ Box operator()(int p) const {
        return Box{tb[AB_X1 * Tcap + p], tb[AB_Y1 * Tcap + p], tb[AB_X2 * Tcap + p], tb[AB_Y2 * Tcap + p]};
    }
};
struct FetchLast {     // last observations (global state)
    const double* lastobs; const int* slot_of; int Tcap;
    __device__ Box operator()(int p) const {
        const int sl = slot_of[p];
        return Box{lastobs[0 * Tcap + sl], lastobs[1 * Tcap + sl], lastobs[2 * Tcap + sl], lastobs[3 * Tcap + sl]};
    }
};

template <bool DEEP>
__device__ __forceinline__ void ocs_assoc_cta(const OcsState& st, const OcsParams& pr, const float* __restrict__ dets,
                                              int s, unsigned char* smem_raw, OcsSmem& sm_out) {
    const int tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    OcsSmem sm = ocs_carve(smem_raw, Tcap, Dcap, pr.use_byte);
    sm_out = sm;
    OcsView v(st, s);
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    int* upd_slot = st.upd + (size_t)s * 2 * Dcap;
    int* upd_det = upd_slot + Dcap;
    double* ab = st.abuf + (size_t)s * AB_COUNT * Tcap;     // operands written by the predict kernel

    long long* prof = st.prof ? st.prof + (size_t)s * 8 : nullptr;
#define BM_STAMP(k) do { if (prof && tid == 0) prof[k] = clock64(); } while (0)
    BM_STAMP(0);
    if (tid < M_COUNT) sm.misc[tid] = 0;
    __syncthreads();
    // ---- the confidence split (ocsort.py:265-275) was done by the predict launch
    const int nD = v.hdr[OH_ND], nS = pr.use_byte ? v.hdr[OH_NS] : 0;
    int nT = v.hdr[OH_NLIST];
    {
        const int* hs = st.hsel + (size_t)s * 2 * Dcap;
        for (int d = tid; d < nD; d += K2_NT) {
            const int i = hs[d];
            sm.dbox[d] = make_float4(dsrc[i * 6 + 0], dsrc[i * 6 + 1], dsrc[i * 6 + 2], dsrc[i * 6 + 3]);
            sm.dind[d] = i;
            sm.d2t[d] = -1;
            sm.rmin[d] = CUDART_INF; sm.rarg[d] = 0x7fffffff; sm.rowcnt[d] = 0; sm.hit_t[d] = -1;
        }
        for (int r = tid; r < nS; r += K2_NT) {
            const int i = hs[Dcap + r];
            sm.sbox[r] = make_float4(dsrc[i * 6 + 0], dsrc[i * 6 + 1], dsrc[i * 6 + 2], dsrc[i * 6 + 3]);
            sm.sind[r] = i;
        }
    }
    // ---- screening tables of the predicted tracks
    const OcmCost ocm = {sm.dbox, ab, Tcap, pr.asso_func, pr.diag, pr.inertia};
    OcmScreen scr;
    scr.boxf = reinterpret_cast<float4*>(sm.work);
    scr.kc = scr.boxf + Tcap;
    scr.mvf = reinterpret_cast<float2*>(scr.kc + Tcap);
    scr.cs = reinterpret_cast<float*>(scr.mvf + Tcap);
    // rows already screened by ocs_screen_kernel (which also saw that no prediction is NaN)
    const bool prescreened = !DEEP && pr.prescreen && pr.asso_func == ASSO_IOU && st.scan != nullptr && v.hdr[OH_DEEP] == 1;
    for (int p = tid; p < nT; p += K2_NT) {
        sm.slot_of[p] = v.order[p];
        sm.t2d[p] = -1;
        sm.colcnt[p] = 0;
        if (!prescreened && ocm_screen_fill(ocm, scr, p)) sm.misc[M_ANYBAD] = 1;
    }
    __syncthreads();
    if (sm.misc[M_ANYBAD]) {   // NaN prediction: the reference drops the track (ocsort.py:284-288)
        if (tid == 0) {
            int k = 0;
            for (int p = 0; p < nT; p++) {
                bool nan = false;
                for (int f = 0; f < 4; f++) nan |= ab[f * Tcap + p] != ab[f * Tcap + p];
                if (nan) continue;
                for (int f = 0; f < AB_COUNT; f++) ab[f * Tcap + k] = ab[f * Tcap + p];
                if (DEEP) st.epos[(size_t)s * Tcap + k] = p;
                sm.slot_of[k] = sm.slot_of[p];
                v.order[k] = sm.slot_of[k];
                k++;
            }
            v.hdr[OH_NLIST] = k;
            sm.misc[M_NT] = k;
        }
        __syncthreads();
        nT = sm.misc[M_NT];
        for (int p = tid; p < nT; p += K2_NT) ocm_screen_fill(ocm, scr, p);
        __syncthreads();
    }

    BM_STAMP(1);
    // ---- first round: cost[d][t] = -(iou + OCM), det-major like enhanced_associate
    // (association.py:427-461), evaluated on demand; permutation fast path (:501-503) or
    // exact LAP (:505-507)
    __shared__ double s_cmx[K2_NT / 32];
    if (nD > 0 && nT > 0) {
        if (DEEP) {
            if (v.hdr[OH_DEEP]) {       // built by deep_cost_kernel: fetch the row / column summaries
                const double* od = st.deep_d + (size_t)s * (Dcap + 1);
                const int* oi = st.deep_i + (size_t)s * (3 * Dcap + Tcap);
                for (int d = tid; d < nD; d += K2_NT) {
                    sm.rmin[d] = od[d];
                    sm.rarg[d] = oi[d]; sm.rowcnt[d] = oi[Dcap + d]; sm.hit_t[d] = oi[2 * Dcap + d];
                }
                for (int p = tid; p < nT; p += K2_NT) sm.colcnt[p] = oi[3 * Dcap + p];
                if (tid < K2_NT / 32) s_cmx[tid] = od[Dcap];
                __syncthreads();
            } else {
                deep_build_cost<K2_NT>(st, pr, s, ocm, sm.dind, nD, nT, sm.misc[M_ANYBAD] != 0, sm.colcnt, sm.rmin,
                                       sm.rarg, sm.rowcnt, sm.hit_t, s_cmx);
            }
        } else if (pr.asso_func == ASSO_IOU)
            ocm_scan_rows_lane(ocm, scr, nD, nT, pr.iou_thr, sm.colcnt, sm.rmin, sm.rarg, sm.rowcnt, sm.hit_t,
                               sm.cand + ocs_cand_ints(Tcap) * wid, prescreened ? st.scan + (size_t)s * Dcap : nullptr);
        else
            ocm_scan_rows(ocm, scr, nD, nT, pr.iou_thr, sm.cand + ocs_cand_ints(Tcap) * wid, sm.colcnt, sm.rmin, sm.rarg,
                          sm.rowcnt, sm.hit_t, sm.um_d);
        __syncthreads();       // the screening tables are dead from here on (sm.work is reused)
        BM_STAMP(2);
        // permutation fast path: every row/column has at most one IoU hit, at least one exists
        int rmax = 0, cmax_ = 0;
        for (int d = lane; d < nD; d += 32) rmax = max(rmax, sm.rowcnt[d]);
        for (int t = lane; t < nT; t += 32) cmax_ = max(cmax_, sm.colcnt[t]);
        for (int o = 16; o > 0; o >>= 1) {
            rmax = max(rmax, __shfl_xor_sync(BM_FULL, rmax, o));
            cmax_ = max(cmax_, __shfl_xor_sync(BM_FULL, cmax_, o));
        }
        const bool fast = rmax == 1 && cmax_ == 1;           // uniform over the CTA
        double* cost = st.scratch + (size_t)s * st.scratch_elems;
        if (!DEEP && !fast && nD > nT) {
            // WHICH zero-gain detections stay unassigned when there are more detections than
            // tracks is decided by lapjv's scan order (the reference's unmatched-detection ORDER,
            // and with it the new ids, depends on it): materialise the matrix for the
            // step-for-step lapjv emulation.
            double cmx = -CUDART_INF;
            for (int d = wid; d < nD; d += K2_NT / 32)
                for (int t = lane; t < nT; t += 32) {
                    const double cc = ocm(d, t);
                    cost[(size_t)d * nT + t] = cc;
                    cmx = fmax(cmx, cc);
                }
            for (int o = 16; o > 0; o >>= 1) cmx = fmax(cmx, shfl_xor_d(cmx, o));
            if (lane == 0) s_cmx[wid] = cmx;
            __syncthreads();
        }
        if (wid == 0) {
            if (fast) {
                for (int d = lane; d < nD; d += 32) {
                    const int t = sm.rowcnt[d] == 1 ? sm.hit_t[d] : -1;
                    sm.d2t[d] = t;
                    sm.rowcnt[d] = t >= 0 ? 1 : 0;           // 1 = accepted match
                    if (t >= 0) sm.t2d[t] = d;
                }
            } else {
                int bad;
                int* xsol;
                if (nD <= nT) {
                    // Every detection is assigned in every optimal solution (any real entry beats
                    // the padding), so the outcome does not depend on lapjv's tie-breaking and the
                    // fast exact solver applies; the null option is never taken (cost +inf).
                    RectWork w = rect_carve(sm.work, nD, nT);
                    if (DEEP) {
                        bad = rect_solve_warp(DenseCost{cost, nT}, nD, nT, CUDART_INF, w, sm.rmin, sm.rarg);
                    } else {
                        const OcmCostStaged oc2 = {ocm, reinterpret_cast<double*>(sm.cand), &sm.misc[M_ROW], w.spc, nT};
                        bad = rect_solve_warp(oc2, nD, nT, CUDART_INF, w, sm.rmin, sm.rarg);
                        oc2.release();
                    }
                    xsol = w.x;
                } else {
                    double pad = s_cmx[0];
                    for (int i = 1; i < K2_NT / 32; i++) pad = fmax(pad, s_cmx[i]);
                    LapMat m = {cost, nT, nD, nT, nD + nT, 1, pad + 1.0};
                    LapWork w = lap_carve(sm.work, nD + nT);
                    bad = lap_solve_warp(m, w);
                    xsol = w.x;
                }
                if (bad && lane == 0) sm.misc[M_LAPBAD] = 1;
                for (int d = lane; d < nD; d += 32) {
                    int t = xsol[d];
                    if (t >= nT) t = -1;
                    int stt = 0;                              // 0 never assigned
                    if (t >= 0) {
                        // iou re-evaluated for the assigned pair only (same arithmetic, same value)
                        const double iou = asso_pair(pr.asso_func, ocm.det(d), ocm.trk(t), pr.diag);
                        if (iou < pr.iou_thr) { stt = 2; sm.colcnt[t] = -1 - d; }   // rejected (remember the row)
                        else { stt = 1; sm.t2d[t] = d; }
                    }
                    sm.d2t[d] = stt == 1 ? t : -1;
                    sm.rowcnt[d] = stt;
                    sm.hit_t[d] = t;                          // lapjv's column, accepted or not
                }
                if (lane == 0) atomicAdd(&v.hdr[OH_LAPS], 1);
            }
            if (lane == 0) sm.misc[M_FAST] = fast ? 1 : 0;
        } else if (!DEEP && wid == 1 && !fast && nD <= nT) {
            RectWork w = rect_carve(sm.work, nD, nT);
            const OcmCostStaged oc2 = {ocm, reinterpret_cast<double*>(sm.cand), &sm.misc[M_ROW], w.spc, nT};
            oc2.helper();
        }
        __syncthreads();
    } else {
        for (int d = tid; d < nD; d += K2_NT) { sm.rowcnt[d] = 0; sm.hit_t[d] = -1; }
        __syncthreads();
    }
    BM_STAMP(3);
    // ---- unmatched lists in the reference's order (repair R3): never-assigned ascending,
    // then IoU-rejected pairs in detection-row order; matched pairs go to the update list
    if (wid == 0) {
        const bool lap_ran = sm.misc[M_FAST] == 0 && nD > 0;
        int nud = warp_compact(nD, sm.um_d, 0, [&](int d) { return sm.rowcnt[d] == 0; }, [&](int d) { return d; });
        int nut = warp_compact(nT, sm.um_t, 0,
                               [&](int t) { return sm.t2d[t] < 0 && !(sm.colcnt[t] < 0 && lap_ran); },
                               [&](int t) { return t; });
        if (nT > 0) {
            const int nud0 = nud;
            nud = warp_compact(nD, sm.um_d, nud0, [&](int d) { return sm.rowcnt[d] == 2; }, [&](int d) { return d; });
            nut = warp_compact(nD, sm.um_t, nut, [&](int d) { return sm.rowcnt[d] == 2; }, [&](int d) { return sm.hit_t[d]; });
        }
        const int nup = warp_compact(nD, upd_slot, 0, [&](int d) { return sm.d2t[d] >= 0; },
                                     [&](int d) { return sm.slot_of[sm.d2t[d]]; });
        warp_compact(nD, upd_det, 0, [&](int d) { return sm.d2t[d] >= 0; }, [&](int d) { return sm.dind[d]; });
        if (lane == 0) { sm.misc[M_NUD] = nud; sm.misc[M_NUT] = nut; sm.misc[M_NUPD] = nup; }
    }
    __syncthreads();

    BM_STAMP(4);
    // ---- BYTE round on the low-score detections (ocsort.py:330-356)
    int nUD = sm.misc[M_NUD], nUT = sm.misc[M_NUT];
    if (pr.use_byte && nS > 0 && nUT > 0) {
        for (int r = tid; r < nS; r += K2_NT) sm.d2t[r] = r;
        for (int t = tid; t < nT; t += K2_NT) sm.colcnt[t] = 0;
        __syncthreads();
        const bool ran = ocs_second_round(st, pr, s, sm, Tcap, Dcap, sm.sbox, sm.d2t, nS, sm.um_t, nUT,
                                          FetchPred{ab, Tcap}, sm.hit_t);
        if (ran) {
            if (wid == 0) {
                const int n0 = sm.misc[M_NUPD];
                for (int r = lane; r < nS; r += 32) if (sm.hit_t[r] >= 0) sm.colcnt[sm.um_t[sm.hit_t[r]]] = 2;   // taken
                __syncwarp();
                const int n1 = warp_compact(nS, upd_slot, n0, [&](int r) { return sm.hit_t[r] >= 0; },
                                            [&](int r) { return sm.slot_of[sm.um_t[sm.hit_t[r]]]; });
                warp_compact(nS, upd_det, n0, [&](int r) { return sm.hit_t[r] >= 0; }, [&](int r) { return sm.sind[r]; });
                // np.setdiff1d: sorted ascending, taken tracks removed
                for (int c = lane; c < nUT; c += 32) { const int t = sm.um_t[c]; if (sm.colcnt[t] != 2) sm.colcnt[t] = 1; }
                __syncwarp();
                const int k = warp_compact(nT, sm.um_t, 0, [&](int t) { return sm.colcnt[t] == 1; }, [&](int t) { return t; });
                if (lane == 0) { sm.misc[M_NUT] = k; sm.misc[M_NUPD] = n1; }
            }
            __syncthreads();
            nUT = sm.misc[M_NUT];
        }
    }
    // ---- OCR: leftover detections vs the tracks' last observations (ocsort.py:358-386)
    if (nUD > 0 && nUT > 0) {
        for (int t = tid; t < nT; t += K2_NT) sm.colcnt[t] = 0;
        for (int d = tid; d < nD; d += K2_NT) sm.rowcnt[d] = 0;
        __syncthreads();
        const bool ran = ocs_second_round(st, pr, s, sm, Tcap, Dcap, sm.dbox, sm.um_d, nUD, sm.um_t, nUT,
                                          FetchLast{v.lastobs, sm.slot_of, Tcap}, sm.hit_t);
        if (ran) {
            if (wid == 0) {
                const int n0 = sm.misc[M_NUPD];
                for (int r = lane; r < nUD; r += 32) {
                    const int c = sm.hit_t[r];
                    if (c >= 0) { sm.colcnt[sm.um_t[c]] = 2; sm.rowcnt[sm.um_d[r]] = 2; }
                }
                __syncwarp();
                const int n1 = warp_compact(nUD, upd_slot, n0, [&](int r) { return sm.hit_t[r] >= 0; },
                                            [&](int r) { return sm.slot_of[sm.um_t[sm.hit_t[r]]]; });
                warp_compact(nUD, upd_det, n0, [&](int r) { return sm.hit_t[r] >= 0; },
                             [&](int r) { return sm.dind[sm.um_d[r]]; });
                for (int c = lane; c < nUT; c += 32) { const int t = sm.um_t[c]; if (sm.colcnt[t] != 2) sm.colcnt[t] = 1; }
                for (int r = lane; r < nUD; r += 32) { const int d = sm.um_d[r]; if (sm.rowcnt[d] != 2) sm.rowcnt[d] = 1; }
                __syncwarp();
                const int k = warp_compact(nT, sm.um_t, 0, [&](int t) { return sm.colcnt[t] == 1; }, [&](int t) { return t; });
                const int kd = warp_compact(nD, sm.um_d, 0, [&](int d) { return sm.rowcnt[d] == 1; }, [&](int d) { return d; });
                if (lane == 0) { sm.misc[M_NUT] = k; sm.misc[M_NUD] = kd; sm.misc[M_NUPD] = n1; }
            }
            __syncthreads();
            nUD = sm.misc[M_NUD];
        }
    }
    BM_STAMP(5);
    // ---- births: ids are handed out in unmatched-detection list order (ocsort.py:392-413)
    int* born = st.born + (size_t)s * Dcap;
    for (int r = tid; r < nUD; r += K2_NT) born[r] = sm.dind[sm.um_d[r]];
    if (tid == 0) {
        v.hdr[OH_NUPD] = sm.misc[M_NUPD];
        v.hdr[OH_NBORN] = nUD;
        v.hdr[OH_MATCH] += sm.misc[M_NUPD];
        if (sm.misc[M_LAPBAD]) atomicOr(&v.hdr[OH_ERR], 8);
        BM_STAMP(6);
#undef BM_STAMP
    }
}

// ============================================================================ K2a
// fp32 screen of the first-round rows as its own launch (asso_func = iou): one lane per detection
// row walks all tracks of its stream against shared-memory tables and leaves the few columns that
// need an exact fp64 evaluation in st.scan.  <= 64 registers x 128 threads: every stream of a
// 1024-stream batch is resident at once with all rows of a frame in ONE pass, instead of two
// passes on the two warps of the register-heavy association CTA.  Phase 1 (every row) only tests
// box overlap; the rows left without a strong overlap (false positives, new objects: ~10 %) are
// packed onto the first lanes for the OCM band search (phase 2).  A stream with a NaN prediction
// (the association kernel drops that track first) or nothing to match is left alone:
// hdr[OH_DEEP] = 0 and the association kernel screens for itself.
constexpr int KS_NT = 128;
__host__ __device__ inline size_t ocs_screen_smem(int Tcap, int Dcap) {
    return ocm_screen_bytes(Tcap, 0) + 16 + 4 * (size_t)Dcap;
}
__global__ void __launch_bounds__(KS_NT, 8)
ocs_screen_kernel(OcsState st, OcsParams pr, const float* __restrict__ dets) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x;
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    int* hdr = st.hdr + (size_t)s * OH_COUNT;
    const int nD = hdr[OH_ND], nT = hdr[OH_NLIST];
    OcmScreen scr;
    scr.boxf = reinterpret_cast<float4*>(smem_raw);
    scr.kc = scr.boxf + Tcap;
    scr.mvf = reinterpret_cast<float2*>(scr.kc + Tcap);
    scr.cs = nullptr;
    int* bad = reinterpret_cast<int*>(scr.mvf + Tcap);     // [0] NaN prediction seen, [1] queued rows
    int* nweak = bad + 1;
    int* weak = bad + 4;
    if (tid == 0) { bad[0] = 0; bad[1] = 0; }
    __syncthreads();
    const OcmCost ocm = {nullptr, st.abuf + (size_t)s * AB_COUNT * Tcap, Tcap, pr.asso_func, pr.diag, pr.inertia};
    for (int p = tid; p < nT; p += KS_NT)
        if (ocm_screen_fill(ocm, scr, p)) bad[0] = 1;
    __syncthreads();
    const bool ok = nD > 0 && nT > 0 && bad[0] == 0;
    if (tid == 0) hdr[OH_DEEP] = ok ? 1 : 0;
    if (!ok) return;
    const int* hs = st.hsel + (size_t)s * 2 * Dcap;
    const float* dsrc = dets + (size_t)s * Dcap * 6;
    ScanRow* out = st.scan + (size_t)s * Dcap;
    const float thr = ocm_strong_thr(pr.inertia);
    for (int d = tid; d < nD; d += KS_NT) {
        const int i = hs[d];
        const float4 af = make_float4(dsrc[i * 6 + 0], dsrc[i * 6 + 1], dsrc[i * 6 + 2], dsrc[i * 6 + 3]);
        ScanRow r;
        if (!ocm_lane_overlaps(af, scr, nT, thr, r) && !r.dense) weak[atomicAdd(nweak, 1)] = d;
        out[d] = r;
    }
    __syncthreads();
    const int nw = *nweak;
    for (int k = tid; k < nw; k += KS_NT) {
        const int d = weak[k], i = hs[d];
        const float4 af = make_float4(dsrc[i * 6 + 0], dsrc[i * 6 + 1], dsrc[i * 6 + 2], dsrc[i * 6 + 3]);
        ScanRow r = out[d];
        ocm_lane_band(af, scr, nT, r);
        out[d] = r;
    }
}

template <bool DEEP>
__global__ void __launch_bounds__(K2_NT, 8)
ocs_assoc_kernel(OcsState st, OcsParams pr, const float* __restrict__ dets) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    OcsSmem sm;
    ocs_assoc_cta<DEEP>(st, pr, dets, blockIdx.x, smem_raw, sm);
}

// ============================================================================ K3
// One thread per (track, detection) pair of the frame: KalmanBoxTracker.update with a
// real detection (ocsort.py:136-169) + KalmanFilterXYSR.update incl. ORU.
template <bool DENSE>
__global__ void __launch_bounds__(K3_NT, DENSE ? 1 : 8)
ocs_update_kernel_t(OcsState st, OcsParams pr, const float* __restrict__ dets) {
    const int chunks = (st.Dcap + K3_NT - 1) / K3_NT;
    const int s = blockIdx.x / chunks;
    const int k = (blockIdx.x % chunks) * K3_NT + threadIdx.x;
    OcsView v(st, s);
    if (k >= v.hdr[OH_NUPD]) return;
    const int* upd_slot = st.upd + (size_t)s * 2 * st.Dcap;
    const int slot = upd_slot[k], di = upd_slot[st.Dcap + k];
    const float* dr = dets + ((size_t)s * st.Dcap + di) * 6;
    const double det[5] = {(double)dr[0], (double)dr[1], (double)dr[2], (double)dr[3], (double)dr[4]};
    const XysrConst kc = {pr.q_xy, pr.q_s};
    ocs_track_update<DENSE>(v, slot, det, (double)dr[5], di, pr.delta_t, kc);
}

// ============================================================================ K4
// Exclusive block scan of one int per thread (blockDim <= 1024); total on all threads.
__device__ __forceinline__ int block_excl_scan(int val, int* total, int* ws /*[33]*/) {
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
        if (lane < nw) ws[lane] = w;     // inclusive warp totals
        if (lane == 31) ws[32] = w;
    }
    __syncthreads();
    const int base = wid > 0 ? ws[wid - 1] : 0;
    *total = ws[32];
    __syncthreads();
    return base + inc - val;
}

// One CTA per stream, one thread per slot / list position (blockDim = Tcap rounded up
// to a warp, Tcap <= 1024).  OC-SORT: held to 32 registers -- a chain of block-wide phases that each start with a
// global load, so eight 256-thread CTAs per SM (every stream of the headline batch resident at once) beat its 84 bytes of
// spills: 36 -> 27 us.
template <bool DENSE>
__global__ void __launch_bounds__(1024, DENSE ? 1 : 2)
ocs_finalize_kernel_t(OcsState st, OcsParams pr, const float* __restrict__ dets, double* __restrict__ out,
                    int* __restrict__ nout) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int* inuse = reinterpret_cast<int*>(smem_raw);     // [Tcap]
    int* bslot = inuse + st.Tcap;                       // [Tcap]
    __shared__ int ws[33];
    const int s = blockIdx.x, tid = threadIdx.x;
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    OcsView v(st, s);
    const int nT = v.hdr[OH_NLIST];
    int born = v.hdr[OH_NBORN];
    const int id0 = v.hdr[OH_NEXT_ID];
    const int frame = v.hdr[OH_FRAME] + 1;
    const XysrConst kc = {pr.q_xy, pr.q_s};
    (void)kc;
    // slots in use, misses
    if (tid < Tcap) inuse[tid] = 0;
    __syncthreads();
    int slot = -1;
    if (tid < nT) {
        slot = v.order[tid];
        inuse[slot] = 1;
        if (v.I(OI_TSU, slot) != 0) ocs_track_miss<DENSE>(v, slot);   // update(None) (ocsort.py:388-389)
    }
    __syncthreads();
    // births take the free slots in ascending order
    const int isfree = (tid < Tcap && !inuse[tid]) ? 1 : 0;
    int nfree;
    const int frank = block_excl_scan(isfree, &nfree, ws);
    if (born > nfree) {
        if (tid == 0) atomicOr(&v.hdr[OH_ERR], BM_ERR_TRACK_OVERFLOW);
        born = nfree;
    }
    if (isfree && frank < born) bslot[frank] = tid;
    __syncthreads();
    if (tid < born) {
        const int di = st.born[(size_t)s * Dcap + tid];
        const float* dr = dets + ((size_t)s * Dcap + di) * 6;
        const double det[5] = {(double)dr[0], (double)dr[1], (double)dr[2], (double)dr[3], (double)dr[4]};
        ocs_track_init(v, bslot[tid], det, (double)dr[5], di, id0 + tid);
        if (st.bornslot) st.bornslot[(size_t)s * Dcap + tid] = bslot[tid];
    }
    if (tid >= nT && tid < nT + born) slot = bslot[tid - nT];
    __syncthreads();
    // output rows walk the list backwards; dead tracks leave the list (ocsort.py:414-436)
    const int nAll = nT + born;
    int emit = 0, keep = 0;
    if (tid < nAll) {
        const int tsu = v.I(OI_TSU, slot);
        emit = (tsu < 1 && (v.I(OI_STREAK, slot) >= pr.min_hits || frame <= pr.min_hits)) ? 1 : 0;
        keep = tsu > pr.max_age ? 0 : 1;
    }
    int n_emit, n_keep;
    const int erank = block_excl_scan(emit, &n_emit, ws);
    const int krank = block_excl_scan(keep, &n_keep, ws);
    if (emit) {
        const int k = n_emit - 1 - erank;              // rows are newest-first
        double b[4], lo[5];
#pragma unroll
        for (int c = 0; c < 5; c++) lo[c] = v.lastobs[c * Tcap + slot];
        const double lsum = (((lo[0] + lo[1]) + lo[2]) + lo[3]) + lo[4];
        if (lsum < 0) {
            double x[7];
#pragma unroll
            for (int c = 0; c < 7; c++) x[c] = v.xP[c * Tcap + slot];
            x_to_xyxy(x, b);
        } else {
#pragma unroll
            for (int c = 0; c < 4; c++) b[c] = lo[c];
        }
        // every emitted row belongs to a track that received a detection this frame (or was born
        // from one), so min(Tcap, Dcap) rows per stream always suffice
        double* o = out + ((size_t)s * min(Tcap, Dcap) + k) * 8;
        o[0] = b[0]; o[1] = b[1]; o[2] = b[2]; o[3] = b[3];
        o[4] = (double)(v.I(OI_ID, slot) + 1);
        o[5] = v.confcls[0 * Tcap + slot];
        o[6] = v.confcls[1 * Tcap + slot];
        o[7] = (double)v.I(OI_DETIND, slot);
    }
    if (keep) v.order[krank] = slot;      // every thread read its slot before the scans' barriers
    if (tid == 0) {
        nout[s] = n_emit;
        v.hdr[OH_NBORN] = born;                // clamped to the free slots (read by docs_emb_kernel)
        v.hdr[OH_NLIST] = n_keep;
        v.hdr[OH_NEXT_ID] = id0 + born;
        v.hdr[OH_FRAME] = frame;
        v.hdr[OH_BIRTHS] += born;
    }
}

// ============================================================================ DeepOCSORT
// Embedding side of the frame: update_emb for every (track, detection) update of the frame
// (deepocsort.py:184-186, :383-384, :418-421: emb = alpha emb + (1 - alpha) det_emb, then
// / ||emb||, alpha = dets_alpha from the detection confidence :357-360) and the raw copy for
// births (:436-445).  One warp per pair; fp64 arithmetic, fp32 storage (the contraction that
// consumes it is fp32-accurate).
constexpr int K5_NT = 256;
__global__ void __launch_bounds__(K5_NT)
docs_emb_kernel(OcsState st, OcsParams pr, const float* __restrict__ dets, const float* __restrict__ embs) {
    const int s = blockIdx.x, lane = lane_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap, E = pr.E;
    const int* hdr = st.hdr + (size_t)s * OH_COUNT;
    const int nupd = hdr[OH_NUPD], nborn = hdr[OH_NBORN];
    const int k = blockIdx.y * (K5_NT / 32) + warp_id();
    if (k >= nupd + nborn) return;
    int slot, di;
    const bool birth = k >= nupd;
    if (!birth) {
        const int* upd_slot = st.upd + (size_t)s * 2 * Dcap;
        slot = upd_slot[k]; di = upd_slot[Dcap + k];
    } else {
        slot = st.bornslot[(size_t)s * Dcap + (k - nupd)];
        di = st.born[(size_t)s * Dcap + (k - nupd)];
    }
    float* dst = st.emb + ((size_t)s * Tcap + slot) * E;
    const float* src = embs + ((size_t)s * Dcap + di) * E;
    const bool vec = (E & 3) == 0;
    if (birth) {
        if (vec) for (int e = lane * 4; e < E; e += 128) *reinterpret_cast<float4*>(dst + e) = __ldg(reinterpret_cast<const float4*>(src + e));
        else for (int e = lane; e < E; e += 32) dst[e] = src[e];
        return;
    }
    const double conf = (double)dets[((size_t)s * Dcap + di) * 6 + 4];
    const double trust = (conf - pr.det_thresh) / (1.0 - pr.det_thresh);
    const double af = pr.alpha_fixed;
    const double alpha = af + (1.0 - af) * (1.0 - trust);
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

// dets_embs @ trk_embs.T of one stream for cosine_tc_kernel (raw dot products): A = the
// frame's detection embeddings gathered through the high-score list, B = the track embeddings
// gathered through the list order; out[detection row][list position].
struct DeepDot {
    OcsState st;
    const float* embs;
    int E;
    __device__ CosineProblem operator()(int s) const {
        const int* hdr = st.hdr + (size_t)s * OH_COUNT;
        CosineProblem p;
        p.A = embs + (size_t)s * st.Dcap * E; p.ra = st.hsel + (size_t)s * 2 * st.Dcap; p.na = nullptr; p.nA = hdr[OH_ND];
        p.B = st.emb + (size_t)s * st.Tcap * E; p.rb = st.order + (size_t)s * st.Tcap; p.nb = nullptr; p.nB = hdr[OH_NLIST];
        p.E = E;
        p.out = st.embsim + (size_t)s * st.Dcap * st.Tcap; p.out_ld = st.Tcap;
        p.raw = 1;
        return p;
    }
};

}  // namespace bm

namespace bm {

size_t ocsort_step_smem(const OcsState& st, const OcsParams& pr) {
    return ocs_smem_bytes(st.Tcap, st.Dcap, pr.use_byte);
}

cudaError_t ocsort_step_configure(size_t) {
    cudaError_t e = smem_optin(ocs_assoc_kernel<false>);
    if (e != cudaSuccess) return e;
    e = smem_optin(ocs_assoc_kernel<true>);
    if (e != cudaSuccess) return e;
    e = smem_optin(ocs_screen_kernel);
    if (e != cudaSuccess) return e;
    e = smem_optin(deep_cost_kernel);
    if (e != cudaSuccess) return e;
    return smem_optin(cosine_tc_kernel<DeepDot>);
}

// BM_DEBUG_SYNC=1 in the environment synchronises after every kernel and names the one that
// faulted (compute-sanitizer is not available on every pool).
static cudaError_t dbg_check(const char* name, cudaStream_t stream) {
    static const bool on = [] { const char* e = getenv("BM_DEBUG_SYNC"); return e && e[0] == '1'; }();
    if (!on) return cudaSuccess;
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) fprintf(stderr, "boxmot_b200: kernel %s failed: %s\n", name, cudaGetErrorString(e));
    return e;
}

cudaError_t ocsort_step_launch(const OcsState& st, const OcsParams& pr, const float* dets, const int* ndets,
                               double* out, int* nout, cudaStream_t stream, cudaEvent_t* ev) {
    const int c1 = (st.Tcap + K1_NT - 1) / K1_NT, c3 = (st.Dcap + K3_NT - 1) / K3_NT;
    cudaError_t e;
    // optional per-kernel timing: five events bracket the four launches (bm_ocsort_kernel_times)
#define BM_EV(k) do { if (ev && (e = cudaEventRecord(ev[k], stream)) != cudaSuccess) return e; } while (0)
    BM_EV(0);
    ocs_predict_kernel<false><<<st.S * c1 + (st.S + K1_NT / 32 - 1) / (K1_NT / 32), K1_NT, 0, stream>>>(st, pr, dets, ndets, nullptr);
    if ((e = dbg_check("ocs_predict_kernel", stream)) != cudaSuccess) return e;
    BM_EV(1);
    if (pr.prescreen && pr.asso_func == ASSO_IOU && st.scan) {
        ocs_screen_kernel<<<st.S, KS_NT, ocs_screen_smem(st.Tcap, st.Dcap), stream>>>(st, pr, dets);
        if ((e = dbg_check("ocs_screen_kernel", stream)) != cudaSuccess) return e;
    }
    ocs_assoc_kernel<false><<<st.S, K2_NT, ocsort_step_smem(st, pr), stream>>>(st, pr, dets);
    if ((e = dbg_check("ocs_assoc_kernel", stream)) != cudaSuccess) return e;
    BM_EV(2);
    ocs_update_kernel_t<false><<<st.S * c3, K3_NT, 0, stream>>>(st, pr, dets);
    if ((e = dbg_check("ocs_update_kernel", stream)) != cudaSuccess) return e;
    BM_EV(3);
    const int nt4 = (st.Tcap + 31) / 32 * 32;
    ocs_finalize_kernel_t<false><<<st.S, nt4, (size_t)st.Tcap * 8, stream>>>(st, pr, dets, out, nout);
    if ((e = dbg_check("ocs_finalize_kernel", stream)) != cudaSuccess) return e;
    BM_EV(4);
#undef BM_EV
    return cudaGetLastError();
}

// DeepOCSORT frame step (DeepOcSort.update, deepocsort.py:316-498): CMC correction + predict,
// embedding contraction on the tensor cores, association with the materialised first round,
// Kalman update, finalize, embedding EMA / births.  ev: optional 7 events around the 6 launches.
cudaError_t deepocsort_step_launch(const OcsState& st, const OcsParams& pr, const float* dets, const int* ndets,
                                   const float* embs, const double* warps, double* out, int* nout,
                                   cudaStream_t stream, cudaEvent_t* ev) {
    const int c1 = (st.Tcap + K1_NT - 1) / K1_NT, c3 = (st.Dcap + K3_NT - 1) / K3_NT;
    cudaError_t e;
#define BM_EV(k) do { if (ev && (e = cudaEventRecord(ev[k], stream)) != cudaSuccess) return e; } while (0)
    BM_EV(0);
    ocs_predict_kernel<true><<<st.S * c1 + (st.S + K1_NT / 32 - 1) / (K1_NT / 32), K1_NT, 0, stream>>>(st, pr, dets, ndets, warps);
    if ((e = dbg_check("ocs_predict_kernel<deep>", stream)) != cudaSuccess) return e;
    BM_EV(1);
    if (pr.emb_on) {
        cosine_tc_kernel<DeepDot><<<dim3((st.Dcap + TC_M - 1) / TC_M, st.S), TC_NT, cosine_tc_smem_bytes(), stream>>>(
            DeepDot{st, embs, pr.E});
        if ((e = dbg_check("cosine_tc_kernel<DeepDot>", stream)) != cudaSuccess) return e;
    }
    deep_cost_kernel<<<st.S, KD_NT, deep_cost_smem(st), stream>>>(st, pr, dets);
    if ((e = dbg_check("deep_cost_kernel", stream)) != cudaSuccess) return e;
    BM_EV(2);
    ocs_assoc_kernel<true><<<st.S, K2_NT, ocsort_step_smem(st, pr), stream>>>(st, pr, dets);
    if ((e = dbg_check("ocs_assoc_kernel<deep>", stream)) != cudaSuccess) return e;
    BM_EV(3);
    ocs_update_kernel_t<true><<<st.S * c3, K3_NT, 0, stream>>>(st, pr, dets);
    if ((e = dbg_check("ocs_update_kernel", stream)) != cudaSuccess) return e;
    BM_EV(4);
    const int nt4 = (st.Tcap + 31) / 32 * 32;
    ocs_finalize_kernel_t<true><<<st.S, nt4, (size_t)st.Tcap * 8, stream>>>(st, pr, dets, out, nout);
    if ((e = dbg_check("ocs_finalize_kernel", stream)) != cudaSuccess) return e;
    BM_EV(5);
    if (pr.emb_on) {
        docs_emb_kernel<<<dim3(st.S, (st.Dcap + K5_NT / 32 - 1) / (K5_NT / 32)), K5_NT, 0, stream>>>(st, pr, dets, embs);
        if ((e = dbg_check("docs_emb_kernel", stream)) != cudaSuccess) return e;
    }
    BM_EV(6);
#undef BM_EV
    return cudaGetLastError();
}

}  // namespace bm
