// Device-resident, stream-batched OC-SORT engine: state layout and per-track helpers.
//
// One engine holds S independent streams (one reference `OcSort` instance each,
// boxmot/trackers/ocsort/ocsort.py:195-439).  Track state is SoA in HBM,
// [stream][field][slot] with `slot` fastest, so a warp that walks consecutive slots
// issues fully coalesced 8-byte loads.  Slots are stable for the life of a track; the
// reference's ordered `active_tracks` list is the per-stream `order` array
// (list position -> slot), which is what keeps cost-matrix column order -- and with it
// lapjv's tie-breaking -- identical to the reference.
#pragma once
#include "common.cuh"
#include "kf_xysr.cuh"
#include "iou.cuh"
#include "lapjv.cuh"

namespace bm {

enum { OI_ID = 0, OI_TSU, OI_HITS, OI_STREAK, OI_AGE, OI_DETIND, OI_FLAGS, OI_NNONE, OI_COUNT };
enum { OF_OBSERVED = 1, OF_SAVED = 2,
       OF_DENSE = 4 };   // the covariance (live and frozen copy) lives in the dense side array dP (after a rotating warp)
enum { OH_NLIST = 0, OH_NEXT_ID, OH_FRAME, OH_ERR, OH_LAPS, OH_MATCH, OH_BIRTHS, OH_NUPD, OH_NBORN, OH_ND, OH_NS, OH_DEEP, OH_COUNT = 12 };
// association operands written by the predict kernel, indexed by LIST POSITION
enum { AB_X1 = 0, AB_Y1, AB_X2, AB_Y2, AB_VY, AB_VX, AB_KCX, AB_KCY, AB_KVALID, AB_COUNT };

struct OcsParams {
    double det_thresh, min_conf, iou_thr, inertia, q_xy, q_s, diag;
    int max_age, min_hits, delta_t, asso_func, use_byte;
    // DeepOCSORT (deepocsort.py:265-314): embedding term of the first round and its EMA
    int emb_on, aw_off, E;         // emb_on = 0 <=> embedding_off
    double w_emb, alpha_fixed, aw_param;
    int prescreen;                 // run the first-round fp32 screen as its own launch (ocs_screen_kernel)
};

// Result of the fp32 screen of one first-round row (ocs_screen_kernel -> ocs_assoc_kernel): the
// columns that must be evaluated exactly, 10 bits each.
struct ScanRow {
    unsigned long long ov0, ov1;   // columns whose outward-rounded float box overlaps the detection (<= 12)
    unsigned long long bf;         // columns whose fp32 OCM cosine is within the band of the row's best (<= 6)
    int nov, nbf;
    int zero_t;                    // lowest zero-motion, non-overlapping column (INT_MAX = none)
    int dense;                     // 1 = evaluate the whole row exactly (degenerate box / list overflow)
};

struct OcsState {
    int S, Tcap, Dcap, R;
    size_t scratch_elems;          // doubles of global cost scratch per stream
    int smem_cost_elems;           // doubles of shared-memory cost region per CTA
    double* xP;                    // [S][20][Tcap]   x(7) then the 13 non-zero covariance entries (kf_xysr.cuh block form)
    double* sxP;                   // [S][20][Tcap]   state frozen at the first miss (ORU)
    double* lastz;                 // [S][4][Tcap]    last real measurement (xc,yc,s,r)
    double* lastobs;               // [S][5][Tcap]    last observation x1,y1,x2,y2,conf (-1 = none)
    double* vel;                   // [S][2][Tcap]    (dy, dx) unit direction
    double* confcls;               // [S][2][Tcap]
    double* ring;                  // [S][R][4][Tcap] observations of the last delta_t ages
    int* ring_age;                 // [S][R][Tcap]
    int* ints;                     // [S][OI_COUNT][Tcap]
    int* order;                    // [S][Tcap]
    int* hdr;                      // [S][OH_COUNT]
    unsigned long long* n_updates; // [S] live tracks that entered predict (bench unit)
    double* scratch;               // [S][scratch_elems]
    double* abuf;                  // [S][AB_COUNT][Tcap] predicted boxes / OCM operands per list position
    int* upd;                      // [S][2][Dcap] this frame's (slot, detection row) Kalman updates
    int* born;                     // [S][Dcap]    detection rows that start a track, in id order
    long long* prof;               // optional [S][8] clock64 phase stamps of the association kernel (debug)
    int* hsel;                     // [S][2][Dcap] detection rows of the high-score set, then of the BYTE set
    ScanRow* scan;                 // [S][Dcap]    screened first-round rows (asso_func = iou; valid when hdr[OH_DEEP] == 1)
    // DeepOCSORT only (nullptr in an OC-SORT engine)
    float* emb;                    // [S][Tcap][E]   track embedding by slot (KalmanBoxTracker.emb)
    float* embsim;                 // [S][Dcap][Tcap] dets_embs @ trk_embs.T by (detection row, list position)
    int* bornslot;                 // [S][Dcap]      slot given to the k-th birth of the frame
    int* epos;                     // [S][Tcap]      list position at contraction time (after a NaN-track removal)
    double* dP;                    // [S][2][49][Tcap] dense 7x7 covariance, live then frozen copy, of the tracks with OF_DENSE
    double* wbuf;                  // [S][Dcap + Tcap] row / column weights of compute_aw_max_metric
    double* deep_d;                // [S][Dcap + 1]  row minima of the materialised first round + the matrix maximum
    int* deep_i;                   // [S][3 Dcap + Tcap] row arg-minima, IoU-hit counts per row, hit column, hits per column
};

// The state of streams [s0, s0 + n) as an engine of its own: every array is stream-major, so a
// group of streams is the same layout behind offset base pointers.  Used to step independent
// groups of streams on separate CUDA streams (capi.cu), E = embedding length (0 without).
inline OcsState ocs_substate(const OcsState& st, int s0, int n, int E) {
    OcsState g = st;
    const size_t o = (size_t)s0, T = st.Tcap, D = st.Dcap;
    g.S = n;
    g.xP += o * XB_N * T;  g.sxP += o * XB_N * T;
    g.lastz += o * 4 * T;  g.lastobs += o * 5 * T;  g.vel += o * 2 * T;  g.confcls += o * 2 * T;
    g.ring += o * st.R * 4 * T;  g.ring_age += o * st.R * T;
    g.ints += o * OI_COUNT * T;  g.order += o * T;  g.hdr += o * OH_COUNT;  g.n_updates += o;
    g.scratch += o * st.scratch_elems;  g.abuf += o * AB_COUNT * T;
    g.upd += o * 2 * D;  g.born += o * D;  g.hsel += o * 2 * D;
    if (st.scan) g.scan += o * D;
    if (st.prof) g.prof += o * 8;
    if (st.emb) g.emb += o * T * E;
    if (st.embsim) g.embsim += o * D * T;
    if (st.bornslot) g.bornslot += o * D;
    if (st.epos) g.epos += o * T;
    if (st.dP) g.dP += o * 98 * T;
    if (st.wbuf) g.wbuf += o * (D + T);
    if (st.deep_d) g.deep_d += o * (D + 1);
    if (st.deep_i) g.deep_i += o * (3 * D + T);
    return g;
}

struct OcsView {   // per-stream base pointers
    double *xP, *sxP, *lastz, *lastobs, *vel, *confcls, *ring, *dP;
    int *ring_age, *ints, *order, *hdr;
    int Tcap, R;
    __device__ OcsView(const OcsState& st, int s) : Tcap(st.Tcap), R(st.R) {
        const size_t T = st.Tcap;
        xP = st.xP + (size_t)s * XB_N * T;
        sxP = st.sxP + (size_t)s * XB_N * T;
        lastz = st.lastz + (size_t)s * 4 * T;
        lastobs = st.lastobs + (size_t)s * 5 * T;
        vel = st.vel + (size_t)s * 2 * T;
        confcls = st.confcls + (size_t)s * 2 * T;
        ring = st.ring + (size_t)s * st.R * 4 * T;
        ring_age = st.ring_age + (size_t)s * st.R * T;
        ints = st.ints + (size_t)s * OI_COUNT * T;
        order = st.order + (size_t)s * T;
        hdr = st.hdr + (size_t)s * OH_COUNT;
        dP = st.dP ? st.dP + (size_t)s * 98 * T : nullptr;
    }
    __device__ __forceinline__ int& I(int f, int slot) const { return ints[f * Tcap + slot]; }
};

__device__ __forceinline__ void load_xP(const double* base, int Tcap, int slot, double* x, double* Pb) {
#pragma unroll
    for (int c = 0; c < 7; c++) x[c] = base[c * Tcap + slot];
#pragma unroll
    for (int c = 0; c < 13; c++) Pb[c] = base[(7 + c) * Tcap + slot];
}
__device__ __forceinline__ void store_xP(double* base, int Tcap, int slot, const double* x, const double* Pb) {
#pragma unroll
    for (int c = 0; c < 7; c++) base[c * Tcap + slot] = x[c];
#pragma unroll
    for (int c = 0; c < 13; c++) base[(7 + c) * Tcap + slot] = Pb[c];
}

// Dense 7x7 covariance of a track with OF_DENSE (which = 0 live, 1 frozen copy); the mean stays in xP / sxP.
__device__ __forceinline__ void load_dense(const OcsView& v, int which, int slot, double* P) {
    const double* b = v.dP + (size_t)which * 49 * v.Tcap + slot;
    for (int c = 0; c < 49; c++) P[c] = b[(size_t)c * v.Tcap];
}
__device__ __forceinline__ void store_dense(const OcsView& v, int which, int slot, const double* P) {
    double* b = v.dP + (size_t)which * 49 * v.Tcap + slot;
    for (int c = 0; c < 49; c++) b[(size_t)c * v.Tcap] = P[c];
}
// Switch a track to the dense covariance (both copies), expanding the block form.
static __device__ __noinline__ void ocs_make_dense(const OcsView& v, int slot) {
    const int T = v.Tcap;
    for (int which = 0; which < 2; which++) {
        const double* src = which ? v.sxP : v.xP;
        double pb[13], P[49];
        for (int c = 0; c < 13; c++) pb[c] = src[(7 + c) * T + slot];
        xysrb_expand(pb, P);
        store_dense(v, which, slot, P);
    }
    v.I(OI_FLAGS, slot) |= OF_DENSE;
}

// Observation `delta_t` steps back (ocsort.py:17-28 / :146-152): the stored box of age
// cur_age-dt for dt = delta_t..1, else the last observation.  Returns false when the
// track has never been observed.
__device__ __forceinline__ bool ocs_lookback(const OcsView& v, int slot, int cur_age, int delta_t,
                                             double* box) {
    if (v.I(OI_HITS, slot) == 0) return false;
    for (int dt = delta_t; dt >= 1; dt--) {
        const int a = cur_age - dt;
        if (a < 0) continue;
        const int r = a % v.R;
        if (v.ring_age[r * v.Tcap + slot] == a) {
#pragma unroll
            for (int c = 0; c < 4; c++) box[c] = v.ring[(r * 4 + c) * v.Tcap + slot];
            return true;
        }
    }
#pragma unroll
    for (int c = 0; c < 4; c++) box[c] = v.lastobs[c * v.Tcap + slot];
    return true;
}

// Dense slow path (tracks with OF_DENSE), kept OUT OF LINE so that the 49-entry covariance never enters the
// register allocation of the common block-form path.
static __device__ __noinline__ void ocs_dense_predict(const OcsView& v, int slot, XysrConst kc, double* x_out) {
    const int T = v.Tcap;
    double x[7], Pd[49];
    for (int c = 0; c < 7; c++) x[c] = v.xP[c * T + slot];
    load_dense(v, 0, slot, Pd);
    if (x[6] + x[2] <= 0) x[6] *= 0.0;
    xysr_predict(x, Pd, kc);
    for (int c = 0; c < 7; c++) { v.xP[c * T + slot] = x[c]; x_out[c] = x[c]; }
    store_dense(v, 0, slot, Pd);
}
static __device__ __noinline__ void ocs_dense_update(const OcsView& v, int slot, const double* z, XysrConst kc) {
    const int T = v.Tcap;
    const int flags = v.I(OI_FLAGS, slot);
    double xd[7], Pd[49];
    if (!(flags & OF_OBSERVED) && (flags & OF_SAVED)) {
        for (int c = 0; c < 7; c++) xd[c] = v.sxP[c * T + slot];
        load_dense(v, 1, slot, Pd);
        double z1[4];
        for (int c = 0; c < 4; c++) z1[c] = v.lastz[c * T + slot];
        xysr_oru_replay(xd, Pd, z1, z, v.I(OI_NNONE, slot) + 1, kc);
    } else {
        for (int c = 0; c < 7; c++) xd[c] = v.xP[c * T + slot];
        load_dense(v, 0, slot, Pd);
    }
    xysr_update(xd, Pd, z);
    for (int c = 0; c < 7; c++) v.xP[c * T + slot] = xd[c];
    store_dense(v, 0, slot, Pd);
    for (int c = 0; c < 4; c++) v.lastz[c * T + slot] = z[c];
    v.I(OI_FLAGS, slot) = flags | OF_OBSERVED;
    v.I(OI_NNONE, slot) = 0;
}
static __device__ __noinline__ void ocs_dense_freeze_copy(const OcsView& v, int slot) {
    double P[49];
    load_dense(v, 0, slot, P);
    store_dense(v, 1, slot, P);
}

// KalmanBoxTracker.update(bbox, cls, det_ind) with a real detection (ocsort.py:136-169)
// followed by KalmanFilterXYSR.update (xysr_kf.py:211-291, ORU at :249-254).
// DENSE: compiled into the DeepOCSORT kernels only (a rotating CMC warp can switch a track to the dense covariance)
template <bool DENSE = false>
__device__ inline void ocs_track_update(const OcsView& v, int slot, const double* det /*x1,y1,x2,y2,conf*/,
                                        double cls, int det_ind, int delta_t, const XysrConst& kc) {
    const int T = v.Tcap;
    v.I(OI_DETIND, slot) = det_ind;
    v.confcls[0 * T + slot] = det[4];
    v.confcls[1 * T + slot] = cls;
    const int age = v.I(OI_AGE, slot);
    // `if self.last_observation.sum() >= 0` -- numpy sums 5 values left to right
    double lo[5];
#pragma unroll
    for (int c = 0; c < 5; c++) lo[c] = v.lastobs[c * T + slot];
    const double lsum = (((lo[0] + lo[1]) + lo[2]) + lo[3]) + lo[4];
    if (lsum >= 0) {
        double pb[4];
        bool found = false;
        for (int dt = delta_t; dt >= 1 && !found; dt--) {
            const int a = age - dt;
            if (a < 0) continue;
            const int r = a % v.R;
            if (v.ring_age[r * T + slot] == a) {
#pragma unroll
                for (int c = 0; c < 4; c++) pb[c] = v.ring[(r * 4 + c) * T + slot];
                found = true;
            }
        }
        if (!found) {
#pragma unroll
            for (int c = 0; c < 4; c++) pb[c] = lo[c];
        }
        // speed_direction(previous_box, bbox)  (ocsort.py:48-53)
        const double cx1 = (pb[0] + pb[2]) / 2.0, cy1 = (pb[1] + pb[3]) / 2.0;
        const double cx2 = (det[0] + det[2]) / 2.0, cy2 = (det[1] + det[3]) / 2.0;
        const double dy = cy2 - cy1, dx = cx2 - cx1;
        const double norm = sqrt(dy * dy + dx * dx) + 1e-6;
        v.vel[0 * T + slot] = dy / norm;
        v.vel[1 * T + slot] = dx / norm;
    }
#pragma unroll
    for (int c = 0; c < 5; c++) v.lastobs[c * T + slot] = det[c];
    {
        const int r = age % v.R;
        v.ring_age[r * T + slot] = age;
#pragma unroll
        for (int c = 0; c < 4; c++) v.ring[(r * 4 + c) * T + slot] = det[c];
    }
    v.I(OI_TSU, slot) = 0;
    v.I(OI_HITS, slot) += 1;
    v.I(OI_STREAK, slot) += 1;

    double z[4];
    xyxy_to_z(det, z);
    int flags = v.I(OI_FLAGS, slot);
    if (DENSE && (flags & OF_DENSE)) { ocs_dense_update(v, slot, z, kc); return; }   // after a rotating CMC warp (slow path)
    double x[7], P[13];
    if (!(flags & OF_OBSERVED) && (flags & OF_SAVED)) {
        load_xP(v.sxP, T, slot, x, P);
        double z1[4];
#pragma unroll
        for (int c = 0; c < 4; c++) z1[c] = v.lastz[c * T + slot];
        xysrb_oru_replay(x, P, z1, z, v.I(OI_NNONE, slot) + 1, kc);
    } else {
        load_xP(v.xP, T, slot, x, P);
    }
    xysrb_update(x, P, z);
    store_xP(v.xP, T, slot, x, P);
#pragma unroll
    for (int c = 0; c < 4; c++) v.lastz[c * T + slot] = z[c];
    v.I(OI_FLAGS, slot) = flags | OF_OBSERVED;
    v.I(OI_NNONE, slot) = 0;
}

// KalmanBoxTracker.update(None, ...) (ocsort.py:170-171 -> xysr_kf.py:233-246)
template <bool DENSE = false>
__device__ inline void ocs_track_miss(const OcsView& v, int slot) {
    const int T = v.Tcap;
    int flags = v.I(OI_FLAGS, slot);
    if (flags & OF_OBSERVED) {
        for (int c = 0; c < XB_N; c++) v.sxP[c * T + slot] = v.xP[c * T + slot];
        if (DENSE && (flags & OF_DENSE)) ocs_dense_freeze_copy(v, slot);
        flags |= OF_SAVED;
    }
    v.I(OI_FLAGS, slot) = flags & ~OF_OBSERVED;
    v.I(OI_NNONE, slot) += 1;
}

// KalmanBoxTracker.__init__ (ocsort.py:63-134)
__device__ inline void ocs_track_init(const OcsView& v, int slot, const double* det, double cls,
                                      int det_ind, int id) {
    const int T = v.Tcap;
    double z[4], x[7], P[13];
    xyxy_to_z(det, z);
    xysrb_init(x, P, z);
    store_xP(v.xP, T, slot, x, P);
    v.I(OI_ID, slot) = id;
    v.I(OI_TSU, slot) = 0;
    v.I(OI_HITS, slot) = 0;
    v.I(OI_STREAK, slot) = 0;
    v.I(OI_AGE, slot) = 0;
    v.I(OI_DETIND, slot) = det_ind;
    v.I(OI_FLAGS, slot) = 0;
    v.I(OI_NNONE, slot) = 0;
    v.confcls[0 * T + slot] = det[4];
    v.confcls[1 * T + slot] = cls;
#pragma unroll
    for (int c = 0; c < 5; c++) v.lastobs[c * T + slot] = -1.0;
    v.vel[0 * T + slot] = 0.0;
    v.vel[1 * T + slot] = 0.0;
    for (int r = 0; r < v.R; r++) v.ring_age[r * T + slot] = -1;
}

}  // namespace bm

namespace bm {
size_t ocsort_step_smem(const OcsState& st, const OcsParams& pr);
cudaError_t ocsort_step_configure(size_t smem);
cudaError_t ocsort_step_launch(const OcsState& st, const OcsParams& pr, const float* dets, const int* ndets,
                               double* out, int* nout, cudaStream_t stream, cudaEvent_t* ev = nullptr);
// DeepOCSORT frame step: embs [S][Dcap][E] float32 (ignored when embedding is off), warps
// [S][6] float64 row-major 2x3 or nullptr (no CMC correction this frame).
cudaError_t deepocsort_step_launch(const OcsState& st, const OcsParams& pr, const float* dets, const int* ndets,
                                   const float* embs, const double* warps, double* out, int* nout,
                                   cudaStream_t stream, cudaEvent_t* ev = nullptr);
}  // namespace bm
