// Device-resident, stream-batched StrongSORT engine (the fork's "enhanced" StrongSORT in its
// runnable configuration: SURVEY 8(c) R4 / M1 / M2 / R5) and its C-ABI (bm_strongsort_*).
//
// Replaces StrongSort.update (boxmot/trackers/strongsort/strongsort.py:115-205, :301-327, :329-359),
// Tracker.predict / update / _enhanced_match / _attempt_id_recovery / _initiate_track
// (sort/tracker.py:102-111, :118-181, :183-281, :300-351, :395-410), matching_cascade /
// min_cost_matching / gate_cost_matrix and helpers, NearestNeighborDistanceMetric.distance /
// partial_fit (sort/linear_assignment.py:14-87, :90-171, :174-379, :538-618), iou_cost
// (sort/iou_matching.py:49-87) and Track (sort/track.py:78-317).
//
// One frame of all S streams = six launches, no host synchronisation:
//   K1 ss_prep_kernel     CTA per stream: min_conf gate, detection quality (:301-327), stable sort
//                         by quality (:190), camera_update + Track.predict for every track
//   K2 cosine_tc_kernel   tcgen05 contraction: every stored bank feature x every detection
//   K3 ss_cost_kernel     gated_metric + _enhance_cost_matrix for (confirmed track, detection)
//   K4 ss_match_kernel    warp per stream: the three-stage cascade; every assignment is solved by a
//                         step-for-step emulation of scipy's rectangular linear_sum_assignment
//   K5 ss_update_kernel   warp per updated track: adaptive feature EMA + NSA Kalman update + scores
//   K6 ss_finalize_kernel CTA per stream: mark_missed, id recovery, births, lost buffer,
//                         partial_fit's first pass (which targets have real appends), output rows
//   K6b ss_bank_kernel    warp per listed target: partial_fit's appends / evictions on a register-resident list
//   K6c ss_live_kernel    CTA per stream: list of live bank features for the next frame
//
// Feature storage.  The reference re-adds all (<= 10) features of every confirmed track to the
// target's sample list EVERY frame, so a sample list of `budget` entries holds ~budget/10 distinct
// vectors.  Here a feature vector is stored once, in a per-stream POOL of slots (fp64 for the EMA
// chain and the quality norm, an fp32 shadow for the contraction); the track's ring of 10 and
// its sample list hold slot numbers.  partial_fit is then integer bookkeeping, and the
// contraction reads every distinct vector once per frame instead of once per copy.
//
// partial_fit's "keep the best `budget` by np.linalg.norm(feature)" ranks unit vectors by the last
// ulp of a BLAS dot product.  The engine evaluates that norm in the summation order of OpenBLAS'
// SkylakeX ddot kernel (warp_ddot_skx; the same order as oracle/numerics.c), which is what the
// reference executed when the fixtures were generated; within one quality value list order is
// insertion order (Python's sort is stable), so "drop the worst" = lowest quality, newest entry.
#include "../../include/boxmot_b200.h"
#include "strongsort_units.cuh"
#include "cosine_tc.cuh"
#include <algorithm>
#include <cstdio>
#include <new>
#include <vector>

namespace bm {

void set_last_error(const char* msg);

enum { SH_NTRACKS = 0, SH_NLOST, SH_NTFREE, SH_NPFREE, SH_NPFREED, SH_NEXT_ID, SH_FRAME, SH_ERR, SH_NHIST, SH_NLIVE,
       SH_NDET, SH_NUPD, SH_STAGE3, SH_NUNM, SH_LSA, SH_MATCH, SH_BIRTHS, SH_HADDETS, SH_NWORK, SH_COUNT = 20 };
enum { TI_ID = 0, TI_NRING, TI_NBANK, TI_NDIST, TI_LOSTFRAME, TI_SEQ, TI_NUPD, TI_MISS, TI_FEAS, TI_COUNT };
// sorted detection table
enum { DT_TLWH = 0, DT_XYAH = 4, DT_CONF = 8, DT_CLS, DT_DETIND, DT_QUALITY, DT_NORM, DT_COUNT };
constexpr int SS_RING = 10;          // track.py:109 max_features
constexpr int SS_LOST = 30;          // tracker.py:100 id_recovery_buffer_size
constexpr int SS_MAXUPD = 3;         // a track can be matched by the cascade and by both of its IoU-stage rows
enum { SS_ERR_TRACKS = 1, SS_ERR_DETS = 2, SS_ERR_POOL = 8, SS_ERR_LSA = 16 };

struct SsParams {
    double min_conf, max_cos_dist, max_iou_dist, mc_lambda, ema_alpha, conf_hi, conf_lo, id_weight;
    int max_age, n_init, budget, born_confirmed, E;
};

struct SsState {
    int s0 = 0;                                // first stream of this launch (stream groups: S = streams of the group)
    int S, Tcap, Dcap, Pcap, B, M;       // M = 2 * Tcap + Dcap: bound of an assignment problem's side
    double *mean, *covb, *tab;           // [S][Tcap][8], [S][Tcap][16], [S][Tcap][ST_FIELDS]
    int *tint, *ring;                    // [S][Tcap][TI_COUNT], [S][Tcap][SS_RING]
    int *bslot, *bseq, *dslot, *dcnt;    // [S][Tcap][B] sample list entries (slot, seq); distinct slots and counts
    double* bminq;                       // [S][Tcap]
    int *upd, *order, *lost, *tfree, *hdr;   // [S][Tcap][3], [S][Tcap], [S][32], [S][Tcap], [S][SH_COUNT]
    unsigned long long* n_updates;       // [S]
    double* pool; float* pool32;         // [S][Pcap][E]
    double* pq;                          // [S][Pcap] np.linalg.norm of the stored vector
    int *pbank, *pring, *pfree, *pfreed, *live;   // [S][Pcap]
    float *lnorm, *dist;                 // [S][Pcap], [S][Pcap][Dcap]
    double* cost;                        // [S][Tcap][Dcap]
    double* dtab; int* dord; float* dnormf;   // [S][Dcap][DT_COUNT], [S][Dcap], [S][Dcap]
    int *ulist, *unm, *wki;              // [S][Dcap], [S][Dcap], [S][WKI]
    double *wkd, *sims;                  // [S][3*M], [S][32][Dcap]
    size_t wki_stride;
};

struct SsView {
    const SsState& st; int s;
    __device__ SsView(const SsState& st_, int s_) : st(st_), s(s_) {}
    __device__ int* hdr() const { return st.hdr + (size_t)s * SH_COUNT; }
    __device__ int* order() const { return st.order + (size_t)s * st.Tcap; }
    __device__ double* mean(int t) const { return st.mean + ((size_t)s * st.Tcap + t) * 8; }
    __device__ double* covb(int t) const { return st.covb + ((size_t)s * st.Tcap + t) * XA_P; }
    __device__ double* tab(int t) const { return st.tab + ((size_t)s * st.Tcap + t) * ST_FIELDS; }
    __device__ int* tint(int t) const { return st.tint + ((size_t)s * st.Tcap + t) * TI_COUNT; }
    __device__ int* ring(int t) const { return st.ring + ((size_t)s * st.Tcap + t) * SS_RING; }
    __device__ int* bslot(int t) const { return st.bslot + ((size_t)s * st.Tcap + t) * st.B; }
    __device__ int* bseq(int t) const { return st.bseq + ((size_t)s * st.Tcap + t) * st.B; }
    __device__ int* dslot(int t) const { return st.dslot + ((size_t)s * st.Tcap + t) * st.B; }
    __device__ int* dcnt(int t) const { return st.dcnt + ((size_t)s * st.Tcap + t) * st.B; }
    __device__ double& bminq(int t) const { return st.bminq[(size_t)s * st.Tcap + t]; }
    __device__ int* upd(int t) const { return st.upd + ((size_t)s * st.Tcap + t) * SS_MAXUPD; }
    __device__ double* pool(int p, int E) const { return st.pool + ((size_t)s * st.Pcap + p) * E; }
    __device__ float* pool32(int p, int E) const { return st.pool32 + ((size_t)s * st.Pcap + p) * E; }
    __device__ double* pq() const { return st.pq + (size_t)s * st.Pcap; }
    __device__ int* pbank() const { return st.pbank + (size_t)s * st.Pcap; }
    __device__ int* pring() const { return st.pring + (size_t)s * st.Pcap; }
    __device__ int* pfree() const { return st.pfree + (size_t)s * st.Pcap; }
    __device__ int* pfreed() const { return st.pfreed + (size_t)s * st.Pcap; }
    __device__ double* dtab(int k) const { return st.dtab + ((size_t)s * st.Dcap + k) * DT_COUNT; }
    __device__ int* dord() const { return st.dord + (size_t)s * st.Dcap; }
    __device__ double* cost() const { return st.cost + (size_t)s * st.Tcap * st.Dcap; }
    __device__ float* dist() const { return st.dist + (size_t)s * st.Pcap * st.Dcap; }
};

// ---------------------------------------------------------------------------------------------
// x . y in the summation order of OpenBLAS' SkylakeX ddot (see the file header): lane L owns the
// FMA accumulator of elements i = L (mod 32); lanes l and l + 4 of each group of eight are
// folded, the four groups added left to right, then (s0 + s2) + (s1 + s3); the tail beyond the
// last multiple of 32 is folded in element by element.  All lanes return the result.
template <typename FX, typename FY>
__device__ __forceinline__ double warp_ddot_skx(FX x, FY y, int n) {
    const int lane = lane_id();
    const int n1 = n & ~31;
    double acc = 0.0;
    for (int i = lane; i < n1; i += 32) acc = __fma_rn(x(i), y(i), acc);
    double dot = 0.0;
    if (n1) {
        const double a = acc + __shfl_down_sync(BM_FULL, acc, 4);       // valid on lanes 8k + l, l < 4
        const double a1 = __shfl_sync(BM_FULL, a, (lane & 3) + 8), a2 = __shfl_sync(BM_FULL, a, (lane & 3) + 16),
                     a3 = __shfl_sync(BM_FULL, a, (lane & 3) + 24), a0 = __shfl_sync(BM_FULL, a, lane & 3);
        const double sl = ((a0 + a1) + a2) + a3;                        // s[lane & 3]
        const double s0 = __shfl_sync(BM_FULL, sl, 0), s1 = __shfl_sync(BM_FULL, sl, 1),
                     s2 = __shfl_sync(BM_FULL, sl, 2), s3 = __shfl_sync(BM_FULL, sl, 3);
        dot = (s0 + s2) + (s1 + s3);
    }
    for (int i = n1; i < n; i++) dot = __fma_rn(y(i), x(i), dot);
    return dot;
}

__device__ __forceinline__ void ss_set_err(int* hdr, int bit) { atomicOr(&hdr[SH_ERR], bit); }

// Pool slot allocation (pop only; frees are deferred to the finalize kernel's merge, so pops and
// pushes never run concurrently).  Returns -1 and sets the error bit when the pool is exhausted.
__device__ __forceinline__ int ss_pool_pop(const SsView& v) {
    const int n = atomicSub(&v.hdr()[SH_NPFREE], 1);
    if (n <= 0) { atomicAdd(&v.hdr()[SH_NPFREE], 1); ss_set_err(v.hdr(), SS_ERR_POOL); return -1; }
    return v.pfree()[n - 1];
}
__device__ __forceinline__ void ss_pool_release(const SsView& v, int p) {      // deferred
    const int k = atomicAdd(&v.hdr()[SH_NPFREED], 1);
    v.pfreed()[k] = p;
}

// ---------------------------------------------------------------------------------------------
// K1: detections (gate, quality, stable sort by quality) and tracks (camera_update, predict)
__global__ void __launch_bounds__(256)
ss_prep_kernel(SsState st, SsParams pr, const float* __restrict__ dets, const int* __restrict__ ndets,
               const float* __restrict__ embs, const double* __restrict__ warps) {
    extern __shared__ __align__(16) unsigned char ss_smem[];
    const int s = st.s0 + blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    SsView v(st, s);
    int* hdr = v.hdr();
    double* q = reinterpret_cast<double*>(ss_smem);           // [Dcap]
    double* nrm = q + st.Dcap;                                // [Dcap]
    int* sel = reinterpret_cast<int*>(nrm + st.Dcap);         // [Dcap]
    __shared__ int s_n;
    int nd = ndets[s];
    if (nd > st.Dcap) { if (tid == 0) ss_set_err(hdr, SS_ERR_DETS); nd = st.Dcap; }
    const float* D = dets + (size_t)s * st.Dcap * 6;
    if (wid == 0) {
        const int n = warp_compact(nd, sel, 0, [&](int i) { return (double)D[i * 6 + 4] >= pr.min_conf; },
                                   [&](int i) { return i; });
        if (lane == 0) s_n = n;
    }
    __syncthreads();
    const int n = s_n;
    // feature norms (np.linalg.norm of the raw feature: strongsort.py:308, track.py:114, :240)
    for (int k = wid; k < n; k += blockDim.x >> 5) {
        const float* f = embs + ((size_t)s * st.Dcap + sel[k]) * pr.E;
        auto fx = [&](int i) { return (double)f[i]; };
        const double d = warp_ddot_skx(fx, fx, pr.E);
        if (lane == 0) nrm[k] = sqrt(d);
    }
    __syncthreads();
    for (int k = tid; k < n; k += blockDim.x) {              // _compute_detection_quality (:301-327)
        const float* b = D + sel[k] * 6;
        const double conf = (double)b[4];
        const double w = (double)b[2] - (double)b[0], h = (double)b[3] - (double)b[1];
        double ql = 0.7 * conf + 0.3 * fmin(nrm[k] / 10.0, 1.0);
        if (h > 0.0) ql = 0.9 * ql + 0.1 * fmax(0.1, 1.0 - fabs(w / h - 0.5) / 2.0);
        q[k] = ql;
    }
    __syncthreads();
    for (int k = tid; k < n; k += blockDim.x) {              // detections.sort(key=quality, reverse=True): stable
        const double qk = q[k];
        int rank = 0;
        for (int j = 0; j < n; j++) rank += (q[j] > qk || (q[j] == qk && j < k)) ? 1 : 0;
        const float* b = D + sel[k] * 6;
        double* d = v.dtab(rank);
        const double x1 = b[0], y1 = b[1], w = (double)b[2] - (double)b[0], h = (double)b[3] - (double)b[1];
        d[DT_TLWH] = x1; d[DT_TLWH + 1] = y1; d[DT_TLWH + 2] = w; d[DT_TLWH + 3] = h;
        d[DT_XYAH] = x1 + w / 2.0; d[DT_XYAH + 1] = y1 + h / 2.0; d[DT_XYAH + 2] = w / h; d[DT_XYAH + 3] = h;
        d[DT_CONF] = b[4]; d[DT_CLS] = b[5]; d[DT_DETIND] = sel[k]; d[DT_QUALITY] = qk; d[DT_NORM] = nrm[k];
        v.dord()[rank] = sel[k];
        st.dnormf[(size_t)s * st.Dcap + rank] = (float)(nrm[k] + 1e-8);
    }
    // tracks: camera_update (strongsort.py:168-172, only with detections) and Track.predict
    const int nT = hdr[SH_NTRACKS];
    const double ident[6] = {1, 0, 0, 0, 1, 0};
    for (int p = tid; p < nT; p += blockDim.x) {
        const int t = v.order()[p];
        double* a = v.tab(t);
        double* m = v.mean(t);
        if (n > 0) {
            double w[6];
            for (int c = 0; c < 6; c++) w[c] = warps ? warps[(size_t)s * 6 + c] : ident[c];
            ss_track_camera(m, a, w);
        }
        double x[8], P[XA_P];
        double* C = v.covb(t);
        for (int c = 0; c < 8; c++) x[c] = m[c];
        for (int c = 0; c < XA_P; c++) P[c] = C[c];
        ss_track_predict(x, P, a);
        for (int c = 0; c < 8; c++) m[c] = x[c];
        for (int c = 0; c < XA_P; c++) C[c] = P[c];
        int* ti = v.tint(t);
        ti[TI_NUPD] = 0; ti[TI_MISS] = 0; ti[TI_FEAS] = 0;
    }
    if (tid == 0) {
        hdr[SH_NDET] = n;
        hdr[SH_FRAME] += 1;
        hdr[SH_NUPD] = 0; hdr[SH_STAGE3] = 0; hdr[SH_NUNM] = 0;
        st.n_updates[s] += (unsigned long long)nT;
    }
}

// K2 problem: every live bank feature (rows gathered through `live`) against the sorted detections
struct SsBankProblem {
    SsState st; int E; const float* embs;
    __device__ CosineProblem operator()(int ls) const {
        CosineProblem p;
        const int s = st.s0 + ls;
        const int* hdr = st.hdr + (size_t)s * SH_COUNT;
        p.A = st.pool32 + (size_t)s * st.Pcap * E; p.ra = st.live + (size_t)s * st.Pcap;
        p.na = st.lnorm + (size_t)s * st.Pcap; p.nA = hdr[SH_NLIVE];
        p.B = embs + (size_t)s * st.Dcap * E; p.rb = st.dord + (size_t)s * st.Dcap;
        p.nb = st.dnormf + (size_t)s * st.Dcap; p.nB = hdr[SH_NDET];
        p.E = E; p.out = st.dist + (size_t)s * st.Pcap * st.Dcap; p.out_ld = st.Dcap; p.raw = 0;
        return p;
    }
};

// K3: gated_metric (tracker.py:186-206) followed by _enhance_cost_matrix (linear_assignment.py:251-273,
// the first thing min_cost_matching does to the metric's matrix) for every (confirmed track, detection).
//
// Exact screen.  An entry is lam * app + (1 - lam) * motion, motion >= 0, times factors of which the track's
// are known and the detection's are >= 0.8 * 0.9 * 0.9; it is only ever compared with the stage thresholds
// (<= max_cos_dist) and clamped above them (min_cost_matching, :67).  So a pair whose appearance term alone puts
// it beyond the threshold needs no Mahalanobis / motion arithmetic -- any value above the threshold is the same
// entry -- and a row without a pair that survives the screen needs none of the track's Kalman state either.
// (1 - 1e-9 covers the roundings of the real chain.)
__device__ __forceinline__ double ss_track_screen(const double* a, double mc_lambda, double id_weight) {
    double lam = mc_lambda + (1.0 - mc_lambda) * fmin(a[SS_AGE] / 10.0, 1.0) * 0.1;
    lam = lam * (0.8 + 0.4 * a[SS_MOTION]);
    if (a[SS_APPEAR] < 0.5) lam = fmin(lam * 1.2, 0.99);
    lam = fmin(fmax(lam, 0.1), 0.99);
    double f = lam * (0.8 * 0.9 * 0.9) * (1.0 - 1e-9);
    if (a[SS_QUALITY] > 0.8) f *= 0.95;
    if (a[SS_HITS] > 10.0 && a[SS_TSU] == 0.0) f *= 0.98;
    if (a[SS_HI_STREAK] > 3.0) f *= 0.97;
    if (a[SS_LO_STREAK] > 2.0) f *= 1.05;
    if (id_weight > 0.0) f *= 1.0 - id_weight * fmin(a[SS_HITS] / 10.0, 1.0);
    return f;                                                    // <= 0 (id_weight > 1): no screening
}

// One row in full: the projection of the Kalman state, its reciprocal square roots, the adaptive lambda and the
// track's adjustment flags are evaluated once per row (ss_track_terms), lanes over the detections.  Out of line:
// the kernel is held to 40 registers for six CTAs per SM (screened workload 0.66 -> 0.54 ms; a workload whose rows
// are all open pays 4 % for this function's spills: 1.36 -> 1.42 ms).
__device__ __noinline__ void ss_cost_row_full(const SsState& st, const SsParams& pr, int s, int p, int t, int n,
                                              const double* dz, double screen) {
    SsView v(st, s);
    const int Dc = st.Dcap, lane = lane_id();
    const double* dq = dz + 4 * Dc;
    const double* dcf = dq + Dc;
    const double* dcl = dcf + Dc;
    const float* dist = v.dist();
    double* cost = v.cost();
    int* ti = v.tint(t);
    const SsTrackTerms tt = ss_track_terms(v.mean(t), v.covb(t), v.tab(t), pr.mc_lambda, pr.id_weight);
    const int nbank = ti[TI_NBANK], ndist = ti[TI_NDIST];
    const int* ds = v.dslot(t);
    int bits = 0;
    for (int k = lane; k < n; k += 32) {
        double app = SS_INFTY_COST;                          // metric.distance: no samples (:616-617)
        if (nbank > 0) {
            float mn = CUDART_INF_F;
            for (int i = 0; i < ndist; i++) mn = fminf(mn, dist[(size_t)ds[i] * Dc + k]);
            app = (double)mn;
        }
        if (screen > 0.0 && app * screen > pr.max_cos_dist) { cost[(size_t)p * Dc + k] = SS_INFTY_COST; continue; }
        const double z[4] = {dz[k], dz[Dc + k], dz[2 * Dc + k], dz[3 * Dc + k]};
        const double cf = dcf[k];
        const double c = ss_pair_cost(tt, app, z, dq[k], cf, dcl[k]);
        cost[(size_t)p * Dc + k] = c;
        // bit 0 / 1: some high- / medium-confidence detection is within the stage's threshold of this track
        // (the cascade skips the assignment of a level none of whose tracks has the stage's bit)
        if (cf >= pr.conf_hi) { if (c <= pr.max_cos_dist * 0.8) bits |= 1; }
        else if (cf >= pr.conf_lo) { if (c <= pr.max_cos_dist) bits |= 2; }
    }
    for (int o = 16; o > 0; o >>= 1) bits |= __shfl_xor_sync(BM_FULL, bits, o);
    if (lane == 0) ti[TI_FEAS] |= bits | 4;                      // bit 2: this row of the matrix is stored
}

// A warp takes 32 rows at a time: each lane first fetches ONE row's list slot, state, screen factor and first
// sample slot (32 rows' dependent loads in flight at once), then the warp walks the rows with the lanes over the
// detections; the frame's detection columns sit in shared memory as seven arrays.  A row without a surviving pair
// is not written at all: the cascade reads it as the clamp constant (TI_FEAS bit 2 clear, SsClampCost::stored).
__global__ void __launch_bounds__(256, 6)
ss_cost_kernel(SsState st, SsParams pr) {
    extern __shared__ double ss_cost_sm[];
    const int s = st.s0 + blockIdx.y;
    SsView v(st, s);
    const int nT = v.hdr()[SH_NTRACKS], n = v.hdr()[SH_NDET];
    if (nT <= 0 || n <= 0) return;
    const int Dc = st.Dcap;
    double* dz = ss_cost_sm;                     // [4][Dc] xyah, then quality, conf, cls
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        const double* d = v.dtab(k);
        dz[k] = d[DT_XYAH]; dz[Dc + k] = d[DT_XYAH + 1]; dz[2 * Dc + k] = d[DT_XYAH + 2]; dz[3 * Dc + k] = d[DT_XYAH + 3];
        dz[4 * Dc + k] = d[DT_QUALITY]; dz[5 * Dc + k] = d[DT_CONF]; dz[6 * Dc + k] = d[DT_CLS];
    }
    __syncthreads();
    const float* dist = v.dist();
    const int lane = lane_id(), wpb = blockDim.x >> 5;
    for (int base = (blockIdx.x * wpb + (threadIdx.x >> 5)) * 32; base < nT; base += gridDim.x * wpb * 32) {
        const int p = base + lane;
        int t = -1, nbank = 0, ndist = 0, ds0 = 0;
        double screen = 0.0;
        bool conf = false;
        if (p < nT) {
            t = v.order()[p];
            const double* a = v.tab(t);
            conf = a[ST_STATE] == (double)ST_CONFIRMED_STATE;
            if (conf) {
                const int* ti = v.tint(t);
                nbank = ti[TI_NBANK]; ndist = ti[TI_NDIST];
                if (nbank > 0 && ndist > 0) ds0 = v.dslot(t)[0];
                screen = ss_track_screen(a, pr.mc_lambda, pr.id_weight);
            }
        }
        unsigned cm = __ballot_sync(BM_FULL, conf);
        while (cm) {
            const int r = __ffs(cm) - 1;
            cm &= cm - 1;
            const int rt = __shfl_sync(BM_FULL, t, r), rnb = __shfl_sync(BM_FULL, nbank, r),
                      rnd = __shfl_sync(BM_FULL, ndist, r), rds0 = __shfl_sync(BM_FULL, ds0, r);
            const double rsc = shfl_d(screen, r);
            const int rp = base + r;
            bool open = !(rsc > 0.0) || (rnb > 0 && rnd <= 0);
            if (!open) {
                const int* ds = v.dslot(rt);
                for (int k = lane; k < n; k += 32) {
                    double app = SS_INFTY_COST;
                    if (rnb > 0) {
                        float mn = dist[(size_t)rds0 * Dc + k];
                        for (int i = 1; i < rnd; i++) mn = fminf(mn, dist[(size_t)ds[i] * Dc + k]);
                        app = (double)mn;
                    }
                    if (!(app * rsc > pr.max_cos_dist)) open = true;
                }
            }
            if (__any_sync(BM_FULL, open)) ss_cost_row_full(st, pr, s, rp, rt, n, dz, rsc);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// scipy.optimize.linear_sum_assignment (rectangular_lsap.cpp: Crouse's shortest augmenting path),
// emulated step for step by one warp so that ties -- duplicated candidate rows of the IoU stage,
// clamped entries -- resolve as in the reference.  The caller passes nr <= nc (it transposes as
// scipy does); col4row[i] is the column of row i on return.
struct LsaWork {
    double *u, *v, *spc;                       // [nr] [nc] [nc]
    int *path, *col4row, *row4col, *remaining, *srl, *scl;
};
__device__ __forceinline__ LsaWork lsa_carve(double* wd, int* wi, int M) {
    LsaWork w;
    w.u = wd; w.v = wd + M; w.spc = wd + 2 * M;
    w.path = wi; w.col4row = wi + M; w.row4col = wi + 2 * M; w.remaining = wi + 3 * M; w.srl = wi + 4 * M;
    w.scl = wi + 5 * M;
    return w;
}

template <typename CostFn>
__device__ int lsa_scipy_warp(const CostFn& cost, int nr, int nc, const LsaWork& w) {
    const int lane = lane_id();
    for (int i = lane; i < nr; i += 32) { w.u[i] = 0.0; w.col4row[i] = -1; }
    for (int j = lane; j < nc; j += 32) { w.v[j] = 0.0; w.row4col[j] = -1; }
    __syncwarp();
    for (int cur = 0; cur < nr; cur++) {
        for (int j = lane; j < nc; j += 32) { w.remaining[j] = nc - j - 1; w.spc[j] = CUDART_INF; }
        __syncwarp();
        double min_val = 0.0;
        int num_remaining = nc, i = cur, nsr = 0, nsc = 0, sink = -1;
        while (sink == -1) {
            if (lane == 0) w.srl[nsr] = i;
            nsr++;
            const double ui = w.u[i];
            double bv = CUDART_INF;
            int bit = 0x7fffffff, bun = 0;
#pragma unroll 4
            for (int it = lane; it < num_remaining; it += 32) {
                const int j = w.remaining[it];
                const double r = ((min_val + cost(i, j)) - ui) - w.v[j];
                double sj = w.spc[j];
                if (r < sj) { sj = r; w.spc[j] = r; w.path[j] = i; }
                const int un = w.row4col[j] == -1;
                if (sj < bv || (sj == bv && un)) { bv = sj; bit = it; bun = un; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ov = shfl_xor_d(bv, o);
                const int oit = __shfl_xor_sync(BM_FULL, bit, o), oun = __shfl_xor_sync(BM_FULL, bun, o);
                bool take;
                if (ov != bv) take = ov < bv;
                else if (oun != bun) take = oun > bun;                 // an unassigned column wins a tie ...
                else take = oun ? (oit > bit) : (oit < bit);           // ... the last such; else the first column
                if (take) { bv = ov; bit = oit; bun = oun; }
            }
            if (!(bv < CUDART_INF)) return 1;                          // infeasible (NaN costs)
            min_val = bv;
            const int j = w.remaining[bit];
            __syncwarp();
            if (lane == 0) { w.remaining[bit] = w.remaining[num_remaining - 1]; w.scl[nsc] = j; }
            num_remaining--;
            nsc++;
            const int rj = w.row4col[j];
            __syncwarp();
            if (rj == -1) sink = j; else i = rj;
        }
        // dual variables
        for (int k = lane; k < nsr; k += 32) {
            const int r = w.srl[k];
            if (r == cur) w.u[r] += min_val;
            else w.u[r] += min_val - w.spc[w.col4row[r]];
        }
        for (int k = lane; k < nsc; k += 32) { const int j = w.scl[k]; w.v[j] -= min_val - w.spc[j]; }
        __syncwarp();
        if (lane == 0) {                                               // augment
            int j = sink;
            while (true) {
                const int r = w.path[j];
                w.row4col[j] = r;
                const int t = w.col4row[r];
                w.col4row[r] = j;
                j = t;
                if (r == cur) break;
            }
        }
        __syncwarp();
    }
    return 0;
}

// min_cost_matching (linear_assignment.py:14-87) on rows `rows[0..nr)` (track list positions, duplicates
// allowed) and columns `cols[0..nc)` (sorted detection indices) of the stored matrix `c` (by list
// position and detection): clamp at max_distance + 1e-5, assign, reject pairs above max_distance.
// match_col[r] = accepted column position of row r or -1; the unmatched detections are returned in
// `cols` in scipy's order (never-assigned columns ascending, then the rejected ones in row order).
struct SsClampCost {
    const double* c; int ld; const int* rows; const int* cols; double maxd; bool tr;
    const int* far;      // optional, by list position: the whole row is above max_distance (never stored)
    const int* stored;   // optional, pinfo by list position: bit 22 clear = the cost kernel screened the whole row out
    __device__ __forceinline__ double at(int r, int cpos) const {
        const int p = rows[r];
        if (far && far[p]) return maxd + 1e-5;
        if (stored && !((stored[p] >> 22) & 1)) return maxd + 1e-5;
        const double x = c[(size_t)p * ld + cols[cpos]];
        return x > maxd ? maxd + 1e-5 : x;
    }
    __device__ __forceinline__ double operator()(int i, int j) const { return tr ? at(j, i) : at(i, j); }
};

__device__ int ss_min_cost_matching(const double* c, int ld, const int* rows, int nr, int* cols, int& nc, double maxd,
                                    const LsaWork& w, int* match_col, int* tmp, const int* far = nullptr,
                                    const int* stored = nullptr) {
    const int lane = lane_id();
    const bool tr = nc < nr;
    SsClampCost cf{c, ld, rows, cols, maxd, tr, far, stored};
    if (lsa_scipy_warp(cf, tr ? nc : nr, tr ? nr : nc, w)) return 1;
    // row -> column of the original orientation, into match_col (w.srl is free after the solve)
    for (int r = lane; r < nr; r += 32) match_col[r] = tr ? w.row4col[r] : w.col4row[r];
    __syncwarp();
    // unmatched detections: columns without a row, ascending; then rejected assignments in row order
    int* colrow = w.scl;                                               // col position -> row or -1
    for (int j = lane; j < nc; j += 32) colrow[j] = tr ? w.col4row[j] : w.row4col[j];
    __syncwarp();
    int k = warp_compact(nc, tmp, 0, [&](int j) { return colrow[j] < 0; }, [&](int j) { return cols[j]; });
    k = warp_compact(nr, tmp, k, [&](int r) { return match_col[r] >= 0 && cf.at(r, match_col[r]) > maxd; },
                     [&](int r) { return cols[match_col[r]]; });
    for (int r = lane; r < nr; r += 32)
        if (match_col[r] >= 0) match_col[r] = cf.at(r, match_col[r]) > maxd ? -1 : cols[match_col[r]];
    __syncwarp();
    for (int j = lane; j < k; j += 32) cols[j] = tmp[j];
    nc = k;
    __syncwarp();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// The IoU stage's assignment (rows = a few dozen detections after scipy's transposition, columns = every
// non-tentative track plus the duplicated ones, most of them "far": time_since_update > 1, i.e. the same
// clamped constant for every detection) with the far columns handled IMPLICITLY.  Same result as
// lsa_scipy_warp on the full matrix -- including which of two tied columns is taken -- because:
//   * a far column is never scanned: its dual stays 0 and all far columns share one path cost, so it
//     can only end a search as the sink, and scipy's tie rule (an unassigned column beats an assigned
//     one, the LAST unassigned one in `remaining` order wins) makes an assigned far column lose against
//     any unassigned one -- there always is one, #far >= #rows being a precondition;
//   * `remaining` starts as C-1 .. 0 and only changes by "the last entry moves into the scanned
//     column's place": the position of every column is C-1-j unless it is one of the few that moved in
//     this search (kept in a small list), so the tie rule can be evaluated on the stored (near) columns
//     plus ONE representative of the far class: the unassigned far column at the highest position.
// Validated against scipy on random matrices of this shape (tests/test_gpu_strongsort.py).
constexpr int SSF_NMAX = 352, SSF_RMAX = 128, SSF_MV = 32;
constexpr int SSF_CMAX = 8192;                 // candidate columns (bit sets)
struct FarWork {                               // carved from one per-warp shared-memory block
    double *v, *spc, *u;                       // [N] [N] [R]
    int *pos, *path, *row4col, *col4row, *srl, *scl, *mvc, *mvp;   // [N] [N] [N] [R] [R] [N] [MV] [MV]
    unsigned *near_bits, *asg_bits;            // [CMAX / 32]
};
constexpr size_t SSF_BYTES = (size_t)(2 * SSF_NMAX + SSF_RMAX) * 8 + (size_t)(4 * SSF_NMAX + 2 * SSF_RMAX + 2 * SSF_MV) * 4 +
                             2 * (SSF_CMAX / 32) * 4;
__device__ __forceinline__ FarWork far_carve(unsigned char* b) {
    FarWork w;
    w.v = reinterpret_cast<double*>(b); w.spc = w.v + SSF_NMAX; w.u = w.spc + SSF_NMAX;
    w.pos = reinterpret_cast<int*>(w.u + SSF_RMAX); w.path = w.pos + SSF_NMAX; w.row4col = w.path + SSF_NMAX;
    w.scl = w.row4col + SSF_NMAX; w.col4row = w.scl + SSF_NMAX; w.srl = w.col4row + SSF_RMAX;
    w.mvc = w.srl + SSF_RMAX; w.mvp = w.mvc + SSF_MV;
    w.near_bits = reinterpret_cast<unsigned*>(w.mvp + SSF_MV); w.asg_bits = w.near_bits + SSF_CMAX / 32;
    return w;
}

// cost(n, i): clamped cost of near column n (candidate index nearcol[n], ascending) and row i.
// Returns 0 with w.row4col[n] = row of near column n or -1; 1 = needs the explicit solver (too many moves).
template <typename CostFn>
__device__ int lsa_far_warp(const CostFn& cost, const int* nearcol, int N, int C, int R, double cstar, const FarWork& w) {
    const int lane = lane_id();
    for (int k = lane; k < (C + 31) / 32; k += 32) { w.near_bits[k] = 0u; w.asg_bits[k] = 0u; }
    __syncwarp();
    for (int n = lane; n < N; n += 32) { atomicOr(&w.near_bits[nearcol[n] >> 5], 1u << (nearcol[n] & 31)); w.v[n] = 0.0; w.row4col[n] = -1; }
    for (int i = lane; i < R; i += 32) { w.u[i] = 0.0; w.col4row[i] = -1; }
    __syncwarp();
    for (int cur = 0; cur < R; cur++) {
        for (int n = lane; n < N; n += 32) { w.pos[n] = C - 1 - nearcol[n]; w.spc[n] = CUDART_INF; }
        __syncwarp();
        double min_val = 0.0, spcF = CUDART_INF;
        int num = C, i = cur, nsr = 0, nsc = 0, nmv = 0, pathF = -1;
        int sink_near = -1, sink_far = -1;
        while (true) {
            if (lane == 0) w.srl[nsr] = i;
            nsr++;
            const double ui = w.u[i];
            double bv = CUDART_INF;
            int bit = 0x7fffffff, bun = 0, bidx = -1;
            for (int n = lane; n < N; n += 32) {
                const int ps = w.pos[n];
                if (ps < 0) continue;                                   // already scanned in this search
                const double r = ((min_val + cost(n, i)) - ui) - w.v[n];
                double sj = w.spc[n];
                if (r < sj) { sj = r; w.spc[n] = r; w.path[n] = i; }
                const int un = w.row4col[n] == -1;
                bool take;
                if (sj != bv) take = sj < bv;
                else if (un != bun) take = un > bun;
                else take = un ? (ps > bit) : (ps < bit);
                if (take) { bv = sj; bit = ps; bun = un; bidx = n; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ov = shfl_xor_d(bv, o);
                const int oit = __shfl_xor_sync(BM_FULL, bit, o), oun = __shfl_xor_sync(BM_FULL, bun, o);
                const int oidx = __shfl_xor_sync(BM_FULL, bidx, o);
                bool take;
                if (oidx < 0) take = false;
                else if (bidx < 0) take = true;
                else if (ov != bv) take = ov < bv;
                else if (oun != bun) take = oun > bun;
                else take = oun ? (oit > bit) : (oit < bit);
                if (take) { bv = ov; bit = oit; bun = oun; bidx = oidx; }
            }
            // the far class: one value, represented by its unassigned member at the highest position
            const double rF = ((min_val + cstar) - ui) - 0.0;
            if (rF < spcF) { spcF = rF; pathF = i; }
            int pF = -1, cF = -1;
            {
                const int lo = C - num;                                  // unmoved columns still listed have j >= lo
                int found = -1;
                for (int wbase = lo >> 5; wbase < (C + 31) / 32 && found < 0; wbase += 32) {
                    const int wi_ = wbase + lane;
                    unsigned freeb = 0u;
                    if (wi_ < (C + 31) / 32) {
                        freeb = ~(w.near_bits[wi_] | w.asg_bits[wi_]);
                        if (wi_ == (lo >> 5)) freeb &= ~0u << (lo & 31);
                        if (wi_ == (C >> 5)) freeb &= (C & 31) ? ((1u << (C & 31)) - 1u) : 0u;
                    }
                    const unsigned any = __ballot_sync(BM_FULL, freeb != 0u);
                    if (any) {
                        const int l0 = __ffs(any) - 1;
                        const unsigned fb = __shfl_sync(BM_FULL, freeb, l0);
                        found = ((wbase + l0) << 5) + (__ffs(fb) - 1);
                    }
                }
                if (found >= 0) { pF = C - 1 - found; cF = found; }
                for (int k = 0; k < nmv; k++) {                          // moved far columns
                    const int c = w.mvc[k];
                    if (!((w.near_bits[c >> 5] >> (c & 31)) & 1u) && !((w.asg_bits[c >> 5] >> (c & 31)) & 1u) && w.mvp[k] > pF) {
                        pF = w.mvp[k]; cF = c;
                    }
                }
            }
            bool far_wins;
            if (bidx < 0) far_wins = true;
            else if (spcF != bv) far_wins = spcF < bv;
            else if (!bun) far_wins = true;                              // unassigned beats assigned
            else far_wins = pF > bit;
            if (far_wins) { min_val = spcF; sink_far = cF; break; }
            min_val = bv;
            const int n = bidx;
            // the scanned column leaves `remaining`: the entry at the last position moves into its place
            const int h = bit, L = num - 1;
            __syncwarp();
            if (h != L) {
                int colL = C - 1 - L, k_at = -1;
                for (int k = 0; k < nmv; k++) if (w.mvp[k] == L) { colL = w.mvc[k]; k_at = k; }
                if (k_at < 0) {
                    if (nmv >= SSF_MV) return 1;
                    k_at = nmv++;
                    if (lane == 0) w.mvc[k_at] = colL;
                }
                if (lane == 0) w.mvp[k_at] = h;
                if ((w.near_bits[colL >> 5] >> (colL & 31)) & 1u) {       // a near column moved: binary search its index
                    int a = 0, b = N - 1;
                    while (a < b) { const int m = (a + b) >> 1; if (nearcol[m] < colL) a = m + 1; else b = m; }
                    if (lane == 0) w.pos[a] = h;
                }
            }
            __syncwarp();
            {   // forget a list entry of the scanned column itself
                const int cn = nearcol[n];
                int k_rm = -1;
                for (int k = 0; k < nmv; k++) if (w.mvc[k] == cn) k_rm = k;
                if (k_rm >= 0) {
                    if (lane == 0) { w.mvc[k_rm] = w.mvc[nmv - 1]; w.mvp[k_rm] = w.mvp[nmv - 1]; }
                    nmv--;
                }
            }
            num--;
            if (lane == 0) { w.pos[n] = -1; w.scl[nsc] = n; }
            nsc++;
            const int rj = w.row4col[n];
            __syncwarp();
            if (rj == -1) { sink_near = n; break; }
            i = rj;
        }
        __syncwarp();
        // dual variables
        for (int k = lane; k < nsr; k += 32) {
            const int r = w.srl[k];
            if (r == cur) w.u[r] += min_val;
            else w.u[r] += min_val - w.spc[w.col4row[r]];
        }
        for (int k = lane; k < nsc; k += 32) { const int n = w.scl[k]; w.v[n] -= min_val - w.spc[n]; }
        __syncwarp();
        if (lane == 0) {                                                 // augment
            int r, j;
            if (sink_far >= 0) {
                w.asg_bits[sink_far >> 5] |= 1u << (sink_far & 31);
                r = pathF;
                j = w.col4row[r]; w.col4row[r] = -2;
            } else {
                r = w.path[sink_near]; w.row4col[sink_near] = r;
                j = w.col4row[r]; w.col4row[r] = sink_near;
            }
            while (r != cur) {
                const int n = j;
                r = w.path[n]; w.row4col[n] = r;
                j = w.col4row[r]; w.col4row[r] = n;
            }
        }
        __syncwarp();
    }
    return 0;
}

// Unit-testable form: one warp per problem on stored matrices, scipy's own transposition rule.
struct LsaDense {
    const double* c; int ld; bool tr;
    __device__ __forceinline__ double operator()(int i, int j) const { return tr ? c[(size_t)j * ld + i] : c[(size_t)i * ld + j]; }
};
__global__ void __launch_bounds__(128)
lsa_batch_kernel(const double* __restrict__ cost, int n_problems, int nr_max, int nc_max, const int* __restrict__ nr_,
                 const int* __restrict__ nc_, int* __restrict__ x_out, double* wd, int* wi, int M) {
    const int b = blockIdx.x * 4 + (threadIdx.x >> 5), lane = lane_id();
    if (b >= n_problems) return;
    const int nr = nr_[b], nc = nc_[b];
    int* x = x_out + (size_t)b * nr_max;
    for (int i = lane; i < nr_max; i += 32) x[i] = -1;
    if (nr <= 0 || nc <= 0) return;
    LsaWork w = lsa_carve(wd + (size_t)b * 3 * M, wi + (size_t)b * 6 * M, M);
    const bool tr = nc < nr;
    LsaDense cf{cost + (size_t)b * nr_max * nc_max, nc_max, tr};
    if (lsa_scipy_warp(cf, tr ? nc : nr, tr ? nr : nc, w)) return;
    for (int r = lane; r < nr; r += 32) x[r] = tr ? w.row4col[r] : w.col4row[r];
}

// Unit-testable form of lsa_far_warp: rows flagged in `far_rows` are constant (cstar) for every column.
struct FarDense {
    const double* c; int ld; const int* nearcol;
    __device__ __forceinline__ double operator()(int n, int i) const { return c[(size_t)nearcol[n] * ld + i]; }
};
__global__ void __launch_bounds__(64)
lsa_far_batch_kernel(const double* __restrict__ cost, int n_problems, int nr_max, int nc_max, const int* __restrict__ nr_,
                     const int* __restrict__ nc_, const int* __restrict__ far_rows, double cstar, int* __restrict__ x_out,
                     int* __restrict__ used_far, double* wd, int* wi, int M) {
    __shared__ __align__(16) unsigned char s_far[2][SSF_BYTES];
    const int b = blockIdx.x * 2 + (threadIdx.x >> 5), lane = lane_id();
    if (b >= n_problems) return;
    const int nr = nr_[b], nc = nc_[b];
    int* x = x_out + (size_t)b * nr_max;
    for (int i = lane; i < nr_max; i += 32) x[i] = -1;
    if (nr <= 0 || nc <= 0) return;
    const int* fr = far_rows + (size_t)b * nr_max;
    int* nearcol = wi + (size_t)b * 6 * M;                      // scratch: list of near rows
    const int N = warp_compact(nr, nearcol, 0, [&](int r) { return fr[r] == 0; }, [&](int r) { return r; });
    const bool ok = nc < nr && N <= SSF_NMAX && nc <= SSF_RMAX && nr <= SSF_CMAX && nr - N >= nc;
    int fell_back = 1;
    if (ok) {
        FarWork w = far_carve(s_far[threadIdx.x >> 5]);
        FarDense cf{cost + (size_t)b * nr_max * nc_max, nc_max, nearcol};
        if (lsa_far_warp(cf, nearcol, N, nr, nc, cstar, w) == 0) {
            for (int n = lane; n < N; n += 32) x[nearcol[n]] = w.row4col[n];
            fell_back = 0;
        }
        __syncwarp();
    }
    if (fell_back) {
        LsaWork w = lsa_carve(wd + (size_t)b * 3 * M, wi + (size_t)b * 6 * M, M);
        const bool tr = nc < nr;
        LsaDense cf{cost + (size_t)b * nr_max * nc_max, nc_max, tr};
        if (lsa_scipy_warp(cf, tr ? nc : nr, tr ? nr : nc, w)) return;
        for (int r = lane; r < nr; r += 32) x[r] = tr ? w.row4col[r] : w.col4row[r];
    }
    if (lane == 0 && used_far) used_far[b] = fell_back ? 0 : 1;
}

// K4: Tracker._enhanced_match (tracker.py:183-281), one warp per stream
constexpr int SS_MATCH_WPB = 2;
constexpr int SS_LSA_SMEM_M = 160;        // assignment problems with both sides <= this run out of shared memory
__global__ void __launch_bounds__(SS_MATCH_WPB * 32)
ss_match_kernel(SsState st, SsParams pr) {
    // per warp: the explicit solver's work area (3 M doubles + 6 M ints) or the implicit-far solver's (never both)
    static_assert(SSF_BYTES >= (size_t)SS_LSA_SMEM_M * (3 * 8 + 6 * 4), "shared work area");
    __shared__ __align__(16) unsigned char s_lsa[SS_MATCH_WPB][SSF_BYTES];
    const int ls = blockIdx.x * SS_MATCH_WPB + (threadIdx.x >> 5), lane = lane_id();
    if (ls >= st.S) return;
    const int s = st.s0 + ls;
    SsView v(st, s);
    int* hdr = v.hdr();
    const int nT = hdr[SH_NTRACKS], n = hdr[SH_NDET];
    const int Tc = st.Tcap, Dc = st.Dcap, M = st.M;
    // integer work area
    int* wi = st.wki + (size_t)s * st.wki_stride;
    int* tmatched = wi;               wi += Tc;        // by list position: matched in stage 1 / 2
    int* byage = wi;                  wi += Tc;        // confirmed unmatched positions grouped by time_since_update
    int* lvl = wi;                    wi += Tc;        // the level's rows ordered by priority
    int* rows3 = wi;                  wi += 2 * Tc;    // IoU-stage candidate rows
    int* mcol = wi;                   wi += 2 * Tc;    // match column of a row
    int* cols = wi;                   wi += Dc;
    int* tmp = wi;                    wi += 2 * Tc + Dc;
    int* dmatched = wi;               wi += Dc;
    int* agecnt = wi;                 wi += pr.max_age + 2;
    const LsaWork w_glob = lsa_carve(st.wkd + (size_t)s * 3 * M, wi, M);
    unsigned char* s_my = s_lsa[threadIdx.x >> 5];
    const LsaWork w_smem = lsa_carve(reinterpret_cast<double*>(s_my),
                                     reinterpret_cast<int*>(s_my + (size_t)SS_LSA_SMEM_M * 3 * 8), SS_LSA_SMEM_M);
    int* ulist = st.ulist + (size_t)s * Dc;
    int* unm = st.unm + (size_t)s * Dc;
    double* cost = v.cost();

    // per list position, packed once: time_since_update (16 bits) | state << 16 | feasibility bits << 20 |
    // matched-in-stage-1/2 << 24 -- the passes below read this one int instead of chasing order -> table
    int* pinfo = tmatched;
    for (int p = lane; p < nT; p += 32) {
        const int t = v.order()[p];
        const double* a = v.tab(t);
        const int tsu = min((int)a[SS_TSU], 0xffff);
        pinfo[p] = tsu | ((int)a[ST_STATE] << 16) | ((v.tint(t)[TI_FEAS] & 7) << 20);      // bit 22: cost row stored
    }
    for (int k = lane; k < n; k += 32) dmatched[k] = 0;
    __syncwarp();
#define PI_TSU(x) ((x) & 0xffff)
#define PI_STATE(x) (((x) >> 16) & 3)
#define PI_FEAS(x) (((x) >> 20) & 3)
#define PI_MATCHED(x) (((x) >> 24) & 1)
    if (n == 0) {                                   // tracker.update([]): nothing matches, the IoU stage is skipped
        if (lane == 0) { hdr[SH_NUNM] = 0; hdr[SH_STAGE3] = 0; hdr[SH_NUPD] = 0; }
        return;
    }
    int nupd = 0, bad = 0, nlsa = 0, nmatch = 0;
    auto record = [&](int p, int k) {               // lane 0: match (track at list position p, detection k)
        int* ti = v.tint(v.order()[p]);
        const int c = ti[TI_NUPD];
        if (c == 0) ulist[nupd++] = p;
        if (c < SS_MAXUPD) v.upd(v.order()[p])[c] = k;
        ti[TI_NUPD] = c + 1;
        nmatch++;
    };
    // ---- stages 1 and 2: matching_cascade over the appearance metric
    for (int stage = 0; stage < 2; stage++) {
        const double thr = stage == 0 ? pr.max_cos_dist * 0.8 : pr.max_cos_dist;
        int nc = warp_compact(n, cols, 0, [&](int k) {
            const double c = v.dtab(k)[DT_CONF];
            const bool in = stage == 0 ? c >= pr.conf_hi : (c >= pr.conf_lo && c < pr.conf_hi);
            return in && !dmatched[k];
        }, [&](int k) { return k; });
        if (nc == 0) continue;
        // group the confirmed, still unmatched tracks by time_since_update (stable in list order)
        for (int a = lane; a <= pr.max_age + 1; a += 32) agecnt[a] = 0;
        __syncwarp();
        for (int p = lane; p < nT; p += 32) {
            const int pi = pinfo[p], age = PI_TSU(pi);
            if (PI_STATE(pi) == ST_CONFIRMED_STATE && !PI_MATCHED(pi) && age <= pr.max_age) atomicAdd(&agecnt[age], 1);
        }
        __syncwarp();
        {                                           // exclusive prefix (warp scan); agecnt[a] becomes the write cursor of age a
            int run = 0;
            for (int base = 0; base <= pr.max_age; base += 32) {
                const int a = base + lane;
                const int c = a <= pr.max_age ? agecnt[a] : 0;
                int inc = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(BM_FULL, inc, o); if (lane >= o) inc += t; }
                if (a <= pr.max_age) agecnt[a] = run + inc - c;
                run += __shfl_sync(BM_FULL, inc, 31);
            }
            if (lane == 0) agecnt[pr.max_age + 1] = run;
        }
        __syncwarp();
        for (int base = 0; base < nT; base += 32) {
            const int p = base + lane;
            int age = -1;
            if (p < nT) {
                const int pi = pinfo[p], g = PI_TSU(pi);
                if (PI_STATE(pi) == ST_CONFIRMED_STATE && !PI_MATCHED(pi) && g <= pr.max_age) age = g;
            }
            const unsigned peers = __match_any_sync(BM_FULL, age);
            if (age >= 0) {
                const int rank = __popc(peers & ((1u << lane) - 1));
                byage[agecnt[age] + rank] = p;
            }
            __syncwarp();
            if (age >= 0 && (peers & ((1u << lane) - 1)) == 0) agecnt[age] += __popc(peers);
            __syncwarp();
        }
        // agecnt[a] is now the END of age a's run; its start is the end of the previous non-empty age
        int start = 0;
        for (int age = 0; age <= pr.max_age && nc > 0; age++) {
            const int end = agecnt[age];
            const int m = end - start;
            if (m <= 0) continue;
            // A level none of whose tracks is within the threshold of any detection of the stage is a matrix of
            // equal (clamped) entries: scipy assigns row i to column i (transposed: column j to row j), every
            // pair is rejected, and the detections are handed on as [m, nc) followed by [0, m) when m < nc.
            int feas = 0;
            for (int i = lane; i < m; i += 32) feas |= PI_FEAS(pinfo[byage[start + i]]) & (1 << stage);
            feas = __any_sync(BM_FULL, feas != 0);
            if (!feas) {
                if (m < nc) {
                    for (int j = lane; j < nc; j += 32) tmp[j] = cols[j < nc - m ? j + m : j - (nc - m)];
                    __syncwarp();
                    for (int j = lane; j < nc; j += 32) cols[j] = tmp[j];
                    __syncwarp();
                }
                start = end;
                continue;
            }
            // _prioritize_tracks_by_quality (:276-285): stable sort by -(quality + stability)
            for (int i = lane; i < m; i += 32) {
                const int p = byage[start + i];
                const double* a = v.tab(v.order()[p]);
                const double key = -(a[SS_QUALITY] + a[ST_STABILITY]);
                int rank = 0;
                for (int j = 0; j < m; j++) {
                    const double* b = v.tab(v.order()[byage[start + j]]);
                    const double kj = -(b[SS_QUALITY] + b[ST_STABILITY]);
                    rank += (kj < key || (kj == key && j < i)) ? 1 : 0;
                }
                lvl[rank] = p;
            }
            __syncwarp();
            bad |= ss_min_cost_matching(cost, Dc, lvl, m, cols, nc, thr, (m <= SS_LSA_SMEM_M && nc <= SS_LSA_SMEM_M) ? w_smem : w_glob,
                                        mcol, tmp, nullptr, pinfo);
            nlsa++;
            if (lane == 0)
                for (int r = 0; r < m; r++)
                    if (mcol[r] >= 0) { pinfo[lvl[r]] |= 1 << 24; dmatched[mcol[r]] = 1; record(lvl[r], mcol[r]); }
            __syncwarp();
            start = end;
        }
        nupd = __shfl_sync(BM_FULL, nupd, 0);
        nmatch = __shfl_sync(BM_FULL, nmatch, 0);
    }
    // ---- stage 3: IoU matching.  Candidates = every non-tentative track (matched or not), then
    // AGAIN the unmatched ones with time_since_update == 1 (tracker.py:257-259: the list is a concatenation)
    int nr3 = warp_compact(nT, rows3, 0, [&](int p) { return PI_STATE(pinfo[p]) != ST_TENTATIVE; }, [&](int p) { return p; });
    nr3 = warp_compact(nT, rows3, nr3, [&](int p) {
        const int pi = pinfo[p];
        return PI_STATE(pi) == ST_CONFIRMED_STATE && !PI_MATCHED(pi) && PI_TSU(pi) == 1;
    }, [&](int p) { return p; });
    int nc3 = warp_compact(n, cols, 0, [&](int k) { return !dmatched[k] && !(v.dtab(k)[DT_CONF] < pr.conf_lo); },
                           [&](int k) { return k; });
    if (nr3 > 0 && nc3 > 0) {
        // iou_cost (iou_matching.py:49-87) + _enhance_cost_matrix into the stage matrix (stages 1 / 2 are done).
        // Rows of tracks with time_since_update > 1 are INFTY_COST times positive factors: above any
        // threshold, so the clamped entry is the same constant whatever the factors are.
        int nnear = warp_compact(nT, tmp, 0, [&](int p) {
            const int pi = pinfo[p];
            return PI_STATE(pi) != ST_TENTATIVE && PI_TSU(pi) <= 1;
        }, [&](int p) { return p; });
        for (int e = lane; e < nnear; e += 32) {          // a lane per near track: its operands stay in registers
            const int p = tmp[e];
            const int t = v.order()[p];
            const double* a = v.tab(t);
            const double* m = v.mean(t);
            const double m4[4] = {m[0], m[1], m[2], m[3]};
            for (int j = 0; j < nc3; j++) {
                const int k = cols[j];
                const double* d = v.dtab(k);
                cost[(size_t)p * Dc + k] = ss_enhance_entry(ss_iou_entry(m4, d + DT_TLWH), a, d[DT_QUALITY], d[DT_CONF], d[DT_CLS]);
            }
        }
        for (int p = lane; p < nT; p += 32) lvl[p] = PI_TSU(pinfo[p]) > 1 ? 1 : 0;
        __syncwarp();
        // most candidate rows are far (time_since_update > 1: one constant for every detection): solve with the
        // far columns implicit when the shape allows it, else on the full matrix -- same result either way
        int* nearcol = tmp;                                    // candidate indices of the near rows, ascending
        int* nearpos = byage;                                  // their list positions (the age groups are done with)
        const int nN = warp_compact(nr3, nearcol, 0, [&](int r) { return lvl[rows3[r]] == 0; }, [&](int r) { return r; });
        bool solved = false;
        if (nc3 < nr3 && nN <= SSF_NMAX && nc3 <= SSF_RMAX && nr3 <= SSF_CMAX && nr3 - nN >= nc3) {
            for (int k = lane; k < nN; k += 32) nearpos[k] = rows3[nearcol[k]];
            for (int r = lane; r < nr3; r += 32) mcol[r] = -1;
            __syncwarp();
            const double maxd = pr.max_iou_dist;
            auto cf = [&](int nidx, int i) {
                const double x = cost[(size_t)nearpos[nidx] * Dc + cols[i]];
                return x > maxd ? maxd + 1e-5 : x;
            };
            const FarWork fw = far_carve(s_my);
            if (lsa_far_warp(cf, nearcol, nN, nr3, nc3, maxd + 1e-5, fw) == 0) {
                for (int k = lane; k < nN; k += 32) {
                    const int i = fw.row4col[k];
                    mcol[nearcol[k]] = (i >= 0 && !(cf(k, i) > maxd)) ? cols[i] : -1;
                }
                solved = true;
            }
            __syncwarp();
        }
        if (!solved)
            bad |= ss_min_cost_matching(cost, Dc, rows3, nr3, cols, nc3, pr.max_iou_dist,
                                        (nr3 <= SS_LSA_SMEM_M && nc3 <= SS_LSA_SMEM_M) ? w_smem : w_glob, mcol, tmp, lvl);
        nlsa++;
        for (int r = lane; r < nr3; r += 32)
            if (mcol[r] < 0) v.tint(v.order()[rows3[r]])[TI_MISS] = 1;        // unmatched_tracks_3 -> mark_missed
        const int nm3 = warp_compact(nr3, tmp, 0, [&](int r) { return mcol[r] >= 0; }, [&](int r) { return r; });
        if (lane == 0) {
            for (int q = 0; q < nm3; q++) {                                      // matches in row order
                const int r = tmp[q];
                dmatched[mcol[r]] = 1;
                record(rows3[r], mcol[r]);
            }
            hdr[SH_STAGE3] = 1;
        }
        __syncwarp();
        nupd = __shfl_sync(BM_FULL, nupd, 0);
        nmatch = __shfl_sync(BM_FULL, nmatch, 0);
    }
    // all_unmatched_detections: ascending (sorted) detection index
    const int nunm = warp_compact(n, unm, 0, [&](int k) { return !dmatched[k]; }, [&](int k) { return k; });
    if (lane == 0) {
        hdr[SH_NUNM] = nunm; hdr[SH_NUPD] = nupd;
        hdr[SH_LSA] += nlsa; hdr[SH_MATCH] += nmatch;
        if (bad) ss_set_err(hdr, SS_ERR_LSA);
    }
}

// ---------------------------------------------------------------------------------------------
// Feature step of Track.update (track.py:238-268 with _calculate_adaptive_ema_alpha :319-335) by one
// warp: the new smoothed feature goes to a fresh pool slot that is pushed on the track's ring.
__device__ void ss_feature_step(const SsView& v, const SsParams& pr, int t, int k, const float* embs) {
    const int lane = lane_id(), E = pr.E;
    int* ti = v.tint(t);
    int* ring = v.ring(t);
    double* a = v.tab(t);
    const double* d = v.dtab(k);
    const float* f = embs + ((size_t)v.s * v.st.Dcap + v.dord()[k]) * E;
    const double fn = d[DT_NORM] + 1e-8;
    const int nring = ti[TI_NRING];
    int slot = -1;
    if (lane == 0) slot = ss_pool_pop(v);
    slot = __shfl_sync(BM_FULL, slot, 0);
    if (slot < 0) return;
    double* out = v.pool(slot, E);
    float* out32 = v.pool32(slot, E);
    auto nw = [&](int i) { return (double)f[i] / fn; };
    if (nring == 0) {                                           // first feature (:261-263)
        for (int i = lane; i < E; i += 32) { const double x = nw(i); out[i] = x; out32[i] = (float)x; }
    } else {
        const int ls = ring[nring - 1];
        const double* last = v.pool(ls, E);
        auto lf = [&](int i) { return last[i]; };
        const double dot = warp_ddot_skx(nw, lf, E);
        const double nn = sqrt(warp_ddot_skx(nw, nw, E));
        const double nl = v.pq()[ls];
        const double sim = dot / (nn * nl + 1e-8);             // :245-247
        const double conf = d[DT_CONF];
        const double cf = conf > 0.7 ? 1.0 : (conf > 0.3 ? 0.5 : 0.2);
        const double af = sim > 0.7 ? 1.0 : (sim > 0.4 ? 0.7 : 0.4);
        const double alpha = fmin(fmax(pr.ema_alpha * cf * af, 0.1), 0.95);
        const double om = 1.0 - alpha;
        auto sm = [&](int i) { return alpha * last[i] + om * nw(i); };
        const double sn = sqrt(warp_ddot_skx(sm, sm, E)) + 1e-8;   // :257
        for (int i = lane; i < E; i += 32) { const double x = sm(i) / sn; out[i] = x; out32[i] = (float)x; }
        if (lane == 0) a[SS_APPEAR] = 0.9 * a[SS_APPEAR] + 0.1 * sim;   // :248
    }
    __syncwarp();
    auto of = [&](int i) { return out[i]; };
    const double qn = sqrt(warp_ddot_skx(of, of, E));           // partial_fit's quality of this vector (:560)
    if (lane == 0) {
        v.pq()[slot] = qn;
        v.pbank()[slot] = 0;
        v.pring()[slot] = 1;
        if (nring == SS_RING) {                                 // features[-max_features:] (:266-268)
            const int old = ring[0];
            for (int i = 1; i < SS_RING; i++) ring[i - 1] = ring[i];
            ring[SS_RING - 1] = slot;
            v.pring()[old] = 0;
            if (v.pbank()[old] == 0) ss_pool_release(v, old);
        } else {
            ring[nring] = slot;
            ti[TI_NRING] = nring + 1;
        }
    }
    __syncwarp();
}

__device__ void ss_apply_update(const SsView& v, const SsParams& pr, int t, int k, const float* embs) {
    ss_feature_step(v, pr, t, k, embs);
    if (lane_id() == 0) {
        const double* d = v.dtab(k);
        double* m = v.mean(t);
        double* C = v.covb(t);
        double x[8], P[XA_P];
        for (int c = 0; c < 8; c++) x[c] = m[c];
        for (int c = 0; c < XA_P; c++) P[c] = C[c];
        const double d7[7] = {d[DT_XYAH], d[DT_XYAH + 1], d[DT_XYAH + 2], d[DT_XYAH + 3], d[DT_CONF], d[DT_CLS],
                              d[DT_DETIND]};
        ss_track_update(x, P, v.tab(t), d7, pr.n_init);
        for (int c = 0; c < 8; c++) m[c] = x[c];
        for (int c = 0; c < XA_P; c++) C[c] = P[c];
    }
    __syncwarp();
}

// K5: Track.update for every matched track, in match order (a track matched twice is updated twice).  Held to 64
// registers: a chain of dependent loads per warp (update list -> track -> ring -> pool vector), so eight CTAs per
// SM beat the 250 bytes of spills (0.75 -> 0.50 ms at 512 streams; 80 registers: 0.57)
__global__ void __launch_bounds__(128, 8)
ss_update_kernel(SsState st, SsParams pr, const float* __restrict__ embs) {
    const int s = st.s0 + blockIdx.y, u = blockIdx.x * 4 + (threadIdx.x >> 5);
    SsView v(st, s);
    if (u >= v.hdr()[SH_NUPD]) return;
    const int t = v.order()[st.ulist[(size_t)s * st.Dcap + u]];
    const int c = min(v.tint(t)[TI_NUPD], SS_MAXUPD);
    for (int i = 0; i < c; i++) ss_apply_update(v, pr, t, v.upd(t)[i], embs);
}

// ---------------------------------------------------------------------------------------------
// Index of `slot` in the target's distinct-slot list, or -1; all lanes converged, same result on all.
__device__ __forceinline__ int ss_dist_find(const int* ds, int nd, int slot) {
    const int lane = lane_id();
    for (int base = 0; base < nd; base += 32) {
        const int i = base + lane;
        const unsigned b = __ballot_sync(BM_FULL, i < nd && ds[i] == slot);
        if (b) return base + __ffs(b) - 1;
    }
    return -1;
}

// Sample list of one target (NearestNeighborDistanceMetric.samples[target]) as slot references.
// Removes the worst entry = lowest quality, newest (largest seq); all lanes converged.
__device__ void ss_bank_evict(const SsView& v, int t) {
    const int lane = lane_id();
    int* ti = v.tint(t);
    int* bs = v.bslot(t);
    int* bq = v.bseq(t);
    const int n = ti[TI_NBANK];
    const double* pq = v.pq();
    double wq = CUDART_INF;
    int wseq = -1, wi = -1;
    for (int i = lane; i < n; i += 32) {
        const double q = pq[bs[i]];
        if (q < wq || (q == wq && bq[i] > wseq)) { wq = q; wseq = bq[i]; wi = i; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double oq = shfl_xor_d(wq, o);
        const int os = __shfl_xor_sync(BM_FULL, wseq, o), oi = __shfl_xor_sync(BM_FULL, wi, o);
        if (oq < wq || (oq == wq && os > wseq)) { wq = oq; wseq = os; wi = oi; }
    }
    {
        const int slot = bs[wi];
        int* ds = v.dslot(t);
        int* dc = v.dcnt(t);
        const int nd = ti[TI_NDIST];
        const int di = ss_dist_find(ds, nd, slot);
        __syncwarp();
        if (lane == 0) {
            bs[wi] = bs[n - 1]; bq[wi] = bq[n - 1];
            ti[TI_NBANK] = n - 1;
            if (di >= 0 && --dc[di] == 0) { ds[di] = ds[nd - 1]; dc[di] = dc[nd - 1]; ti[TI_NDIST] = nd - 1; }
            if (--v.pbank()[slot] == 0 && v.pring()[slot] == 0) ss_pool_release(v, slot);
        }
    }
    __syncwarp();
    // new minimum quality
    double mq = CUDART_INF;
    for (int i = lane; i < n - 1; i += 32) mq = fmin(mq, pq[bs[i]]);
    for (int o = 16; o > 0; o >>= 1) mq = fmin(mq, shfl_xor_d(mq, o));
    if (lane == 0) v.bminq(t) = mq;
    __syncwarp();
}

// samples[target].append(feature) + the budget rule (linear_assignment.py:553-573)
__device__ void ss_bank_append(const SsView& v, const SsParams& pr, int t, int slot) {
    const int lane = lane_id();
    int* ti = v.tint(t);
    const double q = v.pq()[slot];
    const int n = ti[TI_NBANK];
    if (n >= pr.budget && !(q > v.bminq(t))) return;            // it would be the entry that is dropped again
    if (n >= pr.budget) ss_bank_evict(v, t);
    int* ds = v.dslot(t);
    int* dc = v.dcnt(t);
    const int nd = ti[TI_NDIST];
    const int di = ss_dist_find(ds, nd, slot);
    __syncwarp();
    if (lane == 0) {
        const int m = ti[TI_NBANK];
        v.bslot(t)[m] = slot;
        v.bseq(t)[m] = ti[TI_SEQ]++;
        ti[TI_NBANK] = m + 1;
        v.bminq(t) = m == 0 ? q : fmin(v.bminq(t), q);
        v.pbank()[slot] += 1;
        if (di >= 0) dc[di]++;
        else { ds[nd] = slot; dc[nd] = 1; ti[TI_NDIST] = nd + 1; }
    }
    __syncwarp();
}

// partial_fit for ONE target on a register-resident copy of its sample list (position i of the list lives in
// lane i % 32, register i / 32; the distinct-slot list one entry per lane): the same appends and evictions, in
// the same order and with the same tie rule as ss_bank_append / ss_bank_evict, but an eviction is a handful of
// shuffles instead of two scans of global memory and a round trip through lane 0.  `confirmed`: append the
// ring's features in ring order (linear_assignment.py:553-573); else trim an inactive target to `keep`
// (:575-593).  Returns false (nothing touched) when the lists do not fit the registers; the caller then takes
// the global-memory path.  All lanes converged.
template <int J>
__device__ __noinline__ bool ss_bank_pass2_regs(const SsView& v, const SsParams& pr, int t, bool confirmed, int keep) {
    const int lane = lane_id();
    int* ti = v.tint(t);
    int n = ti[TI_NBANK], nd = ti[TI_NDIST];
    const int nr = confirmed ? ti[TI_NRING] : 0;
    if (n > 32 * J || pr.budget > 32 * J || nd + nr > 32) return false;
    const double* pq = v.pq();
    int* bs = v.bslot(t);
    int* bq = v.bseq(t);
    int slot[J], seq[J];
    double q[J];
    unsigned alive = 0;
#pragma unroll
    for (int j = 0; j < J; j++) {
        const int i = j * 32 + lane;
        slot[j] = -1; seq[j] = 0; q[j] = CUDART_INF;
        if (i < n) { slot[j] = bs[i]; seq[j] = bq[i]; q[j] = pq[slot[j]]; alive |= 1u << j; }
    }
    int ds = -1, dc = 0, dc0 = 0;                              // distinct slot, its count now and at load
    if (lane < nd) { ds = v.dslot(t)[lane]; dc = dc0 = v.dcnt(t)[lane]; }
    double minq = n ? v.bminq(t) : CUDART_INF;
    int nseq = ti[TI_SEQ];

    // removes the worst entry = lowest quality, newest (largest seq); returns its position
    auto evict = [&]() -> int {
        double wq = CUDART_INF;
        int ws = -1, wp = -1, wslot = -1;
#pragma unroll
        for (int j = 0; j < J; j++)
            if ((alive >> j & 1) && (q[j] < wq || (q[j] == wq && seq[j] > ws))) { wq = q[j]; ws = seq[j]; wp = j * 32 + lane; wslot = slot[j]; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double oq = shfl_xor_d(wq, o);
            const int os = __shfl_xor_sync(BM_FULL, ws, o), op = __shfl_xor_sync(BM_FULL, wp, o),
                      osl = __shfl_xor_sync(BM_FULL, wslot, o);
            if (oq < wq || (oq == wq && os > ws)) { wq = oq; ws = os; wp = op; wslot = osl; }
        }
#pragma unroll
        for (int j = 0; j < J; j++) if (wp == j * 32 + lane) alive &= ~(1u << j);
        n--;
        const unsigned b = __ballot_sync(BM_FULL, lane < nd && ds == wslot);
        if (b) {
            const int di = __ffs(b) - 1;
            const int cnt = __shfl_sync(BM_FULL, dc, di) - 1;
            if (lane == di) dc = cnt;
            if (cnt == 0) {                                    // the slot leaves the list: the last entry takes its place
                const int c0 = __shfl_sync(BM_FULL, dc0, di);
                const int lds = __shfl_sync(BM_FULL, ds, nd - 1), ldc = __shfl_sync(BM_FULL, dc, nd - 1),
                          ldc0 = __shfl_sync(BM_FULL, dc0, nd - 1);
                if (lane == di) { ds = lds; dc = ldc; dc0 = ldc0; }
                if (lane == nd - 1) { ds = -1; dc = 0; dc0 = 0; }
                nd--;
                if (lane == 0) {
                    const int nv = v.pbank()[wslot] - c0;
                    v.pbank()[wslot] = nv;
                    if (nv == 0 && v.pring()[wslot] == 0) ss_pool_release(v, wslot);
                }
            }
        }
        double m = CUDART_INF;
#pragma unroll
        for (int j = 0; j < J; j++) if (alive >> j & 1) m = fmin(m, q[j]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmin(m, shfl_xor_d(m, o));
        minq = m;
        return wp;
    };

    if (confirmed) {
        // the ring's slots and their qualities, one per lane, fetched ONCE (a load pair per loop iteration was a
        // serial chain of 2 nr memory round trips: most of this pass's time)
        const int* rg = v.ring(t);
        const int rsl = lane < nr ? rg[lane] : 0;
        const double rqi = lane < nr ? pq[rsl] : 0.0;
        for (int i = 0; i < nr; i++) {
            const int sl = __shfl_sync(BM_FULL, rsl, i);
            const double qi = shfl_d(rqi, i);
            if (n >= pr.budget && !(qi > minq)) continue;      // it would be the entry that is dropped again
            const int h = n >= pr.budget ? evict() : n;        // the list has no holes here: an eviction's is refilled at once
#pragma unroll
            for (int j = 0; j < J; j++)
                if (h == j * 32 + lane) { slot[j] = sl; seq[j] = nseq; q[j] = qi; alive |= 1u << j; }
            nseq++;
            minq = n == 0 ? qi : fmin(minq, qi);
            n++;
            const unsigned b = __ballot_sync(BM_FULL, lane < nd && ds == sl);
            if (b) { if (lane == __ffs(b) - 1) dc++; }
            else { if (lane == nd) { ds = sl; dc = 1; dc0 = 0; } nd++; }
        }
    } else if (n > keep) {
        // Trim of an inactive target (:575-593): n - keep evictions of the worst entry leave the `keep` BEST entries
        // under the same order (quality, then older first) -- select those directly (a dying zombie's list is 50
        // copies of one vector: 45 evictions of ~500 cycles each were most of this pass)
        unsigned sel = 0;
        for (int it = 0; it < keep; it++) {
            double bq2 = -CUDART_INF;
            int bs2 = 0x7fffffff, bp = -1;
#pragma unroll
            for (int j = 0; j < J; j++)
                if ((alive >> j & 1) && !(sel >> j & 1) && (q[j] > bq2 || (q[j] == bq2 && seq[j] < bs2))) { bq2 = q[j]; bs2 = seq[j]; bp = j * 32 + lane; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double oq = shfl_xor_d(bq2, o);
                const int os = __shfl_xor_sync(BM_FULL, bs2, o), op = __shfl_xor_sync(BM_FULL, bp, o);
                if (op >= 0 && (bp < 0 || oq > bq2 || (oq == bq2 && os < bs2))) { bq2 = oq; bs2 = os; bp = op; }
            }
#pragma unroll
            for (int j = 0; j < J; j++) if (bp == j * 32 + lane) sel |= 1u << j;
        }
        alive = sel;
        n = keep;
        // distinct list: the count of each old slot among the survivors; a slot without one leaves (and goes back to
        // the pool when the ring does not hold it either)
        int dcn = 0;
#pragma unroll
        for (int j = 0; j < J; j++) {
            unsigned m = __ballot_sync(BM_FULL, alive >> j & 1);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const int sl = __shfl_sync(BM_FULL, slot[j], src);
                if (lane < nd && ds == sl) dcn++;
            }
        }
        const bool gone = lane < nd && dcn == 0;
        if (gone) {
            const int nv = v.pbank()[ds] - dc0;
            v.pbank()[ds] = nv;
            if (nv == 0 && v.pring()[ds] == 0) ss_pool_release(v, ds);
        }
        // the kept entries move to the front: lane i takes the i-th kept lane's (slot, new count, count at load); the
        // write-back below applies pbank[slot] += new - old
        const unsigned km = __ballot_sync(BM_FULL, lane < nd && dcn > 0);
        const int nd2 = __popc(km);
        const unsigned srcl = __fns(km, 0, lane + 1);
        const int from = srcl < 32u ? (int)srcl : 0;
        const int nds = __shfl_sync(BM_FULL, ds, from), ndc = __shfl_sync(BM_FULL, dcn, from),
                  ndc0 = __shfl_sync(BM_FULL, dc0, from);
        if (lane < nd2) { ds = nds; dc = ndc; dc0 = ndc0; } else { ds = -1; dc = 0; dc0 = 0; }
        nd = nd2;
        double m2 = CUDART_INF;
#pragma unroll
        for (int j = 0; j < J; j++) if (alive >> j & 1) m2 = fmin(m2, q[j]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m2 = fmin(m2, shfl_xor_d(m2, o));
        minq = m2;
    }
    // write back, compacting the holes a trim leaves
    int base = 0;
#pragma unroll
    for (int j = 0; j < J; j++) {
        const bool a = alive >> j & 1;
        const unsigned m = __ballot_sync(BM_FULL, a);
        if (a) {
            const int pos = base + __popc(m & ((1u << lane) - 1u));
            bs[pos] = slot[j]; bq[pos] = seq[j];
        }
        base += __popc(m);
    }
    if (lane < nd) {
        v.dslot(t)[lane] = ds; v.dcnt(t)[lane] = dc;
        if (dc != dc0) v.pbank()[ds] += dc - dc0;
    }
    if (lane == 0) { ti[TI_NBANK] = n; ti[TI_NDIST] = nd; ti[TI_SEQ] = nseq; v.bminq(t) = minq; }
    __syncwarp();
    return true;
}

// Everything a track holds goes back to the pool (the track left the lost buffer or never entered it); one
// warp, lanes over the ring / sample-list entries (several entries may reference one slot: atomic counts).
__device__ void ss_track_release_warp(const SsView& v, int t) {
    const int lane = lane_id();
    int* ti = v.tint(t);
    const int nr = ti[TI_NRING], nb = ti[TI_NBANK];
    for (int i = lane; i < nr; i += 32) {
        const int p = v.ring(t)[i];
        v.pring()[p] = 0;
    }
    __syncwarp();
    for (int i = lane; i < nb; i += 32) {
        const int p = v.bslot(t)[i];
        if (atomicSub(&v.pbank()[p], 1) == 1 && v.pring()[p] == 0) ss_pool_release(v, p);
    }
    __syncwarp();
    for (int i = lane; i < nr; i += 32) {          // ring features no sample list references (any more)
        const int p = v.ring(t)[i];
        if (v.pbank()[p] == 0) {
            // a slot referenced by the list was released above when its count reached 0; one that never was is released here
            bool in_bank = false;
            for (int k = 0; k < nb && !in_bank; k++) in_bank = v.bslot(t)[k] == p;
            if (!in_bank) ss_pool_release(v, p);
        }
    }
    __syncwarp();
    if (lane == 0) {
        ti[TI_NRING] = 0; ti[TI_NBANK] = 0; ti[TI_NDIST] = 0;
        const int k = atomicAdd(&v.hdr()[SH_NTFREE], 1);
        v.st.tfree[(size_t)v.s * v.st.Tcap + k] = t;
    }
    __syncwarp();
}

// Stable compaction by the whole CTA: calls emit(i, rank) for every i in [0, n) with pred(i), rank = number of
// earlier i with pred.  Returns the total on all threads.  ws: shared int[33].
template <typename Pred, typename Emit>
__device__ int block_compact(int n, int* ws, Pred pred, Emit emit) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarp = blockDim.x >> 5;
    int base_total = 0;
    for (int base = 0; base < n; base += blockDim.x) {
        const int i = base + tid;
        const bool ok = i < n && pred(i);
        const unsigned b = __ballot_sync(BM_FULL, ok);
        __syncthreads();
        if (lane == 0) ws[wid] = __popc(b);
        __syncthreads();
        int before = 0, chunk = 0;
        for (int w = 0; w < nwarp; w++) { const int c = ws[w]; if (w < wid) before += c; chunk += c; }
        if (ok) emit(i, base_total + before + __popc(b & ((1u << lane) - 1)));
        base_total += chunk;
    }
    return base_total;
}

// K6: the rest of Tracker.update (tracker.py:137-181) and StrongSort._format_outputs
__global__ void __launch_bounds__(512, 2)
ss_finalize_kernel(SsState st, SsParams pr, const float* __restrict__ embs, double* __restrict__ out,
                   int* __restrict__ nout, int out_rows) {
    const int s = st.s0 + blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
    SsView v(st, s);
    int* hdr = v.hdr();
    int* order = v.order();
    int* unm = st.unm + (size_t)s * st.Dcap;
    int* lost = st.lost + (size_t)s * 32;
    double* sims = st.sims + (size_t)s * 32 * st.Dcap;
    int* born = st.ulist + (size_t)s * st.Dcap;          // reused: track slots of this frame's births
    const int E = pr.E;
    __shared__ int s_rec_l, s_rec_u, s_nb;
    long long* stamp = reinterpret_cast<long long*>(st.wkd + (size_t)s * 3 * st.M);     // debug: phase clocks
#define SS_STAMP(k) do { if (tid == 0) stamp[k] = clock64(); } while (0)
    SS_STAMP(0);
    int nT = hdr[SH_NTRACKS];
    const int n = hdr[SH_NDET];
    // (a) mark_missed for the IoU stage's unmatched rows (tracker.py:141-142)
    for (int p = tid; p < nT; p += blockDim.x) {
        const int t = order[p];
        if (v.tint(t)[TI_MISS]) ss_track_mark_missed(v.tab(t), pr.max_age);
    }
    __syncthreads();
    SS_STAMP(1);
    // (b) _attempt_id_recovery (tracker.py:300-351)
    const int nlost = hdr[SH_NLOST];
    int nunm = hdr[SH_NUNM];
    if (nlost > 0 && nunm > 0) {
        // Only "is the best similarity above 0.7, and which detection has it" is used (:325-331).  Every pair is
        // screened in fp32 against the feature's fp32 shadow; a pair within 0.02 of the threshold -- none, for
        // unrelated identities -- is recomputed in fp64 as the reference does, so the decision is the reference's.
        const int ngrp = (nunm + 3) / 4;
        for (int e = wid; e < nlost * ngrp; e += nw) {            // a warp per (lost track, 4 detections): 4 dot products in flight
            const int li = e / ngrp, u0 = (e % ngrp) * 4;
            const int t = lost[li];
            const int nr = v.tint(t)[TI_NRING];
            double sim[4] = {-CUDART_INF, -CUDART_INF, -CUDART_INF, -CUDART_INF};
            if (nr > 0) {
                const int ls = v.ring(t)[nr - 1];
                const float* tf = v.pool32(ls, E);
                const float* f[4];
                double den[4];
                for (int q = 0; q < 4; q++) {
                    const int uj = min(u0 + q, nunm - 1);
                    const int k = unm[uj];
                    f[q] = embs + ((size_t)s * st.Dcap + v.dord()[k]) * E;
                    den[q] = v.dtab(k)[DT_NORM] * v.pq()[ls];
                }
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                if ((E & 3) == 0) {
                    const float4* t4 = reinterpret_cast<const float4*>(tf);
                    for (int i = lane; i < E / 4; i += 32) {
                        const float4 b = t4[i];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            const float4 a = reinterpret_cast<const float4*>(f[q])[i];
                            acc[q] += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
                        }
                    }
                } else {
                    for (int i = lane; i < E; i += 32)
#pragma unroll
                        for (int q = 0; q < 4; q++) acc[q] += f[q][i] * tf[i];
                }
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    for (int o = 16; o > 0; o >>= 1) acc[q] += __shfl_xor_sync(BM_FULL, acc[q], o);
                    acc[q] = __shfl_sync(BM_FULL, acc[q], 0);     // one value for the warp: the branch below is uniform
                    sim[q] = (double)acc[q] / den[q];
                    if (sim[q] > 0.68) {                          // near or above the threshold: exact
                        const double* td = v.pool(ls, E);
                        double ad = 0.0;
                        for (int i = lane; i < E; i += 32) ad += (double)f[q][i] * td[i];
                        for (int o = 16; o > 0; o >>= 1) ad += shfl_xor_d(ad, o);
                        sim[q] = shfl_d(ad, 0) / den[q];
                    }
                }
            }
            if (lane < 4 && u0 + lane < nunm) sims[li * st.Dcap + u0 + lane] = sim[lane];
        }
    }
    __syncthreads();
    if (tid == 0) {
        s_rec_l = -1; s_rec_u = -1;
        if (nlost > 0 && nunm > 0)
            for (int li = 0; li < nlost && s_rec_l < 0; li++) {
                if (v.tint(lost[li])[TI_NRING] == 0) continue;
                int b = 0;
                for (int uj = 1; uj < nunm; uj++) if (sims[li * st.Dcap + uj] > sims[li * st.Dcap + b]) b = uj;
                if (sims[li * st.Dcap + b] > 0.7) { s_rec_l = li; s_rec_u = b; }
            }
    }
    __syncthreads();
    if (s_rec_l >= 0) {
        const int t = lost[s_rec_l], k = unm[s_rec_u];
        if (wid == 0) {
            if (lane == 0) { v.tab(t)[ST_STATE] = ST_TENTATIVE; v.tab(t)[SS_TSU] = 0.0; }   // "state = 1" (:337)
            __syncwarp();
            ss_apply_update(v, pr, t, k, embs);
            if (lane == 0) {
                if (nT < st.Tcap) order[nT] = t; else ss_set_err(hdr, SS_ERR_TRACKS);
                for (int i = s_rec_l; i + 1 < nlost; i++) lost[i] = lost[i + 1];
                hdr[SH_NLOST] = nlost - 1;
                for (int i = s_rec_u; i + 1 < nunm; i++) unm[i] = unm[i + 1];
                hdr[SH_MATCH] += 1;
            }
        }
        __syncthreads();
        if (nT < st.Tcap) nT++;
        nunm--;
    }
    __syncthreads();
    SS_STAMP(2);
    // (c) _initiate_track for the remaining unmatched detections, in order (tracker.py:147-150, :395-410)
    if (tid == 0) {
        int nb = 0;
        for (int j = 0; j < nunm; j++) {
            if (hdr[SH_NTFREE] <= 0 || nT + nb >= st.Tcap) { ss_set_err(hdr, SS_ERR_TRACKS); break; }
            const int t = st.tfree[(size_t)s * st.Tcap + --hdr[SH_NTFREE]];
            born[j] = t;
            v.tint(t)[TI_ID] = hdr[SH_NEXT_ID]++;
            order[nT + nb] = t;
            nb++;
        }
        s_nb = nb;
        hdr[SH_BIRTHS] += nb;
    }
    __syncthreads();
    const int nb = s_nb;
    for (int j = wid; j < nb; j += nw) {
        const int t = born[j], k = unm[j];
        const double* d = v.dtab(k);
        int* ti = v.tint(t);
        if (lane == 0) {
            double x[8], P[XA_P];
            cv8_initiate<CV8_XYAH>(d + DT_XYAH, x, P);
            for (int c = 0; c < 8; c++) v.mean(t)[c] = x[c];
            for (int c = 0; c < XA_P; c++) v.covb(t)[c] = P[c];
            double* a = v.tab(t);
            for (int c = 0; c < ST_FIELDS; c++) a[c] = 0.0;
            a[SS_AGE] = 1.0; a[SS_HITS] = 1.0; a[SS_MOTION] = 1.0; a[SS_APPEAR] = 1.0;
            a[SS_QUALITY] = d[DT_QUALITY];
            a[SS_HI_STREAK] = d[DT_CONF] > 0.7 ? 1.0 : 0.0;
            a[SS_CONF] = d[DT_CONF]; a[SS_CLS] = d[DT_CLS];
            a[ST_STATE] = pr.born_confirmed ? (double)ST_CONFIRMED_STATE : (double)ST_TENTATIVE;
            a[ST_CONFIRMED] = 1.0; a[ST_NPOS] = 1.0; a[ST_PX] = d[DT_XYAH]; a[ST_PY] = d[DT_XYAH + 1];
            a[ST_NCONF] = 1.0; a[ST_CONF0] = d[DT_CONF]; a[ST_DETIND] = d[DT_DETIND];
            ti[TI_NRING] = 0; ti[TI_NBANK] = 0; ti[TI_NDIST] = 0; ti[TI_SEQ] = 0; ti[TI_NUPD] = 0; ti[TI_MISS] = 0;
            ti[TI_LOSTFRAME] = 0;
        }
        __syncwarp();
        ss_feature_step(v, pr, t, k, embs);                 // first feature: the normalised detection feature
    }
    __syncthreads();
    nT += nb;
    SS_STAMP(3);
    // (d) deleted tracks -> lost buffer (tracker.py:152-163), compaction of the list
    __shared__ int s_ndel, s_w;
    int* dellist = st.wki + (size_t)s * st.wki_stride;   // the match kernel's scratch (>= Tcap ints) is free now
    __shared__ int s_ws[33];
    {
        const int k = block_compact(nT, s_ws, [&](int p) { return v.tab(order[p])[ST_STATE] == (double)ST_DELETED; },
                                    [&](int p, int r) { dellist[r] = order[p]; });
        if (tid == 0) s_ndel = k;
    }
    __syncthreads();
    const int ndel = s_ndel;
    if (ndel > 0) {
        // the lost buffer's bookkeeping on shared-memory copies (<= 30 entries), one thread; the tracks that
        // drop out are collected and released in parallel afterwards
        __shared__ int s_lost[SS_LOST + 2], s_lframe[SS_LOST + 2], s_nrel;
        int* rellist = dellist + st.Tcap;                    // second Tcap ints of the match kernel's scratch
        const int nl0 = hdr[SH_NLOST];
        if (tid < nl0) { s_lost[tid] = lost[tid]; s_lframe[tid] = v.tint(lost[tid])[TI_LOSTFRAME]; }
        __syncthreads();
        if (tid == 0) {
            const int now = n > 0 ? hdr[SH_FRAME] : hdr[SH_NHIST];      // frame_id is None for tracker.update([])
            int nl = nl0, nrel = 0;
            for (int i0 = 0; i0 < ndel; i0++) {
                const int t = dellist[i0];
                bool kept = false;
                if (nl < SS_LOST) { s_lost[nl] = t; s_lframe[nl] = now; nl++; kept = true; v.tint(t)[TI_LOSTFRAME] = now; }
                int k2 = 0;
                for (int i = 0; i < nl; i++) {
                    if (now - s_lframe[i] < pr.max_age) { s_lost[k2] = s_lost[i]; s_lframe[k2] = s_lframe[i]; k2++; }
                    else rellist[nrel++] = s_lost[i];
                }
                nl = k2;
                if (!kept) rellist[nrel++] = t;
            }
            for (int i = 0; i < nl; i++) lost[i] = s_lost[i];
            hdr[SH_NLOST] = nl;
            s_nrel = nrel;
        }
        __syncthreads();
        for (int i = wid; i < s_nrel; i += nw) ss_track_release_warp(v, rellist[i]);
        __syncthreads();
        {                                                // stable compaction of the list (through a scratch copy)
            int* keep_l = dellist + 2 * (size_t)st.Tcap;     // third Tcap ints of the scratch
            const int k = block_compact(nT, s_ws, [&](int p) { return v.tab(order[p])[ST_STATE] != (double)ST_DELETED; },
                                        [&](int p, int r) { keep_l[r] = order[p]; });
            __syncthreads();
            for (int i = tid; i < k; i += blockDim.x) order[i] = keep_l[i];
            if (tid == 0) s_w = k;
        }
        __syncthreads();
        nT = s_w;
    }
    if (tid == 0) {
        hdr[SH_NTRACKS] = nT;
        hdr[SH_NHIST] = min(hdr[SH_NHIST] + 1, 100);
    }
    int anyc = 0;
    for (int p = tid; p < nT; p += blockDim.x) anyc |= v.tab(order[p])[ST_STATE] == (double)ST_CONFIRMED_STATE;
    anyc = __syncthreads_or(anyc);
    SS_STAMP(4);
    // (e) metric.partial_fit (tracker.py:165-178, linear_assignment.py:538-593)
    if (anyc) {
        const int keep = pr.budget / 4 < 5 ? pr.budget / 4 : 5;
        const int nl = hdr[SH_NLOST];
        // pass 1, a thread per target: a full sample list takes a feature only if its quality beats the list's
        // minimum -- when none of the ring's features does (the steady state) all <= 10 appends are no-ops.
        // The targets with real work are collected for pass 2.
        __shared__ int s_nwork;
        int* work = st.wki + (size_t)s * st.wki_stride + 2 * (size_t)st.Tcap;     // third Tcap ints of the scratch
        if (tid == 0) s_nwork = 0;
        __syncthreads();
        for (int p = tid; p < nT + nl; p += blockDim.x) {
            const int t = p < nT ? order[p] : lost[p - nT];
            const int* ti = v.tint(t);
            bool todo;
            if (p < nT && v.tab(t)[ST_STATE] == (double)ST_CONFIRMED_STATE) {
                const int nb0 = ti[TI_NBANK], nr = ti[TI_NRING];
                const int* rg = v.ring(t);
                if (nb0 + nr <= pr.budget) {
                    // every append fits under the budget: plain appends by this thread (no other thread touches
                    // this target; a pool slot belongs to one target)
                    int* tw = v.tint(t);
                    int* bs = v.bslot(t);
                    int* bq = v.bseq(t);
                    int* ds = v.dslot(t);
                    int* dc = v.dcnt(t);
                    int nb = nb0, ndl = tw[TI_NDIST], seq = tw[TI_SEQ];
                    double mq = nb ? v.bminq(t) : CUDART_INF;
                    for (int i = 0; i < nr; i++) {
                        const int slot = rg[i];
                        bs[nb] = slot; bq[nb] = seq++; nb++;
                        mq = fmin(mq, v.pq()[slot]);
                        v.pbank()[slot] += 1;
                        int j = 0;
                        for (; j < ndl; j++) if (ds[j] == slot) { dc[j]++; break; }
                        if (j == ndl) { ds[ndl] = slot; dc[ndl] = 1; ndl++; }
                    }
                    tw[TI_NBANK] = nb; tw[TI_NDIST] = ndl; tw[TI_SEQ] = seq;
                    v.bminq(t) = mq;
                    todo = false;
                } else {
                    todo = nb0 < pr.budget;
                    if (!todo) {
                        const double mq = v.bminq(t);
                        for (int i = 0; i < nr && !todo; i++) todo = v.pq()[rg[i]] > mq;
                    }
                }
            } else {
                todo = ti[TI_NBANK] > keep;                              // inactive target (:575-593)
            }
            if (todo) work[atomicAdd(&s_nwork, 1)] = p;
        }
        __syncthreads();
        if (tid == 0) { stamp[8] = clock64(); stamp[9] = s_nwork; hdr[SH_NWORK] = s_nwork; }   // pass 2: ss_bank_kernel
    } else if (tid == 0) hdr[SH_NWORK] = 0;
    __syncthreads();
    SS_STAMP(5);
    // (f) _format_outputs (strongsort.py:329-359): confirmed tracks updated this frame, in list order
    {
        double* o = out + (size_t)s * out_rows * 8;
        const int k = block_compact(nT, s_ws, [&](int p) {
            const double* a = v.tab(order[p]);
            return a[ST_STATE] == (double)ST_CONFIRMED_STATE && a[SS_TSU] < 1.0;
        }, [&](int p, int r) {
            if (r >= out_rows) return;
            const int t = order[p];
            const double* m = v.mean(t);
            const double* a = v.tab(t);
            const double w = m[2] * m[3];
            const double x1 = m[0] - w / 2.0, y1 = m[1] - m[3] / 2.0;       // to_tlbr (track.py:144-171)
            double* row = o + (size_t)r * 8;
            row[0] = x1; row[1] = y1; row[2] = x1 + w; row[3] = y1 + m[3];
            row[4] = v.tint(t)[TI_ID]; row[5] = a[SS_CONF]; row[6] = a[SS_CLS]; row[7] = a[ST_DETIND];
        });
        if (tid == 0) nout[s] = k < out_rows ? k : out_rows;
    }
    __syncthreads();
    SS_STAMP(6);
#undef SS_STAMP
}

// K6b: partial_fit's second pass, a warp per target with real appends / trims (the work list of K6), spread over
// several CTAs per stream: in K6 it was 43 % of the CTA's time at eight warps per stream
__global__ void __launch_bounds__(256, 5)
ss_bank_kernel(SsState st, SsParams pr) {
    const int s = st.s0 + blockIdx.y, lane = threadIdx.x & 31;
    SsView v(st, s);
    const int* hdr = v.hdr();
    const int nwork = hdr[SH_NWORK];
    if (nwork <= 0) return;
    const int nT = hdr[SH_NTRACKS];
    const int* order = v.order();
    const int* lost = st.lost + (size_t)s * 32;
    const int* work = st.wki + (size_t)s * st.wki_stride + 2 * (size_t)st.Tcap;
    const int keep = pr.budget / 4 < 5 ? pr.budget / 4 : 5;
    const int nw = (blockDim.x >> 5) * gridDim.x;
    for (int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); q < nwork; q += nw) {
        const int p = work[q];
        const int t = p < nT ? order[p] : lost[p - nT];
        int* ti = v.tint(t);
        const bool conf = p < nT && v.tab(t)[ST_STATE] == (double)ST_CONFIRMED_STATE;
        // the list lives in J registers per lane: the smallest J that holds the budget (the pass is instruction-bound,
        // ~200 warp instructions per append at J = 5)
        if (pr.budget <= 64 ? ss_bank_pass2_regs<2>(v, pr, t, conf, keep) : ss_bank_pass2_regs<5>(v, pr, t, conf, keep)) continue;
        if (conf) {
            const int nr = ti[TI_NRING];
            // the list's minimum quality only rises while entries are replaced: a ring feature that does not
            // beat it now cannot enter later in this loop either (while there is room, every feature enters)
            const bool cand = lane < nr && (ti[TI_NBANK] < pr.budget || v.pq()[v.ring(t)[lane]] > v.bminq(t));
            unsigned m = __ballot_sync(BM_FULL, cand);
            if (ti[TI_NBANK] < pr.budget) m = nr >= 32 ? 0xffffffffu : ((1u << nr) - 1u);
            while (m) {
                const int i = __ffs(m) - 1;
                m &= m - 1;
                ss_bank_append(v, pr, t, v.ring(t)[i]);
            }
        } else {
            while (ti[TI_NBANK] > keep) ss_bank_evict(v, t);
        }
    }
}

// K6c: merge the deferred frees, rebuild the list of features some sample list references (the rows of the next
// frame's contraction)
__global__ void __launch_bounds__(256)
ss_live_kernel(SsState st) {
    const int s = st.s0 + blockIdx.x, tid = threadIdx.x;
    SsView v(st, s);
    int* hdr = v.hdr();
    {
        const int nf = hdr[SH_NPFREED], np = hdr[SH_NPFREE];
        for (int i = tid; i < nf; i += blockDim.x) v.pfree()[np + i] = v.pfreed()[i];
        __syncthreads();
        if (tid == 0) { hdr[SH_NPFREE] = np + nf; hdr[SH_NPFREED] = 0; hdr[SH_NLIVE] = 0; }
    }
    __syncthreads();
    int* live = st.live + (size_t)s * st.Pcap;
    float* lnorm = st.lnorm + (size_t)s * st.Pcap;
    for (int p = tid; p < st.Pcap; p += blockDim.x)
        if (v.pbank()[p] > 0) {
            const int k = atomicAdd(&hdr[SH_NLIVE], 1);
            live[k] = p;
            lnorm[k] = (float)(v.pq()[p] + 1e-8);
        }
}

__global__ void ss_reset_kernel(SsState st) {
    const int s = blockIdx.x;
    SsView v(st, s);
    int* hdr = v.hdr();
    for (int i = threadIdx.x; i < st.Tcap; i += blockDim.x) st.tfree[(size_t)s * st.Tcap + i] = st.Tcap - 1 - i;
    for (int i = threadIdx.x; i < st.Pcap; i += blockDim.x) {
        v.pfree()[i] = st.Pcap - 1 - i; v.pbank()[i] = 0; v.pring()[i] = 0;
    }
    for (int i = threadIdx.x; i < st.Tcap * TI_COUNT; i += blockDim.x) st.tint[(size_t)s * st.Tcap * TI_COUNT + i] = 0;
    if (threadIdx.x == 0) {
        for (int i = 0; i < SH_COUNT; i++) hdr[i] = 0;
        hdr[SH_NTFREE] = st.Tcap; hdr[SH_NPFREE] = st.Pcap; hdr[SH_NEXT_ID] = 1;
        st.n_updates[s] = 0;
    }
}

}  // namespace bm

// ---------------------------------------------------------------------------------------------
namespace {

struct SsEngine {
    bm::SsState st;
    bm::SsParams pr;
    std::vector<void*> allocs;
    int device = 0, n_sm = 148;
    cudaEvent_t ev[7];
    bool timing = false, timed = false;
    // stream groups (bm_strongsort_set_groups): group g = streams [g S / G, (g + 1) S / G) on its own CUDA stream
    std::vector<cudaStream_t> gstreams;
    std::vector<cudaEvent_t> gdone;
    cudaEvent_t fork = nullptr;
};

int ss_fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    char msg[256];
    if (e != cudaSuccess) snprintf(msg, sizeof(msg), "%s: %s", what, cudaGetErrorString(e));
    else snprintf(msg, sizeof(msg), "%s", what);
    bm::set_last_error(msg);
    return code;
}

template <typename T>
cudaError_t ss_alloc(SsEngine* e, T** p, size_t n) {
    cudaError_t r = cudaMalloc(reinterpret_cast<void**>(p), n * sizeof(T));
    if (r == cudaSuccess) e->allocs.push_back(*p);
    return r;
}

#define SS_CUDA(call)                                            \
    do {                                                         \
        cudaError_t e_ = (call);                                 \
        if (e_ != cudaSuccess) return ss_fail(-2, #call, e_);    \
    } while (0)

size_t ss_prep_smem(int Dcap) { return (size_t)Dcap * (8 + 8 + 4) + 16; }

}  // namespace

extern "C" {

int bm_strongsort_create(const bm_strongsort_config* cfg, int n_streams, int track_cap, int det_cap, int pool_cap,
                         void** handle) {
    if (!cfg || !handle || n_streams <= 0 || track_cap <= 0 || det_cap <= 0 || pool_cap <= 0 || cfg->emb_dim <= 0)
        return ss_fail(-1, "bm_strongsort_create: bad argument");
    if (cfg->nn_budget < 1 || cfg->nn_budget > 1024) return ss_fail(-1, "bm_strongsort_create: nn_budget must be in 1..1024");
    if (cfg->max_age < 1 || cfg->max_age > 4096) return ss_fail(-1, "bm_strongsort_create: max_age out of range");
    SsEngine* e = new (std::nothrow) SsEngine();
    if (!e) return ss_fail(-3, "out of host memory");
    bm::SsState& st = e->st;
    bm::SsParams& pr = e->pr;
    pr.min_conf = cfg->min_conf; pr.max_cos_dist = cfg->max_cos_dist; pr.max_iou_dist = cfg->max_iou_dist;
    pr.mc_lambda = cfg->mc_lambda; pr.ema_alpha = cfg->ema_alpha; pr.conf_hi = cfg->conf_thresh_high;
    pr.conf_lo = cfg->conf_thresh_low; pr.id_weight = cfg->id_preservation_weight;
    pr.max_age = cfg->max_age; pr.n_init = cfg->n_init; pr.budget = cfg->nn_budget;
    pr.born_confirmed = cfg->born_confirmed; pr.E = cfg->emb_dim;
    st.S = n_streams; st.Tcap = track_cap; st.Dcap = det_cap; st.Pcap = pool_cap; st.B = cfg->nn_budget;
    st.M = 2 * track_cap + det_cap;
    const size_t S = n_streams, T = track_cap, D = det_cap, P = pool_cap, B = st.B, M = st.M, E = pr.E;
    st.wki_stride = T * 3 + 2 * T + 2 * T + D + (2 * T + D) + D + (size_t)pr.max_age + 2 + 6 * M + 8;
    cudaError_t r = cudaGetDevice(&e->device);
    if (r == cudaSuccess) r = cudaDeviceGetAttribute(&e->n_sm, cudaDevAttrMultiProcessorCount, e->device);
    auto A = [&](auto** p, size_t n) { if (r == cudaSuccess) r = ss_alloc(e, p, n); };
    A(&st.mean, S * T * 8); A(&st.covb, S * T * bm::XA_P); A(&st.tab, S * T * bm::ST_FIELDS);
    A(&st.tint, S * T * bm::TI_COUNT); A(&st.ring, S * T * bm::SS_RING);
    A(&st.bslot, S * T * B); A(&st.bseq, S * T * B); A(&st.dslot, S * T * B); A(&st.dcnt, S * T * B);
    A(&st.bminq, S * T); A(&st.upd, S * T * bm::SS_MAXUPD); A(&st.order, S * T); A(&st.lost, S * 32);
    A(&st.tfree, S * T); A(&st.hdr, S * bm::SH_COUNT); A(&st.n_updates, S);
    A(&st.pool, S * P * E); A(&st.pool32, S * P * E); A(&st.pq, S * P);
    A(&st.pbank, S * P); A(&st.pring, S * P); A(&st.pfree, S * P); A(&st.pfreed, S * P); A(&st.live, S * P);
    A(&st.lnorm, S * P); A(&st.dist, S * P * D); A(&st.cost, S * T * D);
    A(&st.dtab, S * D * bm::DT_COUNT); A(&st.dord, S * D); A(&st.dnormf, S * D);
    A(&st.ulist, S * D); A(&st.unm, S * D); A(&st.wki, S * st.wki_stride);
    A(&st.wkd, S * 3 * M); A(&st.sims, S * 32 * D);
    for (int i = 0; i < 7 && r == cudaSuccess; i++) r = cudaEventCreate(&e->ev[i]);
    if (r == cudaSuccess) r = bm::smem_optin(bm::cosine_tc_kernel<bm::SsBankProblem>);
    if (r == cudaSuccess) r = bm::smem_optin(bm::ss_prep_kernel);
    if (r == cudaSuccess) r = bm::smem_optin(bm::ss_cost_kernel);
    if (r == cudaSuccess) {
        bm::ss_reset_kernel<<<n_streams, 256>>>(st);
        r = cudaGetLastError();
        if (r == cudaSuccess) r = cudaDeviceSynchronize();
    }
    if (r != cudaSuccess) {
        for (void* p : e->allocs) cudaFree(p);
        delete e;
        return ss_fail(-2, "bm_strongsort_create", r);
    }
    *handle = e;
    return 0;
}

int bm_linear_sum_assignment_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                                   const int* d_nc, int* d_x, void* d_work, void* stream) {
    if (n_problems <= 0) return 0;
    if (!d_cost || !d_nr || !d_nc || !d_x || !d_work || nr_max <= 0 || nc_max <= 0)
        return ss_fail(-1, "bm_linear_sum_assignment_batch: bad argument");
    const int M = nr_max > nc_max ? nr_max : nc_max;
    double* wd = static_cast<double*>(d_work);
    int* wi = reinterpret_cast<int*>(wd + (size_t)n_problems * 3 * M);
    bm::lsa_batch_kernel<<<(n_problems + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        d_cost, n_problems, nr_max, nc_max, d_nr, d_nc, d_x, wd, wi, M);
    SS_CUDA(cudaGetLastError());
    return 0;
}

int bm_linear_sum_assignment_far_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                                       const int* d_nc, const int* d_far_rows, double cstar, int* d_x, int* d_used_far,
                                       void* d_work, void* stream) {
    if (n_problems <= 0) return 0;
    if (!d_cost || !d_nr || !d_nc || !d_far_rows || !d_x || !d_work || nr_max <= 0 || nc_max <= 0)
        return ss_fail(-1, "bm_linear_sum_assignment_far_batch: bad argument");
    const int M = nr_max > nc_max ? nr_max : nc_max;
    double* wd = static_cast<double*>(d_work);
    int* wi = reinterpret_cast<int*>(wd + (size_t)n_problems * 3 * M);
    bm::lsa_far_batch_kernel<<<(n_problems + 1) / 2, 64, 0, static_cast<cudaStream_t>(stream)>>>(
        d_cost, n_problems, nr_max, nc_max, d_nr, d_nc, d_far_rows, cstar, d_x, d_used_far, wd, wi, M);
    SS_CUDA(cudaGetLastError());
    return 0;
}

int bm_strongsort_destroy(void* handle) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e) return 0;
    for (void* p : e->allocs) cudaFree(p);
    for (int i = 0; i < 7; i++) cudaEventDestroy(e->ev[i]);
    for (cudaStream_t g : e->gstreams) cudaStreamDestroy(g);
    for (cudaEvent_t g : e->gdone) cudaEventDestroy(g);
    if (e->fork) cudaEventDestroy(e->fork);
    delete e;
    return 0;
}

int bm_strongsort_reset(void* handle, void* stream) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e) return ss_fail(-1, "bm_strongsort_reset: null handle");
    bm::ss_reset_kernel<<<e->st.S, 256, 0, static_cast<cudaStream_t>(stream)>>>(e->st);
    SS_CUDA(cudaGetLastError());
    return 0;
}

// The frame's six launches for the streams [st.s0, st.s0 + st.S) on one CUDA stream
static cudaError_t ss_launch_chain(SsEngine* e, const bm::SsState& st, const float* d_dets, const int* d_ndets,
                                   const float* d_embs, const double* d_warps, double* d_out, int* d_nout,
                                   cudaStream_t s, bool tm) {
    const bm::SsParams& pr = e->pr;
    if (tm) cudaEventRecord(e->ev[0], s);
    bm::ss_prep_kernel<<<st.S, 256, ss_prep_smem(st.Dcap), s>>>(st, pr, d_dets, d_ndets, d_embs, d_warps);
    if (tm) cudaEventRecord(e->ev[1], s);
    bm::SsBankProblem pb{st, pr.E, d_embs};
    bm::cosine_tc_kernel<bm::SsBankProblem><<<dim3((st.Pcap + bm::TC_M - 1) / bm::TC_M, st.S), bm::TC_NT,
                                              bm::cosine_tc_smem_bytes(), s>>>(pb);
    if (tm) cudaEventRecord(e->ev[2], s);
    bm::ss_cost_kernel<<<dim3(8, st.S), 256, (size_t)st.Dcap * 7 * sizeof(double), s>>>(st, pr);
    if (tm) cudaEventRecord(e->ev[3], s);
    bm::ss_match_kernel<<<(st.S + bm::SS_MATCH_WPB - 1) / bm::SS_MATCH_WPB, bm::SS_MATCH_WPB * 32, 0, s>>>(st, pr);
    if (tm) cudaEventRecord(e->ev[4], s);
    bm::ss_update_kernel<<<dim3((st.Dcap + 3) / 4, st.S), 128, 0, s>>>(st, pr, d_embs);
    if (tm) cudaEventRecord(e->ev[5], s);
    // one CTA per stream at 64 registers: two 512-thread CTAs per SM while every stream gets a slot at once, else
    // four of 256 threads (the kernel is a chain of dependent global-memory phases: residency, not width, hides it)
    bm::ss_finalize_kernel<<<st.S, e->st.S > 2 * e->n_sm ? 256 : 512, 0, s>>>(st, pr, d_embs, d_out, d_nout, st.Dcap);
    bm::ss_bank_kernel<<<dim3(4, st.S), 256, 0, s>>>(st, pr);
    bm::ss_live_kernel<<<st.S, 256, 0, s>>>(st);
    if (tm) { cudaEventRecord(e->ev[6], s); e->timed = true; }
    return cudaGetLastError();
}

int bm_strongsort_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                       const double* d_warps, double* d_out, int* d_nout, void* stream) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e || !d_dets || !d_ndets || !d_embs || !d_out || !d_nout) return ss_fail(-1, "bm_strongsort_step: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int G = (int)e->gstreams.size();
    if (G <= 1 || e->timing) {                     // per-kernel event timing always uses the single chain
        SS_CUDA(ss_launch_chain(e, e->st, d_dets, d_ndets, d_embs, d_warps, d_out, d_nout, s, e->timing));
        return 0;
    }
    // fork from the caller's stream into the group streams and join back: no host synchronisation.  Measured at
    // configs[3] (512 streams): 6.58 ms with one chain, 6.89 with 2-8 groups, 7.5 with group g held back until group
    // g - 1 enters its cascade -- a cascade CTA's 31 KB of shared memory takes a contraction CTA's place on its SM,
    // so the overlap costs what it gains.  Kept for hosts that partition their streams anyway; default 1.
    SS_CUDA(cudaEventRecord(e->fork, s));
    for (int g = 0; g < G; g++) {
        bm::SsState sub = e->st;
        sub.s0 = (int)((long long)e->st.S * g / G);
        sub.S = (int)((long long)e->st.S * (g + 1) / G) - sub.s0;
        if (sub.S <= 0) continue;
        SS_CUDA(cudaStreamWaitEvent(e->gstreams[g], e->fork, 0));
        SS_CUDA(ss_launch_chain(e, sub, d_dets, d_ndets, d_embs, d_warps, d_out, d_nout, e->gstreams[g], false));
        SS_CUDA(cudaEventRecord(e->gdone[g], e->gstreams[g]));
        SS_CUDA(cudaStreamWaitEvent(s, e->gdone[g], 0));
    }
    return 0;
}

int bm_strongsort_set_groups(void* handle, int n_groups) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e || n_groups < 1) return ss_fail(-1, "bm_strongsort_set_groups: bad argument");
    n_groups = std::min(n_groups, std::min(e->st.S, 16));
    SS_CUDA(cudaSetDevice(e->device));
    for (cudaStream_t g : e->gstreams) cudaStreamDestroy(g);
    for (cudaEvent_t g : e->gdone) cudaEventDestroy(g);
    e->gstreams.clear();
    e->gdone.clear();
    if (!e->fork) SS_CUDA(cudaEventCreateWithFlags(&e->fork, cudaEventDisableTiming));
    if (n_groups > 1)
        for (int g = 0; g < n_groups; g++) {
            cudaStream_t st;
            cudaEvent_t ev;
            SS_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            e->gstreams.push_back(st);
            SS_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            e->gdone.push_back(ev);
        }
    return n_groups;
}

int bm_strongsort_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e) return ss_fail(-1, "bm_strongsort_counters: null handle");
    const bm::SsState& st = e->st;
    SS_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    std::vector<int> hdr((size_t)st.S * bm::SH_COUNT);
    std::vector<unsigned long long> nu(st.S);
    SS_CUDA(cudaMemcpy(hdr.data(), st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost));
    SS_CUDA(cudaMemcpy(nu.data(), st.n_updates, nu.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    int err = 0;
    unsigned long long o[4] = {0, 0, 0, 0};
    for (int s = 0; s < st.S; s++) {
        const int* h = hdr.data() + (size_t)s * bm::SH_COUNT;
        err |= h[bm::SH_ERR];
        o[0] += nu[s]; o[1] += (unsigned)h[bm::SH_LSA]; o[2] += (unsigned)h[bm::SH_MATCH]; o[3] += (unsigned)h[bm::SH_BIRTHS];
    }
    if (err_bits) *err_bits = err;
    if (out4) for (int i = 0; i < 4; i++) out4[i] = o[i];
    return 0;
}

int bm_strongsort_usage(void* handle, int* max_tracks, int* max_pool_used, int* max_live, void* stream) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e) return ss_fail(-1, "bm_strongsort_usage: null handle");
    const bm::SsState& st = e->st;
    SS_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    std::vector<int> hdr((size_t)st.S * bm::SH_COUNT);
    SS_CUDA(cudaMemcpy(hdr.data(), st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost));
    int mt = 0, mp = 0, ml = 0;
    for (int s = 0; s < st.S; s++) {
        const int* h = hdr.data() + (size_t)s * bm::SH_COUNT;
        mt = std::max(mt, h[bm::SH_NTRACKS] + h[bm::SH_NLOST]);
        mp = std::max(mp, st.Pcap - h[bm::SH_NPFREE]);
        ml = std::max(ml, h[bm::SH_NLIVE]);
    }
    if (max_tracks) *max_tracks = mt;
    if (max_pool_used) *max_pool_used = mp;
    if (max_live) *max_live = ml;
    return 0;
}

int bm_strongsort_debug_copy(void* handle, int which, int stream_index, void* h_dst, size_t nbytes, void* stream) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e || stream_index < 0 || stream_index >= e->st.S || !h_dst) return ss_fail(-1, "bm_strongsort_debug_copy: bad argument");
    const bm::SsState& st = e->st;
    const size_t s = stream_index;
    const void* src = nullptr;
    size_t avail = 0;
    if (which == 0) { src = st.wki + s * st.wki_stride; avail = st.wki_stride * 4; }
    else if (which == 1) { src = st.cost + s * st.Tcap * st.Dcap; avail = (size_t)st.Tcap * st.Dcap * 8; }
    else if (which == 2) { src = st.dtab + s * st.Dcap * bm::DT_COUNT; avail = (size_t)st.Dcap * bm::DT_COUNT * 8; }
    else if (which == 3) { src = st.wkd + s * 3 * st.M; avail = (size_t)3 * st.M * 8; }
    else return ss_fail(-1, "bm_strongsort_debug_copy: unknown buffer");
    SS_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    SS_CUDA(cudaMemcpy(h_dst, src, nbytes < avail ? nbytes : avail, cudaMemcpyDeviceToHost));
    return 0;
}

int bm_strongsort_time_kernels(void* handle, int enable) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e) return ss_fail(-1, "bm_strongsort_time_kernels: null handle");
    e->timing = enable != 0; e->timed = false;
    return 0;
}

int bm_strongsort_kernel_times(void* handle, float* ms6) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e || !ms6) return ss_fail(-1, "bm_strongsort_kernel_times: bad argument");
    if (!e->timed) return ss_fail(-1, "bm_strongsort_kernel_times: no timed step");
    SS_CUDA(cudaEventSynchronize(e->ev[6]));
    for (int i = 0; i < 6; i++) SS_CUDA(cudaEventElapsedTime(&ms6[i], e->ev[i], e->ev[i + 1]));
    return 0;
}

int bm_strongsort_peek(void* handle, int stream_index, int* n_tracks, int* n_lost, double* h_mean, double* h_cov,
                       double* h_tab, int* h_ids, int* h_nbank, void* stream) {
    SsEngine* e = static_cast<SsEngine*>(handle);
    if (!e || stream_index < 0 || stream_index >= e->st.S) return ss_fail(-1, "bm_strongsort_peek: bad argument");
    const bm::SsState& st = e->st;
    const size_t s = stream_index, T = st.Tcap;
    SS_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    std::vector<int> hdr(bm::SH_COUNT), order(T), tint(T * bm::TI_COUNT);
    std::vector<double> mean(T * 8), covb(T * bm::XA_P), tab(T * bm::ST_FIELDS);
    SS_CUDA(cudaMemcpy(hdr.data(), st.hdr + s * bm::SH_COUNT, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost));
    SS_CUDA(cudaMemcpy(order.data(), st.order + s * T, T * sizeof(int), cudaMemcpyDeviceToHost));
    SS_CUDA(cudaMemcpy(tint.data(), st.tint + s * T * bm::TI_COUNT, tint.size() * sizeof(int), cudaMemcpyDeviceToHost));
    SS_CUDA(cudaMemcpy(mean.data(), st.mean + s * T * 8, mean.size() * sizeof(double), cudaMemcpyDeviceToHost));
    SS_CUDA(cudaMemcpy(covb.data(), st.covb + s * T * bm::XA_P, covb.size() * sizeof(double), cudaMemcpyDeviceToHost));
    SS_CUDA(cudaMemcpy(tab.data(), st.tab + s * T * bm::ST_FIELDS, tab.size() * sizeof(double), cudaMemcpyDeviceToHost));
    const int n = hdr[bm::SH_NTRACKS];
    if (n_tracks) *n_tracks = n;
    if (n_lost) *n_lost = hdr[bm::SH_NLOST];
    for (int p = 0; p < n; p++) {
        const int t = order[p];
        if (h_mean) for (int c = 0; c < 8; c++) h_mean[(size_t)p * 8 + c] = mean[(size_t)t * 8 + c];
        if (h_cov) bm::xyah_expand(covb.data() + (size_t)t * bm::XA_P, h_cov + (size_t)p * 64);
        if (h_tab) for (int c = 0; c < bm::ST_FIELDS; c++) h_tab[(size_t)p * bm::ST_FIELDS + c] = tab[(size_t)t * bm::ST_FIELDS + c];
        if (h_ids) h_ids[p] = tint[(size_t)t * bm::TI_COUNT + bm::TI_ID];
        if (h_nbank) h_nbank[p] = tint[(size_t)t * bm::TI_COUNT + bm::TI_NBANK];
    }
    return 0;
}

}  // extern "C"
