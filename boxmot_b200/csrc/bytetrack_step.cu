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
#include <cstdio>

namespace bm {

constexpr int BT_NT = 64;
enum { TF_MATCHED = 1, TF_NEWLOST = 2, TF_REFIND = 4, TF_DUP = 8, TF_USED = 16, TF_AGED = 32 };
enum { BM_NH = 0, BM_NLO, BM_NTR, BM_NP, BM_NU, BM_NUPD, BM_NR, BM_NREST, BM_NBORN, BM_NNL, BM_NRF, BM_NA2, BM_NL2,
       BM_BAD, BM_MISC = 16 };

struct ByteSmem {
    float4 *hbox, *lbox;      // [Dcap] raw high / low score detections
    float* hconf;             // [Dcap]
    void* work;               // solver work area
    int *hind, *lind, *dflag, *rest, *rflag, *born;  // [Dcap]
    int *pool, *unc, *rtr, *xs, *nlost, *refind, *lista, *listb, *tflag, *upd_slot, *upd_det;   // [Tcap]
    int* misc;
};

__host__ __device__ inline size_t byte_smem_bytes(int Tcap, int Dcap) {
    size_t b = 2 * 16 * (size_t)Dcap + (rect_work_bytes(Tcap, Dcap) + 15) / 16 * 16;
    b += 4 * ((size_t)Dcap + 6 * (size_t)Dcap + 11 * (size_t)Tcap + BM_MISC);
    return (b + 15) / 16 * 16;
}

__device__ inline ByteSmem byte_carve(unsigned char* base, int Tcap, int Dcap) {
    ByteSmem m;
    m.hbox = reinterpret_cast<float4*>(base); base += 16 * (size_t)Dcap;
    m.lbox = reinterpret_cast<float4*>(base); base += 16 * (size_t)Dcap;
    m.work = base; base += (rect_work_bytes(Tcap, Dcap) + 15) / 16 * 16;
    m.hconf = reinterpret_cast<float*>(base); base += 4 * (size_t)Dcap;
    int* q = reinterpret_cast<int*>(base);
    m.hind = q; q += Dcap; m.lind = q; q += Dcap; m.dflag = q; q += Dcap; m.rest = q; q += Dcap; m.rflag = q; q += Dcap; m.born = q; q += Dcap;
    m.pool = q; q += Tcap; m.unc = q; q += Tcap; m.rtr = q; q += Tcap; m.xs = q; q += Tcap; m.nlost = q; q += Tcap;
    m.refind = q; q += Tcap; m.lista = q; q += Tcap; m.listb = q; q += Tcap; m.tflag = q; q += Tcap;
    m.upd_slot = q; q += Tcap; m.upd_det = q; q += Tcap;
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

struct ByteView {
    double *xP, *confcls, *tbox;
    int *ints, *act, *lost, *hdr;
    int Tcap;
    __device__ ByteView(const ByteState& st, int s) : Tcap(st.Tcap) {
        const size_t T = st.Tcap;
        xP = st.xP + (size_t)s * XA_N * T;
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
        for (int c = 0; c < 12; c++) P[c] = xP[(8 + c) * Tcap + slot];
    }
    __device__ __forceinline__ void store(int slot, const double* x, const double* P) const {
#pragma unroll
        for (int c = 0; c < 8; c++) xP[c * Tcap + slot] = x[c];
#pragma unroll
        for (int c = 0; c < 12; c++) xP[(8 + c) * Tcap + slot] = P[c];
    }
};

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
    const int bad = rect_solve_warp(DenseCost{cost, nc}, nr, nc, thresh, w, nullptr, nullptr);
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

// <= 146 registers: seven CTAs per SM keep all 1024 streams of the bench in one wave on 148 SMs
__global__ void __launch_bounds__(BT_NT, 7)
byte_step_kernel(ByteState st, ByteParams pr, const float* __restrict__ dets, const int* __restrict__ ndets,
                 double* __restrict__ out, int* __restrict__ nout) {
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
    for (int p = tid; p < nP; p += BT_NT) {
        const int slot = sm.pool[p];
        double x[8], P[12];
        v.load(slot, x, P);
        if (v.I(BI_STATE, slot) != BT_TRACKED) x[7] = 0.0;
        xyah_predict(x, P);
        v.store(slot, x, P);
        v.set_box(slot, xyah_to_xyxy(x));
        sm.tflag[slot] |= TF_USED;
    }
    for (int p = tid; p < nU; p += BT_NT) sm.tflag[sm.unc[p]] |= TF_USED;   // box is current (set at birth)
    if (tid == 0) st.n_updates[s] += (unsigned long long)nP;
    __syncthreads();

    // ---- first association: pool x high-score detections, fuse_score, thresh = match_thresh (:199-219)
    for (int e = tid; e < nP * nH; e += BT_NT) {
        const int t = e / nH, d = e % nH;
        cost[e] = fuse_entry(1.0 - iou_pair(v.box(sm.pool[t]), det_xyxy(sm.hbox[d])), (double)sm.hconf[d]);
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
        if (lane == 0) { sm.misc[BM_NUPD] = n1; sm.misc[BM_NRF] = nrf; sm.misc[BM_NR] = nr; sm.misc[BM_NREST] = nrest; }
    }
    __syncthreads();

    // ---- second association: still-tracked leftovers x low-score detections, thresh 0.5 (:221-250)
    const int nR = sm.misc[BM_NR], nRest = sm.misc[BM_NREST];
    for (int e = tid; e < nR * nLo; e += BT_NT) {
        const int t = e / nLo, d = e % nLo;
        cost[e] = 1.0 - iou_pair(v.box(sm.rtr[t]), det_xyxy(sm.lbox[d]));
    }
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
        if (lane == 0) { sm.misc[BM_NUPD] = n1; sm.misc[BM_NNL] = nnl; }
    }
    __syncthreads();

    // ---- unconfirmed tracks x remaining high-score detections, fuse_score, thresh 0.7 (:252-264)
    for (int e = tid; e < nU * nRest; e += BT_NT) {
        const int t = e / nRest, k = e % nRest, d = sm.rest[k];
        cost[e] = fuse_entry(1.0 - iou_pair(v.box(sm.unc[t]), det_xyxy(sm.hbox[d])), (double)sm.hconf[d]);
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
        double x[8], P[12], z[4];
        int slot, drow;
        if (k < nUpd) {
            slot = sm.upd_slot[k]; drow = sm.upd_det[k];
        } else {
            slot = sm.rtr[k - nUpd]; drow = sm.hind[sm.born[k - nUpd]];
        }
        const float* r = dsrc + drow * 6;
        det_xyah(make_float4(r[0], r[1], r[2], r[3]), z);
        if (k < nUpd) {
            v.load(slot, x, P);
            xyah_update(x, P, z);
            v.I(BI_ACTIVATED, slot) = 1;
        } else {
            xyah_initiate(z, x, P);
            v.I(BI_ID, slot) = id0 + (k - nUpd) + 1;
            v.I(BI_ACTIVATED, slot) = frame == 1 ? 1 : 0;
            v.I(BI_START, slot) = frame;
            v.I(BI_INREMOVED, slot) = 0;
        }
        v.store(slot, x, P);
        v.set_box(slot, xyah_to_xyxy(x));
        v.I(BI_STATE, slot) = BT_TRACKED;
        v.I(BI_FRAME, slot) = frame;
        v.I(BI_DETIND, slot) = drow;
        v.confcls[slot] = (double)r[4];
        v.confcls[Tcap + slot] = (double)r[5];
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
        if (1.0 - iou_pair(v.box(a), v.box(b)) < 0.15) {
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

size_t bytetrack_step_smem(const ByteState& st) { return byte_smem_bytes(st.Tcap, st.Dcap); }

cudaError_t bytetrack_step_configure(size_t smem) {
    return cudaFuncSetAttribute(byte_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

cudaError_t bytetrack_step_launch(const ByteState& st, const ByteParams& pr, const float* dets, const int* ndets,
                                  double* out, int* nout, cudaStream_t stream) {
    byte_step_kernel<<<st.S, BT_NT, bytetrack_step_smem(st), stream>>>(st, pr, dets, ndets, out, nout);
    return cudaGetLastError();
}

}  // namespace bm
