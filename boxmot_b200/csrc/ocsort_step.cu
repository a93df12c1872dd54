// OC-SORT frame step: ONE launch advances every stream of the engine by one frame
// (one CTA per stream).  Mirrors OcSort.update (boxmot/trackers/ocsort/ocsort.py:246-439)
// and enhanced_associate (boxmot/utils/association.py:377-533 with repairs R1-R3).
//
// Phases inside the CTA (all state decisions are made on the device, no host sync):
//   A  split detections by confidence (ocsort.py:265-275)
//   B  Kalman predict of every live track (ocsort.py:281-288, :173-186)
//   C  IoU + velocity-direction (OCM) cost matrix, permutation fast path or lapjv
//   D  Kalman update (+ORU) of matched tracks (ocsort.py:321-324)
//   E  optional BYTE round (ocsort.py:330-356) and OCR round (ocsort.py:358-386)
//   F  misses, births, output rows and dead-track removal (ocsort.py:388-439)
#include "ocsort_engine.cuh"
#include "lap_rect.cuh"

namespace bm {

constexpr int OCS_NT = 128;

struct OcsSmem {
    double *dbox, *dconf, *dcls, *sbox, *sconf, *scls, *tbox, *kcx, *kcy, *vy, *vx, *wred, *rmin;
    int *dind, *sind, *slot_of, *kvalid, *d2t, *t2d, *um_d, *um_t, *rowcnt, *colcnt, *hit_t, *misc, *rarg;
    void* lapw;
    double* cost;
};

enum { M_NT = 0, M_ND, M_NS, M_NUD, M_NUT, M_FAST, M_ANYBAD, M_OCR_RAN, M_NREJ, M_LAPBAD, M_NMATCH, M_COUNT = 16 };

__host__ __device__ inline size_t ocs_smem_bytes(int Tcap, int Dcap, int use_byte, int cost_elems) {
    size_t d = (size_t)Dcap, t = (size_t)Tcap;
    size_t dbl = 7 * d + (use_byte ? 6 * d : 0) + 8 * t + 8;
    size_t ints = 2 * d + (use_byte ? d : 0) + 2 * t + d + t + d + t + d + t + d + M_COUNT;
    size_t b = dbl * 8 + ((ints * 4 + 7) / 8) * 8;
    b += (lap_work_bytes((int)(d + t)) + 7) / 8 * 8;
    b += (size_t)cost_elems * 8;
    return b;
}

__device__ inline OcsSmem ocs_carve(unsigned char* base, int Tcap, int Dcap, int use_byte, int cost_elems) {
    OcsSmem m;
    double* p = reinterpret_cast<double*>(base);
    m.dbox = p; p += 4 * Dcap;
    m.dconf = p; p += Dcap;
    m.dcls = p; p += Dcap;
    if (use_byte) { m.sbox = p; p += 4 * Dcap; m.sconf = p; p += Dcap; m.scls = p; p += Dcap; }
    else { m.sbox = m.sconf = m.scls = nullptr; }
    m.tbox = p; p += 4 * Tcap;
    m.kcx = p; p += Tcap;
    m.kcy = p; p += Tcap;
    m.vy = p; p += Tcap;
    m.vx = p; p += Tcap;
    m.wred = p; p += 8;
    m.rmin = p; p += Dcap;
    int* q = reinterpret_cast<int*>(p);
    int* q0 = q;
    m.dind = q; q += Dcap;
    if (use_byte) { m.sind = q; q += Dcap; } else m.sind = nullptr;
    m.slot_of = q; q += Tcap;
    m.kvalid = q; q += Tcap;
    m.d2t = q; q += Dcap;
    m.t2d = q; q += Tcap;
    m.um_d = q; q += Dcap;
    m.um_t = q; q += Tcap;
    m.rowcnt = q; q += Dcap;
    m.colcnt = q; q += Tcap;
    m.hit_t = q; q += Dcap;
    m.misc = q; q += M_COUNT;
    m.rarg = q; q += Dcap;
    size_t ib = ((size_t)(q - q0) * 4 + 7) / 8 * 8;
    unsigned char* r = reinterpret_cast<unsigned char*>(q0) + ib;
    m.lapw = r;
    r += (lap_work_bytes(Dcap + Tcap) + 7) / 8 * 8;
    m.cost = cost_elems > 0 ? reinterpret_cast<double*>(r) : nullptr;
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

// Second-round association shared by BYTE and OCR: rows = the nA detections listed in
// `alist` (indices into a detection set `abox`), cols = the nB list positions in
// `blist` with boxes fetched by `bfetch`.  lapjv on -iou, accept pairs with iou >= thr
// (ocsort.py:336-353, :363-380).  Writes pair_b[r] = column index or -1.  Returns
// whether the block ran (max iou > thr and no NaN).
template <typename BFetch>
__device__ bool ocs_second_round(const OcsState& st, const OcsParams& pr, int s, OcsSmem& sm,
                                 const double* abox, int Dld, const int* alist, int nA,
                                 const int* blist, int nB, BFetch bfetch, int* pair_b) {
    const int tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const size_t need = (size_t)nA * nB;
    double* cost = (need <= (size_t)st.smem_cost_elems) ? sm.cost : st.scratch + (size_t)s * st.scratch_elems;
    if (need > st.scratch_elems && need > (size_t)st.smem_cost_elems) {
        if (tid == 0) atomicOr(&st.hdr[(size_t)s * OH_COUNT + OH_ERR], BM_ERR_SCRATCH);
        for (int r = tid; r < nA; r += OCS_NT) pair_b[r] = -1;
        __syncthreads();
        return false;
    }
    double mx = -CUDART_INF;   // max of iou_left
    double cmx = -CUDART_INF;  // max of -iou_left
    int nanf = 0;
    for (size_t e = tid; e < need; e += OCS_NT) {
        const int r = (int)(e / nB), c = (int)(e % nB);
        const int d = alist[r];
        Box a = {abox[0 * Dld + d], abox[1 * Dld + d], abox[2 * Dld + d], abox[3 * Dld + d]};
        Box b = bfetch(blist[c]);
        const double iou = asso_pair(pr.asso_func, a, b, pr.diag);
        if (iou != iou) nanf = 1;
        mx = fmax(mx, iou);
        const double cc = -iou;
        cmx = fmax(cmx, cc);
        cost[e] = cc;
    }
    // block reductions (max, max, or)
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, shfl_xor_d(mx, o));
        cmx = fmax(cmx, shfl_xor_d(cmx, o));
        nanf |= __shfl_xor_sync(BM_FULL, nanf, o);
    }
    __shared__ double s_mx[OCS_NT / 32], s_cmx[OCS_NT / 32];
    __shared__ int s_nan[OCS_NT / 32];
    if (lane == 0) { s_mx[wid] = mx; s_cmx[wid] = cmx; s_nan[wid] = nanf; }
    __syncthreads();
    mx = s_mx[0]; cmx = s_cmx[0]; nanf = s_nan[0];
    for (int i = 1; i < OCS_NT / 32; i++) { mx = fmax(mx, s_mx[i]); cmx = fmax(cmx, s_cmx[i]); nanf |= s_nan[i]; }
    __syncthreads();
    const bool run = !nanf && mx > pr.iou_thr;
    if (!run) {
        for (int r = tid; r < nA; r += OCS_NT) pair_b[r] = -1;
        __syncthreads();
        return false;
    }
    if (wid == 0) {
        // Only pairs with iou >= thr survive, and the caller re-sorts what is left: any
        // optimal solution gives the same result, so the fast exact solver is used.
        RectWork w = rect_carve(sm.lapw, nA, nB);
        const int bad = rect_solve_warp(cost, nB, nA, nB, 2.0 * (cmx + 1.0), w, nullptr, nullptr);
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

}  // namespace bm

namespace bm {

struct FetchPred {
    const double* tbox; int Tcap;
    __device__ Box operator()(int p) const {
        return Box{tbox[0 * Tcap + p], tbox[1 * Tcap + p], tbox[2 * Tcap + p], tbox[3 * Tcap + p]};
    }
};
struct FetchLast {
    const double* lastobs; const int* slot_of; int Tcap;
    __device__ Box operator()(int p) const {
        const int sl = slot_of[p];
        return Box{lastobs[0 * Tcap + sl], lastobs[1 * Tcap + sl], lastobs[2 * Tcap + sl], lastobs[3 * Tcap + sl]};
    }
};

// dets: [S][Dcap][6] float32 (x1,y1,x2,y2,conf,cls) -- float32 IS the reference's
// precision: BaseTracker.setup_decorator round-trips every ndarray through float32
// (boxmot/trackers/basetracker.py:122-128).  out: [S][Tcap][8] float64 rows
// (x1,y1,x2,y2,id,conf,cls,det_ind) in the reference's row order; nout[S].
__global__ void __launch_bounds__(OCS_NT)
ocsort_step_kernel(OcsState st, OcsParams pr, const float* __restrict__ dets,
                   const int* __restrict__ ndets, double* __restrict__ out, int* __restrict__ nout) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int s = blockIdx.x, tid = threadIdx.x, lane = lane_id(), wid = warp_id();
    const int Tcap = st.Tcap, Dcap = st.Dcap;
    OcsSmem sm = ocs_carve(smem_raw, Tcap, Dcap, pr.use_byte, st.smem_cost_elems);
    OcsView v(st, s);
    const XysrConst kc = {pr.q_xy, pr.q_s};
    const float* dsrc = dets + (size_t)s * Dcap * 6;

    // ------------------------------------------------------------------ phase A
    int nD_in = ndets[s];
    if (nD_in > Dcap) { if (tid == 0) atomicOr(&v.hdr[OH_ERR], BM_ERR_DET_OVERFLOW); nD_in = Dcap; }
    if (tid < M_COUNT) sm.misc[tid] = 0;
    __syncthreads();
    if (wid == 0) {   // stable compaction of the high (and second-round) detections
        int nh = 0, ns = 0;
        for (int base = 0; base < nD_in; base += 32) {
            const int i = base + lane;
            double conf = -1.0;
            if (i < nD_in) conf = (double)dsrc[i * 6 + 4];
            const bool hi = i < nD_in && conf > pr.det_thresh;
            const bool se = pr.use_byte && i < nD_in && conf > pr.min_conf && conf < pr.det_thresh;
            const unsigned bh = __ballot_sync(BM_FULL, hi), bs = __ballot_sync(BM_FULL, se);
            const unsigned lt = (1u << lane) - 1;
            if (hi) {
                const int p = nh + __popc(bh & lt);
                for (int c = 0; c < 4; c++) sm.dbox[c * Dcap + p] = (double)dsrc[i * 6 + c];
                sm.dconf[p] = conf; sm.dcls[p] = (double)dsrc[i * 6 + 5]; sm.dind[p] = i;
            }
            if (se) {
                const int p = ns + __popc(bs & lt);
                for (int c = 0; c < 4; c++) sm.sbox[c * Dcap + p] = (double)dsrc[i * 6 + c];
                sm.sconf[p] = conf; sm.scls[p] = (double)dsrc[i * 6 + 5]; sm.sind[p] = i;
            }
            nh += __popc(bh); ns += __popc(bs);
        }
        if (lane == 0) { sm.misc[M_ND] = nh; sm.misc[M_NS] = ns; }
    }
    int nT = v.hdr[OH_NLIST];
    const int frame = v.hdr[OH_FRAME] + 1;
    __syncthreads();
    const int nD = sm.misc[M_ND], nS = sm.misc[M_NS];
    if (tid == 0) st.n_updates[s] += (unsigned long long)nT;

    // ------------------------------------------------------------------ phase B
    for (int p = tid; p < nT; p += OCS_NT) {
        const int slot = v.order[p];
        double x[7], P[49];
        load_xP(v.xP, Tcap, slot, x, P);
        if (x[6] + x[2] <= 0) x[6] *= 0.0;
        xysr_predict(x, P, kc);
        store_xP(v.xP, Tcap, slot, x, P);
        v.I(OI_AGE, slot) += 1;
        const int tsu = v.I(OI_TSU, slot);
        if (tsu > 0) v.I(OI_STREAK, slot) = 0;
        v.I(OI_TSU, slot) = tsu + 1;
        double b[4];
        x_to_xyxy(x, b);
        if (b[0] != b[0] || b[1] != b[1] || b[2] != b[2] || b[3] != b[3]) {
            sm.t2d[p] = -2;   // NaN prediction: the reference drops the track (ocsort.py:284-288)
            sm.misc[M_ANYBAD] = 1;
        } else sm.t2d[p] = -1;
    }
    __syncthreads();
    if (sm.misc[M_ANYBAD]) {
        if (tid == 0) {
            int k = 0;
            for (int p = 0; p < nT; p++) if (sm.t2d[p] != -2) v.order[k++] = v.order[p];
            sm.misc[M_NT] = k;
        }
        __syncthreads();
        nT = sm.misc[M_NT];
    }
    // per-position operands of the association
    for (int p = tid; p < nT; p += OCS_NT) {
        const int slot = v.order[p];
        sm.slot_of[p] = slot;
        double x[7];
#pragma unroll
        for (int c = 0; c < 7; c++) x[c] = v.xP[c * Tcap + slot];
        double b[4];
        x_to_xyxy(x, b);
#pragma unroll
        for (int c = 0; c < 4; c++) sm.tbox[c * Tcap + p] = b[c];
        sm.vy[p] = v.vel[0 * Tcap + slot];
        sm.vx[p] = v.vel[1 * Tcap + slot];
        double kb[4];
        const bool ok = ocs_lookback(v, slot, v.I(OI_AGE, slot), pr.delta_t, kb);
        sm.kvalid[p] = ok ? 1 : 0;
        // placeholder observation is [-1,-1,-1,-1,-1] -> centre (-1,-1)
        sm.kcx[p] = ok ? (kb[0] + kb[2]) / 2.0 : -1.0;
        sm.kcy[p] = ok ? (kb[1] + kb[3]) / 2.0 : -1.0;
        sm.t2d[p] = -1;
        sm.colcnt[p] = 0;
    }
    for (int d = tid; d < nD; d += OCS_NT) { sm.d2t[d] = -1; sm.rowcnt[d] = 0; sm.hit_t[d] = -1; }
    __syncthreads();

    // ------------------------------------------------------------------ phase C
    // cost[d][t] = -(iou + OCM), det-major like enhanced_associate (association.py:427-461)
    const size_t need = (size_t)nD * nT;
    double* cost = (need <= (size_t)st.smem_cost_elems) ? sm.cost : st.scratch + (size_t)s * st.scratch_elems;
    bool assoc_ok = true;
    if (need > (size_t)st.smem_cost_elems && need > st.scratch_elems) {
        if (tid == 0) atomicOr(&v.hdr[OH_ERR], BM_ERR_SCRATCH);
        assoc_ok = false;
    }
    if (nD > 0 && nT > 0 && assoc_ok) {
        double cmx = -CUDART_INF;
        for (int d = wid; d < nD; d += OCS_NT / 32) {      // one warp per detection row
            const Box a = {sm.dbox[0 * Dcap + d], sm.dbox[1 * Dcap + d], sm.dbox[2 * Dcap + d], sm.dbox[3 * Dcap + d]};
            const double cxd = (a.x1 + a.x2) / 2.0, cyd = (a.y1 + a.y2) / 2.0;
            double rm = CUDART_INF;
            int ra = 0x7fffffff, nhit = 0, lasthit = -1;
            for (int t = lane; t < nT; t += 32) {
                const Box b = {sm.tbox[0 * Tcap + t], sm.tbox[1 * Tcap + t], sm.tbox[2 * Tcap + t], sm.tbox[3 * Tcap + t]};
                const double iou = asso_pair(pr.asso_func, a, b, pr.diag);
                // speed_direction_batch(dets, previous_obs)  (association.py:10-20)
                double dx = cxd - sm.kcx[t], dy = cyd - sm.kcy[t];
                const double norm = sqrt(dx * dx + dy * dy) + 1e-6;
                dx = dx / norm; dy = dy / norm;
                double cs = sm.vx[t] * dx + sm.vy[t] * dy;
                cs = fmin(1.0, fmax(-1.0, cs));
                const double ang = (CUDART_PI / 2.0 - fabs(acos_pinned(cs))) / CUDART_PI;
                const double motion = ((sm.kvalid[t] ? 1.0 : 0.0) * ang) * pr.inertia;
                const double cc = -(iou + motion);
                cost[(size_t)d * nT + t] = cc;
                cmx = fmax(cmx, cc);
                if (cc < rm) { rm = cc; ra = t; }
                if (iou > pr.iou_thr) { nhit++; lasthit = t; atomicAdd(&sm.colcnt[t], 1); }
            }
            warp_argmin(rm, ra);
            for (int o = 16; o > 0; o >>= 1) {
                nhit += __shfl_xor_sync(BM_FULL, nhit, o);
                lasthit = max(lasthit, __shfl_xor_sync(BM_FULL, lasthit, o));
            }
            if (lane == 0) { sm.rmin[d] = rm; sm.rarg[d] = ra; sm.rowcnt[d] = nhit; sm.hit_t[d] = lasthit; }
        }
        for (int o = 16; o > 0; o >>= 1) cmx = fmax(cmx, shfl_xor_d(cmx, o));
        if (lane == 0) sm.wred[wid] = cmx;
        __syncthreads();
        // permutation fast path: every row/column has at most one IoU hit, at least one exists
        if (wid == 0) {
            int rmax = 0, cmax_ = 0;
            for (int d = lane; d < nD; d += 32) rmax = max(rmax, sm.rowcnt[d]);
            for (int t = lane; t < nT; t += 32) cmax_ = max(cmax_, sm.colcnt[t]);
            for (int o = 16; o > 0; o >>= 1) {
                rmax = max(rmax, __shfl_xor_sync(BM_FULL, rmax, o));
                cmax_ = max(cmax_, __shfl_xor_sync(BM_FULL, cmax_, o));
            }
            const bool fast = rmax == 1 && cmax_ == 1;
            if (fast) {
                for (int d = lane; d < nD; d += 32) {
                    const int t = sm.rowcnt[d] == 1 ? sm.hit_t[d] : -1;
                    sm.d2t[d] = t;
                    sm.rowcnt[d] = t >= 0 ? 1 : 0;           // 1 = accepted match
                    if (t >= 0) sm.t2d[t] = d;
                }
            } else {
                double pad = sm.wred[0];
                for (int i = 1; i < OCS_NT / 32; i++) pad = fmax(pad, sm.wred[i]);
                // With nD <= nT every detection is assigned in every optimal solution, so the
                // reference's unmatched-detection ORDER (never-assigned first, then rejected,
                // ocsort ids follow it) does not depend on lapjv's tie-breaking and the fast
                // exact solver applies.  With nD > nT WHICH zero-gain detections stay
                // unassigned is decided by lapjv's scan order: emulate it step for step.
                int bad;
                int* xsol;
                if (nD <= nT) {
                    RectWork w = rect_carve(sm.lapw, nD, nT);
                    bad = rect_solve_warp(cost, nT, nD, nT, 2.0 * (pad + 1.0), w, sm.rmin, sm.rarg);
                    xsol = w.x;
                } else {
                    LapMat m = {cost, nT, nD, nT, nD + nT, 1, pad + 1.0};
                    LapWork w = lap_carve(sm.lapw, nD + nT);
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
                        Box a = {sm.dbox[0 * Dcap + d], sm.dbox[1 * Dcap + d], sm.dbox[2 * Dcap + d], sm.dbox[3 * Dcap + d]};
                        Box b = {sm.tbox[0 * Tcap + t], sm.tbox[1 * Tcap + t], sm.tbox[2 * Tcap + t], sm.tbox[3 * Tcap + t]};
                        const double iou = asso_pair(pr.asso_func, a, b, pr.diag);
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
        }
        __syncthreads();
    } else {
        for (int d = tid; d < nD; d += OCS_NT) { sm.rowcnt[d] = 0; sm.hit_t[d] = -1; }
        __syncthreads();
    }
    // unmatched lists in the reference's order (repair R3): never-assigned ascending,
    // then IoU-rejected pairs in detection-row order.
    if (tid == 0) {
        int nud = 0, nut = 0;
        for (int d = 0; d < nD; d++) if (sm.rowcnt[d] == 0) sm.um_d[nud++] = d;
        if (nT > 0) {
            for (int t = 0; t < nT; t++) {
                // a track is "never assigned" when no detection row holds it (accepted or rejected)
                const bool rejected_col = sm.colcnt[t] < 0 && sm.misc[M_FAST] == 0 && nD > 0;
                if (sm.t2d[t] < 0 && !rejected_col) sm.um_t[nut++] = t;
            }
            for (int d = 0; d < nD; d++) if (sm.rowcnt[d] == 2) { sm.um_d[nud++] = d; sm.um_t[nut++] = sm.hit_t[d]; }
        }
        sm.misc[M_NUD] = nud; sm.misc[M_NUT] = nut;
    }
    __syncthreads();

    // ------------------------------------------------------------------ phase D
    for (int d = tid; d < nD; d += OCS_NT) {
        const int t = sm.d2t[d];
        if (t >= 0) {
            const double det[5] = {sm.dbox[0 * Dcap + d], sm.dbox[1 * Dcap + d], sm.dbox[2 * Dcap + d],
                                   sm.dbox[3 * Dcap + d], sm.dconf[d]};
            ocs_track_update(v, sm.slot_of[t], det, sm.dcls[d], sm.dind[d], pr.delta_t, kc);
        }
    }
    __syncthreads();

    // ------------------------------------------------------------------ phase E
    int nUD = sm.misc[M_NUD], nUT = sm.misc[M_NUT];
    if (pr.use_byte && nS > 0 && nUT > 0) {   // BYTE round on the low-score detections
        for (int r = tid; r < nS; r += OCS_NT) sm.d2t[r] = r;
        for (int t = tid; t < nT; t += OCS_NT) sm.colcnt[t] = 0;
        __syncthreads();
        const bool ran = ocs_second_round(st, pr, s, sm, sm.sbox, Dcap, sm.d2t, nS, sm.um_t, nUT,
                                          FetchPred{sm.tbox, Tcap}, sm.hit_t);
        if (ran) {
            for (int r = tid; r < nS; r += OCS_NT) {
                const int c = sm.hit_t[r];
                if (c >= 0) {
                    const int t = sm.um_t[c];
                    const double det[5] = {sm.sbox[0 * Dcap + r], sm.sbox[1 * Dcap + r], sm.sbox[2 * Dcap + r],
                                           sm.sbox[3 * Dcap + r], sm.sconf[r]};
                    ocs_track_update(v, sm.slot_of[t], det, sm.scls[r], sm.sind[r], pr.delta_t, kc);
                    sm.colcnt[t] = 2;   // taken
                }
            }
            __syncthreads();
            if (tid == 0) {   // np.setdiff1d: sorted ascending, taken tracks removed
                for (int c = 0; c < nUT; c++) { const int t = sm.um_t[c]; if (sm.colcnt[t] != 2) sm.colcnt[t] = 1; }
                int k = 0;
                for (int t = 0; t < nT; t++) if (sm.colcnt[t] == 1) sm.um_t[k++] = t;
                sm.misc[M_NUT] = k;
            }
            __syncthreads();
            nUT = sm.misc[M_NUT];
        }
    }
    if (nUD > 0 && nUT > 0) {                  // OCR: leftover detections vs last observations
        for (int t = tid; t < nT; t += OCS_NT) sm.colcnt[t] = 0;
        for (int d = tid; d < nD; d += OCS_NT) sm.rowcnt[d] = 0;
        __syncthreads();
        const bool ran = ocs_second_round(st, pr, s, sm, sm.dbox, Dcap, sm.um_d, nUD, sm.um_t, nUT,
                                          FetchLast{v.lastobs, sm.slot_of, Tcap}, sm.hit_t);
        if (ran) {
            for (int r = tid; r < nUD; r += OCS_NT) {
                const int c = sm.hit_t[r];
                if (c >= 0) {
                    const int d = sm.um_d[r], t = sm.um_t[c];
                    const double det[5] = {sm.dbox[0 * Dcap + d], sm.dbox[1 * Dcap + d], sm.dbox[2 * Dcap + d],
                                           sm.dbox[3 * Dcap + d], sm.dconf[d]};
                    ocs_track_update(v, sm.slot_of[t], det, sm.dcls[d], sm.dind[d], pr.delta_t, kc);
                    sm.colcnt[t] = 2;
                    sm.rowcnt[d] = 2;
                }
            }
            __syncthreads();
            if (tid == 0) {
                for (int c = 0; c < nUT; c++) { const int t = sm.um_t[c]; if (sm.colcnt[t] != 2) sm.colcnt[t] = 1; }
                for (int r = 0; r < nUD; r++) { const int d = sm.um_d[r]; if (sm.rowcnt[d] != 2) sm.rowcnt[d] = 1; }
                int k = 0;
                for (int t = 0; t < nT; t++) if (sm.colcnt[t] == 1) sm.um_t[k++] = t;
                sm.misc[M_NUT] = k;
                k = 0;
                for (int d = 0; d < nD; d++) if (sm.rowcnt[d] == 1) sm.um_d[k++] = d;
                sm.misc[M_NUD] = k;
            }
            __syncthreads();
            nUT = sm.misc[M_NUT];
            nUD = sm.misc[M_NUD];
        }
    }

    // ------------------------------------------------------------------ phase F
    for (int c = tid; c < nUT; c += OCS_NT) ocs_track_miss(v, sm.slot_of[sm.um_t[c]]);
    // births: ids and slots are handed out in unmatched-detection list order (ocsort.py:392-413)
    if (tid == 0) {
        for (int t = 0; t < Tcap; t++) sm.colcnt[t] = 0;            // slot-in-use map
        for (int p = 0; p < nT; p++) sm.colcnt[sm.slot_of[p]] = 1;
        int next_free = 0, born = 0;
        for (int r = 0; r < nUD; r++) {
            while (next_free < Tcap && sm.colcnt[next_free]) next_free++;
            if (next_free >= Tcap) { atomicOr(&v.hdr[OH_ERR], BM_ERR_TRACK_OVERFLOW); break; }
            sm.colcnt[next_free] = 1;
            sm.hit_t[r] = next_free;                                 // slot of the r-th birth
            v.order[nT + born] = next_free;
            born++;
        }
        sm.misc[M_NREJ] = born;
    }
    __syncthreads();
    const int born = sm.misc[M_NREJ];
    const int id0 = v.hdr[OH_NEXT_ID];
    for (int r = tid; r < born; r += OCS_NT) {
        const int d = sm.um_d[r];
        const double det[5] = {sm.dbox[0 * Dcap + d], sm.dbox[1 * Dcap + d], sm.dbox[2 * Dcap + d],
                               sm.dbox[3 * Dcap + d], sm.dconf[d]};
        ocs_track_init(v, sm.hit_t[r], det, sm.dcls[d], sm.dind[d], id0 + r);
    }
    __syncthreads();
    const int nAll = nT + born;
    // output rows walk the list backwards; dead tracks leave the list (ocsort.py:414-436)
    double* orow = out + (size_t)s * Tcap * 8;
    for (int p = tid; p < nAll; p += OCS_NT) {
        const int slot = v.order[p];
        const int tsu = v.I(OI_TSU, slot);
        const bool emit = tsu < 1 && (v.I(OI_STREAK, slot) >= pr.min_hits || frame <= pr.min_hits);
        sm.kvalid[p] = (emit ? 1 : 0) | (tsu > pr.max_age ? 2 : 0);
        if (p < nT && tsu == 0) atomicAdd(&sm.misc[M_NMATCH], 1);   // matched this frame (bench counter)
    }
    __syncthreads();
    if (tid == 0) {
        int k = 0;
        for (int p = nAll - 1; p >= 0; p--) if (sm.kvalid[p] & 1) sm.um_t[p] = k++; else sm.um_t[p] = -1;
        nout[s] = k;
        int keep = 0;
        for (int p = 0; p < nAll; p++) if (!(sm.kvalid[p] & 2)) sm.slot_of[keep++] = v.order[p];
        sm.misc[M_NT] = keep;
    }
    __syncthreads();
    for (int p = tid; p < nAll; p += OCS_NT) {
        const int k = sm.um_t[p];
        if (k < 0) continue;
        const int slot = v.order[p];
        double b[4];
        double lo[5];
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
        double* o = orow + (size_t)k * 8;
        o[0] = b[0]; o[1] = b[1]; o[2] = b[2]; o[3] = b[3];
        o[4] = (double)(v.I(OI_ID, slot) + 1);
        o[5] = v.confcls[0 * Tcap + slot];
        o[6] = v.confcls[1 * Tcap + slot];
        o[7] = (double)v.I(OI_DETIND, slot);
    }
    __syncthreads();
    const int keep = sm.misc[M_NT];
    for (int p = tid; p < keep; p += OCS_NT) v.order[p] = sm.slot_of[p];
    if (tid == 0) {
        v.hdr[OH_NLIST] = keep;
        v.hdr[OH_NEXT_ID] = id0 + born;
        v.hdr[OH_FRAME] = frame;
        v.hdr[OH_MATCH] += sm.misc[M_NMATCH];
        v.hdr[OH_BIRTHS] += born;
        if (sm.misc[M_LAPBAD]) atomicOr(&v.hdr[OH_ERR], 8);
    }
}

}  // namespace bm

namespace bm {

size_t ocsort_step_smem(const OcsState& st, const OcsParams& pr) {
    return ocs_smem_bytes(st.Tcap, st.Dcap, pr.use_byte, st.smem_cost_elems);
}

cudaError_t ocsort_step_configure(size_t smem) {
    return cudaFuncSetAttribute(ocsort_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

cudaError_t ocsort_step_launch(const OcsState& st, const OcsParams& pr, const float* dets, const int* ndets,
                               double* out, int* nout, cudaStream_t stream) {
    const size_t smem = ocsort_step_smem(st, pr);
    ocsort_step_kernel<<<st.S, OCS_NT, smem, stream>>>(st, pr, dets, ndets, out, nout);
    return cudaGetLastError();
}

}  // namespace bm
