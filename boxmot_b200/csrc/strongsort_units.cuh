// StrongSORT per-pair cost pieces and the per-track state machine as device functions, shared by
// the unit-testable kernels (unit_kernels.cu) and the StrongSORT engine (strongsort_step.cu).
// Reference: boxmot/trackers/strongsort/sort/{linear_assignment,tracker,track,iou_matching}.py.
#pragma once
#include "common.cuh"
#include "kf_xyah.cuh"

namespace bm {

// Per-track attribute table: SS_* columns are the ones the cost functions read.
enum { SS_AGE = 0, SS_HITS, SS_TSU, SS_MOTION, SS_APPEAR, SS_QUALITY, SS_HI_STREAK, SS_LO_STREAK, SS_HAS_VEL, SS_VX,
       SS_VY, SS_CONF, SS_CLS, SS_FIELDS };
constexpr double SS_INFTY_COST = 1e5;        // linear_assignment.py:11
constexpr double SS_CHI2INV95_4 = 9.4877;    // matching.py chi2inv95[4]

// Track.predict / update / mark_missed (track.py:195-317) table: columns 0..12 are the SS_* columns
// (SS_HAS_VEL = len(velocity_history), capped at 10, SS_VX/VY = velocity_history[-1]); then:
enum { ST_STATE = 13, ST_STABILITY, ST_CONFIRMED, ST_MISSED, ST_NPOS, ST_PX, ST_PY, ST_NCONF, ST_CONF0,
       ST_CONF_CAP = 20, ST_DETIND = ST_CONF0 + ST_CONF_CAP, ST_FIELDS = ST_DETIND + 1 };
enum { ST_TENTATIVE = 1, ST_CONFIRMED_STATE = 2, ST_DELETED = 3 };

__device__ __forceinline__ double ss_norm2(double x, double y) { return sqrt(x * x + y * y); }

// gated_metric after metric.distance (tracker.py:186-206): gate_cost_matrix (linear_assignment.py:174-248)
// + _apply_id_preservation_weighting (tracker.py:283-298).  c = appearance distance of the pair, x / P =
// the track's Kalman state (block covariance), a = its table row, z = the detection as xyah.
__device__ __forceinline__ double ss_gated_entry(double c, const double* x, const double* P, const double* a,
                                                 const double* z, double mc_lambda, double id_weight) {
    const double gd = cv8_gating<CV8_XYAH>(x, P, z, false, false);
    if (gd > SS_CHI2INV95_4) c = SS_INFTY_COST;                              // :232
    // _compute_enhanced_motion_cost (:288-307)
    double motion = gd * (2.0 - a[SS_MOTION]);
    if (a[SS_HAS_VEL] != 0.0) {
        const double dx = z[0] - (x[0] + a[SS_VX]), dy = z[1] - (x[1] + a[SS_VY]);
        const double err = sqrt(dx * dx + dy * dy);
        motion = motion * (1.0 + fmin(err / 50.0, 1.0));
    }
    // _compute_adaptive_lambda (:310-329)
    double lam = mc_lambda + (1.0 - mc_lambda) * fmin(a[SS_AGE] / 10.0, 1.0) * 0.1;
    lam = lam * (0.8 + 0.4 * a[SS_MOTION]);
    if (a[SS_APPEAR] < 0.5) lam = fmin(lam * 1.2, 0.99);
    lam = fmin(fmax(lam, 0.1), 0.99);
    c = lam * c + (1.0 - lam) * motion;                                      // :241-243
    // _apply_track_specific_adjustments (:332-352)
    if (a[SS_QUALITY] > 0.8) c = c * 0.95;
    if (a[SS_HITS] > 10.0 && a[SS_TSU] == 0.0) c = c * 0.98;
    if (a[SS_HI_STREAK] > 3.0) c = c * 0.97;
    if (a[SS_LO_STREAK] > 2.0) c = c * 1.05;
    if (id_weight > 0.0) c = c * (1.0 - id_weight * fmin(a[SS_HITS] / 10.0, 1.0));   // tracker.py:294-296
    return c;
}

// The same entry split into its per-track part and its per-pair part for a kernel that walks one track's row:
// every hoisted value is a subexpression of ss_gated_entry / ss_enhance_entry that depends on the track alone,
// evaluated by the same operations (the library is compiled without FMA contraction), so the pair function
// returns ss_enhance_entry(ss_gated_entry(...)) bit for bit.
struct SsTrackTerms {
    double x[4], inv[4];            // mean and 1 / sqrt(S_ii) of the projected state
    double mfac, px, py, lam, oml, idf, quality, conf, cls;
    bool has_vel, q95, h98, s97, s105, use_id;
};
__device__ __forceinline__ SsTrackTerms ss_track_terms(const double* x, const double* P, const double* a,
                                                       double mc_lambda, double id_weight) {
    SsTrackTerms t;
    double S[4];
    cv8_project<CV8_XYAH>(x, P, 0.0, S);
#pragma unroll
    for (int i = 0; i < 4; i++) { t.x[i] = x[i]; t.inv[i] = 1.0 / sqrt(S[i]); }
    t.mfac = 2.0 - a[SS_MOTION];
    t.has_vel = a[SS_HAS_VEL] != 0.0;
    t.px = x[0] + a[SS_VX]; t.py = x[1] + a[SS_VY];
    double lam = mc_lambda + (1.0 - mc_lambda) * fmin(a[SS_AGE] / 10.0, 1.0) * 0.1;
    lam = lam * (0.8 + 0.4 * a[SS_MOTION]);
    if (a[SS_APPEAR] < 0.5) lam = fmin(lam * 1.2, 0.99);
    lam = fmin(fmax(lam, 0.1), 0.99);
    t.lam = lam; t.oml = 1.0 - lam;
    t.q95 = a[SS_QUALITY] > 0.8;
    t.h98 = a[SS_HITS] > 10.0 && a[SS_TSU] == 0.0;
    t.s97 = a[SS_HI_STREAK] > 3.0;
    t.s105 = a[SS_LO_STREAK] > 2.0;
    t.use_id = id_weight > 0.0;
    t.idf = 1.0 - id_weight * fmin(a[SS_HITS] / 10.0, 1.0);
    t.quality = a[SS_QUALITY]; t.conf = a[SS_CONF]; t.cls = a[SS_CLS];
    return t;
}
// c = appearance distance of the pair, z = the detection as xyah
__device__ __forceinline__ double ss_pair_cost(const SsTrackTerms& t, double c, const double* z, double dq, double dconf,
                                               double dcls) {
    double gd = 0.0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double d = z[i] - t.x[i];
        const double q = d * t.inv[i];
        gd = i == 0 ? q * q : gd + q * q;
    }
    if (gd > SS_CHI2INV95_4) c = SS_INFTY_COST;
    double motion = gd * t.mfac;
    if (t.has_vel) {
        const double dx = z[0] - t.px, dy = z[1] - t.py;
        const double err = sqrt(dx * dx + dy * dy);
        motion = motion * (1.0 + fmin(err / 50.0, 1.0));
    }
    c = t.lam * c + t.oml * motion;
    if (t.q95) c = c * 0.95;
    if (t.h98) c = c * 0.98;
    if (t.s97) c = c * 0.97;
    if (t.s105) c = c * 1.05;
    if (t.use_id) c = c * t.idf;
    const double q = 1.0 - ((t.quality + dq) / 2.0 - 0.5) * 0.2;
    c = c * fmin(fmax(q, 0.8), 1.2);
    if (t.cls == dcls) c = c * 0.9;
    if (t.conf > 0.7 && dconf > 0.7) c = c * 0.9;
    else if (t.conf < 0.3 || dconf < 0.3) c = c * 1.1;
    return c;
}

// _enhance_cost_matrix (linear_assignment.py:251-273) for one entry; detection quality_score, conf, cls
__device__ __forceinline__ double ss_enhance_entry(double c, const double* a, double dq, double dconf, double dcls) {
    const double q = 1.0 - ((a[SS_QUALITY] + dq) / 2.0 - 0.5) * 0.2;         // :355-364
    c = c * fmin(fmax(q, 0.8), 1.2);
    if (a[SS_CLS] == dcls) c = c * 0.9;                                      // :265-267
    if (a[SS_CONF] > 0.7 && dconf > 0.7) c = c * 0.9;                        // :367-379
    else if (a[SS_CONF] < 0.3 || dconf < 0.3) c = c * 1.1;
    return c;
}

// iou_cost (iou_matching.py:49-87) for one entry: 1 - IoU of Track.to_tlwh() with the detection's tlwh
__device__ __forceinline__ double ss_iou_entry(const double* m, const double* b) {
    const double h = m[3], w = m[2] * h;                 // track.py:154-157
    const double x1 = m[0] - w / 2.0, y1 = m[1] - h / 2.0;
    const double tlx = fmax(x1, b[0]), tly = fmax(y1, b[1]);
    const double brx = fmin(x1 + w, b[0] + b[2]), bry = fmin(y1 + h, b[1] + b[3]);
    const double inter = fmax(0.0, brx - tlx) * fmax(0.0, bry - tly);
    return 1.0 - inter / (w * h + b[2] * b[3] - inter);
}

// Track.mark_missed (:303-317)
__device__ __forceinline__ void ss_track_mark_missed(double* a, int max_age) {
    a[ST_MISSED] += 1.0;
    int thr = max_age;
    if (a[SS_QUALITY] > 0.8) thr = (int)((double)max_age * 1.5);
    else if (a[SS_QUALITY] < 0.3) thr = (int)((double)max_age * 0.5);
    if (a[ST_STATE] == (double)ST_TENTATIVE) a[ST_STATE] = ST_DELETED;
    else if (a[SS_TSU] > (double)thr) a[ST_STATE] = ST_DELETED;
}

// _update_motion_consistency (:379-399) with velocity_history[-1] = (a[SS_VX], a[SS_VY])
__device__ __forceinline__ void ss_motion_consistency(double* a, double px, double py, double cx, double cy) {
    if (a[SS_HAS_VEL] >= 2.0) {
        const double ax = cx - px, ay = cy - py;
        const double mx = (px + a[SS_VX]) - px, my = (py + a[SS_VY]) - py;
        const double nm = ss_norm2(mx, my);
        double cons;
        if (nm > 0.0) cons = fmax(0.0, 1.0 - ss_norm2(ax - mx, ay - my) / fmax(nm * 0.5, 10.0));
        else cons = ss_norm2(ax, ay) < 5.0 ? 1.0 : 0.5;
        a[SS_MOTION] = 0.8 * a[SS_MOTION] + 0.2 * cons;
    }
}

// Track.camera_update (:173-189): m = mean (first four entries rewritten), w = row-major 2x3 warp
__device__ __forceinline__ void ss_track_camera(double* m, double* a, const double* w) {
    const double bw0 = m[2] * m[3];
    const double x1 = m[0] - bw0 / 2.0, y1 = m[1] - m[3] / 2.0, x2 = x1 + bw0, y2 = y1 + m[3];   // to_tlbr (:159-171)
    const double x1w = w[0] * x1 + w[1] * y1 + w[2], y1w = w[3] * x1 + w[4] * y1 + w[5];
    const double x2w = w[0] * x2 + w[1] * y2 + w[2], y2w = w[3] * x2 + w[4] * y2 + w[5];
    const double bw = x2w - x1w, bh = y2w - y1w;
    const double cx = x1w + bw / 2.0, cy = y1w + bh / 2.0;
    const double px = m[0], py = m[1];
    m[0] = cx; m[1] = cy; m[2] = bw / bh; m[3] = bh;
    ss_motion_consistency(a, px, py, cx, cy);
}

// Track.predict (:195-219) on the state in registers
__device__ __forceinline__ void ss_track_predict(double* x, double* P, double* a) {
    cv8_predict<CV8_XYAH>(x, P);
    a[SS_AGE] += 1.0;
    a[SS_TSU] += 1.0;
    a[SS_VX] = x[4]; a[SS_VY] = x[5];
    a[SS_HAS_VEL] = fmin(a[SS_HAS_VEL] + 1.0, 10.0);
    const double px = a[ST_PX], py = a[ST_PY], had = a[ST_NPOS];
    a[ST_PX] = x[0]; a[ST_PY] = x[1];
    a[ST_NPOS] = fmin(had + 1.0, 10.0);
    if (had >= 1.0) ss_motion_consistency(a, px, py, x[0], x[1]);
}

// Track.update (:221-301) without its feature step (the appearance consistency of that step must
// already be in a[SS_APPEAR]); d = x, y, a, h, conf, cls, det_ind
__device__ __forceinline__ void ss_track_update(double* x, double* P, double* a, const double* d, int n_init) {
    const double conf = d[4];
    a[SS_CONF] = conf; a[SS_CLS] = d[5]; a[ST_DETIND] = d[6];
    double z[4] = {d[0], d[1], d[2], d[3]};
    cv8_update<CV8_XYAH>(x, P, z, conf);
    int nh = (int)a[ST_NCONF];
    if (nh >= ST_CONF_CAP) {
        for (int k = 1; k < ST_CONF_CAP; k++) a[ST_CONF0 + k - 1] = a[ST_CONF0 + k];
        nh = ST_CONF_CAP - 1;
    }
    a[ST_CONF0 + nh] = conf;
    nh++;
    a[ST_NCONF] = nh;
    if (conf > 0.7) { a[SS_HI_STREAK] += 1.0; a[SS_LO_STREAK] = 0.0; }
    else if (conf < 0.3) { a[SS_LO_STREAK] += 1.0; a[SS_HI_STREAK] = 0.0; }
    else { a[SS_LO_STREAK] = 0.0; a[SS_HI_STREAK] = 0.0; }
    a[SS_HITS] += 1.0;
    a[ST_CONFIRMED] += 1.0;
    a[SS_TSU] = 0.0;
    double sum = 0.0;
    for (int k = 0; k < nh; k++) sum += a[ST_CONF0 + k];
    const double avg = sum / nh;
    double var = 0.0;
    for (int k = 0; k < nh; k++) { const double e = a[ST_CONF0 + k] - avg; var += e * e; }
    const double sd = sqrt(var / nh);
    double cq = conf;                                            // _update_quality_score (:337-359)
    if (nh > 1) cq = 0.7 * cq + 0.3 * avg * (1.0 - sd);
    const double q = cq + fmin(a[SS_HITS] / 20.0, 0.2) + fmax(0.0, (a[SS_APPEAR] - 0.5) * 0.2) +
                     fmax(0.0, (a[SS_MOTION] - 0.5) * 0.1);
    a[SS_QUALITY] = fmin(fmax(q, 0.0), 1.0);
    const double cs = nh > 3 ? 1.0 - fmin(sd, 1.0) : 0.5;        // _update_stability_score (:361-377)
    const double hit_rate = a[ST_CONFIRMED] / fmax(a[SS_AGE], 1.0);
    const double cons = 0.4 * a[SS_APPEAR] + 0.3 * a[SS_MOTION] + 0.3 * cs;
    a[ST_STABILITY] = fmin(fmax(0.5 * hit_rate + 0.5 * cons, 0.0), 1.0);
    if (a[ST_STATE] == (double)ST_TENTATIVE &&
        (a[SS_HITS] >= (double)n_init || (a[SS_HITS] >= 1.0 && a[SS_QUALITY] > 0.8)))
        a[ST_STATE] = ST_CONFIRMED_STATE;
}

}  // namespace bm
