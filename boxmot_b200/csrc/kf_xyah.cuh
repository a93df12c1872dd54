// KalmanFilterXYAH (boxmot/motion/kalman_filters/aabb/xyah_kf.py over
// base_kalman_filter.py:24-155) in block form.
//
// State (x, y, a, h, vx, vy, va, vh); F = I + shift, H = [I4 0], P0 / Q / R diagonal
// (xyah_kf.py:17-62), so the 8x8 covariance only ever couples (pos_i, vel_i): it is stored as
// 12 numbers  P[i] = var(pos_i), P[4+i] = cov(pos_i, vel_i), P[8+i] = var(vel_i)  and every
// formula below is the reference's dense expression restricted to the entries that can be
// non-zero (the dropped terms are exact zeros).  Differences to the reference are at the
// rounding level only: the dense update is not bit-symmetric (K (S K^T)) and scipy's
// cho_solve multiplies by reciprocals inside OpenBLAS; tests bound the state at 1e-9.
#pragma once
#include "common.cuh"
#include "iou.cuh"

namespace bm {

constexpr int XA_N = 20;                       // 8 mean + 12 covariance
constexpr double XA_W_POS = 1.0 / 20, XA_W_VEL = 1.0 / 160;   // base_kalman_filter.py:40-41

// initiate (base_kalman_filter.py:43-53, xyah_kf.py:17-29); z = (xc, yc, a, h)
__device__ __forceinline__ void xyah_initiate(const double* z, double* x, double* P) {
    const double h = z[3];
    const double sp = 2 * XA_W_POS * h, sv = 10 * XA_W_VEL * h;
#pragma unroll
    for (int i = 0; i < 4; i++) { x[i] = z[i]; x[4 + i] = 0.0; P[4 + i] = 0.0; }
    P[0] = sp * sp; P[1] = sp * sp; P[2] = 1e-2 * 1e-2; P[3] = sp * sp;
    P[8] = sv * sv; P[9] = sv * sv; P[10] = 1e-5 * 1e-5; P[11] = sv * sv;
}

// predict / multi_predict (base_kalman_filter.py:62-77, :111-127; noise xyah_kf.py:31-48)
__device__ __forceinline__ void xyah_predict(double* x, double* P) {
    const double h = x[3];
    const double sp = XA_W_POS * h, sv = XA_W_VEL * h;
    const double qp[4] = {sp * sp, sp * sp, 1e-2 * 1e-2, sp * sp};
    const double qv[4] = {sv * sv, sv * sv, 1e-5 * 1e-5, sv * sv};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        x[i] = x[i] + x[4 + i];
        const double pp = P[i], pv = P[4 + i], vv = P[8 + i];
        // (F P)[i][i] = pp + pv, (F P)[i][i+4] = pv + vv; then (F P) F^T
        P[i] = ((pp + pv) + (pv + vv)) + qp[i];
        P[4 + i] = pv + vv;
        P[8 + i] = vv + qv[i];
    }
}

// update (base_kalman_filter.py:129-155) with project (:86-109, confidence = 0) and the
// measurement noise of xyah_kf.py:50-62
__device__ __forceinline__ void xyah_update(double* x, double* P, const double* z) {
    const double sr = XA_W_POS * x[3];
    const double r[4] = {sr * sr, sr * sr, 1e-1 * 1e-1, sr * sr};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double pp = P[i], pv = P[4 + i], vv = P[8 + i];
        const double S = pp + r[i];
        const double l = sqrt(S);                    // cho_factor of a diagonal matrix
        const double kp = (pp / l) / l, kv = (pv / l) / l;   // cho_solve: forward, then back substitution
        const double innov = z[i] - x[i];
        x[i] = x[i] + innov * kp;
        x[4 + i] = x[4 + i] + innov * kv;
        // covariance - K (S K^T)   (multi_dot picks A(BC) for these shapes)
        P[i] = pp - kp * (S * kp);
        P[4 + i] = pv - kp * (S * kv);
        P[8 + i] = vv - kv * (S * kv);
    }
}

// STrack.xyxy (bytetrack.py:105-116) from the mean
__device__ __forceinline__ Box xyah_to_xyxy(const double* x) {
    const double w = x[2] * x[3];
    return Box{x[0] - w / 2, x[1] - x[3] / 2, x[0] + w / 2, x[1] + x[3] / 2};
}

// dense 8x8 covariance for tests / peek
__host__ __device__ inline void xyah_expand(const double* P, double* out64) {
    for (int i = 0; i < 64; i++) out64[i] = 0.0;
    for (int i = 0; i < 4; i++) {
        out64[i * 8 + i] = P[i];
        out64[i * 8 + 4 + i] = P[4 + i];
        out64[(4 + i) * 8 + i] = P[4 + i];
        out64[(4 + i) * 8 + 4 + i] = P[8 + i];
    }
}

}  // namespace bm
