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

// Noise standard deviations of the two constant-velocity box filters:
//   KIND 0  KalmanFilterXYAH  (xyah_kf.py:17-62):  everything scales with h, a-terms constant
//   KIND 1  KalmanFilterXYWH  (xywh_kf.py:17-51):  x / w terms scale with w, y / h terms with h
enum { CV8_XYAH = 0, CV8_XYWH = 1 };
template <int KIND>
__device__ __forceinline__ void cv8_std(const double* m, double pos_mult, double vel_mult, double a_pos, double a_vel,
                                        double* sp, double* sv) {
    if (KIND == CV8_XYAH) {
        const double p = pos_mult * XA_W_POS * m[3], v = vel_mult * XA_W_VEL * m[3];
        sp[0] = p; sp[1] = p; sp[2] = a_pos; sp[3] = p;
        sv[0] = v; sv[1] = v; sv[2] = a_vel; sv[3] = v;
    } else {
        const double pw = pos_mult * XA_W_POS * m[2], ph = pos_mult * XA_W_POS * m[3];
        const double vw = vel_mult * XA_W_VEL * m[2], vh = vel_mult * XA_W_VEL * m[3];
        sp[0] = pw; sp[1] = ph; sp[2] = pw; sp[3] = ph;
        sv[0] = vw; sv[1] = vh; sv[2] = vw; sv[3] = vh;
    }
}

// initiate (base_kalman_filter.py:43-53); z = (xc, yc, a|w, h)
template <int KIND>
__device__ __forceinline__ void cv8_initiate(const double* z, double* x, double* P) {
    double sp[4], sv[4];
    cv8_std<KIND>(z, 2.0, 10.0, 1e-2, 1e-5, sp, sv);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        x[i] = z[i]; x[4 + i] = 0.0;
        P[i] = sp[i] * sp[i]; P[4 + i] = 0.0; P[8 + i] = sv[i] * sv[i];
    }
}

// predict / multi_predict (base_kalman_filter.py:62-77, :111-127)
template <int KIND>
__device__ __forceinline__ void cv8_predict(double* x, double* P) {
    double sp[4], sv[4];
    cv8_std<KIND>(x, 1.0, 1.0, 1e-2, 1e-5, sp, sv);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        x[i] = x[i] + x[4 + i];
        const double pp = P[i], pv = P[4 + i], vv = P[8 + i];
        // (F P)[i][i] = pp + pv, (F P)[i][i+4] = pv + vv; then (F P) F^T
        P[i] = ((pp + pv) + (pv + vv)) + sp[i] * sp[i];
        P[4 + i] = pv + vv;
        P[8 + i] = vv + sv[i] * sv[i];
    }
}

// project (base_kalman_filter.py:86-109): S = diag(P_pos) + diag(((1 - confidence) std)^2)
template <int KIND>
__device__ __forceinline__ void cv8_project(const double* x, const double* P, double confidence, double* S) {
    double sp[4], sv[4];
    cv8_std<KIND>(x, 1.0, 1.0, 1e-1, 0.0, sp, sv);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double sd = (1.0 - confidence) * sp[i];
        S[i] = P[i] + sd * sd;
    }
}

// update (base_kalman_filter.py:129-155)
template <int KIND>
__device__ __forceinline__ void cv8_update(double* x, double* P, const double* z, double confidence) {
    double S[4];
    cv8_project<KIND>(x, P, confidence, S);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double pp = P[i], pv = P[4 + i], vv = P[8 + i];
        const double l = sqrt(S[i]);                 // cho_factor of a diagonal matrix
        const double kp = (pp / l) / l, kv = (pv / l) / l;   // cho_solve: forward, then back substitution
        const double innov = z[i] - x[i];
        x[i] = x[i] + innov * kp;
        x[4 + i] = x[4 + i] + innov * kv;
        // covariance - K (S K^T)   (multi_dot picks A(BC) for these shapes)
        P[i] = pp - kp * (S[i] * kp);
        P[4 + i] = pv - kp * (S[i] * kv);
        P[8 + i] = vv - kv * (S[i] * kv);
    }
}

// gating_distance (base_kalman_filter.py:166-194) of one measurement: squared Mahalanobis
// distance (cholesky + triangular solve of a diagonal S) or the "gaussian" squared distance
template <int KIND>
__device__ __forceinline__ double cv8_gating(const double* x, const double* P, const double* z, bool only_position,
                                             bool gaussian) {
    double S[4];
    cv8_project<KIND>(x, P, 0.0, S);
    const int n = only_position ? 2 : 4;
    double acc = 0.0;
    for (int i = 0; i < n; i++) {
        const double d = z[i] - x[i];
        const double q = gaussian ? d : d / sqrt(S[i]);
        acc = i == 0 ? q * q : acc + q * q;
    }
    return acc;
}

__device__ __forceinline__ void xyah_initiate(const double* z, double* x, double* P) { cv8_initiate<CV8_XYAH>(z, x, P); }
__device__ __forceinline__ void xyah_predict(double* x, double* P) { cv8_predict<CV8_XYAH>(x, P); }
__device__ __forceinline__ void xyah_update(double* x, double* P, const double* z) { cv8_update<CV8_XYAH>(x, P, z, 0.0); }

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
