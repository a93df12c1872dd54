// KalmanFilterXYAH / KalmanFilterXYWH (boxmot/motion/kalman_filters/aabb/xyah_kf.py, xywh_kf.py over
// base_kalman_filter.py:24-194) in block form, BIT-IDENTICAL to the reference's numpy / scipy arithmetic.
//
// State (x, y, a|w, h, vx, vy, va|vw, vh); F = I + shift, H = [I4 0], P0 / Q / R diagonal, so the 8x8
// covariance only ever couples (pos_i, vel_i).  It is stored as 16 numbers
//     P[i] = P[i][i],  P[4+i] = P[i][i+4],  P[8+i] = P[i+4][i],  P[12+i] = P[i+4][i+4]
// -- BOTH off-diagonal entries, because the reference's update `P - K (S K^T)` is not bit-symmetric.
// Every formula below is the reference's dense expression restricted to the entries that can be
// non-zero, in the reference's own operation order: each matrix-product entry has at most one
// non-zero term (exact whatever FMA the BLAS kernel uses) or two terms of coefficient 1; scipy's
// cho_solve / solve_triangular on the (diagonal) innovation covariance multiply by the RECIPROCAL of
// the Cholesky diagonal (OpenBLAS' trsm kernels invert the diagonal once), which is reproduced.
// Checked bit for bit against the scipy-based oracle (tests/test_gpu_units.py).
#pragma once
#include "common.cuh"
#include "iou.cuh"

namespace bm {

constexpr int XA_N = 24;                       // 8 mean + 16 covariance
constexpr int XA_P = 16;
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
        P[i] = sp[i] * sp[i]; P[4 + i] = 0.0; P[8 + i] = 0.0; P[12 + i] = sv[i] * sv[i];
    }
}

// predict (base_kalman_filter.py:62-77): multi_dot((F, P, F^T)) evaluates F (P F^T) for square factors;
// multi_predict (:111-127, MULTI = true): np.dot(np.dot(F, P), F^T).  The two differ in the order of
// the four-term sum of P[i][i] once the two off-diagonal entries differ in their last bits.
template <int KIND, bool MULTI = false>
__device__ __forceinline__ void cv8_predict(double* x, double* P) {
    double sp[4], sv[4];
    cv8_std<KIND>(x, 1.0, 1.0, 1e-2, 1e-5, sp, sv);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        x[i] = x[i] + x[4 + i];
        const double a = P[i], b = P[4 + i], c = P[8 + i], d = P[12 + i];
        P[i] = (MULTI ? (a + c) + (b + d) : (a + b) + (c + d)) + sp[i] * sp[i];
        P[4 + i] = b + d;
        P[8 + i] = c + d;
        P[12 + i] = d + sv[i] * sv[i];
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

// update (base_kalman_filter.py:129-155): K = cho_solve(cho_factor(S), (P H^T)^T)^T uses column i of P
// (P[i][i], P[i+4][i]) and the reciprocal of sqrt(S_ii) twice; P <- P - K (S K^T)
template <int KIND>
__device__ __forceinline__ void cv8_update(double* x, double* P, const double* z, double confidence) {
    double S[4];
    cv8_project<KIND>(x, P, confidence, S);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double a = P[i], b = P[4 + i], c = P[8 + i], d = P[12 + i];
        const double inv = 1.0 / sqrt(S[i]);         // cho_factor of a diagonal matrix; trsm inverts the diagonal
        const double kp = (a * inv) * inv, kv = (c * inv) * inv;
        const double innov = z[i] - x[i];
        x[i] = x[i] + innov * kp;
        x[4 + i] = x[4 + i] + innov * kv;
        P[i] = a - kp * (S[i] * kp);
        P[4 + i] = b - kp * (S[i] * kv);
        P[8 + i] = c - kv * (S[i] * kp);
        P[12 + i] = d - kv * (S[i] * kv);
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
        const double q = gaussian ? d : d * (1.0 / sqrt(S[i]));     // solve_triangular: reciprocal of the diagonal
        acc = i == 0 ? q * q : acc + q * q;
    }
    return acc;
}

__device__ __forceinline__ void xyah_initiate(const double* z, double* x, double* P) { cv8_initiate<CV8_XYAH>(z, x, P); }
__device__ __forceinline__ void xyah_predict(double* x, double* P) { cv8_predict<CV8_XYAH, true>(x, P); }
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
        out64[(4 + i) * 8 + i] = P[8 + i];
        out64[(4 + i) * 8 + 4 + i] = P[12 + i];
    }
}

}  // namespace bm
