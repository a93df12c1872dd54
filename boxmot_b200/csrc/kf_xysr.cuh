// 7-state (x, y, s, r, vx, vy, vs) Kalman filter of the OC-SORT family, fp64, one
// thread per track with the state in registers/local arrays.
//
// Replaces boxmot/motion/kalman_filters/aabb/xysr_kf.py:
//   predict :137-175, update :211-291 (explicit inverse of S, Joseph form),
//   freeze/unfreeze = OC-SORT observation-centric re-update (ORU) :177-209,
// with the constants boxmot/trackers/ocsort/ocsort.py:83-113 installs
// (F = I + shift, H = [I4 0], R = diag(1,1,10,10), Q = diag(1,1,1,1,qxy,qxy,qs),
// P0 = diag(10,10,10,10,1e4,1e4,1e4)).  F and H are 0/1 matrices, so their products
// are written as the adds they reduce to (bit-identical to the BLAS result).
#pragma once
#include "common.cuh"

namespace bm {

struct XysrConst {
    double q_xy, q_s;   // Q[4,4] = Q[5,5] = q_xy ; Q[6,6] = q_s ; Q[0..3] = 1
};

__device__ __forceinline__ void xysr_init(double* x, double* P, const double* z) {
#pragma unroll
    for (int i = 0; i < 49; i++) P[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 7; i++) {
        x[i] = i < 4 ? z[i] : 0.0;
        P[i * 7 + i] = i < 4 ? 10.0 : 10000.0;
    }
}

// x <- F x ; P <- F P F' + Q
__device__ __forceinline__ void xysr_predict(double* x, double* P, const XysrConst& kc) {
    x[0] = x[0] + x[4];
    x[1] = x[1] + x[5];
    x[2] = x[2] + x[6];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 7; j++) P[i * 7 + j] = P[i * 7 + j] + P[(i + 4) * 7 + j];
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) P[i * 7 + j] = P[i * 7 + j] + P[i * 7 + j + 4];
#pragma unroll
    for (int i = 0; i < 4; i++) P[i * 7 + i] = P[i * 7 + i] + 1.0;
    P[4 * 7 + 4] = P[4 * 7 + 4] + kc.q_xy;
    P[5 * 7 + 5] = P[5 * 7 + 5] + kc.q_xy;
    P[6 * 7 + 6] = P[6 * 7 + 6] + kc.q_s;
}

// 4x4 inverse by LU with partial pivoting and substitution against the identity
// (the gesv route numpy.linalg.inv takes).
__device__ __forceinline__ void inv4(const double* S, double* SI) {
    double a[16];
    int piv[4];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = S[i];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int p = k;
        double best = fabs(a[k * 4 + k]);
#pragma unroll
        for (int i = k + 1; i < 4; i++) {
            double t = fabs(a[i * 4 + k]);
            if (t > best) { best = t; p = i; }
        }
        piv[k] = p;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            // swap rows k and p (p is dynamic: select without dynamic indexing)
            double rk = a[k * 4 + j], rp = rk;
#pragma unroll
            for (int i = k + 1; i < 4; i++) if (i == p) rp = a[i * 4 + j];
#pragma unroll
            for (int i = k + 1; i < 4; i++) if (i == p) a[i * 4 + j] = rk;
            a[k * 4 + j] = rp;
        }
        const double r = 1.0 / a[k * 4 + k];
#pragma unroll
        for (int i = k + 1; i < 4; i++) a[i * 4 + k] = a[i * 4 + k] * r;
#pragma unroll
        for (int i = k + 1; i < 4; i++)
#pragma unroll
            for (int j = k + 1; j < 4; j++) a[i * 4 + j] = a[i * 4 + j] - a[i * 4 + k] * a[k * 4 + j];
    }
#pragma unroll
    for (int c = 0; c < 4; c++) {
        double b[4];
#pragma unroll
        for (int i = 0; i < 4; i++) b[i] = i == c ? 1.0 : 0.0;
#pragma unroll
        for (int k = 0; k < 4; k++) {   // apply the row interchanges
            double bk = b[k], bp = bk;
#pragma unroll
            for (int i = k + 1; i < 4; i++) if (i == piv[k]) bp = b[i];
#pragma unroll
            for (int i = k + 1; i < 4; i++) if (i == piv[k]) b[i] = bk;
            b[k] = bp;
        }
#pragma unroll
        for (int k = 0; k < 4; k++)      // L y = b (unit lower)
#pragma unroll
            for (int i = k + 1; i < 4; i++) b[i] = b[i] - a[i * 4 + k] * b[k];
#pragma unroll
        for (int k = 3; k >= 0; k--) {   // U x = y
            b[k] = b[k] / a[k * 4 + k];
#pragma unroll
            for (int i = 0; i < k; i++) b[i] = b[i] - a[i * 4 + k] * b[k];
        }
#pragma unroll
        for (int i = 0; i < 4; i++) SI[i * 4 + c] = b[i];
    }
}

// Measurement update with z = (xc, yc, s, r); R = diag(1, 1, 10, 10).
__device__ __forceinline__ void xysr_update(double* x, double* P, const double* z) {
    const double R[4] = {1.0, 1.0, 10.0, 10.0};
    double y[4], S[16], SI[16], K[28];
#pragma unroll
    for (int i = 0; i < 4; i++) y[i] = z[i] - x[i];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) S[i * 4 + j] = P[i * 7 + j] + (i == j ? R[i] : 0.0);
    inv4(S, SI);
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < 4; k++) acc = acc + P[i * 7 + k] * SI[k * 4 + j];
            K[i * 4 + j] = acc;
        }
#pragma unroll
    for (int i = 0; i < 7; i++) {
        double acc = 0.0;
#pragma unroll
        for (int k = 0; k < 4; k++) acc = acc + K[i * 4 + k] * y[k];
        x[i] = x[i] + acc;
    }
    // A = I - K H ; AP = A P ; P <- AP A' + (K R) K'
    double AP[49];
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 7; j++) {
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < 7; k++) {
                if (k < 4) {
                    const double a = (i == k ? 1.0 : 0.0) - K[i * 4 + k];
                    acc = acc + a * P[k * 7 + j];
                } else if (k == i) {
                    acc = acc + P[k * 7 + j];
                }
            }
            AP[i * 7 + j] = acc;
        }
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 7; j++) {
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < 7; k++) {
                if (k < 4) {
                    const double a = (j == k ? 1.0 : 0.0) - K[j * 4 + k];
                    acc = acc + AP[i * 7 + k] * a;
                } else if (k == j) {
                    acc = acc + AP[i * 7 + k];
                }
            }
            double krk = 0.0;
#pragma unroll
            for (int k = 0; k < 4; k++) krk = krk + (K[i * 4 + k] * R[k]) * K[j * 4 + k];
            P[i * 7 + j] = acc + krk;
        }
}

// [x1,y1,x2,y2] -> (xc, yc, s, r) with r = w / (h + 1e-6)   (repair R0)
__device__ __forceinline__ void xyxy_to_z(const double* b, double* z) {
    const double w = b[2] - b[0], h = b[3] - b[1];
    z[0] = b[0] + w / 2.0;
    z[1] = b[1] + h / 2.0;
    z[2] = w * h;
    z[3] = w / (h + 1e-6);
}

// boxmot/trackers/ocsort/ocsort.py:31-45 convert_x_to_bbox
__device__ __forceinline__ void x_to_xyxy(const double* x, double* b) {
    const double w = sqrt(x[2] * x[3]);
    const double h = x[2] / w;
    b[0] = x[0] - w / 2.0;
    b[1] = x[1] - h / 2.0;
    b[2] = x[0] + w / 2.0;
    b[3] = x[1] + h / 2.0;
}

// ORU (xysr_kf.py:183-209): starting from the state frozen at the first miss, replay
// update(+predict) along the straight line (in x, y, w, h) from the last measurement
// z1 to the new one z2 over `gap` = misses + 1 steps.
__device__ inline void xysr_oru_replay(double* x, double* P, const double* z1, const double* z2,
                                       int gap, const XysrConst& kc) {
    const double x1 = z1[0], y1 = z1[1];
    const double w1 = sqrt(z1[2] * z1[3]), h1 = sqrt(z1[2] / z1[3]);
    const double x2 = z2[0], y2 = z2[1];
    const double w2 = sqrt(z2[2] * z2[3]), h2 = sqrt(z2[2] / z2[3]);
    const double g = (double)gap;
    const double dx = (x2 - x1) / g, dy = (y2 - y1) / g;
    const double dw = (w2 - w1) / g, dh = (h2 - h1) / g;
    for (int i = 0; i < gap; i++) {
        const double f = (double)(i + 1);
        const double xx = x1 + f * dx, yy = y1 + f * dy;
        const double ww = w1 + f * dw, hh = h1 + f * dh;
        double zi[4] = {xx, yy, ww * hh, ww / hh};
        xysr_update(x, P, zi);
        if (i != gap - 1) xysr_predict(x, P, kc);
    }
}

// ---------------------------------------------------------------------------------------
// Block form used by the engine.  With F = I + shift, H = [I4 0] and DIAGONAL Q, R, P0 the
// covariance only ever couples (x,vx), (y,vy), (s,vs); r stands alone.  Every entry outside
// those blocks is an exact zero in the reference's own arithmetic (a product with an exact
// zero factor), and adding an exact zero never changes a sum, so evaluating only the
// non-zero terms IN THE SAME ORDER is bit-identical to the dense 7x7 code above (checked
// on the GPU by tests/test_gpu_units.py::test_xysr_block_form_is_bit_identical).  State per
// track: x[7] and Pb[13] = {P[b][b], P[b][b+4], P[b+4][b], P[b+4][b+4]} for b = 0,1,2, then
// P[3][3]: 20 doubles instead of 56, ~60 flops per update instead of ~3000.
constexpr int XB_N = 20;   // doubles of Kalman state per track (x[7] + Pb[13])

__device__ __forceinline__ void xysrb_init(double* x, double* Pb, const double* z) {
#pragma unroll
    for (int i = 0; i < 7; i++) x[i] = i < 4 ? z[i] : 0.0;
#pragma unroll
    for (int b = 0; b < 3; b++) { Pb[4 * b] = 10.0; Pb[4 * b + 1] = 0.0; Pb[4 * b + 2] = 0.0; Pb[4 * b + 3] = 10000.0; }
    Pb[12] = 10.0;
}

__device__ __forceinline__ void xysrb_predict(double* x, double* Pb, const XysrConst& kc) {
#pragma unroll
    for (int b = 0; b < 3; b++) {
        x[b] = x[b] + x[b + 4];
        const double a = Pb[4 * b], bb = Pb[4 * b + 1], c = Pb[4 * b + 2], d = Pb[4 * b + 3];
        const double a1 = a + c, b1 = bb + d;
        Pb[4 * b] = (a1 + b1) + 1.0;
        Pb[4 * b + 1] = b1;
        Pb[4 * b + 2] = c + d;
        Pb[4 * b + 3] = d + (b < 2 ? kc.q_xy : kc.q_s);
    }
    Pb[12] = Pb[12] + 1.0;
}

__device__ __forceinline__ void xysrb_update(double* x, double* Pb, const double* z) {
#pragma unroll
    for (int b = 0; b < 3; b++) {
        const double r = b < 2 ? 1.0 : 10.0;
        const double a = Pb[4 * b], bb = Pb[4 * b + 1], c = Pb[4 * b + 2], d = Pb[4 * b + 3];
        const double y = z[b] - x[b];
        const double si = 1.0 / (a + r);
        const double Kp = a * si, Kq = c * si;
        x[b] = x[b] + Kp * y;
        x[b + 4] = x[b + 4] + Kq * y;
        const double Ap = 1.0 - Kp, Aq = 0.0 - Kq;
        const double APpp = Ap * a, APpq = Ap * bb, APqp = Aq * a + c, APqq = Aq * bb + d;
        Pb[4 * b] = APpp * Ap + (Kp * r) * Kp;
        Pb[4 * b + 1] = (APpp * Aq + APpq) + (Kp * r) * Kq;
        Pb[4 * b + 2] = APqp * Ap + (Kq * r) * Kp;
        Pb[4 * b + 3] = (APqp * Aq + APqq) + (Kq * r) * Kq;
    }
    const double P33 = Pb[12];
    const double y = z[3] - x[3];
    const double si = 1.0 / (P33 + 10.0);
    const double K = P33 * si;
    x[3] = x[3] + K * y;
    const double A = 1.0 - K;
    Pb[12] = (A * P33) * A + (K * 10.0) * K;
}

// ORU in block form (same schedule as xysr_oru_replay).
__device__ inline void xysrb_oru_replay(double* x, double* Pb, const double* z1, const double* z2,
                                        int gap, const XysrConst& kc) {
    const double x1 = z1[0], y1 = z1[1];
    const double w1 = sqrt(z1[2] * z1[3]), h1 = sqrt(z1[2] / z1[3]);
    const double x2 = z2[0], y2 = z2[1];
    const double w2 = sqrt(z2[2] * z2[3]), h2 = sqrt(z2[2] / z2[3]);
    const double g = (double)gap;
    const double dx = (x2 - x1) / g, dy = (y2 - y1) / g;
    const double dw = (w2 - w1) / g, dh = (h2 - h1) / g;
    for (int i = 0; i < gap; i++) {
        const double f = (double)(i + 1);
        const double xx = x1 + f * dx, yy = y1 + f * dy;
        const double ww = w1 + f * dw, hh = h1 + f * dh;
        double zi[4] = {xx, yy, ww * hh, ww / hh};
        xysrb_update(x, Pb, zi);
        if (i != gap - 1) xysrb_predict(x, Pb, kc);
    }
}

// expand the block form to the dense row-major 7x7 (debug / unit ABI)
__host__ __device__ inline void xysrb_expand(const double* Pb, double* P) {
    for (int i = 0; i < 49; i++) P[i] = 0.0;
    for (int b = 0; b < 3; b++) {
        P[b * 7 + b] = Pb[4 * b];
        P[b * 7 + b + 4] = Pb[4 * b + 1];
        P[(b + 4) * 7 + b] = Pb[4 * b + 2];
        P[(b + 4) * 7 + b + 4] = Pb[4 * b + 3];
    }
    P[3 * 7 + 3] = Pb[12];
}

}  // namespace bm
