// IoU-family pair similarities in fp64, operation order as the numpy reference
// (boxmot/utils/iou.py: iou :50-67, hmiou :79-127, giou :129-169, centroid :171-184,
// ciou :199-264 (fork: eps 1e-7, aspect term ADDED), diou :266-307).  The library is
// built with -fmad=false so every product and sum rounds as numpy's does.
#pragma once
#include "common.cuh"

namespace bm {

enum AssoFunc : int { ASSO_IOU = 0, ASSO_GIOU = 1, ASSO_DIOU = 2, ASSO_CIOU = 3, ASSO_HMIOU = 4, ASSO_CENTROID = 5 };

struct Box { double x1, y1, x2, y2; };

__device__ __forceinline__ double iou_pair(const Box& a, const Box& b) {
    const double w = fmax(0.0, fmin(a.x2, b.x2) - fmax(a.x1, b.x1));
    const double h = fmax(0.0, fmin(a.y2, b.y2) - fmax(a.y1, b.y1));
    const double wh = w * h;
    return wh / ((a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1) - wh);
}

__device__ __forceinline__ double hmiou_pair(const Box& a, const Box& b) {
    const double ih = fmax(0.0, fmin(a.y2, b.y2) - fmax(a.y1, b.y1));
    const double uh = fmax(1e-10, fmax(a.y2, b.y2) - fmin(a.y1, b.y1));
    const double o = ih / uh;
    const double iw = fmax(0.0, fmin(a.x2, b.x2) - fmax(a.x1, b.x1));
    const double inter = iw * ih;
    const double uni = (a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1) - inter;
    return inter / (uni + 1e-10) * o;
}

__device__ __forceinline__ double giou_pair(const Box& a, const Box& b) {
    const double w = fmax(0.0, fmin(a.x2, b.x2) - fmax(a.x1, b.x1));
    const double h = fmax(0.0, fmin(a.y2, b.y2) - fmax(a.y1, b.y1));
    const double wh = w * h;
    const double uni = (a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1) - wh;
    const double iou = wh / uni;
    const double wc = fmax(a.x2, b.x2) - fmin(a.x1, b.x1);
    const double hc = fmax(a.y2, b.y2) - fmin(a.y1, b.y1);
    const double ac = wc * hc;
    const double g = iou - (ac - uni) / ac;
    return (g + 1.0) / 2.0;
}

__device__ __forceinline__ double diou_pair(const Box& a, const Box& b) {
    const double iou = iou_pair(a, b);
    const double dx = (a.x1 + a.x2) / 2.0 - (b.x1 + b.x2) / 2.0;
    const double dy = (a.y1 + a.y2) / 2.0 - (b.y1 + b.y2) / 2.0;
    const double inner = dx * dx + dy * dy;
    const double ow = fmax(a.x2, b.x2) - fmin(a.x1, b.x1);
    const double oh = fmax(a.y2, b.y2) - fmin(a.y1, b.y1);
    const double outer = ow * ow + oh * oh;
    return (iou - inner / outer + 1.0) / 2.0;
}

__device__ __forceinline__ double ciou_pair(const Box& a, const Box& b) {
    const double eps = 1e-7;
    const double w = fmax(0.0, fmin(a.x2, b.x2) - fmax(a.x1, b.x1));
    const double h = fmax(0.0, fmin(a.y2, b.y2) - fmax(a.y1, b.y1));
    const double wh = w * h;
    const double iou = wh / ((a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1) - wh + eps);
    const double dx = (a.x1 + a.x2) / 2.0 - (b.x1 + b.x2) / 2.0;
    const double dy = (a.y1 + a.y2) / 2.0 - (b.y1 + b.y2) / 2.0;
    const double inner = dx * dx + dy * dy;
    const double ow = fmax(a.x2, b.x2) - fmin(a.x1, b.x1);
    const double oh = fmax(a.y2, b.y2) - fmin(a.y1, b.y1);
    const double outer = ow * ow + oh * oh + eps;
    const double w1 = a.x2 - a.x1, h1 = a.y2 - a.y1 + eps;
    const double w2 = b.x2 - b.x1, h2 = b.y2 - b.y1 + eps;
    const double at = atan(w2 / h2) - atan(w1 / h1);
    const double v = (4.0 / (CUDART_PI * CUDART_PI)) * (at * at);
    const double alpha = v / ((1.0 - iou) + v + eps);
    return (iou - inner / outer + alpha * v + 1.0) / 2.0;
}

__device__ __forceinline__ double centroid_pair(const Box& a, const Box& b, double inv_norm_den) {
    const double dx = (a.x1 + a.x2) / 2.0 - (b.x1 + b.x2) / 2.0;
    const double dy = (a.y1 + a.y2) / 2.0 - (b.y1 + b.y2) / 2.0;
    return 1.0 - sqrt(dx * dx + dy * dy) / inv_norm_den;   // inv_norm_den = sqrt(w^2+h^2)
}

__device__ __forceinline__ double asso_pair(int func, const Box& a, const Box& b, double diag) {
    switch (func) {
        case ASSO_GIOU: return giou_pair(a, b);
        case ASSO_DIOU: return diou_pair(a, b);
        case ASSO_CIOU: return ciou_pair(a, b);
        case ASSO_HMIOU: return hmiou_pair(a, b);
        case ASSO_CENTROID: return centroid_pair(a, b, diag);
        default: return iou_pair(a, b);
    }
}

}  // namespace bm
