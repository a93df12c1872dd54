/* CPU oracle helper -- TEST INFRASTRUCTURE ONLY (never linked into the product path).
 *
 * oracle_ddot_skx: x . y in the summation order of OpenBLAS 0.3.30's SkylakeX ddot kernel
 * (kernel/x86_64/ddot.c driver + ddot_microk_skylakex-2.c), which is what numpy's np.dot /
 * np.linalg.norm of float64 vectors execute in the authoring container (threadpoolctl:
 * libscipy_openblas 0.3.30, architecture SkylakeX).  OpenBLAS is a third-party dependency
 * that is absent from /root/reference; its algorithm is restated here from its published
 * source: 32 interleaved FMA accumulators (four 8-lane vectors), lanes l and l+4 folded,
 * the four vectors added left to right, the two 128-bit halves added, then the horizontal
 * add; elements beyond the last multiple of 32 are folded in one by one with FMA.
 *
 * Why it matters: the fork's NearestNeighborDistanceMetric.partial_fit
 * (boxmot/trackers/strongsort/sort/linear_assignment.py:564-573) ranks the samples of a
 * target by np.linalg.norm(feature) -- all features are unit vectors, so the ranking is
 * decided by the last-ulp rounding of this dot product.  Pinned against np.dot in the
 * authoring container (tests/test_oracle.py: bit-equal on random unit vectors for every
 * length >= 32 tried); on a CPU for which OpenBLAS dispatches another kernel the reference
 * itself ranks differently. */
#include <math.h>

double oracle_ddot_skx(const double* x, const double* y, long n) {
    const long n1 = n & -32L;
    double dot = 0.0;
    if (n1) {
        double acc[32];
        for (int l = 0; l < 32; l++) acc[l] = 0.0;
        for (long i = 0; i < n1; i += 32)
            for (int l = 0; l < 32; l++) acc[l] = fma(x[i + l], y[i + l], acc[l]);
        double s[4];
        for (int l = 0; l < 4; l++) {
            const double a0 = acc[l] + acc[l + 4], a1 = acc[8 + l] + acc[8 + l + 4];
            const double a2 = acc[16 + l] + acc[16 + l + 4], a3 = acc[24 + l] + acc[24 + l + 4];
            s[l] = ((a0 + a1) + a2) + a3;
        }
        dot = (s[0] + s[2]) + (s[1] + s[3]);
    }
    for (long i = n1; i < n; i++) dot = fma(y[i], x[i], dot);
    return dot;
}
