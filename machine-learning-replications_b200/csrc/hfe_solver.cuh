// hfe_solver.cuh -- libsvm's pairwise-coupling probability for two classes, shared by host (edge finder) and device.
#pragma once

#include <cmath>
#include <vector>

#include <cuda_runtime.h>

namespace hfe {

// svm.cpp:2035-2043 sigmoid_predict, :2932-2941 clamp, :2046-2107 multiclass_probability for k = 2, as a function of
// the Platt argument t = f*A + B.  Returns P(class 1); *iters = sweeps done before the exit test passed, *gap = min
// over the evaluated exit tests of |max_error - eps|.  The result is a smooth function of t except where *iters
// changes (six places on the whole real line, found at model-create time by find_solver_edges).
__host__ __device__ __noinline__ inline double svc_probability_t(double t, int* iters, double* gap) {
    double e = exp(-fabs(t));
    double r01 = t >= 0 ? e / (1.0 + e) : 1.0 / (1.0 + e);
    r01 = fmin(fmax(r01, 1e-7), 1.0 - 1e-7);
    double r10 = 1.0 - r01;
    double q00 = r10 * r10, q11 = r01 * r01, q01 = -r10 * r01;
    double p0 = 0.5, p1 = 0.5;
    const double eps = 0.005 / 2;
    double g = 1e300;
    int iter = 0;
#pragma unroll 1
    for (; iter < 100; ++iter) {
        double qp0 = q00 * p0 + q01 * p1;
        double qp1 = q01 * p0 + q11 * p1;
        double pqp = p0 * qp0 + p1 * qp1;
        double err = fmax(fabs(qp0 - pqp), fabs(qp1 - pqp));
        g = fmin(g, fabs(err - eps));
        if (err < eps) break;
        double diff = (-qp0 + pqp) / q00;
        p0 += diff;
        pqp = (pqp + diff * (diff * q00 + 2 * qp0)) / (1 + diff) / (1 + diff);
        qp0 = (qp0 + diff * q00) / (1 + diff);
        qp1 = (qp1 + diff * q01) / (1 + diff);
        p0 /= (1 + diff);
        p1 /= (1 + diff);
        diff = (-qp1 + pqp) / q11;
        p1 += diff;
        p0 /= (1 + diff);
        p1 /= (1 + diff);
    }
    *gap = g;
    *iters = iter;
    return p1;
}

// Where, on the t axis, does the solver's sweep count change?  Those are the only discontinuities of
// svc_probability_t; rows whose float32-derived t falls within a guard band of one are re-evaluated in float64.
// The edges depend only on libsvm's constants (eps = 0.005/k), not on the model.  Adaptive scan: a cell is split
// while an exit test at either end came closer to flipping than the test can move across the cell
// (|d max_error / d t| <= 1), so a double crossing inside one cell is not missed.
inline void scan_cell(double a, int ia, double ga, double b, int ib, double gb, int depth, std::vector<double>* out) {
    double h = b - a;
    if (ia != ib) {
        if (h < 1e-12 || depth > 60) { out->push_back(0.5 * (a + b)); return; }
    } else if (fmin(ga, gb) > h || h < 1e-7) {
        return;
    }
    double m = 0.5 * (a + b), gm;
    int im;
    svc_probability_t(m, &im, &gm);
    scan_cell(a, ia, ga, m, im, gm, depth + 1, out);
    scan_cell(m, im, gm, b, ib, gb, depth + 1, out);
}

inline std::vector<double> find_solver_edges() {
    std::vector<double> edges;
    const double lo = -17.5, hi = 17.5, step = 1.0 / 256;     // r is clamped to [1e-7, 1-1e-7] beyond |t| = 16.2
    double a = lo, ga;
    int ia;
    svc_probability_t(a, &ia, &ga);
    for (double b = lo + step; b <= hi + 1e-9; b += step) {
        double gb;
        int ib;
        svc_probability_t(b, &ib, &gb);
        scan_cell(a, ia, ga, b, ib, gb, 0, &edges);
        a = b; ia = ib; ga = gb;
    }
    // merge duplicates produced by adjacent cells
    std::vector<double> merged;
    for (double e : edges)
        if (merged.empty() || e - merged.back() > 1e-9) merged.push_back(e);
    return merged;
}

}  // namespace hfe
