// hfe_solver.cuh -- libsvm's pairwise-coupling probability for two classes, shared by host (edge finder) and device.
#pragma once

#include <cmath>
#include <mutex>
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


// ---- the coupling map as a table ---------------------------------------------------------------------------------------------
// P(class 1) as a function of the Platt argument t is one fixed scalar function (it depends on libsvm's constants only): smooth
// between the six sweep-count edges, constant beyond |t| = 16.2 (the clamp).  The fast passes evaluate it from a table instead
// of running the sweeps in float64 per row: [-17.5, 17.5] in cells of 1/128, per cell one cubic in the local coordinate
// u in [0, 1) for the part left of the edge the cell contains and one for the part right of it (the same cubic twice in the
// 4474 cells without an edge), fitted through Chebyshev nodes of the function with the sweep count held fixed (which is
// smooth across the edge).  Fit error is checked against the function itself when the table is built (<= 1e-9 asked for,
// ~1e-12 found); rows within the guard band of an edge are re-evaluated by the exact path as before.
constexpr double COUPLING_T_LO = -17.5, COUPLING_T_HI = 17.5;
constexpr int COUPLING_CELLS_PER_UNIT = 128;
constexpr int COUPLING_CELLS = 35 * COUPLING_CELLS_PER_UNIT;

// the sweeps of svc_probability_t with the exit test replaced by a fixed count
inline double svc_probability_fixed(double t, int sweeps) {
    double e = std::exp(-std::fabs(t));
    double r01 = t >= 0 ? e / (1.0 + e) : 1.0 / (1.0 + e);
    r01 = std::fmin(std::fmax(r01, 1e-7), 1.0 - 1e-7);
    double r10 = 1.0 - r01;
    double q00 = r10 * r10, q11 = r01 * r01, q01 = -r10 * r01;
    double p0 = 0.5, p1 = 0.5;
    for (int iter = 0; iter < sweeps; ++iter) {
        double qp0 = q00 * p0 + q01 * p1;
        double qp1 = q01 * p0 + q11 * p1;
        double pqp = p0 * qp0 + p1 * qp1;
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
    return p1;
}

struct CouplingTable {
    std::vector<double> coef;        // [COUPLING_CELLS][2][4]: c0..c3 of the left / right cubic
    std::vector<double> edge;        // [COUPLING_CELLS]: local coordinate of the edge in (or just outside) the cell, 1e9 if none
    double max_fit_error = 0.0;
    bool ok = false;
};

inline void fit_cubic(double t0, double h, int sweeps, double* c) {
    // interpolate through the four Chebyshev nodes of [0, 1] and convert to monomials (Newton form -> expansion)
    double u[4], v[4];
    for (int k = 0; k < 4; ++k) {
        u[k] = 0.5 - 0.5 * std::cos((2 * k + 1) * 3.14159265358979323846 / 8.0);
        v[k] = svc_probability_fixed(t0 + u[k] * h, sweeps);
    }
    double d[4] = {v[0], v[1], v[2], v[3]};
    for (int lvl = 1; lvl < 4; ++lvl)
        for (int k = 3; k >= lvl; --k) d[k] = (d[k] - d[k - 1]) / (u[k] - u[k - lvl]);
    // p(x) = d0 + d1 (x-u0) + d2 (x-u0)(x-u1) + d3 (x-u0)(x-u1)(x-u2)
    double poly[4] = {d[3], 0, 0, 0};                  // Horner from the top, expanding (x - u_k)
    int deg = 0;
    for (int k = 2; k >= 0; --k) {
        double next[4] = {0, 0, 0, 0};
        for (int i = 0; i <= deg; ++i) { next[i + 1] += poly[i]; next[i] -= poly[i] * u[k]; }
        next[0] += d[k];
        ++deg;
        for (int i = 0; i < 4; ++i) poly[i] = next[i];
    }
    for (int i = 0; i < 4; ++i) c[i] = poly[i];
}

inline const CouplingTable& coupling_table(const std::vector<double>& edges) {
    static CouplingTable tab;
    static std::once_flag once;
    std::call_once(once, [&] {
        const double h = 1.0 / COUPLING_CELLS_PER_UNIT;
        tab.coef.assign((size_t)COUPLING_CELLS * 8, 0.0);
        tab.edge.assign((size_t)COUPLING_CELLS, 1e9);
        bool ok = true;
        for (int c = 0; c < COUPLING_CELLS; ++c) {
            const double t0 = COUPLING_T_LO + c * h, t1 = t0 + h;
            int n_in = 0;
            double e_in = 0;
            for (double e : edges) {
                if (e >= t0 && e < t1) { ++n_in; e_in = e; }
                // an edge just outside the cell still has to trip the guard band of rows just inside it
                else if (e < t0 && t0 - e < 1e-3 * h) tab.edge[c] = (e - t0) / h;
                else if (e >= t1 && e - t1 < 1e-3 * h) tab.edge[c] = (e - t0) / h;
            }
            if (n_in > 1) { ok = false; break; }
            int it_l, it_r;
            double gap;
            if (n_in == 1) {
                tab.edge[c] = (e_in - t0) / h;
                const double tl = e_in > t0 ? t0 + 0.5 * (e_in - t0) : std::nextafter(t0, -1e300);
                svc_probability_t(tl, &it_l, &gap);                              // sweep count of the part left of the edge
                svc_probability_t(e_in + 0.5 * (t1 - e_in), &it_r, &gap);
            } else {
                svc_probability_t(t0 + 0.5 * h, &it_l, &gap);
                it_r = it_l;
            }
            fit_cubic(t0, h, it_l, &tab.coef[(size_t)c * 8]);
            fit_cubic(t0, h, it_r, &tab.coef[(size_t)c * 8 + 4]);
            // check against the function itself at points the fit did not see
            for (int k = 0; k < 12; ++k) {
                const double u = (k + 0.37) / 12.0, t = t0 + u * h;
                const bool right = u >= tab.edge[c];
                const double* q = &tab.coef[(size_t)c * 8 + (right ? 4 : 0)];
                int it;
                const double want = svc_probability_t(t, &it, &gap);
                if (it != (right ? it_r : it_l)) continue;                     // (within rounding of the edge)
                const double got = ((q[3] * u + q[2]) * u + q[1]) * u + q[0];
                tab.max_fit_error = std::fmax(tab.max_fit_error, std::fabs(got - want));
            }
        }
        tab.ok = ok && tab.max_fit_error <= 1e-9;
    });
    return tab;
}

}  // namespace hfe
