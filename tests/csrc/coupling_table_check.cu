// Host-only check of the coupling table (hfe_solver.cuh) against the exact libsvm procedure; built and run by tests/test_coupling_table.py.
#include <cstdio>
#include <random>
#include "hfe_solver.cuh"
int main() {
    auto edges = hfe::find_solver_edges();
    printf("edges %zu:", edges.size()); for (double e : edges) printf(" %.9f", e); printf("\n");
    const auto& tab = hfe::coupling_table(edges);
    printf("ok %d max fit error %.3e\n", (int)tab.ok, tab.max_fit_error);
    std::mt19937_64 rng(1); std::uniform_real_distribution<double> U(-18.5, 18.5);
    double worst = 0; int skipped = 0;
    for (int i = 0; i < 2000000; ++i) {
        double t = U(rng);
        double tq = std::fmin(std::fmax(t, hfe::COUPLING_T_LO), hfe::COUPLING_T_HI);
        double u = (tq - hfe::COUPLING_T_LO) * hfe::COUPLING_CELLS_PER_UNIT; int c = (int)u; if (c > hfe::COUPLING_CELLS - 1) c = hfe::COUPLING_CELLS - 1;
        double fr = u - c; bool right = fr >= tab.edge[c];
        if (std::fabs(fr - tab.edge[c]) / hfe::COUPLING_CELLS_PER_UNIT < 1e-7) { ++skipped; continue; }
        const double* q = &tab.coef[(size_t)c * 8 + (right ? 4 : 0)];
        double got = ((q[3] * fr + q[2]) * fr + q[1]) * fr + q[0];
        int it; double gap; double want = hfe::svc_probability_t(t, &it, &gap);
        worst = std::fmax(worst, std::fabs(got - want));
    }
    printf("random check worst %.3e skipped %d\n", worst, skipped);
}
