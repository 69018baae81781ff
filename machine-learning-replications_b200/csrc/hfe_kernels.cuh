// hfe_kernels.cuh -- sm_100a kernels of the scoring path.
//
// hfe_score_kernel<SVC, TREE_MODE, PFX> is the fused pass: one thread owns ROWS_PER_THREAD (=4) consecutive rows, fetched
// from the column-major float32 matrix with one 128-bit load per feature (a warp reads 512 contiguous bytes of each
// column).  Per row it evaluates, without leaving registers,
//   a6/a7  StandardScaler + RBF-SVC decision + Platt + clamp + 2-class coupling   (svm.cpp:461-515, 2821-2958, 2046-2107)
//   a8     boosted trees (merged stump table or general node tables) + expit        (_gradient_boosting.pyx:25-73)
//   a10    logistic member on the raw features                                       (linear_model/_base.py:366-451)
//   a11    combiner + hard label (z > 0)                                             (linear_model/_base.py:398-451)
// and writes [1-p, p] (8 bytes per row).  The RBF sum -- ~98 % of the reference's CPU time -- runs in float32 on the
// FMA / MUFU pipes out of a shared-memory support-vector table (staged by a TMA bulk copy, read with warp-broadcast
// LDS.128); with the prefix-shared table layout (PFX, hfe_model.cuh) partial dot products over the leading features are
// computed once per distinct prefix.  Everything after the sum runs in float64 on B200's full-rate FP64 pipe.
// Rows whose result could differ from a float64 evaluation by a branch (coupling-solver exit test, label sign) are
// appended to the per-call work list; hfe_fixup_kernel, launched right behind, re-evaluates them in float64.
//
// hfe_stream_kernel<TREE_MODE> scores stacks without the SVC member, which are bandwidth-bound: persistent warps,
// feature tiles staged in shared memory by 2-D TMA tensor copies, float32 links.
// The remaining kernels are the data-format and evaluation steps either side of the path (ingest, synthetic rows,
// leaf indices, 1-NN imputation, confusion counts / probability histograms).
#pragma once

#include <cuda.h>          // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)

#include "hfe_model.cuh"

namespace hfe {

#ifndef HFE_BLOCK_THREADS
#define HFE_BLOCK_THREADS 128
#endif
constexpr int BLOCK_THREADS = HFE_BLOCK_THREADS;
constexpr int ROWS_PER_THREAD = 4;
constexpr int BLOCK_ROWS = BLOCK_THREADS * ROWS_PER_THREAD;
constexpr int MAX_SV_TABLE_BYTES = (384 * 18 + 2) * 16;   // staged support-vector table (flat layout: 1536 vectors; 108 KB, 2 blocks/SM)
#ifndef HFE_STREAM_WINDOW
#define HFE_STREAM_WINDOW 17
#endif
constexpr int STREAM_WINDOW = HFE_STREAM_WINDOW;          // feature columns in flight per thread in the SVC-free streaming prologue
constexpr int TREE_CHUNK = 16;            // float32 partial sums over trees are folded into float64 every 16 trees
#ifndef HFE_PERSISTENT
#define HFE_PERSISTENT 0                  // 1: resident blocks loop over tiles; 0: one 128-row tile per warp (measured faster)
#endif
#ifndef HFE_MIN_BLOCKS_NOSVC
#define HFE_MIN_BLOCKS_NOSVC 4            // SVC-free streaming pass: all 17 column loads of a thread in flight at once (measured best)
#endif
#ifndef HFE_PFX_FLUSH
#define HFE_PFX_FLUSH 1                   // float32 partial sums go to float64 after every leaf step (1) or every 2nd (2)
#endif
#ifndef HFE_MIN_BLOCKS_STREAM
#define HFE_MIN_BLOCKS_STREAM 5           // streaming pass: 5 blocks x 4 warps, one 8.5 KB tile buffer per warp (measured: 5 blocks 72 G rows/s, 6 blocks at 80 registers 68 G)
#endif
#ifndef HFE_STREAM_TMA
#define HFE_STREAM_TMA 1                  // SVC-free stacks with stumps / no trees go through hfe_stream_kernel
#endif
#ifndef HFE_MIN_BLOCKS_PFX
#define HFE_MIN_BLOCKS_PFX 4              // same, for the prefix-shared support-vector layout
#endif
#ifndef HFE_MIN_BLOCKS
#define HFE_MIN_BLOCKS 4                  // resident blocks per SM the fused kernel is compiled for (128-register cap)
#endif

// -DHFE_BOUNDS_CHECK builds a library whose every access to the feature matrix, the permutation index and the outputs
// is checked against the extents the caller declared; a violation traps (the launch then fails with an error).  The
// test-suite runs its ragged / permuted / tiny cases against that build (compute-sanitizer is not available on the
// test pool).
#ifdef HFE_BOUNDS_CHECK
#define HFE_ASSERT_IN(cond) do { if (!(cond)) __trap(); } while (0)
#else
#define HFE_ASSERT_IN(cond) do { } while (0)
#endif

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __noinline__ double expit_f64(double x) {
    // scipy.special.expit (what SK/_loss/link.py:180 and SK/linear_model/_base.py:429-451 call), sign-split
    double e = exp(-fabs(x));
    return x >= 0 ? 1.0 / (1.0 + e) : e / (1.0 + e);
}

// ---- fast float64 helpers for the per-row epilogue of the fused kernel -----------------------------------------------
// The epilogue runs once per row but in float64, so the library exp() / division (special-case branches, ~30-40
// instructions each) would cost as much as a third of the whole pass.  These versions are branch-free, accurate to
// ~1e-15 relative on the ranges they are given, and are only used where a guard band protects every branch that
// depends on them; the cooperative float64 re-evaluation path keeps the library versions.
__device__ __forceinline__ double rcp_fast(double x) {          // x in [1e-16, 1e16], normal
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));       // MUFU.RCP64H, ~2^-23 relative
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}

__device__ __forceinline__ double exp_neg_fast(double x) {      // exp(x) for x <= 0
    x = fmax(x, -700.0);
    const double shifter = 6755399441055744.0;                  // 1.5 * 2^52: rint() by addition
    double t = fma(x, 1.4426950408889634074, shifter);
    int k = __double2loint(t);
    double kd = t - shifter;
    double r = fma(kd, -6.93147180369123816490e-01, x);
    r = fma(kd, -1.90821492927058770002e-10, r);                // |r| <= ln2/2
    double p = 2.50521083854417187751e-08;                      // 1/11!
    p = fma(p, r, 2.75573192239858906526e-07);
    p = fma(p, r, 2.75573192239858906526e-06);
    p = fma(p, r, 2.48015873015873015873e-05);
    p = fma(p, r, 1.98412698412698412698e-04);
    p = fma(p, r, 1.38888888888888888889e-03);
    p = fma(p, r, 8.33333333333333333333e-03);
    p = fma(p, r, 4.16666666666666666667e-02);
    p = fma(p, r, 1.66666666666666666667e-01);
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));   // p * 2^k, k >= -1010
}

__device__ __forceinline__ double expit_fast(double x) {
    double e = exp_neg_fast(-fabs(x));
    double inv = rcp_fast(1.0 + e);
    return x >= 0 ? inv : e * inv;
}

// float32 logistic link for the streaming pass: e = 2^(-|x| log2 e) in (0, 1] by MUFU.EX2, 1/(1+e) by MUFU.RCP + one
// Newton step; |error| <= ~1.8e-7 absolute.
__device__ __forceinline__ float expit32(float x) {
    const float e = ex2_approx(-fabsf(x) * 1.4426950408889634f);
    const float d = 1.f + e;
    float inv;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(d));
    inv = fmaf(inv, fmaf(-d, inv, 1.f), inv);
    return x >= 0.f ? inv : e * inv;
}

// Links of the members that do not depend on the RBF sum, and their share of the combiner's logit:
// zpart = intercept + sum over {trees, logistic} of final_coef * P(class 1).
template <int R>
__device__ __forceinline__ void partial_logit_fast(const ScoreParams& P, bool has_gbt, bool has_lin,
                                                   const double (&raw)[R], const double (&zl)[R], double (&zpart)[R],
                                                   double (&member)[3][R]) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        member[0][r] = 0.0;
        member[1][r] = has_gbt ? expit_fast(raw[r]) : 0.0;          // SK/_loss/loss.py:980-998
        member[2][r] = has_lin ? expit_fast(zl[r]) : 0.0;           // SK/linear_model/_base.py:429-451
        double z = P.final_intercept;
        if (has_gbt) z = fma(P.final_coef[P.slot_gbt], member[1][r], z);
        if (has_lin) z = fma(P.final_coef[P.slot_lin], member[2][r], z);
        zpart[r] = z;
    }
}

// Platt + clamp + 2-class coupling for the SVC member, then the combiner, for R rows at once with the R dependency
// chains interleaved.
template <int R>
__device__ __forceinline__ void finish_rows_fast(const ScoreParams& P, bool has_svc, const double (&f_svc)[R],
                                                 const double (&zpart)[R], double (&p1)[R], double (&zfin)[R],
                                                 double (&p_svc)[R], bool (&uncertain)[R]) {
#pragma unroll
    for (int r = 0; r < R; ++r) { uncertain[r] = false; p_svc[r] = 0.0; }
    if (has_svc) {
        // Platt, clamp and the two-class coupling sweeps (svm.cpp:2035-2107, 2932-2941) are one fixed function of
        // t = f*A + B: it is read from the table build_model uploads (hfe_solver.cuh: a cubic per 1/128 of t and side of the
        // sweep-count edge, fit error ~3e-11) instead of being iterated per row.  Rows within the guard band of an edge go to
        // the exact path as before.
        double fr[R];
        int cell[R];
        bool nan_t[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double t = fma(f_svc[r], P.prob_a, P.prob_b);
            nan_t[r] = !(t == t);
            const double u = (fmin(fmax(t, COUPLING_T_LO), COUPLING_T_HI) - COUPLING_T_LO) * (double)COUPLING_CELLS_PER_UNIT;
            cell[r] = min(__double2int_rz(u), COUPLING_CELLS - 1);            // (NaN -> cell 0; such rows are flagged elsewhere)
            fr[r] = u - (double)cell[r];
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double edge = __ldg(P.coupling_edge + cell[r]);
            const double dist = fr[r] - edge;
            uncertain[r] = nan_t[r] || !(fabs(dist) * (1.0 / COUPLING_CELLS_PER_UNIT) >= P.solver_guard);
            const double2* q = reinterpret_cast<const double2*>(P.coupling_coef + (size_t)cell[r] * 8 + (dist >= 0.0 ? 4 : 0));
            const double2 c01 = __ldg(q), c23 = __ldg(q + 1);
            p_svc[r] = fma(fma(fma(c23.y, fr[r], c23.x), fr[r], c01.y), fr[r], c01.x);
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        double z = zpart[r];
        if (has_svc) z = fma(P.final_coef[P.slot_svc], p_svc[r], z);
        zfin[r] = z;
        p1[r] = expit_fast(z);
        uncertain[r] = uncertain[r] || !(fabs(z) >= P.label_guard);
    }
}

struct RowResult {
    double p1, z;
    double member[3];   // svc, gbt, lin
    bool uncertain;
};

// Everything after the member decision values, in float64.
__device__ __noinline__ RowResult finish_row(const ScoreParams& P, bool has_svc, bool has_gbt, bool has_lin,
                                                double f_svc, double raw_gbt, double z_lin) {
    RowResult r;
    r.uncertain = false;
    double z = P.final_intercept;
    r.member[0] = r.member[1] = r.member[2] = 0.0;
    // meta-features are combined in estimators_ order (SK/ensemble/_stacking.py:87-135)
#pragma unroll
    for (int slot = 0; slot < MAX_MEMBERS; ++slot) {
        if (has_svc && slot == P.slot_svc) {
            double gap, t = f_svc * P.prob_a + P.prob_b;
            int iters;
            double p = svc_probability_t(t, &iters, &gap);
            // a solver branch is uncertain iff t lies within the guard band of one of the exit-test edges
            double nearest = 1e300;
            for (int e = 0; e < P.n_edges; ++e) nearest = fmin(nearest, fabs(t - P.edges[e]));
            r.uncertain = r.uncertain || (nearest < P.solver_guard);
            r.member[0] = p;
            z += P.final_coef[slot] * p;
        } else if (has_gbt && slot == P.slot_gbt) {
            double p = expit_f64(raw_gbt);
            r.member[1] = p;
            z += P.final_coef[slot] * p;
        } else if (has_lin && slot == P.slot_lin) {
            double p = expit_f64(z_lin);
            r.member[2] = p;
            z += P.final_coef[slot] * p;
        }
    }
    r.z = z;
    r.p1 = expit_f64(z);
    r.uncertain = r.uncertain || (fabs(z) < P.label_guard);
    return r;
}

__device__ __forceinline__ void store_row(const ScoreParams& P, long long row, const RowResult& r,
                                          bool write_out = true) {
    HFE_ASSERT_IN(row >= 0 && row < P.n);
    if (P.out && write_out) {
        P.out[2 * row] = (float)(1.0 - r.p1);
        P.out[2 * row + 1] = (float)r.p1;
    }
    if (P.out64) {
        P.out64[2 * row] = 1.0 - r.p1;
        P.out64[2 * row + 1] = r.p1;
    }
    if (P.labels) P.labels[row] = r.z > 0 ? 1 : 0;
    if (P.decision) P.decision[row] = r.z;
    if (P.meta) {
        if (P.slot_svc >= 0) P.meta[row * P.n_members + P.slot_svc] = r.member[0];
        if (P.slot_gbt >= 0) P.meta[row * P.n_members + P.slot_gbt] = r.member[1];
        if (P.slot_lin >= 0) P.meta[row * P.n_members + P.slot_lin] = r.member[2];
    }
}

__device__ __forceinline__ float load_feature(const ScoreParams& P, int f, long long row) {
    HFE_ASSERT_IN(row >= 0 && row < P.n && f >= 0 && f < F);
    long long src = (f == P.perm_col) ? (long long)P.perm_idx[row] : row;
    HFE_ASSERT_IN(src >= 0 && src < P.n);
    return __ldg(P.x + P.col_off[f] + src);
}

// Four consecutive rows of one feature column.  The 128-bit load is always in bounds: row0 is a multiple of 4 below n
// and ld is a multiple of 4 not below n.  It is kept free of control flow so that a thread's 17 column loads can all
// be in flight at once; fix_rows4 patches the rare cases afterwards.
__device__ __forceinline__ float4 load_rows4(const ScoreParams& P, int f, long long row0) {
    HFE_ASSERT_IN(row0 >= 0 && row0 + 3 < P.ld && row0 < P.n && (row0 & 3) == 0);
    return __ldg(reinterpret_cast<const float4*>(P.x + P.col_off[f] + row0));
}

// Rows at or beyond n (ragged last tile) read padding and are zeroed; a permuted column is gathered through perm_idx.
// ``row_ld`` is the row the loads use (threads whose whole quad lies beyond n load quad 0 instead, so every address
// stays inside the matrix); ``row0`` is the thread's real first row.
__device__ __forceinline__ float4 fix_rows4(const ScoreParams& P, int f, long long row0, long long row_ld, bool full,
                                            float4 v) {
    if (f == P.perm_col) {
        const float* col = P.x + P.col_off[f];
        const int* idx = P.perm_idx + row_ld;
        HFE_ASSERT_IN(row_ld >= 0 && row_ld < P.n && __ldg(idx) >= 0 && __ldg(idx) < P.n);
        v.x = __ldg(col + __ldg(idx));
        if (full || row_ld + 1 < P.n) v.y = __ldg(col + __ldg(idx + 1));
        if (full || row_ld + 2 < P.n) v.z = __ldg(col + __ldg(idx + 2));
        if (full || row_ld + 3 < P.n) v.w = __ldg(col + __ldg(idx + 3));
    }
    if (!full) {
        if (row0 + 0 >= P.n) v.x = 0.f;
        if (row0 + 1 >= P.n) v.y = 0.f;
        if (row0 + 2 >= P.n) v.z = 0.f;
        if (row0 + 3 >= P.n) v.w = 0.f;
    }
    return v;
}

// Merged depth-1 stumps of feature f (compile-time index): the first split of every feature sits in the parameter
// block as immediate constant operands (threshold -inf, delta 0 where a feature has none), further splits of the same
// feature in a table behind it.
template <int R>
__device__ __forceinline__ void apply_splits(const ScoreParams& P, int f, const float (&x)[R], float (&acc)[R]) {
    const float thr0 = P.split1_thr[f], dl0 = P.split1_delta[f];
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (x[r] <= thr0) acc[r] += dl0;
    if (P.split_begin[f + 1] != P.split_begin[f]) {
#pragma unroll 1
        for (int s = P.split_begin[f]; s < P.split_begin[f + 1]; ++s) {
            const float thr = P.split_thr[s], dl = P.split_delta[s];
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (x[r] <= thr) acc[r] += dl;
        }
    }
}

// Walk one tree for one row: x values come from `xrow` (F floats, any address space).  Returns the re-laid leaf index
// (global node index).
template <typename XGet>
__device__ __forceinline__ int walk_tree(const ScoreParams& P, int t, XGet xget) {
    int depth = P.tree_depth[t];
    unsigned node = (unsigned)P.tree_off[t];
    for (int lvl = 0; lvl < depth; ++lvl) {
        uint2 nd = __ldg(P.nodes + node);
        float thr = __uint_as_float(nd.x);
        float xv = xget(nd.y >> 24);
        node = (nd.y & 0xFFFFFFu) + (xv > thr ? 1u : 0u);
    }
    return (int)node;
}

// Cooperative float64 evaluation of one row by one warp -- the exact-arithmetic path.  Follows the reference's
// operation order within each term (scale: SK/preprocessing/_data.py:1132-1134; kernel: svm.cpp:461-515); only the
// order in which the n_sv kernel values / the trees are summed differs (lane-strided partials + shuffle reduction).
// All 32 lanes must call it together.  Lane f (< F) holds feature f of the row; other lanes fetch it by shuffle.
// Returns (on every lane) whether the hard label equals y[row] when y is given.
__device__ __noinline__ bool eval_row_f64_warp(const ScoreParams& P, long long row,
                                               const double* x_rowmajor = nullptr, unsigned* nonfinite_flag = nullptr) {
    const unsigned full_mask = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const bool has_svc = P.slot_svc >= 0, has_gbt = P.slot_gbt >= 0, has_lin = P.slot_lin >= 0;
    // x_rowmajor: a row-major float64 [n, F] source (col_off[] then holds plain column numbers: ld = 1), cast to
    // float32 like the ingest kernel does
    float x_mine = lane < F ? (x_rowmajor ? (float)x_rowmajor[row * F + P.col_off[lane]] : load_feature(P, lane, row)) : 0.f;
    double part_svc = 0.0, part_gbt = 0.0;
    if (has_svc) {
        double z[F];
#pragma unroll
        for (int d = 0; d < F; ++d) z[d] = ((double)__shfl_sync(full_mask, x_mine, d) - P.mean[d]) / P.scale[d];
        for (int j = lane; j < P.n_sv; j += 32) {
            const double* s = P.sv64 + (long long)j * F;
            double sum = 0.0;
#pragma unroll
            for (int d = 0; d < F; ++d) {
                double diff = z[d] - s[d];
                sum += diff * diff;
            }
            part_svc += P.dual64[j] * exp(-P.gamma * sum);
        }
    }
    if (has_gbt) {
        for (int t0 = 0; t0 < P.n_trees; t0 += 32) {          // uniform trip counts: every lane joins every shuffle
            const int t = min(t0 + lane, P.n_trees - 1);
            unsigned node = (unsigned)P.tree_off[t];
            for (int lvl = 0; lvl < P.max_depth; ++lvl) {       // leaves self-loop, so walking past a leaf is harmless
                uint2 nd = __ldg(P.nodes + node);
                float xv = __shfl_sync(full_mask, x_mine, (int)(nd.y >> 24));
                node = (nd.y & 0xFFFFFFu) + (xv > __uint_as_float(nd.x) ? 1u : 0u);
            }
            if (t0 + lane < P.n_trees) part_gbt += P.value64[node];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        part_svc += __shfl_down_sync(full_mask, part_svc, o);
        part_gbt += __shfl_down_sync(full_mask, part_gbt, o);
    }
    double zl = 0.0;
    bool finite = true;
#pragma unroll
    for (int d = 0; d < F; ++d) {
        float xd = __shfl_sync(full_mask, x_mine, d);
        zl += (double)xd * P.lin_coef[d];
        finite = finite && (fabsf(xd) <= 3.402823466e38f);
    }
    zl += P.lin_intercept;
    int label = 0;
    if (lane == 0) {
        RowResult r = finish_row(P, has_svc, has_gbt, has_lin, part_svc + P.svc_intercept, part_gbt + P.gbt_init, zl);
        if (!finite) {
            r.p1 = r.z = __longlong_as_double(0x7FF8000000000000LL);
            r.member[0] = r.member[1] = r.member[2] = r.p1;
            if (nonfinite_flag) atomicOr(nonfinite_flag, HFE_ROW_NONFINITE);
        }
        store_row(P, row, r);
        if (P.svc_dec) P.svc_dec[row] = part_svc + P.svc_intercept;
        label = r.z > 0 ? 1 : 0;
    }
    label = __shfl_sync(full_mask, label, 0);
    return P.y ? (label == (int)P.y[row]) : false;
}

// A row whose result hinges on a branch goes on the call's work list.  The checked build (-DHFE_BOUNDS_CHECK) also counts
// the appends whose entry has been written in the reserved header word; hfe_fixup_kernel traps if that count differs from
// the list length it sees -- the stand-in for a race checker on the hand-off between the two launches.
__device__ __forceinline__ void worklist_append(const ScoreParams& P, long long row) {
    const unsigned slot = atomicAdd(P.status + HFE_STATUS_LIST_LEN, 1u);
    if (slot < (unsigned)HFE_WORKLIST_CAPACITY) P.status[HFE_STATUS_HEADER_WORDS + slot] = (unsigned)row;
    else atomicOr(P.status, HFE_WORKLIST_OVERFLOW);
#ifdef HFE_BOUNDS_CHECK
    __threadfence();
    atomicAdd(P.status + 3, 1u);
#endif
}

// Outputs of one thread's R rows, the work list and the accuracy count -- shared by the fused pass and the streaming pass.
template <int R, typename T>
__device__ __forceinline__ void emit_rows(const ScoreParams& P, long long row0, int lane, bool full, bool nonfinite,
                                          unsigned fixmask, const T (&p1)[R], const T (&zfin)[R],
                                          const T (&p_svc)[R], const bool (&unc)[R]) {
    unsigned char lab[R];
    float po[2 * R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        po[2 * r] = (float)(T(1) - p1[r]);
        po[2 * r + 1] = (float)p1[r];
        lab[r] = zfin[r] > 0 ? 1 : 0;
        if (unc[r]) fixmask |= 1u << r;
        if (row0 + r >= P.n) continue;
        if (P.out && !full) { P.out[2 * (row0 + r)] = po[2 * r]; P.out[2 * (row0 + r) + 1] = po[2 * r + 1]; }
        if (P.out64) { P.out64[2 * (row0 + r)] = (double)(T(1) - p1[r]); P.out64[2 * (row0 + r) + 1] = (double)p1[r]; }
        if (P.labels) P.labels[row0 + r] = lab[r];
        if (P.decision) P.decision[row0 + r] = zfin[r];
        if (P.meta && P.slot_svc >= 0) P.meta[(row0 + r) * P.n_members + P.slot_svc] = p_svc[r];
    }
    if (P.out && full) {                        // 4 rows x [p0, p1] = two 128-bit stores, 32 contiguous bytes per thread
        float4* o = reinterpret_cast<float4*>(P.out + 2 * row0);
        o[0] = make_float4(po[0], po[1], po[2], po[3]);
        o[1] = make_float4(po[4], po[5], po[6], po[7]);
    }
    if (!full) fixmask &= (1u << (unsigned)max(0LL, min((long long)R, P.n - row0))) - 1u;
    if (nonfinite) fixmask = full ? (1u << R) - 1 : fixmask | ((1u << (unsigned)max(0LL, min((long long)R, P.n - row0))) - 1u);

    // ---- rows whose result hinges on a branch go to the work list; hfe_fixup_kernel re-evaluates them in float64 -------
    if (fixmask) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if ((fixmask >> r) & 1u) {
                worklist_append(P, row0 + r);
            }
        }
    }
    if (nonfinite) atomicOr(P.status, HFE_ROW_NONFINITE);

    if (P.y && P.score) {                                         // accuracy count; flagged rows are counted by the fix-up
        unsigned hits = 0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (row0 + r < P.n && !((fixmask >> r) & 1u)) hits += (lab[r] == P.y[row0 + r]) ? 1u : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hits += __shfl_down_sync(0xFFFFFFFFu, hits, o);
        if (lane == 0 && hits) atomicAdd(P.score, (unsigned long long)hits);
    }
}

// ---- the fused pass ----------------------------------------------------------------------------------------------------

constexpr int WARP_ROWS = 32 * ROWS_PER_THREAD;          // rows one warp scores per tile
constexpr int BLOCK_WARPS = BLOCK_THREADS / 32;

template <bool SVC, int TREE_MODE, bool PFX>
__global__ void __launch_bounds__(BLOCK_THREADS, SVC ? (PFX ? HFE_MIN_BLOCKS_PFX : HFE_MIN_BLOCKS) : HFE_MIN_BLOCKS_NOSVC)
hfe_score_kernel(const __grid_constant__ ScoreParams P, int sv_bytes) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* s_sv = reinterpret_cast<float4*>(smem_raw);                       // the support-vector table (sv_bytes)
    float* s_xt = reinterpret_cast<float*>(smem_raw) + (SVC ? (size_t)sv_bytes / 4 : 0);      // [F][BLOCK_ROWS]

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int n_tiles = (int)((P.n + WARP_ROWS - 1) / WARP_ROWS);
    // The support-vector table is staged in shared memory once per block (the only block-level barrier); after that
    // every warp scores its own 128-row tile(s) with no further block-level synchronisation.  With HFE_PERSISTENT the
    // grid is sized to the resident block count and warps loop over tiles; measured on B200 the plain one-tile-per-warp
    // grid is ~15 % faster (the hardware block scheduler staggers blocks better than a static tile assignment).
    // TMA bulk copy (cp.async.bulk, SASS UBLKCP): one thread asks the copy engine for the whole table (plus its zero
    // pad) and every thread goes straight on to load and standardise its rows; the mbarrier is only waited on right
    // before the RBF loop, by which time the bytes have normally landed.
    __shared__ __align__(8) unsigned long long s_table_ready;
    if (SVC) {
        const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_table_ready);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (tid == 0) {
            const unsigned bytes = (unsigned)sv_bytes;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"((unsigned)__cvta_generic_to_shared(s_sv)), "l"(P.sv_tab), "r"(bytes), "r"(mbar)
                         : "memory");
        }
    }
    // SVC-free instances are short streaming passes: resident blocks loop over tiles (grid-stride) so the code stays
    // warm in the instruction cache.  With the RBF loop the one-tile-per-warp grid measured ~15 % faster.
    constexpr bool kPersistent = HFE_PERSISTENT || !SVC;
    for (int tile = blockIdx.x * BLOCK_WARPS + (tid >> 5); tile < n_tiles;
         tile += kPersistent ? gridDim.x * BLOCK_WARPS : n_tiles) {
    const long long row0 = (long long)tile * WARP_ROWS + lane * ROWS_PER_THREAD;
    const bool has_lin = P.slot_lin >= 0;
    constexpr int R = ROWS_PER_THREAD;

    const bool full = (row0 + R <= P.n);
    const long long row_ld = row0 < P.n ? row0 : 0;      // out-of-range threads of the last tile load quad 0
    double raw[R], zl[R];
    float zz32[R];
    float zs[F][R];                 // standardised features, one scalar register per (feature, row)
    bool nonfinite;
#pragma unroll
    for (int r = 0; r < R; ++r) { raw[r] = 0.0; zl[r] = P.lin_intercept; zz32[r] = 0.f; }
    if (!SVC) {
        // ---- streaming prologue for stacks without the SVC member (the HBM-bound configuration): one feature column at
        //      a time -- load, finiteness probe, logistic dot product, that feature's stump splits -- so only a few
        //      registers are live and twice as many warps fit on an SM to keep loads in flight ------------------------
        float finite_probe = 0.f, acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
        // a window of STREAM_WINDOW feature columns is kept in flight: column f+W is requested before column f is used
        // (measured on B200, 10 M rows: W=6/32 warps 50 %, W=12/20 warps 52 %, W=17/16 warps 56 % of HBM copy bandwidth)
        constexpr int W = STREAM_WINDOW;
        auto load_col = [&](int f) -> float4 { return load_rows4(P, f, row_ld); };
        const bool needs_fix = !full || P.perm_col >= 0;
        float4 win[W];
#pragma unroll
        for (int i = 0; i < W; ++i) win[i] = load_col(i);
#pragma unroll
        for (int f = 0; f < F; ++f) {
            float4 v = win[f % W];
            if (f + W < F) win[f % W] = load_col(f + W);
            if (needs_fix) v = fix_rows4(P, f, row0, row_ld, full, v);
            float xf[R];
            xf[0] = v.x; xf[1] = v.y; xf[2] = v.z; xf[3] = v.w;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (has_lin) zl[r] = fma((double)xf[r], P.lin_coef[f], zl[r]);   // NaN/Inf in x makes zl non-finite
                else finite_probe = fmaf(xf[r], 0.f, finite_probe);
                if (TREE_MODE == 2) s_xt[f * BLOCK_ROWS + r * BLOCK_THREADS + tid] = xf[r];
            }
            if (TREE_MODE == 1) apply_splits<R>(P, f, xf, acc);
        }
        nonfinite = !(finite_probe == 0.f);
        if (has_lin) {
#pragma unroll
            for (int r = 0; r < R; ++r) nonfinite = nonfinite || !(fabs(zl[r]) <= 1.7976931348623157e308);
        }
        if (TREE_MODE == 1) {
#pragma unroll
            for (int r = 0; r < R; ++r) raw[r] = P.gbt_split_base + (double)acc[r];
        }
    } else {
    // ---- load 4 rows x 17 features (128-bit loads; guarded scalar loads on the ragged tail / permuted column) ------
    float xv[F][R];
#pragma unroll
    for (int f = 0; f < F; ++f) {
        float4 v = load_rows4(P, f, row_ld);
        xv[f][0] = v.x; xv[f][1] = v.y; xv[f][2] = v.z; xv[f][3] = v.w;
    }
    if (!full || P.perm_col >= 0) {              // rare: ragged last tile, or a permutation-importance pass
#pragma unroll
        for (int f = 0; f < F; ++f) {
            float4 v = fix_rows4(P, f, row0, row_ld, full, make_float4(xv[f][0], xv[f][1], xv[f][2], xv[f][3]));
            xv[f][0] = v.x; xv[f][1] = v.y; xv[f][2] = v.z; xv[f][3] = v.w;
        }
    }
    float finite_probe = 0.f;                       // x * 0 is NaN exactly when x is NaN or +-Inf
#pragma unroll
    for (int f = 0; f < F; ++f)
#pragma unroll
        for (int r = 0; r < R; ++r) finite_probe = fmaf(xv[f][r], 0.f, finite_probe);
    nonfinite = !(finite_probe == 0.f);

    if (TREE_MODE == 2) {
#pragma unroll
        for (int f = 0; f < F; ++f)
#pragma unroll
            for (int r = 0; r < R; ++r) s_xt[f * BLOCK_ROWS + r * BLOCK_THREADS + tid] = xv[f][r];
    }

    // ---- boosted trees ---------------------------------------------------------------------------------------------
    if (TREE_MODE == 1) {
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll
        for (int f = 0; f < F; ++f) apply_splits<R>(P, f, xv[f], acc);
#pragma unroll
        for (int r = 0; r < R; ++r) raw[r] = P.gbt_split_base + (double)acc[r];
    }

    // ---- logistic member (float64) and standardisation (float32, mean split hi+lo so x - mean stays exact-ish) -------
#pragma unroll
    for (int f = 0; f < F; ++f) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (has_lin) zl[r] = fma((double)xv[f][r], P.lin_coef[f], zl[r]);
            zs[f][r] = 0.f;
            if (SVC) {
                zs[f][r] = ((xv[f][r] - P.mean_hi[f]) - P.mean_lo[f]) * P.inv_scale32[f];
                zz32[r] = fmaf(zs[f][r], zs[f][r], zz32[r]);
            }
        }
    }
    }   // full prologue

    // ---- general boosted trees: predicated fixed-depth walk, x from this thread's shared-memory slots ----------------
    if (TREE_MODE == 2) {
        // own rows only: no barrier needed between the s_xt stores above and these loads.  Two trees are walked at a
        // time, so a thread keeps 2 x R independent node chains in flight (the walk is bound by L1 latency).
#pragma unroll
        for (int r = 0; r < R; ++r) raw[r] = P.gbt_init;
        const float* xs = s_xt + tid;
        for (int t0 = 0; t0 < P.n_trees; t0 += TREE_CHUNK) {
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
            const int t1 = min(t0 + TREE_CHUNK, P.n_trees);
            for (int t = t0; t < t1; t += 2) {
                const bool two = t + 1 < t1;
                const int tb = two ? t + 1 : t;
                const unsigned root_a = (unsigned)P.tree_off[t], root_b = (unsigned)P.tree_off[tb];
                const int depth = max((int)P.tree_depth[t], (int)P.tree_depth[tb]);
                unsigned ia[R], ib[R];           // global node indices (children are stored as global indices too)
#pragma unroll
                for (int r = 0; r < R; ++r) { ia[r] = root_a; ib[r] = root_b; }
                for (int lvl = 0; lvl < depth; ++lvl) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const uint2 da = __ldg(P.nodes + ia[r]);
                        const uint2 db = __ldg(P.nodes + ib[r]);
                        const float xa = xs[(da.y >> 24) * BLOCK_ROWS + r * BLOCK_THREADS];
                        const float xb = xs[(db.y >> 24) * BLOCK_ROWS + r * BLOCK_THREADS];
                        ia[r] = (da.y & 0xFFFFFFu) + (xa > __uint_as_float(da.x) ? 1u : 0u);
                        ib[r] = (db.y & 0xFFFFFFu) + (xb > __uint_as_float(db.x) ? 1u : 0u);
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc[r] += __ldg(P.value32 + ia[r]);
                    if (two) acc[r] += __ldg(P.value32 + ib[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) raw[r] += (double)acc[r];
        }
    }

    // Links of the tree and logistic members now, so that only their share of the combiner's logit (one double per
    // row) has to stay in registers across the RBF loop.
    double zpart[R], member[3][R];
    partial_logit_fast<R>(P, TREE_MODE != 0, has_lin, raw, zl, zpart, member);
    if (P.meta) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (row0 + r >= P.n) continue;
            if (P.slot_gbt >= 0) P.meta[(row0 + r) * P.n_members + P.slot_gbt] = member[1][r];
            if (P.slot_lin >= 0) P.meta[(row0 + r) * P.n_members + P.slot_lin] = member[2][r];
        }
    }

    // ---- RBF sum: 4 support vectors x R rows per tile, packed FFMA2 ------------------------------------------------------
    // Each table step is one float4 = the weights of 4 support vectors for one feature (step F: their folded
    // coefficients).  An FFMA2 covers two support vectors: its 64-bit operands are a weight pair (.xy or .zw of the
    // LDS.128 result) and the pair's running exponents, the row's standardised feature enters as a broadcast scalar
    // (SASS ``Rz.F32``).  The four rows that share a weight pair are issued back to back, and the second pair walks
    // the rows in reverse, so consecutive FFMA2s always share one source operand (register-file reuse cache) --
    // measured +4 points of FMA-pipe utilisation over the row-paired form (tools/microbench3.cu).
    // The table streams from shared memory through a 3-deep register ring (LDS.128 two steps ahead of first use).
    double acc64[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc64[r] = 0.0;
    if (SVC) {
        {
            const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_table_ready);
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "TABLE_WAIT_%=:\n"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                "@p bra TABLE_DONE_%=;\n"
                "bra TABLE_WAIT_%=;\n"
                "TABLE_DONE_%=:\n"
                "}\n" ::"r"(mbar) : "memory");
        }
        const float4* tab = s_sv;
        float4 ring[3];
        ring[0] = tab[0]; ring[1] = tab[1];
        if (PFX) {
            // Prefix-shared layout (hfe_model.cuh, build_svc): the table is one stream of float4 steps -- per quad
            // SV_PREFIX prefix-weight steps, then count[q] leaf steps of SV_LEAF_STEPS.  The partial exponents of the
            // prefix stay in p2 and seed every leaf step's FMA chains, so each support vector costs
            // F - SV_PREFIX + 1 FMAs instead of F + 1.
            //
            // A leaf step covers two support-vector pairs, skewed by half a step so that one warp keeps both the FMA
            // and the MUFU pipe busy: while the FMA chains of pair 0 run (phase A), the exponents pair 1 finished in
            // the previous step go through ex2 and their coefficient FMAs; while pair 1's chains run (phase B), the
            // same happens for the pair-0 exponents just finished.  Step layout: [coefficients][pair 0: 4 steps of two
            // features][pair 1: 4 steps of two features].
            static_assert((F - SV_PREFIX) == 2 * R && R == 4, "leaf step: one ex2 row per two-feature table step");
            const unsigned char* quad_steps =
                reinterpret_cast<const unsigned char*>(s_sv) + (size_t)sv_bytes - (((size_t)P.sv_quads + 15) / 16) * 16;
            float2 a2[R], t2[SV_GROUP / 2][R];
            float2 c_pending = make_float2(0.f, 0.f);         // coefficients of the pair-1 exponents waiting in t2[1]
#pragma unroll
            for (int r = 0; r < R; ++r) { a2[r] = make_float2(0.f, 0.f); t2[0][r] = t2[1][r] = make_float2(0.f, 0.f); }
            auto flush = [&]() {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc64[r] += (double)(a2[r].x + a2[r].y);
                    a2[r] = make_float2(0.f, 0.f);
                }
            };
#pragma unroll 1
            for (int q = 0; q < P.sv_quads; ++q) {
                float2 p2[SV_GROUP / 2][R];
#pragma unroll
                for (int st = 0; st < SV_PREFIX; ++st) {
                    ring[(st + 2) % 3] = tab[st + 2];
                    const float4 w4 = ring[st % 3];
                    const float2 wp[SV_GROUP / 2] = {make_float2(w4.x, w4.y), make_float2(w4.z, w4.w)};
#pragma unroll
                    for (int h = 0; h < SV_GROUP / 2; ++h)
#pragma unroll
                        for (int rr = 0; rr < R; ++rr) {
                            const int r = (h & 1) ? R - 1 - rr : rr;
                            const float2 zb = make_float2(zs[st][r], zs[st][r]);
                            p2[h][r] = st == 0 ? __fmul2_rn(wp[h], zb) : __ffma2_rn(wp[h], zb, p2[h][r]);
                        }
                }
                tab += SV_PREFIX;
#pragma unroll 1
                for (int c = quad_steps[q]; c > 0; --c) {
                    ring[2] = tab[2];
                    const float4 cq = ring[0];                // this step's folded coefficients: pair 0 .xy, pair 1 .zw
                    const float2 c_now = make_float2(cq.x, cq.y);
#pragma unroll
                    for (int ph = 0; ph < 2; ++ph) {          // phase A: chains of pair 0; phase B: chains of pair 1
                        const float2 c_ex = ph == 0 ? c_pending : c_now;     // coefficients of the exponents in t2[1 - ph]
#pragma unroll
                        for (int j = 0; j < R; ++j) {
                            const int st = 1 + ph * R + j;
                            ring[(st + 2) % 3] = tab[st + 2];
                            const float4 w4 = ring[st % 3];
                            const float2 wa = make_float2(w4.x, w4.y), wb = make_float2(w4.z, w4.w);
                            const int fa = SV_PREFIX + 2 * j, fb = fa + 1;
#pragma unroll
                            for (int r = 0; r < R; ++r)
                                t2[ph][r] = __ffma2_rn(wa, make_float2(zs[fa][r], zs[fa][r]), j == 0 ? p2[ph][r] : t2[ph][r]);
                            t2[1 - ph][j].x = ex2_approx(t2[1 - ph][j].x);
                            t2[1 - ph][j].y = ex2_approx(t2[1 - ph][j].y);
#pragma unroll
                            for (int r = 0; r < R; ++r)      // same row order: dependent FMAs stay 4 instructions apart
                                t2[ph][r] = __ffma2_rn(wb, make_float2(zs[fb][r], zs[fb][r]), t2[ph][r]);
                            if (j > 0) a2[j - 1] = __ffma2_rn(t2[1 - ph][j - 1], c_ex, a2[j - 1]);
                        }
                        a2[R - 1] = __ffma2_rn(t2[1 - ph][R - 1], c_ex, a2[R - 1]);
                        // after phase A, a2 holds the four vectors of the previous leaf step: float32 partial sums never
                        // span more than one step (HFE_PFX_FLUSH 2: two steps) before they go to the float64 accumulator
                        if (ph == 0 && (HFE_PFX_FLUSH == 1 || (HFE_PFX_FLUSH == 2 && (c & 1)))) flush();
                    }
                    c_pending = make_float2(cq.z, cq.w);
                    tab += SV_LEAF_STEPS;
                }
            }
            // drain: the last step's pair-1 exponents
#pragma unroll
            for (int r = 0; r < R; ++r) {
                t2[1][r].x = ex2_approx(t2[1][r].x);
                t2[1][r].y = ex2_approx(t2[1][r].y);
                a2[r] = __ffma2_rn(t2[1][r], c_pending, a2[r]);
            }
            flush();
        } else
#pragma unroll 1
        for (int g = 0; g < P.n_groups; ++g) {
            float2 t2[SV_GROUP / 2][R];                   // [support-vector pair][row]
#pragma unroll
            for (int st = 0; st < SV_STEPS; ++st) {
                ring[(st + 2) % 3] = tab[st + 2];         // SV_STEPS % 3 == 0: same ring slot in every group
                const float4 w4 = ring[st % 3];
                const float2 wp[SV_GROUP / 2] = {make_float2(w4.x, w4.y), make_float2(w4.z, w4.w)};
                if (st == 0) {
#pragma unroll
                    for (int h = 0; h < SV_GROUP / 2; ++h)
#pragma unroll
                        for (int r = 0; r < R; ++r) t2[h][r] = __fmul2_rn(wp[h], make_float2(zs[0][r], zs[0][r]));
                } else if (st < F) {
#pragma unroll
                    for (int h = 0; h < SV_GROUP / 2; ++h)
#pragma unroll
                        for (int rr = 0; rr < R; ++rr) {
                            const int r = (h & 1) ? R - 1 - rr : rr;
                            t2[h][r] = __ffma2_rn(wp[h], make_float2(zs[st][r], zs[st][r]), t2[h][r]);
                        }
                } else {
                    // exp2 of the 4 x R exponents, weighted by the folded coefficients (wp holds them in this step);
                    // the float32 partial sum of these 4 vectors goes straight into the float64 accumulator
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float2 a2 = make_float2(0.f, 0.f);
#pragma unroll
                        for (int h = 0; h < SV_GROUP / 2; ++h) {
                            float2 e = make_float2(ex2_approx(t2[h][r].x), ex2_approx(t2[h][r].y));
                            a2 = __ffma2_rn(e, wp[h], a2);
                        }
                        acc64[r] += (double)(a2.x + a2.y);
                    }
                }
            }
            tab += SV_STEPS;
        }
    }

    // ---- float64 epilogue: Platt, clamp, coupling, combiner (R rows interleaved) -------------------------------------------
    double f_svc[R], p1[R], zfin[R];
    bool unc[R];
    unsigned fixmask = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        f_svc[r] = 0.0;
        if (SVC) {
            double gzz = P.gamma * (double)zz32[r];                     // gamma |z|^2
            f_svc[r] = fma(acc64[r], exp_neg_fast(-gzz), P.svc_intercept);
            if (!(gzz <= P.zz_limit)) fixmask |= 1u << r;               // float32 2^x range guard (or NaN)
        }
    }
    finish_rows_fast<R>(P, SVC, f_svc, zpart, p1, zfin, member[0], unc);
    if (SVC && P.svc_dec) {                                             // creation-time guard probe only
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (row0 + r < P.n) P.svc_dec[row0 + r] = f_svc[r];
    }
    emit_rows<R, double>(P, row0, lane, full, nonfinite, fixmask, p1, zfin, member[0], unc);
    }   // tile loop
}

// ---- the streaming pass: stacks without the SVC member (HBM-bound) -------------------------------------------------------
// Persistent warps, each with its own feature-tile buffer(s) in shared memory.  One lane asks the copy engine for the
// warp's next 128-row x 17-column tile (cp.async.bulk.tensor.2d + mbarrier complete_tx; SASS UTMALDG) as soon as every
// lane has read the current one, so no registers are spent on loads in flight and 24 warps fit on an SM; the links and
// the combiner run in float32.  Measured on B200, 10 M rows: 72 G rows/s = 84 % of the HBM copy bandwidth (ncu: the kernel alone
// moves 6.0 TB/s); with 17 one-dimensional bulk copies per tile instead of the tensor copy 56 G rows/s (a fifth of all
// instructions went into issuing them); the register-window version of hfe_score_kernel<false, ..> 47 G rows/s.  The
// hot loop is kept small (1.5 k SASS instructions): rare work (permuted column, further splits of a feature) is
// rolled up and reads the tile with a run-time feature index, which only the shared-memory copy allows.
#ifndef HFE_STREAM_STAGES
#define HFE_STREAM_STAGES 1               // measured: 1 stage x 24 warps/SM 51 G rows/s, 2 stages x 12 warps/SM 38 G rows/s
#endif
constexpr int STREAM_STAGES = HFE_STREAM_STAGES;
constexpr int STREAM_TILE_FLOATS = F * WARP_ROWS;
constexpr size_t STREAM_SMEM_BYTES = (size_t)BLOCK_WARPS * STREAM_STAGES * STREAM_TILE_FLOATS * sizeof(float);

template <int TREE_MODE>
__global__ void __launch_bounds__(BLOCK_THREADS, HFE_MIN_BLOCKS_STREAM)
hfe_stream_kernel(const __grid_constant__ ScoreParams P, const __grid_constant__ CUtensorMap x_map) {
    static_assert(TREE_MODE == 0 || TREE_MODE == 1, "general trees keep the x tile of hfe_score_kernel");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) unsigned long long s_full[BLOCK_WARPS][STREAM_STAGES];
    constexpr int R = ROWS_PER_THREAD;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* tiles = reinterpret_cast<float*>(smem_raw) + (size_t)warp * STREAM_STAGES * STREAM_TILE_FLOATS;
    const int n_tiles = (int)((P.n + WARP_ROWS - 1) / WARP_ROWS);
    const int stride = gridDim.x * BLOCK_WARPS;
    const int first = blockIdx.x * BLOCK_WARPS + warp;
    const bool has_lin = P.slot_lin >= 0;

    // one lane asks for a whole tile with a single 2-D tensor copy (SASS UTMALDG): box = 128 rows x 17 columns of the
    // [17][ld] matrix, rows at or beyond n are filled with zeros by the copy engine
    auto issue = [&](int tile, int stage) {
        const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_full[warp][stage]);
        const unsigned dst = (unsigned)__cvta_generic_to_shared(tiles + stage * STREAM_TILE_FLOATS);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                     ::"r"(mbar), "r"((unsigned)(STREAM_TILE_FLOATS * sizeof(float))) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(dst), "l"(&x_map), "r"(tile * WARP_ROWS), "r"(0), "r"(mbar)
                     : "memory");
    };
    if (lane == 0) {
#pragma unroll
        for (int st = 0; st < STREAM_STAGES; ++st)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;"
                         ::"r"((unsigned)__cvta_generic_to_shared(&s_full[warp][st])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
        for (int st = 0; st < STREAM_STAGES; ++st)
            if (first + st * stride < n_tiles) issue(first + st * stride, st);
    }
    __syncwarp();

    int k = 0;
    for (int tile = first; tile < n_tiles; tile += stride, ++k) {
        const int stage = k % STREAM_STAGES;
        const unsigned parity = (unsigned)(k / STREAM_STAGES) & 1u;
        {
            const unsigned mbar = (unsigned)__cvta_generic_to_shared(&s_full[warp][stage]);
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "TILE_WAIT_%=:\n"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                "@p bra TILE_DONE_%=;\n"
                "bra TILE_WAIT_%=;\n"
                "TILE_DONE_%=:\n"
                "}\n" ::"r"(mbar), "r"(parity) : "memory");
        }
        float4* buf = reinterpret_cast<float4*>(tiles + stage * STREAM_TILE_FLOATS) + lane;   // this lane's 4 rows
        const long long row0 = (long long)tile * WARP_ROWS + lane * R;
        const bool full = (row0 + R <= P.n);
        if (P.perm_col >= 0) {
            // rare: a permutation-importance pass -- the permuted column is gathered through perm_idx into the tile
            // (own slots only, so no synchronisation); the ragged tail needs nothing, those rows arrive as zeros
            const long long row_ld = row0 < P.n ? row0 : 0;
            buf[P.perm_col * (WARP_ROWS / 4)] =
                fix_rows4(P, P.perm_col, row0, row_ld, full, make_float4(0.f, 0.f, 0.f, 0.f));
        }
        double raw[R], zl[R];
        float finite_probe = 0.f, acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) { raw[r] = 0.0; zl[r] = P.lin_intercept; acc[r] = 0.f; }
#pragma unroll
        for (int f = 0; f < F; ++f) {
            const float4 v = buf[f * (WARP_ROWS / 4)];
            const float xf[R] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (has_lin) zl[r] = fma((double)xf[r], P.lin_coef[f], zl[r]);   // NaN/Inf in x makes zl non-finite
                else finite_probe = fmaf(xf[r], 0.f, finite_probe);
            }
            if (TREE_MODE == 1) {                       // first merged split of this feature: immediate operands
                const float thr0 = P.split1_thr[f], dl0 = P.split1_delta[f];
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (xf[r] <= thr0) acc[r] += dl0;
            }
        }
        if (TREE_MODE == 1) {                           // the remaining splits, any feature: one rolled loop over the tile
#pragma unroll 1
            for (int s = 0; s < P.n_extra_splits; ++s) {
                const float4 v = buf[P.split_feat[s] * (WARP_ROWS / 4)];
                const float thr = P.split_thr[s], dl = P.split_delta[s];
                if (v.x <= thr) acc[0] += dl;
                if (v.y <= thr) acc[1] += dl;
                if (v.z <= thr) acc[2] += dl;
                if (v.w <= thr) acc[3] += dl;
            }
        }
        // every lane has read its part of this stage: hand the buffer back to the copy engine for the tile after next
#ifdef HFE_BOUNDS_CHECK
        // checked build: the tile is poisoned once every lane has read it and before the copy engine may refill it, so a
        // read that runs ahead of the next tile's arrival (a wrong parity wait) sees NaN and trips the non-finite flag
        __syncwarp();
#pragma unroll
        for (int f = 0; f < F; ++f) buf[f * (WARP_ROWS / 4)] = make_float4(__int_as_float(0x7FC00000), __int_as_float(0x7FC00000),
                                                                         __int_as_float(0x7FC00000), __int_as_float(0x7FC00000));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
        if (P.perm_col >= 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // our stores before the engine's
        __syncwarp();
        if (lane == 0 && tile + STREAM_STAGES * stride < n_tiles) issue(tile + STREAM_STAGES * stride, stage);

        bool nonfinite = !(finite_probe == 0.f);
        if (has_lin) {
#pragma unroll
            for (int r = 0; r < R; ++r) nonfinite = nonfinite || !(fabs(zl[r]) <= 1.7976931348623157e308);
        }
        if (TREE_MODE == 1) {
#pragma unroll
            for (int r = 0; r < R; ++r) raw[r] = P.gbt_split_base + (double)acc[r];
        }
        // Links and combiner in float32 (MUFU.EX2 / MUFU.RCP + one Newton step): with no SVC member the pass would
        // otherwise spend most of its instructions on three float64 exponentials per row.  |error| of a link
        // <= 1.8e-7, of the final probability <= ~4e-7 (bar 1e-6); rows whose |z| is within the (widened) label
        // guard go to the float64 fix-up like everywhere else.
        float p1[R], zfin[R], unused_svc[R];
        bool unc[R];
        const float b32 = (float)P.final_intercept;
        const float wg32 = TREE_MODE != 0 ? (float)P.final_coef[TREE_MODE != 0 ? max(P.slot_gbt, 0) : 0] : 0.f;
        const float wl32 = has_lin ? (float)P.final_coef[max(P.slot_lin, 0)] : 0.f;
        const float guard32 = (float)P.label_guard;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float pg = TREE_MODE != 0 ? expit32((float)raw[r]) : 0.f;          // SK/_loss/loss.py:980-998
            const float pl = has_lin ? expit32((float)zl[r]) : 0.f;                   // SK/linear_model/_base.py:429-451
            float z = b32;
            // meta-features are combined in estimators_ order (SK/ensemble/_stacking.py:87-135)
            if (TREE_MODE != 0 && has_lin && P.slot_lin < P.slot_gbt) { z = fmaf(wl32, pl, z); z = fmaf(wg32, pg, z); }
            else { if (TREE_MODE != 0) z = fmaf(wg32, pg, z); if (has_lin) z = fmaf(wl32, pl, z); }
            zfin[r] = z;
            p1[r] = expit32(z);
            unc[r] = !(fabsf(z) >= guard32);
            unused_svc[r] = 0.f;
            if (P.meta && row0 + r < P.n) {
                if (P.slot_gbt >= 0) P.meta[(row0 + r) * P.n_members + P.slot_gbt] = (double)pg;
                if (P.slot_lin >= 0) P.meta[(row0 + r) * P.n_members + P.slot_lin] = (double)pl;
            }
        }
        emit_rows<R, float>(P, row0, lane, full, nonfinite, 0u, p1, zfin, unused_svc, unc);
    }
}

// Second (tiny) launch of every scoring call: one warp per work-list entry re-evaluates that row in float64 and
// overwrites the fused kernel's outputs for it.
__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_fixup_kernel(const __grid_constant__ ScoreParams P) {
    const unsigned count = min(P.status[HFE_STATUS_LIST_LEN], (unsigned)HFE_WORKLIST_CAPACITY);
    HFE_ASSERT_IN(P.status[3] == P.status[HFE_STATUS_LIST_LEN]);        // every append had landed before this launch began
    const unsigned warp = (blockIdx.x * BLOCK_THREADS + threadIdx.x) >> 5;
    const unsigned n_warps = (gridDim.x * BLOCK_THREADS) >> 5;
    unsigned hits = 0;
    for (unsigned i = warp; i < count; i += n_warps) {
        bool hit = eval_row_f64_warp(P, (long long)P.status[HFE_STATUS_HEADER_WORDS + i]);
        hits += hit ? 1u : 0u;
    }
    if ((threadIdx.x & 31) == 0) {
        if (hits && P.score) atomicAdd(P.score, (unsigned long long)hits);
        if (warp == 0) P.status[1] = count;
    }
}

// A handful of host rows (hfe_predict_host, n <= TINY_HOST_ROWS): one warp per row on the float64 path, reading the
// row-major float64 input from mapped pinned host memory and writing the outputs back into it.
__global__ void __launch_bounds__(32)
hfe_tiny_kernel(const __grid_constant__ ScoreParams P, const double* __restrict__ x_rowmajor, unsigned* status) {
    eval_row_f64_warp(P, (long long)blockIdx.x, x_rowmajor, status);
}

// Every row through the float64 path (calibration / test aid; also what hfe_predict_proba_f64 runs).
__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_score_f64_kernel(const __grid_constant__ ScoreParams P) {
    const long long warp = ((long long)blockIdx.x * BLOCK_THREADS + threadIdx.x) >> 5;
    const long long n_warps = ((long long)gridDim.x * BLOCK_THREADS) >> 5;
    for (long long row = warp; row < P.n; row += n_warps) eval_row_f64_warp(P, row);
}

// ---- GradientBoostingClassifier.apply ------------------------------------------------------------------------------------

__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_apply_kernel(const __grid_constant__ ScoreParams P, unsigned short* leaves) {
    long long row = (long long)blockIdx.x * BLOCK_THREADS + threadIdx.x;
    if (row >= P.n) return;
    float xr[F];
#pragma unroll
    for (int f = 0; f < F; ++f) xr[f] = __ldg(P.x + P.col_off[f] + row);
    for (int t = 0; t < P.n_trees; ++t) {
        int leaf = walk_tree(P, t, [&](unsigned f) {
            float v = xr[0];
#pragma unroll
            for (int k = 1; k < F; ++k) v = (f == (unsigned)k) ? xr[k] : v;
            return v;
        });
        leaves[row * P.n_trees + t] = P.sk_id[leaf];
    }
}

// ---- synthetic rows --------------------------------------------------------------------------------------------------------

__device__ __forceinline__ unsigned int hfe_synth_hash(unsigned long long seed, unsigned long long row, int col_id) {
    unsigned long long x = (row << 8) | (unsigned long long)(col_id & 0xFF);
    x ^= seed * 0x9E3779B97F4A7C15ULL;
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    x ^= x >> 31;
    return (unsigned int)(x >> 32);
}

__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_synth_kernel(unsigned long long seed, long long row0, long long n, int n_cols, const int* col_id,
                 const float* base, const float* step, const int* thr_off, const unsigned int* thr, float* x,
                 long long ld) {
    for (long long i = (long long)blockIdx.x * BLOCK_THREADS + threadIdx.x; i < n;
         i += (long long)gridDim.x * BLOCK_THREADS) {
        for (int c = 0; c < n_cols; ++c) {
            unsigned int u = hfe_synth_hash(seed, (unsigned long long)(row0 + i), col_id[c]);
            int cnt = 0;
            for (int k = thr_off[c]; k < thr_off[c + 1]; ++k) cnt += (u >= __ldg(thr + k)) ? 1 : 0;
            x[(long long)c * ld + i] = __fadd_rn(base[c], __fmul_rn(step[c], (float)cnt));
        }
    }
}

// ---- ingest: row-major float64 -> column-major float32 -----------------------------------------------------------------------

template <typename T>
__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_ingest_kernel(const T* __restrict__ src, long long n, int nf, float* __restrict__ dst, long long ld,
                  unsigned int* status) {
    extern __shared__ float s_tile[];     // [BLOCK_THREADS][nf + 1]
    const long long r0 = (long long)blockIdx.x * BLOCK_THREADS;
    const int rows = (int)min((long long)BLOCK_THREADS, n - r0);
    const int pitch = nf + 1;
    bool bad = false;
    for (int i = threadIdx.x; i < rows * nf; i += BLOCK_THREADS) {
        T v = src[r0 * nf + i];
        float fv = (float)(double)v;               // via float64, like validate_data + astype(float32) (matters for int64)
        bad = bad || !(fabsf(fv) <= 3.402823466e38f);   // NaN/Inf, or a float64 too large for float32
        s_tile[(i / nf) * pitch + (i % nf)] = fv;
    }
    __syncthreads();
    for (int f = 0; f < nf; ++f)
        if ((int)threadIdx.x < rows) dst[(long long)f * ld + r0 + threadIdx.x] = s_tile[threadIdx.x * pitch + f];
    if (bad) atomicOr(status, HFE_ROW_NONFINITE);
}

// ---- 1-nearest-neighbour imputation (train_ensemble_public.py:37-40: KNNImputer(n_neighbors=1)) ---------------------------
// One block per receiver row (rows without a missing value return at once).  Phase 1: threads stride over the fit rows
// and compute the nan-euclidean distance sqrt(F / present * sum_present (x - y)^2) in float64, in the operation order
// of SK/metrics/pairwise.py (NaN = no shared coordinate), into shared memory.  Phase 2, per missing column: block-wide
// arg-min over the fit rows that have that column, ties to the lowest row index; the column mean when no distance is
// defined.  Everything is float64 and row-major, as scikit-learn holds it; this is training-time preprocessing on a few
// hundred fit rows, so it is written for clarity, not for a roofline.
constexpr int KNN_THREADS = 128;

__global__ void __launch_bounds__(KNN_THREADS)
hfe_knn1_kernel(const double* __restrict__ fit, long long n_fit, int nf, const double* __restrict__ col_mean,
                double* __restrict__ x, long long n) {
    extern __shared__ double s_knn[];                 // [nf] the receiver row, then [n_fit] distances
    double* s_row = s_knn;
    double* s_dist = s_knn + nf;
    __shared__ double s_best[KNN_THREADS / 32];
    __shared__ long long s_arg[KNN_THREADS / 32];
    __shared__ int s_any;
    const int tid = threadIdx.x;
    for (long long row = blockIdx.x; row < n; row += gridDim.x) {
        if (tid == 0) s_any = 0;
        __syncthreads();
        bool miss = false;
        for (int f = tid; f < nf; f += KNN_THREADS) {
            const double v = x[row * nf + f];
            s_row[f] = v;
            miss = miss || (v != v && col_mean[f] == col_mean[f]);      // columns that are all-missing in fit are not imputed
        }
        if (miss) s_any = 1;
        __syncthreads();
        const int any = s_any;
        __syncthreads();                               // everyone has read the flag before the next row resets it
        if (!any) continue;
        for (long long j = tid; j < n_fit; j += KNN_THREADS) {
            const double* y = fit + j * nf;
            double sum = 0.0;
            int present = 0;
            for (int f = 0; f < nf; ++f) {
                const double a = s_row[f], b = y[f];
                if (a == a && b == b) {
                    const double d = a - b;
                    sum = __dadd_rn(sum, __dmul_rn(d, d));
                    ++present;
                }
            }
            s_dist[j] = present ? __dsqrt_rn(__dmul_rn(__ddiv_rn(sum, (double)present), (double)nf)) : nan("");
        }
        __syncthreads();
        for (int c = 0; c < nf; ++c) {
            const double v = s_row[c];
            if (v == v || !(col_mean[c] == col_mean[c])) continue;      // uniform across the block
            double best = INFINITY;
            long long arg = -1;
            for (long long j = tid; j < n_fit; j += KNN_THREADS) {
                const double d = s_dist[j], yc = fit[j * nf + c];
                if (yc == yc && d == d && (d < best || (d == best && (arg < 0 || j < arg)))) { best = d; arg = j; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_down_sync(0xFFFFFFFFu, best, o);
                const long long oa = __shfl_down_sync(0xFFFFFFFFu, arg, o);
                if (oa >= 0 && (arg < 0 || ob < best || (ob == best && oa < arg))) { best = ob; arg = oa; }
            }
            if ((tid & 31) == 0) { s_best[tid >> 5] = best; s_arg[tid >> 5] = arg; }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < KNN_THREADS / 32; ++w)
                    if (s_arg[w] >= 0 && (arg < 0 || s_best[w] < best || (s_best[w] == best && s_arg[w] < arg))) {
                        best = s_best[w];
                        arg = s_arg[w];
                    }
                x[row * nf + c] = arg >= 0 ? fit[arg * nf + c] : col_mean[c];
            }
            __syncthreads();
        }
    }
}

// ---- evaluation counts (what train_ensemble_public.py:63-64,67,79 computes next from the probabilities) ---------------------
// confusion[0..3] = tn, fp, fn, tp at `threshold` (the script's ``y_pred[:,1] > 0.5``); hist[c*NB + b] counts rows of true
// class c whose P(class 1) falls in bin b = floor(p * (NB-1) + 0.5): ROC / precision-recall curves and their areas follow
// from the two histograms on the host.
constexpr int METRIC_BINS = 65536;

__global__ void __launch_bounds__(256)
hfe_metrics_kernel(const float* __restrict__ proba, const unsigned char* __restrict__ y, long long n, float threshold,
                   unsigned long long* confusion, unsigned int* hist) {
    unsigned cnt[4] = {0, 0, 0, 0};
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float p = __ldg(proba + 2 * i + 1);
        int c = y[i] ? 1 : 0;
        int pred = p > threshold ? 1 : 0;
        cnt[c * 2 + pred] += 1;
        if (hist) {
            int b = (int)fminf(fmaxf(p * (float)(METRIC_BINS - 1) + 0.5f, 0.f), (float)(METRIC_BINS - 1));
            atomicAdd(hist + c * METRIC_BINS + b, 1u);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        unsigned v = cnt[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(confusion + k, (unsigned long long)v);
    }
}

}  // namespace hfe
