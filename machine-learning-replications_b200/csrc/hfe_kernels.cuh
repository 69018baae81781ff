// hfe_kernels.cuh -- sm_100a kernels of the scoring path.
//
// hfe_score_kernel<SVC, TREE_MODE> is the fused pass: one thread owns ROWS_PER_THREAD (=4) consecutive rows, fetched
// from the column-major float32 matrix with one 128-bit load per feature (a warp reads 512 contiguous bytes of each
// column).  Per row it evaluates, without leaving registers,
//   a6/a7  StandardScaler + RBF-SVC decision + Platt + clamp + 2-class coupling   (svm.cpp:461-515, 2821-2958, 2046-2107)
//   a8     boosted trees (merged stump table or general node tables) + expit        (_gradient_boosting.pyx:25-73)
//   a10    logistic member on the raw features                                       (linear_model/_base.py:366-451)
//   a11    combiner + hard label (z > 0)                                             (linear_model/_base.py:398-451)
// and writes [1-p, p] (8 bytes per row).  The RBF sum -- 434 x 17 multiply-adds + 434 exp2 per row, ~98 % of the
// reference's CPU time -- runs in float32 on the FMA / MUFU pipes out of a shared-memory support-vector table read
// with warp-broadcast LDS.128; everything after it runs in float64 on B200's full-rate FP64 pipe.
// Rows whose result could differ from a float64 evaluation by a branch (coupling-solver exit test, label sign) are
// collected per block and re-evaluated cooperatively in float64 (eval_row_f64) before the block retires.
#pragma once

#include "hfe_model.cuh"

namespace hfe {

constexpr int BLOCK_THREADS = 256;
constexpr int ROWS_PER_THREAD = 4;
constexpr int BLOCK_ROWS = BLOCK_THREADS * ROWS_PER_THREAD;
constexpr int MAX_SMEM_GROUPS = 160;      // support-vector groups staged per pass (160*36*16 B = 90 KB)
constexpr int TREE_CHUNK = 16;            // float32 partial sums over trees are folded into float64 every 16 trees
#ifndef HFE_MIN_BLOCKS
#define HFE_MIN_BLOCKS 2                  // resident blocks per SM the fused kernel is compiled for (register cap)
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
    long long src = (f == P.perm_col) ? (long long)P.perm_idx[row] : row;
    return __ldg(P.x + (long long)f * P.ld + src);
}

// Walk one tree for one row: x values come from `xrow` (F floats, any address space).  Returns the re-laid leaf index
// (global node index).
template <typename XGet>
__device__ __forceinline__ int walk_tree(const ScoreParams& P, int t, XGet xget) {
    int base = P.tree_off[t];
    int depth = P.tree_depth[t];
    int node = 0;
    for (int lvl = 0; lvl < depth; ++lvl) {
        uint2 nd = __ldg(P.nodes + base + node);
        float thr = __uint_as_float(nd.x);
        float xv = xget(nd.y >> 24);
        node = (int)(nd.y & 0xFFFFFFu) + (xv > thr ? 1 : 0);
    }
    return base + node;
}

// Cooperative float64 evaluation of one row by the whole block -- the exact-arithmetic path.  Follows the reference's
// operation order within each term (scale: SK/preprocessing/_data.py:1132-1134; kernel: svm.cpp:461-515); only the
// order in which the n_sv kernel values / the trees are summed differs (strided partials + fixed-shape reduction).
__device__ __noinline__ void eval_row_f64(const ScoreParams& P, long long row, double* s_x, double* s_z, double* s_red,
                             unsigned int* s_hits) {
    const int tid = threadIdx.x;
    const bool has_svc = P.slot_svc >= 0, has_gbt = P.slot_gbt >= 0, has_lin = P.slot_lin >= 0;
    if (tid < F) {
        float xv = load_feature(P, tid, row);
        s_x[tid] = (double)xv;
        s_z[tid] = ((double)xv - P.mean[tid]) / P.scale[tid];
    }
    __syncthreads();
    double part_svc = 0.0, part_gbt = 0.0;
    if (has_svc) {
        for (int j = tid; j < P.n_sv; j += BLOCK_THREADS) {
            const double* s = P.sv64 + (long long)j * F;
            double sum = 0.0;
#pragma unroll
            for (int d = 0; d < F; ++d) {
                double diff = s_z[d] - s[d];
                sum += diff * diff;
            }
            part_svc += P.dual64[j] * exp(-P.gamma * sum);
        }
    }
    if (has_gbt) {
        for (int t = tid; t < P.n_trees; t += BLOCK_THREADS) {
            int leaf = walk_tree(P, t, [&](unsigned f) { return (float)s_x[f]; });
            part_gbt += P.value64[leaf];
        }
    }
    // fixed-shape reduction: lanes by shuffle, then warps in index order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        part_svc += __shfl_down_sync(0xFFFFFFFFu, part_svc, o);
        part_gbt += __shfl_down_sync(0xFFFFFFFFu, part_gbt, o);
    }
    if ((tid & 31) == 0) {
        s_red[(tid >> 5) * 2] = part_svc;
        s_red[(tid >> 5) * 2 + 1] = part_gbt;
    }
    __syncthreads();
    if (tid == 0) {
        double f_svc = 0.0, raw = 0.0, zl = P.lin_intercept;
        for (int w = 0; w < BLOCK_THREADS / 32; ++w) {
            f_svc += s_red[2 * w];
            raw += s_red[2 * w + 1];
        }
        f_svc += P.svc_intercept;
        raw += P.gbt_init;
        if (has_lin) {
            zl = 0.0;
            for (int d = 0; d < F; ++d) zl += s_x[d] * P.lin_coef[d];
            zl += P.lin_intercept;
        }
        RowResult r = finish_row(P, has_svc, has_gbt, has_lin, f_svc, raw, zl);
        bool finite = true;
        for (int d = 0; d < F; ++d) finite = finite && (fabs(s_x[d]) <= 3.4028234663852886e38);
        if (!finite) {
            r.p1 = r.z = __longlong_as_double(0x7FF8000000000000LL);
            r.member[0] = r.member[1] = r.member[2] = r.p1;
        }
        store_row(P, row, r);
        if (s_hits && P.y && P.score) *s_hits += ((r.z > 0 ? 1 : 0) == P.y[row]) ? 1u : 0u;
    }
    __syncthreads();
}

// ---- the fused pass ----------------------------------------------------------------------------------------------------

template <bool SVC, int TREE_MODE>
__global__ void __launch_bounds__(BLOCK_THREADS, HFE_MIN_BLOCKS)
hfe_score_kernel(const __grid_constant__ ScoreParams P, int smem_groups) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* s_sv = reinterpret_cast<float4*>(smem_raw);                       // [smem_groups][SV_ROWS]
    float* s_xt = reinterpret_cast<float*>(smem_raw) + (SVC ? ((size_t)smem_groups * SV_ROWS + SV_PAD_ROWS) * 4 : 0);  // [F][BLOCK_ROWS]
    __shared__ double s_x[F], s_z[F], s_red[2 * BLOCK_THREADS / 32];
    __shared__ unsigned int s_nfix, s_nbad, s_hits;
    __shared__ unsigned short s_fix[BLOCK_ROWS];

    const int tid = threadIdx.x;
    const long long row0 = ((long long)blockIdx.x * BLOCK_THREADS + tid) * ROWS_PER_THREAD;
    const bool has_lin = P.slot_lin >= 0;
    constexpr int R = ROWS_PER_THREAD;
    if (tid == 0) { s_nfix = 0; s_nbad = 0; s_hits = 0; }

    // ---- load 4 rows x 17 features (128-bit loads; guarded scalar loads on the ragged tail / permuted column) ------
    float xv[F][R];
    const bool full = (row0 + R <= P.n);
#pragma unroll
    for (int f = 0; f < F; ++f) {
        if (full && f != P.perm_col) {
            float4 v = __ldg(reinterpret_cast<const float4*>(P.x + (long long)f * P.ld + row0));
            xv[f][0] = v.x; xv[f][1] = v.y; xv[f][2] = v.z; xv[f][3] = v.w;
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) xv[f][r] = (row0 + r < P.n) ? load_feature(P, f, row0 + r) : 0.0f;
        }
    }
    bool nonfinite = false;
#pragma unroll
    for (int f = 0; f < F; ++f)
#pragma unroll
        for (int r = 0; r < R; ++r) nonfinite = nonfinite || !(fabsf(xv[f][r]) <= 3.402823466e38f);

    if (TREE_MODE == 2) {
#pragma unroll
        for (int f = 0; f < F; ++f)
#pragma unroll
            for (int r = 0; r < R; ++r) s_xt[f * BLOCK_ROWS + r * BLOCK_THREADS + tid] = xv[f][r];
    }

    // ---- boosted trees ---------------------------------------------------------------------------------------------
    double raw[R];
#pragma unroll
    for (int r = 0; r < R; ++r) raw[r] = 0.0;
    if (TREE_MODE == 1) {
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll
        for (int f = 0; f < F; ++f) {
            for (int s = P.split_begin[f]; s < P.split_begin[f + 1]; ++s) {
                float thr = P.split_thr[s], dl = P.split_delta[s];
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] += (xv[f][r] <= thr) ? dl : 0.f;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) raw[r] = P.gbt_split_base + (double)acc[r];
    } else if (TREE_MODE == 2) {
        // own rows only: no barrier needed between the s_xt stores above and these loads
#pragma unroll
        for (int r = 0; r < R; ++r) raw[r] = P.gbt_init;
        for (int t0 = 0; t0 < P.n_trees; t0 += TREE_CHUNK) {
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
            int t1 = min(t0 + TREE_CHUNK, P.n_trees);
            for (int t = t0; t < t1; ++t) {
                int base = P.tree_off[t];
                int depth = P.tree_depth[t];
                int node[R];
#pragma unroll
                for (int r = 0; r < R; ++r) node[r] = 0;
                for (int lvl = 0; lvl < depth; ++lvl) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        uint2 nd = __ldg(P.nodes + base + node[r]);
                        float x = s_xt[(nd.y >> 24) * BLOCK_ROWS + r * BLOCK_THREADS + tid];
                        node[r] = (int)(nd.y & 0xFFFFFFu) + (x > __uint_as_float(nd.x) ? 1 : 0);
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] += __ldg(P.value32 + base + node[r]);
            }
#pragma unroll
            for (int r = 0; r < R; ++r) raw[r] += (double)acc[r];
        }
    }

    // ---- logistic member + standardisation (float64), z rounded once to float32 for the RBF sum -----------------------
    // z2[f][p] packs rows (2p, 2p+1) of feature f into one 64-bit register pair for FFMA2.
    double zl[R], zz[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { zl[r] = P.lin_intercept; zz[r] = 0.0; }
    float2 z2[F][R / 2];
#pragma unroll
    for (int f = 0; f < F; ++f) {
        float zf[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            double xd = (double)xv[f][r];
            if (has_lin) zl[r] = fma(xd, P.lin_coef[f], zl[r]);
            zf[r] = 0.f;
            if (SVC) {
                double zd = (xd - P.mean[f]) * P.inv_scale[f];
                zz[r] = fma(zd, zd, zz[r]);
                zf[r] = (float)zd;
            }
        }
#pragma unroll
        for (int p = 0; p < R / 2; ++p) z2[f][p] = make_float2(zf[2 * p], zf[2 * p + 1]);
    }

    // ---- RBF sum: 4 support vectors x 4 rows per tile, packed FFMA2, table streamed from shared memory ------------------
    // The table is a linear stream of steps (two float4 each); a 3-deep register ring keeps the LDS.128 two steps
    // ahead of their first use.
    double acc64[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc64[r] = 0.0;
    if (SVC) {
        for (int g0 = 0; g0 < P.n_groups; g0 += smem_groups) {
            int ng = min(smem_groups, P.n_groups - g0);
            __syncthreads();
            for (int i = tid; i < ng * SV_ROWS + SV_PAD_ROWS; i += BLOCK_THREADS)
                s_sv[i] = (i < ng * SV_ROWS) ? __ldg(P.sv_tab + (size_t)g0 * SV_ROWS + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            __syncthreads();
            const float4* tab = s_sv;
            float4 ring_a[3], ring_b[3];
            ring_a[0] = tab[0]; ring_b[0] = tab[1];
            ring_a[1] = tab[2]; ring_b[1] = tab[3];
#pragma unroll 1
            for (int g = 0; g < ng; g += SV_CHUNK / SV_GROUP) {
                float2 acc2[R / 2];
#pragma unroll
                for (int p = 0; p < R / 2; ++p) acc2[p] = make_float2(0.f, 0.f);
#pragma unroll
                for (int h = 0; h < SV_CHUNK / SV_GROUP; ++h) {
                    float2 t2[SV_GROUP][R / 2];
#pragma unroll
                    for (int st = 0; st < SV_STEPS; ++st) {
                        // SV_STEPS % 3 == 0, so the ring slot of a step is the same in every group
                        ring_a[(st + 2) % 3] = tab[2 * (st + 2)];
                        ring_b[(st + 2) % 3] = tab[2 * (st + 2) + 1];
                        const float4 a = ring_a[st % 3], b = ring_b[st % 3];
                        const float2 w[SV_GROUP] = {make_float2(a.x, a.y), make_float2(a.z, a.w),
                                                    make_float2(b.x, b.y), make_float2(b.z, b.w)};
                        if (st == 0) {
#pragma unroll
                            for (int q = 0; q < SV_GROUP; ++q)
#pragma unroll
                                for (int p = 0; p < R / 2; ++p) t2[q][p] = __fmul2_rn(z2[0][p], w[q]);
                        } else if (st < F) {
#pragma unroll
                            for (int q = 0; q < SV_GROUP; ++q)
#pragma unroll
                                for (int p = 0; p < R / 2; ++p) t2[q][p] = __ffma2_rn(z2[st][p], w[q], t2[q][p]);
                        } else {
#pragma unroll
                            for (int q = 0; q < SV_GROUP; ++q)
#pragma unroll
                                for (int p = 0; p < R / 2; ++p) {
                                    float2 e = make_float2(ex2_approx(t2[q][p].x), ex2_approx(t2[q][p].y));
                                    acc2[p] = __ffma2_rn(w[q], e, acc2[p]);
                                }
                        }
                    }
                    tab += SV_ROWS;
                }
#pragma unroll
                for (int p = 0; p < R / 2; ++p) {
                    acc64[2 * p] += (double)acc2[p].x;
                    acc64[2 * p + 1] += (double)acc2[p].y;
                }
            }
        }
    }

    // ---- float64 epilogue: Platt, coupling, links, combiner -------------------------------------------------------------
    unsigned fixmask = 0;
    unsigned char lab[R];
    float po[2 * R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        lab[r] = 0;
        po[2 * r] = po[2 * r + 1] = 0.f;
        if (row0 + r >= P.n) continue;
        double f_svc = 0.0;
        bool uncertain = false;
        if (SVC) {
            double czz = P.c_log2 * zz[r];
            f_svc = acc64[r] * exp2(-czz) + P.svc_intercept;
            uncertain = !(czz <= P.zz_limit);
        }
        RowResult res = finish_row(P, SVC, TREE_MODE != 0, has_lin, f_svc, raw[r], zl[r]);
        store_row(P, row0 + r, res, !full);
        po[2 * r] = (float)(1.0 - res.p1);
        po[2 * r + 1] = (float)res.p1;
        lab[r] = res.z > 0 ? 1 : 0;
        if (res.uncertain || uncertain) fixmask |= 1u << r;
    }
    if (P.out && full) {                        // 4 rows x [p0, p1] = two 128-bit stores, 32 contiguous bytes per thread
        float4* o = reinterpret_cast<float4*>(P.out + 2 * row0);
        o[0] = make_float4(po[0], po[1], po[2], po[3]);
        o[1] = make_float4(po[4], po[5], po[6], po[7]);
    }
    if (nonfinite) fixmask = (1u << R) - 1;     // the float64 path writes NaN for those rows and nothing else touches them
    if (nonfinite) atomicOr(&s_nbad, 1u);

    // ---- per-block float64 re-evaluation of the flagged rows -----------------------------------------------------------
    __syncthreads();
    if (fixmask) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            if ((fixmask >> r) & 1u)
                if (row0 + r < P.n) s_fix[atomicAdd(&s_nfix, 1u)] = (unsigned short)(tid * R + r);
    }
    __syncthreads();
    const unsigned nfix = s_nfix;
    if (nfix) {
        const long long block_row0 = (long long)blockIdx.x * BLOCK_ROWS;
        for (unsigned i = 0; i < nfix; ++i) eval_row_f64(P, block_row0 + s_fix[i], s_x, s_z, s_red, &s_hits);
        if (tid == 0) {
            atomicAdd(P.status + 1, nfix);
            if (s_nbad) atomicOr(P.status, HFE_ROW_NONFINITE);
        }
    }

    // ---- optional accuracy count for permutation importance (label vs y; flagged rows were counted by the f64 path) ----
    if (P.y && P.score) {
        unsigned hits = 0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (row0 + r < P.n && !((fixmask >> r) & 1u)) hits += (lab[r] == P.y[row0 + r]) ? 1u : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hits += __shfl_down_sync(0xFFFFFFFFu, hits, o);
        if ((tid & 31) == 0 && hits) atomicAdd(&s_hits, hits);
        __syncthreads();
        if (tid == 0 && s_hits) atomicAdd(P.score, (unsigned long long)s_hits);
    }
}

// Every row through the float64 path (calibration / test aid; also what hfe_predict_proba_f64 runs).
__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_score_f64_kernel(const __grid_constant__ ScoreParams P) {
    __shared__ double s_x[F], s_z[F], s_red[2 * BLOCK_THREADS / 32];
    for (long long row = blockIdx.x; row < P.n; row += gridDim.x) eval_row_f64(P, row, s_x, s_z, s_red, nullptr);
}

// ---- GradientBoostingClassifier.apply ------------------------------------------------------------------------------------

__global__ void __launch_bounds__(BLOCK_THREADS)
hfe_apply_kernel(const __grid_constant__ ScoreParams P, unsigned short* leaves) {
    long long row = (long long)blockIdx.x * BLOCK_THREADS + threadIdx.x;
    if (row >= P.n) return;
    float xr[F];
#pragma unroll
    for (int f = 0; f < F; ++f) xr[f] = __ldg(P.x + (long long)f * P.ld + row);
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
        float fv = (float)v;                       // round-to-nearest, like numpy's astype(float32)
        bad = bad || !(fabsf(fv) <= 3.402823466e38f);   // NaN/Inf, or a float64 too large for float32
        s_tile[(i / nf) * pitch + (i % nf)] = fv;
    }
    __syncthreads();
    for (int f = 0; f < nf; ++f)
        if ((int)threadIdx.x < rows) dst[(long long)f * ld + r0 + threadIdx.x] = s_tile[threadIdx.x * pitch + f];
    if (bad) atomicOr(status, HFE_ROW_NONFINITE);
}

}  // namespace hfe
