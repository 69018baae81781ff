// hfe_forest.cuh -- general boosted trees walked out of shared memory (BASELINE config 4: 2000 trees, depth 16).
//
// The reference walks tree-major over 64-byte Node records (SK/ensemble/_gradient_boosting.pyx:25-73, SK/tree/_tree.pxd:
// 15-25).  The fused kernel's general-tree walk reads 8-byte nodes through L1 (__ldg): the lanes of a warp sit at different
// nodes of the same tree, and ncu shows 7.9 sectors per warp-wide load with the L1 tag stage as the bound.  Here
//   * a row first turns into a bit set: one bit per distinct (feature, threshold) pair of the whole forest, "x[feature] >
//     threshold" (every schema feature takes a handful of values, so a 2000-tree forest has ~90 pairs);
//   * a node is 4 bytes: pair index | in-tree index of the left child; one level of the walk is one shared-memory load, a
//     bit test in registers and an add (leaves point at themselves through pair 0, which is never set, so a fixed-depth
//     predicated walk needs no exit test);
//   * the node table streams through shared memory in chunks of up to 16 K nodes (64 KB, ~16 trees of the stress forest),
//     fetched by TMA bulk copies (cp.async.bulk + mbarrier) into two buffers: the whole block walks one chunk while the
//     copy engine brings the next.  Breadth-first node order puts the first levels of a tree -- the ones all lanes visit --
//     into distinct banks.
// Leaf values stay in global memory (one load per row and tree).  Stacks without an SVC member only; with one, the trees
// are a side show and stay in the fused kernel.
#pragma once

#include "hfe_kernels.cuh"

namespace hfe {

constexpr int FOREST_THREADS = 512;
constexpr int FOREST_ROWS = FOREST_THREADS * ROWS_PER_THREAD;          // 2048 rows per block
constexpr int FOREST_BUF_BYTES = (FOREST_CHUNK_NODES + 4) * 4;
constexpr size_t FOREST_SMEM_BYTES = 2 * (size_t)FOREST_BUF_BYTES + 64;

template <int W>                                                        // 32-bit words of the per-row bit set
__global__ void __launch_bounds__(FOREST_THREADS, 1)
hfe_forest_kernel(const __grid_constant__ ScoreParams P, const __grid_constant__ ForestParams Q) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned* s_buf[2] = {reinterpret_cast<unsigned*>(smem_raw), reinterpret_cast<unsigned*>(smem_raw + FOREST_BUF_BYTES)};
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(smem_raw + 2 * FOREST_BUF_BYTES);
    constexpr int R = ROWS_PER_THREAD;
    const int tid = threadIdx.x, lane = tid & 31;
    const long long row0 = ((long long)blockIdx.x * FOREST_THREADS + tid) * R;
    const bool full = row0 + R <= P.n;
    const long long row_ld = row0 < P.n ? row0 : 0;
    const bool has_lin = P.slot_lin >= 0;

    auto bar_addr = [&](int b) { return (unsigned)__cvta_generic_to_shared(s_bar + b); };
    // chunk c of the node table: trees [chunk_tree[c], chunk_tree[c + 1]); the copy starts 16-byte aligned
    auto chunk_start = [&](int c) { return (long long)__ldg(P.tree_off + __ldg(Q.chunk_tree + c)) & ~3LL; };
    auto issue_chunk = [&](int c) {
        const long long start = chunk_start(c);
        const long long end = (long long)__ldg(P.tree_off + __ldg(Q.chunk_tree + c + 1));
        const unsigned bytes = (unsigned)(((end - start + 3) & ~3LL) * 4);
        const unsigned bar = bar_addr(c & 1);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"((unsigned)__cvta_generic_to_shared(s_buf[c & 1])), "l"(Q.nodes4 + start), "r"(bytes), "r"(bar)
                     : "memory");
    };
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr(0)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr(1)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        issue_chunk(0);
        if (Q.n_chunks > 1) issue_chunk(1);
    }

    // ---- rows -> bit sets (and the logistic member's dot product) while the first chunks are on their way --------------------
    unsigned w[R][W];
    double zl[R], raw[R];
    float finite_probe = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        zl[r] = P.lin_intercept;
        raw[r] = P.gbt_init;
#pragma unroll
        for (int k = 0; k < W; ++k) w[r][k] = 0u;
    }
    const bool needs_fix = !full || P.perm_col >= 0;
#pragma unroll
    for (int f = 0; f < F; ++f) {
        float4 v = load_rows4(P, f, row_ld);
        if (needs_fix) v = fix_rows4(P, f, row0, row_ld, full, v);
        const float xf[R] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (has_lin) zl[r] = fma((double)xf[r], P.lin_coef[f], zl[r]);
            finite_probe = fmaf(xf[r], 0.f, finite_probe);
        }
        // the pairs of this feature, word by word of the bit set (warp-uniform bounds: a handful of pairs per feature)
        const int p_lo = Q.pair_begin[f], p_hi = Q.pair_begin[f + 1];
#pragma unroll
        for (int k = 0; k < W; ++k) {
            const int lo = max(p_lo, 32 * k), hi = min(p_hi, 32 * k + 32);
            unsigned m[R];
#pragma unroll
            for (int r = 0; r < R; ++r) m[r] = 0u;
#pragma unroll 1
            for (int p = lo; p < hi; ++p) {
                const float thr = Q.pair_thr[p];
                const unsigned bit = 1u << (p - 32 * k);
#pragma unroll
                for (int r = 0; r < R; ++r) m[r] |= xf[r] > thr ? bit : 0u;
            }
#pragma unroll
            for (int r = 0; r < R; ++r) w[r][k] |= m[r];
        }
    }
    const bool nonfinite = !(finite_probe == 0.f);
    // the bit set as two 64-bit halves held in scalars: a bit test is a select and a funnel shift, never an indexed array
    unsigned long long b_lo[R], b_hi[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        b_lo[r] = (unsigned long long)w[r][0] | (W > 1 ? (unsigned long long)w[r][W > 1 ? 1 : 0] << 32 : 0ull);
        b_hi[r] = W > 2 ? ((unsigned long long)w[r][W > 2 ? 2 : 0] | ((unsigned long long)w[r][W > 3 ? 3 : 0] << 32)) : 0ull;
    }
    __syncthreads();                                                      // the barriers are initialised

    // ---- the walk: chunk by chunk, two trees at a time (8 independent chains per thread) ---------------------------------------
    for (int c = 0; c < Q.n_chunks; ++c) {
        const unsigned* nodes = reinterpret_cast<const unsigned*>(smem_raw + (c & 1) * FOREST_BUF_BYTES);
        {
            const unsigned bar = bar_addr(c & 1), parity = (unsigned)(c >> 1) & 1u;
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "FOREST_WAIT_%=:\n"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                "@p bra FOREST_DONE_%=;\n"
                "bra FOREST_WAIT_%=;\n"
                "FOREST_DONE_%=:\n"
                "}\n" ::"r"(bar), "r"(parity) : "memory");
        }
        const int t_begin = __ldg(Q.chunk_tree + c), t_end = __ldg(Q.chunk_tree + c + 1);
        const int origin = (int)chunk_start(c);
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
        int folded = 0;
        for (int t = t_begin; t < t_end; t += 2) {
            const bool two = t + 1 < t_end;
            const int tb = two ? t + 1 : t;
            const int off_a = __ldg(P.tree_off + t), off_b = __ldg(P.tree_off + tb);
            const unsigned* na = nodes + (off_a - origin);
            const unsigned* nbp = nodes + (off_b - origin);
            const int depth = max((int)__ldg(P.tree_depth + t), (int)__ldg(P.tree_depth + tb));
            unsigned ia[R], ib[R];
#pragma unroll
            for (int r = 0; r < R; ++r) { ia[r] = 0u; ib[r] = 0u; }
            for (int lvl = 0; lvl < depth; ++lvl) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const unsigned da = na[ia[r]], db = nbp[ib[r]];
                    const unsigned pa = da >> 20, pb = db >> 20;
                    const unsigned long long sa = (W > 2 && pa >= 64u) ? b_hi[r] : b_lo[r];
                    const unsigned long long sb = (W > 2 && pb >= 64u) ? b_hi[r] : b_lo[r];
                    ia[r] = (da & 0xFFFFFu) + ((unsigned)(sa >> (pa & 63u)) & 1u);
                    ib[r] = (db & 0xFFFFFu) + ((unsigned)(sb >> (pb & 63u)) & 1u);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc[r] += __ldg(P.value32 + off_a + ia[r]);
                if (two) acc[r] += __ldg(P.value32 + off_b + ib[r]);
            }
            folded += 2;
            if (folded >= TREE_CHUNK) {                                   // float32 partial sums never span more than 16 trees
#pragma unroll
                for (int r = 0; r < R; ++r) { raw[r] += (double)acc[r]; acc[r] = 0.f; }
                folded = 0;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) raw[r] += (double)acc[r];
        __syncthreads();                                                  // everyone is done with this buffer
        if (tid == 0 && c + 2 < Q.n_chunks) issue_chunk(c + 2);
    }

    // ---- links, combiner, outputs: the fused kernel's epilogue without an SVC member ---------------------------------------------
    double zpart[R], member[3][R];
    partial_logit_fast<R>(P, true, has_lin, raw, zl, zpart, member);
    if (P.meta) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (row0 + r >= P.n) continue;
            if (P.slot_gbt >= 0) P.meta[(row0 + r) * P.n_members + P.slot_gbt] = member[1][r];
            if (P.slot_lin >= 0) P.meta[(row0 + r) * P.n_members + P.slot_lin] = member[2][r];
        }
    }
    double f_svc[R], p1[R], zfin[R];
    bool unc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) f_svc[r] = 0.0;
    finish_rows_fast<R>(P, false, f_svc, zpart, p1, zfin, member[0], unc);
    emit_rows<R, double>(P, row0, lane, full, nonfinite, 0u, p1, zfin, member[0], unc);
}

}  // namespace hfe
