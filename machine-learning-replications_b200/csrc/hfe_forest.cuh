// hfe_forest.cuh -- general boosted trees walked out of shared memory (BASELINE config 4: 2000 trees, depth 16).
//
// The reference walks tree-major over 64-byte Node records (SK/ensemble/_gradient_boosting.pyx:25-73, SK/tree/_tree.pxd:
// 15-25).  The fused kernel's general-tree walk reads 8-byte nodes through L1 (__ldg): the lanes of a warp sit at different
// nodes of the same tree, and ncu shows 7.9 sectors per warp-wide load with the L1 tag stage as the bound.  Here
//   * per feature the distinct thresholds of the whole forest are sorted (every schema feature takes a handful of values:
//     a 2000-tree forest has ~90 thresholds in all); a row is turned into 17 bytes, per feature the number of thresholds
//     below its value, kept in shared memory (one 32-bit word per thread and feature = its four rows);
//   * a node is 4 bytes: feature | index of its threshold | in-tree index of the left child.  One level of the walk is a
//     shared-memory load of the node, a byte load of the row's count for that feature, a compare and an add (leaves carry
//     index 255, which no count exceeds, and point at themselves: a fixed-depth predicated walk needs no exit test).  The
//     integer pipe runs at half rate on this part, so the step is kept to five integer instructions; address arithmetic
//     goes to the FMA pipe (IMAD).  A first version that carried the comparisons as a 128-bit set in registers needed
//     nine and was bound by that pipe at 55-58 M rows/s;
//   * the node table streams through shared memory in chunks of up to 16 K nodes (64 KB, ~16 trees of the stress forest),
//     fetched by TMA bulk copies (cp.async.bulk + mbarrier) into two buffers: the whole block walks one chunk while the
//     copy engine brings the next.  Breadth-first node order puts the first levels of a tree -- the ones all lanes visit --
//     into distinct banks.
// Leaf values stay in global memory (one load per row and tree).  Stacks without an SVC member only; with one, the trees
// are a side show and stay in the fused kernel.
#pragma once

#include "hfe_kernels.cuh"

namespace hfe {

#ifndef HFE_FOREST_TREES
#define HFE_FOREST_TREES 4
#endif
constexpr int FOREST_TREES = HFE_FOREST_TREES;                         // trees walked side by side by one thread
constexpr int FOREST_THREADS = 512;
constexpr int FOREST_ROWS = FOREST_THREADS * ROWS_PER_THREAD;          // 2048 rows per block
constexpr int FOREST_BUF_BYTES = (FOREST_CHUNK_NODES + 4) * 4;
constexpr int FOREST_BIN_BYTES = F * FOREST_ROWS;                     // one byte per (feature, row of the block)
constexpr size_t FOREST_SMEM_BYTES = 2 * (size_t)FOREST_BUF_BYTES + FOREST_BIN_BYTES + 64;

__global__ void __launch_bounds__(FOREST_THREADS, 1)
hfe_forest_kernel(const __grid_constant__ ScoreParams P, const __grid_constant__ ForestParams Q) {
    static_assert(FOREST_THREADS * 4 == FOREST_ROWS, "one 32-bit word of counts per thread and feature");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned* s_buf[2] = {reinterpret_cast<unsigned*>(smem_raw), reinterpret_cast<unsigned*>(smem_raw + FOREST_BUF_BYTES)};
    unsigned* s_bins = reinterpret_cast<unsigned*>(smem_raw + 2 * FOREST_BUF_BYTES);          // [F][FOREST_THREADS] words
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(smem_raw + 2 * FOREST_BUF_BYTES + FOREST_BIN_BYTES);
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

    // ---- rows -> per-feature threshold counts (and the logistic member's dot product) while the first chunks are on their way
    double zl[R], raw[R];
    float finite_probe = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        zl[r] = P.lin_intercept;
        raw[r] = P.gbt_init;
    }
    const bool needs_fix = !full || P.perm_col >= 0;
#pragma unroll
    for (int f = 0; f < F; ++f) {
        float4 v = load_rows4(P, f, row_ld);
        if (needs_fix) v = fix_rows4(P, f, row0, row_ld, full, v);
        const float xf[R] = {v.x, v.y, v.z, v.w};
        unsigned cnt[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (has_lin) zl[r] = fma((double)xf[r], P.lin_coef[f], zl[r]);
            finite_probe = fmaf(xf[r], 0.f, finite_probe);
            cnt[r] = 0u;
        }
#pragma unroll 1
        for (int k = Q.thr_begin[f]; k < Q.thr_begin[f + 1]; ++k) {         // warp-uniform: a handful of thresholds per feature
            const float thr = Q.thr[k];
#pragma unroll
            for (int r = 0; r < R; ++r) cnt[r] += xf[r] > thr ? 1u : 0u;
        }
        s_bins[f * FOREST_THREADS + tid] = cnt[0] | (cnt[1] << 8) | (cnt[2] << 16) | (cnt[3] << 24);   // own slot: no barrier needed
    }
    const bool nonfinite = !(finite_probe == 0.f);
    __syncthreads();                                                      // the barriers are initialised

    // ---- the walk: chunk by chunk, FOREST_TREES trees at a time ---------------------------------------
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
        for (int t = t_begin; t < t_end; t += FOREST_TREES) {
            // FOREST_TREES trees at a time: FOREST_TREES x R independent chains per thread (the walk is bound by the latency of
            // one level: shared-memory load -> bit test -> add); a short last group repeats its last tree and drops the copy
            unsigned base[FOREST_TREES];
            int off[FOREST_TREES], depth = 0;
#pragma unroll
            for (int k = 0; k < FOREST_TREES; ++k) {
                const int tk = min(t + k, t_end - 1);
                off[k] = __ldg(P.tree_off + tk);
                base[k] = (unsigned)__cvta_generic_to_shared(nodes) + (unsigned)(off[k] - origin) * 4u;
                depth = max(depth, (int)__ldg(P.tree_depth + tk));
            }
            unsigned at[FOREST_TREES][R];                 // byte address of each chain's node
#pragma unroll
            for (int k = 0; k < FOREST_TREES; ++k)
#pragma unroll
                for (int r = 0; r < R; ++r) at[k][r] = base[k];
            const unsigned bins0 = (unsigned)__cvta_generic_to_shared(s_bins) + (unsigned)tid * 4u;
            for (int lvl = 0; lvl < depth; ++lvl) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
#pragma unroll
                    for (int k = 0; k < FOREST_TREES; ++k) {
                        unsigned d, c;
                        asm("ld.shared.u32 %0, [%1];" : "=r"(d) : "r"(at[k][r]));
                        // node: feature | threshold index | in-tree index of the left child; the row's count for that feature is
                        // byte r of the thread's word in row `feature` of the table
                        asm("ld.shared.u8 %0, [%1];" : "=r"(c) : "r"((d >> 24) * (unsigned)(FOREST_THREADS * 4) + (bins0 + r)));
                        const unsigned step = c > __byte_perm(d, 0u, 0x4442u) ? 4u : 0u;
                        at[k][r] = (d & 0xFFFFu) * 4u + base[k] + step;
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < FOREST_TREES; ++k) {
                const bool live = t + k < t_end;
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (live) acc[r] += __ldg(P.value32 + off[k] + (int)((at[k][r] - base[k]) >> 2));
            }
            folded += FOREST_TREES;
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
