// hfe_forest.cuh -- general boosted trees walked out of shared memory (BASELINE config 4: 2000 trees, depth 16).
//
// The reference walks tree-major over 64-byte Node records (SK/ensemble/_gradient_boosting.pyx:25-73, SK/tree/_tree.pxd:
// 15-25).  The fused kernel's general-tree walk reads 8-byte nodes through L1 (__ldg): the lanes of a warp sit at different
// nodes of the same tree, and ncu shows 7.9 sectors per warp-wide load with the L1 tag stage as the bound.  Here
//   * per feature the distinct thresholds of the whole forest are sorted (every schema feature takes a handful of values:
//     a 2000-tree forest has ~90 thresholds in all); a row is turned into 17 counts, per feature the number of thresholds
//     below its value, kept in shared memory as one 32-bit word each with the count in the top byte;
//   * a node is 4 bytes: index of its threshold << 24 | feature << 16 | byte offset of the left child inside the staged
//     chunk.  One level of the walk is: load the node; PRMT + IMAD make the address of the row's count for that feature;
//     load it; ONE unsigned compare of the two words decides (count << 24 > node  <=>  count > threshold index, whatever
//     the lower bits of the node are); one LOP3 turns the node into the child's address (the chunk buffers sit on 32 KB
//     boundaries of the shared window, so base | offset); a predicated add takes the right child.  Leaves carry index 255,
//     which no count exceeds, and their own offset: a fixed-depth walk needs no exit test.  Four instructions on the
//     half-rate integer pipe, one on the FMA pipe, two loads.  Earlier versions, all bit-identical: counts as bytes with an
//     in-tree child index (six integer + four IMAD instructions per level, 70 M rows/s on the stress forest); the
//     comparisons carried as a 128-bit set in registers (nine integer instructions, 55-58 M rows/s); 8-byte nodes through
//     L1 (48.7 M rows/s, and the same with the node table pinned in an L2 persisting window);
//   * the node table streams through shared memory in chunks of up to 8 K nodes (32 KB, 8 trees of the stress forest),
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
constexpr int FOREST_BUF_BYTES = FOREST_CHUNK_NODES * 4;               // 32 KB, and the buffers are aligned to it
constexpr int FOREST_BIN_BYTES = F * FOREST_ROWS * 4;                  // one word per (feature, row of the block)
constexpr size_t FOREST_SMEM_BYTES = 227 * 1024;                       // bins | barriers | pad to a 32 KB boundary | two buffers
static_assert((FOREST_BUF_BYTES & (FOREST_BUF_BYTES - 1)) == 0 && FOREST_BUF_BYTES <= 65536, "offset field: 16 bits, base | offset");
// dynamic shared memory starts 1 KB into the shared window (the reserved kilobyte); the kernel checks what it finds
static_assert(((1024 + FOREST_BIN_BYTES + 64 + FOREST_BUF_BYTES - 1) / FOREST_BUF_BYTES + 2) * FOREST_BUF_BYTES <= 1024 + (int)FOREST_SMEM_BYTES,
              "bins, barriers and two aligned buffers fit");

__global__ void __launch_bounds__(FOREST_THREADS, 1)
hfe_forest_kernel(const __grid_constant__ ScoreParams P, const __grid_constant__ ForestParams Q) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned* s_bins = reinterpret_cast<unsigned*>(smem_raw);                                   // [F][R][FOREST_THREADS] words
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(smem_raw + FOREST_BIN_BYTES);
    constexpr int R = ROWS_PER_THREAD;
    const int tid = threadIdx.x, lane = tid & 31;
    const long long row0 = ((long long)blockIdx.x * FOREST_THREADS + tid) * R;
    const bool full = row0 + R <= P.n;
    const long long row_ld = row0 < P.n ? row0 : 0;
    const bool has_lin = P.slot_lin >= 0;
    // the two chunk buffers: first 32 KB boundary of the shared window behind the barriers
    const unsigned smem0 = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned buf0 = (smem0 + (unsigned)FOREST_BIN_BYTES + 64u + (unsigned)FOREST_BUF_BYTES - 1u) & ~((unsigned)FOREST_BUF_BYTES - 1u);
    if (buf0 + 2u * FOREST_BUF_BYTES > smem0 + (unsigned)FOREST_SMEM_BYTES) __trap();

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
                     ::"r"(buf0 + (unsigned)(c & 1) * FOREST_BUF_BYTES), "l"(Q.nodes4 + start), "r"(bytes), "r"(bar)
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
#pragma unroll
        for (int r = 0; r < R; ++r) s_bins[(f * R + r) * FOREST_THREADS + tid] = cnt[r] << 24;   // own slots: no barrier needed
    }
    const bool nonfinite = !(finite_probe == 0.f);
    __syncthreads();                                                      // the barriers are initialised

    // ---- the walk: chunk by chunk, FOREST_TREES trees at a time ---------------------------------------
    const unsigned bins0 = (unsigned)__cvta_generic_to_shared(s_bins) + (unsigned)tid * 4u;
    float acc[R];                                   // float32 partial sums over groups of TREE_CHUNK trees, as the fused kernel forms them
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.f;
    for (int c = 0; c < Q.n_chunks; ++c) {
        const unsigned buf = buf0 + (unsigned)(c & 1) * FOREST_BUF_BYTES;
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
        for (int t = t_begin; t < t_end; t += FOREST_TREES) {
            // FOREST_TREES trees at a time: FOREST_TREES x R independent chains per thread (the walk is bound by the latency of
            // one level: node load -> count load -> compare -> add); a short last group repeats its last tree and drops the copy
            int depth = 0;
            unsigned at[FOREST_TREES][R];                 // shared-window address of each chain's node
#pragma unroll
            for (int k = 0; k < FOREST_TREES; ++k) {
                const int tk = min(t + k, t_end - 1);
                const unsigned root = buf + (unsigned)(__ldg(P.tree_off + tk) - origin) * 4u;
                depth = max(depth, (int)__ldg(P.tree_depth + tk));
#pragma unroll
                for (int r = 0; r < R; ++r) at[k][r] = root;
            }
            for (int lvl = 0; lvl < depth; ++lvl) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
#pragma unroll
                    for (int k = 0; k < FOREST_TREES; ++k) {
                        unsigned d, w;
                        asm("ld.shared.u32 %0, [%1];" : "=r"(d) : "r"(at[k][r]));
                        asm("ld.shared.u32 %0, [%1];" : "=r"(w)
                            : "r"(__byte_perm(d, 0u, 0x4442u) * (unsigned)(R * FOREST_THREADS * 4) + bins0 + (unsigned)(r * FOREST_THREADS * 4)));
                        unsigned next = (d & 0xFFFFu) | buf;
                        if (w > d) next += 4u;
                        at[k][r] = next;
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < FOREST_TREES; ++k) {
                const bool live = t + k < t_end;
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (live) acc[r] += __ldg(P.value32 + origin + (int)((at[k][r] - buf) >> 2));
            }
            if ((min(t + FOREST_TREES, t_end) / TREE_CHUNK) != (t / TREE_CHUNK)) {   // float32 partial sums never span more than 16 trees
#pragma unroll
                for (int r = 0; r < R; ++r) { raw[r] += (double)acc[r]; acc[r] = 0.f; }
            }
        }
        __syncthreads();                                                  // everyone is done with this buffer
        if (tid == 0 && c + 2 < Q.n_chunks) issue_chunk(c + 2);
    }

#pragma unroll
    for (int r = 0; r < R; ++r) raw[r] += (double)acc[r];

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
