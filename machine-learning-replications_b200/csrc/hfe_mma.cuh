// hfe_mma.cuh -- the RBF-SVC member on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// What libsvm computes per row (svm.cpp:461-515, 2821-2904) is f = sum_j coef_j * exp(-gamma |z - s_j|^2) + b, and the
// exponent is one row of a small dense contraction:
//     t_j = -c|z - s_j|^2 = sum_d xc_d * W_dj + B_j + r          (log2 units, c = gamma * log2 e)
//     xc_d = x_d - m_d            raw feature minus a fixed centre (the scaler mean rounded to an integer)
//     W_dj = 2c s_jd / scale_d    B_j = sum_d (m_d - mean_d) W_dj - c|s_j|^2        r = -c|z|^2   (per row)
// i.e. T[128 rows x n_sv] = A[128 x K] * Wt[K x n_sv] with the row term and the bias carried in spare K slots.
// The operands are bfloat16 but nothing is rounded away: every W_dj, B_j and r is split into three bfloat16 parts
// (hi + mid + lo = 24 significant bits) and so is xc_d; the products of 8-bit significands are exact in the float32
// accumulator.  For cohorts in the load_data_public.py schema every feature is a small integer, xc_d is one exact
// bfloat16 and the base block (5 K-steps of 16) is all there is; when a tile holds a value that needs more bits the
// issuing thread appends the extension block (mid and lo parts of xc, 5 more K-steps) for that tile only.
// The tensor core aligns the products of one instruction (and the accumulator) to the largest exponent among them and
// drops what falls below ~2^-24 of it (measured with tools/mma_probe.cu: one term of 144 next to 79 terms of ~1e-4 loses
// 3e-4).  So terms of different magnitude classes never share an accumulator: per support vector one TMEM column takes
// the hi x hi products (8-bit x 8-bit significands around 2^3: exact), a second one everything 2^-8 and below (mid and
// lo parts, extension block); the epilogue adds the two in float32.
//
// One persistent CTA per SM, 14 warps:
//   warps 0-3   producers: load the tile's 128 rows (coalesced column reads), centre and split them, write the A operand
//               into shared memory in the canonical K-major core-matrix layout; they also evaluate the tree and logistic
//               members (their share of the combiner's logit travels to the epilogue through shared memory)
//   warps 4-7   epilogue of stream 0, warps 8-11 epilogue of stream 1: thread = row (TMEM lane); tcgen05.ld 16 columns
//               at a time, ex2, coefficient FFMA2, float64 fold; then Platt + coupling + combiner (finish_rows_fast)
//   warps 12,13 one lane each issues the tcgen05.mma instructions of its stream and commits them to mbarriers
// Tiles alternate between the two streams.  Support vectors are taken in blocks of 64; TMEM (512 columns) is cut into
// four regions of 128 columns ([hi accumulator | rest accumulator] of one block), two per stream: while a stream's
// epilogue warps drain one block from one region, the tensor core fills the other, and the other stream's epilogue
// keeps the MUFU pipe busy in the meantime.
#pragma once

#include <cuda_bf16.h>

#include "hfe_kernels.cuh"

namespace hfe {

constexpr int MMA_TILE_ROWS = 128;
constexpr int MMA_PRODUCER_WARPS = 4, MMA_EPI_WARPS = 8, MMA_ISSUE_WARPS = 2;
constexpr int MMA_THREADS = 32 * (MMA_PRODUCER_WARPS + MMA_EPI_WARPS + MMA_ISSUE_WARPS);      // 448
constexpr int MMA_A_STAGES = 3;
constexpr int MMA_A_CHUNKS = 20;                         // 16-byte K chunks per row: 10 base + 10 extension
constexpr int MMA_A_STAGE_BYTES = (MMA_TILE_ROWS / 8) * MMA_A_CHUNKS * 128;                   // 40 960
constexpr int MMA_SIDE_BYTES = MMA_TILE_ROWS * 32;       // per (stream, slot): zpart, p_gbt, p_lin, flags per row

// ---- PTX helpers ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Waits for the phase with the given parity to complete.  A stall of about a second can only be a protocol bug; it is
// turned into a trap (the launch fails) instead of a hung GPU.
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned done = 0;
    long long t0 = 0;
    for (unsigned spin = 0; !done; ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && (spin & 1023u) == 1023u) {
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > 4000000000LL) __trap();
        }
    }
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor: K-major, no swizzle; LBO = bytes between the two 16-byte K chunks of one K-step,
// SBO = bytes between 8-row groups (cute::UMMA::SmemDescriptor: start [0,14), LBO [16,30), SBO [32,46), version 1 at 46)
__device__ __forceinline__ unsigned long long umma_desc(unsigned addr, unsigned lbo, unsigned sbo) {
    return (unsigned long long)((addr >> 4) & 0x3FFF) | ((unsigned long long)((lbo >> 4) & 0x3FFF) << 16) |
           ((unsigned long long)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_bf16(unsigned d_tmem, unsigned long long da, unsigned long long db, unsigned idesc,
                                          unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// 16 consecutive float32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(unsigned taddr, unsigned (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
}
// tcgen05.wait::ld; the registers pass through the statement so that no use of them can be scheduled above it
__device__ __forceinline__ void tmem_ld_wait(unsigned (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                   "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
                 :: "memory");
}

__device__ __forceinline__ void tmem_ld_wait2(unsigned (&v)[16], unsigned (&w)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                   "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
                 :: "memory");
    asm volatile("" : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]), "+r"(w[6]), "+r"(w[7]),
                      "+r"(w[8]), "+r"(w[9]), "+r"(w[10]), "+r"(w[11]), "+r"(w[12]), "+r"(w[13]), "+r"(w[14]), "+r"(w[15]));
}

__device__ __forceinline__ unsigned pack_bf16(float lo, float hi) {       // lo in bits [0,16), hi in [16,32)
    __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<unsigned*>(&p);
}
__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// float32 -> float64 with integer instructions (the F2F conversion shares the MUFU pipe, which this kernel saturates).
// Exact for normal values; a float32 zero / denormal comes out below 2^-126, far under the sum's resolution.
__device__ __forceinline__ double widen_f32(float v) {
    const unsigned b = __float_as_uint(v);
    return __hiloint2double((int)((b & 0x80000000u) | (((b & 0x7FFFFFFFu) >> 3) + 0x38000000u)), (int)(b << 29));
}

// ---- the kernel ------------------------------------------------------------------------------------------------------------
template <int TREE_MODE>
__global__ void __launch_bounds__(MMA_THREADS, 1)
hfe_mma_kernel(const __grid_constant__ ScoreParams P, const __grid_constant__ MmaParams Q) {
    static_assert(TREE_MODE == 0 || TREE_MODE == 1, "general tree tables keep the SIMT kernel");
    extern __shared__ __align__(1024) unsigned char smem[];
    // [W image + coef | A stages | side data | barriers]
    const int w_region = (Q.w_bytes + 1023) & ~1023;
    unsigned char* s_w = smem;
    unsigned char* s_a = smem + w_region;
    unsigned char* s_side = s_a + MMA_A_STAGES * MMA_A_STAGE_BYTES;
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_side + 4 * MMA_SIDE_BYTES);
    unsigned* s_misc = reinterpret_cast<unsigned*>(s_bar + 32);
    // barrier indices
    enum { B_W = 0, B_AFULL = 1, B_AEMPTY = 4, B_ACCFULL = 7, B_ACCEMPTY = 11, B_SIDEFULL = 15, B_SIDEEMPTY = 19, B_COUNT = 23 };
    auto bar = [&](int i) { return smem_addr(s_bar + i); };
    volatile unsigned* s_tmem = s_misc;                    // TMEM base address
    volatile unsigned* s_ext = s_misc + 4;                 // [stage][producer warp]: tile needs the extension block

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long n_tiles = (P.n + MMA_TILE_ROWS - 1) / MMA_TILE_ROWS;
    const int nb = Q.nb;

    // ---- set-up -------------------------------------------------------------------------------------------------------------
    {   // zero the A stages once: the padding chunk of every K block is never written again
        uint4* z = reinterpret_cast<uint4*>(s_a);
        for (int i = tid; i < MMA_A_STAGES * MMA_A_STAGE_BYTES / 16; i += MMA_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (tid == 0) {
        mbar_init(bar(B_W), 1);
        for (int i = 0; i < MMA_A_STAGES; ++i) { mbar_init(bar(B_AFULL + i), MMA_PRODUCER_WARPS); mbar_init(bar(B_AEMPTY + i), 1); }
        for (int i = 0; i < 4; ++i) {
            mbar_init(bar(B_ACCFULL + i), 1);
            mbar_init(bar(B_ACCEMPTY + i), 4);
            mbar_init(bar(B_SIDEFULL + i), MMA_PRODUCER_WARPS);
            mbar_init(bar(B_SIDEEMPTY + i), 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 12) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr((const void*)s_tmem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const unsigned tmem_base = *s_tmem;
    if (tid == 0) {                                        // the B operand and the coefficient table: one bulk copy
        const unsigned bytes = (unsigned)Q.w_bytes;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar(B_W)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_addr(s_w)), "l"(Q.w_img), "r"(bytes), "r"(bar(B_W)) : "memory");
    }

    if (warp < MMA_PRODUCER_WARPS) {
        // =========================== producers ==================================================================================
        const bool has_lin = P.slot_lin >= 0;
        long long i = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
            const int stage = (int)(i % MMA_A_STAGES);
            const int s = (int)(i & 1), slot = (int)((i >> 1) & 1);
            const long long row = tile * MMA_TILE_ROWS + tid;
            const bool valid = row < P.n;
            const long long row_ld = valid ? row : 0;
            // this row's 17 features: one coalesced 128-byte read per warp and column, all in flight at once
            float x[F];
#pragma unroll
            for (int f = 0; f < F; ++f) x[f] = __ldg(P.x + P.col_off[f] + row_ld);
            if (P.perm_col >= 0) {
                HFE_ASSERT_IN(__ldg(P.perm_idx + row_ld) >= 0 && __ldg(P.perm_idx + row_ld) < P.n);
#pragma unroll
                for (int f = 0; f < F; ++f)
                    if (f == P.perm_col) x[f] = __ldg(P.x + P.col_off[f] + __ldg(P.perm_idx + row_ld));
            }
            float probe = 0.f;
#pragma unroll
            for (int f = 0; f < F; ++f) probe = fmaf(x[f], 0.f, probe);          // NaN exactly when some x is NaN / Inf
            const bool nonfinite = valid && !(probe == 0.f);
            if (!valid || nonfinite) {
#pragma unroll
                for (int f = 0; f < F; ++f) x[f] = Q.center[f];                      // keeps NaN out of the tensor core
            }
            // centre, split, row term
            float xc[F], hi[F], rem[F];
            double zz = 0.0, zl = P.lin_intercept;
            bool inexact = false;
#pragma unroll
            for (int f = 0; f < F; ++f) {
                xc[f] = x[f] - Q.center[f];
                hi[f] = bf16_round(xc[f]);
                rem[f] = xc[f] - hi[f];
                inexact = inexact || (rem[f] != 0.f);
                const double zd = (double)xc[f] + Q.dmean[f];
                zz = fma(zd * zd, Q.inv_var[f], zz);
                if (has_lin) zl = fma((double)x[f], P.lin_coef[f], zl);
            }
            const float r = (float)(Q.neg_c * zz);
            const float r_hi = bf16_round(r), r_mid = bf16_round(r - r_hi), r_lo = bf16_round((r - r_hi) - r_mid);
            // the other members (their links, and their share of the combiner's logit)
            double raw1[1] = {0.0}, zl1[1] = {zl}, zpart[1], member[3][1];
            if (TREE_MODE == 1) {
                float acc[1] = {0.f};
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    const float xf[1] = {x[f]};
                    apply_splits<1>(P, f, xf, acc);
                }
                raw1[0] = P.gbt_split_base + (double)acc[0];
            }
            partial_logit_fast<1>(P, TREE_MODE != 0, has_lin, raw1, zl1, zpart, member);

            mbar_wait(bar(B_AEMPTY + stage), (unsigned)(((i / MMA_A_STAGES) & 1) ^ 1));
            // A operand: [row group][chunk][row in group][8 bf16]; base chunks H0 H1 H2 0 | H0 H1 H2 H0 H1 H2 (chunk 3 stays 0):
            // K-steps 0,1 meet the hi parts of W (-> hi accumulator), K-steps 2-4 the mid and lo parts (-> rest accumulator)
            unsigned char* a_row = s_a + stage * MMA_A_STAGE_BYTES + (tid >> 3) * (MMA_A_CHUNKS * 128) + (tid & 7) * 16;
            const uint4 h0 = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]), pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
            const uint4 h1 = make_uint4(pack_bf16(hi[8], hi[9]), pack_bf16(hi[10], hi[11]), pack_bf16(hi[12], hi[13]), pack_bf16(hi[14], hi[15]));
            const uint4 h2 = make_uint4(pack_bf16(hi[16], 1.f), pack_bf16(r_hi, r_mid), pack_bf16(r_lo, 0.f), 0u);
            *reinterpret_cast<uint4*>(a_row + 0 * 128) = h0; *reinterpret_cast<uint4*>(a_row + 1 * 128) = h1;
            *reinterpret_cast<uint4*>(a_row + 2 * 128) = h2;
            *reinterpret_cast<uint4*>(a_row + 4 * 128) = h0; *reinterpret_cast<uint4*>(a_row + 5 * 128) = h1;
            *reinterpret_cast<uint4*>(a_row + 6 * 128) = h2; *reinterpret_cast<uint4*>(a_row + 7 * 128) = h0;
            *reinterpret_cast<uint4*>(a_row + 8 * 128) = h1; *reinterpret_cast<uint4*>(a_row + 9 * 128) = h2;
            // extension block M0 M1 M2 M0 M1 M2 L0 L1 L2: only tiles holding a value that one bfloat16 cannot carry use it,
            // but then all four warps' rows must be there
            const bool ext = __any_sync(0xFFFFFFFFu, inexact);
            uint4 m0 = make_uint4(0, 0, 0, 0), m1 = m0, m2 = m0, l0 = m0, l1 = m0, l2 = m0;
            if (ext) {
                float mid[F], lo[F];
#pragma unroll
                for (int f = 0; f < F; ++f) { mid[f] = bf16_round(rem[f]); lo[f] = bf16_round(rem[f] - mid[f]); }
                m0 = make_uint4(pack_bf16(mid[0], mid[1]), pack_bf16(mid[2], mid[3]), pack_bf16(mid[4], mid[5]), pack_bf16(mid[6], mid[7]));
                m1 = make_uint4(pack_bf16(mid[8], mid[9]), pack_bf16(mid[10], mid[11]), pack_bf16(mid[12], mid[13]), pack_bf16(mid[14], mid[15]));
                m2 = make_uint4(pack_bf16(mid[16], 0.f), 0u, 0u, 0u);
                l0 = make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
                l1 = make_uint4(pack_bf16(lo[8], lo[9]), pack_bf16(lo[10], lo[11]), pack_bf16(lo[12], lo[13]), pack_bf16(lo[14], lo[15]));
                l2 = make_uint4(pack_bf16(lo[16], 0.f), 0u, 0u, 0u);
            }
            unsigned char* e_row = a_row + 10 * 128;
            *reinterpret_cast<uint4*>(e_row + 0 * 128) = m0; *reinterpret_cast<uint4*>(e_row + 1 * 128) = m1;
            *reinterpret_cast<uint4*>(e_row + 2 * 128) = m2; *reinterpret_cast<uint4*>(e_row + 3 * 128) = m0;
            *reinterpret_cast<uint4*>(e_row + 4 * 128) = m1; *reinterpret_cast<uint4*>(e_row + 5 * 128) = m2;
            *reinterpret_cast<uint4*>(e_row + 6 * 128) = l0; *reinterpret_cast<uint4*>(e_row + 7 * 128) = l1;
            *reinterpret_cast<uint4*>(e_row + 8 * 128) = l2;
            if (lane == 0) s_ext[stage * 4 + warp] = ext ? 1u : 0u;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");           // these stores before the tensor core's reads
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_AFULL + stage));

            // side data for the epilogue thread of this row
            mbar_wait(bar(B_SIDEEMPTY + 2 * s + slot), (unsigned)(((i >> 2) & 1) ^ 1));
            double* side = reinterpret_cast<double*>(s_side + (2 * s + slot) * MMA_SIDE_BYTES) + tid * 4;
            side[0] = zpart[0];
            side[1] = member[1][0];
            side[2] = member[2][0];
            side[3] = nonfinite ? 1.0 : 0.0;
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_SIDEFULL + 2 * s + slot));
        }
        // the last commits onto the A-stage barriers must have landed before the CTA may retire
        for (long long k = (i > MMA_A_STAGES ? i - MMA_A_STAGES : 0); k < i; ++k)
            mbar_wait(bar(B_AEMPTY + (int)(k % MMA_A_STAGES)), (unsigned)((k / MMA_A_STAGES) & 1));
    } else if (warp < MMA_PRODUCER_WARPS + MMA_EPI_WARPS) {
        // =========================== epilogue warps ==============================================================================
        const int s = (warp - MMA_PRODUCER_WARPS) >> 2;              // stream
        const int row_in_tile = (warp & 3) * 32 + lane;              // TMEM lane = row of the tile (warp % 4 picks the lane quarter)
        const unsigned lane_base = tmem_base + ((unsigned)((warp & 3) * 32) << 16);
        mbar_wait(bar(B_W), 0);
        const float* s_coef = reinterpret_cast<const float*>(s_w + (size_t)nb * MMA_BLOCK_SV * (MMA_B_CHUNKS * 16));
        long long j = 0;
        for (long long i = s, tile = blockIdx.x + (long long)s * gridDim.x; tile < n_tiles; i += 2, tile += 2LL * gridDim.x, ++j) {
            const int slot = (int)(j & 1);
            mbar_wait(bar(B_SIDEFULL + 2 * s + slot), (unsigned)((j >> 1) & 1));
            const double* side = reinterpret_cast<const double*>(s_side + (2 * s + slot) * MMA_SIDE_BYTES) + row_in_tile * 4;
            const double zpart0 = side[0], p_gbt = side[1], p_lin = side[2];
            const bool nonfinite = side[3] != 0.0;
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_SIDEEMPTY + 2 * s + slot));

            double acc64 = 0.0;
#pragma unroll 1
            for (int blk = 0; blk < nb; ++blk) {
                const long long use = j * nb + blk;                  // this stream's running count of blocks
                const int reg = (int)(use & 1);
                mbar_wait(bar(B_ACCFULL + 2 * s + reg), (unsigned)((use >> 1) & 1));
                tc_fence_after();
                const unsigned taddr = lane_base + (unsigned)((2 * s + reg) * (2 * MMA_BLOCK_SV));   // hi accumulator; rest at +64
                const float4* coef4 = reinterpret_cast<const float4*>(s_coef + blk * MMA_BLOCK_SV);
                // 16 support vectors per step: exponent = hi + rest, 2^x on the MUFU pipe, coefficient FFMA2 (two vectors per
                // instruction), float32 partial sums of 4 vectors each folded into the float64 accumulator; the TMEM loads of
                // the next step are in flight while this one is consumed
                unsigned ha[16], ra[16], hb[16], rb[16];
                tmem_ld16(taddr, ha);
                tmem_ld16(taddr + MMA_BLOCK_SV, ra);
                auto consume = [&](const unsigned (&h)[16], const unsigned (&r)[16], int c0) {
                    float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float4 c = coef4[(c0 >> 2) + k];
                        const float2 t0 = __fadd2_rn(make_float2(__uint_as_float(h[4 * k]), __uint_as_float(h[4 * k + 1])),
                                                     make_float2(__uint_as_float(r[4 * k]), __uint_as_float(r[4 * k + 1])));
                        const float2 t1 = __fadd2_rn(make_float2(__uint_as_float(h[4 * k + 2]), __uint_as_float(h[4 * k + 3])),
                                                     make_float2(__uint_as_float(r[4 * k + 2]), __uint_as_float(r[4 * k + 3])));
                        a0 = __ffma2_rn(make_float2(ex2_approx(t0.x), ex2_approx(t0.y)), make_float2(c.x, c.y), a0);
                        a1 = __ffma2_rn(make_float2(ex2_approx(t1.x), ex2_approx(t1.y)), make_float2(c.z, c.w), a1);
                    }
                    acc64 += widen_f32((a0.x + a0.y) + (a1.x + a1.y));
                };
#pragma unroll
                for (int c0 = 0; c0 < MMA_BLOCK_SV; c0 += 32) {
                    tmem_ld_wait2(ha, ra);
                    tmem_ld16(taddr + c0 + 16, hb);
                    tmem_ld16(taddr + MMA_BLOCK_SV + c0 + 16, rb);
                    consume(ha, ra, c0);
                    tmem_ld_wait2(hb, rb);
                    if (c0 + 32 < MMA_BLOCK_SV) {
                        tmem_ld16(taddr + c0 + 32, ha);
                        tmem_ld16(taddr + MMA_BLOCK_SV + c0 + 32, ra);
                    }
                    consume(hb, rb, c0 + 16);
                }
                // this region may be overwritten by the MMAs of the block after next
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_ACCEMPTY + 2 * s + reg));
            }

            // ---- float64 epilogue of this row: Platt, clamp, coupling, combiner (the same code as the SIMT kernel) --------
            const long long row = tile * MMA_TILE_ROWS + row_in_tile;
            const double f_svc[1] = {acc64 + P.svc_intercept};
            const double zp[1] = {zpart0};
            double p1[1], zfin[1], p_svc[1];
            bool unc[1];
            finish_rows_fast<1>(P, true, f_svc, zp, p1, zfin, p_svc, unc);
            const bool valid = row < P.n;
            const unsigned char lab = zfin[0] > 0 ? 1 : 0;
            const bool fix = valid && (unc[0] || nonfinite);
            if (valid) {
                if (P.out) *reinterpret_cast<float2*>(P.out + 2 * row) = make_float2((float)(1.0 - p1[0]), (float)p1[0]);
                if (P.out64) { P.out64[2 * row] = 1.0 - p1[0]; P.out64[2 * row + 1] = p1[0]; }
                if (P.labels) P.labels[row] = lab;
                if (P.decision) P.decision[row] = zfin[0];
                if (P.meta) {
                    P.meta[row * P.n_members + P.slot_svc] = p_svc[0];
                    if (P.slot_gbt >= 0) P.meta[row * P.n_members + P.slot_gbt] = p_gbt;
                    if (P.slot_lin >= 0) P.meta[row * P.n_members + P.slot_lin] = p_lin;
                }
            }
            if (fix) {          // result hinges on a branch (or the row is not finite): the float64 fix-up kernel redoes it
                const unsigned at = atomicAdd(P.status + HFE_STATUS_LIST_LEN, 1u);
                if (at < (unsigned)HFE_WORKLIST_CAPACITY) P.status[HFE_STATUS_HEADER_WORDS + at] = (unsigned)row;
                else atomicOr(P.status, HFE_WORKLIST_OVERFLOW);
            }
            if (nonfinite && valid) atomicOr(P.status, HFE_ROW_NONFINITE);
            if (P.y && P.score) {                                       // accuracy count; flagged rows are counted by the fix-up
                unsigned hits = (valid && !fix && lab == P.y[row]) ? 1u : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) hits += __shfl_down_sync(0xFFFFFFFFu, hits, o);
                if (lane == 0 && hits) atomicAdd(P.score, (unsigned long long)hits);
            }
        }
    } else {
        // =========================== MMA issuers: one lane per stream =============================================================
        const int s = warp - (MMA_PRODUCER_WARPS + MMA_EPI_WARPS);
        if (lane == 0) {
            mbar_wait(bar(B_W), 0);
            // instruction descriptor (cute::UMMA::InstrDescriptor): D float32, A and B bfloat16, both K-major, N = 64, M = 128
            const unsigned idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(MMA_BLOCK_SV >> 3) << 17) | ((128u >> 4) << 24);
            const unsigned a_sbo = MMA_A_CHUNKS * 128, b_sbo = MMA_B_CHUNKS * 128;
            long long j = 0;
            for (long long i = s, tile = blockIdx.x + (long long)s * gridDim.x; tile < n_tiles; i += 2, tile += 2LL * gridDim.x, ++j) {
                const int stage = (int)(i % MMA_A_STAGES);
                mbar_wait(bar(B_AFULL + stage), (unsigned)((i / MMA_A_STAGES) & 1));
                const bool ext = (s_ext[stage * 4 + 0] | s_ext[stage * 4 + 1] | s_ext[stage * 4 + 2] | s_ext[stage * 4 + 3]) != 0;
                const unsigned a_base = smem_addr(s_a + stage * MMA_A_STAGE_BYTES);
#pragma unroll 1
                for (int blk = 0; blk < nb; ++blk) {
                    const long long use = j * nb + blk;
                    const int reg = (int)(use & 1);
                    mbar_wait(bar(B_ACCEMPTY + 2 * s + reg), (unsigned)(((use >> 1) & 1) ^ 1));
                    tc_fence_after();
                    const unsigned d_hi = tmem_base + (unsigned)((2 * s + reg) * (2 * MMA_BLOCK_SV)), d_rest = d_hi + MMA_BLOCK_SV;
                    const unsigned b_base = smem_addr(s_w) + (unsigned)(blk * (MMA_BLOCK_SV / 8)) * b_sbo;
                    auto a_desc = [&](int ks) { return umma_desc(a_base + ks * 256, 128, a_sbo); };
                    auto b_desc = [&](int chunk, int lbo_chunks) { return umma_desc(b_base + chunk * 128, lbo_chunks * 128, b_sbo); };
                    // A chunks  H0 H1 H2 0 | H0 H1 H2 H0 H1 H2        B chunks  WH0 WH1 WH2 0 | WM0 WM1 WM2 WL0 WL1 WL2
                    umma_bf16(d_hi, a_desc(0), b_desc(0, 1), idesc, 0u);          // (H0 H1).(WH0 WH1)
                    umma_bf16(d_hi, a_desc(1), b_desc(2, 1), idesc, 1u);          // (H2 0).(WH2 0)
                    umma_bf16(d_rest, a_desc(2), b_desc(4, 1), idesc, 0u);        // (H0 H1).(WM0 WM1)
                    umma_bf16(d_rest, a_desc(3), b_desc(6, 1), idesc, 1u);        // (H2 H0).(WM2 WL0)
                    umma_bf16(d_rest, a_desc(4), b_desc(8, 1), idesc, 1u);        // (H1 H2).(WL1 WL2)
                    if (ext) {
                        // extension chunks  M0 M1 M2 M0 M1 M2 L0 L1 L2 0  (mid and lo parts of the centred features): all of it
                        // is 2^-8 and below, so it joins the rest accumulator
                        umma_bf16(d_rest, a_desc(5), b_desc(0, 1), idesc, 1u);    // (M0 M1).(WH0 WH1)
                        umma_bf16(d_rest, a_desc(6), b_desc(2, 2), idesc, 1u);    // (M2 M0).(WH2 WM0)
                        umma_bf16(d_rest, a_desc(7), b_desc(5, 1), idesc, 1u);    // (M1 M2).(WM1 WM2)
                        umma_bf16(d_rest, a_desc(8), b_desc(0, 1), idesc, 1u);    // (L0 L1).(WH0 WH1)
                        umma_bf16(d_rest, a_desc(9), b_desc(2, 1), idesc, 1u);    // (L2 0).(WH2 0)
                    }
                    tc_commit(bar(B_ACCFULL + 2 * s + reg));
                }
                tc_commit(bar(B_AEMPTY + stage));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 12) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

}  // namespace hfe
