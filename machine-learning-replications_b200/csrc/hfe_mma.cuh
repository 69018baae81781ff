// hfe_mma.cuh -- the RBF-SVC member on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// What libsvm computes per row (svm.cpp:461-515, 2821-2904) is f = sum_j coef_j * exp(-gamma |z - s_j|^2) + b.  With
// c = gamma * log2 e, a fixed centre m_d (the scaler mean rounded to an integer) and xc_d = x_d - m_d,
//     exp(-gamma |z - s_j|^2) = 2^(T_j) * 2^(B_j) * 2^(-c|z|^2)
//     T_j = sum_d xc_d * W_dj,   W_dj = 2c s_jd / scale_d,   B_j = sum_d (m_d - mean_d) W_dj - c|s_j|^2
// so the exponents of one tile are a small dense contraction T[128 rows x n_sv] = A[128 x K] * W[K x n_sv]; 2^(B_j) is
// folded into the coefficient and the row factor is applied once, in float64, to the finished sum.
// The operands are bfloat16 but nothing is rounded away: every W_dj is split into three bfloat16 parts (hi + mid + lo =
// 24 significant bits) and so is xc_d; products of 8-bit significands are exact in the float32 accumulator.  For
// cohorts in the load_data_public.py schema every feature is a small integer, xc_d is one exact bfloat16 and the base
// block (6 K-steps of 16) is all there is; when a tile holds a value that needs more bits the issuing thread puts the
// extension block (mid and lo parts of xc, 6 more K-steps) in front of it for that tile only.
// The tensor core aligns the products of one instruction and the accumulator to the largest exponent among them and
// drops what falls below ~2^-24 of it (measured with tools/mma_probe.cu).  Hence: the large constant terms (B_j, the
// row term) stay out of the accumulator, and K-steps run from the smallest parts to the largest, so each accumulation
// loses at most the float32 rounding of the exponent itself.
//
// One persistent CTA per SM, 16 warps (512 threads, 128 registers each):
//   warps 0-3   producers: load the tile's 128 rows (coalesced column reads, the next tile's loads in flight while this
//               one is processed), centre and split them, write the A operand into shared memory in the canonical
//               K-major core-matrix layout (five stages), evaluate the merged stump splits
//   warps 4-7   finishing warps: thread = row; when the RBF sums of four consecutive tiles have come back through the
//               shared-memory ring the thread finishes its four rows together -- links, Platt, clamp, coupling, combiner
//               (partial_logit_fast / finish_rows_fast, the SIMT kernel's code), four float64 chains interleaved
//   warps 8-11  exponent warps of stream 0, warps 12-15 of stream 1: thread = row (TMEM lane); tcgen05.ld 16 columns at a
//               time, ex2, coefficient FFMA2, two-float fold -- nothing else, so that the MUFU pipe always has work.
//               In turn the four warps of a stream also issue the stream's tcgen05.mma instructions (converged warp,
//               elect.sync): as soon as a region has been drained it is refilled with the block two ahead.
// Tiles alternate between the two streams.  Support vectors are taken in blocks of up to 128; TMEM (512 columns) is cut
// into four regions, two per stream: while a stream's warps drain one block from one region the tensor core fills the
// other, and the other stream keeps the MUFU pipe busy in the meantime.  The logistic member's dot product rides along as
// the last column of the B operand.  Both the MUFU pipe and the TMEM read path move 16 values per clock and SM: 28
// clocks per row for 448 columns is the floor of this design (10.4 G rows/s at 1965 MHz).
// What was measured on the way (B200, 10 M rows, published model; DESIGN.md section 10 has the table): MMAs issued from
// a divergent `if (lane == 0)` cost ~150 clocks each (operand broadcast loops); a fixed issuing warp makes its three
// peers wait for it every block; the float64 epilogue needs four warps of its own (two were the bottleneck, and inside
// the exponent warps it kept them off the MUFU pipe a quarter of the time).
#pragma once

#ifndef HFE_ISSUE_MODE
#define HFE_ISSUE_MODE 2          // what the issuing warp does when neither stream is ready: 0 spin, 1 nanosleep, 2 block on the stream further behind (measured: 1.584 / 1.581 / 1.548 ms)
#endif

#include <cuda_bf16.h>

#include "hfe_kernels.cuh"

namespace hfe {

constexpr int MMA_TILE_ROWS = 128;
constexpr int MMA_PRODUCER_WARPS = 4, MMA_FINISH_WARPS = 4, MMA_EPI_WARPS = 8;
constexpr int MMA_EPI_WARP0 = MMA_PRODUCER_WARPS + MMA_FINISH_WARPS;             // exponent warps 8..15
constexpr int MMA_ISSUE_WARP = MMA_PRODUCER_WARPS + MMA_FINISH_WARPS + MMA_EPI_WARPS;                     // warp 16: tcgen05.mma issue
constexpr int MMA_THREADS = 32 * (MMA_ISSUE_WARP + 4);          // 640: the issuing warp brings its warpgroup (three idle warps)
// Registers per thread and warpgroup (setmaxnreg; the kernel starts every thread at 96 = what 640 threads leave of 64 K):
// the float64 finishing code wants the most, the issuing warpgroup next to none.  What one group gives up another can take:
// the sum over the five groups must not exceed 5 x 96 (the pool is what the CTA was launched with; asking for more blocks forever).
#ifndef HFE_MMA_REG_P
#define HFE_MMA_REG_P 80
#define HFE_MMA_REG_F 144
#define HFE_MMA_REG_E 104
#define HFE_MMA_REG_I 48
#endif
constexpr int MMA_REGS[4] = {HFE_MMA_REG_P, HFE_MMA_REG_F, HFE_MMA_REG_E, HFE_MMA_REG_I};               // producers, finishers, exponent warps (each of two groups), issue group
static_assert(MMA_REGS[0] + MMA_REGS[1] + 2 * MMA_REGS[2] + MMA_REGS[3] <= 5 * 96, "register pool of the CTA");
constexpr int MMA_A_STAGES = 5;                          // two tiles in the tensor core / exponent warps, three produced ahead
constexpr int MMA_A_CHUNKS = 12;                         // 16-byte K chunks per row: H0 H1 H2 0 | M0 M1 M2 0 | L0 L1 L2 0
constexpr int MMA_A_STAGE_BYTES = (MMA_TILE_ROWS / 8) * MMA_A_CHUNKS * 128;                   // 24 576
constexpr int MMA_RING = 12;                             // tiles between a row's production and its float64 epilogue
constexpr int MMA_RING_SLOT_BYTES = MMA_TILE_ROWS * 20;  // per tile and row: RBF sum (f64), logistic dot, stump score, gamma|z|^2 (f32)
// MMA_REGIONS / MMA_REGION_COLS (hfe_model.cuh): TMEM regions per stream and their width
static_assert(MMA_A_STAGES * MMA_A_STAGE_BYTES == 5 * 24576 && MMA_RING * MMA_RING_SLOT_BYTES == 12 * 2560,
              "hfe_model.cuh sizes the support-vector limit of this kernel from these");
constexpr int MMA_TAIL_ROWS = 4;                         // tiles whose float64 epilogue one thread runs interleaved

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
__device__ __forceinline__ bool mbar_test(unsigned bar, unsigned parity) {       // has the phase with this parity completed?
    unsigned done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done != 0;
}
template <int REGS, bool INC>
__device__ __forceinline__ void set_max_regs() {          // all four warps of a warpgroup, converged
    static_assert(REGS % 8 == 0 && REGS >= 24 && REGS <= 256, "setmaxnreg operand");
    if (INC) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS));
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS));
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// tcgen05.commit / tcgen05.mma are executed by a converged warp; one elected lane issues.  (Issued from inside a divergent
// `if (lane == 0)` the compiler wraps every operand in an elect / broadcast loop to get it into uniform registers, which
// made six small MMAs cost ~900 clocks.)
__device__ __forceinline__ void tc_commit(unsigned bar) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "elect.sync _|q, 0xffffffff;\n"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
        "}\n" ::"r"(bar) : "memory");
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
        ".reg .pred p, q;\n"
        "elect.sync _|q, 0xffffffff;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
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

// (hi, lo) += v without rounding error (Knuth's TwoSum; lo collects what hi cannot hold)
__device__ __forceinline__ void two_sum_add(float& hi, float& lo, float v) {
    const float s = __fadd_rn(hi, v);
    const float z = __fsub_rn(s, hi);
    lo = __fadd_rn(lo, __fadd_rn(__fsub_rn(hi, __fsub_rn(s, z)), __fsub_rn(v, z)));
    hi = s;
}

#ifdef HFE_MMA_TRACE
// development builds: where the warps wait (clock cycles summed over lane 0 of the waiting warps)
__device__ unsigned long long g_mma_trace[32];
#define MMA_TRACE_WAIT(slot, stmt) do { const long long t0_ = clock64(); stmt; if (lane == 0) atomicAdd(&g_mma_trace[slot], (unsigned long long)(clock64() - t0_)); } while (0)
#else
#define MMA_TRACE_WAIT(slot, stmt) do { stmt; } while (0)
#endif

// ---- the kernel ------------------------------------------------------------------------------------------------------------
template <int TREE_MODE>
__global__ void __launch_bounds__(MMA_THREADS, 1)
hfe_mma_kernel(const __grid_constant__ ScoreParams P, const __grid_constant__ MmaParams Q) {
    static_assert(TREE_MODE == 0 || TREE_MODE == 1, "general tree tables keep the SIMT kernel");
    extern __shared__ __align__(1024) unsigned char smem[];
    // [W image + coef | A stages | ring: sums, logit shares, gamma|z|^2 | barriers]
    const int w_region = (Q.w_bytes + 1023) & ~1023;
    unsigned char* s_w = smem;
    unsigned char* s_a = smem + w_region;
    unsigned char* s_ring = s_a + MMA_A_STAGES * MMA_A_STAGE_BYTES;
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_ring + MMA_RING * MMA_RING_SLOT_BYTES);
    unsigned* s_misc = reinterpret_cast<unsigned*>(s_bar + MMA_BARRIERS);
    enum { B_W = 0, B_AFULL = 1, B_AEMPTY = B_AFULL + MMA_A_STAGES, B_ACCFULL = B_AEMPTY + MMA_A_STAGES, B_ACCEMPTY = B_ACCFULL + 2 * MMA_REGIONS,
           B_SUMFULL = B_ACCEMPTY + 2 * MMA_REGIONS, B_RINGFREE = B_SUMFULL + MMA_RING, B_COUNT = B_RINGFREE + MMA_RING };
    static_assert(B_COUNT <= MMA_BARRIERS, "barrier block");
    auto bar = [&](int i) { return smem_addr(s_bar + i); };
    auto ring_sum = [&](int slot) { return reinterpret_cast<double*>(s_ring + slot * MMA_RING_SLOT_BYTES); };
    auto ring_lin = [&](int slot) { return reinterpret_cast<float*>(s_ring + slot * MMA_RING_SLOT_BYTES + MMA_TILE_ROWS * 8); };
    auto ring_raw = [&](int slot) { return reinterpret_cast<float*>(s_ring + slot * MMA_RING_SLOT_BYTES + MMA_TILE_ROWS * 12); };
    auto ring_gzz = [&](int slot) { return reinterpret_cast<float*>(s_ring + slot * MMA_RING_SLOT_BYTES + MMA_TILE_ROWS * 16); };
    volatile unsigned* s_tmem = s_misc;                    // TMEM base address
    volatile unsigned* s_ext = s_misc + 4;                 // [stage][producer warp]: tile needs the extension block

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xFFFFFFFFu, tid >> 5, 0);          // tells the compiler that the warp index is warp-uniform
    const long long n_tiles = (P.n + MMA_TILE_ROWS - 1) / MMA_TILE_ROWS;
    const int nb = Q.nb, bsv = Q.block_sv;

    // ---- set-up -------------------------------------------------------------------------------------------------------------
    {   // zero the A stages once: the padding chunk of each K block is never written again
        uint4* z = reinterpret_cast<uint4*>(s_a);
        for (int i = tid; i < MMA_A_STAGES * MMA_A_STAGE_BYTES / 16; i += MMA_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (tid == 0) {
        mbar_init(bar(B_W), 1);
        for (int i = 0; i < MMA_A_STAGES; ++i) { mbar_init(bar(B_AFULL + i), 4); mbar_init(bar(B_AEMPTY + i), 1); }
        for (int i = 0; i < 2 * MMA_REGIONS; ++i) { mbar_init(bar(B_ACCFULL + i), 1); mbar_init(bar(B_ACCEMPTY + i), 4); }
        for (int i = 0; i < MMA_RING; ++i) { mbar_init(bar(B_SUMFULL + i), 4); mbar_init(bar(B_RINGFREE + i), MMA_FINISH_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == MMA_EPI_WARP0) {
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
        // =========================== producers ====================================================================================
        set_max_regs<MMA_REGS[0], MMA_REGS[0] >= 96>();
        const int prow = tid;                                  // this thread's row of every tile
        auto fetch = [&](long long tile, float (&x)[F]) {
            // this row's 17 features: one coalesced 128-byte read per warp and column, all in flight at once
            const long long row = tile * MMA_TILE_ROWS + prow;
            const long long row_ld = row < P.n ? row : 0;
#pragma unroll
            for (int f = 0; f < F; ++f) x[f] = __ldg(P.x + P.col_off[f] + row_ld);
            if (P.perm_col >= 0) {                             // a permutation-importance pass: one column through the index
                const int src = __ldg(P.perm_idx + row_ld);
                HFE_ASSERT_IN(src >= 0 && src < P.n);
                const float xp = __ldg(P.x + P.col_off[P.perm_col] + src);
#pragma unroll
                for (int f = 0; f < F; ++f) x[f] = (f == P.perm_col) ? xp : x[f];    // selects: x[] must stay in registers
            }
        };
        float xn[F];
        if ((long long)blockIdx.x < n_tiles) fetch(blockIdx.x, xn);
        long long i = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
            const int stage = (int)(i % MMA_A_STAGES);
            const long long row = tile * MMA_TILE_ROWS + prow;
            const bool valid = row < P.n;
            float x[F];
#pragma unroll
            for (int f = 0; f < F; ++f) x[f] = xn[f];
            if (tile + gridDim.x < n_tiles) fetch(tile + gridDim.x, xn);       // in flight while this tile is processed
            float probe = 0.f;
#pragma unroll
            for (int f = 0; f < F; ++f) probe = fmaf(x[f], 0.f, probe);          // NaN exactly when some x is NaN / Inf
            const bool nonfinite = valid && !(probe == 0.f);
            // centre; |z|^2 as sum_d xc (a xc + b) + c; the exponent of no support vector may leave the float32 2^x range:
            // |T_j| <= max |W| * sum_d |xc_d|
            float xc[F], sum_abs = 0.f, zz = Q.zz_c;
#pragma unroll
            for (int f = 0; f < F; ++f) {
                xc[f] = x[f] - Q.center[f];
                sum_abs += fabsf(xc[f]);
                zz = fmaf(xc[f], fmaf(Q.zz_a[f], xc[f], Q.zz_b[f]), zz);
            }
            const bool fix_row = valid && (nonfinite || !(sum_abs * Q.wmax_all <= 96.f));
            const bool keep = valid && !fix_row;
            // bfloat16 hi parts, two features per conversion; is there anything one bfloat16 cannot carry?
            unsigned hp[9];
#pragma unroll
            for (int k = 0; k < 8; ++k) hp[k] = pack_bf16(xc[2 * k], xc[2 * k + 1]);
            hp[8] = pack_bf16(xc[16], 0.f);
            unsigned rem_bits = 0;
#pragma unroll
            for (int f = 0; f < F; ++f) {
                const float hi = __uint_as_float((f & 1) ? (hp[f >> 1] & 0xFFFF0000u) : (hp[f >> 1] << 16));
                rem_bits |= __float_as_uint(xc[f] - hi);
            }
            const bool inexact = keep && (rem_bits << 1) != 0;
            // boosted stumps: the raw score goes to the finishing warps through the ring, the links are evaluated there
            float raw32 = 0.f;
            if (TREE_MODE == 1) {
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    if (x[f] <= P.split1_thr[f]) raw32 += P.split1_delta[f];
                    if ((Q.extra_split_mask >> f) & 1u) {              // further splits of this feature (warp-uniform, rare)
#pragma unroll 1
                        for (int sp = P.split_begin[f]; sp < P.split_begin[f + 1]; ++sp)
                            if (x[f] <= P.split_thr[sp]) raw32 += P.split_delta[sp];
                    }
                }
            }
            if (nonfinite) atomicOr(P.status, HFE_ROW_NONFINITE);
            // what the finishing warps need besides the RBF sum (they have released slot i % 8 of the ring after tile i - 8)
            MMA_TRACE_WAIT(12, mbar_wait(bar(B_RINGFREE + (int)(i % MMA_RING)), (unsigned)(((i / MMA_RING) & 1) ^ 1)));
            ring_raw((int)(i % MMA_RING))[prow] = raw32;
            ring_gzz((int)(i % MMA_RING))[prow] = fix_row ? -1.f : (float)P.gamma * zz;

            MMA_TRACE_WAIT(11, mbar_wait(bar(B_AEMPTY + stage), (unsigned)(((i / MMA_A_STAGES) & 1) ^ 1)));
            // A operand: [row group][chunk][row in group][8 bf16], chunks H0 H1 H2 0 | M0 M1 M2 0 | L0 L1 L2 0 (hi, mid, lo parts
            // of the centred features; the zero chunks are never written).  Only tiles holding a value that one bfloat16
            // cannot carry use the mid / lo chunks, but then the rows of all four warps must be there.
            unsigned char* a_row = s_a + stage * MMA_A_STAGE_BYTES + (prow >> 3) * (MMA_A_CHUNKS * 128) + (prow & 7) * 16;
            const unsigned km = keep ? 0xFFFFFFFFu : 0u;     // keeps NaN / Inf (and rows beyond n) out of the tensor core
            *reinterpret_cast<uint4*>(a_row + 0 * 128) = make_uint4(hp[0] & km, hp[1] & km, hp[2] & km, hp[3] & km);
            *reinterpret_cast<uint4*>(a_row + 1 * 128) = make_uint4(hp[4] & km, hp[5] & km, hp[6] & km, hp[7] & km);
            *reinterpret_cast<uint4*>(a_row + 2 * 128) = make_uint4(hp[8] & km, 0u, 0u, 0u);
            const bool ext = __any_sync(0xFFFFFFFFu, inexact);
            uint4 m0 = make_uint4(0, 0, 0, 0), m1 = m0, m2 = m0, l0 = m0, l1 = m0, l2 = m0;
            if (ext) {
                unsigned mp[9];
                float rem[F], lo[F];
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    const float hi = __uint_as_float((f & 1) ? (hp[f >> 1] & 0xFFFF0000u) : (hp[f >> 1] << 16));
                    rem[f] = keep ? (x[f] - Q.center[f]) - hi : 0.f;
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) mp[k] = pack_bf16(rem[2 * k], rem[2 * k + 1]);
                mp[8] = pack_bf16(rem[16], 0.f);
#pragma unroll
                for (int f = 0; f < F; ++f)
                    lo[f] = rem[f] - __uint_as_float((f & 1) ? (mp[f >> 1] & 0xFFFF0000u) : (mp[f >> 1] << 16));
                m0 = make_uint4(mp[0], mp[1], mp[2], mp[3]);
                m1 = make_uint4(mp[4], mp[5], mp[6], mp[7]);
                m2 = make_uint4(mp[8], 0u, 0u, 0u);
                l0 = make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
                l1 = make_uint4(pack_bf16(lo[8], lo[9]), pack_bf16(lo[10], lo[11]), pack_bf16(lo[12], lo[13]), pack_bf16(lo[14], lo[15]));
                l2 = make_uint4(pack_bf16(lo[16], 0.f), 0u, 0u, 0u);
            }
            *reinterpret_cast<uint4*>(a_row + 4 * 128) = m0; *reinterpret_cast<uint4*>(a_row + 5 * 128) = m1;
            *reinterpret_cast<uint4*>(a_row + 6 * 128) = m2;
            *reinterpret_cast<uint4*>(a_row + 8 * 128) = l0; *reinterpret_cast<uint4*>(a_row + 9 * 128) = l1;
            *reinterpret_cast<uint4*>(a_row + 10 * 128) = l2;
            if (lane == 0) s_ext[stage * 4 + warp] = ext ? 1u : 0u;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");           // these stores before the tensor core's reads
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_AFULL + stage));

        }
        // the last commits onto the A-stage barriers must have landed before the CTA may retire
        for (long long k = (i > MMA_A_STAGES ? i - MMA_A_STAGES : 0); k < i; ++k)
            mbar_wait(bar(B_AEMPTY + (int)(k % MMA_A_STAGES)), (unsigned)((k / MMA_A_STAGES) & 1));
    } else if (warp < MMA_EPI_WARP0) {
        // =========================== finishing warps: the float64 epilogue of every row ===========================================
        set_max_regs<MMA_REGS[1], MMA_REGS[1] >= 96>();
        // thread = row; the rows of four consecutive tiles are finished together, so four independent float64 chains
        // (Platt, clamp, coupling, links, combiner: the SIMT kernel's epilogue code) are interleaved
        constexpr int R = MMA_TAIL_ROWS;
        const bool has_lin = P.slot_lin >= 0;
        const int prow = tid - 32 * MMA_PRODUCER_WARPS;        // 0..127
        const long long my_tiles = (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
        for (long long t0 = 0; t0 < my_tiles; t0 += R) {
            const int count = (int)(my_tiles - t0 >= R ? R : my_tiles - t0);
            double f_svc[R], zpart[R], p1[R], zfin[R], p_svc[R], raw[R], zl[R], member[3][R];
            float gzz[R];
            long long rows[R];
            bool unc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const long long t = t0 + r;
                rows[r] = P.n;
                f_svc[r] = P.svc_intercept;
                raw[r] = zl[r] = 0.0;
                gzz[r] = 0.f;
                if (r < count) {
                    const int slot = (int)(t % MMA_RING);
                    MMA_TRACE_WAIT(13, mbar_wait(bar(B_SUMFULL + slot), (unsigned)((t / MMA_RING) & 1)));
                    rows[r] = ((long long)blockIdx.x + t * gridDim.x) * MMA_TILE_ROWS + prow;
                    gzz[r] = ring_gzz(slot)[prow];
                    raw[r] = P.gbt_split_base + (double)ring_raw(slot)[prow];
                    zl[r] = Q.lin_base + (double)ring_lin(slot)[prow];      // logistic member: the dot product came off the tensor core
                    f_svc[r] = fma(ring_sum(slot)[prow], exp_neg_fast(-(double)fmaxf(gzz[r], 0.f)), P.svc_intercept);
                }
            }
            __syncwarp();
            if (lane == 0) {                                    // the ring slots may be rewritten
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (r < count) mbar_arrive(bar(B_RINGFREE + (int)((t0 + r) % MMA_RING)));
            }
            partial_logit_fast<R>(P, TREE_MODE != 0, has_lin, raw, zl, zpart, member);
            finish_rows_fast<R>(P, true, f_svc, zpart, p1, zfin, p_svc, unc);
            unsigned hits = 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const long long row = rows[r];
                const bool valid = row < P.n;
                const unsigned char lab = zfin[r] > 0 ? 1 : 0;
                const bool fix = valid && (unc[r] || gzz[r] < 0.f);
                if (valid) {
                    if (P.out) *reinterpret_cast<float2*>(P.out + 2 * row) = make_float2((float)(1.0 - p1[r]), (float)p1[r]);
                    if (P.out64) { P.out64[2 * row] = 1.0 - p1[r]; P.out64[2 * row + 1] = p1[r]; }
                    if (P.labels) P.labels[row] = lab;
                    if (P.decision) P.decision[row] = zfin[r];
                    if (P.svc_dec) P.svc_dec[row] = f_svc[r];
                    if (P.meta) {
                        P.meta[row * P.n_members + P.slot_svc] = p_svc[r];
                        if (P.slot_gbt >= 0) P.meta[row * P.n_members + P.slot_gbt] = member[1][r];
                        if (P.slot_lin >= 0) P.meta[row * P.n_members + P.slot_lin] = member[2][r];
                    }
                }
                if (fix) {      // result hinges on a branch (or the row could not be scored here): the float64 fix-up redoes it
                    worklist_append(P, row);
                }
                if (P.y && P.score) hits += (valid && !fix && lab == P.y[row]) ? 1u : 0u;
            }
            if (P.y && P.score) {                                       // accuracy count; flagged rows are counted by the fix-up
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) hits += __shfl_down_sync(0xFFFFFFFFu, hits, o);
                if (lane == 0 && hits) atomicAdd(P.score, (unsigned long long)hits);
            }
        }
    } else if (warp >= MMA_ISSUE_WARP) {
        // =========================== MMA issue: one warp, converged, one elected lane ============================================
        set_max_regs<MMA_REGS[3], false>();               // the whole warpgroup gives its registers away; three warps then idle
        if (warp == MMA_ISSUE_WARP) {
        // Blocks of a stream's tiles are issued in order, each into the region the stream's four exponent warps have just
        // drained (the first two of a stream go into its still empty regions); the two streams are served as their regions
        // come free.  Kept apart from the exponent warps: issued by them in turn, the issuing warp sat 17 % of its time
        // waiting for its slowest peer and started its own next block late -- so the next issuer waited for it.
        mbar_wait(bar(B_W), 0);
        // instruction descriptor (cute::UMMA::InstrDescriptor): D float32, A and B bfloat16, both K-major, N = bsv, M = 128
        const unsigned idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(bsv >> 3) << 17) | ((128u >> 4) << 24);
        // descriptors are built once; per instruction only the 14-bit start-address field (16-byte units, low word) moves
        const unsigned a_sbo = MMA_A_CHUNKS * 128, b_sbo = MMA_B_CHUNKS * 128;
        const unsigned long long a_desc0 = umma_desc(smem_addr(s_a), 128, a_sbo);
        const unsigned long long b_desc0 = umma_desc(smem_addr(s_w), 128, b_sbo);
        const unsigned b_blk_units = (unsigned)(bsv / 8) * (b_sbo / 16);          // descriptor units per block of support vectors
        struct IssueState {
            long long tile;
            unsigned i, use, stage, phase;      // CTA-local tile number, blocks issued so far, A stage and the phase of its a_full
            int blk;
            bool ext;
        } st[2];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            st[s].tile = blockIdx.x + (long long)s * gridDim.x;
            st[s].i = (unsigned)s;
            st[s].use = 0;
            st[s].stage = (unsigned)s % MMA_A_STAGES;
            st[s].phase = 0;
            st[s].blk = 0;
            st[s].ext = false;
        }
        auto try_issue = [&](IssueState& q, const int s) -> bool {
            const unsigned reg = q.use % MMA_REGIONS;
            // the region must have been drained (block use - 2) and, for the first block of a tile, the A stage filled
            if (q.use >= MMA_REGIONS && !mbar_test(bar(B_ACCEMPTY + MMA_REGIONS * s + reg), ((q.use / MMA_REGIONS) - 1u) & 1u)) return false;
            if (q.blk == 0) {
                if (!mbar_test(bar(B_AFULL + q.stage), q.phase)) return false;
                q.ext = (s_ext[q.stage * 4 + 0] | s_ext[q.stage * 4 + 1] | s_ext[q.stage * 4 + 2] | s_ext[q.stage * 4 + 3]) != 0;
            }
            tc_fence_after();
            const unsigned d = tmem_base + ((unsigned)MMA_REGIONS * s + reg) * MMA_REGION_COLS;
            const unsigned long long a0 = a_desc0 + q.stage * (MMA_A_STAGE_BYTES / 16);
            const unsigned long long b0 = b_desc0 + (unsigned)q.blk * b_blk_units;
            // a K-step takes two adjacent 16-byte chunks of each operand (chunks are 128 bytes = 8 descriptor units apart).
            // A chunks  H0 H1 H2 0 | M0 M1 M2 0 | L0 L1 L2 0      B chunks  WL0 WL1 WL2 WM0 WM1 WM2 WH0 WH1 WH2 0
            // The second chunk of the (H2 0) steps is zero on the A side, so whatever follows on the B side does not matter.
            // Smallest products first.
            if (q.ext) {
                umma_bf16(d, a0 + 8 * 8, b0 + 6 * 8, idesc, 0u);              // (L0 L1).(WH0 WH1)        2^-16
                umma_bf16(d, a0 + 10 * 8, b0 + 8 * 8, idesc, 1u);             // (L2 0).(WH2 .)
                umma_bf16(d, a0 + 4 * 8, b0 + 3 * 8, idesc, 1u);              // (M0 M1).(WM0 WM1)        2^-16
                umma_bf16(d, a0 + 6 * 8, b0 + 5 * 8, idesc, 1u);              // (M2 0).(WM2 .)
                umma_bf16(d, a0 + 4 * 8, b0 + 6 * 8, idesc, 1u);              // (M0 M1).(WH0 WH1)        2^-8
                umma_bf16(d, a0 + 6 * 8, b0 + 8 * 8, idesc, 1u);              // (M2 0).(WH2 .)
            }
            umma_bf16(d, a0 + 0 * 8, b0 + 0 * 8, idesc, q.ext ? 1u : 0u);     // (H0 H1).(WL0 WL1)        2^-16
            umma_bf16(d, a0 + 2 * 8, b0 + 2 * 8, idesc, 1u);                  // (H2 0).(WL2 .)
            umma_bf16(d, a0 + 0 * 8, b0 + 3 * 8, idesc, 1u);                  // (H0 H1).(WM0 WM1)        2^-8
            umma_bf16(d, a0 + 2 * 8, b0 + 5 * 8, idesc, 1u);                  // (H2 0).(WM2 .)
            umma_bf16(d, a0 + 0 * 8, b0 + 6 * 8, idesc, 1u);                  // (H0 H1).(WH0 WH1)        1
            umma_bf16(d, a0 + 2 * 8, b0 + 8 * 8, idesc, 1u);                  // (H2 0).(WH2 .)
            tc_commit(bar(B_ACCFULL + MMA_REGIONS * s + reg));
            ++q.use;
            if (++q.blk == nb) {
                tc_commit(bar(B_AEMPTY + q.stage));      // every MMA that read this A stage came from this thread: one commit covers them
                q.blk = 0;
                q.tile += 2LL * gridDim.x;
                // the stream's next tile is two further on in the CTA's sequence
                q.i += 2;
                q.stage = q.i % MMA_A_STAGES;
                q.phase = (q.i / MMA_A_STAGES) & 1u;
            }
            return true;
        };
        long long idle_since = 0;
        unsigned idle = 0;
        while (st[0].tile < n_tiles || st[1].tile < n_tiles) {
            bool any = false;
#pragma unroll
            for (int s = 0; s < 2; ++s)
                if (st[s].tile < n_tiles) any = try_issue(st[s], s) || any;
            if (any) { idle = 0; idle_since = 0; continue; }
#if HFE_ISSUE_MODE == 1
            __nanosleep(32);
#elif HFE_ISSUE_MODE == 2
            {   // block (hardware-suspended try_wait) on what the stream that is further behind needs next
                const int s = (st[1].tile < n_tiles && (st[0].tile >= n_tiles || st[1].use < st[0].use)) ? 1 : 0;
                const IssueState& q = s ? st[1] : st[0];
                if (q.use >= MMA_REGIONS) mbar_wait(bar(B_ACCEMPTY + MMA_REGIONS * s + (q.use % MMA_REGIONS)), ((q.use / MMA_REGIONS) - 1u) & 1u);
            }
#endif
            if ((++idle & 0xFFFFu) == 0) {               // about a second without progress can only be a protocol bug
                if (idle_since == 0) idle_since = clock64();
                else if (clock64() - idle_since > 4000000000LL) __trap();
            }
        }
        }
    } else {
        // =========================== exponent warps ================================================================================
        set_max_regs<MMA_REGS[2], MMA_REGS[2] >= 96>();
        const int s = (warp - MMA_EPI_WARP0) >> 2;                   // stream
        const int row_in_tile = (warp & 3) * 32 + lane;              // TMEM lane = row of the tile (warp % 4 picks the lane quarter)
        const unsigned lane_base = tmem_base + ((unsigned)((warp & 3) * 32) << 16);
        mbar_wait(bar(B_W), 0);
        const float* s_coef = reinterpret_cast<const float*>(s_w + (size_t)nb * bsv * (MMA_B_CHUNKS * 16));
        const int steps = bsv >> 4;
#ifdef HFE_MMA_TRACE
        const long long t_loop0 = clock64();
#endif
        long long j = 0;
        for (long long i = s, tile = blockIdx.x + (long long)s * gridDim.x; tile < n_tiles; i += 2, tile += 2LL * gridDim.x, ++j) {
            float acc_hi = 0.f, acc_lo = 0.f; // two-float accumulator (error-free TwoSum): keeps the float64 pipe out of the hot loop
            unsigned lin_dot = 0;             // last column of the last block: x . (logistic coefficients), not an exponent
#pragma unroll 1
            for (int blk = 0; blk < nb; ++blk) {
                const long long use = j * nb + blk;                  // this stream's running count of blocks
                const int reg = (int)(use % MMA_REGIONS);
                MMA_TRACE_WAIT(blk < 7 ? blk : 7, mbar_wait(bar(B_ACCFULL + MMA_REGIONS * s + reg), (unsigned)((use / MMA_REGIONS) & 1)));
                tc_fence_after();
                const unsigned taddr = lane_base + (unsigned)((MMA_REGIONS * s + reg) * MMA_REGION_COLS);
                const float4* coef4 = reinterpret_cast<const float4*>(s_coef + blk * bsv);
                // 16 support vectors per step: 2^T on the MUFU pipe, coefficient FFMA2 (two vectors per instruction), float32
                // partial sums of 4 vectors each folded into the float64 accumulator; the TMEM load of the next step is in
                // flight while this one is consumed
                unsigned va[16], vb[16];
                auto consume = [&](const unsigned (&v)[16], const float4* cp) -> float {
                    float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float4 c = cp[k];
                        const float2 e0 = make_float2(ex2_approx(__uint_as_float(v[4 * k])), ex2_approx(__uint_as_float(v[4 * k + 1])));
                        const float2 e1 = make_float2(ex2_approx(__uint_as_float(v[4 * k + 2])), ex2_approx(__uint_as_float(v[4 * k + 3])));
                        a0 = __ffma2_rn(e0, make_float2(c.x, c.y), a0);
                        a1 = __ffma2_rn(e1, make_float2(c.z, c.w), a1);
                    }
                    return (a0.x + a0.y) + (a1.x + a1.y);
                };
                tmem_ld16(taddr, va);
                int st = 0;
#pragma unroll 1
                for (; st + 2 <= steps; st += 2) {
                    tmem_ld_wait(va);
                    tmem_ld16(taddr + (st + 1) * 16, vb);
                    const float pa = consume(va, coef4 + st * 4);
                    tmem_ld_wait(vb);
                    if (st + 2 < steps) tmem_ld16(taddr + (st + 2) * 16, va);
                    const float pb = consume(vb, coef4 + (st + 1) * 4);
                    two_sum_add(acc_hi, acc_lo, pa + pb);            // one error-free fold per 32 support vectors
                }
                if (st < steps) {
                    tmem_ld_wait(va);
                    two_sum_add(acc_hi, acc_lo, consume(va, coef4 + st * 4));
                    lin_dot = va[15];
                } else {
                    lin_dot = vb[15];
                }
                // this region may be overwritten: once all four warps have said so, the issuing warp refills it with the block
                // two ahead
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_ACCEMPTY + MMA_REGIONS * s + reg));
            }
            // hand the row's sum to the producer thread that finishes it
            ring_sum((int)(i % MMA_RING))[row_in_tile] = widen_f32(acc_hi) + widen_f32(acc_lo);
            ring_lin((int)(i % MMA_RING))[row_in_tile] = __uint_as_float(lin_dot);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_SUMFULL + (int)(i % MMA_RING)));
        }
#ifdef HFE_MMA_TRACE
        if (lane == 0) atomicAdd(&g_mma_trace[14], (unsigned long long)(clock64() - t_loop0));
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_EPI_WARP0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

}  // namespace hfe
