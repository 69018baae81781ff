// The RBF inner loop in isolation (development aid): which ingredient keeps the FMA pipe below its peak?
// Variants drop one ingredient at a time: MUFU.EX2, the float64 fold (F2F+DADD), the LDS.128 table reads.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int F = 17, STEPS = 18, R = 4, GROUPS = 110;

__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <bool MUFU, bool FOLD64, bool LDS, int RING, int ORDER = 0>
__global__ void __launch_bounds__(128, 4) k_loop(const float4* __restrict__ tabg, const float* __restrict__ zin, double* out, int groups, int reps) {
    extern __shared__ float4 s_sv[];
    for (int i = threadIdx.x; i < groups * STEPS + 4; i += blockDim.x) s_sv[i] = tabg[i];
    __syncthreads();
    float2 z2[F][R / 2];
    for (int f = 0; f < F; ++f) for (int p = 0; p < R / 2; ++p) z2[f][p] = make_float2(zin[(f * 2 + p) * 2 + (threadIdx.x & 1)], zin[(f * 2 + p) * 2 + 1]);
    double acc64[R] = {0, 0, 0, 0};
    float accf[R] = {0, 0, 0, 0};
    for (int rep = 0; rep < reps; ++rep) {
        const float4* tab = s_sv;
        float4 ring[RING];
#pragma unroll
        for (int i = 0; i < RING - 1; ++i) ring[i] = tab[i];
        float4 wreg = tabg[rep & 3];
#pragma unroll 1
        for (int g = 0; g < groups; ++g) {
            float2 t2[4][R / 2];
#pragma unroll
            for (int st = 0; st < STEPS; ++st) {
                float4 w4;
                if (LDS) { ring[(st + RING - 1) % RING] = tab[st + RING - 1]; w4 = ring[st % RING]; }
                else { w4 = wreg; wreg.x += 1e-9f; }
                const float w[4] = {w4.x, w4.y, w4.z, w4.w};
                if (st == 0) {
#pragma unroll
                    for (int p = 0; p < R / 2; ++p)
#pragma unroll
                        for (int q = 0; q < 4; ++q) t2[q][p] = __fmul2_rn(z2[0][p], make_float2(w[q], w[q]));
                } else if (st < F) {
                    if (ORDER == 0) {
#pragma unroll
                        for (int p = 0; p < R / 2; ++p)
#pragma unroll
                            for (int q = 0; q < 4; ++q) t2[q][p] = __ffma2_rn(z2[st][p], make_float2(w[q], w[q]), t2[q][p]);
                    } else {
#pragma unroll
                        for (int p = 0; p < R / 2; ++p)
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                unsigned long long& t = *reinterpret_cast<unsigned long long*>(&t2[q][p]);
                                const unsigned long long z = *reinterpret_cast<const unsigned long long*>(&z2[st][p]);
                                asm volatile("{\n.reg .b64 wb;\nmov.b64 wb, {%2, %2};\nfma.rn.f32x2 %0, %1, wb, %0;\n}" : "+l"(t) : "l"(z), "f"(w[q]));
                            }
                    }
                } else {
#pragma unroll
                    for (int p = 0; p < R / 2; ++p) {
                        float2 a2 = make_float2(0.f, 0.f);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float2 e = MUFU ? make_float2(ex2(t2[q][p].x), ex2(t2[q][p].y)) : __fmul2_rn(t2[q][p], make_float2(0.5f, 0.5f));
                            a2 = __ffma2_rn(e, make_float2(w[q], w[q]), a2);
                        }
                        if (FOLD64) { acc64[2 * p] += (double)a2.x; acc64[2 * p + 1] += (double)a2.y; }
                        else { accf[2 * p] += a2.x; accf[2 * p + 1] += a2.y; }
                    }
                }
            }
            tab += STEPS;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc64[0] + acc64[1] + acc64[2] + acc64[3] + accf[0] + accf[1] + accf[2] + accf[3];
}

// SV-paired formulation: the FFMA2 pair runs over two support vectors (natural .xy/.zw halves of the LDS.128), the row's
// z value is the broadcast scalar, so four consecutive FFMA2s (the four rows) share the 64-bit weight operand.
template <int RING, bool SNAKE, bool NOLDS = false, bool NOMUFU = false>
__global__ void __launch_bounds__(128, 4) k_loop_svpair(const float4* __restrict__ tabg, const float* __restrict__ zin, double* out, int groups, int reps) {
    extern __shared__ float4 s_sv[];
    for (int i = threadIdx.x; i < groups * STEPS + 4; i += blockDim.x) s_sv[i] = tabg[i];
    __syncthreads();
    float z[F][R];
    for (int f = 0; f < F; ++f) for (int r = 0; r < R; ++r) z[f][r] = zin[f * 4 + r + (threadIdx.x & 1)];
    double acc64[R] = {0, 0, 0, 0};
    for (int rep = 0; rep < reps; ++rep) {
        const float4* tab = s_sv;
        float4 ring[RING];
#pragma unroll
        for (int i = 0; i < RING - 1; ++i) ring[i] = tab[i];
        if (NOLDS) ring[RING - 1] = tab[RING - 1];
#pragma unroll 1
        for (int g = 0; g < groups; ++g) {
            float2 t2[2][R];          // [sv pair][row]
#pragma unroll
            for (int st = 0; st < STEPS; ++st) {
                if (!NOLDS) ring[(st + RING - 1) % RING] = tab[st + RING - 1];
                const float4 w4 = ring[st % RING];
                const float2 wp[2] = {make_float2(w4.x, w4.y), make_float2(w4.z, w4.w)};
                if (st == 0) {
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int r = 0; r < R; ++r) t2[h][r] = __fmul2_rn(wp[h], make_float2(z[0][r], z[0][r]));
                } else if (st < F) {
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int rr = 0; rr < R; ++rr) {
                            const int r = (SNAKE && h == 1) ? R - 1 - rr : rr;
                            t2[h][r] = __ffma2_rn(wp[h], make_float2(z[st][r], z[st][r]), t2[h][r]);
                        }
                } else {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float2 a2 = make_float2(0.f, 0.f);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            float2 e = NOMUFU ? __fmul2_rn(t2[h][r], make_float2(0.5f, 0.5f)) : make_float2(ex2(t2[h][r].x), ex2(t2[h][r].y));
                            a2 = __ffma2_rn(e, wp[h], a2);
                        }
                        acc64[r] += (double)(a2.x + a2.y);
                    }
                }
            }
            tab += STEPS;
            if (NOLDS) { ring[0].x += 1e-7f; ring[1].y += 1e-7f; ring[2].z += 1e-7f; }   // keep the group loop from being hoisted
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc64[0] + acc64[1] + acc64[2] + acc64[3];
}

// SV-paired formulation: the FFMA2 pair runs over two support vectors (natural .xy/.zw halves of the LDS.128), the row's
// z value is the broadcast scalar, so four consecutive FFMA2s (the four rows) share the 64-bit weight operand.
template <int RING, bool SNAKE>
__global__ void __launch_bounds__(128, 2) k_loop_svpair8(const float4* __restrict__ tabg, const float* __restrict__ zin, double* out, int groups, int reps) {
    extern __shared__ float4 s_sv[];
    for (int i = threadIdx.x; i < groups * STEPS + 4; i += blockDim.x) s_sv[i] = tabg[i];
    __syncthreads();
    constexpr int R = 8;
    float z[F][R];
    for (int f = 0; f < F; ++f) for (int r = 0; r < R; ++r) z[f][r] = zin[f * 8 + r + (threadIdx.x & 1)];
    double acc64[R] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int rep = 0; rep < reps; ++rep) {
        const float4* tab = s_sv;
        float4 ring[RING];
#pragma unroll
        for (int i = 0; i < RING - 1; ++i) ring[i] = tab[i];
#pragma unroll 1
        for (int g = 0; g < groups; ++g) {
            float2 t2[2][R];          // [sv pair][row]
#pragma unroll
            for (int st = 0; st < STEPS; ++st) {
                ring[(st + RING - 1) % RING] = tab[st + RING - 1];
                const float4 w4 = ring[st % RING];
                const float2 wp[2] = {make_float2(w4.x, w4.y), make_float2(w4.z, w4.w)};
                if (st == 0) {
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int r = 0; r < R; ++r) t2[h][r] = __fmul2_rn(wp[h], make_float2(z[0][r], z[0][r]));
                } else if (st < F) {
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int rr = 0; rr < R; ++rr) {
                            const int r = (SNAKE && h == 1) ? R - 1 - rr : rr;
                            t2[h][r] = __ffma2_rn(wp[h], make_float2(z[st][r], z[st][r]), t2[h][r]);
                        }
                } else {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float2 a2 = make_float2(0.f, 0.f);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            float2 e = make_float2(ex2(t2[h][r].x), ex2(t2[h][r].y));
                            a2 = __ffma2_rn(e, wp[h], a2);
                        }
                        acc64[r] += (double)(a2.x + a2.y);
                    }
                }
            }
            tab += STEPS;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc64[0] + acc64[1] + acc64[2] + acc64[3] + acc64[4] + acc64[5] + acc64[6] + acc64[7];
}

template <typename K>
void run(const char* name, K k, float4* tab, float* z, double* out) {
    const int blocks = 148 * 4, threads = 128, reps = 40;
    size_t smem = (GROUPS * STEPS + 4) * sizeof(float4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<<<blocks, threads, smem>>>(tab, z, out, GROUPS, 2); cudaDeviceSynchronize();
    cudaEventRecord(a); k<<<blocks, threads, smem>>>(tab, z, out, GROUPS, reps); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double ffma2 = (double)blocks * threads * reps * GROUPS * 144.0;        // FFMA2/FMUL2 per thread
    printf("%-34s %.3f ms  FMA pipe %.1f %% of 128 lanes/clk/SM @1.92GHz  (%s)\n", name, ms,
           100.0 * ffma2 * 2 / (ms * 1e-3) / 1.92e9 / 148 / 128, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    float4* tab; float* z; double* out;
    cudaMalloc(&tab, (GROUPS * STEPS + 8) * sizeof(float4)); cudaMemset(tab, 0, (GROUPS * STEPS + 8) * sizeof(float4));
    cudaMalloc(&z, 4096); cudaMemset(z, 0, 4096); cudaMalloc(&out, 148 * 4 * 128 * 8);
    run("full (MUFU + fold64 + LDS ring3)", k_loop<true, true, true, 3>, tab, z, out);
    run("no MUFU", k_loop<false, true, true, 3>, tab, z, out);
    run("no fold64 (f32 accumulate)", k_loop<true, false, true, 3>, tab, z, out);
    run("no LDS (weights in registers)", k_loop<true, true, false, 3>, tab, z, out);
    run("no MUFU, no fold64", k_loop<false, false, true, 3>, tab, z, out);
    run("no MUFU, no fold64, no LDS", k_loop<false, false, false, 3>, tab, z, out);
    run("full, asm-ordered (z shared x4)", k_loop<true, true, true, 3, 1>, tab, z, out);
    run("SV-paired, ring 3", k_loop_svpair<3, false>, tab, z, out);
    run("SV-paired, ring 3, snake", k_loop_svpair<3, true>, tab, z, out);
    run("SV-paired, ring 2", k_loop_svpair<2, false>, tab, z, out);
    run("SV-paired snake, NO LDS in loop", k_loop_svpair<3, true, true>, tab, z, out);
    run("SV-paired snake, NO LDS, NO MUFU", k_loop_svpair<3, true, true, true>, tab, z, out);
    run("SV-paired snake, LDS, NO MUFU", k_loop_svpair<3, true, false, true>, tab, z, out);
    {
        auto k = k_loop_svpair8<3, true>;
        const int blocks = 148 * 2, threads = 128, reps = 40;
        size_t smem = (GROUPS * STEPS + 4) * sizeof(float4);
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        k<<<blocks, threads, smem>>>(tab, z, out, GROUPS, 2); cudaDeviceSynchronize();
        cudaEventRecord(a); k<<<blocks, threads, smem>>>(tab, z, out, GROUPS, reps); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        double ffma2 = (double)blocks * threads * reps * GROUPS * 288.0;
        printf("%-34s %.3f ms  FMA pipe %.1f %%  (%s)\n", "SV-paired R=8, 8 warps/SM", ms, 100.0 * ffma2 * 2 / (ms * 1e-3) / 1.92e9 / 148 / 128, cudaGetErrorString(cudaGetLastError()));
    }
    run("full, ring 2", k_loop<true, true, true, 2>, tab, z, out);
    run("full, ring 4", k_loop<true, true, true, 4>, tab, z, out);
    return 0;
}
