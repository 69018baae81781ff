// mma_probe.cu -- bring-up check for the tcgen05 path of the RBF contraction (development aid, not shipped).
// One CTA: A [128 x K] and B [N x K] bf16, K-major, no swizzle, core matrices of 8 rows x 16 bytes laid out as
// [row group][K chunk][8 rows][16 B]; D = A * B^T accumulated in TMEM by tcgen05.mma.kind::f16, read back with
// tcgen05.ld.32x32b and compared with a float64 product on the host.  Also reports how the tensor core accumulates
// (exact-product terms of very different magnitude) -- the precision question of DESIGN.md section 10.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o mma_probe mma_probe.cu && ./mma_probe [N] [ksteps] [mode]
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 2; } } while (0)

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ unsigned long long make_desc(unsigned addr, unsigned lbo, unsigned sbo) {
    unsigned long long d = 0;
    d |= (unsigned long long)((addr >> 4) & 0x3FFF);
    d |= (unsigned long long)((lbo >> 4) & 0x3FFF) << 16;
    d |= (unsigned long long)((sbo >> 4) & 0x3FFF) << 32;
    d |= 1ull << 46;                        // descriptor version (Blackwell)
    return d;                               // base offset 0, layout type 0 = no swizzle
}

__global__ void __launch_bounds__(128) probe_kernel(const unsigned char* a_img, const unsigned char* b_img, float* d_out,
                                                    int n_cols, int ksteps, int nc, int mode, int* err, int reps, long long* cycles) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long mbar;
    __shared__ unsigned tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int a_bytes = 16 * nc * 128, b_bytes = (n_cols / 8) * nc * 128;
    unsigned char* sa = smem;
    unsigned char* sb = smem + ((a_bytes + 1023) & ~1023);
    for (int i = tid * 16; i < a_bytes; i += 128 * 16) *reinterpret_cast<uint4*>(sa + i) = *reinterpret_cast<const uint4*>(a_img + i);
    for (int i = tid * 16; i < b_bytes; i += 128 * 16) *reinterpret_cast<uint4*>(sb + i) = *reinterpret_cast<const uint4*>(b_img + i);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");          // generic-proxy smem writes -> async proxy (MMA)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tbase = tmem_base;
    if (tid == 0) {
        // instruction descriptor: D f32, A/B bf16, both K-major, N, M = 128
        const unsigned idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(n_cols >> 3) << 17) | ((128u >> 4) << 24);
        const unsigned lbo = mode == 0 ? 128u : (unsigned)nc * 128u;
        const unsigned sbo = mode == 0 ? (unsigned)nc * 128u : 128u;
        const long long t_start = clock64();
        for (int rep = 0; rep < reps; ++rep)
        for (int s = 0; s < ksteps; ++s) {
            const unsigned long long da = make_desc(smem_u32(sa) + s * 256, lbo, sbo);
            const unsigned long long db = make_desc(smem_u32(sb) + s * 256, lbo, sbo);
            const unsigned acc = (s > 0 || rep > 0) ? 1u : 0u;
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "setp.ne.b32 p, %4, 0;\n"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
                "}\n" ::"r"(tbase), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
        cycles[0] = clock64() - t_start;          // issue time
        unsigned done = 0;
        while (!done)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smem_u32(&mbar)) : "memory");
        cycles[1] = clock64() - t_start;          // issue -> completion visible on the mbarrier
    }
    // wait for the MMAs (bounded spin: a wrong descriptor must not hang the box)
    {
        unsigned done = 0;
        for (long long spin = 0; spin < 20000000 && !done; ++spin) {
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n"
                "selp.u32 %0, 1, 0, p;\n"
                "}\n" : "=r"(done) : "r"(smem_u32(&mbar)) : "memory");
        }
        if (!done) { if (tid == 0) *err = 1; }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // TMEM -> registers: warp w reads lanes 32w..32w+31, 16 columns at a time
    const int row = tid;
    for (int c0 = 0; c0 < n_cols; c0 += 16) {
        unsigned v[16];
        const unsigned taddr = tbase + ((unsigned)(warp * 32) << 16) + (unsigned)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 16; ++j) d_out[row * n_cols + c0 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(512));
}

static unsigned short bf16_bits(float f) {
    __nv_bfloat16 h = __float2bfloat16_rn(f);
    unsigned short b;
    std::memcpy(&b, &h, 2);
    return b;
}
static float bf16_val(float f) { return __bfloat162float(__float2bfloat16_rn(f)); }

int main(int argc, char** argv) {
    const int n_cols = argc > 1 ? atoi(argv[1]) : 112;
    const int ksteps = argc > 2 ? atoi(argv[2]) : 5;
    const int mode = argc > 3 ? atoi(argv[3]) : 0;
    const int test = argc > 4 ? atoi(argv[4]) : 0;          // 0: random small values; 1: accumulation-precision pattern
    const int reps = argc > 5 ? atoi(argv[5]) : 1;
    const int nc = 2 * ksteps, K = 8 * nc;
    std::vector<float> A(128 * K), B((size_t)n_cols * K);
    srand(7);
    for (auto& v : A) v = bf16_val((float)((rand() % 17) - 8) * 0.25f);
    for (auto& v : B) v = bf16_val((float)((rand() % 33) - 16) * 0.125f);
    if (test == 1) {
        // one big term and many small ones per output: does the accumulator keep the small terms' low bits?
        for (int r = 0; r < 128; ++r)
            for (int k = 0; k < K; ++k) A[r * K + k] = bf16_val(k == 0 ? 96.f + r : (float)((r * 7 + k * 3) % 13 + 1));
        for (int c = 0; c < n_cols; ++c)
            for (int k = 0; k < K; ++k) B[(size_t)c * K + k] = bf16_val(k == 0 ? 1.5f : ldexpf((float)((c + k) % 7 + 1), -18 - (k % 5)));
    }
    auto image = [&](const std::vector<float>& M, int rows) {
        std::vector<unsigned char> img((size_t)(rows / 8) * nc * 128, 0);
        for (int r = 0; r < rows; ++r)
            for (int k = 0; k < K; ++k) {
                size_t off = (size_t)(r / 8) * nc * 128 + (size_t)(k / 8) * 128 + (r % 8) * 16 + (k % 8) * 2;
                unsigned short b = bf16_bits(M[(size_t)r * K + k]);
                std::memcpy(&img[off], &b, 2);
            }
        return img;
    };
    auto ia = image(A, 128), ib = image(B, n_cols);
    unsigned char *da, *db;
    float* dd;
    int* derr;
    long long* dcyc;
    CK(cudaMalloc(&da, ia.size()));
    CK(cudaMalloc(&db, ib.size()));
    CK(cudaMalloc(&dd, sizeof(float) * 128 * n_cols));
    CK(cudaMalloc(&derr, 4));
    CK(cudaMalloc(&dcyc, 16));
    CK(cudaMemset(derr, 0, 4));
    CK(cudaMemset(dd, 0xFF, sizeof(float) * 128 * n_cols));
    CK(cudaMemcpy(da, ia.data(), ia.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db, ib.data(), ib.size(), cudaMemcpyHostToDevice));
    size_t smem = ((ia.size() + 1023) & ~1023) + ib.size() + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_kernel<<<1, 128, smem>>>(da, db, dd, n_cols, ksteps, nc, mode, derr, reps, dcyc);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    int err = 0;
    CK(cudaMemcpy(&err, derr, 4, cudaMemcpyDeviceToHost));
    long long cyc[2];
    CK(cudaMemcpy(cyc, dcyc, 16, cudaMemcpyDeviceToHost));
    printf("reps=%d: issue %lld clk, issue->complete %lld clk = %.1f clk per MMA instruction\n", reps, cyc[0], cyc[1], (double)cyc[1] / (reps * ksteps));
    std::vector<float> D(128 * n_cols);
    CK(cudaMemcpy(D.data(), dd, D.size() * 4, cudaMemcpyDeviceToHost));
    double max_abs = 0, max_rel = 0;
    int bad = 0;
    for (int r = 0; r < 128; ++r)
        for (int c = 0; c < n_cols; ++c) {
            double ref = 0;
            for (int k = 0; k < K; ++k) ref += (double)A[r * K + k] * (double)B[(size_t)c * K + k];
            double e = std::fabs((double)D[r * n_cols + c] - ref);
            max_abs = std::max(max_abs, e);
            max_rel = std::max(max_rel, e / std::max(std::fabs(ref), 1e-30));
            if (!(e <= 1e-3 * std::max(1.0, std::fabs(ref)))) ++bad;
        }
    printf("N=%d ksteps=%d mode=%d test=%d timeout=%d  max_abs_err=%.3e max_rel_err=%.3e (fp32 eps 5.96e-8) mismatches=%d  D[0][0..3]=%g %g %g %g  D[127][N-1]=%g\n",
           n_cols, ksteps, mode, test, err, max_abs, max_rel, bad, D[0], D[1], D[2], D[3], D[127 * n_cols + n_cols - 1]);
    return bad == 0 && err == 0 ? 0 : 1;
}
