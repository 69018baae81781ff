// hfe_api.cu -- the extern "C" boundary declared in include/hfe.h.  Plain pointers and sizes in, status codes out.
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <new>

#include "hfe_kernels.cuh"
#include "hfe_mma.cuh"
#include "hfe_forest.cuh"
#include "hfe_host.cuh"

using namespace hfe;

struct hfe_model {
    Model* impl;
};

namespace {

// SM count of the current device (entry points without a model handle run on whatever device is current).
int sm_count_current() {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) {
        cudaGetLastError();
        return 148;
    }
    return sms;
}

int check_matrix(const hfe_model* model, const float* x, int64_t n, int64_t ld) {
    if (!model || !model->impl) return fail(HFE_ERR_INVALID_ARGUMENT, "null model handle");
    if (n < 0) return fail(HFE_ERR_INVALID_ARGUMENT, "negative row count");
    if (n > 0 && !x) return fail(HFE_ERR_INVALID_ARGUMENT, "null feature matrix");
    if (ld < n) return fail(HFE_ERR_INVALID_ARGUMENT, "leading dimension smaller than the row count");
    if ((ld & 3) || (reinterpret_cast<uintptr_t>(x) & 15))
        return fail(HFE_ERR_INVALID_ARGUMENT, "feature matrix must be 16-byte aligned with ld % 4 == 0");
    if (n >= ((int64_t)1 << 32)) return fail(HFE_ERR_TOO_LARGE, "at most 2^32 - 1 rows per call");
    return HFE_OK;
}

// layout of the mapped pinned block of the few-row host path
constexpr size_t TINY_OFF_X = 0;
constexpr size_t TINY_OFF_OUT = TINY_OFF_X + (size_t)TINY_HOST_ROWS * F * sizeof(double);
constexpr size_t TINY_OFF_DECISION = TINY_OFF_OUT + (size_t)TINY_HOST_ROWS * 2 * sizeof(double);
constexpr size_t TINY_OFF_META = TINY_OFF_DECISION + (size_t)TINY_HOST_ROWS * sizeof(double);
constexpr size_t TINY_OFF_LABELS = TINY_OFF_META + (size_t)TINY_HOST_ROWS * MAX_MEMBERS * sizeof(double);
constexpr size_t TINY_OFF_STATUS = TINY_OFF_LABELS + (((size_t)TINY_HOST_ROWS + 63) & ~(size_t)63);
constexpr size_t TINY_BYTES = TINY_OFF_STATUS + 64;

template <bool SVC, int TM, bool PFX>
int launch_score(const Model* m, const ScoreParams& P, cudaStream_t stream) {
    if (SVC && P.sv_bytes > MAX_SV_TABLE_BYTES)
        return fail(HFE_ERR_UNSUPPORTED, "the support-vector table does not fit one block's shared memory");
    size_t smem = (SVC ? (size_t)P.sv_bytes : 0) + (TM == 2 ? (size_t)F * BLOCK_ROWS * sizeof(float) : 0);
    auto kernel = hfe_score_kernel<SVC, TM, PFX>;
    // per device and shared-memory size: opt in to large dynamic shared memory, ask how many blocks stay resident
    static std::mutex mu;
    static std::map<std::pair<int, size_t>, int> resident;
    int per_sm = 0, sms = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        auto key = std::make_pair(m->device, smem);
        auto it = resident.find(key);
        if (it == resident.end()) {
            HFE_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            HFE_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, BLOCK_THREADS, smem));
            if (per_sm < 1) return fail(HFE_ERR_UNSUPPORTED, "model tables do not fit one SM's shared memory");
            resident[key] = per_sm;
        } else {
            per_sm = it->second;
        }
    }
    HFE_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, m->device));
    // persistent blocks: every warp walks its own sequence of 128-row tiles
    long long tiles = (P.n + WARP_ROWS - 1) / WARP_ROWS;
    long long blocks = (tiles + BLOCK_WARPS - 1) / BLOCK_WARPS;
    if (HFE_PERSISTENT || !SVC) blocks = std::min<long long>(blocks, (long long)sms * per_sm);
    kernel<<<(unsigned)blocks, BLOCK_THREADS, smem, stream>>>(P, P.sv_bytes);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

// Tensor map of the caller's column-major matrix for the streaming pass: dimension 0 = rows (contiguous, extent n so
// that the copy engine zero-fills the ragged tail), dimension 1 = the 17 feature columns (stride ld floats), box =
// one warp tile.  cuTensorMapEncodeTiled is a host-only driver call; it is looked up through the runtime so that the
// library does not link against libcuda (it must still load on a machine without a driver).
int make_x_map(const ScoreParams& P, CUtensorMap* map) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            encode = reinterpret_cast<EncodeFn>(fn);
    });
    if (!encode) return fail(HFE_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[2] = {(cuuint64_t)P.n, (cuuint64_t)F};
    const cuuint64_t strides[1] = {(cuuint64_t)P.ld * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)WARP_ROWS, (cuuint32_t)F};
    const cuuint32_t elem[2] = {1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(P.x), dims, strides, box, elem,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(HFE_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    return HFE_OK;
}

template <int TM>
int launch_stream(const Model* m, const ScoreParams& P, cudaStream_t stream) {
    CUtensorMap x_map;
    int rc = make_x_map(P, &x_map);
    if (rc) return rc;
    auto kernel = hfe_stream_kernel<TM>;
    static std::mutex mu;
    static std::map<int, int> resident;
    int per_sm = 0, sms = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = resident.find(m->device);
        if (it == resident.end()) {
            HFE_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STREAM_SMEM_BYTES));
            HFE_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, BLOCK_THREADS, STREAM_SMEM_BYTES));
            if (per_sm < 1) return fail(HFE_ERR_UNSUPPORTED, "streaming pass does not fit one SM's shared memory");
            resident[m->device] = per_sm;
        } else {
            per_sm = it->second;
        }
    }
    HFE_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, m->device));
    long long tiles = (P.n + WARP_ROWS - 1) / WARP_ROWS;
    long long blocks = std::min<long long>((tiles + BLOCK_WARPS - 1) / BLOCK_WARPS, (long long)sms * per_sm);
    kernel<<<(unsigned)blocks, BLOCK_THREADS, STREAM_SMEM_BYTES, stream>>>(P, x_map);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

// The RBF sum on the tensor cores (hfe_mma.cuh): one persistent CTA per SM.
template <int TM>
int launch_mma(const Model* m, const ScoreParams& P, cudaStream_t stream) {
    const MmaParams& Q = m->mma;
    const size_t smem = (((size_t)Q.w_bytes + 1023) & ~(size_t)1023) + (size_t)MMA_A_STAGES * MMA_A_STAGE_BYTES +
                        (size_t)MMA_RING * MMA_RING_SLOT_BYTES + MMA_BARRIERS * sizeof(unsigned long long) + 128;
    auto kernel = hfe_mma_kernel<TM>;
    static std::mutex mu;
    static std::map<int, int> ready;
    int sms = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!ready.count(m->device)) {
            // the warpgroups re-deal their registers with setmaxnreg; what they ask for in all must fit what the CTA is
            // launched with, or the first setmaxnreg.inc would wait forever -- refuse instead
            cudaFuncAttributes attr{};
            HFE_CUDA_TRY(cudaFuncGetAttributes(&attr, kernel));
            if (attr.numRegs * 5 < MMA_REGS[0] + MMA_REGS[1] + 2 * MMA_REGS[2] + MMA_REGS[3])
                return fail(HFE_ERR_UNSUPPORTED, "tensor-core kernel was built with fewer registers per thread than its warpgroups re-deal");
            HFE_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
            ready[m->device] = 1;
        }
    }
    HFE_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, m->device));
    const long long tiles = (P.n + MMA_TILE_ROWS - 1) / MMA_TILE_ROWS;
    const long long blocks = std::min<long long>(tiles, sms);
    kernel<<<(unsigned)blocks, MMA_THREADS, smem, stream>>>(P, Q);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

// General boosted trees out of shared memory (hfe_forest.cuh).
int launch_forest(const Model* m, const ScoreParams& P, cudaStream_t stream) {
    auto kernel = hfe_forest_kernel;
    static std::mutex mu;
    static std::map<int, int> ready;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!ready.count(m->device)) {
            HFE_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FOREST_SMEM_BYTES));
            ready[m->device] = 1;
        }
    }
    const long long blocks = (P.n + FOREST_ROWS - 1) / FOREST_ROWS;
    kernel<<<(unsigned)blocks, FOREST_THREADS, FOREST_SMEM_BYTES, stream>>>(P, m->forest);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

int dispatch_score_inner(const Model* m, const ScoreParams& P, cudaStream_t stream);

// Measurement aid (tree_engine 3): the L1 walk with the 8-byte node table pinned in an L2 persisting window for the
// duration of the launch (the stream's access-policy window is set before and cleared after it).
int dispatch_score(const Model* m, const ScoreParams& P, cudaStream_t stream) {
    if (m->tree_mode == 2 && m->tree_engine == 3 && P.nodes) {
        cudaStreamAttrValue attr{};
        attr.accessPolicyWindow.base_ptr = const_cast<uint2*>(P.nodes);
        attr.accessPolicyWindow.num_bytes = std::min<size_t>((size_t)m->n_nodes * sizeof(uint2), 32u << 20);
        attr.accessPolicyWindow.hitRatio = 1.0f;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        HFE_CUDA_TRY(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 32u << 20));
        HFE_CUDA_TRY(cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr));
        int rc = dispatch_score_inner(m, P, stream);
        attr.accessPolicyWindow.num_bytes = 0;
        cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr);
        return rc;
    }
    return dispatch_score_inner(m, P, stream);
}

int dispatch_score_inner(const Model* m, const ScoreParams& P, cudaStream_t stream) {
    bool svc = P.slot_svc >= 0;
    if (!svc && m->tree_mode == 2 && m->forest.n_chunks > 0 && (m->tree_engine == 0 || m->tree_engine == 2)) {
        return launch_forest(m, P, stream);
    }
    if (m->tree_mode == 2 && m->tree_engine == 2)
        return fail(HFE_ERR_UNSUPPORTED, "the shared-memory tree walk needs a stack without an SVC member, at most 256 "
                                         "distinct (feature, threshold) pairs (255 per feature) and trees of at most 8 188 nodes");
    if (svc && m->mma.nb > 0 && m->tree_mode != 2 && m->rbf_engine != 1)
        return m->tree_mode == 0 ? launch_mma<0>(m, P, stream) : launch_mma<1>(m, P, stream);
    if (svc && m->rbf_engine == 2)
        return fail(HFE_ERR_UNSUPPORTED, "the tensor-core RBF path needs at most ~460 support vectors (shared memory) and no general tree tables");
    bool identity = true;                          // the feature order is only changed for stacks with an SVC member
    for (int f = 0; f < F; ++f) identity = identity && m->col_of[f] == f;
    // (tensor-copy coordinates are signed 32-bit: beyond 2^31 rows the register-window instance takes over)
    if (HFE_STREAM_TMA && !svc && m->tree_mode != 2 && identity && P.n < (1LL << 31))
        return m->tree_mode == 0 ? launch_stream<0>(m, P, stream) : launch_stream<1>(m, P, stream);
    if (svc && P.sv_quads > 0) {
        switch (m->tree_mode) {
            case 0: return launch_score<true, 0, true>(m, P, stream);
            case 1: return launch_score<true, 1, true>(m, P, stream);
            default: return launch_score<true, 2, true>(m, P, stream);
        }
    }
    switch (m->tree_mode * 2 + (svc ? 1 : 0)) {
        case 0: return launch_score<false, 0, false>(m, P, stream);
        case 1: return launch_score<true, 0, false>(m, P, stream);
        case 2: return launch_score<false, 1, false>(m, P, stream);
        case 3: return launch_score<true, 1, false>(m, P, stream);
        case 4: return launch_score<false, 2, false>(m, P, stream);
        default: return launch_score<true, 2, false>(m, P, stream);
    }
}

int score_common(const hfe_model* model, const float* x, int64_t n, int64_t ld, int32_t perm_col,
                 const int32_t* perm_idx, float* out, double* out64, uint8_t* labels, double* decision, double* meta,
                 const uint8_t* y, unsigned long long* score, uint32_t* status, void* stream, bool all_f64,
                 double* svc_dec = nullptr) {
    int rc = check_matrix(model, x, n, ld);
    if (rc) return rc;
    if (!status) return fail(HFE_ERR_INVALID_ARGUMENT, "null status array");
    if (perm_col >= F || (perm_col >= 0 && !perm_idx))
        return fail(HFE_ERR_INVALID_ARGUMENT, "bad permutation column / index array");
    if ((y != nullptr) != (score != nullptr)) return fail(HFE_ERR_INVALID_ARGUMENT, "y and score go together");
    if (reinterpret_cast<uintptr_t>(out) & 15)
        return fail(HFE_ERR_INVALID_ARGUMENT, "probability output must be 16-byte aligned");
    const Model* m = model->impl;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    HFE_CUDA_TRY(cudaSetDevice(m->device));
    HFE_CUDA_TRY(cudaMemsetAsync(status, 0, HFE_STATUS_HEADER_WORDS * sizeof(uint32_t), s));
    if (n == 0) return HFE_OK;
    ScoreParams P = m->base;
    P.x = x;
    P.n = n;
    P.ld = ld;
    for (int f = 0; f < F; ++f) P.col_off[f] = (long long)m->col_of[f] * ld;
    P.perm_col = perm_col < 0 ? -1 : m->internal_of[perm_col];
    P.perm_idx = perm_idx;
    P.out = out;
    P.out64 = out64;
    P.labels = labels;
    P.decision = decision;
    P.meta = meta;
    P.svc_dec = svc_dec;
    P.y = y;
    P.score = score;
    P.status = status;
    int sms = 0;
    HFE_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, m->device));
    if (all_f64) {
        long long blocks = std::min<long long>(n, (long long)sms * 64);
        hfe_score_f64_kernel<<<(unsigned)blocks, BLOCK_THREADS, 0, s>>>(P);
        HFE_CUDA_TRY(cudaGetLastError());
        return HFE_OK;
    }
    rc = dispatch_score(m, P, s);
    if (rc) return rc;
    P.svc_dec = nullptr;                // the probe wants the fast pass's decision values, not the re-evaluated ones
    hfe_fixup_kernel<<<(unsigned)sms, BLOCK_THREADS, 0, s>>>(P);       // re-evaluates the rows the fused kernel put on the work list
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

// ---- creation-time guard probe ---------------------------------------------------------------------------------------------
// The guard bands are derived from a formula (build_svc: ~20 sigma of the error measured while the kernels were written).
// This turns the formula into a checked contract: a fixed probe batch -- the model's own support vectors mapped back to
// raw units, and crossovers between pairs of them, so the rows sit where the RBF terms are largest and stay on the integer
// lattice the cohort lives on -- is scored by every fast engine the model can use and by the all-float64 kernel; the largest
// |fast - float64| of the SVC decision value and of the combiner's decision value over the rows the fast pass did NOT hand to
// the fix-up is kept in the handle (hfe_model_info), and a band narrower than 8 x what was observed is widened to that.
constexpr int PROBE_ROWS = 65536;

int probe_guards(hfe_model* h) {
    Model* m = h->impl;
    m->svc_observed_err = m->label_observed_err = 0.0;
    m->svc_observed_by_engine[0] = m->svc_observed_by_engine[1] = 0.0;
    m->probe_rows = 0;
    m->guard_widened = 0;
    m->svc_guard = m->svc_guard_formula * m->guard_scale;
    derive_guards(m);
    const ScoreParams& B = m->base;
    if (B.slot_svc < 0 || B.n_sv <= 0) return HFE_OK;
    HFE_CUDA_TRY(cudaSetDevice(m->device));
    const int64_t n = PROBE_ROWS, ld = PROBE_ROWS;
    std::vector<double> sv((size_t)B.n_sv * F);
    HFE_CUDA_TRY(cudaMemcpy(sv.data(), B.sv64, sv.size() * sizeof(double), cudaMemcpyDeviceToHost));
    std::vector<float> hx((size_t)F * ld);
    for (int64_t i = 0; i < n; ++i) {
        uint64_t z = (uint64_t)i * 0x9E3779B97F4A7C15ull + 0xD1B54A32D192ED03ull;        // splitmix64 of the row number
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        const int j = (int)(i % B.n_sv), j2 = (int)((z >> 1) % (uint64_t)B.n_sv);
        const unsigned cross = i < B.n_sv ? 0u : (unsigned)((z >> 20) & (z >> 40));         // each feature with probability 1/4
        for (int d = 0; d < F; ++d) {
            const double s = sv[(size_t)(((cross >> d) & 1u) ? j2 : j) * F + d];
            double v = B.mean[d] + B.scale[d] * s;
            const double r = std::nearbyint(v);
            if (std::fabs(v - r) <= 1e-6 * std::max(1.0, std::fabs(v))) v = r;
            hx[(size_t)m->col_of[d] * ld + i] = (float)v;
        }
    }
    const size_t x_bytes = hx.size() * sizeof(float), col_bytes = (size_t)n * sizeof(double);
    const size_t status_bytes = HFE_STATUS_WORDS * sizeof(uint32_t);
    unsigned char* arena = nullptr;
    HFE_CUDA_TRY(cudaMalloc(&arena, x_bytes + 4 * col_bytes + status_bytes));
    float* dx = reinterpret_cast<float*>(arena);
    double* d_dec[2] = {reinterpret_cast<double*>(arena + x_bytes), reinterpret_cast<double*>(arena + x_bytes + col_bytes)};
    double* d_svc[2] = {reinterpret_cast<double*>(arena + x_bytes + 2 * col_bytes),
                        reinterpret_cast<double*>(arena + x_bytes + 3 * col_bytes)};
    uint32_t* d_status = reinterpret_cast<uint32_t*>(arena + x_bytes + 4 * col_bytes);
    cudaStream_t s = cudaStreamPerThread;
    std::vector<double> dec64(n), svc64(n), dec(n), svc(n);
    std::vector<uint32_t> status(HFE_STATUS_WORDS);
    std::vector<unsigned char> flagged(n);
    const int engine_before = m->rbf_engine;
    int rc = HFE_OK;
    auto run = [&]() -> int {
        HFE_CUDA_TRY(cudaMemcpyAsync(dx, hx.data(), x_bytes, cudaMemcpyHostToDevice, s));
        int r = score_common(h, dx, n, ld, -1, nullptr, nullptr, nullptr, nullptr, d_dec[1], nullptr, nullptr, nullptr,
                             d_status, s, true, d_svc[1]);
        if (r) return r;
        HFE_CUDA_TRY(cudaMemcpyAsync(dec64.data(), d_dec[1], col_bytes, cudaMemcpyDeviceToHost, s));
        HFE_CUDA_TRY(cudaMemcpyAsync(svc64.data(), d_svc[1], col_bytes, cudaMemcpyDeviceToHost, s));
        double e_f = 0.0, e_z = 0.0;
        bool probed = false;
        for (int engine = 1; engine <= 2; ++engine) {
            if (engine == 2 && !(m->mma.nb > 0 && m->tree_mode != 2)) continue;
            m->rbf_engine = engine;
            r = score_common(h, dx, n, ld, -1, nullptr, nullptr, nullptr, nullptr, d_dec[0], nullptr, nullptr, nullptr,
                             d_status, s, false, d_svc[0]);
            if (r == HFE_ERR_UNSUPPORTED) continue;         // this engine cannot score the model (table too large): nothing to check
            if (r) return r;
            HFE_CUDA_TRY(cudaMemcpyAsync(dec.data(), d_dec[0], col_bytes, cudaMemcpyDeviceToHost, s));
            HFE_CUDA_TRY(cudaMemcpyAsync(svc.data(), d_svc[0], col_bytes, cudaMemcpyDeviceToHost, s));
            HFE_CUDA_TRY(cudaMemcpyAsync(status.data(), d_status, status_bytes, cudaMemcpyDeviceToHost, s));
            HFE_CUDA_TRY(cudaStreamSynchronize(s));
            probed = true;
            std::fill(flagged.begin(), flagged.end(), 0);
            const uint32_t len = std::min<uint32_t>(status[HFE_STATUS_LIST_LEN], HFE_WORKLIST_CAPACITY);
            for (uint32_t k = 0; k < len; ++k) flagged[status[HFE_STATUS_HEADER_WORDS + k]] = 1;
            for (int64_t i = 0; i < n; ++i) {
                if (flagged[i]) continue;
                const double df = std::fabs(svc[i] - svc64[i]), dz = std::fabs(dec[i] - dec64[i]);
                if (df == df) {
                    e_f = std::max(e_f, df);
                    m->svc_observed_by_engine[engine - 1] = std::max(m->svc_observed_by_engine[engine - 1], df);
                }
                if (dz == dz) e_z = std::max(e_z, dz);
            }
        }
        m->svc_observed_err = e_f;
        m->label_observed_err = e_z;
        m->probe_rows = probed ? (int)n : 0;
        return HFE_OK;
    };
    rc = run();
    m->rbf_engine = engine_before;
    cudaStreamSynchronize(s);
    cudaFree(arena);
    if (rc) return rc;
    if (m->svc_guard < 8.0 * m->svc_observed_err) {
        m->svc_guard = 8.0 * m->svc_observed_err;
        m->guard_widened |= 1;
        derive_guards(m);
    }
    if (m->base.label_guard < 8.0 * m->label_observed_err) {
        m->base.label_guard = 8.0 * m->label_observed_err;
        m->guard_widened |= 2;
    }
    return HFE_OK;
}

}  // namespace

namespace {
template <typename T>
int ingest_common(const T* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld, uint32_t* status,
                  void* stream) {
    if (n < 0 || n_features < 1 || n_features > 64 || ld < n)
        return fail(HFE_ERR_INVALID_ARGUMENT, "bad ingest sizes");
    if (!status) return fail(HFE_ERR_INVALID_ARGUMENT, "null status array");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    HFE_CUDA_TRY(cudaMemsetAsync(status, 0, HFE_STATUS_HEADER_WORDS * sizeof(uint32_t), s));
    if (n == 0) return HFE_OK;
    if (!x_rowmajor || !x) return fail(HFE_ERR_INVALID_ARGUMENT, "null ingest buffer");
    long long blocks = (n + BLOCK_THREADS - 1) / BLOCK_THREADS;
    static_assert((size_t)BLOCK_THREADS * (64 + 1) * sizeof(float) <= 48 * 1024,
                  "ingest tile of the widest accepted matrix must fit the default dynamic shared-memory limit");
    size_t smem = (size_t)BLOCK_THREADS * (n_features + 1) * sizeof(float);
    hfe_ingest_kernel<T><<<(unsigned)blocks, BLOCK_THREADS, smem, s>>>(x_rowmajor, n, n_features, x, ld, status);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}
}  // namespace

extern "C" {

int hfe_abi_version(void) { return 2; }

const char* hfe_last_error_string(void) { return g_last_error.c_str(); }

int hfe_model_create(const void* blob, size_t blob_bytes, int device, hfe_model** out_model) {
    try {
    if (!blob || !out_model) return fail(HFE_ERR_INVALID_ARGUMENT, "null blob or output pointer");
    *out_model = nullptr;
    if (device < 0) return fail(HFE_ERR_INVALID_ARGUMENT, "negative device ordinal");
    // aligned private copy: sections are read through double* / int32_t*
    std::vector<uint64_t> copy((blob_bytes + 7) / 8 + 1, 0);
    std::memcpy(copy.data(), blob, blob_bytes);
    Model* impl = nullptr;
    int rc = build_model(reinterpret_cast<const uint8_t*>(copy.data()), blob_bytes, device, &impl);
    if (rc) return rc;
    {   // staging arena for small host batches: all optional outputs of SMALL_HOST_ROWS rows + both status arrays
        auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
        size_t rows = SMALL_HOST_ROWS;
        size_t bytes = up(rows * F * sizeof(double)) + up(rows * F * sizeof(float)) + up(rows * 2 * sizeof(float)) +
                       up(rows * sizeof(double)) + up(rows * MAX_MEMBERS * sizeof(double)) + up(rows) +
                       up((HFE_STATUS_WORDS + HFE_STATUS_HEADER_WORDS) * sizeof(uint32_t)) + 256;
        void* p = nullptr;
        if (cudaMalloc(&p, bytes) == cudaSuccess) {
            impl->small_arena = static_cast<unsigned char*>(p);
            impl->small_arena_bytes = bytes;
            impl->allocations.push_back(p);
            impl->device_bytes += bytes;
        } else {
            cudaGetLastError();                 // no arena: small calls fall back to per-call allocation
        }
    }
    {   // mapped pinned block for the few-row path (inputs, every optional output, one status word)
        void* hp = nullptr;
        void* dp = nullptr;
        if (cudaHostAlloc(&hp, TINY_BYTES, cudaHostAllocMapped) == cudaSuccess &&
            cudaHostGetDevicePointer(&dp, hp, 0) == cudaSuccess) {
            impl->tiny_host = static_cast<unsigned char*>(hp);
            impl->tiny_dev = static_cast<unsigned char*>(dp);
        } else {
            if (hp) cudaFreeHost(hp);
            cudaGetLastError();                 // no mapped block: few-row calls take the staged path
        }
    }
    hfe_model* h = new (std::nothrow) hfe_model{impl};
    if (!h) {
        for (void* p : impl->allocations) cudaFree(p);
        if (impl->tiny_host) cudaFreeHost(impl->tiny_host);
        delete impl;
        return fail(HFE_ERR_INVALID_ARGUMENT, "out of host memory");
    }
    rc = probe_guards(h);                   // the guard bands as a checked contract (see probe_guards)
    if (rc) {
        hfe_model_destroy(h);
        return rc;
    }
    *out_model = h;
    return HFE_OK;
    } catch (const std::exception& e) {       // nothing may unwind through the C boundary
        return fail(HFE_ERR_INVALID_ARGUMENT, std::string("out of host memory or internal error: ") + e.what());
    }
}

int hfe_model_calibrate_guards(hfe_model* model, double formula_scale) {
    try {
    if (!model) return fail(HFE_ERR_INVALID_ARGUMENT, "null model");
    if (!(formula_scale > 0.0) || !(formula_scale <= 1e12)) return fail(HFE_ERR_INVALID_ARGUMENT, "formula_scale must be positive and finite");
    model->impl->guard_scale = formula_scale;
    return probe_guards(model);
    } catch (const std::exception& e) {       // nothing may unwind through the C boundary
        return fail(HFE_ERR_INVALID_ARGUMENT, std::string("out of host memory or internal error: ") + e.what());
    }
}

void hfe_model_destroy(hfe_model* model) {
    if (!model) return;
    if (model->impl) {
        cudaSetDevice(model->impl->device);
        for (void* p : model->impl->allocations) cudaFree(p);
        if (model->impl->tiny_host) cudaFreeHost(model->impl->tiny_host);
        delete model->impl->pipe;
        delete model->impl;
    }
    delete model;
}

static void fill_info(const Model* m, hfe_model_info* info);

int hfe_model_get_info(const hfe_model* model, hfe_model_info* info) {
    if (!model || !model->impl || !info) return fail(HFE_ERR_INVALID_ARGUMENT, "null model or info pointer");
    fill_info(model->impl, info);
    return HFE_OK;
}

#ifdef HFE_MMA_TRACE
int hfe_debug_trace(unsigned long long* out32, int reset) {
    if (out32) HFE_CUDA_TRY(cudaMemcpyFromSymbol(out32, g_mma_trace, sizeof(unsigned long long) * 32));
    if (reset) { unsigned long long z[32] = {}; HFE_CUDA_TRY(cudaMemcpyToSymbol(g_mma_trace, z, sizeof(z))); }
    return HFE_OK;
}
#endif

int hfe_model_set_rbf_engine(hfe_model* model, int engine) {
    if (!model || !model->impl) return fail(HFE_ERR_INVALID_ARGUMENT, "null model handle");
    if (engine < 0 || engine > 2) return fail(HFE_ERR_INVALID_ARGUMENT, "engine must be 0 (automatic), 1 (SIMT) or 2 (tensor cores)");
    model->impl->rbf_engine = engine;
    return HFE_OK;
}

int hfe_model_set_tree_engine(hfe_model* model, int engine) {
    if (!model || !model->impl) return fail(HFE_ERR_INVALID_ARGUMENT, "null model handle");
    if (engine < 0 || engine > 3) return fail(HFE_ERR_INVALID_ARGUMENT, "engine must be 0 (automatic), 1 (L1), 2 (shared memory) or 3 (L1 + L2 window)");
    model->impl->tree_engine = engine;
    return HFE_OK;
}

int hfe_model_plan(const void* blob, size_t blob_bytes, hfe_model_info* out_info, int32_t* feature_order) {
    try {
    if (!blob || !out_info) return fail(HFE_ERR_INVALID_ARGUMENT, "null blob or info pointer");
    std::vector<uint64_t> copy((blob_bytes + 7) / 8 + 1, 0);
    std::memcpy(copy.data(), blob, blob_bytes);
    Model* impl = nullptr;
    int rc = build_model(reinterpret_cast<const uint8_t*>(copy.data()), blob_bytes, -1, &impl);   // -1: no device work
    if (rc) return rc;
    fill_info(impl, out_info);
    if (feature_order)
        for (int f = 0; f < F; ++f) feature_order[f] = impl->col_of[f];
    delete impl;
    return HFE_OK;
    } catch (const std::exception& e) {       // nothing may unwind through the C boundary
        return fail(HFE_ERR_INVALID_ARGUMENT, std::string("out of host memory or internal error: ") + e.what());
    }
}

static void fill_info(const Model* m, hfe_model_info* info) {
    std::memset(info, 0, sizeof(*info));
    info->n_features = F;
    info->n_members = m->n_members;
    for (int i = 0; i < 4; ++i) info->member_kind[i] = m->kinds[i];
    info->n_support_vectors = m->n_sv;
    info->n_trees = m->n_trees;
    info->n_tree_nodes = m->n_nodes;
    info->max_tree_depth = m->max_depth;
    info->tree_mode = m->tree_mode;
    info->n_merged_splits = m->n_splits;
    info->device = m->device;
    info->rbf_fma_per_row = m->n_sv == 0 ? 0
                            : m->base.sv_quads > 0 ? 4 * (m->base.sv_quads * SV_PREFIX + m->sv_steps * SV_LEAF_STEPS)
                                                   : 4 * m->base.n_groups * SV_STEPS;
    info->svc_guard = m->svc_guard;
    info->svc_observed_err = m->svc_observed_err;
    info->svc_observed_by_engine[0] = m->svc_observed_by_engine[0];
    info->svc_observed_by_engine[1] = m->svc_observed_by_engine[1];
    info->label_guard = m->base.label_guard;
    info->label_observed_err = m->label_observed_err;
    info->probe_rows = m->probe_rows;
    info->guard_widened = m->guard_widened;
    info->device_bytes = m->device_bytes;
    info->rbf_engine = m->n_sv == 0 ? 0 : (m->mma.nb > 0 && m->tree_mode != 2 && m->rbf_engine != 1) ? 2 : 1;
    info->rbf_exp_per_row = info->rbf_engine == 2 ? m->mma.nb * m->mma.block_sv : ((m->n_sv + 3) / 4) * 4;
}

int hfe_predict_proba(const hfe_model* model, const float* x, int64_t n, int64_t ld, float* out_n2, uint32_t* status,
                      void* stream) {
    if (n > 0 && !out_n2) return fail(HFE_ERR_INVALID_ARGUMENT, "null output");
    return score_common(model, x, n, ld, -1, nullptr, out_n2, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                        status, stream, false);
}

int hfe_predict(const hfe_model* model, const float* x, int64_t n, int64_t ld, uint8_t* labels, float* out_n2,
                uint32_t* status, void* stream) {
    if (n > 0 && !labels) return fail(HFE_ERR_INVALID_ARGUMENT, "null labels");
    return score_common(model, x, n, ld, -1, nullptr, out_n2, nullptr, labels, nullptr, nullptr, nullptr, nullptr,
                        status, stream, false);
}

int hfe_score(const hfe_model* model, const float* x, int64_t n, int64_t ld, int32_t perm_col,
              const int32_t* perm_idx, float* out_n2, uint8_t* labels, double* decision, double* meta,
              const uint8_t* y, unsigned long long* score, uint32_t* status, void* stream) {
    return score_common(model, x, n, ld, perm_col, perm_idx, out_n2, nullptr, labels, decision, meta, y, score, status,
                        stream, false);
}

int hfe_predict_proba_f64(const hfe_model* model, const float* x, int64_t n, int64_t ld, double* out_n2,
                          uint32_t* status, void* stream) {
    if (n > 0 && !out_n2) return fail(HFE_ERR_INVALID_ARGUMENT, "null output");
    return score_common(model, x, n, ld, -1, nullptr, nullptr, out_n2, nullptr, nullptr, nullptr, nullptr, nullptr,
                        status, stream, true);
}

int hfe_score_f64(const hfe_model* model, const float* x, int64_t n, int64_t ld, double* out_n2, uint8_t* labels,
                  double* decision, double* meta, uint32_t* status, void* stream) {
    return score_common(model, x, n, ld, -1, nullptr, nullptr, out_n2, labels, decision, meta, nullptr, nullptr, status,
                        stream, true);
}

int hfe_score_permuted(const hfe_model* model, const float* x, int64_t n, int64_t ld, int32_t perm_col,
                       const int32_t* perm_idx, int64_t perm_stride, int32_t n_sel, const int32_t* sel,
                       const uint8_t* y, unsigned long long* scores, uint32_t* flags, uint32_t* status, void* stream) {
    if (n_sel < 0 || (n_sel > 0 && (!sel || !scores || !y))) return fail(HFE_ERR_INVALID_ARGUMENT, "bad permuted-pass selection");
    if (perm_col >= 0 && perm_stride < n) return fail(HFE_ERR_INVALID_ARGUMENT, "index vectors overlap");
    for (int32_t k = 0; k < n_sel; ++k) {
        if (sel[k] < 0) return fail(HFE_ERR_INVALID_ARGUMENT, "negative index-vector number");
        int rc = score_common(model, x, n, ld, perm_col, perm_col >= 0 ? perm_idx + (int64_t)sel[k] * perm_stride : nullptr,
                              nullptr, nullptr, nullptr, nullptr, nullptr, y, scores + k, status, stream, false);
        if (rc) return rc;
        if (flags)      // the status array is reused by the next pass: keep this pass's error bits
            HFE_CUDA_TRY(cudaMemcpyAsync(flags + k, status, sizeof(uint32_t), cudaMemcpyDeviceToDevice,
                                         static_cast<cudaStream_t>(stream)));
    }
    return HFE_OK;
}

namespace {
// numpy's legacy generator (RandomState): MT19937 seeded by init_genrand, 32-bit draws
struct Mt19937 {
    uint32_t mt[624];
    int pos = 624;
    explicit Mt19937(uint32_t seed) {
        mt[0] = seed;
        for (int i = 1; i < 624; ++i) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
    }
    static uint32_t twist(uint32_t hi, uint32_t lo, uint32_t far) {
        const uint32_t yv = (hi & 0x80000000u) | (lo & 0x7FFFFFFFu);
        return far ^ (yv >> 1) ^ ((yv & 1u) ? 0x9908B0DFu : 0u);
    }
    void refill() {                      // genrand_int32's block update, split so that no index wraps inside a loop
        for (int k = 0; k < 227; ++k) mt[k] = twist(mt[k], mt[k + 1], mt[k + 397]);
        for (int k = 227; k < 623; ++k) mt[k] = twist(mt[k], mt[k + 1], mt[k - 227]);
        mt[623] = twist(mt[623], mt[0], mt[396]);
        pos = 0;
    }
    void next_block(uint32_t* out) {            // the next 624 outputs at once (pos must be at a block boundary)
        refill();
        for (int k = 0; k < 624; ++k) {
            uint32_t v = mt[k];
            v ^= v >> 11;
            v ^= (v << 7) & 0x9D2C5680u;
            v ^= (v << 15) & 0xEFC60000u;
            v ^= v >> 18;
            out[k] = v;
        }
        pos = 624;
    }
};
}  // namespace

namespace {
// scikit-learn's index vectors for permutation importance (SK/inspection/_permutation_importance.py:66-74): RandomState(seed)
// shuffles ONE index vector n_repeats times and applies it each time to the already permuted column, so the effective
// source index composes: e_r = e_{r-1}[idx_r].  Three threads: one makes the draws (masked rejection on the MT19937 stream:
// they do not depend on the array), the caller's applies the swaps -- a serial chain, fed a block ahead with its targets
// prefetched -- and one composes a finished shuffle with the previous vector while the next shuffle proceeds.  `ready`
// counts the vectors that are complete in `out`, in order.
struct ShuffleJob {
    uint32_t seed = 0;
    int64_t n = 0;
    int32_t reps = 0;
    int32_t* out = nullptr;
    std::atomic<int> ready{0};
    std::atomic<int> failed{0};

    static constexpr int BLK = 4096, NBLK = 8;

    void run() {
        try {
            run_impl();
        } catch (const std::exception&) {
            failed.store(1);
        }
        if (failed.load()) ready.store(reps);          // never leave a waiter spinning
    }

private:
    void run_impl() {
        if (n <= 0 || reps <= 0) { ready.store(reps); return; }
        std::vector<uint32_t> ring((size_t)BLK * NBLK);
        std::atomic<uint64_t> produced{0}, consumed{0};
        const uint64_t per_rep = (uint64_t)(n - 1), total = per_rep * (uint64_t)reps;
        std::thread drawer([&] {
            // random_interval(top): 32-bit draws masked to the bits of top, rejected while above it.  The generator's outputs
            // are made 624 at a time, and the rejection is branch-free: a draw is written to the next slot and the slot only
            // advances if it was accepted (the mask changes only when top passes a power of two, so runs share it).
            Mt19937 rng(seed);
            uint32_t words[624];
            int wpos = 624;
            uint64_t made = 0;
            uint32_t top = (uint32_t)per_rep;                         // i runs n-1 .. 1 in every repeat
            while (made < total) {
                const uint64_t b = produced.load(std::memory_order_relaxed);
                while (b - consumed.load(std::memory_order_acquire) >= (uint64_t)NBLK) std::this_thread::yield();
                uint32_t* dst = &ring[(size_t)(b % NBLK) * BLK];
                const uint32_t cnt = (uint32_t)std::min<uint64_t>(BLK, total - made);
                uint32_t filled = 0;
                while (filled < cnt) {
                    const uint32_t mask = 0xFFFFFFFFu >> __builtin_clz(top);
                    const uint32_t floor_top = (mask >> 1) + 1u;       // smallest top with this mask
                    uint32_t need = std::min<uint32_t>(cnt - filled, top - floor_top + 1u);
                    while (need > 0) {
                        if (wpos == 624) { rng.next_block(words); wpos = 0; }
                        const uint32_t avail = std::min<uint32_t>((uint32_t)(624 - wpos), need * 2u + 8u);
                        for (uint32_t k = 0; k < avail && need > 0; ++k) {
                            const uint32_t jv = words[wpos++] & mask;
                            dst[filled] = jv;
                            const uint32_t ok = jv <= top ? 1u : 0u;
                            filled += ok;
                            top -= ok;
                            need -= ok;
                        }
                    }
                    if (top == 0) top = (uint32_t)per_rep;             // next repeat
                }
                made += cnt;
                produced.store(b + 1, std::memory_order_release);
            }
        });
        std::vector<int32_t> idx((size_t)n), snap[2];
        for (int64_t i = 0; i < n; ++i) idx[(size_t)i] = (int32_t)i;
        std::thread composer;
        uint64_t done = 0, blk = 0;
        auto wait_block = [&](uint64_t b) { while (produced.load(std::memory_order_acquire) <= b) std::this_thread::yield(); };
        for (int32_t r = 0; r < reps; ++r) {
            uint64_t left = per_rep;
            int64_t i = n - 1;
            while (left > 0) {
                wait_block(blk);
                const uint32_t* js = &ring[(size_t)(blk % NBLK) * BLK];
                const uint64_t in_blk = std::min<uint64_t>(BLK, total - blk * BLK);
                const uint64_t at = done - blk * BLK;                       // position inside the block (a repeat may end mid-block)
                const uint64_t cnt = std::min<uint64_t>(in_blk - at, left);
                // the swap chain is serial and its targets are random: each target is prefetched PD steps ahead (a burst for
                // a whole block would overflow the core's handful of fill buffers)
                constexpr uint64_t PD = 24;
                if (at == 0)
                    for (uint64_t k = 0; k < std::min<uint64_t>(PD, in_blk); ++k) __builtin_prefetch(&idx[js[k]], 1);
                for (uint64_t k = 0; k < cnt; ++k, --i) {
                    if (at + k + PD < in_blk) __builtin_prefetch(&idx[js[at + k + PD]], 1);
                    std::swap(idx[(size_t)i], idx[js[at + k]]);
                }
                done += cnt;
                left -= cnt;
                if (at + cnt == in_blk) {
                    ++blk;
                    consumed.store(blk, std::memory_order_release);
                }
            }
            int32_t* cur = out + (size_t)r * (size_t)n;
            if (composer.joinable()) composer.join();
            if (r == 0) {
                std::memcpy(cur, idx.data(), (size_t)n * sizeof(int32_t));
                ready.store(1, std::memory_order_release);
            } else {
                std::vector<int32_t>& sn = snap[r & 1];
                sn = idx;
                const int32_t* prev = cur - n;
                const int32_t* src = sn.data();
                const int64_t nn = n;
                std::atomic<int>* flag = &ready;
                composer = std::thread([cur, prev, src, nn, flag, r] {
                    for (int64_t k = 0; k < nn; ++k) cur[k] = prev[src[k]];
                    flag->store(r + 1, std::memory_order_release);
                });
            }
        }
        if (composer.joinable()) composer.join();
        drawer.join();
    }
};
}  // namespace

int hfe_perm_indices(uint32_t seed, int64_t n, int32_t n_repeats, int32_t* out) {
    if (n < 0 || n > 0x7FFFFFFF || n_repeats < 0 || (!out && n > 0 && n_repeats > 0))
        return fail(HFE_ERR_INVALID_ARGUMENT, "bad permutation request");
    ShuffleJob job;
    job.seed = seed;
    job.n = n;
    job.reps = n_repeats;
    job.out = out;
    job.run();
    if (job.failed.load()) return fail(HFE_ERR_INVALID_ARGUMENT, "permutation vectors: out of host memory");
    return HFE_OK;
}

int hfe_permutation_passes(const hfe_model* model, const float* x, int64_t n, int64_t ld, const uint8_t* y, uint32_t seed,
                           int32_t n_repeats, int32_t n_jobs, const int32_t* job_col, const int32_t* job_rep,
                           int32_t* perm_host, int32_t* perm_dev, unsigned long long* scores, uint32_t* flags,
                           uint32_t* status, void* stream) {
    if (n_jobs < 0 || n_repeats < 0 || (n_jobs > 0 && (!job_col || !job_rep || !scores || !y)))
        return fail(HFE_ERR_INVALID_ARGUMENT, "bad permutation job list");
    if (n < 0 || n > 0x7FFFFFFF) return fail(HFE_ERR_INVALID_ARGUMENT, "bad row count");
    int last = -1;
    for (int32_t k = 0; k < n_jobs; ++k) {
        if (job_col[k] >= F) return fail(HFE_ERR_INVALID_ARGUMENT, "bad permutation column");
        if (job_col[k] < 0) continue;
        if (job_rep[k] < 0 || job_rep[k] >= n_repeats || job_rep[k] < last)
            return fail(HFE_ERR_INVALID_ARGUMENT, "permutation jobs must come in ascending repeat order");
        last = job_rep[k];
    }
    if (last >= 0 && (!perm_host || !perm_dev)) return fail(HFE_ERR_INVALID_ARGUMENT, "null index-vector buffers");
    try {
        ShuffleJob job;
        job.seed = seed;
        job.n = n;
        job.reps = last + 1;                       // vectors beyond the last one asked for are not made
        job.out = perm_host;
        std::thread maker([&job] { job.run(); });
        struct Join { std::thread& t; ~Join() { if (t.joinable()) t.join(); } } join{maker};
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        int uploaded = 0;
        for (int32_t k = 0; k < n_jobs; ++k) {
            const int col = job_col[k], r = job_rep[k];
            if (col >= 0) {
                while (job.ready.load(std::memory_order_acquire) <= r) std::this_thread::yield();
                if (job.failed.load()) return fail(HFE_ERR_INVALID_ARGUMENT, "permutation vectors: out of host memory");
                for (; uploaded <= r; ++uploaded)
                    HFE_CUDA_TRY(cudaMemcpyAsync(perm_dev + (size_t)uploaded * (size_t)n, perm_host + (size_t)uploaded * (size_t)n,
                                                 (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, s));
            }
            int rc = score_common(model, x, n, ld, col, col >= 0 ? perm_dev + (size_t)r * (size_t)n : nullptr, nullptr, nullptr,
                                  nullptr, nullptr, nullptr, y, scores + k, status, stream, false);
            if (rc) return rc;
            if (flags)
                HFE_CUDA_TRY(cudaMemcpyAsync(flags + k, status, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
        }
    } catch (const std::exception& e) {
        return fail(HFE_ERR_INVALID_ARGUMENT, std::string("permutation passes: ") + e.what());
    }
    return HFE_OK;
}

int hfe_apply(const hfe_model* model, const float* x, int64_t n, int64_t ld, uint16_t* leaves, void* stream) {
    int rc = check_matrix(model, x, n, ld);
    if (rc) return rc;
    const Model* m = model->impl;
    if (m->n_trees == 0) return fail(HFE_ERR_UNSUPPORTED, "model has no boosted-tree member");
    if (n == 0) return HFE_OK;
    if (!leaves) return fail(HFE_ERR_INVALID_ARGUMENT, "null output");
    HFE_CUDA_TRY(cudaSetDevice(m->device));
    ScoreParams P = m->base;
    P.x = x;
    P.n = n;
    P.ld = ld;
    for (int f = 0; f < F; ++f) P.col_off[f] = (long long)m->col_of[f] * ld;
    long long blocks = (n + BLOCK_THREADS - 1) / BLOCK_THREADS;
    hfe_apply_kernel<<<(unsigned)blocks, BLOCK_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(P, leaves);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

int hfe_synth_rows(uint64_t seed, int64_t row0, int64_t n, int32_t n_cols, const int32_t* col_id, const float* base,
                   const float* step, const int32_t* thr_off, const uint32_t* thr, float* x, int64_t ld,
                   void* stream) {
    if (n < 0 || row0 < 0 || n_cols < 1 || ld < n) return fail(HFE_ERR_INVALID_ARGUMENT, "bad synth_rows sizes");
    if (n == 0) return HFE_OK;
    if (!col_id || !base || !step || !thr_off || !thr || !x)
        return fail(HFE_ERR_INVALID_ARGUMENT, "null synth_rows table");
    long long blocks = std::min<long long>((n + BLOCK_THREADS - 1) / BLOCK_THREADS, (long long)sm_count_current() * 16);
    hfe_synth_kernel<<<(unsigned)blocks, BLOCK_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
        seed, row0, n, n_cols, col_id, base, step, thr_off, thr, x, ld);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

int hfe_ingest_rowmajor_f64(const double* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                            uint32_t* status, void* stream) {
    return ingest_common(x_rowmajor, n, n_features, x, ld, status, stream);
}

int hfe_ingest_rowmajor_f32(const float* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                            uint32_t* status, void* stream) {
    return ingest_common(x_rowmajor, n, n_features, x, ld, status, stream);
}

int hfe_ingest_rowmajor_i64(const int64_t* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                            uint32_t* status, void* stream) {
    return ingest_common(reinterpret_cast<const long long*>(x_rowmajor), n, n_features, x, ld, status, stream);
}

int hfe_ingest_rowmajor_u8(const uint8_t* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                           uint32_t* status, void* stream) {
    return ingest_common(x_rowmajor, n, n_features, x, ld, status, stream);
}

int hfe_knn1_impute(const double* fit_x, int64_t n_fit, int32_t n_features, const double* col_mean, double* x,
                    int64_t n, void* stream) {
    if (n < 0 || n_fit < 1 || n_features < 1 || n_features > 4096)
        return fail(HFE_ERR_INVALID_ARGUMENT, "bad imputer sizes");
    if (n == 0) return HFE_OK;
    if (!fit_x || !col_mean || !x) return fail(HFE_ERR_INVALID_ARGUMENT, "null imputer buffer");
    size_t smem = ((size_t)n_features + (size_t)n_fit) * sizeof(double);
    if (smem > 200 * 1024) return fail(HFE_ERR_TOO_LARGE, "imputer: more fit rows than one block's shared memory holds");
    if (smem > 48 * 1024)         // per-device attribute; this is training-time preprocessing, so simply set it per call
        HFE_CUDA_TRY(cudaFuncSetAttribute(hfe_knn1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    long long blocks = std::min<long long>(n, (long long)sm_count_current() * 8);
    hfe_knn1_kernel<<<(unsigned)blocks, KNN_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(fit_x, n_fit, n_features,
                                                                                               col_mean, x, n);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

int hfe_binary_metrics(const float* proba_n2, const uint8_t* y, int64_t n, float threshold,
                       unsigned long long* confusion, uint32_t* hist, void* stream) {
    if (n < 0 || !confusion) return fail(HFE_ERR_INVALID_ARGUMENT, "bad metrics arguments");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    HFE_CUDA_TRY(cudaMemsetAsync(confusion, 0, 4 * sizeof(unsigned long long), s));
    if (hist) HFE_CUDA_TRY(cudaMemsetAsync(hist, 0, 2 * METRIC_BINS * sizeof(uint32_t), s));
    if (n == 0) return HFE_OK;
    if (!proba_n2 || !y) return fail(HFE_ERR_INVALID_ARGUMENT, "null metrics input");
    long long blocks = std::min<long long>((n + 255) / 256, (long long)sm_count_current() * 8);
    hfe_metrics_kernel<<<(unsigned)blocks, 256, 0, s>>>(proba_n2, y, n, threshold, confusion, hist);
    HFE_CUDA_TRY(cudaGetLastError());
    return HFE_OK;
}

// ---- allocation-free variant ------------------------------------------------------------------------------------------
// Scratch for hfe_predict_proba_noalloc: the status header + work list that the fused kernel and the fix-up kernel
// share, kept in the library image (256 KiB of device memory per GPU that loads this library) so that the call never
// has to ask anyone for memory.  In host code this name is only a handle for cudaGetSymbolAddress -- it is not a
// device address and must not be passed to a kernel as one.
static __device__ __align__(256) unsigned int g_noalloc_scratch[HFE_STATUS_WORDS];

namespace {
// Device address of g_noalloc_scratch on the current device.  Asked for on every call: after its first use per device
// the lookup is a table read inside the runtime (no allocation, no synchronisation), and an address cached here would
// go stale if the embedding process reset the device or re-created its primary context.
int noalloc_scratch(unsigned int** out) {
    void* p = nullptr;
    HFE_CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    *out = static_cast<unsigned int*>(p);
    return HFE_OK;
}
}  // namespace

int hfe_predict_proba_noalloc(const hfe_model* model, const float* x, int64_t n, int64_t ld, float* out_n2,
                              uint32_t* flags, void* stream) {
    if (!model || !model->impl) return fail(HFE_ERR_INVALID_ARGUMENT, "null model handle");
    if (n > (int64_t)HFE_NOALLOC_MAX_ROWS)
        return fail(HFE_ERR_TOO_LARGE, "hfe_predict_proba_noalloc scores at most HFE_NOALLOC_MAX_ROWS (2^26) rows per "
                                       "call: split the batch");
    if (n > 0 && !out_n2) return fail(HFE_ERR_INVALID_ARGUMENT, "null output");
    HFE_CUDA_TRY(cudaSetDevice(model->impl->device));
    unsigned int* scratch = nullptr;
    int rc = noalloc_scratch(&scratch);
    if (rc) return rc;
    // score_common clears the header words (error bits, work-list length) on `stream` before the kernels run, so
    // whatever the previous call left in the block cannot leak into this one
    rc = score_common(model, x, n, ld, -1, nullptr, out_n2, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                      scratch, stream, false);
    if (rc) return rc;
    if (flags)
        HFE_CUDA_TRY(cudaMemcpyAsync(flags, scratch, sizeof(uint32_t), cudaMemcpyDeviceToDevice,
                                     static_cast<cudaStream_t>(stream)));
    return HFE_OK;
}

}  // extern "C" (reopened below)

namespace {
// Large host batches: see hfe_host.cuh.  `all_f64` reruns the batch on the float64 kernel (work-list overflow).
//
// One parallel region per chunk k, every thread of the pool in it:
//   * thread 0 first hands chunk k-1 (narrowed in the previous region) to its slot's stream -- copy in, ingest, score, copies
//     out, event -- so no thread waits while the CUDA calls are made;
//   * everybody then takes work units off one counter: blocks of chunk k to narrow into slot k % 3, and blocks of the
//     finished chunk k-3 (the slot's previous tenant, its event already waited for) to widen into the caller's arrays.
// Units are small (32 K values / 16 K rows), so a thread that starts late or runs on a busy CPU just takes fewer.
int predict_host_pipelined(const hfe_model* model, const double* x_rowmajor, int64_t n, double* proba_n2,
                           uint8_t* labels, double* decision, double* meta, bool all_f64) {
    const Model* m = model->impl;
    std::lock_guard<std::mutex> lock(m->pipe_mutex);
    if (!m->pipe) {
        int rc = host_pipeline_create(m->device, m->n_members, &m->pipe);
        if (rc) return rc;
    }
    HostPipeline& hp = *m->pipe;
    const int64_t C = hp.chunk;
    const int nm = m->n_members;
    const int64_t n_chunks = (n + C - 1) / C;
    uint32_t flags = 0;
    struct Drain {                       // whatever happens, no stream still touches the slots when we leave
        HostPipeline& hp;
        ~Drain() { for (auto& s : hp.slot) { if (s.busy) cudaStreamSynchronize(s.stream); s.busy = false; } }
    } drain{hp};
#ifdef HFE_DEV
    struct Report {
        HostPool* pool; double t0, busy0, wait = 0;
        ~Report() { std::fprintf(stderr, "[hfe host pipeline] total %.2f ms, threads busy %.2f ms, waiting for the device %.2f ms\n",
                                 (HostPool::now() - t0) * 1e3, (pool->busy_seconds - busy0) * 1e3, wait * 1e3); }
    } report{hp.pool, HostPool::now(), hp.pool->busy_seconds};
#endif

    // thread 0, inside a region: chunk `k` (narrowed, in its slot) goes to the device
    auto enqueue = [&](int64_t k) -> int {
        HostSlot& s = hp.slot[k % HOST_SLOTS];
        const int64_t rows = s.rows;
        float* d_raw = reinterpret_cast<float*>(s.d_buf + hp.off_raw);
        float* d_col = reinterpret_cast<float*>(s.d_buf + hp.off_col);
        float* d_out = reinterpret_cast<float*>(s.d_buf + hp.off_out);
        uint8_t* d_lab = labels ? s.d_buf + hp.off_lab : nullptr;
        double* d_dec = decision ? reinterpret_cast<double*>(s.d_buf + hp.off_dec) : nullptr;
        double* d_meta = meta ? reinterpret_cast<double*>(s.d_buf + hp.off_meta) : nullptr;
        uint32_t* d_status = reinterpret_cast<uint32_t*>(s.d_buf + hp.off_status);     // [ingest header | scoring status]
        int rc;
        if (s.as_bytes) {
            HFE_CUDA_TRY(cudaMemcpyAsync(d_raw, s.h_in, (size_t)rows * F, cudaMemcpyHostToDevice, s.stream));
            rc = hfe_ingest_rowmajor_u8(reinterpret_cast<const uint8_t*>(d_raw), rows, F, d_col, C, d_status, s.stream);
        } else {
            HFE_CUDA_TRY(cudaMemcpyAsync(d_raw, s.h_in, (size_t)rows * F * sizeof(float), cudaMemcpyHostToDevice, s.stream));
            rc = hfe_ingest_rowmajor_f32(d_raw, rows, F, d_col, C, d_status, s.stream);
        }
        if (rc) return rc;
        rc = score_common(model, d_col, rows, C, -1, nullptr, d_out, nullptr, d_lab, d_dec, d_meta, nullptr, nullptr,
                          d_status + HFE_STATUS_HEADER_WORDS, s.stream, all_f64);
        if (rc) return rc;
        HFE_CUDA_TRY(cudaMemcpyAsync(s.h_out, d_out, (size_t)rows * 2 * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        if (labels) HFE_CUDA_TRY(cudaMemcpyAsync(s.h_out + hp.h_off_lab, d_lab, (size_t)rows, cudaMemcpyDeviceToHost, s.stream));
        if (decision)
            HFE_CUDA_TRY(cudaMemcpyAsync(s.h_out + hp.h_off_dec, d_dec, (size_t)rows * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
        if (meta)
            HFE_CUDA_TRY(cudaMemcpyAsync(s.h_out + hp.h_off_meta, d_meta, (size_t)rows * nm * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
        HFE_CUDA_TRY(cudaMemcpyAsync(s.h_out + hp.h_off_status, d_status, 2 * HFE_STATUS_HEADER_WORDS * sizeof(uint32_t),
                                     cudaMemcpyDeviceToHost, s.stream));
        HFE_CUDA_TRY(cudaEventRecord(s.done, s.stream));
        s.busy = true;
        return HFE_OK;
    };

#ifndef HFE_UNIT_VALUES
#define HFE_UNIT_VALUES (1 << 15)
#endif
    constexpr int64_t UNIT_VALUES = HFE_UNIT_VALUES, UNIT_ROWS = 1 << 14;
    // Wire format, chosen per chunk: while every value so far in this call was an integer in [0, 255] (cohort variables are
    // counts, classes and yes/no answers) a chunk crosses PCIe as bytes -- 17 instead of 68 per row, lossless, so the scores
    // are the same bits; the first chunk that holds anything else is narrowed again to float32 and the rest of the call does
    // not try bytes any more.
    bool bytes_ok = true;
    for (int64_t k = 0; k < n_chunks + HOST_SLOTS; ++k) {
        const bool has_narrow = k < n_chunks;                                     // chunk k -> slot k % 3
        const bool has_enqueue = k >= 1 && k - 1 < n_chunks;                      // chunk k-1, narrowed last time round
        const int64_t w = k - HOST_SLOTS;                                         // chunk to widen (shares the slot with chunk k)
        const bool has_widen = w >= 0 && w < n_chunks;
        HostSlot& s = hp.slot[k % HOST_SLOTS];
        int64_t w_begin = 0, w_rows = 0;
        if (has_widen) {                                                          // its results must have arrived
#ifdef HFE_DEV
            const double w0 = HostPool::now();
#endif
            HFE_CUDA_TRY(cudaEventSynchronize(s.done));
#ifdef HFE_DEV
            report.wait += HostPool::now() - w0;
#endif
            s.busy = false;
            const uint32_t* st = reinterpret_cast<const uint32_t*>(s.h_out + hp.h_off_status);
            flags |= st[0] | st[HFE_STATUS_HEADER_WORDS];
            w_begin = s.begin;
            w_rows = s.rows;
        }
        const int64_t begin = k * C, rows = has_narrow ? std::min(C, n - begin) : 0;
        const double* src = x_rowmajor + begin * F;
        const int64_t total = rows * F;
        const int64_t narrow_units = (total + UNIT_VALUES - 1) / UNIT_VALUES, widen_units = (w_rows + UNIT_ROWS - 1) / UNIT_ROWS;
        const bool try_bytes = bytes_ok;
        std::atomic<int64_t> next{0};
        std::atomic<int> not_bytes{0}, enqueue_rc{HFE_OK};
        unsigned char* h_u8 = reinterpret_cast<unsigned char*>(s.h_in);
        const unsigned char* w_out = s.h_out;        // (the slot's output block is only rewritten by chunk k, enqueued next round)
        hp.pool->run([&](int t, int) {
            if (t == 0 && has_enqueue) enqueue_rc.store(enqueue(k - 1));
            for (;;) {
                const int64_t u = next.fetch_add(1, std::memory_order_relaxed);
                if (u < narrow_units) {
                    const int64_t e0 = u * UNIT_VALUES, e1 = std::min(total, e0 + UNIT_VALUES);
                    if (try_bytes) {
                        if (!not_bytes.load(std::memory_order_relaxed) && !narrow_rows_u8(src + e0, h_u8 + e0, (size_t)(e1 - e0)))
                            not_bytes.store(1, std::memory_order_relaxed);
                    } else {
                        narrow_rows(src + e0, s.h_in + e0, (size_t)(e1 - e0));
                    }
                } else if (u < narrow_units + widen_units) {
                    const int64_t r0 = (u - narrow_units) * UNIT_ROWS, r1 = std::min(w_rows, r0 + UNIT_ROWS);
                    if (proba_n2)
                        widen_proba(reinterpret_cast<const float*>(w_out) + 2 * r0, proba_n2 + 2 * (w_begin + r0), (size_t)(r1 - r0));
                    if (labels) std::memcpy(labels + w_begin + r0, w_out + hp.h_off_lab + r0, (size_t)(r1 - r0));
                    if (decision)
                        std::memcpy(decision + w_begin + r0, w_out + hp.h_off_dec + r0 * sizeof(double), (size_t)(r1 - r0) * sizeof(double));
                    if (meta)
                        std::memcpy(meta + (w_begin + r0) * nm, w_out + hp.h_off_meta + r0 * nm * sizeof(double),
                                    (size_t)(r1 - r0) * nm * sizeof(double));
                } else {
                    break;
                }
            }
        });
        if (enqueue_rc.load()) return enqueue_rc.load();
        if (has_narrow) {
            if (try_bytes && not_bytes.load()) {                                  // something that is not a small integer: float32 from here on
                bytes_ok = false;
                std::atomic<int64_t> again{0};
                hp.pool->run([&](int, int) {
                    for (;;) {
                        const int64_t u = again.fetch_add(1, std::memory_order_relaxed);
                        if (u >= narrow_units) break;
                        const int64_t e0 = u * UNIT_VALUES, e1 = std::min(total, e0 + UNIT_VALUES);
                        narrow_rows(src + e0, s.h_in + e0, (size_t)(e1 - e0));
                    }
                });
            }
            s.as_bytes = bytes_ok;
            s.begin = begin;
            s.rows = rows;
        }
    }
    if (flags & HFE_ROW_NONFINITE) return fail(HFE_ERR_NONFINITE_INPUT, "Input X contains NaN or infinity");
    if ((flags & HFE_WORKLIST_OVERFLOW) && !all_f64) return -1;      // caller reruns on the float64 kernel
    return HFE_OK;
}
}  // namespace

extern "C" {

int hfe_predict_host(const hfe_model* model, const double* x_rowmajor, int64_t n, double* proba_n2, uint8_t* labels,
                     double* decision, double* meta) {
    try {
    if (!model || !model->impl) return fail(HFE_ERR_INVALID_ARGUMENT, "null model handle");
    if (n < 0) return fail(HFE_ERR_INVALID_ARGUMENT, "negative row count");
    if (n == 0) return HFE_OK;
    if (!x_rowmajor) return fail(HFE_ERR_INVALID_ARGUMENT, "null input");
    const Model* m = model->impl;
    HFE_CUDA_TRY(cudaSetDevice(m->device));
    if (n <= TINY_HOST_ROWS && m->tiny_host) {
        // zero-copy: the rows go into mapped pinned memory, one kernel (a warp per row, float64 path) reads them over
        // PCIe and writes the results back; what is left is one launch and one synchronisation
        std::lock_guard<std::mutex> lock(m->small_mutex);
        unsigned char* hb = m->tiny_host;
        unsigned char* db = m->tiny_dev;
        std::memcpy(hb + TINY_OFF_X, x_rowmajor, (size_t)n * F * sizeof(double));
        *reinterpret_cast<uint32_t*>(hb + TINY_OFF_STATUS) = 0;
        ScoreParams P = m->base;
        P.n = n;
        P.ld = 1;
        for (int f = 0; f < F; ++f) P.col_off[f] = m->col_of[f];
        P.out = nullptr;
        P.out64 = reinterpret_cast<double*>(db + TINY_OFF_OUT);      // the kernel evaluates in float64: hand that back
        P.labels = labels ? db + TINY_OFF_LABELS : nullptr;
        P.decision = decision ? reinterpret_cast<double*>(db + TINY_OFF_DECISION) : nullptr;
        P.meta = meta ? reinterpret_cast<double*>(db + TINY_OFF_META) : nullptr;
        hfe_tiny_kernel<<<(unsigned)n, 32, 0, nullptr>>>(P, reinterpret_cast<const double*>(db + TINY_OFF_X),
                                                             reinterpret_cast<unsigned*>(db + TINY_OFF_STATUS));
        HFE_CUDA_TRY(cudaGetLastError());
        HFE_CUDA_TRY(cudaStreamSynchronize(nullptr));
        if (*reinterpret_cast<volatile uint32_t*>(hb + TINY_OFF_STATUS) & HFE_ROW_NONFINITE)
            return fail(HFE_ERR_NONFINITE_INPUT, "Input X contains NaN or infinity");
        if (proba_n2) std::memcpy(proba_n2, hb + TINY_OFF_OUT, (size_t)n * 2 * sizeof(double));
        if (labels) std::memcpy(labels, hb + TINY_OFF_LABELS, (size_t)n);
        if (decision) std::memcpy(decision, hb + TINY_OFF_DECISION, (size_t)n * sizeof(double));
        if (meta) std::memcpy(meta, hb + TINY_OFF_META, (size_t)n * m->n_members * sizeof(double));
        return HFE_OK;
    }
    if (n >= HOST_PIPE_MIN_ROWS) {
        int rc = predict_host_pipelined(model, x_rowmajor, n, proba_n2, labels, decision, meta, false);
        if (rc == -1) rc = predict_host_pipelined(model, x_rowmajor, n, proba_n2, labels, decision, meta, true);
        return rc;
    }
    int64_t ld = (n + 3) & ~(int64_t)3;
    size_t in_bytes = (size_t)n * F * sizeof(double);
    size_t col_bytes = (size_t)ld * F * sizeof(float);
    size_t out_bytes = (size_t)n * 2 * sizeof(float);
    size_t dec_bytes = decision ? (size_t)n * sizeof(double) : 0;
    size_t meta_bytes = meta ? (size_t)n * m->n_members * sizeof(double) : 0;
    size_t lab_bytes = labels ? (size_t)n : 0;
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    size_t status_bytes = (HFE_STATUS_WORDS + HFE_STATUS_HEADER_WORDS) * sizeof(uint32_t);
    size_t total = up(in_bytes) + up(col_bytes) + up(out_bytes) + up(dec_bytes) + up(meta_bytes) + up(lab_bytes) +
                   up(status_bytes) + 256;
    unsigned char* buf = nullptr;
    std::unique_lock<std::mutex> small_lock(m->small_mutex, std::defer_lock);
    struct Guard { unsigned char* p; ~Guard() { if (p) cudaFreeAsync(p, nullptr); } } guard{nullptr};
    if (n <= SMALL_HOST_ROWS && m->small_arena && total <= m->small_arena_bytes) {
        small_lock.lock();                      // the arena is shared by all small host calls on this handle
        buf = m->small_arena;
    } else {
        HFE_CUDA_TRY(cudaMallocAsync(&buf, total, nullptr));     // stream-ordered pool: no device-wide sync per call
        guard.p = buf;
    }
    unsigned char* cur = buf;
    auto take = [&](size_t b) { unsigned char* p = cur; cur += up(b); return p; };
    double* d_in = reinterpret_cast<double*>(take(in_bytes));
    float* d_x = reinterpret_cast<float*>(take(col_bytes));
    float* d_out = reinterpret_cast<float*>(take(out_bytes));
    double* d_dec = decision ? reinterpret_cast<double*>(take(dec_bytes)) : nullptr;
    double* d_meta = meta ? reinterpret_cast<double*>(take(meta_bytes)) : nullptr;
    uint8_t* d_lab = labels ? take(lab_bytes) : nullptr;
    uint32_t* d_status = reinterpret_cast<uint32_t*>(take(status_bytes));   // [ingest header | scoring status]
    cudaStream_t s = nullptr;
    HFE_CUDA_TRY(cudaMemcpyAsync(d_in, x_rowmajor, in_bytes, cudaMemcpyHostToDevice, s));
    int rc = hfe_ingest_rowmajor_f64(d_in, n, F, d_x, ld, d_status, s);
    if (rc) return rc;
    rc = score_common(model, d_x, n, ld, -1, nullptr, d_out, nullptr, d_lab, d_dec, d_meta, nullptr, nullptr,
                      d_status + HFE_STATUS_HEADER_WORDS, s, false);
    if (rc) return rc;
    std::vector<float> host_out((size_t)n * 2);
    uint32_t host_status[2 * HFE_STATUS_HEADER_WORDS];
    HFE_CUDA_TRY(cudaMemcpyAsync(host_out.data(), d_out, out_bytes, cudaMemcpyDeviceToHost, s));
    if (decision) HFE_CUDA_TRY(cudaMemcpyAsync(decision, d_dec, dec_bytes, cudaMemcpyDeviceToHost, s));
    if (meta) HFE_CUDA_TRY(cudaMemcpyAsync(meta, d_meta, meta_bytes, cudaMemcpyDeviceToHost, s));
    if (labels) HFE_CUDA_TRY(cudaMemcpyAsync(labels, d_lab, lab_bytes, cudaMemcpyDeviceToHost, s));
    HFE_CUDA_TRY(cudaMemcpyAsync(host_status, d_status, sizeof(host_status), cudaMemcpyDeviceToHost, s));
    HFE_CUDA_TRY(cudaStreamSynchronize(s));
    if ((host_status[0] | host_status[HFE_STATUS_HEADER_WORDS]) & HFE_ROW_NONFINITE)
        return fail(HFE_ERR_NONFINITE_INPUT, "Input X contains NaN or infinity");
    if (host_status[HFE_STATUS_HEADER_WORDS] & HFE_WORKLIST_OVERFLOW) {
        // more rows than the work list holds needed float64: score the whole batch on the float64 path instead
        rc = score_common(model, d_x, n, ld, -1, nullptr, d_out, nullptr, d_lab, d_dec, d_meta, nullptr, nullptr,
                          d_status + HFE_STATUS_HEADER_WORDS, s, true);
        if (rc) return rc;
        HFE_CUDA_TRY(cudaMemcpyAsync(host_out.data(), d_out, out_bytes, cudaMemcpyDeviceToHost, s));
        if (decision) HFE_CUDA_TRY(cudaMemcpyAsync(decision, d_dec, dec_bytes, cudaMemcpyDeviceToHost, s));
        if (meta) HFE_CUDA_TRY(cudaMemcpyAsync(meta, d_meta, meta_bytes, cudaMemcpyDeviceToHost, s));
        if (labels) HFE_CUDA_TRY(cudaMemcpyAsync(labels, d_lab, lab_bytes, cudaMemcpyDeviceToHost, s));
        HFE_CUDA_TRY(cudaStreamSynchronize(s));
    }
    if (proba_n2) {
        for (int64_t i = 0; i < n; ++i) {
            double p1 = (double)host_out[2 * i + 1];
            proba_n2[2 * i] = 1.0 - p1;
            proba_n2[2 * i + 1] = p1;
        }
    }
    return HFE_OK;
    } catch (const std::exception& e) {       // nothing may unwind through the C boundary
        return fail(HFE_ERR_INVALID_ARGUMENT, std::string("out of host memory or internal error: ") + e.what());
    }
}

}  // extern "C"
