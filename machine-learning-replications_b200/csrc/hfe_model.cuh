// hfe_model.cuh -- host side of the model handle: blob validation and the derived device tables.
//
// The blob (export/blob.py) carries scikit-learn's canonical float64 parameters.  Everything the kernels read is
// derived here, once, so that every rounding decision lives in one place:
//   * RBF-SVC (SK/svm/src/libsvm/svm.cpp:461-515, 2821-2904): exp(-g|z-s|^2) = 2^(2c z.s) * 2^(-c|s|^2) * 2^(-c|z|^2),
//     c = g*log2(e).  Per support vector the float32 table holds w = 2c*s (17 values) and coef' = coef * 2^(-c|s|^2);
//     the row factor 2^(-c|z|^2) is applied once in float64.  Support vectors are re-ordered so positive and negative
//     coefficients alternate (running sum near zero): the float32 partial sums stay small, which is what bounds the
//     accumulation error (DESIGN.md "error budget").
//   * boosted trees (SK/ensemble/_gradient_boosting.pyx:25-73): thresholds rounded toward -inf to float32 so that
//     (x32 <= thr32) == ((double)x32 <= thr64) for every float32 x -- leaf indices stay bit-exact; nodes re-laid in
//     breadth-first order with siblings adjacent (next = child + (x > thr)), leaves made self-looping so a fixed-depth
//     predicated walk needs no per-lane exit test; leaf values pre-multiplied by the learning rate in float64.
//   * depth-1 stumps with few distinct (feature, threshold) pairs are merged into a split table (100 published
//     stumps -> 17 splits).
#pragma once

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/hfe.h"
#include "hfe_solver.cuh"

namespace hfe {

constexpr int F = HFE_N_FEATURES;
constexpr int MAX_MEMBERS = 4;
constexpr int MAX_SPLITS = 128;       // merged stump table capacity (kernel parameter space)
constexpr int SV_GROUP = 4;           // support vectors per inner tile
constexpr int SV_STEPS = F + 1;       // float4 steps per group: F weight steps + one folded-coefficient step
constexpr int SV_PAD_ROWS = 2;        // zero steps after the table so the 2-step-ahead prefetch may over-read
constexpr int SV_PREFIX = 9;          // prefix-shared layout: leading (re-ordered) features whose partial exponent is
                                      // computed once per prefix quad and shared by the support vectors below it
constexpr int SV_LEAF_STEPS = F - SV_PREFIX + 1;   // per leaf step: the remaining weights + the folded coefficients
static_assert(SV_PREFIX % 3 == 0 && SV_LEAF_STEPS % 3 == 0 && SV_STEPS % 3 == 0, "3-deep prefetch ring phase");
static_assert((F - SV_PREFIX) % 2 == 0, "leaf steps hold two features per float4");
#ifndef HFE_TINY_HOST_ROWS
#define HFE_TINY_HOST_ROWS 4096           // measured: 1 row 45 us, 1024 rows 58 us (staged path 109 us), 4096 rows 116 us (155 us)
#endif
constexpr int TINY_HOST_ROWS = HFE_TINY_HOST_ROWS;      // host batches up to this size take the zero-copy single-kernel path
constexpr int SMALL_HOST_ROWS = 4096;   // host batches up to this size use the model's preallocated staging arena
constexpr int MAX_EDGES = 16;          // coupling-solver discontinuities (6 exist for eps = 0.0025)
#ifndef HFE_SV_CHUNK
#define HFE_SV_CHUNK 4
#endif
constexpr int SV_CHUNK = HFE_SV_CHUNK;           // float32 partial sums are folded into float64 every SV_CHUNK vectors

enum MemberKind { KIND_NONE = 0, KIND_SVC = 1, KIND_GBT = 2, KIND_LIN = 3 };

// Parameter block handed to every scoring kernel by value (lives in the constant bank; fully unrolled feature loops
// turn the small tables into immediate constant operands).
struct ScoreParams {
    // request
    const float* x;
    long long n, ld;
    long long col_off[F];       // element offset of internal feature f's column in x: (caller's column) * ld -- the
                                // model re-orders features (build_model); set per call
    const int* perm_idx;
    int perm_col;
    int n_members;
    float* out;                 // [n,2] f32 or null
    double* out64;              // [n,2] f64 or null
    unsigned char* labels;      // [n] or null
    double* decision;           // [n] or null
    double* meta;               // [n, n_members] or null
    double* svc_dec;            // [n] or null: the SVC member's decision value as this pass computed it (guard probe)
    const unsigned char* y;     // [n] or null
    unsigned long long* score;  // [1] or null
    unsigned int* status;       // [HFE_STATUS_WORDS]
    // SVC member
    const float4* sv_tab;       // [n_groups][SV_STEPS] float4: step d = weight d of 4 vectors (+ SV_PAD_ROWS zeros)
    const double* sv64;         // [n_sv][F] original order
    const double* dual64;       // [n_sv]
    int n_groups;               // multiple of SV_CHUNK / SV_GROUP
    int n_sv;
    int sv_quads;               // > 0: prefix-shared layout (build_svc), else the flat [n_groups][SV_STEPS] layout
    int sv_bytes;               // bytes staged in shared memory: table + pad steps (+ the per-quad step counts)
    double mean[F], scale[F];
    float mean_hi[F], mean_lo[F], inv_scale32[F];   // mean = hi + lo in float32; 1/scale rounded once
    double c_log2;              // gamma * log2(e)
    double gamma, svc_intercept, prob_a, prob_b;
    double zz_limit;            // rows with gamma*|z|^2 above this go to the float64 path (float32 2^x range)
    double solver_guard;        // Platt argument within this of a solver edge => the solver's branch is uncertain
    double edges[MAX_EDGES];    // values of t = f*A+B where the coupling solver's sweep count changes
    const double* coupling_coef;    // the coupling map as a table (hfe_solver.cuh): [cells][2][4] cubic coefficients
    const double* coupling_edge;    // [cells] local coordinate of the cell's edge (1e9: none)
    int n_edges;
    int pad_edges_;
    double label_guard;         // |z_final| below this => hard label is uncertain
    // boosted trees
    int n_trees;
    int n_splits;
    double gbt_init;            // init raw prediction
    double gbt_split_base;      // init + sum of right-leaf values (merged stump mode)
    float split1_thr[F], split1_delta[F];   // first merged split of each feature (-inf / 0 where there is none)
    int split_begin[F + 1];                 // the remaining splits of feature f: [split_begin[f], split_begin[f+1])
    float split_thr[MAX_SPLITS];
    float split_delta[MAX_SPLITS];
    unsigned char split_feat[MAX_SPLITS];   // feature of each remaining split (the streaming pass walks them in one loop)
    int n_extra_splits;
    int pad_splits_;
    const uint2* nodes;         // {thr32 bits, feature<<24 | child}: child = GLOBAL index of the left child (right = +1)
    const float* value32;       // lr * value, per node
    const double* value64;
    const int* tree_off;        // [n_trees+1]
    const unsigned char* tree_depth;  // [n_trees]
    const unsigned short* sk_id;      // scikit-learn node id per re-laid node
    int max_depth;                    // deepest tree (uniform loop bound of the warp-cooperative float64 walk)
    int pad_depth_;
    // logistic member
    double lin_coef[F];
    double lin_intercept;
    // combiner
    double final_coef[MAX_MEMBERS];
    double final_intercept;
    int slot_svc, slot_gbt, slot_lin;
    int pad_;
};

// Tensor-core path of the RBF sum (hfe_mma.cuh): the B operand image and what the producers need to build the A operand.
constexpr int MMA_MAX_SV = 512;       // the 512 TMEM columns hold the blocks in flight
#ifndef HFE_MMA_REGIONS
#define HFE_MMA_REGIONS 2             // TMEM regions per stream: 2 x 128 columns (1.55 ms) or 4 x 64 (1.72 ms: the per-block bookkeeping outweighs the deeper pipeline)
#endif
constexpr int MMA_REGIONS = HFE_MMA_REGIONS;             // blocks of a stream that can be in TMEM at once
constexpr int MMA_REGION_COLS = 256 / MMA_REGIONS;       // TMEM columns per region = the largest block of support vectors
constexpr int MMA_BARRIERS = 56;                         // mbarrier slots in shared memory
constexpr int MMA_B_CHUNKS = 10;      // 16-byte K chunks per support vector: WL0 WL1 WL2 WM0 WM1 WM2 WH0 WH1 WH2 0

struct MmaParams {
    const unsigned char* w_img;     // [n_pad/8][MMA_B_CHUNKS][8][16 B] bfloat16 B operand, then float coef'[n_pad]
    int w_bytes;                    // image + coefficient table
    int nb;                         // blocks of block_sv support vectors; n_pad = nb * block_sv; 0 = path not available
    int block_sv;                   // multiple of 16, at most 128 (N of one MMA instruction = columns of one TMEM region)
    int pad_;
    float center[F];                // m_d: the scaler mean rounded to an integer
    float zz_a[F], zz_b[F];         // |z|^2 = sum_d xc_d (a_d xc_d + b_d) + c,  xc = x - m
    float zz_c;
    unsigned extra_split_mask;      // bit f: feature f has more than one merged stump split
    float wmax_all;                 // max_dj |W_dj| (slightly enlarged): |exponent of a row| <= wmax_all * sum_d |xc_d|
    double lin_base;                // logistic member: intercept + sum_d m_d * coef_d; the dot product over x - m is the last
                                    // column of the B operand (coefficient 0 in the RBF sum) and comes off the tensor core
};

// General boosted trees walked out of shared memory (hfe_forest.cuh).  Per feature the distinct thresholds of the whole
// forest are sorted; a row carries, per feature, the number of thresholds below its value (one byte), and a node is 4
// bytes -- index of its threshold in that feature's list, feature, byte offset of the left child inside the staged chunk -- so
// a level of the walk is two shared-memory loads and a handful of integer instructions.
constexpr int FOREST_MAX_THR = 256;            // distinct (feature, threshold) pairs of the forest, at most 255 per feature
constexpr int FOREST_CHUNK_NODES = 8192;       // nodes staged per chunk (32 KB); two chunks are resident

struct ForestParams {
    const unsigned* nodes4;         // per node: threshold index << 24 | feature << 16 | byte offset of the left child from
                                    // the start of its chunk's copy (leaves: index 255 -- never exceeded -- and their own offset)
    const int* chunk_tree;          // [n_chunks + 1] first tree of each chunk
    int n_chunks;
    int n_thr;
    int thr_begin[F + 1];           // thresholds of feature f: [thr_begin[f], thr_begin[f + 1]), ascending
    float thr[FOREST_MAX_THR];
};

struct HostPipeline;            // hfe_host.cuh: pinned / device slots and the thread pool of large host calls

struct Model {
    int device = 0;
    int n_members = 0;
    int kinds[MAX_MEMBERS] = {0, 0, 0, 0};
    int n_sv = 0, n_trees = 0, n_nodes = 0, max_depth = 0, tree_mode = 0, n_splits = 0;
    int sv_steps = 0;             // prefix-shared layout: leaf steps (4 support-vector lanes each); 0 = flat layout
    int col_of[F];                // caller's column of internal feature f
    int internal_of[F];           // internal feature of caller's column c
    double svc_guard = 0.0;       // assumed |error| of the float32 SVC decision value
    double gbt_guard = 0.0;       // assumed |error| of the float32 boosted raw score
    double svc_guard_formula = 0.0;   // svc_guard as build_svc derives it from the dual coefficients (before any widening)
    double guard_scale = 1.0;     // multiplies svc_guard_formula (hfe_model_calibrate_guards; 1 at creation)
    double svc_observed_err = 0.0;    // creation-time probe: max |float32 - float64| of the SVC decision value
    double svc_observed_by_engine[2] = {0.0, 0.0};   // ... per engine: SIMT loop, tensor cores
    double label_observed_err = 0.0;  // ... and of the combiner's decision value, over rows the fast pass did not flag
    int probe_rows = 0;           // rows the probe scored (0: no SVC member, or no device at plan time)
    int guard_widened = 0;        // 1: the probe found svc_guard < 8 x observed and replaced it by 8 x observed
    ScoreParams base{};           // request fields zero; table pointers filled
    ForestParams forest{};        // n_chunks > 0: general trees can be walked out of shared memory
    int tree_engine = 0;          // 0 = automatic, 1 = node tables through L1 (__ldg), 2 = shared-memory chunks,
                                  // 3 = like 1 with the node table pinned in an L2 persisting window (measurement aid)
    MmaParams mma{};              // nb > 0: the RBF sum can run on the tensor cores
    std::vector<unsigned char> mma_img_host;   // B operand image while the model is being built (uploaded by build_model)
    int rbf_engine = 0;           // 0 = automatic (tensor cores when available), 1 = SIMT FFMA2 loop, 2 = tensor cores
    std::vector<void*> allocations;
    uint64_t device_bytes = 0;
    // staging for small host batches (hfe_predict_host with n <= SMALL_HOST_ROWS): allocated once with the model so a
    // single-patient call (predict_hf.py) pays no cudaMalloc / cudaFree; one caller at a time (mutex)
    mutable std::mutex small_mutex;
    unsigned char* small_arena = nullptr;
    size_t small_arena_bytes = 0;
    // a few rows (predict_hf.py scores one patient): pinned host memory mapped into the device address space; one
    // kernel reads the float64 rows from it and writes every output back into it -- no copies, no second launch
    unsigned char* tiny_host = nullptr;          // host view
    unsigned char* tiny_dev = nullptr;           // the same bytes as the device sees them
    // large host batches (train_ensemble_public.py:62 on a big cohort): chunk pipeline, built on first use
    mutable std::mutex pipe_mutex;
    mutable HostPipeline* pipe = nullptr;
};

inline thread_local std::string g_last_error;

inline int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define HFE_CUDA_TRY(expr)                                                                              \
    do {                                                                                                \
        cudaError_t err__ = (expr);                                                                     \
        if (err__ != cudaSuccess)                                                                       \
            return ::hfe::fail(HFE_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(err__));    \
    } while (0)

// ---- blob parsing ----------------------------------------------------------------------------------------------------

inline uint32_t crc32_zlib(const uint8_t* p, size_t n) {
    static uint32_t table[8][256];
    static std::once_flag once;                 // two threads may make their first hfe_model_create call together
    std::call_once(once, [] {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1) ? (0xEDB88320u ^ (c >> 1)) : (c >> 1);
            table[0][i] = c;
        }
        for (uint32_t i = 0; i < 256; ++i)
            for (int t = 1; t < 8; ++t) table[t][i] = (table[t - 1][i] >> 8) ^ table[0][table[t - 1][i] & 0xFF];
    });
    uint32_t c = 0xFFFFFFFFu;
    while (n >= 8) {
        uint32_t a, b;
        std::memcpy(&a, p, 4);
        std::memcpy(&b, p + 4, 4);
        a ^= c;
        c = table[7][a & 0xFF] ^ table[6][(a >> 8) & 0xFF] ^ table[5][(a >> 16) & 0xFF] ^ table[4][a >> 24] ^
            table[3][b & 0xFF] ^ table[2][(b >> 8) & 0xFF] ^ table[1][(b >> 16) & 0xFF] ^ table[0][b >> 24];
        p += 8;
        n -= 8;
    }
    while (n--) c = table[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
    return c ^ 0xFFFFFFFFu;
}

struct Section {
    const uint8_t* data = nullptr;
    uint64_t count = 0;
    const double* f64() const { return reinterpret_cast<const double*>(data); }
    const int32_t* i32() const { return reinterpret_cast<const int32_t*>(data); }
    const uint32_t* u32() const { return reinterpret_cast<const uint32_t*>(data); }
};

struct ParsedBlob {
    uint32_t n_features = 0, n_members = 0;
    std::map<std::pair<uint32_t, uint32_t>, Section> sections;
    const Section* find(uint32_t kind, uint32_t member, uint64_t count) const {
        auto it = sections.find({kind, member});
        if (it == sections.end() || it->second.count != count) return nullptr;
        return &it->second;
    }
    bool has(uint32_t kind, uint32_t member) const { return sections.count({kind, member}) != 0; }
};

enum : uint32_t {
    K_FINAL = 1, K_CLASSES = 2, K_SVC_PARAMS = 10, K_SVC_MEAN = 11, K_SVC_SCALE = 12, K_SVC_SV = 13, K_SVC_DUAL = 14,
    K_GBT_PARAMS = 20, K_GBT_OFFSETS = 21, K_GBT_FEATURE = 22, K_GBT_LEFT = 23, K_GBT_RIGHT = 24,
    K_GBT_THRESHOLD = 25, K_GBT_VALUE = 26, K_LIN_PARAMS = 30, STACK_LEVEL = 0xFFFFFFFFu
};

inline int elem_size(uint32_t kind) {
    switch (kind) {
        case K_GBT_OFFSETS: case K_GBT_FEATURE: case K_GBT_LEFT: case K_GBT_RIGHT: return 4;
        case K_FINAL: case K_CLASSES: case K_SVC_PARAMS: case K_SVC_MEAN: case K_SVC_SCALE: case K_SVC_SV:
        case K_SVC_DUAL: case K_GBT_PARAMS: case K_GBT_THRESHOLD: case K_GBT_VALUE: case K_LIN_PARAMS: return 8;
        default: return 0;
    }
}

inline int parse_blob(const uint8_t* b, size_t len, ParsedBlob* out) {
    if (len < 64 || std::memcmp(b, "HFEBLOB1", 8) != 0) return fail(HFE_ERR_BAD_BLOB, "blob: bad magic or truncated");
    uint32_t version, hbytes, crc, nf, nm, ns;
    uint64_t total;
    std::memcpy(&version, b + 8, 4);
    std::memcpy(&hbytes, b + 12, 4);
    std::memcpy(&total, b + 16, 8);
    std::memcpy(&crc, b + 24, 4);
    std::memcpy(&nf, b + 28, 4);
    std::memcpy(&nm, b + 32, 4);
    std::memcpy(&ns, b + 36, 4);
    if (version != 1 || hbytes != 64) return fail(HFE_ERR_BAD_BLOB, "blob: unknown version");
    if (total != len) return fail(HFE_ERR_BAD_BLOB, "blob: total_bytes does not match the buffer length");
    if (nf < 1 || nf > 4096 || nm < 1 || nm > 16 || ns < 2 || ns > 256)
        return fail(HFE_ERR_BAD_BLOB, "blob: implausible header counts");
    uint64_t pay0 = (64 + 32ull * ns + 15) & ~15ull;
    if (pay0 > len) return fail(HFE_ERR_BAD_BLOB, "blob: section directory runs past the end");
    if (crc32_zlib(b + pay0, len - pay0) != crc) return fail(HFE_ERR_BAD_BLOB, "blob: checksum mismatch");
    out->n_features = nf;
    out->n_members = nm;
    for (uint32_t i = 0; i < ns; ++i) {
        const uint8_t* d = b + 64 + 32ull * i;
        uint32_t kind, member;
        uint64_t off, nbytes, count;
        std::memcpy(&kind, d, 4);
        std::memcpy(&member, d + 4, 4);
        std::memcpy(&off, d + 8, 8);
        std::memcpy(&nbytes, d + 16, 8);
        std::memcpy(&count, d + 24, 8);
        int es = elem_size(kind);
        if (es == 0) return fail(HFE_ERR_BAD_BLOB, "blob: unknown section kind");
        if ((off & 15) || off < pay0 || off > len || nbytes > len - off || count > (1ull << 40) ||
            nbytes != count * (uint64_t)es)
            return fail(HFE_ERR_BAD_BLOB, "blob: section out of bounds or mis-sized");
        if (out->sections.count({kind, member})) return fail(HFE_ERR_BAD_BLOB, "blob: duplicate section");
        out->sections[{kind, member}] = Section{b + off, count};
    }
    return HFE_OK;
}

inline bool all_finite(const double* p, uint64_t n) {
    for (uint64_t i = 0; i < n; ++i)
        if (!std::isfinite(p[i])) return false;
    return true;
}

// largest float32 <= v  (so that x32 <= result  <=>  (double)x32 <= v)
inline float floor_to_f32(double v) {
    float f = (float)v;
    if ((double)f > v) f = std::nextafterf(f, -INFINITY);
    return f;
}

template <typename T>
inline int upload(Model* m, const std::vector<T>& host, const T** dev) {
    void* p = nullptr;
    size_t bytes = std::max<size_t>(host.size() * sizeof(T), 16);
    if (m->device < 0) {                        // planning only (hfe_model_plan): sizes are counted, nothing is uploaded
        m->device_bytes += bytes;
        *dev = nullptr;
        return HFE_OK;
    }
    HFE_CUDA_TRY(cudaMalloc(&p, bytes));
    m->allocations.push_back(p);
    m->device_bytes += bytes;
    if (!host.empty()) HFE_CUDA_TRY(cudaMemcpy(p, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice));
    *dev = static_cast<const T*>(p);
    return HFE_OK;
}

// ---- SVC member --------------------------------------------------------------------------------------------------------

inline int build_svc(Model* m, const ParsedBlob& pb, uint32_t idx) {
    const Section* ps = pb.find(K_SVC_PARAMS, idx, 8);
    if (!ps) return fail(HFE_ERR_BAD_BLOB, "blob: SVC params missing");
    const double* p = ps->f64();
    if (!all_finite(p, 8)) return fail(HFE_ERR_BAD_BLOB, "blob: non-finite SVC parameter");
    double n_sv_d = p[4];
    if (!(n_sv_d >= 1 && n_sv_d <= 1e6) || n_sv_d != std::floor(n_sv_d))
        return fail(HFE_ERR_BAD_BLOB, "blob: bad support vector count");
    int n_sv = (int)n_sv_d;
    // the consistency check scikit-learn applies before handing arrays to libsvm (SK/svm/_base.py:680-686)
    if (p[5] < 0 || p[6] < 0 || p[5] + p[6] != n_sv_d)
        return fail(HFE_ERR_BAD_BLOB, "blob: SVC n_support does not add up to the support vector count");
    const Section* mean = pb.find(K_SVC_MEAN, idx, F);
    const Section* scale = pb.find(K_SVC_SCALE, idx, F);
    const Section* sv = pb.find(K_SVC_SV, idx, (uint64_t)n_sv * F);
    const Section* dual = pb.find(K_SVC_DUAL, idx, n_sv);
    if (!mean || !scale || !sv || !dual) return fail(HFE_ERR_BAD_BLOB, "blob: SVC section missing or mis-sized");
    if (!all_finite(mean->f64(), F) || !all_finite(scale->f64(), F) || !all_finite(sv->f64(), (uint64_t)n_sv * F) ||
        !all_finite(dual->f64(), n_sv))
        return fail(HFE_ERR_BAD_BLOB, "blob: non-finite SVC table entry");
    double gamma = p[0];
    if (!(gamma > 0)) return fail(HFE_ERR_BAD_BLOB, "blob: gamma must be positive");
    ScoreParams& P = m->base;
    for (int d = 0; d < F; ++d) {
        if (scale->f64()[d] == 0.0) return fail(HFE_ERR_BAD_BLOB, "blob: scaler scale_ holds a zero");
        P.mean[d] = mean->f64()[d];
        P.scale[d] = scale->f64()[d];
        P.mean_hi[d] = (float)P.mean[d];
        P.mean_lo[d] = (float)(P.mean[d] - (double)P.mean_hi[d]);
        P.inv_scale32[d] = (float)(1.0 / P.scale[d]);
    }
    P.gamma = gamma;
    P.c_log2 = gamma * 1.4426950408889634074;
    P.svc_intercept = p[1];
    P.prob_a = p[2];
    P.prob_b = p[3];
    P.n_sv = n_sv;
    m->n_sv = n_sv;

    // alternate signs so the running coefficient sum hovers around zero
    const double* coef = dual->f64();
    std::vector<int> pos, neg, order;
    for (int j = 0; j < n_sv; ++j) (coef[j] > 0 ? pos : neg).push_back(j);
    double run = 0;
    size_t ip = 0, in = 0;
    while (ip < pos.size() || in < neg.size()) {
        int j;
        if ((run <= 0 && ip < pos.size()) || in >= neg.size()) j = pos[ip++];
        else j = neg[in++];
        run += coef[j];
        order.push_back(j);
    }
    int padded = ((n_sv + SV_CHUNK - 1) / SV_CHUNK) * SV_CHUNK;
    int n_groups = padded / SV_GROUP;
    std::vector<float4> tab((size_t)n_groups * SV_STEPS + SV_PAD_ROWS, make_float4(0.f, 0.f, 0.f, 0.f));
    double max_norm = 0, coef_l2 = 0;
    for (int k = 0; k < n_sv; ++k) {
        int j = order[k];
        const double* s = sv->f64() + (size_t)j * F;
        double ss = 0;
        for (int d = 0; d < F; ++d) ss += s[d] * s[d];
        max_norm = std::max(max_norm, std::sqrt(ss));
        coef_l2 += coef[j] * coef[j];
        float* base = reinterpret_cast<float*>(&tab[(size_t)(k / SV_GROUP) * SV_STEPS]);
        int lane = k % SV_GROUP;
        for (int d = 0; d < F; ++d) base[d * 4 + lane] = (float)(2.0 * P.c_log2 * s[d]);
        base[F * 4 + lane] = (float)(coef[j] * std::exp2(-P.c_log2 * ss));
    }
    // float32 2^x must not overflow: 2c|z||s| <= 100  =>  c|z|^2 <= c*(100/(2c*max|s|))^2
    double zmax = 100.0 / (2.0 * P.c_log2 * std::max(max_norm, 1e-30));
    P.zz_limit = std::min(gamma * zmax * zmax, 600.0);          // compared against gamma*|z|^2
    // assumed uncertainty of the float32 decision value (DESIGN.md "error budget"): ~40 sigma of the error measured on
    // cohort rows, and >= 8 x the worst error the creation-time probe sees AT the support vectors (where the kernel values,
    // hence the absolute rounding errors, are largest) for both fixture models -- the probe checks it for every model
    m->svc_guard_formula = 8e-7 * std::sqrt(coef_l2) + 1e-7;
    m->svc_guard = m->svc_guard_formula;
    P.n_groups = n_groups;
    P.sv_quads = 0;
    P.sv_bytes = (int)(tab.size() * sizeof(float4));

    // ---- prefix-shared layout -------------------------------------------------------------------------------------
    // Support vectors are training rows, and most cohort features are binary, so many vectors agree on their first
    // SV_PREFIX (re-ordered) features.  The partial exponent over those features is then computed once per distinct
    // prefix and continued by each vector below it: the same FMA chain as the flat layout, feature by feature, so the
    // float32 result is the one the flat loop would give for this feature order.  The four SIMD lanes of a table step
    // walk four independent (prefix, vectors) chunks in lock step; a "quad" is four chunks, its step count the longest
    // of them, shorter lanes are padded with zero-weight, zero-coefficient vectors.  Large prefix classes are split
    // into chunks of at most T vectors (the prefix is then recomputed per chunk); T minimises a two-pipe cost model.
    if (n_sv <= 65536) {
        std::vector<int> idx(n_sv);
        for (int j = 0; j < n_sv; ++j) idx[j] = j;
        auto prefix_less = [&](int a, int b) {
            const double* sa = sv->f64() + (size_t)a * F;
            const double* sb = sv->f64() + (size_t)b * F;
            for (int d = 0; d < SV_PREFIX; ++d)
                if (sa[d] != sb[d]) return sa[d] < sb[d];
            return false;
        };
        std::stable_sort(idx.begin(), idx.end(), prefix_less);
        std::vector<std::pair<int, int>> classes;          // (first position in idx, size)
        for (int k = 0; k < n_sv; ++k) {
            if (k == 0 || prefix_less(idx[k - 1], idx[k])) classes.push_back({k, 1});
            else classes.back().second++;
        }
        auto pipe_cost = [](double fma_lane_ops, double xu_ops) {
            double a = fma_lane_ops / 128.0, b = xu_ops / 16.0;       // clocks per row on the FMA and MUFU pipes of one SM
            return std::max(a, b) + 0.3 * std::min(a, b);
        };
        auto split_sizes = [&](int T, std::vector<int>* sizes) {
            sizes->clear();
            for (auto& c : classes) {
                int parts = (c.second + T - 1) / T, base = c.second / parts, extra = c.second % parts;
                for (int i = 0; i < parts; ++i) sizes->push_back(base + (i < extra ? 1 : 0));
            }
            std::sort(sizes->begin(), sizes->end(), std::greater<int>());
        };
        int largest = 0;
        for (auto& c : classes) largest = std::max(largest, c.second);
        int best_T = 0;
        double best_cost = 0;
        std::vector<int> sizes;
        for (int T = 1; T <= std::min(largest, 255); ++T) {
            split_sizes(T, &sizes);
            long long quads = ((long long)sizes.size() + 3) / 4, steps = 0;
            for (size_t q = 0; q < sizes.size(); q += 4) steps += sizes[q];
            double cost = pipe_cost(4.0 * (quads * SV_PREFIX + steps * SV_LEAF_STEPS), 5.0 * steps);
            if (best_T == 0 || cost < best_cost) { best_T = T; best_cost = cost; }
        }
#ifdef HFE_DEV
        if (const char* t_env = std::getenv("HFE_PFX_CHUNK")) {         // development builds only: force the chunk size
            int t = std::atoi(t_env);
            if (t >= 1 && t <= 255) { best_T = t; best_cost = 0.0; }
        }
#endif
        double flat_cost = pipe_cost(4.0 * n_groups * SV_STEPS, 5.0 * n_groups);
        if (best_T > 0 && best_cost < 0.9 * flat_cost) {
            struct Chunk { int first, count; };
            std::vector<Chunk> chunks;
            for (auto& c : classes) {
                int parts = (c.second + best_T - 1) / best_T, base = c.second / parts, extra = c.second % parts, at = c.first;
                for (int i = 0; i < parts; ++i) {
                    int cnt = base + (i < extra ? 1 : 0);
                    chunks.push_back({at, cnt});
                    at += cnt;
                }
            }
            std::stable_sort(chunks.begin(), chunks.end(), [](const Chunk& a, const Chunk& b) { return a.count > b.count; });
            int quads = ((int)chunks.size() + 3) / 4;
            long long steps = 0;
            for (size_t q = 0; q < chunks.size(); q += 4) steps += chunks[q].count;
            size_t slots = (size_t)quads * SV_PREFIX + (size_t)steps * SV_LEAF_STEPS;
            size_t count_slots = ((size_t)quads + 15) / 16;
            std::vector<float4> ptab(slots + SV_PAD_ROWS + count_slots, make_float4(0.f, 0.f, 0.f, 0.f));
            unsigned char* counts = reinterpret_cast<unsigned char*>(&ptab[slots + SV_PAD_ROWS]);
            size_t at = 0;
            for (int q = 0; q < quads; ++q) {
                int qsteps = chunks[(size_t)q * 4].count;
                counts[q] = (unsigned char)qsteps;
                float* pre = reinterpret_cast<float*>(&ptab[at]);
                float* leaf = reinterpret_cast<float*>(&ptab[at + SV_PREFIX]);
                for (int lane = 0; lane < 4; ++lane) {
                    size_t ci = (size_t)q * 4 + lane;
                    if (ci >= chunks.size()) continue;                       // whole lane stays zero
                    const double* s0 = sv->f64() + (size_t)idx[chunks[ci].first] * F;
                    for (int d = 0; d < SV_PREFIX; ++d) pre[d * 4 + lane] = (float)(2.0 * P.c_log2 * s0[d]);
                    for (int k = 0; k < chunks[ci].count; ++k) {
                        int j = idx[chunks[ci].first + k];
                        const double* s = sv->f64() + (size_t)j * F;
                        double ss = 0;
                        for (int d = 0; d < F; ++d) ss += s[d] * s[d];
                        // leaf step: [coefficients of the 4 lanes][pair 0 (lanes 0,1): 4 float4, each two features x two
                        // lanes][pair 1 (lanes 2,3): the same]
                        float* step = leaf + (size_t)k * SV_LEAF_STEPS * 4;
                        step[lane] = (float)(coef[j] * std::exp2(-P.c_log2 * ss));
                        float* pair = step + 4 + (lane / 2) * ((F - SV_PREFIX) / 2) * 4;
                        for (int d = SV_PREFIX; d < F; ++d) {
                            int e = d - SV_PREFIX;
                            pair[(e / 2) * 4 + (e % 2) * 2 + (lane % 2)] = (float)(2.0 * P.c_log2 * s[d]);
                        }
                    }
                }
                at += SV_PREFIX + (size_t)qsteps * SV_LEAF_STEPS;
            }
            tab.swap(ptab);
            P.sv_quads = quads;
            P.sv_bytes = (int)(tab.size() * sizeof(float4));
            m->sv_steps = (int)steps;
        }
    }
    // ---- tensor-core operand image (hfe_mma.cuh) -------------------------------------------------------------------------
    // T_j = sum_d xc_d W_dj with W_dj carried as three bfloat16 parts; coef'_j = coef_j 2^(B_j).  Column j of the image is
    // support vector order[j] (signs alternating, so the float32 partial sums of the epilogue stay small); pad columns have
    // zero weights and a zero coefficient.
    std::vector<unsigned char> mma_img;
    if (n_sv < MMA_MAX_SV) {
        MmaParams& Q = m->mma;
        auto bf16_bits = [](float v) -> uint16_t {              // round to nearest even, like cvt.rn.bf16.f32
            uint32_t b;
            std::memcpy(&b, &v, 4);
            b += 0x7FFFu + ((b >> 16) & 1u);
            return (uint16_t)(b >> 16);
        };
        auto bf16_val = [](uint16_t h) { uint32_t b = (uint32_t)h << 16; float v; std::memcpy(&v, &b, 4); return v; };
        const int n_cols = n_sv + 1;                        // the last column carries the logistic member's dot product
        const int nb = (n_cols + MMA_REGION_COLS - 1) / MMA_REGION_COLS, per_block = (n_cols + nb - 1) / nb;
        const int bsv = ((per_block + 15) / 16) * 16, n_pad = nb * bsv;
        const size_t img_bytes = (size_t)n_pad * MMA_B_CHUNKS * 16;
        mma_img.assign(img_bytes + (size_t)n_pad * sizeof(float), 0);
        float* coef_tab = reinterpret_cast<float*>(mma_img.data() + img_bytes);
        double zz_c = 0.0;
        for (int d = 0; d < F; ++d) {
            Q.center[d] = (float)std::nearbyint(P.mean[d]);
            const double dm = (double)Q.center[d] - P.mean[d], iv = 1.0 / (P.scale[d] * P.scale[d]);
            Q.zz_a[d] = (float)iv;
            Q.zz_b[d] = (float)(2.0 * dm * iv);
            zz_c += dm * dm * iv;
        }
        Q.zz_c = (float)zz_c;
        Q.wmax_all = 0.f;
        bool usable = true;
        for (int col = 0; col < n_sv; ++col) {
            const int j = order[col];
            const double* s = sv->f64() + (size_t)j * F;
            uint16_t* row = reinterpret_cast<uint16_t*>(mma_img.data() + (size_t)(col / 8) * (MMA_B_CHUNKS * 128) + (col % 8) * 16);
            double bias = 0.0, ss = 0.0;
            for (int d = 0; d < F; ++d) {
                const double w = 2.0 * P.c_log2 * s[d] / P.scale[d];
                double rest = w;
                for (int part = 0; part < 3; ++part) {           // hi, mid, lo; chunks are 128 bytes (64 elements) apart
                    const uint16_t h = bf16_bits((float)rest);
                    rest -= (double)bf16_val(h);
                    row[(6 - 3 * part + d / 8) * 64 + d % 8] = h;       // lo parts in chunks 0-2, mid 3-5, hi 6-8
                }
                Q.wmax_all = std::max(Q.wmax_all, (float)(std::fabs(w) * 1.01));
                bias += ((double)Q.center[d] - P.mean[d]) * w;
                ss += s[d] * s[d];
            }
            bias -= P.c_log2 * ss;
            const double folded = coef[j] * std::exp2(bias);
            usable = usable && std::isfinite(folded) && std::fabs(bias) <= 100.0;
            coef_tab[col] = (float)folded;
        }
        // [B image + coefficients | 5 A stages | 12 ring slots | barriers] must fit one block's 227 KB (hfe_mma.cuh)
        usable = usable && ((mma_img.size() + 1023) & ~(size_t)1023) + 5 * 24576 + 12 * 2560 + MMA_BARRIERS * 8 + 128 <= 227 * 1024;
        if (usable) {
            Q.nb = nb;
            Q.block_sv = bsv;
            Q.w_bytes = (int)mma_img.size();
        } else {
            mma_img.clear();
        }
    }
    std::vector<double> sv_host(sv->f64(), sv->f64() + (size_t)n_sv * F), dual_host(coef, coef + n_sv);
    int rc;
    m->mma_img_host.swap(mma_img);          // the logistic member's column is added once that member has been read
    if ((rc = upload(m, tab, &P.sv_tab))) return rc;
    if ((rc = upload(m, sv_host, &P.sv64))) return rc;
    if ((rc = upload(m, dual_host, &P.dual64))) return rc;
    return HFE_OK;
}

// ---- boosted trees -----------------------------------------------------------------------------------------------------

inline int build_gbt(Model* m, const ParsedBlob& pb, uint32_t idx) {
    const Section* ps = pb.find(K_GBT_PARAMS, idx, 4);
    if (!ps || !all_finite(ps->f64(), 4)) return fail(HFE_ERR_BAD_BLOB, "blob: boosting params missing");
    const double* p = ps->f64();
    double init_raw = p[0], lr = p[1];
    if (!(p[2] >= 1 && p[2] <= 1e6 && p[3] >= 1 && p[3] < 16777216.0))      // 24-bit global node indices
        return fail(HFE_ERR_BAD_BLOB, "blob: bad tree counts");
    int n_trees = (int)p[2];
    int64_t n_nodes = (int64_t)p[3];
    const Section* offs = pb.find(K_GBT_OFFSETS, idx, (uint64_t)n_trees + 1);
    const Section* feat = pb.find(K_GBT_FEATURE, idx, n_nodes);
    const Section* left = pb.find(K_GBT_LEFT, idx, n_nodes);
    const Section* right = pb.find(K_GBT_RIGHT, idx, n_nodes);
    const Section* thr = pb.find(K_GBT_THRESHOLD, idx, n_nodes);
    const Section* val = pb.find(K_GBT_VALUE, idx, n_nodes);
    if (!offs || !feat || !left || !right || !thr || !val)
        return fail(HFE_ERR_BAD_BLOB, "blob: boosting section missing or mis-sized");
    const uint32_t* off = offs->u32();
    if (off[0] != 0 || off[n_trees] != (uint64_t)n_nodes) return fail(HFE_ERR_BAD_BLOB, "blob: bad tree offsets");

    std::vector<uint2> nodes((size_t)n_nodes);
    std::vector<float> value32((size_t)n_nodes);
    std::vector<double> value64((size_t)n_nodes);
    std::vector<unsigned short> sk_id((size_t)n_nodes);
    std::vector<int> tree_off(n_trees + 1);
    std::vector<unsigned char> depth(n_trees);
    bool all_stumps = true;
    int max_depth = 0;
    struct Split { int f; float thr; };
    std::map<std::pair<int, uint32_t>, double> merged;   // (feature, thr32 bits) -> sum lr*(vL - vR)
    double right_sum = 0, sum_max_abs = 0;
    std::vector<int> newpos, queue, dep;
    for (int t = 0; t < n_trees; ++t) {
        if (off[t + 1] <= off[t]) return fail(HFE_ERR_BAD_BLOB, "blob: empty tree");
        int base = (int)off[t], cnt = (int)(off[t + 1] - off[t]);
        if (cnt > 65535) return fail(HFE_ERR_UNSUPPORTED, "tree with more than 65535 nodes");
        tree_off[t] = base;
        const int32_t* L = left->i32() + base;
        const int32_t* R = right->i32() + base;
        const int32_t* Fe = feat->i32() + base;
        const double* T = thr->f64() + base;
        const double* V = val->f64() + base;
        // validate topology: children strictly after parent, in range, each node reached at most once
        std::vector<char> seen(cnt, 0);
        seen[0] = 1;
        for (int i = 0; i < cnt; ++i) {
            bool leaf = (L[i] == -1);
            if (leaf != (R[i] == -1)) return fail(HFE_ERR_BAD_BLOB, "blob: half-leaf tree node");
            if (!std::isfinite(V[i])) return fail(HFE_ERR_BAD_BLOB, "blob: non-finite leaf value");
            if (leaf) continue;
            if (L[i] <= i || R[i] <= i || L[i] >= cnt || R[i] >= cnt || L[i] == R[i])
                return fail(HFE_ERR_BAD_BLOB, "blob: tree child index out of order or range");
            if (seen[L[i]] || seen[R[i]]) return fail(HFE_ERR_BAD_BLOB, "blob: tree node has two parents");
            seen[L[i]] = seen[R[i]] = 1;
            if (Fe[i] < 0 || Fe[i] >= F) return fail(HFE_ERR_BAD_BLOB, "blob: split feature out of range");
            if (!std::isfinite(T[i])) return fail(HFE_ERR_BAD_BLOB, "blob: non-finite threshold");
        }
        for (int i = 0; i < cnt; ++i)
            if (!seen[i]) return fail(HFE_ERR_BAD_BLOB, "blob: unreachable tree node");
        // breadth-first re-layout, siblings adjacent
        newpos.assign(cnt, -1);
        queue.clear();
        dep.assign(cnt, 0);
        queue.push_back(0);
        newpos[0] = 0;
        int next_free = 1, tdepth = 0;
        for (size_t qi = 0; qi < queue.size(); ++qi) {
            int i = queue[qi];
            if (L[i] == -1) continue;
            newpos[L[i]] = next_free++;
            newpos[R[i]] = next_free++;
            dep[L[i]] = dep[R[i]] = dep[i] + 1;
            tdepth = std::max(tdepth, dep[i] + 1);
            queue.push_back(L[i]);
            queue.push_back(R[i]);
        }
        if (tdepth > 255) return fail(HFE_ERR_UNSUPPORTED, "tree deeper than 255 levels");
        depth[t] = (unsigned char)tdepth;
        max_depth = std::max(max_depth, tdepth);
        double tree_max = 0;
        for (int i = 0; i < cnt; ++i) tree_max = std::max(tree_max, std::fabs(lr * V[i]));
        sum_max_abs += tree_max;
        for (int i = 0; i < cnt; ++i) {
            int q = base + newpos[i];
            sk_id[q] = (unsigned short)i;
            value64[q] = lr * V[i];
            value32[q] = (float)value64[q];
            if (L[i] == -1) {
                float inf = INFINITY;
                uint32_t bits;
                std::memcpy(&bits, &inf, 4);
                nodes[q] = make_uint2(bits, (uint32_t)q);            // self loop: x > +inf is never true
            } else {
                float t32 = floor_to_f32(T[i]);
                uint32_t bits;
                std::memcpy(&bits, &t32, 4);
                nodes[q] = make_uint2(bits, ((uint32_t)Fe[i] << 24) | (uint32_t)(base + newpos[L[i]]));
            }
        }
        bool stump = (cnt == 3 && L[0] != -1 && L[L[0]] == -1 && L[R[0]] == -1);
        all_stumps = all_stumps && stump;
        if (stump) {
            float t32 = floor_to_f32(T[0]);
            uint32_t bits;
            std::memcpy(&bits, &t32, 4);
            merged[{Fe[0], bits}] += lr * V[L[0]] - lr * V[R[0]];
            right_sum += lr * V[R[0]];
        }
    }
    tree_off[n_trees] = (int)n_nodes;
    ScoreParams& P = m->base;
    P.n_trees = n_trees;
    P.gbt_init = init_raw;
    m->n_trees = n_trees;
    m->n_nodes = (int)n_nodes;
    m->max_depth = max_depth;
    P.max_depth = max_depth;
    m->gbt_guard = 2.4e-7 * sum_max_abs + 1e-9;       // float32 leaf values and float32 partial sums
    if (all_stumps && (int)merged.size() <= MAX_SPLITS) {
        m->tree_mode = 1;
        int s = 0, f_cur = 0;
        bool have_first[F] = {};
        for (int f = 0; f < F; ++f) { P.split1_thr[f] = -INFINITY; P.split1_delta[f] = 0.f; }
        P.split_begin[0] = 0;
        for (auto& kv : merged) {               // std::map iterates sorted by (feature, bits)
            while (f_cur < kv.first.first) P.split_begin[++f_cur] = s;
            float t32;
            std::memcpy(&t32, &kv.first.second, 4);
            if (!have_first[kv.first.first]) {
                have_first[kv.first.first] = true;
                P.split1_thr[kv.first.first] = t32;
                P.split1_delta[kv.first.first] = (float)kv.second;
                continue;
            }
            P.split_thr[s] = t32;
            P.split_delta[s] = (float)kv.second;
            P.split_feat[s] = (unsigned char)kv.first.first;
            ++s;
        }
        P.n_extra_splits = s;
        while (f_cur < F) P.split_begin[++f_cur] = s;
        P.n_splits = (int)merged.size();
        P.gbt_split_base = init_raw + right_sum;
        m->n_splits = (int)merged.size();
    } else {
        m->tree_mode = 2;
    }
    int rc;
    if (m->tree_mode == 2) {
        // ---- 4-byte nodes + per-feature threshold lists for the shared-memory walk (hfe_forest.cuh) -------------------------
        std::map<std::pair<int, uint32_t>, int> thr_id;                  // (feature, ordered threshold key) -> index in the feature's list
        auto key_of = [](uint32_t bits) { return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u); };   // float order
        for (int64_t q = 0; q < n_nodes; ++q) {
            const uint32_t child = nodes[q].y & 0xFFFFFFu;
            if (child == (uint32_t)q) continue;                          // leaf (self loop)
            thr_id[{(int)(nodes[q].y >> 24), key_of(nodes[q].x)}] = 0;
        }
        bool ok = (int)thr_id.size() <= FOREST_MAX_THR;
        ForestParams& Fp = m->forest;
        std::vector<unsigned> nodes4;
        std::vector<int> chunk_tree;
        if (ok) {
            int next = 0, f_cur = 0;
            Fp.thr_begin[0] = 0;
            for (auto& kv : thr_id) {                                    // sorted by (feature, threshold)
                while (f_cur < kv.first.first) Fp.thr_begin[++f_cur] = next;
                const uint32_t key = kv.first.second, bits = (key & 0x80000000u) ? (key & 0x7FFFFFFFu) : ~key;
                std::memcpy(&Fp.thr[next], &bits, 4);
                kv.second = next - Fp.thr_begin[f_cur];                  // index inside the feature's list
                ok = ok && kv.second < 255;
                ++next;
            }
            while (f_cur < F) Fp.thr_begin[++f_cur] = next;
            Fp.n_thr = next;
        }
        if (ok) {
            nodes4.assign((size_t)n_nodes + 8, 0u);
            chunk_tree.push_back(0);
            int64_t chunk_start = 0;
            for (int t = 0; t < n_trees && ok; ++t) {
                const int base = tree_off[t], cnt = tree_off[t + 1] - tree_off[t];
                if (cnt + 4 > FOREST_CHUNK_NODES) { ok = false; break; }
                if ((base + cnt) - (chunk_start & ~(int64_t)3) > FOREST_CHUNK_NODES) {   // copies start 16-byte aligned
                    chunk_tree.push_back(t);
                    chunk_start = base;
                }
                const int64_t origin = chunk_start & ~(int64_t)3;                          // first node of the chunk's copy
                for (int i = 0; i < cnt; ++i) {
                    const int64_t q = base + i;
                    const uint32_t child = nodes[q].y & 0xFFFFFFu;
                    if (child == (uint32_t)q) {
                        nodes4[q] = (255u << 24) | (uint32_t)((q - origin) * 4);          // leaf: the test never holds
                    } else {
                        const uint32_t f = nodes[q].y >> 24;
                        nodes4[q] = ((uint32_t)thr_id[{(int)f, key_of(nodes[q].x)}] << 24) | (f << 16) |
                                    (uint32_t)(((int64_t)child - origin) * 4);
                    }
                }
            }
            chunk_tree.push_back(n_trees);
        }
        if (ok) {
            Fp.n_chunks = (int)chunk_tree.size() - 1;
            if ((rc = upload(m, nodes4, &Fp.nodes4))) return rc;
            if ((rc = upload(m, chunk_tree, &Fp.chunk_tree))) return rc;
        } else {
            Fp.n_chunks = 0;
        }
    }
    if ((rc = upload(m, nodes, &P.nodes))) return rc;
    if ((rc = upload(m, value32, &P.value32))) return rc;
    if ((rc = upload(m, value64, &P.value64))) return rc;
    if ((rc = upload(m, tree_off, &P.tree_off))) return rc;
    if ((rc = upload(m, depth, &P.tree_depth))) return rc;
    if ((rc = upload(m, sk_id, &P.sk_id))) return rc;
    return HFE_OK;
}

// ---- internal feature order --------------------------------------------------------------------------------------------
// The prefix-shared support-vector layout (build_svc) pays off when many vectors agree on their leading features, so the
// model is rebuilt in a feature order that puts the columns with the fewest distinct prefixes first (greedy, one column
// at a time).  Every per-feature table of every member is permuted the same way before the members are built; the
// kernels read caller column ScoreParams::col[f] for internal feature f, so nothing changes at the C boundary.
inline void choose_feature_order(const double* sv, int n_sv, int order[F]) {
    bool used[F] = {};
    std::vector<int> cls(n_sv, 0);
    std::vector<std::pair<std::pair<int, double>, int>> keys(n_sv);
    auto refine = [&](int col, bool assign) -> size_t {
        for (int j = 0; j < n_sv; ++j) keys[j] = {{cls[j], sv[(size_t)j * F + col] + 0.0}, j};
        std::sort(keys.begin(), keys.end());
        size_t distinct = 0;
        for (int k = 0; k < n_sv; ++k) {
            if (k == 0 || keys[k].first != keys[k - 1].first) ++distinct;
            if (assign) cls[keys[k].second] = (int)distinct - 1;
        }
        return distinct;
    };
    for (int pos = 0; pos < F; ++pos) {
        int best = -1;
        size_t best_count = 0;
        for (int c = 0; c < F; ++c) {
            if (used[c]) continue;
            if (pos >= SV_PREFIX) { best = c; break; }               // beyond the prefix the order does not matter
            size_t count = refine(c, false);
            if (best < 0 || count < best_count) { best = c; best_count = count; }
        }
        used[best] = true;
        order[pos] = best;
        if (pos < SV_PREFIX) refine(best, true);
    }
}

// Copies of the per-feature blob sections in internal feature order (ParsedBlob sections otherwise point into the
// caller's blob).  Sections that are missing or mis-sized are left alone: the member builders report them.
inline void permute_sections(ParsedBlob* pb, const int order[F], std::vector<std::vector<uint8_t>>* owned) {
    int internal_of[F];
    for (int f = 0; f < F; ++f) internal_of[order[f]] = f;
    auto own = [&](Section* s, size_t bytes) -> uint8_t* {
        owned->emplace_back(s->data, s->data + bytes);
        s->data = owned->back().data();
        return owned->back().data();
    };
    for (auto& kv : pb->sections) {
        uint32_t kind = kv.first.first;
        Section& s = kv.second;
        if ((kind == K_SVC_MEAN || kind == K_SVC_SCALE || kind == K_SVC_SV) && s.count >= (uint64_t)F && s.count % F == 0) {
            std::vector<double> src(s.f64(), s.f64() + s.count);
            double* dst = reinterpret_cast<double*>(own(&s, s.count * 8));
            for (uint64_t r = 0; r < s.count / F; ++r)
                for (int f = 0; f < F; ++f) dst[r * F + f] = src[r * F + order[f]];
        } else if (kind == K_LIN_PARAMS && s.count == (uint64_t)F + 1) {
            std::vector<double> src(s.f64(), s.f64() + s.count);
            double* dst = reinterpret_cast<double*>(own(&s, s.count * 8));
            for (int f = 0; f < F; ++f) dst[f] = src[order[f]];
        } else if (kind == K_GBT_FEATURE) {
            int32_t* dst = reinterpret_cast<int32_t*>(own(&s, s.count * 4));
            for (uint64_t i = 0; i < s.count; ++i)
                if (dst[i] >= 0 && dst[i] < F) dst[i] = internal_of[dst[i]];
        }
    }
}

// Guard bands: how far the float32 decision values may be off, pushed through the Platt slope and the
// combiner.  Called by build_model and again whenever the creation-time probe (hfe_api.cu: probe_guards) widens svc_guard.
inline void derive_guards(Model* m) {
    ScoreParams& P = m->base;
    double slope = std::fabs(P.prob_a) * 0.25;                       // |d r / d f| <= |A|/4
    P.solver_guard = std::fabs(P.prob_a) * m->svc_guard;             // in units of t = f*A + B
    double w_svc = P.slot_svc >= 0 ? std::fabs(P.final_coef[P.slot_svc]) : 0.0;
    P.label_guard = 2.0 * w_svc * slope * m->svc_guard + 1e-9 +
                    (P.slot_gbt >= 0 ? std::fabs(P.final_coef[P.slot_gbt]) * 0.25 * m->gbt_guard : 0.0);
    if (m->mma.nb > 0 && P.slot_lin >= 0) {
        // tensor-core engine: the logistic member's dot product comes out of a float32 accumulator (|error| <~ 2^-23 of
        // the largest partial sum) instead of a float64 FMA chain
        double mag = std::fabs(m->mma.lin_base);
        for (int d = 0; d < F; ++d) mag += 64.0 * std::fabs(P.lin_coef[d]);
        P.label_guard += std::fabs(P.final_coef[P.slot_lin]) * 0.25 * 1.2e-7 * mag;
    }
    if (P.slot_svc < 0 && m->tree_mode != 2) {
        // stacks scored by the streaming pass: float32 links (|error| <= 1.8e-7 each) and a float32 combiner logit
        double w_sum = 0.0;
        for (int i = 0; i < m->n_members; ++i) w_sum += std::fabs(P.final_coef[i]);
        P.label_guard += w_sum * 1.8e-7 + (std::fabs(P.final_intercept) + w_sum) * 1.2e-7;
    }
}

inline int build_model(const uint8_t* blob, size_t len, int device, Model** out) {
    ParsedBlob pb;
    int rc = parse_blob(blob, len, &pb);
    if (rc) return rc;
    int order[F];
    for (int f = 0; f < F; ++f) order[f] = f;
    std::vector<std::vector<uint8_t>> owned_sections;
    if (pb.n_features == (uint32_t)F) {
        for (uint32_t i = 0; i < pb.n_members && i < (uint32_t)MAX_MEMBERS; ++i) {
            const Section* ps = pb.find(K_SVC_PARAMS, i, 8);
            if (!ps) continue;
            double n_sv_d = ps->f64()[4];
            if (!(n_sv_d >= 1 && n_sv_d <= 65536) || n_sv_d != std::floor(n_sv_d)) break;
            const Section* sv = pb.find(K_SVC_SV, i, (uint64_t)n_sv_d * F);
            if (!sv || !all_finite(sv->f64(), sv->count)) break;
            choose_feature_order(sv->f64(), (int)n_sv_d, order);
            permute_sections(&pb, order, &owned_sections);
            break;
        }
    }
    if (pb.n_features != (uint32_t)F)
        return fail(HFE_ERR_UNSUPPORTED, "this build scores the 17-feature schema only (HFE_N_FEATURES)");
    if (pb.n_members > (uint32_t)MAX_MEMBERS) return fail(HFE_ERR_UNSUPPORTED, "more than 4 stack members");
    const Section* fin = pb.find(K_FINAL, STACK_LEVEL, pb.n_members + 1);
    const Section* cls = pb.find(K_CLASSES, STACK_LEVEL, 2);
    if (!fin || !cls || !all_finite(fin->f64(), pb.n_members + 1))
        return fail(HFE_ERR_BAD_BLOB, "blob: combiner section missing or non-finite");
    if (device >= 0) HFE_CUDA_TRY(cudaSetDevice(device));
    Model* m = new Model();
    m->device = device;
    m->n_members = (int)pb.n_members;
    ScoreParams& P = m->base;
    std::memset(&P, 0, sizeof(P));
    P.slot_svc = P.slot_gbt = P.slot_lin = -1;
    P.n_members = m->n_members;
    P.perm_col = -1;
    for (int f = 0; f < F; ++f) {
        m->col_of[f] = order[f];
        m->internal_of[order[f]] = f;
    }
    for (uint32_t i = 0; i < pb.n_members && rc == HFE_OK; ++i) {
        P.final_coef[i] = fin->f64()[i];
        if (pb.has(K_SVC_PARAMS, i)) {
            if (P.slot_svc >= 0) rc = fail(HFE_ERR_UNSUPPORTED, "more than one SVC member");
            else { P.slot_svc = (int)i; m->kinds[i] = KIND_SVC; rc = build_svc(m, pb, i); }
        } else if (pb.has(K_GBT_PARAMS, i)) {
            if (P.slot_gbt >= 0) rc = fail(HFE_ERR_UNSUPPORTED, "more than one boosting member");
            else { P.slot_gbt = (int)i; m->kinds[i] = KIND_GBT; rc = build_gbt(m, pb, i); }
        } else if (pb.has(K_LIN_PARAMS, i)) {
            const Section* lp = pb.find(K_LIN_PARAMS, i, F + 1);
            if (P.slot_lin >= 0) rc = fail(HFE_ERR_UNSUPPORTED, "more than one logistic member");
            else if (!lp || !all_finite(lp->f64(), F + 1)) rc = fail(HFE_ERR_BAD_BLOB, "blob: logistic params bad");
            else {
                P.slot_lin = (int)i;
                m->kinds[i] = KIND_LIN;
                for (int d = 0; d < F; ++d) P.lin_coef[d] = lp->f64()[d];
                P.lin_intercept = lp->f64()[F];
            }
        } else {
            rc = fail(HFE_ERR_BAD_BLOB, "blob: member without parameter section");
        }
    }
    if (rc == HFE_OK) {
        if (m->mma.nb > 0 && !m->mma_img_host.empty()) {
            // logistic member on the tensor core: its coefficients (three bfloat16 parts each, like the W columns) are the
            // last column of the B operand; that column's RBF coefficient is zero
            MmaParams& Q = m->mma;
            Q.lin_base = P.lin_intercept;
            Q.extra_split_mask = 0;
            if (m->tree_mode == 1)
                for (int d = 0; d < F; ++d)
                    if (P.split_begin[d + 1] != P.split_begin[d]) Q.extra_split_mask |= 1u << d;
            const int col = Q.nb * Q.block_sv - 1;
            uint16_t* row = reinterpret_cast<uint16_t*>(m->mma_img_host.data() + (size_t)(col / 8) * (MMA_B_CHUNKS * 128) + (col % 8) * 16);
            for (int d = 0; d < F; ++d) {
                Q.lin_base += (double)Q.center[d] * P.lin_coef[d];
                double rest = P.slot_lin >= 0 ? P.lin_coef[d] : 0.0;
                for (int part = 0; part < 3; ++part) {
                    float v = (float)rest;
                    uint32_t b;
                    std::memcpy(&b, &v, 4);
                    b += 0x7FFFu + ((b >> 16) & 1u);
                    const uint16_t h = (uint16_t)(b >> 16);
                    const uint32_t hb = (uint32_t)h << 16;
                    float hv;
                    std::memcpy(&hv, &hb, 4);
                    rest -= (double)hv;
                    row[(6 - 3 * part + d / 8) * 64 + d % 8] = h;
                }
            }
            rc = upload(m, m->mma_img_host, &Q.w_img);
            m->mma_img_host.clear();
            m->mma_img_host.shrink_to_fit();
        }
        P.final_intercept = fin->f64()[pb.n_members];
        if (P.slot_svc >= 0) {
            std::vector<double> edges = find_solver_edges();
            if (edges.size() > (size_t)MAX_EDGES) rc = fail(HFE_ERR_UNSUPPORTED, "coupling solver has too many edges");
            else {
                P.n_edges = (int)edges.size();
                for (size_t e = 0; e < edges.size(); ++e) P.edges[e] = edges[e];
                const CouplingTable& tab = coupling_table(edges);
                if (!tab.ok) rc = fail(HFE_ERR_UNSUPPORTED, "coupling table: fit error above 1e-9 or two solver edges in one cell");
                if (rc == HFE_OK) rc = upload(m, tab.coef, &P.coupling_coef);
                if (rc == HFE_OK) rc = upload(m, tab.edge, &P.coupling_edge);
            }
        }
        derive_guards(m);
    }
    if (rc != HFE_OK) {
        for (void* p : m->allocations) cudaFree(p);
        delete m;
        return rc;
    }
    *out = m;
    return HFE_OK;
}

}  // namespace hfe
