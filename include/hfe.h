/*
 * hfe.h -- C ABI of the B200 (sm_100a) scorer for the heart-failure-progression stack.
 *
 * The reference has no FFI boundary of its own: predict_hf.py:33-36 and train_ensemble_public.py:61-63 call the
 * scikit-learn estimator protocol (predict_proba / predict / decision_function / transform / apply) on an unpickled
 * StackingClassifier.  Each entry point below names the reference call (file:line) or scikit-learn method
 * (SK/ = site-packages/sklearn, version 1.9.0) whose arithmetic it replaces.  The Python facade
 * (machine-learning-replications_b200/facade.py) binds these with ctypes and presents the estimator protocol again.
 *
 * Conventions
 *   - every function returns an hfe_status (0 = HFE_OK); nothing throws across the boundary;
 *     hfe_last_error_string() gives the text for the calling thread's last failure.
 *   - pointers named x, out, labels, ... in the *device* entry points are CUDA device pointers on the model's
 *     device; `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Device entry points
 *     enqueue work on `stream` and return: no device memory is allocated and nothing is waited for.  Host-side set-up
 *     that is not free: the first scoring call per (device, kernel variant) takes a mutex, opts the kernel in to large
 *     dynamic shared memory (cudaFuncSetAttribute) and asks the occupancy calculator how many blocks stay resident;
 *     every call to a stack without the SVC member encodes a tensor map of the caller's matrix on the host
 *     (cuTensorMapEncodeTiled, ~1 us).  Entry points that take a model handle make the model's device current
 *     (cudaSetDevice) and leave it current; those without a handle (ingest, synth_rows, knn1_impute, binary_metrics)
 *     run on whatever device is current, which must be the one that owns the pointers and the stream.
 *   - feature matrices are COLUMN-major float32: feature f of row i lives at x[f*ld + i], ld >= n, ld % 4 == 0 and x
 *     16-byte aligned so that a thread can fetch four consecutive rows of one column with one 128-bit load.
 *   - `status` is a device array of HFE_STATUS_WORDS uint32 (256 KiB) owned by the caller; the call zeroes its header
 *     and then fills it:
 *       [0] error bits: HFE_ROW_NONFINITE if any input value was NaN/Inf (scikit-learn raises ValueError there,
 *           SK/svm/_base.py:652-660); HFE_WORKLIST_OVERFLOW if more than HFE_WORKLIST_CAPACITY rows needed the float64
 *           re-evaluation (then the rows beyond the capacity keep their float32-path result: rerun the batch through
 *           hfe_predict_proba_f64, as the Python facade does);
 *       [1] number of rows that were re-evaluated in float64 ("guard bands");
 *       [2] work-list length, [3] reserved (the checked build counts completed appends there), [4..] the work list
 *           itself (row indices).  The work list is scratch
 *           memory of the call: the fused kernel appends the rows whose result hinges on a branch (coupling-solver
 *           exit test, label sign), a second small kernel re-evaluates exactly those rows in float64.
 *     Two calls that overlap on the device must not share a status array.  n must be < 2^32 rows per call.
 *   - a model handle is immutable after creation and may be shared by host threads and streams.
 */
#ifndef HFE_H_
#define HFE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum hfe_status {
    HFE_OK = 0,
    HFE_ERR_INVALID_ARGUMENT = 1,  /* null pointer, negative size, misaligned matrix */
    HFE_ERR_BAD_BLOB = 2,          /* malformed / corrupted model blob (magic, sizes, checksum, tree shape, ...) */
    HFE_ERR_UNSUPPORTED = 3,       /* well-formed model this build does not implement (feature count, member mix) */
    HFE_ERR_CUDA = 4,              /* a CUDA runtime call failed; see hfe_last_error_string() */
    HFE_ERR_NONFINITE_INPUT = 5,   /* host entry points: X holds NaN/Inf (what scikit-learn reports as ValueError) */
    HFE_ERR_TOO_LARGE = 6          /* request exceeds a documented fixed limit */
} hfe_status;

#define HFE_N_FEATURES 17          /* load_data_public.py schema after feature selection; predict_hf.py:6-26 */
#define HFE_STATUS_HEADER_WORDS 4
#define HFE_WORKLIST_CAPACITY 65532   /* rows one call can hand to the float64 re-evaluation */
#define HFE_STATUS_WORDS (HFE_STATUS_HEADER_WORDS + HFE_WORKLIST_CAPACITY)   /* 65536 uint32 = 256 KiB */
#define HFE_STATUS_LIST_LEN 2
#define HFE_ROW_NONFINITE 1u
#define HFE_WORKLIST_OVERFLOW 2u

typedef struct hfe_model hfe_model;

typedef struct hfe_model_info {
    int32_t n_features;
    int32_t n_members;             /* meta-feature columns, in estimators_ order */
    int32_t member_kind[4];        /* 1 = scaler+RBF-SVC, 2 = boosted trees, 3 = logistic regression, 0 = unused */
    int32_t n_support_vectors;
    int32_t n_trees;
    int32_t n_tree_nodes;
    int32_t max_tree_depth;
    int32_t tree_mode;             /* 0 none, 1 depth-1 stumps merged into a split table, 2 general node tables */
    int32_t n_merged_splits;
    int32_t device;
    int32_t rbf_fma_per_row;       /* float32 lane-FMAs the fused kernel executes per row in its RBF loop (table steps x 4
                                    * lanes); below n_support_vectors * 18 when the prefix-shared layout is in use */
    double  svc_guard;             /* |decision value| uncertainty assumed for the float32 RBF sum */
    uint64_t device_bytes;         /* bytes of device memory the handle owns */
    int32_t rbf_engine;            /* what scores the RBF sum: 0 no SVC member, 1 SIMT FFMA2 loop, 2 tensor cores (tcgen05) */
    int32_t rbf_exp_per_row;       /* exponentials the active engine evaluates per row (tensor cores: padded column count) */
    /* the creation-time guard probe (hfe_model_create): the band in use, and what the probe batch showed */
    double  svc_observed_err;      /* max |fast - float64| of the SVC decision value over the probe rows the fast engines
                                    * (SIMT and, where usable, tensor cores) did not hand to the float64 re-evaluation */
    double  label_guard;           /* |combiner decision value| below this => the row is re-evaluated in float64 */
    double  label_observed_err;    /* max |fast - float64| of the combiner's decision value over the same rows */
    double  svc_observed_by_engine[2];   /* the same per engine: [0] SIMT loop, [1] tensor cores (0 where not usable) */
    int32_t probe_rows;            /* rows in the probe batch (0: no SVC member, or hfe_model_plan) */
    int32_t guard_widened;         /* bit 0: svc_guard was below 8 x svc_observed_err and was raised to it; bit 1: the same
                                    * for label_guard */
} hfe_model_info;

/* Replaces pickle.load('./hf_predict_model.pkl') (predict_hf.py:33-34): validates the flat blob written by
 * export/blob.py (header, CRC-32, section shapes, tree topology, the n_support consistency check of
 * SK/svm/_base.py:680-686), derives the device tables and uploads them to `device`, then checks the float32 guard bands
 * against this model (hfe_model_calibrate_guards(model, 1): 65 536 probe rows, ~10 ms).  Allocates and synchronises; call
 * once. */
int hfe_model_create(const void* blob, size_t blob_bytes, int device, hfe_model** out_model);
void hfe_model_destroy(hfe_model* model);
int hfe_model_get_info(const hfe_model* model, hfe_model_info* out_info);
/* Re-runs the guard probe hfe_model_create ends with.  The bands are first derived from the model's coefficients by a formula
 * (DESIGN.md "error budget"), multiplied by `formula_scale` (1 at creation); then 65 536 probe rows -- the support vectors in
 * raw units and crossovers of pairs of them -- go through every fast engine and through the all-float64 kernel, and a band
 * below 8 x the observed |fast - float64| is raised to that.  formula_scale > 1 buys wider bands (more rows re-evaluated in
 * float64); < 1 exists for the test that shows the probe catching a band that is too narrow.  Allocates and synchronises;
 * not to be called while other threads score with the handle. */
int hfe_model_calibrate_guards(hfe_model* model, double formula_scale);
/* Which engine evaluates the RBF sum of the SVC member (svm.cpp:461-515, 2821-2904): 0 = automatic (the tensor-core kernel
 * when the model fits it: at most ~460 support vectors -- its operand tables share one SM's shared memory -- and no
 * general tree tables), 1 = the SIMT FFMA2 loop, 2 = the
 * tensor-core kernel or HFE_ERR_UNSUPPORTED at scoring time.  Both meet the same parity bars; the switch exists so that
 * the two can be measured and tested side by side.  Set it before the handle is shared between threads. */
int hfe_model_set_rbf_engine(hfe_model* model, int engine);

/* How general (depth > 1) boosted trees are walked (SK/ensemble/_gradient_boosting.pyx:25-73): 0 = automatic (4-byte nodes
 * streamed through shared memory when the stack has no SVC member and the forest has at most 256 distinct (feature,
 * threshold) pairs, else 8-byte nodes through L1), 1 = always through L1, 2 = shared memory or HFE_ERR_UNSUPPORTED at
 * scoring time, 3 = through L1 with the node table pinned in an L2 persisting window (sets and clears the stream's
 * access-policy window around each launch; kept for measurements).  Same results whichever is used. */
int hfe_model_set_tree_engine(hfe_model* model, int engine);

/* Host-only twin of hfe_model_create: validates the blob and derives every table exactly as hfe_model_create does, but
 * touches no device (nothing is uploaded; device_bytes reports what would be).  feature_order (17 entries, may be
 * NULL) receives the caller's column of each internal feature -- the order the prefix-shared support-vector table was
 * built for.  For tooling and for tests on machines without a GPU. */
int hfe_model_plan(const void* blob, size_t blob_bytes, hfe_model_info* out_info, int32_t* feature_order);

/* StackingClassifier.predict_proba (SK/ensemble/_stacking.py:768-794) as called at predict_hf.py:36 and
 * train_ensemble_public.py:62.  out_n2[2*i] = P(class 0) = 1 - p1, out_n2[2*i+1] = p1, float32.
 * One fused kernel: scaler -> RBF-SVC decision -> Platt -> clamp -> 2-class coupling (svm.cpp:2921-2958, 2046-2107),
 * boosted trees (_gradient_boosting.pyx:25-73) -> expit, logistic member (_base.py:366-451), combiner.  Stacks without
 * the SVC member are bandwidth-bound and take a streaming kernel instead (tiles staged in shared memory by the copy
 * engine); either way one scoring launch plus the small float64 fix-up launch over the work list. */
int hfe_predict_proba(const hfe_model* model, const float* x, int64_t n, int64_t ld, float* out_n2,
                      uint32_t* status, void* stream);

/* StackingClassifier.predict (SK/ensemble/_stacking.py:711-766 -> SK/linear_model/_base.py:398-425): labels[i] = 1
 * iff the combiner's decision value is strictly > 0 (index into classes_).  out_n2 may be NULL. */
int hfe_predict(const hfe_model* model, const float* x, int64_t n, int64_t ld, uint8_t* labels, float* out_n2,
                uint32_t* status, void* stream);

/* Everything at once; any output pointer may be NULL.  decision[i] is the combiner's decision value
 * (StackingClassifier.decision_function), meta[i*n_members + m] the P(class 1) of member m
 * (StackingClassifier.transform, SK/ensemble/_stacking.py:814-830, 87-135), both stored as float64.  Their accuracy is
 * that of the pass that produced them: with the SVC member the links and the combiner run in float64 and only the RBF
 * sum is float32 (|error| of decision <~ 1e-6); stacks without the SVC member take the streaming pass, whose links and
 * combiner run in float32 (~1.8e-7 per link, ~5e-7 on decision) and are widened on store.  Rows inside a guard band
 * are re-evaluated in float64 either way; hfe_score_f64 gives float64 values for every row.
 * perm_col >= 0 scores the matrix with column perm_col read through perm_idx (int32 row indices, length n): one pass
 * of permutation importance (SK/inspection/_permutation_importance.py:31-76) without materialising the shuffled
 * matrix.  If y (uint8 class indices) and score are given, score[0] += #{labels == y} (uint64) is accumulated. */
int hfe_score(const hfe_model* model, const float* x, int64_t n, int64_t ld, int32_t perm_col,
              const int32_t* perm_idx, float* out_n2, uint8_t* labels, double* decision, double* meta,
              const uint8_t* y, unsigned long long* score, uint32_t* status, void* stream);

/* Permutation importance, one column (SK/inspection/_permutation_importance.py:31-76): for each selected index vector
 * perm_idx + sel[k] * perm_stride (int32 row indices, length n) the matrix is scored with column perm_col read through it
 * and scores[k] += #{labels == y} -- the hfe_score passes of that column enqueued back to back without returning to the
 * caller in between (perm_col < 0: the baseline pass, sel only counts).  flags[k] (may be NULL) receives that pass's
 * status[0] error bits; `status` itself is reused by every pass.  sel is a HOST array, everything else device memory. */
int hfe_score_permuted(const hfe_model* model, const float* x, int64_t n, int64_t ld, int32_t perm_col,
                       const int32_t* perm_idx, int64_t perm_stride, int32_t n_sel, const int32_t* sel,
                       const uint8_t* y, unsigned long long* scores, uint32_t* flags, uint32_t* status, void* stream);

/* Permutation importance with scikit-learn's own index vectors, shuffles and scoring overlapped
 * (SK/inspection/_permutation_importance.py:31-76, 278-313): job k scores the matrix with column job_col[k] read through
 * the job_rep[k]-th index vector (job_col[k] < 0: unpermuted, the baseline) and scores[k] += #{labels == y}; flags[k] (may be
 * NULL) receives that pass's status[0].  The vectors are what hfe_perm_indices(seed, n, .) returns; they are made on host
 * threads while earlier passes run on the device -- vector r is written to perm_host + r*n (HOST int32 [n_repeats][n], kept
 * for the caller), copied to perm_dev + r*n (DEVICE int32 [n_repeats][n]) and used as soon as it is complete -- so the
 * jobs must be listed with ascending job_rep.  job_col / job_rep are HOST arrays.  Returns when everything is enqueued
 * and the last vector is made. */
int hfe_permutation_passes(const hfe_model* model, const float* x, int64_t n, int64_t ld, const uint8_t* y, uint32_t seed,
                           int32_t n_repeats, int32_t n_jobs, const int32_t* job_col, const int32_t* job_rep,
                           int32_t* perm_host, int32_t* perm_dev, unsigned long long* scores, uint32_t* flags,
                           uint32_t* status, void* stream);

/* The index vectors scikit-learn's permutation_importance uses for one column (same file, lines 66-74, 278): a
 * RandomState(seed) -- numpy's legacy MT19937 stream -- shuffles one index vector n_repeats times in a row and applies it
 * each time to the already permuted column; out[r*n + i] is the row whose value lands in row i at repeat r.  Host
 * memory only, no device work; bit-for-bit numpy's RandomState.shuffle. */
int hfe_perm_indices(uint32_t seed, int64_t n, int32_t n_repeats, int32_t* out);

/* GradientBoostingClassifier.apply (SK/ensemble/_gb.py:1104-1135; SK/tree/_tree.pyx:954-996): scikit-learn node id of
 * the leaf each row reaches in each tree, leaves[i*n_trees + t], uint16 (parity mode; not on the timed path). */
int hfe_apply(const hfe_model* model, const float* x, int64_t n, int64_t ld, uint16_t* leaves, void* stream);

/* Same as evaluating every row through the float64 re-evaluation path (test / calibration aid). */
int hfe_predict_proba_f64(const hfe_model* model, const float* x, int64_t n, int64_t ld, double* out_n2,
                          uint32_t* status, void* stream);

/* hfe_score's outputs with every row on the float64 path (no float32 RBF sum, no guard bands, no work list): labels,
 * decision values and member probabilities as an all-float64 evaluation gives them.  The parity harness compares the
 * fused pass with this on every row (labels must be identical); any output pointer may be NULL. */
int hfe_score_f64(const hfe_model* model, const float* x, int64_t n, int64_t ld, double* out_n2, uint8_t* labels,
                  double* decision, double* meta, uint32_t* status, void* stream);

/* Synthetic cohort rows on the device (the reference's .mat cohorts are unpublished): column c of row (row0+i) is
 * base[c] + step[c] * #{k : hash32(seed,row,col_id[c]) >= thr[k]}, k in [thr_off[c], thr_off[c+1]).  All table
 * pointers are DEVICE pointers.  Bit-identical to synth.host_rows(). */
int hfe_synth_rows(uint64_t seed, int64_t row0, int64_t n, int32_t n_cols, const int32_t* col_id, const float* base,
                   const float* step, const int32_t* thr_off, const uint32_t* thr, float* x, int64_t ld,
                   void* stream);

/* Row-major float64 [n, n_features] (what load_data_public.py:11-14 yields and predict_hf.py:36 passes) ->
 * column-major float32 with round-to-nearest, the cast SK/ensemble/_gb.py:1579-1604 applies; sets HFE_ROW_NONFINITE. */
int hfe_ingest_rowmajor_f64(const double* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                            uint32_t* status, void* stream);
/* Same for a row-major float32 matrix (a cohort already held in single precision): transpose only. */
int hfe_ingest_rowmajor_f32(const float* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                            uint32_t* status, void* stream);

/* Same for the int64 matrix predict_hf.py:29-36 builds (a Python list of ints through np.array): each value goes
 * through float64 to float32, as scikit-learn's validate_data followed by the float32 cast does. */
int hfe_ingest_rowmajor_i64(const int64_t* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                            uint32_t* status, void* stream);
/* Same for a cohort stored as unsigned bytes.  No counterpart in the reference: every variable of the
 * load_data_public.py schema is a small non-negative integer (13 yes/no flags, NYHA class, wall thickness in mm,
 * SAM / regurgitation grade, ejection fraction in %), so a large cohort can be kept and shipped to the GPU at 17 bytes
 * per patient instead of 68 -- the host-to-device copy is what bounds end-to-end throughput. */
int hfe_ingest_rowmajor_u8(const uint8_t* x_rowmajor, int64_t n, int32_t n_features, float* x, int64_t ld,
                           uint32_t* status, void* stream);

/* KNNImputer(missing_values=nan, n_neighbors=1).transform (train_ensemble_public.py:37-40; SK/impute/_knn.py,
 * SK/metrics/pairwise.py nan_euclidean_distances) on the candidate matrix, before feature selection: x is a row-major
 * float64 [n, n_features] device matrix imputed in place from the row-major float64 fit matrix fit_x [n_fit,
 * n_features] (NaN = missing in both).  col_mean[f] = mean of the present values of fit column f, NaN for a column that
 * is all-missing in fit_x (such columns are left alone; the caller drops them as scikit-learn does).  Among equidistant
 * donors -- which scikit-learn leaves to np.argpartition, i.e. undefined -- the lowest fit row wins.  At most ~25 000
 * fit rows (the distances of one receiver row live in one block's shared memory). */
int hfe_knn1_impute(const double* fit_x, int64_t n_fit, int32_t n_features, const double* col_mean, double* x,
                    int64_t n, void* stream);

/* The evaluation train_ensemble_public.py:63-64,67,79 performs next on the probabilities (classification_report of
 * ``y_pred[:,1] > 0.5``, ROC and precision-recall curves): confusion[0..3] = tn, fp, fn, tp at `threshold` (uint64,
 * zeroed by the call); hist (may be NULL) = 2 x HFE_METRIC_BINS uint32, per true class the histogram of P(class 1) in
 * bins floor(p*(BINS-1)+0.5), from which the curves and their areas follow on the host. */
#define HFE_METRIC_BINS 65536
int hfe_binary_metrics(const float* proba_n2, const uint8_t* y, int64_t n, float threshold,
                       unsigned long long* confusion, uint32_t* hist, void* stream);

/* Host-buffer convenience used by the facade for numpy inputs: copies X (row-major float64, host) to the device,
 * ingests, scores, copies P back (float64 [n,2]; p0 = 1 - p1) and synchronises.  From 65 536 rows on (train_ensemble_public.py:62
 * on a big cohort) the batch goes through a chunk pipeline the handle owns and builds on first use -- three pinned /
 * device slots of 262 144 rows with a stream each and a pool of host threads (one per CPU the process may run on, at most
 * 16; environment HFE_HOST_THREADS): the threads narrow the caller's float64 rows into a pinned chunk (float32, or bytes
 * while every value is an integer in [0, 255] -- lossless), the slot's stream copies, ingests, scores and copies back, the
 * threads widen the probabilities into the caller's array; one such call at a time per handle.  Up to 4096 rows (predict_hf.py:36 scores one patient) skip the copies altogether: the rows are placed
 * in pinned memory mapped into the device address space (a block owned by the handle, one caller at a time) and one
 * kernel -- a warp per row, on the float64 path -- evaluates them and writes the results back.
 * labels / decision / meta may be NULL.  Returns HFE_ERR_NONFINITE_INPUT when X holds NaN/Inf. */
int hfe_predict_host(const hfe_model* model, const double* x_rowmajor, int64_t n, double* proba_n2, uint8_t* labels,
                     double* decision, double* meta);

/* No counterpart in the reference (added for embedders; nothing in predict_hf.py / scikit-learn corresponds to it).
 *
 * hfe_predict_proba for callers that cannot obtain device memory of their own.  The ordinary entry points need a
 * caller-supplied 256 KiB `status` array in device memory (error bits + the work list of rows that get re-evaluated
 * in float64).  A caller that lives inside someone else's process -- a scoring plug-in in a database or an inference
 * server that has already handed every byte of device memory to its own pool, or a service whose policy forbids any
 * allocation once it is serving -- receives x and out_n2 from its host but has nowhere to take that array from and
 * may not call cudaMalloc.  This entry point keeps the array inside the library instead: a fixed 256 KiB block of
 * device memory that is part of the loaded library image on every GPU that loads it, for the life of the process.
 * It calls no cudaMalloc / cudaFree and no allocating library routine.
 *
 *   - n may be at most HFE_NOALLOC_MAX_ROWS; larger batches are refused with HFE_ERR_TOO_LARGE (split them).
 *   - flags (device pointer to one uint32, may be NULL) receives the error bits the `status[0]` word would hold:
 *     HFE_ROW_NONFINITE (those rows' outputs are NaN) and HFE_WORKLIST_OVERFLOW, which means what it means for
 *     hfe_predict_proba -- more than HFE_WORKLIST_CAPACITY rows needed the float64 re-evaluation, the rows beyond the
 *     capacity kept their float32-path result, and the batch should be rerun through hfe_predict_proba_f64 (which
 *     needs a caller-owned status array) or in smaller pieces.  A caller that passes NULL never learns of either.
 *   - Every call asks the runtime for the block's device address (cudaGetSymbolAddress) with the model's device
 *     current, so the address is never stale after a device reset in the embedding process; the first such lookup on
 *     a device loads the library's module there, so make that first call outside any stream capture.  Apart from the
 *     lookup, calls only enqueue work on `stream`.
 *   - All calls on one GPU share the block.  Each call clears the counters it needs before use, so results never depend
 *     on an earlier call, but two calls whose kernels overlap on the GPU (different streams, or different host
 *     threads) would race on it: keep calls to this entry point on one stream per device, or serialise them.
 */
#define HFE_NOALLOC_MAX_ROWS 67108864   /* 2^26 rows per call */
int hfe_predict_proba_noalloc(const hfe_model* model, const float* x, int64_t n, int64_t ld, float* out_n2,
                              uint32_t* flags, void* stream);

const char* hfe_last_error_string(void);
int hfe_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* HFE_H_ */
