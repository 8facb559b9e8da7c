/*
 * fdb200.h - C ABI of libfdb200.so, the B200 (sm_100a) implementation of the
 * fraud-detection batch scoring + SHAP hot path.
 *
 * The reference (wtfashwin/fraud-detection) is pure Python: its "plugin interface" for
 * this path is duck-typed sklearn/shap objects held in module globals
 * (SCALER.transform, MODEL.predict/predict_proba, EXPL.shap_values).  The Python
 * classes in fraud_detection_b200/ mirror that surface and bind the entry points below
 * through ctypes; each declaration cites the reference call it replaces.  Paths are
 * relative to the reference repository; "SK:" is scikit-learn's source tree.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross this boundary;
 *   - every function returns FDB_OK (0) or a negative FDB_ERR_* code and never aborts;
 *     fdb_last_error() gives the message for the calling thread;
 *   - "_dev" entry points take DEVICE pointers and a cudaStream_t passed as void*
 *     (NULL = legacy default stream); they enqueue work and return without synchronising;
 *   - "_host" entry points take HOST pointers (pinned or pageable), run the
 *     H2D -> kernel -> D2H pipeline themselves and return when the outputs are complete
 *     (up to 16,384 rows without any copy command: the kernel works on a mapped, page-locked
 *     staging buffer; FDB_ZERO_COPY_ROWS=0 in the environment restores the copies);
 *   - matrices are row-major, 30 features per row (Time, V1..V28, Amount);
 *   - dtype: FDB_F32 or FDB_F64 selects the element type of every x / out / phi / proba
 *     buffer of that call (parameters are always passed as host doubles).
 *   - there is no CPU fallback: without a CUDA device every compute call fails with
 *     FDB_ERR_CUDA.
 */
#ifndef FDB200_H
#define FDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDB_N_FEATURES 30

enum { FDB_F32 = 0, FDB_F64 = 1 };

enum {
    FDB_OK = 0,
    FDB_ERR_ARG = -1,     /* bad argument (null, negative size, unknown dtype) -> ValueError */
    FDB_ERR_ALIGN = -2,   /* device buffer not 16-byte aligned                  -> ValueError */
    FDB_ERR_CUDA = -3,    /* CUDA runtime error / no device                     -> RuntimeError */
    FDB_ERR_NOMEM = -4,   /* allocation failed                                  -> MemoryError */
    FDB_ERR_STATE = -5    /* plan/context used in the wrong state               -> RuntimeError */
};

typedef struct fdb_ctx fdb_ctx;   /* one per (process, device); owns streams and workspaces */

int fdb_version(void);
const char* fdb_last_error(void);

/* Context.  Create lazily, after fork (api/app.py runs under gunicorn prefork workers and
 * Celery prefork, SURVEY 8(b) "Threading / processes"). */
int fdb_ctx_create(int device, fdb_ctx** out);
int fdb_ctx_destroy(fdb_ctx* ctx);
int fdb_ctx_device(const fdb_ctx* ctx);
int fdb_ctx_sm_count(const fdb_ctx* ctx);

/* Pinned host memory helpers for callers that want DMA-able buffers. */
int fdb_pinned_alloc(size_t bytes, void** out);
int fdb_pinned_free(void* p);

/* ---------------------------------------------------------------------------------
 * K1a  StandardScaler.transform:  out = (x - mean) / scale
 * Replaces SCALER.transform(X) at api/app.py:194, explain_model.py:19,
 * predict_single.py:25, deploy.py:33 (body SK:preprocessing/_data.py:1092-1135).
 * ------------------------------------------------------------------------------- */
int fdb_transform_dev(fdb_ctx* ctx, void* stream, int dtype, int64_t n_rows, const void* x,
                      const double* mean, const double* scale, void* out);
int fdb_transform_host(fdb_ctx* ctx, int dtype, int64_t n_rows, const void* x,
                       const double* mean, const double* scale, void* out);

/* ---------------------------------------------------------------------------------
 * K1b  fused linear pass over rows (one read of x, every requested output in one pass):
 *        phi[r][k] = (x[r][k] - sub[k]) * mul[k]                 exact linear SHAP
 *        logit[r]  = bias + sum_k phi[r][k]
 *        proba[r]  = 1 / (1 + exp(-logit[r]))
 *        cls[r]    = logit[r] > 0
 *        abs_sum[k] = sum_r |phi[r][k]|  (k < 30), abs_sum[30] = n_rows, abs_sum[31] = 0
 * Any of phi / logit / proba / cls / abs_sum may be NULL (not written).
 * With sub = 0, mul = coef_, bias = intercept_ this is
 *   decision_function / predict_proba / predict     (api/app.py:209,223; xai_tasks.py:97;
 *   evaluate_model.py:26-27; SK:linear_model/_base.py:366-440);
 * with sub = background mean, mul = coef_ it is LinearExplainer.shap_values
 *   (explain_model.py:27, api/worker.py:75) and coef*x of xai_tasks.py:106-107;
 * with the scaler folded in (sub = mean_ + scale_*mu_bg, mul = coef_/scale_,
 * bias = expected_value) it scores and explains RAW rows in one pass.
 * abs_sum is the device-side input of the global mean|SHAP| summary
 * (explain_model.py:39) that ranks all-reduce across GPUs.
 * proba_cols = 1 writes proba[r] = p; proba_cols = 2 writes the predict_proba layout
 * proba[r][0] = 1 - p, proba[r][1] = p.  cls is uint8; abs_sum is 32 doubles on the device
 * (dev) / host (host).  A _dev call that asks for abs_sum uses a per-context scratch set of
 * its own (the _host entry points and every Kernel SHAP plan have theirs, so those may run
 * at the same time): at most one such _dev launch per context may be in flight.  A
 * non-finite abs_sum[k] means column k of the input held a NaN or infinity.
 * ------------------------------------------------------------------------------- */
int fdb_linear_pass_dev(fdb_ctx* ctx, void* stream, int dtype, int64_t n_rows, const void* x,
                        const double* sub, const double* mul, double bias,
                        void* proba, int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum);
int fdb_linear_pass_host(fdb_ctx* ctx, int dtype, int64_t n_rows, const void* x,
                         const double* sub, const double* mul, double bias,
                         void* proba, int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum);

/* ---------------------------------------------------------------------------------
 * Kernel SHAP (K2 evaluator + K3 solve).  Replaces
 *   shap.KernelExplainer(f, background).shap_values(X, nsamples=..., l1_reg=False)
 * with f = model.predict_proba(.)[:, 1] of the logistic model (BASELINE.json configs 3-4;
 * shap==0.49.1 is pinned at requirements.txt:21, the reference holds no call site).
 *
 * A plan fixes the model (coef, intercept in the space of the rows being explained), the
 * background (n_bg x 30, host doubles, at most 1024 rows; at most 800 when some |log2e * coef * bg| reaches 16384 - the
 * bf16 operand format - and the plan has several coalition sets: creation fails with FDB_ERR_ARG beyond that) and one or more coalition mask
 * sets: bits[s][m] has bit k set when feature k takes the explained row's value, kw[s][m]
 * is the Shapley-kernel weight (maskMatrix / kernelWeights of the reference run).  With
 * n_sets == 1 all rows share the set; otherwise row r uses set set_idx[r] (or r mod n_sets
 * when set_idx is NULL), which reproduces the reference's per-row sampling.  set_idx lives on the
 * device (explain_dev) and is not checked: every entry must lie in [0, n_sets).
 * link: 0 identity, 1 logit (the evaluators then also sum the complements 1 - f, so that
 * logit(ey) stays exact to float32 for expectations arbitrarily close to 0 or 1).
 * evaluator: FDB_KSHAP_EVAL_TC (tcgen05 tensor cores, any plan),
 * FDB_KSHAP_EVAL_FACTORED (single-set plans only: background part of every coalition is
 * precomputed, one exponential per (row, coalition)), FDB_KSHAP_EVAL_SIMT (plain fp32 CUDA
 * cores; the cross-check) or FDB_KSHAP_EVAL_AUTO (FACTORED for single-set plans, else TC).
 * explain: x is n_rows x 30 float32; phi is n_rows x 30 float32 with
 * sum_k phi[r][k] = link(f(x_r)) - expected_value exactly by construction; ey_out
 * (optional, n_rows x n_masks) returns the coalition expectations; abs_sum as in K1b.
 * A singular X'WX (too few or dependent coalitions: a pivot that lost 11 digits against its diagonal
 * entry) makes plan creation fail with FDB_ERR_ARG.
 * A plan owns its scratch (coalition expectations of the batch in flight): explain calls on the
 * SAME plan must not overlap on the GPU; use one plan per stream for concurrent work.  Memory per
 * coalition set: 8 bytes per coalition + 3.8 KB (a 64-byte-per-coalition operand image is kept for
 * single-set plans only).
 * ------------------------------------------------------------------------------- */
typedef struct fdb_kshap_plan fdb_kshap_plan;
enum { FDB_KSHAP_EVAL_AUTO = -1, FDB_KSHAP_EVAL_TC = 0, FDB_KSHAP_EVAL_SIMT = 1, FDB_KSHAP_EVAL_FACTORED = 2 };
int fdb_kshap_plan_create(fdb_ctx* ctx, const double* coef, double intercept, int n_bg, const double* bg,
                          int n_sets, int n_masks, const uint32_t* bits, const double* kw, int link,
                          fdb_kshap_plan** out);
/* Rows for which only some features differ from the background (shap's varyingInds / M < 30, SURVEY.md 8(a) a9
 * step 1: a feature "varies" when the row's value differs from at least one background row).  varying_mask has
 * bit k set for each varying feature; it must contain bit 29 (the caller permutes the feature axis of coef, bg and
 * the rows so that the LAST varying feature sits at index 29 - the one the constraint eliminates, as the reference
 * eliminates the last varying column) and at least 2 bits.  Coalition bits of the other features are ignored; those
 * features get phi = 0, the schedule constants (M - s, num_features) use M = popcount(varying_mask). */
int fdb_kshap_plan_create_varying(fdb_ctx* ctx, const double* coef, double intercept, int n_bg, const double* bg,
                                  int n_sets, int n_masks, const uint32_t* bits, const double* kw, int link,
                                  uint32_t varying_mask, fdb_kshap_plan** out);
/* Same plan, but the n_sets coalition sets are drawn ON THE DEVICE with the reference's schedule (enumerated
 * sizes, paired random draws, duplicate merging, weight renormalisation; Philox stream keyed by (seed, set)):
 * statistically, not bitwise, equivalent to shap's per-row sampling.  nsamples <= 0 means "auto" (2*30 + 2048);
 * 64 <= nsamples <= 4096.  fdb_kshap_plan_get_masks copies one set back (n_masks entries each). */
int fdb_kshap_plan_create_sampled(fdb_ctx* ctx, const double* coef, double intercept, int n_bg, const double* bg,
                                  int n_sets, int nsamples, uint64_t seed, int link, fdb_kshap_plan** out);
int fdb_kshap_plan_get_masks(const fdb_kshap_plan* plan, int set, uint32_t* bits, float* kw);
/* Redraw every coalition set of a sampled plan in place (no allocation of the large buffers): fresh coalitions for
 * the next batch of rows, as the reference draws them per explain() call. */
int fdb_kshap_plan_resample(fdb_kshap_plan* plan, uint64_t seed);
/* The same for plans with more than one set, ENQUEUED on a stream (void* = cudaStream_t) with no host synchronisation:
 * two kernels (draw the sets, per-set (X'WX)^-1), so that a row stream can alternate two plans on two streams and draw
 * the sets of batch i + 1 while batch i is being explained.  A set whose X'WX turns out singular (practically impossible
 * at >= 256 samples over 30 features) cannot raise here: its rows get phi = 0 except for the last feature, and
 * fdb_kshap_plan_status reports (and clears) how many such sets were seen since the last call. */
int fdb_kshap_plan_resample_dev(fdb_kshap_plan* plan, void* stream, uint64_t seed);
int fdb_kshap_plan_status(fdb_kshap_plan* plan, int* n_singular);
int fdb_kshap_plan_destroy(fdb_kshap_plan* plan);
double fdb_kshap_expected_value(const fdb_kshap_plan* plan);
int fdb_kshap_plan_info(const fdb_kshap_plan* plan, int* info8);
/* l1_reg = "num_features(k)" of KernelExplainer.shap_values: k > 0 selects k features per row with LARS on the
 * augmented system (sklearn lars_path semantics, float64) and solves the constrained WLS over them only;
 * k = 0 (default) is l1_reg=False.  k <= 16. */
int fdb_kshap_plan_set_num_features(fdb_kshap_plan* plan, int k);
/* The other l1_reg values of KernelExplainer.shap_values (shap/explainers/_kernel.py::solve [recalled]):
 *   FDB_KSHAP_L1_NONE          l1_reg=False / 0                       param ignored
 *   FDB_KSHAP_L1_NUM_FEATURES  "num_features(k)"                      param = k (same as fdb_kshap_plan_set_num_features)
 *   FDB_KSHAP_L1_AIC / _BIC    "aic" / "bic": make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion)) on
 *                              the augmented system; the features with a nonzero coefficient at the criterion's minimum
 *                              along the LARS-lasso path are kept           param ignored
 *   FDB_KSHAP_L1_ALPHA         a float: Lasso(alpha=param) on the augmented system, nonzero coefficients are kept
 * "auto" is the caller's decision (AIC when nsamples / (2^M - 2) < 0.2, none otherwise).  Selection runs per explained
 * row on the device in float64 (csrc/kshap_lasso.cu); the constrained WLS is then solved over the kept features. */
enum { FDB_KSHAP_L1_NONE = 0, FDB_KSHAP_L1_NUM_FEATURES = 1, FDB_KSHAP_L1_AIC = 2, FDB_KSHAP_L1_BIC = 3, FDB_KSHAP_L1_ALPHA = 4 };
int fdb_kshap_plan_set_l1(fdb_kshap_plan* plan, int mode, double param);
int fdb_kshap_explain_dev(fdb_kshap_plan* plan, void* stream, int64_t n_rows, const float* x,
                          const int32_t* set_idx, int evaluator, float* phi, float* ey_out, double* abs_sum);
int fdb_kshap_explain_host(fdb_kshap_plan* plan, int64_t n_rows, const float* x, const int32_t* set_idx,
                           int evaluator, float* phi, double* abs_sum);

/* ---------------------------------------------------------------------------------
 * The fused linear pass of K1b for callers that are not allowed to allocate device memory
 * (no counterpart in the reference - it exists for embedders of this library, not for
 * parity).  Same arguments and results as fdb_linear_pass_dev, all pointers are DEVICE
 * pointers, but there is no fdb_ctx: the kernel's scratch is a fixed array built into the
 * library image, so the call performs no cudaMalloc/cudaFree and calls nothing that does.
 * It runs on the current device.  Limits and rules:
 *   - at most FDB_NOALLOC_MAX_ROWS rows per call; more returns FDB_ERR_ARG;
 *   - one scratch array per device, shared by all calls: calls that request abs_sum and
 *     may overlap on the same GPU race with each other - serialise them;
 *   - the first call on a device looks the array up (host-side set-up); make it during
 *     start-up, before any stream capture.
 * ------------------------------------------------------------------------------- */
#define FDB_NOALLOC_MAX_ROWS 16384
int fdb_linear_pass_noalloc(void* stream, int dtype, int64_t n_rows, const void* x,
                            const double* sub, const double* mul, double bias,
                            void* proba, int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum);

/* ---------------------------------------------------------------------------------
 * Result sink (api/worker.py:82-88): the task stores json.dumps(shap_values[0].tolist()).
 * fdb_format_json_rows writes exactly that text for every row of a (n_rows, n_cols) float64
 * HOST array - shortest round-trip digits, Python's fixed / exponent layout, "NaN" /
 * "Infinity" like json.dumps, ", " separators - back to back into `out`; row r occupies
 * [offsets[r], offsets[r + 1]) (offsets has n_rows + 1 entries).  out_cap >=
 * n_rows * (n_cols * 28 + 2) always fits; a smaller buffer may return FDB_ERR_ARG.
 * Host code only (no GPU needed); rows are independent, callers may split them over threads.
 * ------------------------------------------------------------------------------- */
int fdb_format_json_rows(const double* values, int64_t n_rows, int n_cols, char* out, int64_t out_cap, int64_t* offsets);

/* ---------------------------------------------------------------------------------
 * Measurement aid (no counterpart in the reference): instruction-pipe rates of this GPU,
 * measured with small saturating kernels at the clock the device runs them at.  They are
 * the denominators of the Kernel SHAP rooflines in bench.py (MEASURED_PEAKS.json has no
 * SFU / FP32-issue figure).  out8: [0] ex2.approx results/s, [1] rcp.approx results/s,
 * [2] independent FFMA results/s, [3] sigmoids/s of K2f's clamp-free quad loop (broadcast
 * LDS.128 reads included) at 16 resident warps per SM, [4] the same at 64 warps per SM,
 * [5] 0, [6] SM count, [7] 0.  Synchronous; allocates 256 bytes while it runs.
 * ------------------------------------------------------------------------------- */
int fdb_probe_pipes(fdb_ctx* ctx, double* out8);

#ifdef __cplusplus
}
#endif
#endif /* FDB200_H */
