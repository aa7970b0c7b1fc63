/*
 * hivemall_b200.h -- C-ABI of the B200-native Hivemall training hot path.
 *
 * One shared library (hivemall_b200/libhivemall_b200.so, plain CUDA C++, no torch
 * types) exports everything a JNI / Panama binding of the path would call.
 *
 * Reference citations.  The mounted reference (/root/reference, maropu/hivemall)
 * is a README-only deprecation stub (README.md:20 "development moved to the ASF
 * repository", README.md:22 "This repository became deprecated"); it has no
 * source file to cite.  Each entry point therefore names the interface that
 * BASELINE.json's north_star says it stands in for (the Hive UDTF lifecycle of
 * train_classifier / train_regressor / train_fm, and the mhash / feature_hashing
 * UDFs).  PARITY IS UNPINNED until real sources are mounted; every fact that
 * would have to be read from them is a run-time parameter here, not a constant.
 *
 * Conventions
 *  - every function returns HM_OK (0) or an hm_status error code and never throws;
 *    hm_last_error() returns the calling thread's last message.
 *  - "host" pointers are ordinary (ideally pinned, see hm_host_alloc) CPU memory;
 *    "device" pointers (prefix d_) are CUDA device memory on the learner's device.
 *  - a learner handle is not thread-safe; use one handle per thread / per GPU
 *    (a Hive task owns exactly one UDTF instance, which is the same rule).
 *  - there is no CPU fallback: without a CUDA device every compute entry point
 *    fails with HM_ERR_CUDA.
 */
#ifndef HIVEMALL_B200_H
#define HIVEMALL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HM_ABI_VERSION 1

typedef enum hm_status {
    HM_OK = 0,
    HM_ERR_INVALID_ARGUMENT = 1, /* bad option string / bad argument (UDFArgumentException role) */
    HM_ERR_CUDA = 2,             /* CUDA runtime failure or no device */
    HM_ERR_STATE = 3,            /* call out of lifecycle order (HiveException role) */
    HM_ERR_OOM = 4,              /* device or host allocation failed */
    HM_ERR_UNSUPPORTED = 5       /* valid option that this build does not implement */
} hm_status;

enum { HM_TASK_CLASSIFIER = 0, HM_TASK_REGRESSOR = 1, HM_TASK_FM = 2 };
enum { HM_LOSS_LOG = 0, HM_LOSS_HINGE = 1, HM_LOSS_SQUARED = 2,
       /* the online learners that share the mini-batch skeleton (train_perceptron, train_pa, train_pa1, train_pa2,
        * train_pa1_regr, train_pa2_regr): the per-row coefficient is the learner's step size times the sign of
        * the correction instead of a loss derivative, and the optimizer is plain SGD with eta = 1 */
       HM_LOSS_PERCEPTRON = 3, HM_LOSS_PA = 4, HM_LOSS_PA1 = 5, HM_LOSS_PA2 = 6,
       HM_LOSS_PA1_REGR = 7, HM_LOSS_PA2_REGR = 8,
       /* learners that keep a per-feature variance beside the weight (train_arow, train_arow_regr,
        * train_arowe_regr): the forward pass also yields the row's variance, and a mini-batch ends in two
        * reduce-and-step launches, one for the weights and one for the variances */
       HM_LOSS_AROW = 9, HM_LOSS_AROW_REGR = 10, HM_LOSS_AROWE_REGR = 11,
       /* ... and the confidence-weighted family on the same two passes (train_cw, train_scw, train_scw2) */
       HM_LOSS_CW = 12, HM_LOSS_SCW1 = 13, HM_LOSS_SCW2 = 14,
       /* further -loss values of train_classifier (squared hinge, modified Huber) and train_regressor (Huber,
        * quantile, epsilon-insensitive and its square): plain loss derivatives like the first three */
       HM_LOSS_SQUARED_HINGE = 15, HM_LOSS_MODIFIED_HUBER = 16, HM_LOSS_HUBER = 17, HM_LOSS_QUANTILE = 18,
       HM_LOSS_EPS_INSENSITIVE = 19, HM_LOSS_SQUARED_EPS_INSENSITIVE = 20 };
enum { HM_OPT_SGD = 0, HM_OPT_ADAGRAD = 1, HM_OPT_ADADELTA = 2, HM_OPT_ADAM = 3, HM_OPT_FTRL = 4,
       HM_OPT_COV = 5 /* set by the variance-keeping learners, not by -opt: state = one variance per feature, 1 at the start */ };
enum { HM_ETA_FIXED = 0, HM_ETA_INVERSE = 1, HM_ETA_SIMPLE = 2 };
enum { HM_REG_NONE = 0, HM_REG_L1 = 1, HM_REG_L2 = 2 };
enum { HM_NORM_BATCH = 0, HM_NORM_COUNT = 1, HM_NORM_SUM = 2 };

/* Numeric form of a UDTF option string (hm_parse_options fills it). */
typedef struct hm_config {
    int32_t task, loss, opt, eta_sched, reg, grad_norm;
    int32_t mini_batch, iters;
    int64_t dims;
    double  total_steps;
    float   eta0, power_t;
    float   lambda;
    float   eps;
    float   rho;
    float   beta1, beta2;
    float   alpha, beta, lambda1, lambda2;
    int32_t factors;
    int32_t fm_classification;
    float   sigma;
    float   lambda_w0, lambda_w, lambda_v;
    uint32_t seed;
    int32_t mix_interval;    /* mini-batches between model averagings, for the CALLER's MIX loop (0 = its choice):
                              * the library itself never communicates, see hm_udtf_array / hm_udtf_scale_replica */
    uint32_t hash_seed;      /* MurmurHash3 seed used by feature hashing */
    float   pa_c;            /* passive-aggressive learners: aggressiveness C (PA-I cap, PA-II slack) */
    float   pa_epsilon;      /* ... regression: width of the insensitive zone (also train_arowe_regr) */
    float   cov_r;           /* train_arow*: regularisation r of the variance update */
    float   cov_phi;         /* train_cw, train_scw, train_scw2: confidence phi = probit(eta) */
    float   loss_param;      /* -loss huber: threshold c (-huber_c, default 1); -loss quantile: tau (-quantile, default 0.5) */
    int32_t reserved[3];
} hm_config;

typedef struct hm_udtf hm_udtf; /* opaque learner = one UDTF instance on one GPU */

/* Counters since creation. */
typedef struct hm_stats {
    int64_t rows, nnz, mini_batches, kernel_launches;
    int64_t h2d_bytes, d2h_bytes;
    double  reserved0;            /* always 0 in this version */
} hm_stats;

/* ---- library ------------------------------------------------------------- */

/* Thread-local text of the most recent error.  Never NULL. */
const char *hm_last_error(void);
int hm_abi_version(void);
/* Number of CUDA devices visible; 0 when there is none (compute calls then fail). */
int hm_device_count(void);

/* Pinned host memory for the host-buffer entry points (JNI: wrap in a direct
 * ByteBuffer).  Pageable memory is accepted everywhere but copies slower. */
hm_status hm_host_alloc(void **out, int64_t bytes);
hm_status hm_host_free(void *p);

/* ---- feature hashing  (stands in for the mhash / feature_hashing UDFs) ---- */

/* mhash over n UTF-8 strings packed back to back in `bytes`; string i is
 * bytes[offsets[i] .. offsets[i+1]).  out[i] = MurmurHash3_x86_32(string, seed)
 * folded into [0, num_features) with Java % semantics (negative remainders are
 * shifted up by num_features).  Host buffers; runs the CUDA kernel on `device`. */
hm_status hm_mhash(const uint8_t *bytes, const int64_t *offsets, int64_t n,
                   uint32_t seed, int32_t num_features, int32_t *out, int device);
/* Same, buffers already on the device, asynchronous on `stream` (a cudaStream_t). */
hm_status hm_mhash_device(const uint8_t *d_bytes, const int64_t *d_offsets, int64_t n,
                          uint32_t seed, int32_t num_features, int32_t *d_out, void *stream);
/* feature_hashing over "name:value" / "name" feature strings: the name part (up
 * to the LAST ':' when what follows parses as a number) is hashed as above, the
 * value part is parsed as a float (1.0 when absent).  A name made only of decimal
 * digits is taken as an index already and passed through unhashed. */
hm_status hm_feature_hashing(const uint8_t *bytes, const int64_t *offsets, int64_t n,
                             uint32_t seed, int32_t num_features,
                             int32_t *out_index, float *out_value, int device);

/* ---- option strings (stands in for the UDTF's getOptions/processOptions) --- */

/* -loss of train_classifier: logloss | hinge | squared_hinge | modified_huber | squared; of train_regressor:
 * squared | huber (-huber_c, default 1) | quantile (-quantile / -tau, default 0.5) | epsilon_insensitive |
 * squared_epsilon_insensitive (-epsilon, default 0.1).
 * function_name: "train_classifier" | "train_regressor" | "train_fm", or one of the online learners on the
 * same skeleton: "train_perceptron" | "train_pa" | "train_pa1" | "train_pa2" (labels as train_classifier) |
 * "train_pa1_regr" | "train_pa2_regr" (targets as train_regressor) | "train_arow" (labels) |
 * "train_arow_regr" | "train_arowe_regr" (targets) | "train_cw" | "train_scw" | "train_scw2" (labels).  Those fix loss, optimizer, eta (1) and regularisation
 * (none) themselves and take -mini_batch, -iters, -dims, -grad_norm, -mix_interval, -hash_seed, plus
 * -c / -aggressiveness (train_pa1, train_pa2, train_pa*_regr; default 1), -epsilon / -e (train_pa*_regr,
 * train_arowe_regr; default 0.1), -r (train_arow*; default 0.1) and -phi (train_cw, train_scw, train_scw2;
 * default 1; or -eta, a probability in (0.5, 1), with phi = probit(eta); train_scw and train_scw2 also take -c).
 * Unknown options, missing values and out-of-range numbers are
 * HM_ERR_INVALID_ARGUMENT with the offending token in hm_last_error(). */
hm_status hm_parse_options(const char *function_name, const char *options, hm_config *out);

/* ---- UDTF lifecycle ------------------------------------------------------- */

/* initialize(): parse options, allocate the model replica on `device`. */
hm_status hm_udtf_create(const char *function_name, const char *options, int device, hm_udtf **out);
/* Same from an already-numeric config. */
hm_status hm_udtf_create_from_config(const hm_config *cfg, int device, hm_udtf **out);

/* process() for n_rows rows at once, HOST buffers in CSR form: row i owns
 * feature[rowptr[i] .. rowptr[i+1]) / value[...]; rowptr[0] may be non-zero.
 * Rows are consumed in order; a mini-batch is trained as soon as -mini_batch rows
 * are buffered, exactly as if process() had been called row by row.
 * `value` may be NULL: every value is then 1.0 (Hivemall's index-only features,
 * "name" without ":value") and half of the feature bytes stay off the bus.
 * Copies host->device inside the call (asynchronously when the buffers are pinned).
 * LIFETIME: with pinned buffers (hm_host_alloc) the call returns while DMA is still reading
 * them -- rowptr, feature, value and label must stay valid AND unmodified until hm_udtf_sync,
 * hm_udtf_last_margins or hm_udtf_close returns (a JNI caller must not refill a direct
 * ByteBuffer before that).  The same holds for hm_udtf_process_split and
 * hm_udtf_process_strings.  A malformed row pointer is HM_ERR_INVALID_ARGUMENT; slices of the
 * call that were consumed before the bad row was met have been trained. */
hm_status hm_udtf_process(hm_udtf *u, int64_t n_rows, const int64_t *rowptr,
                          const int32_t *feature, const float *value, const float *label);
/* process() for rows whose features arrive in two groups, the way Hivemall's feature
 * engineering produces them: a quantitative group (index and value, as from
 * quantitative_features) and an index-only group (value 1.0 implied, as from
 * categorical_features).  Row i is its quantitative elements
 * q_feature/q_value[q_rowptr[i] .. q_rowptr[i+1]) followed by its index-only elements
 * c_feature[c_rowptr[i] .. c_rowptr[i+1]); the two groups are merged into one CSR on the
 * device, so the implied 1.0s never cross the bus.  Otherwise exactly hm_udtf_process
 * (same mini-batch boundaries, same result as pushing the merged rows). */
hm_status hm_udtf_process_split(hm_udtf *u, int64_t n_rows,
                                const int64_t *q_rowptr, const int32_t *q_feature, const float *q_value,
                                const int64_t *c_rowptr, const int32_t *c_feature, const float *label);
/* process() for rows whose quantitative features are the same n_q indices in every row -- what
 * quantitative_features(array("age", "height", ...), age, height, ...) produces: fixed names, one value
 * each -- followed by n_c index-only features (categorical_features: one per field).  The indices of
 * the quantitative block are given ONCE (q_feature[n_q]); q_value is the row-major matrix
 * [n_rows][n_q], c_feature the row-major matrix [n_rows][n_c].  Row i = (q_feature[j], q_value[i][j])
 * for j < n_q, then (c_feature[i][k], 1.0) for k < n_c.  Neither the repeated indices nor the implied
 * 1.0s nor a row pointer cross the bus (Criteo-shaped rows: 160 instead of 228 bytes per row), and
 * the rows are expanded into the learner's CSR on the device.  Otherwise exactly hm_udtf_process
 * (same mini-batch boundaries, same result as pushing the expanded rows). */
hm_status hm_udtf_process_columns(hm_udtf *u, int64_t n_rows, int32_t n_q, const int32_t *q_feature,
                                  const float *q_value, int32_t n_c, const int32_t *c_feature,
                                  const float *label);
/* process() for rows whose features are still strings ("name:value"); hashes on
 * the GPU with the learner's hash_seed into [0, dims) and then behaves as above.
 * rowptr indexes strings; string s is bytes[offsets[s] .. offsets[s+1]). */
hm_status hm_udtf_process_strings(hm_udtf *u, int64_t n_rows, const int64_t *rowptr,
                                  const uint8_t *bytes, const int64_t *offsets, const float *label);
/* process() for rows already resident in device memory (no copies).  n_rows must
 * be a multiple of -mini_batch unless this is the last call before close(), and
 * no host-buffered remainder may be pending.  nnz = d_rowptr[n_rows]-d_rowptr[0].
 * LIFETIME / ORDER: the arrays are used in place, by kernels on the learner's own streams
 * that do not wait for the caller's streams: finish writing them (and synchronise the stream
 * that wrote them) before the call, and keep them alive and unmodified until hm_udtf_sync or
 * hm_udtf_close returns; with -iters > 1 they are read again by close() (the later passes
 * replay them), so keep them until close() has returned. */
hm_status hm_udtf_process_device(hm_udtf *u, int64_t n_rows, int64_t nnz, const int64_t *d_rowptr,
                                 const int32_t *d_feature, const float *d_value, const float *d_label);

/* close(): train the last short mini-batch, run the remaining -iters passes over
 * the retained rows, wait for the GPU, and report the model-table row count. */
hm_status hm_udtf_close(hm_udtf *u, int64_t *n_model_rows);
/* forward(): copy the model table to host arrays of capacity `cap` rows.
 * Linear learners emit (feature, weight) for every weight != 0 in ascending
 * feature order.  train_fm emits a first row with feature = -1 holding w0, then
 * (feature, Wi, Vif[factors]) for every feature seen in training, ascending;
 * vif must hold cap * factors floats (NULL for linear learners). */
hm_status hm_udtf_forward(hm_udtf *u, int32_t *feature, float *weight, float *vif,
                          int64_t cap, int64_t *n_written);
hm_status hm_udtf_destroy(hm_udtf *u);

/* ---- model replica access (MIX averaging, warm start, tests) -------------- */

/* Device memory that makes up the replica (what MIX averaging all-reduces).  which = 0..n-1.
 * A linear learner has ONE array, "replica": `width` floats per feature, interleaved so that
 * a feature's weight and optimizer slots share a memory sector: width 1 = (w) for SGD,
 * 2 = (w, G) for AdaGrad and (z, n) for FTRL -- FTRL stores no weight, it is the closed form of
 * (z, n) --, 4 = (w, slot1, slot2, unused) for AdaDelta (E[g^2], E[dx^2]) and Adam (m, v).  n_floats = dims * width.  train_fm has four: "w0" (w0 and its
 * AdaGrad accumulator), "w" (width 1 or 2: (w[, acc]) per feature), "V" (width k or 2k: the k
 * factors of a feature, then their accumulators) and "touched" (width 0: a byte per feature,
 * 1 = seen in training; combine it with MAX, not with a sum, and do not scale it).  The
 * pointers stay valid until destroy; call hm_udtf_sync before touching them from another
 * stream.
 * MIX semantics: averaging means averaging THESE arrays.  For every optimizer but FTRL that
 * is the mean of the ranks' weights (and of their slots).  FTRL's replica is (z, n): the
 * mean of (z, n) is taken and the weight is then derived from it -- because the closed form
 * holds a soft threshold and 1/sqrt(n), that is NOT the mean of the ranks' weights (it can
 * zero or un-zero a feature); it is the usual way to combine FTRL replicas and keeps them
 * consistent state-wise, but a MIX server that averages weights would differ.  What the
 * reference averages is unknown (no source is mounted). */
int hm_udtf_num_arrays(const hm_udtf *u);
hm_status hm_udtf_array(hm_udtf *u, int which, const char **name, void **d_ptr,
                        int64_t *n_floats, int32_t *width);
/* Copy a LOGICAL array of the model to / from a host buffer of n floats.  Linear learners:
 * which = 0 is the dense weight vector w (n = dims); 1 is the optimizer state (n = dims *
 * slots, the slots of a feature adjacent).  train_fm: 0 = w (dims), 1 = V (dims * factors,
 * one row per feature), 2 = w0 (1), and with -opt adagrad 3 / 4 / 5 = the accumulators of
 * w / V / w0.  Synchronous.  write_array is the warm-start path (for FTRL write the state:
 * its weights are derived, array 0 is read-only). */
hm_status hm_udtf_read_array(hm_udtf *u, int which, float *host, int64_t n);
hm_status hm_udtf_write_array(hm_udtf *u, int which, const float *host, int64_t n);
/* Update counter t (mini-batches applied so far); settable for warm start / MIX. */
hm_status hm_udtf_get_step(const hm_udtf *u, int64_t *t);
hm_status hm_udtf_set_step(hm_udtf *u, int64_t t);
/* Wait until all work queued on the learner's streams has finished. */
hm_status hm_udtf_sync(hm_udtf *u);
/* The learner's compute stream as a cudaStream_t (for event timing by callers). */
void *hm_udtf_stream(hm_udtf *u);
/* Multiply every replica array by `scale` on the learner's stream (the 1/P step
 * of sum-then-scale averaging after an all-reduce). */
hm_status hm_udtf_scale_replica(hm_udtf *u, float scale);

/* ---- prediction (held-out AUC / logloss for parity) ----------------------- */

/* out[i] = margin of row i under the current model (linear: w.x ; fm: y_hat).  Host buffers;
 * waits for the result.  An element whose index is outside [0, dims) contributes 0 and makes the
 * call return HM_ERR_INVALID_ARGUMENT (after writing `out`); it does not taint close(). */
hm_status hm_udtf_predict(hm_udtf *u, int64_t n_rows, const int64_t *rowptr,
                          const int32_t *feature, const float *value, float *out);
/* Same for device buffers: ASYNCHRONOUS on hm_udtf_stream -- the call returns before d_out is
 * written; order your reads after that stream or call hm_udtf_sync.  Out-of-range indices
 * contribute 0 and are not reported by this variant. */
hm_status hm_udtf_predict_device(hm_udtf *u, int64_t n_rows, const int64_t *d_rowptr,
                                 const int32_t *d_feature, const float *d_value, float *d_out);

/* Margins the rows of the most recently trained mini-batch had when they were scored
 * (before that mini-batch's update): the per-step training signal a caller logs or turns
 * into a loss.  Copies min(cap, rows) floats to `host`, waits for the GPU, sets *n. */
hm_status hm_udtf_last_margins(hm_udtf *u, float *host, int64_t cap, int64_t *n);

/* ---- the model table as a file, warm start from one ------------------------------- */

enum { HM_TABLE_TEXT = 0, HM_TABLE_BINARY = 1 };

/* close() + forward() into a file: what a Hive job does with the rows a training UDTF forwards
 * (INSERT into a table (feature, weight[, Vif])).  HM_TABLE_TEXT: one row per line,
 * feature <delimiter> weight [<delimiter> v1,v2,...,vk], numbers printed so that they read back
 * bit for bit; delimiter 0 means TAB (Hive's text tables use \t or ^A: the reference's own
 * choice is unknown, hence the parameter).  HM_TABLE_BINARY: a 32-byte header ("HMTABLE1",
 * factors, dims, rows) followed by feature int32[rows], weight f32[rows], vif f32[rows*factors],
 * little endian.  Rows are in forward() order (ascending feature; train_fm: the w0 row, feature
 * -1, first).  *n_rows (nullable) receives the row count. */
hm_status hm_udtf_save_model_table(hm_udtf *u, const char *path, int format, char delimiter, int64_t *n_rows);
/* Warm start: set the replica of a fresh (or running) learner from such a file.  Weights (and
 * train_fm's w0 and factor rows; factors of features absent from the file keep their initial
 * values) are restored exactly.  Optimizer slots are NOT in a model table: they restart from
 * zero, so continuing is exact for -opt sgd and approximate for the others; for -opt ftrl, which
 * stores (z, n) and derives the weight, z is rebuilt from the weight with n = 0.  Call
 * hm_udtf_set_step to continue a learning-rate schedule. */
hm_status hm_udtf_load_model_table(hm_udtf *u, const char *path, int format, char delimiter, int64_t *n_rows);
/* The learner's numeric configuration (what hm_parse_options made of its option string). */
hm_status hm_udtf_config(const hm_udtf *u, hm_config *out);

/* Hot features of the most recently sorted ingest block (diagnostics, tests).  A few features carry
 * a large share of a mini-batch's elements: the quantitative features every row has, the head of
 * each categorical field.  Per ingest block the linear learners pick up to 64 of them by their
 * frequency in a sample of the block's rows and reduce them in the forward kernel instead of sorting
 * them (same arithmetic contract, other -- fixed -- summation order).  The choice affects speed
 * only.  Writes up to `cap` feature indices, most frequent first, sets *n; waits for the GPU.
 * HM_HOT_FEATURES=0 in the environment before hm_udtf_create turns the mechanism off (every
 * element is then sorted).  Stands in for nothing in the reference. */
hm_status hm_udtf_hot_features(hm_udtf *u, int32_t *feature, int32_t cap, int32_t *n);

/* ---- model table without device allocation ---------------------------------- */

/* Largest model hm_model_table_noalloc accepts (its built-in scratch is 64 KiB per GPU). */
#define HM_MODEL_TABLE_NOALLOC_MAX_DIMS (1 << 26)

/* Has no counterpart in the reference: it exists for hosts of this library that may not
 * allocate device memory at the moment they need the model table -- a task that was given a
 * fixed share of a GPU it shares with other tasks, or a loop running beside other work on
 * the GPU that cannot absorb the device-wide stall cudaMalloc / cudaFree bring with them.
 * (hm_udtf_forward allocates its scratch and staging buffers.)
 *
 * Writes the model table of the dense weight array d_w[dims] (device memory, one float per
 * feature): (feature, weight) for every weight != 0,
 * ascending feature order, into the caller's device arrays d_feature / d_weight of `cap`
 * rows, and the row count into *d_n_rows (device memory).  If *d_n_rows > cap only the
 * first `cap` rows were written.  All work is queued on `stream` (a cudaStream_t) of the
 * CURRENT device; nothing is copied to the host and the call does not wait.
 *
 * It calls no cudaMalloc / cudaFree and no library routine that allocates.  Its scratch is
 * one array compiled into the library (every GPU that loads the library holds those 64 KiB
 * until the process exits), hence the limit: dims > HM_MODEL_TABLE_NOALLOC_MAX_DIMS is
 * rejected with HM_ERR_INVALID_ARGUMENT.  The first call on each device looks up that
 * array's device address; make that first call before any stream capture begins.  The
 * array is shared by every call on the device and keeps whatever the last call left in
 * it; each call overwrites the part it reads before reading it, so nothing needs clearing,
 * but two calls that overlap on the GPU (different streams, or different threads) race on
 * it and produce garbage -- serialise them. */
hm_status hm_model_table_noalloc(const float *d_w, int64_t dims, int32_t *d_feature, float *d_weight,
                                 int64_t cap, int64_t *d_n_rows, void *stream);

/* ---- measurement ----------------------------------------------------------- */

hm_status hm_udtf_stats(const hm_udtf *u, hm_stats *out);
/* When on, every kernel launch is bracketed by CUDA events on the launching
 * stream and accumulated per kernel name (slows the run: use a separate pass). */
hm_status hm_udtf_set_profiling(hm_udtf *u, int on);
/* Per-kernel totals collected while profiling was on.  i = 0..n-1. */
int hm_udtf_num_kernels(const hm_udtf *u);
hm_status hm_udtf_kernel_time(hm_udtf *u, int i, const char **name, int64_t *launches,
                              double *total_ms, double *algorithmic_bytes);

#ifdef __cplusplus
}
#endif
#endif /* HIVEMALL_B200_H */
