/*
 * hm_oracle.h -- CPU oracle for the hivemall_b200 hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED.  The mounted reference (/root/reference, maropu/hivemall) is a
 * 22-line deprecation README (README.md:20-22) with no source, tests or golden
 * vectors, and no JVM exists in this image.  This oracle therefore restates the
 * PUBLISHED algorithms that BASELINE.json's north_star names (MurmurHash3_x86_32,
 * logistic / hinge / squared loss, SGD, AdaGrad, AdaDelta, Adam, FTRL-Proximal,
 * Rendle's factorization machine) under the arithmetic contract written in
 * DESIGN.md section 3.  It is NOT a restatement of Hivemall's Java code and must
 * not be described as one.  Only MurmurHash3 is pinned, against the public
 * known-answer vectors (tests/golden/murmur3_kat.json).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (hivemall_b200/) never does.
 */
#ifndef HM_ORACLE_H
#define HM_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { HMO_TASK_CLASSIFIER = 0, HMO_TASK_REGRESSOR = 1 };
enum { HMO_LOSS_LOG = 0, HMO_LOSS_HINGE = 1, HMO_LOSS_SQUARED = 2,
       /* online learners on the mini-batch skeleton: perceptron (Rosenblatt 1958) and the
        * passive-aggressive family (Crammer, Dekel, Keshet, Shalev-Shwartz, Singer, JMLR 2006) */
       HMO_LOSS_PERCEPTRON = 3, HMO_LOSS_PA = 4, HMO_LOSS_PA1 = 5, HMO_LOSS_PA2 = 6,
       HMO_LOSS_PA1_REGR = 7, HMO_LOSS_PA2_REGR = 8,
       /* adaptive regularization of weights (Crammer, Kulesza, Dredze, NIPS 2009), diagonal covariance:
        * classification, regression, regression with an insensitive zone */
       HMO_LOSS_AROW = 9, HMO_LOSS_AROW_REGR = 10, HMO_LOSS_AROWE_REGR = 11,
       /* confidence-weighted (Dredze, Crammer, Pereira, ICML 2008) and soft confidence-weighted learning
        * (Wang, Zhao, Hoi, ICML 2012), diagonal covariance */
       HMO_LOSS_CW = 12, HMO_LOSS_SCW1 = 13, HMO_LOSS_SCW2 = 14,
       /* further plain losses: squared hinge, modified Huber (classification); Huber, quantile (pinball),
        * epsilon-insensitive and its square (regression) */
       HMO_LOSS_SQUARED_HINGE = 15, HMO_LOSS_MODIFIED_HUBER = 16, HMO_LOSS_HUBER = 17, HMO_LOSS_QUANTILE = 18,
       HMO_LOSS_EPS_INSENSITIVE = 19, HMO_LOSS_SQUARED_EPS_INSENSITIVE = 20 };
enum { HMO_OPT_SGD = 0, HMO_OPT_ADAGRAD = 1, HMO_OPT_ADADELTA = 2, HMO_OPT_ADAM = 3, HMO_OPT_FTRL = 4,
       HMO_OPT_COV = 5 /* the AROW learners: s1 = per-feature variance, 1 at the start (the CALLER fills it) */ };
enum { HMO_ETA_FIXED = 0, HMO_ETA_INVERSE = 1, HMO_ETA_SIMPLE = 2 };
enum { HMO_REG_NONE = 0, HMO_REG_L1 = 1, HMO_REG_L2 = 2 };
enum { HMO_NORM_BATCH = 0, HMO_NORM_COUNT = 1, HMO_NORM_SUM = 2 };

/* Numeric training configuration (the product parses UDTF option strings into
 * an equivalent struct; the oracle takes numbers only). */
typedef struct hmo_config {
    int32_t task, loss, opt, eta_sched, reg, grad_norm;
    int32_t mini_batch, iters;
    int64_t dims;
    double  total_steps;          /* HMO_ETA_SIMPLE denominator */
    float   eta0, power_t;
    float   lambda;               /* L1 / L2 strength */
    float   eps;                  /* AdaGrad / AdaDelta / Adam epsilon */
    float   rho;                  /* AdaDelta decay */
    float   beta1, beta2;         /* Adam */
    float   alpha, beta, lambda1, lambda2; /* FTRL-Proximal */
    /* factorization machine only */
    int32_t factors;
    int32_t fm_classification;
    float   sigma;                /* V init half-width */
    float   lambda_w0, lambda_w, lambda_v;
    uint32_t seed;
    /* passive-aggressive learners */
    float   pa_c;                 /* aggressiveness C */
    float   pa_epsilon;           /* regression: insensitive zone */
    float   cov_r;                /* AROW: regularisation r */
    float   cov_phi;              /* CW / SCW: confidence phi */
    float   loss_param;           /* Huber's threshold / the quantile */
} hmo_config;

/* MurmurHash3_x86_32 (Appleby, public domain algorithm). */
uint32_t hmo_murmur3_x86_32(const void *key, int64_t len, uint32_t seed);
/* Fold a hash into [0, num_features): Java-style signed remainder made non-negative. */
int32_t hmo_fold(uint32_t h, int32_t num_features);
/* Hash n UTF-8 strings packed in `bytes` with offsets[n+1]. */
void hmo_mhash_batch(const uint8_t *bytes, const int64_t *offsets, int64_t n,
                     uint32_t seed, int32_t num_features, int32_t *out);

/* exp() with the exact operation sequence of DESIGN.md section 3.2. */
float hmo_exp(float x);
/* dloss/dmargin for one row; y is +-1 for classifier losses, the target otherwise. */
float hmo_dloss(int32_t loss, float margin, float y);
/* The number that takes dloss's place for a row, any loss id: hmo_dloss for the three losses, else the
 * online learner's -(step size) * (direction); xx = |x|^2 of the row (fp32 squares, fp64 sum, rounded once). */
float hmo_row_coef(const hmo_config *c, float margin, float y, float xx);
/* AROW: the two numbers that take dloss's place for a row -- c1 for the weight sums (times x), c2 for the
 * variance sums (times x^2) -- from the margin, the target and the row's variance v = sum sigma_j x_j^2. */
void hmo_cov_coefs(const hmo_config *c, float margin, float y, float v, float *c1, float *c2);
/* loss value in double (reporting only). */
double hmo_loss(int32_t loss, float margin, float y);

/* Learning rate for 1-based update count t. */
float hmo_eta(const hmo_config *c, int64_t t);

/*
 * Linear learner (train_classifier / train_regressor).  Trains `iters` passes over
 * the CSR rows in order, one optimizer update per mini-batch of `mini_batch` rows
 * (the last batch of a pass may be short).  w, s1, s2 are caller-owned arrays of
 * `dims` floats, read and updated in place (s1/s2 meaning depends on opt:
 * AdaGrad s1=G; AdaDelta s1=E[g^2], s2=E[dx^2]; Adam s1=m, s2=v; FTRL s1=z, s2=n).
 * *t is the update counter (in/out).  margins (nullable) receives the margin each
 * row had when it was scored in the LAST pass.  Returns 0 or a negative error.
 */
int hmo_train_linear(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                     const uint32_t *idx, const float *val, const float *label,
                     float *w, float *s1, float *s2, int64_t *t,
                     float *margins, double *loss_sum);

/* Same result bit for bit, rows and features spread over `threads` OpenMP threads
 * (the timed CPU baseline of bench.py). */
int hmo_train_linear_mt(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                        const uint32_t *idx, const float *val, const float *label,
                        float *w, float *s1, float *s2, int64_t *t, int threads);

/* margin[i] = sum_j w[idx]*val, DESIGN.md section 3.1 arithmetic. */
void hmo_predict_linear(int64_t n_rows, const int64_t *rowptr, const uint32_t *idx,
                        const float *val, const float *w, float *margin);

/* Factorization machine. V is dims x factors row-major.  gw/gv are AdaGrad
 * accumulators (dims, dims x factors), used when opt == HMO_OPT_ADAGRAD.
 * touched (dims bytes, nullable) is set to 1 for every feature seen. */
void hmo_fm_init_v(const hmo_config *c, float *V);
int hmo_train_fm(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                 const uint32_t *idx, const float *val, const float *label,
                 float *w0, float *w, float *V, float *gw0, float *gw, float *gv,
                 uint8_t *touched, int64_t *t, float *preds, double *loss_sum);
void hmo_predict_fm(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                    const uint32_t *idx, const float *val, float w0, const float *w,
                    const float *V, float *pred);

/* Model table of a linear model: features with w != 0 in ascending order.
 * Returns the number of rows; writes at most cap. */
int64_t hmo_model_table(const float *w, int64_t dims, int32_t *feature, float *weight, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif
