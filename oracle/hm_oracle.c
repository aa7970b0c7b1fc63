/*
 * hm_oracle.c -- CPU oracle for the hivemall_b200 hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED (see hm_oracle.h): the reference mount holds no code
 * (/root/reference/README.md:20-22), so every function below follows a published
 * algorithm plus the arithmetic contract of DESIGN.md section 3, not Hivemall source.
 * Plain scalar loops, hash-map-free: a dense double accumulator per feature.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (oracle/Makefile).
 * -ffp-contract=off matters: DESIGN.md section 3 rounds every fp32 operation separately.
 */
#include "hm_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------- hashing */

static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

static inline uint32_t fmix32(uint32_t h)
{
    h ^= h >> 16; h *= 0x85ebca6bu;
    h ^= h >> 13; h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}

/* MurmurHash3_x86_32, Austin Appleby's published algorithm (SMHasher, public domain).
 * Pinned by tests/golden/murmur3_kat.json incl. the SMHasher verification value 0xB0F57EE3. */
uint32_t hmo_murmur3_x86_32(const void *key, int64_t len, uint32_t seed)
{
    const uint8_t *data = (const uint8_t *)key;
    const int64_t nblocks = len / 4;
    uint32_t h1 = seed;
    const uint32_t c1 = 0xcc9e2d51u, c2 = 0x1b873593u;
    for (int64_t i = 0; i < nblocks; i++) {
        uint32_t k1 = (uint32_t)data[4 * i] | ((uint32_t)data[4 * i + 1] << 8) |
                      ((uint32_t)data[4 * i + 2] << 16) | ((uint32_t)data[4 * i + 3] << 24);
        k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2;
        h1 ^= k1; h1 = rotl32(h1, 13); h1 = h1 * 5u + 0xe6546b64u;
    }
    const uint8_t *tail = data + nblocks * 4;
    uint32_t k1 = 0;
    switch (len & 3) {
    case 3: k1 ^= (uint32_t)tail[2] << 16; /* fallthrough */
    case 2: k1 ^= (uint32_t)tail[1] << 8;  /* fallthrough */
    case 1: k1 ^= (uint32_t)tail[0];
        k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2; h1 ^= k1;
    }
    h1 ^= (uint32_t)len;
    return fmix32(h1);
}

int32_t hmo_fold(uint32_t h, int32_t num_features)
{
    int32_t r = (int32_t)h % num_features; /* C99 truncating remainder == Java % */
    if (r < 0) r += num_features;
    return r;
}

void hmo_mhash_batch(const uint8_t *bytes, const int64_t *offsets, int64_t n,
                     uint32_t seed, int32_t num_features, int32_t *out)
{
    for (int64_t i = 0; i < n; i++)
        out[i] = hmo_fold(hmo_murmur3_x86_32(bytes + offsets[i], offsets[i + 1] - offsets[i], seed),
                          num_features);
}

/* ---------------------------------------------------------------- scalar math */

/* Cody-Waite reduction + degree-6 Horner, every op rounded to fp32. */
float hmo_exp(float x)
{
    x = fminf(fmaxf(x, -87.0f), 88.0f);
    float n = rintf(x * 1.44269504f);
    float r = x - n * 0.693359375f;
    r = r - n * -2.12194440e-4f;
    float p = 1.3888889e-3f;
    p = p * r + 8.3333333e-3f;
    p = p * r + 4.1666667e-2f;
    p = p * r + 1.6666667e-1f;
    p = p * r + 0.5f;
    p = p * r + 1.0f;
    p = p * r + 1.0f;
    union { uint32_t u; float f; } s;
    s.u = (uint32_t)((int32_t)n + 127) << 23;
    return p * s.f;
}

float hmo_dloss(int32_t loss, float m, float y)
{
    switch (loss) {
    case HMO_LOSS_LOG:   return -y / (1.0f + hmo_exp(y * m));
    case HMO_LOSS_HINGE: return (y * m < 1.0f) ? -y : 0.0f;
    default:             return m - y;
    }
}

static inline int loss_is_pa(int32_t loss) { return loss >= HMO_LOSS_PA && loss <= HMO_LOSS_PA2_REGR; }
static inline int loss_is_cov(int32_t loss) { return loss >= HMO_LOSS_AROW && loss <= HMO_LOSS_SCW2; }

/* derivative of the further plain losses with respect to the margin */
static float dloss_more(const hmo_config *c, float m, float y)
{
    const int32_t loss = c->loss;
    const float p = c->loss_param, eps = c->pa_epsilon;
    if (loss == HMO_LOSS_SQUARED_HINGE) {
        const float z = y * m;
        return z < 1.0f ? (-2.0f * y) * (1.0f - z) : 0.0f;
    }
    if (loss == HMO_LOSS_MODIFIED_HUBER) {
        const float z = y * m;
        if (z >= 1.0f) return 0.0f;
        if (z >= -1.0f) return (-2.0f * y) * (1.0f - z);
        return -4.0f * y;
    }
    const float e = m - y;
    if (loss == HMO_LOSS_HUBER) return fabsf(e) <= p ? e : (e < 0.0f ? -p : p);
    if (loss == HMO_LOSS_QUANTILE) return e < 0.0f ? -p : 1.0f - p;
    const float a = fabsf(e) - eps;
    if (!(a > 0.0f)) return 0.0f;
    if (loss == HMO_LOSS_EPS_INSENSITIVE) return e < 0.0f ? -1.0f : 1.0f;
    return (e < 0.0f ? -2.0f : 2.0f) * a;
}

float hmo_row_coef(const hmo_config *c, float m, float y, float xx)
{
    const int32_t loss = c->loss;
    if (loss <= HMO_LOSS_SQUARED) return hmo_dloss(loss, m, y);
    if (loss >= HMO_LOSS_SQUARED_HINGE) return dloss_more(c, m, y);
    if (loss == HMO_LOSS_PERCEPTRON) return (y * m <= 0.0f) ? -y : 0.0f;
    float l, dir;
    if (loss == HMO_LOSS_PA1_REGR || loss == HMO_LOSS_PA2_REGR) {
        const float e = y - m;                      /* epsilon-insensitive loss */
        l = fabsf(e) - c->pa_epsilon;
        dir = e < 0.0f ? -1.0f : 1.0f;
    } else {
        l = 1.0f - y * m;                           /* hinge loss */
        dir = y;
    }
    if (!(l > 0.0f)) return 0.0f;
    float tau;
    if (loss == HMO_LOSS_PA2 || loss == HMO_LOSS_PA2_REGR) tau = l / (xx + 0.5f / c->pa_c);
    else {
        if (!(xx > 0.0f)) return 0.0f;
        tau = l / xx;
        if (loss != HMO_LOSS_PA) tau = fminf(c->pa_c, tau);
    }
    return -(tau * dir);
}

/* AROW (Crammer, Kulesza, Dredze 2009), diagonal: beta = 1 / (v + r); classification: l = 1 - y m, step when
 * l > 0, alpha = l beta, w_j += alpha y sigma_j x_j, sigma_j -= beta sigma_j^2 x_j^2.  Regression: e = y - m
 * in place of y l (every row steps); with an insensitive zone: l = |e| - epsilon, step when l > 0, along sgn(e).
 * c1 = -(alpha * direction) so that the common "w -= sigma * mean(c1 x)" adds the step; c2 = beta. */
void hmo_cov_coefs(const hmo_config *c, float m, float y, float v, float *c1, float *c2)
{
    *c1 = 0.0f; *c2 = 0.0f;
    if (c->loss >= HMO_LOSS_CW) {
        /* CW: alpha = max(0, (-(1 + 2 phi M) + sqrt((1 + 2 phi M)^2 - 8 phi (M - phi v))) / (4 phi v)), M = y m;
         * the precision of feature j grows by 2 alpha phi x_j^2.
         * SCW (loss phi sqrt(v) - M > 0): SCW-I alpha = min(C, (sqrt(M^2 phi^4 / 4 + v phi^2 zeta) - M psi) / (v zeta)),
         * psi = 1 + phi^2 / 2, zeta = 1 + phi^2; SCW-II alpha = (gamma - (2 M n + phi^2 M v)) / (2 (n^2 + n v phi^2)),
         * n = v + 1 / (2C), gamma = phi sqrt(phi^2 M^2 v^2 + 4 n v (n + v phi^2));
         * sqrt(u) = (sqrt(alpha^2 v^2 phi^2 + 4 v) - alpha v phi) / 2, beta = alpha phi / (sqrt(u) + alpha v phi). */
        const float phi = c->cov_phi, cc = c->pa_c;
        if (!(v > 0.0f)) return;
        const float my = y * m, p2 = phi * phi;
        float alpha;
        if (c->loss == HMO_LOSS_CW) {
            const float q = 1.0f + (2.0f * phi) * my;
            const float rad = q * q - (8.0f * phi) * (my - phi * v);
            if (!(rad >= 0.0f)) return;
            alpha = (sqrtf(rad) - q) / ((4.0f * phi) * v);
            if (!(alpha > 0.0f)) return;
            *c1 = -(alpha * y);
            *c2 = (2.0f * alpha) * phi;
            return;
        }
        if (!(phi * sqrtf(v) - my > 0.0f)) return;
        const float t1 = my * my;
        if (c->loss == HMO_LOSS_SCW1) {
            const float psi = 1.0f + p2 * 0.5f, zeta = 1.0f + p2;
            const float rad = (t1 * (p2 * p2)) * 0.25f + (v * p2) * zeta;
            alpha = (sqrtf(rad) - my * psi) / (v * zeta);
            alpha = fminf(cc, alpha);
        } else {
            const float n = v + 0.5f / cc;
            const float vp2 = v * p2;
            const float g1 = ((p2 * t1) * v) * v;
            const float g2 = ((4.0f * n) * v) * (n + vp2);
            const float gamma = phi * sqrtf(g1 + g2);
            const float num = gamma - ((2.0f * my) * n + (p2 * my) * v);
            const float den = 2.0f * (n * n + (n * v) * p2);
            alpha = num / den;
        }
        if (!(alpha > 0.0f)) return;
        const float avp = (alpha * v) * phi;
        const float usq = 0.5f * (sqrtf(avp * avp + 4.0f * v) - avp);
        *c1 = -(alpha * y);
        *c2 = (alpha * phi) / (usq + avp);
        return;
    }
    float l, dir;
    if (c->loss == HMO_LOSS_AROW) { l = 1.0f - y * m; dir = y; if (!(l > 0.0f)) return; }
    else if (c->loss == HMO_LOSS_AROW_REGR) { const float e = y - m; l = fabsf(e); dir = e < 0.0f ? -1.0f : 1.0f; }
    else { const float e = y - m; l = fabsf(e) - c->pa_epsilon; dir = e < 0.0f ? -1.0f : 1.0f; if (!(l > 0.0f)) return; }
    const float beta = 1.0f / (v + c->cov_r);
    const float alpha = l * beta;
    *c1 = -(alpha * dir);
    *c2 = beta;
}

static inline float target_of(const hmo_config *c, float label);
static inline float mean_grad(const hmo_config *c, double sum, uint32_t cnt, double inv_nb);

static int train_cov(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                     const uint32_t *idx, const float *val, const float *label,
                     float *w, float *sg, int64_t *t, float *margins, double *loss_sum)
{
    if (!sg) return -4;
    const int64_t B = c->mini_batch;
    double *g1 = (double *)calloc((size_t)c->dims, sizeof(double));
    double *g2 = (double *)calloc((size_t)c->dims, sizeof(double));
    uint32_t *cnt = (uint32_t *)calloc((size_t)c->dims, sizeof(uint32_t));
    float *c1 = (float *)malloc((size_t)B * sizeof(float)), *c2 = (float *)malloc((size_t)B * sizeof(float));
    int64_t max_touch = 1;
    for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
        int64_t e = b0 + B < n_rows ? b0 + B : n_rows;
        if (rowptr[e] - rowptr[b0] > max_touch) max_touch = rowptr[e] - rowptr[b0];
    }
    uint32_t *touch = (uint32_t *)malloc((size_t)max_touch * sizeof(uint32_t));
    if (!g1 || !g2 || !cnt || !c1 || !c2 || !touch) return -2;
    double lsum = 0.0;
    int rc = 0;
    for (int32_t it = 0; it < c->iters && rc == 0; it++) {
        lsum = 0.0;
        for (int64_t b0 = 0; b0 < n_rows && rc == 0; b0 += B) {
            const int64_t b1 = b0 + B < n_rows ? b0 + B : n_rows;
            /* margin and variance of every row with the model as it stands */
            for (int64_t i = b0; i < b1; i++) {
                double am = 0.0, av = 0.0;
                for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                    if ((int64_t)idx[p] >= c->dims) { rc = -3; break; }
                    const float x = val[p];
                    am += (double)(w[idx[p]] * x);
                    const float xx = x * x;
                    av += (double)(sg[idx[p]] * xx);
                }
                if (rc) break;
                const float m = (float)am, y = target_of(c, label[i]);
                hmo_cov_coefs(c, m, y, (float)av, &c1[i - b0], &c2[i - b0]);
                lsum += hmo_loss(c->loss, m, y);
                if (margins) margins[i] = m;
            }
            if (rc) break;
            int64_t nt = 0;
            for (int64_t i = b0; i < b1; i++)
                for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                    const uint32_t j = idx[p];
                    if (cnt[j]++ == 0) touch[nt++] = j;
                    const float x = val[p];
                    g1[j] += (double)(c1[i - b0] * x);
                    const float xx = x * x;
                    g2[j] += (double)(c2[i - b0] * xx);
                }
            *t += 1;
            const double inv_nb = 1.0 / (double)(b1 - b0);
            for (int64_t k = 0; k < nt; k++) {
                const uint32_t j = touch[k];
                const float m1 = mean_grad(c, g1[j], cnt[j], inv_nb), m2 = mean_grad(c, g2[j], cnt[j], inv_nb);
                const float s = sg[j];
                w[j] = w[j] - s * m1;            /* both steps use the variance of the start of the mini-batch */
                if (c->loss == HMO_LOSS_CW) sg[j] = 1.0f / (1.0f / s + m2);       /* the precision grows */
                else { const float ss = s * s; sg[j] = s - ss * m2; }
                g1[j] = 0.0; g2[j] = 0.0; cnt[j] = 0;
            }
        }
    }
    if (loss_sum) *loss_sum = lsum;
    free(g1); free(g2); free(cnt); free(c1); free(c2); free(touch);
    return rc;
}

static inline float row_sqnorm(const int64_t *rowptr, const float *val, int64_t i)
{
    double acc = 0.0;
    for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) acc += (double)(val[p] * val[p]);
    return (float)acc;
}

double hmo_loss(int32_t loss, float m, float y)
{
    double z = (double)y * (double)m;
    if (loss == HMO_LOSS_PERCEPTRON) return z <= 0.0 ? -z : 0.0;
    if (loss == HMO_LOSS_PA || loss == HMO_LOSS_PA1 || loss == HMO_LOSS_PA2) loss = HMO_LOSS_HINGE;
    if (loss >= HMO_LOSS_SQUARED_HINGE) {       /* (reporting only; parameters not known here: unit Huber threshold, median, eps 0) */
        if (loss == HMO_LOSS_SQUARED_HINGE) return z < 1.0 ? (1.0 - z) * (1.0 - z) : 0.0;
        if (loss == HMO_LOSS_MODIFIED_HUBER) return z >= 1.0 ? 0.0 : (z >= -1.0 ? (1.0 - z) * (1.0 - z) : -4.0 * z);
        const double e = fabs((double)m - (double)y);
        if (loss == HMO_LOSS_HUBER) return e <= 1.0 ? 0.5 * e * e : e - 0.5;
        if (loss == HMO_LOSS_SQUARED_EPS_INSENSITIVE) return e * e;
        return loss == HMO_LOSS_QUANTILE ? 0.5 * e : e;
    }
    if (loss == HMO_LOSS_AROW || loss >= HMO_LOSS_CW) loss = HMO_LOSS_HINGE;
    if (loss == HMO_LOSS_AROW_REGR || loss == HMO_LOSS_AROWE_REGR) return fabs((double)y - (double)m);
    if (loss == HMO_LOSS_PA1_REGR || loss == HMO_LOSS_PA2_REGR) return fabs((double)y - (double)m);   /* (epsilon not known here) */
    switch (loss) {
    case HMO_LOSS_LOG:   return z > 0 ? log1p(exp(-z)) : -z + log1p(exp(z));
    case HMO_LOSS_HINGE: return z < 1.0 ? 1.0 - z : 0.0;
    default: { double d = (double)m - (double)y; return 0.5 * d * d; }
    }
}

float hmo_eta(const hmo_config *c, int64_t t)
{
    switch (c->eta_sched) {
    case HMO_ETA_INVERSE: return (float)((double)c->eta0 / pow((double)t, (double)c->power_t));
    case HMO_ETA_SIMPLE:  return (float)((double)c->eta0 / (1.0 + (double)t / c->total_steps));
    default:              return c->eta0;
    }
}

static inline float target_of(const hmo_config *c, float label)
{
    if (c->task == HMO_TASK_CLASSIFIER) return label > 0.0f ? 1.0f : -1.0f;
    return label;
}

/* ---------------------------------------------------------------- optimizers */

/* One coordinate update, DESIGN.md 3.6.  alpha_t is Adam's bias-corrected step. */
static void update_coord(const hmo_config *c, float eta, float alpha_t, float g,
                         float *w, float *s1, float *s2)
{
    float wv = *w;
    if (c->reg == HMO_REG_L2 && c->opt != HMO_OPT_FTRL) g = g + c->lambda * wv;
    switch (c->opt) {
    case HMO_OPT_SGD:
        wv = wv - eta * g;
        break;
    case HMO_OPT_ADAGRAD: {
        float G = *s1 + g * g;
        *s1 = G;
        wv = wv - (eta * g) / (sqrtf(G) + c->eps);
        break;
    }
    case HMO_OPT_ADADELTA: {
        float E = c->rho * *s1 + ((1.0f - c->rho) * g) * g;
        float dx = (sqrtf(*s2 + c->eps) / sqrtf(E + c->eps)) * g;
        *s2 = c->rho * *s2 + ((1.0f - c->rho) * dx) * dx;
        *s1 = E;
        wv = wv - dx;
        break;
    }
    case HMO_OPT_ADAM: {
        float m = c->beta1 * *s1 + (1.0f - c->beta1) * g;
        float v = c->beta2 * *s2 + ((1.0f - c->beta2) * g) * g;
        *s1 = m; *s2 = v;
        wv = wv - (alpha_t * m) / (sqrtf(v) + c->eps);
        break;
    }
    case HMO_OPT_FTRL: {
        float n0 = *s2;
        float n1 = n0 + g * g;
        float sigma = (sqrtf(n1) - sqrtf(n0)) / c->alpha;
        float z = *s1 + g - sigma * wv;
        *s1 = z; *s2 = n1;
        if (fabsf(z) <= c->lambda1) wv = 0.0f;
        else {
            float sgn = z < 0.0f ? -1.0f : 1.0f;
            wv = -(z - sgn * c->lambda1) / ((c->beta + sqrtf(n1)) / c->alpha + c->lambda2);
        }
        break;
    }
    }
    if (c->reg == HMO_REG_L1 && c->opt != HMO_OPT_FTRL) {
        float a = fabsf(wv) - eta * c->lambda;
        wv = a > 0.0f ? copysignf(a, wv) : 0.0f;
    }
    *w = wv;
}

static float adam_alpha(const hmo_config *c, float eta, int64_t t)
{
    double b1 = pow((double)c->beta1, (double)t), b2 = pow((double)c->beta2, (double)t);
    return (float)((double)eta * sqrt(1.0 - b2) / (1.0 - b1));
}

/* the per-feature fp64 sum becomes the fp32 mean gradient.  Batch
 * normalisation multiplies by the fp64 reciprocal of the row count (one rounding). */
static inline float mean_grad(const hmo_config *c, double sum, uint32_t cnt, double inv_nb)
{
    if (c->grad_norm == HMO_NORM_BATCH) return (float)(sum * inv_nb);
    if (c->grad_norm == HMO_NORM_COUNT) return (float)(sum / (double)cnt);
    return (float)sum;
}

/* ---------------------------------------------------------------- linear learner */

void hmo_predict_linear(int64_t n_rows, const int64_t *rowptr, const uint32_t *idx,
                        const float *val, const float *w, float *margin)
{
    for (int64_t i = 0; i < n_rows; i++) {
        double acc = 0.0;
        for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++)
            acc += (double)(w[idx[p]] * val[p]);   /* fp32 product, fp64 sum */
        margin[i] = (float)acc;
    }
}

int hmo_train_linear(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                     const uint32_t *idx, const float *val, const float *label,
                     float *w, float *s1, float *s2, int64_t *t,
                     float *margins, double *loss_sum)
{
    if (c->mini_batch < 1 || c->dims < 1) return -1;
    if (loss_is_cov(c->loss)) return train_cov(c, n_rows, rowptr, idx, val, label, w, s1, t, margins, loss_sum);
    const int64_t B = c->mini_batch;
    double *gsum = (double *)calloc((size_t)c->dims, sizeof(double));
    uint32_t *cnt = (uint32_t *)calloc((size_t)c->dims, sizeof(uint32_t));
    float *dl = (float *)malloc((size_t)B * sizeof(float));
    /* features touched by the current batch, in first-touch order */
    int64_t max_touch = 0;
    for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
        int64_t e = b0 + B < n_rows ? b0 + B : n_rows;
        if (rowptr[e] - rowptr[b0] > max_touch) max_touch = rowptr[e] - rowptr[b0];
    }
    uint32_t *touch = (uint32_t *)malloc((size_t)(max_touch > 0 ? max_touch : 1) * sizeof(uint32_t));
    if (!gsum || !cnt || !dl || !touch) { free(gsum); free(cnt); free(dl); free(touch); return -2; }
    double lsum = 0.0;
    float dummy = 0.0f;
    for (int32_t it = 0; it < c->iters; it++) {
        lsum = 0.0;
        for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
            const int64_t b1 = b0 + B < n_rows ? b0 + B : n_rows;
            const int64_t nb = b1 - b0;
            /* score every row of the batch with the weights as they stand */
            for (int64_t i = b0; i < b1; i++) {
                double acc = 0.0;
                for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                    if ((int64_t)idx[p] >= c->dims) { free(gsum); free(cnt); free(dl); free(touch); return -3; }
                    acc += (double)(w[idx[p]] * val[p]);   /* fp32 product, fp64 sum */
                }
                float m = (float)acc;
                float y = target_of(c, label[i]);
                dl[i - b0] = hmo_row_coef(c, m, y, loss_is_pa(c->loss) ? row_sqnorm(rowptr, val, i) : 0.0f);
                lsum += hmo_loss(c->loss, m, y);
                if (margins) margins[i] = m;
            }
            /* per-feature gradient sums in double */
            int64_t nt = 0;
            for (int64_t i = b0; i < b1; i++)
                for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                    uint32_t j = idx[p];
                    if (cnt[j]++ == 0) touch[nt++] = j;
                    gsum[j] += (double)(dl[i - b0] * val[p]);   /* fp32 product, fp64 sum (3.4) */
                }
            /* one optimizer update per touched feature */
            *t += 1;
            const float eta = hmo_eta(c, *t);
            const float alpha_t = c->opt == HMO_OPT_ADAM ? adam_alpha(c, eta, *t) : 0.0f;
            const double inv_nb = 1.0 / (double)nb;
            for (int64_t k = 0; k < nt; k++) {
                uint32_t j = touch[k];
                float g = mean_grad(c, gsum[j], cnt[j], inv_nb);
                update_coord(c, eta, alpha_t, g, &w[j], s1 ? &s1[j] : &dummy, s2 ? &s2[j] : &dummy);
                gsum[j] = 0.0; cnt[j] = 0;
            }
        }
    }
    if (loss_sum) *loss_sum = lsum;
    free(gsum); free(cnt); free(dl); free(touch);
    return 0;
}

/* Multi-threaded variant for the timed CPU baseline: same arithmetic and the same
 * per-feature summation order (rows in order), so the result is bit-identical to
 * hmo_train_linear.  Per mini-batch: (1) rows are scored in parallel, each thread owning
 * a contiguous row slice; (2) each thread buckets its slice's (feature, dloss*x) pairs by
 * the thread that owns the feature (contiguous feature ranges); (3) each owner walks the
 * buckets addressed to it in source-thread order -- i.e. in row order -- accumulates in
 * fp64 and applies the optimizer to its features. */
typedef struct { uint32_t j; float c; } hmo_pair;

int hmo_train_linear_mt(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                        const uint32_t *idx, const float *val, const float *label,
                        float *w, float *s1, float *s2, int64_t *t, int threads)
{
    if (c->mini_batch < 1 || c->dims < 1 || threads < 1) return -1;
    if (loss_is_cov(c->loss)) return train_cov(c, n_rows, rowptr, idx, val, label, w, s1, t, NULL, NULL);   /* (single thread) */
    if (threads > 256) threads = 256;
    const int64_t B = c->mini_batch;
    const int T = threads;
    double *gsum = (double *)calloc((size_t)c->dims, sizeof(double));
    uint32_t *cnt = (uint32_t *)calloc((size_t)c->dims, sizeof(uint32_t));
    float *dl = (float *)malloc((size_t)B * sizeof(float));
    int64_t max_touch = 1;
    for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
        int64_t e = b0 + B < n_rows ? b0 + B : n_rows;
        if (rowptr[e] - rowptr[b0] > max_touch) max_touch = rowptr[e] - rowptr[b0];
    }
    hmo_pair *pairs = (hmo_pair *)malloc((size_t)max_touch * sizeof(hmo_pair));
    uint32_t *touch = (uint32_t *)malloc((size_t)max_touch * sizeof(uint32_t));
    int64_t *off = (int64_t *)calloc((size_t)T * (T + 1), sizeof(int64_t));  /* off[src][owner] */
    if (!gsum || !cnt || !dl || !pairs || !touch || !off) return -2;
    const int64_t per_owner = (c->dims + T - 1) / T;
    float dummy[256];
    for (int32_t it = 0; it < c->iters; it++) {
        for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
            const int64_t b1 = b0 + B < n_rows ? b0 + B : n_rows;
            const int64_t nb = b1 - b0;
            const int64_t e0 = rowptr[b0];
            const double inv_nb = 1.0 / (double)nb;
            *t += 1;
            const float eta = hmo_eta(c, *t);
            const float alpha_t = c->opt == HMO_OPT_ADAM ? adam_alpha(c, eta, *t) : 0.0f;
#pragma omp parallel num_threads(T)
            {
#ifdef _OPENMP
                const int tid = omp_get_thread_num(), nth = omp_get_num_threads();
#else
                const int tid = 0, nth = 1;
#endif
                /* this thread's contiguous row slice */
                const int64_t r0 = b0 + nb * tid / nth, r1 = b0 + nb * (tid + 1) / nth;
                int64_t *my = off + (int64_t)tid * (T + 1);
                for (int k = 0; k <= T; k++) my[k] = 0;
                for (int64_t i = r0; i < r1; i++) {
                    double acc = 0.0;
                    for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                        acc += (double)(w[idx[p]] * val[p]);
                        my[idx[p] / per_owner + 1]++;
                    }
                    dl[i - b0] = hmo_row_coef(c, (float)acc, target_of(c, label[i]),
                                              loss_is_pa(c->loss) ? row_sqnorm(rowptr, val, i) : 0.0f);
                }
                /* bucket offsets inside this thread's region of `pairs` (region = its element range) */
                const int64_t base = rowptr[r0] - e0;
                int64_t run = base;
                for (int k = 0; k < nth; k++) { int64_t n = my[k + 1]; my[k] = run; run += n; }
                my[nth] = run;
                int64_t cur[256];
                for (int k = 0; k < nth; k++) cur[k] = my[k];
                for (int64_t i = r0; i < r1; i++)
                    for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                        const uint32_t j = idx[p];
                        hmo_pair *q = &pairs[cur[j / per_owner]++];
                        q->j = j;
                        q->c = dl[i - b0] * val[p];
                    }
#pragma omp barrier
                /* owner phase.  Owner k receives recv[k] elements, so it touches at most recv[k]
                 * features: its touched list is the slice of `touch` after all lower owners'. */
                int64_t tbase = 0;
                for (int k = 0; k < tid; k++)
                    for (int src = 0; src < nth; src++) {
                        const int64_t *so = off + (int64_t)src * (T + 1);
                        tbase += so[k + 1] - so[k];
                    }
                uint32_t *mine = touch + tbase;
                int64_t nt = 0;
                for (int src = 0; src < nth; src++) {
                    const int64_t *so = off + (int64_t)src * (T + 1);
                    for (int64_t q = so[tid]; q < so[tid + 1]; q++) {
                        const uint32_t j = pairs[q].j;
                        if (cnt[j]++ == 0) mine[nt++] = j;
                        gsum[j] += (double)pairs[q].c;
                    }
                }
                for (int64_t k = 0; k < nt; k++) {
                    const uint32_t j = mine[k];
                    const float g = mean_grad(c, gsum[j], cnt[j], inv_nb);
                    update_coord(c, eta, alpha_t, g, &w[j], s1 ? &s1[j] : &dummy[tid], s2 ? &s2[j] : &dummy[tid]);
                    gsum[j] = 0.0; cnt[j] = 0;
                }
            }
        }
    }
    free(pairs); free(touch); free(off); free(gsum); free(cnt); free(dl);
    return 0;
}

int64_t hmo_model_table(const float *w, int64_t dims, int32_t *feature, float *weight, int64_t cap)
{
    int64_t n = 0;
    for (int64_t j = 0; j < dims; j++)
        if (w[j] != 0.0f) {
            if (n < cap) { feature[n] = (int32_t)j; weight[n] = w[j]; }
            n++;
        }
    return n;
}

/* ---------------------------------------------------------------- factorization machine */

/* V[j][f] = sigma * u, u in [-1,1) from a 24-bit integer hash. */
void hmo_fm_init_v(const hmo_config *c, float *V)
{
    const int64_t n = c->dims * (int64_t)c->factors;
    for (int64_t q = 0; q < n; q++) {
        uint32_t h = fmix32((uint32_t)q * 0x9E3779B1u + c->seed);
        float u = ((float)(h >> 8) * 5.9604644775390625e-08f - 0.5f) * 2.0f;
        V[q] = c->sigma * u;
    }
}

static float fm_score(const hmo_config *c, const int64_t *rowptr, const uint32_t *idx,
                      const float *val, int64_t i, float w0, const float *w, const float *V,
                      float *S32 /* factors, nullable */)
{
    /* fp32 products, fp64 sums, S_f rounded to fp32 before it is squared */
    const int k = c->factors;
    double lin = 0.0;
    for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++)
        lin += (double)(w[idx[p]] * val[p]);
    double inter = 0.0;
    for (int f = 0; f < k; f++) {
        double S = 0.0, Q = 0.0;
        for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
            float vx = V[(int64_t)idx[p] * k + f] * val[p];
            S += (double)vx;
            Q += (double)(vx * vx);
        }
        float s32 = (float)S;
        if (S32) S32[f] = s32;
        inter += (double)(s32 * s32) - Q;
    }
    return (float)((double)w0 + lin + 0.5 * inter);
}

void hmo_predict_fm(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                    const uint32_t *idx, const float *val, float w0, const float *w,
                    const float *V, float *pred)
{
    for (int64_t i = 0; i < n_rows; i++)
        pred[i] = fm_score(c, rowptr, idx, val, i, w0, w, V, NULL);
}

/* SGD or AdaGrad on one FM parameter with its own L2 strength. */
static inline void fm_step(const hmo_config *c, float eta, float g, float lam, float *th, float *acc)
{
    float v = *th;
    g = g + lam * v;
    if (c->opt == HMO_OPT_ADAGRAD) {
        float G = *acc + g * g;
        *acc = G;
        v = v - (eta * g) / (sqrtf(G) + c->eps);
    } else {
        v = v - eta * g;
    }
    *th = v;
}

int hmo_train_fm(const hmo_config *c, int64_t n_rows, const int64_t *rowptr,
                 const uint32_t *idx, const float *val, const float *label,
                 float *w0, float *w, float *V, float *gw0, float *gw, float *gv,
                 uint8_t *touched, int64_t *t, float *preds, double *loss_sum)
{
    const int k = c->factors;
    const int64_t B = c->mini_batch;
    if (B < 1 || k < 1) return -1;
    /* per-feature fp64 sums over the batch (c1 = dl*x in fp32):
     * A1 = sum c1, A2 = sum (c1*x), GS[f] = sum (c1*S32[row][f]) */
    double *A1 = (double *)calloc((size_t)c->dims, sizeof(double));
    double *A2 = (double *)calloc((size_t)c->dims, sizeof(double));
    uint32_t *cnt = (uint32_t *)calloc((size_t)c->dims, sizeof(uint32_t));
    int64_t max_touch = 1;
    for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
        int64_t e = b0 + B < n_rows ? b0 + B : n_rows;
        if (rowptr[e] - rowptr[b0] > max_touch) max_touch = rowptr[e] - rowptr[b0];
    }
    uint32_t *touch = (uint32_t *)malloc((size_t)max_touch * sizeof(uint32_t));
    uint32_t *slot = (uint32_t *)malloc((size_t)c->dims * sizeof(uint32_t));
    double *GS = (double *)malloc((size_t)max_touch * k * sizeof(double));
    float *dl = (float *)malloc((size_t)B * sizeof(float));
    float *S = (float *)malloc((size_t)B * k * sizeof(float));
    if (!A1 || !A2 || !cnt || !touch || !slot || !GS || !dl || !S) return -2;
    float dummy = 0.0f;
    double lsum = 0.0;
    for (int32_t it = 0; it < c->iters; it++) {
        lsum = 0.0;
        for (int64_t b0 = 0; b0 < n_rows; b0 += B) {
            const int64_t b1 = b0 + B < n_rows ? b0 + B : n_rows;
            const int64_t nb = b1 - b0;
            const double inv_nb = 1.0 / (double)nb;
            double g0 = 0.0;
            for (int64_t i = b0; i < b1; i++) {
                float p = fm_score(c, rowptr, idx, val, i, *w0, w, V, S + (i - b0) * k);
                float d;
                if (c->fm_classification) {
                    float y = label[i] > 0.0f ? 1.0f : -1.0f;
                    d = hmo_dloss(HMO_LOSS_LOG, p, y);
                    lsum += hmo_loss(HMO_LOSS_LOG, p, y);
                } else {
                    d = p - label[i];
                    lsum += hmo_loss(HMO_LOSS_SQUARED, p, label[i]);
                }
                dl[i - b0] = d;
                g0 += (double)d;
                if (preds) preds[i] = p;
            }
            int64_t nt = 0;
            for (int64_t i = b0; i < b1; i++)
                for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
                    uint32_t j = idx[p];
                    if ((int64_t)j >= c->dims) return -3;
                    if (cnt[j]++ == 0) {
                        slot[j] = (uint32_t)nt;
                        for (int f = 0; f < k; f++) GS[nt * k + f] = 0.0;
                        touch[nt++] = j;
                    }
                    const float c1 = dl[i - b0] * val[p];
                    A1[j] += (double)c1;
                    A2[j] += (double)(c1 * val[p]);
                    double *gs = GS + (int64_t)slot[j] * k;
                    const float *Si = S + (i - b0) * k;
                    for (int f = 0; f < k; f++) gs[f] += (double)(c1 * Si[f]);
                }
            *t += 1;
            const float eta = hmo_eta(c, *t);
            fm_step(c, eta, c->grad_norm == HMO_NORM_SUM ? (float)g0 : (float)(g0 * inv_nb), c->lambda_w0, w0,
                    gw0 ? gw0 : &dummy);
            for (int64_t q = 0; q < nt; q++) {
                uint32_t j = touch[q];
                fm_step(c, eta, mean_grad(c, A1[j], cnt[j], inv_nb), c->lambda_w, &w[j], gw ? &gw[j] : &dummy);
                for (int f = 0; f < k; f++) {
                    float *v = &V[(int64_t)j * k + f];
                    double G = GS[q * k + f] - (double)*v * A2[j];
                    fm_step(c, eta, mean_grad(c, G, cnt[j], inv_nb), c->lambda_v, v,
                            gv ? &gv[(int64_t)j * k + f] : &dummy);
                }
                if (touched) touched[j] = 1;
                A1[j] = 0.0; A2[j] = 0.0; cnt[j] = 0;
            }
        }
    }
    if (loss_sum) *loss_sum = lsum;
    free(A1); free(A2); free(cnt); free(touch); free(slot); free(GS); free(dl); free(S);
    return 0;
}
