// hm_options.cpp -- UDTF option-string parser.
//
// Stands in for the option handling of the train_classifier / train_regressor /
// train_fm UDTFs that BASELINE.json names ("keeping the UDTF option strings").
// The reference mount has no source (README.md:20-22), so the grammar below is this
// library's own: whitespace-separated "-name value" pairs and bare "-flag"s, names
// case-sensitive, enum values case-insensitive.  The option names that BASELINE.json
// itself spells out (-loss logloss, -opt AdaGrad, -mini_batch 128) are accepted
// verbatim.  Anything unknown is an error, never ignored.
#include "hm_internal.h"
#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

namespace hm {

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const char *last_error_cstr() { return g_err; }

static std::string lower(std::string s)
{
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    return s;
}

static bool to_f(const std::string &v, double *out)
{
    errno = 0;
    char *end = nullptr;
    double d = std::strtod(v.c_str(), &end);
    if (end == v.c_str() || *end != '\0' || errno == ERANGE || !std::isfinite(d)) return false;
    *out = d;
    return true;
}

static bool to_i(const std::string &v, int64_t *out)
{
    errno = 0;
    char *end = nullptr;
    long long d = std::strtoll(v.c_str(), &end, 10);
    if (end == v.c_str() || *end != '\0' || errno == ERANGE) return false;
    *out = (int64_t)d;
    return true;
}

// Inverse of the standard normal distribution function (P. J. Acklam's rational approximation, relative
// error < 1.2e-9), in double: the confidence-weighted learners' -eta (a probability in (0.5, 1)) becomes
// phi = probit(eta).  tests/_oracle.py restates it with the same operations.
static double probit(double p)
{
    static const double a[6] = {-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02,
                                1.383577518672690e+02, -3.066479806614716e+01, 2.506628277459239e+00};
    static const double b[5] = {-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02,
                                6.680131188771972e+01, -1.328068155288572e+01};
    static const double c[6] = {-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00,
                                -2.549732539343734e+00, 4.374664141464968e+00, 2.938163982698783e+00};
    static const double d[4] = {7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00,
                                3.754408661907416e+00};
    const double plow = 0.02425, phigh = 1.0 - plow;
    if (p < plow) {
        const double q = std::sqrt(-2.0 * std::log(p));
        return (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0);
    }
    if (p > phigh) {
        const double q = std::sqrt(-2.0 * std::log(1.0 - p));
        return -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0);
    }
    const double q = p - 0.5, r = q * q;
    return (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q /
           (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1.0);
}

static void defaults(hm_config *c, int task)
{
    std::memset(c, 0, sizeof(*c));
    c->task = task;
    c->loss = task == HM_TASK_REGRESSOR ? HM_LOSS_SQUARED : HM_LOSS_LOG;
    c->opt = task == HM_TASK_FM ? HM_OPT_SGD : HM_OPT_ADAGRAD;
    c->eta_sched = HM_ETA_FIXED;
    c->reg = HM_REG_NONE;
    c->grad_norm = HM_NORM_BATCH;
    c->mini_batch = 1;
    c->iters = 1;
    c->dims = (int64_t)1 << 24;
    c->total_steps = 10000.0;
    c->eta0 = task == HM_TASK_FM ? 0.05f : 0.1f;
    c->power_t = 0.1f;
    c->lambda = 1e-4f;
    c->eps = 1e-6f;
    c->rho = 0.95f;
    c->beta1 = 0.9f;
    c->beta2 = 0.999f;
    c->alpha = 0.1f;
    c->beta = 1.0f;
    c->lambda1 = 0.001f;
    c->lambda2 = 0.0f;
    c->factors = 5;
    c->fm_classification = 0;
    c->sigma = 0.1f;
    c->lambda_w0 = 0.01f;
    c->lambda_w = 0.01f;
    c->lambda_v = 0.01f;
    c->seed = 31;
    c->mix_interval = 0;
    c->hash_seed = 0x9747b28cu;
    c->pa_c = 1.0f;
    c->pa_epsilon = 0.1f;
    c->cov_r = 0.1f;
    c->cov_phi = 1.0f;
    c->loss_param = 0.0f;
}

hm_status parse_options(const char *function_name, const char *options, hm_config *out)
{
    if (!function_name || !out) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "function_name and out must not be NULL");
    std::string fn = function_name;
    int task;
    int online = -1;        // the loss id of an online learner (its function fixes loss and optimizer)
    if (fn == "train_classifier") task = HM_TASK_CLASSIFIER;
    else if (fn == "train_regressor") task = HM_TASK_REGRESSOR;
    else if (fn == "train_fm") task = HM_TASK_FM;
    else if (fn == "train_perceptron") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_PERCEPTRON; }
    else if (fn == "train_pa") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_PA; }
    else if (fn == "train_pa1") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_PA1; }
    else if (fn == "train_pa2") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_PA2; }
    else if (fn == "train_pa1_regr") { task = HM_TASK_REGRESSOR; online = HM_LOSS_PA1_REGR; }
    else if (fn == "train_pa2_regr") { task = HM_TASK_REGRESSOR; online = HM_LOSS_PA2_REGR; }
    else if (fn == "train_arow") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_AROW; }
    else if (fn == "train_arow_regr") { task = HM_TASK_REGRESSOR; online = HM_LOSS_AROW_REGR; }
    else if (fn == "train_arowe_regr") { task = HM_TASK_REGRESSOR; online = HM_LOSS_AROWE_REGR; }
    else if (fn == "train_cw") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_CW; }
    else if (fn == "train_scw") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_SCW1; }
    else if (fn == "train_scw2") { task = HM_TASK_CLASSIFIER; online = HM_LOSS_SCW2; }
    else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "unknown function '%s' (expected train_classifier, train_regressor, train_fm, "
                 "train_perceptron, train_pa, train_pa1, train_pa2, train_pa1_regr, train_pa2_regr, train_arow, train_arow_regr, "
                 "train_arowe_regr, train_cw, train_scw or train_scw2)", function_name);
    hm_config c;
    defaults(&c, task);
    if (online >= 0) {
        // the step size lives in the per-row coefficient: the "optimizer" only subtracts the mean of coef * x
        c.loss = online; c.opt = HM_OPT_SGD; c.eta_sched = HM_ETA_FIXED; c.eta0 = 1.0f; c.reg = HM_REG_NONE;
        if (online >= HM_LOSS_AROW) c.opt = HM_OPT_COV;
    }
    const bool takes_r = online >= HM_LOSS_AROW && online <= HM_LOSS_AROWE_REGR;
    const bool takes_phi = online >= HM_LOSS_CW;
    const bool takes_c = online == HM_LOSS_PA1 || online == HM_LOSS_PA2 || online == HM_LOSS_PA1_REGR || online == HM_LOSS_PA2_REGR ||
                         online == HM_LOSS_SCW1 || online == HM_LOSS_SCW2;
    const bool takes_eps = online == HM_LOSS_PA1_REGR || online == HM_LOSS_PA2_REGR || online == HM_LOSS_AROWE_REGR;

    std::vector<std::string> tok;
    {
        std::istringstream is(options ? options : "");
        std::string t;
        while (is >> t) tok.push_back(t);
    }
    bool loss_given = false;
    float huber_c = 1.0f, quantile = 0.5f;
    for (size_t i = 0; i < tok.size(); i++) {
        const std::string &name = tok[i];
        if (name.size() < 2 || name[0] != '-')
            HM_FAIL(HM_ERR_INVALID_ARGUMENT, "expected an option starting with '-', got '%s'", name.c_str());
        // ---- flags
        if (name == "-classification" || (name == "-c" && online < 0)) {
            if (task != HM_TASK_FM) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "%s is a train_fm option", name.c_str());
            c.fm_classification = 1;
            continue;
        }
        if (name == "-dense" || name == "-densemodel") continue;  // the model replica is always dense
        // ---- valued options
        if (i + 1 >= tok.size()) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "option '%s' needs a value", name.c_str());
        const std::string &v = tok[++i];
        const std::string lv = lower(v);
        double d = 0;
        int64_t n = 0;
        auto need_f = [&](double lo, double hi) -> bool {
            if (!to_f(v, &d) || d < lo || d > hi) {
                set_error("option '%s': '%s' is not a number in [%g, %g]", name.c_str(), v.c_str(), lo, hi);
                return false;
            }
            return true;
        };
        auto need_i = [&](int64_t lo, int64_t hi) -> bool {
            if (!to_i(v, &n) || n < lo || n > hi) {
                set_error("option '%s': '%s' is not an integer in [%lld, %lld]", name.c_str(), v.c_str(),
                          (long long)lo, (long long)hi);
                return false;
            }
            return true;
        };
#define NEED_F(lo, hi) if (!need_f((lo), (hi))) return HM_ERR_INVALID_ARGUMENT
#define NEED_I(lo, hi) if (!need_i((lo), (hi))) return HM_ERR_INVALID_ARGUMENT
        if (takes_phi && name == "-eta") {       // the confidence as a probability: phi = probit(eta)
            NEED_F(0.5000001, 0.9999999); c.cov_phi = (float)probit(d);
            continue;
        }
        if (online >= 0 && (name == "-loss" || name == "-loss_function" || name == "-opt" || name == "-optimizer" ||
                            name == "-eta" || name == "-eta0" || name == "-eta_0" || name == "-power_t" ||
                            name == "-total_steps" || name == "-t" || name == "-reg" || name == "-regularization" ||
                            name == "-lambda"))
            HM_FAIL(HM_ERR_INVALID_ARGUMENT, "%s is fixed by %s", name.c_str(), function_name);
        if (name == "-c" || name == "-aggressiveness") {
            if (!takes_c) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "%s is an option of train_pa1, train_pa2, their _regr forms, train_scw and train_scw2", name.c_str());
            NEED_F(1e-12, 1e12); c.pa_c = (float)d;
        } else if (name == "-epsilon" || name == "-e") {
            if (!takes_eps && !(online < 0 && task == HM_TASK_REGRESSOR))
                HM_FAIL(HM_ERR_INVALID_ARGUMENT, "%s is an option of train_pa1_regr, train_pa2_regr, train_arowe_regr and of train_regressor "
                        "with an epsilon-insensitive loss", name.c_str());
            NEED_F(0.0, 1e12); c.pa_epsilon = (float)d;
        } else if (name == "-huber_c") {
            if (online >= 0 || task == HM_TASK_FM) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "-huber_c goes with -loss huber");
            NEED_F(1e-12, 1e12); huber_c = (float)d;
        } else if (name == "-quantile" || name == "-tau") {
            if (online >= 0 || task == HM_TASK_FM) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "%s goes with -loss quantile", name.c_str());
            NEED_F(0.0, 1.0); quantile = (float)d;
        } else if (name == "-phi") {
            if (!takes_phi) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "-phi is an option of train_cw, train_scw and train_scw2");
            NEED_F(1e-6, 1e6); c.cov_phi = (float)d;
        } else if (name == "-r") {
            if (!takes_r) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "-r is an option of train_arow, train_arow_regr and train_arowe_regr");
            NEED_F(1e-12, 1e12); c.cov_r = (float)d;
        } else if (name == "-loss" || name == "-loss_function") {
            if (task == HM_TASK_FM) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "-loss is not a train_fm option (use -classification)");
            if (lv == "logloss" || lv == "logistic" || lv == "log") c.loss = HM_LOSS_LOG;
            else if (lv == "hinge" || lv == "hingeloss") c.loss = HM_LOSS_HINGE;
            else if (lv == "squared" || lv == "squaredloss" || lv == "squared_loss") c.loss = HM_LOSS_SQUARED;
            else if (lv == "squared_hinge" || lv == "squaredhinge" || lv == "squaredhingeloss") c.loss = HM_LOSS_SQUARED_HINGE;
            else if (lv == "modified_huber" || lv == "modifiedhuber" || lv == "modifiedhuberloss") c.loss = HM_LOSS_MODIFIED_HUBER;
            else if (lv == "huber" || lv == "huberloss") c.loss = HM_LOSS_HUBER;
            else if (lv == "quantile" || lv == "quantileloss") c.loss = HM_LOSS_QUANTILE;
            else if (lv == "epsilon_insensitive" || lv == "epsiloninsensitive" || lv == "epsiloninsensitiveloss") c.loss = HM_LOSS_EPS_INSENSITIVE;
            else if (lv == "squared_epsilon_insensitive" || lv == "squaredepsiloninsensitive" || lv == "squaredepsiloninsensitiveloss")
                c.loss = HM_LOSS_SQUARED_EPS_INSENSITIVE;
            else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "option -loss: unknown loss '%s' (logloss, hinge, squared_hinge, modified_huber, "
                         "squared, huber, quantile, epsilon_insensitive, squared_epsilon_insensitive)", v.c_str());
            loss_given = true;
        } else if (name == "-opt" || name == "-optimizer") {
            if (lv == "sgd") c.opt = HM_OPT_SGD;
            else if (lv == "adagrad") c.opt = HM_OPT_ADAGRAD;
            else if (lv == "adadelta") c.opt = HM_OPT_ADADELTA;
            else if (lv == "adam") c.opt = HM_OPT_ADAM;
            else if (lv == "ftrl") c.opt = HM_OPT_FTRL;
            else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "option -opt: unknown optimizer '%s' (sgd, adagrad, adadelta, adam, ftrl)", v.c_str());
        } else if (name == "-mini_batch" || name == "-mini_batch_size") {
            NEED_I(1, (int64_t)1 << 26); c.mini_batch = (int32_t)n;
        } else if (name == "-iters" || name == "-iter" || name == "-iterations") {
            NEED_I(1, 1000000); c.iters = (int32_t)n;
        } else if (name == "-dims" || name == "-feature_dimensions" || name == "-p" || name == "-num_features") {
            NEED_I(1, (int64_t)1 << 31); c.dims = n;
        } else if (name == "-eta0" || name == "-eta_0") {
            NEED_F(0.0, 1e6); c.eta0 = (float)d;
        } else if (name == "-eta") {
            if (lv == "fixed") c.eta_sched = HM_ETA_FIXED;
            else if (lv == "inverse" || lv == "inv" || lv == "invscaling") c.eta_sched = HM_ETA_INVERSE;
            else if (lv == "simple") c.eta_sched = HM_ETA_SIMPLE;
            else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "option -eta: unknown schedule '%s' (fixed, inverse, simple)", v.c_str());
        } else if (name == "-power_t") {
            NEED_F(0.0, 10.0); c.power_t = (float)d;
        } else if (name == "-total_steps" || name == "-t") {
            NEED_F(1.0, 1e18); c.total_steps = d;
        } else if (name == "-reg" || name == "-regularization") {
            if (lv == "no" || lv == "none") c.reg = HM_REG_NONE;
            else if (lv == "l1") c.reg = HM_REG_L1;
            else if (lv == "l2") c.reg = HM_REG_L2;
            else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "option -reg: unknown regularization '%s' (no, l1, l2)", v.c_str());
        } else if (name == "-lambda") {
            NEED_F(0.0, 1e6); c.lambda = (float)d;
        } else if (name == "-eps") {
            NEED_F(0.0, 1.0); c.eps = (float)d;
        } else if (name == "-rho" || name == "-decay") {
            NEED_F(0.0, 1.0); c.rho = (float)d;
        } else if (name == "-beta1") {
            NEED_F(0.0, 0.999999); c.beta1 = (float)d;
        } else if (name == "-beta2") {
            NEED_F(0.0, 0.9999999); c.beta2 = (float)d;
        } else if (name == "-alpha") {
            NEED_F(1e-12, 1e6); c.alpha = (float)d;
        } else if (name == "-beta") {
            NEED_F(0.0, 1e6); c.beta = (float)d;
        } else if (name == "-lambda1" || name == "-l1") {
            NEED_F(0.0, 1e6); c.lambda1 = (float)d;
        } else if (name == "-lambda2" || name == "-l2") {
            NEED_F(0.0, 1e6); c.lambda2 = (float)d;
        } else if (name == "-grad_norm") {
            if (lv == "batch") c.grad_norm = HM_NORM_BATCH;
            else if (lv == "count") c.grad_norm = HM_NORM_COUNT;
            else if (lv == "sum") c.grad_norm = HM_NORM_SUM;
            else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "option -grad_norm: unknown mode '%s' (batch, count, sum)", v.c_str());
        } else if (name == "-factors" || name == "-factor" || name == "-f" || name == "-k") {
            if (task != HM_TASK_FM) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "%s is a train_fm option", name.c_str());
            NEED_I(1, 256); c.factors = (int32_t)n;
        } else if (name == "-sigma") {
            NEED_F(0.0, 1e3); c.sigma = (float)d;
        } else if (name == "-lambda0" || name == "-lambda_w0") {
            NEED_F(0.0, 1e6); c.lambda_w0 = (float)d;
        } else if (name == "-lambdaW" || name == "-lambda_w") {
            NEED_F(0.0, 1e6); c.lambda_w = (float)d;
        } else if (name == "-lambdaV" || name == "-lambda_v") {
            NEED_F(0.0, 1e6); c.lambda_v = (float)d;
        } else if (name == "-seed") {
            NEED_I(0, 0xffffffffLL); c.seed = (uint32_t)n;
        } else if (name == "-hash_seed") {
            NEED_I(0, 0xffffffffLL); c.hash_seed = (uint32_t)n;
        } else if (name == "-mix_interval") {
            NEED_I(0, (int64_t)1 << 30); c.mix_interval = (int32_t)n;
        } else {
            HM_FAIL(HM_ERR_INVALID_ARGUMENT, "unknown option '%s' for %s", name.c_str(), function_name);
        }
#undef NEED_F
#undef NEED_I
    }
    // cross-option checks
    if (task == HM_TASK_CLASSIFIER && c.loss == HM_LOSS_SQUARED && loss_given)
        ;  // allowed: squared loss on +-1 targets
    const bool regr_loss = c.loss == HM_LOSS_SQUARED || (c.loss >= HM_LOSS_HUBER && c.loss <= HM_LOSS_SQUARED_EPS_INSENSITIVE);
    if (task == HM_TASK_REGRESSOR && online < 0 && !regr_loss)
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "train_regressor takes -loss squared, huber, quantile, epsilon_insensitive or squared_epsilon_insensitive");
    if (task == HM_TASK_CLASSIFIER && online < 0 && c.loss >= HM_LOSS_HUBER)
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "train_classifier takes -loss logloss, hinge, squared_hinge, modified_huber or squared");
    if (online < 0) c.loss_param = c.loss == HM_LOSS_HUBER ? huber_c : (c.loss == HM_LOSS_QUANTILE ? quantile : 0.0f);
    if (task == HM_TASK_FM && c.opt != HM_OPT_SGD && c.opt != HM_OPT_ADAGRAD)
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "train_fm supports -opt sgd or adagrad");
    if (task == HM_TASK_FM) c.loss = c.fm_classification ? HM_LOSS_LOG : HM_LOSS_SQUARED;
    *out = c;
    return HM_OK;
}

}  // namespace hm
