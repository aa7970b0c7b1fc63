// hm_table.cpp -- the model table as a file, and warm start from one.
//
// Stands in for what a Hive job does with the rows a training UDTF forwards -- INSERT them into a
// table (feature, weight[, Vif]) -- and for starting a later job from that table.  The reference
// mount has no source (README.md:20-22), so the layout is this library's own and its separators
// are parameters: text = one row per line, "feature <delim> weight [<delim> v1,v2,...,vk]" with
// weights printed so that they read back bit for bit (%.9g), which is what Hive's text SerDe
// holds for (int, float[, array<float>]) with delimiter \t and collection separator ','; binary =
// a small header followed by the three arrays, little endian.
//
// Host code only: it moves rows between a file and the entry points of hm_learner.cu.
#include "hm_internal.h"
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace hm {
hm_status learner_info(const hm_udtf *u, int *is_fm, int *factors, int64_t *dims, int *opt);
hm_status learner_set_touched(hm_udtf *u, const int32_t *feature, int64_t n);
}

namespace {

struct BinHeader {
    char magic[8];          // "HMTABLE1"
    int32_t factors;        // 0: linear learner
    int32_t reserved;
    int64_t dims;
    int64_t rows;
};

struct Table {
    int32_t factors = 0;
    int64_t dims = 0;
    std::vector<int32_t> feature;
    std::vector<float> weight;
    std::vector<float> vif;     // rows x factors
};

hm_status write_table(const char *path, int format, char delim, const Table &t)
{
    FILE *f = std::fopen(path, format == HM_TABLE_BINARY ? "wb" : "w");
    if (!f) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "cannot open '%s' for writing: %s", path, std::strerror(errno));
    bool ok = true;
    const size_t n = t.feature.size();
    if (format == HM_TABLE_BINARY) {
        BinHeader h;
        std::memcpy(h.magic, "HMTABLE1", 8);
        h.factors = t.factors; h.reserved = 0; h.dims = t.dims; h.rows = (int64_t)n;
        ok = std::fwrite(&h, sizeof(h), 1, f) == 1;
        if (n) {
            ok = ok && std::fwrite(t.feature.data(), 4, n, f) == n && std::fwrite(t.weight.data(), 4, n, f) == n;
            if (t.factors) ok = ok && std::fwrite(t.vif.data(), 4, n * t.factors, f) == n * t.factors;
        }
    } else {
        for (size_t i = 0; i < n && ok; i++) {
            ok = std::fprintf(f, "%d%c%.9g", t.feature[i], delim, (double)t.weight[i]) > 0;
            if (t.factors) {
                std::fputc(delim, f);
                for (int q = 0; q < t.factors; q++)
                    ok = ok && std::fprintf(f, q ? ",%.9g" : "%.9g", (double)t.vif[i * t.factors + q]) > 0;
            }
            std::fputc('\n', f);
        }
    }
    ok = std::fclose(f) == 0 && ok;
    if (!ok) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "short write to '%s'", path);
    return HM_OK;
}

hm_status read_table(const char *path, int format, char delim, int32_t want_factors, Table *t)
{
    FILE *f = std::fopen(path, format == HM_TABLE_BINARY ? "rb" : "r");
    if (!f) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "cannot open '%s': %s", path, std::strerror(errno));
    hm_status st = HM_OK;
    if (format == HM_TABLE_BINARY) {
        BinHeader h;
        if (std::fread(&h, sizeof(h), 1, f) != 1 || std::memcmp(h.magic, "HMTABLE1", 8) != 0 || h.rows < 0 || h.factors < 0) {
            hm::set_error("'%s' is not a binary model table", path);
            st = HM_ERR_INVALID_ARGUMENT;
        } else {
            t->factors = h.factors; t->dims = h.dims;
            const size_t n = (size_t)h.rows;
            t->feature.resize(n); t->weight.resize(n); t->vif.resize(n * h.factors);
            bool ok = !n || (std::fread(t->feature.data(), 4, n, f) == n && std::fread(t->weight.data(), 4, n, f) == n);
            if (ok && n && h.factors) ok = std::fread(t->vif.data(), 4, n * h.factors, f) == n * h.factors;
            if (!ok) { hm::set_error("'%s' is truncated", path); st = HM_ERR_INVALID_ARGUMENT; }
        }
    } else {
        t->factors = want_factors;
        std::string line;
        char buf[1 << 16];
        int64_t lineno = 0;
        while (st == HM_OK && std::fgets(buf, sizeof(buf), f)) {
            line = buf;
            while (!line.empty() && line.back() != '\n' && std::fgets(buf, sizeof(buf), f)) line += buf;
            lineno++;
            while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
            if (line.empty()) continue;
            const char *p = line.c_str();
            char *end = nullptr;
            errno = 0;
            const long long feat = std::strtoll(p, &end, 10);
            bool ok = end != p && *end == delim && errno == 0 && feat >= -1 && feat <= 0x7fffffffLL;
            float w = 0.f;
            if (ok) { p = end + 1; w = std::strtof(p, &end); ok = end != p; }
            std::vector<float> v;
            if (ok && want_factors) {
                ok = *end == delim;
                p = end + 1;
                for (int q = 0; ok && q < want_factors; q++) {
                    v.push_back(std::strtof(p, &end));
                    ok = end != p && (q + 1 == want_factors ? *end == '\0' : *end == ',');
                    p = end + 1;
                }
            } else if (ok) ok = *end == '\0';
            if (!ok) {
                hm::set_error("'%s' line %lld is not a model-table row (%s)", path, (long long)lineno,
                              want_factors ? "feature, weight, factors" : "feature, weight");
                st = HM_ERR_INVALID_ARGUMENT;
                break;
            }
            t->feature.push_back((int32_t)feat); t->weight.push_back(w);
            t->vif.insert(t->vif.end(), v.begin(), v.end());
        }
    }
    std::fclose(f);
    return st;
}

}  // namespace

extern "C" {

hm_status hm_udtf_save_model_table(hm_udtf *u, const char *path, int format, char delimiter, int64_t *n_rows)
{
    if (!u || !path) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    if (format != HM_TABLE_TEXT && format != HM_TABLE_BINARY) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "unknown model-table format %d", format);
    int is_fm = 0, k = 0, opt = 0; int64_t dims = 0;
    HM_TRY(hm::learner_info(u, &is_fm, &k, &dims, &opt));
    int64_t n = 0;
    HM_TRY(hm_udtf_close(u, &n));
    Table t;
    t.factors = is_fm ? k : 0; t.dims = dims;
    t.feature.resize((size_t)n); t.weight.resize((size_t)n); t.vif.resize((size_t)n * t.factors);
    int64_t got = 0;
    HM_TRY(hm_udtf_forward(u, t.feature.data(), t.weight.data(), t.factors ? t.vif.data() : nullptr, n, &got));
    t.feature.resize((size_t)got); t.weight.resize((size_t)got); t.vif.resize((size_t)got * t.factors);
    if (n_rows) *n_rows = got;
    return write_table(path, format, delimiter ? delimiter : '\t', t);
}

hm_status hm_udtf_load_model_table(hm_udtf *u, const char *path, int format, char delimiter, int64_t *n_rows)
{
    if (!u || !path) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    if (format != HM_TABLE_TEXT && format != HM_TABLE_BINARY) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "unknown model-table format %d", format);
    int is_fm = 0, k = 0, opt = 0; int64_t dims = 0;
    HM_TRY(hm::learner_info(u, &is_fm, &k, &dims, &opt));
    Table t;
    HM_TRY(read_table(path, format, delimiter ? delimiter : '\t', is_fm ? k : 0, &t));
    if (t.factors != (is_fm ? k : 0))
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "'%s' holds %d factors per row, the learner has %d", path, t.factors, is_fm ? k : 0);
    const size_t n = t.feature.size();
    for (size_t i = 0; i < n; i++)
        if (t.feature[i] >= dims || (t.feature[i] < 0 && !(is_fm && t.feature[i] == -1)))
            HM_FAIL(HM_ERR_INVALID_ARGUMENT, "'%s': feature %d does not fit dims = %lld", path, t.feature[i], (long long)dims);
    std::vector<float> w((size_t)dims, 0.f);
    if (!is_fm) {
        if (opt == HM_OPT_FTRL) {
            // FTRL keeps (z, n) and derives the weight.  With n = 0 the closed form is
            // w = -(z - sgn(z) l1) / (beta/alpha + l2), so z = -w (beta/alpha + l2) - sgn(w) l1 reproduces the
            // table's weights (to rounding); the accumulated n of the run that wrote the table is not in it.
            hm_config cfg;
            HM_TRY(hm_udtf_config(u, &cfg));
            std::vector<float> zn((size_t)dims * 2, 0.f);
            const float scale = cfg.beta / cfg.alpha + cfg.lambda2;
            for (size_t i = 0; i < n; i++) {
                const float wi = t.weight[i];
                if (wi != 0.f) zn[(size_t)t.feature[i] * 2] = -wi * scale - (wi > 0.f ? cfg.lambda1 : -cfg.lambda1);
            }
            HM_TRY(hm_udtf_write_array(u, 1, zn.data(), dims * 2));
        } else {
            for (size_t i = 0; i < n; i++) w[(size_t)t.feature[i]] = t.weight[i];
            HM_TRY(hm_udtf_write_array(u, 0, w.data(), dims));
        }
    } else {
        // factors of features that are not in the table keep their initial values
        std::vector<float> V((size_t)dims * k);
        HM_TRY(hm_udtf_read_array(u, 1, V.data(), dims * k));
        std::vector<int32_t> seen;
        for (size_t i = 0; i < n; i++) {
            if (t.feature[i] == -1) { HM_TRY(hm_udtf_write_array(u, 2, &t.weight[i], 1)); continue; }
            w[(size_t)t.feature[i]] = t.weight[i];
            std::memcpy(&V[(size_t)t.feature[i] * k], &t.vif[i * k], (size_t)k * 4);
            seen.push_back(t.feature[i]);
        }
        HM_TRY(hm_udtf_write_array(u, 0, w.data(), dims));
        HM_TRY(hm_udtf_write_array(u, 1, V.data(), dims * k));
        HM_TRY(hm::learner_set_touched(u, seen.data(), (int64_t)seen.size()));
    }
    if (n_rows) *n_rows = (int64_t)n;
    return HM_OK;
}

}  // extern "C"
