// hm_internal.h -- shared declarations of the hivemall_b200 CUDA library (not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/hivemall_b200.h"

namespace hm {

// ---- errors -----------------------------------------------------------------
void set_error(const char *fmt, ...);
#define HM_FAIL(code, ...) do { ::hm::set_error(__VA_ARGS__); return (code); } while (0)
#define HM_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    ::hm::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    return e_ == cudaErrorMemoryAllocation ? HM_ERR_OOM : HM_ERR_CUDA; } } while (0)
#define HM_TRY(expr) do { hm_status s_ = (expr); if (s_ != HM_OK) return s_; } while (0)

// ---- kernel ids for launch accounting -----------------------------------------
enum KernelId {
    K_HASH = 0, K_SEG_PREP, K_SORT_HIST, K_SORT_SCAN, K_SORT_SCATTER,
    K_LIN_FORWARD, K_LIN_UPDATE, K_FM_FORWARD, K_FM_UPDATE, K_PREDICT,
    K_EXPORT, K_MISC, K_HOT_SELECT, K_NUM_KERNELS
};
extern const char *const kKernelNames[K_NUM_KERNELS];

// Launch bookkeeping: counts every launch; when profiling is on, brackets each
// launch with events on the launching stream and accumulates per-kernel time.
struct Launcher {
    cudaStream_t stream = nullptr;
    bool profiling = false;
    int64_t launches = 0;
    int64_t count[K_NUM_KERNELS] = {0};
    double ms[K_NUM_KERNELS] = {0};
    double bytes[K_NUM_KERNELS] = {0};
    struct Pending { int id; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> pool;

    void begin(int id, double alg_bytes);
    void end(int id);
    void drain();      // resolve pending event pairs into ms[] (synchronises)
    void release();
};

#define HM_LAUNCH(L, id, alg_bytes, ...) do { (L).begin((id), (alg_bytes)); __VA_ARGS__; (L).end(id); } while (0)

// SMs of the current device (cudaDevAttrMultiProcessorCount, cached per device); grids of the
// persistent kernels are sized in multiples of it.
int num_sms();

// ---- hot features (hm_sort.cu: hot_select_kernel) -----------------------------------
// A handful of features carry a large share of a mini-batch's elements: the quantitative features
// ("name:value" with a fixed name) that every row has, and the head of each categorical field.
// Sorting them is wasted work -- their gradient is a long plain sum -- and so is gathering their
// weight.  Per ingest block up to kMaxHot features are picked by their frequency in a sample of
// the block's rows.  Their elements never enter the sort (dropped by KEY in the first pass and its
// histogram); the forward kernel reads their weights from shared memory and accumulates their
// gradient sums itself, in a fixed order; extra CTAs of the update kernel step them.  The choice
// only affects speed: whatever set is picked, every occurrence of a picked feature goes one way
// and every other element the other, so no property of the rows has to be verified.
#ifndef HM_MAX_HOT
#define HM_MAX_HOT 64
#endif
constexpr int kMaxHot = HM_MAX_HOT;
// The kernels test "is this feature hot?" once per element, so the test is one multiply, one shift, one
// shared-memory load and one compare: a direct-mapped table, index = hot_index(feature), entry = the
// feature.  The selection kernel only picks features whose indices are pairwise different.
constexpr int kHotSlotBits = 11;
constexpr int kHotSlots = 1 << kHotSlotBits;
constexpr uint32_t kHotEmpty = 0xffffffffu;
__host__ __device__ inline uint32_t hot_index(uint32_t key) { return (key * 0x9E3779B1u) >> (32 - kHotSlotBits); }
struct HotInfo {
    uint32_t count;                    // hot features of this block (0: none)
    uint32_t pad[3];
    uint32_t key[kMaxHot];             // ordered by (sample count descending, feature ascending)
};

// ---- segmented radix sort (hm_sort.cu) ----------------------------------------
constexpr int kMaxSegs = 256;          // mini-batches sorted per launch
constexpr int kSortTile = 4096;        // elements per sort tile
struct SortScratch {
    // device
    uint32_t *key[2] = {nullptr, nullptr};   // feature ids, ping-pong
    uint2 *pay[2] = {nullptr, nullptr};      // (row inside the mini-batch, value bits), ping-pong
    uint32_t *tile_hist = nullptr;     // [seg][digit][tile_in_seg]
    int32_t *tile_row0 = nullptr;      // first row (inside its mini-batch) of every tile
    uint32_t *row_sum = nullptr;       // [seg*256 + digit] (+1)
    // Two coordinate systems.  in_*: the block's CSR as it lies in the chunk (what the first pass
    // reads).  seg_*: the SORTED arrays, from which hot-feature and out-of-range elements are
    // gone; written on the device after the first pass's histogram (the host never learns them).
    uint32_t *in_off = nullptr;        // [kMaxSegs+1] element offset of each mini-batch in the block's CSR
    uint32_t *in_tile_off = nullptr;   // [kMaxSegs+1]
    uint32_t *seg_off = nullptr;       // [kMaxSegs+1] element offset of each mini-batch in the sorted arrays
    uint32_t *seg_tile_off = nullptr;  // [kMaxSegs+1]
    HotInfo *hot = nullptr;            // hot features of the block (count 0: none)
    int64_t cap_elems = 0;
    int64_t cap_tiles = 0;
};
hm_status sort_scratch_reserve(SortScratch &s, int64_t elems);
void sort_scratch_free(SortScratch &s);
// Fill seg_off / seg_tile_off for rows [row0, row0+n_rows) split every `mini_batch` rows.
hm_status launch_seg_prep(Launcher &L, SortScratch &s, const int64_t *d_rowptr, int64_t row0,
                          int64_t n_rows, int64_t n_elems, int32_t mini_batch, int n_segs);
// Pick the hot features of rows [row0, row0+n_rows) (s.hot); `enable` false: none.
hm_status launch_hot_select(Launcher &L, SortScratch &s, const int64_t *d_rowptr, int64_t row0, int64_t n_rows,
                            const uint32_t *d_feature, int64_t dims, bool enable);
// Stable sort of each segment by the low `key_bits` bits of the feature id.
// Result: s.key[*out_buf], s.pay[*out_buf] hold (feature, (row, value)) sorted per segment.
hm_status launch_segmented_sort(Launcher &L, SortScratch &s, const uint32_t *d_feature_block,
                                const float *d_value_block, const int64_t *d_rowptr, int64_t block_row0,
                                int64_t block_rows, int32_t mini_batch, int64_t n_elems, int n_segs, int key_bits,
                                int64_t dims, uint32_t *d_bad, int *out_buf, int ctas_per_sm,
                                std::vector<uint32_t> *seg_off_host = nullptr);   // filled while profiling

// ---- hashing (hm_hash.cu) -----------------------------------------------------
hm_status launch_mhash(Launcher &L, const uint8_t *d_bytes, const int64_t *d_offsets, int64_t n,
                       uint32_t seed, int32_t num_features, int32_t *d_out);
hm_status launch_feature_hashing(Launcher &L, const uint8_t *d_bytes, const int64_t *d_offsets, int64_t n,
                                 uint32_t seed, int32_t num_features, int32_t *d_index, float *d_value,
                                 double alg_bytes = 0.0);   // bytes + 8 (offset) + 8 (index, value) per string

// ---- linear learner kernels (hm_linear.cu) ------------------------------------
struct StepScalars {      // per mini-batch scalars computed on the host in double
    float eta;
    float adam_alpha;
    int32_t n_rows;       // rows in this mini-batch
};
// Where a feature's weight is read from.  Plain: base[j * stride].  FTRL keeps only (z, n) per
// feature (8-byte records, a 2^24-dim replica is 128 MiB and largely stays in L2) and the weight
// is the closed form of DESIGN.md 3.6 evaluated on the fly -- bit-identical to what a stored
// weight would hold, because a stored weight is always that same function of the same (z, n).
struct WeightView {
    const float *base;
    int32_t stride;
    int32_t ftrl;         // 1: base holds (z, n) pairs and the weight is derived
    float alpha, beta, lambda1, lambda2;
};
struct LinearArgs {
    hm_config cfg;
    float *rec;           // replica: rec_width floats per feature
    int rec_width;        // 1 (SGD: w), 2 (AdaGrad: w, G; FTRL: z, n), 4 (AdaDelta, Adam: w, s1, s2, unused)
    WeightView wv;
    const float *sqnorm = nullptr;   // |x|^2 per row of the block being trained (online learners), else NULL
    int cov_pass = 0;                // HM_OPT_COV: 1 = the update launch steps the weights, 2 = the variances
};
inline bool loss_keeps_variance(int loss) { return loss >= HM_LOSS_AROW && loss <= HM_LOSS_SCW2; }
// margin, variance and the two per-row coefficients of the variance-keeping learners (AROW)
hm_status launch_linear_forward_cov(Launcher &L, const LinearArgs &a, const int64_t *d_rowptr, int64_t row0, int32_t n_rows,
                                    int64_t block_row0, const uint32_t *d_feature, const float *d_value, const float *d_label,
                                    float *d_margin, float *d_c1, float *d_c2, uint32_t *d_bad, int64_t nnz_hint);
// |x|^2 of rows [row0, row0 + n_rows) of a chunk into out[0 .. n_rows)
hm_status launch_row_sqnorm(Launcher &L, const int64_t *d_rowptr, int64_t row0, int64_t n_rows, const uint32_t *d_feature,
                            const float *d_value, int64_t dims, float *d_out, int64_t nnz_hint);
inline bool loss_needs_sqnorm(int loss) { return loss >= HM_LOSS_PA && loss <= HM_LOSS_PA2_REGR; }
// Hot features of the block being trained, and where their gradient sums travel from the forward
// kernel to the update kernel.  info == NULL: the block was sorted whole.
struct HotCols {
    const HotInfo *info = nullptr;
    double *sum = nullptr;        // [kMaxHot][stride]: per forward CTA, the sum of dloss * value over its rows
    uint32_t *cnt = nullptr;      // [kMaxHot][stride]: ... and the number of occurrences
    int stride = 0;
};
int linear_forward_grid(int32_t n_rows);
// margin (per block row) and dloss (per mini-batch row) for rows [row0, row0+n) of the chunk
hm_status launch_linear_forward(Launcher &L, const LinearArgs &a, const int64_t *d_rowptr, int64_t row0,
                                int32_t n_rows, int64_t block_row0, const uint32_t *d_feature,
                                const float *d_value, const float *d_label, float *d_margin,
                                float *d_dloss, uint32_t *d_bad, int64_t nnz_hint, const HotCols &hc);
// reduce-by-feature over the sorted segment `seg` and apply the optimizer
struct UpdateScratch {
    void *boundary = nullptr;     // per-CTA open-segment summaries
    uint32_t *done_counter = nullptr;
    int grid = 0;
    double *hot_sum = nullptr;         // see HotCols
    uint32_t *hot_cnt = nullptr;
    int hot_stride = 0;
};
hm_status update_scratch_reserve(UpdateScratch &u);
void update_scratch_free(UpdateScratch &u);
hm_status launch_linear_update(Launcher &L, const LinearArgs &a, const StepScalars &sc, UpdateScratch &us,
                               const uint32_t *d_sorted_key, const uint2 *d_sorted_pay,
                               const uint32_t *d_seg_off, int seg, int64_t seg_begin, const float *d_dloss,
                               int64_t nnz_hint, const HotCols &hc, int64_t sorted_elems = -1);   // (byte accounting)
// small mini-batches (<= 1024 rows and <= one update tile of elements): up to 64 consecutive
// mini-batches of a sorted block trained by one persistent CTA
bool small_batch_eligible(int32_t mini_batch, int64_t max_mb_elems);
hm_status launch_linear_small_batches(Launcher &L, const LinearArgs &a, const int64_t *d_rowptr, int64_t block_row0,
                                      int first_seg, int n_segs, const int32_t *rows, const float *eta,
                                      const float *adam_alpha, const uint32_t *d_feature, const float *d_value,
                                      const float *d_label, float *d_margin, float *d_dloss, uint32_t *d_bad,
                                      const uint32_t *d_sorted_key, const uint2 *d_sorted_pay, const uint32_t *d_seg_off,
                                      int64_t nnz_hint);
hm_status launch_predict_linear(Launcher &L, const WeightView &wv, int64_t dims, int64_t n_rows, const int64_t *d_rowptr,
                                const uint32_t *d_feature, const float *d_value, float *d_out, uint32_t *d_bad);
// model table: features with w != 0, ascending
struct ExportScratch { uint32_t *counts = nullptr; int64_t cap = 0; };
hm_status launch_export_count(Launcher &L, ExportScratch &es, const WeightView &wv, const uint8_t *touched,
                              int64_t dims, int64_t *d_total);
hm_status launch_export_write(Launcher &L, ExportScratch &es, const WeightView &wv, const uint8_t *touched,
                              int64_t dims, int32_t *d_feature, float *d_weight);
hm_status launch_scale(Launcher &L, float *p, int64_t n, float scale);
hm_status launch_fill_u32(Launcher &L, uint32_t *p, int64_t n, uint32_t v);
hm_status launch_fill_strided(Launcher &L, float *p, int64_t n, int stride, float v);
hm_status launch_dense_weights(Launcher &L, const WeightView &wv, int64_t dims, float *d_out);
hm_status launch_strided_copy(Launcher &L, float *dense, float *strided, int64_t n, int stride, int off, int width,
                              bool gather);

// ---- factorization machine kernels (hm_fm.cu) ------------------------------------
struct FmArgs {
    hm_config cfg;
    float *w0rec;        // (w0, accumulator)
    float *wrec;         // ww floats per feature: (w[, accumulator])
    float *vrec;         // vw floats per feature: (V_0..V_k-1[, accumulators])
    uint8_t *touched;    // 1 for every feature seen in training
    int ww, vw, k;
};
struct FmScratch {       // pieces of features whose run straddles update chunks
    void *hdr = nullptr;
    double *gs = nullptr;
    int64_t cap_chunks = 0;
    int k = 0;
};
hm_status fm_scratch_reserve(FmScratch &s, int64_t max_mb_elems, int k);
void fm_scratch_free(FmScratch &s);
hm_status launch_fm_init_v(Launcher &L, const FmArgs &a);
hm_status launch_fm_forward(Launcher &L, const FmArgs &a, const int64_t *d_rowptr, int64_t row0, int32_t n_rows,
                            int64_t block_row0, const uint32_t *d_feature, const float *d_value, const float *d_label,
                            float *d_margin, float *d_dloss, float *d_stab, uint32_t *d_bad, int64_t nnz_hint, bool train);
hm_status launch_fm_update(Launcher &L, const FmArgs &a, const StepScalars &sc, FmScratch &fs,
                           const uint32_t *d_sorted_key, const uint2 *d_sorted_pay, const uint32_t *d_seg_off, int seg,
                           const float *d_dloss, const float *d_stab, int64_t nnz);
hm_status launch_fm_gather_vif(Launcher &L, const FmArgs &a, const int32_t *d_feature, int64_t n, float *d_vif);

// ---- options (hm_options.cpp) -------------------------------------------------
hm_status parse_options(const char *function_name, const char *options, hm_config *out);

}  // namespace hm
