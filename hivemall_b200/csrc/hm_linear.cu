// hm_linear.cu -- kernels of the linear learner (train_classifier / train_regressor).
//
// Stands in for the mini-batch path of the GeneralLearnerBaseUDTF family that
// BASELINE.json's north_star names (logistic / hinge / squared loss; SGD, AdaGrad,
// AdaDelta, Adam, FTRL under -mini_batch).  The reference mount has no source
// (README.md:20-22); the arithmetic contract is DESIGN.md section 3 and the CPU oracle
// (oracle/hm_oracle.c) restates the same contract independently.
//
//   linear_forward_kernel      warp per row: gather the weight out of the feature's record (FTRL:
//                              derive it from (z, n)), margin in fp64, dloss[row]; the block's hot features
//                              (HotInfo) take their weight from shared memory and have their gradient sums
//                              accumulated here, in a fixed order, instead of being sorted
//   linear_update_kernel       one tile of the feature-sorted mini-batch per CTA: dloss[row]*value,
//                              reduce-by-feature in fp64 in a fixed order (no float atomics), finished
//                              features stepped in batches; only segments that really straddle CTAs
//                              leave records, summed by the last CTA to finish (reduce_tile, apply_list);
//                              extra CTAs step the hot features from the forward kernel's partial sums
//   (mini-batches up to 400 K elements launch the two as programmatic dependents: pdl_wait)
//   linear_small_batch_kernel  mini-batches of a few hundred rows: one persistent CTA runs forward and
//                              reduce+step for up to 64 mini-batches per launch
//   predict / export / scale helpers
#include "hm_internal.h"
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace hm {

namespace {

constexpr int kFwdThreads = 256;
#ifndef HM_UPD_THREADS
#define HM_UPD_THREADS 256
#endif
constexpr int kUpdThreads = HM_UPD_THREADS;
constexpr int kUpdCtasPerSm = 1024 / kUpdThreads;     // resident CTAs of the update kernel per SM (64 registers per thread)
#ifndef HM_UPD_ITEMS
#define HM_UPD_ITEMS 8
#endif

// ---- scalar math of DESIGN.md section 3 (every fp32 op rounded on its own) ------

__device__ __forceinline__ float hm_exp(float x)
{
    x = fminf(fmaxf(x, -87.0f), 88.0f);
    const float n = rintf(__fmul_rn(x, 1.44269504f));
    float r = __fsub_rn(x, __fmul_rn(n, 0.693359375f));
    r = __fsub_rn(r, __fmul_rn(n, -2.12194440e-4f));
    float p = 1.3888889e-3f;
    p = __fadd_rn(__fmul_rn(p, r), 8.3333333e-3f);
    p = __fadd_rn(__fmul_rn(p, r), 4.1666667e-2f);
    p = __fadd_rn(__fmul_rn(p, r), 1.6666667e-1f);
    p = __fadd_rn(__fmul_rn(p, r), 0.5f);
    p = __fadd_rn(__fmul_rn(p, r), 1.0f);
    p = __fadd_rn(__fmul_rn(p, r), 1.0f);
    return __fmul_rn(p, __int_as_float(((int)n + 127) << 23));
}

__device__ __forceinline__ float dloss_of(int loss, float m, float y)
{
    if (loss == HM_LOSS_LOG) return __fdiv_rn(-y, __fadd_rn(1.0f, hm_exp(__fmul_rn(y, m))));
    if (loss == HM_LOSS_HINGE) return __fmul_rn(y, m) < 1.0f ? -y : 0.0f;
    return __fsub_rn(m, y);
}

// The other plain losses: derivative with respect to the margin.  p = Huber's threshold or the
// quantile; eps = width of the insensitive zone.  y is +-1 for the two classification losses.
__device__ __forceinline__ float dloss_more(int loss, float m, float y, float p, float eps)
{
    if (loss == HM_LOSS_SQUARED_HINGE) {               // max(0, 1 - z)^2
        const float z = __fmul_rn(y, m);
        return z < 1.0f ? __fmul_rn(__fmul_rn(-2.0f, y), __fsub_rn(1.0f, z)) : 0.0f;
    }
    if (loss == HM_LOSS_MODIFIED_HUBER) {              // z >= 1: 0; -1 <= z < 1: (1 - z)^2; z < -1: -4 z
        const float z = __fmul_rn(y, m);
        if (z >= 1.0f) return 0.0f;
        if (z >= -1.0f) return __fmul_rn(__fmul_rn(-2.0f, y), __fsub_rn(1.0f, z));
        return __fmul_rn(-4.0f, y);
    }
    const float e = __fsub_rn(m, y);
    if (loss == HM_LOSS_HUBER) return fabsf(e) <= p ? e : (e < 0.0f ? -p : p);        // e^2 / 2 inside, c (|e| - c / 2) outside
    if (loss == HM_LOSS_QUANTILE) return e < 0.0f ? -p : __fsub_rn(1.0f, p);          // tau (y - m) below, (1 - tau)(m - y) above
    const float a = __fsub_rn(fabsf(e), eps);
    if (!(a > 0.0f)) return 0.0f;
    if (loss == HM_LOSS_EPS_INSENSITIVE) return e < 0.0f ? -1.0f : 1.0f;              // max(0, |e| - eps)
    return __fmul_rn(e < 0.0f ? -2.0f : 2.0f, a);                                       // max(0, |e| - eps)^2
}

// Online learners on the same skeleton: the number that takes dloss's place for a row is
// -(step size) * (direction of the correction); xx = |x|^2 of the row (row_sqnorm_kernel).  Perceptron
// (Rosenblatt): step 1 on a mistake.  Passive-aggressive (Crammer et al. 2006): PA tau = l / xx, PA-I
// tau = min(C, l / xx), PA-II tau = l / (xx + 1 / (2C)), with l the hinge loss (classification) or the
// epsilon-insensitive loss (regression).  A row without valid elements (xx == 0) gives 0.
__device__ __forceinline__ float online_coef(int loss, float m, float y, float xx, float c, float eps)
{
    if (loss == HM_LOSS_PERCEPTRON) return __fmul_rn(y, m) <= 0.0f ? -y : 0.0f;
    float l, dir;
    if (loss == HM_LOSS_PA1_REGR || loss == HM_LOSS_PA2_REGR) {
        const float e = __fsub_rn(y, m);
        l = __fsub_rn(fabsf(e), eps);
        dir = e < 0.0f ? -1.0f : 1.0f;
    } else {
        l = __fsub_rn(1.0f, __fmul_rn(y, m));
        dir = y;
    }
    if (!(l > 0.0f)) return 0.0f;
    float tau;
    if (loss == HM_LOSS_PA2 || loss == HM_LOSS_PA2_REGR) tau = __fdiv_rn(l, __fadd_rn(xx, __fdiv_rn(0.5f, c)));
    else {
        if (!(xx > 0.0f)) return 0.0f;
        tau = __fdiv_rn(l, xx);
        if (loss != HM_LOSS_PA) tau = fminf(c, tau);
    }
    return -__fmul_rn(tau, dir);
}

// FTRL-Proximal closed form; the same operation sequence as the update step.
__device__ __forceinline__ float ftrl_weight(float z, float n, float alpha, float beta, float l1, float l2)
{
    if (fabsf(z) <= l1) return 0.0f;
    const float sgn = z < 0.0f ? -1.0f : 1.0f;
    const float num = __fsub_rn(z, __fmul_rn(sgn, l1));
    const float den = __fadd_rn(__fdiv_rn(__fadd_rn(beta, __fsqrt_rn(n)), alpha), l2);
    return -__fdiv_rn(num, den);
}

// RO: the replica is not written by the running kernel, so the read-only path may be used.
// How the replica (and dloss) may be read.  kLdRO: the running kernel and nothing alive beside it writes the
// data: read-only path.  kLdPlain: the running CTA itself writes it (small-batch kernel).  kLdCG: the
// predecessor of a programmatically dependent launch wrote it while this kernel was already alive: the
// read-only path would be illegal (the data is not constant over the kernel's lifetime), and an L1 line cached
// on this SM before the predecessor finished must not be hit: read at L2.
enum LdMode { kLdRO = 0, kLdPlain = 1, kLdCG = 2 };
template <int LD> __device__ __forceinline__ float ld_mode(const float *p) { return LD == kLdRO ? __ldg(p) : (LD == kLdCG ? __ldcg(p) : *p); }
template <int LD> __device__ __forceinline__ float2 ld_mode(const float2 *p) { return LD == kLdRO ? __ldg(p) : (LD == kLdCG ? __ldcg(p) : *p); }

template <int LD = kLdRO>
__device__ __forceinline__ float weight_at(const WeightView &v, size_t j)
{
    if (!v.ftrl) return ld_mode<LD>(v.base + j * v.stride);
    const float2 zn = ld_mode<LD>(reinterpret_cast<const float2 *>(v.base) + j);
    return ftrl_weight(zn.x, zn.y, v.alpha, v.beta, v.lambda1, v.lambda2);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- programmatic dependent launch (mid-size mini-batches only) -----------------------------
// Forward and update of a mini-batch are dependent launches, and a mini-batch of a few hundred thousand
// elements does not fill the GPU: what a kernel may do before its predecessor has finished (clear its
// tables, load the sorted tile -- everything that reads only the rows) then overlaps the predecessor's
// tail and its own launch latency.  pdl_wait() returns once the preceding kernel of the stream has completed
// and its writes are visible (at once when the launch carried no such dependency).
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- per-CTA timeline stamps: compiled in with -DHM_STAMP only (scripts/upd_timeline.py, scripts/build_variant.sh) ---------
#ifdef HM_STAMP
__device__ unsigned long long hm_dbg_stamps[4096 * 8];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define STAMP(i) do { if (threadIdx.x == 0 && blockIdx.x < 4096) hm_dbg_stamps[blockIdx.x * 8 + (i)] = gtime(); } while (0)
__device__ unsigned long long hm_dbg_fstamps[4096 * 8];      // the same for the forward kernel
#define FSTAMP(i) do { if (threadIdx.x == 0 && blockIdx.x < 4096) hm_dbg_fstamps[blockIdx.x * 8 + (i)] = gtime(); } while (0)
__device__ unsigned long long hm_dbg_seq[256 * 8];           // per mini-batch of the last block: CTA 0 of each launch
#define SEQSTAMP(seg, i) do { if (threadIdx.x == 0 && blockIdx.x == 0) hm_dbg_seq[((seg) & 255) * 8 + (i)] = gtime(); } while (0)
#define SEQSTAMP_ANY(seg, i) do { if (threadIdx.x == 0) hm_dbg_seq[((seg) & 255) * 8 + (i)] = gtime(); } while (0)
#else
#define SEQSTAMP(seg, i) do { } while (0)
#define SEQSTAMP_ANY(seg, i) do { } while (0)
#define STAMP(i) do { } while (0)
#define FSTAMP(i) do { } while (0)
#endif
#define USTAMP(i) STAMP(i)

// ---- forward ---------------------------------------------------------------------

struct FwdParams {
    WeightView wv;           // where the weight of a feature is read from
    const int64_t *rowptr;   // chunk row pointer (absolute element positions in the chunk)
    const uint32_t *feature;
    const float *value;
    const float *label;
    float *margin;           // indexed by row of the BLOCK (kept for hm_udtf_last_margins)
    float *dloss;            // indexed by row of the MINI-BATCH (table read by the update kernel)
    int64_t block_row0;      // first row of the block in the chunk
    int64_t row0;            // first row of this mini-batch in the chunk
    int32_t n_rows;
    int32_t task, loss;
    int32_t train;           // 0: prediction only
    int32_t lpr;             // lanes per row: 32, 16 or 8 (chosen by the host from the mean row length)
    const float *sqnorm;     // online learners other than the perceptron: |x|^2 per row of the BLOCK, else NULL
    float pa_c, pa_eps, loss_param;
    uint32_t dims;
    uint32_t *bad;           // set to 1 when a feature index >= dims is met (that element scores 0)
    // hot features of the block (training only; NULL: none): their weights are read from shared memory
    // instead of being gathered, and their gradient sums are accumulated right here, in a fixed order,
    // instead of being sorted.
    const HotInfo *hot;
    double *hot_sum;         // [kMaxHot][hot_stride]: per forward CTA, sum of dloss * value over its rows
    uint32_t *hot_cnt;       // [kMaxHot][hot_stride]: ... and the number of occurrences
    int32_t hot_stride;
};

// What the update multiplies a row's values with: the loss derivative, or an online learner's signed step.
template <int LD>
__device__ __forceinline__ float row_coef(const FwdParams &P, float m, float y, int64_t block_row)
{
    if (P.loss <= HM_LOSS_SQUARED) return dloss_of(P.loss, m, y);
    if (P.loss >= HM_LOSS_SQUARED_HINGE) return dloss_more(P.loss, m, y, P.loss_param, P.pa_eps);
    const float xx = P.sqnorm ? ld_mode<LD>(P.sqnorm + block_row) : 0.0f;
    return online_coef(P.loss, m, y, xx, P.pa_c, P.pa_eps);
}

// |x|^2 of every row of a block: fp32 squares, fp64 sum, rounded to fp32 once (elements with an index out of
// range do not count, as everywhere).  One warp per row.  The rows do not change: once per block, before its
// first mini-batch.
__global__ void __launch_bounds__(256) row_sqnorm_kernel(const int64_t *__restrict__ rowptr, int64_t row0, int64_t n_rows,
                                                         const uint32_t *__restrict__ feature, const float *__restrict__ value,
                                                         uint32_t dims, float *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), stride = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = warp; r < n_rows; r += stride) {
        const int64_t p0 = rowptr[row0 + r], p1 = rowptr[row0 + r + 1];
        double acc = 0.0;
        for (int64_t p = p0 + lane; p < p1; p += 32) {
            const float x = __ldcs(value + p);
            if (__ldcs(feature + p) < dims) acc += (double)__fmul_rn(x, x);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) out[r] = (float)acc;
    }
}

// Margins (and dloss) of rows r = warp_first, warp_first + warp_stride, ... of the mini-batch: one
// warp per row, lanes over the row's elements, fp64 shuffle reduction.
template <int LD>
__device__ __forceinline__ void forward_rows(const FwdParams &P, int warp_first, int warp_stride)
{
    const int lane = threadIdx.x & 31;
    for (int r = warp_first; r < P.n_rows; r += warp_stride) {
        const int64_t row = P.row0 + r;
        const int64_t p0 = P.rowptr[row], p1 = P.rowptr[row + 1];
        double acc = 0.0;
        for (int64_t p = p0 + lane; p < p1; p += 32) {
            const uint32_t f = __ldcs(P.feature + p);
            const float x = __ldcs(P.value + p);
            if (f < P.dims) acc += (double)__fmul_rn(weight_at<LD>(P.wv, f), x);
            else atomicOr(P.bad, 1u);
        }
        acc = warp_sum(acc);
        const float m = (float)acc;
        if (lane == 0) {
            P.margin[row - P.block_row0] = m;
            if (P.train) {
                const float lab = P.label[row];
                const float y = P.task == HM_TASK_CLASSIFIER ? (lab > 0.0f ? 1.0f : -1.0f) : lab;
                P.dloss[r] = row_coef<LD>(P, m, y, row - P.block_row0);
            }
        }
    }
}

// The same with LPR lanes per row (32 / LPR rows per warp side by side): rows of a few dozen elements leave
// fewer lanes idle in the gather loop, and two (or four) rows' dloss are worked out at once.  Same fp64 sums,
// another (fixed) tree order.
template <int LD, int LPR>
__device__ __forceinline__ void forward_rows_lpr(const FwdParams &P, int warp_first, int warp_stride)
{
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31, sub = lane & (LPR - 1), part = lane / LPR;
    for (int rb = warp_first * RPW; rb < P.n_rows; rb += warp_stride * RPW) {
        const int r = rb + part;
        const bool live = r < P.n_rows;
        const int64_t row = P.row0 + (live ? r : rb);
        const int64_t p0 = live ? P.rowptr[row] : 0, p1 = live ? P.rowptr[row + 1] : 0;
        double acc = 0.0;
        for (int64_t p = p0 + sub; p < p1; p += LPR) {
            const uint32_t f = __ldcs(P.feature + p);
            const float x = __ldcs(P.value + p);
            if (f < P.dims) acc += (double)__fmul_rn(weight_at<LD>(P.wv, f), x);
            else atomicOr(P.bad, 1u);
        }
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (live && sub == 0) {
            const float m = (float)acc;
            P.margin[row - P.block_row0] = m;
            if (P.train) {
                const float lab = P.label[row];
                const float y = P.task == HM_TASK_CLASSIFIER ? (lab > 0.0f ? 1.0f : -1.0f) : lab;
                P.dloss[r] = row_coef<LD>(P, m, y, row - P.block_row0);
            }
        }
    }
}

// ---- hot features in the forward kernel ------------------------------------------------------
constexpr uint32_t kHotNone = 0xffffffffu;

// Shared memory of a forward CTA that handles hot features (dynamic: a launch without hot features asks for
// none and leaves the whole carve-out to L1).  UNITS = row groups per CTA: 8 with a warp per row, 16 with half.
template <int UNITS>
struct HotSharedT {
    // accumulators per UNIT = group of lanes that works on one row at a time (a warp, or half a warp when rows
    // are short): sum of dloss * value per hot feature, occurrences, and the claim bytes
    static constexpr int kUnits = UNITS;
    double acc[UNITS][kMaxHot];
    uint32_t cnt[UNITS][kMaxHot];
    uint32_t key[kMaxHot];                  // HotInfo::key
    float w[kMaxHot];                       // the hot features' weights at the start of the mini-batch
    uint8_t tab_slot[kHotSlots];            // direct-mapped (hot_index): -> index into key[], kHotFree if none
    uint8_t claim[UNITS][kMaxHot];
};
// The footprint decides how much of the SM's 256 KB is left as L1 and how many CTAs of the sort stream fit beside
// this kernel: with 27 KB per CTA (a 4-byte key per table slot, 4-byte claims) the half-warp variant ran 50 us
// against 40 us (L1 hit rate 8 % against 17 %), and 4 KB of padding alone cost 2 % of the whole step.
static_assert(sizeof(HotSharedT<16>) <= 16 * 1024 && sizeof(HotSharedT<8>) <= 10 * 1024, "forward CTA shared memory");
static_assert(kMaxHot < 255, "slot numbers are bytes");
constexpr uint8_t kHotFree = 0xff;

template <class HS>
__device__ __forceinline__ uint32_t hot_slot_of(const HS &H, uint32_t key)
{
    const uint32_t s = H.tab_slot[hot_index(key)];
    return (s != kHotFree && H.key[s] == key) ? s : kHotNone;
}

// Add this warp's contributions c (lanes with slot != kHotNone) to the warp's accumulators.  Lanes that
// hit DIFFERENT slots add at once; when two lanes hit the same slot (a feature twice in one row) the adds
// are done one lane at a time in lane order, so that the sum's order is fixed.
template <class HS>
__device__ __forceinline__ void hot_accumulate(HS &H, int unit, int lane, uint32_t slot, float c)
{
    const bool hot = slot != kHotNone;
    const uint32_t hot_lanes = __ballot_sync(0xffffffffu, hot);
    if (hot_lanes == 0u) return;
    if (hot) H.claim[unit][slot] = (uint8_t)lane;
    __syncwarp();
    const bool lost = hot && H.claim[unit][slot] != (uint8_t)lane;
    if (!__any_sync(0xffffffffu, lost)) {
        if (hot) { H.acc[unit][slot] += (double)c; H.cnt[unit][slot] += 1u; }
    } else {
        for (uint32_t m = hot_lanes; m != 0u; m &= m - 1u) {
            if (lane == __ffs(m) - 1) { H.acc[unit][slot] += (double)c; H.cnt[unit][slot] += 1u; }
            __syncwarp();
        }
    }
    __syncwarp();
}

// Before the forward pass of a mini-batch: the block's hot keys into the CTA's direct-mapped table, their
// current weights, cleared accumulators.  Two barriers and two dependent round trips (count and key are
// loaded side by side; the weight gather needs the key).
template <int LD, class HS>
__device__ __forceinline__ void hot_prologue(const FwdParams &P, HS &H)
{
    const int tid = threadIdx.x;
    const uint32_t n = P.hot->count;
    const uint32_t key = tid < kMaxHot ? P.hot->key[tid] : 0u;
    for (int q = tid; q < kHotSlots / 4; q += kFwdThreads) reinterpret_cast<uint32_t *>(H.tab_slot)[q] = 0xffffffffu;
    for (int q = tid; q < HS::kUnits * kMaxHot; q += kFwdThreads) { (&H.acc[0][0])[q] = 0.0; (&H.cnt[0][0])[q] = 0u; }
    __syncthreads();
    FSTAMP(1);
    if (LD == kLdCG) pdl_wait();     // the weights below are the previous mini-batch's update
    FSTAMP(2);
    SEQSTAMP((int)((P.row0 - P.block_row0) / max(P.n_rows, 1)), 1);
    if (tid < (int)n) {
        H.tab_slot[hot_index(key)] = (uint8_t)tid;     // (the index is pairwise different for the keys of a HotInfo)
        H.key[tid] = key;
        H.w[tid] = key < P.dims ? weight_at<LD>(P.wv, key) : 0.0f;
    }
    __syncthreads();
}

// After it: this CTA's sums (units added in a fixed order) to slot `cta` of the per-feature arrays.
template <class HS>
__device__ __forceinline__ void hot_end(const FwdParams &P, HS &H, int cta)
{
    const int tid = threadIdx.x;
    __syncthreads();
    if (tid < (int)P.hot->count) {
        double sum = 0.0; uint32_t cnt = 0;
#pragma unroll
        for (int w2 = 0; w2 < HS::kUnits; w2++) { sum += H.acc[w2][tid]; cnt += H.cnt[w2][tid]; }
        P.hot_sum[(size_t)tid * P.hot_stride + cta] = sum;
        P.hot_cnt[(size_t)tid * P.hot_stride + cta] = cnt;
    }
}

// forward_rows_lpr with hot features: LPR lanes per row, 32 / LPR rows per warp side by side.  The margin adds
// the same fp32 products as the plain path; only the source of a hot feature's weight differs (shared memory
// instead of the gather; it is the same number).  After a row's dloss is known, the contributions dloss * value
// of its hot elements (what the update kernel would otherwise have found in the sorted stream) go to the
// accumulators of the row's unit; a lane keeps its first KEEP elements in registers for that, longer rows
// re-read their tail.
// What a warp can fetch of its FIRST row before the model is final (a programmatic dependent waits for the
// previous update at griddepcontrol.wait; the rows do not depend on it): the row's bounds, its label and the
// first 64 elements.  With one row per warp -- the mid-size mini-batches this is for -- that takes the row
// pointer -> feature/value chain and the label's round trip out of the time between two updates.
struct RowPre {
    int64_t p0, p1;
    uint32_t f[2];
    float x[2];
    float lab;
};
__device__ __forceinline__ void preload_first_row(const FwdParams &P, int warp_first, RowPre &pre)
{
    const int lane = threadIdx.x & 31;
    pre.p0 = 0; pre.p1 = 0; pre.lab = 0.0f;
    pre.f[0] = pre.f[1] = 0u; pre.x[0] = pre.x[1] = 0.0f;
    if (warp_first < P.n_rows) {
        const int64_t row = P.row0 + warp_first;
        pre.p0 = P.rowptr[row]; pre.p1 = P.rowptr[row + 1];
        if (lane == 0) pre.lab = P.label[row];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const int64_t p = pre.p0 + lane + i * 32;
            if (p < pre.p1) { pre.f[i] = __ldcs(P.feature + p); pre.x[i] = __ldcs(P.value + p); }
        }
    }
}

template <int LD, int LPR, bool PRE = false>
__device__ __forceinline__ void forward_rows_hot(const FwdParams &P, int warp_first, int warp_stride,
                                                 HotSharedT<256 / LPR> &H, const RowPre *pre = nullptr)
{
    static_assert(!PRE || LPR == 32, "the preloaded first row is a warp's row");
    constexpr int RPW = 32 / LPR;
    constexpr int KEEP = LPR == 32 ? 2 : 3;          // 64 or 48 elements of a row stay in registers
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane & (LPR - 1), part = lane / LPR;
    const int unit = RPW * warp + part;
    bool first = PRE;
    for (int rb = warp_first * RPW; rb < P.n_rows; rb += warp_stride * RPW) {
        const int r = rb + part;
        const bool live = r < P.n_rows;
        const int64_t row = P.row0 + (live ? r : rb);
        int64_t p0, p1;
        float lab;
        if (PRE && first) { p0 = pre->p0; p1 = pre->p1; lab = pre->lab; }
        else {
            p0 = live ? P.rowptr[row] : 0; p1 = live ? P.rowptr[row + 1] : 0;
            lab = (live && sub == 0) ? P.label[row] : 0.0f;       // (requested now, used after the sums)
        }
        double acc = 0.0;
        uint32_t slot[KEEP]; float xk[KEEP];
#pragma unroll
        for (int i = 0; i < KEEP; i++) {
            slot[i] = kHotNone; xk[i] = 0.0f;
            const int64_t p = p0 + sub + i * LPR;
            if (p < p1) {
                uint32_t f; float x;
                if (PRE && first) { f = pre->f[i < 2 ? i : 0]; x = pre->x[i < 2 ? i : 0]; }
                else { f = __ldcs(P.feature + p); x = __ldcs(P.value + p); }
                if (f < P.dims) {
                    const uint32_t sl = hot_slot_of(H, f);
                    const float w = sl != kHotNone ? H.w[sl] : weight_at<LD>(P.wv, f);
                    acc += (double)__fmul_rn(w, x);
                    slot[i] = sl; xk[i] = x;
                } else atomicOr(P.bad, 1u);
            }
        }
        for (int64_t p = p0 + sub + KEEP * LPR; p < p1; p += LPR) {
            const uint32_t f = __ldcs(P.feature + p);
            const float x = __ldcs(P.value + p);
            if (f < P.dims) {
                const uint32_t sl = hot_slot_of(H, f);
                acc += (double)__fmul_rn(sl != kHotNone ? H.w[sl] : weight_at<LD>(P.wv, f), x);
            } else atomicOr(P.bad, 1u);
        }
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        float dl = 0.0f;
        if (live && sub == 0) {
            const float m = (float)acc;
            P.margin[row - P.block_row0] = m;
            const float y = P.task == HM_TASK_CLASSIFIER ? (lab > 0.0f ? 1.0f : -1.0f) : lab;
            dl = row_coef<LD>(P, m, y, row - P.block_row0);
            P.dloss[r] = dl;
        }
        first = false;
        dl = __shfl_sync(0xffffffffu, dl, part * LPR);
        const int len = (int)min(p1 - p0, (int64_t)0x7fffffff);
        const int maxlen = __reduce_max_sync(0xffffffffu, len);          // (uniform over the warp)
#pragma unroll
        for (int i = 0; i < KEEP; i++)
            if (maxlen > i * LPR) hot_accumulate(H, unit, lane, slot[i], __fmul_rn(dl, xk[i]));
        for (int q0 = KEEP * LPR; q0 < maxlen; q0 += LPR) {
            const int64_t p = p0 + q0 + sub;
            uint32_t sl = kHotNone; float x = 0.0f;
            if (p < p1) {
                const uint32_t f = __ldcs(P.feature + p);
                x = __ldcs(P.value + p);
                if (f < P.dims) sl = hot_slot_of(H, f);
            }
            hot_accumulate(H, unit, lane, sl, __fmul_rn(dl, x));
        }
    }
}

// Forward for a small mini-batch inside one CTA: 8 lanes per row and four rows in flight per group,
// so that the three dependent loads of a row (row pointer -> feature/value -> weight) overlap
// across rows instead of queueing behind each other.  Same fp64 sums, other (fixed) tree order.
__device__ __forceinline__ void forward_rows_small(const FwdParams &P)
{
    constexpr int G = 8, R = 4;
    const int sub = threadIdx.x & (G - 1);
    const int group = threadIdx.x / G, n_groups = blockDim.x / G;
    const uint32_t gmask = 0xffu << ((threadIdx.x & 31) & ~(G - 1));
    for (int r0 = group * R; r0 < P.n_rows; r0 += n_groups * R) {
        int64_t p0[R], p1[R];
#pragma unroll
        for (int u = 0; u < R; u++) {
            const bool ok = r0 + u < P.n_rows;
            p0[u] = ok ? P.rowptr[P.row0 + r0 + u] : 0;
            p1[u] = ok ? P.rowptr[P.row0 + r0 + u + 1] : 0;
        }
        double acc[R];
        float lab[R];   // requested now, used after the sums
#pragma unroll
        for (int u = 0; u < R; u++) {
            acc[u] = 0.0;
            lab[u] = (sub == 0 && r0 + u < P.n_rows) ? P.label[P.row0 + r0 + u] : 0.0f;
        }
        int64_t longest = 0;
#pragma unroll
        for (int u = 0; u < R; u++) longest = max(longest, p1[u] - p0[u]);
        for (int64_t o = sub; o < longest; o += G) {
            uint32_t f[R]; float x[R]; bool in[R];
#pragma unroll
            for (int u = 0; u < R; u++) {
                in[u] = p0[u] + o < p1[u];
                f[u] = in[u] ? __ldcs(P.feature + p0[u] + o) : 0u;
                x[u] = in[u] ? __ldcs(P.value + p0[u] + o) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < R; u++)
                if (in[u]) {
                    if (f[u] < P.dims) acc[u] += (double)__fmul_rn(weight_at<kLdPlain>(P.wv, f[u]), x[u]);
                    else atomicOr(P.bad, 1u);
                }
        }
#pragma unroll
        for (int u = 0; u < R; u++) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) acc[u] += __shfl_xor_sync(gmask, acc[u], o);
            if (sub == 0 && r0 + u < P.n_rows) {
                const int64_t row = P.row0 + r0 + u;
                const float m = (float)acc[u];
                P.margin[row - P.block_row0] = m;
                const float y = P.task == HM_TASK_CLASSIFIER ? (lab[u] > 0.0f ? 1.0f : -1.0f) : lab[u];
                P.dloss[r0 + u] = row_coef<kLdPlain>(P, m, y, row - P.block_row0);
            }
        }
    }
}

// PDL: launched as a programmatic dependent of the previous mini-batch's update kernel (mid-size mini-batches).
template <bool PDL>
__global__ void __launch_bounds__(kFwdThreads, 8) linear_forward_kernel(FwdParams P)
{
    constexpr int LD = PDL ? kLdCG : kLdRO;
    extern __shared__ __align__(16) unsigned char fwd_dyn[];
    FSTAMP(0);
    SEQSTAMP((int)((P.row0 - P.block_row0) / max(P.n_rows, 1)), 0);
    if (PDL) pdl_launch_dependents();
    const int warp_first = blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5), warp_stride = gridDim.x * (kFwdThreads / 32);
    if (P.hot && P.hot->count != 0u) {     // (HotInfo is the sort's output: complete before this launch)
        if (P.lpr <= 16) {
            auto &H = *reinterpret_cast<HotSharedT<16> *>(fwd_dyn);
            hot_prologue<LD>(P, H);
            forward_rows_hot<LD, 16>(P, warp_first, warp_stride, H);
            hot_end(P, H, (int)blockIdx.x);
        } else {
            auto &H = *reinterpret_cast<HotSharedT<8> *>(fwd_dyn);
            RowPre pre;
            if (PDL) preload_first_row(P, warp_first, pre);     // (before the prologue's griddepcontrol.wait)
            hot_prologue<LD>(P, H);
            FSTAMP(3);
            forward_rows_hot<LD, 32, PDL>(P, warp_first, warp_stride, H, &pre);
            FSTAMP(4);
            hot_end(P, H, (int)blockIdx.x);
            FSTAMP(5);
            SEQSTAMP((int)((P.row0 - P.block_row0) / max(P.n_rows, 1)), 2);
        }
    } else {
        if (PDL) pdl_wait();
        // lanes per row by the mini-batch's mean row length (measured on 39-element rows: 32.7 us with a warp
        // per row, 28.8 with half a warp, 27.7 with a quarter)
        if (P.lpr == 8) forward_rows_lpr<LD, 8>(P, warp_first, warp_stride);
        else if (P.lpr == 16) forward_rows_lpr<LD, 16>(P, warp_first, warp_stride);
        else forward_rows<LD>(P, warp_first, warp_stride);
    }
}

// ---- learners that keep a variance per feature (AROW) -----------------------
// Forward: margin m = sum w_j x_j AND variance v = sum sigma_j x_j^2 of a row from one 8-byte gather per
// element (the record is (w, sigma)); then the two numbers that stand in dloss's place: c1 for the weight
// sums (times x), c2 = beta for the variance sums (times x^2).  Crammer, Kulesza, Dredze 2009, diagonal.
__device__ __forceinline__ void cov_coefs(int loss, float m, float y, float v, float r, float eps, float phi, float cc,
                                          float *c1, float *c2)
{
    *c1 = 0.0f; *c2 = 0.0f;
    if (loss >= HM_LOSS_CW) {
        // Confidence-weighted (Dredze, Crammer, Pereira 2008, variance form) and soft confidence-weighted
        // learning (Wang, Zhao, Hoi 2012), diagonal: alpha from the closed forms, c1 = -alpha y;
        // CW: the PRECISION grows by 2 alpha phi x^2 (c2 = 2 alpha phi); SCW: sigma -= beta sigma^2 x^2 (c2 = beta).
        if (!(v > 0.0f)) return;
        const float my = __fmul_rn(y, m), p2 = __fmul_rn(phi, phi);
        float alpha;
        if (loss == HM_LOSS_CW) {
            const float q = __fadd_rn(1.0f, __fmul_rn(__fmul_rn(2.0f, phi), my));
            const float rad = __fsub_rn(__fmul_rn(q, q), __fmul_rn(__fmul_rn(8.0f, phi), __fsub_rn(my, __fmul_rn(phi, v))));
            if (!(rad >= 0.0f)) return;
            alpha = __fdiv_rn(__fsub_rn(__fsqrt_rn(rad), q), __fmul_rn(__fmul_rn(4.0f, phi), v));
            if (!(alpha > 0.0f)) return;
            *c1 = -__fmul_rn(alpha, y);
            *c2 = __fmul_rn(__fmul_rn(2.0f, alpha), phi);
            return;
        }
        if (!(__fsub_rn(__fmul_rn(phi, __fsqrt_rn(v)), my) > 0.0f)) return;      // loss = phi sqrt(v) - y m
        const float t1 = __fmul_rn(my, my);
        if (loss == HM_LOSS_SCW1) {
            const float psi = __fadd_rn(1.0f, __fmul_rn(p2, 0.5f)), zeta = __fadd_rn(1.0f, p2);
            const float rad = __fadd_rn(__fmul_rn(__fmul_rn(t1, __fmul_rn(p2, p2)), 0.25f), __fmul_rn(__fmul_rn(v, p2), zeta));
            alpha = __fdiv_rn(__fsub_rn(__fsqrt_rn(rad), __fmul_rn(my, psi)), __fmul_rn(v, zeta));
            alpha = fminf(cc, alpha);
        } else {
            const float n = __fadd_rn(v, __fdiv_rn(0.5f, cc));
            const float vp2 = __fmul_rn(v, p2);
            const float g1 = __fmul_rn(__fmul_rn(__fmul_rn(p2, t1), v), v);
            const float g2 = __fmul_rn(__fmul_rn(__fmul_rn(4.0f, n), v), __fadd_rn(n, vp2));
            const float gamma = __fmul_rn(phi, __fsqrt_rn(__fadd_rn(g1, g2)));
            const float num = __fsub_rn(gamma, __fadd_rn(__fmul_rn(__fmul_rn(2.0f, my), n), __fmul_rn(__fmul_rn(p2, my), v)));
            const float den = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(n, n), __fmul_rn(__fmul_rn(n, v), p2)));
            alpha = __fdiv_rn(num, den);
        }
        if (!(alpha > 0.0f)) return;
        const float avp = __fmul_rn(__fmul_rn(alpha, v), phi);
        const float usq = __fmul_rn(0.5f, __fsub_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(avp, avp), __fmul_rn(4.0f, v))), avp));
        *c1 = -__fmul_rn(alpha, y);
        *c2 = __fdiv_rn(__fmul_rn(alpha, phi), __fadd_rn(usq, avp));
        return;
    }
    float l, dir;
    if (loss == HM_LOSS_AROW) { l = __fsub_rn(1.0f, __fmul_rn(y, m)); dir = y; if (!(l > 0.0f)) return; }
    else if (loss == HM_LOSS_AROW_REGR) { const float e = __fsub_rn(y, m); l = fabsf(e); dir = e < 0.0f ? -1.0f : 1.0f; }
    else { const float e = __fsub_rn(y, m); l = __fsub_rn(fabsf(e), eps); dir = e < 0.0f ? -1.0f : 1.0f; if (!(l > 0.0f)) return; }
    const float beta = __fdiv_rn(1.0f, __fadd_rn(v, r));
    const float alpha = __fmul_rn(l, beta);
    *c1 = -__fmul_rn(alpha, dir);
    *c2 = beta;
}

struct FwdCovParams {
    const float *rec;        // (w, sigma) per feature
    const int64_t *rowptr;
    const uint32_t *feature;
    const float *value, *label;
    float *margin, *c1, *c2;
    int64_t block_row0, row0;
    int32_t n_rows, task, loss;
    float r, eps, phi, cc;
    uint32_t dims;
    uint32_t *bad;
};

__global__ void __launch_bounds__(kFwdThreads) linear_forward_cov_kernel(FwdCovParams P)
{
    const int lane = threadIdx.x & 31;
    const int warp_first = blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5), warp_stride = gridDim.x * (kFwdThreads / 32);
    const float2 *rec = reinterpret_cast<const float2 *>(P.rec);
    for (int r = warp_first; r < P.n_rows; r += warp_stride) {
        const int64_t row = P.row0 + r;
        const int64_t p0 = P.rowptr[row], p1 = P.rowptr[row + 1];
        double am = 0.0, av = 0.0;
        for (int64_t p = p0 + lane; p < p1; p += 32) {
            const uint32_t f = __ldcs(P.feature + p);
            const float x = __ldcs(P.value + p);
            if (f < P.dims) {
                const float2 ws = __ldg(rec + f);
                am += (double)__fmul_rn(ws.x, x);
                av += (double)__fmul_rn(ws.y, __fmul_rn(x, x));
            } else atomicOr(P.bad, 1u);
        }
        am = warp_sum(am); av = warp_sum(av);
        if (lane == 0) {
            const float m = (float)am;
            P.margin[row - P.block_row0] = m;
            const float lab = P.label[row];
            const float y = P.task == HM_TASK_CLASSIFIER ? (lab > 0.0f ? 1.0f : -1.0f) : lab;
            float c1, c2;
            cov_coefs(P.loss, m, y, (float)av, P.r, P.eps, P.phi, P.cc, &c1, &c2);
            P.c1[r] = c1; P.c2[r] = c2;
        }
    }
}

// ---- update ------------------------------------------------------------------------

struct SegRec {           // a (feature, partial sum) record; cnt == 0 means empty
    double sum;
    uint32_t key, cnt;
};
static_assert(sizeof(SegRec) == 16, "boundary records travel as 16-byte vectors");

#ifndef HM_UPD_MAXGRID
#define HM_UPD_MAXGRID 1536
#endif
constexpr int kUpdMaxGrid = HM_UPD_MAXGRID;     // CTAs per launch; beyond that each CTA walks consecutive tiles

// Segments that really straddle CTAs are the hot features' long runs.  Each CTA leaves two
// 16-byte records: `first` = its piece of a segment that began in an earlier CTA (or the
// whole range, when the range lies inside one segment), `last` = its trailing segment when
// that one continues into the next CTA.  Pieces of one segment therefore sit in CONSECUTIVE
// CTAs: last[c], first[c+1], first[c+2], ... -- the last CTA to finish sums them in that order.
struct BoundaryRecs {
    uint4 *first, *last;   // SegRec as bits, one per CTA
    uint32_t *ticket;
};
// What changes from one mini-batch to the next; apart from this block an UpdParams is the same for every
// mini-batch of an ingest block (the small-batch kernel keeps one UpdParams and fills an UpdDyn per mini-batch).
struct UpdDyn {
    int32_t seg;             // mini-batch of the sorted block
    int32_t n_active;        // CTAs that own tiles (all own at least one unless the mini-batch is empty)
    int32_t fwd_grid;        // CTAs of the forward pass (partial sums of the hot features)
    float eta, adam_alpha;
    double inv_rows;         // 1 / rows of the mini-batch, computed in fp64
};

struct UpdParams {
    float *rec;              // replica: rec_width floats per feature = (w[, slot1[, slot2, unused]])
    int32_t rec_width;       // 1, 2 or 4
    int32_t ftrl_derived;    // records are (z, n); the weight is derived
    uint8_t *touched;        // nullable
    const uint32_t *skey;    // sorted features of the block
    const uint2 *spay;       // sorted (row in mini-batch, value bits)
    const float *dloss;      // dloss of every row of this mini-batch
    const uint32_t *seg_off;
    BoundaryRecs br;
    UpdDyn dyn;              // the mini-batch (kernels that train one mini-batch per launch read it from here)
    uint32_t dims;
    int32_t opt, reg, grad_norm;
    int32_t square;          // HM_OPT_COV, variance pass: the sums run over dloss * value^2 and step the variance
                             // (1: sigma -= sigma^2 g; 2: 1/sigma += g, the confidence-weighted form)
    float lambda, eps, rho, beta1, beta2, alpha, beta, lambda1, lambda2;
    // (the mini-batch's bounds in the sorted arrays are read from seg_off: the sorted arrays are shorter
    // than the CSR by the hot-feature and out-of-range elements, a number only the device knows)
    // hot features of the block (NULL: the block was sorted whole)
    const HotInfo *hot;
    const double *hot_sum;   // [kMaxHot][hot_stride] gradient sums left by the forward kernel's CTAs
    const uint32_t *hot_cnt; // ... and occurrence counts
    int32_t hot_stride;
};

// Replica record of one feature while it is being updated.
struct Rec { float w, s1, s2; };

// One 4-, 8- or 16-byte access per feature: weight and optimizer slots share a 32-byte sector,
// so the forward kernel's gather has already pulled the whole record into L2.
// (CG: at L2 -- the kernel is a programmatic dependent, see LdMode)
template <bool CG = false>
__device__ __forceinline__ Rec load_rec(const UpdParams &P, uint32_t j)
{
    Rec r; r.s1 = 0.f; r.s2 = 0.f;
    if (P.rec_width == 4) {
        const float4 *p = reinterpret_cast<const float4 *>(P.rec) + j;
        const float4 v = CG ? __ldcg(p) : *p; r.w = v.x; r.s1 = v.y; r.s2 = v.z;
    }
    else if (P.ftrl_derived) {
        const float2 *p = reinterpret_cast<const float2 *>(P.rec) + j;
        const float2 v = CG ? __ldcg(p) : *p;
        r.s1 = v.x; r.s2 = v.y;
        r.w = ftrl_weight(v.x, v.y, P.alpha, P.beta, P.lambda1, P.lambda2);
    }
    else if (P.rec_width == 2) {
        const float2 *p = reinterpret_cast<const float2 *>(P.rec) + j;
        const float2 v = CG ? __ldcg(p) : *p; r.w = v.x; r.s1 = v.y;
    }
    else r.w = CG ? __ldcg(P.rec + j) : P.rec[j];
    return r;
}

__device__ __forceinline__ void store_rec(const UpdParams &P, uint32_t j, const Rec &r)
{
    if (P.rec_width == 4) reinterpret_cast<float4 *>(P.rec)[j] = make_float4(r.w, r.s1, r.s2, 0.f);
    else if (P.ftrl_derived) reinterpret_cast<float2 *>(P.rec)[j] = make_float2(r.s1, r.s2);
    else if (P.rec_width == 2) reinterpret_cast<float2 *>(P.rec)[j] = make_float2(r.w, r.s1);
    else P.rec[j] = r.w;
    if (P.touched) P.touched[j] = 1;
}

// mean gradient of one feature from its fp64 sum: DESIGN.md 3.4
__device__ __forceinline__ float grad_of(const UpdParams &P, const UpdDyn &D, double sum, uint32_t cnt)
{
    if (P.grad_norm == HM_NORM_BATCH) return (float)(sum * D.inv_rows);
    if (P.grad_norm == HM_NORM_COUNT) return (float)(sum / (double)cnt);
    return (float)sum;
}

// One coordinate step: DESIGN.md 3.5-3.6 (every fp32 operation rounded on its own).
__device__ __forceinline__ void step_rec(const UpdParams &P, const UpdDyn &D, Rec &r, float g)
{
    float wv = r.w;
    if (P.reg == HM_REG_L2 && P.opt != HM_OPT_FTRL) g = __fadd_rn(g, __fmul_rn(P.lambda, wv));
    switch (P.opt) {
    case HM_OPT_SGD:
        wv = __fsub_rn(wv, __fmul_rn(D.eta, g));
        break;
    case HM_OPT_ADAGRAD: {
        const float G = __fadd_rn(r.s1, __fmul_rn(g, g));
        r.s1 = G;
        wv = __fsub_rn(wv, __fdiv_rn(__fmul_rn(D.eta, g), __fadd_rn(__fsqrt_rn(G), P.eps)));
        break;
    }
    case HM_OPT_ADADELTA: {
        const float omr = __fsub_rn(1.0f, P.rho);
        const float E = __fadd_rn(__fmul_rn(P.rho, r.s1), __fmul_rn(__fmul_rn(omr, g), g));
        const float dx = __fmul_rn(__fdiv_rn(__fsqrt_rn(__fadd_rn(r.s2, P.eps)), __fsqrt_rn(__fadd_rn(E, P.eps))), g);
        r.s2 = __fadd_rn(__fmul_rn(P.rho, r.s2), __fmul_rn(__fmul_rn(omr, dx), dx));
        r.s1 = E;
        wv = __fsub_rn(wv, dx);
        break;
    }
    case HM_OPT_ADAM: {
        const float m = __fadd_rn(__fmul_rn(P.beta1, r.s1), __fmul_rn(__fsub_rn(1.0f, P.beta1), g));
        const float v = __fadd_rn(__fmul_rn(P.beta2, r.s2), __fmul_rn(__fmul_rn(__fsub_rn(1.0f, P.beta2), g), g));
        r.s1 = m; r.s2 = v;
        wv = __fsub_rn(wv, __fdiv_rn(__fmul_rn(D.adam_alpha, m), __fadd_rn(__fsqrt_rn(v), P.eps)));
        break;
    }
    case HM_OPT_COV: {    // s1 = the feature's variance; both passes use its value of the start of the mini-batch
        if (P.square == 2) r.s1 = __fdiv_rn(1.0f, __fadd_rn(__fdiv_rn(1.0f, r.s1), g));     // CW: the precision grows
        else if (P.square) r.s1 = __fsub_rn(r.s1, __fmul_rn(__fmul_rn(r.s1, r.s1), g));
        else wv = __fsub_rn(wv, __fmul_rn(r.s1, g));
        break;
    }
    case HM_OPT_FTRL: {   // (s1, s2) = (z, n)
        const float n1 = __fadd_rn(r.s2, __fmul_rn(g, g));
        const float sq1 = __fsqrt_rn(n1);
        const float sigma = __fdiv_rn(__fsub_rn(sq1, __fsqrt_rn(r.s2)), P.alpha);
        const float z = __fsub_rn(__fadd_rn(r.s1, g), __fmul_rn(sigma, wv));
        r.s1 = z; r.s2 = n1;
        if (fabsf(z) <= P.lambda1) wv = 0.0f;
        else {
            const float sgn = z < 0.0f ? -1.0f : 1.0f;
            const float num = __fsub_rn(z, __fmul_rn(sgn, P.lambda1));
            const float den = __fadd_rn(__fdiv_rn(__fadd_rn(P.beta, sq1), P.alpha), P.lambda2);
            wv = -__fdiv_rn(num, den);
        }
        break;
    }
    }
    if (P.reg == HM_REG_L1 && P.opt != HM_OPT_FTRL) {
        const float a = __fsub_rn(fabsf(wv), __fmul_rn(D.eta, P.lambda));
        wv = a > 0.0f ? copysignf(a, wv) : 0.0f;
    }
    r.w = wv;
}

#ifndef HM_UPD_ITEMS_MID
#define HM_UPD_ITEMS_MID 4
#endif
namespace upd8 {            // 2048-element tiles: mini-batches that fill the GPU, and the small-batch kernel
constexpr int kUpdItems = HM_UPD_ITEMS;
constexpr bool kUpdCov = false;
#include "hm_update.cuh"
}  // namespace upd8
namespace upd4 {            // 1024-element tiles: mid-size mini-batches
constexpr int kUpdItems = HM_UPD_ITEMS_MID;
constexpr bool kUpdCov = false;
#include "hm_update.cuh"
}  // namespace upd4
namespace upd2 {            // 512-element tiles: mini-batches of up to 100 K elements
constexpr int kUpdItems = 2;
constexpr bool kUpdCov = false;
#include "hm_update.cuh"
}  // namespace upd2
namespace updc {            // the variance-keeping learners (AROW): 1024-element tiles, sums over value or value^2
constexpr int kUpdItems = HM_UPD_ITEMS_MID;
constexpr bool kUpdCov = true;
#include "hm_update.cuh"
}  // namespace updc
using namespace upd8;

// ---- small mini-batches: one persistent CTA ----------------------------------------
//
// When a mini-batch is a few hundred rows (-mini_batch 128 is BASELINE.json's configs[0]) two
// launches per mini-batch are all latency.  This kernel runs up to kSmallSegs consecutive
// mini-batches of a sorted block in ONE CTA: forward of mini-batch s, barrier, reduce-by-feature and
// optimizer step of mini-batch s (a single tile, so no piece ever leaves the CTA), barrier, s+1.
// Same arithmetic, same order as the two-kernel path: the results are bit-identical.
constexpr int kSmallSegs = 64;

constexpr int kSmallMaxRows = 1024;                 // rows per mini-batch on this path
constexpr size_t kSmallModelMax = 128 * 1024;       // a replica up to this size lives in shared memory for the launch

struct SmallParams {
    FwdParams F;             // row0 / n_rows are filled per mini-batch
    UpdParams U;             // seg / eta / adam_alpha / inv_rows are filled per mini-batch
    int32_t first_seg, n_segs, mini_batch;
    uint32_t model_floats;   // > 0: the replica (dims * rec_width floats) is staged in shared memory
    float eta[kSmallSegs], adam_alpha[kSmallSegs];
    int32_t rows[kSmallSegs];
};

__global__ void __launch_bounds__(kUpdThreads, 1) linear_small_batch_kernel(const __grid_constant__ SmallParams P)
{
    __shared__ __align__(16) unsigned char smem_raw[kUpdSmemTile];
    __shared__ Scan s_warp[kUpdThreads / 32];
    __shared__ SegRec s_carry, s_first;
    __shared__ uint32_t s_carry_open, s_edge[2];
    __shared__ int s_nlist;
    UpdShared S;
    S.s_key = reinterpret_cast<uint32_t *>(smem_raw);
    S.s_val = reinterpret_cast<float *>(smem_raw + (size_t)kUpdPadded * 4);
    S.l_key = reinterpret_cast<uint32_t *>(smem_raw + (size_t)kUpdPadded * 8);
    S.l_g = reinterpret_cast<float *>(smem_raw + (size_t)kUpdPadded * 8 + (size_t)kUpdList * 4);
    S.s_warp = s_warp; S.s_carry = &s_carry; S.s_first = &s_first; S.s_carry_open = &s_carry_open;
    S.s_edge = s_edge; S.s_nlist = &s_nlist;
    const int tid = threadIdx.x;
    // The dloss table of a mini-batch (<= 1024 rows) never leaves shared memory, and neither does a small
    // replica (a9a-shaped models are a kilobyte): every mini-batch of the launch then costs three dependent
    // global round trips (row pointer, features/values, sorted tile) instead of six.
    __shared__ float s_dloss[kSmallMaxRows];
    extern __shared__ __align__(16) float s_model[];
    float *const g_rec = P.U.rec;
    const bool staged = P.model_floats != 0;
    if (staged) {
        for (uint32_t i = tid; i < P.model_floats; i += kUpdThreads) s_model[i] = g_rec[i];
        __syncthreads();
    }
    for (int s = 0; s < P.n_segs; s++) {
        FwdParams F = P.F;
        F.row0 = P.F.block_row0 + (int64_t)(P.first_seg + s) * P.mini_batch;
        F.n_rows = P.rows[s];
        F.dloss = s_dloss;
        if (staged) F.wv.base = s_model;
        forward_rows_small(F);
        if (tid == 0) { s_carry.cnt = 0; s_carry_open = 0; s_first.cnt = 0; s_edge[0] = 1; s_edge[1] = 1; s_nlist = 0; }
        __syncthreads();     // dloss of every row is written; the whole mini-batch is one closed range
        UpdParams U = P.U;
        UpdDyn D;
        D.seg = P.first_seg + s; D.eta = P.eta[s]; D.adam_alpha = P.adam_alpha[s];
        D.inv_rows = 1.0 / (double)P.rows[s]; D.n_active = 1; D.fwd_grid = 0;
        U.dloss = s_dloss;
        if (staged) U.rec = s_model;
        const uint32_t s0 = U.seg_off[D.seg];
        const int n = (int)(U.seg_off[D.seg + 1] - s0);       // <= kUpdTile, checked by the host
        if (n > 0) {
            reduce_tile<kLdPlain>(U, D, S, s0, n, true, true);
            apply_list(U, D, S.l_key, S.l_g, s_nlist, tid);
        }
        __syncthreads();     // the step is complete before the next forward reads the replica
    }
    if (staged)
        for (uint32_t i = tid; i < P.model_floats; i += kUpdThreads) g_rec[i] = s_model[i];
}

// ---- model table --------------------------------------------------------------------

constexpr int kExpThreads = 256;
constexpr int kExpTile = 4096;

__device__ __forceinline__ bool emit(const WeightView &wv, const uint8_t *touched, int64_t j)
{
    return touched ? touched[j] != 0 : weight_at(wv, (size_t)j) != 0.0f;
}

__global__ void __launch_bounds__(kExpThreads)
export_count_kernel(WeightView wv, const uint8_t *__restrict__ touched, int64_t dims,
                    uint32_t *__restrict__ counts)
{
    __shared__ uint32_t s_tot;
    if (threadIdx.x == 0) s_tot = 0;
    __syncthreads();
    const int64_t j0 = (int64_t)blockIdx.x * kExpTile;
    uint32_t c = 0;
    for (int i = 0; i < kExpTile / kExpThreads; i++) {
        const int64_t j = j0 + i * kExpThreads + threadIdx.x;
        if (j < dims && emit(wv, touched, j)) c++;
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_tot, c);
    __syncthreads();
    if (threadIdx.x == 0) counts[blockIdx.x] = s_tot;
}

__global__ void __launch_bounds__(1024) export_scan_kernel(uint32_t *__restrict__ counts, int n, int64_t *total)
{
    // exclusive scan of counts[0..n) by one CTA; counts[n] and *total receive the sum
    __shared__ uint32_t warp_tot[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int per = (n + 1023) / 1024;
    const int lo = t * per, hi = min(lo + per, n);
    uint32_t s = 0;
    for (int i = lo; i < hi; i++) s += counts[i];
    uint32_t inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t v = warp_tot[lane], iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += u; }
        warp_tot[lane] = iv - v;
        if (lane == 31) *total = (int64_t)iv;
    }
    __syncthreads();
    uint32_t run = warp_tot[warp] + inc - s;
    for (int i = lo; i < hi; i++) { uint32_t v = counts[i]; counts[i] = run; run += v; }
}

__global__ void __launch_bounds__(kExpThreads)
export_write_kernel(WeightView wv, const uint8_t *__restrict__ touched, int64_t dims,
                    const uint32_t *__restrict__ offsets, int32_t *__restrict__ feature, float *__restrict__ weight)
{
    __shared__ uint32_t warp_tot[kExpThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t j0 = (int64_t)blockIdx.x * kExpTile;
    uint32_t base = offsets[blockIdx.x];
    // blocked arrangement keeps the output ascending: thread t owns 16 consecutive features
    const int per = kExpTile / kExpThreads;
    uint32_t flags = 0;
    for (int i = 0; i < per; i++) {
        const int64_t j = j0 + (int64_t)threadIdx.x * per + i;
        if (j < dims && emit(wv, touched, j)) flags |= 1u << i;
    }
    const uint32_t c = __popc(flags);
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0;
    for (int w2 = 0; w2 < warp; w2++) wbase += warp_tot[w2];
    uint32_t pos = base + wbase + inc - c;
    for (int i = 0; i < per; i++) {
        if (flags & (1u << i)) {
            const int64_t j = j0 + (int64_t)threadIdx.x * per + i;
            feature[pos] = (int32_t)j;
            weight[pos] = weight_at(wv, (size_t)j);
            pos++;
        }
    }
}

__global__ void scale_kernel(float *__restrict__ p, int64_t n, float scale)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = __fmul_rn(p[i], scale);
}

// dst[i*width + c] = src[i*stride + off + c]  (gather == true)  or the reverse (scatter)
__global__ void strided_copy_kernel(float *__restrict__ dense, float *__restrict__ strided, int64_t n, int stride,
                                    int off, int width, bool gather)
{
    const int64_t total = n * width;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = q / width;
        const int c = (int)(q - i * width);
        if (gather) dense[q] = strided[i * stride + off + c];
        else strided[i * stride + off + c] = dense[q];
    }
}

__global__ void dense_weights_kernel(WeightView wv, int64_t dims, float *__restrict__ out)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < dims; j += (int64_t)gridDim.x * blockDim.x)
        out[j] = weight_at(wv, (size_t)j);
}

__global__ void fill_u32_kernel(uint32_t *__restrict__ p, int64_t n, uint32_t v)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}

__global__ void fill_strided_kernel(float *__restrict__ p, int64_t n, int stride, float v)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i * stride] = v;
}

int grid_for(int64_t ctas_wanted, int per_sm)
{
    const int64_t cap = (int64_t)num_sms() * per_sm;
    if (ctas_wanted < 1) ctas_wanted = 1;
    return (int)(ctas_wanted < cap ? ctas_wanted : cap);
}

}  // namespace

// Launch with programmatic stream serialisation when `pdl`: the kernel may start while its predecessor in the
// stream is still running and synchronises with it at its own pdl_wait().  Used for mid-size mini-batches only:
// with grids that fill the GPU the early CTAs take slots from the running kernel's tail (measured slower).
constexpr int64_t kSmallTileMaxElems = 100000;   // 512-element update tiles up to here (measured: +12 % at 20 K and 80 K, -2 % at 160 K)
constexpr int64_t kPdlMaxElems = 400000;      // elements per mini-batch (measured: +8 % at 320 K, -10 % at 640 K)
template <class P>
cudaError_t launch_maybe_pdl(void (*kernel)(P), int grid, int block, cudaStream_t stream, bool pdl, const P &params,
                             size_t dyn_smem = 0)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)block); cfg.dynamicSmemBytes = dyn_smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, params);
}

// HM_PDL=0 in the environment turns programmatic dependent launches off (A/B measurements)
bool pdl_enabled()
{
    const char *e = std::getenv("HM_PDL");
    return !(e && e[0] == '0');
}

int linear_forward_grid(int32_t n_rows) { return grid_for(((int64_t)n_rows + 7) / 8, 8); }

hm_status launch_linear_forward(Launcher &L, const LinearArgs &a, const int64_t *d_rowptr, int64_t row0,
                                int32_t n_rows, int64_t block_row0, const uint32_t *d_feature,
                                const float *d_value, const float *d_label, float *d_margin,
                                float *d_dloss, uint32_t *d_bad, int64_t nnz_hint, const HotCols &hc)
{
    if (n_rows <= 0) return HM_OK;
    FwdParams P;
    P.hot = hc.info; P.hot_sum = hc.sum; P.hot_cnt = hc.cnt; P.hot_stride = hc.stride;
    {
        // Narrow row groups only when the mini-batch is large enough; a smaller one keeps a warp per row (more
        // CTAs in flight matter more there than instructions per row).  Measured on 39-element rows, examples/s
        // with a warp / half a warp per row: 8192 rows 262 / 225 M, 16384 rows 381 / 395 M, 32768 rows 448 / 474 M.
        const int64_t mean = nnz_hint / std::max<int32_t>(n_rows, 1);
        const bool fills = (int64_t)n_rows >= (int64_t)num_sms() * 100;
        P.lpr = (!fills || nnz_hint <= 0 || mean > 48) ? 32 : (mean <= 24 && !hc.info ? 8 : 16);
        // HM_FWD_LPR=32|16|8 in the environment overrides the choice (A/B measurements)
        if (const char *e = std::getenv("HM_FWD_LPR")) {
            const int v = std::atoi(e);
            if (v == 32 || v == 16 || (v == 8 && !hc.info)) P.lpr = v;
        }
    }
    const size_t dyn = !hc.info ? 0 : (P.lpr <= 16 ? sizeof(HotSharedT<16>) : sizeof(HotSharedT<8>));
    P.wv = a.wv; P.rowptr = d_rowptr; P.feature = d_feature; P.value = d_value; P.label = d_label;
    P.margin = d_margin; P.dloss = d_dloss;
    P.block_row0 = block_row0; P.row0 = row0; P.n_rows = n_rows;
    P.task = a.cfg.task; P.loss = a.cfg.loss; P.train = 1;
    P.sqnorm = a.sqnorm; P.pa_c = a.cfg.pa_c; P.pa_eps = a.cfg.pa_epsilon; P.loss_param = a.cfg.loss_param;
    P.dims = (uint32_t)a.cfg.dims; P.bad = d_bad;
    const int grid = linear_forward_grid(n_rows);
    if (hc.info && grid > hc.stride) HM_FAIL(HM_ERR_STATE, "hot-feature scratch too small (%d > %d)", grid, hc.stride);
    // algorithmic bytes: feature+value read, w gather per element; rowptr, label, margin, dloss per row
    const double bytes = 12.0 * (double)nnz_hint + 20.0 * (double)n_rows;
    if (pdl_enabled() && nnz_hint <= kPdlMaxElems)
        HM_LAUNCH(L, K_LIN_FORWARD, bytes, (launch_maybe_pdl(linear_forward_kernel<true>, grid, kFwdThreads, L.stream, true, P, dyn)));
    else
        HM_LAUNCH(L, K_LIN_FORWARD, bytes, (linear_forward_kernel<false><<<grid, kFwdThreads, dyn, L.stream>>>(P)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_linear_forward_cov(Launcher &L, const LinearArgs &a, const int64_t *d_rowptr, int64_t row0, int32_t n_rows,
                                    int64_t block_row0, const uint32_t *d_feature, const float *d_value, const float *d_label,
                                    float *d_margin, float *d_c1, float *d_c2, uint32_t *d_bad, int64_t nnz_hint)
{
    if (n_rows <= 0) return HM_OK;
    if (a.rec_width != 2) HM_FAIL(HM_ERR_STATE, "variance-keeping learners use (w, sigma) records");
    FwdCovParams P;
    P.rec = a.rec; P.rowptr = d_rowptr; P.feature = d_feature; P.value = d_value; P.label = d_label;
    P.margin = d_margin; P.c1 = d_c1; P.c2 = d_c2; P.block_row0 = block_row0; P.row0 = row0; P.n_rows = n_rows;
    P.task = a.cfg.task; P.loss = a.cfg.loss; P.r = a.cfg.cov_r; P.eps = a.cfg.pa_epsilon; P.phi = a.cfg.cov_phi; P.cc = a.cfg.pa_c;
    P.dims = (uint32_t)a.cfg.dims; P.bad = d_bad;
    HM_LAUNCH(L, K_LIN_FORWARD, 16.0 * (double)nnz_hint + 24.0 * (double)n_rows,
              (linear_forward_cov_kernel<<<linear_forward_grid(n_rows), kFwdThreads, 0, L.stream>>>(P)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_row_sqnorm(Launcher &L, const int64_t *d_rowptr, int64_t row0, int64_t n_rows, const uint32_t *d_feature,
                            const float *d_value, int64_t dims, float *d_out, int64_t nnz_hint)
{
    if (n_rows <= 0) return HM_OK;
    HM_LAUNCH(L, K_MISC, 8.0 * (double)nnz_hint + 12.0 * (double)n_rows,
              (row_sqnorm_kernel<<<grid_for((n_rows + 7) / 8, 8), 256, 0, L.stream>>>(d_rowptr, row0, n_rows, d_feature, d_value,
                                                                                       (uint32_t)dims, d_out)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_predict_linear(Launcher &L, const WeightView &wv, int64_t dims, int64_t n_rows, const int64_t *d_rowptr,
                                const uint32_t *d_feature, const float *d_value, float *d_out, uint32_t *d_bad)
{
    if (n_rows <= 0) return HM_OK;
    if (n_rows > 0x7fffffff) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "predict: too many rows in one call");
    FwdParams P;
    P.wv = wv; P.rowptr = d_rowptr; P.feature = d_feature; P.value = d_value; P.label = nullptr;
    P.margin = d_out; P.dloss = nullptr;
    P.block_row0 = 0; P.row0 = 0; P.n_rows = (int32_t)n_rows;
    P.task = 0; P.loss = 0; P.train = 0; P.sqnorm = nullptr; P.pa_c = 0.f; P.pa_eps = 0.f; P.loss_param = 0.f;
    P.dims = (uint32_t)dims; P.bad = d_bad;
    P.hot = nullptr; P.hot_sum = nullptr; P.hot_cnt = nullptr; P.hot_stride = 0; P.lpr = 32;
    const int grid = grid_for((n_rows + 7) / 8, 8);
    HM_LAUNCH(L, K_PREDICT, 0.0, (linear_forward_kernel<false><<<grid, kFwdThreads, 0, L.stream>>>(P)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status update_scratch_reserve(UpdateScratch &u)
{
    if (u.boundary) return HM_OK;
    // one tile per CTA while the CTAs fit on the device at once (four per SM); beyond that each CTA walks
    // consecutive tiles -- a second, mostly empty wave of CTAs costs more than the walk (measured)
    u.grid = std::min(kUpdMaxGrid, num_sms() * kUpdCtasPerSm);
    // [ first: grid x 16 B | last: grid x 16 B | ticket ]
    const size_t bytes = (size_t)2 * u.grid * 16 + 16;
    HM_CUDA(cudaMalloc(&u.boundary, bytes));
    HM_CUDA(cudaMemset(u.boundary, 0, bytes));
    u.done_counter = reinterpret_cast<uint32_t *>(static_cast<char *>(u.boundary) + (size_t)2 * u.grid * 16);
    // gradient sums of the hot features: one double and one count per (feature, forward CTA)
    u.hot_stride = num_sms() * 8;
    HM_CUDA(cudaMalloc(&u.hot_sum, (size_t)kMaxHot * u.hot_stride * sizeof(double)));
    HM_CUDA(cudaMalloc(&u.hot_cnt, (size_t)kMaxHot * u.hot_stride * sizeof(uint32_t)));
    HM_CUDA(cudaMemset(u.hot_sum, 0, (size_t)kMaxHot * u.hot_stride * sizeof(double)));
    HM_CUDA(cudaMemset(u.hot_cnt, 0, (size_t)kMaxHot * u.hot_stride * sizeof(uint32_t)));
    return HM_OK;
}

void update_scratch_free(UpdateScratch &u)
{
    cudaFree(u.boundary);
    cudaFree(u.hot_sum); cudaFree(u.hot_cnt);
    u.boundary = nullptr; u.done_counter = nullptr; u.grid = 0; u.hot_sum = nullptr; u.hot_cnt = nullptr; u.hot_stride = 0;
}

hm_status launch_linear_update(Launcher &L, const LinearArgs &a, const StepScalars &sc, UpdateScratch &us,
                               const uint32_t *d_sorted_key, const uint2 *d_sorted_pay,
                               const uint32_t *d_seg_off, int seg, int64_t seg_begin, const float *d_dloss, int64_t nnz_hint,
                               const HotCols &hc, int64_t sorted_elems)
{
    UpdParams P;
    (void)seg_begin;
    P.hot = hc.info; P.hot_sum = hc.sum; P.hot_cnt = hc.cnt; P.hot_stride = hc.stride;
    P.dyn.fwd_grid = linear_forward_grid(sc.n_rows);
    P.rec = a.rec; P.rec_width = a.rec_width; P.ftrl_derived = a.wv.ftrl; P.touched = nullptr;
    P.skey = d_sorted_key; P.spay = d_sorted_pay; P.dloss = d_dloss; P.seg_off = d_seg_off;
    P.br.first = reinterpret_cast<uint4 *>(us.boundary);
    P.br.last = P.br.first + us.grid;
    P.br.ticket = us.done_counter;
    P.dyn.seg = seg; P.dims = (uint32_t)a.cfg.dims;
    P.opt = a.cfg.opt; P.reg = a.cfg.reg; P.grad_norm = a.cfg.grad_norm; P.square = a.cov_pass == 2 ? (a.cfg.loss == HM_LOSS_CW ? 2 : 1) : 0;
    P.dyn.eta = sc.eta; P.dyn.adam_alpha = sc.adam_alpha; P.lambda = a.cfg.lambda; P.eps = a.cfg.eps; P.rho = a.cfg.rho;
    P.beta1 = a.cfg.beta1; P.beta2 = a.cfg.beta2; P.alpha = a.cfg.alpha; P.beta = a.cfg.beta;
    P.lambda1 = a.cfg.lambda1; P.lambda2 = a.cfg.lambda2;
    P.dyn.inv_rows = 1.0 / (double)sc.n_rows;
    // Tile size by the mini-batch's element count (hm_update.cuh): 512 elements up to 100 K, 1024 up to the
    // bound for dependent launches, 2048 above.
    const bool cov = a.cov_pass != 0;
    const int tier = cov ? 0 : (nnz_hint <= kSmallTileMaxElems ? 2 : (nnz_hint <= kPdlMaxElems ? 4 : 8));
    const int64_t tile = cov ? updc::kUpdTile : (tier == 2 ? upd2::kUpdTile : (tier == 4 ? upd4::kUpdTile : upd8::kUpdTile));
    // one tile per CTA up to kUpdMaxGrid CTAs; beyond that each CTA walks consecutive tiles.
    // nnz_hint is exact (the host knows rowptr at every mini-batch boundary).
    const int64_t tiles = std::max<int64_t>((nnz_hint + tile - 1) / tile, 1);
    const int64_t tiles_per_cta = (tiles + us.grid - 1) / us.grid;
    const int grid = (int)((tiles + tiles_per_cta - 1) / tiles_per_cta);
    P.dyn.n_active = grid;
    // algorithmic bytes: sorted feature + (row, value) stream; dloss table and the per-feature
    // state traffic (number of distinct features is data dependent) are not counted
    const double bytes = 12.0 * (double)(sorted_elems >= 0 ? sorted_elems : nnz_hint);
    // (nnz_hint bounds the sorted mini-batch from above: CTAs past its end find nothing to do; kHotCtas
    // more CTAs step the hot features)
    const int ctas = grid + (hc.info ? kHotCtas : 0);
    const bool pdl = pdl_enabled();
    if (cov) {
        HM_LAUNCH(L, K_LIN_UPDATE, bytes, (updc::linear_update_kernel<false><<<ctas, kUpdThreads, 0, L.stream>>>(P)));
    } else if (tier == 2) {
        if (pdl) HM_LAUNCH(L, K_LIN_UPDATE, bytes, (launch_maybe_pdl(upd2::linear_update_kernel<true>, ctas, kUpdThreads, L.stream, true, P)));
        else HM_LAUNCH(L, K_LIN_UPDATE, bytes, (upd2::linear_update_kernel<false><<<ctas, kUpdThreads, 0, L.stream>>>(P)));
    } else if (tier == 4) {
        if (pdl) HM_LAUNCH(L, K_LIN_UPDATE, bytes, (launch_maybe_pdl(upd4::linear_update_kernel<true>, ctas, kUpdThreads, L.stream, true, P)));
        else HM_LAUNCH(L, K_LIN_UPDATE, bytes, (upd4::linear_update_kernel<false><<<ctas, kUpdThreads, 0, L.stream>>>(P)));
    } else {
        HM_LAUNCH(L, K_LIN_UPDATE, bytes, (upd8::linear_update_kernel<false><<<ctas, kUpdThreads, 0, L.stream>>>(P)));
    }
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

bool small_batch_eligible(int32_t mini_batch, int64_t max_mb_elems)
{
    return mini_batch <= 1024 && max_mb_elems <= kUpdTile;
}

hm_status launch_linear_small_batches(Launcher &L, const LinearArgs &a, const int64_t *d_rowptr, int64_t block_row0,
                                      int first_seg, int n_segs, const int32_t *rows, const float *eta,
                                      const float *adam_alpha, const uint32_t *d_feature, const float *d_value,
                                      const float *d_label, float *d_margin, float *d_dloss, uint32_t *d_bad,
                                      const uint32_t *d_sorted_key, const uint2 *d_sorted_pay, const uint32_t *d_seg_off,
                                      int64_t nnz_hint)
{
    if (n_segs < 1 || n_segs > kSmallSegs) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "small-batch launch of %d mini-batches", n_segs);
    SmallParams P;
    std::memset(&P, 0, sizeof(P));
    P.F.wv = a.wv; P.F.rowptr = d_rowptr; P.F.feature = d_feature; P.F.value = d_value; P.F.label = d_label;
    P.F.margin = d_margin; P.F.dloss = d_dloss; P.F.block_row0 = block_row0;
    P.F.task = a.cfg.task; P.F.loss = a.cfg.loss; P.F.train = 1; P.F.dims = (uint32_t)a.cfg.dims; P.F.bad = d_bad;
    P.F.sqnorm = a.sqnorm; P.F.pa_c = a.cfg.pa_c; P.F.pa_eps = a.cfg.pa_epsilon; P.F.loss_param = a.cfg.loss_param;
    P.U.rec = a.rec; P.U.rec_width = a.rec_width; P.U.ftrl_derived = a.wv.ftrl; P.U.touched = nullptr;
    P.U.skey = d_sorted_key; P.U.spay = d_sorted_pay; P.U.dloss = d_dloss; P.U.seg_off = d_seg_off;
    P.U.dims = (uint32_t)a.cfg.dims; P.U.opt = a.cfg.opt; P.U.reg = a.cfg.reg; P.U.grad_norm = a.cfg.grad_norm;
    P.U.lambda = a.cfg.lambda; P.U.eps = a.cfg.eps; P.U.rho = a.cfg.rho; P.U.beta1 = a.cfg.beta1; P.U.beta2 = a.cfg.beta2;
    P.U.alpha = a.cfg.alpha; P.U.beta = a.cfg.beta; P.U.lambda1 = a.cfg.lambda1; P.U.lambda2 = a.cfg.lambda2;
    P.first_seg = first_seg; P.n_segs = n_segs; P.mini_batch = a.cfg.mini_batch;
    for (int s = 0; s < n_segs; s++) { P.eta[s] = eta[s]; P.adam_alpha[s] = adam_alpha[s]; P.rows[s] = rows[s]; }
    const size_t model_bytes = (size_t)a.cfg.dims * a.rec_width * sizeof(float);
    size_t dyn = 0;
    if (model_bytes <= kSmallModelMax) {
        static std::atomic<bool> attr_set[64];   // > 48 KB of dynamic shared memory needs opting in, once per device
        int dev = 0;
        HM_CUDA(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
            HM_CUDA(cudaFuncSetAttribute(linear_small_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallModelMax));
            if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
        }
        P.model_floats = (uint32_t)(model_bytes / sizeof(float));
        dyn = model_bytes;
    }
    HM_LAUNCH(L, K_LIN_UPDATE, 24.0 * (double)nnz_hint, (linear_small_batch_kernel<<<1, kUpdThreads, dyn, L.stream>>>(P)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_export_count(Launcher &L, ExportScratch &es, const WeightView &wv, const uint8_t *touched,
                              int64_t dims, int64_t *d_total)
{
    const int64_t ctas = (dims + kExpTile - 1) / kExpTile;
    if (ctas + 1 > es.cap) {
        if (es.counts) cudaFree(es.counts);
        es.counts = nullptr;
        HM_CUDA(cudaMalloc(&es.counts, (size_t)(ctas + 1) * sizeof(uint32_t)));
        es.cap = ctas + 1;
    }
    HM_LAUNCH(L, K_EXPORT, 4.0 * (double)dims,
              (export_count_kernel<<<(unsigned)ctas, kExpThreads, 0, L.stream>>>(wv, touched, dims, es.counts)));
    HM_LAUNCH(L, K_EXPORT, 0.0, (export_scan_kernel<<<1, 1024, 0, L.stream>>>(es.counts, (int)ctas, d_total)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_export_write(Launcher &L, ExportScratch &es, const WeightView &wv, const uint8_t *touched,
                              int64_t dims, int32_t *d_feature, float *d_weight)
{
    const int64_t ctas = (dims + kExpTile - 1) / kExpTile;
    HM_LAUNCH(L, K_EXPORT, 4.0 * (double)dims,
              (export_write_kernel<<<(unsigned)ctas, kExpThreads, 0, L.stream>>>(wv, touched, dims, es.counts,
                                                                                 d_feature, d_weight)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_scale(Launcher &L, float *p, int64_t n, float scale)
{
    if (n <= 0) return HM_OK;
    HM_LAUNCH(L, K_MISC, 8.0 * (double)n,
              (scale_kernel<<<grid_for((n + 1023) / 1024, 8), 256, 0, L.stream>>>(p, n, scale)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_strided_copy(Launcher &L, float *dense, float *strided, int64_t n, int stride, int off, int width,
                              bool gather)
{
    if (n <= 0 || width <= 0) return HM_OK;
    HM_LAUNCH(L, K_MISC, 8.0 * (double)n * width,
              (strided_copy_kernel<<<grid_for((n * width + 1023) / 1024, 8), 256, 0, L.stream>>>(dense, strided, n, stride,
                                                                                                off, width, gather)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_dense_weights(Launcher &L, const WeightView &wv, int64_t dims, float *d_out)
{
    HM_LAUNCH(L, K_MISC, 12.0 * (double)dims,
              (dense_weights_kernel<<<grid_for((dims + 1023) / 1024, 8), 256, 0, L.stream>>>(wv, dims, d_out)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_fill_u32(Launcher &L, uint32_t *p, int64_t n, uint32_t v)
{
    if (n <= 0) return HM_OK;
    HM_LAUNCH(L, K_MISC, 4.0 * (double)n,
              (fill_u32_kernel<<<grid_for((n + 1023) / 1024, 8), 256, 0, L.stream>>>(p, n, v)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

// p[i * stride] = v for i in [0, n)
hm_status launch_fill_strided(Launcher &L, float *p, int64_t n, int stride, float v)
{
    if (n <= 0) return HM_OK;
    HM_LAUNCH(L, K_MISC, 4.0 * (double)n,
              (fill_strided_kernel<<<grid_for((n + 1023) / 1024, 8), 256, 0, L.stream>>>(p, n, stride, v)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

}  // namespace hm

// =====================================================================================
//  hm_model_table_noalloc -- model table with no device allocation
// =====================================================================================
//
// hm_udtf_forward() sizes and allocates its compaction scratch (one counter per 4096
// features) and its output staging with cudaMalloc.  Some hosts of this library must not do
// that: an executor that was handed a fixed slice of a GPU shared with other tasks and has
// already spent it on its own buffers, or a serving loop next to other tenants' streams, for
// which the device-wide synchronisation hidden inside cudaMalloc / cudaFree is a stall it
// cannot afford.  For them the counters live in the file-scope array below, which every GPU
// that loads this library carries until the process exits (64 KiB), and the caller brings
// the input and all outputs.
//
// The array's name is only a host-side handle in host code, so the device address is looked
// up once per device with cudaGetSymbolAddress while that device is current (first call on
// the device; make that call before starting any stream capture) and that address is what
// the kernels receive.

namespace {
constexpr int64_t kNoAllocMaxDims = HM_MODEL_TABLE_NOALLOC_MAX_DIMS;
constexpr int kNoAllocMaxTiles = (int)(kNoAllocMaxDims / hm::kExpTile);
constexpr int kNoAllocMaxDevices = 64;
}  // namespace

static __device__ __align__(256) uint32_t hm_noalloc_tile_counts[kNoAllocMaxTiles + 64];

namespace {

std::mutex g_noalloc_mutex;
uint32_t *g_noalloc_counts[kNoAllocMaxDevices] = {nullptr};

// Same ordered compaction as export_write_kernel, but never writes past `cap` rows: the row
// count is only known on the device, so the guard has to be here.
__global__ void __launch_bounds__(hm::kExpThreads)
export_write_capped_kernel(const float *__restrict__ w, int64_t dims, const uint32_t *__restrict__ offsets,
                           int32_t *__restrict__ feature, float *__restrict__ weight, uint32_t cap)
{
    __shared__ uint32_t warp_tot[hm::kExpThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int per = hm::kExpTile / hm::kExpThreads;
    const int64_t j0 = (int64_t)blockIdx.x * hm::kExpTile + (int64_t)threadIdx.x * per;
    uint32_t flags = 0;
    for (int i = 0; i < per; i++)
        if (j0 + i < dims && w[j0 + i] != 0.0f) flags |= 1u << i;
    const uint32_t c = __popc(flags);
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t pos = offsets[blockIdx.x] + inc - c;
    for (int w2 = 0; w2 < warp; w2++) pos += warp_tot[w2];
    for (int i = 0; i < per; i++)
        if (flags & (1u << i)) {
            if (pos < cap) { feature[pos] = (int32_t)(j0 + i); weight[pos] = w[j0 + i]; }
            pos++;
        }
}

}  // namespace

extern "C" hm_status hm_model_table_noalloc(const float *d_w, int64_t dims, int32_t *d_feature, float *d_weight,
                                            int64_t cap, int64_t *d_n_rows, void *stream)
{
    using namespace hm;
    if (!d_w || !d_n_rows || cap < 0 || (cap > 0 && (!d_feature || !d_weight)))
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "hm_model_table_noalloc: NULL buffer or negative capacity");
    if (dims < 1 || dims > kNoAllocMaxDims)
        HM_FAIL(HM_ERR_INVALID_ARGUMENT,
                "hm_model_table_noalloc: dims = %lld is outside [1, %lld]; the built-in scratch is sized for at most "
                "%lld features (use hm_udtf_forward, which allocates, for larger models)",
                (long long)dims, (long long)kNoAllocMaxDims, (long long)kNoAllocMaxDims);
    int dev = -1;
    HM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) HM_FAIL(HM_ERR_UNSUPPORTED, "hm_model_table_noalloc: device ordinal %d", dev);
    uint32_t *counts = nullptr;
    {
        // once per device, with that device current: the symbol's address on THIS device
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_counts[dev]) {
            void *p = nullptr;
            HM_CUDA(cudaGetSymbolAddress(&p, hm_noalloc_tile_counts));
            g_noalloc_counts[dev] = static_cast<uint32_t *>(p);
        }
        counts = g_noalloc_counts[dev];
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int tiles = (int)((dims + kExpTile - 1) / kExpTile);
    const uint32_t cap32 = (uint32_t)std::min<int64_t>(cap, 0xffffffffLL);
    // counts[0 .. tiles) are all rewritten by the first kernel before anything reads them, and
    // *d_n_rows is written by the second: nothing left over from an earlier call is ever read.
    WeightView plain;
    plain.base = d_w; plain.stride = 1; plain.ftrl = 0; plain.alpha = plain.beta = plain.lambda1 = plain.lambda2 = 0.f;
    export_count_kernel<<<tiles, kExpThreads, 0, st>>>(plain, nullptr, dims, counts);
    export_scan_kernel<<<1, 1024, 0, st>>>(counts, tiles, d_n_rows);
    export_write_capped_kernel<<<tiles, kExpThreads, 0, st>>>(d_w, dims, counts, d_feature, d_weight, cap32);
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

#ifdef HM_STAMP
extern "C" int hm_debug_read_stamps(unsigned long long *host, size_t n)
{
    cudaDeviceSynchronize();
    return (int)cudaMemcpyFromSymbol(host, hm::hm_dbg_stamps, n * sizeof(unsigned long long));
}
extern "C" int hm_debug_read_seq(unsigned long long *host, size_t n)
{
    cudaDeviceSynchronize();
    return (int)cudaMemcpyFromSymbol(host, hm::hm_dbg_seq, n * sizeof(unsigned long long));
}
extern "C" int hm_debug_read_fstamps(unsigned long long *host, size_t n)
{
    cudaDeviceSynchronize();
    return (int)cudaMemcpyFromSymbol(host, hm::hm_dbg_fstamps, n * sizeof(unsigned long long));
}
#endif
