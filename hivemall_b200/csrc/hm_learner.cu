// hm_learner.cu -- host side of the UDTF lifecycle and the extern "C" boundary.
//
// Mirrors the Hive UDTF contract that BASELINE.json's north_star fixes for this path:
// initialize(options) -> process(row)* -> close() -> forward(model rows).  The reference
// mount has no source to cite (README.md:20-22).  What happens between process() and the
// kernels:
//
//   host rows slice --H2D (copy stream; truly async when the caller used hm_host_alloc)--> chunk
//     (plain CSR; or two groups / column matrices / feature strings, which a kernel on the sort
//      stream merges, expands or hashes into the chunk's CSR)
//   chunk = [rows left over from the previous slice | new rows]          (device resident)
//   full mini-batches of the chunk are grouped into blocks (<= kMaxSegs mini-batches,
//   <= kBlockElemCap elements); per block, on the sort stream: hot-feature selection, seg_prep,
//   segmented sort (once, data only; kept for the later passes of -iters > 1), then
//   per mini-batch: linear_forward + linear_update on the compute stream.
//   Rows that do not fill a mini-batch wait at the end of the chunk for the next slice or close().
//
// There is no CPU compute path: every entry point that trains or predicts needs a device.
#include "hm_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <new>

namespace hm {

const char *last_error_cstr();

// ---- Launcher --------------------------------------------------------------------

void Launcher::begin(int id, double alg_bytes)
{
    launches++;
    if (!profiling) return;
    count[id]++;            // per-kernel statistics cover profiled launches only
    bytes[id] += alg_bytes;
    Pending p;
    p.id = id;
    for (cudaEvent_t *e : {&p.a, &p.b}) {
        if (!pool.empty()) { *e = pool.back(); pool.pop_back(); }
        else cudaEventCreate(e);
    }
    cudaEventRecord(p.a, stream);
    pending.push_back(p);
}

void Launcher::end(int)
{
    if (!profiling) return;
    cudaEventRecord(pending.back().b, stream);
    if (pending.size() >= 4096) drain();
}

void Launcher::drain()
{
    if (pending.empty()) return;
    cudaStreamSynchronize(stream);
    for (const Pending &p : pending) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) ms[p.id] += (double)t;
        pool.push_back(p.a);
        pool.push_back(p.b);
    }
    pending.clear();
}

void Launcher::release()
{
    drain();
    for (cudaEvent_t e : pool) cudaEventDestroy(e);
    pool.clear();
}

namespace {

constexpr int64_t kBlockElemCap = (int64_t)1 << 26;   // hard limit of elements sorted per launch set (scratch size)
// Elements per block aimed at.  The sort of block k+1 and the training of block k share the SMs by taking
// turns, and they interleave better in short turns: measured on configs[1] (64 mini-batches of 2.56 M elements
// per call): 67 M (26 mini-batches per block) 489 M examples/s, 33 M 500 M, 26 M 505 M, 21 M (8 per block) 513 M,
// 16.8 M 509 M, 8.4 M 457 M.
#ifndef HM_BLOCK_TARGET
#define HM_BLOCK_TARGET 21000000
#endif
constexpr int64_t kBlockElemTarget = HM_BLOCK_TARGET;
// Host rows copied per H2D slice: the first slices of a call are kSliceRows long, the last ones shrink
// to kMinSliceRows (see hm_udtf_process).  With one fixed size, measured on B200 (bench.py e2e,
// mini-batch 65536): 2^21 -> 127 M rows/s, 2^20 -> 142, 2^19 -> 151, 2^18 -> 154; with the descending
// schedule 159 M for (2^20 | 2^21, 2^16 | 2^17), 157.5 M for (2^20, 2^18), 155.5 M for (2^22, 2^17).
constexpr int64_t kSliceRows = (int64_t)1 << 19;
constexpr int64_t kMinSliceRows = (int64_t)1 << 17;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    hm_status reserve(size_t bytes)
    {
        if (bytes <= cap) return HM_OK;
        if (p) cudaFree(p);   // cudaFree waits for work that may still use the old buffer
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { cudaGetLastError(); want = bytes; e = cudaMalloc(&p, want); }
        if (e != cudaSuccess) { cudaGetLastError(); p = nullptr; HM_FAIL(HM_ERR_OOM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e)); }
        cap = want;
        return HM_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// The sort result of one ingest block, kept for the later passes of -iters > 1: the sort depends
// on the rows only, so a replay of retained rows needs no sort_* launch at all.
struct SortedBlock {
    int64_t row0 = 0, rows = 0, nnz = 0;
    int n_segs = 0;
    void *mem = nullptr;                // one allocation: key | pay | seg_off | hot
    uint32_t *key = nullptr; uint2 *pay = nullptr; uint32_t *seg_off = nullptr; HotInfo *hot = nullptr;
    bool hot_on = false;
};

// A run of rows resident on the device.  Owned chunks hold copies of host data (rowptr is
// 0-based inside the chunk); views alias caller memory (hm_udtf_process_device).
struct Chunk {
    bool view = false;
    std::vector<SortedBlock> sorted;    // -iters > 1: the blocks of the first pass, in row order
    bool sorted_valid = true;           // false: a block could not be kept (memory): replays sort again
    int64_t rows = 0;                   // rows whose mini-batches are listed in mb_abs
    DevBuf rowptr, feat, val, label;    // owned storage
    DevBuf sq_rp, sc_rp, sq_feat, sq_val, sc_feat;   // split pushes: the two groups as they arrive, merged on the device
    DevBuf str_bytes, str_off;                       // string pushes: the feature strings as they arrive, hashed on the device
    const int64_t *v_rowptr = nullptr;  // view storage
    const uint32_t *v_feat = nullptr;
    const float *v_val = nullptr, *v_label = nullptr;
    std::vector<int64_t> mb_abs;        // rowptr value at every mini-batch boundary of rows [0, rows)
    cudaEvent_t free_ev = nullptr;      // recorded on the compute stream when the chunk may be overwritten
    const int64_t *rp() const { return view ? v_rowptr : rowptr.as<int64_t>(); }
    const uint32_t *ft() const { return view ? v_feat : feat.as<uint32_t>(); }
    const float *vl() const { return view ? v_val : val.as<float>(); }
    const float *lb() const { return view ? v_label : label.as<float>(); }
    void drop_sorted()
    {
        for (SortedBlock &b : sorted) if (b.mem) cudaFree(b.mem);
        sorted.clear();
    }
    void release()
    {
        rowptr.release(); feat.release(); val.release(); label.release();
        sq_rp.release(); sc_rp.release(); sq_feat.release(); sq_val.release(); sc_feat.release();
        str_bytes.release(); str_off.release();
        drop_sorted();
        if (free_ev) cudaEventDestroy(free_ev);
        free_ev = nullptr;
    }
};

// One slice of host rows: plain CSR (value may be NULL = all ones), or split into a quantitative group
// (index and value) and an index-only group (value 1.0); row i = its quantitative elements, then its
// index-only elements; or feature STRINGS ("name:value"), one per element, hashed on the device:
// rowptr then indexes strings, string s is s_bytes[s_offsets[s] .. s_offsets[s+1]).
struct HostRows {
    int64_t n = 0;
    const int64_t *rowptr = nullptr; const int32_t *feature = nullptr; const float *value = nullptr;
    bool strings = false;
    const uint8_t *s_bytes = nullptr; const int64_t *s_offsets = nullptr;
    // columns: every row = n_q quantitative features whose indices are the same in every row (q_cols[n_q], given
    // once) with a value each (cq_value[n][n_q]), then n_c index-only features (cc_feature[n][n_c]); no row pointer
    bool columns = false;
    int32_t n_q = 0, n_c = 0;
    const int32_t *q_cols = nullptr, *cc_feature = nullptr;
    const float *cq_value = nullptr;
    bool split = false;
    const int64_t *q_rowptr = nullptr, *c_rowptr = nullptr;
    const int32_t *q_feature = nullptr, *c_feature = nullptr;
    const float *q_value = nullptr;
    const float *label = nullptr;
    int64_t nnz() const
    {
        if (columns) return n * (int64_t)(n_q + n_c);
        return split ? (q_rowptr[n] - q_rowptr[0]) + (c_rowptr[n] - c_rowptr[0]) : rowptr[n] - rowptr[0];
    }
    int64_t row_end(int64_t i) const   // elements before row i, relative to the slice
    {
        return split ? (q_rowptr[i] - q_rowptr[0]) + (c_rowptr[i] - c_rowptr[0]) : rowptr[i] - rowptr[0];
    }
};

// Split rows -> one CSR on the device: a warp per row copies the quantitative elements, then the
// index-only ones with value 1.0, and writes the row pointer.  Only the bytes that carry information
// crossed the bus.
__global__ void __launch_bounds__(256)
merge_split_kernel(int64_t n, const int64_t *__restrict__ q_rp, int64_t q_base, const int64_t *__restrict__ c_rp,
                   int64_t c_base, const uint32_t *__restrict__ q_feat, const float *__restrict__ q_val,
                   const uint32_t *__restrict__ c_feat, int64_t out_base, int64_t *__restrict__ out_rp,
                   uint32_t *__restrict__ out_feat, float *__restrict__ out_val)
{
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < n; row += warps) {
        const int64_t q0 = q_rp[row] - q_base, q1 = q_rp[row + 1] - q_base;
        const int64_t c0 = c_rp[row] - c_base, c1 = c_rp[row + 1] - c_base;
        const int64_t o = out_base + q0 + c0;
        for (int64_t i = lane; i < q1 - q0; i += 32) { out_feat[o + i] = q_feat[q0 + i]; out_val[o + i] = q_val[q0 + i]; }
        const int64_t o2 = o + (q1 - q0);
        for (int64_t i = lane; i < c1 - c0; i += 32) { out_feat[o2 + i] = c_feat[c0 + i]; out_val[o2 + i] = 1.0f; }
        if (lane == 0) {
            out_rp[row] = o;
            if (row == n - 1) out_rp[n] = out_base + q1 + c1;
        }
    }
}

// Column rows -> one CSR on the device: element e of the slice belongs to row e / (n_q + n_c); the first n_q
// elements of a row take their index from q_cols and their value from the value matrix, the others are
// index-only (value 1.0).  Fully regular: a thread per element, every access coalesced.
__global__ void __launch_bounds__(256)
expand_columns_kernel(int64_t n, int32_t n_q, int32_t n_c, const uint32_t *__restrict__ q_cols,
                      const float *__restrict__ q_val, const uint32_t *__restrict__ c_feat, int64_t out_base,
                      int64_t *__restrict__ out_rp, uint32_t *__restrict__ out_feat, float *__restrict__ out_val)
{
    const int64_t per = (int64_t)n_q + n_c, total = n * per;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t row = e / per;
        const int32_t j = (int32_t)(e - row * per);
        if (j < n_q) { out_feat[out_base + e] = q_cols[j]; out_val[out_base + e] = q_val[row * n_q + j]; }
        else { out_feat[out_base + e] = c_feat[row * n_c + (j - n_q)]; out_val[out_base + e] = 1.0f; }
    }
    for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row <= n; row += stride) out_rp[row] = out_base + row * per;
}

__global__ void gather_rowptr_kernel(const int64_t *__restrict__ rowptr, int64_t n_rows, int32_t step,
                                     int64_t n_out, int64_t *__restrict__ out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_out; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i * step;
        if (r > n_rows) r = n_rows;
        out[i] = rowptr[r];
    }
}

int bits_for(int64_t dims)
{
    int b = 0;
    while (((int64_t)1 << b) < dims) b++;
    return b < 1 ? 1 : b;
}

}  // namespace
}  // namespace hm

using namespace hm;

struct hm_udtf {
    hm_config cfg;
    int device = 0;
    Launcher L, Ls, Lc;              // compute stream / sort stream / copy stream (fills only)
    cudaEvent_t merged_ev = nullptr;        // split pushes: the merge kernel of the latest slice has run
    int64_t hint_rows = 0, hint_elems = 0, hint_q = 0, hint_c = 0, hint_bytes = 0;   // largest slice of the running process() call: chunk buffers are sized once
    cudaStream_t compute = nullptr, copy = nullptr, sorter = nullptr;
    cudaEvent_t sorted_ev[2] = {nullptr, nullptr}, trained_ev[2] = {nullptr, nullptr}, ready_ev = nullptr;
    bool trained_valid[2] = {false, false};
    size_t sort_turn = 0;            // blocks sorted so far: picks the scratch set, across calls
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};
    // model replica
    float *rec = nullptr;            // dims x rec_width floats: (w[, slot1[, slot2, unused]]) per feature
    int rec_width = 1;
    int state_width = 0;             // optimizer slots per feature (0, 1 or 2)
    // train_fm replica
    bool fm = false;
    float *w0rec = nullptr, *wrec = nullptr, *vrec = nullptr;
    uint8_t *touched = nullptr;
    int ww = 1, vw = 0, k = 0;
    FmScratch fms;
    DevBuf stab, exp_vif;
    int64_t t = 0;
    // scratch
    SortScratch sort[2];             // block k+1 is sorted (sort stream) while block k trains (compute stream)
    UpdateScratch upd;
    ExportScratch exp;
    DevBuf sqnorm;        // online learners: |x|^2 per row of the block being trained
    DevBuf dloss2;        // variance-keeping learners: the variance pass's per-row coefficients
    DevBuf margin, dloss, mb_gather, exp_feat, exp_weight, tmp_a, tmp_b, tmp_c, tmp_d;
    uint32_t *d_bad = nullptr;       // [0]: a training row held an index outside [0, dims); [1]: a scored row did
    int64_t *d_total = nullptr;
    // data
    std::vector<Chunk *> retained;   // pass-1 chunks kept for -iters > 1
    std::vector<Chunk *> pool;       // recycled chunk storage (oldest first)
    int chunks_alive = 0;            // owned chunks in circulation (pool + tail + in flight)
    Chunk *tail = nullptr;           // chunk whose last rows do not fill a mini-batch yet
    int64_t tail_row0 = 0;           // first untrained row of `tail`
    std::vector<int64_t> tail_rp;    // host copy of the untrained rows' rowptr, 0-based (size tail_rows+1)
    int64_t *h_stage[2] = {nullptr, nullptr};
    int64_t h_stage_cap[2] = {0, 0};
    int stage_turn = 0;
    int64_t *h_small = nullptr;      // pinned scratch for small D2H reads
    int64_t h_small_cap = 0;
    int64_t last_margin_off = 0, last_margin_n = 0;   // last mini-batch's slice of `margin`
    bool device_rows_short = false;  // a process_device call ended on a short mini-batch
    bool hot_enabled = true;         // HM_HOT_FEATURES=0 in the environment sorts every element (A/B measurements)
    hm_stats stats;
    bool closed = false;
    int64_t model_rows = -1;

    int64_t tail_rows() const { return tail_rp.empty() ? 0 : (int64_t)tail_rp.size() - 1; }
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; ok = false; cudaGetLastError(); }
        if (cudaSetDevice(dev) != cudaSuccess) { ok = false; cudaGetLastError(); }
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

hm_status ensure_small(hm_udtf *u, int64_t n)
{
    if (n <= u->h_small_cap) return HM_OK;
    if (u->h_small) cudaFreeHost(u->h_small);
    u->h_small = nullptr; u->h_small_cap = 0;
    HM_CUDA(cudaMallocHost(&u->h_small, (size_t)(n + 64) * sizeof(int64_t)));
    u->h_small_cap = n + 64;
    return HM_OK;
}

float eta_at(const hm_config &c, int64_t t)
{
    switch (c.eta_sched) {
    case HM_ETA_INVERSE: return (float)((double)c.eta0 / std::pow((double)t, (double)c.power_t));
    case HM_ETA_SIMPLE: return (float)((double)c.eta0 / (1.0 + (double)t / c.total_steps));
    default: return c.eta0;
    }
}

float adam_alpha_at(const hm_config &c, float eta, int64_t t)
{
    const double b1 = std::pow((double)c.beta1, (double)t), b2 = std::pow((double)c.beta2, (double)t);
    return (float)((double)eta * std::sqrt(1.0 - b2) / (1.0 - b1));
}

FmArgs fm_args(const hm_udtf *u)
{
    FmArgs a;
    a.cfg = u->cfg; a.w0rec = u->w0rec; a.wrec = u->wrec; a.vrec = u->vrec; a.touched = u->touched;
    a.ww = u->ww; a.vw = u->vw; a.k = u->k;
    return a;
}

// where the weights are read from (forward, prediction, model table)
WeightView weight_view(const hm_udtf *u)
{
    WeightView v;
    v.base = u->fm ? u->wrec : u->rec;
    v.stride = u->fm ? u->ww : u->rec_width;
    v.ftrl = (!u->fm && u->cfg.opt == HM_OPT_FTRL) ? 1 : 0;
    v.alpha = u->cfg.alpha; v.beta = u->cfg.beta; v.lambda1 = u->cfg.lambda1; v.lambda2 = u->cfg.lambda2;
    return v;
}

// Train rows [row0, row0+n_rows) of a chunk.  mb_abs[i] is the chunk rowptr value at
// row0 + min(i*B, n_rows), i = 0..n_mb.
//
// The rows are cut into blocks (<= kMaxSegs mini-batches, <= kBlockElemCap elements).  Sorting a
// block needs only the data, so block k+1 is sorted on the sort stream while block k trains on
// the compute stream (two scratch sets, two events each way).
//
// -iters > 1: the first pass keeps every block's sort result with the chunk (kTrainKeep) when device memory
// allows, and the later passes (kTrainReplay, rows [0, c.rows) with the chunk's own mb_abs) train straight
// from those: no sort_* kernel is launched again.
enum TrainMode { kTrainPlain = 0, kTrainKeep = 1, kTrainReplay = 2 };
constexpr size_t kKeepFreeFloor = (size_t)6 << 30;     // stop keeping sort results when less than this stays free

hm_status train_rows(hm_udtf *u, Chunk &c, int64_t row0, int64_t n_rows, const int64_t *mb_abs, int64_t n_mb,
                     bool rows_ready = false, TrainMode mode = kTrainPlain)
{
    if (n_rows <= 0) return HM_OK;
    const int32_t B = u->cfg.mini_batch;
    const int key_bits = bits_for(u->cfg.dims);
    LinearArgs a;
    a.cfg = u->cfg; a.rec = u->rec; a.rec_width = u->rec_width; a.wv = weight_view(u);
    struct Plan { int64_t m0, m1, row0, rows, nnz; int n_segs; };
    std::vector<Plan> plan;
    // A replay trains from the kept blocks when they tile the chunk's rows, in the first pass's order.
    bool replay = false;
    if (mode == kTrainReplay && c.sorted_valid && !c.sorted.empty() && row0 == 0) {
        int64_t next = 0;
        for (const SortedBlock &b : c.sorted) { if (b.row0 != next) { next = -1; break; } next += b.rows; }
        replay = next == n_rows;
    }
    if (replay) {
        for (const SortedBlock &b : c.sorted) {
            Plan p;
            p.m0 = b.row0 / B; p.m1 = p.m0 + b.n_segs; p.n_segs = b.n_segs;     // (only a chunk's last mini-batch is short)
            p.row0 = b.row0; p.rows = b.rows; p.nnz = b.nnz;
            plan.push_back(p);
        }
        // the kept blocks were written by copies on the sort stream
        HM_CUDA(cudaEventRecord(u->ready_ev, u->sorter));
        HM_CUDA(cudaStreamWaitEvent(u->compute, u->ready_ev, 0));
    } else {
    // Blocks of about kBlockElemTarget elements, evenly sized: the call's elements are cut into
    // ceil(total / target) blocks at the mini-batch boundaries nearest to equal shares (never more than
    // kMaxSegs mini-batches or kBlockElemCap elements, at least one mini-batch).
    const int64_t total_elems = mb_abs[n_mb] - mb_abs[0];
    const int64_t n_blocks = std::max<int64_t>(1, (total_elems + kBlockElemTarget - 1) / kBlockElemTarget);
    int64_t block_no = 0;
    for (int64_t m0 = 0; m0 < n_mb;) {
        block_no++;
        const int64_t share_end = mb_abs[0] + (int64_t)((double)total_elems * (double)block_no / (double)n_blocks);
        int64_t m1 = m0 + 1;
        while (m1 < n_mb && m1 - m0 < kMaxSegs && mb_abs[m1 + 1] - mb_abs[m0] <= kBlockElemCap &&
               (block_no >= n_blocks || mb_abs[m1 + 1] <= share_end + (mb_abs[m1 + 1] - mb_abs[m1]) / 2))
            m1++;
        Plan p;
        p.m0 = m0; p.m1 = m1; p.n_segs = (int)(m1 - m0);
        p.row0 = row0 + m0 * B;
        p.rows = std::min<int64_t>(n_rows - m0 * B, (int64_t)p.n_segs * B);
        p.nnz = mb_abs[m1] - mb_abs[m0];
        if (p.nnz >= ((int64_t)1 << 31))
            HM_FAIL(HM_ERR_UNSUPPORTED, "a mini-batch of %lld elements exceeds the 2^31 block limit", (long long)p.nnz);
        plan.push_back(p);
        m0 = m1;
    }
    }
    // Rows that were copied in by this library: everything queued on the compute stream so far
    // (their copies included) precedes the sorts.  Rows that are already resident (device views,
    // replays of retained chunks) need no such wait: the sort only reads data, so it may start
    // while the previous call's last block is still training or being averaged.
    if (!rows_ready || u->L.profiling) {
        HM_CUDA(cudaEventRecord(u->ready_ev, u->compute));
        HM_CUDA(cudaStreamWaitEvent(u->sorter, u->ready_ev, 0));
    }
    int out_buf[2] = {0, 0};
    std::vector<uint32_t> seg_host[2];      // while profiling: the sorted arrays' mini-batch offsets (byte accounting)
    int64_t call_max_mb = 0;
    for (int64_t m = 0; m < n_mb; m++) call_max_mb = std::max(call_max_mb, mb_abs[m + 1] - mb_abs[m]);
    // Hot features (the quantitative features every row has, the head of each categorical field) are taken out
    // of the sort and reduced by the forward kernel: linear learners on the two-kernel path only.
    // (the variance-keeping learners take the plain two-kernel path: no hot features, no persistent small-batch kernel)
    const bool cov = !u->fm && loss_keeps_variance(u->cfg.loss);
    const bool small_path = !u->fm && !cov && small_batch_eligible(B, call_max_mb);
    const bool hot_on = !u->fm && !cov && u->hot_enabled && !small_path;
    auto enqueue_sort = [&](size_t bi) -> hm_status {
        const Plan &p = plan[bi];
        const int set = (int)((u->sort_turn + bi) & 1);
        // the scratch set is free once the block that used it last has been trained
        if (u->trained_valid[set]) HM_CUDA(cudaStreamWaitEvent(u->sorter, u->trained_ev[set], 0));
        HM_TRY(sort_scratch_reserve(u->sort[set], p.nnz));
        HM_TRY(launch_hot_select(u->Ls, u->sort[set], c.rp(), p.row0, p.rows, c.ft(), u->cfg.dims, hot_on));
        HM_TRY(launch_seg_prep(u->Ls, u->sort[set], c.rp(), p.row0, p.rows, p.nnz, B, p.n_segs));
        HM_TRY(launch_segmented_sort(u->Ls, u->sort[set], c.ft() + mb_abs[p.m0], c.vl() + mb_abs[p.m0], c.rp(), p.row0,
                                     p.rows, B, p.nnz, p.n_segs, key_bits, u->cfg.dims, u->d_bad, &out_buf[set], 0,
                                     &seg_host[set]));
        HM_CUDA(cudaEventRecord(u->sorted_ev[set], u->sorter));
        if (mode == kTrainKeep && c.sorted_valid) {
            // keep the result for the later passes (device-to-device, on the sort stream, behind the event
            // training waits for)
            const size_t key_b = ((size_t)p.nnz * 4 + 255) & ~(size_t)255, pay_b = ((size_t)p.nnz * 8 + 255) & ~(size_t)255;
            const size_t off_b = ((size_t)(p.n_segs + 1) * 4 + 255) & ~(size_t)255;
            const size_t need = key_b + pay_b + off_b + sizeof(HotInfo);
            size_t free_b = 0, total_b = 0;
            SortedBlock sb;
            if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && free_b > need + kKeepFreeFloor &&
                cudaMalloc(&sb.mem, need) == cudaSuccess) {
                char *base = static_cast<char *>(sb.mem);
                sb.key = reinterpret_cast<uint32_t *>(base);
                sb.pay = reinterpret_cast<uint2 *>(base + key_b);
                sb.seg_off = reinterpret_cast<uint32_t *>(base + key_b + pay_b);
                sb.hot = reinterpret_cast<HotInfo *>(base + key_b + pay_b + off_b);
                sb.row0 = p.row0; sb.rows = p.rows; sb.nnz = p.nnz; sb.n_segs = p.n_segs; sb.hot_on = hot_on;
                const SortScratch &ss = u->sort[set];
                // (the sorted arrays are at most nnz long; the tail past the sorted length is never read)
                HM_CUDA(cudaMemcpyAsync(sb.key, ss.key[out_buf[set]], (size_t)p.nnz * 4, cudaMemcpyDeviceToDevice, u->sorter));
                HM_CUDA(cudaMemcpyAsync(sb.pay, ss.pay[out_buf[set]], (size_t)p.nnz * 8, cudaMemcpyDeviceToDevice, u->sorter));
                HM_CUDA(cudaMemcpyAsync(sb.seg_off, ss.seg_off, (size_t)(p.n_segs + 1) * 4, cudaMemcpyDeviceToDevice, u->sorter));
                HM_CUDA(cudaMemcpyAsync(sb.hot, ss.hot, sizeof(HotInfo), cudaMemcpyDeviceToDevice, u->sorter));
                c.sorted.push_back(sb);
            } else {
                cudaGetLastError();
                c.sorted_valid = false;      // not enough memory: this chunk's replays sort again
                c.drop_sorted();
            }
        }
        return HM_OK;
    };
    // While per-kernel profiling is on, the two streams are serialised so that every kernel is
    // timed alone (overlapped kernels stretch each other's event-to-event time).
    const bool overlap = !u->L.profiling;
    if (!replay) HM_TRY(enqueue_sort(0));
    for (size_t bi = 0; bi < plan.size(); bi++) {
        const Plan &p = plan[bi];
        const int set = (int)((u->sort_turn + bi) & 1);
        if (!replay && bi + 1 < plan.size()) {
            if (!overlap) {     // the next sort starts only after everything queued for training so far
                HM_CUDA(cudaEventRecord(u->ready_ev, u->compute));
                HM_CUDA(cudaStreamWaitEvent(u->sorter, u->ready_ev, 0));
            }
            HM_TRY(enqueue_sort(bi + 1));
            if (!overlap) {     // ... and training waits for it
                HM_CUDA(cudaStreamWaitEvent(u->compute, u->sorted_ev[(u->sort_turn + bi + 1) & 1], 0));
            }
        }
        if (!replay) HM_CUDA(cudaStreamWaitEvent(u->compute, u->sorted_ev[set], 0));
        // where this block's sorted arrays are: the scratch set, or the block kept by the first pass
        SortScratch view;
        const int buf = replay ? 0 : out_buf[set];
        if (replay) {
            const SortedBlock &sb = c.sorted[bi];
            view.key[0] = sb.key; view.pay[0] = sb.pay; view.seg_off = sb.seg_off; view.hot = sb.hot;
        }
        SortScratch &ss = replay ? view : u->sort[set];
        const bool block_hot = replay ? c.sorted[bi].hot_on : hot_on;
        HM_TRY(u->margin.reserve((size_t)p.rows * sizeof(float)));
        if (!u->fm && loss_needs_sqnorm(u->cfg.loss)) {
            HM_TRY(u->sqnorm.reserve((size_t)p.rows * sizeof(float)));
            HM_TRY(launch_row_sqnorm(u->L, c.rp(), p.row0, p.rows, c.ft(), c.vl(), u->cfg.dims, u->sqnorm.as<float>(), p.nnz));
            a.sqnorm = u->sqnorm.as<float>();
        }
        HM_TRY(u->dloss.reserve((size_t)std::min<int64_t>(p.rows, B) * sizeof(float)));
        if (cov) HM_TRY(u->dloss2.reserve((size_t)std::min<int64_t>(p.rows, B) * sizeof(float)));
        if (u->fm) {
            int64_t max_mb = 0;
            for (int64_t m = p.m0; m < p.m1; m++) max_mb = std::max(max_mb, mb_abs[m + 1] - mb_abs[m]);
            HM_TRY(fm_scratch_reserve(u->fms, max_mb, u->k));
            HM_TRY(u->stab.reserve((size_t)std::min<int64_t>(p.rows, B) * u->k * sizeof(float)));
        }
        HotCols hc;
        if (block_hot) { hc.info = ss.hot; hc.sum = u->upd.hot_sum; hc.cnt = u->upd.hot_cnt; hc.stride = u->upd.hot_stride; }
        if (small_path) {
            // a few hundred rows per mini-batch: up to 64 mini-batches per launch of one persistent CTA
            for (int s0 = 0; s0 < p.n_segs; s0 += 64) {
                const int ns = std::min(64, p.n_segs - s0);
                int32_t rows[64]; float eta[64], adam[64];
                for (int s = 0; s < ns; s++) {
                    const int64_t mrow0 = p.row0 + (int64_t)(s0 + s) * B;
                    rows[s] = (int32_t)std::min<int64_t>(B, row0 + n_rows - mrow0);
                    u->t += 1;
                    eta[s] = eta_at(u->cfg, u->t);
                    adam[s] = u->cfg.opt == HM_OPT_ADAM ? adam_alpha_at(u->cfg, eta[s], u->t) : 0.f;
                    u->stats.mini_batches++;
                    u->last_margin_off = mrow0 - p.row0;
                    u->last_margin_n = rows[s];
                }
                HM_TRY(launch_linear_small_batches(u->L, a, c.rp(), p.row0, s0, ns, rows, eta, adam, c.ft(), c.vl(), c.lb(),
                                                   u->margin.as<float>(), u->dloss.as<float>(), u->d_bad, ss.key[buf],
                                                   ss.pay[buf], ss.seg_off, mb_abs[p.m0 + s0 + ns] - mb_abs[p.m0 + s0]));
            }
        } else
        for (int s = 0; s < p.n_segs; s++) {
            const int64_t mrow0 = p.row0 + (int64_t)s * B;
            const int32_t mrows = (int32_t)std::min<int64_t>(B, row0 + n_rows - mrow0);
            const int64_t mnnz = mb_abs[p.m0 + s + 1] - mb_abs[p.m0 + s];
            u->t += 1;
            StepScalars sc;
            sc.eta = eta_at(u->cfg, u->t);
            sc.adam_alpha = u->cfg.opt == HM_OPT_ADAM ? adam_alpha_at(u->cfg, sc.eta, u->t) : 0.f;
            sc.n_rows = mrows;
            if (u->fm) {
                const FmArgs fa = fm_args(u);
                HM_TRY(launch_fm_forward(u->L, fa, c.rp(), mrow0, mrows, p.row0, c.ft(), c.vl(), c.lb(),
                                         u->margin.as<float>(), u->dloss.as<float>(), u->stab.as<float>(), u->d_bad, mnnz,
                                         true));
                HM_TRY(launch_fm_update(u->L, fa, sc, u->fms, ss.key[buf], ss.pay[buf], ss.seg_off, s,
                                        u->dloss.as<float>(), u->stab.as<float>(), mnnz));
            } else if (cov) {
                // margin + variance per row, then one reduce-and-step launch for the weights and one for the
                // variances over the same sorted mini-batch (both read the variances of its start)
                HM_TRY(launch_linear_forward_cov(u->L, a, c.rp(), mrow0, mrows, p.row0, c.ft(), c.vl(), c.lb(),
                                                 u->margin.as<float>(), u->dloss.as<float>(), u->dloss2.as<float>(),
                                                 u->d_bad, mnnz));
                for (int pass = 1; pass <= 2; pass++) {
                    a.cov_pass = pass;
                    HM_TRY(launch_linear_update(u->L, a, sc, u->upd, ss.key[buf], ss.pay[buf], ss.seg_off, s,
                                                mb_abs[p.m0 + s] - mb_abs[p.m0],
                                                pass == 1 ? u->dloss.as<float>() : u->dloss2.as<float>(), mnnz, hc, -1));
                }
                a.cov_pass = 0;
            } else {
                HM_TRY(launch_linear_forward(u->L, a, c.rp(), mrow0, mrows, p.row0, c.ft(), c.vl(), c.lb(),
                                             u->margin.as<float>(), u->dloss.as<float>(), u->d_bad, mnnz, hc));
                HM_TRY(launch_linear_update(u->L, a, sc, u->upd, ss.key[buf], ss.pay[buf], ss.seg_off, s,
                                            mb_abs[p.m0 + s] - mb_abs[p.m0], u->dloss.as<float>(), mnnz, hc,
                                            (!replay && (int)seg_host[set].size() == p.n_segs + 1)
                                                ? (int64_t)(seg_host[set][s + 1] - seg_host[set][s]) : -1));
            }
            u->stats.mini_batches++;
            u->last_margin_off = mrow0 - p.row0;
            u->last_margin_n = mrows;
        }
        if (!replay) {
            HM_CUDA(cudaEventRecord(u->trained_ev[set], u->compute));
            u->trained_valid[set] = true;
        }
        u->stats.rows += p.rows;
        u->stats.nnz += p.nnz;
    }
    if (!replay) u->sort_turn += plan.size();
    return HM_OK;
}

// A chunk for the next slice of host rows.  Prefer a recycled chunk the device is already done with;
// otherwise open another one while fewer than kMaxChunks circulate (so that the copy stream can run
// ahead of training by more than one slice when the two are equally fast); otherwise take the one
// that was retired first (the copy stream will wait for its free_ev).
constexpr int kMaxChunks = 4;

Chunk *acquire_chunk(hm_udtf *u)
{
    for (size_t i = 0; i < u->pool.size(); i++) {
        Chunk *c = u->pool[i];
        if (!c->free_ev || cudaEventQuery(c->free_ev) == cudaSuccess) {
            u->pool.erase(u->pool.begin() + (long)i);
            return c;
        }
    }
    cudaGetLastError();   // (cudaErrorNotReady from the queries is not an error)
    if (u->pool.empty() || u->chunks_alive < kMaxChunks) {
        Chunk *c = new (std::nothrow) Chunk();
        if (c) u->chunks_alive++;
        return c;
    }
    Chunk *c = u->pool.front();
    u->pool.erase(u->pool.begin());
    return c;
}

// The compute stream is (or will be, in stream order) done with `c`: recycle or retain it.
hm_status retire_chunk(hm_udtf *u, Chunk *c, bool retain)
{
    if (retain) { u->retained.push_back(c); return HM_OK; }
    if (c->view) { delete c; return HM_OK; }
    if (!c->free_ev) HM_CUDA(cudaEventCreateWithFlags(&c->free_ev, cudaEventDisableTiming));
    HM_CUDA(cudaEventRecord(c->free_ev, u->compute));
    c->rows = 0;
    c->mb_abs.clear();
    c->drop_sorted();          // (cudaFree waits for the device: only ever non-empty with -iters > 1)
    c->sorted_valid = true;
    u->pool.push_back(c);
    return HM_OK;
}

// A chunk that was acquired for a slice which then turned out to be malformed: nothing was trained
// from it, copies into it may be queued.  Back to the pool; its next user waits on the copy stream.
void abandon_chunk(hm_udtf *u, Chunk *c)
{
    if (!c->free_ev && cudaEventCreateWithFlags(&c->free_ev, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); c->free_ev = nullptr; }
    if (c->free_ev) cudaEventRecord(c->free_ev, u->copy);
    c->rows = 0;
    c->mb_abs.clear();
    u->pool.push_back(c);
}

// One slice of host rows: build [tail | slice] on the device, train the full mini-batches.
hm_status process_slice(hm_udtf *u, const HostRows &h)
{
    const int32_t B = u->cfg.mini_batch;
    const int64_t n = h.n;
    const int64_t t_rows = u->tail_rows();
    const int64_t t_nnz = t_rows ? u->tail_rp.back() : 0;
    const int64_t nnz = h.nnz();
    const int64_t rows = t_rows + n, elems = t_nnz + nnz;
    if (nnz < 0) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "rowptr is not non-decreasing");

    Chunk *c = acquire_chunk(u);
    if (!c) HM_FAIL(HM_ERR_OOM, "out of host memory");
    // A recycled chunk may still be read by queued compute work: make the copy stream wait for it.
    if (c->free_ev) HM_CUDA(cudaStreamWaitEvent(u->copy, c->free_ev, 0));
    // (growing a buffer frees the old one; cudaFree itself waits for outstanding work)
    // (the slices of a call grow: size the buffers for the largest one right away)
    const int64_t cap_rows = std::max(rows, u->hint_rows), cap_elems = std::max(elems, u->hint_elems);
    HM_TRY(c->rowptr.reserve((size_t)(cap_rows + 1) * sizeof(int64_t)));
    HM_TRY(c->feat.reserve((size_t)std::max<int64_t>(cap_elems, 1) * sizeof(uint32_t)));
    HM_TRY(c->val.reserve((size_t)std::max<int64_t>(cap_elems, 1) * sizeof(float)));
    HM_TRY(c->label.reserve((size_t)std::max<int64_t>(cap_rows, 1) * sizeof(float)));

    // new rows: host-to-device on the copy stream, queued before the host-side loop below so that the
    // bus is busy while it runs
    int64_t str_b0 = 0, str_nbytes = 0;
    if (h.strings) {
        // the strings of the slice: their bytes and their (absolute) offsets; the hash kernel writes the
        // chunk's feature / value arrays (below, on the sort stream)
        const int64_t s0 = h.rowptr[0];
        str_b0 = h.s_offsets[s0];
        str_nbytes = h.s_offsets[s0 + nnz] - str_b0;
        if (str_nbytes < 0) { abandon_chunk(u, c); HM_FAIL(HM_ERR_INVALID_ARGUMENT, "string offsets are not non-decreasing"); }
        HM_TRY(c->str_bytes.reserve((size_t)std::max<int64_t>(std::max(str_nbytes, u->hint_bytes), 1) + 64));
        HM_TRY(c->str_off.reserve((size_t)(std::max<int64_t>(nnz, u->hint_elems) + 1) * 8));
        if (nnz) {
            if (str_nbytes) HM_CUDA(cudaMemcpyAsync(c->str_bytes.p, h.s_bytes + str_b0, (size_t)str_nbytes, cudaMemcpyHostToDevice, u->copy));
            HM_CUDA(cudaMemcpyAsync(c->str_off.p, h.s_offsets + s0, (size_t)(nnz + 1) * 8, cudaMemcpyHostToDevice, u->copy));
        }
    } else if (h.columns) {
        const int64_t hr = std::max(n, u->hint_rows);
        HM_TRY(c->sq_feat.reserve((size_t)std::max<int32_t>(h.n_q, 1) * 4));
        HM_TRY(c->sq_val.reserve((size_t)std::max<int64_t>(hr * h.n_q, 1) * 4));
        HM_TRY(c->sc_feat.reserve((size_t)std::max<int64_t>(hr * h.n_c, 1) * 4));
        if (h.n_q) {
            HM_CUDA(cudaMemcpyAsync(c->sq_feat.p, h.q_cols, (size_t)h.n_q * 4, cudaMemcpyHostToDevice, u->copy));
            HM_CUDA(cudaMemcpyAsync(c->sq_val.p, h.cq_value, (size_t)n * h.n_q * 4, cudaMemcpyHostToDevice, u->copy));
        }
        if (h.n_c) HM_CUDA(cudaMemcpyAsync(c->sc_feat.p, h.cc_feature, (size_t)n * h.n_c * 4, cudaMemcpyHostToDevice, u->copy));
    } else if (!h.split) {
        const int64_t base = h.rowptr[0];
        if (nnz) {
            HM_CUDA(cudaMemcpyAsync(c->feat.as<uint32_t>() + t_nnz, h.feature + base, (size_t)nnz * 4, cudaMemcpyHostToDevice, u->copy));
            if (h.value) HM_CUDA(cudaMemcpyAsync(c->val.as<float>() + t_nnz, h.value + base, (size_t)nnz * 4, cudaMemcpyHostToDevice, u->copy));
            else HM_TRY(launch_fill_u32(u->Lc, c->val.as<uint32_t>() + t_nnz, nnz, 0x3f800000u));   // index-only features: 1.0f
        }
    } else {
        const int64_t qb = h.q_rowptr[0], cb = h.c_rowptr[0];
        const int64_t qn = h.q_rowptr[n] - qb, cn = h.c_rowptr[n] - cb;
        if (qn < 0 || cn < 0) { abandon_chunk(u, c); HM_FAIL(HM_ERR_INVALID_ARGUMENT, "rowptr is not non-decreasing"); }
        const int64_t hr = std::max(n, u->hint_rows);
        HM_TRY(c->sq_rp.reserve((size_t)(hr + 1) * 8));
        HM_TRY(c->sc_rp.reserve((size_t)(hr + 1) * 8));
        HM_TRY(c->sq_feat.reserve((size_t)std::max<int64_t>(std::max(qn, u->hint_q), 1) * 4));
        HM_TRY(c->sq_val.reserve((size_t)std::max<int64_t>(std::max(qn, u->hint_q), 1) * 4));
        HM_TRY(c->sc_feat.reserve((size_t)std::max<int64_t>(std::max(cn, u->hint_c), 1) * 4));
        HM_CUDA(cudaMemcpyAsync(c->sq_rp.p, h.q_rowptr, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, u->copy));
        HM_CUDA(cudaMemcpyAsync(c->sc_rp.p, h.c_rowptr, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, u->copy));
        if (qn) {
            HM_CUDA(cudaMemcpyAsync(c->sq_feat.p, h.q_feature + qb, (size_t)qn * 4, cudaMemcpyHostToDevice, u->copy));
            HM_CUDA(cudaMemcpyAsync(c->sq_val.p, h.q_value + qb, (size_t)qn * 4, cudaMemcpyHostToDevice, u->copy));
        }
        if (cn) HM_CUDA(cudaMemcpyAsync(c->sc_feat.p, h.c_feature + cb, (size_t)cn * 4, cudaMemcpyHostToDevice, u->copy));
        // (the merge kernel is launched below, on the sort stream: a kernel on the copy stream would hold up
        // the next slice's copies)
    }
    HM_CUDA(cudaMemcpyAsync(c->label.as<float>() + t_rows, h.label, (size_t)n * 4, cudaMemcpyHostToDevice, u->copy));

    // rowptr of the new chunk, rebased on the host into pinned staging
    const int turn = u->stage_turn;
    u->stage_turn ^= 1;
    if (u->stage_ev[turn]) HM_CUDA(cudaEventSynchronize(u->stage_ev[turn]));
    if (rows + 1 > u->h_stage_cap[turn]) {
        if (u->h_stage[turn]) cudaFreeHost(u->h_stage[turn]);
        u->h_stage[turn] = nullptr; u->h_stage_cap[turn] = 0;
        HM_CUDA(cudaMallocHost(&u->h_stage[turn], (size_t)(rows + 1 + rows / 4) * sizeof(int64_t)));
        u->h_stage_cap[turn] = rows + 1 + rows / 4;
    }
    int64_t *st = u->h_stage[turn];
    for (int64_t i = 0; i < t_rows; i++) st[i] = u->tail_rp[i];
    const int64_t per_row = (int64_t)h.n_q + h.n_c;
    // element position of row i of [tail | slice] (column rows need no table for their own part)
    auto st_at = [&](int64_t i) -> int64_t { return (h.columns && i >= t_rows) ? t_nnz + (i - t_rows) * per_row : st[i]; };
    if (h.columns) {
        // (nothing to rebase: the row pointer of the new rows is implied and written on the device)
    } else if (!h.split) {
        const int64_t base = h.rowptr[0];
        for (int64_t i = 0; i <= n; i++) {
            if (i > 0 && h.rowptr[i] < h.rowptr[i - 1]) {
                abandon_chunk(u, c);
                HM_FAIL(HM_ERR_INVALID_ARGUMENT, "rowptr decreases at row %lld (rows pushed by earlier slices of this call were trained)", (long long)i);
            }
            st[t_rows + i] = t_nnz + (h.rowptr[i] - base);
        }
    } else {
        const int64_t qb = h.q_rowptr[0], cb = h.c_rowptr[0];
        for (int64_t i = 0; i <= n; i++) {
            if (i > 0 && (h.q_rowptr[i] < h.q_rowptr[i - 1] || h.c_rowptr[i] < h.c_rowptr[i - 1])) {
                abandon_chunk(u, c);
                HM_FAIL(HM_ERR_INVALID_ARGUMENT, "rowptr decreases at row %lld (rows pushed by earlier slices of this call were trained)", (long long)i);
            }
            st[t_rows + i] = t_nnz + (h.q_rowptr[i] - qb) + (h.c_rowptr[i] - cb);
        }
    }

    // carried rows: device-to-device, also on the copy stream.  Their source is input data that this very
    // stream copied in earlier (and the old chunk is only overwritten by a later copy on this stream), so
    // nothing here waits for the compute stream: the sort of this slice can run while the previous slice
    // is still training.
    if (t_rows) {
        const Chunk *o = u->tail;
        const int64_t o_e0 = o->mb_abs.empty() ? 0 : o->mb_abs.back();
        HM_CUDA(cudaMemcpyAsync(c->feat.p, o->ft() + o_e0, (size_t)t_nnz * 4, cudaMemcpyDeviceToDevice, u->copy));
        HM_CUDA(cudaMemcpyAsync(c->val.p, o->vl() + o_e0, (size_t)t_nnz * 4, cudaMemcpyDeviceToDevice, u->copy));
        HM_CUDA(cudaMemcpyAsync(c->label.p, o->lb() + u->tail_row0, (size_t)t_rows * 4, cudaMemcpyDeviceToDevice, u->copy));
    }
    if (u->tail) {
        // the old tail chunk is finished once the copies above have run
        Chunk *o = u->tail;
        u->tail = nullptr;
        HM_TRY(retire_chunk(u, o, u->cfg.iters > 1 && o->rows > 0));
        if (u->cfg.iters > 1 && o->rows == 0) { /* nothing trained from it: already pooled */ }
    }
    // the row pointer last: it is the only copy that needs the host loop above (split rows: the merge
    // kernel has written the new rows' part, only the carried rows' part is copied)
    const int64_t rp_copy = (h.split || h.columns) ? t_rows : rows + 1;
    if (rp_copy) HM_CUDA(cudaMemcpyAsync(c->rowptr.p, st, (size_t)rp_copy * 8, cudaMemcpyHostToDevice, u->copy));
    if (!u->stage_ev[turn]) HM_CUDA(cudaEventCreateWithFlags(&u->stage_ev[turn], cudaEventDisableTiming));
    HM_CUDA(cudaEventRecord(u->stage_ev[turn], u->copy));
    HM_CUDA(cudaStreamWaitEvent(u->compute, u->stage_ev[turn], 0));
    HM_CUDA(cudaStreamWaitEvent(u->sorter, u->stage_ev[turn], 0));     // the sort needs the rows, not the model
    if (h.split) {
        // the two groups -> the chunk's CSR, on the sort stream (its first consumer); training waits for it too
        const int64_t qb = h.q_rowptr[0], cb = h.c_rowptr[0];
        const int grid = (int)std::min<int64_t>((n + 7) / 8, (int64_t)num_sms() * 8);
        merge_split_kernel<<<grid, 256, 0, u->sorter>>>(n, c->sq_rp.as<int64_t>(), qb, c->sc_rp.as<int64_t>(), cb,
                                                        c->sq_feat.as<uint32_t>(), c->sq_val.as<float>(),
                                                        c->sc_feat.as<uint32_t>(), t_nnz, c->rowptr.as<int64_t>() + t_rows,
                                                        c->feat.as<uint32_t>(), c->val.as<float>());
        u->Ls.launches++;
        HM_CUDA(cudaGetLastError());
        if (!u->merged_ev) HM_CUDA(cudaEventCreateWithFlags(&u->merged_ev, cudaEventDisableTiming));
        HM_CUDA(cudaEventRecord(u->merged_ev, u->sorter));
        HM_CUDA(cudaStreamWaitEvent(u->compute, u->merged_ev, 0));
    }
    if (h.columns) {
        const int grid = (int)std::min<int64_t>((nnz + 1023) / 1024 + 1, (int64_t)num_sms() * 8);
        expand_columns_kernel<<<grid, 256, 0, u->sorter>>>(n, h.n_q, h.n_c, c->sq_feat.as<uint32_t>(), c->sq_val.as<float>(),
                                                          c->sc_feat.as<uint32_t>(), t_nnz, c->rowptr.as<int64_t>() + t_rows,
                                                          c->feat.as<uint32_t>(), c->val.as<float>());
        u->Ls.launches++;
        HM_CUDA(cudaGetLastError());
        if (!u->merged_ev) HM_CUDA(cudaEventCreateWithFlags(&u->merged_ev, cudaEventDisableTiming));
        HM_CUDA(cudaEventRecord(u->merged_ev, u->sorter));
        HM_CUDA(cudaStreamWaitEvent(u->compute, u->merged_ev, 0));
    }
    if (h.strings && nnz) {
        // feature strings -> the chunk's feature / value arrays, on the sort stream (their first consumer);
        // training waits for it too.  The kernel takes absolute offsets: the byte buffer is passed shifted.
        HM_TRY(launch_feature_hashing(u->Ls, c->str_bytes.as<uint8_t>() - str_b0, c->str_off.as<int64_t>(), nnz,
                                      u->cfg.hash_seed, (int32_t)u->cfg.dims, reinterpret_cast<int32_t *>(c->feat.as<uint32_t>() + t_nnz),
                                      c->val.as<float>() + t_nnz, (double)str_nbytes + 16.0 * (double)nnz));
        if (!u->merged_ev) HM_CUDA(cudaEventCreateWithFlags(&u->merged_ev, cudaEventDisableTiming));
        HM_CUDA(cudaEventRecord(u->merged_ev, u->sorter));
        HM_CUDA(cudaStreamWaitEvent(u->compute, u->merged_ev, 0));
    }
    if (h.columns) u->stats.h2d_bytes += n * (per_row * 4 + 4) + (int64_t)h.n_q * 4 + t_rows * 8;
    else if (h.strings) u->stats.h2d_bytes += (rows + 1) * 8 + str_nbytes + (nnz + 1) * 8 + n * 4;
    else if (!h.split) u->stats.h2d_bytes += (rows + 1) * 8 + nnz * (h.value ? 8 : 4) + n * 4;
    else u->stats.h2d_bytes += (n + 1) * 16 + (h.q_rowptr[n] - h.q_rowptr[0]) * 8 + (h.c_rowptr[n] - h.c_rowptr[0]) * 4 + n * 4 + t_rows * 8;

    // full mini-batches
    const int64_t n_mb = rows / B;
    const int64_t train = n_mb * B;
    c->rows = train;
    c->mb_abs.resize((size_t)n_mb + 1);
    for (int64_t i = 0; i <= n_mb; i++) c->mb_abs[(size_t)i] = st_at(i * B);
    HM_TRY(train_rows(u, *c, 0, train, c->mb_abs.data(), n_mb, true, u->cfg.iters > 1 ? kTrainKeep : kTrainPlain));

    // what is left waits in this chunk
    u->tail = c;
    u->tail_row0 = train;
    u->tail_rp.resize((size_t)(rows - train) + 1);
    for (int64_t i = 0; i <= rows - train; i++) u->tail_rp[(size_t)i] = st_at(train + i) - st_at(train);
    if (rows == train) u->tail_rp.clear();
    return HM_OK;
}

// The slices of one process() call (see hm_udtf_process for the schedule).
hm_status process_host_rows(hm_udtf *u, const HostRows &all)
{
    const int64_t n_rows = all.n;
    const int64_t B = u->cfg.mini_batch;
    const int64_t smallest = std::max<int64_t>(kMinSliceRows / B * B, 2 * B);
    // (split rows arrive as more, smaller arrays: twice the rows per slice keep the copies long -- for the
    // linear learners, where the bus is the bottleneck; train_fm trains about as fast as the bus delivers and
    // wants the shorter slices that keep copy and training in step: measured 131 M vs 141 M examples/s)
    const int64_t largest = std::max<int64_t>(((all.split || all.columns) && !u->fm ? 2 * kSliceRows : kSliceRows) / B * B, smallest);
    int64_t ramp = smallest;
    {
        const int64_t big = std::min(n_rows, largest);
        const double frac = (double)(big + B) / (double)n_rows * 1.05;
        u->hint_rows = big + B;
        u->hint_elems = (int64_t)((double)all.nnz() * frac) + 1024;
        u->hint_q = all.split ? (int64_t)((double)(all.q_rowptr[n_rows] - all.q_rowptr[0]) * frac) + 1024 : 0;
        u->hint_c = all.split ? (int64_t)((double)(all.c_rowptr[n_rows] - all.c_rowptr[0]) * frac) + 1024 : 0;
        u->hint_bytes = all.strings ? (int64_t)((double)(all.s_offsets[all.rowptr[n_rows]] - all.s_offsets[all.rowptr[0]]) * frac) + 4096 : 0;
    }
    hm_status st = HM_OK;
    for (int64_t r = 0; r < n_rows && st == HM_OK;) {
        int64_t slice = ramp;
        while (slice > smallest && n_rows - r <= 2 * slice) slice = std::max<int64_t>((slice / 2) / B * B, smallest);
        HostRows h = all;
        h.n = std::min(slice, n_rows - r);
        if (all.columns) { h.cq_value = all.cq_value + r * all.n_q; h.cc_feature = all.cc_feature + r * all.n_c; }
        else if (all.split) { h.q_rowptr = all.q_rowptr + r; h.c_rowptr = all.c_rowptr + r; }
        else h.rowptr = all.rowptr + r;
        h.label = all.label + r;
        st = process_slice(u, h);
        r += h.n;
        ramp = largest;
    }
    u->hint_rows = 0; u->hint_elems = 0; u->hint_q = 0; u->hint_c = 0; u->hint_bytes = 0;
    return st;
}

hm_status check_bad(hm_udtf *u)
{
    uint32_t bad = 0;
    HM_CUDA(cudaMemcpyAsync(u->h_small, u->d_bad, sizeof(uint32_t), cudaMemcpyDeviceToHost, u->compute));
    HM_CUDA(cudaStreamSynchronize(u->compute));
    std::memcpy(&bad, u->h_small, sizeof(bad));
    if (bad) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "a training row held a feature index that was negative or >= dims (%lld); "
                     "those elements took no part in the margins or the updates", (long long)u->cfg.dims);
    return HM_OK;
}

hm_status create_impl(const hm_config &cfg, int device, hm_udtf **out)
{
    if (!out) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "out must not be NULL");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        HM_FAIL(HM_ERR_CUDA, "no CUDA device: hivemall_b200 has no CPU fallback");
    }
    if (device < 0 || device >= ndev) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "device %d out of range [0, %d)", device, ndev);
    if (cfg.task == HM_TASK_FM && (cfg.factors < 1 || cfg.factors > 64))
        HM_FAIL(HM_ERR_UNSUPPORTED, "train_fm supports 1..64 factors in this build (got %d)", cfg.factors);
    if (cfg.dims < 1 || cfg.dims > ((int64_t)1 << 31)) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "dims out of range");
    if (cfg.mini_batch < 1) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "mini_batch must be >= 1");
    DeviceGuard g(device);
    hm_udtf *u = new (std::nothrow) hm_udtf();
    if (!u) HM_FAIL(HM_ERR_OOM, "out of host memory");
    u->cfg = cfg;
    u->device = device;
    std::memset(&u->stats, 0, sizeof(u->stats));
    if (const char *e = std::getenv("HM_HOT_FEATURES")) u->hot_enabled = std::atoi(e) != 0;
    hm_status s = [&]() -> hm_status {
        {
            // the training stream's short kernels go first whenever both streams have CTAs to place
            // (measured: +0.35 %; the sort above training: -1.7 %)
            int lo = 0, hi = 0;
            HM_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            HM_CUDA(cudaStreamCreateWithPriority(&u->compute, cudaStreamNonBlocking, hi));
            HM_CUDA(cudaStreamCreateWithFlags(&u->copy, cudaStreamNonBlocking));
            HM_CUDA(cudaStreamCreateWithPriority(&u->sorter, cudaStreamNonBlocking, lo));
        }
        u->L.stream = u->compute;
        u->Ls.stream = u->sorter;
        u->Lc.stream = u->copy;
        for (int i = 0; i < 2; i++) {
            HM_CUDA(cudaEventCreateWithFlags(&u->sorted_ev[i], cudaEventDisableTiming));
            HM_CUDA(cudaEventCreateWithFlags(&u->trained_ev[i], cudaEventDisableTiming));
        }
        HM_CUDA(cudaEventCreateWithFlags(&u->ready_ev, cudaEventDisableTiming));
        u->state_width = cfg.opt == HM_OPT_SGD ? 0 : ((cfg.opt == HM_OPT_ADAGRAD || cfg.opt == HM_OPT_COV) ? 1 : 2);
        if (cfg.task == HM_TASK_FM) {
            u->fm = true;
            u->k = cfg.factors;
            const bool ada = cfg.opt == HM_OPT_ADAGRAD;
            u->ww = ada ? 2 : 1;
            u->vw = ada ? 2 * u->k : u->k;
            const size_t tb = ((size_t)cfg.dims + 3) / 4 * 4;
            HM_CUDA(cudaMalloc(&u->w0rec, 2 * sizeof(float)));
            HM_CUDA(cudaMalloc(&u->wrec, (size_t)cfg.dims * u->ww * sizeof(float)));
            HM_CUDA(cudaMalloc(&u->vrec, (size_t)cfg.dims * u->vw * sizeof(float)));
            HM_CUDA(cudaMalloc(&u->touched, tb));
            HM_CUDA(cudaMemsetAsync(u->w0rec, 0, 2 * sizeof(float), u->compute));
            HM_CUDA(cudaMemsetAsync(u->wrec, 0, (size_t)cfg.dims * u->ww * sizeof(float), u->compute));
            HM_CUDA(cudaMemsetAsync(u->vrec, 0, (size_t)cfg.dims * u->vw * sizeof(float), u->compute));
            HM_CUDA(cudaMemsetAsync(u->touched, 0, tb, u->compute));
            HM_TRY(launch_fm_init_v(u->L, fm_args(u)));
        } else {
            // FTRL keeps (z, n) only: the weight is derived (see WeightView)
            u->rec_width = u->state_width == 0 ? 1 : ((u->state_width == 1 || cfg.opt == HM_OPT_FTRL) ? 2 : 4);
            HM_CUDA(cudaMalloc(&u->rec, (size_t)cfg.dims * u->rec_width * sizeof(float)));
            HM_CUDA(cudaMemsetAsync(u->rec, 0, (size_t)cfg.dims * u->rec_width * sizeof(float), u->compute));
            // the variance-keeping learners start from sigma = 1 for every feature (slot 1 of the (w, sigma) records)
            if (cfg.opt == HM_OPT_COV) HM_TRY(launch_fill_strided(u->L, u->rec + 1, cfg.dims, u->rec_width, 1.0f));
        }
        HM_CUDA(cudaMalloc(&u->d_bad, 2 * sizeof(uint32_t)));
        HM_CUDA(cudaMemsetAsync(u->d_bad, 0, 2 * sizeof(uint32_t), u->compute));
        HM_CUDA(cudaMalloc(&u->d_total, sizeof(int64_t)));
        HM_TRY(update_scratch_reserve(u->upd));
        HM_TRY(ensure_small(u, 1024));
        HM_CUDA(cudaStreamSynchronize(u->compute));
        return HM_OK;
    }();
    if (s != HM_OK) { hm_udtf_destroy(u); return s; }
    *out = u;
    return HM_OK;
}

}  // namespace

// =================================================================================
//                                   extern "C"
// =================================================================================

extern "C" {

const char *hm_last_error(void) { return hm::last_error_cstr(); }
int hm_abi_version(void) { return HM_ABI_VERSION; }

int hm_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

hm_status hm_host_alloc(void **out, int64_t bytes)
{
    if (!out || bytes < 0) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to hm_host_alloc");
    *out = nullptr;
    HM_CUDA(cudaMallocHost(out, (size_t)std::max<int64_t>(bytes, 1)));
    return HM_OK;
}

hm_status hm_host_free(void *p)
{
    if (p) HM_CUDA(cudaFreeHost(p));
    return HM_OK;
}

hm_status hm_parse_options(const char *function_name, const char *options, hm_config *out)
{
    return hm::parse_options(function_name, options, out);
}

// ---- hashing -----------------------------------------------------------------------

static hm_status hash_host(const uint8_t *bytes, const int64_t *offsets, int64_t n, uint32_t seed,
                           int32_t num_features, int32_t *out_index, float *out_value, int device, bool with_value)
{
    if (n < 0 || (n > 0 && (!bytes || !offsets || !out_index)) || (with_value && n > 0 && !out_value))
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL buffer passed to feature hashing");
    if (num_features < 1) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "num_features must be >= 1");
    if (n == 0) return HM_OK;
    int ndev = hm_device_count();
    if (ndev == 0) HM_FAIL(HM_ERR_CUDA, "no CUDA device: hivemall_b200 has no CPU fallback");
    if (device < 0 || device >= ndev) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "device %d out of range", device);
    DeviceGuard g(device);
    const int64_t nbytes = offsets[n];
    if (offsets[0] != 0 || nbytes < 0) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "offsets must start at 0 and be non-decreasing");
    uint8_t *d_bytes = nullptr; int64_t *d_off = nullptr; int32_t *d_idx = nullptr; float *d_val = nullptr;
    cudaStream_t st = nullptr;
    Launcher L;
    hm_status s = [&]() -> hm_status {
        HM_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        L.stream = st;
        HM_CUDA(cudaMalloc(&d_bytes, (size_t)nbytes + 32));
        HM_CUDA(cudaMalloc(&d_off, (size_t)(n + 1) * 8));
        HM_CUDA(cudaMalloc(&d_idx, (size_t)n * 4));
        if (with_value) HM_CUDA(cudaMalloc(&d_val, (size_t)n * 4));
        HM_CUDA(cudaMemcpyAsync(d_bytes, bytes, (size_t)nbytes, cudaMemcpyHostToDevice, st));
        HM_CUDA(cudaMemcpyAsync(d_off, offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st));
        if (with_value) HM_TRY(launch_feature_hashing(L, d_bytes, d_off, n, seed, num_features, d_idx, d_val));
        else HM_TRY(launch_mhash(L, d_bytes, d_off, n, seed, num_features, d_idx));
        HM_CUDA(cudaMemcpyAsync(out_index, d_idx, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        if (with_value) HM_CUDA(cudaMemcpyAsync(out_value, d_val, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        HM_CUDA(cudaStreamSynchronize(st));
        return HM_OK;
    }();
    cudaFree(d_bytes); cudaFree(d_off); cudaFree(d_idx); cudaFree(d_val);
    if (st) cudaStreamDestroy(st);
    return s;
}

hm_status hm_mhash(const uint8_t *bytes, const int64_t *offsets, int64_t n, uint32_t seed, int32_t num_features,
                   int32_t *out, int device)
{
    return hash_host(bytes, offsets, n, seed, num_features, out, nullptr, device, false);
}

hm_status hm_feature_hashing(const uint8_t *bytes, const int64_t *offsets, int64_t n, uint32_t seed,
                             int32_t num_features, int32_t *out_index, float *out_value, int device)
{
    return hash_host(bytes, offsets, n, seed, num_features, out_index, out_value, device, true);
}

hm_status hm_mhash_device(const uint8_t *d_bytes, const int64_t *d_offsets, int64_t n, uint32_t seed,
                          int32_t num_features, int32_t *d_out, void *stream)
{
    if (n < 0 || (n > 0 && (!d_bytes || !d_offsets || !d_out))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL buffer passed to hm_mhash_device");
    if (num_features < 1) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "num_features must be >= 1");
    Launcher L;
    L.stream = reinterpret_cast<cudaStream_t>(stream);
    return launch_mhash(L, d_bytes, d_offsets, n, seed, num_features, d_out);
}

// ---- lifecycle ---------------------------------------------------------------------

hm_status hm_udtf_create(const char *function_name, const char *options, int device, hm_udtf **out)
{
    hm_config cfg;
    HM_TRY(hm::parse_options(function_name, options, &cfg));
    return create_impl(cfg, device, out);
}

hm_status hm_udtf_create_from_config(const hm_config *cfg, int device, hm_udtf **out)
{
    if (!cfg) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "cfg must not be NULL");
    return create_impl(*cfg, device, out);
}

hm_status hm_udtf_process(hm_udtf *u, int64_t n_rows, const int64_t *rowptr, const int32_t *feature,
                          const float *value, const float *label)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (u->closed) HM_FAIL(HM_ERR_STATE, "process() after close()");
    if (u->device_rows_short) HM_FAIL(HM_ERR_STATE, "process() after a device call that ended on a short mini-batch");
    if (n_rows < 0 || (n_rows > 0 && (!rowptr || !label))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL buffer passed to process()");
    if (n_rows == 0) return HM_OK;
    if (rowptr[n_rows] > rowptr[0] && !feature) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL feature buffer");
    DeviceGuard g(u->device);
    // Slices of whole mini-batches, so that copies and training overlap (copy of slice k+1 beside the training
    // of slice k, two chunk buffers).  A call takes
    //   (copy of the FIRST slice, during which the device idles) + sum over slices of max(copy, training)
    //   + (training of the LAST slice, which no copy overlaps):
    // one small slice first (short fill), then slices of kSliceRows (copies of >= 160 MB reach the full
    // 55 GB/s; equal sizes keep the two stages in step when training is as slow as the bus, as for train_fm),
    // halving towards the end of the call (short drain: matters when the bus is the bottleneck).
    HostRows all;
    all.n = n_rows; all.rowptr = rowptr; all.feature = feature; all.value = value; all.label = label;
    return process_host_rows(u, all);
}

hm_status hm_udtf_process_split(hm_udtf *u, int64_t n_rows, const int64_t *q_rowptr, const int32_t *q_feature,
                                const float *q_value, const int64_t *c_rowptr, const int32_t *c_feature,
                                const float *label)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (u->closed) HM_FAIL(HM_ERR_STATE, "process() after close()");
    if (u->device_rows_short) HM_FAIL(HM_ERR_STATE, "process() after a device call that ended on a short mini-batch");
    if (n_rows < 0 || (n_rows > 0 && (!q_rowptr || !c_rowptr || !label)))
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL buffer passed to process_split()");
    if (n_rows == 0) return HM_OK;
    if (q_rowptr[n_rows] > q_rowptr[0] && (!q_feature || !q_value)) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL quantitative feature/value buffer");
    if (c_rowptr[n_rows] > c_rowptr[0] && !c_feature) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL index-only feature buffer");
    DeviceGuard g(u->device);
    HostRows all;
    all.n = n_rows; all.split = true;
    all.q_rowptr = q_rowptr; all.q_feature = q_feature; all.q_value = q_value;
    all.c_rowptr = c_rowptr; all.c_feature = c_feature; all.label = label;
    return process_host_rows(u, all);
}

hm_status hm_udtf_process_columns(hm_udtf *u, int64_t n_rows, int32_t n_q, const int32_t *q_feature, const float *q_value,
                                  int32_t n_c, const int32_t *c_feature, const float *label)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (u->closed) HM_FAIL(HM_ERR_STATE, "process() after close()");
    if (u->device_rows_short) HM_FAIL(HM_ERR_STATE, "process() after a device call that ended on a short mini-batch");
    if (n_rows < 0 || n_q < 0 || n_c < 0 || (n_rows > 0 && !label)) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to process_columns()");
    if (n_rows > 0 && ((n_q > 0 && (!q_feature || !q_value)) || (n_c > 0 && !c_feature)))
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL buffer passed to process_columns()");
    if (n_rows == 0) return HM_OK;
    DeviceGuard g(u->device);
    HostRows all;
    all.n = n_rows; all.columns = true; all.n_q = n_q; all.n_c = n_c;
    all.q_cols = q_feature; all.cq_value = q_value; all.cc_feature = c_feature; all.label = label;
    return process_host_rows(u, all);
}

hm_status hm_udtf_process_strings(hm_udtf *u, int64_t n_rows, const int64_t *rowptr, const uint8_t *bytes,
                                  const int64_t *offsets, const float *label)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (u->closed) HM_FAIL(HM_ERR_STATE, "process() after close()");
    if (u->device_rows_short) HM_FAIL(HM_ERR_STATE, "process() after a device call that ended on a short mini-batch");
    if (n_rows < 0 || (n_rows > 0 && (!rowptr || !label || !offsets))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL buffer passed to process_strings()");
    if (n_rows == 0) return HM_OK;
    const int64_t n = rowptr[n_rows] - rowptr[0];
    if (n < 0) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "rowptr is not non-decreasing");
    if (n > 0 && !bytes) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL bytes buffer");
    if (u->cfg.dims > 0x7fffffffLL) HM_FAIL(HM_ERR_UNSUPPORTED, "feature hashing needs dims < 2^31");
    DeviceGuard g(u->device);
    // The same slices as hm_udtf_process: the strings' bytes and offsets cross the bus on the copy stream and
    // are hashed on the device straight into the chunk's CSR (no index ever visits the host).
    HostRows all;
    all.n = n_rows; all.rowptr = rowptr; all.strings = true; all.s_bytes = bytes; all.s_offsets = offsets; all.label = label;
    return process_host_rows(u, all);
}

hm_status hm_udtf_process_device(hm_udtf *u, int64_t n_rows, int64_t nnz, const int64_t *d_rowptr,
                                 const int32_t *d_feature, const float *d_value, const float *d_label)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (u->closed) HM_FAIL(HM_ERR_STATE, "process() after close()");
    if (u->tail_rows() > 0) HM_FAIL(HM_ERR_STATE, "process_device() while %lld host rows wait for a full mini-batch", (long long)u->tail_rows());
    if (u->device_rows_short) HM_FAIL(HM_ERR_STATE, "process_device() after a call that ended on a short mini-batch");
    if (n_rows < 0 || nnz < 0 || (n_rows > 0 && (!d_rowptr || !d_label)) || (nnz > 0 && (!d_feature || !d_value)))
        HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to process_device()");
    if (n_rows == 0) return HM_OK;
    DeviceGuard g(u->device);
    const int64_t B = u->cfg.mini_batch;
    const int64_t n_mb = (n_rows + B - 1) / B;
    Chunk *c = new (std::nothrow) Chunk();
    if (!c) HM_FAIL(HM_ERR_OOM, "out of host memory");
    c->view = true;
    c->rows = n_rows;
    c->v_rowptr = d_rowptr; c->v_feat = reinterpret_cast<const uint32_t *>(d_feature);
    c->v_val = d_value; c->v_label = d_label;
    // rowptr at the mini-batch boundaries, read back on the copy stream (the rows are resident already)
    hm_status s = [&]() -> hm_status {
        HM_TRY(u->mb_gather.reserve((size_t)(n_mb + 1) * sizeof(int64_t)));
        HM_TRY(ensure_small(u, n_mb + 1));
        const int grid = (int)std::min<int64_t>((n_mb + 256) / 256, 1024);
        gather_rowptr_kernel<<<grid, 256, 0, u->copy>>>(d_rowptr, n_rows, (int32_t)B, n_mb + 1, u->mb_gather.as<int64_t>());
        u->L.launches++;
        HM_CUDA(cudaGetLastError());
        HM_CUDA(cudaMemcpyAsync(u->h_small, u->mb_gather.p, (size_t)(n_mb + 1) * 8, cudaMemcpyDeviceToHost, u->copy));
        HM_CUDA(cudaStreamSynchronize(u->copy));
        u->stats.d2h_bytes += (n_mb + 1) * 8;
        c->mb_abs.assign(u->h_small, u->h_small + n_mb + 1);
        if (c->mb_abs.back() - c->mb_abs.front() != nnz)
            HM_FAIL(HM_ERR_INVALID_ARGUMENT, "nnz (%lld) does not match d_rowptr (%lld)", (long long)nnz,
                    (long long)(c->mb_abs.back() - c->mb_abs.front()));
        return train_rows(u, *c, 0, n_rows, c->mb_abs.data(), n_mb, true, u->cfg.iters > 1 ? kTrainKeep : kTrainPlain);
    }();
    if (s != HM_OK) { delete c; return s; }
    if (n_rows % B) u->device_rows_short = true;
    return retire_chunk(u, c, u->cfg.iters > 1);
}

hm_status hm_udtf_close(hm_udtf *u, int64_t *n_model_rows)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    DeviceGuard g(u->device);
    if (!u->closed) {
        // last short mini-batch
        if (u->tail) {
            Chunk *c = u->tail;
            const int64_t t_rows = u->tail_rows();
            if (t_rows > 0) {
                const int64_t e0 = c->mb_abs.empty() ? 0 : c->mb_abs.back();
                int64_t mb[2] = {e0, e0 + u->tail_rp.back()};
                HM_TRY(train_rows(u, *c, u->tail_row0, t_rows, mb, 1, false, u->cfg.iters > 1 ? kTrainKeep : kTrainPlain));
                if (c->mb_abs.empty()) c->mb_abs.push_back(e0);
                c->mb_abs.push_back(mb[1]);
                c->rows += t_rows;
            }
            u->tail = nullptr;
            u->tail_rp.clear();
            HM_TRY(retire_chunk(u, c, u->cfg.iters > 1 && c->rows > 0));
        }
        // remaining passes over the retained rows, same order, same mini-batch boundaries
        for (int it = 1; it < u->cfg.iters; it++)
            for (Chunk *c : u->retained) {
                const int64_t B = u->cfg.mini_batch;
                const int64_t n_mb = (c->rows + B - 1) / B;
                HM_TRY(train_rows(u, *c, 0, c->rows, c->mb_abs.data(), n_mb, true, kTrainReplay));
            }
        u->closed = true;
    }
    HM_TRY(check_bad(u));
    HM_TRY(launch_export_count(u->L, u->exp, weight_view(u), u->touched, u->cfg.dims, u->d_total));
    HM_CUDA(cudaMemcpyAsync(u->h_small, u->d_total, sizeof(int64_t), cudaMemcpyDeviceToHost, u->compute));
    HM_CUDA(cudaStreamSynchronize(u->compute));
    u->stats.d2h_bytes += 8;
    u->model_rows = u->h_small[0] + (u->fm ? 1 : 0);   // train_fm: plus the w0 row
    u->L.drain();
    u->Ls.drain();
    if (n_model_rows) *n_model_rows = u->model_rows;
    return HM_OK;
}

hm_status hm_udtf_forward(hm_udtf *u, int32_t *feature, float *weight, float *vif, int64_t cap, int64_t *n_written)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (!u->closed) HM_FAIL(HM_ERR_STATE, "forward() before close()");
    if (cap < 0 || (cap > 0 && (!feature || !weight))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to forward()");
    if (u->fm && cap > 0 && !vif) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "train_fm forward() needs a vif buffer");
    DeviceGuard g(u->device);
    // recount: the replica may have been averaged since close()
    HM_TRY(launch_export_count(u->L, u->exp, weight_view(u), u->touched, u->cfg.dims, u->d_total));
    HM_CUDA(cudaMemcpyAsync(u->h_small, u->d_total, sizeof(int64_t), cudaMemcpyDeviceToHost, u->compute));
    HM_CUDA(cudaStreamSynchronize(u->compute));
    const int64_t n = u->h_small[0];
    const int64_t head = u->fm ? 1 : 0;     // train_fm: row 0 is (feature -1, w0, zeros)
    u->model_rows = n + head;
    if (n + head > cap) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "model table has %lld rows but capacity is %lld", (long long)(n + head), (long long)cap);
    if (u->fm) {
        feature[0] = -1;
        HM_CUDA(cudaMemcpyAsync(weight, u->w0rec, 4, cudaMemcpyDeviceToHost, u->compute));
        std::memset(vif, 0, (size_t)u->k * 4);
    }
    if (n > 0) {
        HM_TRY(u->exp_feat.reserve((size_t)n * 4));
        HM_TRY(u->exp_weight.reserve((size_t)n * 4));
        HM_TRY(launch_export_write(u->L, u->exp, weight_view(u), u->touched, u->cfg.dims,
                                   u->exp_feat.as<int32_t>(), u->exp_weight.as<float>()));
        HM_CUDA(cudaMemcpyAsync(feature + head, u->exp_feat.p, (size_t)n * 4, cudaMemcpyDeviceToHost, u->compute));
        HM_CUDA(cudaMemcpyAsync(weight + head, u->exp_weight.p, (size_t)n * 4, cudaMemcpyDeviceToHost, u->compute));
        u->stats.d2h_bytes += n * 8;
        if (u->fm) {
            HM_TRY(u->exp_vif.reserve((size_t)n * u->k * 4));
            HM_TRY(launch_fm_gather_vif(u->L, fm_args(u), u->exp_feat.as<int32_t>(), n, u->exp_vif.as<float>()));
            HM_CUDA(cudaMemcpyAsync(vif + head * u->k, u->exp_vif.p, (size_t)n * u->k * 4, cudaMemcpyDeviceToHost, u->compute));
            u->stats.d2h_bytes += n * u->k * 4;
        }
    }
    HM_CUDA(cudaStreamSynchronize(u->compute));
    if (n_written) *n_written = n + head;
    return HM_OK;
}

hm_status hm_udtf_destroy(hm_udtf *u)
{
    if (!u) return HM_OK;
    DeviceGuard g(u->device);
    if (u->compute) cudaStreamSynchronize(u->compute);
    if (u->copy) cudaStreamSynchronize(u->copy);
    if (u->sorter) cudaStreamSynchronize(u->sorter);
    u->L.release();
    u->Ls.release();
    u->Lc.release();
    for (Chunk *c : u->retained) { c->release(); delete c; }
    for (Chunk *c : u->pool) { c->release(); delete c; }
    if (u->tail) { u->tail->release(); delete u->tail; }
    sort_scratch_free(u->sort[0]);
    sort_scratch_free(u->sort[1]);
    for (int i = 0; i < 2; i++) {
        if (u->sorted_ev[i]) cudaEventDestroy(u->sorted_ev[i]);
        if (u->trained_ev[i]) cudaEventDestroy(u->trained_ev[i]);
    }
    if (u->ready_ev) cudaEventDestroy(u->ready_ev);
    if (u->merged_ev) cudaEventDestroy(u->merged_ev);
    update_scratch_free(u->upd);
    cudaFree(u->exp.counts);
    for (DevBuf *b : {&u->sqnorm, &u->dloss2, &u->margin, &u->dloss, &u->mb_gather, &u->exp_feat, &u->exp_weight, &u->tmp_a,
                      &u->tmp_b, &u->tmp_c, &u->tmp_d})
        b->release();
    cudaFree(u->rec); cudaFree(u->d_bad); cudaFree(u->d_total);
    cudaFree(u->w0rec); cudaFree(u->wrec); cudaFree(u->vrec); cudaFree(u->touched);
    fm_scratch_free(u->fms);
    u->stab.release(); u->exp_vif.release();
    for (int i = 0; i < 2; i++) {
        if (u->h_stage[i]) cudaFreeHost(u->h_stage[i]);
        if (u->stage_ev[i]) cudaEventDestroy(u->stage_ev[i]);
    }
    if (u->h_small) cudaFreeHost(u->h_small);
    if (u->compute) cudaStreamDestroy(u->compute);
    if (u->copy) cudaStreamDestroy(u->copy);
    if (u->sorter) cudaStreamDestroy(u->sorter);
    cudaGetLastError();
    delete u;
    return HM_OK;
}

// ---- what hm_table.cpp needs to know about a learner ---------------------------------------
}  // extern "C"

namespace hm {

hm_status learner_info(const hm_udtf *u, int *is_fm, int *factors, int64_t *dims, int *opt)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    *is_fm = u->fm ? 1 : 0; *factors = u->k; *dims = u->cfg.dims; *opt = u->cfg.opt;
    return HM_OK;
}

__global__ void set_touched_kernel(uint8_t *__restrict__ touched, const int32_t *__restrict__ feature, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        touched[feature[i]] = 1;
}

// train_fm emits the features seen in training: features restored from a model table count as seen.
hm_status learner_set_touched(hm_udtf *u, const int32_t *feature, int64_t n)
{
    if (!u || !u->fm || n <= 0) return HM_OK;
    DeviceGuard g(u->device);
    HM_TRY(u->tmp_b.reserve((size_t)n * 4));
    HM_CUDA(cudaMemcpyAsync(u->tmp_b.p, feature, (size_t)n * 4, cudaMemcpyHostToDevice, u->compute));
    set_touched_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, 1024), 256, 0, u->compute>>>(u->touched, u->tmp_b.as<int32_t>(), n);
    u->L.launches++;
    HM_CUDA(cudaGetLastError());
    HM_CUDA(cudaStreamSynchronize(u->compute));
    return HM_OK;
}

}  // namespace hm

extern "C" {

hm_status hm_udtf_config(const hm_udtf *u, hm_config *out)
{
    if (!u || !out) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    *out = u->cfg;
    return HM_OK;
}

// ---- replica access ----------------------------------------------------------------

int hm_udtf_num_arrays(const hm_udtf *u) { return u ? (u->fm ? 4 : 1) : 0; }

hm_status hm_udtf_array(hm_udtf *u, int which, const char **name, void **d_ptr, int64_t *n_floats, int32_t *width)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    const char *nm = nullptr; void *p = nullptr; int64_t n = 0; int32_t w = 0;
    if (!u->fm && which == 0) { nm = "replica"; p = u->rec; n = u->cfg.dims * u->rec_width; w = u->rec_width; }
    else if (u->fm && which == 0) { nm = "w0"; p = u->w0rec; n = 2; w = 2; }
    else if (u->fm && which == 1) { nm = "w"; p = u->wrec; n = u->cfg.dims * u->ww; w = u->ww; }
    else if (u->fm && which == 2) { nm = "V"; p = u->vrec; n = u->cfg.dims * u->vw; w = u->vw; }
    else if (u->fm && which == 3) { nm = "touched"; p = u->touched; n = (u->cfg.dims + 3) / 4; w = 0; }
    else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "array index %d out of range", which);
    if (name) *name = nm;
    if (d_ptr) *d_ptr = p;
    if (n_floats) *n_floats = n;
    if (width) *width = w;
    return HM_OK;
}

// Logical arrays inside the interleaved replica.  Linear: 0 = w (dims), 1 = state (dims x slots).
// train_fm: 0 = w (dims), 1 = V (dims x factors), 2 = w0 (1), 3 = AdaGrad accumulators of w,
// 4 = of V, 5 = of w0.
static hm_status array_copy(hm_udtf *u, int which, float *host, const float *host_in, int64_t n)
{
    float *base = nullptr; int64_t rows = u->cfg.dims; int stride = 1, off = 0, width = 1;
    const bool ada = u->cfg.opt == HM_OPT_ADAGRAD;
    const bool ftrl = !u->fm && u->cfg.opt == HM_OPT_FTRL;
    if (ftrl && which == 0) {
        // the weight is not stored: evaluate it (reads only; a warm start writes (z, n) instead)
        if (!host) HM_FAIL(HM_ERR_UNSUPPORTED, "FTRL keeps (z, n) only: warm-start array 1 (state), the weights follow from it");
        if (n != u->cfg.dims) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "array 0 holds %lld floats, caller passed %lld", (long long)u->cfg.dims, (long long)n);
        DeviceGuard g(u->device);
        HM_CUDA(cudaStreamSynchronize(u->copy));
        HM_TRY(u->tmp_a.reserve((size_t)n * 4));
        HM_TRY(launch_dense_weights(u->L, weight_view(u), u->cfg.dims, u->tmp_a.as<float>()));
        HM_CUDA(cudaMemcpyAsync(host, u->tmp_a.p, (size_t)n * 4, cudaMemcpyDeviceToHost, u->compute));
        HM_CUDA(cudaStreamSynchronize(u->compute));
        u->stats.d2h_bytes += n * 4;
        return HM_OK;
    }
    if (!u->fm && which == 0) { base = u->rec; stride = u->rec_width; }
    else if (ftrl && which == 1) { base = u->rec; stride = 2; off = 0; width = 2; }
    else if (!u->fm && which == 1 && u->state_width > 0) { base = u->rec; stride = u->rec_width; off = 1; width = u->state_width; }
    else if (u->fm && which == 0) { base = u->wrec; stride = u->ww; }
    else if (u->fm && which == 1) { base = u->vrec; stride = u->vw; width = u->k; }
    else if (u->fm && which == 2) { base = u->w0rec; rows = 1; stride = 2; }
    else if (u->fm && ada && which == 3) { base = u->wrec; stride = u->ww; off = 1; }
    else if (u->fm && ada && which == 4) { base = u->vrec; stride = u->vw; off = u->k; width = u->k; }
    else if (u->fm && ada && which == 5) { base = u->w0rec; rows = 1; stride = 2; off = 1; }
    else HM_FAIL(HM_ERR_INVALID_ARGUMENT, "array index %d out of range", which);
    const int64_t len = rows * width;
    if (n != len) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "array %d holds %lld floats, caller passed %lld", which, (long long)len, (long long)n);
    DeviceGuard g(u->device);
    HM_CUDA(cudaStreamSynchronize(u->copy));
    HM_TRY(u->tmp_a.reserve((size_t)n * 4));
    if (host) {
        HM_TRY(launch_strided_copy(u->L, u->tmp_a.as<float>(), base, rows, stride, off, width, true));
        HM_CUDA(cudaMemcpyAsync(host, u->tmp_a.p, (size_t)n * 4, cudaMemcpyDeviceToHost, u->compute));
        u->stats.d2h_bytes += n * 4;
    } else {
        HM_CUDA(cudaMemcpyAsync(u->tmp_a.p, host_in, (size_t)n * 4, cudaMemcpyHostToDevice, u->compute));
        HM_TRY(launch_strided_copy(u->L, u->tmp_a.as<float>(), base, rows, stride, off, width, false));
        u->stats.h2d_bytes += n * 4;
    }
    HM_CUDA(cudaStreamSynchronize(u->compute));
    return HM_OK;
}

hm_status hm_udtf_read_array(hm_udtf *u, int which, float *host, int64_t n)
{
    if (!u || !host) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    return array_copy(u, which, host, nullptr, n);
}

hm_status hm_udtf_write_array(hm_udtf *u, int which, const float *host, int64_t n)
{
    if (!u || !host) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    return array_copy(u, which, nullptr, host, n);
}

hm_status hm_udtf_get_step(const hm_udtf *u, int64_t *t)
{
    if (!u || !t) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    *t = u->t;
    return HM_OK;
}

hm_status hm_udtf_set_step(hm_udtf *u, int64_t t)
{
    if (!u || t < 0) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad argument");
    u->t = t;
    return HM_OK;
}

hm_status hm_udtf_sync(hm_udtf *u)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    DeviceGuard g(u->device);
    HM_CUDA(cudaStreamSynchronize(u->copy));
    HM_CUDA(cudaStreamSynchronize(u->sorter));
    HM_CUDA(cudaStreamSynchronize(u->compute));
    return HM_OK;
}

void *hm_udtf_stream(hm_udtf *u) { return u ? (void *)u->compute : nullptr; }

hm_status hm_udtf_scale_replica(hm_udtf *u, float scale)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    DeviceGuard g(u->device);
    if (u->fm) {
        HM_TRY(launch_scale(u->L, u->w0rec, 2, scale));
        HM_TRY(launch_scale(u->L, u->wrec, u->cfg.dims * u->ww, scale));
        HM_TRY(launch_scale(u->L, u->vrec, u->cfg.dims * u->vw, scale));
    } else {
        HM_TRY(launch_scale(u->L, u->rec, u->cfg.dims * u->rec_width, scale));
    }
    return HM_OK;
}

// ---- prediction --------------------------------------------------------------------

hm_status hm_udtf_predict_device(hm_udtf *u, int64_t n_rows, const int64_t *d_rowptr, const int32_t *d_feature,
                                 const float *d_value, float *d_out)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (n_rows < 0 || (n_rows > 0 && (!d_rowptr || !d_out))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to predict");
    DeviceGuard g(u->device);
    if (u->fm) {
        if (n_rows > 0x7fffffff) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "predict: too many rows in one call");
        return launch_fm_forward(u->L, fm_args(u), d_rowptr, 0, (int32_t)n_rows, 0, reinterpret_cast<const uint32_t *>(d_feature),
                                 d_value, nullptr, d_out, nullptr, nullptr, u->d_bad + 1, 0, false);
    }
    return launch_predict_linear(u->L, weight_view(u), u->cfg.dims, n_rows, d_rowptr, reinterpret_cast<const uint32_t *>(d_feature),
                                 d_value, d_out, u->d_bad + 1);
}

hm_status hm_udtf_predict(hm_udtf *u, int64_t n_rows, const int64_t *rowptr, const int32_t *feature,
                          const float *value, float *out)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    if (n_rows < 0 || (n_rows > 0 && (!rowptr || !out))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to predict");
    if (n_rows == 0) return HM_OK;
    DeviceGuard g(u->device);
    const int64_t base = rowptr[0], nnz = rowptr[n_rows] - base;
    if (nnz < 0 || (nnz > 0 && (!feature || !value))) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad CSR passed to predict");
    HM_TRY(u->tmp_a.reserve((size_t)(n_rows + 1) * 8));
    HM_TRY(u->tmp_b.reserve((size_t)std::max<int64_t>(nnz, 1) * 4));
    HM_TRY(u->tmp_c.reserve((size_t)std::max<int64_t>(nnz, 1) * 4));
    HM_TRY(u->tmp_d.reserve((size_t)n_rows * 4));
    std::vector<int64_t> rp((size_t)n_rows + 1);
    for (int64_t i = 0; i <= n_rows; i++) rp[(size_t)i] = rowptr[i] - base;
    HM_CUDA(cudaMemcpyAsync(u->tmp_a.p, rp.data(), (size_t)(n_rows + 1) * 8, cudaMemcpyHostToDevice, u->compute));
    if (nnz) {
        HM_CUDA(cudaMemcpyAsync(u->tmp_b.p, feature + base, (size_t)nnz * 4, cudaMemcpyHostToDevice, u->compute));
        HM_CUDA(cudaMemcpyAsync(u->tmp_c.p, value + base, (size_t)nnz * 4, cudaMemcpyHostToDevice, u->compute));
    }
    if (u->fm) {
        if (n_rows > 0x7fffffff) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "predict: too many rows in one call");
        HM_TRY(launch_fm_forward(u->L, fm_args(u), u->tmp_a.as<int64_t>(), 0, (int32_t)n_rows, 0, u->tmp_b.as<uint32_t>(),
                                 u->tmp_c.as<float>(), nullptr, u->tmp_d.as<float>(), nullptr, nullptr, u->d_bad + 1, nnz, false));
    } else {
        HM_TRY(launch_predict_linear(u->L, weight_view(u), u->cfg.dims, n_rows, u->tmp_a.as<int64_t>(), u->tmp_b.as<uint32_t>(),
                                     u->tmp_c.as<float>(), u->tmp_d.as<float>(), u->d_bad + 1));
    }
    HM_CUDA(cudaMemcpyAsync(out, u->tmp_d.p, (size_t)n_rows * 4, cudaMemcpyDeviceToHost, u->compute));
    // the scoring flag is this call's own: read it, clear it, report it here (never at close())
    HM_CUDA(cudaMemcpyAsync(u->h_small, u->d_bad + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, u->compute));
    HM_CUDA(cudaMemsetAsync(u->d_bad + 1, 0, sizeof(uint32_t), u->compute));
    HM_CUDA(cudaStreamSynchronize(u->compute));   // also keeps `rp` alive long enough
    u->stats.h2d_bytes += (n_rows + 1) * 8 + nnz * 8;
    u->stats.d2h_bytes += n_rows * 4;
    uint32_t bad = 0;
    std::memcpy(&bad, u->h_small, sizeof(bad));
    if (bad) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "a scored row held a feature index that was negative or >= dims (%lld); "
                     "those elements contributed 0 to the margins written to `out`", (long long)u->cfg.dims);
    return HM_OK;
}

hm_status hm_udtf_last_margins(hm_udtf *u, float *host, int64_t cap, int64_t *n)
{
    if (!u || !n || cap < 0 || (cap > 0 && !host)) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to last_margins");
    DeviceGuard g(u->device);
    const int64_t m = std::min(cap, u->last_margin_n);
    if (m > 0) {
        HM_CUDA(cudaMemcpyAsync(host, u->margin.as<float>() + u->last_margin_off, (size_t)m * 4, cudaMemcpyDeviceToHost,
                                u->compute));
        u->stats.d2h_bytes += m * 4;
    }
    HM_CUDA(cudaStreamSynchronize(u->compute));
    *n = m;
    return HM_OK;
}

hm_status hm_udtf_hot_features(hm_udtf *u, int32_t *feature, int32_t cap, int32_t *n)
{
    if (!u || !n || cap < 0 || (cap > 0 && !feature)) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad arguments to hot_features");
    *n = 0;
    if (u->sort_turn == 0) return HM_OK;
    DeviceGuard g(u->device);
    const SortScratch &ss = u->sort[(u->sort_turn - 1) & 1];
    if (!ss.hot) return HM_OK;
    HM_CUDA(cudaStreamSynchronize(u->sorter));
    HotInfo hi;
    HM_CUDA(cudaMemcpy(&hi, ss.hot, sizeof(hi), cudaMemcpyDeviceToHost));
    const int32_t m = (int32_t)std::min<uint32_t>(hi.count, (uint32_t)kMaxHot);
    for (int32_t i = 0; i < m && i < cap; i++) feature[i] = (int32_t)hi.key[i];
    *n = std::min(m, cap);
    return HM_OK;
}

// ---- measurement -------------------------------------------------------------------

hm_status hm_udtf_stats(const hm_udtf *u, hm_stats *out)
{
    if (!u || !out) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL argument");
    *out = u->stats;
    out->kernel_launches = u->L.launches + u->Ls.launches + u->Lc.launches;
    return HM_OK;
}

hm_status hm_udtf_set_profiling(hm_udtf *u, int on)
{
    if (!u) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "NULL learner");
    DeviceGuard g(u->device);
    if (!on) { u->L.drain(); u->Ls.drain(); }
    u->L.profiling = on != 0;
    u->Ls.profiling = on != 0;
    return HM_OK;
}

int hm_udtf_num_kernels(const hm_udtf *) { return K_NUM_KERNELS; }

hm_status hm_udtf_kernel_time(hm_udtf *u, int i, const char **name, int64_t *launches, double *total_ms,
                              double *algorithmic_bytes)
{
    if (!u || i < 0 || i >= K_NUM_KERNELS) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "bad argument");
    DeviceGuard g(u->device);
    u->L.drain();
    u->Ls.drain();
    if (name) *name = kKernelNames[i];
    if (launches) *launches = u->L.count[i] + u->Ls.count[i];
    if (total_ms) *total_ms = u->L.ms[i] + u->Ls.ms[i];
    if (algorithmic_bytes) *algorithmic_bytes = u->L.bytes[i] + u->Ls.bytes[i];
    return HM_OK;
}

}  // extern "C"
