// hm_sort.cu -- stable segmented LSD radix sort: the "sort-by-feature" half of the
// deterministic gradient reduction that BASELINE.json's north_star asks for ("the loss
// gradient fused with a deterministic sort-by-feature segmented reduction (no float
// atomics)").  No reference source exists to cite (README.md:20-22).
//
// One launch set sorts up to kMaxSegs mini-batches ("segments") of an ingest block at
// once.  Segment s is the element range [seg_off[s], seg_off[s+1]) of the block's CSR;
// inside each segment elements are ordered by feature id, ties by original position
// (stable), so the later per-feature sum always adds rows in row order.
//
// Payload of every element is (row inside its mini-batch, value): with it the update kernel
// streams the sorted mini-batch and only looks up dloss[row] in a table of mini_batch floats,
// instead of gathering 4 bytes per element at random (profiles/r01a_summary.md).
//
// Per 8-bit digit pass:  hist (reads keys, 4 B/elem)  ->  two tiny scans  ->
// scatter (reads key+payload, writes key+payload, 24 B/elem).  The first pass reads the
// CSR feature/value arrays directly and looks the row ids up in the row pointer (20 B/elem),
// so no copy of the input is made.  Integer counting only; atomics are integer (order-independent).
//
// Two kinds of element never enter the sort: those of the block's hot features (HotInfo: their
// gradient sums are taken by the forward kernel, see hm_linear.cu) and those whose feature index
// is out of range (they are flagged and ignored everywhere).  Both are dropped by key in the
// first pass and its histogram, so the sorted arrays are shorter than the block's CSR:
// in_off / in_tile_off describe the CSR, seg_off / seg_tile_off the sorted arrays, and the latter
// are written on the device from the first histogram (scan_rows_kernel).
#include "hm_internal.h"
#include <algorithm>
#include <atomic>

namespace hm {

int num_sms()
{
    static std::atomic<int> cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 148; }
    if (dev >= 0 && dev < 64) { const int c = cache[dev].load(std::memory_order_relaxed); if (c > 0) return c; }
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) { cudaGetLastError(); n = 148; }
    if (dev >= 0 && dev < 64) cache[dev].store(n, std::memory_order_relaxed);
    return n;
}

const char *const kKernelNames[K_NUM_KERNELS] = {
    "hash", "seg_prep", "sort_hist", "sort_scan", "sort_scatter", "linear_forward",
    "linear_update", "fm_forward", "fm_update", "predict", "export", "misc", "hot_select"};

namespace {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = kSortTile / kSortThreads;  // 16
constexpr int kMaxRadixBits = 9;          // digits are 8 bits wide, or 9 where that saves a pass (17-18, 25-27 key bits)
constexpr int kMaxRadix = 1 << kMaxRadixBits;
// the scatter kernel covers the same 4096-element tile with twice the threads: more warps per SM
constexpr int kScatThreads = 512;
constexpr int kScatWarps = kScatThreads / 32;
constexpr int kScatItems = kSortTile / kScatThreads;  // 8
constexpr int kScatCtasPerSm = 2;

// Offsets of the block's mini-batches in its CSR (the first pass's coordinates).
__global__ void seg_prep_kernel(const int64_t *__restrict__ rowptr, int64_t row0, int64_t n_rows,
                                int32_t mini_batch, int n_segs, uint32_t *__restrict__ seg_off,
                                uint32_t *__restrict__ seg_tile_off)
{
    __shared__ uint32_t off[kMaxSegs + 1];
    const int t = threadIdx.x;
    const int64_t base = rowptr[row0];
    for (int s = t; s <= n_segs; s += blockDim.x) {
        int64_t r = (int64_t)s * mini_batch;
        if (r > n_rows) r = n_rows;
        off[s] = (uint32_t)(rowptr[row0 + r] - base);
        seg_off[s] = off[s];
    }
    __syncthreads();
    if (t == 0) {
        uint32_t run = 0;
        for (int s = 0; s < n_segs; s++) {
            seg_tile_off[s] = run;
            run += (off[s + 1] - off[s] + kSortTile - 1) / kSortTile;
        }
        seg_tile_off[n_segs] = run;
    }
}

// Which segment does global tile index `tile` belong to?  (last s with tile_off[s] <= tile)
__device__ __forceinline__ int find_seg(const uint32_t *tile_off, int n_segs, uint32_t tile)
{
    int lo = 0, hi = n_segs;  // invariant: tile_off[lo] <= tile < tile_off[hi]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (tile_off[mid] <= tile) lo = mid; else hi = mid;
    }
    return lo;
}

// ---- hot features -----------------------------------------------------------------------
__device__ __forceinline__ bool hot_hit(const uint32_t *tab, uint32_t key) { return tab[hot_index(key)] == key; }

// Called by every thread of the CTA (two barriers): the block's hot keys into the direct-mapped table.
__device__ __forceinline__ uint32_t hot_tab_build(uint32_t *tab, const HotInfo *hot, int nthreads)
{
    const uint32_t n = hot->count;
    for (int q = threadIdx.x; q < kHotSlots; q += nthreads) tab[q] = kHotEmpty;
    __syncthreads();
    if (threadIdx.x < n) { const uint32_t key = hot->key[threadIdx.x]; tab[hot_index(key)] = key; }
    __syncthreads();
    return n;
}

// Pick the block's hot features: count the features of a sample of its rows (kSelRows rows spread
// evenly over the block, the first kSelPerRow elements of each) in a shared-memory hash table
// (integer atomics: the counts do not depend on the order), keep those seen in at least 1/32 of the
// sampled rows, order them by (count descending, feature ascending), drop those whose table index
// (hot_index) is taken by a better-placed one, and keep the first kMaxHot.
// One CTA; a few microseconds on the sort stream.
constexpr int kSelThreads = 1024;
constexpr int kSelRows = 256, kSelPerRow = 32;
constexpr int kSelSlots = 16384;                       // >= 2 x the sample
constexpr int kSelCand = 1024;
constexpr size_t kSelSmem = (size_t)kSelSlots * 8 + (size_t)kSelCand * 12;

__global__ void __launch_bounds__(kSelThreads)
hot_select_kernel(const int64_t *__restrict__ rowptr, int64_t row0, int64_t n_rows,
                  const uint32_t *__restrict__ feature, uint32_t dims, HotInfo *__restrict__ out)
{
    extern __shared__ __align__(16) unsigned char sel_smem[];
    uint32_t *t_key = reinterpret_cast<uint32_t *>(sel_smem);
    uint32_t *t_cnt = t_key + kSelSlots;
    uint32_t *c_key = t_cnt + kSelSlots;
    uint32_t *c_cnt = c_key + kSelCand;
    uint32_t *c_rank = c_cnt + kSelCand;     // rank by (count descending, feature ascending); 0xffffffff: dropped
    __shared__ int s_ncand, s_nhot;
    const int tid = threadIdx.x;
    for (int q = tid; q < kSelSlots; q += kSelThreads) { t_key[q] = kHotEmpty; t_cnt[q] = 0; }
    if (tid == 0) { s_ncand = 0; s_nhot = 0; }
    __syncthreads();
    const int rows = (int)min((int64_t)kSelRows, n_rows);
    for (int q = tid; q < rows * kSelPerRow; q += kSelThreads) {
        const int i = q / kSelPerRow, e = q % kSelPerRow;
        const int64_t r = row0 + (n_rows * i) / rows;            // evenly spaced rows of the block
        const int64_t p = rowptr[r] + e;
        if (p < rowptr[r + 1]) {
            const uint32_t key = __ldg(feature + p);
            if (key < dims) {
                uint32_t h = (key * 0x9E3779B1u) >> 18;          // kSelSlots = 2^14
                while (true) {
                    const uint32_t prev = atomicCAS(&t_key[h], kHotEmpty, key);
                    if (prev == kHotEmpty || prev == key) { atomicAdd(&t_cnt[h], 1u); break; }
                    h = (h + 1) & (kSelSlots - 1);
                }
            }
        }
    }
    __syncthreads();
#ifndef HM_HOT_DIV
#define HM_HOT_DIV 32
#endif
    const uint32_t threshold = (uint32_t)max(2, rows / HM_HOT_DIV);
    for (int q = tid; q < kSelSlots; q += kSelThreads) {
        if (t_key[q] != kHotEmpty && t_cnt[q] >= threshold) {
            const int c = atomicAdd(&s_ncand, 1);
            if (c < kSelCand) { c_key[c] = t_key[q]; c_cnt[c] = t_cnt[q]; }
        }
    }
    __syncthreads();
    const int nc = min(s_ncand, kSelCand);      // (more than kSelCand candidates cannot happen: 8192 / 8)
    for (int c = tid; c < nc; c += kSelThreads) {
        const uint32_t k = c_key[c], n = c_cnt[c];
        uint32_t rank = 0;
        for (int o = 0; o < nc; o++) rank += (c_cnt[o] > n || (c_cnt[o] == n && c_key[o] < k)) ? 1u : 0u;
        c_rank[c] = rank;
    }
    __syncthreads();
    // a candidate whose table index belongs to a better-ranked candidate is dropped (the hot test is a
    // single probe); decided against ALL better-ranked candidates, so the outcome needs no ordering
    bool keep = false;
    uint32_t my_rank = 0;
    if (tid < nc) {
        my_rank = c_rank[tid];
        const uint32_t idx = hot_index(c_key[tid]);
        keep = true;
        for (int o = 0; o < nc; o++) if (c_rank[o] < my_rank && hot_index(c_key[o]) == idx) keep = false;
    }
    __syncthreads();
    if (tid < nc && !keep) c_rank[tid] = 0xffffffffu;
    __syncthreads();
    if (tid < nc && keep) {
        int pos = 0;
        for (int o = 0; o < nc; o++) pos += (c_rank[o] < my_rank) ? 1 : 0;      // kept candidates ahead of me
        if (pos < kMaxHot) { out->key[pos] = c_key[tid]; atomicAdd(&s_nhot, 1); }
    }
    __syncthreads();
    if (tid == 0) out->count = (uint32_t)s_nhot;
}

// One thread per sort tile: the row of the mini-batch that holds the tile's first element.
__global__ void tile_row0_kernel(const int64_t *__restrict__ rowptr, int64_t block_row0, int64_t block_rows,
                                 int32_t mini_batch, const uint32_t *__restrict__ seg_off,
                                 const uint32_t *__restrict__ seg_tile_off, int n_segs, int32_t *__restrict__ tile_row0)
{
    const uint32_t total = seg_tile_off[n_segs];
    const int64_t base = rowptr[block_row0];
    for (uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x; tile < total; tile += gridDim.x * blockDim.x) {
        const int seg = find_seg(seg_tile_off, n_segs, tile);
        const uint32_t e0 = seg_off[seg] + (tile - seg_tile_off[seg]) * kSortTile;
        const int64_t *rp = rowptr + block_row0 + (int64_t)seg * mini_batch;
        int lo = 0, hi = (int)min((int64_t)mini_batch, block_rows - (int64_t)seg * mini_batch);
        while (hi - lo > 1) {   // rp[lo] - base <= e0 < rp[hi] - base
            const int mid = (lo + hi) >> 1;
            if ((uint32_t)(rp[mid] - base) <= e0) lo = mid; else hi = mid;
        }
        tile_row0[tile] = lo;
    }
}

// FIRST: the keys are the block's CSR features; hot-feature and out-of-range elements are not counted.
template <int RB, bool FIRST>
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ seg_off,
                 const uint32_t *__restrict__ seg_tile_off, int n_segs, int shift,
                 uint32_t *__restrict__ tile_hist, uint32_t *__restrict__ row_sum,
                 uint32_t dims, const HotInfo *__restrict__ hot)
{
    constexpr int R = 1 << RB;
    __shared__ uint32_t s_off[kMaxSegs + 1], s_toff[kMaxSegs + 1];
    __shared__ uint32_t hist[kSortWarps][R];
    __shared__ uint32_t s_htab[FIRST ? kHotSlots : 1];
    uint32_t n_hot = 0;
    if (FIRST) n_hot = hot_tab_build(s_htab, hot, kSortThreads);
    for (int s = threadIdx.x; s <= n_segs; s += kSortThreads) { s_off[s] = seg_off[s]; s_toff[s] = seg_tile_off[s]; }
    __syncthreads();
    const uint32_t total = s_toff[n_segs];
    const int warp = threadIdx.x >> 5;
    for (uint32_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int seg = find_seg(s_toff, n_segs, tile);
        const uint32_t tis = tile - s_toff[seg], nt = s_toff[seg + 1] - s_toff[seg];
        const uint32_t e0 = s_off[seg] + tis * kSortTile;
        const uint32_t e1 = min(e0 + (uint32_t)kSortTile, s_off[seg + 1]);
        for (int q = threadIdx.x; q < kSortWarps * R; q += kSortThreads) (&hist[0][0])[q] = 0;
        __syncthreads();
        uint32_t k[kSortItems];
#pragma unroll
        for (int i = 0; i < kSortItems; i++) {
            const uint32_t e = e0 + i * kSortThreads + threadIdx.x;
            k[i] = e < e1 ? __ldg(keys + e) : 0u;
        }
#pragma unroll
        for (int i = 0; i < kSortItems; i++) {
            const uint32_t e = e0 + i * kSortThreads + threadIdx.x;
            bool count = e < e1;
            if (FIRST) count = count && k[i] < dims && !(n_hot && hot_hit(s_htab, k[i]));
            if (count) atomicAdd(&hist[warp][(k[i] >> shift) & (R - 1)], 1u);
        }
        __syncthreads();
        for (int d = threadIdx.x; d < R; d += kSortThreads) {
            uint32_t c = 0;
#pragma unroll
            for (int w = 0; w < kSortWarps; w++) c += hist[w][d];
            tile_hist[(size_t)s_toff[seg] * R + (size_t)d * nt + tis] = c;
            if (c) atomicAdd(&row_sum[seg * R + d], c);
        }
        __syncthreads();
    }
}

// In-place exclusive scan of row_sum[0..n) by one CTA of 1024 threads.  After the first pass's
// histogram (seg_off != NULL) the scan also yields where every mini-batch starts in the sorted
// arrays: row_sum[s * radix] = elements counted before mini-batch s.
__global__ void __launch_bounds__(1024) scan_rows_kernel(uint32_t *__restrict__ row_sum, int n,
                                                         uint32_t *__restrict__ seg_off, uint32_t *__restrict__ seg_tile_off,
                                                         int n_segs, int radix)
{
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t s_total, s_so[kMaxSegs + 1];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int per = (n + 1023) / 1024;
    const int lo = t * per, hi = min(lo + per, n);
    uint32_t s = 0;
    for (int i = lo; i < hi; i++) s += row_sum[i];
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
    }
    __syncthreads();
    uint32_t run = warp_tot[warp] + inc - s;
    for (int i = lo; i < hi; i++) { uint32_t v = row_sum[i]; row_sum[i] = run; run += v; }
    if (!seg_off) return;
    if (t == 1023) s_total = run;       // (threads past the end carry the grand total)
    __syncthreads();
    for (int sg = t; sg <= n_segs; sg += 1024) {
        const uint32_t o = sg < n_segs ? row_sum[sg * radix] : s_total;
        s_so[sg] = o;
        seg_off[sg] = o;
    }
    __syncthreads();
    if (t == 0) {
        uint32_t tiles = 0;
        for (int sg = 0; sg < n_segs; sg++) {
            seg_tile_off[sg] = tiles;
            tiles += (s_so[sg + 1] - s_so[sg] + kSortTile - 1) / kSortTile;
        }
        seg_tile_off[n_segs] = tiles;
    }
}

// One warp per (segment, digit) row of tile_hist: exclusive scan along the row, plus row base.
__global__ void __launch_bounds__(kSortThreads)
scan_tiles_kernel(uint32_t *__restrict__ tile_hist, const uint32_t *__restrict__ row_base,
                  const uint32_t *__restrict__ seg_tile_off, int n_segs, int radix)
{
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * kSortWarps + (threadIdx.x >> 5);
    if (row >= n_segs * radix) return;
    const int seg = row / radix, d = row % radix;
    const uint32_t t0 = seg_tile_off[seg], nt = seg_tile_off[seg + 1] - t0;
    uint32_t *p = tile_hist + (size_t)t0 * radix + (size_t)d * nt;
    uint32_t run = row_base[row];
    for (uint32_t k0 = 0; k0 < nt; k0 += 32) {
        const uint32_t k = k0 + lane;
        const uint32_t v = k < nt ? p[k] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (k < nt) p[k] = run + inc - v;
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
}

// First pass only: the row of an element is not stored anywhere, it is looked up in the CSR row
// pointer.  Per tile one warp finds the row holding the tile's first element (32-ary search, a
// handful of L2 round trips), the CTA stages the row-pointer slice that covers the tile in shared
// memory, and every element binary-searches that slice.
struct RowLookup {
    const int64_t *rowptr;   // chunk row pointer
    int64_t block_row0;      // first row of the block
    int64_t block_rows;      // rows in the block
    int32_t mini_batch;
    const int32_t *tile_row0; // first row (inside its mini-batch) of every sort tile
};
constexpr int kRowSlice = 1024;   // staged row-pointer entries; tiles that need more fall back to global searches

constexpr size_t kScatterSmem = (size_t)kSortTile * 4 + (size_t)kSortTile * 8;   // keys + payloads

template <bool FIRST, int RB>
__global__ void __launch_bounds__(kScatThreads, kScatCtasPerSm)
sort_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint2 *__restrict__ pay_in,
                    RowLookup rl, const float *__restrict__ val_in,
                    uint32_t *__restrict__ keys_out, uint2 *__restrict__ pay_out,
                    const uint32_t *__restrict__ seg_off, const uint32_t *__restrict__ seg_tile_off,
                    int n_segs, int shift, const uint32_t *__restrict__ tile_hist,
                    uint32_t dims, const HotInfo *__restrict__ hot, uint32_t *__restrict__ bad)
{
    // FIRST: seg_off / seg_tile_off are the CSR's (in_off / in_tile_off); elements of hot features and
    // out-of-range features take no part (the histogram did not count them either).
    constexpr int R = 1 << RB;
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    uint32_t *s_key = reinterpret_cast<uint32_t *>(dyn_smem);
    uint2 *s_pay = reinterpret_cast<uint2 *>(dyn_smem + (size_t)kSortTile * 4);
    __shared__ uint32_t s_off[kMaxSegs + 1], s_toff[kMaxSegs + 1];
    __shared__ uint32_t wcnt[kScatWarps][R];
    __shared__ uint32_t dig_base[R], gofs[R], warp_tot[R / 32];
    __shared__ uint32_t s_rows[FIRST ? kRowSlice + 1 : 1];
    __shared__ uint32_t s_mark[FIRST ? kSortTile / 32 : 1];   // bit p: a row starts at tile position p
    __shared__ uint32_t s_dup;                                // two rows start at one position (an empty row)
    __shared__ int s_nr;
    __shared__ uint32_t s_ntile;                              // elements of the tile that are sorted
    uint32_t *s_htab = reinterpret_cast<uint32_t *>(dyn_smem + kScatterSmem);     // (first pass only: kHotSlots entries)
    uint32_t n_hot = 0;
    if (FIRST) {
        n_hot = hot_tab_build(s_htab, hot, kScatThreads);
        for (int q = threadIdx.x; q < kSortTile / 32; q += kScatThreads) s_mark[q] = 0;
        if (threadIdx.x == 0) s_dup = 0;
    }
    for (int s = threadIdx.x; s <= n_segs; s += kScatThreads) { s_off[s] = seg_off[s]; s_toff[s] = seg_tile_off[s]; }
    __syncthreads();
    const uint32_t total = s_toff[n_segs];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    // keys of the first tile; later tiles are prefetched while the previous one is written out
    uint32_t key[kScatItems], rank[kScatItems];
    uint32_t tile = blockIdx.x;
    int seg = 0; uint32_t tis = 0, nt = 0, e0 = 0, e1 = 0;
    int cur_ra = 0; uint32_t slice_v = 0;   // first pass: the tile's first row and one row-pointer entry per thread
    // row-pointer entry `ra + tid` of a tile, relative to the block (first pass only)
    auto load_slice = [&](int sg, uint32_t tl, int *ra_out) -> uint32_t {
        const int ra = rl.tile_row0[tl];
        *ra_out = ra;
        const int mb_rows = (int)min((int64_t)rl.mini_batch, rl.block_rows - (int64_t)sg * rl.mini_batch);
        const int64_t r = (int64_t)ra + threadIdx.x;
        if (r > mb_rows) return 0xffffffffu;
        return (uint32_t)(rl.rowptr[rl.block_row0 + (int64_t)sg * rl.mini_batch + r] - rl.rowptr[rl.block_row0]);
    };
    if (tile < total) {
        seg = find_seg(s_toff, n_segs, tile);
        tis = tile - s_toff[seg]; nt = s_toff[seg + 1] - s_toff[seg];
        e0 = s_off[seg] + tis * kSortTile;
        e1 = min(e0 + (uint32_t)kSortTile, s_off[seg + 1]);
        if (FIRST) slice_v = load_slice(seg, tile, &cur_ra);
#pragma unroll
        for (int i = 0; i < kScatItems; i++) {
            const uint32_t e = e0 + warp * (32 * kScatItems) + lane + i * 32;
            key[i] = e < e1 ? __ldcs(keys_in + e) : 0xffffffffu;
        }
    }
    for (; tile < total; tile += gridDim.x) {
        // warp-striped: warp w owns tile positions [w*256, (w+1)*256), round i, lane
        const uint32_t ew = e0 + warp * (32 * kScatItems) + lane;
        // payloads are requested now and consumed after the ranking
        uint2 pay[kScatItems];
        int nr = 0; bool covered = false;
        if (FIRST) {
            // the values are requested before the row search, so that their latency hides behind it
#pragma unroll
            for (int i = 0; i < kScatItems; i++) {
                const uint32_t e = ew + i * 32;
                if (e < e1) pay[i].y = __float_as_uint(__ldcs(val_in + e));
            }
            const int mb_rows = (int)min((int64_t)rl.mini_batch, rl.block_rows - (int64_t)seg * rl.mini_batch);
            // stage rp[ra ..] - base.  The first kScatThreads entries were prefetched into `slice_v`
            // with the keys and almost always cover the tile (one barrier); otherwise keep loading.
            nr = min(kScatThreads - 1, mb_rows - cur_ra);
            if ((int)threadIdx.x <= nr) s_rows[threadIdx.x] = slice_v;
            __syncthreads();
            covered = s_rows[nr] >= e1;       // uniform
            if (covered) {
                // Fast path: every row that starts inside the tile sets the bit of its start position; the row
                // of an element is then (tile's first row) + (bits at or below its position).  Two rows on one
                // position -- an empty row -- cannot be told apart that way: the tile takes the search path.
                if (threadIdx.x >= 1 && (int)threadIdx.x <= nr && slice_v > e0 && slice_v < e1) {
                    const uint32_t p = slice_v - e0, bit = 1u << (p & 31);
                    if (atomicOr(&s_mark[p >> 5], bit) & bit) s_dup = 1;
                }
            } else {
                const int64_t base = rl.rowptr[rl.block_row0];
                const int64_t *rp = rl.rowptr + rl.block_row0 + (int64_t)seg * rl.mini_batch;
                const int ra = cur_ra;
                nr = -1;
                for (int i0 = kScatThreads; i0 <= kRowSlice; i0 += kScatThreads) {
                    if (threadIdx.x == 0) s_nr = 0x7fffffff;
                    __syncthreads();
                    const int i = i0 + threadIdx.x;
                    int hit = 0x7fffffff;
                    if (i <= kRowSlice && ra + i <= mb_rows) {
                        const uint32_t v = (uint32_t)(rp[ra + i] - base);
                        s_rows[i] = v;
                        if (v >= e1) hit = i;
                    }
                    const int wmin = __reduce_min_sync(0xffffffffu, hit);
                    if (lane == 0 && wmin != 0x7fffffff) atomicMin(&s_nr, wmin);
                    __syncthreads();
                    const int got = s_nr;
                    __syncthreads();
                    if (got != 0x7fffffff) { nr = got; break; }
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < kScatItems; i++) {
                const uint32_t e = ew + i * 32;
                if (e < e1) pay[i] = __ldcs(pay_in + e);
            }
        }
        for (int q = threadIdx.x; q < kScatWarps * R; q += kScatThreads) (&wcnt[0][0])[q] = 0;
        __syncthreads();
        if (FIRST) {
            const int ra = cur_ra;
            if (covered && !s_dup) {
                // rows that start before this warp's 256 positions, then one word of marks per round
                uint32_t before = 0;
#pragma unroll
                for (int j = 0; j < kSortTile / 32 / 32; j++) {
                    const int wd = lane + 32 * j;
                    if (wd < warp * kScatItems) before += __popc(s_mark[wd]);
                }
                before = __reduce_add_sync(0xffffffffu, before);
                const uint32_t le_mask = 0xffffffffu >> (31 - lane);
#pragma unroll
                for (int i = 0; i < kScatItems; i++) {
                    const uint32_t m = s_mark[warp * kScatItems + i];
                    pay[i].x = (uint32_t)ra + before + __popc(m & le_mask);
                    before += __popc(m);
                }
            } else {
                const int64_t base = rl.rowptr[rl.block_row0];
                const int mb_rows = (int)min((int64_t)rl.mini_batch, rl.block_rows - (int64_t)seg * rl.mini_batch);
                const int64_t *rp = rl.rowptr + rl.block_row0 + (int64_t)seg * rl.mini_batch;
                // a thread's elements are 32 apart and ascending: search once, then walk forward
                int idx = 0;
                bool searched = false;
#pragma unroll
                for (int i = 0; i < kScatItems; i++) {
                    const uint32_t e = ew + i * 32;
                    if (e < e1) {
                        int row;
                        if (nr >= 0) {
                            if (!searched) {
                                int lo = 0, hi = nr;        // s_rows[lo] <= e < s_rows[hi]
                                while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_rows[mid] <= e) lo = mid; else hi = mid; }
                                idx = lo; searched = true;
                            } else {
                                while (s_rows[idx + 1] <= e) idx++;     // s_rows[nr] >= e1 > e bounds the walk
                            }
                            row = ra + idx;
                        } else {
                            int lo = ra, hi = mb_rows;  // slice too small (very short rows): search the row pointer itself
                            while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if ((uint32_t)(rp[mid] - base) <= e) lo = mid; else hi = mid; }
                            row = lo;
                        }
                        pay[i].x = (uint32_t)row;
                    }
                }
            }
        }
        // stable rank of every element among equal digits inside its warp
        uint32_t vbits = 0;      // bit i: item i takes part in the sort
#pragma unroll
        for (int i = 0; i < kScatItems; i++) {
            bool valid = ew + i * 32 < e1;
            if (FIRST && valid) {
                if (key[i] >= dims) { valid = false; atomicOr(bad, 1u); }
                else if (n_hot && hot_hit(s_htab, key[i])) valid = false;
            }
            if (valid) vbits |= 1u << i;
            const uint32_t d = (key[i] >> shift) & (R - 1);
            const uint32_t active = __ballot_sync(0xffffffffu, valid);
            uint32_t peers = 0, prev = 0;
            if (FIRST && n_hot == 0u) {
                // match.any costs time in proportion to the number of distinct values in the warp.  In the
                // first pass the elements are still in row order (32 consecutive elements = 32 different
                // hashed features): one ballot per digit bit -- lanes that agree with mine in every bit -- is
                // cheaper there.  Later passes see long runs of equal digits and keep match.any -- and so does
                // the first pass of a block with hot features: half of its lanes are dropped before the
                // ranking, which halves what match.any costs (measured: 134 -> 128 us per launch on average).
                peers = active;
#pragma unroll
                for (int b = 0; b < RB; b++) {
                    const uint32_t m = __ballot_sync(0xffffffffu, valid && ((d >> b) & 1u));
                    peers &= ((d >> b) & 1u) ? m : ~m;
                }
                if (valid) prev = wcnt[warp][d];
            } else
            if (valid) { peers = __match_any_sync(active, d); prev = wcnt[warp][d]; }
            __syncwarp();
            if (valid) {
                const uint32_t r = __popc(peers & lt_mask);
                if (r == 0) wcnt[warp][d] = prev + __popc(peers);
                rank[i] = prev + r;
            }
            __syncwarp();
        }
        __syncthreads();
        // per-digit: exclusive scan over warps, then exclusive scan over digits
        {
            // the first R threads own one digit each
            const int d = threadIdx.x;
            const bool owner = d < R;
            uint32_t run = 0;
            if (owner) {
#pragma unroll
                for (int w = 0; w < kScatWarps; w++) { uint32_t c = wcnt[w][d]; wcnt[w][d] = run; run += c; }
            }
            uint32_t inc = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
            if (lane == 31 && owner) warp_tot[warp] = inc;
            __syncthreads();
            if (owner) {
                uint32_t wbase = 0;
#pragma unroll
                for (int w = 0; w < R / 32; w++) if (w < warp) wbase += warp_tot[w];
                const uint32_t ex = wbase + inc - run;
                dig_base[d] = ex;
                gofs[d] = tile_hist[(size_t)s_toff[seg] * R + (size_t)d * nt + tis] - ex;
                if (d == R - 1) s_ntile = ex + run;
            }
        }
        __syncthreads();
        // place keys and payloads at their tile-local sorted position (payload loaded only now:
        // sixteen independent loads in flight per thread, no registers held across the ranking)
#pragma unroll
        for (int i = 0; i < kScatItems; i++) {
            if ((vbits >> i) & 1u) {
                const uint32_t d = (key[i] >> shift) & (R - 1);
                const uint32_t pos = dig_base[d] + wcnt[warp][d] + rank[i];
                s_key[pos] = key[i];
                s_pay[pos] = pay[i];
            }
        }
        __syncthreads();
        if (FIRST) {      // marks are consumed: clear them for the next tile (its marks come after its slice barrier)
            for (int q = threadIdx.x; q < kSortTile / 32; q += kScatThreads) s_mark[q] = 0;
            if (threadIdx.x == 0) s_dup = 0;
        }
        const uint32_t n_tile = s_ntile;
        // next tile's coordinates and keys (the key registers are free from here on)
        const uint32_t next = tile + gridDim.x;
        int nseg = 0; uint32_t ntis = 0, nnt = 0, ne0 = 0, ne1 = 0;
        if (next < total) {
            nseg = find_seg(s_toff, n_segs, next);
            ntis = next - s_toff[nseg]; nnt = s_toff[nseg + 1] - s_toff[nseg];
            ne0 = s_off[nseg] + ntis * kSortTile;
            ne1 = min(ne0 + (uint32_t)kSortTile, s_off[nseg + 1]);
            if (FIRST) slice_v = load_slice(nseg, next, &cur_ra);
#pragma unroll
            for (int i = 0; i < kScatItems; i++) {
                const uint32_t e = ne0 + warp * (32 * kScatItems) + lane + i * 32;
                key[i] = e < ne1 ? __ldcs(keys_in + e) : 0xffffffffu;
            }
        }
        for (uint32_t p = threadIdx.x; p < n_tile; p += kScatThreads) {
            const uint32_t k = s_key[p];
            const uint32_t g = gofs[(k >> shift) & (R - 1)] + p;
            keys_out[g] = k;
            pay_out[g] = s_pay[p];
        }
        __syncthreads();
        seg = nseg; tis = ntis; nt = nnt; e0 = ne0; e1 = ne1;
    }
}

int persistent_grid(int64_t work_items, int per_sm)
{
    int64_t cap = (int64_t)num_sms() * per_sm;
    if (work_items < 1) work_items = 1;
    return (int)(work_items < cap ? work_items : cap);
}

}  // namespace

hm_status sort_scratch_reserve(SortScratch &s, int64_t elems)
{
    if (elems < 1) elems = 1;
    if (!s.seg_off) {
        HM_CUDA(cudaMalloc(&s.in_off, (kMaxSegs + 1) * sizeof(uint32_t)));
        HM_CUDA(cudaMalloc(&s.in_tile_off, (kMaxSegs + 1) * sizeof(uint32_t)));
        HM_CUDA(cudaMalloc(&s.seg_off, (kMaxSegs + 1) * sizeof(uint32_t)));
        HM_CUDA(cudaMalloc(&s.seg_tile_off, (kMaxSegs + 1) * sizeof(uint32_t)));
        HM_CUDA(cudaMalloc(&s.row_sum, (size_t)kMaxSegs * kMaxRadix * sizeof(uint32_t)));
        HM_CUDA(cudaMalloc(&s.hot, sizeof(HotInfo)));
        HM_CUDA(cudaMemset(s.hot, 0, sizeof(HotInfo)));
    }
    if (elems > s.cap_elems) {
        for (int b = 0; b < 2; b++) {
            if (s.key[b]) cudaFree(s.key[b]);
            if (s.pay[b]) cudaFree(s.pay[b]);
            s.key[b] = nullptr; s.pay[b] = nullptr;
        }
        if (s.tile_hist) cudaFree(s.tile_hist);
        if (s.tile_row0) cudaFree(s.tile_row0);
        s.tile_hist = nullptr; s.tile_row0 = nullptr;
        s.cap_elems = 0;
        int64_t cap = elems + elems / 8;
        for (int b = 0; b < 2; b++) {
            HM_CUDA(cudaMalloc(&s.key[b], (size_t)cap * sizeof(uint32_t)));
            HM_CUDA(cudaMalloc(&s.pay[b], (size_t)cap * sizeof(uint2)));
        }
        s.cap_tiles = cap / kSortTile + kMaxSegs + 1;
        HM_CUDA(cudaMalloc(&s.tile_hist, (size_t)s.cap_tiles * kMaxRadix * sizeof(uint32_t)));
        HM_CUDA(cudaMalloc(&s.tile_row0, (size_t)s.cap_tiles * sizeof(int32_t)));
        s.cap_elems = cap;
    }
    return HM_OK;
}

void sort_scratch_free(SortScratch &s)
{
    for (int b = 0; b < 2; b++) { cudaFree(s.key[b]); cudaFree(s.pay[b]); s.key[b] = nullptr; s.pay[b] = nullptr; }
    cudaFree(s.tile_hist); cudaFree(s.row_sum); cudaFree(s.seg_off); cudaFree(s.seg_tile_off); cudaFree(s.tile_row0);
    cudaFree(s.in_off); cudaFree(s.in_tile_off); cudaFree(s.hot);
    s.tile_hist = s.row_sum = s.seg_off = s.seg_tile_off = s.in_off = s.in_tile_off = nullptr; s.tile_row0 = nullptr;
    s.hot = nullptr;
    s.cap_elems = s.cap_tiles = 0;
}

hm_status launch_seg_prep(Launcher &L, SortScratch &s, const int64_t *d_rowptr, int64_t row0,
                          int64_t n_rows, int64_t n_elems, int32_t mini_batch, int n_segs)
{
    if (n_segs < 1 || n_segs > kMaxSegs) HM_FAIL(HM_ERR_INVALID_ARGUMENT, "segment count %d out of range", n_segs);
    HM_LAUNCH(L, K_SEG_PREP, 0.0,
              (seg_prep_kernel<<<1, 256, 0, L.stream>>>(d_rowptr, row0, n_rows, mini_batch, n_segs, s.in_off,
                                                        s.in_tile_off)));
    // first row of every sort tile of the CSR (one thread per tile; used by the first scatter pass)
    const int64_t max_tiles = n_elems / kSortTile + n_segs;
    HM_LAUNCH(L, K_SEG_PREP, 0.0,
              (tile_row0_kernel<<<persistent_grid((max_tiles + 127) / 128, 8), 128, 0, L.stream>>>(
                  d_rowptr, row0, n_rows, mini_batch, s.in_off, s.in_tile_off, n_segs, s.tile_row0)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_hot_select(Launcher &L, SortScratch &s, const int64_t *d_rowptr, int64_t row0, int64_t n_rows,
                            const uint32_t *d_feature, int64_t dims, bool enable)
{
    if (!enable || n_rows < 1) {
        HM_CUDA(cudaMemsetAsync(s.hot, 0, sizeof(HotInfo), L.stream));
        return HM_OK;
    }
    static std::atomic<bool> attr_set[64];   // > 48 KB of dynamic shared memory needs opting in, once per device
    int dev = 0;
    HM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        HM_CUDA(cudaFuncSetAttribute(hot_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSelSmem));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    HM_LAUNCH(L, K_HOT_SELECT, 0.0,
              (hot_select_kernel<<<1, kSelThreads, kSelSmem, L.stream>>>(d_rowptr, row0, n_rows, d_feature,
                                                                          (uint32_t)std::min<int64_t>(dims, 0xffffffffLL), s.hot)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

namespace {
template <class K>
hm_status opt_in_smem(K kernel, size_t bytes)
{
    HM_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return HM_OK;
}
}  // namespace

hm_status launch_segmented_sort(Launcher &L, SortScratch &s, const uint32_t *d_feature_block,
                                const float *d_value_block, const int64_t *d_rowptr, int64_t block_row0,
                                int64_t block_rows, int32_t mini_batch, int64_t n_elems, int n_segs, int key_bits,
                                int64_t dims, uint32_t *d_bad, int *out_buf, int ctas_per_sm, std::vector<uint32_t> *seg_off_host)
{
    // > 48 KB of dynamic shared memory needs opting in, once per device (the attribute is per
    // device; the flags are atomics because learners on different threads may get here together)
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    HM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        HM_TRY(opt_in_smem(sort_scatter_kernel<true, 8>, kScatterSmem + kHotSlots * 4));
        HM_TRY(opt_in_smem(sort_scatter_kernel<false, 8>, kScatterSmem));
        HM_TRY(opt_in_smem(sort_scatter_kernel<true, 9>, kScatterSmem + kHotSlots * 4));
        HM_TRY(opt_in_smem(sort_scatter_kernel<false, 9>, kScatterSmem));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    if (n_elems > s.cap_elems) HM_FAIL(HM_ERR_STATE, "sort scratch too small (%lld > %lld)", (long long)n_elems,
                                       (long long)s.cap_elems);
    // 8-bit digits, or 9-bit ones where that saves a whole pass (17-18 and 25-27 key bits)
    if (key_bits < 1) key_bits = 1;
    int passes = (key_bits + 7) / 8;
    int wide = 0;                        // how many of the passes use 9 bits
    if ((key_bits + 8) / 9 < passes) { passes = (key_bits + 8) / 9; wide = std::max(0, key_bits - 8 * passes); }
    const int64_t max_tiles = n_elems / kSortTile + n_segs;
    // ctas_per_sm == 0: take the whole GPU; otherwise leave room for kernels of another stream
    const int grid = persistent_grid(max_tiles, ctas_per_sm > 0 ? ctas_per_sm : 6);
    const int hist_grid = persistent_grid(max_tiles, ctas_per_sm > 0 ? 2 * ctas_per_sm : 6);
    const uint32_t dims32 = (uint32_t)std::min<int64_t>(dims, 0xffffffffLL);
    RowLookup rl;
    rl.rowptr = d_rowptr; rl.block_row0 = block_row0; rl.block_rows = block_rows; rl.mini_batch = mini_batch;
    rl.tile_row0 = s.tile_row0;
    int shift = 0;
    double sorted_elems = (double)n_elems;     // (an upper bound unless profiling)
    for (int p = 0; p < passes; p++) {
        const int rb = p < wide ? 9 : 8;
        const int rows = n_segs << rb;
        const uint32_t *kin = p == 0 ? d_feature_block : s.key[(p - 1) & 1];
        const uint2 *pin = p == 0 ? nullptr : s.pay[(p - 1) & 1];
        uint32_t *kout = s.key[p & 1];
        uint2 *pout = s.pay[p & 1];
        // the first pass walks the block's CSR, later passes the (shorter) sorted arrays
        const uint32_t *off = p == 0 ? s.in_off : s.seg_off;
        const uint32_t *toff = p == 0 ? s.in_tile_off : s.seg_tile_off;
        HM_CUDA(cudaMemsetAsync(s.row_sum, 0, (size_t)rows * sizeof(uint32_t), L.stream));
#define HM_HIST(RB_, FIRST_)                                                                                            \
    HM_LAUNCH(L, K_SORT_HIST, 4.0 * (p == 0 ? (double)n_elems : sorted_elems),                                           \
              (sort_hist_kernel<RB_, FIRST_><<<hist_grid, kSortThreads, 0, L.stream>>>(kin, off, toff, n_segs, shift,  \
                                                                                       s.tile_hist, s.row_sum, dims32, s.hot)))
        if (p == 0) { if (rb == 9) HM_HIST(9, true); else HM_HIST(8, true); }
        else { if (rb == 9) HM_HIST(9, false); else HM_HIST(8, false); }
#undef HM_HIST
        // (the first scan also writes the sorted arrays' mini-batch and tile offsets)
        HM_LAUNCH(L, K_SORT_SCAN, 0.0, (scan_rows_kernel<<<1, 1024, 0, L.stream>>>(
                                           s.row_sum, rows, p == 0 ? s.seg_off : nullptr, s.seg_tile_off, n_segs, 1 << rb)));
        HM_LAUNCH(L, K_SORT_SCAN, 0.0,
                  (scan_tiles_kernel<<<(rows + kSortWarps - 1) / kSortWarps, kSortThreads, 0, L.stream>>>(
                      s.tile_hist, s.row_sum, toff, n_segs, 1 << rb)));
        // While profiling, the per-kernel byte counts use the SORTED element count, which only the device knows
        // (hot and out-of-range elements are gone): read the offsets back once the first scan has run.
        if (p == 0 && L.profiling && seg_off_host) {
            seg_off_host->assign((size_t)n_segs + 1, 0u);
            HM_CUDA(cudaMemcpyAsync(seg_off_host->data(), s.seg_off, (size_t)(n_segs + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, L.stream));
            HM_CUDA(cudaStreamSynchronize(L.stream));
            sorted_elems = (double)seg_off_host->back();
        }
        const double bytes = p == 0 ? 8.0 * (double)n_elems + 8.0 * (double)block_rows + 12.0 * sorted_elems : 24.0 * sorted_elems;
#define HM_SCATTER(FIRST_, RB_)                                                                                        \
    HM_LAUNCH(L, K_SORT_SCATTER, bytes,                                                                                 \
              (sort_scatter_kernel<FIRST_, RB_><<<grid, kScatThreads, kScatterSmem + (FIRST_ ? kHotSlots * 4 : 0), L.stream>>>( \
                  kin, pin, rl, p == 0 ? d_value_block : nullptr, kout, pout, off, toff, n_segs, shift,                 \
                  s.tile_hist, dims32, s.hot, d_bad)))
        if (p == 0) { if (rb == 9) HM_SCATTER(true, 9); else HM_SCATTER(true, 8); }
        else { if (rb == 9) HM_SCATTER(false, 9); else HM_SCATTER(false, 8); }
#undef HM_SCATTER
        shift += rb;
    }
    HM_CUDA(cudaGetLastError());
    *out_buf = (passes - 1) & 1;
    return HM_OK;
}

}  // namespace hm
