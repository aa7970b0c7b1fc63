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
// Two kinds of element never enter the sort: those of the block's dense columns (DenseInfo:
// their gradient is a column sum taken by the forward kernel, see hm_linear.cu) and those whose
// feature index is out of range (they are flagged and ignored everywhere).  Both are dropped by
// key in the first pass and its histogram, so the sorted arrays are shorter than the block's CSR:
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
    "linear_update", "fm_forward", "fm_update", "predict", "export", "misc", "dense_check"};

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
#ifndef HM_SCAT_CTAS
#define HM_SCAT_CTAS 2
#endif
constexpr int kScatCtasPerSm = HM_SCAT_CTAS;

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

// ---- dense columns --------------------------------------------------------------------
constexpr uint32_t kDenseEmpty = 0xffffffffu;
constexpr int kDenseSlots = 128;       // open-addressing table of at most 32 keys

__device__ __forceinline__ uint32_t dense_hash(uint32_t key) { return (key * 0x9E3779B1u) >> 25; }

// Index (0..31) of `key` among the table's keys, or -1.
__device__ __forceinline__ int dense_find(const uint32_t *tab, const uint32_t *tab_idx, uint32_t key)
{
    uint32_t h = dense_hash(key);
    while (true) {
        const uint32_t v = tab[h];
        if (v == key) return (int)tab_idx[h];
        if (v == kDenseEmpty) return -1;
        h = (h + 1) & (kDenseSlots - 1);
    }
}
__device__ __forceinline__ bool dense_hit(const uint32_t *tab, uint32_t key)
{
    uint32_t h = dense_hash(key);
    while (true) {
        const uint32_t v = tab[h];
        if (v == key) return true;
        if (v == kDenseEmpty) return false;
        h = (h + 1) & (kDenseSlots - 1);
    }
}
// Called by every thread of the CTA (two barriers).  Lane i of warp 0 inserts key i.
__device__ __forceinline__ void dense_tab_build(uint32_t *tab, uint32_t *tab_idx, uint32_t mask, uint32_t my_key, int nthreads)
{
    for (int q = threadIdx.x; q < kDenseSlots; q += nthreads) tab[q] = kDenseEmpty;
    __syncthreads();
    if (threadIdx.x < 32 && ((mask >> threadIdx.x) & 1u)) {
        uint32_t h = dense_hash(my_key);
        while (atomicCAS(&tab[h], kDenseEmpty, my_key) != kDenseEmpty) h = (h + 1) & (kDenseSlots - 1);
        if (tab_idx) tab_idx[h] = threadIdx.x;
    }
    __syncthreads();
}

constexpr int kChkThreads = 256;

// Which of the first 32 positions of the block's rows are dense columns (see DenseInfo)?  A warp
// per row reads the whole row once (4 B per element): position i stays a candidate while every
// row holds the first row's feature there, and loses the bit as soon as that feature turns up
// anywhere else.  The last CTA to finish publishes the result.
__global__ void __launch_bounds__(kChkThreads)
dense_check_kernel(const int64_t *__restrict__ rowptr, int64_t row0, int64_t n_rows,
                   const uint32_t *__restrict__ feature, const float *__restrict__ value, uint32_t dims,
                   DenseInfo *__restrict__ out)
{
    __shared__ uint32_t s_ref[32], s_tab[kDenseSlots], s_tab_idx[kDenseSlots];
    __shared__ uint32_t s_init, s_noteq, s_collide, s_nupos, s_nuother;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp == 0) {
        const int64_t p0 = rowptr[row0], p1 = rowptr[row0 + 1];
        const bool have = p0 + lane < p1;
        const uint32_t f = have ? __ldg(feature + p0 + lane) : kDenseEmpty;
        const uint32_t peers = __match_any_sync(0xffffffffu, f);
        const bool ok = have && f < dims && __popc(peers) == 1;     // in range, and once in the row
        const uint32_t init = __ballot_sync(0xffffffffu, ok);
        s_ref[lane] = f;
        if (lane == 0) { s_init = init; s_noteq = 0; s_collide = 0; s_nupos = 0; s_nuother = 0; }
    }
    __syncthreads();
    const uint32_t init = s_init;
    dense_tab_build(s_tab, s_tab_idx, init, s_ref[lane], kChkThreads);
    // (a block without a candidate column still learns whether its values are all 1.0)
    {
        const uint32_t ref = s_ref[lane];
        const bool cand = (init >> lane) & 1u;
        uint32_t noteq = 0, coll = 0;
        bool nu_pos = false, nu_other = false;     // a value != 1.0 at my candidate position / anywhere else
        // Four consecutive rows per warp at a time: their loads are all issued before any is used (the
        // kernel is bound by the latency of the rowptr -> feature chain, not by bytes).
        constexpr int RW = 4;
        const int64_t stride = (int64_t)gridDim.x * (kChkThreads / 32) * RW;
        for (int64_t r = ((int64_t)blockIdx.x * (kChkThreads / 32) + warp) * RW; r < n_rows; r += stride) {
            const int64_t rp = (lane <= RW && r + lane <= n_rows) ? rowptr[row0 + r + lane] : 0;
            int64_t p0[RW], p1[RW];
            bool in[RW], in2[RW];
            uint32_t f[RW], g[RW];
            float xv[RW], xg[RW];
#pragma unroll
            for (int u = 0; u < RW; u++) {
                p0[u] = __shfl_sync(0xffffffffu, rp, u);
                p1[u] = __shfl_sync(0xffffffffu, rp, u + 1);
                const bool row = r + u < n_rows;
                in[u] = row && p0[u] + lane < p1[u];
                in2[u] = row && p0[u] + 32 + lane < p1[u];
            }
#pragma unroll
            for (int u = 0; u < RW; u++) {
                f[u] = in[u] ? __ldg(feature + p0[u] + lane) : kDenseEmpty;
                xv[u] = in[u] ? __ldg(value + p0[u] + lane) : 1.0f;
            }
#pragma unroll
            for (int u = 0; u < RW; u++) {
                g[u] = in2[u] ? __ldg(feature + p0[u] + 32 + lane) : kDenseEmpty;
                xg[u] = in2[u] ? __ldg(value + p0[u] + 32 + lane) : 1.0f;
            }
#pragma unroll
            for (int u = 0; u < RW; u++) {
                if (r + u < n_rows) {      // (uniform over the warp)
                    const bool eq = in[u] && f[u] == ref;
                    noteq |= init & ~__ballot_sync(0xffffffffu, eq);
                    if (xv[u] != 1.0f) { if (eq && cand) nu_pos = true; else nu_other = true; }
                    if (xg[u] != 1.0f) nu_other = true;
                    if (init != 0) {
                        if (in[u] && !eq) { const int j = dense_find(s_tab, s_tab_idx, f[u]); if (j >= 0) coll |= 1u << j; }
                        if (in2[u]) { const int j = dense_find(s_tab, s_tab_idx, g[u]); if (j >= 0) coll |= 1u << j; }
                    }
                    for (int64_t p = p0[u] + 64 + lane; p < p1[u]; p += 32) {
                        if (init != 0) { const int j = dense_find(s_tab, s_tab_idx, __ldg(feature + p)); if (j >= 0) coll |= 1u << j; }
                        if (__ldg(value + p) != 1.0f) nu_other = true;
                    }
                }
            }
        }
        const uint32_t nupos = __ballot_sync(0xffffffffu, nu_pos);
        const bool nuother = __any_sync(0xffffffffu, nu_other);
        if (lane == 0) { if (nupos) atomicOr(&s_nupos, nupos); if (nuother) atomicOr(&s_nuother, 1u); }
        coll = __reduce_or_sync(0xffffffffu, coll);
        if (lane == 0) { if (noteq) atomicOr(&s_noteq, noteq); if (coll) atomicOr(&s_collide, coll); }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_noteq) atomicOr(&out->noteq, s_noteq);
        if (s_collide) atomicOr(&out->collide, s_collide);
        if (s_nupos) atomicOr(&out->nonunit_pos, s_nupos);
        if (s_nuother) atomicOr(&out->nonunit_other, 1u);
        __threadfence();
        if (atomicAdd(&out->ticket, 1u) == gridDim.x - 1) {
            __threadfence();
            const uint32_t mask = init & ~atomicOr(&out->noteq, 0u) & ~atomicOr(&out->collide, 0u);
            for (int i = 0; i < 32; i++) out->key[i] = s_ref[i];
            out->mask = mask;
            out->count = (uint32_t)__popc(mask);
            // all ones among what is sorted: nothing != 1.0 outside the candidate columns, and none in a
            // candidate column that did not become a dense column after all
            out->unit = (atomicOr(&out->nonunit_other, 0u) == 0u &&
                         ((init & ~mask) & atomicOr(&out->nonunit_pos, 0u)) == 0u) ? 1u : 0u;
            out->noteq = 0; out->collide = 0; out->ticket = 0; out->nonunit_pos = 0; out->nonunit_other = 0;   // ready for the next block
        }
    }
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

// FIRST: the keys are the block's CSR features; dense-column and out-of-range elements are not counted.
template <int RB, bool FIRST>
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const uint32_t *__restrict__ keys, const uint2 *__restrict__ recs,
                 const uint32_t *__restrict__ seg_off,
                 const uint32_t *__restrict__ seg_tile_off, int n_segs, int shift,
                 uint32_t *__restrict__ tile_hist, uint32_t *__restrict__ row_sum,
                 uint32_t dims, const DenseInfo *__restrict__ dense)
{
    // later passes of a block whose sorted values are all 1.0: the keys sit in packed (feature, row) records
    const bool unit = !FIRST && dense->unit != 0u;
    constexpr int R = 1 << RB;
    __shared__ uint32_t s_off[kMaxSegs + 1], s_toff[kMaxSegs + 1];
    __shared__ uint32_t hist[kSortWarps][R];
    __shared__ uint32_t s_dtab[FIRST ? kDenseSlots : 1];
    uint32_t dmask = 0;
    if (FIRST) {
        dmask = dense->mask;
        dense_tab_build(s_dtab, nullptr, dmask, dense->key[threadIdx.x & 31], kSortThreads);
    }
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
            k[i] = e < e1 ? (unit ? __ldg(&recs[e].x) : __ldg(keys + e)) : 0u;
        }
#pragma unroll
        for (int i = 0; i < kSortItems; i++) {
            const uint32_t e = e0 + i * kSortThreads + threadIdx.x;
            bool count = e < e1;
            if (FIRST) count = count && k[i] < dims && !(dmask && dense_hit(s_dtab, k[i]));
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

// Shared memory of one scatter CTA (declared once in the kernel; the two bodies below share it).
template <bool FIRST, int RB>
struct ScatSmem {
    static constexpr int R = 1 << RB;
    uint32_t off[kMaxSegs + 1], toff[kMaxSegs + 1];
    uint32_t wcnt[kScatWarps][R];
    uint32_t dig_base[R], gofs[R], warp_tot[R / 32];
    uint32_t rows[FIRST ? kRowSlice + 1 : 1];
    uint32_t mark[FIRST ? kSortTile / 32 : 1];   // bit p: a row starts at tile position p
    uint32_t dup;                                // two rows start at one position (an empty row)
    int nr;
    uint32_t ntile;                              // elements of the tile that are sorted
    uint32_t dtab[FIRST ? kDenseSlots : 1];
};

// One digit pass over the tiles of a block.
//   FIRST: the input is the block's CSR (features in keys_in, values in val_in, rows looked up in the row
//          pointer); seg_off / seg_tile_off are the CSR's (in_off / in_tile_off); elements of dense columns and
//          out-of-range features take no part (the histogram did not count them either).
//   UNIT : every sorted element has value 1.0 (DenseInfo::unit): input and output of the later passes are packed
//          (feature, row) records of 8 bytes in the payload arrays; the key arrays and the values are not touched.
template <bool FIRST, int RB, bool UNIT>
__device__ __forceinline__ void
scatter_body(ScatSmem<FIRST, RB> &sm, unsigned char *dyn_smem, const uint32_t *__restrict__ keys_in,
             const uint2 *__restrict__ pay_in, const RowLookup &rl, const float *__restrict__ val_in,
             uint32_t *__restrict__ keys_out, uint2 *__restrict__ pay_out, int n_segs, int shift,
             const uint32_t *__restrict__ tile_hist, uint32_t dims, uint32_t dmask, uint32_t *__restrict__ bad)
{
    constexpr int R = 1 << RB;
    // staging of the tile in sorted order: keys + payloads, or packed records
    uint32_t *s_key = reinterpret_cast<uint32_t *>(dyn_smem);
    uint2 *s_pay = reinterpret_cast<uint2 *>(dyn_smem + (UNIT ? 0 : (size_t)kSortTile * 4));
    const uint32_t total = sm.toff[n_segs];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    // keys of the first tile; later tiles are prefetched while the previous one is written out
    uint32_t key[kScatItems], rank[kScatItems];
    uint2 pay[kScatItems];      // (row, value bits); UNIT: .x = row only
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
    // keys of a tile (UNIT, later passes: the whole record -- its row comes along)
    auto load_keys = [&](uint32_t a0, uint32_t a1) {
#pragma unroll
        for (int i = 0; i < kScatItems; i++) {
            const uint32_t e = a0 + warp * (32 * kScatItems) + lane + i * 32;
            if (UNIT && !FIRST) {
                const uint2 r = e < a1 ? __ldcs(pay_in + e) : make_uint2(0xffffffffu, 0u);
                key[i] = r.x; pay[i].x = r.y;
            } else {
                key[i] = e < a1 ? __ldcs(keys_in + e) : 0xffffffffu;
            }
        }
    };
    if (tile < total) {
        seg = find_seg(sm.toff, n_segs, tile);
        tis = tile - sm.toff[seg]; nt = sm.toff[seg + 1] - sm.toff[seg];
        e0 = sm.off[seg] + tis * kSortTile;
        e1 = min(e0 + (uint32_t)kSortTile, sm.off[seg + 1]);
        if (FIRST) slice_v = load_slice(seg, tile, &cur_ra);
        load_keys(e0, e1);
    }
    for (; tile < total; tile += gridDim.x) {
        // warp-striped: warp w owns tile positions [w*256, (w+1)*256), round i, lane
        const uint32_t ew = e0 + warp * (32 * kScatItems) + lane;
        int nr = 0; bool covered = false;
        if (FIRST) {
            // the values are requested before the row search, so that their latency hides behind it
            if (!UNIT) {
#pragma unroll
                for (int i = 0; i < kScatItems; i++) {
                    const uint32_t e = ew + i * 32;
                    if (e < e1) pay[i].y = __float_as_uint(__ldcs(val_in + e));
                }
            }
            const int mb_rows = (int)min((int64_t)rl.mini_batch, rl.block_rows - (int64_t)seg * rl.mini_batch);
            // stage rp[ra ..] - base.  The first kScatThreads entries were prefetched into `slice_v`
            // with the keys and almost always cover the tile (one barrier); otherwise keep loading.
            nr = min(kScatThreads - 1, mb_rows - cur_ra);
            if ((int)threadIdx.x <= nr) sm.rows[threadIdx.x] = slice_v;
            __syncthreads();
            covered = sm.rows[nr] >= e1;       // uniform
            if (covered) {
                // Fast path: every row that starts inside the tile sets the bit of its start position; the row
                // of an element is then (tile's first row) + (bits at or below its position).  Two rows on one
                // position -- an empty row -- cannot be told apart that way: the tile takes the search path.
                if (threadIdx.x >= 1 && (int)threadIdx.x <= nr && slice_v > e0 && slice_v < e1) {
                    const uint32_t p = slice_v - e0, bit = 1u << (p & 31);
                    if (atomicOr(&sm.mark[p >> 5], bit) & bit) sm.dup = 1;
                }
            } else {
                const int64_t base = rl.rowptr[rl.block_row0];
                const int64_t *rp = rl.rowptr + rl.block_row0 + (int64_t)seg * rl.mini_batch;
                const int ra = cur_ra;
                nr = -1;
                for (int i0 = kScatThreads; i0 <= kRowSlice; i0 += kScatThreads) {
                    if (threadIdx.x == 0) sm.nr = 0x7fffffff;
                    __syncthreads();
                    const int i = i0 + threadIdx.x;
                    int hit = 0x7fffffff;
                    if (i <= kRowSlice && ra + i <= mb_rows) {
                        const uint32_t v = (uint32_t)(rp[ra + i] - base);
                        sm.rows[i] = v;
                        if (v >= e1) hit = i;
                    }
                    const int wmin = __reduce_min_sync(0xffffffffu, hit);
                    if (lane == 0 && wmin != 0x7fffffff) atomicMin(&sm.nr, wmin);
                    __syncthreads();
                    const int got = sm.nr;
                    __syncthreads();
                    if (got != 0x7fffffff) { nr = got; break; }
                }
            }
        } else if (!UNIT) {
            // payloads are requested now and consumed after the ranking
#pragma unroll
            for (int i = 0; i < kScatItems; i++) {
                const uint32_t e = ew + i * 32;
                if (e < e1) pay[i] = __ldcs(pay_in + e);
            }
        }
        for (int q = threadIdx.x; q < kScatWarps * R; q += kScatThreads) (&sm.wcnt[0][0])[q] = 0;
        __syncthreads();
        if (FIRST) {
            const int ra = cur_ra;
            if (covered && !sm.dup) {
                // rows that start before this warp's 256 positions, then one word of marks per round
                uint32_t before = 0;
#pragma unroll
                for (int j = 0; j < kSortTile / 32 / 32; j++) {
                    const int wd = lane + 32 * j;
                    if (wd < warp * kScatItems) before += __popc(sm.mark[wd]);
                }
                before = __reduce_add_sync(0xffffffffu, before);
                const uint32_t le_mask = 0xffffffffu >> (31 - lane);
#pragma unroll
                for (int i = 0; i < kScatItems; i++) {
                    const uint32_t m = sm.mark[warp * kScatItems + i];
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
                                int lo = 0, hi = nr;        // rows[lo] <= e < rows[hi]
                                while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (sm.rows[mid] <= e) lo = mid; else hi = mid; }
                                idx = lo; searched = true;
                            } else {
                                while (sm.rows[idx + 1] <= e) idx++;     // rows[nr] >= e1 > e bounds the walk
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
                else if (dmask && dense_hit(sm.dtab, key[i])) valid = false;
            }
            if (valid) vbits |= 1u << i;
            const uint32_t d = (key[i] >> shift) & (R - 1);
            const uint32_t active = __ballot_sync(0xffffffffu, valid);
            uint32_t peers = 0, prev = 0;
            if (FIRST) {
                // match.any costs time in proportion to the number of distinct values in the warp.  In the
                // first pass the elements are still in row order (32 consecutive elements = 32 different
                // hashed features): one ballot per digit bit -- lanes that agree with mine in every bit -- is
                // cheaper there.  Later passes see long runs of equal digits and keep match.any.
                peers = active;
#pragma unroll
                for (int b = 0; b < RB; b++) {
                    const uint32_t m = __ballot_sync(0xffffffffu, valid && ((d >> b) & 1u));
                    peers &= ((d >> b) & 1u) ? m : ~m;
                }
                if (valid) prev = sm.wcnt[warp][d];
            } else
            if (valid) { peers = __match_any_sync(active, d); prev = sm.wcnt[warp][d]; }
            __syncwarp();
            if (valid) {
                const uint32_t r = __popc(peers & lt_mask);
                if (r == 0) sm.wcnt[warp][d] = prev + __popc(peers);
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
                for (int w = 0; w < kScatWarps; w++) { uint32_t c = sm.wcnt[w][d]; sm.wcnt[w][d] = run; run += c; }
            }
            uint32_t inc = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
            if (lane == 31 && owner) sm.warp_tot[warp] = inc;
            __syncthreads();
            if (owner) {
                uint32_t wbase = 0;
#pragma unroll
                for (int w = 0; w < R / 32; w++) if (w < warp) wbase += sm.warp_tot[w];
                const uint32_t ex = wbase + inc - run;
                sm.dig_base[d] = ex;
                sm.gofs[d] = tile_hist[(size_t)sm.toff[seg] * R + (size_t)d * nt + tis] - ex;
                if (d == R - 1) sm.ntile = ex + run;
            }
        }
        __syncthreads();
        // place keys and payloads (or packed records) at their tile-local sorted position
#pragma unroll
        for (int i = 0; i < kScatItems; i++) {
            if ((vbits >> i) & 1u) {
                const uint32_t d = (key[i] >> shift) & (R - 1);
                const uint32_t pos = sm.dig_base[d] + sm.wcnt[warp][d] + rank[i];
                if (UNIT) s_pay[pos] = make_uint2(key[i], pay[i].x);
                else { s_key[pos] = key[i]; s_pay[pos] = pay[i]; }
            }
        }
        __syncthreads();
        if (FIRST) {      // marks are consumed: clear them for the next tile (its marks come after its slice barrier)
            for (int q = threadIdx.x; q < kSortTile / 32; q += kScatThreads) sm.mark[q] = 0;
            if (threadIdx.x == 0) sm.dup = 0;
        }
        const uint32_t n_tile = sm.ntile;
        // next tile's coordinates and keys (the key registers are free from here on)
        const uint32_t next = tile + gridDim.x;
        int nseg = 0; uint32_t ntis = 0, nnt = 0, ne0 = 0, ne1 = 0;
        if (next < total) {
            nseg = find_seg(sm.toff, n_segs, next);
            ntis = next - sm.toff[nseg]; nnt = sm.toff[nseg + 1] - sm.toff[nseg];
            ne0 = sm.off[nseg] + ntis * kSortTile;
            ne1 = min(ne0 + (uint32_t)kSortTile, sm.off[nseg + 1]);
            if (FIRST) slice_v = load_slice(nseg, next, &cur_ra);
            load_keys(ne0, ne1);
        }
        for (uint32_t p = threadIdx.x; p < n_tile; p += kScatThreads) {
            if (UNIT) {
                const uint2 r = s_pay[p];
                pay_out[sm.gofs[(r.x >> shift) & (R - 1)] + p] = r;
            } else {
                const uint32_t k = s_key[p];
                const uint32_t g = sm.gofs[(k >> shift) & (R - 1)] + p;
                keys_out[g] = k;
                pay_out[g] = s_pay[p];
            }
        }
        __syncthreads();
        seg = nseg; tis = ntis; nt = nnt; e0 = ne0; e1 = ne1;
    }
}

template <bool FIRST, int RB>
__global__ void __launch_bounds__(kScatThreads, kScatCtasPerSm)
sort_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint2 *__restrict__ pay_in,
                    RowLookup rl, const float *__restrict__ val_in,
                    uint32_t *__restrict__ keys_out, uint2 *__restrict__ pay_out,
                    const uint32_t *__restrict__ seg_off, const uint32_t *__restrict__ seg_tile_off,
                    int n_segs, int shift, const uint32_t *__restrict__ tile_hist,
                    uint32_t dims, const DenseInfo *__restrict__ dense, uint32_t *__restrict__ bad)
{
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    __shared__ ScatSmem<FIRST, RB> sm;
    uint32_t dmask = 0;
    if (FIRST) {
        dmask = dense->mask;
        dense_tab_build(sm.dtab, nullptr, dmask, dense->key[threadIdx.x & 31], kScatThreads);
        for (int q = threadIdx.x; q < kSortTile / 32; q += kScatThreads) sm.mark[q] = 0;
        if (threadIdx.x == 0) sm.dup = 0;
    }
    for (int s = threadIdx.x; s <= n_segs; s += kScatThreads) { sm.off[s] = seg_off[s]; sm.toff[s] = seg_tile_off[s]; }
    const bool unit = dense->unit != 0u;     // uniform over the grid
    __syncthreads();
    if (unit) scatter_body<FIRST, RB, true>(sm, dyn_smem, keys_in, pay_in, rl, val_in, keys_out, pay_out, n_segs, shift,
                                            tile_hist, dims, dmask, bad);
    else scatter_body<FIRST, RB, false>(sm, dyn_smem, keys_in, pay_in, rl, val_in, keys_out, pay_out, n_segs, shift,
                                        tile_hist, dims, dmask, bad);
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
        HM_CUDA(cudaMalloc(&s.dense, sizeof(DenseInfo)));
        HM_CUDA(cudaMemset(s.dense, 0, sizeof(DenseInfo)));
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
    cudaFree(s.in_off); cudaFree(s.in_tile_off); cudaFree(s.dense);
    s.tile_hist = s.row_sum = s.seg_off = s.seg_tile_off = s.in_off = s.in_tile_off = nullptr; s.tile_row0 = nullptr;
    s.dense = nullptr;
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

hm_status launch_dense_check(Launcher &L, SortScratch &s, const int64_t *d_rowptr, int64_t row0, int64_t n_rows,
                             const uint32_t *d_feature, const float *d_value, int64_t n_elems, int64_t dims, bool enable)
{
    if (!enable || n_rows < 1) {
        HM_CUDA(cudaMemsetAsync(s.dense, 0, sizeof(DenseInfo), L.stream));
        return HM_OK;
    }
    const int grid = persistent_grid((n_rows + kChkThreads / 8 - 1) / (kChkThreads / 8), 8);
    HM_LAUNCH(L, K_DENSE_CHECK, 8.0 * (double)n_elems + 8.0 * (double)n_rows,
              (dense_check_kernel<<<grid, kChkThreads, 0, L.stream>>>(d_rowptr, row0, n_rows, d_feature, d_value,
                                                                     (uint32_t)std::min<int64_t>(dims, 0xffffffffLL), s.dense)));
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
                                int64_t dims, uint32_t *d_bad, int *out_buf, int ctas_per_sm)
{
    // > 48 KB of dynamic shared memory needs opting in, once per device (the attribute is per
    // device; the flags are atomics because learners on different threads may get here together)
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    HM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        HM_TRY(opt_in_smem(sort_scatter_kernel<true, 8>, kScatterSmem));
        HM_TRY(opt_in_smem(sort_scatter_kernel<false, 8>, kScatterSmem));
        HM_TRY(opt_in_smem(sort_scatter_kernel<true, 9>, kScatterSmem));
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
    HM_LAUNCH(L, K_SORT_HIST, 4.0 * (double)n_elems,                                                                     \
              (sort_hist_kernel<RB_, FIRST_><<<hist_grid, kSortThreads, 0, L.stream>>>(kin, pin, off, toff, n_segs, shift, \
                                                                                       s.tile_hist, s.row_sum, dims32, s.dense)))
        if (p == 0) { if (rb == 9) HM_HIST(9, true); else HM_HIST(8, true); }
        else { if (rb == 9) HM_HIST(9, false); else HM_HIST(8, false); }
#undef HM_HIST
        // (the first scan also writes the sorted arrays' mini-batch and tile offsets)
        HM_LAUNCH(L, K_SORT_SCAN, 0.0, (scan_rows_kernel<<<1, 1024, 0, L.stream>>>(
                                           s.row_sum, rows, p == 0 ? s.seg_off : nullptr, s.seg_tile_off, n_segs, 1 << rb)));
        HM_LAUNCH(L, K_SORT_SCAN, 0.0,
                  (scan_tiles_kernel<<<(rows + kSortWarps - 1) / kSortWarps, kSortThreads, 0, L.stream>>>(
                      s.tile_hist, s.row_sum, toff, n_segs, 1 << rb)));
        const double bytes = p == 0 ? 20.0 * (double)n_elems + 8.0 * (double)block_rows : 24.0 * (double)n_elems;
#define HM_SCATTER(FIRST_, RB_)                                                                                        \
    HM_LAUNCH(L, K_SORT_SCATTER, bytes,                                                                                 \
              (sort_scatter_kernel<FIRST_, RB_><<<grid, kScatThreads, kScatterSmem, L.stream>>>(                        \
                  kin, pin, rl, p == 0 ? d_value_block : nullptr, kout, pout, off, toff, n_segs, shift,                 \
                  s.tile_hist, dims32, s.dense, d_bad)))
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
