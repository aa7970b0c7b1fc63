// hm_update.cuh -- the reduce-and-step half of a mini-batch (linear_update_kernel and what it is made of).
//
// Included by hm_linear.cu THREE times, inside namespaces that differ in kUpdItems only (elements per thread
// of a tile): the mini-batches that fill the GPU use 2048-element tiles, mid-size mini-batches (a few hundred
// thousand elements: fewer tiles than SMs) 1024-element tiles, those of up to 100 K elements 512 -- every phase of a tile (load, two scans, step)
// is a latency chain whose length grows with the elements per thread, and with few tiles nothing else hides
// it.  Measured (examples/s, 8 / 4 elements per thread): 2048 rows per mini-batch 75 / 97 M, 4096 rows
// 151 / 173 M, 8192 rows 261 / 295 M, 16,384 rows 395 / 372 M; 4 / 2 elements per thread: 512 rows 29 / 32.5 M,
// 2048 rows 96 / 108 M, 4096 rows 173 / 170 M, 8192 rows 300 / 278 M (profiles/r02e_update_tile_midsize.txt).
// A fourth inclusion (kUpdCov = true) serves the learners that keep a variance per feature: its sums can run
// over dloss * value^2 (UpdParams::square); the other three compile that choice out.
// Needs kUpdItems (constexpr int) and kUpdCov (constexpr bool) defined by the including namespace; everything else it uses (UpdParams,
// step_rec, load_rec, ...) is defined before the inclusion.  Not a stand-alone header.

constexpr int kUpdTile = kUpdThreads * kUpdItems;  // 2048 / 1024 / 512

__device__ __forceinline__ int pad_idx(int q) { return q + (q >> 5); }

// Segmented-scan element: f = "a segment head lies in the span", (v, c) = sum / count of the
// span's trailing run, k = number of segments closed in the span (plain sum).
struct Scan { double v; uint32_t c; uint32_t f; uint32_t k; };
__device__ __forceinline__ Scan combine(const Scan &a, const Scan &b)
{
    Scan r;
    r.f = a.f | b.f;
    r.v = b.f ? b.v : a.v + b.v;
    r.c = b.f ? b.c : a.c + b.c;
    r.k = a.k + b.k;
    return r;
}

constexpr int kUpdPadded = kUpdTile + kUpdTile / 32 + 2;
constexpr int kUpdList = kUpdTile + 8;
// independent record updates in flight per thread.  (The kernel sits at the 64-register limit of four CTAs
// per SM: 4 in flight spilled 12 bytes and measured 48.8 us per mini-batch, 2 with the segment bounds
// passed as parameters spills nothing: 44.6 us.)
#ifndef HM_UPD_BATCH
#define HM_UPD_BATCH 2
#endif
constexpr int kUpdBatch = HM_UPD_BATCH;
// shared memory is used twice: tile + closed-segment list while streaming, record buffers in the merge
constexpr size_t kUpdSmemTile = (size_t)kUpdPadded * 8 + (size_t)kUpdList * 8;
constexpr size_t kUpdSmemMerge = (size_t)kUpdMaxGrid * 24;   // staged `first` records + output list
constexpr size_t kUpdSmem = kUpdSmemTile > kUpdSmemMerge ? kUpdSmemTile : kUpdSmemMerge;

// Apply the optimizer to list entries [0, n): kUpdBatch independent read-modify-writes per thread
// are issued together so that their DRAM latencies overlap.  Entries are distinct features;
// 0xffffffff marks a slot to skip.
template <bool CG = false>
__device__ __forceinline__ void apply_list(const UpdParams &P, const UpdDyn &D, const uint32_t *l_key, const float *l_g, int n, int tid)
{
    for (int i0 = 0; i0 < n; i0 += kUpdThreads * kUpdBatch) {
        uint32_t j[kUpdBatch]; float g[kUpdBatch]; Rec r[kUpdBatch]; bool ok[kUpdBatch];
#pragma unroll
        for (int u = 0; u < kUpdBatch; u++) {
            const int i = i0 + u * kUpdThreads + tid;
            ok[u] = i < n;
            j[u] = ok[u] ? l_key[i] : 0u;
            g[u] = ok[u] ? l_g[i] : 0.f;
            ok[u] = ok[u] && j[u] < P.dims;
        }
#pragma unroll
        for (int u = 0; u < kUpdBatch; u++) if (ok[u]) r[u] = load_rec<CG>(P, j[u]);
#pragma unroll
        for (int u = 0; u < kUpdBatch; u++) if (ok[u]) step_rec(P, D, r[u], g[u]);
#pragma unroll
        for (int u = 0; u < kUpdBatch; u++) if (ok[u]) store_rec(P, j[u], r[u]);
    }
}

__device__ __forceinline__ uint4 rec_bits(const SegRec &r)
{
    return make_uint4((uint32_t)__double2loint(r.sum), (uint32_t)__double2hiint(r.sum), r.key, r.cnt);
}

// Shared-memory working set of one CTA of the update kernels.
struct UpdShared {
    uint32_t *s_key; float *s_val;       // the tile: features and contributions (padded)
    uint32_t *l_key; float *l_g;         // finished features of the tile, in feature order
    Scan *s_warp;
    SegRec *s_carry;                     // open trailing segment of the previous tile
    SegRec *s_first;                     // the CTA's piece of a segment that began before its range
    uint32_t *s_carry_open;
    const uint32_t *s_edge;              // does the CTA's range start / end on a true segment boundary?
    int *s_nlist;
};

// Reduce-by-feature of one tile [base, base + n_tile) of the sorted mini-batch: contributions, fp64
// segmented scan in a fixed order, finished features compacted into S.l_key / S.l_g (*S.s_nlist
// entries, to be stepped by apply_list), open segments left in S.s_carry / S.s_first.  Ends on a
// barrier.  RO: dloss was written by an earlier kernel and may use the read-only path.
template <int LD>
__device__ __forceinline__ void reduce_tile(const UpdParams &P, const UpdDyn &D, const UpdShared &S, uint32_t base, int n_tile,
                                            bool first_tile, bool last_tile)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- coalesced striped load; contribution = dloss[row] * value (fp32 product)
    uint32_t k[kUpdItems]; uint2 pv[kUpdItems];
#pragma unroll
    for (int i = 0; i < kUpdItems; i++) {
        const int q = i * kUpdThreads + tid;
        if (q < n_tile) { k[i] = __ldcs(P.skey + base + q); pv[i] = __ldcs(P.spay + base + q); }
    }
    if (LD == kLdCG) pdl_wait();     // the sorted tile is on its way; dloss is the forward kernel's output
    USTAMP(7);
    SEQSTAMP(D.seg, 4);
#pragma unroll
    for (int i = 0; i < kUpdItems; i++) {
        const int q = i * kUpdThreads + tid;
        if (q < n_tile) {
            S.s_key[pad_idx(q)] = k[i];
            const float x = __uint_as_float(pv[i].y);
            // (kUpdCov: only the instance for the variance-keeping learners carries the second kind of sum)
            S.s_val[pad_idx(q)] = __fmul_rn(ld_mode<LD>(P.dloss + pv[i].x), (kUpdCov && P.square) ? __fmul_rn(x, x) : x);
        }
    }
    __syncthreads();
    USTAMP(1);
    const bool left_closed = S.s_edge[0] != 0, right_closed = S.s_edge[1] != 0;
    const SegRec carry = *S.s_carry;
    const uint32_t carry_open = *S.s_carry_open;
            const uint32_t key0 = S.s_key[pad_idx(0)];
    const bool edge_continues = carry.cnt != 0 && carry.key == key0;
    // does a segment touching this tile's left edge reach past the CTA's left edge?
    const bool edge_open = first_tile ? !left_closed : (edge_continues && carry_open != 0);
    // the carried segment ended exactly at the tile edge and is complete: thread 0 lists it
    const bool retire_carry = carry.cnt != 0 && !edge_continues && !carry_open;
    // the tile's last item closes its segment when the CTA's range ends on a true boundary
    const bool close_tail = last_tile && right_closed;
    __syncthreads();  // everyone has read the carry before it is overwritten below

    // ---- a tile that holds ONE feature (a hot feature's run spans dozens of tiles: a third of the tiles of a
    // Criteo-shaped mini-batch) needs no segmented scan: a plain fixed-order reduction, and one piece.
    if (first_tile && last_tile && key0 == S.s_key[pad_idx(n_tile - 1)]) {
        double part = 0.0;
#pragma unroll
        for (int i = 0; i < kUpdItems; i++) {
            const int q = i * kUpdThreads + tid;
            if (q < n_tile) part += (double)S.s_val[pad_idx(q)];
        }
        part = warp_sum(part);
        if (lane == 0) S.s_warp[warp].v = part;
        __syncthreads();
        if (tid == 0) {
            double sum = 0.0;
#pragma unroll
            for (int w2 = 0; w2 < kUpdThreads / 32; w2++) sum += S.s_warp[w2].v;
            SegRec r; r.sum = sum; r.key = key0; r.cnt = (uint32_t)n_tile;
            *S.s_nlist = 0;
            S.s_carry->cnt = 0; *S.s_carry_open = 0;
            if (!left_closed && !right_closed) { *S.s_carry = r; *S.s_carry_open = 1; }   // piece of a longer run
            else if (!left_closed) *S.s_first = r;                                          // the run's last piece
            else if (!right_closed) *S.s_carry = r;                                         // the run's first piece
            else { S.l_key[0] = key0; S.l_g[0] = grad_of(P, D, sum, r.cnt); *S.s_nlist = 1; }  // the whole run
        }
        __syncthreads();
        return;
    }

    // ---- pass 1: thread summaries (items [8*tid, 8*tid+8)).  A "head" is an item at tile
    // position >= 1 whose feature differs from its predecessor; position 0 is never a head:
    // it either continues the carried segment or starts afresh with an empty running sum.
    // An item "closes" its segment when the next item has another feature.
    const int q0 = tid * kUpdItems;
    Scan mine; mine.v = 0.0; mine.c = 0; mine.f = 0; mine.k = (tid == 0 && retire_carry) ? 1u : 0u;
    if (tid == 0 && edge_continues) { mine.v = carry.sum; mine.c = carry.cnt; }
    {
        uint32_t kk[kUpdItems + 1];
        const uint32_t prev_key = (q0 > 0 && q0 < n_tile) ? S.s_key[pad_idx(q0 - 1)] : 0u;
#pragma unroll
        for (int i = 0; i <= kUpdItems; i++) kk[i] = q0 + i < n_tile ? S.s_key[pad_idx(q0 + i)] : 0u;
#pragma unroll
        for (int i = 0; i < kUpdItems; i++) {
            const int q = q0 + i;
            if (q < n_tile) {
                const bool head = q > 0 && kk[i] != (i == 0 ? prev_key : kk[i - 1]);
                if (head) { mine.v = 0.0; mine.c = 0; mine.f = 1; }
                mine.v += (double)S.s_val[pad_idx(q)]; mine.c += 1;
                const bool closes = q < n_tile - 1 ? kk[i + 1] != kk[i] : close_tail;
                if (closes) mine.k += 1;
            }
        }
    }
    // ---- exclusive segmented scan of the summaries over the CTA
    Scan inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        Scan up;
        up.v = __shfl_up_sync(0xffffffffu, inc.v, o);
        up.c = __shfl_up_sync(0xffffffffu, inc.c, o);
        up.f = __shfl_up_sync(0xffffffffu, inc.f, o);
        up.k = __shfl_up_sync(0xffffffffu, inc.k, o);
        if (lane >= o) inc = combine(up, inc);
    }
    if (lane == 31) S.s_warp[warp] = inc;
    Scan ex;
    ex.v = __shfl_up_sync(0xffffffffu, inc.v, 1);
    ex.c = __shfl_up_sync(0xffffffffu, inc.c, 1);
    ex.f = __shfl_up_sync(0xffffffffu, inc.f, 1);
    ex.k = __shfl_up_sync(0xffffffffu, inc.k, 1);
    if (lane == 0) { ex.v = 0.0; ex.c = 0; ex.f = 0; ex.k = 0; }
    __syncthreads();
    USTAMP(2);
    Scan wp; wp.v = 0.0; wp.c = 0; wp.f = 0; wp.k = 0;
    for (int w2 = 0; w2 < warp; w2++) wp = combine(wp, S.s_warp[w2]);
    const Scan cin = combine(wp, ex);
    if (tid == kUpdThreads - 1) *S.s_nlist = (int)(cin.k + mine.k);

    // ---- pass 2: walk again; every closed segment goes to the list, in feature order
    double run = cin.v; uint32_t cnt = cin.c; bool seen_head = cin.f != 0;
    int slot = (int)cin.k;
    if (tid == 0 && edge_continues) { run = carry.sum; cnt = carry.cnt; }
    if (tid == 0 && carry.cnt != 0 && !edge_continues) {
        if (carry_open) *S.s_first = carry;
        else { S.l_key[slot] = carry.key; S.l_g[slot] = grad_of(P, D, carry.sum, carry.cnt); slot++; }
    }
    // (the tile is read again from shared memory: keeping sixteen values in registers across the
    // scan costs a fourth resident CTA per SM)
    uint32_t kprev = (q0 > 0 && q0 < n_tile) ? S.s_key[pad_idx(q0 - 1)] : 0u;
    uint32_t kcur = q0 < n_tile ? S.s_key[pad_idx(q0)] : 0u;
#pragma unroll
    for (int i = 0; i < kUpdItems; i++) {
        const int q = q0 + i;
        if (q < n_tile) {
            const uint32_t knext = q + 1 < n_tile ? S.s_key[pad_idx(q + 1)] : 0u;
            const bool head = q > 0 && kcur != kprev;
            if (head) { run = 0.0; cnt = 0; seen_head = true; }
            run += (double)S.s_val[pad_idx(q)]; cnt += 1;
            // open: the segment reaches the tile's left edge and, through it, past the CTA's left edge
            const bool open = !seen_head && edge_open;
            const bool closes = q < n_tile - 1 ? knext != kcur : close_tail;
            if (closes) {
                if (open) {     // at most one such segment per CTA
                    S.s_first->sum = run; S.s_first->key = kcur; S.s_first->cnt = cnt;
                    S.l_key[slot] = 0xffffffffu; S.l_g[slot] = 0.f;   // placeholder: skipped by apply_list
                } else {
                    S.l_key[slot] = kcur; S.l_g[slot] = grad_of(P, D, run, cnt);
                }
                slot++;
                if (q == n_tile - 1) { S.s_carry->cnt = 0; *S.s_carry_open = 0; }
            } else if (q == n_tile - 1) {
                S.s_carry->sum = run; S.s_carry->key = kcur; S.s_carry->cnt = cnt;
                *S.s_carry_open = open ? 1u : 0u;
            }
            kprev = kcur; kcur = knext;
        }
    }
    __syncthreads();
}

// Optimizer step of the block's hot features: one warp per feature adds the forward CTAs' partial sums
// in a fixed order.  Runs in kHotCtas extra CTAs of the update kernel; the features are disjoint from
// everything the sorted stream holds (the sort dropped them by key), so nothing else touches their
// records.  A hot feature that does not occur in this mini-batch is left alone, like any other.
constexpr int kHotCtas = kMaxHot / (kUpdThreads / 32);

__device__ __noinline__ void hot_features_step(const UpdParams &P, const UpdDyn &D, int part)
{
    const int lane = threadIdx.x & 31;
    const int i = part * (kUpdThreads / 32) + (threadIdx.x >> 5);
    if (i >= (int)P.hot->count) return;
    const double *col = P.hot_sum + (size_t)i * P.hot_stride;
    const uint32_t *ccol = P.hot_cnt + (size_t)i * P.hot_stride;
    double sum = 0.0; uint32_t cnt = 0;
    for (int c = lane; c < D.fwd_grid; c += 32) { sum += __ldcg(col + c); cnt += __ldcg(ccol + c); }
    sum = warp_sum(sum);
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0 && cnt != 0u) {
        const uint32_t j = P.hot->key[i];
        Rec r = load_rec<true>(P, j);
        step_rec(P, D, r, grad_of(P, D, sum, cnt));
        store_rec(P, j, r);
    }
}

// Shared memory of one CTA of the update kernels.
struct UpdCtaSmem {
    __align__(16) unsigned char raw[kUpdSmem];
    Scan warp[kUpdThreads / 32];
    SegRec carry;            // open trailing segment of the previous tile
    SegRec first;            // the CTA's piece of a segment that began before its range
    uint32_t carry_open;
    uint32_t is_last;
    uint32_t edge[2];        // does the CTA's range start / end on a true segment boundary?
    int nlist;
};

// Reduce-by-feature and optimizer step of mini-batch P.seg of a sorted block, as CTA `cta` of P.n_active.
// RO: nothing the running kernel writes is read here through the read-only path.
template <int LD>
__device__ __forceinline__ void update_segment(const UpdParams &P, const UpdDyn &D, UpdCtaSmem &sm, const int cta)
{
    unsigned char *smem_raw = sm.raw;
    Scan *s_warp = sm.warp;
    SegRec &s_carry = sm.carry;
    SegRec &s_first = sm.first;
    uint32_t &s_carry_open = sm.carry_open;
    uint32_t &s_is_last = sm.is_last;
    uint32_t *s_edge = sm.edge;
    int &s_nlist = sm.nlist;
    uint32_t *s_key = reinterpret_cast<uint32_t *>(smem_raw);
    float *s_val = reinterpret_cast<float *>(smem_raw + (size_t)kUpdPadded * 4);
    uint32_t *l_key = reinterpret_cast<uint32_t *>(smem_raw + (size_t)kUpdPadded * 8);
    float *l_g = reinterpret_cast<float *>(smem_raw + (size_t)kUpdPadded * 8 + (size_t)kUpdList * 4);
    const int tid = threadIdx.x;
    const uint32_t s0 = P.seg_off[D.seg], n = P.seg_off[D.seg + 1] - s0;
    // contiguous split of the sorted mini-batch over the grid, in whole tiles
    const uint32_t tiles = (n + kUpdTile - 1) / kUpdTile;
    const uint32_t tiles_per_cta = (tiles + D.n_active - 1) / D.n_active;
    const uint32_t lo = min(n, (uint32_t)cta * tiles_per_cta * kUpdTile);
    const uint32_t hi = min(n, lo + tiles_per_cta * kUpdTile);

    if (tid == 0) { s_carry.cnt = 0; s_carry_open = 0; s_first.cnt = 0; s_is_last = 0; }
    // Peek at the neighbours of the range: a first segment that really starts at `lo`, or a last
    // segment that really ends at `hi`, is complete inside this CTA and never reaches the merge.
    // The four loads are issued here and only consumed after the first tile's barrier, so they
    // overlap the tile loads instead of adding a round trip.
    if (tid == 32 && lo < hi) {
        const uint32_t a0 = lo == 0 ? 0u : __ldg(P.skey + s0 + lo - 1), a1 = __ldg(P.skey + s0 + lo);
        const uint32_t b0 = __ldg(P.skey + s0 + hi - 1), b1 = hi == n ? 0u : __ldg(P.skey + s0 + hi);
        s_edge[0] = lo == 0 || a0 != a1;
        s_edge[1] = hi == n || b0 != b1;
    }
    if (lo >= hi) __syncthreads();   // (the tile loop below has its own barriers)

    UpdShared S;
    S.s_key = s_key; S.s_val = s_val; S.l_key = l_key; S.l_g = l_g; S.s_warp = s_warp; S.s_carry = &sm.carry;
    S.s_first = &sm.first; S.s_carry_open = &sm.carry_open; S.s_edge = s_edge; S.s_nlist = &sm.nlist;
    for (uint32_t t0 = lo; t0 < hi; t0 += kUpdTile) {
        const int n_tile = (int)min((uint32_t)kUpdTile, hi - t0);
        const bool last_tile = t0 + kUpdTile >= hi;
        reduce_tile<LD>(P, D, S, s0 + t0, n_tile, t0 == lo, last_tile);
        USTAMP(3);
        // ---- after the last tile the CTA's boundary pieces are final: publish them and take the
        // completion ticket now, so that both round trips overlap the optimizer step below
        if (last_tile && tid == 0) {
            const uint4 none = make_uint4(0u, 0u, 0u, 0u);
            const bool whole = s_carry.cnt != 0 && s_carry_open;     // the range lies inside one segment
            P.br.first[cta] = whole ? rec_bits(s_carry) : (s_first.cnt != 0 ? rec_bits(s_first) : none);
            P.br.last[cta] = (!whole && s_carry.cnt != 0) ? rec_bits(s_carry) : none;
            __threadfence();
            s_is_last = atomicAdd(P.br.ticket, 1u) == (uint32_t)D.n_active - 1;
        }
        // ---- optimizer step for the listed features, latencies overlapped
        USTAMP(4);
        apply_list<LD == kLdCG>(P, D, l_key, l_g, s_nlist, tid);
        __syncthreads();
        USTAMP(5);
        SEQSTAMP(D.seg, 5);
    }
    if (lo >= hi) {   // a CTA without elements leaves empty records and still takes its ticket
        if (LD == kLdCG) pdl_wait();   // (the records and the ticket belong to the previous update until it has finished)
        if (tid == 0) {
            P.br.first[cta] = make_uint4(0u, 0u, 0u, 0u);
            P.br.last[cta] = make_uint4(0u, 0u, 0u, 0u);
            __threadfence();
            s_is_last = atomicAdd(P.br.ticket, 1u) == (uint32_t)D.n_active - 1;
        }
        __syncthreads();
    }
    if (!s_is_last) return;

    // ---- the last CTA to finish resolves the straddling segments.  `first` records are staged
    // in shared memory; the thread that owns CTA c starts a run at last[c] and walks first[c+1],
    // first[c+2], ... while the feature matches (fixed order, fp64).  No sorting is needed.
    __threadfence();
    const int G = D.n_active;
    uint4 *f_sm = reinterpret_cast<uint4 *>(smem_raw);
    uint32_t *ml_key = reinterpret_cast<uint32_t *>(smem_raw + (size_t)kUpdMaxGrid * 16);
    float *ml_g = reinterpret_cast<float *>(ml_key + kUpdMaxGrid);
    constexpr int kPer = (kUpdMaxGrid + kUpdThreads - 1) / kUpdThreads;
    uint4 mylast[kPer];
    if (tid == 0) s_nlist = 0;
#pragma unroll
    for (int u = 0; u < kPer; u++) {
        const int c = u * kUpdThreads + tid;
        if (c < G) { f_sm[c] = __ldcg(P.br.first + c); mylast[u] = __ldcg(P.br.last + c); }
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < kPer; u++) {
        const int c = u * kUpdThreads + tid;
        if (c < G && mylast[u].w != 0u) {
            const uint32_t key = mylast[u].z;
            double sum = __hiloint2double((int)mylast[u].y, (int)mylast[u].x);
            uint32_t cnt = mylast[u].w;
            for (int j = c + 1; j < G; j++) {
                const uint4 f = f_sm[j];
                if (f.w == 0u || f.z != key) break;
                sum += __hiloint2double((int)f.y, (int)f.x); cnt += f.w;
            }
            const int sl = atomicAdd(&s_nlist, 1);     // distinct features: list order is irrelevant
            ml_key[sl] = key; ml_g[sl] = grad_of(P, D, sum, cnt);
        }
    }
    __syncthreads();
    apply_list<LD == kLdCG>(P, D, ml_key, ml_g, s_nlist, tid);
    USTAMP(6);
    SEQSTAMP_ANY(D.seg, 6);
    if (tid == 0) *P.br.ticket = 0;   // ready for the next launch on this stream
}

template <bool PDL>
__global__ void __launch_bounds__(kUpdThreads, kUpdCtasPerSm) linear_update_kernel(const __grid_constant__ UpdParams P)
{
    __shared__ UpdCtaSmem sm;
    USTAMP(0);
    SEQSTAMP(P.dyn.seg, 3);
    if (PDL) pdl_launch_dependents();
    if ((int)blockIdx.x >= P.dyn.n_active) {
        if (PDL) pdl_wait();
        hot_features_step(P, P.dyn, (int)blockIdx.x - P.dyn.n_active);
        return;
    }
    update_segment<PDL ? kLdCG : kLdRO>(P, P.dyn, sm, (int)blockIdx.x);
}
