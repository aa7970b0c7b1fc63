// hm_fm.cu -- kernels of the factorization-machine learner (train_fm).
//
// Stands in for FactorizationMachineUDTF, which BASELINE.json's north_star names ("train_fm's
// FactorizationMachineUDTF", "FM O(k*nnz) pairwise term", model rows carrying Vif).  The
// reference mount has no source (README.md:20-22); the model is Rendle's published FM and the
// arithmetic is DESIGN.md 3.8 (fp32 products, fp64 sums), restated independently in
// oracle/hm_oracle.c.
//
//   fm_forward_kernel  warp per row, lanes over factors: S_f = sum_j v_jf x_j accumulated in
//                      ROW ORDER in fp64 (bit-identical to a sequential loop), y_hat, dloss,
//                      and the table S32[row][f] the update needs
//   fm_update_kernel   CTA per 1024-element chunk of the feature-sorted mini-batch, a segmented scan
//                      like the linear learner's: the segment structure is worked out once, then the
//                      per-feature fp64 sums A1, A2, GS_f are formed four factors per pass (16-byte
//                      loads of S rows and factor records) and finished features are stepped from a
//                      shared-memory list, two entries per thread at a time
//   fm_merge_kernel    features whose run straddles chunks (hot features): pieces sit in
//                      consecutive chunks, one warp walks them in order; one extra CTA sums dloss
//                      for the w0 step
//
// Replica layout: w0rec = (w0, acc); wrec[j] = (w_j[, acc]); vrec[j] = (V_j0..V_j,k-1[, acc_0..acc_k-1]).
// With k = 16 and AdaGrad a feature's factor record is exactly one 128-byte line.
#include "hm_internal.h"
#include <algorithm>
#include <atomic>

namespace hm {

namespace {

constexpr int kFmChunk = 1024;      // sorted elements per warp
constexpr int kFmWarps = 4;         // warps per CTA in update / merge

__device__ __forceinline__ uint32_t fmix32(uint32_t h)
{
    h ^= h >> 16; h *= 0x85ebca6bu;
    h ^= h >> 13; h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}

__device__ __forceinline__ float hm_exp(float x)   // DESIGN.md 3.2, same sequence as hm_linear.cu
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

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void fm_init_v_kernel(float *__restrict__ vrec, int64_t dims, int k, int vw, float sigma, uint32_t seed)
{
    const int64_t n = dims * k;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t h = fmix32((uint32_t)q * 0x9E3779B1u + seed);
        const float u = __fmul_rn(__fsub_rn(__fmul_rn((float)(h >> 8), 5.9604644775390625e-08f), 0.5f), 2.0f);
        vrec[(q / k) * vw + (q % k)] = __fmul_rn(sigma, u);
    }
}

// ---- forward ---------------------------------------------------------------------

struct FmFwdParams {
    const float *w0rec, *wrec, *vrec;
    int32_t ww, vw, k;
    const int64_t *rowptr;
    const uint32_t *feature;
    const float *value;
    const float *label;
    float *margin;       // y_hat, indexed by row of the block (prediction: row of the call)
    float *dloss;        // indexed by row of the mini-batch
    float *stab;         // S32[row of the mini-batch][k]
    int64_t block_row0, row0;
    int32_t n_rows, classification, train;
    uint32_t dims;
    uint32_t *bad;
};

// LPR = lanes per row: 32 (a warp per row, NF factors per lane) or 16 (k <= 16: two rows per warp,
// so that no lane idles).
template <int NF, int LPR>
__global__ void __launch_bounds__(256) fm_forward_kernel(FmFwdParams P)
{
    constexpr int RPW = 32 / LPR;                 // rows per warp
    const int lane = threadIdx.x & 31;
    const int sub = lane & (LPR - 1), half = lane / LPR;
    const int k = P.k;
    const float w0 = P.w0rec[0];
    for (int rb = (blockIdx.x * 8 + (threadIdx.x >> 5)) * RPW; rb < P.n_rows; rb += gridDim.x * 8 * RPW) {
        const int r = rb + half;
        const bool live = r < P.n_rows;
        const int64_t row = P.row0 + (live ? r : rb);
        const int64_t p0 = live ? P.rowptr[row] : 0, p1 = live ? P.rowptr[row + 1] : 0;
        // linear term: lanes of the row over its elements, fp64 tree
        double lin = 0.0;
        for (int64_t p = p0 + sub; p < p1; p += LPR) {
            const uint32_t f = P.feature[p];
            if (f < P.dims) lin += (double)__fmul_rn(__ldg(P.wrec + (size_t)f * P.ww), P.value[p]);
            else atomicOr(P.bad, 1u);
        }
        // factor sums: lanes over f, elements in row order (4 factor-row loads in flight)
        double S[NF], Q[NF];
#pragma unroll
        for (int q = 0; q < NF; q++) { S[q] = 0.0; Q[q] = 0.0; }
        for (int64_t p = p0; p < p1; p += 4) {
            float v[4][NF], x[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool ok = p + u < p1;
                const uint32_t f = ok ? P.feature[p + u] : 0u;
                x[u] = ok ? P.value[p + u] : 0.0f;
                const bool in = ok && f < P.dims;
#pragma unroll
                for (int q = 0; q < NF; q++) {
                    const int fi = sub + LPR * q;
                    v[u][q] = (in && fi < k) ? __ldg(P.vrec + (size_t)f * P.vw + fi) : 0.0f;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (p + u < p1) {
#pragma unroll
                    for (int q = 0; q < NF; q++) {
                        const float px = __fmul_rn(v[u][q], x[u]);
                        S[q] += (double)px;
                        Q[q] += (double)__fmul_rn(px, px);
                    }
                }
        }
        double inter = 0.0;
#pragma unroll
        for (int q = 0; q < NF; q++) {
            const int fi = sub + LPR * q;
            const float s32 = (float)S[q];
            if (live && fi < k) {
                inter += (double)__fmul_rn(s32, s32) - Q[q];
                if (P.train) P.stab[(size_t)r * k + fi] = s32;
            }
        }
        // sums over the lanes of the row (the warp has reconverged here)
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) {
            lin += __shfl_xor_sync(0xffffffffu, lin, o);
            inter += __shfl_xor_sync(0xffffffffu, inter, o);
        }
        const float yhat = (float)((double)w0 + lin + 0.5 * inter);
        if (live && sub == 0) {
            P.margin[row - P.block_row0] = yhat;
            if (P.train) {
                const float lab = P.label[row];
                float d;
                if (P.classification) {
                    const float y = lab > 0.0f ? 1.0f : -1.0f;
                    d = __fdiv_rn(-y, __fadd_rn(1.0f, hm_exp(__fmul_rn(y, yhat))));
                } else {
                    d = __fsub_rn(yhat, lab);
                }
                P.dloss[r] = d;
            }
        }
    }
}

// ---- update ------------------------------------------------------------------------

struct FmPieceHdr { double a1, a2; uint32_t key, cnt; uint32_t pad[2]; };   // 32 bytes
static_assert(sizeof(FmPieceHdr) == 32, "piece headers are 32 bytes");

struct FmUpdParams {
    float *w0rec, *wrec, *vrec;
    uint8_t *touched;
    int32_t ww, vw, k;
    const uint32_t *skey;
    const uint2 *spay;
    const float *dloss;
    const float *stab;
    const uint32_t *seg_off;
    int32_t seg;
    uint32_t dims;
    int32_t adagrad, grad_norm;
    float eta, eps, lambda_w0, lambda_w, lambda_v;
    double inv_rows;
    int32_t n_rows;
    // pieces of features whose run straddles chunks: one `first` and one `last` per chunk
    FmPieceHdr *first_hdr, *last_hdr;
    double *first_gs, *last_gs;      // [chunk][k]
    int32_t n_chunks;
};

__device__ __forceinline__ float fm_grad(const FmUpdParams &P, double sum, uint32_t cnt)
{
    if (P.grad_norm == HM_NORM_BATCH) return (float)(sum * P.inv_rows);
    if (P.grad_norm == HM_NORM_COUNT) return (float)(sum / (double)cnt);
    return (float)sum;
}

// SGD or AdaGrad on one parameter with its own L2 strength.
__device__ __forceinline__ void fm_step(const FmUpdParams &P, float g, float lam, float &th, float &acc)
{
    g = __fadd_rn(g, __fmul_rn(lam, th));
    if (P.adagrad) {
        acc = __fadd_rn(acc, __fmul_rn(g, g));
        th = __fsub_rn(th, __fdiv_rn(__fmul_rn(P.eta, g), __fadd_rn(__fsqrt_rn(acc), P.eps)));
    } else {
        th = __fsub_rn(th, __fmul_rn(P.eta, g));
    }
}

// Step ONE finished feature (the merge kernel's: a feature whose run straddled chunks): its factor row is
// loaded, stepped and stored by the warp, lanes over factors; lane 0 also steps w.
template <int NF>
__device__ __forceinline__ void fm_apply_one(const FmUpdParams &P, const FmPieceHdr &h, const double *gs, int lane)
{
    const int k = P.k;
    const uint32_t j = h.key;
    if (j >= P.dims) return;
#pragma unroll
    for (int q = 0; q < NF; q++) {
        const int fi = lane + 32 * q;
        if (fi < k) {
            float v = P.vrec[(size_t)j * P.vw + fi];
            float a = P.adagrad ? P.vrec[(size_t)j * P.vw + k + fi] : 0.0f;
            const double G = gs[fi] - (double)v * h.a2;
            fm_step(P, fm_grad(P, G, h.cnt), P.lambda_v, v, a);
            P.vrec[(size_t)j * P.vw + fi] = v;
            if (P.adagrad) P.vrec[(size_t)j * P.vw + k + fi] = a;
        }
    }
    if (lane == 0) {
        float wv = P.wrec[(size_t)j * P.ww], wa = P.adagrad ? P.wrec[(size_t)j * P.ww + 1] : 0.0f;
        fm_step(P, fm_grad(P, h.a1, h.cnt), P.lambda_w, wv, wa);
        P.wrec[(size_t)j * P.ww] = wv;
        if (P.adagrad) P.wrec[(size_t)j * P.ww + 1] = wa;
        P.touched[j] = 1;
    }
}

// ---- update: one 1024-element chunk of the feature-sorted mini-batch per CTA ------------
//
// Every thread owns four consecutive elements.  The segment structure (where a feature's run
// starts and ends, which list slot a finished feature gets, whether a run continues outside the
// chunk) is computed once; then the per-feature fp64 sums are formed kFmNV values at a time with
// the same thread-serial + warp-shuffle + cross-warp segmented scan as the linear learner: group
// 0 carries (A1, A2), the following groups four factors each (GS_f).  After each group the
// finished features' parameters are stepped (loads batched per thread).  Runs that straddle
// chunks leave their piece in first_* / last_* for fm_merge_kernel, exactly as before.
constexpr int kFmThreads = 256;
constexpr int kFmItems = kFmChunk / kFmThreads;     // 4
#ifndef HM_FM_NV
#define HM_FM_NV 4
#endif
constexpr int kFmNV = HM_FM_NV;                     // values reduced per pass (a multiple of 4)
static_assert(kFmNV % 4 == 0, "factor groups are moved as 16-byte vectors");
constexpr int kFmPad = kFmChunk + kFmChunk / 32 + 2;
__device__ __forceinline__ int fm_pad(int q) { return q + (q >> 5); }

struct FmScan { double v[kFmNV]; uint32_t c; uint32_t f; };
__device__ __forceinline__ FmScan fm_combine(const FmScan &a, const FmScan &b)
{
    FmScan r;
    r.f = a.f | b.f;
    r.c = b.f ? b.c : a.c + b.c;
#pragma unroll
    for (int v = 0; v < kFmNV; v++) r.v[v] = b.f ? b.v[v] : a.v[v] + b.v[v];
    return r;
}

constexpr size_t kFmSmem =
    (size_t)kFmPad * 4 * 4                       // key, row, c1, c2
    + (size_t)kFmChunk * (4 + 4 + 8 + 8)         // list: key, cnt, A1, A2
    + (size_t)kFmChunk * kFmNV * 8;              // list: the current group's sums

#ifndef HM_FM_STEPB
#define HM_FM_STEPB 2
#endif
#ifndef HM_FM_MINB
#define HM_FM_MINB 2
#endif
__global__ void __launch_bounds__(kFmThreads, HM_FM_MINB) fm_update_kernel(FmUpdParams P)
{
    extern __shared__ __align__(16) unsigned char fm_smem[];
    uint32_t *s_key = reinterpret_cast<uint32_t *>(fm_smem);
    uint32_t *s_row = s_key + kFmPad;
    float *s_c1 = reinterpret_cast<float *>(s_row + kFmPad);
    float *s_c2 = s_c1 + kFmPad;
    uint32_t *l_key = reinterpret_cast<uint32_t *>(s_c2 + kFmPad);
    uint32_t *l_cnt = l_key + kFmChunk;
    double *l_a1 = reinterpret_cast<double *>(l_cnt + kFmChunk);
    double *l_a2 = l_a1 + kFmChunk;
    double *l_gs = l_a2 + kFmChunk;              // [kFmNV][kFmChunk]
    __shared__ double s_warp_v[kFmThreads / 32][kFmNV];
    __shared__ uint32_t s_warp_c[kFmThreads / 32];
    __shared__ int s_wk[kFmThreads / 32];
    __shared__ uint32_t s_edge[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = P.k;
    const uint32_t s0 = P.seg_off[P.seg], n = P.seg_off[P.seg + 1] - s0;
    const int chunk = blockIdx.x;
    const uint32_t c0 = (uint32_t)chunk * kFmChunk, c1e = min(n, c0 + (uint32_t)kFmChunk);
    FmPieceHdr none; none.a1 = 0; none.a2 = 0; none.key = 0; none.cnt = 0; none.pad[0] = none.pad[1] = 0;
    if (c0 >= c1e) {
        if (tid == 0) { P.first_hdr[chunk] = none; P.last_hdr[chunk] = none; }
        return;
    }
    const int nt = (int)(c1e - c0);
    // does the chunk's first / last feature run continue outside the chunk?  (consumed after the barrier)
    if (tid == 32) {
        s_edge[0] = c0 > 0 && __ldg(P.skey + s0 + c0 - 1) == __ldg(P.skey + s0 + c0);
        s_edge[1] = c1e < n && __ldg(P.skey + s0 + c1e) == __ldg(P.skey + s0 + c1e - 1);
    }
    // ---- load: coalesced, striped; c1 = dloss*x, c2 = c1*x (fp32 products)
    {
        uint32_t kx[kFmItems]; uint2 pv[kFmItems];
#pragma unroll
        for (int i = 0; i < kFmItems; i++) {
            const int q = i * kFmThreads + tid;
            if (q < nt) { kx[i] = __ldcs(P.skey + s0 + c0 + q); pv[i] = __ldcs(P.spay + s0 + c0 + q); }
        }
#pragma unroll
        for (int i = 0; i < kFmItems; i++) {
            const int q = i * kFmThreads + tid;
            if (q < nt) {
                const float x = __uint_as_float(pv[i].y);
                const float c1 = __fmul_rn(__ldg(P.dloss + pv[i].x), x);
                s_key[fm_pad(q)] = kx[i]; s_row[fm_pad(q)] = pv[i].x;
                s_c1[fm_pad(q)] = c1; s_c2[fm_pad(q)] = __fmul_rn(c1, x);
            }
        }
    }
    __syncthreads();
    const bool left_open = s_edge[0] != 0, right_open = s_edge[1] != 0;

    // ---- structure of this thread's items [4*tid, 4*tid+4)
    const int q0 = tid * kFmItems;
    uint32_t kk[kFmItems + 1], rr[kFmItems];
    const uint32_t prev_key = (q0 > 0 && q0 < nt) ? s_key[fm_pad(q0 - 1)] : 0u;
#pragma unroll
    for (int i = 0; i <= kFmItems; i++) kk[i] = q0 + i < nt ? s_key[fm_pad(q0 + i)] : 0u;
#pragma unroll
    for (int i = 0; i < kFmItems; i++) rr[i] = q0 + i < nt ? s_row[fm_pad(q0 + i)] : 0u;
    bool head[kFmItems], closes[kFmItems];
    int my_closes = 0; uint32_t my_head = 0;
#pragma unroll
    for (int i = 0; i < kFmItems; i++) {
        const int q = q0 + i;
        head[i] = q < nt && q > 0 && kk[i] != (i == 0 ? prev_key : kk[i - 1]);
        closes[i] = q < nt && (q < nt - 1 ? kk[i + 1] != kk[i] : !right_open);
        my_closes += closes[i]; my_head |= head[i];
    }
    // exclusive scans over threads: list slot base, and "is there a head before my first item"
    __shared__ uint32_t s_anyhead[kFmThreads / 32];
    int inc = my_closes;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    const uint32_t warp_heads = __ballot_sync(0xffffffffu, my_head != 0);
    if (lane == 31) s_wk[warp] = inc;
    if (lane == 0) s_anyhead[warp] = warp_heads != 0;
    __syncthreads();
    int slot0 = inc - my_closes;
    // The segment structure does not depend on the factor group, so the flags of the segmented scans are
    // worked out once: `reach` = how many lanes to my left belong to the run that is open at my last item
    // (the inclusive scan adds lane - o iff o <= reach), `head_in_warp_before` = a run starts in a lower lane
    // of my warp (then nothing is carried in from earlier warps), `carry_from` = first warp whose total is
    // carried into mine (the last earlier warp that holds a head, or warp 0).
    const uint32_t heads_le = warp_heads & (0xffffffffu >> (31 - lane));
    const int reach = heads_le ? lane - (31 - __clz(heads_le)) : lane;
    const bool head_in_warp_before = (warp_heads & ((1u << lane) - 1u)) != 0;
    bool seen_head_before = head_in_warp_before;
    int carry_from = 0;
    for (int w2 = 0; w2 < warp; w2++) {
        slot0 += s_wk[w2];
        if (s_anyhead[w2] != 0) { seen_head_before = true; carry_from = w2; }
    }
    int n_list = 0;
    for (int w2 = 0; w2 < kFmThreads / 32; w2++) n_list += s_wk[w2];
    // length of the run that is open when my first item begins (the same segmented scan, on counts)
    uint32_t cnt_in = 0;
    {
        uint32_t c = 0;
#pragma unroll
        for (int i = 0; i < kFmItems; i++)
            if (q0 + i < nt) { if (head[i]) c = 0; c += 1; }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t uc = __shfl_up_sync(0xffffffffu, c, o); if (o <= reach) c += uc; }
        if (lane == 31) s_warp_c[warp] = c;
        cnt_in = __shfl_up_sync(0xffffffffu, c, 1);
        if (lane == 0) cnt_in = 0;
        __syncthreads();
        if (!head_in_warp_before)
            for (int w2 = carry_from; w2 < warp; w2++) cnt_in += s_warp_c[w2];
    }

    const int n_groups = 1 + (k + kFmNV - 1) / kFmNV;
#pragma unroll 1
    for (int g = 0; g < n_groups; g++) {
        const int f0 = (g - 1) * kFmNV;          // first factor of the group (g >= 1)
        // ---- values of my items
        float val[kFmItems][kFmNV];
#pragma unroll
        for (int i = 0; i < kFmItems; i++) {
            const int q = q0 + i;
#pragma unroll
            for (int v = 0; v < kFmNV; v++) val[i][v] = 0.0f;
            if (q < nt) {
                if (g == 0) { val[i][0] = s_c1[fm_pad(q)]; val[i][1] = s_c2[fm_pad(q)]; }
                else {
                    const float c1 = s_c1[fm_pad(q)];
                    if ((k & 3) == 0) {     // rows of S are 16-byte aligned: one load per four factors
#pragma unroll
                        for (int h = 0; h < kFmNV / 4; h++)
                            if (f0 + 4 * h < k) {
                                const float4 sv = __ldg(reinterpret_cast<const float4 *>(P.stab + (size_t)rr[i] * k + f0 + 4 * h));
                                val[i][4 * h + 0] = __fmul_rn(c1, sv.x); val[i][4 * h + 1] = __fmul_rn(c1, sv.y);
                                val[i][4 * h + 2] = __fmul_rn(c1, sv.z); val[i][4 * h + 3] = __fmul_rn(c1, sv.w);
                            }
                    } else {
#pragma unroll
                        for (int v = 0; v < kFmNV; v++)
                            if (f0 + v < k) val[i][v] = __fmul_rn(c1, __ldg(P.stab + (size_t)rr[i] * k + f0 + v));
                    }
                }
            }
        }
        // ---- thread summary (sum and count of my trailing run), inclusive segmented scan over the warp with
        // the precomputed flags, carry from the earlier warps
        double sc[kFmNV];
#pragma unroll
        for (int v = 0; v < kFmNV; v++) sc[v] = 0.0;
#pragma unroll
        for (int i = 0; i < kFmItems; i++)
            if (q0 + i < nt) {
                if (head[i]) {
#pragma unroll
                    for (int v = 0; v < kFmNV; v++) sc[v] = 0.0; }
#pragma unroll
                for (int v = 0; v < kFmNV; v++) sc[v] += (double)val[i][v];
            }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const bool take = o <= reach;
#pragma unroll
            for (int v = 0; v < kFmNV; v++) {
                const double up = __shfl_up_sync(0xffffffffu, sc[v], o);
                if (take) sc[v] = up + sc[v];
            }
        }
        __syncthreads();                           // s_warp_v / l_gs of the previous group are free
        if (lane == 31) {
#pragma unroll
            for (int v = 0; v < kFmNV; v++) s_warp_v[warp][v] = sc[v];
        }
        double run[kFmNV]; uint32_t cnt = cnt_in;  // the run that is open when my first item begins
#pragma unroll
        for (int v = 0; v < kFmNV; v++) { run[v] = __shfl_up_sync(0xffffffffu, sc[v], 1); if (lane == 0) run[v] = 0.0; }
        __syncthreads();
        if (!head_in_warp_before && warp > 0) {
            double wp[kFmNV];
#pragma unroll
            for (int v = 0; v < kFmNV; v++) wp[v] = 0.0;
            for (int w2 = carry_from; w2 < warp; w2++) {
#pragma unroll
                for (int v = 0; v < kFmNV; v++) wp[v] += s_warp_v[w2][v];
            }
#pragma unroll
            for (int v = 0; v < kFmNV; v++) run[v] = wp[v] + run[v];
        }
        // ---- second walk: totals of the runs that end in my items
        bool seen_head = seen_head_before;
        int slot = slot0;
#pragma unroll
        for (int i = 0; i < kFmItems; i++) {
            const int q = q0 + i;
            if (q < nt) {
                if (head[i]) { seen_head = true; cnt = 0;
#pragma unroll
                    for (int v = 0; v < kFmNV; v++) run[v] = 0.0; }
                cnt += 1;
#pragma unroll
                for (int v = 0; v < kFmNV; v++) run[v] += (double)val[i][v];
                const bool open_left = !seen_head && left_open;      // the run began before the chunk
                const bool tail_piece = q == nt - 1 && right_open;   // the run continues after the chunk
                if (closes[i] || tail_piece) {
                    // a piece for fm_merge_kernel (first: open to the left, or the whole chunk; last: open to the right)
                    const bool to_first = open_left, to_last = !open_left && tail_piece;
                    if (to_first || to_last) {
                        FmPieceHdr *hd = to_first ? P.first_hdr : P.last_hdr;
                        double *gsp = to_first ? P.first_gs : P.last_gs;
                        if (g == 0) { FmPieceHdr h = none; h.a1 = run[0]; h.a2 = run[1]; h.key = kk[i]; h.cnt = cnt; hd[chunk] = h; }
                        else {
#pragma unroll
                            for (int v = 0; v < kFmNV; v++) if (f0 + v < k) gsp[(size_t)chunk * k + f0 + v] = run[v];
                        }
                    }
                    if (closes[i]) {
                        if (to_first) { if (g == 0) { l_key[slot] = 0xffffffffu; l_cnt[slot] = 0; } }
                        else if (g == 0) { l_key[slot] = kk[i]; l_cnt[slot] = cnt; l_a1[slot] = run[0]; l_a2[slot] = run[1]; }
                        else {
#pragma unroll
                            for (int v = 0; v < kFmNV; v++) l_gs[v * kFmChunk + slot] = run[v];
                        }
                        slot++;
                    }
                }
            }
        }
        __syncthreads();
        // ---- step the finished features: group 0 -> w (and touched), group g -> factors f0..f0+3
        constexpr int kPerT = kFmChunk / kFmThreads;
        if (g == 0) {
            float wv[kPerT], wa[kPerT]; bool ok[kPerT];
#pragma unroll
            for (int u = 0; u < kPerT; u++) {
                const int e = u * kFmThreads + tid;
                ok[u] = e < n_list && l_key[e] < P.dims;
                if (ok[u]) { const uint32_t j = l_key[e]; wv[u] = P.wrec[(size_t)j * P.ww]; wa[u] = P.adagrad ? P.wrec[(size_t)j * P.ww + 1] : 0.0f; }
            }
#pragma unroll
            for (int u = 0; u < kPerT; u++)
                if (ok[u]) {
                    const int e = u * kFmThreads + tid;
                    const uint32_t j = l_key[e];
                    fm_step(P, fm_grad(P, l_a1[e], l_cnt[e]), P.lambda_w, wv[u], wa[u]);
                    P.wrec[(size_t)j * P.ww] = wv[u];
                    if (P.adagrad) P.wrec[(size_t)j * P.ww + 1] = wa[u];
                    P.touched[j] = 1;
                }
        } else {
            // two list entries per thread at a time: all four would hold 32 more registers across the step
            constexpr int kStepB = HM_FM_STEPB;
#pragma unroll 1
            for (int ub = 0; ub < kPerT; ub += kStepB) {
                float vv[kStepB][kFmNV], va[kStepB][kFmNV]; bool ok[kStepB];
                const bool vec = (k & 3) == 0;          // factor records are 16-byte aligned: one load / store per four factors
    #pragma unroll
                for (int u = 0; u < kStepB; u++) {
                    const int e = (ub + u) * kFmThreads + tid;
                    ok[u] = e < n_list && l_key[e] < P.dims;
                    if (ok[u]) {
                        const uint32_t j = l_key[e];
                        if (vec) {
    #pragma unroll
                            for (int h = 0; h < kFmNV / 4; h++) {
                                va[u][4 * h] = va[u][4 * h + 1] = va[u][4 * h + 2] = va[u][4 * h + 3] = 0.0f;
                                if (f0 + 4 * h < k) {
                                    const float4 a4 = *reinterpret_cast<const float4 *>(P.vrec + (size_t)j * P.vw + f0 + 4 * h);
                                    vv[u][4 * h] = a4.x; vv[u][4 * h + 1] = a4.y; vv[u][4 * h + 2] = a4.z; vv[u][4 * h + 3] = a4.w;
                                    if (P.adagrad) {
                                        const float4 b4 = *reinterpret_cast<const float4 *>(P.vrec + (size_t)j * P.vw + k + f0 + 4 * h);
                                        va[u][4 * h] = b4.x; va[u][4 * h + 1] = b4.y; va[u][4 * h + 2] = b4.z; va[u][4 * h + 3] = b4.w;
                                    }
                                }
                            }
                        } else {
    #pragma unroll
                            for (int v = 0; v < kFmNV; v++)
                                if (f0 + v < k) {
                                    vv[u][v] = P.vrec[(size_t)j * P.vw + f0 + v];
                                    va[u][v] = P.adagrad ? P.vrec[(size_t)j * P.vw + k + f0 + v] : 0.0f;
                                }
                        }
                    }
                }
    #pragma unroll
                for (int u = 0; u < kStepB; u++)
                    if (ok[u]) {
                        const int e = (ub + u) * kFmThreads + tid;
                        const uint32_t j = l_key[e];
                        const double a2 = l_a2[e]; const uint32_t cn = l_cnt[e];
    #pragma unroll
                        for (int v = 0; v < kFmNV; v++)
                            if (f0 + v < k) {
                                const double G = l_gs[v * kFmChunk + e] - (double)vv[u][v] * a2;
                                fm_step(P, fm_grad(P, G, cn), P.lambda_v, vv[u][v], va[u][v]);
                            }
                        if (vec) {
    #pragma unroll
                            for (int h = 0; h < kFmNV / 4; h++)
                                if (f0 + 4 * h < k) {
                                    *reinterpret_cast<float4 *>(P.vrec + (size_t)j * P.vw + f0 + 4 * h) =
                                        make_float4(vv[u][4 * h], vv[u][4 * h + 1], vv[u][4 * h + 2], vv[u][4 * h + 3]);
                                    if (P.adagrad)
                                        *reinterpret_cast<float4 *>(P.vrec + (size_t)j * P.vw + k + f0 + 4 * h) =
                                            make_float4(va[u][4 * h], va[u][4 * h + 1], va[u][4 * h + 2], va[u][4 * h + 3]);
                                }
                        } else {
    #pragma unroll
                            for (int v = 0; v < kFmNV; v++)
                                if (f0 + v < k) {
                                    P.vrec[(size_t)j * P.vw + f0 + v] = vv[u][v];
                                    if (P.adagrad) P.vrec[(size_t)j * P.vw + k + f0 + v] = va[u][v];
                                }
                        }
                    }
        }
        }
    }
    // a chunk without a piece on one side says so (group 0 wrote the headers that exist)
    __syncthreads();
    if (tid == 0) {
        // first piece exists iff the first run is open to the left; last piece iff the trailing run is
        // open to the right and did not begin before the chunk
        const bool whole = left_open && right_open && s_key[fm_pad(0)] == s_key[fm_pad(nt - 1)];
        if (!left_open) P.first_hdr[chunk] = none;
        if (!right_open || whole) P.last_hdr[chunk] = none;
    }
}

// One warp per chunk: a non-empty `last` piece starts a run that continues through the `first`
// pieces of the following chunks while the feature matches.  The extra last CTA steps w0.
template <int NF>
__global__ void __launch_bounds__(kFmWarps * 32) fm_merge_kernel(FmUpdParams P)
{
    __shared__ __align__(16) unsigned char m_smem[kFmWarps * (sizeof(FmPieceHdr) + 256 * sizeof(double))];
    __shared__ double s_part[kFmWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k = P.k;
    if ((int)blockIdx.x == (P.n_chunks + kFmWarps - 1) / kFmWarps) {
        // ---- w0: fp64 sum of dloss over the mini-batch in a fixed order
        double s = 0.0;
        for (int i = threadIdx.x; i < P.n_rows; i += kFmWarps * 32) s += (double)P.dloss[i];
        s = warp_sum(s);
        if (lane == 0) s_part[warp] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            double tot = 0.0;
            for (int w = 0; w < kFmWarps; w++) tot += s_part[w];
            float w0 = P.w0rec[0], acc = P.w0rec[1];
            const float g = P.grad_norm == HM_NORM_SUM ? (float)tot : (float)(tot * P.inv_rows);
            fm_step(P, g, P.lambda_w0, w0, acc);
            P.w0rec[0] = w0; P.w0rec[1] = acc;
        }
        return;
    }
    const int c = blockIdx.x * kFmWarps + warp;
    if (c >= P.n_chunks) return;
    const FmPieceHdr start = P.last_hdr[c];
    if (start.cnt == 0) return;
    double a1 = start.a1, a2 = start.a2, gs[NF];
    uint32_t cnt = start.cnt;
#pragma unroll
    for (int q = 0; q < NF; q++) { const int fi = lane + 32 * q; gs[q] = fi < k ? P.last_gs[(size_t)c * k + fi] : 0.0; }
    // walk forward four pieces at a time: the loads are issued together, the adds stay in order
    int j = c + 1;
    bool more = true;
    while (more && j < P.n_chunks) {
        FmPieceHdr h[4]; double g4[4][NF];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int jj = min(j + u, P.n_chunks - 1);
            h[u] = P.first_hdr[jj];
#pragma unroll
            for (int q = 0; q < NF; q++) { const int fi = lane + 32 * q; g4[u][q] = fi < k ? P.first_gs[(size_t)jj * k + fi] : 0.0; }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (more && j + u < P.n_chunks && h[u].cnt != 0 && h[u].key == start.key) {
                a1 += h[u].a1; a2 += h[u].a2; cnt += h[u].cnt;
#pragma unroll
                for (int q = 0; q < NF; q++) gs[q] += g4[u][q];
            } else {
                more = false;
            }
        }
        j += 4;
    }
    // one finished feature
    FmPieceHdr *ph = reinterpret_cast<FmPieceHdr *>(m_smem + (size_t)warp * (sizeof(FmPieceHdr) + 256 * sizeof(double)));
    double *pgs = reinterpret_cast<double *>(reinterpret_cast<unsigned char *>(ph) + sizeof(FmPieceHdr));
    if (lane == 0) { ph[0].a1 = a1; ph[0].a2 = a2; ph[0].key = start.key; ph[0].cnt = cnt; }
#pragma unroll
    for (int q = 0; q < NF; q++) { const int fi = lane + 32 * q; if (fi < k) pgs[fi] = gs[q]; }
    __syncwarp();
    fm_apply_one<NF>(P, ph[0], pgs, lane);
}

__global__ void fm_gather_vif_kernel(const float *__restrict__ vrec, int vw, int k, const int32_t *__restrict__ feature,
                                     int64_t n, float *__restrict__ vif)
{
    const int64_t total = n * k;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int64_t m = q / k;
        vif[q] = vrec[(size_t)feature[m] * vw + (q - m * k)];
    }
}

int fm_grid(int64_t ctas_wanted, int per_sm)
{
    const int64_t cap = (int64_t)num_sms() * per_sm;
    if (ctas_wanted < 1) ctas_wanted = 1;
    return (int)(ctas_wanted < cap ? ctas_wanted : cap);
}


}  // namespace

hm_status launch_fm_init_v(Launcher &L, const FmArgs &a)
{
    const int64_t n = a.cfg.dims * a.k;
    HM_LAUNCH(L, K_MISC, 4.0 * (double)n,
              (fm_init_v_kernel<<<fm_grid((n + 1023) / 1024, 8), 256, 0, L.stream>>>(a.vrec, a.cfg.dims, a.k, a.vw,
                                                                                    a.cfg.sigma, a.cfg.seed)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_fm_forward(Launcher &L, const FmArgs &a, const int64_t *d_rowptr, int64_t row0, int32_t n_rows,
                            int64_t block_row0, const uint32_t *d_feature, const float *d_value, const float *d_label,
                            float *d_margin, float *d_dloss, float *d_stab, uint32_t *d_bad, int64_t nnz_hint, bool train)
{
    if (n_rows <= 0) return HM_OK;
    FmFwdParams P;
    P.w0rec = a.w0rec; P.wrec = a.wrec; P.vrec = a.vrec; P.ww = a.ww; P.vw = a.vw; P.k = a.k;
    P.rowptr = d_rowptr; P.feature = d_feature; P.value = d_value; P.label = d_label;
    P.margin = d_margin; P.dloss = d_dloss; P.stab = d_stab;
    P.block_row0 = block_row0; P.row0 = row0; P.n_rows = n_rows;
    P.classification = a.cfg.fm_classification; P.train = train ? 1 : 0;
    P.dims = (uint32_t)a.cfg.dims; P.bad = d_bad;
    const int grid = fm_grid(((int64_t)n_rows + 7) / 8, 8);
    // algorithmic bytes: feature+value, w and the k factors per element; S table row + scalars per row
    const double bytes = (12.0 + 4.0 * a.k) * (double)nnz_hint + (20.0 + 4.0 * a.k) * (double)n_rows;
    const int id = train ? K_FM_FORWARD : K_PREDICT;
    if (a.k <= 16) HM_LAUNCH(L, id, bytes, (fm_forward_kernel<1, 16><<<fm_grid(((int64_t)n_rows + 15) / 16, 8), 256, 0, L.stream>>>(P)));
    else if (a.k <= 32) HM_LAUNCH(L, id, bytes, (fm_forward_kernel<1, 32><<<grid, 256, 0, L.stream>>>(P)));
    else HM_LAUNCH(L, id, bytes, (fm_forward_kernel<2, 32><<<grid, 256, 0, L.stream>>>(P)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status fm_scratch_reserve(FmScratch &s, int64_t max_mb_elems, int k)
{
    const int64_t chunks = (max_mb_elems + kFmChunk - 1) / kFmChunk + 1;
    if (chunks <= s.cap_chunks && k == s.k) return HM_OK;
    fm_scratch_free(s);
    HM_CUDA(cudaMalloc(&s.hdr, (size_t)chunks * 2 * sizeof(FmPieceHdr)));
    HM_CUDA(cudaMalloc(&s.gs, (size_t)chunks * 2 * k * sizeof(double)));
    s.cap_chunks = chunks; s.k = k;
    return HM_OK;
}

void fm_scratch_free(FmScratch &s)
{
    cudaFree(s.hdr); cudaFree(s.gs);
    s.hdr = nullptr; s.gs = nullptr; s.cap_chunks = 0; s.k = 0;
}

hm_status launch_fm_update(Launcher &L, const FmArgs &a, const StepScalars &sc, FmScratch &fs,
                           const uint32_t *d_sorted_key, const uint2 *d_sorted_pay, const uint32_t *d_seg_off, int seg,
                           const float *d_dloss, const float *d_stab, int64_t nnz)
{
    // > 48 KB of dynamic shared memory needs opting in, once per device
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    HM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        HM_CUDA(cudaFuncSetAttribute(fm_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFmSmem));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    const int n_chunks = (int)std::max<int64_t>((nnz + kFmChunk - 1) / kFmChunk, 1);
    if (n_chunks > fs.cap_chunks) HM_FAIL(HM_ERR_STATE, "fm scratch too small");
    FmUpdParams P;
    P.w0rec = a.w0rec; P.wrec = a.wrec; P.vrec = a.vrec; P.touched = a.touched;
    P.ww = a.ww; P.vw = a.vw; P.k = a.k;
    P.skey = d_sorted_key; P.spay = d_sorted_pay; P.dloss = d_dloss; P.stab = d_stab; P.seg_off = d_seg_off; P.seg = seg;
    P.dims = (uint32_t)a.cfg.dims;
    P.adagrad = a.cfg.opt == HM_OPT_ADAGRAD; P.grad_norm = a.cfg.grad_norm;
    P.eta = sc.eta; P.eps = a.cfg.eps; P.lambda_w0 = a.cfg.lambda_w0; P.lambda_w = a.cfg.lambda_w; P.lambda_v = a.cfg.lambda_v;
    P.inv_rows = 1.0 / (double)sc.n_rows; P.n_rows = sc.n_rows;
    P.first_hdr = reinterpret_cast<FmPieceHdr *>(fs.hdr); P.last_hdr = P.first_hdr + fs.cap_chunks;
    P.first_gs = fs.gs; P.last_gs = fs.gs + (size_t)fs.cap_chunks * a.k;
    P.n_chunks = n_chunks;
    const int ctas = (n_chunks + kFmWarps - 1) / kFmWarps;
    // algorithmic bytes: sorted stream, S row per element (the per-feature record traffic is data dependent)
    const double bytes = (12.0 + 4.0 * a.k) * (double)nnz;
    HM_LAUNCH(L, K_FM_UPDATE, bytes, (fm_update_kernel<<<n_chunks, kFmThreads, kFmSmem, L.stream>>>(P)));
    if (a.k <= 32) HM_LAUNCH(L, K_FM_UPDATE, 0.0, (fm_merge_kernel<1><<<ctas + 1, kFmWarps * 32, 0, L.stream>>>(P)));
    else HM_LAUNCH(L, K_FM_UPDATE, 0.0, (fm_merge_kernel<2><<<ctas + 1, kFmWarps * 32, 0, L.stream>>>(P)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_fm_gather_vif(Launcher &L, const FmArgs &a, const int32_t *d_feature, int64_t n, float *d_vif)
{
    if (n <= 0) return HM_OK;
    HM_LAUNCH(L, K_EXPORT, 8.0 * (double)n * a.k,
              (fm_gather_vif_kernel<<<fm_grid((n * a.k + 1023) / 1024, 8), 256, 0, L.stream>>>(a.vrec, a.vw, a.k, d_feature,
                                                                                            n, d_vif)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

}  // namespace hm
