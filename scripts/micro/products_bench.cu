// Microbenchmark (not part of the product): forward from the feature-sorted stream.
// p[pos] = w[key] * value for sorted (key, (pos, value)) elements, then row sums over p.
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <numeric>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int B>
__global__ void __launch_bounds__(256) products_kernel(const float2 *__restrict__ table, const uint32_t *__restrict__ key,
                                                       const uint2 *__restrict__ pay, int64_t n, float *__restrict__ p)
{
    for (int64_t q0 = (int64_t)blockIdx.x * 256 * B + threadIdx.x; q0 < n; q0 += (int64_t)gridDim.x * 256 * B) {
        uint32_t f[B]; uint2 pv[B]; float2 w[B];
#pragma unroll
        for (int u = 0; u < B; u++) if (q0 + u * 256 < n) { f[u] = key[q0 + u * 256]; pv[u] = pay[q0 + u * 256]; }
#pragma unroll
        for (int u = 0; u < B; u++) if (q0 + u * 256 < n) w[u] = __ldg(table + f[u]);
#pragma unroll
        for (int u = 0; u < B; u++) if (q0 + u * 256 < n) p[pv[u].x] = w[u].x * __uint_as_float(pv[u].y);
    }
}
// 4 lanes per row, 39 elements per row
__global__ void __launch_bounds__(256) rowsum_kernel(float *__restrict__ p, int n_rows, int per_row, float *__restrict__ margin)
{
    const int row = blockIdx.x * 64 + threadIdx.x / 4, sub = threadIdx.x % 4;
    double acc = 0.0;
    if (row < n_rows) for (int q = sub; q < per_row; q += 4) acc += (double)p[(int64_t)row * per_row + q];
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (row < n_rows) {
        const float d = (float)acc * 0.5f;
        if (sub == 0) margin[row] = (float)acc;
        for (int q = sub; q < per_row; q += 4) p[(int64_t)row * per_row + q] = d;
    }
}
__global__ void fill_kernel(float *p, int64_t n, float v) { for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < n; i += gridDim.x * 256ll) p[i] = v; }

int main()
{
    const int rows = 65536, per_row = 39; const int64_t n = (int64_t)rows * per_row; const int64_t dims = 1 << 24;
    uint64_t s = 88172645463325252ull;
    auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
    // criteo-like: 13 hot features in every row; 26 categorical with u^3 * card skew
    std::vector<uint32_t> feat(n);
    for (int r = 0; r < rows; r++) for (int c = 0; c < per_row; c++) {
        uint32_t v;
        if (c < 13) v = (uint32_t)(c * 2654435761u) % dims;
        else { double u = (rnd() >> 11) * (1.0 / 9007199254740992.0); double card = 10.0 * pow(10.0, (c - 13) * 6.0 / 25.0);
               uint64_t cat = (uint64_t)(u * u * u * card); v = (uint32_t)((cat * 0x9E3779B1ull + c * 0x7FEB352Dull) * 0x2545F4914F6CDD1Dull >> 40) % dims; }
        feat[(int64_t)r * per_row + c] = v;
    }
    std::vector<uint32_t> order(n); std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return feat[a] < feat[b]; });
    std::vector<uint32_t> skey(n); std::vector<uint2> spay(n);
    int64_t distinct = 0;
    for (int64_t i = 0; i < n; i++) { skey[i] = feat[order[i]]; spay[i] = make_uint2(order[i], 0x3f800000u); if (i == 0 || skey[i] != skey[i - 1]) distinct++; }
    printf("elements %lld distinct %lld\n", (long long)n, (long long)distinct);
    float2 *table; uint32_t *dk; uint2 *dp; float *p, *margin, *flush; const int64_t nflush = 96ll << 20;
    CK(cudaMalloc(&table, dims * sizeof(float2))); CK(cudaMalloc(&dk, n * 4)); CK(cudaMalloc(&dp, n * 8)); CK(cudaMalloc(&p, n * 4));
    CK(cudaMalloc(&margin, rows * 4)); CK(cudaMalloc(&flush, nflush * 4));
    CK(cudaMemset(table, 0, dims * sizeof(float2)));
    CK(cudaMemcpy(dk, skey.data(), n * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dp, spay.data(), n * 8, cudaMemcpyHostToDevice));
    cudaEvent_t a, b, c; cudaEventCreate(&a); cudaEventCreate(&b); cudaEventCreate(&c);
    for (int fl = 0; fl < 2; fl++) for (int B = 4; B <= 8; B += 4) {
        float t1 = 0, t2 = 0; const int reps = 10;
        for (int r = 0; r < reps + 2; r++) {
            if (fl) fill_kernel<<<148 * 8, 256>>>(flush, nflush, (float)r);
            cudaEventRecord(a);
            if (B == 4) products_kernel<4><<<(int)((n + 1023) / 1024), 256>>>(table, dk, dp, n, p);
            else products_kernel<8><<<(int)((n + 2047) / 2048), 256>>>(table, dk, dp, n, p);
            cudaEventRecord(b);
            rowsum_kernel<<<rows / 64, 256>>>(p, rows, per_row, margin);
            cudaEventRecord(c); cudaEventSynchronize(c);
            float m1, m2; cudaEventElapsedTime(&m1, a, b); cudaEventElapsedTime(&m2, b, c);
            if (r >= 2) { t1 += m1; t2 += m2; }
        }
        printf("%s B=%d  products %6.1f us   rowsum %6.1f us\n", fl ? "L2 flushed" : "L2 warm   ", B, t1 / reps * 1000, t2 / reps * 1000);
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
