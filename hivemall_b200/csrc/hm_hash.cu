// hm_hash.cu -- MurmurHash3_x86_32 feature hashing on the GPU.
//
// Stands in for the mhash / feature_hashing UDFs named by BASELINE.json's north_star
// ("MurmurHash3 feature hashing (mhash / feature_hashing)", bit-exact).  The
// reference mount has no source (README.md:20-22): the hash is Appleby's published
// MurmurHash3_x86_32; seed and folding are run-time parameters.
//
// Layout: n strings packed back to back, offsets[n+1].  One thread hashes one string;
// a CTA first stages the contiguous byte span of its 256 strings into shared memory
// with coalesced 16-byte loads (HBM traffic = the bytes once + 8 B offsets + 4 B out),
// then threads walk their strings out of shared memory.
#include "hm_internal.h"

namespace hm {

namespace {

constexpr int kHashThreads = 256;
constexpr int kHashSmemBytes = 40 * 1024;  // staged span; longer spans read global memory directly

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

__device__ __forceinline__ uint32_t fmix32(uint32_t h)
{
    h ^= h >> 16; h *= 0x85ebca6bu;
    h ^= h >> 13; h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}

// Byte source: shared memory when staged, else global.
struct Bytes {
    const uint8_t *p;
    int64_t base;  // byte offset that p[0] corresponds to
    __device__ __forceinline__ uint32_t operator[](int64_t i) const { return p[i - base]; }
};

__device__ __forceinline__ uint32_t murmur3_32(const Bytes &b, int64_t start, int64_t len, uint32_t seed)
{
    const uint32_t c1 = 0xcc9e2d51u, c2 = 0x1b873593u;
    uint32_t h1 = seed;
    const int64_t nblocks = len >> 2;
    for (int64_t i = 0; i < nblocks; i++) {
        const int64_t o = start + 4 * i;
        uint32_t k1 = b[o] | (b[o + 1] << 8) | (b[o + 2] << 16) | (b[o + 3] << 24);
        k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2;
        h1 ^= k1; h1 = rotl32(h1, 13); h1 = h1 * 5u + 0xe6546b64u;
    }
    const int64_t t = start + 4 * nblocks;
    uint32_t k1 = 0;
    const int rem = (int)(len & 3);
    if (rem == 3) k1 ^= b[t + 2] << 16;
    if (rem >= 2) k1 ^= b[t + 1] << 8;
    if (rem >= 1) {
        k1 ^= b[t];
        k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2; h1 ^= k1;
    }
    h1 ^= (uint32_t)len;
    return fmix32(h1);
}

__device__ __forceinline__ int32_t fold(uint32_t h, int32_t num_features)
{
    int32_t r = (int32_t)h % num_features;  // truncating remainder, as Java's %
    if (r < 0) r += num_features;
    return r;
}

__device__ __forceinline__ bool is_digit(uint32_t c) { return c >= '0' && c <= '9'; }

// Parse bytes [s, e) as a decimal float: [+-]digits[.digits][(e|E)[+-]digits].
// Mantissa kept to 19 digits in u64, scaled in double, rounded once to float.
__device__ bool parse_float(const Bytes &b, int64_t s, int64_t e, float *out)
{
    if (s >= e) return false;
    bool neg = false;
    if (b[s] == '-' || b[s] == '+') { neg = b[s] == '-'; s++; }
    uint64_t mant = 0;
    int nd = 0, dexp = 0, digits = 0;
    while (s < e && is_digit(b[s])) {
        if (nd < 19) { mant = mant * 10 + (b[s] - '0'); if (mant) nd++; } else dexp++;
        s++; digits++;
    }
    if (s < e && b[s] == '.') {
        s++;
        while (s < e && is_digit(b[s])) {
            if (nd < 19) { mant = mant * 10 + (b[s] - '0'); if (mant) nd++; dexp--; }
            s++; digits++;
        }
    }
    if (digits == 0) return false;
    if (s < e && (b[s] == 'e' || b[s] == 'E')) {
        s++;
        bool eneg = false;
        if (s < e && (b[s] == '-' || b[s] == '+')) { eneg = b[s] == '-'; s++; }
        if (s >= e) return false;
        int ex = 0;
        while (s < e && is_digit(b[s])) { if (ex < 10000) ex = ex * 10 + (int)(b[s] - '0'); s++; }
        dexp += eneg ? -ex : ex;
    }
    if (s != e) return false;
    double v = (double)mant;
    // exact powers of ten up to 1e22 keep short decimals correctly rounded
    const double p10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14,
                            1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    while (dexp > 0) { int k = dexp > 22 ? 22 : dexp; v *= p10[k]; dexp -= k; }
    while (dexp < 0) { int k = -dexp > 22 ? 22 : -dexp; v /= p10[k]; dexp += k; }
    *out = (float)(neg ? -v : v);
    return true;
}

// MODE 0: mhash (index only).  MODE 1: feature_hashing ("name[:value]" -> index, value).
template <int MODE>
__global__ void __launch_bounds__(kHashThreads)
hash_kernel(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ offsets, int64_t n,
            uint32_t seed, int32_t num_features, int32_t *__restrict__ out_index, float *__restrict__ out_value)
{
    __shared__ __align__(16) uint8_t stage[kHashSmemBytes];
    const int64_t tiles = (n + kHashThreads - 1) / kHashThreads;
    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int64_t s0 = tile * kHashThreads;
        const int64_t s1 = min(s0 + (int64_t)kHashThreads, n);
        const int64_t b0 = offsets[s0], b1 = offsets[s1];
        // start the staged window at a 16 B-aligned ADDRESS so vector loads are legal
        const int64_t a0 = b0 - (int64_t)(reinterpret_cast<uintptr_t>(bytes + b0) & 15);
        const bool staged = (b1 - a0) <= kHashSmemBytes;
        __syncthreads();  // previous tile's readers are done
        if (staged) {
            const int64_t nvec = (b1 - a0 + 15) >> 4;
            uint4 *dst = reinterpret_cast<uint4 *>(stage);
            for (int64_t v = threadIdx.x; v < nvec; v += kHashThreads) {
                const int64_t lo = a0 + 16 * v;
                if (lo >= b0 && lo + 16 <= b1) dst[v] = *reinterpret_cast<const uint4 *>(bytes + lo);
                else for (int q = 0; q < 16; q++) { int64_t o = lo + q; stage[16 * v + q] = (o >= b0 && o < b1) ? bytes[o] : 0; }
            }
        }
        __syncthreads();
        const int64_t s = s0 + threadIdx.x;
        if (s < s1) {
            Bytes b{staged ? stage : bytes, staged ? a0 : 0};
            const int64_t start = offsets[s], end = offsets[s + 1];
            if (MODE == 0) {
                out_index[s] = fold(murmur3_32(b, start, end - start, seed), num_features);
            } else {
                // split at the LAST ':' if the tail is a number
                int64_t colon = -1;
                for (int64_t p = end - 1; p >= start; p--) if (b[p] == ':') { colon = p; break; }
                float value = 1.0f;
                int64_t name_end = end;
                if (colon >= 0 && parse_float(b, colon + 1, end, &value)) name_end = colon; else value = 1.0f;
                // an all-digit name is an index already
                bool digits = name_end > start && (name_end - start) <= 10;
                uint64_t asint = 0;
                for (int64_t p = start; digits && p < name_end; p++) {
                    if (is_digit(b[p])) asint = asint * 10 + (b[p] - '0'); else digits = false;
                }
                if (digits && asint < (uint64_t)num_features) out_index[s] = (int32_t)asint;
                else out_index[s] = fold(murmur3_32(b, start, name_end - start, seed), num_features);
                out_value[s] = value;
            }
        }
    }
}

int hash_grid(int64_t n)
{
    int64_t tiles = (n + kHashThreads - 1) / kHashThreads;
    int64_t cap = (int64_t)num_sms() * 4;
    return (int)(tiles < 1 ? 1 : (tiles < cap ? tiles : cap));
}

}  // namespace

hm_status launch_mhash(Launcher &L, const uint8_t *d_bytes, const int64_t *d_offsets, int64_t n,
                       uint32_t seed, int32_t num_features, int32_t *d_out)
{
    if (n == 0) return HM_OK;
    HM_LAUNCH(L, K_HASH, 0.0,
              (hash_kernel<0><<<hash_grid(n), kHashThreads, 0, L.stream>>>(d_bytes, d_offsets, n, seed, num_features,
                                                                           d_out, nullptr)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

hm_status launch_feature_hashing(Launcher &L, const uint8_t *d_bytes, const int64_t *d_offsets, int64_t n,
                                 uint32_t seed, int32_t num_features, int32_t *d_index, float *d_value, double alg_bytes)
{
    if (n == 0) return HM_OK;
    HM_LAUNCH(L, K_HASH, alg_bytes,
              (hash_kernel<1><<<hash_grid(n), kHashThreads, 0, L.stream>>>(d_bytes, d_offsets, n, seed, num_features,
                                                                           d_index, d_value)));
    HM_CUDA(cudaGetLastError());
    return HM_OK;
}

}  // namespace hm
