// Phase 1 — solid k-mer counting in HBM (replaces DSK as used by Leon, SURVEY.md 8a-1;
// K/SortingCountAlgorithm.cpp:499-644, K/Model.hpp:725-764,876-883, K/PartitionsCommand.cpp:1206-1460):
//   extract_kmers : one warp per read, canonical k-mer of every window free of invalid characters
//   radix sort    : LSD, 8-bit digits, stable (tile histogram -> digit-major scan -> ranked scatter)
//   run stats     : run heads by scan, abundance histogram (u16 truncation + clamp), solid count
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"
#include "scan.cuh"

namespace lbcount {

// ---- packing helpers -------------------------------------------------------------------------------------
// spread the 32 bits of x to the even bit positions of a 64-bit word
LB_D uint64_t spread_bits(uint32_t x)
{
    uint64_t v = x;
    v = (v | (v << 16)) & 0x0000FFFF0000FFFFULL;
    v = (v | (v << 8)) & 0x00FF00FF00FF00FFULL;
    v = (v | (v << 4)) & 0x0F0F0F0F0F0F0F0FULL;
    v = (v | (v << 2)) & 0x3333333333333333ULL;
    v = (v | (v << 1)) & 0x5555555555555555ULL;
    return v;
}

// One warp packs 32 consecutive bases (lane t holds 2-bit `code` of base t) into a word with base 0 in the two
// most significant bits, i.e. the layout of a 32-mer.
LB_D uint64_t pack32(unsigned code)
{
    uint32_t b0 = __brev(__ballot_sync(0xffffffffu, code & 1));
    uint32_t b1 = __brev(__ballot_sync(0xffffffffu, (code >> 1) & 1));
    return (spread_bits(b1) << 1) | spread_bits(b0);
}

// k-mer starting t bases into the window (w0,w1,w2) of packed 32-base words
template <typename K> LB_D K window_kmer(uint64_t w0, uint64_t w1, uint64_t w2, int t, int k);
template <> LB_D uint64_t window_kmer<uint64_t>(uint64_t w0, uint64_t w1, uint64_t, int t, int k)
{
    uint64_t hi = t ? ((w0 << (2 * t)) | (w1 >> (64 - 2 * t))) : w0;
    return hi >> (64 - 2 * k);
}
template <> LB_D u128 window_kmer<u128>(uint64_t w0, uint64_t w1, uint64_t w2, int t, int k)
{
    uint64_t h = t ? ((w0 << (2 * t)) | (w1 >> (64 - 2 * t))) : w0;
    uint64_t m = t ? ((w1 << (2 * t)) | (w2 >> (64 - 2 * t))) : w1;
    return u128(m, h) >> (128 - 2 * k);
}

// number of k-mers of a read of length L
LB_HD uint64_t nkmers_of(uint64_t L, int k) { return L >= (uint64_t)k ? L - k + 1 : 0; }

struct NkIn {
    const uint64_t *offsets; int k;
    __device__ uint64_t operator()(uint64_t i) const { return nkmers_of(offsets[i + 1] - offsets[i], k); }
};
struct NkOut {
    uint64_t *koff;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t) const { koff[i] = ex; }
};

// ---- extraction ------------------------------------------------------------------------------------------
// keys[koff[r] + p] = canonical k-mer at position p of read r, or the sentinel when the window holds an invalid
// character ((c>>3)&1, MI/api/Data.hpp:176).  One warp per read, 32 positions per step.
template <typename K>
__global__ void __launch_bounds__(256) k_extract(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                 const uint64_t *__restrict__ koff, uint64_t nreads, int k, K *__restrict__ keys)
{
    const int lane = threadIdx.x & 31;
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t r = warp; r < nreads; r += nwarps) {
        uint64_t o = offsets[r], L = offsets[r + 1] - o;
        if (L < (uint64_t)k) continue;
        uint64_t n = L - k + 1;
        const uint8_t *s = bases + o;
        K *out = keys + koff[r];
        uint64_t w0, w1 = 0, w2 = 0;
        uint32_t n0, n1 = 0, n2 = 0;
        // prime the window with chunks 0 and 1
        {
            uint8_t c = (uint64_t)lane < L ? s[lane] : (uint8_t)'A';
            w0 = pack32(base_code(c));
            n0 = __ballot_sync(0xffffffffu, base_invalid(c));
            uint64_t i1 = 32 + lane;
            c = i1 < L ? s[i1] : (uint8_t)'A';
            w1 = pack32(base_code(c));
            n1 = __ballot_sync(0xffffffffu, base_invalid(c));
        }
        for (uint64_t base = 0; base < n; base += 32) {
            uint64_t i2 = base + 64 + lane;
            uint8_t c = i2 < L ? s[i2] : (uint8_t)'A';
            w2 = pack32(base_code(c));
            n2 = __ballot_sync(0xffffffffu, base_invalid(c));
            uint64_t p = base + lane;
            if (p < n) {
                K fwd = window_kmer<K>(w0, w1, w2, lane, k);
                // invalid characters in [lane, lane+k) of the 96-bit mask n0 | n1<<32 | n2<<64 (k <= 63)
                uint64_t nlo = (uint64_t)n0 | ((uint64_t)n1 << 32);
                uint64_t win = lane ? ((nlo >> lane) | ((uint64_t)n2 << (64 - lane))) : nlo;
                win &= (1ULL << k) - 1;
                out[p] = win ? kmer_sentinel<K>() : kmer_canon(fwd, k);
            }
            w0 = w1; w1 = w2; n0 = n1; n1 = n2;
        }
    }
}

// ---- radix sort --------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;     // 4096 keys per CTA
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_WCHUNK = 32 * RS_ITEMS;           // 512 consecutive keys per warp

template <typename K> LB_D unsigned digit_of(K key, int shift) { return (unsigned)(key >> shift) & 255u; }

// tile_hist[d * ntiles + tile] = number of keys of the tile whose digit is d
template <typename K>
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const K *__restrict__ keys, uint64_t n, int shift,
                                                        uint32_t *__restrict__ tile_hist, uint32_t ntiles)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    uint64_t base = (uint64_t)blockIdx.x * RS_TILE;
#pragma unroll
    for (int j = 0; j < RS_ITEMS; j++) {
        uint64_t i = base + (uint64_t)j * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[digit_of(keys[i], shift)], 1u);
    }
    __syncthreads();
    tile_hist[(uint64_t)threadIdx.x * ntiles + blockIdx.x] = h[threadIdx.x];
}

struct HistIn {
    const uint32_t *h;
    __device__ uint64_t operator()(uint64_t i) const { return h[i]; }
};
struct HistOut {
    uint64_t *o;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t) const { o[i] = ex; }
};

// stable scatter: rank of a key among equal digits = (tile offset for the digit) + (keys of earlier warps of the
// tile with that digit) + (earlier keys of this warp with that digit).  Warp w owns keys [w*512, w*512+512) of the
// tile, visited row by row (32 consecutive keys per row), so ranks follow the input order.
template <typename K>
__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const K *__restrict__ in, K *__restrict__ out, uint64_t n, int shift,
                                                           const uint64_t *__restrict__ tile_off, uint32_t ntiles)
{
    __shared__ uint32_t whist[RS_WARPS][256];
    __shared__ uint64_t wbase[RS_WARPS][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();

    K key[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
    uint64_t wstart = (uint64_t)blockIdx.x * RS_TILE + (uint64_t)warp * RS_WCHUNK;
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
        uint64_t i = wstart + (uint64_t)r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? in[i] : kmer_sentinel<K>();
        unsigned d = valid ? digit_of(key[r], shift) : 256u;   // 256 groups the out-of-range lanes together
        uint32_t peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader && valid) {
            old = whist[warp][d];
            whist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[r] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    {
        // thread d turns the per-warp counts of digit d into per-warp global bases
        unsigned d = threadIdx.x;
        uint64_t b = tile_off[(uint64_t)d * ntiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            wbase[w][d] = b;
            b += whist[w][d];
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
        uint64_t i = wstart + (uint64_t)r * 32 + lane;
        if (i < n) out[wbase[warp][digit_of(key[r], shift)] + rank[r]] = key[r];
    }
}

// ---- run-length statistics -----------------------------------------------------------------------------------
template <typename K> struct HeadIn {
    const K *keys;
    __device__ uint32_t operator()(uint64_t i) const { return (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u; }
};
struct HeadOut {
    uint64_t *heads;
    __device__ void operator()(uint64_t i, uint32_t ex, uint32_t v) const { if (v) heads[ex] = i; }
};
// (run count can exceed 2^32 only if n does; the scan type is widened by the caller in that case)
template <typename K> struct HeadIn64 {
    const K *keys;
    __device__ uint64_t operator()(uint64_t i) const { return (i == 0 || keys[i] != keys[i - 1]) ? 1ull : 0ull; }
};
struct HeadOut64 {
    uint64_t *heads;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t v) const { if (v) heads[ex] = i; }
};

struct CountStats { unsigned long long nb_solid_total, nb_distinct, nb_valid; };

// abundance histogram over ALL distinct k-mers: bin = count truncated to u16, clamped to 10000
// (MI/impl/Histogram.hpp:87); solid = abundance >= 2 (K/SortingCountAlgorithm.cpp:210)
template <typename K>
__global__ void __launch_bounds__(256) k_run_stats(const K *__restrict__ keys, const uint64_t *__restrict__ heads, uint64_t nruns,
                                                   uint64_t n, unsigned long long *__restrict__ histo, CountStats *stats)
{
    extern __shared__ uint32_t sh[];  // 10001 bins
    for (int i = threadIdx.x; i < 10001; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    unsigned long long solid = 0, distinct = 0, valid = 0;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < nruns; j += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t s = heads[j], e = j + 1 < nruns ? heads[j + 1] : n;
        if (keys[s] == kmer_sentinel<K>()) continue;
        uint64_t cnt = e - s;
        unsigned idx = (unsigned)(cnt & 0xffffu);
        atomicAdd(&sh[idx >= 10000u ? 10000u : idx], 1u);
        distinct++;
        valid += cnt;
        if (cnt >= 2) solid++;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 10001; i += blockDim.x)
        if (sh[i]) atomicAdd(&histo[i], (unsigned long long)sh[i]);
    // block reduction of the three counters through warp shuffles + atomics
    for (int d = 16; d; d >>= 1) {
        solid += __shfl_down_sync(0xffffffffu, solid, d);
        distinct += __shfl_down_sync(0xffffffffu, distinct, d);
        valid += __shfl_down_sync(0xffffffffu, valid, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (solid) atomicAdd(&stats->nb_solid_total, solid);
        if (distinct) atomicAdd(&stats->nb_distinct, distinct);
        if (valid) atomicAdd(&stats->nb_valid, valid);
    }
}

}  // namespace lbcount
