// Phase 2 — Bloom filter build, bit-identical with BloomNeighborCoherent::insert (COL/Bloom.hpp:555-588) fed by
// BloomBuilder (K/BloomBuilder.hpp:173-180): every solid k-mer (abundance >= 2) whose abundance is >= min_abundance.
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"

namespace lbbloom {

// solid list = (canonical k-mer, abundance) pairs with abundance >= 2, in any order (OR is order-independent)
template <typename K>
__global__ void __launch_bounds__(256) k_insert_solid(const K *__restrict__ keys, const uint32_t *__restrict__ cnt, uint64_t nsolid,
                                                      uint32_t min_abundance, BloomDev b, const uint64_t *__restrict__ tbl_g)
{
    __shared__ uint64_t tbl[256];
    tbl[threadIdx.x] = tbl_g[threadIdx.x];
    __syncthreads();
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < nsolid; j += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t c = cnt[j];
        uint32_t c31 = c > 0x7fffffffu ? 0x7fffffffu : c;
        if ((int32_t)c31 < (int32_t)min_abundance) continue;   // (int)abundance >= _min_abundance, K/BloomBuilder.hpp:173-180
        K key = keys[j];
        K hp = bloom_hashpart(key, b.k);
        uint64_t h0 = bloom_racine(hp, b) + bloom_prefval(key, b.k);
        atomicOr(b.words + (h0 >> 5), 1u << (h0 & 31));
#pragma unroll
        for (int i = 1; i < 7; i++) {
            uint64_t h1 = h0 + (kmer_simplehash16(hp, i, tbl) & 4095);
            atomicOr(b.words + (h1 >> 5), 1u << (h1 & 31));
        }
    }
}

// N GPUs: every rank inserted the solid k-mers it owns into a zeroed array of the full size; the filter is the OR of
// the N arrays (all-gathered over NVLink into parts[rank][nwords]).
__global__ void __launch_bounds__(256) k_or_reduce(const uint32_t *__restrict__ parts, int n, uint64_t nwords, uint32_t *__restrict__ out)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t v = 0;
        for (int r = 0; r < n; r++) v |= parts[(uint64_t)r * nwords + i];
        out[i] = v;
    }
}

// contains4 of COL/Bloom.hpp:645-798 executed by one warp: lane c*7+i tests hash i of candidate nucleotide c, so the
// 28 probes of a step are ONE warp-wide gather inside the <= 514-byte window [racine, racine+13+4095].
// All lanes must call it with the same `kmer`; returns the 4-bit mask (bit nt = neighbour nt present).
template <typename K> LB_D unsigned warp_contains4(const BloomDev &b, K kmer, bool right, const uint64_t *tbl, int lane)
{
    const int k = b.k;
    K elem = right ? ((kmer << 2) & kmer_mask(K(0), k)) : (kmer >> 2);
    K hp = bloom_hashpart(elem, k);
    uint64_t racine = bloom_racine(hp, b);
    int c = lane / 7, i = lane - c * 7;
    unsigned first = right ? (unsigned)((elem >> (2 * (k - 1))) & 3) : (unsigned)c;
    unsigned last = right ? (unsigned)c : (unsigned)(elem & 3);
    uint64_t h = racine + bloom_cano2((first << 2) + last);
    if (i) h += kmer_simplehash16(hp, i, tbl) & 4095;
    unsigned set = 1;
    if (lane < 28) set = (__ldg(b.words + (h >> 5)) >> (h & 31)) & 1u;
    unsigned bal = __ballot_sync(0xffffffffu, set);
    unsigned res = 0;
#pragma unroll
    for (int nt = 0; nt < 4; nt++)
        if (((bal >> (7 * nt)) & 0x7fu) == 0x7fu) res |= 1u << nt;
    return res;
}

// test hook: one warp per query k-mer
template <typename K>
__global__ void __launch_bounds__(256) k_debug_query(const uint64_t *__restrict__ kmers, uint64_t n, int right, BloomDev b,
                                                     const uint64_t *__restrict__ tbl_g, uint8_t *out4, uint8_t *out1)
{
    __shared__ uint64_t tbl[256];
    tbl[threadIdx.x] = tbl_g[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t q = warp; q < n; q += nwarps) {
        K x;
        kmer_make(x, kmers[2 * q], kmers[2 * q + 1]);
        unsigned r4 = warp_contains4<K>(b, x, right != 0, tbl, lane);
        if (lane == 0) {
            out4[q] = (uint8_t)r4;
            out1[q] = bloom_contains<K>(b, kmer_canon(x, b.k), tbl) ? 1 : 0;
        }
    }
}

// test hook: raw hash values of the Bloom addressing for each query k-mer (12 u64 per k-mer:
// canon lo/hi, hashpart lo/hi, racine, prefval, simplehash16(hashpart, 1..6))
template <typename K>
__global__ void k_debug_hashes(const uint64_t *__restrict__ kmers, uint64_t n, BloomDev b, const uint64_t *__restrict__ tbl,
                               uint64_t *out)
{
    uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    K x;
    kmer_make(x, kmers[2 * q], kmers[2 * q + 1]);
    K c = kmer_canon(x, b.k);
    K hp = bloom_hashpart(c, b.k);
    uint64_t *o = out + 12 * q;
    o[0] = kmer_low64(c); o[1] = kmer_high64(c);
    o[2] = kmer_low64(hp); o[3] = kmer_high64(hp);
    o[4] = bloom_racine(hp, b);
    o[5] = bloom_prefval(c, b.k);
    for (int i = 1; i < 7; i++) o[5 + i] = kmer_simplehash16(hp, i, tbl);
}

}  // namespace lbbloom
