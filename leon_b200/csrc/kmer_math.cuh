// k-mer integer arithmetic shared by all kernels: 2-bit codes (A0 C1 T2 G3), reverse complement, the two Bloom
// hash families and the neighbour-coherent Bloom addressing.  Bit-exact with the reference functions cited on
// each item (paths relative to gatb-core/src/gatb/: M/ = tools/math/, COL/ = tools/collections/impl/).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

#define LB_HD __host__ __device__ __forceinline__
#define LB_D __device__ __forceinline__

// 128-bit k-mer word for 32 <= k <= 63.  An explicit two-word struct: nvcc 12.9's device code for variable shifts
// of `unsigned __int128` dropped a bit across the word boundary (seen on sm_100a in revcomp), so the compiler's
// __int128 is not used in device code at all.
struct u128 {
    uint64_t lo, hi;
    u128() = default;
    LB_HD explicit u128(uint64_t l) : lo(l), hi(0) {}
    LB_HD u128(uint64_t l, uint64_t h) : lo(l), hi(h) {}
    LB_HD explicit operator uint64_t() const { return lo; }
    LB_HD explicit operator unsigned() const { return (unsigned)lo; }
    LB_HD explicit operator bool() const { return (lo | hi) != 0; }
};
LB_HD u128 operator<<(u128 x, int s)
{
    if (s == 0) return x;
    if (s >= 128) return u128(0, 0);
    if (s >= 64) return u128(0, x.lo << (s - 64));
    return u128(x.lo << s, (x.hi << s) | (x.lo >> (64 - s)));
}
LB_HD u128 operator>>(u128 x, int s)
{
    if (s == 0) return x;
    if (s >= 128) return u128(0, 0);
    if (s >= 64) return u128(x.hi >> (s - 64), 0);
    return u128((x.lo >> s) | (x.hi << (64 - s)), x.hi >> s);
}
LB_HD u128 operator&(u128 a, u128 b) { return u128(a.lo & b.lo, a.hi & b.hi); }
LB_HD u128 operator|(u128 a, u128 b) { return u128(a.lo | b.lo, a.hi | b.hi); }
LB_HD u128 operator^(u128 a, u128 b) { return u128(a.lo ^ b.lo, a.hi ^ b.hi); }
LB_HD u128 operator&(u128 a, uint64_t b) { return u128(a.lo & b, 0); }
LB_HD u128 operator~(u128 a) { return u128(~a.lo, ~a.hi); }
LB_HD bool operator==(u128 a, u128 b) { return a.lo == b.lo && a.hi == b.hi; }
LB_HD bool operator!=(u128 a, u128 b) { return a.lo != b.lo || a.hi != b.hi; }
LB_HD bool operator<(u128 a, u128 b) { return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo); }
LB_HD u128 &operator>>=(u128 &a, int s) { a = a >> s; return a; }

// KW = number of 64-bit words of a k-mer: 1 for k <= 31 (reference LargeInt<1>), 2 for 32 <= k <= 63 (LargeInt<2>).
template <int KW> struct KmerOf;
template <> struct KmerOf<1> { typedef uint64_t type; };
template <> struct KmerOf<2> { typedef u128 type; };

LB_HD uint64_t rc64(uint64_t x)  // M/LargeInt1.pri:137-154: reverse the 32 fields, complement = ^2
{
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
    x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
    x = (x >> 32) | (x << 32);
    return x ^ 0xAAAAAAAAAAAAAAAAULL;
}

LB_HD uint64_t kmer_revcomp(uint64_t x, int k) { return rc64(x) >> (2 * (32 - k)); }
LB_HD u128 kmer_revcomp(u128 x, int k)  // M/LargeInt.hpp:722-734
{
    u128 r(rc64(x.hi), rc64(x.lo));
    return r >> (2 * (64 - k));
}
LB_HD uint64_t kmer_mask(uint64_t, int k) { return k >= 32 ? ~0ULL : ((1ULL << (2 * k)) - 1); }
LB_HD u128 kmer_mask(u128, int k)
{
    if (k >= 64) return u128(~0ULL, ~0ULL);
    if (k >= 32) return u128(~0ULL, k == 32 ? 0ULL : ((1ULL << (2 * (k - 32))) - 1));
    return u128((1ULL << (2 * k)) - 1, 0);
}
template <typename K> LB_HD K kmer_canon(K x, int k) { K r = kmer_revcomp(x, k); return r < x ? r : x; }
LB_HD uint64_t kmer_low64(uint64_t x) { return x; }
LB_HD uint64_t kmer_low64(u128 x) { return x.lo; }
LB_HD uint64_t kmer_fold64(uint64_t x) { return x; }
LB_HD uint64_t kmer_fold64(u128 x) { return x.lo ^ x.hi; }
LB_HD uint64_t kmer_high64(uint64_t) { return 0; }
LB_HD uint64_t kmer_high64(u128 x) { return x.hi; }
LB_HD void kmer_make(uint64_t &out, uint64_t lo, uint64_t) { out = lo; }
LB_HD void kmer_make(u128 &out, uint64_t lo, uint64_t hi) { out = u128(lo, hi); }

// "no k-mer here" marker: never a canonical k-mer (a canonical k-mer has at least its top two bits below 0b11).
template <typename K> LB_HD K kmer_sentinel() { return ~(K)0; }
// Windows with an invalid character are written as INVALID keys: top bit set (a k-mer of k <= 31 / k <= 63 bases stays
// below 2^62 / 2^126) over pseudo-random low bits, so that the hash partition spreads them over all buckets instead of
// piling every one of them into the sentinel's bucket (at 0.1 % N bases that bucket held 3.5 % of all keys).
LB_HD bool kmer_invalid(uint64_t x) { return (x >> 63) != 0; }
LB_HD bool kmer_invalid(const u128 &x) { return (x.hi >> 63) != 0; }
LB_HD void kmer_make_invalid(uint64_t &out, uint64_t salt) { out = (1ull << 63) | (salt * 0x9E3779B97F4A7C15ULL >> 1); }
LB_HD void kmer_make_invalid(u128 &out, uint64_t salt) { out = u128(salt * 0x9E3779B97F4A7C15ULL, (1ull << 63) | (salt >> 1)); }

LB_HD uint64_t hash64(uint64_t key, uint64_t seed)  // M/LargeInt1.pri:157-170
{
    uint64_t hash = seed;
    hash ^= (hash << 7) ^ key * (hash >> 3) ^ (~((hash << 11) + (key ^ (hash >> 5))));
    hash = (~hash) + (hash << 21);
    hash = hash ^ (hash >> 24);
    hash = (hash + (hash << 3)) + (hash << 8);
    hash = hash ^ (hash >> 14);
    hash = (hash + (hash << 2)) + (hash << 4);
    hash = hash ^ (hash >> 28);
    hash = hash + (hash << 31);
    return hash;
}
LB_HD uint64_t kmer_hash1(uint64_t x, uint64_t seed) { return hash64(x, seed); }
LB_HD uint64_t kmer_hash1(u128 x, uint64_t seed)  // M/LargeInt.hpp:738-748
{
    return hash64(x.lo, seed) ^ hash64(x.hi, seed);
}

// simplehash16: one word -> three table XORs (M/LargeInt1.pri:190-211); two words -> two XORs on the low word
// (M/LargeInt.hpp:792-801 -> M/NativeInt64.hpp:211-221).  tbl = the 256-entry table (shared or global memory).
LB_HD uint64_t kmer_simplehash16(uint64_t x, int shift, const uint64_t *tbl)
{
    uint64_t input = x >> shift;
    uint64_t res = tbl[input & 255];
    input >>= 8;
    res ^= tbl[input & 255];
    res ^= tbl[x & 255];
    return res;
}
LB_HD uint64_t kmer_simplehash16(u128 x, int shift, const uint64_t *tbl)
{
    uint64_t input = x.lo >> shift;
    uint64_t res = tbl[input & 255];
    input >>= 8;
    res ^= tbl[input & 255];
    return res;
}

// seed_tab[0] of HashFunctors (COL/Bloom.hpp:83-95): rbase[0] * rbase[3] mod 2^64
#define LB_BLOOM_SEED0 0xffaa54ffe6e6e6e7ULL

// BloomNeighborCoherent geometry (COL/Bloom.hpp:184-199, 437-442, 523-552): 7 hashes, 4096-bit block.
struct BloomDev {
    uint32_t *words;        // bit h lives in byte h>>3, mask 1<<(h&7) == word h>>5, bit h&31 (little endian)
    uint64_t reduced_tai;   // modulus of the root hash ("size" attribute)
    uint64_t magic;         // floor((2^64-1) / reduced_tai) for the multiply-high modulo
    uint64_t nbytes;        // 1 + (reduced_tai + 8192 [+1 if power of two]) / 8
    int k;
};

// One filter word of a probe whose result is consumed later: a volatile asm keeps the loads of a probe (and of the
// probes a caller wants in flight together) in program order ahead of their uses.  With plain __ldg the compiler sank
// each pair of loads down to its funnel shift in k_walk -- 5 dependent round trips per step instead of 1 -- which cost
// 3x at configs[3] scale, where every probe goes to DRAM (profiles/r1g_c4_walk_ncu_metrics.json).
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t bloom_ldw(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint2 bloom_ldw2(const uint2 *p)      // 8-byte aligned
{
    uint2 v;
    asm volatile("ld.global.nc.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
#endif

LB_HD uint64_t umulhi64(uint64_t a, uint64_t b)
{
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

// x mod d with the precomputed magic: the estimate is never above q and at most 2 below it.
LB_HD uint64_t fastmod64(uint64_t x, uint64_t d, uint64_t magic)
{
    uint64_t q = umulhi64(x, magic);
    uint64_t r = x - q * d;
    while (r >= d) r -= d;      // the estimate is at most 2 short
    return r;
}

// cano2 of COL/Bloom.hpp:526-541 packed 4 bits per entry (index = first_nt*4 + last_nt)
LB_HD unsigned bloom_cano2(unsigned idx)
{
    // entries: 0,1,2,3,4,5,3,7,8,9,0,4,9,13,1,5
    const uint64_t packed = 0x51D9409873543210ULL;
    return (unsigned)((packed >> (4 * (idx & 15))) & 15);
}

template <typename K> LB_HD K bloom_hashpart(K item, int k)  // middle k-2 nucleotides, canonical
{
    K hp = (item >> 2) & kmer_mask(K(0), k - 2);
    K rev = kmer_revcomp(hp, k - 2);
    return rev < hp ? rev : hp;
}
template <typename K> LB_HD unsigned bloom_prefval(K item, int k)
{
    unsigned suffix = (unsigned)(item & 3);
    unsigned prefix = (unsigned)((item >> (2 * (k - 1))) & 3);
    return bloom_cano2((prefix << 2) + suffix);
}
template <typename K> LB_HD uint64_t bloom_racine(K hashpart, const BloomDev &b)
{
    return fastmod64(kmer_hash1(hashpart, LB_BLOOM_SEED0), b.reduced_tai, b.magic);
}

// contains() of COL/Bloom.hpp:597-642, one thread per k-mer.
template <typename K> LB_D bool bloom_contains(const BloomDev &b, K item, const uint64_t *tbl)
{
    K hp = bloom_hashpart(item, b.k);
    uint64_t h0 = bloom_racine(hp, b) + bloom_prefval(item, b.k);
    if (!((__ldg(b.words + (h0 >> 5)) >> (h0 & 31)) & 1)) return false;
#pragma unroll
    for (int i = 1; i < 7; i++) {
        uint64_t h1 = h0 + (kmer_simplehash16(hp, i, tbl) & 4095);
        if (!((__ldg(b.words + (h1 >> 5)) >> (h1 & 31)) & 1)) return false;
    }
    return true;
}

// contains() for a k-mer given as its forward and reverse-complement words (rolled by the caller), in two halves so
// that a caller can have the probes of several k-mers in flight: the canonical middle k-2 nucleotides are the smaller
// of the two middles; all 7 probe words are loaded (no early exit), their bit positions packed 5 bits each in `sh`.
template <typename K> LB_D void bloom_probe_issue(const BloomDev &b, K fwd, K rc, const uint64_t *tbl, uint32_t (&w)[7], uint64_t &sh)
{
    const K m2 = kmer_mask(K(0), b.k - 2);
    const K mf = (fwd >> 2) & m2, mr = (rc >> 2) & m2;
    const K hp = mr < mf ? mr : mf;
    const K item = rc < fwd ? rc : fwd;
    // branch-free modulo (the estimate is at most 2 short) so that the chains of several k-mers can be interleaved
    const uint64_t x = kmer_hash1(hp, LB_BLOOM_SEED0);
    uint64_t rr = x - umulhi64(x, b.magic) * b.reduced_tai;
    rr -= rr >= b.reduced_tai ? b.reduced_tai : 0;
    rr -= rr >= b.reduced_tai ? b.reduced_tai : 0;
    const uint64_t h0 = rr + bloom_prefval(item, b.k);
    w[0] = bloom_ldw(b.words + (h0 >> 5));
    sh = h0 & 31u;
#pragma unroll
    for (int i = 1; i < 7; i++) {
        const uint64_t h1 = h0 + (kmer_simplehash16(hp, i, tbl) & 4095);
        w[i] = bloom_ldw(b.words + (h1 >> 5));
        sh |= (h1 & 31u) << (5 * i);
    }
}
LB_D bool bloom_probe_test(const uint32_t (&w)[7], uint64_t sh)
{
    uint32_t all = 1u;
#pragma unroll
    for (int i = 0; i < 7; i++) all &= w[i] >> (unsigned)((sh >> (5 * i)) & 31u);
    return (all & 1u) != 0;
}

// 2-bit code of an ASCII base as the k-mer models see it (MI/api/Data.hpp:176); 'N' is coded as 'A' because the
// encoder rewrites N to A before building k-mers (C/DnaCoder.cpp:519-528).
LB_HD unsigned base_code_n_as_a(uint8_t c) { return c == 'N' ? 0u : ((c >> 1) & 3u); }
LB_HD unsigned base_code(uint8_t c) { return (c >> 1) & 3u; }
LB_HD unsigned base_invalid(uint8_t c) { return (c >> 3) & 1u; }
// Leon::nt2bin (C/Leon.cpp:44-56): A0 C1 T2 G3 N4, anything else 0
LB_HD unsigned leon_nt2bin(uint8_t c)
{
    return c == 'C' ? 1u : c == 'T' ? 2u : c == 'G' ? 3u : c == 'N' ? 4u : 0u;
}
