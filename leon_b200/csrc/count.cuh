// Phase 1 — solid k-mer counting in HBM (replaces DSK as used by Leon, SURVEY.md 8a-1;
// K/SortingCountAlgorithm.cpp:499-644, K/Model.hpp:725-764,876-883, K/PartitionsCommand.cpp:1206-1460).
// Only the multiset of canonical k-mers matters to Leon (histogram, solid list), not their order, so instead of a
// full sort the keys are PARTITIONED by a mixing hash and each bucket is counted in a shared-memory hash table:
//   k_extract      : one warp per read, canonical k-mer of every window free of invalid characters
//   k_part_hist /
//   k_part_scatter : stable 8-bit partition pass (tile histogram -> digit-major scan -> ranks -> shared-memory
//                    staging -> coalesced runs); 1-3 passes on hash digits give buckets of ~5 k keys.  The same
//                    two kernels, with a plain shift digit, are the radix sort of entropy.cuh.
//   k_bucket_bounds: first key of every bucket
//   k_bucket_count : one CTA per bucket; open-addressing table in shared memory (CAS on the key, atomicAdd on the
//                    count); abundance histogram (u16 truncation + clamp, MI/impl/Histogram.hpp:87), solid list
//                    (abundance >= 2, K/SortingCountAlgorithm.cpp:210).  A bucket whose distinct keys overflow the
//                    table is split by further hash bits and redone, so any input is handled.
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
// Stages the bases of reads [r0, r1) of a CTA in shared memory with one bulk asynchronous copy (TMA, 1-D), so that
// the per-thread walks read their bases from shared memory.  Returns the shared-memory address of bases[offsets[r0]],
// or nullptr when the span does not fit (very long reads): callers then read from global memory.
LB_D const uint8_t *stage_reads(const uint8_t *__restrict__ bases, uint64_t o0, uint64_t o1, uint8_t *smem, uint32_t smem_bytes,
                                uint64_t *bar)
{
    const uint8_t *src = bases + (o0 & ~15ull);
    uint64_t bytes = ((o1 - (o0 & ~15ull)) + 15ull) & ~15ull;
    if (bytes > smem_bytes) return nullptr;        // CTA-uniform
    if (bytes == 0) return smem;
    uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t dst_a = (uint32_t)__cvta_generic_to_shared(smem);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"((uint32_t)bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst_a), "l"(src), "r"((uint32_t)bytes), "r"(bar_a) : "memory");
    }
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar_a) : "memory");
    }
    return smem + (o0 & 15ull);
}

// The same in two halves, for callers with work to overlap with the copy: begin() returns nullptr (and starts nothing)
// when the span does not fit, else the address as above; wait() must then be called by every thread before the bases are
// read (and before the next __syncthreads-separated use of the barrier).
LB_D const uint8_t *stage_reads_begin(const uint8_t *__restrict__ bases, uint64_t o0, uint64_t o1, uint8_t *smem, uint32_t smem_bytes,
                                      uint64_t *bar)
{
    const uint8_t *src = bases + (o0 & ~15ull);
    uint64_t bytes = ((o1 - (o0 & ~15ull)) + 15ull) & ~15ull;
    if (bytes > smem_bytes) return nullptr;        // CTA-uniform
    uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t dst_a = (uint32_t)__cvta_generic_to_shared(smem);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"((uint32_t)bytes) : "memory");
        if (bytes)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst_a), "l"(src), "r"((uint32_t)bytes), "r"(bar_a) : "memory");
    }
    return smem + (o0 & 15ull);
}
LB_D void stage_reads_wait(uint64_t *bar)
{
    uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar_a) : "memory");
    }
}

// Canonical k-mer of every window of every read; an INVALID key when the window holds an invalid character ((c>>3)&1,
// MI/api/Data.hpp:176).  One THREAD per read: the bases of the 256 reads of a CTA are staged in shared memory by one
// bulk copy, the forward k-mer and its reverse complement are rolled base by base (4 operations each, K/Model.hpp:725-764)
// and the canonical one is their minimum (:876-883).  Only the multiset of keys matters downstream, so when all reads
// of the CTA have the same length the keys of the tile are written position-major (key of read t, window p at
// tile_base + p * reads + t: every store instruction of a warp is one contiguous run); otherwise at koff[r] + p.
template <typename K>
__global__ void __launch_bounds__(256) k_extract(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                 const uint64_t *__restrict__ koff, uint64_t nreads, int k, K *__restrict__ keys,
                                                 uint32_t stage_bytes)
{
    extern __shared__ __align__(128) uint8_t stage[];
    __shared__ __align__(8) uint64_t bar;
    const uint64_t r0 = (uint64_t)blockIdx.x * 256;
    const uint64_t r1 = r0 + 256 < nreads ? r0 + 256 : nreads;
    const uint64_t o0 = offsets[r0];
    const uint8_t *sbase = stage_reads(bases, o0, offsets[r1], stage, stage_bytes, &bar);
    const uint64_t r = r0 + threadIdx.x;
    const bool have = r < nreads;
    const uint64_t o = have ? offsets[r] : 0, L = have ? offsets[r + 1] - o : 0;
    const uint64_t L0 = offsets[r0 + 1] - o0;
    const int uniform = __syncthreads_and(!have || L == L0);      // also orders the staged bytes before their use
    if (!have || L < (uint64_t)k) return;
    const uint8_t *s = sbase ? sbase + (o - o0) : bases + o;
    const uint64_t n = L - k + 1;
    const uint64_t stride = uniform ? (r1 - r0) : 1;
    K *out = uniform ? keys + koff[r0] + threadIdx.x : keys + koff[r];
    const K mask = kmer_mask(K(0), k);
    const int top = 2 * (k - 1);
    K fwd = K(0), rc = K(0);
    uint32_t clean = 0;                    // valid characters since the last invalid one (saturates at k)
    for (uint64_t i = 0; i < L; i++) {
        const uint8_t c = s[i];
        const unsigned code = base_code(c);
        fwd = ((fwd << 2) | K((uint64_t)code)) & mask;
        rc = (rc >> 2) | (K((uint64_t)(code ^ 2u)) << top);
        clean = base_invalid(c) ? 0u : (clean < (uint32_t)k ? clean + 1 : clean);
        if (i + 1 >= (uint64_t)k) {
            const uint64_t p = i + 1 - k;
            K key;
            if (clean >= (uint32_t)k) key = rc < fwd ? rc : fwd;
            else kmer_make_invalid(key, (r << 24) ^ p);
            out[p * stride] = key;
        }
    }
    (void)n;
}

// The same for read sets of VARIABLE length (the host knows: min != max).  koff[r] + p is then the only layout there is, and a
// thread that stores its own read's keys one by one makes every store instruction of its warp touch 32 lines (measured on
// configs[3], 2 % trimmed reads so that hardly any CTA is uniform: 0.69 TB/s against 4.4 TB/s of the position-major stores
// above).  Here every warp transposes through shared memory: XT positions of its 32 reads are rolled into a tile
// [32 reads][XT + 1], then each group of XT lanes writes one read's XT consecutive keys (128 contiguous bytes per group).
template <typename K> struct ExtractRagged { static constexpr int XT = sizeof(K) == 8 ? 16 : 8; };
template <typename K> LB_HD constexpr uint32_t extract_ragged_tile_bytes() { return 8u * 32u * (ExtractRagged<K>::XT + 1) * (uint32_t)sizeof(K); }

template <typename K>
__global__ void __launch_bounds__(256) k_extract_ragged(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                        const uint64_t *__restrict__ koff, uint64_t nreads, int k, K *__restrict__ keys,
                                                        uint32_t stage_bytes)
{
    constexpr int XT = ExtractRagged<K>::XT;
    extern __shared__ __align__(128) uint8_t stage[];          // [tiles of the 8 warps][staged bases]
    __shared__ __align__(8) uint64_t bar;
    K *tile = reinterpret_cast<K *>(stage) + (size_t)(threadIdx.x >> 5) * 32 * (XT + 1);
    uint8_t *sbytes = stage + extract_ragged_tile_bytes<K>();
    const uint64_t r0 = (uint64_t)blockIdx.x * 256;
    const uint64_t r1 = r0 + 256 < nreads ? r0 + 256 : nreads;
    const uint64_t o0 = offsets[r0];
    const uint8_t *sbase = stage_reads(bases, o0, offsets[r1], sbytes, stage_bytes, &bar);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint64_t r = r0 + threadIdx.x;
    const bool have = r < nreads;
    const uint64_t o = have ? offsets[r] : 0, L = have ? offsets[r + 1] - o : 0;
    const uint64_t n = L >= (uint64_t)k ? L - k + 1 : 0;
    const uint64_t kbase = have ? koff[r] : 0;
    // the warp walks to the end of its longest read (reads of 2^32 bases and more do not exist here: leon_upload checks)
    const uint32_t nmax = __reduce_max_sync(0xffffffffu, (uint32_t)n);
    if (!nmax) return;
    const uint8_t *s = sbase ? sbase + (o - o0) : bases + o;
    const K mask = kmer_mask(K(0), k);
    const int top = 2 * (k - 1);
    K fwd = K(0), rc = K(0);
    uint32_t clean = 0;
    const uint32_t steps = nmax + (uint32_t)k - 1;
    const int row0 = lane / XT, j = lane % XT;
    for (uint32_t i = 0; i < steps; i++) {
        if (i < L) {
            const uint8_t c = s[i];
            const unsigned code = base_code(c);
            fwd = ((fwd << 2) | K((uint64_t)code)) & mask;
            rc = (rc >> 2) | (K((uint64_t)(code ^ 2u)) << top);
            clean = base_invalid(c) ? 0u : (clean < (uint32_t)k ? clean + 1 : clean);
            if (i + 1 >= (uint32_t)k) {
                const uint32_t p = i + 1 - k;
                K key;
                if (clean >= (uint32_t)k) key = rc < fwd ? rc : fwd;
                else kmer_make_invalid(key, (r << 24) ^ (uint64_t)p);
                tile[lane * (XT + 1) + (p % XT)] = key;
            }
        }
        if (i + 1 >= (uint32_t)k) {
            const uint32_t p = i + 1 - k;                      // warp-uniform
            if (p % XT == XT - 1 || i + 1 == steps) {
                const uint32_t pbase = p - p % XT;
                __syncwarp();
#pragma unroll 4
                for (int row = row0; row < 32; row += 32 / XT) {
                    const uint64_t n_row = __shfl_sync(0xffffffffu, n, row);
                    const uint64_t k_row = __shfl_sync(0xffffffffu, kbase, row);
                    if ((uint64_t)pbase + j < n_row) keys[k_row + pbase + j] = tile[row * (XT + 1) + j];
                }
                __syncwarp();
            }
        }
    }
}

// ---- partition pass (also the LSD radix sort pass of entropy.cuh) ------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
template <typename K> struct RsCfg { static constexpr int ITEMS = sizeof(K) == 8 ? 16 : 8; static constexpr int TILE = RS_THREADS * ITEMS; };

// murmur3 finaliser: every output bit depends on every input bit
LB_HD uint64_t mix64(uint64_t x)
{
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33;
    return x;
}
LB_HD uint64_t key_mix(uint64_t k) { return mix64(k); }
LB_HD uint64_t key_mix(u128 k) { return mix64(k.lo ^ mix64(k.hi + 0x9E3779B97F4A7C15ULL)); }

template <typename K> struct ShiftDigit {     // radix sort on key bits [shift, shift + 8)
    int shift;
    LB_D unsigned operator()(K key) const { return (unsigned)(key >> shift) & 255u; }
};
template <typename K> struct HashDigit {      // partition on hash bits [shift, shift + 8) & mask
    int shift; unsigned mask;
    LB_D unsigned operator()(K key) const { return (unsigned)(key_mix(key) >> shift) & mask; }
};

// tile_hist[d * ntiles + tile] = number of keys of the tile whose digit is d
template <typename K, typename D>
__global__ void __launch_bounds__(RS_THREADS) k_part_hist(const K *__restrict__ keys, uint64_t n, D digit,
                                                          uint32_t *__restrict__ tile_hist, uint32_t ntiles)
{
    constexpr int ITEMS = RsCfg<K>::ITEMS, TILE = RsCfg<K>::TILE;
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    uint64_t base = (uint64_t)blockIdx.x * TILE;
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = base + (uint64_t)j * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[digit(keys[i])], 1u);
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

// Stable scatter.  Rank of a key among the keys of its digit = (tile offset of the digit) + (keys of earlier warps of
// the tile with that digit) + (earlier keys of this warp with that digit); warp w owns keys [w*32*ITEMS, ...) of the
// tile, row by row, so ranks follow the input order.  Keys are first placed at their tile-local sorted position in
// shared memory, then written out: consecutive threads write consecutive addresses of a digit's run.
template <typename K, typename D>
__global__ void __launch_bounds__(RS_THREADS) k_part_scatter(const K *__restrict__ in, K *__restrict__ out, uint64_t n, D digit,
                                                             const uint64_t *__restrict__ tile_off, uint32_t ntiles)
{
    constexpr int ITEMS = RsCfg<K>::ITEMS, TILE = RsCfg<K>::TILE;
    __shared__ uint32_t whist[RS_WARPS][256];
    __shared__ uint32_t dstart[256];
    __shared__ uint64_t gbase[256];
    __shared__ uint32_t wsum[RS_WARPS];
    __shared__ K stage[TILE];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();

    K key[ITEMS];
    uint16_t rank[ITEMS], dig[ITEMS];
    const uint64_t tbase = (uint64_t)blockIdx.x * TILE;
    const uint64_t wstart = tbase + (uint64_t)warp * (32 * ITEMS);
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        uint64_t i = wstart + (uint64_t)r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? in[i] : kmer_sentinel<K>();
        unsigned d = valid ? digit(key[r]) : 256u;   // 256 groups the out-of-range lanes together
        uint32_t peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader && valid) {
            old = whist[warp][d];
            whist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[r] = (uint16_t)(old + __popc(peers & lt_mask));
        dig[r] = (uint16_t)d;
        __syncwarp();
    }
    __syncthreads();
    {
        // thread d: per-warp counts of digit d -> exclusive prefix over the warps; digit totals -> tile-local starts
        const unsigned d = threadIdx.x;
        uint32_t tot = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            uint32_t c = whist[w][d];
            whist[w][d] = tot;
            tot += c;
        }
        uint32_t incl = tot;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, incl, s);
            if (lane >= s) incl += o;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        uint32_t base = 0;
        for (int w = 0; w < warp; w++) base += wsum[w];
        dstart[d] = base + incl - tot;
        gbase[d] = tile_off[(uint64_t)d * ntiles + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < ITEMS; r++)
        if (dig[r] < 256) stage[dstart[dig[r]] + whist[warp][dig[r]] + rank[r]] = key[r];
    __syncthreads();
    const uint32_t nt = n - tbase < (uint64_t)TILE ? (uint32_t)(n - tbase) : (uint32_t)TILE;
    for (uint32_t i = threadIdx.x; i < nt; i += RS_THREADS) {
        K kx = stage[i];
        unsigned d = digit(kx);
        out[gbase[d] + (i - dstart[d])] = kx;
    }
}

// ---- hash partition of the k-mer keys (MSD, not stable) ---------------------------------------------------------------
// Counting only needs every bucket's MULTISET of keys, so the partition on the top `bits` bits of key_mix does not have
// to be stable: most significant digit first, one level per 8 bits.  A level subdivides each of the P partitions of
// the previous one; tiles are aligned to partition starts (tstart[p] = first tile of partition p), a CTA finds its
// partition by binary search.  k_msd_hist counts (partition, digit) pairs (shared-memory histogram, then one global
// atomicAdd per non-empty digit); their exclusive scan is both the partition table of the next level and, after the
// last level, the bucket boundaries bstart.  k_msd_scatter ranks the keys of a tile inside their digit with
// shared-memory atomics (the returned old value is the rank), reserves the tile's run of every digit with one global
// atomicAdd on that digit's cursor, orders the tile by digit in shared memory and writes the runs out coalesced.
struct MsdLevel {
    const uint64_t *pstart;   // [P + 1] key index of the first key of every partition
    const uint64_t *tstart;   // [P + 1] first tile of every partition
    uint64_t P;
    int shift;                // digit = (key_mix >> shift) & mask
    unsigned mask;
    int wbits;
};

// partition of tile c (the last p with tstart[p] <= c); false when c is past the last tile
LB_D bool msd_locate(const MsdLevel &lv, uint64_t c, uint64_t tile, uint64_t &p, uint64_t &beg, uint64_t &end)
{
    if (c >= lv.tstart[lv.P]) return false;
    uint64_t lo = 0, hi = lv.P;      // invariant: tstart[lo] <= c < tstart[hi]
    while (hi - lo > 1) {
        uint64_t mid = (lo + hi) >> 1;
        if (lv.tstart[mid] <= c) lo = mid; else hi = mid;
    }
    p = lo;
    beg = lv.pstart[lo] + (c - lv.tstart[lo]) * tile;
    end = lv.pstart[lo + 1];
    if (end > beg + tile) end = beg + tile;
    return true;
}

struct MsdTilesIn {
    const uint64_t *pstart; uint64_t tile;
    __device__ uint64_t operator()(uint64_t i) const { return (pstart[i + 1] - pstart[i] + tile - 1) / tile; }
};
struct MsdTilesOut {
    uint64_t *tstart;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t) const { tstart[i] = ex; }
};
struct MsdCntIn {
    const unsigned long long *cnt;
    __device__ uint64_t operator()(uint64_t i) const { return cnt[i]; }
};
struct MsdCntOut {
    uint64_t *pstart; unsigned long long *cursor;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t) const { pstart[i] = ex; cursor[i] = ex; }
};

template <typename K>
__global__ void __launch_bounds__(RS_THREADS) k_msd_hist(const K *__restrict__ keys, const MsdLevel lv,
                                                         unsigned long long *__restrict__ cnt)
{
    constexpr int ITEMS = RsCfg<K>::ITEMS, TILE = RsCfg<K>::TILE;
    __shared__ uint32_t h[256];
    __shared__ uint64_t sp, sbeg, send;
    __shared__ int live;
    if (threadIdx.x == 0) {
        uint64_t p = 0, b = 0, e = 0;
        live = msd_locate(lv, blockIdx.x, TILE, p, b, e);
        sp = p; sbeg = b; send = e;
    }
    h[threadIdx.x] = 0;
    __syncthreads();
    if (!live) return;
    const uint64_t beg = sbeg, end = send;
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = beg + (uint64_t)j * RS_THREADS + threadIdx.x;
        if (i < end) atomicAdd(&h[(unsigned)(key_mix(keys[i]) >> lv.shift) & lv.mask], 1u);
    }
    __syncthreads();
    if (threadIdx.x <= lv.mask && h[threadIdx.x]) atomicAdd(&cnt[(sp << lv.wbits) + threadIdx.x], (unsigned long long)h[threadIdx.x]);
}

// THREADS = 256: 16 keys per thread (64 registers, 4 CTAs = 32 warps per SM); THREADS = 512: 8 keys per thread, 40 registers,
// 3 CTAs = 48 warps per SM over the same 4096-key tile (the kernel is latency-bound: three barrier-separated phases and one
// global atomic round trip per tile; ncu: dram 27 %, issue 27 %, occupancy 49 % with 256 threads).  Measured at configs[2]
// (2 x 322 M keys): 4.84 ms with 256 threads, 3.99 ms with 512, 4.47 ms with 1024 (2 CTAs, 32 registers; profiles/r3c_msd_threads.txt):
// 512 is the default.
template <typename K, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == 256 ? 4 : 3) k_msd_scatter(const K *__restrict__ in, K *__restrict__ out, const MsdLevel lv,
                                                                               unsigned long long *__restrict__ cursor)
{
    constexpr int TILE = RsCfg<K>::TILE, ITEMS = TILE / THREADS, WARPS = THREADS / 32;
    static_assert(WARPS >= 8, "the digit scan runs on the first 8 warps");
    __shared__ uint32_t h[256], dstart[256], wsum[8];
    __shared__ uint64_t gbase[256];
    __shared__ K stage[TILE];
    __shared__ uint8_t sdig[TILE];
    __shared__ uint64_t sp, sbeg, send;
    __shared__ int live;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        uint64_t p = 0, b = 0, e = 0;
        live = msd_locate(lv, blockIdx.x, TILE, p, b, e);
        sp = p; sbeg = b; send = e;
    }
    if (threadIdx.x < 256) h[threadIdx.x] = 0;
    __syncthreads();
    if (!live) return;
    const uint64_t beg = sbeg, end = send;
    K key[ITEMS];
    uint16_t rank[ITEMS];
    uint8_t dig[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = beg + (uint64_t)j * THREADS + threadIdx.x;
        if (i < end) {
            key[j] = in[i];
            unsigned d = (unsigned)(key_mix(key[j]) >> lv.shift) & lv.mask;
            dig[j] = (uint8_t)d;
            rank[j] = (uint16_t)atomicAdd(&h[d], 1u);
        }
    }
    __syncthreads();
    {
        const unsigned d = threadIdx.x & 255u;
        const bool scanner = threadIdx.x < 256;
        const uint32_t tot = scanner ? h[d] : 0u;
        uint32_t incl = tot;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, incl, s);
            if (lane >= s) incl += o;
        }
        if (scanner && lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (scanner) {
            uint32_t base = 0;
            for (int w = 0; w < warp; w++) base += wsum[w];
            dstart[d] = base + incl - tot;
            gbase[d] = tot ? atomicAdd(&cursor[(sp << lv.wbits) + d], (unsigned long long)tot) : 0ull;
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = beg + (uint64_t)j * THREADS + threadIdx.x;
        if (i < end) {
            uint32_t pos = dstart[dig[j]] + rank[j];
            stage[pos] = key[j];
            sdig[pos] = dig[j];
        }
    }
    __syncthreads();
    const uint32_t nt = (uint32_t)(end - beg);
    for (uint32_t i = threadIdx.x; i < nt; i += THREADS) {
        unsigned d = sdig[i];
        out[gbase[d] + (i - dstart[d])] = stage[i];
    }
}

// ---- buckets ------------------------------------------------------------------------------------------------------
// After the partition passes the keys are ordered by bucket id = top `bits` bits of key_mix.
// bstart[b] = index of the first key of bucket b (b = 0 .. nbuckets, bstart[nbuckets] = n).
template <typename K>
__global__ void __launch_bounds__(256) k_bucket_bounds(const K *__restrict__ keys, uint64_t n, int bits, uint64_t nbuckets,
                                                       uint64_t *__restrict__ bstart)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t b = bits ? key_mix(keys[i]) >> (64 - bits) : 0;
        uint64_t first = i ? (bits ? key_mix(keys[i - 1]) >> (64 - bits) : 0) + 1 : 0;
        for (uint64_t x = first; x <= b; x++) bstart[x] = i;
        if (i == n - 1)
            for (uint64_t x = b + 1; x <= nbuckets; x++) bstart[x] = n;
    }
}

struct CountStats { unsigned long long nb_solid_total, nb_distinct, nb_valid; };

// Where the keys of a bucket are.  One GPU: one source, its own partitioned array.  N GPUs: rank p owns the buckets
// [lo, lo + gridDim.x) and received, from every rank s, that rank's keys of those buckets as one contiguous chunk
// (still ordered by bucket): the keys of bucket b that came from s are [base[s] + bstart[s][b], base[s] + bstart[s][b+1])
// with base[s] = (offset of the chunk in `keys`) - bstart[s][lo]  (unsigned wrap-around arithmetic).
constexpr int MAX_RANKS = 8;
struct BucketSrc {
    int nsrc;
    uint64_t lo;
    const uint64_t *bstart[MAX_RANKS];
    uint64_t base[MAX_RANKS];
};

constexpr int BC_THREADS = 256;
constexpr int BC_STACK = 48;
template <typename K> struct BcCfg {
    static constexpr int SLOTS = sizeof(K) == 8 ? 8192 : 4096;     // table slots per CTA (96 KB / 112 KB: 2 CTAs per SM)
    static constexpr int SLOTS_SMALL = SLOTS / 4;                  // first attempt: 24 / 28 KB, 8 CTAs per SM
    static constexpr int TARGET = SLOTS * 3 / 4;                   // keys per bucket the partition aims for
    static constexpr size_t slot_bytes = 12 + (sizeof(K) == 8 ? 0 : sizeof(K));
    static constexpr size_t SMEM = (size_t)SLOTS * slot_bytes;
    static constexpr size_t SMEM_SMALL = (size_t)SLOTS_SMALL * slot_bytes;
};

template <typename K> LB_D uint64_t table_tag(K key, uint64_t salt);
template <> LB_D uint64_t table_tag<uint64_t>(uint64_t key, uint64_t) { return key; }     // never the sentinel ~0
template <> LB_D uint64_t table_tag<u128>(u128 key, uint64_t salt)
{
    uint64_t t = mix64(key.lo + salt) ^ key.hi;
    return t == ~0ull ? 0x5bd1e995ull : t;
}

// One CTA per bucket.  `sel/nsplit`: a table pass admits the keys whose selector bits equal `sel`; it starts with
// the whole bucket (or ceil(size / BC_TARGET) parts) and halves a part whenever its distinct keys overflow the table.
// Two launches: at sequencing coverage a bucket of ~5 k keys holds a few hundred distinct ones, so the first launch
// gives every bucket a small table (slots = SLOTS_SMALL: a quarter of the clearing and scanning, four times the
// resident CTAs) and only lists the buckets whose distinct keys do not fit (retry != nullptr); the second launch
// (retry_list != nullptr, one CTA per listed bucket) redoes those with the full-size table and the splitting logic.
// 128-bit keys are claimed through a 64-bit tag (CAS) and verified against the full key after the pass; a tag shared
// by two different keys (probability ~2^-40 per run) makes the pass repeat with another salt.
template <typename K>
__global__ void __launch_bounds__(BC_THREADS) k_bucket_count(const K *__restrict__ keys, const BucketSrc src,
                                                             unsigned long long *__restrict__ histo,
                                                             CountStats *__restrict__ stats, K *__restrict__ solid_keys,
                                                             uint32_t *__restrict__ solid_cnt,
                                                             unsigned long long *__restrict__ nsolid_out, const int BC_SLOTS,
                                                             uint32_t *__restrict__ retry, uint32_t *__restrict__ retry_count,
                                                             const uint32_t *__restrict__ retry_list)
{
    constexpr int BC_TARGET = BcCfg<K>::TARGET;
    const int BC_FILL_MAX = BC_SLOTS * 3 / 4;                      // distinct keys admitted per table
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t *tags = reinterpret_cast<uint64_t *>(smem_raw);                       // [BC_SLOTS]
    uint32_t *cnts = reinterpret_cast<uint32_t *>(tags + BC_SLOTS);                 // [BC_SLOTS]
    K *full = reinterpret_cast<K *>(cnts + BC_SLOTS);                               // [BC_SLOTS], 128-bit keys only
    __shared__ uint32_t st_sel[BC_STACK], st_ns[BC_STACK];
    __shared__ int sp;
    __shared__ uint32_t ndistinct, overflow, collision, nsol;
    __shared__ unsigned long long sol_base;
    __shared__ unsigned long long acc[3];
    __shared__ uint32_t shist[256];
    __shared__ uint64_t seg_beg[MAX_RANKS], seg_end[MAX_RANKS];
    const uint64_t b = src.lo + (retry_list ? retry_list[blockIdx.x] : blockIdx.x);
    const int tid = threadIdx.x;
    if (tid < src.nsrc) {
        seg_beg[tid] = src.base[tid] + src.bstart[tid][b];
        seg_end[tid] = src.base[tid] + src.bstart[tid][b + 1];
    }
    __syncthreads();
    uint64_t size = 0;
    for (int sx = 0; sx < src.nsrc; sx++) size += seg_end[sx] - seg_beg[sx];
    if (size == 0) return;
    if (retry && size > (uint64_t)BC_TARGET) {      // would need several parts: straight to the full-size launch
        if (tid == 0) retry[atomicAdd(retry_count, 1u)] = (uint32_t)(b - src.lo);
        return;
    }
    if (tid == 0) {
        uint32_t ns = 1;
        while ((uint64_t)ns * BC_TARGET < size && ns < 32) ns <<= 1;
        sp = 0;
        for (uint32_t x = 0; x < ns; x++) { st_sel[sp] = x; st_ns[sp] = ns; sp++; }
        acc[0] = acc[1] = acc[2] = 0;
    }
    shist[tid] = 0;
    __syncthreads();
    uint64_t salt = 0x1234567;
    while (true) {
        __syncthreads();
        if (sp == 0) break;
        const uint32_t sel = st_sel[sp - 1], ns = st_ns[sp - 1];
        __syncthreads();
        if (tid == 0) { sp--; ndistinct = 0; overflow = 0; collision = 0; nsol = 0; }
        for (int i = tid; i < BC_SLOTS; i += BC_THREADS) { tags[i] = ~0ull; cnts[i] = 0; }
        __syncthreads();
        // insert
        for (int sx = 0; sx < src.nsrc; sx++)
        for (uint64_t i = seg_beg[sx] + tid; i < seg_end[sx]; i += BC_THREADS) {
            K key = keys[i];
            if (kmer_invalid(key)) continue;
            uint64_t mx = key_mix(key);
            if (((uint32_t)(mx >> 20) & (ns - 1)) != sel) continue;
            uint64_t tag = table_tag<K>(key, salt);
            uint32_t slot = (uint32_t)(mx ^ (mx >> 37)) & (BC_SLOTS - 1);
            for (int probe = 0; probe < BC_SLOTS; probe++) {
                // (a plain load of the tag before the compare-and-swap, for the keys that repeat one already in the table, was
                // measured: 2.10 -> 2.14 ms; the kernel is bound by shared-memory wavefronts, not by the atomic's latency)
                unsigned long long old = atomicCAS((unsigned long long *)&tags[slot], ~0ull, (unsigned long long)tag);
                if (old == ~0ull) {
                    if (sizeof(K) > 8) full[slot] = key;
                    if (atomicAdd(&ndistinct, 1u) >= BC_FILL_MAX) overflow = 1;
                    atomicAdd(&cnts[slot], 1u);
                    break;
                }
                if (old == tag) { atomicAdd(&cnts[slot], 1u); break; }
                slot = (slot + 1) & (BC_SLOTS - 1);
            }
            if (*(volatile uint32_t *)&overflow) break;
        }
        __syncthreads();
        if (overflow && retry) {   // first launch: leave this bucket to the launch with full-size tables (nothing emitted yet)
            if (tid == 0) retry[atomicAdd(retry_count, 1u)] = (uint32_t)(b - src.lo);
            return;
        }
        if (overflow) {   // too many distinct keys for one table: split this part in two and redo
            if (tid == 0) {
                if (sp + 2 <= BC_STACK && ns < (1u << 24)) {
                    st_sel[sp] = sel; st_ns[sp] = ns * 2; sp++;
                    st_sel[sp] = sel + ns; st_ns[sp] = ns * 2; sp++;
                } else atomicAdd(&stats->nb_valid, 1ull << 62);   // cannot happen for distinct keys; poisons the result
            }
            continue;
        }
        if (sizeof(K) > 8) {   // verify the tags of 128-bit keys
            for (int sx = 0; sx < src.nsrc; sx++)
            for (uint64_t i = seg_beg[sx] + tid; i < seg_end[sx]; i += BC_THREADS) {
                K key = keys[i];
                if (kmer_invalid(key)) continue;
                uint64_t mx = key_mix(key);
                if (((uint32_t)(mx >> 20) & (ns - 1)) != sel) continue;
                uint64_t tag = table_tag<K>(key, salt);
                uint32_t slot = (uint32_t)(mx ^ (mx >> 37)) & (BC_SLOTS - 1);
                while (tags[slot] != tag) slot = (slot + 1) & (BC_SLOTS - 1);
                if (full[slot] != key) collision = 1;
            }
            __syncthreads();
            if (collision) {
                if (tid == 0) { st_sel[sp] = sel; st_ns[sp] = ns; sp++; }
                salt = salt * 6364136223846793005ULL + 1442695040888963407ULL;
                continue;
            }
        }
        // emit: histogram over all distinct keys, solid list for abundance >= 2
        unsigned long long distinct = 0, valid = 0, solid = 0;
        for (int i = tid; i < BC_SLOTS; i += BC_THREADS) {
            uint32_t c = cnts[i];
            if (!c) continue;
            distinct++;
            valid += c;
            unsigned idx = c & 0xffffu;
            if (idx < 256u) atomicAdd(&shist[idx], 1u);
            else atomicAdd(&histo[idx >= 10000u ? 10000u : idx], 1ull);
            if (c >= 2) { solid++; atomicAdd(&nsol, 1u); }
        }
        for (int d = 16; d; d >>= 1) {
            distinct += __shfl_down_sync(0xffffffffu, distinct, d);
            valid += __shfl_down_sync(0xffffffffu, valid, d);
            solid += __shfl_down_sync(0xffffffffu, solid, d);
        }
        if ((tid & 31) == 0) { atomicAdd(&acc[0], solid); atomicAdd(&acc[1], distinct); atomicAdd(&acc[2], valid); }
        __syncthreads();
        if (tid == 0) { sol_base = nsol ? atomicAdd(nsolid_out, (unsigned long long)nsol) : 0; nsol = 0; }
        __syncthreads();
        for (int i = tid; i < BC_SLOTS; i += BC_THREADS) {
            uint32_t c = cnts[i];
            if (c < 2) continue;
            uint64_t o = sol_base + atomicAdd(&nsol, 1u);
            K key;
            if (sizeof(K) > 8) key = full[i];
            else kmer_make(key, tags[i], 0);
            solid_keys[o] = key;
            solid_cnt[o] = c;
        }
    }
    __syncthreads();
    if (shist[tid]) atomicAdd(&histo[tid], (unsigned long long)shist[tid]);
    if (tid == 0) {
        if (acc[0]) atomicAdd(&stats->nb_solid_total, acc[0]);
        if (acc[1]) atomicAdd(&stats->nb_distinct, acc[1]);
        if (acc[2]) atomicAdd(&stats->nb_valid, acc[2]);
    }
}

}  // namespace lbcount
