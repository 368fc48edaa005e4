// Phase 3a — anchor selection with the reference's sequential semantics, computed in parallel.
//
// Reference (one thread, reads in file order; C/DnaCoder.cpp:563-576 findExistingAnchor, C/Leon.cpp:1105-1243
// findAndInsertAnchor, COL/Hash16.hpp): read i takes the FIRST position whose canonical k-mer is already in the
// dictionary D; if none, the first Bloom-positive k-mer in the order [n/2, n/2+10) -> [0, n/2) -> [n/2+10, n) is
// inserted with address |D| and becomes the anchor; if none, the read has no anchor.
//
// Here reads are handled in ordered batches.  Every dictionary entry carries the index of the read that inserted it
// ("time"); read i may only see entries with time < i.  Per batch:
//   A  lookup against D (entries of earlier batches)            -> reads with a hit can never insert
//   P  the no-hit reads, in order, with their Bloom candidate c_j
//   R  resolve who inserts: j inserts  <=>  no earlier inserter l < j has c_l among the k-mers of read j.
//      Rounds of deterministic reservations: j is decided once every earlier reader of each of its k-mers is
//      decided; the smallest undecided j always is, so it terminates with exactly the serial answer.
//   I  inserters get addresses |D| + rank, in read order, and enter D with time = read index
//   B  second lookup for the other reads of the batch, restricted to positions before their current hit (the
//      reference returns the first POSITION, so an anchor inserted by an earlier read of the same batch can win)
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"
#include "count.cuh"

namespace lbanchor {

using lbcount::pack32;
using lbcount::window_kmer;

#define LB_EMPTY32 0xffffffffu
#define LB_EMPTY64 0xffffffffffffffffull

template <typename K> struct DictDev {
    K *keys;
    uint32_t *addr;
    uint32_t *time;      // LB_EMPTY32 = free slot
    uint8_t *tag;        // 0 = free slot, else dict_tag of the key: one byte per slot, the only array a failed probe touches
    uint64_t mask;       // capacity - 1 (power of two)
};

// The dictionary is private to this implementation (the reference keeps a std::unordered_map-like container,
// C/Leon.cpp:1105-1243), so its hash is free: one multiply, the high half folded into the low one (slot = low bits, tag = top
// byte).  The reference's hash64 used here at first cost ~60 of the ~120 instructions the lookups spend per position, and the
// lookups are issue-bound (ncu: sm 76 %, dram 4 %; profiles/r3e_ncu_full.csv).
template <typename K> LB_HD uint64_t dict_hash(K key)
{
    const uint64_t x = kmer_fold64(key) * 0x9E3779B97F4A7C15ULL;
    return x ^ (x >> 32);
}
// The lookups probe every k-mer of a read until the first hit, so most probes fail, and at the scale of configs[3] (2.8 M
// anchors, 16 M slots) the 4-byte and 8-byte arrays of the table are 64 + 128 MB: every failed probe was a DRAM access.  The
// tags of the same table are 16 MB and stay in L2; keys and times are only read behind a matching tag (1 in 255 by chance).
LB_HD uint8_t dict_tag(uint64_t h) { uint8_t t = (uint8_t)(h >> 56); return t ? t : (uint8_t)1; }

// forward-strand k-mers of positions [start, end) of a read, 32 per step, N read as A (C/DnaCoder.cpp:519-535).
// f(p, kmer, active) is called by all 32 lanes together and returns a warp-uniform "stop".
template <typename K, typename F>
LB_D void warp_for_kmers(const uint8_t *__restrict__ s, uint64_t L, int k, uint64_t start, uint64_t end, int lane, F &f)
{
    if (start >= end) return;
    uint64_t i0 = start + lane, i1 = start + 32 + lane;
    uint64_t w0 = pack32(i0 < L ? base_code_n_as_a(s[i0]) : 0u);
    uint64_t w1 = pack32(i1 < L ? base_code_n_as_a(s[i1]) : 0u);
    for (uint64_t base = start; base < end; base += 32) {
        uint64_t i2 = base + 64 + lane;
        uint64_t w2 = pack32(i2 < L ? base_code_n_as_a(s[i2]) : 0u);
        uint64_t p = base + lane;
        K kmer = window_kmer<K>(w0, w1, w2, lane, k);
        if (f(p, kmer, p < end)) return;
        w0 = w1; w1 = w2;
    }
}

template <typename K> LB_D bool dict_find(const DictDev<K> &d, K key, uint32_t before, uint32_t *addr)
{
    const uint64_t h = dict_hash(key);
    const uint8_t tg = dict_tag(h);
    uint64_t s = h & d.mask;
    while (true) {
        const uint8_t t8 = d.tag[s];
        if (t8 == 0) return false;
        if (t8 == tg && d.keys[s] == key) {
            if (d.time[s] < before) { *addr = d.addr[s]; return true; }
            return false;   // keys are unique in D
        }
        s = (s + 1) & d.mask;
    }
}

template <typename K> struct LookupFn {
    const DictDev<K> &d; int k; uint32_t r; uint64_t lim; int lane;
    int64_t pos; uint32_t addr;
    LB_D bool operator()(uint64_t p, K kmer, bool active)
    {
        uint32_t a = 0;
        bool hit = active && p < lim && dict_find(d, kmer_canon(kmer, k), r, &a);
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (bal) {
            int first = __ffs(bal) - 1;
            pos = (int64_t)(p - lane + first);
            addr = __shfl_sync(0xffffffffu, a, first);
            return true;
        }
        return false;
    }
};

// passes A and B.  apos[r] < 0: search all positions; apos[r] >= 0: only positions before it.
// One thread per read (the bases of the CTA's 256 reads staged in shared memory by one bulk copy): the forward k-mer
// and its reverse complement are rolled base by base, N read as A (C/DnaCoder.cpp:519-535), and every canonical k-mer is
// probed until the first hit -- a few tens of instructions per position instead of a warp-wide pack / shift / reverse.
template <typename K>
__global__ void __launch_bounds__(256) k_anchor_lookup(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                       uint64_t lo, uint64_t hi, uint64_t delta, int k, DictDev<K> d,
                                                       int32_t *apos, uint32_t *aaddr, const uint8_t *__restrict__ skip,
                                                       uint32_t stage_bytes)
{
    extern __shared__ __align__(128) uint8_t stage[];
    __shared__ __align__(8) uint64_t bar;
    const uint64_t r0 = lo + (uint64_t)blockIdx.x * 256;
    const uint64_t r1 = r0 + 256 < hi ? r0 + 256 : hi;
    const uint64_t o0 = offsets[r0];
    const uint8_t *sbase = lbcount::stage_reads(bases, o0, offsets[r1], stage, stage_bytes, &bar);
    __syncthreads();
    const uint64_t r = r0 + threadIdx.x;
    if (r >= hi) return;
    if (skip && skip[r]) return;
    const uint64_t o = offsets[r], L = offsets[r + 1] - o;
    if (L < (uint64_t)k) return;
    const int32_t cur = apos[r];
    const uint64_t lim = cur >= 0 ? (uint64_t)cur : L - k + 1;
    if (lim == 0) return;
    const uint8_t *s = sbase ? sbase + (o - o0) : bases + o;
    const K mask = kmer_mask(K(0), k);
    const int top = 2 * (k - 1);
    const uint32_t before = (uint32_t)(r + delta);      // index of the read in the whole set
    K fwd = K(0), rc = K(0);
    // (rolling LU = 4 positions and requesting their tags together before looking at them in order was measured: anchor stage
    // 8.8 -> 9.5 ms at configs[2], no change at configs[3] -- the probes are not what the threads wait for; profiles/r3b_*)
    for (uint64_t i = 0; i < lim + k - 1; i++) {
        const unsigned code = base_code_n_as_a(s[i]);
        fwd = ((fwd << 2) | K((uint64_t)code)) & mask;
        rc = (rc >> 2) | (K((uint64_t)(code ^ 2u)) << top);
        if (i + 1 >= (uint64_t)k) {
            uint32_t a = 0;
            if (dict_find(d, rc < fwd ? rc : fwd, before, &a)) {
                apos[r] = (int32_t)(i + 1 - k);
                aaddr[r] = a;
                return;
            }
        }
    }
}

// P list: reads of the batch that have k-mers but no hit yet
struct PIn {
    const uint64_t *offsets; const int32_t *apos; uint64_t lo; int k;
    __device__ uint32_t operator()(uint64_t i) const
    {
        uint64_t r = lo + i;
        return (offsets[r + 1] - offsets[r] >= (uint64_t)k && apos[r] < 0) ? 1u : 0u;
    }
};
struct POut {
    uint32_t *plist; uint64_t lo;
    __device__ void operator()(uint64_t i, uint32_t ex, uint32_t v) const { if (v) plist[ex] = (uint32_t)(lo + i); }
};
// index of the P reads in the whole read set (local index + delta of the batch)
__global__ void k_fill_gid(const uint32_t *__restrict__ plist, uint32_t m, uint64_t delta, uint32_t *__restrict__ gid)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) gid[j] = (uint32_t)(plist[j] + delta);
}

template <typename K> struct CandFn {
    const BloomDev &b; const uint64_t *tbl; int k; int lane;
    int64_t pos; K key;
    LB_D bool operator()(uint64_t p, K kmer, bool active)
    {
        K c = kmer_canon(kmer, k);
        bool hit = active && bloom_contains<K>(b, c, tbl);
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (bal) {
            int first = __ffs(bal) - 1;
            pos = (int64_t)(p - lane + first);
            uint64_t l = __shfl_sync(0xffffffffu, kmer_low64(c), first);
            uint64_t h = __shfl_sync(0xffffffffu, kmer_high64(c), first);
            kmer_make(key, l, h);
            return true;
        }
        return false;
    }
};

// Bloom candidate of every P read, search order of C/Leon.cpp:1139-1187
template <typename K>
__global__ void __launch_bounds__(256) k_candidates(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                    const uint32_t *__restrict__ plist, uint32_t m, int k, BloomDev b,
                                                    const uint64_t *__restrict__ tbl_g, int32_t *cand_pos, K *cand_key)
{
    __shared__ uint64_t tbl[256];
    tbl[threadIdx.x] = tbl_g[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t j = warp; j < m; j += nwarps) {
        uint64_t r = plist[j];
        uint64_t o = offsets[r], L = offsets[r + 1] - o;
        uint64_t n = L - k + 1;
        uint64_t iMin = n / 2, iMax = n / 2 + 10;
        if (iMax > n) iMax = n;
        CandFn<K> f{b, tbl, k, lane, -1, K(0)};
        warp_for_kmers<K>(bases + o, L, k, iMin, iMax, lane, f);
        if (f.pos < 0) warp_for_kmers<K>(bases + o, L, k, 0, iMin, lane, f);
        if (f.pos < 0) warp_for_kmers<K>(bases + o, L, k, iMax, n, lane, f);
        if (lane == 0) { cand_pos[j] = (int32_t)f.pos; cand_key[j] = f.key; }
    }
}

// ---- batch-local table over candidate values ------------------------------------------------------------------
template <typename K> struct CandTable {
    uint64_t *proxy;     // 64-bit stand-in used for the CAS (the key itself for 64-bit k-mers); LB_EMPTY64 = free
    K *keys;
    uint32_t *ins;       // P index of the decided inserter of this value (at most one), LB_EMPTY32 if none yet
    uint32_t *fu;        // smallest undecided P index having this value as candidate (per round)
    uint64_t mask;
};

template <typename K> LB_HD uint64_t cand_proxy(K key, uint64_t salt)
{
    uint64_t p = sizeof(K) == 8 ? kmer_low64(key) : (hash64(kmer_low64(key), salt) ^ kmer_high64(key));
    return p == LB_EMPTY64 ? 0x5bd1e995ull : p;
}

// state of a P entry
enum { ST_UNDECIDED = 0, ST_INSERTER = 1, ST_BLOCKED = 2, ST_NOCAND = 3 };

template <typename K>
__global__ void k_cand_build(const int32_t *__restrict__ cand_pos, const K *__restrict__ cand_key, uint32_t m, CandTable<K> t,
                             uint64_t salt, uint32_t *cand_slot, uint8_t *state)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    if (cand_pos[j] < 0) { state[j] = ST_NOCAND; cand_slot[j] = LB_EMPTY32; return; }
    state[j] = ST_UNDECIDED;
    K key = cand_key[j];
    uint64_t px = cand_proxy(key, salt);
    uint64_t s = hash64(px, 0x2545F4914F6CDD1DULL) & t.mask;
    while (true) {
        unsigned long long old = atomicCAS((unsigned long long *)&t.proxy[s], (unsigned long long)LB_EMPTY64, (unsigned long long)px);
        if (old == LB_EMPTY64) { t.keys[s] = key; break; }
        if (old == px) break;
        s = (s + 1) & t.mask;
    }
    cand_slot[j] = (uint32_t)s;
}

// two different wide k-mers with the same proxy would share a slot: detected here, the build is redone with a new salt
template <typename K>
__global__ void k_cand_verify(const K *__restrict__ cand_key, const uint32_t *__restrict__ cand_slot, uint32_t m, CandTable<K> t,
                              uint32_t *collision)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m || cand_slot[j] == LB_EMPTY32) return;
    if (t.keys[cand_slot[j]] != cand_key[j]) *collision = 1;
}

__global__ void k_round_fu(const uint8_t *__restrict__ state, const uint32_t *__restrict__ cand_slot, uint32_t m, uint32_t *fu)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m || state[j] != ST_UNDECIDED) return;
    atomicMin(&fu[cand_slot[j]], j);
}

template <typename K> struct DecideFn {
    const CandTable<K> &t; int k; uint32_t j; uint64_t salt;
    bool blocked, wait;
    LB_D bool operator()(uint64_t, K kmer, bool active)
    {
        bool b = false, w = false;
        if (active) {
            K c = kmer_canon(kmer, k);
            uint64_t px = cand_proxy(c, salt);
            uint64_t s = hash64(px, 0x2545F4914F6CDD1DULL) & t.mask;
            while (true) {
                uint64_t q = t.proxy[s];
                if (q == LB_EMPTY64) break;
                if (q == px && t.keys[s] == c) {
                    b = t.ins[s] < j;
                    w = t.fu[s] < j;
                    break;
                }
                s = (s + 1) & t.mask;
            }
        }
        blocked |= __any_sync(0xffffffffu, b);
        wait |= __any_sync(0xffffffffu, w);
        return blocked;
    }
};

template <typename K>
__global__ void __launch_bounds__(256) k_round_decide(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                      const uint32_t *__restrict__ plist, uint32_t m, uint32_t joff, int k,
                                                      CandTable<K> t, uint64_t salt, const uint32_t *__restrict__ cand_slot,
                                                      uint8_t *state, uint32_t *remaining)
{
    const int lane = threadIdx.x & 31;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    // this rank's P reads are entries [joff, joff + m) of the batch-wide P list; plist holds their local read indices
    for (uint32_t jl = warp; jl < m; jl += nwarps) {
        const uint32_t j = joff + jl;
        if (state[j] != ST_UNDECIDED) continue;
        uint64_t r = plist[jl];
        uint64_t o = offsets[r], L = offsets[r + 1] - o;
        uint64_t n = L - k + 1;
        DecideFn<K> f{t, k, j, salt, false, false};
        warp_for_kmers<K>(bases + o, L, k, 0, n, lane, f);
        if (lane == 0) {
            if (f.blocked) state[j] = ST_BLOCKED;
            else if (!f.wait) { state[j] = ST_INSERTER; atomicMin(&t.ins[cand_slot[j]], j); }
            else atomicAdd(remaining, 1u);
        }
    }
}

// after the states of the other ranks arrived: their new inserters enter the table
__global__ void k_ins_update(const uint8_t *__restrict__ state, const uint32_t *__restrict__ cand_slot, uint32_t m, uint32_t *ins)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m && state[j] == ST_INSERTER) atomicMin(&ins[cand_slot[j]], j);
}

struct InsIn {
    const uint8_t *state;
    __device__ uint32_t operator()(uint64_t j) const { return state[j] == ST_INSERTER ? 1u : 0u; }
};
template <typename K> struct InsOut {
    const uint32_t *gid; const int32_t *cand_pos; const K *cand_key;
    DictDev<K> d; uint32_t base_addr; K *anchors; int32_t *apos; uint32_t *aaddr; uint8_t *inserter;
    uint32_t own_lo, own_hi; uint64_t delta;      // entries [own_lo, own_hi) are reads of this rank: local index = gid - delta
    __device__ void operator()(uint64_t j, uint32_t ex, uint32_t v) const
    {
        if (!v) return;
        uint32_t g = gid[j];
        uint32_t a = base_addr + ex;
        K key = cand_key[j];
        anchors[a] = key;
        if (j >= own_lo && j < own_hi) {
            uint64_t r = g - delta;
            apos[r] = cand_pos[j];
            aaddr[r] = a;
            inserter[r] = 1;
        }
        const uint64_t h = dict_hash(key);
        uint64_t s = h & d.mask;
        while (atomicCAS(&d.time[s], LB_EMPTY32, g) != LB_EMPTY32) s = (s + 1) & d.mask;   // keys are distinct
        d.keys[s] = key;
        d.addr[s] = a;
        d.tag[s] = dict_tag(h);        // lookups run in later kernels
    }
};

// rehash into a larger dictionary
template <typename K> __global__ void k_dict_rehash(DictDev<K> from, DictDev<K> to)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > from.mask) return;
    uint32_t t = from.time[i];
    if (t == LB_EMPTY32) return;
    K key = from.keys[i];
    const uint64_t h = dict_hash(key);
    uint64_t s = h & to.mask;
    while (atomicCAS(&to.time[s], LB_EMPTY32, t) != LB_EMPTY32) s = (s + 1) & to.mask;
    to.keys[s] = key;
    to.addr[s] = from.addr[i];
    to.tag[s] = dict_tag(h);
}

}  // namespace lbanchor
