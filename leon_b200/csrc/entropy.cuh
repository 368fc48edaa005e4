// Phase 4 — per-block adaptive order-0 modelling + 64-bit carry-less range coding, bit-exact with
// C/RangeCoder.cpp:29-145, C/CompressionUtils.hpp:75-103 and the symbol order of C/DnaCoder.cpp:599-757,1158-1211.
//
// The adaptive model is evaluated WITHOUT serial model updates.  Before symbol t of model m the cumulative table of
// Order0Model is [0,1,..,A] plus the counts of the earlier symbols of m in the block (RangeCoder.cpp:47-51; the halving
// at 2^48 never triggers below 2^32 symbols), so for a symbol of value v
//     rangeLow = v + #{earlier, value < v}   rangeHigh - rangeLow = 1 + #{earlier, value == v}   total = A + #{earlier}
// which is a ranking problem:
//   k_emit_syms      one thread per read lists its symbols in coding order as (block, model, position, value) records
//   radix sort       stable sort of the records by (block, model)  [the kernels of count.cuh]
//   k_chunk_hist     512-symbol chunks of every (block, model) segment: 256-bin value histogram
//   k_seg_scan       exclusive prefix of those histograms along the chunks of a segment
//   k_chunk_ranks    per symbol: prefix over values + in-chunk ranks -> (low, width, total) and the exact reciprocal of
//                    total, written back at the symbol's coding position
//   k_range_encode   one warp per stream: the low/range recurrence (the only serial part) and byte emission
// The anchor dictionary stream (one never-reset Order0Model(5), C/Leon.cpp:1245-1260) gets its records from a prefix
// sum of nucleotide counts (DictCountIn/DictRecOut) and runs through the same k_range_encode on a second stream.
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"
#include "walk.cuh"
#include "scan.cuh"

namespace lbent {

using lbwalk::ReadRec;

constexpr int NB = 5;             // varint bytes modelled per numeric field: values < 2^35
constexpr int NUM_SETS = 8;
enum { S_ANCHOR_ADDR = 0, S_ANCHOR_POS, S_NOANCHOR_SIZE, S_READ_SIZE, S_NPOS, S_ERRPOS, S_NUMERIC, S_LEFTERR };
enum { M_READTYPE = 0, M_NOANCHOR, M_BIF, M_BIFBIN, M_REVCOMP, NUM_SMALL };
constexpr int NUM_MODELS = NUM_SMALL + NUM_SETS * NB;   // 45

LB_HD uint32_t model_alphabet(unsigned m)
{
    return m == M_READTYPE ? 2u : m == M_NOANCHOR ? 5u : m == M_BIF ? 5u : m == M_BIFBIN ? 2u : m == M_REVCOMP ? 2u : 256u;
}

// symbol record: [block:18][model:6][position:32][value:8]; position = index of the symbol in coding order
constexpr int REC_KEY_SHIFT = 40;
constexpr int REC_MODEL_BITS = 6;
LB_HD uint64_t make_rec(uint64_t block, unsigned model, uint64_t pos, unsigned value)
{
    return (block << (REC_KEY_SHIFT + REC_MODEL_BITS)) | ((uint64_t)model << REC_KEY_SHIFT) | (pos << 8) | value;
}

#define RC_TOP (1ull << 56)
#define RC_BOTTOM (1ull << 48)

LB_HD int varint_len(uint64_t v)
{
    int n = 1;
    while (v >>= 7) n++;
    return n;
}

// Walks the symbols of read r in coding order; f(model, value) is called once per symbol.
template <typename F>
LB_D void for_each_symbol(uint64_t r, const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                          const uint64_t *__restrict__ koff, const int32_t *__restrict__ apos,
                          const uint32_t *__restrict__ aaddr, const ReadRec *__restrict__ recs,
                          const uint8_t *__restrict__ bif, const uint32_t *__restrict__ err, F &f)
{
    auto numeric = [&](int set, uint64_t v) {
        int j = 0;
        do {
            unsigned byte = (unsigned)(v & 127);
            v >>= 7;
            if (v) byte |= 128;
            f((unsigned)(NUM_SMALL + set * NB + (j < NB ? j : NB - 1)), byte);
            j++;
        } while (v);
    };
    uint64_t o = offsets[r], L = offsets[r + 1] - o;
    const uint8_t *s = bases + o;
    int32_t ap = apos[r];
    if (ap < 0) {  // encodeNoAnchorRead, C/DnaCoder.cpp:1158-1211
        f(M_READTYPE, 1u);
        numeric(S_NOANCHOR_SIZE, L);
        for (uint64_t i = 0; i < L; i++) f(M_NOANCHOR, leon_nt2bin(s[i]));
        return;
    }
    ReadRec rec = recs[r];
    f(M_READTYPE, 0u);
    numeric(S_READ_SIZE, L);
    numeric(S_ANCHOR_POS, (uint64_t)ap);
    numeric(S_ANCHOR_ADDR, aaddr[r]);
    f(M_REVCOMP, rec.revcomp);
    numeric(S_NUMERIC, rec.nN);          // N positions, delta from 0 (C/DnaCoder.cpp:702-716)
    if (rec.nN) {
        uint64_t prev = 0;
        for (uint64_t i = 0; i < L; i++)
            if (s[i] == 'N') { numeric(S_NPOS, i - prev); prev = i; }
    }
    // error positions, sorted ascending = reversed left-walk part then right-walk part (C/DnaCoder.cpp:718-729)
    numeric(S_LEFTERR, rec.nErr);
    {
        const uint32_t *e = err + koff[r];
        uint64_t prev = 0;
        for (uint32_t x = 0; x < rec.nErr; x++) {
            uint32_t idx = x < rec.nErrLeft ? rec.nErrLeft - 1 - x : x;
            uint64_t pos = e[idx];
            numeric(S_ERRPOS, pos - prev);
            prev = pos;
        }
    }
    // bifurcation symbols in walk order, each with the model of its type (C/DnaCoder.cpp:735-753)
    const uint8_t *bp = bif + koff[r];
    for (uint32_t i = 0; i < rec.nBif; i++) {
        unsigned sym = bp[i];
        if (sym & 0x80u) f(M_BIFBIN, sym & 1u);
        else f(M_BIF, sym & 7u);
    }
}

struct SymCounter {
    uint64_t n;
    LB_D void operator()(unsigned, unsigned) { n++; }
};

// exact number of symbols of every read (scan input) and their first position (scan output)
struct SymCountIn {
    const uint8_t *bases; const uint64_t *offsets; const uint64_t *koff; const int32_t *apos; const uint32_t *aaddr;
    const ReadRec *recs; const uint8_t *bif; const uint32_t *err;
    __device__ uint64_t operator()(uint64_t r) const
    {
        SymCounter c{0};
        for_each_symbol(r, bases, offsets, koff, apos, aaddr, recs, bif, err, c);
        return c.n;
    }
};
struct SymOffOut {
    uint64_t *symoff;
    __device__ void operator()(uint64_t r, uint64_t ex, uint64_t) const { symoff[r] = ex; }
};

struct SymWriter {
    uint64_t *out; uint64_t pos; uint64_t block;
    LB_D void operator()(unsigned model, unsigned value) { out[pos] = make_rec(block, model, pos, value); pos++; }
};

__global__ void __launch_bounds__(256) k_emit_syms(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                   const uint64_t *__restrict__ koff, uint64_t nreads, uint32_t reads_per_block,
                                                   const int32_t *__restrict__ apos, const uint32_t *__restrict__ aaddr,
                                                   const ReadRec *__restrict__ recs, const uint8_t *__restrict__ bif,
                                                   const uint32_t *__restrict__ err, const uint64_t *__restrict__ symoff,
                                                   uint64_t *__restrict__ out)
{
    uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nreads) return;
    SymWriter w{out, symoff[r], r / reads_per_block};
    for_each_symbol(r, bases, offsets, koff, apos, aaddr, recs, bif, err, w);
}

// ---- segments (runs of equal (block, model) in the sorted records) and their 512-symbol chunks -----------------
constexpr int CHUNK = 512;

struct SegHeadIn {
    const uint64_t *recs;
    __device__ uint64_t operator()(uint64_t i) const
    {
        return (i == 0 || (recs[i] >> REC_KEY_SHIFT) != (recs[i - 1] >> REC_KEY_SHIFT)) ? 1ull : 0ull;
    }
};
struct SegHeadOut {
    uint64_t *seg_start;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t v) const { if (v) seg_start[ex] = i; }
};
// chunks per segment; seg_start has nseg + 1 entries (the last one = number of records)
struct SegChunksIn {
    const uint64_t *seg_start;
    __device__ uint64_t operator()(uint64_t s) const { return (seg_start[s + 1] - seg_start[s] + CHUNK - 1) / CHUNK; }
};
struct SegChunksOut {
    uint64_t *chunk_off;
    __device__ void operator()(uint64_t s, uint64_t ex, uint64_t) const { chunk_off[s] = ex; }
};

// segment of chunk c: the last s with chunk_off[s] <= c (segments are never empty, so chunk_off is strictly increasing)
LB_D uint64_t seg_of_chunk(const uint64_t *__restrict__ chunk_off, uint64_t nseg, uint64_t c)
{
    uint64_t lo = 0, hi = nseg;      // invariant: chunk_off[lo] <= c, hi == nseg or chunk_off[hi] > c
    while (hi - lo > 1) {
        uint64_t mid = (lo + hi) >> 1;
        if (chunk_off[mid] <= c) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(256) k_chunk_hist(const uint64_t *__restrict__ recs, const uint64_t *__restrict__ seg_start,
                                                    const uint64_t *__restrict__ chunk_off, uint64_t nseg,
                                                    uint32_t *__restrict__ H)
{
    __shared__ uint32_t h[256];
    __shared__ uint64_t sh_beg, sh_end;
    const uint64_t c = blockIdx.x;
    h[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        uint64_t s = seg_of_chunk(chunk_off, nseg, c);
        uint64_t beg = seg_start[s] + (c - chunk_off[s]) * CHUNK;
        uint64_t end = seg_start[s + 1];
        sh_beg = beg;
        sh_end = end < beg + CHUNK ? end : beg + CHUNK;
    }
    __syncthreads();
    for (uint64_t i = sh_beg + threadIdx.x; i < sh_end; i += 256) atomicAdd(&h[recs[i] & 255u], 1u);
    __syncthreads();
    H[c * 256 + threadIdx.x] = h[threadIdx.x];
}

// one CTA per segment, thread v: H[chunk][v] becomes the number of symbols of value v in the earlier chunks of the segment
__global__ void __launch_bounds__(256) k_seg_scan(const uint64_t *__restrict__ chunk_off, uint64_t nseg, uint64_t nchunks,
                                                  uint32_t *__restrict__ H)
{
    const uint64_t s = blockIdx.x;
    uint64_t c0 = chunk_off[s], c1 = s + 1 < nseg ? chunk_off[s + 1] : nchunks;
    uint32_t run = 0;
    uint32_t *p = H + c0 * 256 + threadIdx.x;
    uint64_t n = c1 - c0, i = 0;
    for (; i + 4 <= n; i += 4) {
        uint32_t a = p[0], b = p[256], c = p[512], d = p[768];
        p[0] = run; run += a;
        p[256] = run; run += b;
        p[512] = run; run += c;
        p[768] = run; run += d;
        p += 1024;
    }
    for (; i < n; i++) { uint32_t a = p[0]; p[0] = run; run += a; p += 256; }
}

// Exact n / d for every 64-bit n by one multiply-add-high and a shift:  n / d == (hi64(n * m + a)) >> s.
// Round-up magic when it fits in 64 bits, else round-down magic with the numerator incremented (a = m), which also
// covers powers of two with m = 2^64 - 1.
LB_D void div_magic(uint32_t d, uint64_t &m, uint32_t &s, uint32_t &inc)
{
    int L = 31 - __clz(d);
    s = (uint32_t)L;
    if ((d & (d - 1)) == 0) { m = ~0ull; inc = 1; return; }
    uint64_t t = (1ull << L) << 32;
    uint64_t q1 = t / d, r1 = t - q1 * d;
    uint64_t t2 = r1 << 32;
    uint64_t q0 = t2 / d, r2 = t2 - q0 * d;
    uint64_t mdown = (q1 << 32) | q0;            // floor(2^(64+L) / d)
    if ((uint64_t)d - r2 < (1ull << L)) { m = mdown + 1; inc = 0; }
    else { m = mdown; inc = 1; }
}

// what k_range_encode consumes per symbol: (low, width, magic lo, magic hi) + a byte (shift | increment << 7)
LB_D void store_symrec(uint4 *__restrict__ rec, uint8_t *__restrict__ sh, uint64_t pos, uint32_t lo, uint32_t w, uint32_t tot)
{
    uint64_t m; uint32_t s, inc;
    div_magic(tot, m, s, inc);
    rec[pos] = make_uint4(lo, w, (uint32_t)m, (uint32_t)(m >> 32));
    sh[pos] = (uint8_t)(s | (inc << 7));
}

__global__ void __launch_bounds__(CHUNK) k_chunk_ranks(const uint64_t *__restrict__ recs, const uint64_t *__restrict__ seg_start,
                                                       const uint64_t *__restrict__ chunk_off, uint64_t nseg,
                                                       const uint32_t *__restrict__ H, uint4 *__restrict__ rec_out,
                                                       uint8_t *__restrict__ sh_out)
{
    __shared__ uint8_t vals[CHUNK];
    __shared__ uint32_t cnt[256], cum[256], wsum[8];
    __shared__ uint64_t sh_beg, sh_end, sh_ci;
    const uint64_t c = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        uint64_t s = seg_of_chunk(chunk_off, nseg, c);
        uint64_t ci = c - chunk_off[s];
        uint64_t beg = seg_start[s] + ci * CHUNK, end = seg_start[s + 1];
        sh_beg = beg;
        sh_end = end < beg + CHUNK ? end : beg + CHUNK;
        sh_ci = ci;
    }
    __syncthreads();
    const uint64_t beg = sh_beg;
    const int n = (int)(sh_end - beg);
    uint64_t rec = 0;
    unsigned v = 0;
    if (tid < n) { rec = recs[beg + tid]; v = (unsigned)(rec & 255u); vals[tid] = (uint8_t)v; }
    // exclusive prefix over the 256 values of the counts of the earlier chunks
    uint32_t hv = 0, incl = 0;
    if (tid < 256) {
        hv = H[c * 256 + tid];
        cnt[tid] = hv;
        incl = hv;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) wsum[warp] = incl;
    }
    __syncthreads();
    if (tid < 256) {
        uint32_t base = 0;
        for (int w = 0; w < warp; w++) base += wsum[w];
        cum[tid] = base + incl - hv;
    }
    __syncthreads();
    if (tid >= n) return;
    // ranks among the earlier symbols of this chunk
    uint32_t less = 0, eq = 0;
    const int bound = n - 1 < warp * 32 + 31 ? n - 1 : warp * 32 + 31;   // largest tid of the warp that has a symbol
    for (int j = 0; j < bound; j++) {
        unsigned u = vals[j];
        if (j < tid) { less += u < v; eq += u == v; }
    }
    unsigned model = (unsigned)(rec >> REC_KEY_SHIFT) & ((1u << REC_MODEL_BITS) - 1u);
    uint64_t pos = (rec >> 8) & 0xffffffffull;
    uint32_t lo = v + cum[v] + less;
    uint32_t w = 1u + cnt[v] + eq;
    uint32_t tot = model_alphabet(model) + (uint32_t)(sh_ci * CHUNK) + (uint32_t)tid;
    store_symrec(rec_out, sh_out, pos, lo, w, tot);
}

// ---- range coder -------------------------------------------------------------------------------------------------
struct StreamDesc {
    uint64_t rec_base;   // first symbol record of the stream
    uint64_t nsym;
    uint64_t out_base;   // byte offset of the stream's output region
    uint64_t out_cap;
};

// one stream per block: its symbols are [symoff[first read], symoff[first read of the next block])
__global__ void k_block_descs(const uint64_t *__restrict__ symoff, uint64_t nreads, uint32_t reads_per_block,
                              uint64_t nblocks, StreamDesc *descs)
{
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nblocks) return;
    uint64_t r0 = b * reads_per_block, r1 = r0 + reads_per_block;
    if (r1 > nreads) r1 = nreads;
    uint64_t base = symoff[r0], end = symoff[r1];
    StreamDesc d;
    d.rec_base = base;
    d.nsym = end - base;
    d.out_base = 2 * base + 1024 * b;
    d.out_cap = 2 * (end - base) + 1024;
    descs[b] = d;
}

// One coding step (RangeEncoder::encode, C/RangeCoder.cpp:106-134) with range / total done as
// hi64((range + inc) * magic) >> shift (div_magic).  Bytes shifted out go to *op (op advances); when `store` is
// false (output region exhausted, reported as overflow by the caller) they are only counted.
LB_D void rc_step(uint64_t &low, uint64_t &range, const uint4 t, const uint32_t sh, uint8_t *&op, const bool store)
{
    const uint64_t m = ((uint64_t)t.w << 32) | t.z;
    const uint64_t q = __umul64hi(range + (sh >> 7), m) >> (sh & 63u);
    low += (uint64_t)t.x * q;
    range = q * (uint64_t)t.y;
    while (true) {
        const uint32_t lh = (uint32_t)(low >> 32);
        const uint32_t sh32 = (uint32_t)((low + range) >> 32);
        if ((lh ^ sh32) >= (1u << 24)) {                       // top bytes of low and low + range differ
            if ((uint32_t)(range >> 32) >= (1u << 16)) break;  // range >= 2^48
            range = (0 - low) & (RC_BOTTOM - 1);
        }
        if (store) *op = (uint8_t)(lh >> 24);
        op++;
        range <<= 8;
        low <<= 8;
    }
}

constexpr uint32_t RC_MAX_BYTES_PER_SYMBOL = 12;   // <= 4 bytes of information + one clamp and its 7 refill shifts

// RangeEncoder::encode / flush, C/RangeCoder.cpp:106-145.  One warp per stream: the 32 lanes stage the next 32 symbol
// records in shared memory (coalesced loads, issued one batch ahead), lane 0 runs the low/range recurrence over them
// with the record of the next symbol already in registers.
__global__ void __launch_bounds__(32) k_range_encode(const uint4 *__restrict__ rec_all, const uint8_t *__restrict__ sh_all,
                                                     const StreamDesc *__restrict__ descs, uint8_t *__restrict__ out_all,
                                                     uint64_t *__restrict__ out_size, int always_flush, uint32_t *overflow)
{
    __shared__ uint4 srec[2][34];
    __shared__ uint32_t ssh[2][34];
    const int lane = threadIdx.x;
    const StreamDesc d = descs[blockIdx.x];
    const uint4 *rec = rec_all + d.rec_base;
    const uint8_t *sh = sh_all + d.rec_base;
    uint8_t *const out0 = out_all + d.out_base;
    uint8_t *op = out0;
    bool ovf = false;
    uint64_t low = 0, range = ~0ull;
    uint4 nx = make_uint4(0u, 1u, 0xffffffffu, 0xffffffffu);
    uint32_t nsx = 0x80u;
    if ((uint64_t)lane < d.nsym) { nx = rec[lane]; nsx = sh[lane]; }
    for (uint64_t base = 0; base < d.nsym; base += 32) {
        const int buf = (int)((base >> 5) & 1);
        srec[buf][lane] = nx;
        ssh[buf][lane] = nsx;
        uint64_t in = base + 32 + lane;
        if (in < d.nsym) { nx = rec[in]; nsx = sh[in]; }
        __syncwarp();
        if (lane == 0) {
            const int steps = d.nsym - base < 32 ? (int)(d.nsym - base) : 32;
            // room for the worst case of this batch, checked once instead of per byte
            const bool store = (uint64_t)(op - out0) + 32 * RC_MAX_BYTES_PER_SYMBOL + 8 <= d.out_cap;
            ovf |= !store;
            const uint4 *sr = srec[buf];
            const uint32_t *ss = ssh[buf];
            int j = 0;
            if (base == 0) {
                // first symbol: range is 2^64 - 1, so range + inc may wrap; (2^64 * m) >> 64 == m
                uint4 t = sr[0];
                uint32_t s = ss[0];
                uint64_t m = ((uint64_t)t.w << 32) | t.z;
                uint64_t q = ((s >> 7) ? m : __umul64hi(range, m)) >> (s & 63u);
                low = (uint64_t)t.x * q;
                range = q * (uint64_t)t.y;
                while ((low ^ (low + range)) < RC_TOP || range < RC_BOTTOM) {
                    if (range < RC_BOTTOM && (low ^ (low + range)) >= RC_TOP) range = (0 - low) & (RC_BOTTOM - 1);
                    if (store) *op = (uint8_t)(low >> 56);
                    op++;
                    range <<= 8;
                    low <<= 8;
                }
                j = 1;
            }
            uint4 ta = sr[j];
            uint32_t sa = ss[j];
            for (; j + 1 < steps; j += 2) {          // two symbols per iteration: no register rotation
                uint4 tb = sr[j + 1];
                uint32_t sb = ss[j + 1];
                rc_step(low, range, ta, sa, op, store);
                ta = sr[j + 2];                       // slots 32, 33 exist: the arrays have 34 entries
                sa = ss[j + 2];
                rc_step(low, range, tb, sb, op, store);
            }
            if (j < steps) rc_step(low, range, ta, sa, op, store);
        }
        __syncwarp();
    }
    if (lane == 0) {
        uint64_t nout = (uint64_t)(op - out0);
        if (always_flush || nout > 0) {                       // DnaEncoder::writeBlock flushes only a non-empty buffer
            for (int x = 0; x < 8; x++) {
                if (nout < d.out_cap) out0[nout] = (uint8_t)(low >> 56);
                else ovf = true;
                nout++;
                low <<= 8;
            }
        }
        out_size[blockIdx.x] = nout;
        if (ovf) *overflow = 1;
    }
}

// ---- gather the per-stream regions into one contiguous payload ---------------------------------------------
struct SizeIn {
    const uint64_t *sz;
    __device__ uint64_t operator()(uint64_t i) const { return sz[i]; }
};
struct SizeOut {
    uint64_t *off;
    __device__ void operator()(uint64_t i, uint64_t ex, uint64_t) const { off[i] = ex; }
};
__global__ void __launch_bounds__(256) k_gather_streams(const uint8_t *__restrict__ out_all, const StreamDesc *__restrict__ descs,
                                                        const uint64_t *__restrict__ sizes, const uint64_t *__restrict__ dst_off,
                                                        uint8_t *__restrict__ dst)
{
    const uint8_t *src = out_all + descs[blockIdx.x].out_base;
    uint8_t *d = dst + dst_off[blockIdx.x];
    uint64_t n = sizes[blockIdx.x];
    for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) d[i] = src[i];
}

// blocksizes[2b] = bytes, [2b+1] = reads in block b (C/Leon.cpp:699-705)
__global__ void k_blocksizes(const uint64_t *__restrict__ sizes, uint64_t nreads, uint32_t reads_per_block, uint64_t nblocks,
                             uint64_t *__restrict__ bs)
{
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nblocks) return;
    uint64_t r0 = b * reads_per_block, r1 = r0 + reads_per_block;
    if (r1 > nreads) r1 = nreads;
    bs[2 * b] = sizes[b];
    bs[2 * b + 1] = r1 - r0;
}

// ---- anchor dictionary stream ----------------------------------------------------------------------------------
template <typename K> struct DictCountIn {
    const K *anchors; int k;
    __device__ lbscan::U32x4 operator()(uint64_t a) const
    {
        K x = anchors[a];
        uint32_t c[4] = {0, 0, 0, 0};
        for (int i = 0; i < k; i++) { c[(unsigned)(x & 3)]++; x >>= 2; }
        return lbscan::U32x4{c[0], c[1], c[2], c[3]};
    }
};
// k symbols per anchor, most significant nucleotide first (toString order, C/Leon.cpp:1245-1253), Order0Model(5)
template <typename K> struct DictRecOut {
    const K *anchors; int k; uint4 *rec; uint8_t *sh;
    __device__ void operator()(uint64_t a, lbscan::U32x4 ex, lbscan::U32x4) const
    {
        K x = anchors[a];
        uint32_t c[4] = {ex.a, ex.b, ex.c, ex.d};
        uint32_t tot = 5u + ex.a + ex.b + ex.c + ex.d;
        uint64_t o = a * (uint64_t)k;
        for (int i = k - 1; i >= 0; i--) {
            unsigned v = (unsigned)(x >> (2 * i)) & 3u;
            uint32_t lo = v;
            for (unsigned u = 0; u < v; u++) lo += c[u];
            store_symrec(rec, sh, o + (uint64_t)(k - 1 - i), lo, 1u + c[v], tot);
            c[v]++;
            tot++;
        }
    }
};

}  // namespace lbent
