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
#include <vector>

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
        const uint32_t *e = err + (koff[r] - koff[0]);    // bif / err hold the k-mer slots of the super-block koff points at
        uint64_t prev = 0;
        for (uint32_t x = 0; x < rec.nErr; x++) {
            uint32_t idx = x < rec.nErrLeft ? rec.nErrLeft - 1 - x : x;
            uint64_t pos = e[idx];
            numeric(S_ERRPOS, pos - prev);
            prev = pos;
        }
    }
    // bifurcation symbols in walk order, each with the model of its type (C/DnaCoder.cpp:735-753)
    const uint8_t *bp = bif + (koff[r] - koff[0]);
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

// floor(num / d) and the remainder for num < d * 2^32 (32-bit quotient), d < 2^32: quotient estimate in double
// precision, biased down so that it is the exact quotient or one less, then one correction.
LB_D uint32_t div_64_32(uint64_t num, uint32_t d, double rcp_d, uint64_t &rem)
{
    // |relative error| of double(num) * rcp_d is < 2^-50; the quotient is < 2^32, so after the bias the estimate is
    // within (q - 1, q]: its floor is q or q - 1
    uint64_t q = __double2ull_rz(__ull2double_rz(num) * rcp_d * (1.0 - 1.0 / 1125899906842624.0));
    uint64_t r = num - q * d;
    if (r >= d) { q++; r -= d; }
    rem = r;
    return (uint32_t)q;
}

// Exact n / d for every 64-bit n by one multiply-add-high and a shift:  n / d == (hi64(n * m + a)) >> s.
// Round-up magic when it fits in 64 bits, else round-down magic with the numerator incremented (a = m), which also
// covers powers of two with m = 2^64 - 1.
LB_D void div_magic(uint32_t d, uint64_t &m, uint32_t &s, uint32_t &inc)
{
    int L = 31 - __clz(d);
    s = (uint32_t)L;
    if ((d & (d - 1)) == 0) { m = ~0ull; inc = 1; return; }
    const double rcp = 1.0 / (double)d;
    uint64_t r1, r2;
    uint64_t q1 = div_64_32((1ull << L) << 32, d, rcp, r1);      // in (2^31, 2^32): d is not a power of two
    uint64_t q0 = div_64_32(r1 << 32, d, rcp, r2);
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

// Ranks of the symbols of one 512-symbol chunk among the earlier symbols of their (block, model) segment.  The counts of
// the earlier chunks come from H (k_seg_scan); inside the chunk the 16 warps first count their own values (hw), a pass
// over the warps turns that into "occurrences of value v before warp w" (T), each warp prefix-sums its row of T over the
// values (P), and the ranks among the earlier lanes of the own warp come from one ballot per value bit.
__global__ void __launch_bounds__(CHUNK) k_chunk_ranks(const uint64_t *__restrict__ recs, const uint64_t *__restrict__ seg_start,
                                                       const uint64_t *__restrict__ chunk_off, uint64_t nseg,
                                                       const uint32_t *__restrict__ H, uint4 *__restrict__ rec_out,
                                                       uint8_t *__restrict__ sh_out)
{
    constexpr int NW = CHUNK / 32;
    __shared__ uint32_t T[NW][256];
    __shared__ uint32_t P[NW][256];
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
#pragma unroll
    for (int i = 0; i < NW * 256 / CHUNK; i++) (&T[0][0])[tid + i * CHUNK] = 0;
    __syncthreads();
    const uint64_t beg = sh_beg;
    const int n = (int)(sh_end - beg);
    const bool act = tid < n;
    uint64_t rec = 0;
    unsigned v = 0;
    if (act) { rec = recs[beg + tid]; v = (unsigned)(rec & 255u); atomicAdd(&T[warp][v], 1u); }
    const unsigned model = (unsigned)(__shfl_sync(0xffffffffu, rec, 0) >> REC_KEY_SHIFT) & ((1u << REC_MODEL_BITS) - 1u);
    __syncthreads();
    if (tid < 256) {      // occurrences of value tid before warp w (earlier chunks + earlier warps)
        uint32_t run = H[c * 256 + tid];
#pragma unroll
        for (int w = 0; w < NW; w++) { uint32_t t = T[w][tid]; T[w][tid] = run; run += t; }
    }
    __syncthreads();
    {   // P[warp][v] = sum over u < v of T[warp][u]
        uint32_t a[8], sum = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) { a[i] = T[warp][lane * 8 + i]; sum += a[i]; }
        uint32_t incl = sum;
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, incl, dlt);
            if (lane >= dlt) incl += o;
        }
        uint32_t run = incl - sum;
#pragma unroll
        for (int i = 0; i < 8; i++) { P[warp][lane * 8 + i] = run; run += a[i]; }
    }
    __syncwarp();
    // ranks among the earlier lanes of the warp: lanes that agree with v on the bits above b and have bit b clear while
    // v has it set are smaller
    const uint32_t alphabet = model_alphabet(model);        // a chunk holds one (block, model) segment
    const unsigned amask = __ballot_sync(0xffffffffu, act);
    const unsigned lt = (1u << lane) - 1u;
    unsigned same = amask, less = 0;
#pragma unroll
    for (int b = 7; b >= 0; b--) {
        const unsigned bb = __ballot_sync(0xffffffffu, act && ((v >> b) & 1u));
        if ((v >> b) & 1u) { less += __popc(same & ~bb & lt); same &= bb; }
        else same &= ~bb;
    }
    const unsigned eq = __popc(same & lt);
    if (!act) return;
    uint64_t pos = (rec >> 8) & 0xffffffffull;
    uint32_t lo = v + P[warp][v] + less;
    uint32_t w = 1u + T[warp][v] + eq;
    uint32_t tot = alphabet + (uint32_t)(sh_ci * CHUNK) + (uint32_t)tid;
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

constexpr uint32_t RC_MAX_BYTES_PER_SYMBOL = 12;   // <= 4 bytes of information + one clamp and its 7 refill shifts

// ---- range coder, current kernel ------------------------------------------------------------------------------
// A single warp issues at most one instruction every other cycle per pipe, so the serial recurrence is bound by the
// NUMBER of instructions per symbol as much as by their dependent chain.  Everything that does not depend on
// (low, range) is therefore done by the 31 other lanes: they unpack the records into shared memory in the form lane 0
// consumes (LDS.128 + LDS.64 per symbol), pad the last batch with identity symbols so that lane 0 runs fixed groups of
// four steps without bounds checks, and copy the output bytes, which lane 0 leaves in shared memory, to global memory.

// the two halves of a 64-bit value as separate 32-bit registers (keeps the compiler from re-fusing 32-bit tests on
// the high word into 64-bit compares)
LB_D void split64(uint64_t v, uint32_t &lo, uint32_t &hi) { asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }

// shared-memory accesses by 32-bit address (plain LDS / STS without generic-address arithmetic)
LB_D uint4 lds128(uint32_t a)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
LB_D uint2 lds64(uint32_t a)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
LB_D void sts8(uint32_t a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

// One generic coding step (RangeEncoder::encode, C/RangeCoder.cpp:106-134): any symbol, any state.  Bytes go to shared
// memory at address `on` (advanced).
LB_D void rc_step_any(uint64_t &low, uint64_t &range, const uint4 t, const uint32_t shift, const uint32_t inc, uint32_t &on)
{
    const uint64_t m = ((uint64_t)t.w << 32) | t.z;
    const uint64_t q = (uint64_t)(((unsigned __int128)range * m + (inc ? m : 0ull)) >> 64) >> shift;
    low += (uint64_t)t.x * q;
    range = q * (uint64_t)t.y;
#pragma unroll 1
    for (int guard = 0; guard < (int)RC_MAX_BYTES_PER_SYMBOL; guard++) {   // the bound only matters for corrupt records (range 0)
        const uint32_t lh = (uint32_t)(low >> 32);
        if ((lh ^ (uint32_t)((low + range) >> 32)) >= (1u << 24)) {     // top bytes of low and low + range differ
            if ((uint32_t)(range >> 32) >= (1u << 16)) break;           // range >= 2^48
            range = (0 - low) & (RC_BOTTOM - 1);
        }
        sts8(on++, lh >> 24);
        range <<= 8;
        low <<= 8;
    }
}

// The common case of a coding step as straight-line code without any predicate or branch (ISETP -> predicated use and
// ISETP -> BRA each cost ~13 cycles of a single warp's time, measured with ncu's per-instruction stall samples).
// State: low, and rp = range + inc of THIS symbol (div_magic: range / total == hi64((range + inc) * magic) >> shift; rp
// fits in 64 bits because range < 2^64 - 1 once a real symbol has been coded; the first symbols of a stream go through
// rc_step_any).  With x = hi32(low) ^ hi32(low + range) the renormalisation loop of RangeEncoder::encode exits after
// 0 byte (x >= 2^24 and range >= 2^48) or after exactly 1 byte (x in [2^16, 2^24): the second byte differs once shifted
// in; shifted range >= 2^48) for 98.6 % of the symbols.  The "one byte settled" decision is an arithmetic mask that
// selects between the two candidates of (low, rp); the byte is stored unconditionally and the output index advances by
// 0 or 1.  Whether the step was one of the rare ones (two bytes or more, or a range clamp) is only accumulated (rmin):
// the caller tests it once per group of four steps and, if so, recodes the group from its saved start state with
// rc_step_any.  After a
// rare symbol the remaining steps of the group run on a meaningless state, which is harmless: pure arithmetic, and the
// output index advances by at most one per step.
LB_D void rc_step_spec(uint64_t &low, uint64_t &rp, uint32_t &on, uint32_t &rmin, const uint4 t, const uint32_t shift,
                       const uint32_t inc_next)
{
    const uint64_t m = ((uint64_t)t.w << 32) | t.z;
    const uint64_t q = __umul64hi(rp, m) >> shift;
    const uint64_t lw = low + (uint64_t)t.x * q;
    const uint64_t rg = q * (uint64_t)t.y;
    const uint64_t top = low + (uint64_t)(t.x + t.y) * q;     // == lw + rg, formed beside them instead of after them
    uint32_t ll, lh, rl, rh, sl, shi;
    split64(lw, ll, lh);
    split64(rg, rl, rh);
    split64(top, sl, shi);
    const uint32_t x = lh ^ shi;
    uint32_t mask;                                            // all ones when x < 2^24 (top byte settled), else 0
    asm("{ .reg .u32 t; shr.u32 t, %1, 24; sub.u32 t, t, 1; shr.s32 %0, t, 31; }" : "=r"(mask) : "r"(x));
    sts8(on, lh >> 24);
    on -= mask;
    // candidates: no byte -> (lw, rg + inc_next); one byte -> (lw << 8, (rg << 8) | inc_next)
    const uint32_t l1l = ll << 8, l1h = (lh << 8) | (ll >> 24);
    const uint64_t c0 = rg + inc_next;
    uint32_t c0l, c0h;
    split64(c0, c0l, c0h);
    const uint32_t c1l = (rl << 8) | inc_next, c1h = (rh << 8) | (rl >> 24);
    // rare unless the second byte is not settled as well (x >= 2^16) and the range after the 0 or 1 byte is >= 2^48
    rmin = min(rmin, min(x, rh ^ ((rh ^ ((rh << 8) | (rl >> 24))) & mask)));
    const uint32_t nl = ll ^ ((ll ^ l1l) & mask), nh = lh ^ ((lh ^ l1h) & mask);
    const uint32_t pl = c0l ^ ((c0l ^ c1l) & mask), ph = c0h ^ ((c0h ^ c1h) & mask);
    low = ((uint64_t)nh << 32) | nl;
    rp = ((uint64_t)ph << 32) | pl;
}

constexpr int RC_BATCH = 64;       // symbol records staged per round (two per lane)
constexpr int RC_OBUF = RC_BATCH * RC_MAX_BYTES_PER_SYMBOL + 32;   // + scratch word at the end

// RangeEncoder::encode / flush, C/RangeCoder.cpp:106-145.  One warp per stream.
__global__ void __launch_bounds__(32) k_range_encode(const uint4 *__restrict__ rec_all, const uint8_t *__restrict__ sh_all,
                                                     const StreamDesc *__restrict__ descs, uint8_t *__restrict__ out_all,
                                                     uint64_t *__restrict__ out_size, int always_flush, uint32_t *overflow)
{
    __shared__ uint4 srec[2][RC_BATCH + 4];
    __shared__ uint2 ssh[2][RC_BATCH + 4];       // (shift, inc)
    __shared__ uint8_t sout[RC_OBUF];
    const int lane = threadIdx.x;
    const StreamDesc d = descs[blockIdx.x];
    const uint4 *rec = rec_all + d.rec_base;
    const uint8_t *sh = sh_all + d.rec_base;
    uint8_t *const out0 = out_all + d.out_base;
    uint64_t opos = 0;                            // bytes produced so far (all lanes)
    uint64_t low = 0, range = ~0ull;              // lane 0
    // identity symbol (total 1: magic 2^64 - 1, inc 1, shift 0 gives q = range; low += 0; range = q * 1): pads the last batch
    const uint4 idle = make_uint4(0u, 1u, 0xffffffffu, 0xffffffffu);
    uint4 na = idle, nb = idle;
    uint32_t nsa = 0x80u, nsb = 0x80u;
    if ((uint64_t)lane < d.nsym) { na = rec[lane]; nsa = sh[lane]; }
    if ((uint64_t)lane + 32 < d.nsym) { nb = rec[lane + 32]; nsb = sh[lane + 32]; }
    if (lane < 4) {     // slots past the batch: only their inc (0) is read, as the "next inc" of the last step
        srec[0][RC_BATCH + lane] = idle; srec[1][RC_BATCH + lane] = idle;
        ssh[0][RC_BATCH + lane] = make_uint2(0u, 0u); ssh[1][RC_BATCH + lane] = make_uint2(0u, 0u);
    }
    for (uint64_t base = 0; base < d.nsym; base += RC_BATCH) {
        const int buf = (int)((base / RC_BATCH) & 1);
        srec[buf][lane] = na;
        srec[buf][lane + 32] = nb;
        ssh[buf][lane] = make_uint2(nsa & 63u, nsa >> 7);
        ssh[buf][lane + 32] = make_uint2(nsb & 63u, nsb >> 7);
        const uint64_t in = base + RC_BATCH + lane;
        na = idle; nb = idle; nsa = 0x80u; nsb = 0x80u;
        if (in < d.nsym) { na = rec[in]; nsa = sh[in]; }
        if (in + 32 < d.nsym) { nb = rec[in + 32]; nsb = sh[in + 32]; }
        __syncwarp();
        const uint32_t sout_a = (uint32_t)__cvta_generic_to_shared(sout);
        uint32_t on = sout_a;
        if (lane == 0) {
            const uint32_t sr = (uint32_t)__cvta_generic_to_shared(srec[buf]);     // record j at sr + 16 j
            const uint32_t ss = (uint32_t)__cvta_generic_to_shared(ssh[buf]);      // (shift, inc) of j at ss + 8 j
            int j = 0;
            if (base == 0) {      // range = 2^64 - 1: range + inc does not fit in 64 bits yet
#pragma unroll 1
                for (; j < 4; j++) {
                    const uint2 sx = lds64(ss + 8 * j);
                    rc_step_any(low, range, lds128(sr + 16 * j), sx.x, sx.y, on);
                }
            }
            uint4 t0 = lds128(sr + 16 * j);
            uint2 s0 = lds64(ss + 8 * j);
            uint64_t rp = range + s0.y;
#pragma unroll 1
            for (; j < RC_BATCH; j += 4) {
                const uint4 t1 = lds128(sr + 16 * j + 16), t2 = lds128(sr + 16 * j + 32), t3 = lds128(sr + 16 * j + 48);
                const uint4 t4 = lds128(sr + 16 * j + 64);
                const uint2 s1 = lds64(ss + 8 * j + 8), s2 = lds64(ss + 8 * j + 16), s3 = lds64(ss + 8 * j + 24);
                const uint2 s4 = lds64(ss + 8 * j + 32);
                const uint64_t low_g = low, rp_g = rp;          // state at the start of the group
                const uint32_t on_g = on;
                uint32_t rmin = 0xffffffffu;
                rc_step_spec(low, rp, on, rmin, t0, s0.x, s1.y);
                rc_step_spec(low, rp, on, rmin, t1, s1.x, s2.y);
                rc_step_spec(low, rp, on, rmin, t2, s2.x, s3.y);
                rc_step_spec(low, rp, on, rmin, t3, s3.x, s4.y);
                if (__builtin_expect(rmin < (1u << 16), 0)) {
                    // a step needed more than one byte or a range clamp (5.5 % of the groups): recode the group from its start
                    // state, the common steps again with the straight-line step, only the rare ones with the generic step
                    low = low_g;
                    rp = rp_g;
                    on = on_g;
#pragma unroll 1
                    for (int i = j; i < j + 4; i++) {
                        const uint2 sx = lds64(ss + 8 * i), sn = lds64(ss + 8 * i + 8);
                        const uint4 t = lds128(sr + 16 * i);
                        uint64_t l2 = low, r2 = rp;
                        uint32_t o2 = on, rm = 0xffffffffu;
                        rc_step_spec(l2, r2, o2, rm, t, sx.x, sn.y);
                        if (rm >= (1u << 16)) { low = l2; rp = r2; on = o2; }
                        else {
                            range = rp - sx.y;
                            rc_step_any(low, range, t, sx.x, sx.y, on);
                            rp = range + sn.y;
                        }
                    }
                }
                t0 = t4;
                s0 = s4;
            }
            range = rp;           // the slot past the batch has inc 0
        }
        on -= sout_a;
        // all lanes: copy the bytes of this batch to the stream's output region
        on = __shfl_sync(0xffffffffu, on, 0);
        for (uint32_t i = lane; i < on; i += 32)
            if (opos + i < d.out_cap) out0[opos + i] = sout[i];
        opos += on;
        __syncwarp();
    }
    if (lane == 0) {
        uint64_t nout = opos;
        if (always_flush || nout > 0) {                       // DnaEncoder::writeBlock flushes only a non-empty buffer
            for (int x = 0; x < 8; x++) {
                if (nout < d.out_cap) out0[nout] = (uint8_t)(low >> 56);
                nout++;
                low <<= 8;
            }
        }
        out_size[blockIdx.x] = nout;
        if (nout > d.out_cap) *overflow = 1;
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
// The dictionary grows batch by batch; slices [a0, a0 + n) of it are turned into records as soon as they exist, `base`
// = nucleotide counts of the anchors before the slice (the total of the previous slices' scans).
template <typename K> struct DictCountIn {
    const K *anchors; int k;       // anchors: first anchor of the slice
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
    const K *anchors; int k; uint4 *rec; uint8_t *sh; lbscan::U32x4 base;
    __device__ void operator()(uint64_t a, lbscan::U32x4 ex, lbscan::U32x4) const
    {
        K x = anchors[a];
        uint32_t c[4] = {base.a + ex.a, base.b + ex.b, base.c + ex.c, base.d + ex.d};
        uint32_t tot = 5u + c[0] + c[1] + c[2] + c[3];
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

// ---- host side of the anchor dictionary stream: the serial recurrence over the records DictRecOut made ----------------
// One slice of records through the recurrence; same step as rc_step_spec (entropy.cuh): 0 or 1 output byte without a
// data-dependent branch, the rare rest through the generic loop of RangeEncoder::encode.
inline void dict_code_slice(const uint32_t *rec, const uint8_t *sh, uint64_t n, bool first, uint64_t &low, uint64_t &range,
                     std::vector<uint8_t> &out)
{
    const uint64_t TOP = RC_TOP, BOTTOM = RC_BOTTOM;
    const size_t pos0 = out.size();
    out.resize(pos0 + n * 2 + 64);          // the common step writes <= 1 byte per symbol; the rare path re-checks
    uint8_t *op = out.data() + pos0;
    const uint8_t *op_end = out.data() + out.size();
    uint64_t j = 0;
    if (first && n) {                       // first symbol: range = 2^64 - 1, range + inc needs 65 bits
        const uint64_t m = ((uint64_t)rec[3] << 32) | rec[2];
        const uint32_t s = sh[0];
        const uint64_t q = (uint64_t)(((unsigned __int128)range * m + ((s & 0x80u) ? m : 0ull)) >> 64) >> (s & 63u);
        low += (uint64_t)rec[0] * q;
        range = q * (uint64_t)rec[1];
        while ((low ^ (low + range)) < TOP || range < BOTTOM) {
            if (range < BOTTOM && (low ^ (low + range)) >= TOP) range = (0 - low) & (BOTTOM - 1);
            *op++ = (uint8_t)(low >> 56);
            range <<= 8;
            low <<= 8;
        }
        j = 1;
    }
    uint64_t rp = range + (j < n ? (sh[j] >> 7) : 0u);      // range + inc of symbol j
    for (; j < n; j++) {
        const uint32_t lo = rec[4 * j], w = rec[4 * j + 1];
        const uint64_t m = ((uint64_t)rec[4 * j + 3] << 32) | rec[4 * j + 2];
        const uint32_t inc_next = j + 1 < n ? (sh[j + 1] >> 7) : 0u;
        const uint64_t q = (uint64_t)(((unsigned __int128)rp * m) >> 64) >> (sh[j] & 63u);   // == range / total
        low += (uint64_t)lo * q;
        range = q * (uint64_t)w;
        const uint32_t x = (uint32_t)((low ^ (low + range)) >> 32), rh = (uint32_t)(range >> 32);
        if (__builtin_expect((x < rh ? x : rh) < 65536u, 0)) {
            while ((low ^ (low + range)) < TOP || range < BOTTOM) {
                if (range < BOTTOM && (low ^ (low + range)) >= TOP) range = (0 - low) & (BOTTOM - 1);
                *op++ = (uint8_t)(low >> 56);
                range <<= 8;
                low <<= 8;
            }
            if (op + (n - j) + 32 > op_end) {      // keep one byte per remaining symbol plus a rare step in reserve
                const size_t used = op - out.data();
                out.resize(out.size() * 2 + 64);
                op = out.data() + used;
                op_end = out.data() + out.size();
            }
            rp = range + inc_next;
            continue;
        }
        // both candidates of (low, rp) beside the test, selected by conditional moves: the dependent chain of a step is
        // multiply-high, shift, multiply, add, xor, compare, select
        const bool one = x < (1u << 24);
        const uint64_t rp0 = range + inc_next, rp1 = (range << 8) | inc_next;
        const uint64_t low1 = low << 8;
        *op = (uint8_t)(low >> 56);
        op += one;
        rp = one ? rp1 : rp0;
        low = one ? low1 : low;
    }
    range = rp;      // inc_next of the last symbol is 0
    out.resize(op - out.data());
}
// (A variant that forms the next quotient for both outcomes of the renormalisation decision and only selects -- a shorter dependent
// chain on paper -- is byte-identical but 15 % SLOWER on x86: 7.4 vs 6.4 ns per symbol, five multiplies per symbol on one multiply
// port; kept as dict_slice_spec in tools/rc_bench.cu, `rc_bench --dict`.)

}  // namespace lbent
