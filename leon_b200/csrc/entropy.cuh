// Phase 4 — per-block adaptive order-0 modelling + 64-bit carry-less range coding, bit-exact with
// C/RangeCoder.cpp:29-145, C/CompressionUtils.hpp:75-103 and the symbol order of C/DnaCoder.cpp:599-757,1158-1211.
//
// Split in two so that the only strictly serial part is the low/range recurrence:
//   k_model_scan   : one warp per block walks the block's symbols in coding order, keeps every Order0Model as a
//                    cumulative table in shared memory (lanes update it in parallel) and writes, per symbol, the
//                    triple (rangeLow, rangeHigh-rangeLow, totalRange) the coder needs.
//   k_range_encode : one warp per stream; 32 triples are fetched (and their reciprocals computed) in parallel, then
//                    the recurrence runs over them and emits bytes.
// The anchor dictionary stream (one never-reset Order0Model(5), C/Leon.cpp:1245-1260) gets its triples from a
// prefix sum of nucleotide counts instead (k_dict_counts + scan + k_dict_triples).
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

#define RC_TOP (1ull << 56)
#define RC_BOTTOM (1ull << 48)

// upper bound of the number of symbols a read contributes (exact count is produced by k_model_scan)
struct SymBoundIn {
    const uint64_t *offsets; const int32_t *apos; const ReadRec *recs;
    __device__ uint64_t operator()(uint64_t r) const
    {
        if (apos[r] < 0) return 8 + (offsets[r + 1] - offsets[r]);
        const ReadRec &x = recs[r];
        return 32ull + 5ull * ((uint64_t)x.nN + x.nErr) + x.nBif;
    }
};
struct SymBoundOut {
    uint64_t *symoff;
    __device__ void operator()(uint64_t r, uint64_t ex, uint64_t) const { symoff[r] = ex; }
};

struct ModelTables {
    uint32_t small[NUM_SMALL][8];
    uint32_t num[NUM_SETS][NB][257];
};

LB_D void model_emit(uint32_t *cum, int A, unsigned c, uint4 *tri, uint64_t &ns, int lane)
{
    uint32_t lo = cum[c], hi = cum[c + 1], tot = cum[A];
    if (lane == 0) tri[ns] = make_uint4(lo, hi - lo, tot, 0u);
    ns++;
    __syncwarp();
    for (int i = (int)c + 1 + lane; i <= A; i += 32) cum[i]++;     // Order0Model::update, RangeCoder.cpp:47-51
    __syncwarp();
}

LB_D void model_emit_numeric(ModelTables *m, int set, uint64_t v, uint4 *tri, uint64_t &ns, int lane)
{
    int j = 0;
    do {
        unsigned byte = (unsigned)(v & 127);
        v >>= 7;
        if (v) byte |= 128;
        model_emit(m->num[set][j < NB ? j : NB - 1], 256, byte, tri, ns, lane);
        j++;
    } while (v);
}

__global__ void __launch_bounds__(32) k_model_scan(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                                   const uint64_t *__restrict__ koff, uint64_t nreads, uint32_t reads_per_block,
                                                   const int32_t *__restrict__ apos, const uint32_t *__restrict__ aaddr,
                                                   const ReadRec *__restrict__ recs, const uint8_t *__restrict__ bif,
                                                   const uint32_t *__restrict__ err, const uint64_t *__restrict__ symoff,
                                                   uint4 *__restrict__ tri_all, uint64_t *__restrict__ nsym_out)
{
    __shared__ ModelTables m;
    const int lane = threadIdx.x;
    const uint64_t blk = blockIdx.x;
    const uint64_t r0 = blk * reads_per_block;
    uint64_t r1 = r0 + reads_per_block;
    if (r1 > nreads) r1 = nreads;
    // startBlock: every model back to [0,1,...,n] (C/DnaCoder.cpp:78-112)
    for (int i = lane; i < NUM_SMALL * 8; i += 32) (&m.small[0][0])[i] = (uint32_t)(i & 7);
    for (int i = lane; i < NUM_SETS * NB * 257; i += 32) (&m.num[0][0][0])[i] = (uint32_t)(i % 257);
    __syncwarp();
    uint4 *tri = tri_all + symoff[r0];
    uint64_t ns = 0;
    for (uint64_t r = r0; r < r1; r++) {
        uint64_t o = offsets[r], L = offsets[r + 1] - o;
        const uint8_t *s = bases + o;
        int32_t ap = apos[r];
        if (ap < 0) {
            // encodeNoAnchorRead, C/DnaCoder.cpp:1158-1211
            model_emit(m.small[M_READTYPE], 2, 1u, tri, ns, lane);
            model_emit_numeric(&m, S_NOANCHOR_SIZE, L, tri, ns, lane);
            for (uint64_t base = 0; base < L; base += 32) {
                uint64_t i = base + lane;
                unsigned cb = i < L ? s[i] : 0u;
                int steps = L - base < 32 ? (int)(L - base) : 32;
                for (int t = 0; t < steps; t++)
                    model_emit(m.small[M_NOANCHOR], 5, leon_nt2bin((uint8_t)__shfl_sync(0xffffffffu, cb, t)), tri, ns, lane);
            }
            continue;
        }
        ReadRec rec = recs[r];
        model_emit(m.small[M_READTYPE], 2, 0u, tri, ns, lane);
        model_emit_numeric(&m, S_READ_SIZE, L, tri, ns, lane);
        model_emit_numeric(&m, S_ANCHOR_POS, (uint64_t)ap, tri, ns, lane);
        model_emit_numeric(&m, S_ANCHOR_ADDR, aaddr[r], tri, ns, lane);
        model_emit(m.small[M_REVCOMP], 2, rec.revcomp, tri, ns, lane);
        // N positions, delta from 0 (C/DnaCoder.cpp:702-716)
        model_emit_numeric(&m, S_NUMERIC, rec.nN, tri, ns, lane);
        if (rec.nN) {
            uint64_t prev = 0;
            for (uint64_t base = 0; base < L; base += 32) {
                uint64_t i = base + lane;
                unsigned bal = __ballot_sync(0xffffffffu, i < L && s[i] == 'N');
                while (bal) {
                    uint64_t pos = base + (__ffs(bal) - 1);
                    model_emit_numeric(&m, S_NPOS, pos - prev, tri, ns, lane);
                    prev = pos;
                    bal &= bal - 1;
                }
            }
        }
        // error positions, sorted ascending = reversed left-walk part then right-walk part (C/DnaCoder.cpp:718-729)
        model_emit_numeric(&m, S_LEFTERR, rec.nErr, tri, ns, lane);
        {
            const uint32_t *e = err + koff[r];
            uint64_t prev = 0;
            for (uint32_t x = 0; x < rec.nErr; x++) {
                uint32_t idx = x < rec.nErrLeft ? rec.nErrLeft - 1 - x : x;
                uint64_t pos = e[idx];
                model_emit_numeric(&m, S_ERRPOS, pos - prev, tri, ns, lane);
                prev = pos;
            }
        }
        // bifurcation symbols in walk order, each with the model of its type (C/DnaCoder.cpp:735-753)
        {
            const uint8_t *bp = bif + koff[r];
            for (uint32_t base = 0; base < rec.nBif; base += 32) {
                uint32_t i = base + lane;
                unsigned sb = i < rec.nBif ? bp[i] : 0u;
                int steps = rec.nBif - base < 32 ? (int)(rec.nBif - base) : 32;
                for (int t = 0; t < steps; t++) {
                    unsigned sym = __shfl_sync(0xffffffffu, sb, t);
                    if (sym & 0x80u) model_emit(m.small[M_BIFBIN], 2, sym & 1u, tri, ns, lane);
                    else model_emit(m.small[M_BIF], 5, sym & 7u, tri, ns, lane);
                }
            }
        }
    }
    if (lane == 0) nsym_out[blk] = ns;
}

struct StreamDesc {
    uint64_t tri_base;   // first triple of the stream
    uint64_t nsym;
    uint64_t out_base;   // byte offset of the stream's output region
    uint64_t out_cap;
};

// descriptors of the per-block streams: triples start at symoff[first read], output region sized from the bound
__global__ void k_block_descs(const uint64_t *__restrict__ symoff, const uint64_t *__restrict__ symtotal,
                              const uint64_t *__restrict__ nsym, uint64_t nreads, uint32_t reads_per_block, uint64_t nblocks,
                              StreamDesc *descs)
{
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nblocks) return;
    uint64_t r0 = b * reads_per_block;
    uint64_t r1 = r0 + reads_per_block;
    uint64_t base = symoff[r0];
    uint64_t end = r1 < nreads ? symoff[r1] : *symtotal;
    StreamDesc d;
    d.tri_base = base;
    d.nsym = nsym[b];
    d.out_base = 2 * base + 16 * b;
    d.out_cap = 2 * (end - base) + 16;
    descs[b] = d;
}

// RangeEncoder::encode / flush, C/RangeCoder.cpp:106-145.  One warp per stream; all lanes run the same
// recurrence (uniform), lane 0 stores.  range / total uses the per-symbol reciprocal computed 32 at a time.
__global__ void __launch_bounds__(32) k_range_encode(const uint4 *__restrict__ tri_all, const StreamDesc *__restrict__ descs,
                                                     uint8_t *__restrict__ out_all, uint64_t *__restrict__ out_size,
                                                     int always_flush, uint32_t *overflow)
{
    const int lane = threadIdx.x;
    StreamDesc d = descs[blockIdx.x];
    const uint4 *tri = tri_all + d.tri_base;
    uint8_t *out = out_all + d.out_base;
    uint64_t low = 0, range = ~0ull, nout = 0;
    bool ovf = false;
    for (uint64_t base = 0; base < d.nsym; base += 32) {
        uint64_t i = base + lane;
        uint4 t = i < d.nsym ? tri[i] : make_uint4(0u, 1u, 1u, 0u);
        uint64_t magic = ~0ull / (uint64_t)t.z;
        int steps = d.nsym - base < 32 ? (int)(d.nsym - base) : 32;
        for (int j = 0; j < steps; j++) {
            uint64_t lo = __shfl_sync(0xffffffffu, t.x, j);
            uint64_t w = __shfl_sync(0xffffffffu, t.y, j);
            uint64_t tot = __shfl_sync(0xffffffffu, t.z, j);
            uint64_t mg = __shfl_sync(0xffffffffu, magic, j);
            uint64_t q = __umul64hi(range, mg);
            uint64_t rem = range - q * tot;                     // estimate is never above, at most 2 below
            if (rem >= tot) { q++; rem -= tot; }
            if (rem >= tot) q++;                                // q = range / tot exactly
            low += lo * q;
            range = q * w;
            while ((low ^ (low + range)) < RC_TOP || range < RC_BOTTOM) {
                if (range < RC_BOTTOM && (low ^ (low + range)) >= RC_TOP) range = (0 - low) & (RC_BOTTOM - 1);
                if (nout < d.out_cap) { if (lane == 0) out[nout] = (uint8_t)(low >> 56); }
                else ovf = true;
                nout++;
                range <<= 8;
                low <<= 8;
            }
        }
    }
    if (always_flush || nout > 0) {                             // DnaEncoder::writeBlock flushes only a non-empty buffer
        for (int x = 0; x < 8; x++) {
            if (nout < d.out_cap) { if (lane == 0) out[nout] = (uint8_t)(low >> 56); }
            else ovf = true;
            nout++;
            low <<= 8;
        }
    }
    if (lane == 0) {
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
template <typename K> struct DictTriOut {
    const K *anchors; int k; uint4 *tri;
    __device__ void operator()(uint64_t a, lbscan::U32x4 ex, lbscan::U32x4) const
    {
        K x = anchors[a];
        uint32_t c[4] = {ex.a, ex.b, ex.c, ex.d};
        uint32_t tot = 5u + ex.a + ex.b + ex.c + ex.d;
        uint4 *o = tri + a * (uint64_t)k;
        for (int i = k - 1; i >= 0; i--) {
            unsigned v = (unsigned)(x >> (2 * i)) & 3u;
            uint32_t lo = v;
            for (unsigned u = 0; u < v; u++) lo += c[u];
            o[k - 1 - i] = make_uint4(lo, 1u + c[v], tot, 0u);
            c[v]++;
            tot++;
        }
    }
};

}  // namespace lbent
