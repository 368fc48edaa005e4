// Decompression of the sequence stream on the GPU (SURVEY.md 8f-3 "GPU decompressor"): one warp per dna_<id> block.
// Follows C/DnaCoder.cpp:1400-1480 (decodeAnchorRead), :1482-1605 (extendAnchor), :1607-1622 (decodeNoAnchorRead),
// C/RangeCoder.cpp:179-244 (RangeDecoder), C/CompressionUtils.hpp:114-137 (decodeNumeric), C/DnaCoder.cpp:43-112 (the
// model set), COL/Bloom.hpp:645-798 (contains4).
//
// A block is one dependent chain: the model a symbol is decoded with depends on the symbols before it, and how many
// bifurcation symbols a read consumes is only known while walking it through the filter.  So the parallelism across
// blocks is one warp per block (a grid-stride loop when there are more blocks than resident warps), and inside the
// warp the lanes share the work of each step of the chain:
//  * the coder state (low, range, code) and the k-mer are warp-uniform registers;
//  * a 256-symbol model is 8 counts per lane: the total and the symbol holding a cumulative count come from one warp
//    prefix sum and one ballot instead of a walk over the table (Order0Model keeps cumulative sums and updates O(n));
//  * the 28 bit tests of contains4 (7 hashes x 4 candidate nucleotides) are one warp-wide gather inside the <= 514-byte
//    window of the filter;
//  * N and error positions of the read are bitmaps in shared memory, the read's characters are stored as they are found.
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"

namespace lbdd {

constexpr uint64_t RC_TOP = 1ull << 56, RC_BOTTOM = 1ull << 48;
constexpr int NB_NUMERIC = 20;
// numeric model vectors of DnaDecoder (C/DnaCoder.cpp:43-112), 20 models of 256 symbols each
enum { V_ADDR = 0, V_APOS, V_NOANCHOR_SIZE, V_READSIZE, V_NPOS, V_ERRPOS, V_NUMERIC, V_LEFTERR, NVEC };
constexpr size_t MODEL_WORDS = (size_t)NVEC * NB_NUMERIC * 256;      // u32 counts per warp (160 KB)

// why a block stopped (the host turns them into the messages of the host decoder)
enum { DE_OK = 0, DE_RANGE = 1, DE_NCOUNT = 2, DE_ECOUNT = 3, DE_NOSUCC = 4, DE_NOANCHOR_LEN = 5 };

struct BlockIn { uint64_t in_off, in_bytes, nreads, out_off; };

struct Rc {
    uint64_t low, range, code;
    const uint8_t *p, *end;
};
LB_D uint32_t rc_next(Rc &rc) { return rc.p < rc.end ? (uint32_t)__ldg(rc.p++) : 0xffu; }   // past the end the reference reads EOF
LB_D void rc_start(Rc &rc, const uint8_t *data, uint64_t n)
{
    rc.low = 0; rc.range = ~0ull; rc.code = 0; rc.p = data; rc.end = data + n;
    for (int i = 0; i < 8; i++) rc.code = (rc.code << 8) | rc_next(rc);
}
LB_D void rc_consume(Rc &rc, uint64_t lo, uint32_t width)      // RangeDecoder::removeRange + normalize
{
    rc.low += lo * rc.range;
    rc.range *= width;
    while ((rc.low ^ (rc.low + rc.range)) < RC_TOP || rc.range < RC_BOTTOM) {
        if (rc.range < RC_BOTTOM && (rc.low ^ (rc.low + rc.range)) >= RC_TOP) rc.range = (0 - rc.low) & (RC_BOTTOM - 1);
        rc.code = (rc.code << 8) | rc_next(rc);
        rc.range <<= 8;
        rc.low <<= 8;
    }
}

// models of 2 or 5 symbols: counts in warp-uniform registers
template <int N> LB_D int decode_small(Rc &rc, uint32_t (&c)[N])
{
    uint32_t total = 0;
#pragma unroll
    for (int i = 0; i < N; i++) total += c[i];
    rc.range /= total;
    const uint64_t count = (rc.code - rc.low) / rc.range;
    uint64_t s = 0;
    int sym = 0;
#pragma unroll
    for (int i = 0; i < N - 1; i++)
        if (sym == i && s + c[i] <= count) { s += c[i]; sym = i + 1; }
    uint32_t w = 1;
#pragma unroll
    for (int i = 0; i < N; i++)
        if (sym == i) { w = c[i]; c[i]++; }
    rc_consume(rc, s, w);
    return sym;
}

// model of 256 symbols: lane l holds the counts of symbols 8l .. 8l+7 (global memory, one 32-byte sector per lane)
LB_D int decode256(Rc &rc, uint32_t *cnt, int lane)
{
    uint4 *mine = (uint4 *)cnt + 2 * lane;
    const uint4 a = mine[0], b = mine[1];
    const uint32_t c[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t sum = (a.x + a.y) + (a.z + a.w) + (b.x + b.y) + (b.z + b.w);
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    rc.range /= total;
    const uint64_t count = (rc.code - rc.low) / rc.range;
    const uint32_t excl = incl - sum;
    const unsigned bal = __ballot_sync(0xffffffffu, (uint64_t)excl <= count);      // lane 0 always votes
    const int owner = 31 - __clz((int)bal);
    uint64_t s = excl;
    int j = 0;
#pragma unroll
    for (int i = 0; i < 7; i++)
        if (j == i && s + c[i] <= count) { s += c[i]; j = i + 1; }
    uint32_t w = 1;
#pragma unroll
    for (int i = 0; i < 8; i++)
        if (j == i) w = c[i];
    if (lane == owner) cnt[8 * lane + j] = w + 1;
    const int sym = __shfl_sync(0xffffffffu, 8 * lane + j, owner);
    const uint32_t lo = __shfl_sync(0xffffffffu, (uint32_t)s, owner);
    w = __shfl_sync(0xffffffffu, w, owner);
    rc_consume(rc, lo, w);
    return sym;
}

LB_D uint64_t decode_numeric(Rc &rc, uint32_t *vec, int lane)          // C/CompressionUtils.hpp:114-137
{
    uint64_t value = 0;
    int shift = 0, i = 0, byte;
    do {
        byte = decode256(rc, vec + 256 * (i < NB_NUMERIC ? i : NB_NUMERIC - 1), lane);
        if (shift < 64) value += (uint64_t)(byte & 127) << shift;
        shift += 7;
        i++;
    } while (byte > 127 && i < 64);      // a real value ends within 10 bytes; the bound only stops garbage
    return value;
}

// contains4 (COL/Bloom.hpp:645-798) by one warp: lane c*7+i tests hash i of candidate nucleotide c.  Warp-uniform kmer.
template <typename K> LB_D unsigned warp_res4(const BloomDev &b, K kmer, bool right, const uint64_t *tbl, int lane)
{
    const int k = b.k;
    const K elem = right ? ((kmer << 2) & kmer_mask(K(0), k)) : (kmer >> 2);
    const K hp = bloom_hashpart(elem, k);
    const uint64_t racine = bloom_racine(hp, b);
    const int c = lane / 7, i = lane - c * 7;
    const unsigned first = right ? (unsigned)((elem >> (2 * (k - 1))) & 3) : (unsigned)c;
    const unsigned last = right ? (unsigned)c : (unsigned)(elem & 3);
    uint64_t h = racine + bloom_cano2((first << 2) + last);
    if (i) h += kmer_simplehash16(hp, i, tbl) & 4095;
    unsigned set = 1;
    if (lane < 28) set = (__ldg(b.words + (h >> 5)) >> (h & 31)) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, set);
    unsigned res = 0;
#pragma unroll
    for (int nt = 0; nt < 4; nt++)
        if (((bal >> (7 * nt)) & 0x7fu) == 0x7fu) res |= 1u << nt;
    return res;
}

template <typename K> LB_D K advance(K kmer, unsigned nt, bool right, int k)       // codeSeedBin / codeSeedNT
{
    if (right) return ((kmer << 2) | K((uint64_t)nt)) & kmer_mask(K(0), k);
    return (kmer >> 2) | (K((uint64_t)nt) << (2 * (k - 1)));
}

LB_D char bin2nt(unsigned c) { return (char)((0x4E47544341ull >> (8 * c)) & 0xff); }   // "ACTGN", Leon::bin2nt (C/Leon.cpp:58-66)

// One CTA = one warp; it decodes blocks blockIdx.x, blockIdx.x + gridDim.x, ...  Dynamic shared memory: the simplehash
// table (2 KB) and two bitmaps of map_words words (N positions, error positions of the current read).
template <typename K>
__global__ void __launch_bounds__(32) k_decode_blocks(BloomDev bloom, const uint64_t *__restrict__ tbl_g, const K *__restrict__ anchors,
                                                      uint64_t nanchors, const uint8_t *__restrict__ dna, const BlockIn *__restrict__ blocks,
                                                      uint64_t nblocks, uint32_t *__restrict__ model_ws, uint8_t *__restrict__ out,
                                                      uint64_t *__restrict__ out_len, uint32_t *__restrict__ status, uint32_t max_read,
                                                      uint32_t map_words)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint64_t *tbl = (uint64_t *)smem;
    uint32_t *nmap = (uint32_t *)(smem + 2048), *emap = nmap + map_words;
    const int lane = threadIdx.x;
    for (int i = lane; i < 256; i += 32) tbl[i] = tbl_g[i];
    uint32_t *ws = model_ws + (size_t)blockIdx.x * MODEL_WORDS;
    const int k = bloom.k;

    for (uint64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const BlockIn bi = blocks[blk];
        // DnaDecoder::startBlock: every model back to one occurrence per symbol
        for (size_t i = lane; i < MODEL_WORDS / 4; i += 32) ((uint4 *)ws)[i] = make_uint4(1u, 1u, 1u, 1u);
        uint32_t m_type[2] = {1, 1}, m_bin[2] = {1, 1}, m_rev[2] = {1, 1};
        uint32_t m_noanchor[5] = {1, 1, 1, 1, 1}, m_bif[5] = {1, 1, 1, 1, 1};
        __syncwarp();
        Rc rc;
        rc_start(rc, dna + bi.in_off, bi.in_bytes);
        uint8_t *o = out + bi.out_off;
        uint32_t err = DE_OK;

        // one step of DnaDecoder::extendAnchor (C/DnaCoder.cpp:1482-1605): the character at `pos`, kmer advanced
        auto extend = [&](K &kmer, uint32_t pos, bool right) -> char {
            if ((nmap[pos >> 5] >> (pos & 31)) & 1u) {                  // an N advances the walk as 'A'
                kmer = advance(kmer, 0u, right, k);
                return 'N';
            }
            if ((emap[pos >> 5] >> (pos & 31)) & 1u) {
                const int nt = decode_small<5>(rc, m_bif);
                const unsigned res4 = warp_res4<K>(bloom, kmer, right, tbl, lane);
                if (res4) kmer = advance(kmer, (unsigned)(__ffs((int)res4) - 1), right, k);   // the walk follows the graph
                else err = DE_NOSUCC;
                return bin2nt((unsigned)nt);
            }
            const unsigned res4 = warp_res4<K>(bloom, kmer, right, tbl, lane);
            const int cnt = __popc(res4);
            unsigned nt;
            if (cnt == 1) nt = (unsigned)(__ffs((int)res4) - 1);
            else if (cnt == 2) {
                const unsigned nt1 = (unsigned)(__ffs((int)res4) - 1), nt2 = (unsigned)(__ffs((int)(res4 & (res4 - 1))) - 1);
                nt = decode_small<2>(rc, m_bin) == 0 ? nt1 : nt2;
            } else nt = (unsigned)decode_small<5>(rc, m_bif);
            if (nt > 3) { kmer = advance(kmer, 0u, right, k); return 'N'; }      // Leon::nt2bin('N') = 0 in codeSeedNT
            kmer = advance(kmer, nt, right, k);
            return bin2nt(nt);
        };

        for (uint64_t r = 0; r < bi.nreads && err == DE_OK; r++) {
            const int type = decode_small<2>(rc, m_type);
            if (type == 0) {
                const uint64_t L = decode_numeric(rc, ws + (size_t)V_READSIZE * NB_NUMERIC * 256, lane);
                const uint64_t apos = decode_numeric(rc, ws + (size_t)V_APOS * NB_NUMERIC * 256, lane);
                const uint64_t addr = decode_numeric(rc, ws + (size_t)V_ADDR * NB_NUMERIC * 256, lane);
                if (addr >= nanchors || apos > L || apos + (uint64_t)k > L || L > (uint64_t)max_read) { err = DE_RANGE; break; }
                K anchor = anchors[addr];
                if (decode_small<2>(rc, m_rev) == 1) anchor = kmer_revcomp(anchor, k);
                const uint32_t words = ((uint32_t)L + 31) >> 5;
                for (uint32_t i = lane; i < words; i += 32) { nmap[i] = 0; emap[i] = 0; }
                __syncwarp();
                const uint64_t nn = decode_numeric(rc, ws + (size_t)V_NUMERIC * NB_NUMERIC * 256, lane);
                if (nn > L) { err = DE_NCOUNT; break; }
                uint64_t prev = 0;
                for (uint64_t i = 0; i < nn; i++) {
                    prev += decode_numeric(rc, ws + (size_t)V_NPOS * NB_NUMERIC * 256, lane);
                    if (prev < L && lane == 0) nmap[prev >> 5] |= 1u << (prev & 31);
                }
                const uint64_t ne = decode_numeric(rc, ws + (size_t)V_LEFTERR * NB_NUMERIC * 256, lane);
                if (ne > L) { err = DE_ECOUNT; break; }
                prev = 0;
                for (uint64_t i = 0; i < ne; i++) {
                    prev += decode_numeric(rc, ws + (size_t)V_ERRPOS * NB_NUMERIC * 256, lane);
                    if (prev < L && lane == 0) emap[prev >> 5] |= 1u << (prev & 31);
                }
                __syncwarp();
                for (uint32_t j = lane; j < (uint32_t)k; j += 32) {
                    const uint32_t pos = (uint32_t)apos + j;
                    const char ch = bin2nt((unsigned)((anchor >> (2 * (k - 1 - (int)j))) & 3));
                    o[pos] = ((nmap[pos >> 5] >> (pos & 31)) & 1u) ? 'N' : ch;
                }
                K kmer = anchor;
                for (uint32_t i = (uint32_t)apos; i-- > 0 && err == DE_OK;) {
                    const char ch = extend(kmer, i, false);
                    if (lane == 0) o[i] = ch;
                }
                kmer = anchor;
                for (uint32_t i = (uint32_t)apos + (uint32_t)k; i < (uint32_t)L && err == DE_OK; i++) {
                    const char ch = extend(kmer, i, true);
                    if (lane == 0) o[i] = ch;
                }
                if (lane == 0) o[L] = '\n';
                o += L + 1;
                __syncwarp();
            } else {
                const uint64_t L = decode_numeric(rc, ws + (size_t)V_NOANCHOR_SIZE * NB_NUMERIC * 256, lane);
                if (L > (uint64_t)max_read) { err = DE_NOANCHOR_LEN; break; }
                for (uint64_t i = 0; i < L; i++) {
                    const char ch = bin2nt((unsigned)decode_small<5>(rc, m_noanchor));
                    if (lane == 0) o[i] = ch;
                }
                if (lane == 0) o[L] = '\n';
                o += L + 1;
            }
        }
        if (lane == 0) {
            out_len[blk] = (uint64_t)(o - (out + bi.out_off));
            status[blk] = err;
        }
        __syncwarp();
    }
}

}  // namespace lbdd
