// Phase 3b — per-read de Bruijn walk (C/DnaCoder.cpp:599-757 encodeAnchorRead, :759-969 buildBifurcationList) and
// the lossy quality smoothing (C/DnaCoder.cpp:428-516).  One THREAD per read (the bases of a CTA's reads staged in shared
// memory by one TMA bulk copy); each extension step is thread_contains4: 7 independent 64-bit probes (+ a predicated spill
// load) that cover the 14-bit field of the four candidates of every hash.
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"
#include "bloom.cuh"
#include "anchors.cuh"

namespace lbwalk {

using lbcount::pack32;
using lbcount::stage_reads;
using lbcount::stage_reads_begin;
using lbcount::stage_reads_wait;
using lbcount::window_kmer;

// what the walk leaves per read for the entropy stage
struct ReadRec {
    uint32_t nN;        // number of 'N' in the read
    uint32_t nErr;      // recorded error positions (both walks)
    uint32_t nErrLeft;  // how many of them came from the left walk (stored in descending order)
    uint32_t nBif;      // bifurcation symbols in walk order (left walk first)
    uint32_t revcomp;   // anchor k-mer of the read is the reverse complement of its canonical form
};

template <typename K> LB_D K seed_step(K kmer, unsigned nt, bool right, int k)  // C/DnaCoder.cpp:118-136
{
    if (right) return ((kmer << 2) | (K)nt) & kmer_mask(K(0), k);
    return (kmer >> 2) | ((K)nt << (2 * (k - 1)));
}

// ---- Bloom neighbourhood query, one thread per query ------------------------------------------------------------
// contains4 of COL/Bloom.hpp:645-798 for the k-mer `fwd` (with `rc` its reverse complement, kept incrementally by the
// caller).  The 28 tested bits are racine + cano2(c) + {0, simplehash16(i) & 4095}: for each of the 7 hashes the four
// candidates sit inside one 14-bit field, so a query is 7 independent 2-word gathers (no early exit: in a walk one of
// the four neighbours is almost always present and needs all 7).  tbl12 = low 12 bits of the simplehash16 table.
template <typename K>
LB_D unsigned thread_contains4(const BloomDev &b, K fwd, K rc, bool right, const uint16_t *tbl12)
{
    const int k = b.k;
    const K m2 = kmer_mask(K(0), k - 2);
    K mid = right ? (fwd & m2) : (fwd >> 4);
    K rmid = right ? (rc >> 4) : (rc & m2);
    K hp = rmid < mid ? rmid : mid;
    uint64_t racine = bloom_racine(hp, b);
    // cano2 of (first nt, last nt) for the four candidates, 4 bits each
    unsigned fixed = right ? (unsigned)((fwd >> (2 * (k - 2))) & 3) : (unsigned)((fwd >> 2) & 3);
    unsigned p0, p1, p2, p3;
    if (right) {
        p0 = bloom_cano2((fixed << 2) + 0); p1 = bloom_cano2((fixed << 2) + 1);
        p2 = bloom_cano2((fixed << 2) + 2); p3 = bloom_cano2((fixed << 2) + 3);
    } else {
        p0 = bloom_cano2(0 + fixed); p1 = bloom_cano2(4 + fixed); p2 = bloom_cano2(8 + fixed); p3 = bloom_cano2(12 + fixed);
    }
    const uint64_t hlow = kmer_low64(hp);
    const unsigned t0 = sizeof(K) == 8 ? tbl12[hlow & 255] : 0u;     // third table term of the one-word variant
    // Per hash: the aligned 64-bit word holding bit racine + off, plus (13 positions out of 64) the low half of the next
    // one when the 14-bit field of the four candidates runs past it: 7 LDG.64 and 7 mostly predicated-off LDG per step
    // instead of 14 LDG (the kernel is L1-request bound when the filter sits in L2).
    uint32_t lo[7], hi[7], nx[7];
    unsigned sh[7];
#pragma unroll
    for (int i = 0; i < 7; i++) {
        unsigned off = 0;
        if (i) off = (tbl12[(hlow >> i) & 255] ^ tbl12[(hlow >> (i + 8)) & 255] ^ t0) & 4095u;
        uint64_t bit = racine + off;
        const uint2 *w = (const uint2 *)b.words + (bit >> 6);
        sh[i] = (unsigned)bit & 63u;
        const uint2 q = bloom_ldw2(w);
        lo[i] = q.x;
        hi[i] = q.y;
        nx[i] = 0;
        if (sh[i] > 50u) nx[i] = bloom_ldw((const uint32_t *)(w + 1));
    }
    // All loads of a step must be in flight together: a warp issues in order, and ptxas otherwise puts each funnel shift
    // right behind its own loads (5 dependent round trips per step; 3x slower where the probes go to DRAM).  z is always 0
    // (nbytes < 2^63) but only at run time, and it depends on every loaded word, so no shift can issue before the last load.
    uint32_t any = 0;
#pragma unroll
    for (int i = 0; i < 7; i++) any |= lo[i] | hi[i] | nx[i];
    const uint32_t z = any & (uint32_t)(b.nbytes >> 63);
    unsigned alive = 15u;
#pragma unroll
    for (int i = 0; i < 7; i++) {
        const bool up = sh[i] >= 32u;
        const uint32_t f = __funnelshift_r(up ? hi[i] : lo[i], up ? nx[i] : hi[i], (sh[i] & 31u) | z);   // filter bits from racine + off on
        alive &= ((f >> p0) & 1u) | (((f >> p1) & 1u) << 1) | (((f >> p2) & 1u) << 2) | (((f >> p3) & 1u) << 3);
    }
    return alive;
}

// One thread per read.  Left walk (positions anchor-1 .. 0) then right walk (anchor+k .. L-1), as one loop of L-k
// direction-generic steps so that the threads of a warp stay converged whatever their anchor positions.
template <typename K>
__global__ void __launch_bounds__(256, sizeof(K) == 8 ? 4 : 3) k_walk(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                              const uint64_t *__restrict__ koff, uint64_t nreads, BloomDev b,
                                              const uint64_t *__restrict__ tbl_g, const int32_t *__restrict__ apos,
                                              uint8_t *__restrict__ bif, uint32_t *__restrict__ err, ReadRec *__restrict__ recs,
                                              uint32_t stage_bytes)
{
    extern __shared__ __align__(128) uint8_t stage[];
    __shared__ uint16_t tbl12[256];
    __shared__ __align__(8) uint64_t bar;
    tbl12[threadIdx.x] = (uint16_t)(tbl_g[threadIdx.x] & 4095u);
    const int k = b.k;
    const uint64_t r0 = (uint64_t)blockIdx.x * 256;
    const uint64_t r1 = r0 + 256 < nreads ? r0 + 256 : nreads;
    const uint64_t o0 = offsets[r0];
    const uint8_t *sbase = stage_reads(bases, o0, offsets[r1], stage, stage_bytes, &bar);
    __syncthreads();
    const uint64_t r = r0 + threadIdx.x;
    if (r >= nreads) return;
    const uint64_t o = offsets[r], L = offsets[r + 1] - o;
    const uint8_t *s = sbase ? sbase + (o - o0) : bases + o;
    uint32_t nN = 0;
    for (uint64_t i = 0; i < L; i++) nN += s[i] == 'N';
    ReadRec rec = {nN, 0, 0, 0, 0};
    const int32_t ap = apos[r];
    if (ap >= 0 && L >= (uint64_t)k) {
        K anchor = K(0);
        for (int i = 0; i < k; i++) anchor = (anchor << 2) | K((uint64_t)base_code_n_as_a(s[ap + i]));
        const K arc = kmer_revcomp(anchor, k);
        rec.revcomp = arc < anchor;
        const K mask = kmer_mask(K(0), k);
        const int top = 2 * (k - 1);
        K fwd = anchor, rc = arc;
        const uint64_t kslot = koff[r] - koff[0];     // bif / err hold the k-mer slots of the super-block koff points at
        uint8_t *bp = bif + kslot;
        uint32_t *ep = err + kslot;
        uint32_t nBif = 0, nErr = 0;
        const uint32_t nsteps = (uint32_t)(L - k), nleft = (uint32_t)ap;
        for (uint32_t t = 0; t < nsteps; t++) {
            const bool right = t >= nleft;
            if (t == nleft) { fwd = anchor; rc = arc; rec.nErrLeft = nErr; }
            const uint32_t pos = right ? (uint32_t)ap + k + (t - nleft) : nleft - 1 - t;
            const uint8_t c = s[pos];
            unsigned nt = 0;
            if (c != 'N') {   // at an N the walk advances with 'A' and emits nothing (C/DnaCoder.cpp:772-790)
                const unsigned ntb = leon_nt2bin(c);
                const unsigned res4 = thread_contains4<K>(b, fwd, rc, right, tbl12);
                const int cnt = __popc(res4);
                const unsigned first = res4 ? (unsigned)(__ffs(res4) - 1) : 0u;
                const bool solid = (res4 >> ntb) & 1u;
                nt = ntb;
                int sym = -1;          // symbol to emit: bit 7 = binary model
                bool e = false;
                if (solid) {
                    if (cnt == 2) sym = 0x80 | (ntb == first ? 0 : 1);
                    else if (cnt > 2) sym = (int)ntb;
                } else {
                    sym = (int)ntb;
                    if (cnt == 1 || cnt == 2) { e = true; nt = first; }
                }
                if (sym >= 0) bp[nBif++] = (uint8_t)sym;
                if (e) ep[nErr++] = pos;
            }
            if (right) {
                fwd = ((fwd << 2) | K((uint64_t)nt)) & mask;
                rc = (rc >> 2) | (K((uint64_t)(nt ^ 2u)) << top);
            } else {
                fwd = (fwd >> 2) | (K((uint64_t)nt) << top);
                rc = ((rc << 2) | K((uint64_t)(nt ^ 2u))) & mask;
            }
        }
        if (nleft >= nsteps) rec.nErrLeft = nErr;
        rec.nErr = nErr;
        rec.nBif = nBif;
    }
    recs[r] = rec;
}

// ---- qualities -----------------------------------------------------------------------------------------------
// nb_solids[p] = number of Bloom-positive (canonical) k-mers of the read covering position p
// (storeSolidCoverageInfo, C/DnaCoder.cpp:490-516), then smoothQuals/apply_smoothing_at_pos (:428-486).
// Output text: quality string + '\n' per read at qout[offsets[r] + r].  `pref` = scratch, one u32 per k-mer slot.
// One THREAD per read (like k_walk): the bases of the CTA's QT reads are staged in shared memory by one bulk copy and
// their qualities are copied straight to where the output text of the CTA (one contiguous range of qout, a newline
// after every read) is assembled in shared memory; the canonical k-mer is rolled base by base, a k-bit shift register
// of the last k contains() answers gives nb_solids at every position, the qualities are smoothed in place and the
// text is written out with 16-byte stores.  stage_bytes = size of the bases area, the text area has stage_bytes + QT
// + 32.  The 7 Bloom probes of QB consecutive k-mers are issued together, branch-free, so that QB hash chains
// interleave and 7 * QB loads are in flight per thread (most k-mers of a read are solid: the early exit of contains()
// would only serialise 7 L2 round trips).
constexpr int QT = 128;
template <typename K, int QB, int MINB>
__global__ void __launch_bounds__(QT, MINB) k_quals(const uint8_t *__restrict__ bases, const uint8_t *__restrict__ quals,
                                               const uint64_t *__restrict__ offsets, uint64_t nreads, BloomDev b,
                                               const uint64_t *__restrict__ tbl_g, int lossless, uint8_t *__restrict__ qout,
                                               uint32_t stage_bytes)
{
    extern __shared__ __align__(128) uint8_t qstage[];
    __shared__ uint64_t tbl[256];
    __shared__ __align__(8) uint64_t bar;
    for (int i = threadIdx.x; i < 256; i += QT) tbl[i] = tbl_g[i];
    const int k = b.k;
    const uint64_t r0 = (uint64_t)blockIdx.x * QT;
    const uint64_t r1 = r0 + QT < nreads ? r0 + QT : nreads;
    const uint64_t o0 = offsets[r0], o1 = offsets[r1];
    // the CTA's text goes to qout[o0 + r0, o1 + r1): staged at the same offset modulo 16
    uint8_t *gout = qout + o0 + r0;
    const uint64_t out_bytes = (o1 - o0) + (r1 - r0);
    const uint32_t mis = (uint32_t)((uintptr_t)gout & 15u);
    // the bulk copy of the bases is in flight while the warps copy the qualities, one read per warp at a time
    const uint8_t *sb = stage_reads_begin(bases, o0, o1, qstage, stage_bytes, &bar);
    uint8_t *so = sb ? qstage + (uint64_t)stage_bytes + mis : nullptr;
    if (so) {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        for (uint64_t t = warp; t < r1 - r0; t += QT / 32) {
            const uint64_t o = offsets[r0 + t], L = offsets[r0 + t + 1] - o;
            uint8_t *dst = so + (o - o0) + t;
            for (uint64_t p = lane; p < L; p += 32) dst[p] = quals[o + p];
        }
        stage_reads_wait(&bar);
    }
    __syncthreads();
    const uint64_t r = r0 + threadIdx.x;
    if (r < r1) {
        const uint64_t o = offsets[r], L = offsets[r + 1] - o;
        const uint8_t *s = sb ? sb + (o - o0) : bases + o;
        uint8_t *out = so ? so + (o - o0) + (r - r0) : qout + o + r;
        const uint8_t *q = so ? out : quals + o;                  // in place when staged
        const bool smooth = !lossless && L >= (uint64_t)k;
        if (!smooth) {
            if (!so)
                for (uint64_t p = 0; p < L; p++) out[p] = q[p];
        } else {
            const K mask = kmer_mask(K(0), k);
            const int top = 2 * (k - 1);
            const uint64_t wmask = (k >= 64) ? ~0ull : ((1ull << k) - 1ull);
            K fwd = K(0), rc = K(0);
            uint64_t window = 0;
            auto roll = [&](uint8_t ch) {
                const unsigned code = base_code_n_as_a(ch);
                fwd = ((fwd << 2) | K((uint64_t)code)) & mask;
                rc = (rc >> 2) | (K((uint64_t)(code ^ 2u)) << top);
            };
            auto emit = [&](uint64_t p, bool solid) {            // all k-mers covering position p are known now
                window = ((window << 1) | (solid ? 1ull : 0ull)) & wmask;
                const int nb = __popcll(window);
                uint8_t c = q[p];
                if (nb >= 2 || c > (uint8_t)'@') {
                    const int phred = (int)c - 33;
                    if (!(phred == 0 || phred == 2)) {
                        const int diff = (int)'@' - (int)c;
                        if (!(diff > 10 && !(nb > diff - 5))) c = '@';
                    }
                }
                out[p] = c;
            };
            for (int i = 0; i < k - 1; i++) roll(s[i]);
            const uint64_t n = L - (uint64_t)k + 1;
            for (uint64_t j0 = 0; j0 < n; j0 += QB) {             // QB k-mers = 7 * QB probes in flight per thread
                uint32_t w[QB][7];
                uint64_t sh[QB];
#pragma unroll
                for (int jj = 0; jj < QB; jj++) {                 // past the last k-mer: the last base again, result unused
                    const uint64_t bi = (uint64_t)k - 1 + j0 + jj;
                    roll(s[bi < L ? bi : L - 1]);
                    bloom_probe_issue<K>(b, fwd, rc, tbl, w[jj], sh[jj]);
                }
#pragma unroll
                for (int jj = 0; jj < QB; jj++)
                    if (j0 + jj < n) emit(j0 + jj, bloom_probe_test(w[jj], sh[jj]));
            }
            for (uint64_t p = n; p < L; p++) emit(p, false);
        }
        out[L] = '\n';
    }
    if (so) {                                                       // CTA-uniform
        __syncthreads();
        // head bytes up to the first 16-byte boundary, 16-byte body, tail bytes
        const uint64_t head = mis ? (16 - mis < out_bytes ? 16 - mis : out_bytes) : 0;
        const uint64_t body = (out_bytes - head) / 16;
        if (threadIdx.x < head) gout[threadIdx.x] = so[threadIdx.x];
        const uint4 *s4 = (const uint4 *)(so + head);
        uint4 *g4 = (uint4 *)(gout + head);
        for (uint64_t i = threadIdx.x; i < body; i += QT) g4[i] = s4[i];
        const uint64_t done = head + body * 16;
        if (threadIdx.x < out_bytes - done) gout[done + threadIdx.x] = so[done + threadIdx.x];
    }
}

}  // namespace lbwalk
