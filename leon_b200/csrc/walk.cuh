// Phase 3b — per-read de Bruijn walk (C/DnaCoder.cpp:599-757 encodeAnchorRead, :759-969 buildBifurcationList) and
// the lossy quality smoothing (C/DnaCoder.cpp:428-516).  One warp per read; each extension step is one
// warp-cooperative contains4 (28 probes = one gather, see bloom.cuh).
#pragma once
#include "common.cuh"
#include "kmer_math.cuh"
#include "bloom.cuh"
#include "anchors.cuh"

namespace lbwalk {

using lbcount::pack32;
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

template <typename K> struct WalkState {
    K kmer;
    uint32_t nBif, nErr;
    uint8_t *bif;
    uint32_t *err;
};

// one extension step; everything is warp-uniform except the probes inside warp_contains4
template <typename K>
LB_D void walk_step(WalkState<K> &w, uint8_t c, uint32_t pos, bool right, const BloomDev &b, const uint64_t *tbl, int lane)
{
    const int k = b.k;
    if (c == 'N') { w.kmer = seed_step(w.kmer, 0u, right, k); return; }   // N was rewritten to A, nothing is emitted
    unsigned ntb = leon_nt2bin(c);
    unsigned res4 = lbbloom::warp_contains4<K>(b, w.kmer, right, tbl, lane);
    int cnt = __popc(res4);
    unsigned first = res4 ? (unsigned)(__ffs(res4) - 1) : 0u;
    bool solid = (res4 >> ntb) & 1u;
    unsigned nt = ntb;
    int sym = -1;          // symbol to emit: bit 7 = binary model
    bool err = false;
    if (solid) {
        if (cnt == 1) nt = first;
        else if (cnt == 2) sym = 0x80 | (ntb == first ? 0 : 1);
        else sym = (int)ntb;
    } else {
        sym = (int)ntb;
        if (cnt == 1 || cnt == 2) { err = true; nt = first; }
    }
    if (sym >= 0) { if (lane == 0) w.bif[w.nBif] = (uint8_t)sym; w.nBif++; }
    if (err) { if (lane == 0) w.err[w.nErr] = pos; w.nErr++; }
    w.kmer = seed_step(w.kmer, nt, right, k);
}

template <typename K>
__global__ void __launch_bounds__(256) k_walk(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ offsets,
                                              const uint64_t *__restrict__ koff, uint64_t nreads, BloomDev b,
                                              const uint64_t *__restrict__ tbl_g, const int32_t *__restrict__ apos,
                                              uint8_t *__restrict__ bif, uint32_t *__restrict__ err, ReadRec *__restrict__ recs)
{
    __shared__ uint64_t tbl[256];
    tbl[threadIdx.x] = tbl_g[threadIdx.x];
    __syncthreads();
    const int k = b.k;
    const int lane = threadIdx.x & 31;
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t r = warp; r < nreads; r += nwarps) {
        uint64_t o = offsets[r], L = offsets[r + 1] - o;
        const uint8_t *s = bases + o;
        uint32_t nN = 0;
        for (uint64_t base = 0; base < L; base += 32) {
            uint64_t i = base + lane;
            nN += __popc(__ballot_sync(0xffffffffu, i < L && s[i] == 'N'));
        }
        ReadRec rec = {nN, 0, 0, 0, 0};
        int32_t ap = apos[r];
        if (ap >= 0 && L >= (uint64_t)k) {
            uint64_t i0 = (uint64_t)ap + lane, i1 = (uint64_t)ap + 32 + lane;
            uint64_t w0 = pack32(i0 < L ? base_code_n_as_a(s[i0]) : 0u);
            uint64_t w1 = pack32(i1 < L ? base_code_n_as_a(s[i1]) : 0u);
            K anchor = window_kmer<K>(w0, w1, 0, 0, k);
            rec.revcomp = anchor != kmer_canon(anchor, k);
            WalkState<K> w{anchor, 0, 0, bif + koff[r], err + koff[r]};
            // left walk: positions ap-1 .. 0
            for (int64_t i = (int64_t)ap - 1; i >= 0; i -= 32) {
                int64_t mine = i - lane;
                uint8_t cb = mine >= 0 ? s[mine] : (uint8_t)0;
                int steps = i + 1 < 32 ? (int)(i + 1) : 32;
                for (int t = 0; t < steps; t++) {
                    uint8_t c = (uint8_t)__shfl_sync(0xffffffffu, (unsigned)cb, t);
                    walk_step<K>(w, c, (uint32_t)(i - t), false, b, tbl, lane);
                }
            }
            rec.nErrLeft = w.nErr;
            // right walk: positions ap+k .. L-1
            w.kmer = anchor;
            for (uint64_t i = (uint64_t)ap + k; i < L; i += 32) {
                uint64_t mine = i + lane;
                uint8_t cb = mine < L ? s[mine] : (uint8_t)0;
                int steps = L - i < 32 ? (int)(L - i) : 32;
                for (int t = 0; t < steps; t++) {
                    uint8_t c = (uint8_t)__shfl_sync(0xffffffffu, (unsigned)cb, t);
                    walk_step<K>(w, c, (uint32_t)(i + t), true, b, tbl, lane);
                }
            }
            rec.nErr = w.nErr;
            rec.nBif = w.nBif;
        }
        if (lane == 0) recs[r] = rec;
    }
}

// ---- qualities -----------------------------------------------------------------------------------------------
// nb_solids[p] = number of Bloom-positive (canonical) k-mers of the read covering position p
// (storeSolidCoverageInfo, C/DnaCoder.cpp:490-516), then smoothQuals/apply_smoothing_at_pos (:428-486).
// Output text: quality string + '\n' per read at qout[offsets[r] + r].  `pref` = scratch, one u32 per k-mer slot.
template <typename K> struct SolidFn {
    const BloomDev &b; const uint64_t *tbl; int k; int lane;
    uint32_t carry; uint32_t *pref;
    LB_D bool operator()(uint64_t p, K kmer, bool active)
    {
        bool solid = active && bloom_contains<K>(b, kmer_canon(kmer, k), tbl);
        unsigned bal = __ballot_sync(0xffffffffu, solid);
        if (active) pref[p] = carry + __popc(bal & (0xffffffffu >> (31 - lane)));   // inclusive prefix count
        carry += __popc(bal);
        return false;
    }
};

template <typename K>
__global__ void __launch_bounds__(256) k_quals(const uint8_t *__restrict__ bases, const uint8_t *__restrict__ quals,
                                               const uint64_t *__restrict__ offsets, const uint64_t *__restrict__ koff,
                                               uint64_t nreads, BloomDev b, const uint64_t *__restrict__ tbl_g, int lossless,
                                               uint32_t *__restrict__ pref_all, uint8_t *__restrict__ qout)
{
    __shared__ uint64_t tbl[256];
    tbl[threadIdx.x] = tbl_g[threadIdx.x];
    __syncthreads();
    const int k = b.k;
    const int lane = threadIdx.x & 31;
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t r = warp; r < nreads; r += nwarps) {
        uint64_t o = offsets[r], L = offsets[r + 1] - o;
        const uint8_t *q = quals + o;
        uint8_t *out = qout + o + r;
        bool smooth = !lossless && L >= (uint64_t)k;
        uint32_t *pref = pref_all + koff[r];
        uint64_t n = 0;
        if (smooth) {
            n = L - k + 1;
            SolidFn<K> f{b, tbl, k, lane, 0, pref};
            lbanchor::warp_for_kmers<K>(bases + o, L, k, 0, n, lane, f);
            __syncwarp();
        }
        for (uint64_t p = lane; p < L; p += 32) {
            uint8_t c = q[p];
            if (smooth) {
                uint64_t hi_i = p < n ? p : n - 1;
                uint32_t nb = pref[hi_i] - (p >= (uint64_t)k ? pref[p - k] : 0u);
                if (nb >= 2 || c > (uint8_t)'@') {
                    int phred = (int)c - 33;
                    if (!(phred == 0 || phred == 2)) {
                        int diff = (int)'@' - (int)c;
                        if (!(diff > 10 && !((int)nb > diff - 5))) c = '@';
                    }
                }
            }
            out[p] = c;
        }
        if (lane == 0) out[L] = '\n';
        __syncwarp();
    }
}

}  // namespace lbwalk
