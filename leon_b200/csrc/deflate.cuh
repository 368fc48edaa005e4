// Quality entropy stage on the device: zlib streams of the quality text of every block, written by
// the GPU instead of host zlib.  The reference hands each block's text to deflate at Z_BEST_COMPRESSION
// (C/Leon.cpp:581-607) and its decoder inflates whatever zlib stream it finds (C/Leon.cpp:1630-1690), so any valid
// stream is a drop-in; the bytes are not the reference's (they depend on the zlib version there as well).
//
// Smoothed qualities are long runs of one character broken by a few original values, so the LZ77 stage only looks for
// runs (matches at distance 1, like zlib's Z_RLE strategy, which on this data is within ~4 % of level 9) and the
// entropy stage is a per-chunk dynamic Huffman code (RFC 1951 block type 2).  A block's text is cut into chunks of
// DF_CHUNK bytes, one CTA per chunk; every chunk is a complete deflate block followed by an empty stored block, which
// byte-aligns the stream so that chunks can be concatenated; a second kernel lays the chunks of a block end to end
// between the 2-byte zlib header and the Adler-32 of the text.
//
// Three kernels per run, so that the serial part of a chunk (Huffman construction, ~10^4 dependent instructions) is
// done by ONE THREAD PER CHUNK, thousands of chunks at once, instead of stalling a CTA:
//   k_deflate_count  CTA per chunk: run structure of the text, symbol frequencies, Adler-32 pieces
//   k_deflate_codes  thread per chunk: code lengths, canonical codes, dynamic-block header
//   k_deflate_emit   CTA per chunk: bits of every thread's tokens, prefix sum, bits into the chunk's slot
// then k_deflate_layout / k_deflate_gather lay the chunks of a block end to end behind the zlib header.
//
// The per-thread phases are plain functions over small state structs so that tools/deflate_sim.cpp can run the very
// same code on the host, thread by thread, as a development check; the library only ever runs them in the kernels.
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define DF_HD __host__ __device__ __forceinline__
#else
#define DF_HD inline
#endif

namespace lbdeflate {

constexpr int DF_THREADS = 256;
constexpr int DF_SLICE = 252;                       // bytes per thread (63 words)
constexpr int DF_CHUNK = DF_THREADS * DF_SLICE;     // 64512 <= 65535, so a chunk also fits one stored block
constexpr int DF_SLOT = DF_CHUNK + 64;              // bytes reserved per chunk in the scratch buffer
constexpr int DF_NLIT = 286, DF_NCL = 19;
constexpr uint32_t DF_ADLER = 65521u;
constexpr int DF_HDR_WORDS = 160;                   // dynamic block header: <= 14 + 19 * 3 + 287 * 14 bits
constexpr int DF_EQ_WORDS = 8;                      // 252 "equals its predecessor" bits per slice
constexpr int DF_CODES_CHUNKS = 8;                  // k_deflate_codes: chunks (warps) per CTA

struct ChunkMeta {            // per chunk
    uint32_t nbytes;          // compressed bytes in the slot
    uint32_t sum_a, sum_b;    // byte sum and position-weighted sum of the chunk's text (Adler-32 pieces), mod 65521
    uint32_t n;               // text bytes
};

// Per chunk, in global memory: what k_deflate_count hands to k_deflate_codes and that one to k_deflate_emit.
struct HuffState {
    uint32_t freq[DF_NLIT + 2];
    uint32_t nmatch;
    uint32_t lc[DF_NLIT];                      // code | length << 16
    uint32_t hdr_w[DF_HDR_WORDS];
    uint32_t hdr_bits;
    uint16_t lcode[DF_NLIT];   uint8_t llen[DF_NLIT];
    uint16_t clcode[DF_NCL];   uint8_t cllen[DF_NCL];
    // scratch of the Huffman construction
    uint32_t node_w[2 * DF_NLIT];
    uint16_t node_parent[2 * DF_NLIT], order[DF_NLIT];
    uint8_t  node_depth[2 * DF_NLIT];
    uint8_t  cl_sym[DF_NLIT + 34], cl_extra[DF_NLIT + 34];
};

// Per CTA, in shared memory: the run structure of the chunk's 256 slices.
struct ChunkState {
    const uint8_t *text;                       // the chunk's text (global memory)
    uint32_t eqm[DF_EQ_WORDS * DF_THREADS];    // [word][thread]
    uint32_t freq[DF_NLIT + 2];
    uint32_t nmatch;
    uint32_t lc[DF_NLIT];
    uint8_t  first[DF_THREADS], last[DF_THREADS];
    uint16_t slen[DF_THREADS], lead[DF_THREADS], trail[DF_THREADS];
    uint32_t bits[DF_THREADS];
    uint32_t sa[DF_THREADS], sb[DF_THREADS];
    uint32_t hdr_bits, end_bits, stored;
};

DF_HD int df_min(int a, int b) { return a < b ? a : b; }
DF_HD uint32_t df_byte(const ChunkState &S, int i) { return S.text[i]; }

// Word k of the bytes at p (any alignment).  The device reads two aligned words (the text is padded by 8 bytes).
DF_HD uint32_t df_word(const uint8_t *p, int k) {
#ifdef __CUDA_ARCH__
    const uintptr_t a = (uintptr_t)p;
    const uint32_t *g = (const uint32_t *)(a & ~(uintptr_t)3) + k;
    const int sh = (int)(a & 3) * 8;
    return sh ? __funnelshift_r(g[0], g[1], sh) : g[0];
#else
    uint32_t w;
    memcpy(&w, p + 4 * k, 4);
    return w;
#endif
}

// byte-wise helpers: device intrinsics, plain C on the host (tools/deflate_sim.cpp)
DF_HD uint32_t df_cmpeq4(uint32_t a, uint32_t b) {          // 0xff in every byte where a and b agree
#ifdef __CUDA_ARCH__
    return __vcmpeq4(a, b);
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) if (((a >> (8 * i)) & 255u) == ((b >> (8 * i)) & 255u)) r |= 255u << (8 * i);
    return r;
#endif
}
DF_HD uint32_t df_bytesum(uint32_t w) {
#ifdef __CUDA_ARCH__
    return __vsadu4(w, 0u);
#else
    return (w & 255u) + ((w >> 8) & 255u) + ((w >> 16) & 255u) + (w >> 24);
#endif
}
DF_HD uint32_t df_dot0123(uint32_t w) {                     // 0 * b0 + 1 * b1 + 2 * b2 + 3 * b3
#ifdef __CUDA_ARCH__
    return __dp4a(w, 0x03020100u, 0u);
#else
    return ((w >> 8) & 255u) + 2u * ((w >> 16) & 255u) + 3u * (w >> 24);
#endif
}
DF_HD int df_ffs(uint32_t w) {                              // 1-based position of the lowest set bit, w != 0
#ifdef __CUDA_ARCH__
    return __ffs((int)w);
#else
    return __builtin_ffs((int)w);
#endif
}
DF_HD int df_clz(uint32_t w) {
#ifdef __CUDA_ARCH__
    return __clz((int)w);
#else
    return w ? __builtin_clz(w) : 32;
#endif
}

// RFC 1951 3.2.5: length 3..258 -> symbol 257..285, extra bits
DF_HD void df_len_code(int len, int *sym, int *ebits, int *eval) {
    if (len == 258) { *sym = 285; *ebits = 0; *eval = 0; return; }
    int l = len - 3;
    if (l < 8) { *sym = 257 + l; *ebits = 0; *eval = 0; return; }
    const int e = 29 - df_clz((uint32_t)l);                 // floor(log2(l)) - 2
    *ebits = e;
    *sym = 257 + 4 * e + 4 + ((l >> e) & 3);
    *eval = l & ((1 << e) - 1);
}

DF_HD uint32_t df_reverse(uint32_t code, int len) {
    uint32_t r = 0;
    for (int i = 0; i < len; i++) { r = (r << 1) | (code & 1); code >>= 1; }
    return r;
}

DF_HD uint32_t &df_eq(ChunkState &S, int t, int w) { return S.eqm[w * DF_THREADS + t]; }
DF_HD uint32_t df_eq(const ChunkState &S, int t, int w) { return S.eqm[w * DF_THREADS + t]; }

// First position >= pos whose eq bit is clear (a run start), or len.
DF_HD int df_next_start(const ChunkState &S, int t, int pos, int len) {
    while (pos < len) {
        const uint32_t w = ~df_eq(S, t, pos >> 5) >> (pos & 31);
        if (w) { pos += df_ffs(w) - 1; return pos < len ? pos : len; }
        pos = (pos | 31) + 1;
    }
    return len;
}

// ---- phase 1: every thread summarises its slice, a word at a time ---------------------------------------------------
// eq bit i (i >= 1) = byte i of the slice equals byte i - 1; sums for the Adler-32; first / last byte and the lengths of
// the first and last run, through which runs are followed across slices.
DF_HD void df_summary(ChunkState &S, int t, int n) {
    const int base = t * DF_SLICE;
    int len = n - base;
    len = len > DF_SLICE ? DF_SLICE : (len < 0 ? 0 : len);
    S.slen[t] = (uint16_t)len;
    uint32_t a = 0, b = 0, prevb = 0;
    const int nw = (len + 3) >> 2;
    const uint8_t *p = S.text + base;
    for (int q = 0; q < DF_EQ_WORDS; q++) {                 // 8 words of text -> one word of eq bits
        uint32_t cur = 0;
        for (int kk = 0; kk < 8; kk++) {
            const int k = 8 * q + kk;
            if (k >= nw) break;
            uint32_t w = df_word(p, k);
            const int rem = len - 4 * k;
            if (rem < 4) w &= (1u << (8 * rem)) - 1u;
            const uint32_t m = df_cmpeq4(w, (w << 8) | prevb);
            uint32_t nib = (((m & 0x01010101u) * 0x01020408u) >> 24) & 15u;
            if (k == 0) nib &= ~1u;
            if (rem < 4) nib &= (1u << rem) - 1u;
            cur |= nib << (4 * kk);
            prevb = w >> 24;
            const uint32_t sum = df_bytesum(w);
            a += sum;
            b += (uint32_t)(len - 4 * k) * sum - df_dot0123(w);
        }
        df_eq(S, t, q) = cur;
    }
    S.sa[t] = a; S.sb[t] = b;
    if (len) {
        S.first[t] = (uint8_t)df_byte(S, base);
        S.last[t] = (uint8_t)df_byte(S, base + len - 1);
        const int lead = df_next_start(S, t, 1, len);        // bytes equal to the first one
        int trail = 1;
        while (trail < len && ((df_eq(S, t, (len - trail) >> 5) >> ((len - trail) & 31)) & 1u)) trail++;
        S.lead[t] = (uint16_t)lead; S.trail[t] = (uint16_t)trail;
    }
}

// Tokens of the part [j0, j0 + cnt) of a run of L equal bytes: position 0 is a literal, the rest is cut into matches of
// up to 258 bytes at distance 1; a tail shorter than 3 becomes literals.
template <class Sink>
DF_HD void df_run_tokens(int j0, int cnt, int L, uint32_t v, Sink &sink) {
    int j = j0;
    const int jend = j0 + cnt, rem = L - 1;
    if (j == 0) { sink.literal(v); j = 1; }
    while (j < jend) {
        const int seg = (j - 1) / 258, off = (j - 1) - seg * 258;
        const int seglen = df_min(258, rem - seg * 258);
        if (seglen >= 3) {
            if (off == 0) sink.match(seglen);
            j += seglen - off;
        } else {
            sink.literal(v);
            j++;
        }
    }
}

// Walks the runs of slice t (extended into the neighbouring slices through the summaries) and feeds the sink.
template <class Sink>
DF_HD void df_slice_tokens(const ChunkState &S, int t, int nact, Sink &sink) {
    const int len = S.slen[t], base = t * DF_SLICE;
    if (!len) return;
    int back = 0, fwd = 0;
    for (int p = t - 1; p >= 0; p--) {
        if (S.last[p] != S.first[t]) break;
        back += S.trail[p];
        if (S.trail[p] != S.slen[p]) break;
    }
    for (int p = t + 1; p < nact; p++) {
        if (S.first[p] != S.last[t]) break;
        fwd += S.lead[p];
        if (S.lead[p] != S.slen[p]) break;
    }
    int i = 0;
    while (i < len) {
        const uint32_t v = df_byte(S, base + i);
        const int e = df_next_start(S, t, i + 1, len);
        const int j0 = i == 0 ? back : 0;
        const int L = (e - i) + j0 + (e == len ? fwd : 0);
        if (L == 1) sink.literal(v);
        else df_run_tokens(j0, e - i, L, v, sink);
        i = e;
    }
}

#ifdef __CUDA_ARCH__
#define DF_ATOMIC_ADD(p, v) atomicAdd((p), (v))
#define DF_ATOMIC_OR(p, v) atomicOr((p), (v))
#else
#define DF_ATOMIC_ADD(p, v) (*(p) += (v))
#define DF_ATOMIC_OR(p, v) (*(p) |= (v))
#endif

struct CountSink {            // phase 2: symbol frequencies
    ChunkState &S;
    DF_HD void literal(uint32_t v) { DF_ATOMIC_ADD(&S.freq[v], 1u); }
    DF_HD void match(int len) {
        int sym, eb, ev;
        df_len_code(len, &sym, &eb, &ev);
        DF_ATOMIC_ADD(&S.freq[sym], 1u);
        DF_ATOMIC_ADD(&S.nmatch, 1u);
    }
};

struct BitCountSink {         // phase 4: bits of a thread's tokens
    const ChunkState &S;
    uint32_t bits;
    DF_HD void literal(uint32_t v) { bits += S.lc[v] >> 16; }
    DF_HD void match(int len) {
        int sym, eb, ev;
        df_len_code(len, &sym, &eb, &ev);
        bits += (S.lc[sym] >> 16) + eb + 1;     // + the 1-bit distance code
    }
};

// LSB-first bit packing into zero-initialised words.  A thread owns the bits [start, end) of the stream: the words in
// the middle are its own (plain stores), the first and the last one are shared with its neighbours (atomic OR).
struct BitWriter {
    uint32_t *w;
    uint64_t acc;
    int nbits;
    uint32_t word;
    bool shared;
    DF_HD void start(uint32_t *out, uint32_t bitpos) { w = out; acc = 0; nbits = (int)(bitpos & 31u); word = bitpos >> 5; shared = true; }
    DF_HD void put(uint32_t v, int n) {
        acc |= (uint64_t)v << nbits;
        nbits += n;
        if (nbits >= 32) {
            if (shared) { DF_ATOMIC_OR(&w[word], (uint32_t)acc); shared = false; }
            else w[word] = (uint32_t)acc;
            word++; acc >>= 32; nbits -= 32;
        }
    }
    DF_HD void finish() { if (nbits > 0) DF_ATOMIC_OR(&w[word], (uint32_t)acc); }
    DF_HD uint32_t bitpos() const { return word * 32u + (uint32_t)nbits; }
};

struct EmitSink {             // phase 5: the codes themselves
    const ChunkState &S;
    BitWriter bw;
    DF_HD void literal(uint32_t v) { const uint32_t c = S.lc[v]; bw.put(c & 0xffffu, (int)(c >> 16)); }
    DF_HD void match(int len) {
        int sym, eb, ev;
        df_len_code(len, &sym, &eb, &ev);
        const uint32_t c = S.lc[sym];
        // code, extra bits and the 1-bit distance code 0 (distance 1) in one go: at most 15 + 5 + 1 bits
        bw.put((c & 0xffffu) | ((uint32_t)ev << (c >> 16)), (int)(c >> 16) + eb + 1);
    }
};

// ---- phase 3 (one thread): Huffman code lengths, canonical codes, block header ------------------------------------
// Lengths of an optimal prefix code limited to maxbits: Huffman by two queues over the sorted leaves, depths clipped
// to maxbits, then the Kraft sum repaired by lengthening the deepest leaves that are still shorter than maxbits.
DF_HD void df_code_lengths(HuffState &S, const uint32_t *freq, int n, int maxbits, uint8_t *len) {
    int m = 0;
    for (int s = 0; s < n; s++) {
        len[s] = 0;
        if (freq[s]) {                          // insertion sort by (frequency, symbol)
            int k = m++;
            while (k > 0 && freq[S.order[k - 1]] > freq[s]) { S.order[k] = S.order[k - 1]; k--; }
            S.order[k] = (uint16_t)s;
        }
    }
    if (m == 0) return;
    if (m == 1) { len[S.order[0]] = 1; return; }
    for (int k = 0; k < m; k++) S.node_w[k] = freq[S.order[k]];
    int leaf = 0, inner = m, next = m;
    for (int k = 0; k < m - 1; k++) {
        uint32_t w = 0;
        for (int pick = 0; pick < 2; pick++) {
            int x;
            if (leaf < m && (inner >= next || S.node_w[leaf] <= S.node_w[inner])) x = leaf++;
            else x = inner++;
            w += S.node_w[x];
            S.node_parent[x] = (uint16_t)next;
        }
        S.node_w[next++] = w;
    }
    const int root = 2 * m - 2;
    S.node_depth[root] = 0;
    for (int x = root - 1; x >= 0; x--) {
        int d = S.node_depth[S.node_parent[x]] + 1;
        S.node_depth[x] = (uint8_t)(d > 60 ? 60 : d);
    }
    uint64_t kraft = 0;                          // in units of 2^-maxbits
    for (int k = 0; k < m; k++) {
        int d = S.node_depth[k] > maxbits ? maxbits : S.node_depth[k];
        len[S.order[k]] = (uint8_t)d;
        kraft += 1ull << (maxbits - d);
    }
    const uint64_t one = 1ull << maxbits;
    while (kraft > one) {                        // lengthen the least frequent leaf among the deepest ones below maxbits
        int best = -1;
        for (int k = 0; k < m; k++) {
            int l = len[S.order[k]];
            if (l < maxbits && (best < 0 || l > len[S.order[best]])) best = k;
        }
        const int l = len[S.order[best]];
        kraft -= 1ull << (maxbits - l - 1);
        len[S.order[best]] = (uint8_t)(l + 1);
    }
    // inflate rejects incomplete codes: hand the slack back until the Kraft sum is exactly one.  The deepest leaf's
    // own weight divides both the sum and one, hence the deficit, so shortening it always fits.
    while (kraft < one) {
        int best = 0;
        for (int k = 1; k < m; k++) if (len[S.order[k]] >= len[S.order[best]]) best = k;   // deepest, most frequent
        const int l = len[S.order[best]];
        kraft += 1ull << (maxbits - l);
        len[S.order[best]] = (uint8_t)(l - 1);
    }
}

DF_HD void df_canonical(const uint8_t *len, int n, int maxbits, uint16_t *code) {
    uint32_t count[16], next[16];
    for (int b = 0; b <= maxbits; b++) count[b] = 0;
    for (int s = 0; s < n; s++) count[len[s]]++;
    count[0] = 0;
    uint32_t c = 0;
    for (int b = 1; b <= maxbits; b++) { c = (c + count[b - 1]) << 1; next[b] = c; }
    for (int s = 0; s < n; s++) code[s] = len[s] ? (uint16_t)df_reverse(next[len[s]]++, len[s]) : 0;
}

// One thread per chunk: codes from the frequencies, then the dynamic-block header into hdr_w (RFC 1951 3.2.7).
DF_HD void df_build_block_header(HuffState &S) {
    S.freq[256] = 1;                                           // end of block
    df_code_lengths(S, S.freq, DF_NLIT, 15, S.llen);
    df_canonical(S.llen, DF_NLIT, 15, S.lcode);
    for (int k = 0; k < DF_NLIT; k++) S.lc[k] = (uint32_t)S.lcode[k] | ((uint32_t)S.llen[k] << 16);
    int hlit = DF_NLIT;
    while (hlit > 257 && S.llen[hlit - 1] == 0) hlit--;
    const int dlen0 = S.nmatch ? 1 : 0;                        // one distance code; 0 bits = "no distances at all"
    // run-length code of the hlit + 1 code lengths
    int ntok = 0;
    uint32_t clfreq[DF_NCL];
    for (int s = 0; s < DF_NCL; s++) clfreq[s] = 0;
    const int total = hlit + 1;
    int i = 0;
    while (i < total) {
        const int v = i < hlit ? S.llen[i] : dlen0;
        int c = 1;
        while (i + c < total && (i + c < hlit ? S.llen[i + c] : dlen0) == v) c++;
        i += c;
        if (v == 0) {
            while (c >= 11) { int r = df_min(c, 138); S.cl_sym[ntok] = 18; S.cl_extra[ntok++] = (uint8_t)(r - 11); c -= r; }
            if (c >= 3) { S.cl_sym[ntok] = 17; S.cl_extra[ntok++] = (uint8_t)(c - 3); c = 0; }
            while (c-- > 0) { S.cl_sym[ntok] = 0; S.cl_extra[ntok++] = 0; }
        } else {
            S.cl_sym[ntok] = (uint8_t)v; S.cl_extra[ntok++] = 0; c--;
            while (c >= 3) { int r = df_min(c, 6); S.cl_sym[ntok] = 16; S.cl_extra[ntok++] = (uint8_t)(r - 3); c -= r; }
            while (c-- > 0) { S.cl_sym[ntok] = (uint8_t)v; S.cl_extra[ntok++] = 0; }
        }
    }
    for (int k = 0; k < ntok; k++) clfreq[S.cl_sym[k]]++;
    df_code_lengths(S, clfreq, DF_NCL, 7, S.cllen);
    {                                                          // a lone code-length symbol still needs a complete code
        int used = 0, spare = -1;
        for (int s = 0; s < DF_NCL; s++) { if (S.cllen[s]) used++; else if (spare < 0) spare = s; }
        if (used == 1) S.cllen[spare] = 1;
    }
    df_canonical(S.cllen, DF_NCL, 7, S.clcode);
    const uint8_t order[DF_NCL] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    int hclen = DF_NCL;
    while (hclen > 4 && S.cllen[order[hclen - 1]] == 0) hclen--;
    for (int k = 0; k < DF_HDR_WORDS; k++) S.hdr_w[k] = 0;
    BitWriter bw;
    bw.start(S.hdr_w, 0);
    bw.put(0u, 1);                                             // BFINAL = 0: the stream is closed by the stored block
    bw.put(2u, 2);                                             // BTYPE = 10, dynamic Huffman
    bw.put((uint32_t)(hlit - 257), 5);
    bw.put(0u, 5);                                             // HDIST = 1 code
    bw.put((uint32_t)(hclen - 4), 4);
    for (int k = 0; k < hclen; k++) bw.put(S.cllen[order[k]], 3);
    for (int k = 0; k < ntok; k++) {
        const int s = S.cl_sym[k];
        bw.put(S.clcode[s], S.cllen[s]);
        if (s == 16) bw.put(S.cl_extra[k], 2);
        else if (s == 17) bw.put(S.cl_extra[k], 3);
        else if (s == 18) bw.put(S.cl_extra[k], 7);
    }
    bw.finish();
    S.hdr_bits = bw.bitpos();
}

// Thread 0, after the prefix sums: total size, stored-block fallback decision, end-of-stream marker position.
DF_HD void df_plan(ChunkState &S, int n, uint32_t token_bits) {
    S.end_bits = S.hdr_bits + token_bits;                       // token_bits includes the end-of-block code
    const uint32_t coded = (S.end_bits + 3 + 7) / 8 + 4;        // + empty stored block: 3 bits, pad, 00 00 FF FF
    S.stored = coded > (uint32_t)n + 5u ? 1u : 0u;
}

// Thread 0, last: the empty stored block that byte-aligns the chunk (BFINAL on the last chunk of a text block).
DF_HD uint32_t df_close(const ChunkState &S, uint32_t *out, int is_last) {
    BitWriter bw;
    bw.start(out, S.end_bits);
    bw.put(is_last ? 1u : 0u, 1);
    bw.put(0u, 2);
    uint32_t pos = bw.bitpos();
    const int pad = (int)((8 - (pos & 7)) & 7);
    if (pad) bw.put(0u, pad);
    bw.put(0x0000u, 16);
    bw.put(0xFFFFu, 16);
    if (bw.nbits > 0) DF_ATOMIC_OR(&bw.w[bw.word], (uint32_t)bw.acc);
    return bw.bitpos() / 8;
}

// Thread 0: Adler-32 pieces of the chunk from the slices' sums.
DF_HD void df_chunk_sums(const ChunkState &S, int n, int nact, uint32_t *sum_a, uint32_t *sum_b) {
    uint64_t a = 0, b = 0;                                     // <= 256 * 2^32 each: no overflow, one reduction at the end
    for (int t = 0; t < nact; t++) {
        const uint64_t after = (uint64_t)(n - t * DF_SLICE - S.slen[t]);
        a += S.sa[t];
        b += S.sb[t] + after * S.sa[t];
    }
    *sum_a = (uint32_t)(a % DF_ADLER);
    *sum_b = (uint32_t)(b % DF_ADLER);
}

// Stored block (BTYPE 00) for a chunk that does not compress: 5 bytes + the text, straight into the slot.
DF_HD void df_stored_header(uint8_t *o, int n, int is_last) {
    o[0] = is_last ? 1 : 0;
    o[1] = (uint8_t)(n & 255); o[2] = (uint8_t)(n >> 8);
    o[3] = (uint8_t)(~n & 255); o[4] = (uint8_t)((~n >> 8) & 255);
}
DF_HD void df_stored_copy(const ChunkState &S, uint8_t *o, int t) {
    const int base = t * DF_SLICE;
    for (int i = 0; i < S.slen[t]; i++) o[5 + base + i] = (uint8_t)df_byte(S, base + i);
}

#ifdef __CUDACC__
struct ChunkGeom { const uint8_t *text; int n, is_last, nact; };
__device__ __forceinline__ ChunkGeom df_geom(const uint8_t *text, const uint64_t *blk_off, const uint32_t *chunk_first,
                                             const uint32_t *chunk_block, uint32_t c) {
    const uint32_t b = chunk_block[c], k = c - chunk_first[b];
    const uint64_t start = blk_off[b] + (uint64_t)k * DF_CHUNK;
    ChunkGeom g;
    g.text = text + start;
    g.n = (int)min((uint64_t)DF_CHUNK, blk_off[b + 1] - start);
    g.is_last = (c + 1 == chunk_first[b + 1]);
    g.nact = (g.n + DF_SLICE - 1) / DF_SLICE;
    return g;
}

// text: the quality text of all blocks (padded by 8 bytes); blk_off[nblocks + 1]: byte offsets of the blocks in it;
// chunk_first[nblocks + 1]: first chunk of every block; chunk_block[nchunks].
__global__ void __launch_bounds__(DF_THREADS) k_deflate_count(const uint8_t *__restrict__ text, const uint64_t *__restrict__ blk_off,
                                                               const uint32_t *__restrict__ chunk_first,
                                                               const uint32_t *__restrict__ chunk_block, HuffState *__restrict__ huff,
                                                               ChunkMeta *__restrict__ meta) {
    __shared__ ChunkState S;
    const int t = threadIdx.x;
    const uint32_t c = blockIdx.x;
    const ChunkGeom g = df_geom(text, blk_off, chunk_first, chunk_block, c);
    if (t == 0) { S.text = g.text; S.nmatch = 0; }
    for (int w = t; w < DF_NLIT + 2; w += DF_THREADS) S.freq[w] = 0;
    __syncthreads();
    df_summary(S, t, g.n);
    __syncthreads();
    { CountSink sink{S}; df_slice_tokens(S, t, g.nact, sink); }
    __syncthreads();
    HuffState &H = huff[c];
    for (int w = t; w < DF_NLIT + 2; w += DF_THREADS) H.freq[w] = S.freq[w];
    if (t == 0) {
        H.nmatch = S.nmatch;
        ChunkMeta m;
        m.nbytes = 0;
        m.n = (uint32_t)g.n;
        df_chunk_sums(S, g.n, g.nact, &m.sum_a, &m.sum_b);
        meta[c] = m;
    }
}

// One warp per chunk, 8 chunks per CTA: the warp stages the chunk's frequencies in shared memory, lane 0 runs the serial
// construction there (in global memory every load that follows a store to the same line goes to L2: 1.8 ms per
// run; in shared memory 0.3 ms), the warp copies the code table and the header back.
__global__ void __launch_bounds__(DF_CODES_CHUNKS * 32) k_deflate_codes(HuffState *__restrict__ huff, uint32_t nchunks) {
    extern __shared__ __align__(16) unsigned char df_codes_smem[];
    HuffState *sh = reinterpret_cast<HuffState *>(df_codes_smem);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t c = blockIdx.x * DF_CODES_CHUNKS + w;
    if (c >= nchunks) return;
    HuffState &G = huff[c], &L = sh[w];
    for (int k = lane; k < DF_NLIT + 2; k += 32) L.freq[k] = G.freq[k];
    if (lane == 0) L.nmatch = G.nmatch;
    __syncwarp();
    if (lane == 0) df_build_block_header(L);
    __syncwarp();
    for (int k = lane; k < DF_NLIT; k += 32) G.lc[k] = L.lc[k];
    for (int k = lane; k < DF_HDR_WORDS; k += 32) G.hdr_w[k] = L.hdr_w[k];
    if (lane == 0) G.hdr_bits = L.hdr_bits;
}

// slots: nchunks x DF_SLOT bytes.  The bits go straight to the chunk's slot, whose used part is zeroed first.
__global__ void __launch_bounds__(DF_THREADS) k_deflate_emit(const uint8_t *__restrict__ text, const uint64_t *__restrict__ blk_off,
                                                              const uint32_t *__restrict__ chunk_first,
                                                              const uint32_t *__restrict__ chunk_block, const HuffState *__restrict__ huff,
                                                              uint8_t *__restrict__ slots, ChunkMeta *__restrict__ meta) {
    __shared__ ChunkState S;
    const int t = threadIdx.x;
    const uint32_t c = blockIdx.x;
    const ChunkGeom g = df_geom(text, blk_off, chunk_first, chunk_block, c);
    const HuffState &H = huff[c];
    uint32_t *slot = (uint32_t *)(slots + (uint64_t)c * DF_SLOT);
    if (t == 0) { S.text = g.text; S.hdr_bits = H.hdr_bits; }
    for (int w = t; w < DF_NLIT; w += DF_THREADS) S.lc[w] = H.lc[w];
    __syncthreads();
    df_summary(S, t, g.n);
    __syncthreads();
    uint32_t mybits;
    {
        BitCountSink sink{S, 0u};
        df_slice_tokens(S, t, g.nact, sink);
        mybits = sink.bits + (t == g.nact - 1 ? (S.lc[256] >> 16) : 0u);
    }
    uint32_t incl = mybits;                                    // exclusive prefix of the per-thread bit counts
    for (int d = 1; d < 32; d <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= d) incl += v; }
    if ((t & 31) == 31) S.bits[t >> 5] = incl;
    __syncthreads();
    uint32_t warp_base = 0, total = 0;
    for (int w = 0; w < DF_THREADS / 32; w++) { uint32_t v = S.bits[w]; if (w < (t >> 5)) warp_base += v; total += v; }
    const uint32_t excl = warp_base + incl - mybits;
    if (t == 0) df_plan(S, g.n, total);
    __syncthreads();
    if (!S.stored) {
        const int nw = (int)((S.end_bits + 3 + 7 + 32) / 32 + 1);          // tokens + closing stored block
        for (int w = t; w < nw; w += DF_THREADS) slot[w] = 0;
        __syncthreads();
        for (int w = t; w < (int)((S.hdr_bits + 31) / 32); w += DF_THREADS) atomicOr(&slot[w], H.hdr_w[w]);
        EmitSink sink{S};
        sink.bw.start(slot, S.hdr_bits + excl);
        df_slice_tokens(S, t, g.nact, sink);
        if (t == g.nact - 1) sink.bw.put(S.lc[256] & 0xffffu, (int)(S.lc[256] >> 16));
        sink.bw.finish();
    } else {
        if (t == 0) df_stored_header((uint8_t *)slot, g.n, g.is_last);
        df_stored_copy(S, (uint8_t *)slot, t);
    }
    __syncthreads();
    if (t == 0) meta[c].nbytes = S.stored ? (uint32_t)g.n + 5u : df_close(S, slot, g.is_last);
}

// Layout of the zlib streams: per block the offsets of its chunks, its size (2 + chunks + 4) and its Adler-32; then
// the exclusive prefix of the block sizes.  One CTA; blocks are few (<= a few ten thousand).
__global__ void __launch_bounds__(1024) k_deflate_layout(const ChunkMeta *__restrict__ meta, const uint32_t *__restrict__ chunk_first,
                                                         uint32_t nblocks, uint32_t *__restrict__ chunk_zoff,
                                                         uint64_t *__restrict__ blk_zsize, uint64_t *__restrict__ blk_zoff,
                                                         uint32_t *__restrict__ blk_adler) {
    for (uint32_t b = threadIdx.x; b < nblocks; b += blockDim.x) {
        uint64_t A = 1, B = 0, z = 2;
        for (uint32_t c = chunk_first[b]; c < chunk_first[b + 1]; c++) {
            const ChunkMeta m = meta[c];
            chunk_zoff[c] = (uint32_t)z;
            z += m.nbytes;
            B = (B + (uint64_t)m.n % DF_ADLER * A + m.sum_b) % DF_ADLER;
            A = (A + m.sum_a) % DF_ADLER;
        }
        blk_zsize[b] = z + 4;
        blk_adler[b] = (uint32_t)((B << 16) | A);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t off = 0;
        for (uint32_t b = 0; b < nblocks; b++) { blk_zoff[b] = off; off += blk_zsize[b]; }
        blk_zoff[nblocks] = off;
    }
}

// One CTA per chunk: slot -> its place in the block's stream; the first / last chunk of a block also write the zlib
// header (CM = 8, 32 KiB window, no preset dictionary, check bits) and the big-endian Adler-32.
__global__ void __launch_bounds__(256) k_deflate_gather(const uint8_t *__restrict__ slots, const ChunkMeta *__restrict__ meta,
                                                        const uint32_t *__restrict__ chunk_first, const uint32_t *__restrict__ chunk_block,
                                                        const uint32_t *__restrict__ chunk_zoff, const uint64_t *__restrict__ blk_zoff,
                                                        const uint32_t *__restrict__ blk_adler, uint8_t *__restrict__ out) {
    const uint32_t c = blockIdx.x, b = chunk_block[c];
    const uint8_t *src = slots + (uint64_t)c * DF_SLOT;
    uint8_t *blk = out + blk_zoff[b], *dst = blk + chunk_zoff[c];
    const uint32_t nb = meta[c].nbytes;
    for (uint32_t i = threadIdx.x; i < nb; i += blockDim.x) dst[i] = src[i];
    if (threadIdx.x == 0) {
        if (c == chunk_first[b]) { blk[0] = 0x78; blk[1] = 0x01; }
        if (c + 1 == chunk_first[b + 1]) {
            const uint32_t a = blk_adler[b];
            uint8_t *e = dst + nb;
            e[0] = (uint8_t)(a >> 24); e[1] = (uint8_t)(a >> 16); e[2] = (uint8_t)(a >> 8); e[3] = (uint8_t)a;
        }
    }
}
#endif  // __CUDACC__

}  // namespace lbdeflate
