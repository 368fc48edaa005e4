/* TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.  See leon_oracle.h for scope, citations shorthand and
 * parity status (PINNED against the compiled reference, tests/golden/).
 *
 * Plain-C restatement of the reference `leon -c` path (single thread, == `-nb-cores 1`).
 */
#include "leon_oracle.h"
#include "leon_tables.h"
#include <stdlib.h>
#include <string.h>
#include <limits.h>

/* ------------------------------------------------------------------------------------------------
 * k-mer integer primitives
 * ---------------------------------------------------------------------------------------------- */

/* reverse-complement of all 32 nucleotides of a 64-bit word; codes A0 C1 T2 G3 so complement = ^2.
 * M/LargeInt1.pri:137-154 */
static uint64_t rc64(uint64_t x)
{
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
    x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
    x = (x >> 32) | (x << 32);
    return x ^ 0xAAAAAAAAAAAAAAAAULL;
}

/* M/LargeInt1.pri:137-154 (k<=31, one word) and M/LargeInt.hpp:722-734 (two words: byte-reverse the whole
 * 16-byte value through revcomp_4NT, then shift right by 2*(64-k)); both are "reverse the k fields, ^2". */
lo_kmer lo_revcomp(lo_kmer x, int k)
{
    uint64_t lo = (uint64_t)x, hi = (uint64_t)(x >> 64);
    lo_kmer r = ((lo_kmer)rc64(lo) << 64) | (lo_kmer)rc64(hi);
    return r >> (2 * (64 - k));
}

/* M/LargeInt1.pri:157-170 == M/NativeInt64.hpp:175-188 */
uint64_t lo_hash64(uint64_t key, uint64_t seed)
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

/* k<=31 is compiled as LargeInt<1> in the stock reference build (KSIZE_LIST starts at 32) and 32<=k<=63 as
 * LargeInt<2> in the "64" build: hash1 XORs hash64 of every 64-bit word. M/LargeInt.hpp:738-748 */
uint64_t lo_hash1(lo_kmer x, int k, uint64_t seed)
{
    if (k <= 31) return lo_hash64((uint64_t)x, seed);
    return lo_hash64((uint64_t)x, seed) ^ lo_hash64((uint64_t)(x >> 64), seed);
}

/* One word: three table XORs (M/LargeInt1.pri:190-211).  Two words: two table XORs on the LOW word only
 * (M/LargeInt.hpp:792-801 -> M/NativeInt64.hpp:211-221). */
uint64_t lo_simplehash16(lo_kmer x, int k, int shift)
{
    uint64_t key = (uint64_t)x;
    uint64_t input = key >> shift;
    uint64_t res = lo_random_values[input & 255];
    input >>= 8;
    res ^= lo_random_values[input & 255];
    if (k <= 31) res ^= lo_random_values[key & 255];
    return res;
}

/* COL/Bloom.hpp:83-95: seed_tab[i] = rbase[i]; then in place seed_tab[i] = seed_tab[i]*seed_tab[(i+3)%10] + user_seed,
 * i ascending (so late entries see already-updated early ones). Only seed 0 is used by BloomNeighborCoherent. */
uint64_t lo_bloom_seed(int idx)
{
    static const uint64_t rbase[10] = {
        0xAAAAAAAA55555555ULL, 0x33333333CCCCCCCCULL, 0x6666666699999999ULL, 0xB5B5B5B54B4B4B4BULL,
        0xAA55AA5555335533ULL, 0x33CC33CCCC66CC66ULL, 0x6699669999B599B5ULL, 0xB54BB54B4BAA4BAAULL,
        0xAA33AA3355CC55CCULL, 0x33663366CC99CC99ULL };
    uint64_t s[10];
    for (int i = 0; i < 10; i++) s[i] = rbase[i];
    for (int i = 0; i < 10; i++) s[i] = s[i] * s[(i + 3) % 10] + 0;
    return s[idx];
}

/* K/Model.hpp:636-658 (polynom) with MI/api/Data.hpp:176 ConvertASCII: code = (c>>1)&3 : A0 C1 T2 G3 */
lo_kmer lo_encode_kmer(const char *s, int k)
{
    lo_kmer v = 0;
    for (int i = 0; i < k; i++) v = (v << 2) | (lo_kmer)((s[i] >> 1) & 3);
    return v;
}

static inline lo_kmer kmask(int k) { return (((lo_kmer)1) << (2 * k)) - 1; }

/* ------------------------------------------------------------------------------------------------
 * BloomNeighborCoherent — COL/Bloom.hpp:184-199 (BloomContainer), :437-442 (BloomCacheCoherent), :523-552
 * ---------------------------------------------------------------------------------------------- */
static const uint8_t cano2[16] = { 0, 1, 2, 3, 4, 5, 3, 7, 8, 9, 0, 4, 9, 13, 1, 5 };
#define LO_BLOCK_MASK 4095u

int lo_bloom_init(lo_bloom *b, uint64_t tai_bloom, int k, int nhash)
{
    uint64_t tai = tai_bloom + 2 * 4096;       /* Bloom.hpp:438 */
    b->nbytes = 1 + tai / 8;                   /* Bloom.hpp:187 */
    b->bits = (uint8_t *)calloc(b->nbytes, 1);
    if (!b->bits) return -1;
    if (tai && !(tai & (tai - 1))) tai--;      /* Bloom.hpp:193-198 */
    b->tai = tai;
    b->reduced_tai = tai - 2 * 4096;           /* Bloom.hpp:441 */
    b->k = k;
    b->nhash = nhash;
    return 0;
}

void lo_bloom_free(lo_bloom *b) { free(b->bits); b->bits = NULL; }

static inline lo_kmer bloom_hashpart(const lo_bloom *b, lo_kmer item)
{
    lo_kmer hp = (item >> 2) & kmask(b->k - 2);
    lo_kmer rev = lo_revcomp(hp, b->k - 2);
    return rev < hp ? rev : hp;
}

static inline unsigned bloom_prefval(const lo_bloom *b, lo_kmer item)
{
    unsigned suffix = (unsigned)(item & 3);
    unsigned prefix = (unsigned)((item >> (2 * (b->k - 1))) & 3);   /* (item & _prefmask) >> ((k-2)*2) == first_nt<<2 */
    return cano2[((prefix << 2) + suffix) & 15];
}

void lo_bloom_insert(lo_bloom *b, lo_kmer item)
{
    lo_kmer hp = bloom_hashpart(b, item);
    uint64_t racine = lo_hash1(hp, b->k, lo_bloom_seed(0)) % b->reduced_tai;
    uint64_t h0 = racine + bloom_prefval(b, item);
    b->bits[h0 >> 3] |= (uint8_t)(1u << (h0 & 7));
    for (int i = 1; i < b->nhash; i++) {
        uint64_t h1 = h0 + (lo_simplehash16(hp, b->k, i) & LO_BLOCK_MASK);
        b->bits[h1 >> 3] |= (uint8_t)(1u << (h1 & 7));
    }
}

int lo_bloom_contains(const lo_bloom *b, lo_kmer item)
{
    lo_kmer hp = bloom_hashpart(b, item);
    uint64_t racine = lo_hash1(hp, b->k, lo_bloom_seed(0)) % b->reduced_tai;
    uint64_t h0 = racine + bloom_prefval(b, item);
    if (!(b->bits[h0 >> 3] & (1u << (h0 & 7)))) return 0;
    for (int i = 1; i < b->nhash; i++) {
        uint64_t h1 = h0 + (lo_simplehash16(hp, b->k, i) & LO_BLOCK_MASK);
        if (!(b->bits[h1 >> 3] & (1u << (h1 & 7)))) return 0;
    }
    return 1;
}

/* COL/Bloom.hpp:645-798.  `item` is used as given (forward strand of the walk); returns a 4-bit mask, bit nt set
 * iff the neighbour obtained by appending (right) / prepending (left) nucleotide nt tests positive. */
int lo_bloom_contains4(const lo_bloom *b, lo_kmer item, int right)
{
    int k = b->k;
    lo_kmer elem = right ? ((item << 2) & kmask(k)) : (item >> 2);
    lo_kmer hp = bloom_hashpart(b, elem);
    uint64_t racine = lo_hash1(hp, k, lo_bloom_seed(0)) % b->reduced_tai;
    uint64_t th[8];
    for (int i = 1; i < b->nhash; i++) th[i] = lo_simplehash16(hp, k, i) & LO_BLOCK_MASK;
    int res = 0;
    for (int nt = 0; nt < 4; nt++) {
        lo_kmer tmp = right ? (elem + (lo_kmer)nt) : (elem + ((lo_kmer)nt << (2 * (k - 1))));
        uint64_t h0 = racine + (bloom_prefval(b, tmp) & LO_BLOCK_MASK);
        int ok = (b->bits[h0 >> 3] >> (h0 & 7)) & 1;
        for (int i = 1; ok && i < b->nhash; i++) {
            uint64_t h1 = h0 + th[i];
            ok = (b->bits[h1 >> 3] >> (h1 & 7)) & 1;
        }
        if (ok) res |= 1 << nt;
    }
    return res;
}

/* ------------------------------------------------------------------------------------------------
 * Order0Model + RangeEncoder — C/RangeCoder.cpp:29-145, constants C/RangeCoder.hpp:37-39
 * ---------------------------------------------------------------------------------------------- */
#define RC_TOP    ((uint64_t)1 << 56)
#define RC_BOTTOM ((uint64_t)1 << 48)

typedef struct { uint64_t *r; int n; /* n = charCount+1 entries */ } o0model;

static void o0_init(o0model *m, int char_count)
{
    m->n = char_count + 1;
    m->r = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)m->n);
    for (int i = 0; i < m->n; i++) m->r[i] = (uint64_t)i;
}
static void o0_clear(o0model *m) { for (int i = 0; i < m->n; i++) m->r[i] = (uint64_t)i; }
static void o0_free(o0model *m) { free(m->r); m->r = NULL; }
static void o0_update(o0model *m, uint8_t c)
{
    for (int i = c + 1; i < m->n; i++) m->r[i] += 1;
    if (m->r[m->n - 1] >= RC_BOTTOM) {            /* rescale, RangeCoder.cpp:57-64 */
        for (int i = 1; i < m->n; i++) {
            m->r[i] /= 2;
            if (m->r[i] <= m->r[i - 1]) m->r[i] = m->r[i - 1] + 1;
        }
    }
}

typedef struct { uint64_t low, range; uint8_t *buf; uint64_t n, cap; } rcenc;

static void rc_init(rcenc *e) { e->low = 0; e->range = (uint64_t)-1; e->buf = NULL; e->n = e->cap = 0; }
static void rc_put(rcenc *e, uint8_t b)
{
    if (e->n == e->cap) { e->cap = e->cap ? e->cap * 2 : 4096; e->buf = (uint8_t *)realloc(e->buf, e->cap); }
    e->buf[e->n++] = b;
}
static void rc_encode(rcenc *e, o0model *m, uint8_t c)
{
    e->range /= m->r[m->n - 1];
    e->low += m->r[c] * e->range;
    e->range *= m->r[c + 1] - m->r[c];
    while ((e->low ^ (e->low + e->range)) < RC_TOP || e->range < RC_BOTTOM) {
        if (e->range < RC_BOTTOM && (e->low ^ (e->low + e->range)) >= RC_TOP)
            e->range = (0 - e->low) & (RC_BOTTOM - 1);
        rc_put(e, (uint8_t)(e->low >> 56));
        e->range <<= 8;
        e->low <<= 8;
    }
    o0_update(m, c);
}
static void rc_flush(rcenc *e)
{
    for (int i = 0; i < 8; i++) { rc_put(e, (uint8_t)(e->low >> 56)); e->low <<= 8; }
    e->low = 0; e->range = (uint64_t)-1;
}
static void rc_clear(rcenc *e) { e->low = 0; e->range = (uint64_t)-1; e->n = 0; }

/* C/CompressionUtils.hpp:75-103: 7 bits per byte LSB first, bit 7 = "more", byte j coded with model j */
#define NB_MODELS_PER_NUMERIC 20
static void rc_numeric(rcenc *e, o0model *models, uint64_t value)
{
    int i = 0;
    do {
        uint8_t byte = (uint8_t)(value & 127);
        value >>= 7;
        if (value != 0) byte |= 128;
        rc_encode(e, &models[i], byte);
        i++;
    } while (value != 0);
}

/* known-answer helper: symbols 0,1,2,3,0,0,1,4,2,2 with one Order0Model(5), then varints 100, 150,
 * 300000 over a fresh 20 x Order0Model(256) set, then flush -> 17 bytes 16dcd51bc15c990793871000eb56600000 */
uint64_t lo_rangecoder_kat(uint8_t *out, uint64_t cap)
{
    static const uint8_t syms[10] = { 0, 1, 2, 3, 0, 0, 1, 4, 2, 2 };
    rcenc e; rc_init(&e);
    o0model m5; o0_init(&m5, 5);
    o0model num[NB_MODELS_PER_NUMERIC];
    for (int i = 0; i < NB_MODELS_PER_NUMERIC; i++) o0_init(&num[i], 256);
    for (int i = 0; i < 10; i++) rc_encode(&e, &m5, syms[i]);
    rc_numeric(&e, num, 100); rc_numeric(&e, num, 150); rc_numeric(&e, num, 300000);
    rc_flush(&e);
    uint64_t n = e.n < cap ? e.n : cap;
    memcpy(out, e.buf, n);
    free(e.buf); o0_free(&m5);
    for (int i = 0; i < NB_MODELS_PER_NUMERIC; i++) o0_free(&num[i]);
    return n;
}

/* ------------------------------------------------------------------------------------------------
 * auto cutoff — MI/impl/Histogram.cpp:62-195 with _length = 10000 (-histo-max)
 * ---------------------------------------------------------------------------------------------- */
void lo_compute_threshold(const uint64_t *h, int min_auto_threshold, int32_t *cutoff_out, uint64_t *nbsolids_out)
{
    const size_t length = 10000;
    static uint64_t sm[10001];
    memset(sm, 0, sizeof sm);
    int cutoff = 0;                 /* Histogram ctor: _cutoff(0), _nbsolids(0) */
    uint64_t nbsolids = 0;
    uint64_t sum_allk = 0;

    sm[1] = (uint64_t)(0.6 * (double)h[1] + 0.4 * (double)h[2]);
    sum_allk += h[1] * 1;

    int index_first_increase = -1, index_maxval_after_first_increase = -1;
    uint64_t max_val = 0;
    for (size_t i = 2; i < length; i++) {
        sum_allk += h[i] * i;
        sm[i] = (uint64_t)(0.2 * (double)h[i - 1] + 0.6 * (double)h[i] + 0.2 * (double)h[i + 1]);
        if (index_first_increase == -1 && sm[i - 1] < sm[i]) index_first_increase = (int)i - 1;
        if (index_first_increase > 0 && sm[i] > max_val) { max_val = sm[i]; index_maxval_after_first_increase = (int)i; }
    }
    sum_allk += h[length] * length;

    if (index_first_increase == -1) { *cutoff_out = min_auto_threshold; *nbsolids_out = 0; return; }

    uint64_t min_val = 10000000000ULL;
    int index_minval = -1;
    for (int i = index_first_increase; i <= index_maxval_after_first_increase; i++)
        if (sm[i] < min_val) { min_val = sm[i]; index_minval = i; }
    if (index_minval != -1) cutoff = index_minval;

    uint64_t sum_elim = 0;
    int max_cutoff = 0;
    for (size_t i = 0; i < length + 1; i++) {
        sum_elim += h[i] * i;
        double ratio = (double)sum_elim / sum_allk;
        if (ratio >= 0.25) { max_cutoff = (int)i + 1; break; }
    }
    if (cutoff > max_cutoff) cutoff = max_cutoff;
    if (cutoff < min_auto_threshold) cutoff = min_auto_threshold;

    for (size_t i = (size_t)cutoff; i < length + 1; i++) nbsolids += h[i];
    *cutoff_out = cutoff;
    *nbsolids_out = nbsolids;
}

/* ------------------------------------------------------------------------------------------------
 * anchor dictionary: exact map canonical k-mer -> address (COL/Hash16.hpp:163-195,262-285 semantics)
 * ---------------------------------------------------------------------------------------------- */
typedef struct { lo_kmer *keys; uint32_t *vals; uint8_t *used; uint64_t cap, n; } kmap;

static void kmap_init(kmap *m, uint64_t cap) {
    m->cap = 1024; while (m->cap < cap) m->cap <<= 1;
    m->keys = (lo_kmer *)malloc(sizeof(lo_kmer) * m->cap);
    m->vals = (uint32_t *)malloc(sizeof(uint32_t) * m->cap);
    m->used = (uint8_t *)calloc(m->cap, 1);
    m->n = 0;
}
static void kmap_free(kmap *m) { free(m->keys); free(m->vals); free(m->used); }
static uint64_t kmap_slot(const kmap *m, lo_kmer key) {
    uint64_t h = lo_hash64((uint64_t)key ^ (uint64_t)(key >> 64), 0x9E3779B97F4A7C15ULL) & (m->cap - 1);
    while (m->used[h] && m->keys[h] != key) h = (h + 1) & (m->cap - 1);
    return h;
}
static void kmap_put(kmap *m, lo_kmer key, uint32_t val);
static void kmap_grow(kmap *m) {
    kmap o = *m;
    kmap_init(m, o.cap * 2);
    for (uint64_t i = 0; i < o.cap; i++) if (o.used[i]) kmap_put(m, o.keys[i], o.vals[i]);
    kmap_free(&o);
}
static void kmap_put(kmap *m, lo_kmer key, uint32_t val) {
    if ((m->n + 1) * 2 > m->cap) kmap_grow(m);
    uint64_t h = kmap_slot(m, key);
    if (!m->used[h]) { m->used[h] = 1; m->keys[h] = key; m->n++; }
    m->vals[h] = val;
}
static int kmap_get(const kmap *m, lo_kmer key, uint32_t *val) {
    uint64_t h = kmap_slot(m, key);
    if (!m->used[h]) return 0;
    *val = m->vals[h];
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * DSK semantics as used by Leon: multiset of canonical k-mers over all windows free of
 * invalid characters ((c>>3)&1, MI/api/Data.hpp:176), K/Model.hpp:725-764,876-883.
 * ---------------------------------------------------------------------------------------------- */
static int cmp_kmer(const void *a, const void *b)
{
    lo_kmer x = *(const lo_kmer *)a, y = *(const lo_kmer *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

static inline int nt_code(uint8_t c) { return (c >> 1) & 3; }
static inline int nt_invalid(uint8_t c) { return (c >> 3) & 1; }

/* C/Leon.cpp:44-56 nt2binTab */
static inline int leon_nt2bin(uint8_t c)
{
    switch (c) { case 'A': return 0; case 'C': return 1; case 'T': return 2; case 'G': return 3; case 'N': return 4; default: return 0; }
}

typedef struct {
    o0model readType, noAnchorRead, bifurcation, bifurcationBinary, readAnchorRevcomp;
    o0model anchorAddress[NB_MODELS_PER_NUMERIC], anchorPos[NB_MODELS_PER_NUMERIC],
            noAnchorReadSize[NB_MODELS_PER_NUMERIC], readSize[NB_MODELS_PER_NUMERIC], Npos[NB_MODELS_PER_NUMERIC],
            leftErrorPos[NB_MODELS_PER_NUMERIC], numeric[NB_MODELS_PER_NUMERIC], leftError[NB_MODELS_PER_NUMERIC];
} dna_models;

/* C/DnaCoder.cpp:43-75 */
static void models_init(dna_models *m)
{
    o0_init(&m->readType, 2); o0_init(&m->noAnchorRead, 5); o0_init(&m->bifurcation, 5);
    o0_init(&m->bifurcationBinary, 2); o0_init(&m->readAnchorRevcomp, 2);
    for (int i = 0; i < NB_MODELS_PER_NUMERIC; i++) {
        o0_init(&m->anchorAddress[i], 256); o0_init(&m->anchorPos[i], 256); o0_init(&m->noAnchorReadSize[i], 256);
        o0_init(&m->readSize[i], 256); o0_init(&m->Npos[i], 256); o0_init(&m->leftErrorPos[i], 256);
        o0_init(&m->numeric[i], 256); o0_init(&m->leftError[i], 256);
    }
}
/* C/DnaCoder.cpp:78-112 startBlock */
static void models_clear(dna_models *m)
{
    o0_clear(&m->readType); o0_clear(&m->noAnchorRead); o0_clear(&m->bifurcation);
    o0_clear(&m->bifurcationBinary); o0_clear(&m->readAnchorRevcomp);
    for (int i = 0; i < NB_MODELS_PER_NUMERIC; i++) {
        o0_clear(&m->anchorAddress[i]); o0_clear(&m->anchorPos[i]); o0_clear(&m->noAnchorReadSize[i]);
        o0_clear(&m->readSize[i]); o0_clear(&m->Npos[i]); o0_clear(&m->leftErrorPos[i]);
        o0_clear(&m->numeric[i]); o0_clear(&m->leftError[i]);
    }
}
static void models_free(dna_models *m)
{
    o0_free(&m->readType); o0_free(&m->noAnchorRead); o0_free(&m->bifurcation);
    o0_free(&m->bifurcationBinary); o0_free(&m->readAnchorRevcomp);
    for (int i = 0; i < NB_MODELS_PER_NUMERIC; i++) {
        o0_free(&m->anchorAddress[i]); o0_free(&m->anchorPos[i]); o0_free(&m->noAnchorReadSize[i]);
        o0_free(&m->readSize[i]); o0_free(&m->Npos[i]); o0_free(&m->leftErrorPos[i]);
        o0_free(&m->numeric[i]); o0_free(&m->leftError[i]);
    }
}

typedef struct { uint8_t *p; uint64_t n, cap; } bytebuf;
static void bb_append(bytebuf *b, const void *src, uint64_t n)
{
    if (b->n + n > b->cap) { while (b->n + n > b->cap) b->cap = b->cap ? b->cap * 2 : 65536; b->p = (uint8_t *)realloc(b->p, b->cap); }
    if (n) memcpy(b->p + b->n, src, n);
    b->n += n;
}

/* ------------------------------------------------------------------------------------------------
 * per-read encoder state (C/DnaCoder.cpp:350-405 execute and callees)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int k;
    const lo_bloom *bloom;
    dna_models m;
    rcenc rc;
    uint8_t *seq;           /* working copy of the read (N -> A in place, restored for no-anchor reads) */
    lo_kmer *kmers;         /* forward-strand k-mers of the working copy */
    int *Npos; int nN;
    int *errPos; int nErr;
    uint8_t *bif, *bifType; int nBif;   /* symbols in walk order; type 0 -> bifurcation(5), 1 -> binary(2) */
    int *nb_solids;
    uint64_t cap;           /* capacity of the per-read arrays */
    lo_result *res;
} encoder;

static void enc_reserve(encoder *e, uint64_t L)
{
    if (L + 2 <= e->cap) return;
    e->cap = L + 1024;
    e->seq = (uint8_t *)realloc(e->seq, e->cap);
    e->kmers = (lo_kmer *)realloc(e->kmers, sizeof(lo_kmer) * e->cap);
    e->Npos = (int *)realloc(e->Npos, sizeof(int) * e->cap);
    e->errPos = (int *)realloc(e->errPos, sizeof(int) * e->cap);
    e->bif = (uint8_t *)realloc(e->bif, e->cap);
    e->bifType = (uint8_t *)realloc(e->bifType, e->cap);
    e->nb_solids = (int *)realloc(e->nb_solids, sizeof(int) * e->cap);
}

static inline lo_kmer canon(lo_kmer x, int k) { lo_kmer r = lo_revcomp(x, k); return r < x ? r : x; }

/* C/DnaCoder.cpp:118-136 codeSeedBin: right = shift in at the low end; left = revcomp/shift/revcomp == shift in at the top */
static inline lo_kmer seed_step(lo_kmer kmer, int nt, int right, int k)
{
    if (right) return ((kmer << 2) | (lo_kmer)nt) & kmask(k);
    return (kmer >> 2) | ((lo_kmer)nt << (2 * (k - 1)));
}

static int cmp_int(const void *a, const void *b) { int x = *(const int *)a, y = *(const int *)b; return x < y ? -1 : x > y; }

/* C/DnaCoder.cpp:759-969 buildBifurcationList */
static lo_kmer walk_step(encoder *e, int pos, lo_kmer kmer, int right)
{
    int k = e->k;
    uint8_t nextNt = e->seq[pos];
    int nextNtBin = leon_nt2bin(nextNt);
    for (int i = 0; i < e->nN; i++)
        if (e->Npos[i] == pos) return seed_step(kmer, nextNtBin, right, k);     /* N was rewritten to 'A' */

    int res4 = lo_bloom_contains4(e->bloom, kmer, right);
    int count = 0, firstNt = -1, solid = 0;
    for (int nt = 0; nt < 4; nt++)
        if (res4 & (1 << nt)) { count++; if (firstNt < 0) firstNt = nt; if (nt == nextNtBin) solid = 1; }
    lo_kmer uniq = firstNt >= 0 ? seed_step(kmer, firstNt, right, k) : 0;
    e->res->mc_total++;

    if (solid) {
        if (count == 1) { e->res->mc_uniq_solid++; return uniq; }
        if (count == 2) {
            int nt1 = firstNt;                                   /* lowest set nucleotide */
            e->bif[e->nBif] = (nt1 == nextNtBin) ? 0 : 1; e->bifType[e->nBif++] = 1;
            e->res->mc_multiple_solid++;
            return seed_step(kmer, nextNtBin, right, k);
        }
        e->bif[e->nBif] = (uint8_t)nextNtBin; e->bifType[e->nBif++] = 0;
        return seed_step(kmer, nextNtBin, right, k);
    }
    if (count == 0) {
        e->res->mc_no_alternative++;
    } else if (count == 1) {
        e->res->mc_uniq_nosolid++;
        e->bif[e->nBif] = (uint8_t)nextNtBin; e->bifType[e->nBif++] = 0;
        e->errPos[e->nErr++] = pos;
        return uniq;
    } else if (count == 2) {
        e->res->mc_uniq_nosolid++;
        e->bif[e->nBif] = (uint8_t)nextNtBin; e->bifType[e->nBif++] = 0;
        e->errPos[e->nErr++] = pos;
        return seed_step(kmer, firstNt, right, k);               /* first path of the bifurcation */
    } else {
        e->res->mc_uniq_nosolid++;
    }
    e->bif[e->nBif] = (uint8_t)nextNtBin; e->bifType[e->nBif++] = 0;
    return seed_step(kmer, nextNtBin, right, k);
}

/* C/DnaCoder.cpp:599-757 */
static void encode_anchor_read(encoder *e, uint64_t L, int anchorPos, uint32_t anchorAddress)
{
    int k = e->k;
    rc_encode(&e->rc, &e->m.readType, 0);
    rc_numeric(&e->rc, e->m.readSize, L);
    rc_numeric(&e->rc, e->m.anchorPos, (uint64_t)anchorPos);
    rc_numeric(&e->rc, e->m.anchorAddress, anchorAddress);
    lo_kmer anchor = e->kmers[anchorPos];
    rc_encode(&e->rc, &e->m.readAnchorRevcomp, anchor == canon(anchor, k) ? 0 : 1);

    e->nBif = 0; e->nErr = 0;
    lo_kmer kmer = anchor;
    for (int i = anchorPos - 1; i >= 0; i--) kmer = walk_step(e, i, kmer, 0);
    kmer = anchor;
    for (uint64_t i = (uint64_t)anchorPos + k; i < L; i++) kmer = walk_step(e, (int)i, kmer, 1);

    rc_numeric(&e->rc, e->m.numeric, (uint64_t)e->nN);
    int prev = 0;
    for (int i = 0; i < e->nN; i++) { rc_numeric(&e->rc, e->m.Npos, (uint64_t)(e->Npos[i] - prev)); prev = e->Npos[i]; }

    rc_numeric(&e->rc, e->m.leftError, (uint64_t)e->nErr);
    qsort(e->errPos, (size_t)e->nErr, sizeof(int), cmp_int);
    prev = 0;
    for (int i = 0; i < e->nErr; i++) { rc_numeric(&e->rc, e->m.leftErrorPos, (uint64_t)(e->errPos[i] - prev)); prev = e->errPos[i]; }

    for (int i = 0; i < e->nBif; i++)
        rc_encode(&e->rc, e->bifType[i] == 0 ? &e->m.bifurcation : &e->m.bifurcationBinary, e->bif[i]);
}

/* C/DnaCoder.cpp:1158-1211 */
static void encode_noanchor_read(encoder *e, uint64_t L)
{
    for (int i = 0; i < e->nN; i++) e->seq[e->Npos[i]] = 'N';
    rc_encode(&e->rc, &e->m.readType, 1);
    e->res->reads_without_anchor++;
    rc_numeric(&e->rc, e->m.noAnchorReadSize, L);
    for (uint64_t i = 0; i < L; i++) rc_encode(&e->rc, &e->m.noAnchorRead, (uint8_t)leon_nt2bin(e->seq[i]));
}

/* C/DnaCoder.cpp:428-486 smoothQuals + apply_smoothing_at_pos; appends "qual\n" to the block text */
static void smooth_quals(encoder *e, const uint8_t *q, uint64_t L, int lossless, bytebuf *out)
{
    uint64_t base = out->n;
    bb_append(out, q, L);
    static const uint8_t nl = '\n';
    bb_append(out, &nl, 1);
    if (lossless || L < (uint64_t)e->k) return;
    uint8_t *qs = out->p + base;
    for (uint64_t i = 0; i < L; i++) {
        if (e->nb_solids[i] >= 2 || (int)qs[i] > (int)'@') {
            int phred = (int)qs[i] - 33;
            if (phred == 0 || phred == 2) continue;
            int diff = '@' - (int)qs[i];
            if (diff > 10 && !(e->nb_solids[i] > diff - 5)) continue;
            qs[i] = '@';
        }
    }
}

/* C/DnaCoder.cpp:319-348 writeBlock + C/Leon.cpp:653-710: flush only when some byte was already emitted */
static void write_block(encoder *e, uint64_t nreads_in_block, bytebuf *dna, uint64_t **bs, uint64_t *nblocks, uint64_t *bscap)
{
    if (nreads_in_block == 0) return;
    if (e->rc.n > 0) rc_flush(&e->rc);
    if (*nblocks == *bscap) { *bscap = *bscap ? *bscap * 2 : 64; *bs = (uint64_t *)realloc(*bs, sizeof(uint64_t) * 2 * *bscap); }
    (*bs)[2 * *nblocks] = e->rc.n;
    (*bs)[2 * *nblocks + 1] = nreads_in_block;
    (*nblocks)++;
    bb_append(dna, e->rc.buf, e->rc.n);
    rc_clear(&e->rc);
}

int lo_compress(const uint8_t *bases, const uint64_t *offsets, uint64_t nreads, const uint8_t *quals,
                const lo_params *p, lo_result *r)
{
    const int k = p->kmer_size;
    if (k < 2 || k > 63 || p->reads_per_block <= 0) return -1;
    memset(r, 0, sizeof *r);

    /* ---------- phase 1: DSK semantics ---------- */
    uint64_t nk_max = 0;
    for (uint64_t i = 0; i < nreads; i++) { uint64_t L = offsets[i + 1] - offsets[i]; if (L >= (uint64_t)k) nk_max += L - k + 1; }
    lo_kmer *all = (lo_kmer *)malloc(sizeof(lo_kmer) * (nk_max ? nk_max : 1));
    uint64_t nk = 0;
    for (uint64_t i = 0; i < nreads; i++) {
        const uint8_t *s = bases + offsets[i];
        uint64_t L = offsets[i + 1] - offsets[i];
        if (L < (uint64_t)k) continue;
        lo_kmer fwd = 0; int valid_run = 0;
        for (uint64_t j = 0; j < L; j++) {
            fwd = ((fwd << 2) | (lo_kmer)nt_code(s[j])) & kmask(k);
            valid_run = nt_invalid(s[j]) ? 0 : valid_run + 1;
            if (valid_run >= k) all[nk++] = canon(fwd, k);
        }
    }
    qsort(all, nk, sizeof(lo_kmer), cmp_kmer);
    /* run-length: histogram over ALL distinct k-mers (index truncated to u16 then clamped, MI/impl/Histogram.hpp:87),
     * solid list = count >= 2 (K/SortingCountAlgorithm.cpp:210, CountProcessorSolidity) */
    uint64_t nsolid = 0;
    lo_kmer *solid = (lo_kmer *)malloc(sizeof(lo_kmer) * (nk ? nk : 1));
    uint32_t *solid_cnt = (uint32_t *)malloc(sizeof(uint32_t) * (nk ? nk : 1));
    for (uint64_t i = 0; i < nk;) {
        uint64_t j = i + 1;
        while (j < nk && all[j] == all[i]) j++;
        uint64_t cnt = j - i;
        uint16_t idx = (uint16_t)cnt;
        r->histogram[idx >= 10000 ? 10000 : idx]++;
        if (cnt >= 2) { solid[nsolid] = all[i]; solid_cnt[nsolid] = (uint32_t)(cnt > 0x7fffffff ? 0x7fffffff : cnt); nsolid++; }
        i = j;
    }
    free(all);
    r->nb_solid_total = nsolid;

    /* ---------- phase 2: cutoff + Bloom (C/Leon.cpp:226-338, min threshold 4 forced at C/Leon.cpp:456) ---------- */
    int32_t auto_cutoff = 0; uint64_t nbs = 0;
    lo_compute_threshold(r->histogram, 4, &r->cutoff, &r->nb_solids_cutoff);
    if (!p->abundance_set) { auto_cutoff = r->cutoff; nbs = r->nb_solids_cutoff; }
    else { auto_cutoff = 0; nbs = nsolid; }
    uint64_t est = (uint64_t)((double)nbs * 12);
    if (est == 0) est = 1000;
    lo_bloom bloom;
    if (lo_bloom_init(&bloom, est, k, 7)) return -2;
    for (uint64_t i = 0; i < nsolid; i++)
        if ((int)solid_cnt[i] >= auto_cutoff) lo_bloom_insert(&bloom, solid[i]);   /* K/BloomBuilder.hpp:176 */
    free(solid); free(solid_cnt);
    r->bloom = bloom.bits; r->bloom_nbytes = bloom.nbytes; r->bloom_bits = bloom.reduced_tai;

    /* ---------- phase 3+4: per-read encoding, per-block range coding ---------- */
    encoder e; memset(&e, 0, sizeof e);
    e.k = k; e.bloom = &bloom; e.res = r;
    models_init(&e.m); rc_init(&e.rc);
    kmap dict; kmap_init(&dict, 1024);
    rcenc anchor_rc; rc_init(&anchor_rc);
    o0model anchor_model; o0_init(&anchor_model, 5);
    uint64_t anchor_cap = 1024; r->anchor_kmers = (lo_kmer *)malloc(sizeof(lo_kmer) * anchor_cap);
    r->anchor_pos = (int32_t *)malloc(sizeof(int32_t) * (nreads ? nreads : 1));
    r->anchor_addr = (uint32_t *)calloc(nreads ? nreads : 1, sizeof(uint32_t));
    bytebuf dna = {0}, qtext = {0};
    uint64_t *bs = NULL, nblocks = 0, bscap = 0;
    uint64_t *qsizes = NULL; uint64_t qsz_cap = 0;
    uint64_t in_block = 0, qblock_start = 0;
    r->min_read_size = INT_MAX; r->max_read_size = 0;

    for (uint64_t ri = 0; ri < nreads; ri++) {
        uint64_t L = offsets[ri + 1] - offsets[ri];
        enc_reserve(&e, L);
        memcpy(e.seq, bases + offsets[ri], L);
        r->total_dna_size += L; r->read_count++;
        if ((int)L < r->min_read_size) r->min_read_size = (int)L;
        if ((int)L > r->max_read_size) r->max_read_size = (int)L;
        e.nN = 0;
        const uint8_t *q = (p->has_qual && quals) ? quals + offsets[ri] : NULL;

        if (L < (uint64_t)k) {                                            /* C/DnaCoder.cpp:363-371 */
            encode_noanchor_read(&e, L);
            r->anchor_pos[ri] = -1;
            if (q) smooth_quals(&e, q, L, p->lossless, &qtext);
        } else {
            /* buildKmers, C/DnaCoder.cpp:519-535 */
            for (uint64_t i = 0; i < L; i++) if (e.seq[i] == 'N') { e.Npos[e.nN++] = (int)i; e.seq[i] = 'A'; }
            uint64_t n = L - k + 1;
            lo_kmer fwd = 0;
            for (uint64_t j = 0; j < L; j++) {
                fwd = ((fwd << 2) | (lo_kmer)nt_code(e.seq[j])) & kmask(k);
                if (j + 1 >= (uint64_t)k) e.kmers[j + 1 - k] = fwd;
            }
            if (q) {
                /* storeSolidCoverageInfo, C/DnaCoder.cpp:490-516 */
                memset(e.nb_solids, 0, sizeof(int) * (L + 1));
                for (uint64_t i = 0; i < n; i++)
                    if (lo_bloom_contains(&bloom, canon(e.kmers[i], k)))
                        for (int j = 0; j < k; j++) e.nb_solids[i + j]++;
                smooth_quals(&e, q, L, p->lossless, &qtext);
            }
            /* findExistingAnchor, C/DnaCoder.cpp:563-576 */
            int anchorPos = -1; uint32_t addr = 0;
            for (uint64_t i = 0; i < n; i++)
                if (kmap_get(&dict, canon(e.kmers[i], k), &addr)) { anchorPos = (int)i; break; }
            if (anchorPos < 0) {
                /* findAndInsertAnchor, C/Leon.cpp:1105-1243: [n/2, n/2+10) then [0, n/2) then [n/2+10, n) */
                int iMin = (int)(n / 2), iMax = (int)(n / 2) + 10;
                if (iMax > (int)n) iMax = (int)n;
                int best = -1;
                for (int i = iMin; i < iMax && best < 0; i++) if (lo_bloom_contains(&bloom, canon(e.kmers[i], k))) best = i;
                for (int i = 0; i < iMin && best < 0; i++) if (lo_bloom_contains(&bloom, canon(e.kmers[i], k))) best = i;
                for (int i = iMax; i < (int)n && best < 0; i++) if (lo_bloom_contains(&bloom, canon(e.kmers[i], k))) best = i;
                if (best >= 0) {
                    lo_kmer bk = canon(e.kmers[best], k);
                    /* encodeInsertedAnchor, C/Leon.cpp:1245-1260: k symbols, most significant nucleotide first */
                    for (int i = k - 1; i >= 0; i--) rc_encode(&anchor_rc, &anchor_model, (uint8_t)((bk >> (2 * i)) & 3));
                    addr = (uint32_t)r->anchor_count;
                    kmap_put(&dict, bk, addr);
                    if (r->anchor_count == anchor_cap) { anchor_cap *= 2; r->anchor_kmers = (lo_kmer *)realloc(r->anchor_kmers, sizeof(lo_kmer) * anchor_cap); }
                    r->anchor_kmers[r->anchor_count++] = bk;
                    anchorPos = best;
                }
            }
            r->anchor_pos[ri] = anchorPos; r->anchor_addr[ri] = addr;
            if (anchorPos < 0) encode_noanchor_read(&e, L);
            else encode_anchor_read(&e, L, anchorPos, addr);
        }
        in_block++;
        if (in_block >= (uint64_t)p->reads_per_block || ri + 1 == nreads) {
            write_block(&e, in_block, &dna, &bs, &nblocks, &bscap);
            if (nblocks > qsz_cap) { qsz_cap = qsz_cap ? qsz_cap * 2 : 64; while (qsz_cap < nblocks) qsz_cap *= 2; qsizes = (uint64_t *)realloc(qsizes, sizeof(uint64_t) * qsz_cap); }
            qsizes[nblocks - 1] = qtext.n - qblock_start; qblock_start = qtext.n;
            models_clear(&e.m);
            in_block = 0;
        }
    }
    rc_flush(&anchor_rc);                                               /* C/Leon.cpp:1034-1042 */
    r->anchors_dict = anchor_rc.buf; r->anchors_dict_nbytes = anchor_rc.n;
    r->nblocks = nblocks; r->blocksizes = bs;
    r->dna = dna.p; r->dna_nbytes = dna.n;
    r->qual_text = qtext.p; r->qual_nbytes = qtext.n; r->qual_block_sizes = qsizes;
    if (nreads == 0) r->min_read_size = INT_MAX;

    models_free(&e.m); free(e.rc.buf); o0_free(&anchor_model); kmap_free(&dict);
    free(e.seq); free(e.kmers); free(e.Npos); free(e.errPos); free(e.bif); free(e.bifType); free(e.nb_solids);
    return 0;
}

void lo_free_result(lo_result *r)
{
    free(r->bloom); free(r->anchors_dict); free(r->anchor_kmers); free(r->blocksizes); free(r->dna);
    free(r->qual_text); free(r->qual_block_sizes); free(r->anchor_pos); free(r->anchor_addr);
    memset(r, 0, sizeof *r);
}
