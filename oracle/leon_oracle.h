/* TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
 *
 * CPU restatement (plain C, gcc) of the reference's `leon -c` hot path, used ONLY by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker for the CUDA path.
 * The product (leon_b200/, libleon_b200.so) never includes, links or calls anything in oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this restatement byte-for-byte against
 * outputs of the compiled reference (`leon -c -nb-cores 1`, fixtures under tests/golden/, generator
 * tests/golden/make_golden.py) and against the known-answer vectors of SURVEY.md section 8c.
 *
 * Path shorthand in citations (all under /root/reference/thirdparty/gatb-core/gatb-core/src/gatb/):
 *   C/ = tools/compression/   COL/ = tools/collections/impl/   M/ = tools/math/   K/ = kmer/impl/
 *   MI/ = tools/misc/
 */
#ifndef LEON_ORACLE_H
#define LEON_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef unsigned __int128 lo_kmer;

typedef struct {
    int32_t kmer_size;        /* -kmer-size, 1 < k < 64 (k<=31: LargeInt<1> arithmetic, else LargeInt<2>) */
    int32_t reads_per_block;  /* -reads, default 50000 */
    int32_t has_qual;         /* FASTQ input with qualities kept (not -noqual / -seq-only) */
    int32_t lossless;         /* -lossless */
    int32_t abundance_set;    /* -abundance given on the command line (value itself is unused) */
} lo_params;

typedef struct {
    int32_t  cutoff;             /* histogram/cutoff (auto) — 0 when abundance_set */
    uint64_t nb_solids_cutoff;   /* histogram/nbsolidsforcutoff */
    uint64_t nb_solid_total;     /* #distinct k-mers with count >= 2 (DSK solid list length) */
    uint64_t histogram[10001];
    uint64_t bloom_bits;         /* reduced_tai ("size" attribute) */
    uint64_t bloom_nbytes;
    uint8_t *bloom;
    uint64_t anchor_count;
    uint64_t anchors_dict_nbytes;
    uint8_t *anchors_dict;
    lo_kmer *anchor_kmers;       /* canonical anchor k-mers in insertion order (debug) */
    uint64_t nblocks;
    uint64_t *blocksizes;        /* [2*nblocks]: bytes, nreads */
    uint64_t dna_nbytes;
    uint8_t *dna;                /* concatenated dna_<id> payloads */
    uint64_t qual_nbytes;
    uint8_t *qual_text;          /* concatenated (smoothed) "qual\n" text, block after block */
    uint64_t *qual_block_sizes;  /* [nblocks] */
    int32_t *anchor_pos;         /* per read: anchor position, -1 = read without anchor (debug/parity) */
    uint32_t *anchor_addr;       /* per read */
    uint64_t read_count, total_dna_size;
    int32_t  min_read_size, max_read_size;
    uint64_t mc_total, mc_uniq_solid, mc_uniq_nosolid, mc_no_alternative, mc_multiple_solid, reads_without_anchor;
} lo_result;

/* primitives (exported for the known-answer tests) */
lo_kmer  lo_revcomp(lo_kmer x, int k);                         /* M/LargeInt1.pri:137-154, M/LargeInt.hpp:722-734 */
uint64_t lo_hash64(uint64_t key, uint64_t seed);               /* M/LargeInt1.pri:157-170 */
uint64_t lo_hash1(lo_kmer x, int k, uint64_t seed);            /* M/LargeInt1.pri:255 / M/LargeInt.hpp:738-748 */
uint64_t lo_simplehash16(lo_kmer x, int k, int shift);         /* M/LargeInt1.pri:190-211 / M/NativeInt64.hpp:211-221 */
uint64_t lo_bloom_seed(int idx);                               /* COL/Bloom.hpp:83-95 */
lo_kmer  lo_encode_kmer(const char *s, int k);                 /* K/Model.hpp:636-658 with (c>>1)&3 */

/* Bloom filter (BloomNeighborCoherent, 7 hashes, block 4096 bits) */
typedef struct { uint8_t *bits; uint64_t nbytes; uint64_t tai; uint64_t reduced_tai; int k; int nhash; } lo_bloom;
int  lo_bloom_init(lo_bloom *b, uint64_t tai_bloom, int k, int nhash);   /* COL/Bloom.hpp:184-199,437-442,523-552 */
void lo_bloom_free(lo_bloom *b);
void lo_bloom_insert(lo_bloom *b, lo_kmer item);                          /* COL/Bloom.hpp:555-588 */
int  lo_bloom_contains(const lo_bloom *b, lo_kmer item);                  /* COL/Bloom.hpp:597-642 */
int  lo_bloom_contains4(const lo_bloom *b, lo_kmer item, int right);      /* COL/Bloom.hpp:645-798; bit nt set iff neighbour nt present */

/* range coder known-answer helper: codes `n` (model_id, symbol) pairs; model ids 0..4 = Order0Model(alpha[id]),
 * ids >= 16: varint sets, see leon_oracle.c.  Used by the KAT in tests. */
uint64_t lo_rangecoder_kat(uint8_t *out, uint64_t cap);

/* auto cutoff, MI/impl/Histogram.cpp:62-195 (length 10000) */
void lo_compute_threshold(const uint64_t *histo /*[10001]*/, int min_auto_threshold, int32_t *cutoff, uint64_t *nbsolids);

/* whole path: DSK semantics -> cutoff -> Bloom -> anchors -> per-block DNA streams (+ smoothed qualities) */
int  lo_compress(const uint8_t *bases, const uint64_t *offsets, uint64_t nreads, const uint8_t *quals,
                 const lo_params *p, lo_result *r);
void lo_free_result(lo_result *r);

#ifdef __cplusplus
}
#endif
#endif
