/* leon_b200 — C ABI of the B200-native Leon compression hot path (libleon_b200.so).
 *
 * The reference (GATB/leon) has no FFI: its boundary is the C++ class `Leon : Tool` and the functors it hands to
 * `Dispatcher::iterate`.  Every entry point below replaces one of those call sites and is
 * what a cgo/JNI/ctypes/C++ host would bind.  Citations are relative to
 * /root/reference/thirdparty/gatb-core/gatb-core/src/gatb/ (C/ = tools/compression/, K/ = kmer/impl/).
 *
 * Conventions: plain pointers and sizes only; every call returns 0 or a negative LEON_E_* code and leaves a
 * message retrievable with leon_last_error(); nothing throws or aborts across the boundary.  A context is
 * single-owner (not re-entrant).  Inputs are owned by the caller and must stay valid until the call returns;
 * outputs are owned by the context (pinned host memory) and stay valid until the next leon_compress() /
 * leon_encode() on that context or leon_destroy().  There is NO CPU fallback: without a CUDA device or with
 * the extension missing every compute call fails with LEON_E_CUDA.
 */
#ifndef LEON_B200_H
#define LEON_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LEON_B200_ABI_VERSION 3

enum {
    LEON_OK = 0,
    LEON_E_ARG = -1,      /* bad argument / unsupported option value */
    LEON_E_CUDA = -2,     /* CUDA runtime error (message has the CUDA string) */
    LEON_E_NOMEM = -3,    /* host or device allocation failed */
    LEON_E_STATE = -4,    /* phase called out of order */
    LEON_E_OVERFLOW = -5, /* an internal capacity bound was exceeded (message says which) */
    LEON_E_IO = -6        /* the container file could not be written */
};

/* Option subset of the `leon -c` command line that reaches the hot path (C/Leon.cpp:102-140, 342-431). */
typedef struct leon_opts {
    int32_t kmer_size;        /* -kmer-size, default 31; 2..31 uses 64-bit keys, 32..63 uses 128-bit keys */
    int32_t reads_per_block;  /* -reads (hidden), default 50000: block = that many consecutive reads */
    int32_t has_qual;         /* 1: FASTQ with qualities kept; 0: FASTA, -noqual or -seq-only */
    int32_t lossless;         /* -lossless: qualities copied, no smoothing */
    int32_t abundance_set;    /* -abundance X given: disables the auto cutoff (its value is unused) */
    int32_t device;           /* CUDA device ordinal */
    int32_t anchor_batch;     /* reads per anchor-resolution batch (0 = default); any value gives identical output */
    int32_t dict_on_device;   /* 1: run the serial low/range recurrence of the anchor dictionary dump (C/Leon.cpp:1245-1260) as a
                                 single-warp kernel instead of on the host thread (same bytes, ~20x slower per step) */
    int32_t qual_zlib_on_device; /* 1: the GPU also writes the zlib streams of the quality blocks (leon_result.qual_z, what
                                 C/Leon.cpp:581-607 makes with host zlib; run matches + dynamic Huffman, leon_b200/csrc/deflate.cuh) */
    int32_t skip_qual_text;   /* 1: do not copy the quality text to the host (qual_text stays NULL; needs qual_zlib_on_device) */
    int32_t skip_read_arrays; /* 1: do not copy the per-read parity arrays (anchor_pos / anchor_addr, 8 bytes per read) to the host */
    int32_t reserved[5];
} leon_opts;

/* A set of reads: concatenated ASCII bases (upper-case ACGTN for reference parity) + R+1 offsets
 * (what K/../bank Sequence::getDataBuffer()/getDataSize() give the reference, C/DnaCoder.cpp:279-296),
 * and, when has_qual, the quality characters at the same offsets. */
typedef struct leon_reads {
    const uint8_t  *bases;
    const uint64_t *offsets;   /* offsets[0] = 0, offsets[R] = total bases */
    const uint8_t  *quals;     /* NULL when qualities are not kept */
    uint64_t        nreads;
} leon_reads;

/* Result of phase 1 (replaces SortingCountAlgorithm<>::execute as used by C/Leon.cpp:449-459 and the two values
 * Leon reads back from it at C/Leon.cpp:247-280). */
typedef struct leon_count_result {
    int32_t  cutoff;              /* histogram/cutoff (auto cutoff, floor 4) */
    uint64_t nb_solids_cutoff;    /* histogram/nbsolidsforcutoff */
    uint64_t nb_solid_total;      /* #distinct canonical k-mers with abundance >= 2 */
    uint64_t nb_distinct;         /* #distinct canonical k-mers */
    uint64_t nb_kmers;            /* #valid k-mer occurrences counted */
    const uint64_t *histogram;    /* [10001] abundance histogram (bin 10000 = clamp), host memory */
} leon_count_result;

/* Everything `leon -c` stores for the sequence stream (C/Leon.cpp:965-1081 endDnaCompression, writeBloom,
 * writeAnchorDict; C/Leon.cpp:653-710 writeBlock; C/Leon.cpp:581-651 writeBlockLena before zlib). */
typedef struct leon_result {
    leon_count_result count;
    /* dataset leon/bloom + its "size" attribute */
    const uint8_t *bloom;          uint64_t bloom_nbytes;   uint64_t bloom_bits;
    /* datasets leon/anchors/anchorsDict, anchorAdress */
    const uint8_t *anchors_dict;   uint64_t anchors_dict_nbytes;   uint64_t anchor_count;
    /* datasets leon/dna/dna_<id> (concatenated), blocksizes ([2*nblocks]: bytes, nreads), nb_blocks = 2*nblocks */
    uint64_t nblocks;
    const uint64_t *blocksizes;
    const uint8_t *dna;            uint64_t dna_nbytes;
    /* text handed to zlib for leon/qual/qual_<id> (concatenated), and its per-block sizes; empty without qualities */
    const uint8_t *qual_text;      uint64_t qual_nbytes;
    const uint64_t *qual_block_sizes;
    /* metadata datasets */
    uint64_t read_count, total_dna_size;
    int32_t  min_read_size, max_read_size;
    /* per-read encoding decisions, for parity checks: anchor position (-1 = read coded without anchor), address */
    const int32_t  *anchor_pos;
    const uint32_t *anchor_addr;
    uint64_t reads_without_anchor;
    /* leon/qual/qual_<id> as stored: one zlib stream per block, concatenated (only with opts.qual_zlib_on_device) */
    const uint8_t *qual_z;         uint64_t qual_z_nbytes;
    const uint64_t *qual_z_block_sizes;
} leon_result;

/* Device-side phase timings of the last leon_compress()/phase calls, CUDA events on the context's stream. */
typedef struct leon_timings {
    /* anchor_ms: the anchor batches on the main stream; the walk / entropy stage of a super-block runs beside the later
     * batches on its own streams, so walk_ms = how long the main stream then still waited for that stage (its exposed
     * tail: mostly the range coder chains of the last super-block) and entropy_ms = the wait for the dictionary coder and
     * the quality streams after that */
    float h2d_ms, count_ms, sort_ms, bloom_ms, anchor_ms, walk_ms, qual_ms, entropy_ms, d2h_ms, total_ms;
    uint64_t kernel_launches;     /* kernels launched by the library since leon_create */
    uint32_t sort_passes;         /* radix passes executed by the counting sort */
    uint32_t anchor_batches, anchor_rounds;
    float sort_pass_ms;           /* average duration of one scatter pass of the counting sort */
    uint64_t sort_keys;           /* keys per pass */
    float dict_host_ms;           /* wall time of the host thread coding the anchor dictionary (0 when on device) */
    float dict_wait_ms;           /* how long the stream's work had been finished before that thread was */
    float dict_copy_wait_ms;      /* part of dict_host_ms spent waiting for the records to arrive from the device */
    float reserved[3];
    uint64_t nsym_total;          /* symbols coded by the per-block range coders in the last run */
} leon_timings;

typedef struct leon_ctx leon_ctx;

int         leon_b200_abi_version(void);
/* Creates a context on opts->device.  Replaces Leon::Leon + Leon::execute option capture (C/Leon.cpp:81-224). */
int         leon_create(const leon_opts *opts, leon_ctx **out);
void        leon_destroy(leon_ctx *ctx);
const char *leon_last_error(const leon_ctx *ctx);   /* ctx may be NULL: error of a failed leon_create */

/* Pinned host memory for inputs, so that uploads are true DMA (optional: any host pointer is accepted). */
void       *leon_alloc_pinned(size_t nbytes);
void        leon_free_pinned(void *p);

/* One-shot: phases 1-4 for an in-memory read set.  Replaces Leon::executeCompression's sequence stream
 * (C/Leon.cpp:449-459 DSK, :903-962 startDnaCompression, :965-1081 endDnaCompression). */
int leon_compress(leon_ctx *ctx, const leon_reads *reads, leon_result *out);

/* The same work phase by phase (what the CLI and multi-step hosts call): */
int leon_upload(leon_ctx *ctx, const leon_reads *reads);                 /* H2D of the read set */
int leon_count_resident(leon_ctx *ctx, leon_count_result *out);          /* K/SortingCountAlgorithm.cpp:499-644 as used by Leon, on the uploaded set */
int leon_bloom_build(leon_ctx *ctx, const uint8_t **bits, uint64_t *nbytes, uint64_t *reduced_bits); /* C/Leon.cpp:226-338 */
int leon_encode(leon_ctx *ctx, leon_result *out);                        /* C/DnaCoder.cpp:279-405 over all blocks + range coding */
/* Device-only run on the already uploaded read set (phases 1-4, outputs left in HBM); used to time the path with
 * inputs resident in HBM.  leon_fetch() then copies the outputs to host. */
int leon_run_device(leon_ctx *ctx);
int leon_fetch(leon_ctx *ctx, leon_result *out);

int leon_get_timings(const leon_ctx *ctx, leon_timings *out);

/* Device time of the main kernels of the last run (CUDA events on the context's stream around each launch). */
/* ---- Streaming path: read sets that do not fit on the device (BASELINE configs[4]) -------------------------------------
 * The reference never holds the read set: its Dispatcher hands the functors chunks of `-reads` consecutive sequences
 * (DP/api/ICommand.hpp:304-331, driven from C/Leon.cpp:863 and :954) and DSK counts in passes over partitions
 * (K/SortingCountAlgorithm.cpp:545-552).  These entry points are what those two call sites map to: the caller feeds
 * the reads in batches, twice -- once to count, once to encode -- and the device only ever holds one batch, the k-mers
 * of one counting pass, the Bloom filter and the anchor dictionary: its footprint does not depend on the number of
 * reads.  Batches are read sets like leon_upload takes (offsets[0] = 0).
 *
 *   leon_count_begin(plan) -> npasses
 *   repeat npasses times: leon_count(batch, last = 0) ... leon_count(batch, last = 1)       (the same batches every pass)
 *   leon_count_finish(&count)          cutoff, histogram (Histogram::compute_threshold)
 *   leon_bloom_build(...)              C/Leon.cpp:226-338
 *   leon_encode_block(first_block, batch, &out) ... for every batch, in file order          DnaEncoder::operator() + writeBlock
 *   leon_encode_finish(&result)        anchor dictionary, Bloom filter, totals (C/Leon.cpp:965-1081)
 *
 * Every batch but the last must hold whole blocks (a multiple of reads_per_block reads) and, with several ranks, whole
 * rounds (a multiple of leon_comm_batch_reads(); every rank makes the same number of calls, with empty batches if it has
 * run out of reads).  Output bytes equal those of the one-shot path whatever the batching and the number of passes. */
typedef struct leon_stream_plan {
    uint64_t total_reads;     /* reads this rank will feed (per pass) */
    uint64_t total_bases;     /* their bases: bounds the number of k-mers */
    uint32_t count_passes;    /* 0: chosen from the free device memory; n: n hash-partitioned passes over the batches */
    uint32_t reserved[3];
} leon_stream_plan;
/* What one leon_encode_block call produced: the blocks [first_block, first_block + nblocks) of this rank. */
typedef struct leon_block_out {
    uint64_t first_block, nblocks, read_count;
    const uint64_t *blocksizes;        /* [2 * nblocks]: bytes, reads */
    const uint8_t  *dna;               uint64_t dna_nbytes;          /* dna_<id> of these blocks, concatenated */
    const uint8_t  *qual_text;         uint64_t qual_nbytes;         /* as leon_result, for these blocks */
    const uint64_t *qual_block_sizes;
    const uint8_t  *qual_z;            uint64_t qual_z_nbytes;
    const uint64_t *qual_z_block_sizes;
    const int32_t  *anchor_pos;        /* per read of the batch (NULL with skip_read_arrays) */
    const uint32_t *anchor_addr;
    uint64_t reserved[4];
} leon_block_out;
int leon_count_begin(leon_ctx *ctx, const leon_stream_plan *plan, uint32_t *npasses);
int leon_count(leon_ctx *ctx, const leon_reads *batch, int last);        /* Dispatcher::iterate(it, DSK functor, chunk) */
int leon_count_finish(leon_ctx *ctx, leon_count_result *out);
/* Optional, any time between two leon_count / leon_encode_block calls: starts copying the batch that will be submitted NEXT
 * into spare device buffers, so that its upload runs beside the work on the current one (the reference's bank iterator
 * reads ahead the same way, DP/api/ICommand.hpp:304-331 hands chunks to free threads).  The batch's memory must stay
 * valid and unchanged until the call it is then submitted to returns.  Since submissions are synchronous, the usual order is
 * "announce batch i + 1, submit batch i": two announcements may be open at a time -- the second one's copy starts the moment
 * the first one is swapped in by its submission and runs beside that call (a third announcement replaces the second).  A batch
 * that was not announced is uploaded by its submission; an announced batch that is never submitted is dropped by
 * leon_count_begin. */
int leon_stream_prefetch(leon_ctx *ctx, const leon_reads *next_batch, int with_quals);
/* Codes the blocks of `batch`; first_block must be the number of blocks this rank has coded so far.  The outputs are on
 * the host when it returns and stay valid until the next leon_encode_block / leon_encode_finish on the context. */
int leon_encode_block(leon_ctx *ctx, uint64_t first_block, const leon_reads *batch, leon_block_out *out);
/* Outputs of the last leon_encode_block again (the call is synchronous today; kept for hosts written against an
 * asynchronous submission). */
int leon_block_wait(leon_ctx *ctx, uint64_t first_block, leon_block_out *out);
/* Closes the encoding pass: result holds the Bloom filter, the anchor dictionary stream, the counts and the totals of the
 * whole set; its per-block fields (dna, blocksizes, quality streams) are empty -- they were handed out block by block. */
int leon_encode_finish(leon_ctx *ctx, leon_result *out);

typedef struct leon_kernel_time {
    char     name[32];
    float    ms_total;
    uint32_t launches;
} leon_kernel_time;
int leon_get_kernel_times(const leon_ctx *ctx, leon_kernel_time *out, int cap, int *n);

/* Multi-GPU (one process and one context per GPU of a node, NCCL over NVLink).  The reference has no
 * counterpart (its parallelism is the Dispatcher's threads, DP/impl/Command.cpp:119).  Rank 0 calls
 * leon_comm_unique_id and hands the 128 bytes to the other ranks (any host channel), every rank then calls
 * leon_comm_init before leon_upload.  The read set is dealt in rounds of nranks * S consecutive reads, S to each rank
 * in rank order, S = leon_comm_batch_reads() (a multiple of reads_per_block); every rank uploads its own reads in
 * file order.  Counting exchanges k-mers by hash owner (all-to-all), the Bloom filter is the OR of the ranks' parts
 * (all-gather), anchor candidates are all-gathered per batch so that addresses equal the serial reference's.  Each
 * rank returns its own blocks (block j of round g on rank r is block (g * nranks + r) * S / reads_per_block + j of the
 * file); the Bloom filter, the counts and (on rank 0) the anchor dictionary are those of the whole set.
 * 1 <= nranks <= 8 (one NVSwitch node; the per-bucket source table of the counting kernel is sized for it): more ranks are
 * refused with LEON_E_ARG and a message, nothing is capped silently. */
int leon_comm_unique_id(void *id128);
int leon_comm_init(leon_ctx *ctx, int rank, int nranks, const void *id128);
int leon_comm_batch_reads(const leon_ctx *ctx, uint64_t *reads);

/* Makes the context launch on the caller's CUDA stream (a cudaStream_t; e.g. the host framework's current stream)
 * instead of its own, so that the caller's events bracket the work.  The reference has no equivalent: its
 * Dispatcher owns its threads (DP/impl/Command.cpp:119). */
int leon_set_stream(leon_ctx *ctx, void *stream);

/* Header stream, host only.  Codes the read names (FASTA/FASTQ header lines without the leading
 * '>' / '@'; text = concatenated, nreads + 1 offsets) exactly like C/HeaderCoder.cpp:257-524 driven by C/Leon.cpp:786-899
 * with -nb-cores 1: field-wise comparison with the previous header, per-column adaptive models, one range-coded
 * stream per block of reads_per_block headers.  Blocks are coded by nthreads host threads (0 = all cores); the bytes
 * do not depend on nthreads (see leon_b200/csrc/headers.cpp on the reference's left-over state).  The result is
 * owned by the returned object until leon_headers_free(). */
typedef struct leon_header_stream {
    const uint8_t  *first_header;  uint64_t first_header_nbytes;  /* leon/header/firstheader: first name minus its first character */
    uint64_t        nblocks;
    const uint64_t *blocksizes;    /* [2 * nblocks]: bytes, headers (leon/header/blocksizes) */
    const uint8_t  *data;          uint64_t data_nbytes;          /* header_<id> streams, concatenated */
    int32_t         coded_serially;  /* 1: left-over state of the reference was consulted, the file was coded by one serial coder */
    int32_t         reserved;
} leon_header_stream;
int  leon_headers_encode(const uint8_t *text, const uint64_t *offsets, uint64_t nreads, int32_t reads_per_block,
                         int32_t nthreads, leon_header_stream **out);
void leon_headers_free(leon_header_stream *stream);

/* `.leon` container.  Writes the outputs of leon_compress()/leon_fetch() as the HDF5 file the reference's
 * `leon -d` opens: groups /metadata, /leon/{anchors,dna,qual,header} and dataset /leon/bloom with the datasets, byte
 * layouts and string attributes of C/Leon.cpp:485-531 (infobyte, kmerSize), :581-651 (qual_<id>: zlib stream of the
 * block's quality text, level 9), :653-710 (dna_<id>, header_<id>, attribute "size"), :965-1081 (bloom via
 * storage/impl/StorageTools.hpp:93-123, anchorsDict/size/anchorAdress, nb_blocks, blocksizes, read statistics).
 * Host only (no device work, usable without a GPU); the HDF5 1.8 file format is emitted directly (superblock v0,
 * symbol-table groups, contiguous u8 datasets, variable-length string attributes), the HDF5 library is not needed.
 * The header stream comes from the host header coder (C/HeaderCoder.cpp); with has_header = 0 the container is what
 * `-noheader` / `-seq-only` produce and the decoder numbers the reads.  Errors: leon_last_error(NULL). */
typedef struct leon_container_info {
    int32_t kmer_size;          /* metadata/kmerSize and the Bloom filter's kmer_size attribute */
    int32_t is_fasta;           /* 1: decompresses to FASTA, no leon/qual group (FASTA input, -noqual, -seq-only) */
    int32_t has_header;         /* 0: -noheader / -seq-only */
    int32_t reserved0;
    const uint8_t  *first_header;       uint64_t first_header_nbytes;   /* leon/header/firstheader */
    uint64_t        header_nblocks;
    const uint64_t *header_blocksizes;  /* [2 * header_nblocks]: bytes, reads (leon/header/blocksizes) */
    const uint8_t  *header_data;        /* header_<id> blocks, concatenated */
    uint64_t reserved[4];
} leon_container_info;
int leon_write_container(const char *path, const leon_result *result, const leon_container_info *info);

/* Reading a `.leon` container (the decompression side; host only).  Handles the files written by leon_write_container
 * and the ones the reference writes through the HDF5 library (chunked, extendible, unfiltered datasets,
 * storage/impl/CollectionHDF5Patch.hpp:164-200).  Dataset paths are like "leon/dna/dna_0"; the returned pointers stay
 * valid until leon_container_close(). */
typedef struct leon_container leon_container;
int  leon_container_open(const char *path, leon_container **out);
void leon_container_close(leon_container *c);
int  leon_container_has(leon_container *c, const char *path);                       /* 1 when the group or dataset exists */
int  leon_container_dataset(leon_container *c, const char *path, const uint8_t **data, uint64_t *nbytes);
int  leon_container_attr(leon_container *c, const char *dataset_path, const char *name, const char **value);

/* Decompression of the sequence stream (C/DnaCoder.cpp:1221-1622, C/Leon.cpp:1948-2015): the datasets
 * of a container (leon_container_dataset) in, the reads of every block out as text, one sequence per line.  Blocks are
 * decoded by nthreads host threads (0 = all cores): range decoding and the Bloom walk of a block are one sequential
 * chain (see leon_b200/csrc/decode.cu). */
typedef struct leon_dna_stream {
    int32_t kmer_size, max_read_size;      /* max_read_size: metadata/maxSequenceSize, bounds what a corrupt block can claim (-1: unchecked) */
    const uint8_t *bloom;         uint64_t bloom_nbytes;  uint64_t bloom_bits;      /* leon/bloom + its "size" attribute */
    const uint8_t *anchors_dict;  uint64_t anchors_dict_nbytes;  uint64_t anchor_count;
    uint64_t nblocks;  const uint64_t *blocksizes;         /* leon/dna/blocksizes: bytes, reads per block */
    const uint8_t *dna;           uint64_t dna_nbytes;     /* dna_<id> blocks, concatenated */
} leon_dna_stream;
typedef struct leon_decoded leon_decoded;     /* decoded text of a stream, block by block */
/* Inverse of leon_headers_encode (C/HeaderCoder.cpp:526-789): the names of every block, one per line. */
int  leon_headers_decode(const leon_header_stream *stream, int32_t nthreads, leon_decoded **out);
int  leon_decode_dna(const leon_dna_stream *stream, int32_t nthreads, leon_decoded **out);
/* The same stream decoded on GPU `device` (mirrors DnaDecoder::execute, C/DnaCoder.cpp:1221-1622): one
 * warp per block runs the range decoder and the walk through the filter, the lanes share each step (model search, the 28
 * bit tests of contains4).  Needs max_read_size >= 0.  kernel_ms (may be NULL): device time of the decoding kernel.
 * Fails with LEON_E_CUDA when there is no usable device; there is no silent host substitute. */
int  leon_decode_dna_device(const leon_dna_stream *stream, int32_t device, leon_decoded **out, float *kernel_ms);
int  leon_decoded_block(const leon_decoded *d, uint64_t block, const uint8_t **text, uint64_t *nbytes);
void leon_decoded_free(leon_decoded *d);

/* File level: `leon -file X -c` end to end (C/Leon.cpp:342-470 executeCompression).  Reads the FASTA/FASTQ file like the
 * reference's bank (bank/impl/BankFasta.cpp:484-570; format from the file name as C/Leon.cpp:392-416), codes the header
 * stream on host threads while the GPU codes the sequence stream, writes `out_path` (NULL: in_path + ".leon").
 * Option names are the reference's command line (C/Leon.cpp:102-140). */
typedef struct leon_file_opts {
    int32_t kmer_size;        /* -kmer-size (0 = 31) */
    int32_t reads_per_block;  /* -reads (0 = 50000) */
    int32_t lossless;         /* -lossless */
    int32_t abundance_set;    /* -abundance given */
    int32_t noheader;         /* -noheader */
    int32_t noqual;           /* -noqual */
    int32_t seq_only;         /* -seq-only = -noheader -noqual */
    int32_t device;           /* CUDA device */
    int32_t nthreads;         /* -nb-cores: host threads of the header coder and of zlib (0 = all) */
    int32_t stream_reads;     /* > 0: feed the device batches of this many reads through the streaming entry points (0: the
                                 whole set at once when it fits the device, else batches sized from the free device memory) */
    int32_t count_passes;     /* streaming: number of counting passes (0 = automatic) */
    int32_t reserved[5];
} leon_file_opts;
typedef struct leon_file_stats {
    uint64_t nreads, total_bases, nblocks, input_bytes, output_bytes;
    uint64_t header_bytes_in, header_bytes_out, dna_bytes_out;
    int32_t  is_fasta, has_header, headers_coded_serially, reserved0;
    float    parse_ms, compress_ms, header_ms, write_ms, total_ms;   /* host wall clock; compress_ms covers the GPU path and, in parallel, the header coder */
    float    reserved1[3];
    leon_timings gpu;
    /* what the reference prints in its "Abundance threshold" / "End Sequence compression" sections (C/Leon.cpp:970-1015) */
    int32_t  cutoff, abundance_auto;
    uint64_t nb_solid_kmers, anchor_count, reads_without_anchor;
    uint64_t bloom_bytes, anchors_dict_bytes, seq_bytes_out, qual_bytes_out;
} leon_file_stats;
int leon_compress_file(const char *in_path, const char *out_path, const leon_file_opts *opts, leon_file_stats *stats);
/* `leon -file X.leon -d` (C/Leon.cpp:1311-1470, 1697-1860): decodes the three streams of a container (ours or the
 * reference's) on host threads and writes <dir>/<prefix>.fasta.d or .fastq.d (out_path NULL) in the reference's
 * output layout.  stats: parse_ms = container, compress_ms = decoding, reserved1[0..1] = sequence / quality stream. */
int leon_decompress_file(const char *in_path, const char *out_path, int32_t nthreads, leon_file_stats *stats);
/* The same with the sequence stream decoded on GPU `device` (leon_decode_dna_device) while host threads decode the names
 * and the qualities; stats->reserved1[2] = device time of the decoding kernel. */
int leon_decompress_file_device(const char *in_path, const char *out_path, int32_t nthreads, int32_t device, leon_file_stats *stats);

/* The record reader of leon_compress_file on its own (bank/impl/BankFasta.cpp:484-570): reads a FASTA/FASTQ file (multi-line
 * sequences, blank lines and CR LF accepted) into the layout leon_compress takes (exactly sized pageable host memory).
 * Regular files (FASTA without quality sections, 4-line FASTQ) are read by several host threads over ranges cut at
 * record starts; anything else by one thread, record by record like the reference.  headers = the name lines
 * without '>' / '@', concatenated. */
typedef struct leon_file_reads {
    leon_reads      reads;
    const uint8_t  *headers;          /* NULL unless want_headers */
    const uint64_t *header_offsets;   /* [nreads + 1] */
    int32_t is_fastq, threads_used;
} leon_file_reads;
int  leon_read_file(const char *path, int want_quals, int want_headers, leon_file_reads **out);
void leon_file_reads_free(leon_file_reads *reads);

/* NO COUNTERPART IN THE REFERENCE.  An entry point that allocates nothing, for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the
 * k-mer slot scan of the counting stage.  d_offsets: DEVICE pointer to nreads + 1 read offsets (as leon_reads.offsets);
 * d_koff: DEVICE pointer to nreads + 1 values, d_koff[r] = number of k-mer slots (max(L - k + 1, 0) per read) of the reads
 * before r, d_koff[nreads] = their total -- what leon_upload + leon_count_resident compute as their first step
 * (leon_debug_fetch "koff").  The work is enqueued on `stream` (a cudaStream_t, NULL = default stream) of `device`.
 * The scan's scratch (tile sums) is a static device array of 64 KB per GPU, which bounds a call to
 * LEON_NOALLOC_MAX_READS reads: more is refused with LEON_E_OVERFLOW (message via leon_last_error(NULL)).  The call makes
 * no cudaMalloc / cudaFree / allocating library call; the first call on a device looks the array's address up once
 * (cudaGetSymbolAddress), so make one call outside a capture.  Every call on a device shares the array: calls that overlap
 * on the GPU (different streams without ordering) race. */
#define LEON_NOALLOC_MAX_READS (1ull << 24)
int leon_kmer_offsets_noalloc(int32_t device, const uint64_t *d_offsets, uint64_t nreads, int32_t kmer_size,
                              uint64_t *d_koff, void *stream);

/* Test hooks (parity tests only): Bloom queries on the device filter for host-provided forward k-mers.
 * kmers: n x (lo, hi) u64 pairs.  out4[i] = BloomNeighborCoherent::contains4 mask (COL Bloom.hpp:645-798),
 * out1[i] = contains() of the canonical form (Bloom.hpp:597-642). */
int leon_debug_bloom_query(leon_ctx *ctx, const uint64_t *kmers, uint64_t n, int right, uint8_t *out4, uint8_t *out1);
/* out12[i*12..]: canon lo/hi, hashpart lo/hi, racine, prefval, simplehash16(hashpart,1..6) as computed on the device. */
int leon_debug_hashes(leon_ctx *ctx, const uint64_t *kmers, uint64_t n, uint64_t *out12);
/* Copies a named device buffer of the last run to host (names listed in DESIGN.md); returns bytes via *nbytes. */
int leon_debug_fetch(leon_ctx *ctx, const char *name, void *dst, uint64_t cap, uint64_t *nbytes);

#ifdef __cplusplus
}
#endif
#endif
