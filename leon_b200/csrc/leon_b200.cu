// libleon_b200.so — context, phase orchestration and the C ABI of include/leon_b200.h.
// All compute is in the sm_100a kernels of count.cuh / bloom.cuh / anchors.cuh / walk.cuh / entropy.cuh; the host
// only sizes buffers, runs the (exact, double precision) auto-cutoff of MI/impl/Histogram.cpp:62-195 and copies
// results.  There is no CPU implementation of any kernel here: without a device every call fails.
#include "common.cuh"
#include "kmer_math.cuh"
#include "scan.cuh"
#include "count.cuh"
#include "bloom.cuh"
#include "anchors.cuh"
#include "walk.cuh"
#include "deflate.cuh"
#include "entropy.cuh"
#include "simplehash_table.h"
#include "comm.cuh"
#include <string.h>
#include <vector>
#include <algorithm>
#include <thread>
#include <atomic>
#include <chrono>
#include <deque>
#include <mutex>

namespace {

constexpr int NUM_SMS = 148;

// Histogram::compute_threshold (MI/impl/Histogram.cpp:62-195) with length 10000; min threshold forced to 4 by
// Leon (C/Leon.cpp:456).  Pure host arithmetic on 10001 bins.
void compute_threshold(const uint64_t *h, int min_auto_threshold, int32_t *cutoff_out, uint64_t *nbsolids_out)
{
    const size_t length = 10000;
    std::vector<uint64_t> sm(length + 1, 0);
    int cutoff = 0;
    uint64_t nbsolids = 0, sum_allk = 0;
    sm[1] = (uint64_t)(0.6 * (double)h[1] + 0.4 * (double)h[2]);
    sum_allk += h[1];
    int first_inc = -1, idx_max = -1;
    uint64_t max_val = 0;
    for (size_t i = 2; i < length; i++) {
        sum_allk += h[i] * i;
        sm[i] = (uint64_t)(0.2 * (double)h[i - 1] + 0.6 * (double)h[i] + 0.2 * (double)h[i + 1]);
        if (first_inc == -1 && sm[i - 1] < sm[i]) first_inc = (int)i - 1;
        if (first_inc > 0 && sm[i] > max_val) { max_val = sm[i]; idx_max = (int)i; }
    }
    sum_allk += h[length] * length;
    if (first_inc == -1) { *cutoff_out = min_auto_threshold; *nbsolids_out = 0; return; }
    uint64_t min_val = 10000000000ULL;
    int idx_min = -1;
    for (int i = first_inc; i <= idx_max; i++)
        if (sm[i] < min_val) { min_val = sm[i]; idx_min = i; }
    if (idx_min != -1) cutoff = idx_min;
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

enum Phase { PH_NONE = 0, PH_UPLOADED, PH_COUNTED, PH_BLOOM, PH_ENCODED };

}  // namespace

// tuning knobs read once from the environment (experiments; the defaults are what was measured best)
static uint64_t env_u64(const char *name, uint64_t dflt)
{
    const char *v = getenv(name);
    return v && *v ? strtoull(v, nullptr, 10) : dflt;
}

// One host thread's private launch context: stream, scratch of the scan / partition helpers, pinned scalar slot,
// launch counter and kernel timer (none of them is shared between threads).
// LEON_B200_TRACE: host-clock marks of the sections of a call, printed to stderr (no extra synchronisation: a mark shows when
// the HOST got there)
struct TraceClock {
    bool on;
    std::chrono::steady_clock::time_point t0, last;
    std::string line;
    explicit TraceClock(const char *what) : on(getenv("LEON_B200_TRACE") != nullptr)
    {
        if (on) { t0 = last = std::chrono::steady_clock::now(); line = std::string("[leon_b200 trace] ") + what + ":"; }
    }
    void mark(const char *name)
    {
        if (!on) return;
        auto t = std::chrono::steady_clock::now();
        char b[96];
        snprintf(b, sizeof b, " %s %.2f", name, std::chrono::duration<double, std::milli>(t - last).count());
        line += b;
        last = t;
    }
    ~TraceClock()
    {
        if (!on) return;
        fprintf(stderr, "%s | total %.2f ms\n", line.c_str(), std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
};

struct Lane {
    cudaStream_t st = nullptr;
    DevBuf d_scan, d_tile_hist, d_tile_off;
    PinBuf h_small;
    LaunchCounter lc;
    KernelTimer kt;
    // transient buffers of the walk / entropy stage of the SECOND lane (the first one uses the context's, which the debug
    // fetches read)
    DevBuf d_bif, d_err, d_recs, d_symoff, d_srec, d_srec_alt, d_segstart, d_chunkoff, d_chist;
};

// A super-block of the walk / entropy stage: whole blocks [first_block, first_block + nblocks) = reads [lo, hi) of this
// rank.  What its range coders read and write lives here until they have finished; the rest of the stage's buffers are
// transient and shared by all super-blocks (used in stream order on the worker's lane).
struct SbState {
    uint64_t lo = 0, hi = 0, nk = 0, nbases = 0, first_block = 0, nblocks = 0, nsym = 0;
    bool launched = false, fetched = false;
    DevBuf d_rec, d_sh, d_descs, d_outall, d_outsize;
    PinBuf h_sizes, h_descs;
    cudaStream_t rc_stream = nullptr;
    cudaEvent_t ev_anchored = nullptr, ev_records = nullptr, ev_coded = nullptr;
    KernelTimer kt;
    void ensure()
    {
        if (rc_stream) return;
        int lo_pri = 0, hi_pri = 0;
        CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
        CUDA_TRY(cudaStreamCreateWithPriority(&rc_stream, cudaStreamNonBlocking, hi_pri));   // one warp per block: never queue behind bulk kernels
        CUDA_TRY(cudaEventCreateWithFlags(&ev_anchored, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&ev_records, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&ev_coded, cudaEventDisableTiming));
    }
    void destroy()
    {
        if (rc_stream) { cudaStreamSynchronize(rc_stream); cudaStreamDestroy(rc_stream); rc_stream = nullptr; }
        if (ev_anchored) cudaEventDestroy(ev_anchored);
        if (ev_records) cudaEventDestroy(ev_records);
        if (ev_coded) cudaEventDestroy(ev_coded);
        ev_anchored = ev_records = ev_coded = nullptr;
        kt.release();
    }
};

struct leon_ctx {
    leon_opts opts;
    std::string err;
    cudaStream_t st = nullptr, st2 = nullptr, st3 = nullptr;   // st3: quality smoothing beside the anchor resolution
    cudaEvent_t ev_q[2] = {nullptr, nullptr};   // st2: the anchor dictionary coder runs beside the per-block work
    LaunchCounter lc;
    KernelTimer kt;
    bool own_stream = true;
    Phase phase = PH_NONE;
    leon_timings tm;
    cudaEvent_t ev[16];
    std::vector<uint64_t> qsizes;   // per-block size of the quality text (host)

    // read set
    uint64_t nreads = 0, nbases = 0, nkmers = 0;
    DevBuf d_bases, d_offsets, d_quals, d_koff;
    // counting
    DevBuf d_keys, d_keys_alt, d_tile_hist, d_tile_off, d_scan, d_histo, d_stats, d_tbl;
    void *solid_keys = nullptr;      // solid list (abundance >= 2): keys in d_keys or d_keys_alt, counts in d_solid_cnt
    uint64_t nsolid = 0;
    DevBuf d_retry, d_pstart_alt, d_tstart, d_mcnt, d_mcur, d_bstart, d_solid_cnt, d_bstart_all, d_keys_recv, d_solid_keys, d_pick, d_bloom_all, d_xchg;
    Comm comm;
    uint64_t nkmers_total = 0;
    // counting in pieces (count_begin / count_add / count_close / count_finish)
    uint32_t count_npasses = 1, count_pass = 0;
    uint64_t keys_fill = 0, keys_cap = 0, count_reads_total = 0, count_bases_total = 0;
    DevBuf d_keys_batch;
    PinBuf h_histo, h_small;
    leon_count_result cres;
    int32_t min_abundance = 0;
    // bloom
    DevBuf d_bloom;
    BloomDev bloom;
    PinBuf h_bloom;
    // anchors
    DevBuf d_apos, d_aaddr, d_insflag, d_plist, d_cand_pos, d_cand_key, d_cand_gid, d_cand_slot, d_state;
    DevBuf d_ct_proxy, d_ct_keys, d_ct_ins, d_ct_fu;
    DevBuf d_dict_keys[2], d_dict_addr[2], d_dict_time[2], d_dict_tag[2];
    int dict_cur = 0;
    uint64_t dict_cap = 0, dict_count = 0;
    DevBuf d_anchors, d_counters;
    // walk + entropy
    DevBuf d_bif, d_err, d_recs, d_symoff, d_srec, d_srec_alt, d_segstart, d_chunkoff, d_chist, d_rec, d_sh, d_descs, d_outall,
        d_outsize, d_dnaoff, d_dna, d_bs, d_qout;
    DevBuf d_dict_rec, d_dict_sh, d_dict_desc, d_dict_out, d_dict_size, d_scan2;
    PinBuf h_desc, h_dstage_rec[2], h_dstage_sh[2], h_dict_tot;
    DevBuf d_dict_rec2[2], d_dict_sh2[2];
    cudaEvent_t ev_dstage[2] = {nullptr, nullptr}, ev_dtot = nullptr;
    std::atomic<uint64_t> dict_pub{0};     // anchors inserted so far (published by the main thread after every batch)
    std::atomic<bool> dict_fin{false}, dict_abort{false};
    LaunchCounter lc2;                     // launches of the dictionary thread
    std::thread dict_thread;             // host side of the anchor dictionary coder (see dict_host_worker)
    std::vector<uint8_t> dict_host_out;
    std::string dict_thread_err;
    double dict_host_ms = 0, dict_copy_wait_ms = 0;
    uint64_t nblocks = 0, dna_nbytes = 0, dict_nbytes = 0, nsym_total = 0;
    PinBuf h_dna, h_bs, h_dict, h_qual, h_qsizes, h_apos, h_aaddr;
    // zlib streams of the quality blocks written on the device (deflate.cuh)
    DevBuf d_qz_huff, d_qz_blkoff, d_qz_first, d_qz_cblock, d_qz_slots, d_qz_meta, d_qz_coff, d_qz_bsize, d_qz_boff, d_qz_adler, d_qz_out;
    PinBuf h_qz_tables, h_qz_boff, h_qz_out, h_qz_sizes;
    cudaEvent_t ev_qz = nullptr;
    uint32_t qz_nchunks = 0;
    bool qz_done = false;
    uint64_t reads_without_anchor = 0;
    int32_t min_len = 0, max_len = 0;
    // encoding pass state across leon_encode_block calls (encode_begin / encode_batch / encode_end)
    bool enc_open = false;
    uint64_t enc_rounds_done = 0, enc_reads_done = 0, enc_blocks_done = 0, enc_noanchor = 0, enc_total_reads = 0;
    uint64_t enc_total_bases = 0;
    bool enc_partial_seen = false, stream_counting = false;
    uint64_t stream_reads_seen = 0;
    leon_block_out last_block_out = {};
    // chunked upload of the one-shot path (do_upload): read ranges of the pieces, their events ([8] = qualities)
    std::vector<std::pair<uint64_t, uint64_t>> up_chunks;
    cudaEvent_t ev_up[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // [9] = offsets
    bool up_quals_pending = false;
    // leon_stream_prefetch: spare read buffers and what is known about the batch in them
    DevBuf d_bases_next, d_offsets_next, d_quals_next;
    cudaStream_t st_copy = nullptr;
    cudaEvent_t ev_copy = nullptr;
    bool pf_valid = false, pf_quals = false;
    leon_reads pf_reads = {};
    int32_t pf_min = 0, pf_max = 0;
    std::vector<uint64_t> pf_qsizes;
    // one more batch announced while the prefetched one has not been submitted yet: its copy starts when that one is swapped in
    bool pfq_valid = false, pfq_quals = false;
    leon_reads pfq_reads = {};
    int32_t pfq_min = 0, pfq_max = 0;
    std::vector<uint64_t> pfq_qsizes;
    cudaEvent_t ev_pf_guard = nullptr;
    // walk / entropy stage (stage_b_worker)
    Lane wk;
    Lane wk2;                            // second lane: super-blocks are prepared alternately on two streams by two threads
    std::atomic<bool> lane_abort{false};
    std::string sb_err2; int sb_err2_code = 0;
    std::deque<SbState> sbs;
    uint64_t sbs_used = 0;
    std::atomic<uint64_t> sb_pub{0};
    std::atomic<bool> sb_abort{false}, sb_walks_done{false}, sb_failed{false};
    std::atomic<uint64_t> sb_walked{0};     // super-blocks whose records are made (sb_prepare returned)
    bool quals_late_active = false;      // this batch: quality kernels start after the last walk (encode_batch)
    std::thread sb_thread;
    std::string sb_err;
    int sb_err_code = 0;
    // caller's stream (leon_set_stream): the library's own streams are ordered after it at entry and it after them at exit
    cudaStream_t user_st = nullptr;
    cudaEvent_t ev_user_in = nullptr, ev_user_out = nullptr;
};

namespace {

template <typename T> T fetch_scalar(leon_ctx *c, const T *dptr)
{
    c->h_small.reserve(256);
    CUDA_TRY(cudaMemcpyAsync(c->h_small.p, dptr, sizeof(T), cudaMemcpyDeviceToHost, c->st));
    CUDA_TRY(cudaStreamSynchronize(c->st));
    T v;
    memcpy(&v, c->h_small.p, sizeof(T));
    return v;
}

unsigned warp_grid(uint64_t nwarps_wanted)
{
    uint64_t blocks = (nwarps_wanted + 7) / 8;
    uint64_t cap = (uint64_t)NUM_SMS * 16;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    return (unsigned)blocks;
}

constexpr int UP_CHUNKS = 8;

// chunked = the one-shot path from host buffers (leon_compress): the bases travel in UP_CHUNKS pieces on the copy stream and
// k_extract starts on each piece as it lands (count_add); the qualities follow the bases on the same stream, so the two copies
// do not share the link, and k_quals waits for them on its own stream.
void do_upload(leon_ctx *c, const leon_reads *rd, bool chunked = false)
{
    if (!rd || !rd->offsets || (rd->nreads && !rd->bases)) throw LeonError(LEON_E_ARG, "leon_upload: null read set");
    if (rd->nreads >= 0xfffffff0ull) throw LeonError(LEON_E_ARG, "leon_upload: more than 2^32-16 reads in one set");
    if (c->opts.has_qual && !rd->quals) throw LeonError(LEON_E_ARG, "leon_upload: has_qual set but quals is NULL");
    c->nreads = rd->nreads;
    c->nbases = rd->offsets[rd->nreads];
    CUDA_TRY(cudaEventRecord(c->ev[0], c->st));
    c->d_bases.reserve(c->nbases + 128);
    c->d_offsets.reserve((c->nreads + 1) * 8);
    c->up_chunks.clear();
    c->up_quals_pending = false;
    if (c->opts.has_qual) c->d_quals.reserve(c->nbases + 128);
    static const uint64_t up_chunked = env_u64("LEON_B200_UPLOAD_CHUNKED", 1);
    if (chunked && up_chunked && c->nreads >= 65536) {
        if (!c->st_copy) {
            CUDA_TRY(cudaStreamCreateWithFlags(&c->st_copy, cudaStreamNonBlocking));
            CUDA_TRY(cudaEventCreateWithFlags(&c->ev_copy, cudaEventDisableTiming));
        }
        if (!c->ev_up[0])
            for (int i = 0; i <= UP_CHUNKS + 1; i++) CUDA_TRY(cudaEventCreateWithFlags(&c->ev_up[i], cudaEventDisableTiming));
        // the offsets FIRST and on the same stream as the rest: issued on the main stream they were measured to land only after
        // everything queued on the copy stream (920 MB, 17 ms), holding up the whole counting stage
        CUDA_TRY(cudaMemcpyAsync(c->d_offsets.p, rd->offsets, (c->nreads + 1) * 8, cudaMemcpyHostToDevice, c->st_copy));
        CUDA_TRY(cudaEventRecord(c->ev_up[UP_CHUNKS + 1], c->st_copy));
        CUDA_TRY(cudaStreamWaitEvent(c->st, c->ev_up[UP_CHUNKS + 1], 0));
        const uint64_t per = ((c->nreads + UP_CHUNKS - 1) / UP_CHUNKS + 255) & ~255ull;       // whole 256-read tiles
        int i = 0;
        for (uint64_t r0 = 0; r0 < c->nreads; r0 += per, i++) {
            const uint64_t r1 = std::min(c->nreads, r0 + per), b0 = rd->offsets[r0], b1 = rd->offsets[r1];
            if (b1 > b0) CUDA_TRY(cudaMemcpyAsync(c->d_bases.as<uint8_t>() + b0, rd->bases + b0, b1 - b0, cudaMemcpyHostToDevice, c->st_copy));
            CUDA_TRY(cudaEventRecord(c->ev_up[i], c->st_copy));
            c->up_chunks.push_back({r0, r1});
        }
        if (c->opts.has_qual) {
            CUDA_TRY(cudaMemcpyAsync(c->d_quals.p, rd->quals, c->nbases, cudaMemcpyHostToDevice, c->st_copy));
            CUDA_TRY(cudaEventRecord(c->ev_up[UP_CHUNKS], c->st_copy));
            c->up_quals_pending = true;
        }
    } else {
        CUDA_TRY(cudaMemcpyAsync(c->d_bases.p, rd->bases, c->nbases, cudaMemcpyHostToDevice, c->st));
        CUDA_TRY(cudaMemcpyAsync(c->d_offsets.p, rd->offsets, (c->nreads + 1) * 8, cudaMemcpyHostToDevice, c->st));
        if (c->opts.has_qual) {
            // Qualities are first needed by k_quals on the third stream, long after counting has started: their copy goes
            // there, beside the counting kernels, instead of in front of them (one-shot calls from pinned memory: -9 ms).
            CUDA_TRY(cudaStreamWaitEvent(c->st3, c->ev[0], 0));
            CUDA_TRY(cudaMemcpyAsync(c->d_quals.p, rd->quals, c->nbases, cudaMemcpyHostToDevice, c->st3));
        }
    }
    CUDA_TRY(cudaEventRecord(c->ev[1], c->st));
    // min / max read length for the metadata datasets (host side, offsets are already here)
    int32_t mn = 0x7fffffff, mx = 0;
    for (uint64_t i = 0; i < rd->nreads; i++) {
        uint64_t L = rd->offsets[i + 1] - rd->offsets[i];
        if (L >= 0x7fffffffull) throw LeonError(LEON_E_ARG, "leon_upload: read longer than 2^31-1");
        if ((int32_t)L < mn) mn = (int32_t)L;
        if ((int32_t)L > mx) mx = (int32_t)L;
    }
    c->min_len = mn; c->max_len = mx;
    // size of the quality text of every block: sum of (L + 1) over its reads (C/DnaCoder.cpp:428-459)
    c->qsizes.clear();
    if (c->opts.has_qual) {
        uint64_t rpb = (uint64_t)c->opts.reads_per_block;
        for (uint64_t r0 = 0; r0 < rd->nreads; r0 += rpb) {
            uint64_t r1 = std::min(rd->nreads, r0 + rpb);
            c->qsizes.push_back(rd->offsets[r1] - rd->offsets[r0] + (r1 - r0));
        }
    }
    c->phase = PH_UPLOADED;
}

// out[s * (n + 1) + p] = tables[s * stride + los[p]]: the bucket boundaries the all-to-all needs, N x (N + 1) values
__global__ void k_pick_bounds(const uint64_t *__restrict__ tables, uint64_t stride, int n, const uint64_t *__restrict__ los,
                              uint64_t *__restrict__ out)
{
    int i = threadIdx.x;
    if (i >= n * (n + 1)) return;
    int s = i / (n + 1), p = i % (n + 1);
    out[i] = tables[(uint64_t)s * stride + los[p]];
}

// One stable 8-bit partition pass (tile histogram -> digit-major scan -> scatter); keys/alt are swapped so that `keys`
// points at the result.
template <typename K, typename D> void partition_pass(leon_ctx *c, K *&keys, K *&alt, uint64_t n, D digit)
{
    using namespace lbcount;
    cudaStream_t st = c->st;
    uint32_t ntiles = cdiv_u(n, RsCfg<K>::TILE);
    if (ntiles == 0) ntiles = 1;
    c->d_tile_hist.reserve((uint64_t)ntiles * 256 * 4);
    c->d_tile_off.reserve((uint64_t)ntiles * 256 * 8);
    LB_TIMED(c->kt, "k_part_hist", st, LB_LAUNCH(c->lc, (k_part_hist<K, D>), ntiles, RS_THREADS, 0, st, keys, n, digit, c->d_tile_hist.as<uint32_t>(), ntiles));
    lbscan::exclusive_scan<uint64_t>((uint64_t)ntiles * 256, HistIn{c->d_tile_hist.as<uint32_t>()},
                                     HistOut{c->d_tile_off.as<uint64_t>()}, c->d_scan, st, c->lc);
    LB_TIMED(c->kt, "k_part_scatter", st, LB_LAUNCH(c->lc, (k_part_scatter<K, D>), ntiles, RS_THREADS, 0, st, keys, alt, n, digit, c->d_tile_off.as<uint64_t>(), ntiles));
    std::swap(keys, alt);
}

// Stable LSD radix sort of n keys on the `passes` 8-bit digits starting at bit `shift0`.
template <typename K> void radix_sort(leon_ctx *c, K *&keys, K *&alt, uint64_t n, int shift0, int passes)
{
    for (int p = 0; p < passes && n > 1; p++) partition_pass(c, keys, alt, n, lbcount::ShiftDigit<K>{shift0 + 8 * p});
}

// keys of the current counting pass whose pass id (a hash of the key, independent of the partition bits) is `pass`:
// appended to dst in any order (only the multiset matters); INVALID keys are dropped
template <typename K>
__global__ void __launch_bounds__(256) k_select_pass(const K *__restrict__ src, uint64_t n, uint32_t npasses, uint32_t pass,
                                                     K *__restrict__ dst, unsigned long long *__restrict__ fill)
{
    __shared__ unsigned long long base;
    __shared__ uint32_t cnt;
    const uint64_t i0 = (uint64_t)blockIdx.x * 256 * 8;
    if (threadIdx.x == 0) cnt = 0;
    __syncthreads();
    K keep[8];
    uint32_t slot[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint64_t i = i0 + (uint64_t)j * 256 + threadIdx.x;
        slot[j] = 0xffffffffu;
        if (i < n) {
            const K key = src[i];
            if (!kmer_invalid(key) && (uint32_t)(((lbcount::mix64(lbcount::key_mix(key) + 0x51ed270b0f1fULL) >> 32) * npasses) >> 32) == pass) {
                keep[j] = key;
                slot[j] = atomicAdd(&cnt, 1u);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) base = atomicAdd(fill, (unsigned long long)cnt);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; j++)
        if (slot[j] != 0xffffffffu) dst[base + slot[j]] = keep[j];
}

// ---- counting in pieces: begin -> add (once per batch of reads) -> close (once per pass) -> finish ---------------------
// In-memory read sets use one pass and one batch (do_count).  The streaming path (leon_count(batch, last)) adds the k-mers
// of one batch at a time; with more than one pass every pass sees all batches again and keeps the keys whose pass id is
// its own, so the key buffer is 1 / npasses of the k-mers (K/SortingCountAlgorithm.cpp:545-552: DSK's passes over
// partitions, here without the disk).
template <typename K> void count_begin(leon_ctx *c, uint32_t npasses, uint64_t key_capacity)
{
    cudaStream_t st = c->st;
    c->kt.reset();
    if (!c->d_tbl.p) {
        c->d_tbl.reserve(256 * 8);
        CUDA_TRY(cudaMemcpyAsync(c->d_tbl.p, kSimpleHashTable, 256 * 8, cudaMemcpyHostToDevice, st));
    }
    c->count_npasses = npasses ? npasses : 1;
    c->count_pass = 0;
    c->keys_fill = 0;
    c->keys_cap = key_capacity;
    c->count_reads_total = 0;
    c->count_bases_total = 0;
    c->d_keys.reserve((key_capacity + 1) * sizeof(K));
    c->d_keys_alt.reserve((key_capacity + 1) * (sizeof(K) > 8 ? sizeof(K) : 8));
    c->d_histo.reserve(10001 * 8);
    c->d_stats.reserve(sizeof(lbcount::CountStats) + 16);
    CUDA_TRY(cudaMemsetAsync(c->d_histo.p, 0, 10001 * 8, st));
    CUDA_TRY(cudaMemsetAsync(c->d_stats.p, 0, sizeof(lbcount::CountStats) + 16, st));
    c->nsolid = 0;
    c->tm.sort_passes = 0;
    c->tm.sort_keys = 0;
}

// k-mer slot offsets of the reads in d_bases / d_offsets (c->nreads of them); returns the number of slots
uint64_t count_koff(leon_ctx *c)
{
    using namespace lbcount;
    cudaStream_t st = c->st;
    c->d_koff.reserve((c->nreads + 1) * 8);
    uint64_t *koff = c->d_koff.as<uint64_t>();
    uint64_t *tot = lbscan::exclusive_scan<uint64_t>(c->nreads, NkIn{c->d_offsets.as<uint64_t>(), c->opts.kmer_size}, NkOut{koff}, c->d_scan, st, c->lc);
    uint64_t nk = c->nreads ? fetch_scalar(c, tot) : 0;
    c->nkmers = nk;
    CUDA_TRY(cudaMemcpyAsync(koff + c->nreads, tot, 8, cudaMemcpyDeviceToDevice, st));
    if (c->nreads == 0) CUDA_TRY(cudaMemsetAsync(koff, 0, 8, st));
    return nk;
}

// the keys of those reads appended to the pass buffer (count_koff has run for them)
template <typename K> void count_add(leon_ctx *c)
{
    using namespace lbcount;
    const int k = c->opts.kmer_size;
    cudaStream_t st = c->st;
    uint64_t *koff = c->d_koff.as<uint64_t>();
    const uint64_t nk = c->nkmers;
    if (c->count_pass == 0) { c->count_reads_total += c->nreads; c->count_bases_total += c->nbases; }
    if (!nk) return;
    // bases of the 256 reads of a CTA are staged in shared memory when they fit (40 KB), else read from global
    uint64_t need = ((256ull * (uint64_t)c->max_len + 31) & ~15ull) + 16;
    uint32_t stage_bytes = need <= 40960 ? (uint32_t)need : 16u;
    // reads of one length: position-major stores; variable lengths: per-warp transposition (k_extract_ragged)
    static const uint64_t ragged_knob = env_u64("LEON_B200_EXTRACT_RAGGED", 1);
    const bool ragged = ragged_knob && c->min_len != c->max_len;
    auto extract = [&](uint64_t r0, uint64_t nr, K *keys) {
        if (ragged) {
            const uint32_t smem = extract_ragged_tile_bytes<K>() + stage_bytes;
            CUDA_TRY(cudaFuncSetAttribute(k_extract_ragged<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(extract_ragged_tile_bytes<K>() + 40960 + 16)));
            LB_TIMED(c->kt, "k_extract", st, LB_LAUNCH(c->lc, (k_extract_ragged<K>), cdiv_u(nr, 256), 256, smem, st, c->d_bases.as<uint8_t>(),
                      c->d_offsets.as<uint64_t>() + r0, koff + r0, nr, k, keys, stage_bytes));
        } else {
            LB_TIMED(c->kt, "k_extract", st, LB_LAUNCH(c->lc, (k_extract<K>), cdiv_u(nr, 256), 256, stage_bytes, st, c->d_bases.as<uint8_t>(),
                      c->d_offsets.as<uint64_t>() + r0, koff + r0, nr, k, keys, stage_bytes));
        }
    };
    if (c->count_npasses == 1) {
        if (c->keys_fill + nk > c->keys_cap) throw LeonError(LEON_E_OVERFLOW, "k-mer counting: more k-mers than the plan announced");
        K *keys = c->d_keys.as<K>() + c->keys_fill;
        if (!c->up_chunks.empty()) {
            // bases still arriving in pieces (do_upload, chunked): the k-mers of every piece as soon as it is there
            for (size_t i = 0; i < c->up_chunks.size(); i++) {
                const uint64_t r0 = c->up_chunks[i].first, r1 = c->up_chunks[i].second;
                CUDA_TRY(cudaStreamWaitEvent(st, c->ev_up[i], 0));
                extract(r0, r1 - r0, keys);
            }
            c->up_chunks.clear();
        } else {
            extract(0, c->nreads, keys);
        }
        c->keys_fill += nk;
    } else {
        // all keys of the batch, then the ones of this pass appended to the pass buffer
        c->d_keys_batch.reserve((nk + 1) * sizeof(K));
        K *tmp = c->d_keys_batch.as<K>();
        extract(0, c->nreads, tmp);
        unsigned long long *fill = (unsigned long long *)(c->d_stats.as<uint8_t>() + sizeof(CountStats) + 8);
        LB_TIMED(c->kt, "k_select_pass", st, LB_LAUNCH(c->lc, (k_select_pass<K>), cdiv_u(nk, 2048), 256, 0, st, tmp, nk, c->count_npasses, c->count_pass,
                  c->d_keys.as<K>(), fill));
        c->keys_fill = fetch_scalar(c, fill);
        if (c->keys_fill > c->keys_cap) throw LeonError(LEON_E_OVERFLOW, "k-mer counting: a pass holds more keys than its buffer (raise the number of passes)");
    }
}

// Partition, exchange (multi-GPU) and count the keys gathered for the current pass; solid k-mers are appended to the list.
template <typename K> void count_close(leon_ctx *c)
{
    using namespace lbcount;
    cudaStream_t st = c->st;
    const uint64_t nk = c->keys_fill;
    K *keys = c->d_keys.as<K>(), *alt = c->d_keys_alt.as<K>();
    const bool append = c->count_npasses > 1;          // several passes: the solid list lives in its own buffer and grows
    CUDA_TRY(cudaEventRecord(c->ev[3], st));

    // partition by hash into buckets of ~BcCfg::TARGET keys: LSD passes over the top `bits` bits of key_mix.
    // With N GPUs the buckets are sized on the total number of keys and rank p owns a contiguous range of them.
    const int N = c->comm.n, me = c->comm.rank;
    uint64_t nk_total = nk;
    if (c->comm.active()) {
        c->d_xchg.reserve(256);
        CUDA_TRY(cudaMemcpyAsync(c->d_xchg.p, &nk, 8, cudaMemcpyHostToDevice, st));
        NCCL_TRY(nccl_api().AllReduce(c->d_xchg.p, c->d_xchg.p, 1, ncclUint64, ncclSum, c->comm.comm, st));
        nk_total = fetch_scalar(c, c->d_xchg.as<uint64_t>());
    }
    c->nkmers_total = nk_total;
    int bits = 0;
    while (bits < 24 && ((nk_total >> bits) > (uint64_t)BcCfg<K>::TARGET || (1 << bits) < N)) bits++;
    const uint64_t nbuckets = 1ull << bits;
    // two tables of partition starts (ping-pong over the levels); the last one is the bucket table bstart
    c->d_bstart.reserve((nbuckets + 2) * 8);
    c->d_pstart_alt.reserve((nbuckets + 2) * 8);
    c->d_tstart.reserve((nbuckets + 2) * 8);
    c->d_mcnt.reserve((nbuckets + 2) * 8);
    c->d_mcur.reserve((nbuckets + 2) * 8);
    uint64_t *bstart = c->d_bstart.as<uint64_t>();
    {
        // level l handles hash bits [64 - off - w, 64 - off): most significant digit first
        int nlevels = (bits + 7) / 8;
        // the table written by the last level must be d_bstart: start so that the ping-pong ends there
        uint64_t *pcur = (nlevels & 1) ? c->d_pstart_alt.as<uint64_t>() : c->d_bstart.as<uint64_t>();
        uint64_t *pnext = (nlevels & 1) ? c->d_bstart.as<uint64_t>() : c->d_pstart_alt.as<uint64_t>();
        if (bits == 0 || nk == 0) pcur = bstart;
        // pinned staging slots of in-flight H2D copies must not be reused before the copy has run: distinct offsets of
        // h_small (sized once here, so no reallocation in between): +0 scalars read back, +1024 this table, +2048 / +4096 below
        c->h_small.reserve(8192);
        uint64_t init[2] = {0, nk};
        memcpy((char *)c->h_small.p + 1024, init, 16);
        CUDA_TRY(cudaMemcpyAsync(pcur, (char *)c->h_small.p + 1024, 16, cudaMemcpyHostToDevice, st));
        if (nk == 0 && bits) CUDA_TRY(cudaMemsetAsync(bstart, 0, (nbuckets + 1) * 8, st));
        uint64_t P = 1;
        int off = 0;
        const uint64_t TILE = RsCfg<K>::TILE;
        while (off < bits && nk > 0) {
            const int w = bits - off < 8 ? bits - off : 8;
            uint64_t *tstart = c->d_tstart.as<uint64_t>();
            uint64_t *ttot = lbscan::exclusive_scan<uint64_t>(P, MsdTilesIn{pcur, TILE}, MsdTilesOut{tstart}, c->d_scan, st, c->lc);
            CUDA_TRY(cudaMemcpyAsync(tstart + P, ttot, 8, cudaMemcpyDeviceToDevice, st));
            MsdLevel lv{pcur, tstart, P, 64 - off - w, (1u << w) - 1u, w};
            const unsigned grid = (unsigned)(nk / TILE + P + 1);      // >= number of tiles; the surplus CTAs exit at once
            unsigned long long *cnt = c->d_mcnt.as<unsigned long long>(), *cur = c->d_mcur.as<unsigned long long>();
            CUDA_TRY(cudaMemsetAsync(cnt, 0, (P << w) * 8, st));
            LB_TIMED(c->kt, "k_msd_hist", st, LB_LAUNCH(c->lc, (k_msd_hist<K>), grid, RS_THREADS, 0, st, keys, lv, cnt));
            uint64_t *ptot = lbscan::exclusive_scan<uint64_t>(P << w, MsdCntIn{cnt}, MsdCntOut{pnext, cur}, c->d_scan, st, c->lc);
            CUDA_TRY(cudaMemcpyAsync(pnext + (P << w), ptot, 8, cudaMemcpyDeviceToDevice, st));
            static const uint64_t msd_threads = env_u64("LEON_B200_MSD_THREADS", 512);
            if (msd_threads == 512) {
                LB_TIMED(c->kt, "k_msd_scatter", st, LB_LAUNCH(c->lc, (k_msd_scatter<K, 512>), grid, 512, 0, st, keys, alt, lv, cur));
            } else {
                LB_TIMED(c->kt, "k_msd_scatter", st, LB_LAUNCH(c->lc, (k_msd_scatter<K, 256>), grid, 256, 0, st, keys, alt, lv, cur));
            }
            std::swap(keys, alt);
            std::swap(pcur, pnext);
            P <<= w;
            off += w;
            c->tm.sort_passes++;
        }
    }
    c->tm.sort_keys += nk;
    CUDA_TRY(cudaEventRecord(c->ev[4], st));

    // per-bucket counting in shared-memory tables -> abundance histogram, solid list
    if (!append) c->solid_keys = alt;
    BucketSrc src;
    memset(&src, 0, sizeof src);
    const K *count_keys = keys;
    uint64_t count_buckets = nbuckets, count_n = nk;
    if (!c->comm.active()) {
        src.nsrc = 1; src.lo = 0; src.bstart[0] = bstart; src.base[0] = 0;
    } else {
        // exchange: every rank gets, from every rank, the keys of the buckets it owns (all-to-all over NVLink)
        auto lo_of = [&](int p) { return ((uint64_t)p * nbuckets + N - 1) / N; };
        c->d_bstart_all.reserve((uint64_t)N * (nbuckets + 1) * 8);
        uint64_t *ball = c->d_bstart_all.as<uint64_t>();
        NCCL_TRY(nccl_api().AllGather(bstart, ball, nbuckets + 1, ncclUint64, c->comm.comm, st));
        // boundary values bstart_s[lo_p] for all s, p
        c->d_pick.reserve((uint64_t)N * (N + 1) * 8);
        std::vector<uint64_t> los(N + 1);
        for (int p = 0; p <= N; p++) los[p] = lo_of(p);
        memcpy((char *)c->h_small.p + 2048, los.data(), (N + 1) * 8);
        CUDA_TRY(cudaMemcpyAsync(c->d_xchg.as<uint8_t>() + 64, (char *)c->h_small.p + 2048, (N + 1) * 8, cudaMemcpyHostToDevice, st));
        LB_LAUNCH(c->lc, k_pick_bounds, 1, 128, 0, st, ball, nbuckets + 1, N, (const uint64_t *)(c->d_xchg.as<uint8_t>() + 64),
                  c->d_pick.as<uint64_t>());
        std::vector<uint64_t> pick((size_t)N * (N + 1));
        CUDA_TRY(cudaMemcpyAsync((char *)c->h_small.p + 4096, c->d_pick.p, pick.size() * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        memcpy(pick.data(), (char *)c->h_small.p + 4096, pick.size() * 8);
        auto B = [&](int s_, int p) { return pick[(size_t)s_ * (N + 1) + p]; };
        uint64_t recv_total = 0;
        std::vector<uint64_t> recv_off(N);
        for (int s_ = 0; s_ < N; s_++) { recv_off[s_] = recv_total; recv_total += B(s_, me + 1) - B(s_, me); }
        c->d_keys_recv.reserve((recv_total + 1) * sizeof(K));
        K *recv = c->d_keys_recv.as<K>();
        const size_t W = sizeof(K) / 8;
        NCCL_TRY(nccl_api().GroupStart());
        for (int p = 0; p < N; p++) {
            uint64_t scnt = B(me, p + 1) - B(me, p), rcnt = B(p, me + 1) - B(p, me);
            if (p == me) continue;
            if (scnt) NCCL_TRY(nccl_api().Send(keys + B(me, p), scnt * W, ncclUint64, p, c->comm.comm, st));
            if (rcnt) NCCL_TRY(nccl_api().Recv(recv + recv_off[p], rcnt * W, ncclUint64, p, c->comm.comm, st));
        }
        NCCL_TRY(nccl_api().GroupEnd());
        {
            uint64_t cnt = B(me, me + 1) - B(me, me);
            if (cnt) CUDA_TRY(cudaMemcpyAsync(recv + recv_off[me], keys + B(me, me), cnt * sizeof(K), cudaMemcpyDeviceToDevice, st));
        }
        src.nsrc = N; src.lo = lo_of(me);
        for (int s_ = 0; s_ < N; s_++) { src.bstart[s_] = ball + (uint64_t)s_ * (nbuckets + 1); src.base[s_] = recv_off[s_] - B(s_, me); }
        count_keys = recv;
        count_buckets = lo_of(me + 1) - lo_of(me);
        count_n = recv_total;
        if (!append) {
            c->d_solid_keys.reserve((recv_total / 2 + 2) * sizeof(K));
            c->solid_keys = c->d_solid_keys.p;
        }
    }
    if (append) {     // room for the solid k-mers of this pass behind those of the earlier ones
        const uint64_t want = c->nsolid + count_n / 2 + 2;
        if (want * sizeof(K) > c->d_solid_keys.cap || want * 4 > c->d_solid_cnt.cap) {
            DevBuf nk_, nc_;
            nk_.reserve(want * sizeof(K) * 3 / 2);
            nc_.reserve(want * 4 * 3 / 2);
            if (c->nsolid) {
                CUDA_TRY(cudaMemcpyAsync(nk_.p, c->d_solid_keys.p, c->nsolid * sizeof(K), cudaMemcpyDeviceToDevice, st));
                CUDA_TRY(cudaMemcpyAsync(nc_.p, c->d_solid_cnt.p, c->nsolid * 4, cudaMemcpyDeviceToDevice, st));
            }
            CUDA_TRY(cudaStreamSynchronize(st));
            std::swap(c->d_solid_keys.p, nk_.p); std::swap(c->d_solid_keys.cap, nk_.cap);
            std::swap(c->d_solid_cnt.p, nc_.p); std::swap(c->d_solid_cnt.cap, nc_.cap);
        }
        c->solid_keys = c->d_solid_keys.p;
    }
    if (count_n && count_buckets) {
        if (!append) c->d_solid_cnt.reserve((count_n / 2 + 2) * 4);
        CUDA_TRY(cudaFuncSetAttribute(k_bucket_count<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BcCfg<K>::SMEM));
        unsigned long long *nsolid_d = (unsigned long long *)(c->d_stats.as<uint8_t>() + sizeof(CountStats));
        // first launch: small tables; buckets whose distinct keys do not fit are listed and redone with full-size tables
        c->d_retry.reserve((count_buckets + 2) * 4);
        uint32_t *retry_cnt = c->d_retry.as<uint32_t>(), *retry_list = retry_cnt + 1;
        CUDA_TRY(cudaMemsetAsync(retry_cnt, 0, 4, st));
        LB_TIMED(c->kt, "k_bucket_count", st, LB_LAUNCH(c->lc, (k_bucket_count<K>), (unsigned)count_buckets, BC_THREADS, BcCfg<K>::SMEM_SMALL, st, count_keys,
                  src, (unsigned long long *)c->d_histo.p, c->d_stats.as<CountStats>(), (K *)c->solid_keys,
                  c->d_solid_cnt.as<uint32_t>(), nsolid_d, BcCfg<K>::SLOTS_SMALL, retry_list, retry_cnt, (const uint32_t *)nullptr));
        const uint32_t nretry = fetch_scalar(c, retry_cnt);
        if (nretry)
            LB_TIMED(c->kt, "k_bucket_count", st, LB_LAUNCH(c->lc, (k_bucket_count<K>), nretry, BC_THREADS, BcCfg<K>::SMEM, st, count_keys,
                      src, (unsigned long long *)c->d_histo.p, c->d_stats.as<CountStats>(), (K *)c->solid_keys,
                      c->d_solid_cnt.as<uint32_t>(), nsolid_d, BcCfg<K>::SLOTS, (uint32_t *)nullptr, (uint32_t *)nullptr,
                      (const uint32_t *)retry_list));
        c->nsolid = fetch_scalar(c, nsolid_d);
    }
    CUDA_TRY(cudaEventRecord(c->ev[4], st));
    c->count_pass++;
    c->keys_fill = 0;
    if (c->count_npasses > 1) {
        unsigned long long *fill = (unsigned long long *)(c->d_stats.as<uint8_t>() + sizeof(CountStats) + 8);
        CUDA_TRY(cudaMemsetAsync(fill, 0, 8, st));
    }
}

// Histogram and counters of the whole set (all ranks), cutoff (Histogram::compute_threshold)
void count_finish(leon_ctx *c)
{
    using namespace lbcount;
    cudaStream_t st = c->st;
    if (c->comm.active()) {   // global histogram and counters
        NCCL_TRY(nccl_api().AllReduce(c->d_histo.p, c->d_histo.p, 10001, ncclUint64, ncclSum, c->comm.comm, st));
        NCCL_TRY(nccl_api().AllReduce(c->d_stats.p, c->d_stats.p, 3, ncclUint64, ncclSum, c->comm.comm, st));
    }
    CountStats stats = fetch_scalar(c, c->d_stats.as<CountStats>());
    if (stats.nb_valid >> 62) throw LeonError(LEON_E_OVERFLOW, "k-mer counting: a bucket could not be split any further");
    c->h_histo.reserve(10001 * 8);
    CUDA_TRY(cudaMemcpyAsync(c->h_histo.p, c->d_histo.p, 10001 * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    CUDA_TRY(cudaEventRecord(c->ev[5], st));

    leon_count_result &r = c->cres;
    r.histogram = c->h_histo.as<uint64_t>();
    r.nb_solid_total = stats.nb_solid_total;
    r.nb_distinct = stats.nb_distinct;
    r.nb_kmers = stats.nb_valid;
    compute_threshold(r.histogram, 4, &r.cutoff, &r.nb_solids_cutoff);
    c->phase = PH_COUNTED;
}

// in-memory read set: one pass, one batch
template <typename K> void do_count(leon_ctx *c)
{
    const int k = c->opts.kmer_size;
    // exact number of k-mer slots: bases - (k - 1) per read, reads shorter than k have none (upper bound used as capacity)
    CUDA_TRY(cudaEventRecord(c->ev[2], c->st));
    const uint64_t nk = count_koff(c);
    count_begin<K>(c, 1, nk + 1);
    count_add<K>(c);
    count_close<K>(c);
    count_finish(c);
    (void)k;
}

template <typename K> void do_bloom(leon_ctx *c)
{
    cudaStream_t st = c->st;
    const leon_count_result &r = c->cres;
    // C/Leon.cpp:247-335: auto cutoff unless -abundance was given; 12 bits per solid k-mer, 1000 if none
    uint64_t nbs = c->opts.abundance_set ? r.nb_solid_total : r.nb_solids_cutoff;
    c->min_abundance = c->opts.abundance_set ? 0 : r.cutoff;
    uint64_t est = (uint64_t)((double)nbs * 12);
    if (est == 0) est = 1000;
    uint64_t tai = est + 2 * 4096;                 // COL/Bloom.hpp:438
    uint64_t nbytes = 1 + tai / 8;                 // COL/Bloom.hpp:187
    if (tai && !(tai & (tai - 1))) tai--;          // COL/Bloom.hpp:193-198
    BloomDev &b = c->bloom;
    b.reduced_tai = tai - 2 * 4096;                // COL/Bloom.hpp:441
    b.magic = ~0ull / b.reduced_tai;
    b.nbytes = nbytes;
    b.k = c->opts.kmer_size;
    // with N ranks the array is cut into N equal parts of whole 16-byte units for the exchange below
    const uint64_t N_ = (uint64_t)std::max(1, c->comm.n);
    uint64_t part_words = (((nbytes + 3) / 4 + 64 / 4 + N_ - 1) / N_ + 3) & ~3ull;
    uint64_t alloc = part_words * N_ * 4;
    c->d_bloom.reserve(alloc);
    b.words = c->d_bloom.as<uint32_t>();
    CUDA_TRY(cudaMemsetAsync(b.words, 0, alloc, st));
    if (c->nsolid) {
        unsigned grid = (unsigned)std::min<uint64_t>((c->nsolid + 255) / 256, (uint64_t)NUM_SMS * 8);
        LB_TIMED(c->kt, "k_insert_solid", st, LB_LAUNCH(c->lc, (lbbloom::k_insert_solid<K>), grid, 256, 0, st, (const K *)c->solid_keys,
                  c->d_solid_cnt.as<uint32_t>(), c->nsolid, (uint32_t)c->min_abundance, b, c->d_tbl.as<uint64_t>()));
    }
    if (c->comm.active()) {
        // Every rank inserted the solid k-mers it owns into a zeroed array of the full size; the filter is the OR of the N
        // arrays.  NCCL has no bitwise reduction, so: reduce-scatter by hand (rank p receives part p of every rank's array over
        // NVLink and ORs them), then an in-place all-gather of the reduced parts.  Traffic and receive buffer are 1x the
        // filter each, whatever N (the all-gather of N whole arrays used before needed N x: 36 GB per rank at 3 Gbp).
        const int N = c->comm.n, me = c->comm.rank;
        c->d_bloom_all.reserve(alloc);
        uint32_t *recv = c->d_bloom_all.as<uint32_t>();
        NCCL_TRY(nccl_api().GroupStart());
        for (int p = 0; p < N; p++) {
            if (p == me) continue;
            NCCL_TRY(nccl_api().Send(b.words + (uint64_t)p * part_words, part_words, ncclUint32, p, c->comm.comm, st));
            NCCL_TRY(nccl_api().Recv(recv + (uint64_t)p * part_words, part_words, ncclUint32, p, c->comm.comm, st));
        }
        NCCL_TRY(nccl_api().GroupEnd());
        CUDA_TRY(cudaMemcpyAsync(recv + (uint64_t)me * part_words, b.words + (uint64_t)me * part_words, part_words * 4, cudaMemcpyDeviceToDevice, st));
        unsigned grid = (unsigned)std::min<uint64_t>((part_words + 255) / 256, (uint64_t)NUM_SMS * 8);
        LB_TIMED(c->kt, "k_or_reduce", st, LB_LAUNCH(c->lc, lbbloom::k_or_reduce, grid, 256, 0, st, recv, N, part_words, b.words + (uint64_t)me * part_words));
        NCCL_TRY(nccl_api().AllGather(b.words + (uint64_t)me * part_words, b.words, part_words, ncclUint32, c->comm.comm, st));
    }
    CUDA_TRY(cudaEventRecord(c->ev[6], st));
    c->phase = PH_BLOOM;
}

}  // namespace

namespace {

// Anchor dictionary dump (C/Leon.cpp:1034-1042,1245-1260): ONE never-reset range coder over all anchors.  The
// data-parallel part (order-0 model state and the exact reciprocal of `total` for every symbol) is computed on the
// GPU (DictCountIn/DictRecOut); what is left is the strictly serial low/range recurrence, 3 multiplies per symbol.
// A GPU thread needs ~70 ns per step of it (measured, profiles/), a host core ~6 ns, and the dump belongs to the
// orchestrator (Leon::encodeInsertedAnchor), so by default this recurrence runs on a host thread that streams the
// records out of HBM while the GPU resolves the later anchor batches, walks and codes the blocks: the dictionary is
// append-only, so the thread starts with the first batch and follows the published anchor count.
// opts.dict_on_device = 1 runs it as a kernel instead (k_range_encode on the second stream); both produce the same bytes.
constexpr uint64_t DICT_STAGE_SYMS = 1ull << 19;

template <typename K> void dict_host_worker(leon_ctx *c)
{
    auto t0 = std::chrono::steady_clock::now();
    double copy_wait = 0, idle = 0;
    try {
        CUDA_TRY(cudaSetDevice(c->opts.device));
        const int k = c->opts.kmer_size;
        const uint64_t slice_anchors = std::max<uint64_t>(1, DICT_STAGE_SYMS / (uint64_t)k);
        const uint64_t slice_syms = slice_anchors * (uint64_t)k;
        cudaStream_t s2 = c->st2;
        for (int i = 0; i < 2; i++) {
            c->d_dict_rec2[i].reserve((slice_syms + 1) * sizeof(uint4));
            c->d_dict_sh2[i].reserve(slice_syms + 1);
            c->h_dstage_rec[i].reserve(slice_syms * sizeof(uint4));
            c->h_dstage_sh[i].reserve(slice_syms);
            if (!c->ev_dstage[i]) CUDA_TRY(cudaEventCreateWithFlags(&c->ev_dstage[i], cudaEventDisableTiming));
        }
        if (!c->ev_dtot) CUDA_TRY(cudaEventCreateWithFlags(&c->ev_dtot, cudaEventDisableTiming));
        c->h_dict_tot.reserve(64);
        std::vector<uint8_t> &out = c->dict_host_out;
        out.clear();
        uint64_t low = 0, range = ~0ull;
        lbscan::U32x4 base{0, 0, 0, 0};
        uint64_t launched = 0;                 // anchors handed to the GPU so far
        struct Pending { uint64_t n; int slot; } pend[2];
        int npend = 0, next_slot = 0;
        bool first = true;
        while (true) {
            if (c->dict_abort.load()) break;
            const bool fin = c->dict_fin.load(std::memory_order_acquire);
            const uint64_t avail = c->dict_pub.load(std::memory_order_acquire);
            while (npend < 2 && launched < avail) {     // records of the next slice: scan + D2H on the second stream
                const uint64_t n = std::min(slice_anchors, avail - launched);
                const int slot = next_slot;
                next_slot ^= 1;
                const K *an = c->d_anchors.as<K>() + launched;
                lbscan::U32x4 *tot = lbscan::exclusive_scan<lbscan::U32x4>(
                    n, lbent::DictCountIn<K>{an, k},
                    lbent::DictRecOut<K>{an, k, c->d_dict_rec2[slot].as<uint4>(), c->d_dict_sh2[slot].as<uint8_t>(), base},
                    c->d_scan2, s2, c->lc2);
                CUDA_TRY(cudaMemcpyAsync(c->h_dict_tot.p, tot, sizeof(lbscan::U32x4), cudaMemcpyDeviceToHost, s2));
                CUDA_TRY(cudaEventRecord(c->ev_dtot, s2));
                CUDA_TRY(cudaMemcpyAsync(c->h_dstage_rec[slot].p, c->d_dict_rec2[slot].p, n * k * sizeof(uint4), cudaMemcpyDeviceToHost, s2));
                CUDA_TRY(cudaMemcpyAsync(c->h_dstage_sh[slot].p, c->d_dict_sh2[slot].p, n * k, cudaMemcpyDeviceToHost, s2));
                CUDA_TRY(cudaEventRecord(c->ev_dstage[slot], s2));
                auto w1 = std::chrono::steady_clock::now();
                CUDA_TRY(cudaEventSynchronize(c->ev_dtot));      // counts of this slice: base of the next one
                copy_wait += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w1).count();
                base = base + *c->h_dict_tot.as<lbscan::U32x4>();
                pend[npend++] = Pending{n, slot};
                launched += n;
            }
            if (npend == 0) {
                if (fin) break;             // fin is set after the last publication, and everything published is coded
                auto w0 = std::chrono::steady_clock::now();
                std::this_thread::sleep_for(std::chrono::microseconds(20));
                idle += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
                continue;
            }
            const Pending p = pend[0];
            pend[0] = pend[1];
            npend--;
            auto w0 = std::chrono::steady_clock::now();
            CUDA_TRY(cudaEventSynchronize(c->ev_dstage[p.slot]));
            copy_wait += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
            lbent::dict_code_slice(c->h_dstage_rec[p.slot].as<uint32_t>(), c->h_dstage_sh[p.slot].as<uint8_t>(), p.n * (uint64_t)k, first,
                            low, range, out);
            first = false;
        }
        for (int x = 0; x < 8; x++) { out.push_back((uint8_t)(low >> 56)); low <<= 8; }   // flush, RangeCoder.cpp:136-145
    } catch (const LeonError &e) {
        c->dict_thread_err = e.msg;
    } catch (const std::exception &e) {
        c->dict_thread_err = e.what();
    }
    c->dict_copy_wait_ms = copy_wait;
    c->dict_host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() - idle;
}

template <typename K> lbanchor::DictDev<K> dict_view(leon_ctx *c, int which)
{
    lbanchor::DictDev<K> d;
    d.keys = c->d_dict_keys[which].as<K>();
    d.addr = c->d_dict_addr[which].as<uint32_t>();
    d.time = c->d_dict_time[which].as<uint32_t>();
    d.tag = c->d_dict_tag[which].as<uint8_t>();
    d.mask = c->dict_cap - 1;
    return d;
}

template <typename K> void dict_alloc(leon_ctx *c, int which, uint64_t cap)
{
    c->d_dict_keys[which].reserve(cap * sizeof(K));
    c->d_dict_addr[which].reserve(cap * 4);
    c->d_dict_time[which].reserve(cap * 4);
    CUDA_TRY(cudaMemsetAsync(c->d_dict_time[which].p, 0xff, cap * 4, c->st));
    c->d_dict_tag[which].reserve(cap);
    CUDA_TRY(cudaMemsetAsync(c->d_dict_tag[which].p, 0, cap, c->st));
}

template <typename K> void dict_ensure(leon_ctx *c, uint64_t want_entries)
{
    if (want_entries * 2 <= c->dict_cap) return;
    uint64_t ncap = c->dict_cap;
    while (want_entries * 2 > ncap) ncap *= 4;
    lbanchor::DictDev<K> from = dict_view<K>(c, c->dict_cur);
    int other = c->dict_cur ^ 1;
    uint64_t old_cap = c->dict_cap;
    c->dict_cap = ncap;
    dict_alloc<K>(c, other, ncap);
    lbanchor::DictDev<K> to = dict_view<K>(c, other);
    LB_LAUNCH(c->lc, (lbanchor::k_dict_rehash<K>), cdiv_u(old_cap, 256), 256, 0, c->st, from, to);
    c->dict_cur = other;
}

// In-place variable-size all-gather: rank p's part is buf[off[p] .. off[p] + cnt[p]) (elements of `esize` bytes).
void allgather_v(leon_ctx *c, void *buf, const std::vector<uint32_t> &off, const std::vector<uint32_t> &cnt, size_t esize)
{
    NCCL_TRY(nccl_api().GroupStart());
    for (int p = 0; p < c->comm.n; p++) {
        if (!cnt[p]) continue;
        uint8_t *part = (uint8_t *)buf + (size_t)off[p] * esize;
        NCCL_TRY(nccl_api().Broadcast(part, part, (size_t)cnt[p] * esize, ncclUint8, p, c->comm.comm, c->st));
    }
    NCCL_TRY(nccl_api().GroupEnd());
}

// Reads per rank in one anchor batch.  With N ranks the read set is dealt in rounds of N * S consecutive reads, S to
// each rank in rank order (leon_comm_shard): batch g holds the g-th such round, so batches follow the file order and
// every rank works in every batch.  S is then a multiple of the block size so that blocks stay whole.
uint64_t anchor_batch_reads(const leon_ctx *c)
{
    if (c->opts.anchor_batch > 0) return (uint64_t)c->opts.anchor_batch;
    if (c->comm.n > 1) return 10ull * (uint64_t)c->opts.reads_per_block;
    // measured on B200 (4.6 M reads): 32 k -> 30 ms, 256 k -> 12 ms, 1 M -> 11 ms, one batch -> 27 ms; a whole number of
    // blocks, so that a batch is also a super-block of the walk / entropy stage
    const uint64_t rpb = (uint64_t)c->opts.reads_per_block;
    return rpb >= 524288 ? rpb : (524288 / rpb) * rpb;
}

// Reads per super-block of the walk / entropy stage: whole blocks, about one anchor batch.
uint64_t superblock_reads(const leon_ctx *c)
{
    const uint64_t rpb = (uint64_t)c->opts.reads_per_block, S = anchor_batch_reads(c);
    return S <= rpb ? rpb : (S / rpb) * rpb;
}

// ---- pipeline between the anchor resolution (main thread, in file order) and the walk / entropy stage (worker) -----
// A super-block can be walked and coded as soon as the anchors of all reads up to its end are known; its blocks'
// range coders (one warp each, tens of milliseconds of a dependent chain) then run beside everything that follows.
// The main thread records an event when the anchor batches have passed a super-block boundary and publishes it; the
// worker thread owns its own streams, scratch buffers and launch counter.
template <typename T> T fetch_scalar_on(Lane &ln, const T *dptr)
{
    ln.h_small.reserve(256);
    CUDA_TRY(cudaMemcpyAsync(ln.h_small.p, dptr, sizeof(T), cudaMemcpyDeviceToHost, ln.st));
    CUDA_TRY(cudaStreamSynchronize(ln.st));
    T v;
    memcpy(&v, ln.h_small.p, sizeof(T));
    return v;
}

// One stable 8-bit partition pass on a lane (the symbol records of a super-block)
template <typename K, typename D> void partition_pass_on(Lane &ln, K *&keys, K *&alt, uint64_t n, D digit)
{
    using namespace lbcount;
    uint32_t ntiles = cdiv_u(n, RsCfg<K>::TILE);
    if (ntiles == 0) ntiles = 1;
    ln.d_tile_hist.reserve((uint64_t)ntiles * 256 * 4);
    ln.d_tile_off.reserve((uint64_t)ntiles * 256 * 8);
    LB_TIMED(ln.kt, "k_part_hist", ln.st, LB_LAUNCH(ln.lc, (k_part_hist<K, D>), ntiles, RS_THREADS, 0, ln.st, keys, n, digit, ln.d_tile_hist.as<uint32_t>(), ntiles));
    lbscan::exclusive_scan<uint64_t>((uint64_t)ntiles * 256, HistIn{ln.d_tile_hist.as<uint32_t>()},
                                     HistOut{ln.d_tile_off.as<uint64_t>()}, ln.d_scan, ln.st, ln.lc);
    LB_TIMED(ln.kt, "k_part_scatter", ln.st, LB_LAUNCH(ln.lc, (k_part_scatter<K, D>), ntiles, RS_THREADS, 0, ln.st, keys, alt, n, digit, ln.d_tile_off.as<uint64_t>(), ntiles));
    std::swap(keys, alt);
}

__global__ void k_sb_info(const uint64_t *__restrict__ offsets, const uint64_t *__restrict__ koff, uint64_t nreads, uint64_t S,
                          uint64_t nsb, uint64_t *__restrict__ out)
{
    uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nsb) return;
    uint64_t lo = g * S, hi = lo + S < nreads ? lo + S : nreads;
    out[2 * g] = koff[hi] - koff[lo];
    out[2 * g + 1] = offsets[hi] - offsets[lo];
}

// Walk, symbol records, per-(block, model) ranks and the launch of the range coder for super-block g.
template <typename K> void sb_prepare(leon_ctx *c, uint64_t g, int lane = 0)
{
    using namespace lbent;
    Lane &ln = lane ? c->wk2 : c->wk;
    // transient buffers: the context's on lane 0, the lane's own on lane 1
    DevBuf &d_bif = lane ? ln.d_bif : c->d_bif, &d_err = lane ? ln.d_err : c->d_err, &d_recs = lane ? ln.d_recs : c->d_recs;
    DevBuf &d_symoff = lane ? ln.d_symoff : c->d_symoff, &d_srec = lane ? ln.d_srec : c->d_srec, &d_srec_alt = lane ? ln.d_srec_alt : c->d_srec_alt;
    DevBuf &d_segstart = lane ? ln.d_segstart : c->d_segstart, &d_chunkoff = lane ? ln.d_chunkoff : c->d_chunkoff, &d_chist = lane ? ln.d_chist : c->d_chist;
    cudaStream_t st = ln.st;
    SbState &sb = c->sbs[g];
    const int k = c->opts.kmer_size;
    const uint32_t rpb = (uint32_t)c->opts.reads_per_block;
    const uint64_t R = sb.hi - sb.lo;
    if (!R) return;
    const uint8_t *bases = c->d_bases.as<uint8_t>();
    const uint64_t *offsets = c->d_offsets.as<uint64_t>() + sb.lo;
    const uint64_t *koff = c->d_koff.as<uint64_t>() + sb.lo;
    const int32_t *apos = c->d_apos.as<int32_t>() + sb.lo;
    const uint32_t *aaddr = c->d_aaddr.as<uint32_t>() + sb.lo;
    // Waits are taken on the HOST, never as a stream wait sitting at the head of a queue: with more streams than hardware
    // connections (8 by default) a stream that waits for an event holds up every other stream mapped to its connection
    // (measured: the anchor kernels and the chains of other super-blocks queued behind such waits, 83 -> 199 ms per run).
    CUDA_TRY(cudaEventSynchronize(sb.ev_anchored));
    // the transient buffers of the stage are reused by the next super-block: in stream order on this lane
    d_bif.reserve(sb.nk + 1);
    d_err.reserve((sb.nk + 1) * 4);
    d_recs.reserve((R + 1) * sizeof(lbwalk::ReadRec));
    ReadRec *recs = d_recs.as<ReadRec>();
    {
        // bases of the 256 reads of a CTA are staged in shared memory when they fit (40 KB), else read from global
        uint64_t need = ((256ull * (uint64_t)c->max_len + 31) & ~15ull) + 16;
        uint32_t stage_bytes = need <= 40960 ? (uint32_t)need : 16u;
        // LEON_B200_WALK_SMEM: dynamic shared memory claimed per CTA beyond the staging need (limits the CTAs per SM, so that the
        // small kernels of the anchor stage find room beside a running walk)
        static const uint64_t walk_smem = env_u64("LEON_B200_WALK_SMEM", 0);
        size_t smem_claim = std::max<size_t>(stage_bytes, walk_smem);
        if (walk_smem) CUDA_TRY(cudaFuncSetAttribute(lbwalk::k_walk<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_claim));
        LB_TIMED(ln.kt, "k_walk", st, LB_LAUNCH(ln.lc, (lbwalk::k_walk<K>), cdiv_u(R, 256), 256, smem_claim, st, bases, offsets, koff, R, c->bloom,
                  c->d_tbl.as<uint64_t>(), apos, d_bif.as<uint8_t>(), d_err.as<uint32_t>(), recs, stage_bytes));
    }
    sb.nblocks = (R + rpb - 1) / rpb;
    if (sb.nblocks >= (1ull << 18)) throw LeonError(LEON_E_OVERFLOW, "more than 2^18 blocks in one super-block");
    d_symoff.reserve((R + 2) * 8);
    uint64_t *symoff = d_symoff.as<uint64_t>();
    SymCountIn cin{bases, offsets, koff, apos, aaddr, recs, d_bif.as<uint8_t>(), d_err.as<uint32_t>()};
    uint64_t *symtot = lbscan::exclusive_scan<uint64_t>(R, cin, SymOffOut{symoff}, ln.d_scan, st, ln.lc);
    CUDA_TRY(cudaMemcpyAsync(symoff + R, symtot, 8, cudaMemcpyDeviceToDevice, st));
    const uint64_t nsym = fetch_scalar_on(ln, symtot);
    if (nsym >= 0xffffff00ull) throw LeonError(LEON_E_OVERFLOW, "more than 2^32 symbols in one super-block");
    sb.nsym = nsym;
    d_srec.reserve((nsym + 1) * 8);
    d_srec_alt.reserve((nsym + 1) * 8);
    uint64_t *srec = d_srec.as<uint64_t>(), *salt = d_srec_alt.as<uint64_t>();
    LB_TIMED(ln.kt, "k_emit_syms", st, LB_LAUNCH(ln.lc, k_emit_syms, cdiv_u(R, 256), 256, 0, st, bases, offsets, koff, R, rpb, apos, aaddr, recs,
              d_bif.as<uint8_t>(), d_err.as<uint32_t>(), symoff, srec));
    // stable sort by (block, model): key bits [40, 46 + bits(nblocks - 1))
    int keybits = REC_MODEL_BITS;
    while (keybits < 24 && ((sb.nblocks - 1) >> (keybits - REC_MODEL_BITS))) keybits++;
    for (int p = 0; p < (keybits + 7) / 8 && nsym > 1; p++)
        partition_pass_on(ln, srec, salt, nsym, lbcount::ShiftDigit<uint64_t>{REC_KEY_SHIFT + 8 * p});
    // segments and chunks
    d_segstart.reserve((sb.nblocks * NUM_MODELS + 2) * 8);
    d_chunkoff.reserve((sb.nblocks * NUM_MODELS + 2) * 8);
    uint64_t *segstart = d_segstart.as<uint64_t>(), *chunkoff = d_chunkoff.as<uint64_t>();
    uint64_t nseg = 0, nchunks = 0;
    if (nsym) {
        uint64_t *segtot = lbscan::exclusive_scan<uint64_t>(nsym, SegHeadIn{srec}, SegHeadOut{segstart}, ln.d_scan, st, ln.lc);
        nseg = fetch_scalar_on(ln, segtot);
        CUDA_TRY(cudaMemcpyAsync(segstart + nseg, symoff + R, 8, cudaMemcpyDeviceToDevice, st));
        uint64_t *chtot = lbscan::exclusive_scan<uint64_t>(nseg, SegChunksIn{segstart}, SegChunksOut{chunkoff}, ln.d_scan, st, ln.lc);
        nchunks = fetch_scalar_on(ln, chtot);
    }
    // what the range coder reads stays with the super-block until its chains have finished
    sb.d_rec.reserve((nsym + 1) * sizeof(uint4));
    sb.d_sh.reserve(nsym + 1);
    if (nchunks) {
        d_chist.reserve(nchunks * 256 * 4);
        uint32_t *H = d_chist.as<uint32_t>();
        LB_TIMED(ln.kt, "k_chunk_hist", st, LB_LAUNCH(ln.lc, k_chunk_hist, (unsigned)nchunks, 256, 0, st, srec, segstart, chunkoff, nseg, H));
        LB_TIMED(ln.kt, "k_seg_scan", st, LB_LAUNCH(ln.lc, k_seg_scan, (unsigned)nseg, 256, 0, st, chunkoff, nseg, nchunks, H));
        LB_TIMED(ln.kt, "k_chunk_ranks", st, LB_LAUNCH(ln.lc, k_chunk_ranks, (unsigned)nchunks, CHUNK, 0, st, srec, segstart, chunkoff, nseg, H,
                  sb.d_rec.as<uint4>(), sb.d_sh.as<uint8_t>()));
    }
    sb.d_descs.reserve(sb.nblocks * sizeof(StreamDesc));
    sb.d_outall.reserve(2 * nsym + 1024 * sb.nblocks + 1024);
    sb.d_outsize.reserve(sb.nblocks * 8 + 8);
    LB_LAUNCH(ln.lc, k_block_descs, cdiv_u(sb.nblocks, 128), 128, 0, st, symoff, R, rpb, sb.nblocks, sb.d_descs.as<StreamDesc>());
    sb.h_descs.reserve(sb.nblocks * sizeof(StreamDesc) + 8);
    CUDA_TRY(cudaMemcpyAsync(sb.h_descs.p, sb.d_descs.p, sb.nblocks * sizeof(StreamDesc), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(sb.ev_records, st));
    // the chains of this super-block on their own stream: they do not hold up the next super-block's preparation
    CUDA_TRY(cudaEventSynchronize(sb.ev_records));
    // The chains that are still running when the quality kernels start (quals_late: they start after the last walk) get an SM
    // each to themselves: a CTA that claims all the shared memory of an SM shares it with no other CTA.  A lone warp per
    // scheduler is what the chain's 130 cycles per symbol were measured with; beside the warps of k_quals it ran 31 % slower
    // (153 against 138 cycles per symbol over the whole chain, +4.4 ms on the run).  Only the last RC_EXCLUSIVE super-blocks
    // (configs[2], 10 super-blocks: 0 -> 74.4 ms, 3 -> 72.8, 5 -> 71.9, 10 -> 74.5 with the walks slowed from 9.4 to 11.9 ms):
    // earlier chains have that much slack before the last one ends, and their SMs stay open to the walks still to come.
    static const uint64_t rc_exclusive = env_u64("LEON_B200_RC_EXCLUSIVE", 5);
    size_t rc_smem = 0;
    if (c->quals_late_active && g + rc_exclusive >= c->sbs_used) {
        static std::mutex mu;
        static int optin[64];                 // per device: 0 = not asked yet, -1 = refused
        std::lock_guard<std::mutex> lk(mu);
        int &max_optin = optin[c->opts.device & 63];
        if (max_optin == 0) {
            int v = 0;
            cudaFuncAttributes fa{};
            cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->opts.device);
            if (cudaFuncGetAttributes(&fa, k_range_encode) == cudaSuccess) v -= (int)fa.sharedSizeBytes;    // its own static arrays
            if (v <= 0 || cudaFuncSetAttribute(k_range_encode, cudaFuncAttributeMaxDynamicSharedMemorySize, v) != cudaSuccess) {
                cudaGetLastError();
                v = -1;
            }
            max_optin = v;
        }
        rc_smem = max_optin > 0 ? (size_t)max_optin : 0;
    }
    LB_TIMED(sb.kt, "k_range_encode", sb.rc_stream, LB_LAUNCH(ln.lc, k_range_encode, (unsigned)sb.nblocks, 32, rc_smem, sb.rc_stream, sb.d_rec.as<uint4>(), sb.d_sh.as<uint8_t>(),
              sb.d_descs.as<StreamDesc>(), sb.d_outall.as<uint8_t>(), sb.d_outsize.as<uint64_t>(), 0, c->d_counters.as<uint32_t>() + 2));
    sb.h_sizes.reserve(sb.nblocks * 8 + 8);
    CUDA_TRY(cudaMemcpyAsync(sb.h_sizes.p, sb.d_outsize.p, sb.nblocks * 8, cudaMemcpyDeviceToHost, sb.rc_stream));
    CUDA_TRY(cudaEventRecord(sb.ev_coded, sb.rc_stream));
    sb.launched = true;
}

// Output of super-block g: sizes -> block table, streams copied back to back into the host payload.
void sb_finish(leon_ctx *c, uint64_t g)
{
    using namespace lbent;
    SbState &sb = c->sbs[g];
    if (!sb.launched) return;
    const uint32_t rpb = (uint32_t)c->opts.reads_per_block;
    CUDA_TRY(cudaEventSynchronize(sb.ev_coded));
    const uint64_t *sz = sb.h_sizes.as<uint64_t>();
    const StreamDesc *descs = sb.h_descs.as<StreamDesc>();      // out_base of every block's region (k_block_descs)
    uint64_t *bs = c->h_bs.as<uint64_t>();
    uint8_t *dna = c->h_dna.as<uint8_t>();
    const uint64_t R = sb.hi - sb.lo;
    for (uint64_t b = 0; b < sb.nblocks; b++) {
        const uint64_t gb = sb.first_block + b;
        const uint64_t r0 = b * rpb, r1 = std::min<uint64_t>(R, r0 + rpb);
        if (sz[b] > descs[b].out_cap) throw LeonError(LEON_E_OVERFLOW, "range coder output exceeded its capacity bound");
        if (c->dna_nbytes + sz[b] > c->h_dna.cap) {
            // the payload outgrew the pinned buffer (sized for typical reads): the copies issued so far must have landed, then
            // a buffer twice as large takes over
            for (uint64_t e = 0; e < g; e++)
                if (c->sbs[e].fetched) CUDA_TRY(cudaEventSynchronize(c->sbs[e].ev_coded));
            CUDA_TRY(cudaStreamSynchronize(sb.rc_stream));
            PinBuf bigger;
            bigger.reserve(2 * (c->dna_nbytes + sz[b]) + (1u << 20));
            memcpy(bigger.p, c->h_dna.p, c->dna_nbytes);
            std::swap(c->h_dna.p, bigger.p); std::swap(c->h_dna.cap, bigger.cap);
            dna = c->h_dna.as<uint8_t>();
        }
        bs[2 * gb] = sz[b];
        bs[2 * gb + 1] = r1 - r0;
        if (sz[b]) CUDA_TRY(cudaMemcpyAsync(dna + c->dna_nbytes, sb.d_outall.as<uint8_t>() + descs[b].out_base, sz[b], cudaMemcpyDeviceToHost, sb.rc_stream));
        c->dna_nbytes += sz[b];
    }
    CUDA_TRY(cudaEventRecord(sb.ev_coded, sb.rc_stream));       // now: "outputs are on the host"
    sb.launched = false;
    sb.fetched = true;
}

constexpr uint64_t SB_PIPELINE_MIN_DEFAULT = 16;
static uint64_t sb_pipeline_min()
{
    static const uint64_t v = env_u64("LEON_B200_SB_PIPELINE_MIN", SB_PIPELINE_MIN_DEFAULT);
    return v;
}


// super-blocks lane, lane + nlanes, ... in turn: each waits for its anchors (all anchors when the stage is deferred)
template <typename K> void stage_b_lane(leon_ctx *c, int lane, int nlanes, uint64_t gate)
{
    const uint64_t nsb = c->sbs_used;
    for (uint64_t g = (uint64_t)lane; g < nsb; g += (uint64_t)nlanes) {
        while (c->sb_pub.load(std::memory_order_acquire) <= std::max(g, gate)) {
            if (c->sb_abort.load() || c->lane_abort.load()) return;
            std::this_thread::sleep_for(std::chrono::microseconds(10));
        }
        if (c->lane_abort.load()) return;
        sb_prepare<K>(c, g, lane);
        c->sb_walked.fetch_add(1, std::memory_order_acq_rel);
    }
}

template <typename K> void stage_b_worker(leon_ctx *c)
{
    std::thread helper;
    try {
        CUDA_TRY(cudaSetDevice(c->opts.device));
        const uint64_t nsb = c->sbs_used;
        // Small read sets (up to SB_PIPELINE_MIN super-blocks, i.e. about one wave of range coder warps): the stage starts when
        // the last anchor batch is done.  Walking super-block g beside the anchor batches g+1.. was measured on B200 at
        // 4.6 M reads (10 super-blocks): the bulk kernels take the SMs from the small latency-critical anchor kernels
        // (anchor stage 8.6 -> 15.5 ms) and that costs what the overlap gains (72.4 vs 72.2 ms; FASTQ 87.4 vs 82.8 ms), since
        // the run ends one chain length (40 ms) after the LAST chain is launched either way.  Larger sets are pipelined.
        const uint64_t gate = nsb <= sb_pipeline_min() ? nsb - 1 : 0;
        // Two lanes (LEON_B200_WALK_LANES=2; default 1): a super-block's preparation is a walk (0.8 ms at configs[2]) and a dozen
        // small kernels with four host round trips (0.4 ms); on two streams, driven by two threads, the walk of one super-block
        // runs beside the small kernels of the other.  Measured (profiles/r3b_walk_lanes.txt): configs[2] 72.2 -> 93.0 ms -- the
        // chains that claim a whole SM cannot start while the other lane's walk holds every SM -- and the configs[3] encode pass
        // 1.21 -> 1.23 s (the GPU is busy either way), so one lane it is.
        static const uint64_t lanes_knob = env_u64("LEON_B200_WALK_LANES", 1);
        const int nlanes = (lanes_knob >= 2 && nsb >= 2 && c->wk2.st) ? 2 : 1;
        c->lane_abort.store(false);
        c->sb_err2.clear();
        if (nlanes == 2)
            helper = std::thread([c, gate]() {
                try {
                    CUDA_TRY(cudaSetDevice(c->opts.device));
                    stage_b_lane<K>(c, 1, 2, gate);
                } catch (const LeonError &e) {
                    c->sb_err2 = e.msg; c->sb_err2_code = e.code; c->lane_abort.store(true);
                } catch (const std::exception &e) {
                    c->sb_err2 = e.what(); c->sb_err2_code = LEON_E_NOMEM; c->lane_abort.store(true);
                }
            });
        stage_b_lane<K>(c, 0, nlanes, gate);
        if (helper.joinable()) helper.join();
        if (!c->sb_err2.empty()) throw LeonError(c->sb_err2_code, c->sb_err2);
        if (c->sb_abort.load()) { c->sb_walks_done.store(true); return; }
        // sb_prepare returns once the records of its super-block are made (it waits for them before it launches the chains):
        // from here on only range coder chains run
        c->sb_walks_done.store(true, std::memory_order_release);
        // outputs in file order; by the time the last super-block is launched the first chains have long finished
        for (uint64_t g = 0; g < nsb; g++) sb_finish(c, g);
        for (uint64_t g = 0; g < nsb; g++)
            if (c->sbs[g].fetched) CUDA_TRY(cudaEventSynchronize(c->sbs[g].ev_coded));
    } catch (const LeonError &e) {
        c->sb_err = e.msg; c->sb_err_code = e.code; c->sb_failed.store(true);
    } catch (const std::exception &e) {
        c->sb_err = e.what(); c->sb_err_code = LEON_E_NOMEM; c->sb_failed.store(true);
    }
    if (helper.joinable()) { c->lane_abort.store(true); helper.join(); }
    c->sb_walks_done.store(true, std::memory_order_release);       // also after an error: nobody waits for ever
}

// Start of the encoding pass over a read set of R_total reads (all ranks): empty anchor dictionary, the list the anchors
// are appended to (every read inserts at most one, so it is sized once and the dictionary thread can read it while it
// grows), the host thread that codes the dictionary stream.
template <typename K> void encode_begin(leon_ctx *c, uint64_t R_total)
{
    const int k = c->opts.kmer_size;
    if (R_total >= 0xfffffff0ull) throw LeonError(LEON_E_ARG, "more than 2^32-16 reads in one read set");
    if (c->opts.dict_on_device && (uint64_t)k * R_total >= 0xffffff00ull)
        throw LeonError(LEON_E_OVERFLOW, "dict_on_device: the anchor dictionary stream could exceed 2^32 symbols");
    c->dict_cap = 1 << 16;
    c->dict_cur = 0;
    c->dict_count = 0;
    dict_alloc<K>(c, 0, c->dict_cap);
    c->tm.anchor_batches = 0;
    c->tm.anchor_rounds = 0;
    c->enc_rounds_done = 0;
    c->enc_reads_done = 0;
    c->enc_blocks_done = 0;
    c->enc_noanchor = 0;
    c->enc_total_reads = R_total;
    c->d_anchors.reserve((R_total + 1) * sizeof(K));
    if (c->comm.rank == 0 && !c->opts.dict_on_device) {
        if (c->dict_thread.joinable()) c->dict_thread.join();
        c->dict_thread_err.clear();
        c->dict_pub.store(0);
        c->dict_fin.store(false);
        c->dict_abort.store(false);
        c->dict_thread = std::thread(dict_host_worker<K>, c);
    }
    c->enc_open = true;
}

// Phase 3a: anchors for all reads, in ordered batches (see anchors.cuh); every time the batches pass the end of a
// super-block it is handed to the walk / entropy stage.
template <typename K> void do_anchors(leon_ctx *c)
{
    using namespace lbanchor;
    cudaStream_t st = c->st;
    const int k = c->opts.kmer_size;
    const uint64_t R = c->nreads;
    const int N = c->comm.n, me = c->comm.rank;
    const uint8_t *bases = c->d_bases.as<uint8_t>();
    const uint64_t *offsets = c->d_offsets.as<uint64_t>();

    int32_t *apos = c->d_apos.as<int32_t>();
    uint32_t *aaddr = c->d_aaddr.as<uint32_t>();
    uint8_t *insflag = c->d_insflag.as<uint8_t>();
    uint32_t *counters = c->d_counters.as<uint32_t>();   // [0] collision, [1] remaining, [2], [3] range coder overflow flags

    // the lookups stage the bases of the 256 reads of a CTA in shared memory when they fit (40 KB)
    const uint64_t look_need = ((256ull * (uint64_t)c->max_len + 31) & ~15ull) + 16;
    const uint32_t look_stage = look_need <= 40960 ? (uint32_t)look_need : 16u;
    const uint64_t S = anchor_batch_reads(c);
    // number of batches: the same on every rank
    uint64_t nbatches = (R + S - 1) / S;
    if (N > 1) {
        c->d_xchg.reserve(256);
        uint64_t v[2] = {nbatches, 0};
        CUDA_TRY(cudaMemcpyAsync(c->d_xchg.p, v, 16, cudaMemcpyHostToDevice, st));
        NCCL_TRY(nccl_api().AllReduce(c->d_xchg.p, c->d_xchg.p, 1, ncclUint64, ncclMax, c->comm.comm, st));
        nbatches = fetch_scalar(c, c->d_xchg.as<uint64_t>());
    }
    const uint64_t anchors_cap = c->d_anchors.cap / sizeof(K);
    const bool dict_thread_on = c->comm.rank == 0 && !c->opts.dict_on_device;
    // index in the whole read set of local read 0 of anchor batch g: the reads are dealt in rounds of N * S, S to each rank;
    // enc_rounds_done rounds (single GPU: enc_reads_done reads) were coded by earlier leon_encode_block calls
    const uint64_t delta0 = N > 1 ? c->enc_rounds_done * (uint64_t)N * S + (uint64_t)me * S : c->enc_reads_done;
    std::vector<uint32_t> m_all(N, 0), m_off(N, 0);
    uint64_t sb_next = 0;                                    // first super-block not handed over yet
    // LEON_B200_TRACE: host clock per section of the batches (the stream is drained at every mark, so the run is slower)
    static const bool trace = getenv("LEON_B200_TRACE") != nullptr;
    double tsec[6] = {0, 0, 0, 0, 0, 0};
    uint64_t m_sum = 0, ins_sum = 0, rounds0 = (uint64_t)c->tm.anchor_rounds;
    auto tnow = std::chrono::steady_clock::now();
    auto mark = [&](int i) {
        if (!trace) return;
        cudaStreamSynchronize(st);
        auto t = std::chrono::steady_clock::now();
        tsec[i] += std::chrono::duration<double, std::milli>(t - tnow).count();
        tnow = t;
    };

    for (uint64_t g = 0; g < nbatches; g++) {
        const uint64_t lo = std::min(R, g * S), hi = std::min(R, (g + 1) * S);
        const uint64_t nb = hi - lo;
        const uint64_t delta = delta0 + g * (uint64_t)(N - 1) * S;             // local read index + delta = index in the whole set
        c->tm.anchor_batches++;
        // A: lookup against the dictionary of earlier batches
        if (c->dict_count && nb)
            LB_LAUNCH(c->lc, (k_anchor_lookup<K>), cdiv_u(nb, 256), 256, look_stage, st, bases, offsets, lo, hi, delta, k,
                      dict_view<K>(c, c->dict_cur), apos, aaddr, (const uint8_t *)nullptr, look_stage);
        mark(0);
        // P: no-hit reads in order
        c->d_plist.reserve((nb + 1) * 4);
        uint32_t *plist = c->d_plist.as<uint32_t>();
        uint32_t m_own = 0;
        if (nb) {
            uint32_t *mtot = lbscan::exclusive_scan<uint32_t>(nb, PIn{offsets, apos, lo, k}, POut{plist, lo}, c->d_scan, st, c->lc);
            m_own = fetch_scalar(c, mtot);
        }
        uint32_t m = m_own, joff = 0;
        if (N > 1) {   // sizes of everybody's P list
            CUDA_TRY(cudaMemcpyAsync(c->d_xchg.as<uint32_t>() + 16 + me, &m_own, 4, cudaMemcpyHostToDevice, st));
            NCCL_TRY(nccl_api().AllGather(c->d_xchg.as<uint32_t>() + 16 + me, c->d_xchg.as<uint32_t>() + 16, 1, ncclUint32, c->comm.comm, st));
            CUDA_TRY(cudaMemcpyAsync(c->h_small.p, c->d_xchg.as<uint32_t>() + 16, 4 * N, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            memcpy(m_all.data(), c->h_small.p, 4 * N);
            uint64_t tot = 0;
            for (int p = 0; p < N; p++) { m_off[p] = (uint32_t)tot; tot += m_all[p]; }
            if (tot >= 0xfffffff0ull) throw LeonError(LEON_E_OVERFLOW, "anchor batch with more than 2^32 candidate reads");
            m = (uint32_t)tot;
            joff = m_off[me];
        }
        uint32_t nins = 0;
        mark(1);
        m_sum += m;
        if (m) {
            c->d_cand_pos.reserve((uint64_t)m * 4);
            c->d_cand_key.reserve((uint64_t)m * sizeof(K));
            c->d_cand_gid.reserve((uint64_t)m * 4);
            c->d_cand_slot.reserve((uint64_t)m * 4);
            c->d_state.reserve(m);
            int32_t *cand_pos = c->d_cand_pos.as<int32_t>();
            K *cand_key = c->d_cand_key.as<K>();
            uint32_t *cand_gid = c->d_cand_gid.as<uint32_t>();
            uint32_t *cand_slot = c->d_cand_slot.as<uint32_t>();
            uint8_t *state = c->d_state.as<uint8_t>();
            if (m_own) {
                LB_LAUNCH(c->lc, k_fill_gid, cdiv_u(m_own, 256), 256, 0, st, plist, m_own, delta, cand_gid + joff);
                LB_LAUNCH(c->lc, (k_candidates<K>), warp_grid(m_own), 256, 0, st, bases, offsets, plist, m_own, k, c->bloom,
                          c->d_tbl.as<uint64_t>(), cand_pos + joff, cand_key + joff);
            }
            if (N > 1) {   // candidates of all ranks, in the order of the whole read set
                allgather_v(c, cand_pos, m_off, m_all, 4);
                allgather_v(c, cand_key, m_off, m_all, sizeof(K));
                allgather_v(c, cand_gid, m_off, m_all, 4);
            }
            // table over candidate values
            uint64_t tcap = 1024;
            while (tcap < 2ull * m) tcap <<= 1;
            c->d_ct_proxy.reserve(tcap * 8);
            c->d_ct_keys.reserve(tcap * sizeof(K));
            c->d_ct_ins.reserve(tcap * 4);
            c->d_ct_fu.reserve(tcap * 4);
            CandTable<K> t{c->d_ct_proxy.as<uint64_t>(), c->d_ct_keys.as<K>(), c->d_ct_ins.as<uint32_t>(),
                           c->d_ct_fu.as<uint32_t>(), tcap - 1};
            uint64_t salt = 0x1234567;
            for (int attempt = 0;; attempt++) {
                CUDA_TRY(cudaMemsetAsync(t.proxy, 0xff, tcap * 8, st));
                CUDA_TRY(cudaMemsetAsync(t.ins, 0xff, tcap * 4, st));
                LB_LAUNCH(c->lc, (k_cand_build<K>), cdiv_u(m, 256), 256, 0, st, cand_pos, cand_key, m, t, salt, cand_slot, state);
                if (sizeof(K) == 8) break;
                CUDA_TRY(cudaMemsetAsync(counters, 0, 4, st));
                LB_LAUNCH(c->lc, (k_cand_verify<K>), cdiv_u(m, 256), 256, 0, st, cand_key, cand_slot, m, t, counters);
                if (!fetch_scalar(c, counters)) break;
                if (attempt > 8) throw LeonError(LEON_E_OVERFLOW, "anchor candidate table: repeated proxy collisions");
                salt = salt * 6364136223846793005ULL + 1442695040888963407ULL;
            }
            // R: rounds of deterministic reservations; every rank decides its own reads, states are shared per round.
            // A round after the last undecided read changes nothing, so rounds are launched in groups and the "undecided"
            // counter is only read once per group: fewer host round trips than one per round.
            const int group = N > 1 ? 1 : 3;
            mark(2);
            while (true) {
                for (int it = 0; it < group; it++) {
                    c->tm.anchor_rounds++;
                    CUDA_TRY(cudaMemsetAsync(t.fu, 0xff, tcap * 4, st));
                    CUDA_TRY(cudaMemsetAsync(counters + 1, 0, 4, st));
                    LB_LAUNCH(c->lc, k_round_fu, cdiv_u(m, 256), 256, 0, st, state, cand_slot, m, t.fu);
                    if (m_own)
                        LB_LAUNCH(c->lc, (k_round_decide<K>), warp_grid(m_own), 256, 0, st, bases, offsets, plist, m_own, joff, k, t, salt,
                                  cand_slot, state, counters + 1);
                    if (N > 1) {
                        allgather_v(c, state, m_off, m_all, 1);
                        LB_LAUNCH(c->lc, k_ins_update, cdiv_u(m, 256), 256, 0, st, state, cand_slot, m, t.ins);
                        NCCL_TRY(nccl_api().AllReduce(counters + 1, counters + 1, 1, ncclUint32, ncclSum, c->comm.comm, st));
                    }
                }
                if (!fetch_scalar(c, counters + 1)) break;
            }
            mark(3);
            // I: addresses in read order, insertion into D (every rank inserts every new anchor)
            dict_ensure<K>(c, c->dict_count + m);
            if (c->dict_count + m + 1 > anchors_cap) throw LeonError(LEON_E_OVERFLOW, "anchor list: more anchors than reads");
            InsOut<K> io{cand_gid, cand_pos, cand_key, dict_view<K>(c, c->dict_cur), (uint32_t)c->dict_count,
                         c->d_anchors.as<K>(), apos, aaddr, insflag, joff, joff + m_own, delta};
            uint32_t *itot = lbscan::exclusive_scan<uint32_t>(m, InsIn{state}, io, c->d_scan, st, c->lc);
            nins = fetch_scalar(c, itot);
            c->dict_count += nins;
            ins_sum += nins;
            mark(4);
            if (dict_thread_on) c->dict_pub.store(c->dict_count, std::memory_order_release);   // the stream is idle: the new anchors are in HBM
            // B: anchors of this batch can pre-empt later reads of the batch at earlier positions
            if (nins && nb)
                LB_LAUNCH(c->lc, (k_anchor_lookup<K>), cdiv_u(nb, 256), 256, look_stage, st, bases, offsets, lo, hi, delta, k,
                          dict_view<K>(c, c->dict_cur), apos, aaddr, (const uint8_t *)insflag, look_stage);
        }
        mark(5);
        // hand over every super-block whose reads are all anchored now
        while (sb_next < c->sbs_used && (c->sbs[sb_next].hi <= hi || g + 1 == nbatches)) {
            CUDA_TRY(cudaEventRecord(c->sbs[sb_next].ev_anchored, st));
            sb_next++;
            c->sb_pub.store(sb_next, std::memory_order_release);
        }
    }
    while (sb_next < c->sbs_used) {      // no batches at all on this rank
        CUDA_TRY(cudaEventRecord(c->sbs[sb_next].ev_anchored, st));
        sb_next++;
        c->sb_pub.store(sb_next, std::memory_order_release);
    }
    c->enc_rounds_done += nbatches;
    c->enc_reads_done += R;
    if (trace)
        fprintf(stderr, "[leon_b200 trace] anchors: %llu reads, %llu batches, %llu rounds, %llu no-hit reads, %llu inserted, dictionary %llu | ms: "
                        "lookup A %.2f, no-hit list %.2f, candidates+table %.2f, rounds %.2f, insert %.2f, lookup B %.2f\n",
                (unsigned long long)R, (unsigned long long)nbatches, (unsigned long long)((uint64_t)c->tm.anchor_rounds - rounds0),
                (unsigned long long)m_sum, (unsigned long long)ins_sum, (unsigned long long)c->dict_count,
                tsec[0], tsec[1], tsec[2], tsec[3], tsec[4], tsec[5]);
    CUDA_TRY(cudaEventRecord(c->ev[7], st));
}

__global__ void k_count_noanchor(const int32_t *__restrict__ apos, uint64_t n, unsigned long long *out)
{
    unsigned long long v = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        v += apos[i] < 0;
    for (int d = 16; d; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(out, v);
}

// Quality blocks -> zlib streams on the third stream, right behind k_quals (deflate.cuh).  The chunk tables are built on
// the host from the block sizes known since leon_upload.
void deflate_quals(leon_ctx *c)
{
    using namespace lbdeflate;
    const uint64_t nb = c->qsizes.size();
    uint64_t nchunks = 0;
    for (uint64_t b = 0; b < nb; b++) nchunks += (c->qsizes[b] + DF_CHUNK - 1) / DF_CHUNK;
    if (nchunks >= 0xffffffffull) throw LeonError(LEON_E_OVERFLOW, "deflate_quals: too many chunks");
    c->qz_nchunks = (uint32_t)nchunks;
    // one pinned table: blk_off[nb + 1] (u64) | chunk_first[nb + 1] (u32) | chunk_block[nchunks] (u32)
    const size_t off_first = (nb + 1) * 8, off_cblock = off_first + (nb + 1) * 4, tbytes = off_cblock + nchunks * 4;
    c->h_qz_tables.reserve(tbytes + 16);
    uint64_t *blk_off = (uint64_t *)c->h_qz_tables.p;
    uint32_t *first = (uint32_t *)((char *)c->h_qz_tables.p + off_first), *cblock = (uint32_t *)((char *)c->h_qz_tables.p + off_cblock);
    uint64_t at = 0, ch = 0;
    for (uint64_t b = 0; b < nb; b++) {
        blk_off[b] = at; first[b] = (uint32_t)ch;
        const uint64_t k = (c->qsizes[b] + DF_CHUNK - 1) / DF_CHUNK;
        for (uint64_t i = 0; i < k; i++) cblock[ch++] = (uint32_t)b;
        at += c->qsizes[b];
    }
    blk_off[nb] = at; first[nb] = (uint32_t)ch;
    cudaStream_t s3 = c->st3;
    c->d_qz_blkoff.reserve(tbytes + 16);
    CUDA_TRY(cudaMemcpyAsync(c->d_qz_blkoff.p, c->h_qz_tables.p, tbytes, cudaMemcpyHostToDevice, s3));
    const uint64_t *d_blk_off = c->d_qz_blkoff.as<uint64_t>();
    const uint32_t *d_first = (const uint32_t *)((const char *)c->d_qz_blkoff.p + off_first);
    const uint32_t *d_cblock = (const uint32_t *)((const char *)c->d_qz_blkoff.p + off_cblock);
    c->d_qz_slots.reserve(nchunks * (uint64_t)DF_SLOT + 16);
    c->d_qz_meta.reserve(nchunks * sizeof(ChunkMeta) + 16);
    c->d_qz_coff.reserve(nchunks * 4 + 16);
    c->d_qz_bsize.reserve(nb * 8 + 16);
    c->d_qz_boff.reserve((nb + 1) * 8 + 16);
    c->d_qz_adler.reserve(nb * 4 + 16);
    c->d_qz_out.reserve(at + nchunks * 16 + nb * 8 + 64);       // every chunk is at worst stored: text + 5 bytes
    c->d_qz_huff.reserve(nchunks * sizeof(HuffState) + 16);
    if (nchunks) {
        LB_TIMED(c->kt, "k_deflate_count", s3, LB_LAUNCH(c->lc, k_deflate_count, (unsigned)nchunks, DF_THREADS, 0, s3,
                 c->d_qout.as<uint8_t>(), d_blk_off, d_first, d_cblock, c->d_qz_huff.as<HuffState>(), c->d_qz_meta.as<ChunkMeta>()));
        CUDA_TRY(cudaFuncSetAttribute(k_deflate_codes, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(DF_CODES_CHUNKS * sizeof(HuffState))));
        LB_TIMED(c->kt, "k_deflate_codes", s3, LB_LAUNCH(c->lc, k_deflate_codes, (unsigned)((nchunks + DF_CODES_CHUNKS - 1) / DF_CODES_CHUNKS),
                 DF_CODES_CHUNKS * 32, DF_CODES_CHUNKS * sizeof(HuffState), s3, c->d_qz_huff.as<HuffState>(), (uint32_t)nchunks));
        LB_TIMED(c->kt, "k_deflate_emit", s3, LB_LAUNCH(c->lc, k_deflate_emit, (unsigned)nchunks, DF_THREADS, 0, s3,
                 c->d_qout.as<uint8_t>(), d_blk_off, d_first, d_cblock, c->d_qz_huff.as<HuffState>(), c->d_qz_slots.as<uint8_t>(),
                 c->d_qz_meta.as<ChunkMeta>()));
        LB_LAUNCH(c->lc, k_deflate_layout, 1, 1024, 0, s3, c->d_qz_meta.as<ChunkMeta>(), d_first, (uint32_t)nb, c->d_qz_coff.as<uint32_t>(),
                  c->d_qz_bsize.as<uint64_t>(), c->d_qz_boff.as<uint64_t>(), c->d_qz_adler.as<uint32_t>());
        LB_TIMED(c->kt, "k_deflate_gather", s3, LB_LAUNCH(c->lc, k_deflate_gather, (unsigned)nchunks, 256, 0, s3, c->d_qz_slots.as<uint8_t>(),
                 c->d_qz_meta.as<ChunkMeta>(), d_first, d_cblock, c->d_qz_coff.as<uint32_t>(), c->d_qz_boff.as<uint64_t>(),
                 c->d_qz_adler.as<uint32_t>(), c->d_qz_out.as<uint8_t>()));
        // the layout goes to the host at once so that leon_fetch knows how many bytes to copy
        c->h_qz_boff.reserve((nb + 1) * 8 + 16);
        CUDA_TRY(cudaMemcpyAsync(c->h_qz_boff.p, c->d_qz_boff.p, (nb + 1) * 8, cudaMemcpyDeviceToHost, s3));
    }
    CUDA_TRY(cudaEventRecord(c->ev_qz, s3));
    c->qz_done = true;
}

// Quality stream, anchors, walk and entropy coding of the reads in d_bases / d_offsets (/ d_quals): all blocks of these
// reads are coded and on the host when it returns.  encode_begin has run; count_koff has run for these reads.
template <typename K> void encode_batch(leon_ctx *c)
{
    using namespace lbent;
    cudaStream_t st = c->st;
    const int k = c->opts.kmer_size;
    const uint64_t R = c->nreads;
    const uint32_t rpb = (uint32_t)c->opts.reads_per_block;
    const uint8_t *bases = c->d_bases.as<uint8_t>();
    const uint64_t *offsets = c->d_offsets.as<uint64_t>();

    // Quality smoothing only needs the Bloom filter: it runs on a third (low priority) stream; the run ends when its zlib
    // streams are written too.  WHEN it starts: with up to SB_PIPELINE_MIN super-blocks (deferred start of the walk / entropy
    // stage, see stage_b_worker) the run ends one range coder chain (38 ms of ~100 lone warps on an otherwise idle GPU) after
    // the last walk, so the quality kernels (8 ms + 5 ms of deflate at configs[2]) go THERE; started beside the anchor stage
    // they took the SMs from its small latency-critical kernels (k_quals 7.9 ms alone, 15 ms beside them, and the anchor
    // stage as much longer).  Larger sets (pipelined stage) have no idle tail to speak of: the kernels start at once.
    // LEON_B200_QUALS_LATE=0 restores the early start.
    c->qz_done = false;
    static const uint64_t quals_late_knob = env_u64("LEON_B200_QUALS_LATE", 1);
    auto launch_quals = [&]() {
        if (!c->opts.has_qual) return;
        c->d_qout.reserve(c->nbases + R + 16);
        if (!R) return;
        CUDA_TRY(cudaEventRecord(c->ev_q[0], st));
        CUDA_TRY(cudaStreamWaitEvent(c->st3, c->ev_q[0], 0));
        if (c->up_quals_pending) {      // qualities copied behind the bases on the copy stream (do_upload, chunked)
            CUDA_TRY(cudaStreamWaitEvent(c->st3, c->ev_up[UP_CHUNKS], 0));
            c->up_quals_pending = false;
        }
        {
            // bases and the output text of the QT reads of a CTA are staged in shared memory when they fit
            uint64_t need = (((uint64_t)lbwalk::QT * (uint64_t)c->max_len + 31) & ~15ull) + 16;
            uint32_t stage_bytes = need <= 100000 ? (uint32_t)need : 16u;
            size_t smem = 2 * (size_t)stage_bytes + lbwalk::QT + 32;
            // 4 k-mers in flight per thread, 4 CTAs per SM: the best of the (2..8, 2..8) combinations measured on the B200
            CUDA_TRY(cudaFuncSetAttribute(lbwalk::k_quals<K, 4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 100000 + lbwalk::QT + 32));
            LB_TIMED(c->kt, "k_quals", c->st3, LB_LAUNCH(c->lc, (lbwalk::k_quals<K, 4, 4>), cdiv_u(R, lbwalk::QT), lbwalk::QT, smem, c->st3, bases,
                      c->d_quals.as<uint8_t>(), offsets, R, c->bloom, c->d_tbl.as<uint64_t>(), c->opts.lossless, c->d_qout.as<uint8_t>(), stage_bytes));
        }
        CUDA_TRY(cudaEventRecord(c->ev_q[1], c->st3));
        if (c->opts.qual_zlib_on_device) deflate_quals(c);
    };
    const uint64_t nsb_q = R ? (R + superblock_reads(c) - 1) / superblock_reads(c) : 0;
    const bool deferred = nsb_q <= sb_pipeline_min();
    // pipelined stage: LEON_B200_QUALS_LATE_PIPE = per cent of the super-blocks walked before the quality kernels start, 0 = at
    // once (default).  Measured on configs[3] (batches of 8 M reads = 32 super-blocks, the GPU busy throughout): encode pass
    // 1.202 s at once, 1.182 / 1.192 / 1.197 s at 60 / 75 / 90 % -- within the noise, so the simple order stays.
    static const uint64_t late_pipe_pct = std::min<uint64_t>(100, env_u64("LEON_B200_QUALS_LATE_PIPE", 0));
    const bool quals_late = quals_late_knob && c->opts.has_qual && R && (deferred || late_pipe_pct > 0);
    const uint64_t quals_after = deferred ? nsb_q : nsb_q * late_pipe_pct / 100;      // super-blocks walked before they start
    c->quals_late_active = quals_late && deferred;
    TraceClock tc("encode_batch");
    if (!quals_late) launch_quals();
    tc.mark("quals_early");

    // ---- super-blocks of the walk / entropy stage
    c->d_apos.reserve((R + 1) * 4);
    c->d_aaddr.reserve((R + 1) * 4);
    c->d_insflag.reserve(R + 1);
    c->d_counters.reserve(64);
    CUDA_TRY(cudaMemsetAsync(c->d_apos.p, 0xff, (R + 1) * 4, st));     // -1
    CUDA_TRY(cudaMemsetAsync(c->d_aaddr.p, 0, (R + 1) * 4, st));
    CUDA_TRY(cudaMemsetAsync(c->d_insflag.p, 0, R + 1, st));
    CUDA_TRY(cudaMemsetAsync(c->d_counters.p, 0, 64, st));
    const uint64_t SB = superblock_reads(c);
    const uint64_t nsb = R ? (R + SB - 1) / SB : 0;
    c->nblocks = R ? (R + rpb - 1) / rpb : 0;
    c->dna_nbytes = 0;
    c->nsym_total = 0;
    if (c->sbs.size() < nsb) c->sbs.resize(nsb);
    c->sbs_used = nsb;
    std::vector<uint64_t> info(2 * nsb + 2, 0);
    if (nsb) {
        c->d_xchg.reserve(std::max<uint64_t>(256, nsb * 16 + 256));
        uint64_t *dinfo = c->d_xchg.as<uint64_t>() + 16;
        LB_LAUNCH(c->lc, k_sb_info, cdiv_u(nsb, 128), 128, 0, st, offsets, c->d_koff.as<uint64_t>(), R, SB, nsb, dinfo);
        CUDA_TRY(cudaMemcpyAsync(info.data(), dinfo, nsb * 16, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
    }
    for (uint64_t g = 0; g < nsb; g++) {
        SbState &sb = c->sbs[g];
        sb.lo = g * SB; sb.hi = std::min(R, sb.lo + SB);
        sb.nk = info[2 * g]; sb.nbases = info[2 * g + 1];
        sb.first_block = sb.lo / rpb;
        sb.nblocks = 0; sb.nsym = 0; sb.launched = false; sb.fetched = false;
        sb.kt.reset();
        sb.ensure();
    }
    // host payload: sized for typical reads (4-7 bytes per read measured on the BASELINE shapes); sb_finish grows it on demand
    // (pinned memory is slow to allocate: 377 MB for the worst case of configs[2] cost more than the GPU path of a CLI run)
    c->h_dna.reserve(c->nbases / 16 + 8 * R + 1024 * c->nblocks + 4096);
    c->h_bs.reserve(c->nblocks * 16 + 16);
    c->wk.kt.reset();
    c->wk2.kt.reset();
    c->sb_pub.store(0);
    c->sb_abort.store(false);
    c->sb_err.clear();
    if (c->sb_thread.joinable()) c->sb_thread.join();
    c->sb_walks_done.store(nsb == 0);
    c->sb_walked.store(0);
    c->sb_failed.store(false);
    if (nsb) c->sb_thread = std::thread(stage_b_worker<K>, c);
    tc.mark("prepare");

    do_anchors<K>(c);
    tc.mark("anchors");
    int32_t *apos = c->d_apos.as<int32_t>();
    uint32_t *ovf = c->d_counters.as<uint32_t>() + 2;
    CUDA_TRY(cudaEventRecord(c->ev[8], st));
    {
        unsigned long long *cnt = (unsigned long long *)(c->d_counters.as<uint8_t>() + 32);
        if (R) LB_LAUNCH(c->lc, k_count_noanchor, (unsigned)std::min<uint64_t>((R + 255) / 256, 1024), 256, 0, st, apos, R, cnt);
        CUDA_TRY(cudaMemsetAsync(cnt, 0, 8, st));
        if (R) LB_LAUNCH(c->lc, k_count_noanchor, (unsigned)std::min<uint64_t>((R + 255) / 256, 1024), 256, 0, st, apos, R, cnt);
        c->enc_noanchor += fetch_scalar(c, cnt);
        c->reads_without_anchor = c->enc_noanchor;
    }
    tc.mark("noanchor");
    if (quals_late) {
        // the last walk has finished (or the stage has failed): the quality kernels run beside the tail of range coder chains
        while (!c->sb_walks_done.load(std::memory_order_acquire) && c->sb_walked.load(std::memory_order_acquire) < quals_after)
            std::this_thread::sleep_for(std::chrono::microseconds(20));
        tc.mark("walks");
        if (!c->sb_failed.load()) launch_quals();
        tc.mark("quals_late");
    }
    // join the walk / entropy stage: every block is coded and on the host
    if (c->sb_thread.joinable()) c->sb_thread.join();
    tc.mark("join");
    if (!c->sb_err.empty()) throw LeonError(c->sb_err_code, "walk / entropy stage: " + c->sb_err);
    for (uint64_t g = 0; g < nsb; g++) c->nsym_total += c->sbs[g].nsym;
    CUDA_TRY(cudaEventRecord(c->ev[9], st));
    if (fetch_scalar(c, ovf)) throw LeonError(LEON_E_OVERFLOW, "range coder output exceeded its capacity bound");
    tc.mark("ovf");
    if (c->qz_done) CUDA_TRY(cudaEventSynchronize(c->ev_qz));   // the run ends when the quality streams are written too
    else if (c->opts.has_qual && R) CUDA_TRY(cudaEventSynchronize(c->ev_q[1]));
    tc.mark("quals_wait");
    c->enc_blocks_done += c->nblocks;
}

// End of the encoding pass: the anchor dictionary stream (one coder over all anchors, flushed once; C/Leon.cpp:1034-1042,
// 1245-1260).  By default the host thread started by encode_begin has been coding it all along; dict_on_device runs it
// as one more stream of k_range_encode.
template <typename K> void encode_end(leon_ctx *c)
{
    using namespace lbent;
    cudaStream_t st = c->st;
    const int k = c->opts.kmer_size;
    uint32_t *ovf = c->d_counters.as<uint32_t>() + 2;
    if (c->comm.rank == 0 && !c->opts.dict_on_device) c->dict_fin.store(true, std::memory_order_release);
    // anchor dictionary stream: one coder over all anchors, flushed once (C/Leon.cpp:1034-1042,1245-1260).  By default
    // the host thread started by do_anchors is already coding it; dict_on_device runs it as one more stream of
    // k_range_encode on the second stream next to everything below.
    if (c->comm.rank == 0 && c->opts.dict_on_device) {
        cudaStream_t s2 = c->st2;
        uint64_t A = c->dict_count;
        uint64_t nsym = A * (uint64_t)k;
        c->d_dict_rec.reserve((nsym + 1) * sizeof(uint4));
        c->d_dict_sh.reserve(nsym + 1);
        c->d_dict_desc.reserve(sizeof(StreamDesc));
        c->d_dict_out.reserve(2 * nsym + 1024);
        c->d_dict_size.reserve(16);
        CUDA_TRY(cudaEventRecord(c->ev[13], st));
        CUDA_TRY(cudaStreamWaitEvent(s2, c->ev[13], 0));
        if (A)
            lbscan::exclusive_scan<lbscan::U32x4>(A, DictCountIn<K>{c->d_anchors.as<K>(), k},
                                                  DictRecOut<K>{c->d_anchors.as<K>(), k, c->d_dict_rec.as<uint4>(), c->d_dict_sh.as<uint8_t>(),
                                                                lbscan::U32x4{0, 0, 0, 0}},
                                                  c->d_scan2, s2, c->lc);
        StreamDesc d{0, nsym, 0, 2 * nsym + 1024};
        c->h_desc.reserve(sizeof d);
        memcpy(c->h_desc.p, &d, sizeof d);
        CUDA_TRY(cudaMemcpyAsync(c->d_dict_desc.p, c->h_desc.p, sizeof d, cudaMemcpyHostToDevice, s2));
        LB_TIMED(c->kt, "k_range_encode_dict", s2, LB_LAUNCH(c->lc, k_range_encode, 1, 32, 0, s2, c->d_dict_rec.as<uint4>(), c->d_dict_sh.as<uint8_t>(),
                  c->d_dict_desc.as<StreamDesc>(), c->d_dict_out.as<uint8_t>(), c->d_dict_size.as<uint64_t>(), 1, ovf + 1));
        CUDA_TRY(cudaEventRecord(c->ev[14], s2));
    }
    // join the dictionary coder
    if (c->comm.rank != 0) {
        c->dict_nbytes = 0;
    } else if (c->opts.dict_on_device) {
        CUDA_TRY(cudaEventSynchronize(c->ev[14]));
        c->dict_nbytes = fetch_scalar(c, c->d_dict_size.as<uint64_t>());
        if (fetch_scalar(c, ovf + 1)) throw LeonError(LEON_E_OVERFLOW, "range coder output of the anchor dictionary exceeded its capacity bound");
    } else {
        CUDA_TRY(cudaEventRecord(c->ev[15], st));
        c->dict_thread.join();
        CUDA_TRY(cudaEventRecord(c->ev[14], st));      // ev[15]..ev[14]: how long the stream waited for the host coder
        if (!c->dict_thread_err.empty()) throw LeonError(LEON_E_CUDA, "anchor dictionary coder: " + c->dict_thread_err);
        c->dict_nbytes = c->dict_host_out.size();
    }
    CUDA_TRY(cudaEventRecord(c->ev[10], st));
    c->enc_open = false;
    c->phase = PH_ENCODED;
}

template <typename K> void do_encode_body(leon_ctx *c)
{
    uint64_t R_total = c->nreads;
    if (c->comm.n > 1) {
        c->d_xchg.reserve(256);
        CUDA_TRY(cudaMemcpyAsync(c->d_xchg.p, &R_total, 8, cudaMemcpyHostToDevice, c->st));
        NCCL_TRY(nccl_api().AllReduce(c->d_xchg.p, c->d_xchg.p, 1, ncclUint64, ncclSum, c->comm.comm, c->st));
        R_total = fetch_scalar(c, c->d_xchg.as<uint64_t>());
    }
    encode_begin<K>(c, R_total);
    encode_batch<K>(c);
    encode_end<K>(c);
}

// do_encode_body with the helper threads stopped on every error path
template <typename K> void do_encode(leon_ctx *c)
{
    try {
        do_encode_body<K>(c);
    } catch (...) {
        c->dict_abort.store(true);
        c->sb_abort.store(true);
        if (c->dict_thread.joinable()) c->dict_thread.join();
        if (c->sb_thread.joinable()) c->sb_thread.join();
        throw;
    }
}

void do_fetch(leon_ctx *c, leon_result *out, bool want_global = true)
{
    cudaStream_t st = c->st;
    const uint64_t R = c->nreads;
    CUDA_TRY(cudaEventRecord(c->ev[11], st));
    if (want_global) {
        c->h_bloom.reserve(c->bloom.nbytes + 8);
        CUDA_TRY(cudaMemcpyAsync(c->h_bloom.p, c->bloom.words, c->bloom.nbytes, cudaMemcpyDeviceToHost, st));
        c->h_dict.reserve(c->dict_nbytes + 8);
        if (c->comm.rank != 0) { }
        else if (c->opts.dict_on_device) CUDA_TRY(cudaMemcpyAsync(c->h_dict.p, c->d_dict_out.p, c->dict_nbytes, cudaMemcpyDeviceToHost, st));
        else memcpy(c->h_dict.p, c->dict_host_out.data(), c->dict_nbytes);
    }
    // dna / blocksizes were copied to the host by the walk / entropy stage as the blocks finished
    c->h_apos.reserve((R + 1) * 4);
    c->h_aaddr.reserve((R + 1) * 4);
    const bool want_arrays = !c->opts.skip_read_arrays;
    if (R && want_arrays) {
        CUDA_TRY(cudaMemcpyAsync(c->h_apos.p, c->d_apos.p, R * 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(c->h_aaddr.p, c->d_aaddr.p, R * 4, cudaMemcpyDeviceToHost, st));
    }
    uint64_t qbytes = 0;
    c->h_qsizes.reserve(c->nblocks * 8 + 8);
    const bool want_text = !(c->opts.skip_qual_text && c->opts.qual_zlib_on_device);
    if (c->opts.has_qual && R && want_text) {
        qbytes = c->nbases + R;
        c->h_qual.reserve(qbytes + 8);
        CUDA_TRY(cudaMemcpyAsync(c->h_qual.p, c->d_qout.p, qbytes, cudaMemcpyDeviceToHost, st));
    }
    uint64_t zbytes = 0;
    if (c->opts.has_qual && R && c->opts.qual_zlib_on_device && c->qz_done) {
        CUDA_TRY(cudaEventSynchronize(c->ev_qz));               // layout table is on the host now
        const uint64_t *boff = c->h_qz_boff.as<uint64_t>();
        zbytes = boff[c->nblocks];
        c->h_qz_out.reserve(zbytes + 8);
        c->h_qz_sizes.reserve(c->nblocks * 8 + 8);
        for (uint64_t b = 0; b < c->nblocks; b++) c->h_qz_sizes.as<uint64_t>()[b] = boff[b + 1] - boff[b];
        CUDA_TRY(cudaMemcpyAsync(c->h_qz_out.p, c->d_qz_out.p, zbytes, cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(cudaStreamSynchronize(st));
    memset(out, 0, sizeof *out);
    out->count = c->cres;
    if (want_global) {
        out->bloom = c->h_bloom.as<uint8_t>(); out->bloom_nbytes = c->bloom.nbytes; out->bloom_bits = c->bloom.reduced_tai;
        out->anchors_dict = c->h_dict.as<uint8_t>(); out->anchors_dict_nbytes = c->dict_nbytes; out->anchor_count = c->dict_count;
    }
    out->nblocks = c->nblocks; out->blocksizes = c->h_bs.as<uint64_t>();
    out->dna = c->h_dna.as<uint8_t>(); out->dna_nbytes = c->dna_nbytes;
    out->qual_text = qbytes ? c->h_qual.as<uint8_t>() : nullptr; out->qual_nbytes = qbytes; out->qual_block_sizes = c->h_qsizes.as<uint64_t>();
    if (zbytes) { out->qual_z = c->h_qz_out.as<uint8_t>(); out->qual_z_nbytes = zbytes; out->qual_z_block_sizes = c->h_qz_sizes.as<uint64_t>(); }
    out->read_count = R; out->total_dna_size = c->nbases;
    out->min_read_size = c->min_len; out->max_read_size = c->max_len;
    out->anchor_pos = want_arrays ? c->h_apos.as<int32_t>() : nullptr; out->anchor_addr = want_arrays ? c->h_aaddr.as<uint32_t>() : nullptr;
    out->reads_without_anchor = c->reads_without_anchor;
    if (c->opts.has_qual && !c->qsizes.empty()) memcpy(c->h_qsizes.p, c->qsizes.data(), c->qsizes.size() * 8);
    CUDA_TRY(cudaEventRecord(c->ev[12], st));
    CUDA_TRY(cudaEventSynchronize(c->ev[12]));
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------------------
static thread_local std::string g_create_error;

#define LB_API_BEGIN try {
#define LB_API_END(ctx)                                                      \
    }                                                                        \
    catch (const LeonError &e) { (ctx)->err = e.msg; return e.code; }        \
    catch (const std::exception &e) { (ctx)->err = e.what(); return LEON_E_NOMEM; } \
    return LEON_OK;

#define LB_DISPATCH(ctx, fn)                                                  \
    do {                                                                      \
        if ((ctx)->opts.kmer_size <= 31) fn<uint64_t>(ctx); else fn<u128>(ctx); \
    } while (0)

// caller's stream (leon_set_stream): order the library's main stream after it / it after the main stream
static void user_stream_in(leon_ctx *c)
{
    if (!c->user_st) return;
    CUDA_TRY(cudaEventRecord(c->ev_user_in, c->user_st));
    CUDA_TRY(cudaStreamWaitEvent(c->st, c->ev_user_in, 0));
    CUDA_TRY(cudaStreamWaitEvent(c->st3, c->ev_user_in, 0));
}
static void user_stream_out(leon_ctx *c)
{
    if (!c->user_st) return;
    CUDA_TRY(cudaEventRecord(c->ev_user_out, c->st));
    CUDA_TRY(cudaStreamWaitEvent(c->user_st, c->ev_user_out, 0));
}

static void fill_timings(leon_ctx *c)
{
    auto el = [&](int a, int b) { float ms = 0; if (cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]) != cudaSuccess) { cudaGetLastError(); ms = 0; } return ms; };
    c->tm.h2d_ms = el(0, 1);
    c->tm.count_ms = el(2, 5);
    c->tm.sort_ms = el(3, 4);
    c->tm.bloom_ms = el(5, 6);
    c->tm.anchor_ms = el(6, 7);
    c->tm.qual_ms = el(7, 8);
    c->tm.walk_ms = el(8, 9);
    c->tm.entropy_ms = el(9, 10);
    c->tm.d2h_ms = el(11, 12);
    c->tm.total_ms = el(2, 10);
    c->tm.kernel_launches = c->lc.n + c->lc2.n;
    c->tm.dict_host_ms = c->opts.dict_on_device ? 0.f : (float)c->dict_host_ms;
    c->tm.dict_copy_wait_ms = c->opts.dict_on_device ? 0.f : (float)c->dict_copy_wait_ms;
    c->tm.dict_wait_ms = c->opts.dict_on_device ? 0.f : el(15, 14);
    c->kt.resolve();
    c->wk.kt.resolve();
    c->wk2.kt.resolve();
    for (uint64_t g = 0; g < c->sbs_used; g++) c->sbs[g].kt.resolve();
    c->tm.kernel_launches += c->wk.lc.n + c->wk2.lc.n;
    c->tm.nsym_total = c->nsym_total;
    c->tm.sort_pass_ms = 0;
    for (auto &sl : c->kt.slots)
        if (!strcmp(sl.name, "k_rs_scatter") && sl.n) c->tm.sort_pass_ms = (float)(sl.ms / sl.n);
}

extern "C" {

int leon_b200_abi_version(void) { return LEON_B200_ABI_VERSION; }

int leon_create(const leon_opts *opts, leon_ctx **out)
{
    if (!opts || !out) { g_create_error = "leon_create: null argument"; return LEON_E_ARG; }
    *out = nullptr;
    if (opts->kmer_size < 2 || opts->kmer_size > 63) { g_create_error = "leon_create: kmer_size must be in [2,63]"; return LEON_E_ARG; }
    if (opts->reads_per_block <= 0) { g_create_error = "leon_create: reads_per_block must be positive"; return LEON_E_ARG; }
    if (opts->skip_qual_text && !opts->qual_zlib_on_device) { g_create_error = "leon_create: skip_qual_text needs qual_zlib_on_device"; return LEON_E_ARG; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("leon_create: no CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback";
        return LEON_E_CUDA;
    }
    if (opts->device < 0 || opts->device >= ndev) { g_create_error = "leon_create: bad device ordinal"; return LEON_E_ARG; }
    leon_ctx *c = new (std::nothrow) leon_ctx();
    if (!c) { g_create_error = "leon_create: out of memory"; return LEON_E_NOMEM; }
    c->opts = *opts;
    memset(&c->tm, 0, sizeof c->tm);
    memset(&c->cres, 0, sizeof c->cres);
    memset(&c->bloom, 0, sizeof c->bloom);
    try {
        CUDA_TRY(cudaSetDevice(opts->device));
        {   // Priorities (lower number = more urgent).  The anchor resolution on the main stream is a train of small,
            // latency-critical kernels: it must not queue behind the bulk kernels of the walk / entropy stage, which in turn
            // go before the quality stream.  The second stream carries the small record kernels and copies of the dictionary
            // thread: most urgent, so that they slot in between the thread blocks of the big kernels.
            int lo_pri = 0, hi_pri = 0;
            CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
            const int mid_pri = hi_pri < lo_pri - 1 ? lo_pri - 1 : lo_pri;
            CUDA_TRY(cudaStreamCreateWithPriority(&c->st, cudaStreamNonBlocking, hi_pri < lo_pri - 2 ? lo_pri - 2 : hi_pri));
            CUDA_TRY(cudaStreamCreateWithPriority(&c->st2, cudaStreamNonBlocking, hi_pri));
            CUDA_TRY(cudaStreamCreateWithPriority(&c->st3, cudaStreamNonBlocking, mid_pri));   // quality kernels: issued first, so they run first and are out of the way of the last chains
            CUDA_TRY(cudaStreamCreateWithPriority(&c->wk.st, cudaStreamNonBlocking, mid_pri));
            CUDA_TRY(cudaStreamCreateWithPriority(&c->wk2.st, cudaStreamNonBlocking, mid_pri));
        }
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_user_in, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_user_out, cudaEventDisableTiming));
        for (int i = 0; i < 2; i++) CUDA_TRY(cudaEventCreateWithFlags(&c->ev_q[i], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_qz, cudaEventDisableTiming));
        for (int i = 0; i < 16; i++) CUDA_TRY(cudaEventCreate(&c->ev[i]));
        for (int i = 0; i < 16; i++) CUDA_TRY(cudaEventRecord(c->ev[i], c->st));
        c->h_small.reserve(8192);      // staging slots at fixed offsets (do_count): never reallocated while a copy is in flight
    } catch (const LeonError &er) {
        g_create_error = er.msg;
        delete c;
        return er.code;
    }
    *out = c;
    return LEON_OK;
}

void leon_destroy(leon_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->opts.device);
    c->dict_abort.store(true);
    c->sb_abort.store(true);
    if (c->dict_thread.joinable()) c->dict_thread.join();
    if (c->sb_thread.joinable()) c->sb_thread.join();
    if (c->st) cudaStreamSynchronize(c->st);
    if (c->wk.st) { cudaStreamSynchronize(c->wk.st); }
    if (c->wk2.st) { cudaStreamSynchronize(c->wk2.st); }
    for (auto &sb : c->sbs) sb.destroy();
    c->wk.kt.release();
    c->wk2.kt.release();
    if (c->comm.comm) nccl_api().CommDestroy(c->comm.comm);
    DevBuf *dev[] = {&c->d_retry, &c->d_pstart_alt, &c->d_tstart, &c->d_mcnt, &c->d_mcur, &c->d_bases, &c->d_offsets, &c->d_quals, &c->d_koff, &c->d_keys, &c->d_keys_alt, &c->d_tile_hist, &c->d_tile_off,
                     &c->d_scan, &c->d_bstart, &c->d_solid_cnt, &c->d_bstart_all, &c->d_keys_recv, &c->d_solid_keys, &c->d_pick,
                     &c->d_bloom_all, &c->d_xchg, &c->d_cand_gid, &c->d_histo, &c->d_stats, &c->d_tbl, &c->d_bloom, &c->d_apos, &c->d_aaddr, &c->d_insflag, &c->d_plist,
                     &c->d_cand_pos, &c->d_cand_key, &c->d_cand_slot, &c->d_state, &c->d_ct_proxy, &c->d_ct_keys, &c->d_ct_ins,
                     &c->d_ct_fu, &c->d_dict_keys[0], &c->d_dict_keys[1], &c->d_dict_addr[0], &c->d_dict_addr[1], &c->d_dict_time[0],
                     &c->d_dict_time[1], &c->d_dict_tag[0], &c->d_dict_tag[1], &c->d_anchors, &c->d_counters, &c->d_bif, &c->d_err, &c->d_recs, &c->d_symoff, &c->d_srec, &c->d_srec_alt, &c->d_segstart,
                     &c->d_chunkoff, &c->d_chist, &c->d_rec, &c->d_sh, &c->d_scan2, &c->d_descs, &c->d_outall, &c->d_outsize, &c->d_dnaoff, &c->d_dna, &c->d_bs, &c->d_qout,
                     &c->d_dict_rec, &c->d_dict_sh, &c->d_dict_desc, &c->d_dict_out, &c->d_dict_size,
                     &c->d_dict_rec2[0], &c->d_dict_rec2[1], &c->d_dict_sh2[0], &c->d_dict_sh2[1]};
    for (DevBuf *b : dev) b->release();
    PinBuf *pin[] = {&c->h_histo, &c->h_small, &c->h_bloom, &c->h_dna, &c->h_bs, &c->h_dict, &c->h_qual, &c->h_qsizes, &c->h_apos, &c->h_aaddr, &c->h_desc, &c->h_dstage_rec[0], &c->h_dstage_rec[1],
                     &c->h_dstage_sh[0], &c->h_dstage_sh[1], &c->h_dict_tot};
    for (PinBuf *b : pin) b->release();
    for (int i = 0; i < 16; i++) cudaEventDestroy(c->ev[i]);
    c->kt.release();
    for (int i = 0; i < 2; i++) if (c->ev_dstage[i]) cudaEventDestroy(c->ev_dstage[i]);
    if (c->ev_dtot) cudaEventDestroy(c->ev_dtot);
    if (c->st) cudaStreamDestroy(c->st);
    if (c->wk.st) cudaStreamDestroy(c->wk.st);
    if (c->wk2.st) cudaStreamDestroy(c->wk2.st);
    if (c->st_copy) { cudaStreamSynchronize(c->st_copy); cudaStreamDestroy(c->st_copy); }
    if (c->ev_copy) cudaEventDestroy(c->ev_copy);
    if (c->ev_pf_guard) cudaEventDestroy(c->ev_pf_guard);
    for (int i = 0; i < 10; i++) if (c->ev_up[i]) cudaEventDestroy(c->ev_up[i]);
    if (c->ev_user_in) cudaEventDestroy(c->ev_user_in);
    if (c->ev_user_out) cudaEventDestroy(c->ev_user_out);
    if (c->st2) { cudaStreamSynchronize(c->st2); cudaStreamDestroy(c->st2); }
    if (c->st3) { cudaStreamSynchronize(c->st3); cudaStreamDestroy(c->st3); }
    for (int i = 0; i < 2; i++) if (c->ev_q[i]) cudaEventDestroy(c->ev_q[i]);
    if (c->ev_qz) cudaEventDestroy(c->ev_qz);
    delete c;
}

const char *leon_last_error(const leon_ctx *c) { return c ? c->err.c_str() : g_create_error.c_str(); }
void leon_internal_set_error(const char *msg) { g_create_error = msg ? msg : ""; }   // used by container.cpp

void *leon_alloc_pinned(size_t nbytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, nbytes ? nbytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void leon_free_pinned(void *p) { if (p) cudaFreeHost(p); }

int leon_upload(leon_ctx *c, const leon_reads *reads)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    do_upload(c, reads);
    CUDA_TRY(cudaStreamSynchronize(c->st));          // the caller's buffers are free again when this call returns
    CUDA_TRY(cudaStreamSynchronize(c->st3));
    user_stream_out(c);
    LB_API_END(c)
}

int leon_count_resident(leon_ctx *c, leon_count_result *out)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_UPLOADED) throw LeonError(LEON_E_STATE, "leon_count_resident: no read set uploaded");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    LB_DISPATCH(c, do_count);
    if (out) *out = c->cres;
    user_stream_out(c);
    LB_API_END(c)
}

int leon_bloom_build(leon_ctx *c, const uint8_t **bits, uint64_t *nbytes, uint64_t *reduced_bits)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_COUNTED) throw LeonError(LEON_E_STATE, "leon_bloom_build: leon_count has not run");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    LB_DISPATCH(c, do_bloom);
    if (bits) {
        c->h_bloom.reserve(c->bloom.nbytes + 8);
        CUDA_TRY(cudaMemcpyAsync(c->h_bloom.p, c->bloom.words, c->bloom.nbytes, cudaMemcpyDeviceToHost, c->st));
        CUDA_TRY(cudaStreamSynchronize(c->st));
        *bits = c->h_bloom.as<uint8_t>();
    }
    if (nbytes) *nbytes = c->bloom.nbytes;
    if (reduced_bits) *reduced_bits = c->bloom.reduced_tai;
    user_stream_out(c);
    LB_API_END(c)
}


}  // extern "C"

// Host-side facts of a batch the stages need: length range and the size of every block's quality text
static void batch_host_scan(const leon_reads *rd, uint64_t rpb, bool with_quals, int32_t &mn, int32_t &mx, std::vector<uint64_t> &qsizes)
{
    mn = 0x7fffffff; mx = 0;
    for (uint64_t i = 0; i < rd->nreads; i++) {
        uint64_t L = rd->offsets[i + 1] - rd->offsets[i];
        if (L >= 0x7fffffffull) throw LeonError(LEON_E_ARG, "read longer than 2^31-1");
        if ((int32_t)L < mn) mn = (int32_t)L;
        if ((int32_t)L > mx) mx = (int32_t)L;
    }
    qsizes.clear();
    if (with_quals)
        for (uint64_t r0 = 0; r0 < rd->nreads; r0 += rpb) {
            uint64_t r1 = std::min(rd->nreads, r0 + rpb);
            qsizes.push_back(rd->offsets[r1] - rd->offsets[r0] + (r1 - r0));
        }
}

static void batch_check(const leon_reads *rd, bool with_quals)
{
    if (!rd || !rd->offsets || (rd->nreads && !rd->bases)) throw LeonError(LEON_E_ARG, "null read batch");
    if (rd->offsets[0] != 0) throw LeonError(LEON_E_ARG, "a batch's offsets must start at 0");
    if (with_quals && !rd->quals) throw LeonError(LEON_E_ARG, "has_qual set but the batch has no qualities");
}

// leon_stream_prefetch: the NEXT batch on its way into the spare read buffers while the current one is being processed
static bool same_batch(const leon_reads &a, const leon_reads &b, bool a_quals, bool want_quals)
{
    return a.bases == b.bases && a.offsets == b.offsets && a.nreads == b.nreads && (!want_quals || (a_quals && a.quals == b.quals));
}

// copies of batch `rd` into the spare buffers on the copy stream; the spare buffers held the batch before the current one, and
// whatever the main stream still does with them comes first
static void prefetch_start(leon_ctx *c, const leon_reads *rd, bool with_quals)
{
    if (!c->st_copy) CUDA_TRY(cudaStreamCreateWithFlags(&c->st_copy, cudaStreamNonBlocking));
    if (!c->ev_copy) CUDA_TRY(cudaEventCreateWithFlags(&c->ev_copy, cudaEventDisableTiming));
    if (!c->ev_pf_guard) CUDA_TRY(cudaEventCreateWithFlags(&c->ev_pf_guard, cudaEventDisableTiming));
    const uint64_t nbases = rd->offsets[rd->nreads];
    c->d_bases_next.reserve(nbases + 128);
    c->d_offsets_next.reserve((rd->nreads + 1) * 8);
    if (with_quals) c->d_quals_next.reserve(nbases + 128);
    CUDA_TRY(cudaEventRecord(c->ev_pf_guard, c->st));
    CUDA_TRY(cudaStreamWaitEvent(c->st_copy, c->ev_pf_guard, 0));
    // the offsets FIRST: callers tend to build them per batch in pageable memory, and a copy from pageable memory holds the
    // host until everything queued before it on the stream is done (measured: 27 ms per batch behind 1.2 GB of bases)
    CUDA_TRY(cudaMemcpyAsync(c->d_offsets_next.p, rd->offsets, (rd->nreads + 1) * 8, cudaMemcpyHostToDevice, c->st_copy));
    CUDA_TRY(cudaMemcpyAsync(c->d_bases_next.p, rd->bases, nbases, cudaMemcpyHostToDevice, c->st_copy));
    if (with_quals) CUDA_TRY(cudaMemcpyAsync(c->d_quals_next.p, rd->quals, nbases, cudaMemcpyHostToDevice, c->st_copy));
    CUDA_TRY(cudaEventRecord(c->ev_copy, c->st_copy));
}

// leon_stream_prefetch.  Callers announce batch i + 1 just before they submit batch i (the submission is synchronous), so two
// announcements can be open at a time: the prefetched batch (copy started) and one queued behind it, whose copy starts the
// moment the prefetched one is swapped in by its submission -- and then runs beside that whole call.
static void stream_prefetch(leon_ctx *c, const leon_reads *rd, bool with_quals)
{
    TraceClock tc("leon_stream_prefetch");
    batch_check(rd, with_quals);
    tc.mark("check");
    if (c->pf_valid && same_batch(c->pf_reads, *rd, c->pf_quals, with_quals)) return;      // already on its way
    if (c->pf_valid) {
        // the prefetched batch has not been submitted yet: this one waits behind it (a later announcement replaces it)
        batch_host_scan(rd, (uint64_t)c->opts.reads_per_block, with_quals, c->pfq_min, c->pfq_max, c->pfq_qsizes);
        c->pfq_reads = *rd;
        c->pfq_quals = with_quals;
        c->pfq_valid = true;
        tc.mark("queued_host_scan");
        return;
    }
    prefetch_start(c, rd, with_quals);
    tc.mark("enqueue");
    batch_host_scan(rd, (uint64_t)c->opts.reads_per_block, with_quals, c->pf_min, c->pf_max, c->pf_qsizes);   // while the copies run
    tc.mark("host_scan");
    c->pf_reads = *rd;
    c->pf_quals = with_quals;
    c->pf_valid = true;
}

// One batch of the streaming path into the read buffers (or the prefetched copy of it swapped in); the caller's buffers
// are free again on return.
static void stream_upload(leon_ctx *c, const leon_reads *rd, bool with_quals, bool first_of_set)
{
    batch_check(rd, with_quals);
    c->nreads = rd->nreads;
    c->nbases = rd->offsets[rd->nreads];
    int32_t bmn, bmx;
    const bool hit = c->pf_valid && same_batch(c->pf_reads, *rd, c->pf_quals, with_quals);
    if (hit) {
        CUDA_TRY(cudaEventSynchronize(c->ev_copy));
        std::swap(c->d_bases.p, c->d_bases_next.p); std::swap(c->d_bases.cap, c->d_bases_next.cap);
        std::swap(c->d_offsets.p, c->d_offsets_next.p); std::swap(c->d_offsets.cap, c->d_offsets_next.cap);
        if (with_quals) { std::swap(c->d_quals.p, c->d_quals_next.p); std::swap(c->d_quals.cap, c->d_quals_next.cap); }
        bmn = c->pf_min; bmx = c->pf_max;
        c->qsizes.swap(c->pf_qsizes);
        c->pf_valid = false;
        if (c->pfq_valid) {        // the batch announced behind this one: its copy runs beside the call that starts now
            c->pfq_valid = false;
            prefetch_start(c, &c->pfq_reads, c->pfq_quals);
            c->pf_reads = c->pfq_reads; c->pf_quals = c->pfq_quals;
            c->pf_min = c->pfq_min; c->pf_max = c->pfq_max;
            c->pf_qsizes.swap(c->pfq_qsizes);
            c->pf_valid = true;
        }
    } else {
        c->d_bases.reserve(c->nbases + 128);
        c->d_offsets.reserve((c->nreads + 1) * 8);
        if (with_quals) c->d_quals.reserve(c->nbases + 128);
        CUDA_TRY(cudaMemcpyAsync(c->d_offsets.p, rd->offsets, (c->nreads + 1) * 8, cudaMemcpyHostToDevice, c->st));     // first: see prefetch_start
        CUDA_TRY(cudaMemcpyAsync(c->d_bases.p, rd->bases, c->nbases, cudaMemcpyHostToDevice, c->st));
        if (with_quals) CUDA_TRY(cudaMemcpyAsync(c->d_quals.p, rd->quals, c->nbases, cudaMemcpyHostToDevice, c->st));
        batch_host_scan(rd, (uint64_t)c->opts.reads_per_block, with_quals, bmn, bmx, c->qsizes);
        CUDA_TRY(cudaStreamSynchronize(c->st));
    }
    c->min_len = first_of_set ? bmn : std::min(c->min_len, bmn);        // running over the set
    c->max_len = first_of_set ? bmx : std::max(c->max_len, bmx);
}

static double count_budget_frac()
{
    static const uint64_t pct = std::min<uint64_t>(90, std::max<uint64_t>(5, env_u64("LEON_B200_COUNT_BUDGET", 50)));
    return (double)pct / 100.0;
}

template <typename K> static void stream_count_begin(leon_ctx *c, const leon_stream_plan *plan, uint32_t *npasses_out)
{
    const uint64_t kb = sizeof(K);
    uint32_t np = plan->count_passes;
    // upper bound on the k-mer slots this rank extracts: one per base
    const uint64_t nk_bound = plan->total_bases + 1;
    if (np == 0) {
        size_t free_b = 0, total_b = 0;
        CUDA_TRY(cudaMemGetInfo(&free_b, &total_b));
        // key buffer + its ping-pong twin (+ the receive buffer of the exchange with several ranks) within half of what is free
        // (LEON_B200_COUNT_BUDGET, per cent): the buffers stay allocated (grow-only, like every buffer of a context: freeing and
        // re-allocating tens of GB per run cost more than a second counting pass), so they must leave room for the encoding
        // stage, whose buffers grow with the caller's batch (5 KB per read of 150 bp measured: 43 GB for batches of 8 M reads).
        // Measured on configs[3], 1 GPU: 70 % with release 1.65 s, 30 % = two passes 0.82 s, one pass with the buffers kept 0.48 s.
        const double per_key = (double)kb * (c->comm.n > 1 ? 3.2 : 2.1);
        const double budget = count_budget_frac() * (double)(free_b + c->d_keys.cap + c->d_keys_alt.cap + c->d_keys_recv.cap + c->d_keys_batch.cap);
        np = (uint32_t)std::max(1.0, std::ceil((double)nk_bound * per_key / budget));
    }
    uint64_t cap = nk_bound;
    if (np > 1) cap = (uint64_t)((double)nk_bound / np * 1.06) + (1u << 20);     // hash split: a few per cent of imbalance at most
    count_begin<K>(c, np, cap);
    CUDA_TRY(cudaEventRecord(c->ev[2], c->st));
    if (c->pf_valid) CUDA_TRY(cudaEventSynchronize(c->ev_copy));      // announcements of an earlier read set are void
    c->pf_valid = c->pfq_valid = false;
    c->stream_reads_seen = 0;
    c->stream_counting = true;
    if (npasses_out) *npasses_out = np;
}

template <typename K> static void stream_count_batch(leon_ctx *c, const leon_reads *batch, int last)
{
    if (!c->stream_counting) throw LeonError(LEON_E_STATE, "leon_count: leon_count_begin has not run");
    if (c->count_pass >= c->count_npasses) throw LeonError(LEON_E_STATE, "leon_count: all counting passes are closed");
    stream_upload(c, batch, false, c->count_pass == 0 && c->stream_reads_seen == 0);
    c->stream_reads_seen += batch->nreads;
    count_koff(c);
    count_add<K>(c);
    if (last) {
        count_close<K>(c);
        c->stream_reads_seen = 0;
    }
}

template <typename K> static void stream_encode_block(leon_ctx *c, uint64_t first_block, const leon_reads *batch, leon_block_out *out)
{
    if (c->phase < PH_BLOOM) throw LeonError(LEON_E_STATE, "leon_encode_block: leon_bloom_build has not run");
    if (!c->enc_open) {
        uint64_t R_total = c->count_reads_total;
        if (c->comm.n > 1) {
            c->d_xchg.reserve(256);
            CUDA_TRY(cudaMemcpyAsync(c->d_xchg.p, &R_total, 8, cudaMemcpyHostToDevice, c->st));
            NCCL_TRY(nccl_api().AllReduce(c->d_xchg.p, c->d_xchg.p, 1, ncclUint64, ncclSum, c->comm.comm, c->st));
            R_total = fetch_scalar(c, c->d_xchg.as<uint64_t>());
        }
        encode_begin<K>(c, R_total);
        c->enc_total_bases = 0;
        c->enc_partial_seen = false;
    }
    if (first_block != c->enc_blocks_done) throw LeonError(LEON_E_ARG, "leon_encode_block: blocks must be submitted in order (first_block = blocks coded so far)");
    if (c->enc_partial_seen && batch->nreads) throw LeonError(LEON_E_ARG, "leon_encode_block: only the last batch may hold a partial block");
    if (batch->nreads % (uint64_t)c->opts.reads_per_block) c->enc_partial_seen = true;
    if (c->comm.n > 1 && batch->nreads % anchor_batch_reads(c)) c->enc_partial_seen = true;
    const int32_t mn = c->min_len, mx = c->max_len;
    TraceClock tc("leon_encode_block");
    stream_upload(c, batch, c->opts.has_qual != 0, false);
    tc.mark("upload");
    c->min_len = std::min(mn, c->min_len); c->max_len = std::max(mx, c->max_len);     // of the whole set (counting pass)
    count_koff(c);
    tc.mark("koff");
    encode_batch<K>(c);
    tc.mark("encode_batch");
    c->enc_total_bases += c->nbases;
    leon_result r;
    do_fetch(c, &r, false);
    tc.mark("fetch");
    memset(out, 0, sizeof *out);
    out->first_block = first_block; out->nblocks = r.nblocks; out->read_count = r.read_count;
    out->blocksizes = r.blocksizes; out->dna = r.dna; out->dna_nbytes = r.dna_nbytes;
    out->qual_text = r.qual_text; out->qual_nbytes = r.qual_nbytes; out->qual_block_sizes = r.qual_block_sizes;
    out->qual_z = r.qual_z; out->qual_z_nbytes = r.qual_z_nbytes; out->qual_z_block_sizes = r.qual_z_block_sizes;
    out->anchor_pos = r.anchor_pos; out->anchor_addr = r.anchor_addr;
    c->last_block_out = *out;
}

template <typename K> static void stream_encode_finish(leon_ctx *c, leon_result *out)
{
    if (!c->enc_open) {      // no batch at all: an empty set
        encode_begin<K>(c, 0);
        c->enc_total_bases = 0;
        CUDA_TRY(cudaMemsetAsync(c->d_counters.p, 0, 64, c->st));
    }
    encode_end<K>(c);
    // whole-set fields only
    const uint64_t nreads = c->nreads, nbases = c->nbases, nblocks = c->nblocks;
    c->nreads = 0; c->nblocks = 0;
    do_fetch(c, out, true);
    c->nreads = nreads; c->nbases = nbases; c->nblocks = nblocks;
    out->nblocks = c->enc_blocks_done;
    out->blocksizes = nullptr; out->dna = nullptr; out->dna_nbytes = 0;
    out->qual_text = nullptr; out->qual_nbytes = 0; out->qual_z = nullptr; out->qual_z_nbytes = 0;
    out->qual_block_sizes = nullptr; out->qual_z_block_sizes = nullptr;
    out->anchor_pos = nullptr; out->anchor_addr = nullptr;
    out->read_count = c->enc_reads_done; out->total_dna_size = c->enc_total_bases;
    out->reads_without_anchor = c->enc_noanchor;
    out->min_read_size = c->min_len; out->max_read_size = c->max_len;
}

extern "C" {

int leon_count_begin(leon_ctx *c, const leon_stream_plan *plan, uint32_t *npasses)
{
    if (!c || !plan) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    if (c->opts.kmer_size <= 31) stream_count_begin<uint64_t>(c, plan, npasses); else stream_count_begin<u128>(c, plan, npasses);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_count(leon_ctx *c, const leon_reads *batch, int last)
{
    if (!c || !batch) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    if (c->opts.kmer_size <= 31) stream_count_batch<uint64_t>(c, batch, last); else stream_count_batch<u128>(c, batch, last);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_stream_prefetch(leon_ctx *c, const leon_reads *next_batch, int with_quals)
{
    if (!c || !next_batch) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    stream_prefetch(c, next_batch, with_quals != 0 && c->opts.has_qual);
    LB_API_END(c)
}

int leon_count_finish(leon_ctx *c, leon_count_result *out)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    if (!c->stream_counting) throw LeonError(LEON_E_STATE, "leon_count_finish: leon_count_begin has not run");
    if (c->count_pass < c->count_npasses) throw LeonError(LEON_E_STATE, "leon_count_finish: not every counting pass has been closed (last = 1)");
    user_stream_in(c);
    count_finish(c);
    c->stream_counting = false;
    if (out) *out = c->cres;
    user_stream_out(c);
    LB_API_END(c)
}

int leon_encode_block(leon_ctx *c, uint64_t first_block, const leon_reads *batch, leon_block_out *out)
{
    if (!c || !batch || !out) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    try {
        if (c->opts.kmer_size <= 31) stream_encode_block<uint64_t>(c, first_block, batch, out); else stream_encode_block<u128>(c, first_block, batch, out);
    } catch (...) {
        c->dict_abort.store(true);
        c->sb_abort.store(true);
        if (c->dict_thread.joinable()) c->dict_thread.join();
        if (c->sb_thread.joinable()) c->sb_thread.join();
        c->enc_open = false;
        throw;
    }
    fill_timings(c);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_block_wait(leon_ctx *c, uint64_t first_block, leon_block_out *out)
{
    if (!c || !out) return LEON_E_ARG;
    if (c->last_block_out.first_block != first_block || !c->last_block_out.blocksizes) { c->err = "leon_block_wait: not the last submitted block"; return LEON_E_ARG; }
    *out = c->last_block_out;
    return LEON_OK;
}

int leon_encode_finish(leon_ctx *c, leon_result *out)
{
    if (!c || !out) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    if (c->phase < PH_BLOOM) throw LeonError(LEON_E_STATE, "leon_encode_finish: leon_bloom_build has not run");
    user_stream_in(c);
    try {
        if (c->opts.kmer_size <= 31) stream_encode_finish<uint64_t>(c, out); else stream_encode_finish<u128>(c, out);
    } catch (...) {
        c->dict_abort.store(true);
        if (c->dict_thread.joinable()) c->dict_thread.join();
        c->enc_open = false;
        throw;
    }
    memset(&c->last_block_out, 0, sizeof c->last_block_out);
    fill_timings(c);
    user_stream_out(c);
    LB_API_END(c)
}

}  // extern "C"

extern "C" {

int leon_fetch(leon_ctx *c, leon_result *out)
{
    if (!c || !out) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_ENCODED) throw LeonError(LEON_E_STATE, "leon_fetch: nothing encoded yet");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    do_fetch(c, out);
    fill_timings(c);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_encode(leon_ctx *c, leon_result *out)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_BLOOM) throw LeonError(LEON_E_STATE, "leon_encode: leon_bloom_build has not run");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    LB_DISPATCH(c, do_encode);
    if (out) do_fetch(c, out);
    fill_timings(c);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_run_device(leon_ctx *c)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_UPLOADED) throw LeonError(LEON_E_STATE, "leon_run_device: no read set uploaded");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    LB_DISPATCH(c, do_count);
    LB_DISPATCH(c, do_bloom);
    LB_DISPATCH(c, do_encode);
    CUDA_TRY(cudaStreamSynchronize(c->st));
    fill_timings(c);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_compress(leon_ctx *c, const leon_reads *reads, leon_result *out)
{
    if (!c || !out) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    user_stream_in(c);
    static const bool trace = getenv("LEON_B200_TRACE") != nullptr;
    auto now = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = trace ? now() : 0;
    double t1 = 0, t2 = 0, t3 = 0, t4 = 0;
    try {
        do_upload(c, reads, true);
        if (trace) t1 = now();
        LB_DISPATCH(c, do_count);
        if (trace) t2 = now();
        LB_DISPATCH(c, do_bloom);
        if (trace) t3 = now();
        LB_DISPATCH(c, do_encode);
        if (trace) t4 = now();
    } catch (...) {
        if (c->st_copy) cudaStreamSynchronize(c->st_copy);      // the caller's buffers are free again when the call returns
        cudaStreamSynchronize(c->st3);
        throw;
    }
    do_fetch(c, out);
    if (trace) fprintf(stderr, "[leon_b200] leon_compress host clock: upload %.1f ms, count %.1f, bloom %.1f, encode %.1f, fetch %.1f\n",
                       t1 - t0, t2 - t1, t3 - t2, t4 - t3, now() - t4);
    fill_timings(c);
    user_stream_out(c);
    LB_API_END(c)
}

int leon_get_timings(const leon_ctx *c, leon_timings *out)
{
    if (!c || !out) return LEON_E_ARG;
    *out = c->tm;
    return LEON_OK;
}

int leon_comm_unique_id(void *id128)
{
    if (!id128) return LEON_E_ARG;
    std::string err;
    if (!nccl_api().load(err)) { g_create_error = err; return LEON_E_CUDA; }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (nccl_api().GetUniqueId(&id) != ncclSuccess) { g_create_error = "ncclGetUniqueId failed"; return LEON_E_CUDA; }
    memcpy(id128, &id, 128);
    return LEON_OK;
}

int leon_comm_init(leon_ctx *c, int rank, int nranks, const void *id128)
{
    if (!c || !id128) return LEON_E_ARG;
    LB_API_BEGIN
    if (nranks < 1 || nranks > lbcount::MAX_RANKS || rank < 0 || rank >= nranks)
        throw LeonError(LEON_E_ARG, "leon_comm_init: 1 <= nranks <= 8 and 0 <= rank < nranks");
    if (c->comm.comm) throw LeonError(LEON_E_STATE, "leon_comm_init: communicator already set");
    if (nranks > 1 && c->opts.anchor_batch > 0 && c->opts.anchor_batch % c->opts.reads_per_block)
        throw LeonError(LEON_E_ARG, "leon_comm_init: anchor_batch must be a multiple of reads_per_block");
    std::string err;
    if (!nccl_api().load(err)) throw LeonError(LEON_E_CUDA, err);
    CUDA_TRY(cudaSetDevice(c->opts.device));
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    NCCL_TRY(nccl_api().CommInitRank(&c->comm.comm, nranks, id, rank));
    c->comm.rank = rank;
    c->comm.n = nranks;
    LB_API_END(c)
}

int leon_comm_batch_reads(const leon_ctx *c, uint64_t *reads)
{
    if (!c || !reads) return LEON_E_ARG;
    *reads = anchor_batch_reads(c);
    return LEON_OK;
}

int leon_set_stream(leon_ctx *c, void *stream)
{
    if (!c) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    // The library keeps its own streams (their priorities order the stages of the path); the caller's stream is a
    // synchronisation point: every call starts after what was enqueued on it and it waits for the call's device work.
    c->user_st = (cudaStream_t)stream;
    c->own_stream = stream == nullptr;
    LB_API_END(c)
}

int leon_get_kernel_times(const leon_ctx *c, leon_kernel_time *out, int cap, int *n)
{
    if (!c || !n || (cap > 0 && !out)) return LEON_E_ARG;
    // slots of the main thread, of the walk / entropy worker and of every super-block's range coder stream, merged by name
    std::vector<KernelTimer::Slot> all;
    auto add = [&](const KernelTimer &kt) {
        for (auto &sl : kt.slots) {
            bool found = false;
            for (auto &a : all)
                if (!strcmp(a.name, sl.name)) { a.ms += sl.ms; a.n += sl.n; found = true; break; }
            if (!found) all.push_back(sl);
        }
    };
    add(c->kt);
    add(c->wk.kt);
    add(c->wk2.kt);
    for (uint64_t g = 0; g < c->sbs_used && g < c->sbs.size(); g++) add(c->sbs[g].kt);
    int m = 0;
    for (auto &sl : all) {
        if (m < cap) {
            memset(&out[m], 0, sizeof out[m]);
            strncpy(out[m].name, sl.name, sizeof(out[m].name) - 1);
            out[m].ms_total = (float)sl.ms;
            out[m].launches = sl.n;
        }
        m++;
    }
    *n = m;
    return LEON_OK;
}

int leon_debug_bloom_query(leon_ctx *c, const uint64_t *kmers, uint64_t n, int right, uint8_t *out4, uint8_t *out1)
{
    if (!c || !kmers || !out4 || !out1) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_BLOOM) throw LeonError(LEON_E_STATE, "leon_debug_bloom_query: no Bloom filter built");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    DevBuf dk, d4, d1;
    dk.reserve(n * 16 + 16); d4.reserve(n + 16); d1.reserve(n + 16);
    try {
        CUDA_TRY(cudaMemcpyAsync(dk.p, kmers, n * 16, cudaMemcpyHostToDevice, c->st));
        if (n) {
            if (c->opts.kmer_size <= 31)
                LB_LAUNCH(c->lc, (lbbloom::k_debug_query<uint64_t>), warp_grid(n), 256, 0, c->st, dk.as<uint64_t>(), n, right, c->bloom,
                          c->d_tbl.as<uint64_t>(), d4.as<uint8_t>(), d1.as<uint8_t>());
            else
                LB_LAUNCH(c->lc, (lbbloom::k_debug_query<u128>), warp_grid(n), 256, 0, c->st, dk.as<uint64_t>(), n, right, c->bloom,
                          c->d_tbl.as<uint64_t>(), d4.as<uint8_t>(), d1.as<uint8_t>());
        }
        CUDA_TRY(cudaMemcpyAsync(out4, d4.p, n, cudaMemcpyDeviceToHost, c->st));
        CUDA_TRY(cudaMemcpyAsync(out1, d1.p, n, cudaMemcpyDeviceToHost, c->st));
        CUDA_TRY(cudaStreamSynchronize(c->st));
    } catch (...) { dk.release(); d4.release(); d1.release(); throw; }
    dk.release(); d4.release(); d1.release();
    LB_API_END(c)
}

int leon_debug_hashes(leon_ctx *c, const uint64_t *kmers, uint64_t n, uint64_t *out12)
{
    if (!c || !kmers || !out12) return LEON_E_ARG;
    LB_API_BEGIN
    if (c->phase < PH_BLOOM) throw LeonError(LEON_E_STATE, "leon_debug_hashes: no Bloom filter built");
    CUDA_TRY(cudaSetDevice(c->opts.device));
    DevBuf dk, d_o;
    dk.reserve(n * 16 + 16); d_o.reserve(n * 96 + 16);
    try {
        CUDA_TRY(cudaMemcpyAsync(dk.p, kmers, n * 16, cudaMemcpyHostToDevice, c->st));
        if (n) {
            if (c->opts.kmer_size <= 31)
                LB_LAUNCH(c->lc, (lbbloom::k_debug_hashes<uint64_t>), cdiv_u(n, 128), 128, 0, c->st, dk.as<uint64_t>(), n, c->bloom,
                          c->d_tbl.as<uint64_t>(), d_o.as<uint64_t>());
            else
                LB_LAUNCH(c->lc, (lbbloom::k_debug_hashes<u128>), cdiv_u(n, 128), 128, 0, c->st, dk.as<uint64_t>(), n, c->bloom,
                          c->d_tbl.as<uint64_t>(), d_o.as<uint64_t>());
        }
        CUDA_TRY(cudaMemcpyAsync(out12, d_o.p, n * 96, cudaMemcpyDeviceToHost, c->st));
        CUDA_TRY(cudaStreamSynchronize(c->st));
    } catch (...) { dk.release(); d_o.release(); throw; }
    dk.release(); d_o.release();
    LB_API_END(c)
}

int leon_debug_fetch(leon_ctx *c, const char *name, void *dst, uint64_t cap, uint64_t *nbytes)
{
    if (!c || !name || !nbytes) return LEON_E_ARG;
    LB_API_BEGIN
    CUDA_TRY(cudaSetDevice(c->opts.device));
    const void *src = nullptr;
    uint64_t n = 0;
    std::string s(name);
    size_t ksz = c->opts.kmer_size <= 31 ? 8 : 16;
    if (s == "solid_keys") { src = c->solid_keys; n = c->nsolid * ksz; }
    else if (s == "solid_cnt") { src = c->d_solid_cnt.p; n = c->nsolid * 4; }
    else if (s == "anchors") { src = c->d_anchors.p; n = c->dict_count * ksz; }
    else if (s == "recs") { src = c->d_recs.p; n = c->nreads * sizeof(lbwalk::ReadRec); }
    else if (s == "koff") { src = c->d_koff.p; n = (c->nreads + 1) * 8; }
    else if (s == "bif") { src = c->d_bif.p; n = c->nkmers; }
    else if (s == "err") { src = c->d_err.p; n = c->nkmers * 4; }
    else if (s == "nsym") { src = c->d_symoff.as<uint64_t>() + c->nreads; n = c->nblocks ? 8 : 0; }
    else throw LeonError(LEON_E_ARG, "leon_debug_fetch: unknown buffer name " + s);
    *nbytes = n;
    if (dst && src) {
        if (cap < n) throw LeonError(LEON_E_ARG, "leon_debug_fetch: destination too small");
        CUDA_TRY(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, c->st));
        CUDA_TRY(cudaStreamSynchronize(c->st));
    }
    LB_API_END(c)
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// Entry point that allocates nothing (no counterpart in the reference).
// Some callers cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
// leon_kmer_offsets_noalloc is the k-mer slot scan of the counting stage (count_koff above: koff[r] = number of k-mers of
// the reads before r, K/Model.hpp:725-764 gives L - k + 1 windows per read) with the scratch of the device-wide scan -- the
// tile sums -- in a file-scope device array instead of a grow-only DevBuf.  The array is shared by every call on a device
// and keeps the last call's contents; nothing in it needs clearing, because every slot a call reads (tile sums 0..nb-1 and
// the total behind them) is written by that same call first.  Calls that overlap on one GPU race on it.
// ------------------------------------------------------------------------------------------------------------
namespace {

constexpr uint64_t NOALLOC_MAX_READS = LEON_NOALLOC_MAX_READS;                             // 2^24 reads per call
constexpr uint64_t NOALLOC_MAX_TILES = (NOALLOC_MAX_READS + lbscan::TILE - 1) / lbscan::TILE;
// tile sums + the grand total: 8193 x 8 bytes = 64 KB, held by every GPU that loads the library until exit
static __device__ __align__(256) unsigned long long g_noalloc_scan_scratch[NOALLOC_MAX_TILES + 1];

constexpr int NOALLOC_MAX_DEVICES = 64;
uint64_t *g_noalloc_ptr[NOALLOC_MAX_DEVICES];       // device address of the array on every device (set-up: once per device)
std::mutex g_noalloc_mu;

}  // namespace

extern "C" int leon_kmer_offsets_noalloc(int32_t device, const uint64_t *d_offsets, uint64_t nreads, int32_t kmer_size,
                                         uint64_t *d_koff, void *stream)
{
    if (!d_offsets || !d_koff) { g_create_error = "leon_kmer_offsets_noalloc: null argument"; return LEON_E_ARG; }
    if (kmer_size < 2 || kmer_size > 63) { g_create_error = "leon_kmer_offsets_noalloc: kmer_size must be in [2,63]"; return LEON_E_ARG; }
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) { g_create_error = "leon_kmer_offsets_noalloc: bad device ordinal"; return LEON_E_ARG; }
    if (nreads > NOALLOC_MAX_READS) {
        g_create_error = "leon_kmer_offsets_noalloc: " + std::to_string(nreads) + " reads exceed the static scratch of this entry point (maximum " +
                         std::to_string(NOALLOC_MAX_READS) + " reads per call); split the set or use leon_upload + leon_count_resident";
        return LEON_E_OVERFLOW;
    }
    try {
        CUDA_TRY(cudaSetDevice(device));
        uint64_t *partials = nullptr;
        {   // In host code the array's name is a host-side stand-in, not a device address: ask the runtime for the address on
            // THIS device, once, with the device current.
            std::lock_guard<std::mutex> g(g_noalloc_mu);
            if (!g_noalloc_ptr[device]) {
                void *p = nullptr;
                CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scan_scratch));
                g_noalloc_ptr[device] = (uint64_t *)p;
            }
            partials = g_noalloc_ptr[device];
        }
        cudaStream_t st = (cudaStream_t)stream;
        uint64_t nb = (nreads + lbscan::TILE - 1) / lbscan::TILE;
        if (nb == 0) nb = 1;
        uint64_t *total = partials + nb;
        LaunchCounter lc;
        lbcount::NkIn in{d_offsets, kmer_size};
        lbcount::NkOut out{d_koff};
        LB_LAUNCH(lc, (lbscan::k_reduce<uint64_t, lbcount::NkIn>), (unsigned)nb, lbscan::THREADS, 0, st, nreads, in, partials);
        LB_LAUNCH(lc, (lbscan::k_scan_partials<uint64_t>), 1, lbscan::THREADS, 0, st, partials, nb, total);
        LB_LAUNCH(lc, (lbscan::k_apply<uint64_t, lbcount::NkIn, lbcount::NkOut>), (unsigned)nb, lbscan::THREADS, 0, st, nreads, in, partials, out);
        CUDA_TRY(cudaMemcpyAsync(d_koff + nreads, total, 8, cudaMemcpyDeviceToDevice, st));      // koff[nreads] = number of k-mer slots
    } catch (const LeonError &e) {
        g_create_error = e.msg;
        return e.code;
    }
    return LEON_OK;
}
