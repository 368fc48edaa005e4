// Decompression of the sequence stream, host side: per-block range decoding interleaved with the
// de-Bruijn walk through the Bloom filter, the mirror of csrc/walk.cuh + csrc/entropy.cuh.  Follows
// C/DnaCoder.cpp:1221-1398 (DnaDecoder setup / execute), :1400-1480 decodeAnchorRead, :1482-1605 extendAnchor,
// :1607-1622 decodeNoAnchorRead, C/RangeCoder.cpp:179-244 (RangeDecoder), C/CompressionUtils.hpp:114-137
// (decodeNumeric), C/Leon.cpp:1948-2015 (anchor dictionary), COL/Bloom.hpp:645-798 (contains4).
// A block is inherently sequential (every symbol's model depends on the symbols before it, and how many bifurcation
// symbols a read consumes is only known while walking it), so the parallelism is one host thread per block; the k-mer
// arithmetic and the Bloom hashing are the __host__ __device__ functions of kmer_math.cuh that the kernels use.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/leon_b200.h"
#include "decoded.h"
#include "kmer_math.cuh"
#include "simplehash_table.h"
#include "decode_device.cuh"

extern "C" void leon_internal_set_error(const char *msg);

namespace lbdecode {

constexpr uint64_t RC_TOP = 1ull << 56, RC_BOTTOM = 1ull << 48;
constexpr int NB_NUMERIC = 20;

// Adaptive order-0 model (C/RangeCoder.cpp:29-81) as per-symbol counts with group sums: total, low(c) and the
// symbol holding a cumulative count in ~32 steps.  Halving at 2^48 cannot be reached inside a block.
struct Model {
    int n = 0;
    uint32_t cnt[256], grp[16];
    uint64_t total = 0;
    void init(int symbols) {
        n = symbols;
        for (int i = 0; i < 256; i++) cnt[i] = i < n ? 1u : 0u;
        for (int g = 0; g < 16; g++) { uint32_t s = 0; for (int j = 0; j < 16; j++) s += cnt[g * 16 + j]; grp[g] = s; }
        total = (uint64_t)n;
    }
    // largest symbol c with low(c) <= count (RangeDecoder::nextByte); *low gets low(c)
    int find(uint64_t count, uint64_t *low) const {
        uint64_t s = 0;
        int g = 0;
        const int ng = (n + 15) >> 4;
        while (g + 1 < ng && s + grp[g] <= count) { s += grp[g]; g++; }
        int c = g * 16;
        const int last = n - 1;
        while (c < last && s + cnt[c] <= count) { s += cnt[c]; c++; }
        *low = s;
        return c;
    }
    void update(int c) { cnt[c]++; grp[c >> 4]++; total++; }
};

struct RangeDecoder {
    uint64_t low = 0, range = ~0ull, code = 0;
    const uint8_t *p = nullptr, *end = nullptr;
    uint8_t next() { return p < end ? *p++ : (uint8_t)0xff; }       // past the end the reference reads EOF (0xff)
    void start(const uint8_t *data, uint64_t n) {
        low = 0; range = ~0ull; code = 0; p = data; end = data + n;
        for (int i = 0; i < 8; i++) code = (code << 8) | next();
    }
    int decode(Model &m) {
        range /= m.total;
        const uint64_t count = (code - low) / range;
        uint64_t lo;
        const int c = m.find(count, &lo);
        low += lo * range;
        range *= m.cnt[c];
        while ((low ^ (low + range)) < RC_TOP || range < RC_BOTTOM) {
            if (range < RC_BOTTOM && (low ^ (low + range)) >= RC_TOP) range = (0 - low) & (RC_BOTTOM - 1);
            code = (code << 8) | next();
            range <<= 8;
            low <<= 8;
        }
        m.update(c);
        return c;
    }
    uint64_t numeric(Model *models) {                                 // C/CompressionUtils.hpp:114-137
        uint64_t value = 0;
        int shift = 0, i = 0, byte;
        do {
            byte = decode(models[i < NB_NUMERIC ? i : NB_NUMERIC - 1]);
            if (shift < 64) value += (uint64_t)(byte & 127) << shift;
            shift += 7;
            i++;
        } while (byte > 127);
        return value;
    }
};

struct DnaModels {                                                    // C/DnaCoder.cpp:43-112
    Model readType, noAnchorRead, bifurcation, bifurcationBinary, readAnchorRevcomp;
    Model anchorAddress[NB_NUMERIC], anchorPos[NB_NUMERIC], noAnchorReadSize[NB_NUMERIC], readSize[NB_NUMERIC], Npos[NB_NUMERIC],
        leftErrorPos[NB_NUMERIC], numeric[NB_NUMERIC], leftError[NB_NUMERIC];
    void clear() {
        readType.init(2); noAnchorRead.init(5); bifurcation.init(5); bifurcationBinary.init(2); readAnchorRevcomp.init(2);
        for (int i = 0; i < NB_NUMERIC; i++) {
            anchorAddress[i].init(256); anchorPos[i].init(256); noAnchorReadSize[i].init(256); readSize[i].init(256);
            Npos[i].init(256); leftErrorPos[i].init(256); numeric[i].init(256); leftError[i].init(256);
        }
    }
};

static const char BIN2NT[5] = {'A', 'C', 'T', 'G', 'N'};              // Leon::bin2nt, C/Leon.cpp:58-66

// The 4 successors (right) or predecessors (left) of a k-mer that the filter holds, bit nt = "ACTG"[nt]
// (BloomNeighborCoherent::contains4, COL/Bloom.hpp:645-798): the k - 2 shared nucleotides give the window start,
// the two end nucleotides of each candidate the bit inside it, for each of the 7 hash functions.
template <typename K> unsigned contains4(const BloomDev &b, K fwd, bool right) {
    const int k = b.k;
    const K rc = kmer_revcomp(fwd, k);
    const K m2 = kmer_mask(K(0), k - 2);
    const K mid = right ? (fwd & m2) : (fwd >> 4);
    const K rmid = right ? (rc >> 4) : (rc & m2);
    const K hp = rmid < mid ? rmid : mid;
    const uint64_t racine = bloom_racine(hp, b);
    const unsigned fixed = right ? (unsigned)((fwd >> (2 * (k - 2))) & 3) : (unsigned)((fwd >> 2) & 3);
    unsigned pos[4];
    for (unsigned nt = 0; nt < 4; nt++) pos[nt] = right ? bloom_cano2((fixed << 2) + nt) : bloom_cano2((nt << 2) + fixed);
    unsigned alive = 15u;
    for (int i = 0; i < 7 && alive; i++) {
        const uint64_t bit = racine + (i ? (kmer_simplehash16(hp, i, kSimpleHashTable) & 4095) : 0);
        uint64_t w;
        memcpy(&w, (const uint8_t *)b.words + (bit >> 5) * 4, 8);     // 64 filter bits from the word holding `bit`
        const uint32_t f = (uint32_t)(w >> (bit & 31));
        unsigned m = 0;
        for (unsigned nt = 0; nt < 4; nt++) m |= ((f >> pos[nt]) & 1u) << nt;
        alive &= m;
    }
    return alive;
}

template <typename K> K advance(K kmer, unsigned nt, bool right, int k) {   // codeSeedBin / codeSeedNT
    if (right) return ((kmer << 2) | K((uint64_t)nt)) & kmer_mask(K(0), k);
    return (kmer >> 2) | (K((uint64_t)nt) << (2 * (k - 1)));
}

template <typename K> struct BlockDecoder {
    const BloomDev &bloom;
    const std::vector<K> &anchors;
    int k;
    DnaModels m;
    RangeDecoder rc;
    std::vector<uint64_t> npos, errpos;
    std::string err;

    static constexpr uint64_t MAX_SANE_READ = 1ull << 31;     // leon_upload refuses longer reads; containers without maxSequenceSize
    BlockDecoder(const BloomDev &b, const std::vector<K> &a) : bloom(b), anchors(a), k(b.k) {}

    // one step of DnaDecoder::extendAnchor (C/DnaCoder.cpp:1482-1605); returns the character at `pos`
    char extend(K &kmer, uint64_t pos, bool right) {
        if (std::find(npos.begin(), npos.end(), pos) != npos.end()) {
            kmer = advance(kmer, 0u, right, k);                       // an N advances the walk as 'A'
            return 'A';
        }
        if (std::find(errpos.begin(), errpos.end(), pos) != errpos.end()) {
            const int nt = rc.decode(m.bifurcation);
            const unsigned res4 = contains4(bloom, kmer, right);
            if (res4) {                                               // the walk follows the graph, the read keeps its base
                unsigned first = 0;
                while (!((res4 >> first) & 1u)) first++;
                kmer = advance(kmer, first, right, k);
                return BIN2NT[nt];
            }
            // (the reference falls through to the general case here; the encoder never produces it)
            err = "sequencing error recorded where the graph has no successor";
            return BIN2NT[nt];
        }
        const unsigned res4 = contains4(bloom, kmer, right);
        const int cnt = __builtin_popcount(res4);
        unsigned nt;
        if (cnt == 1) {
            nt = (unsigned)__builtin_ctz(res4);
        } else if (cnt == 2) {
            const unsigned nt1 = (unsigned)__builtin_ctz(res4), nt2 = (unsigned)__builtin_ctz(res4 & (res4 - 1));
            nt = rc.decode(m.bifurcationBinary) == 0 ? nt1 : nt2;
        } else {
            nt = (unsigned)rc.decode(m.bifurcation);
        }
        if (nt > 3) { kmer = advance(kmer, 0u, right, k); return BIN2NT[4]; }   // Leon::nt2bin('N') = 0 in codeSeedNT
        kmer = advance(kmer, nt, right, k);
        return BIN2NT[nt];
    }

    // Decodes `nreads` reads of one dna_<id> block; sequences go to out, each followed by '\n' (DnaDecoder::_buffer).
    bool run(const uint8_t *data, uint64_t nbytes, uint64_t nreads, int64_t max_read, std::string &out) {
        m.clear();
        rc.start(data, nbytes);
        std::string seq;
        for (uint64_t r = 0; r < nreads; r++) {
            const int type = rc.decode(m.readType);
            seq.clear();
            if (type == 0) {
                const uint64_t L = rc.numeric(m.readSize);
                const uint64_t apos = rc.numeric(m.anchorPos);
                const uint64_t addr = rc.numeric(m.anchorAddress);
                if (addr >= anchors.size() || apos + (uint64_t)k > L || L > (max_read >= 0 ? (uint64_t)max_read : MAX_SANE_READ)) {
                    err = "corrupt block: anchor address, anchor position or read length out of range";
                    return false;
                }
                K anchor = anchors[addr];
                if (rc.decode(m.readAnchorRevcomp) == 1) anchor = kmer_revcomp(anchor, k);
                npos.clear(); errpos.clear();
                uint64_t prev = 0;
                const uint64_t nn = rc.numeric(m.numeric);
                if (nn > L) { err = "corrupt block: more N positions than bases"; return false; }
                for (uint64_t i = 0; i < nn; i++) { prev += rc.numeric(m.Npos); npos.push_back(prev); }
                const uint64_t ne = rc.numeric(m.leftError);
                if (ne > L) { err = "corrupt block: more error positions than bases"; return false; }
                prev = 0;
                for (uint64_t i = 0; i < ne; i++) { prev += rc.numeric(m.leftErrorPos); errpos.push_back(prev); }
                seq.assign((size_t)L, 'A');
                for (int i = 0; i < k; i++) seq[apos + i] = BIN2NT[(unsigned)((anchor >> (2 * (k - 1 - i))) & 3)];
                K kmer = anchor;
                for (uint64_t i = apos; i-- > 0;) seq[i] = extend(kmer, i, false);
                kmer = anchor;
                for (uint64_t i = apos + k; i < L; i++) seq[i] = extend(kmer, i, true);
                for (uint64_t p : npos) if (p < L) seq[p] = 'N';
            } else {
                const uint64_t L = rc.numeric(m.noAnchorReadSize);
                if (L > (max_read >= 0 ? (uint64_t)max_read : MAX_SANE_READ)) { err = "corrupt block: read length out of range"; return false; }
                seq.resize((size_t)L);
                for (uint64_t i = 0; i < L; i++) seq[i] = BIN2NT[rc.decode(m.noAnchorRead)];
            }
            if (!err.empty()) return false;
            out += seq;
            out += '\n';
        }
        return true;
    }
};

// leon/anchors/anchorsDict -> k-mers (C/Leon.cpp:1948-2007): one coder, one 5-symbol model, k symbols per anchor.
template <typename K> bool decode_anchor_dict(const uint8_t *data, uint64_t nbytes, uint64_t count, int k, std::vector<K> &out) {
    Model model;
    model.init(5);
    RangeDecoder rc;
    rc.start(data, nbytes);
    out.resize(count);
    for (uint64_t a = 0; a < count; a++) {
        K kmer(0);
        for (int i = 0; i < k; i++) {
            const int c = rc.decode(model);
            kmer = (kmer << 2) | K((uint64_t)(c & 3));                // codeSeed of the text: N (4) cannot occur in an anchor
        }
        out[a] = kmer;
    }
    return true;
}

template <typename K>
int decode_dna(const leon_dna_stream *s, int nthreads, std::vector<std::string> &blocks, std::string &err) {
    BloomDev bloom;
    memset(&bloom, 0, sizeof bloom);
    bloom.k = s->kmer_size;
    bloom.words = (uint32_t *)s->bloom;
    bloom.reduced_tai = s->bloom_bits;
    bloom.magic = s->bloom_bits ? ~0ull / s->bloom_bits : 0;
    // contains4 probes bits up to reduced_tai + 4095 + 13 with 8-byte loads: what BloomNeighborCoherent allocates
    // (1 + (bits + 2 * 4096) / 8, COL/Bloom.hpp:187,438) plus the load width
    if (s->bloom_nbytes < (s->bloom_bits + 2 * 4096) / 8 + 1 + 8) { err = "Bloom array shorter than its size attribute says"; return LEON_E_ARG; }
    // an anchor takes at least 2 * k / 8 bytes of information only loosely; bound the count by what the stream can hold
    if (s->anchor_count > s->anchors_dict_nbytes * 64 + 64) { err = "anchor count exceeds what the dictionary stream can hold"; return LEON_E_ARG; }
    std::vector<K> anchors;
    decode_anchor_dict<K>(s->anchors_dict, s->anchors_dict_nbytes, s->anchor_count, s->kmer_size, anchors);
    blocks.assign(s->nblocks, std::string());
    std::vector<uint64_t> start(s->nblocks + 1, 0);
    for (uint64_t b = 0; b < s->nblocks; b++) start[b + 1] = start[b] + s->blocksizes[2 * b];
    if (start[s->nblocks] > s->dna_nbytes) { err = "blocksizes exceed dna_nbytes"; return LEON_E_ARG; }
    std::atomic<uint64_t> next(0);
    std::atomic<int> failed(0);
    std::string first_err;
    std::mutex *mu = new std::mutex();
    auto work = [&]() {
        // nothing may leave a pool thread as an exception (std::terminate): sizes come from an untrusted container
        uint64_t b = 0;
        try {
            BlockDecoder<K> d(bloom, anchors);
            for (;;) {
                b = next.fetch_add(1);
                if (b >= s->nblocks || failed.load()) return;
                // a coded read takes at least a few bits: cap the reservation by what the block's bytes can expand to
                const uint64_t want = s->blocksizes[2 * b + 1] * 102, cap = s->blocksizes[2 * b] * 64 + 4096;
                blocks[b].reserve((size_t)std::min(want, cap));
                if (!d.run(s->dna + start[b], s->blocksizes[2 * b], s->blocksizes[2 * b + 1], s->max_read_size, blocks[b])) {
                    std::lock_guard<std::mutex> g(*mu);
                    if (!failed.exchange(1)) first_err = "block " + std::to_string(b) + ": " + d.err;
                    return;
                }
            }
        } catch (const std::exception &e) {
            std::lock_guard<std::mutex> g(*mu);
            if (!failed.exchange(1)) first_err = "block " + std::to_string(b) + ": " + e.what();
        }
    };
    unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::max(1u, std::thread::hardware_concurrency());
    nt = (unsigned)std::min<uint64_t>(nt, std::max<uint64_t>(s->nblocks, 1));
    std::vector<std::thread> pool;
    for (unsigned i = 1; i < nt; i++) pool.emplace_back(work);
    work();
    for (auto &t : pool) t.join();
    delete mu;
    if (failed.load()) { err = first_err; return LEON_E_ARG; }
    return LEON_OK;
}

// ---- the same stream decoded on the GPU: k_decode_blocks (decode_device.cuh), one warp per block --------------------
// The anchor dictionary is one never-reset coder over all anchors (C/Leon.cpp:1948-2007), a single dependent chain: it
// is decoded by the calling host thread (like its encoder) while the filter and the blocks are uploaded.
template <typename K>
int decode_dna_device(const leon_dna_stream *s, int device, std::vector<std::string> &blocks, float *kernel_ms, std::string &err) {
    if (s->max_read_size < 0) { err = "the device decoder needs max_read_size (metadata/maxSequenceSize) to size its output"; return LEON_E_ARG; }
    if (s->bloom_nbytes < (s->bloom_bits + 2 * 4096) / 8 + 1 + 8) { err = "Bloom array shorter than its size attribute says"; return LEON_E_ARG; }
    std::vector<uint64_t> start(s->nblocks + 1, 0);
    for (uint64_t b = 0; b < s->nblocks; b++) start[b + 1] = start[b] + s->blocksizes[2 * b];
    if (start[s->nblocks] > s->dna_nbytes) { err = "blocksizes exceed dna_nbytes"; return LEON_E_ARG; }
    blocks.assign(s->nblocks, std::string());
    if (!s->nblocks) return LEON_OK;
    DevBuf d_bloom, d_tbl, d_anchors, d_dna, d_blocks, d_ws, d_out, d_len, d_status;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = LEON_OK;
    try {
        CUDA_TRY(cudaSetDevice(device));
        CUDA_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        int sms = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        const uint64_t max_read = (uint64_t)s->max_read_size;
        std::vector<lbdd::BlockIn> bin(s->nblocks);
        uint64_t out_total = 0;
        for (uint64_t b = 0; b < s->nblocks; b++) {
            bin[b] = lbdd::BlockIn{start[b], s->blocksizes[2 * b], s->blocksizes[2 * b + 1], out_total};
            out_total += s->blocksizes[2 * b + 1] * (max_read + 1);
        }
        const uint32_t map_words = (uint32_t)((max_read + 31) / 32 + 1);
        const size_t smem = 2048 + 2 * (size_t)map_words * 4;
        if (smem > 200 * 1024) throw LeonError(LEON_E_ARG, "reads longer than the device decoder's position bitmaps hold");
        const unsigned grid = (unsigned)std::min<uint64_t>(s->nblocks, (uint64_t)sms * 8);
        d_bloom.reserve(s->bloom_nbytes + 16);
        d_tbl.reserve(256 * 8);
        d_dna.reserve(s->dna_nbytes + 16);
        d_blocks.reserve(bin.size() * sizeof(lbdd::BlockIn));
        d_ws.reserve((size_t)grid * lbdd::MODEL_WORDS * 4);
        d_out.reserve(out_total + 16);
        d_len.reserve(s->nblocks * 8);
        d_status.reserve(s->nblocks * 4);
        CUDA_TRY(cudaMemsetAsync(d_bloom.p, 0, s->bloom_nbytes + 16, st));
        CUDA_TRY(cudaMemcpyAsync(d_bloom.p, s->bloom, s->bloom_nbytes, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_tbl.p, kSimpleHashTable, 256 * 8, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_dna.p, s->dna, s->dna_nbytes, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_blocks.p, bin.data(), bin.size() * sizeof(lbdd::BlockIn), cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemsetAsync(d_status.p, 0xff, s->nblocks * 4, st));
        std::vector<K> anchors;
        decode_anchor_dict<K>(s->anchors_dict, s->anchors_dict_nbytes, s->anchor_count, s->kmer_size, anchors);
        d_anchors.reserve(anchors.size() * sizeof(K) + 16);
        CUDA_TRY(cudaMemcpyAsync(d_anchors.p, anchors.data(), anchors.size() * sizeof(K), cudaMemcpyHostToDevice, st));
        BloomDev bloom;
        memset(&bloom, 0, sizeof bloom);
        bloom.k = s->kmer_size;
        bloom.words = d_bloom.as<uint32_t>();
        bloom.reduced_tai = s->bloom_bits;
        bloom.magic = s->bloom_bits ? ~0ull / s->bloom_bits : 0;
        bloom.nbytes = s->bloom_nbytes;
        CUDA_TRY(cudaFuncSetAttribute(lbdd::k_decode_blocks<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CUDA_TRY(cudaEventCreate(&e0));
        CUDA_TRY(cudaEventCreate(&e1));
        CUDA_TRY(cudaEventRecord(e0, st));
        lbdd::k_decode_blocks<K><<<grid, 32, smem, st>>>(bloom, d_tbl.as<uint64_t>(), d_anchors.as<K>(), (uint64_t)anchors.size(),
                                                          d_dna.as<uint8_t>(), d_blocks.as<lbdd::BlockIn>(), s->nblocks, d_ws.as<uint32_t>(),
                                                          d_out.as<uint8_t>(), d_len.as<uint64_t>(), d_status.as<uint32_t>(),
                                                          (uint32_t)max_read, map_words);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(e1, st));
        std::vector<uint64_t> len(s->nblocks);
        std::vector<uint32_t> status(s->nblocks);
        CUDA_TRY(cudaMemcpyAsync(len.data(), d_len.p, s->nblocks * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(status.data(), d_status.p, s->nblocks * 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (kernel_ms) CUDA_TRY(cudaEventElapsedTime(kernel_ms, e0, e1));
        static const char *const why[] = {"", "corrupt block: anchor address, anchor position or read length out of range",
                                          "corrupt block: more N positions than bases", "corrupt block: more error positions than bases",
                                          "sequencing error recorded where the graph has no successor", "corrupt block: read length out of range"};
        for (uint64_t b = 0; b < s->nblocks && rc == LEON_OK; b++) {
            if (status[b] != lbdd::DE_OK) {
                err = "block " + std::to_string(b) + ": " + (status[b] < 6 ? why[status[b]] : "not decoded");
                rc = LEON_E_ARG;
            }
        }
        if (rc == LEON_OK) {
            for (uint64_t b = 0; b < s->nblocks; b++) {
                blocks[b].resize((size_t)len[b]);
                if (len[b]) CUDA_TRY(cudaMemcpyAsync(&blocks[b][0], d_out.as<uint8_t>() + bin[b].out_off, len[b], cudaMemcpyDeviceToHost, st));
            }
            CUDA_TRY(cudaStreamSynchronize(st));
        }
    } catch (const LeonError &e) {
        err = e.msg;
        rc = e.code;
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (st) cudaStreamDestroy(st);
    for (DevBuf *b : {&d_bloom, &d_tbl, &d_anchors, &d_dna, &d_blocks, &d_ws, &d_out, &d_len, &d_status}) b->release();
    return rc;
}

}  // namespace lbdecode


extern "C" int leon_decode_dna(const leon_dna_stream *s, int32_t nthreads, leon_decoded **out) {
    try {
        if (!s || !out || !s->bloom || (s->nblocks && (!s->dna || !s->blocksizes))) { leon_internal_set_error("leon_decode_dna: null argument"); return LEON_E_ARG; }
        if (s->kmer_size < 2 || s->kmer_size > 63) { leon_internal_set_error("leon_decode_dna: bad kmer_size"); return LEON_E_ARG; }
        std::unique_ptr<leon_decoded> d(new leon_decoded());
        std::string err;
        const int rc = s->kmer_size <= 31 ? lbdecode::decode_dna<uint64_t>(s, nthreads, d->blocks, err)
                                          : lbdecode::decode_dna<u128>(s, nthreads, d->blocks, err);
        if (rc) { leon_internal_set_error(("leon_decode_dna: " + err).c_str()); return rc; }
        *out = d.release();
        return LEON_OK;
    } catch (const std::bad_alloc &) {
        leon_internal_set_error("leon_decode_dna: out of host memory");
        return LEON_E_NOMEM;
    }
}

extern "C" int leon_decode_dna_device(const leon_dna_stream *s, int32_t device, leon_decoded **out, float *kernel_ms) {
    try {
        if (!s || !out || !s->bloom || (s->nblocks && (!s->dna || !s->blocksizes))) { leon_internal_set_error("leon_decode_dna_device: null argument"); return LEON_E_ARG; }
        if (s->kmer_size < 2 || s->kmer_size > 63) { leon_internal_set_error("leon_decode_dna_device: bad kmer_size"); return LEON_E_ARG; }
        std::unique_ptr<leon_decoded> d(new leon_decoded());
        std::string err;
        const int rc = s->kmer_size <= 31 ? lbdecode::decode_dna_device<uint64_t>(s, device, d->blocks, kernel_ms, err)
                                          : lbdecode::decode_dna_device<u128>(s, device, d->blocks, kernel_ms, err);
        if (rc) { leon_internal_set_error(("leon_decode_dna_device: " + err).c_str()); return rc; }
        *out = d.release();
        return LEON_OK;
    } catch (const std::bad_alloc &) {
        leon_internal_set_error("leon_decode_dna_device: out of host memory");
        return LEON_E_NOMEM;
    }
}

extern "C" int leon_decoded_block(const leon_decoded *d, uint64_t block, const uint8_t **text, uint64_t *nbytes) {
    if (!d || block >= d->blocks.size() || !text || !nbytes) { leon_internal_set_error("leon_decoded_block: bad argument"); return LEON_E_ARG; }
    *text = (const uint8_t *)d->blocks[block].data();
    *nbytes = d->blocks[block].size();
    return LEON_OK;
}

extern "C" void leon_decoded_free(leon_decoded *d) { delete d; }
