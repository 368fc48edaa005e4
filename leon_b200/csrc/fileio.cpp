// File-level entry point: what `leon -file X -c` does end to end (C/Leon.cpp:342-470 executeCompression): read the
// FASTA/FASTQ file, code the header stream on host threads while the GPU codes the sequence stream, write the `.leon`
// container.  The record reader follows the reference's bank (bank/impl/BankFasta.cpp:484-570): a record starts at a
// line beginning with '>' or '@', the name is the rest of that line, the sequence is every following line up to one
// beginning with '>', '@' or '+'; after '+' the quality is read line by line until it is as long as the sequence.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/leon_b200.h"

extern "C" void leon_internal_set_error(const char *msg);

namespace {

struct Mapped {
    const uint8_t *p = nullptr;
    size_t n = 0;
    int fd = -1;
    size_t file_bytes = 0;                   // size of the file itself (== n unless it was gzipped)
    std::vector<uint8_t> inflated;           // gzipped input: the text, p points here
    const uint8_t *map_p = nullptr;
    size_t map_n = 0;
    ~Mapped() {
        if (map_p && map_n) munmap((void *)map_p, map_n);
        if (fd >= 0) close(fd);
    }
};

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

inline const uint8_t *line_end(const uint8_t *p, const uint8_t *end) {
    const uint8_t *e = (const uint8_t *)memchr(p, '\n', (size_t)(end - p));
    return e ? e : end;
}
inline size_t trimmed(const uint8_t *p, const uint8_t *e) { return (size_t)((e > p && e[-1] == '\r' ? e - 1 : e) - p); }
inline const uint8_t *next_line(const uint8_t *e, const uint8_t *end) { return e < end ? e + 1 : end; }

// Walks the records of [p, end) with the reference's grammar (bank/impl/BankFasta.cpp:484-570), p at a record marker, and
// stops at the first record that starts at or after `stop`.  The sink sees every name line, sequence line and quality
// line.  `strict` comes back false when the range holds anything but one-line sequences, one-line qualities and
// back-to-back records (FASTQ) or holds a quality section at all (FASTA): the conditions under which ranges cut at
// guessed record starts are guaranteed to parse like the whole file does.
template <class Sink>
int walk_records(const uint8_t *p, const uint8_t *end, const uint8_t *stop, bool fastq_mode, Sink &s, const uint8_t **stopped,
                 bool *strict, std::string &err) {
    *strict = true;
    while (p < end && p < stop) {
        const uint8_t *e = line_end(p, end);                 // name line, marker excluded
        s.header(p + 1, trimmed(p + 1, e));
        p = next_line(e, end);
        uint64_t len = 0;
        int seq_lines = 0;
        while (p < end && *p != '>' && *p != '@' && *p != '+') {
            e = line_end(p, end);
            size_t n = trimmed(p, e);
            s.seq(p, n);
            len += n;
            seq_lines++;
            p = next_line(e, end);
        }
        s.end_seq();
        if (p < end && *p == '+') {
            if (!fastq_mode) *strict = false;
            e = line_end(p, end);
            p = next_line(e, end);
            uint64_t got = 0;
            int qual_lines = 0;
            do {                                             // at least one line, then until as long as the sequence
                if (p >= end) break;
                e = line_end(p, end);
                size_t n = trimmed(p, e);
                if (!s.qual(p, n, got, len)) { err = "record " + std::to_string(s.record()) + ": quality longer than sequence"; return LEON_E_ARG; }
                got += n;
                qual_lines++;
                p = next_line(e, end);
            } while (got < len);
            if (s.wants_quals() && got < len) { err = "record " + std::to_string(s.record()) + ": quality shorter than sequence"; return LEON_E_ARG; }
            if (seq_lines != 1 || qual_lines != 1 || (p < end && *p != '@' && *p != '>')) *strict = false;
            while (p < end && *p != '>' && *p != '@') p++;   // the reference scans for the next marker character
        } else {
            if (fastq_mode) *strict = false;
            if (s.wants_quals()) { err = "record " + std::to_string(s.record()) + " has no quality line (FASTQ expected)"; return LEON_E_ARG; }
        }
        s.end_record();
    }
    *stopped = p;
    return LEON_OK;
}

struct CountSink {
    uint64_t records = 0, bases = 0, header_bytes = 0, first_record = 0;
    bool any_qual = false, quals = false;
    void header(const uint8_t *, size_t n) { header_bytes += n; }
    void seq(const uint8_t *, size_t n) { bases += n; }
    void end_seq() {}
    bool qual(const uint8_t *, size_t n, uint64_t got, uint64_t len) { any_qual = true; return !quals || got + n <= len; }
    void end_record() { records++; }
    bool wants_quals() const { return quals; }
    uint64_t record() const { return first_record + records; }
};

struct WriteSink {
    uint8_t *bases, *quals, *headers;                      // bases of the whole set; quals / headers may be null
    uint64_t *offsets, *header_offsets;
    uint64_t rec, nb, nh, start = 0;
    void header(const uint8_t *p, size_t n) { if (headers) { memcpy(headers + nh, p, n); nh += n; header_offsets[rec + 1] = nh; } }
    void seq(const uint8_t *p, size_t n) { memcpy(bases + nb, p, n); nb += n; }
    void end_seq() { offsets[rec + 1] = nb; }
    bool qual(const uint8_t *p, size_t n, uint64_t got, uint64_t len) {
        if (!quals) return true;
        if (got + n > len) return false;
        memcpy(quals + (nb - len) + got, p, n);
        return true;
    }
    void end_record() { rec++; }
    bool wants_quals() const { return quals != nullptr; }
    uint64_t record() const { return rec; }
};

// Exactly sized anonymous mapping, transparent huge pages requested: the first touch of a fresh gigabyte by the writer
// threads is page-fault bound with 4 KiB pages.
struct HostBuf {
    uint8_t *p = nullptr;
    size_t cap = 0;
    bool alloc(size_t n) {
        cap = ((n ? n : 1) + (2u << 20) - 1) & ~(size_t)((2u << 20) - 1);
        void *q = mmap(nullptr, cap, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (q == MAP_FAILED) { cap = 0; return false; }
        madvise(q, cap, MADV_HUGEPAGE);
        p = (uint8_t *)q;
        return true;
    }
    ~HostBuf() { if (p) munmap(p, cap); }
};

struct ReadSet {
    HostBuf bases, quals, headers;
    std::vector<uint64_t> offsets, header_offsets;
    uint64_t nreads = 0, header_bytes = 0;
    bool any_qual = false;
    unsigned threads_used = 1;
};

// Record start at line start c?  FASTA: '>'.  FASTQ: '@' whose line after next starts with '+' (a quality line that
// begins with '@' is followed, two lines on, by a sequence line, which cannot begin with '+').
bool record_start(const uint8_t *c, const uint8_t *end, bool fastq_mode) {
    if (c >= end) return true;
    if (!fastq_mode) return *c == '>';
    if (*c != '@') return false;
    const uint8_t *l2 = next_line(line_end(next_line(line_end(c, end), end), end), end);
    return l2 < end && *l2 == '+';
}

// Two passes (count, then write into exactly sized buffers) over up to nthreads ranges cut at record starts; when any
// range is not strictly regular the file is parsed as one range, i.e. serially like the reference.
int parse_reads(const Mapped &m, bool want_quals, bool want_headers, unsigned nthreads, ReadSet &rs, std::string &err) {
    const uint8_t *p = m.p, *end = m.p + m.n;
    while (p < end && *p != '>' && *p != '@') p++;
    const bool fastq_mode = p < end && *p == '@';
    unsigned T = nthreads ? nthreads : std::max(1u, std::thread::hardware_concurrency());
    T = std::min<unsigned>(T, 64);
    if ((size_t)(end - p) < (4u << 20)) T = 1;
    const bool trace = getenv("LEON_B200_TRACE") != nullptr;
    for (;;) {
        const double ta = now_ms();
        std::vector<const uint8_t *> cut(T + 1, end);
        cut[0] = p;
        for (unsigned i = 1; i < T; i++) {
            const uint8_t *c = p + (size_t)(end - p) / T * i;
            c = next_line(line_end(c, end), end);
            while (c < end && !record_start(c, end, fastq_mode)) c = next_line(line_end(c, end), end);
            cut[i] = std::max(c, cut[i - 1]);
        }
        std::vector<CountSink> cnt(T);
        std::vector<int> rcs(T, LEON_OK);
        std::vector<std::string> errs(T);
        std::vector<uint8_t> ok(T, 1);
        auto run = [&](auto &&fn) {
            std::vector<std::thread> pool;
            for (unsigned i = 1; i < T; i++) pool.emplace_back(fn, i);
            fn(0u);
            for (auto &t : pool) t.join();
        };
        run([&](unsigned i) {
            const uint8_t *stopped = cut[i];
            bool strict = true;
            cnt[i].quals = want_quals;
            if (cut[i] < cut[i + 1]) rcs[i] = walk_records(cut[i], end, cut[i + 1], fastq_mode, cnt[i], &stopped, &strict, errs[i]);
            ok[i] = (strict && stopped == cut[i + 1]) ? 1 : 0;
        });
        const double tb = now_ms();
        bool regular = true;
        for (unsigned i = 0; i < T; i++) regular = regular && ok[i];
        if (T > 1 && !regular) { T = 1; continue; }          // irregular layout: one serial range
        for (unsigned i = 0; i < T; i++) if (rcs[i]) {       // (record numbers inside a message are range-relative when T > 1)
            if (T > 1) { T = 1; break; }
            err = errs[i];
            return rcs[i];
        }
        if (T == 1 && rcs[0]) continue;
        std::vector<uint64_t> rec0(T + 1, 0), nb0(T + 1, 0), nh0(T + 1, 0);
        for (unsigned i = 0; i < T; i++) {
            rec0[i + 1] = rec0[i] + cnt[i].records;
            nb0[i + 1] = nb0[i] + cnt[i].bases;
            nh0[i + 1] = nh0[i] + cnt[i].header_bytes;
            rs.any_qual = rs.any_qual || cnt[i].any_qual;
        }
        rs.nreads = rec0[T];
        rs.header_bytes = nh0[T];
        rs.threads_used = T;
        rs.offsets.assign(rs.nreads + 1, 0);
        if (want_headers) rs.header_offsets.assign(rs.nreads + 1, 0);
        if (!rs.bases.alloc(nb0[T]) || (want_quals && !rs.quals.alloc(nb0[T])) || (want_headers && !rs.headers.alloc(nh0[T]))) {
            err = "cannot allocate host memory for the reads";
            return LEON_E_NOMEM;
        }
        const double tc = now_ms();
        run([&](unsigned i) {
            WriteSink w{rs.bases.p, want_quals ? rs.quals.p : nullptr, want_headers ? rs.headers.p : nullptr, rs.offsets.data(),
                        want_headers ? rs.header_offsets.data() : nullptr, rec0[i], nb0[i], nh0[i]};
            const uint8_t *stopped;
            bool strict;
            if (cut[i] < cut[i + 1]) rcs[i] = walk_records(cut[i], end, cut[i + 1], fastq_mode, w, &stopped, &strict, errs[i]);
        });
        for (unsigned i = 0; i < T; i++) if (rcs[i]) { err = errs[i]; return rcs[i]; }
        if (trace) fprintf(stderr, "[leon_b200] read: %u threads, count pass %.1f ms, alloc %.1f ms, write pass %.1f ms\n", T, tb - ta, tc - tb, now_ms() - tc);
        return LEON_OK;
    }
}

int map_file(const char *path, Mapped &m, std::string &err) {
    m.fd = open(path, O_RDONLY);
    struct stat st;
    if (m.fd < 0 || fstat(m.fd, &st) != 0) { err = std::string("cannot open ") + path + ": " + strerror(errno); return LEON_E_IO; }
    m.n = m.file_bytes = (size_t)st.st_size;
    if (m.n) {
        void *q = mmap(nullptr, m.n, PROT_READ, MAP_PRIVATE, m.fd, 0);
        if (q == MAP_FAILED) { m.n = 0; err = std::string("cannot map ") + path + ": " + strerror(errno); return LEON_E_IO; }
        m.p = m.map_p = (const uint8_t *)q;
        m.map_n = m.n;
        madvise(q, m.n, MADV_SEQUENTIAL);
    }
    // The reference reads every input through gzopen / gzread (bank/impl/BankFasta.cpp:652), so `leon -file x.fastq.gz -c`
    // works there: a gzip file (magic 1f 8b; several members as bgzip writes them included) is inflated here first.
    if (m.n >= 2 && m.p[0] == 0x1f && m.p[1] == 0x8b) {
        z_stream zs;
        memset(&zs, 0, sizeof zs);
        if (inflateInit2(&zs, 15 + 16) != Z_OK) { err = "zlib: inflateInit2 failed"; return LEON_E_IO; }
        std::vector<uint8_t> &o = m.inflated;
        o.resize(std::max<size_t>(m.n * 4, 1u << 20));
        size_t in_at = 0, out_at = 0;
        int zr = Z_OK;
        while (in_at < m.n) {
            zs.next_in = (Bytef *)(m.p + in_at);
            const size_t in_chunk = std::min<size_t>(m.n - in_at, 1u << 30);
            zs.avail_in = (uInt)in_chunk;
            if (out_at == o.size()) o.resize(o.size() * 2);
            const size_t out_chunk = std::min<size_t>(o.size() - out_at, 1u << 30);
            zs.next_out = o.data() + out_at;
            zs.avail_out = (uInt)out_chunk;
            zr = inflate(&zs, Z_NO_FLUSH);
            in_at += in_chunk - zs.avail_in;
            out_at += out_chunk - zs.avail_out;
            if (zr == Z_STREAM_END) {                    // next member, if any
                if (in_at >= m.n) break;
                if (inflateReset(&zs) != Z_OK) break;
                zr = Z_OK;
                continue;
            }
            if (zr != Z_OK && zr != Z_BUF_ERROR) break;
            if (zr == Z_BUF_ERROR && zs.avail_in == 0 && in_at >= m.n) break;
        }
        inflateEnd(&zs);
        if (zr != Z_STREAM_END) { err = std::string("cannot inflate ") + path + ": corrupt or truncated gzip data"; return LEON_E_IO; }
        o.resize(out_at);
        m.p = o.data();
        m.n = o.size();
    }
    return LEON_OK;
}

struct FileReads {
    leon_file_reads pub;
    ReadSet rs;
};

struct Joiner {                                          // joins on every exit path, exceptions included
    std::thread &t;
    ~Joiner() { if (t.joinable()) t.join(); }
};

bool has_any(const std::string &s, const char *a, const char *b) { return s.find(a) != std::string::npos || s.find(b) != std::string::npos; }

}  // namespace

extern "C" int leon_read_file(const char *path, int want_quals, int want_headers, leon_file_reads **out) {
    std::string err;
    try {
        if (!path || !out) { leon_internal_set_error("leon_read_file: null argument"); return LEON_E_ARG; }
        Mapped m;
        int rc = map_file(path, m, err);
        if (rc) { leon_internal_set_error(err.c_str()); return rc; }
        std::unique_ptr<FileReads> f(new FileReads());
        rc = parse_reads(m, want_quals != 0, want_headers != 0, 0, f->rs, err);
        if (rc) { leon_internal_set_error(("leon_read_file: " + err).c_str()); return rc; }
        memset(&f->pub, 0, sizeof f->pub);
        f->pub.reads.bases = f->rs.bases.p;
        f->pub.reads.offsets = f->rs.offsets.data();
        f->pub.reads.quals = want_quals ? f->rs.quals.p : nullptr;
        f->pub.reads.nreads = f->rs.nreads;
        f->pub.headers = want_headers ? f->rs.headers.p : nullptr;
        f->pub.header_offsets = want_headers ? f->rs.header_offsets.data() : nullptr;
        f->pub.is_fastq = f->rs.any_qual ? 1 : 0;
        f->pub.threads_used = (int32_t)f->rs.threads_used;
        *out = &f.release()->pub;
        return LEON_OK;
    } catch (const std::bad_alloc &) {
        leon_internal_set_error("leon_read_file: out of host memory");
        return LEON_E_NOMEM;
    }
}

extern "C" void leon_file_reads_free(leon_file_reads *p) {
    if (p) delete reinterpret_cast<FileReads *>(p);      // pub is the first member
}

extern "C" int leon_compress_file(const char *in_path, const char *out_path, const leon_file_opts *o, leon_file_stats *stats) {
    std::string err;
    leon_ctx *ctx = nullptr;
    leon_header_stream *hs = nullptr;
    int rc = LEON_OK;
    try {
        if (!in_path || !o) { leon_internal_set_error("leon_compress_file: null argument"); return LEON_E_ARG; }
        const double t0 = now_ms();
        const std::string in(in_path);
        // output name, C/Leon.cpp:466-479: X.leon, or <dir>/<name minus its last suffix>.leon when the extension holds "gz"
        std::string out = out_path ? std::string(out_path) : in + ".leon";
        if (!out_path) {
            const size_t dot = in.find_last_of('.');
            if (dot != std::string::npos && in.substr(dot + 1).find("gz") != std::string::npos) {
                const size_t slash = in.find_last_of('/');
                const std::string dir = slash == std::string::npos ? "." : in.substr(0, slash);
                std::string name = slash == std::string::npos ? in : in.substr(slash + 1);
                const size_t nd = name.find_last_of('.');
                if (nd != std::string::npos) name = name.substr(0, nd);
                out = dir + "/" + name + ".leon";
            }
        }
        // format from the file name, as C/Leon.cpp:392-416
        bool fastq_input;
        if (has_any(in, ".fq", ".fastq")) fastq_input = true;
        else if (has_any(in, ".fa", ".fasta")) fastq_input = false;
        else { leon_internal_set_error("input extension must be one among fasta (.fa, .fasta) or fastq (.fq, .fastq)"); return LEON_E_ARG; }
        const bool noheader = o->noheader || o->seq_only;
        const bool is_fasta = !fastq_input || o->noqual || o->seq_only;     // container "type": decompresses to FASTA

        leon_opts lo;
        memset(&lo, 0, sizeof lo);
        lo.kmer_size = o->kmer_size ? o->kmer_size : 31;
        lo.reads_per_block = o->reads_per_block ? o->reads_per_block : 50000;
        lo.has_qual = is_fasta ? 0 : 1;
        lo.lossless = o->lossless;
        lo.abundance_set = o->abundance_set;
        lo.device = o->device;
        lo.qual_zlib_on_device = lo.has_qual;                 // quality blocks leave the GPU as zlib streams
        lo.skip_qual_text = lo.has_qual;
        lo.skip_read_arrays = 1;                              // per-read parity arrays are not stored in the container
        // the CUDA context comes up (a few hundred ms) while the host threads read the file
        int crc = LEON_OK;
        std::string cerr_msg;
        std::thread create_thread([&]() {
            crc = leon_create(&lo, &ctx);
            if (crc) cerr_msg = leon_last_error(nullptr);
        });
        Joiner join_create{create_thread};
        Mapped m;
        ReadSet rs;
        rc = map_file(in_path, m, err);
        if (!rc) rc = parse_reads(m, !is_fasta, !noheader, (unsigned)std::max(0, o->nthreads), rs, err);
        const double t_parsed = now_ms();
        create_thread.join();
        if (getenv("LEON_B200_TRACE")) fprintf(stderr, "[leon_b200] file read %.1f ms, then waited %.1f ms for the CUDA context\n", t_parsed - t0, now_ms() - t_parsed);
        if (rc) { if (ctx) leon_destroy(ctx); leon_internal_set_error(("leon_compress_file: " + err).c_str()); return rc; }
        if (crc) { leon_internal_set_error(cerr_msg.c_str()); return crc; }
        const uint64_t nreads = rs.nreads;
        const double t1 = now_ms();

        // header stream on host threads while the GPU works on the sequence stream
        int hrc = LEON_OK;
        std::string herr;
        double header_ms = 0;
        std::thread header_thread;
        Joiner join_header{header_thread};
        if (!noheader) {
            header_thread = std::thread([&]() {
                const double a = now_ms();
                hrc = leon_headers_encode(rs.headers.p, rs.header_offsets.data(), nreads, lo.reads_per_block, o->nthreads, &hs);
                if (hrc) herr = leon_last_error(nullptr);
                header_ms = now_ms() - a;
            });
        }
        leon_reads rd;
        rd.bases = rs.bases.p;
        rd.offsets = rs.offsets.data();
        rd.quals = is_fasta ? nullptr : rs.quals.p;
        rd.nreads = nreads;
        leon_result res;
        // Read sets beyond what the device holds at once (or on request) go through the streaming entry points: counted in
        // hash-partitioned passes and encoded batch by batch, the per-block outputs gathered here for the container.
        uint64_t batch_reads = o->stream_reads > 0 ? (uint64_t)o->stream_reads : 0;
        if (!batch_reads && rs.offsets.back() > 3000000000ull) batch_reads = 8000000;
        std::vector<uint8_t> s_dna, s_qz;
        std::vector<uint64_t> s_bs, s_qzs, s_qs;
        if (batch_reads) {
            const uint64_t rpb = (uint64_t)lo.reads_per_block;
            batch_reads = std::max<uint64_t>(rpb, batch_reads / rpb * rpb);          // whole blocks
            std::vector<uint64_t> boffs2[2];
            int flip = 0;
            auto batch_of = [&](uint64_t r0, leon_reads &b) {
                std::vector<uint64_t> &boffs = boffs2[flip];          // two in turn: the batch prefetched and the one in work
                flip ^= 1;
                const uint64_t r1 = std::min(nreads, r0 + batch_reads), base = rs.offsets[r0];
                boffs.resize(r1 - r0 + 1);
                for (uint64_t i = r0; i <= r1; i++) boffs[i - r0] = rs.offsets[i] - base;
                b.bases = rs.bases.p + base;
                b.quals = is_fasta ? nullptr : rs.quals.p + base;
                b.offsets = boffs.data();
                b.nreads = r1 - r0;
            };
            leon_stream_plan plan;
            memset(&plan, 0, sizeof plan);
            plan.total_reads = nreads;
            plan.total_bases = rs.offsets.back();
            plan.count_passes = o->count_passes > 0 ? (uint32_t)o->count_passes : 0;
            uint32_t npasses = 1;
            rc = leon_count_begin(ctx, &plan, &npasses);
            leon_reads b, nextb;
            batch_of(0, b);
            for (uint32_t p = 0; !rc && p < npasses; p++) {
                uint64_t r0 = 0;
                do {
                    r0 += batch_reads;
                    // the next batch is announced before this one is submitted: its upload runs beside the counting of b.  Behind
                    // the last batch of a pass that is the first batch again, of the next pass or (with qualities) of the encoding pass
                    if (r0 < nreads) { batch_of(r0, nextb); leon_stream_prefetch(ctx, &nextb, 0); }
                    else { batch_of(0, nextb); leon_stream_prefetch(ctx, &nextb, p + 1 < npasses ? 0 : 1); }
                    rc = leon_count(ctx, &b, r0 >= nreads);
                    b = nextb;
                } while (!rc && r0 < nreads);
            }
            leon_count_result cr;
            if (!rc) rc = leon_count_finish(ctx, &cr);
            if (!rc) rc = leon_bloom_build(ctx, nullptr, nullptr, nullptr);
            uint64_t first_block = 0, r0 = 0;
            do {
                if (rc) break;
                r0 += batch_reads;
                if (r0 < nreads) { batch_of(r0, nextb); leon_stream_prefetch(ctx, &nextb, 1); }
                leon_block_out bo;
                rc = leon_encode_block(ctx, first_block, &b, &bo);
                if (rc) break;
                s_dna.insert(s_dna.end(), bo.dna, bo.dna + bo.dna_nbytes);
                s_bs.insert(s_bs.end(), bo.blocksizes, bo.blocksizes + 2 * bo.nblocks);
                if (bo.qual_z_nbytes) {
                    s_qz.insert(s_qz.end(), bo.qual_z, bo.qual_z + bo.qual_z_nbytes);
                    s_qzs.insert(s_qzs.end(), bo.qual_z_block_sizes, bo.qual_z_block_sizes + bo.nblocks);
                    s_qs.insert(s_qs.end(), bo.qual_block_sizes, bo.qual_block_sizes + bo.nblocks);
                }
                first_block += bo.nblocks;
                b = nextb;
            } while (r0 < nreads);
            if (!rc) rc = leon_encode_finish(ctx, &res);
            if (!rc) {
                res.nblocks = first_block;
                res.blocksizes = s_bs.data();
                res.dna = s_dna.data(); res.dna_nbytes = s_dna.size();
                if (!s_qzs.empty()) {
                    res.qual_z = s_qz.data(); res.qual_z_nbytes = s_qz.size();
                    res.qual_z_block_sizes = s_qzs.data(); res.qual_block_sizes = s_qs.data();
                }
            }
        } else {
            rc = leon_compress(ctx, &rd, &res);
        }
        if (rc) err = leon_last_error(ctx);
        if (header_thread.joinable()) header_thread.join();
        if (!rc && hrc) { rc = hrc; err = herr; }
        const double t2 = now_ms();
        if (!rc) {
            leon_container_info ci;
            memset(&ci, 0, sizeof ci);
            ci.kmer_size = lo.kmer_size;
            ci.is_fasta = is_fasta;
            ci.has_header = !noheader;
            if (hs) {
                ci.first_header = hs->first_header; ci.first_header_nbytes = hs->first_header_nbytes;
                ci.header_nblocks = hs->nblocks; ci.header_blocksizes = hs->blocksizes; ci.header_data = hs->data;
            }
            rc = leon_write_container(out.c_str(), &res, &ci);
            if (rc) err = leon_last_error(nullptr);
        }
        const double t3 = now_ms();
        if (!rc && stats) {
            memset(stats, 0, sizeof *stats);
            stats->nreads = nreads;
            stats->total_bases = rs.offsets.back();
            stats->input_bytes = m.file_bytes;
            struct stat so;
            if (stat(out.c_str(), &so) == 0) stats->output_bytes = (uint64_t)so.st_size;
            stats->header_bytes_in = rs.header_bytes;
            stats->header_bytes_out = hs ? hs->data_nbytes : 0;
            stats->dna_bytes_out = res.dna_nbytes + res.bloom_nbytes + res.anchors_dict_nbytes;
            stats->nblocks = res.nblocks;
            stats->is_fasta = is_fasta;
            stats->has_header = !noheader;
            stats->headers_coded_serially = hs ? hs->coded_serially : 0;
            stats->parse_ms = (float)(t1 - t0);
            stats->compress_ms = (float)(t2 - t1);
            stats->header_ms = (float)header_ms;
            stats->write_ms = (float)(t3 - t2);
            stats->total_ms = (float)(t3 - t0);
            leon_get_timings(ctx, &stats->gpu);
            stats->cutoff = res.count.cutoff;
            stats->abundance_auto = lo.abundance_set ? 0 : 1;
            stats->nb_solid_kmers = lo.abundance_set ? res.count.nb_solid_total : res.count.nb_solids_cutoff;
            stats->anchor_count = res.anchor_count;
            stats->reads_without_anchor = res.reads_without_anchor;
            stats->bloom_bytes = res.bloom_nbytes;
            stats->anchors_dict_bytes = res.anchors_dict_nbytes;
            stats->seq_bytes_out = res.dna_nbytes;
            stats->qual_bytes_out = res.qual_z_nbytes;
        }
    } catch (const std::bad_alloc &) {
        rc = LEON_E_NOMEM;
        err = "leon_compress_file: out of host memory";
    } catch (const std::exception &e) {
        rc = LEON_E_IO;
        err = e.what();
    }
    if (hs) leon_headers_free(hs);
    if (ctx) leon_destroy(ctx);
    if (rc) leon_internal_set_error(err.c_str());
    return rc;
}
