// File-level entry point: what `leon -file X -c` does end to end (C/Leon.cpp:342-470 executeCompression): read the
// FASTA/FASTQ file, code the header stream on host threads while the GPU codes the sequence stream, write the `.leon`
// container.  The record reader follows the reference's bank (bank/impl/BankFasta.cpp:484-570): a record starts at a
// line beginning with '>' or '@', the name is the rest of that line, the sequence is every following line up to one
// beginning with '>', '@' or '+'; after '+' the quality is read line by line until it is as long as the sequence.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
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
    ~Mapped() {
        if (p && n) munmap((void *)p, n);
        if (fd >= 0) close(fd);
    }
};

// Host buffer for bases / qualities: pinned when a CUDA device is there (true DMA uploads), plain memory otherwise
// (reading a file is not compute; leon_compress still fails without a device).
struct Pinned {
    uint8_t *p = nullptr;
    bool pinned = false;
    bool alloc(size_t n) {
        p = (uint8_t *)leon_alloc_pinned(n);
        pinned = p != nullptr;
        if (!p) p = (uint8_t *)malloc(n);
        return p != nullptr;
    }
    ~Pinned() { if (p) { if (pinned) leon_free_pinned(p); else free(p); } }
};

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

inline const uint8_t *line_end(const uint8_t *p, const uint8_t *end) {
    const uint8_t *e = (const uint8_t *)memchr(p, '\n', (size_t)(end - p));
    return e ? e : end;
}
inline size_t trimmed(const uint8_t *p, const uint8_t *e) { return (size_t)((e > p && e[-1] == '\r' ? e - 1 : e) - p); }

struct ReadSet {
    Pinned bases, quals;
    std::vector<uint64_t> offsets{0}, header_offsets{0};
    std::vector<uint8_t> headers;
    bool any_qual = false;
};

// Returns 0 or a LEON_E_* code; err gets the message.
int parse_reads(const Mapped &m, bool want_quals, bool want_headers, ReadSet &rs, std::string &err) {
    if (!rs.bases.alloc(m.n + 1) || (want_quals && !rs.quals.alloc(m.n + 1))) { err = "cannot allocate host memory for the reads"; return LEON_E_NOMEM; }
    const uint8_t *p = m.p, *end = m.p + m.n;
    uint64_t nb = 0;
    rs.offsets.reserve(m.n / 200 + 16);
    while (p < end && *p != '>' && *p != '@') p++;
    while (p < end) {
        const uint8_t *e = line_end(p, end);                 // name line, marker excluded
        if (want_headers) {
            rs.headers.insert(rs.headers.end(), p + 1, p + 1 + trimmed(p + 1, e));
            rs.header_offsets.push_back(rs.headers.size());
        }
        p = e < end ? e + 1 : end;
        const uint64_t start = nb;
        while (p < end && *p != '>' && *p != '@' && *p != '+') {
            e = line_end(p, end);
            size_t n = trimmed(p, e);
            memcpy(rs.bases.p + nb, p, n);
            nb += n;
            p = e < end ? e + 1 : end;
        }
        const uint64_t len = nb - start;
        rs.offsets.push_back(nb);
        if (p < end && *p == '+') {
            rs.any_qual = true;
            e = line_end(p, end);
            p = e < end ? e + 1 : end;
            uint64_t got = 0;
            do {                                             // at least one line, then until as long as the sequence
                if (p >= end) break;
                e = line_end(p, end);
                size_t n = trimmed(p, e);
                if (want_quals) {
                    // a quality longer than its sequence cannot be laid out at the sequence's offsets
                    if (got + n > len) { err = "record " + std::to_string(rs.offsets.size() - 1) + ": quality longer than sequence"; return LEON_E_ARG; }
                    memcpy(rs.quals.p + start + got, p, n);
                }
                got += n;
                p = e < end ? e + 1 : end;
            } while (got < len);
            if (want_quals && got < len) { err = "record " + std::to_string(rs.offsets.size() - 1) + ": quality shorter than sequence"; return LEON_E_ARG; }
            while (p < end && *p != '>' && *p != '@') p++;   // the reference scans for the next marker character
        } else if (want_quals) {
            err = "record " + std::to_string(rs.offsets.size() - 1) + " has no quality line (FASTQ expected)";
            return LEON_E_ARG;
        }
    }
    return LEON_OK;
}

int map_file(const char *path, Mapped &m, std::string &err) {
    m.fd = open(path, O_RDONLY);
    struct stat st;
    if (m.fd < 0 || fstat(m.fd, &st) != 0) { err = std::string("cannot open ") + path + ": " + strerror(errno); return LEON_E_IO; }
    m.n = (size_t)st.st_size;
    if (m.n) {
        void *q = mmap(nullptr, m.n, PROT_READ, MAP_PRIVATE, m.fd, 0);
        if (q == MAP_FAILED) { m.n = 0; err = std::string("cannot map ") + path + ": " + strerror(errno); return LEON_E_IO; }
        m.p = (const uint8_t *)q;
        madvise(q, m.n, MADV_SEQUENTIAL);
    }
    return LEON_OK;
}

struct FileReads {
    leon_file_reads pub;
    ReadSet rs;
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
        rc = parse_reads(m, want_quals != 0, want_headers != 0, f->rs, err);
        if (rc) { leon_internal_set_error(("leon_read_file: " + err).c_str()); return rc; }
        memset(&f->pub, 0, sizeof f->pub);
        f->pub.reads.bases = f->rs.bases.p;
        f->pub.reads.offsets = f->rs.offsets.data();
        f->pub.reads.quals = want_quals ? f->rs.quals.p : nullptr;
        f->pub.reads.nreads = f->rs.offsets.size() - 1;
        f->pub.headers = want_headers ? f->rs.headers.data() : nullptr;
        f->pub.header_offsets = want_headers ? f->rs.header_offsets.data() : nullptr;
        f->pub.is_fastq = f->rs.any_qual ? 1 : 0;
        f->pub.pinned = f->rs.bases.pinned ? 1 : 0;
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
        const std::string in(in_path), out = out_path ? std::string(out_path) : in + ".leon";
        // format from the file name, as C/Leon.cpp:392-416
        bool fastq_input;
        if (has_any(in, ".fq", ".fastq")) fastq_input = true;
        else if (has_any(in, ".fa", ".fasta")) fastq_input = false;
        else { leon_internal_set_error("input extension must be one among fasta (.fa, .fasta) or fastq (.fq, .fastq)"); return LEON_E_ARG; }
        const bool noheader = o->noheader || o->seq_only;
        const bool is_fasta = !fastq_input || o->noqual || o->seq_only;     // container "type": decompresses to FASTA

        Mapped m;
        rc = map_file(in_path, m, err);
        if (rc) { leon_internal_set_error(err.c_str()); return rc; }
        ReadSet rs;
        rc = parse_reads(m, !is_fasta, !noheader, rs, err);
        if (rc) { leon_internal_set_error(("leon_compress_file: " + err).c_str()); return rc; }
        const uint64_t nreads = rs.offsets.size() - 1;
        const double t1 = now_ms();

        leon_opts lo;
        memset(&lo, 0, sizeof lo);
        lo.kmer_size = o->kmer_size ? o->kmer_size : 31;
        lo.reads_per_block = o->reads_per_block ? o->reads_per_block : 50000;
        lo.has_qual = is_fasta ? 0 : 1;
        lo.lossless = o->lossless;
        lo.abundance_set = o->abundance_set;
        lo.device = o->device;
        rc = leon_create(&lo, &ctx);
        if (rc) return rc;                                    // message already set by leon_create

        // header stream on host threads while the GPU works on the sequence stream
        int hrc = LEON_OK;
        std::string herr;
        double header_ms = 0;
        std::thread header_thread;
        if (!noheader) {
            header_thread = std::thread([&]() {
                const double a = now_ms();
                hrc = leon_headers_encode(rs.headers.data(), rs.header_offsets.data(), nreads, lo.reads_per_block, o->nthreads, &hs);
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
        rc = leon_compress(ctx, &rd, &res);
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
            stats->input_bytes = m.n;
            struct stat so;
            if (stat(out.c_str(), &so) == 0) stats->output_bytes = (uint64_t)so.st_size;
            stats->header_bytes_in = rs.headers.size();
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
