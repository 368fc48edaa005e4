// File-level decompression: what `leon -file X.leon -d` does (C/Leon.cpp:1311-1470 executeDecompression, :1510-1580
// setup, :1697-1860 startDecompressionAllStreams): open the container, decode the sequence stream (decode.cu), the
// header stream (headers.cpp) and inflate the quality blocks, then interleave the three streams block by block into
// <dir>/<prefix>.fasta.d or .fastq.d exactly like the reference's writer loop (numbered pseudo-headers "> n" in
// header-less files, one extra numbered line at the end of every block included).
#include <zlib.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/leon_b200.h"
#include "decoded.h"

extern "C" void leon_internal_set_error(const char *msg);

namespace {

struct Fail {
    int code;
    std::string msg;
};

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Ds { const uint8_t *p = nullptr; uint64_t n = 0; };
Ds dataset(leon_container *c, const std::string &path) {
    Ds d;
    if (leon_container_dataset(c, path.c_str(), &d.p, &d.n) != LEON_OK) throw Fail{LEON_E_IO, leon_last_error(nullptr)};
    return d;
}
std::string attr(leon_container *c, const char *path, const char *name) {
    const char *v = nullptr;
    if (leon_container_attr(c, path, name, &v) != LEON_OK) throw Fail{LEON_E_IO, leon_last_error(nullptr)};
    return v;
}
template <typename T> T scalar(leon_container *c, const char *path) {
    Ds d = dataset(c, path);
    if (d.n < sizeof(T)) throw Fail{LEON_E_IO, std::string("dataset ") + path + " is too short"};
    T v;
    memcpy(&v, d.p, sizeof(T));
    return v;
}

bool inflate_block(const uint8_t *z, uint64_t n, std::string &out) {
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (inflateInit(&zs) != Z_OK) return false;
    zs.next_in = (Bytef *)z;
    zs.avail_in = (uInt)n;
    std::vector<char> buf(1 << 20);
    int rc;
    do {
        zs.next_out = (Bytef *)buf.data();
        zs.avail_out = (uInt)buf.size();
        rc = inflate(&zs, Z_NO_FLUSH);
        if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&zs); return false; }
        out.append(buf.data(), buf.size() - zs.avail_out);
    } while (rc != Z_STREAM_END);
    inflateEnd(&zs);
    return true;
}

// next line of a '\n'-separated text (std::getline on the decoders' buffers in the reference)
struct Joiner {                                          // joins on every exit path, exceptions included
    std::thread &t;
    ~Joiner() { if (t.joinable()) t.join(); }
};

inline bool next_line(const std::string &s, size_t &pos, const char **line, size_t *len) {
    if (pos >= s.size()) return false;
    const char *p = s.data() + pos;
    const char *e = (const char *)memchr(p, '\n', s.size() - pos);
    const size_t n = e ? (size_t)(e - p) : s.size() - pos;
    *line = p;
    *len = n;
    pos += n + (e ? 1 : 0);
    return true;
}

}  // namespace

// device < 0: the sequence stream on host threads (leon_decode_dna); device >= 0: on that GPU (leon_decode_dna_device)
static int decompress_file_impl(const char *in_path, const char *out_path, int32_t nthreads, int32_t device, leon_file_stats *stats) {
    leon_container *c = nullptr;
    leon_decoded *dna = nullptr, *names = nullptr;
    int rc = LEON_OK;
    std::string err;
    try {
        if (!in_path) throw Fail{LEON_E_ARG, "leon_decompress_file: null argument"};
        const double t0 = now_ms();
        if (leon_container_open(in_path, &c) != LEON_OK) throw Fail{LEON_E_IO, leon_last_error(nullptr)};
        const bool fasta = attr(c, "metadata/infobyte", "type") == "fasta";
        const bool has_header = attr(c, "metadata/infobyte", "header") == "true";
        const uint64_t ksize = scalar<uint64_t>(c, "metadata/kmerSize");

        // output name: <dir>/<name without ".leon" and without its last extension>.fasta.d / .fastq.d (C/Leon.cpp:1357-1448)
        std::string out;
        if (out_path) out = out_path;
        else {
            std::string in(in_path);
            const size_t slash = in.find_last_of('/');
            std::string dir = slash == std::string::npos ? "." : in.substr(0, slash), base = slash == std::string::npos ? in : in.substr(slash + 1);
            for (int i = 0; i < 2; i++) { const size_t dot = base.find_last_of('.'); if (dot != std::string::npos) base.erase(dot); }
            out = dir + "/" + base + (fasta ? ".fasta.d" : ".fastq.d");
        }

        // sequence stream
        leon_dna_stream ds;
        memset(&ds, 0, sizeof ds);
        ds.kmer_size = (int32_t)ksize;
        ds.max_read_size = leon_container_has(c, "metadata/maxSequenceSize") ? scalar<int32_t>(c, "metadata/maxSequenceSize") : -1;
        const Ds bloom = dataset(c, "leon/bloom");
        std::vector<uint8_t> bloom_padded(bloom.n + 16, 0);              // the filter is read 8 bytes at a time
        memcpy(bloom_padded.data(), bloom.p, bloom.n);
        ds.bloom = bloom_padded.data(); ds.bloom_nbytes = bloom_padded.size();
        ds.bloom_bits = std::stoull(attr(c, "leon/bloom", "size"));
        const Ds dict = dataset(c, "leon/anchors/anchorsDict");
        ds.anchors_dict = dict.p; ds.anchors_dict_nbytes = dict.n;
        ds.anchor_count = scalar<uint32_t>(c, "leon/anchors/anchorAdress");
        const Ds bs = dataset(c, "leon/dna/blocksizes");
        const uint64_t nblocks = bs.n / 16;
        std::vector<uint64_t> blocksizes(2 * nblocks);
        memcpy(blocksizes.data(), bs.p, nblocks * 16);
        std::vector<uint8_t> dna_all;
        for (uint64_t b = 0; b < nblocks; b++) {
            if (!blocksizes[2 * b]) continue;
            const Ds d = dataset(c, "leon/dna/dna_" + std::to_string(b));
            if (d.n < blocksizes[2 * b]) throw Fail{LEON_E_IO, "dna_" + std::to_string(b) + " is shorter than its block size"};
            dna_all.insert(dna_all.end(), d.p, d.p + blocksizes[2 * b]);
        }
        ds.nblocks = nblocks; ds.blocksizes = blocksizes.data(); ds.dna = dna_all.data(); ds.dna_nbytes = dna_all.size();

        // the three streams side by side
        int hrc = LEON_OK, qrc = LEON_OK;
        std::string herr;
        double dna_ms = 0, header_ms = 0, qual_ms = 0;
        std::vector<uint64_t> hbs;
        std::vector<uint8_t> hdata;
        Ds first;
        std::vector<std::string> quals(fasta ? 0 : nblocks);
        std::vector<Ds> zq(fasta ? 0 : nblocks);
        std::thread header_thread, qual_thread;                // declared after everything they touch: joined first
        Joiner join_header{header_thread}, join_qual{qual_thread};
        if (has_header) {
            first = dataset(c, "leon/header/firstheader");
            const uint64_t first_size = scalar<uint64_t>(c, "leon/header/firstheadersize");
            if (first.n < first_size) throw Fail{LEON_E_IO, "firstheader is shorter than firstheadersize"};
            first.n = first_size;
            const Ds hb = dataset(c, "leon/header/blocksizes");
            hbs.resize(hb.n / 8);
            memcpy(hbs.data(), hb.p, hbs.size() * 8);
            for (uint64_t b = 0; b < hbs.size() / 2; b++) {
                if (!hbs[2 * b]) continue;
                const Ds d = dataset(c, "leon/header/header_" + std::to_string(b));
                if (d.n < hbs[2 * b]) throw Fail{LEON_E_IO, "header_" + std::to_string(b) + " is shorter than its block size"};
                hdata.insert(hdata.end(), d.p, d.p + hbs[2 * b]);
            }
            header_thread = std::thread([&]() {
                const double a = now_ms();
                leon_header_stream hs;
                memset(&hs, 0, sizeof hs);
                hs.first_header = first.p; hs.first_header_nbytes = first.n;
                hs.nblocks = hbs.size() / 2; hs.blocksizes = hbs.data(); hs.data = hdata.data(); hs.data_nbytes = hdata.size();
                hrc = leon_headers_decode(&hs, nthreads, &names);
                if (hrc) herr = leon_last_error(nullptr);
                header_ms = now_ms() - a;
            });
        }
        if (!fasta) {
            for (uint64_t b = 0; b < nblocks; b++) zq[b] = dataset(c, "leon/qual/qual_" + std::to_string(b));
            qual_thread = std::thread([&]() {
                const double a = now_ms();
                std::atomic<uint64_t> next(0);
                std::atomic<int> bad(0);
                auto work = [&]() {
                    try {                                       // bad_alloc / length_error in a pool thread would be std::terminate
                        for (;;) {
                            const uint64_t b = next.fetch_add(1);
                            if (b >= nblocks) return;
                            if (!inflate_block(zq[b].p, zq[b].n, quals[b])) bad.store(1);
                        }
                    } catch (const std::exception &) { bad.store(1); }
                };
                unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::max(1u, std::thread::hardware_concurrency());
                nt = (unsigned)std::min<uint64_t>(nt, std::max<uint64_t>(nblocks, 1));
                std::vector<std::thread> pool;
                for (unsigned i = 1; i < nt; i++) pool.emplace_back(work);
                work();
                for (auto &t : pool) t.join();
                if (bad.load()) qrc = LEON_E_IO;
                qual_ms = now_ms() - a;
            });
        }
        const double t1 = now_ms();
        float dna_kernel_ms = 0;
        const int drc = device >= 0 ? leon_decode_dna_device(&ds, device, &dna, &dna_kernel_ms) : leon_decode_dna(&ds, nthreads, &dna);
        std::string derr = drc ? leon_last_error(nullptr) : "";
        dna_ms = now_ms() - t1;
        if (header_thread.joinable()) header_thread.join();
        if (qual_thread.joinable()) qual_thread.join();
        if (drc) throw Fail{drc, derr};
        if (hrc) throw Fail{hrc, herr};
        if (qrc) throw Fail{qrc, "leon_decompress_file: a quality block does not inflate"};
        const double t2 = now_ms();

        // writer loop of C/Leon.cpp:1745-1822, block by block
        FILE *f = fopen(out.c_str(), "wb");
        if (!f) throw Fail{LEON_E_IO, "cannot create " + out + ": " + strerror(errno)};
        std::vector<char> iobuf(1 << 22);
        setvbuf(f, iobuf.data(), _IOFBF, iobuf.size());
        uint64_t nreads = 0, nbases = 0;
        std::string block_out;
        bool ok = true;
        for (uint64_t b = 0; b < nblocks && ok; b++) {
            const std::string &sd = dna->blocks[b];
            static const std::string empty;
            const std::string &sh = has_header && b < names->blocks.size() ? names->blocks[b] : empty;
            const std::string &sq = fasta ? empty : quals[b];
            size_t pd = 0, ph = 0, pq = 0;
            uint64_t readid = 0;
            block_out.clear();
            bool reading = true;
            while (reading) {
                const char *line;
                size_t len;
                if (has_header) {
                    if (next_line(sh, ph, &line, &len)) { block_out += fasta ? '>' : '@'; block_out.append(line, len); block_out += '\n'; }
                    else reading = false;
                } else {
                    block_out += fasta ? "> " : "@ ";
                    block_out += std::to_string(readid++);
                    block_out += '\n';
                }
                if (next_line(sd, pd, &line, &len)) { block_out.append(line, len); block_out += '\n'; nreads++; nbases += len; }
                else reading = false;
                if (!fasta) {
                    if (next_line(sq, pq, &line, &len)) { block_out += "+\n"; block_out.append(line, len); block_out += '\n'; }
                    else reading = false;
                }
            }
            ok = fwrite(block_out.data(), 1, block_out.size(), f) == block_out.size();
        }
        if (fclose(f) != 0) ok = false;
        if (!ok) throw Fail{LEON_E_IO, "write to " + out + " failed"};
        const double t3 = now_ms();
        if (stats) {
            memset(stats, 0, sizeof *stats);
            stats->nreads = nreads;
            stats->total_bases = nbases;
            stats->nblocks = nblocks;
            stats->is_fasta = fasta;
            stats->has_header = has_header;
            stats->parse_ms = (float)(t1 - t0);          // container open + dataset assembly
            stats->compress_ms = (float)(t2 - t1);       // the three decoders side by side
            stats->header_ms = (float)header_ms;
            stats->write_ms = (float)(t3 - t2);
            stats->total_ms = (float)(t3 - t0);
            stats->reserved1[0] = (float)dna_ms;
            stats->reserved1[1] = (float)qual_ms;
            stats->reserved1[2] = dna_kernel_ms;         // device time of k_decode_blocks (0 on host threads)
        }
    } catch (const Fail &fl) {
        rc = fl.code;
        err = fl.msg;
    } catch (const std::bad_alloc &) {
        rc = LEON_E_NOMEM;
        err = "leon_decompress_file: out of host memory";
    } catch (const std::exception &e) {
        rc = LEON_E_IO;
        err = e.what();
    }
    if (dna) leon_decoded_free(dna);
    if (names) leon_decoded_free(names);
    if (c) leon_container_close(c);
    if (rc) leon_internal_set_error(err.c_str());
    return rc;
}

extern "C" int leon_decompress_file(const char *in_path, const char *out_path, int32_t nthreads, leon_file_stats *stats) {
    return decompress_file_impl(in_path, out_path, nthreads, -1, stats);
}

extern "C" int leon_decompress_file_device(const char *in_path, const char *out_path, int32_t nthreads, int32_t device, leon_file_stats *stats) {
    if (device < 0) { leon_internal_set_error("leon_decompress_file_device: device must be >= 0"); return LEON_E_ARG; }
    return decompress_file_impl(in_path, out_path, nthreads, device, stats);
}
