// `leon` on the B200 path: the command line of the reference (C/Leon.cpp:102-140, src/main.cpp:37-51), option for option.
//   leon -file reads.fastq -c [-kmer-size 31] [-abundance N] [-nb-cores N] [-lossless] [-seq-only] [-noheader] [-noqual]
//        [-reads 50000] [-verbose 0|1]            writes reads.fastq.leon (reads.fastq.gz -> <dir>/reads.fastq.leon)
//   leon -file reads.fastq.leon -d [-nb-cores N] [-test-file]
// Same messages and exit codes as the reference where scripts look at them: neither / both of -c -d prints "Choose one
// option ..." and exits 0 (C/Leon.cpp:177-180); an unknown extension prints "Input format: unknown ..." and exits 0
// (:427-431); failures print "EXCEPTION: ..." and exit 1 (src/main.cpp:43-49); the summary keeps the reference's keys
// (Input format, cut-off, nb solid kmers, # blocks, # sequences, # nucleotides, Time:, Speed:, Rates; :726-767, :970-1015).
// Additions of this build: -out FILE, -device N, -gpu-decode (sequence stream of -d on the GPU), -stream-reads N and
// -count-passes N (streaming path).  Installed as `leon` and as `leon_b200` (the same program).
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <string>
#include <thread>
#include <sys/stat.h>

#include "../../include/leon_b200.h"

static void usage() {
    fprintf(stderr,
            "usage: leon -file <in.fa|in.fasta|in.fq|in.fastq[.gz]> -c [options]\n"
            "       leon -file <in.leon> -d [-nb-cores N] [-test-file] [-out FILE] [-gpu-decode] [-device N]\n"
            "  -kmer-size N   k-mer size, 2..63 (default 31)\n"
            "  -abundance N   abundance threshold given explicitly (disables the automatic cutoff)\n"
            "  -nb-cores N    host threads for the header stream and zlib (default: all)\n"
            "  -lossless      keep the qualities unchanged\n"
            "  -seq-only      store the sequences only (same as -noheader -noqual)\n"
            "  -noheader      do not store the read names\n"
            "  -noqual        do not store the qualities\n"
            "  -reads N       reads per block (default 50000)\n"
            "  -test-file     with -d: check the decompressed file against the original (same folder)\n"
            "  -verbose N     0: quiet, 1: summary (default 1)\n"
            "  -out FILE      output file (default: the reference's naming)\n"
            "  -device N      CUDA device (default 0)\n"
            "  -gpu-decode    with -d: decode the sequence stream on the GPU (one warp per block)\n"
            "  -host-decode   with -d: decode the sequence stream on host threads (default: chosen from the number of blocks)\n"
            "  -stream-reads N  feed the GPU batches of N reads (streaming path; default: whole set when it fits the device)\n"
            "  -count-passes N  streaming: number of k-mer counting passes (default: from the free device memory)\n");
}

// the reference's info tree (tools/misc/impl/Property: key padded to a 40-column cell, 4 spaces per level)
static void info(int level, const char *key, const char *fmt = nullptr, ...) {
    char val[1024] = "";
    if (fmt) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(val, sizeof val, fmt, ap);
        va_end(ap);
    }
    printf("%*s%-40s%s%s\n", 4 * level, "", key, fmt ? " : " : "", val);
}

static unsigned long long file_size(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0 ? (unsigned long long)st.st_size : 0ull;
}

static std::string dir_of(const std::string &p) {
    const size_t s = p.find_last_of('/');
    return s == std::string::npos ? "." : p.substr(0, s);
}

// -test-file (C/Leon.cpp:2035-2110): names and sequences of the decompressed file against <dir>/<prefix>.fasta|.fastq
static void test_file(const std::string &leon_in, const std::string &decoded, bool fasta) {
    info(1, "Checking decompressed file");
    std::string name = leon_in.substr(leon_in.find_last_of('/') == std::string::npos ? 0 : leon_in.find_last_of('/') + 1);
    const size_t dot = name.find('.');
    if (dot != std::string::npos) name = name.substr(0, dot);
    const std::string original = dir_of(leon_in) + "/" + name + (fasta ? ".fasta" : ".fastq");
    info(2, "Original file", "%s", original.c_str());
    info(2, "New file", "%s", decoded.c_str());
    leon_file_reads *a = nullptr, *b = nullptr;
    if (leon_read_file(original.c_str(), 0, 1, &a) || leon_read_file(decoded.c_str(), 0, 1, &b)) {
        info(1, "cannot read one of the files", "%s", leon_last_error(nullptr));
        if (a) leon_file_reads_free(a);
        return;
    }
    const uint64_t n = a->reads.nreads < b->reads.nreads ? a->reads.nreads : b->reads.nreads;
    bool ok = true;
    for (uint64_t i = 0; i < n && ok; i++) {
        const uint64_t ha = a->header_offsets[i + 1] - a->header_offsets[i], hb = b->header_offsets[i + 1] - b->header_offsets[i];
        const uint64_t la = a->reads.offsets[i + 1] - a->reads.offsets[i], lb = b->reads.offsets[i + 1] - b->reads.offsets[i];
        if (ha != hb || memcmp(a->headers + a->header_offsets[i], b->headers + b->header_offsets[i], ha)) {
            info(1, "Sequence with a different header", "%llu", (unsigned long long)i);
            ok = false;
        } else if (la != lb || memcmp(a->reads.bases + a->reads.offsets[i], b->reads.bases + b->reads.offsets[i], la)) {
            info(1, "Sequence with a different DNA", "%llu", (unsigned long long)i);
            ok = false;
        }
    }
    if (ok) {
        if (a->reads.nreads == b->reads.nreads) info(1, "OK");
        else if (b->reads.nreads < a->reads.nreads) info(1, "Decompressed file end but not the original file");
        else info(1, "Original file end but not the decomrpessed file");
    }
    leon_file_reads_free(a);
    leon_file_reads_free(b);
}

int main(int argc, char **argv) {
    // CUDA_DEVICE_MAX_CONNECTIONS is left at its default here: 32 hardware queues take 25 ms off the GPU path of a run but add
    // 0.5-1 s to the creation of the CUDA context, which a one-shot process pays every time (measured); long-lived
    // hosts (the Python package) set 32.
    const auto t_start = std::chrono::steady_clock::now();
    leon_file_opts o;
    memset(&o, 0, sizeof o);
    std::string in, out;
    bool compress = false, decompress = false, gpu_decode = false, host_decode = false, test = false;
    int verbose = 1;
    if (argc > 1 && (!strcmp(argv[1], "-version") || !strcmp(argv[1], "-v"))) {     // src/main.cpp:38-41
        printf("* Leon on the B200 path (leon_b200, ABI %d); container format of Leon 1.1.0\n", leon_b200_abi_version());
        return EXIT_FAILURE;
    }
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto value = [&](const char *name) -> const char * {
            if (i + 1 >= argc) { fprintf(stderr, "ERROR: option %s needs a value\n", name); usage(); exit(EXIT_FAILURE); }
            return argv[++i];
        };
        if (a == "-file") in = value("-file");
        else if (a == "-c") compress = true;
        else if (a == "-d") decompress = true;
        else if (a == "-kmer-size") o.kmer_size = atoi(value("-kmer-size"));
        else if (a == "-abundance") { value("-abundance"); o.abundance_set = 1; }   // only its presence matters
        else if (a == "-nb-cores") o.nthreads = atoi(value("-nb-cores"));
        else if (a == "-stream-reads") o.stream_reads = atoi(value("-stream-reads"));
        else if (a == "-count-passes") o.count_passes = atoi(value("-count-passes"));
        else if (a == "-lossless") o.lossless = 1;
        else if (a == "-seq-only") o.seq_only = 1;
        else if (a == "-noheader") o.noheader = 1;
        else if (a == "-noqual") o.noqual = 1;
        else if (a == "-reads") o.reads_per_block = atoi(value("-reads"));
        else if (a == "-out") out = value("-out");
        else if (a == "-device") o.device = atoi(value("-device"));
        else if (a == "-gpu-decode") gpu_decode = true;
        else if (a == "-host-decode") host_decode = true;
        else if (a == "-test-file") test = true;
        else if (a == "-verbose") verbose = atoi(value("-verbose"));
        else if (a == "-h" || a == "-help" || a == "--help") { usage(); return EXIT_SUCCESS; }
        else { fprintf(stderr, "ERROR: Unknown parameter '%s'\n", a.c_str()); usage(); return EXIT_SUCCESS; }   // Tool::run: option failure is reported, not an error status
    }
    if (in.empty()) { fprintf(stderr, "ERROR: Option '-file' is mandatory\n"); usage(); return EXIT_SUCCESS; }
    if (compress == decompress) {                                    // C/Leon.cpp:177-180
        printf("Choose one option among -c (compress) or -d (decompress)\n\n");
        return EXIT_SUCCESS;
    }
    auto seconds = [&]() { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count(); };
    leon_file_stats s;
    if (decompress) {
        bool auto_choice = false;
        if (!gpu_decode && !host_decode) {
            // Default of -d, from the measured cost of the two decoders (profiles/r1g_decode_bench_*.json, B200 + 16
            // host cores): the device kernel runs one warp per block, all blocks side by side, ~50 us per read of the LARGEST block
            // whatever the number of blocks, plus ~1 s to bring up a CUDA context in a fresh process; the host threads need ~5.6 us
            // per read and thread.  92 blocks of 50000 reads: host (1.6 s vs 3.5 s); 920 blocks of 5000: device (0.3 + 1 s vs 1.6 s).
            leon_container *ct = nullptr;
            if (leon_container_open(in.c_str(), &ct) == LEON_OK) {
                const uint8_t *bs = nullptr;
                uint64_t nbytes = 0;
                if (leon_container_dataset(ct, "leon/dna/blocksizes", &bs, &nbytes) == LEON_OK && nbytes >= 16) {
                    const uint64_t nblocks = nbytes / 16;
                    uint64_t total = 0, largest = 0;
                    for (uint64_t b = 0; b < nblocks; b++) {
                        uint64_t n;
                        memcpy(&n, bs + 16 * b + 8, 8);
                        total += n;
                        if (n > largest) largest = n;
                    }
                    unsigned threads = o.nthreads > 0 ? (unsigned)o.nthreads : std::max(1u, std::thread::hardware_concurrency());
                    const double t_dev = 1.0 + 50e-6 * (double)largest * (double)((nblocks + 4095) / 4096);
                    const double t_host = 5.6e-6 * (double)total / threads;
                    gpu_decode = t_dev < 0.8 * t_host;
                    auto_choice = gpu_decode;
                }
                leon_container_close(ct);
            }
        }
        int rc = gpu_decode ? leon_decompress_file_device(in.c_str(), out.empty() ? nullptr : out.c_str(), o.nthreads, o.device, &s)
                            : leon_decompress_file(in.c_str(), out.empty() ? nullptr : out.c_str(), o.nthreads, &s);
        if (rc == LEON_E_CUDA && auto_choice) {     // no usable device: the automatic choice (only) goes back to the host threads
            gpu_decode = false;
            rc = leon_decompress_file(in.c_str(), out.empty() ? nullptr : out.c_str(), o.nthreads, &s);
        }
        if (rc) {
            fprintf(stderr, "EXCEPTION: %s\n", leon_last_error(nullptr));
            return EXIT_FAILURE;
        }
        std::string outname = out;
        if (outname.empty()) {         // <dir>/<name minus its two last suffixes>.fasta.d | .fastq.d (C/Leon.cpp:1357-1448)
            std::string name = in.substr(in.find_last_of('/') == std::string::npos ? 0 : in.find_last_of('/') + 1);
            for (int cut = 0; cut < 2; cut++) {
                const size_t d = name.find_last_of('.');
                if (d != std::string::npos) name = name.substr(0, d);
            }
            outname = dir_of(in) + "/" + name + (s.is_fasta ? ".fasta.d" : ".fastq.d");
        }
        if (verbose) {
            info(0, "Input format", "%s", s.is_fasta ? "FastA" : "FastQ");
            info(0, "End decompression");
            info(1, "# blocks", "%llu", (unsigned long long)s.nblocks);
            info(1, "# sequences", "%llu", (unsigned long long)s.nreads);
            info(1, "# nucleotides", "%llu", (unsigned long long)s.total_bases);
            info(1, "Output filename", "%s", outname.c_str());
            info(1, "Time", "%.2f seconds", seconds());
            info(1, "Speed", "%.2f Mo/seconds", file_size(outname) / 1000000.0 / seconds());
            info(1, "Stages (ms)", "container %.1f, decode %.1f (sequences %.1f, headers %.1f, qualities %.1f side by side), write %.1f",
                 s.parse_ms, s.compress_ms, s.reserved1[0], s.header_ms, s.reserved1[1], s.write_ms);
            if (gpu_decode) info(1, "Sequences on the GPU", "k_decode_blocks %.1f ms", s.reserved1[2]);
        }
        if (test) test_file(in, outname, s.is_fasta != 0);
        return EXIT_SUCCESS;
    }
    // format from the file name, as C/Leon.cpp:402-431: an unknown extension is reported and the run ends without error
    const bool fq = in.find(".fq") != std::string::npos || in.find(".fastq") != std::string::npos;
    const bool fa = in.find(".fa") != std::string::npos || in.find(".fasta") != std::string::npos;
    if (!fq && !fa) {
        info(0, "Input format: unknown. Input extension must be one among fasta (.fa, .fasta) or fastq (.fq, .fastq)");
        return EXIT_SUCCESS;
    }
    int rc = leon_compress_file(in.c_str(), out.empty() ? nullptr : out.c_str(), &o, &s);
    if (rc) {
        fprintf(stderr, "EXCEPTION: %s\n", leon_last_error(nullptr));
        return EXIT_FAILURE;
    }
    if (verbose) {
        std::string outname = out;
        if (outname.empty()) {
            outname = in + ".leon";
            const size_t dot = in.find_last_of('.');
            if (dot != std::string::npos && in.substr(dot + 1).find("gz") != std::string::npos) {
                std::string name = in.substr(in.find_last_of('/') == std::string::npos ? 0 : in.find_last_of('/') + 1);
                name = name.substr(0, name.find_last_of('.'));
                outname = dir_of(in) + "/" + name + ".leon";
            }
        }
        info(0, fq ? "Input format: FastQ" : "Input format: FastA");
        if (fq && !o.seq_only && !o.noqual)
            info(0, o.lossless ? "Quality compression: LOSSLESS mode" : "Quality compression: lossy mode (use '-lossless' for lossless compression)");
        if (s.has_header) {
            info(0, "End Header compression");
            info(1, "# blocks", "%llu", (unsigned long long)s.nblocks);
            info(1, "headers size", "%llu", (unsigned long long)s.header_bytes_in);
            info(1, "headers compressed size", "%llu", (unsigned long long)s.header_bytes_out);
            info(1, "compression rate", "%.4f", s.header_bytes_in ? (double)s.header_bytes_out / (double)s.header_bytes_in : 0.0);
        }
        info(0, "Abundance threshold");
        info(1, s.abundance_auto ? "cut-off (auto)" : "cut-off", "%d", s.abundance_auto ? s.cutoff : 2);
        info(1, "nb solid kmers", "%llu", (unsigned long long)s.nb_solid_kmers);
        info(0, "End Sequence compression");
        info(1, "# blocks", "%llu", (unsigned long long)s.nblocks);
        info(1, "# sequences", "%llu", (unsigned long long)s.nreads);
        info(1, "# nucleotides", "%llu", (unsigned long long)s.total_bases);
        info(1, "Compression rates");
        const double nt = s.total_bases ? (double)s.total_bases : 1.0;
        info(2, "overall", "%.4f (%llu)", (double)s.dna_bytes_out / nt, (unsigned long long)s.dna_bytes_out);
        info(2, "Bloom", "%.2f (%llu)", 100.0 * s.bloom_bytes / (s.dna_bytes_out ? (double)s.dna_bytes_out : 1.0), (unsigned long long)s.bloom_bytes);
        info(2, "Anchors dict", "%.2f (%llu) (%llu entries)", 100.0 * s.anchors_dict_bytes / (s.dna_bytes_out ? (double)s.dna_bytes_out : 1.0),
             (unsigned long long)s.anchors_dict_bytes, (unsigned long long)s.anchor_count);
        info(2, "Reads", "%.2f (%llu)", 100.0 * s.seq_bytes_out / (s.dna_bytes_out ? (double)s.dna_bytes_out : 1.0), (unsigned long long)s.seq_bytes_out);
        info(1, "Reads per anchor", "%llu", (unsigned long long)(s.anchor_count ? s.nreads / s.anchor_count : 0));
        info(1, "Read without anchor", "%.2f", s.nreads ? 100.0 * (double)s.reads_without_anchor / (double)s.nreads : 0.0);
        info(0, "End compression");
        info(1, "Input file");
        info(2, "name", "%s", in.c_str());
        info(2, "size", "%llu bytes (%llu Mb)", (unsigned long long)s.input_bytes, (unsigned long long)s.input_bytes / 1024ull / 1024ull);
        info(1, "Output file");
        info(2, "name", "%s", outname.c_str());
        info(2, "size", "%llu bytes (%llu Mb)", (unsigned long long)s.output_bytes, (unsigned long long)s.output_bytes / 1024ull / 1024ull);
        info(1, "Compression");
        info(2, "Time:", "%.2f seconds", seconds());
        info(2, "Speed:", "%.2f Mb/seconds", s.input_bytes / 1000000.0 / seconds());
        info(2, "Rates");
        info(3, "overall", "%.4f (%.4f)", s.input_bytes ? (double)s.output_bytes / (double)s.input_bytes : 0.0,
             s.output_bytes ? (double)s.input_bytes / (double)s.output_bytes : 0.0);
        if (s.has_header)
            info(3, "header only", "%.4f (%.4f)", s.header_bytes_in ? (double)s.header_bytes_out / (double)s.header_bytes_in : 0.0,
                 s.header_bytes_out ? (double)s.header_bytes_in / (double)s.header_bytes_out : 0.0);
        else
            info(3, "header completely discarded in '-seq-only' mode");
        info(3, "Sequence only", "%.4f (%.4f)", (double)s.dna_bytes_out / nt, nt / (s.dna_bytes_out ? (double)s.dna_bytes_out : 1.0));
        if (!s.is_fasta)
            info(3, "Quality only", "%.4f (%.4f) [%s mode]", (double)s.qual_bytes_out / nt, nt / (s.qual_bytes_out ? (double)s.qual_bytes_out : 1.0),
                 o.lossless ? "lossless" : "lossy");
        info(1, "Stages (ms)", "parse %.1f, compress %.1f (GPU path %.1f, header coder %.1f in parallel), write %.1f, total %.1f%s",
             s.parse_ms, s.compress_ms, s.gpu.total_ms, s.header_ms, s.write_ms, s.total_ms, s.headers_coded_serially ? " (headers: serial coder)" : "");
    }
    return EXIT_SUCCESS;
}
