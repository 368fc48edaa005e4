// Command line of the B200 path, option for option the `leon -c` of the reference (C/Leon.cpp:102-140):
//   leon_b200 -file reads.fastq -c [-kmer-size 31] [-abundance N] [-nb-cores N] [-lossless] [-seq-only] [-noheader]
//             [-noqual] [-reads 50000] [-out file.leon] [-device 0] [-verbose 0|1]
//   leon_b200 -file reads.fastq.leon -d [-nb-cores N] [-out file] [-gpu-decode] [-device 0]
// -c writes reads.fastq.leon, which the reference's `leon -d` decodes too; -d reads containers of either tool: the
// sequence stream on host threads, or with -gpu-decode one warp per block on the GPU (k_decode_blocks).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/leon_b200.h"

static void usage() {
    fprintf(stderr,
            "usage: leon_b200 -file <in.fa|in.fasta|in.fq|in.fastq> -c [options]\n"
            "       leon_b200 -file <in.leon> -d [-nb-cores N] [-out FILE] [-gpu-decode] [-device N]\n"
            "  -kmer-size N   k-mer size, 2..63 (default 31)\n"
            "  -abundance N   abundance threshold given explicitly (disables the automatic cutoff)\n"
            "  -nb-cores N    host threads for the header stream and zlib (default: all)\n"
            "  -stream-reads N  feed the GPU batches of N reads (streaming path; default: whole set when it fits the device)\n"
            "  -count-passes N  streaming: number of k-mer counting passes (default: from the free device memory)\n"
            "  -lossless      keep the qualities unchanged\n"
            "  -seq-only      store the sequences only (same as -noheader -noqual)\n"
            "  -noheader      do not store the read names\n"
            "  -noqual        do not store the qualities\n"
            "  -reads N       reads per block (default 50000)\n"
            "  -out FILE      output file (default: <in>.leon)\n"
            "  -device N      CUDA device (default 0)\n"
            "  -gpu-decode    with -d: decode the sequence stream on the GPU (one warp per block)\n"
            "  -verbose N     0: quiet, 1: summary (default 1)\n");
}

int main(int argc, char **argv) {
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);     // the path uses more streams than the default 8 hardware queues
    leon_file_opts o;
    memset(&o, 0, sizeof o);
    std::string in, out;
    bool compress = false, decompress = false, gpu_decode = false;
    int verbose = 1;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto value = [&](const char *name) -> const char * {
            if (i + 1 >= argc) { fprintf(stderr, "option %s needs a value\n", name); exit(2); }
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
        else if (a == "-verbose") verbose = atoi(value("-verbose"));
        else if (a == "-h" || a == "-help" || a == "--help") { usage(); return 0; }
        else { fprintf(stderr, "unknown option %s\n", a.c_str()); usage(); return 2; }
    }
    if (in.empty() || compress == decompress) { usage(); return 2; }
    leon_file_stats s;
    if (decompress) {
        int rc = gpu_decode ? leon_decompress_file_device(in.c_str(), out.empty() ? nullptr : out.c_str(), o.nthreads, o.device, &s)
                            : leon_decompress_file(in.c_str(), out.empty() ? nullptr : out.c_str(), o.nthreads, &s);
        if (rc) {
            fprintf(stderr, "leon_b200: error %d: %s\n", rc, leon_last_error(nullptr));
            return 1;
        }
        if (verbose) {
            printf("reads %llu  bases %llu  blocks %llu  %s, headers %s\n", (unsigned long long)s.nreads, (unsigned long long)s.total_bases,
                   (unsigned long long)s.nblocks, s.is_fasta ? "FASTA" : "FASTQ", s.has_header ? "stored" : "numbered");
            printf("time: container %.1f ms, decode %.1f ms (sequences %.1f ms, headers %.1f ms, qualities %.1f ms, side by side), write %.1f ms, total %.1f ms\n",
                   s.parse_ms, s.compress_ms, s.reserved1[0], s.header_ms, s.reserved1[1], s.write_ms, s.total_ms);
            if (gpu_decode) printf("sequences on the GPU: k_decode_blocks %.1f ms\n", s.reserved1[2]);
        }
        return 0;
    }
    int rc = leon_compress_file(in.c_str(), out.empty() ? nullptr : out.c_str(), &o, &s);
    if (rc) {
        fprintf(stderr, "leon_b200: error %d: %s\n", rc, leon_last_error(nullptr));
        return 1;
    }
    if (verbose) {
        printf("reads %llu  bases %llu  blocks %llu  %s, headers %s\n", (unsigned long long)s.nreads, (unsigned long long)s.total_bases,
               (unsigned long long)s.nblocks, s.is_fasta ? "FASTA" : "FASTQ", s.has_header ? "kept" : "discarded");
        printf("input %llu bytes -> output %llu bytes (rate %.4f)\n", (unsigned long long)s.input_bytes, (unsigned long long)s.output_bytes,
               s.input_bytes ? (double)s.output_bytes / (double)s.input_bytes : 0.0);
        if (s.has_header)
            printf("headers %llu -> %llu bytes%s\n", (unsigned long long)s.header_bytes_in, (unsigned long long)s.header_bytes_out,
                   s.headers_coded_serially ? " (serial coder)" : "");
        printf("time: parse %.1f ms, compress %.1f ms (GPU path %.1f ms, header coder %.1f ms in parallel), write %.1f ms, total %.1f ms\n",
               s.parse_ms, s.compress_ms, s.gpu.total_ms, s.header_ms, s.write_ms, s.total_ms);
    }
    return 0;
}
