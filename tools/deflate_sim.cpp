// Development check of leon_b200/csrc/deflate.cuh without a GPU: runs the per-thread phases of k_deflate_chunks on
// the host, thread after thread, and assembles the zlib stream like k_deflate_layout / k_deflate_gather do.
// usage: deflate_sim <text file> <out.zlib>      (verify with python: zlib.decompress(open(out).read()) == text)
// Not part of the library and never used by it.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "../leon_b200/csrc/deflate.cuh"
using namespace lbdeflate;

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: deflate_sim <in> <out>\n"); return 2; }
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    std::vector<uint8_t> text;
    uint8_t buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, f)) > 0) text.insert(text.end(), buf, buf + got);
    fclose(f);
    std::vector<uint8_t> out = {0x78, 0x01};
    std::unique_ptr<ChunkState> Sp(new ChunkState());
    std::unique_ptr<HuffState> Hp(new HuffState());
    ChunkState &S = *Sp;
    HuffState &H = *Hp;
    text.resize(text.size() + 8);                       // the device text is padded too
    const size_t text_size = text.size() - 8;
    uint64_t A = 1, B = 0;
    const size_t nchunks = (text_size + DF_CHUNK - 1) / DF_CHUNK;
    size_t stored_chunks = 0;
    for (size_t c = 0; c < nchunks; c++) {
        const int n = (int)std::min<size_t>(DF_CHUNK, text_size - c * DF_CHUNK);
        const int is_last = c + 1 == nchunks, nact = (n + DF_SLICE - 1) / DF_SLICE;
        // k_deflate_count
        memset(&S, 0xA5, sizeof S);
        memset(&H, 0xA5, sizeof H);
        S.text = text.data() + c * DF_CHUNK;
        S.nmatch = 0;
        for (int w = 0; w < DF_NLIT + 2; w++) S.freq[w] = 0;
        for (int t = 0; t < DF_THREADS; t++) df_summary(S, t, n);
        for (int t = 0; t < DF_THREADS; t++) { CountSink sink{S}; df_slice_tokens(S, t, nact, sink); }
        for (int w = 0; w < DF_NLIT + 2; w++) H.freq[w] = S.freq[w];
        H.nmatch = S.nmatch;
        uint32_t sa, sb;
        df_chunk_sums(S, n, nact, &sa, &sb);
        // k_deflate_codes
        df_build_block_header(H);
        // k_deflate_emit
        memset(&S, 0x5A, sizeof S);
        S.text = text.data() + c * DF_CHUNK;
        S.hdr_bits = H.hdr_bits;
        for (int w = 0; w < DF_NLIT; w++) S.lc[w] = H.lc[w];
        for (int t = 0; t < DF_THREADS; t++) df_summary(S, t, n);
        std::vector<uint32_t> slotv(DF_SLOT / 4, 0xdeadbeefu);         // like the device slot: garbage until zeroed
        uint32_t *slot = slotv.data();
        uint32_t total = 0;
        for (int t = 0; t < DF_THREADS; t++) {
            BitCountSink sink{S, 0u};
            df_slice_tokens(S, t, nact, sink);
            uint32_t mine = sink.bits + (t == nact - 1 ? (S.lc[256] >> 16) : 0u);
            S.bits[t] = total;
            total += mine;
        }
        df_plan(S, n, total);
        if (!S.stored) {
            const int nw = (int)((S.end_bits + 3 + 7 + 32) / 32 + 1);
            for (int w = 0; w < nw; w++) slot[w] = 0;
            for (int w = 0; w < (int)((S.hdr_bits + 31) / 32); w++) slot[w] |= H.hdr_w[w];
            for (int t = 0; t < DF_THREADS; t++) {
                EmitSink sink{S};
                sink.bw.start(slot, S.hdr_bits + S.bits[t]);
                df_slice_tokens(S, t, nact, sink);
                if (t == nact - 1) sink.bw.put(S.lc[256] & 0xffffu, (int)(S.lc[256] >> 16));
                sink.bw.finish();
            }
        } else {
            df_stored_header((uint8_t *)slot, n, is_last);
            for (int t = 0; t < DF_THREADS; t++) df_stored_copy(S, (uint8_t *)slot, t);
            stored_chunks++;
        }
        const uint32_t nbytes = S.stored ? (uint32_t)n + 5u : df_close(S, slot, is_last);
        B = (B + (uint64_t)n % DF_ADLER * A + sb) % DF_ADLER;
        A = (A + sa) % DF_ADLER;
        out.insert(out.end(), (uint8_t *)slot, (uint8_t *)slot + nbytes);
    }
    if (nchunks == 0) { const uint8_t e[5] = {1, 0, 0, 0xff, 0xff}; out.insert(out.end(), e, e + 5); }
    const uint32_t ad = (uint32_t)((B << 16) | A);
    out.push_back((uint8_t)(ad >> 24)); out.push_back((uint8_t)(ad >> 16)); out.push_back((uint8_t)(ad >> 8)); out.push_back((uint8_t)ad);
    f = fopen(argv[2], "wb");
    fwrite(out.data(), 1, out.size(), f);
    fclose(f);
    printf("%zu -> %zu bytes, %zu chunks (%zu stored)\n", text_size, out.size(), nchunks, stored_chunks);
    return 0;
}
