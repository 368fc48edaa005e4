// Micro-benchmark + self-check of the range coder kernels of entropy.cuh on synthetic symbol records shaped like a
// C2 block (about 10.8 symbols and 3.5-4 output bytes per read).  Development tool, not part of the library:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I leon_b200/csrc -I tools -o tools/bin/rc_bench tools/rc_bench.cu
//   ./rc_bench [streams=92] [reads_per_stream=50000]
// Every variant must reproduce the bytes of a plain host coder that divides (RangeCoder.cpp:106-145).
#include "entropy.cuh"
#include "rc_v2_experiment.cuh"
#include <vector>
#include <random>
#include <chrono>

struct HostModel {
    std::vector<uint32_t> cnt; uint32_t total;
    explicit HostModel(int n) : cnt(n, 1), total(n) {}
    void get(unsigned v, uint32_t &lo, uint32_t &w, uint32_t &tot) {
        lo = 0; for (unsigned u = 0; u < v; u++) lo += cnt[u];
        w = cnt[v]; tot = total; cnt[v]++; total++;
    }
};
static void host_magic(uint32_t d, uint64_t &m, uint32_t &s, uint32_t &inc)
{
    int L = 31 - __builtin_clz(d);
    s = L;
    if ((d & (d - 1)) == 0) { m = ~0ull; inc = 1; return; }
    unsigned __int128 t = (unsigned __int128)1 << (64 + L);
    uint64_t mdown = (uint64_t)(t / d);
    uint64_t r2 = (uint64_t)(t - (unsigned __int128)mdown * d);
    if ((uint64_t)d - r2 < (1ull << L)) { m = mdown + 1; inc = 0; } else { m = mdown; inc = 1; }
}
struct Sym { uint32_t lo, w, tot; };

static void gen_stream(uint64_t seed, int reads, std::vector<Sym> &out)
{
    std::mt19937_64 rng(seed);
    HostModel type(2), size0(256), pos0(256), a0(256), a1(256), a2(256), rev(2), nn(256), ne(256), ep(256), bif(5), bifb(2);
    auto put = [&](HostModel &m, unsigned v) { Sym s; m.get(v, s.lo, s.w, s.tot); out.push_back(s); };
    for (int r = 0; r < reads; r++) {
        put(type, 0); put(size0, 100); put(pos0, rng() % 70);
        uint32_t addr = rng() % 258028;
        put(a0, (addr & 127) | 128); put(a1, ((addr >> 7) & 127) | 128); put(a2, addr >> 14);
        put(rev, rng() & 1); put(nn, 0);
        unsigned nerr = (rng() % 100) < 30 ? 1 + (rng() % 100 < 15) : 0;
        put(ne, nerr);
        for (unsigned e = 0; e < nerr; e++) put(ep, rng() % 69);
        for (unsigned e = 0; e < nerr; e++) put(bif, rng() % 4);
        if (rng() % 100 < 8) put(bifb, rng() & 1);
    }
}

static uint64_t g_stat[6];   // symbols, 0 byte, 1 byte, 2 bytes, >= 3 bytes, clamps
static void host_encode(const std::vector<Sym> &syms, std::vector<uint8_t> &out)
{
    uint64_t low = 0, range = ~0ull;
    for (const Sym &s : syms) {
        range /= s.tot; low += (uint64_t)s.lo * range; range *= s.w;
        {
            uint64_t l = low, r = range; int nb = 0; bool clamp = false;
            while ((l ^ (l + r)) < RC_TOP || r < RC_BOTTOM) {
                if (r < RC_BOTTOM && (l ^ (l + r)) >= RC_TOP) { r = (0 - l) & (RC_BOTTOM - 1); clamp = true; }
                nb++; r <<= 8; l <<= 8;
            }
            g_stat[0]++; if (clamp) g_stat[5]++; else g_stat[1 + (nb > 3 ? 3 : nb)]++;
        }
        while ((low ^ (low + range)) < RC_TOP || range < RC_BOTTOM) {
            if (range < RC_BOTTOM && (low ^ (low + range)) >= RC_TOP) range = (0 - low) & (RC_BOTTOM - 1);
            out.push_back((uint8_t)(low >> 56)); range <<= 8; low <<= 8;
        }
    }
    for (int i = 0; i < 8; i++) { out.push_back((uint8_t)(low >> 56)); low <<= 8; }
}


// ---- version 2 on the HOST: the same LB_HD step functions the kernel runs, same grouping and fallback -------------
struct HostStb { std::vector<uint8_t> *o; void operator()(uint32_t at, uint32_t v) const { if (o->size() <= at) o->resize(at + 1); (*o)[at] = (uint8_t)v; } };
struct HostPut { std::vector<uint8_t> *o; uint32_t &on; void operator()(uint8_t b) { if (o->size() <= on) o->resize(on + 1); (*o)[on++] = b; } };
static uint64_t g_rare_groups = 0, g_groups = 0;
static void host_encode_v2(const std::vector<Sym> &syms, std::vector<uint8_t> &out)
{
    using namespace lbent;
    const size_t n = syms.size();
    std::vector<uint64_t> mp(n + 8, 1); std::vector<uint32_t> ell(n + 8, 0), lo(n + 8, 0), w(n + 8, 1);
    for (size_t i = 0; i < n; i++) { div_magic65(syms[i].tot, mp[i], ell[i]); lo[i] = syms[i].lo; w[i] = syms[i].w; }
    uint64_t low = 0, q = div65(~0ull, mp[0], ell[0]);
    uint32_t on = 0;
    out.clear();
    HostStb stb{&out};
    for (size_t j = 0; j < n; j += 4) {
        bool wide = false;
        for (int i = 0; i <= 4; i++) wide |= ell[j + i] > RC2_MAX_ELL;
        const uint64_t low_g = low, q_g = q; const uint32_t on_g = on;
        uint32_t rmin = 0xffffffffu;
        g_groups++;
        if (!wide) {
            for (int i = 0; i < 4; i++)
                rc_step_fast(low, q, on, rmin, lo[j + i], w[j + i], lo[j + i] + w[j + i], (uint32_t)mp[j + i + 1], (uint32_t)(mp[j + i + 1] >> 32), ell[j + i + 1], stb);
        }
        if (wide || rmin < (1u << 16)) {
            g_rare_groups++;
            low = low_g; q = q_g; on = on_g;
            HostPut put{&out, on};
            for (int i = 0; i < 4; i++) rc_step_gen(low, q, lo[j + i], w[j + i], mp[j + i + 1], ell[j + i + 1], put);
        }
    }
    out.resize(on);
    for (int i = 0; i < 8; i++) { out.push_back((uint8_t)(low >> 56)); low <<= 8; }
}

static int host_selftest(int streams, int reads)
{
    using namespace lbent;
    // div65 == n / d: every d in [2, 2^20] with adversarial n, then random (n, d) up to 2^32 - 1
    std::mt19937_64 rng(7);
    uint64_t bad = 0, checked = 0;
    for (uint32_t d = 1; d < (1u << 20); d++) {
        uint64_t mp; uint32_t ell; div_magic65(d, mp, ell);
        const uint64_t ns[6] = {~0ull, ~0ull - d, (~0ull / d) * d, (~0ull / d) * d - 1, rng(), rng() >> (rng() % 40)};
        for (uint64_t nn : ns) { checked++; if (div65(nn, mp, ell) != nn / d) bad++; }
    }
    for (int i = 0; i < 20000000; i++) {
        uint32_t d = (uint32_t)(rng() >> (32 + rng() % 31)); if (d < 1) d = 1;
        uint64_t nn = rng() >> (rng() % 17);
        uint64_t mp; uint32_t ell; div_magic65(d, mp, ell);
        checked++; if (div65(nn, mp, ell) != nn / d) bad++;
    }
    printf("host: div65 mismatches %llu of %llu\n", (unsigned long long)bad, (unsigned long long)checked);
    int badstreams = 0; uint64_t nsym = 0;
    for (int s = 0; s < streams; s++) {
        std::vector<Sym> syms; gen_stream(1000 + s, reads, syms);
        if (s % 3 == 1) {        // adversarial: huge totals (ell > 23) and tiny alphabets mixed in
            HostModel big(5); big.total = (s % 2) ? 40000000u : 9000000u; big.cnt.assign(5, big.total / 5);
            std::vector<Sym> t2;
            for (size_t i = 0; i < syms.size(); i++) { t2.push_back(syms[i]); if (i % 7 == 0) { Sym y; big.get(rng() % 5, y.lo, y.w, y.tot); t2.push_back(y); } }
            syms.swap(t2);
        }
        std::vector<uint8_t> want, got;
        host_encode(syms, want);
        host_encode_v2(syms, got);
        nsym += syms.size();
        if (want != got) { badstreams++; printf("host: stream %d differs (%zu vs %zu bytes)\n", s, want.size(), got.size()); }
    }
    printf("host: v2 coder on %d streams, %llu symbols: %d mismatching streams; %llu of %llu groups recoded (%.3f %%)\n", streams,
           (unsigned long long)nsym, badstreams, (unsigned long long)g_rare_groups, (unsigned long long)g_groups, 100.0 * g_rare_groups / g_groups);
    return (bad || badstreams) ? 1 : 0;
}

// ---- the host coder of the anchor dictionary stream (lbent::dict_code_slice) against the dividing coder, and its speed ----
// REJECTED EXPERIMENT: next quotient formed for both outcomes of the renormalisation decision (see entropy.cuh)
static void dict_slice_spec(const uint32_t *rec, const uint8_t *sh, uint64_t n, bool first, uint64_t &low, uint64_t &range,
                     std::vector<uint8_t> &out)
{
    const uint64_t TOP = (1ull << 56), BOTTOM = (1ull << 48);
    const size_t pos0 = out.size();
    out.resize(pos0 + n * 2 + 64);          // the common step writes <= 1 byte per symbol; the rare path re-checks
    uint8_t *op = out.data() + pos0;
    const uint8_t *op_end = out.data() + out.size();
    uint64_t j = 0;
    if (first && n) {                       // first symbol: range = 2^64 - 1, range + inc needs 65 bits
        const uint64_t m = ((uint64_t)rec[3] << 32) | rec[2];
        const uint32_t s = sh[0];
        const uint64_t q = (uint64_t)(((unsigned __int128)range * m + ((s & 0x80u) ? m : 0ull)) >> 64) >> (s & 63u);
        low += (uint64_t)rec[0] * q;
        range = q * (uint64_t)rec[1];
        while ((low ^ (low + range)) < TOP || range < BOTTOM) {
            if (range < BOTTOM && (low ^ (low + range)) >= TOP) range = (0 - low) & (BOTTOM - 1);
            *op++ = (uint8_t)(low >> 56);
            range <<= 8;
            low <<= 8;
        }
        j = 1;
    }
    // The chain of a step on a CPU is latency, not issue: multiply-high (4-5 cycles), shift, multiply (3), add, xor, compare,
    // select -- about 15 cycles when the next quotient can only start after the renormalisation decision.  So the quotient of
    // the NEXT symbol is formed for both outcomes of that decision (no byte: range; one byte: range << 8) as soon as the
    // new range exists, beside the computation of low and the decision itself, and the decision only selects one of them:
    // the dependent chain shrinks to multiply, multiply-high, shift, select (~10 cycles).  State carried: (low, q) with
    // q = range / total of the symbol to code.
    auto quotient = [](uint64_t rp, uint64_t m, uint32_t s) { return (uint64_t)(((unsigned __int128)rp * m) >> 64) >> (s & 63u); };
    auto magic = [&](uint64_t i) { return ((uint64_t)rec[4 * i + 3] << 32) | rec[4 * i + 2]; };
    uint64_t q = j < n ? quotient(range + (sh[j] >> 7), magic(j), sh[j]) : 0;
    for (; j < n; j++) {
        const uint32_t lo = rec[4 * j], w = rec[4 * j + 1];
        const bool more = j + 1 < n;
        const uint64_t mn = more ? magic(j + 1) : 0ull;                 // record of the next symbol: independent of the chain
        const uint32_t sn = more ? sh[j + 1] : 0u, inc_next = sn >> 7;
        low += (uint64_t)lo * q;
        range = q * (uint64_t)w;
        const uint64_t qa = quotient(range + inc_next, mn, sn);          // next quotient if no byte leaves ...
        const uint64_t qb = quotient((range << 8) | inc_next, mn, sn);   // ... and if exactly one does
        const uint32_t x = (uint32_t)((low ^ (low + range)) >> 32), rh = (uint32_t)(range >> 32);
        if (__builtin_expect((x < rh ? x : rh) < 65536u, 0)) {           // two bytes or more, or a range clamp: 1.4 % of the symbols
            while ((low ^ (low + range)) < TOP || range < BOTTOM) {
                if (range < BOTTOM && (low ^ (low + range)) >= TOP) range = (0 - low) & (BOTTOM - 1);
                *op++ = (uint8_t)(low >> 56);
                range <<= 8;
                low <<= 8;
            }
            if (op + (n - j) + 32 > op_end) {      // keep one byte per remaining symbol plus a rare step in reserve
                const size_t used = op - out.data();
                out.resize(out.size() * 2 + 64);
                op = out.data() + used;
                op_end = out.data() + out.size();
            }
            q = quotient(range + inc_next, mn, sn);
            continue;
        }
        const bool one = x < (1u << 24);
        *op = (uint8_t)(low >> 56);
        op += one;
        q = one ? qb : qa;
        range = one ? range << 8 : range;
        low = one ? low << 8 : low;
    }
    out.resize(op - out.data());
}



static int dict_selftest(int anchors, int k)
{
    // a dictionary-shaped stream: one Order0Model(5), symbols 0..3, never reset
    std::mt19937_64 rng(11);
    HostModel model(5);
    std::vector<Sym> syms;
    for (int a = 0; a < anchors; a++)
        for (int i = 0; i < k; i++) { Sym y; model.get((unsigned)(rng() % 4), y.lo, y.w, y.tot); syms.push_back(y); }
    std::vector<uint8_t> want;
    host_encode(syms, want);
    std::vector<uint32_t> rec(4 * syms.size()); std::vector<uint8_t> sh(syms.size());
    for (size_t i = 0; i < syms.size(); i++) {
        uint64_t m; uint32_t s, inc; host_magic(syms[i].tot, m, s, inc);
        rec[4 * i] = syms[i].lo; rec[4 * i + 1] = syms[i].w; rec[4 * i + 2] = (uint32_t)m; rec[4 * i + 3] = (uint32_t)(m >> 32);
        sh[i] = (uint8_t)(s | (inc << 7));
    }
    int bad = 0;
    double ns[2] = {0, 0};
    for (int variant = 0; variant < 2; variant++) {
        for (int rep = 0; rep < 3; rep++) {
            std::vector<uint8_t> got;
            uint64_t low = 0, range = ~0ull;
            auto t0 = std::chrono::steady_clock::now();
            size_t at = 0;
            bool first = true;
            while (at < syms.size()) {          // slices of uneven sizes, like the worker's
                size_t nsl = std::min<size_t>(syms.size() - at, rep == 0 ? 1 + rng() % 5000 : (1u << 19));
                if (variant == 0) lbent::dict_code_slice(rec.data() + 4 * at, sh.data() + at, nsl, first, low, range, got);
                else dict_slice_spec(rec.data() + 4 * at, sh.data() + at, nsl, first, low, range, got);
                first = false;
                at += nsl;
            }
            double dt = std::chrono::duration<double, std::nano>(std::chrono::steady_clock::now() - t0).count() / syms.size();
            if (rep == 0 || dt < ns[variant]) ns[variant] = dt;
            for (int x = 0; x < 8; x++) { got.push_back((uint8_t)(low >> 56)); low <<= 8; }
            if (got != want) bad++;
        }
    }
    printf("host: dictionary coder, %zu symbols: library step %.2f ns/symbol, speculative-quotient experiment %.2f ns/symbol, mismatches %d\n", syms.size(), ns[0], ns[1], bad);
    return bad;
}

__global__ void k_magic_check(uint32_t d0, uint32_t n, uint32_t stride, const uint64_t *want_m, const uint8_t *want_s, uint32_t *bad)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t d = d0 + i * stride;
    uint64_t m; uint32_t s, inc;
    lbent::div_magic(d, m, s, inc);
    if (m != want_m[i] || (uint8_t)(s | (inc << 7)) != want_s[i]) atomicAdd(bad, 1u);
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

int main(int argc, char **argv)
{
    if (argc > 1 && !strcmp(argv[1], "--dict")) return dict_selftest(argc > 2 ? atoi(argv[2]) : 300000, argc > 3 ? atoi(argv[3]) : 31);
    if (argc > 1 && !strcmp(argv[1], "--host")) return host_selftest(argc > 2 ? atoi(argv[2]) : 12, argc > 3 ? atoi(argv[3]) : 20000);
    int streams = argc > 1 ? atoi(argv[1]) : 92, reads = argc > 2 ? atoi(argv[2]) : 50000;
    using namespace lbent;
    {   // div_magic on the device == exact host arithmetic: every d in [2, 2^22], then a stride through the 32-bit range
        const uint32_t n = 1u << 22;
        std::vector<uint64_t> wm(n); std::vector<uint8_t> ws(n);
        uint64_t *d_m; uint8_t *d_s; uint32_t *d_bad;
        CK(cudaMalloc(&d_m, n * 8)); CK(cudaMalloc(&d_s, n)); CK(cudaMalloc(&d_bad, 4));
        for (int pass = 0; pass < 2; pass++) {
            uint32_t d0 = 2, stride = pass ? 1021 : 1;
            for (uint32_t i = 0; i < n; i++) { uint32_t sft, inc; host_magic(d0 + i * stride, wm[i], sft, inc); ws[i] = (uint8_t)(sft | (inc << 7)); }
            CK(cudaMemcpy(d_m, wm.data(), n * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_s, ws.data(), n, cudaMemcpyHostToDevice));
            CK(cudaMemset(d_bad, 0, 4));
            k_magic_check<<<n / 256, 256>>>(d0, n, stride, d_m, d_s, d_bad);
            uint32_t bad; CK(cudaMemcpy(&bad, d_bad, 4, cudaMemcpyDeviceToHost));
            printf("div_magic check, stride %u: %u mismatches of %u\n", stride, bad, n);
        }
        cudaFree(d_m); cudaFree(d_s); cudaFree(d_bad);
    }
    std::vector<uint4> rec, rec2; std::vector<uint8_t> sh, sh2; std::vector<StreamDesc> descs; std::vector<std::vector<uint8_t>> want(streams);
    uint64_t base = 0, rare = 0;
    for (int s = 0; s < streams; s++) {
        std::vector<Sym> syms; gen_stream(1000 + s, reads, syms);
        host_encode(syms, want[s]);
        for (const Sym &y : syms) {
            uint64_t m; uint32_t sft, inc; host_magic(y.tot, m, sft, inc);
            rec.push_back(make_uint4(y.lo, y.w, (uint32_t)m, (uint32_t)(m >> 32))); sh.push_back((uint8_t)(sft | (inc << 7)));
            uint64_t mp; uint32_t ell; div_magic65(y.tot, mp, ell);
            rec2.push_back(make_uint4(y.lo, y.w, (uint32_t)mp, (uint32_t)(mp >> 32))); sh2.push_back((uint8_t)ell);
        }
        StreamDesc d; d.rec_base = base; d.nsym = syms.size(); d.out_base = 2 * base + 1024 * s; d.out_cap = 2 * syms.size() + 1024;
        descs.push_back(d); base += syms.size();
    }
    uint64_t nbytes = 0; for (auto &w : want) nbytes += w.size();
    printf("%d streams, %llu symbols (%.1f per read), %llu bytes (%.2f per read)\n", streams, (unsigned long long)base,
           (double)base / streams / reads, (unsigned long long)nbytes, (double)nbytes / streams / reads);
    (void)rare;
    printf("renormalisation per symbol: 0 byte %.4f, 1 byte %.4f, 2 bytes %.4f, >= 3 bytes %.4f, clamp %.5f\n", (double)g_stat[1] / g_stat[0],
           (double)g_stat[2] / g_stat[0], (double)g_stat[3] / g_stat[0], (double)g_stat[4] / g_stat[0], (double)g_stat[5] / g_stat[0]);
    uint4 *d_rec; uint8_t *d_sh, *d_out; StreamDesc *d_desc; uint64_t *d_size; uint32_t *d_ovf;
    uint64_t outcap = 2 * base + 1024ull * streams + 1024;
    CK(cudaMalloc(&d_rec, (base + 64) * 16)); CK(cudaMalloc(&d_sh, base + 64)); CK(cudaMalloc(&d_out, outcap));
    CK(cudaMalloc(&d_desc, streams * sizeof(StreamDesc))); CK(cudaMalloc(&d_size, streams * 8)); CK(cudaMalloc(&d_ovf, 4));
    CK(cudaMemcpy(d_rec, rec.data(), base * 16, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_sh, sh.data(), base, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_desc, descs.data(), streams * sizeof(StreamDesc), cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::vector<uint8_t> got(outcap); std::vector<uint64_t> sizes(streams);
    for (int variant = 1; variant <= 2; variant++) {
        float best = 1e9f;
        CK(cudaMemcpy(d_rec, variant == 1 ? rec.data() : rec2.data(), base * 16, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_sh, variant == 1 ? sh.data() : sh2.data(), base, cudaMemcpyHostToDevice));
        for (int it = 0; it < 4; it++) {
            CK(cudaMemset(d_out, 0xEE, outcap)); CK(cudaMemset(d_ovf, 0, 4));
            CK(cudaEventRecord(e0));
            if (variant == 1) k_range_encode<<<streams, 32>>>(d_rec, d_sh, d_desc, d_out, d_size, 0, d_ovf);
            else k_range_encode2<<<streams, 32>>>(d_rec, d_sh, d_desc, d_out, d_size, 0, d_ovf);
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        CK(cudaMemcpy(got.data(), d_out, outcap, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(sizes.data(), d_size, streams * 8, cudaMemcpyDeviceToHost));
        uint32_t ovf; CK(cudaMemcpy(&ovf, d_ovf, 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (int s = 0; s < streams; s++) {
            if (sizes[s] != want[s].size() || memcmp(got.data() + descs[s].out_base, want[s].data(), want[s].size())) bad++;
        }
        double nsps = best * 1e6 / ((double)base / streams);
        printf("variant %d: %.3f ms  %.1f ns/symbol/stream (%.0f cycles @1.965GHz)  mismatching streams %d  overflow %u\n", variant, best, nsps, nsps * 1.965, bad, ovf);
    }
    return 0;
}
