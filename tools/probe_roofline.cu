// Roofline of the Bloom probes: what the memory system delivers for the access pattern of
// BloomNeighborCoherent::contains4 -- 7 reads inside one 4109-bit (514-byte) window at a random position of the filter --
// and for plain random 32-byte sectors, on filters that are L2-resident (C2: 6.9 MB) or not (C4: 150 MB, C5: 4.5 GB).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/probe_roofline tools/probe_roofline.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x)
{
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

// how a word is loaded: 0 plain read-only load, 1 / 2 / 3 ld.global.nc with an L2 cache hint (evict_last / evict_normal /
// evict_first policy), 4 ld.global.cg (L2 only)
template <int MODE> __device__ __forceinline__ uint64_t load2(const uint32_t *p, uint64_t pol)
{
    uint32_t a, b;
    if (MODE >= 1 && MODE <= 3) {
        asm("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(a) : "l"(p), "l"(pol));
        asm("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(b) : "l"(p + 1), "l"(pol));
    } else if (MODE == 4) {
        asm("ld.global.cg.u32 %0, [%1];" : "=r"(a) : "l"(p));
        asm("ld.global.cg.u32 %0, [%1];" : "=r"(b) : "l"(p + 1));
    } else { a = __ldg(p); b = __ldg(p + 1); }
    return (uint64_t)a ^ ((uint64_t)b << 32);
}

// every thread: `per` independent items; an item = NP reads of 8 bytes at (window base + offset_j), offsets < span bytes
template <int NP, int MODE>
__global__ void __launch_bounds__(256) k_probe(const uint8_t *__restrict__ buf, uint64_t nbytes, uint32_t span, int per, uint64_t *sink)
{
    uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t acc = 0, pol = 0;
    if (MODE == 1) asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    if (MODE == 2) asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    if (MODE == 3) asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    for (int it = 0; it < per; it++) {
        uint64_t h = mix(tid * 1315423911ull + it);
        uint64_t base = (h % (nbytes - span - 8)) & ~3ull;
        uint64_t h2 = mix(h);
#pragma unroll
        for (int j = 0; j < NP; j++) {
            uint32_t off = NP == 1 ? 0u : (uint32_t)((h2 >> (9 * j)) & 511u) % span;
            const uint32_t *p = (const uint32_t *)(buf + ((base + off) & ~3ull));
            acc += load2<MODE>(p, pol);
        }
    }
    if (acc == 0x1234567ull) *sink = acc;
}

int main()
{
    const uint64_t sizes[] = {6900000ull, 150000000ull, 1000000000ull, 4500000000ull};
    uint8_t *buf; uint64_t *sink;
    if (cudaMalloc(&buf, sizes[3] + 4096) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMalloc(&sink, 8);
    cudaMemset(buf, 0x5a, sizes[3] + 4096);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * 16, per = 64;
    const double items = (double)grid * 256 * per;
    printf("%-12s %-44s %10s %12s %12s\n", "filter MB", "pattern, load", "ms", "G items/s", "G reads/s");
    const char *mname[] = {"plain", "hint evict_last", "hint evict_normal", "hint evict_first", "ld.cg"};
    for (uint64_t sz : sizes) {
        for (int mode = 0; mode < 5; mode++) {
            for (int pat = 0; pat < 2; pat++) {
                float best = 1e9f;
                for (int rep = 0; rep < 4; rep++) {
                    cudaEventRecord(e0);
#define RUN(M) do { if (pat == 0) k_probe<1, M><<<grid, 256>>>(buf, sz, 8, per * 7, sink); else k_probe<7, M><<<grid, 256>>>(buf, sz, 514, per, sink); } while (0)
                    switch (mode) { case 0: RUN(0); break; case 1: RUN(1); break; case 2: RUN(2); break; case 3: RUN(3); break; default: RUN(4); }
                    cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
                }
                double n_items = pat == 0 ? items * 7 : items;
                double reads = items * 7;
                char name[96];
                snprintf(name, sizeof name, "%s, %s", pat == 0 ? "1 random sector" : "7 reads / 514 B window", mname[mode]);
                printf("%-12.1f %-44s %10.3f %12.2f %12.2f\n", sz / 1e6, name, best, n_items / best / 1e6, reads / best / 1e6);
            }
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
