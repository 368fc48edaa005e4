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

// every thread: `per` independent items; an item = NP reads of 8 bytes at (window base + offset_j), offsets < span bytes
template <int NP>
__global__ void __launch_bounds__(256) k_probe(const uint8_t *__restrict__ buf, uint64_t nbytes, uint32_t span, int per, uint64_t *sink)
{
    uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t acc = 0;
    for (int it = 0; it < per; it++) {
        uint64_t h = mix(tid * 1315423911ull + it);
        uint64_t base = (h % (nbytes - span - 8)) & ~3ull;
        uint64_t h2 = mix(h);
#pragma unroll
        for (int j = 0; j < NP; j++) {
            uint32_t off = NP == 1 ? 0u : (uint32_t)((h2 >> (9 * j)) & 511u) % span;
            const uint32_t *p = (const uint32_t *)(buf + ((base + off) & ~3ull));
            acc += (uint64_t)p[0] ^ ((uint64_t)p[1] << 32);
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
    printf("%-12s %-28s %12s %14s %12s\n", "filter", "pattern", "ms", "G items/s", "G reads/s");
    for (uint64_t sz : sizes) {
        for (int pat = 0; pat < 2; pat++) {
            float best = 1e9f;
            for (int rep = 0; rep < 4; rep++) {
                cudaEventRecord(e0);
                if (pat == 0) k_probe<1><<<grid, 256>>>(buf, sz, 8, per * 7, sink);
                else k_probe<7><<<grid, 256>>>(buf, sz, 514, per, sink);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
            }
            double n_items = pat == 0 ? items * 7 : items;
            double reads = items * 7;
            printf("%-12.1f %-28s %12.3f %14.2f %12.2f\n", sz / 1e6, pat == 0 ? "random 8 B reads (1 sector)" : "7 reads in a 514 B window", best,
                   n_items / best / 1e6, reads / best / 1e6);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
