// Independent roofline of the Bloom probes (north_star "random-sector rate for Bloom probes"): what the memory
// system of a B200 delivers for the ACCESS SHAPE of one contains4 window as k_walk issues it -- nothing of k_walk's code:
//   an item = one window at a uniformly random bit position of the filter:
//             7 independent, 8-byte aligned 64-bit loads at (window + off_i), off_i < 4096 (+13), and for every probe whose
//             14-bit candidate field crosses its 64-bit word (13 bit positions of 64) one more 32-bit load of the next word;
//   no address depends on loaded data, the position comes from a multiply-high (no 64-bit modulo), the loaded words are
//   only XOR-ed into an accumulator.
// The figure reported as the roofline is the PLATEAU over the loads kept in flight (items per thread x resident warps), for
// filters that are L2-resident (6.9 MB, configs[1]/[2]) or not (150 MB = configs[3], 1 GB, 4.5 GB = configs[4]).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/probe_roofline tools/probe_roofline.cu
//   ./tools/bin/probe_roofline > profiles/r2_probe_roofline.txt        (a JSON summary is the last line)
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x)
{
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

// U items in flight per thread, `per` rounds of them
template <int U>
__global__ void __launch_bounds__(256) k_window(const uint2 *__restrict__ words, uint64_t nbits, int per, uint64_t *sink)
{
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    uint64_t h = mix(tid + 0x9E3779B97F4A7C15ULL);
    for (int it = 0; it < per; it++) {
        uint2 q[U][7];
        uint32_t nx[U][7];
#pragma unroll
        for (int u = 0; u < U; u++) {
            h = h * 6364136223846793005ULL + 1442695040888963407ULL;       // next position: 2 integer ops, off the loads' path
            const uint64_t base = __umul64hi(h, nbits);                     // uniform in [0, nbits)
            const uint64_t o = mix(h);
#pragma unroll
            for (int i = 0; i < 7; i++) {
                const uint64_t bit = base + (i ? ((o >> (9 * i)) & 4095u) : 0u);
                const uint2 *w = words + (bit >> 6);
                q[u][i] = __ldg(w);
                nx[u][i] = 0;
                if (((unsigned)bit & 63u) > 50u) nx[u][i] = __ldg((const uint32_t *)(w + 1));
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int i = 0; i < 7; i++) acc ^= q[u][i].x ^ q[u][i].y ^ nx[u][i];
    }
    if (acc == 0x12345u) *sink = acc;
}

// plain random 32-byte sectors (one 8-byte load each): the memory system's sector rate without any window locality
template <int U>
__global__ void __launch_bounds__(256) k_sector(const uint2 *__restrict__ words, uint64_t nwords, int per, uint64_t *sink)
{
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    uint64_t h = mix(tid + 77);
    for (int it = 0; it < per; it++) {
        uint2 q[U * 7];
#pragma unroll
        for (int u = 0; u < U * 7; u++) {
            h = h * 6364136223846793005ULL + 1442695040888963407ULL;
            q[u] = __ldg(words + __umul64hi(h, nwords));
        }
#pragma unroll
        for (int u = 0; u < U * 7; u++) acc ^= q[u].x ^ q[u].y;
    }
    if (acc == 0x12345u) *sink = acc;
}

template <int U> static float time_window(const uint2 *buf, uint64_t nbits, int grid, int per, uint64_t *sink, bool sector)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0);
        if (sector) k_sector<U><<<grid, 256>>>(buf, nbits / 64, per, sink);
        else k_window<U><<<grid, 256>>>(buf, nbits, per, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return best;
}

int main()
{
    const uint64_t sizes[] = {6900953ull, 150000905ull, 1000000000ull, 4500000000ull};
    const char *names[] = {"l2_resident_6.9MB", "150MB", "dram_1GB", "dram_4.5GB"};
    uint8_t *buf; uint64_t *sink;
    if (cudaMalloc(&buf, sizes[3] + 8192) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMalloc(&sink, 8);
    cudaMemset(buf, 0x5a, sizes[3] + 8192);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("%-20s %-10s %5s %6s %10s %14s %14s\n", "filter", "pattern", "U", "CTA/SM", "ms", "G windows/s", "G loads/s");
    double plateau[4] = {0, 0, 0, 0}, plateau_sector[4] = {0, 0, 0, 0};
    char best_cfg[4][64];
    for (int s = 0; s < 4; s++) {
        const uint64_t nbits = (sizes[s] - 1024) * 8;                 // windows stay inside the buffer
        for (int sector = 0; sector < 2; sector++)
            for (int cps : {2, 4, 8}) {
                for (int U : {1, 2, 4}) {
                    const int grid = sms * cps * 4, per = 48 / U;
                    const double items = (double)grid * 256 * per * U;
                    float ms = U == 1 ? time_window<1>((const uint2 *)buf, nbits, grid, per, sink, sector)
                             : U == 2 ? time_window<2>((const uint2 *)buf, nbits, grid, per, sink, sector)
                                      : time_window<4>((const uint2 *)buf, nbits, grid, per, sink, sector);
                    const double wps = items / ms / 1e6, lps = items * (sector ? 7.0 : 7.0 + 7.0 * 13 / 64) / ms / 1e6;
                    printf("%-20s %-10s %5d %6d %10.3f %14.2f %14.2f\n", names[s], sector ? "sectors" : "window", U, cps, ms, wps, lps);
                    if (!sector && wps > plateau[s]) { plateau[s] = wps; snprintf(best_cfg[s], sizeof best_cfg[s], "U=%d, %d CTAs/SM", U, cps); }
                    if (sector && wps > plateau_sector[s]) plateau_sector[s] = wps;
                }
            }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    printf("{\"windows_per_s\": {");
    for (int s = 0; s < 4; s++) printf("%s\"%s\": %.4g", s ? ", " : "", names[s], plateau[s] * 1e9);
    printf("}, \"random_sector_groups_of_7_per_s\": {");
    for (int s = 0; s < 4; s++) printf("%s\"%s\": %.4g", s ? ", " : "", names[s], plateau_sector[s] * 1e9);
    printf("}, \"best_config\": {");
    for (int s = 0; s < 4; s++) printf("%s\"%s\": \"%s\"", s ? ", " : "", names[s], best_cfg[s]);
    printf("}}\n");
    return 0;
}
