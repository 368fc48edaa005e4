// Dependent-issue latencies seen by ONE warp on sm_100a (what bounds the serial range-coder recurrence).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/lat_bench tools/lat_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define N_IT 4096
template <int KIND> __global__ void k_lat(uint64_t *out, uint64_t seed, uint32_t a, uint32_t b, int active_lanes)
{
    if ((int)threadIdx.x >= active_lanes) return;
    uint64_t x = seed + threadIdx.x;
    uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
    __shared__ uint32_t sm[64];
    sm[threadIdx.x & 63] = threadIdx.x & 63; sm[(threadIdx.x + 32) & 63] = (threadIdx.x + 32) & 63;
    __syncwarp();
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_IT; i++) {
        if (KIND == 0) { asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x) : "r"((uint32_t)x), "r"(a)); }            // IMAD.WIDE chain (lo feeds multiplier)
        if (KIND == 1) { asm volatile("add.u32 %0, %0, %1;" : "+r"(lo) : "r"(a)); }                                       // IADD chain
        if (KIND == 2) { asm volatile("add.u64 %0, %0, %1;" : "+l"(x) : "l"((uint64_t)a << 20)); }                        // 64-bit add chain
        if (KIND == 3) { x = __umul64hi(x, ((uint64_t)a << 32) | b) | (1ull << 63); }                                     // mulhi64 chain (+1 LOP)
        if (KIND == 4) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(a), "r"(b)); }                        // IMAD chain
        if (KIND == 5) { asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(lo) : "r"(hi), "r"(a & 7)); }               // SHF chain
        if (KIND == 6) { asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lo) : "r"((uint32_t)__cvta_generic_to_shared(sm) + 4 * (lo & 63))); }  // LDS pointer chase
        if (KIND == 7) { lo += a; if (__builtin_expect(lo == 0x12345u, 0)) { out[1] = lo; } }                              // add + not-taken branch
        if (KIND == 8) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(lo) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(a), "r"(b)); } // ALU->FMA alternating
        if (KIND == 9) { uint64_t q = __umul64hi(x, ((uint64_t)a << 32) | b) >> (a & 7); x = q * (uint64_t)b; x |= 1ull << 60; } // mulhi + shift + 64x32 mul
    }
    uint32_t r0 = lo, r1 = hi, r2 = a, r3 = b, r4 = lo ^ a, r5 = hi ^ b, r6 = lo + 7, r7 = hi + 9;
    uint64_t w0 = x, w1 = x + 1, w2 = x + 2, w3 = x + 3;
    if (KIND >= 10) {
#pragma unroll 8
        for (int i = 0; i < N_IT; i++) {
            if (KIND == 10) {   // 8 independent ALU chains: single-warp issue rate (per instruction = /8)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r0) : "r"(a), "r"(b)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r1) : "r"(a), "r"(b));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r2) : "r"(a), "r"(b)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r3) : "r"(a), "r"(b));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r4) : "r"(a), "r"(b)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r5) : "r"(a), "r"(b));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r6) : "r"(a), "r"(b)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r7) : "r"(a), "r"(b));
            }
            if (KIND == 11) {   // 4 ALU + 4 FMA-pipe independent chains
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r0) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r1) : "r"(a), "r"(b));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r2) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r3) : "r"(a), "r"(b));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r4) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r5) : "r"(a), "r"(b));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r6) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r7) : "r"(a), "r"(b));
            }
            if (KIND == 12) {   // 4 independent IMAD.WIDE chains
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w0) : "r"(r0), "r"(a)); asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w1) : "r"(r1), "r"(a));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w2) : "r"(r2), "r"(a)); asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w3) : "r"(r3), "r"(a));
            }
            if (KIND == 13) {   // 8 independent IMAD.lo chains
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r0) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r1) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r2) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r3) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r4) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r5) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r6) : "r"(a), "r"(b)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r7) : "r"(a), "r"(b));
            }
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) { out[0] = x + lo + hi + r0 + r1 + r2 + r3 + r4 + r5 + r6 + r7 + w0 + w1 + w2 + w3; out[2 + KIND] = (uint64_t)(t1 - t0); }
}

int main()
{
    uint64_t *d; cudaMalloc(&d, 64 * 8); cudaMemset(d, 0, 64 * 8);
    const char *names[] = {"IMAD.WIDE chain", "IADD chain", "add.u64 chain", "umul64hi|or chain", "IMAD.lo chain", "SHF chain", "LDS chase",
                           "IADD + not-taken BRA", "LOP3->IMAD alternating (per pair)", "mulhi>>s * w | bit",
                           "8 indep LOP3 (per 8)", "4 LOP3 + 4 IMAD indep (per 8)", "4 indep IMAD.WIDE (per 4)", "8 indep IMAD.lo (per 8)"};
    for (int lanes = 1; lanes <= 1; lanes += 31) {
        k_lat<0><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<1><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<2><<<1, 32>>>(d, 12345, 77777, 991, lanes);
        k_lat<3><<<1, 32>>>(d, 12345, 0x9e3779b9u, 991, lanes); k_lat<4><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<5><<<1, 32>>>(d, 12345, 77777, 991, lanes);
        k_lat<6><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<7><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<8><<<1, 32>>>(d, 12345, 77777, 991, lanes);
        k_lat<9><<<1, 32>>>(d, 12345, 0x9e3779b9u, 991, lanes);
        k_lat<10><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<11><<<1, 32>>>(d, 12345, 77777, 991, lanes);
        k_lat<12><<<1, 32>>>(d, 12345, 77777, 991, lanes); k_lat<13><<<1, 32>>>(d, 12345, 77777, 991, lanes);
        cudaDeviceSynchronize();
        uint64_t h[64]; cudaMemcpy(h, d, 64 * 8, cudaMemcpyDeviceToHost);
        printf("active lanes %d (%s)\n", lanes, cudaGetErrorString(cudaGetLastError()));
        for (int k = 0; k < 14; k++) printf("  %-36s %.2f cycles/iter\n", names[k], (double)h[2 + k] / N_IT);
    }
    return 0;
}
