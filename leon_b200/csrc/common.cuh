// Host-side plumbing shared by the kernels' launchers: error type, grow-only device buffers, launch counter.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <stdexcept>
#include "../../include/leon_b200.h"

struct LeonError {
    int code;
    std::string msg;
    LeonError(int c, const std::string &m) : code(c), msg(m) {}
};

#define CUDA_TRY(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t _e = (expr);                                                                         \
        if (_e != cudaSuccess) {                                                                         \
            char _b[512];                                                                                \
            snprintf(_b, sizeof _b, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
            throw LeonError(_e == cudaErrorMemoryAllocation ? LEON_E_NOMEM : LEON_E_CUDA, _b);           \
        }                                                                                                \
    } while (0)

// every kernel launch of the library goes through this so that leon_timings.kernel_launches is a real count
struct LaunchCounter { uint64_t n = 0; };
#define LB_LAUNCH(counter, kernel, grid, block, smem, stream, ...)                   \
    do {                                                                             \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                  \
        (counter).n++;                                                               \
        CUDA_TRY(cudaGetLastError());                                                \
    } while (0)

// grow-only device allocation
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    void reserve(size_t n)
    {
        if (n <= cap) return;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = n + n / 16 + 256;
        CUDA_TRY(cudaMalloc(&p, want));
        cap = want;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *)p; }
};

// grow-only pinned host allocation
struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    void reserve(size_t n)
    {
        if (n <= cap) return;
        if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
        size_t want = n + n / 16 + 256;
        CUDA_TRY(cudaHostAlloc(&p, want, cudaHostAllocDefault));
        cap = want;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *)p; }
};

static inline unsigned cdiv_u(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }
