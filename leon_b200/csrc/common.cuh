// Host-side plumbing shared by the kernels' launchers: error type, grow-only device buffers, launch counter.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <stdexcept>
#include <vector>
#include <string.h>
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
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }              // a member of a context can never be forgotten in a release list
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
    PinBuf() = default;
    PinBuf(const PinBuf &) = delete;
    PinBuf &operator=(const PinBuf &) = delete;
    ~PinBuf() { release(); }
};

// Per-kernel device time, CUDA events on the launching stream around selected launches.  Slots are named after
// the kernel; leon_get_kernel_times() reports total ms and launch count per slot for the last run.
struct KernelTimer {
    struct Slot { const char *name; double ms; uint32_t n; };
    struct Pending { int slot; cudaEvent_t a, b; };
    std::vector<Slot> slots;
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> pool;
    bool enabled = true;
    int slot_of(const char *name)
    {
        for (size_t i = 0; i < slots.size(); i++) if (slots[i].name == name || !strcmp(slots[i].name, name)) return (int)i;
        slots.push_back(Slot{name, 0.0, 0});
        return (int)slots.size() - 1;
    }
    cudaEvent_t get()
    {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreate(&e));
        return e;
    }
    cudaEvent_t begin(cudaStream_t st)
    {
        cudaEvent_t a = get();
        CUDA_TRY(cudaEventRecord(a, st));
        return a;
    }
    void end(const char *name, cudaEvent_t a, cudaStream_t st)
    {
        cudaEvent_t b = get();
        CUDA_TRY(cudaEventRecord(b, st));
        pending.push_back(Pending{slot_of(name), a, b});
    }
    void reset()
    {
        for (auto &p : pending) { pool.push_back(p.a); pool.push_back(p.b); }
        pending.clear();
        slots.clear();
    }
    void resolve()   // after the stream has been synchronised
    {
        for (auto &p : pending) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) { slots[p.slot].ms += ms; slots[p.slot].n++; }
            else cudaGetLastError();
            pool.push_back(p.a); pool.push_back(p.b);
        }
        pending.clear();
    }
    void release()
    {
        reset();
        for (cudaEvent_t e : pool) cudaEventDestroy(e);
        pool.clear();
    }
};
#define LB_TIMED(timer, name, stream, ...)                                           \
    do {                                                                             \
        if ((timer).enabled) {                                                       \
            cudaEvent_t _ta = (timer).begin(stream);                                 \
            __VA_ARGS__;                                                             \
            (timer).end(name, _ta, stream);                                          \
        } else { __VA_ARGS__; }                                                      \
    } while (0)

static inline unsigned cdiv_u(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }
