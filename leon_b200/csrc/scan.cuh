// Generic device-wide exclusive scan (reduce -> scan of tile sums -> apply), hand-written.
// in(i) produces element i, out(i, exclusive_prefix, element) consumes it; both are device functors, so head
// flags, varint lengths etc. are computed on the fly and results are scattered without an intermediate array.
#pragma once
#include "common.cuh"

namespace lbscan {

constexpr int THREADS = 256;
constexpr int ITEMS = 8;
constexpr int TILE = THREADS * ITEMS;

struct U32x4 { uint32_t a, b, c, d; };
__host__ __device__ inline U32x4 operator+(const U32x4 &x, const U32x4 &y) { return U32x4{x.a + y.a, x.b + y.b, x.c + y.c, x.d + y.d}; }

template <typename T> __device__ __forceinline__ T shfl_up_any(T v, int delta)
{
    static_assert(sizeof(T) % 4 == 0, "scan element must be a multiple of 4 bytes");
    constexpr int W = sizeof(T) / 4;
    uint32_t w[W];
    memcpy(w, &v, sizeof(T));
#pragma unroll
    for (int i = 0; i < W; i++) w[i] = __shfl_up_sync(0xffffffffu, w[i], delta);
    T r;
    memcpy(&r, w, sizeof(T));
    return r;
}

// exclusive scan of one value per thread across the block; returns the exclusive prefix, *total = block sum
template <typename T> __device__ T block_exclusive(T v, T *total)
{
    __shared__ unsigned char smem_raw[(THREADS / 32 + 1) * sizeof(T)];
    T *wsum = reinterpret_cast<T *>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = shfl_up_any(incl, d);
        if (lane >= d) incl = o + incl;
    }
    __syncthreads();  // protect wsum reuse across calls
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        T w = lane < THREADS / 32 ? wsum[lane] : T();
        T wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            T o = shfl_up_any(wi, d);
            if (lane >= d) wi = o + wi;
        }
        if (lane < THREADS / 32) wsum[lane] = wi;          // inclusive warp sums
        if (lane == THREADS / 32 - 1) wsum[THREADS / 32] = wi;  // block total
    }
    __syncthreads();
    T base = warp ? wsum[warp - 1] : T();
    *total = wsum[THREADS / 32];
    T excl_in_warp = shfl_up_any(incl, 1);
    if (lane == 0) excl_in_warp = T();
    return base + excl_in_warp;
}

template <typename T, typename InFn> __global__ void __launch_bounds__(THREADS) k_reduce(uint64_t n, InFn in, T *partials)
{
    uint64_t base = (uint64_t)blockIdx.x * TILE + (uint64_t)threadIdx.x * ITEMS;
    T s = T();
#pragma unroll
    for (int j = 0; j < ITEMS; j++)
        if (base + j < n) s = s + in(base + j);
    T total;
    block_exclusive(s, &total);
    if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// single block: in-place exclusive scan of nb tile sums, grand total to *total
template <typename T> __global__ void __launch_bounds__(THREADS) k_scan_partials(T *partials, uint64_t nb, T *total)
{
    T carry = T();
    for (uint64_t base = 0; base < nb; base += THREADS) {
        uint64_t i = base + threadIdx.x;
        T v = i < nb ? partials[i] : T();
        T blk;
        T ex = block_exclusive(v, &blk);
        if (i < nb) partials[i] = carry + ex;
        carry = carry + blk;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

template <typename T, typename InFn, typename OutFn>
__global__ void __launch_bounds__(THREADS) k_apply(uint64_t n, InFn in, const T *partials, OutFn out)
{
    uint64_t base = (uint64_t)blockIdx.x * TILE + (uint64_t)threadIdx.x * ITEMS;
    T v[ITEMS];
    T s = T();
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        v[j] = base + j < n ? in(base + j) : T();
        s = s + v[j];
    }
    T total;
    T ex = block_exclusive(s, &total) + partials[blockIdx.x];
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        if (base + j < n) out(base + j, ex, v[j]);
        ex = ex + v[j];
    }
}

// scratch: device buffer for tile sums (+1 slot for the total).  Returns a device pointer to the grand total.
template <typename T, typename InFn, typename OutFn>
T *exclusive_scan(uint64_t n, InFn in, OutFn out, DevBuf &scratch, cudaStream_t st, LaunchCounter &lc)
{
    uint64_t nb = (n + TILE - 1) / TILE;
    if (nb == 0) nb = 1;
    scratch.reserve((nb + 1) * sizeof(T));
    T *partials = scratch.as<T>();
    T *total = partials + nb;
    LB_LAUNCH(lc, (k_reduce<T, InFn>), (unsigned)nb, THREADS, 0, st, n, in, partials);
    LB_LAUNCH(lc, (k_scan_partials<T>), 1, THREADS, 0, st, partials, nb, total);
    LB_LAUNCH(lc, (k_apply<T, InFn, OutFn>), (unsigned)nb, THREADS, 0, st, n, in, partials, out);
    return total;
}

}  // namespace lbscan
