// Block-level scan helpers shared by the radix sort and the run-length (segments) kernels.
#pragma once
#include "common.cuh"

namespace dlrmx {

__device__ __forceinline__ int warp_inclusive_scan(int v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}

// Exclusive scan of one int per thread over the whole block (blockDim.x multiple of 32, <= 1024).
// `warp_sums` is a __shared__ int[33]; returns the exclusive prefix, *total gets the block sum.
__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int* total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int inc = warp_inclusive_scan(v);
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        int s = (lane < nw) ? warp_sums[lane] : 0;
        int si = warp_inclusive_scan(s);
        warp_sums[lane] = si - s;
        if (lane == 31) warp_sums[32] = si;
    }
    __syncthreads();
    const int excl = warp_sums[w] + inc - v;
    *total = warp_sums[32];
    __syncthreads();
    return excl;
}

// In-place exclusive scan of a[0..n) by ONE block (any n); optionally writes the total to *total_out.
static __global__ void __launch_bounds__(1024) single_block_exclusive_scan(int* a, int64_t n, int* total_out) {
    __shared__ int warp_sums[33];
    const int nt = blockDim.x;
    const int64_t per = ceil_div(n, nt);
    const int64_t lo = min(n, (int64_t)threadIdx.x * per), hi = min(n, lo + per);
    int s = 0;
    for (int64_t i = lo; i < hi; ++i) s += a[i];
    int total;
    int run = block_exclusive_scan(s, warp_sums, &total);
    for (int64_t i = lo; i < hi; ++i) {
        int t = a[i];
        a[i] = run;
        run += t;
    }
    if (total_out != nullptr && threadIdx.x == 0) *total_out = total;
}

}  // namespace dlrmx
