// Backward step 2: run-length structure of the sorted keys (unique rows + run offsets).
// Needed only by the non-fused parity seam (coalesced COO gradient for the reference's own optimizers)
// and by the tests that pin the integer artefacts against sort/unique_consecutive; the fused update
// finds run boundaries on the fly.  What it replaces: the index half of grad.coalesce()
// (optim/rwsadagrad.py:98-100).
#include "common.cuh"
#include "scan.cuh"

namespace dlrmx {

constexpr int kSegThreads = 256;
constexpr int kSegItems = 8;
constexpr int kSegTile = kSegThreads * kSegItems;

__device__ __forceinline__ bool is_head(const uint32_t* __restrict__ k, int64_t p) {
    return p == 0 || k[p] != k[p - 1];
}

__global__ void __launch_bounds__(kSegThreads)
count_heads_kernel(const uint32_t* __restrict__ skey, int64_t N, int* __restrict__ block_counts) {
    __shared__ int ws[33];
    const int64_t base = (int64_t)blockIdx.x * kSegTile + (int64_t)threadIdx.x * kSegItems;
    int c = 0;
#pragma unroll
    for (int i = 0; i < kSegItems; ++i) {
        const int64_t p = base + i;
        if (p < N && is_head(skey, p)) ++c;
    }
    int total;
    block_exclusive_scan(c, ws, &total);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kSegThreads)
write_runs_kernel(const uint32_t* __restrict__ skey, int64_t N, const int* __restrict__ block_excl,
                  const int* __restrict__ total_heads, uint32_t* __restrict__ unique_key,
                  int32_t* __restrict__ run_off, int32_t* __restrict__ num_unique) {
    __shared__ int ws[33];
    const int64_t base = (int64_t)blockIdx.x * kSegTile + (int64_t)threadIdx.x * kSegItems;
    bool h[kSegItems];
    int c = 0;
#pragma unroll
    for (int i = 0; i < kSegItems; ++i) {
        const int64_t p = base + i;
        h[i] = p < N && is_head(skey, p);
        c += h[i];
    }
    int total;
    int r = block_excl[blockIdx.x] + block_exclusive_scan(c, ws, &total);
#pragma unroll
    for (int i = 0; i < kSegItems; ++i)
        if (h[i]) {
            unique_key[r] = skey[base + i];
            run_off[r] = (int32_t)(base + i);
            ++r;
        }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const int U = *total_heads;
        run_off[U] = (int32_t)N;
        *num_unique = U;
    }
}

static size_t segments_workspace_bytes(int64_t N) {
    return align_up((size_t)(ceil_div(N > 0 ? N : 1, kSegTile) + 1) * 4, 256);
}

}  // namespace dlrmx

extern "C" {
size_t dlrmx_bwd_segments_workspace_bytes(int64_t N) { return dlrmx::segments_workspace_bytes(N); }

int dlrmx_bwd_segments(const uint32_t* sorted_key, int64_t N, uint32_t* unique_key, int32_t* run_off,
                       int32_t* num_unique, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(N >= 0 && N < (1ll << 31), "bwd_segments: N must be < 2^31");
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) {
        // run_off[0] = 0, num_unique = 0
        if (cudaMemsetAsync(run_off, 0, 4, st) != cudaSuccess || cudaMemsetAsync(num_unique, 0, 4, st) != cudaSuccess) {
            set_error("bwd_segments: memset failed");
            return DLRMX_E_CUDA;
        }
        return DLRMX_OK;
    }
    if (workspace_bytes < segments_workspace_bytes(N)) {
        set_error("bwd_segments: workspace %zu < %zu", workspace_bytes, segments_workspace_bytes(N));
        return DLRMX_E_WORKSPACE;
    }
    const int nblk = (int)ceil_div(N, kSegTile);
    int* block_counts = (int*)workspace;  // [nblk] + total at [nblk]
    count_heads_kernel<<<nblk, kSegThreads, 0, st>>>(sorted_key, N, block_counts);
    int rc = check_launch("count_heads");
    if (rc) return rc;
    single_block_exclusive_scan<<<1, 1024, 0, st>>>(block_counts, nblk, block_counts + nblk);
    if ((rc = check_launch("scan_heads"))) return rc;
    write_runs_kernel<<<nblk, kSegThreads, 0, st>>>(sorted_key, N, block_counts, block_counts + nblk, unique_key,
                                                   run_off, num_unique);
    return check_launch("write_runs");
}
}
