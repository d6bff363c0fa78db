// Backward step 1: (global row id, bag id) pairs, stably radix-sorted by row id.  sm_100a.
//
// What this replaces: EmbeddingBag.backward emits one COO entry per looked-up index and
// grad.coalesce() (optim/rwsadagrad.py:98) sorts + merges them.  Here only 8 bytes per looked-up
// index are sorted (uint32 key, int32 bag id); the [N, D] gradient expansion never exists.
//
// Design (everything is L2-resident on B200: N ~ 1e6 pairs = 8 MB vs 126 MB of L2):
//   build_keys   one CTA per 2048 sorted-input positions: finds each position's bag by binary search in
//                the offsets (first restricted to the CTA's bag range), writes key/bag, the per-tile digit
//                counts of pass 0, and the digit histograms of ALL passes (they are permutation-invariant)
//   per pass     row_scan   256 CTAs (one per digit): exclusive scan of that digit's per-tile counts
//                scatter    one CTA per tile: warp-private match_any ranking (stable), tile-local reorder
//                           in shared memory, coalesced scatter; it also counts the NEXT pass's per-tile
//                           digits at the destination (warp-aggregated atomics), so there is no separate
//                           histogram kernel per pass
//   passes = ceil(bits(total_rows-1) / 8): Terabyte (188 M rows) needs 4, the 8x1M bench shape 3.
// LSD radix sort is stable, so equal rows keep ascending original position == ascending bag id:
// the order torch.sort(stable=True) gives, which tests/ pin bit-exactly.
#include "common.cuh"
#include "scan.cuh"

namespace dlrmx {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;
constexpr int kSortTile = kSortThreads * kSortItems;  // 2048
constexpr int kMaxPasses = 4;

__device__ __forceinline__ unsigned digit_of(uint32_t key, int pass) { return (key >> (8 * pass)) & 255u; }

template <typename IdxT>
__global__ void __launch_bounds__(kSortThreads)
build_keys_kernel(const IdxT* __restrict__ indices, const IdxT* __restrict__ offsets, int64_t N,
                  int64_t total_bags, int B, const int64_t* __restrict__ table_row_base, int passes,
                  int n_tiles, uint32_t* __restrict__ keys, int32_t* __restrict__ bags,
                  int* __restrict__ tile_counts0, int* __restrict__ hist, uint32_t invalid_key,
                  int* __restrict__ err_flag) {
    __shared__ int s_hist[kMaxPasses][256];
    __shared__ int64_t s_range[2];
    const int tile = blockIdx.x;
    const int64_t p0 = (int64_t)tile * kSortTile;
    const int64_t p1 = min(N, p0 + kSortTile);
    for (int i = threadIdx.x; i < kMaxPasses * 256; i += kSortThreads) (&s_hist[0][0])[i] = 0;
    if (threadIdx.x < 2) {
        // upper_bound(offsets, p) - 1 for p = p0 and p = p1-1: the CTA's bag range
        const int64_t p = threadIdx.x == 0 ? p0 : p1 - 1;
        int64_t lo = 0, hi = total_bags;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if ((int64_t)offsets[mid] <= p) lo = mid + 1; else hi = mid;
        }
        s_range[threadIdx.x] = lo - 1;
    }
    __syncthreads();
    const int64_t bag_lo = s_range[0], bag_hi = s_range[1];
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int64_t p = p0 + i * kSortThreads + threadIdx.x;
        if (p < p1) {
            int64_t lo = bag_lo, hi = bag_hi + 1;  // answer-1 in [bag_lo, bag_hi]
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if ((int64_t)offsets[mid] <= p) lo = mid + 1; else hi = mid;
            }
            const int64_t bag = lo - 1;
            const int t = (int)(bag / B);
            // an index outside [0, rows_t) would alias a row of another table (or run past the last one): it gets the
            // key one past the last valid row -- sorted to the very end, skipped by the update -- and raises the flag
            // (nn.EmbeddingBag raises / device-asserts on the same input)
            const int64_t base_t = table_row_base[t], ix = (int64_t)indices[p];
            uint32_t key = (uint32_t)(base_t + ix);
            if (ix < 0 || ix >= table_row_base[t + 1] - base_t) {
                key = invalid_key;
                if (err_flag != nullptr) atomicOr(err_flag, 2);
            }
            keys[p] = key;
            bags[p] = (int32_t)bag;
            for (int q = 0; q < passes; ++q) atomicAdd(&s_hist[q][digit_of(key, q)], 1);
        }
    }
    __syncthreads();
    tile_counts0[(int64_t)threadIdx.x * n_tiles + tile] = s_hist[0][threadIdx.x];
    for (int q = 0; q < passes; ++q) {
        const int c = s_hist[q][threadIdx.x];
        if (c) atomicAdd(&hist[q * 256 + threadIdx.x], c);
    }
}

// One CTA per digit: in-place exclusive scan of counts[d][0..n_tiles).
__global__ void __launch_bounds__(256) row_scan_kernel(int* __restrict__ counts, int n_tiles) {
    __shared__ int warp_sums[33];
    int* row = counts + (int64_t)blockIdx.x * n_tiles;
    int carry = 0;
    for (int base = 0; base < n_tiles; base += 256) {
        const int i = base + threadIdx.x;
        const int v = (i < n_tiles) ? row[i] : 0;
        int total;
        const int ex = block_exclusive_scan(v, warp_sums, &total);
        if (i < n_tiles) row[i] = carry + ex;
        carry += total;
    }
}

__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const int32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, int32_t* __restrict__ vals_out, int64_t N, int pass,
                     int n_tiles, const int* __restrict__ tile_excl /*[256][n_tiles] scanned*/,
                     const int* __restrict__ hist /*[256] of this pass*/,
                     int* __restrict__ next_counts /*[256][n_tiles] or null*/) {
    __shared__ int s_warp_hist[kSortThreads / 32][256];
    __shared__ int s_tile_excl[256];
    __shared__ int s_goff[256];
    __shared__ int s_ws[33];
    __shared__ uint32_t s_key[kSortTile];
    __shared__ int32_t s_val[kSortTile];

    const int tile = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t p0 = (int64_t)tile * kSortTile;
    const int valid = (int)min((int64_t)kSortTile, N - p0);
    for (int i = threadIdx.x; i < (kSortThreads / 32) * 256; i += kSortThreads) (&s_warp_hist[0][0])[i] = 0;
    __syncthreads();

    uint32_t key[kSortItems];
    int32_t val[kSortItems];
    int rank[kSortItems];
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {  // warp-striped: original order = (warp, item, lane)
        const int local = w * (32 * kSortItems) + i * 32 + lane;
        const bool ok = local < valid;
        key[i] = ok ? keys_in[p0 + local] : 0xffffffffu;
        val[i] = ok ? vals_in[p0 + local] : 0;
    }
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const unsigned d = digit_of(key[i], pass);
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        int old = 0;
        if (lane == leader) {
            old = s_warp_hist[w][d];
            s_warp_hist[w][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();
    {   // thread d: prefix over warps, then over digits
        const int d = threadIdx.x;
        int run = 0;
#pragma unroll
        for (int ww = 0; ww < kSortThreads / 32; ++ww) {
            const int c = s_warp_hist[ww][d];
            s_warp_hist[ww][d] = run;
            run += c;
        }
        int total;
        const int ex = block_exclusive_scan(run, s_ws, &total);
        s_tile_excl[d] = ex;
        // global start of digit d = sum of hist[0..d) ; block scan again over the pass histogram
        const int h = hist[d];
        const int hex = block_exclusive_scan(h, s_ws, &total);
        s_goff[d] = hex + tile_excl[(int64_t)d * n_tiles + tile] - ex;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const unsigned d = digit_of(key[i], pass);
        const int pos = s_tile_excl[d] + s_warp_hist[w][d] + rank[i];
        s_key[pos] = key[i];
        s_val[pos] = val[i];
    }
    __syncthreads();
    for (int jb = 0; jb < kSortTile; jb += kSortThreads) {
        const int j = jb + threadIdx.x;
        const bool ok = j < valid;
        uint32_t k = 0;
        int dst = 0;
        if (ok) {
            k = s_key[j];
            dst = s_goff[digit_of(k, pass)] + j;
            keys_out[dst] = k;
            vals_out[dst] = s_val[j];
        }
        if (next_counts != nullptr) {
            // per-tile digit count of the next pass at the destination, warp-aggregated
            const unsigned cell = ok ? (digit_of(k, pass + 1) * (unsigned)n_tiles + (unsigned)(dst / kSortTile))
                                     : 0xffffffffu;
            const unsigned peers = __match_any_sync(0xffffffffu, cell);
            if (ok && lane == __ffs(peers) - 1) atomicAdd(&next_counts[cell], __popc(peers));
        }
    }
}

size_t sort_workspace_bytes(int64_t N) {
    const int64_t n_tiles = ceil_div(N > 0 ? N : 1, kSortTile);
    size_t b = 0;
    b += align_up((size_t)N * 4, 256) * 2;                           // ping-pong keys / bags
    b += align_up((size_t)kMaxPasses * 256 * n_tiles * 4, 256);      // per-pass tile counts
    b += align_up((size_t)kMaxPasses * 256 * 4, 256);                // per-pass histograms
    return b;
}

template <typename IdxT>
static int bwd_sort(const IdxT* indices, const IdxT* offsets, int64_t N, int T, int B,
                    const int64_t* table_row_base, int64_t total_rows, uint32_t* sorted_key,
                    int32_t* sorted_bag, int* err_flag, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    DLRMX_REQUIRE(N >= 0 && N < (1ll << 31), "bwd_sort: N must be < 2^31");
    DLRMX_REQUIRE(total_rows >= 0 && total_rows < (1ll << 32) - 1, "bwd_sort: total rows must be < 2^32 - 1");
    DLRMX_REQUIRE((int64_t)T * B < (1ll << 31), "bwd_sort: T*B must be < 2^31");
    if (N == 0) return DLRMX_OK;
    if (workspace_bytes < sort_workspace_bytes(N)) {
        set_error("bwd_sort: workspace %zu < %zu", workspace_bytes, sort_workspace_bytes(N));
        return DLRMX_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int bits = 1;
    while (bits < 32 && ((int64_t)1 << bits) < total_rows + 1) ++bits;   // keys 0 .. total_rows (= the invalid key)
    const int passes = (bits + 7) / 8;
    const int n_tiles = (int)ceil_div(N, kSortTile);

    char* ws = (char*)workspace;
    uint32_t* tmp_key = (uint32_t*)ws; ws += align_up((size_t)N * 4, 256);
    int32_t* tmp_bag = (int32_t*)ws;   ws += align_up((size_t)N * 4, 256);
    int* counts = (int*)ws;            ws += align_up((size_t)kMaxPasses * 256 * n_tiles * 4, 256);
    int* hist = (int*)ws;
    const size_t counts_stride = (size_t)256 * n_tiles;

    // zero the tile counts of passes 1.. and all histograms (pass-0 counts are fully overwritten)
    if (cudaMemsetAsync(counts + counts_stride, 0,
                        (size_t)(((char*)hist + kMaxPasses * 256 * 4) - (char*)(counts + counts_stride)),
                        st) != cudaSuccess) {
        set_error("bwd_sort: memset failed");
        return DLRMX_E_CUDA;
    }
    // buffer X_i holds the input of pass i; X_passes must be the user's output
    uint32_t* kbuf[2] = {(passes % 2) ? tmp_key : sorted_key, (passes % 2) ? sorted_key : tmp_key};
    int32_t* vbuf[2] = {(passes % 2) ? tmp_bag : sorted_bag, (passes % 2) ? sorted_bag : tmp_bag};

    build_keys_kernel<IdxT><<<n_tiles, kSortThreads, 0, st>>>(indices, offsets, N, (int64_t)T * B, B,
                                                             table_row_base, passes, n_tiles, kbuf[0],
                                                             vbuf[0], counts, hist, (uint32_t)total_rows,
                                                             err_flag);
    int rc = check_launch("build_keys");
    if (rc) return rc;
    for (int p = 0; p < passes; ++p) {
        int* c = counts + (size_t)p * counts_stride;
        row_scan_kernel<<<256, 256, 0, st>>>(c, n_tiles);
        if ((rc = check_launch("row_scan"))) return rc;
        radix_scatter_kernel<<<n_tiles, kSortThreads, 0, st>>>(
            kbuf[p & 1], vbuf[p & 1], kbuf[(p + 1) & 1], vbuf[(p + 1) & 1], N, p, n_tiles, c, hist + p * 256,
            (p + 1 < passes) ? c + counts_stride : nullptr);
        if ((rc = check_launch("radix_scatter"))) return rc;
    }
    return DLRMX_OK;
}

}  // namespace dlrmx

extern "C" {
size_t dlrmx_bwd_sort_workspace_bytes(int64_t N, int64_t total_bags) {
    (void)total_bags;
    return dlrmx::sort_workspace_bytes(N);
}
int dlrmx_bwd_sort_i32(const int32_t* indices, const int32_t* offsets, int64_t N, int T, int B,
                       const int64_t* table_row_base, int64_t total_rows, uint32_t* sorted_key,
                       int32_t* sorted_bag, int* err_flag, void* workspace, size_t workspace_bytes,
                       dlrmx_stream_t stream) {
    return dlrmx::bwd_sort<int32_t>(indices, offsets, N, T, B, table_row_base, total_rows, sorted_key,
                                    sorted_bag, err_flag, workspace, workspace_bytes, stream);
}
int dlrmx_bwd_sort_i64(const int64_t* indices, const int64_t* offsets, int64_t N, int T, int B,
                       const int64_t* table_row_base, int64_t total_rows, uint32_t* sorted_key,
                       int32_t* sorted_bag, int* err_flag, void* workspace, size_t workspace_bytes,
                       dlrmx_stream_t stream) {
    return dlrmx::bwd_sort<int64_t>(indices, offsets, N, T, B, table_row_base, total_rows, sorted_key,
                                    sorted_bag, err_flag, workspace, workspace_bytes, stream);
}
}
