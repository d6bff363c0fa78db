// Table-batched embedding-bag forward (sum pooling) for sm_100a.
//
// Replaces nn.EmbeddingBag(mode="sum") as driven by DLRM_Net.apply_emb (dlrm_s_pytorch.py:526-587),
// the batched module call of dlrm_s_pytorch.py:589-591 and the yx kernel
// (yx_modfs/table_batched_embeddings_cuda_yx.cu:251-315).  Not a translation of the latter (one 16-B load in
// flight per lane, 32-bit addressing, a serial offsets -> index -> row dependent chain per bag).  Here:
//   * a group of G = D/4 lanes (a full warp at D=128) owns a TILE of G consecutive bags.  The tile's metadata is
//     fetched lane-parallel with coalesced loads -- lane i reads offsets[bag_i], offsets[bag_i+1], the table
//     pointer / row count of its bag and the bag's FIRST index -- so the offsets -> index -> pointer chain is
//     paid once per G bags (two memory latencies), not once per bag;
//   * the tile is then walked in batches of K bags in LOCKSTEP: the metadata of the K bags is broadcast with
//     warp shuffles, round r issues the r-th row of all K bags before any is consumed (K independent 16-B
//     `ld.global.nc.L1::no_allocate` loads in flight per lane) and the indices of round r+1 are requested
//     before the rows of round r are consumed (index stream prefetch, mostly L1 hits: 8 indices per sector);
//   * each bag is summed sequentially in index order with plain fp32 adds -> bit-exact with ATen;
//   * 64-bit row addressing, per-table base pointers, strided output so pooled rows land directly in the
//     [B, 1+T, D] interaction input (no torch.cat afterwards).
// HBM-bound: algorithmic bytes = N*D*4 (rows) + N*ib + T*B*ob + T*B*D*4 (pooled rows).
#pragma once
#include <type_traits>
#include "common.cuh"

namespace dlrmx {

template <int G>
__device__ __forceinline__ unsigned fwd_group_mask() {
    if constexpr (G == 32) {
        return 0xffffffffu;
    } else {
        return ((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1));
    }
}

template <typename T>
__device__ __forceinline__ T shfl64(unsigned mask, T v, int src, int width) {
    static_assert(sizeof(T) == 8, "shfl64 moves 8-byte values");
    unsigned long long u = (unsigned long long)v;
    unsigned lo = __shfl_sync(mask, (unsigned)u, src, width);
    unsigned hi = __shfl_sync(mask, (unsigned)(u >> 32), src, width);
    return (T)(((unsigned long long)hi << 32) | lo);
}
__device__ __forceinline__ int64_t shfl_idx(unsigned mask, int64_t v, int src, int width) {
    return shfl64<int64_t>(mask, v, src, width);
}
__device__ __forceinline__ int32_t shfl_idx(unsigned mask, int32_t v, int src, int width) {
    return __shfl_sync(mask, v, src, width);
}

// K bags in lockstep x U consecutive rows of each bag in flight; FULL: a tile is G bags (short bags, e.g. the
// one-index-per-bag Criteo layout), otherwise a tile is one batch of K bags (long bags: more groups, and the
// memory-level parallelism comes from unrolling ALONG the bag -- loads are issued U at a time, the adds still
// happen one by one in index order).
// OCC > 0 overrides the CTAs-per-SM target of the register allocation.
// EXACT: D == VEC*G*CH, every lane owns columns of the row (no per-lane guard).
// THREADS: CTA size (224 x 4 CTAs per SM = 28 warps at up to 72 registers: between the 24 warps of 256 x 3 and the 64-register
// budget of 256 x 4).
// TabT: element type of the table rows (float, or __nv_bfloat16 widened on load; sums, outputs and weights stay fp32).
template <int VEC, int G, int CH, int K, int U, bool FULL, bool SCATTER, typename IdxT, int OCC = 0, bool EXACT = false,
          int THREADS = 256, typename TabT = float>
__global__ void __launch_bounds__(THREADS, OCC > 0 ? OCC : ((K * U * CH <= 4) ? 4 : ((K * U * CH <= 8) ? 3 : 2)))
embbag_fwd_kernel(const TabT* const* __restrict__ tables, const int64_t* __restrict__ table_rows,
                  const IdxT* __restrict__ indices, const IdxT* __restrict__ offsets, int64_t N,
                  int B, int D, int64_t ld, int64_t total_bags, const float* __restrict__ psw,
                  float* __restrict__ out, int64_t stride_b, int64_t stride_t, int* err_flag,
                  float* const* __restrict__ out_peers, int lb) {
    constexpr int KB = (K < G) ? K : G;       // bags walked in lockstep
    constexpr int TILE = FULL ? G : KB;       // bags per group
    const int gl = threadIdx.x & (G - 1);
    const unsigned gmask = fwd_group_mask<G>();
    const int64_t tile = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    const int64_t tile_bag0 = tile * TILE;
    if (tile_bag0 >= total_bags) return;      // whole group leaves together
    const int col0 = gl * VEC;
    // ---- lane-parallel metadata of the tile: lane i (< TILE) describes bag tile_bag0 + i ---------------
    IdxT my_s = 0, my_ix = -1;
    int my_len = 0;
    const TabT* my_base = nullptr;
    int64_t my_rows = 0;
    {
        const int64_t q = tile_bag0 + gl;
        if (gl < TILE && q < total_bags) {
            const int64_t bag = q;
            const IdxT s = offsets[bag];
            const IdxT e = (bag + 1 < total_bags) ? offsets[bag + 1] : (IdxT)N;
            const int t = (int)(bag / B);
            my_base = tables[t];
            my_rows = (table_rows != nullptr) ? table_rows[t] : (int64_t)0x7fffffffffffffffll;
            my_s = s;
            my_len = (int)(e - s);
            if (my_len > 0) my_ix = indices[s];
        }
    }

#pragma unroll 1
    for (int b0 = 0; b0 < TILE; b0 += KB) {
        if (tile_bag0 + b0 >= total_bags) break;
        IdxT start[KB], ix[KB][U];
        int len[KB];
        const TabT* base[KB];
        int64_t nrows[KB];
        int maxlen = 0;
#pragma unroll
        for (int k = 0; k < KB; ++k) {
            start[k] = shfl_idx(gmask, my_s, b0 + k, G);
            ix[k][0] = shfl_idx(gmask, my_ix, b0 + k, G);
            len[k] = __shfl_sync(gmask, my_len, b0 + k, G);
            base[k] = shfl64<const TabT*>(gmask, my_base, b0 + k, G);
            nrows[k] = shfl64<int64_t>(gmask, my_rows, b0 + k, G);
            maxlen = max(maxlen, len[k]);
        }
#pragma unroll
        for (int k = 0; k < KB; ++k)
#pragma unroll
            for (int u = 1; u < U; ++u) ix[k][u] = (u < len[k]) ? indices[start[k] + u] : (IdxT)-1;

        RowVec<VEC> acc[KB][CH];
#pragma unroll
        for (int k = 0; k < KB; ++k)
#pragma unroll
            for (int c = 0; c < CH; ++c) acc[k][c].zero();

        for (int r = 0; r < maxlen; r += U) {
            // rows in flight: bf16 rows of the 8-element lane geometry stay packed (4 registers, widened when added)
            typename std::conditional<(VEC == 8 && sizeof(TabT) == 2), PackedRow8, RowVec<VEC>>::type row[KB][U][CH];
            bool live[KB][U];
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int k = 0; k < KB; ++k) {
                    live[k][u] = r + u < len[k];
                    if (live[k][u] && ((int64_t)ix[k][u] < 0 || (int64_t)ix[k][u] >= nrows[k])) {   // bad index
                        if (err_flag != nullptr && gl == 0) atomicOr(err_flag, 1);
                        live[k][u] = false;
                    }
                    if (live[k][u]) {
                        const TabT* p = base[k] + (int64_t)ix[k][u] * ld + col0;
#pragma unroll
                        for (int c = 0; c < CH; ++c)
                            if (EXACT || col0 + c * G * VEC < D) row[k][u][c].load_stream(p + c * G * VEC);
                    }
                }
            // request the next rounds' indices before this round's rows are consumed
            IdxT nix[KB][U];
#pragma unroll
            for (int k = 0; k < KB; ++k)
#pragma unroll
                for (int u = 0; u < U; ++u)
                    nix[k][u] = (r + U + u < len[k]) ? indices[start[k] + r + U + u] : (IdxT)-1;
#pragma unroll
            for (int k = 0; k < KB; ++k)
#pragma unroll
                for (int u = 0; u < U; ++u)            // sequential in index order: bit-exact with ATen
                    if (live[k][u]) {
                        if (psw == nullptr) {
#pragma unroll
                            for (int c = 0; c < CH; ++c)
                                if (EXACT || col0 + c * G * VEC < D) acc[k][c].add(row[k][u][c]);
                        } else {
                            const float w = __ldg(psw + start[k] + r + u);
#pragma unroll
                            for (int c = 0; c < CH; ++c)
                                if (EXACT || col0 + c * G * VEC < D) acc[k][c].add_scaled(row[k][u][c], w);
                        }
                    }
#pragma unroll
            for (int k = 0; k < KB; ++k)
#pragma unroll
                for (int u = 0; u < U; ++u) ix[k][u] = nix[k][u];
        }

#pragma unroll
        for (int k = 0; k < KB; ++k) {
            const int64_t q = tile_bag0 + b0 + k;
            if (q < total_bags) {
                const int64_t bag = q;
                const int64_t t = bag / B;
                const int64_t b = bag - t * B;
                // model-parallel exchange fused into the lookup: sample b belongs to rank b / lb, whose (peer-mapped)
                // interaction input receives the pooled row directly over NVLink -- 512-byte row stores
                float* o = SCATTER ? out_peers[b / lb] + (b % lb) * stride_b + t * stride_t + col0
                                   : out + b * stride_b + t * stride_t + col0;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (EXACT || col0 + c * G * VEC < D) acc[k][c].store(o + c * G * VEC);
            }
        }
    }
}

template <int VEC, int G, int CH, int K, int U, bool FULL, typename IdxT, int OCC = 0, int THREADS = 256, typename TabT = float>
static int launch_fwd(const TabT* const* tables, const int64_t* table_rows, const IdxT* indices,
                      const IdxT* offsets, int64_t N, int T, int B, int D, int64_t ld, const float* psw, float* out,
                      int64_t stride_b, int64_t stride_t, int* err_flag, float* const* out_peers, int lb,
                      cudaStream_t st) {
    const int64_t total_bags = (int64_t)T * B;
    constexpr int TILE = FULL ? G : ((K < G) ? K : G);
    const int64_t groups = ceil_div(total_bags, TILE);   // one group of G lanes per tile of TILE bags
    const int threads = THREADS;
    const int64_t blocks = ceil_div(groups * G, threads);
    if (blocks == 0) return DLRMX_OK;
#define DLRMX_FWD_LAUNCH(SC, EX, ...) \
    embbag_fwd_kernel<VEC, G, CH, K, U, FULL, SC, IdxT, OCC, EX, THREADS, TabT><<<(unsigned)blocks, threads, 0, st>>>(__VA_ARGS__)
    const bool exact = D == VEC * G * CH;
    if (out_peers != nullptr) {
        if (exact) DLRMX_FWD_LAUNCH(true, true, tables, table_rows, indices, offsets, N, B, D, ld, total_bags, psw, out, stride_b, stride_t, err_flag, out_peers, lb);
        else DLRMX_FWD_LAUNCH(true, false, tables, table_rows, indices, offsets, N, B, D, ld, total_bags, psw, out, stride_b, stride_t, err_flag, out_peers, lb);
    } else {
        if (exact) DLRMX_FWD_LAUNCH(false, true, tables, table_rows, indices, offsets, N, B, D, ld, total_bags, psw, out, stride_b, stride_t, err_flag, nullptr, 1);
        else DLRMX_FWD_LAUNCH(false, false, tables, table_rows, indices, offsets, N, B, D, ld, total_bags, psw, out, stride_b, stride_t, err_flag, nullptr, 1);
    }
#undef DLRMX_FWD_LAUNCH
    return check_launch("embbag_fwd");
}

template <typename IdxT, typename TabT = float>
static int embbag_fwd(const TabT* const* tables, const int64_t* table_rows, const IdxT* indices,
                      const IdxT* offsets, int64_t N, int T, int B, int D, int64_t table_ld, const float* psw, float* out,
                      int64_t stride_b, int64_t stride_t, int* err_flag, float* const* out_peers, int lb,
                      dlrmx_stream_t stream) {
    DLRMX_REQUIRE(T >= 0 && B >= 0 && N >= 0, "embbag_fwd: negative size");
    DLRMX_REQUIRE(out_peers == nullptr || (lb > 0 && B % lb == 0), "embbag_fwd_scatter: B must be a multiple of lb > 0");
    DLRMX_REQUIRE((int64_t)T * B < (1ll << 31), "embbag_fwd: T*B must be < 2^31");
    RowGeom g;
    constexpr bool kBf16 = sizeof(TabT) == 2;     // bf16 rows: 8 elements (16 bytes) per lane when D % 8 == 0, else one
    DLRMX_REQUIRE(row_geometry(D, &g, kBf16 ? 8 : 4), "embbag_fwd: unsupported D=%d (need D<=512 if D%%4==0, else D<=128)", D);
    const int64_t ld = table_ld > 0 ? table_ld : D;      // row pitch of the tables in floats
    DLRMX_REQUIRE(ld >= D && (g.vec == 1 || ld % g.vec == 0), "embbag_fwd: table_ld must be >= D (and a multiple of the lane width %d)", g.vec);
    if (g.vec >= 4)
        DLRMX_REQUIRE(stride_b % 4 == 0 && stride_t % 4 == 0 && (out_peers != nullptr || ((uintptr_t)out & 15) == 0),
                      "embbag_fwd: out must be 16-B aligned with strides multiple of 4 when D%%4==0");
    if ((int64_t)T * B == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // Short bags (<= 2 indices on average, e.g. Criteo's one per bag): tiles of G bags, K in lockstep.
    // Long bags: ONE bag per group with 8 of its rows in flight, in CTAs of 224 threads at 4 CTAs per SM (28 warps, 69-72
    // registers, no spills) -- measured on B200 at 100 rows per bag (fraction of the measured copy peak; C3 shape / 8x the
    // bags / 512-byte rows): 0.80 / 0.87 / 1.06.  The variants that lost (256 x 3 CTAs: 0.61 / 0.78 / 0.99; two bags in
    // lockstep: 0.52 / 0.62 / 0.50; 16 rows in flight: 0.56 / 0.67 / 0.84; L2 bulk prefetch; register double buffer) are
    // recorded in DESIGN.md section 4 and no longer built.
    const bool short_bags = N <= 2 * (int64_t)T * B;
#define DLRMX_FWD_ARGS \
    tables, table_rows, indices, offsets, N, T, B, D, ld, psw, out, stride_b, stride_t, err_flag, out_peers, lb, st
#define DLRMX_FWD_CASE(VEC_, G_, CH_)                                                                       \
    if (g.vec == VEC_ && g.g == G_ && g.ch == CH_) {                                                        \
        if (short_bags) return launch_fwd<VEC_, G_, CH_, ((CH_ == 1 && VEC_ != 8) ? 4 : 2), 1, true, IdxT, 0, 256, TabT>(DLRMX_FWD_ARGS);  \
        if constexpr (CH_ == 1) return launch_fwd<VEC_, G_, CH_, 1, 8, false, IdxT, 4, 224, TabT>(DLRMX_FWD_ARGS); \
        return launch_fwd<VEC_, G_, CH_, 1, (CH_ == 2 ? 4 : 2), false, IdxT, 3, 256, TabT>(DLRMX_FWD_ARGS);           \
    }
    if constexpr (kBf16) {
        DLRMX_FWD_CASE(8, 1, 1) DLRMX_FWD_CASE(8, 2, 1) DLRMX_FWD_CASE(8, 4, 1) DLRMX_FWD_CASE(8, 8, 1)
        DLRMX_FWD_CASE(8, 16, 1) DLRMX_FWD_CASE(8, 32, 1) DLRMX_FWD_CASE(8, 32, 2)
    } else {
        DLRMX_FWD_CASE(4, 1, 1) DLRMX_FWD_CASE(4, 2, 1) DLRMX_FWD_CASE(4, 4, 1) DLRMX_FWD_CASE(4, 8, 1)
        DLRMX_FWD_CASE(4, 16, 1) DLRMX_FWD_CASE(4, 32, 1) DLRMX_FWD_CASE(4, 32, 2) DLRMX_FWD_CASE(4, 32, 4)
    }
    DLRMX_FWD_CASE(1, 1, 1) DLRMX_FWD_CASE(1, 2, 1) DLRMX_FWD_CASE(1, 4, 1) DLRMX_FWD_CASE(1, 8, 1)
    DLRMX_FWD_CASE(1, 16, 1) DLRMX_FWD_CASE(1, 32, 1) DLRMX_FWD_CASE(1, 32, 2) DLRMX_FWD_CASE(1, 32, 4)
#undef DLRMX_FWD_CASE
    set_error("embbag_fwd: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}

}  // namespace dlrmx

