// Table-batched embedding-bag forward (sum pooling) for sm_100a.
//
// Replaces nn.EmbeddingBag(mode="sum") as driven by DLRM_Net.apply_emb (dlrm_s_pytorch.py:526-587),
// the batched module call of dlrm_s_pytorch.py:589-591 and the yx kernel
// (yx_modfs/table_batched_embeddings_cuda_yx.cu:251-315).  Not a translation of the latter: that
// kernel gives one 16-B load in flight per lane and 32-bit addressing; here
//   * a group of G = D/4 lanes (a full warp at D=128) owns K consecutive bags and walks them in
//     LOCKSTEP: round r issues the r-th row of all K bags before any of them is consumed, so every
//     lane keeps K independent 16-B `ld.global.nc.L1::no_allocate` row loads in flight -- the
//     memory-level parallelism a random 512-B gather needs to approach HBM bandwidth (Criteo has one
//     index per bag, so unrolling along the bag would give exactly one load in flight);
//   * offsets / indices of the K bags are read with group-uniform (broadcast) loads: one 32-B sector
//     serves 8 bags, and they stay L1-resident across rounds;
//   * each bag is summed sequentially in index order with plain fp32 adds -> bit-exact with ATen;
//   * 64-bit row addressing, per-table base pointers, strided output so pooled rows land directly
//     in the [B, 1+T, D] interaction input (no torch.cat afterwards).
// HBM-bound: algorithmic bytes = N*D*4 (rows) + N*ib + T*B*ob + T*B*D*4 (pooled rows).
#include "common.cuh"

namespace dlrmx {

template <int VEC, int G, int CH, int K, typename IdxT>
__global__ void __launch_bounds__(256, (K * CH >= 8) ? 2 : 3)
embbag_fwd_kernel(const float* const* __restrict__ tables, const int64_t* __restrict__ table_rows,
                  const IdxT* __restrict__ indices, const IdxT* __restrict__ offsets, int64_t N,
                  int B, int D, int64_t total_bags, const float* __restrict__ psw,
                  float* __restrict__ out, int64_t stride_b, int64_t stride_t, int* err_flag) {
    const int gl = threadIdx.x & (G - 1);
    const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    const int64_t bag0 = group * K;
    if (bag0 >= total_bags) return;
    const int col0 = gl * VEC;

    IdxT start[K];
    int len[K];
    int tk[K];
    int maxlen = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int64_t bag = bag0 + k;
        if (bag < total_bags) {
            const IdxT s = offsets[bag];
            const IdxT e = (bag + 1 < total_bags) ? offsets[bag + 1] : (IdxT)N;
            start[k] = s;
            len[k] = (int)(e - s);
            tk[k] = (int)(bag / B);
        } else {
            start[k] = 0;
            len[k] = 0;
            tk[k] = 0;
        }
        maxlen = max(maxlen, len[k]);
    }

    RowVec<VEC> acc[K][CH];
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
        for (int c = 0; c < CH; ++c) acc[k][c].zero();

    for (int r = 0; r < maxlen; ++r) {
        IdxT ix[K];
#pragma unroll
        for (int k = 0; k < K; ++k) ix[k] = (r < len[k]) ? indices[start[k] + r] : (IdxT)-1;

        RowVec<VEC> row[K][CH];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (ix[k] >= 0) {
                if (table_rows != nullptr && (int64_t)ix[k] >= table_rows[tk[k]]) {
                    if (err_flag != nullptr && gl == 0) atomicOr(err_flag, 1);
                    ix[k] = -1;
                }
            } else if (r < len[k]) {  // negative index in the data
                if (err_flag != nullptr && gl == 0) atomicOr(err_flag, 1);
            }
            if (ix[k] >= 0) {
                const float* p = tables[tk[k]] + (int64_t)ix[k] * D + col0;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (CH == 1 || col0 + c * G * VEC < D) row[k][c].load_stream(p + c * G * VEC);
            }
        }
        if (psw == nullptr) {
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (ix[k] >= 0) {
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (CH == 1 || col0 + c * G * VEC < D) acc[k][c].add(row[k][c]);
                }
        } else {
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (ix[k] >= 0) {
                    const float w = __ldg(psw + start[k] + r);
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (CH == 1 || col0 + c * G * VEC < D) acc[k][c].add_scaled(row[k][c], w);
                }
        }
    }

#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int64_t bag = bag0 + k;
        if (bag < total_bags) {
            const int64_t b = bag - (int64_t)tk[k] * B;
            float* o = out + b * stride_b + (int64_t)tk[k] * stride_t + col0;
#pragma unroll
            for (int c = 0; c < CH; ++c)
                if (col0 + c * G * VEC < D) acc[k][c].store(o + c * G * VEC);
        }
    }
}

template <int VEC, int G, int CH, int K, typename IdxT>
static int launch_fwd(const float* const* tables, const int64_t* table_rows, const IdxT* indices,
                      const IdxT* offsets, int64_t N, int T, int B, int D, const float* psw, float* out,
                      int64_t stride_b, int64_t stride_t, int* err_flag, cudaStream_t st) {
    const int64_t total_bags = (int64_t)T * B;
    const int64_t groups = ceil_div(total_bags, K);
    const int threads = 256;
    const int64_t blocks = ceil_div(groups * G, threads);
    if (blocks == 0) return DLRMX_OK;
    embbag_fwd_kernel<VEC, G, CH, K, IdxT><<<(unsigned)blocks, threads, 0, st>>>(
        tables, table_rows, indices, offsets, N, B, D, total_bags, psw, out, stride_b, stride_t, err_flag);
    return check_launch("embbag_fwd");
}

template <typename IdxT>
static int embbag_fwd(const float* const* tables, const int64_t* table_rows, const IdxT* indices,
                      const IdxT* offsets, int64_t N, int T, int B, int D, const float* psw, float* out,
                      int64_t stride_b, int64_t stride_t, int* err_flag, dlrmx_stream_t stream) {
    DLRMX_REQUIRE(T >= 0 && B >= 0 && N >= 0, "embbag_fwd: negative size");
    DLRMX_REQUIRE((int64_t)T * B < (1ll << 31), "embbag_fwd: T*B must be < 2^31");
    RowGeom g;
    DLRMX_REQUIRE(row_geometry(D, &g), "embbag_fwd: unsupported D=%d (need D<=512 if D%%4==0, else D<=128)", D);
    if (g.vec == 4)
        DLRMX_REQUIRE(stride_b % 4 == 0 && stride_t % 4 == 0 && ((uintptr_t)out & 15) == 0,
                      "embbag_fwd: out must be 16-B aligned with strides multiple of 4 when D%%4==0");
    if ((int64_t)T * B == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
#define DLRMX_FWD_CASE(VEC_, G_, CH_)                                                                  \
    if (g.vec == VEC_ && g.g == G_ && g.ch == CH_) {                                                   \
        if (CH_ == 1 && get_tunable(DLRMX_TUNE_FWD_K) == 4)                                            \
            return launch_fwd<VEC_, G_, CH_, 4, IdxT>(tables, table_rows, indices, offsets, N, T, B, D, psw, \
                                                      out, stride_b, stride_t, err_flag, st);          \
        return launch_fwd<VEC_, G_, CH_, 8 / CH_, IdxT>(tables, table_rows, indices, offsets, N, T, B, D, \
                                                        psw, out, stride_b, stride_t, err_flag, st);   \
    }
    DLRMX_FWD_CASE(4, 1, 1) DLRMX_FWD_CASE(4, 2, 1) DLRMX_FWD_CASE(4, 4, 1) DLRMX_FWD_CASE(4, 8, 1)
    DLRMX_FWD_CASE(4, 16, 1) DLRMX_FWD_CASE(4, 32, 1) DLRMX_FWD_CASE(4, 32, 2) DLRMX_FWD_CASE(4, 32, 4)
    DLRMX_FWD_CASE(1, 1, 1) DLRMX_FWD_CASE(1, 2, 1) DLRMX_FWD_CASE(1, 4, 1) DLRMX_FWD_CASE(1, 8, 1)
    DLRMX_FWD_CASE(1, 16, 1) DLRMX_FWD_CASE(1, 32, 1) DLRMX_FWD_CASE(1, 32, 2) DLRMX_FWD_CASE(1, 32, 4)
#undef DLRMX_FWD_CASE
    set_error("embbag_fwd: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}

}  // namespace dlrmx

extern "C" {
int dlrmx_embbag_fwd_i32(const float* const* tables, const int64_t* table_rows, const int32_t* indices,
                         const int32_t* offsets, int64_t N, int T, int B, int D, const float* psw,
                         float* out, int64_t stride_b, int64_t stride_t, int* err_flag,
                         dlrmx_stream_t stream) {
    return dlrmx::embbag_fwd<int32_t>(tables, table_rows, indices, offsets, N, T, B, D, psw, out, stride_b,
                                      stride_t, err_flag, stream);
}
int dlrmx_embbag_fwd_i64(const float* const* tables, const int64_t* table_rows, const int64_t* indices,
                         const int64_t* offsets, int64_t N, int T, int B, int D, const float* psw,
                         float* out, int64_t stride_b, int64_t stride_t, int* err_flag,
                         dlrmx_stream_t stream) {
    return dlrmx::embbag_fwd<int64_t>(tables, table_rows, indices, offsets, N, T, B, D, psw, out, stride_b,
                                      stride_t, err_flag, stream);
}
}
