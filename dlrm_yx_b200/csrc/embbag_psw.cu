// Backward of the per-sample-weighted embedding bag (weighted pooling).  sm_100a.
//
// Reference: DLRM_Net.apply_emb gathers one weight per looked-up index from a per-row vector,
// `per_sample_weights = v_W_l[k].gather(0, indices)` (dlrm_s_pytorch.py:544-547), and nn.EmbeddingBag(mode="sum")
// pools `out[b] = sum_l psw[l] * W[idx[l]]`; with weighted_pooling="learned" v_W_l[k] is a Parameter (475-478), so
// autograd needs BOTH halves of the backward:
//   d psw[l]   = < grad_out[bag(l)], W[idx[l]] >                              -> psw_grad_kernel
//   d W[r]     = sum over lookups l of row r of  psw[l] * grad_out[bag(l)]     -> bwd_coalesce_weighted_kernel
// The second is the weighted form of bwd_coalesce_kernel: lookups are sorted by row with their ORIGINAL POSITION as the
// payload (the sort is run with every lookup as its own bag), a position -> bag map finds the gradient row, and runs
// of equal rows are summed sequentially in sorted order -- no [N, D] expansion of the gradient is ever written.
// Per-table path only (the reference refuses weighted pooling for the batched modules, dlrm_s_pytorch.py:322, 337).
#include "common.cuh"

namespace dlrmx {

// one warp per bag: the gradient row of the bag stays in registers, one table row per lookup is read and dotted with it
template <typename IdxT>
__global__ void __launch_bounds__(256)
psw_grad_kernel(const float* __restrict__ table, int64_t ld, int64_t rows, const IdxT* __restrict__ indices,
                const IdxT* __restrict__ offsets, int64_t N, int B, int D, const float* __restrict__ grad,
                int64_t stride_b, float* __restrict__ dpsw) {
    const int lane = threadIdx.x & 31;
    const int64_t bag = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (bag >= B) return;
    const int64_t s = (int64_t)offsets[bag];
    const int64_t e = (bag + 1 < B) ? (int64_t)offsets[bag + 1] : N;
    const float* g = grad + bag * stride_b;
    for (int64_t l = s; l < e; ++l) {
        const int64_t ix = (int64_t)indices[l];
        float acc = 0.f;
        if (ix >= 0 && ix < rows) {
            const float* w = table + ix * ld;
            for (int c = lane; c < D; c += 32) acc = fmaf(__ldg(g + c), __ldg(w + c), acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) dpsw[l] = acc;
    }
}

// one warp per run of equal rows: gsum[u, :] = sum_p psw[pos_p] * grad[bag(pos_p), :], sequential in sorted order
__global__ void __launch_bounds__(256)
bwd_coalesce_weighted_kernel(int D, const float* __restrict__ grad, int64_t stride_b,
                             const int32_t* __restrict__ sorted_pos, const int32_t* __restrict__ pos2bag,
                             const float* __restrict__ psw, const int32_t* __restrict__ run_off,
                             const int32_t* __restrict__ num_unique, float* __restrict__ gsum) {
    const int lane = threadIdx.x & 31;
    const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (u >= *num_unique) return;
    const int p0 = run_off[u], p1 = run_off[u + 1];
    for (int c0 = 0; c0 < D; c0 += 128) {           // 4 columns per lane and pass
        float a[4] = {0.f, 0.f, 0.f, 0.f};
        for (int p = p0; p < p1; ++p) {
            const int pos = sorted_pos[p];
            const float w = __ldg(psw + pos);
            const float* g = grad + (int64_t)pos2bag[pos] * stride_b;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = c0 + lane + 32 * q;
                if (c < D) a[q] = __fadd_rn(a[q], __fmul_rn(w, __ldg(g + c)));     // (w * g) rounded, then added: ATen's order
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int c = c0 + lane + 32 * q;
            if (c < D) gsum[u * D + c] = a[q];
        }
    }
}

}  // namespace dlrmx

extern "C" {
int dlrmx_embbag_psw_grad_i32(const float* table, int64_t table_ld, int64_t rows, const int32_t* indices,
                              const int32_t* offsets, int64_t N, int B, int D, const float* grad, int64_t stride_b,
                              float* dpsw, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && N >= 0 && D > 0, "embbag_psw_grad: bad sizes");
    if (B == 0 || N == 0) return DLRMX_OK;
    psw_grad_kernel<int32_t><<<(unsigned)ceil_div((int64_t)B * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        table, table_ld > 0 ? table_ld : D, rows, indices, offsets, N, B, D, grad, stride_b, dpsw);
    return check_launch("embbag_psw_grad");
}
int dlrmx_embbag_psw_grad_i64(const float* table, int64_t table_ld, int64_t rows, const int64_t* indices,
                              const int64_t* offsets, int64_t N, int B, int D, const float* grad, int64_t stride_b,
                              float* dpsw, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && N >= 0 && D > 0, "embbag_psw_grad: bad sizes");
    if (B == 0 || N == 0) return DLRMX_OK;
    psw_grad_kernel<int64_t><<<(unsigned)ceil_div((int64_t)B * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        table, table_ld > 0 ? table_ld : D, rows, indices, offsets, N, B, D, grad, stride_b, dpsw);
    return check_launch("embbag_psw_grad");
}
int dlrmx_bwd_coalesce_weighted(int D, const float* grad, int64_t stride_b, const int32_t* sorted_pos,
                                const int32_t* pos2bag, const float* psw, const int32_t* run_off,
                                const int32_t* num_unique, int64_t N, float* gsum, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(D > 0 && N >= 0 && N < (1ll << 31), "bwd_coalesce_weighted: bad sizes");
    if (N == 0) return DLRMX_OK;
    bwd_coalesce_weighted_kernel<<<(unsigned)ceil_div(N * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        D, grad, stride_b, sorted_pos, pos2bag, psw, run_off, num_unique, gsum);
    return check_launch("bwd_coalesce_weighted");
}
}
