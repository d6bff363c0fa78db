// Binary cross-entropy with mean reduction, forward and backward in one kernel each.  sm_100a.
//
// Replaces dlrm.loss_fn = torch.nn.BCELoss(reduction="mean") (dlrm_s_pytorch.py:505-507, called at 170-173) and its
// autograd backward, which on the GPU are five launches per step (element-wise loss, mean reduce, fill, scale,
// element-wise gradient) at the tail of a step whose prediction is a [B, 1] vector.  Same arithmetic as ATen:
//   loss_i = -( y * max(log p, -100) + (1 - y) * max(log(1 - p), -100) ),   loss = sum_i loss_i / B
//   dp_i   = dloss * (p - y) / max((1 - p) * p, 1e-12) / B
// The sum is deterministic: fixed tree inside a CTA, per-CTA partials added in index order by the CTA that finishes last.
#include "common.cuh"

namespace dlrmx {

constexpr int kLossThreads = 256, kLossPerThread = 4;

__global__ void __launch_bounds__(kLossThreads)
bce_mean_fwd_kernel(const float* __restrict__ p, const float* __restrict__ y, int64_t n, float* __restrict__ partial,
                    unsigned int* __restrict__ counter, float* __restrict__ loss) {
    __shared__ float red[kLossThreads / 32];
    __shared__ bool is_last;
    const int64_t i0 = ((int64_t)blockIdx.x * kLossThreads + threadIdx.x) * kLossPerThread;
    float s = 0.f;
#pragma unroll
    for (int u = 0; u < kLossPerThread; ++u) {
        const int64_t i = i0 + u;
        if (i < n) {
            const float pi = p[i], yi = y[i];
            const float l1 = fmaxf(logf(pi), -100.f), l0 = fmaxf(logf(1.f - pi), -100.f);
            s += -(yi * l1 + (1.f - yi) * l0);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kLossThreads / 32; ++w) t += red[w];
        partial[blockIdx.x] = t;
        __threadfence();
        const unsigned int done = atomicAdd(counter, 1u);
        is_last = (done == gridDim.x - 1);
        if (is_last) *counter = 0u;            // ready for the next call
    }
    __syncthreads();
    if (!is_last || threadIdx.x != 0) return;
    __threadfence();
    float tot = 0.f;
    for (unsigned int b = 0; b < gridDim.x; ++b) tot += __ldcg(partial + b);
    loss[0] = tot / (float)n;
}

__global__ void __launch_bounds__(kLossThreads)
bce_mean_bwd_kernel(const float* __restrict__ p, const float* __restrict__ y, const float* __restrict__ dloss,
                    int64_t n, float* __restrict__ dp) {
    const int64_t i = (int64_t)blockIdx.x * kLossThreads + threadIdx.x;
    if (i >= n) return;
    const float pi = p[i];
    const float g = __ldg(dloss);
    dp[i] = g * (pi - y[i]) / fmaxf((1.f - pi) * pi, 1e-12f) / (float)n;
}

}  // namespace dlrmx

extern "C" {
size_t dlrmx_bce_mean_workspace_bytes(int64_t n) {
    return (size_t)(dlrmx::ceil_div(n > 0 ? n : 1, dlrmx::kLossThreads * dlrmx::kLossPerThread)) * sizeof(float);
}

int dlrmx_bce_mean_fwd(const float* p, const float* target, int64_t n, float* workspace, unsigned int* counter,
                       float* loss, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(n > 0 && p != nullptr && target != nullptr && workspace != nullptr && counter != nullptr && loss != nullptr,
                  "bce_mean_fwd: bad arguments");
    const unsigned blocks = (unsigned)ceil_div(n, kLossThreads * kLossPerThread);
    bce_mean_fwd_kernel<<<blocks, kLossThreads, 0, (cudaStream_t)stream>>>(p, target, n, workspace, counter, loss);
    return check_launch("bce_mean_fwd");
}

int dlrmx_bce_mean_bwd(const float* p, const float* target, const float* dloss, int64_t n, float* dp,
                       dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(n > 0 && p != nullptr && target != nullptr && dloss != nullptr && dp != nullptr,
                  "bce_mean_bwd: bad arguments");
    bce_mean_bwd_kernel<<<(unsigned)ceil_div(n, kLossThreads), kLossThreads, 0, (cudaStream_t)stream>>>(p, target, dloss,
                                                                                                      n, dp);
    return check_launch("bce_mean_bwd");
}
}
