// Dot feature interaction, forward and backward, fp32 (parity mode).  sm_100a.
//
// Replaces DLRM_Net.interact_features('dot') (dlrm_s_pytorch.py:627-659): cat -> bmm(T, T^t) ->
// tril_indices gather -> cat with x, i.e. 5 kernels and 3 intermediates, by ONE kernel each way that
// reads the [B, F, D] feature block once (feature 0 = bottom-MLP output, features 1.. = pooled rows that
// the lookup kernel already wrote in place) and writes the top-MLP input [B, D + F(F-1)/2] directly.
//   forward : per sample the F x D block is staged in shared memory (row stride D+VEC to spread banks);
//             each lane owns a 4x4 block of the lower triangle of Z = T T^t and runs the k loop with
//             16 accumulators (8 shared loads per 16 FMAs); results are staged and written coalesced.
//   backward: dT = (dZ + dZ^t) T.  The symmetric F x F coefficient block S lives in shared memory (read
//             with broadcast 16-B loads), each lane owns one 16-B column slice of all F rows.
// This is the fp32 CUDA-core version that meets the 1e-5 / 1e-6 parity bound; HBM traffic is
// B*F*D*4 in + B*(D+nz)*4 out.
#include "common.cuh"

namespace dlrmx {

constexpr int kIxThreads = 128;

__host__ __device__ inline int tri_count(int F, int itself) { return itself ? F * (F + 1) / 2 : F * (F - 1) / 2; }
// position of pair (i, j), j <= i (itself) or j < i, in torch.tril_indices row-major order
__device__ __forceinline__ int tri_index(int i, int j, int itself) {
    return itself ? i * (i + 1) / 2 + j : i * (i - 1) / 2 + j;
}

template <int VEC, int TPS>
__global__ void __launch_bounds__(kIxThreads)
interact_fwd_kernel(const float* __restrict__ feat, int64_t fsb, int B, int F, int D, int itself,
                    float* __restrict__ out, int64_t ostride) {
    extern __shared__ __align__(16) float smem[];
    constexpr int SPB = kIxThreads / TPS;
    const int DS = D + VEC;
    const int nz = tri_count(F, itself);
    const int per_sample = F * DS + ((nz + 3) & ~3);
    const int sl = threadIdx.x / TPS, tl = threadIdx.x % TPS;
    const int64_t b = (int64_t)blockIdx.x * SPB + sl;
    float* tile = smem + (size_t)sl * per_sample;
    float* zrow = tile + F * DS;
    const bool live = b < B;

    if (live) {
        const float* src = feat + b * fsb;
        for (int e = tl * VEC; e < F * D; e += TPS * VEC) {
            const int f = e / D, k = e - f * D;
            RowVec<VEC> v;
            v.load_ro(src + e);
            v.store(tile + f * DS + k);
        }
    }
    __syncthreads();
    if (live) {
        const int nb = (F + 3) / 4;
        const int nblocks = nb * (nb + 1) / 2;
        for (int blk = tl; blk < nblocks; blk += TPS) {
            int bi = 0;
            while ((bi + 1) * (bi + 2) / 2 <= blk) ++bi;
            const int bj = blk - bi * (bi + 1) / 2;
            const float* ar[4];
            const float* br[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                ar[r] = tile + min(4 * bi + r, F - 1) * DS;
                br[r] = tile + min(4 * bj + r, F - 1) * DS;
            }
            float c[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int s = 0; s < 4; ++s) c[r][s] = 0.f;
            for (int k = 0; k < D; k += VEC) {
                float a[4][VEC], bb[4][VEC];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if constexpr (VEC == 4) {
                        const float4 va = *reinterpret_cast<const float4*>(ar[r] + k);
                        const float4 vb = *reinterpret_cast<const float4*>(br[r] + k);
                        a[r][0] = va.x; a[r][1 % VEC] = va.y; a[r][2 % VEC] = va.z; a[r][3 % VEC] = va.w;
                        bb[r][0] = vb.x; bb[r][1 % VEC] = vb.y; bb[r][2 % VEC] = vb.z; bb[r][3 % VEC] = vb.w;
                    } else {
                        a[r][0] = ar[r][k];
                        bb[r][0] = br[r][k];
                    }
                }
#pragma unroll
                for (int v = 0; v < VEC; ++v)
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int s = 0; s < 4; ++s) c[r][s] = __fmaf_rn(a[r][v], bb[s][v], c[r][s]);
            }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    const int i = 4 * bi + r, j = 4 * bj + s;
                    if (i < F && (itself ? j <= i : j < i)) zrow[tri_index(i, j, itself)] = c[r][s];
                }
        }
    }
    __syncthreads();
    if (live) {
        float* o = out + b * ostride;
        for (int k = tl; k < D; k += TPS) o[k] = tile[k];
        for (int q = tl; q < nz; q += TPS) o[D + q] = zrow[q];
    }
}

template <int VEC, int G, int IB>
__global__ void __launch_bounds__(kIxThreads)
interact_bwd_kernel(const float* __restrict__ feat, int64_t fsb, int B, int F, int D, int itself,
                    const float* __restrict__ dout, int64_t dstride, float* __restrict__ dfeat, int64_t dfsb) {
    extern __shared__ __align__(16) float smem[];
    constexpr int SPB = kIxThreads / G;
    const int FP = (F + 3) & ~3;
    const int sl = threadIdx.x / G, gl = threadIdx.x % G;
    const int64_t b = (int64_t)blockIdx.x * SPB + sl;
    float* S = smem + (size_t)sl * F * FP;  // S[i][j], row stride FP
    const bool live = b < B;
    const int col0 = gl * VEC;
    const bool colok = col0 < D;

    if (live) {
        const float* dz = dout + b * dstride + D;
        for (int e = gl; e < F * FP; e += G) {
            const int i = e / FP, j = e - i * FP;
            float v = 0.f;
            if (j < F) {
                if (i == j) v = itself ? 2.f * dz[tri_index(i, i, 1)] : 0.f;
                else v = dz[tri_index(max(i, j), min(i, j), itself)];
            }
            S[e] = v;
        }
    }
    __syncthreads();
    if (!live || !colok) return;
    const float* src = feat + b * fsb + col0;
    float* dst = dfeat + b * dfsb + col0;
    for (int i0 = 0; i0 < F; i0 += IB) {
        RowVec<VEC> acc[IB];
#pragma unroll
        for (int i = 0; i < IB; ++i) acc[i].zero();
        for (int j4 = 0; j4 < F; j4 += 4) {
            RowVec<VEC> t[4];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                if (j4 + jj < F) t[jj].load_ro(src + (int64_t)(j4 + jj) * D);
                else t[jj].zero();
            }
#pragma unroll
            for (int i = 0; i < IB; ++i) {
                if (i0 + i < F) {
                    const float4 s4 = *reinterpret_cast<const float4*>(S + (i0 + i) * FP + j4);
                    acc[i].map(t[0], [s4](float a, float tv) { return __fmaf_rn(s4.x, tv, a); });
                    acc[i].map(t[1], [s4](float a, float tv) { return __fmaf_rn(s4.y, tv, a); });
                    acc[i].map(t[2], [s4](float a, float tv) { return __fmaf_rn(s4.z, tv, a); });
                    acc[i].map(t[3], [s4](float a, float tv) { return __fmaf_rn(s4.w, tv, a); });
                }
            }
        }
#pragma unroll
        for (int i = 0; i < IB; ++i) {
            if (i0 + i < F) {
                if (i0 + i == 0) {  // dense pass-through: R[:, 0:D] = x
                    RowVec<VEC> dx;
                    dx.load_unaligned(dout + b * dstride + col0);
                    acc[i].add(dx);
                }
                acc[i].store(dst + (int64_t)(i0 + i) * D);
            }
        }
    }
}

template <int VEC, int TPS>
static int launch_ix_fwd(const float* feat, int64_t fsb, int B, int F, int D, int itself, float* out,
                         int64_t ostride, cudaStream_t st) {
    constexpr int SPB = kIxThreads / TPS;
    const int nz = tri_count(F, itself);
    const size_t smem = (size_t)SPB * (F * (D + VEC) + ((nz + 3) & ~3)) * 4;
    DLRMX_REQUIRE(smem <= 200 * 1024, "interact_dot_fwd: F=%d D=%d needs %zu B of shared memory", F, D, smem);
    auto kern = interact_fwd_kernel<VEC, TPS>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<(unsigned)ceil_div(B, SPB), kIxThreads, smem, st>>>(feat, fsb, B, F, D, itself, out, ostride);
    return check_launch("interact_dot_fwd");
}

template <int VEC, int G>
static int launch_ix_bwd(const float* feat, int64_t fsb, int B, int F, int D, int itself, const float* dout,
                         int64_t dstride, float* dfeat, int64_t dfsb, cudaStream_t st) {
    constexpr int SPB = kIxThreads / G;
    constexpr int IB = 14;
    const int FP = (F + 3) & ~3;
    const size_t smem = (size_t)SPB * F * FP * 4;
    DLRMX_REQUIRE(smem <= 200 * 1024, "interact_dot_bwd: F=%d needs %zu B of shared memory", F, smem);
    auto kern = interact_bwd_kernel<VEC, G, IB>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<(unsigned)ceil_div(B, SPB), kIxThreads, smem, st>>>(feat, fsb, B, F, D, itself, dout, dstride, dfeat, dfsb);
    return check_launch("interact_dot_bwd");
}

// tensor-core (3xTF32) kernels of interact_mma.cu; return 1 when they took the shape
int interact_fwd_mma(const float* feat, int64_t fsb, int B, int F, int D, int itself, float* out, int64_t ostride,
                     cudaStream_t st, int* rc);
int interact_bwd_mma(const float* feat, int64_t fsb, int B, int F, int D, int itself, const float* dout,
                     int64_t dstride, float* dfeat, int64_t dfsb, cudaStream_t st, int* rc);
int interact_fwd_mixed(const float* feat, int64_t fsb, int64_t fst, const void* x_bf16, int B, int F, int D, int itself,
                       void* out_bf16, int64_t ostride, cudaStream_t st);
int interact_bwd_mixed(const float* feat, int64_t fsb, int64_t fst, const void* x_bf16, int B, int F, int D, int itself,
                       const void* dout_bf16, int64_t dstride, float* dfeat, int64_t dfsb, void* dx_bf16, int relu_x,
                       cudaStream_t st);
enum { DLRMX_TUNE_IX_CUDACORE = 2 };   // tunable 2 = 1 forces the CUDA-core kernels (tests compare both)

}  // namespace dlrmx

extern "C" {
int dlrmx_interact_dot_fwd_mixed(const float* feat, int64_t feat_stride_b, int64_t feat_stride_f, const void* x_bf16, int B, int F, int D,
                                 int itself, void* out_bf16, int64_t out_stride, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && F >= 2 && D >= 1 && feat != nullptr && x_bf16 != nullptr && out_bf16 != nullptr,
                  "interact_dot_fwd_mixed: bad arguments");
    if (B == 0) return DLRMX_OK;
    return interact_fwd_mixed(feat, feat_stride_b, feat_stride_f, x_bf16, B, F, D, itself, out_bf16, out_stride, (cudaStream_t)stream);
}

int dlrmx_interact_dot_bwd_mixed(const float* feat, int64_t feat_stride_b, int64_t feat_stride_f, const void* x_bf16, int B, int F, int D,
                                 int itself, const void* dout_bf16, int64_t dout_stride, float* dfeat,
                                 int64_t dfeat_stride_b, void* dx_bf16, int relu_x, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && F >= 2 && D >= 1 && feat != nullptr && x_bf16 != nullptr && dout_bf16 != nullptr &&
                      dfeat != nullptr && dx_bf16 != nullptr, "interact_dot_bwd_mixed: bad arguments");
    if (B == 0) return DLRMX_OK;
    return interact_bwd_mixed(feat, feat_stride_b, feat_stride_f, x_bf16, B, F, D, itself, dout_bf16, dout_stride, dfeat,
                              dfeat_stride_b, dx_bf16, relu_x, (cudaStream_t)stream);
}

int dlrmx_interact_dot_fwd(const float* feat, int64_t feat_stride_b, int B, int F, int D, int itself,
                           float* out, int64_t out_stride, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && F >= 1 && D >= 1, "interact_dot_fwd: bad sizes");
    DLRMX_REQUIRE(out_stride >= D + tri_count(F, itself), "interact_dot_fwd: out_stride too small");
    if (B == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (get_tunable(DLRMX_TUNE_IX_CUDACORE) == 0) {
        int rc = 0;
        if (interact_fwd_mma(feat, feat_stride_b, B, F, D, itself, out, out_stride, st, &rc)) return rc;
    }
    const int nb = (F + 3) / 4, nblocks = nb * (nb + 1) / 2;
    if (D % 4 == 0) {
        DLRMX_REQUIRE(feat_stride_b % 4 == 0 && ((uintptr_t)feat & 15) == 0,
                      "interact_dot_fwd: feat must be 16-B aligned, stride multiple of 4");
        if (nblocks <= 4) return launch_ix_fwd<4, 4>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
        if (nblocks <= 8) return launch_ix_fwd<4, 8>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
        if (nblocks <= 16) return launch_ix_fwd<4, 16>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
        return launch_ix_fwd<4, 32>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
    }
    if (nblocks <= 4) return launch_ix_fwd<1, 4>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
    if (nblocks <= 8) return launch_ix_fwd<1, 8>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
    if (nblocks <= 16) return launch_ix_fwd<1, 16>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
    return launch_ix_fwd<1, 32>(feat, feat_stride_b, B, F, D, itself, out, out_stride, st);
}

int dlrmx_interact_dot_bwd(const float* feat, int64_t feat_stride_b, int B, int F, int D, int itself,
                           const float* dout, int64_t dout_stride, float* dfeat, int64_t dfeat_stride_b,
                           dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && F >= 1 && D >= 1, "interact_dot_bwd: bad sizes");
    if (B == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (get_tunable(DLRMX_TUNE_IX_CUDACORE) == 0) {
        int rc = 0;
        if (interact_bwd_mma(feat, feat_stride_b, B, F, D, itself, dout, dout_stride, dfeat, dfeat_stride_b, st, &rc))
            return rc;
    }
    RowGeom g;
    DLRMX_REQUIRE(row_geometry(D, &g) && g.ch == 1, "interact_dot_bwd: unsupported D=%d (D%%4==0 and D<=128, or D<=32)", D);
    if (g.vec == 4)
        DLRMX_REQUIRE(feat_stride_b % 4 == 0 && dfeat_stride_b % 4 == 0 && ((uintptr_t)feat & 15) == 0 &&
                          ((uintptr_t)dfeat & 15) == 0,
                      "interact_dot_bwd: feat/dfeat must be 16-B aligned with strides multiple of 4");
#define DLRMX_IXB(V, G_)                                                                              \
    if (g.vec == V && g.g == G_)                                                                      \
        return launch_ix_bwd<V, G_>(feat, feat_stride_b, B, F, D, itself, dout, dout_stride, dfeat, dfeat_stride_b, st);
    DLRMX_IXB(4, 1) DLRMX_IXB(4, 2) DLRMX_IXB(4, 4) DLRMX_IXB(4, 8) DLRMX_IXB(4, 16) DLRMX_IXB(4, 32)
    DLRMX_IXB(1, 1) DLRMX_IXB(1, 2) DLRMX_IXB(1, 4) DLRMX_IXB(1, 8) DLRMX_IXB(1, 16) DLRMX_IXB(1, 32)
#undef DLRMX_IXB
    set_error("interact_dot_bwd: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}
}
