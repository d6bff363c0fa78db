// Elementwise / reduction kernels around the MLP-tower GEMMs (create_mlp / apply_mlp, dlrm_s_pytorch.py:227-265,
// 518-524) and the dense half of the optimizers (torch.optim.SGD, dlrm_s_pytorch.py:1640; RWSAdagrad.step dense
// branch, optim/rwsadagrad.py:115-118).  sm_100a.
//
// The GEMMs themselves are plain library GEMMs on the tensor cores (bf16 in, fp32 accumulate, bias + ReLU in the
// cuBLASLt epilogue); what PyTorch would add around them -- ReLU backward, bias-gradient reduction, dtype casts, 4
// optimizer kernels per parameter tensor -- is fused here into two HBM-bound passes:
//   relu_bwd_bias_grad : dz = dy * (y > 0) (bf16) and the column sums of dz (the bias gradient) in ONE read of dy, y:
//                        algorithmic bytes 3 * B*N*2; column sums are per-row-chunk partials added later in a fixed
//                        order (deterministic).
//   dense_adagrad/sgd  : ONE launch over the flat parameter buffer of all layers: acc += g^2; p -= lr*g/(sqrt(acc)+eps)
//                        (or p -= lr*g), refreshes the bf16 shadow the GEMMs read, and clears g for the next step:
//                        20 bytes per parameter.
#include <cuda_bf16.h>

#include "common.cuh"

namespace dlrmx {

constexpr int kRbThreads = 256;
constexpr int kRbRowLanes = 8;                     // row lanes per CTA
constexpr int kRbColVecs = kRbThreads / kRbRowLanes;  // 32 column vectors of 8 bf16 = 256 columns per CTA

struct alignas(16) Bf16x8 {
    __nv_bfloat162 v[4];
};

__device__ __forceinline__ Bf16x8 ld_stream16(const __nv_bfloat16* p) {   // read-once data: do not pollute L1
    Bf16x8 r;
    uint32_t* u = reinterpret_cast<uint32_t*>(&r);
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "l"(p));
    return r;
}

__global__ void __launch_bounds__(kRbThreads)
relu_bwd_bias_grad_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                          __nv_bfloat16* __restrict__ dz, float* __restrict__ db_partial, float* __restrict__ db,
                          unsigned int* __restrict__ counters, int B, int N, int rows_per_chunk, int relu) {
    __shared__ float red[kRbRowLanes][kRbColVecs * 8 + 8];
    __shared__ bool is_last;
    const int cv = threadIdx.x % kRbColVecs, rl = threadIdx.x / kRbColVecs;
    const int col = (blockIdx.x * kRbColVecs + cv) * 8;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(B, r0 + rows_per_chunk);
    float s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = 0.f;
    if (col < N) {
        constexpr int U = 4;                        // rows in flight per thread: 2*U 16-B loads before the first use
        for (int rb = r0 + rl; rb < r1; rb += kRbRowLanes * U) {
            Bf16x8 g[U], a[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = rb + u * kRbRowLanes;
                if (r < r1) {
                    const size_t off = (size_t)r * N + col;
                    g[u] = ld_stream16(dy + off);
                    if (relu) a[u] = ld_stream16(y + off);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = rb + u * kRbRowLanes;
                if (r < r1) {
                    if (relu) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float2 av = __bfloat1622float2(a[u].v[i]);
                            float2 gv = __bfloat1622float2(g[u].v[i]);
                            gv.x = av.x > 0.f ? gv.x : 0.f;
                            gv.y = av.y > 0.f ? gv.y : 0.f;
                            g[u].v[i] = __float22bfloat162_rn(gv);
                        }
                        *reinterpret_cast<Bf16x8*>(dz + (size_t)r * N + col) = g[u];
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float2 gv = __bfloat1622float2(g[u].v[i]);
                        s[2 * i] += gv.x;
                        s[2 * i + 1] += gv.y;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) red[rl][cv * 8 + i] = s[i];
    __syncthreads();
    for (int c = threadIdx.x; c < kRbColVecs * 8; c += kRbThreads) {
        const int gc = blockIdx.x * kRbColVecs * 8 + c;
        if (gc < N) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < kRbRowLanes; ++q) t += red[q][c];
            db_partial[(size_t)blockIdx.y * N + gc] = t;
        }
    }
    if (db == nullptr) return;
    // Second stage without a second launch: the CTA that finishes LAST for this column block adds the per-chunk
    // partials in a fixed order (row lane q takes chunks q, q+8, ...; the 8 lane sums are then added in lane
    // order), so the result does not depend on which CTA that is.  The counter is left at zero for the next call.
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int done = atomicAdd(&counters[blockIdx.x], 1u);
        is_last = (done == gridDim.y - 1);
        if (is_last) counters[blockIdx.x] = 0u;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = 0.f;
    if (col < N) {
        constexpr int UP = 8;                 // partial rows in flight per thread: this tail is pure L2 latency
        for (int kb = rl; kb < (int)gridDim.y; kb += kRbRowLanes * UP) {
            float4 a[UP], b4[UP];
#pragma unroll
            for (int u = 0; u < UP; ++u) {
                const int k = kb + u * kRbRowLanes;
                if (k < (int)gridDim.y) {
                    const float4* pp = reinterpret_cast<const float4*>(db_partial + (size_t)k * N + col);
                    a[u] = __ldcg(pp);
                    b4[u] = __ldcg(pp + 1);
                }
            }
#pragma unroll
            for (int u = 0; u < UP; ++u) {
                if (kb + u * kRbRowLanes < (int)gridDim.y) {
                    s[0] += a[u].x; s[1] += a[u].y; s[2] += a[u].z; s[3] += a[u].w;
                    s[4] += b4[u].x; s[5] += b4[u].y; s[6] += b4[u].z; s[7] += b4[u].w;
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) red[rl][cv * 8 + i] = s[i];
    __syncthreads();
    for (int c = threadIdx.x; c < kRbColVecs * 8; c += kRbThreads) {
        const int gc = blockIdx.x * kRbColVecs * 8 + c;
        if (gc < N) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < kRbRowLanes; ++q) t += red[q][c];
            db[gc] = t;
        }
    }
}

// mode 0: SGD  p -= lr * g ;  mode 1: Adagrad  acc += g*g ; p -= lr * g / (sqrt(acc) + eps)
__global__ void __launch_bounds__(256)
dense_opt_kernel(int mode, float* __restrict__ p, float* __restrict__ acc, float* __restrict__ g,
                 __nv_bfloat16* __restrict__ shadow, int64_t n, float lr, float eps, float gscale) {
    const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 >= n) return;
    if (i4 + 4 <= n) {
        float4 pv = *reinterpret_cast<float4*>(p + i4);
        float4 gv = *reinterpret_cast<float4*>(g + i4);
        gv.x *= gscale; gv.y *= gscale; gv.z *= gscale; gv.w *= gscale;
        if (mode == 1) {
            float4 av = *reinterpret_cast<float4*>(acc + i4);
            av.x = __fmaf_rn(gv.x, gv.x, av.x); av.y = __fmaf_rn(gv.y, gv.y, av.y);
            av.z = __fmaf_rn(gv.z, gv.z, av.z); av.w = __fmaf_rn(gv.w, gv.w, av.w);
            *reinterpret_cast<float4*>(acc + i4) = av;
            pv.x -= lr * __fdiv_rn(gv.x, __fsqrt_rn(av.x) + eps); pv.y -= lr * __fdiv_rn(gv.y, __fsqrt_rn(av.y) + eps);
            pv.z -= lr * __fdiv_rn(gv.z, __fsqrt_rn(av.z) + eps); pv.w -= lr * __fdiv_rn(gv.w, __fsqrt_rn(av.w) + eps);
        } else {
            pv.x -= lr * gv.x; pv.y -= lr * gv.y; pv.z -= lr * gv.z; pv.w -= lr * gv.w;
        }
        *reinterpret_cast<float4*>(p + i4) = pv;
        *reinterpret_cast<float4*>(g + i4) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (shadow != nullptr) {
            *reinterpret_cast<__nv_bfloat162*>(shadow + i4) = __floats2bfloat162_rn(pv.x, pv.y);
            *reinterpret_cast<__nv_bfloat162*>(shadow + i4 + 2) = __floats2bfloat162_rn(pv.z, pv.w);
        }
    } else {
        for (int64_t i = i4; i < n; ++i) {
            const float gv = g[i] * gscale;
            float pv = p[i];
            if (mode == 1) {
                const float av = __fmaf_rn(gv, gv, acc[i]);
                acc[i] = av;
                pv -= lr * __fdiv_rn(gv, __fsqrt_rn(av) + eps);
            } else {
                pv -= lr * gv;
            }
            p[i] = pv;
            g[i] = 0.f;
            if (shadow != nullptr) shadow[i] = __float2bfloat16_rn(pv);
        }
    }
}


// ---------------------------------------------------------------------------------------------------------
// Head of the top tower: Linear(K -> 1) + Sigmoid (the last create_mlp layer, dlrm_s_pytorch.py:227-265 with
// sigmoid_layer = last) on a bf16 activation, fp32 master weights.  A 1-column GEMM is a row-wise dot product: HBM-bound
// on reading h once (forward) and reading h + writing dh once (backward); torch runs it as ~10 kernels (fp32 copy of h,
// gemv, sigmoid, their backward, casts).
//   forward : p[b] = sigmoid(sum_k h[b,k] w[k] + bias), one warp per row, 16-byte loads.
//   backward: dz = dp * p * (1 - p);  dh[b,:] = bf16(dz * w);  gw[k] = sum_b dz[b] h[b,k];  gb = sum_b dz[b]
//             -- per-chunk partials, finished in a fixed order by the CTA that ends last (deterministic).
// ---------------------------------------------------------------------------------------------------------
constexpr int kHeadRows = 4;     // rows per warp: their loads are all issued before the first is consumed

__global__ void __launch_bounds__(256)
head_fwd_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ w, const float* __restrict__ bias,
                float* __restrict__ p, int B, int K, int sigmoid) {
    const int lane = threadIdx.x & 31;
    const int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * kHeadRows;
    if (row0 >= B) return;
    float acc[kHeadRows];
#pragma unroll
    for (int r = 0; r < kHeadRows; ++r) acc[r] = 0.f;
    for (int k8 = lane; k8 < (K >> 3); k8 += 32) {
        Bf16x8 hv[kHeadRows];
#pragma unroll
        for (int r = 0; r < kHeadRows; ++r)
            if (row0 + r < B) hv[r] = ld_stream16(h + (row0 + r) * K + 8 * k8);
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + 8 * k8));
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + 8 * k8) + 1);
#pragma unroll
        for (int r = 0; r < kHeadRows; ++r) {
            if (row0 + r >= B) break;
            const float2 a = __bfloat1622float2(hv[r].v[0]), b2 = __bfloat1622float2(hv[r].v[1]);
            const float2 c = __bfloat1622float2(hv[r].v[2]), d = __bfloat1622float2(hv[r].v[3]);
            float t = acc[r];
            t = fmaf(a.x, w0.x, t); t = fmaf(a.y, w0.y, t); t = fmaf(b2.x, w0.z, t); t = fmaf(b2.y, w0.w, t);
            t = fmaf(c.x, w1.x, t); t = fmaf(c.y, w1.y, t); t = fmaf(d.x, w1.z, t); t = fmaf(d.y, w1.w, t);
            acc[r] = t;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < kHeadRows; ++r) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], o);
    }
    if (lane < kHeadRows && row0 + lane < B) {
        float z = acc[0];
#pragma unroll
        for (int r = 1; r < kHeadRows; ++r) z = (lane == r) ? acc[r] : z;
        z += bias[0];
        p[row0 + lane] = sigmoid ? 1.f / (1.f + expf(-z)) : z;
    }
}

__global__ void __launch_bounds__(kRbThreads)
head_bwd_kernel(const float* __restrict__ dp, const float* __restrict__ p, const __nv_bfloat16* __restrict__ h,
                const float* __restrict__ w, __nv_bfloat16* __restrict__ dh, float* __restrict__ gw_partial,
                float* __restrict__ gw, float* __restrict__ gb, unsigned int* __restrict__ counters, int B, int K,
                int rows_per_chunk, int sigmoid, float* __restrict__ gbp_partial, float* __restrict__ gbp) {
    // gbp != nullptr: h is the output of a Linear + ReLU layer whose backward is fused in here -- dh leaves as that layer's
    // dz = dh * (h > 0) and gbp [K] receives its bias gradient (column sums of the bf16-rounded dz), so the layer needs no
    // dlrmx_relu_bwd_bias_grad pass.  gbp_partial: [chunks, K] floats.
    const bool fuse_relu = gbp != nullptr;
    float sb[8];
    // partial layout: [chunks, K + 8]; column K holds the chunk's sum of dz
    __shared__ __align__(16) float red[kRbRowLanes][kRbColVecs * 8 + 8];
    __shared__ float redz[kRbRowLanes];
    __shared__ bool is_last;
    const int cv = threadIdx.x % kRbColVecs, rl = threadIdx.x / kRbColVecs;
    const int col = (blockIdx.x * kRbColVecs + cv) * 8;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(B, r0 + rows_per_chunk);
    const int KP = K + 8;
    float s[8], sz = 0.f, wv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i] = 0.f; sb[i] = 0.f; wv[i] = (col < K) ? w[col + i] : 0.f; }
    if (col < K) {
        constexpr int U = 4;
        for (int rb = r0 + rl; rb < r1; rb += kRbRowLanes * U) {
            Bf16x8 hv[U];
            float dz[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = rb + u * kRbRowLanes;
                dz[u] = 0.f;
                if (r < r1) {
                    hv[u] = ld_stream16(h + (size_t)r * K + col);
                    const float pr = __ldg(p + r);
                    dz[u] = sigmoid ? __ldg(dp + r) * pr * (1.f - pr) : __ldg(dp + r);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = rb + u * kRbRowLanes;
                if (r < r1) {
                    Bf16x8 o;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float2 hf = __bfloat1622float2(hv[u].v[i]);
                        s[2 * i] = fmaf(dz[u], hf.x, s[2 * i]);
                        s[2 * i + 1] = fmaf(dz[u], hf.y, s[2 * i + 1]);
                        float ox = dz[u] * wv[2 * i], oy = dz[u] * wv[2 * i + 1];
                        if (fuse_relu) {
                            ox = hf.x > 0.f ? ox : 0.f;
                            oy = hf.y > 0.f ? oy : 0.f;
                        }
                        o.v[i] = __floats2bfloat162_rn(ox, oy);
                        if (fuse_relu) {
                            const float2 of = __bfloat1622float2(o.v[i]);
                            sb[2 * i] += of.x;
                            sb[2 * i + 1] += of.y;
                        }
                    }
                    *reinterpret_cast<Bf16x8*>(dh + (size_t)r * K + col) = o;
                    sz += dz[u];
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) red[rl][cv * 8 + i] = s[i];
    if (cv == 0) redz[rl] = sz;
    __syncthreads();
    for (int c = threadIdx.x; c < kRbColVecs * 8; c += kRbThreads) {
        const int gc = blockIdx.x * kRbColVecs * 8 + c;
        if (gc < K) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < kRbRowLanes; ++q) t += red[q][c];
            gw_partial[(size_t)blockIdx.y * KP + gc] = t;
        }
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int q = 0; q < kRbRowLanes; ++q) t += redz[q];
        gw_partial[(size_t)blockIdx.y * KP + K] = t;
    }
    if (fuse_relu) {          // same fixed-order column sums for the fused layer's bias gradient
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 8; ++i) red[rl][cv * 8 + i] = sb[i];
        __syncthreads();
        for (int c = threadIdx.x; c < kRbColVecs * 8; c += kRbThreads) {
            const int gc = blockIdx.x * kRbColVecs * 8 + c;
            if (gc < K) {
                float t = 0.f;
#pragma unroll
                for (int q = 0; q < kRbRowLanes; ++q) t += red[q][c];
                gbp_partial[(size_t)blockIdx.y * K + gc] = t;
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int done = atomicAdd(&counters[blockIdx.x], 1u);
        is_last = (done == gridDim.y - 1);
        if (is_last) counters[blockIdx.x] = 0u;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // fixed-order final sums.  The tail is pure L2 latency (one CTA, chunks x (K+8) floats), so it is spread over the
    // whole CTA: thread (cg, sl) sums float4 column group cg (the last group holds the dz column) over chunks
    // sl, sl + S, ... with 16 independent loads in flight, and slice 0 then adds the S slice sums in slice order.
    // (One thread per column walking all chunks 4 at a time took 45 us of this kernel's 53 at 592 chunks.)
    {
        constexpr int S = 3, UNR = 16;
        const int ngroups = (kRbColVecs * 8 + 8) / 4;                 // 66 float4 groups per column block incl. the dz group
        float4* red4 = reinterpret_cast<float4*>(&red[0][0]);         // [S][ngroups] float4 fits in the staging array
        const int cg = threadIdx.x % ngroups, sl = threadIdx.x / ngroups;
        const int gc0 = (cg == ngroups - 2) ? K : blockIdx.x * kRbColVecs * 8 + 4 * cg;   // group 64 = columns K..K+3
        const bool live = sl < S && cg < ngroups - 1 && (cg == ngroups - 2 ? blockIdx.x == 0 : gc0 < K);
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f), t2 = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool live2 = fuse_relu && live && cg < ngroups - 2;
        if (live2) {
            for (int k0 = sl; k0 < (int)gridDim.y; k0 += S * UNR) {
                float4 v[UNR];
#pragma unroll
                for (int u = 0; u < UNR; ++u) {
                    const int k = k0 + u * S;
                    v[u] = (k < (int)gridDim.y) ? __ldcg(reinterpret_cast<const float4*>(gbp_partial + (size_t)k * K + gc0))
                                                : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < UNR; ++u) { t2.x += v[u].x; t2.y += v[u].y; t2.z += v[u].z; t2.w += v[u].w; }
            }
        }
        if (live) {
            for (int k0 = sl; k0 < (int)gridDim.y; k0 += S * UNR) {
                float4 v[UNR];
#pragma unroll
                for (int u = 0; u < UNR; ++u) {
                    const int k = k0 + u * S;
                    v[u] = (k < (int)gridDim.y) ? __ldcg(reinterpret_cast<const float4*>(gw_partial + (size_t)k * KP + gc0))
                                                : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < UNR; ++u) { t.x += v[u].x; t.y += v[u].y; t.z += v[u].z; t.w += v[u].w; }
            }
        }
        __syncthreads();                                              // everyone is done with `red` from the first stage
        if (sl < S) red4[sl * ngroups + cg] = t;
        __syncthreads();
        if (live && sl == 0) {
            float4 a = red4[cg];
#pragma unroll
            for (int q = 1; q < S; ++q) {
                const float4 o = red4[q * ngroups + cg];
                a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
            }
            if (cg == ngroups - 2) gb[0] = a.x;
            else { gw[gc0] = a.x; gw[gc0 + 1] = a.y; gw[gc0 + 2] = a.z; gw[gc0 + 3] = a.w; }
        }
        if (fuse_relu) {
            __syncthreads();
            if (sl < S) red4[sl * ngroups + cg] = t2;
            __syncthreads();
            if (live2 && sl == 0) {
                float4 a = red4[cg];
#pragma unroll
                for (int q = 1; q < S; ++q) {
                    const float4 o = red4[q * ngroups + cg];
                    a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
                }
                gbp[gc0] = a.x; gbp[gc0 + 1] = a.y; gbp[gc0 + 2] = a.z; gbp[gc0 + 3] = a.w;
            }
        }
    }
}

}  // namespace dlrmx

extern "C" {
int dlrmx_relu_bwd_bias_grad(const void* dy, const void* y, void* dz, float* db_partial, float* db,
                             unsigned int* counters, int B, int N, int chunks, int relu, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && N > 0 && N % 8 == 0 && chunks >= 1, "relu_bwd_bias_grad: need N %% 8 == 0, chunks >= 1");
    DLRMX_REQUIRE((((uintptr_t)dy | (uintptr_t)y | (uintptr_t)dz | (uintptr_t)db_partial) & 15) == 0,
                  "relu_bwd_bias_grad: 16-B alignment");
    DLRMX_REQUIRE(db == nullptr || counters != nullptr, "relu_bwd_bias_grad: db needs the counters workspace");
    const int rows_per_chunk = (int)ceil_div(B > 0 ? B : 1, chunks);
    dim3 grid((unsigned)ceil_div(N, kRbColVecs * 8), (unsigned)chunks);
    relu_bwd_bias_grad_kernel<<<grid, kRbThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (__nv_bfloat16*)dz, db_partial, db, counters, B, N,
        rows_per_chunk, relu);
    return check_launch("relu_bwd_bias_grad");
}

// ---- relu_bwd_bias_grad without any allocation ---------------------------------------------------------------------
// Same kernel, same grid as the Python wrapper of dlrmx_relu_bwd_bias_grad picks (so the sums are bit-identical), but its
// scratch (chunk partials + last-CTA counters) is this file-scope device array: nothing is allocated, by the caller or
// here.  Every GPU that loads the library holds the array (304 KB) until the process exits; every call on a device shares
// it, so calls that overlap on one GPU race on it.
namespace dlrmx {
constexpr int kNoAllocMaxN = 2048;                                     // widest layer accepted
constexpr int kNoAllocCtas = 2 * kSMs;                                 // chunks x column blocks ~ 2 CTAs per SM
constexpr int kNoAllocPartialFloats = kNoAllocCtas * kRbColVecs * 8 + kNoAllocMaxN;   // bound of chunks * N, see below
constexpr int kNoAllocCounterBytes = 256;
static __device__ __align__(256) unsigned char g_rb_noalloc_scratch[kNoAllocPartialFloats * 4 + kNoAllocCounterBytes];

// Device address of the scratch array on the CURRENT device (the symbol's name is only a host-side stand-in): looked up
// once per device with cudaGetSymbolAddress, with that device current.
static unsigned char* rb_noalloc_scratch() {
    static std::atomic<unsigned char*> ptrs[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("relu_bwd_bias_grad_noalloc: cudaGetDevice failed or device index >= 64");
        return nullptr;
    }
    unsigned char* p = ptrs[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        void* q = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&q, g_rb_noalloc_scratch);
        if (e != cudaSuccess || q == nullptr) {
            set_error("relu_bwd_bias_grad_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return nullptr;
        }
        p = static_cast<unsigned char*>(q);
        ptrs[dev].store(p, std::memory_order_release);
    }
    return p;
}
}  // namespace dlrmx

int dlrmx_relu_bwd_bias_grad_noalloc(const void* dy, const void* y, void* dz, float* db, int B, int N, int relu,
                                     dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && N > 0 && N % 8 == 0, "relu_bwd_bias_grad_noalloc: need N %% 8 == 0 (B=%d N=%d)", B, N);
    DLRMX_REQUIRE(N <= kNoAllocMaxN, "relu_bwd_bias_grad_noalloc: N = %d exceeds the static scratch (N <= %d); use "
                  "dlrmx_relu_bwd_bias_grad with a caller-provided workspace", N, kNoAllocMaxN);
    DLRMX_REQUIRE(dy != nullptr && db != nullptr && (relu == 0 || (y != nullptr && dz != nullptr)),
                  "relu_bwd_bias_grad_noalloc: dy, db (and y, dz when relu != 0) must be given");
    DLRMX_REQUIRE((((uintptr_t)dy | (uintptr_t)y | (uintptr_t)dz) & 15) == 0, "relu_bwd_bias_grad_noalloc: 16-B alignment");
    const int col_blocks = (int)ceil_div(N, kRbColVecs * 8);
    // as ops.relu_bwd_bias_grad(chunks=None): ~2 CTAs per SM, at least 32 rows per chunk.
    // chunks <= kNoAllocCtas / col_blocks + 1 and N <= 256 * col_blocks, so chunks * N <= kNoAllocPartialFloats.
    int chunks = (int)ceil_div(kNoAllocCtas, col_blocks);
    const int by_rows = (int)ceil_div(B > 0 ? B : 1, 32);
    if (chunks > by_rows) chunks = by_rows;
    if (chunks < 1) chunks = 1;
    DLRMX_REQUIRE((int64_t)chunks * N <= kNoAllocPartialFloats, "relu_bwd_bias_grad_noalloc: internal scratch bound");
    unsigned char* scratch = rb_noalloc_scratch();
    if (scratch == nullptr) return DLRMX_E_CUDA;
    float* partial = reinterpret_cast<float*>(scratch);
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch + (size_t)kNoAllocPartialFloats * 4);
    // the array keeps the previous call's contents: the counters must start at zero (the partials are fully overwritten)
    if (cudaMemsetAsync(counters, 0, kNoAllocCounterBytes, (cudaStream_t)stream) != cudaSuccess) {
        set_error("relu_bwd_bias_grad_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(cudaGetLastError()));
        return DLRMX_E_CUDA;
    }
    const int rows_per_chunk = (int)ceil_div(B > 0 ? B : 1, chunks);
    dim3 grid((unsigned)col_blocks, (unsigned)chunks);
    relu_bwd_bias_grad_kernel<<<grid, kRbThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (__nv_bfloat16*)dz, partial, db, counters, B, N, rows_per_chunk,
        relu);
    return check_launch("relu_bwd_bias_grad_noalloc");
}

int dlrmx_dense_opt(int mode, float* params, float* acc, float* grads, void* shadow_bf16, int64_t n, float lr,
                    float eps, float grad_scale, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(mode == 0 || (mode == 1 && acc != nullptr), "dense_opt: mode 0 (sgd) or 1 (adagrad, needs acc)");
    DLRMX_REQUIRE(n >= 0 && (((uintptr_t)params | (uintptr_t)grads | (uintptr_t)acc | (uintptr_t)shadow_bf16) & 15) == 0,
                  "dense_opt: buffers must be 16-B aligned");
    if (n == 0) return DLRMX_OK;
    const unsigned blocks = (unsigned)ceil_div(ceil_div(n, 4), 256);
    dense_opt_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(mode, params, acc, grads, (__nv_bfloat16*)shadow_bf16, n,
                                                             lr, eps, grad_scale);
    return check_launch("dense_opt");
}

int dlrmx_head_fwd(const void* h_bf16, const float* w, const float* bias, float* p, int B, int K, int sigmoid,
                   dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && K > 0 && K % 8 == 0, "head_fwd: need K %% 8 == 0");
    DLRMX_REQUIRE((((uintptr_t)h_bf16 | (uintptr_t)w) & 15) == 0, "head_fwd: 16-B alignment");
    if (B == 0) return DLRMX_OK;
    head_fwd_kernel<<<(unsigned)ceil_div(ceil_div((int64_t)B, kHeadRows) * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)h_bf16, w, bias, p, B, K, sigmoid);
    return check_launch("head_fwd");
}

size_t dlrmx_head_bwd_workspace_bytes(int K, int chunks) { return (size_t)chunks * (2 * K + 8) * sizeof(float); }

int dlrmx_head_bwd(const float* dp, const float* p, const void* h_bf16, const float* w, void* dh_bf16, float* gw,
                   float* gb, float* workspace, unsigned int* counters, int B, int K, int chunks, int sigmoid,
                   float* gb_prev_relu, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(B >= 0 && K > 0 && K % 8 == 0 && chunks >= 1, "head_bwd: need K %% 8 == 0, chunks >= 1");
    DLRMX_REQUIRE((((uintptr_t)h_bf16 | (uintptr_t)dh_bf16) & 15) == 0, "head_bwd: 16-B alignment");
    DLRMX_REQUIRE(workspace != nullptr && counters != nullptr && gw != nullptr && gb != nullptr, "head_bwd: null buffer");
    const int rows_per_chunk = (int)ceil_div(B > 0 ? B : 1, chunks);
    dim3 grid((unsigned)ceil_div(K, kRbColVecs * 8), (unsigned)chunks);
    head_bwd_kernel<<<grid, kRbThreads, 0, (cudaStream_t)stream>>>(dp, p, (const __nv_bfloat16*)h_bf16, w,
                                                                   (__nv_bfloat16*)dh_bf16, workspace, gw, gb, counters,
                                                                   B, K, rows_per_chunk, sigmoid,
                                                                   workspace + (size_t)chunks * (K + 8), gb_prev_relu);
    return check_launch("head_bwd");
}
}
