// bf16-facing dot interaction, ONE WARP PER SAMPLE.  sm_100a.
//
// Same contract as the MIXED kernels of interact_mma.cu (dlrmx_interact_dot_fwd_mixed / _bwd_mixed; reference:
// DLRM_Net.interact_features('dot'), dlrm_s_pytorch.py:627-659, and its autograd backward), different decomposition.
// The CTA-per-sample kernels were instruction-bound (ncu r01 v4: ~2900 / ~3970 warp instructions per sample, 15-17 %
// tensor pipe, 24 % warps active): four warps shared one 27x27x128 problem, exchanged partial triangles through shared
// memory and met at two block barriers per sample.  Here a warp owns a sample end to end:
//   * its F x D tile arrives by cp.async into the warp's own double buffer (the next sample streams in while this one
//     is contracted -- that is the memory-level parallelism; 6 warps per SM keep ~90 KB of loads in flight);
//   * all k-steps accumulate in the warp's registers (single-pass TF32 mma.sync m16n8k8, fp32 accumulate), no partial
//     sums, no block barriers -- only __syncwarp;
//   * forward: the accumulator fragments are dropped into a bf16 row image in shared memory ([x | tril | 0 pad]) through
//     a per-lane destination table and leave as 16-byte coalesced stores;
//   * backward: the A fragments of S = dZ + dZ^t are gathered straight from the staged bf16 dR row through a per-lane
//     index table (no S matrix is ever built), kept in registers for the whole sample; results go from the accumulator
//     fragments to global memory as full 32-byte sectors (fp32 embedding gradients) / bf16 pairs (dx).
// ~3x fewer instructions per sample -- but MEASURED SLOWER on B200 (Terabyte shape: forward 0.230 ms vs 0.146 ms,
// backward 0.344 ms vs 0.213 ms for the CTA-per-sample kernels): a warp needs a double-buffered 30 KB of shared memory,
// so only 6 warps fit on an SM (1.5 per scheduler) and every dependent instruction of the per-sample chain (fragment
// loads -> mma -> epilogue) is exposed; the 16 warps per SM of the CTA-per-sample layout hide that latency.  Kept as an
// experiment behind tunable 4 (= 1), parity-tested like the default; the way forward is fewer instructions in the
// CTA-per-sample kernels (bulk-copy staging, 128-bit fragment loads through a permuted contraction index).
// Shapes: 2 <= F <= 32, D % 8 == 0, per-warp shared memory <= ~100 KB.
#include <cuda_bf16.h>

#include "common.cuh"

namespace dlrmx {

namespace {

constexpr int kIwWarps = 2;          // warps per CTA (each works alone; they only share the lookup table)

__device__ __forceinline__ uint32_t iw_round_tf32(float x) { return __float_as_uint(x) + 0x1000u; }
__device__ __forceinline__ void iw_cp16(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void iw_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void iw_wait_prev() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void iw_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                       uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__host__ __device__ inline int iw_tri(int F, int itself) { return itself ? F * (F + 1) / 2 : F * (F - 1) / 2; }
__device__ __forceinline__ float iw_bf(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// bf16 x row (staged) -> fp32 row 0 of the tile, so the contraction loops need no special case for feature 0
__device__ __forceinline__ void iw_x_to_row0(float* tile, const __nv_bfloat16* xs, int D, int lane) {
    for (int k4 = lane; k4 < (D >> 2); k4 += 32) {
        const uint2 raw = *reinterpret_cast<const uint2*>(xs + 4 * k4);
        float4 v;
        v.x = __uint_as_float(raw.x << 16);
        v.y = __uint_as_float(raw.x & 0xffff0000u);
        v.z = __uint_as_float(raw.y << 16);
        v.w = __uint_as_float(raw.y & 0xffff0000u);
        *reinterpret_cast<float4*>(tile + 4 * k4) = v;
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kIwWarps * 32)
interact_fwd_warp_kernel(const float* __restrict__ feat, int64_t fsb, const __nv_bfloat16* __restrict__ xb, int B, int F,
                         int D, int itself, __nv_bfloat16* __restrict__ outb, int64_t ostride, int per_warp_bytes) {
    extern __shared__ __align__(16) unsigned char iw_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int DS = D + 4;            // fragment loads T[g + 8r][k0 + t + 4h]: banks 4g + t, all distinct
    const int nz = iw_tri(F, itself);
    // accumulator tiles (m, n) of the 32 x 32 product that touch the lower triangle: (0,0) (0,1) (1,0) (1,1) (1,2) (1,3)
    // table: where accumulator element q of tile ti of this lane goes inside the output row (0xffff = dropped)
    uint16_t* ztab = reinterpret_cast<uint16_t*>(iw_smem);            // [24][32]
    for (int e = threadIdx.x; e < 24 * 32; e += blockDim.x) {
        const int k = e >> 5, ln = e & 31, ti = k >> 2, q = k & 3;
        const int m = ti < 2 ? 0 : 1, n = ti < 2 ? ti : ti - 2;
        const int i = 16 * m + (ln >> 2) + 8 * (q >> 1), j = 8 * n + 2 * (ln & 3) + (q & 1);
        const bool keep = i < F && (itself ? j <= i : j < i);
        ztab[e] = keep ? (uint16_t)(D + (itself ? i * (i + 1) / 2 + j : i * (i - 1) / 2 + j)) : (uint16_t)0xffff;
    }
    __syncthreads();
    unsigned char* wbase = iw_smem + 24 * 32 * 2 + (size_t)warp * per_warp_bytes;
    const int tile_bytes = F * DS * 4, stage_bytes = tile_bytes + D * 2;
    __nv_bfloat16* zs = reinterpret_cast<__nv_bfloat16*>(wbase + 2 * stage_bytes);       // output row image
    for (int k = D + nz + lane; k < ostride; k += 32) zs[k] = __float2bfloat16(0.f);     // pad columns stay zero
    const int ntiles = F > 16 ? 6 : 2;
    int rowoff[4];                   // rows >= F only feed dropped outputs: alias row F-1
#pragma unroll
    for (int r = 0; r < 4; ++r) rowoff[r] = min(g + 8 * r, F - 1) * DS;

    const int64_t nwarps = (int64_t)gridDim.x * kIwWarps;
    const int64_t first = (int64_t)blockIdx.x * kIwWarps + warp;
    auto stage = [&](int64_t s, int which) {
        if (s < B) {
            unsigned char* st = wbase + which * stage_bytes;
            float* tl = reinterpret_cast<float*>(st);
            const float* src = feat + s * fsb;
            for (int f = 1; f < F; ++f)
                for (int k4 = lane; k4 < (D >> 2); k4 += 32) iw_cp16(tl + f * DS + 4 * k4, src + (int64_t)f * D + 4 * k4);
            __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(st + tile_bytes);
            for (int k8 = lane; k8 < (D >> 3); k8 += 32) iw_cp16(xs + 8 * k8, xb + s * D + 8 * k8);
        }
        iw_commit();
    };
    int buf = 0;
    stage(first, 0);
    for (int64_t s = first; s < B; s += nwarps, buf ^= 1) {
        stage(s + nwarps, buf ^ 1);          // the other buffer's readers (this warp) finished in the last iteration
        iw_wait_prev();
        __syncwarp();
        float* tile = reinterpret_cast<float*>(wbase + buf * stage_bytes);
        const __nv_bfloat16* xs = reinterpret_cast<const __nv_bfloat16*>(wbase + buf * stage_bytes + tile_bytes);
        iw_x_to_row0(tile, xs, D, lane);
        for (int k8 = lane; k8 < (D >> 3); k8 += 32)                                      // R[:, 0:D] = x
            *reinterpret_cast<uint4*>(zs + 8 * k8) = *reinterpret_cast<const uint4*>(xs + 8 * k8);
        __syncwarp();

        float c[6][4];
#pragma unroll
        for (int ti = 0; ti < 6; ++ti)
#pragma unroll
            for (int q = 0; q < 4; ++q) c[ti][q] = 0.f;
#pragma unroll 4
        for (int k0 = 0; k0 < D; k0 += 8) {
            uint32_t v[4][2];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int h = 0; h < 2; ++h) v[r][h] = iw_round_tf32(tile[rowoff[r] + k0 + t + 4 * h]);
            // A of m-tile m = rows {2m, 2m+1} x both halves; B of n-tile n = row group n: the same registers
            iw_mma(c[0], v[0][0], v[1][0], v[0][1], v[1][1], v[0][0], v[0][1]);
            iw_mma(c[1], v[0][0], v[1][0], v[0][1], v[1][1], v[1][0], v[1][1]);
            if (ntiles == 6) {
                iw_mma(c[2], v[2][0], v[3][0], v[2][1], v[3][1], v[0][0], v[0][1]);
                iw_mma(c[3], v[2][0], v[3][0], v[2][1], v[3][1], v[1][0], v[1][1]);
                iw_mma(c[4], v[2][0], v[3][0], v[2][1], v[3][1], v[2][0], v[2][1]);
                iw_mma(c[5], v[2][0], v[3][0], v[2][1], v[3][1], v[3][0], v[3][1]);
            }
        }
#pragma unroll
        for (int ti = 0; ti < 6; ++ti) {
            if (ti >= ntiles) break;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint16_t dst = ztab[((ti << 2) + q) * 32 + lane];
                if (dst != 0xffff) zs[dst] = __float2bfloat16_rn(c[ti][q]);
            }
        }
        __syncwarp();
        __nv_bfloat16* o = outb + s * ostride;
        for (int ch = lane; ch < (int)(ostride >> 3); ch += 32)
            *reinterpret_cast<uint4*>(o + 8 * ch) = *reinterpret_cast<const uint4*>(zs + 8 * ch);
        __syncwarp();                        // zs is rewritten by the next sample
    }
}

// ---------------------------------------------------------------------------------------------------------
// backward:  dfeat[s, i, :] = sum_j S[i][j] * feat[s, j, :]  (+ dR[s, :D] on i = 0),  S = dZ + dZ^t
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kIwWarps * 32)
interact_bwd_warp_kernel(const float* __restrict__ feat, int64_t fsb, const __nv_bfloat16* __restrict__ xb, int B, int F,
                         int D, int itself, const __nv_bfloat16* __restrict__ doutb, int64_t dstride,
                         float* __restrict__ dfeat, int64_t dfsb, __nv_bfloat16* __restrict__ dxb, int per_warp_bytes) {
    extern __shared__ __align__(16) unsigned char iw_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int DS = D + 8;            // B fragments T[k0 + t (+4)][n0 + g]: banks 8t + g, all distinct
    const int nz = iw_tri(F, itself);
    const int nchunk = (D + nz + 7) >> 3;
    const int zero_slot = 8 * nchunk;
    // table: A-fragment register k = (m*4 + ks)*4 + r of this lane = S[16m + g + 8(r&1)][8ks + t + 4(r>>1)], given as
    // the index of that dZ entry inside the staged dR row (bit 15: diagonal, counts twice), zero slot outside F x F
    uint16_t* atab = reinterpret_cast<uint16_t*>(iw_smem);            // [32][32]
    for (int e = threadIdx.x; e < 32 * 32; e += blockDim.x) {
        const int k = e >> 5, ln = e & 31, m = k >> 4, ks = (k >> 2) & 3, r = k & 3;
        const int i = 16 * m + (ln >> 2) + 8 * (r & 1), j = 8 * ks + (ln & 3) + 4 * (r >> 1);
        uint16_t v = (uint16_t)zero_slot;
        if (i < F && j < F) {
            if (i == j) {
                if (itself) v = (uint16_t)((D + i * (i + 1) / 2 + i) | 0x8000);
            } else {
                const int hi = max(i, j), lo = min(i, j);
                v = (uint16_t)(D + (itself ? hi * (hi + 1) / 2 + lo : hi * (hi - 1) / 2 + lo));
            }
        }
        atab[e] = v;
    }
    __syncthreads();
    unsigned char* wbase = iw_smem + 32 * 32 * 2 + (size_t)warp * per_warp_bytes;
    const int tile_bytes = F * DS * 4, zraw_bytes = (8 * nchunk + 8) * 2, stage_bytes = tile_bytes + zraw_bytes + D * 2;
    if (lane < 2) *reinterpret_cast<uint4*>(wbase + lane * stage_bytes + tile_bytes + 16 * nchunk) = make_uint4(0, 0, 0, 0);
    const bool two_m = F > 16;
    const int ksteps = (F + 7) >> 3;
    int jr0[4], jr1[4];              // feature rows >= F meet zero coefficients: alias row F-1 (finite data)
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
        jr0[ks] = min(8 * ks + t, F - 1) * DS;
        jr1[ks] = min(8 * ks + t + 4, F - 1) * DS;
    }

    const int64_t nwarps = (int64_t)gridDim.x * kIwWarps;
    const int64_t first = (int64_t)blockIdx.x * kIwWarps + warp;
    auto stage = [&](int64_t s, int which) {
        if (s < B) {
            unsigned char* st = wbase + which * stage_bytes;
            float* tl = reinterpret_cast<float*>(st);
            const float* src = feat + s * fsb;
            for (int f = 1; f < F; ++f)
                for (int k4 = lane; k4 < (D >> 2); k4 += 32) iw_cp16(tl + f * DS + 4 * k4, src + (int64_t)f * D + 4 * k4);
            __nv_bfloat16* zr = reinterpret_cast<__nv_bfloat16*>(st + tile_bytes);
            for (int q = lane; q < nchunk; q += 32) iw_cp16(zr + 8 * q, doutb + s * dstride + 8 * q);
            __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(st + tile_bytes + zraw_bytes);
            for (int k8 = lane; k8 < (D >> 3); k8 += 32) iw_cp16(xs + 8 * k8, xb + s * D + 8 * k8);
        }
        iw_commit();
    };
    int buf = 0;
    stage(first, 0);
    for (int64_t s = first; s < B; s += nwarps, buf ^= 1) {
        stage(s + nwarps, buf ^ 1);
        iw_wait_prev();
        __syncwarp();
        float* tile = reinterpret_cast<float*>(wbase + buf * stage_bytes);
        const __nv_bfloat16* zraw = reinterpret_cast<const __nv_bfloat16*>(wbase + buf * stage_bytes + tile_bytes);
        const __nv_bfloat16* xs = reinterpret_cast<const __nv_bfloat16*>(wbase + buf * stage_bytes + tile_bytes + zraw_bytes);
        iw_x_to_row0(tile, xs, D, lane);

        uint32_t a[32];              // all A fragments of S for this sample (bf16 values are exact in TF32)
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const uint16_t e = atab[k * 32 + lane];
            float v = iw_bf(zraw + (e & 0x7fff));
            if (e & 0x8000) v *= 2.f;
            a[k] = __float_as_uint(v);
        }
        __syncwarp();                // row 0 of the tile is complete

        float* dst = dfeat + s * dfsb;
        for (int nb = 0; nb < D; nb += 32) {
            float c[4][2][4];
#pragma unroll
            for (int n = 0; n < 4; ++n)
#pragma unroll
                for (int m = 0; m < 2; ++m)
#pragma unroll
                    for (int q = 0; q < 4; ++q) c[n][m][q] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                if (ks >= ksteps) break;
#pragma unroll
                for (int n = 0; n < 4; ++n) {
                    const int n0 = nb + 8 * n;
                    if (n0 >= D) break;
                    const uint32_t b0 = iw_round_tf32(tile[jr0[ks] + n0 + g]);
                    const uint32_t b1 = iw_round_tf32(tile[jr1[ks] + n0 + g]);
                    iw_mma(c[n][0], a[ks * 4 + 0], a[ks * 4 + 1], a[ks * 4 + 2], a[ks * 4 + 3], b0, b1);
                    if (two_m)
                        iw_mma(c[n][1], a[16 + ks * 4 + 0], a[16 + ks * 4 + 1], a[16 + ks * 4 + 2], a[16 + ks * 4 + 3], b0, b1);
                }
            }
#pragma unroll
            for (int n = 0; n < 4; ++n) {
                const int n0 = nb + 8 * n;
                if (n0 >= D) break;
                const int col = n0 + 2 * t;
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    if (m == 1 && !two_m) break;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int i = 16 * m + g + 8 * h;
                        if (i >= F) continue;
                        float2 v = make_float2(c[n][m][2 * h], c[n][m][2 * h + 1]);
                        if (i == 0) {      // dense pass-through R[:, 0:D] = x; the gradient of x leaves as bf16
                            v.x = __fadd_rn(v.x, iw_bf(zraw + col));
                            v.y = __fadd_rn(v.y, iw_bf(zraw + col + 1));
                            *reinterpret_cast<__nv_bfloat162*>(dxb + s * D + col) = __floats2bfloat162_rn(v.x, v.y);
                        } else {
                            *reinterpret_cast<float2*>(dst + (int64_t)i * D + col) = v;
                        }
                    }
                }
            }
        }
        __syncwarp();                // every lane is done with this stage before it is refilled (next-next sample)
    }
}

// Launchers: return 1 when the shape is handled here, 0 when the caller should use the CTA-per-sample kernels.
static int iw_grid(int B, int ctas_per_sm) {
    const int64_t cap = (int64_t)kSMs * ctas_per_sm;
    const int64_t need = ((int64_t)B + kIwWarps - 1) / kIwWarps;
    return (int)(need < cap ? need : cap);
}

int interact_fwd_warp(const float* feat, int64_t fsb, const void* x_bf16, int B, int F, int D, int itself,
                      void* out_bf16, int64_t ostride, cudaStream_t st, int* rc) {
    if (ostride % 8 != 0 || ((uintptr_t)out_bf16 & 15) != 0) return 0;
    const size_t per_warp = 2 * ((size_t)F * (D + 4) * 4 + (size_t)D * 2) + (size_t)ostride * 2;
    const size_t smem = 24 * 32 * 2 + kIwWarps * per_warp;
    if (per_warp > 100 * 1024 || smem > 220 * 1024) return 0;
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(interact_fwd_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        // several CTAs per SM only fit if the L1/shared split is at its shared-memory maximum
        cudaFuncSetAttribute(interact_fwd_warp_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
        attr_done = true;
    }
    const int per_sm = (int)((220 * 1024) / (smem + 1024));
    interact_fwd_warp_kernel<<<iw_grid(B, per_sm < 1 ? 1 : per_sm), kIwWarps * 32, smem, st>>>(
        feat, fsb, (const __nv_bfloat16*)x_bf16, B, F, D, itself, (__nv_bfloat16*)out_bf16, ostride, (int)per_warp);
    *rc = check_launch("interact_dot_fwd_mixed(warp)");
    return 1;
}

int interact_bwd_warp(const float* feat, int64_t fsb, const void* x_bf16, int B, int F, int D, int itself,
                      const void* dout_bf16, int64_t dstride, float* dfeat, int64_t dfsb, void* dx_bf16,
                      cudaStream_t st, int* rc) {
    const int nz = iw_tri(F, itself);
    const int nchunk = (D + nz + 7) >> 3;
    const size_t per_warp = 2 * ((size_t)F * (D + 8) * 4 + (size_t)(8 * nchunk + 8) * 2 + (size_t)D * 2);
    const size_t smem = 32 * 32 * 2 + kIwWarps * per_warp;
    if (per_warp > 100 * 1024 || smem > 220 * 1024 || 8 * nchunk >= 0x8000) return 0;
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(interact_bwd_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        cudaFuncSetAttribute(interact_bwd_warp_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
        attr_done = true;
    }
    const int per_sm = (int)((220 * 1024) / (smem + 1024));
    interact_bwd_warp_kernel<<<iw_grid(B, per_sm < 1 ? 1 : per_sm), kIwWarps * 32, smem, st>>>(
        feat, fsb, (const __nv_bfloat16*)x_bf16, B, F, D, itself, (const __nv_bfloat16*)dout_bf16, dstride, dfeat, dfsb,
        (__nv_bfloat16*)dx_bf16, (int)per_warp);
    *rc = check_launch("interact_dot_bwd_mixed(warp)");
    return 1;
}

}  // namespace dlrmx
