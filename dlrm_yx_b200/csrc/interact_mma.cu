// Dot feature interaction on the tensor cores, fp32-accurate (3xTF32 split).  sm_100a.
//
// Replaces DLRM_Net.interact_features('dot') (dlrm_s_pytorch.py:627-659: cat -> bmm(T, T^t) -> tril_indices gather
// -> cat) and its autograd backward, for F <= 32 features and D % 8 == 0 (the Kaggle / benchmark / Terabyte shapes:
// F = 27, 9, 27; D = 16, 64, 128).  Other shapes take the CUDA-core kernels of interact.cu.
//
// One CTA of 4 warps owns one sample at a time (persistent loop over samples):
//   forward : the F x D feature block is staged into shared memory with 16-byte cp.async copies (all in flight
//             at once -- that is the memory-level parallelism), row stride D+4 floats so that the mma fragment
//             loads T[g + 8r][k0 + t + 4h] hit 32 distinct banks.  Z = T T^t is a 32x32x D contraction done as
//             m16n8k8 TF32 mma tiles, lower-triangle tiles only (6 of 8).  Because A and B are the SAME matrix,
//             the 8 values a lane loads per k-step serve as both its A and its B fragments.
//             fp32 accuracy: x = hi + lo with hi = tf32(x), lo = tf32(x - hi) (round-half-away);  Z += a_lo*b_hi + a_hi*b_lo + a_hi*b_hi
//             (the dropped lo*lo term is 2^-22 relative), accumulated in fp32 -- meets the 1e-5 / 1e-6 bar.
//   backward: dT = (dZ + dZ^t) T is a 32 x D x 32 contraction: the symmetric coefficient block S is built in shared
//             memory from dR[:, D:], its fragments (hi/lo) are kept in registers for the whole sample, T is staged
//             with row stride D+8 (bank-conflict-free B fragments), the result goes straight from the accumulator
//             fragments to global memory (full 32-byte sectors), with the pass-through dR[:, :D] added to feature 0.
// HBM traffic is the algorithmic minimum: B*F*D*4 in + B*(D+nz)*4 out (forward).
//
// interact_fwd_bulk_kernel (below the generic forward) is the forward kernel of the training step: same arithmetic as
// interact_fwd_mma_kernel<true, true>, a third of its instructions (ldmatrix fragment loads, constant-offset epilogue).
//
// MIXED variants (template flag) serve the bf16 MLP towers: feature 0 (the bottom-MLP output) is read as bf16, the
// result [x | tril] is written as bf16 with its row padded with zeros to the top MLP's K (a multiple of 8), dR is read
// as bf16 and the feature-0 gradient is written as bf16; pooled embeddings and their gradients stay fp32.  Since the
// consumer rounds to bf16 (2^-9) anyway, the products are single-pass TF32 (operands rounded to nearest, 2^-11) with
// fp32 accumulation -- one mma per tile instead of three, which makes both kernels HBM-bound.
#include <cuda_bf16.h>

#include "common.cuh"

namespace dlrmx {

__device__ __forceinline__ uint32_t round_tf32(float x) { return __float_as_uint(x) + 0x1000u; }   // low 13 bits ignored
__device__ __forceinline__ void cp_async16_raw(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc));
}

// x = hi + lo with hi = x rounded to TF32 (10 explicit mantissa bits, round-half-away: add half an ulp to the
// magnitude, clear the 13 low bits) and lo = (x - hi) rounded the same way.  Written with integer ops because
// `cvt.rna.tf32.f32` expands to a 4-instruction sequence with an Inf/NaN guard on sm_100; the tensor core ignores
// the 13 low operand bits, so `lo` needs only the add.  4 instructions per split instead of 9.  Finite inputs only.
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi)) + 0x1000u;
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// all but the most recent group have landed
__device__ __forceinline__ void cp_async_wait_prev() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// ---- TMA bulk-copy staging (cp.async.bulk global -> shared, completion counted in bytes on an mbarrier) ----------
// The staging loop of per-thread 16-byte cp.async copies is a third of the forward kernel's issued instructions (ncu r01
// v6 source view) in a kernel that is instruction-bound.  When the interaction input is laid out in HBM with the SAME row
// pitch as the shared-memory tile (D + 8 floats: the lookup kernel writes pooled rows with that
// stride), a sample's F-1 embedding rows are ONE contiguous block and ONE bulk copy issued by one thread stages them,
// padding included; the bf16 x row (and the dR row in the backward) are one more bulk copy each.  (One bulk copy per
// 512-byte row into padded rows from an unpadded input was tried first and was slower than cp.async: the copy
// instruction is issued from the uniform datapath, lane-parallel issue becomes a 27-iteration loop in one warp.)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t ok, tries = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (!ok && ++tries > (1u << 24)) __trap();       // a copy that never lands must not hang the GPU
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__host__ __device__ inline int tri_count_mma(int F, int itself) { return itself ? F * (F + 1) / 2 : F * (F - 1) / 2; }

constexpr int kMmaWarps = 4;

// ---------------------------------------------------------------------------------------------------------
// forward: one CTA (4 warps) per sample, persistent loop.  The contraction dimension D is split across the
// warps (warp w takes k-steps w, w+4, ...): every warp owns all six lower-triangle tiles for its k-steps,
// writes its partial triangle to shared memory, and the four partials are added in a fixed order with
// round-to-nearest adds (short accumulation chains inside the tensor core, deterministic result).
// ---------------------------------------------------------------------------------------------------------
template <bool MIXED, bool BULK>
__global__ void __launch_bounds__(kMmaWarps * 32, MIXED ? 6 : 4)
interact_fwd_mma_kernel(const float* __restrict__ feat, int64_t fsb, int64_t fst, int B, int F, int D, int itself,
                        void* __restrict__ out_, int64_t ostride, const __nv_bfloat16* __restrict__ xb) {
    static_assert(!BULK || MIXED, "bulk staging serves the bf16-facing kernels");
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    // fragment loads T[g + 8r][k0 + t + 4h]: with a row pitch of D + 4 the banks are 4g + t, all distinct.  BULK: the tile
    // has the input's own pitch D + 8 (one bulk copy per sample; rows of the input must start on 32-byte sectors or the
    // lookup's row stores become partial-sector writes: a pitch of D + 4 cost the lookup kernel 25 %), where rows g and
    // g + 4 share a bank -- 2-way conflicts on 8 loads per k-step, far cheaper than the staging loop they replace.
    const int DS = BULK ? D + 8 : D + 4;
    const int nz = tri_count_mma(F, itself);
    const int nzp = (nz + 3) & ~3;
    // two tile buffers [F][DS]: the next sample is copied in (cp.async) while this one is being contracted
    float* zpart = smem + 2 * F * DS + warp * nzp;        // this warp's partial triangle
    // MIXED: the bf16 x row of each stage is parked behind the partials and converted at fragment-load time
    __nv_bfloat16* xs_base = reinterpret_cast<__nv_bfloat16*>(smem + 2 * F * DS + kMmaWarps * nzp);
    const int dv = D >> 2;
    const int dv_shift = (dv & (dv - 1)) == 0 ? 31 - __clz(dv) : -1;
    const bool two_m = F > 16;
    int rowoff[4];                                        // rows >= F only feed discarded outputs: alias row F-1
#pragma unroll
    for (int r = 0; r < 4; ++r) rowoff[r] = min(g + 8 * r, F - 1) * DS;
    // where each accumulator element goes in the lower-triangle vector (torch.tril_indices order), -1 = dropped;
    // fixed per lane, so computed once instead of per sample
    int zdst[2][4][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int i = 16 * m + g + 8 * (q >> 1), j = 8 * n + 2 * t + (q & 1);
                const bool keep = (n <= 2 * m + 1) && i < F && (itself ? j <= i : j < i);
                zdst[m][n][q] = keep ? (itself ? i * (i + 1) / 2 + j : i * (i - 1) / 2 + j) : -1;
            }

    __shared__ uint64_t mbar[2];            // BULK: one per stage, a phase completes when the stage's bytes have landed
    if (BULK) {
        if (threadIdx.x == 0) {
            mbar_init(&mbar[0], 1);
            mbar_init(&mbar[1], 1);
            mbar_init_fence();
        }
        __syncthreads();
    }
    auto stage = [&](int64_t bb, int which) {
        if (BULK) {
            // fst == DS: rows 1..F-1 of the sample are one contiguous block with the tile's own row pitch
            if (threadIdx.x == 0 && bb < B) {
                const uint32_t nb_rows = (uint32_t)((F - 1) * DS * 4);
                mbar_expect_tx(&mbar[which], nb_rows + (uint32_t)(D * 2));
                bulk_g2s(smem + which * F * DS + DS, feat + bb * fsb + fst, nb_rows, &mbar[which]);
                bulk_g2s(xs_base + which * D, xb + bb * D, (uint32_t)(D * 2), &mbar[which]);
            }
            return;
        }
        if (bb < B) {
            const float* src = feat + bb * fsb;
            float* dstt = smem + which * F * DS;
            // the F x D block is contiguous in global memory: 16-byte chunk e lands at 4e + (row pad) * (e / dv);
            // a flat loop with a shift instead of nested row / column loops (the staging loop was ~30 % of the
            // kernel's issued instructions, ncu r01 v4 source view)
#pragma unroll 4
            for (int e = (MIXED ? dv : 0) + threadIdx.x; e < F * dv; e += kMmaWarps * 32) {
                const int f = dv_shift >= 0 ? (e >> dv_shift) : (e / dv);
                cp_async16(dstt + 4 * e + 4 * f, src + 4 * e);
            }
            if (MIXED) {
                for (int k8 = threadIdx.x; k8 < (D >> 3); k8 += kMmaWarps * 32)
                    cp_async16_raw(xs_base + which * D + 8 * k8, xb + bb * D + 8 * k8);
            }
        }
        cp_async_commit();
    };
    int buf = 0;
    uint32_t it = 0;
    stage(blockIdx.x, 0);
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x, buf ^= 1, ++it) {
        stage(b + gridDim.x, buf ^ 1);       // that buffer's last readers finished before the previous barrier
        if (BULK) mbar_wait(&mbar[buf], (it >> 1) & 1u);
        else cp_async_wait_prev();
        __syncthreads();
        const float* tile = smem + buf * F * DS;
        const __nv_bfloat16* xs = xs_base + buf * D;
        float* o = nullptr;
        __nv_bfloat16* ob = nullptr;
        if (MIXED) {
            ob = reinterpret_cast<__nv_bfloat16*>(out_) + b * ostride;
            for (int k = threadIdx.x; k < (D >> 1); k += kMmaWarps * 32)          // R[:, 0:D] = x
                reinterpret_cast<__nv_bfloat162*>(ob)[k] = reinterpret_cast<const __nv_bfloat162*>(xs)[k];
        } else {
            o = reinterpret_cast<float*>(out_) + b * ostride;
            for (int k = threadIdx.x; k < D; k += kMmaWarps * 32) o[k] = tile[k];     // R[:, 0:D] = x
        }

        float c[2][4][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int n = 0; n < 4; ++n)
#pragma unroll
                for (int q = 0; q < 4; ++q) c[m][n][q] = 0.f;

        for (int k0 = 8 * warp; k0 < D; k0 += 8 * kMmaWarps) {
            uint32_t hi[4][2], lo[4][2];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float x = tile[rowoff[r] + k0 + t + 4 * h];
                    if (MIXED) {
                        // feature 0 arrives as bf16 (a warp-local pre-conversion into row 0 was measured slower)
                        if (r == 0 && g == 0) x = __bfloat162float(xs[k0 + t + 4 * h]);
                        hi[r][h] = round_tf32(x);
                    } else {
                        split_tf32(x, hi[r][h], lo[r][h]);
                    }
                }
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                if (m == 1 && !two_m) break;
                const uint32_t ah[4] = {hi[2 * m][0], hi[2 * m + 1][0], hi[2 * m][1], hi[2 * m + 1][1]};
                // lower-triangle tiles only (n <= 2m+1); term-major order keeps the dependent mmas apart
                if (!MIXED) {
                    const uint32_t al[4] = {lo[2 * m][0], lo[2 * m + 1][0], lo[2 * m][1], lo[2 * m + 1][1]};
#pragma unroll
                    for (int n = 0; n < 4; ++n)
                        if (n <= 2 * m + 1) mma_tf32(c[m][n], al, hi[n][0], hi[n][1]);
#pragma unroll
                    for (int n = 0; n < 4; ++n)
                        if (n <= 2 * m + 1) mma_tf32(c[m][n], ah, lo[n][0], lo[n][1]);
                }
#pragma unroll
                for (int n = 0; n < 4; ++n)
                    if (n <= 2 * m + 1) mma_tf32(c[m][n], ah, hi[n][0], hi[n][1]);
            }
        }
        // accumulator fragments -> this warp's partial lower-triangle vector
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int n = 0; n < 4; ++n) {
                if (n > 2 * m + 1) continue;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (zdst[m][n][q] >= 0) zpart[zdst[m][n][q]] = c[m][n][q];
            }
        __syncthreads();
        const float* z0 = smem + 2 * F * DS;
        if (MIXED) {       // bf16 pairs; the tail of the row up to ostride (the top MLP's padded K) is zero-filled
            const int npair = (int)((ostride - D) >> 1);
            for (int q2 = threadIdx.x; q2 < npair; q2 += kMmaWarps * 32) {
                float v[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int q = 2 * q2 + e;
                    v[e] = q < nz ? __fadd_rn(__fadd_rn(z0[q], z0[nzp + q]), __fadd_rn(z0[2 * nzp + q], z0[3 * nzp + q]))
                                  : 0.f;
                }
                reinterpret_cast<__nv_bfloat162*>(ob + D)[q2] = __floats2bfloat162_rn(v[0], v[1]);
            }
        } else {
            for (int q = threadIdx.x; q < nz; q += kMmaWarps * 32)
                o[D + q] = __fadd_rn(__fadd_rn(z0[q], z0[nzp + q]), __fadd_rn(z0[2 * nzp + q], z0[3 * nzp + q]));
        }
        // no barrier here: the partials are rewritten only after the next iteration's barrier
    }
}

// ---------------------------------------------------------------------------------------------------------
// forward, bf16-facing, bulk-staged input: the kernel the training step runs.  Same arithmetic as
// interact_fwd_mma_kernel<true, true> (bit-identical results: same k-steps per warp, same MMA order per tile, same
// order of the four partial sums) with a third of the issued instructions.  The ncu source view of that kernel
// (profiles/r02_ncu_full_step_summary.txt: 78.7 M warp instructions per launch = 600 per warp and sample, 4 % of
// them MMAs) put 43 % of the instructions into the k-loop's address arithmetic and scalar fragment loads and 35 % into the
// predicated scatter of the accumulators into the lower-triangle vector.  Here
//   * fragments come from `ldmatrix.x4` (an 8 x 8 b16 matrix row = four fp32 of one feature row: lane (g, t) receives
//     T[8r + g][k0 + 4h + t], which is the m16n8k8 TF32 A fragment AND, for the transposed product, the B fragment):
//     two instructions per k-step instead of eight loads, per-lane row addresses computed once per kernel;
//   * the bf16 x row is widened into row 0 of the tile by the warp that contracts those columns (one load, one shift, one
//     store per lane) instead of a select inside the loop;
//   * accumulators leave as 12 64-bit stores at constant offsets into a [tile][16][8] partial buffer; which slot of it
//     an output element sums is computed once per thread, not per sample.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
constexpr int kFwdTileFloats = 16 * 8;                    // one m16n8 accumulator tile
constexpr int kFwdPartFloats = 6 * kFwdTileFloats;        // lower-triangle tiles: (m0: n0 n1) (m1: n0 n1 n2 n3)
constexpr int kFwdMaxIt = 3;                              // output pairs per thread: (out_stride - D) / 2 <= 3 * 128

template <int NST>      // stages: NST - 1 samples are in flight beside the one being contracted
__global__ void __launch_bounds__(kMmaWarps * 32, NST == 2 ? 5 : 4)
interact_fwd_bulk_kernel(const float* __restrict__ feat, int64_t fsb, int B, int F, int D, int itself,
                         __nv_bfloat16* __restrict__ out, int64_t ostride, const __nv_bfloat16* __restrict__ xb) {
    extern __shared__ __align__(16) float smem[];
    __shared__ uint64_t mbar[NST];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int DS = D + 8;                                 // the input's own row pitch (see interact_fwd_mma_kernel)
    const int tile_floats = F * DS;
    const int nk = D >> 5;                                // k-steps per warp: warp w contracts columns 8w + 32i .. + 7
    const bool two_m = F > 16;
    float* zpart = smem + NST * tile_floats;                // [kMmaWarps][kFwdPartFloats]
    __nv_bfloat16* xs_base = reinterpret_cast<__nv_bfloat16*>(zpart + kMmaWarps * kFwdPartFloats);
    const int nz = tri_count_mma(F, itself);

    // ldmatrix row addresses (bytes from the start of a tile buffer): lane l supplies row l % 8 of matrix l / 8; matrix
    // (r, h) = rows 8r .. 8r+7, columns k0 + 4h .. + 3; instruction p loads (2p, 0) (2p+1, 0) (2p, 1) (2p+1, 1) = the A
    // fragment of row block p.  Rows >= F only feed discarded outputs: they alias row F - 1.
    uint32_t ldoff[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int mi = lane >> 3, r = 2 * p + (mi & 1), h = mi >> 1;
        ldoff[p] = (uint32_t)((min(8 * r + (lane & 7), F - 1) * DS + 4 * h + 8 * warp) * 4);
    }
    // which slot of a partial buffer each of this thread's output elements sums (-1: zero padding of the row)
    int zsrc[kFwdMaxIt][2];
#pragma unroll
    for (int itn = 0; itn < kFwdMaxIt; ++itn)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int q = 2 * ((int)threadIdx.x + itn * kMmaWarps * 32) + e;
            int off = -1;
            if (q < nz) {
                int i = itself ? 0 : 1;
                while (tri_count_mma(i + 1, itself) <= q) ++i;           // row of the lower triangle holding element q
                const int j = q - tri_count_mma(i, itself);
                const int tile = (i >> 4) == 0 ? (j >> 3) : 2 + (j >> 3);
                off = tile * kFwdTileFloats + (i & 15) * 8 + (j & 7);
            }
            zsrc[itn][e] = off;
        }
    const int npair = (int)((ostride - D) >> 1);

    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NST; ++i) mbar_init(&mbar[i], 1);
        mbar_init_fence();
    }
    __syncthreads();
    auto stage = [&](int64_t bb, int which) {             // one thread: rows 1 .. F-1 are one contiguous block
        const uint32_t nb_rows = (uint32_t)((F - 1) * DS * 4);
        mbar_expect_tx(&mbar[which], nb_rows + (uint32_t)(D * 2));
        bulk_g2s(smem + which * tile_floats + DS, feat + bb * fsb + DS, nb_rows, &mbar[which]);
        bulk_g2s(xs_base + which * D, xb + bb * D, (uint32_t)(D * 2), &mbar[which]);
    };
    const uint32_t smem_base = smem_u32(smem);
    float* zp = zpart + warp * kFwdPartFloats + g * 8 + 2 * t;
    int buf = 0;
    uint32_t phase = 0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NST - 1; ++i)
            if ((int64_t)blockIdx.x + (int64_t)i * gridDim.x < B) stage((int64_t)blockIdx.x + (int64_t)i * gridDim.x, i);
    }
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        // the buffer consumed last: its readers (ldmatrix, the x copy) all finished before the previous sample's barriers
        if (threadIdx.x == 0 && b + (int64_t)(NST - 1) * gridDim.x < B)
            stage(b + (int64_t)(NST - 1) * gridDim.x, buf == 0 ? NST - 1 : buf - 1);
        mbar_wait(&mbar[buf], phase);
        float* tile = smem + buf * tile_floats;
        const __nv_bfloat16* xs = xs_base + buf * D;
        __nv_bfloat16* ob = out + b * ostride;
        for (int k = threadIdx.x; k < (D >> 1); k += kMmaWarps * 32)          // R[:, 0:D] = x
            reinterpret_cast<__nv_bfloat162*>(ob)[k] = reinterpret_cast<const __nv_bfloat162*>(xs)[k];
        // feature 0 arrives as bf16: this warp's columns of it -> fp32 row 0 of the tile
        for (int idx = lane; idx < 8 * nk; idx += 32) {
            const int k = 32 * (idx >> 3) + 8 * warp + (idx & 7);
            tile[k] = __uint_as_float((uint32_t)reinterpret_cast<const unsigned short*>(xs)[k] << 16);
        }
        __syncwarp();

        float c[6][4];
#pragma unroll
        for (int n = 0; n < 6; ++n)
#pragma unroll
            for (int q = 0; q < 4; ++q) c[n][q] = 0.f;
        const uint32_t tb = smem_base + (uint32_t)(buf * tile_floats * 4);
        uint32_t ad0 = tb + ldoff[0], ad1 = tb + ldoff[1];
#pragma unroll 4
        for (int i = 0; i < nk; ++i, ad0 += 128, ad1 += 128) {
            uint32_t a0[4], a1[4];
            ldsm_x4(ad0, a0);
#pragma unroll
            for (int q = 0; q < 4; ++q) a0[q] += 0x1000u;             // round to nearest TF32 (low 13 bits are ignored)
            mma_tf32(c[0], a0, a0[0], a0[2]);
            mma_tf32(c[1], a0, a0[1], a0[3]);
            if (two_m) {
                ldsm_x4(ad1, a1);
#pragma unroll
                for (int q = 0; q < 4; ++q) a1[q] += 0x1000u;
                mma_tf32(c[2], a1, a0[0], a0[2]);
                mma_tf32(c[3], a1, a0[1], a0[3]);
                mma_tf32(c[4], a1, a1[0], a1[2]);
                mma_tf32(c[5], a1, a1[1], a1[3]);
            }
        }
        __syncthreads();            // the previous sample's partials have been summed by every thread
#pragma unroll
        for (int n = 0; n < 6; ++n) {
            if (n >= 2 && !two_m) break;
            *reinterpret_cast<float2*>(zp + n * kFwdTileFloats) = make_float2(c[n][0], c[n][1]);
            *reinterpret_cast<float2*>(zp + n * kFwdTileFloats + 64) = make_float2(c[n][2], c[n][3]);
        }
        __syncthreads();
        // bf16 pairs; the tail of the row up to ostride (the top MLP's padded K) is zero-filled
#pragma unroll
        for (int itn = 0; itn < kFwdMaxIt; ++itn) {
            const int q2 = (int)threadIdx.x + itn * kMmaWarps * 32;
            if (q2 >= npair) break;
            float v[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int o = zsrc[itn][e];
                v[e] = o >= 0 ? __fadd_rn(__fadd_rn(zpart[o], zpart[kFwdPartFloats + o]),
                                          __fadd_rn(zpart[2 * kFwdPartFloats + o], zpart[3 * kFwdPartFloats + o]))
                              : 0.f;
            }
            reinterpret_cast<__nv_bfloat162*>(ob + D)[q2] = __floats2bfloat162_rn(v[0], v[1]);
        }
        if (++buf == NST) { buf = 0; phase ^= 1u; }
    }
}

// ---------------------------------------------------------------------------------------------------------
// backward:  dfeat[b, i, :] = sum_j S[i][j] * feat[b, j, :]  (+ dR[b, :D] on i = 0),  S = dZ + dZ^t
// One CTA per sample; the D output columns are split across the 4 warps in 8-column tiles.
// ---------------------------------------------------------------------------------------------------------
template <bool MIXED, bool BULK>
__global__ void __launch_bounds__(kMmaWarps * 32, 4)
interact_bwd_mma_kernel(const float* __restrict__ feat, int64_t fsb, int64_t fst, int B, int F, int D, int itself,
                        const void* __restrict__ dout_, int64_t dstride, float* __restrict__ dfeat, int64_t dfsb,
                        const __nv_bfloat16* __restrict__ xb, __nv_bfloat16* __restrict__ dxb, int relu_x) {
    static_assert(!BULK || MIXED, "bulk staging serves the bf16-facing kernels");
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int DS = D + 8;          // B fragments T[k0 + t (+4)][n0 + g]: banks 8t + g, all distinct (BULK: the input's own pitch)
    constexpr int SS = 36;         // A fragments S[g (+8)][k0 + t (+4)]: banks 4g + t, all distinct
    constexpr int NT = 4;          // 8-column output tiles per warp and pass
    const int nz = tri_count_mma(F, itself);
    const int nzp = (nz + 3) & ~3;
    // two stages of { feature tile [F][DS], dR[b, D:] (+ zero slot at [nzp]) }: the next sample streams in via
    // cp.async while this one is contracted; S [32][SS] is single (rebuilt between the two barriers)
    // MIXED: the stage holds the raw bf16 row dR[b, 0:D+nz] (16-byte chunks, zero slot behind it) and the bf16 x row
    const int nchunk = (D + nz + 7) >> 3;
    const int stage_floats = MIXED ? F * DS + 4 * nchunk + 4 + (D >> 1) : F * DS + nzp + 4;
    const int zero_slot = MIXED ? 8 * nchunk : nzp;
    const int zbase = MIXED ? D : 0;             // index of dZ[0] inside the staged row
    float* S = smem + 2 * stage_floats;
    const int dv = D >> 2;
    const int dv_shift = (dv & (dv - 1)) == 0 ? 31 - __clz(dv) : -1;
    const bool two_m = F > 16;
    const int ksteps = (F + 7) >> 3;

    // S[i][j] = dZ[max][min] (off-diagonal), 2 dZ[i][i] (diagonal, `itself` only), 0 outside F x F.  Each thread
    // fills the same 8 entries for every sample: source slot and scale are computed once.
    int ssrc[8];
    float sscale[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const int e = threadIdx.x + u * kMmaWarps * 32;
        const int i = e >> 5, j = e & 31;
        ssrc[u] = zero_slot;
        sscale[u] = 1.f;
        if (i < F && j < F) {
            if (i == j) {
                if (itself) { ssrc[u] = zbase + i * (i + 1) / 2 + i; sscale[u] = 2.f; }
            } else {
                const int hi_ = max(i, j), lo_ = min(i, j);
                ssrc[u] = zbase + (itself ? hi_ * (hi_ + 1) / 2 + lo_ : hi_ * (hi_ - 1) / 2 + lo_);
            }
        }
    }
    if (threadIdx.x < 2) smem[threadIdx.x * stage_floats + F * DS + (MIXED ? 4 * nchunk : nzp)] = 0.f;
    const float* dout = reinterpret_cast<const float*>(dout_);
    const __nv_bfloat16* doutb = reinterpret_cast<const __nv_bfloat16*>(dout_);

    __shared__ uint64_t mbar[2];
    if (BULK) {
        if (threadIdx.x == 0) {
            mbar_init(&mbar[0], 1);
            mbar_init(&mbar[1], 1);
            mbar_init_fence();
        }
        __syncthreads();
    }
    auto stage = [&](int64_t bb, int which) {
        if (BULK) {
            if (threadIdx.x == 0 && bb < B) {
                float* dstt = smem + which * stage_floats;
                __nv_bfloat16* zr = reinterpret_cast<__nv_bfloat16*>(dstt + F * DS);
                const uint32_t nb_rows = (uint32_t)((F - 1) * DS * 4);
                mbar_expect_tx(&mbar[which], nb_rows + (uint32_t)(nchunk * 16 + D * 2));
                bulk_g2s(dstt + DS, feat + bb * fsb + fst, nb_rows, &mbar[which]);
                bulk_g2s(zr, doutb + bb * dstride, (uint32_t)(nchunk * 16), &mbar[which]);
                bulk_g2s(zr + 8 * nchunk + 8, xb + bb * D, (uint32_t)(D * 2), &mbar[which]);
            }
            return;
        }
        if (bb < B) {
            const float* src = feat + bb * fsb;
            float* dstt = smem + which * stage_floats;
#pragma unroll 4
            for (int e = (MIXED ? dv : 0) + threadIdx.x; e < F * dv; e += kMmaWarps * 32) {
                const int f = dv_shift >= 0 ? (e >> dv_shift) : (e / dv);
                cp_async16(dstt + 4 * e + 8 * f, src + 4 * e);         // row stride D + 8
            }
            if (MIXED) {
                __nv_bfloat16* zr = reinterpret_cast<__nv_bfloat16*>(dstt + F * DS);
                for (int q = threadIdx.x; q < nchunk; q += kMmaWarps * 32)
                    cp_async16_raw(zr + 8 * q, doutb + bb * dstride + 8 * q);
                __nv_bfloat16* xr = zr + 8 * nchunk + 8;
                for (int q = threadIdx.x; q < (D >> 3); q += kMmaWarps * 32)
                    cp_async16_raw(xr + 8 * q, xb + bb * D + 8 * q);
            } else {
                const float* dz = dout + bb * dstride + D;         // rows of dR are only 4-byte aligned
                for (int q = threadIdx.x; q < nz; q += kMmaWarps * 32) cp_async4(dstt + F * DS + q, dz + q);
            }
        }
        cp_async_commit();
    };
    int buf = 0;
    uint32_t it = 0;
    stage(blockIdx.x, 0);
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x, buf ^= 1, ++it) {
        if (BULK) mbar_wait(&mbar[buf], (it >> 1) & 1u);
        if (!BULK) cp_async_wait_all();      // this sample's stage has landed (the only group in flight)
        __syncthreads();                     // ... and every warp is done contracting the previous sample,
        stage(b + gridDim.x, buf ^ 1);       // so its stage can be refilled while this one is worked on
        const float* tile = smem + buf * stage_floats;
        const float* zbuf = tile + F * DS;
        const __nv_bfloat16* zraw = reinterpret_cast<const __nv_bfloat16*>(zbuf);
        const __nv_bfloat16* xs = zraw + 8 * nchunk + 8;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = threadIdx.x + u * kMmaWarps * 32;
            const float zv = MIXED ? __bfloat162float(zraw[ssrc[u]]) : zbuf[ssrc[u]];
            S[(e >> 5) * SS + (e & 31)] = zv * sscale[u];
        }
        __syncthreads();

        float* dst = dfeat + b * dfsb;
        const float* dx = dout + b * dstride;
        for (int nb = 8 * NT * warp; nb < D; nb += 8 * NT * kMmaWarps) {      // NT column tiles per pass
            if (MIXED) {     // feature 0 arrives as bf16: this warp's 32 columns of it -> fp32 row 0 of the tile
                float* row0 = smem + buf * stage_floats;
                if (nb + lane < D) row0[nb + lane] = __bfloat162float(xs[nb + lane]);
                __syncwarp();
            }
            float c[NT][2][4];
#pragma unroll
            for (int n = 0; n < NT; ++n)
#pragma unroll
                for (int m = 0; m < 2; ++m)
#pragma unroll
                    for (int q = 0; q < 4; ++q) c[n][m][q] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                if (ks >= ksteps) break;
                uint32_t ah[2][4], al[2][4];                 // A fragments of S for this k-step, split once
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const int r0 = 16 * m + g, k0 = 8 * ks + t;
                    if (MIXED) {
                        ah[m][0] = round_tf32(S[r0 * SS + k0]);
                        ah[m][1] = round_tf32(S[(r0 + 8) * SS + k0]);
                        ah[m][2] = round_tf32(S[r0 * SS + k0 + 4]);
                        ah[m][3] = round_tf32(S[(r0 + 8) * SS + k0 + 4]);
                    } else {
                        split_tf32(S[r0 * SS + k0], ah[m][0], al[m][0]);
                        split_tf32(S[(r0 + 8) * SS + k0], ah[m][1], al[m][1]);
                        split_tf32(S[r0 * SS + k0 + 4], ah[m][2], al[m][2]);
                        split_tf32(S[(r0 + 8) * SS + k0 + 4], ah[m][3], al[m][3]);
                    }
                }
                const int j0 = 8 * ks + t, j1 = j0 + 4;      // feature rows >= F contribute zero
                const bool ok0 = j0 < F, ok1 = j1 < F;
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    const int n0 = nb + 8 * n;
                    if (n0 >= D) break;
                    uint32_t bh0, bl0 = 0, bh1, bl1 = 0;
                    if (MIXED) {
                        // rows >= F meet zero coefficients of S: alias the last real row (finite) instead of a select
                        bh0 = round_tf32(tile[min(j0, F - 1) * DS + n0 + g]);
                        bh1 = round_tf32(tile[min(j1, F - 1) * DS + n0 + g]);
                    } else {
                        split_tf32(ok0 ? tile[j0 * DS + n0 + g] : 0.f, bh0, bl0);
                        split_tf32(ok1 ? tile[j1 * DS + n0 + g] : 0.f, bh1, bl1);
                    }
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        if (m == 1 && !two_m) break;
                        if (!MIXED) {
                            mma_tf32(c[n][m], al[m], bh0, bh1);
                            mma_tf32(c[n][m], ah[m], bl0, bl1);
                        }
                        mma_tf32(c[n][m], ah[m], bh0, bh1);
                    }
                }
            }
#pragma unroll
            for (int n = 0; n < NT; ++n) {
                const int n0 = nb + 8 * n;
                if (n0 >= D) break;
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    if (m == 1 && !two_m) break;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int i = 16 * m + g + 8 * h;
                        if (i < F) {
                            float2 v = make_float2(c[n][m][2 * h], c[n][m][2 * h + 1]);
                            const int col = n0 + 2 * t;
                            if (MIXED) {
                                if (i == 0) {      // dense pass-through R[:, 0:D] = x; gradient of x leaves as bf16
                                    v.x = __fadd_rn(v.x, __bfloat162float(zraw[col]));
                                    v.y = __fadd_rn(v.y, __bfloat162float(zraw[col + 1]));
                                    if (relu_x) {      // x is the output of a Linear + ReLU layer: its ReLU backward rides along
                                        if (!(__bfloat162float(xs[col]) > 0.f)) v.x = 0.f;
                                        if (!(__bfloat162float(xs[col + 1]) > 0.f)) v.y = 0.f;
                                    }
                                    *reinterpret_cast<__nv_bfloat162*>(dxb + b * D + col) = __floats2bfloat162_rn(v.x, v.y);
                                } else {
                                    *reinterpret_cast<float2*>(dst + (i * D + col)) = v;     // 32-bit offset: F*D < 2^15
                                }
                                continue;
                            }
                            if (i == 0) {          // dense pass-through: R[:, 0:D] = x
                                v.x = __fadd_rn(v.x, __ldg(dx + col));
                                v.y = __fadd_rn(v.y, __ldg(dx + col + 1));
                            }
                            *reinterpret_cast<float2*>(dst + (int64_t)i * D + col) = v;
                        }
                    }
                }
            }
        }
        // no barrier here: S and the other stage are only written after the next iteration's first barrier
    }
}

// tunable 10 = 1: per-thread cp.async staging instead of TMA bulk copies; tunable 11: CTAs per SM of the persistent grid
// tunable 12 = 1: the bulk-staged forward runs interact_fwd_mma_kernel<true, true> instead of interact_fwd_bulk_kernel (tests
// compare); 2: interact_fwd_bulk_kernel with two stages even where three fit
enum { DLRMX_TUNE_IX_STAGE = 10, DLRMX_TUNE_IX_GRID = 11, DLRMX_TUNE_IX_FWD = 12 };
static bool use_bulk() { return get_tunable(DLRMX_TUNE_IX_STAGE) != 1; }

static int mma_grid(int B) {
    const int per_sm = get_tunable(DLRMX_TUNE_IX_GRID) > 0 ? get_tunable(DLRMX_TUNE_IX_GRID) : 12;
    const int64_t cap = (int64_t)kSMs * per_sm;  // persistent: a few CTAs per SM, each loops over samples
    return (int)(B < cap ? B : cap);
}

// interact_fwd_bulk_kernel: persistent, as many CTAs as stay resident (5 per SM at the Terabyte shape's 42 KB)
static int fwd_bulk_grid(int B, size_t smem, int max_per_sm) {
    int per_sm = get_tunable(DLRMX_TUNE_IX_GRID);
    if (per_sm <= 0) {
        per_sm = (int)((size_t)228 * 1024 / (smem + 1024 + 64));
        per_sm = per_sm < 1 ? 1 : (per_sm > max_per_sm ? max_per_sm : per_sm);
    }
    const int64_t cap = (int64_t)kSMs * per_sm;
    return (int)(B < cap ? B : cap);
}

static bool mma_shape_ok(int F, int D) { return F <= 32 && F >= 2 && D % 8 == 0 && D <= 512; }

// Returns 1 if the shape is handled here (and launches), 0 if the caller should use the CUDA-core kernel.
int interact_fwd_mma(const float* feat, int64_t fsb, int B, int F, int D, int itself, float* out, int64_t ostride,
                     cudaStream_t st, int* rc) {
    if (!mma_shape_ok(F, D) || fsb % 4 != 0 || ((uintptr_t)feat & 15) != 0) return 0;
    const int nz = tri_count_mma(F, itself);
    const size_t smem = (2 * (size_t)F * (D + 4) + (size_t)kMmaWarps * ((nz + 3) & ~3)) * 4;
    if (smem > 200 * 1024) return 0;
    static DeviceOnce once;         // the opt-in is per device, not per process
    if (once.first()) {
        cudaFuncSetAttribute(interact_fwd_mma_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    interact_fwd_mma_kernel<false, false><<<mma_grid(B), kMmaWarps * 32, smem, st>>>(feat, fsb, D, B, F, D, itself, out,
                                                                                     ostride, nullptr);
    *rc = check_launch("interact_dot_fwd(mma)");
    return 1;
}

int interact_bwd_mma(const float* feat, int64_t fsb, int B, int F, int D, int itself, const float* dout,
                     int64_t dstride, float* dfeat, int64_t dfsb, cudaStream_t st, int* rc) {
    if (!mma_shape_ok(F, D) || fsb % 4 != 0 || dfsb % 2 != 0 || ((uintptr_t)feat & 15) != 0 ||
        ((uintptr_t)dfeat & 7) != 0)
        return 0;
    const int nz = tri_count_mma(F, itself);
    const size_t smem = (2 * ((size_t)F * (D + 8) + ((nz + 3) & ~3) + 4) + 32 * 36) * 4;
    if (smem > 200 * 1024) return 0;
    static DeviceOnce once;         // the opt-in is per device, not per process
    if (once.first()) {
        cudaFuncSetAttribute(interact_bwd_mma_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    interact_bwd_mma_kernel<false, false><<<mma_grid(B), kMmaWarps * 32, smem, st>>>(
        feat, fsb, D, B, F, D, itself, dout, dstride, dfeat, dfsb, nullptr, nullptr, 0);
    *rc = check_launch("interact_dot_bwd(mma)");
    return 1;
}

// bf16-facing variants.  Return 0 on success, a negative code (message set) when the shape is not supported.
int interact_fwd_mixed(const float* feat, int64_t fsb, int64_t fst, const void* x_bf16, int B, int F, int D, int itself,
                       void* out_bf16, int64_t ostride, cudaStream_t st) {
    const int nz = tri_count_mma(F, itself);
    if (fst <= 0) fst = D;
    if (!mma_shape_ok(F, D) || fsb % 4 != 0 || ((uintptr_t)feat & 15) != 0 || ((uintptr_t)x_bf16 & 15) != 0 ||
        ((uintptr_t)out_bf16 & 3) != 0 || ostride % 2 != 0 || ostride < D + nz || (fst != D && fst != D + 8)) {
        set_error("interact_dot_fwd_mixed: needs 2 <= F <= 32, D %% 8 == 0, D <= 512, 16-B aligned inputs, feature pitch D "
                  "or D + 8, even out_stride >= D + nz (F=%d D=%d pitch=%lld out_stride=%lld)", F, D, (long long)fst,
                  (long long)ostride);
        return DLRMX_E_ARG;
    }
    const bool padded = fst == D + 8;          // the input has the shared-memory tile's row pitch: bulk-copy staging
    if (padded && !(use_bulk() && D % 32 == 0)) {
        set_error("interact_dot_fwd_mixed: a feature pitch of D + 8 needs D %% 32 == 0 (bulk-copy staging)");
        return DLRMX_E_ARG;
    }
    const size_t smem = (2 * (size_t)F * (D + (padded ? 8 : 4)) + (size_t)kMmaWarps * ((nz + 3) & ~3)) * 4 + 2 * (size_t)D * 2;
    if (smem > 200 * 1024) { set_error("interact_dot_fwd_mixed: F*D too large for shared memory"); return DLRMX_E_ARG; }
    static DeviceOnce once;         // the opt-in is per device, not per process
    if (once.first()) {
        cudaFuncSetAttribute(interact_fwd_mma_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(interact_fwd_mma_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    if (padded && get_tunable(DLRMX_TUNE_IX_FWD) != 1 && (ostride - D) / 2 <= kFwdMaxIt * kMmaWarps * 32) {
        // three stages (two samples in flight per CTA) where four such CTAs fit an SM, else two
        const size_t per_stage = (size_t)F * (D + 8) * 4 + (size_t)D * 2;
        const size_t fixed = (size_t)kMmaWarps * kFwdPartFloats * 4;
        const bool three = get_tunable(DLRMX_TUNE_IX_FWD) != 2 && 4 * (3 * per_stage + fixed + 1024 + 64) <= (size_t)228 * 1024;
        const size_t smem2 = (three ? 3 : 2) * per_stage + fixed;
        static DeviceOnce once2;
        if (once2.first()) {
            cudaFuncSetAttribute(interact_fwd_bulk_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            cudaFuncSetAttribute(interact_fwd_bulk_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        }
        if (smem2 <= 200 * 1024) {
            const int grid = fwd_bulk_grid(B, smem2, three ? 4 : 5);
            if (three)
                interact_fwd_bulk_kernel<3><<<grid, kMmaWarps * 32, smem2, st>>>(
                    feat, fsb, B, F, D, itself, (__nv_bfloat16*)out_bf16, ostride, (const __nv_bfloat16*)x_bf16);
            else
                interact_fwd_bulk_kernel<2><<<grid, kMmaWarps * 32, smem2, st>>>(
                    feat, fsb, B, F, D, itself, (__nv_bfloat16*)out_bf16, ostride, (const __nv_bfloat16*)x_bf16);
            return check_launch("interact_dot_fwd_mixed");
        }
    }
    if (padded)
        interact_fwd_mma_kernel<true, true><<<mma_grid(B), kMmaWarps * 32, smem, st>>>(
            feat, fsb, fst, B, F, D, itself, out_bf16, ostride, (const __nv_bfloat16*)x_bf16);
    else
        interact_fwd_mma_kernel<true, false><<<mma_grid(B), kMmaWarps * 32, smem, st>>>(
            feat, fsb, fst, B, F, D, itself, out_bf16, ostride, (const __nv_bfloat16*)x_bf16);
    return check_launch("interact_dot_fwd_mixed");
}

int interact_bwd_mixed(const float* feat, int64_t fsb, int64_t fst, const void* x_bf16, int B, int F, int D, int itself,
                       const void* dout_bf16, int64_t dstride, float* dfeat, int64_t dfsb, void* dx_bf16, int relu_x,
                       cudaStream_t st) {
    const int nz = tri_count_mma(F, itself);
    if (fst <= 0) fst = D;
    if (!mma_shape_ok(F, D) || fsb % 4 != 0 || dfsb % 2 != 0 || ((uintptr_t)feat & 15) != 0 ||
        ((uintptr_t)dfeat & 7) != 0 || ((uintptr_t)x_bf16 & 15) != 0 || ((uintptr_t)dout_bf16 & 15) != 0 ||
        ((uintptr_t)dx_bf16 & 3) != 0 || dstride % 8 != 0 || dstride < ((D + nz + 7) & ~7) || (fst != D && fst != D + 8)) {
        set_error("interact_dot_bwd_mixed: needs 2 <= F <= 32, D %% 8 == 0, D <= 512, 16-B aligned inputs, feature pitch D "
                  "or D + 8, dout_stride %% 8 == 0 and >= D + nz rounded up to 8 (F=%d D=%d pitch=%lld dout_stride=%lld)",
                  F, D, (long long)fst, (long long)dstride);
        return DLRMX_E_ARG;
    }
    const bool padded = fst == D + 8;
    if (padded && !(use_bulk() && D % 32 == 0)) {
        set_error("interact_dot_bwd_mixed: a feature pitch of D + 8 needs D %% 32 == 0 (bulk-copy staging)");
        return DLRMX_E_ARG;
    }
    const int nchunk = (D + nz + 7) >> 3;
    const size_t smem = (2 * ((size_t)F * (D + 8) + 4 * nchunk + 4 + (D >> 1)) + 32 * 36) * 4;
    if (smem > 200 * 1024) { set_error("interact_dot_bwd_mixed: F*D too large for shared memory"); return DLRMX_E_ARG; }
    static DeviceOnce once;         // the opt-in is per device, not per process
    if (once.first()) {
        cudaFuncSetAttribute(interact_bwd_mma_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(interact_bwd_mma_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    if (padded)
        interact_bwd_mma_kernel<true, true><<<mma_grid(B), kMmaWarps * 32, smem, st>>>(
            feat, fsb, fst, B, F, D, itself, dout_bf16, dstride, dfeat, dfsb, (const __nv_bfloat16*)x_bf16,
            (__nv_bfloat16*)dx_bf16, relu_x);
    else
        interact_bwd_mma_kernel<true, false><<<mma_grid(B), kMmaWarps * 32, smem, st>>>(
            feat, fsb, fst, B, F, D, itself, dout_bf16, dstride, dfeat, dfsb, (const __nv_bfloat16*)x_bf16,
            (__nv_bfloat16*)dx_bf16, relu_x);
    return check_launch("interact_dot_bwd_mixed");
}

}  // namespace dlrmx
