// MLP-tower GEMMs on the 5th-generation tensor cores: tcgen05.mma with TMEM accumulators, operands staged by TMA.  sm_100a.
//
// Replaces the cuBLAS GEMMs behind create_mlp / apply_mlp (dlrm_s_pytorch.py:227-265, 518-524: nn.Linear + ReLU stacks)
// and their autograd backward, for bf16 operands with fp32 accumulation:
//   forward   Y  = relu(X W^t + b)          gemm_bf16_kernel<BN, /*B_MN=*/false>, epilogue BIAS_RELU -> bf16
//   dgrad     dX = (dZ W) * (Y_prev > 0)    gemm_bf16_kernel<BN, /*B_MN=*/true>: W [N, K] row-major is the MN-major B operand
//                                           (no transposed weight copy), epilogue applies the PREVIOUS layer's ReLU mask, so
//                                           its result is that layer's dZ -- no separate ReLU-backward pass
//   wgrad     dW = dZ^t X,  db = dZ^t 1     wgrad_streamk_kernel: ALL layers of a tower in ONE launch.  Both operands are
//                                           MN-major (the contraction runs over the batch), tiles [128 n] x [<=256 k]; the
//                                           (tile, k-step) space of all layers is cut into equal contiguous ranges, one per
//                                           SM (stream-K); tiles cut into several segments leave fp32 partials that
//                                           wgrad_reduce_kernel sums in segment order (deterministic, ONE reduce launch
//                                           for the whole tower, element-parallel over all SMs); the bias gradient
//                                           rides along as one extra N = 16 MMA against a tile of ones.
// Kernel shape (both): one CTA per SM, persistent; warp 0 = TMA producer (cp.async.bulk.tensor into 128-byte-swizzled
// stages, mbarrier expect_tx), warp 1 = MMA issuer (one elected lane issues tcgen05.mma, tcgen05.commit releases stages),
// warps 2-5 = epilogue (tcgen05.ld 32x32b from the warp's TMEM lane quarter -> registers -> global).  The forward / dgrad
// kernel double-buffers the accumulator in TMEM (2 x BN columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace dlrmx {
namespace tc {

constexpr int BM = 128;   // UMMA M (cta_group::1): output rows per tile = TMEM lanes
constexpr int BK = 64;    // contraction elements per stage = one 128-byte swizzle atom of bf16
constexpr int UK = 16;    // UMMA K for 16-bit operands
constexpr int kThreads = 192;
constexpr uint32_t kAtomBytes = 64 * BK * 2;   // one [64 mn][64 k] (or [64 k][64 mn]) bf16 block: 8 KB

// ------------------------------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(saddr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = saddr(bar);
    uint32_t ok;
    const long long t0 = clock64();
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (!ok && clock64() - t0 > 4000000000ll) __trap();     // ~2 s: a lost arrival must not hang the GPU
    } while (!ok);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
// 2-D tiled TMA load: box (c0 = innermost element coordinate, c1 = row coordinate) -> shared memory, bytes counted on `bar`
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(saddr(dst)), "l"(map), "r"(saddr(bar)), "r"(c0), "r"(c1) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(saddr(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16 inputs with fp32 accumulation
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on `bar` once every MMA issued so far by this thread has completed (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(saddr(bar)) : "memory");
}
// 32 lanes x 32 columns of fp32: thread i of the warp gets TMEM lane (base lane + i), columns c .. c+31
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- shared-memory matrix descriptors (SWIZZLE_128B, version 1 = Blackwell) -------------------------------------------
// K-major operand tile [rows][64 bf16]: a row is one 128-byte swizzle line, 8 rows form a 1024-byte atom; SBO = distance
// between 8-row groups; one MMA (K = 16) reads 32 bytes of every row: the next k-step starts 32 bytes further.
// MN-major operand tile [64 k-rows][64 mn] blocks of 8 KB: the mn extent of one block is the 128-byte line, 8 k-rows form an
// atom (SBO = 1024 bytes between k-atoms), the next 64 mn live in the next block (LBO = 8192 bytes); the next k-step (16
// k-rows) starts 2048 bytes further.
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_byte_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_byte_addr >> 4) & 0x3fffu);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= 1ull << 46;          // descriptor version (sm_100)
    d |= 2ull << 61;          // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_addr, int k) { return make_desc(tile_addr + k * (UK * 2), 16, 1024); }
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_addr, int k) { return make_desc(tile_addr + k * (UK * 128), kAtomBytes, 1024); }

// instruction descriptor: fp32 accumulate, bf16 x bf16, M = 128
__host__ __device__ constexpr uint32_t make_idesc(int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) |
           ((uint32_t)(BM >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// ------------------------------------------------------------------------------------------------------------------
// forward / dgrad GEMM:  C[M, N] (bf16) = epilogue( A[M, K] . op(B) )
//   B_MN = false: B is [N, K] row-major (contraction contiguous)  -> y = x W^t
//   B_MN = true : B is [K, N] row-major (output dim contiguous)   -> dx = dz W with W [N_layer, K_layer] as stored
// ------------------------------------------------------------------------------------------------------------------
enum { EPI_NONE = 0, EPI_BIAS_RELU = 1, EPI_RELU_MASK = 2, EPI_BIAS = 3 };

struct GemmParams {
    int M, N, K;
    __nv_bfloat16* C;
    int64_t ldc;
    const float* bias;              // [N]            (EPI_BIAS_RELU / EPI_BIAS)
    const __nv_bfloat16* mask;      // [M, N] bf16: C = acc where mask > 0, else 0   (EPI_RELU_MASK)
    int64_t ld_mask;
    int epilogue;
};

template <int BN>
struct GemmSmem {
    static constexpr int kStages = (BN == 256) ? 4 : 6;
    static constexpr uint32_t kABytes = BM * BK * 2, kBBytes = BN * BK * 2;
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;
    static constexpr uint32_t kBarOff = kStages * kStageBytes;
    static constexpr uint32_t kBytes = kBarOff + 256 + 1024;      // barriers + the alignment slack of the dynamic window
};

template <int BN, bool B_MN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const GemmParams P) {
    using S = GemmSmem<BN>;
    constexpr int STAGES = S::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
    uint64_t* empty = full + STAGES;
    uint64_t* acc_full = empty + STAGES;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (P.M + BM - 1) / BM, n_tiles = (P.N + BN - 1) / BN;
    const int tiles = m_tiles * n_tiles;
    const int kblocks = (P.K + BK - 1) / BK;

    if (threadIdx.x == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 2 * BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1u);
                    uint8_t* sa = smem + stage * S::kStageBytes;
                    uint8_t* sb = sa + S::kABytes;
                    mbar_arrive_expect_tx(&full[stage], S::kStageBytes);
                    tma_load_2d(sa, &map_a, &full[stage], kb * BK, m0);
                    if (!B_MN) {
                        tma_load_2d(sb, &map_b, &full[stage], kb * BK, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(sb + j * kAtomBytes, &map_b, &full[stage], n0 + 64 * j, kb * BK);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = make_idesc(BN, false, B_MN);
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int as = it & 1;
            const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
            mbar_wait(&acc_empty[as], aphase ^ 1u);          // the epilogue has drained this accumulator
            tc_fence_after();
            for (int kb = 0; kb < kblocks; ++kb) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = saddr(smem + stage * S::kStageBytes);
                    const uint32_t sb = sa + S::kABytes;
#pragma unroll
                    for (int k = 0; k < BK / UK; ++k) {
                        const uint64_t ad = desc_kmajor(sa, k);
                        const uint64_t bd = B_MN ? desc_mnmajor(sb, k) : desc_kmajor(sb, k);
                        umma(tmem_base + as * BN, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty[stage]);              // the stage is free once these MMAs have read it
                    if (kb == kblocks - 1) umma_commit(&acc_full[as]);
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else {
        // ===== epilogue: 4 warps, warp w reads TMEM lanes 32*(w % 4) .. +31 =====
        const int q = warp & 3;
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
            const int as = it & 1;
            const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < P.M;
            __nv_bfloat16* crow = P.C + (int64_t)row * P.ldc;
            const __nv_bfloat16* mrow = P.mask + (int64_t)row * P.ld_mask;
            const bool masked = P.epilogue == EPI_RELU_MASK && row_ok;
            // the ReLU mask is read one 32-column chunk ahead (the first chunk before the accumulator is even complete):
            // its global-load latency overlaps the MMAs / the previous chunk instead of stalling every chunk
            uint4 mk[4], mk_next[4];
            auto load_mask = [&](uint4 (&dst)[4], int col0) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (col0 + 8 * j < P.N) dst[j] = __ldg(reinterpret_cast<const uint4*>(mrow + col0 + 8 * j));
            };
            if (masked) load_mask(mk_next, n0);
            mbar_wait(&acc_full[as], aphase);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                const int col0 = n0 + c * 32;
                if (col0 >= P.N) break;                      // warp-uniform
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c * 32), v);
                if (masked) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) mk[j] = mk_next[j];
                    if (c + 1 < BN / 32 && col0 + 32 < P.N) load_mask(mk_next, col0 + 32);
                }
                tmem_ld_wait();
                float f[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
                if (P.epilogue == EPI_BIAS_RELU || P.epilogue == EPI_BIAS) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        if (col0 + j < P.N) {
                            const float4 b = __ldg(reinterpret_cast<const float4*>(P.bias + col0 + j));
                            f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
                        }
                    }
                    if (P.epilogue == EPI_BIAS_RELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
                    }
                } else if (P.epilogue == EPI_RELU_MASK && row_ok) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t w4[4] = {mk[j].x, mk[j].y, mk[j].z, mk[j].w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            // bf16 > 0  <=>  sign bit clear and magnitude non-zero
                            const uint32_t lo = w4[e] & 0xffffu, hi = w4[e] >> 16;
                            if (!(lo != 0u && lo < 0x8000u)) f[8 * j + 2 * e] = 0.f;
                            if (!(hi != 0u && hi < 0x8000u)) f[8 * j + 2 * e + 1] = 0.f;
                        }
                    }
                }
                if (row_ok) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (col0 + 8 * j < P.N) {
                            uint4 o;
                            o.x = pack_bf16(f[8 * j], f[8 * j + 1]);
                            o.y = pack_bf16(f[8 * j + 2], f[8 * j + 3]);
                            o.z = pack_bf16(f[8 * j + 4], f[8 * j + 5]);
                            o.w = pack_bf16(f[8 * j + 6], f[8 * j + 7]);
                            *reinterpret_cast<uint4*>(crow + col0 + 8 * j) = o;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[as]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * BN);
    }
}

// ------------------------------------------------------------------------------------------------------------------
// wgrad, all layers of a tower in one launch (stream-K)
// ------------------------------------------------------------------------------------------------------------------
constexpr int kMaxLayers = 8;
constexpr int WN = 256;                    // k-columns of a dW tile (UMMA N), at most
constexpr int kWStages = 4;

struct WgradLayer {
    int N, K;                 // dW is [N, K] (row pitch ld_dw floats)
    int n_tiles, k_tiles;     // tiles of 128 n x up to 256 k
    int tile_base;            // first tile id of this layer
    float* dW;
    int64_t ld_dw;
    float* db;                // [N] or nullptr
};
struct WgradParams {
    int n_layers, M, ksteps, total_tiles, max_segs;
    WgradLayer L[kMaxLayers];
    float* partials;          // [total_tiles][max_segs][128][WN + 32] fp32
};
struct WgradMaps {
    CUtensorMap dz[kMaxLayers];   // [M, N] bf16, box {64 n, 64 m}
    CUtensorMap x[kMaxLayers];    // [M, K] bf16, box {64 k, 64 m}
};
constexpr int kPartialPitch = WN + 32;     // columns 0..255: the dW tile; column 256: the bias-gradient column

struct Segment { int layer, n0, k0, bn, tile, ks0, ks1, seg, nsegs; };

// CTA c of G owns the global k-steps [c*total/G, (c+1)*total/G) of the flattened (tile, k-step) space
__device__ __forceinline__ int64_t range_begin(int64_t c, int64_t total, int G) { return c * total / G; }
__device__ __forceinline__ int owner_of(int64_t g, int64_t total, int G) {
    int c = (int)(g * G / total);
    while (c + 1 < G && range_begin(c + 1, total, G) <= g) ++c;
    while (c > 0 && range_begin(c, total, G) > g) --c;
    return c;
}
__device__ __forceinline__ Segment segment_at(const WgradParams& P, int64_t g, int64_t g_end, int64_t total, int G, int cta) {
    Segment s;
    s.tile = (int)(g / P.ksteps);
    s.ks0 = (int)(g - (int64_t)s.tile * P.ksteps);
    s.ks1 = (int)min((int64_t)P.ksteps, s.ks0 + (g_end - g));
    int l = 0;
    while (l + 1 < P.n_layers && P.L[l + 1].tile_base <= s.tile) ++l;
    s.layer = l;
    const int local = s.tile - P.L[l].tile_base;
    s.n0 = (local / P.L[l].k_tiles) * BM;
    s.k0 = (local % P.L[l].k_tiles) * WN;
    const int kw = min(WN, P.L[l].K - s.k0);
    s.bn = (kw + 63) & ~63;                               // UMMA N / TMA blocks in 64-column units
    const int c_first = owner_of((int64_t)s.tile * P.ksteps, total, G);
    const int c_last = owner_of((int64_t)(s.tile + 1) * P.ksteps - 1, total, G);
    s.seg = cta - c_first;
    s.nsegs = c_last - c_first + 1;
    return s;
}

struct WgradSmem {
    static constexpr uint32_t kABytes = BM * BK * 2, kBBytes = WN * BK * 2;
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;            // 48 KB
    static constexpr uint32_t kOnesOff = kWStages * kStageBytes;          // 16 rows x 128 B of bf16 ones (K-major B operand)
    static constexpr uint32_t kBarOff = kOnesOff + 2048;
    static constexpr uint32_t kBytes = kBarOff + 256 + 1024;
};

__global__ void __launch_bounds__(kThreads, 1)
wgrad_streamk_kernel(const __grid_constant__ WgradMaps maps, const WgradParams P) {
    using S = WgradSmem;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
    uint64_t* empty = full + kWStages;
    uint64_t* acc_full = empty + kWStages;
    uint64_t* acc_empty = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x, cta = blockIdx.x;
    const int64_t total = (int64_t)P.total_tiles * P.ksteps;
    const int64_t g_begin = range_begin(cta, total, G), g_end = range_begin(cta + 1, total, G);

    if (threadIdx.x == 0) {
        for (int l = 0; l < P.n_layers; ++l) { prefetch_tmap(&maps.dz[l]); prefetch_tmap(&maps.x[l]); }
        for (int i = 0; i < kWStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, 4);
        fence_barrier_init();
    }
    // the tile of ones behind the bias gradient: every element 1.0 (bf16 0x3f80), so its layout does not matter
    for (int i = threadIdx.x; i < 2048 / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem + S::kOnesOff)[i] = 0x3f803f80u;
    fence_proxy_async();           // generic-proxy writes -> visible to the tensor core's async proxy
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t g = g_begin; g < g_end;) {
                const Segment s = segment_at(P, g, g_end, total, G, cta);
                const uint32_t bytes = S::kABytes + (uint32_t)s.bn * (BK * 2);
                for (int ks = s.ks0; ks < s.ks1; ++ks) {
                    mbar_wait(&empty[stage], phase ^ 1u);
                    uint8_t* sa = smem + stage * S::kStageBytes;
                    uint8_t* sb = sa + S::kABytes;
                    mbar_arrive_expect_tx(&full[stage], bytes);
                    // A = dZ^t: [64 batch rows][64 n] blocks, two of them for the tile's 128 n
                    tma_load_2d(sa, &maps.dz[s.layer], &full[stage], s.n0, ks * BK);
                    tma_load_2d(sa + kAtomBytes, &maps.dz[s.layer], &full[stage], s.n0 + 64, ks * BK);
                    // B = X: [64 batch rows][64 k] blocks
                    for (int j = 0; j < s.bn / 64; ++j)
                        tma_load_2d(sb + j * kAtomBytes, &maps.x[s.layer], &full[stage], s.k0 + 64 * j, ks * BK);
                    if (++stage == kWStages) { stage = 0; phase ^= 1u; }
                }
                g += s.ks1 - s.ks0;
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        const uint32_t ones = saddr(smem + S::kOnesOff);
        for (int64_t g = g_begin; g < g_end; ++it) {
            const Segment s = segment_at(P, g, g_end, total, G, cta);
            const uint32_t idesc = make_idesc(s.bn, true, true);
            const uint32_t idesc_b = make_idesc(16, true, false);
            const bool with_bias = s.k0 == 0 && P.L[s.layer].db != nullptr;
            mbar_wait(acc_empty, ((uint32_t)it & 1u) ^ 1u);
            tc_fence_after();
            for (int ks = s.ks0; ks < s.ks1; ++ks) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = saddr(smem + stage * S::kStageBytes);
                    const uint32_t sb = sa + S::kABytes;
#pragma unroll
                    for (int k = 0; k < BK / UK; ++k) {
                        const uint32_t acc = (ks > s.ks0 || k > 0) ? 1u : 0u;
                        umma(tmem_base, desc_mnmajor(sa, k), desc_mnmajor(sb, k), idesc, acc);
                        if (with_bias) umma(tmem_base + WN, desc_mnmajor(sa, k), make_desc(ones, 16, 1024), idesc_b, acc);
                    }
                    umma_commit(&empty[stage]);
                    if (ks == s.ks1 - 1) umma_commit(acc_full);
                }
                __syncwarp();
                if (++stage == kWStages) { stage = 0; phase ^= 1u; }
            }
            g += s.ks1 - s.ks0;
        }
    } else {
        // ===== epilogue =====
        const int q = warp & 3;
        int it = 0;
        for (int64_t g = g_begin; g < g_end; ++it) {
            const Segment s = segment_at(P, g, g_end, total, G, cta);
            g += s.ks1 - s.ks0;
            const WgradLayer& L = P.L[s.layer];
            const bool with_bias = s.k0 == 0 && L.db != nullptr;
            const int r = q * 32 + lane;                  // row of the tile = n index
            const int n = s.n0 + r;
            const bool single = s.nsegs == 1;
            mbar_wait(acc_full, (uint32_t)it & 1u);
            tc_fence_after();
            float* prow = P.partials + (((int64_t)s.tile * P.max_segs + s.seg) * BM + r) * kPartialPitch;
            float* wrow = L.dW + (int64_t)n * L.ld_dw + s.k0;
            const int kw = min(WN, L.K - s.k0);
#pragma unroll 1
            for (int c = 0; c < s.bn / 32; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 32), v);
                tmem_ld_wait();
                float* dst = single ? wrow + c * 32 : prow + c * 32;
                if (!single || n < L.N) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        if (!single || c * 32 + j < kw)
                            *reinterpret_cast<float4*>(dst + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                              __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                }
            }
            if (with_bias) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)WN, v);     // 16 valid columns, all equal
                tmem_ld_wait();
                if (single) { if (n < L.N) L.db[n] = __uint_as_float(v[0]); }
                else prow[WN] = __uint_as_float(v[0]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty);        // the next segment's MMAs may overwrite the accumulator
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// Second pass of the stream-K wgrad: every tile that was cut into several segments is summed in segment order (a fixed
// order: deterministic) into dW / db.  Element-parallel over all SMs: CTA (tile, slice of 16 rows), 256 threads, float4
// columns, coalesced.  Tiles computed by one CTA were written by wgrad_streamk_kernel itself.
constexpr int kReduceRows = 16;
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const WgradParams P, const int G) {
    const int tile = blockIdx.x / (BM / kReduceRows);
    const int r0 = (blockIdx.x % (BM / kReduceRows)) * kReduceRows;
    const int64_t total = (int64_t)P.total_tiles * P.ksteps;
    const int c_first = owner_of((int64_t)tile * P.ksteps, total, G);
    const int c_last = owner_of((int64_t)(tile + 1) * P.ksteps - 1, total, G);
    const int nsegs = c_last - c_first + 1;
    if (nsegs == 1) return;
    int l = 0;
    while (l + 1 < P.n_layers && P.L[l + 1].tile_base <= tile) ++l;
    const WgradLayer& L = P.L[l];
    const int local = tile - L.tile_base;
    const int n0 = (local / L.k_tiles) * BM, k0 = (local % L.k_tiles) * WN;
    const int kw = min(WN, L.K - k0);
    const int q4 = kw >> 2;                                   // float4 columns of the tile (K % 8 == 0)
    const float* p0 = P.partials + (int64_t)tile * P.max_segs * BM * kPartialPitch;
    const int64_t seg_stride = (int64_t)BM * kPartialPitch;
    const int rows = min(kReduceRows, L.N - n0 - r0);
    for (int idx = threadIdx.x; idx < rows * q4; idx += 256) {
        const int r = r0 + idx / q4, c = (idx % q4) << 2;
        const float* src = p0 + (int64_t)r * kPartialPitch + c;
        float4 a = __ldcg(reinterpret_cast<const float4*>(src));
#pragma unroll 4
        for (int sgi = 1; sgi < nsegs; ++sgi) {
            const float4 b = __ldcg(reinterpret_cast<const float4*>(src + sgi * seg_stride));
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        *reinterpret_cast<float4*>(L.dW + (int64_t)(n0 + r) * L.ld_dw + k0 + c) = a;
    }
    if (k0 == 0 && L.db != nullptr && (int)threadIdx.x < rows) {
        const int r = r0 + threadIdx.x;
        const float* src = p0 + (int64_t)r * kPartialPitch + WN;
        float a = __ldcg(src);
        for (int sgi = 1; sgi < nsegs; ++sgi) a += __ldcg(src + sgi * seg_stride);
        L.db[n0 + r] = a;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// bf16 matrix [rows, cols] with row pitch ld (elements); box = {box_cols (<= 64: one 128-byte swizzle line), box_rows}
static bool make_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_cols, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return false; }
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (ld * 2) % 16 != 0) {
        set_error("tensor-core GEMM operands must be 16-byte aligned with row pitches that are multiples of 8 elements");
        return false;
    }
    const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return false; }
    return true;
}

template <int BN, bool B_MN>
static int launch_gemm(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& P, cudaStream_t st) {
    using S = GemmSmem<BN>;
    static DeviceOnce once;
    if (once.first())
        cudaFuncSetAttribute(gemm_bf16_kernel<BN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::kBytes);
    const int tiles = (int)(ceil_div(P.M, BM) * ceil_div(P.N, BN));
    const int grid = tiles < kSMs ? tiles : kSMs;
    gemm_bf16_kernel<BN, B_MN><<<grid, kThreads, S::kBytes, st>>>(ma, mb, P);
    return check_launch("gemm_bf16");
}

}  // namespace tc
}  // namespace dlrmx

extern "C" {

int dlrmx_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int b_is_kn, void* C, int64_t ldc, int M, int N,
                    int K, int epilogue, const float* bias, const void* relu_mask, int64_t ld_mask, dlrmx_stream_t stream) {
    using namespace dlrmx;
    using namespace dlrmx::tc;
    DLRMX_REQUIRE(M >= 0 && N > 0 && K > 0 && N % 8 == 0 && K % 8 == 0, "gemm_bf16: need N %% 8 == 0 and K %% 8 == 0 (M=%d N=%d K=%d)", M, N, K);
    DLRMX_REQUIRE(A != nullptr && B != nullptr && C != nullptr && ldc % 8 == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0,
                  "gemm_bf16: C must be 16-byte aligned with ldc %% 8 == 0");
    DLRMX_REQUIRE(epilogue >= EPI_NONE && epilogue <= EPI_BIAS, "gemm_bf16: bad epilogue %d", epilogue);
    DLRMX_REQUIRE((epilogue != EPI_BIAS_RELU && epilogue != EPI_BIAS) || (bias != nullptr && (reinterpret_cast<uintptr_t>(bias) & 15) == 0),
                  "gemm_bf16: the bias epilogues need a 16-byte aligned bias");
    DLRMX_REQUIRE(epilogue != EPI_RELU_MASK || (relu_mask != nullptr && ld_mask % 8 == 0 && (reinterpret_cast<uintptr_t>(relu_mask) & 15) == 0),
                  "gemm_bf16: the ReLU-mask epilogue needs a 16-byte aligned mask with ld_mask %% 8 == 0");
    if (M == 0) return DLRMX_OK;
    const int BN = N > 128 ? 256 : (N > 64 ? 128 : 64);
    CUtensorMap ma, mb;
    if (!make_map(&ma, A, M, K, lda, BK, BM)) return DLRMX_E_ARG;
    if (b_is_kn) {
        if (!make_map(&mb, B, K, N, ldb, 64, BK)) return DLRMX_E_ARG;          // [K, N]: blocks of 64 k-rows x 64 n
    } else {
        if (!make_map(&mb, B, N, K, ldb, BK, BN)) return DLRMX_E_ARG;          // [N, K]: BN rows x 64 k
    }
    GemmParams P;
    P.M = M; P.N = N; P.K = K; P.C = (__nv_bfloat16*)C; P.ldc = ldc; P.bias = bias;
    P.mask = (const __nv_bfloat16*)relu_mask; P.ld_mask = ld_mask; P.epilogue = epilogue;
    cudaStream_t st = (cudaStream_t)stream;
    if (b_is_kn) {
        if (BN == 256) return launch_gemm<256, true>(ma, mb, P, st);
        if (BN == 128) return launch_gemm<128, true>(ma, mb, P, st);
        return launch_gemm<64, true>(ma, mb, P, st);
    }
    if (BN == 256) return launch_gemm<256, false>(ma, mb, P, st);
    if (BN == 128) return launch_gemm<128, false>(ma, mb, P, st);
    return launch_gemm<64, false>(ma, mb, P, st);
}

static int wgrad_tiles(int n_layers, const int* N, const int* K) {
    int t = 0;
    for (int l = 0; l < n_layers; ++l) t += (int)(dlrmx::ceil_div(N[l], dlrmx::tc::BM) * dlrmx::ceil_div(K[l], dlrmx::tc::WN));
    return t;
}
static int wgrad_grid(int total_tiles, int M) {
    const int64_t total = (int64_t)total_tiles * dlrmx::ceil_div(M, dlrmx::tc::BK);
    int64_t g = total / 8;                      // at least 8 k-steps per CTA
    if (g < 1) g = 1;
    return (int)(g < dlrmx::kSMs ? g : dlrmx::kSMs);
}
static int wgrad_max_segs(int total_tiles, int M) {
    const int ksteps = (int)dlrmx::ceil_div(M, dlrmx::tc::BK);
    const int G = wgrad_grid(total_tiles, M);
    const int64_t total = (int64_t)total_tiles * ksteps;
    const int64_t per = total / G > 0 ? total / G : 1;
    int64_t s = ksteps / per + 2;
    return (int)(s < G ? s : G);
}

size_t dlrmx_gemm_wgrad_workspace_bytes(int n_layers, const int* N, const int* K, int M) {
    if (n_layers <= 0 || M <= 0) return 256;
    const int tiles = wgrad_tiles(n_layers, N, K);
    return (size_t)tiles * wgrad_max_segs(tiles, M) * dlrmx::tc::BM * dlrmx::tc::kPartialPitch * 4 + dlrmx::align_up((size_t)tiles * 4, 256);
}

int dlrmx_gemm_wgrad_bf16(int n_layers, const void* const* dZ, const int64_t* ld_dz, const void* const* X, const int64_t* ld_x,
                          float* const* dW, const int64_t* ld_dw, float* const* db, const int* N, const int* K, int M,
                          void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    using namespace dlrmx;
    using namespace dlrmx::tc;
    DLRMX_REQUIRE(n_layers >= 1 && n_layers <= kMaxLayers, "gemm_wgrad: 1..%d layers per launch", kMaxLayers);
    DLRMX_REQUIRE(M > 0, "gemm_wgrad: empty batch");
    if (workspace_bytes < dlrmx_gemm_wgrad_workspace_bytes(n_layers, N, K, M)) {
        set_error("gemm_wgrad: workspace %zu < %zu", workspace_bytes, dlrmx_gemm_wgrad_workspace_bytes(n_layers, N, K, M));
        return DLRMX_E_WORKSPACE;
    }
    WgradMaps maps;
    WgradParams P;
    P.n_layers = n_layers; P.M = M; P.ksteps = (int)ceil_div(M, BK);
    int tiles = 0;
    for (int l = 0; l < n_layers; ++l) {
        DLRMX_REQUIRE(N[l] > 0 && K[l] > 0 && N[l] % 8 == 0 && K[l] % 8 == 0 && ld_dw[l] % 4 == 0 &&
                          (reinterpret_cast<uintptr_t>(dW[l]) & 15) == 0,
                      "gemm_wgrad: layer %d needs N %% 8 == 0, K %% 8 == 0 and a 16-byte aligned dW", l);
        if (!make_map(&maps.dz[l], dZ[l], M, N[l], ld_dz[l], 64, BK)) return DLRMX_E_ARG;
        if (!make_map(&maps.x[l], X[l], M, K[l], ld_x[l], 64, BK)) return DLRMX_E_ARG;
        WgradLayer& L = P.L[l];
        L.N = N[l]; L.K = K[l]; L.n_tiles = (int)ceil_div(N[l], BM); L.k_tiles = (int)ceil_div(K[l], WN);
        L.tile_base = tiles; L.dW = dW[l]; L.ld_dw = ld_dw[l]; L.db = db != nullptr ? db[l] : nullptr;
        tiles += L.n_tiles * L.k_tiles;
    }
    P.total_tiles = tiles;
    P.max_segs = wgrad_max_segs(tiles, M);
    P.partials = (float*)workspace;
    static DeviceOnce once;
    if (once.first())
        cudaFuncSetAttribute(wgrad_streamk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WgradSmem::kBytes);
    const int G = wgrad_grid(tiles, M);
    wgrad_streamk_kernel<<<G, kThreads, WgradSmem::kBytes, (cudaStream_t)stream>>>(maps, P);
    if (int rc = check_launch("gemm_wgrad")) return rc;
    if (G > tiles || (int64_t)tiles * P.ksteps % G != 0 || ((int64_t)tiles * P.ksteps / G) % P.ksteps != 0) {   // some tile is split
        wgrad_reduce_kernel<<<tiles * (BM / kReduceRows), 256, 0, (cudaStream_t)stream>>>(P, G);
        return check_launch("gemm_wgrad_reduce");
    }
    return DLRMX_OK;
}
}
