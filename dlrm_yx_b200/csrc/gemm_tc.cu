// MLP-tower GEMMs on the 5th-generation tensor cores: tcgen05.mma.cta_group::2 (CTA pairs, M = 256 per MMA) with TMEM
// accumulators, operands staged by TMA.  sm_100a.
//
// Replaces the cuBLAS GEMMs behind create_mlp / apply_mlp (dlrm_s_pytorch.py:227-265, 518-524: nn.Linear + ReLU stacks)
// and their autograd backward, for bf16 operands with fp32 accumulation:
//   forward   Y  = relu(X W^t + b)          gemm_pair_kernel<BN, /*B_MN=*/false>, epilogue BIAS_RELU -> bf16 (+ 1 gate bit per
//                                           element for the backward)
//   dgrad     dX = (dZ W) * (Y_prev > 0)    gemm_pair_kernel<BN, /*B_MN=*/true>: W [N, K] row-major is the MN-major B operand
//                                           (no transposed weight copy), epilogue applies the PREVIOUS layer's ReLU gates, so
//                                           its result is that layer's dZ -- no separate ReLU-backward pass
//   wgrad     dW = dZ^t X,  db = dZ^t 1     wgrad_splitk_kernel: ALL layers of a tower in ONE launch.  Both operands are
//                                           MN-major (the contraction runs over the batch), tiles [256 n] x [<=256 k]; the
//                                           batch range of every tile is cut into `split` chunks, work units (chunk, tile)
//                                           go to the CTA pairs chunk-major (concurrent pairs walk the same batch rows:
//                                           operands come from DRAM once); the chunks leave fp32 partials that
//                                           wgrad_reduce_kernel sums in chunk order (deterministic, ONE reduce launch
//                                           for the whole tower, element-parallel over all SMs); the bias gradient
//                                           rides along as one extra N = 16 MMA against a tile of ones.
// Kernel shape (both): persistent, one CTA per SM, two CTAs of a cluster (the SMs of a TPC) per tile; warp 0 = TMA producer
// (cp.async.bulk.tensor.cta_group::2 into 128-byte-swizzled stages, bytes counted on the leader's mbarrier), warp 1 of the
// leader = MMA issuer (one elected lane issues tcgen05.mma, multicast tcgen05.commit releases the stages of both CTAs), the
// other warps = epilogue (tcgen05.ld 32x32b from the warp's TMEM lane quarter -> registers -> shared memory -> TMA store, or
// fp32 rows straight to global in the wgrad).  The forward / dgrad kernel double-buffers the accumulator in TMEM (2 x BN
// columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>

#include <utility>

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

// ---- CTA pairs (cta_group::2): two CTAs of one cluster, on the two SMs of a TPC, run ONE MMA of M = 256 -------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory variable in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t cta_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
    return r;
}
// (default .release.cta: an explicit .release.cluster costs a MEMBAR.ALL.GPU per arrive -- measured: it paced the whole pipeline)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load issued by either CTA of a pair into its OWN shared memory; the bytes are counted on the barrier at cluster address
// `bar_cluster` (the leader CTA's)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(saddr(dst)), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(saddr(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// one MMA of M = 256 over the pair: rows 0..127 of D accumulate in the leader's TMEM, rows 128..255 in the peer's; A rows and
// the two halves of B's N extent are read from the two CTAs' shared memory at the SAME offsets
__device__ __forceinline__ void umma_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(saddr(bar)), "h"((uint16_t)3) : "memory");
}
// TMA store of a shared-memory box (128-byte swizzled) into the tensor; rows / columns beyond the tensor are clipped
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(saddr(src)), "r"(c0), "r"(c1) : "memory");
}
// Programmatic dependent launch: the kernels of a tower run back to back on one stream.  Each lets the next one be launched
// as soon as all its CTAs have got past their prologue (launch_dependents) and itself waits -- after ITS prologue: barrier
// init, TMEM allocation, cluster sync, tensor-map prefetch, none of which touch global memory -- until the grid before it has
// completed and flushed (grid_dependency_wait).  The launch latency and the prologue of kernel i+1 overlap the tail of
// kernel i.  Both are no-ops for a kernel launched without the attribute.
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

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
__host__ __device__ constexpr uint32_t make_idesc(int n, bool a_mn, bool b_mn, int m = BM) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) |
           ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// ------------------------------------------------------------------------------------------------------------------
// forward / dgrad GEMM:  C[M, N] (bf16) = epilogue( A[M, K] . op(B) )
//   B_MN = false: B is [N, K] row-major (contraction contiguous)  -> y = x W^t
//   B_MN = true : B is [K, N] row-major (output dim contiguous)   -> dx = dz W with W [N_layer, K_layer] as stored
// (Round 2's first version ran one CTA per SM with M = 128 MMAs: 48 KB from L2 per 128 x 256 x 64 block of MMA work put it
// on the ~12 TB/s L2 -> SM limit, 1.02 PFLOP/s at 1024 x 1024 against cuBLAS' 1.18, and its 16-byte row stores made the
// epilogue the bottleneck of every layer with K <= 512.  Replaced by the CTA-pair kernel below; numbers in DESIGN.md.)
// ------------------------------------------------------------------------------------------------------------------
enum { EPI_NONE = 0, EPI_BIAS_RELU = 1, EPI_BIAS = 3, EPI_RELU_BITS = 4 };

// ------------------------------------------------------------------------------------------------------------------
// forward / dgrad GEMM on CTA pairs.  The single-CTA kernel above moves 48 KB from L2 per 128 x 256 x 64 block of MMA work
// and is bound by the L2 -> SM fabric (~12 TB/s over the chip); here the two CTAs of a cluster (the two SMs of a TPC) run
// ONE tcgen05.mma.cta_group::2 of M = 256: each CTA stages its own 128 rows of A and only HALF of the B tile (32 KB per
// the same MMA work), the tensor core reads both halves.  The epilogue goes registers -> 128-byte-swizzled shared memory
// -> TMA store (full-line writes instead of 32 scattered 16-byte stores per instruction), and the ReLU gates travel as one
// bit per element: the forward writes them next to Y (relu_bits), the dgrad epilogue reads 4 bytes per 32 columns.
//   pair tile [256 m] x [BN n];  CTA r: A rows m0 + 128 r .. +127, B rows (N) n0 + r BN/2 .. (K-major) or the same columns
//   of the [K, N] matrix (MN-major, 64-column blocks), D rows 128 r .. in its own TMEM (2 x BN columns, double-buffered).
//   Barriers: full[s] lives in the leader (its arrive.expect_tx counts both CTAs' bytes; the peer's TMA signals it remotely),
//   empty[s] / acc_full[a] in both CTAs (multicast tcgen05.commit from the leader), acc_empty[a] in the leader (16 epilogue
//   warps arrive, the peer's remotely).
// ------------------------------------------------------------------------------------------------------------------
struct PairParams {
    int M, N, K;
    const float* bias;              // [N] fp32 (EPI_BIAS_RELU / EPI_BIAS)
    uint32_t* bits_out;             // EPI_BIAS_RELU, optional: bit (n & 31) of word [m, n >> 5] = (Y[m, n] > 0)
    const uint32_t* bits_in;        // EPI_RELU_BITS: C = acc where the bit is set, else 0
    int64_t ld_bits;                // words per row of either bit matrix
    int epilogue;
};

template <int BN>
struct PairSmem {
    static constexpr uint32_t kABytes = BM * BK * 2, kBBytes = (BN / 2) * BK * 2;
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;
    static constexpr uint32_t kOutBytes = 8 * 2 * 4096;                       // 8 epilogue warps x 2 buffers x [32 rows][128 B]
    static constexpr int kStages = (227 * 1024 - 2048 - (int)kOutBytes) / (int)kStageBytes > 8
                                       ? 8 : (227 * 1024 - 2048 - (int)kOutBytes) / (int)kStageBytes;
    static constexpr uint32_t kOutOff = kStages * kStageBytes;
    static constexpr uint32_t kBarOff = kOutOff + kOutBytes;
    static constexpr uint32_t kBytes = kBarOff + 512 + 1024;                  // barriers + the alignment slack
};

constexpr int kPairThreads = 320;          // warp 0: TMA producer, warp 1: MMA issuer, warps 2-9: epilogue

template <int BN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_c, const PairParams P) {
    using S = PairSmem<BN>;
    constexpr int STAGES = S::kStages;
    static_assert(!B_MN || BN >= 128, "an MN-major B half must span whole 64-column swizzle blocks");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
    uint64_t* empty = full + STAGES;
    uint64_t* acc_full = empty + STAGES;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int m_tiles = (P.M + 2 * BM - 1) / (2 * BM), n_tiles = (P.N + BN - 1) / BN;
    const int tiles = m_tiles * n_tiles;
    const int kblocks = (P.K + BK - 1) / BK;
    const int pair = blockIdx.x >> 1, pairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        prefetch_tmap(&map_c);
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 16); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, 2 * BN);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();            // the peer's barriers are initialised before anything arrives on them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    grid_launch_dependents();
    grid_dependency_wait();        // everything below reads or writes global memory

    if (warp == 0) {
        // ===== TMA producer (both CTAs: own A rows, own half of B) =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = pair; tile < tiles; tile += pairs) {
                const int m0 = (tile / n_tiles) * (2 * BM) + (int)rank * BM;
                const int n0 = (tile % n_tiles) * BN + (int)rank * (BN / 2);
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1u);
                    uint8_t* sa = smem + stage * S::kStageBytes;
                    uint8_t* sb = sa + S::kABytes;
                    const uint32_t bar = mapa(saddr(&full[stage]), 0);
                    // the leader alone arrives, expecting both CTAs' bytes: the peer cannot reach this phase of the barrier early,
                    // its empty[stage] is released by the same commit that follows the leader's wait on the previous phase
                    if (leader) mbar_arrive_expect_tx(&full[stage], 2 * S::kStageBytes);
                    tma_load_2d_pair(sa, &map_a, bar, kb * BK, m0);
                    if (!B_MN) {
                        tma_load_2d_pair(sb, &map_b, bar, kb * BK, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 128; ++j)
                            tma_load_2d_pair(sb + j * kAtomBytes, &map_b, bar, n0 + 64 * j, kb * BK);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (leader) {
            constexpr uint32_t idesc = make_idesc(BN, false, B_MN, 2 * BM);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = pair; tile < tiles; tile += pairs, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
                mbar_wait(&acc_empty[as], aphase ^ 1u);          // both CTAs' epilogues have drained this accumulator
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
                            umma_pair(tmem_base + as * BN, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
                        }
                        umma_commit_pair(&empty[stage]);         // frees the stage in both CTAs
                        if (kb == kblocks - 1) umma_commit_pair(&acc_full[as]);
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else {
        // ===== epilogue: 8 warps per CTA.  Warp w reads TMEM lanes 32*(w % 4) .. +31 (rows of this CTA's half of the tile);
        // the two warps of a lane quarter split the tile's 64-column boxes (even / odd), each with two staging buffers =====
        const int q = warp & 3, h = (warp - 2) >> 2;
        constexpr int BOXES = BN / 64, MYB = (BOXES + 1) / 2;      // boxes per tile, and per warp at most
        uint8_t* obuf = smem + S::kOutOff + (warp - 2) * 8192;     // two [32 rows][128 B] swizzled boxes
        const uint32_t acc_empty_leader0 = mapa(saddr(&acc_empty[0]), 0), acc_empty_leader1 = mapa(saddr(&acc_empty[1]), 0);
        const bool with_bias = P.epilogue == EPI_BIAS_RELU || P.epilogue == EPI_BIAS;
        int it = 0;
        int ob = 0;
        for (int tile = pair; tile < tiles; tile += pairs, ++it) {
            const int m0 = (tile / n_tiles) * (2 * BM) + (int)rank * BM, n0 = (tile % n_tiles) * BN;
            const int as = it & 1;
            const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < P.M;
            uint32_t gate[2 * MYB], outw[2 * MYB];
            if (P.epilogue == EPI_RELU_BITS) {                      // this warp's gates of the tile row, before the accumulator is ready
#pragma unroll
                for (int i = 0; i < 2 * MYB; ++i) {
                    const int c = 2 * (h + 2 * (i >> 1)) + (i & 1);
                    gate[i] = (row_ok && c < BN / 32 && n0 + c * 32 < P.N) ? __ldg(P.bits_in + (int64_t)row * P.ld_bits + ((n0 >> 5) + c)) : 0u;
                }
            }
            mbar_wait(&acc_full[as], aphase);
            tc_fence_after();
#pragma unroll
            for (int bi = 0; bi < MYB; ++bi) {
                const int bx = h + 2 * bi;                          // box of 64 columns
                if (bx < BOXES && n0 + bx * 64 < P.N) {             // warp-uniform
                    if (lane == 0) bulk_wait_read<1>();             // the store that last used this buffer has read it
                    __syncwarp();
                    uint8_t* line = obuf + ob * 4096 + lane * 128;
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc) {
                        const int c = 2 * bx + cc;
                        const int col0 = n0 + c * 32;
                        if (col0 < P.N) {
                            uint32_t v[32];
                            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c * 32), v);
                            float4 bv[8];                           // the bias loads fly while the TMEM load does
                            if (with_bias) {
#pragma unroll
                                for (int j = 0; j < 8; ++j)
                                    bv[j] = (col0 + 4 * j < P.N) ? __ldg(reinterpret_cast<const float4*>(P.bias + col0 + 4 * j))
                                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
                            }
                            tmem_ld_wait();
                            float f[32];
#pragma unroll
                            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
                            if (with_bias) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    f[4 * j] += bv[j].x; f[4 * j + 1] += bv[j].y; f[4 * j + 2] += bv[j].z; f[4 * j + 3] += bv[j].w;
                                }
                                if (P.epilogue == EPI_BIAS_RELU) {
#pragma unroll
                                    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
                                }
                            } else if (P.epilogue == EPI_RELU_BITS) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) f[j] = ((gate[2 * bi + cc] >> j) & 1u) ? f[j] : 0.f;
                            }
                            uint32_t pk[16], bits = 0u;
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                pk[j] = pack_bf16(f[2 * j], f[2 * j + 1]);
                                bits |= ((pk[j] & 0x7fffu) != 0u ? 1u : 0u) << (2 * j);
                                bits |= ((pk[j] & 0x7fff0000u) != 0u ? 1u : 0u) << (2 * j + 1);
                            }
                            outw[2 * bi + cc] = bits;               // after ReLU: non-zero == positive
                            // 16-byte chunk j of this thread's 64-byte half row goes to chunk (cc * 4 + j) ^ (row & 7) of its line
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int ch = ((cc << 2) + j) ^ (lane & 7);
                                *reinterpret_cast<uint4*>(line + (ch << 4)) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
                            }
                        }
                    }
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(&map_c, obuf + ob * 4096, n0 + bx * 64, m0 + q * 32);      // clipped at M and N
                        bulk_commit();
                    }
                    ob ^= 1;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(as ? acc_empty_leader1 : acc_empty_leader0);   // the accumulator is drained
            if (P.epilogue == EPI_BIAS_RELU && P.bits_out != nullptr && row_ok) {   // the gate words, off the critical path
#pragma unroll
                for (int i = 0; i < 2 * MYB; ++i) {
                    const int c = 2 * (h + 2 * (i >> 1)) + (i & 1);
                    if (c < BN / 32 && n0 + c * 32 < P.N) P.bits_out[(int64_t)row * P.ld_bits + ((n0 >> 5) + c)] = outw[i];
                }
            }
        }
        if (lane == 0) bulk_wait_all();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();            // the leader's MMAs read the peer's shared memory; nobody leaves before everybody is done
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 2 * BN);
    }
}

// ------------------------------------------------------------------------------------------------------------------
// wgrad, all layers of a tower in one launch (split-K over CTA pairs)
//   pair tile [256 n] x [<= 256 k] of dW; one tcgen05.mma.cta_group::2 of M = 256 per 16 batch rows.  CTA r of the pair
//   stages dZ columns n0 + 128 r .. +127 (A, MN-major) and X columns k0 + r bn/2 .. (B, MN-major), 64 batch rows per stage.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kMaxLayers = 8;
constexpr int WM = 2 * BM;                 // n-rows of a dW tile (UMMA M of the pair)
constexpr int WN = 256;                    // k-columns of a dW tile (UMMA N), at most
constexpr int kWStages = 6;

struct WgradLayer {
    int N, K;                 // dW is [N, K] (row pitch ld_dw floats)
    int n_tiles, k_tiles;     // tiles of 256 n x up to 256 k
    int tile_base;            // first tile id of this layer
    float* dW;
    int64_t ld_dw;
    float* db;                // [N] or nullptr
};
struct WgradParams {
    int n_layers, M, ksteps, total_tiles;
    int split, per;           // every tile's batch (k-step) range is cut into `split` chunks of `per` k-steps
    WgradLayer L[kMaxLayers];
    float* partials;          // [total_tiles][split][256][WN + 32] fp32 (used when split > 1)
};
struct WgradMaps {
    CUtensorMap dz[kMaxLayers];   // [M, N] bf16, box {64 n, 64 m}
    CUtensorMap x[kMaxLayers];    // [M, K] bf16, box {64 k, 64 m}
};
constexpr int kPartialPitch = WN + 32;     // columns 0..255: the dW tile; column 256: the bias-gradient column

struct Segment { int layer, n0, k0, bn, tile, ks0, ks1, seg; };

// Work unit u of a launch = (batch chunk u / total_tiles, tile u % total_tiles): chunk-major, so the pairs that run at the same
// time walk the SAME batch rows of dZ and X -- every operand block comes from DRAM once and is shared through L2 by the tiles
// that need it.  (The first version cut the tile-major (tile, k-step) space into equal ranges, classic stream-K: every pair
// started somewhere else in the batch and the Terabyte top tower read 1.1 GB from DRAM instead of 0.38 GB -- HBM-bound.)
__device__ __forceinline__ Segment segment_of(const WgradParams& P, int u) {
    Segment s;
    s.seg = u / P.total_tiles;
    s.tile = u - s.seg * P.total_tiles;
    s.ks0 = s.seg * P.per;
    s.ks1 = min(P.ksteps, s.ks0 + P.per);
    int l = 0;
    while (l + 1 < P.n_layers && P.L[l + 1].tile_base <= s.tile) ++l;
    s.layer = l;
    const int local = s.tile - P.L[l].tile_base;
    s.n0 = (local / P.L[l].k_tiles) * WM;
    s.k0 = (local % P.L[l].k_tiles) * WN;
    s.bn = (P.L[l].K - s.k0) > 128 ? 256 : 128;           // UMMA N; each CTA stages bn / 2 columns (whole 64-column blocks)
    return s;
}

struct WgradSmem {
    static constexpr uint32_t kABytes = BM * BK * 2, kBBytes = (WN / 2) * BK * 2;
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;            // 32 KB
    static constexpr uint32_t kOnesOff = kWStages * kStageBytes;          // 16 rows x 128 B of bf16 ones (K-major B operand)
    static constexpr uint32_t kBarOff = kOnesOff + 2048;
    static constexpr uint32_t kBytes = kBarOff + 256 + 1024;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
wgrad_splitk_kernel(const __grid_constant__ WgradMaps maps, const WgradParams P) {
    using S = WgradSmem;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
    uint64_t* empty = full + kWStages;
    uint64_t* acc_full = empty + kWStages;
    uint64_t* acc_empty = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int G = gridDim.x >> 1, pair = blockIdx.x >> 1;
    const int units = P.total_tiles * P.split;

    if (threadIdx.x == 0) {
        for (int l = 0; l < P.n_layers; ++l) { prefetch_tmap(&maps.dz[l]); prefetch_tmap(&maps.x[l]); }
        for (int i = 0; i < kWStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, 8);
        fence_barrier_init();
    }
    // the tile of ones behind the bias gradient: every element 1.0 (bf16 0x3f80), so its layout does not matter
    for (int i = threadIdx.x; i < 2048 / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem + S::kOnesOff)[i] = 0x3f803f80u;
    fence_proxy_async();           // generic-proxy writes -> visible to the tensor core's async proxy
    if (warp == 1) tmem_alloc_pair(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    grid_launch_dependents();
    grid_dependency_wait();

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int u = pair; u < units; u += G) {
                const Segment s = segment_of(P, u);
                const uint32_t bytes = S::kABytes + (uint32_t)(s.bn / 2) * (BK * 2);
                const int na = s.n0 + (int)rank * BM, kb0 = s.k0 + (int)rank * (s.bn / 2);
                for (int ks = s.ks0; ks < s.ks1; ++ks) {
                    mbar_wait(&empty[stage], phase ^ 1u);
                    uint8_t* sa = smem + stage * S::kStageBytes;
                    uint8_t* sb = sa + S::kABytes;
                    const uint32_t bar = mapa(saddr(&full[stage]), 0);
                    if (leader) mbar_arrive_expect_tx(&full[stage], 2 * bytes);
                    // A = dZ^t: [64 batch rows][64 n] blocks, two of them for this CTA's 128 n
                    tma_load_2d_pair(sa, &maps.dz[s.layer], bar, na, ks * BK);
                    tma_load_2d_pair(sa + kAtomBytes, &maps.dz[s.layer], bar, na + 64, ks * BK);
                    // B = X: [64 batch rows][64 k] blocks, bn / 128 of them for this CTA's half
                    for (int j = 0; j < s.bn / 128; ++j)
                        tma_load_2d_pair(sb + j * kAtomBytes, &maps.x[s.layer], bar, kb0 + 64 * j, ks * BK);
                    if (++stage == kWStages) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (leader) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            const uint32_t ones = saddr(smem + S::kOnesOff);
            for (int u = pair; u < units; u += G, ++it) {
                const Segment s = segment_of(P, u);
                const uint32_t idesc = make_idesc(s.bn, true, true, WM);
                const uint32_t idesc_b = make_idesc(16, true, false, WM);
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
                            umma_pair(tmem_base, desc_mnmajor(sa, k), desc_mnmajor(sb, k), idesc, acc);
                            if (with_bias) umma_pair(tmem_base + WN, desc_mnmajor(sa, k), make_desc(ones, 16, 1024), idesc_b, acc);
                        }
                        umma_commit_pair(&empty[stage]);
                        if (ks == s.ks1 - 1) umma_commit_pair(acc_full);
                    }
                    __syncwarp();
                    if (++stage == kWStages) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else {
        // ===== epilogue (both CTAs: their own 128 n-rows of the tile) =====
        const int q = warp & 3;
        const uint32_t acc_empty_leader = mapa(saddr(acc_empty), 0);
        const bool single = P.split == 1;
        int it = 0;
        for (int u = pair; u < units; u += G, ++it) {
            const Segment s = segment_of(P, u);
            const WgradLayer& L = P.L[s.layer];
            const bool with_bias = s.k0 == 0 && L.db != nullptr;
            const int r = (int)rank * BM + q * 32 + lane;  // row of the tile = n index
            const int n = s.n0 + r;
            mbar_wait(acc_full, (uint32_t)it & 1u);
            tc_fence_after();
            float* prow = P.partials + (((int64_t)s.tile * P.split + s.seg) * WM + r) * kPartialPitch;
            float* wrow = L.dW + (int64_t)n * L.ld_dw + s.k0;
            const int kw = min(WN, L.K - s.k0);
#pragma unroll 1
            for (int c = 0; c < s.bn / 32; ++c) {
                if (c * 32 >= kw) break;                   // columns past K are zero padding of the tile
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
            if (lane == 0) mbar_arrive_cluster(acc_empty_leader);     // the next unit's MMAs may overwrite the accumulator
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 512);
    }
}

// Second pass of the split-K wgrad: the `split` partials of every tile are summed in chunk order (a fixed order:
// deterministic) into dW / db.  Element-parallel over all SMs: CTA (tile, slice of 8 rows), 256 threads, float4 columns,
// coalesced, up to 8 partials in flight per thread.
constexpr int kReduceRows = 8;
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const WgradParams P) {
    grid_launch_dependents();
    grid_dependency_wait();
    const int tile = blockIdx.x / (WM / kReduceRows);
    const int r0 = (blockIdx.x % (WM / kReduceRows)) * kReduceRows;
    const int nsegs = P.split;
    int l = 0;
    while (l + 1 < P.n_layers && P.L[l + 1].tile_base <= tile) ++l;
    const WgradLayer& L = P.L[l];
    const int local = tile - L.tile_base;
    const int n0 = (local / L.k_tiles) * WM, k0 = (local % L.k_tiles) * WN;
    const int kw = min(WN, L.K - k0);
    const int q4 = kw >> 2;                                   // float4 columns of the tile (K % 8 == 0)
    const float* p0 = P.partials + (int64_t)tile * P.split * WM * kPartialPitch;
    const int64_t seg_stride = (int64_t)WM * kPartialPitch;
    const int rows = min(kReduceRows, L.N - n0 - r0);
    for (int idx = threadIdx.x; idx < rows * q4; idx += 256) {
        const int r = r0 + idx / q4, c = (idx % q4) << 2;
        const float* src = p0 + (int64_t)r * kPartialPitch + c;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s0 = 0; s0 < nsegs; s0 += 8) {
            float4 b[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (s0 + i < nsegs) b[i] = __ldcg(reinterpret_cast<const float4*>(src + (s0 + i) * seg_stride));
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (s0 + i < nsegs) { a.x += b[i].x; a.y += b[i].y; a.z += b[i].z; a.w += b[i].w; }
        }
        *reinterpret_cast<float4*>(L.dW + (int64_t)(n0 + r) * L.ld_dw + k0 + c) = a;
    }
    if (k0 == 0 && L.db != nullptr && (int)threadIdx.x < rows) {
        const int r = r0 + threadIdx.x;
        const float* src = p0 + (int64_t)r * kPartialPitch + WN;
        float a = 0.f;
        for (int sgi = 0; sgi < nsegs; ++sgi) a += __ldcg(src + sgi * seg_stride);
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

// Launch with programmatic stream serialization allowed (see grid_dependency_wait) when the launch is SHORT: at most
// kPdlMaxRows batch rows.  Measured inside the whole step (profiles/r02_pdl.txt): the 2048-row step gains 5% and the 128-row
// step 5%, because the next kernel's prologue (barrier init, TMEM allocation, descriptor prefetch) hides behind the
// previous kernel; the 32768-row step LOSES 2%, because the early-resident CTAs of the dependent spin in
// griddepcontrol.wait on the SMs whose tile queue ran dry, where the side stream's sparse-update CTAs used to run.
// DLRMX_NO_PDL=1 in the environment turns it off everywhere.
constexpr long long kPdlMaxRows = 8192;
template <typename... KArgs, typename... Args>
static void launch_pdl(bool allow, void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st,
                       Args&&... args) {
    static const bool env_off = getenv("DLRMX_NO_PDL") != nullptr && getenv("DLRMX_NO_PDL")[0] == '1';
    const bool off = env_off || !allow;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = off ? 0 : 1;
    cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

template <int BN, bool B_MN>
static int launch_pair(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const PairParams& P, cudaStream_t st) {
    using S = PairSmem<BN>;
    static DeviceOnce once;
    if (once.first())
        cudaFuncSetAttribute(gemm_pair_kernel<BN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::kBytes);
    const int tiles = (int)(ceil_div(P.M, 2 * BM) * ceil_div(P.N, BN));
    const int pairs = tiles < kSMs / 2 ? tiles : kSMs / 2;
    launch_pdl(P.M <= kPdlMaxRows, gemm_pair_kernel<BN, B_MN>, 2 * pairs, kPairThreads, S::kBytes, st, ma, mb, mc, P);
    return check_launch("gemm_bf16");
}

}  // namespace tc
}  // namespace dlrmx

extern "C" {

int dlrmx_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int b_is_kn, void* C, int64_t ldc, int M, int N,
                    int K, int epilogue, const float* bias, const void* relu_mask, int64_t ld_mask, uint32_t* relu_bits_out,
                    dlrmx_stream_t stream) {
    using namespace dlrmx;
    using namespace dlrmx::tc;
    DLRMX_REQUIRE(M >= 0 && N > 0 && K > 0 && N % 8 == 0 && K % 8 == 0, "gemm_bf16: need N %% 8 == 0 and K %% 8 == 0 (M=%d N=%d K=%d)", M, N, K);
    DLRMX_REQUIRE(A != nullptr && B != nullptr && C != nullptr && ldc % 8 == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0,
                  "gemm_bf16: C must be 16-byte aligned with ldc %% 8 == 0");
    DLRMX_REQUIRE(epilogue == EPI_NONE || epilogue == EPI_BIAS_RELU || epilogue == EPI_BIAS || epilogue == EPI_RELU_BITS,
                  "gemm_bf16: bad epilogue %d (0 none, 1 bias + ReLU, 3 bias, 4 ReLU gate bits)", epilogue);
    DLRMX_REQUIRE((epilogue != EPI_BIAS_RELU && epilogue != EPI_BIAS) || (bias != nullptr && (reinterpret_cast<uintptr_t>(bias) & 15) == 0),
                  "gemm_bf16: the bias epilogues need a 16-byte aligned bias");
    DLRMX_REQUIRE(epilogue != EPI_RELU_BITS || (relu_mask != nullptr && ld_mask >= (N + 31) / 32 && (reinterpret_cast<uintptr_t>(relu_mask) & 3) == 0),
                  "gemm_bf16: the ReLU-bits epilogue needs the gate words with ld_mask >= ceil(N / 32)");
    DLRMX_REQUIRE(relu_bits_out == nullptr || (epilogue == EPI_BIAS_RELU && ld_mask >= (N + 31) / 32),
                  "gemm_bf16: relu_bits_out goes with the bias + ReLU epilogue and ld_mask >= ceil(N / 32) words");
    if (M == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    CUtensorMap ma, mb;
    if (!make_map(&ma, A, M, K, lda, BK, BM)) return DLRMX_E_ARG;
    // CTA-pair kernel.  An MN-major B half must cover whole 64-column blocks: BN >= 128 there (columns past N are zero-filled)
    const int BN = N > 128 ? 256 : ((N > 64 || b_is_kn) ? 128 : 64);
    CUtensorMap mc;
    if (!make_map(&mc, C, M, N, ldc, 64, 32)) return DLRMX_E_ARG;                  // store boxes: 32 rows x 64 columns
    if (b_is_kn) {
        if (!make_map(&mb, B, K, N, ldb, 64, BK)) return DLRMX_E_ARG;
    } else {
        if (!make_map(&mb, B, N, K, ldb, BK, BN / 2)) return DLRMX_E_ARG;          // each CTA of the pair: BN / 2 rows x 64 k
    }
    PairParams P;
    P.M = M; P.N = N; P.K = K; P.bias = bias; P.bits_out = relu_bits_out;
    P.bits_in = (const uint32_t*)relu_mask; P.ld_bits = ld_mask; P.epilogue = epilogue;
    if (b_is_kn) {
        if (BN == 256) return launch_pair<256, true>(ma, mb, mc, P, st);
        return launch_pair<128, true>(ma, mb, mc, P, st);
    }
    if (BN == 256) return launch_pair<256, false>(ma, mb, mc, P, st);
    if (BN == 128) return launch_pair<128, false>(ma, mb, mc, P, st);
    return launch_pair<64, false>(ma, mb, mc, P, st);
}

static int wgrad_tiles(int n_layers, const int* N, const int* K) {
    int t = 0;
    for (int l = 0; l < n_layers; ++l) t += (int)(dlrmx::ceil_div(N[l], dlrmx::tc::WM) * dlrmx::ceil_div(K[l], dlrmx::tc::WN));
    return t;
}
// Split factor of the batch (k-step) range: the launch runs ceil(tiles * S / pairs) waves of units of ceil(ksteps / S) k-steps,
// each followed by an accumulator drain worth ~10 k-steps (single-buffered); the S with the least such time wins.
static void wgrad_split(int total_tiles, int M, int* split, int* per) {
    const int ksteps = (int)dlrmx::ceil_div(M, dlrmx::tc::BK), pairs = dlrmx::kSMs / 2;
    int best_s = 1;
    int64_t best = -1;
    for (int s = 1; s <= 32 && s <= ksteps; ++s) {
        const int64_t waves = dlrmx::ceil_div((int64_t)total_tiles * s, pairs);
        const int64_t cost = waves * (dlrmx::ceil_div(ksteps, s) + 10);
        if (best < 0 || cost < best) { best = cost; best_s = s; }
    }
    *per = (int)dlrmx::ceil_div(ksteps, best_s);
    *split = (int)dlrmx::ceil_div(ksteps, *per);          // no empty chunks
}

size_t dlrmx_gemm_wgrad_workspace_bytes(int n_layers, const int* N, const int* K, int M) {
    if (n_layers <= 0 || M <= 0) return 256;
    const int tiles = wgrad_tiles(n_layers, N, K);
    int split, per;
    wgrad_split(tiles, M, &split, &per);
    return (size_t)tiles * split * dlrmx::tc::WM * dlrmx::tc::kPartialPitch * 4 + 256;
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
        L.N = N[l]; L.K = K[l]; L.n_tiles = (int)ceil_div(N[l], WM); L.k_tiles = (int)ceil_div(K[l], WN);
        L.tile_base = tiles; L.dW = dW[l]; L.ld_dw = ld_dw[l]; L.db = db != nullptr ? db[l] : nullptr;
        tiles += L.n_tiles * L.k_tiles;
    }
    P.total_tiles = tiles;
    wgrad_split(tiles, M, &P.split, &P.per);
    P.partials = (float*)workspace;
    static DeviceOnce once;
    if (once.first())
        cudaFuncSetAttribute(wgrad_splitk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WgradSmem::kBytes);
    const int units = tiles * P.split;
    const int G = units < kSMs / 2 ? units : kSMs / 2;      // CTA pairs
    launch_pdl(M <= kPdlMaxRows, wgrad_splitk_kernel, 2 * G, kThreads, WgradSmem::kBytes, (cudaStream_t)stream, maps, P);
    if (int rc = check_launch("gemm_wgrad")) return rc;
    if (P.split > 1) {
        launch_pdl(M <= kPdlMaxRows, wgrad_reduce_kernel, tiles * (WM / kReduceRows), 256, 0, (cudaStream_t)stream, P);
        return check_launch("gemm_wgrad_reduce");
    }
    return DLRMX_OK;
}
}
