// All-reduce (sum) of a symmetric buffer over NVLink peer memory, in place, two-shot.  sm_100a.
//
// Replaces the data-parallel gradient all-reduce that DistributedDataParallel runs for the reference's MLP towers
// (dlrm_s_pytorch.py:1606-1613: DDP(bot_l) / DDP(top_l) on the model-parallel path) and that this repo adds for the
// replicated small tables' dense row gradients.  With NCCL those were three ring all-reduces per
// step, run one behind the other in issue order: 304 + 158 + 33 us on 8 GPUs (profiles/r02_timeline_n8_rank0.txt) -- the end
// of the step waited for them.
//
// Every rank holds its operand in a peer-mapped (symmetric) allocation of the same size.  Rank r owns the r-th slice of
// the buffer: it loads that slice from every rank (W independent 16-byte loads in flight per thread), adds them in rank
// order 0 .. W-1 in fp32 -- the same order on whichever rank does it, so the result does not depend on r -- and stores the
// sum back into that slice of EVERY rank's buffer.  Slices are disjoint, so no rank reads what another one writes.  The
// caller brackets the launch with two device barriers over the same ranks: before (every operand is complete, nobody
// still reads the previous result) and after (every remote store has landed).  bf16 buffers are widened, added in fp32
// and rounded once (a ring rounds after every hop).
#include <cuda_bf16.h>

#include "common.cuh"

namespace dlrmx {

constexpr int kMaxPeerRanks = 16;
struct PeerPtrs {
    uint4* p[kMaxPeerRanks];
};

__device__ __forceinline__ uint4 ld_sys(const uint4* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

template <bool BF16>
__device__ __forceinline__ void add_into(float (&acc)[8], const uint4& v) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    if (BF16) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[2 * i] = __fadd_rn(acc[2 * i], __uint_as_float(w[i] << 16));
            acc[2 * i + 1] = __fadd_rn(acc[2 * i + 1], __uint_as_float(w[i] & 0xffff0000u));
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i] = __fadd_rn(acc[i], __uint_as_float(w[i]));
    }
}

// n16: 16-byte vectors in the buffer; slice of rank r = [r * per, min(n16, (r + 1) * per))
template <bool BF16, int W>
__global__ void __launch_bounds__(256) peer_allreduce_kernel(PeerPtrs P, int rank, int64_t n16, int64_t per) {
    const int64_t lo = (int64_t)rank * per;
    const int64_t hi = lo + per < n16 ? lo + per : n16;
    for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x) {
        uint4 v[W];
#pragma unroll
        for (int p = 0; p < W; ++p) v[p] = ld_sys(P.p[p] + i);
        float acc[8];
        {   // the first operand starts the sum (x + 0 would turn -0 into +0)
            const uint32_t w[4] = {v[0].x, v[0].y, v[0].z, v[0].w};
            if (BF16) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    acc[2 * k] = __uint_as_float(w[k] << 16);
                    acc[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u);
                }
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[k] = __uint_as_float(w[k]);
            }
        }
#pragma unroll
        for (int p = 1; p < W; ++p) add_into<BF16>(acc, v[p]);
        uint4 o;
        if (BF16) {
            uint32_t w[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * k], acc[2 * k + 1]);
                w[k] = *reinterpret_cast<const uint32_t*>(&h);
            }
            o = make_uint4(w[0], w[1], w[2], w[3]);
        } else {
            o = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
        }
#pragma unroll
        for (int p = 0; p < W; ++p) P.p[(rank + p) % W][i] = o;      // every rank starts with a different destination
    }
}

template <bool BF16>
static int launch_peer_allreduce(const PeerPtrs& P, int W, int rank, int64_t n16, cudaStream_t st) {
    const int64_t per = ceil_div(n16, (int64_t)W);
    int64_t blocks = ceil_div(per, (int64_t)256);
    if (blocks > (int64_t)kSMs * 4) blocks = (int64_t)kSMs * 4;
    if (blocks < 1) blocks = 1;
    const unsigned g = (unsigned)blocks;
    switch (W) {
#define DLRMX_PAR_CASE(w) case w: peer_allreduce_kernel<BF16, w><<<g, 256, 0, st>>>(P, rank, n16, per); break;
        DLRMX_PAR_CASE(1) DLRMX_PAR_CASE(2) DLRMX_PAR_CASE(3) DLRMX_PAR_CASE(4) DLRMX_PAR_CASE(5) DLRMX_PAR_CASE(6)
        DLRMX_PAR_CASE(7) DLRMX_PAR_CASE(8) DLRMX_PAR_CASE(16)
#undef DLRMX_PAR_CASE
        default:
            set_error("peer_allreduce: W=%d not supported (1..8, 16)", W);
            return DLRMX_E_ARG;
    }
    return check_launch("peer_allreduce");
}

}  // namespace dlrmx

extern "C" int dlrmx_peer_allreduce(void* const* host_peer_buffers, int W, int rank, int64_t numel, int is_bf16,
                                    dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(W >= 1 && W <= kMaxPeerRanks && rank >= 0 && rank < W && host_peer_buffers != nullptr,
                  "peer_allreduce: bad ranks (W=%d rank=%d)", W, rank);
    const int per16 = is_bf16 ? 8 : 4;
    DLRMX_REQUIRE(numel >= 0 && numel % per16 == 0, "peer_allreduce: numel must be a multiple of %d (16-byte vectors)", per16);
    PeerPtrs P;
    for (int p = 0; p < W; ++p) {
        DLRMX_REQUIRE(host_peer_buffers[p] != nullptr && ((uintptr_t)host_peer_buffers[p] & 15) == 0,
                      "peer_allreduce: peer buffer %d null or not 16-byte aligned", p);
        P.p[p] = (uint4*)host_peer_buffers[p];
    }
    if (numel == 0) return DLRMX_OK;
    const int64_t n16 = numel / per16;
    return is_bf16 ? launch_peer_allreduce<true>(P, W, rank, n16, (cudaStream_t)stream)
                   : launch_peer_allreduce<false>(P, W, rank, n16, (cudaStream_t)stream);
}
