// Layout kernels either side of the model-parallel exchange.  sm_100a.
//
// Replaces the torch.cat / split / .contiguous() copies of All2All_Req / All2All_Wait
// (extend_distributed.py:424, 460, 482-487, 494-495) and the torch.cat of interact_features
// (dlrm_s_pytorch.py:629-637): the send side needs no packing at all because the lookup kernel already
// writes [B, T_r, D], which IS the concatenation over destination ranks of [B/W, T_r, D]; the receive
// side is one pass that drops every source rank's [lb, T_p, D] chunk into its feature slots of the
// [lb, 1+T, D] interaction input (rank-major table order, as the reference's cat produces), and the
// backward is the mirror image.  Pure copy: 2 * lb*T*D*4 bytes of HBM traffic.
#include <cuda_bf16.h>

#include "common.cuh"

namespace dlrmx {

constexpr int kMaxRanks = 64;
enum { DLRMX_TUNE_PUSH_GRID = 13 };
struct RankSplit {
    int W;
    int first_table[kMaxRanks + 1];  // exclusive prefix of tables per rank
};

// dir 0: recv chunks -> feature block ; dir 1: feature-block gradient -> send chunks
template <int VEC>
__global__ void __launch_bounds__(256)
a2a_layout_kernel(float* __restrict__ chunks, float* __restrict__ block, int64_t bsb, int64_t bsf, int feat_off,
                  RankSplit rs, int lb, int D, int dir) {
    const int T = rs.first_table[rs.W];
    const int dv = D / VEC;
    const int64_t total = (int64_t)lb * T * dv;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int d = (int)(e % dv);
        const int64_t r = e / dv;
        const int t = (int)(r % T);
        const int b = (int)(r / T);
        int p = 0;
        while (t >= rs.first_table[p + 1]) ++p;
        const int tp = rs.first_table[p + 1] - rs.first_table[p];
        // chunk p starts at lb*first_table[p]*D and is [lb, tp, D]
        float* c = chunks + ((int64_t)lb * rs.first_table[p] + (int64_t)b * tp + (t - rs.first_table[p])) * D + d * VEC;
        float* f = block + (int64_t)b * bsb + (int64_t)(feat_off + t) * bsf + d * VEC;      // bsf: feature pitch (>= D)
        RowVec<VEC> v;
        if (dir == 0) { v.load(c); v.store(f); }
        else          { v.load(f); v.store(c); }
    }
}

// Gradient half of the exchange WITHOUT a send buffer or a collective: each rank's slice of the feature-block
// gradient is stored straight into the owning rank's (peer-mapped) gradient buffer over NVLink, one 16-byte
// lane-store per thread, whole D-float rows contiguous -> full-line NVLink writes.
struct PeerBases {
    float* base[kMaxRanks];      // where this rank's [lb, T_p, D] chunk starts inside rank p's buffer
};

// OUT_BF16: the rows cross NVLink as bf16 (rounded to nearest even here, widened exactly by the update kernel that reads
// them: dlrmx_bwd_update_ex, grad_is_bf16) -- half the bytes of the one transfer the backward pass waits for.
template <int VEC, bool OUT_BF16 = false>
__global__ void __launch_bounds__(256)
a2a_push_kernel(const float* __restrict__ block, int64_t bsb, int64_t bsf, int feat_off, RankSplit rs, PeerBases pb, int lb,
                int D) {
    const int T = rs.first_table[rs.W];
    const int dv = D / VEC;
    const int64_t total = (int64_t)lb * T * dv;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int d = (int)(e % dv);
        const int64_t r = e / dv;
        const int t = (int)(r % T);
        const int b = (int)(r / T);
        int p = 0;
        while (t >= rs.first_table[p + 1]) ++p;
        const int tp = rs.first_table[p + 1] - rs.first_table[p];
        const int64_t coff = ((int64_t)b * tp + (t - rs.first_table[p])) * D + d * VEC;
        const float* f = block + (int64_t)b * bsb + (int64_t)(feat_off + t) * bsf + d * VEC;
        RowVec<VEC> v;
        v.load_stream(f);
        if (OUT_BF16) v.store(reinterpret_cast<__nv_bfloat16*>(pb.base[p]) + coff);
        else v.store(pb.base[p] + coff);
    }
}

static int a2a_layout(float* chunks, float* block, int64_t bsb, int64_t bsf, int feat_off, int W, const int* tpr, int lb,
                      int D, int dir, dlrmx_stream_t stream) {
    DLRMX_REQUIRE(W >= 1 && W <= kMaxRanks, "a2a layout: W=%d out of range", W);
    DLRMX_REQUIRE(lb >= 0 && D >= 1 && bsf >= D, "a2a layout: bad sizes (feature pitch %lld < D = %d)", (long long)bsf, D);
    RankSplit rs;
    rs.W = W;
    rs.first_table[0] = 0;
    for (int p = 0; p < W; ++p) {
        DLRMX_REQUIRE(tpr[p] >= 0, "a2a layout: negative table count");
        rs.first_table[p + 1] = rs.first_table[p] + tpr[p];
    }
    const int T = rs.first_table[W];
    if (lb == 0 || T == 0) return DLRMX_OK;
    const bool vec = (D % 4 == 0) && (bsb % 4 == 0) && (bsf % 4 == 0) && (((uintptr_t)chunks | (uintptr_t)block) & 15) == 0;
    const int64_t total = (int64_t)lb * T * (vec ? D / 4 : D);
    int64_t nblk = ceil_div(total, 256);
    if (nblk > (int64_t)kSMs * 16) nblk = (int64_t)kSMs * 16;
    const unsigned blocks = (unsigned)nblk;
    cudaStream_t st = (cudaStream_t)stream;
    if (vec) a2a_layout_kernel<4><<<blocks, 256, 0, st>>>(chunks, block, bsb, bsf, feat_off, rs, lb, D, dir);
    else     a2a_layout_kernel<1><<<blocks, 256, 0, st>>>(chunks, block, bsb, bsf, feat_off, rs, lb, D, dir);
    return check_launch(dir == 0 ? "a2a_unpack" : "a2a_pack_grad");
}

}  // namespace dlrmx

extern "C" {
static int a2a_push_impl(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                         const int* host_tables_per_rank, int lb, int D, float* const* host_peer_chunk_bases, bool out_bf16,
                         dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(W >= 1 && W <= kMaxRanks, "a2a_push_grad: W=%d out of range", W);
    DLRMX_REQUIRE(lb >= 0 && D >= 1 && dfeat_stride_f >= D && host_peer_chunk_bases != nullptr, "a2a_push_grad: bad arguments");
    RankSplit rs;
    PeerBases pb;
    rs.W = W;
    rs.first_table[0] = 0;
    bool vec = (D % 4 == 0) && (dfeat_stride_b % 4 == 0) && (dfeat_stride_f % 4 == 0) && ((uintptr_t)dfeat & 15) == 0;
    for (int p = 0; p < W; ++p) {
        DLRMX_REQUIRE(host_tables_per_rank[p] >= 0, "a2a_push_grad: negative table count");
        rs.first_table[p + 1] = rs.first_table[p] + host_tables_per_rank[p];
        pb.base[p] = host_peer_chunk_bases[p];
        DLRMX_REQUIRE(host_tables_per_rank[p] == 0 || pb.base[p] != nullptr, "a2a_push_grad: null peer base");
        vec = vec && ((uintptr_t)pb.base[p] & 15) == 0;
    }
    const int T = rs.first_table[W];
    if (lb == 0 || T == 0) return DLRMX_OK;
    const int64_t total = (int64_t)lb * T * (vec ? D / 4 : D);
    // tunable 13 = 1 (dlrm_net.set_mgpu_overlap): many short-lived CTAs (one or two elements per thread) rather than a looping
    // grid: the gradient all-reduces then run beside this kernel at a higher stream priority and get an SM slot whenever one
    // of these CTAs retires (with 16 CTAs per SM looping, the all-reduce kernel waited ~0.1 ms for its first slot:
    // profiles/r02_timeline_n8_peer_ar_rank0.txt).  Default: the looping grid every multi-GPU run so far was measured with.
    const int64_t cap = (int64_t)kSMs * (get_tunable(DLRMX_TUNE_PUSH_GRID) == 1 ? 512 : 16);
    int64_t nblk = ceil_div(total, 256);
    if (nblk > cap) nblk = cap;
    cudaStream_t st = (cudaStream_t)stream;
    if (out_bf16) {
        DLRMX_REQUIRE(vec && D % 8 == 0, "a2a_push_grad_bf16: needs D %% 8 == 0 and 16-byte aligned buffers and strides");
        const int64_t total8 = (int64_t)lb * T * (D / 8);
        int64_t nb8 = ceil_div(total8, 256);
        if (nb8 > cap) nb8 = cap;
        a2a_push_kernel<8, true><<<(unsigned)nb8, 256, 0, st>>>(dfeat, dfeat_stride_b, dfeat_stride_f, feat_off, rs, pb, lb, D);
        return check_launch("a2a_push_grad_bf16");
    }
    if (vec) a2a_push_kernel<4><<<(unsigned)nblk, 256, 0, st>>>(dfeat, dfeat_stride_b, dfeat_stride_f, feat_off, rs, pb, lb, D);
    else     a2a_push_kernel<1><<<(unsigned)nblk, 256, 0, st>>>(dfeat, dfeat_stride_b, dfeat_stride_f, feat_off, rs, pb, lb, D);
    return check_launch("a2a_push_grad");
}

int dlrmx_a2a_push_grad(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                        const int* host_tables_per_rank, int lb, int D, float* const* host_peer_chunk_bases,
                        dlrmx_stream_t stream) {
    return a2a_push_impl(dfeat, dfeat_stride_b, dfeat_stride_f, feat_off, W, host_tables_per_rank, lb, D, host_peer_chunk_bases,
                         false, stream);
}
int dlrmx_a2a_push_grad_bf16(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                             const int* host_tables_per_rank, int lb, int D, void* const* host_peer_chunk_bases_bf16,
                             dlrmx_stream_t stream) {
    return a2a_push_impl(dfeat, dfeat_stride_b, dfeat_stride_f, feat_off, W, host_tables_per_rank, lb, D,
                         reinterpret_cast<float* const*>(host_peer_chunk_bases_bf16), true, stream);
}

int dlrmx_a2a_unpack(const float* recv, int W, const int* host_tables_per_rank, int lb, int D, float* dst,
                     int64_t dst_stride_b, int64_t dst_stride_f, int feat_off, dlrmx_stream_t stream) {
    return dlrmx::a2a_layout(const_cast<float*>(recv), dst, dst_stride_b, dst_stride_f, feat_off, W, host_tables_per_rank, lb,
                             D, 0, stream);
}
int dlrmx_a2a_pack_grad(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                        const int* host_tables_per_rank, int lb, int D, float* send, dlrmx_stream_t stream) {
    return dlrmx::a2a_layout(send, const_cast<float*>(dfeat), dfeat_stride_b, dfeat_stride_f, feat_off, W,
                             host_tables_per_rank, lb, D, 1, stream);
}
}
