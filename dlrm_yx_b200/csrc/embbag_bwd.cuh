// Backward step 3: segmented reduce of pooled-row gradients over runs of equal (sorted) row ids, with
// the sparse optimizer update fused into the same pass.  sm_100a.
//
// Replaces EmbeddingBag.backward (sparse COO grad, one [D] entry per looked-up index) followed by
// torch.optim.SGD.step on sparse grads (dlrm_s_pytorch.py:1640, 1933) or RWSAdagrad.step's sparse
// branch (optim/rwsadagrad.py:97-113: coalesce, mom[r] += mean_d(g^2), W[r] -= clr*g/(sqrt(mom[r])+eps)).
//
// Work decomposition is by SORTED POSITION, not by row: a group of G = D/4 lanes owns one segment of
// 32 consecutive sorted positions, whatever runs they contain.  That keeps the load balanced under the
// Criteo skew (3-row tables with ~11 K duplicates per row next to 40 M-row tables with none):
//   * runs that start and end inside the segment are reduced in registers and the table row is
//     updated on the spot (read W row, read accumulator, write both) -- the common case;
//   * a run that crosses a segment boundary leaves per-segment partial sums (head/tail partials) and is
//     finished by `bwd_fixup_kernel`: the segment where the run STARTS adds the partials of the
//     following segments in ascending order and applies the update.  Fixed order -> deterministic.
// Within a batch of K positions the K gradient rows and the W rows of the runs ending there are all
// issued before any is consumed (K 16-B loads in flight per lane); run boundaries come from comparing
// adjacent keys (group-uniform broadcast loads, no run-offset array needed).
// HBM-bound: algorithmic bytes = T*B*D*4 (pooled-row gradients, re-reads hit L2) + N*8 (key, bag)
//            + 2*U*D*4 (table rows RMW) [+ 2*U*4 accumulators].
#pragma once
#include "common.cuh"

namespace dlrmx {

constexpr int kSeg = 32;  // sorted positions per group

struct BwdParams {
    int optimizer;
    int tab_bf16;         // 1: the table rows are bf16 (row pointers below still travel as float*, see tab_off): a row is
                          // widened to fp32 when it is read, updated in fp32 and rounded to nearest even when it is written
    float* const* tables;
    float* const* moms;
    const int64_t* row_base;
    int B, D;
    int64_t ld;           // row pitch of the tables in floats (>= D); the row-wise accumulator of a row may live in its
                          // padding (moms == nullptr: column D of the row), i.e. in the DRAM page the update opens anyway
    const float* grad;
    int grad_bf16;        // 1: the pooled-row gradients are bf16 (the pointer still travels as float*, see grad_off; strides in
                          // elements): the multi-GPU exchange moved them over NVLink in half the bytes; widened exactly when read
    int64_t stride_b, stride_t;
    const uint32_t* skey;
    const int32_t* sbag;
    int64_t N, nseg;
    int seg_len;          // sorted positions per group (multiple of 32)
    float lr, eps;
    float inv_D;          // 1/D when D is a power of two, else 0 (row_mean)
    float* head_partial;  // [nseg, D]
    float* tail_partial;  // [nseg, D]
    uint32_t invalid_key; // key of lookups whose index was out of range (dlrmx_bwd_sort): their run, sorted to the very end,
    float* scratch_row;   // is "applied" to this scratch row [D + 1] instead of a table row -- no branch in the hot loop
};

// Element arithmetic on a table-row pointer: fp32 rows advance 4 bytes per element, bf16 rows 2.
__device__ __forceinline__ float* tab_off(const BwdParams& P, float* p, int64_t elems) {
    return P.tab_bf16 ? reinterpret_cast<float*>(reinterpret_cast<__nv_bfloat16*>(p) + elems) : p + elems;
}
template <int VEC>
__device__ __forceinline__ void load_row(const BwdParams& P, RowVec<VEC>& w, const float* p) {
    if (P.tab_bf16) w.load(reinterpret_cast<const __nv_bfloat16*>(p));
    else w.load(p);
}
template <int VEC>
__device__ __forceinline__ void store_row(const BwdParams& P, const RowVec<VEC>& w, float* p) {
    if (P.tab_bf16) w.store(reinterpret_cast<__nv_bfloat16*>(p));
    else w.store(p);
}

// The same for the pooled-row gradients (read-only for the kernel's lifetime: non-coherent loads).
// GBF is a template flag of the update kernel (BwdParams::grad_bf16 selects the instantiation): as a run-time flag it
// cost the row-wise Adagrad kernel of 512-byte rows its spill-free 64 registers.
template <bool GBF>
__device__ __forceinline__ const float* grad_off(const float* p, int64_t elems) {
    return GBF ? reinterpret_cast<const float*>(reinterpret_cast<const __nv_bfloat16*>(p) + elems) : p + elems;
}
template <bool GBF, int VEC>
__device__ __forceinline__ void load_grad(RowVec<VEC>& g, const float* p) {
    if (GBF) g.load_stream(reinterpret_cast<const __nv_bfloat16*>(p));
    else g.load_ro(p);
}

// Where the run of `key` (looked up through bag `bag`) lives: table row and, for row-wise Adagrad, its accumulator.
__device__ __forceinline__ void run_target(const BwdParams& P, uint32_t key, int bag, float*& wp, float*& mp) {
    if (key == P.invalid_key) {
        wp = P.scratch_row;
        mp = P.scratch_row + P.D;
        return;
    }
    const int t = bag / P.B;
    const int64_t row = (int64_t)key - P.row_base[t];
    wp = tab_off(P, P.tables[t], row * P.ld);
    mp = nullptr;
    if (P.optimizer == DLRMX_OPT_RWSADAGRAD) mp = P.moms != nullptr ? P.moms[t] + row : wp + P.D;
}

template <int G>
__device__ __forceinline__ unsigned group_mask() {
    if constexpr (G == 32) {
        return 0xffffffffu;
    } else {
        return ((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1));
    }
}

// mean over the row of g^2 from its sum: a multiplication when D is a power of two (bit-identical to the division).
__device__ __forceinline__ float row_mean(const BwdParams& P, float ss) {
    return P.inv_D != 0.f ? __fmul_rn(ss, P.inv_D) : __fdiv_rn(ss, (float)P.D);
}

// Row-wise Adagrad step size  -lr / (sqrt(mom) + eps), IEEE sqrt and division, once per row.  (The MUFU.RSQ / MUFU.RCP
// approximations were tried: 16 fewer instructions per row and no measurable gain -- with Adagrad the kernel is bound
// by the extra random DRAM accesses of the 4-byte accumulators, not by this arithmetic.)
__device__ __forceinline__ float adagrad_step(float lr, float mom, float eps) {
    return __fdiv_rn(-lr, __fadd_rn(__fsqrt_rn(mom), eps));
}

// One table row: w <- optimizer(w, acc).  `w` was loaded by the caller (prefetched), m_old likewise.
template <int VEC, int G, int CH>
__device__ __forceinline__ void apply_row(const BwdParams& P, float* wp, float* mp, float m_old,
                                          RowVec<VEC> (&w)[CH], const RowVec<VEC> (&acc)[CH], int col0) {
    if (P.optimizer == DLRMX_OPT_SGD) {
        const float a = -P.lr;
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (col0 + c * G * VEC < P.D) {
                w[c].map(acc[c], [a](float wv, float gv) { return __fmaf_rn(a, gv, wv); });
                store_row<VEC>(P, w[c], tab_off(P, wp, c * G * VEC));
            }
    } else {
        float ss = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (col0 + c * G * VEC < P.D) ss = __fadd_rn(ss, acc[c].sumsq());
        const unsigned m = group_mask<G>();
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) ss = __fadd_rn(ss, __shfl_xor_sync(m, ss, o));
        const float mom = __fadd_rn(m_old, row_mean(P, ss));
        // ONE step size per row (-lr / std), then one FMA per element: w - (lr/std)*g differs from the reference's
        // w - lr*(g/std) by a few ulp of the update term, far inside the 1e-5 bar, and removes 4 IEEE divisions per lane
        // (the kernel was issue-bound: 164 warp instructions per sorted position, ncu r01 v2).
        const float a = adagrad_step(P.lr, mom, P.eps);
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (col0 + c * G * VEC < P.D) {
                w[c].map(acc[c], [a](float wv, float gv) { return __fmaf_rn(a, gv, wv); });
                store_row<VEC>(P, w[c], tab_off(P, wp, c * G * VEC));
            }
        if ((threadIdx.x & (G - 1)) == 0) *mp = mom;
    }
}

template <int VEC, int G, int CH>
__device__ __forceinline__ void store_partial(float* base, int64_t seg, int D, int col0,
                                              const RowVec<VEC> (&acc)[CH]) {
    float* p = base + seg * D + col0;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (col0 + c * G * VEC < D) acc[c].store(p + c * G * VEC);
}
}  // namespace dlrmx
