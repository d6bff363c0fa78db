// Kernels + C ABI of the fused backward/update (design notes in embbag_bwd.cuh).
#include "embbag_bwd.cuh"

namespace dlrmx {

template <typename T, int G>
__device__ __forceinline__ T bshfl64(unsigned mask, T v, int src) {
    unsigned long long u = (unsigned long long)v;
    unsigned lo = __shfl_sync(mask, (unsigned)u, src, G);
    unsigned hi = __shfl_sync(mask, (unsigned)(u >> 32), src, G);
    return (T)(((unsigned long long)hi << 32) | lo);
}

// One group of G lanes per segment of kSeg sorted positions, walked as kSeg/G sub-tiles.  Per sub-tile the
// metadata is fetched LANE-PARALLEL with coalesced loads (lane i: key, bag, is-last-of-run, gradient pointer and,
// for run ends, the table-row / accumulator pointers), so the key -> bag -> pointer chain costs one round of
// memory latency per G positions; positions are then consumed K at a time with K gradient rows (and the table
// rows of the runs ending there) in flight per lane.
//
// Runs that cross a segment boundary INSIDE the CTA (its 256/G groups own consecutive segments) are finished by the CTA
// itself: every group parks its head / tail partial in shared memory, and after one barrier the group where such a run
// starts adds the following groups' head partials in ascending order and applies the update.  Only runs that cross a
// CTA boundary leave partials in global memory (one head / one tail row per CTA) for `bwd_fixup_kernel`, whose chains
// are 256/G times shorter than with per-segment partials.
enum { kHasHead = 1, kHeadClosed = 2, kHasTail = 4, kActive = 8 };

// EXACT: D == VEC*G*CH, every lane owns columns of the row (no per-lane guard in the loop).
template <int VEC, int G, int CH, int K, int OCC = 0, bool EXACT = false, bool GBF = false>
__global__ void __launch_bounds__(256, OCC > 0 ? OCC : ((K * CH <= 4) ? 3 : 2)) bwd_update_kernel(const BwdParams P) {
    constexpr int GPC = 256 / G;             // groups (= segments) per CTA
    extern __shared__ __align__(16) float sh_part[];     // [2][GPC][D]: head partials, tail partials
    __shared__ unsigned char sh_flag[GPC];
    const int gl = threadIdx.x & (G - 1);
    const int gi = threadIdx.x / G;
    const unsigned gmask = group_mask<G>();
    const int gshift = (threadIdx.x & 31) & ~(G - 1);
    const int64_t seg = (int64_t)blockIdx.x * GPC + gi;
    const bool active = seg < P.nseg;
    const int64_t p0 = seg * P.seg_len, p1 = active ? min(P.N, p0 + P.seg_len) : p0;
    const int col0 = gl * VEC;
    const uint32_t* __restrict__ skey = P.skey;
    const int32_t* __restrict__ sbag = P.sbag;
    float* sh_head = sh_part + (size_t)gi * P.D;
    float* sh_tail = sh_part + (size_t)(GPC + gi) * P.D;
    unsigned flags = active ? kActive : 0;

    bool owned = true;                       // does the first run START in this segment?
    if (active && p0 > 0) owned = skey[p0 - 1] != skey[p0];
    if (!owned) flags |= kHasHead;
    RowVec<VEC> acc[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c].zero();
    bool open = false;                       // does the position walked last belong to a run that is not finished?

#pragma unroll 1
    for (int64_t pt = p0; pt < p1; pt += G) {
        const int nvalid = (int)min((int64_t)G, p1 - pt);
        // ---- lane-parallel metadata ---------------------------------------------------------------------
        const int64_t p = pt + gl;
        bool last = false;
        const float* gptr = nullptr;
        float* wptr = nullptr;
        float* mptr = nullptr;
        if (gl < nvalid) {
            const uint32_t key = skey[p];
            const uint32_t knext = (p + 1 < P.N) ? skey[p + 1] : ~key;
            const int bag = sbag[p];
            last = knext != key;
            const int t = bag / P.B;
            const int b = bag - t * P.B;
            gptr = grad_off<GBF>(P.grad, (int64_t)b * P.stride_b + (int64_t)t * P.stride_t);    // row start; col0 added after broadcast
            if (last) run_target(P, key, bag, wptr, mptr);
        }
        const unsigned lastmask = (__ballot_sync(gmask, last) >> gshift);

        for (int j0 = 0; j0 < nvalid; j0 += K) {
            RowVec<VEC> g[K][CH], w[K][CH];
            float* wp[K];
            float* mp[K];
            float mo[K];
            bool v[K], l[K];
            // ---- stage 1: every load of the batch is issued before anything is consumed -------------------
#pragma unroll
            for (int j = 0; j < K; ++j) {
                const int jj = j0 + j;
                v[j] = jj < nvalid;
                l[j] = v[j] && ((lastmask >> jj) & 1u);
                wp[j] = nullptr;
                mp[j] = nullptr;
                mo[j] = 0.f;
                const float* gp = grad_off<GBF>(bshfl64<const float*, G>(gmask, gptr, jj & (G - 1)), col0);
                if (v[j]) {
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (EXACT || col0 + c * G * VEC < P.D) load_grad<GBF, VEC>(g[j][c], grad_off<GBF>(gp, c * G * VEC));
                }
                if (l[j]) {
                    wp[j] = tab_off(P, bshfl64<float*, G>(gmask, wptr, jj & (G - 1)), col0);
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (EXACT || col0 + c * G * VEC < P.D) load_row<VEC>(P, w[j][c], tab_off(P, wp[j], c * G * VEC));
                    if (P.optimizer == DLRMX_OPT_RWSADAGRAD) {
                        mp[j] = bshfl64<float*, G>(gmask, mptr, jj & (G - 1));
                        mo[j] = *mp[j];
                    }
                }
            }
            // ---- stage 2: running sums in place, g[j] <- sum of the run up to position j (same order of
            //      additions as a sequential walk: acc + g0 + g1 + ...) ---------------------------------------
#pragma unroll
            for (int j = 0; j < K; ++j) {
                if (v[j]) {
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (EXACT || col0 + c * G * VEC < P.D) {
                            if (j == 0) g[0][c].add(acc[c]);
                            else if (!l[j > 0 ? j - 1 : 0]) g[j][c].add(g[j > 0 ? j - 1 : 0][c]);
                        }
                }
            }
            // ---- stage 3: per-row step sizes for all run ends of the batch at once (the K butterflies, square
            //      roots and divisions are independent -> they overlap instead of forming one long chain) ------
            const unsigned ends = (lastmask >> j0) & ((1u << K) - 1u);
            float a[K], mnew[K];
#pragma unroll
            for (int j = 0; j < K; ++j) { a[j] = -P.lr; mnew[j] = 0.f; }
            if (P.optimizer == DLRMX_OPT_RWSADAGRAD && ends != 0u) {
                float ss[K];
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    ss[j] = 0.f;
                    if (l[j]) {
#pragma unroll
                        for (int c = 0; c < CH; ++c)
                            if (EXACT || col0 + c * G * VEC < P.D) ss[j] = __fadd_rn(ss[j], g[j][c].sumsq());
                    }
                }
#pragma unroll
                for (int o = G / 2; o > 0; o >>= 1) {
#pragma unroll
                    for (int j = 0; j < K; ++j) ss[j] = __fadd_rn(ss[j], __shfl_xor_sync(gmask, ss[j], o));
                }
#pragma unroll
                for (int j = 0; j < K; ++j)
                    if (l[j]) {
                        mnew[j] = __fadd_rn(mo[j], row_mean(P, ss[j]));
                        a[j] = adagrad_step(P.lr, mnew[j], P.eps);
                    }
            }
            // ---- stage 4: apply / park the finished runs -----------------------------------------------------
#pragma unroll
            for (int j = 0; j < K; ++j) {
                if (l[j]) {
                    if (owned) {
                        const float aj = a[j];
#pragma unroll
                        for (int c = 0; c < CH; ++c)
                            if (EXACT || col0 + c * G * VEC < P.D) {
                                w[j][c].map(g[j][c], [aj](float wv, float gv) { return __fmaf_rn(aj, gv, wv); });
                                store_row<VEC>(P, w[j][c], tab_off(P, wp[j], c * G * VEC));
                            }
                        if (P.optimizer == DLRMX_OPT_RWSADAGRAD && gl == 0) *mp[j] = mnew[j];
                    } else {
                        store_partial<VEC, G, CH>(sh_head, 0, P.D, col0, g[j]);
                        flags |= kHeadClosed;
                    }
                    owned = true;
                }
            }
            // ---- stage 5: carry the open run into the next batch ----------------------------------------------
#pragma unroll
            for (int j = 0; j < K; ++j) {
                if (v[j]) {
#pragma unroll
                    for (int c = 0; c < CH; ++c) {
                        acc[c] = g[j][c];
                        if (l[j]) acc[c].zero();
                    }
                    open = !l[j];
                }
            }
        }
    }
    if (open) {   // the segment ends inside a run
        store_partial<VEC, G, CH>(owned ? sh_tail : sh_head, 0, P.D, col0, acc);
        if (owned) flags |= kHasTail;
    }
    if (gl == 0) sh_flag[gi] = (unsigned char)flags;
    __syncthreads();

    // ---- runs crossing segment boundaries inside this CTA ------------------------------------------------------
    if ((flags & kHasTail) != 0) {
        // acc still holds my tail partial; add the head partials of the groups after me, in order
        int v = gi + 1;
        bool closed = false;
        for (; v < GPC && (sh_flag[v] & kActive); ++v) {
            const float* hp = sh_part + (size_t)v * P.D + col0;
#pragma unroll
            for (int c = 0; c < CH; ++c)
                if (EXACT || col0 + c * G * VEC < P.D) {
                    RowVec<VEC> h;
                    h.load(hp + c * G * VEC);
                    acc[c].add(h);
                }
            if (sh_flag[v] & kHeadClosed) { closed = true; break; }
        }
        if (closed) {
            float* wp;
            float* mp;
            run_target(P, skey[p1 - 1], sbag[p1 - 1], wp, mp);
            wp = tab_off(P, wp, col0);
            RowVec<VEC> w[CH];
#pragma unroll
            for (int c = 0; c < CH; ++c)
                if (EXACT || col0 + c * G * VEC < P.D) load_row<VEC>(P, w[c], tab_off(P, wp, c * G * VEC));
            const float mo = (P.optimizer == DLRMX_OPT_RWSADAGRAD) ? *mp : 0.f;
            apply_row<VEC, G, CH>(P, wp, mp, mo, w, acc, col0);
        } else {                                   // the run leaves the CTA: one tail partial per CTA
            store_partial<VEC, G, CH>(P.tail_partial, blockIdx.x, P.D, col0, acc);
        }
    }
    if (gi == 0 && (flags & kHasHead) != 0) {      // the run that entered the CTA from the left: one head partial per CTA
        RowVec<VEC> hacc[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            hacc[c].zero();
            if (EXACT || col0 + c * G * VEC < P.D) hacc[c].load(sh_part + col0 + c * G * VEC);
        }
        if ((flags & kHeadClosed) == 0) {
            for (int v = 1; v < GPC && (sh_flag[v] & kActive); ++v) {
                const float* hp = sh_part + (size_t)v * P.D + col0;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (EXACT || col0 + c * G * VEC < P.D) {
                        RowVec<VEC> h;
                        h.load(hp + c * G * VEC);
                        hacc[c].add(h);
                    }
                if (sh_flag[v] & kHeadClosed) break;
            }
        }
        store_partial<VEC, G, CH>(P.head_partial, blockIdx.x, P.D, col0, hacc);
    }
}

// Finishes runs that cross segment boundaries.  One group per segment; only the segment where a crossing run
// STARTS does work: it adds, in ascending order, the head partials of the following segments the run covers.
// Which segments those are is found lane-parallel: lane i tests whether segment s+i lies entirely inside the run
// (two key loads), a ballot gives the length of the covered stretch, and the partial rows are then fetched
// up to 32 at a time -- so a hot row spanning hundreds of segments costs a few memory latencies per G segments.
template <int VEC, int G, int CH>
__global__ void __launch_bounds__(256) bwd_fixup_kernel(const BwdParams P) {
    constexpr int KF = (CH == 1) ? 32 : 8;   // partial rows in flight: the chain walk is pure latency
    const int gl = threadIdx.x & (G - 1);
    const unsigned gmask = group_mask<G>();
    const int gshift = (threadIdx.x & 31) & ~(G - 1);
    const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (seg >= P.nseg) return;
    const int64_t p0 = seg * P.seg_len, p1 = min(P.N, p0 + P.seg_len);
    const uint32_t* __restrict__ skey = P.skey;
    if (p1 >= P.N) return;
    const uint32_t key = skey[p1 - 1];
    const uint32_t kafter = skey[p1];
    const uint32_t kfirst = skey[p0];
    const uint32_t kbefore = skey[p0 > 0 ? p0 - 1 : 0];
    if (kafter != key) return;                                        // no run crosses my end
    if (kfirst == key && p0 > 0 && kbefore == key) return;            // I am a middle segment
    const int col0 = gl * VEC;

    RowVec<VEC> acc[CH];
    {
        const float* tp = P.tail_partial + seg * P.D + col0;
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (col0 + c * G * VEC < P.D) acc[c].load(tp + c * G * VEC);
    }
    bool done = false;
    int64_t s = seg + 1;
    while (!done) {
        // lane i: is segment s+i entirely inside the run, with the run continuing past its end?
        const int64_t sj = s + gl;
        const int64_t e = min(P.N, (sj + 1) * P.seg_len);
        bool full = false;
        if (sj < P.nseg && e < P.N) {
            const uint32_t a = skey[e - 1], bnext = skey[e];
            full = (a == key) & (bnext == key);
        }
        const unsigned fullmask = (__ballot_sync(gmask, full) >> gshift) & ((G == 32) ? 0xffffffffu : ((1u << G) - 1u));
        const int nfull = (fullmask == ((G == 32) ? 0xffffffffu : ((1u << G) - 1u))) ? G : (__ffs(~fullmask) - 1);
        done = nfull < G;
        const int cnt = done ? nfull + 1 : G;      // the stretch ends with the segment holding the run's tail piece
        for (int q0 = 0; q0 < cnt; q0 += KF) {
            RowVec<VEC> h[KF][CH];
#pragma unroll
            for (int j = 0; j < KF; ++j)
                if (q0 + j < cnt) {
                    const float* hp = P.head_partial + (s + q0 + j) * P.D + col0;
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (col0 + c * G * VEC < P.D) h[j][c].load(hp + c * G * VEC);
                }
#pragma unroll
            for (int j = 0; j < KF; ++j)
                if (q0 + j < cnt) {
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (col0 + c * G * VEC < P.D) acc[c].add(h[j][c]);
                }
        }
        s += cnt;
    }
    float* wp;
    float* mp;
    run_target(P, key, P.sbag[p1 - 1], wp, mp);
    wp = tab_off(P, wp, col0);
    RowVec<VEC> w[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (col0 + c * G * VEC < P.D) load_row<VEC>(P, w[c], tab_off(P, wp, c * G * VEC));
    const float mo = (P.optimizer == DLRMX_OPT_RWSADAGRAD) ? *mp : 0.f;
    apply_row<VEC, G, CH>(P, wp, mp, mo, w, acc, col0);
}

// Parity seam: gsum[u,:] = sum of the gradient rows of run u, sequential in sorted order.
template <int VEC, int G, int CH>
__global__ void __launch_bounds__(256)
bwd_coalesce_kernel(int B, int D, const float* __restrict__ grad, int64_t stride_b, int64_t stride_t,
                    const int32_t* __restrict__ sbag, const int32_t* __restrict__ run_off,
                    const int32_t* __restrict__ num_unique, float* __restrict__ gsum) {
    constexpr int K = 4;
    const int gl = threadIdx.x & (G - 1);
    const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (u >= *num_unique) return;
    const int col0 = gl * VEC;
    const int p0 = run_off[u], p1 = run_off[u + 1];
    RowVec<VEC> acc[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c].zero();
    for (int pb = p0; pb < p1; pb += K) {
        RowVec<VEC> g[K][CH];
#pragma unroll
        for (int j = 0; j < K; ++j)
            if (pb + j < p1) {
                const int bag = sbag[pb + j];
                const int t = bag / B, b = bag - t * B;
                const float* gp = grad + (int64_t)b * stride_b + (int64_t)t * stride_t + col0;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (col0 + c * G * VEC < D) g[j][c].load_ro(gp + c * G * VEC);
            }
#pragma unroll
        for (int j = 0; j < K; ++j)
            if (pb + j < p1) {
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (col0 + c * G * VEC < D) acc[c].add(g[j][c]);
            }
    }
    float* o = gsum + u * D + col0;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (col0 + c * G * VEC < D) acc[c].store(o + c * G * VEC);
}


// Non-fused optimizer on an already coalesced gradient (unique rows + summed gradient rows): the GPU
// form of torch.optim.SGD / RWSAdagrad.step's sparse branch (optim/rwsadagrad.py:97-113) for ONE table.
// The reference refuses rwsadagrad on GPU (dlrm_s_pytorch.py:1636-1637); this is what it would run.
template <int VEC, int G, int CH>
__global__ void __launch_bounds__(256)
rows_update_kernel(int optimizer, float* __restrict__ W, float* __restrict__ mom,
                   const int64_t* __restrict__ rows, const float* __restrict__ gsum, int64_t U, int D,
                   float lr, float eps, int tab_bf16) {
    const int gl = threadIdx.x & (G - 1);
    const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (u >= U) return;
    const int col0 = gl * VEC;
    const int64_t row = rows[u];
    BwdParams P;
    P.optimizer = optimizer; P.D = D; P.ld = D; P.lr = lr; P.eps = eps; P.inv_D = (D & (D - 1)) == 0 ? 1.0f / (float)D : 0.f;
    P.invalid_key = 0xffffffffu; P.scratch_row = nullptr; P.tab_bf16 = tab_bf16; P.grad_bf16 = 0;
    RowVec<VEC> g[CH], w[CH];
    float* wp = tab_off(P, W, row * D + col0);
    const float* gp = gsum + u * D + col0;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (col0 + c * G * VEC < D) {
            g[c].load_ro(gp + c * G * VEC);
            load_row<VEC>(P, w[c], tab_off(P, wp, c * G * VEC));
        } else {
            g[c].zero();
        }
    float* mp = nullptr;
    float mo = 0.f;
    if (optimizer == DLRMX_OPT_RWSADAGRAD) {
        mp = mom + row;
        mo = *mp;
    }
    apply_row<VEC, G, CH>(P, wp, mp, mo, w, g, col0);
}

// Replicated (data-parallel) tables: gsum [rows, D] holds the pooled-row gradients of the whole global batch summed per
// row (every rank accumulated its local batch into a zeroed buffer, the buffers were all-reduced).  Each group of G lanes
// takes rows r, r + stride, ...: a row whose summed gradient is all zero was not looked up (or cancels exactly) and keeps
// its bits -- as in the reference's sparse step, where only coalesced rows are touched (optim/rwsadagrad.py:97-113) --
// every other row gets the optimizer step, and its gradient row is cleared for the next step.
template <int VEC, int G, int CH>
__global__ void __launch_bounds__(256)
dense_rows_update_kernel(int optimizer, float* __restrict__ W, float* __restrict__ mom, float* __restrict__ gsum,
                         int64_t rows, int D, int64_t ld, float lr, float eps, int tab_bf16) {
    const int gl = threadIdx.x & (G - 1);
    const unsigned gmask = group_mask<G>();
    const int col0 = gl * VEC;
    BwdParams P;
    P.optimizer = optimizer; P.D = D; P.ld = ld; P.lr = lr; P.eps = eps; P.inv_D = (D & (D - 1)) == 0 ? 1.0f / (float)D : 0.f;
    P.invalid_key = 0xffffffffu; P.scratch_row = nullptr; P.tab_bf16 = tab_bf16; P.grad_bf16 = 0;
    const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) / G;
    for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G; r < rows; r += ngroups) {
        RowVec<VEC> g[CH], w[CH];
        float* gp = gsum + r * D + col0;
        bool nz = false;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            g[c].zero();
            if (col0 + c * G * VEC < D) {
                g[c].load(gp + c * G * VEC);
                nz = nz || g[c].sumsq() != 0.f;
            }
        }
        if (__ballot_sync(gmask, nz) == 0u) continue;           // group-uniform: the whole row is zero
        float* wp = tab_off(P, W, r * ld + col0);
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (col0 + c * G * VEC < D) load_row<VEC>(P, w[c], tab_off(P, wp, c * G * VEC));
        float* mp = nullptr;
        float mo = 0.f;
        if (optimizer == DLRMX_OPT_RWSADAGRAD) {
            mp = mom + r;
            mo = *mp;
        }
        apply_row<VEC, G, CH>(P, wp, mp, mo, w, g, col0);
        RowVec<VEC> z;
        z.zero();
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (col0 + c * G * VEC < D) z.store(gp + c * G * VEC);
    }
}

template <int VEC, int G, int CH>
static int launch_dense_rows_update(int optimizer, float* W, float* mom, float* gsum, int64_t rows, int D, int64_t ld,
                                    float lr, float eps, cudaStream_t st) {
    int64_t blocks = ceil_div(rows * G, 256);
    if (blocks > (int64_t)kSMs * 8) blocks = (int64_t)kSMs * 8;
    dense_rows_update_kernel<VEC, G, CH><<<(unsigned)blocks, 256, 0, st>>>(optimizer, W, mom, gsum, rows, D, ld, lr, eps, 0);
    return check_launch("dense_rows_update");
}

template <int VEC, int G, int CH>
static int launch_rows_update(int optimizer, float* W, float* mom, const int64_t* rows, const float* gsum,
                              int64_t U, int D, float lr, float eps, cudaStream_t st) {
    const unsigned blocks = (unsigned)ceil_div(U * G, 256);
    rows_update_kernel<VEC, G, CH><<<blocks, 256, 0, st>>>(optimizer, W, mom, rows, gsum, U, D, lr, eps, 0);
    return check_launch("rows_update");
}

template <int VEC, int G, int CH, int K = (CH == 1 ? 4 : 2), int OCC = 0>
static int launch_update(const BwdParams& P, cudaStream_t st) {
    const int threads = 256;
    constexpr int GPC = 256 / G;
    const unsigned blocks = (unsigned)ceil_div(P.nseg, (int64_t)GPC);
    const size_t smem = (size_t)2 * GPC * P.D * sizeof(float);
    const bool exact = P.D == VEC * G * CH;
    if (P.grad_bf16) {
        if (exact) bwd_update_kernel<VEC, G, CH, K, OCC, true, true><<<blocks, threads, smem, st>>>(P);
        else bwd_update_kernel<VEC, G, CH, K, OCC, false, true><<<blocks, threads, smem, st>>>(P);
    } else {
        if (exact) bwd_update_kernel<VEC, G, CH, K, OCC, true, false><<<blocks, threads, smem, st>>>(P);
        else bwd_update_kernel<VEC, G, CH, K, OCC, false, false><<<blocks, threads, smem, st>>>(P);
    }
    int rc = check_launch("bwd_update");
    if (rc) return rc;
    // the fix-up works on CTA tiles: GPC segments each, one head / tail partial per tile
    BwdParams Q = P;
    Q.seg_len = P.seg_len * GPC;
    Q.nseg = blocks;
    bwd_fixup_kernel<VEC, G, CH><<<(unsigned)ceil_div((int64_t)blocks * G, threads), threads, 0, st>>>(Q);
    return check_launch("bwd_fixup");
}

template <int VEC, int G, int CH>
static int launch_coalesce(int B, int D, const float* grad, int64_t sb, int64_t stt, const int32_t* sbag,
                           const int32_t* run_off, const int32_t* num_unique, int64_t N, float* gsum,
                           cudaStream_t st) {
    const int threads = 256;
    const unsigned blocks = (unsigned)ceil_div(N * G, threads);
    bwd_coalesce_kernel<VEC, G, CH><<<blocks, threads, 0, st>>>(B, D, grad, sb, stt, sbag, run_off, num_unique, gsum);
    return check_launch("bwd_coalesce");
}

#define DLRMX_GEOM_DISPATCH(g, CALL)                                                  \
    do {                                                                              \
        if (g.vec == 4) {                                                             \
            if (g.g == 1 && g.ch == 1) return CALL(4, 1, 1);                          \
            if (g.g == 2 && g.ch == 1) return CALL(4, 2, 1);                          \
            if (g.g == 4 && g.ch == 1) return CALL(4, 4, 1);                          \
            if (g.g == 8 && g.ch == 1) return CALL(4, 8, 1);                          \
            if (g.g == 16 && g.ch == 1) return CALL(4, 16, 1);                        \
            if (g.g == 32 && g.ch == 1) return CALL(4, 32, 1);                        \
            if (g.g == 32 && g.ch == 2) return CALL(4, 32, 2);                        \
            if (g.g == 32 && g.ch == 4) return CALL(4, 32, 4);                        \
        } else {                                                                      \
            if (g.g == 1 && g.ch == 1) return CALL(1, 1, 1);                          \
            if (g.g == 2 && g.ch == 1) return CALL(1, 2, 1);                          \
            if (g.g == 4 && g.ch == 1) return CALL(1, 4, 1);                          \
            if (g.g == 8 && g.ch == 1) return CALL(1, 8, 1);                          \
            if (g.g == 16 && g.ch == 1) return CALL(1, 16, 1);                        \
            if (g.g == 32 && g.ch == 1) return CALL(1, 32, 1);                        \
            if (g.g == 32 && g.ch == 2) return CALL(1, 32, 2);                        \
            if (g.g == 32 && g.ch == 4) return CALL(1, 32, 4);                        \
        }                                                                             \
    } while (0)

// Sorted positions per group.  Longer segments mean fewer partial rows and shorter fix-up chains (the fix-up is pure
// latency: 47 us at 32, 29 at 64, 17 at 128, 12 at 256 on the Terabyte shape) but fewer, longer warps in the main pass;
// measured optimum on B200: 64 with the row-wise Adagrad math in the loop, 128 for SGD.  Small inputs keep 32 so the
// grid still fills the 148 SMs.  Tunable 3 overrides (32/64/128/256).
static int seg_len_for(int64_t N, int optimizer) {
    const int t = get_tunable(DLRMX_TUNE_BWD_SEG);
    if (t == 32 || t == 64 || t == 128 || t == 256) return t;
    if (N < (int64_t)148 * 24 * 64 * 2) return kSeg;
    return optimizer == DLRMX_OPT_SGD ? 128 : 64;
}

static size_t update_workspace_bytes(int64_t N, int D) {
    const int64_t nseg = ceil_div(N > 0 ? N : 1, kSeg);       // sized for the smallest segment length
    return align_up((size_t)nseg * D * 4, 256) * 2 + align_up((size_t)(D + 4) * 4, 256);   // + the scratch row
}

static int bwd_update_impl(int optimizer, int tab_bf16, int grad_bf16, float* const* tables, float* const* moms, const int64_t* table_row_base,
                           int T, int B, int D, int64_t table_ld, const float* grad, int64_t stride_b, int64_t stride_t,
                           const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows, float lr,
                           float eps, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(!tab_bf16 || optimizer == DLRMX_OPT_SGD || moms != nullptr,
                  "bwd_update_bf16: rwsadagrad needs separate fp32 accumulators (moms)");
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || optimizer == DLRMX_OPT_RWSADAGRAD, "bwd_update: bad optimizer %d", optimizer);
    const int64_t ld = table_ld > 0 ? table_ld : D;
    DLRMX_REQUIRE(ld >= D, "bwd_update: table_ld must be >= D");
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || moms != nullptr || ld > D,
                  "bwd_update: rwsadagrad needs accumulators (moms, or a padded row with table_ld > D)");
    DLRMX_REQUIRE(N >= 0 && N < (1ll << 31) && T >= 0 && B > 0, "bwd_update: bad sizes");
    RowGeom g;
    // bf16 rows keep the fp32 lane geometry here (4 elements = 8 bytes per lane): same summation order, so accumulators are
    // bit-identical to the fp32 path's.  An 8-element geometry was measured (Terabyte shape 0.234 ms against 0.194 ms with this
    // one and 0.172 ms on fp32 rows; C3 shape 0.192 / 0.241 / 0.181): the kernel is latency-bound, fewer positions in flight
    // cost more than the halved row bytes return.
    DLRMX_REQUIRE(row_geometry(D, &g), "bwd_update: unsupported D=%d", D);
    if (g.vec >= 4)
        DLRMX_REQUIRE(stride_b % 4 == 0 && stride_t % 4 == 0 && ((uintptr_t)grad & (grad_bf16 ? 7 : 15)) == 0 && ld % g.vec == 0,
                      "bwd_update: grad must be 16-B aligned (bf16: 8-B) with strides (and table_ld) multiple of the lane width");
    if (N == 0) return DLRMX_OK;
    if (workspace_bytes < update_workspace_bytes(N, D)) {
        set_error("bwd_update: workspace %zu < %zu", workspace_bytes, update_workspace_bytes(N, D));
        return DLRMX_E_WORKSPACE;
    }
    BwdParams P;
    P.optimizer = optimizer; P.tab_bf16 = tab_bf16; P.grad_bf16 = grad_bf16; P.tables = tables; P.moms = moms; P.row_base = table_row_base;
    P.B = B; P.D = D; P.ld = ld; P.grad = grad; P.stride_b = stride_b; P.stride_t = stride_t;
    P.skey = sorted_key; P.sbag = sorted_bag; P.N = N; P.seg_len = seg_len_for(N, optimizer); P.nseg = ceil_div(N, (int64_t)P.seg_len);
    P.lr = lr; P.eps = eps; P.inv_D = (D & (D - 1)) == 0 ? 1.0f / (float)D : 0.f;
    P.head_partial = (float*)workspace;
    P.tail_partial = (float*)((char*)workspace + align_up((size_t)P.nseg * D * 4, 256));
    P.scratch_row = (float*)((char*)workspace + update_workspace_bytes(N, D) - align_up((size_t)(D + 4) * 4, 256));
    P.invalid_key = (uint32_t)total_rows;
    cudaStream_t st = (cudaStream_t)stream;
#define DLRMX_CALL_UPDATE(V, G_, C) launch_update<V, G_, C>(P, st)
    // Row-wise Adagrad on 512-byte rows: 2 positions in flight per warp at 4 CTAs per SM (64 registers, no spills) beat 4
    // positions at 3 CTAs per SM (0.175 vs 0.183 ms on the Terabyte shape: the sum-of-squares butterfly, square root and
    // division lengthen every batch, and more resident warps hide that better than deeper batches); plain SGD prefers 4.
    if (g.vec == 4 && g.g == 32 && g.ch == 1 && optimizer == DLRMX_OPT_RWSADAGRAD)
        return launch_update<4, 32, 1, 2, 4>(P, st);
    // the same holds for 256-byte rows (C3 shape, 1.48 M distinct rows: 0.261 -> 0.224 ms)
    if (g.vec == 4 && g.g == 16 && g.ch == 1 && optimizer == DLRMX_OPT_RWSADAGRAD)
        return launch_update<4, 16, 1, 2, 4>(P, st);
    DLRMX_GEOM_DISPATCH(g, DLRMX_CALL_UPDATE);
    set_error("bwd_update: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}

}  // namespace dlrmx

extern "C" {
size_t dlrmx_bwd_update_workspace_bytes(int64_t N, int D) { return dlrmx::update_workspace_bytes(N, D); }

int dlrmx_bwd_update(int optimizer, float* const* tables, float* const* moms, const int64_t* table_row_base,
                     int T, int B, int D, int64_t table_ld, const float* grad, int64_t stride_b, int64_t stride_t,
                     const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows, float lr,
                     float eps, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    return dlrmx::bwd_update_impl(optimizer, 0, 0, tables, moms, table_row_base, T, B, D, table_ld, grad, stride_b, stride_t,
                                  sorted_key, sorted_bag, N, total_rows, lr, eps, workspace, workspace_bytes, stream);
}

int dlrmx_bwd_update_bf16(int optimizer, void* const* tables_bf16, float* const* moms, const int64_t* table_row_base,
                          int T, int B, int D, int64_t table_ld, const float* grad, int64_t stride_b, int64_t stride_t,
                          const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows, float lr,
                          float eps, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    return dlrmx::bwd_update_impl(optimizer, 1, 0, reinterpret_cast<float* const*>(tables_bf16), moms, table_row_base, T, B, D,
                                  table_ld, grad, stride_b, stride_t, sorted_key, sorted_bag, N, total_rows, lr, eps, workspace,
                                  workspace_bytes, stream);
}

int dlrmx_bwd_update_ex(int optimizer, int tables_are_bf16, int grad_is_bf16, void* const* tables, float* const* moms,
                        const int64_t* table_row_base, int T, int B, int D, int64_t table_ld, const void* grad, int64_t stride_b,
                        int64_t stride_t, const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows,
                        float lr, float eps, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    return dlrmx::bwd_update_impl(optimizer, tables_are_bf16 ? 1 : 0, grad_is_bf16 ? 1 : 0, reinterpret_cast<float* const*>(tables),
                                  moms, table_row_base, T, B, D, table_ld, reinterpret_cast<const float*>(grad), stride_b, stride_t,
                                  sorted_key, sorted_bag, N, total_rows, lr, eps, workspace, workspace_bytes, stream);
}

int dlrmx_bwd_coalesce(int B, int D, const float* grad, int64_t stride_b, int64_t stride_t,
                       const int32_t* sorted_bag, const int32_t* run_off, const int32_t* num_unique, int64_t N,
                       float* gsum, dlrmx_stream_t stream) {
    using namespace dlrmx;
    RowGeom g;
    DLRMX_REQUIRE(row_geometry(D, &g), "bwd_coalesce: unsupported D=%d", D);
    DLRMX_REQUIRE(B > 0 && N >= 0 && N < (1ll << 31), "bwd_coalesce: bad sizes");
    if (g.vec == 4)
        DLRMX_REQUIRE(stride_b % 4 == 0 && stride_t % 4 == 0 && ((uintptr_t)grad & 15) == 0 && ((uintptr_t)gsum & 15) == 0,
                      "bwd_coalesce: buffers must be 16-B aligned with strides multiple of 4 when D%%4==0");
    if (N == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
#define DLRMX_CALL_COAL(V, G_, C) \
    launch_coalesce<V, G_, C>(B, D, grad, stride_b, stride_t, sorted_bag, run_off, num_unique, N, gsum, st)
    DLRMX_GEOM_DISPATCH(g, DLRMX_CALL_COAL);
    set_error("bwd_coalesce: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}

int dlrmx_rows_update(int optimizer, float* table, float* mom, const int64_t* rows, const float* gsum,
                      int64_t U, int D, float lr, float eps, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || optimizer == DLRMX_OPT_RWSADAGRAD, "rows_update: bad optimizer %d", optimizer);
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || mom != nullptr, "rows_update: rwsadagrad needs the accumulator");
    RowGeom g;
    DLRMX_REQUIRE(row_geometry(D, &g), "rows_update: unsupported D=%d", D);
    DLRMX_REQUIRE(U >= 0 && U < (1ll << 31), "rows_update: bad U");
    if (g.vec == 4)
        DLRMX_REQUIRE((((uintptr_t)table | (uintptr_t)gsum) & 15) == 0, "rows_update: table/gsum must be 16-B aligned");
    if (U == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
#define DLRMX_CALL_ROWS(V, G_, C) launch_rows_update<V, G_, C>(optimizer, table, mom, rows, gsum, U, D, lr, eps, st)
    DLRMX_GEOM_DISPATCH(g, DLRMX_CALL_ROWS);
    set_error("rows_update: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}

int dlrmx_dense_rows_update(int optimizer, float* table, float* mom, float* gsum, int64_t rows, int D, int64_t table_ld,
                            float lr, float eps, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || optimizer == DLRMX_OPT_RWSADAGRAD, "dense_rows_update: bad optimizer %d", optimizer);
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || mom != nullptr, "dense_rows_update: rwsadagrad needs the accumulator");
    RowGeom g;
    DLRMX_REQUIRE(row_geometry(D, &g), "dense_rows_update: unsupported D=%d", D);
    const int64_t ld = table_ld > 0 ? table_ld : D;
    DLRMX_REQUIRE(rows >= 0 && ld >= D, "dense_rows_update: bad sizes");
    if (g.vec == 4)
        DLRMX_REQUIRE((((uintptr_t)table | (uintptr_t)gsum) & 15) == 0 && ld % 4 == 0,
                      "dense_rows_update: table/gsum must be 16-B aligned, table_ld a multiple of 4");
    if (rows == 0) return DLRMX_OK;
    cudaStream_t st = (cudaStream_t)stream;
#define DLRMX_CALL_DENSE(V, G_, C) launch_dense_rows_update<V, G_, C>(optimizer, table, mom, gsum, rows, D, ld, lr, eps, st)
    DLRMX_GEOM_DISPATCH(g, DLRMX_CALL_DENSE);
    set_error("dense_rows_update: no kernel for D=%d", D);
    return DLRMX_E_ARG;
}
}
