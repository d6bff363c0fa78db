// Kernels + C ABI of the fused backward/update (design notes in embbag_bwd.cuh).
#include "embbag_bwd.cuh"

namespace dlrmx {

template <int VEC, int G, int CH, int K>
__global__ void __launch_bounds__(256, 2) bwd_update_kernel(const BwdParams P) {
    const int gl = threadIdx.x & (G - 1);
    const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (seg >= P.nseg) return;
    const int64_t p0 = seg * kSeg, p1 = min(P.N, p0 + kSeg);
    const int col0 = gl * VEC;
    const uint32_t* __restrict__ skey = P.skey;
    const int32_t* __restrict__ sbag = P.sbag;

    bool owned = !(p0 > 0 && skey[p0 - 1] == skey[p0]);  // does the first run start in this segment?
    RowVec<VEC> acc[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c].zero();

    for (int64_t pb = p0; pb < p1; pb += K) {
        uint32_t kk[K + 1];
        kk[0] = skey[pb];
#pragma unroll
        for (int j = 1; j <= K; ++j) kk[j] = (pb + j < P.N) ? skey[pb + j] : ~kk[j - 1];

        RowVec<VEC> g[K][CH], w[K][CH];
        float* wp[K];
        float* mp[K];
        float mo[K];
        bool valid[K], last[K];
#pragma unroll
        for (int j = 0; j < K; ++j) {
            valid[j] = pb + j < p1;
            last[j] = valid[j] && kk[j + 1] != kk[j];
            wp[j] = nullptr;
            mp[j] = nullptr;
            mo[j] = 0.f;
            if (valid[j]) {
                const int bag = sbag[pb + j];
                const int t = bag / P.B;
                const int b = bag - t * P.B;
                const float* gp = P.grad + (int64_t)b * P.stride_b + (int64_t)t * P.stride_t + col0;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (CH == 1 || col0 + c * G * VEC < P.D) g[j][c].load_ro(gp + c * G * VEC);
                if (last[j]) {
                    const int64_t row = (int64_t)kk[j] - P.row_base[t];
                    wp[j] = P.tables[t] + row * P.D + col0;
#pragma unroll
                    for (int c = 0; c < CH; ++c)
                        if (CH == 1 || col0 + c * G * VEC < P.D) w[j][c].load(wp[j] + c * G * VEC);
                    if (P.optimizer == DLRMX_OPT_RWSADAGRAD) {
                        mp[j] = P.moms[t] + row;
                        mo[j] = *mp[j];
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if (valid[j]) {
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (CH == 1 || col0 + c * G * VEC < P.D) acc[c].add(g[j][c]);
                if (last[j]) {
                    if (owned) apply_row<VEC, G, CH>(P, wp[j], mp[j], mo[j], w[j], acc, col0);
                    else store_partial<VEC, G, CH>(P.head_partial, seg, P.D, col0, acc);
#pragma unroll
                    for (int c = 0; c < CH; ++c) acc[c].zero();
                    owned = true;
                } else if (pb + j == p1 - 1) {  // the segment ends inside a run
                    store_partial<VEC, G, CH>(owned ? P.tail_partial : P.head_partial, seg, P.D, col0, acc);
                }
            }
        }
    }
}

// Finishes runs that cross segment boundaries; one group per segment, most exit immediately.
template <int VEC, int G, int CH>
__global__ void __launch_bounds__(256) bwd_fixup_kernel(const BwdParams P) {
    constexpr int KF = 8;
    const int gl = threadIdx.x & (G - 1);
    const int64_t seg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (seg >= P.nseg) return;
    const int64_t p0 = seg * kSeg, p1 = min(P.N, p0 + kSeg);
    const uint32_t* __restrict__ skey = P.skey;
    if (p1 >= P.N) return;
    const uint32_t key = skey[p1 - 1];
    if (skey[p1] != key) return;                                     // no run crosses my end
    if (skey[p0] == key && p0 > 0 && skey[p0 - 1] == key) return;     // I am a middle segment
    const int col0 = gl * VEC;

    RowVec<VEC> acc[CH];
    {
        const float* tp = P.tail_partial + seg * P.D + col0;
#pragma unroll
        for (int c = 0; c < CH; ++c)
            if (CH == 1 || col0 + c * G * VEC < P.D) acc[c].load(tp + c * G * VEC);
    }
    bool done = false;
    for (int64_t s = seg + 1; !done; s += KF) {
        RowVec<VEC> h[KF][CH];
        bool cont[KF];
#pragma unroll
        for (int j = 0; j < KF; ++j) {
            const int64_t sj = s + j;
            cont[j] = false;
            if (sj < P.nseg) {
                const float* hp = P.head_partial + sj * P.D + col0;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (CH == 1 || col0 + c * G * VEC < P.D) h[j][c].load(hp + c * G * VEC);
                const int64_t e = min(P.N, (sj + 1) * kSeg);
                cont[j] = e < P.N && skey[e - 1] == key && skey[e] == key;
            }
        }
#pragma unroll
        for (int j = 0; j < KF; ++j) {
            if (!done) {
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (CH == 1 || col0 + c * G * VEC < P.D) acc[c].add(h[j][c]);
                if (!cont[j]) done = true;
            }
        }
    }
    const int bag = P.sbag[p1 - 1];
    const int t = bag / P.B;
    const int64_t row = (int64_t)key - P.row_base[t];
    float* wp = P.tables[t] + row * P.D + col0;
    RowVec<VEC> w[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (CH == 1 || col0 + c * G * VEC < P.D) w[c].load(wp + c * G * VEC);
    float* mp = nullptr;
    float mo = 0.f;
    if (P.optimizer == DLRMX_OPT_RWSADAGRAD) {
        mp = P.moms[t] + row;
        mo = *mp;
    }
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
                    if (CH == 1 || col0 + c * G * VEC < D) g[j][c].load_ro(gp + c * G * VEC);
            }
#pragma unroll
        for (int j = 0; j < K; ++j)
            if (pb + j < p1) {
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (CH == 1 || col0 + c * G * VEC < D) acc[c].add(g[j][c]);
            }
    }
    float* o = gsum + u * D + col0;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (CH == 1 || col0 + c * G * VEC < D) acc[c].store(o + c * G * VEC);
}


// Non-fused optimizer on an already coalesced gradient (unique rows + summed gradient rows): the GPU
// form of torch.optim.SGD / RWSAdagrad.step's sparse branch (optim/rwsadagrad.py:97-113) for ONE table.
// The reference refuses rwsadagrad on GPU (dlrm_s_pytorch.py:1636-1637); this is what it would run.
template <int VEC, int G, int CH>
__global__ void __launch_bounds__(256)
rows_update_kernel(int optimizer, float* __restrict__ W, float* __restrict__ mom,
                   const int64_t* __restrict__ rows, const float* __restrict__ gsum, int64_t U, int D,
                   float lr, float eps) {
    const int gl = threadIdx.x & (G - 1);
    const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (u >= U) return;
    const int col0 = gl * VEC;
    const int64_t row = rows[u];
    BwdParams P;
    P.optimizer = optimizer; P.D = D; P.lr = lr; P.eps = eps;
    RowVec<VEC> g[CH], w[CH];
    float* wp = W + row * D + col0;
    const float* gp = gsum + u * D + col0;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        if (CH == 1 || col0 + c * G * VEC < D) {
            g[c].load_ro(gp + c * G * VEC);
            w[c].load(wp + c * G * VEC);
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

template <int VEC, int G, int CH>
static int launch_rows_update(int optimizer, float* W, float* mom, const int64_t* rows, const float* gsum,
                              int64_t U, int D, float lr, float eps, cudaStream_t st) {
    const unsigned blocks = (unsigned)ceil_div(U * G, 256);
    rows_update_kernel<VEC, G, CH><<<blocks, 256, 0, st>>>(optimizer, W, mom, rows, gsum, U, D, lr, eps);
    return check_launch("rows_update");
}

template <int VEC, int G, int CH>
static int launch_update(const BwdParams& P, cudaStream_t st) {
    const int threads = 256;
    const unsigned blocks = (unsigned)ceil_div(P.nseg * G, threads);
    if (get_tunable(DLRMX_TUNE_BWD_K) == 8 && CH == 1)
        bwd_update_kernel<VEC, G, CH, 8><<<blocks, threads, 0, st>>>(P);
    else
        bwd_update_kernel<VEC, G, CH, (CH == 1 ? 4 : 2)><<<blocks, threads, 0, st>>>(P);
    int rc = check_launch("bwd_update");
    if (rc) return rc;
    bwd_fixup_kernel<VEC, G, CH><<<blocks, threads, 0, st>>>(P);
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

static size_t update_workspace_bytes(int64_t N, int D) {
    const int64_t nseg = ceil_div(N > 0 ? N : 1, kSeg);
    return align_up((size_t)nseg * D * 4, 256) * 2;
}

}  // namespace dlrmx

extern "C" {
size_t dlrmx_bwd_update_workspace_bytes(int64_t N, int D) { return dlrmx::update_workspace_bytes(N, D); }

int dlrmx_bwd_update(int optimizer, float* const* tables, float* const* moms, const int64_t* table_row_base,
                     int T, int B, int D, const float* grad, int64_t stride_b, int64_t stride_t,
                     const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, float lr, float eps,
                     void* workspace, size_t workspace_bytes, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || optimizer == DLRMX_OPT_RWSADAGRAD, "bwd_update: bad optimizer %d", optimizer);
    DLRMX_REQUIRE(optimizer == DLRMX_OPT_SGD || moms != nullptr, "bwd_update: rwsadagrad needs accumulators");
    DLRMX_REQUIRE(N >= 0 && N < (1ll << 31) && T >= 0 && B > 0, "bwd_update: bad sizes");
    RowGeom g;
    DLRMX_REQUIRE(row_geometry(D, &g), "bwd_update: unsupported D=%d", D);
    if (g.vec == 4)
        DLRMX_REQUIRE(stride_b % 4 == 0 && stride_t % 4 == 0 && ((uintptr_t)grad & 15) == 0,
                      "bwd_update: grad must be 16-B aligned with strides multiple of 4 when D%%4==0");
    if (N == 0) return DLRMX_OK;
    if (workspace_bytes < update_workspace_bytes(N, D)) {
        set_error("bwd_update: workspace %zu < %zu", workspace_bytes, update_workspace_bytes(N, D));
        return DLRMX_E_WORKSPACE;
    }
    BwdParams P;
    P.optimizer = optimizer; P.tables = tables; P.moms = moms; P.row_base = table_row_base;
    P.B = B; P.D = D; P.grad = grad; P.stride_b = stride_b; P.stride_t = stride_t;
    P.skey = sorted_key; P.sbag = sorted_bag; P.N = N; P.nseg = ceil_div(N, kSeg);
    P.lr = lr; P.eps = eps;
    P.head_partial = (float*)workspace;
    P.tail_partial = (float*)((char*)workspace + align_up((size_t)P.nseg * D * 4, 256));
    cudaStream_t st = (cudaStream_t)stream;
#define DLRMX_CALL_UPDATE(V, G_, C) launch_update<V, G_, C>(P, st)
    DLRMX_GEOM_DISPATCH(g, DLRMX_CALL_UPDATE);
    set_error("bwd_update: no kernel for D=%d", D);
    return DLRMX_E_ARG;
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
}
