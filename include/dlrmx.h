/*
 * dlrmx.h -- C ABI of libdlrmx.so, the B200 (sm_100a) sparse hot path for DLRM.
 *
 * This is the drop-in boundary.  The reference's only native interface for this
 * path is the torch op of yx_modfs:
 *     TORCH_LIBRARY(batched_forward, m) { m.def("forward", &batched_embedding_forward_cuda); }
 *         -- yx_modfs/batched_forward.cpp:4-13, kernel yx_modfs/table_batched_embeddings_cuda_yx.cu:251-389
 * and everything else on the path is reached through Python call sites of dlrm_s_pytorch.py /
 * optim/rwsadagrad.py / extend_distributed.py.  Every entry point below names the reference
 * interface it replaces.  The Python host layer (dlrm_yx_b200/*.py) binds these with ctypes and
 * re-creates the reference-facing names (torch.ops.batched_forward.forward, nn.EmbeddingBag-like
 * module, TableBatchedEmbeddingBags, interact_features, RWSAdagrad, alltoall).
 *
 * Conventions
 *   - plain pointers and sizes only; all data pointers are DEVICE pointers unless named host_*;
 *   - `stream` is a cudaStream_t passed as void*; every call is stream-ordered and non-blocking,
 *     never switches device, never allocates (workspace sizes are queried and passed in);
 *   - return 0 on success, a negative DLRMX_E_* code on failure; dlrmx_last_error() gives the text;
 *   - "bag" (t,b) = lookup b of table t, bag id = t*B + b (table-major, the packed layout of
 *     dlrm_data_pytorch.py:748-753);  offsets[bag] is the first index position of the bag, its end
 *     is offsets[bag+1] for bag+1 < T*B and N (the index count) for the last bag -- which is both the
 *     nn.EmbeddingBag rule (dlrm_s_pytorch.py:571-577) and the T*B+1 packed rule;
 *   - pooled rows / their gradients are addressed as  base + b*stride_b + t*stride_t  (in floats),
 *     so the same kernels write [B,T,D], [T,B,D] or straight into the [B,1+T,D] interaction input.
 */
#ifndef DLRMX_H_
#define DLRMX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLRMX_VERSION 100

#define DLRMX_OK 0
#define DLRMX_E_ARG (-1)      /* bad argument (shape / alignment / unsupported D) */
#define DLRMX_E_CUDA (-2)     /* CUDA runtime error at launch */
#define DLRMX_E_WORKSPACE (-3)/* workspace too small */

typedef void* dlrmx_stream_t;

/* Sparse optimizers fused into the backward (dlrm_s_pytorch.py:1639-1666). */
#define DLRMX_OPT_SGD 0        /* torch.optim.SGD on sparse grads: W[r] += -lr * g[r]            */
#define DLRMX_OPT_RWSADAGRAD 1 /* optim/rwsadagrad.py:95-113: row-wise accumulator                */

int dlrmx_version(void);
const char* dlrmx_last_error(void);
/* Number of kernels launched by this library since load (bench.py's gpu_launches claim). */
int64_t dlrmx_launch_count(void);
/* Test hooks (0 = default) that select between code paths the library ships anyway, so the tests can compare them bit
 * for bit: 2 = force the CUDA-core interaction kernels, 3 = sorted positions per group of the fused update
 * (32/64/128/256), 10 = per-thread cp.async staging instead of TMA bulk copies in the interaction kernels, 11 = CTAs per
 * SM of their persistent grid.  Not part of the reference-facing surface. */
int dlrmx_set_tunable(int which, int value);

/* ---------------------------------------------------------------------------------------------
 * Forward: table-batched embedding-bag sum pooling.
 * Replaces nn.EmbeddingBag(mode="sum").forward as called by DLRM_Net.apply_emb
 * (dlrm_s_pytorch.py:526-587), TableBatchedEmbeddingBags.forward (dlrm_s_pytorch.py:589-591) and
 * batched_embedding_forward_cuda (yx_modfs/table_batched_embeddings_cuda_yx.cu:317-389).
 *   tables      [T] device array of row-major fp32 table base pointers ([rows_t, D] each, 16-B aligned
 *               when D % 4 == 0)
 *   table_rows  [T] device int64 row counts, or NULL to skip index validation
 *   table_ld    row pitch of the tables in floats (0 = D).  A pitch > D leaves padding after every row; the
 *               fused update can keep the row-wise Adagrad accumulator there (see dlrmx_bwd_update)
 *   indices     [N] row ids local to their table; offsets [>= T*B] bag starts (see conventions)
 *   psw         [N] per-sample weights (dlrm_s_pytorch.py:544-547) or NULL
 *   out         pooled rows, fp32, written for every bag (empty bag -> zeros)
 *   err_flag    device int, or NULL; bit 0 is set if an index is outside [0, rows_t)
 * Sum order: sequential in index order, plain fp32 adds (bit-exact with ATen's EmbeddingBag).
 * ------------------------------------------------------------------------------------------- */
int dlrmx_embbag_fwd_i32(const float* const* tables, const int64_t* table_rows,
                         const int32_t* indices, const int32_t* offsets, int64_t N,
                         int T, int B, int D, int64_t table_ld, const float* psw,
                         float* out, int64_t stride_b, int64_t stride_t,
                         int* err_flag, dlrmx_stream_t stream);
int dlrmx_embbag_fwd_i64(const float* const* tables, const int64_t* table_rows,
                         const int64_t* indices, const int64_t* offsets, int64_t N,
                         int T, int B, int D, int64_t table_ld, const float* psw,
                         float* out, int64_t stride_b, int64_t stride_t,
                         int* err_flag, dlrmx_stream_t stream);

/* The same lookup over bf16 TABLE ROWS (fp16/bf16 row storage of the table-batched module, the `weights_precision` knob of
 * the reference's fbgemm slot, dlrm_s_pytorch.py:350-366): tables_bf16 [T] device array of row-major bf16 table base
 * pointers (8-byte aligned when D % 4 == 0), table_ld in ELEMENTS.  Rows are widened exactly to fp32 when they are read and
 * summed in fp32 in index order; pooled rows, per-sample weights and everything downstream stay fp32.  Half the table bytes
 * of the fp32 entry points, bit-identical to them on a table holding the widened values. */
int dlrmx_embbag_fwd_bf16_i32(const void* const* tables_bf16, const int64_t* table_rows, const int32_t* indices,
                              const int32_t* offsets, int64_t N, int T, int B, int D, int64_t table_ld, const float* psw,
                              float* out, int64_t stride_b, int64_t stride_t, int* err_flag, dlrmx_stream_t stream);
int dlrmx_embbag_fwd_bf16_i64(const void* const* tables_bf16, const int64_t* table_rows, const int64_t* indices,
                              const int64_t* offsets, int64_t N, int T, int B, int D, int64_t table_ld, const float* psw,
                              float* out, int64_t stride_b, int64_t stride_t, int* err_flag, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Backward, step 1: sort the looked-up rows.
 * Replaces what EmbeddingBag backward + grad.coalesce() do implicitly (sparse COO gradient with
 * one entry per looked-up index, then sort + merge: optim/rwsadagrad.py:98); no [N,D] gradient
 * expansion is ever materialised.
 *   key[p]  = table_row_base[t(p)] + indices[p]   (global row id, uint32: sum of rows < 2^32)
 *   bag[p]  = t(p)*B + b(p)
 * sorted by key with an LSD radix sort (8-bit digits, only ceil(log2(total_rows)/8) passes), STABLE
 * in p -- identical to torch.sort(stable=True) on the keys.
 *   table_row_base  [T+1] device int64 exclusive prefix of the table row counts
 *   total_rows      host copy of table_row_base[T] (< 2^32 - 1)
 *   sorted_key [N] uint32, sorted_bag [N] int32: outputs
 *   err_flag        device int, or NULL; bit 1 is set if an index is outside [0, rows_t).  Such a lookup gets the key
 *                   `total_rows` (one past the last valid row): it sorts to the very end and dlrmx_bwd_update skips it, so a
 *                   bad index can neither alias another table's row nor write out of bounds (nn.EmbeddingBag raises /
 *                   device-asserts on the same input; the host layer raises when it reads the flag)
 * ------------------------------------------------------------------------------------------- */
size_t dlrmx_bwd_sort_workspace_bytes(int64_t N, int64_t total_bags);
int dlrmx_bwd_sort_i32(const int32_t* indices, const int32_t* offsets, int64_t N, int T, int B,
                       const int64_t* table_row_base, int64_t total_rows,
                       uint32_t* sorted_key, int32_t* sorted_bag, int* err_flag,
                       void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);
int dlrmx_bwd_sort_i64(const int64_t* indices, const int64_t* offsets, int64_t N, int T, int B,
                       const int64_t* table_row_base, int64_t total_rows,
                       uint32_t* sorted_key, int32_t* sorted_bag, int* err_flag,
                       void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);

/* Backward, step 2 (only needed by the non-fused "parity seam" and the tests): run-length
 * structure of the sorted keys.
 *   unique_key [N] (first num_unique valid), run_off [N+1] (first num_unique+1 valid; run u covers
 *   sorted positions [run_off[u], run_off[u+1])), num_unique: device int32[1].
 * These integer arrays are the bit-exact contract against sort/unique_consecutive. */
size_t dlrmx_bwd_segments_workspace_bytes(int64_t N);
int dlrmx_bwd_segments(const uint32_t* sorted_key, int64_t N,
                       uint32_t* unique_key, int32_t* run_off, int32_t* num_unique,
                       void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Backward, step 3 (fused): segmented reduce of the pooled-row gradients over each run of equal
 * keys + the sparse optimizer update of that table row, one pass.
 * Replaces EmbeddingBag.backward + torch.optim.SGD.step (dlrm_s_pytorch.py:1640, 1933) and
 * RWSAdagrad.step's sparse branch (optim/rwsadagrad.py:97-113).
 *   grad        gradient of the pooled rows, addressed like `out` of the forward
 *   psw         per-sample weights [N] indexed by ORIGINAL position -- not supported when fused: NULL
 *   moms        [T] device array of row-wise accumulator pointers ([rows_t] fp32), RWSADAGRAD only; NULL with
 *               table_ld > D: the accumulator of a row is the float right after it (column D of the padded
 *               row), i.e. in the DRAM page the row update opens anyway -- the separate 4-byte accumulator
 *               arrays of optim/rwsadagrad.py:72-78 cost a DRAM access per row as expensive as the row itself
 *   table_ld    row pitch of the tables in floats (0 = D)
 *   total_rows  as given to dlrmx_bwd_sort: runs with that key (out-of-range lookups) update nothing
 *   lr          SGD: lr.  RWSADAGRAD: clr = lr / (1 + (step-1)*lr_decay), computed by the caller
 * Deterministic: each run is summed in sorted (stable) order; runs longer than one 32-position
 * segment are reduced per segment and the segment partials added in ascending order.
 * ------------------------------------------------------------------------------------------- */
size_t dlrmx_bwd_update_workspace_bytes(int64_t N, int D);
int dlrmx_bwd_update(int optimizer, float* const* tables, float* const* moms,
                     const int64_t* table_row_base, int T, int B, int D, int64_t table_ld,
                     const float* grad, int64_t stride_b, int64_t stride_t,
                     const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows,
                     float lr, float eps,
                     void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);
/* The fused update on bf16 table rows: every touched row is widened to fp32, updated in fp32 exactly as above and rounded to
 * nearest even when it is written back (stochastic_rounding=False of the reference's call, dlrm_s_pytorch.py:333); the
 * row-wise Adagrad accumulators stay fp32 in `moms` (required).  The result equals rounding the fp32 entry point's result
 * on a table holding the widened rows. */
int dlrmx_bwd_update_bf16(int optimizer, void* const* tables_bf16, float* const* moms,
                          const int64_t* table_row_base, int T, int B, int D, int64_t table_ld,
                          const float* grad, int64_t stride_b, int64_t stride_t,
                          const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows,
                          float lr, float eps,
                          void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);

/* The same kernel with the row types as flags: tables fp32 / bf16 (as dlrmx_bwd_update_bf16) and, when grad_is_bf16, pooled-row
 * gradients stored as bf16 (strides in elements; 8-byte aligned rows): what dlrmx_a2a_push_grad_bf16 delivers.  A bf16
 * gradient is widened exactly, so the result equals dlrmx_bwd_update's on an fp32 buffer holding the same values. */
int dlrmx_bwd_update_ex(int optimizer, int tables_are_bf16, int grad_is_bf16, void* const* tables, float* const* moms,
                        const int64_t* table_row_base, int T, int B, int D, int64_t table_ld,
                        const void* grad, int64_t stride_b, int64_t stride_t,
                        const uint32_t* sorted_key, const int32_t* sorted_bag, int64_t N, int64_t total_rows,
                        float lr, float eps,
                        void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);

/* Non-fused variant: emits the coalesced gradient rows  gsum[u,:] = sum over run u  ([num_unique, D],
 * row stride D) so that the reference's own optimizers can consume a coalesced sparse COO gradient
 * ("parity seam").  Needs run_off/num_unique from dlrmx_bwd_segments. */
int dlrmx_bwd_coalesce(int B, int D, const float* grad, int64_t stride_b, int64_t stride_t,
                       const int32_t* sorted_bag, const int32_t* run_off,
                       const int32_t* num_unique, int64_t N, float* gsum, dlrmx_stream_t stream);

/* Non-fused sparse optimizer step of ONE table on an already coalesced gradient (the GPU form of
 * torch.optim.SGD on a sparse grad, dlrm_s_pytorch.py:1640/1933, and of RWSAdagrad.step's sparse
 * branch, optim/rwsadagrad.py:97-113 -- which the reference refuses to run on GPU,
 * dlrm_s_pytorch.py:1636-1637).  rows [U] int64 unique row ids, gsum [U, D] summed gradient rows. */
int dlrmx_rows_update(int optimizer, float* table, float* mom, const int64_t* rows, const float* gsum,
                      int64_t U, int D, float lr, float eps, dlrmx_stream_t stream);

/* Replicated (data-parallel) tables -- a sharding policy next to the reference's table-wise one (sharders.py:23-60,
 * dlrm_s_pytorch.py:444-468): tables with fewer rows than the batch are cheaper to keep on every rank than to exchange.
 * Each rank accumulates its LOCAL batch's pooled-row gradients per row into a zeroed dense buffer gsum [rows, D]
 * (dlrmx_bwd_update with DLRMX_OPT_SGD, lr = -1 and gsum as the "table": w + 1*g on w = 0 is exact), the buffers are
 * summed across ranks (all-reduce), and this call applies the sparse optimizer step of optim/rwsadagrad.py:97-113 /
 * torch.optim.SGD to every row whose summed gradient is non-zero -- untouched rows keep their bits, as in the reference --
 * and clears those rows of gsum for the next step.  table [rows, D] with row pitch table_ld (0 = D), mom [rows]. */
int dlrmx_dense_rows_update(int optimizer, float* table, float* mom, float* gsum, int64_t rows, int D,
                            int64_t table_ld, float lr, float eps, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Dot / cat feature interaction.  Replaces DLRM_Net.interact_features (dlrm_s_pytorch.py:627-673):
 *   feat   [B, F, D] fp32 (feature 0 = bottom-MLP output x, features 1.. = pooled embeddings), row
 *          stride feat_stride_b floats -- the forward kernel above writes pooled rows straight into it;
 *   out    [B, out_stride] fp32:  out[b, 0:D] = x[b];  out[b, D + k] = <feat[b,li[k]], feat[b,lj[k]]>
 *          with (li,lj) = tril_indices(F, F, itself ? 0 : -1) in row-major order (dlrm_s_pytorch.py:646-656).
 * Backward: dfeat[b,i,:] = sum_j (dZ[i,j] + dZ[j,i]) feat[b,j,:]  (+ dout[b,0:D] on feature 0).
 * ------------------------------------------------------------------------------------------- */
int dlrmx_interact_dot_fwd(const float* feat, int64_t feat_stride_b, int B, int F, int D, int itself,
                           float* out, int64_t out_stride, dlrmx_stream_t stream);
int dlrmx_interact_dot_bwd(const float* feat, int64_t feat_stride_b, int B, int F, int D, int itself,
                           const float* dout, int64_t dout_stride,
                           float* dfeat, int64_t dfeat_stride_b, dlrmx_stream_t stream);

/* bf16-facing variants for towers that run in bf16 (the same interact_features call, dlrm_s_pytorch.py:627-673, between
 * a bf16 bottom MLP and a bf16 top MLP):
 *   feature 0 is NOT read from feat but from x_bf16 [B, D] (the bottom-MLP output as the GEMM epilogue wrote it);
 *   out_bf16 [B, out_stride] bf16 = [x | tril], columns D+nz .. out_stride-1 are written as zeros (out_stride = the top
 *   MLP's K rounded up to 8, so the GEMM reads it without a pad/cast pass);
 *   backward reads dout_bf16 [B, dout_stride] bf16 (dout_stride % 8 == 0), writes the pooled-embedding gradients
 *   dfeat[:, 1:, :] in fp32 (feature 0 of dfeat is left untouched) and the gradient of x as dx_bf16 [B, D]; with relu_x != 0
 *   dx is also gated by (x > 0): x is then the output of a Linear + ReLU layer (the bottom tower's last) and dx leaves as
 *   that layer's pre-activation gradient, no separate ReLU-backward pass.
 *   feat_stride_f = row pitch of the features inside a sample, in floats: D (contiguous [F, D] blocks; 0 means D), or
 *   D + 8 when D % 32 == 0 -- the interaction input then has the row pitch of the kernels' shared-memory tile (the
 *   lookup writes pooled rows with stride_t = D + 8: rows stay 32-byte aligned, so its stores stay full sectors), and a
 *   sample is staged by ONE TMA bulk copy (cp.async.bulk + mbarrier) instead of F*D/4 per-thread 16-byte cp.async copies.
 * Products are single-pass TF32 (operands rounded to nearest) with fp32 accumulation.  Needs 2 <= F <= 32, D % 8 == 0,
 * D <= 512; other shapes: use the fp32 entry points and cast. */
int dlrmx_interact_dot_fwd_mixed(const float* feat, int64_t feat_stride_b, int64_t feat_stride_f, const void* x_bf16, int B, int F, int D,
                                 int itself, void* out_bf16, int64_t out_stride, dlrmx_stream_t stream);
int dlrmx_interact_dot_bwd_mixed(const float* feat, int64_t feat_stride_b, int64_t feat_stride_f, const void* x_bf16, int B, int F, int D,
                                 int itself, const void* dout_bf16, int64_t dout_stride, float* dfeat,
                                 int64_t dfeat_stride_b, void* dx_bf16, int relu_x, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * MLP towers (create_mlp / apply_mlp, dlrm_s_pytorch.py:227-265, 518-524): fused passes around the
 * tensor-core GEMMs, and the dense half of the optimizers.
 * relu_bwd_bias_grad: dz = dy * (y > 0) (bf16 [B,N], skipped when relu == 0) and per-row-chunk column
 *   sums of dz into db_partial [chunks, N] fp32 (the bias gradient is their sum over chunks).
 * dense_opt: ONE launch over a flat parameter buffer: mode 0 = torch.optim.SGD (dlrm_s_pytorch.py:1640)
 *   p -= lr*g ; mode 1 = Adagrad, the dense branch of RWSAdagrad.step (optim/rwsadagrad.py:115-118)
 *   db (optional) [N] receives the column sums themselves: the CTA finishing last per 256-column block adds the
 *   chunk partials in a fixed order (deterministic); counters = ceil(N/256) uint32, zero before the first call, left
 *   zero by every call.
 *   acc += g*g ; p -= lr * g / (sqrt(acc) + eps).  g is scaled by grad_scale first and cleared afterwards;
 *   shadow_bf16 (optional) receives the bf16 copy of p that the GEMMs read.
 * ------------------------------------------------------------------------------------------- */
int dlrmx_relu_bwd_bias_grad(const void* dy_bf16, const void* y_bf16, void* dz_bf16, float* db_partial, float* db,
                             unsigned int* counters, int B, int N, int chunks, int relu, dlrmx_stream_t stream);
int dlrmx_dense_opt(int mode, float* params, float* acc, float* grads, void* shadow_bf16, int64_t n, float lr,
                    float eps, float grad_scale, dlrmx_stream_t stream);

/* relu_bwd_bias_grad_noalloc: NO COUNTERPART IN THE REFERENCE -- added for callers
 * that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The same
 * kernel and grid as dlrmx_relu_bwd_bias_grad with the default chunking (results bit-identical), but the scratch (chunk
 * partials + counters) is a static 304 KB __device__ array inside the library: the call allocates nothing and the caller
 * passes only inputs and outputs.  dz [B, N] bf16 = dy * (y > 0) (relu != 0; with relu == 0 dz and y are ignored) and
 * db [N] fp32 = column sums of dz (of dy when relu == 0).  Limit: N <= 2048 (N % 8 == 0); larger N returns DLRMX_E_ARG.
 * Every call on a device shares the one array: calls that overlap on the same GPU (different streams) race on it. */
int dlrmx_relu_bwd_bias_grad_noalloc(const void* dy_bf16, const void* y_bf16, void* dz_bf16, float* db, int B, int N,
                                     int relu, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Tower GEMMs on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands staged by TMA), bf16 in /
 * fp32 accumulate.  Replaces the cuBLAS GEMMs of nn.Linear in create_mlp / apply_mlp (dlrm_s_pytorch.py:227-265, 518-524)
 * and of their autograd backward.
 *
 * gemm_bf16:  C[M, N] (bf16, row pitch ldc) = epilogue( A[M, K] . op(B) ), A bf16 row-major (pitch lda)
 *   b_is_kn = 0: B is [N, K] row-major  (forward:  y = x W^t with W [N, K] as nn.Linear stores it)
 *   b_is_kn = 1: B is [K, N] row-major  (dgrad:    dx = dz W with the same W, no transposed copy)
 *   epilogue 0: none;  1: + bias[N] (fp32), ReLU -- with relu_bits_out the gates (Y > 0) are also written as one bit per
 *   element, word [m, n >> 5] bit (n & 31), row pitch ld_mask words;  3: + bias;
 *   4: C = acc where the gate bit of relu_mask (such a bit matrix, pitch ld_mask words) is set, else 0 -- the ReLU backward
 *   of the layer that produced this GEMM's output gradient, so the result is that layer's dz.
 *   All run on CTA pairs (tcgen05.mma.cta_group::2, M = 256 per MMA, TMA-store epilogue).
 *   N % 8 == 0, K % 8 == 0, all pointers 16-byte aligned, pitches multiples of 8 elements.
 * gemm_wgrad_bf16: for every layer l of a tower, in ONE launch:  dW[l] [N_l, K_l] (fp32, pitch ld_dw) = dZ[l]^t X[l]  and
 *   db[l] [N_l] = column sums of dZ[l]  (dZ[l] [M, N_l], X[l] [M, K_l] bf16 row-major).  The batch range of every
 *   256 x 256 tile of every layer is cut into chunks that go to the CTA pairs chunk-major (split-K); the chunks leave fp32
 *   partials in the workspace that ONE second launch sums in chunk order, so the result is deterministic.  workspace: the size the query
 *   returns (contents need no initialisation).  Host arrays of device pointers / sizes.
 * ------------------------------------------------------------------------------------------- */
int dlrmx_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int b_is_kn, void* C, int64_t ldc, int M, int N,
                    int K, int epilogue, const float* bias, const void* relu_mask, int64_t ld_mask, uint32_t* relu_bits_out,
                    dlrmx_stream_t stream);
size_t dlrmx_gemm_wgrad_workspace_bytes(int n_layers, const int* N, const int* K, int M);
int dlrmx_gemm_wgrad_bf16(int n_layers, const void* const* dZ, const int64_t* ld_dz, const void* const* X,
                          const int64_t* ld_x, float* const* dW, const int64_t* ld_dw, float* const* db, const int* N,
                          const int* K, int M, void* workspace, size_t workspace_bytes, dlrmx_stream_t stream);

/* Head of the top tower: Linear(K -> 1) + Sigmoid (last layer of create_mlp with sigmoid_layer = last,
 * dlrm_s_pytorch.py:227-265) on a bf16 activation h [B, K] with fp32 master weights w [K], bias [1].
 *   head_fwd: p[b] = sigmoid(<h[b,:], w> + bias)            (sigmoid = 0: no activation)
 *   head_bwd: dz = dp * p * (1 - p);  dh_bf16[b,:] = dz * w;  gw[k] = sum_b dz h[b,k];  gb[0] = sum_b dz.
 *             gb_prev_relu != NULL: h is the output of a Linear + ReLU layer whose backward is fused in -- dh leaves as
 *             that layer's dz = dh * (h > 0) and gb_prev_relu [K] receives its bias gradient (no relu_bwd_bias_grad pass).
 *             workspace: dlrmx_head_bwd_workspace_bytes(K, chunks); counters: ceil(K/256) uint32, zero before the
 *             first call, left zero.  Sums are finished in a fixed order (deterministic). */
int dlrmx_head_fwd(const void* h_bf16, const float* w, const float* bias, float* p, int B, int K, int sigmoid,
                   dlrmx_stream_t stream);
size_t dlrmx_head_bwd_workspace_bytes(int K, int chunks);
int dlrmx_head_bwd(const float* dp, const float* p, const void* h_bf16, const float* w, void* dh_bf16, float* gw,
                   float* gb, float* workspace, unsigned int* counters, int B, int K, int chunks, int sigmoid,
                   float* gb_prev_relu, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Model-parallel exchange helpers.  Replaces the cat / split / contiguous copies around
 * dist.all_to_all_single in extend_distributed.py:405-508 (All2All_Req / All2All_Wait).
 * a2a_unpack: recv buffer = concat over source ranks p of [lb, T_p, D]  ->  dst [lb, F, D] features
 *             feat_off + (rank-major table order), i.e. straight into the interaction input.
 * a2a_pack_grad: the transpose, dfeat [lb, F, D] -> send buffer concat over p of [lb, T_p, D].
 *   tables_per_rank: HOST int array [W].  *_stride_b / *_stride_f: floats between samples / between features of the
 *   [lb, F, D] block (stride_f >= D: the interaction input may be pitched, e.g. D + 8 for the bulk-copy staging).
 * ------------------------------------------------------------------------------------------- */
int dlrmx_a2a_unpack(const float* recv, int W, const int* host_tables_per_rank, int lb, int D,
                     float* dst, int64_t dst_stride_b, int64_t dst_stride_f, int feat_off, dlrmx_stream_t stream);
int dlrmx_a2a_pack_grad(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                        const int* host_tables_per_rank, int lb, int D,
                        float* send, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Exchange fused into the producing kernels over NVLink peer memory (no NCCL call, no send / receive
 * staging, no unpack pass).  Replaces the same reference code as the block above (ext_dist.alltoall,
 * extend_distributed.py:405-508, 601-639) when every rank's interaction input and pooled-row gradient
 * buffer are peer-mapped (CUDA VMM / torch symmetric memory); ranks synchronise with a device barrier
 * before the first and after the last remote store (host layer: dlrm_yx_b200/extend_distributed.py).
 *
 * embbag_fwd_scatter: dlrmx_embbag_fwd, but the pooled row of bag (t, b) of the GLOBAL batch is stored
 *   at  out_peers[b / lb] + (b % lb)*stride_b + t*stride_t : out_peers is a DEVICE array [B / lb] of
 *   peer-mapped pointers, each already offset to this rank's first feature slot in that peer's
 *   [lb, 1+T, D] interaction input.
 * a2a_push_grad: dfeat [lb, F, D] (this rank's local batch, all tables) -> for every owner rank p the
 *   chunk [lb, T_p, D] stored at host_peer_chunk_bases[p] (HOST array [W] of peer-mapped device
 *   pointers = rank p's gradient buffer [W*lb, T_p, D] + my_rank*lb*T_p*D).
 * ------------------------------------------------------------------------------------------- */
int dlrmx_embbag_fwd_scatter_i32(const float* const* tables, const int64_t* table_rows,
                                 const int32_t* indices, const int32_t* offsets, int64_t N, int T, int B,
                                 int D, int64_t table_ld, const float* per_sample_weights, float* const* out_peers, int lb,
                                 int64_t stride_b, int64_t stride_t, int* err_flag, dlrmx_stream_t stream);
int dlrmx_embbag_fwd_scatter_i64(const float* const* tables, const int64_t* table_rows,
                                 const int64_t* indices, const int64_t* offsets, int64_t N, int T, int B,
                                 int D, int64_t table_ld, const float* per_sample_weights, float* const* out_peers, int lb,
                                 int64_t stride_b, int64_t stride_t, int* err_flag, dlrmx_stream_t stream);
int dlrmx_a2a_push_grad(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                        const int* host_tables_per_rank, int lb, int D,
                        float* const* host_peer_chunk_bases, dlrmx_stream_t stream);
/* bf16 on the wire (the mixed-precision mode of the towers): the same slices, rounded to nearest even, into bf16 gradient
 * buffers [W*lb, T_p, D] of the owners; dlrmx_bwd_update_ex(grad_is_bf16 = 1) reads them.  D % 8 == 0. */
int dlrmx_a2a_push_grad_bf16(const float* dfeat, int64_t dfeat_stride_b, int64_t dfeat_stride_f, int feat_off, int W,
                             const int* host_tables_per_rank, int lb, int D,
                             void* const* host_peer_chunk_bases_bf16, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Data-parallel gradient all-reduce over NVLink peer memory (csrc/peer_allreduce.cu).  Replaces the
 * all-reduce DistributedDataParallel runs for the reference's replicated MLP towers
 * (dlrm_s_pytorch.py:1606-1613, DDP(bot_l) / DDP(top_l); extend_distributed.py:110-166 sets the
 * process group up) -- and serves the replicated small tables' dense row gradients of this repo.
 * host_peer_buffers = HOST array [W] of peer-mapped device pointers, entry p = rank p's buffer
 * (entry `rank` = the caller's own), all of `numel` elements (fp32, or bf16 when is_bf16), 16-byte
 * aligned, numel a multiple of 4 (fp32) / 8 (bf16).  In place: on return (stream-ordered, and after
 * the caller's closing barrier) every buffer holds the element-wise sum, added in rank order
 * 0 .. W-1 in fp32 (bf16: widened, added, rounded once) -- bit-identical on every rank.
 * The caller brackets the call with device barriers over the W ranks on the same stream: BEFORE
 * (every rank's operand is complete) and AFTER (every remote store has landed).  W = 1: no-op sum.
 * ------------------------------------------------------------------------------------------- */
int dlrmx_peer_allreduce(void* const* host_peer_buffers, int W, int rank, int64_t numel, int is_bf16,
                         dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Loss: binary cross-entropy, mean reduction.  Replaces dlrm.loss_fn = torch.nn.BCELoss(reduction="mean")
 * (dlrm_s_pytorch.py:505-507, called through loss_fn_wrap 170-173) and its backward: one kernel each way instead of
 * five element-wise / reduce launches, same arithmetic as ATen (log clamped at -100, gradient denominator at 1e-12).
 *   p, target [n] fp32;  loss [1] fp32 (device);  dloss [1] fp32 (device: the incoming gradient of the loss);
 *   workspace: dlrmx_bce_mean_workspace_bytes(n); counter: device uint32, zero before the first call (self-resetting).
 * The sum is deterministic (fixed tree per CTA, partials added in index order).
 * ------------------------------------------------------------------------------------------- */
size_t dlrmx_bce_mean_workspace_bytes(int64_t n);
int dlrmx_bce_mean_fwd(const float* p, const float* target, int64_t n, float* workspace, unsigned int* counter,
                       float* loss, dlrmx_stream_t stream);
int dlrmx_bce_mean_bwd(const float* p, const float* target, const float* dloss, int64_t n, float* dp,
                       dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Weighted pooling, backward (per-table path).  The reference gathers one weight per looked-up index from a per-row
 * vector, `per_sample_weights = v_W_l[k].gather(0, indices)` (dlrm_s_pytorch.py:544-547); with
 * weighted_pooling="learned" that vector is a Parameter (475-478), so EmbeddingBag's backward must deliver
 *   dpsw[l] = < grad_out[bag(l), :], W[indices[l], :] >                                   (embbag_psw_grad)
 *   gsum[u] = sum over the lookups l of unique row u of  psw[l] * grad_out[bag(l), :]     (bwd_coalesce_weighted)
 * bwd_coalesce_weighted is dlrmx_bwd_coalesce with weights: the sort is run with every lookup as its own bag, so its
 * payload sorted_pos[p] is the ORIGINAL position of sorted element p; pos2bag [N] maps a position to its bag.
 * grad is [B, D] with row stride stride_b (one table).  No [N, D] gradient expansion is written.
 * ------------------------------------------------------------------------------------------- */
int dlrmx_embbag_psw_grad_i32(const float* table, int64_t table_ld, int64_t rows, const int32_t* indices,
                              const int32_t* offsets, int64_t N, int B, int D, const float* grad, int64_t stride_b,
                              float* dpsw, dlrmx_stream_t stream);
int dlrmx_embbag_psw_grad_i64(const float* table, int64_t table_ld, int64_t rows, const int64_t* indices,
                              const int64_t* offsets, int64_t N, int B, int D, const float* grad, int64_t stride_b,
                              float* dpsw, dlrmx_stream_t stream);
int dlrmx_bwd_coalesce_weighted(int D, const float* grad, int64_t stride_b, const int32_t* sorted_pos,
                                const int32_t* pos2bag, const float* psw, const int32_t* run_off,
                                const int32_t* num_unique, int64_t N, float* gsum, dlrmx_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Criteo Terabyte binary records -> one batch in the hot path's input layout.
 * Replaces CriteoBinDataset.__getitem__ + _transform_features (data_loader_terabyte.py:205-252, 68-95),
 * which the reference runs on the host with torch CPU ops.
 *   records  [B, 40] int32 as they lie in the file: label, 13 dense counts, 26 categorical ids
 *            (device copy of the raw bytes, 16-byte aligned)
 *   dense    [B, 13] fp32 = log(float(count) + 1)
 *   indices  [26*B]  categorical ids, table-major (the reference's x_cat.t()), taken modulo
 *            max_ind_range when max_ind_range > 0 (floor modulo, like torch's % on int32)
 *   offsets  batched != 0: [26*B + 1] = arange (packed layout of the batched op, data_loader_terabyte.py:88-93);
 *            batched == 0: [26, B], every row arange(B) (lS_o of the per-table API)
 *   target   [B] fp32 = float(label)   (the reference views it as [B, 1])
 * Integer outputs are bit-exact with the reference; dense differs by the device logf (<= 2 ulp).
 * ------------------------------------------------------------------------------------------- */
int dlrmx_criteo_bin_batch_i32(const int32_t* records, int B, int max_ind_range, int batched, float* dense,
                               int32_t* indices, int32_t* offsets, float* target, dlrmx_stream_t stream);
int dlrmx_criteo_bin_batch_i64(const int32_t* records, int B, int max_ind_range, int batched, float* dense,
                               int64_t* indices, int64_t* offsets, float* target, dlrmx_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DLRMX_H_ */
