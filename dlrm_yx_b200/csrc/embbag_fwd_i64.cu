// C ABI of the embedding-bag forward, int64 indices (kernels in embbag_fwd.cuh; one translation unit per index type so the
// two halves of the instantiation set compile in parallel).
#include "embbag_fwd.cuh"

extern "C" {
int dlrmx_embbag_fwd_i64(const float* const* tables, const int64_t* table_rows, const int64_t* indices,
                         const int64_t* offsets, int64_t N, int T, int B, int D, int64_t table_ld, const float* psw,
                         float* out, int64_t stride_b, int64_t stride_t, int* err_flag,
                         dlrmx_stream_t stream) {
    return dlrmx::embbag_fwd<int64_t>(tables, table_rows, indices, offsets, N, T, B, D, table_ld, psw, out, stride_b,
                                      stride_t, err_flag, nullptr, 0, stream);
}
int dlrmx_embbag_fwd_scatter_i64(const float* const* tables, const int64_t* table_rows, const int64_t* indices,
                                 const int64_t* offsets, int64_t N, int T, int B, int D, int64_t table_ld, const float* psw,
                                 float* const* out_peers, int lb, int64_t stride_b,
                                 int64_t stride_t, int* err_flag, dlrmx_stream_t stream) {
    using namespace dlrmx;
    DLRMX_REQUIRE(out_peers != nullptr, "embbag_fwd_scatter: out_peers is null");
    return dlrmx::embbag_fwd<int64_t>(tables, table_rows, indices, offsets, N, T, B, D, table_ld, psw,
                                      nullptr, stride_b, stride_t, err_flag, out_peers, lb, stream);
}
}
