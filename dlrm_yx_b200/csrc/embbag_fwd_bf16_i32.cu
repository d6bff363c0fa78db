// C ABI of the embedding-bag forward over bf16 table rows, int32_t indices (kernels in embbag_fwd.cuh, TabT = __nv_bfloat16).
#include "embbag_fwd.cuh"

extern "C" {
int dlrmx_embbag_fwd_bf16_i32(const void* const* tables_bf16, const int64_t* table_rows, const int32_t* indices,
                              const int32_t* offsets, int64_t N, int T, int B, int D, int64_t table_ld, const float* psw,
                              float* out, int64_t stride_b, int64_t stride_t, int* err_flag, dlrmx_stream_t stream) {
    return dlrmx::embbag_fwd<int32_t, __nv_bfloat16>(reinterpret_cast<const __nv_bfloat16* const*>(tables_bf16), table_rows, indices,
                                                 offsets, N, T, B, D, table_ld, psw, out, stride_b, stride_t, err_flag, nullptr, 0,
                                                 stream);
}
}
