// Criteo Terabyte binary records -> one training batch in the hot path's input layout, on the device.  sm_100a.
//
// Replaces CriteoBinDataset.__getitem__ + _transform_features of the reference (data_loader_terabyte.py:205-252 and
// 68-95), which run on the host with torch CPU ops: a batch is B consecutive records of 40 little-endian int32
// [label, 13 dense counts, 26 categorical ids]; the reference produces
//     X      = log(float(dense) + 1)                                  fp32 [B, 13]
//     lS_i   = (cat % max_ind_range  if max_ind_range > 0) transposed  int64 [26, B]  (floor modulo, like torch's %)
//     lS_o   = arange(B) per table                                     int64 [26, B]
//     T      = float(label)                                            fp32 [B, 1]
// and, for the batched embedding op, indices = lS_i flattened table-major as int32 [26*B] and
// offsets = arange(26*B + 1) int32 (data_loader_terabyte.py:88-93).
// Here the raw records are copied to the device as they lie in the file (160 B per sample, the only H2D traffic of a
// batch) and ONE kernel does the split, the modulo, the log, the [B,26] -> [26,B] transpose and the offsets:
// a CTA stages 64 records in shared memory with coalesced 16-byte loads and writes every output coalesced.
// HBM-bound: algorithmic bytes = B*160 (records) + B*13*4 + 26*B*ib (indices) + (26*B + 1)*ib (offsets) + B*4.
#include "common.cuh"

namespace dlrmx {

constexpr int kCriteoFeat = 40;     // 1 label + 13 dense + 26 categorical (data_loader_terabyte.py:210-215)
constexpr int kCriteoDense = 13;
constexpr int kCriteoCat = 26;
constexpr int kRecPerCta = 64;

template <typename IdxT>
__global__ void __launch_bounds__(256)
criteo_bin_kernel(const int32_t* __restrict__ records, int B, int max_ind_range, int batched,
                  float* __restrict__ dense, IdxT* __restrict__ indices, IdxT* __restrict__ offsets,
                  float* __restrict__ target) {
    __shared__ int32_t sh[kRecPerCta * kCriteoFeat];
    const int b0 = blockIdx.x * kRecPerCta;
    const int nrec = min(kRecPerCta, B - b0);
    const int nint = nrec * kCriteoFeat;
    // records of the tile are contiguous: 160 B per record keeps every tile 16-byte aligned
    const int4* src = reinterpret_cast<const int4*>(records + (int64_t)b0 * kCriteoFeat);
    for (int i = threadIdx.x; i < nint / 4; i += blockDim.x) reinterpret_cast<int4*>(sh)[i] = __ldg(src + i);
    __syncthreads();
    // dense block of the tile is contiguous in the output: [nrec, 13]
    for (int i = threadIdx.x; i < nrec * kCriteoDense; i += blockDim.x) {
        const int r = i / kCriteoDense, c = i - r * kCriteoDense;
        const float x = (float)sh[r * kCriteoFeat + 1 + c];
        dense[(int64_t)b0 * kCriteoDense + i] = logf(__fadd_rn(x, 1.0f));
    }
    for (int r = threadIdx.x; r < nrec; r += blockDim.x) target[b0 + r] = (float)sh[r * kCriteoFeat];
    // categorical ids, transposed: table t gets nrec consecutive entries at t*B + b0
    for (int i = threadIdx.x; i < kCriteoCat * kRecPerCta; i += blockDim.x) {
        const int t = i / kRecPerCta, r = i - t * kRecPerCta;
        if (r < nrec) {
            int32_t v = sh[r * kCriteoFeat + 1 + kCriteoDense + t];
            if (max_ind_range > 0) {              // floor modulo (torch `%` on int32: sign of the divisor)
                v %= max_ind_range;
                if (v < 0) v += max_ind_range;
            }
            const int64_t bag = (int64_t)t * B + b0 + r;
            indices[bag] = (IdxT)v;
            // batched op: offsets = arange(26*B + 1); per-table API: arange(B) for every table
            offsets[bag] = batched ? (IdxT)bag : (IdxT)(b0 + r);
        }
    }
    if (batched && blockIdx.x == 0 && threadIdx.x == 0) offsets[(int64_t)kCriteoCat * B] = (IdxT)((int64_t)kCriteoCat * B);
}

template <typename IdxT>
static int criteo_bin(const int32_t* records, int B, int max_ind_range, int batched, float* dense, IdxT* indices,
                      IdxT* offsets, float* target, dlrmx_stream_t stream) {
    DLRMX_REQUIRE(B >= 0, "criteo_bin_batch: negative batch");
    DLRMX_REQUIRE(((uintptr_t)records & 15) == 0, "criteo_bin_batch: records must be 16-byte aligned");
    DLRMX_REQUIRE((int64_t)kCriteoCat * B < (1ll << 31), "criteo_bin_batch: 26*B must be < 2^31");
    if (B == 0) return DLRMX_OK;
    const unsigned blocks = (unsigned)ceil_div(B, kRecPerCta);
    criteo_bin_kernel<IdxT><<<blocks, 256, 0, (cudaStream_t)stream>>>(records, B, max_ind_range, batched, dense,
                                                                      indices, offsets, target);
    return check_launch("criteo_bin_batch");
}

}  // namespace dlrmx

extern "C" {
int dlrmx_criteo_bin_batch_i32(const int32_t* records, int B, int max_ind_range, int batched, float* dense,
                               int32_t* indices, int32_t* offsets, float* target, dlrmx_stream_t stream) {
    return dlrmx::criteo_bin<int32_t>(records, B, max_ind_range, batched, dense, indices, offsets, target, stream);
}
int dlrmx_criteo_bin_batch_i64(const int32_t* records, int B, int max_ind_range, int batched, float* dense,
                               int64_t* indices, int64_t* offsets, float* target, dlrmx_stream_t stream) {
    return dlrmx::criteo_bin<int64_t>(records, B, max_ind_range, batched, dense, indices, offsets, target, stream);
}
}
