// halo.cu -- pack kernel of the halo exchange used by spatially decomposed runs (torchmdnet/distributed.py).
// Rows of a node tensor ([n, row_floats] fp32; row_floats = 9*H for TensorNet features, 3 for positions) owned by
// this rank and needed by a neighbour are gathered into one contiguous send buffer; the receive side needs no
// unpack because ghost rows are the contiguous tail of every local array.
#include "common.cuh"

namespace {
__global__ void gather_rows_kernel(const float* __restrict__ src, const int32_t* __restrict__ idx, long long n_rows,
                                   int row_floats, float* __restrict__ out) {
    pdl_sync();
    const long long total = n_rows * row_floats;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const long long r = t / row_floats;
        const int c = (int)(t - r * row_floats);
        out[t] = src[(size_t)idx[r] * row_floats + c];
    }
}
__global__ void gather_rows4_kernel(const float4* __restrict__ src, const int32_t* __restrict__ idx, long long n_rows,
                                    int row_vec, float4* __restrict__ out) {
    pdl_sync();
    const long long total = n_rows * row_vec;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const long long r = t / row_vec;
        const int c = (int)(t - r * row_vec);
        out[t] = src[(size_t)idx[r] * row_vec + c];
    }
}
}  // namespace

extern "C" int tmdn_halo_pack_f32(const float* src, const int32_t* idx, int64_t n_rows, int64_t row_floats, float* out,
                                  tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_rows == 0) return TMDN_OK;
    TMDN_CHECK_ARG(src && idx && out, "halo_pack: null pointer argument");
    TMDN_CHECK_ARG(n_rows > 0 && row_floats > 0 && row_floats < (1 << 30), "halo_pack: bad sizes");
    KTimer ktimer__("halo_pack", stream);
    const bool v4 = row_floats % 4 == 0 && ((uintptr_t)src % 16 == 0) && ((uintptr_t)out % 16 == 0);
    const long long total = n_rows * (v4 ? row_floats / 4 : row_floats);
    long long blocks = cdiv(total, 256);
    if (blocks > 148 * 16) blocks = 148 * 16;   // grid-stride over 16 CTAs per SM
    if (v4)
        launch_k(gather_rows4_kernel, (unsigned)blocks, 256, 0, stream, (const float4*)src, idx, n_rows, (int)(row_floats / 4),
                                                                  (float4*)out);
    else
        launch_k(gather_rows_kernel, (unsigned)blocks, 256, 0, stream, src, idx, n_rows, (int)row_floats, out);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}
