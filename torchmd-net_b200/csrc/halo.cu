// halo.cu -- pack kernel of the halo exchange used by spatially decomposed runs (torchmdnet/distributed.py).
// Rows of a node tensor ([n, row_floats] fp32; row_floats = 9*H for TensorNet features, 3 for positions) owned by
// this rank and needed by a neighbour are gathered into one contiguous send buffer; the receive side needs no
// unpack because ghost rows are the contiguous tail of every local array.
#include <stdlib.h>
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
    if (blocks > (long long)tmdn_sm_count() * 16) blocks = (long long)tmdn_sm_count() * 16;   // grid-stride over 16 CTAs per SM
    if (v4)
        launch_k(gather_rows4_kernel, (unsigned)blocks, 256, 0, stream, (const float4*)src, idx, n_rows, (int)(row_floats / 4),
                                                                  (float4*)out);
    else
        launch_k(gather_rows_kernel, (unsigned)blocks, 256, 0, stream, src, idx, n_rows, (int)row_floats, out);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

// ------------------------------------------------------------------ push over peer memory
namespace {
// one thread per float4 (or float) of the packed rows; the peer segment of a row is found by a short linear search
// (n_peers <= 26).  Writes are plain stores into the peer's address space: NVLink carries them, nothing is staged.
template <typename V>
__global__ void push_rows_kernel(const V* __restrict__ src, const int32_t* __restrict__ idx,
                                 const long long* __restrict__ seg, V* const* __restrict__ dst, int n_peers,
                                 long long n_rows, int row_vec) {
    pdl_sync();
    const long long total = n_rows * row_vec;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const long long r = t / row_vec;
        const int c = (int)(t - r * row_vec);
        int p = 0;
        while (p + 1 < n_peers && r >= seg[p + 1]) ++p;
        dst[p][(size_t)(r - seg[p]) * row_vec + c] = src[(size_t)idx[r] * row_vec + c];
    }
    __threadfence_system();
}
}  // namespace

extern "C" int tmdn_halo_push_f32(const float* src, const int32_t* idx, const int64_t* seg, float* const* dst, int n_peers,
                                  int64_t n_rows_total, int64_t row_floats, tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_rows_total == 0 || n_peers == 0) return TMDN_OK;
    TMDN_CHECK_ARG(src && idx && seg && dst, "halo_push: null pointer argument");
    TMDN_CHECK_ARG(n_peers > 0 && n_peers <= 64 && n_rows_total > 0 && row_floats > 0 && row_floats < (1 << 30),
                   "halo_push: bad sizes");
    KTimer ktimer__("halo_push", stream);
    const bool v4 = row_floats % 4 == 0 && ((uintptr_t)src % 16 == 0);   // ghost rows of 16-byte-aligned arrays stay aligned
    const long long total = n_rows_total * (v4 ? row_floats / 4 : row_floats);
    long long blocks = cdiv(total, 256);
    // CTAs of the push: it runs beside the interior rows of the producing stage and takes SM slots from them
    // (TMDN_PUSH_CTAS overrides the cap for A/B runs)
    static const long long cap_env = [] { const char* e = getenv("TMDN_PUSH_CTAS"); return e ? atoll(e) : 0LL; }();
    const long long cap = cap_env > 0 ? cap_env : (long long)tmdn_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (v4)
        launch_k(push_rows_kernel<float4>, (unsigned)blocks, 256, 0, stream, (const float4*)src, idx, (const long long*)seg,
                 (float4* const*)dst, n_peers, (long long)n_rows_total, (int)(row_floats / 4));
    else
        launch_k(push_rows_kernel<float>, (unsigned)blocks, 256, 0, stream, src, idx, (const long long*)seg, dst, n_peers,
                 (long long)n_rows_total, (int)row_floats);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}
