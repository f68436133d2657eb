// edge_common.cuh -- thread mapping, reductions and radial helpers shared by the edge kernels of tensornet.cu / et.cu.
#pragma once
#include "common.cuh"

__device__ __forceinline__ void cutoff_fn(float d, float lo, float hi, float& c, float& dc) {
    const float PI = 3.14159265358979323846f;
    if (lo > 0.f) {  // models/utils.py:369-385
        const float arg = PI * (2.0f * (d - lo) / (hi - lo) + 1.0f);
        const float m = (d < hi && d > lo) ? 1.0f : 0.0f;
        c = 0.5f * (cosf(arg) + 1.0f) * m;
        dc = -0.5f * sinf(arg) * (2.0f * PI / (hi - lo)) * m;
    } else {         // models/utils.py:386-390
        const float m = d < hi ? 1.0f : 0.0f;
        c = 0.5f * (cosf(d * PI / hi) + 1.0f) * m;
        dc = -0.5f * sinf(d * PI / hi) * (PI / hi) * m;
    }
}


// ---- edge kernels: block = (H channels, S edge slices, A atoms).  Slice s of an atom's thread group walks the
// edges r0+s, r0+s+S, ... of the row; partial sums are combined through shared memory in slice order, so the
// result depends only on (row content, S) and stays bit-reproducible.  S > 1 shortens the serial gather chain of
// small systems (166 atoms: 8 slices -> 15 dependent iterations instead of 120).
template <int NV> __device__ __forceinline__ void slice_reduce(float (&v)[NV], float* s_part, int H) {
    const int S = blockDim.y;
    if (S == 1) return;
    const int h = threadIdx.x, sl = threadIdx.y, a = threadIdx.z;
    __syncthreads();
#pragma unroll
    for (int c = 0; c < NV; ++c) s_part[(((size_t)a * S + sl) * NV + c) * H + h] = v[c];
    __syncthreads();
    if (sl == 0) {
#pragma unroll
        for (int c = 0; c < NV; ++c) {
            float t = 0.f;
            for (int q = 0; q < S; ++q) t += s_part[(((size_t)a * S + q) * NV + c) * H + h];
            v[c] = t;
        }
    }
}
// Sum over all threads of one atom's group (channels x slices).  Every thread of the block must call.
template <int NV> __device__ __forceinline__ void atom_reduce(float (&v)[NV], float* s_red) {
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int lane = tid & 31, wr = tid >> 5, nw = (blockDim.x * blockDim.y) >> 5, a = threadIdx.z;
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = warp_sum(v[k]);
    __syncthreads();
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < NV; ++k) s_red[(a * nw + wr) * NV + k] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        float t = 0.f;
        for (int w = 0; w < nw; ++w) t += s_red[(a * nw + w) * NV + k];
        v[k] = t;
    }
}
// row handled by this (blockIdx.x, threadIdx.z) of a launch over the rows [row0, N): identity, or through the visiting
// order of small systems (full-range launches only)
__device__ __forceinline__ int edge_row(const int* __restrict__ order, int row0, int N, bool& valid) {
    int i = row0 + blockIdx.x * blockDim.z + threadIdx.z;
    valid = i < N;
    if (order != nullptr && valid) i = order[i];
    return i;
}

struct EdgeLaunch { dim3 block; unsigned grid; size_t smem; };
// s_red (<= 32 warps x 3 values per atom) lives at the start of dynamic shared memory, slice partials after it
constexpr int EDGE_RED_FLOATS = 128;
static inline EdgeLaunch edge_launch(int N, int H, int max_threads, int nv = 9) {
    int smax = max_threads / H;
    if (smax < 1) smax = 1;
    const long long target = (long long)tmdn_sm_count() * 2048;
    int S = 1;
    while (S < smax && (long long)N * H * S < target) S *= 2;
    int A = 128 / (H * S);
    if (A < 1) A = 1;
    EdgeLaunch l;
    l.block = dim3((unsigned)H, (unsigned)S, (unsigned)A);
    l.grid = (unsigned)cdiv(N, A);
    l.smem = sizeof(float) * ((size_t)EDGE_RED_FLOATS * A + (S > 1 ? (size_t)A * S * nv * H : 0));
    return l;
}

