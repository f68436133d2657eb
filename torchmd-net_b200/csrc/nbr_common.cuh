// nbr_common.cuh -- arithmetic and cell-binning helpers shared by the pair-list kernels (neighbors.cu) and the
// direct CSR builder (csr_direct.cu).  All distance arithmetic uses explicit round-to-nearest mul/add/sub/div (no
// FMA contraction) so that deltas and distances are bit-identical to the CPU restatement in oracle/neighbors.py.
#pragma once
#include "common.cuh"
#include <limits.h>

namespace nbr {

constexpr int NB_THREADS = 128;

template <typename T> struct Box3 { T m[3][3]; };

__device__ __forceinline__ float rn_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double rn_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float rn_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double rn_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float rn_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double rn_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ float rn_div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double rn_div(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float sqrt_rn(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ double sqrt_rn(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ float round_(float a) { return roundf(a); }
__device__ __forceinline__ double round_(double a) { return round(a); }
__device__ __forceinline__ float floor_(float a) { return floorf(a); }
__device__ __forceinline__ double floor_(double a) { return floor(a); }

// Number of box lengths to subtract: round(d / L) as the reference computes it (IEEE division, then round), without the
// division for the cases that matter.  Exact: for |d| < L/2 the quotient rounds to a float strictly below 0.5 (the
// neighbour of 0.5 is 2^-25 away, the quotient at least that far), so the result is 0; for L/2 < |d| < 1.25 L it
// rounds into (0.5, 1.5), so the result is +-1.  Ties (|d| == L/2 exactly) and far images take the division.
template <typename T> __device__ __forceinline__ T image_count(T d, T L) {
    const T a = fabs(d), h = T(0.5) * L;      // 0.5 * L is exact
    if (a < h) return T(0);
    if (a > h && a < T(1.25) * L) return d < T(0) ? T(-1) : T(1);
    return round_(rn_div(d, L));
}
// neighbors/common.cuh:180-194 (triclinic minimum image, z -> y -> x)
template <typename T> __device__ __forceinline__ void min_image_tric(T& dx, T& dy, T& dz, const Box3<T>& b) {
    T s = image_count(dz, b.m[2][2]);
    if (s != T(0)) {
        dx = rn_sub(dx, rn_mul(s, b.m[2][0]));
        dy = rn_sub(dy, rn_mul(s, b.m[2][1]));
        dz = rn_sub(dz, rn_mul(s, b.m[2][2]));
    }
    s = image_count(dy, b.m[1][1]);
    if (s != T(0)) {
        dx = rn_sub(dx, rn_mul(s, b.m[1][0]));
        dy = rn_sub(dy, rn_mul(s, b.m[1][1]));
        dz = rn_sub(dz, rn_mul(s, b.m[1][2]));
    }
    s = image_count(dx, b.m[0][0]);
    if (s != T(0)) {
        dx = rn_sub(dx, rn_mul(s, b.m[0][0]));
        dy = rn_sub(dy, rn_mul(s, b.m[0][1]));
        dz = rn_sub(dz, rn_mul(s, b.m[0][2]));
    }
}
// neighbors/common.cuh:142-148 (rectangular minimum image)
template <typename T> __device__ __forceinline__ T min_image_rect1(T d, T L) {
    return rn_sub(d, rn_mul(floor_(rn_add(rn_div(d, L), T(0.5))), L));
}
template <typename T> __device__ __forceinline__ T dist2(T dx, T dy, T dz) {
    return rn_add(rn_add(rn_mul(dx, dx), rn_mul(dy, dy)), rn_mul(dz, dz));
}

// ------------------------------------------------------------------ cell list
template <typename T> struct CellGrid {
    int n[3];       // cells per dimension (cell edge L/n >= cutoff*(1+1e-4))
    int lo[3], hi[3];  // neighbour offsets to visit per dimension (unique cells only)
    T L[3];
    int ncells;
};
template <typename T> struct SortedAtom { T x, y, z; int idx; };

template <typename T> __device__ __forceinline__ int cell_coord(T p, T L, int n) {
    const T w = rn_sub(p, rn_mul(floor_(rn_add(rn_div(p, L), T(0.5))), L));  // wrapped to [-L/2, L/2)
    int c = (int)floor_((w / L + T(0.5)) * T(n));
    return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

template <typename T>
__global__ void cell_assign_kernel(const T* __restrict__ pos, int n, CellGrid<T> g, int* __restrict__ cell_cnt,
                                   int* __restrict__ cell_of, int* __restrict__ slot_of) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cx = cell_coord(pos[3 * (size_t)i], g.L[0], g.n[0]);
    const int cy = cell_coord(pos[3 * (size_t)i + 1], g.L[1], g.n[1]);
    const int cz = cell_coord(pos[3 * (size_t)i + 2], g.L[2], g.n[2]);
    const int c = cx + g.n[0] * (cy + g.n[1] * cz);
    cell_of[i] = c;
    slot_of[i] = atomicAdd(&cell_cnt[c], 1);
}

template <typename T>
__global__ void cell_scatter_kernel(const T* __restrict__ pos, const int64_t* __restrict__ batch, int n,
                                    const int* __restrict__ cell_start, const int* __restrict__ cell_of,
                                    const int* __restrict__ slot_of, SortedAtom<T>* __restrict__ spos,
                                    long long* __restrict__ sbatch) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = cell_start[cell_of[i]] + slot_of[i];
    SortedAtom<T> a;
    a.x = pos[3 * (size_t)i];
    a.y = pos[3 * (size_t)i + 1];
    a.z = pos[3 * (size_t)i + 2];
    a.idx = i;
    spos[p] = a;
    sbatch[p] = batch[i];
}

template <typename T> bool make_grid(const double box[9], double cutoff, CellGrid<T>& g) {
    long long nc = 1;
    for (int k = 0; k < 3; ++k) {
        const double L = box[4 * k];
        int n = (int)(L / (cutoff * (1.0 + 1e-4)));
        if (n < 1) n = 1;
        if (n > 1024) return false;  // neighbors_cuda_cell.cuh:23-25
        g.n[k] = n;
        g.L[k] = (T)L;
        if (n == 1) { g.lo[k] = 0; g.hi[k] = 0; }
        else if (n == 2) { g.lo[k] = 0; g.hi[k] = 1; }
        else { g.lo[k] = -1; g.hi[k] = 1; }
        nc *= n;
    }
    g.ncells = (int)nc;
    return true;
}


}  // namespace nbr
