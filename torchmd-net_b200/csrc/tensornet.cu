// tensornet.cu -- TensorNet energy + analytic forces for sm_100a (fp32).
//
// Replaces TensorNet.forward / TensorEmbedding / Interaction (torchmdnet/models/tensornet.py:200-410), the
// Scalar head + reduce (output_modules.py:27-43,76-77) and the autograd force path of TorchMD_Net.forward
// (models/model.py:286-298) of the reference.  The stage list, buffer names and every adjoint formula are the
// ones of oracle/staged_tensornet.py, which is checked against torch.autograd on the CPU.
//
// Data layout in HBM: node tensors are [N, 9, H] fp32 in irreducible components (tensor3.cuh); a thread owns
// one (atom, channel) pair, so all loads/stores of a warp are 128-byte coalesced rows and no per-edge tensor
// ([E,H,3,3] in the reference) ever exists.  Edges are CSR by target atom, sorted by source; every sum over
// neighbours is an in-order loop in registers (deterministic, no float atomics).  Because the edge set is
// symmetric, the backward pass of a message-passing stage is evaluated inside the SAME row structure: row i
// gathers the adjoints of its neighbours and also accumulates both directions' contribution to dE/dpos_i.
#include "common.cuh"
#include "gemm.cuh"
#include "node_gemm.cuh"
#include "tc_mix.cuh"
#include "tensor3.cuh"
#include "edge_common.cuh"
#include "table_grid.cuh"
#include "scan.cuh"
#include <vector>
#include <string>

namespace {

// ------------------------------------------------------------------ filters f_g(d) (g = I, A, S), f*C(d) and d/dd
struct FilterView {
    const float* F;     // [E,3,H]   (exact mode, CSR edge order)
    const float* dF;    // [E,3,H]
    const float4* tab;  // [n_int,3,H] cubic coefficients (table mode) on the grid `grid` (table_grid.cuh)
    TabGrid grid;
    const float4* hot;  // shared-memory copy of rows [0, n_hot): the far end of the grid (persistent kernels only)
    int n_hot;
};

// coefficient row of filter g of interval k for channel h: shared memory for the hot rows, L2 for the rest
template <bool HOT>
__device__ __forceinline__ float4 filter_row(const FilterView& fv, int k, int g, int H, int h) {
    if (HOT && k < fv.n_hot) return fv.hot[((size_t)k * 3 + g) * H + h];
    return __ldg(&fv.tab[((size_t)k * 3 + g) * H + h]);
}

template <bool TAB, bool HOT = false>
__device__ __forceinline__ void filter_val(const FilterView& fv, size_t e, float d, int H, int h, float f[3]) {
    if (TAB) {
        int k;
        float s, ds;
        tab_locate(fv.grid, d, k, s, ds);
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const float4 c = filter_row<HOT>(fv, k, g, H, h);
            f[g] = c.x + s * (c.y + s * (c.z + s * c.w));
        }
    } else {
#pragma unroll
        for (int g = 0; g < 3; ++g) f[g] = fv.F[(e * 3 + g) * H + h];
    }
}
template <bool TAB, bool HOT = false>
__device__ __forceinline__ void filter_val_grad(const FilterView& fv, size_t e, float d, int H, int h, float f[3],
                                                float df[3]) {
    if (TAB) {
        int k;
        float s, ds;
        tab_locate(fv.grid, d, k, s, ds);
        const float s2 = 2.0f * s, s3 = 3.0f * s * s;      // df/ds = c.y + 2 s c.z + 3 s^2 c.w as two FFMA per filter
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const float4 c = filter_row<HOT>(fv, k, g, H, h);
            f[g] = c.x + s * (c.y + s * (c.z + s * c.w));
            df[g] = fmaf(s3, c.w, fmaf(s2, c.z, c.y)) * ds;
        }
    } else {
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            f[g] = fv.F[(e * 3 + g) * H + h];
            df[g] = fv.dF[(e * 3 + g) * H + h];
        }
    }
}

// ------------------------------------------------------------------ small device utilities
__device__ __forceinline__ void ld9(const float* __restrict__ p, size_t i, int H, int h, float x[9]) {
    const float* q = p + i * 9 * H + h;
#pragma unroll
    for (int c = 0; c < 9; ++c) x[c] = q[(size_t)c * H];
}
__device__ __forceinline__ void st9(float* __restrict__ p, size_t i, int H, int h, const float x[9]) {
    float* q = p + i * 9 * H + h;
#pragma unroll
    for (int c = 0; c < 9; ++c) q[(size_t)c * H] = x[c];
}

// Sum over the channels of one atom row.  Block = (H, APB); every thread of the block must call.
template <int NV> __device__ __forceinline__ void row_reduce(float (&v)[NV], float* s_red) {
    const int lane = threadIdx.x & 31, wr = threadIdx.x >> 5, nw = blockDim.x >> 5, a = threadIdx.y;
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = warp_sum(v[k]);
    __syncthreads();
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < NV; ++k) s_red[(a * nw + wr) * NV + k] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        float s = 0.f;
        for (int w = 0; w < nw; ++w) s += s_red[(a * nw + w) * NV + k];
        v[k] = s;
    }
}

constexpr int LPT_MAX_ROWS = 1024;
// z -> int32 with range check, overflow status; optionally (order != null, n_rows <= LPT_MAX_ROWS) the row visiting
// order of the edge kernels: rows sorted by decreasing degree, ties by index (a pure function of the graph)
__global__ void tn_prep_kernel(const int64_t* __restrict__ z, int* __restrict__ z32, int n, int max_z,
                               const int32_t* __restrict__ num_found, long long max_edges, int32_t* __restrict__ status,
                               const int* __restrict__ row_ptr, int n_rows, int* __restrict__ order,
                               const int64_t* __restrict__ batch) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (order != nullptr) {
        __shared__ int s_deg[LPT_MAX_ROWS];
        for (int j = threadIdx.x; j < n_rows; j += blockDim.x) s_deg[j] = row_ptr[j + 1] - row_ptr[j];
        __syncthreads();
        if (i < n_rows) {
            const int di = s_deg[i];
            int rank = 0;
            for (int j = 0; j < n_rows; ++j) {
                const int dj = s_deg[j];
                rank += (dj > di || (dj == di && j < i)) ? 1 : 0;
            }
            order[rank] = i;
        }
    }
    if (i == 0) {
        const int f = *num_found;
        status[0] = f;
        status[1] = f > max_edges ? 1 : 0;
    }
    if (i >= n) return;
    if (i > 0 && batch[i] < batch[i - 1]) status[3] = 1;   // molecule ranges are found by binary search: batch must be sorted
    long long v = z[i];
    if (v < 0 || v >= max_z) {
        status[2] = 1;
        v = 0;
    }
    z32[i] = (int)v;
}

// ------------------------------------------------------------------ exact-mode filter MLP helpers
// R[e,k] = rbf_k(d_e), dR = d/dd  (models/utils.py:298-300, 339-344); rows e >= E are zero-filled.
__global__ void tn_rbf_kernel(const float4* __restrict__ geom, const int* __restrict__ row_ptr, int n_atoms,
                              long long e_cap, int K, int rbf_type, float lo, float hi, const float* __restrict__ p0,
                              const float* __restrict__ p1, float* __restrict__ R, float* __restrict__ dR) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= e_cap * K) return;
    const long long e = idx / K;
    const int k = (int)(idx % K);
    float r = 0.f, dr = 0.f;
    if (e < row_ptr[n_atoms]) {
        const float d = geom[e].w;
        if (rbf_type == TMDN_RBF_EXPNORM) {
            const float alpha = 5.0f / (hi - lo);
            float c0, dc0;
            cutoff_fn(d, 0.f, hi, c0, dc0);
            const float u = expf(alpha * (lo - d));
            const float mu = p0[k], beta = p1[k];
            const float g = expf(-beta * (u - mu) * (u - mu));
            r = c0 * g;
            dr = dc0 * g + c0 * g * (2.0f * alpha) * beta * u * (u - mu);
        } else {
            const float off = p0[k], coeff = p1[k];
            const float g = expf(coeff * (d - off) * (d - off));
            r = g;
            dr = 2.0f * coeff * (d - off) * g;
        }
    }
    R[idx] = r;
    dR[idx] = dr;
}
// F = a*C(d), dF = t*C(d) + a*C'(d)    (in place on a/t)
__global__ void tn_filter_finalize_kernel(const float4* __restrict__ geom, long long e_cap, int n3, float lo, float hi,
                                          float* __restrict__ a, float* __restrict__ t) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= e_cap * n3) return;
    const long long e = idx / n3;
    float c, dc;
    cutoff_fn(geom[e].w, lo, hi, c, dc);
    const float av = a[idx], tv = t[idx];
    a[idx] = av * c;
    t[idx] = tv * c + av * dc;
}

// ------------------------------------------------------------------ E1: embedding edge kernel (+ init LayerNorm)
// ias_c[i,h] = sum_{e in row i} (P[z_i]+Q[z_j])[h] * W_g(c)(d_e)[h] * geom_c(u_e)      (tensornet.py:295-315)
// nh = LayerNorm_H(||I+A+S||^2)                                                     (tensornet.py:316)
// HC > 0: hidden width known at compile time -> the 9 component rows of a node tensor are immediate offsets from one
// base pointer (saves ~40 address instructions per edge; the kernels are issue-bound once the gathers hit L2)
// partial sums of row i over the edges r0+sl, r0+sl+S, ... (one slice of the row)
template <bool TAB, bool HOT>
__device__ __forceinline__ void embed_fwd_row(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                              const float4* __restrict__ geom, const int* __restrict__ z32,
                                              const float* __restrict__ Pz, const float* __restrict__ Qz,
                                              const FilterView& fv, int i, int sl, int S, int H, int h, float acc[9]) {
    const float pz = Pz[(size_t)z32[i] * H + h];
    const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
    // the index, geometry and element of the NEXT edge of this slice are fetched while the gathers of the current
    // one are in flight: one dependent round trip per edge instead of three (col -> z32 -> Q row)
    int e = r0 + sl, jn = 0, zn = 0;
    float4 gn = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e < r1) { jn = col[e]; gn = geom[e]; zn = z32[jn]; }
    for (; e < r1; e += S) {
        const float4 g = gn;
        const int zj = zn, e2 = e + S;
        if (e2 < r1) { jn = col[e2]; gn = geom[e2]; }
        float f[3];
        filter_val<TAB, HOT>(fv, (size_t)e, g.w, H, h, f);
        const float zv = pz + Qz[(size_t)zj * H + h];
        if (e2 < r1) zn = z32[jn];
        const float w1 = zv * f[0], w2 = zv * f[1], w3 = zv * f[2];
        const float n2 = (g.x * g.x + g.y * g.y + g.z * g.z) * (1.0f / 3.0f);
        acc[0] += w1;
        acc[1] += w2 * g.x; acc[2] += w2 * g.y; acc[3] += w2 * g.z;
        acc[4] += w3 * (g.x * g.x - n2);
        acc[5] += w3 * (g.y * g.y - n2);
        acc[6] += w3 * (g.x * g.y);
        acc[7] += w3 * (g.x * g.z);
        acc[8] += w3 * (g.y * g.z);
    }
}

template <bool TAB, int HC = 0>
__global__ void __launch_bounds__(1024) tn_embed_edge_fwd(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                  const float4* __restrict__ geom, const int* __restrict__ z32,
                                  const float* __restrict__ Pz, const float* __restrict__ Qz, FilterView fv, int N,
                                  int H_, float* __restrict__ ias, const float* __restrict__ ln_w,
                                  const float* __restrict__ ln_b, float* __restrict__ nh,
                                  const int* __restrict__ order, int row0) {
    pdl_sync();
    const int H = HC > 0 ? HC : H_;
    extern __shared__ float smem[];
    float* s_red = smem;
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    bool valid;
    const int i = edge_row(order, row0, N, valid);
    float acc[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) acc[c] = 0.f;
    if (valid) embed_fwd_row<TAB, false>(row_ptr, col, geom, z32, Pz, Qz, fv, i, sl, S, H, h, acc);
    slice_reduce<9>(acc, s_part, H);
    const bool lead = sl == 0;
    if (valid && lead) st9(ias, (size_t)i, H, h, acc);
    const float q = lead ? t3_q(acc) : 0.f;
    float s[1] = {q};
    atom_reduce<1>(s, s_red);
    const float mean = s[0] / (float)H;
    float v[1] = {lead ? (q - mean) * (q - mean) : 0.f};
    atom_reduce<1>(v, s_red);
    const float rstd = rsqrtf(v[0] / (float)H + 1e-5f);
    if (valid && lead) nh[(size_t)i * H + h] = (q - mean) * rstd * ln_w[h] + ln_b[h];
}

// ---- persistent large-system variants of the message-passing kernels (table mode) -------------------------------
// One CTA per SM; a CTA is W "row workers" of H threads (threadIdx.y = worker), worker w of CTA b takes the rows
// b*W + w + k*gridDim.x*W, so the rows in flight on the chip at any time are one contiguous window of the
// (spatially ordered) atom index, exactly as with one-CTA-per-row launches: neighbouring rows gather the same
// source rows and L2 serves them.  Measured at 98 304 atoms (gpurun r02a): tn_msg_fwd 2.21 -> 2.00 ms, tn_msg_bwd
// 4.27 -> 3.53 ms per launch against one CTA per row; the embedding edge kernels did not gain and keep that form.
// Optional (tuning knob "hot_rows", default 0): the first n_hot rows of the filter table -- the far end of the grid
// of table_grid.cuh, where most neighbours of a condensed system are -- are copied once per CTA into shared memory.
// Measured: no gain for 0 / 16 / 24 / 32 rows (23.72 / 23.68 / 23.58 / 23.61 ms per step), because shared memory and
// L1 share one data pipe and that pipe (l1tex 71-75 % busy, L2 only 40 %), not L2, is what bounds these kernels
// once the table is 6 MB.  Sums run in the same order as in the one-row-per-CTA kernels: results are bit-identical.
__device__ __forceinline__ void hot_rows_load(FilterView& fv, float4* s_hot, int n_hot, int H) {
    fv.n_hot = n_hot;
    fv.hot = s_hot;
    const float4* src = fv.tab;
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, nt = blockDim.x * blockDim.y;
    for (int q = tid; q < n_hot * 3 * H; q += nt) s_hot[q] = __ldg(&src[q]);
    __syncthreads();
}
// sum over the H threads of one worker (named barrier 1 + worker id); every thread of the worker must call
template <int NV> __device__ __forceinline__ void worker_reduce(float (&v)[NV], float* s_red_w, int H) {
    const int lane = threadIdx.x & 31, wr = threadIdx.x >> 5, nw = H >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = warp_sum(v[k]);
    asm volatile("bar.sync %0, %1;" ::"r"(1 + (int)threadIdx.y), "r"(H) : "memory");
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < NV; ++k) s_red_w[wr * NV + k] = v[k];
    asm volatile("bar.sync %0, %1;" ::"r"(1 + (int)threadIdx.y), "r"(H) : "memory");
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        float t = 0.f;
        for (int w = 0; w < nw; ++w) t += s_red_w[w * NV + k];
        v[k] = t;
    }
}
constexpr int WORKER_RED_FLOATS = 32;   // per-worker scratch of worker_reduce (<= 8 warps x 3 values)

// ------------------------------------------------------------------ per-(atom,channel) elementwise stages
__global__ void tn_normalize_fwd(const float* __restrict__ X, float* __restrict__ Xh, long long total, int H) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    float x[9], xh[9];
    ld9(X, i, H, h, x);
    t3_normalize(x, xh);
    st9(Xh, i, H, h, xh);
}
// Xn = Xh + dX + (dX)^2 ; optionally also the next layer's normalised input   (tensornet.py:409, :385)
__global__ void tn_update_fwd(const float* Xh, const float* __restrict__ dX, float* __restrict__ Xn,
                              float* Xh_next, long long total, int H) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    float xh[9], d[9], sq[9], xn[9];
    ld9(Xh, i, H, h, xh);
    ld9(dX, i, H, h, d);
    t3_square(d, sq);
#pragma unroll
    for (int c = 0; c < 9; ++c) xn[c] = xh[c] + d[c] + sq[c];
    st9(Xn, i, H, h, xn);
    if (Xh_next) {
        float nx[9];
        t3_normalize(xn, nx);
        st9(Xh_next, i, H, h, nx);
    }
}
// adjoint of the normalisation followed by the adjoint of the update of the layer below (ddX = dXn + adjoint of the square
// term), one pass over the (atom, channel) pairs: the same
// operations in the same order (dXa is still written: it is the residual branch of the next stage)
__global__ void tn_normalize_update_bwd(const float* __restrict__ X, float* __restrict__ dXa,
                                        const float* __restrict__ dXprev, float* __restrict__ ddX, long long total, int H) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    float x[9], dxh[9], g[9], d[9], dd[9];
    ld9(X, i, H, h, x);
    ld9(dXa, i, H, h, dxh);
    t3_normalize_bwd(x, dxh, g);
    st9(dXa, i, H, h, g);
    ld9(dXprev, i, H, h, d);
    t3_square_bwd(d, g, dd);
#pragma unroll
    for (int c = 0; c < 9; ++c) dd[c] += g[c];
    st9(ddX, i, H, h, dd);
}

// recompute tau = product(M,Y); dtau from dtauh; then dM -> dMout, dY(direct) -> dtauh buffer (in place)
__global__ void tn_prod_bwd(const float* __restrict__ Y, const float* __restrict__ M, float* __restrict__ dtauh_dY,
                            float* __restrict__ dMout, long long total, int H, int so3) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    float y[9], m[9], tau[9], dth[9], dtau[9], dm[9], dy[9];
    ld9(Y, i, H, h, y);
    ld9(M, i, H, h, m);
    ld9(dtauh_dY, i, H, h, dth);
    t3_product(m, y, so3 != 0, tau);
    t3_normalize_bwd(tau, dth, dtau);
    t3_product_bwd(m, y, so3 != 0, dtau, dm, dy);
    st9(dMout, i, H, h, dm);
    st9(dtauh_dY, i, H, h, dy);
}

// NB = edges whose loads are issued together (independent register sets): the kernel is bound by the latency of the
// gathers, and the compiler does not overlap the loads of consecutive iterations on its own.
template <bool TAB> struct FilterRaw { float4 c[3]; float s, ds; };
template <> struct FilterRaw<false> { float f[3], df[3]; };
template <bool TAB, bool HOT = false>
__device__ __forceinline__ void filter_load(const FilterView& fv, size_t e, float d, int H, int h, FilterRaw<TAB>& r) {
    if constexpr (TAB) {
        int k;
        tab_locate(fv.grid, d, k, r.s, r.ds);
#pragma unroll
        for (int g = 0; g < 3; ++g) r.c[g] = filter_row<HOT>(fv, k, g, H, h);
    } else {
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            r.f[g] = fv.F[(e * 3 + g) * H + h];
            r.df[g] = fv.dF[(e * 3 + g) * H + h];
        }
    }
}
template <bool TAB> __device__ __forceinline__ void filter_eval_val(const FilterRaw<TAB>& r, float f[3]) {
    if constexpr (TAB) {
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const float4 c = r.c[g];
            f[g] = c.x + r.s * (c.y + r.s * (c.z + r.s * c.w));
        }
    } else {
#pragma unroll
        for (int g = 0; g < 3; ++g) f[g] = r.f[g];
    }
}
template <bool TAB>
__device__ __forceinline__ void filter_eval(const FilterRaw<TAB>& r, float f[3], float df[3]) {
    if constexpr (TAB) {
        const float s2 = 2.0f * r.s, s3 = 3.0f * r.s * r.s;   // df/ds = c.y + 2 s c.z + 3 s^2 c.w as two FFMA per filter
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const float4 c = r.c[g];
            f[g] = c.x + r.s * (c.y + r.s * (c.z + r.s * c.w));
            df[g] = fmaf(s3, c.w, fmaf(s2, c.z, c.y)) * r.ds;
        }
    } else {
#pragma unroll
        for (int g = 0; g < 3; ++g) { f[g] = r.f[g]; df[g] = r.df[g]; }
    }
}

// ------------------------------------------------------------------ E2: interaction message passing (+ product)
// M_c[i,h] = sum_{e in row i} f_g(c)(d_e)[h] * Y_c[j,h]  ;  tauh = normalise(decompose(M Y + Y M))   (tensornet.py:392-405)
constexpr int MSG_FWD_THREADS = 1024;   // 64 registers, one edge per batch: at 98k atoms this runs at 83 % of L2 throughput
// PF: index and distance of the next batch of edges are fetched while the gathers of the current one are in flight
// (small systems: the kernel is a chain of dependent round trips, col -> neighbour row)
// partial sums m[9] of row i over the edges r0+sl, r0+sl+S, ...
template <bool TAB, int NB, bool PF, bool HOT>
__device__ __forceinline__ void msg_fwd_row(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                            const float4* __restrict__ geom, const FilterView& fv,
                                            const float* __restrict__ Y, int i, int sl, int S, int H, int h, float m[9]) {
    const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
    int jn[NB];
    float dn[NB];
    if (PF && r0 + sl < r1) {
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const int e = r0 + sl + b * S, ee = e < r1 ? e : r0 + sl;
            jn[b] = col[ee];
            dn[b] = geom[ee].w;
        }
    }
    for (int e0 = r0 + sl; e0 < r1; e0 += NB * S) {
        FilterRaw<TAB> fr[NB];
        float y[NB][9];
        bool on[NB];
        int jc[NB];
        float dc[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) { jc[b] = jn[b]; dc[b] = dn[b]; }
        if (PF && e0 + NB * S < r1) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const int e = e0 + NB * S + b * S, ee = e < r1 ? e : e0 + NB * S;
                jn[b] = col[ee];
                dn[b] = geom[ee].w;
            }
        }
#pragma unroll
        for (int b = 0; b < NB; ++b) {           // all loads of the batch first
            const int e = e0 + b * S;
            on[b] = e < r1;
            const int ee = on[b] ? e : e0;
            const int j = PF ? jc[b] : col[ee];
            filter_load<TAB, HOT>(fv, (size_t)ee, PF ? dc[b] : geom[ee].w, H, h, fr[b]);
            ld9(Y, (size_t)j, H, h, y[b]);
        }
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            if (!on[b]) continue;
            float f[3];
            filter_eval_val<TAB>(fr[b], f);
            m[0] += f[0] * y[b][0];
            m[1] += f[1] * y[b][1]; m[2] += f[1] * y[b][2]; m[3] += f[1] * y[b][3];
            m[4] += f[2] * y[b][4]; m[5] += f[2] * y[b][5]; m[6] += f[2] * y[b][6]; m[7] += f[2] * y[b][7];
            m[8] += f[2] * y[b][8];
        }
    }
}
// M -> HBM, tauh = normalise(decompose(M Y + Y M))
__device__ __forceinline__ void msg_fwd_finish(const float* __restrict__ Y, int i, int H, int h, int so3, const float m[9],
                                               float* __restrict__ M, float* __restrict__ tauh) {
    st9(M, (size_t)i, H, h, m);
    float yi[9], tau[9], th[9];
    ld9(Y, (size_t)i, H, h, yi);
    t3_product(m, yi, so3 != 0, tau);
    t3_normalize(tau, th);
    st9(tauh, (size_t)i, H, h, th);
}

template <bool TAB, int NB = 1, int HC = 0, bool PF = false, int MAXT = MSG_FWD_THREADS>
__global__ void __launch_bounds__(MAXT) tn_msg_fwd(const int* __restrict__ row_ptr, const int* __restrict__ col,
                           const float4* __restrict__ geom, FilterView fv, const float* __restrict__ Y, int N, int H_,
                           int so3, float* __restrict__ M, float* __restrict__ tauh, const int* __restrict__ order, int row0) {
    pdl_sync();
    const int H = HC > 0 ? HC : H_;
    extern __shared__ float smem[];
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    // blockIdx.y > 0 only in the small-system launch, where a CTA takes a 32-channel slab of a row (finer work units:
    // 166 rows on 148 SMs are two full waves of whole-row CTAs, 664 slabs are 4.5 waves of quarter-length ones)
    const int h = blockIdx.y * blockDim.x + threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    bool valid;
    const int i = edge_row(order, row0, N, valid);
    float m[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) m[c] = 0.f;
    if (valid) msg_fwd_row<TAB, NB, PF, false>(row_ptr, col, geom, fv, Y, i, sl, S, H, h, m);
    slice_reduce<9>(m, s_part, blockDim.x);
    if (!valid || sl != 0) return;
    msg_fwd_finish(Y, i, H, h, so3, m, M, tauh);
}

// persistent large-system variant: W row workers
template <int HC, int NB, int MAXT = 1024, int MINB = 1, bool HOT = true>
__global__ void __launch_bounds__(MAXT, MINB) tn_msg_fwd_p(const int* __restrict__ row_ptr, const int* __restrict__ col,
                             const float4* __restrict__ geom, FilterView fv, const float* __restrict__ Y, int N,
                             int so3, float* __restrict__ M, float* __restrict__ tauh, int n_hot, int row0) {
    pdl_sync();
    constexpr int H = HC;
    extern __shared__ float4 s_hot[];
    if (HOT) hot_rows_load(fv, s_hot, n_hot, H);       // HOT = false (no hot rows requested): the loop carries no shared-memory branch
    const int h = threadIdx.x, W = blockDim.y;
    // a CTA takes blocks of FWD_ROW_BLOCK * W consecutive rows and its workers walk a block with stride W: the next rows of
    // a worker are neighbours of the ones it just did and find their gathers in L1 (water box 3.72 -> 3.62 ms, MD17 batch
    // 3.59 -> 3.37 ms for the two launches; 16 is slower again, and the adjoint kernel loses with any blocking, r03p)
    constexpr int FWD_ROW_BLOCK = 4;
    for (int i0 = row0 + blockIdx.x * (FWD_ROW_BLOCK * W); i0 < N; i0 += gridDim.x * (FWD_ROW_BLOCK * W))
    for (int i = i0 + threadIdx.y; i < min(N, i0 + FWD_ROW_BLOCK * W); i += W) {
        float m[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) m[c] = 0.f;
        msg_fwd_row<true, NB, false, HOT>(row_ptr, col, geom, fv, Y, i, 0, 1, H, h, m);
        msg_fwd_finish(Y, i, H, h, so3, m, M, tauh);
    }
}

// E3: adjoint of the message passing, evaluated in row i of the symmetric edge set:
//   dY_i += sum_e f_e * dM_j
//   dE/dpos_i += u_e * sum_{h,c} f'_g(c)(d_e)[h] * (dM_c[i,h] Y_c[j,h] + dM_c[j,h] Y_c[i,h])
constexpr int MSG_BWD_THREADS = 256;   // 128 registers: two edges' gathers (60 loads) in flight per thread; small systems use 4 slices
// MAXT = 512 (4 edge slices per row) is the small-system instantiation: same registers, more rows' worth of
// gathers in flight when there are fewer rows than SMs can hold
// partial sums of row i over one slice of its edges: acc[9] += f dM_j (acc enters holding dY_i for slice 0), gp[3]
template <bool TAB, int NB, bool PF, bool HOT>
__device__ __forceinline__ void msg_bwd_row(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                            const float4* __restrict__ geom, const FilterView& fv,
                                            const float* __restrict__ Y, const float* __restrict__ dM,
                                            const float* dY /* also written by the caller */, int i, int sl,
                                            int S, int H, int h, float acc[9], float gp[3]) {
    float yi[9], dmi[9];
    ld9(Y, (size_t)i, H, h, yi);
    ld9(dM, (size_t)i, H, h, dmi);
    if (sl == 0) ld9(dY, (size_t)i, H, h, acc);
    const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
    int jn[NB];
    float dn[NB];
    if (PF && r0 + sl < r1) {
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const int e = r0 + sl + b * S, ee = e < r1 ? e : r0 + sl;
            jn[b] = col[ee];
            dn[b] = geom[ee].w;
        }
    }
    for (int e0 = r0 + sl; e0 < r1; e0 += NB * S) {
        float4 g[NB];
        FilterRaw<TAB> fr[NB];
        float yj[NB][9], dmj[NB][9];
        bool on[NB];
        int jc[NB];
        float dc[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) { jc[b] = jn[b]; dc[b] = dn[b]; }
        if (PF && e0 + NB * S < r1) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const int e = e0 + NB * S + b * S, ee = e < r1 ? e : e0 + NB * S;
                jn[b] = col[ee];
                dn[b] = geom[ee].w;
            }
        }
#pragma unroll
        for (int b = 0; b < NB; ++b) {           // all loads of the batch first
            const int e = e0 + b * S;
            on[b] = e < r1;
            const int ee = on[b] ? e : e0;
            const int j = PF ? jc[b] : col[ee];
            g[b] = geom[ee];
            filter_load<TAB, HOT>(fv, (size_t)ee, PF ? dc[b] : g[b].w, H, h, fr[b]);
            ld9(Y, (size_t)j, H, h, yj[b]);
            ld9(dM, (size_t)j, H, h, dmj[b]);
        }
#pragma unroll
        for (int b = 0; b < NB; ++b) {           // then the sums, in edge order
            if (!on[b]) continue;
            float f[3], df[3];
            filter_eval<TAB>(fr[b], f, df);
            acc[0] += f[0] * dmj[b][0];
            acc[1] += f[1] * dmj[b][1]; acc[2] += f[1] * dmj[b][2]; acc[3] += f[1] * dmj[b][3];
            acc[4] += f[2] * dmj[b][4]; acc[5] += f[2] * dmj[b][5]; acc[6] += f[2] * dmj[b][6];
            acc[7] += f[2] * dmj[b][7]; acc[8] += f[2] * dmj[b][8];
            // pure FFMA chains (the loop is issue-bound: a product, an FFMA and an add per component cost half again as much)
            const float s0 = fmaf(dmi[0], yj[b][0], dmj[b][0] * yi[0]);
            float s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int c = 1; c < 4; ++c) s1 = fmaf(dmi[c], yj[b][c], fmaf(dmj[b][c], yi[c], s1));
#pragma unroll
            for (int c = 4; c < 9; ++c) s2 = fmaf(dmi[c], yj[b][c], fmaf(dmj[b][c], yi[c], s2));
            const float ge = df[0] * s0 + df[1] * s1 + df[2] * s2;
            gp[0] += g[b].x * ge; gp[1] += g[b].y * ge; gp[2] += g[b].z * ge;
        }
    }
}

template <bool TAB, int NB = 2, int HC = 0, int MAXT = MSG_BWD_THREADS, bool PF = false>
__global__ void __launch_bounds__(MAXT) tn_msg_bwd(const int* __restrict__ row_ptr, const int* __restrict__ col,
                           const float4* __restrict__ geom, FilterView fv, const float* __restrict__ Y,
                           const float* __restrict__ dM, float* __restrict__ dY, float* __restrict__ gpos, int N,
                           int H_, const int* __restrict__ order, int row0) {
    pdl_sync();
    const int H = HC > 0 ? HC : H_;
    extern __shared__ float smem[];
    float* s_red = smem;
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    bool valid;
    const int i = edge_row(order, row0, N, valid);
    float gp[3] = {0.f, 0.f, 0.f};
    float acc[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) acc[c] = 0.f;
    if (valid) msg_bwd_row<TAB, NB, PF, false>(row_ptr, col, geom, fv, Y, dM, dY, i, sl, S, H, h, acc, gp);
    slice_reduce<9>(acc, s_part, H);
    if (valid && sl == 0) st9(dY, (size_t)i, H, h, acc);
    atom_reduce<3>(gp, s_red);
    if (valid && h == 0 && sl == 0) {
        gpos[3 * (size_t)i] += gp[0];
        gpos[3 * (size_t)i + 1] += gp[1];
        gpos[3 * (size_t)i + 2] += gp[2];
    }
}

// persistent large-system variant: 4 row workers at 128 registers
template <int HC, int NB, int MAXT, bool HOT = true>
__global__ void __launch_bounds__(MAXT, 1) tn_msg_bwd_p(const int* __restrict__ row_ptr, const int* __restrict__ col,
                            const float4* __restrict__ geom, FilterView fv, const float* __restrict__ Y,
                            const float* __restrict__ dM, float* __restrict__ dY, float* __restrict__ gpos, int N,
                            int n_hot, int row0) {
    pdl_sync();
    constexpr int H = HC;
    extern __shared__ float4 s_hot[];
    float* s_red_w = (float*)(s_hot + (size_t)n_hot * 3 * H) + WORKER_RED_FLOATS * threadIdx.y;
    if (HOT) hot_rows_load(fv, s_hot, n_hot, H);
    const int h = threadIdx.x, W = blockDim.y;
    for (int i = row0 + blockIdx.x * W + threadIdx.y; i < N; i += gridDim.x * W) {
        float gp[3] = {0.f, 0.f, 0.f};
        float acc[9];
        msg_bwd_row<true, NB, false, HOT>(row_ptr, col, geom, fv, Y, dM, dY, i, 0, 1, H, h, acc, gp);
        st9(dY, (size_t)i, H, h, acc);
        worker_reduce<3>(gp, s_red_w, H);
        if (h == 0) {
            gpos[3 * (size_t)i] += gp[0];
            gpos[3 * (size_t)i + 1] += gp[1];
            gpos[3 * (size_t)i + 2] += gp[2];
        }
    }
}

// ------------------------------------------------------------------ output norm / head
// x = [||I||^2, ||A||^2, ||S||^2] -> LayerNorm(3H)      (tensornet.py:230-232)
__global__ void tn_out_norm_fwd(const float* __restrict__ X, const float* __restrict__ ln_w,
                                const float* __restrict__ ln_b, float* __restrict__ xl, float* __restrict__ stat,
                                int N, int H) {
    pdl_sync();
    __shared__ float s_red[64];
    const int h = threadIdx.x;
    const int i = blockIdx.x * blockDim.y + threadIdx.y;
    const bool valid = i < N;
    float q[3] = {0.f, 0.f, 0.f};
    if (valid) {
        float x[9];
        ld9(X, (size_t)i, H, h, x);
        t3_qparts(x, q[0], q[1], q[2]);
    }
    float s[1] = {q[0] + q[1] + q[2]};
    row_reduce<1>(s, s_red);
    const float mean = s[0] / (3.0f * H);
    float v[1] = {(q[0] - mean) * (q[0] - mean) + (q[1] - mean) * (q[1] - mean) + (q[2] - mean) * (q[2] - mean)};
    row_reduce<1>(v, s_red);
    const float rstd = rsqrtf(v[0] / (3.0f * H) + 1e-5f);
    if (valid) {
#pragma unroll
        for (int g = 0; g < 3; ++g)
            xl[(size_t)i * 3 * H + g * H + h] = (q[g] - mean) * rstd * ln_w[g * H + h] + ln_b[g * H + h];
        if (h == 0) { stat[2 * (size_t)i] = mean; stat[2 * (size_t)i + 1] = rstd; }
    }
}
__global__ void tn_out_norm_bwd(const float* __restrict__ X, const float* __restrict__ ln_w,
                                const float* __restrict__ stat, const float* __restrict__ dxl,
                                float* __restrict__ dXa, int N, int H, const float* __restrict__ dXprev,
                                float* __restrict__ ddX) {
    pdl_sync();
    __shared__ float s_red[64];
    const int h = threadIdx.x;
    const int i = blockIdx.x * blockDim.y + threadIdx.y;
    const bool valid = i < N;
    float x[9], q[3] = {0.f, 0.f, 0.f}, xh[3] = {0.f, 0.f, 0.f}, dg[3] = {0.f, 0.f, 0.f};
    float mean = 0.f, rstd = 0.f;
#pragma unroll
    for (int c = 0; c < 9; ++c) x[c] = 0.f;
    if (valid) {
        ld9(X, (size_t)i, H, h, x);
        t3_qparts(x, q[0], q[1], q[2]);
        mean = stat[2 * (size_t)i];
        rstd = stat[2 * (size_t)i + 1];
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            xh[g] = (q[g] - mean) * rstd;
            dg[g] = dxl[(size_t)i * 3 * H + g * H + h] * ln_w[g * H + h];
        }
    }
    float r[2] = {dg[0] + dg[1] + dg[2], dg[0] * xh[0] + dg[1] * xh[1] + dg[2] * xh[2]};
    row_reduce<2>(r, s_red);
    if (!valid) return;
    const float m1 = r[0] / (3.0f * H), m2 = r[1] / (3.0f * H);
    float dq[3];
#pragma unroll
    for (int g = 0; g < 3; ++g) dq[g] = (dg[g] - m1 - xh[g] * m2) * rstd;
    float gq[9], out[9];
    t3_qgrad(x, gq);
    out[0] = gq[0] * dq[0];
#pragma unroll
    for (int c = 1; c < 4; ++c) out[c] = gq[c] * dq[1];
#pragma unroll
    for (int c = 4; c < 9; ++c) out[c] = gq[c] * dq[2];
    st9(dXa, (size_t)i, H, h, out);
    if (dXprev != nullptr) {   // adjoint of the last layer's update, fused (same operations in the same order)
        float d[9], dd[9];
        ld9(dXprev, (size_t)i, H, h, d);
        t3_square_bwd(d, out, dd);
#pragma unroll
        for (int c = 0; c < 9; ++c) dd[c] += out[c];
        st9(ddX, (size_t)i, H, h, dd);
    }
}
// e_i = (w2 . a4_i + b2) * std      (output_modules.py:76-77, model.py:262-263); one warp per atom
// d4 != null: also the seed of the backward pass, d4[i,n] = std * w2[n] * act'(h4[i,n])
__global__ void tn_head_energy(const float* __restrict__ a4, const float* __restrict__ w2,
                               const float* __restrict__ b2, float stdv, float* __restrict__ e_atom, int N, int Hh,
                               const float* __restrict__ h4, int act, float* __restrict__ d4) {
    pdl_sync();
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= N) return;
    float s = 0.f;
    for (int k = lane; k < Hh; k += 32) s += a4[(size_t)i * Hh + k] * w2[k];
    s = warp_sum(s);
    if (lane == 0) e_atom[i] = (s + b2[0]) * stdv;
    if (d4 != nullptr)
        for (int k = lane; k < Hh; k += 32) d4[(size_t)i * Hh + k] = stdv * w2[k] * act_grad(act, h4[(size_t)i * Hh + k]);
}
// y[m] = sum_{i: batch[i]==m} e_i + mean  (batch sorted); one CTA per molecule, fixed-order tree => deterministic
// status[1] (edge capacity exceeded -> truncated, asymmetric edge set) or status[2] (z out of range) poison the
// result with NaN on the device: the analytic backward assumes a symmetric edge set, so forces of a truncated list
// would not be -dE/dpos of the returned energy, and nothing on a sync-free path reads the flags.
__global__ void __launch_bounds__(128) tn_mol_reduce(const float* __restrict__ e_atom, const int64_t* __restrict__ batch,
                                                     int N, float mean, float* __restrict__ energy,
                                                     const int32_t* __restrict__ status) {
    pdl_sync();
    __shared__ float s_part[128];
    const long long m = blockIdx.x;
    int a, b;
    block_molecule_range(batch, N, m, a, b);
    float s = 0.f;
    for (int i = a + threadIdx.x; i < b; i += 128) s += e_atom[i];
    s_part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (threadIdx.x < o) s_part[threadIdx.x] += s_part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) energy[m] = (status[1] | status[2] | status[3]) ? __int_as_float(0x7fc00000) : s_part[0] + mean;
}
// ------------------------------------------------------------------ embedding backward (node part)
// X0_c = fias_g(c) * iasm_c :  dh2[g*H+h] = (sum_{c in g} dX_c iasm_c) * act'(h2) ;  diasm_c = dX_c * fias_g  (in place)
__global__ void tn_embed_scale_bwd(float* __restrict__ dXa, const float* __restrict__ iasm,
                                   const float* __restrict__ fias, const float* __restrict__ h2, int act,
                                   float* __restrict__ dh2, long long total, int H, const float* __restrict__ Xn) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    float dx[9], im[9], f[3];
    ld9(dXa, i, H, h, dx);
    if (Xn != nullptr) {   // tn_normalize_bwd of layer 0 fused in front (same operations)
        float x[9], g[9];
        ld9(Xn, i, H, h, x);
        t3_normalize_bwd(x, dx, g);
#pragma unroll
        for (int c = 0; c < 9; ++c) dx[c] = g[c];
    }
    ld9(iasm, i, H, h, im);
#pragma unroll
    for (int g = 0; g < 3; ++g) f[g] = fias[i * 3 * H + g * H + h];
    float s0 = dx[0] * im[0], s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 1; c < 4; ++c) s1 += dx[c] * im[c];
#pragma unroll
    for (int c = 4; c < 9; ++c) s2 += dx[c] * im[c];
    dh2[i * 3 * H + h] = s0 * act_grad(act, h2[i * 3 * H + h]);
    dh2[i * 3 * H + H + h] = s1 * act_grad(act, h2[i * 3 * H + H + h]);
    dh2[i * 3 * H + 2 * H + h] = s2 * act_grad(act, h2[i * 3 * H + 2 * H + h]);
    dx[0] *= f[0];
#pragma unroll
    for (int c = 1; c < 4; ++c) dx[c] *= f[1];
#pragma unroll
    for (int c = 4; c < 9; ++c) dx[c] *= f[2];
    st9(dXa, i, H, h, dx);
}
// Dias_c += LayerNorm^T(dnh)[h] * dq/dias_c
__global__ void tn_embed_norm_bwd(const float* __restrict__ ias, const float* __restrict__ ln_w,
                                  const float* __restrict__ dnh, float* __restrict__ Dias, int N, int H) {
    pdl_sync();
    __shared__ float s_red[64];
    const int h = threadIdx.x;
    const int i = blockIdx.x * blockDim.y + threadIdx.y;
    const bool valid = i < N;
    float x[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) x[c] = 0.f;
    if (valid) ld9(ias, (size_t)i, H, h, x);
    const float q = t3_q(x);
    float s[1] = {q};
    row_reduce<1>(s, s_red);
    const float mean = s[0] / (float)H;
    float v[1] = {(q - mean) * (q - mean)};
    row_reduce<1>(v, s_red);
    const float rstd = rsqrtf(v[0] / (float)H + 1e-5f);
    const float xh = (q - mean) * rstd;
    const float dg = valid ? dnh[(size_t)i * H + h] * ln_w[h] : 0.f;
    float r[2] = {dg, dg * xh};
    row_reduce<2>(r, s_red);
    if (!valid) return;
    const float dq = (dg - r[0] / (float)H - xh * r[1] / (float)H) * rstd;
    float gq[9], d[9];
    t3_qgrad(x, gq);
    ld9(Dias, (size_t)i, H, h, d);
#pragma unroll
    for (int c = 0; c < 9; ++c) d[c] += dq * gq[c];
    st9(Dias, (size_t)i, H, h, d);
}

// E4: adjoint of the embedding edge kernel w.r.t. positions, both directions of every edge inside row i.
//   forces[i] = -(gpos[i] + sum_e [ u (gd_e + gd_e') + perp(gu_e) - perp(gu_e') ])
template <bool TAB, bool HOT>
__device__ __forceinline__ void embed_bwd_row(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                              const float4* __restrict__ geom, const int* __restrict__ z32,
                                              const float* __restrict__ Pz, const float* __restrict__ Qz,
                                              const FilterView& fv, const float* __restrict__ Dias, int i, int sl,
                                              int S, int H, int h, float gp[3]) {
    float di[9];
    ld9(Dias, (size_t)i, H, h, di);
    const int zi = z32[i];
    const float pzi = Pz[(size_t)zi * H + h], qzi = Qz[(size_t)zi * H + h];
    const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
    int e = r0 + sl, jn = 0, zn = 0;     // next edge's index / geometry / element prefetched, as in the forward
    float4 gn = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e < r1) { jn = col[e]; gn = geom[e]; zn = z32[jn]; }
    for (; e < r1; e += S) {
        const int j = jn, zj = zn, e2 = e + S;
        const float4 g = gn;
        if (e2 < r1) { jn = col[e2]; gn = geom[e2]; }
        if (j == i) {          // self loop: d = 0, u = 0, no position dependence
            if (e2 < r1) zn = z32[jn];
            continue;
        }
        float f[3], df[3], dj[9];
        filter_val_grad<TAB, HOT>(fv, (size_t)e, g.w, H, h, f, df);
        ld9(Dias, (size_t)j, H, h, dj);
        const float ze = pzi + Qz[(size_t)zj * H + h];        // Z of edge (i <- j)
        const float zr = Pz[(size_t)zj * H + h] + qzi;        // Z of the reverse edge (j <- i)
        const float ux = g.x, uy = g.y, uz = g.z;
        const float n2 = (ux * ux + uy * uy + uz * uz) * (1.0f / 3.0f);
        const float g4 = ux * ux - n2, g5 = uy * uy - n2, g6 = ux * uy, g7 = ux * uz, g8 = uy * uz;
        // a_c = Dias_c[i]*Z_e ; r_c = Dias_c[j]*Z_r ; geom of the reverse edge: odd components flip sign.
        // p_c = a_c - r_c (c = 1..3) or a_c + r_c (else) is all that is needed of them: one product + one FFMA each
        float p[9];
        const float nzr = -zr;
        p[0] = fmaf(di[0], ze, dj[0] * zr);
#pragma unroll
        for (int c = 1; c < 4; ++c) p[c] = fmaf(di[c], ze, dj[c] * nzr);
#pragma unroll
        for (int c = 4; c < 9; ++c) p[c] = fmaf(di[c], ze, dj[c] * zr);
        const float t1 = fmaf(p[1], ux, fmaf(p[2], uy, p[3] * uz));
        const float t2 = fmaf(p[4], g4, fmaf(p[5], g5, fmaf(p[6], g6, fmaf(p[7], g7, p[8] * g8))));
        const float gd = fmaf(df[0], p[0], fmaf(df[1], t1, df[2] * t2));
        // dL/du of edge e minus dL/du' of the reverse edge (d geom_c/du is even for c=1..3, odd for c=4..8):
        //   gu_e - gu_r  with gu_r evaluated at u' = -u  ==  sum_c (a_c*f - r_c*f*sign_c) dgeom_c(u)
        float w[9];
#pragma unroll
        for (int c = 1; c < 4; ++c) w[c] = p[c] * f[1];
#pragma unroll
        for (int c = 4; c < 9; ++c) w[c] = p[c] * f[2];
        const float tw = 2.0f / 3.0f;
        const float gx = w[1] + w[4] * (2.0f * ux - tw * ux) - w[5] * tw * ux + w[6] * uy + w[7] * uz;
        const float gy = w[2] - w[4] * tw * uy + w[5] * (2.0f * uy - tw * uy) + w[6] * ux + w[8] * uz;
        const float gz = w[3] - (w[4] + w[5]) * tw * uz + w[7] * ux + w[8] * uy;
        const float dot = gx * ux + gy * uy + gz * uz;
        const float inv_d = __frcp_rn(g.w);      // the correctly rounded reciprocal = 1.0f / d without the division's slow path
        gp[0] += ux * gd + (gx - dot * ux) * inv_d;
        gp[1] += uy * gd + (gy - dot * uy) * inv_d;
        gp[2] += uz * gd + (gz - dot * uz) * inv_d;
        if (e2 < r1) zn = z32[jn];   // second level of the prefetch, after this edge's gathers were issued and used
    }

}

__device__ __forceinline__ void embed_bwd_finish(const float* __restrict__ gpos, float* __restrict__ forces, int i,
                                                 const float gp[3], const int32_t* __restrict__ status) {
    const bool bad = (status[1] | status[2] | status[3]) != 0;    // see tn_mol_reduce
    const float nan = __int_as_float(0x7fc00000);
    forces[3 * (size_t)i] = bad ? nan : -(gpos[3 * (size_t)i] + gp[0]);
    forces[3 * (size_t)i + 1] = bad ? nan : -(gpos[3 * (size_t)i + 1] + gp[1]);
    forces[3 * (size_t)i + 2] = bad ? nan : -(gpos[3 * (size_t)i + 2] + gp[2]);
}

template <bool TAB, int HC = 0>
__global__ void __launch_bounds__(1024) tn_embed_edge_bwd(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                  const float4* __restrict__ geom, const int* __restrict__ z32,
                                  const float* __restrict__ Pz, const float* __restrict__ Qz, FilterView fv,
                                  const float* __restrict__ Dias, const float* __restrict__ gpos,
                                  float* __restrict__ forces, int N, int H_, const int* __restrict__ order, int row0,
                                  const int32_t* __restrict__ status) {
    pdl_sync();
    const int H = HC > 0 ? HC : H_;
    extern __shared__ float smem[];
    float* s_red = smem;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    bool valid;
    const int i = edge_row(order, row0, N, valid);
    float gp[3] = {0.f, 0.f, 0.f};
    if (valid) embed_bwd_row<TAB, false>(row_ptr, col, geom, z32, Pz, Qz, fv, Dias, i, sl, S, H, h, gp);
    atom_reduce<3>(gp, s_red);
    if (valid && h == 0 && sl == 0) embed_bwd_finish(gpos, forces, i, gp, status);
}

// ------------------------------------------------------------------ workspace
struct TnWs {
    // graph
    int32_t *num_found, *row_ptr, *col;
    float* geom;
    int* z32;
    int* row_order;
    void* graph_ws;
    size_t graph_ws_bytes;
    // exact-mode filters
    float *R, *dR, *fh[3], *fa[3], *ft[3];
    float *F[TMDN_MAX_LAYERS + 1], *dF[TMDN_MAX_LAYERS + 1];
    // forward node buffers
    float *ias, *iasm, *nh, *h1, *a1, *h2, *fias, *X[TMDN_MAX_LAYERS + 1], *Xh, *Y[TMDN_MAX_LAYERS],
        *M[TMDN_MAX_LAYERS], *tauh, *dX[TMDN_MAX_LAYERS], *xl, *stat, *h3, *a3, *h4, *a4, *e_atom;
    // backward
    float *dXa, *dB, *dB2, *dC, *d4, *d3, *dxl, *dh2, *dh1, *dnh, *gpos;
};
struct WsEntry { std::string name; size_t off, bytes; };

template <typename T> T* carve(Arena& a, std::vector<WsEntry>* names, const char* name, size_t n) {
    const size_t o = a.off;
    T* p = a.get<T>(n);
    if (names) names->push_back({name, o, n * sizeof(T)});
    return p;
}

int carve_tn(Arena& a, const tmdn_tensornet_model* m, int64_t N, int64_t P, int64_t n_mol, TnWs& w,
             std::vector<WsEntry>* names, const double* box, int strategy, double cutoff) {
    const size_t H = m->hidden, K = m->num_rbf, L = m->num_layers;
    const size_t n = (size_t)N, nh9 = n * 9 * H;
    char buf[32];
    w.num_found = carve<int32_t>(a, names, "num_found", 4);
    w.row_ptr = carve<int32_t>(a, names, "row_ptr", n + 1);
    w.col = carve<int32_t>(a, names, "col", (size_t)P);
    w.geom = carve<float>(a, names, "geom", 4 * (size_t)P);
    w.z32 = carve<int>(a, names, "z32", n);
    w.graph_ws_bytes = tmdn_csr_build_workspace_bytes(strategy, N, box, cutoff);
    if (w.graph_ws_bytes == 0) {
        tmdn_set_error("tensornet: bad box for the cell strategy");
        return TMDN_EINVAL;
    }
    w.graph_ws = carve<char>(a, names, "graph_ws", w.graph_ws_bytes);
    if (m->filter_mode == TMDN_FILTER_EXACT) {
        const size_t e = (size_t)P;
        w.R = carve<float>(a, names, "R", e * K);
        w.dR = carve<float>(a, names, "dR", e * K);
        const size_t dims[3] = {H, 2 * H, 3 * H};
        for (int k = 0; k < 2; ++k) {
            snprintf(buf, sizeof buf, "fh%d", k); w.fh[k] = carve<float>(a, names, buf, e * dims[k]);
            snprintf(buf, sizeof buf, "fa%d", k); w.fa[k] = carve<float>(a, names, buf, e * dims[k]);
            snprintf(buf, sizeof buf, "ft%d", k); w.ft[k] = carve<float>(a, names, buf, e * dims[k]);
        }
        w.fh[2] = carve<float>(a, names, "fh2", e * 3 * H);
        for (size_t s = 0; s <= L; ++s) {
            snprintf(buf, sizeof buf, "F%zu", s); w.F[s] = carve<float>(a, names, buf, e * 3 * H);
            snprintf(buf, sizeof buf, "dF%zu", s); w.dF[s] = carve<float>(a, names, buf, e * 3 * H);
        }
    }
    w.ias = carve<float>(a, names, "ias", nh9);
    w.iasm = carve<float>(a, names, "iasm", nh9);
    w.nh = carve<float>(a, names, "nh", n * H);
    w.h1 = carve<float>(a, names, "h1", n * 2 * H);
    w.a1 = carve<float>(a, names, "a1", n * 2 * H);
    w.h2 = carve<float>(a, names, "h2", n * 3 * H);
    w.fias = carve<float>(a, names, "fias", n * 3 * H);
    for (size_t l = 0; l <= L; ++l) { snprintf(buf, sizeof buf, "X%zu", l); w.X[l] = carve<float>(a, names, buf, nh9); }
    w.Xh = carve<float>(a, names, "Xh", nh9);
    for (size_t l = 0; l < L; ++l) {
        snprintf(buf, sizeof buf, "Y%zu", l); w.Y[l] = carve<float>(a, names, buf, nh9);
        snprintf(buf, sizeof buf, "M%zu", l); w.M[l] = carve<float>(a, names, buf, nh9);
        snprintf(buf, sizeof buf, "dX%zu", l); w.dX[l] = carve<float>(a, names, buf, nh9);
    }
    w.tauh = carve<float>(a, names, "tauh", nh9);
    w.xl = carve<float>(a, names, "xl", n * 3 * H);
    w.stat = carve<float>(a, names, "stat", n * 2);
    w.h3 = carve<float>(a, names, "h3", n * H);
    w.a3 = carve<float>(a, names, "a3", n * H);
    w.h4 = carve<float>(a, names, "h4", n * (H / 2));
    w.a4 = carve<float>(a, names, "a4", n * (H / 2));
    w.e_atom = carve<float>(a, names, "e_atom", n);
    w.dXa = carve<float>(a, names, "dXa", nh9);
    w.dB = carve<float>(a, names, "Dias", nh9);    // backward halo points 0, 2, ... (TnRun::dBk); with an even number
    w.dB2 = carve<float>(a, names, "Dias1", nh9);  // of layers the last content of "Dias" is dE/d ias, else of "Dias1"
    w.dC = carve<float>(a, names, "dC", nh9);
    w.d4 = carve<float>(a, names, "d4", n * (H / 2));
    w.d3 = carve<float>(a, names, "d3", n * H);
    w.dxl = carve<float>(a, names, "dxl", n * 3 * H);
    w.dh2 = carve<float>(a, names, "dh2", n * 3 * H);
    w.dh1 = carve<float>(a, names, "dh1", n * 2 * H);
    w.dnh = carve<float>(a, names, "dnh", n * H);
    w.gpos = carve<float>(a, names, "gpos", n * 3);
    w.row_order = carve<int>(a, names, "row_order", LPT_MAX_ROWS);
    (void)n_mol;
    return TMDN_OK;
}

int check_model(const tmdn_tensornet_model* m) {
    TMDN_CHECK_ARG(m, "tensornet: null model");
    TMDN_CHECK_ARG(m->hidden >= 32 && m->hidden <= 1024 && m->hidden % 32 == 0,
                   "tensornet: hidden_channels must be a multiple of 32 in [32,1024] (got %d)", m->hidden);
    TMDN_CHECK_ARG(m->num_layers >= 0 && m->num_layers <= TMDN_MAX_LAYERS, "tensornet: num_layers out of range");
    TMDN_CHECK_ARG(m->num_rbf > 0 && m->max_z > 0, "tensornet: bad num_rbf/max_z");
    TMDN_CHECK_ARG(m->cutoff_upper > m->cutoff_lower && m->cutoff_lower >= 0, "tensornet: bad cutoffs");
    TMDN_CHECK_ARG(m->filter_mode == TMDN_FILTER_EXACT || m->filter_mode == TMDN_FILTER_TABLE,
                   "tensornet: bad filter_mode");
    if (m->filter_mode == TMDN_FILTER_TABLE) {
        TMDN_CHECK_ARG(m->emb_filter.table && m->emb_filter.n_int > 0, "tensornet: missing embedding filter table");
        for (int l = 0; l < m->num_layers; ++l)
            TMDN_CHECK_ARG(m->layer_filter[l].table && m->layer_filter[l].n_int > 0,
                           "tensornet: missing filter table of layer %d", l);
    }
    return TMDN_OK;
}

// grouped channel mix over the rows of an [N,9,H] tensor: problem g uses weight Wg[g] on components of group g
GemmBatch mix_batch(const float* A, float* C, const float* const Wg[3], int N, int H) {
    GemmBatch b;
    memset(&b, 0, sizeof(b));
    const int gs[3] = {1, 3, 5}, c0[3] = {0, 1, 4};
    for (int g = 0; g < 3; ++g) {
        GemmP p = gemm_desc(A, H, Wg[g], nullptr, C, H, N * gs[g], H, H);
        p.gs = gs[g]; p.c0 = c0[g]; p.rpa = 9; p.g = g;
        b.p[g] = p;
    }
    b.n = 3;
    return b;
}

// tensor cores for H = 128 (tc_mix.cuh); TMDN_TC_MIX=0 keeps the FFMA kernel (A/B measurements)
int mix_launch(int epi, const GemmBatch& b, cudaStream_t stream) {
    static const int use_tc = [] { const char* e = getenv("TMDN_TC_MIX"); return e ? atoi(e) : 1; }();
    // small systems (<= ~1000 atoms): a handful of tensor-core tiles on 13 SMs loses to the latency-oriented FFMA
    // kernel spread over the whole machine (166 atoms: 13 us -> see profiles/r01i)
    if (small_gemm_ok(b)) return gemm_launch(epi, b, stream);
    if (use_tc && (epi == EPI_NONE || epi == EPI_SCALE3) && tcmix::supported(b)) return tcmix::launch(epi, b, stream);
    return gemm_launch(epi, b, stream);
}

// exact-mode evaluation of one filter set on all (capacity) edges -> F, dF [E,3,H]
int eval_filter_exact(const tmdn_tensornet_model* m, const tmdn_filter_set& fs, const TnWs& w, int64_t N, int64_t P,
                      float* F, float* dF, cudaStream_t stream) {
    const int K = m->num_rbf;
    const float* a = w.R;
    const float* t = w.dR;
    for (int k = 0; k < fs.n_layers; ++k) {
        const int in = fs.dims[k], out = fs.dims[k + 1];
        const bool last = k == fs.n_layers - 1;
        const bool has_act = !last || fs.final_act;
        float* ha = last ? F : w.fa[k];
        float* tt = last ? dF : w.ft[k];
        if (has_act) {
            GemmP p = gemm_desc(a, in, fs.w[k], fs.b[k], ha, out, (int)P, out, in);
            p.C2 = w.fh[k]; p.act = m->activation;
            TMDN_TRY(gemm1(EPI_ACT, p, stream, "filter_gemm"));
            GemmP q = gemm_desc(t, in, fs.w[k], nullptr, tt, out, (int)P, out, in);
            q.aux = w.fh[k]; q.ldaux = out; q.act = m->activation;
            TMDN_TRY(gemm1(EPI_MULGRAD, q, stream, "filter_gemm"));
        } else {
            GemmP p = gemm_desc(a, in, fs.w[k], fs.b[k], ha, out, (int)P, out, in);
            TMDN_TRY(gemm1(EPI_NONE, p, stream, "filter_gemm"));
            GemmP q = gemm_desc(t, in, fs.w[k], nullptr, tt, out, (int)P, out, in);
            TMDN_TRY(gemm1(EPI_NONE, q, stream, "filter_gemm"));
        }
        a = ha;
        t = tt;
    }
    (void)K; (void)N;
    const long long tot = (long long)P * fs.dims[fs.n_layers];
    {
        KTimer ktimer__("tn_filter_finalize_kernel", stream);
        launch_k(tn_filter_finalize_kernel, (unsigned)cdiv(tot, 256), 256, 0, stream, (const float4*)w.geom, (long long)P,
                                                                                fs.dims[fs.n_layers], m->cutoff_lower,
                                                                                m->cutoff_upper, F, dF);
        TMDN_LAUNCH_CHECK();
    }
    return TMDN_OK;
}

FilterView make_view(const tmdn_tensornet_model* m, const tmdn_filter_set& fs, const float* F, const float* dF) {
    FilterView v;
    v.F = F; v.dF = dF;
    v.tab = (const float4*)fs.table;
    v.grid = make_tab_grid(fs.n_int, m->cutoff_upper, fs.grid_alpha, fs.grid_gamma);
    v.hot = nullptr;
    v.n_hot = 0;
    return v;
}

}  // namespace

extern "C" size_t tmdn_tensornet_workspace_bytes(const tmdn_tensornet_model* m, int64_t n_atoms, int64_t max_edges,
                                                 int64_t n_mol, int strategy, const double box[9]) {
    if (check_model(m) != TMDN_OK || n_atoms <= 0 || max_edges <= 0) return 0;
    Arena a(nullptr, 0);
    TnWs w;
    if (carve_tn(a, m, n_atoms, max_edges, n_mol, w, nullptr, box, strategy, m->cutoff_upper) != TMDN_OK) return 0;
    return a.off + 256;
}

extern "C" int tmdn_tensornet_workspace_lookup(const tmdn_tensornet_model* m, int64_t n_atoms, int64_t max_edges,
                                               int64_t n_mol, int strategy, const double box[9], const char* name,
                                               size_t* offset, size_t* bytes) {
    TMDN_TRY(check_model(m));
    TMDN_CHECK_ARG(name && offset && bytes, "lookup: null argument");
    Arena a(nullptr, 0);
    TnWs w;
    std::vector<WsEntry> names;
    TMDN_TRY(carve_tn(a, m, n_atoms, max_edges, n_mol, w, &names, box, strategy, m->cutoff_upper));
    for (const auto& e : names)
        if (e.name == name) {
            *offset = e.off;
            *bytes = e.bytes;
            return TMDN_OK;
        }
    tmdn_set_error("lookup: unknown buffer '%s'", name);
    return TMDN_EINVAL;
}

namespace {
// launch shape of the persistent edge kernels: one CTA per SM, W workers of H threads, n_hot table rows + worker scratch
struct PersistLaunch { bool ok; unsigned grid; dim3 block; size_t smem; int n_hot; };
template <typename Kern>
PersistLaunch persist_launch(Kern kernel, bool eligible, int N, int H, int workers, const FilterView& fv, int ctas_per_sm = 1) {
    PersistLaunch p;
    memset(&p, 0, sizeof(p));
    if (!eligible || !g_tmdn_tuning[TMDN_TUNE_EDGE_PERSIST] || H != 128 || N < 8192 || fv.grid.n_int <= 0) return p;
    int n_hot = g_tmdn_tuning[TMDN_TUNE_HOT_ROWS];
    if (n_hot > fv.grid.n_int) n_hot = fv.grid.n_int;
    if (n_hot > 32) n_hot = 32;    // 192 KB at H = 128
    if (n_hot < 0) n_hot = 0;
    p.n_hot = n_hot;
    p.smem = (size_t)n_hot * 3 * H * sizeof(float4) + sizeof(float) * WORKER_RED_FLOATS * workers;
    p.block = dim3((unsigned)H, (unsigned)workers, 1);
    p.grid = (unsigned)(tmdn_sm_count() * ctas_per_sm);
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem) != cudaSuccess) {
        cudaGetLastError();
        return p;
    }
    p.ok = true;
    return p;
}
}  // namespace

namespace {
// One evaluation = the stage sequence below.  Halo points (the only places where a spatially decomposed run needs
// data of atoms it does not own; owner -> ghost copies of node rows, see torchmdnet/distributed.py):
//   after EMBED        : Y0            after LAYER_FWD(l) : Y{l+1}
//   after HEAD         : Dias (= dM of layer L-1, or dE/d(ias) if there are no layers)
//   after LAYER_BWD(l) : Dias (= dM of layer l-1, or dE/d(ias) for l = 0)
// Rows (targets) are the first n_rows atoms; the remaining n_atoms - n_rows atoms are ghosts that only appear as
// sources.  Because the edge set is symmetric every adjoint is gathered, never scattered, so the backward pass needs
// the same owner -> ghost direction as the forward pass and forces of owned atoms are complete locally.
struct TnRun {
    const tmdn_tensornet_model* m;
    const int64_t* z; const float* pos; const int64_t* batch;
    int64_t n_atoms, n_rows, n_mol, P;
    const double* box; int use_periodic, strategy;
    float* energy; float* forces; int32_t* status;
    cudaStream_t stream;
    TnWs w;
    FilterView fv_emb, fv_layer[TMDN_MAX_LAYERS];

    int init(void* workspace, size_t workspace_bytes) {
        TMDN_TRY(check_model(m));
        TMDN_CHECK_ARG(z && pos && batch && status && workspace, "tensornet: null pointer argument");
        TMDN_CHECK_ARG(n_atoms > 0 && n_mol > 0 && P > 0, "tensornet: bad sizes");
        TMDN_CHECK_ARG(n_rows > 0 && n_rows <= n_atoms, "tensornet: n_rows must be in [1, n_atoms]");
        TMDN_CHECK_ARG((long long)n_atoms * 9 * m->hidden < (1LL << 40), "tensornet: system too large");
        TMDN_CHECK_ARG(P < (1LL << 31) && n_atoms < (1LL << 31) / 16, "tensornet: sizes exceed int32 indexing");
        TMDN_CHECK_ARG(strategy != TMDN_CELL || box, "tensornet: the cell strategy needs a box");
        Arena arena(workspace, workspace_bytes);
        TMDN_TRY(carve_tn(arena, m, n_atoms, P, n_mol, w, nullptr, box, strategy, m->cutoff_upper));
        if (!arena.ok) {
            tmdn_set_error("tensornet: workspace too small (%zu < %zu)", workspace_bytes, arena.off);
            return TMDN_EWORKSPACE;
        }
        fv_emb = make_view(m, m->emb_filter, w.F[0], w.dF[0]);
        for (int l = 0; l < m->num_layers; ++l) fv_layer[l] = make_view(m, m->layer_filter[l], w.F[l + 1], w.dF[l + 1]);
        return TMDN_OK;
    }

    // node buffers advanced to row r0 (every per-row array is row-major, so a row range is a pointer shift)
    static TnWs shift_rows(const TnWs& a, size_t r0, size_t H, int L) {
        TnWs x = a;
        auto sh = [&](float*& p, size_t row_floats) { if (p) p += r0 * row_floats; };
        sh(x.ias, 9 * H); sh(x.iasm, 9 * H); sh(x.nh, H); sh(x.h1, 2 * H); sh(x.a1, 2 * H); sh(x.h2, 3 * H);
        sh(x.fias, 3 * H); sh(x.Xh, 9 * H); sh(x.tauh, 9 * H); sh(x.xl, 3 * H); sh(x.stat, 2); sh(x.h3, H);
        sh(x.a3, H); sh(x.h4, H / 2); sh(x.a4, H / 2); sh(x.e_atom, 1); sh(x.dXa, 9 * H); sh(x.dB, 9 * H);
        sh(x.dB2, 9 * H); sh(x.dC, 9 * H); sh(x.d4, H / 2); sh(x.d3, H); sh(x.dxl, 3 * H); sh(x.dh2, 3 * H);
        sh(x.dh1, 2 * H); sh(x.dnh, H); sh(x.gpos, 3);
        for (int l = 0; l <= L; ++l) sh(x.X[l], 9 * H);
        for (int l = 0; l < L; ++l) { sh(x.Y[l], 9 * H); sh(x.M[l], 9 * H); sh(x.dX[l], 9 * H); }
        return x;
    }
    // Backward halo point k (k = 0: after HEAD, k: after LAYER_BWD(L-k)) lives in buffer k & 1 ("Dias" / "Dias1"):
    // a stage reads the neighbours' rows of one buffer and produces the rows of the other, so a rank that runs ahead
    // can deliver its rows into a neighbour's ghost rows while that neighbour still reads the previous ones.
    static float* dBk(const TnWs& x, int k) { return (k & 1) ? x.dB2 : x.dB; }

    // adjoint of the (M, Y) product of layer l: dB (= d tau-hat mixed back) -> dB (= dM_l), dC (= direct dY_l)
    // (w: row-shifted buffers, N rows, dB: the rows of this halo point's buffer)
    int pre_bwd(const TnWs& w, int N, float* dB, int l) {
        const int H = m->hidden;
        const long long NH = (long long)N * H;
        const unsigned egrid = (unsigned)cdiv(NH, 256);
        TMDN_TRY(mix_launch(EPI_NONE, mix_batch(dB, w.dC, m->layer_mixT[l] + 3, N, H), stream));
        {
            KTimer ktimer__("tn_prod_bwd", stream);
            launch_k(tn_prod_bwd, egrid, 256, 0, stream, w.Y[l], w.M[l], w.dC, dB, NH, H, m->so3);
            TMDN_LAUNCH_CHECK();
        }
        return TMDN_OK;
    }
    // adjoint of the embedding node stage: dXa -> dB (= dE/d ias)
    int embed_node_bwd(const TnWs& w, int N, float* dB, const float* Xn = nullptr) {
        const int H = m->hidden, act = m->activation;
        const long long NH = (long long)N * H;
        const unsigned egrid = (unsigned)cdiv(NH, 256);
        const int apb = H >= 128 ? 1 : 128 / H;
        {
            KTimer ktimer__("tn_embed_scale_bwd", stream);
            launch_k(tn_embed_scale_bwd, egrid, 256, 0, stream, w.dXa, w.iasm, w.fias, w.h2, act, w.dh2, NH, H, Xn);
            TMDN_LAUNCH_CHECK();
        }
        GemmP p = gemm_desc(w.dh2, 3 * H, m->emb_w2T, nullptr, w.dh1, 2 * H, N, 2 * H, 3 * H);
        p.aux = w.h1; p.ldaux = 2 * H; p.act = act;
        TMDN_TRY(node_gemm1(EPI_MULGRAD, p, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
        GemmP q = gemm_desc(w.dh1, 2 * H, m->emb_w1T, nullptr, w.dnh, H, N, H, 2 * H);
        TMDN_TRY(node_gemm1(EPI_NONE, q, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
        TMDN_TRY(mix_launch(EPI_NONE, mix_batch(w.dXa, dB, m->emb_mixT, N, H), stream));
        {
            KTimer ktimer__("tn_embed_norm_bwd", stream);
            launch_k(tn_embed_norm_bwd, (unsigned)cdiv(N, apb), dim3(H, apb), 0, stream, w.ias, m->emb_ln_w, w.dnh, dB, N, H);
            TMDN_LAUNCH_CHECK();
        }
        return TMDN_OK;
    }

    // One stage over the rows [r0, r1) of the owned atoms.  Row-local kernels run on row-shifted buffers, the edge
    // kernels index rows absolutely (their gathers reach every local atom, ghosts included).
    int stage(int s, int l, int64_t r0_, int64_t r1_) {
        TMDN_CHECK_ARG(s >= TMDN_TN_EMBED && s <= TMDN_TN_GRAPH, "tensornet: bad stage %d", s);
        TMDN_CHECK_ARG(r0_ >= 0 && r0_ < r1_ && r1_ <= n_rows, "tensornet: bad row range [%lld, %lld)", (long long)r0_,
                       (long long)r1_);
        const int r0 = (int)r0_, r1 = (int)r1_, N = r1 - r0, NT = (int)n_atoms, H = m->hidden, L = m->num_layers;
        const int act = m->activation;
        const bool tab = m->filter_mode == TMDN_FILTER_TABLE;
        const bool full = r0 == 0 && r1 == (int)n_rows;
        const TnWs& wa = this->w;                                  // absolute buffers (edge kernels)
        const TnWs w = shift_rows(wa, (size_t)r0, (size_t)H, L);   // buffers of rows [r0, r1) (row-local kernels)
        const float4* geom = (const float4*)wa.geom;
        const int apb = H >= 128 ? 1 : 128 / H;
        const dim3 rblock(H, apb);
        // launch shapes (edge slices per row) follow the size of the whole problem, not of the row range, so that a
        // split evaluation sums in the same order as the unsplit one
        auto edge_shape = [&](int max_threads) {
            EdgeLaunch e = edge_launch((int)n_rows, H, max_threads);
            e.grid = (unsigned)cdiv(N, (int)e.block.z);
            return e;
        };
        const EdgeLaunch el = edge_shape(1024);    // embedding edge kernels: <= 64 registers/thread
        const EdgeLaunch elb = edge_shape(1024);   // 56 registers
        const EdgeLaunch elf2 = edge_shape(MSG_FWD_THREADS);
        // two edges in flight per thread; up to 4 slices per row when the rows alone cannot fill the machine
        const bool bwd4 = tab && (long long)n_rows * H * 2 < (long long)tmdn_sm_count() * 2048;
        const EdgeLaunch elb2 = edge_shape(bwd4 ? 512 : MSG_BWD_THREADS);
        const unsigned rgrid = (unsigned)cdiv(N, apb);
        const long long NH = (long long)N * H;
        const unsigned egrid = (unsigned)cdiv(NH, 256);
        // few rows: one CTA per row leaves a second partial wave (166 rows on 148 SMs); rows are visited in order of
        // decreasing degree so that the short ones fill the tail (longest-processing-time-first)
        const bool lpt = g_tmdn_tuning[TMDN_TUNE_LPT_ROWS] && n_rows <= LPT_MAX_ROWS;
        const int* order = (lpt && full) ? wa.row_order : nullptr;
        const bool small_rows = n_rows <= LPT_MAX_ROWS;
        TMDN_CHECK_ARG((s != TMDN_TN_LAYER_FWD && s != TMDN_TN_LAYER_BWD) || (l >= 0 && l < L), "tensornet: bad layer %d", l);

        if (s == TMDN_TN_GRAPH) {
            // graph: neighbour search straight into CSR by target (= row 0 of the reference's pair list,
            // tensornet.py:313-315), self loops included
            {
                KTimer ktimer__("neighbors_csr", stream);
                TMDN_TRY(tmdn_csr_build_rows_f32(strategy, pos, batch, n_atoms, n_rows, box, use_periodic, m->cutoff_lower,
                                                 m->cutoff_upper, P, wa.row_ptr, wa.col, wa.geom, wa.num_found, wa.graph_ws,
                                                 wa.graph_ws_bytes, (tmdn_stream_t)stream));
            }
            TMDN_CUDA(cudaMemsetAsync(status, 0, 4 * sizeof(int32_t), stream));
            {
                KTimer ktimer__("tn_prep_kernel", stream);
                launch_k(tn_prep_kernel, (unsigned)cdiv(NT, 256), 256, 0, stream, z, wa.z32, NT, m->max_z, wa.num_found,
                         (long long)P, status, wa.row_ptr, (int)n_rows, lpt ? wa.row_order : nullptr, batch);
                TMDN_LAUNCH_CHECK();
            }
            if (!tab) {  // exact filters: rbf once, then one MLP chain per filter set
                const long long tot = (long long)P * m->num_rbf;
                {
                    KTimer ktimer__("tn_rbf_kernel", stream);
                    launch_k(tn_rbf_kernel, (unsigned)cdiv(tot, 256), 256, 0, stream, geom, wa.row_ptr, (int)n_rows,
                             (long long)P, m->num_rbf, m->rbf_type, m->cutoff_lower, m->cutoff_upper, m->rbf_p0, m->rbf_p1,
                             wa.R, wa.dR);
                    TMDN_LAUNCH_CHECK();
                }
                TMDN_TRY(eval_filter_exact(m, m->emb_filter, wa, n_atoms, P, wa.F[0], wa.dF[0], stream));
                for (int k = 0; k < L; ++k)
                    TMDN_TRY(eval_filter_exact(m, m->layer_filter[k], wa, n_atoms, P, wa.F[k + 1], wa.dF[k + 1], stream));
            }
            return TMDN_OK;
        }
        if (s == TMDN_TN_EMBED) {
            {
                KTimer ktimer__("tn_embed_edge_fwd", stream);
                if (tab && H == 128)
                    launch_k(tn_embed_edge_fwd<true, 128>, el.grid, el.block, el.smem, stream, wa.row_ptr, wa.col, geom,
                             wa.z32, m->P, m->Q, fv_emb, r1, H, wa.ias, m->emb_ln_w, m->emb_ln_b, wa.nh, order, r0);
                else if (tab)
                    launch_k(tn_embed_edge_fwd<true>, el.grid, el.block, el.smem, stream, wa.row_ptr, wa.col, geom, wa.z32,
                             m->P, m->Q, fv_emb, r1, H, wa.ias, m->emb_ln_w, m->emb_ln_b, wa.nh, order, r0);
                else
                    launch_k(tn_embed_edge_fwd<false>, el.grid, el.block, el.smem, stream, wa.row_ptr, wa.col, geom, wa.z32,
                             m->P, m->Q, fv_emb, r1, H, wa.ias, m->emb_ln_w, m->emb_ln_b, wa.nh, order, r0);
                TMDN_LAUNCH_CHECK();
            }
            GemmP p = gemm_desc(w.nh, H, m->emb_w1, m->emb_b1, w.a1, 2 * H, N, 2 * H, H);
            p.C2 = w.h1; p.act = act;
            TMDN_TRY(node_gemm1(EPI_ACT, p, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
            GemmP q = gemm_desc(w.a1, 2 * H, m->emb_w2, m->emb_b2, w.fias, 3 * H, N, 3 * H, 2 * H);
            q.C2 = w.h2; q.act = act;
            TMDN_TRY(node_gemm1(EPI_ACT, q, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
            GemmBatch b = mix_batch(w.ias, w.X[0], m->emb_mix, N, H);
            for (int g = 0; g < 3; ++g) { b.p[g].C2 = w.iasm; b.p[g].aux = w.fias; }
            TMDN_TRY(mix_launch(EPI_SCALE3, b, stream));
            if (L > 0) {
                {
                    KTimer ktimer__("tn_normalize_fwd", stream);
                    launch_k(tn_normalize_fwd, egrid, 256, 0, stream, w.X[0], w.Xh, NH, H);
                    TMDN_LAUNCH_CHECK();
                }
                TMDN_TRY(mix_launch(EPI_NONE, mix_batch(w.Xh, w.Y[0], m->layer_mix[0], N, H), stream));
            }
            return TMDN_OK;
        }
        if (s == TMDN_TN_LAYER_FWD) {
            {
                KTimer ktimer__("tn_msg_fwd", stream);
                const int fvar = g_tmdn_tuning[TMDN_TUNE_EDGE_VARIANT];
                // default: two CTAs of 5 workers per SM (1280 threads at 46 registers; 4 % faster than one CTA of 8 workers, r02m)
                const PersistLaunch pl = (fvar & 9) ? persist_launch(tn_msg_fwd_p<128, 1>, tab, N, H, 8, fv_layer[l])
                                                    : persist_launch(tn_msg_fwd_p<128, 1, 640, 2>, tab, N, H, 5, fv_layer[l], 2);
                // small systems: 4 slices x 2 edges per thread at 128 registers instead of 8 slices x 1 at 64 -- the same
                // number of edges in flight per SM, but with 64 registers ptxas reuses the landing registers of the
                // table rows and an iteration becomes 3-4 dependent round trips instead of one (SASS, r01i)
                if (pl.ok && !(fvar & 9) && pl.n_hot == 0)
                    launch_k(tn_msg_fwd_p<128, 1, 640, 2, false>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], r1, m->so3, wa.M[l], wa.tauh, 0, r0);
                else if (pl.ok && !(fvar & 9))
                    launch_k(tn_msg_fwd_p<128, 1, 640, 2>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], r1, m->so3, wa.M[l], wa.tauh, pl.n_hot, r0);
                else if (pl.ok && (fvar & 1))
                    launch_k(tn_msg_fwd_p<128, 2>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom, fv_layer[l],
                             wa.Y[l], r1, m->so3, wa.M[l], wa.tauh, pl.n_hot, r0);
                else if (pl.ok)
                    launch_k(tn_msg_fwd_p<128, 1>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom, fv_layer[l],
                             wa.Y[l], r1, m->so3, wa.M[l], wa.tauh, pl.n_hot, r0);
                else if (tab && small_rows) {
                    const EdgeLaunch e4 = edge_shape(512);
                    launch_k(tn_msg_fwd<true, 2, 0, true, 512>, e4.grid, e4.block, e4.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], r1, H, m->so3, wa.M[l], wa.tauh, order, r0);
                } else if (tab)
                    launch_k(tn_msg_fwd<true>, elf2.grid, elf2.block, elf2.smem, stream, wa.row_ptr, wa.col, geom, fv_layer[l],
                             wa.Y[l], r1, H, m->so3, wa.M[l], wa.tauh, order, r0);
                else
                    launch_k(tn_msg_fwd<false>, elf2.grid, elf2.block, elf2.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], r1, H, m->so3, wa.M[l], wa.tauh, order, r0);
                TMDN_LAUNCH_CHECK();
            }
            TMDN_TRY(mix_launch(EPI_NONE, mix_batch(w.tauh, w.dX[l], m->layer_mix[l] + 3, N, H), stream));
            // Xh is both input (residual) and, for the next layer, output: safe because each thread reads its own
            // (atom, channel) column before writing it.
            {
                KTimer ktimer__("tn_update_fwd", stream);
                launch_k(tn_update_fwd, egrid, 256, 0, stream, w.Xh, w.dX[l], w.X[l + 1], l + 1 < L ? w.Xh : nullptr, NH, H);
                TMDN_LAUNCH_CHECK();
            }
            if (l + 1 < L) TMDN_TRY(mix_launch(EPI_NONE, mix_batch(w.Xh, w.Y[l + 1], m->layer_mix[l + 1], N, H), stream));
            return TMDN_OK;
        }
        if (s == TMDN_TN_HEAD) {
            TMDN_CHECK_ARG(energy, "tensornet: null energy pointer");
            {
                KTimer ktimer__("tn_out_norm_fwd", stream);
                launch_k(tn_out_norm_fwd, rgrid, rblock, 0, stream, w.X[L], m->out_ln_w, m->out_ln_b, w.xl, w.stat, N, H);
                TMDN_LAUNCH_CHECK();
            }
            {
                GemmP p = gemm_desc(w.xl, 3 * H, m->out_w, m->out_b, w.a3, H, N, H, 3 * H);
                p.C2 = w.h3; p.act = act;
                TMDN_TRY(node_gemm1(EPI_ACT, p, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
                GemmP q = gemm_desc(w.a3, H, m->head_w1, m->head_b1, w.a4, H / 2, N, H / 2, H);
                q.C2 = w.h4; q.act = act;
                TMDN_TRY(node_gemm1(EPI_ACT, q, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
            }
            {
                KTimer ktimer__("tn_head_energy", stream);
                launch_k(tn_head_energy, (unsigned)cdiv(N, 4), 128, 0, stream, w.a4, m->head_w2, m->head_b2, m->std, w.e_atom, N,
                         H / 2, (const float*)w.h4, act, forces ? w.d4 : (float*)nullptr);
                TMDN_LAUNCH_CHECK();
            }
            if (r1 == (int)n_rows) {   // the molecular sums run in the call that ends the row range (make it the last one)
                KTimer ktimer__("tn_mol_reduce", stream);
                launch_k(tn_mol_reduce, (unsigned)n_mol, 128, 0, stream, wa.e_atom, batch, (int)n_rows, m->mean, energy, status);
                TMDN_LAUNCH_CHECK();
            }
            if (!forces) return TMDN_OK;
            TMDN_CUDA(cudaMemsetAsync(w.gpos, 0, sizeof(float) * 3 * (size_t)N, stream));
            GemmP p = gemm_desc(w.d4, H / 2, m->head_w1T, nullptr, w.d3, H, N, H, H / 2);
            p.aux = w.h3; p.ldaux = H; p.act = act;
            TMDN_TRY(node_gemm1(EPI_MULGRAD, p, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
            GemmP q = gemm_desc(w.d3, H, m->out_wT, nullptr, w.dxl, 3 * H, N, 3 * H, H);
            TMDN_TRY(node_gemm1(EPI_NONE, q, stream, m->gemm_mode == TMDN_GEMM_TF32X3));
            float* dB = dBk(w, 0);
            {
                KTimer ktimer__("tn_out_norm_bwd", stream);
                launch_k(tn_out_norm_bwd, rgrid, rblock, 0, stream, w.X[L], m->out_ln_w, w.stat, w.dxl, w.dXa, N, H,
                         L > 0 ? (const float*)w.dX[L - 1] : (const float*)nullptr, L > 0 ? dB : (float*)nullptr);
                TMDN_LAUNCH_CHECK();
            }
            return L > 0 ? pre_bwd(w, N, dB, L - 1) : embed_node_bwd(w, N, dB);
        }
        if (s == TMDN_TN_LAYER_BWD) {
            const int k_in = L - 1 - l;
            const float* dM = dBk(wa, k_in);        // gathered over all local atoms
            float* dB = dBk(w, k_in + 1);           // rows of the next halo point
            {
                KTimer ktimer__("tn_msg_bwd", stream);
                const int variant = g_tmdn_tuning[TMDN_TUNE_EDGE_VARIANT];
                const PersistLaunch pl = (variant & 2) ? persist_launch(tn_msg_bwd_p<128, 1, 1024>, tab, N, H, 8, fv_layer[l])
                                                         : persist_launch(tn_msg_bwd_p<128, 2, 512>, tab, N, H, 4, fv_layer[l]);
                // (capping the registers to fit 5 or 6 workers -- 96 / 80 registers -- makes ptxas serialise the gathers:
                // 16.1 / 13.1 ms per step against 7.6 ms, r02m)
                if (pl.ok && (variant & 2))
                    launch_k(tn_msg_bwd_p<128, 1, 1024>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], dM, wa.dC, wa.gpos, r1, pl.n_hot, r0);
                else if (pl.ok && pl.n_hot == 0)
                    launch_k(tn_msg_bwd_p<128, 2, 512, false>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], dM, wa.dC, wa.gpos, r1, 0, r0);
                else if (pl.ok)
                    launch_k(tn_msg_bwd_p<128, 2, 512>, pl.grid, pl.block, pl.smem, stream, wa.row_ptr, wa.col, geom,
                             fv_layer[l], wa.Y[l], dM, wa.dC, wa.gpos, r1, pl.n_hot, r0);
                else if (bwd4)
                    launch_k(tn_msg_bwd<true, 2, 0, 512, true>, elb2.grid, elb2.block, elb2.smem, stream, wa.row_ptr, wa.col,
                             geom, fv_layer[l], wa.Y[l], dM, wa.dC, wa.gpos, r1, H, order, r0);
                else if (tab)
                    launch_k(tn_msg_bwd<true>, elb2.grid, elb2.block, elb2.smem, stream, wa.row_ptr, wa.col, geom, fv_layer[l],
                             wa.Y[l], dM, wa.dC, wa.gpos, r1, H, order, r0);
                else
                    launch_k(tn_msg_bwd<false>, elb2.grid, elb2.block, elb2.smem, stream, wa.row_ptr, wa.col, geom, fv_layer[l],
                             wa.Y[l], dM, wa.dC, wa.gpos, r1, H, order, r0);
                TMDN_LAUNCH_CHECK();
            }
            GemmBatch b = mix_batch(w.dC, w.dXa, m->layer_mixT[l], N, H);
            for (int g = 0; g < 3; ++g) b.p[g].accumulate = 1;
            TMDN_TRY(mix_launch(EPI_NONE, b, stream));
            if (l > 0) {
                {
                    KTimer ktimer__("tn_normalize_bwd", stream);
                    launch_k(tn_normalize_update_bwd, egrid, 256, 0, stream, w.X[l], w.dXa, w.dX[l - 1], dB, NH, H);
                    TMDN_LAUNCH_CHECK();
                }
                return pre_bwd(w, N, dB, l - 1);
            }
            return embed_node_bwd(w, N, dB, w.X[0]);   // layer 0: the normalise adjoint runs inside tn_embed_scale_bwd
        }
        // TMDN_TN_FINAL
        TMDN_CHECK_ARG(forces, "tensornet: null forces pointer");
        {
            const float* Dias = dBk(wa, L);
            KTimer ktimer__("tn_embed_edge_bwd", stream);
            if (tab && H == 128)
                launch_k(tn_embed_edge_bwd<true, 128>, elb.grid, elb.block, elb.smem, stream, wa.row_ptr, wa.col, geom, wa.z32,
                         m->P, m->Q, fv_emb, Dias, wa.gpos, forces, r1, H, order, r0, status);
            else if (tab)
                launch_k(tn_embed_edge_bwd<true>, elb.grid, elb.block, elb.smem, stream, wa.row_ptr, wa.col, geom, wa.z32,
                         m->P, m->Q, fv_emb, Dias, wa.gpos, forces, r1, H, order, r0, status);
            else
                launch_k(tn_embed_edge_bwd<false>, elb.grid, elb.block, elb.smem, stream, wa.row_ptr, wa.col, geom, wa.z32,
                         m->P, m->Q, fv_emb, Dias, wa.gpos, forces, r1, H, order, r0, status);
            TMDN_LAUNCH_CHECK();
        }
        return TMDN_OK;
    }
};
}  // namespace

extern "C" int tmdn_tensornet_stage_rows(const tmdn_tensornet_model* m, int stage, int layer, const int64_t* z,
                                         const float* pos, const int64_t* batch, int64_t n_atoms, int64_t n_rows,
                                         int64_t row_begin, int64_t row_end, int64_t n_mol, const double box[9],
                                         int use_periodic, int strategy, int64_t max_edges, float* energy, float* forces,
                                         int32_t* status, void* workspace, size_t workspace_bytes, tmdn_stream_t stream_) {
    TnRun r{m, z, pos, batch, n_atoms, n_rows, n_mol, max_edges, box, use_periodic, strategy, energy, forces, status,
            (cudaStream_t)stream_};
    TMDN_TRY(r.init(workspace, workspace_bytes));
    if (stage == TMDN_TN_GRAPH) return r.stage(TMDN_TN_GRAPH, 0, 0, n_rows);
    return r.stage(stage, layer, row_begin, row_end);
}

extern "C" int tmdn_tensornet_stage(const tmdn_tensornet_model* m, int stage, int layer, const int64_t* z,
                                    const float* pos, const int64_t* batch, int64_t n_atoms, int64_t n_rows,
                                    int64_t n_mol, const double box[9], int use_periodic, int strategy,
                                    int64_t max_edges, float* energy, float* forces, int32_t* status, void* workspace,
                                    size_t workspace_bytes, tmdn_stream_t stream_) {
    TnRun r{m, z, pos, batch, n_atoms, n_rows, n_mol, max_edges, box, use_periodic, strategy, energy, forces, status,
            (cudaStream_t)stream_};
    TMDN_TRY(r.init(workspace, workspace_bytes));
    if (stage == TMDN_TN_EMBED) TMDN_TRY(r.stage(TMDN_TN_GRAPH, 0, 0, n_rows));   // whole-range form: graph + embedding
    return r.stage(stage, layer, 0, n_rows);
}

extern "C" int tmdn_tensornet_energy_forces(const tmdn_tensornet_model* m, const int64_t* z, const float* pos,
                                            const int64_t* batch, int64_t n_atoms, int64_t n_mol, const double box[9],
                                            int use_periodic, int strategy, int64_t max_edges, float* energy,
                                            float* forces, int32_t* status, void* workspace, size_t workspace_bytes,
                                            tmdn_stream_t stream_) {
    TMDN_CHECK_ARG(energy, "tensornet: null pointer argument");
    TnRun r{m, z, pos, batch, n_atoms, n_atoms, n_mol, max_edges, box, use_periodic, strategy, energy, forces, status,
            (cudaStream_t)stream_};
    TMDN_TRY(r.init(workspace, workspace_bytes));
    const int L = m->num_layers;
    const int64_t N = n_atoms;
    TMDN_TRY(r.stage(TMDN_TN_GRAPH, 0, 0, N));
    TMDN_TRY(r.stage(TMDN_TN_EMBED, 0, 0, N));
    for (int l = 0; l < L; ++l) TMDN_TRY(r.stage(TMDN_TN_LAYER_FWD, l, 0, N));
    TMDN_TRY(r.stage(TMDN_TN_HEAD, 0, 0, N));
    if (!forces) return TMDN_OK;
    for (int l = L - 1; l >= 0; --l) TMDN_TRY(r.stage(TMDN_TN_LAYER_BWD, l, 0, N));
    return r.stage(TMDN_TN_FINAL, 0, 0, N);
}

extern "C" int tmdn_channel_mix_f32(const float* x, const float* w, float* y, int64_t n_atoms, int H, int accumulate,
                                    tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_CHECK_ARG(x && w && y, "channel_mix: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_atoms * 9 < (1LL << 31), "channel_mix: bad atom count");
    TMDN_CHECK_ARG(H > 0 && H % 4 == 0, "channel_mix: H must be a positive multiple of 4");
    const float* Wg[3] = {w, w + (size_t)H * H, w + 2 * (size_t)H * H};
    GemmBatch b = mix_batch(x, y, Wg, (int)n_atoms, H);
    for (int g = 0; g < 3; ++g) { b.p[g].accumulate = accumulate; b.p[g].w_static = 0; }   // caller's tensor
    return mix_launch(EPI_NONE, b, stream);
}

extern "C" int tmdn_node_gemm_f32(const float* a, const float* w, const float* bias, float* c, int64_t M, int64_t N,
                                  int64_t K, int accumulate, tmdn_stream_t stream_) {
    TMDN_CHECK_ARG(a && w && c, "node_gemm: null pointer argument");
    TMDN_CHECK_ARG(M > 0 && N > 0 && K > 0 && M < (1LL << 31) && N < (1 << 20) && K < (1 << 20), "node_gemm: bad sizes");
    GemmP p = gemm_desc(a, (int)K, w, bias, c, (int)N, (int)M, (int)N, (int)K);
    p.accumulate = accumulate;
    p.w_static = 0;   // caller's tensor, possibly written by the kernel just before this one
    return node_gemm1(EPI_NONE, p, (cudaStream_t)stream_);
}
