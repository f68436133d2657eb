// gemm.cuh -- fp32 FFMA GEMM  C[M,N] = epi(A[M,K] * W[N,K]^T + bias)  with the fused epilogues the
// energy/force pipeline needs.  Exact fp32 (no TF32) so results stay within the 1e-5 parity budget.
//
// Up to three independent problems run in one launch (blockIdx.z); that is how the grouped channel mixes of
// TensorNet (one H x H weight per irreducible component I / A / S acting on rows of an [N,9,H] tensor) are
// done without permuting the tensor: each problem sees its rows through the map
//     physical_row(m) = (m / gs) * rpa + c0 + m % gs
// (identity for gs = rpa = 1, c0 = 0).
#pragma once
#include "common.cuh"

enum { EPI_NONE = 0, EPI_ACT = 1, EPI_MULGRAD = 2, EPI_SCALE3 = 3 };

struct GemmP {
    const float* A;
    const float* W;     // [N,K] row-major
    const float* bias;  // [N] or null
    float* C;
    float* C2;          // EPI_ACT: pre-activation; EPI_SCALE3: unscaled result; may be null
    const float* aux;   // EPI_MULGRAD: pre-activation [rows,N] (ld = ldaux); EPI_SCALE3: scale[(m/gs)*3+g][N]
    int M, N, K;
    int lda, ldc, ldaux;
    int gs, c0, rpa, g;
    int accumulate;     // C += result
    int act;
};
struct GemmBatch {
    GemmP p[3];
    int n;
};

constexpr int GBM = 64, GBN = 64, GBK = 16, GTHREADS = 256;

template <int EPI>
__global__ void __launch_bounds__(GTHREADS) sgemm_nt_kernel(GemmBatch batch) {
    const GemmP& P = batch.p[blockIdx.z];
    const int m0 = blockIdx.x * GBM, n0 = blockIdx.y * GBN;
    if (m0 >= P.M || n0 >= P.N) return;
    __shared__ __align__(16) float As[GBK][GBM + 4];
    __shared__ __align__(16) float Bs[GBK][GBN + 4];
    const int t = threadIdx.x;
    const int tx = t & 15, ty = t >> 4;
    // global -> smem loader mapping: 64 rows x 16 k = 256 float4
    const int lrow = t >> 2, lk = (t & 3) * 4;
    const int am = m0 + lrow;
    const bool a_ok = am < P.M;
    const size_t a_row = a_ok ? ((size_t)(am / P.gs) * P.rpa + P.c0 + am % P.gs) : 0;
    const float* a_ptr = P.A + a_row * P.lda;
    const int wn = n0 + lrow;
    const bool w_ok = wn < P.N;
    const float* w_ptr = P.W + (size_t)(w_ok ? wn : 0) * P.K;
    const bool vec4 = (P.K & 3) == 0 && (P.lda & 3) == 0;

    float acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;

    float4 ra, rw;
    auto gload = [&](int k0) {
        const int k = k0 + lk;
        ra = make_float4(0.f, 0.f, 0.f, 0.f);
        rw = ra;
        if (vec4) {
            if (k < P.K) {
                if (a_ok) ra = *reinterpret_cast<const float4*>(a_ptr + k);
                if (w_ok) rw = *reinterpret_cast<const float4*>(w_ptr + k);
            }
        } else {
            float ta[4] = {0.f, 0.f, 0.f, 0.f}, tw[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (k + q < P.K) {
                    if (a_ok) ta[q] = a_ptr[k + q];
                    if (w_ok) tw[q] = w_ptr[k + q];
                }
            ra = make_float4(ta[0], ta[1], ta[2], ta[3]);
            rw = make_float4(tw[0], tw[1], tw[2], tw[3]);
        }
    };
    gload(0);
    for (int k0 = 0; k0 < P.K; k0 += GBK) {
        __syncthreads();
        As[lk + 0][lrow] = ra.x; As[lk + 1][lrow] = ra.y; As[lk + 2][lrow] = ra.z; As[lk + 3][lrow] = ra.w;
        Bs[lk + 0][lrow] = rw.x; Bs[lk + 1][lrow] = rw.y; Bs[lk + 2][lrow] = rw.z; Bs[lk + 3][lrow] = rw.w;
        __syncthreads();
        if (k0 + GBK < P.K) gload(k0 + GBK);
#pragma unroll
        for (int k = 0; k < GBK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], bv[c], acc[r][c]);
        }
    }
    // epilogue
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int m = m0 + ty * 4 + r;
        if (m >= P.M) continue;
        const int atom = m / P.gs;
        const size_t prow = (size_t)atom * P.rpa + P.c0 + m % P.gs;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int n = n0 + tx * 4 + c;
            if (n >= P.N) continue;
            float v = acc[r][c];
            if (P.bias) v += P.bias[n];
            const size_t ci = prow * P.ldc + n;
            if (EPI == EPI_ACT) {
                if (P.C2) P.C2[ci] = v;
                v = act_apply(P.act, v);
            } else if (EPI == EPI_MULGRAD) {
                v *= act_grad(P.act, P.aux[prow * P.ldaux + n]);
            } else if (EPI == EPI_SCALE3) {
                if (P.C2) P.C2[ci] = v;
                v *= P.aux[((size_t)atom * 3 + P.g) * P.N + n];
            }
            if (P.accumulate) v += P.C[ci];
            P.C[ci] = v;
        }
    }
}

static inline GemmP gemm_desc(const float* A, int lda, const float* W, const float* bias, float* C, int ldc, int M,
                              int N, int K) {
    GemmP p;
    memset(&p, 0, sizeof(p));
    p.A = A; p.lda = lda; p.W = W; p.bias = bias; p.C = C; p.ldc = ldc;
    p.M = M; p.N = N; p.K = K;
    p.gs = 1; p.c0 = 0; p.rpa = 1; p.g = 0;
    return p;
}

static inline int gemm_launch(int epi, const GemmBatch& b, cudaStream_t stream, const char* tag = "node_gemm") {
    KTimer ktimer__(tag, stream);
    int maxM = 0, maxN = 0;
    for (int i = 0; i < b.n; ++i) {
        if (b.p[i].M > maxM) maxM = b.p[i].M;
        if (b.p[i].N > maxN) maxN = b.p[i].N;
    }
    if (maxM == 0 || maxN == 0) return TMDN_OK;
    dim3 grid((unsigned)cdiv(maxM, GBM), (unsigned)cdiv(maxN, GBN), (unsigned)b.n);  // M tiles on x (2^31-1 limit)
    switch (epi) {
        case EPI_NONE: sgemm_nt_kernel<EPI_NONE><<<grid, GTHREADS, 0, stream>>>(b); break;
        case EPI_ACT: sgemm_nt_kernel<EPI_ACT><<<grid, GTHREADS, 0, stream>>>(b); break;
        case EPI_MULGRAD: sgemm_nt_kernel<EPI_MULGRAD><<<grid, GTHREADS, 0, stream>>>(b); break;
        case EPI_SCALE3: sgemm_nt_kernel<EPI_SCALE3><<<grid, GTHREADS, 0, stream>>>(b); break;
        default: tmdn_set_error("gemm: bad epilogue"); return TMDN_EINVAL;
    }
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

static inline int gemm1(int epi, const GemmP& p, cudaStream_t stream, const char* tag = "node_gemm") {
    GemmBatch b;
    memset(&b, 0, sizeof(b));
    b.p[0] = p;
    b.n = 1;
    return gemm_launch(epi, b, stream, tag);
}
