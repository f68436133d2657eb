// gemm.cuh -- fp32 FFMA GEMM  C[M,N] = epi(A[M,K] * W[N,K]^T + bias)  with the fused epilogues the
// energy/force pipeline needs.  Exact fp32 (no TF32) so results stay within the 1e-5 parity budget.
//
// Up to three independent problems run in one launch (blockIdx.z); that is how the grouped channel mixes of
// TensorNet (one H x H weight per irreducible component I / A / S acting on rows of an [N,9,H] tensor) are
// done without permuting the tensor: each problem sees its rows through the map
//     physical_row(m) = (m / gs) * rpa + c0 + m % gs
// (identity for gs = rpa = 1, c0 = 0).
#pragma once
#include <stdlib.h>
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
    int w_static;       // W is a model parameter (not written by a preceding kernel): may be fetched before the PDL wait
};
struct GemmBatch {
    GemmP p[3];
    int n;
};

constexpr int GBM = 64, GBN = 64, GBK = 16, GTHREADS = 256;

template <int EPI>
__global__ void __launch_bounds__(GTHREADS) sgemm_nt_kernel(GemmBatch batch) {
    pdl_sync();
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

// ---------------------------------------------------------------------------------------------------------------
// Latency-oriented variant for small problems (a few hundred rows: single molecules, MD of a peptide).  The tiled
// kernel above walks K in 16-wide steps with one global round trip per step (K = 384 -> 24 dependent round trips,
// 11-27 us at 166 atoms, ncu r01h); here a CTA owns a (16*RM) x 32 output tile (RM = 1 in use), fetches its whole A and W panels with
// one burst of 16-byte cp.async copies (a single memory round trip, nothing staged through registers) and then runs
// the K loop out of shared memory.  Same epilogues, same row map, same fp32 FFMA arithmetic (sum order over k
// differs from the tiled kernel only in being strictly sequential).
constexpr int SBN = 32, SMALL_MAX_K = 768;
// Tile and thread tile in use.  Chain of channel mixes inside a CUDA graph, us per launch at 42 / 166 / 500 / 1000 atoms
// (profiles/microbench_small_gemm.cu, r02w): 32 x 32 tiles, 2 x 2 per thread, 256 threads (the first version; it switched
// to 16 x 32, 1 x 2 below 148 tiles) 4.8 / 8.3 / 14.4 / 27.7; 16 x 32, 2 x 2, 128 threads 5.2 / 7.7 / 11.3 / 20.2;
// larger thread tiles (4 x 2 ... 8 x 2) are slower (fewer warps to hide the LDS latency), an empty kernel of the same
// shape takes 0.6 and the panel fetch alone 3.5 us.
constexpr int SMALL_RM = 1, SMALL_TM = 2, SMALL_TN = 2;

// A thread owns TM x TN outputs: rows ty + (BM / TM) r, columns tx + (SBN / TN) c -- interleaved so that the 16-byte
// shared-memory reads of a warp fall on distinct banks or coincide (row stride K + 4 words = 4 banks).  The first version
// (TM x TN = RM x 2, 256 threads) issued 4 LDS.128 per 16 FFMA and was bound by the shared-memory pipe: of the 8.1 us a
// 166-atom channel mix takes inside a graph, 0.6 are the launch, 2.8 the panel fetch and 4.7 the K loop
// (profiles/microbench_small_gemm.cu, gpurun r02w).  Every output is still one strictly sequential fmaf chain over k, so
// all tilings give the same bits.
template <int EPI, int RM, int TM, int TN>
__global__ void __launch_bounds__((16 * RM / TM) * (SBN / TN)) sgemm_small_kernel(GemmBatch batch) {
    extern __shared__ __align__(16) float sg_smem[];
    constexpr int BM = 16 * RM, TROWS = BM / TM, TCOLS = SBN / TN, NT = TROWS * TCOLS, NW = NT / 32;
    static_assert(NT % 32 == 0 && BM % TM == 0 && SBN % TN == 0, "thread tile");
    const GemmP& P = batch.p[blockIdx.z];
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * SBN;
    // the weight panel does not depend on the previous kernel: its copies are issued before the dependency wait, so
    // under programmatic dependent launch they overlap the predecessor's tail
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const bool live = m0 < P.M && n0 < P.N;
    const int K = P.K, ld = K + 4, kv = K >> 2;
    float* As = sg_smem;
    float* Ws = sg_smem + BM * ld;
    const int t = threadIdx.x;
    // copies: a warp takes one row at a time (lane = 16-byte column, strided by 32 for K > 128) -- no integer division
    // per copy (the first version spent half of its instructions on idx / kv)
    const int lane = t & 31, wrp = t >> 5;
    auto load_w = [&]() {
        for (int row = wrp; row < SBN; row += NW) {
            const int n = n0 + row;
            const bool ok = n < P.N;
            const float* src = P.W + (size_t)(ok ? n : 0) * K;
            const unsigned dst = (unsigned)__cvta_generic_to_shared(Ws + row * ld);
            const int bytes = ok ? 16 : 0;   // zero-fill rows beyond the problem
            for (int k4 = lane; k4 < kv; k4 += 32)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4), "r"(bytes) : "memory");
        }
    };
    if (live && P.w_static) load_w();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (!live) return;
    if (!P.w_static) load_w();
    for (int row = wrp; row < BM; row += NW) {
        const int m = m0 + row;
        const bool ok = m < P.M;
        const size_t prow = !ok ? 0 : (P.gs == 1 ? (size_t)m * P.rpa + P.c0 : (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs);
        const float* src = P.A + prow * P.lda;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(As + row * ld);
        const int bytes = ok ? 16 : 0;
        for (int k4 = lane; k4 < kv; k4 += 32)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4), "r"(bytes) : "memory");
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    const int tx = t % TCOLS, ty = t / TCOLS;
    float acc[TM][TN];
#pragma unroll
    for (int r = 0; r < TM; ++r)
#pragma unroll
        for (int c = 0; c < TN; ++c) acc[r][c] = 0.f;
    const float* a0 = As + ty * ld;
    const float* w0 = Ws + tx * ld;
#pragma unroll 2
    for (int k = 0; k < K; k += 4) {
        float4 a[TM], w[TN];
#pragma unroll
        for (int r = 0; r < TM; ++r) a[r] = *reinterpret_cast<const float4*>(a0 + r * TROWS * ld + k);
#pragma unroll
        for (int c = 0; c < TN; ++c) w[c] = *reinterpret_cast<const float4*>(w0 + c * TCOLS * ld + k);
#pragma unroll
        for (int r = 0; r < TM; ++r)
#pragma unroll
            for (int c = 0; c < TN; ++c) {
                acc[r][c] = fmaf(a[r].x, w[c].x, acc[r][c]);
                acc[r][c] = fmaf(a[r].y, w[c].y, acc[r][c]);
                acc[r][c] = fmaf(a[r].z, w[c].z, acc[r][c]);
                acc[r][c] = fmaf(a[r].w, w[c].w, acc[r][c]);
            }
    }
#pragma unroll
    for (int r = 0; r < TM; ++r) {
        const int m = m0 + ty + TROWS * r;
        if (m >= P.M) continue;
        const int atom = m / P.gs;
        const size_t prow = (size_t)atom * P.rpa + P.c0 + m % P.gs;
#pragma unroll
        for (int c = 0; c < TN; ++c) {
            const int n = n0 + tx + TCOLS * c;
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

// true when every problem of the batch fits the small kernel and the launch is small enough to be latency-bound
static inline bool small_gemm_ok(const GemmBatch& b) {
    if (!g_tmdn_tuning[TMDN_TUNE_SMALL_GEMM]) return false;
    long long macs = 0;
    for (int i = 0; i < b.n; ++i) {
        const GemmP& p = b.p[i];
        if (p.K > SMALL_MAX_K || (p.K & 3) || (p.lda & 3) || p.K != b.p[0].K) return false;
        if (((uintptr_t)p.A & 15) || ((uintptr_t)p.W & 15)) return false;
        macs += (long long)p.M * p.N * p.K;
    }
    return macs <= 160LL * 1000 * 1000;   // ~9k rows of a 128 x 128 mix; above that throughput kernels win
}

template <int RM, int TM, int TN>
static inline int small_gemm_launch_cfg(int epi, const GemmBatch& b, dim3 grid, size_t smem, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        const int maxb = (16 * RM + SBN) * (SMALL_MAX_K + 4) * (int)sizeof(float);
        if (cudaFuncSetAttribute(sgemm_small_kernel<EPI_NONE, RM, TM, TN>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb) != cudaSuccess ||
            cudaFuncSetAttribute(sgemm_small_kernel<EPI_ACT, RM, TM, TN>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb) != cudaSuccess ||
            cudaFuncSetAttribute(sgemm_small_kernel<EPI_MULGRAD, RM, TM, TN>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb) != cudaSuccess ||
            cudaFuncSetAttribute(sgemm_small_kernel<EPI_SCALE3, RM, TM, TN>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb) != cudaSuccess) {
            tmdn_set_error("gemm: cudaFuncSetAttribute failed: %s", cudaGetErrorString(cudaGetLastError()));
            return TMDN_ECUDA;
        }
        attr_set = true;
    }
    constexpr unsigned threads = (16 * RM / TM) * (SBN / TN);
    switch (epi) {
        case EPI_NONE: launch_k(sgemm_small_kernel<EPI_NONE, RM, TM, TN>, grid, threads, smem, stream, b); break;
        case EPI_ACT: launch_k(sgemm_small_kernel<EPI_ACT, RM, TM, TN>, grid, threads, smem, stream, b); break;
        case EPI_MULGRAD: launch_k(sgemm_small_kernel<EPI_MULGRAD, RM, TM, TN>, grid, threads, smem, stream, b); break;
        case EPI_SCALE3: launch_k(sgemm_small_kernel<EPI_SCALE3, RM, TM, TN>, grid, threads, smem, stream, b); break;
        default: tmdn_set_error("gemm: bad epilogue"); return TMDN_EINVAL;
    }
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

static inline int small_gemm_launch(int epi, const GemmBatch& b, cudaStream_t stream, const char* tag) {
    KTimer ktimer__(tag, stream);
    int maxM = 0, maxN = 0;
    for (int i = 0; i < b.n; ++i) {
        if (b.p[i].M > maxM) maxM = b.p[i].M;
        if (b.p[i].N > maxN) maxN = b.p[i].N;
    }
    if (maxM == 0 || maxN == 0) return TMDN_OK;
    const dim3 grid((unsigned)cdiv(maxM, 16 * SMALL_RM), (unsigned)cdiv(maxN, SBN), (unsigned)b.n);
    const size_t smem = (size_t)(16 * SMALL_RM + SBN) * (b.p[0].K + 4) * sizeof(float);
    return small_gemm_launch_cfg<SMALL_RM, SMALL_TM, SMALL_TN>(epi, b, grid, smem, stream);
}

static inline GemmP gemm_desc(const float* A, int lda, const float* W, const float* bias, float* C, int ldc, int M,
                              int N, int K) {
    GemmP p;
    memset(&p, 0, sizeof(p));
    p.A = A; p.lda = lda; p.W = W; p.bias = bias; p.C = C; p.ldc = ldc;
    p.M = M; p.N = N; p.K = K;
    p.gs = 1; p.c0 = 0; p.rpa = 1; p.g = 0;
    p.w_static = 1;
    return p;
}

static inline int gemm_launch(int epi, const GemmBatch& b, cudaStream_t stream, const char* tag = "node_gemm") {
    if (small_gemm_ok(b)) return small_gemm_launch(epi, b, stream, strcmp(tag, "node_gemm") == 0 ? "node_gemm_small" : tag);
    KTimer ktimer__(tag, stream);
    int maxM = 0, maxN = 0;
    for (int i = 0; i < b.n; ++i) {
        if (b.p[i].M > maxM) maxM = b.p[i].M;
        if (b.p[i].N > maxN) maxN = b.p[i].N;
    }
    if (maxM == 0 || maxN == 0) return TMDN_OK;
    dim3 grid((unsigned)cdiv(maxM, GBM), (unsigned)cdiv(maxN, GBN), (unsigned)b.n);  // M tiles on x (2^31-1 limit)
    switch (epi) {
        case EPI_NONE: launch_k(sgemm_nt_kernel<EPI_NONE>, grid, GTHREADS, 0, stream, b); break;
        case EPI_ACT: launch_k(sgemm_nt_kernel<EPI_ACT>, grid, GTHREADS, 0, stream, b); break;
        case EPI_MULGRAD: launch_k(sgemm_nt_kernel<EPI_MULGRAD>, grid, GTHREADS, 0, stream, b); break;
        case EPI_SCALE3: launch_k(sgemm_nt_kernel<EPI_SCALE3>, grid, GTHREADS, 0, stream, b); break;
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
