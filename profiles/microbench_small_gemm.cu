// microbench_small_gemm.cu -- where do the ~10 us of a small-system channel mix go?  (measurement aid, not part of the library)
//
// The 166-atom step spends 40 % of its time in 18 GEMM launches of 5-11 us each (profiles/r02i_launches_chignolin166.txt).
// This program times, as a dependent chain inside one CUDA graph (the way the step runs them), on the shape of the
// TensorNet channel mix (three problems I / A / S over an [N,9,128] tensor, 128 x 128 weights):
//   empty     a kernel with the same grid, block and shared memory that does nothing
//   loads     the cp.async panel fetch of sgemm_small_kernel + one store per thread
//   full      the production kernel (gemm_launch -> sgemm_small_kernel<EPI_NONE, 2>)
//   RMxTMxTN  sgemm_small_kernel with 16*RM-row tiles and TM x TN outputs per thread (all give the same bits)
//   algebra   an elementwise pass over the tensor (the shape of the node-algebra kernels: 83 CTAs x 256 threads)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I torchmd-net_b200/csrc profiles/microbench_small_gemm.cu \
//        -L torchmd-net_b200/torchmdnet/lib -ltmdn -Xlinker -rpath -Xlinker $PWD/torchmd-net_b200/torchmdnet/lib -o /tmp/mb_gemm
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <vector>
#include "gemm.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) k_empty(GemmBatch batch) {
    extern __shared__ float sm[];
    if (batch.n < 0) sm[threadIdx.x] = 0.f;
}

template <int RM>
__global__ void __launch_bounds__(256) k_loads(GemmBatch batch) {
    extern __shared__ __align__(16) float sg[];
    constexpr int BM = 16 * RM;
    const GemmP& P = batch.p[blockIdx.z];
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * SBN;
    if (m0 >= P.M || n0 >= P.N) return;
    const int K = P.K, ld = K + 4, kv = K >> 2;
    float* As = sg;
    float* Ws = sg + BM * ld;
    const int t = threadIdx.x, lane = t & 31, wrp = t >> 5;
    for (int row = wrp; row < SBN; row += 8) {
        const float* src = P.W + (size_t)(n0 + row) * K;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(Ws + row * ld);
        for (int k4 = lane; k4 < kv; k4 += 32)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4) : "memory");
    }
    for (int row = wrp; row < BM; row += 8) {
        const int m = m0 + row;
        const bool ok = m < P.M;
        const size_t prow = !ok ? 0 : (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs;
        const float* src = P.A + prow * P.lda;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(As + row * ld);
        for (int k4 = lane; k4 < kv; k4 += 32)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4) : "memory");
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    const int tx = t & 15, ty = t >> 4;
    const int m = m0 + ty;
    if (m < P.M) {
        const size_t prow = (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs;
        P.C[prow * P.ldc + n0 + tx] = As[ty * ld + tx] + Ws[tx * ld + ty];
    }
}


// the production kernel (16 x 32 tile, 2 x 2 per thread, EPI_NONE) with its phases switchable: which one does a launch wait for?
template <bool LOAD, bool COMPUTE, bool FULL_EPI>
__global__ void __launch_bounds__(128) k_probe(GemmBatch batch) {
    extern __shared__ __align__(16) float sg[];
    constexpr int BM = 16, TROWS = 8, TCOLS = 16, NW = 4;
    const GemmP& P = batch.p[blockIdx.z];
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * SBN;
    if (m0 >= P.M || n0 >= P.N) return;
    const int K = P.K, ld = K + 4, kv = K >> 2;
    float* As = sg;
    float* Ws = sg + BM * ld;
    const int t = threadIdx.x, lane = t & 31, wrp = t >> 5;
    if (LOAD) {
        for (int row = wrp; row < SBN; row += NW) {
            const float* src = P.W + (size_t)(n0 + row) * K;
            const unsigned dst = (unsigned)__cvta_generic_to_shared(Ws + row * ld);
            for (int k4 = lane; k4 < kv; k4 += 32)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4) : "memory");
        }
        for (int row = wrp; row < BM; row += NW) {
            const int m = m0 + row;
            const bool ok = m < P.M;
            const size_t prow = !ok ? 0 : (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs;
            const float* src = P.A + prow * P.lda;
            const unsigned dst = (unsigned)__cvta_generic_to_shared(As + row * ld);
            const int bytes = ok ? 16 : 0;
            for (int k4 = lane; k4 < kv; k4 += 32)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4), "r"(bytes) : "memory");
        }
        asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const int tx = t % TCOLS, ty = t / TCOLS;
    float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
    const float* a0 = As + ty * ld;
    const float* w0 = Ws + tx * ld;
    if (COMPUTE) {
#pragma unroll 2
        for (int k = 0; k < K; k += 4) {
            float4 a[2], w[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) a[r] = *reinterpret_cast<const float4*>(a0 + r * TROWS * ld + k);
#pragma unroll
            for (int c = 0; c < 2; ++c) w[c] = *reinterpret_cast<const float4*>(w0 + c * TCOLS * ld + k);
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    acc[r][c] = fmaf(a[r].x, w[c].x, acc[r][c]);
                    acc[r][c] = fmaf(a[r].y, w[c].y, acc[r][c]);
                    acc[r][c] = fmaf(a[r].z, w[c].z, acc[r][c]);
                    acc[r][c] = fmaf(a[r].w, w[c].w, acc[r][c]);
                }
        }
    } else {
        acc[0][0] = a0[0]; acc[0][1] = w0[0]; acc[1][0] = a0[1]; acc[1][1] = w0[1];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int m = m0 + ty + TROWS * r;
        if (m >= P.M) continue;
        const size_t prow = (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            if (!FULL_EPI && (r | c)) continue;
            P.C[prow * P.ldc + n0 + tx + TCOLS * c] = acc[r][c];
        }
    }
}

// tensor-core probe: same tile and panels, warp w = the 16 x 8 block of columns 8w..8w+7; SPLIT = 1: plain tf32, 3: two-term
// split of both operands with three products; partial sums of 32-wide K chunks are added in fp32 outside the MMA
__device__ __forceinline__ void probe_split(float x, unsigned& hi, unsigned& lo) {
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
    const float r = x - __uint_as_float(hi);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}
__device__ __forceinline__ void probe_mma(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
template <int SPLIT>
__global__ void __launch_bounds__(128) k_probe_mma(GemmBatch batch) {
    extern __shared__ __align__(16) float sg[];
    constexpr int BM = 16, NW = 4;
    const GemmP& P = batch.p[blockIdx.z];
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * SBN;
    if (m0 >= P.M || n0 >= P.N) return;
    const int K = P.K, ld = K + 4, kv = K >> 2;
    float* As = sg;
    float* Ws = sg + BM * ld;
    const int t = threadIdx.x, lane = t & 31, wrp = t >> 5;
    for (int row = wrp; row < SBN; row += NW) {
        const float* src = P.W + (size_t)(n0 + row) * K;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(Ws + row * ld);
        for (int k4 = lane; k4 < kv; k4 += 32)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4) : "memory");
    }
    for (int row = wrp; row < BM; row += NW) {
        const int m = m0 + row;
        const bool ok = m < P.M;
        const size_t prow = !ok ? 0 : (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs;
        const float* src = P.A + prow * P.lda;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(As + row * ld);
        const int bytes = ok ? 16 : 0;
        for (int k4 = lane; k4 < kv; k4 += 32)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16u * k4), "l"(src + 4 * k4), "r"(bytes) : "memory");
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    const int g = lane >> 2, q = lane & 3, cb = wrp * 8;
    const float* ap = As + g * ld + q;
    const float* ap8 = ap + 8 * ld;
    const float* wp = Ws + (cb + g) * ld + q;
    float sum[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k0 = 0; k0 < K; k0 += 32) {
        float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f}, c2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = k0; k < k0 + 32; k += 8) {
            unsigned ah[4], al[4], bh[2], bl[2];
            if (SPLIT == 3) {
                probe_split(ap[k], ah[0], al[0]); probe_split(ap8[k], ah[1], al[1]);
                probe_split(ap[k + 4], ah[2], al[2]); probe_split(ap8[k + 4], ah[3], al[3]);
                probe_split(wp[k], bh[0], bl[0]); probe_split(wp[k + 4], bh[1], bl[1]);
                probe_mma(c1, al, bh); probe_mma(c2, ah, bl); probe_mma(c0, ah, bh);
            } else {
                ah[0] = __float_as_uint(ap[k]); ah[1] = __float_as_uint(ap8[k]); ah[2] = __float_as_uint(ap[k + 4]);
                ah[3] = __float_as_uint(ap8[k + 4]); bh[0] = __float_as_uint(wp[k]); bh[1] = __float_as_uint(wp[k + 4]);
                probe_mma(c0, ah, bh);
            }
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) sum[e] += (c1[e] + c2[e]) + c0[e];
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int m = m0 + g + 8 * (e >> 1), n = n0 + cb + 2 * q + (e & 1);
        if (m >= P.M) continue;
        const size_t prow = (size_t)(m / P.gs) * P.rpa + P.c0 + m % P.gs;
        P.C[prow * P.ldc + n] = sum[e];
    }
}

__global__ void __launch_bounds__(256) k_algebra(const float* __restrict__ x, float* __restrict__ y, int n4) {
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n4; i += gridDim.x * 256) {
        float4 v = reinterpret_cast<const float4*>(x)[i];
        v.x = v.x * 1.0001f + 0.5f; v.y *= 0.999f; v.z += 1e-3f; v.w = v.w * v.x;
        reinterpret_cast<float4*>(y)[i] = v;
    }
}

static GemmBatch mix_batch(const float* x, const float* w, float* y, int N, int H) {
    GemmBatch b;
    memset(&b, 0, sizeof(b));
    const int gs[3] = {1, 3, 5}, c0[3] = {0, 1, 4};
    for (int g = 0; g < 3; ++g) {
        GemmP p = gemm_desc(x, H, w + (size_t)g * H * H, nullptr, y, H, N * gs[g], H, H);
        p.gs = gs[g]; p.c0 = c0[g]; p.rpa = 9; p.g = g;
        b.p[g] = p;
    }
    b.n = 3;
    return b;
}

int main(int argc, char** argv) {
    const int N = argc > 1 ? atoi(argv[1]) : 166, H = 128, CHAIN = 64, REPS = 30;
    float *x, *y, *w;
    const size_t nx = (size_t)N * 9 * H;
    CK(cudaMalloc(&x, nx * 4)); CK(cudaMalloc(&y, nx * 4)); CK(cudaMalloc(&w, 3 * H * H * 4));
    std::vector<float> hx(nx), hw(3 * H * H);
    for (auto& v : hx) v = (rand() % 2001 - 1000) * 1e-3f;
    for (auto& v : hw) v = (rand() % 2001 - 1000) * 1e-4f;
    CK(cudaMemcpy(x, hx.data(), nx * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(y, hx.data(), nx * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(w, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice));
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    const size_t smem2 = (size_t)(32 + SBN) * (H + 4) * 4, smem1 = (size_t)(16 + SBN) * (H + 4) * 4;
    CK(cudaFuncSetAttribute(k_empty, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    CK(cudaFuncSetAttribute(k_loads<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    CK(cudaFuncSetAttribute(k_loads<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    // (the probes need 25 KB: below the 48 KB that needs the attribute)
    const dim3 grid2((unsigned)cdiv(5 * N, 32), 4, 3), grid1((unsigned)cdiv(5 * N, 16), 4, 3);
    const char* names[] = {"empty", "loads32", "loads16", "full (production choice)", "algebra 83 CTAs", "empty 1 CTA",
                           "tile32 2x2 (256 thr)", "tile32 4x2 (128 thr)", "tile32 2x4 (128 thr)", "tile32 4x4 (64 thr)",
                           "tile32 8x2 (64 thr)", "tile16 1x2 (256 thr)", "tile16 2x2 (128 thr)", "tile16 4x2 (64 thr)",
                           "tile16 2x4 (64 thr)", "tile16 4x4 (32 thr)", "probe: load + compute + 4 stores", "probe: load + 4 stores",
                           "probe: load + 1 store", "probe: compute + 4 stores (no load)", "probe: 4 stores only", "probe: mma.sync tf32 (1 product)",
                           "probe: mma.sync tf32 split (3 products)"};
    std::vector<float> ref(nx), got(nx);
    for (int variant = 0; variant < 23; ++variant) {
        auto launch = [&](int it) {
            const float* in = (it & 1) ? y : x;
            float* out = (it & 1) ? x : y;
            GemmBatch b = mix_batch(in, w, out, N, H);
            switch (variant) {
                case 0: k_empty<<<grid2, 256, smem2, s>>>(b); break;
                case 1: k_loads<2><<<grid2, 256, smem2, s>>>(b); break;
                case 2: k_loads<1><<<grid1, 256, smem1, s>>>(b); break;
                case 3: gemm_launch(EPI_NONE, b, s); break;
                case 4: k_algebra<<<83, 256, 0, s>>>(in, out, (int)(nx / 4)); break;
                case 5: k_empty<<<1, 256, 0, s>>>(b); break;
                case 6: small_gemm_launch_cfg<2, 2, 2>(EPI_NONE, b, grid2, smem2, s); break;
                case 7: small_gemm_launch_cfg<2, 4, 2>(EPI_NONE, b, grid2, smem2, s); break;
                case 8: small_gemm_launch_cfg<2, 2, 4>(EPI_NONE, b, grid2, smem2, s); break;
                case 9: small_gemm_launch_cfg<2, 4, 4>(EPI_NONE, b, grid2, smem2, s); break;
                case 10: small_gemm_launch_cfg<2, 8, 2>(EPI_NONE, b, grid2, smem2, s); break;
                case 11: small_gemm_launch_cfg<1, 1, 2>(EPI_NONE, b, grid1, smem1, s); break;
                case 12: small_gemm_launch_cfg<1, 2, 2>(EPI_NONE, b, grid1, smem1, s); break;
                case 13: small_gemm_launch_cfg<1, 4, 2>(EPI_NONE, b, grid1, smem1, s); break;
                case 14: small_gemm_launch_cfg<1, 2, 4>(EPI_NONE, b, grid1, smem1, s); break;
                case 15: small_gemm_launch_cfg<1, 4, 4>(EPI_NONE, b, grid1, smem1, s); break;
                case 16: k_probe<true, true, true><<<grid1, 128, smem1, s>>>(b); break;
                case 17: k_probe<true, false, true><<<grid1, 128, smem1, s>>>(b); break;
                case 18: k_probe<true, false, false><<<grid1, 128, smem1, s>>>(b); break;
                case 19: k_probe<false, true, true><<<grid1, 128, smem1, s>>>(b); break;
                case 20: k_probe<false, false, true><<<grid1, 128, smem1, s>>>(b); break;
                case 21: k_probe_mma<1><<<grid1, 128, smem1, s>>>(b); break;
                case 22: k_probe_mma<3><<<grid1, 128, smem1, s>>>(b); break;
            }
        };
        // one launch from the same input: all GEMM variants must give the bits of variant 6 (the first production tiling)
        CK(cudaMemcpy(x, hx.data(), nx * 4, cudaMemcpyHostToDevice));
        launch(0);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(got.data(), y, nx * 4, cudaMemcpyDeviceToHost));
        if (variant == 6) ref = got;
        const bool same = (variant >= 6 && variant <= 16) ? memcmp(ref.data(), got.data(), nx * 4) == 0 : true;
        if (variant >= 21) {
            double worst = 0, scale = 0;
            for (size_t q = 0; q < nx; ++q) { worst = fmax(worst, fabs((double)got[q] - ref[q])); scale = fmax(scale, fabs((double)ref[q])); }
            printf("    max |mma - ffma| = %.3g of max |y| = %.3g\n", worst, scale);
        }
        for (int it = 0; it < 4; ++it) launch(it);
        CK(cudaStreamSynchronize(s));
        cudaGraph_t g;
        cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        for (int it = 0; it < CHAIN; ++it) launch(it);
        CK(cudaStreamEndCapture(s, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        for (int r = 0; r < 3; ++r) CK(cudaGraphLaunch(ge, s));
        CK(cudaEventRecord(e0, s));
        for (int r = 0; r < REPS; ++r) CK(cudaGraphLaunch(ge, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%-28s N=%d  %.2f us per launch%s\n", names[variant], N, 1e3f * ms / (REPS * CHAIN), same ? "" : "  BITS DIFFER");
        CK(cudaGetLastError());
        cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
    }
    return 0;
}
