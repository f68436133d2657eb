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
    const dim3 grid2((unsigned)cdiv(5 * N, 32), 4, 3), grid1((unsigned)cdiv(5 * N, 16), 4, 3);
    const char* names[] = {"empty", "loads32", "loads16", "full (production choice)", "algebra 83 CTAs", "empty 1 CTA",
                           "tile32 2x2 (256 thr)", "tile32 4x2 (128 thr)", "tile32 2x4 (128 thr)", "tile32 4x4 (64 thr)",
                           "tile32 8x2 (64 thr)", "tile16 1x2 (256 thr)", "tile16 2x2 (128 thr)", "tile16 4x2 (64 thr)",
                           "tile16 2x4 (64 thr)", "tile16 4x4 (32 thr)"};
    std::vector<float> ref(nx), got(nx);
    for (int variant = 0; variant < 16; ++variant) {
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
            }
        };
        // one launch from the same input: all GEMM variants must give the bits of variant 6 (the first production tiling)
        CK(cudaMemcpy(x, hx.data(), nx * 4, cudaMemcpyHostToDevice));
        launch(0);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(got.data(), y, nx * 4, cudaMemcpyDeviceToHost));
        if (variant == 6) ref = got;
        const bool same = variant >= 6 ? memcmp(ref.data(), got.data(), nx * 4) == 0 : true;
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
