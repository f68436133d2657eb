// microbench_gather.cu -- the gather loop of the TensorNet edge kernels in isolation (r01i).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/mb_gather.bin profiles/microbench_gather.cu && profiles/mb_gather.bin   (the .bin is git-ignored)
// One CTA per row (R rows of D edges), block = (H channels, S slices); a thread walks the edges r0+sl, r0+sl+S, ...
// and per edge loads  T float4 table entries (row picked by the edge's distance) and  Y  floats of the neighbour row.
// Prints the kernel time for a grid of (S, NB, T, Y) so that the cost of an iteration can be attributed.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int H = 128;

template <int NB, int T, int Y, int MAXT>
__global__ void __launch_bounds__(MAXT) gather_kernel(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                                      const float* __restrict__ dist, const float4* __restrict__ tab,
                                                      int n_int, float inv_h, const float* __restrict__ y,
                                                      float* __restrict__ out, int n_rows) {
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y, i = blockIdx.x;
    if (i >= n_rows) return;
    float acc = 0.f;
    const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
    for (int e0 = r0 + sl; e0 < r1; e0 += NB * S) {
        float4 t[NB][T > 0 ? T : 1];
        float v[NB][Y > 0 ? Y : 1];
        bool on[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const int e = e0 + b * S;
            on[b] = e < r1;
            const int ee = on[b] ? e : e0;
            const int j = col[ee];
            int k = (int)(dist[ee] * inv_h);
            k = k > n_int - 1 ? n_int - 1 : k;
#pragma unroll
            for (int g = 0; g < T; ++g) t[b][g] = __ldg(&tab[((size_t)k * 3 + g) * H + h]);
#pragma unroll
            for (int c = 0; c < Y; ++c) v[b][c] = y[((size_t)j * 9 + c) * H + h];
        }
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            if (!on[b]) continue;
#pragma unroll
            for (int g = 0; g < T; ++g) acc += t[b][g].x + t[b][g].y * t[b][g].z + t[b][g].w;
#pragma unroll
            for (int c = 0; c < Y; ++c) acc += v[b][c];
        }
    }
    // slice combine left out on purpose: the loop is what is being measured
    if (acc == 123.456f) out[(size_t)i * H + h] = acc;
    if (sl == 0) out[(size_t)i * H + h] = acc;
}

template <int NB, int T, int Y, int MAXT>
float run(int S, const int* rp, const int* col, const float* d, const float4* tab, int n_int, float inv_h, const float* y,
          float* out, int R, float* flush, size_t flush_bytes, bool do_flush) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e9f;
    for (int rep = 0; rep < 6; ++rep) {
        if (do_flush) cudaMemsetAsync(flush, 0, flush_bytes);
        cudaEventRecord(a);
        gather_kernel<NB, T, Y, MAXT><<<R, dim3(H, S)>>>(rp, col, d, tab, n_int, inv_h, y, out, R);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    return best * 1e3f;
}

int main(int argc, char** argv) {
    const int R = argc > 1 ? atoi(argv[1]) : 166, D = argc > 2 ? atoi(argv[2]) : 120, n_int = 4096;
    std::vector<int> rp(R + 1), col((size_t)R * D);
    std::vector<float> d((size_t)R * D);
    srand(1);
    for (int i = 0; i <= R; ++i) rp[i] = i * D;
    for (size_t e = 0; e < col.size(); ++e) { col[e] = rand() % R; d[e] = 1.0f + 9.0f * (rand() / (float)RAND_MAX) * 0.999f; }
    int *drp, *dcol;
    float *dd, *dy, *dout, *flush;
    float4* dtab;
    const size_t flush_bytes = 512u << 20;
    cudaMalloc(&drp, rp.size() * 4); cudaMalloc(&dcol, col.size() * 4); cudaMalloc(&dd, d.size() * 4);
    cudaMalloc(&dtab, (size_t)n_int * 3 * H * 16); cudaMalloc(&dy, (size_t)R * 18 * H * 4); cudaMalloc(&dout, (size_t)R * H * 4);
    cudaMalloc(&flush, flush_bytes);
    cudaMemcpy(drp, rp.data(), rp.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dcol, col.data(), col.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dd, d.data(), d.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(dtab, 0, (size_t)n_int * 3 * H * 16);
    cudaMemset(dy, 0, (size_t)R * 18 * H * 4);
    const float inv_h = n_int / 10.0f;
    printf("rows %d, edges/row %d, H %d  (times in us, best of 5; warm = no L2 flush, cold = 512 MiB memset before)\n", R, D, H);
    printf("%-32s %8s %8s %8s %8s\n", "variant", "S=1", "S=2", "S=4", "S=8");
#define ROW(NB, T, Y, MAXT, flushf)                                                                                \
    {                                                                                                              \
        printf("NB=%d tab=%d y=%2d regs<=%3d %-4s", NB, T, Y, 65536 / MAXT > 255 ? 255 : 65536 / MAXT, flushf ? "cold" : "warm"); \
        for (int S = 1; S <= 8; S *= 2)                                                                            \
            if (S * H <= MAXT)                                                                                     \
                printf(" %8.1f", run<NB, T, Y, MAXT>(S, drp, dcol, dd, dtab, n_int, inv_h, dy, dout, R, flush, flush_bytes, flushf)); \
        printf("\n");                                                                                              \
    }
    ROW(1, 0, 0, 1024, false) ROW(1, 3, 0, 1024, false) ROW(1, 0, 9, 1024, false) ROW(1, 3, 9, 1024, false)
    ROW(1, 3, 9, 512, false) ROW(2, 3, 9, 512, false) ROW(4, 3, 9, 512, false) ROW(4, 3, 9, 256, false) ROW(8, 3, 9, 256, false)
    ROW(1, 3, 18, 512, false) ROW(2, 3, 18, 512, false) ROW(4, 3, 18, 256, false)
    ROW(1, 3, 9, 1024, true) ROW(2, 3, 9, 512, true)
    cudaError_t err = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(err));
    return 0;
}
