// common.cuh -- shared host/device helpers for libtmdn (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdarg.h>
#include "../../include/tmdn.h"

// ------------------------------------------------------------------ error plumbing (never throws)
void tmdn_set_error(const char* fmt, ...);
// every kernel launch of the library goes through TMDN_LAUNCH_CHECK, which also counts it (tmdn_launch_count)
void tmdn_note_launch();
// Optional per-kernel-family device timing (tmdn_profile_*): a KTimer brackets the launches issued in its scope
// with a CUDA event pair on `stream` when profiling is enabled and the stream is not being captured.
struct KTimer {
    KTimer(const char* name, cudaStream_t stream);
    ~KTimer();
    int slot;
    cudaStream_t stream;
};

// process-wide tuning knobs (tmdn_set_tuning, defined in scan.cu)
enum { TMDN_TUNE_SMALL_GEMM = 0, TMDN_TUNE_LPT_ROWS = 1, TMDN_TUNE_PDL = 2, TMDN_TUNE_EDGE_PERSIST = 3, TMDN_TUNE_HOT_ROWS = 4,
       TMDN_TUNE_EDGE_VARIANT = 5, TMDN_TUNE_SMALL_CSR = 6, TMDN_TUNE_COUNT = 7 };
extern int g_tmdn_tuning[TMDN_TUNE_COUNT];

#define TMDN_CHECK_ARG(cond, ...)                 \
    do {                                          \
        if (!(cond)) {                            \
            tmdn_set_error(__VA_ARGS__);          \
            return TMDN_EINVAL;                   \
        }                                         \
    } while (0)

#define TMDN_CUDA(call)                                                                         \
    do {                                                                                        \
        cudaError_t err__ = (call);                                                             \
        if (err__ != cudaSuccess) {                                                             \
            tmdn_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,                 \
                           cudaGetErrorString(err__));                                          \
            return TMDN_ECUDA;                                                                  \
        }                                                                                       \
    } while (0)

#define TMDN_LAUNCH_CHECK()                                                                     \
    do {                                                                                        \
        cudaError_t err__ = cudaPeekAtLastError();                                              \
        tmdn_note_launch();                                                                     \
        if (err__ != cudaSuccess) {                                                             \
            tmdn_set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__,             \
                           cudaGetErrorString(err__));                                          \
            return TMDN_ECUDA;                                                                  \
        }                                                                                       \
    } while (0)

#define TMDN_TRY(call)                 \
    do {                               \
        int rc__ = (call);             \
        if (rc__ != TMDN_OK) return rc__; \
    } while (0)

// ------------------------------------------------------------------ kernel launch (programmatic dependent launch)
// Every kernel of the library starts with pdl_sync(): it lets the NEXT kernel of the stream be scheduled as soon as
// all CTAs of this one are resident (griddepcontrol.launch_dependents) and then waits until the PREVIOUS kernel has
// completed and its writes are visible (griddepcontrol.wait).  Launched through launch_k with the programmatic
// stream-serialisation attribute, consecutive kernels overlap launch latency, CTA scheduling and the prologue before
// pdl_sync with the tail of their predecessor; data dependencies are unchanged because nothing is read or written
// before the wait and completion is transitive along the stream.  Without the attribute both instructions are
// no-ops.  A step of a small system is ~45 dependent launches of 3-30 us each, so this is where its time goes.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                   Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_tmdn_tuning[TMDN_TUNE_PDL] ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(static_cast<Args&&>(args))...);
}
#endif

// SM count of the current device (B200: 148); queried once -- one process drives one GPU model
static inline int tmdn_sm_count() {
    static const int n = [] {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
            v = 148;
        return v;
    }();
    return n;
}
static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Bump allocator over the caller's workspace.
struct Arena {
    char* base;
    size_t cap, off;
    bool ok;
    Arena(void* b, size_t c) : base((char*)b), cap(c), off(0), ok(true) {}
    template <typename T> T* get(size_t n) {
        size_t bytes = align_up(n * sizeof(T));
        size_t o = off;
        off += bytes;
        if (base && off > cap) ok = false;
        return base ? (T*)(base + o) : (T*)nullptr;
    }
};

// ------------------------------------------------------------------ device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// [a, b) = the index range of molecule m in the sorted batch vector, found by the whole CTA (128 threads): every round
// probes 128 evenly spaced entries at once, so a range costs ceil(log128 N) dependent memory round trips instead of the
// log2 N of a binary search (the one-CTA-per-molecule reductions were 9 us at 166 atoms, all of it those round trips).
// Must be called by all threads of a 128-thread CTA.
__device__ __forceinline__ int block_lower_bound_128(const int64_t* __restrict__ batch, int lo, int hi, long long v) {
    while (hi > lo) {
        const int step = (hi - lo + 127) >> 7;
        const long long idx = (long long)lo + (long long)threadIdx.x * step;
        const bool less = idx < hi && batch[idx] < v;
        const int T = __syncthreads_count(less);
        if (T == 0) {
            hi = lo;
        } else {
            const long long top = (long long)lo + (long long)T * step;
            lo = lo + (T - 1) * step + 1;
            hi = top < hi ? (int)top : hi;
        }
    }
    return lo;
}
__device__ __forceinline__ void block_molecule_range(const int64_t* __restrict__ batch, int N, long long m, int& a, int& b) {
    a = block_lower_bound_128(batch, 0, N, m);
    b = block_lower_bound_128(batch, a, N, m + 1);
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_incl_scan_i(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// Activations (models/utils.py act_class_mapping) and their derivatives.
// the correctly rounded reciprocal equals 1.0f / y bit for bit (y in [1, inf]) without the division's slow path
__device__ __forceinline__ float sigmoidf_(float x) { return __frcp_rn(1.0f + expf(-x)); }
__device__ __forceinline__ float act_apply(int act, float x) {
    switch (act) {
        case TMDN_ACT_SILU: return x * sigmoidf_(x);
        case TMDN_ACT_TANH: return tanhf(x);
        case TMDN_ACT_SIGMOID: return sigmoidf_(x);
        default: {  // shifted softplus, F.softplus threshold 20
            float sp = x > 20.0f ? x : log1pf(expf(x));
            return sp - 0.69314718055994530942f;
        }
    }
}
// value and derivative from ONE evaluation of the transcendental (same expressions as act_apply / act_grad, so the
// results are bit-identical to calling both)
__device__ __forceinline__ void act_both(int act, float x, float& a, float& g) {
    switch (act) {
        case TMDN_ACT_SILU: { const float s = sigmoidf_(x); a = x * s; g = s * (1.0f + x * (1.0f - s)); break; }
        case TMDN_ACT_TANH: { const float y = tanhf(x); a = y; g = 1.0f - y * y; break; }
        case TMDN_ACT_SIGMOID: { const float s = sigmoidf_(x); a = s; g = s * (1.0f - s); break; }
        default: { a = act_apply(act, x); g = sigmoidf_(x); }
    }
}
__device__ __forceinline__ float act_grad(int act, float x) {
    switch (act) {
        case TMDN_ACT_SILU: { float s = sigmoidf_(x); return s * (1.0f + x * (1.0f - s)); }
        case TMDN_ACT_TANH: { float y = tanhf(x); return 1.0f - y * y; }
        case TMDN_ACT_SIGMOID: { float s = sigmoidf_(x); return s * (1.0f - s); }
        default: return sigmoidf_(x);
    }
}
