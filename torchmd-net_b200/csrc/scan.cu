// scan.cu -- exclusive prefix sum over int32 (cell offsets, CSR row pointers) + error-message storage.
#include "common.cuh"
#include "scan.cuh"
#include <stdlib.h>

static thread_local char g_err[512] = "";

void tmdn_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* tmdn_last_error(void) { return g_err; }
extern "C" int tmdn_version(void) { return TMDN_VERSION; }

extern "C" int tmdn_is_stream_capturing(tmdn_stream_t stream) {
    cudaStreamCaptureStatus st;
    cudaError_t err = cudaStreamIsCapturing((cudaStream_t)stream, &st);
    if (err != cudaSuccess) {
        tmdn_set_error("cudaStreamIsCapturing: %s", cudaGetErrorString(err));
        return -1;
    }
    return st == cudaStreamCaptureStatusActive ? 1 : 0;
}

// ------------------------------------------------------------------ launch counter + per-kernel timing
#include <atomic>
#include <mutex>
#include <vector>
#include <map>
#include <string>
static std::atomic<long long> g_launches{0};
void tmdn_note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long long tmdn_launch_count(void) { return g_launches.load(); }

// ------------------------------------------------------------------ tuning knobs (tmdn_set_tuning)
// Defaults can be overridden at load time by the environment (TMDN_SMALL_GEMM, TMDN_LPT_ROWS, TMDN_PDL, TMDN_EDGE_PERSIST,
// TMDN_HOT_ROWS).  Every setting computes the same function; the knobs select kernels / launch shapes.
//   edge_persist: large systems run the TensorNet message-passing kernels as persistent CTAs (one per SM)
//   hot_rows:     number of table rows (from the far end of the grid) those kernels keep in shared memory
//   edge_variant: A/B instantiations of the persistent kernels (bit 0: forward 2 edges per batch; bit 1: backward 8 workers x 1 edge; bit 3: forward 1 x 8 workers);
//                 same sums in the same order
//   small_csr:    systems of <= 256 atoms build their CSR neighbour structure in ONE kernel (a cluster of 8 CTAs; same rows, same order)
static int env_or(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
int g_tmdn_tuning[TMDN_TUNE_COUNT] = {env_or("TMDN_SMALL_GEMM", 1), env_or("TMDN_LPT_ROWS", 1), env_or("TMDN_PDL", 0),
                                         env_or("TMDN_EDGE_PERSIST", 1), env_or("TMDN_HOT_ROWS", 0),
                                         env_or("TMDN_EDGE_VARIANT", 0), env_or("TMDN_SMALL_CSR", 1)};
extern "C" int tmdn_set_tuning(const char* key, int value) {
    static const char* names[TMDN_TUNE_COUNT] = {"small_gemm", "lpt_rows", "pdl", "edge_persist", "hot_rows", "edge_variant",
                                                 "small_csr"};
    if (!key) { tmdn_set_error("set_tuning: null key"); return -1; }
    for (int i = 0; i < TMDN_TUNE_COUNT; ++i)
        if (strcmp(key, names[i]) == 0) {
            const int old = g_tmdn_tuning[i];
            if (value >= 0) g_tmdn_tuning[i] = value;
            return old;
        }
    tmdn_set_error("set_tuning: unknown key '%s'", key);
    return -1;
}

namespace {
struct TimedSpan { std::string name; cudaEvent_t e0, e1; };
std::mutex g_prof_mu;
std::atomic<int> g_prof_on{0};
std::vector<TimedSpan> g_spans;
}  // namespace

KTimer::KTimer(const char* name, cudaStream_t s) : slot(-1), stream(s) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess || st != cudaStreamCaptureStatusNone) return;
    TimedSpan sp;
    sp.name = name;
    if (cudaEventCreate(&sp.e0) != cudaSuccess) return;
    if (cudaEventCreate(&sp.e1) != cudaSuccess) { cudaEventDestroy(sp.e0); return; }
    cudaEventRecord(sp.e0, s);
    std::lock_guard<std::mutex> lock(g_prof_mu);
    g_spans.push_back(sp);
    slot = (int)g_spans.size() - 1;
}
KTimer::~KTimer() {
    if (slot < 0) return;
    std::lock_guard<std::mutex> lock(g_prof_mu);
    if (slot < (int)g_spans.size()) cudaEventRecord(g_spans[slot].e1, stream);
}

extern "C" int tmdn_profile_enable(int on) {
    std::lock_guard<std::mutex> lock(g_prof_mu);
    for (auto& sp : g_spans) { cudaEventDestroy(sp.e0); cudaEventDestroy(sp.e1); }
    g_spans.clear();
    g_prof_on.store(on ? 1 : 0);
    return TMDN_OK;
}
// Writes "name count total_ms\n" per kernel family (sorted by name) into buf; synchronises the recorded events.
extern "C" int tmdn_profile_report(char* buf, size_t buf_bytes) {
    TMDN_CHECK_ARG(buf && buf_bytes > 0, "profile_report: null buffer");
    std::lock_guard<std::mutex> lock(g_prof_mu);
    std::map<std::string, std::pair<long long, double>> agg;
    for (auto& sp : g_spans) {
        float ms = 0.f;
        TMDN_CUDA(cudaEventSynchronize(sp.e1));
        TMDN_CUDA(cudaEventElapsedTime(&ms, sp.e0, sp.e1));
        auto& a = agg[sp.name];
        a.first += 1;
        a.second += ms;
    }
    size_t off = 0;
    buf[0] = 0;
    for (auto& kv : agg) {
        int n = snprintf(buf + off, buf_bytes - off, "%s %lld %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
        if (n < 0 || (size_t)n >= buf_bytes - off) break;
        off += (size_t)n;
    }
    return TMDN_OK;
}

namespace {
constexpr int SCAN_THREADS = 1024;
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// out[i] = sum_{k<i} in[k] within the tile (+ optional carry added later); tile total -> sums[blockIdx.x].
// Elements with index >= n_in read as 0; outputs are written for index < n_out.
__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_kernel(const int* in, int* out,
                                                                 int64_t n_in, int64_t n_out, int* __restrict__ sums) {
    pdl_sync();
    __shared__ int s_warp[32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)t * SCAN_ITEMS;
    int v[SCAN_ITEMS];
    int tot = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        int64_t i = base + k;
        v[k] = (i < n_in) ? in[i] : 0;
        tot += v[k];
    }
    int incl = warp_incl_scan_i(tot, lane);
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    if (w == 0) {
        int x = s_warp[lane];
        int xi = warp_incl_scan_i(x, lane);
        s_warp[lane] = xi - x;  // exclusive warp offsets
        if (lane == 31 && sums) sums[blockIdx.x] = xi;
    }
    __syncthreads();
    int run = s_warp[w] + incl - tot;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        int64_t i = base + k;
        if (i < n_out) out[i] = run;
        run += v[k];
    }
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_add_kernel(int* __restrict__ out, int64_t n_out,
                                                                const int* __restrict__ carry) {
    pdl_sync();
    const int c = carry[blockIdx.x];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        int64_t i = base + k;
        if (i < n_out) out[i] += c;
    }
}
}  // namespace

size_t tmdn_scan_tmp_ints(int64_t n_out) {
    size_t total = 0;
    int64_t n = n_out;
    while (n > SCAN_TILE) {
        n = cdiv(n, SCAN_TILE);
        total += align_up((size_t)n * sizeof(int)) / sizeof(int);
    }
    return total + 64;
}

// out[0..n_in] (n_in+1 entries) = exclusive scan of in[0..n_in); out[n_in] = total.  in may alias out.
int tmdn_scan_exclusive_i32(const int* in, int* out, int64_t n_in, int* tmp, cudaStream_t stream) {
    const int64_t n_out = n_in + 1;
    const int64_t nb = cdiv(n_out, SCAN_TILE);
    if (nb == 1) {
        launch_k(scan_tile_kernel, 1, SCAN_THREADS, 0, stream, in, out, n_in, n_out, nullptr);
        TMDN_LAUNCH_CHECK();
        return TMDN_OK;
    }
    int* sums = tmp;
    int* next_tmp = tmp + align_up((size_t)nb * sizeof(int)) / sizeof(int);
    launch_k(scan_tile_kernel, (unsigned)nb, SCAN_THREADS, 0, stream, in, out, n_in, n_out, sums);
    TMDN_LAUNCH_CHECK();
    // exclusive scan of the tile totals, in place (only nb entries are needed)
    {
        const int64_t nb2 = cdiv(nb, SCAN_TILE);
        if (nb2 == 1) {
            launch_k(scan_tile_kernel, 1, SCAN_THREADS, 0, stream, sums, sums, nb, nb, nullptr);
            TMDN_LAUNCH_CHECK();
        } else {
            int* sums2 = next_tmp;
            launch_k(scan_tile_kernel, (unsigned)nb2, SCAN_THREADS, 0, stream, sums, sums, nb, nb, sums2);
            TMDN_LAUNCH_CHECK();
            TMDN_CHECK_ARG(nb2 <= SCAN_TILE, "scan: too many elements (%lld)", (long long)n_in);
            launch_k(scan_tile_kernel, 1, SCAN_THREADS, 0, stream, sums2, sums2, nb2, nb2, nullptr);
            TMDN_LAUNCH_CHECK();
            launch_k(scan_add_kernel, (unsigned)nb2, SCAN_THREADS, 0, stream, sums, nb, sums2);
            TMDN_LAUNCH_CHECK();
        }
    }
    launch_k(scan_add_kernel, (unsigned)nb, SCAN_THREADS, 0, stream, out, n_out, sums);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}
