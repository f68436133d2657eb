// csr_direct.cu -- neighbour search straight into the CSR-by-target edge structure the model kernels consume.
//
// The reference materialises a padded pair list (neighbors/common.cuh:64-116) that its models then index with
// gathers and atomics.  The fused models here never need that list: one warp per target atom walks its candidate
// sources (its molecule's index range for brute, the 27 neighbouring cells for cell), counts the hits with
// ballot/popc, a scan turns counts into row pointers, and a second identical walk writes (col, geom) in place.
// No atomics on the output, no sort, and the edge order inside a row is a pure function of the positions
// (ascending source index for brute; cell-traversal order of nbr_cell_tile.cuh with cells internally sorted by index for cell), so
// every downstream segmented sum is bit-reproducible.  Same pair set and same rounding as tmdn_neighbors_f32.
#include <cooperative_groups.h>
#include <mutex>
#include "common.cuh"
#include "nbr_cell_tile.cuh"
#include "scan.cuh"

namespace {
using namespace nbr;

constexpr int CD_WARPS = 8;  // warps (target atoms) per CTA

struct EdgeOut {
    const int* row_ptr;  // null in the counting pass
    int* row_cnt;        // counting pass output
    int32_t* col;
    float4* geom;
    long long cap;
};

// first index in [0,n) with batch[idx] >= v (batch sorted ascending)
__device__ __forceinline__ int lower_bound_batch(const int64_t* __restrict__ batch, int n, long long v) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (batch[mid] < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

template <bool FILL>
__device__ __forceinline__ void emit(bool hit, int lane, int j, float dx, float dy, float dz, float d2, int& run,
                                     long long base, const EdgeOut& o) {
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (FILL && hit) {
        const long long slot = base + run + __popc(m & ((1u << lane) - 1u));
        if (slot < o.cap) {
            float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
            if (d2 > 0.f) {
                const float d = __fsqrt_rn(d2);
                const float inv = 1.0f / d;
                g = make_float4(dx * inv, dy * inv, dz * inv, d);
            }
            o.col[slot] = j;
            o.geom[slot] = g;
        }
    }
    run += __popc(m);
}

// ------------------------------------------------------------------ brute: sources = the target's own molecule
// one warp walks the sources [m0, m1) of target i in ascending order; -> number of hits
template <bool FILL>
__device__ __forceinline__ int brute_row(const float* __restrict__ pos, int i, int lane, int m0, int m1, const Box3<float>& box,
                                         int periodic, float lo2, float hi2, long long base, const EdgeOut& out) {
    const float xi = pos[3 * (size_t)i], yi = pos[3 * (size_t)i + 1], zi = pos[3 * (size_t)i + 2];
    int run = 0;
    for (int j0 = m0; j0 < m1; j0 += 32) {
        const int j = j0 + lane;
        bool hit = false;
        float dx = 0.f, dy = 0.f, dz = 0.f, d2 = 0.f;
        if (j < m1) {
            if (j == i) {
                hit = true;  // self loop (loop=True in both models)
            } else {
                dx = rn_sub(xi, pos[3 * (size_t)j]);
                dy = rn_sub(yi, pos[3 * (size_t)j + 1]);
                dz = rn_sub(zi, pos[3 * (size_t)j + 2]);
                if (periodic) min_image_tric(dx, dy, dz, box);
                d2 = dist2(dx, dy, dz);
                hit = d2 < hi2 && d2 >= lo2;
            }
        }
        emit<FILL>(hit, lane, j, dx, dy, dz, d2, run, base, out);
    }
    return run;
}

template <bool FILL>
__global__ void __launch_bounds__(CD_WARPS * 32)
csr_brute_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int n, int n_rows, Box3<float> box,
                 int periodic, float lo2, float hi2, EdgeOut out) {
    pdl_sync();
    const int i = blockIdx.x * CD_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= n_rows) return;
    const long long bi = batch[i];
    const int m0 = lower_bound_batch(batch, n, bi), m1 = lower_bound_batch(batch, n, bi + 1);
    const int run = brute_row<FILL>(pos, i, lane, m0, m1, box, periodic, lo2, hi2, FILL ? out.row_ptr[i] : 0, out);
    if (!FILL && lane == 0) out.row_cnt[i] = run;
}

// Small systems (n <= CS_MAX_ATOMS: a peptide in an MD loop): count, scan, fill and the capacity clamp of
// csr_finish_kernel in ONE launch instead of four dependent ones.  One thread-block cluster of 8 CTAs: CTA r owns the
// contiguous rows [r * per, (r + 1) * per), a warp walks a row (brute_row, as above), the CTAs exchange their edge totals
// through distributed shared memory between the counting and the filling pass.  Same walk, same order inside a row,
// same clamping: bitwise the same structure.  (A single CTA was tried first: 35 us at 166 atoms against 22 us for the
// four launches -- 2 x 27 k pair tests are instruction-bound on one SM, gpurun r02v.)
constexpr int CS_MAX_ATOMS = 256, CS_CLUSTER = 8, CS_ROWS = CS_MAX_ATOMS / CS_CLUSTER, CS_THREADS = 512;
__global__ void __cluster_dims__(CS_CLUSTER, 1, 1) __launch_bounds__(CS_THREADS)
csr_brute_small_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int n, Box3<float> box,
                       int periodic, float lo2, float hi2, int* __restrict__ row_ptr, int32_t* __restrict__ col,
                       float4* __restrict__ geom, long long cap, int32_t* __restrict__ num_found) {
    pdl_sync();
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ int s_m0[CS_ROWS], s_m1[CS_ROWS], s_start[CS_ROWS], s_total, s_base;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int rank = (int)cluster.block_rank();
    const int per = (n + CS_CLUSTER - 1) / CS_CLUSTER;           // <= CS_ROWS = 32
    const int r0 = rank * per;
    const EdgeOut out{nullptr, nullptr, col, geom, cap};
    for (int li = wrp; li < per; li += CS_THREADS / 32) {
        const int i = r0 + li;
        int run = 0, m0 = 0, m1 = 0;
        if (i < n) {
            const long long bi = batch[i];
            m0 = lower_bound_batch(batch, n, bi);
            m1 = lower_bound_batch(batch, n, bi + 1);
            run = brute_row<false>(pos, i, lane, m0, m1, box, periodic, lo2, hi2, 0, out);
        }
        if (lane == 0) { s_m0[li] = m0; s_m1[li] = m1; s_start[li] = run; }
    }
    __syncthreads();
    if (wrp == 0) {                                              // exclusive scan of this CTA's <= 32 counts
        const int v = lane < per ? s_start[lane] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (lane < per) s_start[lane] = incl - v;
        if (lane == 31) s_total = incl;
    }
    cluster.sync();
    if (wrp == 0) {                                              // edges of the CTAs before this one, and of all of them
        const int t = lane < CS_CLUSTER ? *cluster.map_shared_rank(&s_total, lane) : 0;
        const int before = warp_sum_i(lane < rank ? t : 0);
        const int all = warp_sum_i(t);
        if (lane == 0) {
            s_base = before;
            if (rank == 0) {
                row_ptr[n] = all > cap ? (int)cap : all;
                *num_found = all;
            }
        }
    }
    __syncthreads();
    const int base = s_base;
    if (threadIdx.x < per && r0 + threadIdx.x < n) {
        const int start = base + s_start[threadIdx.x];
        row_ptr[r0 + threadIdx.x] = start > cap ? (int)cap : start;
    }
    for (int li = wrp; li < per; li += CS_THREADS / 32) {
        const int i = r0 + li;
        if (i < n) brute_row<true>(pos, i, lane, s_m0[li], s_m1[li], box, periodic, lo2, hi2, base + s_start[li], out);
    }
    cluster.sync();                                              // nobody leaves while its shared memory may still be read
}

// ------------------------------------------------------------------ cell list
// cells are filled by atomics (arbitrary order); re-rank every cell by atom index so traversal order is deterministic
__global__ void cell_rank_kernel(const SortedAtom<float>* __restrict__ tmp, const long long* __restrict__ tmp_batch,
                                 const int* __restrict__ cell_start, const int* __restrict__ cell_of,
                                 const int* __restrict__ slot_of, int n, SortedAtom<float>* __restrict__ spos,
                                 long long* __restrict__ sbatch) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = cell_of[i];
    const int cs = cell_start[c], ce = cell_start[c + 1];
    int rank = 0;
    for (int k = cs; k < ce; ++k) rank += tmp[k].idx < i ? 1 : 0;
    const int src = cs + slot_of[i];
    spos[cs + rank] = tmp[src];
    sbatch[cs + rank] = tmp_batch[src];
}

// num_found = edges found; row pointers are clamped to the capacity so that an overflowing list is truncated
// (rows past the capacity become empty) instead of read out of bounds by the model kernels
__global__ void csr_finish_kernel(int* __restrict__ row_ptr, int n, long long cap, int32_t* __restrict__ num_found) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    const int v = row_ptr[i];
    if (i == n) *num_found = v;
    if (v > cap) row_ptr[i] = (int)cap;
}

struct DirectWs {
    int *row_cnt, *scan_tmp;
    // cell only
    int *cell_cnt, *cell_start, *cell_of, *slot_of, *cell_scan_tmp;
    SortedAtom<float>*tmp, *spos;
    long long *tmp_batch, *sbatch;
};
void carve_direct(Arena& a, int strategy, int64_t n, int ncells, DirectWs& w) {
    w.row_cnt = a.get<int>(n + 1);
    w.scan_tmp = a.get<int>(tmdn_scan_tmp_ints(n + 1));
    if (strategy == TMDN_CELL) {
        w.cell_cnt = a.get<int>(ncells + 1);
        w.cell_start = a.get<int>(ncells + 1);
        w.cell_of = a.get<int>(n);
        w.slot_of = a.get<int>(n);
        w.cell_scan_tmp = a.get<int>(tmdn_scan_tmp_ints(ncells + 1));
        w.tmp = a.get<SortedAtom<float>>(n);
        w.spos = a.get<SortedAtom<float>>(n);
        w.tmp_batch = a.get<long long>(n);
        w.sbatch = a.get<long long>(n);
    }
}
}  // namespace

extern "C" size_t tmdn_csr_build_workspace_bytes(int strategy, int64_t n_atoms, const double box[9],
                                                 double cutoff_upper) {
    Arena a(nullptr, 0);
    DirectWs w;
    int ncells = 0;
    if (strategy == TMDN_CELL) {
        CellGrid<float> g;
        if (!box || cutoff_upper <= 0 || !make_grid<float>(box, cutoff_upper, g)) return 0;
        ncells = g.ncells;
    }
    carve_direct(a, strategy, n_atoms, ncells, w);
    return a.off + 256;
}

extern "C" int tmdn_csr_build_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms,
                                  const double box[9], int use_periodic, double cutoff_lower, double cutoff_upper,
                                  int64_t max_edges, int32_t* row_ptr, int32_t* col, float* geom, int32_t* num_found,
                                  void* workspace, size_t workspace_bytes, tmdn_stream_t stream_) {
    return tmdn_csr_build_rows_f32(strategy, pos, batch, n_atoms, n_atoms, box, use_periodic, cutoff_lower, cutoff_upper,
                                   max_edges, row_ptr, col, geom, num_found, workspace, workspace_bytes, stream_);
}

extern "C" int tmdn_csr_build_rows_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms,
                                       int64_t n_rows, const double box[9], int use_periodic, double cutoff_lower,
                                       double cutoff_upper, int64_t max_edges, int32_t* row_ptr, int32_t* col,
                                       float* geom, int32_t* num_found, void* workspace, size_t workspace_bytes,
                                       tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_CHECK_ARG(n_rows > 0 && n_rows <= n_atoms, "csr_build: n_rows must be in [1, n_atoms]");
    TMDN_CHECK_ARG(pos && batch && row_ptr && col && geom && num_found, "csr_build: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_atoms < (1LL << 31) - 256, "csr_build: bad atom count");
    TMDN_CHECK_ARG(max_edges > 0 && max_edges < (1LL << 31), "csr_build: max_edges must be positive and fit int32");
    TMDN_CHECK_ARG(cutoff_upper > 0, "Expected \"cutoff\" to be positive");
    TMDN_CHECK_ARG(strategy == TMDN_BRUTE || strategy == TMDN_SHARED || strategy == TMDN_CELL, "Unknown kernel name");
    TMDN_CHECK_ARG(!(use_periodic || strategy == TMDN_CELL) || box, "csr_build: box required");
    const int n = (int)n_atoms;
    const float lo = (float)cutoff_lower, hi = (float)cutoff_upper;
    const float lo2 = lo * lo, hi2 = hi * hi;
    Arena arena(workspace, workspace_bytes);
    DirectWs w;
    CellGrid<float> g;
    memset(&g, 0, sizeof(g));
    if (strategy == TMDN_CELL) {
        TMDN_CHECK_ARG(box[1] == 0 && box[2] == 0 && box[3] == 0 && box[5] == 0 && box[6] == 0 && box[7] == 0,
                       "Expected \"box_size\" to be diagonal");
        TMDN_CHECK_ARG(box[0] > 0 && box[4] > 0 && box[8] > 0, "csr_build: box lengths must be positive");
        TMDN_CHECK_ARG(make_grid<float>(box, cutoff_upper, g), "Too many cells in one dimension. Maximum is 1024");
    }
    carve_direct(arena, strategy, n_atoms, g.ncells, w);
    if (!arena.ok || !workspace) {
        tmdn_set_error("csr_build: workspace too small (%zu < %zu)", workspace_bytes, arena.off);
        return TMDN_EWORKSPACE;
    }
    EdgeOut cnt{nullptr, w.row_cnt, col, (float4*)geom, (long long)max_edges};
    EdgeOut fill{row_ptr, nullptr, col, (float4*)geom, (long long)max_edges};
    const int nr = (int)n_rows;
    const unsigned rgrid = (unsigned)cdiv(nr, CD_WARPS);
    if (nr < n || strategy == TMDN_CELL) TMDN_CUDA(cudaMemsetAsync(w.row_cnt, 0, sizeof(int) * ((size_t)n + 1), stream));
    if (strategy == TMDN_CELL) {
        TMDN_CUDA(cudaMemsetAsync(w.cell_cnt, 0, sizeof(int) * (g.ncells + 1), stream));
        const int blocks = (int)cdiv(n, 256);
        launch_k(cell_assign_kernel<float>, blocks, 256, 0, stream, pos, n, g, w.cell_cnt, w.cell_of, w.slot_of);
        TMDN_LAUNCH_CHECK();
        TMDN_TRY(tmdn_scan_exclusive_i32(w.cell_cnt, w.cell_start, g.ncells, w.cell_scan_tmp, stream));
        launch_k(cell_scatter_kernel<float>, blocks, 256, 0, stream, pos, batch, n, w.cell_start, w.cell_of, w.slot_of, w.tmp,
                                                               w.tmp_batch);
        TMDN_LAUNCH_CHECK();
        launch_k(cell_rank_kernel, blocks, 256, 0, stream, w.tmp, w.tmp_batch, w.cell_start, w.cell_of, w.slot_of, n, w.spos,
                                                     w.sbatch);
        TMDN_LAUNCH_CHECK();
        CellTileOut<float> to{};
        to.row_cnt = w.row_cnt; to.row_ptr = row_ptr; to.col = col; to.geom = (float4*)geom;
        to.cap = (long long)max_edges; to.n_rows = nr;
        launch_k(cell_tile_kernel<float, CT_CSR_COUNT>, g.ncells, CT_THREADS, 0, stream, w.spos, w.sbatch, w.cell_start, g,
                                                                                  use_periodic, lo2, hi2, to);
        TMDN_LAUNCH_CHECK();
        TMDN_TRY(tmdn_scan_exclusive_i32(w.row_cnt, row_ptr, n_atoms, w.scan_tmp, stream));
        TMDN_CUDA(cudaMemsetAsync(w.row_cnt, 0, sizeof(int) * ((size_t)n + 1), stream));   // now the fill cursor
        launch_k(cell_tile_kernel<float, CT_CSR_FILL>, g.ncells, CT_THREADS, 0, stream, w.spos, w.sbatch, w.cell_start, g,
                                                                                 use_periodic, lo2, hi2, to);
        TMDN_LAUNCH_CHECK();
    } else {
        Box3<float> b;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) b.m[r][c] = (use_periodic && box) ? (float)box[3 * r + c] : 0.f;
        if (nr == n && n <= CS_MAX_ATOMS && g_tmdn_tuning[TMDN_TUNE_SMALL_CSR]) {
            launch_k(csr_brute_small_kernel, CS_CLUSTER, CS_THREADS, 0, stream, pos, batch, n, b, use_periodic, lo2, hi2, row_ptr, col,
                     (float4*)geom, (long long)max_edges, num_found);
            TMDN_LAUNCH_CHECK();
            return TMDN_OK;
        }
        launch_k(csr_brute_kernel<false>, rgrid, CD_WARPS * 32, 0, stream, pos, batch, n, nr, b, use_periodic, lo2, hi2, cnt);
        TMDN_LAUNCH_CHECK();
        TMDN_TRY(tmdn_scan_exclusive_i32(w.row_cnt, row_ptr, n_atoms, w.scan_tmp, stream));
        launch_k(csr_brute_kernel<true>, rgrid, CD_WARPS * 32, 0, stream, pos, batch, n, nr, b, use_periodic, lo2, hi2, fill);
        TMDN_LAUNCH_CHECK();
    }
    launch_k(csr_finish_kernel, (unsigned)cdiv(n + 1, 256), 256, 0, stream, row_ptr, n, (long long)max_edges, num_found);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

// ------------------------------------------------------------------ allocation-free variant (static device scratch)
// tmdn_csr_build_static_f32: tmdn_csr_build_f32 with its scratch (row counters, scan partials, the cell-sorted copies of
// the atoms) in a file-scope device array instead of a caller-provided workspace, for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds
// the array until exit.  g_csr_scratch is a host-side stand-in in host code, not a device address: its address is
// taken once per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_csr_scratch[TMDN_CSR_STATIC_SCRATCH_BYTES];

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mu;
void* g_static_ptr[STATIC_MAX_DEVICES] = {nullptr};
}  // namespace

extern "C" int tmdn_csr_build_static_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms,
                                         const double box[9], int use_periodic, double cutoff_lower, double cutoff_upper,
                                         int64_t max_edges, int32_t* row_ptr, int32_t* col, float* geom,
                                         int32_t* num_found, tmdn_stream_t stream_) {
    TMDN_CHECK_ARG(n_atoms > 0, "csr_build_static: bad atom count");
    TMDN_CHECK_ARG(strategy == TMDN_BRUTE || strategy == TMDN_SHARED || strategy == TMDN_CELL, "Unknown kernel name");
    const size_t need = tmdn_csr_build_workspace_bytes(strategy, n_atoms, box, cutoff_upper);
    TMDN_CHECK_ARG(need > 0, "csr_build_static: bad box for the cell strategy");
    if (need > (size_t)TMDN_CSR_STATIC_SCRATCH_BYTES) {
        tmdn_set_error("csr_build_static: %lld atoms need %zu bytes of scratch, the static array holds %d "
                       "(TMDN_CSR_STATIC_SCRATCH_BYTES); use tmdn_csr_build_f32 with a caller-provided workspace",
                       (long long)n_atoms, need, TMDN_CSR_STATIC_SCRATCH_BYTES);
        return TMDN_EWORKSPACE;
    }
    int dev = -1;
    TMDN_CUDA(cudaGetDevice(&dev));
    TMDN_CHECK_ARG(dev >= 0 && dev < STATIC_MAX_DEVICES, "csr_build_static: device index %d out of range", dev);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (g_static_ptr[dev] == nullptr) {   // once per device, with that device current
            void* p = nullptr;
            TMDN_CUDA(cudaGetSymbolAddress(&p, g_csr_scratch));
            g_static_ptr[dev] = p;
        }
        scratch = g_static_ptr[dev];
    }
    // the builder clears the counters it accumulates into and overwrites everything else it reads, so the contents
    // left by the previous call do not matter (the caller-provided workspaces are uninitialised memory as well)
    return tmdn_csr_build_rows_f32(strategy, pos, batch, n_atoms, n_atoms, box, use_periodic, cutoff_lower, cutoff_upper,
                                   max_edges, row_ptr, col, geom, num_found, scratch, (size_t)TMDN_CSR_STATIC_SCRATCH_BYTES,
                                   stream_);
}
