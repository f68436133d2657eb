// neighbors.cu -- neighbor-pair search for sm_100a: tiled brute force (batch-aware) and cell list.
//
// Replaces torchmdnet/neighbors/neighbors_cuda_{brute,shared,cell}.cuh of the reference (same pair set,
// same padded outputs, same found-count semantics; see include/tmdn.h).  Design differences:
//   * brute/shared: one CTA per 128 target atoms streams 128-atom source tiles through shared memory and
//     skips every tile whose batch-id range cannot overlap its own, so a batch of small molecules costs
//     O(N * molecule size) instead of N^2/2 tests; no N < 32768 limit.
//   * cell: counting sort into cells (integer atomics + scan, no radix sort), then ONE CTA per cell stages the
//     27-cell neighbourhood in shared memory and emits the cell's pairs.
//   * in both, pairs are appended with one global atomicAdd per (CTA, tile): hits are counted, block-scanned,
//     a contiguous range is reserved, and the range is filled -- instead of one contended atomic per pair.
//   * arithmetic is written with explicit round-to-nearest mul/add/sub (no FMA contraction) so deltas and
//     distances are bit-identical to the CPU restatement in oracle/neighbors.py.
#include "common.cuh"
#include "nbr_cell_tile.cuh"
#include "scan.cuh"

namespace {
using namespace nbr;

template <typename T> struct PairOut {
    int32_t* nbr;   // [2, max_pairs]
    T* deltas;      // [max_pairs, 3]
    T* dist;        // [max_pairs]
    int32_t* counter;
    long long max_pairs;
};

template <typename T>
__device__ __forceinline__ void write_pair(const PairOut<T>& o, long long slot, int a, int b, T dx, T dy, T dz, T d) {
    if (slot < o.max_pairs) {
        o.nbr[slot] = a;
        o.nbr[o.max_pairs + slot] = b;
        o.deltas[3 * slot + 0] = dx;
        o.deltas[3 * slot + 1] = dy;
        o.deltas[3 * slot + 2] = dz;
        o.dist[slot] = d;
    }
}

// Block-wide exclusive scan of one int per thread (NB_THREADS threads); returns the thread's offset and the total.
__device__ __forceinline__ int block_excl_scan(int v, int& total, int* s_warp /* [NB_THREADS/32 + 1] */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int incl = warp_incl_scan_i(v, lane);
    __syncthreads();  // protect s_warp reuse
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < NB_THREADS / 32; ++k) {
        int x = s_warp[k];
        if (k < w) woff += x;
        tot += x;
    }
    total = tot;
    return woff + incl - v;
}

// ------------------------------------------------------------------ batch range of every 128-atom tile
__global__ void __launch_bounds__(NB_THREADS) tile_batch_range_kernel(const int64_t* __restrict__ batch, int n,
                                                                      long long* __restrict__ tile_mm) {
    pdl_sync();
    __shared__ long long s_min[NB_THREADS / 32], s_max[NB_THREADS / 32];
    const int i = blockIdx.x * NB_THREADS + threadIdx.x;
    long long lo = LLONG_MAX, hi = LLONG_MIN;
    if (i < n) lo = hi = batch[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        long long a = __shfl_xor_sync(0xffffffffu, lo, o), b = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = a < lo ? a : lo;
        hi = b > hi ? b : hi;
    }
    if ((threadIdx.x & 31) == 0) { s_min[threadIdx.x >> 5] = lo; s_max[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < NB_THREADS / 32; ++k) {
            lo = s_min[k] < lo ? s_min[k] : lo;
            hi = s_max[k] > hi ? s_max[k] : hi;
        }
        tile_mm[2 * blockIdx.x] = lo;
        tile_mm[2 * blockIdx.x + 1] = hi;
    }
}

// ------------------------------------------------------------------ tiled brute force
template <typename T>
__global__ void __launch_bounds__(NB_THREADS)
nbr_brute_kernel(const T* __restrict__ pos, const int64_t* __restrict__ batch, int n, Box3<T> box, int periodic, T lo2,
                 T hi2, int loop, int transpose, PairOut<T> out, const long long* __restrict__ tile_mm) {
    pdl_sync();
    __shared__ T sx[NB_THREADS], sy[NB_THREADS], sz[NB_THREADS];
    __shared__ long long sb[NB_THREADS];
    __shared__ int s_warp[NB_THREADS / 32 + 1];
    __shared__ int s_base;
    const int t = threadIdx.x;
    const int i = blockIdx.x * NB_THREADS + t;
    const bool vi = i < n;
    T xi = 0, yi = 0, zi = 0;
    long long bi = 0;
    if (vi) {
        xi = pos[3 * (size_t)i];
        yi = pos[3 * (size_t)i + 1];
        zi = pos[3 * (size_t)i + 2];
        bi = batch[i];
    }
    // tile_mm == null (few tiles): no batch-range skipping, one launch less
    const long long bmin = tile_mm ? tile_mm[2 * blockIdx.x] : LLONG_MIN, bmax = tile_mm ? tile_mm[2 * blockIdx.x + 1] : LLONG_MAX;
    const int per_hit = transpose ? 2 : 1;
    // 2-D grid: the source tiles jt <= blockIdx.x of this target tile are dealt round-robin to the gridDim.y CTAs of the
    // row, so a single system of 1k atoms (8 target tiles) still fills the chip and the triangular imbalance between
    // the first and the last target tile is spread over all CTAs (the reference uses one THREAD per pair:
    // neighbors_cuda_brute.cuh:81-92).  Pair order in the output depends on CTA scheduling, as in the reference.
    for (int jt = blockIdx.y; jt <= (int)blockIdx.x; jt += gridDim.y) {
        if (tile_mm && (tile_mm[2 * jt + 1] < bmin || tile_mm[2 * jt] > bmax)) continue;  // block-uniform
        __syncthreads();
        const int j0 = jt * NB_THREADS;
        const int nj = min(NB_THREADS, n - j0);
        if (t < nj) {
            sx[t] = pos[3 * (size_t)(j0 + t)];
            sy[t] = pos[3 * (size_t)(j0 + t) + 1];
            sz[t] = pos[3 * (size_t)(j0 + t) + 2];
            sb[t] = batch[j0 + t];
        }
        __syncthreads();
        const int jend = vi ? min(nj, i - j0) : 0;  // only j < i
        // pass 1: test every candidate once; the hits of this thread (<= 128 per tile) are kept as a bit mask in four
        // registers, so that the write pass below only revisits hits (0.3 % of the tests at liquid density)
        unsigned hit[NB_THREADS / 32];
        int cnt = 0;
#pragma unroll
        for (int w = 0; w < NB_THREADS / 32; ++w) {
            unsigned m = 0;
            const int je = min(32, jend - 32 * w);
            for (int bq = 0; bq < je; ++bq) {
                const int jj = 32 * w + bq;
                if (sb[jj] != bi) continue;
                T dx = rn_sub(xi, sx[jj]), dy = rn_sub(yi, sy[jj]), dz = rn_sub(zi, sz[jj]);
                if (periodic) min_image_tric(dx, dy, dz, box);
                const T d2 = dist2(dx, dy, dz);
                if (d2 < hi2 && d2 >= lo2) m |= 1u << bq;
            }
            hit[w] = m;
            cnt += __popc(m);
        }
        const int self = (loop && vi && jt == (int)blockIdx.x) ? 1 : 0;
        const int items = cnt * per_hit + self;
        int total;
        const int off = block_excl_scan(items, total, s_warp);
        if (total == 0) continue;  // block-uniform
        if (t == 0) s_base = atomicAdd(out.counter, total);
        __syncthreads();
        long long slot = (long long)s_base + off;
#pragma unroll
        for (int w = 0; w < NB_THREADS / 32; ++w) {
            unsigned m = hit[w];
            while (m) {
                const int jj = 32 * w + __ffs(m) - 1;
                m &= m - 1;
                T dx = rn_sub(xi, sx[jj]), dy = rn_sub(yi, sy[jj]), dz = rn_sub(zi, sz[jj]);
                if (periodic) min_image_tric(dx, dy, dz, box);
                const T d = sqrt_rn(dist2(dx, dy, dz));
                write_pair(out, slot++, i, j0 + jj, dx, dy, dz, d);
                if (transpose) write_pair(out, slot++, j0 + jj, i, -dx, -dy, -dz, d);
            }
        }
        if (self) write_pair(out, slot, i, i, T(0), T(0), T(0), T(0));
    }
}

// slots in [min(found, max_pairs), max_pairs) get the reference's padding values (common.cuh:70-76).  A CTA covers
// PAD_SPAN * 256 consecutive slots in PAD_SPAN coalesced passes, so the (usually much larger) filled part of the list
// costs one early exit per CTA of 2048 slots instead of one thread per slot.
constexpr int PAD_SPAN = 8;
template <typename T> __global__ void pad_fill_kernel(PairOut<T> out) {
    pdl_sync();
    const long long base = (long long)blockIdx.x * blockDim.x * PAD_SPAN;
    const long long found = *out.counter;
    if (base + (long long)blockDim.x * PAD_SPAN <= found) return;
#pragma unroll
    for (int k = 0; k < PAD_SPAN; ++k) {
        const long long s = base + (long long)k * blockDim.x + threadIdx.x;
        if (s < found || s >= out.max_pairs) continue;
        out.nbr[s] = -1;
        out.nbr[out.max_pairs + s] = -1;
        out.deltas[3 * s] = T(0);
        out.deltas[3 * s + 1] = T(0);
        out.deltas[3 * s + 2] = T(0);
        out.dist[s] = T(0);
    }
}

// ------------------------------------------------------------------ cell list: binning in nbr_common.cuh, traversal in
// nbr_cell_tile.cuh (one CTA per cell, shared-memory candidate tile, hit lists, one atomic per centre group)

// ------------------------------------------------------------------ backward (neighbors_cuda.cu:43-71)
template <typename T>
__global__ void nbr_backward_kernel(const int32_t* __restrict__ nbr, const T* __restrict__ deltas,
                                    const T* __restrict__ dist, const T* __restrict__ gdel, const T* __restrict__ gdist,
                                    long long max_pairs, T* __restrict__ gpos) {
    pdl_sync();
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= max_pairs) return;
    const T d = dist[e];
    const int i = nbr[e], j = nbr[max_pairs + e];
    if (d == T(0) || i < 0) return;
    T g[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        T v = gdel ? gdel[3 * e + k] : T(0);
        if (gdist) v += deltas[3 * e + k] / d * gdist[e];
        g[k] = v;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        atomicAdd(&gpos[3 * (size_t)i + k], g[k]);
        atomicAdd(&gpos[3 * (size_t)j + k], -g[k]);
    }
}

struct CellWs {
    int *cell_cnt, *cell_start, *cell_of, *slot_of, *scan_tmp;
    void* spos;
    long long* sbatch;
};
template <typename T> void carve_cell(Arena& a, int64_t n, int ncells, CellWs& w) {
    w.cell_cnt = a.get<int>(ncells + 1);
    w.cell_start = a.get<int>(ncells + 1);
    w.cell_of = a.get<int>(n);
    w.slot_of = a.get<int>(n);
    w.scan_tmp = a.get<int>(tmdn_scan_tmp_ints(ncells + 1));
    w.spos = a.get<SortedAtom<T>>(n);
    w.sbatch = a.get<long long>(n);
}

template <typename T>
int neighbors_impl(int strategy, const T* pos, const int64_t* batch, int64_t n_atoms, const double box[9],
                   int use_periodic, double cutoff_lower, double cutoff_upper, int loop, int include_transpose,
                   int64_t max_pairs, int32_t* neighbors, T* deltas, T* distances, int32_t* num_found, void* workspace,
                   size_t workspace_bytes, cudaStream_t stream) {
    TMDN_CHECK_ARG(pos && batch && neighbors && deltas && distances && num_found, "neighbors: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0, "Expected the 1nd dimension size of \"positions\" to be more than 0");
    TMDN_CHECK_ARG(n_atoms < (1LL << 31) - 256, "neighbors: too many atoms");
    TMDN_CHECK_ARG(max_pairs > 0, "Expected \"max_num_neighbors\" to be positive");
    TMDN_CHECK_ARG(max_pairs < (1LL << 31), "neighbors: max_num_pairs must fit int32");
    TMDN_CHECK_ARG(cutoff_upper > 0, "Expected \"cutoff\" to be positive");
    TMDN_CHECK_ARG(strategy == TMDN_BRUTE || strategy == TMDN_SHARED || strategy == TMDN_CELL, "Unknown kernel name");
    TMDN_CHECK_ARG(!(use_periodic || strategy == TMDN_CELL) || box, "neighbors: box required");
    const int n = (int)n_atoms;
    PairOut<T> out{neighbors, deltas, distances, num_found, (long long)max_pairs};
    const T lo = (T)cutoff_lower, hi = (T)cutoff_upper;
    const T lo2 = lo * lo, hi2 = hi * hi;
    TMDN_CUDA(cudaMemsetAsync(num_found, 0, sizeof(int32_t), stream));
    Arena arena(workspace, workspace_bytes);
    if (strategy == TMDN_CELL) {
        TMDN_CHECK_ARG(box[1] == 0 && box[2] == 0 && box[3] == 0 && box[5] == 0 && box[6] == 0 && box[7] == 0,
                       "Expected \"box_size\" to be diagonal");
        TMDN_CHECK_ARG(box[0] > 0 && box[4] > 0 && box[8] > 0, "neighbors: box lengths must be positive");
        CellGrid<T> g;
        TMDN_CHECK_ARG(make_grid<T>(box, cutoff_upper, g), "Too many cells in one dimension. Maximum is 1024");
        CellWs w;
        carve_cell<T>(arena, n_atoms, g.ncells, w);
        if (!arena.ok || !workspace) {
            tmdn_set_error("neighbors(cell): workspace too small (%zu < %zu)", workspace_bytes, arena.off);
            return TMDN_EWORKSPACE;
        }
        TMDN_CUDA(cudaMemsetAsync(w.cell_cnt, 0, sizeof(int) * (g.ncells + 1), stream));
        const int blocks = (int)cdiv(n, 256);
        launch_k(cell_assign_kernel<T>, blocks, 256, 0, stream, pos, n, g, w.cell_cnt, w.cell_of, w.slot_of);
        TMDN_LAUNCH_CHECK();
        TMDN_TRY(tmdn_scan_exclusive_i32(w.cell_cnt, w.cell_start, g.ncells, w.scan_tmp, stream));
        launch_k(cell_scatter_kernel<T>, blocks, 256, 0, stream, pos, batch, n, w.cell_start, w.cell_of, w.slot_of,
                                                           (SortedAtom<T>*)w.spos, w.sbatch);
        TMDN_LAUNCH_CHECK();
        CellTileOut<T> to{};
        to.nbr = neighbors; to.deltas = deltas; to.dist = distances; to.counter = num_found;
        to.max_pairs = (long long)max_pairs; to.loop = loop; to.transpose = include_transpose;
        launch_k(cell_tile_kernel<T, CT_PAIRS>, g.ncells, CT_THREADS, 0, stream, (const SortedAtom<T>*)w.spos, w.sbatch,
                                                                          w.cell_start, g, use_periodic, lo2, hi2, to);
        TMDN_LAUNCH_CHECK();
    } else {
        Box3<T> b;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) b.m[r][c] = (use_periodic && box) ? (T)box[3 * r + c] : T(0);
        const int ntiles = (int)cdiv(n, NB_THREADS);
        long long* tile_mm = arena.get<long long>(2 * (size_t)ntiles);
        if (!arena.ok || !workspace) {
            tmdn_set_error("neighbors(brute): workspace too small (%zu < %zu)", workspace_bytes, arena.off);
            return TMDN_EWORKSPACE;
        }
        if (ntiles > 32) {      // skipping tiles of other molecules pays from a few thousand atoms on
            launch_k(tile_batch_range_kernel, ntiles, NB_THREADS, 0, stream, batch, n, tile_mm);
            TMDN_LAUNCH_CHECK();
        } else {
            tile_mm = nullptr;
        }
        // source-tile chunks per target tile: enough CTAs for ~one full wave of 16 CTAs per SM
        long long ny = cdiv((long long)tmdn_sm_count() * 16, ntiles);
        if (ny > ntiles) ny = ntiles;
        if (ny < 1) ny = 1;
        launch_k(nbr_brute_kernel<T>, dim3((unsigned)ntiles, (unsigned)ny), NB_THREADS, 0, stream, pos, batch, n, b,
                 use_periodic, lo2, hi2, loop, include_transpose, out, tile_mm);
        TMDN_LAUNCH_CHECK();
    }
    launch_k(pad_fill_kernel<T>, (unsigned)cdiv(cdiv(max_pairs, PAD_SPAN), 256), 256, 0, stream, out);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

template <typename T>
int backward_impl(const int32_t* neighbors, const T* deltas, const T* distances, const T* grad_deltas,
                  const T* grad_distances, int64_t max_pairs, int64_t n_atoms, T* grad_pos, cudaStream_t stream) {
    TMDN_CHECK_ARG(neighbors && deltas && distances && grad_pos, "neighbors_backward: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && max_pairs >= 0, "neighbors_backward: bad sizes");
    TMDN_CUDA(cudaMemsetAsync(grad_pos, 0, sizeof(T) * 3 * (size_t)n_atoms, stream));
    if (max_pairs == 0 || (!grad_deltas && !grad_distances)) return TMDN_OK;
    launch_k(nbr_backward_kernel<T>, (unsigned)cdiv(max_pairs, 256), 256, 0, stream, 
        neighbors, deltas, distances, grad_deltas, grad_distances, (long long)max_pairs, grad_pos);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

}  // namespace

extern "C" size_t tmdn_neighbors_workspace_bytes(int strategy, int64_t n_atoms, const double box[9],
                                                 double cutoff_upper, int elem_bytes) {
    Arena a(nullptr, 0);
    if (strategy == TMDN_CELL) {
        if (!box || cutoff_upper <= 0) return 0;
        CellGrid<double> g;
        if (!make_grid<double>(box, cutoff_upper, g)) return 0;
        CellWs w;
        if (elem_bytes == 8) carve_cell<double>(a, n_atoms, g.ncells, w);
        else carve_cell<float>(a, n_atoms, g.ncells, w);
    } else {
        a.get<long long>(2 * (size_t)cdiv(n_atoms, NB_THREADS));
    }
    return a.off + 256;
}

extern "C" int tmdn_neighbors_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms,
                                  const double box[9], int use_periodic, double cutoff_lower, double cutoff_upper,
                                  int loop, int include_transpose, int64_t max_pairs, int32_t* neighbors, float* deltas,
                                  float* distances, int32_t* num_found, void* workspace, size_t workspace_bytes,
                                  tmdn_stream_t stream) {
    return neighbors_impl<float>(strategy, pos, batch, n_atoms, box, use_periodic, cutoff_lower, cutoff_upper, loop,
                                 include_transpose, max_pairs, neighbors, deltas, distances, num_found, workspace,
                                 workspace_bytes, (cudaStream_t)stream);
}
extern "C" int tmdn_neighbors_f64(int strategy, const double* pos, const int64_t* batch, int64_t n_atoms,
                                  const double box[9], int use_periodic, double cutoff_lower, double cutoff_upper,
                                  int loop, int include_transpose, int64_t max_pairs, int32_t* neighbors,
                                  double* deltas, double* distances, int32_t* num_found, void* workspace,
                                  size_t workspace_bytes, tmdn_stream_t stream) {
    return neighbors_impl<double>(strategy, pos, batch, n_atoms, box, use_periodic, cutoff_lower, cutoff_upper, loop,
                                  include_transpose, max_pairs, neighbors, deltas, distances, num_found, workspace,
                                  workspace_bytes, (cudaStream_t)stream);
}
extern "C" int tmdn_neighbors_backward_f32(const int32_t* neighbors, const float* deltas, const float* distances,
                                           const float* grad_deltas, const float* grad_distances, int64_t max_pairs,
                                           int64_t n_atoms, float* grad_pos, tmdn_stream_t stream) {
    return backward_impl<float>(neighbors, deltas, distances, grad_deltas, grad_distances, max_pairs, n_atoms, grad_pos,
                                (cudaStream_t)stream);
}
extern "C" int tmdn_neighbors_backward_f64(const int32_t* neighbors, const double* deltas, const double* distances,
                                           const double* grad_deltas, const double* grad_distances, int64_t max_pairs,
                                           int64_t n_atoms, double* grad_pos, tmdn_stream_t stream) {
    return backward_impl<double>(neighbors, deltas, distances, grad_deltas, grad_distances, max_pairs, n_atoms,
                                 grad_pos, (cudaStream_t)stream);
}
