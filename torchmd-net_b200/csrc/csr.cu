// csr.cu -- CSR-by-target view of a (possibly padded, unordered) neighbor pair list.
//
// The model kernels reduce over the incoming edges of each atom with plain in-order loops (no float atomics),
// so rows must be contiguous; sorting each row by source index additionally makes every sum bit-reproducible
// from run to run and between a sharded and an unsharded evaluation.
#include "common.cuh"
#include "scan.cuh"

namespace {

__global__ void csr_count_kernel(const int32_t* __restrict__ nbr, const int32_t* __restrict__ num_found,
                                 long long max_pairs, int target_row, int* __restrict__ row_cnt) {
    pdl_sync();
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long ne = *num_found;
    ne = ne < max_pairs ? ne : max_pairs;
    if (e >= ne) return;
    const int t = nbr[(long long)target_row * max_pairs + e];
    if (t >= 0) atomicAdd(&row_cnt[t], 1);
}

__global__ void csr_fill_kernel(const int32_t* __restrict__ nbr, const int32_t* __restrict__ num_found,
                                long long max_pairs, int target_row, const int* __restrict__ row_ptr,
                                int* __restrict__ cursor, int* __restrict__ col_tmp, int* __restrict__ eid_tmp) {
    pdl_sync();
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long ne = *num_found;
    ne = ne < max_pairs ? ne : max_pairs;
    if (e >= ne) return;
    const int t = nbr[(long long)target_row * max_pairs + e];
    if (t < 0) return;
    const int s = nbr[(long long)(1 - target_row) * max_pairs + e];
    const int p = row_ptr[t] + atomicAdd(&cursor[t], 1);
    col_tmp[p] = s;
    eid_tmp[p] = (int)e;
}

// one warp per row: rank-sort the row by source index and emit (col, geom)
__global__ void __launch_bounds__(128)
csr_sort_rows_kernel(const int* __restrict__ row_ptr, int n_atoms, const int* __restrict__ col_tmp,
                     const int* __restrict__ eid_tmp, const float* __restrict__ deltas,
                     const float* __restrict__ dist, int target_row, int* __restrict__ col,
                     float4* __restrict__ geom) {
    pdl_sync();
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n_atoms) return;
    const int r0 = row_ptr[row], r1 = row_ptr[row + 1];
    for (int k = r0 + lane; k < r1; k += 32) {
        const int c = col_tmp[k], e = eid_tmp[k];
        int rank = 0;
        for (int q = r0; q < r1; ++q) {
            const int cq = col_tmp[q];
            rank += (cq < c || (cq == c && eid_tmp[q] < e)) ? 1 : 0;
        }
        const float d = dist[e];
        float4 g = make_float4(0.f, 0.f, 0.f, d);
        if (d > 0.f) {
            const float sgn = target_row == 0 ? 1.0f : -1.0f;  // deltas = pos[row0] - pos[row1]
            const float inv = sgn / d;
            g.x = deltas[3 * (size_t)e] * inv;
            g.y = deltas[3 * (size_t)e + 1] * inv;
            g.z = deltas[3 * (size_t)e + 2] * inv;
        }
        col[r0 + rank] = c;
        geom[r0 + rank] = g;
    }
}

struct CsrWs { int *row_cnt, *cursor, *col_tmp, *eid_tmp, *scan_tmp; };
void carve_csr(Arena& a, int64_t n_atoms, int64_t max_pairs, CsrWs& w) {
    w.row_cnt = a.get<int>(n_atoms + 1);
    w.cursor = a.get<int>(n_atoms + 1);
    w.col_tmp = a.get<int>(max_pairs);
    w.eid_tmp = a.get<int>(max_pairs);
    w.scan_tmp = a.get<int>(tmdn_scan_tmp_ints(n_atoms + 1));
}
}  // namespace

extern "C" size_t tmdn_csr_workspace_bytes(int64_t n_atoms, int64_t max_pairs) {
    Arena a(nullptr, 0);
    CsrWs w;
    carve_csr(a, n_atoms, max_pairs, w);
    return a.off + 256;
}

extern "C" int tmdn_csr_from_pairs(const int32_t* neighbors, const float* deltas, const float* distances,
                                   const int32_t* num_found, int64_t max_pairs, int64_t n_atoms, int target_row,
                                   int32_t* row_ptr, int32_t* col, float* geom, void* workspace,
                                   size_t workspace_bytes, tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_CHECK_ARG(neighbors && deltas && distances && num_found && row_ptr && col && geom, "csr: null pointer");
    TMDN_CHECK_ARG(n_atoms > 0 && max_pairs > 0, "csr: bad sizes");
    TMDN_CHECK_ARG(target_row == 0 || target_row == 1, "csr: target_row must be 0 or 1");
    Arena a(workspace, workspace_bytes);
    CsrWs w;
    carve_csr(a, n_atoms, max_pairs, w);
    if (!a.ok || !workspace) {
        tmdn_set_error("csr: workspace too small (%zu < %zu)", workspace_bytes, a.off);
        return TMDN_EWORKSPACE;
    }
    // row_cnt and cursor are adjacent in the arena: one memset clears both
    TMDN_CUDA(cudaMemsetAsync(w.row_cnt, 0, (size_t)((char*)w.col_tmp - (char*)w.row_cnt), stream));
    const unsigned eb = (unsigned)cdiv(max_pairs, 256);
    launch_k(csr_count_kernel, eb, 256, 0, stream, neighbors, num_found, (long long)max_pairs, target_row, w.row_cnt);
    TMDN_LAUNCH_CHECK();
    TMDN_TRY(tmdn_scan_exclusive_i32(w.row_cnt, row_ptr, n_atoms, w.scan_tmp, stream));
    launch_k(csr_fill_kernel, eb, 256, 0, stream, neighbors, num_found, (long long)max_pairs, target_row, row_ptr, w.cursor,
                                            w.col_tmp, w.eid_tmp);
    TMDN_LAUNCH_CHECK();
    launch_k(csr_sort_rows_kernel, (unsigned)cdiv(n_atoms, 4), 128, 0, stream, row_ptr, (int)n_atoms, w.col_tmp, w.eid_tmp,
                                                                         deltas, distances, target_row, col,
                                                                         (float4*)geom);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}
