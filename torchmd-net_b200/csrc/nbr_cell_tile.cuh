// nbr_cell_tile.cuh -- the cell-list traversal shared by the pair-list op (neighbors.cu) and the direct CSR builder
// (csr_direct.cu).  Replaces traverseCellList / addNeighborsForCell of the reference
// (torchmdnet/neighbors/neighbors_cuda_cell.cuh:264-336: one thread per atom walking 27 cells with scalar global
// loads and one global atomic per pair).
//
// One CTA per cell.  The 27-cell neighbourhood is <= 18 contiguous ranges of the cell-sorted atom array (the three
// x-neighbours of a cell row are adjacent); their atoms are staged once in shared memory.  A warp owns a centre atom:
//   pass 1  lanes test 32 candidates at a time from shared memory, `ballot` compacts the hits into a per-centre
//           list of candidate slots (uint16) -- no output yet, ~20 instructions per test;
//   reserve (pair list only) warp scan of the per-centre counts, ONE global atomicAdd per (chunk, centre group);
//   pass 2  dense loop over the hit list: recompute the delta with the same rounding, sqrt, coalesced stores.
// The pair list tests each unordered pair once and writes both directions: with >= 3 cells per dimension it walks a
// HALF stencil -- the 13 neighbour cells that precede the centre cell in (z, y, x) order plus the centre cell itself
// (there: original index i > j) -- so a centre atom meets ~175 candidates instead of the 337 of the full 27 cells
// (rho = 0.1 A^-3, r_c = 5 A); the pair is emitted as (larger index, smaller index) with delta = pos[larger] -
// pos[smaller] evaluated in that orientation, exactly what the reference emits.  Smaller grids walk all cells with
// the i > j filter.  The CSR builder tests all candidates of a row (rows must be complete).  The rectangular minimum image of the reference,
// d - floor(d/L + 0.5) L (common.cuh:142-148), is applied only to components with |d| >= 0.499 L: below that the
// floor is 0 and the result is d itself, bit for bit, so the IEEE division is skipped for almost every test.
#pragma once
#include "nbr_common.cuh"

namespace nbr {

constexpr int CT_THREADS = 128;
constexpr int CT_WARPS = CT_THREADS / 32;
constexpr int CT_CAND = 512;    // candidates staged per chunk
constexpr int CT_GROUP = 8;     // centre atoms per group (2 per warp)
enum { CT_PAIRS = 0, CT_CSR_COUNT = 1, CT_CSR_FILL = 2 };

template <typename T> struct CellTileOut {
    // pair list
    int32_t* nbr; T* deltas; T* dist; int32_t* counter; long long max_pairs;
    int loop, transpose;
    // csr
    int* row_cnt; const int* row_ptr; int32_t* col; float4* geom; long long cap; int n_rows;
};

template <typename T> __device__ __forceinline__ T min_image_fast(T d, T L) {
    return (fabs(d) < T(0.499) * L) ? d : min_image_rect1(d, L);
}

template <typename T, int MODE>
__global__ void __launch_bounds__(CT_THREADS)
cell_tile_kernel(const SortedAtom<T>* __restrict__ spos, const long long* __restrict__ sbatch,
                 const int* __restrict__ cell_start, CellGrid<T> g, int periodic, T lo2, T hi2, CellTileOut<T> out) {
    pdl_sync();
    __shared__ T cx[CT_CAND], cy[CT_CAND], cz[CT_CAND];
    __shared__ long long cb[CT_CAND];
    __shared__ int cidx[CT_CAND];
    __shared__ unsigned short hits[CT_GROUP][CT_CAND];
    __shared__ int r_start[18], r_pref[19];
    __shared__ int s_cnt[CT_GROUP], s_off[CT_GROUP];
    __shared__ int s_base;
    __shared__ int s_own[2];                                // candidate range of the centre cell (half stencil)
    const bool half = MODE == CT_PAIRS && g.n[0] >= 3 && g.n[1] >= 3 && g.n[2] >= 3;
    const int cell = blockIdx.x;
    const int cs = cell_start[cell], ce = cell_start[cell + 1];
    if (cs == ce) return;                                   // empty cell (block-uniform)
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int ccx = cell % g.n[0], ccy = (cell / g.n[0]) % g.n[1], ccz = cell / (g.n[0] * g.n[1]);
    // ---- neighbourhood as contiguous ranges: lanes 0..17 of warp 0, two candidate ranges per (y,z) row
    if (w == 0) {
        const int ny = g.hi[1] - g.lo[1] + 1, nz = g.hi[2] - g.lo[2] + 1;
        int len = 0, start = 0;
        const int row = lane >> 1, part = lane & 1;
        if (row < ny * nz) {
            const int oy = g.lo[1] + row % ny, oz = g.lo[2] + row / ny;
            const int y = (ccy + oy + g.n[1]) % g.n[1], z = (ccz + oz + g.n[2]) % g.n[2];
            const int row0 = g.n[0] * (y + g.n[1] * z);
            int x0, x1;                                      // inclusive cell range of this part; x0 > x1 = empty
            const bool before = oz < 0 || (oz == 0 && oy < 0);
            if (half && !before && !(oy == 0 && oz == 0)) { x0 = 1; x1 = 0; }          // row after the centre: skipped
            else if (half && !before) {                      // the centre's own row: the cell before it and itself
                x0 = x1 = part == 0 ? ccx : (ccx + g.n[0] - 1) % g.n[0];
            }
            else if (g.n[0] <= 3) { x0 = part == 0 ? 0 : 1; x1 = part == 0 ? g.n[0] - 1 : 0; }
            else if (ccx == 0) { x0 = part == 0 ? 0 : g.n[0] - 1; x1 = part == 0 ? 1 : g.n[0] - 1; }
            else if (ccx == g.n[0] - 1) { x0 = part == 0 ? ccx - 1 : 0; x1 = part == 0 ? ccx : 0; }
            else { x0 = part == 0 ? ccx - 1 : 1; x1 = part == 0 ? ccx + 1 : 0; }
            if (x0 <= x1) {
                start = cell_start[row0 + x0];
                len = cell_start[row0 + x1 + 1] - start;
            }
        }
        const int incl = warp_incl_scan_i(len, lane);
        if (half && row < ny * nz && g.lo[1] + row % ny == 0 && g.lo[2] + row / ny == 0 && part == 0) {
            s_own[0] = incl - len;
            s_own[1] = incl;
        }
        if (lane < 18) { r_start[lane] = start; r_pref[lane] = incl - len; }
        if (lane == 17) r_pref[18] = incl;                   // lanes >= 18 contribute 0
    }
    __syncthreads();
    const int ncand = r_pref[18];
    const int nchunks = (ncand + CT_CAND - 1) / CT_CAND;
    const int per_hit = (MODE == CT_PAIRS && out.transpose) ? 2 : 1;
    for (int ch = 0; ch < nchunks; ++ch) {
        const int k0 = ch * CT_CAND, kn = min(CT_CAND, ncand - k0);
        // ---- stage the candidates of this chunk: a warp per range, coalesced
        for (int r = w; r < 18; r += CT_WARPS) {
            const int p0 = r_pref[r], len = r_pref[r + 1] - p0, st = r_start[r];
            const int kb = max(0, k0 - p0), ke = min(len, k0 + CT_CAND - p0);   // part of the range in this chunk
            for (int k = kb + lane; k < ke; k += 32) {
                const int kk = p0 + k - k0;
                const SortedAtom<T> a = spos[st + k];
                cx[kk] = a.x; cy[kk] = a.y; cz[kk] = a.z; cidx[kk] = a.idx;
                cb[kk] = sbatch[st + k];
            }
        }
        __syncthreads();
        for (int g0 = cs; g0 < ce; g0 += CT_GROUP) {
            const int gn = min(CT_GROUP, ce - g0);
            // ---- pass 1: hit lists
#pragma unroll
            for (int q = 0; q < CT_GROUP / CT_WARPS; ++q) {
                const int al = w + q * CT_WARPS;
                if (al >= gn) continue;                      // warp-uniform
                const SortedAtom<T> a = spos[g0 + al];
                const long long ba = sbatch[g0 + al];
                const bool row_ok = MODE == CT_PAIRS || a.idx < out.n_rows;
                int cnt = 0;
                if (row_ok)
                    for (int kb = 0; kb < kn; kb += 32) {
                        const int k = kb + lane;
                        bool hit = false;
                        if (k < kn) {
                            const int iq = cidx[k];
                            if (MODE != CT_PAIRS && iq == a.idx) {
                                hit = true;                  // self loop closes the row whatever the cutoffs are
                            } else if (cb[k] == ba && (MODE != CT_PAIRS || iq < a.idx ||
                                                       (half && (k0 + k < s_own[0] || k0 + k >= s_own[1])))) {
                                // delta in the orientation it is emitted in: pos[larger index] - pos[smaller index]
                                const bool sw = MODE == CT_PAIRS && iq > a.idx;
                                T dx = sw ? rn_sub(cx[k], a.x) : rn_sub(a.x, cx[k]);
                                T dy = sw ? rn_sub(cy[k], a.y) : rn_sub(a.y, cy[k]);
                                T dz = sw ? rn_sub(cz[k], a.z) : rn_sub(a.z, cz[k]);
                                if (periodic) {
                                    dx = min_image_fast(dx, g.L[0]);
                                    dy = min_image_fast(dy, g.L[1]);
                                    dz = min_image_fast(dz, g.L[2]);
                                }
                                const T d2 = dist2(dx, dy, dz);
                                hit = d2 < hi2 && d2 >= lo2;
                            }
                        }
                        const unsigned m = __ballot_sync(0xffffffffu, hit);
                        if (hit) hits[al][cnt + __popc(m & ((1u << lane) - 1u))] = (unsigned short)k;
                        cnt += __popc(m);
                    }
                if (lane == 0) s_cnt[al] = cnt;
            }
            __syncthreads();
            // ---- reserve output slots: one global atomic per (chunk, centre group)
            if (MODE == CT_PAIRS) {
                if (w == 0) {
                    int items = 0;
                    if (lane < gn) items = s_cnt[lane] * per_hit + ((out.loop && ch == 0) ? 1 : 0);
                    const int incl = warp_incl_scan_i(items, lane);
                    if (lane < CT_GROUP) s_off[lane] = incl - items;
                    const int total = __shfl_sync(0xffffffffu, incl, 31);
                    if (lane == 0) s_base = total > 0 ? atomicAdd(out.counter, total) : 0;
                }
                __syncthreads();
            }
            // ---- pass 2: dense output
#pragma unroll
            for (int q = 0; q < CT_GROUP / CT_WARPS; ++q) {
                const int al = w + q * CT_WARPS;
                if (al >= gn) continue;
                const int cnt = s_cnt[al];
                const SortedAtom<T> a = spos[g0 + al];
                long long base = 0;
                if (MODE == CT_PAIRS) {
                    base = (long long)s_base + s_off[al];
                } else {
                    // the row's running length across chunks lives in row_cnt (zeroed by the host before each
                    // pass); only this warp touches the entry
                    if (a.idx >= out.n_rows || cnt == 0) continue;
                    int run = 0;
                    if (lane == 0) { run = out.row_cnt[a.idx]; out.row_cnt[a.idx] = run + cnt; }
                    if (MODE == CT_CSR_COUNT) continue;
                    run = __shfl_sync(0xffffffffu, run, 0);
                    base = (long long)out.row_ptr[a.idx] + run;
                }
                for (int hb = 0; hb < cnt; hb += 32) {
                    const int hI = hb + lane;
                    if (hI < cnt) {
                        const int k = hits[al][hI];
                        const bool sw = MODE == CT_PAIRS && cidx[k] > a.idx;
                        T dx = sw ? rn_sub(cx[k], a.x) : rn_sub(a.x, cx[k]);
                        T dy = sw ? rn_sub(cy[k], a.y) : rn_sub(a.y, cy[k]);
                        T dz = sw ? rn_sub(cz[k], a.z) : rn_sub(a.z, cz[k]);
                        if (periodic) {
                            dx = min_image_fast(dx, g.L[0]);
                            dy = min_image_fast(dy, g.L[1]);
                            dz = min_image_fast(dz, g.L[2]);
                        }
                        const T d = sqrt_rn(dist2(dx, dy, dz));
                        if (MODE == CT_PAIRS) {
                            long long slot = base + (long long)hI * per_hit;
                            const int pi = sw ? cidx[k] : a.idx, pj = sw ? a.idx : cidx[k];
                            if (slot < out.max_pairs) {
                                out.nbr[slot] = pi; out.nbr[out.max_pairs + slot] = pj;
                                out.deltas[3 * slot] = dx; out.deltas[3 * slot + 1] = dy; out.deltas[3 * slot + 2] = dz;
                                out.dist[slot] = d;
                            }
                            if (per_hit == 2 && ++slot < out.max_pairs) {
                                out.nbr[slot] = pj; out.nbr[out.max_pairs + slot] = pi;
                                out.deltas[3 * slot] = -dx; out.deltas[3 * slot + 1] = -dy; out.deltas[3 * slot + 2] = -dz;
                                out.dist[slot] = d;
                            }
                        } else {
                            const long long slot = base + hI;
                            if (slot < out.cap) {
                                float4 gm = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (d > T(0)) {
                                    const float inv = 1.0f / (float)d;
                                    gm = make_float4((float)dx * inv, (float)dy * inv, (float)dz * inv, (float)d);
                                }
                                out.col[slot] = cidx[k];
                                out.geom[slot] = gm;
                            }
                        }
                    }
                }
                if (MODE == CT_PAIRS && out.loop && ch == 0 && lane == 0) {
                    const long long slot = base + (long long)cnt * per_hit;
                    if (slot < out.max_pairs) {
                        out.nbr[slot] = a.idx; out.nbr[out.max_pairs + slot] = a.idx;
                        out.deltas[3 * slot] = T(0); out.deltas[3 * slot + 1] = T(0); out.deltas[3 * slot + 2] = T(0);
                        out.dist[slot] = T(0);
                    }
                }
            }
            __syncthreads();                                 // hit lists and counters are reused by the next group
        }
    }
}

}  // namespace nbr
