// tc_gemm.cuh -- general  C[M,N] = epi(A[M,K] * W[N,K]^T + bias)  on the tcgen05 tensor cores, fp32-level accuracy
// by the three-product tf32 split (see tc_mix.cuh for the scheme and the PTX wrappers).  Used for the node GEMMs that
// are not the grouped channel mix: TensorNet's embedding / output / head MLPs and every ET projection
// (torchmd_et.py:249-262 q/k/v/o/vec projections, their transposes in the backward), K up to 640.
//
// Differences from tc_mix_kernel: the weight does not stay resident (K and N are arbitrary), so a ring stage holds a
// 32-channel K chunk of BOTH operands (raw -> hi in place, lo beside it: 64 KB per stage, 3 stages) and the split warps
// round both; TMEM holds one chunk's partial product at a time and the chunks are summed in fp32 round-to-nearest by two
// epilogue warpgroups (the tensor-core accumulator truncates: see the MMA issuer); tiles are 128 x 128 over a persistent grid, N-tiles of one M-tile adjacent in the schedule so that the
// A rows are fetched from HBM once and re-read from L2.  TMA zero-fills rows beyond M / N and channels beyond K and
// clips the stores, so M, N, K only need to be multiples of 4.  Epilogues as in gemm.cuh (bias, activation with the
// pre-activation saved, multiply by act'(saved), accumulate = TMA reduce-add).
#pragma once
#include "tc_mix.cuh"

namespace tcgemm {
using namespace tcmix;

constexpr int GSTAGES = 3;
constexpr int GSMEM_BYTES = (4 * GSTAGES + CSTAGES) * CHUNK_FLOATS * 4 + 256 + 1024;

struct GParams {
    const float* bias;
    const float* aux;     // EPI_MULGRAD: saved pre-activation [M, ldaux]
    int M, N, K, ldaux;
    int m_tiles, n_tiles;
    int epi, act, accumulate, has_c2;
};
struct GMaps {
    CUtensorMap a, w, c, c2;
};

constexpr int GTHREADS_TC = 512;   // + a second epilogue warpgroup (warps 12-15): each thread keeps 64 fp32 partial sums

static __global__ void __launch_bounds__(GTHREADS_TC, 1)
tc_gemm_kernel(const __grid_constant__ GMaps M, const __grid_constant__ GParams P) {
    pdl_sync();
    extern __shared__ unsigned char tcg_smem_raw[];
    unsigned char* sm = tcg_smem_raw + ((1024u - (s32(tcg_smem_raw) & 1023u)) & 1023u);
    float* ring = reinterpret_cast<float*>(sm);                 // stage s: [Ahi | Alo | Bhi | Blo], 16 KB each
    float* Cst = ring + 4 * GSTAGES * CHUNK_FLOATS;             // one staging buffer per epilogue warpgroup
    uint64_t* bars = reinterpret_cast<uint64_t*>(Cst + CSTAGES * CHUNK_FLOATS);
    uint64_t* full_raw = bars;            // [GSTAGES] TMA -> split
    uint64_t* lo_ready = bars + 3;        // [GSTAGES] split -> MMA
    uint64_t* empty = bars + 6;           // [GSTAGES] MMA -> TMA
    uint64_t* tmem_full = bars + 9;       // [2] MMA -> epilogue (one K chunk's partial product)
    uint64_t* tmem_empty = bars + 11;     // [2] epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = P.m_tiles * P.n_tiles;
    const int rank = blockIdx.x, nctas = gridDim.x;
    const int my_tiles = rank < total ? (total - rank + nctas - 1) / nctas : 0;
    const int kchunks = (P.K + KC - 1) / KC;

    if (threadIdx.x == 0) {
        for (int s = 0; s < GSTAGES; ++s) { mbar_init(&full_raw[s], 1); mbar_init(&lo_ready[s], 128); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], 256); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_async_smem();
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "r"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer: one K chunk of A rows and of W rows per stage
        if (lane == 0) {
            int it = 0;
            for (int t = 0; t < my_tiles; ++t) {
                const int tile = rank + t * nctas;
                const int m0 = (tile / P.n_tiles) * TILE, n0 = (tile % P.n_tiles) * TILE;
                for (int c = 0; c < kchunks; ++c, ++it) {
                    const int s = it % GSTAGES;
                    float* st = ring + 4 * s * CHUNK_FLOATS;
                    mbar_wait(&empty[s], ((it / GSTAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_raw[s], 2 * CHUNK_FLOATS * 4);
                    tma_2d(st, &M.a, c * KC, m0, &full_raw[s]);
                    tma_2d(st + 2 * CHUNK_FLOATS, &M.w, c * KC, n0, &full_raw[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer.  The tensor-core accumulator truncates (round toward zero), which biases long
        // same-sign sums; so TMEM only ever holds the partial product of ONE 32-channel chunk (12 MMAs, the two
        // correction products first), and the chunks are summed in fp32 round-to-nearest by the epilogue warps.
        if (lane == 0) {
            int it = 0;
            for (int t = 0; t < my_tiles; ++t) {
                for (int c = 0; c < kchunks; ++c, ++it) {
                    const int s = it % GSTAGES, b = it & 1;
                    mbar_wait(&tmem_empty[b], ((it >> 1) & 1) ^ 1);
                    mbar_wait(&lo_ready[s], (it / GSTAGES) & 1);
                    tc_fence_after();
                    const uint32_t d = tmem + (uint32_t)(b * TILE);
                    const uint32_t ah = s32(ring + 4 * s * CHUNK_FLOATS), al = ah + CHUNK_FLOATS * 4;
                    const uint32_t bh = al + CHUNK_FLOATS * 4, bl = bh + CHUNK_FLOATS * 4;
#pragma unroll
                    for (int k = 0; k < KC / 8; ++k) {
                        const uint32_t off = (uint32_t)k * 32u;
                        tc_mma_tf32(d, smem_desc(ah + off), smem_desc(bl + off), IDESC, k != 0);
                        tc_mma_tf32(d, smem_desc(al + off), smem_desc(bh + off), IDESC, 1);
                    }
#pragma unroll
                    for (int k = 0; k < KC / 8; ++k) {
                        const uint32_t off = (uint32_t)k * 32u;
                        tc_mma_tf32(d, smem_desc(ah + off), smem_desc(bh + off), IDESC, 1);
                    }
                    tc_commit(&empty[s]);
                    tc_commit(&tmem_full[b]);
                }
            }
        }
    } else if ((warp >= 4 && warp < 8) || warp >= 12) {
        // ================= epilogue: two warpgroups, thread = one row x 64 columns of the tile
        const int grp = warp >= 12 ? 1 : 0;
        const int wq = warp & 3;                         // TMEM lane quarter this warp may read
        const int row = wq * 32 + lane;
        const bool issuer = (threadIdx.x & 127) == 0;    // first thread of the warpgroup
        float* cbuf = Cst + grp * CHUNK_FLOATS;
        auto group_barrier = [&]() {
            if (grp == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
            else asm volatile("bar.sync 2, 128;" ::: "memory");
        };
        auto stage_out = [&](const float* v, const CUtensorMap* map, int col0, int m0, bool add) {
            if (issuer) tma_store_wait_read<0>();        // the previous store of this group has read the buffer
            group_barrier();
            float* buf = cbuf + row * KC;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(buf + ((j ^ (row & 7)) << 2)) =
                    make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            fence_async_smem();
            group_barrier();
            if (issuer) {
                if (add)
                    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
                                 "r"(s32(cbuf)), "r"(col0), "r"(m0)
                                 : "memory");
                else
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
                                 "r"(s32(cbuf)), "r"(col0), "r"(m0)
                                 : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        };
        float acc[64];
        int it = 0;
        for (int t = 0; t < my_tiles; ++t) {
            const int tile = rank + t * nctas;
            const int m0 = (tile / P.n_tiles) * TILE, n0 = (tile % P.n_tiles) * TILE;
            const int m = m0 + row;
#pragma unroll
            for (int j = 0; j < 64; ++j) acc[j] = 0.f;
            for (int c = 0; c < kchunks; ++c, ++it) {
                const int b = it & 1;
                mbar_wait(&tmem_full[b], (it >> 1) & 1);
                tc_fence_after();
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    float v[32];
                    tc_ld32(tmem + ((uint32_t)(wq * 32) << 16) + (uint32_t)(b * TILE + grp * 64 + half * 32), v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc[half * 32 + j] += v[j];
                }
                tc_fence_before();
                mbar_arrive(&tmem_empty[b]);
            }
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int nb = n0 + grp * 64 + half * 32;
                if (nb >= P.N) continue;                 // warpgroup-uniform
                float* v = acc + half * 32;
                if (P.bias) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (nb + j < P.N) v[j] += __ldg(P.bias + nb + j);
                }
                if (P.epi == EPI_ACT) {
                    if (P.has_c2) stage_out(v, &M.c2, nb, m0, false);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = act_apply(P.act, v[j]);
                } else if (P.epi == EPI_MULGRAD) {
                    if (m < P.M) {
                        const float* ax = P.aux + (size_t)m * P.ldaux + nb;
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            if (nb + j < P.N) {   // N % 4 == 0
                                const float4 q = *reinterpret_cast<const float4*>(ax + j);
                                v[j] *= act_grad(P.act, q.x); v[j + 1] *= act_grad(P.act, q.y);
                                v[j + 2] *= act_grad(P.act, q.z); v[j + 3] *= act_grad(P.act, q.w);
                            }
                        }
                    }
                }
                stage_out(v, &M.c, nb, m0, P.accumulate != 0);
            }
        }
        if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    } else if (warp >= 8 && warp < 12) {
        // ================= split warps: A chunk and W chunk of the stage
        const int tid = threadIdx.x - 256;
        int it = 0;
        for (int t = 0; t < my_tiles; ++t)
            for (int c = 0; c < kchunks; ++c, ++it) {
                const int s = it % GSTAGES;
                mbar_wait(&full_raw[s], (it / GSTAGES) & 1);
                float4* st = reinterpret_cast<float4*>(ring + 4 * s * CHUNK_FLOATS);
#pragma unroll
                for (int i = 0; i < CHUNK_FLOATS / 4 / 128; ++i) {
                    split_inplace(st, st + CHUNK_FLOATS / 4, i * 128 + tid);
                    split_inplace(st + 2 * (CHUNK_FLOATS / 4), st + 3 * (CHUNK_FLOATS / 4), i * 128 + tid);
                }
                fence_async_smem();
                mbar_arrive(&lo_ready[s]);
            }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}

// 2-D row-major fp32 matrix [rows, cols] with leading dimension ld; box = 32 columns x 128 rows, 128-byte swizzle
static inline bool make_map_2d_raw(CUtensorMap* m, const float* p, int64_t rows, int64_t cols, int64_t ld);
static inline bool make_map_2d(CUtensorMap* m, const float* p, int64_t rows, int64_t cols, int64_t ld) {
    return tcmix::map_cached(m, tcmix::MapKey{p, rows, cols, ld, 2},
                             [&](CUtensorMap* o) { return make_map_2d_raw(o, p, rows, cols, ld); });
}
static inline bool make_map_2d_raw(CUtensorMap* m, const float* p, int64_t rows, int64_t cols, int64_t ld) {
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)KC, (cuuint32_t)TILE};
    const cuuint32_t es[2] = {1, 1};
    return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(p), dims, strides, box, es,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static inline bool supported(int epi, const GemmP& p) {
    if (epi == EPI_SCALE3) return false;
    if (p.gs != 1 || p.rpa != 1 || p.c0 != 0) return false;
    if ((p.K & 3) || (p.N & 3) || (p.lda & 3) || (p.ldc & 3) || (p.ldaux & 3)) return false;
    if (((uintptr_t)p.A | (uintptr_t)p.W | (uintptr_t)p.C | (uintptr_t)p.C2 | (uintptr_t)p.aux) & 15) return false;
    return encode_fn() != nullptr;
}

static inline int launch(int epi, const GemmP& p, cudaStream_t stream) {
    KTimer ktimer__("node_gemm_tc", stream);
    static bool attr_set = false;
    if (!attr_set) {
        TMDN_CUDA(cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GSMEM_BYTES));
        attr_set = true;
    }
    if (p.M <= 0 || p.N <= 0) return TMDN_OK;
    GParams P;
    memset(&P, 0, sizeof(P));
    GMaps M;
    memset(&M, 0, sizeof(M));
    P.bias = p.bias; P.aux = p.aux; P.M = p.M; P.N = p.N; P.K = p.K; P.ldaux = p.ldaux;
    P.m_tiles = (int)cdiv(p.M, TILE); P.n_tiles = (int)cdiv(p.N, TILE);
    P.epi = epi; P.act = p.act; P.accumulate = p.accumulate; P.has_c2 = (epi == EPI_ACT && p.C2) ? 1 : 0;
    bool ok = make_map_2d(&M.a, p.A, p.M, p.K, p.lda) && make_map_2d(&M.w, p.W, p.N, p.K, p.K) &&
              make_map_2d(&M.c, p.C, p.M, p.N, p.ldc);
    if (ok && P.has_c2) ok = make_map_2d(&M.c2, p.C2, p.M, p.N, p.ldc);
    if (!ok) {
        tmdn_set_error("tc_gemm: cuTensorMapEncodeTiled failed");
        return TMDN_EINVAL;
    }
    const long long tiles = (long long)P.m_tiles * P.n_tiles;
    const unsigned grid = (unsigned)(tiles < tmdn_sm_count() ? tiles : tmdn_sm_count());
    launch_k(tc_gemm_kernel, grid, GTHREADS_TC, GSMEM_BYTES, stream, M, P);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

}  // namespace tcgemm
