// tc_mix.cuh -- the grouped channel mix of TensorNet on the 5th-generation tensor cores (tcgen05, TMEM, TMA).
//
//   Y[a, c, :] = X[a, c, :] * W_g(c)^T     a = atom, c = 0..8 irreducible component, g(c) = I | A | S, W_g [H,H]
//
// (tensornet.py:273-284, 386-408: linears_tensor applied to the I, A and S parts).  With 9 N rows and K = N = H = 128
// this is the one dense contraction of the model; at fp32 FFMA rate it was 31 % of the 98k-atom step.  fp32 accuracy
// is kept with the three-product split  x*w ~ xh*wh + xh*wl + xl*wh  (xh = tf32(x), xl = tf32(x - xh)), every
// product a kind::tf32 MMA with fp32 accumulation in TMEM; the dropped xl*wl term and the rounding of the low parts
// are O(2^-22) relative.
//
// One persistent CTA per SM works on 128-row tiles of one component group (a tile = na atoms x gs components:
// 128x1, 42x3 or 25x5 rows, fetched as ONE 3-D TMA box from the [N,9,H] tensor, so no permuted copy of X exists):
//   warp 0       TMA producer: W once (4 K-chunks of 32), then X chunks [rows, 32] into a 2-stage ring (SWIZZLE_128B)
//   warps 8-11   split: round the landed chunk to tf32 in place, write the low part beside it, fence to the async proxy
//   warp 1       one thread issues 3 x 4 tcgen05.mma (M128 N128 K8) per chunk; tcgen05.commit frees the ring slot and
//                hands the chunk's partial product to the epilogue
//   warps 4-7, 12-15  epilogue: tcgen05.ld each chunk's partial product (double buffered in TMEM, 2 x 128 columns)
//                and sum the chunks in fp32 registers (the tensor-core accumulator truncates), apply the
//                fused scale, stage 32-column slabs in shared memory and TMA-store (or reduce-add) them through
//                the same 3-D box
// Per tile the tensor pipe needs ~1.7 us and HBM (64 KB in, 64 KB out) ~2.9 us, so the kernel is bound by HBM.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include <unordered_map>
#include "common.cuh"
#include "gemm.cuh"

namespace tcmix {

constexpr int TILE = 128;            // rows per tile = MMA M; also N and K (= hidden channels)
constexpr int KC = 32;               // K floats per chunk: one 128-byte swizzle row
constexpr int NCHUNK = TILE / KC;    // 4
constexpr int STAGES = 2;           // X ring (raw/hi + lo per stage)
constexpr int CSTAGES = 2;          // output staging buffers [128 rows x 32 columns]
constexpr int CHUNK_FLOATS = TILE * KC;          // 4096 floats = 16 KB
constexpr int THREADS = 384;          // tc_gemm.cuh
constexpr int MIX_THREADS = 512;      // 16 warps: TMA, MMA, alloc, -, epilogue group 0 (4), split (4), epilogue group 1 (4)
constexpr int SMEM_BYTES = (2 * NCHUNK + 2 * STAGES + CSTAGES) * CHUNK_FLOATS * 4 + 256 + 1024;
constexpr int TMEM_COLS = 256;

struct Problem {
    float* C;
    float* C2;
    const float* aux;
    int gs, c0, na, tiles, g;
};
struct Params {
    Problem p[3];
    int cta_off[4];
    int n_atoms;
    int epi, accumulate;
};
struct Maps {
    CUtensorMap x[3], w[3], c[3], c2[3];   // per component group: input box, weight, output box, second output
};

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(s32(b)), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            s32(dst)),
        "l"(map), "r"(x), "r"(y), "r"(s32(bar))
        : "memory");
}
__device__ __forceinline__ void tma_3d(void* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            s32(dst)),
        "l"(map), "r"(x), "r"(y), "r"(z), "r"(s32(bar))
        : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* src, int x, int y, int z, bool add) {
    if (add)
        asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
                     "r"(s32(src)), "r"(x), "r"(y), "r"(z)
                     : "memory");
    else
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
                     "r"(s32(src)), "r"(x), "r"(y), "r"(z)
                     : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 32 consecutive fp32 columns of the accumulator: thread = lane (row), v[j] = column j
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}
__device__ __forceinline__ float tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
// K-major operand tile, rows of 128 bytes, SWIZZLE_128B: 8-row groups 1024 bytes apart (cute/arch/mma_sm100_desc.hpp
// SmemDescriptor: start >> 4 | LBO << 16 | SBO << 32 | version 1 << 46 | layout SWIZZLE_128B (2) << 61)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// cute InstrDescriptor: D fp32 (1 << 4), A and B tf32 (2 << 7, 2 << 10), both K-major, N >> 3 at 17, M >> 4 at 24
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TILE >> 3) << 17) | ((uint32_t)(TILE >> 4) << 24);

__device__ __forceinline__ void split_inplace(float4* hi, float4* lo, int idx) {
    const float4 v = hi[idx];
    float4 h, l;
    h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
    // x - hi is exact in fp32; round it to tf32 as well (the MMA would truncate it: a one-sided 2^-23 error per product)
    l.x = tf32_rna(v.x - h.x); l.y = tf32_rna(v.y - h.y); l.z = tf32_rna(v.z - h.z); l.w = tf32_rna(v.w - h.w);
    hi[idx] = h;
    lo[idx] = l;
}

static __global__ void __launch_bounds__(MIX_THREADS, 1)
tc_mix_kernel(const __grid_constant__ Maps M, const __grid_constant__ Params P) {
    pdl_sync();
    extern __shared__ unsigned char tc_smem_raw[];
    unsigned char* sm = tc_smem_raw + ((1024u - (s32(tc_smem_raw) & 1023u)) & 1023u);   // SWIZZLE_128B: 1024-byte base
    float* Bhi = reinterpret_cast<float*>(sm);
    float* Blo = Bhi + NCHUNK * CHUNK_FLOATS;
    float* Ahi = Blo + NCHUNK * CHUNK_FLOATS;
    float* Alo = Ahi + STAGES * CHUNK_FLOATS;
    float* Cst = Alo + STAGES * CHUNK_FLOATS;
    uint64_t* bars = reinterpret_cast<uint64_t*>(Cst + CSTAGES * CHUNK_FLOATS);
    uint64_t* full_raw = bars;            // [STAGES] TMA -> split
    uint64_t* lo_ready = bars + 3;        // [STAGES] split -> MMA
    uint64_t* empty = bars + 6;           // [STAGES] MMA -> TMA
    uint64_t* b_full = bars + 9;          // W landed
    uint64_t* b_ready = bars + 10;        // W split
    uint64_t* tmem_full = bars + 11;      // [2] MMA -> epilogue
    uint64_t* tmem_empty = bars + 13;     // [2] epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int pi = 0;
    while (pi < 2 && (int)blockIdx.x >= P.cta_off[pi + 1]) ++pi;
    const int rank = blockIdx.x - P.cta_off[pi], nctas = P.cta_off[pi + 1] - P.cta_off[pi];
    const Problem pr = P.p[pi];
    const CUtensorMap* mA = &M.x[pi];
    const CUtensorMap* mW = &M.w[pi];
    const int my_tiles = rank < pr.tiles ? (pr.tiles - rank + nctas - 1) / nctas : 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_raw[s], 1); mbar_init(&lo_ready[s], 128); mbar_init(&empty[s], 1); }
        mbar_init(b_full, 1);
        mbar_init(b_ready, 128);
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
        // ================= TMA producer
        if (lane == 0) {
            mbar_expect_tx(b_full, NCHUNK * CHUNK_FLOATS * 4);
            for (int c = 0; c < NCHUNK; ++c) tma_2d(Bhi + c * CHUNK_FLOATS, mW, c * KC, 0, b_full);
            const uint32_t bytes = (uint32_t)(KC * 4 * pr.gs * pr.na);
            int it = 0;
            for (int t = 0; t < my_tiles; ++t) {
                const int atom0 = (rank + t * nctas) * pr.na;
                for (int c = 0; c < NCHUNK; ++c, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_raw[s], bytes);
                    tma_3d(Ahi + s * CHUNK_FLOATS, mA, c * KC, pr.c0, atom0, &full_raw[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer.  The tensor-core accumulator truncates (round toward zero), which biases long
        // same-sign sums (measured 1.2e-6 relative on the energy of a 5184-atom box); so TMEM only ever holds the
        // partial product of ONE 32-channel chunk (12 MMAs, the two correction products first), and the four chunks
        // are summed in fp32 round-to-nearest by the epilogue warps.
        if (lane == 0) {
            mbar_wait(b_ready, 0);
            tc_fence_after();
            int it = 0;
            for (int t = 0; t < my_tiles; ++t) {
                for (int c = 0; c < NCHUNK; ++c, ++it) {
                    const int s = it % STAGES, b = it & 1;
                    mbar_wait(&tmem_empty[b], ((it >> 1) & 1) ^ 1);
                    mbar_wait(&lo_ready[s], (it / STAGES) & 1);
                    tc_fence_after();
                    const uint32_t d = tmem + (uint32_t)(b * TILE);
                    const uint32_t ah = s32(Ahi + s * CHUNK_FLOATS), al = s32(Alo + s * CHUNK_FLOATS);
                    const uint32_t bh = s32(Bhi + c * CHUNK_FLOATS), bl = s32(Blo + c * CHUNK_FLOATS);
#pragma unroll
                    for (int k = 0; k < KC / 8; ++k) {
                        const uint32_t off = (uint32_t)k * 32u;   // 8 tf32 = 32 bytes along K inside the swizzle row
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
        // ================= epilogue: two warpgroups, thread = one row x 64 columns of the tile; warp w reads TMEM lanes
        // 32 (w % 4) ...  Per chunk: tcgen05.ld the partial product and add it (fp32, round to nearest).  Per tile: a
        // 32-column slab goes registers -> swizzled staging buffer -> ONE TMA store (or reduce-add) through the same
        // (channel, component, atom) box as the input, so global memory sees whole 128-byte row segments.
        const int grp = warp >= 12 ? 1 : 0;
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const bool issuer = (threadIdx.x & 127) == 0;
        const CUtensorMap* mC = &M.c[pi];
        const CUtensorMap* mC2 = &M.c2[pi];
        const bool scale = P.epi == EPI_SCALE3;
        float* cbuf = Cst + grp * CHUNK_FLOATS;
        auto group_barrier = [&]() {
            if (grp == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
            else asm volatile("bar.sync 2, 128;" ::: "memory");
        };
        auto stage_out = [&](const float* v, const CUtensorMap* map, int col0, int atom0, bool add) {
            if (issuer) tma_store_wait_read<0>();            // the previous store of this group has read the buffer
            group_barrier();
            float* buf = cbuf + row * KC;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(buf + ((j ^ (row & 7)) << 2)) =
                    make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            fence_async_smem();
            group_barrier();
            if (issuer) tma_store_3d(map, cbuf, col0, pr.c0, atom0, add);
        };
        float acc[64];
        int it = 0;
        for (int t = 0; t < my_tiles; ++t) {
            const int atom0 = (rank + t * nctas) * pr.na;
            const int atom = atom0 + row / pr.gs;
            const bool valid = row < pr.gs * pr.na && atom < P.n_atoms;
#pragma unroll
            for (int j = 0; j < 64; ++j) acc[j] = 0.f;
            for (int c = 0; c < NCHUNK; ++c, ++it) {
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
                const int col0 = grp * 64 + half * 32;
                float* v = acc + half * 32;
                if (scale) {
                    if (pr.C2) stage_out(v, mC2, col0, atom0, false);
                    if (valid) {
                        const float4* sc = reinterpret_cast<const float4*>(pr.aux + ((size_t)atom * 3 + pr.g) * TILE + col0);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 q = sc[j];
                            v[4 * j] *= q.x; v[4 * j + 1] *= q.y; v[4 * j + 2] *= q.z; v[4 * j + 3] *= q.w;
                        }
                    }
                }
                stage_out(v, mC, col0, atom0, P.accumulate != 0);
            }
        }
        if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    } else if (warp >= 8 && warp < 12) {
        // ================= split warps
        const int tid = threadIdx.x - 256;
        mbar_wait(b_full, 0);
        for (int i = 0; i < NCHUNK * CHUNK_FLOATS / 4 / 128; ++i)
            split_inplace(reinterpret_cast<float4*>(Bhi), reinterpret_cast<float4*>(Blo), i * 128 + tid);
        fence_async_smem();
        mbar_arrive(b_ready);
        int it = 0;
        for (int t = 0; t < my_tiles; ++t)
            for (int c = 0; c < NCHUNK; ++c, ++it) {
                const int s = it % STAGES;
                mbar_wait(&full_raw[s], (it / STAGES) & 1);
                float4* hi = reinterpret_cast<float4*>(Ahi + s * CHUNK_FLOATS);
                float4* lo = reinterpret_cast<float4*>(Alo + s * CHUNK_FLOATS);
#pragma unroll
                for (int i = 0; i < CHUNK_FLOATS / 4 / 128; ++i) split_inplace(hi, lo, i * 128 + tid);
                fence_async_smem();
                mbar_arrive(&lo_ready[s]);
            }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}
// An encoded map is a pure function of (pointer, extents, box): memoised, because one encode costs about a
// microsecond of host time and a mix launch needs 9-12 of them (invisible inside a CUDA graph, a real cost in
// eager calls and in the staged, spatially decomposed path).
struct MapKey {
    const void* p;
    int64_t a, b, c;
    int kind;
    bool operator==(const MapKey& o) const { return p == o.p && a == o.a && b == o.b && c == o.c && kind == o.kind; }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = (size_t)(uintptr_t)k.p * 0x9E3779B97F4A7C15ull;
        h ^= (size_t)k.a * 0xC2B2AE3D27D4EB4Full + (size_t)k.b * 0x165667B19E3779F9ull + (size_t)k.c * 31 + (size_t)k.kind;
        return h;
    }
};
template <class Build> static inline bool map_cached(CUtensorMap* out, const MapKey& key, Build build) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return true; }
    if (!build(out)) return false;
    if (cache.size() > 8192) cache.clear();
    cache.emplace(key, *out);
    return true;
}
// X viewed as [atoms][9][H] fp32; box = 32 channels x gs components x na atoms, 128-byte swizzle
static inline bool make_map_x_raw(CUtensorMap* m, const float* X, int64_t n_atoms, int H, int gs, int na) {
    const cuuint64_t dims[3] = {(cuuint64_t)H, 9, (cuuint64_t)n_atoms};
    const cuuint64_t strides[2] = {(cuuint64_t)H * 4, (cuuint64_t)H * 36};
    const cuuint32_t box[3] = {(cuuint32_t)KC, (cuuint32_t)gs, (cuuint32_t)na};
    const cuuint32_t es[3] = {1, 1, 1};
    return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(X), dims, strides, box, es,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static inline bool make_map_w_raw(CUtensorMap* m, const float* W, int H);
static inline bool make_map_x(CUtensorMap* m, const float* X, int64_t n_atoms, int H, int gs, int na) {
    return map_cached(m, MapKey{X, n_atoms, H, gs * 1024 + na, 0},
                      [&](CUtensorMap* o) { return make_map_x_raw(o, X, n_atoms, H, gs, na); });
}
static inline bool make_map_w(CUtensorMap* m, const float* W, int H) {
    return map_cached(m, MapKey{W, H, 0, 0, 1}, [&](CUtensorMap* o) { return make_map_w_raw(o, W, H); });
}
static inline bool make_map_w_raw(CUtensorMap* m, const float* W, int H) {
    const cuuint64_t dims[2] = {(cuuint64_t)H, (cuuint64_t)H};
    const cuuint64_t strides[1] = {(cuuint64_t)H * 4};
    const cuuint32_t box[2] = {(cuuint32_t)KC, (cuuint32_t)TILE};
    const cuuint32_t es[2] = {1, 1};
    return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(W), dims, strides, box, es,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static inline bool supported(const GemmBatch& b) {
    if (b.n != 3) return false;
    for (int g = 0; g < 3; ++g) {
        const GemmP& p = b.p[g];
        if (p.N != TILE || p.K != TILE || p.lda != TILE || p.ldc != TILE || p.rpa != 9 || p.bias) return false;
        if (p.A != b.p[0].A || p.C != b.p[0].C) return false;
    }
    return encode_fn() != nullptr;
}

// the three grouped problems of mix_batch() (gemm.cuh conventions: gs, c0, rpa = 9, M = atoms * gs) in one launch
static inline int launch(int epi, const GemmBatch& b, cudaStream_t stream) {
    KTimer ktimer__("node_gemm_tc", stream);
    static bool attr_set = false;
    if (!attr_set) {
        TMDN_CUDA(cudaFuncSetAttribute(tc_mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        attr_set = true;
    }
    const int n_atoms = b.p[0].M / b.p[0].gs;
    if (n_atoms <= 0) return TMDN_OK;
    Params P;
    memset(&P, 0, sizeof(P));
    Maps M;
    memset(&M, 0, sizeof(M));
    long long total_tiles = 0;
    for (int g = 0; g < 3; ++g) {
        const GemmP& p = b.p[g];
        Problem& q = P.p[g];
        q.C = p.C; q.C2 = p.C2; q.aux = p.aux;
        q.gs = p.gs; q.c0 = p.c0; q.g = p.g;
        q.na = TILE / p.gs;
        q.tiles = (int)cdiv(n_atoms, q.na);
        total_tiles += q.tiles;
        bool ok = make_map_x(&M.x[g], p.A, n_atoms, TILE, p.gs, q.na) && make_map_w(&M.w[g], p.W, TILE) &&
                  make_map_x(&M.c[g], p.C, n_atoms, TILE, p.gs, q.na);
        if (ok && epi == EPI_SCALE3 && p.C2) ok = make_map_x(&M.c2[g], p.C2, n_atoms, TILE, p.gs, q.na);
        if (!ok) {
            tmdn_set_error("tc_mix: cuTensorMapEncodeTiled failed");
            return TMDN_EINVAL;
        }
    }
    // CTAs per problem in proportion to its tiles (at least one, at most its tile count)
    const int sms = tmdn_sm_count();
    int ctas[3], used = 0;
    for (int g = 0; g < 3; ++g) {
        long long c = total_tiles <= sms ? P.p[g].tiles : (long long)P.p[g].tiles * sms / total_tiles;
        if (c < 1) c = 1;
        if (c > P.p[g].tiles) c = P.p[g].tiles;
        ctas[g] = (int)c;
        used += ctas[g];
    }
    P.cta_off[0] = 0;
    for (int g = 0; g < 3; ++g) P.cta_off[g + 1] = P.cta_off[g] + ctas[g];
    P.n_atoms = n_atoms;
    P.epi = epi;
    P.accumulate = b.p[0].accumulate;
    launch_k(tc_mix_kernel, (unsigned)used, MIX_THREADS, SMEM_BYTES, stream, M, P);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}

}  // namespace tcmix
