// et.cu -- Equivariant Transformer energy + analytic forces for sm_100a (fp32).
//
// Replaces TorchMD_ET.forward / NeighborEmbedding / EquivariantMultiHeadAttention (torchmdnet/models/torchmd_et.py:
// 154-187, 272-347; models/utils.py:73-108), the EquivariantScalar head (models/utils.py:496-522,
// output_modules.py:111-115), the reduce (output_modules.py:27-43) and the autograd force path of TorchMD_Net.forward
// (models/model.py:286-298).  Stage list, buffer names and adjoint formulas are those of oracle/staged_et.py, which is
// checked against torch.autograd on the CPU in fp64.
//
// Layout: x [N,H]; vec [N,3,H]; q|k|v in one [N,5H] buffer with v part-major [3][H]; a thread owns one (atom,
// channel); per-head sums are segmented warp shuffles over head_dim consecutive lanes.  Edges are CSR by target;
// the edge set is symmetric, so row i also evaluates the reverse edge of every entry and collects the source-side
// adjoints (k_i, v_i, vec_i) -- no transposed graph, no float atomics, bit-reproducible.
// The per-edge [E,H], [E,3H], [E,3,H] message tensors of the reference never exist.
#include "common.cuh"
#include "gemm.cuh"
#include "node_gemm.cuh"
#include "edge_common.cuh"
#include "table_grid.cuh"
#include "scan.cuh"
#include <string>
#include <vector>

namespace {

struct Tab {
    const float4* t;
    TabGrid grid;   // table_grid.cuh
};
// NF filters per interval: value and d/dd from the same cubic.  Split into the loads (tab_load) and the arithmetic
// (tab_poly) so that an edge kernel can put the loads of several edges in flight before it starts to compute.
template <int NF> struct TabRow { float4 c[NF]; float s, ds; };
template <int NF>
__device__ __forceinline__ void tab_load(const Tab& tb, float d, int H, int h, TabRow<NF>& r) {
    int k;
    tab_locate(tb.grid, d, k, r.s, r.ds);
#pragma unroll
    for (int g = 0; g < NF; ++g) r.c[g] = __ldg(&tb.t[((size_t)k * NF + g) * H + h]);
}
template <int NF>
__device__ __forceinline__ void tab_poly(const Tab& tb, const TabRow<NF>& r, float (&f)[NF], float (&df)[NF]) {
    const float s2 = 2.0f * r.s, s3 = 3.0f * r.s * r.s;     // df/ds = c.y + 2 s c.z + 3 s^2 c.w as two FFMA per filter
#pragma unroll
    for (int g = 0; g < NF; ++g) {
        const float4 c = r.c[g];
        f[g] = c.x + r.s * (c.y + r.s * (c.z + r.s * c.w));
        df[g] = fmaf(s3, c.w, fmaf(s2, c.z, c.y)) * r.ds;
    }
}
template <int NF>
__device__ __forceinline__ void tab_eval(const Tab& tb, float d, int H, int h, float (&f)[NF], float (&df)[NF]) {
    TabRow<NF> r;
    tab_load<NF>(tb, d, H, h, r);
    tab_poly<NF>(tb, r, f, df);
}
// sum over the hd consecutive lanes of a head (hd power of two <= 32); result in every lane of the segment
__device__ __forceinline__ float seg_sum(float v, int hd) {
    for (int o = hd >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// HDC > 0: head width known at compile time (fully unrolled butterfly)
template <int HDC> __device__ __forceinline__ float seg_sum_t(float v, int hd) {
    if constexpr (HDC > 0) {
#pragma unroll
        for (int o = HDC >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    } else {
        return seg_sum(v, hd);
    }
}

// Four segmented sums at once (the adjoint needs q.k and g_x.v of both directions of an edge): the first two butterfly
// steps halve the number of live values instead of duplicating them, the remaining steps run on ONE value per lane, four
// shuffles hand every lane all four results: 9 shuffles + 5 adds instead of 16 + 16 at hd = 16 (a warp shuffle costs four
// issue slots of the SM, and this kernel is bound by them).  hd >= 4 only.
template <int HDC>
__device__ __forceinline__ void seg_sum4_t(float& a0, float& a1, float& a2, float& a3, int hd) {
    if constexpr (HDC >= 4) {
        const int lane = threadIdx.x & 31;
        const bool b0 = lane & 1, b1 = lane & 2;
        float x0 = b0 ? a2 : a0, x1 = b0 ? a3 : a1;
        x0 += __shfl_xor_sync(0xffffffffu, b0 ? a0 : a2, 1);
        x1 += __shfl_xor_sync(0xffffffffu, b0 ? a1 : a3, 1);
        float k = b1 ? x1 : x0;
        k += __shfl_xor_sync(0xffffffffu, b1 ? x0 : x1, 2);
#pragma unroll
        for (int o = 4; o < HDC; o <<= 1) k += __shfl_xor_sync(0xffffffffu, k, o);
        // lane (b1, b0) of every group of four holds: (0,0) a0, (1,0) a1, (0,1) a2, (1,1) a3
        const int base = lane & ~3;
        a0 = __shfl_sync(0xffffffffu, k, base);
        a1 = __shfl_sync(0xffffffffu, k, base + 2);
        a2 = __shfl_sync(0xffffffffu, k, base + 1);
        a3 = __shfl_sync(0xffffffffu, k, base + 3);
    } else {
        a0 = seg_sum_t<HDC>(a0, hd); a1 = seg_sum_t<HDC>(a1, hd); a2 = seg_sum_t<HDC>(a2, hd); a3 = seg_sum_t<HDC>(a3, hd);
    }
}

__global__ void et_prep_kernel(const int64_t* __restrict__ z, int* __restrict__ z32, int n, int max_z,
                               const int32_t* __restrict__ num_found, long long max_edges, int32_t* __restrict__ status,
                               const int64_t* __restrict__ batch) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        const int f = *num_found;
        status[0] = f;
        status[1] = f > max_edges ? 1 : 0;
    }
    if (i >= n) return;
    if (i > 0 && batch[i] < batch[i - 1]) status[3] = 1;   // the molecule ranges are found by binary search: batch must be sorted
    long long v = z[i];
    if (v < 0 || v >= max_z) {
        status[2] = 1;
        v = 0;
    }
    z32[i] = (int)v;
}

// ------------------------------------------------------------------ embedding (+ neighbour embedding edge sum)
// out[i, 0:H] = emb[z_i];  if NE: out[i, H:2H] = sum_{e in row i, j != i} ne_emb[z_j] * fne(d_e)   (models/utils.py:92-103)
template <bool NE>
__global__ void __launch_bounds__(1024)
et_embed_fwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom,
             const int* __restrict__ z32, const float* __restrict__ emb, const float* __restrict__ ne_emb, Tab tb, int N,
             int H, float* __restrict__ out, int ldo) {
    pdl_sync();
    extern __shared__ float smem[];
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    float acc[1] = {0.f};
    if (NE && valid) {
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e = r0 + sl; e < r1; e += S) {
            const int j = col[e];
            if (j == i) continue;
            float f[1], df[1];
            tab_eval<1>(tb, geom[e].w, H, h, f, df);
            acc[0] += ne_emb[(size_t)z32[j] * H + h] * f[0];
        }
    }
    if (NE) slice_reduce<1>(acc, s_part, H);
    if (!valid || sl != 0) return;
    out[(size_t)i * ldo + h] = emb[(size_t)z32[i] * H + h];
    if (NE) out[(size_t)i * ldo + H + h] = acc[0];
}

// ------------------------------------------------------------------ LayerNorm over H (block = (H,1,A))
__global__ void et_ln_fwd(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                          float* __restrict__ y, int ldy, float* __restrict__ stat, int N, int H) {
    pdl_sync();
    __shared__ float s_red[EDGE_RED_FLOATS * 4];
    const int h = threadIdx.x;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    const float v = valid ? x[(size_t)i * H + h] : 0.f;
    float s[1] = {v};
    atom_reduce<1>(s, s_red);
    const float mean = s[0] / (float)H;
    float q[1] = {(v - mean) * (v - mean)};
    atom_reduce<1>(q, s_red);
    const float rstd = rsqrtf(q[0] / (float)H + 1e-5f);
    if (!valid) return;
    y[(size_t)i * ldy + h] = (v - mean) * rstd * w[h] + b[h];
    if (h == 0) { stat[2 * (size_t)i] = mean; stat[2 * (size_t)i + 1] = rstd; }
}
// gx (+)= LN^T(gy): gy has row stride ldg
__global__ void et_ln_bwd(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ stat,
                          const float* __restrict__ gy, int ldg, float* __restrict__ gx, int accumulate, int N, int H) {
    pdl_sync();
    __shared__ float s_red[EDGE_RED_FLOATS * 4];
    const int h = threadIdx.x;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    float xh = 0.f, dg = 0.f, rstd = 0.f;
    if (valid) {
        const float mean = stat[2 * (size_t)i];
        rstd = stat[2 * (size_t)i + 1];
        xh = (x[(size_t)i * H + h] - mean) * rstd;
        dg = gy[(size_t)i * ldg + h] * w[h];
    }
    float r[2] = {dg, dg * xh};
    atom_reduce<2>(r, s_red);
    if (!valid) return;
    const float v = (dg - r[0] / (float)H - xh * r[1] / (float)H) * rstd;
    const size_t o = (size_t)i * H + h;
    gx[o] = accumulate ? gx[o] + v : v;
}

// ------------------------------------------------------------------ attention edge kernel (torchmd_et.py:314-347)
// xa[i] = sum_e v0_j dv0 * attn_e ;  va[i] = sum_e vec_j (v1_j dv1) + (v2_j dv2) dhat_e,  dhat = -u
// attn_e(head) = attn_act(sum_{c in head} q_i k_j dk) * C(d_e)
// NB = edges whose loads are issued together.  Unlike the TensorNet messages these kernels are bound by arithmetic
// (segmented shuffles, activation, cutoff), so occupancy wins: NB = 1 (measured: NB = 2 at 128 registers is 40 % slower).
constexpr int ET_FWD_THREADS = 1024, ET_BWD_THREADS = 512;
// HC / HDC > 0: hidden width and head width known at compile time (row offsets become immediates, the head sums unroll)
template <int NB = 1, int HC = 0, int HDC = 0>
__global__ void __launch_bounds__(ET_FWD_THREADS)
et_edge_fwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom, Tab tb,
            float lo, float hi, int attn_act, int hd, const float* __restrict__ qkv, const float* __restrict__ vec,
            int N, int H_, float* __restrict__ xa, float* __restrict__ va) {
    pdl_sync();
    const int H = HC > 0 ? HC : H_;
    extern __shared__ float smem[];
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    const int H5 = 5 * H;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    if (valid) {
        const float qi = qkv[(size_t)i * H5 + h];
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e0 = r0 + sl; e0 < r1; e0 += NB * S) {
            float4 g[NB];
            TabRow<4> tr[NB];
            float kj[NB], v0[NB], v1[NB], v2[NB], vx[NB], vy[NB], vz[NB];
            bool on[NB];
#pragma unroll
            for (int b = 0; b < NB; ++b) {           // all loads of the batch first (warp-uniform predicate)
                const int e = e0 + b * S;
                on[b] = e < r1;
                const int ee = on[b] ? e : e0;
                const int j = col[ee];
                g[b] = geom[ee];
                tab_load<4>(tb, g[b].w, H, h, tr[b]);
                const float* pj = qkv + (size_t)j * H5 + h;
                kj[b] = pj[H]; v0[b] = pj[2 * H]; v1[b] = pj[3 * H]; v2[b] = pj[4 * H];
                const float* vj = vec + (size_t)j * 3 * H + h;
                vx[b] = vj[0]; vy[b] = vj[H]; vz[b] = vj[2 * H];
            }
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                if (!on[b]) continue;
                float f[4], df[4];
                tab_poly<4>(tb, tr[b], f, df);
                const float s = seg_sum_t<HDC>(qi * kj[b] * f[0], hd);
                const float attn = act_apply(attn_act, s);          // f[1] carries the cutoff C(d)
                acc[0] += v0[b] * f[1] * attn;
                const float a1 = v1[b] * f[2], a2 = v2[b] * f[3];
                acc[1] += vx[b] * a1 - a2 * g[b].x;
                acc[2] += vy[b] * a1 - a2 * g[b].y;
                acc[3] += vz[b] * a1 - a2 * g[b].z;
            }
        }
    }
    slice_reduce<4>(acc, s_part, H);
    if (!valid || sl != 0) return;
    xa[(size_t)i * H + h] = acc[0];
    float* o = va + (size_t)i * 3 * H + h;
    o[0] = acc[1]; o[H] = acc[2]; o[2 * H] = acc[3];
}

// x' = x + (sum_xyz vec1 vec2) o2 + o3 ;  vec' = vec + vec3 o1 + va        (torchmd_et.py:278-312, 181-184)
__global__ void et_update_fwd(const float* __restrict__ x, const float* __restrict__ vec, const float* __restrict__ vp,
                              const float* __restrict__ o, const float* __restrict__ va, float* __restrict__ xn,
                              float* __restrict__ vecn, long long total, int H) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    const float* p = vp + i * 9 * H + h;     // [3 xyz][3H]
    const float o1 = o[i * 3 * H + h], o2 = o[i * 3 * H + H + h], o3 = o[i * 3 * H + 2 * H + h];
    float dot = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v1 = p[(size_t)c * 3 * H], v2 = p[(size_t)c * 3 * H + H], v3 = p[(size_t)c * 3 * H + 2 * H];
        dot += v1 * v2;
        const size_t k = (i * 3 + c) * H + h;
        vecn[k] = vec[k] + v3 * o1 + va[k];
    }
    xn[idx] = x[idx] + dot * o2 + o3;
}

// ------------------------------------------------------------------ gated equivariant block pieces (models/utils.py:496-522)
// cat[i, hin + c] = || vb[i,:,c] ||
__global__ void et_gate_norm_fwd(const float* __restrict__ vb, float* __restrict__ cat, long long total, int hin) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / hin;
    const int c = (int)(idx % hin);
    const float* p = vb + i * 3 * hin + c;
    const float a = p[0], b = p[hin], d = p[2 * hin];
    cat[i * 2 * hin + hin + c] = sqrtf(a * a + b * b + d * d);
}
// block with scalar activation: xo = act(h2[:, :out]) -> xout (row stride ldx); vo = h2[:, out:] * vec2
__global__ void et_gate_out_fwd(const float* __restrict__ h2, const float* __restrict__ vec2, int act,
                                float* __restrict__ xout, int ldx, float* __restrict__ vo, long long total, int out) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / out;
    const int c = (int)(idx % out);
    xout[i * ldx + c] = act_apply(act, h2[i * 2 * out + c]);
    const float gate = h2[i * 2 * out + out + c];
#pragma unroll
    for (int k = 0; k < 3; ++k) vo[(i * 3 + k) * out + c] = gate * vec2[(i * 3 + k) * out + c];
}
// e_i = (w . a_i + b) * std ; one warp per atom
__global__ void et_head_energy(const float* __restrict__ a, const float* __restrict__ w, const float* __restrict__ b,
                               float stdv, float* __restrict__ e_atom, int N, int K) {
    pdl_sync();
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= N) return;
    float s = 0.f;
    for (int k = lane; k < K; k += 32) s += a[(size_t)i * K + k] * w[k];
    s = warp_sum(s);
    if (lane == 0) e_atom[i] = (s + b[0]) * stdv;
}
// status[1] / status[2] (edge capacity exceeded, z out of range) poison the result with NaN: see tn_mol_reduce
__global__ void __launch_bounds__(128) et_mol_reduce(const float* __restrict__ e_atom, const int64_t* __restrict__ batch,
                                                     int N, float mean, float* __restrict__ energy,
                                                     const int32_t* __restrict__ status) {
    pdl_sync();
    __shared__ float s_part[128];
    const long long m = blockIdx.x;
    int a, b;
    block_molecule_range(batch, N, m, a, b);
    float s = 0.f;
    for (int i = a + threadIdx.x; i < b; i += 128) s += e_atom[i];
    s_part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (threadIdx.x < o) s_part[threadIdx.x] += s_part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) energy[m] = (status[1] | status[2] | status[3]) ? __int_as_float(0x7fc00000) : s_part[0] + mean;
}

// ================================================================== backward kernels
// d[i,n] = std * w[n] * act'(h0[i,n])      (seed of dE/d(update_net.0 output) of the last gated block)
__global__ void et_head_seed_bwd(const float* __restrict__ h0, const float* __restrict__ w, float stdv, int act,
                                 float* __restrict__ d, long long total, int K) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    d[idx] = stdv * w[idx % K] * act_grad(act, h0[idx]);
}
// g_vb[i,:,c] = g_cat[i,hin+c] * vb[i,:,c] / ||vb[i,:,c]||   (0 where the norm is 0: models/utils.py:499-512)
__global__ void et_gate_norm_bwd(const float* __restrict__ vb, const float* __restrict__ cat,
                                 const float* __restrict__ g_cat, float* __restrict__ g_vb, long long total, int hin) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / hin;
    const int c = (int)(idx % hin);
    const float nrm = cat[i * 2 * hin + hin + c];
    const float s = nrm > 0.f ? g_cat[i * 2 * hin + hin + c] / nrm : 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) g_vb[(i * 3 + k) * hin + c] = s * vb[(i * 3 + k) * hin + c];
}
// adjoint of et_gate_out_fwd: g_h2 = [g_xo * act'(h2_x) | sum_xyz g_vo vec2] ; g_vec2 = g_vo * gate
__global__ void et_gate_out_bwd(const float* __restrict__ h2, const float* __restrict__ vec2, int act,
                                const float* __restrict__ g_xo, int ldx, const float* __restrict__ g_vo,
                                float* __restrict__ g_h2, float* __restrict__ g_vec2, long long total, int out) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / out;
    const int c = (int)(idx % out);
    g_h2[i * 2 * out + c] = g_xo[i * ldx + c] * act_grad(act, h2[i * 2 * out + c]);
    const float gate = h2[i * 2 * out + out + c];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const size_t o = (i * 3 + k) * out + c;
        const float gv = g_vo[o];
        s += gv * vec2[o];
        g_vec2[o] = gv * gate;
    }
    g_h2[i * 2 * out + out + c] = s;
}
// adjoint of et_update_fwd: g_o = [sum_xyz g_vec vec3 | g_x vec_dot | g_x] ; g_vp = [g_x o2 vec2 | g_x o2 vec1 | g_vec o1]
// g_vec_new = g_vec (residual path; the vec_proj^T and edge terms are added afterwards)
__global__ void et_update_bwd(const float* __restrict__ g_x, const float* __restrict__ g_vec,
                              const float* __restrict__ vp, const float* __restrict__ o, float* __restrict__ g_o,
                              float* __restrict__ g_vp, float* __restrict__ g_vec_new, long long total, int H) {
    pdl_sync();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t i = idx / H;
    const int h = (int)(idx % H);
    const float gx = g_x[idx];
    const float o1 = o[i * 3 * H + h], o2 = o[i * 3 * H + H + h];
    const float* p = vp + i * 9 * H + h;
    float* gp = g_vp + i * 9 * H + h;
    float dot = 0.f, go1 = 0.f;
    const float gd = gx * o2;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v1 = p[(size_t)c * 3 * H], v2 = p[(size_t)c * 3 * H + H], v3 = p[(size_t)c * 3 * H + 2 * H];
        const size_t k = (i * 3 + c) * H + h;
        const float gv = g_vec[k];
        dot += v1 * v2;
        go1 += gv * v3;
        gp[(size_t)c * 3 * H] = gd * v2;
        gp[(size_t)c * 3 * H + H] = gd * v1;
        gp[(size_t)c * 3 * H + 2 * H] = gv * o1;
        g_vec_new[k] = gv;
    }
    g_o[i * 3 * H + h] = go1;
    g_o[i * 3 * H + H + h] = gx * dot;
    g_o[i * 3 * H + 2 * H + h] = gx;
}

// Adjoint of et_edge_fwd evaluated in row i for both directions of every entry (i, j):
//   A: edge (i <- j): g_q_i, and -G_A on pos_i          B: reverse edge (j <- i): g_k_i, g_v_i, g_vec_i, +G_B on pos_i
//   G = g_d dhat + (g_dhat - (g_dhat.dhat) dhat)/d  is linear in (g_d, g_dhat), so every channel accumulates its share
template <int NB = 1, int HC = 0, int HDC = 0>
__global__ void __launch_bounds__(ET_BWD_THREADS, 2)
et_edge_bwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom, Tab tb,
            float lo, float hi, int attn_act, int hd, const float* __restrict__ qkv, const float* __restrict__ vec,
            const float* __restrict__ g_xa, const float* __restrict__ g_va, int N, int H_, float* __restrict__ g_qkv,
            float* __restrict__ g_vec_new, float* __restrict__ gpos) {
    pdl_sync();
    const int H = HC > 0 ? HC : H_;
    extern __shared__ float smem[];
    float* s_red = smem;
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    const int H5 = 5 * H;
    float acc[8];   // g_q, g_k, g_v0, g_v1, g_v2, g_vec xyz
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[c] = 0.f;
    float gp[3] = {0.f, 0.f, 0.f};
    if (valid) {
        const float* pi = qkv + (size_t)i * H5 + h;
        const float qi = pi[0], ki = pi[H], vi0 = pi[2 * H], vi1 = pi[3 * H], vi2 = pi[4 * H];
        const float* vci = vec + (size_t)i * 3 * H + h;
        const float vix = vci[0], viy = vci[H], viz = vci[2 * H];
        const float gxi = g_xa[(size_t)i * H + h];
        const float* gvi = g_va + (size_t)i * 3 * H + h;
        const float gvix = gvi[0], gviy = gvi[H], gviz = gvi[2 * H];
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e0 = r0 + sl; e0 < r1; e0 += NB * S) {
            int jj[NB];
            float4 gg[NB];
            TabRow<4> tr[NB];
            float L[NB][12];                         // q k v0 v1 v2 | vec xyz | g_xa | g_va xyz of atom j
            bool on[NB];
#pragma unroll
            for (int b = 0; b < NB; ++b) {           // all loads of the batch first (warp-uniform predicate)
                const int e = e0 + b * S;
                on[b] = e < r1;
                const int ee = on[b] ? e : e0;
                const int j = col[ee];
                jj[b] = j;
                gg[b] = geom[ee];
                tab_load<4>(tb, gg[b].w, H, h, tr[b]);
                const float* pj = qkv + (size_t)j * H5 + h;
#pragma unroll
                for (int q = 0; q < 5; ++q) L[b][q] = pj[q * H];
                const float* vcj = vec + (size_t)j * 3 * H + h;
                L[b][5] = vcj[0]; L[b][6] = vcj[H]; L[b][7] = vcj[2 * H];
                L[b][8] = g_xa[(size_t)j * H + h];
                const float* gvj = g_va + (size_t)j * 3 * H + h;
                L[b][9] = gvj[0]; L[b][10] = gvj[H]; L[b][11] = gvj[2 * H];
            }
#pragma unroll
            for (int b = 0; b < NB; ++b) {
            if (!on[b]) continue;
            const int j = jj[b];
            const float4 g = gg[b];
            const float hx = -g.x, hy = -g.y, hz = -g.z;   // dhat of direction A; direction B uses +u
            float f[4], df[4];                                  // f[1], df[1] carry the cutoff C(d) and its derivative
            tab_poly<4>(tb, tr[b], f, df);
            const float qj = L[b][0], kj = L[b][1], vj0 = L[b][2], vj1 = L[b][3], vj2 = L[b][4];
            const float vjx = L[b][5], vjy = L[b][6], vjz = L[b][7];
            const float gxj = L[b][8];
            const float gvjx = L[b][9], gvjy = L[b][10], gvjz = L[b][11];
            // ---- A (edge i <- j, dhat = -u) and B (edge j <- i, dhat' = +u).  The issue slots bound this kernel: shared
            // products are named once, the position gradient is assembled from its closed form (below)
            const float tA = qi * kj, uA = gxi * vj0, tB = qj * ki, uB = gxj * vi0;
            float sA = tA * f[0], gattA = uA * f[1], sB = tB * f[0], gattB = uB * f[1];
            seg_sum4_t<HDC>(sA, gattA, sB, gattB, hd);
            float attnA, dactA, attnB, dactB;
            act_both(attn_act, sA, attnA, dactA);               // one sigmoid for value and derivative
            act_both(attn_act, sB, attnB, dactB);
            const float gsA = gattA * dactA, gsB = gattB * dactB;
            acc[0] = fmaf(gsA * kj, f[0], acc[0]);
            acc[1] = fmaf(gsB * qj, f[0], acc[1]);
            const float gv1A = fmaf(gvix, vjx, fmaf(gviy, vjy, gviz * vjz));
            const float gv2A = fmaf(gvix, hx, fmaf(gviy, hy, gviz * hz));
            const float gv1B = fmaf(gvjx, vix, fmaf(gvjy, viy, gvjz * viz));
            const float gv2B = fmaf(gvjx, g.x, fmaf(gvjy, g.y, gvjz * g.z));
            acc[2] = fmaf(gxj * attnB, f[1], acc[2]);
            acc[3] = fmaf(gv1B, f[2], acc[3]);
            acc[4] = fmaf(gv2B, f[3], acc[4]);
            const float s1B = vi1 * f[2];
            acc[5] = fmaf(gvjx, s1B, acc[5]); acc[6] = fmaf(gvjy, s1B, acc[6]); acc[7] = fmaf(gvjz, s1B, acc[7]);
            if (j != i) {   // the self loop has d = 0 and no position dependence
                const float gdA = fmaf(gsA * tA, df[0], fmaf(uA * attnA, df[1], fmaf(gv1A * vj1, df[2], gv2A * vj2 * df[3])));
                const float gdB = fmaf(gsB * tB, df[0], fmaf(uB * attnB, df[1], fmaf(gv1B * vi1, df[2], gv2B * vi2 * df[3])));
                // -G_A + G_B with G = g_d dhat + (g_dhat - (g_dhat . dhat) dhat) / d, g_dhat(A) = vj2 f3 g_va_i, g_dhat(B) = vi2 f3 g_va_j:
                //   = u (g_dA + g_dB - wA gv2A - wB gv2B) + wB g_va_j - wA g_va_i,   wA = vj2 f3 / d, wB = vi2 f3 / d
                const float fd = f[3] * __frcp_rn(g.w);
                const float wA = vj2 * fd, wB = vi2 * fd;
                const float cu = (gdA + gdB) - fmaf(wA, gv2A, wB * gv2B);
                gp[0] += fmaf(g.x, cu, fmaf(wB, gvjx, -wA * gvix));
                gp[1] += fmaf(g.y, cu, fmaf(wB, gvjy, -wA * gviy));
                gp[2] += fmaf(g.z, cu, fmaf(wB, gvjz, -wA * gviz));
            }
            }
        }
    }
    slice_reduce<8>(acc, s_part, H);
    if (valid && sl == 0) {
        float* o = g_qkv + (size_t)i * H5 + h;
        o[0] = acc[0]; o[H] = acc[1]; o[2 * H] = acc[2]; o[3 * H] = acc[3]; o[4 * H] = acc[4];
        float* gv = g_vec_new + (size_t)i * 3 * H + h;
        gv[0] += acc[5]; gv[H] += acc[6]; gv[2 * H] += acc[7];
    }
    atom_reduce<3>(gp, s_red);
    if (valid && h == 0 && sl == 0) {
        gpos[3 * (size_t)i] += gp[0];
        gpos[3 * (size_t)i + 1] += gp[1];
        gpos[3 * (size_t)i + 2] += gp[2];
    }
}

// neighbour-embedding adjoint + force write-out: forces[i] = -(gpos[i] + sum_e u_e * g_d(e)),
// g_d(e) = sum_h (g_xn_i ne_emb[z_j] + g_xn_j ne_emb[z_i]) fne'(d_e)       (both directions of the entry; dhat = -u)
template <bool NE>
__global__ void __launch_bounds__(512)
et_final_bwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom,
             const int* __restrict__ z32, const float* __restrict__ ne_emb, Tab tb, const float* __restrict__ g_xn,
             const float* __restrict__ gpos, float* __restrict__ forces, int N, int H,
             const int32_t* __restrict__ status) {
    pdl_sync();
    extern __shared__ float smem[];
    float* s_red = smem;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    float gp[3] = {0.f, 0.f, 0.f};
    if (NE && valid) {
        const float gi = g_xn[(size_t)i * H + h], ei = ne_emb[(size_t)z32[i] * H + h];
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e = r0 + sl; e < r1; e += S) {
            const int j = col[e];
            if (j == i) continue;
            const float4 g = geom[e];
            float f[1], df[1];
            tab_eval<1>(tb, g.w, H, h, f, df);
            const float gd = (gi * ne_emb[(size_t)z32[j] * H + h] + g_xn[(size_t)j * H + h] * ei) * df[0];
            gp[0] += g.x * gd; gp[1] += g.y * gd; gp[2] += g.z * gd;
        }
    }
    atom_reduce<3>(gp, s_red);
    if (valid && h == 0 && sl == 0) {
        const bool bad = (status[1] | status[2] | status[3]) != 0;
        const float nan = __int_as_float(0x7fc00000);
        forces[3 * (size_t)i] = bad ? nan : -(gpos[3 * (size_t)i] + gp[0]);
        forces[3 * (size_t)i + 1] = bad ? nan : -(gpos[3 * (size_t)i + 1] + gp[1]);
        forces[3 * (size_t)i + 2] = bad ? nan : -(gpos[3 * (size_t)i + 2] + gp[2]);
    }
}

// ================================================================== workspace
struct EtWs {
    int32_t *num_found, *row_ptr, *col;
    float* geom;
    int* z32;
    void* graph_ws;
    size_t graph_ws_bytes;
    float *cat0, *X[TMDN_MAX_LAYERS + 1], *VEC[TMDN_MAX_LAYERS + 1], *stat[TMDN_MAX_LAYERS + 1], *xl,
        *QKV[TMDN_MAX_LAYERS], *VP[TMDN_MAX_LAYERS], *O[TMDN_MAX_LAYERS], *xa, *va;
    // head
    float *cat1, *vb1, *vec2_1, *h0_1, *a0_1, *h2_1, *v1, *cat2, *vb2, *h0_2, *a0_2, *e_atom;
    // backward
    float *d_a0_2, *g_cat2, *g_vb2, *g_v1, *g_h2, *g_vec2, *g_h0, *g_cat1, *g_vb, *g_x, *g_vecA, *g_vecB, *g_o, *g_vp,
        *g_xa, *g_qkv, *g_xl, *g_xn, *gpos;
};
struct EtEntry { std::string name; size_t off, bytes; };
template <typename T> T* ecarve(Arena& a, std::vector<EtEntry>* names, const char* name, size_t n) {
    const size_t o = a.off;
    T* p = a.get<T>(n);
    if (names) names->push_back({name, o, n * sizeof(T)});
    return p;
}

int carve_et(Arena& a, const tmdn_et_model* m, int64_t N, int64_t P, EtWs& w, std::vector<EtEntry>* names,
             const double* box, int strategy) {
    const size_t H = m->hidden, L = m->num_layers, n = (size_t)N, H2 = H / 2;
    char buf[32];
    w.num_found = ecarve<int32_t>(a, names, "num_found", 4);
    w.row_ptr = ecarve<int32_t>(a, names, "row_ptr", n + 1);
    w.col = ecarve<int32_t>(a, names, "col", (size_t)P);
    w.geom = ecarve<float>(a, names, "geom", 4 * (size_t)P);
    w.z32 = ecarve<int>(a, names, "z32", n);
    w.graph_ws_bytes = tmdn_csr_build_workspace_bytes(strategy, N, box, m->cutoff_upper);
    if (w.graph_ws_bytes == 0) {
        tmdn_set_error("et: bad box for the cell strategy");
        return TMDN_EINVAL;
    }
    w.graph_ws = ecarve<char>(a, names, "graph_ws", w.graph_ws_bytes);
    w.cat0 = ecarve<float>(a, names, "cat0", n * 2 * H);
    for (size_t l = 0; l <= L; ++l) {
        snprintf(buf, sizeof buf, "x%zu", l); w.X[l] = ecarve<float>(a, names, buf, n * H);
        snprintf(buf, sizeof buf, "vec%zu", l); w.VEC[l] = ecarve<float>(a, names, buf, n * 3 * H);
        snprintf(buf, sizeof buf, "stat%zu", l); w.stat[l] = ecarve<float>(a, names, buf, n * 2);
    }
    w.xl = ecarve<float>(a, names, "xl", n * H);
    for (size_t l = 0; l < L; ++l) {
        snprintf(buf, sizeof buf, "qkv%zu", l); w.QKV[l] = ecarve<float>(a, names, buf, n * 5 * H);
        snprintf(buf, sizeof buf, "vp%zu", l); w.VP[l] = ecarve<float>(a, names, buf, n * 9 * H);
        snprintf(buf, sizeof buf, "o%zu", l); w.O[l] = ecarve<float>(a, names, buf, n * 3 * H);
    }
    w.xa = ecarve<float>(a, names, "xa", n * H);
    w.va = ecarve<float>(a, names, "va", n * 3 * H);
    w.cat1 = ecarve<float>(a, names, "cat1", n * 2 * H);
    w.vb1 = ecarve<float>(a, names, "vb1", n * 3 * H);
    w.vec2_1 = ecarve<float>(a, names, "vec2_1", n * 3 * H2);
    w.h0_1 = ecarve<float>(a, names, "h0_1", n * H);
    w.a0_1 = ecarve<float>(a, names, "a0_1", n * H);
    w.h2_1 = ecarve<float>(a, names, "h2_1", n * H);
    w.v1 = ecarve<float>(a, names, "v1", n * 3 * H2);
    w.cat2 = ecarve<float>(a, names, "cat2", n * H);
    w.vb2 = ecarve<float>(a, names, "vb2", n * 3 * H2);
    w.h0_2 = ecarve<float>(a, names, "h0_2", n * H2);
    w.a0_2 = ecarve<float>(a, names, "a0_2", n * H2);
    w.e_atom = ecarve<float>(a, names, "e_atom", n);
    w.d_a0_2 = ecarve<float>(a, names, "d_a0_2", n * H2);
    w.g_cat2 = ecarve<float>(a, names, "g_cat2", n * H);
    w.g_vb2 = ecarve<float>(a, names, "g_vb2", n * 3 * H2);
    w.g_v1 = ecarve<float>(a, names, "g_v1", n * 3 * H2);
    w.g_h2 = ecarve<float>(a, names, "g_h2", n * H);
    w.g_vec2 = ecarve<float>(a, names, "g_vec2", n * 3 * H2);
    w.g_h0 = ecarve<float>(a, names, "g_h0", n * H);
    w.g_cat1 = ecarve<float>(a, names, "g_cat1", n * 2 * H);
    w.g_vb = ecarve<float>(a, names, "g_vb", n * 3 * H);
    w.g_x = ecarve<float>(a, names, "g_x", n * H);
    w.g_vecA = ecarve<float>(a, names, "g_vecA", n * 3 * H);
    w.g_vecB = ecarve<float>(a, names, "g_vecB", n * 3 * H);
    w.g_o = ecarve<float>(a, names, "g_o", n * 3 * H);
    w.g_vp = ecarve<float>(a, names, "g_vp", n * 9 * H);
    w.g_xa = ecarve<float>(a, names, "g_xa", n * H);
    w.g_qkv = ecarve<float>(a, names, "g_qkv", n * 5 * H);
    w.g_xl = ecarve<float>(a, names, "g_xl", n * H);
    w.g_xn = ecarve<float>(a, names, "g_xn", n * H);
    w.gpos = ecarve<float>(a, names, "gpos", n * 3);
    return TMDN_OK;
}

int check_et(const tmdn_et_model* m) {
    TMDN_CHECK_ARG(m, "et: null model");
    TMDN_CHECK_ARG(m->hidden >= 32 && m->hidden <= 1024 && m->hidden % 32 == 0,
                   "et: hidden_channels must be a multiple of 32 in [32,1024] (got %d)", m->hidden);
    TMDN_CHECK_ARG(m->num_layers >= 0 && m->num_layers <= TMDN_MAX_LAYERS, "et: num_layers out of range");
    TMDN_CHECK_ARG(m->num_heads > 0 && m->hidden % m->num_heads == 0, "et: hidden must be divisible by num_heads");
    const int hd = m->hidden / m->num_heads;
    TMDN_CHECK_ARG(hd <= 32 && (hd & (hd - 1)) == 0, "et: head_dim must be a power of two <= 32 (got %d)", hd);
    TMDN_CHECK_ARG(m->max_z > 0 && m->n_int > 0, "et: bad max_z / table size");
    TMDN_CHECK_ARG(m->cutoff_upper > m->cutoff_lower && m->cutoff_lower >= 0, "et: bad cutoffs");
    TMDN_CHECK_ARG(m->emb && m->out_ln_w && m->out_ln_b, "et: missing weights");
    for (int l = 0; l < m->num_layers; ++l) TMDN_CHECK_ARG(m->layer[l].table, "et: missing filter table of layer %d", l);
    if (m->has_neighbor_embedding) TMDN_CHECK_ARG(m->ne_emb && m->ne_table && m->ne_comb_w, "et: missing NE weights");
    return TMDN_OK;
}
}  // namespace

extern "C" size_t tmdn_et_workspace_bytes(const tmdn_et_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                          int strategy, const double box[9]) {
    (void)n_mol;
    if (check_et(m) != TMDN_OK || n_atoms <= 0 || max_edges <= 0) return 0;
    Arena a(nullptr, 0);
    EtWs w;
    if (carve_et(a, m, n_atoms, max_edges, w, nullptr, box, strategy) != TMDN_OK) return 0;
    return a.off + 256;
}

extern "C" int tmdn_et_workspace_lookup(const tmdn_et_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                        int strategy, const double box[9], const char* name, size_t* offset,
                                        size_t* bytes) {
    (void)n_mol;
    TMDN_TRY(check_et(m));
    TMDN_CHECK_ARG(name && offset && bytes, "lookup: null argument");
    Arena a(nullptr, 0);
    EtWs w;
    std::vector<EtEntry> names;
    TMDN_TRY(carve_et(a, m, n_atoms, max_edges, w, &names, box, strategy));
    for (const auto& e : names)
        if (e.name == name) {
            *offset = e.off;
            *bytes = e.bytes;
            return TMDN_OK;
        }
    tmdn_set_error("lookup: unknown buffer '%s'", name);
    return TMDN_EINVAL;
}

#define ET_K(name, ...)                      \
    do {                                     \
        KTimer ktimer__(name, stream);       \
        __VA_ARGS__;                         \
        TMDN_LAUNCH_CHECK();                 \
    } while (0)

extern "C" int tmdn_et_energy_forces(const tmdn_et_model* m, const int64_t* z, const float* pos, const int64_t* batch,
                                     int64_t n_atoms, int64_t n_mol, const double box[9], int use_periodic,
                                     int strategy, int64_t max_edges, float* energy, float* forces, int32_t* status,
                                     void* workspace, size_t workspace_bytes, tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_TRY(check_et(m));
    TMDN_CHECK_ARG(z && pos && batch && energy && status && workspace, "et: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_mol > 0 && max_edges > 0, "et: bad sizes");
    TMDN_CHECK_ARG(max_edges < (1LL << 31) && n_atoms < (1LL << 31) / 16, "et: sizes exceed int32 indexing");
    TMDN_CHECK_ARG(strategy != TMDN_CELL || box, "et: the cell strategy needs a box");
    const int N = (int)n_atoms, H = m->hidden, L = m->num_layers, act = m->activation, H2 = H / 2;
    const bool tc = m->gemm_mode == TMDN_GEMM_TF32X3;
    const int hd = H / m->num_heads;
    const int64_t P = max_edges;
    const float lo = m->cutoff_lower, hi = m->cutoff_upper;
    Arena arena(workspace, workspace_bytes);
    EtWs w;
    TMDN_TRY(carve_et(arena, m, n_atoms, P, w, nullptr, box, strategy));
    if (!arena.ok) {
        tmdn_set_error("et: workspace too small (%zu < %zu)", workspace_bytes, arena.off);
        return TMDN_EWORKSPACE;
    }
    {
        KTimer ktimer__("neighbors_csr", stream);
        TMDN_TRY(tmdn_csr_build_f32(strategy, pos, batch, n_atoms, box, use_periodic, lo, hi, P, w.row_ptr, w.col, w.geom,
                                    w.num_found, w.graph_ws, w.graph_ws_bytes, stream_));
    }
    TMDN_CUDA(cudaMemsetAsync(status, 0, 4 * sizeof(int32_t), stream));
    ET_K("et_prep", launch_k(et_prep_kernel, (unsigned)cdiv(N, 256), 256, 0, stream, z, w.z32, N, m->max_z, w.num_found,
                                                                                (long long)P, status, batch));
    const float4* geom = (const float4*)w.geom;
    const TabGrid tgrid = make_tab_grid(m->n_int, hi, m->grid_alpha, m->grid_gamma);
    const Tab ne_tab{(const float4*)m->ne_table, tgrid};
    const EdgeLaunch e1 = edge_launch(N, H, 1024, 1), e4 = edge_launch(N, H, ET_FWD_THREADS, 4), e8 = edge_launch(N, H, 512, 8);
    const EdgeLaunch e8b = edge_launch(N, H, ET_BWD_THREADS, 8);
    const int apb = H >= 128 ? 1 : 128 / H;
    const dim3 lblock(H, 1, apb);
    const unsigned lgrid = (unsigned)cdiv(N, apb);
    const long long NH = (long long)N * H;
    const unsigned egrid = (unsigned)cdiv(NH, 256);
    const unsigned hgrid = (unsigned)cdiv((long long)N * H2, 256);

    // ---- embedding (+ neighbour embedding)
    if (m->has_neighbor_embedding) {
        ET_K("et_embed_fwd", launch_k(et_embed_fwd<true>, e1.grid, e1.block, e1.smem, stream, 
                                 w.row_ptr, w.col, geom, w.z32, m->emb, m->ne_emb, ne_tab, N, H, w.cat0, 2 * H));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.cat0, 2 * H, m->ne_comb_w, m->ne_comb_b, w.X[0], H, N, H, 2 * H), stream, tc));
    } else {
        ET_K("et_embed_fwd", launch_k(et_embed_fwd<false>, e1.grid, e1.block, e1.smem, stream, 
                                 w.row_ptr, w.col, geom, w.z32, m->emb, nullptr, ne_tab, N, H, w.X[0], H));
    }
    TMDN_CUDA(cudaMemsetAsync(w.VEC[0], 0, sizeof(float) * 3 * (size_t)NH, stream));
    // ---- attention layers
    for (int l = 0; l < L; ++l) {
        const tmdn_et_layer& ly = m->layer[l];
        const Tab tb{(const float4*)ly.table, tgrid};
        ET_K("et_ln_fwd", launch_k(et_ln_fwd, lgrid, lblock, 0, stream, w.X[l], ly.ln_w, ly.ln_b, w.xl, H, w.stat[l], N, H));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.xl, H, ly.qkv_w, ly.qkv_b, w.QKV[l], 5 * H, N, 5 * H, H), stream, tc));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.VEC[l], H, ly.vp_w, nullptr, w.VP[l], 3 * H, 3 * N, 3 * H, H), stream, tc));
        if (H == 128 && hd == 16)
            ET_K("et_edge_fwd", (launch_k(et_edge_fwd<1, 128, 16>, e4.grid, e4.block, e4.smem, stream, 
                                    w.row_ptr, w.col, geom, tb, lo, hi, m->attn_activation, hd, w.QKV[l], w.VEC[l], N, H, w.xa,
                                    w.va)));
        else
            ET_K("et_edge_fwd", (launch_k(et_edge_fwd<1>, e4.grid, e4.block, e4.smem, stream, 
                                    w.row_ptr, w.col, geom, tb, lo, hi, m->attn_activation, hd, w.QKV[l], w.VEC[l], N, H, w.xa,
                                    w.va)));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.xa, H, ly.o_w, ly.o_b, w.O[l], 3 * H, N, 3 * H, H), stream, tc));
        ET_K("et_update_fwd", launch_k(et_update_fwd, egrid, 256, 0, stream, w.X[l], w.VEC[l], w.VP[l], w.O[l], w.va,
                                                                       w.X[l + 1], w.VEC[l + 1], NH, H));
    }
    // ---- out_norm written straight into the first half of cat1 = [xo | vec1]
    ET_K("et_ln_fwd", launch_k(et_ln_fwd, lgrid, lblock, 0, stream, w.X[L], m->out_ln_w, m->out_ln_b, w.cat1, 2 * H, w.stat[L],
                                                              N, H));
    // ---- gated block 1: H -> H/2 (scalar activation)
    const tmdn_et_gated& g1 = m->head[0];
    const tmdn_et_gated& g2 = m->head[1];
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.VEC[L], H, g1.w1, nullptr, w.vb1, H, 3 * N, H, H), stream, tc));
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.VEC[L], H, g1.w2, nullptr, w.vec2_1, H2, 3 * N, H2, H), stream, tc));
    ET_K("et_gate_norm_fwd", launch_k(et_gate_norm_fwd, egrid, 256, 0, stream, w.vb1, w.cat1, NH, H));
    {
        GemmP p = gemm_desc(w.cat1, 2 * H, g1.u0_w, g1.u0_b, w.a0_1, H, N, H, 2 * H);
        p.C2 = w.h0_1; p.act = act;
        TMDN_TRY(node_gemm1(EPI_ACT, p, stream, tc));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.a0_1, H, g1.u2_w, g1.u2_b, w.h2_1, H, N, H, H), stream, tc));
    }
    ET_K("et_gate_out_fwd", launch_k(et_gate_out_fwd, hgrid, 256, 0, stream, w.h2_1, w.vec2_1, act, w.cat2, H, w.v1,
                                                                       (long long)N * H2, H2));
    // ---- gated block 2: H/2 -> 1 (its vector output does not reach the energy)
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.v1, H2, g2.w1, nullptr, w.vb2, H2, 3 * N, H2, H2), stream, tc));
    ET_K("et_gate_norm_fwd", launch_k(et_gate_norm_fwd, hgrid, 256, 0, stream, w.vb2, w.cat2, (long long)N * H2, H2));
    {
        GemmP p = gemm_desc(w.cat2, H, g2.u0_w, g2.u0_b, w.a0_2, H2, N, H2, H);
        p.C2 = w.h0_2; p.act = act;
        TMDN_TRY(node_gemm1(EPI_ACT, p, stream, tc));
    }
    ET_K("et_head_energy", launch_k(et_head_energy, (unsigned)cdiv(N, 4), 128, 0, stream, w.a0_2, g2.u2_w, g2.u2_b, m->std,
                                                                                   w.e_atom, N, H2));
    ET_K("et_mol_reduce", launch_k(et_mol_reduce, (unsigned)n_mol, 128, 0, stream, w.e_atom, batch, N, m->mean, energy, status));
    if (!forces) return TMDN_OK;

    // ================================================================= backward
    TMDN_CUDA(cudaMemsetAsync(w.gpos, 0, sizeof(float) * 3 * (size_t)N, stream));
    // block 2
    ET_K("et_head_seed_bwd", launch_k(et_head_seed_bwd, hgrid, 256, 0, stream, w.h0_2, g2.u2_w, m->std, act, w.d_a0_2,
                                                                         (long long)N * H2, H2));
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.d_a0_2, H2, g2.u0_wT, nullptr, w.g_cat2, H, N, H, H2), stream, tc));
    ET_K("et_gate_norm_bwd", launch_k(et_gate_norm_bwd, hgrid, 256, 0, stream, w.vb2, w.cat2, w.g_cat2, w.g_vb2,
                                                                         (long long)N * H2, H2));
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_vb2, H2, g2.w1T, nullptr, w.g_v1, H2, 3 * N, H2, H2), stream, tc));
    // block 1
    ET_K("et_gate_out_bwd", launch_k(et_gate_out_bwd, hgrid, 256, 0, stream, w.h2_1, w.vec2_1, act, w.g_cat2, H, w.g_v1, w.g_h2,
                                                                       w.g_vec2, (long long)N * H2, H2));
    {
        GemmP p = gemm_desc(w.g_h2, H, g1.u2_wT, nullptr, w.g_h0, H, N, H, H);
        p.aux = w.h0_1; p.ldaux = H; p.act = act;
        TMDN_TRY(node_gemm1(EPI_MULGRAD, p, stream, tc));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_h0, H, g1.u0_wT, nullptr, w.g_cat1, 2 * H, N, 2 * H, H), stream, tc));
    }
    ET_K("et_gate_norm_bwd", launch_k(et_gate_norm_bwd, egrid, 256, 0, stream, w.vb1, w.cat1, w.g_cat1, w.g_vb, NH, H));
    float* g_vec = w.g_vecA;
    float* g_vec_new = w.g_vecB;
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_vb, H, g1.w1T, nullptr, g_vec, H, 3 * N, H, H), stream, tc));
    {
        GemmP p = gemm_desc(w.g_vec2, H2, g1.w2T, nullptr, g_vec, H, 3 * N, H, H2);
        p.accumulate = 1;
        TMDN_TRY(node_gemm1(EPI_NONE, p, stream, tc));
    }
    ET_K("et_ln_bwd", launch_k(et_ln_bwd, lgrid, lblock, 0, stream, w.X[L], m->out_ln_w, w.stat[L], w.g_cat1, 2 * H, w.g_x, 0, N,
                                                              H));
    for (int l = L - 1; l >= 0; --l) {
        const tmdn_et_layer& ly = m->layer[l];
        const Tab tb{(const float4*)ly.table, tgrid};
        ET_K("et_update_bwd", launch_k(et_update_bwd, egrid, 256, 0, stream, w.g_x, g_vec, w.VP[l], w.O[l], w.g_o, w.g_vp,
                                                                       g_vec_new, NH, H));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_o, 3 * H, ly.o_wT, nullptr, w.g_xa, H, N, H, 3 * H), stream, tc));
        {
            GemmP p = gemm_desc(w.g_vp, 3 * H, ly.vp_wT, nullptr, g_vec_new, H, 3 * N, H, 3 * H);
            p.accumulate = 1;
            TMDN_TRY(node_gemm1(EPI_NONE, p, stream, tc));
        }
        if (H == 128 && hd == 16)
            ET_K("et_edge_bwd", (launch_k(et_edge_bwd<1, 128, 16>, e8b.grid, e8b.block, e8b.smem, stream, 
                                    w.row_ptr, w.col, geom, tb, lo, hi, m->attn_activation, hd, w.QKV[l], w.VEC[l], w.g_xa,
                                    g_vec, N, H, w.g_qkv, g_vec_new, w.gpos)));
        else
            ET_K("et_edge_bwd", (launch_k(et_edge_bwd<1>, e8b.grid, e8b.block, e8b.smem, stream, 
                                    w.row_ptr, w.col, geom, tb, lo, hi, m->attn_activation, hd, w.QKV[l], w.VEC[l], w.g_xa,
                                    g_vec, N, H, w.g_qkv, g_vec_new, w.gpos)));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_qkv, 5 * H, ly.qkv_wT, nullptr, w.g_xl, H, N, H, 5 * H), stream, tc));
        ET_K("et_ln_bwd", launch_k(et_ln_bwd, lgrid, lblock, 0, stream, w.X[l], ly.ln_w, w.stat[l], w.g_xl, H, w.g_x, 1, N, H));
        float* t = g_vec; g_vec = g_vec_new; g_vec_new = t;
    }
    if (m->has_neighbor_embedding) {
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_x, H, m->ne_comb_w2T, nullptr, w.g_xn, H, N, H, H), stream, tc));
        ET_K("et_final_bwd", launch_k(et_final_bwd<true>, e8.grid, e8.block, e8.smem, stream, 
                                 w.row_ptr, w.col, geom, w.z32, m->ne_emb, ne_tab, w.g_xn, w.gpos, forces, N, H, status));
    } else {
        ET_K("et_final_bwd", launch_k(et_final_bwd<false>, e8.grid, e8.block, e8.smem, stream, 
                                 w.row_ptr, w.col, geom, w.z32, nullptr, ne_tab, nullptr, w.gpos, forces, N, H, status));
    }
    return TMDN_OK;
}

// ====================================================================================================================
// Graph network (TorchMD_GN): torchmdnet/models/torchmd_gn.py:148-290.  Shares the embedding / neighbour-embedding
// kernels, the table lookup, the head and the force write-out with the transformer above; what is specific are the two
// CFConv edge kernels.  Stage list (names = workspace buffers):
//   x0 = emb(z) [NE]; per layer l: h_l = lin1 x_l; agg = conv(h_l); a2_l = act(m2_l = lin2 agg + b); x_{l+1} = x_l + lin a2_l + b
//   h3 = W1 x_L + b1; a3 = act(h3); e_i = (w2 . a3 + b2) std
// backward: d3 = std w2 act'(h3); g_x = d3 W1; per layer (reverse): g_m = (g_x lin_w) act'(m2_l); g_agg = g_m lin2_w;
//   conv adjoint -> g_h, gpos; g_x += g_h lin1_w;  then the neighbour-embedding adjoint and the force write-out.
namespace {

// agg[i,f] = sum_{e in row i, j != i} W(d_e)[f] * h[j,f]                          (CFConv.message / aggregate, "add")
__global__ void __launch_bounds__(1024)
gn_conv_fwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom, Tab tb,
            const float* __restrict__ h, int N, int F, float* __restrict__ agg) {
    pdl_sync();
    extern __shared__ float smem[];
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int f = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    float acc[1] = {0.f};
    if (valid) {
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e = r0 + sl; e < r1; e += S) {
            const int j = col[e];
            if (j == i) continue;
            float w[1], dw[1];
            tab_eval<1>(tb, geom[e].w, F, f, w, dw);
            acc[0] += w[0] * h[(size_t)j * F + f];
        }
    }
    slice_reduce<1>(acc, s_part, F);
    if (valid && sl == 0) agg[(size_t)i * F + f] = acc[0];
}

// adjoint, in row i of the symmetric pair set:  g_h[i,f] = sum_e W(d_e)[f] g_agg[j,f]
//   dE/dpos_i += u_e * sum_f W'(d_e)[f] (g_agg[i,f] h[j,f] + g_agg[j,f] h[i,f])       (both directions of the pair)
__global__ void __launch_bounds__(1024)
gn_conv_bwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom, Tab tb,
            const float* __restrict__ h, const float* __restrict__ g_agg, int N, int F, float* __restrict__ g_h,
            float* __restrict__ gpos) {
    pdl_sync();
    extern __shared__ float smem[];
    float* s_red = smem;
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int f = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    float acc[1] = {0.f};
    float gp[3] = {0.f, 0.f, 0.f};
    if (valid) {
        const float hi_ = h[(size_t)i * F + f], gi = g_agg[(size_t)i * F + f];
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e = r0 + sl; e < r1; e += S) {
            const int j = col[e];
            if (j == i) continue;
            const float4 g = geom[e];
            float w[1], dw[1];
            tab_eval<1>(tb, g.w, F, f, w, dw);
            const float gj = g_agg[(size_t)j * F + f];
            acc[0] += w[0] * gj;
            const float gd = dw[0] * (gi * h[(size_t)j * F + f] + gj * hi_);
            gp[0] += g.x * gd; gp[1] += g.y * gd; gp[2] += g.z * gd;
        }
    }
    slice_reduce<1>(acc, s_part, F);
    if (valid && sl == 0) g_h[(size_t)i * F + f] = acc[0];
    atom_reduce<3>(gp, s_red);
    if (valid && f == 0 && sl == 0) {
        gpos[3 * (size_t)i] += gp[0];
        gpos[3 * (size_t)i + 1] += gp[1];
        gpos[3 * (size_t)i + 2] += gp[2];
    }
}

struct GnWs {
    int32_t *num_found, *row_ptr, *col;
    float* geom;
    int* z32;
    void* graph_ws;
    size_t graph_ws_bytes;
    float *cat0, *X[TMDN_MAX_LAYERS + 1], *Hh[TMDN_MAX_LAYERS], *M2[TMDN_MAX_LAYERS], *A2[TMDN_MAX_LAYERS], *agg, *h3, *a3,
        *e_atom, *d3, *g_x, *g_m, *g_agg, *g_h, *g_xn, *gpos;
};

int carve_gn(Arena& a, const tmdn_gn_model* m, int64_t N, int64_t P, GnWs& w, std::vector<EtEntry>* names,
             const double* box, int strategy) {
    const size_t H = m->hidden, F = m->num_filters, L = m->num_layers, n = (size_t)N, H2 = H / 2;
    char buf[32];
    w.num_found = ecarve<int32_t>(a, names, "num_found", 4);
    w.row_ptr = ecarve<int32_t>(a, names, "row_ptr", n + 1);
    w.col = ecarve<int32_t>(a, names, "col", (size_t)P);
    w.geom = ecarve<float>(a, names, "geom", 4 * (size_t)P);
    w.z32 = ecarve<int>(a, names, "z32", n);
    w.graph_ws_bytes = tmdn_csr_build_workspace_bytes(strategy, N, box, m->cutoff_upper);
    if (w.graph_ws_bytes == 0) {
        tmdn_set_error("gn: bad box for the cell strategy");
        return TMDN_EINVAL;
    }
    w.graph_ws = ecarve<char>(a, names, "graph_ws", w.graph_ws_bytes);
    w.cat0 = ecarve<float>(a, names, "cat0", n * 2 * H);
    for (size_t l = 0; l <= L; ++l) { snprintf(buf, sizeof buf, "x%zu", l); w.X[l] = ecarve<float>(a, names, buf, n * H); }
    for (size_t l = 0; l < L; ++l) {
        snprintf(buf, sizeof buf, "h%zu", l); w.Hh[l] = ecarve<float>(a, names, buf, n * F);
        snprintf(buf, sizeof buf, "m2_%zu", l); w.M2[l] = ecarve<float>(a, names, buf, n * H);
        snprintf(buf, sizeof buf, "a2_%zu", l); w.A2[l] = ecarve<float>(a, names, buf, n * H);
    }
    w.agg = ecarve<float>(a, names, "agg", n * F);
    w.h3 = ecarve<float>(a, names, "h3", n * H2);
    w.a3 = ecarve<float>(a, names, "a3", n * H2);
    w.e_atom = ecarve<float>(a, names, "e_atom", n);
    w.d3 = ecarve<float>(a, names, "d3", n * H2);
    w.g_x = ecarve<float>(a, names, "g_x", n * H);
    w.g_m = ecarve<float>(a, names, "g_m", n * H);
    w.g_agg = ecarve<float>(a, names, "g_agg", n * F);
    w.g_h = ecarve<float>(a, names, "g_h", n * F);
    w.g_xn = ecarve<float>(a, names, "g_xn", n * H);
    w.gpos = ecarve<float>(a, names, "gpos", n * 3);
    return TMDN_OK;
}

int check_gn(const tmdn_gn_model* m) {
    TMDN_CHECK_ARG(m, "gn: null model");
    TMDN_CHECK_ARG(m->hidden >= 32 && m->hidden <= 1024 && m->hidden % 32 == 0,
                   "gn: hidden_channels must be a multiple of 32 in [32,1024] (got %d)", m->hidden);
    TMDN_CHECK_ARG(m->num_filters >= 32 && m->num_filters <= 1024 && m->num_filters % 32 == 0,
                   "gn: num_filters must be a multiple of 32 in [32,1024] (got %d)", m->num_filters);
    TMDN_CHECK_ARG(m->num_layers >= 0 && m->num_layers <= TMDN_MAX_LAYERS, "gn: num_layers out of range");
    TMDN_CHECK_ARG(m->max_z > 0 && m->n_int > 0, "gn: bad max_z / table size");
    TMDN_CHECK_ARG(m->cutoff_upper > m->cutoff_lower && m->cutoff_lower >= 0, "gn: bad cutoffs");
    TMDN_CHECK_ARG(m->emb && m->head_w1 && m->head_w2 && m->head_w1T, "gn: missing weights");
    for (int l = 0; l < m->num_layers; ++l)
        TMDN_CHECK_ARG(m->layer[l].table && m->layer[l].lin1_w && m->layer[l].lin2_w && m->layer[l].lin_w,
                       "gn: missing weights of layer %d", l);
    if (m->has_neighbor_embedding) TMDN_CHECK_ARG(m->ne_emb && m->ne_table && m->ne_comb_w, "gn: missing NE weights");
    return TMDN_OK;
}
}  // namespace

extern "C" size_t tmdn_gn_workspace_bytes(const tmdn_gn_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                          int strategy, const double box[9]) {
    (void)n_mol;
    if (check_gn(m) != TMDN_OK || n_atoms <= 0 || max_edges <= 0) return 0;
    Arena a(nullptr, 0);
    GnWs w;
    if (carve_gn(a, m, n_atoms, max_edges, w, nullptr, box, strategy) != TMDN_OK) return 0;
    return a.off + 256;
}

extern "C" int tmdn_gn_workspace_lookup(const tmdn_gn_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                        int strategy, const double box[9], const char* name, size_t* offset,
                                        size_t* bytes) {
    (void)n_mol;
    TMDN_TRY(check_gn(m));
    TMDN_CHECK_ARG(name && offset && bytes, "lookup: null argument");
    Arena a(nullptr, 0);
    GnWs w;
    std::vector<EtEntry> names;
    TMDN_TRY(carve_gn(a, m, n_atoms, max_edges, w, &names, box, strategy));
    for (const auto& e : names)
        if (e.name == name) {
            *offset = e.off;
            *bytes = e.bytes;
            return TMDN_OK;
        }
    tmdn_set_error("lookup: unknown buffer '%s'", name);
    return TMDN_EINVAL;
}

extern "C" int tmdn_gn_energy_forces(const tmdn_gn_model* m, const int64_t* z, const float* pos, const int64_t* batch,
                                     int64_t n_atoms, int64_t n_mol, const double box[9], int use_periodic,
                                     int strategy, int64_t max_edges, float* energy, float* forces, int32_t* status,
                                     void* workspace, size_t workspace_bytes, tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_TRY(check_gn(m));
    TMDN_CHECK_ARG(z && pos && batch && energy && status && workspace, "gn: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_mol > 0 && max_edges > 0, "gn: bad sizes");
    TMDN_CHECK_ARG(max_edges < (1LL << 31) && n_atoms < (1LL << 31) / 16, "gn: sizes exceed int32 indexing");
    TMDN_CHECK_ARG(strategy != TMDN_CELL || box, "gn: the cell strategy needs a box");
    const int N = (int)n_atoms, H = m->hidden, F = m->num_filters, L = m->num_layers, act = m->activation, H2 = H / 2;
    const bool tc = m->gemm_mode == TMDN_GEMM_TF32X3;
    const int64_t P = max_edges;
    const float lo = m->cutoff_lower, hi = m->cutoff_upper;
    Arena arena(workspace, workspace_bytes);
    GnWs w;
    TMDN_TRY(carve_gn(arena, m, n_atoms, P, w, nullptr, box, strategy));
    if (!arena.ok) {
        tmdn_set_error("gn: workspace too small (%zu < %zu)", workspace_bytes, arena.off);
        return TMDN_EWORKSPACE;
    }
    {
        KTimer ktimer__("neighbors_csr", stream);
        TMDN_TRY(tmdn_csr_build_f32(strategy, pos, batch, n_atoms, box, use_periodic, lo, hi, P, w.row_ptr, w.col, w.geom,
                                    w.num_found, w.graph_ws, w.graph_ws_bytes, stream_));
    }
    TMDN_CUDA(cudaMemsetAsync(status, 0, 4 * sizeof(int32_t), stream));
    ET_K("et_prep", launch_k(et_prep_kernel, (unsigned)cdiv(N, 256), 256, 0, stream, z, w.z32, N, m->max_z, w.num_found,
                             (long long)P, status, batch));
    const float4* geom = (const float4*)w.geom;
    const TabGrid tgrid = make_tab_grid(m->n_int, hi, m->grid_alpha, m->grid_gamma);
    const Tab ne_tab{(const float4*)m->ne_table, tgrid};
    const EdgeLaunch e1 = edge_launch(N, H, 1024, 1), ef = edge_launch(N, F, 1024, 1), e8 = edge_launch(N, H, 512, 8);
    const unsigned hgrid = (unsigned)cdiv((long long)N * H2, 256);

    // ---- embedding (+ neighbour embedding)
    if (m->has_neighbor_embedding) {
        ET_K("et_embed_fwd", launch_k(et_embed_fwd<true>, e1.grid, e1.block, e1.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      m->emb, m->ne_emb, ne_tab, N, H, w.cat0, 2 * H));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.cat0, 2 * H, m->ne_comb_w, m->ne_comb_b, w.X[0], H, N, H, 2 * H), stream, tc));
    } else {
        ET_K("et_embed_fwd", launch_k(et_embed_fwd<false>, e1.grid, e1.block, e1.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      m->emb, nullptr, ne_tab, N, H, w.X[0], H));
    }
    // ---- interaction blocks
    for (int l = 0; l < L; ++l) {
        const tmdn_gn_layer& ly = m->layer[l];
        const Tab tb{(const float4*)ly.table, tgrid};
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.X[l], H, ly.lin1_w, nullptr, w.Hh[l], F, N, F, H), stream, tc));
        ET_K("gn_conv_fwd", launch_k(gn_conv_fwd, ef.grid, ef.block, ef.smem, stream, w.row_ptr, w.col, geom, tb, w.Hh[l], N, F,
                                     w.agg));
        {
            GemmP p = gemm_desc(w.agg, F, ly.lin2_w, ly.lin2_b, w.A2[l], H, N, H, F);
            p.C2 = w.M2[l]; p.act = act;
            TMDN_TRY(node_gemm1(EPI_ACT, p, stream, tc));
        }
        TMDN_CUDA(cudaMemcpyAsync(w.X[l + 1], w.X[l], sizeof(float) * (size_t)N * H, cudaMemcpyDeviceToDevice, stream));
        {
            GemmP p = gemm_desc(w.A2[l], H, ly.lin_w, ly.lin_b, w.X[l + 1], H, N, H, H);
            p.accumulate = 1;
            TMDN_TRY(node_gemm1(EPI_NONE, p, stream, tc));
        }
    }
    // ---- Scalar head (output_modules.py:49-77) and molecular sums
    {
        GemmP p = gemm_desc(w.X[L], H, m->head_w1, m->head_b1, w.a3, H2, N, H2, H);
        p.C2 = w.h3; p.act = act;
        TMDN_TRY(node_gemm1(EPI_ACT, p, stream, tc));
    }
    ET_K("et_head_energy", launch_k(et_head_energy, (unsigned)cdiv(N, 4), 128, 0, stream, w.a3, m->head_w2, m->head_b2, m->std,
                                    w.e_atom, N, H2));
    ET_K("et_mol_reduce", launch_k(et_mol_reduce, (unsigned)n_mol, 128, 0, stream, w.e_atom, batch, N, m->mean, energy, status));
    if (!forces) return TMDN_OK;

    // ---- backward
    TMDN_CUDA(cudaMemsetAsync(w.gpos, 0, sizeof(float) * 3 * (size_t)N, stream));
    ET_K("et_head_seed_bwd", launch_k(et_head_seed_bwd, hgrid, 256, 0, stream, w.h3, m->head_w2, m->std, act, w.d3,
                                      (long long)N * H2, H2));
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.d3, H2, m->head_w1T, nullptr, w.g_x, H, N, H, H2), stream, tc));
    for (int l = L - 1; l >= 0; --l) {
        const tmdn_gn_layer& ly = m->layer[l];
        const Tab tb{(const float4*)ly.table, tgrid};
        {
            GemmP p = gemm_desc(w.g_x, H, ly.lin_wT, nullptr, w.g_m, H, N, H, H);
            p.aux = w.M2[l]; p.ldaux = H; p.act = act;
            TMDN_TRY(node_gemm1(EPI_MULGRAD, p, stream, tc));
        }
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_m, H, ly.lin2_wT, nullptr, w.g_agg, F, N, F, H), stream, tc));
        ET_K("gn_conv_bwd", launch_k(gn_conv_bwd, ef.grid, ef.block, ef.smem, stream, w.row_ptr, w.col, geom, tb, w.Hh[l],
                                     w.g_agg, N, F, w.g_h, w.gpos));
        {
            GemmP p = gemm_desc(w.g_h, F, ly.lin1_wT, nullptr, w.g_x, H, N, H, F);
            p.accumulate = 1;
            TMDN_TRY(node_gemm1(EPI_NONE, p, stream, tc));
        }
    }
    if (m->has_neighbor_embedding) {
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_x, H, m->ne_comb_w2T, nullptr, w.g_xn, H, N, H, H), stream, tc));
        ET_K("et_final_bwd", launch_k(et_final_bwd<true>, e8.grid, e8.block, e8.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      m->ne_emb, ne_tab, w.g_xn, w.gpos, forces, N, H, status));
    } else {
        ET_K("et_final_bwd", launch_k(et_final_bwd<false>, e8.grid, e8.block, e8.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      nullptr, ne_tab, nullptr, w.gpos, forces, N, H, status));
    }
    return TMDN_OK;
}

// ====================================================================================================================
// Transformer (TorchMD_T): torchmdnet/models/torchmd_t.py:144-300 -- the attention of the equivariant transformer
// without its vector channel.  Per layer: xl = LN(x); q|k|v = proj(xl); attn_e = attn_act(sum_head q_i k_j dk(d_e)) C(d_e);
// out_i = sum_e v_j dv(d_e) attn_e (self pairs included, loop = True); x += o_proj(out).  Then out_norm and the Scalar head.
// Filter table per layer: [n_int][2][H] float4 = act(dk_proj(rbf)), act(dv_proj(rbf)) * C(d)  (a missing filter is 1 / C).
namespace {

template <int HDC = 0>
__global__ void __launch_bounds__(1024)
t_edge_fwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom, Tab tb,
           int attn_act, int hd, const float* __restrict__ qkv, int N, int H, float* __restrict__ out) {
    pdl_sync();
    extern __shared__ float smem[];
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    const int H3 = 3 * H;
    float acc[1] = {0.f};
    if (valid) {
        const float qi = qkv[(size_t)i * H3 + h];
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e = r0 + sl; e < r1; e += S) {
            const int j = col[e];
            float f[2], df[2];
            tab_eval<2>(tb, geom[e].w, H, h, f, df);
            const float* pj = qkv + (size_t)j * H3 + h;
            const float s = seg_sum_t<HDC>(qi * pj[H] * f[0], hd);
            acc[0] += pj[2 * H] * f[1] * act_apply(attn_act, s);
        }
    }
    slice_reduce<1>(acc, s_part, H);
    if (valid && sl == 0) out[(size_t)i * H + h] = acc[0];
}

// adjoint in row i of the symmetric pair set, both directions of every pair (A: i <- j, B: j <- i):
//   g_q_i = sum gsA k_j dk;  g_k_i = sum gsB q_j dk;  g_v_i = sum g_out_j attnB dv;  dE/dpos_i += u (gdA + gdB)
template <int HDC = 0>
__global__ void __launch_bounds__(512)
t_edge_bwd(const int* __restrict__ row_ptr, const int* __restrict__ col, const float4* __restrict__ geom, Tab tb,
           int attn_act, int hd, const float* __restrict__ qkv, const float* __restrict__ g_out, int N, int H,
           float* __restrict__ g_qkv, float* __restrict__ gpos) {
    pdl_sync();
    extern __shared__ float smem[];
    float* s_red = smem;
    float* s_part = smem + EDGE_RED_FLOATS * blockDim.z;
    const int h = threadIdx.x, sl = threadIdx.y, S = blockDim.y;
    const int i = blockIdx.x * blockDim.z + threadIdx.z;
    const bool valid = i < N;
    const int H3 = 3 * H;
    float acc[3] = {0.f, 0.f, 0.f};   // g_q, g_k, g_v
    float gp[3] = {0.f, 0.f, 0.f};
    if (valid) {
        const float* pi = qkv + (size_t)i * H3 + h;
        const float qi = pi[0], ki = pi[H], vi = pi[2 * H];
        const float gi = g_out[(size_t)i * H + h];
        const int r0 = row_ptr[i], r1 = row_ptr[i + 1];
        for (int e = r0 + sl; e < r1; e += S) {
            const int j = col[e];
            const float4 g = geom[e];
            float f[2], df[2];
            tab_eval<2>(tb, g.w, H, h, f, df);
            const float* pj = qkv + (size_t)j * H3 + h;
            const float qj = pj[0], kj = pj[H], vj = pj[2 * H];
            const float gj = g_out[(size_t)j * H + h];
            float attnA, dactA, attnB, dactB;
            act_both(attn_act, seg_sum_t<HDC>(qi * kj * f[0], hd), attnA, dactA);
            const float gsA = seg_sum_t<HDC>(gi * vj * f[1], hd) * dactA;
            acc[0] += gsA * kj * f[0];
            act_both(attn_act, seg_sum_t<HDC>(qj * ki * f[0], hd), attnB, dactB);
            const float gsB = seg_sum_t<HDC>(gj * vi * f[1], hd) * dactB;
            acc[1] += gsB * qj * f[0];
            acc[2] += gj * attnB * f[1];
            if (j != i) {   // the self pair has d = 0 and no position dependence
                const float gd = gsA * qi * kj * df[0] + gi * attnA * vj * df[1] + gsB * qj * ki * df[0] + gj * attnB * vi * df[1];
                gp[0] += g.x * gd; gp[1] += g.y * gd; gp[2] += g.z * gd;
            }
        }
    }
    slice_reduce<3>(acc, s_part, H);
    if (valid && sl == 0) {
        float* o = g_qkv + (size_t)i * H3 + h;
        o[0] = acc[0]; o[H] = acc[1]; o[2 * H] = acc[2];
    }
    atom_reduce<3>(gp, s_red);
    if (valid && h == 0 && sl == 0) {
        gpos[3 * (size_t)i] += gp[0];
        gpos[3 * (size_t)i + 1] += gp[1];
        gpos[3 * (size_t)i + 2] += gp[2];
    }
}

struct TWs {
    int32_t *num_found, *row_ptr, *col;
    float* geom;
    int* z32;
    void* graph_ws;
    size_t graph_ws_bytes;
    float *cat0, *X[TMDN_MAX_LAYERS + 1], *stat[TMDN_MAX_LAYERS + 1], *QKV[TMDN_MAX_LAYERS], *xl, *out, *xo, *h3, *a3, *e_atom,
        *d3, *g_xo, *g_x, *g_out, *g_qkv, *g_xl, *g_xn, *gpos;
};

int carve_t(Arena& a, const tmdn_t_model* m, int64_t N, int64_t P, TWs& w, std::vector<EtEntry>* names, const double* box,
            int strategy) {
    const size_t H = m->hidden, L = m->num_layers, n = (size_t)N, H2 = H / 2;
    char buf[32];
    w.num_found = ecarve<int32_t>(a, names, "num_found", 4);
    w.row_ptr = ecarve<int32_t>(a, names, "row_ptr", n + 1);
    w.col = ecarve<int32_t>(a, names, "col", (size_t)P);
    w.geom = ecarve<float>(a, names, "geom", 4 * (size_t)P);
    w.z32 = ecarve<int>(a, names, "z32", n);
    w.graph_ws_bytes = tmdn_csr_build_workspace_bytes(strategy, N, box, m->cutoff_upper);
    if (w.graph_ws_bytes == 0) {
        tmdn_set_error("t: bad box for the cell strategy");
        return TMDN_EINVAL;
    }
    w.graph_ws = ecarve<char>(a, names, "graph_ws", w.graph_ws_bytes);
    w.cat0 = ecarve<float>(a, names, "cat0", n * 2 * H);
    for (size_t l = 0; l <= L; ++l) {
        snprintf(buf, sizeof buf, "x%zu", l); w.X[l] = ecarve<float>(a, names, buf, n * H);
        snprintf(buf, sizeof buf, "stat%zu", l); w.stat[l] = ecarve<float>(a, names, buf, n * 2);
    }
    for (size_t l = 0; l < L; ++l) { snprintf(buf, sizeof buf, "qkv%zu", l); w.QKV[l] = ecarve<float>(a, names, buf, n * 3 * H); }
    w.xl = ecarve<float>(a, names, "xl", n * H);
    w.out = ecarve<float>(a, names, "out", n * H);
    w.xo = ecarve<float>(a, names, "xo", n * H);
    w.h3 = ecarve<float>(a, names, "h3", n * H2);
    w.a3 = ecarve<float>(a, names, "a3", n * H2);
    w.e_atom = ecarve<float>(a, names, "e_atom", n);
    w.d3 = ecarve<float>(a, names, "d3", n * H2);
    w.g_xo = ecarve<float>(a, names, "g_xo", n * H);
    w.g_x = ecarve<float>(a, names, "g_x", n * H);
    w.g_out = ecarve<float>(a, names, "g_out", n * H);
    w.g_qkv = ecarve<float>(a, names, "g_qkv", n * 3 * H);
    w.g_xl = ecarve<float>(a, names, "g_xl", n * H);
    w.g_xn = ecarve<float>(a, names, "g_xn", n * H);
    w.gpos = ecarve<float>(a, names, "gpos", n * 3);
    return TMDN_OK;
}

int check_t(const tmdn_t_model* m) {
    TMDN_CHECK_ARG(m, "t: null model");
    TMDN_CHECK_ARG(m->hidden >= 32 && m->hidden <= 1024 && m->hidden % 32 == 0,
                   "t: hidden_channels must be a multiple of 32 in [32,1024] (got %d)", m->hidden);
    TMDN_CHECK_ARG(m->num_layers >= 0 && m->num_layers <= TMDN_MAX_LAYERS, "t: num_layers out of range");
    TMDN_CHECK_ARG(m->num_heads > 0 && m->hidden % m->num_heads == 0, "t: hidden must be divisible by num_heads");
    const int hd = m->hidden / m->num_heads;
    TMDN_CHECK_ARG(hd <= 32 && (hd & (hd - 1)) == 0, "t: head_dim must be a power of two <= 32 (got %d)", hd);
    TMDN_CHECK_ARG(m->max_z > 0 && m->n_int > 0, "t: bad max_z / table size");
    TMDN_CHECK_ARG(m->cutoff_upper > m->cutoff_lower && m->cutoff_lower >= 0, "t: bad cutoffs");
    TMDN_CHECK_ARG(m->emb && m->out_ln_w && m->out_ln_b && m->head_w1 && m->head_w2 && m->head_w1T, "t: missing weights");
    for (int l = 0; l < m->num_layers; ++l)
        TMDN_CHECK_ARG(m->layer[l].table && m->layer[l].qkv_w && m->layer[l].o_w, "t: missing weights of layer %d", l);
    if (m->has_neighbor_embedding) TMDN_CHECK_ARG(m->ne_emb && m->ne_table && m->ne_comb_w, "t: missing NE weights");
    return TMDN_OK;
}
}  // namespace

extern "C" size_t tmdn_t_workspace_bytes(const tmdn_t_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                         int strategy, const double box[9]) {
    (void)n_mol;
    if (check_t(m) != TMDN_OK || n_atoms <= 0 || max_edges <= 0) return 0;
    Arena a(nullptr, 0);
    TWs w;
    if (carve_t(a, m, n_atoms, max_edges, w, nullptr, box, strategy) != TMDN_OK) return 0;
    return a.off + 256;
}

extern "C" int tmdn_t_energy_forces(const tmdn_t_model* m, const int64_t* z, const float* pos, const int64_t* batch,
                                    int64_t n_atoms, int64_t n_mol, const double box[9], int use_periodic, int strategy,
                                    int64_t max_edges, float* energy, float* forces, int32_t* status, void* workspace,
                                    size_t workspace_bytes, tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_TRY(check_t(m));
    TMDN_CHECK_ARG(z && pos && batch && energy && status && workspace, "t: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_mol > 0 && max_edges > 0, "t: bad sizes");
    TMDN_CHECK_ARG(max_edges < (1LL << 31) && n_atoms < (1LL << 31) / 16, "t: sizes exceed int32 indexing");
    TMDN_CHECK_ARG(strategy != TMDN_CELL || box, "t: the cell strategy needs a box");
    const int N = (int)n_atoms, H = m->hidden, L = m->num_layers, act = m->activation, H2 = H / 2;
    const bool tc = m->gemm_mode == TMDN_GEMM_TF32X3;
    const int hd = H / m->num_heads;
    const int64_t P = max_edges;
    const float lo = m->cutoff_lower, hi = m->cutoff_upper;
    Arena arena(workspace, workspace_bytes);
    TWs w;
    TMDN_TRY(carve_t(arena, m, n_atoms, P, w, nullptr, box, strategy));
    if (!arena.ok) {
        tmdn_set_error("t: workspace too small (%zu < %zu)", workspace_bytes, arena.off);
        return TMDN_EWORKSPACE;
    }
    {
        KTimer ktimer__("neighbors_csr", stream);
        TMDN_TRY(tmdn_csr_build_f32(strategy, pos, batch, n_atoms, box, use_periodic, lo, hi, P, w.row_ptr, w.col, w.geom,
                                    w.num_found, w.graph_ws, w.graph_ws_bytes, stream_));
    }
    TMDN_CUDA(cudaMemsetAsync(status, 0, 4 * sizeof(int32_t), stream));
    ET_K("et_prep", launch_k(et_prep_kernel, (unsigned)cdiv(N, 256), 256, 0, stream, z, w.z32, N, m->max_z, w.num_found,
                             (long long)P, status, batch));
    const float4* geom = (const float4*)w.geom;
    const TabGrid tgrid = make_tab_grid(m->n_int, hi, m->grid_alpha, m->grid_gamma);
    const Tab ne_tab{(const float4*)m->ne_table, tgrid};
    const EdgeLaunch e1 = edge_launch(N, H, 1024, 1), e3 = edge_launch(N, H, 512, 3), e8 = edge_launch(N, H, 512, 8);
    const int apb = H >= 128 ? 1 : 128 / H;
    const dim3 lblock(H, 1, apb);
    const unsigned lgrid = (unsigned)cdiv(N, apb);
    const unsigned hgrid = (unsigned)cdiv((long long)N * H2, 256);

    if (m->has_neighbor_embedding) {
        ET_K("et_embed_fwd", launch_k(et_embed_fwd<true>, e1.grid, e1.block, e1.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      m->emb, m->ne_emb, ne_tab, N, H, w.cat0, 2 * H));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.cat0, 2 * H, m->ne_comb_w, m->ne_comb_b, w.X[0], H, N, H, 2 * H), stream, tc));
    } else {
        ET_K("et_embed_fwd", launch_k(et_embed_fwd<false>, e1.grid, e1.block, e1.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      m->emb, nullptr, ne_tab, N, H, w.X[0], H));
    }
    for (int l = 0; l < L; ++l) {
        const tmdn_t_layer& ly = m->layer[l];
        const Tab tb{(const float4*)ly.table, tgrid};
        ET_K("et_ln_fwd", launch_k(et_ln_fwd, lgrid, lblock, 0, stream, w.X[l], ly.ln_w, ly.ln_b, w.xl, H, w.stat[l], N, H));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.xl, H, ly.qkv_w, ly.qkv_b, w.QKV[l], 3 * H, N, 3 * H, H), stream, tc));
        if (hd == 16)
            ET_K("t_edge_fwd", launch_k(t_edge_fwd<16>, e1.grid, e1.block, e1.smem, stream, w.row_ptr, w.col, geom, tb,
                                        m->attn_activation, hd, w.QKV[l], N, H, w.out));
        else
            ET_K("t_edge_fwd", launch_k(t_edge_fwd<0>, e1.grid, e1.block, e1.smem, stream, w.row_ptr, w.col, geom, tb,
                                        m->attn_activation, hd, w.QKV[l], N, H, w.out));
        TMDN_CUDA(cudaMemcpyAsync(w.X[l + 1], w.X[l], sizeof(float) * (size_t)N * H, cudaMemcpyDeviceToDevice, stream));
        {
            GemmP p = gemm_desc(w.out, H, ly.o_w, ly.o_b, w.X[l + 1], H, N, H, H);
            p.accumulate = 1;
            TMDN_TRY(node_gemm1(EPI_NONE, p, stream, tc));
        }
    }
    ET_K("et_ln_fwd", launch_k(et_ln_fwd, lgrid, lblock, 0, stream, w.X[L], m->out_ln_w, m->out_ln_b, w.xo, H, w.stat[L], N, H));
    {
        GemmP p = gemm_desc(w.xo, H, m->head_w1, m->head_b1, w.a3, H2, N, H2, H);
        p.C2 = w.h3; p.act = act;
        TMDN_TRY(node_gemm1(EPI_ACT, p, stream, tc));
    }
    ET_K("et_head_energy", launch_k(et_head_energy, (unsigned)cdiv(N, 4), 128, 0, stream, w.a3, m->head_w2, m->head_b2, m->std,
                                    w.e_atom, N, H2));
    ET_K("et_mol_reduce", launch_k(et_mol_reduce, (unsigned)n_mol, 128, 0, stream, w.e_atom, batch, N, m->mean, energy, status));
    if (!forces) return TMDN_OK;

    TMDN_CUDA(cudaMemsetAsync(w.gpos, 0, sizeof(float) * 3 * (size_t)N, stream));
    ET_K("et_head_seed_bwd", launch_k(et_head_seed_bwd, hgrid, 256, 0, stream, w.h3, m->head_w2, m->std, act, w.d3,
                                      (long long)N * H2, H2));
    TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.d3, H2, m->head_w1T, nullptr, w.g_xo, H, N, H, H2), stream, tc));
    ET_K("et_ln_bwd", launch_k(et_ln_bwd, lgrid, lblock, 0, stream, w.X[L], m->out_ln_w, w.stat[L], w.g_xo, H, w.g_x, 0, N, H));
    for (int l = L - 1; l >= 0; --l) {
        const tmdn_t_layer& ly = m->layer[l];
        const Tab tb{(const float4*)ly.table, tgrid};
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_x, H, ly.o_wT, nullptr, w.g_out, H, N, H, H), stream, tc));
        if (hd == 16)
            ET_K("t_edge_bwd", launch_k(t_edge_bwd<16>, e3.grid, e3.block, e3.smem, stream, w.row_ptr, w.col, geom, tb,
                                        m->attn_activation, hd, w.QKV[l], w.g_out, N, H, w.g_qkv, w.gpos));
        else
            ET_K("t_edge_bwd", launch_k(t_edge_bwd<0>, e3.grid, e3.block, e3.smem, stream, w.row_ptr, w.col, geom, tb,
                                        m->attn_activation, hd, w.QKV[l], w.g_out, N, H, w.g_qkv, w.gpos));
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_qkv, 3 * H, ly.qkv_wT, nullptr, w.g_xl, H, N, H, 3 * H), stream, tc));
        ET_K("et_ln_bwd", launch_k(et_ln_bwd, lgrid, lblock, 0, stream, w.X[l], ly.ln_w, w.stat[l], w.g_xl, H, w.g_x, 1, N, H));
    }
    if (m->has_neighbor_embedding) {
        TMDN_TRY(node_gemm1(EPI_NONE, gemm_desc(w.g_x, H, m->ne_comb_w2T, nullptr, w.g_xn, H, N, H, H), stream, tc));
        ET_K("et_final_bwd", launch_k(et_final_bwd<true>, e8.grid, e8.block, e8.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      m->ne_emb, ne_tab, w.g_xn, w.gpos, forces, N, H, status));
    } else {
        ET_K("et_final_bwd", launch_k(et_final_bwd<false>, e8.grid, e8.block, e8.smem, stream, w.row_ptr, w.col, geom, w.z32,
                                      nullptr, ne_tab, nullptr, w.gpos, forces, N, H, status));
    }
    return TMDN_OK;
}
