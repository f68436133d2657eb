// tn_chain.cuh -- fused node stages of TensorNet for small systems (included by tensornet.cu after its device helpers).
//
// Between two edge kernels the model only does row-local work: channel mixes, the scalar MLPs, LayerNorms and the
// per-(atom, channel) 3x3 algebra (tensornet.py:287-326, 376-410, 230-233; output_modules.py:49-77) and their
// adjoints.  Launched one kernel per operation that is 34 of the 45 launches of a TensorNet-SPICE step, each 3-11 us at
// 166 atoms although it moves a few hundred KB (ncu r01i).  Here a CTA owns `na` atoms (block = (H, na), a thread is one
// (atom, channel) and keeps its 9 irreducible components in registers) and walks a whole chain of those operations:
//   CH_EMBED    scalar MLP (H -> 2H -> 3H), embedding mix + scaling, normalise, first mix            (5 launches -> 1)
//   CH_LAYER    mix, X + dX + dX^2, [normalise, next mix]                                              (3 -> 1)
//   CH_HEAD     output LayerNorm, linear, head MLP, atom energy, and the whole adjoint back to dM, dY  (12 -> 1)
//   CH_BWD      accumulate mix^T, normalise^T, square^T, mix^T, product^T                              (5 -> 1)
//   CH_BWD_EMB  accumulate mix^T, normalise^T, scaling^T, MLP^T, mix^T, LayerNorm^T                    (7 -> 1)
// A GEMM inside a chain reads its A operand from shared memory (the CTA's own rows) and the weight in [K, N] layout
// (every weight exists in both layouts because the adjoint needs the transpose).  The weights of ALL the GEMMs of a
// chain form one stream of 16-row chunks (<= 24 KB: 16 consecutive k rows of a matrix, or of the three matrices of a mix)
// that every thread helps to copy with 16-byte cp.async into a 7-stage shared-memory ring, six chunks ahead of the
// arithmetic and across GEMM boundaries: the weights do not depend on the data.  The k range of a GEMM is split
// across the warps of the CTA and the partial sums are added in slice order, so results differ from the per-operation
// kernels by rounding only (tests: rtol 3e-6 on energies, 1e-5 eV/A on forces, and the reference's numbers).
//
// STATUS: opt-in (tmdn_set_tuning("fused_nodes", 1)).  Correct, 18 launches per step instead of 45, but not faster:
// one SM pulls a chain's 0.4-0.9 MB of cold weights at 21-42 GB/s, while the per-operation kernels spread every
// matrix over 44-312 CTAs.  Versions and measurements: profiles/r01i_node_chain_experiments.md.
#pragma once

enum { CH_EMBED = 0, CH_LAYER = 1, CH_HEAD = 2, CH_BWD = 3, CH_BWD_EMB = 4 };
constexpr int CHAIN_MAX_ROWS = 1024, CHAIN_MAX_H = 128;

constexpr int WS_STAGES = 7, WS_STAGE_FLOATS = 6144, WS_KC = 16, WS_MAX_SEGS = 8;
constexpr int CHAIN_NA = 2;   // atoms per CTA (the dense GEMMs are written for two)
// one GEMM's weight in the stream: ng matrices [K, N] (ng = 3: the I / A / S matrices of a channel mix)
struct WSeg { const float* p[3]; int K, N, ng; };

struct ChainArgs {
    int N, H, act, so3, flag;            // flag: CH_LAYER -> has next layer; CH_HEAD -> backward wanted
    float stdv;
    int nseg;
    int dbg;                             // measurement only: 1 = no weight copies, 2 = no GEMM arithmetic
    WSeg seg[WS_MAX_SEGS];               // weights in the order the chain consumes them
    const float *b1, *b2;
    const float *ln_w, *ln_b, *hw2, *hb2;
    // node buffers (see TnWs)
    const float *in0, *in1, *in2, *in3, *in4, *in5;
    float *out0, *out1, *out2, *out3, *out4, *out5, *out6, *out7;
    float *a3, *a4;                      // CH_HEAD: post-activations (a3 is the representation output)
};

// The weight stream.  Every thread of the CTA holds the same state and makes the same calls.  A chunk is WS_KC = 16
// consecutive k rows of a segment (of each of its ng matrices); CTA b walks the chunks of a segment starting at chunk
// b mod nchunks, so that at any moment the CTAs of the grid pull different lines out of L2 instead of all 83 asking
// for the same 24 KB (the partial sums of a GEMM are combined across warps anyway, see below).
struct WStream {
    const ChainArgs* a;
    float* ring;
    int seg, j;           // next chunk to issue: segment, position in the (rotated) chunk order
    int issued, consumed;
    int cseg, cj;         // chunk the arithmetic is at
    int nthreads, tid;

    __device__ __forceinline__ int chunk_k0(const WSeg& S, int jj) const {
        const int nch = S.K / WS_KC;
        return ((jj + (int)blockIdx.x) % nch) * WS_KC;
    }
    __device__ __forceinline__ void issue() {
        if (seg < a->nseg) {
            const WSeg& S = a->seg[seg];
            const int k0 = chunk_k0(S, j);
            float* dst = ring + (size_t)(issued % WS_STAGES) * WS_STAGE_FLOATS;
            const int n16 = WS_KC * S.N / 4;
            for (int g = 0; g < S.ng; ++g) {
                const float* src = S.p[g] + (size_t)k0 * S.N;
                const unsigned d0 = (unsigned)__cvta_generic_to_shared(dst + (size_t)g * WS_KC * S.N);
                for (int idx = tid; idx < n16; idx += nthreads)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 16u * idx), "l"(src + 4 * idx) : "memory");
            }
            if (++j >= S.K / WS_KC) { ++seg; j = 0; }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        ++issued;
    }
    __device__ __forceinline__ void start(const ChainArgs* args, float* r) {
        a = args; ring = r; seg = 0; j = 0; issued = 0; consumed = 0; cseg = 0; cj = 0;
        nthreads = blockDim.x * blockDim.y;
        tid = threadIdx.y * blockDim.x + threadIdx.x;
        for (int s = 0; s < WS_STAGES - 1; ++s) issue();
    }
    // the oldest chunk in flight has landed for every thread; the stage used before it is free again and refilled.
    // Returns the chunk and the k index of its first row.
    __device__ __forceinline__ const float* acquire(int& k0) {
        asm volatile("cp.async.wait_group %0;" ::"n"(WS_STAGES - 2) : "memory");
        __syncthreads();
        const float* w = ring + (size_t)(consumed % WS_STAGES) * WS_STAGE_FLOATS;
        const WSeg& S = a->seg[cseg];
        k0 = chunk_k0(S, cj);
        if (++cj >= S.K / WS_KC) { ++cseg; cj = 0; }
        issue();
        ++consumed;
        return w;
    }
};

// GEMMs of a chain.  For the arithmetic the CTA's 8 warps are re-mapped: a lane owns 4 consecutive output channels
// (one 16-byte shared-memory read per weight row), a warp owns a slice of every 16-row chunk, partial sums are
// combined through shared memory by the thread that owns (atom, channel) in the element-wise layout.  With a lane
// per output channel (first version) the broadcast reads of the A operand cost one shared-memory pass per FMA and a
// mix took 7 us; here a weight element is read once or twice per CTA and an A element once.
//
// out[c] = sum_k A[y, c, k] * Wt_g(c)[k, h]    (next segment: ng = 3, K = N = H <= 128).
// warp = (atom, quarter of the chunk's rows): 4 rows x 9 components x 4 channels per chunk.
__device__ __forceinline__ void chain_mix(const float x[9], WStream& ws, float* sA, float* sP, int H, float out[9]) {
    const int h = threadIdx.x, y = threadIdx.y;
    const int tid = y * blockDim.x + h, warp = tid >> 5, lane = tid & 31;
    const int nwa = blockDim.x >> 5;                 // warps per atom in the element-wise layout (H / 32)
    const int nq = (blockDim.x * blockDim.y) >> 6;   // k slices per atom = warps / 2
    const int wy = warp / nq, q = warp % nq;         // GEMM role of this warp: atom, slice
    const int rpw = WS_KC / nq;                      // rows of a chunk per warp (nq = 4: 4, nq = 2: 8, nq = 1: 16)
    const int n0 = 4 * lane;
    const bool lane_on = n0 < H;
    (void)nwa;
    float* row = sA + (size_t)y * 9 * H;
    __syncthreads();   // earlier readers of the panel and of the partial-sum area are done
#pragma unroll
    for (int c = 0; c < 9; ++c) row[c * H + h] = x[c];
    float acc[9][4];
#pragma unroll
    for (int c = 0; c < 9; ++c)
#pragma unroll
        for (int jn = 0; jn < 4; ++jn) acc[c][jn] = 0.f;
    const float* arow = sA + (size_t)wy * 9 * H;
    const int nch = H / WS_KC;
    for (int jc = 0; jc < nch; ++jc) {
        int k0;
        const float* W = ws.acquire(k0);   // its __syncthreads also publishes the panel
        if (ws.a->dbg == 2 || !lane_on) continue;
        for (int r = q * rpw; r < (q + 1) * rpw; r += 4) {
            float4 a4[9];
#pragma unroll
            for (int c = 0; c < 9; ++c) a4[c] = *reinterpret_cast<const float4*>(arow + c * H + k0 + r);
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const float4 wi = *reinterpret_cast<const float4*>(W + (size_t)(r + rr) * H + n0);
                const float4 wa = *reinterpret_cast<const float4*>(W + (size_t)(WS_KC + r + rr) * H + n0);
                const float4 wsv = *reinterpret_cast<const float4*>(W + (size_t)(2 * WS_KC + r + rr) * H + n0);
#pragma unroll
                for (int c = 0; c < 9; ++c) {
                    const float4 w = c == 0 ? wi : (c < 4 ? wa : wsv);
                    const float av = rr == 0 ? a4[c].x : (rr == 1 ? a4[c].y : (rr == 2 ? a4[c].z : a4[c].w));
                    acc[c][0] = fmaf(av, w.x, acc[c][0]);
                    acc[c][1] = fmaf(av, w.y, acc[c][1]);
                    acc[c][2] = fmaf(av, w.z, acc[c][2]);
                    acc[c][3] = fmaf(av, w.w, acc[c][3]);
                }
            }
        }
    }
    // partial sums [slice][atom][c][H] -> the owning thread adds the slices in order
    if (lane_on) {
#pragma unroll
        for (int c = 0; c < 9; ++c)
            *reinterpret_cast<float4*>(sP + ((size_t)(q * blockDim.y + wy) * 9 + c) * H + n0) =
                make_float4(acc[c][0], acc[c][1], acc[c][2], acc[c][3]);
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 9; ++c) {
        float sum = 0.f;
        for (int s = 0; s < nq; ++s) sum += sP[((size_t)(s * blockDim.y + y) * 9 + c) * H + h];
        out[c] = sum;
    }
}

// out[j] = sum_k v[y, k] * Wt[k, h + j*H]   for the outputs n = h + j*H < N   (v: shared memory [na, ldv]; next
// segment: ng = 1).  warp = slice of the chunk's rows, both atoms: 2 atoms x NB blocks of 128 channels x 4.
template <int NJ>
__device__ __forceinline__ void chain_dense(const float* sV, int ldv, int K, WStream& ws, float* sP, int N, int H,
                                            float out[NJ]) {
    constexpr int NB = 3;                            // N <= 3 * 128
    const int h = threadIdx.x, y = threadIdx.y;
    const int tid = y * blockDim.x + h, warp = tid >> 5, lane = tid & 31;
    const int nw = (blockDim.x * blockDim.y) >> 5;   // slices = warps (8 at H = 128)
    const int rpw = WS_KC / nw > 0 ? WS_KC / nw : 1; // rows of a chunk per warp
    const bool warp_on = warp * rpw < WS_KC;
    float acc[2][NB][4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int jn = 0; jn < 4; ++jn) acc[a][b][jn] = 0.f;
    const int nch = K / WS_KC;
    for (int jc = 0; jc < nch; ++jc) {
        int k0;
        const float* W = ws.acquire(k0);   // its __syncthreads also publishes v
        if (ws.a->dbg == 2 || !warp_on) continue;
        for (int r = warp * rpw; r < (warp + 1) * rpw; ++r) {
            const float a0 = sV[k0 + r], a1 = sV[ldv + k0 + r];
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const int n0 = b * 128 + 4 * lane;
                if (n0 < N) {
                    const float4 w = *reinterpret_cast<const float4*>(W + (size_t)r * N + n0);
                    acc[0][b][0] = fmaf(a0, w.x, acc[0][b][0]); acc[0][b][1] = fmaf(a0, w.y, acc[0][b][1]);
                    acc[0][b][2] = fmaf(a0, w.z, acc[0][b][2]); acc[0][b][3] = fmaf(a0, w.w, acc[0][b][3]);
                    acc[1][b][0] = fmaf(a1, w.x, acc[1][b][0]); acc[1][b][1] = fmaf(a1, w.y, acc[1][b][1]);
                    acc[1][b][2] = fmaf(a1, w.z, acc[1][b][2]); acc[1][b][3] = fmaf(a1, w.w, acc[1][b][3]);
                }
            }
        }
    }
    __syncthreads();   // earlier readers of the partial-sum area are done (acquire's barrier came before their reads)
    if (warp_on) {
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const int n0 = b * 128 + 4 * lane;
                if (n0 < N)
                    *reinterpret_cast<float4*>(sP + ((size_t)warp * 2 + a) * N + n0) =
                        make_float4(acc[a][b][0], acc[a][b][1], acc[a][b][2], acc[a][b][3]);
            }
    }
    __syncthreads();
    const int nsl = nw * rpw > WS_KC ? WS_KC / rpw : nw;   // slices that hold data
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
        const int n = h + j * H;
        float sum = 0.f;
        if (n < N)
            for (int s = 0; s < nsl; ++s) sum += sP[((size_t)s * 2 + y) * N + n];
        out[j] = sum;
    }
}

__device__ __forceinline__ void chain_scale3(float x[9], const float f[3]) {
    x[0] *= f[0];
#pragma unroll
    for (int c = 1; c < 4; ++c) x[c] *= f[1];
#pragma unroll
    for (int c = 4; c < 9; ++c) x[c] *= f[2];
}

// adjoint of X_{l+1} = Xh + dX + dX^2 followed by mix^T and the adjoint of the (M, Y) product:
// g = dE/dX_{l+1} -> dM (-> out_dM) and the direct part of dY (-> out_dY)        (tn_update_bwd, mix, tn_prod_bwd)
__device__ __forceinline__ void chain_pre_bwd(const ChainArgs& a, const float g[9], const float* dXl, const float* Y,
                                              const float* M, WStream& ws, float* sA, float* sP, size_t i, bool valid,
                                              float* out_dM, float* out_dY) {
    const int H = a.H, h = threadIdx.x;
    float d[9], dd[9], dc[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) d[c] = 0.f;
    if (valid) ld9(dXl, i, H, h, d);
    t3_square_bwd(d, g, dd);
#pragma unroll
    for (int c = 0; c < 9; ++c) dd[c] += g[c];
    chain_mix(dd, ws, sA, sP, H, dc);
    if (!valid) return;
    float y[9], m[9], tau[9], dtau[9], dm[9], dy[9];
    ld9(Y, i, H, h, y);
    ld9(M, i, H, h, m);
    t3_product(m, y, a.so3 != 0, tau);
    t3_normalize_bwd(tau, dc, dtau);
    t3_product_bwd(m, y, a.so3 != 0, dtau, dm, dy);
    st9(out_dM, i, H, h, dm);
    st9(out_dY, i, H, h, dy);
}

// dXa(total) = normalise^T( dXa + mix^T(dY) )                         (accumulating mix, tn_normalize_bwd)
__device__ __forceinline__ void chain_layer_in_bwd(const ChainArgs& a, const float* dY, const float* dXa_in,
                                                   const float* X, WStream& ws, float* sA, float* sP, size_t i,
                                                   bool valid, float g[9]) {
    const int H = a.H, h = threadIdx.x;
    float dy[9], t[9], r[9], x[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) dy[c] = r[c] = x[c] = 0.f;
    if (valid) ld9(dY, i, H, h, dy);
    chain_mix(dy, ws, sA, sP, H, t);
    if (valid) {
        ld9(dXa_in, i, H, h, r);
        ld9(X, i, H, h, x);
    }
#pragma unroll
    for (int c = 0; c < 9; ++c) t[c] += r[c];
    t3_normalize_bwd(x, t, g);
}

template <int CHAIN>
__global__ void __launch_bounds__(256) tn_chain_kernel(const ChainArgs a) {
    pdl_sync();
    extern __shared__ __align__(16) float ch_smem[];
    const int H = a.H, h = threadIdx.x, y = threadIdx.y, na = blockDim.y;
    float* ring = ch_smem;                                // weight stream, WS_STAGES x 24 KB
    float* sA = ring + WS_STAGES * WS_STAGE_FLOATS;       // [na, 9, H]
    float* sV0 = sA + (size_t)na * 9 * H;                 // [na, 3H]
    float* sV1 = sV0 + (size_t)na * 3 * H;                // [na, 3H]
    float* s_red = sV1 + (size_t)na * 3 * H;              // 64
    float* sP = s_red + 64;                               // partial sums of a GEMM: [slices][na][9][H]
    const int ldv = 3 * H;
    const int ii = blockIdx.x * na + y;
    const bool valid = ii < a.N;
    const size_t i = valid ? (size_t)ii : 0;
    float* v0 = sV0 + (size_t)y * ldv;
    float* v1 = sV1 + (size_t)y * ldv;
    WStream ws;
    ws.start(&a, ring);

    if constexpr (CHAIN == CH_EMBED) {
        // in0 = nh, in1 = ias; out0 = h1, out1 = a1, out2 = h2, out3 = fias, out4 = iasm, out5 = X0, out6 = Xh, out7 = Y0
        v0[h] = valid ? a.in0[i * H + h] : 0.f;
        __syncthreads();
        float o2[2], f[3];
        chain_dense<2>(sV0, ldv, H, ws, sP, 2 * H, H, o2);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int n = h + j * H;
            const float pre = o2[j] + a.b1[n];
            const float post = act_apply(a.act, pre);
            if (valid) { a.out0[i * 2 * H + n] = pre; a.out1[i * 2 * H + n] = post; }
            v1[n] = post;
        }
        __syncthreads();
        chain_dense<3>(sV1, ldv, 2 * H, ws, sP, 3 * H, H, f);
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const int n = h + g * H;
            const float pre = f[g] + a.b2[n];
            f[g] = act_apply(a.act, pre);
            if (valid) { a.out2[i * 3 * H + n] = pre; a.out3[i * 3 * H + n] = f[g]; }
        }
        float x[9], im[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) x[c] = 0.f;
        if (valid) ld9(a.in1, i, H, h, x);
        chain_mix(x, ws, sA, sP, H, im);
        if (valid) st9(a.out4, i, H, h, im);
        chain_scale3(im, f);
        if (valid) st9(a.out5, i, H, h, im);
        if (a.flag) {
            float xh[9], yv[9];
            t3_normalize(im, xh);
            if (valid) st9(a.out6, i, H, h, xh);
            chain_mix(xh, ws, sA, sP, H, yv);
            if (valid) st9(a.out7, i, H, h, yv);
        }
    } else if constexpr (CHAIN == CH_LAYER) {
        // in0 = tauh, in1 = Xh; out0 = dX[l], out1 = X[l+1], out2 = Xh (next), out3 = Y[l+1]
        float th[9], d[9], xh[9], sq[9], xn[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) th[c] = xh[c] = 0.f;
        if (valid) { ld9(a.in0, i, H, h, th); ld9(a.in1, i, H, h, xh); }
        chain_mix(th, ws, sA, sP, H, d);
        t3_square(d, sq);
#pragma unroll
        for (int c = 0; c < 9; ++c) xn[c] = xh[c] + d[c] + sq[c];
        if (valid) { st9(a.out0, i, H, h, d); st9(a.out1, i, H, h, xn); }
        if (a.flag) {
            float nx[9], yv[9];
            t3_normalize(xn, nx);
            if (valid) st9(a.out2, i, H, h, nx);
            chain_mix(nx, ws, sA, sP, H, yv);
            if (valid) st9(a.out3, i, H, h, yv);
        }
    } else if constexpr (CHAIN == CH_HEAD) {
        // in0 = X[L], in1 = dX[L-1], in2 = Y[L-1], in3 = M[L-1]
        // out0 = h3, out1 = h4, out2 = e_atom, out3 = dXa, out4 = dM (Dias buffer), out5 = dC, out6 = stat, out7 = xl
        const int Hh = H / 2;
        float x[9], q[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 9; ++c) x[c] = 0.f;
        if (valid) {
            ld9(a.in0, i, H, h, x);
            t3_qparts(x, q[0], q[1], q[2]);
        }
        float s[1] = {q[0] + q[1] + q[2]};
        row_reduce<1>(s, s_red);
        const float mean = s[0] / (3.0f * H);
        float v[1] = {(q[0] - mean) * (q[0] - mean) + (q[1] - mean) * (q[1] - mean) + (q[2] - mean) * (q[2] - mean)};
        row_reduce<1>(v, s_red);
        const float rstd = rsqrtf(v[0] / (3.0f * H) + 1e-5f);
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const float xl = (q[g] - mean) * rstd * a.ln_w[g * H + h] + a.ln_b[g * H + h];
            v0[g * H + h] = valid ? xl : 0.f;
            if (valid) a.out7[i * 3 * H + g * H + h] = xl;
        }
        if (valid && h == 0) { a.out6[2 * i] = mean; a.out6[2 * i + 1] = rstd; }
        __syncthreads();
        float o1[1];
        chain_dense<1>(sV0, ldv, 3 * H, ws, sP, H, H, o1);
        const float h3 = o1[0] + a.b1[h];
        const float a3 = act_apply(a.act, h3);
        if (valid) { a.out0[i * H + h] = h3; a.a3[i * H + h] = a3; }
        v1[h] = a3;
        __syncthreads();
        chain_dense<1>(sV1, ldv, H, ws, sP, Hh, H, o1);
        float h4 = 0.f;
        if (h < Hh) {
            h4 = o1[0] + a.b2[h];
            const float a4 = act_apply(a.act, h4);
            if (valid) { a.out1[i * Hh + h] = h4; a.a4[i * Hh + h] = a4; }
            v0[h] = a4;
        }
        __syncthreads();
        if (h < 32) {   // tn_head_energy's order: lane-strided partial sums, then the xor-shuffle tree
            float e = 0.f;
            for (int k = h; k < Hh; k += 32) e += v0[k] * a.hw2[k];
            e = warp_sum(e);
            if (valid && h == 0) a.out2[i] = (e + a.hb2[0]) * a.stdv;
        }
        if (!a.flag) return;
        __syncthreads();
        if (h < Hh) v1[h] = a.stdv * a.hw2[h] * act_grad(a.act, h4);
        __syncthreads();
        chain_dense<1>(sV1, ldv, Hh, ws, sP, H, H, o1);
        v0[h] = o1[0] * act_grad(a.act, h3);
        __syncthreads();
        float dxl[3];
        chain_dense<3>(sV0, ldv, H, ws, sP, 3 * H, H, dxl);
        // tn_out_norm_bwd
        float xhat[3], dg[3];
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            xhat[g] = valid ? (q[g] - mean) * rstd : 0.f;
            dg[g] = valid ? dxl[g] * a.ln_w[g * H + h] : 0.f;
        }
        float r[2] = {dg[0] + dg[1] + dg[2], dg[0] * xhat[0] + dg[1] * xhat[1] + dg[2] * xhat[2]};
        row_reduce<2>(r, s_red);
        const float m1 = r[0] / (3.0f * H), m2 = r[1] / (3.0f * H);
        float dq[3];
#pragma unroll
        for (int g = 0; g < 3; ++g) dq[g] = (dg[g] - m1 - xhat[g] * m2) * rstd;
        float gq[9], gx[9];
        t3_qgrad(x, gq);
        gx[0] = gq[0] * dq[0];
#pragma unroll
        for (int c = 1; c < 4; ++c) gx[c] = gq[c] * dq[1];
#pragma unroll
        for (int c = 4; c < 9; ++c) gx[c] = gq[c] * dq[2];
        if (valid) st9(a.out3, i, H, h, gx);
        chain_pre_bwd(a, gx, a.in1, a.in2, a.in3, ws, sA, sP, i, valid, a.out4, a.out5);
    } else if constexpr (CHAIN == CH_BWD) {
        // in0 = dC (total dY_l), in1 = dXa, in2 = X[l], in3 = dX[l-1], in4 = Y[l-1], in5 = M[l-1]
        // out0 = dXa, out1 = dM (Dias buffer), out2 = dC
        float g[9];
        chain_layer_in_bwd(a, a.in0, a.in1, a.in2, ws, sA, sP, i, valid, g);
        if (valid) st9(a.out0, i, H, h, g);
        chain_pre_bwd(a, g, a.in3, a.in4, a.in5, ws, sA, sP, i, valid, a.out1, a.out2);
    } else {
        // CH_BWD_EMB: in0 = dC (total dY_0), in1 = dXa, in2 = X[0], in3 = iasm, in4 = fias, in5 = ias
        // w1 = emb_w2 [3H, 2H], w2 = emb_w1 [2H, H]; b1 = h2 (pre-activations), b2 = h1
        // out0 = dXa (scaled), out1 = Dias, out2 = dh2, out3 = dh1, out4 = dnh
        float g[9];
        chain_layer_in_bwd(a, a.in0, a.in1, a.in2, ws, sA, sP, i, valid, g);
        float im[9], f[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 9; ++c) im[c] = 0.f;
        if (valid) {
            ld9(a.in3, i, H, h, im);
#pragma unroll
            for (int k = 0; k < 3; ++k) f[k] = a.in4[i * 3 * H + k * H + h];
        }
        float sg[3] = {g[0] * im[0], 0.f, 0.f};
#pragma unroll
        for (int c = 1; c < 4; ++c) sg[1] += g[c] * im[c];
#pragma unroll
        for (int c = 4; c < 9; ++c) sg[2] += g[c] * im[c];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float dh2 = valid ? sg[k] * act_grad(a.act, a.b1[i * 3 * H + k * H + h]) : 0.f;
            v0[k * H + h] = dh2;
            if (valid) a.out2[i * 3 * H + k * H + h] = dh2;
        }
        chain_scale3(g, f);
        if (valid) st9(a.out0, i, H, h, g);
        __syncthreads();
        float o2[2], o1[1];
        chain_dense<2>(sV0, ldv, 3 * H, ws, sP, 2 * H, H, o2);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int n = h + j * H;
            const float dh1 = valid ? o2[j] * act_grad(a.act, a.b2[i * 2 * H + n]) : 0.f;
            v1[n] = dh1;
            if (valid) a.out3[i * 2 * H + n] = dh1;
        }
        __syncthreads();
        chain_dense<1>(sV1, ldv, 2 * H, ws, sP, H, H, o1);
        const float dnh = o1[0];
        if (valid) a.out4[i * H + h] = dnh;
        float db[9];
        chain_mix(g, ws, sA, sP, H, db);
        // tn_embed_norm_bwd
        float x[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) x[c] = 0.f;
        if (valid) ld9(a.in5, i, H, h, x);
        const float q = t3_q(x);
        float s[1] = {q};
        row_reduce<1>(s, s_red);
        const float mean = s[0] / (float)H;
        float v[1] = {(q - mean) * (q - mean)};
        row_reduce<1>(v, s_red);
        const float rstd = rsqrtf(v[0] / (float)H + 1e-5f);
        const float xhat = (q - mean) * rstd;
        const float dgv = valid ? dnh * a.ln_w[h] : 0.f;
        float r[2] = {dgv, dgv * xhat};
        row_reduce<2>(r, s_red);
        if (!valid) return;
        const float dq = (dgv - r[0] / (float)H - xhat * r[1] / (float)H) * rstd;
        float gq[9];
        t3_qgrad(x, gq);
#pragma unroll
        for (int c = 0; c < 9; ++c) db[c] += dq * gq[c];
        st9(a.out1, i, H, h, db);
    }
}

static inline void chain_seg(ChainArgs& c, int ng, const float* const* p, int K, int N) {
    WSeg& S = c.seg[c.nseg++];
    for (int g = 0; g < 3; ++g) S.p[g] = g < ng ? p[g] : nullptr;
    S.K = K; S.N = N; S.ng = ng;
}
static inline void chain_seg1(ChainArgs& c, const float* w, int K, int N) { chain_seg(c, 1, &w, K, N); }

static inline bool chain_ok(const tmdn_tensornet_model* m, int n_rows) {
    return g_tmdn_tuning[TMDN_TUNE_FUSED_NODES] && g_tmdn_tuning[TMDN_TUNE_SMALL_GEMM] && n_rows <= CHAIN_MAX_ROWS &&
           (m->hidden == 32 || m->hidden == 64 || m->hidden == 128) && m->num_layers >= 1;
}

template <int CHAIN>
static inline int chain_launch(const ChainArgs& a, cudaStream_t stream, const char* tag) {
    KTimer ktimer__(tag, stream);
    // 2 atoms per CTA while that still gives every SM at most one CTA's worth of weight traffic, 4 above
    const int na = CHAIN_NA;
    // ring + panel + two vectors + reduction scratch + partial sums ([warps/2 slices][na][9][H] >= [warps][na][3H])
    const size_t nwarps = (size_t)a.H * na / 32;
    const size_t part = nwarps / 2 > 0 ? nwarps / 2 : 1;
    const size_t psum = (part * na * 9 * a.H > nwarps * na * 3 * a.H ? part * na * 9 * a.H : nwarps * na * 3 * a.H);
    const size_t smem = sizeof(float) * ((size_t)WS_STAGES * WS_STAGE_FLOATS + (size_t)na * 9 * a.H +
                                         2 * (size_t)na * 3 * a.H + 64 + psum);
    static bool attr_set = false;
    if (!attr_set) {
        TMDN_CUDA(cudaFuncSetAttribute(tn_chain_kernel<CHAIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    launch_k(tn_chain_kernel<CHAIN>, dim3((unsigned)cdiv(a.N, na)), dim3((unsigned)a.H, (unsigned)na), smem, stream, a);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}
