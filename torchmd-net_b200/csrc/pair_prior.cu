// pair_prior.cu -- pairwise prior potentials on the CSR neighbour structure, energy AND analytic forces (fp32).
//
// Replaces ZBL.post_reduce (torchmdnet/priors/zbl.py:47-63) and Coulomb.post_reduce (priors/coulomb.py:38-50) together
// with the autograd pass that turns them into forces (models/model.py:286-298).  The reference builds a pair list with
// its legacy Distance module (models/utils.py:393-455, loop = False, both directions), evaluates V(d) per directed
// pair, and adds 0.5 * scale * sum_pairs V to the molecule's energy.  Here the pairs are the rows of the CSR built by
// tmdn_csr_build_f32 with the prior's own cutoff (self loops skipped), a warp owns a target atom:
//     e_i    = 0.5 * scale * sum_{e in row i} V(d_e)                (summed per molecule in a fixed order)
//     F_i   -= scale * sum_{e in row i} V'(d_e) * u_e               (the pair appears in row i and in row j; both depend on pos_i)
//   ZBL      V = phi(d s / a_ij) C(d) Z_i Z_j / d,   a_ij = 0.8854 a0 / (Z_i^0.23 + Z_j^0.23),   phi = sum_k c_k exp(-b_k x)
//   Coulomb  V = erf(alpha d) q_i q_j / (1e9 s d)                   (positions converted to nm, priors/coulomb.py:40-41)
//   D2       V = sqrt(C6_i C6_j) / R^6 / (1 + exp(-dmp (R / (Rr_i + Rr_j) - 1))),  R = 1e9 s d  (nm; priors/d2.py:171-191,
//            Grimme 2006; scale = -s_6 / (energy_scale N_A), no cutoff function: the pair set ends at the cutoff)
#include "common.cuh"
#include "edge_common.cuh"

namespace {

struct PriorParams {
    int kind;            // TMDN_PRIOR_ZBL | TMDN_PRIOR_COULOMB | TMDN_PRIOR_D2
    float scale;         // energy prefactor WITHOUT the 0.5 of the double counting
    float cutoff;        // ZBL: cosine cutoff radius
    float dist_scale;    // distance unit -> metres
    float alpha;         // Coulomb: erf range parameter; D2: steepness of the damping function
};

__device__ __forceinline__ void zbl_pair(float d, float zi, float zj, const PriorParams& p, float& v, float& dv) {
    const float a = 0.8854f * 5.29177210903e-11f / (powf(zi, 0.23f) + powf(zj, 0.23f));
    const float k = p.dist_scale / a;
    const float x = d * k;
    const float e0 = 0.1818f * expf(-3.2f * x), e1 = 0.5099f * expf(-0.9423f * x), e2 = 0.2802f * expf(-0.4029f * x),
                e3 = 0.02817f * expf(-0.2016f * x);
    const float phi = e0 + e1 + e2 + e3;
    const float dphi = -(3.2f * e0 + 0.9423f * e1 + 0.4029f * e2 + 0.2016f * e3) * k;
    float c, dc;
    cutoff_fn(d, 0.f, p.cutoff, c, dc);
    const float zz = zi * zj, inv = 1.0f / d;
    v = phi * c * zz * inv;
    dv = (dphi * c + phi * dc) * zz * inv - v * inv;
}
__device__ __forceinline__ void coulomb_pair(float d, float qi, float qj, const PriorParams& p, float& v, float& dv) {
    const float inv = 1.0f / d, qq = qi * qj / (1e9f * p.dist_scale);
    const float ad = p.alpha * d;
    const float er = erff(ad);
    v = er * qq * inv;
    dv = qq * (1.1283791670955126f * p.alpha * expf(-ad * ad) * inv - er * inv * inv);
}

// wi/wj = sqrt(C_6) of the atoms (J/mol nm^6)^(1/2), ri/rj = their van der Waals radii in nm
__device__ __forceinline__ void d2_pair(float d, float wi, float wj, float ri, float rj, const PriorParams& p, float& v,
                                        float& dv) {
    const float nm = 1e9f * p.dist_scale;
    const float R = d * nm, inv = 1.0f / R, inv2 = inv * inv;
    const float c = wi * wj * inv2 * inv2 * inv2;          // C6_ij / R^6
    const float inv_rr = 1.0f / (ri + rj);
    const float f = 1.0f / (1.0f + expf(-p.alpha * (R * inv_rr - 1.0f)));
    v = c * f;
    dv = nm * c * f * (p.alpha * inv_rr * (1.0f - f) - 6.0f * inv);
}

// one warp per target atom; w[i] = atomic number (ZBL: table[z_i], as float) or partial charge (Coulomb);
// D2: w[i] = sqrt(C_6) and w[n + i] = R_r of atom i
__global__ void __launch_bounds__(128) pair_prior_rows(const int* __restrict__ row_ptr, const int* __restrict__ col,
                                                       const float4* __restrict__ geom, const float* __restrict__ w,
                                                       PriorParams p, int n, float* __restrict__ e_row,
                                                       float* __restrict__ forces) {
    pdl_sync();
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const float wi = w[i];
    const float ri = p.kind == TMDN_PRIOR_D2 ? w[n + i] : 0.f;
    float e = 0.f, fx = 0.f, fy = 0.f, fz = 0.f;
    for (int k = row_ptr[i] + lane; k < row_ptr[i + 1]; k += 32) {
        const int j = col[k];
        if (j == i) continue;
        const float4 g = geom[k];
        float v, dv;
        if (p.kind == TMDN_PRIOR_ZBL) zbl_pair(g.w, wi, w[j], p, v, dv);
        else if (p.kind == TMDN_PRIOR_D2) d2_pair(g.w, wi, w[j], ri, w[n + j], p, v, dv);
        else coulomb_pair(g.w, wi, w[j], p, v, dv);
        e += v;
        fx += dv * g.x; fy += dv * g.y; fz += dv * g.z;
    }
    e = warp_sum(e); fx = warp_sum(fx); fy = warp_sum(fy); fz = warp_sum(fz);
    if (lane == 0) {
        e_row[i] = 0.5f * p.scale * e;
        if (forces) {
            forces[3 * (size_t)i] -= p.scale * fx;
            forces[3 * (size_t)i + 1] -= p.scale * fy;
            forces[3 * (size_t)i + 2] -= p.scale * fz;
        }
    }
}

// energy[m] += sum_{i in molecule m} e_row[i]   (batch sorted; one CTA per molecule, fixed-order tree => deterministic)
__global__ void __launch_bounds__(128) pair_prior_mol_add(const float* __restrict__ e_row, const int64_t* __restrict__ batch,
                                                          int n, float* __restrict__ energy) {
    pdl_sync();
    __shared__ float s_part[128];
    const long long m = blockIdx.x;
    int a, b;
    block_molecule_range(batch, n, m, a, b);
    float s = 0.f;
    for (int i = a + threadIdx.x; i < b; i += 128) s += e_row[i];
    s_part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (threadIdx.x < o) s_part[threadIdx.x] += s_part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) energy[m] += s_part[0];
}
}  // namespace

extern "C" size_t tmdn_pair_prior_workspace_bytes(int64_t n_atoms, int64_t max_edges) {
    if (n_atoms <= 0 || max_edges <= 0) return 0;
    Arena a(nullptr, 0);
    a.get<int32_t>(n_atoms + 1); a.get<int32_t>(max_edges); a.get<float>(4 * (size_t)max_edges); a.get<int32_t>(4);
    a.get<float>(n_atoms);
    a.get<char>(tmdn_csr_build_workspace_bytes(TMDN_BRUTE, n_atoms, nullptr, 1.0));
    return a.off + 256;
}

extern "C" int tmdn_pair_prior_f32(int kind, const float* pos, const int64_t* batch, const float* weight, int64_t n_atoms,
                                   int64_t n_mol, double cutoff, double energy_prefactor, double distance_scale,
                                   double alpha, int64_t max_edges, float* energy, float* forces, int32_t* num_found,
                                   void* workspace, size_t workspace_bytes, tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_CHECK_ARG(kind == TMDN_PRIOR_ZBL || kind == TMDN_PRIOR_COULOMB || kind == TMDN_PRIOR_D2, "pair_prior: unknown kind %d", kind);
    TMDN_CHECK_ARG(pos && batch && weight && energy && num_found && workspace, "pair_prior: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_mol > 0 && max_edges > 0, "pair_prior: bad sizes");
    TMDN_CHECK_ARG(kind == TMDN_PRIOR_COULOMB || cutoff > 0, "pair_prior: ZBL and D2 need a positive cutoff distance");
    Arena a(workspace, workspace_bytes);
    int32_t* row_ptr = a.get<int32_t>(n_atoms + 1);
    int32_t* col = a.get<int32_t>(max_edges);
    float* geom = a.get<float>(4 * (size_t)max_edges);
    int32_t* found = a.get<int32_t>(4);
    float* e_row = a.get<float>(n_atoms);
    const size_t gws = tmdn_csr_build_workspace_bytes(TMDN_BRUTE, n_atoms, nullptr, 1.0);
    void* graph_ws = a.get<char>(gws);
    if (!a.ok) {
        tmdn_set_error("pair_prior: workspace too small (%zu < %zu)", workspace_bytes, a.off);
        return TMDN_EWORKSPACE;
    }
    // Coulomb has no cutoff (Distance(0, inf)): every pair of a molecule; a huge finite radius keeps the builder's d^2 test
    const double rc = kind == TMDN_PRIOR_COULOMB ? 1e18 : cutoff;
    TMDN_TRY(tmdn_csr_build_f32(TMDN_BRUTE, pos, batch, n_atoms, nullptr, 0, 0.0, rc, max_edges, row_ptr, col, geom, found,
                                graph_ws, gws, stream_));
    TMDN_CUDA(cudaMemcpyAsync(num_found, found, sizeof(int32_t), cudaMemcpyDeviceToDevice, stream));
    PriorParams p;
    p.kind = kind;
    p.scale = (float)energy_prefactor;
    p.cutoff = (float)cutoff;
    p.dist_scale = (float)distance_scale;
    p.alpha = (float)alpha;
    {
        KTimer ktimer__("pair_prior_rows", stream);
        launch_k(pair_prior_rows, (unsigned)cdiv(n_atoms, 4), 128, 0, stream, (const int*)row_ptr, (const int*)col,
                 (const float4*)geom, weight, p, (int)n_atoms, e_row, forces);
        TMDN_LAUNCH_CHECK();
    }
    {
        KTimer ktimer__("pair_prior_mol_add", stream);
        launch_k(pair_prior_mol_add, (unsigned)n_mol, 128, 0, stream, (const float*)e_row, batch, (int)n_atoms, energy);
        TMDN_LAUNCH_CHECK();
    }
    return TMDN_OK;
}

// out[m] += sum_{i in molecule m} values[i]  (batch sorted): the molecular sum of the Atomref prior, deterministic
// (torch's index_add_ on CUDA adds with atomics in arbitrary order).  No counterpart call in the reference: there the sum
// is part of the output head's scatter (models/model.py:266-272).
extern "C" int tmdn_segment_add_f32(const float* values, const int64_t* batch, int64_t n_atoms, int64_t n_mol, float* out,
                                    tmdn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    TMDN_CHECK_ARG(values && batch && out, "segment_add: null pointer argument");
    TMDN_CHECK_ARG(n_atoms > 0 && n_mol > 0 && n_atoms < (1LL << 31), "segment_add: bad sizes");
    KTimer ktimer__("segment_add", stream);
    launch_k(pair_prior_mol_add, (unsigned)n_mol, 128, 0, stream, values, batch, (int)n_atoms, out);
    TMDN_LAUNCH_CHECK();
    return TMDN_OK;
}
