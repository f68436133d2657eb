/* tmdn.h -- C ABI of the B200-native TorchMD-NET energy+force hot path (libtmdn.so).
 *
 * Plain pointers and sizes only; no torch types.  Every pointer marked "device" is a CUDA device
 * pointer on the current device; `stream` is a cudaStream_t.  No entry point allocates, frees or
 * synchronises: the caller passes a workspace (size from the matching *_workspace_bytes), so every
 * call is CUDA-graph capturable.  All entry points return TMDN_OK (0) or an error code and never
 * throw; tmdn_last_error() gives the message of the last failure on the calling thread.
 *
 * Reference interfaces replaced (paths relative to the reference repository):
 *   tmdn_neighbors_*            torch op torchmdnet_neighbors::get_neighbor_pairs
 *                               schema  torchmdnet/neighbors/neighbors.cpp:3-5
 *                               brute   torchmdnet/neighbors/neighbors_cuda_brute.cuh:61-101
 *                               shared  torchmdnet/neighbors/neighbors_cuda_shared.cuh:71-106
 *                               cell    torchmdnet/neighbors/neighbors_cuda_cell.cuh:338-378
 *   tmdn_neighbors_backward_*   NeighborAutograd::backward, torchmdnet/neighbors/neighbors_cuda.cu:43-71
 *   tmdn_is_stream_capturing    torch_extension::is_stream_capturing, torchmdnet/models/utils.py:533-555
 *   tmdn_csr_from_pairs         (new) CSR-by-target view of the pair list consumed by the model kernels
 *   tmdn_tensornet_*            TensorNet.forward + Scalar head + TorchMD_Net force path:
 *                               torchmdnet/models/tensornet.py:200-234,287-326,376-410,
 *                               torchmdnet/models/output_modules.py:27-43,76-77,
 *                               torchmdnet/models/model.py:232-300 (autograd.grad replaced by
 *                               hand-written adjoint kernels)
 *   tmdn_et_*                   TorchMD_ET.forward + EquivariantScalar head + force path:
 *                               torchmdnet/models/torchmd_et.py:154-187,272-347,
 *                               torchmdnet/models/utils.py:73-108,496-522,
 *                               torchmdnet/models/output_modules.py:111-115
 */
#ifndef TMDN_H
#define TMDN_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TMDN_VERSION 100
#define TMDN_MAX_LAYERS 16

typedef void* tmdn_stream_t; /* cudaStream_t */

enum tmdn_status {
    TMDN_OK = 0,
    TMDN_EINVAL = 1,      /* bad argument (shape, null pointer, unsupported size) */
    TMDN_ECUDA = 2,       /* a CUDA runtime call / kernel launch failed */
    TMDN_EWORKSPACE = 3,  /* workspace too small */
    TMDN_EUNSUPPORTED = 4 /* valid request this build does not implement */
};

enum tmdn_strategy { TMDN_BRUTE = 0, TMDN_SHARED = 1, TMDN_CELL = 2 };
enum tmdn_rbf { TMDN_RBF_EXPNORM = 0, TMDN_RBF_GAUSS = 1 };
enum tmdn_act { TMDN_ACT_SILU = 0, TMDN_ACT_TANH = 1, TMDN_ACT_SIGMOID = 2, TMDN_ACT_SSP = 3 };
enum tmdn_filter_mode {
    TMDN_FILTER_EXACT = 0, /* per-edge filter MLP evaluated exactly (fp32 FFMA GEMM chain), [E,*] in the workspace */
    TMDN_FILTER_TABLE = 1  /* cubic Hermite table in d (host-built in fp64), evaluated inside the edge kernels */
};

int tmdn_version(void);
const char* tmdn_last_error(void);

/* 1 if `stream` is being captured into a CUDA graph, 0 if not, <0 on error. */
int tmdn_is_stream_capturing(tmdn_stream_t stream);

/* Measurement hooks (new; no reference counterpart).  tmdn_launch_count: kernels launched by this library since
 * load (process-wide).  tmdn_profile_enable(1) makes every entry point bracket each kernel family with a CUDA
 * event pair on the caller's stream (skipped while the stream is being captured); tmdn_profile_report
 * synchronises those events and writes one line "family launches total_ms" per family into buf;
 * tmdn_profile_enable(0|1) also clears what was recorded. */
long long tmdn_launch_count(void);
int tmdn_profile_enable(int on);
int tmdn_profile_report(char* buf, size_t buf_bytes);

/* Kernel-selection knobs (new; for A/B measurements and so that tests can force each path).  Keys:
 *   "small_gemm" 1 = node GEMMs / channel mixes of small systems use the latency-oriented fp32 kernel (default 1)
 *   "lpt_rows"   1 = edge kernels of small systems visit rows in order of decreasing degree (default 1)
 *   "pdl"        1 = dependent kernels are launched with programmatic stream serialisation (default 0: measured
 *                    slower inside CUDA graphs, profiles/r01i_bench_*_pdl_{on,off}.json)
 *   "edge_persist" 1 = on large systems (>= 8192 rows, H = 128, table filters) the TensorNet message-passing kernels
 *                    run as one persistent CTA per SM (default 1; 10-17 % faster than one CTA per row at 98 k atoms)
 *   "hot_rows"     number of filter-table rows (from the cutoff end of the grid) those kernels keep in shared memory
 *                    (default 0: measured no gain, shared memory and L1 share the data pipe that bounds them)
 *   "edge_variant" A/B instantiations of those kernels (bit 0: forward with 2 edges per batch, bit 1: backward with 8
 *                    workers of one edge at 64 registers instead of 4 workers of two edges at 128, bit 3: forward as
 *                    one CTA of 8 workers per SM instead of two CTAs of 5); same results
 *   "small_csr"  1 = brute-force CSR builds of <= 256 atoms run as ONE kernel (count, scan, fill, clamp in a cluster of 8
 *                    CTAs exchanging their row counts through distributed shared memory) instead of four launches; same
 *                    rows in the same order (default 1)
 * Returns the previous value (>= 0); value < 0 only queries; -1 for an unknown key (tmdn_last_error says which).
 * Process-wide. */
int tmdn_set_tuning(const char* key, int value);

/* ---------------------------------------------------------------- neighbor list
 * Semantics = the reference CUDA op: all pairs of atoms of the same batch with
 * cutoff_lower^2 <= d^2 < cutoff_upper^2; delta = pos[i]-pos[j] (minimum image if use_periodic:
 * triclinic round() z->y->x for brute/shared, rectangular floor(x/L+0.5) for cell); emitted as (i,j), i>j,
 * plus (j,i) with -delta if include_transpose, plus (i,i,0,0) if loop.  Order is unspecified.
 * Outputs are sized max_pairs: slots beyond the number found hold neighbors=-1, deltas=0,
 * distances=0.  *num_found (device int32) receives the number FOUND, which may exceed max_pairs (the
 * excess is dropped).  `box` is 9 host doubles, rows a,b,c (ignored unless use_periodic, except that
 * the cell strategy always needs a diagonal box for binning).
 */
size_t tmdn_neighbors_workspace_bytes(int strategy, int64_t n_atoms, const double box[9], double cutoff_upper,
                                      int elem_bytes);
int tmdn_neighbors_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms, const double box[9],
                       int use_periodic, double cutoff_lower, double cutoff_upper, int loop, int include_transpose,
                       int64_t max_pairs, int32_t* neighbors, float* deltas, float* distances, int32_t* num_found,
                       void* workspace, size_t workspace_bytes, tmdn_stream_t stream);
int tmdn_neighbors_f64(int strategy, const double* pos, const int64_t* batch, int64_t n_atoms, const double box[9],
                       int use_periodic, double cutoff_lower, double cutoff_upper, int loop, int include_transpose,
                       int64_t max_pairs, int32_t* neighbors, double* deltas, double* distances, int32_t* num_found,
                       void* workspace, size_t workspace_bytes, tmdn_stream_t stream);
/* grad_pos[n_atoms,3] = sum_e (+g_e at neighbors[0][e], -g_e at neighbors[1][e]),
 * g_e = grad_deltas[e] + deltas[e]/distances[e]*grad_distances[e], rows with distances==0 skipped.
 * grad_pos is overwritten.  Either grad pointer may be NULL (treated as zero). */
int tmdn_neighbors_backward_f32(const int32_t* neighbors, const float* deltas, const float* distances,
                                const float* grad_deltas, const float* grad_distances, int64_t max_pairs,
                                int64_t n_atoms, float* grad_pos, tmdn_stream_t stream);
int tmdn_neighbors_backward_f64(const int32_t* neighbors, const double* deltas, const double* distances,
                                const double* grad_deltas, const double* grad_distances, int64_t max_pairs,
                                int64_t n_atoms, double* grad_pos, tmdn_stream_t stream);

/* ---------------------------------------------------------------- CSR view of the pair list (fp32)
 * Rows = target atoms (neighbors[target_row][e]), columns = the other atom, sorted by column inside each
 * row so that every segmented reduction downstream is deterministic.  geom[e] = (ux,uy,uz,d) with
 * u = (pos_target - pos_source)/d  (0 for self loops).  num_found/max_pairs as produced by tmdn_neighbors_*.
 * row_ptr has n_atoms+1 entries.  If the list overflowed (found > max_pairs) only the stored pairs are used.
 */
size_t tmdn_csr_workspace_bytes(int64_t n_atoms, int64_t max_pairs);
int tmdn_csr_from_pairs(const int32_t* neighbors, const float* deltas, const float* distances,
                        const int32_t* num_found, int64_t max_pairs, int64_t n_atoms, int target_row,
                        int32_t* row_ptr, int32_t* col, float* geom /* [max_pairs,4] */, void* workspace,
                        size_t workspace_bytes, tmdn_stream_t stream);

/* Direct build (new): the same pair set as tmdn_neighbors_f32(loop=1, include_transpose=1), emitted straight as
 * CSR rows (target atom -> sources) without materialising the reference's pair list.  `batch` must be sorted
 * ascending (molecules contiguous); brute/shared search each atom's own molecule, cell searches the 27 cells
 * around it (diagonal box).  Row order: ascending source index (brute) / cell-traversal order with cells
 * internally sorted by atom index (cell) -- a pure function of the input, so downstream sums are reproducible.
 * *num_found = edges found; if it exceeds max_edges the structure is truncated (row_ptr clamped to max_edges).
 */
size_t tmdn_csr_build_workspace_bytes(int strategy, int64_t n_atoms, const double box[9], double cutoff_upper);
int tmdn_csr_build_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms, const double box[9],
                       int use_periodic, double cutoff_lower, double cutoff_upper, int64_t max_edges,
                       int32_t* row_ptr, int32_t* col, float* geom /* [max_edges,4] */, int32_t* num_found,
                       void* workspace, size_t workspace_bytes, tmdn_stream_t stream);

/* Rows only for the first n_rows atoms (the atoms a rank owns); atoms [n_rows, n_atoms) are ghosts that appear as
 * sources only.  row_ptr still has n_atoms+1 entries (ghost rows empty). */
int tmdn_csr_build_rows_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms, int64_t n_rows,
                            const double box[9], int use_periodic, double cutoff_lower, double cutoff_upper,
                            int64_t max_edges, int32_t* row_ptr, int32_t* col, float* geom, int32_t* num_found,
                            void* workspace, size_t workspace_bytes, tmdn_stream_t stream);

/* Allocation-free variant of tmdn_csr_build_f32.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and results as tmdn_csr_build_f32 without the workspace: the scratch (row counters,
 * scan partials, cell-sorted atom copies) is a file-scope static __device__ array of TMDN_CSR_STATIC_SCRATCH_BYTES
 * (4 MiB) that every GPU loading the library holds until exit -- enough for about 1 000 000 atoms with the brute /
 * shared strategies (4 B per atom) and about 65 000 with the cell strategy (60 B per atom + 8 B per cell;
 * tmdn_csr_build_workspace_bytes gives the exact need).  Larger
 * problems are refused with TMDN_EWORKSPACE and a message naming the limit.  The call allocates nothing (no cudaMalloc,
 * cudaFree or allocating library routine) and does not synchronise.  All calls on a device share the one array, which
 * keeps the previous call's contents (the builder clears what must start at zero): calls that overlap on the GPU --
 * different streams without an ordering between them -- race on it; calls on one stream are ordered and safe. */
#define TMDN_CSR_STATIC_SCRATCH_BYTES (4 << 20)
int tmdn_csr_build_static_f32(int strategy, const float* pos, const int64_t* batch, int64_t n_atoms, const double box[9],
                              int use_periodic, double cutoff_lower, double cutoff_upper, int64_t max_edges,
                              int32_t* row_ptr, int32_t* col, float* geom /* [max_edges,4] */, int32_t* num_found,
                              tmdn_stream_t stream);

/* ---------------------------------------------------------------- TensorNet energy + forces (fp32)
 * All weight pointers are device fp32 arrays prepared once by the host side (layouts below; H = hidden,
 * K = num_rbf).  "T" suffix = transposed copy used by the backward kernels.
 */
typedef struct tmdn_filter_set {
    /* EXACT mode: MLP K -> n1 [-> n2 -> n3], weights [out,in] row-major (+ transposed copies unused: the
     * filter derivative is propagated forward as a tangent).  n_layers = 1 (embedding) or 3 (interaction);
     * the last layer's rows are ordered g*H+h (g = I/A/S group).  final_act: activation after the last layer. */
    int32_t n_layers, final_act;
    int32_t dims[4]; /* K, n1, n2, n3 */
    const float* w[3];
    const float* b[3];
    /* TABLE mode: [n_int][3][H] float4 cubic coefficients of f(d)*C(d) on n_int intervals over [0, cutoff_upper],
     * uniform in t(d) = (exp(-grid_alpha d) - exp(-grid_alpha r_c)) + grid_gamma (r_c - d), counted from the cutoff
     * (row 0 ends at d = r_c; grid_alpha = 0: uniform in d, t = r_c - d); local coordinate s = t(d) n_int / t(0) - k,
     * value = c.x + s (c.y + s (c.z + s c.w)).  See csrc/table_grid.cuh. */
    const float* table;
    int32_t n_int;
    float grid_alpha, grid_gamma;
} tmdn_filter_set;

/* Kernel of the plain node GEMMs (MLPs of TensorNet, projections of ET).  TF32X3 = tcgen05 with the three-product tf32
 * split; TMEM holds the partial product of one 32-channel K chunk and the chunks are summed in fp32 round-to-nearest
 * on the CUDA cores, because the tensor-core accumulator truncates (accumulating all of K in TMEM biased same-sign sums:
 * 1.2e-5 relative on the energy of a 5184-atom water box).  Measured rms error 1.0-1.9e-8 of sum|a w| per GEMM, below the
 * 2.8e-8 of an fp32 FFMA GEMM.  FP32 = the FFMA kernel.  The host default is TF32X3. */
enum { TMDN_GEMM_FP32 = 0, TMDN_GEMM_TF32X3 = 1 };

typedef struct tmdn_tensornet_model {
    int32_t hidden, num_rbf, num_layers, max_z;
    int32_t rbf_type, activation, so3, filter_mode;
    int32_t gemm_mode;   /* TMDN_GEMM_FP32 (default) | TMDN_GEMM_TF32X3 */
    float cutoff_lower, cutoff_upper, mean, std;
    const float* rbf_p0; /* expnorm: means[K]; gauss: offset[K] */
    const float* rbf_p1; /* expnorm: betas[K]; gauss: coeff broadcast to [K] */
    const float* P;      /* [max_z,H]  emb @ emb2.weight[:, :H]^T + emb2.bias   (tensornet.py:310-312) */
    const float* Q;      /* [max_z,H]  emb @ emb2.weight[:, H:]^T */
    tmdn_filter_set emb_filter;                  /* distance_proj1..3 stacked -> [3H] */
    const float* emb_mix[3];  const float* emb_mixT[3];   /* tensor_embedding.linears_tensor.k  [H,H] */
    const float* emb_ln_w;    const float* emb_ln_b;      /* init_norm */
    const float* emb_w1;      const float* emb_b1;  const float* emb_w1T;   /* linears_scalar.0 [2H,H] */
    const float* emb_w2;      const float* emb_b2;  const float* emb_w2T;   /* linears_scalar.1 [3H,2H], rows g*H+h */
    tmdn_filter_set layer_filter[TMDN_MAX_LAYERS];        /* layers.l.linears_scalar.0-2 */
    const float* layer_mix[TMDN_MAX_LAYERS][6];           /* layers.l.linears_tensor.0-5 [H,H] */
    const float* layer_mixT[TMDN_MAX_LAYERS][6];
    const float* out_ln_w;    const float* out_ln_b;      /* out_norm [3H] */
    const float* out_w;       const float* out_b;   const float* out_wT;    /* linear [H,3H] */
    const float* head_w1;     const float* head_b1; const float* head_w1T;  /* output_network.0 [H/2,H] */
    const float* head_w2;     const float* head_b2;                         /* output_network.2 [1,H/2] */
} tmdn_tensornet_model;

size_t tmdn_tensornet_workspace_bytes(const tmdn_tensornet_model* m, int64_t n_atoms, int64_t max_edges,
                                      int64_t n_mol, int strategy, const double box[9] /* cell strategy only */);
/* One energy(+force) evaluation.  z,batch int64 device [n_atoms] (batch sorted non-decreasing, values in
 * [0,n_mol)); pos fp32 device [n_atoms,3]; energy fp32 device [n_mol]; forces fp32 device [n_atoms,3] or NULL
 * (forward only).  Neighbors are searched with `strategy` under box/use_periodic; max_edges is the pair
 * capacity INCLUDING self loops (reference: max_num_neighbors * n_atoms, models/utils.py:234-236).
 * status (device int32[4]): [0] edges found, [1] 1 if found > max_edges, [2] 1 if some z was outside [0, max_z),
 * [3] 1 if batch was not sorted.  When any of [1..3] is set, energy and forces are filled with NaN on the device: a
 * truncated edge list is not symmetric, so the analytic forces would not belong to the energy (the reference's
 * autograd forces stay consistent with its truncated energy; this path says "invalid" instead of being subtly wrong).
 */
int tmdn_tensornet_energy_forces(const tmdn_tensornet_model* m, const int64_t* z, const float* pos,
                                 const int64_t* batch, int64_t n_atoms, int64_t n_mol, const double box[9],
                                 int use_periodic, int strategy, int64_t max_edges, float* energy, float* forces,
                                 int32_t* status, void* workspace, size_t workspace_bytes, tmdn_stream_t stream);
/* The same evaluation cut at its halo points, for spatially decomposed runs (one rank owns the first n_rows atoms of
 * its local arrays, the remaining n_atoms - n_rows are ghosts owned by other ranks).  Call order:
 *   EMBED, [exchange "Y0"], LAYER_FWD(0), [exchange "Y1"], ..., LAYER_FWD(L-1), HEAD, [exchange B0],
 *   LAYER_BWD(L-1), [exchange B1], ..., LAYER_BWD(0), [exchange B_L], FINAL
 * where an exchange copies the rows of the named workspace buffer ([n_atoms,9,H] fp32) from their owners into the
 * ghost rows, and B_k = "Dias" for even k, "Dias1" for odd k: consecutive backward halo points alternate between two
 * buffers, so a rank may deliver its rows into a neighbour's ghost rows while that neighbour still reads the previous
 * halo point.  energy receives the sum over the OWNED atoms only (+ mean); forces has n_rows rows.
 * tmdn_tensornet_energy_forces == all stages in sequence with n_rows = n_atoms. */
enum tmdn_tn_stage { TMDN_TN_EMBED = 0, TMDN_TN_LAYER_FWD = 1, TMDN_TN_HEAD = 2, TMDN_TN_LAYER_BWD = 3, TMDN_TN_FINAL = 4,
                     TMDN_TN_GRAPH = 5 };
int tmdn_tensornet_stage(const tmdn_tensornet_model* m, int stage, int layer, const int64_t* z, const float* pos,
                         const int64_t* batch, int64_t n_atoms, int64_t n_rows, int64_t n_mol, const double box[9],
                         int use_periodic, int strategy, int64_t max_edges, float* energy, float* forces,
                         int32_t* status, void* workspace, size_t workspace_bytes, tmdn_stream_t stream);
/* A stage over the owned rows [row_begin, row_end) only -- every operation between two halo points is local to a row
 * (the edge kernels gather from any local atom but write their own row), so a rank can evaluate the rows its neighbours
 * are waiting for first, start their transfer, and evaluate the interior rows while the transfer is in flight.
 * Here GRAPH (neighbour search -> CSR of all owned rows, z conversion, exact-mode filters; row range ignored) is a
 * stage of its own and must precede EMBED; HEAD performs the molecular sums in the call whose range ends at n_rows
 * (make it the last one).  Edge-kernel sums do not depend on the split (launch shapes follow n_rows); the node GEMMs
 * pick their kernel by the size of the range, so a split result is bit-identical when both parts select the same
 * GEMM kernel as the whole (always beyond ~10^4 rows) and within rounding otherwise. */
int tmdn_tensornet_stage_rows(const tmdn_tensornet_model* m, int stage, int layer, const int64_t* z, const float* pos,
                              const int64_t* batch, int64_t n_atoms, int64_t n_rows, int64_t row_begin, int64_t row_end,
                              int64_t n_mol, const double box[9], int use_periodic, int strategy, int64_t max_edges,
                              float* energy, float* forces, int32_t* status, void* workspace, size_t workspace_bytes,
                              tmdn_stream_t stream);
/* Halo pack: out[r, :] = src[idx[r], :] for r < n_rows (rows of row_floats fp32).  The receive side writes straight into
 * the ghost rows (the contiguous tail of the local arrays), so there is no unpack. */
int tmdn_halo_pack_f32(const float* src, const int32_t* idx, int64_t n_rows, int64_t row_floats, float* out,
                       tmdn_stream_t stream);
/* Halo push over peer memory (NVLink): the pack and the transport in one kernel.  For peer p (p < n_peers) the rows
 * idx[seg[p] .. seg[p+1]) of src are written to dst[p], a device pointer into the PEER's address space (its ghost rows;
 * obtained by the host side through CUDA IPC / symmetric memory), row r of the segment at dst[p] + r * row_floats.
 * seg (int64[n_peers+1]), idx (int32) and dst (float*[n_peers]) are device arrays.  The stores are followed by a
 * system-scope fence; the caller orders "data has landed" with a cross-rank barrier on the same stream. */
int tmdn_halo_push_f32(const float* src, const int32_t* idx, const int64_t* seg, float* const* dst, int n_peers,
                       int64_t n_rows_total, int64_t row_floats, tmdn_stream_t stream);
/* Grouped channel mix of TensorNet (tensornet.py:273-284, 386-408: linears_tensor on the I | A | S parts):
 *   y[a, c, :] (+)= x[a, c, :] * w[g(c)]^T,   x, y [n_atoms,9,H] fp32, w [3,H,H] (g = 0 for c = 0, 1 for c = 1..3, 2 for
 *   c = 4..8).  H = 128 runs on the tcgen05 tensor cores with the three-product tf32 split (fp32-level accuracy);
 * other widths on the fp32 FFMA kernel.  This is the operator the model path calls for every mix; exported for tests. */
int tmdn_channel_mix_f32(const float* x, const float* w, float* y, int64_t n_atoms, int H, int accumulate,
                         tmdn_stream_t stream);
/* Plain node GEMM of the path (embedding / output / head MLPs of TensorNet, every ET projection):
 *   c[M,N] (+)= a[M,K] * w[N,K]^T + bias[N] (bias may be null).  M, N, K multiples of 4 run on the tcgen05 tensor cores
 * with the three-product tf32 split (fp32-level accuracy), anything else on the fp32 FFMA kernel.  Exported for tests. */
int tmdn_node_gemm_f32(const float* a, const float* w, const float* bias, float* c, int64_t M, int64_t N, int64_t K,
                       int accumulate, tmdn_stream_t stream);
/* Debug/inspection: byte offset and size of a named intermediate inside the workspace (names as in
 * oracle/staged_tensornet.py: "row_ptr","col","geom","ias","nh","fias","X0","Y0","M0","tauh0","X1",...,"e_atom","Dias","gpos"). */
int tmdn_tensornet_workspace_lookup(const tmdn_tensornet_model* m, int64_t n_atoms, int64_t max_edges,
                                    int64_t n_mol, int strategy, const double box[9], const char* name,
                                    size_t* offset, size_t* bytes);

/* ---------------------------------------------------------------- pairwise priors: energy + analytic forces (fp32)
 * ZBL.post_reduce (torchmdnet/priors/zbl.py:47-63), Coulomb.post_reduce (priors/coulomb.py:38-50) and D2.post_reduce
 * (priors/d2.py:163-193) plus the autograd pass
 * that differentiates them (models/model.py:286-298), on the pair set of the legacy Distance module (models/utils.py:393-455:
 * open boundaries, both directions, no self pairs): the neighbour search runs inside the call (brute strategy, the prior's
 * own cutoff; Coulomb has none: every pair of a molecule).
 *   energy[m] += 0.5 * energy_prefactor * sum_{pairs of m} V(d);   forces[i] -= dE/dpos_i   (forces may be NULL)
 *   ZBL:     V = phi(d * distance_scale / a_ij) * C(d; cutoff) * Z_i Z_j / d,  weight[i] = atomic number of atom i
 *   Coulomb: V = erf(alpha d) q_i q_j / (1e9 distance_scale d),               weight[i] = partial charge of atom i
 *   D2:      V = sqrt(C6_i C6_j) / R^6 / (1 + exp(-alpha (R / (Rr_i + Rr_j) - 1))),  R = 1e9 distance_scale d  (nm);
 *            weight has 2 * n_atoms entries: weight[i] = sqrt(C_6 of atom i) in (J/mol nm^6)^(1/2), weight[n_atoms + i] =
 *            its van der Waals radius in nm; alpha = the damping steepness (20); the pair set ends at `cutoff`
 * energy_prefactor = 2.30707755e-28 / energy_scale / distance_scale for ZBL and Coulomb (the caller's unit conversion,
 * zbl.py:58), -s_6 / (energy_scale * 6.02214076e23) for D2 (d2.py:168,188-191).
 * *num_found = directed pairs + n_atoms self entries found; if it exceeds max_edges the pair set was truncated (the
 * reference asserts in that case, models/utils.py:419-425; the host side raises). */
enum { TMDN_PRIOR_ZBL = 0, TMDN_PRIOR_COULOMB = 1, TMDN_PRIOR_D2 = 2 };
size_t tmdn_pair_prior_workspace_bytes(int64_t n_atoms, int64_t max_edges);
int tmdn_pair_prior_f32(int kind, const float* pos, const int64_t* batch, const float* weight, int64_t n_atoms,
                        int64_t n_mol, double cutoff, double energy_prefactor, double distance_scale, double alpha,
                        int64_t max_edges, float* energy, float* forces, int32_t* num_found, void* workspace,
                        size_t workspace_bytes, tmdn_stream_t stream);

/* out[m] += sum of values[i] over the atoms of molecule m (batch sorted ascending), one CTA per molecule, fixed summation
 * order: the deterministic molecular sum behind the Atomref prior (torchmdnet/priors/atomref.py:35-44 adds atomref[z] per atom
 * before the head's scatter, models/model.py:266-272; torch's index_add_ on CUDA sums with atomics). */
int tmdn_segment_add_f32(const float* values, const int64_t* batch, int64_t n_atoms, int64_t n_mol, float* out,
                         tmdn_stream_t stream);

/* ---------------------------------------------------------------- Equivariant Transformer energy + forces (fp32)
 * TorchMD_ET.forward + NeighborEmbedding + EquivariantMultiHeadAttention + EquivariantScalar head + force path
 * (torchmdnet/models/torchmd_et.py:154-187,272-347; models/utils.py:73-108,496-522; models/output_modules.py:111-115).
 * Weights are device fp32 arrays prepared once by the host side.  "T" = transposed copy for the backward GEMMs.
 * v_proj / dv_proj rows are re-ordered part-major (row p*H + head*hd + c <- reference row head*3hd + p*hd + c).
 * The distance filters are univariate in d and are passed as cubic Hermite tables built in fp64 on the grid
 * (n_int, grid_alpha, grid_gamma) of tmdn_filter_set:
 *   layer table [n_int][4][H] float4 : act(dk_proj(rbf(d))), act(dv_proj(rbf(d))) parts 0..2 (NO cutoff factor;
 *                                      a missing filter (distance_influence) is the constant 1)
 *   ne_table    [n_int][H]    float4 : distance_proj(rbf(d)) * C(d)      (models/utils.py:98-99)
 * head_dim = hidden/num_heads must be a power of two <= 32.
 */
typedef struct tmdn_et_layer {
    const float *ln_w, *ln_b;
    const float *qkv_w, *qkv_b, *qkv_wT;   /* [5H,H] = q | k | v(part-major), [5H], [H,5H] */
    const float *vp_w, *vp_wT;             /* vec_proj [3H,H], [H,3H] */
    const float *o_w, *o_b, *o_wT;         /* o_proj [3H,H], [3H], [H,3H] */
    const float* table;
} tmdn_et_layer;

typedef struct tmdn_et_gated {              /* GatedEquivariantBlock: hin -> out, intermediate = hin */
    int32_t hin, out;
    const float *w1, *w1T;                  /* vec1_proj [hin,hin] */
    const float *w2, *w2T;                  /* vec2_proj [out,hin], [hin,out] */
    const float *u0_w, *u0_b, *u0_wT;       /* update_net.0 [hin,2hin], [hin], [2hin,hin] */
    const float *u2_w, *u2_b, *u2_wT;       /* update_net.2 [2out,hin], [2out], [hin,2out] */
} tmdn_et_gated;

typedef struct tmdn_et_model {
    int32_t hidden, num_layers, num_heads, max_z;
    int32_t activation, attn_activation, has_neighbor_embedding, n_int;
    int32_t gemm_mode;                      /* TMDN_GEMM_FP32 (default, exact fp32 FFMA) | TMDN_GEMM_TF32X3 (tensor cores) */
    float cutoff_lower, cutoff_upper, mean, std;
    float grid_alpha, grid_gamma;           /* interval grid of all tables (see tmdn_filter_set) */
    const float* emb;                       /* embedding [max_z,H] */
    const float *ne_emb, *ne_table;         /* neighbor_embedding.embedding [max_z,H], filter table */
    const float *ne_comb_w, *ne_comb_b;     /* combine [H,2H], [H] */
    const float* ne_comb_w2T;               /* (combine.weight[:, H:])^T  [H,H]: adjoint w.r.t. the neighbour half */
    tmdn_et_layer layer[TMDN_MAX_LAYERS];
    const float *out_ln_w, *out_ln_b;
    tmdn_et_gated head[2];
} tmdn_et_model;

size_t tmdn_et_workspace_bytes(const tmdn_et_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                               int strategy, const double box[9]);
/* Same contract as tmdn_tensornet_energy_forces. */
int tmdn_et_energy_forces(const tmdn_et_model* m, const int64_t* z, const float* pos, const int64_t* batch,
                          int64_t n_atoms, int64_t n_mol, const double box[9], int use_periodic, int strategy,
                          int64_t max_edges, float* energy, float* forces, int32_t* status, void* workspace,
                          size_t workspace_bytes, tmdn_stream_t stream);
int tmdn_et_workspace_lookup(const tmdn_et_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol, int strategy,
                             const double box[9], const char* name, size_t* offset, size_t* bytes);

/* ---------------------------------------------------------------- Graph network (TorchMD_GN) energy + forces (fp32)
 * TorchMD_GN.forward + NeighborEmbedding + InteractionBlock / CFConv + Scalar head + force path
 * (torchmdnet/models/torchmd_gn.py:148-290; models/utils.py:73-108; models/output_modules.py:49-77; model.py:286-298):
 *   x = emb(z) [-> neighbour embedding];  per layer: h = lin1 x;  agg_i = sum_{j != i, d_ij < r_c} W(d_ij) * h_j,
 *   W(d) = mlp(rbf(d)) C(d);  x += lin(act(lin2 agg));  e_i = (w2 . act(W1 x + b1) + b2) std;  y = sum e_i + mean.
 * The pair set is the one of the reference's legacy Distance module (no self pairs); aggr = "add" only.  The filters are
 * univariate in d and are passed as cubic Hermite tables on the grid of tmdn_filter_set:
 *   layer table [n_int][F] float4 : mlp(rbf(d)) * C(d);   ne_table [n_int][H] float4 : distance_proj(rbf(d)) * C(d).
 * "T" = transposed copy for the backward GEMMs.  hidden, num_filters: multiples of 32 in [32,1024]. */
typedef struct tmdn_gn_layer {
    const float *lin1_w, *lin1_wT;          /* conv.lin1 [F,H] (no bias), [H,F] */
    const float* table;
    const float *lin2_w, *lin2_b, *lin2_wT; /* conv.lin2 [H,F], [H], [F,H] */
    const float *lin_w, *lin_b, *lin_wT;    /* lin [H,H], [H], [H,H] */
} tmdn_gn_layer;

typedef struct tmdn_gn_model {
    int32_t hidden, num_filters, num_layers, max_z;
    int32_t activation, has_neighbor_embedding, n_int, gemm_mode;
    float cutoff_lower, cutoff_upper, mean, std;
    float grid_alpha, grid_gamma;
    const float* emb;                       /* embedding [max_z,H] */
    const float *ne_emb, *ne_table;         /* neighbor_embedding.embedding [max_z,H], filter table */
    const float *ne_comb_w, *ne_comb_b;     /* combine [H,2H], [H] */
    const float* ne_comb_w2T;               /* (combine.weight[:, H:])^T [H,H] */
    tmdn_gn_layer layer[TMDN_MAX_LAYERS];
    const float *head_w1, *head_b1, *head_w1T;   /* output_network.0 [H/2,H], [H/2], [H,H/2] */
    const float *head_w2, *head_b2;              /* output_network.2 [1,H/2], [1] */
} tmdn_gn_model;

size_t tmdn_gn_workspace_bytes(const tmdn_gn_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol, int strategy,
                               const double box[9]);
/* Same contract as tmdn_tensornet_energy_forces (status, NaN poisoning, sorted batch, max_edges incl. one self entry
 * per atom: (max_num_neighbors + 1) * n_atoms covers the reference's per-atom budget). */
int tmdn_gn_energy_forces(const tmdn_gn_model* m, const int64_t* z, const float* pos, const int64_t* batch,
                          int64_t n_atoms, int64_t n_mol, const double box[9], int use_periodic, int strategy,
                          int64_t max_edges, float* energy, float* forces, int32_t* status, void* workspace,
                          size_t workspace_bytes, tmdn_stream_t stream);
int tmdn_gn_workspace_lookup(const tmdn_gn_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol, int strategy,
                             const double box[9], const char* name, size_t* offset, size_t* bytes);

/* ---------------------------------------------------------------- Transformer (TorchMD_T) energy + forces (fp32)
 * TorchMD_T.forward + NeighborEmbedding + MultiHeadAttention + Scalar head + force path
 * (torchmdnet/models/torchmd_t.py:144-300): the attention of the equivariant transformer without its vector channel;
 * self pairs are part of the attention (Distance(..., loop=True), torchmd_t.py:102-104).
 *   layer table [n_int][2][H] float4 : act(dk_proj(rbf(d))), act(dv_proj(rbf(d))) * C(d)   (a missing filter is 1, resp. C)
 * qkv = q | k | v rows of one [3H,H] projection.  head_dim = hidden/num_heads must be a power of two <= 32. */
typedef struct tmdn_t_layer {
    const float *ln_w, *ln_b;
    const float *qkv_w, *qkv_b, *qkv_wT;    /* [3H,H], [3H], [H,3H] */
    const float *o_w, *o_b, *o_wT;          /* o_proj [H,H], [H], [H,H] */
    const float* table;
} tmdn_t_layer;

typedef struct tmdn_t_model {
    int32_t hidden, num_layers, num_heads, max_z;
    int32_t activation, attn_activation, has_neighbor_embedding, n_int;
    int32_t gemm_mode;
    float cutoff_lower, cutoff_upper, mean, std;
    float grid_alpha, grid_gamma;
    const float* emb;
    const float *ne_emb, *ne_table, *ne_comb_w, *ne_comb_b, *ne_comb_w2T;
    tmdn_t_layer layer[TMDN_MAX_LAYERS];
    const float *out_ln_w, *out_ln_b;
    const float *head_w1, *head_b1, *head_w1T, *head_w2, *head_b2;   /* Scalar head, as tmdn_gn_model */
} tmdn_t_model;

size_t tmdn_t_workspace_bytes(const tmdn_t_model* m, int64_t n_atoms, int64_t max_edges, int64_t n_mol, int strategy,
                              const double box[9]);
/* Same contract as tmdn_tensornet_energy_forces. */
int tmdn_t_energy_forces(const tmdn_t_model* m, const int64_t* z, const float* pos, const int64_t* batch, int64_t n_atoms,
                         int64_t n_mol, const double box[9], int use_periodic, int strategy, int64_t max_edges,
                         float* energy, float* forces, int32_t* status, void* workspace, size_t workspace_bytes,
                         tmdn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TMDN_H */
