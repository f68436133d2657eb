// torch_shim.cpp -- thin torch C++ extension over the C ABI of libtmdn.so (include/tmdn.h).
//
// Registers, with the reference's namespace / name / schema (torchmdnet/neighbors/neighbors.cpp:3-5):
//     torchmdnet_neighbors::get_neighbor_pairs            (AutogradCUDA; CPU tensors are refused: no CPU fallback)
// the reference's stream-capture probe as a compiled op (torchmdnet/models/utils.py:533-555):
//     torch_extension::is_stream_capturing
// and the fused model entry points:
//     torchmdnet_b200::csr_from_pairs, ::tensornet_create, ::tensornet_energy_forces, ::tensornet_debug_lookup
// Nothing here computes: tensors are validated, outputs/workspaces are taken from the caching allocator, and the
// C ABI is called on the current stream.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/autograd.h>
#include <torch/library.h>

#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "../../include/tmdn.h"

using at::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::tensor_list;

namespace {

#define TMDN_CALL(expr)                                                              \
    do {                                                                             \
        int rc__ = (expr);                                                           \
        TORCH_CHECK(rc__ == TMDN_OK, tmdn_last_error(), " (tmdn status ", rc__, ")"); \
    } while (0)

tmdn_stream_t cur_stream() { return (tmdn_stream_t)c10::cuda::getCurrentCUDAStream().stream(); }

int strategy_id(const std::string& s) {
    if (s == "brute") return TMDN_BRUTE;
    if (s == "shared") return TMDN_SHARED;
    if (s == "cell") return TMDN_CELL;
    throw std::runtime_error("Unknown kernel name");  // neighbors_cuda.cu:17
}

// box_vectors: CPU tensor [3,3] (or empty when not periodic) -> 9 host doubles. Returns false if absent.
bool read_box(const Tensor& box, double out[9]) {
    if (!box.defined() || box.numel() == 0) return false;
    TORCH_CHECK(box.device().is_cpu(), "Expected \"box_vectors\" to be on CPU");
    TORCH_CHECK(box.dim() == 2 && box.size(0) == 3 && box.size(1) == 3, "Expected \"box_vectors\" to have shape (3, 3)");
    const Tensor b = box.to(at::kDouble).contiguous();
    const double* p = b.data_ptr<double>();
    for (int i = 0; i < 9; ++i) out[i] = p[i];
    return true;
}

void check_input(const Tensor& positions, const Tensor& batch) {  // neighbors/common.cuh:118-131
    TORCH_CHECK(positions.is_cuda(),
                "torchmdnet_neighbors::get_neighbor_pairs: this build runs on CUDA (sm_100a) only; got a ",
                positions.device().type(), " tensor and there is no CPU fallback");
    TORCH_CHECK(positions.dim() == 2, "Expected \"positions\" to have two dimensions");
    TORCH_CHECK(positions.size(0) > 0, "Expected the 1nd dimension size of \"positions\" to be more than 0");
    TORCH_CHECK(positions.size(1) == 3, "Expected the 2nd dimension size of \"positions\" to be 3");
    TORCH_CHECK(positions.is_contiguous(), "Expected \"positions\" to be contiguous");
    TORCH_CHECK(batch.dim() == 1, "Expected \"batch\" to have one dimension");
    TORCH_CHECK(batch.size(0) == positions.size(0),
                "Expected the 1st dimension size of \"batch\" to be the same as the 1st dimension size of \"positions\"");
    TORCH_CHECK(batch.is_contiguous(), "Expected \"batch\" to be contiguous");
    TORCH_CHECK(batch.dtype() == at::kLong, "Expected \"batch\" to be of type torch::kLong");
    TORCH_CHECK(batch.device() == positions.device(), "Expected \"batch\" on the device of \"positions\"");
    TORCH_CHECK(positions.scalar_type() == at::kFloat || positions.scalar_type() == at::kDouble,
                "Expected \"positions\" to be float32 or float64");
}

std::vector<Tensor> neighbors_forward(const std::string& strategy, const Tensor& positions, const Tensor& batch,
                                      const Tensor& box_vectors, bool use_periodic, double lo, double hi,
                                      int64_t max_pairs, bool loop, bool include_transpose) {
    check_input(positions, batch);
    TORCH_CHECK(max_pairs > 0, "Expected \"max_num_neighbors\" to be positive");
    TORCH_CHECK(hi > 0, "Expected \"cutoff\" to be positive");
    const int sid = strategy_id(strategy);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TORCH_CHECK(have_box || !(use_periodic || sid == TMDN_CELL), "Expected \"box_vectors\" to have two dimensions");
    const c10::cuda::CUDAGuard guard(positions.device());
    const int64_t n = positions.size(0);
    const auto opts = positions.options();
    Tensor neighbors = at::empty({2, max_pairs}, opts.dtype(at::kInt));
    Tensor deltas = at::empty({max_pairs, 3}, opts);
    Tensor distances = at::empty({max_pairs}, opts);
    Tensor num = at::empty({1}, opts.dtype(at::kInt));
    const bool f64 = positions.scalar_type() == at::kDouble;
    const size_t ws_bytes = tmdn_neighbors_workspace_bytes(sid, n, have_box ? box : nullptr, hi, f64 ? 8 : 4);
    TORCH_CHECK(ws_bytes > 0, sid == TMDN_CELL ? "Too many cells in one dimension. Maximum is 1024" : "workspace");
    Tensor ws = at::empty({(int64_t)ws_bytes}, opts.dtype(at::kByte));
    if (f64)
        TMDN_CALL(tmdn_neighbors_f64(sid, positions.data_ptr<double>(), batch.data_ptr<int64_t>(), n,
                                     have_box ? box : nullptr, use_periodic, lo, hi, loop, include_transpose, max_pairs,
                                     neighbors.data_ptr<int32_t>(), deltas.data_ptr<double>(),
                                     distances.data_ptr<double>(), num.data_ptr<int32_t>(), ws.data_ptr(), ws_bytes,
                                     cur_stream()));
    else
        TMDN_CALL(tmdn_neighbors_f32(sid, positions.data_ptr<float>(), batch.data_ptr<int64_t>(), n,
                                     have_box ? box : nullptr, use_periodic, lo, hi, loop, include_transpose, max_pairs,
                                     neighbors.data_ptr<int32_t>(), deltas.data_ptr<float>(),
                                     distances.data_ptr<float>(), num.data_ptr<int32_t>(), ws.data_ptr(), ws_bytes,
                                     cur_stream()));
    return {neighbors, deltas, distances, num};
}

// dL/dpos from dL/d(deltas), dL/d(distances).  First-order use (inference, forces) runs the CUDA kernel; when the
// backward itself must be differentiable (create_graph=True, training on forces) it is composed of ATen ops.
Tensor neighbors_backward(const Tensor& neighbors, const Tensor& deltas, const Tensor& distances, Tensor grad_deltas,
                          Tensor grad_distances, int64_t num_atoms) {
    const c10::cuda::CUDAGuard guard(deltas.device());
    if (at::GradMode::is_enabled()) {
        const Tensor keep = (distances != 0).to(deltas.dtype());
        const Tensor safe_d = distances + (1 - keep);
        Tensor g = at::zeros_like(deltas);
        if (grad_deltas.defined()) g = g + grad_deltas * keep.unsqueeze(1);
        if (grad_distances.defined()) g = g + deltas / safe_d.unsqueeze(1) * (grad_distances * keep).unsqueeze(1);
        const Tensor idx = neighbors.to(at::kLong).clamp_min(0);
        Tensor gp = at::zeros({num_atoms, 3}, deltas.options());
        gp = gp.index_add(0, idx[0], g).index_add(0, idx[1], -g);
        return gp;
    }
    Tensor gp = at::empty({num_atoms, 3}, deltas.options());
    const int64_t P = distances.size(0);
    if (grad_deltas.defined()) grad_deltas = grad_deltas.contiguous();
    if (grad_distances.defined()) grad_distances = grad_distances.contiguous();
    if (deltas.scalar_type() == at::kDouble)
        TMDN_CALL(tmdn_neighbors_backward_f64(neighbors.data_ptr<int32_t>(), deltas.data_ptr<double>(),
                                              distances.data_ptr<double>(),
                                              grad_deltas.defined() ? grad_deltas.data_ptr<double>() : nullptr,
                                              grad_distances.defined() ? grad_distances.data_ptr<double>() : nullptr, P,
                                              num_atoms, gp.data_ptr<double>(), cur_stream()));
    else
        TMDN_CALL(tmdn_neighbors_backward_f32(neighbors.data_ptr<int32_t>(), deltas.data_ptr<float>(),
                                              distances.data_ptr<float>(),
                                              grad_deltas.defined() ? grad_deltas.data_ptr<float>() : nullptr,
                                              grad_distances.defined() ? grad_distances.data_ptr<float>() : nullptr, P,
                                              num_atoms, gp.data_ptr<float>(), cur_stream()));
    return gp;
}

class NeighborFn : public torch::autograd::Function<NeighborFn> {
public:
    static tensor_list forward(AutogradContext* ctx, const std::string& strategy, const Tensor& positions,
                               const Tensor& batch, const Tensor& box_vectors, bool use_periodic, double lo, double hi,
                               int64_t max_pairs, bool loop, bool include_transpose) {
        auto out = neighbors_forward(strategy, positions, batch, box_vectors, use_periodic, lo, hi, max_pairs, loop,
                                     include_transpose);
        ctx->save_for_backward({out[0], out[1], out[2]});
        ctx->saved_data["num_atoms"] = positions.size(0);
        ctx->mark_non_differentiable({out[0], out[3]});
        return out;
    }
    static tensor_list backward(AutogradContext* ctx, tensor_list grads) {
        const auto saved = ctx->get_saved_variables();
        const int64_t n = ctx->saved_data["num_atoms"].toInt();
        Tensor gp = neighbors_backward(saved[0], saved[1], saved[2], grads[1], grads[2], n);
        Tensor none;
        return {none, gp, none, none, none, none, none, none, none, none};
    }
};

std::tuple<Tensor, Tensor, Tensor, Tensor> get_neighbor_pairs_autograd(
    const std::string& strategy, const Tensor& positions, const Tensor& batch, const Tensor& box_vectors,
    bool use_periodic, const at::Scalar& cutoff_lower, const at::Scalar& cutoff_upper, const at::Scalar& max_num_pairs,
    bool loop, bool include_transpose) {
    auto r = NeighborFn::apply(strategy, positions, batch, box_vectors, use_periodic, cutoff_lower.toDouble(),
                               cutoff_upper.toDouble(), max_num_pairs.toLong(), loop, include_transpose);
    return std::make_tuple(r[0], r[1], r[2], r[3]);
}

std::tuple<Tensor, Tensor, Tensor, Tensor> get_neighbor_pairs_cpu(const std::string&, const Tensor& positions,
                                                                  const Tensor&, const Tensor&, bool, const at::Scalar&,
                                                                  const at::Scalar&, const at::Scalar&, bool, bool) {
    TORCH_CHECK(false, "torchmdnet_neighbors::get_neighbor_pairs: this build runs on CUDA (sm_100a) only; got a ",
                positions.device().type(), " tensor and there is no CPU fallback");
    return {};
}

bool is_stream_capturing() {
    const int r = tmdn_is_stream_capturing(cur_stream());
    TORCH_CHECK(r >= 0, tmdn_last_error());
    return r == 1;
}

std::tuple<Tensor, Tensor, Tensor> csr_from_pairs(const Tensor& neighbors, const Tensor& deltas,
                                                  const Tensor& distances, const Tensor& num_found, int64_t n_atoms,
                                                  int64_t target_row) {
    TORCH_CHECK(neighbors.is_cuda() && neighbors.scalar_type() == at::kInt && neighbors.is_contiguous(),
                "csr_from_pairs: neighbors must be a contiguous CUDA int32 tensor");
    TORCH_CHECK(deltas.scalar_type() == at::kFloat && distances.scalar_type() == at::kFloat,
                "csr_from_pairs: fp32 only");
    const c10::cuda::CUDAGuard guard(neighbors.device());
    const int64_t P = neighbors.size(1);
    const auto opts = neighbors.options();
    Tensor row_ptr = at::empty({n_atoms + 1}, opts);
    Tensor col = at::full({P}, -1, opts);
    Tensor geom = at::zeros({P, 4}, deltas.options());
    const size_t ws_bytes = tmdn_csr_workspace_bytes(n_atoms, P);
    Tensor ws = at::empty({(int64_t)ws_bytes}, opts.dtype(at::kByte));
    TMDN_CALL(tmdn_csr_from_pairs(neighbors.data_ptr<int32_t>(), deltas.contiguous().data_ptr<float>(),
                                  distances.contiguous().data_ptr<float>(), num_found.data_ptr<int32_t>(), P, n_atoms,
                                  (int)target_row, row_ptr.data_ptr<int32_t>(), col.data_ptr<int32_t>(),
                                  geom.data_ptr<float>(), ws.data_ptr(), ws_bytes, cur_stream()));
    return std::make_tuple(row_ptr, col, geom);
}

// -> (row_ptr int32[N+1], col int32[max_edges], geom float32[max_edges,4], num_found int32[1])
std::tuple<Tensor, Tensor, Tensor, Tensor> csr_build(const std::string& strategy, const Tensor& positions,
                                                     const Tensor& batch, const Tensor& box_vectors, bool use_periodic,
                                                     double lo, double hi, int64_t max_edges) {
    check_input(positions, batch);
    TORCH_CHECK(positions.scalar_type() == at::kFloat, "csr_build: fp32 positions only");
    const int sid = strategy_id(strategy);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TORCH_CHECK(have_box || !(use_periodic || sid == TMDN_CELL), "csr_build: box required");
    const c10::cuda::CUDAGuard guard(positions.device());
    const int64_t n = positions.size(0);
    const auto iopts = positions.options().dtype(at::kInt);
    Tensor row_ptr = at::empty({n + 1}, iopts);
    Tensor col = at::full({max_edges}, -1, iopts);
    Tensor geom = at::zeros({max_edges, 4}, positions.options());
    Tensor num = at::empty({1}, iopts);
    const size_t ws_bytes = tmdn_csr_build_workspace_bytes(sid, n, have_box ? box : nullptr, hi);
    TORCH_CHECK(ws_bytes > 0, "csr_build: bad box for the cell strategy");
    Tensor ws = at::empty({(int64_t)ws_bytes}, positions.options().dtype(at::kByte));
    TMDN_CALL(tmdn_csr_build_f32(sid, positions.data_ptr<float>(), batch.data_ptr<int64_t>(), n,
                                 have_box ? box : nullptr, use_periodic, lo, hi, max_edges, row_ptr.data_ptr<int32_t>(),
                                 col.data_ptr<int32_t>(), geom.data_ptr<float>(), num.data_ptr<int32_t>(),
                                 ws.data_ptr(), ws_bytes, cur_stream()));
    return std::make_tuple(row_ptr, col, geom, num);
}

// ------------------------------------------------------------------ TensorNet model handles
struct TnHandle {
    tmdn_tensornet_model m;
    std::vector<Tensor> keep;  // keeps the packed weights alive
};
std::mutex g_mu;
std::unordered_map<int64_t, std::shared_ptr<TnHandle>> g_tn;
int64_t g_next = 1;

const float* fptr(const Tensor& t) {
    if (!t.defined() || t.numel() == 0) return nullptr;
    TORCH_CHECK(t.is_cuda() && t.scalar_type() == at::kFloat && t.is_contiguous(),
                "tensornet_create: weights must be contiguous CUDA float32 tensors");
    return t.data_ptr<float>();
}

// Weight order (must match torchmdnet/models/tensornet.py::_pack):
//   rbf_p0 rbf_p1 P Q | emb_filter: w b table | emb_mix x3 emb_mixT x3 | emb_ln_w emb_ln_b
//   emb_w1 emb_b1 emb_w1T emb_w2 emb_b2 emb_w2T | out_ln_w out_ln_b out_w out_b out_wT
//   head_w1 head_b1 head_w1T head_w2 head_b2 | per layer: w0 b0 w1 b1 w2 b2 table mix x6 mixT x6
// iparams: hidden num_rbf num_layers max_z rbf_type activation so3 filter_mode n_int gemm_mode ;
// fparams: lo hi mean std grid_alpha grid_gamma
int64_t tensornet_create(std::vector<Tensor> w, std::vector<int64_t> ip, std::vector<double> fp) {
    TORCH_CHECK(ip.size() == 10 && fp.size() == 6, "tensornet_create: bad parameter vectors");
    auto h = std::make_shared<TnHandle>();
    tmdn_tensornet_model& m = h->m;
    memset(&m, 0, sizeof(m));
    m.hidden = (int32_t)ip[0]; m.num_rbf = (int32_t)ip[1]; m.num_layers = (int32_t)ip[2]; m.max_z = (int32_t)ip[3];
    m.rbf_type = (int32_t)ip[4]; m.activation = (int32_t)ip[5]; m.so3 = (int32_t)ip[6]; m.filter_mode = (int32_t)ip[7];
    m.gemm_mode = (int32_t)ip[9];
    const int32_t n_int = (int32_t)ip[8];
    m.cutoff_lower = (float)fp[0]; m.cutoff_upper = (float)fp[1]; m.mean = (float)fp[2]; m.std = (float)fp[3];
    TORCH_CHECK(m.num_layers >= 0 && m.num_layers <= TMDN_MAX_LAYERS, "tensornet_create: too many layers");
    const size_t expect = 31 + 19 * (size_t)m.num_layers;
    TORCH_CHECK(w.size() == expect, "tensornet_create: expected ", expect, " weight tensors, got ", w.size());
    const int H = m.hidden, K = m.num_rbf;
    size_t k = 0;
    auto next = [&]() -> const float* { return fptr(w[k++]); };
    m.rbf_p0 = next(); m.rbf_p1 = next(); m.P = next(); m.Q = next();
    m.emb_filter.n_layers = 1; m.emb_filter.final_act = 0;
    m.emb_filter.dims[0] = K; m.emb_filter.dims[1] = 3 * H;
    m.emb_filter.w[0] = next(); m.emb_filter.b[0] = next(); m.emb_filter.table = next(); m.emb_filter.n_int = n_int;
    m.emb_filter.grid_alpha = (float)fp[4]; m.emb_filter.grid_gamma = (float)fp[5];
    for (int g = 0; g < 3; ++g) m.emb_mix[g] = next();
    for (int g = 0; g < 3; ++g) m.emb_mixT[g] = next();
    m.emb_ln_w = next(); m.emb_ln_b = next();
    m.emb_w1 = next(); m.emb_b1 = next(); m.emb_w1T = next();
    m.emb_w2 = next(); m.emb_b2 = next(); m.emb_w2T = next();
    m.out_ln_w = next(); m.out_ln_b = next(); m.out_w = next(); m.out_b = next(); m.out_wT = next();
    m.head_w1 = next(); m.head_b1 = next(); m.head_w1T = next(); m.head_w2 = next(); m.head_b2 = next();
    for (int l = 0; l < m.num_layers; ++l) {
        tmdn_filter_set& f = m.layer_filter[l];
        f.n_layers = 3; f.final_act = 1;
        f.dims[0] = K; f.dims[1] = H; f.dims[2] = 2 * H; f.dims[3] = 3 * H;
        for (int j = 0; j < 3; ++j) { f.w[j] = next(); f.b[j] = next(); }
        f.table = next(); f.n_int = n_int; f.grid_alpha = (float)fp[4]; f.grid_gamma = (float)fp[5];
        for (int j = 0; j < 6; ++j) m.layer_mix[l][j] = next();
        for (int j = 0; j < 6; ++j) m.layer_mixT[l][j] = next();
    }
    h->keep = std::move(w);
    std::lock_guard<std::mutex> lock(g_mu);
    const int64_t id = g_next++;
    g_tn[id] = h;
    return id;
}

void tensornet_destroy(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_tn.erase(id);
}

std::shared_ptr<TnHandle> tn_get(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_tn.find(id);
    TORCH_CHECK(it != g_tn.end(), "tensornet: unknown model handle ", id);
    return it->second;
}

// -> (energy[n_mol,1], forces[n_atoms,3] (empty if !compute_forces), status int32[4], workspace uint8)
std::tuple<Tensor, Tensor, Tensor, Tensor> tensornet_energy_forces(int64_t handle, const Tensor& z, const Tensor& pos,
                                                                   const Tensor& batch, const Tensor& box_vectors,
                                                                   bool use_periodic, const std::string& strategy,
                                                                   int64_t max_edges, int64_t n_mol,
                                                                   bool compute_forces,
                                                                   const c10::optional<Tensor>& workspace) {
    auto h = tn_get(handle);
    TORCH_CHECK(pos.is_cuda(), "tensornet_energy_forces: CUDA tensors required (sm_100a build, no CPU fallback)");
    TORCH_CHECK(pos.scalar_type() == at::kFloat && pos.dim() == 2 && pos.size(1) == 3 && pos.is_contiguous(),
                "tensornet_energy_forces: pos must be a contiguous float32 [N,3] tensor");
    TORCH_CHECK(z.scalar_type() == at::kLong && z.dim() == 1 && z.is_contiguous() && z.size(0) == pos.size(0),
                "tensornet_energy_forces: z must be a contiguous int64 [N] tensor");
    TORCH_CHECK(batch.scalar_type() == at::kLong && batch.dim() == 1 && batch.is_contiguous() &&
                    batch.size(0) == pos.size(0),
                "tensornet_energy_forces: batch must be a contiguous int64 [N] tensor");
    TORCH_CHECK(z.device() == pos.device() && batch.device() == pos.device(), "tensornet_energy_forces: device mismatch");
    TORCH_CHECK(h->keep.empty() || h->keep[0].device() == pos.device(),
                "tensornet_energy_forces: the model was packed on ", h->keep[0].device(), " but pos is on ", pos.device());
    const c10::cuda::CUDAGuard guard(pos.device());
    const int64_t n = pos.size(0);
    const int sid = strategy_id(strategy);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TORCH_CHECK(have_box || !(use_periodic || sid == TMDN_CELL), "tensornet_energy_forces: box required");
    const size_t need = tmdn_tensornet_workspace_bytes(&h->m, n, max_edges, n_mol, sid, have_box ? box : nullptr);
    TORCH_CHECK(need > 0, "tensornet_energy_forces: ", tmdn_last_error());
    Tensor ws;
    if (workspace.has_value() && workspace->defined() && (size_t)workspace->numel() >= need && workspace->is_cuda() &&
        workspace->device() == pos.device() && workspace->scalar_type() == at::kByte)
        ws = *workspace;
    else
        ws = at::empty({(int64_t)need}, pos.options().dtype(at::kByte));
    Tensor energy = at::empty({n_mol, 1}, pos.options());
    Tensor forces = compute_forces ? at::empty({n, 3}, pos.options()) : at::empty({0, 3}, pos.options());
    Tensor status = at::empty({4}, pos.options().dtype(at::kInt));
    TMDN_CALL(tmdn_tensornet_energy_forces(&h->m, z.data_ptr<int64_t>(), pos.data_ptr<float>(),
                                           batch.data_ptr<int64_t>(), n, n_mol, have_box ? box : nullptr, use_periodic,
                                           sid, max_edges, energy.data_ptr<float>(),
                                           compute_forces ? forces.data_ptr<float>() : nullptr,
                                           status.data_ptr<int32_t>(), ws.data_ptr(), (size_t)ws.numel(), cur_stream()));
    return std::make_tuple(energy, forces, status, ws);
}

// -> [offset, bytes] of a named intermediate inside the workspace
std::vector<int64_t> tensornet_debug_lookup(int64_t handle, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                            const std::string& strategy, const Tensor& box_vectors,
                                            const std::string& name) {
    auto h = tn_get(handle);
    size_t off = 0, bytes = 0;
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TMDN_CALL(tmdn_tensornet_workspace_lookup(&h->m, n_atoms, max_edges, n_mol, strategy_id(strategy),
                                              have_box ? box : nullptr, name.c_str(), &off, &bytes));
    return {(int64_t)off, (int64_t)bytes};
}

// ------------------------------------------------------------------ Equivariant Transformer model handles
struct EtHandle {
    tmdn_et_model m;
    std::vector<Tensor> keep;
};
std::unordered_map<int64_t, std::shared_ptr<EtHandle>> g_et;

// Weight order (must match torchmdnet/models/torchmd_et.py::pack_weights):
//   emb ne_emb ne_table ne_comb_w ne_comb_b ne_comb_w2T out_ln_w out_ln_b
//   head block x2: w1 w1T w2 w2T u0_w u0_b u0_wT u2_w u2_b u2_wT
//   per layer: ln_w ln_b qkv_w qkv_b qkv_wT vp_w vp_wT o_w o_b o_wT table
// iparams: hidden num_layers num_heads max_z activation attn_activation has_ne n_int gemm_mode ;
// fparams: lo hi mean std grid_alpha grid_gamma
int64_t et_create(std::vector<Tensor> w, std::vector<int64_t> ip, std::vector<double> fp) {
    TORCH_CHECK(ip.size() == 9 && fp.size() == 6, "et_create: bad parameter vectors");
    auto h = std::make_shared<EtHandle>();
    tmdn_et_model& m = h->m;
    memset(&m, 0, sizeof(m));
    m.hidden = (int32_t)ip[0]; m.num_layers = (int32_t)ip[1]; m.num_heads = (int32_t)ip[2]; m.max_z = (int32_t)ip[3];
    m.activation = (int32_t)ip[4]; m.attn_activation = (int32_t)ip[5]; m.has_neighbor_embedding = (int32_t)ip[6];
    m.n_int = (int32_t)ip[7];
    m.gemm_mode = (int32_t)ip[8];
    m.cutoff_lower = (float)fp[0]; m.cutoff_upper = (float)fp[1]; m.mean = (float)fp[2]; m.std = (float)fp[3];
    m.grid_alpha = (float)fp[4]; m.grid_gamma = (float)fp[5];
    TORCH_CHECK(m.num_layers >= 0 && m.num_layers <= TMDN_MAX_LAYERS, "et_create: too many layers");
    const size_t expect = 28 + 11 * (size_t)m.num_layers;
    TORCH_CHECK(w.size() == expect, "et_create: expected ", expect, " weight tensors, got ", w.size());
    size_t k = 0;
    auto next = [&]() -> const float* { return fptr(w[k++]); };
    m.emb = next(); m.ne_emb = next(); m.ne_table = next(); m.ne_comb_w = next(); m.ne_comb_b = next();
    m.ne_comb_w2T = next(); m.out_ln_w = next(); m.out_ln_b = next();
    for (int b = 0; b < 2; ++b) {
        tmdn_et_gated& g = m.head[b];
        g.hin = b == 0 ? m.hidden : m.hidden / 2;
        g.out = b == 0 ? m.hidden / 2 : 1;
        g.w1 = next(); g.w1T = next(); g.w2 = next(); g.w2T = next();
        g.u0_w = next(); g.u0_b = next(); g.u0_wT = next(); g.u2_w = next(); g.u2_b = next(); g.u2_wT = next();
    }
    for (int l = 0; l < m.num_layers; ++l) {
        tmdn_et_layer& y = m.layer[l];
        y.ln_w = next(); y.ln_b = next(); y.qkv_w = next(); y.qkv_b = next(); y.qkv_wT = next();
        y.vp_w = next(); y.vp_wT = next(); y.o_w = next(); y.o_b = next(); y.o_wT = next(); y.table = next();
    }
    h->keep = std::move(w);
    std::lock_guard<std::mutex> lock(g_mu);
    const int64_t id = g_next++;
    g_et[id] = h;
    return id;
}
void et_destroy(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_et.erase(id);
}
std::shared_ptr<EtHandle> et_get(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_et.find(id);
    TORCH_CHECK(it != g_et.end(), "et: unknown model handle ", id);
    return it->second;
}

std::tuple<Tensor, Tensor, Tensor, Tensor> et_energy_forces(int64_t handle, const Tensor& z, const Tensor& pos,
                                                            const Tensor& batch, const Tensor& box_vectors,
                                                            bool use_periodic, const std::string& strategy,
                                                            int64_t max_edges, int64_t n_mol, bool compute_forces,
                                                            const c10::optional<Tensor>& workspace) {
    auto h = et_get(handle);
    TORCH_CHECK(pos.is_cuda(), "et_energy_forces: CUDA tensors required (sm_100a build, no CPU fallback)");
    TORCH_CHECK(pos.scalar_type() == at::kFloat && pos.dim() == 2 && pos.size(1) == 3 && pos.is_contiguous(),
                "et_energy_forces: pos must be a contiguous float32 [N,3] tensor");
    TORCH_CHECK(z.scalar_type() == at::kLong && z.dim() == 1 && z.is_contiguous() && z.size(0) == pos.size(0),
                "et_energy_forces: z must be a contiguous int64 [N] tensor");
    TORCH_CHECK(batch.scalar_type() == at::kLong && batch.dim() == 1 && batch.is_contiguous() &&
                    batch.size(0) == pos.size(0),
                "et_energy_forces: batch must be a contiguous int64 [N] tensor");
    TORCH_CHECK(z.device() == pos.device() && batch.device() == pos.device(), "et_energy_forces: device mismatch");
    TORCH_CHECK(h->keep.empty() || h->keep[0].device() == pos.device(),
                "et_energy_forces: the model was packed on ", h->keep[0].device(), " but pos is on ", pos.device());
    const c10::cuda::CUDAGuard guard(pos.device());
    const int64_t n = pos.size(0);
    const int sid = strategy_id(strategy);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TORCH_CHECK(have_box || !(use_periodic || sid == TMDN_CELL), "et_energy_forces: box required");
    const size_t need = tmdn_et_workspace_bytes(&h->m, n, max_edges, n_mol, sid, have_box ? box : nullptr);
    TORCH_CHECK(need > 0, "et_energy_forces: ", tmdn_last_error());
    Tensor ws;
    if (workspace.has_value() && workspace->defined() && (size_t)workspace->numel() >= need && workspace->is_cuda() &&
        workspace->device() == pos.device() && workspace->scalar_type() == at::kByte)
        ws = *workspace;
    else
        ws = at::empty({(int64_t)need}, pos.options().dtype(at::kByte));
    Tensor energy = at::empty({n_mol, 1}, pos.options());
    Tensor forces = compute_forces ? at::empty({n, 3}, pos.options()) : at::empty({0, 3}, pos.options());
    Tensor status = at::empty({4}, pos.options().dtype(at::kInt));
    TMDN_CALL(tmdn_et_energy_forces(&h->m, z.data_ptr<int64_t>(), pos.data_ptr<float>(), batch.data_ptr<int64_t>(), n,
                                    n_mol, have_box ? box : nullptr, use_periodic, sid, max_edges,
                                    energy.data_ptr<float>(), compute_forces ? forces.data_ptr<float>() : nullptr,
                                    status.data_ptr<int32_t>(), ws.data_ptr(), (size_t)ws.numel(), cur_stream()));
    return std::make_tuple(energy, forces, status, ws);
}

std::vector<int64_t> et_debug_lookup(int64_t handle, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                     const std::string& strategy, const Tensor& box_vectors, const std::string& name) {
    auto h = et_get(handle);
    size_t off = 0, bytes = 0;
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TMDN_CALL(tmdn_et_workspace_lookup(&h->m, n_atoms, max_edges, n_mol, strategy_id(strategy),
                                       have_box ? box : nullptr, name.c_str(), &off, &bytes));
    return {(int64_t)off, (int64_t)bytes};
}

// ------------------------------------------------------------------ graph network (TorchMD_GN) model handles
struct GnHandle {
    tmdn_gn_model m;
    std::vector<Tensor> keep;
};
std::unordered_map<int64_t, std::shared_ptr<GnHandle>> g_gn;

// Weight order (must match torchmdnet/models/torchmd_gn.py::pack_weights):
//   emb ne_emb ne_table ne_comb_w ne_comb_b ne_comb_w2T head_w1 head_b1 head_w1T head_w2 head_b2
//   per layer: lin1_w lin1_wT table lin2_w lin2_b lin2_wT lin_w lin_b lin_wT
// iparams: hidden num_filters num_layers max_z activation has_ne n_int gemm_mode ; fparams: lo hi mean std grid_alpha grid_gamma
int64_t gn_create(std::vector<Tensor> w, std::vector<int64_t> ip, std::vector<double> fp) {
    TORCH_CHECK(ip.size() == 8 && fp.size() == 6, "gn_create: bad parameter vectors");
    auto h = std::make_shared<GnHandle>();
    tmdn_gn_model& m = h->m;
    memset(&m, 0, sizeof(m));
    m.hidden = (int32_t)ip[0]; m.num_filters = (int32_t)ip[1]; m.num_layers = (int32_t)ip[2]; m.max_z = (int32_t)ip[3];
    m.activation = (int32_t)ip[4]; m.has_neighbor_embedding = (int32_t)ip[5]; m.n_int = (int32_t)ip[6];
    m.gemm_mode = (int32_t)ip[7];
    m.cutoff_lower = (float)fp[0]; m.cutoff_upper = (float)fp[1]; m.mean = (float)fp[2]; m.std = (float)fp[3];
    m.grid_alpha = (float)fp[4]; m.grid_gamma = (float)fp[5];
    TORCH_CHECK(m.num_layers >= 0 && m.num_layers <= TMDN_MAX_LAYERS, "gn_create: too many layers");
    const size_t expect = 11 + 9 * (size_t)m.num_layers;
    TORCH_CHECK(w.size() == expect, "gn_create: expected ", expect, " weight tensors, got ", w.size());
    size_t k = 0;
    auto next = [&]() -> const float* { return fptr(w[k++]); };
    m.emb = next(); m.ne_emb = next(); m.ne_table = next(); m.ne_comb_w = next(); m.ne_comb_b = next();
    m.ne_comb_w2T = next();
    m.head_w1 = next(); m.head_b1 = next(); m.head_w1T = next(); m.head_w2 = next(); m.head_b2 = next();
    for (int l = 0; l < m.num_layers; ++l) {
        tmdn_gn_layer& y = m.layer[l];
        y.lin1_w = next(); y.lin1_wT = next(); y.table = next(); y.lin2_w = next(); y.lin2_b = next(); y.lin2_wT = next();
        y.lin_w = next(); y.lin_b = next(); y.lin_wT = next();
    }
    h->keep = std::move(w);
    std::lock_guard<std::mutex> lock(g_mu);
    const int64_t id = g_next++;
    g_gn[id] = h;
    return id;
}
void gn_destroy(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_gn.erase(id);
}
std::shared_ptr<GnHandle> gn_get(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_gn.find(id);
    TORCH_CHECK(it != g_gn.end(), "gn: unknown model handle ", id);
    return it->second;
}

std::tuple<Tensor, Tensor, Tensor, Tensor> gn_energy_forces(int64_t handle, const Tensor& z, const Tensor& pos,
                                                            const Tensor& batch, const Tensor& box_vectors,
                                                            bool use_periodic, const std::string& strategy,
                                                            int64_t max_edges, int64_t n_mol, bool compute_forces,
                                                            const c10::optional<Tensor>& workspace) {
    auto h = gn_get(handle);
    TORCH_CHECK(pos.is_cuda(), "gn_energy_forces: CUDA tensors required (sm_100a build, no CPU fallback)");
    TORCH_CHECK(pos.scalar_type() == at::kFloat && pos.dim() == 2 && pos.size(1) == 3 && pos.is_contiguous(),
                "gn_energy_forces: pos must be a contiguous float32 [N,3] tensor");
    TORCH_CHECK(z.scalar_type() == at::kLong && z.dim() == 1 && z.is_contiguous() && z.size(0) == pos.size(0),
                "gn_energy_forces: z must be a contiguous int64 [N] tensor");
    TORCH_CHECK(batch.scalar_type() == at::kLong && batch.dim() == 1 && batch.is_contiguous() &&
                    batch.size(0) == pos.size(0),
                "gn_energy_forces: batch must be a contiguous int64 [N] tensor");
    TORCH_CHECK(z.device() == pos.device() && batch.device() == pos.device(), "gn_energy_forces: device mismatch");
    TORCH_CHECK(h->keep.empty() || h->keep[0].device() == pos.device(),
                "gn_energy_forces: the model was packed on ", h->keep[0].device(), " but pos is on ", pos.device());
    const c10::cuda::CUDAGuard guard(pos.device());
    const int64_t n = pos.size(0);
    const int sid = strategy_id(strategy);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TORCH_CHECK(have_box || !(use_periodic || sid == TMDN_CELL), "gn_energy_forces: box required");
    const size_t need = tmdn_gn_workspace_bytes(&h->m, n, max_edges, n_mol, sid, have_box ? box : nullptr);
    TORCH_CHECK(need > 0, "gn_energy_forces: ", tmdn_last_error());
    Tensor ws;
    if (workspace.has_value() && workspace->defined() && (size_t)workspace->numel() >= need && workspace->is_cuda() &&
        workspace->device() == pos.device() && workspace->scalar_type() == at::kByte)
        ws = *workspace;
    else
        ws = at::empty({(int64_t)need}, pos.options().dtype(at::kByte));
    Tensor energy = at::empty({n_mol, 1}, pos.options());
    Tensor forces = compute_forces ? at::empty({n, 3}, pos.options()) : at::empty({0, 3}, pos.options());
    Tensor status = at::empty({4}, pos.options().dtype(at::kInt));
    TMDN_CALL(tmdn_gn_energy_forces(&h->m, z.data_ptr<int64_t>(), pos.data_ptr<float>(), batch.data_ptr<int64_t>(), n,
                                    n_mol, have_box ? box : nullptr, use_periodic, sid, max_edges,
                                    energy.data_ptr<float>(), compute_forces ? forces.data_ptr<float>() : nullptr,
                                    status.data_ptr<int32_t>(), ws.data_ptr(), (size_t)ws.numel(), cur_stream()));
    return std::make_tuple(energy, forces, status, ws);
}

std::vector<int64_t> gn_debug_lookup(int64_t handle, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                     const std::string& strategy, const Tensor& box_vectors, const std::string& name) {
    auto h = gn_get(handle);
    size_t off = 0, bytes = 0;
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TMDN_CALL(tmdn_gn_workspace_lookup(&h->m, n_atoms, max_edges, n_mol, strategy_id(strategy),
                                       have_box ? box : nullptr, name.c_str(), &off, &bytes));
    return {(int64_t)off, (int64_t)bytes};
}

// ------------------------------------------------------------------ transformer (TorchMD_T) model handles
struct TfHandle {
    tmdn_t_model m;
    std::vector<Tensor> keep;
};
std::unordered_map<int64_t, std::shared_ptr<TfHandle>> g_tf;

// Weight order (must match torchmdnet/models/torchmd_t.py::pack_weights):
//   emb ne_emb ne_table ne_comb_w ne_comb_b ne_comb_w2T out_ln_w out_ln_b head_w1 head_b1 head_w1T head_w2 head_b2
//   per layer: ln_w ln_b qkv_w qkv_b qkv_wT o_w o_b o_wT table
// iparams: hidden num_layers num_heads max_z activation attn_activation has_ne n_int gemm_mode ;
// fparams: lo hi mean std grid_alpha grid_gamma
int64_t t_create(std::vector<Tensor> w, std::vector<int64_t> ip, std::vector<double> fp) {
    TORCH_CHECK(ip.size() == 9 && fp.size() == 6, "t_create: bad parameter vectors");
    auto h = std::make_shared<TfHandle>();
    tmdn_t_model& m = h->m;
    memset(&m, 0, sizeof(m));
    m.hidden = (int32_t)ip[0]; m.num_layers = (int32_t)ip[1]; m.num_heads = (int32_t)ip[2]; m.max_z = (int32_t)ip[3];
    m.activation = (int32_t)ip[4]; m.attn_activation = (int32_t)ip[5]; m.has_neighbor_embedding = (int32_t)ip[6];
    m.n_int = (int32_t)ip[7]; m.gemm_mode = (int32_t)ip[8];
    m.cutoff_lower = (float)fp[0]; m.cutoff_upper = (float)fp[1]; m.mean = (float)fp[2]; m.std = (float)fp[3];
    m.grid_alpha = (float)fp[4]; m.grid_gamma = (float)fp[5];
    TORCH_CHECK(m.num_layers >= 0 && m.num_layers <= TMDN_MAX_LAYERS, "t_create: too many layers");
    const size_t expect = 13 + 9 * (size_t)m.num_layers;
    TORCH_CHECK(w.size() == expect, "t_create: expected ", expect, " weight tensors, got ", w.size());
    size_t k = 0;
    auto next = [&]() -> const float* { return fptr(w[k++]); };
    m.emb = next(); m.ne_emb = next(); m.ne_table = next(); m.ne_comb_w = next(); m.ne_comb_b = next();
    m.ne_comb_w2T = next(); m.out_ln_w = next(); m.out_ln_b = next();
    m.head_w1 = next(); m.head_b1 = next(); m.head_w1T = next(); m.head_w2 = next(); m.head_b2 = next();
    for (int l = 0; l < m.num_layers; ++l) {
        tmdn_t_layer& y = m.layer[l];
        y.ln_w = next(); y.ln_b = next(); y.qkv_w = next(); y.qkv_b = next(); y.qkv_wT = next();
        y.o_w = next(); y.o_b = next(); y.o_wT = next(); y.table = next();
    }
    h->keep = std::move(w);
    std::lock_guard<std::mutex> lock(g_mu);
    const int64_t id = g_next++;
    g_tf[id] = h;
    return id;
}
void t_destroy(int64_t id) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_tf.erase(id);
}

std::tuple<Tensor, Tensor, Tensor, Tensor> t_energy_forces(int64_t handle, const Tensor& z, const Tensor& pos,
                                                           const Tensor& batch, const Tensor& box_vectors,
                                                           bool use_periodic, const std::string& strategy,
                                                           int64_t max_edges, int64_t n_mol, bool compute_forces,
                                                           const c10::optional<Tensor>& workspace) {
    std::shared_ptr<TfHandle> h;
    {
        std::lock_guard<std::mutex> lock(g_mu);
        auto it = g_tf.find(handle);
        TORCH_CHECK(it != g_tf.end(), "t: unknown model handle ", handle);
        h = it->second;
    }
    TORCH_CHECK(pos.is_cuda(), "t_energy_forces: CUDA tensors required (sm_100a build, no CPU fallback)");
    TORCH_CHECK(pos.scalar_type() == at::kFloat && pos.dim() == 2 && pos.size(1) == 3 && pos.is_contiguous(),
                "t_energy_forces: pos must be a contiguous float32 [N,3] tensor");
    TORCH_CHECK(z.scalar_type() == at::kLong && z.dim() == 1 && z.is_contiguous() && z.size(0) == pos.size(0),
                "t_energy_forces: z must be a contiguous int64 [N] tensor");
    TORCH_CHECK(batch.scalar_type() == at::kLong && batch.dim() == 1 && batch.is_contiguous() &&
                    batch.size(0) == pos.size(0),
                "t_energy_forces: batch must be a contiguous int64 [N] tensor");
    TORCH_CHECK(z.device() == pos.device() && batch.device() == pos.device(), "t_energy_forces: device mismatch");
    TORCH_CHECK(h->keep.empty() || h->keep[0].device() == pos.device(),
                "t_energy_forces: the model was packed on ", h->keep[0].device(), " but pos is on ", pos.device());
    const c10::cuda::CUDAGuard guard(pos.device());
    const int64_t n = pos.size(0);
    const int sid = strategy_id(strategy);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    TORCH_CHECK(have_box || !(use_periodic || sid == TMDN_CELL), "t_energy_forces: box required");
    const size_t need = tmdn_t_workspace_bytes(&h->m, n, max_edges, n_mol, sid, have_box ? box : nullptr);
    TORCH_CHECK(need > 0, "t_energy_forces: ", tmdn_last_error());
    Tensor ws;
    if (workspace.has_value() && workspace->defined() && (size_t)workspace->numel() >= need && workspace->is_cuda() &&
        workspace->device() == pos.device() && workspace->scalar_type() == at::kByte)
        ws = *workspace;
    else
        ws = at::empty({(int64_t)need}, pos.options().dtype(at::kByte));
    Tensor energy = at::empty({n_mol, 1}, pos.options());
    Tensor forces = compute_forces ? at::empty({n, 3}, pos.options()) : at::empty({0, 3}, pos.options());
    Tensor status = at::empty({4}, pos.options().dtype(at::kInt));
    TMDN_CALL(tmdn_t_energy_forces(&h->m, z.data_ptr<int64_t>(), pos.data_ptr<float>(), batch.data_ptr<int64_t>(), n, n_mol,
                                   have_box ? box : nullptr, use_periodic, sid, max_edges, energy.data_ptr<float>(),
                                   compute_forces ? forces.data_ptr<float>() : nullptr, status.data_ptr<int32_t>(),
                                   ws.data_ptr(), (size_t)ws.numel(), cur_stream()));
    return std::make_tuple(energy, forces, status, ws);
}

// ------------------------------------------------------------------ script-friendly entry points
// A scripted / saved / re-loaded model cannot carry a process-local handle.  It carries ONE flat fp32 buffer with every
// packed weight (segments aligned to 256 bytes; `offsets` has n_weights + 1 entries, in floats) plus the integer / float
// parameter vectors of *_create, and these ops find -- or create -- the native handle from the identity of that buffer
// (address, version counter, device).  The workspace is cached with the handle.  box: 9 floats, or empty.
struct PackedEntry { int64_t handle; Tensor workspace; };
struct PackedKeyT {
    const void* ptr; uint32_t version; int device; std::vector<int64_t> ip; std::vector<double> fp; int kind;
    bool operator<(const PackedKeyT& o) const {
        return std::tie(ptr, version, device, kind, ip, fp) < std::tie(o.ptr, o.version, o.device, o.kind, o.ip, o.fp);
    }
};
std::map<PackedKeyT, PackedEntry> g_packed;

// offsets = [o_0 .. o_n | len_0 .. len_{n-1}]: segment i starts at o_i (multiple of 64 floats) and holds len_i floats
PackedEntry& packed_entry(int kind, const Tensor& flat, const std::vector<int64_t>& layout, const std::vector<int64_t>& ip,
                          const std::vector<double>& fp) {
    TORCH_CHECK(layout.size() % 2 == 0 && !layout.empty(), "packed model: bad layout vector");
    PackedKeyT key{flat.data_ptr(), flat._version(), (int)flat.get_device(), ip, fp, kind};
    std::unique_lock<std::mutex> lock(g_mu);
    auto it = g_packed.find(key);
    if (it != g_packed.end()) return it->second;
    TORCH_CHECK(flat.is_cuda() && flat.scalar_type() == at::kFloat && flat.is_contiguous() && flat.dim() == 1,
                "packed model: the weight buffer must be a contiguous 1-D CUDA float32 tensor (sm_100a build, no CPU fallback)");
    const size_t n = layout.size() / 2;
    std::vector<Tensor> w;
    for (size_t i = 0; i < n; ++i) {
        const int64_t off = layout[i], len = layout[n + i];
        TORCH_CHECK(off >= 0 && len >= 0 && off + len <= flat.numel(), "packed model: segment ", i, " out of range");
        w.push_back(len > 0 ? flat.narrow(0, off, len) : at::empty({0}, flat.options()));
    }
    if (g_packed.size() >= 64) {            // stale versions of re-packed models: start over
        for (auto& kv : g_packed) { g_tn.erase(kv.second.handle); g_et.erase(kv.second.handle); g_gn.erase(kv.second.handle); g_tf.erase(kv.second.handle); }
        g_packed.clear();
    }
    lock.unlock();                          // *_create takes the lock itself
    const int64_t h = kind == 0 ? tensornet_create(w, ip, fp) : kind == 1 ? et_create(w, ip, fp) : kind == 2 ? gn_create(w, ip, fp) : t_create(w, ip, fp);
    lock.lock();
    return g_packed.emplace(key, PackedEntry{h, Tensor()}).first->second;
}

Tensor box_tensor(const std::vector<double>& box) {
    if (box.empty()) return at::empty({0, 0}, at::TensorOptions().dtype(at::kDouble));
    TORCH_CHECK(box.size() == 9, "packed model: box must have 9 entries (row-major 3x3) or none");
    Tensor b = at::empty({3, 3}, at::TensorOptions().dtype(at::kDouble));
    for (int i = 0; i < 9; ++i) b.data_ptr<double>()[i] = box[i];
    return b;
}

std::tuple<Tensor, Tensor, Tensor> tensornet_energy_forces_packed(
    const Tensor& flat, std::vector<int64_t> layout, std::vector<int64_t> ip, std::vector<double> fp, const Tensor& z,
    const Tensor& pos, const Tensor& batch, std::vector<double> box, bool use_periodic, const std::string& strategy,
    int64_t max_edges, int64_t n_mol, bool compute_forces) {
    PackedEntry& e = packed_entry(0, flat, layout, ip, fp);
    auto r = tensornet_energy_forces(e.handle, z, pos, batch, box_tensor(box), use_periodic, strategy, max_edges, n_mol,
                                     compute_forces, e.workspace.defined() ? c10::optional<Tensor>(e.workspace) : c10::nullopt);
    e.workspace = std::get<3>(r);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r));
}

std::tuple<Tensor, Tensor, Tensor> et_energy_forces_packed(
    const Tensor& flat, std::vector<int64_t> layout, std::vector<int64_t> ip, std::vector<double> fp, const Tensor& z,
    const Tensor& pos, const Tensor& batch, std::vector<double> box, bool use_periodic, const std::string& strategy,
    int64_t max_edges, int64_t n_mol, bool compute_forces) {
    PackedEntry& e = packed_entry(1, flat, layout, ip, fp);
    auto r = et_energy_forces(e.handle, z, pos, batch, box_tensor(box), use_periodic, strategy, max_edges, n_mol,
                              compute_forces, e.workspace.defined() ? c10::optional<Tensor>(e.workspace) : c10::nullopt);
    e.workspace = std::get<3>(r);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r));
}

std::tuple<Tensor, Tensor, Tensor> gn_energy_forces_packed(
    const Tensor& flat, std::vector<int64_t> layout, std::vector<int64_t> ip, std::vector<double> fp, const Tensor& z,
    const Tensor& pos, const Tensor& batch, std::vector<double> box, bool use_periodic, const std::string& strategy,
    int64_t max_edges, int64_t n_mol, bool compute_forces) {
    PackedEntry& e = packed_entry(2, flat, layout, ip, fp);
    auto r = gn_energy_forces(e.handle, z, pos, batch, box_tensor(box), use_periodic, strategy, max_edges, n_mol,
                              compute_forces, e.workspace.defined() ? c10::optional<Tensor>(e.workspace) : c10::nullopt);
    e.workspace = std::get<3>(r);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r));
}

std::tuple<Tensor, Tensor, Tensor> t_energy_forces_packed(
    const Tensor& flat, std::vector<int64_t> layout, std::vector<int64_t> ip, std::vector<double> fp, const Tensor& z,
    const Tensor& pos, const Tensor& batch, std::vector<double> box, bool use_periodic, const std::string& strategy,
    int64_t max_edges, int64_t n_mol, bool compute_forces) {
    PackedEntry& e = packed_entry(3, flat, layout, ip, fp);
    auto r = t_energy_forces(e.handle, z, pos, batch, box_tensor(box), use_periodic, strategy, max_edges, n_mol,
                             compute_forces, e.workspace.defined() ? c10::optional<Tensor>(e.workspace) : c10::nullopt);
    e.workspace = std::get<3>(r);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r));
}

// ------------------------------------------------------------------ staged evaluation (spatial decomposition)
int64_t tensornet_workspace_bytes(int64_t handle, int64_t n_atoms, int64_t max_edges, int64_t n_mol,
                                  const std::string& strategy, const Tensor& box_vectors) {
    auto h = tn_get(handle);
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    const size_t need = tmdn_tensornet_workspace_bytes(&h->m, n_atoms, max_edges, n_mol, strategy_id(strategy),
                                                       have_box ? box : nullptr);
    TORCH_CHECK(need > 0, "tensornet_workspace_bytes: ", tmdn_last_error());
    return (int64_t)need;
}

// One stage of the evaluation (include/tmdn.h: tmdn_tensornet_stage) on caller-owned outputs and workspace.
void tensornet_stage(int64_t handle, int64_t stage, int64_t layer, const Tensor& z, const Tensor& pos, const Tensor& batch,
                     int64_t n_rows, const Tensor& box_vectors, bool use_periodic, const std::string& strategy,
                     int64_t max_edges, int64_t n_mol, Tensor energy, Tensor forces, Tensor status, Tensor workspace) {
    auto h = tn_get(handle);
    TORCH_CHECK(pos.is_cuda() && pos.scalar_type() == at::kFloat && pos.is_contiguous() && pos.dim() == 2 &&
                    pos.size(1) == 3, "tensornet_stage: pos must be a contiguous CUDA float32 [N,3] tensor");
    TORCH_CHECK(z.scalar_type() == at::kLong && z.is_contiguous() && z.size(0) == pos.size(0) &&
                    batch.scalar_type() == at::kLong && batch.is_contiguous() && batch.size(0) == pos.size(0),
                "tensornet_stage: z / batch must be contiguous int64 [N] tensors");
    TORCH_CHECK(energy.is_cuda() && energy.scalar_type() == at::kFloat && energy.numel() >= n_mol &&
                    status.scalar_type() == at::kInt && status.numel() >= 4 && workspace.scalar_type() == at::kByte,
                "tensornet_stage: bad output tensors");
    TORCH_CHECK(!forces.defined() || forces.numel() == 0 ||
                    (forces.scalar_type() == at::kFloat && forces.is_contiguous() && forces.numel() >= 3 * n_rows),
                "tensornet_stage: forces must be a contiguous float32 [n_rows,3] tensor");
    const c10::cuda::CUDAGuard guard(pos.device());
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    float* fptr_ = (forces.defined() && forces.numel() > 0) ? forces.data_ptr<float>() : nullptr;
    TMDN_CALL(tmdn_tensornet_stage(&h->m, (int)stage, (int)layer, z.data_ptr<int64_t>(), pos.data_ptr<float>(),
                                   batch.data_ptr<int64_t>(), pos.size(0), n_rows, n_mol, have_box ? box : nullptr,
                                   use_periodic, strategy_id(strategy), max_edges, energy.data_ptr<float>(), fptr_,
                                   status.data_ptr<int32_t>(), workspace.data_ptr(), (size_t)workspace.numel(),
                                   cur_stream()));
}

// One stage over the owned rows [row_begin, row_end) (include/tmdn.h: tmdn_tensornet_stage_rows).
void tensornet_stage_rows(int64_t handle, int64_t stage, int64_t layer, const Tensor& z, const Tensor& pos, const Tensor& batch,
                          int64_t n_rows, int64_t row_begin, int64_t row_end, const Tensor& box_vectors, bool use_periodic,
                          const std::string& strategy, int64_t max_edges, int64_t n_mol, Tensor energy, Tensor forces,
                          Tensor status, Tensor workspace) {
    auto h = tn_get(handle);
    TORCH_CHECK(pos.is_cuda() && pos.scalar_type() == at::kFloat && pos.is_contiguous() && pos.dim() == 2 &&
                    pos.size(1) == 3, "tensornet_stage_rows: pos must be a contiguous CUDA float32 [N,3] tensor");
    TORCH_CHECK(z.scalar_type() == at::kLong && z.is_contiguous() && z.size(0) == pos.size(0) &&
                    batch.scalar_type() == at::kLong && batch.is_contiguous() && batch.size(0) == pos.size(0),
                "tensornet_stage_rows: z / batch must be contiguous int64 [N] tensors");
    TORCH_CHECK(energy.is_cuda() && energy.scalar_type() == at::kFloat && energy.numel() >= n_mol &&
                    status.scalar_type() == at::kInt && status.numel() >= 4 && workspace.scalar_type() == at::kByte,
                "tensornet_stage_rows: bad output tensors");
    TORCH_CHECK(!forces.defined() || forces.numel() == 0 ||
                    (forces.scalar_type() == at::kFloat && forces.is_contiguous() && forces.numel() >= 3 * n_rows),
                "tensornet_stage_rows: forces must be a contiguous float32 [n_rows,3] tensor");
    const c10::cuda::CUDAGuard guard(pos.device());
    double box[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool have_box = read_box(box_vectors, box);
    float* fptr_ = (forces.defined() && forces.numel() > 0) ? forces.data_ptr<float>() : nullptr;
    TMDN_CALL(tmdn_tensornet_stage_rows(&h->m, (int)stage, (int)layer, z.data_ptr<int64_t>(), pos.data_ptr<float>(),
                                        batch.data_ptr<int64_t>(), pos.size(0), n_rows, row_begin, row_end, n_mol,
                                        have_box ? box : nullptr, use_periodic, strategy_id(strategy), max_edges,
                                        energy.data_ptr<float>(), fptr_, status.data_ptr<int32_t>(), workspace.data_ptr(),
                                        (size_t)workspace.numel(), cur_stream()));
}

// Rows idx[seg[p]..seg[p+1]) of src -> the peer address dst[p] (device pointers into other GPUs' memory, as int64)
void halo_push(const Tensor& src, const Tensor& idx, const Tensor& seg, const Tensor& dst) {
    TORCH_CHECK(src.is_cuda() && src.scalar_type() == at::kFloat && src.is_contiguous() && src.dim() == 2,
                "halo_push: src must be a contiguous CUDA float32 [n,row] tensor");
    TORCH_CHECK(idx.is_cuda() && idx.scalar_type() == at::kInt && idx.is_contiguous(), "halo_push: idx must be int32");
    TORCH_CHECK(seg.is_cuda() && seg.scalar_type() == at::kLong && seg.is_contiguous() && seg.numel() >= 2,
                "halo_push: seg must be a CUDA int64 [n_peers+1] tensor");
    TORCH_CHECK(dst.is_cuda() && dst.scalar_type() == at::kLong && dst.is_contiguous() && dst.numel() == seg.numel() - 1,
                "halo_push: dst must be a CUDA int64 [n_peers] tensor of device pointers");
    const c10::cuda::CUDAGuard guard(src.device());
    TMDN_CALL(tmdn_halo_push_f32(src.data_ptr<float>(), idx.data_ptr<int32_t>(), seg.data_ptr<int64_t>(),
                                 reinterpret_cast<float* const*>(dst.data_ptr<int64_t>()), (int)dst.numel(), idx.numel(),
                                 src.size(1), cur_stream()));
}

// out[r,:] = src[idx[r],:]
Tensor halo_pack(const Tensor& src, const Tensor& idx) {
    TORCH_CHECK(src.is_cuda() && src.scalar_type() == at::kFloat && src.is_contiguous() && src.dim() == 2,
                "halo_pack: src must be a contiguous CUDA float32 [n,row] tensor");
    TORCH_CHECK(idx.is_cuda() && idx.scalar_type() == at::kInt && idx.is_contiguous(), "halo_pack: idx must be int32");
    const c10::cuda::CUDAGuard guard(src.device());
    Tensor out = at::empty({idx.numel(), src.size(1)}, src.options());
    TMDN_CALL(tmdn_halo_pack_f32(src.data_ptr<float>(), idx.data_ptr<int32_t>(), idx.numel(), src.size(1),
                                 out.data_ptr<float>(), cur_stream()));
    return out;
}

// y[a,c,:] (+)= x[a,c,:] * w[g(c)]^T  (grouped channel mix of TensorNet)
Tensor channel_mix(const Tensor& x, const Tensor& w, const c10::optional<Tensor>& y_, bool accumulate) {
    TORCH_CHECK(x.is_cuda() && x.scalar_type() == at::kFloat && x.is_contiguous() && x.dim() == 3 && x.size(1) == 9,
                "channel_mix: x must be a contiguous CUDA float32 [n,9,H] tensor");
    const int64_t H = x.size(2);
    TORCH_CHECK(w.is_cuda() && w.scalar_type() == at::kFloat && w.is_contiguous() && w.dim() == 3 && w.size(0) == 3 &&
                    w.size(1) == H && w.size(2) == H,
                "channel_mix: w must be a contiguous CUDA float32 [3,H,H] tensor");
    const c10::cuda::CUDAGuard guard(x.device());
    Tensor y = y_.has_value() ? *y_ : at::zeros_like(x);
    TORCH_CHECK(y.is_cuda() && y.scalar_type() == at::kFloat && y.is_contiguous() && y.sizes() == x.sizes(),
                "channel_mix: y must match x");
    TMDN_CALL(tmdn_channel_mix_f32(x.data_ptr<float>(), w.data_ptr<float>(), y.data_ptr<float>(), x.size(0), (int)H,
                                   accumulate ? 1 : 0, cur_stream()));
    return y;
}

// c (+)= a w^T + bias
Tensor node_gemm(const Tensor& a, const Tensor& w, const c10::optional<Tensor>& bias, const c10::optional<Tensor>& c_,
                 bool accumulate) {
    TORCH_CHECK(a.is_cuda() && a.scalar_type() == at::kFloat && a.is_contiguous() && a.dim() == 2, "node_gemm: a");
    TORCH_CHECK(w.is_cuda() && w.scalar_type() == at::kFloat && w.is_contiguous() && w.dim() == 2 && w.size(1) == a.size(1),
                "node_gemm: w must be [N,K]");
    const c10::cuda::CUDAGuard guard(a.device());
    Tensor c = c_.has_value() ? *c_ : at::zeros({a.size(0), w.size(0)}, a.options());
    TORCH_CHECK(c.is_cuda() && c.scalar_type() == at::kFloat && c.is_contiguous() && c.size(0) == a.size(0) &&
                    c.size(1) == w.size(0), "node_gemm: c");
    const float* b = nullptr;
    if (bias.has_value()) {
        TORCH_CHECK(bias->is_cuda() && bias->scalar_type() == at::kFloat && bias->is_contiguous() && bias->numel() == w.size(0),
                    "node_gemm: bias");
        b = bias->data_ptr<float>();
    }
    TMDN_CALL(tmdn_node_gemm_f32(a.data_ptr<float>(), w.data_ptr<float>(), b, c.data_ptr<float>(), a.size(0), w.size(0),
                                 a.size(1), accumulate ? 1 : 0, cur_stream()));
    return c;
}

// energy[n_mol,1] and forces[N,3] are updated in place; -> num_found int32[1]
Tensor pair_prior(const std::string& kind, const Tensor& pos, const Tensor& batch, const Tensor& weight, double cutoff,
                  double energy_prefactor, double distance_scale, double alpha, int64_t max_edges, Tensor energy,
                  const c10::optional<Tensor>& forces) {
    check_input(pos, batch);
    TORCH_CHECK(pos.scalar_type() == at::kFloat, "pair_prior: fp32 positions only");
    TORCH_CHECK(weight.is_cuda() && weight.scalar_type() == at::kFloat && weight.is_contiguous() &&
                    weight.numel() == (kind == "d2" ? 2 : 1) * pos.size(0),
                "pair_prior: weight must be a contiguous CUDA float32 [N] tensor ([2,N] for d2)");
    TORCH_CHECK(energy.is_cuda() && energy.scalar_type() == at::kFloat && energy.is_contiguous(), "pair_prior: energy");
    const int k = kind == "zbl" ? TMDN_PRIOR_ZBL : kind == "coulomb" ? TMDN_PRIOR_COULOMB : kind == "d2" ? TMDN_PRIOR_D2 : -1;
    TORCH_CHECK(k >= 0, "pair_prior: unknown kind '", kind, "'");
    float* fp = nullptr;
    if (forces.has_value() && forces->defined() && forces->numel() > 0) {
        TORCH_CHECK(forces->is_cuda() && forces->scalar_type() == at::kFloat && forces->is_contiguous() &&
                        forces->numel() == 3 * pos.size(0), "pair_prior: forces must be a contiguous float32 [N,3] tensor");
        fp = forces->data_ptr<float>();
    }
    const c10::cuda::CUDAGuard guard(pos.device());
    const int64_t n = pos.size(0);
    const size_t ws_bytes = tmdn_pair_prior_workspace_bytes(n, max_edges);
    TORCH_CHECK(ws_bytes > 0, "pair_prior: bad sizes");
    Tensor ws = at::empty({(int64_t)ws_bytes}, pos.options().dtype(at::kByte));
    Tensor num = at::empty({1}, pos.options().dtype(at::kInt));
    TMDN_CALL(tmdn_pair_prior_f32(k, pos.data_ptr<float>(), batch.data_ptr<int64_t>(), weight.data_ptr<float>(), n,
                                  energy.numel(), cutoff, energy_prefactor, distance_scale, alpha, max_edges,
                                  energy.data_ptr<float>(), fp, num.data_ptr<int32_t>(), ws.data_ptr(), ws_bytes, cur_stream()));
    return num;
}

// out[m] += sum of values over molecule m (batch sorted); deterministic
void segment_add(const Tensor& values, const Tensor& batch, Tensor out) {
    TORCH_CHECK(values.is_cuda() && values.scalar_type() == at::kFloat && values.is_contiguous() && values.dim() == 1,
                "segment_add: values must be a contiguous CUDA float32 vector");
    TORCH_CHECK(batch.is_cuda() && batch.scalar_type() == at::kLong && batch.is_contiguous() && batch.numel() == values.numel(),
                "segment_add: batch must be a contiguous CUDA int64 vector of the same length");
    TORCH_CHECK(out.is_cuda() && out.scalar_type() == at::kFloat && out.is_contiguous(), "segment_add: out");
    TORCH_CHECK(values.get_device() == batch.get_device() && values.get_device() == out.get_device(), "segment_add: devices differ");
    if (values.numel() == 0 || out.numel() == 0) return;
    const c10::cuda::CUDAGuard guard(values.device());
    TMDN_CALL(tmdn_segment_add_f32(values.data_ptr<float>(), batch.data_ptr<int64_t>(), values.numel(), out.numel(),
                                   out.data_ptr<float>(), cur_stream()));
}

int64_t launch_count() { return (int64_t)tmdn_launch_count(); }
int64_t set_tuning(std::string key, int64_t value) {
    const int old = tmdn_set_tuning(key.c_str(), (int)value);
    TORCH_CHECK(old >= 0, tmdn_last_error());
    return old;
}
void profile_enable(bool on) { TMDN_CALL(tmdn_profile_enable(on ? 1 : 0)); }
std::string profile_report() {
    std::vector<char> buf(1 << 16);
    TMDN_CALL(tmdn_profile_report(buf.data(), buf.size()));
    return std::string(buf.data());
}

}  // namespace

TORCH_LIBRARY(torchmdnet_neighbors, m) {
    m.def("get_neighbor_pairs(str strategy, Tensor positions, Tensor batch, Tensor box_vectors, bool use_periodic, "
          "Scalar cutoff_lower, Scalar cutoff_upper, Scalar max_num_pairs, bool loop, bool include_transpose) -> "
          "(Tensor neighbors, Tensor distances, Tensor distance_vecs, Tensor num_pairs)");
}
TORCH_LIBRARY_IMPL(torchmdnet_neighbors, AutogradCUDA, m) { m.impl("get_neighbor_pairs", get_neighbor_pairs_autograd); }
TORCH_LIBRARY_IMPL(torchmdnet_neighbors, CPU, m) { m.impl("get_neighbor_pairs", get_neighbor_pairs_cpu); }

// CONSERVATIVE alias analysis = "has side effects": TorchScript must not constant-fold or hoist a call that takes no
// arguments (it would freeze the answer of the first, non-capturing run into the graph)
TORCH_LIBRARY(torch_extension, m) {
    m.def(torch::schema("is_stream_capturing() -> bool", c10::AliasAnalysisKind::CONSERVATIVE), is_stream_capturing);
}

TORCH_LIBRARY(torchmdnet_b200, m) {
    m.def("csr_from_pairs(Tensor neighbors, Tensor deltas, Tensor distances, Tensor num_found, int n_atoms, "
          "int target_row) -> (Tensor, Tensor, Tensor)",
          csr_from_pairs);
    m.def("tensornet_create(Tensor[] weights, int[] iparams, float[] fparams) -> int", tensornet_create);
    m.def("tensornet_destroy(int handle) -> ()", tensornet_destroy);
    m.def("tensornet_energy_forces(int handle, Tensor z, Tensor pos, Tensor batch, Tensor box_vectors, "
          "bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces, Tensor? workspace) -> "
          "(Tensor, Tensor, Tensor, Tensor)",
          tensornet_energy_forces);
    m.def("tensornet_energy_forces_packed(Tensor flat, int[] layout, int[] iparams, float[] fparams, Tensor z, Tensor pos, "
          "Tensor batch, float[] box, bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces) -> "
          "(Tensor, Tensor, Tensor)", tensornet_energy_forces_packed);
    m.def("et_energy_forces_packed(Tensor flat, int[] layout, int[] iparams, float[] fparams, Tensor z, Tensor pos, "
          "Tensor batch, float[] box, bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces) -> "
          "(Tensor, Tensor, Tensor)", et_energy_forces_packed);
    m.def("t_create(Tensor[] weights, int[] iparams, float[] fparams) -> int", t_create);
    m.def("t_destroy(int handle) -> ()", t_destroy);
    m.def("t_energy_forces(int handle, Tensor z, Tensor pos, Tensor batch, Tensor box_vectors, "
          "bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces, Tensor? workspace) -> "
          "(Tensor, Tensor, Tensor, Tensor)", t_energy_forces);
    m.def("t_energy_forces_packed(Tensor flat, int[] layout, int[] iparams, float[] fparams, Tensor z, Tensor pos, "
          "Tensor batch, float[] box, bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces) -> "
          "(Tensor, Tensor, Tensor)", t_energy_forces_packed);
    m.def("gn_create(Tensor[] weights, int[] iparams, float[] fparams) -> int", gn_create);
    m.def("gn_destroy(int handle) -> ()", gn_destroy);
    m.def("gn_energy_forces(int handle, Tensor z, Tensor pos, Tensor batch, Tensor box_vectors, "
          "bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces, Tensor? workspace) -> "
          "(Tensor, Tensor, Tensor, Tensor)", gn_energy_forces);
    m.def("gn_debug_lookup(int handle, int n_atoms, int max_edges, int n_mol, str strategy, Tensor box_vectors, "
          "str name) -> int[]", gn_debug_lookup);
    m.def("gn_energy_forces_packed(Tensor flat, int[] layout, int[] iparams, float[] fparams, Tensor z, Tensor pos, "
          "Tensor batch, float[] box, bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces) -> "
          "(Tensor, Tensor, Tensor)", gn_energy_forces_packed);
    m.def("et_create(Tensor[] weights, int[] iparams, float[] fparams) -> int", et_create);
    m.def("et_destroy(int handle) -> ()", et_destroy);
    m.def("et_energy_forces(int handle, Tensor z, Tensor pos, Tensor batch, Tensor box_vectors, "
          "bool use_periodic, str strategy, int max_edges, int n_mol, bool compute_forces, Tensor? workspace) -> "
          "(Tensor, Tensor, Tensor, Tensor)",
          et_energy_forces);
    m.def("et_debug_lookup(int handle, int n_atoms, int max_edges, int n_mol, str strategy, Tensor box_vectors, "
          "str name) -> int[]",
          et_debug_lookup);
    m.def("tensornet_workspace_bytes(int handle, int n_atoms, int max_edges, int n_mol, str strategy, "
          "Tensor box_vectors) -> int", tensornet_workspace_bytes);
    m.def("tensornet_stage(int handle, int stage, int layer, Tensor z, Tensor pos, Tensor batch, int n_rows, "
          "Tensor box_vectors, bool use_periodic, str strategy, int max_edges, int n_mol, Tensor(a!) energy, "
          "Tensor(b!) forces, Tensor(c!) status, Tensor(d!) workspace) -> ()", tensornet_stage);
    m.def("tensornet_stage_rows(int handle, int stage, int layer, Tensor z, Tensor pos, Tensor batch, int n_rows, "
          "int row_begin, int row_end, Tensor box_vectors, bool use_periodic, str strategy, int max_edges, int n_mol, "
          "Tensor(a!) energy, Tensor(b!) forces, Tensor(c!) status, Tensor(d!) workspace) -> ()", tensornet_stage_rows);
    m.def("halo_pack(Tensor src, Tensor idx) -> Tensor", halo_pack);
    m.def("halo_push(Tensor src, Tensor idx, Tensor seg, Tensor dst) -> ()", halo_push);
    m.def("channel_mix(Tensor x, Tensor w, Tensor? y, bool accumulate) -> Tensor", channel_mix);
    m.def("node_gemm(Tensor a, Tensor w, Tensor? bias, Tensor? c, bool accumulate) -> Tensor", node_gemm);
    m.def("pair_prior(str kind, Tensor pos, Tensor batch, Tensor weight, float cutoff, float energy_prefactor, "
          "float distance_scale, float alpha, int max_edges, Tensor(a!) energy, Tensor(b!)? forces) -> Tensor", pair_prior);
    m.def("segment_add(Tensor values, Tensor batch, Tensor(a!) out) -> ()", segment_add);
    m.def("launch_count() -> int", launch_count);
    m.def("set_tuning(str key, int value) -> int", set_tuning);
    m.def("profile_enable(bool on) -> ()", profile_enable);
    m.def("profile_report() -> str", profile_report);
    m.def("csr_build(str strategy, Tensor positions, Tensor batch, Tensor box_vectors, bool use_periodic, "
          "float cutoff_lower, float cutoff_upper, int max_edges) -> (Tensor, Tensor, Tensor, Tensor)",
          csr_build);
    m.def("tensornet_debug_lookup(int handle, int n_atoms, int max_edges, int n_mol, str strategy, Tensor box_vectors, "
          "str name) -> int[]",
          tensornet_debug_lookup);
}
