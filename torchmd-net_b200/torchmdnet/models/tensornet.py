"""TensorNet (reference: torchmdnet/models/tensornet.py) as a parameter tree + weight packer for the fused
CUDA energy/force pipeline of libtmdn (csrc/tensornet.cu).

The module tree reproduces the reference's parameter names, shapes and RNG consumption order
(tensornet.py:104-198, 236-286, 335-375), so ``state_dict``s are interchangeable and a seeded ``create_model``
gives bit-identical initial weights.  No layer has a PyTorch ``forward``: ``pack_weights`` rearranges the
parameters once per weight version into the layouts the kernels read and ``TensorNet.energy_forces`` runs
neighbour search, embedding, interactions, head and the analytic backward in one C-ABI call.

Two deliberate differences from the reference's *CUDA* behaviour (both documented in DESIGN.md):
  * the unpadded edge semantics of the reference CPU path are implemented -- the static-shape padding
    artefact of tensornet.py:215-221 (spurious self-loop messages on atom 0) is not reproduced;
  * matmuls keep fp32-level accuracy (default: three-product tf32 on the tensor cores with fp32 chunk sums,
    ``gemm_mode = "fp32"`` = FFMA); the reference flips plain TF32 on process-wide at import (tensornet.py:13-14).
"""
from typing import Optional, Tuple

import torch
from torch import Tensor, nn

from ._fused import FusedRepresentation
from .utils import (ACT_ID, RBF_ID, CosineCutoff, OptimizedDistance, _fused_only, act_class_mapping,
                    rbf_class_mapping)

__all__ = ["TensorNet"]

FILTER_EXACT, FILTER_TABLE = 0, 1


class TensorEmbedding(nn.Module):
    """Parameters of tensornet.py:236-286."""

    def __init__(self, hidden_channels, num_rbf, activation, cutoff_lower, cutoff_upper, trainable_rbf=False,
                 max_z=128, dtype=torch.float32):
        super().__init__()
        H = hidden_channels
        self.hidden_channels = H
        self.distance_proj1 = nn.Linear(num_rbf, H, dtype=dtype)
        self.distance_proj2 = nn.Linear(num_rbf, H, dtype=dtype)
        self.distance_proj3 = nn.Linear(num_rbf, H, dtype=dtype)
        self.cutoff = CosineCutoff(cutoff_lower, cutoff_upper)
        self.max_z = max_z
        self.emb = nn.Embedding(max_z, H, dtype=dtype)
        self.emb2 = nn.Linear(2 * H, H, dtype=dtype)
        self.act = activation()
        # the reference builds these three without dtype= (tensornet.py:261-264); TorchMD_Net casts afterwards
        self.linears_tensor = nn.ModuleList([nn.Linear(H, H, bias=False) for _ in range(3)])
        self.linears_scalar = nn.ModuleList([nn.Linear(H, 2 * H, bias=True, dtype=dtype),
                                             nn.Linear(2 * H, 3 * H, bias=True, dtype=dtype)])
        self.init_norm = nn.LayerNorm(H, dtype=dtype)
        self.reset_parameters()

    def reset_parameters(self):
        for m in (self.distance_proj1, self.distance_proj2, self.distance_proj3, self.emb, self.emb2,
                  *self.linears_tensor, *self.linears_scalar, self.init_norm):
            m.reset_parameters()

    forward = _fused_only


class Interaction(nn.Module):
    """Parameters of tensornet.py:335-375."""

    def __init__(self, num_rbf, hidden_channels, activation, cutoff_lower, cutoff_upper,
                 equivariance_invariance_group, dtype=torch.float32):
        super().__init__()
        H = hidden_channels
        self.num_rbf = num_rbf
        self.hidden_channels = H
        self.cutoff = CosineCutoff(cutoff_lower, cutoff_upper)
        self.linears_scalar = nn.ModuleList([nn.Linear(num_rbf, H, bias=True, dtype=dtype),
                                             nn.Linear(H, 2 * H, bias=True, dtype=dtype),
                                             nn.Linear(2 * H, 3 * H, bias=True, dtype=dtype)])
        self.linears_tensor = nn.ModuleList([nn.Linear(H, H, bias=False) for _ in range(6)])
        self.act = activation()
        self.equivariance_invariance_group = equivariance_invariance_group
        self.reset_parameters()

    def reset_parameters(self):
        for m in (*self.linears_scalar, *self.linears_tensor):
            m.reset_parameters()

    forward = _fused_only


# ------------------------------------------------------------------------------------------------ fp64 helpers
def _act_and_grad(name, x):
    if name == "silu":
        s = torch.sigmoid(x)
        return x * s, s * (1 + x * (1 - s))
    if name == "tanh":
        y = torch.tanh(x)
        return y, 1 - y * y
    if name == "sigmoid":
        s = torch.sigmoid(x)
        return s, s * (1 - s)
    if name == "ssp":
        return torch.nn.functional.softplus(x) - 0.6931471805599453, torch.sigmoid(x)
    raise ValueError(name)


def _cutoff_and_grad(d, lo, hi):
    pi = 3.141592653589793
    if lo > 0:
        arg = pi * (2 * (d - lo) / (hi - lo) + 1.0)
        m = ((d < hi) & (d > lo)).to(d.dtype)
        return 0.5 * (torch.cos(arg) + 1) * m, -0.5 * torch.sin(arg) * (2 * pi / (hi - lo)) * m
    m = (d < hi).to(d.dtype)
    return 0.5 * (torch.cos(d * pi / hi) + 1) * m, -0.5 * torch.sin(d * pi / hi) * (pi / hi) * m


def filter_table(d, rbf_type, p0, p1, lo, hi, layers, act, final_act, cutoff=True):
    """f(d) = [act o] Lin_n(... act(Lin_1(rbf(d)))) * C(d) and df/dd at the points ``d`` (all in d.dtype;
    called in fp64 on the device).  ``layers`` = [(W[out,in], b[out]), ...].  Reference functions:
    models/utils.py:298-300, 339-344, 368-390; tensornet.py:295-304, 381-382."""
    dd = d.unsqueeze(-1)
    if rbf_type == "expnorm":
        alpha = 5.0 / (hi - lo)
        c0, dc0 = _cutoff_and_grad(dd, 0.0, hi)
        u = torch.exp(alpha * (lo - dd))
        g = torch.exp(-p1 * (u - p0) ** 2)
        h, t = c0 * g, dc0 * g + c0 * g * (2 * alpha) * p1 * u * (u - p0)
    else:
        g = torch.exp(p1 * (dd - p0) ** 2)
        h, t = g, 2 * p1 * (dd - p0) * g
    for k, (W, b) in enumerate(layers):
        h = h @ W.t() + b
        t = t @ W.t()
        if k < len(layers) - 1 or final_act:
            h, gact = _act_and_grad(act, h)
            t = t * gact
    if not cutoff:
        return h, t
    c, dc = _cutoff_and_grad(d, lo, hi)
    return h * c[:, None], t * c[:, None] + h * dc[:, None]


class TableGrid:
    """Interval grid of the filter tables (csrc/table_grid.cuh): ``n_int`` intervals over [0, hi], uniform in
    t(d) = (exp(-alpha d) - exp(-alpha hi)) + gamma (hi - d), counted from the cutoff (row 0 ends at d = hi).
    ExpNormal filters have features of constant width in exp(-alpha d) (models/utils.py:339-344) times factors that
    vary on the scale of the cutoff, so this grid needs far fewer intervals than one uniform in d.  alpha = 0
    (Gaussian smearing, cutoff_lower > 0, or ``warp=False``) is the grid uniform in d."""

    def __init__(self, n_int, lo, hi, rbf_type, warp=True):
        self.n_int, self.hi = int(n_int), float(hi)
        if warp and rbf_type == "expnorm" and float(lo) == 0.0:
            self.alpha, self.gamma = 5.0 / self.hi, 0.05 / self.hi
        else:
            self.alpha = self.gamma = 0.0

    def t(self, d):
        if self.alpha == 0:
            return self.hi - d
        import math
        return (torch.exp(-self.alpha * d) - math.exp(-self.alpha * self.hi)) + self.gamma * (self.hi - d)

    def dt(self, d):
        return -(self.alpha * torch.exp(-self.alpha * d) + self.gamma) if self.alpha > 0 else -torch.ones_like(d)

    @property
    def step(self):
        import math
        T = 1.0 - math.exp(-self.alpha * self.hi) + self.gamma * self.hi if self.alpha > 0 else self.hi
        return T / self.n_int

    def nodes(self, device=None):
        """d_k with t(d_k) = k * step, k = 0..n_int: descending from hi to 0 (fp64; Newton on the monotone t)."""
        tk = torch.arange(self.n_int + 1, device=device, dtype=torch.float64) * self.step
        if self.alpha == 0:
            return self.hi - tk
        d = self.hi * (1.0 - tk / (self.step * self.n_int))
        for _ in range(80):
            d = (d - (self.t(d) - tk) / self.dt(d)).clamp(0.0, self.hi)
        d[0], d[-1] = self.hi, 0.0
        return d

    def coefficients(self, d, f, df):
        """Cubic Hermite coefficients in the local coordinate of each interval from f, df/dd at the nodes ``d``."""
        return hermite_coefficients(f, df / self.dt(d)[:, None], self.step)

    def locate(self, d):
        """-> (interval k, local coordinate s, ds/dd) exactly as the kernels compute them (tab_locate)."""
        q = self.t(d) / self.step
        k = q.floor().clamp(min=0, max=self.n_int - 1).long()
        return k, q - k, self.dt(d) / self.step


def hermite_coefficients(f, df, h):
    """Cubic Hermite coefficients per interval: value(s) = c0 + s(c1 + s(c2 + s c3)), s in [0,1).
    f, df: [n_int+1, C] -> [n_int, C, 4]."""
    f0, f1, d0, d1 = f[:-1], f[1:], df[:-1] * h, df[1:] * h
    return torch.stack([f0, d0, 3 * (f1 - f0) - 2 * d0 - d1, 2 * (f0 - f1) + d0 + d1], -1)


class TensorNet(FusedRepresentation):
    """Drop-in for ``torchmdnet.models.tensornet.TensorNet`` (tensornet.py:70-234): same constructor, same
    parameters.  ``filter_mode``: "exact" evaluates the per-edge filter MLPs with fp32 GEMMs, "table"
    evaluates them from a cubic Hermite table built in fp64 from the weights on the grid of ``TableGrid``
    (``table_intervals`` intervals, default max(1024, 32 * num_rbf); ``table_warp=False`` = uniform in d)."""
    _op = "tensornet"
    _kernel_name = "TensorNet"

    def __init__(self, hidden_channels=128, num_layers=2, num_rbf=32, rbf_type="expnorm", trainable_rbf=False,
                 activation="silu", cutoff_lower=0, cutoff_upper=4.5, max_num_neighbors=64, max_z=128,
                 equivariance_invariance_group="O(3)", static_shapes=True, dtype=torch.float32):
        super().__init__()
        assert rbf_type in rbf_class_mapping, f'Unknown RBF type "{rbf_type}".'
        assert activation in act_class_mapping, f'Unknown activation function "{activation}".'
        assert equivariance_invariance_group in ["O(3)", "SO(3)"], \
            f'Unknown group "{equivariance_invariance_group}". Choose O(3) or SO(3).'
        self.hidden_channels = hidden_channels
        self.equivariance_invariance_group = equivariance_invariance_group
        self.num_layers = num_layers
        self.num_rbf = num_rbf
        self.rbf_type = rbf_type
        self.activation = activation
        self.cutoff_lower = cutoff_lower
        self.cutoff_upper = cutoff_upper
        self.max_z = max_z
        self.max_num_neighbors = max_num_neighbors
        act_class = act_class_mapping[activation]
        self.distance_expansion = rbf_class_mapping[rbf_type](cutoff_lower, cutoff_upper, num_rbf, trainable_rbf)
        self.tensor_embedding = TensorEmbedding(hidden_channels, num_rbf, act_class, cutoff_lower, cutoff_upper,
                                                trainable_rbf, max_z, dtype)
        self.layers = nn.ModuleList([Interaction(num_rbf, hidden_channels, act_class, cutoff_lower, cutoff_upper,
                                                 equivariance_invariance_group, dtype)
                                     for _ in range(num_layers)])
        self.linear = nn.Linear(3 * hidden_channels, hidden_channels, dtype=dtype)
        self.out_norm = nn.LayerNorm(3 * hidden_channels, dtype=dtype)
        self.act = act_class()
        self.static_shapes = static_shapes
        # kept for API compatibility (callers swap in a periodic OptimizedDistance by attribute);
        # the fused op reads cutoffs / box / strategy / capacity from it on every call.
        self.distance = OptimizedDistance(cutoff_lower, cutoff_upper, max_num_pairs=-max_num_neighbors,
                                          return_vecs=True, loop=True, check_errors=False,
                                          resize_to_fit=not static_shapes, long_edge_index=True)
        self._init_fused()
        self.filter_mode = "exact"
        # MLP GEMMs: "tf32x3" (default) = tcgen05 three-product tf32 with the K chunks summed in fp32 outside the tensor
        # core (rms error below an fp32 FFMA GEMM); "fp32" = fp32 FFMA kernel.  The channel mixes always use tcgen05.
        self.gemm_mode = "tf32x3"
        self.reset_parameters()

    def reset_parameters(self):
        self.tensor_embedding.reset_parameters()
        for layer in self.layers:
            layer.reset_parameters()
        self.linear.reset_parameters()
        self.out_norm.reset_parameters()

    # -------------------------------------------------------------------------------------------- packing
    def pack_weights(self, head: Optional[nn.Module], mean: float, std: float):
        """-> (weights list in the order csrc/torch_shim.cpp::tensornet_create expects, iparams, fparams)."""
        dev = self.linear.weight.device
        H, K = self.hidden_channels, self.num_rbf
        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()   # noqa: E731
        f64 = lambda t: t.detach().to(device=dev, dtype=torch.float64)                 # noqa: E731
        tr = lambda t: f32(t.detach().t())                                             # noqa: E731
        te = self.tensor_embedding
        table = self.filter_mode == "table"
        n_int = self.resolved_table_intervals() if table else 0
        empty = torch.empty(0, device=dev, dtype=torch.float32)
        lo, hi = float(self.cutoff_lower), float(self.cutoff_upper)

        if self.rbf_type == "expnorm":
            p0, p1 = self.distance_expansion.means, self.distance_expansion.betas
        else:
            p0 = self.distance_expansion.offset
            p1 = self.distance_expansion.coeff.expand(K)
        grid = TableGrid(max(n_int, 1), lo, hi, self.rbf_type, self.table_warp)
        nodes = grid.nodes(dev) if table else None

        def make_table(layers, final_act):
            if not table:
                return empty
            f, df = filter_table(nodes, self.rbf_type, f64(p0), f64(p1), lo, hi,
                                 [(f64(w), f64(b)) for w, b in layers], self.activation, final_act)
            return f32(grid.coefficients(nodes, f, df).view(n_int, 3, H, 4))

        # Z_ij = emb2([emb(z_i) | emb(z_j)]) = P[z_i] + Q[z_j]      (tensornet.py:310-312)
        emb, w2, b2 = f64(te.emb.weight), f64(te.emb2.weight), f64(te.emb2.bias)
        P, Q = emb @ w2[:, :H].t() + b2, emb @ w2[:, H:].t()
        # reference filter rows are h*3+g (reshape(-1,H,3), tensornet.py:325,383); kernels want g*H+h
        perm = torch.arange(3 * H, device=dev).view(H, 3).t().reshape(-1)
        ew = torch.cat([te.distance_proj1.weight, te.distance_proj2.weight, te.distance_proj3.weight]).detach()
        eb = torch.cat([te.distance_proj1.bias, te.distance_proj2.bias, te.distance_proj3.bias]).detach()
        w = [f32(p0), f32(p1), f32(P), f32(Q), f32(ew), f32(eb), make_table([(ew, eb)], False)]
        w += [f32(l.weight) for l in te.linears_tensor] + [tr(l.weight) for l in te.linears_tensor]
        w += [f32(te.init_norm.weight), f32(te.init_norm.bias)]
        s0, s1 = te.linears_scalar
        w += [f32(s0.weight), f32(s0.bias), tr(s0.weight),
              f32(s1.weight[perm]), f32(s1.bias[perm]), tr(s1.weight[perm])]
        w += [f32(self.out_norm.weight), f32(self.out_norm.bias), f32(self.linear.weight), f32(self.linear.bias),
              tr(self.linear.weight)]
        if head is not None:
            h0, h2 = head.output_network[0], head.output_network[2]
            w += [f32(h0.weight), f32(h0.bias), tr(h0.weight), f32(h2.weight), f32(h2.bias)]
        else:   # features-only use: a zero head
            w += [torch.zeros(H // 2, H, device=dev), torch.zeros(H // 2, device=dev),
                  torch.zeros(H, H // 2, device=dev), torch.zeros(1, H // 2, device=dev), torch.zeros(1, device=dev)]
        for layer in self.layers:
            ls = layer.linears_scalar
            flt = [(ls[0].weight.detach(), ls[0].bias.detach()), (ls[1].weight.detach(), ls[1].bias.detach()),
                   (ls[2].weight.detach()[perm], ls[2].bias.detach()[perm])]
            for wk, bk in flt:
                w += [f32(wk), f32(bk)]
            w.append(make_table(flt, True))
            w += [f32(l.weight) for l in layer.linears_tensor] + [tr(l.weight) for l in layer.linears_tensor]
        iparams = [H, K, self.num_layers, self.max_z, RBF_ID[self.rbf_type], ACT_ID[self.activation],
                   int(self.equivariance_invariance_group == "SO(3)"), FILTER_TABLE if table else FILTER_EXACT,
                   n_int, {"fp32": 0, "tf32x3": 1}[self.gemm_mode]]
        fparams = [lo, hi, float(mean), float(std), grid.alpha, grid.gamma]
        return w, iparams, fparams

    def _extra_key(self):
        return (self.filter_mode, self.gemm_mode, self.table_warp)

    def forward(self, z: Tensor, pos: Tensor, batch: Tensor, q: Optional[Tensor] = None,
                s: Optional[Tensor] = None):
        """Representation only (tensornet.py:200-234): -> (x[N,H], None, z, pos, batch).  Runs the fused forward
        with a zero head and returns the activations of the final Linear+act."""
        n_mol = int(batch.max()) + 1
        self.energy_forces(z, pos, batch, n_mol, None, 0.0, 1.0, compute_forces=False)
        x = self.debug_buffer("a3", pos.shape[0], n_mol).view(pos.shape[0], self.hidden_channels).clone()
        return x, None, z, pos, batch
