"""TorchMD_GN, the SchNet-style graph network (reference: torchmdnet/models/torchmd_gn.py), as a parameter tree + weight
packer for the fused CUDA energy/force pipeline (csrc/et.cu, ``tmdn_gn_energy_forces``).

The module tree reproduces the reference's parameter names (including the filter MLP being registered both as
``interactions.l.mlp`` and as ``interactions.l.conv.net``), shapes and RNG consumption order
(torchmd_gn.py:60-146, 185-258), so ``state_dict``s are interchangeable and a seeded ``create_model`` gives the same
initial weights.  The continuous-filter convolution's filter ``W(d) = mlp(rbf(d)) * C(d)`` is univariate in d and is
evaluated from a cubic Hermite table built in fp64 from the weights (``tensornet.TableGrid``).  ``aggr="add"`` only.
The pair set is that of the reference's legacy ``Distance`` module: every pair within the cutoff, no self pairs."""
from typing import Optional

import torch
from torch import Tensor, nn

from ._fused import FusedRepresentation
from .tensornet import TableGrid, filter_table
from .utils import (ACT_ID, CosineCutoff, NeighborEmbedding, OptimizedDistance, _fused_only, act_class_mapping,
                    rbf_class_mapping)

__all__ = ["TorchMD_GN"]


class CFConv(nn.Module):
    """Parameters of torchmd_gn.py:238-290."""

    def __init__(self, in_channels, out_channels, num_filters, net, cutoff_lower, cutoff_upper, aggr="add",
                 dtype=torch.float32):
        super().__init__()
        self.lin1 = nn.Linear(in_channels, num_filters, bias=False, dtype=dtype)
        self.lin2 = nn.Linear(num_filters, out_channels, bias=True, dtype=dtype)
        self.net = net
        self.cutoff = CosineCutoff(cutoff_lower, cutoff_upper)
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.xavier_uniform_(self.lin1.weight)
        nn.init.xavier_uniform_(self.lin2.weight)
        self.lin2.bias.data.fill_(0)

    def jittable(self):
        return self

    forward = _fused_only


class InteractionBlock(nn.Module):
    """Parameters of torchmd_gn.py:185-236."""

    def __init__(self, hidden_channels, num_rbf, num_filters, activation, cutoff_lower, cutoff_upper, aggr="add",
                 dtype=torch.float32):
        super().__init__()
        self.mlp = nn.Sequential(nn.Linear(num_rbf, num_filters, dtype=dtype), activation(),
                                 nn.Linear(num_filters, num_filters, dtype=dtype))
        self.conv = CFConv(hidden_channels, hidden_channels, num_filters, self.mlp, cutoff_lower, cutoff_upper, aggr=aggr,
                           dtype=dtype).jittable()
        self.act = activation()
        self.lin = nn.Linear(hidden_channels, hidden_channels, dtype=dtype)
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.xavier_uniform_(self.mlp[0].weight)
        self.mlp[0].bias.data.fill_(0)
        nn.init.xavier_uniform_(self.mlp[2].weight)
        self.mlp[2].bias.data.fill_(0)
        self.conv.reset_parameters()
        nn.init.xavier_uniform_(self.lin.weight)
        self.lin.bias.data.fill_(0)

    forward = _fused_only


class TorchMD_GN(FusedRepresentation):
    """Drop-in for ``torchmdnet.models.torchmd_gn.TorchMD_GN`` (torchmd_gn.py:14-183): same constructor, same parameters."""
    _op = "gn"
    _kernel_name = "TorchMD_GN"

    def __init__(self, hidden_channels=128, num_filters=128, num_layers=6, num_rbf=50, rbf_type="expnorm",
                 trainable_rbf=True, activation="silu", neighbor_embedding=True, cutoff_lower=0.0, cutoff_upper=5.0,
                 max_z=100, max_num_neighbors=32, aggr="add", dtype=torch.float32):
        super().__init__()
        assert rbf_type in rbf_class_mapping, f'Unknown RBF type "{rbf_type}".'
        assert activation in act_class_mapping, f'Unknown activation function "{activation}".'
        assert aggr in ["add", "mean", "max"], 'Argument aggr must be one of: "add", "mean", or "max"'
        if aggr != "add":
            raise NotImplementedError('aggr="%s": only "add" is built' % aggr)
        self.hidden_channels = hidden_channels
        self.num_filters = num_filters
        self.num_layers = num_layers
        self.num_rbf = num_rbf
        self.rbf_type = rbf_type
        self.trainable_rbf = trainable_rbf
        self.activation = activation
        self.cutoff_lower = cutoff_lower
        self.cutoff_upper = cutoff_upper
        self.max_z = max_z
        self.max_num_neighbors = max_num_neighbors
        self.aggr = aggr
        act_class = act_class_mapping[activation]
        self.embedding = nn.Embedding(self.max_z, hidden_channels, dtype=dtype)
        # the reference's legacy Distance(cutoff_lower, cutoff_upper, max_num_neighbors) keeps no parameters; this holder
        # gives the fused op its cutoffs / strategy / capacity.  The CSR rows carry one self entry per atom, hence + 1.
        self.distance = OptimizedDistance(cutoff_lower, cutoff_upper, max_num_pairs=-(max_num_neighbors + 1),
                                          return_vecs=True, loop=True, check_errors=True, long_edge_index=True)
        self.distance_expansion = rbf_class_mapping[rbf_type](cutoff_lower, cutoff_upper, num_rbf, trainable_rbf)
        self.neighbor_embedding = (NeighborEmbedding(hidden_channels, num_rbf, cutoff_lower, cutoff_upper, self.max_z,
                                                     dtype).jittable() if neighbor_embedding else None)
        self.interactions = nn.ModuleList()
        for _ in range(num_layers):
            self.interactions.append(InteractionBlock(hidden_channels, num_rbf, num_filters, act_class, cutoff_lower,
                                                      cutoff_upper, aggr=self.aggr, dtype=dtype))
        self._init_fused()
        self.gemm_mode = "tf32x3"
        self.reset_parameters()

    def reset_parameters(self):
        self.embedding.reset_parameters()
        self.distance_expansion.reset_parameters()
        if self.neighbor_embedding is not None:
            self.neighbor_embedding.reset_parameters()
        for interaction in self.interactions:
            interaction.reset_parameters()

    def _extra_key(self):
        return (self.gemm_mode, self.table_warp)

    def pack_weights(self, head: Optional[nn.Module], mean: float, std: float):
        """-> (weights in the order csrc/torch_shim.cpp::gn_create expects, iparams, fparams)."""
        if head is None:
            raise NotImplementedError("TorchMD_GN is evaluated together with its Scalar head")
        dev = self.embedding.weight.device
        H, F, K = self.hidden_channels, self.num_filters, self.num_rbf
        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()   # noqa: E731
        f64 = lambda t: t.detach().to(device=dev, dtype=torch.float64)                 # noqa: E731
        tr = lambda t: f32(t.detach().t())                                             # noqa: E731
        lo, hi = float(self.cutoff_lower), float(self.cutoff_upper)
        n_int = self.resolved_table_intervals()
        if self.rbf_type == "expnorm":
            p0, p1 = f64(self.distance_expansion.means), f64(self.distance_expansion.betas)
        else:
            p0, p1 = f64(self.distance_expansion.offset), f64(self.distance_expansion.coeff).expand(K)
        tgrid = TableGrid(n_int, lo, hi, self.rbf_type, self.table_warp)
        nodes = tgrid.nodes(dev)

        def table(layers, n_out):
            f, df = filter_table(nodes, self.rbf_type, p0, p1, lo, hi, [(f64(w), f64(b)) for w, b in layers],
                                 self.activation, False, cutoff=True)
            return f32(tgrid.coefficients(nodes, f, df).view(n_int, n_out, 4))

        empty = torch.empty(0, device=dev, dtype=torch.float32)
        ne = self.neighbor_embedding
        w = [f32(self.embedding.weight)]
        if ne is not None:
            w += [f32(ne.embedding.weight), table([(ne.distance_proj.weight, ne.distance_proj.bias)], H),
                  f32(ne.combine.weight), f32(ne.combine.bias), tr(ne.combine.weight[:, H:])]
        else:
            w += [empty] * 5
        h0, h2 = head.output_network[0], head.output_network[2]
        w += [f32(h0.weight), f32(h0.bias), tr(h0.weight), f32(h2.weight), f32(h2.bias)]
        for blk in self.interactions:
            mlp, conv = blk.mlp, blk.conv
            w += [f32(conv.lin1.weight), tr(conv.lin1.weight),
                  table([(mlp[0].weight, mlp[0].bias), (mlp[2].weight, mlp[2].bias)], F),
                  f32(conv.lin2.weight), f32(conv.lin2.bias), tr(conv.lin2.weight),
                  f32(blk.lin.weight), f32(blk.lin.bias), tr(blk.lin.weight)]
        iparams = [H, F, self.num_layers, self.max_z, ACT_ID[self.activation], int(ne is not None), n_int,
                   {"fp32": 0, "tf32x3": 1}[self.gemm_mode]]
        return w, iparams, [lo, hi, float(mean), float(std), tgrid.alpha, tgrid.gamma]

    def forward(self, z: Tensor, pos: Tensor, batch: Tensor, s: Optional[Tensor] = None, q: Optional[Tensor] = None):
        raise RuntimeError("TorchMD_GN is evaluated inside the fused CUDA energy+force kernels through "
                           "TorchMD_Net.forward (no standalone representation output, no CPU path)")
