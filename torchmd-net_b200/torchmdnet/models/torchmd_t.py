"""TorchMD_T, the (non-equivariant) transformer (reference: torchmdnet/models/torchmd_t.py), as a parameter tree + weight
packer for the fused CUDA energy/force pipeline (csrc/et.cu, ``tmdn_t_energy_forces``).

Parameter names, shapes and RNG consumption order follow the reference (torchmd_t.py:60-142, 203-258), so ``state_dict``s
are interchangeable and a seeded ``create_model`` gives the same initial weights.  The distance filters ``dk_proj`` /
``dv_proj`` are univariate in d and are evaluated from cubic Hermite tables built in fp64 (``tensornet.TableGrid``)."""
from typing import Optional

import torch
from torch import Tensor, nn

from ._fused import FusedRepresentation
from .tensornet import TableGrid, _cutoff_and_grad, filter_table
from .utils import (ACT_ID, CosineCutoff, NeighborEmbedding, OptimizedDistance, _fused_only, act_class_mapping,
                    rbf_class_mapping)

__all__ = ["TorchMD_T"]


class MultiHeadAttention(nn.Module):
    """Parameters of torchmd_t.py:203-258."""

    def __init__(self, hidden_channels, num_rbf, distance_influence, num_heads, activation, attn_activation, cutoff_lower,
                 cutoff_upper, dtype=torch.float):
        super().__init__()
        assert hidden_channels % num_heads == 0, (
            f"The number of hidden channels ({hidden_channels}) must be evenly divisible by the number of "
            f"attention heads ({num_heads})")
        self.distance_influence = distance_influence
        self.num_heads = num_heads
        self.head_dim = hidden_channels // num_heads
        self.layernorm = nn.LayerNorm(hidden_channels, dtype=dtype)
        self.act = activation()
        self.attn_activation = attn_activation()
        self.cutoff = CosineCutoff(cutoff_lower, cutoff_upper)
        self.q_proj = nn.Linear(hidden_channels, hidden_channels, dtype=dtype)
        self.k_proj = nn.Linear(hidden_channels, hidden_channels, dtype=dtype)
        self.v_proj = nn.Linear(hidden_channels, hidden_channels, dtype=dtype)
        self.o_proj = nn.Linear(hidden_channels, hidden_channels, dtype=dtype)
        self.dk_proj = None
        if distance_influence in ["keys", "both"]:
            self.dk_proj = nn.Linear(num_rbf, hidden_channels, dtype=dtype)
        self.dv_proj = None
        if distance_influence in ["values", "both"]:
            self.dv_proj = nn.Linear(num_rbf, hidden_channels, dtype=dtype)
        self.reset_parameters()

    def reset_parameters(self):
        self.layernorm.reset_parameters()
        for lin in (self.q_proj, self.k_proj, self.v_proj, self.o_proj):
            nn.init.xavier_uniform_(lin.weight)
            lin.bias.data.fill_(0)
        if self.dk_proj:
            nn.init.xavier_uniform_(self.dk_proj.weight)
            self.dk_proj.bias.data.fill_(0)
        if self.dv_proj:
            nn.init.xavier_uniform_(self.dv_proj.weight)
            self.dv_proj.bias.data.fill_(0)

    def jittable(self):
        return self

    forward = _fused_only


class TorchMD_T(FusedRepresentation):
    """Drop-in for ``torchmdnet.models.torchmd_t.TorchMD_T`` (torchmd_t.py:15-183): same constructor, same parameters."""
    _op = "t"
    _kernel_name = "TorchMD_T"

    def __init__(self, hidden_channels=128, num_layers=6, num_rbf=50, rbf_type="expnorm", trainable_rbf=True,
                 activation="silu", attn_activation="silu", neighbor_embedding=True, num_heads=8,
                 distance_influence="both", cutoff_lower=0.0, cutoff_upper=5.0, max_z=100, max_num_neighbors=32,
                 dtype=torch.float):
        super().__init__()
        assert distance_influence in ["keys", "values", "both", "none"]
        assert rbf_type in rbf_class_mapping, f'Unknown RBF type "{rbf_type}".'
        assert activation in act_class_mapping, f'Unknown activation function "{activation}".'
        assert attn_activation in act_class_mapping, f'Unknown attention activation function "{attn_activation}".'
        self.hidden_channels = hidden_channels
        self.num_layers = num_layers
        self.num_rbf = num_rbf
        self.rbf_type = rbf_type
        self.trainable_rbf = trainable_rbf
        self.activation = activation
        self.attn_activation = attn_activation
        self.num_heads = num_heads
        self.distance_influence = distance_influence
        self.cutoff_lower = cutoff_lower
        self.cutoff_upper = cutoff_upper
        self.max_z = max_z
        self.max_num_neighbors = max_num_neighbors
        act_class = act_class_mapping[activation]
        attn_act_class = act_class_mapping[attn_activation]
        self.embedding = nn.Embedding(self.max_z, hidden_channels, dtype=dtype)
        # the reference's legacy Distance(..., loop=True): this holder gives the fused op its cutoffs / capacity (+ the self pair)
        self.distance = OptimizedDistance(cutoff_lower, cutoff_upper, max_num_pairs=-(max_num_neighbors + 1),
                                          return_vecs=True, loop=True, check_errors=True, long_edge_index=True)
        self.distance_expansion = rbf_class_mapping[rbf_type](cutoff_lower, cutoff_upper, num_rbf, trainable_rbf)
        self.neighbor_embedding = (NeighborEmbedding(hidden_channels, num_rbf, cutoff_lower, cutoff_upper, self.max_z,
                                                     dtype).jittable() if neighbor_embedding else None)
        self.attention_layers = nn.ModuleList()
        for _ in range(num_layers):
            self.attention_layers.append(MultiHeadAttention(hidden_channels, num_rbf, distance_influence, num_heads,
                                                            act_class, attn_act_class, cutoff_lower, cutoff_upper,
                                                            dtype=dtype).jittable())
        self.out_norm = nn.LayerNorm(hidden_channels, dtype=dtype)
        self._init_fused()
        self.gemm_mode = "tf32x3"
        self.reset_parameters()

    def reset_parameters(self):
        self.embedding.reset_parameters()
        self.distance_expansion.reset_parameters()
        if self.neighbor_embedding is not None:
            self.neighbor_embedding.reset_parameters()
        for attn in self.attention_layers:
            attn.reset_parameters()
        self.out_norm.reset_parameters()

    def _extra_key(self):
        return (self.gemm_mode, self.table_warp)

    def pack_weights(self, head: Optional[nn.Module], mean: float, std: float):
        """-> (weights in the order csrc/torch_shim.cpp::t_create expects, iparams, fparams)."""
        if head is None:
            raise NotImplementedError("TorchMD_T is evaluated together with its Scalar head")
        dev = self.embedding.weight.device
        H, K = self.hidden_channels, self.num_rbf
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

        def table(lin, act, cutoff):
            """[n_int, H, 4] coefficients of [act o] lin(rbf(d)) [* C(d)]; a missing projection is the constant 1"""
            if lin is None:
                if cutoff:
                    cc, dcc = _cutoff_and_grad(nodes, lo, hi)
                    return tgrid.coefficients(nodes, cc[:, None].expand(-1, H), dcc[:, None].expand(-1, H))
                c = torch.zeros(n_int, H, 4, device=dev, dtype=torch.float64)
                c[..., 0] = 1.0
                return c
            f, df = filter_table(nodes, self.rbf_type, p0, p1, lo, hi, [(f64(lin.weight), f64(lin.bias))], act,
                                 act is not None, cutoff=cutoff)
            return tgrid.coefficients(nodes, f, df)

        empty = torch.empty(0, device=dev, dtype=torch.float32)
        ne = self.neighbor_embedding
        w = [f32(self.embedding.weight)]
        if ne is not None:
            w += [f32(ne.embedding.weight), f32(table(ne.distance_proj, None, True)), f32(ne.combine.weight),
                  f32(ne.combine.bias), tr(ne.combine.weight[:, H:])]
        else:
            w += [empty] * 5
        w += [f32(self.out_norm.weight), f32(self.out_norm.bias)]
        h0, h2 = head.output_network[0], head.output_network[2]
        w += [f32(h0.weight), f32(h0.bias), tr(h0.weight), f32(h2.weight), f32(h2.bias)]
        for ly in self.attention_layers:
            qkv_w = torch.cat([ly.q_proj.weight, ly.k_proj.weight, ly.v_proj.weight]).detach()
            qkv_b = torch.cat([ly.q_proj.bias, ly.k_proj.bias, ly.v_proj.bias]).detach()
            # the cutoff C(d) multiplies the attention weight, i.e. every value term: it is folded into the dv table
            tab = torch.stack([table(ly.dk_proj, self.activation, False), table(ly.dv_proj, self.activation, True)], 1)
            w += [f32(ly.layernorm.weight), f32(ly.layernorm.bias), f32(qkv_w), f32(qkv_b), tr(qkv_w),
                  f32(ly.o_proj.weight), f32(ly.o_proj.bias), tr(ly.o_proj.weight), f32(tab.view(n_int, 2, H, 4))]
        iparams = [H, self.num_layers, self.num_heads, self.max_z, ACT_ID[self.activation], ACT_ID[self.attn_activation],
                   int(ne is not None), n_int, {"fp32": 0, "tf32x3": 1}[self.gemm_mode]]
        return w, iparams, [lo, hi, float(mean), float(std), tgrid.alpha, tgrid.gamma]

    def forward(self, z: Tensor, pos: Tensor, batch: Tensor, s: Optional[Tensor] = None, q: Optional[Tensor] = None):
        raise RuntimeError("TorchMD_T is evaluated inside the fused CUDA energy+force kernels through "
                           "TorchMD_Net.forward (no standalone representation output, no CPU path)")
