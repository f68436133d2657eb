"""Equivariant Transformer (reference: torchmdnet/models/torchmd_et.py) as a parameter tree for the fused
CUDA energy/force pipeline (csrc/et.cu).  Parameter names, shapes and RNG consumption order follow
torchmd_et.py:55-152 and :208-270 so state_dicts are interchangeable."""
from typing import Optional

import torch
from torch import Tensor, nn

from ._fused import FusedRepresentation
from .tensornet import TableGrid, filter_table, hermite_coefficients
from .utils import (ACT_ID, CosineCutoff, NeighborEmbedding, OptimizedDistance, _fused_only, _xavier_zero_bias,
                    act_class_mapping, rbf_class_mapping)

__all__ = ["TorchMD_ET"]


class EquivariantMultiHeadAttention(nn.Module):
    """Parameters of torchmd_et.py:208-270."""

    def __init__(self, hidden_channels, num_rbf, distance_influence, num_heads, activation, attn_activation,
                 cutoff_lower, cutoff_upper, dtype=torch.float32):
        super().__init__()
        assert hidden_channels % num_heads == 0, (
            f"The number of hidden channels ({hidden_channels}) must be evenly divisible by the number of "
            f"attention heads ({num_heads})")
        H = hidden_channels
        self.distance_influence = distance_influence
        self.num_heads = num_heads
        self.hidden_channels = H
        self.head_dim = H // num_heads
        self.layernorm = nn.LayerNorm(H, dtype=dtype)
        self.act = activation()
        self.attn_activation = act_class_mapping[attn_activation]()
        self.cutoff = CosineCutoff(cutoff_lower, cutoff_upper)
        self.q_proj = nn.Linear(H, H, dtype=dtype)
        self.k_proj = nn.Linear(H, H, dtype=dtype)
        self.v_proj = nn.Linear(H, H * 3, dtype=dtype)
        self.o_proj = nn.Linear(H, H * 3, dtype=dtype)
        self.vec_proj = nn.Linear(H, H * 3, bias=False, dtype=dtype)
        self.dk_proj = nn.Linear(num_rbf, H, dtype=dtype) if distance_influence in ["keys", "both"] else None
        self.dv_proj = nn.Linear(num_rbf, H * 3, dtype=dtype) if distance_influence in ["values", "both"] else None
        self.reset_parameters()

    def reset_parameters(self):
        self.layernorm.reset_parameters()
        for lin in (self.q_proj, self.k_proj, self.v_proj, self.o_proj, self.vec_proj):
            _xavier_zero_bias(lin)
        if self.dk_proj is not None:
            _xavier_zero_bias(self.dk_proj)
        if self.dv_proj is not None:
            _xavier_zero_bias(self.dv_proj)

    def jittable(self):
        return self

    forward = _fused_only


class TorchMD_ET(FusedRepresentation):
    """Drop-in for ``torchmdnet.models.torchmd_et.TorchMD_ET`` (torchmd_et.py:55-152): same constructor, same
    parameters; evaluated by the fused kernels of csrc/et.cu.  The distance filters (dk_proj, dv_proj, the
    neighbour-embedding distance_proj) are univariate in d and are evaluated from cubic Hermite tables built in fp64
    from the weights on the grid of ``tensornet.TableGrid`` (``table_intervals`` intervals over [0, cutoff_upper],
    default max(1024, 32 * num_rbf))."""
    _op = "et"
    _kernel_name = "TorchMD_ET"

    def __init__(self, hidden_channels=128, num_layers=6, num_rbf=50, rbf_type="expnorm", trainable_rbf=True,
                 activation="silu", attn_activation="silu", neighbor_embedding=True, num_heads=8,
                 distance_influence="both", cutoff_lower=0.0, cutoff_upper=5.0, max_z=100, max_num_neighbors=32,
                 dtype=torch.float32):
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
        self.dtype = dtype
        act_class = act_class_mapping[activation]
        self.embedding = nn.Embedding(max_z, hidden_channels, dtype=dtype)
        self.distance = OptimizedDistance(cutoff_lower, cutoff_upper, max_num_pairs=-max_num_neighbors,
                                          return_vecs=True, loop=True, long_edge_index=True)
        self.distance_expansion = rbf_class_mapping[rbf_type](cutoff_lower, cutoff_upper, num_rbf, trainable_rbf)
        self.neighbor_embedding = (NeighborEmbedding(hidden_channels, num_rbf, cutoff_lower, cutoff_upper, max_z,
                                                     dtype) if neighbor_embedding else None)
        self.attention_layers = nn.ModuleList([
            EquivariantMultiHeadAttention(hidden_channels, num_rbf, distance_influence, num_heads, act_class,
                                          attn_activation, cutoff_lower, cutoff_upper, dtype)
            for _ in range(num_layers)])
        self.out_norm = nn.LayerNorm(hidden_channels, dtype=dtype)
        self._init_fused()
        # projections: "tf32x3" (default) = tcgen05 three-product tf32 with the K chunks summed in fp32 outside the
        # tensor core (rms error below an fp32 FFMA GEMM); "fp32" = fp32 FFMA kernel
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

    # -------------------------------------------------------------------------------------------- packing
    def _extra_key(self):
        return (self.gemm_mode, self.table_warp)

    def pack_weights(self, head: Optional[nn.Module], mean: float, std: float):
        """-> (weights in the order csrc/torch_shim.cpp::et_create expects, iparams, fparams)."""
        if head is None:
            raise NotImplementedError("TorchMD_ET is evaluated together with its EquivariantScalar head")
        dev = self.embedding.weight.device
        H, nh, K = self.hidden_channels, self.num_heads, self.num_rbf
        hd = H // nh
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
        grid = tgrid.nodes(dev)

        def table(lin, act, n_out, cutoff):
            """[n_int, n_out, 4] Hermite coefficients of [act o] lin(rbf(d)) [* C(d)]; the constant 1 if lin is None"""
            if lin is None:
                if cutoff:   # the bare cutoff C(d)
                    from .tensornet import _cutoff_and_grad
                    cc, dcc = _cutoff_and_grad(grid, lo, hi)
                    return tgrid.coefficients(grid, cc[:, None].expand(-1, n_out), dcc[:, None].expand(-1, n_out))
                c = torch.zeros(n_int, n_out, 4, device=dev, dtype=torch.float64)
                c[..., 0] = 1.0
                return c
            f, df = filter_table(grid, self.rbf_type, p0, p1, lo, hi, [(f64(lin[0]), f64(lin[1]))],
                                 act, act is not None, cutoff=cutoff)
            return tgrid.coefficients(grid, f, df)

        # part-major row order of v / dv: row p*H + head*hd + c <- reference row head*3hd + p*hd + c
        perm = torch.arange(3 * H, device=dev).view(nh, 3, hd).permute(1, 0, 2).reshape(-1)
        empty = torch.empty(0, device=dev, dtype=torch.float32)
        ne = self.neighbor_embedding
        w = [f32(self.embedding.weight)]
        if ne is not None:
            ne_tab = table((ne.distance_proj.weight, ne.distance_proj.bias), None, H, True)
            w += [f32(ne.embedding.weight), f32(ne_tab), f32(ne.combine.weight), f32(ne.combine.bias),
                  tr(ne.combine.weight[:, H:])]
        else:
            w += [empty] * 5
        w += [f32(self.out_norm.weight), f32(self.out_norm.bias)]
        for blk in head.output_network:
            u0, u2 = blk.update_net[0], blk.update_net[2]
            w += [f32(blk.vec1_proj.weight), tr(blk.vec1_proj.weight), f32(blk.vec2_proj.weight),
                  tr(blk.vec2_proj.weight), f32(u0.weight), f32(u0.bias), tr(u0.weight), f32(u2.weight), f32(u2.bias),
                  tr(u2.weight)]
        for ly in self.attention_layers:
            qkv_w = torch.cat([ly.q_proj.weight, ly.k_proj.weight, ly.v_proj.weight[perm]]).detach()
            qkv_b = torch.cat([ly.q_proj.bias, ly.k_proj.bias, ly.v_proj.bias[perm]]).detach()
            dk = None if ly.dk_proj is None else (ly.dk_proj.weight, ly.dk_proj.bias)
            dv = None if ly.dv_proj is None else (ly.dv_proj.weight[perm], ly.dv_proj.bias[perm])
            # the attention cutoff C(d) (torchmd_et.py:297) only ever multiplies the scalar value filter (part 0 of dv),
            # so it is folded into that table: the edge kernels evaluate no cosine
            dv0 = None if dv is None else (dv[0][:H], dv[1][:H])
            dv12 = None if dv is None else (dv[0][H:], dv[1][H:])
            tab = torch.cat([table(dk, self.activation, H, False), table(dv0, self.activation, H, True),
                             table(dv12, self.activation, 2 * H, False)], 1)
            w += [f32(ly.layernorm.weight), f32(ly.layernorm.bias), f32(qkv_w), f32(qkv_b), tr(qkv_w),
                  f32(ly.vec_proj.weight), tr(ly.vec_proj.weight), f32(ly.o_proj.weight), f32(ly.o_proj.bias),
                  tr(ly.o_proj.weight), f32(tab.view(n_int, 4, H, 4))]
        iparams = [H, self.num_layers, nh, self.max_z, ACT_ID[self.activation], ACT_ID[self.attn_activation],
                   int(ne is not None), n_int, {"fp32": 0, "tf32x3": 1}[self.gemm_mode]]
        return w, iparams, [lo, hi, float(mean), float(std), tgrid.alpha, tgrid.gamma]

    def forward(self, z: Tensor, pos: Tensor, batch: Tensor, q: Optional[Tensor] = None,
                s: Optional[Tensor] = None):
        raise RuntimeError("TorchMD_ET is evaluated inside the fused CUDA energy+force kernels through "
                           "TorchMD_Net.forward (no standalone representation output, no CPU path)")
