"""Equivariant Transformer (reference: torchmdnet/models/torchmd_et.py) as a parameter tree for the fused
CUDA energy/force pipeline (csrc/et.cu).  Parameter names, shapes and RNG consumption order follow
torchmd_et.py:55-152 and :208-270 so state_dicts are interchangeable."""
from typing import Optional

import torch
from torch import Tensor, nn

from .utils import (CosineCutoff, NeighborEmbedding, OptimizedDistance, _fused_only, _xavier_zero_bias,
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


class TorchMD_ET(nn.Module):
    """Drop-in constructor for ``torchmdnet.models.torchmd_et.TorchMD_ET`` (torchmd_et.py:55-152)."""

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
        self.reset_parameters()

    def reset_parameters(self):
        self.embedding.reset_parameters()
        self.distance_expansion.reset_parameters()
        if self.neighbor_embedding is not None:
            self.neighbor_embedding.reset_parameters()
        for attn in self.attention_layers:
            attn.reset_parameters()
        self.out_norm.reset_parameters()

    forward = _fused_only

    def energy_forces(self, *a, **k):
        raise NotImplementedError("the fused Equivariant Transformer kernels (csrc/et.cu) are not built yet; "
                                  "there is no PyTorch or CPU fallback")
