"""Energy heads (reference: torchmdnet/models/output_modules.py:13-115) as parameter trees.

``Scalar`` (H -> H/2 -> 1 MLP) is fused into the TensorNet kernels (csrc/tensornet.cu: out linear, head GEMMs,
per-molecule segmented sum); ``EquivariantScalar`` (two gated equivariant blocks) into the ET kernels.  Same
parameter names, shapes and initialisation as the reference.
"""
import torch
from torch import nn

from .utils import GatedEquivariantBlock, _fused_only, _xavier_zero_bias, act_class_mapping

__all__ = ["Scalar", "EquivariantScalar"]


class OutputModel(nn.Module):
    def __init__(self, allow_prior_model, reduce_op):
        super().__init__()
        self.allow_prior_model = allow_prior_model
        self.reduce_op = reduce_op
        self.dim_size = 0

    def reset_parameters(self):
        pass

    pre_reduce = _fused_only
    reduce = _fused_only

    def post_reduce(self, x):
        return x


class Scalar(OutputModel):
    """output_modules.py:49-77."""

    def __init__(self, hidden_channels, activation="silu", allow_prior_model=True, reduce_op="sum",
                 dtype=torch.float):
        super().__init__(allow_prior_model=allow_prior_model, reduce_op=reduce_op)
        act_class = act_class_mapping[activation]
        self.output_network = nn.Sequential(nn.Linear(hidden_channels, hidden_channels // 2, dtype=dtype),
                                            act_class(),
                                            nn.Linear(hidden_channels // 2, 1, dtype=dtype))
        self.reset_parameters()

    def reset_parameters(self):
        _xavier_zero_bias(self.output_network[0])
        _xavier_zero_bias(self.output_network[2])


class EquivariantScalar(OutputModel):
    """output_modules.py:80-115."""

    def __init__(self, hidden_channels, activation="silu", allow_prior_model=True, reduce_op="sum",
                 dtype=torch.float):
        super().__init__(allow_prior_model=allow_prior_model, reduce_op=reduce_op)
        self.output_network = nn.ModuleList([
            GatedEquivariantBlock(hidden_channels, hidden_channels // 2, activation=activation,
                                  scalar_activation=True, dtype=dtype),
            GatedEquivariantBlock(hidden_channels // 2, 1, activation=activation, dtype=dtype)])
        self.reset_parameters()

    def reset_parameters(self):
        for layer in self.output_network:
            layer.reset_parameters()
