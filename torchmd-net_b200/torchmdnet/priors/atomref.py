"""``Atomref`` with the reference's constructor, buffers and ``state_dict`` keys (torchmdnet/priors/atomref.py:8-44):
``pre_reduce`` adds ``atomref[z]`` to every atom's (already std-scaled) energy before the molecular sum
(models/model.py:266-269).  The reference energies do not depend on positions, so forces are unchanged and the prior
is applied to the fused kernels' output as ``y[m] += sum_{i in m} atomref[z_i]``."""
from typing import Dict, Optional

import torch
from torch import Tensor, nn


class BasePrior(nn.Module):
    """torchmdnet/priors/base.py: the hooks a prior may implement."""

    def __init__(self, dataset=None):
        super().__init__()

    def get_init_args(self):
        return {}

    def pre_reduce(self, x, z, pos, batch, extra_args: Optional[Dict[str, Tensor]]):
        return x

    def post_reduce(self, y, z, pos, batch, extra_args: Optional[Dict[str, Tensor]]):
        return y


class Atomref(BasePrior):
    def __init__(self, max_z=None, dataset=None):
        super().__init__()
        if max_z is None and dataset is None:
            raise ValueError("Can't instantiate Atomref prior, all arguments are None.")
        if dataset is None:
            atomref = torch.zeros(max_z, 1)
        else:
            atomref = dataset.get_atomref()
            if atomref is None:
                import warnings
                warnings.warn("The atomref returned by the dataset is None, defaulting to zeros with max. "
                              "atomic number 99. Maybe atomref is not defined for the current target.")
                atomref = torch.zeros(100, 1)
        if atomref.ndim == 1:
            atomref = atomref.view(-1, 1)
        self.register_buffer("initial_atomref", atomref)
        self.atomref = nn.Embedding(len(atomref), 1)
        self.atomref.weight.data.copy_(atomref)

    def reset_parameters(self):
        self.atomref.weight.data.copy_(self.initial_atomref)

    def get_init_args(self):
        return dict(max_z=self.initial_atomref.size(0))

    def pre_reduce(self, x: Tensor, z: Tensor, pos: Tensor, batch: Tensor, extra_args: Optional[Dict[str, Tensor]]):
        return x + self.atomref(z)

    def molecular_sum(self, z: Tensor, batch: Tensor, n_mol: int) -> Tensor:
        """sum_{i in molecule} atomref[z_i] -> [n_mol, 1]: what pre_reduce + reduce add to y."""
        out = torch.zeros(n_mol, 1, dtype=self.atomref.weight.dtype, device=z.device)
        if z.is_cuda and out.dtype == torch.float32:
            # fixed summation order (index_add_ on CUDA adds with atomics): tmdn_segment_add_f32, batch sorted as everywhere
            from .. import _native
            _native.load()
            torch.ops.torchmdnet_b200.segment_add(self.atomref(z).detach().reshape(-1).contiguous(), batch.contiguous(),
                                                  out.view(-1))
            return out
        return out.index_add_(0, batch, self.atomref(z))
