"""``ZBL`` and ``Coulomb`` with the reference's constructors, buffers and ``get_init_args`` (torchmdnet/priors/zbl.py:7-45,
priors/coulomb.py:7-36).  The reference evaluates them in ``post_reduce`` on the pair list of its legacy ``Distance`` module
and obtains their forces from the autograd pass over the whole model; here one fused call (``tmdn_pair_prior_f32``:
neighbour search with the prior's own cutoff, V(d) and V'(d) per pair) adds the energy to ``y`` and the analytic forces
to the forces of the fused model call.  Open boundaries, as the reference (``Distance`` takes no box)."""
from typing import Dict, Optional

import torch
from torch import Tensor

from .atomref import BasePrior


class _PairPrior(BasePrior):
    kind = ""

    def _weight(self, z: Tensor, extra_args: Optional[Dict[str, Tensor]]) -> Tensor:
        raise NotImplementedError

    def _params(self):
        raise NotImplementedError      # -> (cutoff, energy_prefactor, distance_scale, alpha)

    def add_energy_forces(self, y: Tensor, forces: Optional[Tensor], z: Tensor, pos: Tensor, batch: Tensor,
                          extra_args: Optional[Dict[str, Tensor]]) -> Tensor:
        """y[m] += E_prior(m) and forces -= dE_prior/dpos, in place on contiguous fp32 CUDA tensors; -> y."""
        from .. import _native
        _native.load()
        _native.require_cuda(pos, type(self).__name__)
        cutoff, pref, dscale, alpha = self._params()
        n = pos.shape[0]
        max_edges = (int(self.max_num_neighbors) + 1) * n          # + the self entry of every row
        w = self._weight(z, extra_args).to(torch.float32).contiguous()
        f = forces if forces is not None and forces.numel() > 0 else None
        found = torch.ops.torchmdnet_b200.pair_prior(self.kind, pos.detach().to(torch.float32).contiguous(), batch.contiguous(),
                                                     w, cutoff, pref, dscale, alpha, max_edges, y, f)
        if not torch.ops.torch_extension.is_stream_capturing() and int(found) > max_edges:
            # models/utils.py:419-425
            raise AssertionError("The neighbor search missed some atoms due to max_num_neighbors being too low. "
                                 "Please increase this parameter to include the maximum number of atoms within the cutoff.")
        return y


class ZBL(_PairPrior):
    """Ziegler-Biersack-Littmark screened nuclear repulsion (priors/zbl.py)."""
    kind = "zbl"

    def __init__(self, cutoff_distance, max_num_neighbors, atomic_number=None, distance_scale=None, energy_scale=None,
                 dataset=None):
        super().__init__()
        if atomic_number is None:
            atomic_number = dataset.atomic_number
        if distance_scale is None:
            distance_scale = dataset.distance_scale
        if energy_scale is None:
            energy_scale = dataset.energy_scale
        self.register_buffer("atomic_number", torch.as_tensor(atomic_number, dtype=torch.long))
        self.cutoff_distance = cutoff_distance
        self.max_num_neighbors = max_num_neighbors
        self.distance_scale = float(distance_scale)
        self.energy_scale = float(energy_scale)

    def get_init_args(self):
        return {"cutoff_distance": self.cutoff_distance, "max_num_neighbors": self.max_num_neighbors,
                "atomic_number": self.atomic_number.tolist(), "distance_scale": self.distance_scale,
                "energy_scale": self.energy_scale}

    def reset_parameters(self):
        pass

    def _weight(self, z, extra_args):
        return self.atomic_number[z]

    def _params(self):
        return (float(self.cutoff_distance), 2.30707755e-28 / self.energy_scale / self.distance_scale, self.distance_scale,
                0.0)


class Coulomb(_PairPrior):
    """Coulomb interaction of the partial charges in ``extra_args['partial_charges']``, damped by erf(alpha r)
    (priors/coulomb.py; alpha in 1/nm of the reference's internal unit = 1/(1e9 * distance_scale) per distance unit)."""
    kind = "coulomb"

    def __init__(self, alpha, max_num_neighbors, distance_scale=None, energy_scale=None, dataset=None):
        super().__init__()
        if distance_scale is None:
            distance_scale = dataset.distance_scale
        if energy_scale is None:
            energy_scale = dataset.energy_scale
        self.alpha = alpha
        self.max_num_neighbors = max_num_neighbors
        self.distance_scale = float(distance_scale)
        self.energy_scale = float(energy_scale)

    def get_init_args(self):
        return {"alpha": self.alpha, "max_num_neighbors": self.max_num_neighbors, "distance_scale": self.distance_scale,
                "energy_scale": self.energy_scale}

    def reset_parameters(self):
        pass

    def _weight(self, z, extra_args):
        if extra_args is None or "partial_charges" not in extra_args:
            raise KeyError("the Coulomb prior needs extra_args['partial_charges']")
        return extra_args["partial_charges"]

    def _params(self):
        # the reference converts positions to nm and alpha to 1/nm (coulomb.py:40-41): alpha' * D == alpha * d
        return (0.0, 2.30707e-28 / self.energy_scale / self.distance_scale, self.distance_scale, float(self.alpha))
