"""``ZBL``, ``Coulomb`` and ``D2`` with the reference's constructors, buffers and ``get_init_args``
(torchmdnet/priors/zbl.py:7-45, priors/coulomb.py:7-36, priors/d2.py:7-161).  The reference evaluates them in ``post_reduce`` on the pair list of its legacy ``Distance`` module
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

    def post_reduce(self, y: Tensor, z: Tensor, pos: Tensor, batch: Tensor, extra_args: Optional[Dict[str, Tensor]]):
        """The reference's hook (priors/base.py): y + E_prior per molecule, energy only (CUDA tensors; no autograd graph --
        the forces of a model evaluation come from ``add_energy_forces``)."""
        e = torch.zeros(y.numel(), dtype=torch.float32, device=pos.device)
        self.add_energy_forces(e, None, z, pos, batch, extra_args)
        return y + e.view(y.shape).to(y.dtype)


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


# Grimme, J. Comput. Chem. 27 (2006) 1787, Table 1: C_6 in J nm^6 / mol and van der Waals radius R_0 in Angstrom for
# H .. Xe (index = atomic number; the transition metals of a row share one entry).
_D2_C6 = ([float("nan"), 0.14, 0.08, 1.61, 1.61, 3.13, 1.75, 1.23, 0.70, 0.75, 0.63, 5.71, 5.71, 10.79, 9.23, 7.84, 5.57, 5.07,
           4.61] + [10.80] * 12 + [16.99, 17.10, 16.37, 12.64, 12.47, 12.01] + [24.67] * 12
          + [37.32, 38.71, 38.44, 31.74, 31.50, 29.99])
_D2_R0 = ([float("nan"), 1.001, 1.012, 0.825, 1.408, 1.485, 1.452, 1.397, 1.342, 1.287, 1.243, 1.144, 1.364, 1.639, 1.716,
           1.705, 1.683, 1.639, 1.595, 1.485, 1.474] + [1.562] * 10 + [1.650, 1.727, 1.760, 1.771, 1.749, 1.727, 1.628, 1.606]
          + [1.639] * 10 + [1.672, 1.804, 1.881, 1.892, 1.892, 1.881])


class D2(_PairPrior):
    """Grimme's DFT-D2 dispersion correction (priors/d2.py): E = -s_6/2 sum_pairs sqrt(C6_i C6_j) / R^6 * f_damp(R).
    Same constructor, buffers (``Z_map``, ``C_6``, ``R_r``) and ``get_init_args`` as the reference."""
    kind = "d2"

    def __init__(self, cutoff_distance, max_num_neighbors, atomic_number=None, distance_scale=None, energy_scale=None,
                 dataset=None, dtype=torch.float32):
        super().__init__()
        one = torch.tensor(1.0, dtype=dtype).item()
        self.cutoff_distance = cutoff_distance * one
        self.max_num_neighbors = int(max_num_neighbors)
        self.atomic_number = list(dataset.atomic_number if atomic_number is None else atomic_number)
        self.distance_scale = one * (dataset.distance_scale if distance_scale is None else distance_scale)
        self.energy_scale = one * (dataset.energy_scale if energy_scale is None else energy_scale)
        assert len(_D2_C6) == 55 and len(_D2_R0) == 55
        self.register_buffer("Z_map", torch.tensor(self.atomic_number, dtype=torch.long))
        self.register_buffer("C_6", torch.tensor(_D2_C6, dtype=torch.float64).to(dtype))
        self.register_buffer("R_r", (torch.tensor(_D2_R0, dtype=torch.float64) * 0.1).to(dtype))      # Angstrom -> nm
        self.d = 20
        self.s_6 = 1

    def reset_parameters(self):
        pass

    def get_init_args(self):
        return {"cutoff_distance": self.cutoff_distance, "max_num_neighbors": self.max_num_neighbors,
                "atomic_number": self.atomic_number, "distance_scale": self.distance_scale,
                "energy_scale": self.energy_scale}

    def _weight(self, z, extra_args):
        zz = self.Z_map[z]
        return torch.stack([self.C_6[zz].to(torch.float32).sqrt(), self.R_r[zz].to(torch.float32)])

    def _params(self):
        # d2.py:167-168,188-191: internal units nm and J/mol
        return (float(self.cutoff_distance), -float(self.s_6) / (self.energy_scale * 6.02214076e23), self.distance_scale,
                float(self.d))
