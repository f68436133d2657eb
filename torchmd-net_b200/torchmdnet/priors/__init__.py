"""Prior models (reference: torchmdnet/priors/).  Built: ``Atomref`` (atom-wise reference energies; forces are not
affected, one gather + segment sum on the GPU) and the pairwise ``ZBL``, ``Coulomb`` and ``D2`` potentials (one fused CUDA
call for energy and analytic forces, priors/pairwise.py)."""
from .atomref import Atomref, BasePrior
from .pairwise import D2, ZBL, Coulomb

__all__ = ["Atomref", "D2", "ZBL", "Coulomb"]
