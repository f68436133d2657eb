"""Prior models (reference: torchmdnet/priors/).  Built: ``Atomref`` (atom-wise reference energies; forces are not
affected, one gather + segment sum on the GPU) and the pairwise ``ZBL`` and ``Coulomb`` potentials (one fused CUDA call for
energy and analytic forces, priors/pairwise.py).  ``D2`` is outside this build and is refused by ``create_model``."""
from .atomref import Atomref, BasePrior
from .pairwise import ZBL, Coulomb

__all__ = ["Atomref", "ZBL", "Coulomb"]
