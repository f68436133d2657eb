"""Prior models (reference: torchmdnet/priors/).  Built: ``Atomref`` (atom-wise reference energies; forces are not
affected, so it composes with the fused energy+force call as one gather + segment sum on the GPU).  The pairwise priors
(ZBL, D2, Coulomb) are outside this build and are refused by ``create_model``."""
from .atomref import Atomref

__all__ = ["Atomref"]
