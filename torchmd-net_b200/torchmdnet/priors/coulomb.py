"""``Coulomb`` under the reference's module path (torchmdnet/priors/coulomb.py); implemented in priors/pairwise.py."""
from .pairwise import Coulomb

__all__ = ["Coulomb"]
