"""``D2`` under the reference's module path (torchmdnet/priors/d2.py); implemented in priors/pairwise.py."""
from .pairwise import D2

__all__ = ["D2"]
