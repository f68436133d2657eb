"""``BasePrior`` under the reference's module path (torchmdnet/priors/base.py)."""
from .atomref import BasePrior

__all__ = ["BasePrior"]
