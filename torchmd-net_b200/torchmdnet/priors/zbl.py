"""``ZBL`` under the reference's module path (torchmdnet/priors/zbl.py); implemented in priors/pairwise.py."""
from .pairwise import ZBL

__all__ = ["ZBL"]
