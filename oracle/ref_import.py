"""TEST INFRASTRUCTURE ONLY -- imports the UNMODIFIED reference package from /root/reference.

Usable only in the build container (``/root/reference`` does not exist on the GPU box).  It is
used by ``oracle/make_golden.py`` to generate the committed fixtures under ``tests/golden/`` and by
the ``-m "not gpu"`` tests that cross-check the oracle restatement when the reference is present.

Nothing is copied from the reference: the package is imported from where it lies, and the four
third-party modules that are absent from this image are provided as tiny
in-memory stand-ins whose arithmetic is index-gather and sum-scatter only:

* ``torch_scatter.scatter``            (call sites: tensornet.py:313-315,331, torchmd_et.py:345-346,
                                        output_modules.py:43)
* ``torch_geometric.nn.MessagePassing`` (models/utils.py:43, torchmd_et.py:208)
* ``torch_cluster.radius_graph``        (models/utils.py:410, legacy Distance only)
* ``lightning_utilities.core.rank_zero.rank_zero_warn`` (models/model.py:11)

The reference's CPU neighbour op is the .so built by ``oracle/build_ref.sh`` from the reference's
own neighbors.cpp + neighbors_cpu.cpp.  NB: that .so defines TORCH_LIBRARY(torchmdnet_neighbors),
the same namespace the product extension defines, so this module must never be imported in a
process that also loads the product's torch extension.
"""
import inspect
import os
import sys
import types
import warnings

import torch

REF_ROOT = os.environ.get("TMDN_REFERENCE", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "torchmdnet_neighbors_ref.so")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "torchmdnet")) and os.path.isfile(REF_SO)


def _scatter(src, index, dim=-1, out=None, dim_size=None, reduce="sum"):
    if dim < 0:
        dim += src.dim()
    if dim_size is None:
        dim_size = int(index.max()) + 1 if index.numel() else 0
    shape = list(src.shape)
    shape[dim] = dim_size
    res = torch.zeros(shape, dtype=src.dtype, device=src.device).index_add(dim, index, src)
    if reduce in ("mean",):
        cnt = torch.zeros(dim_size, dtype=src.dtype, device=src.device).index_add(
            0, index, torch.ones_like(index, dtype=src.dtype))
        view = [1] * src.dim()
        view[dim] = dim_size
        res = res / cnt.clamp(min=1).view(view)
    elif reduce not in ("sum", "add"):
        raise NotImplementedError(reduce)
    return res


class _MessagePassing(torch.nn.Module):
    """source_to_target flow: `_j` gathered at edge_index[0], `_i` at edge_index[1]; aggregate at [1]."""

    def __init__(self, aggr="add", flow="source_to_target", node_dim=-2):
        super().__init__()
        assert aggr == "add" and flow == "source_to_target"
        self.node_dim = node_dim

    def jittable(self):
        return self

    def propagate(self, edge_index, size=None, **kwargs):
        names = list(inspect.signature(self.message).parameters)
        n = None
        for v in kwargs.values():
            if isinstance(v, torch.Tensor) and n is None:
                n = v.size(self.node_dim)
        args = []
        for name in names:
            if name.endswith("_i") or name.endswith("_j"):
                base, which = name[:-2], name[-1]
                if base in kwargs:
                    t = kwargs[base]
                    n = t.size(self.node_dim)
                    idx = edge_index[1] if which == "i" else edge_index[0]
                    args.append(t.index_select(self.node_dim, idx))
                    continue
            args.append(kwargs[name])
        out = self.message(*args)
        out = self.aggregate(out, edge_index[1], None, n)
        return self.update(out)

    def aggregate(self, inputs, index, ptr=None, dim_size=None):
        return _scatter(inputs, index, dim=self.node_dim, dim_size=dim_size)

    def update(self, inputs):
        return inputs


def _radius_graph(x, r, batch=None, loop=False, max_num_neighbors=32, flow="source_to_target"):
    d = torch.cdist(x, x)
    m = d < r
    if batch is not None:
        m &= batch[:, None] == batch[None, :]
    if not loop:
        m &= ~torch.eye(x.size(0), dtype=torch.bool)
    row, col = m.nonzero(as_tuple=True)
    return torch.stack([col, row])


def _install_shims():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    if "torch_scatter" not in sys.modules:
        mod("torch_scatter", scatter=_scatter)
    if "torch_geometric" not in sys.modules:
        nn = mod("torch_geometric.nn", MessagePassing=_MessagePassing)
        mod("torch_geometric", nn=nn)
    if "torch_cluster" not in sys.modules:
        mod("torch_cluster", radius_graph=_radius_graph)
    if "lightning_utilities" not in sys.modules:
        rz = mod("lightning_utilities.core.rank_zero", rank_zero_warn=warnings.warn)
        core = mod("lightning_utilities.core", rank_zero=rz)
        mod("lightning_utilities", core=core)
    # models/utils.py:524-566 JIT-compiles torch_extension::is_stream_capturing at import unless the op
    # already exists; on CPU it always answers False, so define exactly that.
    if not hasattr(torch.ops.torch_extension, "is_stream_capturing"):
        lib = torch.library.Library("torch_extension", "DEF")
        lib.define("is_stream_capturing() -> bool")
        lib.impl("is_stream_capturing", lambda: False, "CompositeExplicitAutograd")
        _install_shims._lib = lib


def import_reference():
    """Returns the reference `torchmdnet` package (imported from REF_ROOT, unmodified)."""
    if not available():
        raise RuntimeError("reference not available (need %s and %s)" % (REF_ROOT, REF_SO))
    if "torchmdnet" in sys.modules and getattr(sys.modules["torchmdnet"], "__tmdn_ref__", False):
        return sys.modules["torchmdnet"]
    assert "torchmdnet" not in sys.modules, "product torchmdnet already imported in this process"
    _install_shims()
    torch.ops.load_library(REF_SO)
    # neighbors/__init__.py:6-14 searches for the .so inside its own (read-only) directory; provide the
    # module it would have produced instead.
    nb = types.ModuleType("torchmdnet.neighbors")
    nb.get_neighbor_pairs_kernel = torch.ops.torchmdnet_neighbors.get_neighbor_pairs
    sys.modules["torchmdnet.neighbors"] = nb
    sys.path.insert(0, REF_ROOT)
    try:
        import torchmdnet  # noqa
        from torchmdnet.models import model as _m  # noqa
    finally:
        sys.path.remove(REF_ROOT)
    torchmdnet.__tmdn_ref__ = True
    torchmdnet.neighbors = nb
    return torchmdnet


def seed_everything(seed):
    import random
    import numpy as np
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
