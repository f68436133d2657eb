"""``get_neighbor_pairs_kernel`` -- the reference's handle on its native op
(torchmdnet/neighbors/__init__.py:6-17).  Here the op ``torchmdnet_neighbors::get_neighbor_pairs`` is
registered (same namespace, name and schema, torchmdnet/neighbors/neighbors.cpp:3-5) by the torch shim of
libtmdn and resolved lazily, so importing the package does not need the built library; calling does."""
import torch

from .. import _native


def get_neighbor_pairs_kernel(strategy, positions, batch, box_vectors, use_periodic, cutoff_lower, cutoff_upper,
                              max_num_pairs, loop, include_transpose):
    """-> (neighbors int32[2,P], deltas[P,3], distances[P], num_found int32[1]); padded with -1 / 0 / 0."""
    _native.load()
    return torch.ops.torchmdnet_neighbors.get_neighbor_pairs(strategy, positions, batch, box_vectors, use_periodic,
                                                             cutoff_lower, cutoff_upper, max_num_pairs, loop,
                                                             include_transpose)


try:  # keep the op out of dynamo graphs, as the reference does (neighbors/__init__.py:15-17)
    import torch._dynamo as _dynamo

    _dynamo.disallow_in_graph(get_neighbor_pairs_kernel)
except Exception:  # pragma: no cover
    pass
