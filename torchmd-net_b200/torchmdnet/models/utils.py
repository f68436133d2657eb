"""Host-side mirror of the reference's ``torchmdnet/models/utils.py`` for the energy+force path.

* ``OptimizedDistance``  -- same constructor/forward contract as models/utils.py:112-269, over the CUDA op.
* ``ExpNormalSmearing`` / ``GaussianSmearing`` / ``CosineCutoff`` / ``ShiftedSoftplus`` -- parameter holders
  with the reference's buffers and initial values (models/utils.py:272-390).  Inside the models these
  functions are evaluated by the fused kernels; the module ``forward``s are elementwise torch expressions kept
  for callers that use the modules on their own (tests/test_rbfs.py of the reference).
* ``NeighborEmbedding`` / ``GatedEquivariantBlock`` -- parameter trees (same names, shapes, init order as
  models/utils.py:43-108, 456-522) consumed by the fused ET kernels.
"""
import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from .. import _native
from ..neighbors import get_neighbor_pairs_kernel  # noqa: F401  (re-exported, as in the reference)


class OptimizedDistance(nn.Module):
    """Neighbor pairs within ``[cutoff_lower, cutoff_upper)`` (models/utils.py:112-269).

    ``max_num_pairs < 0`` means ``-max_num_pairs`` pairs per atom; ``resize_to_fit`` trims the ``-1`` padding
    (one device->host sync); ``check_errors`` raises when more pairs were found than fit (one sync).  With
    both off the call is CUDA-graph capturable.  ``strategy`` is "brute", "shared" or "cell"; "brute" and
    "shared" run the same batch-aware tiled kernel here and have no atom-count limit.
    """

    def __init__(self, cutoff_lower=0.0, cutoff_upper=5.0, max_num_pairs=-32, return_vecs=False, loop=False,
                 strategy="brute", include_transpose=True, resize_to_fit=True, check_errors=True, box=None,
                 long_edge_index=True):
        super().__init__()
        self.cutoff_upper = cutoff_upper
        self.cutoff_lower = cutoff_lower
        self.max_num_pairs = max_num_pairs
        self.strategy = strategy
        self.box: Optional[Tensor] = box
        self.loop = loop
        self.return_vecs = return_vecs
        self.include_transpose = include_transpose
        self.resize_to_fit = resize_to_fit
        self.use_periodic = box is not None
        if box is None:
            self.box = torch.empty((0, 0))
            if strategy == "cell":      # binning box for the open-boundary cell list (models/utils.py:199-202)
                lbox = cutoff_upper * 3.0
                self.box = torch.tensor([[lbox, 0, 0], [0, lbox, 0], [0, 0, lbox]])
        self.box = self.box.cpu()       # the op reads the box on the host
        self.check_errors = check_errors
        self.long_edge_index = long_edge_index

    def __prepare_scriptable__(self):
        """``torch.jit.script(OptimizedDistance(...))`` (reference: tests/test_neighbors.py:471-537) resolves the op while
        compiling ``forward``: the native library has to be loaded by then."""
        _native.load()
        return self

    def forward(self, pos: Tensor, batch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Optional[Tensor]]:
        if not torch.jit.is_scripting():
            _native.load()
        if not pos.is_cuda:
            raise RuntimeError("OptimizedDistance: this build runs on CUDA (sm_100a) only; got a CPU tensor and there is no "
                               "CPU fallback")
        box = self.box
        assert box is not None
        box = box.to(pos.dtype)
        max_pairs = self.max_num_pairs
        if max_pairs < 0:
            max_pairs = -max_pairs * pos.shape[0]
        if batch is None:
            batch = torch.zeros(pos.shape[0], dtype=torch.long, device=pos.device)
        edge_index, edge_vec, edge_weight, num_pairs = torch.ops.torchmdnet_neighbors.get_neighbor_pairs(
            self.strategy, pos, batch, box, self.use_periodic, self.cutoff_lower, self.cutoff_upper, max_pairs, self.loop,
            self.include_transpose)
        if self.check_errors:
            found = int(num_pairs[0].item())
            if found > max_pairs:
                raise RuntimeError("Found num_pairs(" + str(found) + ") > max_num_pairs(" + str(max_pairs) + ")")
        if self.resize_to_fit:
            keep = edge_index[0] != -1
            edge_index = edge_index[:, keep]
            edge_weight = edge_weight[keep]
            edge_vec = edge_vec[keep, :]
        if self.long_edge_index:
            edge_index = edge_index.to(torch.long)
        if self.return_vecs:
            return edge_index, edge_weight, edge_vec
        return edge_index, edge_weight, None


class Distance(nn.Module):
    """The reference's legacy distance module (models/utils.py:393-455; ``torch_cluster.radius_graph`` there, the module its
    graph network, transformer and pairwise priors are built on): every ordered pair (i, j), i != j, of a molecule with
    ``cutoff_lower <= d < cutoff_upper``, self pairs when ``loop`` (kept whatever the lower cutoff), and the reference's
    guard -- an atom with more than ``max_num_neighbors`` neighbours raises the same ``AssertionError``.  Here over the CUDA
    op of ``OptimizedDistance`` (open boundaries); differentiable in ``pos`` like it."""

    def __init__(self, cutoff_lower, cutoff_upper, max_num_neighbors=32, return_vecs=False, loop=False):
        super().__init__()
        self.cutoff_lower = cutoff_lower
        self.cutoff_upper = cutoff_upper
        self.max_num_neighbors = max_num_neighbors
        self.return_vecs = return_vecs
        self.loop = loop

    def forward(self, pos: Tensor, batch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Optional[Tensor]]:
        _native.load()
        _native.require_cuda(pos, "Distance")
        n = pos.shape[0]
        if batch is None:
            batch = torch.zeros(n, dtype=torch.long, device=pos.device)
        # room for one neighbour more than allowed per atom (+ the self pair): if the pairs found do not fit, some atom
        # is over its budget; if they fit, the list is complete and the budget is checked per atom
        cap = n * (int(self.max_num_neighbors) + 2)
        ei, ev, ew, num = torch.ops.torchmdnet_neighbors.get_neighbor_pairs(
            "brute", pos, batch, torch.empty((0, 0), dtype=pos.dtype), False, float(self.cutoff_lower),
            float(self.cutoff_upper), cap, bool(self.loop), True)
        msg = ("The neighbor search missed some atoms due to max_num_neighbors being too low. Please increase this "
               "parameter to include the maximum number of atoms within the cutoff.")
        assert int(num) <= cap, msg
        keep = ei[0] != -1
        ei = ei[:, keep].to(torch.long)
        ew = ew[keep]
        ev = ev[keep, :]
        counts = torch.bincount(ei[0][ei[0] != ei[1]], minlength=n)
        assert not bool((counts > self.max_num_neighbors).any()), msg
        return ei, ew, (ev if self.return_vecs else None)


class CosineCutoff(nn.Module):
    """models/utils.py:362-390."""

    def __init__(self, cutoff_lower=0.0, cutoff_upper=5.0):
        super().__init__()
        self.cutoff_lower = cutoff_lower
        self.cutoff_upper = cutoff_upper

    def forward(self, distances: Tensor) -> Tensor:
        lo, hi = self.cutoff_lower, self.cutoff_upper
        if lo > 0:
            c = 0.5 * (torch.cos(math.pi * (2 * (distances - lo) / (hi - lo) + 1.0)) + 1.0)
            return c * (distances < hi) * (distances > lo)
        return 0.5 * (torch.cos(distances * math.pi / hi) + 1.0) * (distances < hi)


class _Smearing(nn.Module):
    _names = ()

    def __init__(self, cutoff_lower, cutoff_upper, num_rbf, trainable, dtype):
        super().__init__()
        self.cutoff_lower = cutoff_lower
        self.cutoff_upper = cutoff_upper
        self.num_rbf = num_rbf
        self.trainable = trainable
        self.dtype = dtype
        for name, value in zip(self._names, self._initial_params()):
            if trainable:
                self.register_parameter(name, nn.Parameter(value))
            else:
                self.register_buffer(name, value)

    def reset_parameters(self):
        for name, value in zip(self._names, self._initial_params()):
            getattr(self, name).data.copy_(value)


class GaussianSmearing(_Smearing):
    """exp(coeff (d - offset_k)^2), models/utils.py:272-300.  Buffers: coeff, offset."""
    _names = ("coeff", "offset")

    def __init__(self, cutoff_lower=0.0, cutoff_upper=5.0, num_rbf=50, trainable=True, dtype=torch.float32):
        super().__init__(cutoff_lower, cutoff_upper, num_rbf, trainable, dtype)

    def _initial_params(self):
        offset = torch.linspace(self.cutoff_lower, self.cutoff_upper, self.num_rbf, dtype=self.dtype)
        return -0.5 / (offset[1] - offset[0]) ** 2, offset

    def forward(self, dist):
        return torch.exp(self.coeff * (dist.unsqueeze(-1) - self.offset) ** 2)


class ExpNormalSmearing(_Smearing):
    """C_0(d) exp(-beta_k (exp(alpha (lo - d)) - mu_k)^2), models/utils.py:303-344.  Buffers: means, betas."""
    _names = ("means", "betas")

    def __init__(self, cutoff_lower=0.0, cutoff_upper=5.0, num_rbf=50, trainable=True, dtype=torch.float32):
        super().__init__(cutoff_lower, cutoff_upper, num_rbf, trainable, dtype)
        self.alpha = 5.0 / (cutoff_upper - cutoff_lower)
        self.cutoff_fn = CosineCutoff(0, cutoff_upper)

    def _initial_params(self):
        start = torch.exp(torch.scalar_tensor(-self.cutoff_upper + self.cutoff_lower, dtype=self.dtype))
        means = torch.linspace(start, 1, self.num_rbf, dtype=self.dtype)
        betas = torch.tensor([(2 / self.num_rbf * (1 - start)) ** -2] * self.num_rbf, dtype=self.dtype)
        return means, betas

    def forward(self, dist):
        dist = dist.unsqueeze(-1)
        return self.cutoff_fn(dist) * torch.exp(
            -self.betas * (torch.exp(self.alpha * (-dist + self.cutoff_lower)) - self.means) ** 2)


class ShiftedSoftplus(nn.Module):
    def __init__(self):
        super().__init__()
        self.shift = math.log(2.0)

    def forward(self, x):
        return F.softplus(x) - self.shift


def _xavier_zero_bias(lin):
    nn.init.xavier_uniform_(lin.weight)
    if lin.bias is not None:
        lin.bias.data.fill_(0)


def _fused_only(self, *a, **k):
    raise RuntimeError(type(self).__name__ + " holds parameters only; it is evaluated inside the fused CUDA "
                       "energy+force kernels through TorchMD_Net.forward (no standalone / CPU path)")


class NeighborEmbedding(nn.Module):
    """Parameters of the ET neighbour embedding (models/utils.py:43-108): embedding, distance_proj, combine."""

    def __init__(self, hidden_channels, num_rbf, cutoff_lower, cutoff_upper, max_z=100, dtype=torch.float32):
        super().__init__()
        self.embedding = nn.Embedding(max_z, hidden_channels, dtype=dtype)
        self.distance_proj = nn.Linear(num_rbf, hidden_channels, dtype=dtype)
        self.combine = nn.Linear(hidden_channels * 2, hidden_channels, dtype=dtype)
        self.cutoff = CosineCutoff(cutoff_lower, cutoff_upper)
        self.reset_parameters()

    def reset_parameters(self):
        self.embedding.reset_parameters()
        nn.init.xavier_uniform_(self.distance_proj.weight)
        nn.init.xavier_uniform_(self.combine.weight)
        self.distance_proj.bias.data.fill_(0)
        self.combine.bias.data.fill_(0)

    def jittable(self):
        return self

    forward = _fused_only


class GatedEquivariantBlock(nn.Module):
    """Parameters of the gated equivariant block (models/utils.py:456-522)."""

    def __init__(self, hidden_channels, out_channels, intermediate_channels=None, activation="silu",
                 scalar_activation=False, dtype=torch.float):
        super().__init__()
        self.out_channels = out_channels
        self.scalar_activation = scalar_activation
        if intermediate_channels is None:
            intermediate_channels = hidden_channels
        self.vec1_proj = nn.Linear(hidden_channels, hidden_channels, bias=False, dtype=dtype)
        self.vec2_proj = nn.Linear(hidden_channels, out_channels, bias=False, dtype=dtype)
        act_class = act_class_mapping[activation]
        self.update_net = nn.Sequential(nn.Linear(hidden_channels * 2, intermediate_channels, dtype=dtype),
                                        act_class(),
                                        nn.Linear(intermediate_channels, out_channels * 2, dtype=dtype))
        self.act = act_class() if scalar_activation else None

    def reset_parameters(self):
        nn.init.xavier_uniform_(self.vec1_proj.weight)
        nn.init.xavier_uniform_(self.vec2_proj.weight)
        _xavier_zero_bias(self.update_net[0])
        _xavier_zero_bias(self.update_net[2])

    forward = _fused_only


def check_stream_capturing():
    """True while the current CUDA stream is being captured (torch_extension::is_stream_capturing,
    models/utils.py:533-575; here a compiled op of the shim instead of an import-time JIT build)."""
    _native.load()
    return torch.ops.torch_extension.is_stream_capturing()


rbf_class_mapping = {"gauss": GaussianSmearing, "expnorm": ExpNormalSmearing}
act_class_mapping = {"ssp": ShiftedSoftplus, "silu": nn.SiLU, "tanh": nn.Tanh, "sigmoid": nn.Sigmoid}
dtype_mapping = {16: torch.float16, 32: torch.float, 64: torch.float64}

# ids shared with include/tmdn.h
ACT_ID = {"silu": 0, "tanh": 1, "sigmoid": 2, "ssp": 3}
RBF_ID = {"expnorm": 0, "gauss": 1}
STRATEGY_NAMES = ("brute", "shared", "cell")
