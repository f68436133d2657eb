"""Shared host plumbing of the fused representation models (TensorNet, TorchMD_ET): weight-version tracking,
native model handles, workspace reuse and the single native call per evaluation."""
from typing import Optional, Tuple

import torch
from torch import Tensor, nn

from .. import _native


class FusedRepresentation(nn.Module):
    _op = ""            # op-name prefix inside torch.ops.torchmdnet_b200: "tensornet" | "et"
    _kernel_name = ""

    def _init_fused(self):
        self.table_intervals = None     # None: max(1024, 32 * num_rbf) intervals (TableGrid in tensornet.py)
        self.table_warp = True
        self._handle = None
        self._pack_key = None
        self._workspace = None

    def resolved_table_intervals(self) -> int:
        if self.table_intervals:
            return int(self.table_intervals)
        return max(1024, 32 * int(self.num_rbf)) if self.table_warp else 4096

    # subclasses: pack_weights(head, mean, std) -> (weights, iparams, fparams); _extra_key() -> hashable
    def _extra_key(self):
        return ()

    def _tensors_for_key(self, head):
        mods = [self] if head is None else [self, head]
        for m in mods:
            yield from m.parameters()
            yield from m.buffers()

    def _ops(self, what):
        return getattr(torch.ops.torchmdnet_b200, self._op + "_" + what)

    def native_handle(self, head: Optional[nn.Module] = None, mean: float = 0.0, std: float = 1.0) -> int:
        """Handle of the packed model inside the shim; repacked when any weight changed."""
        _native.load()
        key = (tuple((t.data_ptr(), t._version, t.device) for t in self._tensors_for_key(head)),
               self._extra_key(), self.table_intervals, float(mean), float(std))
        if key != self._pack_key:
            p = next(self.parameters())
            _native.require_cuda(p, self._kernel_name)
            if p.dtype != torch.float32:
                raise NotImplementedError("%s kernels are fp32 (precision: 32); got %s" % (self._kernel_name, p.dtype))
            if self._handle is not None:
                self._ops("destroy")(self._handle)
            self._workspace = None          # may live on the previous device
            w, ip, fp = self.pack_weights(head, mean, std)
            self._handle = self._ops("create")(w, ip, fp)
            self._pack_key = key
        return self._handle

    def __getstate__(self):
        """copy.deepcopy / pickle: the native handle, its key and the workspace belong to THIS instance (a copy that
        kept the integer handle would destroy the original's packed model on its first call)."""
        state = dict(self.__dict__)
        state["_handle"] = state["_pack_key"] = state["_workspace"] = None
        return state

    def invalidate_cache(self):
        """Force re-packing (needed only after in-place edits through ``.data`` that bypass version counters)."""
        self._pack_key = None

    def __del__(self):
        try:
            if self._handle is not None:
                self._ops("destroy")(self._handle)
        except Exception:
            pass

    def _max_edges(self, n_atoms):
        dist = self.distance
        return dist.max_num_pairs if dist.max_num_pairs > 0 else -dist.max_num_pairs * n_atoms

    def energy_forces(self, z: Tensor, pos: Tensor, batch: Tensor, n_mol: int, head: Optional[nn.Module],
                      mean: float, std: float, compute_forces: bool = True) -> Tuple[Tensor, Tensor, Tensor]:
        """One fused evaluation: -> (y[n_mol,1], forces[N,3] (empty if not compute_forces), status int32[4]).
        status[0] = edges found (incl. self loops), status[1] = 1 if they exceeded the capacity
        max_num_neighbors*N (models/utils.py:234-236), status[2] = 1 if some z was outside [0,max_z).
        When status[1] or status[2] is set the kernels write NaN into y and forces (a truncated edge list is not
        symmetric, so the analytic forces would not belong to the returned energy).
        Neighbour search parameters (cutoffs, box, strategy, capacity, check_errors) are read from
        ``self.distance`` on every call, so a periodic ``OptimizedDistance`` can be swapped in by attribute.
        ``batch`` must be sorted ascending (molecules contiguous), as the reference's own collate produces it.
        One instance owns one workspace: do not evaluate the same instance concurrently on two streams."""
        handle = self.native_handle(head, mean, std)
        _native.require_cuda(pos, self._kernel_name)
        if pos.dtype != torch.float32:
            raise NotImplementedError("%s kernels are fp32; got positions of dtype %s" % (self._kernel_name, pos.dtype))
        dist = self.distance
        max_edges = self._max_edges(pos.shape[0])
        y, f, status, ws = self._ops("energy_forces")(
            handle, z.contiguous(), pos.detach().contiguous(), batch.contiguous(), dist.box, dist.use_periodic,
            dist.strategy, max_edges, n_mol, compute_forces, self._workspace)
        self._workspace = ws
        if dist.check_errors and not torch.ops.torch_extension.is_stream_capturing() and int(status[1]) != 0:
            raise RuntimeError("Found num_pairs({}) > max_num_pairs({})".format(int(status[0]), max_edges))
        return y, f, status

    def debug_buffer(self, name: str, n_atoms: int, n_mol: int, dtype=torch.float32) -> Tensor:
        """Flat view of a named intermediate of the LAST call (names as in oracle/staged_*.py)."""
        dist = self.distance
        off, nbytes = self._ops("debug_lookup")(self._handle, n_atoms, self._max_edges(n_atoms), n_mol, dist.strategy,
                                                dist.box, name)
        return self._workspace[off:off + nbytes].view(dtype)
