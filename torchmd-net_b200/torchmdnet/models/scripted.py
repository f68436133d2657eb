"""TorchScript deployment of the fused models (the reference pins it with tests/test_model.py:40-58,
85-125: script -> save -> load -> CUDA graph).

``torch.jit.script(model)`` on a ``TorchMD_Net`` of this package returns the scripted form of ``ScriptableTorchMD_Net``
(through ``TorchMD_Net.__prepare_scriptable__``): a module that holds ONE flat fp32 buffer with every packed weight
(filter tables included) and calls the fused op ``torchmdnet_b200::{tensornet,et,gn}_energy_forces_packed``, which finds or
creates the native model from the identity of that buffer.  The buffer follows ``.to(device)``, is written by
``torch.jit.save`` and restored by ``torch.jit.load`` in another process; nothing process-local is serialised.

The weights are a snapshot taken at scripting time (deployment is inference).  As in the reference, the number of
molecules is ``batch.max() + 1`` (one sync) unless the stream is being captured, in which case the value of the last
eager call is used (output_modules.py:27-43)."""
from typing import Dict, List, Optional, Tuple

import torch
from torch import Tensor, nn


def _flatten(weights: List[Tensor]):
    """-> (flat fp32 tensor, layout = [offset_0..offset_{n-1} | len_0..len_{n-1}]), segments aligned to 256 bytes"""
    offs, lens, cur = [], [], 0
    for w in weights:
        offs.append(cur)
        lens.append(int(w.numel()))
        cur += (int(w.numel()) + 63) // 64 * 64
    dev = weights[0].device
    flat = torch.zeros(max(cur, 64), dtype=torch.float32, device=dev)
    for w, o in zip(weights, offs):
        if w.numel():
            flat[o:o + w.numel()] = w.reshape(-1).to(torch.float32)
    return flat, offs + lens


class ScriptableTorchMD_Net(nn.Module):
    """``forward(z, pos, batch=None, q=None, s=None, extra_args=None) -> (y, neg_dy | None)`` with the signature of
    ``TorchMD_Net.forward`` (models/model.py:232-300), scriptable, saveable, CUDA-graph capturable."""

    layout: List[int]
    iparams: List[int]
    fparams: List[float]
    box: List[float]
    prior_kind: List[int]          # per prior: 0 Atomref, 1 ZBL, 2 Coulomb, 3 D2
    prior_off: List[int]           # offset of the prior's per-element table in ``prior_data``
    prior_len: List[int]           # entries per table (D2: two tables of this length, back to back)
    prior_par: List[float]         # per prior: cutoff, energy prefactor, distance scale, alpha, max_num_neighbors

    def __init__(self, model):
        super().__init__()
        rm = model.representation_model
        mean, std = model._mean_std()
        w, ip, fp = rm.pack_weights(model.output_model, mean, std)
        flat, layout = _flatten(w)
        self.register_buffer("packed", flat)
        self.layout = [int(x) for x in layout]
        self.iparams = [int(x) for x in ip]
        self.fparams = [float(x) for x in fp]
        self.kind: int = {"tensornet": 0, "et": 1, "gn": 2, "t": 3}[rm._op]
        self.derivative: bool = bool(model.derivative)
        dist = rm.distance
        self.max_num_pairs: int = int(dist.max_num_pairs)
        self.use_periodic: bool = bool(dist.use_periodic)
        self.strategy: str = str(dist.strategy)
        box = dist.box
        self.box = [] if box is None or box.numel() == 0 else [float(x) for x in box.double().reshape(-1)]
        self.dim_size: int = 0
        # priors (models/model.py:266-284): per-element tables in one flat buffer, scalars in plain lists
        self.prior_kind, self.prior_off, self.prior_len, self.prior_par = [], [], [], []
        tables = []
        cur = 0
        for prior in (model.prior_model if model.prior_model is not None else []):
            name = type(prior).__name__
            if name == "Atomref":
                kind, tabs, par = 0, [prior.atomref.weight.detach().reshape(-1)], [0.0, 0.0, 0.0, 0.0, 0.0]
            elif name == "ZBL":
                kind, tabs = 1, [prior.atomic_number.detach()]
                par = list(prior._params()) + [float(prior.max_num_neighbors)]
            elif name == "Coulomb":
                kind, tabs = 2, [torch.zeros(1)]
                par = list(prior._params()) + [float(prior.max_num_neighbors)]
            elif name == "D2":
                zz = prior.Z_map
                kind, tabs = 3, [prior.C_6[zz].detach().to(torch.float32).sqrt(), prior.R_r[zz].detach().to(torch.float32)]
                par = list(prior._params()) + [float(prior.max_num_neighbors)]
            else:
                raise NotImplementedError("scripting a model with a %s prior is not built" % name)
            self.prior_kind.append(kind)
            self.prior_off.append(cur)
            self.prior_len.append(int(tabs[0].numel()))
            self.prior_par += [float(x) for x in par]
            for t in tabs:
                tables.append(t.to(device=flat.device, dtype=torch.float32).reshape(-1))
                cur += int(t.numel())
        self.register_buffer("prior_data", torch.cat(tables) if tables else torch.zeros(1, device=flat.device))

    def forward(self, z: Tensor, pos: Tensor, batch: Optional[Tensor] = None, q: Optional[Tensor] = None,
                s: Optional[Tensor] = None, extra_args: Optional[Dict[str, Tensor]] = None
                ) -> Tuple[Tensor, Optional[Tensor]]:
        assert z.dim() == 1 and z.dtype == torch.long
        if batch is None:
            b = torch.zeros_like(z)
            n_mol = 1
        else:
            b = batch
            if torch.ops.torch_extension.is_stream_capturing():
                assert self.dim_size > 0, "Warming up is needed before capturing the model into a CUDA graph"
                n_mol = self.dim_size
            else:
                n_mol = int(batch.max().item()) + 1
                self.dim_size = n_mol
        n = pos.size(0)
        max_edges = self.max_num_pairs if self.max_num_pairs > 0 else -self.max_num_pairs * n
        if self.kind == 0:
            y, f, status = torch.ops.torchmdnet_b200.tensornet_energy_forces_packed(
                self.packed, self.layout, self.iparams, self.fparams, z.contiguous(), pos.detach().contiguous(),
                b.contiguous(), self.box, self.use_periodic, self.strategy, max_edges, n_mol, self.derivative)
        elif self.kind == 1:
            y, f, status = torch.ops.torchmdnet_b200.et_energy_forces_packed(
                self.packed, self.layout, self.iparams, self.fparams, z.contiguous(), pos.detach().contiguous(),
                b.contiguous(), self.box, self.use_periodic, self.strategy, max_edges, n_mol, self.derivative)
        elif self.kind == 3:
            y, f, status = torch.ops.torchmdnet_b200.t_energy_forces_packed(
                self.packed, self.layout, self.iparams, self.fparams, z.contiguous(), pos.detach().contiguous(),
                b.contiguous(), self.box, self.use_periodic, self.strategy, max_edges, n_mol, self.derivative)
        else:
            y, f, status = torch.ops.torchmdnet_b200.gn_energy_forces_packed(
                self.packed, self.layout, self.iparams, self.fparams, z.contiguous(), pos.detach().contiguous(),
                b.contiguous(), self.box, self.use_periodic, self.strategy, max_edges, n_mol, self.derivative)
        for ip in range(len(self.prior_kind)):
            kind, off, ln = self.prior_kind[ip], self.prior_off[ip], self.prior_len[ip]
            if kind == 0:      # Atomref.pre_reduce + the molecular sum: forces stand as they are
                w = self.prior_data[off:off + ln]
                add = torch.zeros(n_mol, 1, dtype=torch.float32, device=y.device)
                torch.ops.torchmdnet_b200.segment_add(w[z].contiguous(), b.contiguous(), add.view(-1))   # fixed summation order
                y = y + add.to(y.dtype)
            else:              # pair priors: energy and analytic forces in place (csrc/pair_prior.cu)
                cutoff, pref, dscale, alpha = (self.prior_par[5 * ip], self.prior_par[5 * ip + 1], self.prior_par[5 * ip + 2],
                                               self.prior_par[5 * ip + 3])
                cap = (int(self.prior_par[5 * ip + 4]) + 1) * n
                if kind == 1:
                    name, w = "zbl", self.prior_data[off:off + ln][z]
                elif kind == 2:
                    assert extra_args is not None and "partial_charges" in extra_args, \
                        "the Coulomb prior needs extra_args['partial_charges']"
                    name, w = "coulomb", extra_args["partial_charges"].to(torch.float32)
                else:
                    name = "d2"
                    w = torch.stack([self.prior_data[off:off + ln][z], self.prior_data[off + ln:off + 2 * ln][z]])
                y = y.contiguous()
                found = torch.ops.torchmdnet_b200.pair_prior(name, pos.detach().to(torch.float32).contiguous(), b.contiguous(),
                                                             w.contiguous(), cutoff, pref, dscale, alpha, cap, y,
                                                             f if self.derivative else None)
                if not torch.ops.torch_extension.is_stream_capturing():
                    assert int(found.item()) <= cap, ("The neighbor search missed some atoms due to max_num_neighbors being "
                                                      "too low. Please increase this parameter to include the maximum "
                                                      "number of atoms within the cutoff.")
        if self.derivative:
            return y, f
        return y, None
