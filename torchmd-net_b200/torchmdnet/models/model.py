"""``create_model`` / ``load_model`` / ``TorchMD_Net`` with the reference's signatures
(torchmdnet/models/model.py:15-300) over the fused CUDA energy+force pipeline.

``TorchMD_Net.forward(z, pos, batch=None, q=None, s=None, extra_args=None) -> (y, neg_dy | None)`` performs ONE
native call: neighbour search, representation, head, per-molecule reduction, ``* std + mean`` and -- when
``derivative`` -- the hand-written backward kernels that give ``neg_dy = -dy/dpos`` (the reference replays the
graph with ``torch.autograd.grad(create_graph=True)``, model.py:286-298).  In evaluation mode the returned tensors carry
no autograd graph.  In training mode (``model.train()``, grad enabled, parameters requiring grad) the forward is the
differentiable composition of ``models/eager.py`` on the GPU -- edge list from this package's neighbour op, forces by
``autograd.grad(create_graph=True)`` -- so energy and force-matching losses train the same parameters.
"""
import re
import warnings
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor, nn

from . import output_modules
from .utils import dtype_mapping


def create_model(args, prior_model=None, mean=None, std=None):
    """Build a model from the flat hparam dict the reference reads (model.py:28-116)."""
    dtype = dtype_mapping[args["precision"]]
    shared_args = dict(hidden_channels=args["embedding_dimension"], num_layers=args["num_layers"],
                       num_rbf=args["num_rbf"], rbf_type=args["rbf_type"], trainable_rbf=args["trainable_rbf"],
                       activation=args["activation"], cutoff_lower=args["cutoff_lower"],
                       cutoff_upper=args["cutoff_upper"], max_z=args["max_z"],
                       max_num_neighbors=args["max_num_neighbors"], dtype=dtype)
    if args["model"] == "equivariant-transformer":
        from .torchmd_et import TorchMD_ET
        is_equivariant = True
        representation_model = TorchMD_ET(attn_activation=args["attn_activation"], num_heads=args["num_heads"],
                                          distance_influence=args["distance_influence"],
                                          neighbor_embedding=args["neighbor_embedding"], **shared_args)
    elif args["model"] == "tensornet":
        from .tensornet import TensorNet
        is_equivariant = False          # TensorNet always takes the plain Scalar head (model.py:79-80)
        representation_model = TensorNet(equivariance_invariance_group=args["equivariance_invariance_group"],
                                         **shared_args)
    elif args["model"] == "graph-network":
        from .torchmd_gn import TorchMD_GN
        is_equivariant = False
        representation_model = TorchMD_GN(num_filters=args["embedding_dimension"], aggr=args["aggr"],
                                          neighbor_embedding=args["neighbor_embedding"], **shared_args)
    elif args["model"] == "transformer":
        from .torchmd_t import TorchMD_T
        is_equivariant = False
        representation_model = TorchMD_T(attn_activation=args["attn_activation"], num_heads=args["num_heads"],
                                         distance_influence=args["distance_influence"],
                                         neighbor_embedding=args["neighbor_embedding"], **shared_args)
    else:
        raise ValueError(f'Unknown architecture: {args["model"]}')
    if args.get("atom_filter", -1) > -1:
        if args["derivative"]:
            raise ValueError("Derivative and atom filter can't be used together")
        raise NotImplementedError("AtomFilter is outside the accelerated energy+force path of this build")
    if prior_model is None:
        prior_model = create_prior_models(args)      # when loading a model (model.py:96-98)
    if args["output_model"] != "Scalar":
        raise NotImplementedError('output_model "%s": only the Scalar energy head is built' % args["output_model"])
    if args.get("reduce_op", "add") not in ("add", "sum"):
        raise NotImplementedError('reduce_op "%s": only add/sum is built' % args["reduce_op"])
    output_prefix = "Equivariant" if is_equivariant else ""
    output_model = getattr(output_modules, output_prefix + args["output_model"])(
        args["embedding_dimension"], activation=args["activation"], reduce_op=args["reduce_op"], dtype=dtype)
    return TorchMD_Net(representation_model, output_model, prior_model=prior_model, mean=mean, std=std,
                       derivative=args["derivative"], dtype=dtype)


def create_prior_models(args, dataset=None):
    """Parse the ``prior_model`` option (str | dict | list of them, with ``prior_args``) as models/model.py:146-177.
    Built here: ``Atomref``, ``ZBL``, ``Coulomb``, ``D2``."""
    from .. import priors
    prior_models = []
    if args.get("prior_model"):
        prior_model = args["prior_model"]
        prior_names, prior_args = [], []
        if not isinstance(prior_model, list):
            prior_model = [prior_model]
        for prior in prior_model:
            if isinstance(prior, dict):
                for key, value in prior.items():
                    prior_names.append(key)
                    prior_args.append({} if value is None else value)
            else:
                prior_names.append(prior)
                prior_args.append({})
        if "prior_args" in args and args["prior_args"] is not None:
            prior_args = args["prior_args"]
            if not isinstance(prior_args, list):
                prior_args = [prior_args]
        for name, arg in zip(prior_names, prior_args):
            if not hasattr(priors, name):
                raise NotImplementedError('prior model "%s" is outside this build (available: %s)'
                                          % (name, ", ".join(priors.__all__)))
            prior_models.append(getattr(priors, name)(dataset=dataset, **arg))
    return prior_models or None


def load_model(filepath, args=None, device="cpu", **kwargs):
    """Rebuild a model from a Lightning checkpoint (``hyper_parameters`` + ``state_dict`` with a ``model.``
    prefix), model.py:121-143."""
    ckpt = torch.load(filepath, map_location="cpu", weights_only=False)
    if args is None:
        args = ckpt["hyper_parameters"]
    for key, value in kwargs.items():
        if key not in args:
            warnings.warn(f"Unknown hyperparameter: {key}={value}")
        args[key] = value
    model = create_model(args)
    state_dict = {re.sub(r"^model\.", "", k): v for k, v in ckpt["state_dict"].items()}
    # backward compatibility with checkpoints written when Atomref was the only prior (model.py:134-141)
    for old, new in (("prior_model.initial_atomref", "prior_model.0.initial_atomref"),
                     ("prior_model.atomref.weight", "prior_model.0.atomref.weight")):
        if old in state_dict:
            state_dict[new] = state_dict.pop(old)
    model.load_state_dict(state_dict)
    return model.to(device)


class TorchMD_Net(nn.Module):
    """Representation model + energy head + (fused) force evaluation (model.py:180-300)."""

    def __init__(self, representation_model, output_model, prior_model=None, mean=None, std=None,
                 derivative=False, dtype=torch.float32):
        super().__init__()
        self.representation_model = representation_model.to(dtype=dtype)
        self.output_model = output_model.to(dtype=dtype)
        from ..priors import D2, Atomref, BasePrior, Coulomb, ZBL
        if isinstance(prior_model, BasePrior):
            prior_model = [prior_model]
        if prior_model is not None and not all(isinstance(p, (Atomref, ZBL, Coulomb, D2)) for p in prior_model):
            raise NotImplementedError("priors built here: Atomref, ZBL, Coulomb, D2")
        self.prior_model = None if not prior_model else nn.ModuleList(prior_model).to(dtype=dtype)
        self.derivative = derivative
        mean = torch.scalar_tensor(0) if mean is None else mean
        self.register_buffer("mean", mean.to(dtype=dtype))
        std = torch.scalar_tensor(1) if std is None else std
        self.register_buffer("std", std.to(dtype=dtype))
        self._scale = None          # (mean, std) as host floats, cached per buffer version (one sync, once)
        self.reset_parameters()

    def reset_parameters(self):
        self.representation_model.reset_parameters()
        self.output_model.reset_parameters()
        if self.prior_model is not None:          # model.py:224-226
            for prior in self.prior_model:
                prior.reset_parameters()

    def _mean_std(self):
        key = (self.mean.data_ptr(), self.mean._version, self.std.data_ptr(), self.std._version)
        if self._scale is None or self._scale[0] != key:
            self._scale = (key, float(self.mean), float(self.std))
        return self._scale[1], self._scale[2]

    def forward(self, z: Tensor, pos: Tensor, batch: Optional[Tensor] = None, q: Optional[Tensor] = None,
                s: Optional[Tensor] = None, extra_args: Optional[Dict[str, Tensor]] = None
                ) -> Tuple[Tensor, Optional[Tensor]]:
        assert z.dim() == 1 and z.dtype == torch.long
        if batch is None:
            batch = torch.zeros_like(z)
            n_mol = 1
        else:
            n_mol = self._num_molecules(batch)
        if self.training and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            return self._forward_training(z, pos, batch, n_mol, extra_args)
        if self.mean.dtype == torch.float64:
            # precision: 64 (models/utils.py:586): the fused kernels are fp32; double-precision models run the differentiable
            # composition (models/eager.py) with the fp64 neighbour op, which is what the reference's fp64 gradient check
            # (tests/test_model.py:191-211) exercises.  Outside training the result is detached, like the fused path's.
            with torch.enable_grad():
                y, neg_dy = self._forward_training(z, pos.detach() if not torch.is_grad_enabled() else pos, batch, n_mol,
                                                   extra_args, quiet=True)
            if not torch.is_grad_enabled():
                y, neg_dy = y.detach(), (None if neg_dy is None else neg_dy.detach())
            return y, neg_dy
        mean, std = self._mean_std()
        y, neg_dy, _ = self.representation_model.energy_forces(z, pos, batch, n_mol, self.output_model, mean, std,
                                                                compute_forces=self.derivative)
        y = self.apply_priors(y, z, batch, n_mol, pos, neg_dy if self.derivative else None, extra_args)
        return y, (neg_dy if self.derivative else None)

    def _forward_training(self, z, pos, batch, n_mol, extra_args, quiet=False):
        """``model.train()`` with grad enabled: the differentiable composition of models/eager.py (training and
        double backward are outside the fused kernels).  Forces come from ``autograd.grad(create_graph=True)`` exactly as in
        the reference (model.py:286-298), so a force-matching loss back-propagates into the parameters.  Pairwise priors
        have no parameters: their energy and analytic forces are added as constants."""
        from . import eager
        if not quiet and not getattr(self, "_warned_training_path", False):
            self._warned_training_path = True
            warnings.warn("TorchMD_Net is in training mode: forward runs the differentiable PyTorch composition "
                          "(models/eager.py) so that losses reach the parameters; call model.eval() for the fused "
                          "CUDA energy+force kernels")
        if self.derivative:
            pos = pos.requires_grad_(True) if pos.is_leaf else pos
            if not pos.requires_grad:
                pos = pos.detach().requires_grad_(True)
        y = eager.energy(self, z, pos, batch, n_mol)
        neg_dy = None
        if self.derivative:
            (dy,) = torch.autograd.grad([y], [pos], grad_outputs=[torch.ones_like(y)], create_graph=True, retain_graph=True)
            neg_dy = -dy
        if self.prior_model is not None:
            from ..priors import Atomref
            for prior in self.prior_model:
                if not isinstance(prior, Atomref):
                    e = torch.zeros_like(y).detach()
                    f = torch.zeros_like(pos).detach() if self.derivative else None
                    prior.add_energy_forces(e, f, z, pos.detach(), batch, extra_args)
                    y = y + e
                    if f is not None:
                        neg_dy = neg_dy + f
        return y, neg_dy

    def apply_priors(self, y: Tensor, z: Tensor, batch: Tensor, n_mol: int, pos: Optional[Tensor] = None,
                     forces: Optional[Tensor] = None, extra_args: Optional[Dict[str, Tensor]] = None) -> Tensor:
        """Priors on top of the fused call's output.  Atom-wise reference energies (Atomref.pre_reduce + reduce,
        model.py:266-272) are position independent: forces stand as they are.  Pairwise priors (post_reduce,
        model.py:281-284) add their energy to ``y`` and their analytic forces to ``forces`` in place."""
        if self.prior_model is not None:
            from ..priors import Atomref
            for prior in self.prior_model:
                if isinstance(prior, Atomref):
                    y = y + prior.molecular_sum(z, batch, n_mol).to(y.dtype)
                else:
                    y = prior.add_energy_forces(y.contiguous(), forces, z, pos, batch, extra_args)
        return y

    def __prepare_scriptable__(self):
        """``torch.jit.script(model)`` compiles this stand-in (models/scripted.py): one flat weight buffer + the fused op,
        priors included (their tables are packed too).  The weights are packed as they are now; script again after
        changing them."""
        from .. import _native
        from .scripted import ScriptableTorchMD_Net
        _native.load()
        return ScriptableTorchMD_Net(self)

    def _num_molecules(self, batch):
        """``batch.max()+1`` (output_modules.py:33); cached while capturing a CUDA graph, as the reference."""
        from .utils import check_stream_capturing
        om = self.output_model
        if batch.is_cuda and check_stream_capturing():
            assert om.dim_size > 0, "Warming up is needed before capturing the model into a CUDA graph"
            return om.dim_size
        om.dim_size = int(batch.max().item() + 1)
        return om.dim_size
