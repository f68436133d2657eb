"""Spatial domain decomposition of one large periodic system across GPUs (BASELINE.json configs[3]; new capability,
the reference has no inference-time multi-GPU path: SURVEY section 5/8e).

One process per GPU.  The box is cut into slabs along one axis (all peers are one NVSwitch hop apart, so the shape
is chosen for halo volume and message count: two neighbours per rank).  A rank owns the atoms whose *base* position
lies in its slab and keeps, as the contiguous tail of every local array, ghost copies of the neighbours' atoms within
``cutoff + 2*skin`` of the shared boundary; ownership and send lists stay fixed while no atom has moved more than
``skin`` from its base position (the usual MD neighbour-skin contract).

Per energy+force evaluation (stages of ``tmdn_tensornet_stage``, include/tmdn.h):
    positions halo -> EMBED -> Y0 halo -> LAYER_FWD(0) -> Y1 halo -> ... -> HEAD -> dM halo -> LAYER_BWD(L-1) -> ...
    -> LAYER_BWD(0) -> Dias halo -> FINAL;   energy = all-reduce of the per-rank sums over owned atoms.
Every halo is an owner -> ghost copy of node rows ([.,9,H] fp32, 4608 B/atom at H=128): because the edge set is
symmetric the backward kernels gather adjoints instead of scattering them, so no reverse (accumulating) halo exists
and the forces of owned atoms are complete locally.  Packing is a gather kernel of libtmdn (``tmdn_halo_pack_f32``);
receives land directly in the ghost rows; transport is grouped NCCL send/recv (torch.distributed) over NVLink.
"""
from dataclasses import dataclass
from typing import Callable, List, Optional

import torch

from . import _native

EMBED, LAYER_FWD, HEAD, LAYER_BWD, FINAL = 0, 1, 2, 3, 4


@dataclass
class SlabPlan:
    rank: int
    world: int
    left: int
    right: int
    own: torch.Tensor          # int64 global indices of owned atoms (ascending)
    send_left: torch.Tensor    # int32 local indices (into own) of atoms the left neighbour needs
    send_right: torch.Tensor
    n_ghost_left: int = 0      # rows received from the left neighbour (= its send_right)
    n_ghost_right: int = 0

    @property
    def n_own(self):
        return int(self.own.numel())

    @property
    def n_local(self):
        return self.n_own + self.n_ghost_left + self.n_ghost_right


def slab_plans(pos: torch.Tensor, box: torch.Tensor, cutoff: float, world: int, skin: float = 0.5,
               axis: int = 0) -> List[SlabPlan]:
    """Plans of ALL ranks from the global base positions (deterministic: every rank computes the same list, so no
    size exchange is needed).  ``pos`` [N,3], ``box`` diagonal [3,3]."""
    L = float(box[axis, axis])
    x = pos[:, axis].double().cpu()
    x = x - torch.floor(x / L + 0.5) * L                      # [-L/2, L/2)
    w = L / world
    halo = cutoff + 2.0 * skin
    if world > 1 and (w < halo or (world == 2 and w < 2 * halo)):
        raise ValueError("slab width %.2f too small for halo %.2f with %d ranks" % (w, halo, world))
    owner = torch.clamp(torch.floor((x + L / 2) / w).long(), 0, world - 1)
    plans = []
    for r in range(world):
        own = torch.nonzero(owner == r).flatten()
        xo = x[own]
        lo = -L / 2 + r * w
        hi = lo + w
        if world == 1:
            sl = sr = torch.zeros(0, dtype=torch.int32)
        else:
            sl = torch.nonzero(xo < lo + halo).flatten().to(torch.int32)
            sr = torch.nonzero(xo >= hi - halo).flatten().to(torch.int32)
        plans.append(SlabPlan(r, world, (r - 1) % world, (r + 1) % world, own, sl, sr))
    for p in plans:
        if world > 1:
            p.n_ghost_left = int(plans[p.left].send_right.numel())
            p.n_ghost_right = int(plans[p.right].send_left.numel())
    return plans


def ghost_global_indices(plans: List[SlabPlan], rank: int) -> torch.Tensor:
    """Global indices of the ghost rows of ``rank`` in local order [from left | from right]."""
    p = plans[rank]
    if p.world == 1:
        return torch.zeros(0, dtype=torch.long)
    gl = plans[p.left].own[plans[p.left].send_right.long()]
    gr = plans[p.right].own[plans[p.right].send_left.long()]
    return torch.cat([gl, gr])


class NcclHalo:
    """Owner -> ghost exchange of the rows of a local [n_local, F] fp32 tensor with the two slab neighbours:
    pack kernel + grouped NCCL send/recv; receives write straight into the ghost rows."""

    def __init__(self, plan: SlabPlan, device, group=None, pack=None):
        self.plan, self.group = plan, group
        self.send_left = plan.send_left.to(device)
        self.send_right = plan.send_right.to(device)
        # pack = the gather kernel of libtmdn; host-logic tests on the gloo backend pass an index_select instead
        self.pack = pack if pack is not None else (lambda buf, idx: torch.ops.torchmdnet_b200.halo_pack(buf, idx))

    def __call__(self, buf: torch.Tensor):
        import torch.distributed as dist
        p = self.plan
        if p.world == 1:
            return
        n0, gl, gr = p.n_own, p.n_ghost_left, p.n_ghost_right
        sl = self.pack(buf, self.send_left)
        sr = self.pack(buf, self.send_right)
        # a peer posts [send_left, send_right]; when both neighbours are the same rank (world == 2) messages between the
        # pair match in posting order, so receive first what the right neighbour sent leftwards, then the other one
        ops = [dist.P2POp(dist.isend, sl, p.left, self.group), dist.P2POp(dist.isend, sr, p.right, self.group),
               dist.P2POp(dist.irecv, buf[n0 + gl:n0 + gl + gr], p.right, self.group),
               dist.P2POp(dist.irecv, buf[n0:n0 + gl], p.left, self.group)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()


class DecomposedTensorNet:
    """Runs a TensorNet ``TorchMD_Net`` on one slab.  ``step_stages(pos_own)`` is a generator that yields every local
    [n_local, F] buffer whose ghost rows must be filled before it is resumed; ``step`` drives it with a halo callable
    (``NcclHalo`` in a real multi-GPU run, direct copies in the single-process emulation used by the tests)."""

    def __init__(self, model, plans: List[SlabPlan], rank: int, z_global: torch.Tensor, box: torch.Tensor, device,
                 halo: Optional[Callable] = None, strategy: str = "cell"):
        _native.load()
        self.model, self.plan, self.device = model, plans[rank], torch.device(device)
        rm = model.representation_model
        if rm._op != "tensornet":
            raise NotImplementedError("spatial decomposition is built for TensorNet")
        self.rm = rm
        p = self.plan
        loc = torch.cat([p.own, ghost_global_indices(plans, rank)])
        self.z = z_global.cpu()[loc].to(self.device).contiguous()
        self.batch = torch.zeros(p.n_local, dtype=torch.long, device=self.device)
        self.box = box.detach().cpu().to(torch.float32)
        self.strategy = strategy
        self.max_edges = int(rm.max_num_neighbors) * p.n_local
        self.halo = halo if halo is not None else NcclHalo(p, self.device)
        _, self.std = model._mean_std()
        # own packed copy of the weights with mean = 0 (the mean is added once, after the all-reduce)
        ops = torch.ops.torchmdnet_b200
        w, ip, fp = rm.pack_weights(model.output_model, 0.0, self.std)
        self.handle = ops.tensornet_create(w, ip, fp)
        nbytes = ops.tensornet_workspace_bytes(self.handle, p.n_local, self.max_edges, 1, strategy, self.box)
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.pos = torch.zeros(p.n_local, 3, device=self.device)
        self.energy = torch.zeros(1, 1, device=self.device)
        self.forces = torch.zeros(p.n_own, 3, device=self.device)
        self.status = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.H, self.L = rm.hidden_channels, rm.num_layers

    def __del__(self):
        try:
            torch.ops.torchmdnet_b200.tensornet_destroy(self.handle)
        except Exception:
            pass

    def _buffer(self, name):
        off, nbytes = torch.ops.torchmdnet_b200.tensornet_debug_lookup(self.handle, self.plan.n_local, self.max_edges, 1,
                                                                       self.strategy, self.box, name)
        return self.ws[off:off + nbytes].view(torch.float32).view(self.plan.n_local, 9 * self.H)

    def _stage(self, stage, layer=0):
        torch.ops.torchmdnet_b200.tensornet_stage(self.handle, stage, layer, self.z, self.pos, self.batch,
                                                  self.plan.n_own, self.box, True, self.strategy, self.max_edges, 1,
                                                  self.energy, self.forces, self.status, self.ws)

    def step_stages(self, pos_own: torch.Tensor):
        L = self.L
        self.pos[:self.plan.n_own].copy_(pos_own)
        yield self.pos
        self._stage(EMBED)
        for l in range(L):
            yield self._buffer("Y%d" % l)
            self._stage(LAYER_FWD, l)
        self._stage(HEAD)
        for l in range(L - 1, -1, -1):
            yield self._buffer("Dias")          # dM of layer l
            self._stage(LAYER_BWD, l)
        yield self._buffer("Dias")              # dE/d(ias) of the embedding
        self._stage(FINAL)

    def step(self, pos_own: torch.Tensor):
        """-> (sum of the owned atoms' energies [1,1] (no mean), forces of the owned atoms [n_own,3])."""
        for buf in self.step_stages(pos_own):
            self.halo(buf)
        return self.energy, self.forces

    def total_energy(self, group=None):
        """All-reduce of the per-rank sums + the model's mean (one scalar per evaluation)."""
        import torch.distributed as dist
        e = self.energy.clone()
        if self.plan.world > 1:
            dist.all_reduce(e, group=group)
        return e + self.model._mean_std()[0]


def emulate_step(runners: List[DecomposedTensorNet], pos_global: torch.Tensor):
    """All ranks of a decomposition in ONE process on one GPU (tests): the stage generators advance in lock step and
    every halo point is served by direct device copies.  -> (energy [1,1] incl. mean, forces [N,3] in global order)."""
    gens = [r.step_stages(pos_global[r.plan.own.to(pos_global.device)]) for r in runners]
    while True:
        bufs = []
        for g in gens:
            try:
                bufs.append(next(g))
            except StopIteration:
                bufs.append(None)
        if bufs[0] is None:
            break
        for r, buf in zip(runners, bufs):
            p = r.plan
            if p.world == 1:
                continue
            n0, gl, gr = p.n_own, p.n_ghost_left, p.n_ghost_right
            lp, rp = runners[p.left].plan, runners[p.right].plan
            buf[n0:n0 + gl] = bufs[p.left][lp.send_right.long().to(buf.device)]
            buf[n0 + gl:n0 + gl + gr] = bufs[p.right][rp.send_left.long().to(buf.device)]
    e = sum(r.energy for r in runners) + runners[0].model._mean_std()[0]
    f = torch.zeros_like(pos_global)
    for r in runners:
        f[r.plan.own.to(f.device)] = r.forces
    return e, f
