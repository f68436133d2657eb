"""Spatial domain decomposition of one large periodic system across GPUs (BASELINE.json configs[3]; new capability,
the reference has no inference-time multi-GPU path: SURVEY section 5/8e).

One process per GPU.  The box is cut into a px x py x pz grid of bricks (slabs are the grid (P,1,1)).  A rank owns
the atoms whose *base* position lies in its brick and keeps, as the contiguous tail of every local array, ghost
copies of the other ranks' atoms within ``cutoff + 2*skin`` of the brick; ownership and send lists stay fixed while
no atom has moved more than ``skin`` from its base position (the usual MD neighbour-skin contract).  Owned atoms are
stored boundary-first: the rows some other rank needs come before the interior rows.

Per energy+force evaluation (stages of ``tmdn_tensornet_stage_rows``, include/tmdn.h):
    positions halo -> GRAPH -> EMBED -> Y0 halo -> LAYER_FWD(0) -> Y1 halo -> ... -> LAYER_FWD(L-1), HEAD -> B0 halo
    -> LAYER_BWD(L-1) -> B1 halo -> ... -> LAYER_BWD(0) -> B_L halo -> FINAL;  energy = all-reduce of the per-rank sums.
Every halo is an owner -> ghost copy of node rows ([.,9,H] fp32, 4608 B/atom at H=128): because the edge set is
symmetric the backward kernels gather adjoints instead of scattering them, so no reverse (accumulating) halo exists
and the forces of owned atoms are complete locally.

Overlap: every operation between two halo points is local to a row, so each stage runs first over the boundary rows,
their transfer starts on a second stream, and the interior rows are evaluated while it is in flight.

Transport (``transport=``):
  "nccl"  pack kernel (``tmdn_halo_pack_f32``) + grouped NCCL send/recv straight into the ghost rows, on the
          communication stream;
  "peer"  the pack and the transfer are ONE kernel (``tmdn_halo_push_f32``) that gathers the rows and stores them
          straight into the neighbours' ghost rows over NVLink peer memory; the workspaces live in symmetric memory
          (torch.distributed._symmetric_memory: CUDA IPC plumbing only) and "data has landed" is one cross-rank
          barrier per halo point on the communication stream.  No NCCL call in the step.
"""
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import torch

from . import _native

EMBED, LAYER_FWD, HEAD, LAYER_BWD, FINAL, GRAPH = 0, 1, 2, 3, 4, 5


# ------------------------------------------------------------------------------------------------ plans
@dataclass
class DomainPlan:
    rank: int
    world: int
    grid: Tuple[int, int, int]
    own: torch.Tensor                 # int64 global indices of the owned atoms in LOCAL order (boundary rows first)
    n_boundary: int                   # rows [0, n_boundary) are needed by some other rank
    peers: List[int] = field(default_factory=list)             # ranks exchanged with, ascending
    send: List[torch.Tensor] = field(default_factory=list)     # per peer: int32 local indices (< n_boundary) it needs
    recv: List[int] = field(default_factory=list)              # per peer: rows received from it
    peer_ghost_offset: List[int] = field(default_factory=list)  # per peer: offset (rows) of MY rows in ITS ghost region

    @property
    def n_own(self):
        return int(self.own.numel())

    @property
    def n_ghost(self):
        return int(sum(self.recv))

    @property
    def n_local(self):
        return self.n_own + self.n_ghost

    def recv_offset(self, k):
        """first ghost row (relative to n_own) filled by peer number k"""
        return int(sum(self.recv[:k]))


def choose_grid(world: int, box: torch.Tensor, halo: float) -> Tuple[int, int, int]:
    """Brick grid with the smallest ghost volume among the factorisations of ``world`` that fit the halo."""
    L = [float(box[d, d]) for d in range(3)]
    best, best_cost = None, None
    for px in range(1, world + 1):
        if world % px:
            continue
        for py in range(1, world // px + 1):
            if (world // px) % py:
                continue
            p = (px, py, world // px // py)
            w = [L[d] / p[d] for d in range(3)]
            if any(p[d] > 1 and (p[d] - 1) * w[d] < 2 * halo for d in range(3)):
                continue
            vol = 1.0
            for d in range(3):
                vol *= (w[d] + 2 * halo) if p[d] > 1 else w[d]
            cost = (vol, sum(1 for d in range(3) if p[d] > 1))
            if best is None or cost < best_cost:
                best, best_cost = p, cost
    if best is None:
        raise ValueError("no %d-rank brick grid fits a halo of %.2f in this box" % (world, halo))
    return best


def domain_plans(pos: torch.Tensor, box: torch.Tensor, cutoff: float, world: int, skin: float = 0.5,
                 grid: Optional[Sequence[int]] = None) -> List[DomainPlan]:
    """Plans of ALL ranks from the global base positions (deterministic: every rank computes the same list, so no
    size exchange is needed).  ``pos`` [N,3], ``box`` diagonal [3,3]; ``grid`` = (px,py,pz) with px*py*pz == world
    (default: the factorisation with the smallest ghost volume)."""
    halo = cutoff + 2.0 * skin
    boxd = box.detach().double().cpu()
    grid = tuple(int(g) for g in grid) if grid is not None else choose_grid(world, boxd, halo)
    assert grid[0] * grid[1] * grid[2] == world, "grid does not match the number of ranks"
    L = torch.tensor([float(boxd[d, d]) for d in range(3)], dtype=torch.float64)
    w = L / torch.tensor(grid, dtype=torch.float64)
    for d in range(3):
        if grid[d] > 1 and (grid[d] - 1) * float(w[d]) < 2 * halo:
            raise ValueError("brick width %.2f too small for halo %.2f with %d ranks along axis %d"
                             % (float(w[d]), halo, grid[d], d))
    x = pos.detach().double().cpu()
    x = x - torch.floor(x / L + 0.5) * L                                   # [-L/2, L/2)
    cell = torch.floor((x + L / 2) / w).long()
    for d in range(3):
        cell[:, d].clamp_(0, grid[d] - 1)
    owner = (cell[:, 0] * grid[1] + cell[:, 1]) * grid[2] + cell[:, 2]
    N = x.shape[0]
    if world == 1:
        return [DomainPlan(0, 1, grid, torch.arange(N), 0)]
    # need[r] = atoms of other ranks within the halo of brick r (box-shaped halo region, periodic distance per axis)
    need = []
    for r in range(world):
        c = torch.tensor([r // (grid[1] * grid[2]), (r // grid[2]) % grid[1], r % grid[2]], dtype=torch.float64)
        centre = -L / 2 + (c + 0.5) * w
        dx = x - centre
        dx = dx - torch.round(dx / L) * L
        dist = (dx.abs() - w / 2).clamp_min(0.0)
        inside = torch.ones(N, dtype=torch.bool)
        for d in range(3):
            if grid[d] > 1:
                inside &= dist[:, d] < halo
        need.append(inside & (owner != r))
    plans = []
    for r in range(world):
        mine = owner == r
        wanted = torch.zeros(N, dtype=torch.bool)
        for q in range(world):
            if q != r:
                wanted |= need[q] & mine
        boundary = torch.nonzero(mine & wanted).flatten()
        interior = torch.nonzero(mine & ~wanted).flatten()
        plans.append(DomainPlan(r, world, grid, torch.cat([boundary, interior]), int(boundary.numel())))
    local_of = torch.full((N,), -1, dtype=torch.long)
    for p in plans:
        local_of[p.own] = torch.arange(p.n_own)
    for r, p in enumerate(plans):
        for q in range(world):
            if q == r:
                continue
            to_q = torch.nonzero(need[q] & (owner == r)).flatten()        # what q needs from r (ascending global index)
            from_q = int((need[r] & (owner == q)).sum())
            if to_q.numel() == 0 and from_q == 0:
                continue
            p.peers.append(q)
            p.send.append(local_of[to_q].to(torch.int32))
            p.recv.append(from_q)
    for r, p in enumerate(plans):
        for q in p.peers:
            pq = plans[q]
            p.peer_ghost_offset.append(pq.recv_offset(pq.peers.index(r)))
    return plans


def slab_plans(pos: torch.Tensor, box: torch.Tensor, cutoff: float, world: int, skin: float = 0.5,
               axis: int = 0) -> List[DomainPlan]:
    """Slab decomposition along ``axis``: the brick grid (world,1,1) / (1,world,1) / (1,1,world)."""
    grid = [1, 1, 1]
    grid[axis] = world
    return domain_plans(pos, box, cutoff, world, skin, grid)


def ghost_global_indices(plans: List[DomainPlan], rank: int) -> torch.Tensor:
    """Global indices of the ghost rows of ``rank`` in local order (peers ascending, each in its owner's send order)."""
    p = plans[rank]
    out = [torch.zeros(0, dtype=torch.long)]
    for q in p.peers:
        pq = plans[q]
        out.append(pq.own[pq.send[pq.peers.index(rank)].long()])
    return torch.cat(out)


# ------------------------------------------------------------------------------------------------ transports
class NcclHalo:
    """Owner -> ghost exchange of the rows of a local [n_local, F] fp32 tensor: pack kernel + grouped send/recv
    (torch.distributed; NCCL over NVLink on GPUs, gloo in the CPU tests); receives write straight into the ghost rows."""

    def __init__(self, plan: DomainPlan, device, group=None, pack=None):
        self.plan, self.group = plan, group
        self.send = [s.to(device) for s in plan.send]
        # pack = the gather kernel of libtmdn; host-logic tests on the gloo backend pass an index_select instead
        self.pack = pack if pack is not None else (lambda buf, idx: torch.ops.torchmdnet_b200.halo_pack(buf, idx))
        self.bytes_per_row_exchange = 0

    def start(self, buf: torch.Tensor, name: str = ""):
        import torch.distributed as dist
        p = self.plan
        if p.world == 1:
            return
        ops, keep = [], []
        for k, q in enumerate(p.peers):
            if self.send[k].numel() > 0:
                s = self.pack(buf, self.send[k])
                keep.append(s)
                ops.append(dist.P2POp(dist.isend, s, q, self.group))
            if p.recv[k] > 0:
                a = p.n_own + p.recv_offset(k)
                ops.append(dist.P2POp(dist.irecv, buf[a:a + p.recv[k]], q, self.group))
        for req in dist.batch_isend_irecv(ops):
            req.wait()

    def finish(self, buf: torch.Tensor, name: str = ""):
        pass

    def __call__(self, buf: torch.Tensor, name: str = ""):
        self.start(buf, name)
        self.finish(buf, name)


class PeerHalo:
    """Owner -> ghost exchange as ONE kernel per halo point: ``tmdn_halo_push_f32`` gathers the rows every peer needs
    and stores them directly into that peer's ghost rows over NVLink peer memory.  All ranks' local arrays live at
    known offsets of one symmetric-memory allocation per rank (same size everywhere), so a destination address is
    ``peer_base + offset_of(buffer, peer's layout) + (peer.n_own + my offset in its ghosts) * row_bytes``, computed once.
    ``finish`` is a cross-rank barrier on the calling stream: when it returns, every push issued before it has landed."""

    def __init__(self, plans: List[DomainPlan], rank: int, device, symm_handle, base_ptrs: List[int],
                 offset_of: Callable[[int, str], int]):
        self.plan, self.hdl, self.device = plans[rank], symm_handle, device
        p = self.plan
        self.idx = torch.cat([s for s in p.send] + [torch.zeros(0, dtype=torch.int32)]).to(device)
        seg = [0]
        for s in p.send:
            seg.append(seg[-1] + int(s.numel()))
        self.seg = torch.tensor(seg, dtype=torch.long, device=device)
        self.n_rows = seg[-1]
        self._plans, self._base, self._offset_of = plans, base_ptrs, offset_of
        self._dst = {}

    def _dst_ptrs(self, name: str, row_bytes: int):
        key = (name, row_bytes)
        if key not in self._dst:
            p = self.plan
            ptrs = []
            for k, q in enumerate(p.peers):
                pq = self._plans[q]
                ptrs.append(self._base[q] + self._offset_of(q, name) + (pq.n_own + p.peer_ghost_offset[k]) * row_bytes)
            self._dst[key] = torch.tensor(ptrs, dtype=torch.long, device=self.device)
        return self._dst[key]

    def start(self, buf: torch.Tensor, name: str):
        if self.plan.world == 1 or self.n_rows == 0:
            return
        torch.ops.torchmdnet_b200.halo_push(buf, self.idx, self.seg, self._dst_ptrs(name, buf.shape[1] * 4))

    def finish(self, buf: torch.Tensor, name: str = ""):
        if self.plan.world > 1:
            self.hdl.barrier(channel=0)

    def __call__(self, buf: torch.Tensor, name: str):
        self.start(buf, name)
        self.finish(buf, name)


# ------------------------------------------------------------------------------------------------ runner
@dataclass
class _Group:
    """Stages evaluated between two halo points and the buffer whose ghost rows the next group needs."""
    stages: List[Tuple[int, int]]
    produces: Optional[str]


def stage_groups(L: int) -> List[_Group]:
    """GRAPH needs the positions halo; B_k = "Dias" (k even) / "Dias1" (k odd), see include/tmdn.h."""
    B = lambda k: "Dias1" if k & 1 else "Dias"    # noqa: E731
    if L == 0:
        return [_Group([(GRAPH, 0), (EMBED, 0), (HEAD, 0)], B(0)), _Group([(FINAL, 0)], None)]
    g = [_Group([(GRAPH, 0), (EMBED, 0)], "Y0")]
    for l in range(L - 1):
        g.append(_Group([(LAYER_FWD, l)], "Y%d" % (l + 1)))
    g.append(_Group([(LAYER_FWD, L - 1), (HEAD, 0)], B(0)))
    for l in range(L - 1, -1, -1):
        g.append(_Group([(LAYER_BWD, l)], B(L - l)))
    g.append(_Group([(FINAL, 0)], None))
    return g


class DecomposedTensorNet:
    """Runs a TensorNet ``TorchMD_Net`` on one brick of a decomposed periodic system.

    ``step(pos_own)`` evaluates energy and forces of the owned atoms (``pos_own`` in the plan's local order,
    ``plan.own``): every group of stages runs over the boundary rows, the halo of the buffer it produces starts on the
    communication stream, and the interior rows are evaluated while it is in flight.  ``transport``: "nccl" | "peer"
    (module docstring); a custom ``halo`` object with start/finish (the single-process emulation, the gloo tests)
    replaces both."""

    def __init__(self, model, plans: List[DomainPlan], rank: int, z_global: torch.Tensor, box: torch.Tensor, device,
                 halo=None, strategy: str = "cell", transport: str = "nccl", group=None, overlap: bool = True):
        _native.load()
        self.model, self.plans, self.plan, self.device = model, plans, plans[rank], torch.device(device)
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
        self.H, self.L = rm.hidden_channels, rm.num_layers
        self.groups = stage_groups(self.L)
        self.overlap = overlap and p.world > 1 and 0 < p.n_boundary < p.n_own
        _, self.std = model._mean_std()
        # own packed copy of the weights with mean = 0 (the mean is added once, after the all-reduce)
        ops = torch.ops.torchmdnet_b200
        w, ip, fp = rm.pack_weights(model.output_model, 0.0, self.std)
        self.handle = ops.tensornet_create(w, ip, fp)
        self._layout = {}
        nbytes = self._ws_bytes(rank)
        pos_bytes = self._pos_bytes()
        self.transport = transport if halo is None else "custom"
        if self.transport == "peer" and p.world > 1:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            total = pos_bytes + max(self._ws_bytes(r) for r in range(p.world))
            self.arena = symm_mem.empty(total, dtype=torch.uint8, device=self.device)
            hdl = symm_mem.rendezvous(self.arena, group if group is not None else dist.group.WORLD)
            base = [int(x) for x in hdl.buffer_ptrs]
            self.halo = PeerHalo(plans, rank, self.device, hdl, base, self._arena_offset)
        else:
            self.arena = torch.empty(pos_bytes + nbytes, dtype=torch.uint8, device=self.device)
            self.halo = halo if halo is not None else NcclHalo(p, self.device, group)
        self.pos = self.arena[:p.n_local * 12].view(torch.float32).view(p.n_local, 3)
        self.pos.zero_()
        self.ws = self.arena[pos_bytes:pos_bytes + nbytes]
        self.energy = torch.zeros(1, 1, device=self.device)
        self.forces = torch.zeros(p.n_own, 3, device=self.device)
        self.status = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.comm_stream = torch.cuda.Stream(device=self.device) if self.overlap else None
        self.exchange_events = None     # set to [] to record CUDA events around every exchange (exchange_report)

    def __del__(self):
        try:
            torch.ops.torchmdnet_b200.tensornet_destroy(self.handle)
        except Exception:
            pass

    # ---- layout of the rank-local arrays inside the arena (identical code on every rank: peers' offsets are computable)
    def _max_edges(self, r):
        return int(self.rm.max_num_neighbors) * self.plans[r].n_local

    def _pos_bytes(self):
        n = max(p.n_local for p in self.plans)
        return (n * 12 + 255) // 256 * 256

    def _ws_bytes(self, r):
        return int(torch.ops.torchmdnet_b200.tensornet_workspace_bytes(self.handle, self.plans[r].n_local, self._max_edges(r),
                                                                       1, self.strategy, self.box))

    def _lookup(self, r, name):
        key = (r, name)
        if key not in self._layout:
            self._layout[key] = tuple(torch.ops.torchmdnet_b200.tensornet_debug_lookup(
                self.handle, self.plans[r].n_local, self._max_edges(r), 1, self.strategy, self.box, name))
        return self._layout[key]

    def _arena_offset(self, r, name):
        """byte offset of rank r's array ``name`` inside ITS arena ("pos" or a workspace buffer)"""
        return 0 if name == "pos" else self._pos_bytes() + self._lookup(r, name)[0]

    def _buffer(self, name):
        if name == "pos":
            return self.pos
        off, nbytes = self._lookup(self.plan.rank, name)
        return self.ws[off:off + nbytes].view(torch.float32).view(self.plan.n_local, 9 * self.H)

    def _stage(self, stage, layer, r0, r1):
        if r1 <= r0 and stage != GRAPH:
            return
        p = self.plan
        torch.ops.torchmdnet_b200.tensornet_stage_rows(self.handle, stage, layer, self.z, self.pos, self.batch, p.n_own,
                                                       r0, r1, self.box, True, self.strategy,
                                                       self._max_edges(p.rank), 1, self.energy, self.forces, self.status,
                                                       self.ws)

    def _run_group(self, g: _Group, r0, r1, with_graph):
        for st, l in g.stages:
            if st == GRAPH:
                if with_graph:
                    self._stage(GRAPH, 0, 0, self.plan.n_own)
            else:
                self._stage(st, l, r0, r1)

    # ---- evaluation
    def step_stages(self, pos_own: torch.Tensor):
        """Generator over the halo points (no overlap): yields (buffer, name) whose ghost rows must be filled before it
        is resumed.  Used by the single-process emulation and the host-logic tests."""
        p = self.plan
        self.pos[:p.n_own].copy_(pos_own)
        yield self.pos, "pos"
        for g in self.groups:
            # boundary rows first, then interior rows: the same split (and therefore the same bits) as the overlapped step
            nb = p.n_boundary if 0 < p.n_boundary < p.n_own else p.n_own
            self._run_group(g, 0, nb, True)
            self._run_group(g, nb, p.n_own, False)
            if g.produces is not None:
                yield self._buffer(g.produces), g.produces

    def _exchange_begin(self, buf, name):
        """pack + transfer on the current (communication) stream; with ``exchange_events`` set, CUDA events bracket it"""
        rec = self.exchange_events is not None
        if rec:
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
        self.halo.start(buf, name)
        if rec:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            self.exchange_events.append({"name": name, "start": e0, "sent": e1, "row_bytes": buf.shape[1] * 4})

    def _exchange_end(self, buf, name, interior_done=None):
        """completion (peer transport: the cross-rank barrier).  ``interior_done``: event recorded on the compute stream
        after the interior rows -- the part of the exchange that ends after it is what the overlap did not hide."""
        self.halo.finish(buf, name)
        if self.exchange_events is not None:
            e2 = torch.cuda.Event(enable_timing=True)
            e2.record()
            self.exchange_events[-1].update(done=e2, interior_done=interior_done)

    def step(self, pos_own: torch.Tensor):
        """-> (sum of the owned atoms' energies [1,1] (no mean), forces of the owned atoms [n_own,3])."""
        p = self.plan
        if not self.overlap:
            for buf, name in self.step_stages(pos_own):
                if p.world > 1:
                    self._exchange_begin(buf, name)
                    self._exchange_end(buf, name)
            return self.energy, self.forces
        cs, ms = torch.cuda.current_stream(self.device), self.comm_stream
        nb = p.n_boundary
        self.pos[:p.n_own].copy_(pos_own)
        ev = torch.cuda.Event()
        ev.record(cs)                       # everything of the previous step and the new positions
        with torch.cuda.stream(ms):
            ms.wait_event(ev)
            self._exchange_begin(self.pos, "pos")
            self._exchange_end(self.pos, "pos")
            ev_c = torch.cuda.Event()
            ev_c.record(ms)
        cs.wait_event(ev_c)
        for g in self.groups:
            self._run_group(g, 0, nb, True)                     # rows the neighbours are waiting for
            if g.produces is not None:
                buf = self._buffer(g.produces)
                ev_b = torch.cuda.Event()
                ev_b.record(cs)
                with torch.cuda.stream(ms):
                    ms.wait_event(ev_b)
                    self._exchange_begin(buf, g.produces)       # in flight during the interior rows
            self._run_group(g, nb, p.n_own, False)
            if g.produces is not None:
                ev_i = torch.cuda.Event(enable_timing=self.exchange_events is not None)
                ev_i.record(cs)
                with torch.cuda.stream(ms):
                    # completion is only signalled once the local readers of the buffer this halo point's successor will
                    # overwrite are done (the ping-pong argument of include/tmdn.h needs every rank past its interior rows)
                    ms.wait_event(ev_i)
                    self._exchange_end(buf, g.produces, ev_i)
                    ev_c = torch.cuda.Event()
                    ev_c.record(ms)
                cs.wait_event(ev_c)
        return self.energy, self.forces

    def total_energy(self, group=None):
        """All-reduce of the per-rank sums + the model's mean (one scalar per evaluation)."""
        import torch.distributed as dist
        e = self.energy.clone()
        if self.plan.world > 1:
            dist.all_reduce(e, group=group)
        return e + self.model._mean_std()[0]

    def exchange_report(self):
        """From the recorded events (after a synchronize): per halo point the duration of pack + transfer on the
        communication stream and the part of the exchange that ended after the interior rows were done (exposed)."""
        rows = []
        for e in self.exchange_events or []:
            r = {"name": e["name"], "transfer_ms": e["start"].elapsed_time(e["sent"])}
            if e.get("interior_done") is not None:
                r["exposed_ms"] = max(0.0, e["interior_done"].elapsed_time(e["done"]))
            elif "done" in e:
                r["exposed_ms"] = e["start"].elapsed_time(e["done"])         # not overlapped: all of it is exposed
            rows.append(r)
        return rows

    def halo_rows_per_step(self):
        """(rows sent per halo point, number of node-row halo points): bytes = rows * (12 + n_points * 36 * H)"""
        return int(sum(int(s.numel()) for s in self.plan.send)), len([g for g in self.groups if g.produces])


def emulate_step(runners: List[DecomposedTensorNet], pos_global: torch.Tensor):
    """All ranks of a decomposition in ONE process on one GPU (tests): the stage generators advance in lock step and
    every halo point is served by direct device copies.  -> (energy [1,1] incl. mean, forces [N,3] in global order)."""
    gens = [r.step_stages(pos_global[r.plan.own.to(pos_global.device)]) for r in runners]
    while True:
        items = []
        for g in gens:
            try:
                items.append(next(g))
            except StopIteration:
                items.append(None)
        if items[0] is None:
            break
        for r, it in zip(runners, items):
            p = r.plan
            buf = it[0]
            for k, q in enumerate(p.peers):
                pq = runners[q].plan
                src = items[q][0][pq.send[pq.peers.index(p.rank)].long().to(buf.device)]
                a = p.n_own + p.recv_offset(k)
                buf[a:a + p.recv[k]] = src
    e = sum(r.energy for r in runners) + runners[0].model._mean_std()[0]
    f = torch.zeros_like(pos_global)
    for r in runners:
        f[r.plan.own.to(f.device)] = r.forces
    return e, f
