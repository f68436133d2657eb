"""Host-side logic of the spatial decomposition (torchmdnet/distributed.py) on CPU: plan construction, ghost
coverage, boundary-first ordering, and the halo exchange protocol with real processes on the gloo backend (slabs at
world_size 2 and 3, a 2x2x1 brick grid at world_size 4)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import synth_inputs as S
from oracle import neighbors as onb


def _system(n_side=6):
    z, pos, batch, box = S.water_box(n_side, seed=4)
    return torch.from_numpy(z), torch.from_numpy(pos), torch.from_numpy(box)


def _check_plans(plans, pos, box, cutoff, skin):
    from torchmdnet.distributed import ghost_global_indices
    own_all = torch.cat([p.own for p in plans])
    assert torch.equal(torch.sort(own_all).values, torch.arange(pos.shape[0]))      # a partition
    nb, _, _ = onb.neighbor_pairs(pos.numpy(), None, box.numpy(), 0.0, cutoff + 2 * skin - 1e-3, rect=True)
    for p in plans:
        ghosts = ghost_global_indices(plans, p.rank)
        assert ghosts.numel() == p.n_ghost == sum(p.recv)
        local = set(p.own.tolist()) | set(ghosts.tolist())
        assert len(local) == p.n_local                                                # no duplicated ghost
        owned = set(p.own.tolist())
        need = {int(j) for i, j in zip(nb[0], nb[1]) if int(i) in owned}
        assert need <= local                                                          # every neighbour is present
        # boundary rows first: everything a peer needs lies in [0, n_boundary), and every boundary row is needed
        sent = torch.cat([s.long() for s in p.send] + [torch.zeros(0, dtype=torch.long)])
        assert sent.numel() == 0 or int(sent.max()) < p.n_boundary
        assert set(sent.tolist()) == set(range(p.n_boundary))
        # what I send to q is what q receives from me, and I know where it lands in q's ghost rows
        for k, q in enumerate(p.peers):
            pq = plans[q]
            kq = pq.peers.index(p.rank)
            assert pq.recv[kq] == p.send[k].numel() and p.peer_ghost_offset[k] == pq.recv_offset(kq)
            gq = ghost_global_indices(plans, q)
            assert torch.equal(gq[pq.recv_offset(kq):pq.recv_offset(kq) + pq.recv[kq]], p.own[p.send[k].long()])


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_plans_partition_and_cover(world):
    from torchmdnet.distributed import slab_plans
    z, pos, box = _system(8)            # 1536 atoms, L = 24.9 A
    cutoff, skin = 4.5, 0.25
    plans = slab_plans(pos, box, cutoff, world, skin=skin)
    assert plans[0].grid == (world, 1, 1)
    _check_plans(plans, pos, box, cutoff, skin)


@pytest.mark.parametrize("grid", [(2, 2, 1), (2, 2, 2), (1, 3, 2)])
def test_brick_plans_partition_and_cover(grid):
    from torchmdnet.distributed import choose_grid, domain_plans
    z, pos, box = _system(10)           # 3000 atoms, L = 31.1 A: bricks of 15.5 A (10.4 A along a 3-way axis)
    cutoff, skin = 4.5, 0.25
    world = grid[0] * grid[1] * grid[2]
    plans = domain_plans(pos, box, cutoff, world, skin=skin, grid=grid)
    _check_plans(plans, pos, box, cutoff, skin)
    if grid == (2, 2, 2):
        assert all(len(p.peers) == 7 for p in plans)
        # the automatic choice is the factorisation with the smallest owned + ghost volume (1 M-atom box: L = 217.5 A)
        halo, Lb = cutoff + 2 * skin, 217.5
        vol = lambda g: np.prod([(Lb / n + 2 * halo) if n > 1 else Lb for n in g])    # noqa: E731
        best = choose_grid(8, torch.eye(3, dtype=torch.float64) * Lb, halo)
        assert all(vol(best) <= vol(g) + 1e-6 for g in [(8, 1, 1), (2, 2, 2), (1, 2, 4), (4, 2, 1)])
        assert vol(best) < vol((8, 1, 1))


def test_stage_groups_cover_every_stage_once():
    from torchmdnet.distributed import EMBED, FINAL, GRAPH, HEAD, LAYER_BWD, LAYER_FWD, stage_groups
    for L in (0, 1, 2, 3):
        g = stage_groups(L)
        flat = [s for grp in g for s in grp.stages]
        expect = [(GRAPH, 0), (EMBED, 0)] + [(LAYER_FWD, l) for l in range(L)] + [(HEAD, 0)] + \
                 [(LAYER_BWD, l) for l in range(L - 1, -1, -1)] + [(FINAL, 0)]
        assert flat == expect
        names = [grp.produces for grp in g]
        assert names[-1] is None and all(n is not None for n in names[:-1])
        back = [n for n in names if n and n.startswith("Dias")]
        assert back == [("Dias1" if k & 1 else "Dias") for k in range(L + 1)]        # ping-pong
        assert [n for n in names if n and n.startswith("Y")] == ["Y%d" % l for l in range(L)]


def test_too_thin_slabs_are_refused():
    """A brick plus its halo on both sides must fit into the box ((p-1) * width >= 2 * halo), otherwise an atom would be
    needed through two periodic images.  Narrower-than-halo bricks that satisfy it are fine: peers are not only the
    adjacent bricks."""
    from torchmdnet.distributed import slab_plans
    z, pos, box = _system(5)            # L = 15.5 A, halo 5.5 A
    with pytest.raises(ValueError):
        slab_plans(pos, box, 4.5, 2)    # 7.8 A slabs: 7.8 < 11
    with pytest.raises(ValueError):
        slab_plans(pos, box, 4.5, 3)    # 5.2 A slabs: 10.3 < 11
    plans = slab_plans(pos, box, 4.5, 4)    # 3.9 A slabs: 11.6 >= 11, every rank exchanges with all three others
    assert all(len(p.peers) == 3 for p in plans)
    _check_plans(plans, pos, box, 4.5, 0.5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from torchmdnet.distributed import NcclHalo, domain_plans, ghost_global_indices, slab_plans
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z, pos, box = _system(10 if world == 4 else 8)
        plans = domain_plans(pos, box, 4.5, 4, skin=0.25, grid=(2, 2, 1)) if world == 4 else \
            slab_plans(pos, box, 4.5, world, skin=0.25)
        p = plans[rank]
        halo = NcclHalo(p, "cpu", pack=lambda buf, idx: buf[idx.long()].contiguous())
        F = 36
        feat = lambda g: (g.double()[:, None] * 0.5 + torch.arange(F)[None, :].double()).float()   # noqa: E731
        buf = torch.full((p.n_local, F), -1.0)
        buf[:p.n_own] = feat(p.own)
        halo(buf)
        expect = feat(torch.cat([p.own, ghost_global_indices(plans, rank)]))
        ok = bool(torch.equal(buf, expect))
        # positions travel the same way (rows of 3 floats)
        pb = torch.zeros(p.n_local, 3)
        pb[:p.n_own] = pos[p.own]
        halo(pb)
        ok = ok and bool(torch.equal(pb, pos[torch.cat([p.own, ghost_global_indices(plans, rank)])]))
        e = torch.tensor([float(rank + 1)])
        dist.all_reduce(e)
        ok = ok and float(e) == world * (world + 1) / 2
        out[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3, 4])
def test_halo_exchange_gloo(world):
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(120)
        assert pr.exitcode == 0
    assert all(out.get(r) for r in range(world)), dict(out)
