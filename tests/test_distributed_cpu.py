"""Host-side logic of the spatial decomposition (torchmdnet/distributed.py) on CPU: plan construction, ghost
coverage, and the halo exchange protocol with two real processes on the gloo backend (world_size 2 and 3)."""
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


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_plans_partition_and_cover(world):
    from torchmdnet.distributed import ghost_global_indices, slab_plans
    z, pos, box = _system(8)            # 1536 atoms, L = 24.9 A
    cutoff, skin = 4.5, 0.25
    plans = slab_plans(pos, box, cutoff, world, skin=skin)
    own_all = torch.cat([p.own for p in plans])
    assert torch.equal(torch.sort(own_all).values, torch.arange(pos.shape[0]))      # a partition
    nb, _, _ = onb.neighbor_pairs(pos.numpy(), None, box.numpy(), 0.0, cutoff + 2 * skin - 1e-3, rect=True)
    for p in plans:
        ghosts = ghost_global_indices(plans, p.rank)
        assert ghosts.numel() == p.n_ghost_left + p.n_ghost_right
        local = set(p.own.tolist()) | set(ghosts.tolist())
        assert len(local) == p.n_local                                                # no duplicated ghost
        owned = set(p.own.tolist())
        need = {int(j) for i, j in zip(nb[0], nb[1]) if int(i) in owned}
        assert need <= local                                                          # every neighbour is present


def test_too_thin_slabs_are_refused():
    from torchmdnet.distributed import slab_plans
    z, pos, box = _system(5)            # L = 15.5 A
    with pytest.raises(ValueError):
        slab_plans(pos, box, 4.5, 4)
    with pytest.raises(ValueError):
        slab_plans(pos, box, 4.5, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from torchmdnet.distributed import NcclHalo, ghost_global_indices, slab_plans
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z, pos, box = _system(8)
        plans = slab_plans(pos, box, 4.5, world, skin=0.25)
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


@pytest.mark.parametrize("world", [2, 3])
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
