"""torchrun helper of tests/test_gpu_distributed.py: every rank evaluates its brick of a periodic water box (transport =
argv[1]: nccl | peer)
and rank 0 compares energy and forces with its own single-domain evaluation of the whole box."""
import os
import sys

import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "torchmd-net_b200"), HERE):
    sys.path.insert(0, p)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from test_gpu_distributed import _model_and_system
    from torchmdnet.distributed import DecomposedTensorNet, domain_plans
    transport = sys.argv[1] if len(sys.argv) > 1 else "nccl"
    m, hp, z, pos, box = _model_and_system(12)          # 5184 atoms, L = 37.3 A
    m = m.to(dev)
    grid = (2, 2, 1) if world == 4 else (world, 1, 1)
    plans = domain_plans(pos, box, hp["cutoff_upper"], world, skin=0.25, grid=grid)
    runner = DecomposedTensorNet(m, plans, rank, z, box, dev, transport=transport)
    assert runner.overlap
    own = plans[rank].own.to(dev)
    for _ in range(3):
        runner.step(pos.to(dev)[own])
    e = runner.total_energy()
    sizes = [p.n_own for p in plans]
    fg = [torch.zeros(s, 3, device=dev) for s in sizes]
    dist.all_gather(fg, runner.forces)
    ok = True
    if rank == 0:
        f = torch.zeros(pos.shape[0], 3, device=dev)
        for p, t in zip(plans, fg):
            f[p.own.to(dev)] = t
        y_ref, f_ref = m(z.to(dev), pos.to(dev))
        de, df = float((e - y_ref).abs().max()), float((f - f_ref).abs().max())
        ok = de <= 2e-5 + 2e-6 * float(y_ref.abs().max()) and df <= 2e-5
        print("world %d %s: |dE| = %.2e, max|dF| = %.2e -> %s" % (world, transport, de, df, "DIST_OK" if ok else "DIST_FAIL"))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
