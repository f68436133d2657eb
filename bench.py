#!/usr/bin/env python
"""bench.py -- atom-steps/s (energy + forces) of the B200-native TorchMD-NET hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload chignolin|md17|water|qm9-et|md17-et] [--filter table|exact]
    python bench.py --impl reference ...      # the reference algorithm's CPU restatement on the host cores

One "step" = one energy+force evaluation (neighbour search included) of the workload on one GPU:
  chignolin (default; BASELINE.json configs[1]): TensorNet, TensorNet-SPICE hparams, one 166-atom system, fresh
            positions every step (MD-step variant);
  md17      (configs[2]): TensorNet-rMD17 hparams, 8192 21-atom molecules per GPU;
  water     (configs[3], single-GPU slice): TensorNet-ANI1X hparams, periodic water box, cell-list neighbours.
With --gpus N > 1 (torchrun, one rank per GPU) every rank evaluates its own shard / replica with no collective in
the step (molecules and replicas are independent): weak scaling, value = N * atoms / max-over-ranks step time.
Every run of the default workload ALSO evaluates configs[3] on all N GPUs together -- one 1 029 000-atom periodic water
box, monolithic at N = 1, brick-decomposed with halo exchange over NVLink at N > 1 -- and reports it as the
"strong_scaling" object of the same JSON line (ms_per_step, atom-steps/s, ghost atoms, halo bytes / duration / exposed
time, launches, roofline of tn_msg_bwd, and the decomposed result checked against rank 0's monolithic evaluation).

Timing: W >= 3 warm-up steps; L2 is flushed (a 512 MiB write) before every timed step, outside the timed span;
each step is bracketed by CUDA events on the launching stream; ms_per_step = sum of the K spans / K, max over
ranks.  `value` runs the whole step as one CUDA graph with inputs resident in HBM; `e2e` goes through the public
calculators.External.calculate() with pinned HOST positions in and host energy+forces out, copies timed.
"""
import argparse
import json
import os
import random
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torchmd-net_b200"))
import synth_inputs as S  # noqa: E402

# rank 0 prints ONE JSON line on stdout.  Libraries write there too (c10d prints "NCCL version ..." when the first
# communicator is created), so file descriptor 1 is pointed at stderr for the whole run and the line goes to a
# duplicate of the original stdout.
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())

METRIC = "atom-steps/sec (energy+forces) TensorNet/ET at 1/2/4/8 B200; neighbor-list ms"
UNIT = "atom-steps/s"

HP_BASE = dict(model="tensornet", precision=32, trainable_rbf=False, activation="silu", cutoff_lower=0.0,
               max_z=128, rbf_type="expnorm", equivariance_invariance_group="O(3)", derivative=True, atom_filter=-1,
               prior_model=None, output_model="Scalar", reduce_op="add", num_layers=2, embedding_dimension=128)
# hparams of the reference's example yamls (examples/TensorNet-SPICE.yaml, TensorNet-rMD17.yaml, TensorNet-ANI1X.yaml)
HPARAMS = {
    "chignolin": dict(HP_BASE, num_rbf=64, cutoff_upper=10.0, max_num_neighbors=128),
    "md17": dict(HP_BASE, num_rbf=32, cutoff_upper=4.5, max_num_neighbors=32),
    "water": dict(HP_BASE, num_rbf=32, cutoff_upper=4.5, max_num_neighbors=128),
}
HP_ET = dict(model="equivariant-transformer", precision=32, trainable_rbf=False, activation="silu",
             attn_activation="silu", cutoff_lower=0.0, cutoff_upper=5.0, max_z=100, rbf_type="expnorm",
             neighbor_embedding=True, distance_influence="both", derivative=True, atom_filter=-1, prior_model=None,
             output_model="Scalar", reduce_op="add", num_heads=8)
# examples/ET-QM9.yaml and examples/ET-MD17.yaml
HPARAMS["qm9-et"] = dict(HP_ET, embedding_dimension=256, num_layers=8, num_rbf=64, max_num_neighbors=64)
HPARAMS["md17-et"] = dict(HP_ET, embedding_dimension=128, num_layers=6, num_rbf=32, max_num_neighbors=32)
WORKLOAD_DESC = {
    "qm9-et": "ET-QM9 hparams, 128 synthetic QM9-size molecules (3-29 atoms) per GPU, batched energy+forces",
    "md17-et": "ET-MD17 hparams, 8192 synthetic 21-atom molecules per GPU, batched energy+forces",
    "chignolin": "TensorNet-SPICE hparams, 166-atom chignolin-size molecule, single-system MD energy+force step",
    "md17": "TensorNet-rMD17 hparams, 8192 synthetic 21-atom molecules per GPU, batched energy+forces",
    "water": "TensorNet-ANI1X hparams, periodic water box, cell-list neighbours, energy+forces",
}


def seed_everything(s):
    random.seed(s)
    np.random.seed(s)
    torch.manual_seed(s)


def make_inputs(workload, rank, args):
    """-> z, pos0, batch, box (numpy).  Rank r draws its own seeded shard/replica."""
    if workload == "chignolin":
        z, pos, batch = S.chignolin_like(seed=2 + 1000 * rank)
        return z, pos, batch, None
    if workload in ("md17", "md17-et"):
        z, pos, batch = S.md17_like_batch(n_mol=args.molecules, seed=3 + 1000 * rank)
        return z, pos, batch, None
    if workload == "qm9-et":
        z, pos, batch = S.qm9_like_batch(n_mol=128, seed=1 + 1000 * rank)
        return z, pos, batch, None
    z, pos, batch, box = S.water_box(args.water_side, seed=4 + 1000 * rank, order=args.water_order)
    return z, pos, batch, box


def step_positions(pos0, n, seed):
    """Fresh positions per step: N(0, 0.05 A) displacements of the base geometry (cfg 2 MD variant)."""
    return [S.md_step_noise(pos0, i, seed=seed) for i in range(n)]


# ------------------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm = sorted(float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(rows)}


# ------------------------------------------------------------------------------------------- CPU (reference) arm
def cpu_reference_steps(workload, hp, args, steps, warmup, wall_budget_s=0.0):
    """Times oracle/model.py (the CPU restatement of the reference's PyTorch path, pinned to the reference's
    goldens) on all host threads.  -> (atom-steps/s, cores, sample description)."""
    from oracle import model as om
    from torchmdnet.models.model import create_model
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    seed_everything(1234)
    sd = {k: v.clone() for k, v in create_model(hp).state_dict().items()}
    if workload in ("chignolin", "qm9-et"):
        z, pos0, batch, box = make_inputs(workload, 0, args)
        sample = "the full %d-atom step" % z.shape[0]
    elif workload in ("md17", "md17-et"):
        z, pos0, batch = S.md17_like_batch(n_mol=64, seed=3)
        box = None
        sample = "64 of the 8192 molecules per step (the reference CPU op is O(N^2) over the whole batch)"
    else:
        z, pos0, batch, box = S.water_box(8, seed=4)
        sample = "a 1536-atom periodic box (8^3 waters) per step"
    tz, tb = torch.from_numpy(z), torch.from_numpy(batch)
    tbox = None if box is None else torch.from_numpy(box)
    times = []
    t_start = time.perf_counter()
    for i in range(steps + warmup):
        p = S.md_step_noise(pos0, i, seed=2)
        t0 = time.perf_counter()
        om.energy_forces(sd, hp, tz, torch.from_numpy(p), tb, box=tbox)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
        # bounded sample: the requested step count is honoured unless the wall budget runs out first
        if wall_budget_s and times and time.perf_counter() - t_start > wall_budget_s:
            break
    per = sum(times) / len(times)
    return (z.shape[0] / per, cores, "%s, %d timed step(s) after %d warm-up, %.3f s/step" % (sample, len(times), warmup, per),
            per, len(times))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    hp = HPARAMS[args.workload]
    # --steps / --warmup are honoured; a wall budget (not a step cap) keeps the run within a few minutes
    warmup = max(1, args.warmup)
    v, cores, sample, per, steps = cpu_reference_steps(args.workload, hp, args, max(1, args.steps), warmup, wall_budget_s=150.0)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": per * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[args.workload], "device": "host CPU, %d threads" % cores},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "note": "oracle/model.py: the CPU restatement of the reference's PyTorch path (pinned to its "
                                     "goldens), not the reference package itself, which cannot be installed here (DESIGN.md 2); "
                                     "the survey measured the reference's own code at 0.97 s/step on 8 threads"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------- roofline
def algorithmic_bytes(family, N, E, H, L):
    """Per-LAUNCH algorithmic HBM bytes of a kernel family (SURVEY 8d per-unit figures)."""
    node = 36 * H * N                       # one [N,9,H] fp32 tensor
    table = {
        "tn_msg_fwd": 2 * node + 8 * E,     # read Y, write M; per edge: i32 column + f32 distance
        "tn_msg_bwd": 4 * node + 16 * E,    # read Y, dM, dY; write dY (backward = 2x forward)
        "tn_embed_edge_fwd": node + 20 * E,
        "tn_embed_edge_bwd": node + 20 * E + 12 * N,
        "neighbors_csr": 16 * N + 20 * E,
        "et_edge_fwd": 48 * H * N + 20 * E,     # read q,k,v,vec once, write xa,va; CSR (SURVEY 8d B_et)
        "et_edge_bwd": 96 * H * N + 40 * E,     # backward = 2x forward
    }
    return table.get(family, 2 * node)


def l2_bytes(family, E, H):
    """Bytes one launch moves L2 -> SM through its gathers: table row + neighbour rows per edge-channel."""
    per = {"tn_msg_fwd": 84, "tn_msg_bwd": 120, "tn_embed_edge_fwd": 48, "tn_embed_edge_bwd": 48,
           "et_edge_fwd": 92, "et_edge_bwd": 112}
    return per.get(family, 0) * H * E


def ncu_capture(workload, family):
    """dram bytes / L2 throughput of `family` from the committed ncu full capture of this workload, if there is one
    (profiles/ncu_traffic.json, written from `ncu --set full` reports by profiles/summarize.py traffic)."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return None
    with open(path) as fh:
        return json.load(fh).get(workload, {}).get(family)


def profile_families(run_step, steps):
    torch.ops.torchmdnet_b200.profile_enable(True)
    for i in range(steps):
        run_step(i)
    torch.cuda.synchronize()
    rep = torch.ops.torchmdnet_b200.profile_report()
    torch.ops.torchmdnet_b200.profile_enable(False)
    fam = {}
    for ln in rep.strip().splitlines():
        name, cnt, ms = ln.split()
        fam[name] = (int(cnt), float(ms))
    return fam


# ------------------------------------------------------------------------------------------- strong-scaling arm
def strong_scaling(args, dev, rank, world, local):
    """BASELINE.json configs[3]: ONE periodic water box (TensorNet-ANI1X hparams, --strong-side^3 molecules; default
    70 -> 1 029 000 atoms), energy + forces.  N = 1: the monolithic call.  N > 1: brick decomposition over N GPUs
    (torchmdnet/distributed.py), halo exchange overlapped with the interior rows; the result is checked inside the
    bench against rank 0's monolithic evaluation of the same positions (north_star tolerances).  Strong scaling: the
    box is fixed, value = atoms / max-over-ranks step time.  -> dict on rank 0, None elsewhere."""
    import torch.distributed as dist
    from torchmdnet.distributed import DecomposedTensorNet, domain_plans
    from torchmdnet.models.model import create_model
    from torchmdnet.models.utils import OptimizedDistance
    hp = HPARAMS["water"]
    H, L = hp["embedding_dimension"], hp["num_layers"]
    seed_everything(1234)
    model = create_model(hp).to(dev).eval()
    rm = model.representation_model
    rm.filter_mode = "table"
    z_np, pos0, batch_np, box = S.water_box(args.strong_side, seed=4)
    N = z_np.shape[0]
    K, W = max(1, min(args.steps, args.strong_steps)), 3
    n_pos = 4
    tbox = torch.from_numpy(box)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    ops = torch.ops.torchmdnet_b200

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks_seen = {}

    def timed(step_fn, k, w, sample_clocks=False):
        for i in range(w):
            step_fn(i)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
        sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
        barrier()
        t_begin = time.perf_counter()
        for i in range(k):
            flush.zero_()
            if world > 1:
                dist.barrier()          # every rank starts the step together (device-side: NCCL on the current stream)
            ev[i][0].record()
            step_fn(w + i)
            ev[i][1].record()
        barrier()
        if sampler is not None:
            clocks_seen.update(sampler.stop(t_begin, time.perf_counter()))
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) / k

    def monolithic():
        rm.distance = OptimizedDistance(hp["cutoff_lower"], hp["cutoff_upper"], max_num_pairs=-hp["max_num_neighbors"],
                                        return_vecs=True, loop=True, check_errors=False, resize_to_fit=False, box=tbox,
                                        strategy="cell")
        return torch.from_numpy(z_np).to(dev), torch.from_numpy(batch_np).to(dev)

    out = {"atoms_total": N, "workload": WORKLOAD_DESC["water"], "box_A": float(box[0, 0]), "n_gpus": world,
           "steps": K, "warmup": W, "l2": "flushed (512 MiB write) before every timed step"}
    if world == 1:
        z, batch = monolithic()
        poss = torch.stack([torch.from_numpy(S.md_step_noise(pos0, i, seed=2)) for i in range(n_pos)]).to(dev)

        def step(i):
            return model(z, poss[i % n_pos], batch)
        step(0)
        torch.cuda.synchronize()
        c0 = ops.launch_count()
        y, f = step(0)
        torch.cuda.synchronize()
        launches = ops.launch_count() - c0
        status = rm.energy_forces(z, poss[0], batch, 1, model.output_model, *model._mean_std())[2].cpu()
        assert status[1:].tolist() == [0, 0, 0], "bad status %s" % status.tolist()
        E = int(status[0])
        ms = timed(step, K, W, sample_clocks=True)
        fam = profile_families(step, 2)
        out.update({"ms_per_step": ms, "value": N / (ms * 1e-3), "unit": UNIT, "parallelism": "one GPU, monolithic call",
                    "ghost_atoms_max": 0, "halo_bytes_per_step": 0, "halo_ms_per_step": 0.0,
                    "halo_exposed_ms_per_step": 0.0, "launches_per_step": int(launches), "edges": E,
                    "correctness": {"reference": "this IS the monolithic evaluation", "max_abs_dF": 0.0, "rel_dE": 0.0},
                    "energy": float(y)})
        n_own, n_prof = N, 2
    else:
        grid = tuple(int(x) for x in args.grid.split("x")) if args.grid else None
        plans = domain_plans(torch.from_numpy(pos0), tbox, hp["cutoff_upper"], world, skin=0.5, grid=grid)
        plan = plans[rank]
        transport, note = args.transport, None
        try:
            runner = DecomposedTensorNet(model, plans, rank, torch.from_numpy(z_np), tbox, dev, transport=transport)
        except Exception as exc:            # symmetric memory unavailable on this box: NCCL transport
            if transport != "peer":
                raise
            note = "peer transport unavailable (%s: %s); NCCL send/recv used" % (type(exc).__name__, str(exc)[:120])
            transport = "nccl"
            runner = DecomposedTensorNet(model, plans, rank, torch.from_numpy(z_np), tbox, dev, transport=transport)
        own = plan.own
        poss = torch.stack([torch.from_numpy(S.md_step_noise(pos0, i, seed=2)[own.numpy()]) for i in range(n_pos)]).to(dev)

        def step(i):
            runner.step(poss[i % n_pos])
            return runner.total_energy()
        step(0)
        torch.cuda.synchronize()
        c0 = ops.launch_count()
        step(0)
        torch.cuda.synchronize()
        launches = ops.launch_count() - c0
        assert runner.status.cpu()[1:].tolist() == [0, 0, 0], "bad status on rank %d" % rank
        ms = timed(step, K, W, sample_clocks=True)
        # exchange durations: CUDA events around the pack + transfer of every halo point (communication stream)
        runner.exchange_events = []
        step(0)
        step(1)
        torch.cuda.synchronize()
        ex = runner.exchange_report()
        runner.exchange_events = None
        halo_ms = torch.tensor([sum(r["transfer_ms"] for r in ex) / 2, sum(r.get("exposed_ms", 0.0) for r in ex) / 2],
                               device=dev, dtype=torch.float64)
        dist.all_reduce(halo_ms, op=dist.ReduceOp.MAX)
        per_point = {}
        for r in ex:
            d = per_point.setdefault(r["name"], {"transfer_ms": 0.0, "exposed_ms": 0.0})
            d["transfer_ms"] += r["transfer_ms"] / 2
            d["exposed_ms"] += r.get("exposed_ms", 0.0) / 2
        # exposed part: the same step with the exchanges switched off (ghost rows keep stale values: timing only)
        halo_obj = runner.halo
        runner.halo = type("NoExchange", (), {"start": lambda self, b, n="": None, "finish": lambda self, b, n="": None})()
        ms_noex = timed(step, K, 1)
        runner.halo = halo_obj
        step(0)                                       # a valid state again
        fam = profile_families(step, 2)
        rows_sent, n_points = runner.halo_rows_per_step()
        gh = torch.tensor([plan.n_ghost, plan.n_boundary, rows_sent], device=dev)
        dist.all_reduce(gh, op=dist.ReduceOp.MAX)
        # ---- correctness against the monolithic evaluation on rank 0 (same positions)
        runner.step(poss[0])
        e_dec = runner.total_energy().clone()
        f_own = runner.forces.clone()
        sizes = [p.n_own for p in plans]
        pad = max(sizes)
        fbuf = torch.zeros(pad, 3, device=dev)
        fbuf[:plan.n_own] = f_own
        gathered = [torch.zeros(pad, 3, device=dev) for _ in range(world)] if rank == 0 else None
        dist.gather(fbuf, gathered, dst=0)
        E = int(runner.status[0])
        et = torch.tensor([E], device=dev)
        dist.all_reduce(et)
        del runner, halo_obj
        torch.cuda.empty_cache()
        corr = None
        if rank == 0:
            f_dec = torch.zeros(N, 3, device=dev)
            for p, t in zip(plans, gathered):
                f_dec[p.own.to(dev)] = t[:p.n_own]
            del gathered
            z, batch = monolithic()
            y_ref, f_ref = model(z, torch.from_numpy(S.md_step_noise(pos0, 0, seed=2)).to(dev), batch)
            d_f = float((f_dec - f_ref).abs().max())
            d_e = float((e_dec - y_ref).abs().max()) / max(1.0, float(y_ref.abs().max()))
            corr = {"reference": "rank 0 monolithic evaluation of the same positions", "max_abs_dF": d_f, "rel_dE": d_e,
                    "tolerance": {"max_abs_dF": 1e-4, "rel_dE": 1e-5}, "energy_decomposed": float(e_dec),
                    "energy_monolithic": float(y_ref)}
            assert d_f <= 1e-4 and d_e <= 1e-5, "decomposed result differs from the monolithic one: %s" % corr
        barrier()
        out.update({"ms_per_step": ms, "value": N / (ms * 1e-3), "unit": UNIT,
                    "parallelism": "%dx%dx%d bricks, %d halo points of [rows,9,%d] fp32 + positions per step, transport %s, "
                                   "boundary rows first and the exchange overlapped with the interior rows; one scalar "
                                   "all-reduce" % (plans[0].grid + (n_points, H, transport)),
                    "transport": transport, "transport_note": note, "grid": list(plans[0].grid),
                    "ghost_atoms_max": int(gh[0]), "boundary_atoms_max": int(gh[1]),
                    "halo_bytes_per_step": int(gh[2]) * (12 + n_points * 36 * H),
                    "halo_ms_per_step": float(halo_ms[0]), "halo_exposed_ms_per_step": float(halo_ms[1]),
                    "halo_per_point_rank0": per_point, "ms_per_step_without_exchange": ms_noex,
                    "ms_per_step_minus_without_exchange": ms - ms_noex, "launches_per_step": int(launches),
                    "halo_note": "halo_ms = pack + transfer time on the communication stream (CUDA events, max over ranks); "
                                 "exposed = the part of each exchange (incl. its completion barrier) that ended after the "
                                 "interior rows of the producing stage were done, plus the positions halo",
                    "edges": int(et), "correctness": corr})
        n_own, n_prof = plan.n_own, 2
        E = int(et) // world
    if rank != 0:
        return None
    if "tn_msg_bwd" in fam:
        cnt, tms = fam["tn_msg_bwd"]
        bpl = algorithmic_bytes("tn_msg_bwd", n_own, E, H, L)
        peak = hbm_peak()
        ach = bpl / (tms / cnt * 1e-3) / 1e9
        out["roofline"] = {"bound": "hbm", "kernel": "tn_msg_bwd", "achieved": ach, "peak": peak[0], "unit": "GB/s",
                           "frac": ach / peak[0], "traffic": None, "peak_source": peak[1],
                           "algorithmic_bytes_per_launch": bpl, "avg_launch_us": tms / cnt * 1e3,
                           "l2_bytes_per_launch": l2_bytes("tn_msg_bwd", E, H),
                           "note": "rank 0's launches (all rows of one stage = boundary + interior launch); CUDA events "
                                   "inside the library over %d eager steps" % n_prof}
    out["clocks"] = dict(clocks_seen) if clocks_seen else None     # nvidia-smi samples of rank 0's GPU during the timed steps
    out["kernel_families"] = {k: {"launches_per_step": c / n_prof, "us_per_step": m_ / n_prof * 1e3} for k, (c, m_) in
                              sorted(fam.items(), key=lambda kv: -kv[1][1])[:8]}
    return out


def hbm_peak():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        with open(peaks_path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------- main arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="chignolin", choices=list(HPARAMS))
    ap.add_argument("--filter", default="table", choices=["table", "exact"])
    ap.add_argument("--molecules", type=int, default=8192)
    ap.add_argument("--gemm", "--et-gemm", dest="et_gemm", default="tf32x3", choices=["fp32", "tf32x3"],
                    help="plain node GEMMs (TensorNet MLPs, ET projections): tcgen05 three-product tf32 with fp32 chunk sums "
                         "(default) or the fp32 FFMA kernel; the TensorNet channel mixes always use the tensor cores")
    ap.add_argument("--water-side", type=int, default=32)
    ap.add_argument("--water-order", default="raster", choices=["raster", "morton", "shuffle"],
                    help="index order of the synthetic water molecules (raster = lattice order)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true",
                    help="skip the strong-scaling arm (the 1 M-atom periodic water box evaluated on all N GPUs together)")
    ap.add_argument("--strong-side", type=int, default=70, help="water molecules per box edge of the strong-scaling arm (70 -> 1 029 000 atoms)")
    ap.add_argument("--strong-steps", type=int, default=5, help="timed steps of the strong-scaling arm (<= --steps)")
    ap.add_argument("--transport", default="peer", choices=["peer", "nccl"],
                    help="halo transport of the decomposed step: fused pack+push kernel over NVLink peer memory, or NCCL send/recv")
    ap.add_argument("--grid", default="", help="brick grid of the decomposed step, e.g. 2x2x2 (default: smallest ghost volume)")
    ap.add_argument("--no-flush", action="store_true",
                    help="experiment only: do not flush L2 between the timed steps of `value` (the line says so; not a bench number)")
    ap.add_argument("--table-intervals", type=int, default=0,
                    help="intervals of the filter tables (0 = the model's default, max(1024, 32 * num_rbf)); A/B runs only")
    ap.add_argument("--no-table-warp", action="store_true",
                    help="A/B runs only: filter tables on a grid uniform in d (round-1 layout; use with --table-intervals 4096)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node %d" % args.gpus

    from torchmdnet import _native
    from torchmdnet.calculators import External
    from torchmdnet.models.model import create_model
    from torchmdnet.models.utils import OptimizedDistance
    _native.load()

    hp = HPARAMS[args.workload]
    seed_everything(1234)
    model = create_model(hp).to(dev).eval()
    if args.workload == "water" and world > 1:      # one box on all GPUs: the strong-scaling arm IS the line
        del model
        args.strong_side, args.strong_steps = args.water_side, args.steps
        ss = strong_scaling(args, dev, rank, world, local)
        if rank == 0:
            emit({"metric": METRIC, "value": ss["value"], "unit": UNIT, "n_gpus": world, "steps": ss["steps"],
                  "warmup": ss["warmup"], "ms_per_step": ss["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                  "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                  "config": {"workload": WORKLOAD_DESC["water"], "atoms_total": ss["atoms_total"],
                             "parallelism": ss["parallelism"], "l2": ss["l2"]},
                  "e2e": None, "gpu_launches": ss["launches_per_step"] * ss["steps"] * world,
                  "launches_per_step": ss["launches_per_step"], "roofline": ss.get("roofline"), "cpu_baseline": None,
                  "strong_scaling": ss})
        return
    rm = model.representation_model
    rm.gemm_mode = args.et_gemm
    if hp["model"] == "tensornet":
        rm.filter_mode = args.filter
        if args.table_intervals > 0:
            rm.table_intervals = args.table_intervals
        rm.table_warp = not args.no_table_warp
    else:
        args.filter = "table"
    z_np, pos0, batch_np, box = make_inputs(args.workload, rank, args)
    if box is not None:
        rm.distance = OptimizedDistance(hp["cutoff_lower"], hp["cutoff_upper"], max_num_pairs=-hp["max_num_neighbors"],
                                        return_vecs=True, loop=True, check_errors=False, resize_to_fit=False,
                                        box=torch.from_numpy(box), strategy="cell")
    N = z_np.shape[0]
    n_mol = int(batch_np.max()) + 1
    K, W = args.steps, args.warmup
    n_pos = min(K + W, 64)                                   # distinct position sets, cycled
    poss_np = step_positions(pos0, n_pos, 2 + rank)
    z, batch = torch.from_numpy(z_np).to(dev), torch.from_numpy(batch_np).to(dev)
    poss = torch.stack([torch.from_numpy(p) for p in poss_np]).to(dev)       # resident in HBM
    static_pos = poss[0].clone()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)           # > 126 MB L2

    def eager_step(i):
        static_pos.copy_(poss[i % n_pos])
        return model(z, static_pos, batch)

    # launches per step (the library counts every kernel it launches)
    eager_step(0)
    torch.cuda.synchronize()
    c0 = torch.ops.torchmdnet_b200.launch_count()
    y, f = eager_step(1)
    torch.cuda.synchronize()
    launches_per_step = torch.ops.torchmdnet_b200.launch_count() - c0
    status = rm.energy_forces(z, static_pos, batch, n_mol, model.output_model, *model._mean_std())[2].cpu()
    E = int(status[0])
    assert int(status[1]) == 0, "pair capacity exceeded: %d edges" % E

    # whole step as one CUDA graph
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(3):
            model(z, static_pos, batch)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        gy, gf = model(z, static_pos, batch)

    def graph_step(i):
        static_pos.copy_(poss[i % n_pos])
        graph.replay()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for i in range(W):
        graph_step(i)
    sampler = ClockSampler(local) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    t_begin = time.perf_counter()
    for i in range(K):
        if not args.no_flush:
            flush.zero_()                               # L2 flush, outside the timed span
        ev[i][0].record()
        graph_step(W + i)
        ev[i][1].record()
    barrier()
    t_end = time.perf_counter()
    clocks = sampler.stop(t_begin, t_end) if sampler else None
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    total_ms = float(t)
    ms_per_step = total_ms / K
    value = world * N / (ms_per_step * 1e-3)

    # ---- e2e: public API, host buffers in and out, copies inside the timed span
    tmp = tempfile.mkdtemp()
    ckpt = os.path.join(tmp, "bench.ckpt")
    torch.save({"hyper_parameters": dict(hp), "state_dict": {"model." + k: v.cpu() for k, v in
                                                              model.state_dict().items()}}, ckpt)
    equal_sizes = n_mol == 1 or (N % n_mol == 0 and bool((np.bincount(batch_np) == N // n_mol).all()))
    emb = torch.from_numpy(z_np).view(n_mol, N // n_mol) if equal_sizes else torch.from_numpy(z_np).view(1, N)
    calc = External(ckpt, emb, dev)
    calc.model.representation_model.gemm_mode = args.et_gemm
    if hp["model"] == "tensornet":
        calc.model.representation_model.filter_mode = args.filter
        if args.table_intervals > 0:
            calc.model.representation_model.table_intervals = args.table_intervals
        calc.model.representation_model.table_warp = not args.no_table_warp
    if box is not None:
        calc.model.representation_model.distance = rm.distance
    if emb.shape[0] != n_mol:
        calc.batch = batch
        calc._n_mol = n_mol
    host_pos = [torch.from_numpy(p).pin_memory() for p in poss_np]
    h_e = torch.empty(n_mol, 1).pin_memory()
    h_f = torch.empty(N, 3).pin_memory()
    Ke = K

    def e2e_step(i):
        e, frc = calc.calculate(host_pos[i % n_pos], None)
        h_e.copy_(e, non_blocking=True)
        h_f.copy_(frc.view(N, 3), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for i in range(W):
        e2e_step(i)
    barrier()
    e2e_t = 0.0
    for i in range(Ke):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e2e_step(W + i)
        e2e_t += time.perf_counter() - t0
    barrier()
    t = torch.tensor([e2e_t], device=dev, dtype=torch.float64)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    e2e_value = world * N / (float(t) / Ke)
    # the e2e result must be the product's result
    assert torch.allclose(h_f, calc.calculate(host_pos[(W + Ke - 1) % n_pos], None)[1].view(N, 3).cpu())

    # ---- strong-scaling arm: the 1 M-atom periodic box on all N GPUs together (every rank takes part)
    strong = None
    if args.workload == "chignolin" and not args.no_strong:
        strong = strong_scaling(args, dev, rank, world, local)
        torch.cuda.set_device(local)
    if rank != 0:
        return
    # ---- per-kernel-family device times (separate eager pass, events inside the library) -> roofline
    fam = profile_families(eager_step, min(K, 20))
    n_prof = min(K, 20)
    step_ms_prof = sum(ms for _, ms in fam.values()) / n_prof
    top = max(fam, key=lambda k: fam[k][1])
    # the roofline is quoted for the neighbour-aggregation kernels north_star names; the table shows all
    agg = [k for k in ("tn_msg_bwd", "tn_msg_fwd", "et_edge_bwd", "et_edge_fwd") if k in fam]
    dom = max(agg, key=lambda k: fam[k][1]) if agg else top
    cnt, ms = fam[dom]
    H, L = hp["embedding_dimension"], hp["num_layers"]
    bytes_per_launch = algorithmic_bytes(dom, N, E, H, L)
    achieved = bytes_per_launch / (ms / cnt * 1e-3) / 1e9
    peak, peak_src = hbm_peak()
    cap = ncu_capture("%s:%d" % (args.workload, N), dom)
    l2b = l2_bytes(dom, E, H)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": cap["dram_bytes"] if cap else None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_us": ms / cnt * 1e3,
                "share_of_step": ms / n_prof / step_ms_prof,
                "l2_bytes_per_launch": l2b, "l2_achieved_gbs": l2b / (ms / cnt * 1e-3) / 1e9,
                "ncu": cap,
                "note": "time of the kernel = CUDA events inside the library over %d eager steps; the gathers of this "
                        "kernel are served by L2 (DESIGN.md 4.3), l2_* is the traffic that bounds it" % n_prof}
    families = {k: {"launches_per_step": c / n_prof, "us_per_step": m / n_prof * 1e3} for k, (c, m) in
                sorted(fam.items(), key=lambda kv: -kv[1][1])}

    # which GEMM kernels this launch size selected (families observed by the library's own timers)
    gemm_names = {"node_gemm_tc": "tcgen05 tf32x3 (three-product split, fp32 chunk sums)",
                  "node_gemm_small": "latency-oriented fp32 FFMA (whole-K panels via cp.async, small systems)",
                  "node_gemm": "tiled fp32 FFMA"}
    gemm_desc = "; ".join("%s: %s x%d/step" % (k, v, round(fam[k][0] / n_prof)) for k, v in gemm_names.items() if k in fam)

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        v, cores, sample, _, _ = cpu_reference_steps(args.workload, hp, args, 3, 1)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[args.workload], "atoms_per_gpu": N, "molecules_per_gpu": n_mol,
                       "edges_per_gpu": E, "filter_mode": args.filter,
                       "gemm": gemm_desc,
                       "neighbor_strategy": rm.distance.strategy,
                       "l2": "NOT flushed (experiment, not a bench number)" if args.no_flush else
                             "flushed (512 MiB write) before every timed step",
                       "step": "one CUDA graph replay: neighbour list + CSR + forward + analytic backward",
                       "parallelism": "independent replicas/shards per GPU, no collective in the step"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": N * 12,
                    "d2h_bytes_per_step": N * 12 + n_mol * 4, "api": "torchmdnet.calculators.External.calculate"},
            "gpu_launches": int(launches_per_step) * K, "launches_per_step": int(launches_per_step),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "kernel_families": families,
            "energy_checksum": float(gy.double().sum()), "strong_scaling": strong}
    emit(line)


if __name__ == "__main__":
    main()
