#!/usr/bin/env python
"""Neighbour-list sweep (BASELINE.json configs[4]; methodology of the reference's benchmarks/neighbors.py:59-120):
OptimizedDistance (pair list in the reference layout) and the direct CSR builder, N = 1k ... 4M atoms at density
0.1 A^-3, cutoff 5 / 7.5 / 10 A, periodic and open boundaries, brute vs cell; check_errors=False, resize_to_fit=False,
max_num_pairs = 1.3 x expected; CUDA events, 3 warm-ups, median of 10; L2 flushed between repetitions.
Prints one JSON line per case: ms, pairs found, achieved GB/s on the algorithmic bytes (20 N + 24 P for the pair list,
16 N + 20 E for the CSR build) and the fraction of the measured HBM peak."""
import argparse
import json
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torchmd-net_b200"))
import synth_inputs as S  # noqa: E402


def timed(fn, flush, reps=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def reference_sweep(args):
    """The same cases through torch.ops.torchmdnet_neighbors.get_neighbor_pairs of the reference's CUDA op.  Runs in a
    process that never loads the product's shim (both register the same op name).  Bench infrastructure only."""
    lib = os.path.join(ROOT, "oracle", "_ref", "torchmdnet_neighbors_ref_cuda.so")
    if not os.path.exists(lib):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/torchmdnet_neighbors_ref_cuda.so not built"}))
        return
    torch.ops.load_library(lib)
    op = torch.ops.torchmdnet_neighbors.get_neighbor_pairs
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for n in [int(x) for x in args.sizes.split(",")]:
        pos_np, box_np = S.uniform_gas(n, seed=5)
        pos = torch.from_numpy(pos_np).cuda()
        batch = torch.zeros(n, dtype=torch.long, device="cuda")
        for rc in [float(x) for x in args.cutoffs.split(",")]:
            if rc == 10.0 and n > 1048576:
                continue
            expected = n * 0.1 * 4.0 / 3.0 * math.pi * rc ** 3
            for periodic in (True, False):
                if periodic and box_np[0, 0] < 2 * rc:
                    continue
                for strategy in ("brute", "shared", "cell"):
                    if strategy != "cell" and n > args.brute_max:
                        continue
                    if strategy == "brute" and n >= 32768:
                        continue                             # neighbors_cuda_brute.cuh:75 refuses N >= 32768
                    cap = int(min(expected * 1.3 + 1000, 2 ** 31 - 2))
                    box = torch.from_numpy(box_np)           # CPU tensor, as the reference requires
                    try:
                        fn = lambda: op(strategy, pos, batch, box, periodic, 0.0, rc, cap, False, True)  # noqa: E731
                        out = fn()
                        found = int(out[3])
                        ms = timed(fn, flush)
                        print(json.dumps({"impl": "reference", "op": "get_neighbor_pairs", "n": n, "cutoff": rc,
                                          "periodic": periodic, "strategy": strategy, "ms": ms, "pairs": found}), flush=True)
                        del out
                    except Exception as ex:      # noqa: BLE001
                        print(json.dumps({"impl": "reference", "n": n, "cutoff": rc, "periodic": periodic,
                                          "strategy": strategy, "error": str(ex)[:200]}), flush=True)
                    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1000,4000,16000,32767,65536,262144,1048576,4194304")
    ap.add_argument("--cutoffs", default="5.0,7.5,10.0")
    ap.add_argument("--brute-max", type=int, default=65536)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"],
                    help="reference = the reference's own CUDA kernels (neighbors_cuda_{brute,shared,cell}.cuh) recompiled "
                         "for sm_100 by BUILD_REF_CUDA=1 oracle/build_ref.sh: the GPU-side bar (SURVEY 2.2)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_sweep(args)
    from torchmdnet import _native
    from torchmdnet.models.utils import OptimizedDistance
    _native.load()
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for n in [int(x) for x in args.sizes.split(",")]:
        pos_np, box_np = S.uniform_gas(n, seed=5)
        pos = torch.from_numpy(pos_np).cuda()
        batch = torch.zeros(n, dtype=torch.long, device="cuda")
        for rc in [float(x) for x in args.cutoffs.split(",")]:
            if rc == 10.0 and n > 1048576:
                continue                                     # 419 neighbours/atom x 4M atoms does not fit int32 pairs
            expected = n * 0.1 * 4.0 / 3.0 * math.pi * rc ** 3
            for periodic in (True, False):
                if periodic and box_np[0, 0] < 2 * rc:
                    continue
                for strategy in ("brute", "cell"):
                    if strategy == "brute" and n > args.brute_max:
                        continue
                    cap = int(min(expected * 1.3 + 1000, 2 ** 31 - 2))
                    box = torch.from_numpy(box_np) if (periodic or strategy == "cell") else None
                    od = OptimizedDistance(0.0, rc, max_num_pairs=cap, return_vecs=True, loop=False, strategy=strategy,
                                           include_transpose=True, resize_to_fit=False, check_errors=False,
                                           box=box if periodic else None, long_edge_index=False)
                    if strategy == "cell" and not periodic:
                        od.box = torch.from_numpy(box_np)    # binning box of the open-boundary cell list
                    try:
                        nb = od(pos, batch)
                        found = int((nb[0][0] >= 0).sum())
                        ms = timed(lambda: od(pos, batch), flush)
                        gb = (20.0 * n + 24.0 * found) / 1e9
                        line = {"op": "OptimizedDistance", "n": n, "cutoff": rc, "periodic": periodic,
                                "strategy": strategy, "ms": ms, "pairs": found, "gbs": gb / (ms * 1e-3),
                                "frac_hbm": gb / (ms * 1e-3) / peak}
                        print(json.dumps(line), flush=True)
                        del nb
                        ecap = int(min(expected * 1.3 + n + 1000, 2 ** 31 - 2))
                        tb = torch.empty(0, 0) if box is None else box
                        fn = lambda: torch.ops.torchmdnet_b200.csr_build(strategy, pos, batch, tb if (periodic or strategy == "cell") else torch.empty(0, 0), periodic, 0.0, rc, ecap)  # noqa: E731
                        e = int(fn()[3])
                        ms = timed(fn, flush)
                        gb = (16.0 * n + 20.0 * e) / 1e9
                        print(json.dumps({"op": "csr_build", "n": n, "cutoff": rc, "periodic": periodic,
                                          "strategy": strategy, "ms": ms, "pairs": e, "gbs": gb / (ms * 1e-3),
                                          "frac_hbm": gb / (ms * 1e-3) / peak}), flush=True)
                    except Exception as ex:      # noqa: BLE001
                        print(json.dumps({"n": n, "cutoff": rc, "periodic": periodic, "strategy": strategy,
                                          "error": str(ex)[:200]}), flush=True)
                    torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
