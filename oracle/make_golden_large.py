"""TEST INFRASTRUCTURE ONLY -- tests/golden/water_<N>_ani1x.npz: the UNMODIFIED reference TensorNet with the full
TensorNet-ANI1X hparams (H = 128, K = 32, 2 layers) on periodic water boxes, OptimizedDistance(box=...) swapped in
(SURVEY 8c), evaluated on the CPU in fp32 (the parity target) and, for the smaller box, in fp64 as well (what the fp32
numbers are rounding errors OF: it bounds the energy tolerance the tests may use).

    python oracle/make_golden_large.py        # build container only; 42 GB of RAM, a few minutes

Sizes: n_side 11 -> 3993 atoms (L = 33.9 A) in fp32; n_side 8 -> 1536 atoms in fp32 and fp64.  The reference keeps about
a hundred [E,H,3,3] temporaries alive under autograd (10.4 MB per atom in fp32, measured: 32 GB peak RSS for the fp64 pass
at 1536 atoms), so the 20 577-atom box SURVEY 8c names (214 GB) and even 8232 atoms (86 GB) do not fit the 62 GB of this
container; 3993 atoms is what does.  It is large enough for the tcgen05 channel mixes and MLP GEMMs on full tiles, the
cell list on a 7^3 grid and fp32 sums over thousands of atoms; beyond it the chain continues with in-tree identities
(persistent == per-row kernels bit for bit at 8232 atoms, decomposed == monolithic at 10^6 atoms inside bench.py)."""
import json
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_import as R  # noqa: E402
from oracle.make_golden import GOLD, hp_of, load_yaml_args, checksums, t  # noqa: E402
import synth_inputs as S  # noqa: E402


def run(n_side, precisions):
    from torchmdnet.models.model import create_model
    from torchmdnet.models.utils import OptimizedDistance
    z, pos, batch, box = S.water_box(n_side, seed=4)
    out, sums = {}, None
    for prec in precisions:
        dt = torch.float32 if prec == 32 else torch.float64
        R.seed_everything(1234)
        m = create_model(load_yaml_args("TensorNet-ANI1X", precision=32))
        if prec == 32:
            sums = checksums(m.state_dict())
        else:               # the same weights, widened
            sd = m.state_dict()
            m = create_model(load_yaml_args("TensorNet-ANI1X", precision=64))
            m.load_state_dict({k: v.to(dt) if v.is_floating_point() else v for k, v in sd.items()})
        a = load_yaml_args("TensorNet-ANI1X")
        m.representation_model.distance = OptimizedDistance(
            a["cutoff_lower"], a["cutoff_upper"], max_num_pairs=-a["max_num_neighbors"], return_vecs=True, loop=True,
            check_errors=False, resize_to_fit=True, box=t(box).to(dt), long_edge_index=True)
        t0 = time.time()
        y, f = m(t(z), t(pos).to(dt).clone(), t(batch))
        out[prec] = (y.detach().double().numpy().copy(), f.detach().double().numpy().copy())
        print("n_side %d precision %d: E = %.10f  max|F| = %.4f  (%.0f s)" %
              (n_side, prec, float(y.detach()), float(f.detach().abs().max()), time.time() - t0), flush=True)
        del m, y, f
    extra = {}
    if 64 in out:
        (y32, f32), (y64, f64) = out[32], out[64]
        print("fp32 vs fp64 of the reference itself: |dE| = %.3e (%.2e per atom), max|dF| = %.3e" %
              (abs(float(y32 - y64)), abs(float(y32 - y64)) / z.shape[0], np.abs(f32 - f64).max()), flush=True)
        extra = dict(y64=y64, forces64=f64.astype(np.float32))
    y32, f32 = out[32]
    np.savez_compressed(os.path.join(GOLD, "water_%d_ani1x.npz" % z.shape[0]), hparams=json.dumps(hp_of(a)), seed=1234,
                        n_side=n_side, z=z, pos=pos, batch=batch, box=box, y=y32.astype(np.float32),
                        forces=f32.astype(np.float32), checksums=json.dumps(sums), **extra)


def main():
    R.import_reference()
    torch.set_float32_matmul_precision("highest")
    torch.set_num_threads(os.cpu_count())
    run(8, (32, 64))
    run(11, (32,))


if __name__ == "__main__":
    main()
