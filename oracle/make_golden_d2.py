"""TEST INFRASTRUCTURE ONLY -- tests/golden/small_tensornet_d2.npz: the UNMODIFIED reference TensorNet with the D2
dispersion prior (torchmdnet/priors/d2.py) on the small synthetic batch of the other small fixtures, plus the prior evaluated
alone (energy from ``D2.post_reduce``, forces from autograd over it) in fp32 and fp64.
``python oracle/make_golden_d2.py`` (build container only)."""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_import as R  # noqa: E402
from oracle.make_golden import GOLD, hp_of, load_yaml_args  # noqa: E402
import synth_inputs as S  # noqa: E402


def main():
    R.import_reference()
    from torchmdnet.models.model import create_model
    from torchmdnet.priors.d2 import D2
    torch.set_float32_matmul_precision("highest")
    z, pos, batch = S.qm9_like_batch(n_mol=5, seed=11)
    # Angstrom / eV
    dargs = dict(cutoff_distance=6.0, max_num_neighbors=40, atomic_number=list(range(55)), distance_scale=1e-10,
                 energy_scale=1.602176634e-19)
    R.seed_everything(7)
    a = load_yaml_args("TensorNet-rMD17", embedding_dimension=32, num_rbf=16, num_layers=2)
    m = create_model(a, prior_model=[D2(**dargs)], mean=torch.tensor(0.37), std=torch.tensor(1.9))
    sd = {k: v.numpy() for k, v in m.state_dict().items()}
    tz, tb = torch.from_numpy(z), torch.from_numpy(batch)
    y, f = m(tz, torch.from_numpy(pos).clone(), tb)
    m.prior_model = None
    y0, f0 = m(tz, torch.from_numpy(pos).clone(), tb)
    n_mol = int(batch.max()) + 1

    def alone(dtype):
        prior = D2(dtype=dtype, **dargs)
        p = torch.from_numpy(pos).to(dtype).clone().requires_grad_(True)
        e = prior.post_reduce(torch.zeros(n_mol, 1, dtype=dtype), tz, p, tb, {})
        (g,) = torch.autograd.grad(e.sum(), p)
        return e.detach().numpy(), (-g).numpy()

    e32, f32 = alone(torch.float32)
    e64, f64 = alone(torch.float64)
    hp = hp_of(a)
    hp["prior_model"] = [{"D2": dargs}]
    hp["prior_args"] = None
    np.savez_compressed(os.path.join(GOLD, "small_tensornet_d2.npz"), hparams=json.dumps(hp), z=z, pos=pos, batch=batch,
                        y=y.detach().numpy(), forces=f.detach().numpy(), y_noprior=y0.detach().numpy(),
                        forces_noprior=f0.detach().numpy(), y_prior=e32, forces_prior=f32, y_prior_f64=e64,
                        forces_prior_f64=f64, **{"sd:" + k: v for k, v in sd.items()})
    print("D2: y", y.detach().flatten().numpy(), "prior alone", e64.flatten(), "max |F_prior|", float(np.abs(f64).max()),
          "fp32 vs fp64 dE", float(np.abs(e32 - e64).max()), "dF", float(np.abs(f32 - f64).max()))


if __name__ == "__main__":
    main()
