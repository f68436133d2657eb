"""TEST INFRASTRUCTURE ONLY -- tests/golden/small_tensornet_atomref.npz: the UNMODIFIED reference TensorNet with an
Atomref prior (torchmdnet/priors/atomref.py) holding non-trivial reference energies, non-trivial mean / std, on the small
synthetic batch of the other small fixtures.  ``python oracle/make_golden_priors.py`` (build container only)."""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_import as R  # noqa: E402
from oracle.make_golden import GOLD, hp_of, load_yaml_args, run_ref  # noqa: E402
import synth_inputs as S  # noqa: E402


def main():
    R.import_reference()
    from torchmdnet.models.model import create_model
    from torchmdnet.priors.atomref import Atomref
    torch.set_float32_matmul_precision("highest")
    z, pos, batch = S.qm9_like_batch(n_mol=5, seed=11)
    R.seed_everything(7)
    a = load_yaml_args("TensorNet-rMD17", embedding_dimension=32, num_rbf=16, num_layers=2)
    prior = Atomref(max_z=a["max_z"])
    m = create_model(a, prior_model=prior, mean=torch.tensor(0.37), std=torch.tensor(1.9))
    # after create_model: TorchMD_Net.__init__ resets the prior to its initial (zero) reference energies
    m.prior_model[0].atomref.weight.data.copy_(torch.linspace(-3.0, 5.0, a["max_z"]).view(-1, 1))
    y, f = run_ref(m, z, pos, batch)
    sd = {k: v.numpy() for k, v in m.state_dict().items()}
    hp = hp_of(a)
    hp["prior_model"] = "Atomref"
    hp["prior_args"] = [prior.get_init_args()]
    np.savez_compressed(os.path.join(GOLD, "small_tensornet_atomref.npz"), hparams=json.dumps(hp), z=z, pos=pos,
                        batch=batch, y=y, forces=f, **{"sd:" + k: v for k, v in sd.items()})
    print("atomref", y.flatten(), [k for k in sd if "prior" in k])

    # ---- pairwise priors: ZBL (priors/zbl.py) + Coulomb (priors/coulomb.py) behind the same small TensorNet; forces from
    # the reference's autograd pass include the priors.  Units: Angstrom / eV (distance_scale 1e-10 m, energy_scale 1 eV in J).
    from torchmdnet.priors.coulomb import Coulomb
    from torchmdnet.priors.zbl import ZBL
    R.seed_everything(7)
    a = load_yaml_args("TensorNet-rMD17", embedding_dimension=32, num_rbf=16, num_layers=2)
    zargs = dict(cutoff_distance=4.0, max_num_neighbors=40, atomic_number=list(range(a["max_z"])), distance_scale=1e-10,
                 energy_scale=1.602176634e-19)
    cargs = dict(alpha=1.8, max_num_neighbors=40, distance_scale=1e-10, energy_scale=1.602176634e-19)
    pri = [ZBL(**zargs), Coulomb(**cargs)]
    m = create_model(a, prior_model=pri, mean=torch.tensor(0.37), std=torch.tensor(1.9))
    rng = np.random.RandomState(3)
    q = (rng.rand(z.shape[0]).astype(np.float32) - 0.5)
    tz, tp, tb, tq = torch.from_numpy(z), torch.from_numpy(pos).clone(), torch.from_numpy(batch), torch.from_numpy(q)
    y, f = m(tz, tp, tb, extra_args={"partial_charges": tq})
    m.prior_model = None
    y0, f0 = m(tz, torch.from_numpy(pos).clone(), tb)
    sd = {k: v.numpy() for k, v in create_model(a, prior_model=pri, mean=torch.tensor(0.37), std=torch.tensor(1.9)).state_dict().items()}
    R.seed_everything(7)
    m2 = create_model(a, prior_model=[ZBL(**zargs), Coulomb(**cargs)], mean=torch.tensor(0.37), std=torch.tensor(1.9))
    sd = {k: v.numpy() for k, v in m2.state_dict().items()}
    hp = hp_of(a)
    hp["prior_model"] = [{"ZBL": zargs}, {"Coulomb": cargs}]
    hp["prior_args"] = None
    np.savez_compressed(os.path.join(GOLD, "small_tensornet_pairpriors.npz"), hparams=json.dumps(hp), z=z, pos=pos,
                        batch=batch, partial_charges=q, y=y.detach().numpy(), forces=f.detach().numpy(),
                        y_noprior=y0.detach().numpy(), forces_noprior=f0.detach().numpy(),
                        **{"sd:" + k: v for k, v in sd.items()})
    print("pair priors: y", y.detach().flatten().numpy(), "prior part", (y - y0).detach().flatten().numpy(),
          "max |dF| from priors", float((f - f0).abs().max()))


if __name__ == "__main__":
    main()
