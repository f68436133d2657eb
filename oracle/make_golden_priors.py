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
    prior.atomref.weight.data.copy_(torch.linspace(-3.0, 5.0, a["max_z"]).view(-1, 1))
    m = create_model(a, prior_model=prior, mean=torch.tensor(0.37), std=torch.tensor(1.9))
    y, f = run_ref(m, z, pos, batch)
    sd = {k: v.numpy() for k, v in m.state_dict().items()}
    hp = hp_of(a)
    hp["prior_model"] = "Atomref"
    hp["prior_args"] = [prior.get_init_args()]
    np.savez_compressed(os.path.join(GOLD, "small_tensornet_atomref.npz"), hparams=json.dumps(hp), z=z, pos=pos,
                        batch=batch, y=y, forces=f, **{"sd:" + k: v for k, v in sd.items()})
    print("atomref", y.flatten(), [k for k in sd if "prior" in k])


if __name__ == "__main__":
    main()
