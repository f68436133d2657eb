"""TEST INFRASTRUCTURE ONLY -- graph-network (TorchMD_GN) and transformer (TorchMD_T) fixtures written by the UNMODIFIED
reference (the T files carry `_t` / `T` in place of `_gn` / `GN`):
  expected_pkl_gn.json   the graph-network/Scalar row of the reference's own golden file tests/expected.pkl with the
                         seeded inputs of its recipe (tests/test_model.py:137-188) and state_dict checksums;
  small_gn_ne.npz / small_gn_plain.npz   small models WITH their weights (with / without neighbour embedding);
  model_GN.npz           seeded create_model at H = 128, 6 layers, 50 RBFs on a synthetic batch (+ checksums).
``python oracle/make_golden_gn.py`` (build container only)."""
import json
import os
import pickle
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_import as R  # noqa: E402
from oracle.make_golden import GOLD, checksums, hp_of, load_yaml_args, run_ref  # noqa: E402
import synth_inputs as S  # noqa: E402

GN = dict(model="graph-network", aggr="add", neighbor_embedding=True, output_model="Scalar")


def main():
    R.import_reference()
    from torchmdnet.models.model import create_model
    torch.set_float32_matmul_precision("highest")
    with open(os.path.join(R.REF_ROOT, "tests", "expected.pkl"), "rb") as f:
        exp = pickle.load(f)
    zs = torch.tensor([1, 6, 7, 8, 9], dtype=torch.long)
    R.seed_everything(1234)
    a = load_yaml_args("ET-QM9", seed=1234, **GN)
    m = create_model(a)
    z = zs[torch.randint(0, len(zs), (5,))]
    pos = torch.randn(5, 3)
    batch = torch.zeros(5, dtype=torch.long)
    batch[2:] = 1
    y, f = m(z, pos.clone(), batch)
    e = exp["graph-network"]["Scalar"]
    assert (y - e["pred"]).abs().max() < 1e-5 and (f - e["deriv"]).abs().max() < 1e-5, (y, e["pred"])
    hp = dict(hp_of(a), aggr="add")
    out = dict(hparams=hp, yaml="ET-QM9 + model=graph-network", seed=1234, z=z.tolist(), pos=pos.tolist(),
               batch=batch.tolist(), expected_pred=e["pred"].tolist(), expected_deriv=e["deriv"].tolist(),
               rerun_pred=y.tolist(), rerun_deriv=f.tolist(), checksums=checksums(m.state_dict()))
    with open(os.path.join(GOLD, "expected_pkl_gn.json"), "w") as fh:
        json.dump({"graph-network": out}, fh, indent=1)
    print("expected.pkl GN row reproduced:", y.flatten().tolist())

    zs_, ps_, bs_ = S.qm9_like_batch(n_mol=5, seed=11)
    for name, kw in (("gn_ne", dict(neighbor_embedding=True)), ("gn_plain", dict(neighbor_embedding=False, activation="tanh"))):
        R.seed_everything(7)
        a = load_yaml_args("ET-MD17", embedding_dimension=32, num_rbf=16, num_layers=2, **dict(GN, **kw))
        m = create_model(a, mean=torch.tensor(0.37), std=torch.tensor(1.9))
        y, f = run_ref(m, zs_, ps_, bs_)
        sd = {k: v.numpy() for k, v in m.state_dict().items()}
        np.savez_compressed(os.path.join(GOLD, "small_%s.npz" % name), hparams=json.dumps(dict(hp_of(a), aggr="add")),
                            z=zs_, pos=ps_, batch=bs_, y=y, forces=f, **{"sd:" + k: v for k, v in sd.items()})
        print(name, y.flatten()[:3], np.abs(f).max())

    # ---- transformer (TorchMD_T): the same three kinds of fixture
    T = dict(model="transformer", output_model="Scalar")
    R.seed_everything(1234)
    a = load_yaml_args("ET-QM9", seed=1234, **T)
    m = create_model(a)
    z = zs[torch.randint(0, len(zs), (5,))]
    pos = torch.randn(5, 3)
    y, f = m(z, pos.clone(), batch)
    e = exp["transformer"]["Scalar"]
    assert (y - e["pred"]).abs().max() < 1e-5 and (f - e["deriv"]).abs().max() < 1e-5, (y, e["pred"])
    out = dict(hparams=hp_of(a), yaml="ET-QM9 + model=transformer", seed=1234, z=z.tolist(), pos=pos.tolist(),
               batch=batch.tolist(), expected_pred=e["pred"].tolist(), expected_deriv=e["deriv"].tolist(),
               rerun_pred=y.tolist(), rerun_deriv=f.tolist(), checksums=checksums(m.state_dict()))
    with open(os.path.join(GOLD, "expected_pkl_t.json"), "w") as fh:
        json.dump({"transformer": out}, fh, indent=1)
    print("expected.pkl T row reproduced:", y.flatten().tolist())
    for name, kw in (("t_both", dict()), ("t_values_nonbr", dict(distance_influence="values", neighbor_embedding=False,
                                                                 attn_activation="tanh"))):
        R.seed_everything(7)
        a = load_yaml_args("ET-MD17", embedding_dimension=32, num_rbf=16, num_layers=2, num_heads=4, **dict(T, **kw))
        m = create_model(a, mean=torch.tensor(0.37), std=torch.tensor(1.9))
        y, f = run_ref(m, zs_, ps_, bs_)
        sd = {k: v.numpy() for k, v in m.state_dict().items()}
        np.savez_compressed(os.path.join(GOLD, "small_%s.npz" % name), hparams=json.dumps(hp_of(a)),
                            z=zs_, pos=ps_, batch=bs_, y=y, forces=f, **{"sd:" + k: v for k, v in sd.items()})
        print(name, y.flatten()[:3], np.abs(f).max())
    z3, p3, b3 = S.md17_like_batch(n_mol=3, seed=3)
    R.seed_everything(1234)
    a = load_yaml_args("ET-MD17", **T)
    m = create_model(a)
    y, f = run_ref(m, z3, p3, b3)
    np.savez_compressed(os.path.join(GOLD, "model_T.npz"), hparams=json.dumps(hp_of(a)), seed=1234,
                        z=z3, pos=p3, batch=b3, y=y, forces=f, checksums=json.dumps(checksums(m.state_dict())))
    print("model_T", y.flatten(), np.abs(f).max())

    R.seed_everything(1234)
    a = load_yaml_args("ET-MD17", embedding_dimension=128, num_layers=6, num_rbf=50, **GN)
    m = create_model(a)
    y, f = run_ref(m, z3, p3, b3)
    np.savez_compressed(os.path.join(GOLD, "model_GN.npz"), hparams=json.dumps(dict(hp_of(a), aggr="add")), seed=1234,
                        z=z3, pos=p3, batch=b3, y=y, forces=f, checksums=json.dumps(checksums(m.state_dict())))
    print("model_GN", y.flatten(), np.abs(f).max())


if __name__ == "__main__":
    main()
