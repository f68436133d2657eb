"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/* by running the UNMODIFIED reference here.

Run in the build container only:  ``python oracle/make_golden.py``  (needs /root/reference and
oracle/_ref/torchmdnet_neighbors_ref.so from oracle/build_ref.sh).  The fixtures it writes are
committed; nothing at test/bench time on the GPU box reads /root/reference.

Fixtures:
  expected_pkl.json        the ET/Scalar and TensorNet/Scalar rows of the reference's own golden file
                           tests/expected.pkl (numbers only) + the seeded inputs of its recipe
                           (tests/test_model.py:137-188, tests/utils.py:7-36) + a re-run of the reference.
  model_<name>.npz         reference energies/forces for seeded create_model(<yaml>) on synthetic inputs,
                           with per-tensor state_dict checksums (pins seed-for-seed init of the product).
  small_<name>.npz         small-hparam models WITH their full state_dict (weights travel with the test).
  neighbors_<case>.npz     outputs of the reference's compiled CPU op.
"""
import json
import os
import pickle
import sys

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_import as R  # noqa: E402
import synth_inputs as S  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
EX = os.path.join(R.REF_ROOT, "examples")


def load_yaml_args(name, **kw):
    with open(os.path.join(EX, name + ".yaml")) as f:
        a = yaml.safe_load(f)
    a.setdefault("precision", 32)
    a["prior_model"] = None
    a["derivative"] = True
    a.update(kw)
    return a


HP_KEYS = ["model", "precision", "embedding_dimension", "num_layers", "num_rbf", "rbf_type", "trainable_rbf",
           "activation", "cutoff_lower", "cutoff_upper", "max_z", "max_num_neighbors", "attn_activation",
           "num_heads", "distance_influence", "neighbor_embedding", "equivariance_invariance_group",
           "derivative", "atom_filter", "prior_model", "output_model", "reduce_op"]


def hp_of(a):
    return {k: a[k] for k in HP_KEYS if k in a}


def checksums(sd):
    return {k: [float(v.double().sum()), float(v.double().abs().sum())] for k, v in sd.items()}


def t(x, dt=None):
    x = torch.from_numpy(np.asarray(x))
    return x if dt is None else x.to(dt)


def run_ref(model, z, pos, batch):
    y, f = model(t(z), t(pos).clone(), t(batch))
    return y.detach().numpy(), f.detach().numpy()


def main():
    R.import_reference()
    from torchmdnet.models.model import create_model
    from torchmdnet.models.utils import OptimizedDistance
    os.makedirs(GOLD, exist_ok=True)
    torch.set_float32_matmul_precision("highest")   # tensornet.py:13 flips this at import; irrelevant on CPU

    # ---- (1) the reference's own golden file
    with open(os.path.join(R.REF_ROOT, "tests", "expected.pkl"), "rb") as f:
        exp = pickle.load(f)
    out = {}
    zs = torch.tensor([1, 6, 7, 8, 9], dtype=torch.long)
    for model_name, cfg in (("equivariant-transformer", "ET-QM9"), ("tensornet", "TensorNet-QM9")):
        R.seed_everything(1234)
        a = load_yaml_args(cfg, model=model_name, output_model="Scalar", seed=1234)
        m = create_model(a)
        z = zs[torch.randint(0, len(zs), (5,))]
        pos = torch.randn(5, 3)
        batch = torch.zeros(5, dtype=torch.long)
        batch[2:] = 1
        y, f = m(z, pos.clone(), batch)
        e = exp[model_name]["Scalar"]
        assert (y - e["pred"]).abs().max() < 1e-6 and (f - e["deriv"]).abs().max() < 1e-6
        out[model_name] = dict(hparams=hp_of(a), yaml=cfg, seed=1234, z=z.tolist(), pos=pos.tolist(),
                               batch=batch.tolist(), expected_pred=e["pred"].tolist(),
                               expected_deriv=e["deriv"].tolist(), rerun_pred=y.tolist(),
                               rerun_deriv=f.tolist(), checksums=checksums(m.state_dict()))
    with open(os.path.join(GOLD, "expected_pkl.json"), "w") as f:
        json.dump(out, f, indent=1)

    # ---- (2) full-size hparams, seeded init, synthetic inputs
    z1, p1, b1 = S.qm9_like_batch(n_mol=6, seed=1)
    z2, p2, b2 = S.chignolin_like()
    z3, p3, b3 = S.md17_like_batch(n_mol=3, seed=3)
    cases = [("ET-QM9", (z1, p1, b1)), ("ET-MD17", (z3, p3, b3)), ("TensorNet-SPICE", (z2, p2, b2)),
             ("TensorNet-rMD17", (z3, p3, b3)), ("TensorNet-ANI1X", (z1, p1, b1)),
             ("TensorNet-QM9", (z1, p1, b1))]
    for cfg, (z, p, b) in cases:
        R.seed_everything(1234)
        a = load_yaml_args(cfg)
        m = create_model(a)
        y, f = run_ref(m, z, p, b)
        np.savez_compressed(os.path.join(GOLD, "model_%s.npz" % cfg), hparams=json.dumps(hp_of(a)), seed=1234,
                            z=z, pos=p, batch=b, y=y, forces=f, checksums=json.dumps(checksums(m.state_dict())))
        print(cfg, y.flatten()[:3], np.abs(f).max())

    # ---- (3) small models with full weights (+ non-trivial mean/std, SO(3) variant, gauss rbf, fp64 input)
    small = [
        ("tensornet_o3", dict(base="TensorNet-rMD17", embedding_dimension=32, num_rbf=16, num_layers=2)),
        ("tensornet_so3", dict(base="TensorNet-rMD17", embedding_dimension=32, num_rbf=16, num_layers=1,
                               equivariance_invariance_group="SO(3)")),
        ("tensornet_gauss", dict(base="TensorNet-rMD17", embedding_dimension=32, num_rbf=8, num_layers=1,
                                 rbf_type="gauss", cutoff_lower=0.0)),
        ("et_both", dict(base="ET-MD17", embedding_dimension=32, num_rbf=16, num_layers=2, num_heads=4)),
        ("et_keys_nonbr", dict(base="ET-MD17", embedding_dimension=32, num_rbf=16, num_layers=1, num_heads=2,
                               distance_influence="keys", neighbor_embedding=False)),
    ]
    zs_, ps_, bs_ = S.qm9_like_batch(n_mol=5, seed=11)
    for name, kw in small:
        kw = dict(kw)
        base = kw.pop("base")
        R.seed_everything(7)
        a = load_yaml_args(base, **kw)
        m = create_model(a, mean=torch.tensor(0.37), std=torch.tensor(1.9))
        y, f = run_ref(m, zs_, ps_, bs_)
        sd = {k: v.numpy() for k, v in m.state_dict().items()}
        np.savez_compressed(os.path.join(GOLD, "small_%s.npz" % name), hparams=json.dumps(hp_of(a)),
                            z=zs_, pos=ps_, batch=bs_, y=y, forces=f,
                            **{"sd:" + k: v for k, v in sd.items()})
        print(name, y.flatten()[:3])

    # ---- (4) periodic TensorNet: reference model with OptimizedDistance(box=...) swapped in
    zw, pw, bw, box = S.water_box(5, seed=4)
    R.seed_everything(1234)
    a = load_yaml_args("TensorNet-ANI1X", embedding_dimension=32, num_rbf=16)
    m = create_model(a)
    rm = m.representation_model
    rm.distance = OptimizedDistance(a["cutoff_lower"], a["cutoff_upper"], max_num_pairs=-a["max_num_neighbors"],
                                    return_vecs=True, loop=True, check_errors=False, resize_to_fit=True,
                                    box=t(box), long_edge_index=True)
    y, f = run_ref(m, zw, pw, bw)
    sd = {k: v.numpy() for k, v in m.state_dict().items()}
    np.savez_compressed(os.path.join(GOLD, "small_tensornet_periodic.npz"), hparams=json.dumps(hp_of(a)),
                        z=zw, pos=pw, batch=bw, box=box, y=y, forces=f, **{"sd:" + k: v for k, v in sd.items()})
    print("periodic", y.flatten(), np.abs(f).max())

    # ---- (5) neighbour op goldens (reference compiled CPU op, neighbors_cpu.cpp)
    op = torch.ops.torchmdnet_neighbors.get_neighbor_pairs
    rng = np.random.RandomState(99)
    ncases = []
    for i, (n, nb_, cutoff, loop, tr, boxkind, dt) in enumerate([
            (40, 1, 3.0, False, True, None, np.float32), (40, 3, 4.9, True, True, None, np.float32),
            (64, 4, 1.0, False, False, "rect", np.float32), (64, 2, 3.0, True, True, "tric", np.float32),
            (100, 1, 4.9, True, False, "tric", np.float64), (30, 128, 3.0, True, True, "rect", np.float64),
            (1, 1, 3.0, True, True, None, np.float32), (2, 1, 0.1, False, True, None, np.float32)]):
        lbox = 10.0
        pos = (rng.rand(n, 3) * 2 * lbox - lbox).astype(dt)      # spans several periodic images
        batch = np.sort(rng.randint(0, nb_, n)).astype(np.int64)
        if boxkind == "rect":
            box = np.diag([lbox, lbox, lbox])
        elif boxkind == "tric":
            box = np.array([[lbox, 0, 0], [0.1 * lbox, lbox, 0], [0.3 * lbox, 0.2 * lbox, lbox]])
        else:
            box = None
        tb = torch.zeros(0, 0) if box is None else torch.from_numpy(box).to(t(pos).dtype)
        nbr, dl, ds, cnt = op("brute", t(pos), t(batch), tb, box is not None, 0.0, cutoff, 10 ** 6, loop, tr)
        np.savez_compressed(os.path.join(GOLD, "neighbors_%d.npz" % i), pos=pos, batch=batch,
                            box=np.zeros((0, 0)) if box is None else box.astype(dt), cutoff=cutoff, loop=loop,
                            transpose=tr, neighbors=nbr.numpy(), deltas=dl.numpy(), distances=ds.numpy(),
                            count=cnt.numpy())
        ncases.append((i, int(cnt[0])))
    print("neighbors", ncases)


if __name__ == "__main__":
    main()
