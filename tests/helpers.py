"""Shared test helpers: golden fixture loading."""
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_npz(name):
    d = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    out = {k: d[k] for k in d.files if not k.startswith("sd:")}
    sd = {k[3:]: torch.from_numpy(d[k]) for k in d.files if k.startswith("sd:")}
    if "hparams" in out:
        out["hparams"] = json.loads(str(out["hparams"]))
    if "checksums" in out:
        out["checksums"] = json.loads(str(out["checksums"]))
    out["sd"] = sd
    return out


def tt(x, dtype=None):
    t = torch.from_numpy(np.asarray(x))
    return t if dtype is None else t.to(dtype)


SMALL = ["small_tensornet_o3.npz", "small_tensornet_so3.npz", "small_tensornet_gauss.npz",
         "small_et_both.npz", "small_et_keys_nonbr.npz"]
