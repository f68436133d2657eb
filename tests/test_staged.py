"""The stage list and hand-derived adjoints that csrc/tensornet.cu transcribes (oracle/staged_tensornet.py)
against torch.autograd on the oracle restatement of the reference (oracle/model.py).  CPU, fp64."""
import pytest
import torch

from helpers import load_npz, tt
from oracle import model as om
from oracle import staged_tensornet as st


@pytest.mark.parametrize("name", ["small_tensornet_o3.npz", "small_tensornet_so3.npz", "small_tensornet_gauss.npz",
                                  "small_tensornet_periodic.npz"])
def test_staged_pipeline_matches_autograd_fp64(name):
    g = load_npz(name)
    sd = {k: (v.double() if v.is_floating_point() else v) for k, v in g["sd"].items()}
    hp = g["hparams"]
    z, pos, batch = tt(g["z"]), tt(g["pos"]).double(), tt(g["batch"])
    box = tt(g["box"]).double() if "box" in g else None
    n_mol = int(batch.max()) + 1
    y0, f0 = om.energy_forces(sd, hp, z, pos, batch, box=box)
    ei, shift = om.build_edges(pos, batch, hp, box)
    y1, f1 = st.energy_forces_staged(sd, hp, z, pos, batch, ei, shift, n_mol)
    torch.testing.assert_close(y1, y0, rtol=1e-10, atol=1e-10)
    torch.testing.assert_close(f1, f0, rtol=1e-8, atol=1e-9)


def test_irreps_roundtrip_and_norm():
    torch.manual_seed(0)
    x = torch.randn(7, 5, 9, dtype=torch.float64)
    T = st.to_mat(x)
    torch.testing.assert_close(st.from_mat(T), x)
    torch.testing.assert_close(st.qnorm(x), (T ** 2).sum((-2, -1)))
    I, A, S = om._decompose(T)
    qi, qa, qs = st.qnorm_parts(x)
    torch.testing.assert_close(qi, (I ** 2).sum((-2, -1)))
    torch.testing.assert_close(qa, (A ** 2).sum((-2, -1)))
    torch.testing.assert_close(qs, (S ** 2).sum((-2, -1)))


@pytest.mark.parametrize("name", ["small_et_both.npz", "small_et_keys_nonbr.npz"])
def test_staged_et_pipeline_matches_autograd_fp64(name):
    from oracle import staged_et as se
    g = load_npz(name)
    sd = {k: (v.double() if v.is_floating_point() else v) for k, v in g["sd"].items()}
    hp = g["hparams"]
    z, pos, batch = tt(g["z"]), tt(g["pos"]).double(), tt(g["batch"])
    n_mol = int(batch.max()) + 1
    y0, f0 = om.energy_forces(sd, hp, z, pos, batch)
    ei, shift = om.build_edges(pos, batch, hp, None)
    y1, f1 = se.energy_forces_staged(sd, hp, z, pos, batch, ei, shift, n_mol)
    torch.testing.assert_close(y1, y0, rtol=1e-10, atol=1e-10)
    torch.testing.assert_close(f1, f0, rtol=1e-8, atol=1e-9)
