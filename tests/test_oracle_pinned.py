"""Pins the oracle (oracle/model.py, oracle/neighbors.py) against fixtures produced by the unmodified
reference (oracle/make_golden.py).  CPU only."""
import glob
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN, SMALL, load_npz, tt
from oracle import model as om
from oracle import neighbors as onb

# the reference's own comparison tolerance for model goldens (torch.testing.assert_allclose fp32 defaults,
# tests/test_model.py:184-188): rtol 1.3e-6, atol 1e-5
RTOL, ATOL = 1.3e-6, 1e-5


@pytest.mark.parametrize("name", SMALL + ["small_tensornet_periodic.npz", "small_gn_ne.npz", "small_gn_plain.npz",
                                          "small_t_both.npz", "small_t_values_nonbr.npz"])
def test_oracle_matches_reference_small_models(name):
    g = load_npz(name)
    box = tt(g["box"]) if "box" in g else None
    y, f = om.energy_forces(g["sd"], g["hparams"], tt(g["z"]), tt(g["pos"]), tt(g["batch"]), box=box)
    torch.testing.assert_close(y, tt(g["y"]), rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(f, tt(g["forces"]), rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize("i", range(8))
def test_oracle_neighbors_match_reference_cpu_op(i):
    g = load_npz("neighbors_%d.npz" % i)
    box = g["box"] if g["box"].size else None
    nb, dl, ds = onb.neighbor_pairs(g["pos"], g["batch"], box, 0.0, float(g["cutoff"]), loop=bool(g["loop"]),
                                    include_transpose=bool(g["transpose"]), d2_rule=False)
    assert nb.shape[1] == int(g["count"][0]) == g["neighbors"].shape[1]
    a = onb.canonical_order(nb, dl, ds)
    b = onb.canonical_order(g["neighbors"], g["deltas"], g["distances"])
    assert np.array_equal(a[0], b[0])                       # pair set bit-exact after sorting
    np.testing.assert_allclose(a[1], b[1], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(a[2], b[2], rtol=1e-5, atol=1e-6)


def test_neighbor_backward_restatement_matches_autograd():
    rng = np.random.RandomState(0)
    pos = rng.randn(30, 3) * 1.5
    nb, dl, ds = onb.neighbor_pairs(pos, None, None, 0.0, 2.5, loop=True)
    gd, gw = rng.randn(*dl.shape), rng.randn(*ds.shape)
    p = torch.tensor(pos, requires_grad=True)
    i, j = torch.from_numpy(nb[0].astype("int64")), torch.from_numpy(nb[1].astype("int64"))
    v = p[i] - p[j]
    nz = torch.from_numpy(ds != 0)
    w = v[nz].norm(dim=1)
    loss = (v[nz] * torch.from_numpy(gd)[nz]).sum() + (w * torch.from_numpy(gw)[nz]).sum()
    loss.backward()
    np.testing.assert_allclose(onb.neighbor_backward(nb, dl, ds, gd, gw, 30), p.grad.numpy(), atol=1e-10)


@pytest.mark.parametrize("arch,tag,fixture", [("graph-network", "gn", "model_GN.npz"), ("transformer", "t", "model_T.npz")])
def test_oracle_matches_reference_gn_and_t_goldens(arch, tag, fixture):
    """The graph-network / transformer restatements against the reference's own golden rows (tests/expected.pkl
    <arch>/Scalar, numbers in tests/golden/expected_pkl_{gn,t}.json) and the full-size fixtures; weights = seeded
    create_model of this package, whose initial values are checked against the reference's checksums."""
    import json
    import random
    from torchmdnet.models.model import create_model
    with open(os.path.join(GOLDEN, "expected_pkl_%s.json" % tag)) as fh:
        e = json.load(fh)[arch]
    for hp, seed, z, pos, batch, y_ref, f_ref, sums in (
            (e["hparams"], 1234, torch.tensor(e["z"]), torch.tensor(e["pos"]), torch.tensor(e["batch"]),
             torch.tensor(e["expected_pred"]), torch.tensor(e["expected_deriv"]), e["checksums"]),) + tuple(
            (g["hparams"], int(g["seed"]), tt(g["z"]), tt(g["pos"]), tt(g["batch"]), tt(g["y"]), tt(g["forces"]), g["checksums"])
            for g in [load_npz(fixture)]):
        random.seed(seed)
        np.random.seed(seed)
        torch.manual_seed(seed)
        m = create_model(hp)
        sd = m.state_dict()
        assert set(sd) == set(sums)
        for k, v in sd.items():
            assert abs(float(v.double().sum()) - sums[k][0]) <= 1e-5 * max(1.0, abs(sums[k][0])), k
            assert abs(float(v.double().abs().sum()) - sums[k][1]) <= 1e-5 * max(1.0, sums[k][1]), k
        y, f = om.energy_forces({k: v.detach() for k, v in sd.items()}, hp, z, pos, batch)
        torch.testing.assert_close(y, y_ref, rtol=RTOL, atol=ATOL)
        torch.testing.assert_close(f, f_ref, rtol=1e-5, atol=ATOL)


def test_pair_prior_oracle_against_reference_fixtures():
    """oracle/priors.py (ZBL, Coulomb, D2 in numpy fp64, forces by central differences) against what the unmodified reference
    wrote: D2 alone in fp64 (energies to 1e-12, forces to 1e-8), D2 against the Psi4 energies of the reference's
    tests/test_prior_d2.py (1e-5), ZBL + Coulomb as the difference of the reference's fp32 runs with and without the priors."""
    from oracle import priors as P
    g = load_npz("small_tensornet_d2.npz")
    da = g["hparams"]["prior_model"][0]["D2"]
    z, pos, batch = g["z"], g["pos"], g["batch"]
    nm = int(batch.max()) + 1
    fn = lambda p: P.d2_energy(p, batch, z, nm, da["atomic_number"], da["cutoff_distance"], da["distance_scale"],  # noqa: E731
                               da["energy_scale"])
    e = fn(pos)
    assert np.abs(e - g["y_prior_f64"].ravel()).max() <= 1e-12 * np.abs(e).max()
    assert np.abs(P.forces_fd(fn, pos) - g["forces_prior_f64"]).max() <= 1e-8
    psi4 = {"h2": ([1, 1], [[0.0, 0.0, 0.0], [1.0, 0.0, 0.0]], [0, 0], [-2.396695338419219e-06]),
            "h2o_n2_ch4": ([1, 8, 1, 7, 7, 6, 1, 1, 1],
                           [[0.0, 0.0, 0.0], [0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [1.0, 2.0, 3.0], [2.0, 1.0, 3.0], [0.0, 0.0, 0.0],
                            [1.0, 1.0, 0.0], [0.0, 1.0, 1.0], [1.0, 0.0, 1.0]], [0, 0, 0, 1, 1, 2, 2, 2, 2],
                           [-4.7815923481300465e-06, -3.0071055390390565e-06, -7.100283425158895e-05])}
    for zz, pp, bb, yy in psi4.values():
        e = P.d2_energy(np.array(pp), np.array(bb), np.array(zz), max(bb) + 1, list(range(100)), 10.0, 1e-10, 4.35974e-18)
        np.testing.assert_allclose(e, yy, rtol=1e-5, atol=0)
    g = load_npz("small_tensornet_pairpriors.npz")
    za, ca = g["hparams"]["prior_model"][0]["ZBL"], g["hparams"]["prior_model"][1]["Coulomb"]
    z, pos, batch, q = g["z"], g["pos"], g["batch"], g["partial_charges"]
    nm = int(batch.max()) + 1
    fn = lambda p: (P.zbl_energy(p, batch, z, nm, za["atomic_number"], za["cutoff_distance"], za["distance_scale"],  # noqa: E731
                                 za["energy_scale"])
                    + P.coulomb_energy(p, batch, q, nm, ca["alpha"], ca["distance_scale"], ca["energy_scale"]))
    ref = (g["y"].astype(np.float64) - g["y_noprior"]).ravel()          # two fp32 runs of the reference
    assert np.abs(fn(pos) - ref).max() <= 3e-6 * np.abs(ref).max()
    fref = g["forces"].astype(np.float64) - g["forces_noprior"]
    assert np.abs(P.forces_fd(fn, pos) - fref).max() <= 2e-6 * np.abs(fref).max()
