"""Fused CUDA Equivariant Transformer energy+forces (TorchMD_Net.forward -> torch op -> C ABI) against the oracle
restatement of the reference CPU path, fixtures written by the unmodified reference, the reference's own golden
row (tests/expected.pkl ET/Scalar) and the staged restatement of the kernel pipeline buffer by buffer.
Tolerances as in test_gpu_tensornet.py (north_star: energies 1e-5 relative, forces 1e-4 eV/A absolute)."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN, load_npz, tt
from oracle import model as om
from oracle import staged_et as se
from test_gpu_tensornet import check, seed_everything

pytestmark = pytest.mark.gpu


def build(g):
    from torchmdnet.models.model import create_model
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    return m.cuda().eval()


@pytest.mark.parametrize("name", ["small_et_both.npz", "small_et_keys_nonbr.npz"])
def test_small_models_match_reference(name):
    g = load_npz(name)
    m = build(g)
    y, f = m(tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda())
    check(y, f, g["y"], g["forces"])


@pytest.mark.parametrize("cfg", ["ET-QM9", "ET-MD17"])
def test_full_size_hparams_match_reference(cfg):
    from torchmdnet.models.model import create_model
    g = load_npz("model_%s.npz" % cfg)
    seed_everything(int(g["seed"]))
    m = create_model(g["hparams"]).cuda().eval()
    y, f = m(tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda())
    check(y, f, g["y"], g["forces"])


@pytest.mark.parametrize("cfg", ["ET-QM9", "ET-MD17"])
def test_table_grid_sweep_against_the_reference(cfg):
    """ET evaluates its distance filters only through tables, so its parity rests on the table grid: the default
    (max(1024, 32 * num_rbf) intervals uniform in t(d), tensornet.TableGrid), coarser and finer ones, and the round-1 grid
    (4096 intervals uniform in d) all meet the reference fixture at the parity tolerances; the spread between the
    grids shows the margin."""
    from torchmdnet.models.model import create_model
    g = load_npz("model_%s.npz" % cfg)
    seed_everything(int(g["seed"]))
    m = create_model(g["hparams"]).cuda().eval()
    rm = m.representation_model
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    K = int(g["hparams"]["num_rbf"])
    out = {}
    for warp, n_int in [(True, None), (True, 16 * K), (True, 64 * K), (True, 4096), (False, 4096), (False, None)]:
        rm.table_warp, rm.table_intervals = warp, n_int
        y, f = m(z, pos, batch)
        check(y, f, g["y"], g["forces"])
        out[warp, n_int] = (y.double().cpu(), f.double().cpu())
    y_ref, f_ref = out[False, 4096]
    y, f = out[True, None]
    assert float((y - y_ref).abs().max()) <= 2e-6 * max(1.0, float(y_ref.abs().max()))
    assert float((f - f_ref).abs().max()) <= 2e-5


def test_reference_expected_pkl_row():
    """tests/expected.pkl, equivariant-transformer/Scalar row (tests/test_model.py:137-188)."""
    from torchmdnet.models.model import create_model
    with open(os.path.join(GOLDEN, "expected_pkl.json")) as fh:
        e = json.load(fh)["equivariant-transformer"]
    seed_everything(1234)
    m = create_model(e["hparams"]).cuda().eval()
    y, f = m(torch.tensor(e["z"]).cuda(), torch.tensor(e["pos"]).cuda(), torch.tensor(e["batch"]).cuda())
    torch.testing.assert_close(y.cpu(), torch.tensor(e["expected_pred"]), rtol=1.3e-6, atol=1e-5)
    torch.testing.assert_close(f.cpu(), torch.tensor(e["expected_deriv"]), rtol=1.3e-6, atol=1e-5)


def test_intermediate_buffers_match_staged_restatement():
    g = load_npz("small_et_both.npz")
    m = build(g)
    z, pos, batch = tt(g["z"]), tt(g["pos"]), tt(g["batch"])
    n, n_mol = z.shape[0], int(batch.max()) + 1
    y, f = m(z.cuda(), pos.cuda(), batch.cuda())
    sd = {k: (v.double() if v.is_floating_point() else v) for k, v in g["sd"].items()}
    hp = g["hparams"]
    ei, shift = om.build_edges(pos.double(), batch, hp, None)
    keep = {}
    ys, fs = se.energy_forces_staged(sd, hp, z, pos.double(), batch, ei, shift, n_mol, keep)
    rm = m.representation_model
    H = hp["embedding_dimension"]
    rp = rm.debug_buffer("row_ptr", n, n_mol, torch.int32)[:n + 1].cpu().long()
    assert torch.equal(rp, keep["row_ptr"])
    for name, shape in [("x0", (n, H)), ("x1", (n, H)), ("vec1", (n, 3, H)), ("x2", (n, H)), ("vec2", (n, 3, H))]:
        got = rm.debug_buffer(name, n, n_mol)[:int(np.prod(shape))].view(shape).cpu().double()
        ref = keep[name]
        tol = 5e-5 * max(1.0, float(ref.abs().max()))
        assert float((got - ref).abs().max()) < tol, (name, float((got - ref).abs().max()), tol)
    qkv = rm.debug_buffer("qkv0", n, n_mol)[:n * 5 * H].view(n, 5 * H).cpu().double()
    torch.testing.assert_close(qkv[:, :H], keep["q0"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(qkv[:, 2 * H:].reshape(n, 3, H), keep["v0"], rtol=1e-4, atol=1e-5)
    check(y, f, ys, fs)


def test_forward_only_and_batch_none_and_repeatability():
    g = load_npz("small_et_both.npz")
    m = build(g)
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    y, f = m(z, pos, batch)
    y2, f2 = m(z, pos, batch)
    assert torch.equal(y, y2) and torch.equal(f, f2)          # no float atomics anywhere
    sel = batch == 0
    y1, f1 = m(z[sel], pos[sel])
    torch.testing.assert_close(y1[0], y[0], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(f1, f[sel], rtol=0, atol=1e-6)
    m.derivative = False
    y3, none = m(z, pos, batch)
    assert none is None and torch.equal(y3, y)


def test_isolated_atom_has_finite_forces():
    """An atom outside every cutoff has vec == 0: the gated block's zero mask (models/utils.py:499-512) must give
    zero, not NaN, gradients (tests/test_model_utils.py:87-104 of the reference)."""
    g = load_npz("small_et_both.npz")
    m = build(g)
    z = tt(g["z"])[:6].cuda()
    pos = tt(g["pos"])[:6].clone()
    pos[5] += 100.0
    y, f = m(z, pos.cuda())
    assert torch.isfinite(y).all() and torch.isfinite(f).all()
    assert float(f[5].abs().max()) == 0.0
    yr, fr = om.energy_forces(g["sd"], g["hparams"], z.cpu(), pos)
    check(y, f, yr, fr)


def test_rotation_invariance_and_force_equivariance():
    g = load_npz("small_et_both.npz")
    m = build(g)
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    R = torch.tensor([[0.9362934, -0.2896295, 0.1986693], [0.3129918, 0.9447025, -0.0978434],
                      [-0.1593451, 0.1537920, 0.9751703]], device="cuda")
    y, f = m(z, pos, batch)
    y2, f2 = m(z, pos @ R.T, batch)
    torch.testing.assert_close(y2, y, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(f2, f @ R.T, rtol=0, atol=1e-4)


def test_qm9_like_batch_of_128_molecules():
    """BASELINE.json configs[0]: ET-QM9 hparams, 128 QM9-size molecules, against the oracle on a 12-molecule slice
    (the oracle is O(N^2) over the batch) and shard-vs-batch consistency for the rest."""
    import synth_inputs as S
    from torchmdnet.models.model import create_model
    g = load_npz("model_ET-QM9.npz")
    seed_everything(int(g["seed"]))
    m = create_model(g["hparams"]).cuda().eval()
    z, pos, batch = (torch.from_numpy(a) for a in S.qm9_like_batch(n_mol=128, seed=1))
    y, f = m(z.cuda(), pos.cuda(), batch.cuda())
    assert y.shape == (128, 1) and torch.isfinite(f).all()
    sel = batch < 12
    sd = {k: v.cpu() for k, v in m.state_dict().items()}
    yr, fr = om.energy_forces(sd, g["hparams"], z[sel], pos[sel], batch[sel])
    check(y[:12], f[sel.cuda()], yr, fr)
    sel2 = batch >= 100
    y2, f2 = m(z[sel2].cuda(), pos[sel2].cuda(), (batch[sel2] - 100).cuda())
    torch.testing.assert_close(y2, y[100:], rtol=2e-6, atol=2e-6)
    torch.testing.assert_close(f2, f[sel2.cuda()], rtol=0, atol=5e-6)


def test_projection_gemm_modes_agree():
    """gemm_mode="tf32x3" (default) runs every ET projection on the tcgen05 tensor cores (three-product tf32 split, K
    chunks summed in fp32), "fp32" on the FFMA kernel: two arithmetic paths, both inside the fp32 parity bar."""
    import synth_inputs as S
    from torchmdnet.models.model import create_model
    g = load_npz("model_ET-QM9.npz")
    seed_everything(int(g["seed"]))
    m = create_model(g["hparams"]).cuda().eval()
    z, pos, batch = (torch.from_numpy(a) for a in S.qm9_like_batch(n_mol=128, seed=1))
    m.representation_model.gemm_mode = "fp32"
    y0, f0 = m(z.cuda(), pos.cuda(), batch.cuda())
    m.representation_model.gemm_mode = "tf32x3"
    n0 = torch.ops.torchmdnet_b200.launch_count()
    y1, f1 = m(z.cuda(), pos.cuda(), batch.cuda())
    assert torch.ops.torchmdnet_b200.launch_count() > n0
    assert not torch.equal(y0, y1)                       # a different arithmetic path really ran
    sel = batch < 12
    sd = {k: v.cpu() for k, v in m.state_dict().items()}
    yr, fr = om.energy_forces(sd, g["hparams"], z[sel], pos[sel], batch[sel])
    dy = (y1[:12].double().cpu() - yr.double()).abs()
    assert float((dy / yr.double().abs().clamp_min(1.0)).max()) <= 1e-5
    assert float((f1[sel.cuda()].double().cpu() - fr.double()).abs().max()) <= 1e-4
    assert float((y1 - y0).abs().max()) <= 1e-5 * float(y0.abs().max())
    m.representation_model.gemm_mode = "fp32"
    y2, _ = m(z.cuda(), pos.cuda(), batch.cuda())
    assert torch.equal(y0, y2)


def test_cuda_graph_and_external(tmp_path):
    from torchmdnet.calculators import External
    g = load_npz("small_et_both.npz")
    ckpt = str(tmp_path / "et.ckpt")
    torch.save({"hyper_parameters": dict(g["hparams"]), "state_dict": {"model." + k: v for k, v in g["sd"].items()}},
               ckpt)
    z, pos = tt(g["z"])[:9], tt(g["pos"])[:9]
    calc = External(ckpt, z.unsqueeze(0), "cuda")
    outs = [calc.calculate(pos + 0.01 * i, None) for i in range(5)]      # calls 3.. replay the captured graph
    m = build(g)
    for i, (e, f) in enumerate(outs):
        er, fr = m(z.cuda(), (pos + 0.01 * i).cuda())
        torch.testing.assert_close(e, er, rtol=0, atol=0)
        torch.testing.assert_close(f[0], fr, rtol=0, atol=0)
