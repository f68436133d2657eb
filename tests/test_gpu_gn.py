"""Fused CUDA graph network (TorchMD_GN) and transformer (TorchMD_T) energy+forces against fixtures written by the unmodified reference, the
reference's own golden rows (tests/expected.pkl graph-network/Scalar, transformer/Scalar) and the oracle restatement; scripting and capture."""
import json
import os

import pytest
import torch

from helpers import GOLDEN, load_npz, tt
from oracle import model as om
from test_gpu_tensornet import check, seed_everything

pytestmark = pytest.mark.gpu


def build(g):
    from torchmdnet.models.model import create_model
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    return m.cuda().eval()


@pytest.mark.parametrize("name", ["small_gn_ne.npz", "small_gn_plain.npz", "small_t_both.npz", "small_t_values_nonbr.npz"])
def test_small_models_match_reference(name):
    g = load_npz(name)
    m = build(g)
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    y, f = m(z, pos, batch)
    check(y, f, g["y"], g["forces"])
    y_or, f_or = om.energy_forces(g["sd"], g["hparams"], tt(g["z"]), tt(g["pos"]), tt(g["batch"]))
    check(y, f, y_or, f_or)
    m.representation_model.gemm_mode = "fp32"
    y2, f2 = m(z, pos, batch)
    check(y2, f2, g["y"], g["forces"])
    # forward only, and one molecule with batch=None
    m.derivative = False
    y3, none = m(z, pos, batch)
    assert none is None
    torch.testing.assert_close(y3, y2, rtol=0, atol=0)
    m.derivative = True
    n0 = int((tt(g["batch"]) == 0).sum())
    ya, fa = m(z[:n0], pos[:n0])
    torch.testing.assert_close(ya, y2[:1], rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(fa, f2[:n0], rtol=0, atol=1e-5)


@pytest.mark.parametrize("arch,tag,fixture", [("graph-network", "gn", "model_GN.npz"), ("transformer", "t", "model_T.npz")])
def test_full_size_and_reference_golden_row(arch, tag, fixture):
    from torchmdnet.models.model import create_model
    g = load_npz(fixture)
    seed_everything(int(g["seed"]))
    m = create_model(g["hparams"]).cuda().eval()
    y, f = m(tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda())
    check(y, f, g["y"], g["forces"])
    with open(os.path.join(GOLDEN, "expected_pkl_%s.json" % tag)) as fh:
        e = json.load(fh)[arch]
    seed_everything(1234)
    m = create_model(e["hparams"]).cuda().eval()
    y, f = m(torch.tensor(e["z"]).cuda(), torch.tensor(e["pos"]).cuda(), torch.tensor(e["batch"]).cuda())
    torch.testing.assert_close(y.cpu(), torch.tensor(e["expected_pred"]), rtol=1.3e-6, atol=1e-5)
    torch.testing.assert_close(f.cpu(), torch.tensor(e["expected_deriv"]), rtol=1.3e-6, atol=1e-5)


@pytest.mark.parametrize("name", ["small_gn_ne.npz", "small_t_both.npz"])
def test_forces_are_minus_energy_gradient_and_rotations(name):
    g = load_npz(name)
    m = build(g)
    m.representation_model.gemm_mode = "fp32"
    z, batch = tt(g["z"]).cuda(), tt(g["batch"]).cuda()
    pos = tt(g["pos"]).double()
    y0, f0 = m(z, pos.float().cuda(), batch)
    h = 2e-3
    for (i, c) in ((0, 0), (5, 1), (20, 2)):
        pp, pm = pos.clone(), pos.clone()
        pp[i, c] += h
        pm[i, c] -= h
        ep = m(z, pp.float().cuda(), batch)[0].double().sum()
        em = m(z, pm.float().cuda(), batch)[0].double().sum()
        assert abs(float(-(ep - em) / (2 * h)) - float(f0[i, c])) <= 5e-3 * max(1.0, abs(float(f0[i, c])))
    th = 0.7
    R = torch.tensor([[1, 0, 0], [0, torch.cos(torch.tensor(th)), -torch.sin(torch.tensor(th))],
                      [0, torch.sin(torch.tensor(th)), torch.cos(torch.tensor(th))]], dtype=torch.float32)
    y1, f1 = m(z, (pos.float() @ R.t()).cuda(), batch)
    torch.testing.assert_close(y1, y0, rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(f1.cpu(), f0.cpu() @ R.t(), rtol=0, atol=5e-5)


@pytest.mark.parametrize("name", ["small_gn_ne.npz", "small_t_both.npz"])
def test_scripted_model_and_capture(name, tmp_path):
    g = load_npz(name)
    m = build(g)
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    y0, f0 = m(z, pos, batch)
    path = str(tmp_path / "gn.pt")
    torch.jit.save(torch.jit.script(m), path)
    lm = torch.jit.load(path, map_location="cuda")
    y1, f1 = lm(z, pos, batch)
    assert torch.equal(y0, y1) and torch.equal(f0, f1)
    static_pos = pos.clone()
    s_ = torch.cuda.Stream()
    s_.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s_):
        lm(z, static_pos, batch)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s_):
            yg, fg = lm(z, static_pos, batch)
    torch.cuda.current_stream().wait_stream(s_)
    gr.replay()
    torch.cuda.synchronize()
    assert torch.equal(yg, y0) and torch.equal(fg, f0)


def test_neighbour_budget_overflow_is_loud():
    g = load_npz("small_gn_ne.npz")
    m = build(g)
    m.representation_model.distance.max_num_pairs = -2
    with pytest.raises(RuntimeError, match="max_num_pairs"):
        m(tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda())
