"""Training mode: model.train() routes TorchMD_Net.forward through the differentiable composition of
models/eager.py on the GPU -- parameter gradients of an energy + force-matching loss (the second-order path of
models/model.py:286-298, caller module.py:130-179) against the oracle's autograd on the CPU, for all four architectures;
eval mode stays on the fused kernels and agrees with the training-mode forward."""
import pytest
import torch

from helpers import load_npz, tt
from oracle import model as om

pytestmark = pytest.mark.gpu

CASES = ["small_tensornet_o3.npz", "small_et_both.npz", "small_t_both.npz", "small_gn_ne.npz"]


def _loss(y, f, wy, wf):
    return (y * wy).sum() + (f * wf).pow(2).sum() + f.sum(0).pow(2).sum()


@pytest.mark.parametrize("name", CASES)
def test_parameter_gradients_of_a_force_matching_loss(name):
    from torchmdnet.models.model import create_model
    g = load_npz(name)
    hp = g["hparams"]
    m = create_model(hp)
    m.load_state_dict(g["sd"])
    m = m.cuda()
    z, pos, batch = tt(g["z"]), tt(g["pos"]), tt(g["batch"])
    n_mol = int(batch.max()) + 1
    gen = torch.Generator().manual_seed(5)
    wy, wf = torch.randn(n_mol, 1, generator=gen), torch.randn(pos.shape[0], 3, generator=gen)
    # ---- eval mode: fused kernels, no graph
    m.eval()
    if hp["model"] == "tensornet":
        m.representation_model.filter_mode = "exact"
    y_e, f_e = m(z.cuda(), pos.cuda(), batch.cuda())
    assert not y_e.requires_grad and not f_e.requires_grad
    # ---- training mode: differentiable, the same numbers
    m.train()
    y, f = m(z.cuda(), pos.cuda(), batch.cuda())
    assert y.requires_grad and f.requires_grad
    torch.testing.assert_close(y.detach(), y_e, rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(f.detach(), f_e, rtol=0, atol=5e-5)
    _loss(y, f, wy.cuda(), wf.cuda()).backward()
    grads = {k: p.grad.detach().cpu() for k, p in m.named_parameters() if p.grad is not None}
    # ---- oracle: the same loss through autograd on the CPU (fp64 weights for a clean reference)
    sd = {k: (v.double().requires_grad_(v.is_floating_point() and k in dict(m.named_parameters())))
          if v.is_floating_point() else v for k, v in g["sd"].items()}
    p64 = pos.double().requires_grad_(True)
    ei, shift = om.build_edges(p64, batch, hp, None)
    if hp["model"] == "graph-network":
        pass                                               # gn_representation drops the self pairs itself
    x = om.atom_energies(sd, hp, z, p64, ei, shift.double())
    y_o = torch.zeros(n_mol, 1, dtype=torch.float64).index_add(0, batch, x) + sd["mean"]
    (dy,) = torch.autograd.grad(y_o.sum(), p64, create_graph=True)
    _loss(y_o, -dy, wy.double(), wf.double()).backward()
    checked = 0
    for k, v in sd.items():
        if not (torch.is_tensor(v) and v.requires_grad) or v.grad is None:
            continue
        if k.startswith("representation_model.") and ".conv.net." in k:
            continue                                        # the GN filter MLP is registered twice; compared under .mlp.
        assert k in grads, k
        scale = max(1.0, float(v.grad.abs().max()))
        assert float((grads[k].double() - v.grad).abs().max()) <= 2e-3 * scale, (k, float((grads[k].double() - v.grad).abs().max()), scale)
        checked += 1
    assert checked >= 10
    torch.testing.assert_close(y.detach().cpu().double(), y_o.detach(), rtol=1e-5, atol=2e-5)


def test_training_mode_with_frozen_parameters_uses_the_fused_path():
    from torchmdnet.models.model import create_model
    g = load_npz("small_tensornet_o3.npz")
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    m = m.cuda().train()
    for p in m.parameters():
        p.requires_grad_(False)
    y, f = m(tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda())
    assert not y.requires_grad
    with torch.no_grad():
        for p in m.parameters():
            p.requires_grad_(True)
        y2, _ = m(tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda())
    assert not y2.requires_grad and torch.equal(y, y2)


@pytest.mark.parametrize("name", CASES)
def test_double_precision_models(name):
    """precision: 64 -- the fused kernels are fp32, double-precision models evaluate through models/eager.py with the fp64
    neighbour op: energies and forces against the oracle in fp64 (1e-9), forces = -dE/dpos by central differences, and the
    result carries a graph only when grad is enabled (reference: tests/test_model.py:191-211 runs gradcheck on such models)."""
    from torchmdnet.models.model import create_model
    g = load_npz(name)
    hp = dict(g["hparams"], precision=64)
    m = create_model(hp)
    m.load_state_dict({k: v.double() if v.is_floating_point() else v for k, v in g["sd"].items()})
    m = m.cuda().eval()
    z, pos, batch = tt(g["z"]), tt(g["pos"]).double(), tt(g["batch"])
    with torch.no_grad():
        y, f = m(z.cuda(), pos.cuda(), batch.cuda())
    assert y.dtype == torch.float64 and not y.requires_grad and not f.requires_grad
    sd = {k: v.double() if v.is_floating_point() else v for k, v in g["sd"].items()}
    y_o, f_o = om.energy_forces(sd, hp, z, pos, batch)
    torch.testing.assert_close(y.cpu(), y_o, rtol=1e-9, atol=1e-9)
    torch.testing.assert_close(f.cpu(), f_o, rtol=0, atol=1e-8)
    h = 1e-5
    for (i, c) in ((1, 0), (7, 2)):
        pp, pm = pos.clone(), pos.clone()
        pp[i, c] += h
        pm[i, c] -= h
        with torch.no_grad():
            ep = m(z.cuda(), pp.cuda(), batch.cuda())[0].sum()
            em = m(z.cuda(), pm.cuda(), batch.cuda())[0].sum()
        assert abs(float(-(ep - em) / (2 * h)) - float(f[i, c])) <= 1e-6 * max(1.0, abs(float(f[i, c])))
    p = pos.cuda().requires_grad_(True)
    y2, f2 = m(z.cuda(), p, batch.cuda())
    assert y2.requires_grad and f2.requires_grad                 # second derivatives are available (create_graph=True)
    (h2,) = torch.autograd.grad(f2.pow(2).sum(), p)
    assert bool(torch.isfinite(h2).all())
