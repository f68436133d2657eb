"""Deployment surface: TorchScript of the fused models -- script, save, load, CUDA-graph capture -- as
the reference pins it (tests/test_model.py:40-58, 85-125), and calculators.External with a periodic box."""
import pytest
import torch

import synth_inputs as S
from helpers import load_npz, tt
from test_gpu_tensornet import check, seed_everything

pytestmark = pytest.mark.gpu


def _small(name, mode="table"):
    from torchmdnet.models.model import create_model
    g = load_npz(name)
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    m = m.cuda().eval()
    if g["hparams"]["model"] == "tensornet":
        m.representation_model.filter_mode = mode
    return g, m


@pytest.mark.parametrize("name", ["small_tensornet_o3.npz", "small_tensornet_so3.npz", "small_et_both.npz",
                                  "small_et_keys_nonbr.npz"])
def test_scripted_model_equals_eager_and_survives_save_load(name, tmp_path):
    g, m = _small(name)
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    y0, f0 = m(z, pos, batch)
    check(y0, f0, g["y"], g["forces"])
    sm = torch.jit.script(m)
    y1, f1 = sm(z, pos, batch)
    assert torch.equal(y0, y1) and torch.equal(f0, f1)          # the same native model behind both
    path = str(tmp_path / "model.pt")
    torch.jit.save(sm, path)
    lm = torch.jit.load(path, map_location="cuda")
    y2, f2 = lm(z, pos, batch)
    assert torch.equal(y0, y2) and torch.equal(f0, f2)
    # scripted on the CPU, moved afterwards: the packed buffer follows .to()
    g2, mc = _small(name)
    sc = torch.jit.script(mc.cpu()).to("cuda")
    y3, f3 = sc(z, pos, batch)
    check(y3, f3, g["y"], g["forces"])
    # batch=None: one molecule
    n0 = int((batch == 0).sum())
    ya, fa = m(z[:n0], pos[:n0])
    yb, fb = lm(z[:n0], pos[:n0], None)
    assert torch.equal(ya, yb) and torch.equal(fa, fb)


def test_scripted_model_in_a_cuda_graph(tmp_path):
    """script -> save -> load -> capture (reference: tests/test_model.py:85-125): warm up eagerly (caches the number of
    molecules), capture, replay on new positions."""
    g, m = _small("small_tensornet_o3.npz")
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    path = str(tmp_path / "model.pt")
    torch.jit.save(torch.jit.script(m), path)
    lm = torch.jit.load(path, map_location="cuda")
    static_pos = pos.clone()
    s_ = torch.cuda.Stream()
    s_.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s_):
        for _ in range(2):
            lm(z, static_pos, batch)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s_):
            yg, fg = lm(z, static_pos, batch)
    torch.cuda.current_stream().wait_stream(s_)
    for k in range(3):
        p = pos + 0.01 * k * torch.sin(torch.arange(pos.numel(), device="cuda").float()).view_as(pos)
        static_pos.copy_(p)
        gr.replay()
        torch.cuda.synchronize()
        y, f = m(z, p, batch)
        assert torch.equal(y, yg) and torch.equal(f, fg), k


def test_scripted_periodic_model_carries_its_box():
    from torchmdnet.models.utils import OptimizedDistance
    g, m = _small("small_tensornet_periodic.npz")
    hp = g["hparams"]
    m.representation_model.distance = OptimizedDistance(
        hp["cutoff_lower"], hp["cutoff_upper"], max_num_pairs=-hp["max_num_neighbors"], return_vecs=True, loop=True,
        check_errors=False, resize_to_fit=False, box=tt(g["box"]), strategy="cell")
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    sm = torch.jit.script(m)
    assert len(sm.box) == 9 and sm.use_periodic and sm.strategy == "cell"
    y, f = sm(z, pos, batch)
    check(y, f, g["y"], g["forces"])


def test_external_uses_its_box(tmp_path):
    """External.calculate(pos, box): the reference ignores box (calculators.py:50-56); here it selects periodic boundaries.
    Checked against the reference-written periodic fixture (375 atoms; cell list) and against open boundaries."""
    from torchmdnet.calculators import External
    g = load_npz("small_tensornet_periodic.npz")
    ckpt = str(tmp_path / "m.ckpt")
    torch.save({"hyper_parameters": dict(g["hparams"]), "state_dict": {"model." + k: v for k, v in g["sd"].items()}}, ckpt)
    n = g["z"].shape[0]
    calc = External(ckpt, tt(g["z"]).view(1, n), "cuda")
    calc.model.representation_model.filter_mode = "table"
    pos, box = tt(g["pos"]), tt(g["box"])
    for k in range(5):                                   # eager calls, graph capture, replays
        e, f = calc.calculate(pos, box)
        check(e, f.view(n, 3), g["y"], g["forces"])
    assert calc.model.representation_model.distance.strategy == "cell" and calc._graph is not None
    e_open, f_open = calc.calculate(pos, None)           # open boundaries: another answer, graph dropped and rebuilt
    assert not calc.model.representation_model.distance.use_periodic
    assert float((e_open - e).abs().max()) > 1e-3
    for k in range(3):
        e2, f2 = calc.calculate(pos, None)
        assert torch.equal(e2, e_open) and torch.equal(f2, f_open)
    e3, f3 = calc.calculate(pos, box.unsqueeze(0))       # [1,3,3], back to periodic
    check(e3, f3.view(n, 3), g["y"], g["forces"])


def test_atomref_prior_matches_reference(tmp_path):
    """TensorNet + Atomref prior against the fixture written by the unmodified reference; through TorchMD_Net.forward and
    through External (legacy checkpoint keys included)."""
    from torchmdnet.calculators import External
    g, m = _small("small_tensornet_atomref.npz")
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    y, f = m(z, pos, batch)
    check(y, f, g["y"], g["forces"])
    m.prior_model = None                      # the prior is what separates the two results
    y0, _ = m(z, pos, batch)
    assert float((y0 - y).abs().min()) > 0.1
    n0 = int((tt(g["batch"]) == 0).sum())
    legacy = {"model." + k.replace("prior_model.0.", "prior_model."): v for k, v in g["sd"].items()}
    path = str(tmp_path / "legacy.ckpt")
    torch.save({"hyper_parameters": dict(g["hparams"]), "state_dict": legacy}, path)
    calc = External(path, tt(g["z"])[:n0].view(1, n0), "cuda")
    for _ in range(4):
        e, frc = calc.calculate(tt(g["pos"])[:n0], None)
    g1, m1 = _small("small_tensornet_atomref.npz", mode="exact")
    y1, f1 = m1(z[:n0], pos[:n0])
    torch.testing.assert_close(e, y1, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(frc.view(n0, 3), f1, rtol=0, atol=1e-5)


def test_pairwise_priors_match_reference():
    """TensorNet + ZBL + Coulomb against the fixture written by the unmodified reference (priors/zbl.py, priors/coulomb.py;
    forces from its autograd pass).  The prior part is two orders of magnitude larger than the network's output in this
    fixture, so the check is on the priors; energies 1e-5 relative, forces 1e-4 absolute + 2e-6 relative (|F| up to 44)."""
    g, m = _small("small_tensornet_pairpriors.npz")
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    q = tt(g["partial_charges"]).cuda()
    y, f = m(z, pos, batch, extra_args={"partial_charges": q})
    y_ref, f_ref = torch.as_tensor(g["y"]).double(), torch.as_tensor(g["forces"]).double()
    assert float(((y.cpu().double() - y_ref).abs() / y_ref.abs().clamp_min(1.0)).max()) <= 1e-5
    assert float(((f.cpu().double() - f_ref).abs() - 2e-6 * f_ref.abs()).max()) <= 1e-4
    # the priors' share, against the reference's run without them
    m.prior_model = None
    y0, f0 = m(z, pos, batch)
    check(y0, f0, g["y_noprior"], g["forces_noprior"])
    dy = (y - y0).cpu().double()
    dy_ref = torch.as_tensor(g["y"]).double() - torch.as_tensor(g["y_noprior"]).double()
    assert float(((dy - dy_ref).abs() / dy_ref.abs().clamp_min(1.0)).max()) <= 1e-5
    # Coulomb without charges is an error, as in the reference (KeyError on extra_args)
    g2, m2 = _small("small_tensornet_pairpriors.npz")
    with pytest.raises((KeyError, TypeError)):
        m2(z, pos, batch)
    # too small a neighbour budget is caught (models/utils.py:419-425 asserts)
    m2.prior_model[0].max_num_neighbors = 1
    with pytest.raises(AssertionError, match="max_num_neighbors"):
        m2(z, pos, batch, extra_args={"partial_charges": q})


# known answers of the reference's own test (tests/test_prior_d2.py:33-62; computed with Psi4, TPSS-D2, Hartree)
D2_CASES = {
    "he": ([2], [[0.0, 0.0, 0.0]], [0], [0.0]),
    "h2": ([1, 1], [[0.0, 0.0, 0.0], [1.0, 0.0, 0.0]], [0, 0], [-2.396695338419219e-06]),
    "h2o_n2_ch4": ([1, 8, 1, 7, 7, 6, 1, 1, 1],
                   [[0.0, 0.0, 0.0], [0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [1.0, 2.0, 3.0], [2.0, 1.0, 3.0], [0.0, 0.0, 0.0],
                    [1.0, 1.0, 0.0], [0.0, 1.0, 1.0], [1.0, 0.0, 1.0]], [0, 0, 0, 1, 1, 2, 2, 2, 2],
                   [-4.7815923481300465e-06, -3.0071055390390565e-06, -7.100283425158895e-05]),
}


@pytest.mark.parametrize("name", list(D2_CASES))
def test_d2_known_answers(name):
    """D2.post_reduce against the Psi4 energies the reference pins (tests/test_prior_d2.py:65-85, same constructor call).
    The reference compares with assert_allclose's fp32 defaults (rtol 1.3e-6, atol 1e-5 -- looser than the values); here
    1e-5 relative, no absolute slack."""
    from torchmdnet.priors import D2
    z, pos, batch, y_ref = D2_CASES[name]
    prior = D2(cutoff_distance=10.0, max_num_neighbors=128, atomic_number=list(range(100)), distance_scale=1e-10,
               energy_scale=4.35974e-18).cuda()
    y_ref = torch.tensor(y_ref)
    y = prior.post_reduce(torch.zeros_like(y_ref).cuda(), torch.tensor(z).cuda(), torch.tensor(pos).cuda(),
                          torch.tensor(batch).cuda(), {})
    torch.testing.assert_close(y.cpu(), y_ref, rtol=1e-5, atol=0)


def test_d2_prior_matches_reference():
    """TensorNet + D2 against the fixture written by the unmodified reference (priors/d2.py; forces from its autograd pass),
    and the prior alone against the reference's fp64 evaluation: energies 1e-5 relative, forces 1e-6 absolute (|F| <= 0.09)."""
    g, m = _small("small_tensornet_d2.npz")
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    y, f = m(z, pos, batch)
    check(y, f, g["y"], g["forces"])
    prior = m.prior_model[0]
    n_mol = int(batch.max()) + 1
    e = torch.zeros(n_mol, 1, device="cuda")
    frc = torch.zeros_like(pos)
    prior.add_energy_forces(e, frc, z, pos, batch, None)
    e_ref, f_ref = torch.as_tensor(g["y_prior_f64"]), torch.as_tensor(g["forces_prior_f64"])
    assert float(((e.cpu().double() - e_ref).abs() / e_ref.abs()).max()) <= 1e-5
    assert float((frc.cpu().double() - f_ref).abs().max()) <= 1e-6
    torch.testing.assert_close(prior.post_reduce(torch.ones(n_mol, 1, device="cuda"), z, pos, batch, {}), e + 1.0)
    # without the prior: the reference's run without it
    m.prior_model = None
    y0, f0 = m(z, pos, batch)
    check(y0, f0, g["y_noprior"], g["forces_noprior"])
    # too small a neighbour budget is caught (models/utils.py:419-425 asserts)
    prior.max_num_neighbors = 1
    with pytest.raises(AssertionError, match="max_num_neighbors"):
        prior.post_reduce(torch.zeros(n_mol, 1, device="cuda"), z, pos, batch, {})


@pytest.mark.parametrize("name", ["small_tensornet_atomref.npz", "small_tensornet_pairpriors.npz", "small_tensornet_d2.npz"])
def test_scripted_model_with_priors(name, tmp_path):
    """torch.jit.script of a model WITH priors (the reference's TorchMD_Net is scripted with its prior_model ModuleList,
    models/model.py:266-284): the priors' tables travel in the scripted module, the result equals the eager model's bit
    for bit, before and after save / load, and matches the reference-written fixture."""
    g, m = _small(name)
    z, pos, batch = tt(g["z"]).cuda(), tt(g["pos"]).cuda(), tt(g["batch"]).cuda()
    extra = {"partial_charges": tt(g["partial_charges"]).cuda()} if "partial_charges" in g else None
    y0, f0 = m(z, pos, batch, extra_args=extra)
    sm = torch.jit.script(m)
    y1, f1 = sm(z, pos, batch, None, None, extra)
    assert torch.equal(y0, y1) and torch.equal(f0, f1)
    path = str(tmp_path / "model_with_priors.pt")
    torch.jit.save(sm, path)
    lm = torch.jit.load(path, map_location="cuda")
    y2, f2 = lm(z, pos, batch, None, None, extra)
    assert torch.equal(y0, y2) and torch.equal(f0, f2)
    y_ref, f_ref = torch.as_tensor(g["y"]).double(), torch.as_tensor(g["forces"]).double()
    assert float(((y2.cpu().double() - y_ref).abs() / y_ref.abs().clamp_min(1.0)).max()) <= 1e-5
    assert float(((f2.cpu().double() - f_ref).abs() - 2e-6 * f_ref.abs()).max()) <= 1e-4
    if extra is not None:            # Coulomb without its charges: an assertion inside the scripted module
        with pytest.raises(Exception, match="partial_charges"):
            lm(z, pos, batch, None, None, None)


@pytest.mark.parametrize("kind", ["zbl", "coulomb", "d2"])
def test_pair_prior_kernels_against_the_oracle(kind):
    """tmdn_pair_prior_f32 against oracle/priors.py (numpy fp64, finite-difference forces; pinned to the reference's fixtures
    in tests/test_oracle_pinned.py) on random molecules and other parameters than the fixtures use: several molecules of
    different size, atoms closer than the fixtures' (down to 0.7), Angstrom / eV and nm / kJ-mol-like unit scales.
    Energies 2e-6 relative, forces 2e-5 of the largest force."""
    import numpy as np
    from oracle import priors as P
    from torchmdnet.priors import D2, ZBL, Coulomb
    rng = np.random.RandomState({"zbl": 1, "coulomb": 2, "d2": 3}[kind])
    sizes = [1, 17, 2, 30, 9]
    batch = np.repeat(np.arange(len(sizes)), sizes)
    n, nm = batch.size, len(sizes)
    for trial, (dscale, escale, box) in enumerate([(1e-10, 1.602176634e-19, 4.0), (1e-9, 1.66054e-21, 0.45)]):
        pos = np.zeros((n, 3))
        for i in range(n):                      # random positions with a minimum distance of 0.7 (0.07 in nm)
            while True:
                p = rng.rand(3) * box * 2
                same = pos[:i][batch[:i] == batch[i]]
                if len(same) == 0 or np.linalg.norm(same - p, axis=1).min() > 0.175 * box:
                    pos[i] = p
                    break
        z = rng.randint(1, 10, n)
        if kind == "zbl":
            cut = 0.9 * box
            prior = ZBL(cutoff_distance=cut, max_num_neighbors=40, atomic_number=list(range(10)), distance_scale=dscale,
                        energy_scale=escale)
            extra = None
            fn = lambda p: P.zbl_energy(p, batch, z, nm, list(range(10)), cut, dscale, escale)      # noqa: E731
        elif kind == "coulomb":
            q = rng.rand(n) - 0.5
            alpha = 1.3 / box
            prior = Coulomb(alpha=alpha, max_num_neighbors=40, distance_scale=dscale, energy_scale=escale)
            extra = {"partial_charges": torch.from_numpy(q).float().cuda()}
            fn = lambda p: P.coulomb_energy(p, batch, q.astype(np.float32), nm, alpha, dscale, escale)      # noqa: E731
        else:
            cut = 1.5 * box
            prior = D2(cutoff_distance=cut, max_num_neighbors=40, atomic_number=list(range(10)), distance_scale=dscale,
                       energy_scale=escale)
            extra = None
            fn = lambda p: P.d2_energy(p, batch, z, nm, list(range(10)), cut, dscale, escale)      # noqa: E731
        prior = prior.cuda()
        pos32 = pos.astype(np.float32)
        e = torch.zeros(nm, 1, device="cuda")
        f = torch.zeros(n, 3, device="cuda")
        prior.add_energy_forces(e, f, torch.from_numpy(z).cuda(), torch.from_numpy(pos32).cuda(),
                                torch.from_numpy(batch).cuda(), extra)
        e_ref = fn(pos32.astype(np.float64))
        f_ref = P.forces_fd(fn, pos32.astype(np.float64), h=1e-5 * box)
        assert np.abs(e.cpu().numpy().ravel().astype(np.float64) - e_ref).max() <= 2e-6 * np.abs(e_ref).max(), (kind, trial)
        assert np.abs(f.cpu().numpy().astype(np.float64) - f_ref).max() <= 2e-5 * np.abs(f_ref).max(), (kind, trial)
        assert e_ref[0] == 0.0 and float(e[0]) == 0.0 and float(f[0].abs().max()) == 0.0      # the single atom


def test_zbl_and_coulomb_known_answers():
    """The reference's own checks of the two priors (tests/test_priors.py:38-93): four atoms, energies against the pair
    formulas written out in the test (Bohr / kJ-mol for ZBL, nm / kJ-mol for Coulomb).  Tolerance 1e-5 relative: the
    test's Coulomb constant 138.935 is the six-digit rounding of the 138.935458 that ZBL / Coulomb use (2.30707755e-28 J m
    converted), a 3.3e-6 difference in itself."""
    from torchmdnet.priors import ZBL, Coulomb
    pos = torch.tensor([[1.0, 0.0, 0.0], [2.5, 0.0, 0.0], [1.0, 1.0, 0.0], [0.0, 0.0, -1.0]])
    types = torch.tensor([0, 1, 2, 1])
    atomic_number = torch.tensor([1, 6, 8], dtype=torch.int8)
    zbl = ZBL(10.0, 5, atomic_number, distance_scale=5.29177210903e-11, energy_scale=1000.0 / 6.02214076e23).cuda()
    energy = zbl.post_reduce(torch.zeros(1, device="cuda"), types.cuda(), pos.cuda(), torch.zeros_like(types).cuda(), {})[0]

    def zbl_pair(p1, p2, z1, z2):
        r = torch.sqrt(torch.dot(p1 - p2, p1 - p2))
        x = r / (0.8854 / (z1 ** 0.23 + z2 ** 0.23))
        phi = (0.1818 * torch.exp(-3.2 * x) + 0.5099 * torch.exp(-0.9423 * x) + 0.2802 * torch.exp(-0.4029 * x)
               + 0.02817 * torch.exp(-0.2016 * x))
        return 0.5 * (torch.cos(r * torch.pi / 10.0) + 1.0) * phi * (138.935 / 5.29177210903e-2) * z1 * z2 / r

    expected = sum(zbl_pair(pos[i], pos[j], atomic_number[types[i]].float(), atomic_number[types[j]].float())
                   for i in range(4) for j in range(i))
    torch.testing.assert_close(energy.cpu(), expected, rtol=1e-5, atol=1e-5)

    pos = torch.tensor([[0.5, 0.0, 0.0], [1.5, 0.0, 0.0], [0.8, 0.8, 0.0], [0.0, 0.0, -0.4]])
    charge = torch.tensor([0.2, -0.1, 0.8, -0.9])
    alpha = 1.8
    coulomb = Coulomb(alpha, 5, distance_scale=1e-9, energy_scale=1000.0 / 6.02214076e23).cuda()
    energy = coulomb.post_reduce(torch.zeros(1, device="cuda"), types.cuda(), pos.cuda(), torch.zeros_like(types).cuda(),
                                 {"partial_charges": charge.cuda()})[0]

    def coulomb_pair(p1, p2, q1, q2):
        r = torch.sqrt(torch.dot(p1 - p2, p1 - p2))
        return torch.erf(alpha * r) * 138.935 * q1 * q2 / r

    expected = sum(coulomb_pair(pos[i], pos[j], charge[i], charge[j]) for i in range(4) for j in range(i))
    torch.testing.assert_close(energy.cpu(), expected, rtol=1e-5, atol=1e-5)
