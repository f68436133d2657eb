"""Host-side logic that needs no GPU: seed-for-seed parameter initialisation against the reference
(checksums recorded by oracle/make_golden.py), state_dict compatibility, C-ABI exports, loud failure on CPU."""
import ctypes
import json
import os
import random
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import GOLDEN, load_npz, tt

CONFIGS = ["TensorNet-SPICE", "TensorNet-rMD17", "TensorNet-ANI1X", "TensorNet-QM9", "ET-QM9", "ET-MD17"]


def seed_everything(s):
    random.seed(s)
    np.random.seed(s)
    torch.manual_seed(s)


@pytest.mark.parametrize("cfg", CONFIGS)
def test_seeded_init_matches_reference(cfg):
    from torchmdnet.models.model import create_model
    g = load_npz("model_%s.npz" % cfg)
    seed_everything(int(g["seed"]))
    sd = create_model(g["hparams"]).state_dict()
    cs = g["checksums"]
    assert list(sd) == list(cs)          # same keys, same order as the reference's state_dict
    for k, v in sd.items():
        assert float(v.double().sum()) == pytest.approx(cs[k][0], abs=1e-9), k
        assert float(v.double().abs().sum()) == pytest.approx(cs[k][1], abs=1e-9), k


@pytest.mark.parametrize("name", ["small_tensornet_o3.npz", "small_et_both.npz", "small_et_keys_nonbr.npz"])
def test_reference_state_dict_loads_strictly(name):
    from torchmdnet.models.model import create_model
    g = load_npz(name)
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"], strict=True)


def test_expected_pkl_recipe_init(tmp_path):
    """The reference's own golden recipe (tests/test_model.py:137-188): seed 1234 then create_model."""
    from torchmdnet.models.model import create_model
    with open(os.path.join(GOLDEN, "expected_pkl.json")) as f:
        exp = json.load(f)
    for key in ("tensornet", "equivariant-transformer"):
        e = exp[key]
        seed_everything(1234)
        sd = create_model(e["hparams"]).state_dict()
        for k, v in sd.items():
            assert float(v.double().sum()) == pytest.approx(e["checksums"][k][0], abs=1e-9), k


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "tmdn.h")) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(tmdn_[a-z0-9_]+)\s*\(", src)))


def test_cabi_exports_every_declared_symbol():
    from torchmdnet import _native
    lib = _native.cabi()
    names = _declared_symbols()
    assert "tmdn_neighbors_f32" in names and "tmdn_tensornet_energy_forces" in names
    for n in names:
        assert hasattr(lib, n), "libtmdn.so does not export %s" % n
    lib.tmdn_version.restype = ctypes.c_int
    assert lib.tmdn_version() == 100
    lib.tmdn_neighbors_workspace_bytes.restype = ctypes.c_size_t
    lib.tmdn_neighbors_workspace_bytes.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_double,
                                                   ctypes.c_int]
    assert lib.tmdn_neighbors_workspace_bytes(0, 1000, None, 5.0, 4) > 0


def test_cabi_rejects_bad_arguments_without_a_gpu():
    from torchmdnet import _native
    lib = _native.cabi()
    lib.tmdn_last_error.restype = ctypes.c_char_p
    rc = lib.tmdn_neighbors_f32(0, None, None, ctypes.c_int64(0), None, 0, ctypes.c_double(0), ctypes.c_double(5),
                                0, 1, ctypes.c_int64(10), None, None, None, None, None, ctypes.c_size_t(0), None)
    assert rc == 1 and b"null pointer" in lib.tmdn_last_error()


def test_cabi_tuning_knobs_without_a_gpu():
    """tmdn_set_tuning (include/tmdn.h): query with a negative value, set, restore; unknown keys are an error."""
    from torchmdnet import _native
    lib = _native.cabi()
    lib.tmdn_last_error.restype = ctypes.c_char_p
    lib.tmdn_set_tuning.restype = ctypes.c_int
    lib.tmdn_set_tuning.argtypes = [ctypes.c_char_p, ctypes.c_int]
    defaults = {b"small_gemm": 1, b"lpt_rows": 1, b"pdl": 0, b"edge_persist": 1, b"small_csr": 1}
    for key, dflt in defaults.items():
        assert lib.tmdn_set_tuning(key, -1) == dflt, key
        assert lib.tmdn_set_tuning(key, 1 - dflt) == dflt
        assert lib.tmdn_set_tuning(key, -1) == 1 - dflt
        assert lib.tmdn_set_tuning(key, dflt) == 1 - dflt
    assert lib.tmdn_set_tuning(b"no_such_knob", 1) == -1 and b"no_such_knob" in lib.tmdn_last_error()
    assert lib.tmdn_set_tuning(None, 1) == -1


def test_cpu_tensors_fail_loudly():
    from torchmdnet.models.model import create_model
    from torchmdnet.models.utils import OptimizedDistance
    with pytest.raises(RuntimeError, match="CUDA"):
        OptimizedDistance(0.0, 5.0)(torch.randn(10, 3))
    g = load_npz("small_tensornet_o3.npz")
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    with pytest.raises(RuntimeError, match="CUDA"):
        m(tt(g["z"]), tt(g["pos"]), tt(g["batch"]))


def test_unsupported_configurations_are_refused():
    from torchmdnet.models.model import create_model
    hp = dict(load_npz("small_tensornet_o3.npz")["hparams"])
    with pytest.raises(NotImplementedError):
        create_model(dict(hp, model="graph-network", aggr="mean", neighbor_embedding=True))
    with pytest.raises(NotImplementedError):
        create_model(dict(hp, output_model="DipoleMoment"))
    with pytest.raises(ValueError):
        create_model(dict(hp, model="nope"))


def test_rbf_modules_match_oracle():
    from oracle import model as om
    from torchmdnet.models.utils import CosineCutoff, ExpNormalSmearing, GaussianSmearing
    d = torch.linspace(0, 6, 50)
    e = ExpNormalSmearing(0.0, 5.0, 16, trainable=False)
    torch.testing.assert_close(e(d), om.expnorm_rbf(d, e.means, e.betas, 0.0, 5.0))
    gs = GaussianSmearing(0.0, 5.0, 16, trainable=False)
    torch.testing.assert_close(gs(d), om.gauss_rbf(d, gs.offset, gs.coeff))
    torch.testing.assert_close(CosineCutoff(0.0, 5.0)(d), om.cosine_cutoff(d, 0.0, 5.0))
    torch.testing.assert_close(CosineCutoff(1.0, 5.0)(d), om.cosine_cutoff(d, 1.0, 5.0))
    # decay outside the cutoffs, as tests/test_rbfs.py:6-30 of the reference
    assert float(e(torch.tensor([5.5])).abs().max()) == 0.0


@pytest.mark.parametrize("name,warp", [("small_tensornet_o3.npz", True), ("small_tensornet_o3.npz", False),
                                       ("small_tensornet_gauss.npz", True)])
def test_filter_table_interpolates_filter(name, warp):
    """The Hermite tables the host builds for filter_mode='table' (TableGrid: max(1024, 32*num_rbf) intervals uniform
    in t(d) = exp(-alpha d) - exp(-alpha r_c) + gamma (r_c - d); uniform in d for Gaussian smearing / warp=False with
    4096 intervals) reproduce the filter MLP and its derivative, evaluated the way the kernels do (tab_locate)."""
    from torchmdnet.models.model import create_model
    from torchmdnet.models.tensornet import TableGrid, filter_table
    g = load_npz(name)
    m = create_model(g["hparams"]).double()
    m.load_state_dict({k: v.double() if v.is_floating_point() else v for k, v in g["sd"].items()})
    rm = m.representation_model
    lo, hi = float(rm.cutoff_lower), float(rm.cutoff_upper)
    ls = rm.layers[0].linears_scalar
    layers = [(l.weight.detach(), l.bias.detach()) for l in ls]
    if rm.rbf_type == "expnorm":
        p0, p1 = rm.distance_expansion.means, rm.distance_expansion.betas
    else:
        p0, p1 = rm.distance_expansion.offset, rm.distance_expansion.coeff.expand(rm.num_rbf)
    rm.table_warp = warp and rm.rbf_type == "expnorm"
    n_int = rm.resolved_table_intervals()
    assert n_int == (1024 if rm.table_warp else 4096)
    grid = TableGrid(n_int, lo, hi, rm.rbf_type, warp)
    assert (grid.alpha > 0) == (warp and rm.rbf_type == "expnorm")
    nodes = grid.nodes()
    assert float(nodes[0]) == hi and float(nodes[-1]) == 0.0 and bool((nodes[1:] < nodes[:-1]).all())
    torch.testing.assert_close(grid.t(nodes) / grid.step, torch.arange(n_int + 1, dtype=torch.float64), rtol=0, atol=1e-9)
    f, df = filter_table(nodes, rm.rbf_type, p0, p1, lo, hi, layers, "silu", True)
    c = grid.coefficients(nodes, f, df)
    d = torch.cat([torch.zeros(1, dtype=torch.float64), 0.5 + torch.rand(4000, dtype=torch.float64) * (hi * 0.999 - 0.5)])
    k, s, ds = grid.locate(d)
    s, ck = s[:, None], c[k]
    approx = ck[..., 0] + s * (ck[..., 1] + s * (ck[..., 2] + s * ck[..., 3]))
    dapprox = (ck[..., 1] + s * (2 * ck[..., 2] + 3 * s * ck[..., 3])) * ds[:, None]
    exact, dexact = filter_table(d, rm.rbf_type, p0, p1, lo, hi, layers, "silu", True)
    assert float((approx - exact).abs().max()) < 2e-8 * max(1.0, float(exact.abs().max()))
    assert float((dapprox - dexact).abs().max()) < 2e-6 * max(1.0, float(dexact.abs().max()))


def test_deepcopy_and_pickle_do_not_share_the_native_handle():
    """A copied model must not carry the original's integer handle / workspace (its first call would destroy the
    original's packed weights): FusedRepresentation.__getstate__ resets them."""
    import copy
    import pickle
    from torchmdnet.models.model import create_model
    g = load_npz("small_tensornet_o3.npz")
    m = create_model(g["hparams"])
    rm = m.representation_model
    rm._handle, rm._pack_key, rm._workspace = 7, ("k",), torch.zeros(4)
    try:
        for c in (copy.deepcopy(m), pickle.loads(pickle.dumps(m))):
            crm = c.representation_model
            assert crm._handle is None and crm._pack_key is None and crm._workspace is None
            assert crm.table_warp is True and crm.filter_mode == rm.filter_mode
            assert all(torch.equal(a, b) for a, b in zip(c.state_dict().values(), m.state_dict().values()))
        assert rm._handle == 7
    finally:
        rm._handle = rm._pack_key = rm._workspace = None


@pytest.mark.parametrize("name", ["small_tensornet_o3.npz", "small_et_both.npz"])
def test_models_are_scriptable_and_saveable(name, tmp_path):
    """torch.jit.script(model) (reference: tests/test_model.py:40-58) compiles the stand-in of models/scripted.py: the
    packed weights travel as one buffer, survive save -> load, and the graph calls the fused op.  No GPU needed to
    script; running the scripted model on CPU tensors is refused by the op (no CPU fallback)."""
    from torchmdnet.models.model import create_model
    g = load_npz(name)
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    if g["hparams"]["model"] == "tensornet":
        m.representation_model.filter_mode = "table"
    sm = torch.jit.script(m)
    op = "tensornet_energy_forces_packed" if g["hparams"]["model"] == "tensornet" else "et_energy_forces_packed"
    assert op in str(sm.graph) or op in str(sm.code)
    path = str(tmp_path / "model.pt")
    torch.jit.save(sm, path)
    lm = torch.jit.load(path)
    assert torch.equal(lm.packed, sm.packed) and lm.layout == sm.layout and lm.iparams == sm.iparams
    assert lm.packed.numel() > 1000 and all(o % 64 == 0 for o in sm.layout[:len(sm.layout) // 2])
    with pytest.raises(RuntimeError, match="CUDA"):
        lm(tt(g["z"]), tt(g["pos"]), tt(g["batch"]))


def test_external_box_routing(tmp_path):
    """External.calculate(pos, box): a non-zero box selects periodic boundaries on the model's distance module (cell list
    for wide rectangular boxes, brute otherwise); None / zeros keep open boundaries; an unchanged box changes nothing."""
    from torchmdnet.calculators import External
    from torchmdnet.models.model import create_model
    g = load_npz("small_tensornet_o3.npz")
    m = create_model(g["hparams"])
    ckpt = str(tmp_path / "m.ckpt")
    torch.save({"hyper_parameters": dict(g["hparams"]), "state_dict": {"model." + k: v for k, v in m.state_dict().items()}}, ckpt)
    calc = External(ckpt, torch.ones(1, 9, dtype=torch.long), "cpu")
    rm = calc.model.representation_model
    rc = float(rm.cutoff_upper)
    assert calc._use_box(None) is False and calc._use_box(torch.zeros(3, 3)) is False and not rm.distance.use_periodic
    wide = torch.eye(3) * (4 * rc)
    assert calc._use_box(wide) is True
    assert rm.distance.use_periodic and rm.distance.strategy == "cell" and torch.equal(rm.distance.box, wide)
    assert calc._use_box(wide) is False and calc._use_box(wide.clone()) is False
    small = torch.eye(3) * (2.5 * rc)
    assert calc._use_box(small.unsqueeze(0)) is True and rm.distance.strategy == "brute"
    tric = wide.clone()
    tric[1, 0] = 1.0
    assert calc._use_box(tric) is True and rm.distance.strategy == "brute" and rm.distance.use_periodic
    assert calc._use_box(torch.tensor([4 * rc] * 3)) is True and rm.distance.strategy == "cell"
    assert calc._use_box(None) is True and not rm.distance.use_periodic
    assert rm.distance.max_num_pairs == -g["hparams"]["max_num_neighbors"] and rm.distance.loop


def test_atomref_prior_and_legacy_checkpoint_keys(tmp_path):
    """create_model / load_model with the Atomref prior (priors/atomref.py, models/model.py:96-98,146-177): the reference's
    state_dict keys load as they are, checkpoints from the time when Atomref was the only prior (un-indexed
    ``prior_model.*`` keys, model.py:134-141) are renamed, the prior adds sum_i atomref[z_i] per molecule; ZBL, Coulomb and D2 are constructible
    with the reference's arguments and an unknown prior is refused by name."""
    from torchmdnet.models.model import create_model, load_model
    g = load_npz("small_tensornet_atomref.npz")
    hp = g["hparams"]
    m = create_model(hp)
    assert set(m.state_dict()) == set(g["sd"]) and "prior_model.0.atomref.weight" in g["sd"]
    m.load_state_dict(g["sd"])
    assert m.prior_model[0].get_init_args() == {"max_z": hp["max_z"]}
    z, batch = tt(g["z"]), tt(g["batch"])
    add = m.apply_priors(torch.zeros(int(batch.max()) + 1, 1), z, batch, int(batch.max()) + 1)
    w = g["sd"]["prior_model.0.atomref.weight"][:, 0]
    expect = torch.zeros(int(batch.max()) + 1).index_add_(0, batch, w[z])
    torch.testing.assert_close(add[:, 0], expect)
    # legacy keys + the Lightning "model." prefix
    legacy = {}
    for k, v in g["sd"].items():
        legacy["model." + k.replace("prior_model.0.", "prior_model.")] = v
    path = str(tmp_path / "legacy.ckpt")
    torch.save({"hyper_parameters": dict(hp), "state_dict": legacy}, path)
    lm = load_model(path)
    assert all(torch.equal(a, b) for a, b in zip(lm.state_dict().values(), m.state_dict().values()))
    # hparams overrides through kwargs (model.py:126-129): unknown keys warn, known keys replace
    with pytest.warns(UserWarning, match="Unknown hyperparameter"):
        lm2 = load_model(path, derivative=False, some_new_flag=1)
    assert lm2.derivative is False
    # the forms the option can take (str | dict | list) and the priors outside this build
    from torchmdnet.models.model import create_prior_models
    assert create_prior_models({"prior_model": None}) is None
    assert create_prior_models({"prior_model": {"Atomref": {"max_z": 7}}})[0].initial_atomref.shape == (7, 1)
    assert len(create_prior_models({"prior_model": ["Atomref"], "prior_args": [{"max_z": 5}]})) == 1
    with pytest.raises(NotImplementedError, match="LearnableAtomref"):
        create_model(dict(hp, prior_model="LearnableAtomref", prior_args=None))
    # the pairwise priors keep the reference's constructor arguments and init-arg round trip (zbl.py:20-45, coulomb.py:17-36)
    zargs = dict(cutoff_distance=4.0, max_num_neighbors=50, atomic_number=list(range(10)), distance_scale=1e-10,
                 energy_scale=1.602e-19)
    pm = create_prior_models({"prior_model": [{"ZBL": zargs}, {"Coulomb": dict(alpha=1.8, max_num_neighbors=50,
                                                                              distance_scale=1e-10, energy_scale=1.602e-19)}]})
    assert pm[0].get_init_args() == zargs and list(pm[0].state_dict()) == ["atomic_number"]
    assert pm[1].get_init_args()["alpha"] == 1.8 and len(pm[1].state_dict()) == 0
    m3 = create_model(dict(hp, prior_model=[{"ZBL": zargs}], prior_args=None))
    assert "prior_model.0.atomic_number" in m3.state_dict()
    # D2 (d2.py:109-161): constructor, buffers and parameter table as the reference's (the fixture holds ITS buffers);
    # the list form of tests/priors.yaml with a dataset supplying the units (tests/test_priors.py:95-127)
    gd = load_npz("small_tensornet_d2.npz")
    dargs = gd["hparams"]["prior_model"][0]["D2"]
    d2 = create_prior_models({"prior_model": [{"D2": dargs}]})[0]
    assert d2.get_init_args() == dargs and list(d2.state_dict()) == ["Z_map", "C_6", "R_r"]
    for k in ("Z_map", "C_6", "R_r"):
        assert torch.equal(d2.state_dict()[k][1:], gd["sd"]["prior_model.0." + k][1:]), k       # entry 0 is NaN
    assert torch.isnan(d2.C_6[0]) and d2.d == 20 and d2.s_6 == 1

    class _Dataset:
        atomic_number = list(range(20))
        distance_scale = 1e-10
        energy_scale = 4.35974e-18

        def get_atomref(self):
            return torch.arange(20.0)

    pm = create_prior_models({"prior_model": [{"ZBL": {"cutoff_distance": 4.0, "max_num_neighbors": 50}},
                                              {"D2": {"cutoff_distance": 10.0, "max_num_neighbors": 100}}, "Atomref"]},
                             dataset=_Dataset())
    assert [type(p).__name__ for p in pm] == ["ZBL", "D2", "Atomref"]
    assert pm[1].cutoff_distance == 10.0 and pm[1].max_num_neighbors == 100 and pm[1].energy_scale == 4.35974e-18


@pytest.mark.parametrize("name,kinds", [("small_tensornet_atomref.npz", [0]), ("small_tensornet_pairpriors.npz", [1, 2]),
                                        ("small_tensornet_d2.npz", [3])])
def test_models_with_priors_script_on_the_cpu(name, kinds, tmp_path):
    """torch.jit.script of a model with priors needs no GPU (compute does): the priors' tables and scalars are packed into the
    scripted module (models/scripted.py) and survive save / load."""
    from torchmdnet.models.model import create_model
    g = load_npz(name)
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    sm = torch.jit.script(m.eval())
    assert list(sm.prior_kind) == kinds and len(sm.prior_par) == 5 * len(kinds)
    path = str(tmp_path / "m.pt")
    torch.jit.save(sm, path)
    lm = torch.jit.load(path)
    assert list(lm.prior_kind) == kinds
    torch.testing.assert_close(lm.prior_data, sm.prior_data, rtol=0, atol=0, equal_nan=True)   # D2: element 0 is NaN
    if kinds == [3]:        # D2: sqrt(C_6) and R_r per element type, back to back
        n = sm.prior_len[0]
        pr = m.prior_model[0]
        torch.testing.assert_close(sm.prior_data[1:n] ** 2, pr.C_6[pr.Z_map][1:].float())
        torch.testing.assert_close(sm.prior_data[n + 1:2 * n], pr.R_r[pr.Z_map][1:].float())
    with pytest.raises(Exception):          # evaluation still needs the CUDA extension and CUDA tensors
        lm(tt(g["z"]), tt(g["pos"]), tt(g["batch"]))


def test_reference_module_paths_resolve():
    """The import paths a user of the reference has in their scripts (torchmdnet/models/__init__.py, priors/{base,atomref,zbl,
    coulomb,d2}.py, models/utils.py) exist here under the same names."""
    import importlib
    from torchmdnet import models, priors
    assert models.__all__ == ["graph-network", "transformer", "equivariant-transformer", "tensornet"]
    assert priors.__all__ == ["Atomref", "D2", "ZBL", "Coulomb"]
    for mod, names in {"torchmdnet.priors.base": ["BasePrior"], "torchmdnet.priors.atomref": ["Atomref"],
                       "torchmdnet.priors.zbl": ["ZBL"], "torchmdnet.priors.coulomb": ["Coulomb"],
                       "torchmdnet.priors.d2": ["D2"],
                       "torchmdnet.models.utils": ["OptimizedDistance", "Distance", "GaussianSmearing", "ExpNormalSmearing",
                                                   "ShiftedSoftplus", "CosineCutoff", "NeighborEmbedding",
                                                   "GatedEquivariantBlock", "rbf_class_mapping", "act_class_mapping",
                                                   "dtype_mapping"],
                       "torchmdnet.models.model": ["create_model", "load_model", "create_prior_models", "TorchMD_Net"],
                       "torchmdnet.models.output_modules": ["Scalar", "EquivariantScalar"],
                       "torchmdnet.models.tensornet": ["TensorNet"], "torchmdnet.models.torchmd_et": ["TorchMD_ET"],
                       "torchmdnet.models.torchmd_gn": ["TorchMD_GN"], "torchmdnet.models.torchmd_t": ["TorchMD_T"],
                       "torchmdnet.calculators": ["External"],
                       "torchmdnet.neighbors": ["get_neighbor_pairs_kernel"]}.items():
        m = importlib.import_module(mod)
        for n in names:
            assert hasattr(m, n), (mod, n)
