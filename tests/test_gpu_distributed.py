"""Spatial decomposition of a periodic TensorNet evaluation: (i) all ranks emulated in one process on one GPU (stage
generators in lock step, halo points served by device copies) against the single-domain evaluation; (ii) the real
thing, one process per GPU over NCCL, when the box has >= 2 GPUs (tests/dist_water_check.py under torchrun)."""
import os
import subprocess
import sys

import pytest
import torch

import synth_inputs as S
from helpers import load_npz, tt

pytestmark = pytest.mark.gpu


def _model_and_system(n_side=8):
    from torchmdnet.models.model import create_model
    from torchmdnet.models.utils import OptimizedDistance
    g = load_npz("small_tensornet_periodic.npz")
    m = create_model(g["hparams"])
    m.load_state_dict(g["sd"])
    m = m.cuda().eval()
    m.representation_model.filter_mode = "table"
    z, pos, batch, box = S.water_box(n_side, seed=4)
    hp = g["hparams"]
    m.representation_model.distance = OptimizedDistance(
        hp["cutoff_lower"], hp["cutoff_upper"], max_num_pairs=-hp["max_num_neighbors"], return_vecs=True, loop=True,
        check_errors=False, resize_to_fit=False, box=torch.from_numpy(box), strategy="cell")
    return m, hp, torch.from_numpy(z), torch.from_numpy(pos), torch.from_numpy(box)


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_emulated_decomposition_matches_single_domain(world):
    from torchmdnet.distributed import DecomposedTensorNet, emulate_step, slab_plans
    m, hp, z, pos, box = _model_and_system(8)          # 1536 atoms, L = 24.9 A
    y_ref, f_ref = m(z.cuda(), pos.cuda())
    plans = slab_plans(pos, box, hp["cutoff_upper"], world, skin=0.25)
    runners = [DecomposedTensorNet(m, plans, r, z, box, "cuda", halo=lambda buf: None) for r in range(world)]
    for step in range(2):                                # second step: atoms moved (< skin), same plan
        p = pos.cuda() if step == 0 else torch.from_numpy(S.md_step_noise(pos.numpy(), 1, sigma=0.03)).cuda()
        y_ref, f_ref = m(z.cuda(), p)
        e, f = emulate_step(runners, p)
        # the ranks are below the 1024-row threshold of the tensor-core MLP GEMMs (FFMA kernel), the single domain is
        # above it: fp32-level differences per atom, summed over N atoms (same 3e-8 N allowance as the parity tests)
        torch.testing.assert_close(e, y_ref, rtol=2e-6, atol=2e-5 + 3e-8 * pos.shape[0])
        torch.testing.assert_close(f, f_ref, rtol=0, atol=2e-5)
    assert sum(r.plan.n_own for r in runners) == pos.shape[0]
    assert all(int(r.status[1]) == 0 for r in runners)


def test_mlp_gemm_modes_agree():
    """gemm_mode="tf32x3" (default: MLP GEMMs on tcgen05, K chunks summed in fp32) against "fp32" (FFMA kernel) on a
    system above the 1024-row threshold of tc_gemm: no one-sided drift of the summed energy."""
    m, hp, z, pos, box = _model_and_system(8)          # 1536 atoms
    y1, f1 = m(z.cuda(), pos.cuda())
    m.representation_model.gemm_mode = "fp32"
    y0, f0 = m(z.cuda(), pos.cuda())
    assert not torch.equal(y0, y1)
    assert float((y1 - y0).abs().max()) <= 1e-5 * max(1.0, float(y0.abs().max())) + 3e-8 * pos.shape[0]
    assert float((f1 - f0).abs().max()) <= 2e-5
    m.representation_model.gemm_mode = "tf32x3"
    y2, _ = m(z.cuda(), pos.cuda())
    assert torch.equal(y1, y2)


def test_stage_api_equals_monolithic_call():
    """world = 1 through the staged C ABI gives bit-identical results to tmdn_tensornet_energy_forces."""
    from torchmdnet.distributed import DecomposedTensorNet, emulate_step, slab_plans
    m, hp, z, pos, box = _model_and_system(6)
    m.mean.fill_(0.0)
    m._scale = None
    y_ref, f_ref = m(z.cuda(), pos.cuda())
    plans = slab_plans(pos, box, hp["cutoff_upper"], 1)
    r = DecomposedTensorNet(m, plans, 0, z, box, "cuda", halo=lambda buf: None)
    e, f = emulate_step([r], pos.cuda())
    assert torch.equal(f, f_ref) and torch.equal(e, y_ref)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_two_processes_over_nccl():
    here = os.path.dirname(os.path.abspath(__file__))
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(here, "dist_water_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "DIST_OK" in out.stdout
