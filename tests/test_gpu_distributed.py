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


@pytest.mark.parametrize("grid", [(1, 1, 1), (2, 1, 1), (3, 1, 1), (4, 1, 1), (2, 2, 1), (2, 2, 2)])
def test_emulated_decomposition_matches_single_domain(grid):
    from torchmdnet.distributed import DecomposedTensorNet, domain_plans, emulate_step
    m, hp, z, pos, box = _model_and_system(10)         # 3000 atoms, L = 31.1 A
    world = grid[0] * grid[1] * grid[2]
    plans = domain_plans(pos, box, hp["cutoff_upper"], world, skin=0.25, grid=grid)
    runners = [DecomposedTensorNet(m, plans, r, z, box, "cuda", halo="emulated") for r in range(world)]
    for step in range(2):                                # second step: atoms moved (< skin), same plan
        p = pos.cuda() if step == 0 else torch.from_numpy(S.md_step_noise(pos.numpy(), 1, sigma=0.03)).cuda()
        y_ref, f_ref = m(z.cuda(), p)
        e, f = emulate_step(runners, p)
        # the ranks are below the 1024-row threshold of the tensor-core MLP GEMMs (FFMA kernel), the single domain is
        # above it: fp32-level differences per atom, summed over N atoms (same 3e-8 N allowance as the parity tests)
        torch.testing.assert_close(e, y_ref, rtol=2e-6, atol=2e-5 + 3e-8 * pos.shape[0])
        torch.testing.assert_close(f, f_ref, rtol=0, atol=2e-5)
    assert sum(r.plan.n_own for r in runners) == pos.shape[0]
    assert all(r.status[1:].tolist() == [0, 0, 0] for r in runners)


def test_row_ranges_do_not_change_the_result():
    """tmdn_tensornet_stage_rows: a stage evaluated as [0, k) then [k, n) -- what the overlapped step does with its
    boundary and interior rows -- against the whole range at once.  The edge kernels are bit-identical by construction;
    the node GEMMs may select another kernel for a shorter range, hence the fp32-level tolerance."""
    from torchmdnet.distributed import DecomposedTensorNet, domain_plans, emulate_step
    m, hp, z, pos, box = _model_and_system(8)
    plans = domain_plans(pos, box, hp["cutoff_upper"], 1)
    whole = DecomposedTensorNet(m, plans, 0, z, box, "cuda", halo="emulated")
    e0, f0 = emulate_step([whole], pos.cuda())
    e0, f0 = e0.clone(), f0.clone()
    for k in (1, 500, 1535):
        plans[0].n_boundary = k                          # world 1 has no boundary: force a split
        part = DecomposedTensorNet(m, plans, 0, z, box, "cuda", halo="emulated")
        e1, f1 = emulate_step([part], pos.cuda())
        torch.testing.assert_close(e1, e0, rtol=2e-6, atol=2e-5)
        torch.testing.assert_close(f1, f0, rtol=0, atol=2e-5)
    plans[0].n_boundary = 0


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
    r = DecomposedTensorNet(m, plans, 0, z, box, "cuda", halo="emulated")
    e, f = emulate_step([r], pos.cuda())
    assert torch.equal(f, f_ref) and torch.equal(e, y_ref)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("transport", ["nccl", "peer"])
def test_processes_over_nvlink(transport):
    """One process per GPU: NCCL send/recv and the fused pack+push kernel over peer memory, overlapped with the interior
    rows; rank 0 compares with its own single-domain evaluation (tests/dist_water_check.py)."""
    here = os.path.dirname(os.path.abspath(__file__))
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(here, "dist_water_check.py"), transport]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "DIST_OK" in out.stdout
