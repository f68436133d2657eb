"""Grouped channel mix (tensornet.py:273-284, 386-408 linears_tensor) through the C ABI: tcgen05 path (H = 128,
three-product tf32 split) and FFMA path (other widths) against an fp64 einsum."""
import pytest
import torch

pytestmark = pytest.mark.gpu

GROUP = [0, 1, 1, 1, 2, 2, 2, 2, 2]


def reference(x, w):
    xd, wd = x.double().cpu(), w.double().cpu()
    return torch.stack([xd[:, c] @ wd[GROUP[c]].T for c in range(9)], 1)


@pytest.fixture(scope="module")
def ops():
    from torchmdnet import _native
    _native.load()
    return torch.ops.torchmdnet_b200


@pytest.fixture(params=[1, 0], ids=["small_gemm", "tiled"])
def small(ops, request):
    """Both kernel selections for small launches: the latency-oriented fp32 kernel (default) and the throughput
    kernels (tcgen05 for H = 128 / tiled FFMA) that large systems use."""
    old = ops.set_tuning("small_gemm", request.param)
    yield request.param
    ops.set_tuning("small_gemm", old)


@pytest.mark.parametrize("n", [1, 7, 25, 42, 43, 128, 166, 1000, 20011])
@pytest.mark.parametrize("H", [128, 64])
def test_channel_mix_matches_fp64(ops, small, n, H):
    g = torch.Generator().manual_seed(n * 7 + H)
    x = (torch.randn(n, 9, H, generator=g) * torch.logspace(-2, 1, 9).view(1, 9, 1)).cuda()
    w = (torch.randn(3, H, H, generator=g) / H ** 0.5).cuda()
    y = ops.channel_mix(x, w, None, False)
    ref = reference(x, w)
    bound = torch.stack([x.double().cpu()[:, c].abs() @ w.double().cpu()[GROUP[c]].abs().T for c in range(9)], 1)
    err = (y.double().cpu() - ref).abs()
    # fp32-level accuracy: 2e-6 of sum_k |x_k w_k| (fp32 FFMA itself is ~1e-6 of that bound at K = 128)
    assert (err <= 2e-6 * bound + 1e-30).all(), float((err / (bound + 1e-30)).max())


def test_channel_mix_accumulates(ops, small):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(333, 9, 128, generator=g).cuda()
    w = (torch.randn(3, 128, 128, generator=g) / 128 ** 0.5).cuda()
    y0 = torch.randn(333, 9, 128, generator=g).cuda()
    y = ops.channel_mix(x, w, y0.clone(), True)
    ref = reference(x, w) + y0.double().cpu()
    assert torch.allclose(y.double().cpu(), ref, rtol=0, atol=2e-5)


def test_channel_mix_is_repeatable(ops):
    g = torch.Generator().manual_seed(6)
    x = torch.randn(5000, 9, 128, generator=g).cuda()
    w = (torch.randn(3, 128, 128, generator=g) / 128 ** 0.5).cuda()
    a = ops.channel_mix(x, w, None, False)
    b = ops.channel_mix(x, w, None, False)
    assert torch.equal(a, b)


# ------------------------------------------------------------------ general tcgen05 GEMM (csrc/tc_gemm.cuh) through the ET model
def test_tc_gemm_against_ffma_through_et(monkeypatch, small):
    """The plain node GEMMs run on tensor cores by default; the same model evaluated with the fp32 FFMA kernels
    (TMDN_TC_GEMM=0 is read once per process, so the comparison is made against the oracle in test_gpu_et.py and,
    here, as a direct operator check on odd shapes through channel_mix's sibling entry point)."""
    ops_ = torch.ops.torchmdnet_b200
    if not hasattr(ops_, "node_gemm"):
        pytest.skip("node_gemm op not exported")
    g = torch.Generator().manual_seed(11)
    for (M, N, K) in [(1, 64, 64), (166, 384, 256), (2306, 640, 128), (5000, 128, 640), (777, 64, 128), (130, 36, 44),
                      (166, 128, 384), (166, 256, 128), (40, 100, 768), (33, 7, 772), (500, 384, 640)]:
        a = torch.randn(M, K, generator=g).cuda()
        w = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
        bias = torch.randn(N, generator=g).cuda()
        c = ops_.node_gemm(a, w, bias, None, False)
        ref = a.double().cpu() @ w.double().cpu().T + bias.double().cpu()
        bound = a.double().cpu().abs() @ w.double().cpu().abs().T + bias.double().cpu().abs()
        err = (c.double().cpu() - ref).abs()
        assert (err <= 2e-6 * bound).all(), ((M, N, K), float((err / bound).max()))
        c0 = torch.randn(M, N, generator=g).cuda()
        c2 = ops_.node_gemm(a, w, None, c0.clone(), True)
        ref2 = a.double().cpu() @ w.double().cpu().T + c0.double().cpu()
        assert torch.allclose(c2.double().cpu(), ref2, rtol=0, atol=3e-5), (M, N, K)
