"""Error of the node GEMM kernels against an fp64 product, in units of sum_k |a_k w_k| (run on the GPU box)."""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "torchmd-net_b200"))
from torchmdnet import _native
_native.load()
ops = torch.ops.torchmdnet_b200
g = torch.Generator().manual_seed(1)
for (M, N, K) in [(4096, 128, 128), (4096, 384, 256), (4096, 128, 640)]:
    a = torch.randn(M, K, generator=g).cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    ref = a.double() @ w.double().T
    bound = a.double().abs() @ w.double().abs().T
    c = ops.node_gemm(a, w, None, None, False)
    f = (a @ w.T)           # cuBLAS fp32 (TF32 off by default)
    e1 = ((c.double() - ref).abs() / bound)
    e2 = ((f.double() - ref).abs() / bound)
    print("M N K", M, N, K, "tcgen05 3xTF32: max %.2e rms %.2e mean-signed %.2e | fp32 FFMA (cuBLAS): max %.2e rms %.2e" %
          (e1.max(), e1.pow(2).mean().sqrt(), ((c.double() - ref) / bound).mean(), e2.max(), e2.pow(2).mean().sqrt()))
