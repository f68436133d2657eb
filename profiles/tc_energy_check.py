"""Energy/force difference of the TensorNet path between the GEMM variants (env TMDN_TC_MIX / TMDN_TC_GEMM, read once per
process) and the fp64 oracle, 1536-atom periodic water box (run on the GPU box)."""
import os
import sys
import numpy as np
import torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torchmd-net_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_distributed as T
m, hp, z, pos, box = T._model_and_system(8)
y, f = m(z.cuda(), pos.cuda())
out = sys.argv[1]
np.savez(out, y=y.double().cpu().numpy(), f=f.double().cpu().numpy())
print(os.environ.get("TMDN_TC_MIX", "1"), os.environ.get("TMDN_TC_GEMM", "1"), float(y))
