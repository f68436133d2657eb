
import torch, sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "torchmd-net_b200"))
from torchmdnet import _native
_native.load()
ops = torch.ops.torchmdnet_b200
n = int(sys.argv[1]) if len(sys.argv) > 1 else 98304
x = torch.randn(n, 9, 128, device="cuda"); w = torch.randn(3, 128, 128, device="cuda") / 11.3
y = torch.empty_like(x)
for _ in range(3): ops.channel_mix(x, w, y, False)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): ops.channel_mix(x, w, y, False)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print("dbg", os.environ.get("TMDN_TC_DBG", "0"), "n", n, "ms", round(ms, 4), "GB/s", round(2 * x.numel() * 4 / ms / 1e6, 1), "TF/s(3x)", round(3 * 2 * n * 9 * 128 * 128 / ms / 1e9, 1))
