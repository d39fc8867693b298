"""One launch of K10 per IMPALA-CNN layer shape (for ncu captures).  LAYER=s1.res|s2.conv|... selects one."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from impala_b200 import ops
CL = torch.channels_last
SH = {"s1.conv": (3, 16, 64), "s1.res": (16, 16, 32), "s2.conv": (16, 32, 32), "s2.res": (32, 32, 16), "s3.res": (32, 32, 8)}
N = int(os.environ.get("N", 5120))
for name in os.environ.get("LAYER", "s1.res,s2.res").split(","):
    cin, cout, hw = SH[name]
    x = torch.randn(N, cin, hw, hw, device="cuda").contiguous(memory_format=CL)
    dy = torch.randn(N, cout, hw, hw, device="cuda").contiguous(memory_format=CL)
    for _ in range(2):
        ops.conv3x3_wgrad(x, dy)
    if hasattr(ops, "conv3x3_fwd"):
        w = torch.randn(cout, cin, 3, 3, device="cuda").contiguous(memory_format=CL)
        for _ in range(2):
            ops.conv3x3_fwd(x, w)
torch.cuda.synchronize()
print("ok")
