"""Two launches of each K9/K10/K11 kernel at one IMPALA-CNN layer shape of the config-5 step (5,120 frames), for ncu
captures.  LAYER=s1.res|s1.conv|s2.conv|s2.res|s3.res (comma separated)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from impala_b200 import ops
CL = torch.channels_last
SH = {"s1.conv": (3, 16, 64), "s1.res": (16, 16, 32), "s2.conv": (16, 32, 32), "s2.res": (32, 32, 16), "s3.res": (32, 32, 8)}
N = int(os.environ.get("N", 5120))
for name in os.environ.get("LAYER", "s1.res").split(","):
    cin, cout, hw = SH[name]
    x = torch.randn(N, cin, hw, hw, device="cuda").contiguous(memory_format=CL)
    dy = torch.randn(N, cout, hw, hw, device="cuda").contiguous(memory_format=CL)
    w = torch.randn(cout, cin, 3, 3, device="cuda").contiguous(memory_format=CL)
    bias = torch.randn(cout, device="cuda")
    for _ in range(2):
        ops.conv3x3_wgrad(x, dy)
        y = ops.conv3x3_fwd(x, w)
        if cin != 3:
            ops.conv3x3_dgrad(dy, w)
        p, r, code = ops.maxpool3x3s2_fwd(y, bias, with_relu=True)
        ops.maxpool3x3s2_bwd(p, code, tuple(y.shape))
        v, rr = ops.bias_act_fwd(y, bias, dy, want_y=True, want_r=True)
        ops.relu_bwd_add(dy, y, rr)
torch.cuda.synchronize()
print("ok")
