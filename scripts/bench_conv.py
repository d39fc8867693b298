"""Per-layer device time of the IMPALA-CNN convolution pieces at the config-5 shape (5,120 frames): K10/K11 vs cuDNN.
Run on a GPU box; diagnostic (CUDA events around 20 back-to-back launches after warm-up; inputs 0.2-1.6 GB >> L2)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from impala_b200 import ops

dev = torch.device("cuda", 0)
CL = torch.channels_last
N = int(os.environ.get("N", 5120))
torch.backends.cudnn.benchmark = os.environ.get("CUDNN_BENCHMARK", "0") == "1"     # the reference sets it (main.py:59)
LAYERS = [("s1.conv", 3, 16, 64), ("s1.res", 16, 16, 32), ("s2.conv", 16, 32, 32), ("s2.res", 32, 32, 16),
          ("s3.conv", 32, 32, 16), ("s3.res", 32, 32, 8)]


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ONLY = os.environ.get("LAYERS")
for name, cin, cout, hw in LAYERS:
    if ONLY and name not in ONLY.split(","):
        continue
    x = torch.randn(N, cin, hw, hw, device=dev).contiguous(memory_format=CL)
    dy = torch.randn(N, cout, hw, hw, device=dev).contiguous(memory_format=CL)
    w = torch.randn(cout, cin, 3, 3, device=dev).contiguous(memory_format=CL)
    gb = (x.numel() + dy.numel()) * 4 / 1e9
    row = {"layer": name, "GB": round(gb, 3), "ideal_ms@6.5TB/s": round(gb / 6.55, 3)}
    row["k10_wgrad_ms"] = round(timeit(lambda: ops.conv3x3_wgrad(x, dy)), 3)
    row["cudnn_wgrad_ms"] = round(timeit(lambda: torch.ops.aten.convolution_backward(
        dy, x, w, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1, (False, True, False))), 3)
    row["cudnn_fprop_ms"] = round(timeit(lambda: F.conv2d(x, w, None, padding=1)), 3)
    row["cudnn_dgrad_ms"] = round(timeit(lambda: torch.ops.aten.convolution_backward(
        dy, x, w, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1, (True, False, False))), 3)
    row["k11_fprop_ms"] = round(timeit(lambda: ops.conv3x3_fwd(x, w)), 3)
    if cin != 3:
        row["k11_dgrad_ms"] = round(timeit(lambda: ops.conv3x3_dgrad(dy, w)), 3)
    print(row, flush=True)
