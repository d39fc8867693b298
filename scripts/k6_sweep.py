import os, sys, json, subprocess
sys.path.insert(0, "/root/repo")
import torch
from impala_b200 import ops
dev = torch.device("cuda", 0)
n = 630_352
sets = [[torch.randn(n, device=dev) for _ in range(2)] + [torch.zeros(n, device=dev), torch.zeros(n, device=dev)] for _ in range(24)]
s = torch.cuda.Stream(dev)
with torch.cuda.stream(s):
    ops.workspace(dev)
    for i in range(3):
        ops.clip_adam_step(*sets[i], i + 1, 1e-4, eps=1e-5, max_norm=0.5)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=s):
    for i in range(96):
        ops.clip_adam_step(*sets[i % 24], i + 1, 1e-4, eps=1e-5, max_norm=0.5)
g.replay(); torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(s):
    a.record(s); g.replay(); b.record(s)
torch.cuda.synchronize()
print(os.environ.get("B200RL_K6_ELEMS"), "us/launch %.2f" % (a.elapsed_time(b) / 96 * 1e3))
