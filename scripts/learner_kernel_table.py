"""Per-kernel device-time table of the IMPALA-CNN forward+backward inside the learner step (config 5 shape,
B=256 x T=20 = 5,120 frames), via torch.profiler (CUPTI).  Run on a GPU box; diagnostic only, not a bench number."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from impala_b200 import models

dev = torch.device("cuda", 0)
cl = os.environ.get("CL", "1") == "1"
torch.manual_seed(0)
fused = cl and os.environ.get("FUSED", "1") == "1"
model = models.ActorCritic((3, 64, 64), 15, body="impala", channels_last=cl, fused_glue=fused).to(dev)
if cl:
    model = model.to(memory_format=torch.channels_last)
obs = torch.randint(0, 256, (5120, 3, 64, 64), device=dev, dtype=torch.uint8)


def step():
    pi, v = model(obs)
    (pi.sum() + v.sum()).backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=90))
