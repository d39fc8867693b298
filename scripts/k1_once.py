"""A handful of K1 launches at BASELINE config 2 on rotating cold buffers (the ncu target of scripts/gpu_profile_k1.sh)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from impala_b200 import ops, synth  # noqa: E402

B, T, A = 4096, 20, 15
dev = torch.device("cuda", 0)
sets = []
for i in range(12):
    d = synth.impala_rollout(B, T, A, seed=100 + i)
    g = {k: v.to(dev) for k, v in d.items()}
    sets.append((g, (torch.empty(8, device=dev), torch.empty_like(g["logits"]), torch.empty_like(g["values"]))))
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 24):
    g, out = sets[i % len(sets)]
    ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"], out=out)
torch.cuda.synchronize()
print("ok", float(sets[0][1][0][0]))
