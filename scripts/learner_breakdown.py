"""Phase breakdown of the full IMPALA learner step (BASELINE config 5) on one GPU.  Run on a GPU box."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from impala_b200 import models, replay, ops, parallel

dev = torch.device("cuda", 0)
B, T, A = 256, 20, 15


def run(channels_last, benchmark, tf32=False):
    torch.backends.cudnn.benchmark = benchmark
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.manual_seed(0)
    model = models.ActorCritic((3, 64, 64), A, body="impala", channels_last=channels_last).to(dev)
    if channels_last:
        model = model.to(memory_format=torch.channels_last)
    opt = torch.optim.Adam(model.parameters(), lr=1e-4, eps=1e-5)
    obs = torch.randint(0, 256, (1000, T, 3, 64, 64), device=dev, dtype=torch.uint8)
    d = {k: v.to(dev) for k, v in __import__("impala_b200.synth", fromlist=["x"]).impala_rollout(B, T, A, 1).items()}
    bucket = parallel.GradBucket(model.parameters())
    ev = lambda: torch.cuda.Event(enable_timing=True)
    names = ["sample+gather", "cnn fwd", "K1 loss", "cnn bwd", "clip", "adam"]
    tot = dict.fromkeys(names, 0.0)
    n = 12
    for it in range(n + 4):
        e = [ev() for _ in range(7)]
        bucket.attach()
        e[0].record()
        idx = ops.sample_uniform(123, 0, it * B, B, 1000, dev)
        s = ops.gather_rows(obs, idx)
        e[1].record()
        pi, v = model(s.flatten(0, 1))
        pi, v = pi.reshape(B, T, -1), v.reshape(B, T)
        e[2].record()
        sc, dl, dv = ops.impala_loss_fwdbwd(pi, v, d["a"], d["r"], d["discount"], d["pi_ref"])
        e[3].record()
        torch.autograd.backward([pi, v], [dl, dv])
        e[4].record()
        bucket.clip_(0.5)
        e[5].record()
        opt.step()
        e[6].record()
        torch.cuda.synchronize()
        if it >= 4:
            for k, nm in enumerate(names):
                tot[nm] += e[k].elapsed_time(e[k + 1])
    total = sum(tot.values()) / n
    print("channels_last=%s cudnn.benchmark=%s tf32=%s: %.2f ms/step = %.0f samples/s | %s" % (
        channels_last, benchmark, tf32, total, B * T / total * 1e3,
        ", ".join("%s %.3f" % (k, v / n) for k, v in tot.items())))


if os.environ.get("TF32_ONLY") != "1":
    for cl in (False, True):
        for bm in (False, True):
            run(cl, bm)
for cl in (False, True):
    for bm in (False, True):
        run(cl, bm, tf32=True)
