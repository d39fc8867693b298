"""K1 launch-configuration sweep (run on a GPU box): per-launch time of a dependent chain of K1 launches on ONE
stream, for the launch knobs the library reads from the environment (B200RL_K1_*).  Prints one line per
configuration: serial = every launch waits for its predecessor before reading (what a learner step sees),
stable = B200RL_INPUTS_STABLE (independent batches, PDL overlap)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from impala_b200 import ops, synth  # noqa: E402

B, T, A = int(os.environ.get("B", 4096)), int(os.environ.get("T", 20)), int(os.environ.get("A", 15))
N_LAUNCH, REPLAYS = 100, 30
dev = torch.device("cuda", 0)
nsets = int(os.environ.get("NSETS", max(4, int(2.5 * 126 * 2**20 // (204 * B * T)) + 1)))   # NSETS=1: L2-warm inputs (diagnostic only)
sets = []
for i in range(nsets):
    d = synth.impala_rollout(B, T, A, seed=100 + i)
    g = {k: v.to(dev) for k, v in d.items()}
    sets.append((g, (torch.empty(8, device=dev), torch.empty_like(g["logits"]), torch.empty_like(g["values"]))))
PEAK = 6547.2


def measure(flags):
    def launch(i):
        g, out = sets[i % nsets]
        ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"], flags=flags, out=out)
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        ops.workspace(dev)
        for i in range(3):
            launch(i)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=s):
        for i in range(N_LAUNCH):
            launch(i)
    ts = []
    with torch.cuda.stream(s):
        gr.replay()
        torch.cuda.synchronize()
        for _ in range(REPLAYS):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            gr.replay()
            e1.record(s)
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3 / N_LAUNCH)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    configs = [dict(x.split("=") for x in c.split(",") if x) for c in sys.argv[1:]] or [{}]
    for cfg in configs:
        for k in list(os.environ):
            if k.startswith("B200RL_K1_") or k == "B200RL_PDL":
                del os.environ[k]
        for k, v in cfg.items():
            os.environ[k] = v
        try:
            (ser, ser_min), (stb, stb_min) = measure(0), measure(ops.INPUTS_STABLE)
        except RuntimeError as e:
            print("%-60s ERROR %s" % (cfg, e), flush=True)
            continue
        by = 204 * B * T
        print("%-60s serial %.3f us (min %.3f) frac %.3f | stable %.3f us (min %.3f) frac %.3f" % (
            ",".join("%s=%s" % (k.replace("B200RL_", ""), v) for k, v in cfg.items()) or "default",
            ser, ser_min, by / ser / 1e3 / PEAK, stb, stb_min, by / stb / 1e3 / PEAK), flush=True)


if __name__ == "__main__":
    main()
