"""Per-CTA phase timeline of K1 (B200RL_TRACE): prints where a launch's time goes.  Run on a GPU box."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from impala_b200 import ops, synth, _cabi

B, T, A = 4096, 20, 15
TM = len(sys.argv) > 1 and sys.argv[1] == "time_major"
dev = torch.device("cuda", 0)
sets = []
for i in range(24):
    d = synth.impala_rollout(B, T, A, seed=100 + i)
    if TM:
        d = {k: v.transpose(0, 1).contiguous() for k, v in d.items()}
    g = {k: v.to(dev) for k, v in d.items()}
    sets.append((g, (torch.empty(8, device=dev), torch.empty_like(g["logits"]), torch.empty_like(g["values"]))))
off = _cabi.load().b200rl_trace_offset_bytes()
flags_list = {"serial": ops.TRACE, "independent": ops.TRACE | ops.INPUTS_STABLE}
names = ["start", "landed", "stats", "scan", "stored", "reduced", "drained"]   # first unit of every warp group (persistent kernel)
flags_list["isolated"] = ops.TRACE      # ONE launch on an idle GPU with a flushed L2 (what ncu's gpu__time_duration sees)
only = sys.argv[1] if len(sys.argv) > 1 and sys.argv[1] in flags_list else None
flush = torch.empty(512 * 2**20, dtype=torch.uint8, device=dev)
for mode, flags in flags_list.items():
    if only and mode != only:
        continue

    def launch(i):
        g, out = sets[i % len(sets)]
        ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"], flags=flags, time_major=TM, out=out)
    for i in range(5):
        launch(i)
    torch.cuda.synchronize()
    if mode == "isolated":
        flush.fill_(1)
        torch.cuda.synchronize()
        for w in ops._workspaces.values():
            w[off:off + 4096 * 128].zero_()
        flush.fill_(2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); launch(7); e1.record()
        torch.cuda.synchronize()
        print("isolated launch, CUDA events around it: %.2f us" % (e0.elapsed_time(e1) * 1e3))
    else:
        cap = torch.cuda.Stream(dev)
        with torch.cuda.stream(cap):
            ops.workspace(dev)          # the capture stream's workspace must exist before capture
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=cap):
            for i in range(20):
                launch(i)
        gr.replay(); torch.cuda.synchronize()
        for w in ops._workspaces.values():
            w[off:off + 4096 * 128].zero_()
        torch.cuda.synchronize()
        gr.replay(); torch.cuda.synchronize()
    ws = [w for w in ops._workspaces.values() if int(w[off:off + 8].view(torch.int64).item()) > 0][0]
    tr = ws[off:off + 4096 * 128].view(torch.int64).cpu().numpy().reshape(-1, 16)
    halves = [tr[2048:3072], tr[3072:4096]]
    halves = [h[h[:, 2] > 0] for h in halves]
    halves.sort(key=lambda h: h[:, 2].min() if len(h) else 0)      # previous launch first
    cta = halves[-1]
    if len(halves[0]) and len(halves[1]):
        prev, cur = halves
        z = prev[:, 2].min()
        f = lambda x: "min %.0f med %.0f max %.0f" % (x.min() - z, np.median(x) - z, x.max() - z)
        print("  two consecutive launches, ns since the first CTA entry of the earlier one:")
        print("    earlier: entry %s | past wait %s | last warp at reduction %s | exit(thread 0) %s" % (f(prev[:, 2]), f(prev[:, 3]), f(prev[:, 0]), f(prev[:, 4])))
        print("    later:   entry %s | past wait %s | last warp at reduction %s | exit(thread 0) %s" % (f(cur[:, 2]), f(cur[:, 3]), f(cur[:, 0]), f(cur[:, 4])))
    tr = tr[:2048]
    tr = tr[tr[:, 1] > 0]          # warp groups that processed at least one unit
    ncta = len(tr)
    gt, ck, sm = tr[:, :7].astype(np.float64), tr[:, 8:15].astype(np.float64), tr[:, 7]
    t0 = gt[:, 0].min()
    print("== %s: last launch of a 20-launch graph, %d warp groups on %d SMs" % (mode, ncta, len(set(sm.tolist()))))
    print("globaltimer (ns since first CTA start): phase  min  median  max")
    for i, n in enumerate(names):
        x = gt[:, i] - t0
        print("  %-8s %8.0f %8.0f %8.0f" % (n, x.min(), np.median(x), x.max()))
    last = tr[tr[:, 6].argmax()]
    print("  latest group: " + " ".join("%s=%d" % (n, last[i] - t0) for i, n in enumerate(names)) + " smid=%d" % last[7])
    if len(cta):
        a, b = cta[:, 0].astype(np.float64) - t0, cta[:, 1].astype(np.float64) - t0
        print("  per CTA, last warp arrives at the reduction: min %.0f median %.0f p90 %.0f max %.0f; published/collected: median %.0f max %.0f (collector CTA %d: arrives %.0f, done %.0f)"
              % (a.min(), np.median(a), np.percentile(a, 90), a.max(), np.median(b), b.max(), len(cta) - 1, a[-1], b[-1]))
    print("clock64 deltas per CTA (cycles): phase  min  median  max")
    for i in range(1, 7):
        x = ck[:, i] - ck[:, i - 1]
        print("  %-8s %8.0f %8.0f %8.0f" % (names[i], x.min(), np.median(x), x.max()))
