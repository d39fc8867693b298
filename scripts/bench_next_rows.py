"""Device timings of the SURVEY sec 8f rows (K6/N1..N5) against their algorithmic bytes.  Run on a GPU box.
CUDA-graph replays of back-to-back launches, CUDA events, rotating buffers where the data would fit L2."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from impala_b200 import ops

dev = torch.device("cuda", 0)
PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists("MEASURED_PEAKS.json") else 6650.0


def timed(fn, steps=200, warm=5, n_streams=4):
    """Independent launches issued round-robin on n_streams side streams inside one graph (as bench.py does)."""
    s = torch.cuda.Stream(dev)
    side = [torch.cuda.Stream(dev) for _ in range(n_streams)]
    for st in [s] + side:
        with torch.cuda.stream(st):
            ops.workspace(dev)
    with torch.cuda.stream(s):
        for i in range(warm):
            fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        for st in side:
            st.wait_stream(s)
        for i in range(steps):
            with torch.cuda.stream(side[i % n_streams]):
                fn(i)
        for st in side:
            s.wait_stream(st)
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        a.record(s); g.replay(); b.record(s)
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps * 1e3      # us


def report(name, us, nbytes, note=""):
    gbs = nbytes / us / 1e3 if nbytes else 0.0
    print(json.dumps({"row": name, "us_per_launch": round(us, 3), "algorithmic_bytes": nbytes,
                      "GBps": round(gbs, 1), "frac_of_measured_hbm_peak": round(gbs / PEAK, 3) if nbytes else None, "note": note}))


# K6 / N1: clip + Adam, 32 B per parameter
for n in (630_352, 50_000_000):
    sets = [[torch.randn(n, device=dev) for _ in range(2)] + [torch.zeros(n, device=dev), torch.zeros(n, device=dev)]
            for _ in range(max(2, int(400e6 // (16 * n)) + 1))]
    report("K6 clip+Adam n=%d" % n, timed(lambda i: ops.clip_adam_step(*sets[i % len(sets)], i + 1, 1e-4, eps=1e-5, max_norm=0.5),
                                          steps=100, n_streams=1), 32 * n, "16 B read + 16 B written per parameter")
    del sets
# N2: ingest 256 rollouts of T=20 (5 fields) + fused discount
n, T, A, cap = 256, 20, 15, 4096
fields = [torch.empty(cap, T, 3, 64, 64, dtype=torch.uint8, device=dev), torch.empty(cap, T, 1, dtype=torch.int64, device=dev),
          torch.empty(cap, T, 1, device=dev), torch.empty(cap, T, 1, device=dev), torch.empty(cap, T, A, device=dev)]
rows = [torch.randint(0, 256, (n, T, 3, 64, 64), dtype=torch.uint8, device=dev), torch.zeros(n, T, 1, dtype=torch.int64, device=dev),
        torch.randn(n, T, 1, device=dev), torch.rand(n, T, 1, device=dev) > .9, torch.randn(n, T, A, device=dev)]
row_b = sum(r[0].numel() * r.element_size() for r in rows) + T * 4      # + the float discount written
report("N2 ingest 256 rollouts (5 fields, fused discount)", timed(lambda i: ops.ring_append_fields(fields, (i * n) % cap, rows, done_field=3), n_streams=1),
       2 * n * row_b, "read + write of every row")
idxs = [ops.sample_uniform(1, 0, k * n, n, cap, dev) for k in range(8)]     # 8 x 63 MB of distinct rows > L2
outs = [[torch.empty((n,) + tuple(f.shape[1:]), dtype=f.dtype, device=dev) for f in fields] for _ in range(4)]
report("N2 gather 256 rollouts (5 fields, one launch)", timed(lambda i: ops.gather_fields(fields, idxs[i % 8], outs[i % 4])), 2 * n * row_b,
       "8 rotating index sets over a 1 GB ring, 4 rotating outputs (252 MB)")
del fields, rows, outs
# N3: sum tree over 1M slots
P = 1 << 20
tree = torch.zeros(2 * P, dtype=torch.float64, device=dev)
ops.sumtree_update(tree, torch.arange(1_000_000, device=dev), torch.rand(1_000_000, device=dev, dtype=torch.float64) + 1e-3)
ui = torch.randint(0, 1_000_000, (4096,), device=dev)
uv = torch.rand(4096, device=dev, dtype=torch.float64)
report("N3 sumtree update 4096 of 1M", timed(lambda i: ops.sumtree_update(tree, ui, uv), n_streams=1), 0, "128 subtree CTAs + one top CTA, latency-bound")
oi, op_ = torch.empty(4096, dtype=torch.int64, device=dev), torch.empty(4096, device=dev)
from impala_b200 import _cabi
report("N3 sumtree sample 4096 of 1M", timed(lambda i: _cabi.call("b200rl_sumtree_sample", tree.data_ptr(), P, 1_000_000, 1, 0, i * 4096,
                                                                 4096, oi.data_ptr(), op_.data_ptr(), torch.cuda.current_stream().cuda_stream)),
       0, "20 dependent L2/HBM loads per draw, latency-bound")
# N4: n-step returns at the C3 shape, time-major
T, B = 256, 8192
xs = [[torch.randn(T, B, device=dev) for _ in range(3)] for _ in range(6)]
report("N4 n-step returns n=3, T=256 B=8192", timed(lambda i: ops.n_step_returns(*[x.t() for x in xs[i % 6]], 3)), 16 * T * B)
outs2 = [torch.empty(T, B, device=dev) for _ in range(6)]
report("K2 lambda-returns T=256 B=8192 (alone)", timed(lambda i: ops.lambda_returns(*[x.t() for x in xs[i % 6]], 0.95)), 16 * T * B)
report("K2 lambda-returns T=256 B=8192 (alone, 1 stream)", timed(lambda i: ops.lambda_returns(*[x.t() for x in xs[i % 6]], 0.95), n_streams=1), 16 * T * B)
xs5 = [[torch.randn(512, 4096, device=dev) for _ in range(3)] for _ in range(6)]
report("K2 lambda-returns T=512 B=4096 (alone)", timed(lambda i: ops.lambda_returns(*[x.t() for x in xs5[i % 6]], 0.95)), 16 * 512 * 4096)
del xs5
# N4b: quantile Huber, B=256 N=M=64 ; N5: act epilogue 128 rows
src, tau, tgt = torch.randn(256, 64, device=dev), torch.rand(256, 64, device=dev), torch.randn(256, 64, device=dev)
report("N4b quantile-Huber B=256 N=M=64", timed(lambda i: ops.quantile_huber_fwdbwd(src, tau, tgt)), 0, "launch-latency bound (262 KB)")
lg = torch.randn(128, 15, device=dev)
det = torch.tensor([False], device=dev)
report("N5 act epilogue 128 rows", timed(lambda i: ops.act_epilogue(lg, det, 1, 0, i * 128)), 0, "launch-latency bound (reference inference batch 128)")

# K7: ppo_loss_is (losses.py:62-89) at the C2 shape, 201 B/sample; rotating sets > L2
from impala_b200 import synth
B, T, A = 4096, 20, 15
nsets = int(2.5 * 126e6 // (201 * B * T)) + 1
sets = []
for i in range(nsets):
    d = {k: v.to(dev) for k, v in synth.ppo_is_rollout(B, T, A, seed=i).items()}
    sets.append((d, (torch.empty(8, device=dev), torch.empty_like(d["logits"]), torch.empty_like(d["values"]))))


def k7(i):
    d, out = sets[i % nsets]
    ops.ppo_is_loss_fwdbwd(d["logits"], d["values"], d["v_boot"], d["a"], d["r"], d["done"], d["pi_ref"],
                           flags=ops.INPUTS_STABLE, out=out)


report("K7 ppo_loss_is T=20 B=4096 A=15", timed(k7, steps=400, n_streams=6), 201 * B * T + 4 * B,
       "6 streams, INPUTS_STABLE; 201 B/sample + 4 B/rollout bootstrap")

# K1 in the same harness and on the same rollouts, for comparison with K7 (204 B/sample)
disc = [(~d["done"]).float() * 0.99 for d, _ in sets]


def k1(i):
    d, out = sets[i % nsets]
    ops.impala_loss_fwdbwd(d["logits"], d["values"], d["a"], d["r"], disc[i % nsets], d["pi_ref"], flags=ops.INPUTS_STABLE, out=out)


report("K1 impala_loss T=20 B=4096 A=15 (same harness)", timed(k1, steps=400, n_streams=6), 204 * B * T, "6 streams, INPUTS_STABLE")
report("K1 impala_loss T=20 B=4096 A=15 (same harness, 4000 steps)", timed(k1, steps=4000, n_streams=6), 204 * B * T, "6 streams, INPUTS_STABLE")
report("K7 ppo_loss_is (4000 steps)", timed(k7, steps=4000, n_streams=6), 201 * B * T + 4 * B, "6 streams, INPUTS_STABLE")

# stand-alone V-trace (A1 drop-in, b200rl_vtrace) at the C2 shape: 5 inputs + 3 outputs = 32 B per step
Bv, Tv = 4096, 19
vs = [[torch.randn(Bv, Tv, device=dev) for _ in range(4)] + [torch.rand(Bv, Tv, device=dev) * 2] for _ in range(60)]
report("A1 stand-alone V-trace fwd B=4096 T-1=19 (batch-major)", timed(lambda i: ops.vtrace_fwd(*vs[i % 60])), 32 * Bv * Tv,
       "20 B read + 12 B written per step; 2.5 MB per launch: launch-bound")
del vs

# N3b: PER-weighted TD loss (agents/dqn/learning.py:161-172): Ape-X batch 512 x 18 actions, and a large batch; next to
# the eager torch ops of the reference's lines on the same tensors
for Np, Ap in ((512, 18), (65536, 18)):
    per = [(torch.randn(Np, Ap, device=dev), torch.randint(0, Ap, (Np,), device=dev), torch.randn(Np, device=dev),
            torch.rand(Np, device=dev) * 0.01 + 1e-5) for _ in range(20)]
    us = timed(lambda i: ops.per_td_loss_fwdbwd(*per[i % 20], 0.4), n_streams=1)
    nbytes = Np * (32 + 8 + 4 + 4 + 4 * Ap + 4)     # one sector of q, action, target, probability in; dq row + priority out
    report("N3b PER TD loss N=%d A=%d" % (Np, Ap), us, nbytes, "dependent chain on one stream")

    def eager(i):
        q, a, tg, p = per[i % 20]
        q = q.clone().requires_grad_(True)
        qa = q.gather(1, a.unsqueeze(-1)).squeeze(-1)
        w = p.pow(-0.4)
        w = w / w.max()
        err = tg - qa
        (0.5 * (err.pow(2) * w).mean()).backward()
        return err.detach().abs()
    for _ in range(5):
        eager(0)
    torch.cuda.synchronize()
    a0, b0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record()
    for i in range(100):
        eager(i)
    b0.record(); torch.cuda.synchronize()
    report("   the reference's eager torch ops for the same lines N=%d" % Np, a0.elapsed_time(b0) * 10, nbytes, "eager, host-launched")
    del per
