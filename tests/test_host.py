"""CPU: host-side logic -- column sharding, the one-bucket gradient allreduce (gloo, world_size 2), and
the property the multi-GPU path rests on: rank-local losses normalised by the GLOBAL batch sum to the
single-process result."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from closed_form import impala_closed_form
from impala_b200 import models, parallel, synth


def test_shard_columns_partition():
    for B, W in [(8192, 8), (10, 4), (3, 8), (4096, 1)]:
        spans = [parallel.shard_columns(B, r, W) for r in range(W)]
        assert spans[0][0] == 0 and spans[-1][1] == B
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


def test_grad_bucket_views_and_clip():
    m = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.Linear(3, 2))
    b = parallel.GradBucket(m.parameters())
    b.attach()
    m(torch.ones(5, 4)).sum().backward()
    flat = torch.cat([p.grad.reshape(-1) for p in m.parameters()])
    assert torch.equal(flat, b.grads) and b.extra.numel() == 8
    want = torch.linalg.vector_norm(flat).item()
    ref = [p.grad.clone() for p in m.parameters()]
    total = b.clip_(0.5)
    assert total.item() == pytest.approx(want)
    for p, g in zip(m.parameters(), ref):
        assert torch.allclose(p.grad, g * min(1.0, 0.5 / (want + 1e-6)))
    b.attach()
    assert float(b.flat.abs().sum()) == 0.0


def test_model_parameter_counts():
    """Same shapes as the reference networks: 344,496 / 630,352 parameters at A=15."""
    assert sum(p.numel() for p in models.ActorCritic((3, 64, 64), 15, body="nature").parameters()) == 344_496
    assert sum(p.numel() for p in models.ActorCritic((3, 64, 64), 15, body="impala").parameters()) == 630_352
    pi, v = models.ActorCritic((3, 64, 64), 15)(torch.zeros(2, 3, 64, 64, dtype=torch.uint8))
    assert pi.shape == (2, 15) and v.shape == (2, 1)


def _worker(rank, world_size, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        B, T, A = 12, 9, 5
        d = {k: (v.double() if v.is_floating_point() else v) for k, v in synth.impala_rollout(B, T, A, 42).items()}
        lo, hi = parallel.shard_columns(B, rank, world_size)
        sh = {k: v[lo:hi] for k, v in d.items()}
        # rank-local loss with GLOBAL normalisation == what the kernel does with n_global_b
        loss, m, dl, dv = impala_closed_form(sh["logits"], sh["values"], sh["a"], sh["r"], sh["discount"], sh["pi_ref"])
        scale_pg = (hi - lo) * (T - 1) / (B * (T - 1))
        scale_all = (hi - lo) * T / (B * T)
        # a "model": logits = W-independent passthrough, so parameter grads = sums of dlogits/dvalues
        w = torch.nn.Parameter(torch.zeros(A, dtype=torch.float32))
        u = torch.nn.Parameter(torch.zeros(1, dtype=torch.float32))
        bucket = parallel.GradBucket([w, u])
        bucket.attach()
        w.grad += (dl * scale_pg).sum((0, 1)).float()
        u.grad += (dv * scale_pg).sum().float()
        bucket.extra[0] = float(m["pg"] * scale_pg)
        bucket.extra[1] = float(m["entropy"] * scale_all)
        bucket.allreduce()
        if rank == 0:
            out.put((bucket.flat.clone(), ))
    finally:
        dist.destroy_process_group()


def test_sharded_sum_equals_global_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    import socket
    with socket.socket() as sk:                                  # a port that is free right now (not pid-derived)
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    (flat,) = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    B, T, A = 12, 9, 5
    d = {k: (v.double() if v.is_floating_point() else v) for k, v in synth.impala_rollout(B, T, A, 42).items()}
    loss, m, dl, dv = impala_closed_form(d["logits"], d["values"], d["a"], d["r"], d["discount"], d["pi_ref"])
    # entropy part of dlogits is normalised by N_all, pg part by N_pg: with equal T per shard both scale alike
    assert torch.allclose(flat[:A].double(), dl.sum((0, 1)), atol=1e-6)
    assert flat[A].item() == pytest.approx(dv.sum().item(), abs=1e-6)
    assert flat[A + 1].item() == pytest.approx(m["pg"].item(), abs=1e-6)
    assert flat[A + 2].item() == pytest.approx(m["entropy"].item(), abs=1e-6)


def test_make_bucket_routes_on_cpu(monkeypatch):
    """``parallel.make_bucket``: CPU parameters / one rank / B200RL_PEER_ALLREDUCE=0 give the plain bucket; the
    peer-mapped bucket (K8) is a CUDA feature and says so."""
    from impala_b200 import parallel
    ps = [torch.nn.Parameter(torch.randn(5)), torch.nn.Parameter(torch.randn(3, 2))]
    b = parallel.make_bucket(ps, None, fused_optimizer=True)
    assert type(b) is parallel.GradBucket and b.flat.numel() == 11 + 8 and b.store.numel() == b.flat.numel()
    b.attach()
    assert all(p.grad.data_ptr() == v.data_ptr() for p, v in zip(b.params, b.views))
    monkeypatch.setenv("B200RL_PEER_ALLREDUCE", "0")
    assert type(parallel.make_bucket(ps, None, fused_optimizer=True)) is parallel.GradBucket
    with pytest.raises(RuntimeError):
        parallel.PeerGradBucket(ps)


def test_flat_view_keeps_memory_format():
    """Flat parameter / gradient buckets keep a channels_last convolution weight channels_last (K10/K11 read and
    write the physical [O][3][3][I] layout), everything else contiguous."""
    from impala_b200.parallel import flat_view
    w = torch.randn(32, 16, 3, 3).contiguous(memory_format=torch.channels_last)
    b = torch.randn(32)
    flat = torch.zeros(5 + w.numel() + b.numel())
    vw, vb = flat_view(flat, 5, w), flat_view(flat, 5 + w.numel(), b)
    vw.copy_(w)
    vb.copy_(b)
    assert torch.equal(vw, w) and vw.stride() == w.stride() and vw.data_ptr() == flat[5:].data_ptr()
    assert torch.equal(flat[5:5 + w.numel()].view(32, 3, 3, 16).permute(0, 3, 1, 2), w)
    assert torch.equal(flat[5 + w.numel():], b) and flat[:5].abs().sum() == 0
    wc = torch.randn(8, 4, 3, 3)                                # plain contiguous weight: plain view
    assert flat_view(torch.zeros(wc.numel()), 0, wc).is_contiguous()


def test_fused_model_host_contract():
    """K9-K11 host side, no kernel calls: the fused model has the eager model's parameters and state dict, refuses CPU
    tensors (no fallback) and non-channels_last construction; the shape predicate matches the C ABI's supported set."""
    from impala_b200 import cnn_glue, ops
    eager = models.ActorCritic((3, 64, 64), 15, body="impala", channels_last=True)
    fused = models.ActorCritic((3, 64, 64), 15, body="impala", channels_last=True, fused_glue=True)
    assert [(n, tuple(p.shape)) for n, p in eager.named_parameters()] == [(n, tuple(p.shape)) for n, p in fused.named_parameters()]
    fused.load_state_dict(eager.state_dict())
    assert sum(p.numel() for p in fused.parameters()) == 630_352
    with pytest.raises(RuntimeError):
        fused(torch.zeros(2, 3, 64, 64, dtype=torch.uint8))                  # CPU tensor: raises, never falls back
    with pytest.raises(RuntimeError):
        models.ActorCritic((3, 64, 64), 15, body="impala", channels_last=False, fused_glue=True)
    with pytest.raises(RuntimeError):
        models.ActorCritic((3, 64, 64), 15, body="nature", channels_last=True, fused_glue=True)
    for shape, cout, want in (((8, 3, 64, 64), 16, True), ((8, 16, 32, 32), 16, True), ((8, 16, 32, 32), 32, True),
                              ((8, 32, 16, 16), 32, True), ((8, 32, 8, 8), 32, True), ((8, 16, 12, 24), 16, False),
                              ((8, 8, 32, 32), 16, False), ((8, 16, 32, 32), 64, False), ((8, 16, 32, 128), 16, False)):
        assert ops.conv3x3_wgrad_supported(shape, cout) is want, (shape, cout)
    w3232, w1616 = torch.zeros(32, 32, 3, 3), torch.zeros(16, 16, 3, 3)
    assert not cnn_glue._k11_wins(w3232) and cnn_glue._k11_wins(w1616)       # 32 -> 32 forward / dgrad stay on cuDNN
