"""GPU: K8 (fused peer-memory allreduce + clip + Adam).  One GPU: the world-size-1 peer path (flags and system-scope
loads against this rank's own bucket) must reproduce K6 bit for bit.  Two or more GPUs: the torchrun check in
scripts/peer_adam_check.py (skipped on a one-GPU box)."""
import os
import subprocess
import sys

import pytest
import torch

from impala_b200 import optim, parallel

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("route", ["0", "1"])      # one-shot / two-shot exchange (B200RL_K8_TWO_SHOT)
@pytest.mark.parametrize("n", [1001, 630_352, 3_000_000])
def test_peer_path_world1_equals_k6(n, route, monkeypatch):
    monkeypatch.setenv("B200RL_K8_TWO_SHOT", route)
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    pa = [torch.nn.Parameter(torch.randn(n - 5, device=dev)), torch.nn.Parameter(torch.randn(5, device=dev))]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    opt_a, opt_b = optim.FusedClipAdam(pa, lr=1e-2, eps=1e-5), optim.FusedClipAdam(pb, lr=1e-2, eps=1e-5)
    bucket, ref = parallel.PeerGradBucket(pa, n_extra=8), parallel.GradBucket(pb, n_extra=8)
    assert bucket.world_size == 1 and bucket.bucket_ptrs == [bucket.store.data_ptr()]
    for step in range(3):
        bucket.attach()
        ref.attach()
        g = torch.randn(n + 8, device=dev) * (0.001 if step == 1 else 1.0)
        bucket.flat.copy_(g)
        ref.flat.copy_(g)
        na = opt_a.step_peer(bucket, 0.5).clone()
        nb = opt_b.step_flat(ref.grads, 0.5).clone()
        assert torch.equal(na, nb)
        assert torch.equal(opt_a.flat, opt_b.flat) and torch.equal(opt_a.exp_avg_sq, opt_b.exp_avg_sq)
        assert torch.equal(bucket.reduced, ref.grads) and torch.equal(bucket.extra_reduced, g[n:])
        assert torch.equal(bucket.grads, g[:n])                     # the shared bucket itself is left untouched
    with pytest.raises(RuntimeError):
        bucket.allreduce()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_allreduce_two_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "scripts", "peer_adam_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0 and '"peer_adam_check": "ok"' in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
