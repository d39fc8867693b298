"""Seeded random-shape cross-checks of the paths that have TWO device implementations (chosen by a size rule at run
time): a shape the rule sends to one kernel must give what the other kernel gives.  The per-sample code is shared, so
gradients / gathered rows / indices are compared bit for bit; scalars (another summation order) to the tolerance.  The
hand-picked shapes live in test_gpu_impala.py / test_gpu_ppo_returns.py / test_gpu_replay_learner.py; these draws are
there for the shapes nobody thought of (a ring-round aliasing bug in the persistent K1 kernel was of that kind)."""
import os
import random

import pytest
import torch

from impala_b200 import ops, synth
from gpu_util import assert_close, oracle_impala, to_cuda

pytestmark = pytest.mark.gpu
_MORE = int(os.environ.get("B200RL_FUZZ_SCALE", "1"))   # e.g. 10 for a one-off long run (scripts/gpu_fuzz.sh)


def _k1_shapes(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        T = rng.choice([2, 3, 5, 16, 20, 20, 20, 33, 64, 128])
        A = rng.choice([1, 2, 3, 6, 15, 15, 18, 32])
        # keep B*T*A bounded (~40 M elements) so one draw stays in the tens of milliseconds
        B = rng.randint(1, max(1, min(70000, 40_000_000 // (T * A))))
        if rng.random() < 0.3:
            B = rng.randint(1, 300)
        out.append((B, T, A))
    return out


@pytest.mark.parametrize("B,T,A", _k1_shapes(24 * _MORE, seed=20261105))
def test_k1_random_shapes_persistent_equals_tile(cuda, B, T, A, monkeypatch):
    g = to_cuda(synth.impala_rollout(B, T, A, seed=B * 7 + T))
    outs = {}
    for impl in ("1", "2"):
        monkeypatch.setenv("B200RL_K1_TILE", impl)
        outs[impl] = [x.clone() for x in
                      ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"])]
        torch.cuda.synchronize()
    assert torch.equal(outs["1"][1], outs["2"][1]), "dlogits differ between the two K1 kernels"
    assert torch.equal(outs["1"][2], outs["2"][2]), "dvalues differ between the two K1 kernels"
    torch.testing.assert_close(outs["1"][0][:6], outs["2"][0][:6], rtol=1e-5, atol=1e-6)
    # oracle on a column sample (columns are independent; n_global_b = B keeps the 1/B scaling)
    n = min(B, 16)
    for sl in (slice(0, n), slice(B - n, B)):
        d = {k: v[sl].cpu() for k, v in g.items()}
        _, _, dl_ref, dv_ref = oracle_impala(d)
        assert_close(outs["2"][1][sl], dl_ref * (n / B), "dlogits (column sample)", rel_to_max=2e-6)
        assert_close(outs["2"][2][sl], dv_ref * (n / B), "dvalues (column sample)", rel_to_max=2e-6)


def _k3_shapes(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        A = rng.choice([1, 2, 6, 15, 15, 18, 32])
        N = rng.randint(1, 2_500_000 if A <= 18 else 1_200_000)
        if rng.random() < 0.3:
            N = rng.randint(1, 5000)
        out.append((N, A))
    return out


@pytest.mark.parametrize("N,A", _k3_shapes(16 * _MORE, seed=31))
def test_k3_random_sizes_streaming_equals_tile(cuda, N, A, monkeypatch):
    d = to_cuda(synth.ppo_minibatch(N, A, seed=N % 1000))
    outs = {}
    for impl in ("1", "2"):
        monkeypatch.setenv("B200RL_K3_STREAM", impl)
        outs[impl] = [x.clone() for x in
                      ops.ppo_loss_fwdbwd(d["logits"], d["v"].reshape(-1), d["a"], d["v_target"], d["pi_ref"])]
        torch.cuda.synchronize()
    assert torch.equal(outs["1"][1], outs["2"][1]), "dlogits differ between the two K3 kernels"
    assert torch.equal(outs["1"][2], outs["2"][2]), "dv differ between the two K3 kernels"
    torch.testing.assert_close(outs["1"][0][:7], outs["2"][0][:7], rtol=1e-5, atol=1e-6)


def _replay_cases(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        cap = rng.randint(1, 5000)
        size = rng.randint(1, cap)
        draws = rng.randint(1, 4000)
        rows = [rng.choice([1, 3, 4, 7, 16, 33, 64, 250, 1024, 4099]) for _ in range(rng.randint(1, 5))]
        out.append((cap, size, draws, tuple(rows), rng.randint(0, 2 ** 40)))
    return out


@pytest.mark.parametrize("cap,size,draws,rows,offset", _replay_cases(16 * _MORE, seed=47))
def test_fused_sample_gather_random_cases_equal_two_launches(cuda, cap, size, draws, rows, offset):
    g = torch.Generator(device=cuda).manual_seed(cap)
    dts = [torch.uint8, torch.float32, torch.int64, torch.int32]
    stores = [torch.randint(0, 250, (cap, r), generator=g, device=cuda).to(dts[i % 4]) for i, r in enumerate(rows)]
    idx = ops.sample_uniform(11, 3, offset, draws, size, cuda)
    want = ops.gather_fields(stores, idx)
    got_idx, got = ops.sample_gather_fields(stores, 11, 3, offset, draws, size)
    torch.cuda.synchronize()
    assert torch.equal(got_idx, idx)
    assert int(idx.max()) < size and int(idx.min()) >= 0
    for a, b in zip(got, want):
        assert torch.equal(a, b)
