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


# ----------------------------------------------------------------------------------------------------------------
# single-implementation kernels against the oracle at drawn shapes (the comparators of the hand-picked tests)
def _scan_cases(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        T = rng.choice([1, 2, 3, 31, 32, 33, 64, 127, 128, 256, 257, 511, 700]) if rng.random() < 0.6 else rng.randint(1, 700)
        B = rng.randint(1, 4096) if rng.random() < 0.7 else rng.randint(1, 40)
        out.append((B, T, rng.choice([0.0, 0.5, 0.95, 1.0]), rng.random() < 0.5))
    return out


@pytest.mark.parametrize("B,T,lam,time_major", _scan_cases(24 * _MORE, seed=5))
def test_k2_random_shapes_against_the_oracle(cuda, B, T, lam, time_major):
    from impala_b200 import vtrace
    from oracle import rlego_ref
    x = synth.returns_rollout(B, T + 1, seed=B * 7 + T)
    r, d, v = x["r"][:, :T], x["discount"][:, :T], x["v"]
    want_lr = rlego_ref.batched(rlego_ref.lambda_returns, lambda_=lam)(r, d, v[:, 1:])
    want_gae = rlego_ref.batched(lambda a, b, c: rlego_ref.truncated_generalized_advantage_estimation(a, b, lam, c))(r, d, v)
    if time_major:
        rg, dg, vg = (t.t().contiguous().cuda() for t in (r, d, v))
        got_lr = vtrace.lambda_returns(rg, dg, vg[1:], lam, time_major=True).t()
        got_gae = vtrace.truncated_generalized_advantage_estimation(rg, dg, lam, vg, time_major=True).t()
    else:
        rg, dg, vg = r.cuda(), d.cuda(), v.cuda()
        got_lr = vtrace.lambda_returns(rg, dg, vg[:, 1:], lam)
        got_gae = vtrace.truncated_generalized_advantage_estimation(rg, dg, lam, vg)
    for name, got, want in (("lambda_returns", got_lr, want_lr), ("gae", got_gae, want_gae)):
        M = max(1.0, want.abs().max().item())   # fp32 recursion: rounding error ~ eps * magnitude (test_scans_parity)
        assert_close(got, want, name, rtol=1e-5, atol=1e-6 * M, rel_to_max=2e-6)
    if not time_major:
        assert torch.equal(got_lr.cpu(), want_lr) and torch.equal(got_gae.cpu(), want_gae)


def _nstep_cases(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        T = rng.randint(1, 300)
        out.append((rng.randint(1, 300), T, rng.randint(1, min(T + 1, 12)), rng.choice([1.0, 0.9, 0.0]), rng.random() < 0.5))
    return out


@pytest.mark.parametrize("B,T,n,lam,time_major", _nstep_cases(16 * _MORE, seed=6))
def test_n4_random_shapes_bit_exact(cuda, B, T, n, lam, time_major):
    from impala_b200 import vtrace
    from oracle import rlego_ref
    d = synth.returns_rollout(B, T, seed=B * 100 + T + n)
    want = rlego_ref.batched_loop(rlego_ref.n_step_bootstrapped_returns, d["r"], d["discount"], d["v"], n=n, lambda_t=lam)
    g = {k: v.cuda() for k, v in d.items()}
    if time_major:
        got = vtrace.n_step_bootstrapped_returns(g["r"].t().contiguous(), g["discount"].t().contiguous(),
                                                 g["v"].t().contiguous(), n, lam, time_major=True).t()
    else:
        got = vtrace.n_step_bootstrapped_returns(g["r"], g["discount"], g["v"], n, lam)
    assert torch.equal(got.cpu(), want)


def _k1_layout_cases(n, seed):
    rng = random.Random(seed)
    return [(rng.randint(1, 6000), rng.choice([2, 5, 20, 20, 33, 64]), rng.choice([1, 4, 6, 15, 18, 32]),
             rng.choice([0.0, 0.9, 1.0])) for _ in range(n)]


@pytest.mark.parametrize("B,T,A,lam", _k1_layout_cases(16 * _MORE, seed=7))
def test_k1_random_shapes_time_major_and_stable_flag_equal_batch_major(cuda, B, T, A, lam):
    """[T,B,...] storage (the reference learner's layout before its transpose) and the B200RL_INPUTS_STABLE contract run
    other load paths of the same per-sample code: gradients bit for bit."""
    g = to_cuda(synth.impala_rollout(B, T, A, seed=B + 13 * T))
    base = [x.clone() for x in ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"],
                                                      lambda_=lam)]
    stable = ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"], lambda_=lam,
                                    flags=ops.INPUTS_STABLE)
    torch.cuda.synchronize()
    assert torch.equal(base[1], stable[1]) and torch.equal(base[2], stable[2])
    tm = {k: v.transpose(0, 1).contiguous() for k, v in g.items()}                      # [T,B,...] storage
    got = ops.impala_loss_fwdbwd(tm["logits"], tm["values"], tm["a"], tm["r"], tm["discount"], tm["pi_ref"], lambda_=lam,
                                 time_major=True)
    torch.cuda.synchronize()
    assert torch.equal(base[1], got[1].transpose(0, 1)) and torch.equal(base[2], got[2].transpose(0, 1))
    torch.testing.assert_close(base[0][:6], got[0][:6], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n", [random.Random(8).randint(1, 3_000_000) for _ in range(6 * _MORE)])
def test_k6_random_sizes_match_torch(cuda, n):
    g = torch.Generator(device=cuda).manual_seed(n)
    p0, g0 = torch.randn(n, generator=g, device=cuda), torch.randn(n, generator=g, device=cuda) * 0.01
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref], lr=1e-4, eps=1e-5)
    p, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    for k in range(2):
        ref.grad = g0 * (k + 1)
        want_norm = torch.nn.utils.clip_grad_norm_([ref], 0.05)
        opt.step()
        norm = ops.clip_adam_step(p, g0 * (k + 1), m, v, k + 1, 1e-4, 0.9, 0.999, 1e-5, 0.05)
        assert_close(norm[0], want_norm, "norm", rtol=1e-5, atol=0)
    assert_close(p, ref.detach(), "params", rtol=1e-6, atol=1e-7)
    assert_close(m, opt.state[ref]["exp_avg"], "exp_avg", rtol=2e-5, atol=1e-8)


# ----------------------------------------------------------------------------------------------------------------
# pointers that are only 4-byte aligned (views one element into an allocation): the bulk-copy / float4 paths must
# step aside for the scalar ones and give the same bits
def _off1(x):
    """Same values in storage that starts one element (4 or 8 bytes) into a fresh allocation."""
    buf = torch.empty(x.numel() + 1, dtype=x.dtype, device=x.device)
    v = buf[1:].view(x.shape)
    v.copy_(x)
    return v


def _misaligned_k1(n, seed):
    rng = random.Random(seed)
    return [(rng.randint(1, 3000), rng.choice([2, 7, 20, 20, 33]), rng.choice([1, 3, 6, 15, 18]), rng.random() < 0.5)
            for _ in range(n)]


@pytest.mark.parametrize("B,T,A,time_major", _misaligned_k1(12 * _MORE, seed=9))
def test_k1_misaligned_pointers_equal_aligned(cuda, B, T, A, time_major):
    g = to_cuda(synth.impala_rollout(B, T, A, seed=3 * B + T))
    if time_major:
        g = {k: v.transpose(0, 1).contiguous() for k, v in g.items()}
    order = ("logits", "values", "a", "r", "discount", "pi_ref")
    want = [x.clone() for x in ops.impala_loss_fwdbwd(*[g[k] for k in order], time_major=time_major)]
    out = (torch.empty(8, device=cuda), _off1(torch.empty_like(g["logits"])), _off1(torch.empty_like(g["values"])))
    got = ops.impala_loss_fwdbwd(*[_off1(g[k]) for k in order], time_major=time_major, out=out)
    torch.cuda.synchronize()
    assert torch.equal(want[1], got[1]) and torch.equal(want[2], got[2])
    torch.testing.assert_close(want[0][:6], got[0][:6], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("N,A", [(random.Random(10 + i).randint(1, 400_000), random.Random(20 + i).choice([1, 3, 6, 15, 18]))
                                 for i in range(8 * _MORE)])
def test_k3_misaligned_pointers_equal_aligned(cuda, N, A):
    d = to_cuda(synth.ppo_minibatch(N, A, seed=N % 977))
    args = (d["logits"], d["v"].reshape(-1), d["a"], d["v_target"], d["pi_ref"])
    want = [x.clone() for x in ops.ppo_loss_fwdbwd(*args)]
    got = ops.ppo_loss_fwdbwd(*[_off1(x) for x in args])
    torch.cuda.synchronize()
    assert torch.equal(want[1], got[1]) and torch.equal(want[2], got[2])
    torch.testing.assert_close(want[0][:7], got[0][:7], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("B,T,time_major", [(random.Random(30 + i).randint(1, 3000), random.Random(40 + i).choice([1, 5, 32, 64, 256, 300]),
                                             i % 2 == 0) for i in range(8 * _MORE)])
def test_k2_misaligned_pointers_equal_aligned(cuda, B, T, time_major):
    from impala_b200 import vtrace
    x = synth.returns_rollout(B, T + 1, seed=B + T)
    r, d, v = (t.cuda() for t in (x["r"][:, :T].contiguous(), x["discount"][:, :T].contiguous(), x["v"]))
    if time_major:
        r, d, v = (t.t().contiguous() for t in (r, d, v))
        f = lambda r, d, v: (vtrace.lambda_returns(r, d, v[1:], 0.95, time_major=True),
                             vtrace.truncated_generalized_advantage_estimation(r, d, 0.95, v, time_major=True))
    else:
        f = lambda r, d, v: (vtrace.lambda_returns(r, d, v[:, 1:], 0.95),
                             vtrace.truncated_generalized_advantage_estimation(r, d, 0.95, v))
    want = [y.clone() for y in f(r, d, v)]
    got = f(_off1(r), _off1(d), _off1(v))
    torch.cuda.synchronize()
    for a, b in zip(want, got):
        if time_major:   # the staged (16-byte aligned) and the chunked scan compose the steps in another order
            M = max(1.0, a.abs().max().item())
            assert_close(b, a, "scan", rtol=1e-5, atol=1e-6 * M, rel_to_max=2e-6)
        else:
            assert torch.equal(a, b)


# ----------------------------------------------------------------------------------------------------------------
# the rows beyond K1-K5 at drawn shapes
@pytest.mark.parametrize("cap,seed", [(random.Random(50 + i).choice([1, 2, 3, 17, 255, 256, 257, 1023, 4097, 70_001, 300_000]), i)
                                      for i in range(6 * _MORE)])
def test_sumtree_random_update_sequences_bit_exact(cuda, cap, seed):
    import numpy as np
    from oracle import replay_ref
    rng = np.random.default_rng(seed)
    ot = replay_ref.SumTree(cap)
    tree = torch.zeros(2 * ot.leaves, dtype=torch.float64, device=cuda)
    for _ in range(5):
        n = int(rng.integers(1, min(cap, 20_000) + 1))
        idx = rng.choice(cap, size=n, replace=False).astype(np.int64)
        val = rng.random(n) ** 3 + 1e-3
        if rng.random() < 0.3:
            val[: n // 3] = 0.0                                   # zero-priority leaves are never drawn
        ot.update(idx, val)
        ops.sumtree_update(tree, torch.from_numpy(idx).to(cuda), torch.from_numpy(val).to(cuda))
        assert np.array_equal(tree.cpu().numpy(), ot.tree)
        if ot.tree[1] > 0:
            off, draws = int(rng.integers(0, 2 ** 40)), int(rng.integers(1, 5000))
            want_idx, want_p = ot.sample(77, 1, off, draws, cap)
            got_idx, got_p = ops.sumtree_sample(tree, cap, 77, 1, off, draws)
            assert np.array_equal(got_idx.cpu().numpy(), want_idx) and np.array_equal(got_p.cpu().numpy(), want_p)


def _k7_cases(n, seed):
    rng = random.Random(seed)
    return [(rng.randint(1, 400), rng.choice([1, 2, 7, 20, 33, 128]), rng.choice([1, 4, 6, 15, 18, 32]),
             rng.choice([0.0, 0.9, 0.95, 1.0]), rng.choice([0.9, 0.99])) for _ in range(n)]


@pytest.mark.parametrize("B,T,A,lam,gamma", _k7_cases(10 * _MORE, seed=11))
def test_k7_random_shapes_against_the_oracle(cuda, B, T, A, lam, gamma):
    """As test_ppo_is_parity, minus the samples whose probability ratio sits within 1e-5 of a clip boundary (0.9, 1, 1.1):
    the gradient is discontinuous there and a last-bit difference of the ratio legitimately picks the other branch
    (one such sample in ~2.4 M drawn ones is what this test met before the mask)."""
    from oracle import losses_ref
    from impala_b200 import losses
    d = synth.ppo_is_rollout(B, T, A, seed=B * 3 + T)
    kw = dict(lambda_=lam, gamma=gamma)
    lg0, v0 = d["logits"].clone().requires_grad_(True), d["values"].clone().requires_grad_(True)
    loss0, m0 = losses_ref.ppo_is_loss_from_outputs(lg0, v0, d["v_boot"], d["a"], d["r"], d["done"], d["pi_ref"], **kw)
    loss0.backward()
    g = to_cuda(d)
    lg, v = g["logits"].requires_grad_(True), g["values"].requires_grad_(True)
    loss, m = losses.ppo_is_loss_from_outputs(lg, v, g["v_boot"], g["a"], g["r"], g["done"], g["pi_ref"], **kw)
    loss.backward()
    for k in m0:
        assert_close(m[k], m0[k], k, rtol=1e-5, atol=1e-6)
    lp = torch.log_softmax(d["logits"].double(), -1).gather(-1, d["a"].unsqueeze(-1)).squeeze(-1)
    lpr = torch.log_softmax(d["pi_ref"].double(), -1).gather(-1, d["a"].unsqueeze(-1)).squeeze(-1)
    ratio = (lp - lpr).exp()
    near = ((ratio - 0.9).abs() < 1e-5) | ((ratio - 1.0).abs() < 1e-5) | ((ratio - 1.1).abs() < 1e-5)
    if A == 1:
        near[:] = False       # ratio == 1 everywhere and every gradient w.r.t. the single logit is 0
    assert int(near.sum()) <= 5 + B * T // 2000        # ~6e-5 of the samples fall in the three windows
    keep = (~near).unsqueeze(-1).float()
    assert_close(lg.grad.cpu() * keep, lg0.grad * keep, "dlogits", rel_to_max=2e-6)
    # the V-trace target of step t-1 carries rho_t: a flipped branch at t also moves dv there -- compare the columns without any
    cols = ~near.any(dim=1)
    assert_close(v.grad.cpu()[cols], v0.grad[cols], "dv", rel_to_max=2e-6)


@pytest.mark.parametrize("B,N,M,kappa", [(random.Random(60 + i).randint(1, 300), random.Random(70 + i).randint(1, 200),
                                          random.Random(80 + i).randint(1, 200), random.Random(90 + i).choice([0.0, 0.3, 1.0]))
                                         for i in range(8 * _MORE)])
def test_quantile_huber_random_shapes(cuda, B, N, M, kappa):
    import importlib
    t = importlib.import_module("test_gpu_ppo_returns")
    t.test_quantile_huber_fwd_bwd(B, N, M, kappa)


def _cnn_shapes(n, seed):
    rng = random.Random(seed)
    return [((rng.randint(1, 40), rng.choice([4, 8, 16, 32, 64]), rng.randint(1, 70), rng.randint(1, 70)),
             rng.choice([None, 1, 2, 3]), rng.random() < 0.5) for _ in range(n)]


@pytest.mark.parametrize("shape,levels,with_bias", _cnn_shapes(10 * _MORE, seed=12))
def test_k9_pool_random_shapes_match_torch(cuda, shape, levels, with_bias):
    import importlib
    importlib.import_module("test_gpu_cnn_glue").test_maxpool_forward_and_backward_match_torch(cuda, shape, levels, with_bias)


@pytest.mark.parametrize("shape,levels,_", _cnn_shapes(8 * _MORE, seed=13))
def test_k9_bias_act_random_shapes_match_torch(cuda, shape, levels, _):
    import importlib
    importlib.import_module("test_gpu_cnn_glue").test_bias_act_and_relu_bwd_add_match_torch(cuda, shape)


@pytest.mark.parametrize("shape", [(random.Random(100 + i).randint(1, 300), 3, random.Random(110 + i).randint(1, 90),
                                    random.Random(120 + i).randint(1, 90)) for i in range(6 * _MORE)])
def test_k9_obs_conversion_random_shapes(cuda, shape):
    g = torch.Generator(device=cuda).manual_seed(shape[0])
    obs = torch.randint(0, 256, shape, device=cuda, dtype=torch.uint8, generator=g)
    got = ops.obs_to_f32_nhwc(obs)
    assert got.is_contiguous(memory_format=torch.channels_last) and torch.equal(got, obs.float() / 255.)


@pytest.mark.parametrize("cap,seed", [(random.Random(130 + i).choice([1, 2, 7, 37, 64, 100, 1000]), i) for i in range(8 * _MORE)])
def test_ring_buffer_random_extend_sequences_match_the_oracle(cuda, cap, seed):
    """Random extend sizes (wraps, exact fills, more rows than the capacity) then a sample: storage, keys, rows and
    probabilities as oracle/replay_ref.py (the rlmeta CircularBuffer + UniformSampler semantics)."""
    import numpy as np
    from impala_b200 import replay
    from oracle import replay_ref
    rng = random.Random(seed)
    rb = replay.ReplayBuffer(replay.CircularBuffer(cap, device=cuda), replay.UniformSampler())
    ob = replay_ref.ReplayBuffer(replay_ref.CircularBuffer(cap), replay_ref.UniformSampler())
    rb.sampler.reset(seed)
    ob.sampler.reset(seed)
    g = torch.Generator().manual_seed(seed)
    for step in range(8):
        n = rng.choice([1, 1, 2, cap, cap + 1, 2 * cap + 3, rng.randint(1, 3 * cap)])
        obs = torch.randint(0, 255, (n, 5, 3), generator=g, dtype=torch.uint8)
        act = torch.randint(0, 15, (n, 1), generator=g)
        rew = torch.randn(n, 2, generator=g)
        rb.extend((obs, act, rew))
        if n <= cap:
            ob.extend([obs.numpy(), act.numpy(), rew.numpy()])
        else:                                                       # oracle: only the survivors matter
            ob.storage.count += n - cap
            ob.extend([obs[n - cap:].numpy(), act[n - cap:].numpy(), rew[n - cap:].numpy()])
        assert rb.size == ob.storage.size and rb.info() == (cap, ob.storage.size)
        occupied = min(rb.storage.count, cap)
        for f, of in zip(rb.storage.fields, ob.storage.fields):
            assert np.array_equal(f.cpu().numpy()[:occupied], of[:occupied])
        draws = rng.randint(1, 200)
        keys, batch, probs = rb.sample(draws)
        okeys, obatch, oprobs = ob.sample(draws)
        assert np.array_equal(keys.cpu().numpy(), okeys)
        assert all(np.array_equal(b.cpu().numpy(), o) for b, o in zip(batch, obatch))
        assert np.allclose(probs.cpu().numpy(), oprobs)


def _conv_cases(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        W = rng.choice([8, 16, 32, 64])
        H = rng.choice([1, 2, 3, 7, 8, 9, 15, 16, 17, 31, 33, 64]) if rng.random() < 0.7 else rng.randint(1, 70)
        N = rng.randint(1, max(1, min(400, 600_000 // (H * W))))
        cin, cout = rng.choice([3, 16, 16, 32]), rng.choice([16, 32])
        if not ops.conv3x3_wgrad_supported((N, cin, H, W), cout):   # 64 columns x many channels: stays on cuDNN
            W = 32
        out.append((N, cin, cout, H, W))
    return out


@pytest.mark.parametrize("N,Cin,Cout,H,W", _conv_cases(12 * _MORE, seed=14))
def test_k10_k11_random_shapes(cuda, N, Cin, Cout, H, W):
    """K10 / K11 (plain and fused epilogue) at drawn image counts and heights: TF32 tolerance vs cuDNN fp32, bit-identical on
    TF32-exact integer data vs fp64 -- the comparators of tests/test_gpu_cnn_glue.py."""
    import importlib
    t = importlib.import_module("test_gpu_cnn_glue")
    t.test_conv3x3_forward_and_dgrad_match_cudnn_fp32(cuda, N, Cin, Cout, H, W)
    t.test_conv3x3_fused_epilogue_exact_on_integers(cuda, N, Cin, Cout, H, W)
    # the wgrad test ends with a W = 4 refusal built by slicing the last axis: it needs W >= 8, which every draw has
    t.test_conv3x3_wgrad_matches_cudnn_fp32(cuda, N, Cin, Cout, H, W)


def _param_sets(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        shapes = []
        for _ in range(rng.randint(1, 9)):
            kind = rng.random()
            if kind < 0.35:
                shapes.append((rng.choice([16, 32]), rng.choice([3, 16, 32]), 3, 3, rng.random() < 0.7))     # conv weight (+ channels_last?)
            elif kind < 0.6:
                shapes.append((rng.randint(1, 64),))
            elif kind < 0.7:
                shapes.append(())
            else:
                shapes.append((rng.randint(1, 300), rng.randint(1, 700)))
        out.append((tuple(shapes), rng.choice([0.0, 0.5, 40.0])))
    return out


@pytest.mark.parametrize("shapes,max_norm", _param_sets(8 * _MORE, seed=15))
def test_fused_clip_adam_class_random_parameter_sets_match_torch(cuda, shapes, max_norm):
    """FusedClipAdam.step() over drawn parameter lists (scalars, vectors, matrices, channels_last convolution weights whose
    gradients arrive in either memory format) == clip_grad_norm_ + torch.optim.Adam on copies."""
    from impala_b200 import optim
    g = torch.Generator(device=cuda).manual_seed(len(shapes))
    mine, ref = [], []
    for sh in shapes:
        cl = len(sh) == 5 and sh[4]
        t = torch.randn(sh[:4] if len(sh) == 5 else sh, device=cuda, generator=g)
        if cl:
            t = t.contiguous(memory_format=torch.channels_last)
        mine.append(torch.nn.Parameter(t.clone(memory_format=torch.preserve_format)))
        ref.append(torch.nn.Parameter(t.clone(memory_format=torch.preserve_format)))
    opt = optim.FusedClipAdam(mine, lr=1e-3, eps=1e-5, max_grad_norm=max_norm)
    ropt = torch.optim.Adam(ref, lr=1e-3, eps=1e-5)
    for k in range(3):
        for i, (a, b) in enumerate(zip(mine, ref)):
            gr = torch.randn(a.shape, device=cuda, generator=g) * (0.1 + k)
            if a.dim() == 4 and (i + k) % 2 == 0:
                gr = gr.contiguous(memory_format=torch.channels_last)
            a.grad, b.grad = gr.clone(memory_format=torch.preserve_format), gr.clone(memory_format=torch.preserve_format)
        if max_norm > 0:
            want_norm = torch.nn.utils.clip_grad_norm_(ref, max_norm)
        ropt.step()
        opt.step()
        if max_norm > 0:
            assert_close(opt.total_norm[0], want_norm, "norm", rtol=1e-5, atol=0)
    for i, (a, b) in enumerate(zip(mine, ref)):
        assert a.shape == b.shape and a.stride() == b.stride()
        assert_close(a, b, "param %d %s" % (i, tuple(a.shape)), rtol=1e-5, atol=1e-6)
    sd = opt.state_dict()                       # and the checkpoint carries the moments in the parameters' own layout
    rsd = ropt.state_dict()
    for i in range(len(mine)):
        assert_close(sd["state"][i]["exp_avg"], rsd["state"][i]["exp_avg"], "exp_avg %d" % i, rtol=2e-5, atol=1e-7)


# ----------------------------------------------------------------------------------------------------------------
def test_empty_inputs_are_no_ops_or_clean_errors(cuda):
    """Zero rows / zero draws / zero columns through the public surface: replay calls are no-ops (rlmeta's extend([]) is),
    the losses and scans refuse with the library's message -- never a CUDA launch error that would poison the context."""
    from impala_b200 import replay, vtrace
    rb = replay.ReplayBuffer(replay.CircularBuffer(16, device=cuda), replay.UniformSampler())
    rb.sampler.reset(1)
    assert len(rb.extend([])) == 0
    rb.extend((torch.arange(5, device=cuda).reshape(5, 1).float(), torch.arange(5, device=cuda).reshape(5, 1)))
    assert len(rb.extend((torch.empty(0, 1, device=cuda), torch.empty(0, 1, dtype=torch.int64, device=cuda)))) == 0
    assert rb.size == 5
    keys, batch, probs = rb.sample(0)
    assert keys.numel() == 0 and batch[0].shape == (0, 1) and probs.numel() == 0
    keys, batch, _ = rb.sample(3)                                   # the stream is where it was: draw offset 0
    assert torch.equal(keys, ops.sample_uniform(1, 0, 0, 3, 5, cuda))
    assert ops.sample_uniform(1, 0, 0, 0, 5, cuda).numel() == 0
    assert ops.gather_rows(torch.zeros(4, 3, device=cuda), torch.empty(0, dtype=torch.int64, device=cuda)).shape == (0, 3)
    tree = torch.zeros(16, dtype=torch.float64, device=cuda)
    ops.sumtree_update(tree, torch.empty(0, dtype=torch.int64, device=cuda), torch.empty(0, dtype=torch.float64, device=cuda))
    assert float(tree.sum()) == 0.0
    assert ops.act_epilogue(torch.empty(0, 6, device=cuda), torch.tensor([False]), 1, 0, 0).numel() == 0
    z = lambda *s: torch.zeros(*s, device=cuda)
    zi = lambda *s: torch.zeros(*s, dtype=torch.int64, device=cuda)
    for call in (lambda: ops.impala_loss_fwdbwd(z(0, 20, 6), z(0, 20), zi(0, 20), z(0, 20), z(0, 20), z(0, 20, 6)),
                 lambda: ops.impala_loss_fwdbwd(z(4, 1, 6), z(4, 1), zi(4, 1), z(4, 1), z(4, 1), z(4, 1, 6)),     # T = 1
                 lambda: ops.ppo_loss_fwdbwd(z(0, 6), z(0), zi(0), z(0), z(0, 6)),
                 lambda: vtrace.lambda_returns(z(0, 8), z(0, 8), z(0, 8), 0.9),
                 lambda: ops.clip_adam_step(z(0), z(0), z(0), z(0), 1, 1e-3)):
        with pytest.raises(RuntimeError):
            call()
    # the context is intact
    torch.cuda.synchronize()
    assert float(ops.impala_loss_fwdbwd(z(2, 3, 4), z(2, 3), zi(2, 3), z(2, 3), z(2, 3), z(2, 3, 4))[0][1]) > 0   # entropy = ln 4


@pytest.mark.parametrize("B,T,lam", [(random.Random(200 + i).randint(1, 600), random.Random(210 + i).randint(2, 300),
                                      random.Random(220 + i).choice([0.0, 0.5, 0.9, 1.0])) for i in range(8 * _MORE)])
def test_stand_alone_vtrace_random_shapes(cuda, B, T, lam):
    """batched_vtrace (forward, backward, output orders, per-sequence form) and its time-major storage path at drawn
    shapes -- the comparators of tests/test_gpu_ppo_returns.py."""
    import importlib
    t = importlib.import_module("test_gpu_ppo_returns")
    t.test_batched_vtrace_dropin_forward_backward(B, T, lam)
    t.test_vtrace_forward_time_major(B, T)


def _k1_switch_cases(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        lam, crho, cpg, ent = rng.choice([0.0, 0.5, 0.9, 1.0]), rng.choice([0.7, 1.0, 1.3]), rng.choice([0.8, 1.0, 1.5]), rng.choice([0.0, 0.01, 0.2])
        da, dr = rng.random() < 0.3, rng.random() < 0.3
        out.append((rng.randint(1, 400), rng.choice([2, 3, 11, 20, 20, 40]), rng.choice([1, 2, 6, 15, 18, 32]), rng.random() < 0.4,
                    rng.choice(["0", "1", "2"]), lam, crho, cpg, ent, da, dr))
    return out


@pytest.mark.parametrize("B,T,A,time_major,impl,lam,crho,cpg,ent,da,dr", _k1_switch_cases(16 * _MORE, seed=16))
def test_k1_random_semantic_switches_against_the_oracle(cuda, B, T, A, time_major, impl, lam, crho, cpg, ent, da, dr, monkeypatch):
    """Every K1 argument drawn at once (lambda, both clip thresholds, entropy coefficient, detach flags, layout, kernel):
    metrics and gradients against the fp32 oracle, fp64-arbitrated (tests/test_gpu_impala.py::check)."""
    import importlib
    monkeypatch.setenv("B200RL_K1_TILE", impl)
    t = importlib.import_module("test_gpu_impala")
    okw = dict(lambda_=lam, clip_rho_threshold=crho, clip_pg_rho_threshold=cpg, entropy_coeff=ent, detach_adv=da, detach_rho=dr)
    kkw = dict(lambda_=lam, clip_rho=crho, clip_pg_rho=cpg, entropy_coeff=ent,
               flags=(ops.DETACH_ADV if da else 0) | (ops.DETACH_RHO if dr else 0))
    t.check(synth.impala_rollout(B, T, A, seed=B * 31 + T), okw, kkw, time_major=time_major)


def _sampler_cases(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        off = rng.choice([0, 2 ** 32 - rng.randint(1, 2000), 2 ** 32, 2 ** 33 + rng.randint(0, 10 ** 6), rng.randint(0, 2 ** 62),
                          2 ** 62 - rng.randint(1, 5000)])
        size = rng.choice([1, 2, 3, rng.randint(1, 10 ** 6), 2 ** 31 - 1, 2 ** 31, 2 ** 32 + 1, rng.randint(1, 2 ** 62)])
        out.append((rng.randint(0, 2 ** 63 - 1), rng.randint(0, 2 ** 31 - 1), off, rng.randint(1, 5000), size))
    return out


@pytest.mark.parametrize("seed,stream,off,n,size", _sampler_cases(16 * _MORE, seed=17))
def test_k4_random_streams_offsets_and_sizes_bit_exact(cuda, seed, stream, off, n, size):
    """Draw offsets across the 2^32 counter-word boundary, 63-bit seeds, buffer sizes beyond 2^32: numpy and C oracles."""
    import numpy as np
    from oracle import c_oracle, replay_ref
    got = ops.sample_uniform(seed, stream, off, n, size, cuda).cpu().numpy()
    assert np.array_equal(got, replay_ref.sample_uniform(seed, stream, off, n, size))
    assert np.array_equal(got, c_oracle.sample_uniform(seed, stream, off, n, size))
    if size <= 5000:                                           # the fused draw + gather uses the same stream
        store = torch.arange(size, device=cuda).reshape(size, 1)
        idx, (rows,) = ops.sample_gather_fields([store], seed, stream, off, n, size)
        assert np.array_equal(idx.cpu().numpy(), got) and torch.equal(rows.reshape(-1), idx)


@pytest.mark.parametrize("N,A,beta", [(random.Random(300 + i).randint(1, 20000), random.Random(310 + i).choice([1, 2, 4, 6, 18, 33, 51, 200]),
                                       random.Random(320 + i).choice([0.0, 0.4, 0.6, 1.0])) for i in range(8 * _MORE)])
def test_per_td_loss_random_shapes(cuda, N, A, beta):
    import importlib
    importlib.import_module("test_gpu_replay_learner").test_per_td_loss_matches_the_reference_lines(cuda, N, A, beta)
