"""GPU: K3 (fused PPO loss fwd/bwd), K2 (lambda-returns / GAE scans) and the stand-alone V-trace drop-in
through the C ABI vs the CPU oracle."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_close, assert_grad_parity, to_cuda
from impala_b200 import losses, ops, synth, vtrace
from impala_b200.agents import ppo as ppo_agent
from oracle import losses_ref, rlego_ref

pytestmark = pytest.mark.gpu
PPO_KEYS = ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio")


def oracle_ppo(d, ec=.01, clip=.1, dtype=torch.float32):
    lg, v = d["logits"].detach().to(dtype).clone().requires_grad_(True), d["v"].detach().to(dtype).clone().requires_grad_(True)
    loss, m = losses_ref.ppo_loss_from_outputs(lg, v, d["a"], d["v_target"].to(dtype), d["pi_ref"].to(dtype), ec, clip)
    loss.backward()
    return m, lg.grad, v.grad


@pytest.mark.parametrize("N,A", [(1000, 15), (1, 1), (255, 6), (256, 15), (257, 15), (4099, 8), (513, 17), (300, 32)])
@pytest.mark.parametrize("ec,clip", [(.01, .1), (0., .3)])
def test_ppo_parity(N, A, ec, clip):
    d = synth.ppo_minibatch(N, A, seed=N + A)
    m, dl_ref, dv_ref = oracle_ppo(d, ec, clip)
    g = to_cuda(d)
    lg, v = g["logits"].requires_grad_(True), g["v"].requires_grad_(True)
    loss, mm = losses.ppo_loss_from_outputs(lg, v, g["a"], g["v_target"], g["pi_ref"], ec, clip)
    assert list(mm) == list(PPO_KEYS)
    loss.backward()
    for k in PPO_KEYS:
        assert_close(mm[k], m[k], k, rtol=1e-5, atol=1e-6)
    assert_close(lg.grad, dl_ref, "dlogits", rel_to_max=2e-6)
    assert_close(v.grad, dv_ref, "dv", rel_to_max=2e-6)
    _, dl64, dv64 = oracle_ppo(d, ec, clip, dtype=torch.float64)
    assert_grad_parity(lg.grad, dl_ref, N, "dlogits x N", dl64)
    assert_grad_parity(v.grad, dv_ref, N, "dv x N", dv64)


def test_ppo_golden_and_clip_branches():
    z = np.load(os.path.join(GOLDEN, "ppo_small.npz"))
    d = {k: torch.from_numpy(z[k]) for k in ("logits", "v", "a", "v_target", "pi_ref")}
    g = to_cuda(d)
    sc, dl, dv = ops.ppo_loss_fwdbwd(g["logits"], g["v"], g["a"], g["v_target"], g["pi_ref"], .01, .1)
    assert sc[ops.S_LOSS].item() == pytest.approx(z["metrics"][0], rel=1e-5, abs=1e-6)
    assert_close(dl, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(dv, torch.from_numpy(z["dv"]).reshape(-1), "dv", rel_to_max=2e-6)
    # ratio far outside the clip range on both sides, and exactly 1 (the tie inside the range)
    d2 = synth.ppo_minibatch(2048, 15, seed=9, ref_noise=2.0)
    d2["pi_ref"][:512] = d2["logits"][:512]
    m, dl_ref, dv_ref = oracle_ppo(d2)
    g = to_cuda(d2)
    sc, dl, dv = ops.ppo_loss_fwdbwd(g["logits"], g["v"], g["a"], g["v_target"], g["pi_ref"], .01, .1)
    assert_close(sc[ops.S_PG], m["train/pg"], "pg", rtol=1e-5, atol=1e-6)
    assert_close(dl, dl_ref, "dlogits", rel_to_max=2e-6)


def test_ppo_full_size_properties():
    """One 8-GPU shard of BASELINE config 3 (T=256, B=8192/8 -> N=262,144): determinism, shard-sum == global."""
    N = 262_144
    d = to_cuda(synth.ppo_minibatch(N, 15, seed=3000))
    a = ops.ppo_loss_fwdbwd(d["logits"], d["v"], d["a"], d["v_target"], d["pi_ref"])
    b = ops.ppo_loss_fwdbwd(d["logits"], d["v"], d["a"], d["v_target"], d["pi_ref"])
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    h = N // 2
    p0 = ops.ppo_loss_fwdbwd(d["logits"][:h], d["v"][:h], d["a"][:h], d["v_target"][:h], d["pi_ref"][:h], n_global=N)
    p1 = ops.ppo_loss_fwdbwd(d["logits"][h:], d["v"][h:], d["a"][h:], d["v_target"][h:], d["pi_ref"][h:], n_global=N)
    assert torch.equal(torch.cat([p0[1], p1[1]]), a[1]) and torch.equal(torch.cat([p0[2], p1[2]]), a[2])
    assert_close((p0[0] + p1[0])[:7], a[0][:7], "scalars", rtol=1e-5, atol=1e-6)
    m, dl_ref, _ = oracle_ppo({k: v.cpu() for k, v in d.items()})
    assert_close(a[0][ops.S_LOSS], m["train/loss"], "loss", rtol=1e-5, atol=1e-6)
    assert_close(a[1], dl_ref, "dlogits", rel_to_max=2e-6)


@pytest.mark.parametrize("n_shards", [1, 2, 4, 8])
def test_config3_shard_shapes(n_shards):
    """BASELINE config 3 (T=256, B=8192, A=15) at the per-GPU shard sizes of 1 / 2 / 4 / 8 GPUs: K2 (lambda-return scan)
    in both layouts against the oracle on a column sample, bit-identical between layouts' storage orders where the
    operation order is the oracle's; K3 (PPO loss) at N = T * B/G: sum of the G shards' scalars and the concatenation of
    their gradients == the un-sharded launch, and oracle parity of one shard."""
    T, Bg = 256, 8192 // n_shards
    x = synth.returns_rollout(Bg, T + 1, seed=3000 + n_shards)
    r, dsc, vn = x["r"][:, :T], x["discount"][:, :T], x["v"][:, 1:]
    cols = torch.arange(0, Bg, max(1, Bg // 64))
    want = oracle_returns(r[cols], dsc[cols], vn[cols], .95)
    got_bm = vtrace.lambda_returns(r.cuda(), dsc.cuda(), vn.cuda(), .95)                               # batch-major
    rt, dt, vt = (t.t().contiguous().cuda() for t in (r, dsc, vn))
    got_tm = vtrace.lambda_returns(rt, dt, vt, .95, time_major=True).t()                               # [T,B] storage
    assert torch.equal(got_bm.cpu()[cols], want), "batch-major scan: the oracle's operation order, bit exact"
    M = max(1.0, want.abs().max().item())
    assert_close(got_tm[cols.cuda()], want, "time-major scan", rtol=1e-5, atol=1e-6 * M, rel_to_max=2e-6)
    # K3 on this shard's N = T * Bg samples
    N = T * Bg
    d = to_cuda(synth.ppo_minibatch(N, 15, seed=3100 + n_shards))
    full = ops.ppo_loss_fwdbwd(d["logits"], d["v"].reshape(-1), d["a"], d["v_target"], d["pi_ref"])
    if n_shards == 8:       # the oracle on the CPU at the smallest shard (262,144 samples); the larger ones by composition
        m, dl_ref, dv_ref = oracle_ppo({k: v.cpu() for k, v in d.items()})
        for k, slot in (("train/loss", ops.S_LOSS), ("train/td", ops.S_TD), ("train/pg", ops.S_PG), ("train/entropy", ops.S_ENTROPY)):
            assert_close(full[0][slot], m[k], k, rtol=1e-5, atol=1e-6)
        assert_close(full[1], dl_ref, "dlogits", rel_to_max=2e-6)
        assert_grad_parity(full[1], dl_ref, N, "dlogits x N", oracle_ppo({k: v.cpu() for k, v in d.items()}, dtype=torch.float64)[1])
    parts = 4
    h = N // parts
    ps = [ops.ppo_loss_fwdbwd(d["logits"][i * h:(i + 1) * h], d["v"].reshape(-1)[i * h:(i + 1) * h], d["a"][i * h:(i + 1) * h],
                              d["v_target"][i * h:(i + 1) * h], d["pi_ref"][i * h:(i + 1) * h], n_global=N) for i in range(parts)]
    assert torch.equal(torch.cat([q[1] for q in ps]), full[1]) and torch.equal(torch.cat([q[2] for q in ps]), full[2])
    assert_close(sum(q[0] for q in ps)[:7], full[0][:7], "sharded scalars", rtol=1e-5, atol=1e-6)


def test_ppo_reference_signature(cuda):
    d = to_cuda(synth.ppo_minibatch(64, 15, seed=1))
    w = torch.nn.Parameter(torch.zeros(16, device=cuda))
    model = lambda s: (d["logits"] + w[:15], d["v"] + w[15:])  # noqa: E731
    loss, m = losses.ppo_loss(model, (None, d["a"], d["v_target"], d["pi_ref"]), entropy_cost=.01, clip_coeff=.1)
    loss.backward()
    c = {k: v.cpu() for k, v in d.items()}
    wc = torch.zeros(16, requires_grad=True)
    l2, _ = losses_ref.ppo_loss(lambda s: (c["logits"] + wc[:15], c["v"] + wc[15:]),
                                (None, c["a"], c["v_target"], c["pi_ref"]))
    l2.backward()
    assert_close(w.grad, wc.grad, "dW", rtol=1e-4, atol=1e-6, rel_to_max=1e-4)


# ------------------------------------------------------------------ K2 scans
def oracle_returns(r, d, v, lam):
    return rlego_ref.batched(rlego_ref.lambda_returns, lambda_=lam)(r, d, v)


def oracle_gae(r, d, vals, lam):
    return rlego_ref.batched(lambda a, b, c: rlego_ref.truncated_generalized_advantage_estimation(a, b, lam, c))(r, d, vals)


@pytest.mark.parametrize("B,T", [(64, 256), (1, 1), (33, 7), (5, 128), (31, 129), (40, 300), (3, 600), (100, 20)])
@pytest.mark.parametrize("lam", [.95, 1.0, 0.0])
@pytest.mark.parametrize("time_major", [False, True])
def test_scans_parity(B, T, lam, time_major):
    x = synth.returns_rollout(B, T + 1, seed=B * 7 + T)
    r, d, v = x["r"][:, :T], x["discount"][:, :T], x["v"]
    want_lr, want_gae = oracle_returns(r, d, v[:, 1:], lam), oracle_gae(r, d, v, lam)
    if time_major:
        rg, dg, vg = (t.t().contiguous().cuda() for t in (r, d, v))          # [T,B] storage
        got_lr = vtrace.lambda_returns(rg, dg, vg[1:], lam, time_major=True).t()
        got_gae = vtrace.truncated_generalized_advantage_estimation(rg, dg, lam, vg, time_major=True).t()
    else:
        rg, dg, vg = r.cuda(), d.cuda(), v.cuda()
        got_lr = vtrace.lambda_returns(rg, dg, vg[:, 1:], lam)
        got_gae = vtrace.truncated_generalized_advantage_estimation(rg, dg, lam, vg)
    # a T-step fp32 recursion whose values reach magnitude M carries rounding error ~ eps * M: the
    # absolute tolerance scales with M (north_star's atol 1e-6 is for O(1) quantities)
    for name, got, want in (("lambda_returns", got_lr, want_lr), ("gae", got_gae, want_gae)):
        M = max(1.0, want.abs().max().item())
        assert_close(got, want, name, rtol=1e-5, atol=1e-6 * M, rel_to_max=2e-6)
    if not time_major:  # sequential kernel, same operation order and roundings as the oracle: bit exact
        assert torch.equal(got_lr.cpu(), want_lr) and torch.equal(got_gae.cpu(), want_gae)


def test_scans_golden_identity_and_ppo_targets():
    z = np.load(os.path.join(GOLDEN, "returns_small.npz"))
    r, d, v = (torch.from_numpy(z[k]).cuda() for k in ("r", "discount", "v"))
    lr = vtrace.lambda_returns(r[:, :-1], d[:, :-1], v[:, 1:], .95)
    gae = vtrace.truncated_generalized_advantage_estimation(r[:, :-1], d[:, :-1], .95, v)
    assert_close(lr, torch.from_numpy(z["lambda_returns"]), "golden lambda_returns", rel_to_max=1e-5)
    assert_close(gae, torch.from_numpy(z["gae"]), "golden gae", rel_to_max=1e-5)
    assert_close(gae + v[:, :-1], lr, "GAE + V == lambda-return", rtol=1e-5, atol=5e-6)
    # one 8-GPU shard of config 3, time-major: T=256, B=1024
    x = synth.returns_rollout(1024, 256, seed=3000)
    done = x["done"].t().contiguous().cuda()
    rt, vt = x["r"].t().contiguous().cuda(), x["v"].t().contiguous().cuda()
    tgt = ppo_agent.make_targets(rt, done, vt, .99, .95, time_major=True)
    assert tgt.shape == (255, 1024)
    want = torch.stack([losses_ref.ppo_make_targets(x["r"][b], x["done"][b], x["v"][b], .99, .95) for b in range(0, 1024, 64)])
    assert_close(tgt.t()[::64], want, "ppo targets", rtol=1e-5, atol=2e-6)
    one = vtrace.lambda_returns(x["r"][0, :-1].cuda(), x["discount"][0, :-1].cuda(), x["v"][0, 1:].cuda(), .95)
    assert_close(one, want[0], "1-D rollout", rtol=1e-5, atol=2e-6)


# ------------------------------------------------------------------ stand-alone V-trace drop-in
@pytest.mark.parametrize("B,T,lam", [(32, 20, 1.), (7, 5, .9), (1, 2, 0.), (300, 33, .5), (4096, 20, 1.), (5, 300, .95),
                                     (3, 257, 1.)])
def test_batched_vtrace_dropin_forward_backward(B, T, lam):
    d = synth.impala_rollout(B, T, 4, seed=B + T)
    rho = torch.exp(.7 * torch.randn(B, T, generator=torch.Generator().manual_seed(1)))
    kw = dict(lambda_=lam, clip_rho_threshold=.9, clip_pg_rho_threshold=1.1)
    vc, rc = d["values"].clone().requires_grad_(True), rho.clone().requires_grad_(True)
    want = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage, **kw)(
        vc[:, :-1], vc[:, 1:], d["r"][:, :-1], d["discount"][:, :-1], rc[:, :-1])
    wts = [torch.randn(B, T - 1, generator=torch.Generator().manual_seed(i)) for i in range(3)]
    sum((o * w).sum() for o, w in zip(want, wts)).backward()
    g = to_cuda(d)
    vg, rg = g["values"].clone().requires_grad_(True), rho.cuda().requires_grad_(True)
    got = vtrace.batched_vtrace(vg[:, :-1], vg[:, 1:], g["r"][:, :-1], g["discount"][:, :-1], rg[:, :-1], **kw)
    sum((o * w.cuda()).sum() for o, w in zip(got, wts)).backward()
    for name, a, b in zip(("adv", "err", "q"), got, want):
        assert_close(a, b, name, rtol=1e-5, atol=2e-6, rel_to_max=1e-5)
    assert_close(vg.grad, vc.grad, "d values", rtol=1e-5, atol=2e-6, rel_to_max=1e-5)
    assert_close(rg.grad, rc.grad, "d rho", rtol=1e-5, atol=2e-6, rel_to_max=1e-5)
    e2, a2, _ = vtrace.batched_vtrace(vg[:, :-1], vg[:, 1:], g["r"][:, :-1], g["discount"][:, :-1], rg[:, :-1],
                                      output_order="rlax", **kw)
    assert torch.equal(e2, got[1]) and torch.equal(a2, got[0])
    # per-sequence form, the function the reference vmaps
    one = vtrace.vtrace_td_error_and_advantage(vg[0, :-1], vg[0, 1:], g["r"][0, :-1], g["discount"][0, :-1], rg[0, :-1], **kw)
    assert_close(one[0], want[0][0], "per-sequence adv", rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("B,T", [(64, 20), (5, 3), (1000, 20), (7, 300)])
def test_vtrace_forward_time_major(B, T):
    """The stand-alone V-trace on time-major storage ([T,B] tensors handed over as [B,T] views) equals the batch-major call."""
    d = to_cuda(synth.impala_rollout(B, T, 4, seed=B + T))
    rho = torch.exp(.7 * torch.randn(B, T, generator=torch.Generator().manual_seed(2))).cuda()
    args = (d["values"][:, :-1], d["values"][:, 1:], d["r"][:, :-1], d["discount"][:, :-1], rho[:, :-1])
    want = ops.vtrace_fwd(*args, lambda_=.9, clip_rho=.9, clip_pg_rho=1.1)[:3]
    tm_args = [x.t().contiguous().t() for x in args]                  # same values, strides (1, B)
    got = ops.vtrace_fwd(*tm_args, lambda_=.9, clip_rho=.9, clip_pg_rho=1.1)[:3]
    for name, a, b in zip(("adv", "err", "q"), got, want):
        assert a.stride() == (1, B) or B == 1 or T == 2
        assert torch.equal(a.contiguous(), b.contiguous()), name


@pytest.mark.parametrize("B,T,n", [(64, 20, 3), (1, 1, 1), (5, 7, 1), (33, 129, 5), (4, 6, 6), (3, 5, 9), (100, 256, 3)])
@pytest.mark.parametrize("lam", [1.0, 0.9])
@pytest.mark.parametrize("time_major", [False, True])
def test_n_step_returns_bit_exact(B, T, n, lam, time_major):
    """N4: rlego.n_step_bootstrapped_returns (reference agents/dqn/learning.py:62), batched; bit-exact vs the oracle."""
    from oracle import rlego_ref
    d = synth.returns_rollout(B, T, seed=B * 100 + T + n)
    want = rlego_ref.batched_loop(rlego_ref.n_step_bootstrapped_returns, d["r"], d["discount"], d["v"], n=n, lambda_t=lam) \
        if n - 1 <= T else None
    if want is None:      # rlax pads with min(n-1, T): every target collapses to the padded window
        pytest.skip("n - 1 > T is outside the oracle restatement")
    g = {k: v.cuda() for k, v in d.items()}
    if time_major:
        got = vtrace.n_step_bootstrapped_returns(g["r"].t().contiguous(), g["discount"].t().contiguous(),
                                                 g["v"].t().contiguous(), n, lam, time_major=True).t()
    else:
        got = vtrace.n_step_bootstrapped_returns(g["r"], g["discount"], g["v"], n, lam)
    assert torch.equal(got.cpu(), want)
    one = vtrace.n_step_bootstrapped_returns(g["r"][0], g["discount"][0], g["v"][0], n, lam)   # reference's [T] form
    assert torch.equal(one.cpu(), want[0])


@pytest.mark.parametrize("B,N,M", [(256, 32, 32), (1, 1, 1), (7, 8, 64), (33, 200, 51), (5, 64, 300)])
@pytest.mark.parametrize("kappa", [1.0, 0.0, 0.3])
def test_quantile_huber_fwd_bwd(B, N, M, kappa):
    """N4b: reference models/quantile_layers.py:83-93, loss and gradient w.r.t. dist_src (as DistributionalApex
    uses it: .mul(weight).mean().backward(), agents/dqn/learning.py:209-210)."""
    g = torch.Generator().manual_seed(B * 31 + N + M)
    src = torch.randn(B, N, generator=g)
    tau = torch.rand(B, N, generator=g)
    tgt = torch.randn(B, M, generator=g)
    wts = torch.rand(B, generator=g)
    s_ref = src.clone().requires_grad_(True)
    want = losses_ref.quantile_regression_loss(s_ref, tau, tgt, kappa)
    want.mul(wts).mean().backward()
    s_gpu = src.cuda().requires_grad_(True)
    got = losses.quantile_regression_loss(s_gpu, tau.cuda(), tgt.cuda(), kappa)
    got.mul(wts.cuda()).mean().backward()
    assert_close(got, want, "loss", rtol=1e-5, atol=1e-6)
    assert_close(s_gpu.grad, s_ref.grad, "d dist_src", rtol=1e-5, atol=1e-7, rel_to_max=1e-5)


# ---- K7: ppo_loss_is (reference losses.py:62-89) ---------------------------------------------------------------
IS_KEYS = ("train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio")


def oracle_ppo_is(d, **kw):
    lg, v = d["logits"].clone().requires_grad_(True), d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_is_loss_from_outputs(lg, v, d["v_boot"], d["a"], d["r"], d["done"], d["pi_ref"], **kw)
    loss.backward()
    return m, lg.grad, v.grad


def test_ppo_is_wide_batch_of_long_rollouts_widens_its_tiles(cuda):
    """B / (columns per tuned CTA tile) > 16384 partial sums of the workspace: the launch takes wider tiles instead of
    refusing.  Columns are independent, so a launch on a column slice (tuned tiles) must give the same gradients."""
    B, T, A = 40000, 128, 6
    d = to_cuda(synth.ppo_is_rollout(B, T, A, seed=9))
    args = lambda sl: (d["logits"][sl], d["values"][sl], d["v_boot"][sl], d["a"][sl], d["r"][sl], d["done"][sl], d["pi_ref"][sl])
    full = [x.clone() for x in ops.ppo_is_loss_fwdbwd(*args(slice(0, B)))]
    for sl in (slice(0, 96), slice(B - 96, B), slice(20000, 20096)):
        part = ops.ppo_is_loss_fwdbwd(*[x.contiguous() for x in args(sl)], n_global_b=B)
        torch.cuda.synchronize()
        assert torch.equal(full[1][sl], part[1]) and torch.equal(full[2][sl], part[2])
    assert torch.isfinite(full[0][:7]).all()
    with pytest.raises(RuntimeError, match="split the batch"):
        ops.ppo_is_loss_fwdbwd(*[x[:, :1].expand(-1, 600, *x.shape[2:]).contiguous() if x.dim() > 1 else x
                                 for x in args(slice(0, 20000))])


@pytest.mark.parametrize("B,T,A", [(12, 20, 15), (1, 1, 1), (7, 1, 6), (33, 20, 15), (5, 33, 8), (9, 64, 17), (3, 256, 4),
                                   (300, 20, 32), (64, 7, 18)])
@pytest.mark.parametrize("kw", [{}, dict(entropy_cost=0., clip_coeff=.3, lambda_=1., gamma=.9)])
def test_ppo_is_parity(B, T, A, kw):
    d = synth.ppo_is_rollout(B, T, A, seed=B + T + A)
    m, dl_ref, dv_ref = oracle_ppo_is(d, **kw)
    g = to_cuda(d)
    lg, v = g["logits"].requires_grad_(True), g["values"].requires_grad_(True)
    loss, mm = losses.ppo_is_loss_from_outputs(lg, v, g["v_boot"], g["a"], g["r"], g["done"], g["pi_ref"], **kw)
    assert list(mm) == list(IS_KEYS)
    loss.backward()
    for k in IS_KEYS:
        assert_close(mm[k], m[k], k, rtol=1e-5, atol=1e-6)
    assert_close(lg.grad, dl_ref, "dlogits", rel_to_max=2e-6)
    assert_close(v.grad, dv_ref, "dv", rel_to_max=2e-6)


def test_ppo_is_golden_edges_and_wrapper():
    z = np.load(os.path.join(GOLDEN, "ppo_is_small.npz"))
    d = {k: torch.from_numpy(z[k]) for k in ("logits", "values", "v_boot", "a", "r", "done", "pi_ref")}
    g = to_cuda(d)
    sc, dl, dv = ops.ppo_is_loss_fwdbwd(g["logits"], g["values"], g["v_boot"], g["a"], g["r"], g["done"], g["pi_ref"])
    for i, slot in enumerate((ops.S_LOSS, ops.S_ENTROPY, ops.S_TD, ops.S_PG, ops.S_KL, ops.S_RATIO)):
        assert sc[slot].item() == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6)
    assert_close(dl, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(dv, torch.from_numpy(z["dvalues"]), "dv", rel_to_max=2e-6)
    # every step done / never done / identical policies (ratio == 1: the tie inside the clip range) / far off-policy
    for name, edit in (("all_done", lambda x: x["done"].fill_(True)), ("no_done", lambda x: x["done"].fill_(False)),
                       ("on_policy", lambda x: x["pi_ref"].copy_(x["logits"])),
                       ("off_policy", lambda x: x["pi_ref"].add_(torch.randn_like(x["pi_ref"]) * 2))):
        d2 = synth.ppo_is_rollout(40, 20, 15, seed=11)
        edit(d2)
        m, dl_ref, dv_ref = oracle_ppo_is(d2)
        g2 = to_cuda(d2)
        sc, dl, dv = ops.ppo_is_loss_fwdbwd(g2["logits"], g2["values"], g2["v_boot"], g2["a"], g2["r"], g2["done"], g2["pi_ref"])
        assert_close(sc[ops.S_LOSS], m["train/loss"], name + " loss", rtol=1e-5, atol=1e-6)
        assert_close(dl, dl_ref, name + " dlogits", rel_to_max=2e-6)
        assert_close(dv, dv_ref, name + " dv", rel_to_max=2e-6)
        if name == "on_policy":
            assert sc[ops.S_RATIO].item() == 1.0
    # float / uint8 done flags are accepted like the bool ones; shards add up to the global batch
    d3 = synth.ppo_is_rollout(64, 20, 15, seed=12)
    g3 = to_cuda(d3)
    args = lambda x, sl: (x["logits"][sl], x["values"][sl], x["v_boot"][sl], x["a"][sl], x["r"][sl])  # noqa: E731
    full = ops.ppo_is_loss_fwdbwd(*args(g3, slice(None)), g3["done"], g3["pi_ref"])
    f2 = ops.ppo_is_loss_fwdbwd(*args(g3, slice(None)), g3["done"].float(), g3["pi_ref"])
    assert all(torch.equal(x, y) for x, y in zip(full, f2))
    lo, hi = slice(0, 24), slice(24, 64)
    p0 = ops.ppo_is_loss_fwdbwd(*args(g3, lo), g3["done"][lo], g3["pi_ref"][lo], n_global_b=64)
    p1 = ops.ppo_is_loss_fwdbwd(*args(g3, hi), g3["done"][hi], g3["pi_ref"][hi], n_global_b=64)
    assert torch.equal(torch.cat([p0[1], p1[1]]), full[1]) and torch.equal(torch.cat([p0[2], p1[2]]), full[2])
    assert_close((p0[0] + p1[0])[:6], full[0][:6], "scalars", rtol=1e-5, atol=1e-6)
    # reference signature: ppo_loss_is(batch, model, ...) with loss.backward() reaching the parameters
    B, T, A = 6, 5, 4
    torch.manual_seed(0)
    net = torch.nn.Linear(8, A + 1).cuda()
    model = lambda s: (lambda o: (o[:, :A], o[:, A:]))(net(s))  # noqa: E731
    s, s1 = torch.randn(B, T, 8, device="cuda"), torch.randn(B, T, 8, device="cuda")
    d4 = to_cuda(synth.ppo_is_rollout(B, T, A, seed=13))
    batch = (s, d4["a"][..., None], d4["r"][..., None], s1, d4["done"], d4["pi_ref"])
    loss, mm = losses.ppo_loss_is(batch, model)
    loss.backward()
    gw = net.weight.grad.clone()
    net.zero_grad()
    cpu_net = torch.nn.Linear(8, A + 1)
    cpu_net.load_state_dict({k: v.cpu() for k, v in net.state_dict().items()})
    cmodel = lambda s: (lambda o: (o[:, :A], o[:, A:]))(cpu_net(s))  # noqa: E731
    l_ref, m_ref = losses_ref.ppo_loss_is(tuple(x.cpu() for x in batch), cmodel)
    l_ref.backward()
    assert_close(loss, l_ref, "loss", rtol=1e-5, atol=1e-6)
    assert_close(gw, cpu_net.weight.grad, "dW", rel_to_max=5e-5)


def test_ppo_is_full_size_properties():
    """T=20, B=4096, A=15 (the C2 shape): determinism, oracle parity of the scalars and of a slice of the gradients."""
    d = synth.ppo_is_rollout(4096, 20, 15, seed=14)
    g = to_cuda(d)
    a = ops.ppo_is_loss_fwdbwd(g["logits"], g["values"], g["v_boot"], g["a"], g["r"], g["done"], g["pi_ref"])
    b = ops.ppo_is_loss_fwdbwd(g["logits"], g["values"], g["v_boot"], g["a"], g["r"], g["done"], g["pi_ref"],
                               flags=ops.INPUTS_STABLE)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    m, dl_ref, dv_ref = oracle_ppo_is(d)
    for k, slot in zip(IS_KEYS, (ops.S_LOSS, ops.S_ENTROPY, ops.S_TD, ops.S_PG, ops.S_KL, ops.S_RATIO)):
        assert_close(a[0][slot], m[k], k, rtol=1e-5, atol=1e-6)
    assert_close(a[1], dl_ref, "dlogits", rel_to_max=2e-6)
    assert_close(a[2], dv_ref, "dv", rel_to_max=2e-6)


@pytest.mark.parametrize("N,A", [(200_004, 15), (151_556, 6), (1_000_000, 15), (262_144 + 252, 18)])
def test_ppo_streaming_kernel_matches_tile_kernel_on_ragged_large_minibatches(cuda, N, A, monkeypatch):
    """The persistent ring kernel (B200RL_K3_STREAM=2) against the tile kernel (=1) where the minibatch does not end on a
    256-sample chunk: gradients bit for bit (same per-sample code), scalars to the tolerance (another summation order)."""
    d = to_cuda(synth.ppo_minibatch(N, A, seed=N % 1000))
    outs = {}
    for impl in ("1", "2"):
        monkeypatch.setenv("B200RL_K3_STREAM", impl)
        outs[impl] = [x.clone() for x in ops.ppo_loss_fwdbwd(d["logits"], d["v"].reshape(-1), d["a"], d["v_target"], d["pi_ref"])]
        torch.cuda.synchronize()
    assert torch.equal(outs["1"][1], outs["2"][1]) and torch.equal(outs["1"][2], outs["2"][2])
    torch.testing.assert_close(outs["1"][0][:7], outs["2"][0][:7], rtol=1e-5, atol=1e-6)


def test_rlax_published_vectors_on_the_kernels(cuda):
    """The stand-alone V-trace kernel, the K2 scans and (through its td / pg metrics) the fused K1 kernel against the known
    answers of rlax's own unit tests (tests/rlax_vectors.py states the provenance)."""
    import rlax_vectors as rv
    C = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(cuda)
    rho, v_t = rv.rho_tm1(), rv.v_t()
    adv, err, q = vtrace.batched_vtrace(C(rv.VALUES), C(v_t), C(rv.REWARDS), C(rv.DISCOUNTS), C(rho), **rv.VTRACE_KW)
    np.testing.assert_allclose(err.cpu().numpy(), rv.EXPECTED_TD, rtol=1e-6, atol=1e-10)
    np.testing.assert_allclose(adv.cpu().numpy(), rv.EXPECTED_PG, rtol=1e-6, atol=1e-10)
    for tm in (False, True):
        f = (lambda x: C(x.T)) if tm else C
        got = vtrace.lambda_returns(f(rv.MS_R), f(rv.MS_DISCOUNT), f(rv.MS_V), rv.MS_LAMBDA, time_major=tm)
        np.testing.assert_allclose((got.t() if tm else got).cpu().numpy(), rv.EXPECTED_LAMBDA_RETURNS, rtol=1e-6)
        for lam, want in rv.EXPECTED_GAE.items():
            got = vtrace.truncated_generalized_advantage_estimation(f(rv.GAE_R), f(rv.GAE_DISCOUNT), lam, f(rv.GAE_V), time_major=tm)
            np.testing.assert_allclose((got.t() if tm else got).cpu().numpy(), want, atol=1e-3)
    # rlax's test_lambda_q_estimate on the kernel: vs = v + min(1, rho) (q - v) at lambda = 0.8 (fixes the q bootstrap)
    adv8, err8, q8 = vtrace.batched_vtrace(C(rv.VALUES), C(v_t), C(rv.REWARDS), C(rv.DISCOUNTS), C(rho), lambda_=0.8)
    vs_from_q = C(rv.VALUES) + torch.clamp(C(rho), max=1.0) * (q8 - C(rv.VALUES))
    np.testing.assert_allclose((err8 + C(rv.VALUES)).cpu().numpy(), vs_from_q.cpu().numpy(), rtol=1e-5, atol=1e-5)
    # K1: the same rollouts as [B, T=5] learner inputs (values carry the bootstrap as their last step; the last step's
    # logits / reward / discount do not enter V-trace).  train/td = mean(err^2), train/pg = mean(log pi(a) * adv), :155-156
    pad = lambda x, fill: np.concatenate([x, np.full_like(x[:, :1], fill)], axis=1)
    logits, pi_ref = pad(rv.TARGET_LOGITS, 0.5), pad(rv.BEHAVIOR_LOGITS, 0.25)
    a = pad(rv.ACTIONS, 0)
    values = np.concatenate([rv.VALUES, rv.BOOTSTRAP[:, None]], axis=1)
    for impl in ("1", "2"):
        os.environ["B200RL_K1_TILE"] = impl
        try:
            sc, _, _ = ops.impala_loss_fwdbwd(C(logits), C(values), C(a), C(pad(rv.REWARDS, 7.0)), C(pad(rv.DISCOUNTS, 0.9)), C(pi_ref))
        finally:
            os.environ.pop("B200RL_K1_TILE")
        lp = torch.log_softmax(torch.from_numpy(rv.TARGET_LOGITS).double(), -1).gather(-1, torch.from_numpy(rv.ACTIONS)[..., None])[..., 0]
        want_td = float((rv.EXPECTED_TD.astype(np.float64) ** 2).mean())
        want_pg = float((lp.numpy() * rv.EXPECTED_PG.astype(np.float64)).mean())
        assert float(sc[ops.S_TD]) == pytest.approx(want_td, rel=2e-6)
        assert float(sc[ops.S_PG]) == pytest.approx(want_pg, rel=5e-6)      # rlax prints EXPECTED_PG to 8 digits; fp32 mean of 8 products
