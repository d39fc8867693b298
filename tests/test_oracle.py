"""CPU: pins the oracle -- Random123 known-answer vectors, hand-derived V-trace / lambda-return cases,
algebraic properties, C restatement vs torch restatement, committed golden vectors."""
import json
import os

import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from closed_form import impala_closed_form, ppo_closed_form
from conftest import GOLDEN
from impala_b200 import synth
from oracle import c_oracle, losses_ref, replay_ref, rlego_ref

# Random123 kat_vectors, philox4x32 10 rounds
PHILOX_KAT = [
    ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
     [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


@pytest.mark.parametrize("ctr,key,want", PHILOX_KAT)
def test_philox_known_answers(ctr, key, want):
    assert list(replay_ref.philox4x32_10(ctr, key)) == want
    assert list(c_oracle.philox4x32_10(ctr, key)) == want
    v = replay_ref.philox4x32_10_vec(*[np.array([c]) for c in ctr], key[0], key[1])
    assert [int(x[0]) for x in v] == want


def test_sampler_python_vs_c_and_range():
    for seed, stream, off, n, size in [(123, 0, 0, 4096, 1_000_000), (2**40 + 7, 5, 2**33, 1000, 3),
                                       (0, 0, 0, 17, 1), (9, 1, 10**12, 100, 2**40)]:
        a = replay_ref.sample_uniform(seed, stream, off, n, size)
        b = c_oracle.sample_uniform(seed, stream, off, n, size)
        assert np.array_equal(a, b)
        assert a.min() >= 0 and a.max() < size
    # a call split in two draws the same stream
    whole = replay_ref.sample_uniform(5, 0, 0, 100, 1000)
    s = replay_ref.UniformSampler()
    s.reset(5)
    assert np.array_equal(np.concatenate([s.sample(40, 1000), s.sample(60, 1000)]), whole)
    # roughly uniform
    big = replay_ref.sample_uniform(1, 0, 0, 200_000, 10)
    assert np.abs(np.bincount(big, minlength=10) / 200_000 - 0.1).max() < 0.005
    assert replay_ref.sample_uniform(1, 0, 0, 0, 10).shape == (0,)


def test_ring_buffer_semantics():
    rb = replay_ref.ReplayBuffer(replay_ref.CircularBuffer(5), replay_ref.UniformSampler())
    rb.sampler.reset(3)
    keys = rb.extend([np.arange(7, dtype=np.int64)[:, None], np.arange(7, dtype=np.float32)])
    assert list(keys) == list(range(7))
    assert rb.storage.size == 5
    assert sorted(rb.storage.fields[0][:, 0].tolist()) == [2, 3, 4, 5, 6]   # oldest two evicted
    assert rb.storage.fields[0][0, 0] == 5 and rb.storage.fields[0][1, 0] == 6  # slot = key % capacity
    k, batch, p = rb.sample(64)
    assert np.array_equal(batch[0][:, 0], k)        # row content == its key
    assert np.allclose(p, 0.2)
    with pytest.raises(RuntimeError):
        replay_ref.sample_uniform(0, 0, 0, 4, 0)


def test_vtrace_hand_cases():
    f = rlego_ref.vtrace_td_error_and_advantage
    t = lambda *x: torch.tensor(x, dtype=torch.float64)  # noqa: E731
    # T=1: td = rho_bar (r + g v' - v)
    adv, err, q = f(t(1.), t(2.), t(.5), t(.9), t(1.))
    assert err.item() == pytest.approx(1.3) and q.item() == pytest.approx(2.3) and adv.item() == pytest.approx(1.3)
    adv, err, q = f(t(1.), t(2.), t(.5), t(.9), t(2.))          # rho > 1 is clipped to 1
    assert err.item() == pytest.approx(1.3) and adv.item() == pytest.approx(1.3)
    adv, err, q = f(t(1.), t(2.), t(.5), t(.9), t(.5))          # rho < 1 scales both
    assert err.item() == pytest.approx(.65) and adv.item() == pytest.approx(.65) and q.item() == pytest.approx(2.3)
    # T=2, on-policy: vs_0 - v_0 = r0 + g r1 + g^2 v2 - v0
    adv, err, q = f(t(1., 2.), t(2., 3.), t(.5, .25), t(.9, .9), t(1., 1.))
    assert err[1].item() == pytest.approx(.25 + .9 * 3 - 2)
    assert err[0].item() == pytest.approx(.5 + .9 * .25 + .81 * 3 - 1)
    assert q[0].item() == pytest.approx(.5 + .9 * (2 + err[1].item()))    # bootstraps vs_1
    assert q[1].item() == pytest.approx(.25 + .9 * 3)                     # last step bootstraps v_t
    # gamma = 0 (episode ends everywhere): no propagation
    adv, err, q = f(t(1., 2.), t(2., 3.), t(.5, .25), t(0., 0.), t(.5, 3.))
    assert err.tolist() == pytest.approx([.5 * (.5 - 1), 1. * (.25 - 2)])
    # lambda = 0: one-step TD errors only
    adv, err, q = f(t(1., 2.), t(2., 3.), t(.5, .25), t(.9, .9), t(1., 1.), lambda_=0.)
    assert err[0].item() == pytest.approx(.5 + .9 * 2 - 1)
    assert q[0].item() == pytest.approx(.5 + .9 * 2)                      # (1-l) v_tm1[1] = 2
    # output order switch
    e2, a2, q2 = f(t(1.), t(2.), t(.5), t(.9), t(.5), output_order="rlax")
    assert e2.item() == pytest.approx(.65) and a2.item() == pytest.approx(.65)


def test_lambda_returns_hand_cases():
    t = lambda *x: torch.tensor(x, dtype=torch.float64)  # noqa: E731
    g = rlego_ref.lambda_returns(t(1., 2.), t(.5, .5), t(10., 20.), 1.0)        # MC + bootstrap
    assert g.tolist() == pytest.approx([1 + .5 * (2 + .5 * 20), 2 + .5 * 20])
    g = rlego_ref.lambda_returns(t(1., 2.), t(.5, .5), t(10., 20.), 0.0)        # one-step
    assert g.tolist() == pytest.approx([1 + .5 * 10, 2 + .5 * 20])
    g = rlego_ref.lambda_returns(t(1., 2.), t(0., .5), t(10., 20.), 0.95)       # done at t=0 cuts the trace
    assert g[0].item() == pytest.approx(1.)


@settings(max_examples=30, deadline=None)
@given(st.integers(1, 12), st.floats(0., 1.), st.integers(0, 2**31 - 1))
def test_gae_plus_v_is_lambda_return(T, lam, seed):
    g = torch.Generator().manual_seed(seed)
    r, v = torch.randn(T, generator=g, dtype=torch.float64), torch.randn(T + 1, generator=g, dtype=torch.float64)
    d = (torch.rand(T, generator=g) > .2).double() * .97
    gae = rlego_ref.truncated_generalized_advantage_estimation(r, d, lam, v)
    lr = rlego_ref.lambda_returns(r, d, v[1:], lam)
    assert torch.allclose(gae + v[:-1], lr, atol=1e-10)


@settings(max_examples=20, deadline=None)
@given(st.integers(2, 10), st.integers(0, 2**31 - 1))
def test_vtrace_onpolicy_is_nstep_and_permutation_equivariant(T, seed):
    g = torch.Generator().manual_seed(seed)
    B = 4
    v = torch.randn(B, T + 1, generator=g, dtype=torch.float64)
    r = torch.randn(B, T, generator=g, dtype=torch.float64)
    d = (torch.rand(B, T, generator=g) > .2).double() * .9
    fn = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage)
    adv, err, q = fn(v[:, :-1], v[:, 1:], r, d, torch.ones(B, T, dtype=torch.float64))
    # rho == 1, lambda == 1: vs_t = lambda-return with lambda 1 (n-step bootstrapped return)
    ret = rlego_ref.batched(rlego_ref.lambda_returns)(r, d, v[:, 1:])
    assert torch.allclose(err + v[:, :-1], ret, atol=1e-10)
    perm = torch.randperm(B, generator=g)
    adv2, err2, q2 = fn(v[perm, :-1], v[perm, 1:], r[perm], d[perm], torch.ones(B, T, dtype=torch.float64))
    assert torch.equal(adv2, adv[perm]) and torch.equal(err2, err[perm]) and torch.equal(q2, q[perm])
    # vmap path == Python-loop fallback (reference learning.py:18 vs :20-26)
    adv3, err3, q3 = rlego_ref.batched_loop(rlego_ref.vtrace_td_error_and_advantage, v[:, :-1], v[:, 1:], r, d,
                                            torch.ones(B, T, dtype=torch.float64))
    assert torch.allclose(adv3, adv) and torch.allclose(err3, err)


def test_c_scans_match_torch_oracle():
    d = synth.impala_rollout(9, 13, 4, 77)
    rho = torch.exp(.5 * torch.randn(9, 12))
    for lam, cr, cp in [(1., 1., 1.), (.9, .8, 1.2)]:
        want = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage, lambda_=lam, clip_rho_threshold=cr,
                                 clip_pg_rho_threshold=cp)(d["values"][:, :-1], d["values"][:, 1:], d["r"][:, :-1],
                                                           d["discount"][:, :-1], rho)
        got = c_oracle.vtrace(d["values"][:, :-1].numpy(), d["values"][:, 1:].numpy(), d["r"][:, :-1].numpy(),
                              d["discount"][:, :-1].numpy(), rho.numpy(), lam, cr, cp)
        for w, g in zip(want, got):
            np.testing.assert_allclose(g, w.numpy(), rtol=2e-6, atol=2e-6)
    x = synth.returns_rollout(5, 40, 3)
    want = rlego_ref.batched(rlego_ref.lambda_returns, lambda_=.95)(x["r"], x["discount"], x["v"])
    np.testing.assert_allclose(c_oracle.lambda_returns(x["r"].numpy(), x["discount"].numpy(), x["v"].numpy(), .95),
                               want.numpy(), rtol=2e-6, atol=2e-6)
    vv = torch.cat([x["v"], x["v"][:, :1]], 1)
    want = rlego_ref.batched(lambda r, g, v: rlego_ref.truncated_generalized_advantage_estimation(r, g, .95, v))(
        x["r"], x["discount"], vv)
    np.testing.assert_allclose(c_oracle.gae(x["r"].numpy(), x["discount"].numpy(), vv.numpy(), .95), want.numpy(),
                               rtol=2e-6, atol=2e-6)
    src = np.random.default_rng(0).integers(0, 255, (50, 7, 3), dtype=np.uint8)
    idx = replay_ref.sample_uniform(1, 0, 0, 33, 50)
    assert np.array_equal(c_oracle.gather_rows(src, idx), src[idx])


@pytest.mark.parametrize("kw", [dict(), dict(lambda_=.9, clip_rho_threshold=.8, clip_pg_rho_threshold=1.2),
                                dict(detach_adv=True), dict(detach_rho=True)])
def test_closed_form_matches_autograd_fp64(kw):
    """The formulas the K1 kernel implements == autograd through the reference-shaped loss."""
    d = {k: (v.double() if v.is_floating_point() else v) for k, v in synth.impala_rollout(6, 8, 5, 11).items()}
    lg, vv = d["logits"].clone().requires_grad_(True), d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.impala_train_loss(lg, vv, d["a"], d["r"], d["discount"], d["pi_ref"], **kw)
    loss.backward()
    l2, m2, dl, dv = impala_closed_form(d["logits"], d["values"], d["a"], d["r"], d["discount"], d["pi_ref"], .01,
                                        kw.get("lambda_", 1.), kw.get("clip_rho_threshold", 1.),
                                        kw.get("clip_pg_rho_threshold", 1.), kw.get("detach_adv", False),
                                        kw.get("detach_rho", False))
    assert abs(loss.item() - l2.item()) < 1e-12
    assert (lg.grad - dl).abs().max() < 1e-14 and (vv.grad - dv).abs().max() < 1e-14
    for k in ("entropy", "td", "pg", "kl", "ratio"):
        assert abs(m["train/" + k].item() - m2[k].item()) < 1e-12


def test_ppo_closed_form_matches_autograd_fp64():
    d = {k: (v.double() if v.is_floating_point() else v) for k, v in synth.ppo_minibatch(400, 6, 5).items()}
    lg, vv = d["logits"].clone().requires_grad_(True), d["v"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_loss_from_outputs(lg, vv, d["a"], d["v_target"], d["pi_ref"])
    loss.backward()
    l2, m2, dl, dv = ppo_closed_form(d["logits"], d["v"], d["a"], d["v_target"], d["pi_ref"])
    assert abs(loss.item() - l2.item()) < 1e-12
    assert (lg.grad - dl).abs().max() < 1e-14 and (vv.grad.reshape(-1) - dv).abs().max() < 1e-14


def test_reference_loss_wrappers():
    """ppo_loss(model, batch, ...) / impala_loss(batch, model, ...) keep the reference signatures."""
    d = synth.ppo_minibatch(64, 5, 1)
    model = lambda s: (d["logits"], d["v"])  # noqa: E731
    loss, m = losses_ref.ppo_loss(model, (None, d["a"], d["v_target"], d["pi_ref"]), entropy_cost=.01, clip_coeff=.1)
    assert list(m) == ["train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio"]
    T, A = 12, 5
    g = torch.Generator().manual_seed(0)
    pi, v = torch.randn(T, A, generator=g), torch.randn(T, 1, generator=g)
    batch = (None, torch.randint(0, A, (T, 1), generator=g), torch.randn(T, 1, generator=g),
             torch.full((T, 1), .99), None)
    loss, m = losses_ref.impala_loss(batch, lambda s: (pi, v))
    assert list(m) == ["train/loss", "train/td", "train/pg", "train/ratio"] and m["train/ratio"].item() == 1.0
    # as written: v_t is the same-step value, so with rho = 1 the one-step TD is r + g v - v
    r, gam = batch[2][:, 0], batch[3][:, 0]
    e = torch.zeros(())
    errs = []
    for t in reversed(range(T)):
        e = (r[t] + gam[t] * v[t, 0] - v[t, 0]) + gam[t] * e
        errs.insert(0, e)
    assert m["train/td"].item() == pytest.approx(.5 * torch.stack(errs).pow(2).mean().item(), rel=1e-5)
    tgt = losses_ref.ppo_make_targets(torch.ones(5), torch.zeros(5, dtype=torch.bool), torch.zeros(5), .5, 1.)
    assert tgt.tolist() == pytest.approx([1.875, 1.75, 1.5, 1.0])


def test_golden_vectors_match_oracle():
    z = np.load(os.path.join(GOLDEN, "impala_c1.npz"))
    t = {k: torch.from_numpy(z[k]) for k in ("logits", "values", "a", "r", "discount", "pi_ref")}
    lg, vv = t["logits"].clone().requires_grad_(True), t["values"].clone().requires_grad_(True)
    loss, m = losses_ref.impala_train_loss(lg, vv, t["a"], t["r"], t["discount"], t["pi_ref"])
    loss.backward()
    np.testing.assert_allclose(lg.grad.numpy(), z["dlogits"], rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(vv.grad.numpy(), z["dvalues"], rtol=1e-5, atol=1e-8)
    assert m["train/loss"].item() == pytest.approx(z["metrics"][0], rel=1e-5)
    z2 = np.load(os.path.join(GOLDEN, "impala_small_lambda.npz"))
    assert json.loads(str(z2["kw"]))["lambda_"] == 0.9
    zr = np.load(os.path.join(GOLDEN, "replay_indices.npz"))
    assert np.array_equal(zr["idx_seed123_1M"], replay_ref.sample_uniform(123, 0, 0, 4096, 1_000_000))
    assert np.array_equal(zr["idx_stream3_big_offset"], c_oracle.sample_uniform(123, 3, 10_000_000_000, 257, 1000))


def test_sumtree_spec_properties():
    """N3 specification (oracle/replay_ref.SumTree): every node is the sum of its children, the root is the
    total, sampling is proportional to the leaves and never returns an empty slot or an index >= size."""
    rng = np.random.default_rng(1)
    t = replay_ref.SumTree(10)
    assert t.leaves == 16
    t.update(np.arange(10), rng.random(10))
    t.update(np.array([3, 7]), np.array([0.0, 5.0]))
    for node in range(1, t.leaves):
        assert t.tree[node] == t.tree[2 * node] + t.tree[2 * node + 1]
    assert np.isclose(t.tree[1], t.tree[t.leaves:].sum())
    idx, prob = t.sample(9, 1, 0, 20000, 10)
    assert idx.min() >= 0 and idx.max() < 10 and not (idx == 3).any()
    freq = np.bincount(idx, minlength=10) / idx.size
    assert np.allclose(freq, t.tree[t.leaves:t.leaves + 10] / t.tree[1], atol=1.5e-2)
    assert np.allclose(prob, (t.tree[t.leaves + idx] / t.tree[1]).astype(np.float32))
    i2, _ = t.sample(9, 1, 0, 100, 10)
    assert np.array_equal(i2, idx[:100])                      # counter-based: same draws, same answers


def test_golden_next_rows_match_oracle():
    z = np.load(os.path.join(GOLDEN, "next_rows.npz"))
    r, d, v = (torch.from_numpy(z[k]) for k in ("r", "discount", "v"))
    ns = rlego_ref.batched_loop(rlego_ref.n_step_bootstrapped_returns, r, d, v, n=3)
    assert np.array_equal(ns.numpy(), z["n_step3"])
    # hand check of one element: 3-step return bootstrapping from v[t+2]
    b, t = 2, 5
    want = r[b, t] + d[b, t] * (r[b, t + 1] + d[b, t + 1] * (r[b, t + 2] + d[b, t + 2] * v[b, t + 2]))
    assert ns[b, t].item() == pytest.approx(want.item(), rel=1e-6)
    tree = replay_ref.SumTree(1000)
    tree.update(np.arange(1000), z["priorities"])
    idx, prob = tree.sample(123, 0, 0, 512, 1000)
    assert np.array_equal(idx, z["st_idx"]) and np.array_equal(prob, z["st_prob"])


def test_ppo_loss_is_restatement():
    """Reference ``losses.ppo_loss_is`` (``losses.py:62-89``) as restated in the oracle: golden fixture, the rolled
    mask, and the degenerate cases that can be worked out by hand."""
    z = np.load(os.path.join(GOLDEN, "ppo_is_small.npz"))
    t = {k: torch.from_numpy(z[k]) for k in ("logits", "values", "v_boot", "a", "r", "done", "pi_ref")}
    lg, vv = t["logits"].clone().requires_grad_(True), t["values"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_is_loss_from_outputs(lg, vv, t["v_boot"], t["a"], t["r"], t["done"], t["pi_ref"])
    loss.backward()
    assert list(m) == ["train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio"]
    np.testing.assert_allclose(np.array([m[k].item() for k in m]), z["metrics"], rtol=1e-5)
    np.testing.assert_allclose(lg.grad.numpy(), z["dlogits"], rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(vv.grad.numpy(), z["dvalues"], rtol=1e-5, atol=1e-8)
    # on-policy, never done, lambda = 1: err_t is the discounted sum of one-step TD errors up to the bootstrap,
    # the advantage equals it, and d loss / d v = -err / N
    B, T, A = 3, 6, 4
    d = synth.ppo_is_rollout(B, T, A, seed=5, p_done=0.)
    v = d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_is_loss_from_outputs(d["logits"], v, d["v_boot"], d["a"], d["r"], d["done"], d["logits"],
                                                  entropy_cost=0., lambda_=1., gamma=.9)
    loss.backward()
    vt = torch.cat([d["values"][:, 1:], d["v_boot"][:, None]], 1)
    delta = d["r"] + .9 * vt - d["values"]
    err = torch.zeros(B, T)
    e = torch.zeros(B)
    for k in reversed(range(T)):
        e = delta[:, k] + .9 * e
        err[:, k] = e
    assert m["train/ratio"].item() == 1.0 and m["train/kl"].item() == 0.0
    assert m["train/td"].item() == pytest.approx(.5 * err.pow(2).mean().item(), rel=1e-5)
    assert m["train/pg"].item() == pytest.approx(-err.mean().item(), rel=1e-4, abs=1e-6)
    np.testing.assert_allclose(v.grad.numpy(), (-err / (B * T)).numpy(), rtol=1e-4, atol=1e-7)
    # the mask is the not-done flag rolled by one step (wrapping): a done at the last step masks step 0
    done = torch.zeros(B, T, dtype=torch.bool)
    done[1, T - 1] = True
    done[2, 2] = True
    v2 = d["values"].clone().requires_grad_(True)
    losses_ref.ppo_is_loss_from_outputs(d["logits"], v2, d["v_boot"], d["a"], d["r"], done, d["pi_ref"])[0].backward()
    assert v2.grad[1, 0].item() == 0. and v2.grad[2, 3].item() == 0. and (v2.grad != 0).sum().item() == B * T - 2
    # reference signature
    model = lambda s: (d["logits"].reshape(-1, A), d["values"].reshape(-1, 1)) if s.shape[0] == B * T \
        else (None, d["v_boot"][:, None])  # noqa: E731
    s = torch.zeros(B, T, 1)
    l2, _ = losses_ref.ppo_loss_is((s, d["a"][..., None], d["r"][..., None], s, done, d["pi_ref"]), model)
    l3, _ = losses_ref.ppo_is_loss_from_outputs(d["logits"], d["values"], d["v_boot"], d["a"], d["r"], done, d["pi_ref"])
    assert l2.item() == l3.item()


def test_rlax_published_vectors():
    """The recursions against the known answers of rlax's own unit tests (tests/rlax_vectors.py states the provenance):
    V-trace TD errors to the last printed digit, policy-gradient advantages and lambda-returns to 1e-6, GAE to the
    rlax test's own atol.  This is an EXTERNAL pin of what `rlego_ref` restates."""
    import rlax_vectors as rv
    T = torch.from_numpy
    rho, v_t = rv.rho_tm1(), rv.v_t()
    for i in range(2):
        adv, err, q = rlego_ref.vtrace_td_error_and_advantage(T(rv.VALUES[i]), T(v_t[i]), T(rv.REWARDS[i]), T(rv.DISCOUNTS[i]),
                                                              T(rho[i]), **rv.VTRACE_KW)
        np.testing.assert_allclose(err.numpy(), rv.EXPECTED_TD[i], rtol=2e-7, atol=1e-11)
        np.testing.assert_allclose(adv.numpy(), rv.EXPECTED_PG[i], rtol=1e-6, atol=1e-10)
        got = rlego_ref.lambda_returns(T(rv.MS_R[i]), T(rv.MS_DISCOUNT[i]), T(rv.MS_V[i]), lambda_=rv.MS_LAMBDA)
        np.testing.assert_allclose(got.numpy(), rv.EXPECTED_LAMBDA_RETURNS[i], rtol=1e-6)
        for lam, want in rv.EXPECTED_GAE.items():
            got = rlego_ref.truncated_generalized_advantage_estimation(T(rv.GAE_R[i]), T(rv.GAE_DISCOUNT[i]), lam, T(rv.GAE_V[i]))
            np.testing.assert_allclose(got.numpy(), want[i], atol=1e-3)
    # batched form (what the reference vmaps) and the C restatement of the scans
    adv, err, q = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage, **rv.VTRACE_KW)(
        T(rv.VALUES), T(v_t), T(rv.REWARDS), T(rv.DISCOUNTS), T(rho))
    np.testing.assert_allclose(err.numpy(), rv.EXPECTED_TD, rtol=2e-7, atol=1e-11)


def test_rlax_lambda_q_estimate_identity():
    """rlax's `test_lambda_q_estimate` (vtrace_test.py), on its data: at lambda = 0.8 and clip_rho = 1 the V-trace targets
    satisfy vs = v + min(1, rho) * (q_estimate - v).  This is what fixes the q bootstrap as lambda * vs' + (1 - lambda) * v'
    (the lambda = 1 vectors above cannot tell)."""
    import rlax_vectors as rv
    T = torch.from_numpy
    rho, v_t = rv.rho_tm1(), rv.v_t()
    adv, err, q = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage, lambda_=0.8, clip_rho_threshold=1.0)(
        T(rv.VALUES), T(v_t), T(rv.REWARDS), T(rv.DISCOUNTS), T(rho))
    vs = err + T(rv.VALUES)
    vs_from_q = T(rv.VALUES) + torch.clamp(T(rho), max=1.0) * (q - T(rv.VALUES))
    np.testing.assert_allclose(vs.numpy(), vs_from_q.numpy(), rtol=1e-5, atol=1e-5)


def test_apex_td_loss_restatement():
    """N3b oracle (reference agents/dqn/learning.py:161-172): weights are max-normalised importance weights of the
    sampler's probabilities, constants for autograd; d loss / d q is -err * w / N at the taken action only."""
    g = torch.Generator().manual_seed(0)
    N, A, beta = 64, 6, 0.4
    q = torch.randn(N, A, generator=g, dtype=torch.float64, requires_grad=True)
    a = torch.randint(0, A, (N,), generator=g)
    tgt = torch.randn(N, generator=g, dtype=torch.float64)
    p = torch.rand(N, generator=g, dtype=torch.float64) * 0.01 + 1e-4
    loss, pri, m = losses_ref.apex_td_loss(q, a, tgt, p, beta)
    loss.backward()
    w = p.pow(-beta)
    w = w / w.max()
    err = tgt - q.detach().gather(1, a[:, None])[:, 0]
    want = torch.zeros(N, A, dtype=torch.float64)
    want[torch.arange(N), a] = -err * w / N
    assert torch.allclose(q.grad, want, rtol=1e-12, atol=1e-15)
    assert torch.equal(pri, err.abs()) and float(w.max()) == 1.0 and float(m["train_step/loss"]) == float(loss.detach())
    assert float(loss.detach()) == pytest.approx(float(0.5 * (err ** 2 * w).mean()), rel=1e-12)
